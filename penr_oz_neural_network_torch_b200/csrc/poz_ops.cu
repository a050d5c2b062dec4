// Unit operators: one entry type of the reference's layer list per call (include/pozmlp.h, "unit operators").
// They are what Layer.forward / LinearLayer.forward / EmbeddingLayer.forward / BatchNormLayer.forward and the
// activation layers' forward of the drop-in class run (neural_net_model.py:48-53,67-68,110-120,158-170,172-188),
// and what NeuralNetworkModel._forward chains when it is asked for every entry's output without a target.
// Stateless: device pointers in, device pointers out, float32 or float64, nothing allocated, nothing kept.
#include "../../include/pozmlp.h"
#include "poz_kernels.cuh"
#include "poz_net_kernels.cuh"
#include "poz_simt_gemm.cuh"

#include <algorithm>

namespace poz {
namespace ops {

// y = x @ W, then += b (neural_net_model.py:117-119)
template <typename T> struct BiasEpi {
    const T* bias;   // may be null
    T* out;
    int ld;
    template <int TN>
    __device__ __forceinline__ void row(long long m, int n, const T (&acc)[TN], int nvalid, int) const {
        T v[TN];
#pragma unroll
        for (int j = 0; j < TN; ++j) v[j] = (bias && j < nvalid) ? acc[j] + bias[n + j] : acc[j];
        store_row<T, TN>(out + m * ld + n, v, nvalid);
    }
};

template <typename T> static Operand<T> operand(const T* p, long long ld, long long rows, int cols) {
    Operand<T> o;
    o.ptr = p; o.ld = ld; o.row_idx = nullptr; o.rows = rows; o.cols = cols; o.ones = 0;
    o.vec_ok = ((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (ld % Vec16<T>::V) == 0) ? 1 : 0;
    return o;
}

template <typename T> struct Tile;
template <> struct Tile<float> { static constexpr int BK = 16, TN_WIDE = 4, TN_THIN = 4; };
template <> struct Tile<double> { static constexpr int BK = 8, TN_WIDE = 2, TN_THIN = 2; };

template <typename T>
static int linear(const T* x, const T* w, const T* b, long long M, int K, int N, T* y, cudaStream_t s) {
    Operand<T> A = operand<T>(x, K, M, K);
    Operand<T> B = operand<T>(w, N, K, N);
    BiasEpi<T> epi{b, y, N};
    constexpr int BK = Tile<T>::BK;
    if (N <= 16) {
        constexpr int TN = Tile<T>::TN_THIN;
        dim3 g((unsigned)((N + 15) / 16), (unsigned)((M + 63) / 64), 1);
        if (g.y > 65535) POZ_FAIL("poz_op_linear: too many rows for one launch");
        gemm_simt_kernel<T, 64, 16, BK, 4, TN, true, false, BiasEpi<T>><<<g, (64 / 4) * (16 / TN), 0, s>>>(A, B, M, N, K, K, epi);
    } else {
        constexpr int BN = sizeof(T) == 8 ? 32 : 64;
        constexpr int TN = Tile<T>::TN_WIDE;
        dim3 g((unsigned)((N + BN - 1) / BN), (unsigned)((M + 31) / 32), 1);
        if (g.y > 65535) POZ_FAIL("poz_op_linear: too many rows for one launch");
        gemm_simt_kernel<T, 32, BN, BK, 4, TN, true, false, BiasEpi<T>><<<g, (32 / 4) * (BN / TN), 0, s>>>(A, B, M, N, K, K, epi);
    }
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

template <typename T> __global__ void act_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, int act) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        y[i] = act_fwd<T>(act, x[i]);
}

// BatchNormLayer.forward (neural_net_model.py:158-170).  Block = 32 columns x 8 row lanes (row-major reads of 32
// consecutive columns coalesce); statistics in double, two passes (mean, then centred sum of squares -- the
// unbiased variance torch.var gives).  scalar != 0: ONE mean / variance over every element, broadcast to all
// columns -- what the reference computes for a 1-D input in training mode (dim = () reduces over everything).
template <typename T>
__global__ void __launch_bounds__(256) bn_unit_kernel(const T* __restrict__ x, long long rows, int cols, const T* __restrict__ gain,
                                                      const T* __restrict__ bias, T* __restrict__ run_mean,
                                                      T* __restrict__ run_var, int training, int scalar, double eps,
                                                      double momentum, T* __restrict__ y) {
    __shared__ double red[8][33];
    __shared__ double s_mean[32], s_var[32];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = blockIdx.x * 32 + tx;
    const bool live = col < cols;
    if (training) {
        if (!scalar) {
            double s = 0.0;
            if (live) for (long long r = ty; r < rows; r += 8) s += (double)x[r * cols + col];
            red[ty][tx] = s;
            __syncthreads();
            if (ty == 0) {
                double t = 0.0;
                for (int q = 0; q < 8; ++q) t += red[q][tx];
                s_mean[tx] = t / (double)rows;
            }
            __syncthreads();
            const double mu = s_mean[tx];
            double v = 0.0;
            if (live) for (long long r = ty; r < rows; r += 8) { const double d = (double)x[r * cols + col] - mu; v += d * d; }
            red[ty][tx] = v;
            __syncthreads();
            if (ty == 0) {
                double t = 0.0;
                for (int q = 0; q < 8; ++q) t += red[q][tx];
                s_var[tx] = t / (double)(rows - 1);          // rows == 1 -> NaN, as torch.var
            }
            __syncthreads();
        } else {
            // one statistic over all rows * cols elements (single block launch)
            const long long n = rows * cols;
            double s = 0.0;
            for (long long i = threadIdx.x; i < n; i += 256) s += (double)x[i];
            red[ty][tx] = s;
            __syncthreads();
            if (threadIdx.x == 0) {
                double t = 0.0;
                for (int a = 0; a < 8; ++a) for (int c = 0; c < 32; ++c) t += red[a][c];
                s_mean[0] = t / (double)n;
            }
            __syncthreads();
            const double mu = s_mean[0];
            double v = 0.0;
            for (long long i = threadIdx.x; i < n; i += 256) { const double d = (double)x[i] - mu; v += d * d; }
            __syncthreads();
            red[ty][tx] = v;
            __syncthreads();
            if (threadIdx.x == 0) {
                double t = 0.0;
                for (int a = 0; a < 8; ++a) for (int c = 0; c < 32; ++c) t += red[a][c];
                s_var[0] = t / (double)(n - 1);
            }
            __syncthreads();
        }
    }
    if (scalar) {
        const double mu = training ? s_mean[0] : 0.0, var = training ? s_var[0] : 1.0;
        for (int c = threadIdx.x; c < cols; c += 256) {
            const double m_use = training ? mu : (double)run_mean[c], v_use = training ? var : (double)run_var[c];
            const double inv = 1.0 / sqrt(v_use + eps);
            for (long long r = 0; r < rows; ++r)
                y[r * cols + c] = (T)((double)gain[c] * ((double)x[r * cols + c] - m_use) * inv + (double)bias[c]);
        }
        __syncthreads();
        if (training)
            for (int c = threadIdx.x; c < cols; c += 256) {
                run_mean[c] = (T)((1.0 - momentum) * (double)run_mean[c] + momentum * mu);
                run_var[c] = (T)((1.0 - momentum) * (double)run_var[c] + momentum * var);
            }
        return;
    }
    if (!live) return;
    const double mu = training ? s_mean[tx] : (double)run_mean[col];
    const double var = training ? s_var[tx] : (double)run_var[col];
    const double inv = 1.0 / sqrt(var + eps);
    const double g = (double)gain[col], b = (double)bias[col];
    for (long long r = ty; r < rows; r += 8)
        y[r * cols + col] = (T)(g * ((double)x[r * cols + col] - mu) * inv + b);
    if (training && ty == 0) {                               // running statistics (:164-166)
        run_mean[col] = (T)((1.0 - momentum) * (double)run_mean[col] + momentum * mu);
        run_var[col] = (T)((1.0 - momentum) * (double)run_var[col] + momentum * var);
    }
}

static int enter(int device, int dtype_code, const char* who) {
    if (dtype_code != POZ_DTYPE_F32 && dtype_code != POZ_DTYPE_F64)
        POZ_FAIL(std::string(who) + ": unit operators take POZ_DTYPE_F32 or POZ_DTYPE_F64");
    int ndev = 0;
    POZ_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) POZ_FAIL(std::string(who) + ": no such CUDA device");
    POZ_CUDA(cudaSetDevice(device));
    return 0;
}

}  // namespace ops
}  // namespace poz

using namespace poz;

extern "C" {

int poz_op_linear(int device, int dtype_code, const void* x_dev, const void* w_dev, const void* b_dev, long long rows,
                  int in, int out, void* y_dev, void* stream) {
    POZ_TRY(ops::enter(device, dtype_code, "poz_op_linear"));
    if (rows < 0 || in < 1 || out < 1) POZ_FAIL("poz_op_linear: bad sizes");
    if (rows == 0) return 0;
    if (!x_dev || !w_dev || !y_dev) POZ_FAIL("poz_op_linear: NULL buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (dtype_code == POZ_DTYPE_F64)
        return ops::linear<double>(static_cast<const double*>(x_dev), static_cast<const double*>(w_dev),
                                   static_cast<const double*>(b_dev), rows, in, out, static_cast<double*>(y_dev), s);
    return ops::linear<float>(static_cast<const float*>(x_dev), static_cast<const float*>(w_dev),
                              static_cast<const float*>(b_dev), rows, in, out, static_cast<float*>(y_dev), s);
}

int poz_op_activation(int device, int dtype_code, int act, const void* x_dev, long long rows, int cols, void* y_dev,
                      void* stream) {
    POZ_TRY(ops::enter(device, dtype_code, "poz_op_activation"));
    if (act < ACT_NONE || act > ACT_SOFTMAX) POZ_FAIL("poz_op_activation: bad activation code");
    if (rows < 0 || cols < 1) POZ_FAIL("poz_op_activation: bad sizes");
    if (rows == 0) return 0;
    if (!x_dev || !y_dev) POZ_FAIL("poz_op_activation: NULL buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const long long n = rows * cols;
    if (act == ACT_SOFTMAX) {                                // over the last dimension (:188)
        const int blocks = (int)std::max<long long>(1, std::min<long long>((rows + 7) / 8, 148 * 8));
        if (dtype_code == POZ_DTYPE_F64)
            softmax_kernel<double><<<blocks, 256, 0, s>>>(static_cast<const double*>(x_dev), static_cast<double*>(y_dev), rows, cols);
        else
            softmax_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(x_dev), static_cast<float*>(y_dev), rows, cols);
    } else {
        const int blocks = (int)std::max<long long>(1, std::min<long long>((n + 255) / 256, 148 * 8));
        if (dtype_code == POZ_DTYPE_F64)
            ops::act_kernel<double><<<blocks, 256, 0, s>>>(static_cast<const double*>(x_dev), static_cast<double*>(y_dev), n, act);
        else
            ops::act_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(x_dev), static_cast<float*>(y_dev), n, act);
    }
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

int poz_op_embedding(int device, int dtype_code, const void* table_dev, int vocab, int dim, const int* idx_dev,
                     long long n, void* y_dev, void* stream) {
    POZ_TRY(ops::enter(device, dtype_code, "poz_op_embedding"));
    if (vocab < 1 || dim < 1 || n < 0) POZ_FAIL("poz_op_embedding: bad sizes");
    if (n == 0) return 0;
    if (!table_dev || !idx_dev || !y_dev) POZ_FAIL("poz_op_embedding: NULL buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int blocks = (int)std::max<long long>(1, std::min<long long>((n * dim + 255) / 256, 148 * 8));
    DropPoint none{};
    if (dtype_code == POZ_DTYPE_F64)
        emb_gather_kernel<double><<<blocks, 256, 0, s>>>(static_cast<const double*>(table_dev), idx_dev, nullptr, n, 1, dim, vocab,
                                                         static_cast<double*>(y_dev), none, nullptr);
    else
        emb_gather_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(table_dev), idx_dev, nullptr, n, 1, dim, vocab,
                                                        static_cast<float*>(y_dev), none, nullptr);
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

int poz_op_batchnorm(int device, int dtype_code, int training, int scalar_stats, const void* x_dev, long long rows, int cols,
                     const void* gain_dev, const void* bias_dev, void* run_mean_dev, void* run_var_dev, double eps,
                     double momentum, void* y_dev, void* stream) {
    POZ_TRY(ops::enter(device, dtype_code, "poz_op_batchnorm"));
    if (rows < 0 || cols < 1) POZ_FAIL("poz_op_batchnorm: bad sizes");
    if (rows == 0) return 0;
    if (!x_dev || !gain_dev || !bias_dev || !run_mean_dev || !run_var_dev || !y_dev) POZ_FAIL("poz_op_batchnorm: NULL buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int blocks = scalar_stats ? 1 : (cols + 31) / 32;
    if (dtype_code == POZ_DTYPE_F64)
        ops::bn_unit_kernel<double><<<blocks, 256, 0, s>>>(static_cast<const double*>(x_dev), rows, cols,
                                                           static_cast<const double*>(gain_dev), static_cast<const double*>(bias_dev),
                                                           static_cast<double*>(run_mean_dev), static_cast<double*>(run_var_dev),
                                                           training, scalar_stats, eps, momentum, static_cast<double*>(y_dev));
    else
        ops::bn_unit_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(x_dev), rows, cols,
                                                          static_cast<const float*>(gain_dev), static_cast<const float*>(bias_dev),
                                                          static_cast<float*>(run_mean_dev), static_cast<float*>(run_var_dev),
                                                          training, scalar_stats, eps, momentum, static_cast<float*>(y_dev));
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
