// Kernels of the entry types beyond Linear + activation:
// EmbeddingLayer (gather forward / scatter-add backward, neural_net_model.py:55-68,
// test_torch_backward.py:97-100) and BatchNormLayer (batch statistics with unbiased variance,
// running-statistics update, closed-form backward, neural_net_model.py:122-170,
// test_torch_backward.py:172-176).  FlattenLayer (:70-91) is a reinterpretation of the same
// row-major memory -- rows / ratio, features * ratio -- and needs no kernel.
#pragma once
#include "poz_common.cuh"

namespace poz {

// out[(b*tokens + t)][e] = C[idx[s(b)*tokens + t]][e], with the dropout of the hidden flatten
// entries that follow folded in.  s(b) = row_idx[b] when the batch is a gather of samples.
template <typename T>
__global__ void emb_gather_kernel(const T* __restrict__ C, const int* __restrict__ idx, const int* __restrict__ row_idx,
                                  long long batch, int tokens, int E, int vocab, T* __restrict__ out, DropPoint d,
                                  const DevState* st) {
    const long long total = batch * tokens * E;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / E;
        const int e = (int)(i - r * E);
        const long long b = r / tokens;
        const int t = (int)(r - b * tokens);
        const long long s = row_idx ? (long long)row_idx[b] : b;
        int v = idx[s * tokens + t];
        v = v < 0 ? 0 : (v >= vocab ? vocab - 1 : v);
        T x = C[(long long)v * E + e];
        if (d.enabled) x = drop_keep(d, st, r, e, E) ? x * T(d.scale) : T(0);
        out[i] = x;
    }
}

// dC[idx[..]][e] += dE[r][e]  (dE is already the gradient of the pre-dropout embedding output)
template <typename T>
__global__ void emb_scatter_kernel(const T* __restrict__ dE, const int* __restrict__ idx,
                                   const int* __restrict__ row_idx, long long batch, int tokens, int E, int vocab,
                                   T* __restrict__ dC) {
    const long long total = batch * tokens * E;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / E;
        const int e = (int)(i - r * E);
        const long long b = r / tokens;
        const int t = (int)(r - b * tokens);
        const long long s = row_idx ? (long long)row_idx[b] : b;
        int v = idx[s * tokens + t];
        v = v < 0 ? 0 : (v >= vocab ? vocab - 1 : v);
        atomicAdd(dC + (long long)v * E + e, dE[i]);
    }
}

// The same for tables small enough for shared memory (vocab * E * sizeof(T) <= EMB_SMEM_BYTES): every block
// accumulates its rows into a private copy of the table with shared-memory atomics and adds that copy to
// dC once -- a character-level table (27 x 10) otherwise takes a million global atomics on 270 addresses.
constexpr int EMB_SMEM_BYTES = 32768;
template <typename T>
__global__ void emb_scatter_small_kernel(const T* __restrict__ dE, const int* __restrict__ idx,
                                         const int* __restrict__ row_idx, long long batch, int tokens, int E, int vocab,
                                         T* __restrict__ dC) {
    __shared__ __align__(16) unsigned char tab_raw[EMB_SMEM_BYTES];
    T* tab = reinterpret_cast<T*>(tab_raw);
    const int ne = vocab * E;
    for (int i = threadIdx.x; i < ne; i += blockDim.x) tab[i] = T(0);
    __syncthreads();
    const long long total = batch * tokens * E;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / E;
        const int e = (int)(i - r * E);
        const long long b = r / tokens;
        const int t = (int)(r - b * tokens);
        const long long s = row_idx ? (long long)row_idx[b] : b;
        int v = idx[s * tokens + t];
        v = v < 0 ? 0 : (v >= vocab ? vocab - 1 : v);
        atomicAdd(tab + v * E + e, dE[i]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ne; i += blockDim.x)
        if (tab[i] != T(0)) atomicAdd(dC + i, tab[i]);
}

constexpr int BN_ROW_SPLITS = 128;   // row splits of the batchnorm column reductions

// Column reductions over the rows of a [R][N] matrix, two deterministic passes.  Pass 1: block
// (x, y) covers 32 columns and the rows y*8+warp, +8*gridDim.y, ...; each warp reads 32 consecutive
// columns of a row.  mode 0: partial sums of z and z^2; mode 1: partial sums of gy and gy * xhat.
template <typename T>
__global__ void bn_col_partial_kernel(const T* __restrict__ a, const T* __restrict__ z, const T* __restrict__ stat,
                                      long long R, int N, int mode, double* __restrict__ partial /*[gridDim.y][2][N]*/) {
    __shared__ double s1[8][32], s2[8][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.x * 32 + lane;
    double p = 0.0, q = 0.0;
    if (n < N) {
        const double mean = mode ? (double)stat[n] : 0.0, inv = mode ? (double)stat[N + n] : 0.0;
        for (long long r = (long long)blockIdx.y * 8 + warp; r < R; r += (long long)gridDim.y * 8) {
            const double v = (double)a[r * N + n];
            p += v;
            q += mode ? v * (((double)z[r * N + n] - mean) * inv) : v * v;
        }
    }
    s1[warp][lane] = p; s2[warp][lane] = q;
    __syncthreads();
    if (warp == 0 && n < N) {
        double sa = 0.0, sb = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { sa += s1[w][lane]; sb += s2[w][lane]; }
        partial[((long long)blockIdx.y * 2) * N + n] = sa;
        partial[((long long)blockIdx.y * 2 + 1) * N + n] = sb;
    }
}

// Pass 2 of the forward statistics: stat[n] = mean, stat[N+n] = 1/sqrt(var_unbiased + eps); with
// `update` the running statistics follow (neural_net_model.py:161-166); with `use_running` the running
// statistics are used instead (inference, :167-168).
template <typename T>
__global__ void bn_stats_final_kernel(const double* __restrict__ partial, int parts, long long R, int N,
                                      T* __restrict__ stat, T* __restrict__ run_mean, T* __restrict__ run_var,
                                      double eps, double momentum, int use_running, int update) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    if (use_running) {
        stat[n] = run_mean[n];
        stat[N + n] = (T)(1.0 / sqrt((double)run_var[n] + eps));
        return;
    }
    double sa = 0.0, sb = 0.0;
    for (int k = 0; k < parts; ++k) { sa += partial[((long long)k * 2) * N + n]; sb += partial[((long long)k * 2 + 1) * N + n]; }
    const double mean = sa / (double)R;
    const double var = (sb - (double)R * mean * mean) / (double)(R - 1);   // unbiased (:162)
    stat[n] = (T)mean;
    stat[N + n] = (T)(1.0 / sqrt(var + eps));
    if (update) {
        run_mean[n] = (T)((1.0 - momentum) * (double)run_mean[n] + momentum * mean);
        run_var[n] = (T)((1.0 - momentum) * (double)run_var[n] + momentum * var);
    }
}

// Pass 2 of the backward reductions: sums[n] = sum_r gy, sums[N+n] = sum_r gy * xhat; they are also
// d_bias and d_gain (test_torch_backward.py:172-173), written to grad_out = [d_gain; d_bias].
template <typename T>
__global__ void bn_bwd_final_kernel(const double* __restrict__ partial, int parts, int N, T* __restrict__ sums,
                                    T* __restrict__ grad_out) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    double sa = 0.0, sb = 0.0;
    for (int k = 0; k < parts; ++k) { sa += partial[((long long)k * 2) * N + n]; sb += partial[((long long)k * 2 + 1) * N + n]; }
    sums[n] = (T)sa;
    sums[N + n] = (T)sb;
    grad_out[n] = (T)sb;          // d_gain
    grad_out[N + n] = (T)sa;      // d_bias
}

// y = gain * (z - mean) * inv + bias -> dropout -> recorded y -> activation -> dropout -> recorded a
template <typename T>
__global__ void bn_apply_kernel(const T* __restrict__ z, const T* __restrict__ stat, const T* __restrict__ gain,
                                const T* __restrict__ bias, long long R, int N, int act, DropPoint dbn, DropPoint da,
                                const DevState* st, T* __restrict__ y_out, T* __restrict__ a_out) {
    const long long total = R * N;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / N;
        const int n = (int)(i - r * N);
        T y = gain[n] * ((z[i] - stat[n]) * stat[N + n]) + bias[n];
        if (dbn.enabled) y = drop_keep(dbn, st, r, n, N) ? y * T(dbn.scale) : T(0);
        if (y_out) y_out[i] = y;
        T a = act_fwd<T>(act, y);
        if (da.enabled) a = drop_keep(da, st, r, n, N) ? a * T(da.scale) : T(0);
        if (a_out) a_out[i] = a;
    }
}

// dz = gain*inv/R * (R*gy - sum(gy) - R/(R-1) * xhat * sum(gy*xhat))   (test_torch_backward.py:174-176),
// then the dropout on the Linear's own output.
template <typename T>
__global__ void bn_bwd_apply_kernel(const T* __restrict__ gy, const T* __restrict__ z, const T* __restrict__ stat,
                                    const T* __restrict__ gain, const T* __restrict__ sums, long long R, int N,
                                    DropPoint dz, const DevState* st, T* __restrict__ dz_rec, T* __restrict__ dz_pre) {
    const long long total = R * N;
    const T Rf = (T)R, corr = (T)((double)R / (double)(R - 1));
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / N;
        const int n = (int)(i - r * N);
        const T xhat = (z[i] - stat[n]) * stat[N + n];
        T g = gain[n] * stat[N + n] / Rf * (Rf * gy[i] - sums[n] - corr * xhat * sums[N + n]);
        if (dz_rec) dz_rec[i] = g;
        if (dz.enabled) g = drop_keep(dz, st, r, n, N) ? g * T(dz.scale) : T(0);
        dz_pre[i] = g;
    }
}

}  // namespace poz
