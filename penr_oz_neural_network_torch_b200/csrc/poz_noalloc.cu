// poz_forward_noalloc: batched inference that touches no allocator.
//
// Who this is for: a caller that is not allowed to obtain device memory at the moment it needs an
// answer.  Typical cases are a serving process whose framework has already handed the whole GPU to
// its own pool, so that a raw cudaMalloc from a library would fail or fight that pool, and a
// latency-bound loop that cannot afford the device-wide stall a cudaMalloc / cudaFree may imply.
// Such a caller cannot create a poz_ctx either (the context allocates parameters and workspace),
// so this entry point takes the weights as device pointers the caller already owns, and keeps the
// one thing the forward pass needs beyond its inputs and outputs -- the two buffers that the
// hidden activations ping-pong through -- in a fixed array that is part of the library image.
//
// Cost of that choice: every GPU on which this library is loaded carries the array
// (POZ_NOALLOC_SCRATCH_BYTES = 8 MiB) from load until the process exits, whether or not the entry
// point is ever called.
//
// Sharing: there is ONE array per device, shared by every call.  It still holds whatever the last
// call left in it.  Nothing here has to start at zero -- every scratch element a kernel reads was
// written earlier in the same call -- so no clearing is done.  Two calls that overlap on the same
// GPU (different streams, or different host threads) would write the same buffers at the same
// time and corrupt each other: the caller must serialise calls per device.
//
// Set-up: the first call on each device resolves the device address of the array.  That first
// call should be made before any stream capture is started on that device.
#include "../../include/pozmlp.h"
#include "poz_simt_gemm.cuh"

#include <mutex>

// The scratch itself, at file scope.  In host code this name is only a handle for the symbol, never
// a device address: the address is fetched once per device with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_poz_noalloc_scratch[POZ_NOALLOC_SCRATCH_BYTES];

namespace poz {
namespace noalloc {

constexpr long long SCRATCH_BYTES = POZ_NOALLOC_SCRATCH_BYTES;
constexpr long long HALF_ELEMS = SCRATCH_BYTES / 2 / (long long)sizeof(float);
constexpr int MAX_DEVICES = 64;

static std::mutex g_mu;
static float* g_scratch_addr[MAX_DEVICES] = {nullptr};

// Resolve (once per device, with that device current) where the scratch array lives on the device.
static int scratch_for_device(int device, float** out) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_poz_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            set_error(std::string("poz_forward_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
            return 1;
        }
        g_scratch_addr[device] = static_cast<float*>(p);
    }
    *out = g_scratch_addr[device];
    return 0;
}

// y = act(x @ W + b): bias added after the K-long sum, as neural_net_model.py:117-119 does.
struct BiasActEpi {
    const float* bias;   // may be null
    float* out;
    int ld;
    int act;
    template <int TN>
    __device__ __forceinline__ void row(long long m, int n, const float (&acc)[TN], int nvalid, int) const {
        float v[TN];
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            float z = acc[j];
            if (bias && j < nvalid) z += bias[n + j];
            v[j] = act_fwd<float>(act, z);
        }
        store_row<float, TN>(out + m * ld + n, v, nvalid);
    }
};

// row softmax, one warp per row (neural_net_model.py:188)
__global__ void row_softmax_kernel(const float* z, float* out, long long M, int N) {
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    long long row = (long long)blockIdx.x * wpb + (threadIdx.x >> 5);
    const long long stride = (long long)gridDim.x * wpb;
    for (; row < M; row += stride) {
        const float* zr = z + row * N;
        float mx = -INFINITY;
        for (int j = lane; j < N; j += 32) mx = fmaxf(mx, zr[j]);
        mx = warp_max(mx);
        float s = 0.f;
        for (int j = lane; j < N; j += 32) s += expf(zr[j] - mx);
        s = warp_sum(s);
        const float inv = 1.f / s;
        for (int j = lane; j < N; j += 32) out[row * N + j] = expf(zr[j] - mx) * inv;
    }
}

static Operand<float> operand(const float* p, long long ld, long long rows, int cols) {
    Operand<float> o;
    o.ptr = p; o.ld = ld; o.row_idx = nullptr; o.rows = rows; o.cols = cols; o.ones = 0;
    o.vec_ok = ((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (ld % 4) == 0) ? 1 : 0;
    return o;
}

static int layer(const float* x, const float* w, const float* b, long long M, int K, int N, int act,
                 float* out, cudaStream_t s) {
    Operand<float> A = operand(x, K, M, K);
    Operand<float> B = operand(w, N, K, N);
    BiasActEpi epi{b, out, N, act};
    if (N <= 16) {
        dim3 g((unsigned)((N + 15) / 16), (unsigned)((M + 63) / 64), 1);
        gemm_simt_kernel<float, 64, 16, 16, 4, 4, true, false, BiasActEpi><<<g, 64, 0, s>>>(A, B, M, N, K, K, epi);
    } else {
        dim3 g((unsigned)((N + 63) / 64), (unsigned)((M + 31) / 32), 1);
        gemm_simt_kernel<float, 32, 64, 16, 4, 4, true, false, BiasActEpi><<<g, 128, 0, s>>>(A, B, M, N, K, K, epi);
    }
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_error(std::string("poz_forward_noalloc: kernel launch failed: ") + cudaGetErrorString(e));
        return 1;
    }
    return 0;
}

}  // namespace noalloc
}  // namespace poz

extern "C" int poz_forward_noalloc(int device, int n_linear, const int* sizes, const int* act_codes,
                                   const float* const* weights_dev, const float* const* biases_dev,
                                   const float* x_dev, long long batch, float* out_dev, void* stream) {
    using namespace poz;
    using namespace poz::noalloc;
    if (!sizes || !act_codes || !weights_dev || !x_dev || !out_dev) POZ_FAIL("poz_forward_noalloc: NULL argument");
    if (n_linear < 1 || n_linear > 16) POZ_FAIL("poz_forward_noalloc: n_linear must be in [1,16]");
    if (device < 0 || device >= MAX_DEVICES) POZ_FAIL("poz_forward_noalloc: device index out of range");
    if (batch < 0) POZ_FAIL("poz_forward_noalloc: negative batch");
    int widest = 0;
    for (int l = 0; l <= n_linear; ++l) {
        if (sizes[l] < 1) POZ_FAIL("poz_forward_noalloc: layer sizes must be positive");
        if (l >= 1 && sizes[l] > widest) widest = sizes[l];
    }
    if (widest > POZ_NOALLOC_MAX_WIDTH)
        POZ_FAIL("poz_forward_noalloc: a layer is wider than POZ_NOALLOC_MAX_WIDTH (" +
                 std::to_string(POZ_NOALLOC_MAX_WIDTH) + " outputs); the fixed " +
                 std::to_string(SCRATCH_BYTES >> 20) + " MiB scratch cannot hold its activations -- use a poz_ctx");
    for (int l = 0; l < n_linear; ++l) {
        if (!weights_dev[l]) POZ_FAIL("poz_forward_noalloc: NULL weight pointer");
        if (act_codes[l] < ACT_NONE || act_codes[l] > ACT_SOFTMAX) POZ_FAIL("poz_forward_noalloc: bad activation code");
        if (act_codes[l] == ACT_SOFTMAX && l != n_linear - 1)
            POZ_FAIL("poz_forward_noalloc: softmax is supported as the last activation only");
    }
    if (batch == 0) return 0;
    // the caller's current device is put back on every way out (the stream must belong to `device`)
    struct DeviceGuard {
        int prev = -1;
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    POZ_CUDA(cudaGetDevice(&guard.prev));
    if (guard.prev == device) guard.prev = -1;
    else POZ_CUDA(cudaSetDevice(device));
    float* scratch = nullptr;
    POZ_TRY(scratch_for_device(device, &scratch));
    float* half[2] = {scratch, scratch + HALF_ELEMS};
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const long long chunk_cap = HALF_ELEMS / widest;          // >= 256 rows by the width limit
    const int n_out = sizes[n_linear];
    const bool softmax_last = act_codes[n_linear - 1] == ACT_SOFTMAX;
    for (long long r0 = 0; r0 < batch; r0 += chunk_cap) {
        const long long m = batch - r0 < chunk_cap ? batch - r0 : chunk_cap;
        const float* cur = x_dev + r0 * sizes[0];
        int flip = 0;
        for (int l = 0; l < n_linear; ++l) {
            const bool last = l == n_linear - 1;
            float* dst = (last && !softmax_last) ? out_dev + r0 * n_out : half[flip];
            const int act = act_codes[l] == ACT_SOFTMAX ? (int)ACT_NONE : act_codes[l];
            POZ_TRY(layer(cur, weights_dev[l], biases_dev ? biases_dev[l] : nullptr, m, sizes[l], sizes[l + 1],
                          act, dst, s));
            cur = dst;
            flip ^= 1;
        }
        if (softmax_last) {
            long long blocks = (m + 7) / 8;
            if (blocks > 148 * 8) blocks = 148 * 8;
            row_softmax_kernel<<<(unsigned)blocks, 256, 0, s>>>(cur, out_dev + r0 * n_out, m, n_out);
            count_launch();
            POZ_CUDA(cudaGetLastError());
        }
    }
    return 0;
}
