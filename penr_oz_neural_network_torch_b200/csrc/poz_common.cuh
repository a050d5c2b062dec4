// Shared device/host helpers for the pozmlp CUDA library (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>

namespace poz {

// ---- error plumbing: every failure becomes a message + non-zero return, never an abort -------
void set_error(const std::string& msg);
extern long long g_launches;   // kernels launched by this library (atomic add on host)
void count_launch(int n = 1);

#define POZ_CUDA(expr)                                                                       \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            poz::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) +       \
                           " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")");          \
            return 1;                                                                        \
        }                                                                                    \
    } while (0)

#define POZ_FAIL(msg)                                                                        \
    do {                                                                                     \
        poz::set_error(std::string(msg));                                                    \
        return 1;                                                                            \
    } while (0)

#define POZ_TRY(expr)                                                                        \
    do {                                                                                     \
        int _r = (expr);                                                                     \
        if (_r != 0) return _r;                                                              \
    } while (0)

enum Act : int { ACT_NONE = 0, ACT_RELU = 1, ACT_SIGMOID = 2, ACT_TANH = 3, ACT_SOFTMAX = 4 };

// ---- device-resident step state -----------------------------------------------------------------
// Everything that changes from step to step lives here so that the kernel arguments of a step are
// constant (the step can be replayed from a CUDA graph).  Written by poz_train_begin (host) and by
// step_begin_kernel (device, one thread, first kernel of every step).
struct DevState {
    // hyper-parameters of the current train() call
    double lr0, decay, beta1, beta2, eps, l2;
    double l2_share;           // share of the L2 penalty this rank adds to the cost (local/global batch)
    unsigned long long seed;
    long long progress_stride;
    int optimizer;
    int pad0;
    // counters
    long long next_epoch;      // epoch index the next step will use
    long long cur_epoch;       // epoch index of the step in flight
    long long adam_t;          // torch Adam `step`, persists across train() calls
    long long cur_record;      // progress record slot of this step, -1 if not recorded
    long long n_records;
    // derived per step
    double lr_t, step_size, bc2_sqrt;
};

// ---- Philox4x32-10 (counter-based RNG for in-kernel dropout) -------------------------------------
// One evaluation yields four 32-bit words: element e of a dropout point takes word e & 3 of the block with counter e >> 2,
// so an epilogue that holds four consecutive elements pays for one evaluation, not four.
__device__ __forceinline__ uint4 philox4(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = __umulhi(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return make_uint4(c0, c1, c2, c3);
}

// One dropout point (the output of one hidden entry).  keep(m, n) is a pure function of
// (seed, epoch, point, element) so the backward pass regenerates the forward mask.
struct DropPoint {
    const uint8_t* mask;   // host-recorded keep mask [M][N] or nullptr
    int enabled;           // 0 = identity
    uint32_t point;        // unique id of this dropout point inside the stack
    double keep_prob;      // 1 - p
    double scale;          // 1 / (1 - p)
};

// Out of line on purpose: the step's kernels are launch-bound and start with a cold instruction cache, and a
// Philox evaluation inlined at every epilogue call site (x 4 elements x 2 points) made each of them tens of
// kilobytes of code that a step without dropout fetches around.  A call keeps the common path compact.
__device__ __noinline__ static uint4 drop_block(const DropPoint& d, const DevState* st, long long q) {
    return philox4((uint32_t)q, (uint32_t)(q >> 32), d.point, (uint32_t)st->cur_epoch, (uint32_t)st->seed,
                   (uint32_t)(st->seed >> 32));
}
__device__ __forceinline__ bool drop_word_keeps(const DropPoint& d, uint32_t r) {
    return (r * 2.3283064365386963e-10f) < (float)d.keep_prob;
}
__device__ __noinline__ static bool drop_keep(const DropPoint& d, const DevState* st, long long m,
                                              int n, int ld) {
    long long e = m * (long long)ld + n;
    if (d.mask) return d.mask[e] != 0;
    const uint4 b = drop_block(d, st, e >> 2);
    const int w = (int)(e & 3);
    return drop_word_keeps(d, w == 0 ? b.x : w == 1 ? b.y : w == 2 ? b.z : b.w);
}
// the same for four consecutive elements (m, n .. n + 3): one Philox evaluation when they are one aligned block
__device__ __forceinline__ void drop_keep4(const DropPoint& d, const DevState* st, long long m, int n, int ld, bool (&keep)[4]) {
    const long long e = m * (long long)ld + n;
    if (!d.mask && (e & 3) == 0) {
        const uint4 b = drop_block(d, st, e >> 2);
        keep[0] = drop_word_keeps(d, b.x); keep[1] = drop_word_keeps(d, b.y);
        keep[2] = drop_word_keeps(d, b.z); keep[3] = drop_word_keeps(d, b.w);
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) keep[j] = drop_keep(d, st, m, n + j, ld);
    }
}

// ---- activations (neural_net_model.py:172-188) and their derivatives from the OUTPUT value -------
// relu / identity inline; the transcendental ones out of line (same reason as drop_keep: compact hot code)
__device__ __noinline__ static double act_fwd_slow(int act, double z) {
    return act == ACT_SIGMOID ? 1.0 / (1.0 + exp(-z)) : tanh(z);
}
__device__ __noinline__ static float act_fwd_slow(int act, float z) {
    return act == ACT_SIGMOID ? 1.f / (1.f + expf(-z)) : tanhf(z);
}
template <typename T> __device__ __forceinline__ T act_fwd(int act, T z) {
    if (act == ACT_RELU) return z < T(0) ? T(0) : z;   // a NaN stays a NaN, as in torch.relu
    if (act == ACT_SIGMOID || act == ACT_TANH) return act_fwd_slow(act, z);
    return z;
}
template <> __device__ __forceinline__ float act_fwd<float>(int act, float z) {
    if (act == ACT_RELU) return z < 0.f ? 0.f : z;      // (not fmaxf: a NaN must stay a NaN, as in torch.relu)
    if (act == ACT_SIGMOID || act == ACT_TANH) return act_fwd_slow(act, z);
    return z;
}
// derivative expressed through the activation output a = act(z)
template <typename T> __device__ __forceinline__ T act_bwd_from_out(int act, T a) {
    switch (act) {
        case ACT_RELU: return a > T(0) ? T(1) : T(0);
        case ACT_SIGMOID: return a * (T(1) - a);
        case ACT_TANH: return T(1) - a * a;     // test_torch_backward.py:170
        default: return T(1);
    }
}

// ---- programmatic dependent launch ------------------------------------------------------------------
// The kernels of a step form a dependent chain of launches that each fill the GPU for a few
// microseconds.  Launched with the programmatic-stream-serialization attribute, a kernel's CTAs are
// scheduled as soon as every CTA of its predecessor has passed pdl_trigger(), run whatever does not
// touch the predecessor's results, and block in pdl_wait() until the predecessor grid has completed
// and its memory is visible.  Both are no-ops for a launch without the attribute.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Debugging aid (POZ_KERNEL_TRACE=1): the CTAs of the step's kernels record %globaltimer at up to six points;
// poz_ctx_destroy prints, per launch, when the CTAs reached them.  The pointer is null in normal runs.  Internal
// linkage: every translation unit has its own copy and only poz_api.cu's is ever set.
namespace fast {
struct KernelTraceRec { int tag, m, n, k; unsigned long long t[6]; };
static __constant__ KernelTraceRec* d_ktrace = nullptr;   // constant bank: the null test at kernel entry costs no memory round trip
static __device__ unsigned int d_ktrace_n = 0;
constexpr unsigned int KTRACE_CAP = 1u << 20;
__device__ __forceinline__ unsigned long long ktrace_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// one record per CTA (thread 0), or null
__device__ __forceinline__ KernelTraceRec* ktrace_open(int tag, int m, int n, int k) {
    if (!d_ktrace || threadIdx.x != 0) return nullptr;
    const unsigned int slot = atomicAdd(&d_ktrace_n, 1u);
    if (slot >= KTRACE_CAP) return nullptr;
    KernelTraceRec* r = d_ktrace + slot;
    r->tag = tag; r->m = m; r->n = n; r->k = k;
    const unsigned long long now = ktrace_now();
    for (int q = 0; q < 6; ++q) r->t[q] = now;
    return r;
}
}  // namespace fast

template <typename T> __device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
template <typename T> __device__ __forceinline__ T warp_max(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        T u = __shfl_xor_sync(0xffffffffu, v, o);
        v = u > v ? u : v;
    }
    return v;
}

}  // namespace poz
