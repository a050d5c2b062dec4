// Non-GEMM kernels of the narrow path: per-step bookkeeping, loss (+ its gradient), softmax,
// fused gradient-reduce + L2 + Adam/SGD update, dtype conversion.
#pragma once
#include <cuda_bf16.h>
#include "poz_common.cuh"

namespace poz {

constexpr int MAX_LAYERS = 16;

constexpr int MAX_SEGS = 2 * MAX_LAYERS + 2;

// Parameter segments of the flat buffers.  Segments 0..n_linear-1 are the Linear layers' [W; b]
// images (segment index == layer index); an embedding table and batchnorm [gain; bias] pairs follow.
struct SegTable {
    int n;                           // segments
    int n_stat;                      // tensors the reference calls `weights` (update-ratio slots)
    long long off[MAX_SEGS];         // first flat element of the segment
    long long n_elems[MAX_SEGS];     // elements of the segment
    long long w_elems[MAX_SEGS];     // leading elements that get L2 and update-ratio statistics
    int splits[MAX_SEGS];            // gradient slabs to sum
    int stat_slot[MAX_SEGS];         // update-ratio slot, -1 if none
};

// element -> double, whatever the storage type of a diagnostics tensor
template <typename T> __device__ __forceinline__ double as_double(T v) { return (double)v; }
template <> __device__ __forceinline__ double as_double<__nv_bfloat16>(__nv_bfloat16 v) { return (double)__bfloat162float(v); }

// First kernel of every step: advances the epoch / Adam counters and derives the step's scalars
// (neural_net_model.py:464 lr decay; torch/optim/adam.py bias corrections).
static __global__ void step_begin_kernel(DevState* st, double* cost_ring, long long cost_cap,
                                  double* stats, long long stats_cap, int n_layers, double l2_share) {
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    long long e = st->next_epoch;
    st->l2_share = l2_share;
    st->cur_epoch = e;
    st->next_epoch = e + 1;
    double lr = st->lr0 * pow(st->decay, (double)e);
    st->lr_t = lr;
    if (st->optimizer == 1) {
        long long t = st->adam_t + 1;
        st->adam_t = t;
        double bc1 = 1.0 - pow(st->beta1, (double)t);
        double bc2 = 1.0 - pow(st->beta2, (double)t);
        st->step_size = lr / bc1;
        st->bc2_sqrt = sqrt(bc2);
    }
    cost_ring[e % cost_cap] = 0.0;
    long long rec = -1;
    if (st->progress_stride > 0 && (e % st->progress_stride) == 0 && st->n_records < stats_cap) {
        rec = st->n_records;
        st->n_records = rec + 1;
        for (int i = 0; i < n_layers * 4; ++i) stats[rec * n_layers * 4 + i] = 0.0;
    }
    st->cur_record = rec;
}

template <typename T> struct LossArgs {
    const T* z;            // CE: recorded logits [M][N];  MSE: recorded final output [M][N]
    const void* target;    // CE: int32 labels;  MSE: T [rows][N]
    const int* row_idx;    // gather of target rows (nullable)
    T* probs;              // CE: softmax output (nullable)
    T* dz_pre;             // gradient w.r.t. the last linear's pre-dropout output (nullable)
    T* dz_rec;             // diagnostics (nullable): gradient of the recorded linear output
    T* da_rec;             // diagnostics (nullable, MSE): gradient of the recorded final output
    long long M;
    int N;
    int act;               // MSE: activation of the last unit
    DropPoint dz;          // dropout point on the last linear's output
    double inv_count;      // CE: 1/global_batch;  MSE: 1/(global_batch*N)
    double* cost_base;
    long long cost_cap;
    const DevState* st;    // nullable: cost goes to cost_base[0]
};

__device__ __forceinline__ double* cost_slot(double* base, long long cap, const DevState* st) {
    return st ? base + (st->cur_epoch % cap) : base;
}

__device__ __forceinline__ void block_add_double(double v, double* dst) {
    __shared__ double red[32];
    v = warp_sum(v);
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) red[w] = v;
    __syncthreads();
    if (w == 0) {
        int nw = (blockDim.x + 31) >> 5;
        double s = l < nw ? red[l] : 0.0;
        s = warp_sum(s);
        if (l == 0 && dst && s != 0.0) atomicAdd(dst, s);   // (hundreds of blocks add to one address: skip the no-ops)
    }
    __syncthreads();
}

// softmax + cross-entropy + gradient, one warp per row
// (neural_net_model.py:188,400-403; gradient test_torch_backward.py:163-164)
template <typename T> __global__ void ce_loss_kernel(LossArgs<T> a) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    long long row = (long long)blockIdx.x * wpb + (threadIdx.x >> 5);
    const long long stride = (long long)gridDim.x * wpb;
    const int* labels = static_cast<const int*>(a.target);
    double local = 0.0;
    const T invb = T(a.inv_count);
    for (; row < a.M; row += stride) {
        const T* z = a.z + row * a.N;
        T mx = -INFINITY;
        for (int j = lane; j < a.N; j += 32) { T v = z[j]; mx = v > mx ? v : mx; }
        mx = warp_max(mx);
        T s = T(0);
        for (int j = lane; j < a.N; j += 32) s += exp(z[j] - mx);
        s = warp_sum(s);
        const T inv_s = T(1) / s;
        int y = -1;
        if (labels) {
            long long lr = a.row_idx ? (long long)a.row_idx[row] : row;
            y = labels[lr];
        }
        for (int j = lane; j < a.N; j += 32) {
            T p = exp(z[j] - mx) * inv_s;
            if (a.probs) a.probs[row * a.N + j] = p;
            if (a.dz_pre || a.dz_rec) {
                T g = (p - (j == y ? T(1) : T(0))) * invb;
                if (a.dz_rec) a.dz_rec[row * a.N + j] = g;
                if (a.dz_pre) {
                    if (a.dz.enabled) g = drop_keep(a.dz, a.st, row, j, a.N) ? g * T(a.dz.scale) : T(0);
                    a.dz_pre[row * a.N + j] = g;
                }
            }
        }
        if (labels && lane == 0 && y >= 0 && y < a.N)
            local += (double)(mx + log(s) - z[y]);
    }
    if (a.cost_base) block_add_double(local * a.inv_count, cost_slot(a.cost_base, a.cost_cap, a.st));
}

// plain row softmax for inference without a target (neural_net_model.py:188)
template <typename T> __global__ void softmax_kernel(const T* z, T* out, long long M, int N) {
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const int wpb = blockDim.x >> 5;
    long long row = (long long)blockIdx.x * wpb + (threadIdx.x >> 5);
    const long long stride = (long long)gridDim.x * wpb;
    for (; row < M; row += stride) {
        const T* zr = z + row * N;
        T mx = -INFINITY;
        for (int j = lane; j < N; j += 32) { T v = zr[j]; mx = v > mx ? v : mx; }
        mx = warp_max(mx);
        T s = T(0);
        for (int j = lane; j < N; j += 32) s += exp(zr[j] - mx);
        s = warp_sum(s);
        const T inv_s = T(1) / s;
        for (int j = lane; j < N; j += 32) out[row * N + j] = exp(zr[j] - mx) * inv_s;
    }
}

// the same for narrow rows (N <= 32): one thread per row, values held in registers
template <typename T> __global__ void softmax_small_kernel(const T* __restrict__ z, T* __restrict__ out, long long M, int N) {
    pdl_trigger();
    pdl_wait();
    const long long row = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= M) return;
    const T* zr = z + row * N;
    T v[32];
    T mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < 32; ++j)
        if (j < N) { v[j] = zr[j]; mx = v[j] > mx ? v[j] : mx; }
    T s = T(0);
#pragma unroll
    for (int j = 0; j < 32; ++j)
        if (j < N) { v[j] = exp(v[j] - mx); s += v[j]; }
    const T inv = T(1) / s;
#pragma unroll
    for (int j = 0; j < 32; ++j)
        if (j < N) out[row * N + j] = v[j] * inv;
}

// mean-squared error + gradient through the last activation (neural_net_model.py:405-406)
template <typename T> __global__ void mse_loss_kernel(LossArgs<T> a) {
    pdl_trigger();
    pdl_wait();
    const long long total = a.M * a.N;
    const T* tgt = static_cast<const T*>(a.target);
    double local = 0.0;
    const T two_inv = T(2.0 * a.inv_count);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        long long row = i / a.N;
        int j = (int)(i - row * a.N);
        long long tr = a.row_idx ? (long long)a.row_idx[row] : row;
        T y = a.z[i];
        T d = y - tgt[tr * a.N + j];
        local += (double)d * (double)d;
        if (a.dz_pre || a.dz_rec || a.da_rec) {
            T ga = two_inv * d;
            if (a.da_rec) a.da_rec[i] = ga;
            T gz = ga * act_bwd_from_out<T>(a.act, y);
            if (a.dz_rec) a.dz_rec[i] = gz;
            if (a.dz_pre) {
                if (a.dz.enabled) gz = drop_keep(a.dz, a.st, row, j, a.N) ? gz * T(a.dz.scale) : T(0);
                a.dz_pre[i] = gz;
            }
        }
    }
    if (a.cost_base) block_add_double(local * a.inv_count, cost_slot(a.cost_base, a.cost_cap, a.st));
}

// Sum the batch-split gradient slabs of every layer into slab 0 (data-parallel path: slab 0 is
// what the host all-reduces).
template <typename T> __global__ void grad_finalize_kernel(T* g, long long slab, SegTable seg) {
    const int l = blockIdx.y;
    const long long n = seg.n_elems[l];
    const int S = seg.splits[l];
    if (S <= 1) return;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        long long idx = seg.off[l] + i;
        T s = g[idx];
        for (int k = 1; k < S; ++k) s += g[(long long)k * slab + idx];
        g[idx] = s;
    }
}

// Reduce gradient slabs, add the L2 gradient, apply Adam / SGD in place, and fold the L2 penalty
// and the weight-update-ratio statistics into the same pass over the parameters.
//   Adam: torch/optim/adam.py _single_tensor_adam as driven by neural_net_model.py:497
//   SGD : neural_net_model.py:499-500
//   L2  : neural_net_model.py:482-483 (weights only)
//   ratio statistics: neural_net_model.py:506-507
template <typename T>
__global__ void update_kernel(T* __restrict__ p, T* __restrict__ am, T* __restrict__ av,
                              const T* __restrict__ g, long long slab, T* __restrict__ g_diag,
                              SegTable seg, const DevState* st, double* cost_base,
                              long long cost_cap, double* stats, __nv_bfloat16* __restrict__ p_bf16) {
    // (one element per thread and pass: with thousands of threads in flight and g marked read-only this
    // measured 11 us on the C2 shape against 15 us for the four-wide body the fused step uses)
    pdl_trigger();
    pdl_wait();
    const int l = blockIdx.y;
    fast::KernelTraceRec* ktr = fast::ktrace_open(128, (int)seg.n_elems[l], l, seg.splits[l]);   // trace tag 128: update
    const long long n = seg.n_elems[l];
    const long long nw = seg.w_elems[l];
    const int S = seg.splits[l];
    const bool adam = st->optimizer == 1;
    const T l2 = T(st->l2);
    const T lr = T(st->lr_t);
    const T b1c = T(1.0 - st->beta1), b2 = T(st->beta2), b2c = T(1.0 - st->beta2);
    const T step_size = T(st->step_size), bc2s = T(st->bc2_sqrt), eps = T(st->eps);
    const long long rec = st->cur_record;
    double l2acc = 0.0, sd = 0.0, sd2 = 0.0, sw = 0.0, sw2 = 0.0;
    if (ktr) ktr->t[1] = fast::ktrace_now();                 // step scalars read
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const long long idx = seg.off[l] + i;
        T gr = g[idx];
        for (int k = 1; k < S; ++k) gr += g[(long long)k * slab + idx];
        const T w = p[idx];
        const bool is_w = i < nw;
        if (is_w && st->l2 > 0.0) {
            l2acc += (double)w * (double)w;
            gr += T(2) * l2 * w;
        }
        if (g_diag) g_diag[idx] = gr;
        T wn;
        if (adam) {
            T m = am[idx], v = av[idx];
            m = m + (gr - m) * b1c;
            v = v * b2 + b2c * gr * gr;
            am[idx] = m;
            av[idx] = v;
            T denom = sqrt(v) / bc2s + eps;
            wn = w - step_size * (m / denom);
        } else {
            wn = w - lr * gr;
        }
        p[idx] = wn;
        if (p_bf16) p_bf16[idx] = __float2bfloat16_rn((float)wn);
        if (rec >= 0 && is_w) {
            double d = (double)wn - (double)w;
            sd += d; sd2 += d * d; sw += (double)wn; sw2 += (double)wn * (double)wn;
        }
    }
    if (ktr) ktr->t[2] = fast::ktrace_now();                 // elements done
    if (st->l2 > 0.0) block_add_double(l2acc * st->l2 * st->l2_share, cost_slot(cost_base, cost_cap, st));
    if (rec >= 0 && seg.stat_slot[l] >= 0) {
        double* s = stats + (rec * seg.n_stat + (seg.stat_slot[l] < 0 ? 0 : seg.stat_slot[l])) * 4;
        block_add_double(sd, s + 0);
        block_add_double(sd2, s + 1);
        block_add_double(sw, s + 2);
        block_add_double(sw2, s + 3);
    }
    if (ktr) { ktr->t[3] = fast::ktrace_now(); ktr->t[4] = ktr->t[3]; ktr->t[5] = ktr->t[3]; }
}

// Bias gradient db[n] = sum_b dz[b][n] for float32 dz (test_torch_backward.py:166), two
// deterministic passes: block (x, r) sums rows r*8+warp, +8*gridDim.y, ... of 64 columns into
// partial[r][n]; the second pass adds the gridDim.y partials in a fixed order.
static __global__ void colsum_f32_partial_kernel(const float* __restrict__ dz, long long rows, int N, float* __restrict__ partial) {
    __shared__ float red[8][64];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = blockIdx.x * 64 + lane, c1 = c0 + 32;
    float s0 = 0.f, s1 = 0.f;
    for (long long r = (long long)blockIdx.y * 8 + warp; r < rows; r += (long long)gridDim.y * 8) {
        if (c0 < N) s0 += dz[r * N + c0];
        if (c1 < N) s1 += dz[r * N + c1];
    }
    red[warp][lane] = s0;
    red[warp][lane + 32] = s1;
    __syncthreads();
    if (threadIdx.x < 64) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
        const int cc = blockIdx.x * 64 + threadIdx.x;
        if (cc < N) partial[(long long)blockIdx.y * N + cc] = t;
    }
}
static __global__ void colsum_f32_final_kernel(const float* __restrict__ partial, int parts, int N, float* __restrict__ out) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float t = 0.f;
    for (int r = 0; r < parts; ++r) t += partial[(long long)r * N + n];
    out[n] = t;
}

// L2 penalty for the inference-with-target path is never needed (compute_output has no L2).

// ---- device-side diagnostics (neural_net_model.py:532-583) -----------------------------------------
// Pass 1: per-block partial moments of a tensor: [sum, sum of squares, min, max, saturated count].
// rule 1: |x| > thr; rule 2: x <= 0; rule 3: row maximum > thr; rule 4: row L2 norm > thr (rules 3-4 count rows).
template <typename T>
__global__ void diag_moments_kernel(const T* __restrict__ x, long long rows, int cols, int rule, double thr,
                                    double* __restrict__ partial /*[gridDim.x][5]*/) {
    const long long n = rows * cols;
    double s = 0.0, s2 = 0.0, mn = INFINITY, mx = -INFINITY, sat = 0.0;
    if (rule == 3 || rule == 4) {
        const int lane = threadIdx.x & 31, wpb = blockDim.x >> 5;
        for (long long r = (long long)blockIdx.x * wpb + (threadIdx.x >> 5); r < rows; r += (long long)gridDim.x * wpb) {
            double rmax = -INFINITY, rsq = 0.0;
            for (int j = lane; j < cols; j += 32) {
                const double v = as_double<T>(x[r * cols + j]);
                s += v; s2 += v * v; mn = v < mn ? v : mn; mx = v > mx ? v : mx; rmax = v > rmax ? v : rmax;
                rsq += v * v;
            }
            rmax = warp_max(rmax);
            rsq = warp_sum(rsq);
            if (lane == 0 && (rule == 3 ? rmax > thr : sqrt(rsq) > thr)) sat += 1.0;
        }
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
            const double v = as_double<T>(x[i]);
            s += v; s2 += v * v; mn = v < mn ? v : mn; mx = v > mx ? v : mx;
            if (rule == 1) sat += (fabs(v) > thr) ? 1.0 : 0.0;
            else if (rule == 2) sat += (v <= 0.0) ? 1.0 : 0.0;
        }
    }
    __shared__ double red[5][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    s = warp_sum(s); s2 = warp_sum(s2); sat = warp_sum(sat);
    mn = -warp_max(-mn); mx = warp_max(mx);
    if (lane == 0) { red[0][w] = s; red[1][w] = s2; red[2][w] = mn; red[3][w] = mx; red[4][w] = sat; }
    __syncthreads();
    if (threadIdx.x == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        for (int k = 1; k < nw; ++k) {
            red[0][0] += red[0][k]; red[1][0] += red[1][k]; red[4][0] += red[4][k];
            red[2][0] = red[2][k] < red[2][0] ? red[2][k] : red[2][0];
            red[3][0] = red[3][k] > red[3][0] ? red[3][k] : red[3][0];
        }
        for (int k = 0; k < 5; ++k) partial[blockIdx.x * 5 + k] = red[k][0];
    }
}

// Pass 2: counts per bin for given bin edges (torch.histogram's own edges, computed by the host):
// bin i = [edges[i], edges[i+1]), the last bin closed on the right.
template <typename T>
__global__ void diag_histogram_kernel(const T* __restrict__ x, long long n, const double* __restrict__ edges, int bins,
                                      unsigned long long* __restrict__ counts) {
    __shared__ double e[129];
    __shared__ unsigned int h[128];
    for (int i = threadIdx.x; i <= bins; i += blockDim.x) e[i] = edges[i];
    for (int i = threadIdx.x; i < bins; i += blockDim.x) h[i] = 0u;
    __syncthreads();
    const double lo = e[0], scale = (double)bins / (e[bins] - e[0]);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double v = as_double<T>(x[i]);
        int pos = (int)((v - lo) * scale);
        pos = pos < 0 ? 0 : (pos > bins - 1 ? bins - 1 : pos);
        while (pos > 0 && v < e[pos]) --pos;
        while (pos < bins - 1 && v >= e[pos + 1]) ++pos;
        atomicAdd(&h[pos], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < bins; i += blockDim.x)
        if (h[i]) atomicAdd(&counts[i], (unsigned long long)h[i]);
}

template <typename S, typename D>
__global__ void convert_kernel(const S* __restrict__ src, D* __restrict__ dst, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        dst[i] = (D)src[i];
}

}  // namespace poz

namespace poz {

// ---- fused classifier head -----------------------------------------------------------------------------
// For stacks that end in a narrow Linear + softmax (out <= 32) one kernel does, per batch row, what would
// otherwise be three launches at a few percent occupancy: the last Linear (neural_net_model.py:117-119),
// softmax + cross-entropy + its gradient (:188,:400-403, test_torch_backward.py:163-164) and the
// backward-x of that Linear pushed through the previous unit's dropout / activation
// (test_torch_backward.py:168-170).  The last layer's [W; b] lives in shared memory (row pitch made odd
// so that the backward's column walk is bank-conflict free); one warp owns one row: lane n accumulates
// logit n over k with the row broadcast from shared memory, then lanes take k = lane, lane+32, ... for dA.
template <typename T> struct HeadArgs {
    const T* a_prev;       // recorded output of the previous unit [M][K]
    const T* wb;           // [K+1][N] = [W; b] of the last layer (bias row only if has_bias)
    const int* labels;     // int32 labels
    const int* row_idx;    // gather of label rows (nullable)
    T* z_rec;              // recorded logits (nullable)
    T* probs;              // softmax output (nullable)
    T* dz_pre;             // gradient of the last Linear's pre-dropout output [M][N]
    T* dz_rec;             // diagnostics (nullable)
    T* dprev_pre;          // gradient of the previous unit's pre-dropout linear output [M][K]
    T* da_rec;             // diagnostics (nullable): gradient of the previous unit's recorded activation
    T* dzprev_rec;         // diagnostics (nullable): gradient of the previous unit's recorded linear output
    long long M;
    int K, N, has_bias, act_prev;
    DropPoint dz_last, da_prev, dz_prev;
    double inv_batch;
    double* cost_base;
    long long cost_cap;
    const DevState* st;
    // optional: dz_pre / dprev_pre also as bf16 planes hi + mid + lo for a tensor-core weight-gradient launch
    // (poz_tcs_kernel.cuh): [3][M][ld], pad columns untouched
    unsigned short* dz_planes = nullptr;
    long long dz_ldp = 0, dz_pstride = 0;
    unsigned short* dprev_planes = nullptr;
    long long dprev_ldp = 0, dprev_pstride = 0;
};

// x -> its three bf16 planes at o, o + stride, o + 2 stride
template <typename T> __device__ __forceinline__ void store_planes3(T x, unsigned short* o, long long stride) {
    const float xv = (float)x;
    const __nv_bfloat16 h = __float2bfloat16_rn(xv);
    const float r1 = xv - __bfloat162float(h);
    const __nv_bfloat16 mm = __float2bfloat16_rn(r1);
    const __nv_bfloat16 ll = __float2bfloat16_rn(r1 - __bfloat162float(mm));
    o[0] = __bfloat16_as_ushort(h);
    o[stride] = __bfloat16_as_ushort(mm);
    o[2 * stride] = __bfloat16_as_ushort(ll);
}

// One row of the head for one warp: arow (shared memory, this warp's own K values, already the recorded output of
// the previous unit), W / bias staged in shared memory; returns the row's contribution to the cost sum in `local`.
template <typename T>
__device__ __forceinline__ void head_ce_row(const HeadArgs<T>& a, const T* W, int Ns, const T* bias, const T* arow,
                                            long long m, int y, int lane, T invb, double& local,
                                            fast::KernelTraceRec* ktr, bool stamp) {
    // logits: lane n < N owns logit n; four partial sums keep the dependent chain short.  (Splitting k over
    // the 32 lanes with a warp reduction per output was measured slower: 2.5 against 1.4 us of this kernel.)
    T z = bias[lane];
    if (lane < a.N) {
        T z1 = T(0), z2 = T(0), z3 = T(0);
        int k = 0;
        for (; k + 4 <= a.K; k += 4) {
            z  = fma(arow[k],     W[k * Ns + lane], z);
            z1 = fma(arow[k + 1], W[(k + 1) * Ns + lane], z1);
            z2 = fma(arow[k + 2], W[(k + 2) * Ns + lane], z2);
            z3 = fma(arow[k + 3], W[(k + 3) * Ns + lane], z3);
        }
        for (; k < a.K; ++k) z = fma(arow[k], W[k * Ns + lane], z);
        z = (z + z1) + (z2 + z3);
    }
    if (ktr && stamp) ktr->t[2] = fast::ktrace_now();   // logits of the first row
    if (a.dz_last.enabled && lane < a.N) z = drop_keep(a.dz_last, a.st, m, lane, a.N) ? z * T(a.dz_last.scale) : T(0);
    if (a.z_rec && lane < a.N) a.z_rec[m * a.N + lane] = z;
    T zm = lane < a.N ? z : T(-INFINITY);
    const T mx = warp_max(zm);
    const T e = lane < a.N ? exp(z - mx) : T(0);
    const T s = warp_sum(e);
    const T p = e / s;
    if (a.probs && lane < a.N) a.probs[m * a.N + lane] = p;
    
    T g = lane < a.N ? (p - (lane == y ? T(1) : T(0))) * invb : T(0);
    if (a.dz_rec && lane < a.N) a.dz_rec[m * a.N + lane] = g;
    if (a.dz_last.enabled && lane < a.N) g = drop_keep(a.dz_last, a.st, m, lane, a.N) ? g * T(a.dz_last.scale) : T(0);
    if (lane < a.N) a.dz_pre[m * a.N + lane] = g;
    if (a.dz_planes && lane < a.N) store_planes3<T>(g, a.dz_planes + m * a.dz_ldp + lane, a.dz_pstride);
    const T zy = __shfl_sync(0xffffffffu, z, (y >= 0 && y < a.N) ? y : 0);
    if (lane == 0 && y >= 0 && y < a.N) local += (double)(mx + log(s) - zy);
    if (ktr && stamp) ktr->t[3] = fast::ktrace_now();   // softmax / loss / dZ of the first row
    // backward-x: lanes take k = lane, lane + 32, ...; g of lane n is broadcast with shuffles, one
    // shuffle per n for four k-chunks at a time (each d[j] still adds its terms in the order n = 0, 1, ...)
    for (int kg = 0; kg < a.K; kg += 128) {
        T dd[4] = {T(0), T(0), T(0), T(0)};
        for (int n = 0; n < a.N; ++n) {
            const T gn = __shfl_sync(0xffffffffu, g, n);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int kk = kg + 32 * j + lane;
                if (kk < a.K) dd[j] = fma(gn, W[kk * Ns + n], dd[j]);
            }
        }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = kg + 32 * j + lane;
        const T d = dd[j];
        if (k < a.K) {
            if (a.da_rec) a.da_rec[m * a.K + k] = d;
            T ga = d, av = arow[k];
            if (a.da_prev.enabled) {
                ga = drop_keep(a.da_prev, a.st, m, k, a.K) ? d * T(a.da_prev.scale) : T(0);
                av = av * T(a.da_prev.keep_prob);
            }
            T gz = ga * act_bwd_from_out<T>(a.act_prev, av);
            if (a.dzprev_rec) a.dzprev_rec[m * a.K + k] = gz;
            if (a.dz_prev.enabled) gz = drop_keep(a.dz_prev, a.st, m, k, a.K) ? gz * T(a.dz_prev.scale) : T(0);
            a.dprev_pre[m * a.K + k] = gz;
            if (a.dprev_planes) store_planes3<T>(gz, a.dprev_planes + m * a.dprev_ldp + k, a.dprev_pstride);
        }
      }
    }
}

template <typename T>
__device__ __forceinline__ void head_ce_body(const HeadArgs<T>& a, unsigned char* head_smem, int part, int parts) {
    T* W = reinterpret_cast<T*>(head_smem);
    const int Ns = a.N | 1;                                  // odd row pitch
    T* bias = W + (long long)a.K * Ns;
    T* rows = bias + 32;                                     // [8 warps][K]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    T* arow = rows + (long long)warp * a.K;
    fast::KernelTraceRec* ktr = fast::ktrace_open(64, (int)a.M, a.N, a.K);    // trace tag 64: head
    // the first row of this warp and its label are requested before the weights, so that the three
    // global round trips overlap
    const long long m_first = (long long)part * wpb + warp;
    int y_next = -1;
    if (m_first < a.M) {
        const T* ap = a.a_prev + m_first * a.K;
        for (int k = lane; k < a.K; k += 32) arow[k] = ap[k];
        y_next = a.labels[a.row_idx ? (long long)a.row_idx[m_first] : m_first];
    }
    for (int i = threadIdx.x; i < a.K * a.N; i += blockDim.x) W[(i / a.N) * Ns + (i % a.N)] = a.wb[i];
    for (int i = threadIdx.x; i < 32; i += blockDim.x) bias[i] = (a.has_bias && i < a.N) ? a.wb[(long long)a.K * a.N + i] : T(0);
    __syncthreads();
    if (ktr) ktr->t[1] = fast::ktrace_now();                 // weights staged
    const T invb = T(a.inv_batch);
    double local = 0.0;
    for (long long m = m_first; m < a.M; m += (long long)parts * wpb) {
        if (m != m_first) {
            const T* ap = a.a_prev + m * a.K;
            for (int k = lane; k < a.K; k += 32) arow[k] = ap[k];
            y_next = a.labels[a.row_idx ? (long long)a.row_idx[m] : m];
        }
        __syncwarp();
        head_ce_row<T>(a, W, Ns, bias, arow, m, y_next, lane, invb, local, ktr, m == m_first);
        __syncwarp();
    }
    if (ktr) ktr->t[4] = fast::ktrace_now();                 // rows done
    if (a.cost_base) block_add_double(local * a.inv_batch, cost_slot(a.cost_base, a.cost_cap, a.st));
    if (ktr) ktr->t[5] = fast::ktrace_now();
}

template <typename T> __global__ void __launch_bounds__(256) head_ce_kernel(HeadArgs<T> a) {
    extern __shared__ __align__(16) unsigned char head_smem[];
    pdl_trigger();
    pdl_wait();
    head_ce_body<T>(a, head_smem, (int)blockIdx.x, (int)gridDim.x);
}

}  // namespace poz
