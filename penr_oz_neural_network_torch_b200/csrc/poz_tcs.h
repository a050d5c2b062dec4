// Small-batch tensor-core GEMM of the narrow train step ("tcs"): interface shared by poz_api.cu (caller) and
// poz_tc3.cu (kernel + launcher).  See poz_tcs_kernel.cuh for the scheme.
#pragma once
#include "poz_kernels.cuh"
#include "poz_simt_fast.cuh"
#include "poz_tc3.h"

namespace poz {
namespace tcs {

constexpr int BM = 128, BN = 64, BK = 64;
constexpr int MAX_KB = 4;                  // 64-deep k-blocks one CTA holds: B planes in shared, A planes in tensor memory, nothing recycled
constexpr int MAX_SPLITS = 8;              // cluster size along k
constexpr int MAX_GROUP = 4;               // problems of one kind in one launch

// C[M][N] = A'[M][K] * B[K][N], A' taken from an fp32 matrix `a` (row pitch lda, optional row gather):
//   a_transposed == 0:  A'[m][k] = a[row(m)][k]                    (forward: a = X, M = batch rows, K = features)
//   a_transposed == 1:  A'[m][k] = a[row(k)][m] for m < a_cols,    (weight gradient: M = features (+1), K = batch rows)
//                       1 for m == a_cols when ones_row (the bias gradient rides along as one more row), else 0
// B = bf16 planes [3][K][ld] (n contiguous).
struct Gemm {
    const float* a;
    long long lda;
    const int* row_idx;
    int a_transposed, a_cols, ones_row;
    long long M;
    int N;
    long long K;
    tc3::PlaneMat b;
    int splits;                            // 1, 2, 4 or 8; ceil(ceil(K / 64) / splits) <= MAX_KB
    int a_dep = 0, b_dep = 0;              // a / b are written by an earlier launch of the same step (see pdl below)
    int pdl = 0;                           // (first problem of a launch) launch with programmatic stream serialization: the CTAs
                                           // may start while the launch before is finishing; they wait before reading its results
};

// x0, x1 -> packed bf16x2 words of the hi / mid / lo planes (x0 in the low half); the same rounding as
// tc3::split3_pair (poz_tc3_gemm.cuh), restated here so that this header does not pull the tensor-core kernels in
__device__ __forceinline__ void split3_pair(float x0, float x1, uint32_t& ph, uint32_t& pm, uint32_t& pl) {
    __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
    ph = *reinterpret_cast<uint32_t*>(&h);
    const float r0 = x0 - __uint_as_float(ph << 16), r1 = x1 - __uint_as_float(ph & 0xffff0000u);
    __nv_bfloat162 m = __floats2bfloat162_rn(r0, r1);
    pm = *reinterpret_cast<uint32_t*>(&m);
    const float s0 = r0 - __uint_as_float(pm << 16), s1 = r1 - __uint_as_float(pm & 0xffff0000u);
    __nv_bfloat162 l = __floats2bfloat162_rn(s0, s1);
    pl = *reinterpret_cast<uint32_t*>(&l);
}

// smallest admissible cluster size for a contraction length, grown while the grid stays within ~one wave
int pick_splits(long long K, long long tiles, int sms);
bool shape_ok(long long M, int N, long long K);

// ---- epilogues: row<4>(m, n, values, nvalid, 0) per finished 4-wide piece of an output row, finish() once per thread ----
// An epilogue is a constant description (it stays in the kernel's parameter space) plus a small per-thread State.
struct NoState {};
struct FwdEpi {
    using State = NoState;
    fast::FwdBiasEpi inner;               // bias, dropout on z, activation, dropout on a, recorded outputs
    template <int TN>
    __device__ __forceinline__ void row(State&, long long m, int n, const float (&v)[TN], int nvalid, int z) const {
        inner.template row<TN>(m, n, v, nvalid, z);
    }
    __device__ __forceinline__ void prefetch(State&, long long, int, int, long long, int) const {}
    __device__ __forceinline__ void finish(State&) const {}
};

// plain store of the finished gradient [in+1][out] into the flat gradient image (slab 0)
struct StoreEpi {
    using State = NoState;
    float* g;
    int ld;
    template <int TN>
    __device__ __forceinline__ void row(State&, long long m, int n, const float (&v)[TN], int nvalid, int) const {
        store_row<float, TN>(g + m * ld + n, v, nvalid);
    }
    __device__ __forceinline__ void prefetch(State&, long long, int, int, long long, int) const {}
    __device__ __forceinline__ void finish(State&) const {}
};

// The optimizer applied where the gradient tile is finished (update_kernel's arithmetic, element for element):
// L2 gradient + penalty, Adam (torch single-tensor order) or SGD in place, update-ratio statistics, the
// diagnostics copy of the gradient, and the bf16 planes of the new weights for the next forward pass.
struct UpdateEpi {
    float* p; float* am; float* av;        // this layer's [W; b] image inside P / AM / AV
    float* g_diag;                         // nullable
    int ld;                                // out features
    long long w_rows;                      // rows < w_rows are weights (L2, statistics, planes); row w_rows is the bias
    const DevState* st;
    double* cost_base; long long cost_cap;
    double* stats; int n_stat, stat_slot;
    __nv_bfloat16* planes;                 // nullable: [3][w_rows][ldp]
    long long ldp, plane_stride;
    static constexpr int PF = 1;           // four-wide pieces per thread whose w / m / v are requested ahead of the gradient
    struct State {
        double l2acc = 0.0, sd = 0.0, sd2 = 0.0, sw = 0.0, sw2 = 0.0;   // per-thread partial sums
        float4 pf_w[PF], pf_m[PF], pf_v[PF];
        unsigned pf_have = 0;
    };

    // The pieces this thread will finish (the kernel's own assignment: piece i = index threadIdx.x + 512 i of the
    // rows_per x 16 pieces starting at row m0, column n0): load their parameters and moments now.
    __device__ __forceinline__ void prefetch(State& s, long long m0, int n0, int rows_per, long long M, int N) const {
        s.pf_have = 0;
        const bool adam = st->optimizer == 1;
#pragma unroll
        for (int i = 0; i < PF; ++i) {
            const int idx = (int)threadIdx.x + 512 * i;
            if (threadIdx.x >= 512 || idx >= rows_per * 16) continue;
            const long long m = m0 + idx / 16;
            const int n = n0 + (idx % 16) * 4;
            const long long e = m * ld + n;
            if (m < M && n + 4 <= N && (e & 3) == 0) {
                s.pf_w[i] = *reinterpret_cast<const float4*>(p + e);
                if (adam) {
                    s.pf_m[i] = *reinterpret_cast<const float4*>(am + e);
                    s.pf_v[i] = *reinterpret_cast<const float4*>(av + e);
                }
                s.pf_have |= 1u << i;
            }
        }
    }

    template <int TN>
    __device__ __forceinline__ void row(State& s, long long m, int n, const float (&gv)[TN], int nvalid, int slot) const {
        const bool adam = st->optimizer == 1;
        const float l2 = (float)st->l2, lr = (float)st->lr_t;
        const float b1c = (float)(1.0 - st->beta1), b2 = (float)st->beta2, b2c = (float)(1.0 - st->beta2);
        const float step_size = (float)st->step_size, bc2s = (float)st->bc2_sqrt, eps = (float)st->eps;
        const bool use_l2 = st->l2 > 0.0, rec = st->cur_record >= 0;
        const bool is_w = m < w_rows;
        const long long idx = m * ld + n;
        float w[TN], mo[TN], vo[TN], wn[TN];
        const bool vec = TN == 4 && nvalid == 4 && ((idx & 3) == 0);
        bool got = false;
        if (TN == 4) {
#pragma unroll
            for (int i = 0; i < PF; ++i)
                if (slot == i && ((s.pf_have >> i) & 1u)) {
                    w[0] = s.pf_w[i].x; w[1] = s.pf_w[i].y; w[2] = s.pf_w[i].z; w[3] = s.pf_w[i].w;
                    if (adam) {
                        mo[0] = s.pf_m[i].x; mo[1] = s.pf_m[i].y; mo[2] = s.pf_m[i].z; mo[3] = s.pf_m[i].w;
                        vo[0] = s.pf_v[i].x; vo[1] = s.pf_v[i].y; vo[2] = s.pf_v[i].z; vo[3] = s.pf_v[i].w;
                    }
                    got = true;
                }
        }
        if (got) {
        } else if (vec) {
            const float4 q = *reinterpret_cast<const float4*>(p + idx);
            w[0] = q.x; w[1] = q.y; w[2] = q.z; w[3] = q.w;
            if (adam) {
                const float4 a = *reinterpret_cast<const float4*>(am + idx);
                const float4 b = *reinterpret_cast<const float4*>(av + idx);
                mo[0] = a.x; mo[1] = a.y; mo[2] = a.z; mo[3] = a.w;
                vo[0] = b.x; vo[1] = b.y; vo[2] = b.z; vo[3] = b.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                w[j] = j < nvalid ? p[idx + j] : 0.f;
                mo[j] = (adam && j < nvalid) ? am[idx + j] : 0.f;
                vo[j] = (adam && j < nvalid) ? av[idx + j] : 0.f;
            }
        }
        float gu[TN];
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            gu[j] = gv[j];
            if (is_w && use_l2 && j < nvalid) {
                s.l2acc += (double)w[j] * (double)w[j];
                gu[j] += 2.f * l2 * w[j];
            }
            if (adam) {
                mo[j] = mo[j] + (gu[j] - mo[j]) * b1c;
                vo[j] = vo[j] * b2 + b2c * gu[j] * gu[j];
                const float denom = sqrtf(vo[j]) / bc2s + eps;
                wn[j] = w[j] - step_size * (mo[j] / denom);
            } else {
                wn[j] = w[j] - lr * gu[j];
            }
            if (rec && is_w && j < nvalid) {
                const double d = (double)wn[j] - (double)w[j];
                s.sd += d; s.sd2 += d * d; s.sw += (double)wn[j]; s.sw2 += (double)wn[j] * (double)wn[j];
            }
        }
        if (g_diag) store_row<float, TN>(g_diag + idx, gu, nvalid);
        store_row<float, TN>(p + idx, wn, nvalid);
        if (adam) {
            store_row<float, TN>(am + idx, mo, nvalid);
            store_row<float, TN>(av + idx, vo, nvalid);
        }
        if (planes && is_w) {
            __nv_bfloat16* o = planes + m * ldp + n;
            if (TN == 4 && nvalid == 4) {
                uint32_t h0, m0, l0, h1, m1, l1;
                split3_pair(wn[0], wn[1], h0, m0, l0);
                split3_pair(wn[2], wn[3], h1, m1, l1);
                *reinterpret_cast<uint2*>(o) = make_uint2(h0, h1);
                *reinterpret_cast<uint2*>(o + plane_stride) = make_uint2(m0, m1);
                *reinterpret_cast<uint2*>(o + 2 * plane_stride) = make_uint2(l0, l1);
            } else {
#pragma unroll
                for (int j = 0; j < TN; ++j)
                    if (j < nvalid) {
                        const __nv_bfloat16 h = __float2bfloat16_rn(wn[j]);
                        const float r1 = wn[j] - __bfloat162float(h);
                        const __nv_bfloat16 mm = __float2bfloat16_rn(r1);
                        const __nv_bfloat16 ll = __float2bfloat16_rn(r1 - __bfloat162float(mm));
                        o[j] = h; o[plane_stride + j] = mm; o[2 * plane_stride + j] = ll;
                    }
            }
        }
    }
    // every thread of the block, once, after its last row()
    __device__ __forceinline__ void finish(State& s) const {
        if (st->l2 > 0.0) block_add_double(s.l2acc * st->l2 * st->l2_share, cost_slot(cost_base, cost_cap, st));
        const long long rec = st->cur_record;
        if (rec >= 0 && stat_slot >= 0) {
            double* q = stats + (rec * n_stat + stat_slot) * 4;
            block_add_double(s.sd, q + 0);
            block_add_double(s.sd2, q + 1);
            block_add_double(s.sw, q + 2);
            block_add_double(s.sw2, q + 3);
        }
    }
};

int trace_enable();      // POZ_KERNEL_TRACE=1
void trace_dump();
int forward(const Gemm& g, const FwdEpi& epi, cudaStream_t stream);
// n <= MAX_GROUP weight gradients of one contraction length (g[i].K and g[i].splits all equal) in ONE launch
int backward_weight_store(const Gemm* g, const StoreEpi* epi, int n, cudaStream_t stream);
int backward_weight_update(const Gemm* g, const UpdateEpi* epi, int n, cudaStream_t stream);

}  // namespace tcs
}  // namespace poz
