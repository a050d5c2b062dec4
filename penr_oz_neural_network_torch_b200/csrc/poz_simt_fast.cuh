// Pipelined fp32 SIMT GEMM for the narrow path (the C2 / C3 shapes): cp.async multi-stage
// shared-memory pipeline, and split-K across a thread-block cluster whose partial tiles are summed
// through distributed shared memory, so that a batch-1024 layer still puts several CTAs on each of
// the 148 SMs without a second kernel or a global scratch buffer.
//
// Same three operand arrangements as poz_simt_gemm.cuh (forward / backward-x / backward-w); the
// generic kernel there stays the fallback for float64 and for shapes this one does not take
// (k-contiguous operands need K % 4 == 0 and 16-byte aligned rows).
#pragma once
#include <cooperative_groups.h>
#include "poz_simt_gemm.cuh"

namespace poz {
namespace fast {

namespace cg = cooperative_groups;

constexpr int BK = 16;
constexpr int STAGES = 4;
constexpr int TM = 4, TN = 4;

struct Op {
    const float* ptr;
    long long ld;
    const int* row_idx;   // gather on the global rows (nullable)
    long long rows;       // valid global rows
    int cols;             // valid global columns
    int vec4;             // rows 16-byte aligned and cols % 4 == 0 -> 16-byte copies, else 4-byte copies
};

__device__ __forceinline__ void cp_async16(void* dst, const void* src, int bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)),
                 "l"(src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_async4(void* dst, const void* src, int bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)),
                 "l"(src), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// One operand tile -> shared memory.
//   KC  (k contiguous in global): S[MN][BK+4], global rows <-> mn, cols <-> k
//   !KC (mn contiguous in global): S[BK][MN+4], global rows <-> k,  cols <-> mn
template <int MN, int NT, bool KC, int PMN = 4>
__device__ __forceinline__ void load_tile(float* S, const Op& op, long long mn0, long long k0, long long k_lim,
                                          int tid) {
    constexpr int LD = KC ? BK + 4 : MN + PMN;
    constexpr int ROWS = KC ? MN : BK;
    constexpr int COLS = KC ? BK : MN;
    const long long row0 = KC ? mn0 : k0, col0 = KC ? k0 : mn0;
    const long long row_lim = KC ? op.rows : (k_lim < op.rows ? k_lim : op.rows);
    const long long col_lim = KC ? (k_lim < op.cols ? k_lim : (long long)op.cols) : (long long)op.cols;
    if (op.vec4) {
        constexpr int CV = COLS / 4;
        for (int c = tid; c < ROWS * CV; c += NT) {
            const int r = c / CV, cc = (c % CV) * 4;
            const long long gr = row0 + r, gc = col0 + cc;
            const bool ok = gr < row_lim && gc < col_lim;
            const float* src = op.ptr;
            if (ok) src = op.ptr + (op.row_idx ? (long long)op.row_idx[gr] : gr) * op.ld + gc;
            cp_async16(S + r * LD + cc, src, ok ? 16 : 0);
        }
    } else {
        for (int c = tid; c < ROWS * COLS; c += NT) {
            const int r = c / COLS, cc = c % COLS;
            const long long gr = row0 + r, gc = col0 + cc;
            const bool ok = gr < row_lim && gc < col_lim;
            const float* src = op.ptr;
            if (ok) src = op.ptr + (op.row_idx ? (long long)op.row_idx[gr] : gr) * op.ld + gc;
            cp_async4(S + r * LD + cc, src, ok ? 4 : 0);
        }
    }
}

// The same tile copy with everything that does not depend on the k-tile hoisted out of the main
// loop: each thread owns a fixed set of 16-byte chunks, so their shared-memory offsets, source
// pointers (at the first k-tile) and row validity are computed once; per k-tile only the k bound
// is re-checked and the pointer advanced.  Used whenever the operand allows 16-byte copies.
template <int MN, int NT, bool KC, int PMN = 4> struct TileCopier {
    static constexpr int LD = KC ? BK + 4 : MN + PMN;
    static constexpr int ROWS = KC ? MN : BK;
    static constexpr int CV = (KC ? BK : MN) / 4;
    static constexpr int N = (ROWS * CV + NT - 1) / NT;
    const float* src[N];
    int soff[N];
    int kofs[N];
    bool ok[N];
    long long step;      // elements to advance per k-tile
    long long k_lim;     // exclusive bound on the global k index
    long long k_first;
    // row gather along k (row-contiguous operand whose rows are the sampled batch rows: the first
    // layer's weight gradient): the row of the next k-tile is looked up while the current one is issued
    const int* gather;
    long long ld;
    int next_row[N];

    __device__ __forceinline__ void init(const Op& op, long long mn0, long long k_begin, long long k_end, int tid) {
        k_first = k_begin;
        gather = KC ? nullptr : op.row_idx;
        ld = op.ld;
        if (KC) { k_lim = k_end < op.cols ? k_end : (long long)op.cols; step = BK; }
        else    { k_lim = k_end < op.rows ? k_end : op.rows; step = (long long)BK * op.ld; }
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const int c = tid + i * NT;
            const int r = c / CV, cc = (c % CV) * 4;
            soff[i] = r * LD + cc;
            ok[i] = c < ROWS * CV;
            if (KC) {
                const long long gr = mn0 + r;
                ok[i] = ok[i] && gr < op.rows;
                kofs[i] = cc;
                src[i] = ok[i] ? op.ptr + (op.row_idx ? (long long)op.row_idx[gr] : gr) * op.ld + k_begin + cc : op.ptr;
            } else {
                const long long gc = mn0 + cc;
                ok[i] = ok[i] && gc < op.cols;
                kofs[i] = r;
                src[i] = ok[i] ? op.ptr + (gather ? gc : (k_begin + r) * op.ld + gc) : op.ptr;
                next_row[i] = (gather && ok[i] && k_begin + r < k_lim) ? gather[k_begin + r] : 0;
            }
        }
    }
    // k-tiles must be issued in order (the gather looks one tile ahead)
    __device__ __forceinline__ void issue(float* S, int t) {
        const long long kt = k_first + (long long)t * BK;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if ((ROWS * CV) % NT != 0 && tid_out_of_range(i)) continue;
            const bool v = ok[i] && (kt + kofs[i] < k_lim);
            const float* p = src[i];
            if (!KC && gather) {
                if (v) p = src[i] + (long long)next_row[i] * ld;
                const long long kn = kt + BK + kofs[i];
                if (ok[i] && kn < k_lim) next_row[i] = gather[kn];
            } else if (v) {
                p = src[i] + (long long)t * step;
            }
            cp_async16(S + soff[i], p, v ? 16 : 0);
        }
    }
    __device__ __forceinline__ bool tid_out_of_range(int i) const { return (int)threadIdx.x + i * NT >= ROWS * CV; }
};

// C[M][N] = A * B over this CTA's k range; z = blockIdx.z selects the range.  When the launch forms
// clusters along z the partial tiles are reduced through distributed shared memory and every CTA
// of the cluster runs the epilogue for its share of the tile rows; otherwise (dW) each z writes its
// own slab through the epilogue.
// COLSUM (backward-w only): CTAs of the first row-tile also produce the column sums of B (the bias
// gradient) as output row `bias_row`.
template <int BM, int BN, bool A_KC, bool B_KC, bool COLSUM, class Epi>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_f32_pipelined(Op A, Op B, long long M, int N, long long Kd, long long k_chunk, long long bias_row, Epi epi) {
    constexpr int NT = (BM / TM) * (BN / TN);
    constexpr int A_LD = A_KC ? BK + 4 : BM + 4;
    constexpr int B_LD = B_KC ? BK + 4 : BN + 4;
    constexpr int A_ELEMS = (A_KC ? BM : BK) * A_LD;
    constexpr int B_ELEMS = (B_KC ? BN : BK) * B_LD;
    constexpr int STAGE_ELEMS = A_ELEMS + B_ELEMS;
    constexpr int RED_ELEMS = BM * BN;
    constexpr int SMEM_ELEMS = STAGES * STAGE_ELEMS > RED_ELEMS ? STAGES * STAGE_ELEMS : RED_ELEMS;
    __shared__ __align__(16) float smem[SMEM_ELEMS];

    const int tid = threadIdx.x;
    KernelTraceRec* ktr = nullptr;
    if (d_ktrace && tid == 0) {
        const unsigned int slot = atomicAdd(&d_ktrace_n, 1u);
        if (slot < KTRACE_CAP) {
            ktr = d_ktrace + slot;
            ktr->tag = (A_KC ? 1 : 0) | (B_KC ? 2 : 0) | (COLSUM ? 4 : 0) | (BM == 64 ? 8 : 0);
            ktr->m = (int)M; ktr->n = N; ktr->k = (int)Kd;
            ktr->t[0] = ktrace_now();
        }
    }
    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    const long long m0 = (long long)blockIdx.y * BM;
    const int n0 = blockIdx.x * BN;
    const long long k_begin = (long long)blockIdx.z * k_chunk;
    const long long k_end = (k_begin + k_chunk < Kd) ? k_begin + k_chunk : Kd;
    const int nk = k_end > k_begin ? (int)((k_end - k_begin + BK - 1) / BK) : 0;

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;
    float bsum[TN] = {0.f, 0.f, 0.f, 0.f};

    pdl_wait();   // everything above is index arithmetic; the operands may come from the previous launch
    // hoisted copy descriptors when the operands allow it (block-uniform conditions)
    const bool a_fast = A.vec4 != 0;
    const bool b_fast = B.vec4 != 0;
    TileCopier<BM, NT, A_KC> ca;
    TileCopier<BN, NT, B_KC> cb;
    if (a_fast) ca.init(A, m0, k_begin, k_end, tid);
    if (b_fast) cb.init(B, n0, k_begin, k_end, tid);
    auto issue = [&](int t) {
        float* sa = smem + (t % STAGES) * STAGE_ELEMS;
        float* sb = sa + A_ELEMS;
        const long long k0 = k_begin + (long long)t * BK;
        if (a_fast) ca.issue(sa, t); else load_tile<BM, NT, A_KC>(sa, A, m0, k0, k_end, tid);
        if (b_fast) cb.issue(sb, t); else load_tile<BN, NT, B_KC>(sb, B, n0, k0, k_end, tid);
    };
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) issue(s);
        cp_commit();
    }
    if (ktr) ktr->t[1] = ktrace_now();
    for (int t = 0; t < nk; ++t) {
        cp_wait<STAGES - 2>();
        __syncthreads();
        if (ktr && t == 0) ktr->t[2] = ktrace_now();
        if (t + STAGES - 1 < nk) issue(t + STAGES - 1);
        cp_commit();
        const float* sa = smem + (t % STAGES) * STAGE_ELEMS;
        const float* sb = sa + A_ELEMS;
        // fragments of four k-values at a time, double-buffered in registers: the reads of block q+1 are
        // issued before the FMAs of block q, so the shared-memory latency sits under 64 FMAs per thread
        auto load_frag = [&](int k4, float (&a)[4][TM], float (&b)[4][TN]) {
            if (A_KC) {
#pragma unroll
                for (int i = 0; i < TM; ++i) {
                    const float4 v = *reinterpret_cast<const float4*>(sa + (ty * TM + i) * A_LD + k4);
                    a[0][i] = v.x; a[1][i] = v.y; a[2][i] = v.z; a[3][i] = v.w;
                }
            } else {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const float4 v = *reinterpret_cast<const float4*>(sa + (k4 + kk) * A_LD + ty * TM);
                    a[kk][0] = v.x; a[kk][1] = v.y; a[kk][2] = v.z; a[kk][3] = v.w;
                }
            }
            if (B_KC) {   // columns of this thread: tx, tx+16, tx+32, ... (conflict-free row reads)
#pragma unroll
                for (int j = 0; j < TN; ++j) {
                    const float4 v = *reinterpret_cast<const float4*>(sb + (tx + (BN / TN) * j) * B_LD + k4);
                    b[0][j] = v.x; b[1][j] = v.y; b[2][j] = v.z; b[3][j] = v.w;
                }
            } else {      // columns tx*4 .. tx*4+3
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const float4 v = *reinterpret_cast<const float4*>(sb + (k4 + kk) * B_LD + tx * TN);
                    b[kk][0] = v.x; b[kk][1] = v.y; b[kk][2] = v.z; b[kk][3] = v.w;
                }
            }
        };
        auto fma_frag = [&](const float (&a)[4][TM], const float (&b)[4][TN]) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[kk][i], b[kk][j], acc[i][j]);
                if (COLSUM) {
#pragma unroll
                    for (int j = 0; j < TN; ++j) bsum[j] += b[kk][j];
                }
            }
        };
        static_assert(BK == 16, "the fragment schedule below is written for four blocks of four k-values");
        float a0[4][TM], b0[4][TN], a1[4][TM], b1[4][TN];
        load_frag(0, a0, b0);
        load_frag(4, a1, b1);
        fma_frag(a0, b0);
        load_frag(8, a0, b0);
        fma_frag(a1, b1);
        load_frag(12, a1, b1);
        fma_frag(a0, b0);
        fma_frag(a1, b1);
    }
    cp_wait<0>();
    pdl_trigger();   // the main loop is done: let the next launch's CTAs be scheduled behind this one's epilogue
    if (ktr) ktr->t[3] = ktrace_now();

    cg::cluster_group cluster = cg::this_cluster();
    const unsigned csize = cluster.num_blocks();
    if (csize > 1) {
        // ---- split-K across the cluster: reduce partial tiles through distributed shared memory ----
        __syncthreads();                       // everyone is done with the pipeline buffers
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                const int col = B_KC ? tx + (BN / TN) * j : tx * TN + j;
                smem[(ty * TM + i) * BN + col] = acc[i][j];
            }
        cluster.sync();
        const unsigned rank = cluster.block_rank();
        const int rows_per = BM / (int)csize;   // csize divides BM (host guarantees)
        for (int idx = tid; idx < rows_per * (BN / 4); idx += NT) {
            const int row = (int)rank * rows_per + idx / (BN / 4);
            const int c4 = (idx % (BN / 4)) * 4;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            for (unsigned r = 0; r < csize; ++r) {
                const float* peer = cluster.map_shared_rank(smem, r);
                const float4 v = *reinterpret_cast<const float4*>(peer + row * BN + c4);
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            const long long m = m0 + row;
            const int n = n0 + c4;
            if (m < M && n < N) {
                const float vals[4] = {s.x, s.y, s.z, s.w};
                epi.template row<4>(m, n, vals, (N - n < 4) ? (N - n) : 4, 0);
            }
        }
        if (ktr) ktr->t[4] = ktrace_now();
        cluster.sync();                        // peers may still be reading this CTA's tile
        if (ktr) ktr->t[5] = ktrace_now();
        return;
    }

#pragma unroll
    for (int i = 0; i < TM; ++i) {
        const long long m = m0 + ty * TM + i;
        if (m >= M) continue;
        if (B_KC) {
#pragma unroll
            for (int j = 0; j < TN; ++j) {
                const int n = n0 + tx + (BN / TN) * j;
                if (n < N) {
                    const float one[1] = {acc[i][j]};
                    epi.template row<1>(m, n, one, 1, (int)blockIdx.z);
                }
            }
        } else {
            const int n = n0 + tx * TN;
            if (n < N) epi.template row<TN>(m, n, acc[i], (N - n < TN) ? (N - n) : TN, (int)blockIdx.z);
        }
    }
    if (COLSUM && blockIdx.y == 0 && ty == 0) {
        const int n = n0 + tx * TN;
        if (n < N) epi.template row<TN>(bias_row, n, bsum, (N - n < TN) ? (N - n) : TN, (int)blockIdx.z);
    }
    if (ktr) { ktr->t[4] = ktrace_now(); ktr->t[5] = ktr->t[4]; }
}

// forward epilogue with an explicit bias (this kernel has no implicit ones column)
struct FwdBiasEpi {
    const float* bias;   // nullable
    FwdEpi<float> inner;
    template <int TNN>
    __device__ __forceinline__ void row(long long m, int n, const float (&acc)[TNN], int nvalid, int z) const {
        float v[TNN];
#pragma unroll
        for (int j = 0; j < TNN; ++j) v[j] = (bias && j < nvalid) ? acc[j] + bias[n + j] : acc[j];
        inner.template row<TNN>(m, n, v, nvalid, z);
    }
};

}  // namespace fast
}  // namespace poz
