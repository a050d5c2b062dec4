// The fused train step of a narrow fp32 stack: ONE persistent kernel per step.
//
// A batch-1024 step of the 784-256-128-10 stack is ~1 GFLOP spread over nine dependent launches,
// each of which fills a fraction of the GPU for a few microseconds: launch gaps, pipeline fill and
// tail effects cost more than the arithmetic.  Here one cooperative grid (one CTA of 8 warps per
// SM) walks a host-built task table phase by phase,
//
//     forward(0) | forward(1) | ... | head: last Linear + softmax/CE + its backward-x
//     | backward-x(l) + the weight gradients whose dZ is ready | ... | weight gradient(0) + updates | update(0)
//
// with a grid-wide barrier between phases instead of a kernel boundary.  Every phase is a list of
// tile tasks (32x64 output tile, or a row range / parameter range for the head and the update) that
// the CTAs take round-robin, so independent GEMMs of a phase share the machine, and the step's
// bookkeeping (epoch / Adam counters, cost slot) is folded into the first phase.
//
// The tile GEMM is a cp.async pipeline (four stages of 64 k-values) like poz_simt_fast.cuh's, re-cut
// so that the shared-memory reads stop being the limit: every warp computes the whole 32x64 tile for
// an eighth of each stage's k range with 8x8 outputs per thread, and the eight partial tiles are
// added through shared memory before the epilogue.  The epilogues are the ones of the separate
// kernels (FwdBiasEpi, BwdDataEpi, BwdWeightEpi), so the arithmetic per output is the same code.
//
// Reference semantics: neural_net_model.py:384-396 (forward), :400-403 (loss), :466-500 (backward,
// L2, optimizer); gradient formulas test_torch_backward.py:163-170.
#pragma once
#include "poz_simt_fast.cuh"
#include "poz_kernels.cuh"

namespace poz {
namespace mk {

constexpr int NT = 256;            // threads per CTA
constexpr int BM = 32, BN = 64;    // output tile
constexpr int BKS = 64;            // k depth of a pipeline stage
constexpr int GROUPS = NT / 32;    // every warp takes a slice of a stage's k range
constexpr int GK = BKS / GROUPS;   // 8 k-values per warp and stage
constexpr int STAGES = 4;
constexpr int UM = 8, UN = 8;      // outputs per thread
constexpr int MAX_LAYERS_MK = 8;
constexpr int MAX_PHASES = 2 * MAX_LAYERS_MK + 4;

constexpr int A_ELEMS_MAX = BKS * (BM + 4);                  // mn-contiguous arrangement is the larger one
constexpr int B_ELEMS_MAX = BKS * (BN + 4);
constexpr int STAGE_ELEMS = A_ELEMS_MAX + B_ELEMS_MAX;
constexpr int PIPE_ELEMS = STAGES * STAGE_ELEMS;
constexpr int RED_TILE = (BM + 1) * BN;                      // one partial tile (+ one row of column sums)
constexpr int RED_ELEMS = GROUPS * RED_TILE;
static_assert(RED_ELEMS <= PIPE_ELEMS, "reduction buffer must fit in the pipeline buffers");
constexpr size_t SMEM_BYTES = (size_t)PIPE_ELEMS * sizeof(float);

enum TaskKind : int { TASK_FWD = 0, TASK_BWD_X = 1, TASK_BWD_W = 2, TASK_HEAD = 3, TASK_UPDATE = 4 };

struct Task {
    int kind, layer;
    int m0, n0;            // tile origin (GEMM tasks)
    long long k0, k1;      // contraction range (backward-w: batch rows of this slab)
    int z;                 // gradient slab (backward-w); part index (head / update)
    int parts;             // cooperating tasks (head / update)
};

struct Layer {
    const float* in;       // layer input [batch][K]; x for layer 0
    const float* wb;       // [K (+1)][N] image of W and b
    float* a_out;          // recorded output of the unit
    float* dz;             // gradient of the unit's pre-dropout linear output
    int K, N, act, pad;
    DropPoint dpz, dpa;
};

struct Args {
    Layer L[MAX_LAYERS_MK];
    int nl, has_bias;
    long long batch;
    const int* row_idx;    // gather of x / label rows (nullable)
    const int* labels;
    float *P, *AM, *AV, *G;
    long long slab;
    SegTable seg;
    DevState* st;
    double* cost_ring; long long cost_cap;
    double* stats; long long stats_cap;
    int n_stat;
    double l2_share, inv_batch;
    const Task* tasks;
    int n_phases;
    int phase_end[MAX_PHASES];
    unsigned int* barrier;  // [0] arrivals of the current barrier, [1] generation
    int trace_reps;              // 1; a tracing experiment may repeat every task to time it warm
    unsigned long long* trace;   // nullable: [phase + 1][grid] %globaltimer stamps (POZ_STEP_TRACE=1), [0][..] = kernel entry
};

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---- grid-wide barrier (the launch is cooperative, so every CTA is resident) -------------------------
__device__ __forceinline__ void grid_barrier(unsigned int* bar) {
    __syncthreads();
    if (threadIdx.x == 0) {
        volatile unsigned int* gen_p = bar + 1;
        const unsigned int gen = *gen_p;
        __threadfence();
        if (atomicAdd(bar, 1u) == gridDim.x - 1) {
            bar[0] = 0u;
            __threadfence();
            atomicAdd(bar + 1, 1u);
        } else {
            while (*gen_p == gen) { }
        }
        __threadfence();
    }
    __syncthreads();
}

// ---- operand tile -> shared memory ---------------------------------------------------------------------
//   KC  (k contiguous in global): S[MN][BKS+4], global rows <-> mn, cols <-> k
//   !KC (mn contiguous in global): S[BKS][MN+4], global rows <-> k,  cols <-> mn
// Each thread owns fixed 16-byte chunks; everything that does not depend on the k-tile is computed once.
// A row gather along k (the sampled batch rows of the first layer's weight gradient) is resolved one
// stage ahead so that the index load is off the copy's issue path.
template <int MN, bool KC> struct Copier {
    static constexpr int LD = KC ? BKS + 4 : MN + 4;
    static constexpr int ROWS = KC ? MN : BKS;
    static constexpr int CV = (KC ? BKS : MN) / 4;
    static constexpr int N = (ROWS * CV) / NT;
    static_assert((ROWS * CV) % NT == 0, "chunks must divide evenly over the CTA");
    const float* src[N];
    int soff[N];
    int kofs[N];
    int next_row[N];
    bool ok[N];
    long long step, k_lim, k_first;
    const int* gather;     // !KC only: rows of the operand are k, looked up through this table
    const float* base;
    long long ld;

    __device__ __forceinline__ void init(const fast::Op& op, long long mn0, long long k_begin, long long k_end, int tid) {
        k_first = k_begin;
        base = op.ptr; ld = op.ld;
        gather = KC ? nullptr : op.row_idx;
        if (KC) { k_lim = k_end < op.cols ? k_end : (long long)op.cols; step = BKS; }
        else    { k_lim = k_end < op.rows ? k_end : op.rows; step = (long long)BKS * op.ld; }
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const int c = tid + i * NT;
            const int r = c / CV, cc = (c % CV) * 4;
            soff[i] = r * LD + cc;
            next_row[i] = 0;
            if (KC) {
                const long long gr = mn0 + r;
                ok[i] = gr < op.rows;
                kofs[i] = cc;
                src[i] = ok[i] ? op.ptr + (op.row_idx ? (long long)op.row_idx[gr] : gr) * op.ld + k_begin + cc : op.ptr;
            } else {
                const long long gc = mn0 + cc;
                ok[i] = gc < op.cols;
                kofs[i] = r;
                src[i] = ok[i] ? op.ptr + (gather ? gc : (k_begin + r) * op.ld + gc) : op.ptr;
                if (gather && ok[i] && k_begin + r < k_lim) next_row[i] = gather[k_begin + r];
            }
        }
    }
    __device__ __forceinline__ void issue(float* S, int t) {
        const long long kt = k_first + (long long)t * BKS;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const bool v = ok[i] && (kt + kofs[i] < k_lim);
            const float* s = src[i];
            if (!KC && gather) {
                if (v) s = src[i] + (long long)next_row[i] * ld;
                const long long kn = kt + BKS + kofs[i];
                if (ok[i] && kn < k_lim) next_row[i] = gather[kn];
            } else if (v) {
                s = src[i] + (long long)t * step;
            }
            fast::cp_async16(S + soff[i], s, v ? 16 : 0);
        }
    }
};

// element-wise copy for operands that do not allow 16-byte chunks
template <int MN, bool KC>
__device__ __forceinline__ void load_tile_slow(float* S, const fast::Op& op, long long mn0, long long k0, long long k_lim,
                                               int tid) {
    constexpr int LD = KC ? BKS + 4 : MN + 4;
    constexpr int ROWS = KC ? MN : BKS;
    constexpr int COLS = KC ? BKS : MN;
    const long long row0 = KC ? mn0 : k0, col0 = KC ? k0 : mn0;
    const long long row_lim = KC ? op.rows : (k_lim < op.rows ? k_lim : op.rows);
    const long long col_lim = KC ? (k_lim < op.cols ? k_lim : (long long)op.cols) : (long long)op.cols;
    for (int c = tid; c < ROWS * COLS; c += NT) {
        const int r = c / COLS, cc = c % COLS;
        const long long gr = row0 + r, gc = col0 + cc;
        const bool ok = gr < row_lim && gc < col_lim;
        const float* src = op.ptr;
        if (ok) src = op.ptr + (op.row_idx ? (long long)op.row_idx[gr] : gr) * op.ld + gc;
        fast::cp_async4(S + r * LD + cc, src, ok ? 4 : 0);
    }
}

// ---- one 32x64 output tile over a contraction range ------------------------------------------------------
// Each of the eight warps computes the WHOLE tile for its eighth of every stage's k range, with an
// 8x8 block of outputs per thread: per four k-values a thread issues 16 128-bit shared-memory reads
// for 256 FMAs, and a half-warp's read never spans more than 128 distinct bytes (one wavefront; the
// hardware serves 128-bit reads per half-warp).  The eight partial tiles are added through shared
// memory before the epilogue.  Output rows/columns of thread (ty, tx), ty = lane / 8, tx = lane % 8:
//   rows  ty + 4 i      when A is k-contiguous in shared memory (adjacent rows sit in distinct banks),
//         ty * 8 + i    when A is row-contiguous (two 128-bit reads fetch the eight rows of one k);
//   cols  tx + 8 j      when B is k-contiguous, else tx * 4 + (j & 3) + 32 (j >> 2).
// COLSUM (backward-w): tiles of the first row block also produce the column sums of B (the bias
// gradient) as output row `bias_row`.
template <bool A_KC> __device__ __forceinline__ int row_of(int ty, int i) { return A_KC ? ty + 4 * i : ty * 8 + i; }
template <bool B_KC> __device__ __forceinline__ int col_of(int tx, int j) { return B_KC ? tx + 8 * j : tx * 4 + (j & 3) + 32 * (j >> 2); }

template <bool A_KC, bool B_KC>
__device__ __forceinline__ void load_frag(const float* sa, const float* sb, int k4, int ty, int tx, float (&a)[4][UM],
                                          float (&b)[4][UN]) {
    constexpr int A_LD = A_KC ? BKS + 4 : BM + 4;
    constexpr int B_LD = B_KC ? BKS + 4 : BN + 4;
    if (A_KC) {
#pragma unroll
        for (int i = 0; i < UM; ++i) {
            const float4 v = *reinterpret_cast<const float4*>(sa + (ty + 4 * i) * A_LD + k4);
            a[0][i] = v.x; a[1][i] = v.y; a[2][i] = v.z; a[3][i] = v.w;
        }
    } else {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const float4 v0 = *reinterpret_cast<const float4*>(sa + (k4 + kk) * A_LD + ty * 8);
            const float4 v1 = *reinterpret_cast<const float4*>(sa + (k4 + kk) * A_LD + ty * 8 + 4);
            a[kk][0] = v0.x; a[kk][1] = v0.y; a[kk][2] = v0.z; a[kk][3] = v0.w;
            a[kk][4] = v1.x; a[kk][5] = v1.y; a[kk][6] = v1.z; a[kk][7] = v1.w;
        }
    }
    if (B_KC) {
#pragma unroll
        for (int j = 0; j < UN; ++j) {
            const float4 v = *reinterpret_cast<const float4*>(sb + (tx + 8 * j) * B_LD + k4);
            b[0][j] = v.x; b[1][j] = v.y; b[2][j] = v.z; b[3][j] = v.w;
        }
    } else {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            const float4 v0 = *reinterpret_cast<const float4*>(sb + (k4 + kk) * B_LD + tx * 4);
            const float4 v1 = *reinterpret_cast<const float4*>(sb + (k4 + kk) * B_LD + 32 + tx * 4);
            b[kk][0] = v0.x; b[kk][1] = v0.y; b[kk][2] = v0.z; b[kk][3] = v0.w;
            b[kk][4] = v1.x; b[kk][5] = v1.y; b[kk][6] = v1.z; b[kk][7] = v1.w;
        }
    }
}

template <bool COLSUM>
__device__ __forceinline__ void fma_frag(float (&acc)[UM][UN], float (&bsum)[UN], const float (&a)[4][UM],
                                         const float (&b)[4][UN]) {
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int i = 0; i < UM; ++i)
#pragma unroll
            for (int j = 0; j < UN; ++j) acc[i][j] = fmaf(a[kk][i], b[kk][j], acc[i][j]);
        if (COLSUM) {
#pragma unroll
            for (int j = 0; j < UN; ++j) bsum[j] += b[kk][j];
        }
    }
}

template <bool A_KC, bool B_KC, bool COLSUM, class Epi>
__device__ __forceinline__ void tile_gemm(float* smem, const fast::Op& A, const fast::Op& B, long long M, int N,
                                          long long m0, int n0, long long k_begin, long long k_end,
                                          long long bias_row, int z, const Epi& epi) {
    const int tid = threadIdx.x;
    const int grp = tid >> 5, lane = tid & 31;
    const int ty = lane >> 3, tx = lane & 7;
    const int nk = k_end > k_begin ? (int)((k_end - k_begin + BKS - 1) / BKS) : 0;

    float acc[UM][UN];
#pragma unroll
    for (int i = 0; i < UM; ++i)
#pragma unroll
        for (int j = 0; j < UN; ++j) acc[i][j] = 0.f;
    float bsum[UN];
#pragma unroll
    for (int j = 0; j < UN; ++j) bsum[j] = 0.f;

    const bool a_fast = A.vec4 != 0;
    const bool b_fast = B.vec4 && (B_KC || !B.row_idx);
    Copier<BM, A_KC> ca;
    Copier<BN, B_KC> cb;
    if (a_fast) ca.init(A, m0, k_begin, k_end, tid);
    if (b_fast) cb.init(B, n0, k_begin, k_end, tid);
    auto issue = [&](int t) {
        float* sa = smem + (t % STAGES) * STAGE_ELEMS;
        float* sb = sa + A_ELEMS_MAX;
        const long long k0 = k_begin + (long long)t * BKS;
        if (a_fast) ca.issue(sa, t); else load_tile_slow<BM, A_KC>(sa, A, m0, k0, k_end, tid);
        if (b_fast) cb.issue(sb, t); else load_tile_slow<BN, B_KC>(sb, B, n0, k0, k_end, tid);
    };
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) issue(s);
        fast::cp_commit();
    }
    for (int t = 0; t < nk; ++t) {
        fast::cp_wait<STAGES - 2>();
        __syncthreads();
        if (t + STAGES - 1 < nk) issue(t + STAGES - 1);
        fast::cp_commit();
        const float* sa = smem + (t % STAGES) * STAGE_ELEMS;
        const float* sb = sa + A_ELEMS_MAX;
        // this warp's k-values of the stage: grp * GK .. + GK, as GK / 4 blocks of four with the next
        // block's fragments in flight while the current one is multiplied
        float a0[4][UM], b0[4][UN], a1[4][UM], b1[4][UN];
        load_frag<A_KC, B_KC>(sa, sb, grp * GK, ty, tx, a0, b0);
#pragma unroll
        for (int kq = 0; kq < GK; kq += 8) {
            if (kq + 4 < GK) load_frag<A_KC, B_KC>(sa, sb, grp * GK + kq + 4, ty, tx, a1, b1);
            fma_frag<COLSUM>(acc, bsum, a0, b0);
            if (kq + 8 < GK) load_frag<A_KC, B_KC>(sa, sb, grp * GK + kq + 8, ty, tx, a0, b0);
            if (kq + 4 < GK) fma_frag<COLSUM>(acc, bsum, a1, b1);
        }
    }
    fast::cp_wait<0>();
    __syncthreads();                           // everyone is done with the pipeline buffers

    // ---- add the warps' partial tiles through shared memory, then the epilogue ----
    float* red = smem + grp * RED_TILE;
#pragma unroll
    for (int i = 0; i < UM; ++i)
#pragma unroll
        for (int j = 0; j < UN; ++j) red[row_of<A_KC>(ty, i) * BN + col_of<B_KC>(tx, j)] = acc[i][j];
    if (COLSUM && ty == 0) {
#pragma unroll
        for (int j = 0; j < UN; ++j) red[BM * BN + col_of<B_KC>(tx, j)] = bsum[j];
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < (BM * BN / 4) / NT; ++q) {
        const int idx = tid + q * NT;
        const int row = idx / (BN / 4), c4 = (idx % (BN / 4)) * 4;
        float4 s = *reinterpret_cast<const float4*>(smem + row * BN + c4);
#pragma unroll
        for (int g = 1; g < GROUPS; ++g) {
            const float4 v = *reinterpret_cast<const float4*>(smem + g * RED_TILE + row * BN + c4);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        const long long m = m0 + row;
        const int n = n0 + c4;
        if (m < M && n < N) {
            const float vals[4] = {s.x, s.y, s.z, s.w};
            epi.template row<4>(m, n, vals, (N - n < 4) ? (N - n) : 4, z);
        }
    }
    if (COLSUM && m0 == 0 && tid < BN / 4) {
        const int c4 = tid * 4;
        float4 s = *reinterpret_cast<const float4*>(smem + BM * BN + c4);
#pragma unroll
        for (int g = 1; g < GROUPS; ++g) {
            const float4 v = *reinterpret_cast<const float4*>(smem + g * RED_TILE + BM * BN + c4);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
        const int n = n0 + c4;
        if (n < N) {
            const float vals[4] = {s.x, s.y, s.z, s.w};
            epi.template row<4>(bias_row, n, vals, (N - n < 4) ? (N - n) : 4, z);
        }
    }
    __syncthreads();                           // the next task reuses the buffers
}

__device__ __forceinline__ fast::Op make_op(const float* ptr, long long ld, const int* row_idx, long long rows, int cols) {
    fast::Op o;
    o.ptr = ptr; o.ld = ld; o.row_idx = row_idx; o.rows = rows; o.cols = cols;
    o.vec4 = ((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld % 4) == 0 && (cols % 4) == 0) ? 1 : 0;
    return o;
}

// The per-step bookkeeping of step_begin_kernel, without publishing next_epoch (the other CTAs read
// it while this runs; it is advanced at the end of the launch).
__device__ __forceinline__ void step_begin_inline(const Args& a, long long e) {
    DevState* st = a.st;
    st->l2_share = a.l2_share;
    st->cur_epoch = e;
    const double lr = st->lr0 * pow(st->decay, (double)e);
    st->lr_t = lr;
    if (st->optimizer == 1) {
        const long long t = st->adam_t + 1;
        st->adam_t = t;
        const double bc1 = 1.0 - pow(st->beta1, (double)t);
        const double bc2 = 1.0 - pow(st->beta2, (double)t);
        st->step_size = lr / bc1;
        st->bc2_sqrt = sqrt(bc2);
    }
    a.cost_ring[e % a.cost_cap] = 0.0;
    long long rec = -1;
    if (st->progress_stride > 0 && (e % st->progress_stride) == 0 && st->n_records < a.stats_cap) {
        rec = st->n_records;
        st->n_records = rec + 1;
        for (int i = 0; i < a.n_stat * 4; ++i) a.stats[rec * a.n_stat * 4 + i] = 0.0;
    }
    st->cur_record = rec;
}

__global__ void __launch_bounds__(NT, 1) narrow_step_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(16) float smem[];
    __shared__ DevState sst;      // what the dropout points and the cost slot need, valid from the first phase
    if (threadIdx.x == 0) {
        const long long e = __ldcg(&a.st->next_epoch);
        sst.cur_epoch = e;
        sst.seed = __ldcg(&a.st->seed);
    }
    __syncthreads();
    if (a.trace && threadIdx.x == 0) a.trace[blockIdx.x] = global_ns();
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) step_begin_inline(a, sst.cur_epoch);

    int lo = 0;
    for (int ph = 0; ph < a.n_phases; ++ph) {
        const int hi = a.phase_end[ph];
        for (int ti = lo + (int)blockIdx.x; ti < hi; ti += (int)gridDim.x) {
            const Task t = a.tasks[ti];
            const Layer& L = a.L[t.layer];
            for (int rep = 0; rep < a.trace_reps; ++rep) {
            if (a.trace && threadIdx.x == 0) a.trace[(size_t)(MAX_PHASES + 1) * gridDim.x + 2 * ti] = global_ns();
            if (t.kind == TASK_FWD) {
                const fast::Op A = make_op(L.in, L.K, t.layer == 0 ? a.row_idx : nullptr, a.batch, L.K);
                const fast::Op B = make_op(L.wb, L.N, nullptr, L.K, L.N);
                fast::FwdBiasEpi epi;
                epi.bias = a.has_bias ? L.wb + (long long)L.K * L.N : nullptr;
                epi.inner.z_out = nullptr; epi.inner.a_out = L.a_out; epi.inner.ld = L.N; epi.inner.act = L.act;
                epi.inner.dz = L.dpz; epi.inner.da = L.dpa; epi.inner.st = &sst;
                tile_gemm<true, false, false>(smem, A, B, a.batch, L.N, t.m0, t.n0, 0, L.K, 0, 0, epi);
            } else if (t.kind == TASK_BWD_X) {
                // dZ(l-1) = (dZ(l) W(l)^T) pushed through unit l-1's dropout / activation
                const Layer& Pv = a.L[t.layer - 1];
                const fast::Op A = make_op(L.dz, L.N, nullptr, a.batch, L.N);
                const fast::Op B = make_op(L.wb, L.N, nullptr, L.K, L.N);
                BwdDataEpi<float> epi;
                epi.a_prev = Pv.a_out; epi.dz_prev = Pv.dz; epi.da_rec = nullptr; epi.dz_rec = nullptr;
                epi.ld = L.K; epi.act = Pv.act; epi.dz = Pv.dpz; epi.da = Pv.dpa; epi.st = &sst;
                tile_gemm<true, true, false>(smem, A, B, a.batch, L.K, t.m0, t.n0, 0, L.N, 0, 0, epi);
            } else if (t.kind == TASK_BWD_W) {
                const fast::Op A = make_op(L.in, L.K, t.layer == 0 ? a.row_idx : nullptr, a.batch, L.K);
                const fast::Op B = make_op(L.dz, L.N, nullptr, a.batch, L.N);
                BwdWeightEpi<float> epi;
                epi.g = a.G; epi.slab = a.slab; epi.off = a.seg.off[t.layer]; epi.ld = L.N;
                if (a.has_bias)
                    tile_gemm<false, false, true>(smem, A, B, L.K, L.N, t.m0, t.n0, t.k0, t.k1, L.K, t.z, epi);
                else
                    tile_gemm<false, false, false>(smem, A, B, L.K, L.N, t.m0, t.n0, t.k0, t.k1, L.K, t.z, epi);
            } else if (t.kind == TASK_HEAD) {
                const Layer& Pv = a.L[t.layer - 1];
                HeadArgs<float> h{};
                h.a_prev = Pv.a_out; h.wb = L.wb; h.labels = a.labels; h.row_idx = a.row_idx;
                h.dz_pre = L.dz; h.dprev_pre = Pv.dz;
                h.M = a.batch; h.K = L.K; h.N = L.N; h.has_bias = a.has_bias; h.act_prev = Pv.act;
                h.dz_last = L.dpz; h.da_prev = Pv.dpa; h.dz_prev = Pv.dpz;
                h.inv_batch = a.inv_batch;
                h.cost_base = a.cost_ring; h.cost_cap = a.cost_cap; h.st = &sst;
                head_ce_body<float>(h, reinterpret_cast<unsigned char*>(smem), t.z, t.parts);
                __syncthreads();
            } else {   // TASK_UPDATE
                update_body<float>(a.P, a.AM, a.AV, a.G, a.slab, nullptr, a.seg, a.st, a.cost_ring, a.cost_cap, a.stats,
                                   nullptr, t.layer, t.z, t.parts);
            }
            if (a.trace && threadIdx.x == 0) a.trace[(size_t)(MAX_PHASES + 1) * gridDim.x + 2 * ti + 1] = global_ns();
            }
        }
        lo = hi;
        if (a.trace) {
            __syncthreads();
            if (threadIdx.x == 0) a.trace[(size_t)(ph + 1) * gridDim.x + blockIdx.x] = global_ns();
        }
        if (ph + 1 < a.n_phases) grid_barrier(a.barrier);
    }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) a.st->next_epoch = sst.cur_epoch + 1;
}

}  // namespace mk
}  // namespace poz
