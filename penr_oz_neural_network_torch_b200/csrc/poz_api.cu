// pozmlp C ABI (include/pozmlp.h): context management and the narrow SIMT path (fp32 / fp64).
#include "../../include/pozmlp.h"
#include "poz_ctx.h"
#include "poz_kernels.cuh"
#include "poz_net_kernels.cuh"
#include "poz_simt_gemm.cuh"
#include "poz_simt_fast.cuh"
#include "poz_infer_small.cuh"
#include "poz_tc.h"
#include "poz_tc3.h"
#include "poz_tcs.h"
#include "poz_tail_kernel.cuh"
#include "poz_wide_kernels.cuh"

#include <algorithm>
#include <atomic>
#include <ctime>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <vector>

namespace poz {

static thread_local std::string t_error;
void set_error(const std::string& msg) { t_error = msg; }
static std::atomic<long long> s_launches{0};
void count_launch(int n) { s_launches.fetch_add(n, std::memory_order_relaxed); }

#define POZ_LAUNCH_CHECK()                                                                    \
    do {                                                                                      \
        poz::count_launch();                                                                  \
        POZ_CUDA(cudaGetLastError());                                                         \
    } while (0)

static int num_sms(int device) {
    static int cached[64] = {0};
    if (device >= 0 && device < 64 && cached[device]) return cached[device];
    int n = 148;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
    if (device >= 0 && device < 64) cached[device] = n;
    return n;
}

// Brackets one launch with CUDA events when the context is in profiling mode (poz_profile_begin).
struct ProfScope {
    poz_ctx* c; cudaStream_t s; size_t idx = 0; bool on;
    ProfScope(poz_ctx* ctx, const std::string& tag, cudaStream_t stream) : c(ctx), s(stream), on(ctx->profiling) {
        if (!on) return;
        poz_ctx::ProfRec r;
        r.tag = tag;
        cudaEventCreate(&r.a);
        cudaEventCreate(&r.b);
        cudaEventRecord(r.a, s);
        c->prof.push_back(r);
        idx = c->prof.size() - 1;
    }
    ~ProfScope() { if (on) cudaEventRecord(c->prof[idx].b, s); }
};

template <typename T> static Operand<T> make_operand(const T* ptr, long long ld, const int* row_idx,
                                                     long long rows, int cols, int ones) {
    Operand<T> o;
    o.ptr = ptr; o.ld = ld; o.row_idx = row_idx; o.rows = rows; o.cols = cols; o.ones = ones;
    constexpr int V = Vec16<T>::V;
    o.vec_ok = ((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld % V) == 0) ? 1 : 0;
    return o;
}

// ---- tile configurations -----------------------------------------------------------------------
template <typename T> struct Tiles;
template <> struct Tiles<float> {
    // small: fills the GPU at batch ~1k; large: better reuse for big batches; thin: out <= 16
    static constexpr int SM_BM = 32, SM_BN = 64, SM_BK = 16, SM_TM = 4, SM_TN = 4;
    static constexpr int LG_BM = 128, LG_BN = 64, LG_BK = 16, LG_TM = 8, LG_TN = 4;
    static constexpr int TH_BM = 64, TH_BN = 16, TH_BK = 16, TH_TM = 4, TH_TN = 4;
};
template <> struct Tiles<double> {
    static constexpr int SM_BM = 32, SM_BN = 32, SM_BK = 8, SM_TM = 4, SM_TN = 2;
    static constexpr int LG_BM = 64, LG_BN = 64, LG_BK = 8, LG_TM = 4, LG_TN = 4;
    static constexpr int TH_BM = 64, TH_BN = 16, TH_BK = 8, TH_TM = 4, TH_TN = 2;
};

enum TileKind { TILE_SMALL, TILE_LARGE, TILE_THIN };

template <typename T, bool A_KC, bool B_KC, class Epi>
static int launch_gemm(TileKind kind, const Operand<T>& A, const Operand<T>& B, long long M, int N,
                       long long Kd, int splits, long long k_chunk, const Epi& epi,
                       cudaStream_t stream) {
    using TL = Tiles<T>;
    if (M <= 0 || N <= 0) return 0;
    auto grid_for = [&](int BM, int BN) {
        return dim3((unsigned)((N + BN - 1) / BN), (unsigned)((M + BM - 1) / BM), (unsigned)splits);
    };
    switch (kind) {
        case TILE_LARGE: {
            dim3 g = grid_for(TL::LG_BM, TL::LG_BN);
            if (g.y > 65535) POZ_FAIL("gemm: too many row tiles for one launch");
            gemm_simt_kernel<T, TL::LG_BM, TL::LG_BN, TL::LG_BK, TL::LG_TM, TL::LG_TN, A_KC, B_KC, Epi>
                <<<g, (TL::LG_BM / TL::LG_TM) * (TL::LG_BN / TL::LG_TN), 0, stream>>>(A, B, M, N, Kd, k_chunk, epi);
            break;
        }
        case TILE_THIN: {
            dim3 g = grid_for(TL::TH_BM, TL::TH_BN);
            if (g.y > 65535) POZ_FAIL("gemm: too many row tiles for one launch");
            gemm_simt_kernel<T, TL::TH_BM, TL::TH_BN, TL::TH_BK, TL::TH_TM, TL::TH_TN, A_KC, B_KC, Epi>
                <<<g, (TL::TH_BM / TL::TH_TM) * (TL::TH_BN / TL::TH_TN), 0, stream>>>(A, B, M, N, Kd, k_chunk, epi);
            break;
        }
        default: {
            dim3 g = grid_for(TL::SM_BM, TL::SM_BN);
            if (g.y > 65535) POZ_FAIL("gemm: too many row tiles for one launch");
            gemm_simt_kernel<T, TL::SM_BM, TL::SM_BN, TL::SM_BK, TL::SM_TM, TL::SM_TN, A_KC, B_KC, Epi>
                <<<g, (TL::SM_BM / TL::SM_TM) * (TL::SM_BN / TL::SM_TN), 0, stream>>>(A, B, M, N, Kd, k_chunk, epi);
            break;
        }
    }
    POZ_LAUNCH_CHECK();
    return 0;
}

template <typename T> static TileKind pick_tile(const poz_ctx* c, long long M, int N) {
    using TL = Tiles<T>;
    if (N <= 16) return TILE_THIN;
    long long small_ctas = ((M + TL::SM_BM - 1) / TL::SM_BM) * ((N + TL::SM_BN - 1) / TL::SM_BN);
    return small_ctas >= 8LL * num_sms(c->device) ? TILE_LARGE : TILE_SMALL;
}

// ---- memory helpers ----------------------------------------------------------------------------
static int dev_alloc(void** p, size_t bytes) {
    *p = nullptr;
    if (bytes == 0) return 0;
    POZ_CUDA(cudaMalloc(p, bytes));
    return 0;
}
static void dev_free(void*& p) {
    if (p) cudaFree(p);
    p = nullptr;
}

static DropPoint make_drop(const poz_ctx* c, int layer, int which, const void* const* masks) {
    DropPoint d{};
    const bool flag = which == 0 ? c->drop_z[layer] : c->drop_a[layer];
    const double p = c->cfg.dropout_rate;
    d.enabled = (flag && c->training_pass && p > 0.0) ? 1 : 0;
    d.mask = (d.enabled && masks) ? static_cast<const uint8_t*>(masks[2 * layer + which]) : nullptr;
    d.point = (uint32_t)(2 * layer + which);
    d.keep_prob = 1.0 - p;
    d.scale = 1.0 / (1.0 - p);
    return d;
}

static int ensure_stage(poz_ctx* c, size_t elems) {
    if (c->stage_elems >= elems) return 0;
    dev_free(reinterpret_cast<void*&>(c->stage));
    POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->stage), elems * sizeof(double)));
    c->stage_elems = elems;
    return 0;
}

// host float64 image -> device compute dtype at flat offset `off`
template <typename T> static int put_f64(poz_ctx* c, void* dst_base, long long off, const double* src, long long n) {
    if (n <= 0) return 0;
    T* dst = static_cast<T*>(dst_base) + off;
    if (sizeof(T) == sizeof(double)) {
        POZ_CUDA(cudaMemcpy(dst, src, n * sizeof(double), cudaMemcpyHostToDevice));
        return 0;
    }
    POZ_TRY(ensure_stage(c, (size_t)n));
    POZ_CUDA(cudaMemcpy(c->stage, src, n * sizeof(double), cudaMemcpyHostToDevice));
    int blocks = (int)std::min<long long>((n + 255) / 256, 4096);
    convert_kernel<double, T><<<blocks, 256>>>(c->stage, dst, n);
    POZ_LAUNCH_CHECK();
    POZ_CUDA(cudaDeviceSynchronize());
    return 0;
}
template <typename T> static int get_f64(poz_ctx* c, const void* src_base, long long off, double* dst, long long n) {
    if (n <= 0) return 0;
    const T* src = static_cast<const T*>(src_base) + off;
    POZ_CUDA(cudaDeviceSynchronize());
    if (sizeof(T) == sizeof(double)) {
        POZ_CUDA(cudaMemcpy(dst, src, n * sizeof(double), cudaMemcpyDeviceToHost));
        return 0;
    }
    POZ_TRY(ensure_stage(c, (size_t)n));
    int blocks = (int)std::min<long long>((n + 255) / 256, 4096);
    convert_kernel<T, double><<<blocks, 256>>>(src, c->stage, n);
    POZ_LAUNCH_CHECK();
    POZ_CUDA(cudaMemcpy(dst, c->stage, n * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

// bf16 device tensor -> float64 host image (activations and their gradients of a bf16 context)
static __global__ void bf16_to_f64_kernel(const __nv_bfloat16* __restrict__ src, double* __restrict__ dst, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dst[i] = (double)__bfloat162float(src[i]);
}
static int get_bf16_f64(poz_ctx* c, const void* src, double* dst, long long n) {
    if (n <= 0) return 0;
    POZ_CUDA(cudaDeviceSynchronize());
    POZ_TRY(ensure_stage(c, (size_t)n));
    int blocks = (int)std::min<long long>((n + 255) / 256, 4096);
    bf16_to_f64_kernel<<<blocks, 256>>>(static_cast<const __nv_bfloat16*>(src), c->stage, n);
    POZ_LAUNCH_CHECK();
    POZ_CUDA(cudaMemcpy(dst, c->stage, n * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}
// activations / activation gradients of a bf16 context are bf16; parameters and their gradients float32
static bool diag_is_bf16(const poz_ctx* c, int source) { return c->dtype == POZ_DTYPE_BF16 && source <= 1; }

#define POZ_DISPATCH(c, fn, ...)                                                              \
    ((c)->dtype == POZ_DTYPE_F64 ? fn<double>(__VA_ARGS__) : fn<float>(__VA_ARGS__))

static void clear_graphs(poz_ctx* c) {
    for (auto& e : c->graphs)
        if (e.exec) cudaGraphExecDestroy(e.exec);
    c->graphs.clear();
    for (auto& e : c->fwd_graphs)
        if (e.exec) cudaGraphExecDestroy(e.exec);
    c->fwd_graphs.clear();
}

// ---- workspace ---------------------------------------------------------------------------------
static void free_workspace(poz_ctx* c) {
    for (auto& p : c->zbuf) dev_free(p);
    for (auto& p : c->abuf) dev_free(p);
    for (auto& p : c->dzbuf) dev_free(p);
    for (auto& p : c->dz_rec) dev_free(p);
    for (auto& p : c->da_rec) dev_free(p);
    dev_free(c->ping);
    dev_free(c->pong);
    c->ws_train = 0;
    c->ws_infer = 0;
    c->ws_diag = 0;
}

static int reserve_train(poz_ctx* c, long long batch, bool diag) {
    const int nl = c->nl;
    if (batch > c->ws_train) {
        for (int l = 0; l < nl; ++l) {
            dev_free(c->zbuf[l]); dev_free(c->abuf[l]); dev_free(c->dzbuf[l]);
            dev_free(c->dz_rec[l]); dev_free(c->da_rec[l]);
        }
        c->ws_diag = 0;
        for (int l = 0; l < nl; ++l) {
            size_t bytes = (size_t)batch * c->sizes[l + 1] * c->act_esz;
            POZ_TRY(dev_alloc(&c->abuf[l], bytes));
            POZ_TRY(dev_alloc(&c->dzbuf[l], bytes));
            if (c->acts[l] != ACT_NONE) POZ_TRY(dev_alloc(&c->zbuf[l], bytes));
        }
        c->ws_train = batch;
        clear_graphs(c);
    }
    if (diag && c->ws_diag < c->ws_train) {
        for (int l = 0; l < nl; ++l) {
            size_t bytes = (size_t)c->ws_train * c->sizes[l + 1] * c->act_esz;
            dev_free(c->dz_rec[l]); dev_free(c->da_rec[l]);
            POZ_TRY(dev_alloc(&c->dz_rec[l], bytes));
            if (c->acts[l] != ACT_NONE) POZ_TRY(dev_alloc(&c->da_rec[l], bytes));
        }
        c->ws_diag = c->ws_train;
        clear_graphs(c);
    }
    return 0;
}

static int reserve_infer(poz_ctx* c, long long chunk) {
    if (chunk <= c->ws_infer) return 0;
    dev_free(c->ping);
    dev_free(c->pong);
    int widest = 0;
    for (int l = 1; l <= c->nl; ++l) widest = std::max(widest, c->sizes[l]);
    size_t bytes = (size_t)chunk * widest * c->act_esz;
    POZ_TRY(dev_alloc(&c->ping, bytes));
    POZ_TRY(dev_alloc(&c->pong, bytes));
    c->ws_infer = chunk;
    clear_graphs(c);
    return 0;
}

static int ensure_grad_slabs(poz_ctx* c, int slabs) {
    if (slabs <= c->g_slabs) return 0;
    dev_free(c->G);
    POZ_TRY(dev_alloc(&c->G, (size_t)slabs * c->p_total * c->esz));
    c->g_slabs = slabs;
    clear_graphs(c);
    return 0;
}


// ---- launch helper ---------------------------------------------------------------------------------
// (Programmatic dependent launches: in round 1, on the ~12 small launches of the step's CUDA graph, they cost 108-114 us per
//  C2 step against 82 us with plain edges, and were removed.  Round 2 brought them back for the two edges that remain on
//  the critical path -- forward -> tail -> grouped weight gradient: 51.3 -> 49.1 us per step; see poz_tcs_kernel.cuh.)
// Set around the launches of a small float32 inference pass (forward_impl): its kernels -- the pipelined SIMT GEMMs and the
// softmax kernels, which all start with index arithmetic and griddepcontrol.wait -- are then launched as programmatic
// dependents of one another, so a launch's CTAs are in place when the one before finishes.
static thread_local bool t_pdl_launches = false;
struct PdlScope {
    bool prev;
    explicit PdlScope(bool on) : prev(t_pdl_launches) { t_pdl_launches = on; }
    ~PdlScope() { t_pdl_launches = prev; }
};

template <typename... P, typename... A>
static int launch_pdl(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    if (t_pdl_launches) {
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
    }
    POZ_CUDA(cudaLaunchKernelEx(&cfg, kern, std::forward<A>(args)...));
    count_launch();
    return 0;
}

// ---- pipelined fp32 kernels (poz_simt_fast.cuh): launch helpers -------------------------------------
static fast::Op make_fast_op(const float* ptr, long long ld, const int* row_idx, long long rows, int cols) {
    fast::Op o;
    o.ptr = ptr; o.ld = ld; o.row_idx = row_idx; o.rows = rows; o.cols = cols;
    o.vec4 = ((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (ld % 4) == 0 && (cols % 4) == 0) ? 1 : 0;
    return o;
}

template <int BM, int BN, bool A_KC, bool B_KC, bool COLSUM, class Epi>
static int launch_fast(const fast::Op& A, const fast::Op& B, long long M, int N, long long Kd, int splits,
                       bool cluster, long long bias_row, const Epi& epi, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return 0;
    long long k_chunk = (Kd + splits - 1) / splits;
    k_chunk = ((k_chunk + fast::BK - 1) / fast::BK) * fast::BK;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)((N + BN - 1) / BN), (unsigned)((M + BM - 1) / BM), (unsigned)splits);
    if (cfg.gridDim.y > 65535) POZ_FAIL("gemm: too many row tiles for one launch");
    cfg.blockDim = dim3((BM / fast::TM) * (BN / fast::TN));
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    unsigned na = 0;
    if (cluster && splits > 1) {
        attr[na].id = cudaLaunchAttributeClusterDimension;
        attr[na].val.clusterDim.x = 1; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = (unsigned)splits;
        ++na;
    }
    if (t_pdl_launches) {
        attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[na].val.programmaticStreamSerializationAllowed = 1;
        ++na;
    }
    if (na) { cfg.attrs = attr; cfg.numAttrs = na; }
        POZ_CUDA(cudaLaunchKernelEx(&cfg, fast::gemm_f32_pipelined<BM, BN, A_KC, B_KC, COLSUM, Epi>, A, B, M, N, Kd, k_chunk,
                                bias_row, epi));
    count_launch();
    return 0;
}

// cluster split-K factor for the forward / backward-x GEMMs: enough CTAs for ~4 per SM, each with
// at least two k-tiles; a power of two that divides the 32-row tile.
static int pick_cluster_split(const poz_ctx* c, long long tiles, long long Kd) {
    const long long sms = num_sms(c->device);
    const long long nk = (Kd + fast::BK - 1) / fast::BK;
    static const long long per_sm = getenv("POZ_FWD_CTAS_PER_SM") ? std::max(1, atoi(getenv("POZ_FWD_CTAS_PER_SM"))) : 4;
    int s = 1;
    while (s < 8 && tiles * s * 2 <= per_sm * sms && nk / (s * 2) >= 2) s *= 2;
    return s;
}

static bool fast_large(const poz_ctx* c, long long M, int N) {
    return ((M + 31) / 32) * ((N + 63) / 64) >= 8LL * num_sms(c->device);
}

// ---- forward of one layer ------------------------------------------------------------------------
template <typename T>
static int layer_forward(poz_ctx* c, int l, const T* x, const int* row_idx, long long x_rows,
                         long long M, T* z_out, T* a_out, const void* const* masks,
                         cudaStream_t stream) {
    const int K = c->sizes[l], N = c->sizes[l + 1];
    const int hb = c->has_bias;
    (void)x_rows;
    if constexpr (std::is_same<T, float>::value) {
        const float* w = static_cast<const float*>(c->P) + c->off[l];
        fast::Op A = make_fast_op(x, K, row_idx, M, K);
        fast::Op B = make_fast_op(w, N, nullptr, K, N);
        fast::FwdBiasEpi epi;
        epi.bias = hb ? w + (long long)K * N : nullptr;
        epi.inner.z_out = z_out; epi.inner.a_out = a_out; epi.inner.ld = N; epi.inner.act = c->acts[l];
        epi.inner.dz = make_drop(c, l, 0, masks);
        epi.inner.da = make_drop(c, l, 1, masks);
        epi.inner.st = c->st;
        ProfScope ps(c, "fwd_l" + std::to_string(l), stream);
        if (fast_large(c, M, N))
            return launch_fast<64, 64, true, false, false>(A, B, M, N, K, 1, false, 0, epi, stream);
        const long long tiles = ((M + 31) / 32) * ((N + 63) / 64);
        const int split = pick_cluster_split(c, tiles, K);
        return launch_fast<32, 64, true, false, false>(A, B, M, N, K, split, true, 0, epi, stream);
    }
    Operand<T> A = make_operand<T>(x, K, row_idx, M, K, hb);
    const T* w = static_cast<const T*>(c->P) + c->off[l];
    Operand<T> B = make_operand<T>(w, N, nullptr, K + hb, N, 0);
    FwdEpi<T> epi;
    epi.z_out = z_out; epi.a_out = a_out; epi.ld = N; epi.act = c->acts[l];
    epi.dz = make_drop(c, l, 0, masks);
    epi.da = make_drop(c, l, 1, masks);
    epi.st = c->st;
    ProfScope ps(c, "fwd_l" + std::to_string(l), stream);
    return launch_gemm<T, true, false>(pick_tile<T>(c, M, N), A, B, M, N, (long long)K + hb, 1,
                                       (long long)K + hb, epi, stream);
}

template <typename T> static int run_softmax(poz_ctx* c, const T* z, T* out, long long M, int N, cudaStream_t s) {
    ProfScope ps(c, "softmax", s);
    if (N <= 32) {
        long long b = (M + 255) / 256;
        return launch_pdl(softmax_small_kernel<T>, dim3((unsigned)std::max<long long>(b, 1)), dim3(256), 0, s, z, out, M, N);
    }
    int blocks = (int)std::min<long long>((M + 7) / 8, 148 * 8);
    return launch_pdl(softmax_kernel<T>, dim3(std::max(blocks, 1)), dim3(256), 0, s, z, out, M, N);
}

template <typename T>
static int run_loss(poz_ctx* c, LossArgs<T>& a, cudaStream_t s) {
    ProfScope ps(c, "loss", s);
    if (c->loss == POZ_LOSS_CE) {
        int blocks = (int)std::min<long long>((a.M + 7) / 8, 148 * 8);
        return launch_pdl(ce_loss_kernel<T>, dim3(std::max(blocks, 1)), dim3(256), 0, s, a);
    } else {
        long long total = a.M * a.N;
        int blocks = (int)std::min<long long>((total + 255) / 256, 148 * 8);
        return launch_pdl(mse_loss_kernel<T>, dim3(std::max(blocks, 1)), dim3(256), 0, s, a);
    }
}

#include "poz_api_infer.inc"

#include "poz_api_train.inc"

#include "poz_api_wide.inc"

#include "poz_general.inc"

template <typename T>
static int forward_general_dispatch(poz_ctx* c, const void* x, const void* target, long long batch, T* out,
                                    double* cost_host, cudaStream_t stream) {
    return forward_general<T>(c, x, target, batch, out, cost_host, stream);
}

template <typename T> static int apply_impl(poz_ctx* c, bool reduced, cudaStream_t stream) {
    if (!c->grads_pending) POZ_FAIL("poz_train_apply: no gradients pending");
    SegTable seg;
    fill_seg(c, seg);
    if (!reduced)
        for (int l = 0; l < c->nl; ++l) seg.splits[l] = c->seg_splits[l];
    // layers whose update already ran in the epilogue of their weight-gradient launch
    const unsigned done = c->upd_done_mask;
    c->upd_done_mask = 0;
    const bool l0_done = (done & 1u) != 0;
    bool anything_left = false;
    for (int k = 0; k < seg.n; ++k) {
        if (k < c->nl && ((done >> k) & 1u)) seg.n_elems[k] = 0;
        anything_left |= seg.n_elems[k] > 0;
    }
    const long long biggest = biggest_seg(seg);
    static const long long blocks_per_sm = getenv("POZ_UPDATE_BLOCKS_PER_SM") ? std::max(1, atoi(getenv("POZ_UPDATE_BLOCKS_PER_SM"))) : 2;
    int bx = (int)std::min<long long>((biggest + 255) / 256, blocks_per_sm * num_sms(c->device));
    T* gdiag = nullptr;
    if (c->diag_pending) {
        if (!c->Gdiag) POZ_TRY(dev_alloc(&c->Gdiag, (size_t)c->p_total * c->esz));
        gdiag = static_cast<T*>(c->Gdiag);
        c->have_diag = true;
    }
    if (anything_left) {
        ProfScope ps(c, "update", stream);
        POZ_TRY(launch_pdl(update_kernel<T>, dim3(std::max(bx, 1), seg.n), dim3(256), 0, stream,
                           static_cast<T*>(c->P), static_cast<T*>(c->AM), static_cast<T*>(c->AV),
                           static_cast<const T*>(c->G), c->p_total, gdiag, seg, static_cast<const DevState*>(c->st), c->cost_ring,
                           c->cost_cap, c->stats, static_cast<__nv_bfloat16*>(c->Pb)));
    }
    c->grads_pending = false;
    c->diag_pending = false;
    c->wp_stale = true; c->w16_stale = true;
    if constexpr (std::is_same<T, float>::value) {
        if (c->tcs_live && !l0_done) POZ_TRY(tcs_split_w0(c, stream));   // a plain update of layer 0: its planes follow
    }
    return 0;
}

template <typename T> static int finalize_impl(poz_ctx* c, cudaStream_t stream) {
    SegTable seg;
    fill_seg(c, seg);
    bool any = false;
    for (int l = 0; l < c->nl; ++l) { seg.splits[l] = c->seg_splits[l]; any |= seg.splits[l] > 1; }
    if (!any) return 0;
    const long long biggest = biggest_seg(seg);
    int bx = (int)std::min<long long>((biggest + 255) / 256, 2LL * num_sms(c->device));
    ProfScope ps(c, "grad_finalize", stream);
    grad_finalize_kernel<T><<<dim3(std::max(bx, 1), seg.n), 256, 0, stream>>>(static_cast<T*>(c->G), c->p_total, seg);
    POZ_LAUNCH_CHECK();
    return 0;
}

static int check_layer(poz_ctx* c, int layer, const char* who) {
    if (!c) POZ_FAIL(std::string(who) + ": NULL context");
    if (layer < 0 || layer >= c->nl) POZ_FAIL(std::string(who) + ": layer out of range");
    POZ_CUDA(cudaSetDevice(c->device));
    return 0;
}

template <typename T>
static int put_layer(poz_ctx* c, void* base, int layer, const double* w, const double* b) {
    long long nw = (long long)c->kin[layer] * c->sizes[layer + 1];
    if (w) POZ_TRY(put_f64<T>(c, base, c->off[layer], w, nw));
    if (b && c->has_bias) POZ_TRY(put_f64<T>(c, base, c->off[layer] + nw, b, c->sizes[layer + 1]));
    return 0;
}
template <typename T>
static int get_layer(poz_ctx* c, const void* base, int layer, double* w, double* b) {
    long long nw = (long long)c->kin[layer] * c->sizes[layer + 1];
    if (w) POZ_TRY(get_f64<T>(c, base, c->off[layer], w, nw));
    if (b && c->has_bias) POZ_TRY(get_f64<T>(c, base, c->off[layer] + nw, b, c->sizes[layer + 1]));
    return 0;
}

}  // namespace poz

using namespace poz;

// =================================================================================================
extern "C" {

const char* poz_last_error(void) { return t_error.c_str(); }
int poz_version(void) { return 100; }
long long poz_launch_count(void) { return s_launches.load(); }

int poz_ctx_create_ex(int device, const poz_stack_desc* d, poz_ctx** out) {
    if (!out) POZ_FAIL("poz_ctx_create: out is NULL");
    *out = nullptr;
    if (!d || !d->layers) POZ_FAIL("poz_ctx_create: NULL stack description");
    const int n_linear = d->n_linear, dtype_code = d->dtype_code, loss_code = d->loss_code;
    if (n_linear < 1 || n_linear > MAX_LAYERS) POZ_FAIL("poz_ctx_create: n_linear must be in [1,16]");
    if (dtype_code != POZ_DTYPE_F32 && dtype_code != POZ_DTYPE_F64 && dtype_code != POZ_DTYPE_BF16)
        POZ_FAIL("poz_ctx_create: unknown dtype code");
    const bool has_emb = d->emb_vocab > 0 || d->emb_dim > 0;
    bool general = has_emb;
    for (int l = 0; l < n_linear; ++l) {
        const poz_layer_desc& L = d->layers[l];
        if (L.in < 1 || L.out < 1) POZ_FAIL("poz_ctx_create: layer sizes must be positive");
        if (l > 0 && L.in * (long long)L.row_mult != d->layers[l - 1].out * (long long)d->layers[l - 1].row_mult)
            POZ_FAIL("poz_ctx_create: layer input does not match the (flattened) output of the previous layer");
        if (L.act < 0 || L.act > ACT_SOFTMAX) POZ_FAIL("poz_ctx_create: bad activation code");
        if (L.act == ACT_SOFTMAX && l != n_linear - 1) POZ_FAIL("poz_ctx_create: softmax is supported as the last activation only");
        if (L.row_mult < 1) POZ_FAIL("poz_ctx_create: row_mult must be >= 1");
        if (L.has_bn && l == n_linear - 1) POZ_FAIL("poz_ctx_create: batchnorm on the last layer is not supported");
        if (L.has_bn || L.row_mult != 1) general = true;
    }
    if (d->layers[n_linear - 1].row_mult != 1) POZ_FAIL("poz_ctx_create: the last layer must produce one row per sample");
    if (has_emb) {
        if (d->emb_vocab < 1 || d->emb_dim < 1 || d->tokens < 1) POZ_FAIL("poz_ctx_create: bad embedding description");
        if ((long long)d->emb_dim * d->tokens != (long long)d->layers[0].in * d->layers[0].row_mult)
            POZ_FAIL("poz_ctx_create: first layer input does not match the flattened embedding output");
    }
    if (general && dtype_code == POZ_DTYPE_BF16)
        POZ_FAIL("poz_ctx_create: embedding / flatten / batchnorm stacks need a float32 or float64 context");
    if (dtype_code == POZ_DTYPE_BF16) {
        // the tcgen05 path: dense layers only (TMA needs 16-byte rows), MSE stacks
        for (int l = 0; l < n_linear; ++l)
            if (d->layers[l].in % 8 || d->layers[l].out % 8)
                POZ_FAIL("poz_ctx_create: POZ_DTYPE_BF16 needs every layer size to be a multiple of 8");
        if (loss_code != POZ_LOSS_MSE) POZ_FAIL("poz_ctx_create: POZ_DTYPE_BF16 supports MSE stacks (no trailing softmax)");
    }
    if ((loss_code == POZ_LOSS_CE) != (d->layers[n_linear - 1].act == ACT_SOFTMAX))
        POZ_FAIL("poz_ctx_create: cross-entropy loss goes with a trailing softmax (neural_net_model.py:400)");
    int ndev = 0;
    POZ_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) POZ_FAIL("poz_ctx_create: no such CUDA device");
    POZ_CUDA(cudaSetDevice(device));

    poz_ctx* c = new poz_ctx();
    c->device = device; c->nl = n_linear; c->has_bias = d->has_bias ? 1 : 0;
    c->loss = loss_code; c->dtype = dtype_code;
    c->esz = dtype_code == POZ_DTYPE_F64 ? 8 : 4;
    c->act_esz = dtype_code == POZ_DTYPE_BF16 ? 2 : c->esz;
    c->general = general;
    c->sizes.assign(n_linear + 1, 0);
    c->acts.assign(n_linear, 0);
    c->drop_z.assign(n_linear, 0); c->drop_a.assign(n_linear, 0);
    c->rmul.assign(n_linear, 1); c->has_bn.assign(n_linear, 0);
    c->kz.assign(n_linear, 0); c->kbn.assign(n_linear, 0); c->ka.assign(n_linear, 0);
    c->bn_eps.assign(n_linear, 1e-5); c->bn_mom.assign(n_linear, 0.1);
    c->off_bn.assign(n_linear, -1);
    c->sizes[0] = d->layers[0].in;
    c->kin.assign(n_linear, 0);
    for (int l = 0; l < n_linear; ++l) {
        const poz_layer_desc& L = d->layers[l];
        c->sizes[l + 1] = L.out;
        c->kin[l] = L.in;
        c->acts[l] = L.act;
        c->kz[l] = L.drop_z; c->kbn[l] = L.drop_bn; c->ka[l] = L.drop_a;
        c->drop_z[l] = L.drop_z > 0; c->drop_a[l] = L.drop_a > 0;
        c->rmul[l] = L.row_mult; c->has_bn[l] = L.has_bn ? 1 : 0;
        c->bn_eps[l] = L.bn_eps; c->bn_mom[l] = L.bn_momentum;
    }
    c->emb_vocab = has_emb ? d->emb_vocab : 0; c->emb_dim = has_emb ? d->emb_dim : 0;
    c->emb_drop = d->emb_drop; c->tokens = has_emb ? d->tokens : 1;
    // flat parameter layout in the reference's `params` order: embedding table, then per Linear
    // layer [W; b] followed by its batchnorm [gain; bias]; every image 16-byte aligned
    auto pad4 = [](long long n) { return (n + 3) & ~3LL; };
    long long cursor = 0;
    if (has_emb) { c->off_emb = 0; cursor = pad4((long long)c->emb_vocab * c->emb_dim); }
    c->off.assign(n_linear + 1, 0);
    for (int l = 0; l < n_linear; ++l) {
        c->off[l] = cursor;
        cursor += pad4(((long long)c->kin[l] + c->has_bias) * c->sizes[l + 1]);
        if (c->has_bn[l]) { c->off_bn[l] = cursor; cursor += pad4(2LL * c->sizes[l + 1]); }
    }
    c->off[n_linear] = cursor;
    c->p_total = cursor;
    c->zbuf.assign(n_linear, nullptr); c->abuf.assign(n_linear, nullptr);
    c->dzbuf.assign(n_linear, nullptr); c->dz_rec.assign(n_linear, nullptr);
    c->da_rec.assign(n_linear, nullptr);
    c->bn_run.assign(n_linear, nullptr); c->bn_stat.assign(n_linear, nullptr); c->bn_sums.assign(n_linear, nullptr);
    c->ybuf.assign(n_linear, nullptr); c->dybuf.assign(n_linear, nullptr); c->dy_rec.assign(n_linear, nullptr);
    // batch splits of a weight-gradient GEMM (one gradient slab each): up to 64 for layers whose output is only a
    // few tiles, bounded so that the slabs (splits x all parameters) stay under 1 GiB; POZ_MAX_SPLITS overrides
    c->max_splits = (int)std::max<long long>(16, std::min<long long>(64, (1LL << 30) / std::max<long long>(1, c->p_total * (long long)c->esz)));
    if (const char* e = getenv("POZ_MAX_SPLITS")) c->max_splits = std::max(1, atoi(e));
    if (const char* e = getenv("POZ_TC3_MIN_BATCH")) c->tc3_min_batch = atoll(e);
    if (const char* e = getenv("POZ_GRAPHS")) c->graphs_enabled = atoi(e) != 0;
    if (const char* e = getenv("POZ_FUSE_HEAD")) c->fuse_head = atoi(e) != 0;
    if (getenv("POZ_KERNEL_TRACE") && atoi(getenv("POZ_KERNEL_TRACE")) != 0) {
        fast::KernelTraceRec* buf = nullptr;
        POZ_CUDA(cudaMemcpyFromSymbol(&buf, fast::d_ktrace, sizeof(buf)));
        if (!buf) {
            POZ_CUDA(cudaMalloc(&buf, sizeof(fast::KernelTraceRec) * (size_t)fast::KTRACE_CAP));
            POZ_CUDA(cudaMemcpyToSymbol(fast::d_ktrace, &buf, sizeof(buf)));
        }
        POZ_TRY(tcs::trace_enable());
    }
    if (const char* e = getenv("POZ_TAIL")) c->tail_fused = atoi(e) != 0;
    if (const char* e = getenv("POZ_TAIL_GROUP")) c->tail_group = atoi(e) != 0;
    if (const char* e = getenv("POZ_TAIL_CLUSTER")) c->tail_cluster = std::max(1, std::min(4, atoi(e)));
    if (const char* e = getenv("POZ_PDL")) c->pdl = atoi(e) != 0;
    if (const char* e = getenv("POZ_TC3_F16")) c->tc3_f16 = atoi(e) != 0;
    if (const char* e = getenv("POZ_TC3_F16_MIN_ROWS")) c->tc3_f16_min_rows = std::max(1LL, atoll(e));
    if (const char* e = getenv("POZ_TCS")) c->tcs_enabled = atoi(e) != 0;
    if (const char* e = getenv("POZ_TCS_FUSE_UPDATE")) c->tcs_fuse_update = atoi(e) != 0;
    if (const char* e = getenv("POZ_TCS_GROUP")) c->tcs_group = atoi(e) != 0;
    if (const char* e = getenv("POZ_TCS_MIN_BATCH")) c->tcs_min_batch = atoll(e);
    if (const char* e = getenv("POZ_TCS_MAX_BATCH")) c->tcs_max_batch = atoll(e);
    if (const char* e = getenv("POZ_TC3_TRAIN")) c->tc3_train = atoi(e) != 0;
    if (const char* e = getenv("POZ_TC3_TRAIN_MIN_BATCH")) c->tc3_train_min_batch = atoll(e);
    auto fail = [&](void) { poz_ctx_destroy(c); return 1; };
    size_t pbytes = (size_t)c->p_total * c->esz;
    if (dev_alloc(&c->P, pbytes) || dev_alloc(&c->AM, pbytes) || dev_alloc(&c->AV, pbytes)) return fail();
    if (dtype_code == POZ_DTYPE_BF16) {
        if (dev_alloc(&c->Pb, (size_t)c->p_total * 2)) return fail();
        if (cudaMemset(c->Pb, 0, (size_t)c->p_total * 2) != cudaSuccess) { set_error("cudaMemset failed"); return fail(); }
    }
    if (cudaMemset(c->P, 0, pbytes) != cudaSuccess || cudaMemset(c->AM, 0, pbytes) != cudaSuccess ||
        cudaMemset(c->AV, 0, pbytes) != cudaSuccess) { set_error("cudaMemset failed"); return fail(); }
    {
        int widest_bn = 0;
        for (int l = 0; l < n_linear; ++l) if (c->has_bn[l]) widest_bn = std::max(widest_bn, c->sizes[l + 1]);
        if (widest_bn && dev_alloc(reinterpret_cast<void**>(&c->bn_part), (size_t)BN_ROW_SPLITS * 2 * widest_bn * sizeof(double)))
            return fail();
    }
    for (int l = 0; l < n_linear; ++l)
        if (c->has_bn[l]) {
            const size_t nb = (size_t)2 * c->sizes[l + 1] * c->esz;
            if (dev_alloc(&c->bn_run[l], nb) || dev_alloc(&c->bn_stat[l], nb) || dev_alloc(&c->bn_sums[l], nb)) return fail();
            if (cudaMemset(c->bn_run[l], 0, nb) != cudaSuccess) { set_error("cudaMemset failed"); return fail(); }
        }
    if (dev_alloc(reinterpret_cast<void**>(&c->st), sizeof(DevState))) return fail();
    if (cudaMemset(c->st, 0, sizeof(DevState)) != cudaSuccess) { set_error("cudaMemset failed"); return fail(); }
    if (dev_alloc(reinterpret_cast<void**>(&c->scalar_cost), sizeof(double))) return fail();
    std::memset(&c->cfg, 0, sizeof(c->cfg));
    for (int l = 0; l < n_linear; ++l)
        if (c->has_bn[l]) {     // BatchNormLayer starts with gain 1, variance 1 (neural_net_model.py:128-131)
            std::vector<double> ones(c->sizes[l + 1], 1.0);
            if (poz_tensor_io(c, 3, l, 0, 0, ones.data()) || poz_tensor_io(c, 6, l, 0, 0, ones.data())) return fail();
        }
    c->grad_ready.assign(n_linear, nullptr);
    c->dz_ready.assign(n_linear, nullptr);
    for (int l = 0; l < n_linear; ++l)
        if (cudaEventCreateWithFlags(&c->grad_ready[l], cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&c->dz_ready[l], cudaEventDisableTiming) != cudaSuccess) {
            set_error("cudaEventCreate failed");
            return fail();
        }
    if (cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&c->side_stream2, cudaStreamNonBlocking) != cudaSuccess) {
        set_error("cudaStreamCreate failed");
        return fail();
    }
    if (cudaEventCreateWithFlags(&c->sb_fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&c->sb_done, cudaEventDisableTiming) != cudaSuccess) {
        set_error("cudaEventCreate failed");
        return fail();
    }
    if (const char* e = getenv("POZ_FORK_DW")) c->fork_dw = atoi(e) != 0;
    *out = c;
    return 0;
}

int poz_ctx_create(int device, int n_linear, const int* sizes, const int* act_codes,
                   const unsigned char* drop_z, const unsigned char* drop_a, int has_bias,
                   int loss_code, int dtype_code, poz_ctx** out) {
    if (!out) POZ_FAIL("poz_ctx_create: out is NULL");
    *out = nullptr;
    if (n_linear < 1 || n_linear > MAX_LAYERS) POZ_FAIL("poz_ctx_create: n_linear must be in [1,16]");
    if (!sizes || !act_codes) POZ_FAIL("poz_ctx_create: NULL argument");
    std::vector<poz_layer_desc> L(n_linear);
    for (int l = 0; l < n_linear; ++l) {
        L[l] = poz_layer_desc{};
        L[l].in = sizes[l]; L[l].out = sizes[l + 1]; L[l].act = act_codes[l];
        L[l].drop_z = drop_z ? (drop_z[l] ? 1 : 0) : 0;
        L[l].drop_a = drop_a ? (drop_a[l] ? 1 : 0) : 0;
        L[l].row_mult = 1; L[l].bn_eps = 1e-5; L[l].bn_momentum = 0.1;
    }
    poz_stack_desc d{};
    d.n_linear = n_linear; d.layers = L.data(); d.has_bias = has_bias; d.loss_code = loss_code; d.dtype_code = dtype_code;
    d.tokens = 1;
    return poz_ctx_create_ex(device, &d, out);
}

// POZ_KERNEL_TRACE=1: print, for the last launch of every pipelined-GEMM shape, when its CTAs reached the
// instrumented points (ns after the first CTA of that launch started).
static void dump_kernel_trace() {
    fast::KernelTraceRec* buf = nullptr;
    if (cudaMemcpyFromSymbol(&buf, fast::d_ktrace, sizeof(buf)) != cudaSuccess || !buf) return;
    cudaDeviceSynchronize();
    unsigned int n = 0;
    cudaMemcpyFromSymbol(&n, fast::d_ktrace_n, sizeof(n));
    n = std::min(n, fast::KTRACE_CAP);
    if (!n) return;
    std::vector<fast::KernelTraceRec> h(n);
    cudaMemcpy(h.data(), buf, sizeof(fast::KernelTraceRec) * n, cudaMemcpyDeviceToHost);
    // the last ~400 us of records, in time order, grouped into launches: records of one kernel shape whose
    // CTAs started within 40 us of the group's first one
    unsigned long long t_end = 0;
    for (auto& r : h) t_end = std::max(t_end, r.t[5]);
    std::vector<const fast::KernelTraceRec*> recent;
    for (auto& r : h)
        if (t_end - r.t[0] < 400000ull) recent.push_back(&r);
    std::sort(recent.begin(), recent.end(), [](const fast::KernelTraceRec* a, const fast::KernelTraceRec* b) { return a->t[0] < b->t[0]; });
    struct Group { int tag, m, n, k; unsigned long long first, last_exit; double sum[6]; unsigned long long mx[6]; size_t cnt; };
    std::vector<Group> groups;
    for (const fast::KernelTraceRec* r : recent) {
        Group* g = nullptr;
        for (size_t i = groups.size(); i-- > 0;)
            if (groups[i].tag == r->tag && groups[i].m == r->m && groups[i].n == r->n && groups[i].k == r->k) {
                if (r->t[0] - groups[i].first < 40000ull) g = &groups[i];
                break;
            }
        if (!g) {
            Group ng{};
            ng.tag = r->tag; ng.m = r->m; ng.n = r->n; ng.k = r->k; ng.first = r->t[0];
            groups.push_back(ng);
            g = &groups.back();
        }
        ++g->cnt;
        g->last_exit = std::max(g->last_exit, r->t[5]);
        for (int q = 0; q < 6; ++q) {
            g->sum[q] += (double)(r->t[q] - g->first);
            g->mx[q] = std::max(g->mx[q], r->t[q] - g->first);
        }
    }
    for (const Group& g : groups) {
        fprintf(stderr, "[poz kernel trace] launch below: first CTA entry at %lld ns, last CTA exit at %lld ns (relative to the latest launch's exit); "
                        "absolute %%globaltimer mod 1e9: %llu .. %llu\n",
                (long long)g.first - (long long)t_end, (long long)g.last_exit - (long long)t_end, g.first % 1000000000ull, g.last_exit % 1000000000ull);
        fprintf(stderr, "[poz kernel trace] tag %d M %d N %d K %d: %zu CTAs; mean (max) ns after first CTA start: entry %.0f (%llu) "
                        "setup+prologue %.0f (%llu) first tile ready %.0f (%llu) main loop done %.0f (%llu) epilogue done %.0f (%llu) exit %.0f (%llu)\n",
                g.tag, g.m, g.n, g.k, g.cnt, g.sum[0] / g.cnt, g.mx[0], g.sum[1] / g.cnt, g.mx[1], g.sum[2] / g.cnt, g.mx[2],
                g.sum[3] / g.cnt, g.mx[3], g.sum[4] / g.cnt, g.mx[4], g.sum[5] / g.cnt, g.mx[5]);
    }
    unsigned int zero = 0;
    cudaMemcpyToSymbol(fast::d_ktrace_n, &zero, sizeof(zero));
}

static void host_trace_dump();

int poz_ctx_destroy(poz_ctx* c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    clear_graphs(c);
    for (int i = 0; i < poz_ctx::HostFeed::SLOTS; ++i) {
        dev_free(c->feed.x[i]); dev_free(c->feed.t[i]);
        if (c->feed.copied[i]) cudaEventDestroy(c->feed.copied[i]);
        if (c->feed.copied_t[i]) cudaEventDestroy(c->feed.copied_t[i]);
        if (c->feed.consumed[i]) cudaEventDestroy(c->feed.consumed[i]);
    }
    if (c->feed.copy_stream) cudaStreamDestroy(c->feed.copy_stream);
    if (c->feed.t_stream) cudaStreamDestroy(c->feed.t_stream);
    if (c->feed.read_stream) cudaStreamDestroy(c->feed.read_stream);
    if (c->feed.cost_pinned) cudaFreeHost(c->feed.cost_pinned);
    if (getenv("POZ_KERNEL_TRACE") && atoi(getenv("POZ_KERNEL_TRACE")) != 0) { dump_kernel_trace(); tcs::trace_dump(); }
    host_trace_dump();
    if (getenv("POZ_X32T_PROBE") && atoi(getenv("POZ_X32T_PROBE")) != 0) tc3::x32t_probe_dump();
    free_workspace(c);
    for (auto e : c->grad_ready) if (e) cudaEventDestroy(e);
    for (auto e : c->dz_ready) if (e) cudaEventDestroy(e);
    if (c->side_stream) cudaStreamDestroy(c->side_stream);
    if (c->side_stream2) cudaStreamDestroy(c->side_stream2);
    if (c->sb_fork) cudaEventDestroy(c->sb_fork);
    if (c->sb_done) cudaEventDestroy(c->sb_done);
    for (auto& p : c->bn_run) dev_free(p);
    for (auto& p : c->bn_stat) dev_free(p);
    for (auto& p : c->bn_sums) dev_free(p);
    for (auto& p : c->ybuf) dev_free(p);
    for (auto& p : c->dybuf) dev_free(p);
    for (auto& p : c->dy_rec) dev_free(p);
    dev_free(c->ebuf); dev_free(c->debuf); dev_free(c->de_rec);
    dev_free(reinterpret_cast<void*&>(c->bn_part));
    dev_free(c->Pb); dev_free(reinterpret_cast<void*&>(c->colsum_scratch)); dev_free(c->wide_xg); dev_free(c->wide_tg);
    dev_free(c->Wp); dev_free(c->Wp16); dev_free(reinterpret_cast<void*&>(c->w16_amax)); dev_free(reinterpret_cast<void*&>(c->f16_flag));
    for (auto& p : c->tc_scr) dev_free(p);
    dev_free(c->tc_tr_x); dev_free(c->tc_tr_dz); for (auto& q : c->tcs_dzp) dev_free(q); dev_free(reinterpret_cast<void*&>(c->tail_part));
    dev_free(reinterpret_cast<void*&>(c->tc_logits));
    dev_free(c->P); dev_free(c->AM); dev_free(c->AV); dev_free(c->G); dev_free(c->Gdiag);
    dev_free(reinterpret_cast<void*&>(c->st));
    dev_free(reinterpret_cast<void*&>(c->cost_ring));
    dev_free(reinterpret_cast<void*&>(c->stats));
    dev_free(reinterpret_cast<void*&>(c->scalar_cost));
    dev_free(reinterpret_cast<void*&>(c->stage));
    delete c;
    return 0;
}

int poz_profile_begin(poz_ctx* c) {
    if (!c) POZ_FAIL("poz_profile_begin: NULL context");
    POZ_CUDA(cudaSetDevice(c->device));
    for (auto& r : c->prof) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    c->prof.clear();
    c->profiling = true;
    return 0;
}

int poz_profile_end(poz_ctx* c, char* json_out, long long cap) {
    if (!c) POZ_FAIL("poz_profile_end: NULL context");
    POZ_CUDA(cudaSetDevice(c->device));
    c->profiling = false;
    POZ_CUDA(cudaDeviceSynchronize());
    std::vector<std::string> tags;
    std::vector<double> ms;
    std::vector<long long> cnt;
    for (auto& r : c->prof) {
        float t = 0.f;
        cudaEventElapsedTime(&t, r.a, r.b);
        size_t i = 0;
        for (; i < tags.size(); ++i) if (tags[i] == r.tag) break;
        if (i == tags.size()) { tags.push_back(r.tag); ms.push_back(0.0); cnt.push_back(0); }
        ms[i] += t; cnt[i] += 1;
        cudaEventDestroy(r.a); cudaEventDestroy(r.b);
    }
    c->prof.clear();
    std::string js = "{";
    for (size_t i = 0; i < tags.size(); ++i) {
        char buf[160];
        snprintf(buf, sizeof(buf), "%s\"%s\": {\"launches\": %lld, \"ms\": %.6f}", i ? ", " : "",
                 tags[i].c_str(), cnt[i], ms[i]);
        js += buf;
    }
    js += "}";
    if (!json_out || cap <= (long long)js.size()) POZ_FAIL("poz_profile_end: output buffer too small");
    std::memcpy(json_out, js.c_str(), js.size() + 1);
    return 0;
}

int poz_ctx_reserve(poz_ctx* c, long long max_batch, int for_training) {
    if (!c) POZ_FAIL("poz_ctx_reserve: NULL context");
    POZ_CUDA(cudaSetDevice(c->device));
    if (for_training) return reserve_train(c, max_batch, false);
    return 0;
}

int poz_params_upload(poz_ctx* c, int layer, const double* w, const double* b) {
    POZ_TRY(check_layer(c, layer, "poz_params_upload"));
    POZ_TRY(POZ_DISPATCH(c, put_layer, c, c->P, layer, w, b));
    c->wp_stale = true; c->w16_stale = true;
    if (c->tcs_live && layer == 0) { POZ_TRY(tcs_split_w0(c, nullptr)); POZ_CUDA(cudaDeviceSynchronize()); }
    if (c->dtype == POZ_DTYPE_BF16) {
        const long long n = c->off[layer + 1] - c->off[layer];
        int blocks = (int)std::min<long long>((n + 255) / 256, 4096);
        convert_kernel<float, __nv_bfloat16><<<blocks, 256>>>(static_cast<const float*>(c->P) + c->off[layer],
                                                              static_cast<__nv_bfloat16*>(c->Pb) + c->off[layer], n);
        POZ_LAUNCH_CHECK();
        POZ_CUDA(cudaDeviceSynchronize());
    }
    return 0;
}
int poz_params_download(poz_ctx* c, int layer, double* w, double* b) {
    POZ_TRY(check_layer(c, layer, "poz_params_download"));
    return POZ_DISPATCH(c, get_layer, c, c->P, layer, w, b);
}
int poz_adam_upload(poz_ctx* c, int layer, const double* m_w, const double* v_w, const double* m_b,
                    const double* v_b) {
    POZ_TRY(check_layer(c, layer, "poz_adam_upload"));
    POZ_TRY(POZ_DISPATCH(c, put_layer, c, c->AM, layer, m_w, m_b));
    return POZ_DISPATCH(c, put_layer, c, c->AV, layer, v_w, v_b);
}
int poz_adam_download(poz_ctx* c, int layer, double* m_w, double* v_w, double* m_b, double* v_b) {
    POZ_TRY(check_layer(c, layer, "poz_adam_download"));
    POZ_TRY(POZ_DISPATCH(c, get_layer, c, c->AM, layer, m_w, m_b));
    return POZ_DISPATCH(c, get_layer, c, c->AV, layer, v_w, v_b);
}
int poz_tensor_io(poz_ctx* c, int kind, int layer, int buffer, int download, double* host) {
    POZ_TRY(check_layer(c, 0, "poz_tensor_io"));
    if (!host) POZ_FAIL("poz_tensor_io: NULL host buffer");
    void* base; long long off, n;
    POZ_TRY(tensor_locate(c, kind, layer, buffer, &base, &off, &n));
    if (download) return POZ_DISPATCH(c, get_f64, c, base, off, host, n);
    if (buffer == 3) POZ_FAIL("poz_tensor_io: the diagnostics gradient is read-only");
    c->wp_stale = true; c->w16_stale = true;
    POZ_TRY(POZ_DISPATCH(c, put_f64, c, base, off, host, n));
    if (c->tcs_live && buffer == 0 && kind == 0 && layer == 0) { POZ_TRY(tcs_split_w0(c, nullptr)); POZ_CUDA(cudaDeviceSynchronize()); }
    if (c->dtype == POZ_DTYPE_BF16 && buffer == 0 && kind <= 1) {
        int blocks = (int)std::min<long long>((n + 255) / 256, 4096);
        convert_kernel<float, __nv_bfloat16><<<blocks, 256>>>(static_cast<const float*>(c->P) + off,
                                                              static_cast<__nv_bfloat16*>(c->Pb) + off, n);
        POZ_LAUNCH_CHECK();
        POZ_CUDA(cudaDeviceSynchronize());
    }
    return 0;
}

int poz_adam_set_step(poz_ctx* c, long long step) {
    POZ_TRY(check_layer(c, 0, "poz_adam_set_step"));
    POZ_CUDA(cudaDeviceSynchronize());
    POZ_CUDA(cudaMemcpy(&c->st->adam_t, &step, sizeof(step), cudaMemcpyHostToDevice));
    return 0;
}
int poz_adam_get_step(poz_ctx* c, long long* step) {
    POZ_TRY(check_layer(c, 0, "poz_adam_get_step"));
    POZ_CUDA(cudaDeviceSynchronize());
    POZ_CUDA(cudaMemcpy(step, &c->st->adam_t, sizeof(*step), cudaMemcpyDeviceToHost));
    return 0;
}

static int forward_launch(poz_ctx* c, const void* x_dev, long long batch, void* out_dev, cudaStream_t s);

// Small inference calls are launch-bound (a handful of tiny kernels), so a call on the same buffers and
// batch size is captured into a CUDA graph the second time it is seen and replayed afterwards.  The
// weight planes / workspace a call needs are set up by the first, eager, call; uploads of new
// parameters go to the same device addresses, so a captured call stays valid.
int poz_forward(poz_ctx* c, const void* x_dev, long long batch, void* out_dev, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_forward"));
    if (batch <= 0) return 0;
    if (!x_dev || !out_dev) POZ_FAIL("poz_forward: NULL buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const bool graphable = c->graphs_enabled && !c->profiling && stream != nullptr && batch <= 4096 &&
                           !(c->dtype == POZ_DTYPE_F32 && batch >= c->tc3_min_batch && !c->general);
    if (!graphable) return forward_launch(c, x_dev, batch, out_dev, s);
    poz_ctx::StepGraph* g = nullptr;
    for (auto& e : c->fwd_graphs)
        if (e.x == x_dev && e.t == out_dev && e.batch == batch) { g = &e; break; }
    if (g && g->exec) {
        POZ_CUDA(cudaGraphLaunch(g->exec, s));
        count_launch(g->launches);
        return 0;
    }
    if (!g) {
        if (c->fwd_graphs.size() >= 64) {
            for (auto& e : c->fwd_graphs) if (e.exec) cudaGraphExecDestroy(e.exec);
            c->fwd_graphs.clear();
        }
        poz_ctx::StepGraph e;
        e.x = x_dev; e.t = out_dev; e.idx = nullptr; e.batch = batch;
        c->fwd_graphs.push_back(e);
        return forward_launch(c, x_dev, batch, out_dev, s);
    }
    const long long before = poz_launch_count();
    POZ_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    int rc = forward_launch(c, x_dev, batch, out_dev, s);
    cudaGraph_t graph = nullptr;
    cudaError_t ce = cudaStreamEndCapture(s, &graph);
    if (rc || ce != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        c->graphs_enabled = false;
        if (rc) return rc;
        return forward_launch(c, x_dev, batch, out_dev, s);
    }
    cudaGraphExec_t exec = nullptr;
    ce = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ce != cudaSuccess) POZ_FAIL(std::string("poz_forward: cudaGraphInstantiate failed: ") + cudaGetErrorString(ce));
    g->exec = exec;
    g->launches = (int)(poz_launch_count() - before);
    POZ_CUDA(cudaGraphLaunch(exec, s));
    return 0;
}

static int forward_launch(poz_ctx* c, const void* x_dev, long long batch, void* out_dev, cudaStream_t s) {
    if (c->dtype == POZ_DTYPE_BF16)
        return forward_wide(c, static_cast<const __nv_bfloat16*>(x_dev), nullptr, batch,
                            static_cast<__nv_bfloat16*>(out_dev), nullptr, s);
    if (c->dtype == POZ_DTYPE_F64)
        return forward_impl<double>(c, static_cast<const double*>(x_dev), nullptr, batch,
                                    static_cast<double*>(out_dev), nullptr, s);
    return forward_impl<float>(c, static_cast<const float*>(x_dev), nullptr, batch,
                               static_cast<float*>(out_dev), nullptr, s);
}

int poz_forward_loss(poz_ctx* c, const void* x_dev, const void* target_dev, long long batch,
                     void* out_dev, double* cost_host, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_forward_loss"));
    if (batch <= 0) POZ_FAIL("poz_forward_loss: empty batch");
    if (!x_dev || !out_dev || !target_dev || !cost_host) POZ_FAIL("poz_forward_loss: NULL buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (c->dtype == POZ_DTYPE_BF16)
        return forward_wide(c, static_cast<const __nv_bfloat16*>(x_dev), target_dev, batch,
                            static_cast<__nv_bfloat16*>(out_dev), cost_host, s);
    if (c->dtype == POZ_DTYPE_F64)
        return forward_impl<double>(c, static_cast<const double*>(x_dev), target_dev, batch,
                                    static_cast<double*>(out_dev), cost_host, s);
    return forward_impl<float>(c, static_cast<const float*>(x_dev), target_dev, batch,
                               static_cast<float*>(out_dev), cost_host, s);
}

int poz_train_begin(poz_ctx* c, const poz_train_config* cfg) {
    POZ_TRY(check_layer(c, 0, "poz_train_begin"));
    if (!cfg) POZ_FAIL("poz_train_begin: NULL config");
    if (cfg->dropout_rate < 0.0 || cfg->dropout_rate >= 1.0) POZ_FAIL("poz_train_begin: dropout_rate must be in [0,1)");
    if (cfg->optimizer != POZ_OPT_SGD && cfg->optimizer != POZ_OPT_ADAM) POZ_FAIL("poz_train_begin: bad optimizer");
    POZ_CUDA(cudaDeviceSynchronize());
    c->cfg = *cfg;
    long long epochs = std::max<long long>(cfg->epochs, 1);
    if (epochs > c->cost_cap) {
        dev_free(reinterpret_cast<void*&>(c->cost_ring));
        POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->cost_ring), (size_t)epochs * sizeof(double)));
        c->cost_cap = epochs;
    }
    long long recs = cfg->progress_stride > 0 ? (epochs + cfg->progress_stride - 1) / cfg->progress_stride + 1 : 1;
    if (recs > c->stats_cap) {
        dev_free(reinterpret_cast<void*&>(c->stats));
        POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->stats), (size_t)recs * n_stat_slots(c) * 4 * sizeof(double)));
        c->stats_cap = recs;
    }
    DevState h;
    POZ_CUDA(cudaMemcpy(&h, c->st, sizeof(h), cudaMemcpyDeviceToHost));
    h.lr0 = cfg->learning_rate; h.decay = cfg->decay_rate; h.beta1 = cfg->beta1; h.beta2 = cfg->beta2;
    h.eps = cfg->epsilon; h.l2 = cfg->l2_lambda; h.seed = cfg->dropout_seed;
    h.progress_stride = cfg->progress_stride; h.optimizer = cfg->optimizer;
    h.next_epoch = 0; h.cur_epoch = 0; h.cur_record = -1; h.n_records = 0;
    POZ_CUDA(cudaMemcpy(c->st, &h, sizeof(h), cudaMemcpyHostToDevice));
    c->train_begun = true;
    c->epoch_host = 0;
    c->feed.calls = 0;
    {   // captured steps bake these into kernel arguments; everything else lives in DevState
        const double sig[6] = {cfg->dropout_rate, (double)cfg->global_batch, (double)c->cost_cap, (double)c->stats_cap,
                               (double)reinterpret_cast<uintptr_t>(c->cost_ring), (double)reinterpret_cast<uintptr_t>(c->stats)};
        if (std::memcmp(sig, c->graph_sig, sizeof(sig)) != 0) {
            clear_graphs(c);
            std::memcpy(c->graph_sig, sig, sizeof(sig));
        }
    }
    c->grads_pending = false;
    return 0;
}

static int train_backward_inner(poz_ctx* c, const void* x_dev, const void* target_dev, const int* row_idx_dev,
                                long long batch, const void* const* keep_masks, int diagnostics, void* stream);

int poz_train_backward(poz_ctx* c, const void* x_dev, const void* target_dev, const int* row_idx_dev,
                       long long batch, const void* const* keep_masks, int diagnostics, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_train_backward"));
    POZ_TRY(train_backward_inner(c, x_dev, target_dev, row_idx_dev, batch, keep_masks, diagnostics, stream));
    c->epoch_host += 1;
    return 0;
}

static int train_backward_inner(poz_ctx* c, const void* x_dev, const void* target_dev, const int* row_idx_dev,
                                long long batch, const void* const* keep_masks, int diagnostics, void* stream) {
    if (!x_dev || !target_dev) POZ_FAIL("poz_train_step: NULL buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (c->dtype == POZ_DTYPE_BF16)
        return backward_wide(c, static_cast<const __nv_bfloat16*>(x_dev), target_dev, row_idx_dev, batch, keep_masks,
                             diagnostics, s);
    if (c->general)
        return c->dtype == POZ_DTYPE_F64 ? backward_general<double>(c, x_dev, target_dev, row_idx_dev, batch, keep_masks, diagnostics, s)
                                         : backward_general<float>(c, x_dev, target_dev, row_idx_dev, batch, keep_masks, diagnostics, s);
    if (c->dtype == POZ_DTYPE_F64) {
        POZ_TRY(backward_impl<double>(c, static_cast<const double*>(x_dev), target_dev, row_idx_dev, batch, keep_masks, diagnostics, s));
        return 0;
    }
    return backward_impl<float>(c, static_cast<const float*>(x_dev), target_dev, row_idx_dev, batch, keep_masks, diagnostics, s);
}

int poz_train_apply(poz_ctx* c, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_train_apply"));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    return c->dtype == POZ_DTYPE_F64 ? apply_impl<double>(c, c->grads_reduced, s)
                                     : apply_impl<float>(c, c->grads_reduced, s);
}

// A step on the same buffers and batch size launches the same kernels with the same arguments
// (everything that changes from step to step lives in DevState on the device), so from the second
// time a (x, target, row_idx, batch) key is seen the step is captured once into a CUDA graph and
// replayed: one graph launch instead of ~11 kernel launches of host work per step.
static int train_step_inner(poz_ctx* c, const void* x_dev, const void* target_dev, const int* row_idx_dev,
                            long long batch, const void* const* keep_masks, int diagnostics, void* stream);

int poz_train_step(poz_ctx* c, const void* x_dev, const void* target_dev, const int* row_idx_dev,
                   long long batch, const void* const* keep_masks, int diagnostics, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_train_step"));
    c->in_step = true;             // backward and update belong to one call: the update may ride in a backward epilogue
    const int rc = train_step_inner(c, x_dev, target_dev, row_idx_dev, batch, keep_masks, diagnostics, stream);
    c->in_step = false;
    return rc;
}

static int train_step_inner(poz_ctx* c, const void* x_dev, const void* target_dev, const int* row_idx_dev,
                            long long batch, const void* const* keep_masks, int diagnostics, void* stream) {
    const bool graphable = c->graphs_enabled && !c->profiling && !diagnostics && !keep_masks &&
                           c->dtype != POZ_DTYPE_BF16 && c->train_begun &&
                           stream != nullptr;   // the legacy default stream cannot be captured
    if (!graphable) {
        POZ_TRY(poz_train_backward(c, x_dev, target_dev, row_idx_dev, batch, keep_masks, diagnostics, stream));
        return poz_train_apply(c, stream);
    }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    poz_ctx::StepGraph* g = nullptr;
    for (auto& e : c->graphs)
        if (e.x == x_dev && e.t == target_dev && e.idx == row_idx_dev && e.batch == batch) { g = &e; break; }
    if (g && g->exec) {
        POZ_CUDA(cudaGraphLaunch(g->exec, s));
        count_launch(g->launches);
        c->epoch_host += 1;
        c->wp_stale = true; c->w16_stale = true;                 // the replayed update changed the parameters under the weight planes
        return 0;
    }
    if (!g) {
        // first sighting: run eagerly (this also performs every allocation the step needs)
        if (c->graphs.size() >= 256) clear_graphs(c);
        poz_ctx::StepGraph e;
        e.x = x_dev; e.t = target_dev; e.idx = row_idx_dev; e.batch = batch;
        c->graphs.push_back(e);
        POZ_TRY(poz_train_backward(c, x_dev, target_dev, row_idx_dev, batch, nullptr, 0, stream));
        return poz_train_apply(c, stream);
    }
    // second sighting: capture, instantiate, launch
    const long long before = poz_launch_count();
    const long long epoch_before = c->epoch_host;
    POZ_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    int rc = poz_train_backward(c, x_dev, target_dev, row_idx_dev, batch, nullptr, 0, stream);
    if (!rc) rc = poz_train_apply(c, stream);
    cudaGraph_t graph = nullptr;
    cudaError_t ce = cudaStreamEndCapture(s, &graph);
    if (rc || ce != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        if (!rc) set_error(std::string("poz_train_step: graph capture failed: ") + cudaGetErrorString(ce));
        c->graphs_enabled = false;          // fall back to plain launches for this context
        cudaGetLastError();
        c->epoch_host = epoch_before;       // the captured (never executed) attempt does not count
        if (rc) return rc;
        POZ_TRY(poz_train_backward(c, x_dev, target_dev, row_idx_dev, batch, nullptr, 0, stream));
        return poz_train_apply(c, stream);
    }
    cudaGraphExec_t exec = nullptr;
    ce = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ce != cudaSuccess) POZ_FAIL(std::string("poz_train_step: cudaGraphInstantiate failed: ") + cudaGetErrorString(ce));
    g->exec = exec;
    g->launches = (int)(poz_launch_count() - before);
    POZ_CUDA(cudaGraphLaunch(exec, s));
    c->wp_stale = true; c->w16_stale = true;
    return 0;
}

#include "poz_api_host.inc"

int poz_grad_buffer(poz_ctx* c, void** dev_ptr, long long* n_elems, int* dtype_code, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_grad_buffer"));
    if (!c->grads_pending) POZ_FAIL("poz_grad_buffer: run poz_train_backward first");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    POZ_TRY(c->dtype == POZ_DTYPE_F64 ? finalize_impl<double>(c, s) : finalize_impl<float>(c, s));
    c->grads_reduced = true;
    if (dev_ptr) *dev_ptr = c->G;
    if (n_elems) *n_elems = c->p_total;
    if (dtype_code) *dtype_code = c->dtype;
    return 0;
}

int poz_grad_layer(poz_ctx* c, int layer, void** dev_ptr, long long* n_elems, void* stream) {
    POZ_TRY(check_layer(c, layer, "poz_grad_layer"));
    if (!c->grads_pending) POZ_FAIL("poz_grad_layer: run poz_train_backward first");
    // per-layer buckets cover the Linear [W;b] images only: an embedding table, batchnorm gain/bias and the batch
    // statistics of a batchnorm would stay per-shard, so data-parallel training of such stacks is refused
    if (c->general) POZ_FAIL("poz_grad_layer: embedding / flatten / batchnorm stacks are not data-parallel (Linear + activation stacks only)");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    POZ_CUDA(cudaStreamWaitEvent(s, c->grad_ready[layer], 0));
    const long long n = ((long long)c->kin[layer] + c->has_bias) * c->sizes[layer + 1];
    if (c->seg_splits[layer] > 1) {
        SegTable seg;
        fill_seg(c, seg);
        for (int l = 0; l < c->nl; ++l) seg.splits[l] = l == layer ? c->seg_splits[l] : 1;
        int bx = (int)std::min<long long>((n + 255) / 256, 2LL * num_sms(c->device));
        ProfScope ps(c, "grad_finalize", s);
        if (c->dtype == POZ_DTYPE_F64)
            grad_finalize_kernel<double><<<dim3(std::max(bx, 1), seg.n), 256, 0, s>>>(static_cast<double*>(c->G), c->p_total, seg);
        else
            grad_finalize_kernel<float><<<dim3(std::max(bx, 1), seg.n), 256, 0, s>>>(static_cast<float*>(c->G), c->p_total, seg);
        POZ_LAUNCH_CHECK();
        c->seg_splits[layer] = 1;
    }
    bool all = true;
    for (int l = 0; l < c->nl; ++l) all &= c->seg_splits[l] == 1;
    if (all) c->grads_reduced = true;
    if (dev_ptr) *dev_ptr = static_cast<char*>(c->G) + c->off[layer] * c->esz;
    if (n_elems) *n_elems = n;
    return 0;
}

int poz_train_costs(poz_ctx* c, long long first, long long count, double* costs_host, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_train_costs"));
    if (first < 0 || count < 0 || first + count > c->cost_cap) POZ_FAIL("poz_train_costs: range outside the cost ring");
    if (count == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    POZ_CUDA(cudaMemcpyAsync(costs_host, c->cost_ring + first, count * sizeof(double), cudaMemcpyDeviceToHost, s));
    POZ_CUDA(cudaStreamSynchronize(s));
    return 0;
}

int poz_train_update_ratios(poz_ctx* c, long long first_record, long long count, double* ratios_host, void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_train_update_ratios"));
    if (first_record < 0 || count < 0 || first_record + count > c->stats_cap) POZ_FAIL("poz_train_update_ratios: range outside the record buffer");
    if (count == 0) return 0;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int ns = n_stat_slots(c), e = c->emb_vocab > 0 ? 1 : 0;
    std::vector<double> raw((size_t)count * ns * 4);
    POZ_CUDA(cudaMemcpyAsync(raw.data(), c->stats + first_record * ns * 4, raw.size() * sizeof(double), cudaMemcpyDeviceToHost, s));
    POZ_CUDA(cudaStreamSynchronize(s));
    for (long long r = 0; r < count; ++r)
        for (int k = 0; k < ns; ++k) {
            const double* q = &raw[(size_t)(r * ns + k) * 4];
            const double n = (e && k == 0) ? (double)c->emb_vocab * c->emb_dim
                                           : (double)c->kin[k - e] * c->sizes[k - e + 1];
            // unbiased std from sums (neural_net_model.py:506)
            double var_d = n > 1 ? (q[1] - q[0] * q[0] / n) / (n - 1) : NAN;
            double var_w = n > 1 ? (q[3] - q[2] * q[2] / n) / (n - 1) : NAN;
            double sd = std::sqrt(std::max(var_d, 0.0)), sw = std::sqrt(std::max(var_w, 0.0));
            if (n <= 1) { sd = NAN; sw = NAN; }
            ratios_host[r * ns + k] = sd / (sw + 1e-8);
        }
    return 0;
}

#include "poz_api_diag.inc"

}  // extern "C"
