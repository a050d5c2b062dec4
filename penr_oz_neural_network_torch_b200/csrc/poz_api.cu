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
// (Programmatic dependent launches were measured inside the step's CUDA graph in round 1: 108-114 us per C2 step against
//  82 us with plain edges, wherever the trigger was placed.  The attribute and its switch are gone; pdl_trigger / pdl_wait
//  in the kernels are no-ops for a launch without it.)
template <typename... P, typename... A>
static int launch_pdl(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
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

// ---- float32 contexts on the tensor cores (split-bf16, csrc/poz_tc3.*) ------------------------------
static long long round8(long long v) { return (v + 7) & ~7LL; }

// every layer deep enough (K >= 64) for a 64-wide k-block to be mostly real data
static bool tc3_stack_ok(const poz_ctx* c) {
    if (c->dtype != POZ_DTYPE_F32) return false;
    for (int l = 0; l < c->nl; ++l)
        if (c->sizes[l] < 64) return false;
    return true;
}

static int tc3_refresh_weights(poz_ctx* c, cudaStream_t stream, int upto = -1) {
    if (upto < 0) upto = c->nl;
    if (!c->Wp) {
        c->wp_off.assign(c->nl + 1, 0);
        c->wp_ld.assign(c->nl, 0);
        for (int l = 0; l < c->nl; ++l) {
            c->wp_ld[l] = round8(c->sizes[l + 1]);
            c->wp_off[l + 1] = c->wp_off[l] + 3LL * c->sizes[l] * c->wp_ld[l];
        }
        c->wp_total = c->wp_off[c->nl];
        POZ_TRY(dev_alloc(&c->Wp, (size_t)c->wp_total * 2));
        c->wp_stale = true; c->w16_stale = true;
    }
    if (c->wp_stale || c->wp_fresh.size() != (size_t)c->nl) {
        c->wp_fresh.assign(c->nl, 0);
        c->wp_stale = false;
    }
    if (c->tcs_live && !c->wp_fresh.empty()) c->wp_fresh[0] = 1;      // maintained by every writer of layer 0
    for (int l = 0; l < upto; ++l) {
        if (c->wp_fresh[l]) continue;
        c->wp_fresh[l] = 1;
        const int K = c->sizes[l], N = c->sizes[l + 1];
        ProfScope ps(c, "split_w", stream);
        POZ_TRY(tc3::split_planes(static_cast<const float*>(c->P) + c->off[l], nullptr, K, N, N,
                                  static_cast<__nv_bfloat16*>(c->Wp) + c->wp_off[l], c->wp_ld[l],
                                  (long long)K * c->wp_ld[l], stream));
    }
    return 0;
}

// Small-batch tensor-core path (csrc/poz_tcs_kernel.cuh): once a context has taken it, the bf16 planes of layer
// 0's weights are kept equal to the parameters at every point where those change -- the fused update writes them,
// every other writer (a plain update kernel, an upload) is followed by this split -- so a step replayed from a
// CUDA graph never depends on host-side freshness flags.
static int tcs_split_w0(poz_ctx* c, cudaStream_t stream) {
    const int K = c->sizes[0], N = c->sizes[1];
    ProfScope ps(c, "split_w0", stream);
    return tc3::split_planes(static_cast<const float*>(c->P) + c->off[0], nullptr, K, N, N,
                             static_cast<__nv_bfloat16*>(c->Wp) + c->wp_off[0], c->wp_ld[0], (long long)K * c->wp_ld[0], stream);
}

// fp16 planes of every layer's weights for the fp16-split inference kernels (made again whenever the parameters changed)
static int tc3_refresh_weights_f16(poz_ctx* c, cudaStream_t stream) {
    if (!c->Wp16) {
        POZ_TRY(dev_alloc(&c->Wp16, (size_t)(c->wp_total / 3 * 2) * 2));
        POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->w16_amax), sizeof(unsigned) * (size_t)c->nl));
        POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->f16_flag), sizeof(int)));
        c->w16_stale = true;
    }
    if (!c->w16_stale) return 0;
    POZ_CUDA(cudaMemsetAsync(c->w16_amax, 0, sizeof(unsigned) * (size_t)c->nl, stream));
    for (int l = 0; l < c->nl; ++l) {
        const int K = c->sizes[l], N = c->sizes[l + 1];
        ProfScope ps(c, "split_w16", stream);
        POZ_TRY(tc3::split_planes_f16(static_cast<const float*>(c->P) + c->off[l], K, N, N,
                                      static_cast<__nv_bfloat16*>(c->Wp16) + c->wp_off[l] / 3 * 2, c->wp_ld[l],
                                      (long long)K * c->wp_ld[l], c->w16_amax + l, stream));
    }
    c->w16_stale = false;
    return 0;
}

static tc3::PlaneMat tc3_weight_f16(const poz_ctx* c, int l) {
    tc3::PlaneMat m;
    m.ptr = static_cast<const __nv_bfloat16*>(c->Wp16) + c->wp_off[l] / 3 * 2;
    m.rows = c->sizes[l]; m.cols = c->sizes[l + 1]; m.ld = c->wp_ld[l];
    m.plane_stride = (long long)c->sizes[l] * c->wp_ld[l];
    m.planes = 2;
    return m;
}

static tc3::PlaneMat tc3_weight(const poz_ctx* c, int l) {
    tc3::PlaneMat m;
    m.ptr = static_cast<const __nv_bfloat16*>(c->Wp) + c->wp_off[l];
    m.rows = c->sizes[l]; m.cols = c->sizes[l + 1]; m.ld = c->wp_ld[l];
    m.plane_stride = (long long)c->sizes[l] * c->wp_ld[l];
    return m;
}

// Inference with every Linear layer on tcgen05: the input is split into bf16 planes once and each
// layer's epilogue writes the next layer's planes directly.
static int forward_tc3(poz_ctx* c, const float* x, const void* target, long long batch, float* out,
                       double* cost_host, cudaStream_t stream) {
    const int nl = c->nl, sms = num_sms(c->device);
    POZ_TRY(tc3_refresh_weights(c, stream));
    long long widest = 0;
    for (int l = 0; l <= nl; ++l) widest = std::max<long long>(widest, round8(c->sizes[l]));
    // Rows are processed in chunks only to bound the scratch planes (POZ_TC3_CHUNK_MB, default 3200 MB of
    // input + widest hidden planes, ~356k rows of the 784-wide stack), and a chunk is a whole number of waves
    // of the first (dominant) layer's 128x128 tiles.  An earlier version kept a chunk inside the L2 (100 MB):
    // measured at 262,144 rows that was 2.16 ms per pass against 1.17 ms with one chunk -- per-chunk launches
    // and tile-wave tails cost far more than the L2 hits saved (none of these kernels is HBM-bound).  Splitting
    // chunk i+1 into planes on a side stream under the GEMMs of chunk i was measured too: 1.24 ms, the two
    // compete for the same SMs and memory system.
    static const long long chunk_mb = getenv("POZ_TC3_CHUNK_MB") ? std::max(1, atoi(getenv("POZ_TC3_CHUNK_MB"))) : 3200;
    long long chunk = (chunk_mb << 20) / (6 * (round8(c->sizes[0]) + widest));
    {
        const long long tiles_n0 = (c->sizes[1] + 127) / 128;
        const long long rows_per_wave = std::max<long long>(1, sms / tiles_n0) * 128;
        chunk = std::max<long long>(1, chunk / rows_per_wave) * rows_per_wave;
    }
    chunk = std::min<long long>(batch, chunk);
    if (chunk > c->tc_chunk) {
        for (auto& p : c->tc_scr) dev_free(p);
        dev_free(reinterpret_cast<void*&>(c->tc_logits));
        for (auto& p : c->tc_scr) POZ_TRY(dev_alloc(&p, (size_t)3 * chunk * widest * 2));
        POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->tc_logits), (size_t)chunk * c->sizes[nl] * sizeof(float)));
        c->tc_chunk = chunk;
    }
    c->training_pass = false;
    if (cost_host) POZ_CUDA(cudaMemsetAsync(c->scalar_cost, 0, sizeof(double), stream));
    const int n_out = c->sizes[nl];
    const bool softmax_last = c->acts[nl - 1] == ACT_SOFTMAX;
    const float* P = static_cast<const float*>(c->P);
    for (long long r0 = 0; r0 < batch; r0 += chunk) {
        const long long m = std::min(chunk, batch - r0);
        // The first layer reads the fp32 rows directly and makes its planes in shared memory when it can
        // (poz_tc3_x32_kernel.cuh); otherwise the chunk is split into planes in global memory first.
        static const bool x32_allowed = !(getenv("POZ_TC3_X32") && atoi(getenv("POZ_TC3_X32")) == 0);
        // (measured, plane path vs this one with the planes in tensor memory: 41.4 vs 37.3 us per call at 2048 rows,
        //  43.9 vs 39.1 us at 4096; the shared-memory-plane version of the kernel only paid from 8192 rows)
        static const long long x32_min_rows = getenv("POZ_TC3_X32_MIN_ROWS") ? atoll(getenv("POZ_TC3_X32_MIN_ROWS")) : 2048;
        const bool x32 = x32_allowed && m >= x32_min_rows && tc3::gemm_forward_x32_ok(x + r0 * c->sizes[0], c->sizes[0], c->sizes[1]);
        tc3::PlaneMat cur;
        cur.ptr = static_cast<const __nv_bfloat16*>(c->tc_scr[0]);
        cur.rows = m; cur.cols = c->sizes[0]; cur.ld = round8(c->sizes[0]); cur.plane_stride = m * cur.ld;
        if (!x32) {
            ProfScope ps(c, "split_x", stream);
            POZ_TRY(tc3::split_planes(x + r0 * c->sizes[0], nullptr, m, c->sizes[0], c->sizes[0],
                                      static_cast<__nv_bfloat16*>(c->tc_scr[0]), cur.ld, cur.plane_stride, stream));
        }
        float* z_last = nullptr;
        // A narrow last unit (<= 16 outputs) behind a 65..128-wide one rides in that unit's epilogue: the tile holds
        // whole rows, so the epilogue warps apply the last Linear in fp32, then its activation / softmax, and the
        // last hidden activation never leaves the SM (POZ_TC3_FUSE_HEAD=0: separate kernels).
        // A hidden unit whose consumer can read fp32 rows (the fp32-input kernel: wide enough, rows 16-byte aligned,
        // enough of them) writes fp32 rows instead of three bf16 planes: 4 instead of 6 bytes per element each way.
        static const bool head_allowed = !(getenv("POZ_TC3_FUSE_HEAD") && atoi(getenv("POZ_TC3_FUSE_HEAD")) == 0);
        static const bool f32_chain = !(getenv("POZ_TC3_F32_CHAIN") && atoi(getenv("POZ_TC3_F32_CHAIN")) == 0);
        bool head_done = false;
        // One pass over the layers.  dry: nothing is launched, the return value says whether every unit would run on the
        // fp32-input kernel.  f16: that kernel in its fp16-split form (two planes, three MMAs per product), which reports
        // operands outside its range through c->f16_flag; run_if: kernels that return at once unless *run_if != 0.
        auto chain = [&](bool dry, bool f16, const int* run_if, bool* all_x32) -> int {
            int flip = 1;
            bool in_f32 = x32;
            const float* in_ptr = x + r0 * c->sizes[0];
            long long in_ld = c->sizes[0];
            tc3::PlaneMat cur_l = cur;
            if (all_x32) *all_x32 = true;
            for (int l = 0; l < nl; ++l) {
                const int K = c->sizes[l], N = c->sizes[l + 1];
                const bool last = l == nl - 1;
                tc3::Epi3 ep{};
                ep.mode = tc3::EPI_BIAS_ACT;
                ep.act = c->acts[l];
                ep.bias = c->has_bias ? P + c->off[l] + (long long)K * N : nullptr;
                ep.run_if = run_if;
                if (f16) { ep.f16_flag = c->f16_flag; ep.f16_w_amax = c->w16_amax + l; }
                tc3::PlaneMat nxt{};
                const bool head_here = head_allowed && !cost_host && l == nl - 2 && c->acts[l] != ACT_SOFTMAX &&
                                       tc3::gemm_forward_head_ok(N, n_out) && (!in_f32 || tc3::gemm_forward_x32_head_ok());
                bool out_f32_rows = false;
                if (head_here) {
                    ep.head_w = P + c->off[nl - 1];
                    ep.head_b = c->has_bias ? P + c->off[nl - 1] + (long long)N * n_out : nullptr;
                    ep.head_out = out + r0 * n_out;
                    ep.head_n = n_out;
                    ep.head_act = c->acts[nl - 1];
                } else if (last) {
                    ep.out_f32 = softmax_last ? c->tc_logits : out + r0 * n_out;
                    ep.ld_f32 = N;
                    if (!dry) z_last = ep.out_f32;
                } else if (f32_chain && x32_allowed && m >= x32_min_rows &&
                           tc3::gemm_forward_x32_ok(static_cast<const float*>(c->tc_scr[flip]), N, c->sizes[l + 2])) {
                    out_f32_rows = true;
                    ep.out_f32 = static_cast<float*>(c->tc_scr[flip]);
                    ep.ld_f32 = N;
                } else {
                    nxt.ptr = static_cast<const __nv_bfloat16*>(c->tc_scr[flip]);
                    nxt.rows = m; nxt.cols = N; nxt.ld = round8(N); nxt.plane_stride = m * nxt.ld;
                    ep.out_planes = static_cast<__nv_bfloat16*>(c->tc_scr[flip]);
                    ep.ld_planes = nxt.ld; ep.plane_stride = nxt.plane_stride;
                }
                if (!in_f32 && all_x32) *all_x32 = false;
                if (!dry) {
                    ProfScope ps(c, std::string(f16 ? "tc3h_fwd_l" : "tc3_fwd_l") + std::to_string(l) + (head_here ? "_head" : ""), stream);
                    if (in_f32)
                        POZ_TRY(tc3::gemm_forward_x32(in_ptr, in_ld, f16 ? tc3_weight_f16(c, l) : tc3_weight(c, l), m, N, K, ep, sms, stream));
                    else
                        POZ_TRY(tc3::gemm_forward(cur_l, tc3_weight(c, l), m, N, K, ep, sms, stream));
                }
                if (head_here) { if (!dry) head_done = true; break; }
                in_f32 = out_f32_rows;
                in_ptr = static_cast<const float*>(c->tc_scr[flip]);
                in_ld = N;
                cur_l = nxt;
                flip = flip == 1 ? 2 : 1;
            }
            return 0;
        };
        bool f16 = false;
        if (c->tc3_f16 && x32 && m >= c->tc3_f16_min_rows) POZ_TRY(chain(true, false, nullptr, &f16));
        if (f16) {
            // the fp16-split pass, then the split-bf16 pass as kernels that only run when the first one raised its flag
            POZ_TRY(tc3_refresh_weights_f16(c, stream));
            POZ_CUDA(cudaMemsetAsync(c->f16_flag, 0, sizeof(int), stream));
            POZ_TRY(chain(false, true, nullptr, nullptr));
            POZ_TRY(chain(false, false, c->f16_flag, nullptr));
        } else {
            POZ_TRY(chain(false, false, nullptr, nullptr));
        }
        if (cost_host) {
            LossArgs<float> a{};
            a.z = z_last; a.M = m; a.N = n_out; a.act = c->acts[nl - 1];
            a.cost_base = c->scalar_cost; a.cost_cap = 1; a.st = nullptr;
            if (c->loss == POZ_LOSS_CE) {
                a.target = static_cast<const int*>(target) + r0;
                a.probs = out + r0 * n_out;
                a.inv_count = 1.0 / (double)batch;
            } else {
                a.target = static_cast<const float*>(target) + r0 * n_out;
                a.inv_count = 1.0 / ((double)batch * n_out);
            }
            POZ_TRY(run_loss<float>(c, a, stream));
        } else if (softmax_last && !head_done) {
            POZ_TRY(run_softmax<float>(c, z_last, out + r0 * n_out, m, n_out, stream));
        }
    }
    if (cost_host) {
        POZ_CUDA(cudaMemcpyAsync(cost_host, c->scalar_cost, sizeof(double), cudaMemcpyDeviceToHost, stream));
        POZ_CUDA(cudaStreamSynchronize(stream));
    }
    return 0;
}

// ---- inference -----------------------------------------------------------------------------------
// Rows are processed in chunks small enough that the intermediate activations of a chunk stay in
// the 126 MB L2 between the per-layer kernels instead of round-tripping through HBM.
template <typename T>
static int forward_general_dispatch(poz_ctx* c, const void* x, const void* target, long long batch, T* out,
                                    double* cost_host, cudaStream_t stream);

// A few rows of a plain float32 stack: the whole forward pass in one cluster launch (poz_infer_small.cuh).
// Returns 1 through *taken when it ran.  POZ_SMALL_INFER=0 keeps such calls on the per-layer kernels.
static int forward_small(poz_ctx* c, const float* x, long long batch, float* out, cudaStream_t stream, bool* taken) {
    *taken = false;
    static const bool enabled = !(getenv("POZ_SMALL_INFER") && atoi(getenv("POZ_SMALL_INFER")) == 0);
    if (!enabled || c->small_infer_cluster < 0 || batch > si::SI_MAX_BATCH || c->nl > si::SI_MAX_LAYERS) return 0;
    for (int l = 0; l + 1 < c->nl; ++l)
        if (c->acts[l] == ACT_SOFTMAX) return 0;
    if (c->small_infer_cluster == 0) {
        // probe once: the largest cluster (16 needs the non-portable opt-in, else 8) this kernel can be launched with
        c->small_infer_cluster = -1;
        POZ_CUDA(cudaFuncSetAttribute(si::infer_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        const bool np = cudaFuncSetAttribute(si::infer_small_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) == cudaSuccess;
        cudaGetLastError();
        const int cs_max = getenv("POZ_SMALL_INFER_CLUSTER") ? atoi(getenv("POZ_SMALL_INFER_CLUSTER")) : 16;
        for (int cs : {16, 8}) {
            if ((cs == 16 && !np) || cs > cs_max) continue;
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(cs); cfg.blockDim = dim3(si::SI_THREADS); cfg.dynamicSmemBytes = 200 * 1024;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, si::infer_small_kernel, &cfg) == cudaSuccess && n >= 1) {
                c->small_infer_cluster = cs;
                break;
            }
            cudaGetLastError();
        }
        if (c->small_infer_cluster < 0) return 0;
    }
    const int C = c->small_infer_cluster;
    si::Args a{};
    a.nl = c->nl; a.has_bias = c->has_bias; a.batch = (int)batch;
    int max_w = 0, cols_cap = 1;
    for (int l = 0; l <= c->nl; ++l) { a.sizes[l] = c->sizes[l]; max_w = std::max(max_w, c->sizes[l]); }
    for (int l = 0; l < c->nl; ++l) {
        a.acts[l] = c->acts[l];
        a.wb[l] = static_cast<const float*>(c->P) + c->off[l];
        cols_cap = std::max(cols_cap, (c->sizes[l + 1] + C - 1) / C);
    }
    if (cols_cap > si::SI_MAX_COLS) return 0;
    a.max_w = max_w;
    a.x = x; a.out = out;
    const size_t smem = ((size_t)2 * batch * max_w + (size_t)si::SI_THREADS * batch) * sizeof(float);
    if (smem > 200 * 1024) return 0;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(C); cfg.blockDim = dim3(si::SI_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = C; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    c->training_pass = false;
    ProfScope ps(c, "infer_small", stream);
    POZ_CUDA(cudaLaunchKernelEx(&cfg, si::infer_small_kernel, a));
    count_launch();
    *taken = true;
    return 0;
}

template <typename T>
static int forward_impl(poz_ctx* c, const T* x, const void* target, long long batch, T* out,
                        double* cost_host, cudaStream_t stream) {
    const int nl = c->nl;
    if (c->general) return forward_general_dispatch<T>(c, x, target, batch, out, cost_host, stream);
    if constexpr (std::is_same<T, float>::value) {
        if (!cost_host && !target) {
            bool taken = false;
            POZ_TRY(forward_small(c, x, batch, out, stream, &taken));
            if (taken) return 0;
        }
    }
    if constexpr (std::is_same<T, float>::value) {
        // big batches of a deep-enough stack: every layer on the tensor cores
        if (batch >= c->tc3_min_batch && tc3_stack_ok(c)) return forward_tc3(c, x, target, batch, out, cost_host, stream);
    }
    int widest = 0;
    for (int l = 1; l <= nl; ++l) widest = std::max(widest, c->sizes[l]);
    long long chunk = std::max<long long>(256, (24LL << 20) / ((long long)widest * sizeof(T)));
    chunk = std::min(chunk, batch);
    POZ_TRY(reserve_infer(c, chunk));
    c->training_pass = false;
    if (cost_host) POZ_CUDA(cudaMemsetAsync(c->scalar_cost, 0, sizeof(double), stream));
    const int n_out = c->sizes[nl];
    for (long long r0 = 0; r0 < batch; r0 += chunk) {
        const long long m = std::min(chunk, batch - r0);
        const T* cur = x + r0 * c->sizes[0];
        T* bufs[2] = {static_cast<T*>(c->ping), static_cast<T*>(c->pong)};
        int flip = 0;
        for (int l = 0; l < nl; ++l) {
            const bool last = l == nl - 1;
            const bool softmax = c->acts[l] == ACT_SOFTMAX;
            T* dst = (last && !softmax) ? out + r0 * n_out : bufs[flip];
            // softmax units keep their logits; everything else only needs the activation output
            POZ_TRY(layer_forward<T>(c, l, cur, nullptr, m, m, softmax ? dst : nullptr,
                                     softmax ? nullptr : dst, nullptr, stream));
            cur = dst;
            flip ^= 1;
        }
        const bool softmax_last = c->acts[nl - 1] == ACT_SOFTMAX;
        if (cost_host) {
            LossArgs<T> a{};
            a.z = cur;
            a.M = m; a.N = n_out; a.act = c->acts[nl - 1];
            a.cost_base = c->scalar_cost; a.cost_cap = 1; a.st = nullptr;
            if (c->loss == POZ_LOSS_CE) {
                a.target = static_cast<const int*>(target) + r0;
                a.probs = out + r0 * n_out;
                a.inv_count = 1.0 / (double)batch;
            } else {
                a.target = static_cast<const T*>(target) + r0 * n_out;
                a.inv_count = 1.0 / ((double)batch * n_out);
            }
            POZ_TRY(run_loss<T>(c, a, stream));
        } else if (softmax_last) {
            POZ_TRY(run_softmax<T>(c, cur, out + r0 * n_out, m, n_out, stream));
        }
    }
    if (cost_host) {
        POZ_CUDA(cudaMemcpyAsync(cost_host, c->scalar_cost, sizeof(double), cudaMemcpyDeviceToHost, stream));
        POZ_CUDA(cudaStreamSynchronize(stream));
    }
    return 0;
}

// ---- training step -------------------------------------------------------------------------------
static void plan_splits(poz_ctx* c, long long batch, SegTable& seg, std::vector<long long>& kchunk) {
    const int sms = num_sms(c->device);
    kchunk.assign(c->nl, batch);
    for (int l = 0; l < c->nl; ++l) {
        const int M = c->sizes[l] + c->has_bias, N = c->sizes[l + 1];
        long long tiles;
        int bk;
        if (c->dtype == POZ_DTYPE_F64) {
            using TL = Tiles<double>;
            tiles = (N <= 16) ? (long long)((M + TL::TH_BM - 1) / TL::TH_BM) * ((N + TL::TH_BN - 1) / TL::TH_BN)
                              : (long long)((M + TL::SM_BM - 1) / TL::SM_BM) * ((N + TL::SM_BN - 1) / TL::SM_BN);
            bk = TL::SM_BK;
        } else {
            using TL = Tiles<float>;
            tiles = (N <= 16) ? (long long)((M + TL::TH_BM - 1) / TL::TH_BM) * ((N + TL::TH_BN - 1) / TL::TH_BN)
                              : (long long)((M + TL::SM_BM - 1) / TL::SM_BM) * ((N + TL::SM_BN - 1) / TL::SM_BN);
            bk = TL::SM_BK;
        }
        static const long long per_sm = getenv("POZ_DW_CTAS_PER_SM") ? std::max(1, atoi(getenv("POZ_DW_CTAS_PER_SM"))) : 4;
        long long want = (per_sm * sms + tiles - 1) / tiles;
        long long max_by_batch = std::max<long long>(1, batch / 64);
        long long s = std::max<long long>(1, std::min<long long>({want, max_by_batch, (long long)c->max_splits}));
        long long kc = (batch + s - 1) / s;
        kc = ((kc + bk - 1) / bk) * bk;
        s = (batch + kc - 1) / kc;
        seg.splits[l] = (int)s;
        kchunk[l] = kc;
    }
}

static int hb0(const poz_ctx* c) { return c->has_bias ? 1 : 0; }
static int n_stat_slots(const poz_ctx* c) { return c->nl + (c->emb_vocab > 0 ? 1 : 0); }

static void fill_seg(const poz_ctx* c, SegTable& seg) {
    const int e = c->emb_vocab > 0 ? 1 : 0;
    seg.n = c->nl;
    seg.n_stat = n_stat_slots(c);
    for (int l = 0; l < c->nl; ++l) {
        seg.off[l] = c->off[l];
        seg.w_elems[l] = (long long)c->kin[l] * c->sizes[l + 1];
        seg.n_elems[l] = ((long long)c->kin[l] + c->has_bias) * c->sizes[l + 1];
        seg.splits[l] = 1;
        seg.stat_slot[l] = e + l;
    }
    if (e) {
        const int k = seg.n++;
        seg.off[k] = c->off_emb;
        seg.n_elems[k] = seg.w_elems[k] = (long long)c->emb_vocab * c->emb_dim;
        seg.splits[k] = 1;
        seg.stat_slot[k] = 0;
    }
    for (int l = 0; l < c->nl; ++l)
        if (!c->has_bn.empty() && c->has_bn[l]) {
            const int k = seg.n++;
            seg.off[k] = c->off_bn[l];
            seg.n_elems[k] = 2LL * c->sizes[l + 1];
            seg.w_elems[k] = 0;
            seg.splits[k] = 1;
            seg.stat_slot[k] = -1;
        }
}

static long long biggest_seg(const SegTable& seg) {
    long long b = 0;
    for (int k = 0; k < seg.n; ++k) b = std::max(b, seg.n_elems[k]);
    return b;
}

template <typename T>
static int backward_impl(poz_ctx* c, const T* x, const void* target, const int* row_idx,
                         long long batch, const void* const* masks, int diag, cudaStream_t stream) {
    const int nl = c->nl;
    if (!c->train_begun) POZ_FAIL("poz_train_step: call poz_train_begin first");
    if (batch <= 0) POZ_FAIL("poz_train_step: empty batch");
    POZ_TRY(reserve_train(c, batch, diag != 0));
    SegTable seg;
    fill_seg(c, seg);
    std::vector<long long> kchunk;
    plan_splits(c, batch, seg, kchunk);
    int slabs = 1;
    for (int l = 0; l < nl; ++l) slabs = std::max(slabs, seg.splits[l]);
    // The first layer (the only deep contraction of a narrow stack) goes to the tensor cores
    // (split-bf16) when nothing needs the SIMT epilogue's extras: no dropout, masks or diagnostics.
    bool tc0 = false;
    int tc0_splits = 1;
    if constexpr (std::is_same<T, float>::value) {
        tc0 = c->tc3_train && nl >= 2 && c->sizes[0] >= 256 && c->sizes[1] >= 64 && batch >= c->tc3_train_min_batch && !diag && !masks &&
              c->cfg.dropout_rate == 0.0 && c->acts[0] != ACT_SOFTMAX;
        if (tc0) {
            const long long tiles = (long long)((c->sizes[0] + 127) / 128) * ((c->sizes[1] + 63) / 64);
            while (tc0_splits < 16 && tiles * tc0_splits * 2 <= num_sms(c->device) && batch / (tc0_splits * 2) >= 128)
                tc0_splits *= 2;
            seg.splits[0] = tc0_splits;
            slabs = std::max(slabs, tc0_splits);
            if (batch > c->tc_tr_batch) {
                dev_free(c->tc_tr_x); dev_free(c->tc_tr_dz);
                POZ_TRY(dev_alloc(&c->tc_tr_x, (size_t)3 * batch * round8(c->sizes[0]) * 2));
                POZ_TRY(dev_alloc(&c->tc_tr_dz, (size_t)3 * batch * round8(c->sizes[1]) * 2));
                c->tc_tr_batch = batch;
                clear_graphs(c);
            }
            if (!c->Wp) POZ_TRY(tc3_refresh_weights(c, stream, 0));   // allocation only
            if (!c->colsum_scratch) POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->colsum_scratch), (size_t)32 * c->sizes[1] * sizeof(float)));
        }
    }
    // Below the batch where the persistent plane kernels pay (tc0), the first layer's two deep contractions run as
    // cluster split-K tcgen05 launches that convert their fp32 operand on the fly (poz_tcs_kernel.cuh).
    bool tcs0 = false, tcs_fused = false, tcs_all = false;
    int tcs_f_splits = 1, tcs_w_splits = 1;
    if constexpr (std::is_same<T, float>::value) {
        const int sms = num_sms(c->device);
        const int K0 = c->sizes[0], N0 = c->sizes[1];
        tcs0 = c->tcs_enabled && !tc0 && nl >= 2 && c->acts[0] != ACT_SOFTMAX && K0 >= 128 && N0 >= 32 && N0 % 4 == 0 &&
               batch >= c->tcs_min_batch && batch <= c->tcs_max_batch;
        if (tcs0) {
            tcs_f_splits = tcs::pick_splits(K0, ((batch + tcs::BM - 1) / tcs::BM) * ((N0 + tcs::BN - 1) / tcs::BN), sms);
            tcs_w_splits = tcs::pick_splits(batch, (long long)((K0 + hb0(c) + tcs::BM - 1) / tcs::BM) * ((N0 + tcs::BN - 1) / tcs::BN), sms);
            tcs0 = tcs_f_splits > 0 && tcs_w_splits > 0;
        }
        if (tcs0) {
            if (!c->Wp) POZ_TRY(tc3_refresh_weights(c, stream, 0));   // allocation only
            if (!c->tcs_live) { POZ_TRY(tcs_split_w0(c, stream)); c->tcs_live = true; }
            // With the fused classifier head every layer's dZ is born in a kernel that can also write its planes, and
            // the weight gradients of ALL layers (same contraction length: the batch) go into ONE grouped launch
            // whose epilogues apply the optimizer: no side streams, no slabs, no update kernel.
            {
                const int hK = c->sizes[nl - 1], hN = c->sizes[nl];
                const size_t hsm = ((size_t)hK * (hN | 1) + 32 + 8 * (size_t)hK) * sizeof(T);
                const bool head = c->fuse_head && c->loss == POZ_LOSS_CE && hN <= 32 && hsm <= 96 * 1024;
                tcs_all = c->tcs_group && head && nl <= tcs::MAX_GROUP;
                if (tcs_all) {
                    long long tiles = 0;
                    for (int l = 0; l < nl; ++l)
                        tiles += (long long)((c->sizes[l] + hb0(c) + tcs::BM - 1) / tcs::BM) * ((c->sizes[l + 1] + tcs::BN - 1) / tcs::BN);
                    tcs_w_splits = tcs::pick_splits(batch, tiles, sms);
                }
            }
            if (c->tcs_dzp.size() != (size_t)nl) c->tcs_dzp.assign(nl, nullptr);
            const bool want_l1 = c->tail_fused && c->tail_group && nl == 3;     // the fused tail hands dZ1's planes to the grouped launch
            if (batch > c->tcs_dz_batch || (tcs_all && !c->tcs_dzp[nl - 1]) || (want_l1 && !c->tcs_dzp[1])) {
                for (int l = 0; l < nl; ++l) {
                    dev_free(c->tcs_dzp[l]);
                    if (l > 0 && !tcs_all && !(want_l1 && l == 1)) continue;
                    const size_t bytes = (size_t)3 * batch * round8(c->sizes[l + 1]) * 2;
                    POZ_TRY(dev_alloc(&c->tcs_dzp[l], bytes));
                    POZ_CUDA(cudaMemsetAsync(c->tcs_dzp[l], 0, bytes, stream));     // pad columns stay zero
                }
                c->tcs_dz_batch = batch;
                clear_graphs(c);
            }
            tcs_fused = c->in_step && c->tcs_fuse_update;
            seg.splits[0] = 1;
            if (tcs_all)
                for (int l = 0; l < nl; ++l) seg.splits[l] = 1;
        }
    }
    POZ_TRY(ensure_grad_slabs(c, slabs));
    c->seg_splits.assign(seg.splits, seg.splits + nl);
    c->training_pass = true;
    c->grads_reduced = false;
    c->diag_batch = diag ? batch : 0;
    c->last_batch = batch;
    tc3::PlaneMat xp{};
    if (tc0) {
        xp.ptr = static_cast<const __nv_bfloat16*>(c->tc_tr_x);
        xp.rows = batch; xp.cols = c->sizes[0]; xp.ld = round8(c->sizes[0]); xp.plane_stride = batch * xp.ld;
    }

    static const bool sb_allowed = !(getenv("POZ_FORK_STEP_BEGIN") && atoi(getenv("POZ_FORK_STEP_BEGIN")) == 0);
    const bool sb_forked = sb_allowed && c->fork_dw && c->side_stream && !masks && c->cfg.dropout_rate == 0.0;
    {
        // The step's bookkeeping (epoch / Adam counters, lr_t, cost slot) is first needed by the loss kernel --
        // the forward kernels read the step state only for dropout -- so without dropout it runs on the side
        // stream beside the forward GEMMs instead of in front of them.
        const double share = c->cfg.global_batch > 0 ? (double)batch / (double)c->cfg.global_batch : 1.0;
        cudaStream_t bs = stream;
        if (sb_forked) {
            bs = c->side_stream;
            POZ_CUDA(cudaEventRecord(c->sb_fork, stream));
            POZ_CUDA(cudaStreamWaitEvent(bs, c->sb_fork, 0));
        }
        {
            ProfScope ps(c, "step_begin", bs);
            POZ_TRY(launch_pdl(step_begin_kernel, dim3(1), dim3(32), 0, bs, c->st, c->cost_ring, c->cost_cap, c->stats,
                               c->stats_cap, n_stat_slots(c), share));
        }
        if (sb_forked) POZ_CUDA(cudaEventRecord(c->sb_done, bs));
    }

    // narrow classifier head: last Linear + softmax/CE + its backward-x in one kernel (head_ce_kernel)
    const int head_K = c->sizes[nl - 1], head_N = c->sizes[nl];
    const size_t head_smem = ((size_t)head_K * (head_N | 1) + 32 + 8 * (size_t)head_K) * sizeof(T);
    const bool fuse_head = c->fuse_head && nl >= 2 && c->loss == POZ_LOSS_CE && head_N <= 32 && head_smem <= 96 * 1024;

    // The tail of the stack -- last hidden unit, head, and the backward-x of both -- as ONE launch (poz_tail_kernel.cuh)
    // when nothing needs the separate kernels' extras (dropout, masks, diagnostics).
    bool tail = false, tail_group = false;
    size_t tail_smem = 0;
    if constexpr (std::is_same<T, float>::value) {
        const int p = nl - 2;
        tail = c->tail_fused && fuse_head && p >= 1 && !diag && !masks && c->cfg.dropout_rate == 0.0 && !tcs_all &&
               batch <= c->tail_max_batch && c->sizes[p] <= TAIL_MAX_K1 && c->sizes[p] % 4 == 0 && c->sizes[p + 1] <= TAIL_MAX_N1 &&
               c->sizes[p + 1] % 4 == 0 && head_N <= TAIL_MAX_N2 && c->acts[p] != ACT_SOFTMAX;
        if (tail) {
            tail_smem = tail_smem_bytes(c->sizes[p], c->sizes[p + 1], head_N);
            tail = tail_smem <= 200 * 1024;
        }
        // ... and then the weight gradients of layers 0 and 1 fit ONE tensor-core launch (layer 1 without its bias row:
        // 28 + 4 clusters of 4 CTAs on the C2 stack, where 33 can be resident), while the head's parameters and b1 are
        // updated from the tail's per-block partial sums: four launches on the step's critical path, no side-stream GEMMs.
        if (tail && tcs0 && tcs_fused && c->tail_group && nl == 3 && c->tcs_dzp.size() == 3 && c->tcs_dzp[1]) {
            const long long t0 = (long long)((c->sizes[0] + hb0(c) + tcs::BM - 1) / tcs::BM) * ((c->sizes[1] + tcs::BN - 1) / tcs::BN);
            const long long t1 = (long long)((c->sizes[1] + tcs::BM - 1) / tcs::BM) * ((c->sizes[2] + tcs::BN - 1) / tcs::BN);
            const int sp = tcs::pick_splits(batch, t0 + t1, num_sms(c->device));
            if (sp == 4 && t0 + t1 <= 33) { tail_group = true; tcs_w_splits = sp; }
        }
    }
    // forward, keeping what the backward needs (and, for diagnostics, every entry's output)
    const T* cur = x;
    const int* cur_idx = row_idx;
    for (int l = 0; l < (fuse_head ? (tail ? nl - 2 : nl - 1) : nl); ++l) {
        T* zb = static_cast<T*>(c->zbuf[l]);
        T* ab = static_cast<T*>(c->abuf[l]);
        const int act = c->acts[l];
        T* z_out = nullptr;
        T* a_out = ab;
        if (act == ACT_SOFTMAX) { z_out = zb; a_out = nullptr; }
        else if (act != ACT_NONE && diag) z_out = zb;
        if constexpr (std::is_same<T, float>::value) {
            if (tc0 && l == 0) {
                const int K = c->sizes[0], N = c->sizes[1];
                // every step re-splits W0: the previous step's update changed it, and a step captured into a
                // graph must record the split whatever the host-side freshness flags said at capture time
                if (c->wp_fresh.size() == (size_t)c->nl) c->wp_fresh[0] = 0;
                POZ_TRY(tc3_refresh_weights(c, stream, 1));
                {
                    ProfScope ps(c, "split_x", stream);
                    POZ_TRY(tc3::split_planes(x, row_idx, batch, K, K, static_cast<__nv_bfloat16*>(c->tc_tr_x), xp.ld,
                                              xp.plane_stride, stream));
                }
                tc3::Epi3 ep{};
                ep.mode = tc3::EPI_BIAS_ACT; ep.act = act;
                ep.bias = c->has_bias ? static_cast<const float*>(c->P) + c->off[0] + (long long)K * N : nullptr;
                ep.out_f32 = ab; ep.ld_f32 = N;
                ProfScope ps(c, "tc3_fwd_l0", stream);
                POZ_TRY(tc3::gemm_forward(xp, tc3_weight(c, 0), batch, N, K, ep, num_sms(c->device), stream));
                cur = ab;
                cur_idx = nullptr;
                continue;
            }
        }
        if constexpr (std::is_same<T, float>::value) {
            if (tcs0 && l == 0) {
                const int K = c->sizes[0], N = c->sizes[1];
                const float* w = static_cast<const float*>(c->P) + c->off[0];
                tcs::Gemm g{};
                g.a = x; g.lda = K; g.row_idx = row_idx; g.a_transposed = 0; g.a_cols = K; g.ones_row = 0;
                g.M = batch; g.N = N; g.K = K; g.b = tc3_weight(c, 0); g.splits = tcs_f_splits;
                tcs::FwdEpi fe{};
                fe.inner.bias = c->has_bias ? w + (long long)K * N : nullptr;
                fe.inner.inner.z_out = z_out; fe.inner.inner.a_out = a_out; fe.inner.inner.ld = N; fe.inner.inner.act = act;
                fe.inner.inner.dz = make_drop(c, 0, 0, masks);
                fe.inner.inner.da = make_drop(c, 0, 1, masks);
                fe.inner.inner.st = c->st;
                ProfScope ps(c, "tcs_fwd_l0", stream);
                POZ_TRY(tcs::forward(g, fe, stream));
                cur = ab;
                cur_idx = nullptr;
                continue;
            }
        }
        POZ_TRY(layer_forward<T>(c, l, cur, cur_idx, batch, batch, z_out, a_out, masks, stream));
        cur = (act == ACT_SOFTMAX) ? zb : ab;
        cur_idx = nullptr;
    }

    if (sb_forked) POZ_CUDA(cudaStreamWaitEvent(stream, c->sb_done, 0));
    // loss and the gradient of the last linear's pre-dropout output
    const long long gb = c->cfg.global_batch > 0 ? c->cfg.global_batch : batch;
    const int last = nl - 1;
    const int n_out = c->sizes[nl];
    bool tail_done = false;
    if constexpr (std::is_same<T, float>::value) {
        if (tail) {
            const int p = last - 1;
            static bool configured[64] = {};                       // a function attribute is per device
            if (c->device < 0 || c->device >= 64 || !configured[c->device]) {
                POZ_CUDA(cudaFuncSetAttribute(tail_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                if (c->device >= 0 && c->device < 64) configured[c->device] = true;
            }
            TailArgs ta{};
            ta.a0 = cur;
            ta.w1b = static_cast<const float*>(c->P) + c->off[p];
            ta.w2b = static_cast<const float*>(c->P) + c->off[last];
            ta.labels = static_cast<const int*>(target); ta.row_idx = row_idx;
            ta.a1_out = static_cast<float*>(c->abuf[p]);
            ta.dz2_out = static_cast<float*>(c->dzbuf[last]);
            ta.dz1_out = static_cast<float*>(c->dzbuf[p]);
            ta.dz0_out = static_cast<float*>(c->dzbuf[p - 1]);
            if (tcs0 && p - 1 == 0) {
                ta.dz0_planes = static_cast<unsigned short*>(c->tcs_dzp[0]);
                ta.dz0_ldp = round8(c->sizes[p]); ta.dz0_pstride = batch * round8(c->sizes[p]);
            }
            const int n_w2 = (c->sizes[p + 1] + c->has_bias) * head_N;              // parameters of the head
            const int pw2 = ((c->sizes[p + 1] + 1) * head_N + 3) & ~3;           // the tail always writes the bias row of the partial too
            const int n_blocks = (int)((batch + TAIL_ROWS - 1) / TAIL_ROWS);
            if (tail_group) {
                const long long need = (long long)n_blocks * (pw2 + c->sizes[p + 1]);
                if (need > c->tail_part_elems) {
                    dev_free(reinterpret_cast<void*&>(c->tail_part));
                    POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->tail_part), (size_t)need * sizeof(float)));
                    c->tail_part_elems = need;
                    clear_graphs(c);
                }
                ta.part = c->tail_part; ta.part_stride = pw2 + c->sizes[p + 1]; ta.part_b1 = pw2;
                ta.dz1_planes = static_cast<unsigned short*>(c->tcs_dzp[p]);
                ta.dz1_ldp = round8(c->sizes[p + 1]); ta.dz1_pstride = batch * round8(c->sizes[p + 1]);
            }
            ta.M = batch; ta.K1 = c->sizes[p]; ta.N1 = c->sizes[p + 1]; ta.N2 = head_N; ta.has_bias = c->has_bias;
            ta.act1 = c->acts[p]; ta.act0 = c->acts[p - 1];
            ta.inv_batch = 1.0 / (double)gb;
            ta.cost_base = c->cost_ring; ta.cost_cap = c->cost_cap; ta.st = c->st;
            {
                ProfScope ps(c, "tail", stream);
                // clusters of 4 (or 2) blocks share the reads of W1 (multicast within the cluster, from this GPU's own memory)
                cudaLaunchConfig_t cfg{};
                cfg.gridDim = dim3((unsigned)n_blocks); cfg.blockDim = dim3(TAIL_THREADS);
                cfg.dynamicSmemBytes = tail_smem; cfg.stream = stream;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = c->tail_cluster > 1 && n_blocks % 4 == 0 && c->tail_cluster >= 4 ? 4
                                         : (c->tail_cluster > 1 && n_blocks % 2 == 0 ? 2 : 1);
                at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                cfg.attrs = at; cfg.numAttrs = 1;
                POZ_CUDA(cudaLaunchKernelEx(&cfg, tail_fused_kernel, ta));
                count_launch();
            }
            tail_done = true;
            if (tail_group) {
                // the head's [W2; b2] and b1: partial sums -> optimizer, on the side stream beside the grouped launch
                cudaStream_t ws = c->side_stream;
                POZ_CUDA(cudaEventRecord(c->dz_ready[last], stream));
                POZ_CUDA(cudaStreamWaitEvent(ws, c->dz_ready[last], 0));
                auto mk = [&](int l) {
                    tcs::UpdateEpi ue{};
                    ue.p = static_cast<float*>(c->P) + c->off[l];
                    ue.am = static_cast<float*>(c->AM) + c->off[l];
                    ue.av = static_cast<float*>(c->AV) + c->off[l];
                    ue.ld = c->sizes[l + 1]; ue.w_rows = c->sizes[l]; ue.st = c->st;
                    ue.cost_base = c->cost_ring; ue.cost_cap = c->cost_cap;
                    ue.stats = c->stats; ue.n_stat = n_stat_slots(c); ue.stat_slot = seg.stat_slot[l];
                    return ue;
                };
                const int n_b1 = c->has_bias ? c->sizes[p + 1] : 0;
                const int total = n_w2 + n_b1;
                ProfScope ps(c, "tail_update", ws);
                POZ_TRY(launch_pdl(tail_update_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, ws,
                                   static_cast<const float*>(c->tail_part), n_blocks, pw2 + c->sizes[p + 1], pw2, n_w2, head_N, mk(last),
                                   n_b1, c->sizes[p], mk(p)));
                POZ_CUDA(cudaEventRecord(c->grad_ready[last], ws));
            }
        }
    }
    if (tail_done) {
    } else if (fuse_head) {
        const int p = last - 1;
        HeadArgs<T> h{};
        h.a_prev = cur;
        h.wb = static_cast<const T*>(c->P) + c->off[last];
        h.labels = static_cast<const int*>(target); h.row_idx = row_idx;
        h.dz_pre = static_cast<T*>(c->dzbuf[last]);
        h.dprev_pre = static_cast<T*>(c->dzbuf[p]);
        if (diag) {
            h.z_rec = static_cast<T*>(c->zbuf[last]); h.probs = static_cast<T*>(c->abuf[last]);
            h.dz_rec = static_cast<T*>(c->dz_rec[last]);
            if (c->acts[p] != ACT_NONE) { h.da_rec = static_cast<T*>(c->da_rec[p]); h.dzprev_rec = static_cast<T*>(c->dz_rec[p]); }
            else h.dzprev_rec = static_cast<T*>(c->dz_rec[p]);
        }
        h.M = batch; h.K = head_K; h.N = head_N; h.has_bias = c->has_bias; h.act_prev = c->acts[p];
        h.dz_last = make_drop(c, last, 0, masks);
        h.da_prev = make_drop(c, p, 1, masks);
        h.dz_prev = make_drop(c, p, 0, masks);
        h.inv_batch = 1.0 / (double)gb;
        h.cost_base = c->cost_ring; h.cost_cap = c->cost_cap; h.st = c->st;
        if constexpr (std::is_same<T, float>::value) {
            if (tcs_all) {
                h.dz_planes = static_cast<unsigned short*>(c->tcs_dzp[last]);
                h.dz_ldp = round8(head_N); h.dz_pstride = batch * round8(head_N);
                h.dprev_planes = static_cast<unsigned short*>(c->tcs_dzp[p]);
                h.dprev_ldp = round8(head_K); h.dprev_pstride = batch * round8(head_K);
            }
        }
        static bool configured[2] = {false, false};
        if (!configured[sizeof(T) == 8]) {
            POZ_CUDA(cudaFuncSetAttribute(head_ce_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
            configured[sizeof(T) == 8] = true;
        }
        const int blocks = (int)std::max<long long>(1, std::min<long long>((batch + 7) / 8, 2LL * num_sms(c->device)));
        {
            ProfScope ps(c, "head", stream);
            POZ_TRY(launch_pdl(head_ce_kernel<T>, dim3(blocks), dim3(256), head_smem, stream, h));
        }
    } else {
        LossArgs<T> a{};
        a.z = cur; a.target = target; a.row_idx = row_idx;
        a.M = batch; a.N = n_out; a.act = c->acts[last];
        a.dz_pre = static_cast<T*>(c->dzbuf[last]);
        a.dz = make_drop(c, last, 0, masks);
        a.cost_base = c->cost_ring; a.cost_cap = c->cost_cap; a.st = c->st;
        if (c->loss == POZ_LOSS_CE) {
            a.inv_count = 1.0 / (double)gb;
            a.probs = static_cast<T*>(c->abuf[last]);
            if (diag) a.dz_rec = static_cast<T*>(c->dz_rec[last]);
        } else {
            a.inv_count = 1.0 / ((double)gb * n_out);
            if (diag) {
                if (c->acts[last] != ACT_NONE) {
                    a.da_rec = static_cast<T*>(c->da_rec[last]);
                    a.dz_rec = static_cast<T*>(c->dz_rec[last]);
                } else {
                    a.dz_rec = static_cast<T*>(c->dz_rec[last]);
                }
            }
        }
        POZ_TRY(run_loss<T>(c, a, stream));
    }

    // backward: weight gradients (batch-split partial sums) and input gradients, last layer first
    // The weight-gradient GEMM of a layer >= 1 only needs that layer's dZ, so it is forked onto a side
    // stream and runs beside the backward-x chain (each of these launches fills a fraction of the GPU);
    // the compute stream joins again before the update.  Inside a graph capture this becomes a fork/join.
    const bool fork = c->fork_dw && c->side_stream && nl >= 2 && !tcs_all && !tail_group;
    bool dz0_planes_done = tail && tcs0 && nl == 3;      // the epilogue that produced layer 0's dZ also wrote its bf16 planes
    for (int l = last; l >= 0; --l) {
        if (tail_group) break;          // nothing left: the tail produced every dZ, the grouped launch below the rest
        const int K = c->sizes[l], N = c->sizes[l + 1], hb = c->has_bias;
        const T* dz = static_cast<const T*>(c->dzbuf[l]);
        cudaStream_t ws = stream;
      if (!tcs_all) {
        if (fork && l >= 1) {
            static const bool two = !(getenv("POZ_FORK_DW_STREAMS") && atoi(getenv("POZ_FORK_DW_STREAMS")) == 1);
            ws = (two && c->side_stream2 && (l & 1)) ? c->side_stream2 : c->side_stream;
            POZ_CUDA(cudaEventRecord(c->dz_ready[l], stream));
            POZ_CUDA(cudaStreamWaitEvent(ws, c->dz_ready[l], 0));
        }
        if constexpr (std::is_same<T, float>::value) {
          if (tc0 && l == 0) {
            tc3::PlaneMat dzp;
            dzp.ptr = static_cast<const __nv_bfloat16*>(c->tc_tr_dz);
            dzp.rows = batch; dzp.cols = N; dzp.ld = round8(N); dzp.plane_stride = batch * dzp.ld;
            {
                ProfScope ps(c, "split_dz", stream);
                POZ_TRY(tc3::split_planes(dz, nullptr, batch, N, N, static_cast<__nv_bfloat16*>(c->tc_tr_dz), dzp.ld,
                                          dzp.plane_stride, stream));
            }
            float* G = static_cast<float*>(c->G);
            tc3::Epi3 ep{};
            ep.mode = tc3::EPI_PLAIN;
            ep.out_f32 = G + c->off[0]; ep.ld_f32 = N; ep.slab = c->p_total;
            {
                ProfScope ps(c, "tc3_bwd_w_l0", stream);
                POZ_TRY(tc3::gemm_backward_weight(xp, dzp, batch, K, N, tc0_splits, ep, num_sms(c->device), stream));
            }
            if (hb) {
                ProfScope ps(c, "bias_grad", stream);
                float* gb_row = G + c->off[0] + (long long)K * N;
                for (int s2 = 1; s2 < tc0_splits; ++s2)
                    POZ_CUDA(cudaMemsetAsync(gb_row + (long long)s2 * c->p_total, 0, (size_t)N * sizeof(float), stream));
                colsum_f32_partial_kernel<<<dim3((N + 63) / 64, 32), 256, 0, stream>>>(dz, batch, N, c->colsum_scratch);
                POZ_LAUNCH_CHECK();
                colsum_f32_final_kernel<<<(N + 255) / 256, 256, 0, stream>>>(c->colsum_scratch, 32, N, gb_row);
                POZ_LAUNCH_CHECK();
            }
            if (c->grad_ready.size() == (size_t)nl) POZ_CUDA(cudaEventRecord(c->grad_ready[l], stream));
          } else if (tcs0 && l == 0) {
            tc3::PlaneMat dzp;
            dzp.ptr = static_cast<const __nv_bfloat16*>(c->tcs_dzp[0]);
            dzp.rows = batch; dzp.cols = N; dzp.ld = round8(N); dzp.plane_stride = batch * dzp.ld;
            if (!dz0_planes_done) {
                ProfScope ps(c, "split_dz0", stream);
                POZ_TRY(tc3::split_planes(dz, nullptr, batch, N, N, static_cast<__nv_bfloat16*>(c->tcs_dzp[0]), dzp.ld, dzp.plane_stride, stream));
            }
            tcs::Gemm g{};
            g.a = x; g.lda = K; g.row_idx = row_idx; g.a_transposed = 1; g.a_cols = K; g.ones_row = hb;
            g.M = (long long)K + hb; g.N = N; g.K = batch; g.b = dzp; g.splits = tcs_w_splits;
            if (tcs_fused) {
                tcs::UpdateEpi ue{};
                ue.p = static_cast<float*>(c->P) + c->off[0];
                ue.am = static_cast<float*>(c->AM) + c->off[0];
                ue.av = static_cast<float*>(c->AV) + c->off[0];
                if (diag) {
                    if (!c->Gdiag) POZ_TRY(dev_alloc(&c->Gdiag, (size_t)c->p_total * c->esz));
                    ue.g_diag = static_cast<float*>(c->Gdiag) + c->off[0];
                }
                ue.ld = N; ue.w_rows = K; ue.st = c->st; ue.cost_base = c->cost_ring; ue.cost_cap = c->cost_cap;
                ue.stats = c->stats; ue.n_stat = n_stat_slots(c); ue.stat_slot = seg.stat_slot[0];
                ue.planes = static_cast<__nv_bfloat16*>(c->Wp) + c->wp_off[0];
                ue.ldp = c->wp_ld[0]; ue.plane_stride = (long long)K * c->wp_ld[0];
                ProfScope ps(c, "tcs_bwd_w_l0_update", stream);
                POZ_TRY(tcs::backward_weight_update(&g, &ue, 1, stream));
                c->upd_done_mask |= 1u;
            } else {
                tcs::StoreEpi se{static_cast<float*>(c->G) + c->off[0], N};
                ProfScope ps(c, "tcs_bwd_w_l0", stream);
                POZ_TRY(tcs::backward_weight_store(&g, &se, 1, stream));
            }
            if (c->grad_ready.size() == (size_t)nl) POZ_CUDA(cudaEventRecord(c->grad_ready[l], stream));
          } else {
            const float* inp = l == 0 ? x : static_cast<const float*>(c->abuf[l - 1]);
            fast::Op A = make_fast_op(inp, K, l == 0 ? row_idx : nullptr, batch, K);
            fast::Op B = make_fast_op(dz, N, nullptr, batch, N);
            BwdWeightEpi<float> epi;
            epi.g = static_cast<float*>(c->G); epi.slab = c->p_total; epi.off = c->off[l]; epi.ld = N;
            {
                ProfScope ps(c, "bwd_w_l" + std::to_string(l), ws);
                if (hb) POZ_TRY((launch_fast<32, 64, false, false, true>(A, B, K, N, batch, seg.splits[l], false, K, epi, ws)));
                else    POZ_TRY((launch_fast<32, 64, false, false, false>(A, B, K, N, batch, seg.splits[l], false, K, epi, ws)));
            }
            if (c->grad_ready.size() == (size_t)nl && !(tcs_fused && fork && l >= 1)) POZ_CUDA(cudaEventRecord(c->grad_ready[l], ws));
          }
        } else {
            const T* inp = l == 0 ? x : static_cast<const T*>(c->abuf[l - 1]);
            Operand<T> A = make_operand<T>(inp, K, l == 0 ? row_idx : nullptr, batch, K, hb);
            Operand<T> B = make_operand<T>(dz, N, nullptr, batch, N, 0);
            BwdWeightEpi<T> epi;
            epi.g = static_cast<T*>(c->G); epi.slab = c->p_total; epi.off = c->off[l]; epi.ld = N;
            TileKind kind = N <= 16 ? TILE_THIN : TILE_SMALL;
            ProfScope ps(c, "bwd_w_l" + std::to_string(l), ws);
            POZ_TRY((launch_gemm<T, false, false>(kind, A, B, (long long)K + hb, N, batch,
                                                  seg.splits[l], kchunk[l], epi, ws)));
            if (c->grad_ready.size() == (size_t)nl) POZ_CUDA(cudaEventRecord(c->grad_ready[l], ws));
        }
      }
        if (l > 0 && !(fuse_head && l == last) && !(tail && l == last - 1)) {
            const T* w = static_cast<const T*>(c->P) + c->off[l];
            Operand<T> A = make_operand<T>(dz, N, nullptr, batch, N, 0);
            Operand<T> B = make_operand<T>(w, N, nullptr, K, N, 0);
            BwdDataEpi<T> epi;
            epi.a_prev = static_cast<const T*>(c->abuf[l - 1]);
            epi.dz_prev = static_cast<T*>(c->dzbuf[l - 1]);
            epi.ld = K;
            epi.act = c->acts[l - 1];
            epi.dz = make_drop(c, l - 1, 0, masks);
            epi.da = make_drop(c, l - 1, 1, masks);
            epi.st = c->st;
            epi.da_rec = nullptr; epi.dz_rec = nullptr;
            if constexpr (std::is_same<T, float>::value) {
                if (tcs0 && (l == 1 || tcs_all)) {   // dZ of the layer below leaves this epilogue as planes too
                    epi.planes = static_cast<unsigned short*>(c->tcs_dzp[l - 1]);
                    epi.ldp = round8(K); epi.plane_stride = batch * round8(K);
                    if (l == 1) dz0_planes_done = true;
                }
            }
            if (diag) {
                if (c->acts[l - 1] != ACT_NONE) {
                    epi.da_rec = static_cast<T*>(c->da_rec[l - 1]);
                    epi.dz_rec = static_cast<T*>(c->dz_rec[l - 1]);
                } else {
                    epi.dz_rec = static_cast<T*>(c->dz_rec[l - 1]);
                }
            }
            ProfScope ps(c, "bwd_x_l" + std::to_string(l), stream);
            if constexpr (std::is_same<T, float>::value) {
                fast::Op FA = make_fast_op(dz, N, nullptr, batch, N);
                fast::Op FB = make_fast_op(w, N, nullptr, K, N);
                const long long tiles = ((batch + 31) / 32) * ((K + 63) / 64);
                const int split = pick_cluster_split(c, tiles, N);
                POZ_TRY((launch_fast<32, 64, true, true, false>(FA, FB, batch, K, N, split, true, 0, epi, stream)));
            } else
            POZ_TRY((launch_gemm<T, true, true>(pick_tile<T>(c, batch, K), A, B, batch, K, (long long)N, 1,
                                                (long long)N, epi, stream)));
        }
        if constexpr (std::is_same<T, float>::value) {
            if (tcs_fused && fork && l >= 1 && !tcs_all) {
                // Layer 0's update rides in its weight-gradient launch; this layer's follows its own weight gradient on
                // the same side stream, so that no update kernel is left on the step's critical path.  It must not
                // overtake the backward-x launch above, which still reads this layer's weights.
                if (!(fuse_head && l == last) && !(tail && l == last - 1)) {
                    POZ_CUDA(cudaEventRecord(c->dz_ready[l], stream));
                    POZ_CUDA(cudaStreamWaitEvent(ws, c->dz_ready[l], 0));
                }
                SegTable one;
                fill_seg(c, one);
                for (int k = 0; k < one.n; ++k) { one.splits[k] = seg.splits[k]; if (k != l) one.n_elems[k] = 0; }
                float* gdiag = nullptr;
                if (diag) {
                    if (!c->Gdiag) POZ_TRY(dev_alloc(&c->Gdiag, (size_t)c->p_total * c->esz));
                    gdiag = static_cast<float*>(c->Gdiag);
                }
                const int bx = (int)std::max<long long>(1, std::min<long long>((one.n_elems[l] + 255) / 256, 2LL * num_sms(c->device)));
                ProfScope ps(c, "update_l" + std::to_string(l), ws);
                POZ_TRY(launch_pdl(update_kernel<float>, dim3(bx, one.n), dim3(256), 0, ws, static_cast<float*>(c->P),
                                   static_cast<float*>(c->AM), static_cast<float*>(c->AV), static_cast<const float*>(c->G), c->p_total,
                                   gdiag, one, static_cast<const DevState*>(c->st), c->cost_ring, c->cost_cap, c->stats,
                                   static_cast<__nv_bfloat16*>(c->Pb)));
                c->upd_done_mask |= 1u << l;
                if (c->grad_ready.size() == (size_t)nl) POZ_CUDA(cudaEventRecord(c->grad_ready[l], ws));
            }
        }
    }
    if (fork)
        for (int l = 1; l < nl; ++l) POZ_CUDA(cudaStreamWaitEvent(stream, c->grad_ready[l], 0));
    if constexpr (std::is_same<T, float>::value) {
        if (tail_group) {
            tcs::Gemm gs[2];
            tcs::UpdateEpi ues[2];
            for (int l = 0; l < 2; ++l) {
                const int K = c->sizes[l], N = c->sizes[l + 1], hb = (l == 0) ? c->has_bias : 0;   // b1 comes from the tail's partial sums
                tcs::Gemm& g = gs[l];
                g = tcs::Gemm{};
                g.a = l == 0 ? x : static_cast<const float*>(c->abuf[0]);
                g.lda = K; g.row_idx = l == 0 ? row_idx : nullptr; g.a_transposed = 1; g.a_cols = K; g.ones_row = hb;
                g.M = (long long)K + hb; g.N = N; g.K = batch; g.splits = tcs_w_splits;
                g.b.ptr = static_cast<const __nv_bfloat16*>(c->tcs_dzp[l]);
                g.b.rows = batch; g.b.cols = N; g.b.ld = round8(N); g.b.plane_stride = batch * round8(N);
                tcs::UpdateEpi& ue = ues[l];
                ue = tcs::UpdateEpi{};
                ue.p = static_cast<float*>(c->P) + c->off[l];
                ue.am = static_cast<float*>(c->AM) + c->off[l];
                ue.av = static_cast<float*>(c->AV) + c->off[l];
                ue.ld = N; ue.w_rows = K; ue.st = c->st; ue.cost_base = c->cost_ring; ue.cost_cap = c->cost_cap;
                ue.stats = c->stats; ue.n_stat = n_stat_slots(c); ue.stat_slot = seg.stat_slot[l];
                if (l == 0) {
                    ue.planes = static_cast<__nv_bfloat16*>(c->Wp) + c->wp_off[0];
                    ue.ldp = c->wp_ld[0]; ue.plane_stride = (long long)K * c->wp_ld[0];
                }
            }
            {
                ProfScope ps(c, "tcs_bwd_w_l0_l1_update", stream);
                POZ_TRY(tcs::backward_weight_update(gs, ues, 2, stream));
            }
            c->upd_done_mask |= 7u;
            if (c->grad_ready.size() == (size_t)nl) {
                POZ_CUDA(cudaEventRecord(c->grad_ready[0], stream));
                POZ_CUDA(cudaEventRecord(c->grad_ready[1], stream));
            }
            POZ_CUDA(cudaStreamWaitEvent(stream, c->grad_ready[last], 0));      // join the side stream (tail_update)
        }
        if (tcs_all) {
            // every layer's [dW; db] = [A 1]^T dZ in one grouped launch; its epilogues are the optimizer
            tcs::Gemm gs[tcs::MAX_GROUP];
            tcs::UpdateEpi ues[tcs::MAX_GROUP];
            tcs::StoreEpi ses[tcs::MAX_GROUP];
            if (diag && tcs_fused && !c->Gdiag) POZ_TRY(dev_alloc(&c->Gdiag, (size_t)c->p_total * c->esz));
            for (int l = 0; l < nl; ++l) {
                const int K = c->sizes[l], N = c->sizes[l + 1], hb = c->has_bias;
                tcs::Gemm& g = gs[l];
                g = tcs::Gemm{};
                g.a = l == 0 ? x : static_cast<const float*>(c->abuf[l - 1]);
                g.lda = K; g.row_idx = l == 0 ? row_idx : nullptr; g.a_transposed = 1; g.a_cols = K; g.ones_row = hb;
                g.M = (long long)K + hb; g.N = N; g.K = batch; g.splits = tcs_w_splits;
                g.b.ptr = static_cast<const __nv_bfloat16*>(c->tcs_dzp[l]);
                g.b.rows = batch; g.b.cols = N; g.b.ld = round8(N); g.b.plane_stride = batch * round8(N);
                tcs::UpdateEpi& ue = ues[l];
                ue = tcs::UpdateEpi{};
                ue.p = static_cast<float*>(c->P) + c->off[l];
                ue.am = static_cast<float*>(c->AM) + c->off[l];
                ue.av = static_cast<float*>(c->AV) + c->off[l];
                ue.g_diag = diag ? static_cast<float*>(c->Gdiag) + c->off[l] : nullptr;
                ue.ld = N; ue.w_rows = K; ue.st = c->st; ue.cost_base = c->cost_ring; ue.cost_cap = c->cost_cap;
                ue.stats = c->stats; ue.n_stat = n_stat_slots(c); ue.stat_slot = seg.stat_slot[l];
                if (l == 0) {
                    ue.planes = static_cast<__nv_bfloat16*>(c->Wp) + c->wp_off[0];
                    ue.ldp = c->wp_ld[0]; ue.plane_stride = (long long)K * c->wp_ld[0];
                }
                ses[l] = tcs::StoreEpi{static_cast<float*>(c->G) + c->off[l], N};
            }
            if (tcs_fused) {
                ProfScope ps(c, "tcs_bwd_w_all_update", stream);
                POZ_TRY(tcs::backward_weight_update(gs, ues, nl, stream));
                c->upd_done_mask |= (1u << nl) - 1u;
            } else {
                ProfScope ps(c, "tcs_bwd_w_all", stream);
                POZ_TRY(tcs::backward_weight_store(gs, ses, nl, stream));
            }
            if (c->grad_ready.size() == (size_t)nl)
                for (int l = 0; l < nl; ++l) POZ_CUDA(cudaEventRecord(c->grad_ready[l], stream));
        }
    }
    c->grads_pending = true;
    c->diag_pending = diag != 0;
    return 0;
}


// ---- wide (bf16 / tcgen05) path -------------------------------------------------------------------
// fp32 master parameters + Adam state (update_kernel<float>), bf16 parameter copy `Pb`, bf16
// activations and activation gradients, fp32 weight gradients straight from tensor memory.
static int forward_wide(poz_ctx* c, const __nv_bfloat16* x, const void* target, long long batch,
                        __nv_bfloat16* out, double* cost_host, cudaStream_t stream) {
    const int nl = c->nl;
    const int sms = num_sms(c->device);
    long long chunk = std::min<long long>(batch, 16384);
    POZ_TRY(reserve_infer(c, chunk));
    c->training_pass = false;
    if (cost_host) POZ_CUDA(cudaMemsetAsync(c->scalar_cost, 0, sizeof(double), stream));
    const int n_out = c->sizes[nl];
    const float* P = static_cast<const float*>(c->P);
    const __nv_bfloat16* Pb = static_cast<const __nv_bfloat16*>(c->Pb);
    for (long long r0 = 0; r0 < batch; r0 += chunk) {
        const long long m = std::min(chunk, batch - r0);
        const __nv_bfloat16* cur = x + r0 * c->sizes[0];
        __nv_bfloat16* bufs[2] = {static_cast<__nv_bfloat16*>(c->ping), static_cast<__nv_bfloat16*>(c->pong)};
        int flip = 0;
        for (int l = 0; l < nl; ++l) {
            const int K = c->sizes[l], N = c->sizes[l + 1];
            __nv_bfloat16* dst = (l == nl - 1) ? out + r0 * n_out : bufs[flip];
            ProfScope ps(c, "tc_fwd_l" + std::to_string(l), stream);
            POZ_TRY(tc::gemm_forward(cur, Pb + c->off[l], m, N, K,
                                     c->has_bias ? P + c->off[l] + (long long)K * N : nullptr, c->acts[l], dst, sms, stream));
            cur = dst;
            flip ^= 1;
        }
        if (cost_host) {
            WideLossArgs a{};
            a.y = cur; a.target = static_cast<const __nv_bfloat16*>(target) + r0 * n_out;
            a.total = m * n_out; a.act = c->acts[nl - 1];
            a.inv_count = 1.0 / ((double)batch * n_out);
            a.cost_base = c->scalar_cost; a.cost_cap = 1; a.st = nullptr;
            ProfScope ps(c, "loss", stream);
            int blocks = (int)std::min<long long>((a.total / 8 + 255) / 256, 148 * 8);
            wide_mse_loss_kernel<<<std::max(blocks, 1), 256, 0, stream>>>(a);
            POZ_LAUNCH_CHECK();
        }
    }
    if (cost_host) {
        POZ_CUDA(cudaMemcpyAsync(cost_host, c->scalar_cost, sizeof(double), cudaMemcpyDeviceToHost, stream));
        POZ_CUDA(cudaStreamSynchronize(stream));
    }
    return 0;
}

static int backward_wide(poz_ctx* c, const __nv_bfloat16* x, const void* target, const int* row_idx,
                         long long batch, const void* const* masks, int diag, cudaStream_t stream) {
    const int nl = c->nl;
    const int sms = num_sms(c->device);
    if (!c->train_begun) POZ_FAIL("poz_train_step: call poz_train_begin first");
    if (batch <= 0) POZ_FAIL("poz_train_step: empty batch");
    if (batch % 8) POZ_FAIL("poz_train_step: bf16 contexts need a batch that is a multiple of 8");
    POZ_TRY(reserve_train(c, batch, diag != 0));
    const int n_out = c->sizes[nl];
    const __nv_bfloat16* tgt = static_cast<const __nv_bfloat16*>(target);
    if (row_idx) {
        // TMA reads dense tiles: the sampled rows (inputs and targets) are materialised first -- 2 x batch x width x 2
        // bytes each way, ~1 % of the step at the C4 shape
        if (batch > c->wide_g_batch) {
            dev_free(c->wide_xg); dev_free(c->wide_tg);
            POZ_TRY(dev_alloc(&c->wide_xg, (size_t)batch * c->sizes[0] * 2));
            POZ_TRY(dev_alloc(&c->wide_tg, (size_t)batch * n_out * 2));
            c->wide_g_batch = batch;
        }
        ProfScope ps(c, "gather_rows", stream);
        const int bx = (int)std::min<long long>((batch * (c->sizes[0] / 8) + 255) / 256, 8LL * sms);
        wide_gather_rows_kernel<<<std::max(bx, 1), 256, 0, stream>>>(x, row_idx, batch, c->sizes[0], static_cast<__nv_bfloat16*>(c->wide_xg));
        POZ_LAUNCH_CHECK();
        const int bt = (int)std::min<long long>((batch * (n_out / 8) + 255) / 256, 8LL * sms);
        wide_gather_rows_kernel<<<std::max(bt, 1), 256, 0, stream>>>(tgt, row_idx, batch, n_out, static_cast<__nv_bfloat16*>(c->wide_tg));
        POZ_LAUNCH_CHECK();
        x = static_cast<const __nv_bfloat16*>(c->wide_xg);
        tgt = static_cast<const __nv_bfloat16*>(c->wide_tg);
    }
    // split the batch reduction of dW only when the output tiles alone cannot fill the GPU
    c->seg_splits.assign(nl, 1);
    int slabs = 1;
    for (int l = 0; l < nl; ++l) {
        long long tiles = (long long)((c->sizes[l] + tc_tile_m() - 1) / tc_tile_m()) * ((c->sizes[l + 1] + tc_tile_n() - 1) / tc_tile_n());
        long long s = 1;
        if (tiles < sms) s = std::min<long long>({(sms + tiles - 1) / tiles, std::max<long long>(1, batch / 2048), 16});
        c->seg_splits[l] = (int)s;
        slabs = std::max<int>(slabs, (int)s);
    }
    POZ_TRY(ensure_grad_slabs(c, slabs));
    int widest = 0;
    for (int l = 1; l <= nl; ++l) widest = std::max(widest, c->sizes[l]);
    if (!c->colsum_scratch) POZ_TRY(dev_alloc(reinterpret_cast<void**>(&c->colsum_scratch), (size_t)COLSUM_ROWS * widest * sizeof(float)));
    c->training_pass = true;
    c->grads_reduced = false;
    c->diag_batch = diag ? batch : 0;
    c->last_batch = batch;
    {
        ProfScope ps(c, "step_begin", stream);
        const double share = c->cfg.global_batch > 0 ? (double)batch / (double)c->cfg.global_batch : 1.0;
        step_begin_kernel<<<1, 32, 0, stream>>>(c->st, c->cost_ring, c->cost_cap, c->stats, c->stats_cap, n_stat_slots(c), share);
        POZ_LAUNCH_CHECK();
    }
    float* P = static_cast<float*>(c->P);
    float* G = static_cast<float*>(c->G);
    const __nv_bfloat16* Pb = static_cast<const __nv_bfloat16*>(c->Pb);
    const __nv_bfloat16* cur = x;
    for (int l = 0; l < nl; ++l) {
        const int K = c->sizes[l], N = c->sizes[l + 1];
        __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(c->abuf[l]);
        tc::Extras ex;
        ex.dz = make_drop(c, l, 0, masks);
        ex.da = make_drop(c, l, 1, masks);
        ex.st = c->st;
        if (diag && c->acts[l] != ACT_NONE) ex.z_rec = static_cast<__nv_bfloat16*>(c->zbuf[l]);
        const bool plain = !ex.dz.enabled && !ex.da.enabled && !ex.z_rec;
        ProfScope ps(c, "tc_fwd_l" + std::to_string(l), stream);
        POZ_TRY(tc::gemm_forward(cur, Pb + c->off[l], batch, N, K,
                                 c->has_bias ? P + c->off[l] + (long long)K * N : nullptr, c->acts[l], dst, sms, stream,
                                 plain ? nullptr : &ex));
        cur = dst;
    }
    const long long gb = c->cfg.global_batch > 0 ? c->cfg.global_batch : batch;
    const int last = nl - 1;
    {
        WideLossArgs a{};
        a.y = cur; a.target = tgt;
        a.dz = static_cast<__nv_bfloat16*>(c->dzbuf[last]);
        a.total = batch * n_out; a.act = c->acts[last];
        a.inv_count = 1.0 / ((double)gb * n_out);
        a.cost_base = c->cost_ring; a.cost_cap = c->cost_cap; a.st = c->st;
        a.N = n_out;
        a.dzp = make_drop(c, last, 0, masks);
        if (diag) {
            a.dz_rec = static_cast<__nv_bfloat16*>(c->dz_rec[last]);
            if (c->acts[last] != ACT_NONE) a.da_rec = static_cast<__nv_bfloat16*>(c->da_rec[last]);
        }
        ProfScope ps(c, "loss", stream);
        int blocks = (int)std::min<long long>((a.total / 8 + 255) / 256, 148 * 8);
        wide_mse_loss_kernel<<<std::max(blocks, 1), 256, 0, stream>>>(a);
        POZ_LAUNCH_CHECK();
    }
    for (int l = last; l >= 0; --l) {
        const int K = c->sizes[l], N = c->sizes[l + 1];
        const __nv_bfloat16* dz = static_cast<const __nv_bfloat16*>(c->dzbuf[l]);
        const __nv_bfloat16* inp = l == 0 ? x : static_cast<const __nv_bfloat16*>(c->abuf[l - 1]);
        {
            ProfScope ps(c, "tc_bwd_w_l" + std::to_string(l), stream);
            POZ_TRY(tc::gemm_backward_weight(inp, dz, batch, K, N, G + c->off[l], c->p_total, c->seg_splits[l], sms, stream));
        }
        if (c->has_bias) {
            ProfScope ps(c, "bias_grad", stream);
            float* gb_row = G + c->off[l] + (long long)K * N;
            for (int s2 = 1; s2 < c->seg_splits[l]; ++s2)
                POZ_CUDA(cudaMemsetAsync(gb_row + (long long)s2 * c->p_total, 0, (size_t)N * sizeof(float), stream));
            dim3 grid((unsigned)((N + 63) / 64), COLSUM_ROWS);
            wide_colsum_partial_kernel<<<grid, 256, 0, stream>>>(dz, batch, N, c->colsum_scratch);
            POZ_LAUNCH_CHECK();
            wide_colsum_final_kernel<<<(N + 255) / 256, 256, 0, stream>>>(c->colsum_scratch, N, gb_row);
            POZ_LAUNCH_CHECK();
        }
        POZ_CUDA(cudaEventRecord(c->grad_ready[l], stream));
        if (l > 0) {
            tc::Extras ex;
            ex.dz = make_drop(c, l - 1, 0, masks);
            ex.da = make_drop(c, l - 1, 1, masks);
            ex.st = c->st;
            if (diag) {
                ex.dz_rec = static_cast<__nv_bfloat16*>(c->dz_rec[l - 1]);
                if (c->acts[l - 1] != ACT_NONE) ex.da_rec = static_cast<__nv_bfloat16*>(c->da_rec[l - 1]);
            }
            const bool plain = !ex.dz.enabled && !ex.da.enabled && !ex.dz_rec && !ex.da_rec;
            ProfScope ps(c, "tc_bwd_x_l" + std::to_string(l), stream);
            POZ_TRY(tc::gemm_backward_data(dz, Pb + c->off[l], batch, K, N, static_cast<const __nv_bfloat16*>(c->abuf[l - 1]),
                                           c->acts[l - 1], static_cast<__nv_bfloat16*>(c->dzbuf[l - 1]), sms, stream,
                                           plain ? nullptr : &ex));
        }
    }
    c->grads_pending = true;
    c->diag_pending = diag != 0;
    return 0;
}

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
        if (c->feed.consumed[i]) cudaEventDestroy(c->feed.consumed[i]);
    }
    if (c->feed.copy_stream) cudaStreamDestroy(c->feed.copy_stream);
    if (c->feed.read_stream) cudaStreamDestroy(c->feed.read_stream);
    if (c->feed.cost_pinned) cudaFreeHost(c->feed.cost_pinned);
    if (getenv("POZ_KERNEL_TRACE") && atoi(getenv("POZ_KERNEL_TRACE")) != 0) { dump_kernel_trace(); tcs::trace_dump(); }
    host_trace_dump();
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

// ---- host-buffer stepping ---------------------------------------------------------------------------
static long long feed_x_row_bytes(const poz_ctx* c) {
    if (c->emb_vocab > 0) return (long long)c->tokens * 4;            // int32 token indices
    return (long long)c->sizes[0] * (long long)c->act_esz;
}
static long long feed_t_row_bytes(const poz_ctx* c) {
    return c->loss == POZ_LOSS_CE ? 4 : (long long)c->sizes[c->nl] * (long long)c->act_esz;
}

static int feed_reserve(poz_ctx* c, long long batch) {
    poz_ctx::HostFeed& f = c->feed;
    constexpr int SLOTS = poz_ctx::HostFeed::SLOTS;
    if (!f.copy_stream) {
        POZ_CUDA(cudaStreamCreateWithFlags(&f.copy_stream, cudaStreamNonBlocking));
        POZ_CUDA(cudaStreamCreateWithFlags(&f.read_stream, cudaStreamNonBlocking));
        for (int i = 0; i < SLOTS; ++i) {
            POZ_CUDA(cudaEventCreateWithFlags(&f.copied[i], cudaEventDisableTiming));
            POZ_CUDA(cudaEventCreateWithFlags(&f.consumed[i], cudaEventDisableTiming));
        }
        POZ_CUDA(cudaMallocHost(reinterpret_cast<void**>(&f.cost_pinned), sizeof(double)));
    }
    if (batch > f.batch) {
        POZ_CUDA(cudaDeviceSynchronize());
        for (int i = 0; i < SLOTS; ++i) {
            dev_free(f.x[i]); dev_free(f.t[i]);
            POZ_TRY(dev_alloc(&f.x[i], (size_t)(batch * feed_x_row_bytes(c))));
            POZ_TRY(dev_alloc(&f.t[i], (size_t)(batch * feed_t_row_bytes(c))));
        }
        f.batch = batch;
        f.calls = 0;
        clear_graphs(c);
    }
    return 0;
}

// cost of the step that last used `slot`: waits for that step only
static int feed_read_cost(poz_ctx* c, int slot, double* out) {
    poz_ctx::HostFeed& f = c->feed;
    // The pinned word is preset to a NaN payload no kernel produces and polled until the copy overwrites it: the
    // wait costs no driver call (the host side of this entry point is what bounds it on a slow host).  If nothing
    // arrives for a long time, fall back to a stream synchronise, which also surfaces a launch failure.
    static const unsigned long long SENTINEL = 0x7ff8dead0badc0deULL;
    volatile unsigned long long* word = reinterpret_cast<volatile unsigned long long*>(f.cost_pinned);
    *word = SENTINEL;
    POZ_CUDA(cudaStreamWaitEvent(f.read_stream, f.consumed[slot], 0));
    POZ_CUDA(cudaMemcpyAsync(f.cost_pinned, c->cost_ring + (f.epoch_of[slot] % c->cost_cap), sizeof(double),
                             cudaMemcpyDeviceToHost, f.read_stream));
    for (unsigned long long spins = 0; *word == SENTINEL; ++spins) {
        if (spins > 200ull * 1000ull * 1000ull) {
            POZ_CUDA(cudaStreamSynchronize(f.read_stream));
            break;
        }
    }
    unsigned long long bits = *word;
    if (bits == SENTINEL) POZ_FAIL("poz_train_step_host: the step's cost did not arrive");
    memcpy(out, &bits, sizeof(double));
    return 0;
}

// POZ_HOST_TRACE=1: host time spent in the segments of poz_train_step_host, printed when the context is destroyed
static bool host_trace_on() {
    static const bool on = getenv("POZ_HOST_TRACE") && atoi(getenv("POZ_HOST_TRACE")) != 0;
    return on;
}
static double g_host_seg[6] = {0, 0, 0, 0, 0, 0};
static long long g_host_calls = 0;
static inline double host_now() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
}
static void host_trace_dump() {
    if (!host_trace_on() || !g_host_calls) return;
    fprintf(stderr, "[poz host trace] poz_train_step_host, us per call over %lld calls: copies %.2f, copied-event + wait %.2f, "
                    "step launch %.2f, consumed-event %.2f, cost read-back %.2f\n", g_host_calls, g_host_seg[0] / g_host_calls,
            g_host_seg[1] / g_host_calls, g_host_seg[2] / g_host_calls, g_host_seg[3] / g_host_calls, g_host_seg[4] / g_host_calls);
    g_host_calls = 0;
    for (double& v : g_host_seg) v = 0;
}

int poz_train_step_host(poz_ctx* c, const void* x_host, const void* target_host, long long batch, double* done_cost,
                        void* stream) {
    POZ_TRY(check_layer(c, 0, "poz_train_step_host"));
    if (!x_host || !target_host) POZ_FAIL("poz_train_step_host: NULL buffer");
    if (!c->train_begun) POZ_FAIL("poz_train_step_host: call poz_train_begin first");
    if (batch <= 0) POZ_FAIL("poz_train_step_host: empty batch");
    POZ_TRY(feed_reserve(c, batch));
    poz_ctx::HostFeed& f = c->feed;
    constexpr int SLOTS = poz_ctx::HostFeed::SLOTS;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int slot = (int)(f.calls % SLOTS);
    // host -> device on the copy stream.  The slot is free: the step that last used it was issued SLOTS calls ago,
    // and the previous call did not return before the step issued POZ_HOST_LAG = SLOTS - 1 calls before IT had
    // finished (its cost read or the event wait below) -- no stream-side wait needed, one driver call fewer per step.
    const bool ht = host_trace_on();
    double t0 = ht ? host_now() : 0.0, t1;
    POZ_CUDA(cudaMemcpyAsync(f.x[slot], x_host, (size_t)(batch * feed_x_row_bytes(c)), cudaMemcpyHostToDevice, f.copy_stream));
    POZ_CUDA(cudaMemcpyAsync(f.t[slot], target_host, (size_t)(batch * feed_t_row_bytes(c)), cudaMemcpyHostToDevice, f.copy_stream));
    if (ht) { t1 = host_now(); g_host_seg[0] += t1 - t0; t0 = t1; }
    POZ_CUDA(cudaEventRecord(f.copied[slot], f.copy_stream));
    POZ_CUDA(cudaStreamWaitEvent(s, f.copied[slot], 0));
    if (ht) { t1 = host_now(); g_host_seg[1] += t1 - t0; t0 = t1; }
    f.epoch_of[slot] = c->epoch_host;
    POZ_TRY(poz_train_step(c, f.x[slot], f.t[slot], nullptr, batch, nullptr, 0, stream));
    if (ht) { t1 = host_now(); g_host_seg[2] += t1 - t0; t0 = t1; }
    POZ_CUDA(cudaEventRecord(f.consumed[slot], s));
    if (ht) { t1 = host_now(); g_host_seg[3] += t1 - t0; t0 = t1; }
    f.calls += 1;
    // the cost of the step issued POZ_HOST_LAG calls ago: only that step is waited for
    if (done_cost) {
        *done_cost = NAN;
        if (f.calls > POZ_HOST_LAG) POZ_TRY(feed_read_cost(c, (int)((f.calls - 1 - POZ_HOST_LAG) % SLOTS), done_cost));
    } else if (f.calls > POZ_HOST_LAG) {
        POZ_CUDA(cudaEventSynchronize(f.consumed[(f.calls - 1 - POZ_HOST_LAG) % SLOTS]));   // keep the host POZ_HOST_LAG steps ahead, no further
    }
    if (ht) { g_host_seg[4] += host_now() - t0; g_host_calls += 1; }
    return 0;
}

int poz_train_host_flush(poz_ctx* c, double* last_costs) {
    POZ_TRY(check_layer(c, 0, "poz_train_host_flush"));
    poz_ctx::HostFeed& f = c->feed;
    constexpr int SLOTS = poz_ctx::HostFeed::SLOTS;
    if (last_costs)
        for (int i = 0; i < POZ_HOST_LAG; ++i) last_costs[i] = NAN;
    if (!f.copy_stream || f.calls == 0) return 0;
    const long long outstanding = std::min<long long>(f.calls, POZ_HOST_LAG);
    for (long long i = 0; i < outstanding; ++i) {
        const long long call = f.calls - outstanding + i;          // 0-based index of the host step
        const int slot = (int)(call % SLOTS);
        if (last_costs) POZ_TRY(feed_read_cost(c, slot, last_costs + (POZ_HOST_LAG - outstanding + i)));
        else POZ_CUDA(cudaEventSynchronize(f.consumed[slot]));
    }
    f.calls = 0;
    return 0;
}

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

static int diag_ready(poz_ctx* c, int layer, const char* who) {
    POZ_TRY(check_layer(c, layer < 0 ? 0 : layer, who));
    if (c->diag_batch <= 0) POZ_FAIL(std::string(who) + ": no step was run with diagnostics enabled");
    return 0;
}

static int diag_tensor(poz_ctx* c, int layer, int which, int source, const void** ptr, long long* rows, int* cols) {
    *ptr = nullptr;
    if (layer < -1 || layer >= c->nl) POZ_FAIL("diagnostics: layer out of range");
    if ((source == 0 || source == 1) && c->diag_batch <= 0) POZ_FAIL("diagnostics: no step was run with diagnostics enabled");
    if (source == 3 && (!c->have_diag || !c->Gdiag)) POZ_FAIL("diagnostics: no weight gradients recorded");
    if (layer == -1) {                                  // the embedding head (and the flatten entries after it)
        if (c->emb_vocab <= 0) POZ_FAIL("diagnostics: the stack has no embedding");
        if (source <= 1) {
            *rows = c->diag_batch * c->tokens; *cols = c->emb_dim;
            *ptr = source == 0 ? c->ebuf : c->de_rec;
        } else {
            *rows = c->emb_vocab; *cols = c->emb_dim;
            *ptr = static_cast<const char*>(source == 2 ? c->P : c->Gdiag) + c->off_emb * c->esz;
        }
        return 0;
    }
    const int act = c->acts[layer];
    const bool bn = !c->has_bn.empty() && c->has_bn[layer];
    if (source <= 1) {
        *rows = c->diag_batch * (c->rmul.empty() ? 1 : c->rmul[layer]); *cols = c->sizes[layer + 1];
        if (source == 0) {
            if (which == 0) *ptr = (act == ACT_NONE && !bn) ? c->abuf[layer] : c->zbuf[layer];
            else if (which == 1 && act != ACT_NONE) *ptr = c->abuf[layer];
            else if (which == 2 && bn) *ptr = act == ACT_NONE ? c->abuf[layer] : c->ybuf[layer];
        } else {
            if (which == 0) *ptr = c->dz_rec[layer];
            else if (which == 1 && act != ACT_NONE && act != ACT_SOFTMAX) *ptr = c->da_rec[layer];
            else if (which == 2 && bn) *ptr = c->dy_rec[layer];
        }
    } else {
        *rows = c->kin[layer]; *cols = c->sizes[layer + 1];
        if (which == 2) {                                 // batchnorm [gain; bias]
            if (!bn) POZ_FAIL("diagnostics: the layer has no batchnorm");
            *rows = 2;
            *ptr = static_cast<const char*>(source == 2 ? c->P : c->Gdiag) + c->off_bn[layer] * c->esz;
        } else {
            *ptr = static_cast<const char*>(source == 2 ? c->P : c->Gdiag) + c->off[layer] * c->esz;
        }
    }
    return 0;
}

int poz_diag_activation(poz_ctx* c, int layer, int which, double* host_out) {
    POZ_TRY(diag_ready(c, layer, "poz_diag_activation"));
    const void* src; long long rows; int cols;
    POZ_TRY(diag_tensor(c, layer, which, 0, &src, &rows, &cols));
    if (!src) POZ_FAIL("poz_diag_activation: no such entry");
    if (diag_is_bf16(c, 0)) return get_bf16_f64(c, src, host_out, rows * cols);
    return POZ_DISPATCH(c, get_f64, c, src, 0, host_out, rows * cols);
}

int poz_diag_activation_grad(poz_ctx* c, int layer, int which, double* host_out, int* has_grad) {
    POZ_TRY(diag_ready(c, layer, "poz_diag_activation_grad"));
    const void* src; long long rows; int cols;
    POZ_TRY(diag_tensor(c, layer, which, 1, &src, &rows, &cols));
    // the trailing softmax is not on the loss graph: its .grad is None (neural_net_model.py:545)
    if (has_grad) *has_grad = src ? 1 : 0;
    if (!src) return 0;
    if (diag_is_bf16(c, 1)) return get_bf16_f64(c, src, host_out, rows * cols);
    return POZ_DISPATCH(c, get_f64, c, src, 0, host_out, rows * cols);
}

// which tensor a diagnostics request means: activation / its gradient of an entry, weights, weight gradient
int poz_diag_moments(poz_ctx* c, int layer, int which, int source, int rule, double threshold, double* out6) {
    POZ_TRY(check_layer(c, layer < 0 ? 0 : layer, "poz_diag_moments"));
    const void* ptr; long long rows; int cols;
    POZ_TRY(diag_tensor(c, layer, which, source, &ptr, &rows, &cols));
    out6[0] = 0.0;
    if (!ptr) return 0;                      // no such tensor (e.g. the trailing softmax has no gradient)
    const long long n = rows * cols;
    const int blocks = (int)std::min<long long>((n + 255) / 256, 512);
    POZ_TRY(ensure_stage(c, (size_t)blocks * 5));
    if (diag_is_bf16(c, source))
        diag_moments_kernel<__nv_bfloat16><<<blocks, 256>>>(static_cast<const __nv_bfloat16*>(ptr), rows, cols, rule, threshold, c->stage);
    else if (c->dtype == POZ_DTYPE_F64)
        diag_moments_kernel<double><<<blocks, 256>>>(static_cast<const double*>(ptr), rows, cols, rule, threshold, c->stage);
    else
        diag_moments_kernel<float><<<blocks, 256>>>(static_cast<const float*>(ptr), rows, cols, rule, threshold, c->stage);
    POZ_LAUNCH_CHECK();
    std::vector<double> part((size_t)blocks * 5);
    POZ_CUDA(cudaMemcpy(part.data(), c->stage, part.size() * sizeof(double), cudaMemcpyDeviceToHost));
    double s = 0, s2 = 0, mn = INFINITY, mx = -INFINITY, sat = 0;
    for (int b = 0; b < blocks; ++b) {
        s += part[b * 5]; s2 += part[b * 5 + 1]; sat += part[b * 5 + 4];
        mn = std::min(mn, part[b * 5 + 2]); mx = std::max(mx, part[b * 5 + 3]);
    }
    out6[0] = (double)n; out6[1] = s; out6[2] = s2; out6[3] = mn; out6[4] = mx;
    out6[5] = sat / ((rule == 3 || rule == 4) ? (double)rows : (double)n);
    return 0;
}

int poz_diag_histogram(poz_ctx* c, int layer, int which, int source, const double* edges, int bins, double* counts) {
    POZ_TRY(check_layer(c, layer < 0 ? 0 : layer, "poz_diag_histogram"));
    if (bins < 1 || bins > 128) POZ_FAIL("poz_diag_histogram: bins must be in [1,128]");
    const void* ptr; long long rows; int cols;
    POZ_TRY(diag_tensor(c, layer, which, source, &ptr, &rows, &cols));
    if (!ptr) POZ_FAIL("poz_diag_histogram: no such tensor");
    const long long n = rows * cols;
    POZ_TRY(ensure_stage(c, 512));
    double* d_edges = c->stage;
    unsigned long long* d_counts = reinterpret_cast<unsigned long long*>(c->stage + 256);
    POZ_CUDA(cudaMemcpy(d_edges, edges, (size_t)(bins + 1) * sizeof(double), cudaMemcpyHostToDevice));
    POZ_CUDA(cudaMemset(d_counts, 0, (size_t)bins * sizeof(unsigned long long)));
    const int blocks = (int)std::min<long long>((n + 255) / 256, 1184);
    if (diag_is_bf16(c, source))
        diag_histogram_kernel<__nv_bfloat16><<<blocks, 256>>>(static_cast<const __nv_bfloat16*>(ptr), n, d_edges, bins, d_counts);
    else if (c->dtype == POZ_DTYPE_F64)
        diag_histogram_kernel<double><<<blocks, 256>>>(static_cast<const double*>(ptr), n, d_edges, bins, d_counts);
    else
        diag_histogram_kernel<float><<<blocks, 256>>>(static_cast<const float*>(ptr), n, d_edges, bins, d_counts);
    POZ_LAUNCH_CHECK();
    std::vector<unsigned long long> h(bins);
    POZ_CUDA(cudaMemcpy(h.data(), d_counts, (size_t)bins * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    for (int i = 0; i < bins; ++i) counts[i] = (double)h[i];
    return 0;
}

int poz_diag_weight_grad(poz_ctx* c, int layer, double* w_grad, double* b_grad) {
    POZ_TRY(diag_ready(c, layer, "poz_diag_weight_grad"));
    if (!c->have_diag || !c->Gdiag) POZ_FAIL("poz_diag_weight_grad: no diagnostics gradients recorded");
    return POZ_DISPATCH(c, get_layer, c, c->Gdiag, layer, w_grad, b_grad);
}

}  // extern "C"
