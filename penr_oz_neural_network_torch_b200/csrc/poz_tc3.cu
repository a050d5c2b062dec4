// Host side of the split-bf16 tcgen05 GEMM: 3-D tensor maps over plane stacks, launchers, and the
// fp32-in / fp32-out building-block entry point poz_tc3_gemm.
#include "../../include/pozmlp.h"
#include "poz_tc3.h"
#include "poz_tc3_kernel.cuh"
#include "poz_tc3_x32t_kernel.cuh"
#include "poz_tcs_kernel.cuh"
#include "poz_tc.h"

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

namespace poz {
namespace tc3 {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

static int make_map3(CUtensorMap* map, const PlaneMat& m, int box_cols, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) POZ_FAIL("tcgen05 path: cuTensorMapEncodeTiled is not available from the driver");
    if ((reinterpret_cast<uintptr_t>(m.ptr) & 15) != 0 || (m.ld * 2) % 16 != 0 || (m.plane_stride * 2) % 16 != 0)
        POZ_FAIL("tcgen05 path: plane base, row pitch and plane stride must be 16-byte aligned");
    cuuint64_t gdim[3] = {(cuuint64_t)m.cols, (cuuint64_t)m.rows, (cuuint64_t)m.planes};
    cuuint64_t gstride[2] = {(cuuint64_t)m.ld * 2, (cuuint64_t)m.plane_stride * 2};
    cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<__nv_bfloat16*>(m.ptr), gdim, gstride, box,
                    estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) POZ_FAIL("tcgen05 path: cuTensorMapEncodeTiled (3-D) failed with code " + std::to_string((int)r));
    return 0;
}

template <int A_MAJOR, int B_MAJOR, int BN>
static int launch(const PlaneMat& a, const PlaneMat& b, long long M, int N, long long K, int splits, const Epi3& ep,
                  int sms, cudaStream_t stream) {
    using C = Cfg<BN>;
    static bool configured = false;
    auto kern = tc3_gemm_kernel<A_MAJOR, B_MAJOR, BN>;
    constexpr bool head_capable = (A_MAJOR == MAJOR_K && B_MAJOR == MAJOR_MN && BN == 128);
    constexpr int smem_max = C::SMEM_BYTES + (head_capable ? HEAD_SMEM_BYTES : 0);
    static_assert(smem_max <= 227 * 1024, "shared memory of the tc3 GEMM (with the fused head's buffers)");
    if (ep.head_w && (!head_capable || N > BN || splits != 1 || ep.head_n < 1 || ep.head_n > HEAD_MAX_N || !ep.head_out))
        POZ_FAIL("tc3 forward: fused head not available for this shape");
    if (!configured) {
        POZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
        configured = true;
    }
    const int smem_bytes = C::SMEM_BYTES + (ep.head_w ? HEAD_SMEM_BYTES : 0);
    CUtensorMap ma, mb;
    POZ_TRY(make_map3(&ma, a, 64, A_MAJOR == MAJOR_K ? BM : 64));
    POZ_TRY(make_map3(&mb, b, 64, B_MAJOR == MAJOR_K ? BN : 64));
    long long k_per = (K + splits - 1) / splits;
    k_per = ((k_per + BK - 1) / BK) * BK;
    const long long tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN) * splits;
    const int grid = (int)(tiles < sms ? tiles : sms);
    kern<<<grid, NUM_THREADS, smem_bytes, stream>>>(ma, mb, M, N, K, splits, k_per, ep);
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

template <int A_MAJOR, int B_MAJOR>
static int launch_bn(const PlaneMat& a, const PlaneMat& b, long long M, int N, long long K, int splits,
                     const Epi3& ep, int sms, cudaStream_t stream) {
    const long long tiles128 = ((M + BM - 1) / BM) * ((N + 127) / 128) * splits;
    if (!ep.head_w && (N <= 64 || tiles128 * 2 < sms)) return launch<A_MAJOR, B_MAJOR, 64>(a, b, M, N, K, splits, ep, sms, stream);
    return launch<A_MAJOR, B_MAJOR, 128>(a, b, M, N, K, splits, ep, sms, stream);
}

int split_planes(const float* src, const int* row_idx, long long rows, int cols, long long ld, __nv_bfloat16* dst,
                 long long ldp, long long plane_stride, cudaStream_t stream) {
    if (rows <= 0) return 0;
    const long long total = rows * (ldp / 8);
    int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    split_planes_kernel<<<blocks, 256, 0, stream>>>(src, row_idx, rows, cols, ld, dst, ldp, plane_stride);
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

int gemm_forward(const PlaneMat& x, const PlaneMat& w, long long M, int N, int K, const Epi3& ep, int sms, cudaStream_t s) {
    return launch_bn<MAJOR_K, MAJOR_MN>(x, w, M, N, K, 1, ep, sms, s);
}
bool gemm_forward_head_ok(int N, int head_n) { return N > 64 && N <= 128 && head_n >= 1 && head_n <= HEAD_MAX_N; }
bool gemm_forward_x32_head_ok() { return true; }
bool gemm_forward_x32_ok(const float* x, long long ld, int N) {
    return N > 64 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (ld % 4) == 0;
}

int split_planes_f16(const float* src, long long rows, int cols, long long ld, __nv_bfloat16* dst, long long ldp,
                     long long plane_stride, unsigned* amax_bits, cudaStream_t stream) {
    if (rows <= 0) return 0;
    const long long total = rows * (ldp / 8);
    int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
    split_planes_f16_kernel<<<blocks, 256, 0, stream>>>(src, rows, cols, ld, dst, ldp, plane_stride, amax_bits);
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

int gemm_forward_x32(const float* x, long long ld, const PlaneMat& w, long long M, int N, int K, const Epi3& ep, int sms,
                     cudaStream_t stream) {
    constexpr int BN = 128;
    using CT = CfgX32T<BN, 3, false>;
    using CH = CfgX32T<BN, 2, true>;                  // with the fused head: one W stage fewer, its buffers instead
    using FT = CfgX32T<BN, 3, false, true, 7>;        // fp16 planes: two thirds of a stage; the room goes to the X ring
    using FH = CfgX32T<BN, 3, true, true>;
    if (!gemm_forward_x32_ok(x, ld, N)) POZ_FAIL("tc3 fp32-input forward: operand not eligible");
    if (ep.head_w && (N > BN || ep.head_n < 1 || ep.head_n > HEAD_MAX_N || !ep.head_out))
        POZ_FAIL("tc3 fp32-input forward: fused head not available for this shape");
    static bool configured_dev[64] = {};              // function attributes are per device
    int dev = 0;
    POZ_CUDA(cudaGetDevice(&dev));
    bool& configured = configured_dev[dev & 63];
    if (!configured) {
        POZ_CUDA(cudaFuncSetAttribute(tc3_fwd_x32t_kernel<BN, 3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, CT::SMEM_BYTES));
        POZ_CUDA(cudaFuncSetAttribute(tc3_fwd_x32t_kernel<BN, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, CH::SMEM_BYTES));
        POZ_CUDA(cudaFuncSetAttribute(tc3_fwd_x32t_kernel<BN, 3, false, true, 7>, cudaFuncAttributeMaxDynamicSharedMemorySize, FT::SMEM_BYTES));
        POZ_CUDA(cudaFuncSetAttribute(tc3_fwd_x32t_kernel<BN, 3, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, FH::SMEM_BYTES));
        configured = true;
    }
    CUtensorMap mx, mb;
    POZ_TRY(tc::make_map_f32(&mx, x, M, K, ld, BM));     // boxes of 32 floats x 128 rows, 128-byte swizzle
    POZ_TRY(make_map3(&mb, w, 64, 64));
    const long long tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    const int grid = (int)(tiles < sms ? tiles : sms);
    if (w.planes == 2) {
        if (ep.head_w)
            tc3_fwd_x32t_kernel<BN, 3, true, true><<<grid, X32T_THREADS, FH::SMEM_BYTES, stream>>>(mx, mb, M, N, (long long)K, ep);
        else
            tc3_fwd_x32t_kernel<BN, 3, false, true, 7><<<grid, x32t_threads(2), FT::SMEM_BYTES, stream>>>(mx, mb, M, N, (long long)K, ep);
    } else if (ep.head_w)
        tc3_fwd_x32t_kernel<BN, 2, true><<<grid, X32T_THREADS, CH::SMEM_BYTES, stream>>>(mx, mb, M, N, (long long)K, ep);
    else
        tc3_fwd_x32t_kernel<BN, 3, false><<<grid, X32T_THREADS, CT::SMEM_BYTES, stream>>>(mx, mb, M, N, (long long)K, ep);
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

void x32t_probe_dump() {
    unsigned long long h[16] = {0};
    if (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpyFromSymbol(h, x32t_probe, sizeof(h)) != cudaSuccess || !h[5]) return;
    const double n = (double)h[5];
    fprintf(stderr, "[poz x32t probe] converter warp 10 of block 0, %llu half-blocks, cycles per half-block: wait for fp32 data %.0f, "
                    "loads + conversion %.0f, wait for the tensor-memory slot %.0f, tcgen05.st + wait %.0f, fences + arrivals %.0f; "
                    "whole loop %.0f\n", h[5], h[0] / n, h[1] / n, h[2] / n, h[3] / n, h[4] / n, h[6] / n);
    if (h[12]) {
        const double m = (double)h[12];
        fprintf(stderr, "[poz x32t probe] MMA issuer of block 0, %llu k-blocks, cycles per k-block: wait for the accumulators (tile boundary) %.0f, "
                        "for W's planes %.0f, for X's planes %.0f, issuing %.0f\n", h[12], h[8] / m, h[9] / m, h[10] / m, h[11] / m);
    }
    unsigned long long z[16] = {0};
    cudaMemcpyToSymbol(x32t_probe, z, sizeof(z));
}

int gemm_backward_data(const PlaneMat& dz, const PlaneMat& w, long long M, int K_in, int N_out, const Epi3& ep, int sms,
                       cudaStream_t s) {
    return launch_bn<MAJOR_K, MAJOR_K>(dz, w, M, K_in, N_out, 1, ep, sms, s);
}
int gemm_backward_weight(const PlaneMat& x, const PlaneMat& dz, long long batch, int K_in, int N_out, int splits,
                         const Epi3& ep, int sms, cudaStream_t s) {
    return launch_bn<MAJOR_MN, MAJOR_MN>(x, dz, K_in, N_out, batch, splits, ep, sms, s);
}

}  // namespace tc3

// ---- small-batch cluster split-K GEMM of the narrow train step (poz_tcs_kernel.cuh) ---------------------------
namespace tcs {

int pick_splits(long long K, long long tiles, int sms) {
    const long long nkb = (K + BK - 1) / BK;
    int s = 1;
    while (s < MAX_SPLITS && (nkb + s - 1) / s > MAX_KB) s *= 2;
    if ((nkb + s - 1) / s > MAX_KB) return 0;                                  // deeper than a cluster can hold
    while (s < MAX_SPLITS && tiles * s * 2 <= sms && nkb >= 2 * s) s *= 2;     // grow while the grid stays within one wave
    return s;
}

bool shape_ok(long long M, int N, long long K) {
    return M > 0 && N > 0 && K > 0 && (M + BM - 1) / BM <= 65535 && pick_splits(K, 1, 1) > 0;
}

template <bool A_T, class Epi> static int launch(const Gemm* gs, const Epi* epis, int n, cudaStream_t stream) {
    auto kern = tcs_gemm_kernel<A_T, Epi>;
    static bool configured = false;
    if (!configured) {
        POZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        configured = true;
    }
    if (n < 1 || n > MAX_GROUP) POZ_FAIL("tcs gemm: between 1 and MAX_GROUP problems per launch");
    const int splits = gs[0].splits;
    const long long nkb = (gs[0].K + BK - 1) / BK;
    if (splits < 1 || splits > MAX_SPLITS || (splits & (splits - 1)) || (nkb + splits - 1) / splits > MAX_KB)
        POZ_FAIL("tcs gemm: bad split count for this contraction length");
    GroupArgs ga{};
    GroupEpi<Epi> ge{};
    CUtensorMap maps[MAX_GROUP];
    ga.n = n;
    long long tiles = 0;
    for (int i = 0; i < n; ++i) {
        const Gemm& g = gs[i];
        if (g.K != gs[0].K || g.splits != splits) POZ_FAIL("tcs gemm: the problems of one launch share K and the split count");
        POZ_TRY(tc3::make_map3(&maps[i], g.b, 64, 64));
        Args& a = ga.a[i];
        a.a = g.a; a.lda = g.lda; a.row_idx = g.row_idx; a.a_cols = g.a_cols; a.ones_row = g.ones_row;
        a.M = g.M; a.N = g.N; a.K = g.K; a.a_dep = g.a_dep; a.b_dep = g.b_dep;
        a.k_per_split = ((nkb + splits - 1) / splits) * BK;
        ga.tiles_n[i] = (g.N + BN - 1) / BN;
        tiles += ((g.M + BM - 1) / BM) * ga.tiles_n[i];
        if (tiles > 0x7fffffffLL / 2) POZ_FAIL("tcs gemm: too many tiles");
        ga.tile_end[i] = (int)tiles;
        ge.e[i] = epis[i];
    }
    for (int i = n; i < MAX_GROUP; ++i) { maps[i] = maps[0]; ga.tile_end[i] = (int)tiles; ga.tiles_n[i] = 1; ga.a[i] = ga.a[0]; ge.e[i] = epis[0]; }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)tiles, 1, (unsigned)splits);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = (unsigned)splits;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = gs[0].pdl ? 2 : 1;
    POZ_CUDA(cudaLaunchKernelEx(&cfg, kern, ga, maps[0], maps[1], maps[2], maps[3], ge));
    count_launch();
    return 0;
}

// POZ_KERNEL_TRACE=1 (this translation unit's own copy of the trace symbols: they have internal linkage)
int trace_enable() {
    fast::KernelTraceRec* buf = nullptr;
    POZ_CUDA(cudaMemcpyFromSymbol(&buf, fast::d_ktrace, sizeof(buf)));
    if (!buf) {
        POZ_CUDA(cudaMalloc(&buf, sizeof(fast::KernelTraceRec) * (size_t)fast::KTRACE_CAP));
        POZ_CUDA(cudaMemcpyToSymbol(fast::d_ktrace, &buf, sizeof(buf)));
    }
    return 0;
}

void trace_dump() {
    fast::KernelTraceRec* buf = nullptr;
    if (cudaMemcpyFromSymbol(&buf, fast::d_ktrace, sizeof(buf)) != cudaSuccess || !buf) return;
    cudaDeviceSynchronize();
    unsigned int n = 0;
    cudaMemcpyFromSymbol(&n, fast::d_ktrace_n, sizeof(n));
    n = n < fast::KTRACE_CAP ? n : fast::KTRACE_CAP;
    if (!n) return;
    std::vector<fast::KernelTraceRec> h(n);
    cudaMemcpy(h.data(), buf, sizeof(fast::KernelTraceRec) * n, cudaMemcpyDeviceToHost);
    for (int tag : {200, 201}) {
        // the CTAs of the last launch with this tag: entries within 60 us of the latest entry
        unsigned long long latest = 0;
        for (auto& r : h) if (r.tag == tag && r.t[0] > latest) latest = r.t[0];
        if (!latest) continue;
        unsigned long long first = latest;
        for (auto& r : h) if (r.tag == tag && latest - r.t[0] < 60000ull && r.t[0] < first) first = r.t[0];
        double sum[6] = {0}; unsigned long long mx[6] = {0}; size_t cnt = 0; int m = 0, nn = 0, k = 0;
        for (auto& r : h)
            if (r.tag == tag && latest - r.t[0] < 60000ull) {
                ++cnt; m = r.m; nn = r.n; k = r.k;
                for (int q = 0; q < 6; ++q) { sum[q] += (double)(r.t[q] - first); if (r.t[q] - first > mx[q]) mx[q] = r.t[q] - first; }
            }
        unsigned long long last_exit = 0;
        for (auto& r : h) if (r.tag == tag && latest - r.t[0] < 60000ull && r.t[5] > last_exit) last_exit = r.t[5];
        fprintf(stderr, "[poz tcs trace] tag %d: absolute %%globaltimer mod 1e9 of the last launch: %llu .. %llu\n", tag, first % 1000000000ull,
                last_exit % 1000000000ull);
        fprintf(stderr, "[poz tcs trace] tag %d (M %d N %d K %d): %zu CTAs; mean (max) ns after the first CTA's entry: entry %.0f (%llu) "
                        "set-up %.0f (%llu) first half-block in tensor memory %.0f (%llu) accumulators complete %.0f (%llu) "
                        "partials exchanged %.0f (%llu) exit %.0f (%llu)\n",
                tag, m, nn, k, cnt, sum[0] / cnt, mx[0], sum[1] / cnt, mx[1], sum[2] / cnt, mx[2], sum[3] / cnt, mx[3],
                sum[4] / cnt, mx[4], sum[5] / cnt, mx[5]);
    }
    unsigned int zero = 0;
    cudaMemcpyToSymbol(fast::d_ktrace_n, &zero, sizeof(zero));
}

int forward(const Gemm& g, const FwdEpi& epi, cudaStream_t stream) { return launch<false>(&g, &epi, 1, stream); }
int backward_weight_store(const Gemm* g, const StoreEpi* epi, int n, cudaStream_t stream) { return launch<true>(g, epi, n, stream); }
int backward_weight_update(const Gemm* g, const UpdateEpi* epi, int n, cudaStream_t stream) { return launch<true>(g, epi, n, stream); }

}  // namespace tcs
}  // namespace poz

// fp32 matrices in, fp32 out: splits both operands into temporary plane stacks (allocated and freed
// here -- this is a test / benchmark building block, not the context path).
extern "C" int poz_tc3_gemm(int device, int variant, const float* a, const float* b, long long m, int n, long long k,
                            float* out, const float* bias, const float* aux, int act, int splits, void* stream) {
    using namespace poz;
    using namespace poz::tc3;
    if (!a || !b || !out) POZ_FAIL("poz_tc3_gemm: NULL operand");
    if (m <= 0 || n <= 0 || k <= 0) POZ_FAIL("poz_tc3_gemm: empty problem");
    if (variant < 0 || variant > 2) POZ_FAIL("poz_tc3_gemm: unknown variant");
    if (variant == 2 && (splits < 1 || splits > 64)) POZ_FAIL("poz_tc3_gemm: splits must be in [1,64]");
    POZ_CUDA(cudaSetDevice(device));
    int sms = 148;
    POZ_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // global shapes of the two operands
    long long a_rows, b_rows; int a_cols, b_cols;
    if (variant == 0) { a_rows = m; a_cols = (int)k; b_rows = k; b_cols = n; }          // a[m][k], b[k][n]
    else if (variant == 1) { a_rows = m; a_cols = (int)k; b_rows = n; b_cols = (int)k; } // a[m][k], b[n][k]
    else { a_rows = k; a_cols = (int)m; b_rows = k; b_cols = n; }                        // a[k][m], b[k][n]
    const long long lda = (a_cols + 7) & ~7, ldb = (b_cols + 7) & ~7;
    __nv_bfloat16 *pa = nullptr, *pb = nullptr;
    POZ_CUDA(cudaMalloc(&pa, (size_t)3 * a_rows * lda * 2));
    if (cudaMalloc(&pb, (size_t)3 * b_rows * ldb * 2) != cudaSuccess) { cudaFree(pa); POZ_FAIL("poz_tc3_gemm: cudaMalloc failed"); }
    int rc = split_planes(a, nullptr, a_rows, a_cols, a_cols, pa, lda, a_rows * lda, s);
    if (!rc) rc = split_planes(b, nullptr, b_rows, b_cols, b_cols, pb, ldb, b_rows * ldb, s);
    PlaneMat A{pa, a_rows, a_cols, lda, a_rows * lda}, B{pb, b_rows, b_cols, ldb, b_rows * ldb};
    Epi3 ep{};
    ep.out_f32 = out; ep.ld_f32 = n; ep.slab = m * (long long)n; ep.bias = bias; ep.aux = aux; ep.ld_aux = n; ep.act = act;
    if (!rc) {
        if (variant == 0) { ep.mode = EPI_BIAS_ACT; rc = gemm_forward(A, B, m, n, (int)k, ep, sms, s); }
        else if (variant == 1) { ep.mode = aux ? EPI_ACTGRAD : EPI_PLAIN; rc = gemm_backward_data(A, B, m, n, (int)k, ep, sms, s); }
        else { ep.mode = EPI_PLAIN; rc = gemm_backward_weight(A, B, k, (int)m, n, splits, ep, sms, s); }
    }
    cudaStreamSynchronize(s);
    cudaFree(pa);
    cudaFree(pb);
    return rc;
}
