// Host side of the tcgen05 GEMM: TMA tensor-map construction and launches (wide bf16 path).
#include "../../include/pozmlp.h"
#include "poz_tc.h"
#include "poz_tc_gemm.cuh"

#include <mutex>

namespace poz {
namespace tc {

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

// Row-major bf16 matrix [rows][cols] (pitch ld elements); box = box_cols x box_rows, 128B swizzle.
static int make_map(CUtensorMap* map, const void* ptr, long long rows, long long cols, long long ld,
                    int box_cols, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) POZ_FAIL("tcgen05 path: cuTensorMapEncodeTiled is not available from the driver");
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (ld * 2) % 16 != 0)
        POZ_FAIL("tcgen05 path: operand base and row pitch must be 16-byte aligned");
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) POZ_FAIL("tcgen05 path: cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return 0;
}

int make_map_f32(void* map_out, const void* ptr, long long rows, long long cols, long long ld, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) POZ_FAIL("TMA path: cuTensorMapEncodeTiled is not available from the driver");
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (ld * 4) % 16 != 0)
        POZ_FAIL("TMA path: operand base and row pitch must be 16-byte aligned");
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {32u, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(static_cast<CUtensorMap*>(map_out), CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), gdim,
                    gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) POZ_FAIL("TMA path: cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
    return 0;
}

template <int A_MAJOR, int B_MAJOR, int EPI>
static int launch(const CUtensorMap& ma, const CUtensorMap& mb, long long M, int N, long long K, int splits,
                  const EpiParams& ep, int sms, cudaStream_t stream) {
    static bool configured = false;
    auto kern = tc_gemm_kernel<A_MAJOR, B_MAJOR, EPI>;
    if (!configured) {
        POZ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        configured = true;
    }
    long long k_per = (K + splits - 1) / splits;
    k_per = ((k_per + BK - 1) / BK) * BK;
    const long long tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN) * splits;
    const int grid = (int)(tiles < sms ? tiles : sms);
    kern<<<grid, NUM_THREADS, SMEM_BYTES, stream>>>(ma, mb, M, N, K, splits, k_per, ep);
    count_launch();
    POZ_CUDA(cudaGetLastError());
    return 0;
}

static void apply_extras(EpiParams& ep, const Extras* ex) {
    if (!ex) return;
    ep.dz = ex->dz; ep.da = ex->da; ep.st = ex->st;
    ep.z_rec = ex->z_rec; ep.da_rec = ex->da_rec; ep.dz_rec = ex->dz_rec;
}

int gemm_forward(const __nv_bfloat16* x, const __nv_bfloat16* w, long long M, int N, int K, const float* bias,
                 int act, __nv_bfloat16* out, int sms, cudaStream_t s, const Extras* ex) {
    CUtensorMap ma, mb;
    POZ_TRY(make_map(&ma, x, M, K, K, 64, 256));     // A = X [M][K], k-contiguous: box {64 k, 256 m}
    POZ_TRY(make_map(&mb, w, K, N, N, 64, 64));      // B = W [K][N], n-contiguous: boxes {64 n, 64 k}
    EpiParams ep{};
    ep.out = out; ep.ldo = N; ep.bias = bias; ep.act = act;
    apply_extras(ep, ex);
    return launch<MAJOR_K, MAJOR_MN, EPI_BIAS_ACT_BF16>(ma, mb, M, N, K, 1, ep, sms, s);
}

int gemm_backward_data(const __nv_bfloat16* dz, const __nv_bfloat16* w, long long M, int K_in, int N_out,
                       const __nv_bfloat16* a_prev, int act_prev, __nv_bfloat16* dz_prev, int sms,
                       cudaStream_t s, const Extras* ex) {
    CUtensorMap ma, mb;
    POZ_TRY(make_map(&ma, dz, M, N_out, N_out, 64, 256));        // A = dZ [M][N_out], k-contiguous
    POZ_TRY(make_map(&mb, w, K_in, N_out, N_out, 64, 256));      // B = W [K_in][N_out]: rows = mma-N, k-contiguous
    EpiParams ep{};
    ep.out = dz_prev; ep.ldo = K_in; ep.aux = a_prev; ep.act = act_prev;
    apply_extras(ep, ex);
    return launch<MAJOR_K, MAJOR_K, EPI_ACTGRAD_BF16>(ma, mb, M, K_in, N_out, 1, ep, sms, s);
}

int gemm_backward_weight(const __nv_bfloat16* x, const __nv_bfloat16* dz, long long batch, int K_in, int N_out,
                         float* g, long long slab, int splits, int sms, cudaStream_t s) {
    CUtensorMap ma, mb;
    POZ_TRY(make_map(&ma, x, batch, K_in, K_in, 64, 64));        // A = X [batch][K_in], m-contiguous
    POZ_TRY(make_map(&mb, dz, batch, N_out, N_out, 64, 64));     // B = dZ [batch][N_out], n-contiguous
    EpiParams ep{};
    ep.out = g; ep.ldo = N_out; ep.slab = slab;
    return launch<MAJOR_MN, MAJOR_MN, EPI_F32>(ma, mb, K_in, N_out, batch, splits, ep, sms, s);
}

}  // namespace tc
}  // namespace poz

extern "C" int poz_tc_gemm(int device, int variant, const void* a, const void* b, long long m, int n, long long k,
                           void* out, const float* bias, const void* aux, int act, int splits, void* stream) {
    using namespace poz;
    if (!a || !b || !out) POZ_FAIL("poz_tc_gemm: NULL operand");
    if (m <= 0 || n <= 0 || k <= 0) POZ_FAIL("poz_tc_gemm: empty problem");
    if (n % 8 || k % 8 || (variant == 2 && m % 8)) POZ_FAIL("poz_tc_gemm: dimensions must be multiples of 8");
    POZ_CUDA(cudaSetDevice(device));
    int sms = 148;
    POZ_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const __nv_bfloat16* A = static_cast<const __nv_bfloat16*>(a);
    const __nv_bfloat16* B = static_cast<const __nv_bfloat16*>(b);
    switch (variant) {
        case 0: return tc::gemm_forward(A, B, m, n, (int)k, bias, act, static_cast<__nv_bfloat16*>(out), sms, s);
        case 1:
            if (!aux) POZ_FAIL("poz_tc_gemm: variant 1 needs aux");
            return tc::gemm_backward_data(A, B, m, n, (int)k, static_cast<const __nv_bfloat16*>(aux), act,
                                          static_cast<__nv_bfloat16*>(out), sms, s);
        case 2:
            if (splits < 1 || splits > 64) POZ_FAIL("poz_tc_gemm: splits must be in [1,64]");
            return tc::gemm_backward_weight(A, B, k, (int)m, n, static_cast<float*>(out), m * (long long)n, splits,
                                            sms, s);
        default: POZ_FAIL("poz_tc_gemm: unknown variant");
    }
}
