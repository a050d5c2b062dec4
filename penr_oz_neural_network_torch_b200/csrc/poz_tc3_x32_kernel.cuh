// Forward split-bf16 tcgen05 GEMM whose A operand is the fp32 matrix itself: out = epilogue(X[M][K] * W[K][N]).
//
// The plain tc3 path first converts X into three bf16 planes in HBM (read 4 B, write 6 B per element) and the
// GEMM reads those 6 B back: 16 B of HBM traffic per input element where the input itself is 4 B -- measured
// at 262,144 rows of the 784-wide stack the split kernel alone is 0.36 ms of a 1.17 ms pass, at 88 % of HBM
// bandwidth.  Here TMA brings fp32 half-blocks (128 rows x 32 k, 128-byte swizzle) into a two-slot ring,
// four converter warps (one thread per row) split them into the hi / mid / lo planes directly in the
// K-major 128-byte-swizzled shared-memory layout the MMA descriptors expect, make them visible to the async
// proxy and hand the stage to the MMA warp; W's planes still arrive by TMA.  The planes of X never exist in
// global memory.
//   warps 0-7 : epilogue          warp 8 : TMA producer (W planes, fp32 X half-blocks)
//   warp  9   : TMEM + MMA        warps 10-13 : converters
#pragma once
#include "poz_tc3_kernel.cuh"

namespace poz {
namespace tc3 {

constexpr int X32_CONV_WARPS = 4;
constexpr int X32_THREADS = (NUM_EPI_WARPS + 2 + X32_CONV_WARPS) * 32;
constexpr int X32_RAW_BYTES = BM * 32 * 4;          // one half-block: 128 rows x 32 floats

template <int BN> struct CfgX32 {
    static constexpr int STAGES = 2;
    static constexpr int A_PLANE = BM * BK * 2;
    static constexpr int B_PLANE = BN * BK * 2;
    static constexpr int STAGE_BYTES = PLANES * (A_PLANE + B_PLANE);
    static constexpr int RAW_OFF = STAGES * STAGE_BYTES;
    static constexpr int BAR_OFF = RAW_OFF + 2 * X32_RAW_BYTES;
    static constexpr int SMEM_BYTES = BAR_OFF + 1024 + 256;
    static constexpr int TMEM_COLS = 4 * BN;
};

__device__ __forceinline__ void tma_load_2d_f32(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(tc::smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}

template <int BN>
__global__ void __launch_bounds__(X32_THREADS, 1)
tc3_fwd_x32_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b,
                   long long M, int N, long long K, Epi3 ep) {
    using C = CfgX32<BN>;
    constexpr int STAGES = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* raw = smem + C::RAW_OFF;                                  // [2][128 rows][32 floats], swizzled
    uint64_t* b_full = reinterpret_cast<uint64_t*>(smem + C::BAR_OFF); // [STAGES] W planes landed
    uint64_t* a_full = b_full + STAGES;                                // [STAGES] X planes written
    uint64_t* empty_bar = a_full + STAGES;                             // [STAGES] MMAs of the stage retired
    uint64_t* raw_full = empty_bar + STAGES;                           // [2]
    uint64_t* raw_empty = raw_full + 2;                                // [2]
    uint64_t* tmem_full_bar = raw_empty + 2;                           // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;                      // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long tiles_m = (M + BM - 1) / BM;
    const int tiles_n = (N + BN - 1) / BN;
    const long long num_tiles = tiles_m * tiles_n;
    const int nkb = (int)((K + BK - 1) / BK);

    if (warp == 8 && lane == 0) {
        tc::tma_prefetch_desc(&map_x);
        tc::tma_prefetch_desc(&map_b);
        for (int s = 0; s < STAGES; ++s) {
            tc::mbar_init(&b_full[s], 1);
            tc::mbar_init(&a_full[s], X32_CONV_WARPS);
            tc::mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            tc::mbar_init(&raw_full[b], 1);
            tc::mbar_init(&raw_empty[b], X32_CONV_WARPS);
            tc::mbar_init(&tmem_full_bar[b], 1);
            tc::mbar_init(&tmem_empty_bar[b], NUM_EPI_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        // ================= TMA producer: fp32 half-blocks of X into the raw ring, W planes into the stage =================
        if (tc::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            uint32_t rc = 0;                               // half-blocks issued so far (raw slot = rc & 1)
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int nb = (int)(t % tiles_n);
                const long long mb = t / tiles_n;
                for (int kb = 0; kb < nkb; ++kb) {
                    const int kc = kb * BK;
#pragma unroll
                    for (int h = 0; h < 2; ++h, ++rc) {
                        const uint32_t slot = rc & 1, rphase = (rc >> 1) & 1;
                        tc::mbar_wait(&raw_empty[slot], rphase ^ 1);
                        tc::mbar_expect_tx(&raw_full[slot], X32_RAW_BYTES);
                        tma_load_2d_f32(&map_x, &raw_full[slot], raw + slot * X32_RAW_BYTES, kc + 32 * h, (int)(mb * BM));
                    }
                    tc::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sb = smem + stage * C::STAGE_BYTES + PLANES * C::A_PLANE;
                    tc::mbar_expect_tx(&b_full[stage], PLANES * C::B_PLANE);
#pragma unroll
                    for (int p = 0; p < PLANES; ++p)
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_3d(&map_b, &b_full[stage], sb + p * C::B_PLANE + j * 8192, nb * BN + 64 * j, kc, p);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 9) {
        // ================= MMA issuer =================
        if (tc::elect_one()) {
            constexpr uint32_t idesc = make_idesc3<MAJOR_K, MAJOR_MN, BN>();
            int stage = 0;
            uint32_t phase = 0;
            uint32_t it = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
                const uint32_t ab = it & 1, aphase = (it >> 1) & 1;
                tc::mbar_wait(&tmem_empty_bar[ab], aphase ^ 1);
                tc::tc_fence_after();
                const uint32_t acc = tmem_base + ab * (2 * BN);      // main (hi*hi) accumulator
                const uint32_t acs = acc + BN;                       // small-terms accumulator
                for (int kb = 0; kb < nkb; ++kb) {
                    tc::mbar_wait(&b_full[stage], phase);
                    tc::mbar_wait(&a_full[stage], phase);
                    tc::tc_fence_after();
                    const uint32_t sa = tc::smem_u32(smem + stage * C::STAGE_BYTES);
                    const uint32_t sb = sa + PLANES * C::A_PLANE;
#pragma unroll
                    for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                        uint64_t ad[PLANES], bd[PLANES];
#pragma unroll
                        for (int p = 0; p < PLANES; ++p) {
                            ad[p] = tc::make_desc(sa + p * C::A_PLANE + kk * 32, 16, 1024);
                            bd[p] = tc::make_desc(sb + p * C::B_PLANE + kk * 2048, 8192, 1024);
                        }
                        const uint32_t first = (kb | kk) != 0 ? 1u : 0u;
                        tc::umma_bf16(acs, ad[0], bd[2], idesc, first);
                        tc::umma_bf16(acs, ad[2], bd[0], idesc, 1u);
                        tc::umma_bf16(acs, ad[1], bd[1], idesc, 1u);
                        tc::umma_bf16(acs, ad[0], bd[1], idesc, 1u);
                        tc::umma_bf16(acs, ad[1], bd[0], idesc, 1u);
                        tc::umma_bf16(acc, ad[0], bd[0], idesc, first);
                    }
                    tc::umma_commit(&empty_bar[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                tc::umma_commit(&tmem_full_bar[ab]);
            }
        }
    } else if (warp >= NUM_EPI_WARPS + 2) {
        // ================= converters: one thread per tile row =================
        const int r = (warp - (NUM_EPI_WARPS + 2)) * 32 + lane;       // 0 .. 127
        const uint32_t sw = (uint32_t)(r & 7);                          // the 128-byte swizzle XORs the 16-byte chunk index with row % 8
        int stage = 0;
        uint32_t phase = 0;
        uint32_t rc = 0;
        for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            for (int kb = 0; kb < nkb; ++kb) {
                tc::mbar_wait(&empty_bar[stage], phase ^ 1);            // the MMAs that read this stage's planes have retired
                uint8_t* sa = smem + stage * C::STAGE_BYTES;
#pragma unroll
                for (int h = 0; h < 2; ++h, ++rc) {
                    const uint32_t slot = rc & 1, rphase = (rc >> 1) & 1;
                    tc::mbar_wait(&raw_full[slot], rphase);
                    const uint8_t* src = raw + slot * X32_RAW_BYTES + r * 128;
                    float x[32];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        const float4 v = *reinterpret_cast<const float4*>(src + ((c ^ sw) << 4));
                        x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
                    }
                    uint32_t ph[16], pm[16], pl[16];
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        __nv_bfloat16 h0, m0, l0, h1, m1, l1;
                        split3(x[j], h0, m0, l0);
                        split3(x[j + 1], h1, m1, l1);
                        ph[j / 2] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
                        pm[j / 2] = (uint32_t)__bfloat16_as_ushort(m0) | ((uint32_t)__bfloat16_as_ushort(m1) << 16);
                        pl[j / 2] = (uint32_t)__bfloat16_as_ushort(l0) | ((uint32_t)__bfloat16_as_ushort(l1) << 16);
                    }
                    // 32 k-values = 64 bytes = chunks 4h .. 4h+3 of this row in each plane's [128][64 bf16] tile
                    uint8_t* dst = sa + r * 128;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t off = (((uint32_t)(4 * h + q)) ^ sw) << 4;
                        *reinterpret_cast<uint4*>(dst + off) = make_uint4(ph[4 * q], ph[4 * q + 1], ph[4 * q + 2], ph[4 * q + 3]);
                        *reinterpret_cast<uint4*>(dst + C::A_PLANE + off) = make_uint4(pm[4 * q], pm[4 * q + 1], pm[4 * q + 2], pm[4 * q + 3]);
                        *reinterpret_cast<uint4*>(dst + 2 * C::A_PLANE + off) = make_uint4(pl[4 * q], pl[4 * q + 1], pl[4 * q + 2], pl[4 * q + 3]);
                    }
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(&raw_empty[slot]);   // the half-block has been consumed by this warp
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to tcgen05.mma
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&a_full[stage]);
                if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else {
        // ================= epilogue =================
        const int quarter = warp & 3, chalf = warp >> 2;
        uint32_t it = 0;
        for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
            const int nb = (int)(t % tiles_n);
            const long long mb = t / tiles_n;
            const uint32_t ab = it & 1, aphase = (it >> 1) & 1;
            tc::mbar_wait(&tmem_full_bar[ab], aphase);
            tc::tc_fence_after();
            epilogue_tile<BN>(ep, M, N, mb, nb, 0, false, tmem_base, ab, quarter, chalf, lane);
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&tmem_empty_bar[ab]);
        }
    }

    tc::tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

}  // namespace tc3
}  // namespace poz
