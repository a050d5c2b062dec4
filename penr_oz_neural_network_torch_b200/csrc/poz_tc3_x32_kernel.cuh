// Forward split-bf16 tcgen05 GEMM whose A operand is the fp32 matrix itself: out = epilogue(X[M][K] * W[K][N]).
//
// The plain tc3 path first converts X into three bf16 planes in HBM (read 4 B, write 6 B per element) and the
// GEMM reads those 6 B back: 16 B of HBM traffic per input element where the input itself is 4 B -- measured
// at 262,144 rows of the 784-wide stack the split kernel alone is 0.36 ms of a 1.17 ms pass, at 88 % of HBM
// bandwidth.  Here TMA brings fp32 half-blocks (128 rows x 32 k, 128-byte swizzle) into a five-slot ring,
// eight converter warps (two threads per row) split them into the hi / mid / lo planes directly in the
// K-major 128-byte-swizzled shared-memory layout the MMA descriptors expect, make them visible to the async
// proxy and hand the stage to the MMA warp; W's planes still arrive by TMA.  The planes of X never exist in
// global memory.
//   warps 0-7 : epilogue          warp 8 : TMA producer (W planes, fp32 X half-blocks)
//   warp  9   : TMEM + MMA        warps 10-17 : converters
#pragma once
#include "poz_tc3_kernel.cuh"

namespace poz {
namespace tc3 {

constexpr int X32_CONV_WARPS = 8;                   // two threads per tile row
constexpr int X32_THREADS = (NUM_EPI_WARPS + 2 + X32_CONV_WARPS) * 32;
constexpr int X32_RAW_BYTES = BM * 32 * 4;          // one half-block: 128 rows x 32 floats

constexpr int X32_RAW_SLOTS = 5;                    // fp32 half-blocks in flight (2.5 k-blocks of look-ahead)

// Shared memory: ONE set of X planes (3 x 16 KB; its two 32-k halves are handed over separately, so the converters
// refill half h of the next k-block while the MMAs still read the other half), two stages of W planes, and the ring
// of fp32 half-blocks -- 48 + 96 + 80 KB for BN = 128.
template <int BN> struct CfgX32 {
    static constexpr int STAGES = 2;                // W-plane stages
    static constexpr int A_PLANE = BM * BK * 2;
    static constexpr int B_PLANE = BN * BK * 2;
    static constexpr int B_OFF = PLANES * A_PLANE;
    static constexpr int B_STAGE = PLANES * B_PLANE;
    static constexpr int RAW_OFF = B_OFF + STAGES * B_STAGE;
    static constexpr int BAR_OFF = RAW_OFF + X32_RAW_SLOTS * X32_RAW_BYTES;
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
    uint8_t* raw = smem + C::RAW_OFF;                                  // [X32_RAW_SLOTS][128 rows][32 floats], swizzled
    uint64_t* b_full = reinterpret_cast<uint64_t*>(smem + C::BAR_OFF); // [STAGES] W planes landed
    uint64_t* b_empty = b_full + STAGES;                               // [STAGES] MMAs of the stage retired
    uint64_t* a_full = b_empty + STAGES;                               // [2] half h of the X planes written
    uint64_t* a_empty = a_full + 2;                                    // [2] MMAs that read half h retired
    uint64_t* raw_full = a_empty + 2;                                  // [X32_RAW_SLOTS]
    uint64_t* raw_empty = raw_full + X32_RAW_SLOTS;                    // [X32_RAW_SLOTS]
    uint64_t* tmem_full_bar = raw_empty + X32_RAW_SLOTS;               // [2]
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
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&b_full[s], 1); tc::mbar_init(&b_empty[s], 1); }
        for (int h = 0; h < 2; ++h) { tc::mbar_init(&a_full[h], X32_CONV_WARPS); tc::mbar_init(&a_empty[h], 1); }
        for (int q = 0; q < X32_RAW_SLOTS; ++q) { tc::mbar_init(&raw_full[q], 1); tc::mbar_init(&raw_empty[q], X32_CONV_WARPS); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&tmem_full_bar[b], 1); tc::mbar_init(&tmem_empty_bar[b], NUM_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        // ================= TMA producer: fp32 half-blocks of X into the raw ring, W planes into their stage =================
        if (tc::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            uint32_t slot = 0, rphase = 0;                 // raw ring position
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int nb = (int)(t % tiles_n);
                const long long mb = t / tiles_n;
                for (int kb = 0; kb < nkb; ++kb) {
                    const int kc = kb * BK;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        tc::mbar_wait(&raw_empty[slot], rphase ^ 1);
                        tc::mbar_expect_tx(&raw_full[slot], X32_RAW_BYTES);
                        tma_load_2d_f32(&map_x, &raw_full[slot], raw + slot * X32_RAW_BYTES, kc + 32 * h, (int)(mb * BM));
                        if (++slot == X32_RAW_SLOTS) { slot = 0; rphase ^= 1; }
                    }
                    tc::mbar_wait(&b_empty[stage], phase ^ 1);
                    uint8_t* sb = smem + C::B_OFF + stage * C::B_STAGE;
                    tc::mbar_expect_tx(&b_full[stage], C::B_STAGE);
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
            uint32_t kbc = 0;                              // k-blocks done: parity of the per-half barriers
            uint32_t it = 0;
            const uint32_t sa = tc::smem_u32(smem);
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
                const uint32_t ab = it & 1, aphase = (it >> 1) & 1;
                tc::mbar_wait(&tmem_empty_bar[ab], aphase ^ 1);
                tc::tc_fence_after();
                const uint32_t acc = tmem_base + ab * (2 * BN);      // main (hi*hi) accumulator
                const uint32_t acs = acc + BN;                       // small-terms accumulator
                for (int kb = 0; kb < nkb; ++kb, ++kbc) {
                    tc::mbar_wait(&b_full[stage], phase);
                    const uint32_t sb = tc::smem_u32(smem + C::B_OFF + stage * C::B_STAGE);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        tc::mbar_wait(&a_full[h], kbc & 1);
                        tc::tc_fence_after();
#pragma unroll
                        for (int kk = 2 * h; kk < 2 * h + 2; ++kk) {
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
                        tc::umma_commit(&a_empty[h]);       // this half of the X planes may be refilled
                    }
                    tc::umma_commit(&b_empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                tc::umma_commit(&tmem_full_bar[ab]);
            }
        }
    } else if (warp >= NUM_EPI_WARPS + 2) {
        // ================= converters: two threads per tile row, 16 of a half-block's 32 k-values each =================
        // a warp takes 16 rows; lanes 0-15 the first 16 k-values of each, lanes 16-31 the second: every group of eight
        // lanes is eight consecutive rows at one chunk position, so the swizzle spreads its 16-byte loads AND stores
        // over all banks (with (row, part) = (lane / 2, lane % 2) the plane stores were two-way conflicted)
        const int r = (warp - (NUM_EPI_WARPS + 2)) * 16 + (lane & 15), part = lane >> 4;
        const uint32_t sw = (uint32_t)(r & 7);                          // the 128-byte swizzle XORs the 16-byte chunk index with row % 8
        uint32_t slot = 0, rphase = 0, kbc = 0;
        uint8_t* dst = smem + r * 128;
        for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            for (int kb = 0; kb < nkb; ++kb, ++kbc) {
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    tc::mbar_wait(&raw_full[slot], rphase);
                    const uint8_t* src = raw + slot * X32_RAW_BYTES + r * 128;
                    float x[16];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const float4 v = *reinterpret_cast<const float4*>(src + (((uint32_t)(4 * part + c) ^ sw) << 4));
                        x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
                    }
                    uint32_t ph[8], pm[8], pl[8];
#pragma unroll
                    for (int j = 0; j < 16; j += 2) split3_pair(x[j], x[j + 1], ph[j / 2], pm[j / 2], pl[j / 2]);
                    // the MMAs that read this half of the planes (previous k-block) must have retired
                    tc::mbar_wait(&a_empty[h], (kbc & 1) ^ 1);
                    // this thread's 16 k-values = 32 bytes = chunks 4h + 2 part, + 1 of the row in each plane's [128][64 bf16] tile
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const uint32_t off = (((uint32_t)(4 * h + 2 * part + q)) ^ sw) << 4;
                        *reinterpret_cast<uint4*>(dst + off) = make_uint4(ph[4 * q], ph[4 * q + 1], ph[4 * q + 2], ph[4 * q + 3]);
                        *reinterpret_cast<uint4*>(dst + C::A_PLANE + off) = make_uint4(pm[4 * q], pm[4 * q + 1], pm[4 * q + 2], pm[4 * q + 3]);
                        *reinterpret_cast<uint4*>(dst + 2 * C::A_PLANE + off) = make_uint4(pl[4 * q], pl[4 * q + 1], pl[4 * q + 2], pl[4 * q + 3]);
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to tcgen05.mma
                    __syncwarp();
                    if (lane == 0) {
                        tc::mbar_arrive(&raw_empty[slot]);  // the half-block has been consumed by this warp
                        tc::mbar_arrive(&a_full[h]);
                    }
                    if (++slot == X32_RAW_SLOTS) { slot = 0; rphase ^= 1; }
                }
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
