// tcgen05 / TMEM / TMA GEMM for the wide (bf16) path, hand-written for sm_100a.
//
//   C[M][N] = op(A) * op(B),  bf16 operands, fp32 accumulation in tensor memory.
//
// One persistent CTA per SM works on 256x256 output tiles: two M=128 tcgen05.mma accumulators
// (2 x 256 TMEM columns = the SM's whole tensor memory) that share every B tile, so a CTA moves
// (256+256) x 64 x 2 B = 64 KB per k-block for 8.4 MFLOP (128 FLOP per L2 byte).  Warp roles:
//   warps 0-7 : epilogue  (warp w reads TMEM lanes 32*(w%4).., accumulator half w/4)
//   warp  8   : TMA producer (one elected lane)
//   warp  9   : TMEM allocator + MMA issuer (one elected lane)
// Operands are staged by TMA into 128-byte-swizzled shared memory (3 stages x 64 KB); both operand
// majors are supported so that no tensor is ever transposed in memory:
//   forward   Y  = X  * W      A = X  [M][K]   k-contiguous,   B = W  [K][N]  n-contiguous
//   backward  dA = dZ * W^T    A = dZ [M][K']  k-contiguous,   B = W  [N'][K'] k-contiguous
//   backward  dW = X^T * dZ    A = X  [K''][M''] m-contiguous, B = dZ [K''][N] n-contiguous
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include "poz_common.cuh"

namespace poz {
namespace tc {

constexpr int BM = 256, BN = 256, BK = 64;
constexpr int UMMA_M = 128, UMMA_N = 256, UMMA_K = 16;
constexpr int STAGES = 3;
constexpr int A_STAGE_BYTES = BM * BK * 2;                  // 32 KB
constexpr int B_STAGE_BYTES = BN * BK * 2;                  // 32 KB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;  // 64 KB
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = (NUM_EPI_WARPS + 2) * 32;       // 320
constexpr int TMEM_COLS = 512;

enum Major { MAJOR_K = 0, MAJOR_MN = 1 };
enum EpiKind { EPI_BIAS_ACT_BF16 = 0, EPI_ACTGRAD_BF16 = 1, EPI_F32 = 2 };

struct EpiParams {
    void* out;                 // bf16 [M][ldo] or float [M][ldo] (+ split * slab)
    long long ldo;
    long long slab;            // EPI_F32: elements between split-K slabs
    const float* bias;         // EPI_BIAS_ACT_BF16: [N] or null
    const __nv_bfloat16* aux;  // EPI_ACTGRAD_BF16: recorded activation of the previous unit [M][ldo]
    int act;
    // dropout points (neural_net_model.py:393-395) of this unit (forward: on z, then on a) or of the previous unit
    // (backward-x: the masks are regenerated), both optional; diagnostics outputs (:491-493), all nullable
    DropPoint dz, da;
    const DevState* st;
    __nv_bfloat16* z_rec;      // forward: the recorded (post-dropout) linear output
    __nv_bfloat16* da_rec;     // backward-x: gradient of the previous unit's recorded activation output
    __nv_bfloat16* dz_rec;     // backward-x: gradient of its recorded linear output
};

// ---- PTX wrappers ---------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, P1;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug must surface as a trapped launch, never as a hung GPU.
__device__ __noinline__ static void mbar_timeout() {
    printf("poz tcgen05 kernel: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
    __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (uint32_t spins = 0; !mbar_try_wait(bar, parity); ++spins)
        if (spins > 50u * 1000u * 1000u) mbar_timeout();
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
        "elect.sync rx|px, %1;\n\t"
        "selp.b32 %0, 1, 0, px;\n\t}"
        : "=r"(pred)
        : "r"(0xffffffffu));
    return pred != 0;
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "r"(cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols));
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor (sm_100 format): start address, leading / stride byte offsets
// (all >> 4), version 1, 128-byte swizzle.
//   K-major tile  [rows][64 bf16]: rows 128 B apart, 8-row swizzle atoms 1024 B apart (SBO);
//                 the next UMMA_K slice is +32 B inside the atom.
//   MN-major tile [64 k][64 mn] boxes: k rows 128 B apart, 8-k atoms 1024 B apart (SBO), the next
//                 64 mn elements one box (8192 B) further (LBO); the next UMMA_K slice is +2048 B.
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
    d |= (uint64_t)2 << 61;   // SWIZZLE_128B
    return d;
}

template <int A_MAJOR, int B_MAJOR> __device__ __forceinline__ constexpr uint32_t make_idesc() {
    return (1u << 4)                        // accumulator format F32
           | (1u << 7) | (1u << 10)         // A, B format BF16
           | ((uint32_t)A_MAJOR << 15) | ((uint32_t)B_MAJOR << 16)
           | ((uint32_t)(UMMA_N >> 3) << 17) | ((uint32_t)(UMMA_M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// ---- the kernel -------------------------------------------------------------------------------------
// tiles_m x tiles_n output tiles x `splits` K ranges; tile index t -> n fastest (CTAs running together
// share the A row block through L2), then m, then split.
template <int A_MAJOR, int B_MAJOR, int EPI>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               long long M, int N, long long K, int splits, long long k_per_split, EpiParams ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;
    uint64_t* tmem_empty_bar = tmem_full_bar + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const long long tiles_m = (M + BM - 1) / BM;
    const int tiles_n = (N + BN - 1) / BN;
    const long long num_tiles = tiles_m * tiles_n * splits;

    if (warp == 8 && lane == 0) {
        tma_prefetch_desc(&map_a);
        tma_prefetch_desc(&map_b);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(tmem_full_bar, 1);
        mbar_init(tmem_empty_bar, NUM_EPI_WARPS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) tmem_alloc(tmem_slot, TMEM_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        // ================= TMA producer =================
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int nb = (int)(t % tiles_n);
                const long long mb = (t / tiles_n) % tiles_m;
                const int sp = (int)(t / (tiles_n * tiles_m));
                const long long k0 = (long long)sp * k_per_split;
                const long long k1 = (k0 + k_per_split < K) ? k0 + k_per_split : K;
                const int nkb = (int)((k1 - k0 + BK - 1) / BK);
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    uint8_t* sb = sa + A_STAGE_BYTES;
                    mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
                    const int kc = (int)(k0 + (long long)kb * BK);
                    if (A_MAJOR == MAJOR_K) {
                        tma_load_2d(&map_a, &full_bar[stage], sa, kc, (int)(mb * BM));               // box {64 k, 256 m}
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)                                           // 4 boxes {64 m, 64 k}
                            tma_load_2d(&map_a, &full_bar[stage], sa + j * 8192, (int)(mb * BM) + 64 * j, kc);
                    }
                    if (B_MAJOR == MAJOR_K) {
                        tma_load_2d(&map_b, &full_bar[stage], sb, kc, nb * BN);                      // box {64 k, 256 n}
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)                                           // 4 boxes {64 n, 64 k}
                            tma_load_2d(&map_b, &full_bar[stage], sb + j * 8192, nb * BN + 64 * j, kc);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 9) {
        // ================= MMA issuer =================
        if (elect_one()) {
            constexpr uint32_t idesc = make_idesc<A_MAJOR, B_MAJOR>();
            int stage = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int sp = (int)(t / (tiles_n * tiles_m));
                const long long k0 = (long long)sp * k_per_split;
                const long long k1 = (k0 + k_per_split < K) ? k0 + k_per_split : K;
                const int nkb = (int)((k1 - k0 + BK - 1) / BK);
                mbar_wait(tmem_empty_bar, acc_phase ^ 1);      // epilogue has drained the accumulators
                tc_fence_after();
                for (int kb = 0; kb < nkb; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * STAGE_BYTES);
                    const uint32_t sb = sa + A_STAGE_BYTES;
#pragma unroll
                    for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                        const uint64_t bdesc = (B_MAJOR == MAJOR_K) ? make_desc(sb + kk * 32, 16, 1024)
                                                                    : make_desc(sb + kk * 2048, 8192, 1024);
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const uint64_t adesc = (A_MAJOR == MAJOR_K)
                                                       ? make_desc(sa + h * 16384 + kk * 32, 16, 1024)
                                                       : make_desc(sa + h * 16384 + kk * 2048, 8192, 1024);
                            umma_bf16(tmem_base + h * UMMA_N, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
                        }
                    }
                    umma_commit(&empty_bar[stage]);            // frees the smem slot when these MMAs retire
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit(tmem_full_bar);                    // accumulators complete -> epilogue
                acc_phase ^= 1;
            }
        }
    } else {
        // ================= epilogue (warps 0..7) =================
        const int quarter = warp & 3, half = warp >> 2;
        uint32_t acc_phase = 0;
        for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int nb = (int)(t % tiles_n);
            const long long mb = (t / tiles_n) % tiles_m;
            const int sp = (int)(t / (tiles_n * tiles_m));
            mbar_wait(tmem_full_bar, acc_phase);
            tc_fence_after();
            const long long m = mb * BM + half * UMMA_M + quarter * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(half * UMMA_N);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t v[32];
                tmem_ld32(taddr + c, v);
                tmem_ld_wait();
                const int n = nb * BN + c;
                if (m < M && n < N) {
                    // every register array below is indexed with compile-time constants only (fully
                    // unrolled, predicated tails) so that it stays in registers
                    const bool full = (n + 32 <= N);
                    if (EPI == EPI_F32) {
                        float* out = static_cast<float*>(ep.out) + (long long)sp * ep.slab + m * ep.ldo + n;
                        if (full && ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
#pragma unroll
                            for (int j = 0; j < 32; j += 4)
                                *reinterpret_cast<uint4*>(out + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (n + j < N) out[j] = __uint_as_float(v[j]);
                        }
                    } else {
                        float f[32];
#pragma unroll
                        for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
                        __nv_bfloat16* out = static_cast<__nv_bfloat16*>(ep.out) + m * ep.ldo + n;
                        if (EPI == EPI_BIAS_ACT_BF16) {
                            if (ep.bias) {
                                const float* bp = ep.bias + n;
                                if (full && ((reinterpret_cast<uintptr_t>(bp) & 15) == 0)) {
#pragma unroll
                                    for (int j = 0; j < 32; j += 4) {
                                        const float4 b4 = __ldg(reinterpret_cast<const float4*>(bp + j));
                                        f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
                                    }
                                } else {
#pragma unroll
                                    for (int j = 0; j < 32; ++j)
                                        if (n + j < N) f[j] += __ldg(bp + j);
                                }
                            }
                            if (ep.dz.enabled) {
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (n + j < N) f[j] = drop_keep(ep.dz, ep.st, m, n + j, (int)ep.ldo) ? f[j] * (float)ep.dz.scale : 0.f;
                            }
                            if (ep.z_rec) {
                                __nv_bfloat16* zr = ep.z_rec + m * ep.ldo + n;
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (n + j < N) zr[j] = __float2bfloat16_rn(f[j]);
                            }
                            if (ep.act == ACT_RELU) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) f[j] = f[j] < 0.f ? 0.f : f[j];   // (keeps a NaN a NaN)
                            } else if (ep.act != ACT_NONE) {
#pragma unroll
                                for (int j = 0; j < 32; ++j) f[j] = act_fwd<float>(ep.act, f[j]);
                            }
                            if (ep.da.enabled) {
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (n + j < N) f[j] = drop_keep(ep.da, ep.st, m, n + j, (int)ep.ldo) ? f[j] * (float)ep.da.scale : 0.f;
                            }
                        } else if (ep.da.enabled || ep.dz.enabled || ep.da_rec || ep.dz_rec) {
                            // EPI_ACTGRAD_BF16, general form (test_torch_backward.py:168-170 with the dropout of :393-395):
                            // f = gradient of the previous unit's RECORDED (dropped, scaled) activation output
                            const __nv_bfloat16* aux = ep.aux + m * ep.ldo + n;
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                if (n + j >= N) continue;
                                float g = f[j], a = __bfloat162float(aux[j]);
                                if (ep.da_rec) ep.da_rec[m * ep.ldo + n + j] = __float2bfloat16_rn(g);
                                if (ep.da.enabled) {
                                    g = drop_keep(ep.da, ep.st, m, n + j, (int)ep.ldo) ? g * (float)ep.da.scale : 0.f;
                                    a *= (float)ep.da.keep_prob;
                                }
                                g *= act_bwd_from_out<float>(ep.act, a);
                                if (ep.dz_rec) ep.dz_rec[m * ep.ldo + n + j] = __float2bfloat16_rn(g);
                                if (ep.dz.enabled) g = drop_keep(ep.dz, ep.st, m, n + j, (int)ep.ldo) ? g * (float)ep.dz.scale : 0.f;
                                f[j] = g;
                            }
                        } else {   // EPI_ACTGRAD_BF16: gradient through the previous unit's activation
                            const __nv_bfloat16* aux = ep.aux + m * ep.ldo + n;
                            if (full && ((reinterpret_cast<uintptr_t>(aux) & 15) == 0)) {
#pragma unroll
                                for (int j = 0; j < 32; j += 8) {
                                    uint4 q = *reinterpret_cast<const uint4*>(aux + j);
                                    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
                                    for (int e = 0; e < 4; ++e) {
                                        float2 a2 = __bfloat1622float2(h2[e]);
                                        f[j + 2 * e] *= act_bwd_from_out<float>(ep.act, a2.x);
                                        f[j + 2 * e + 1] *= act_bwd_from_out<float>(ep.act, a2.y);
                                    }
                                }
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j)
                                    if (n + j < N) f[j] *= act_bwd_from_out<float>(ep.act, __bfloat162float(aux[j]));
                            }
                        }
                        if (full && ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
#pragma unroll
                            for (int j = 0; j < 32; j += 8)
                                *reinterpret_cast<uint4*>(out + j) =
                                    make_uint4(pack_bf16(f[j], f[j + 1]), pack_bf16(f[j + 2], f[j + 3]),
                                               pack_bf16(f[j + 4], f[j + 5]), pack_bf16(f[j + 6], f[j + 7]));
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (n + j < N) out[j] = __float2bfloat16_rn(f[j]);
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty_bar);        // this warp no longer needs the accumulators
            acc_phase ^= 1;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 9) {
        tc_fence_after();
        tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace tc
}  // namespace poz
