// The 3xTF32 GEMM of poz_tf32x3_gemm.cuh with its operand tiles brought in by TMA.
//
// Why: tile_delivery.cu shows the operand tiles of these GEMMs arriving 2-4 us
// sooner per launch as 32-wide TMA boxes than as per-thread 16-byte cp.async copies.  So: one elected
// thread issues one tensor copy per operand tile (two for a 64-column row-contiguous B tile, whose rows
// exceed the 128-byte swizzle span), completion is counted on an mbarrier per stage, and the 128-byte
// swizzle the copy applies makes the k-contiguous fragment reads of mma.sync conflict-free without
// padding (two-way conflicts remain on the row-contiguous arrangement).  Operands that TMA cannot
// describe -- a row gather, a row pitch that is not a multiple of 16 bytes -- stay on the cp.async
// kernel.  Measured: the C2 step goes from 76.0 to 73.7 us; opt-in (POZ_TMA_GEMM=1) because it computes
// in 3xTF32 rather than exact fp32 for that gain.
//
// Tile 32x64, 128 threads, BK = 32, three stages; split-K over a cluster (forward, backward-x) or over
// gradient slabs (backward-w) and the epilogues exactly as in gemm_tf32x3_pipelined.
#pragma once
#include <cuda.h>
#include "poz_tf32x3_gemm.cuh"

namespace poz {
namespace fast {

constexpr int TBK = 32;        // k depth of a TMA stage (32 floats = one 128-byte swizzle row)
constexpr int TSTAGES = 3;

__device__ __forceinline__ uint32_t tm_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tm_bar_init(uint64_t* b, uint32_t n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tm_smem(b)), "r"(n));
}
__device__ __forceinline__ void tm_bar_expect(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tm_smem(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tm_bar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nTMW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra TMD_%=;\nbra TMW_%=;\nTMD_%=:\n}"
        ::"r"(tm_smem(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tm_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* b) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(tm_smem(dst)), "l"(map), "r"(c0), "r"(c1), "r"(tm_smem(b)) : "memory");
}

// float offset of element (row r, column c) of a [rows][32] tile written by TMA with the 128-byte swizzle
__device__ __forceinline__ int swz(int r, int c) { return r * 32 + ((((c >> 2) ^ (r & 7)) << 2) | (c & 3)); }

// A: k-contiguous [BM][TBK] box or row-contiguous [TBK][BM] box.  B: k-contiguous [BN][TBK] box, or two
// row-contiguous [TBK][32] boxes (columns n0.. and n0+32..).
template <bool A_KC, bool B_KC, bool COLSUM, class Epi>
__global__ void __launch_bounds__(128)
gemm_tf32x3_tma(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, long long M, int N,
                long long Kd, long long k_chunk, long long bias_row, Epi epi) {
    constexpr int BM = 32, BN = 64, NT = 128;
    constexpr int A_ELEMS = BM * TBK;                 // 4 KB either way
    constexpr int B_ELEMS = BN * TBK;                 // 8 KB: [64][32] or two [32][32] halves
    constexpr int STAGE_ELEMS = A_ELEMS + B_ELEMS;
    constexpr int RED_ELEMS = (BM + 1) * BN;
    static_assert(RED_ELEMS <= TSTAGES * STAGE_ELEMS, "reduction tile must fit in the stage buffers");
    __shared__ __align__(1024) float smem[TSTAGES * STAGE_ELEMS];
    __shared__ uint64_t full[TSTAGES];

    const int tid = threadIdx.x;
    KernelTraceRec* ktr = nullptr;
    if (d_ktrace && tid == 0) {
        const unsigned int slot = atomicAdd(&d_ktrace_n, 1u);
        if (slot < KTRACE_CAP) {
            ktr = d_ktrace + slot;
            ktr->tag = 32 | (A_KC ? 1 : 0) | (B_KC ? 2 : 0) | (COLSUM ? 4 : 0);
            ktr->m = (int)M; ktr->n = N; ktr->k = (int)Kd;
            ktr->t[0] = ktrace_now();
        }
    }
    const int warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int mb = (warp >> 1) * 16, nb = (warp & 1) * 32;
    const long long m0 = (long long)blockIdx.y * BM;
    const int n0 = blockIdx.x * BN;
    const long long k_begin = (long long)blockIdx.z * k_chunk;
    const long long k_end = (k_begin + k_chunk < Kd) ? k_begin + k_chunk : Kd;
    const int nk = k_end > k_begin ? (int)((k_end - k_begin + TBK - 1) / TBK) : 0;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < TSTAGES; ++s) tm_bar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    float big[4][4], small[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int r = 0; r < 4; ++r) { big[j][r] = 0.f; small[j][r] = 0.f; }
    float bsum[4] = {0.f, 0.f, 0.f, 0.f};

    pdl_wait();   // the operands may come from the previous launch
    auto issue = [&](int s) {     // one thread
        float* sa = smem + (s % TSTAGES) * STAGE_ELEMS;
        float* sb = sa + A_ELEMS;
        uint64_t* bar = &full[s % TSTAGES];
        const int k0 = (int)(k_begin + (long long)s * TBK);
        tm_bar_expect(bar, STAGE_ELEMS * 4);
        if (A_KC) tm_load_2d(sa, &mapA, k0, (int)m0, bar);
        else      tm_load_2d(sa, &mapA, (int)m0, k0, bar);
        if (B_KC) tm_load_2d(sb, &mapB, k0, n0, bar);
        else {
            tm_load_2d(sb, &mapB, n0, k0, bar);
            tm_load_2d(sb + TBK * 32, &mapB, n0 + 32, k0, bar);
        }
    };
    if (tid == 0)
        for (int s = 0; s < TSTAGES - 1; ++s)
            if (s < nk) issue(s);

    if (ktr) ktr->t[1] = ktrace_now();
    for (int s = 0; s < nk; ++s) {
        tm_bar_wait(&full[s % TSTAGES], (uint32_t)((s / TSTAGES) & 1));
        if (ktr && s == 0) ktr->t[2] = ktrace_now();
        const float* sa = smem + (s % TSTAGES) * STAGE_ELEMS;
        const float* sb = sa + A_ELEMS;
        // the k range of the last split may end inside a box: the copy zero-fills beyond the tensor, but
        // inside the tensor the next split's k-values would be read -- k_chunk is a multiple of TBK (host)
#pragma unroll
        for (int ks = 0; ks < TBK; ks += 8) {
            float af[4];
            if (A_KC) {
                af[0] = sa[swz(mb + g, ks + t)];
                af[1] = sa[swz(mb + g + 8, ks + t)];
                af[2] = sa[swz(mb + g, ks + t + 4)];
                af[3] = sa[swz(mb + g + 8, ks + t + 4)];
            } else {
                af[0] = sa[swz(ks + t, mb + g)];
                af[1] = sa[swz(ks + t, mb + g + 8)];
                af[2] = sa[swz(ks + t + 4, mb + g)];
                af[3] = sa[swz(ks + t + 4, mb + g + 8)];
            }
            uint32_t ah[4], al[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) split_tf32(af[r], ah[r], al[r]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float bf[2];
                if (B_KC) {
                    bf[0] = sb[swz(nb + 8 * j + g, ks + t)];
                    bf[1] = sb[swz(nb + 8 * j + g, ks + t + 4)];
                } else {   // half (nb / 32) of the tile, column 8 j + g of it
                    const float* h = sb + (nb >> 5) * (TBK * 32);
                    bf[0] = h[swz(ks + t, 8 * j + g)];
                    bf[1] = h[swz(ks + t + 4, 8 * j + g)];
                }
                if (COLSUM) bsum[j] += bf[0] + bf[1];
                uint32_t bh[2], bl[2];
                split_tf32(bf[0], bh[0], bl[0]);
                split_tf32(bf[1], bh[1], bl[1]);
                mma_tf32(small[j], ah, bl);
                mma_tf32(small[j], al, bh);
                mma_tf32(big[j], ah, bh);
            }
        }
        __syncthreads();                       // every warp is done with this stage's buffers
        if (tid == 0 && s + TSTAGES - 1 < nk) issue(s + TSTAGES - 1);
    }
    pdl_trigger();
    if (ktr) ktr->t[3] = ktrace_now();

    // ---- the tile goes through shared memory: the accumulator fragments become rows ----
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int col = nb + 8 * j + 2 * t;
        *reinterpret_cast<float2*>(smem + (mb + g) * BN + col) = make_float2(big[j][0] + small[j][0], big[j][1] + small[j][1]);
        *reinterpret_cast<float2*>(smem + (mb + g + 8) * BN + col) = make_float2(big[j][2] + small[j][2], big[j][3] + small[j][3]);
    }
    if (COLSUM) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v = bsum[j];
            v += __shfl_xor_sync(0xffffffffu, v, 1);
            v += __shfl_xor_sync(0xffffffffu, v, 2);
            if (warp < 2 && t == 0) smem[BM * BN + nb + 8 * j + g] = v;
        }
    }

    cg::cluster_group cluster = cg::this_cluster();
    const unsigned csize = cluster.num_blocks();
    if (csize > 1) {
        cluster.sync();
        const unsigned rank = cluster.block_rank();
        const int rows_per = BM / (int)csize;
        for (int idx = tid; idx < rows_per * (BN / 4); idx += NT) {
            const int row = (int)rank * rows_per + idx / (BN / 4);
            const int c4 = (idx % (BN / 4)) * 4;
            float4 s4 = make_float4(0.f, 0.f, 0.f, 0.f);
            for (unsigned r = 0; r < csize; ++r) {
                const float* peer = cluster.map_shared_rank(smem, r);
                const float4 v = *reinterpret_cast<const float4*>(peer + row * BN + c4);
                s4.x += v.x; s4.y += v.y; s4.z += v.z; s4.w += v.w;
            }
            const long long m = m0 + row;
            const int n = n0 + c4;
            if (m < M && n < N) {
                const float vals[4] = {s4.x, s4.y, s4.z, s4.w};
                epi.template row<4>(m, n, vals, (N - n < 4) ? (N - n) : 4, 0);
            }
        }
        if (ktr) ktr->t[4] = ktrace_now();
        cluster.sync();
        if (ktr) ktr->t[5] = ktrace_now();
        return;
    }

    __syncthreads();
    for (int idx = tid; idx < BM * (BN / 4); idx += NT) {
        const int row = idx / (BN / 4), c4 = (idx % (BN / 4)) * 4;
        const long long m = m0 + row;
        const int n = n0 + c4;
        if (m < M && n < N) {
            const float4 v = *reinterpret_cast<const float4*>(smem + row * BN + c4);
            const float vals[4] = {v.x, v.y, v.z, v.w};
            epi.template row<4>(m, n, vals, (N - n < 4) ? (N - n) : 4, (int)blockIdx.z);
        }
    }
    if (COLSUM && blockIdx.y == 0 && tid < BN / 4) {
        const int c4 = tid * 4;
        const int n = n0 + c4;
        if (n < N) {
            const float4 v = *reinterpret_cast<const float4*>(smem + BM * BN + c4);
            const float vals[4] = {v.x, v.y, v.z, v.w};
            epi.template row<4>(bias_row, n, vals, (N - n < 4) ? (N - n) : 4, (int)blockIdx.z);
        }
    }
    if (ktr) { ktr->t[4] = ktrace_now(); ktr->t[5] = ktr->t[4]; }
}

}  // namespace fast
}  // namespace poz
