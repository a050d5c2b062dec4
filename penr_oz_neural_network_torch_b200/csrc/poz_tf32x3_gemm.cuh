// fp32 GEMM of the narrow path on the warp-level tensor-core instruction, error-compensated
// ("3xTF32"): every fp32 operand x is split in registers into hi = tf32(x) and lo = tf32(x - hi), and
//     A B  ~=  Ahi Bhi + (Ahi Blo + Alo Bhi)
// is accumulated in fp32, the large and the small terms in separate accumulators (the tensor core's
// accumulation rounds toward zero, so the small terms would otherwise be lost against the large sum).
// Measured on B200 (scripts/micro/mma_rate.cu): mma.sync m16n8k8 tf32 issues at 278 TFLOP/s against
// 70 TFLOP/s for FFMA; three MMAs per product make that ~93 TFLOP/s of fp32-grade work, and -- what
// matters more at these sizes -- a 16x32x8 warp step costs 60 instructions and 12 conflict-free
// 32-bit shared-memory reads where the FFMA kernel needs 144 instructions and 16 128-bit reads.
// (tcgen05 is not used here: its operands come from shared memory, so the hi/lo planes would have to
// be materialised there first; poz_tc3 does that for layers that are large enough.)
//
// Same interface, tile (32x64, 128 threads, BK = 16, four cp.async stages), cluster split-K and
// epilogues as gemm_f32_pipelined in poz_simt_fast.cuh.  Warp w owns the 16x32 block (w / 2, w % 2) of
// the tile as four m16n8 accumulators.
#pragma once
#include "poz_simt_fast.cuh"

namespace poz {
namespace fast {

__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
    const float r = x - __uint_as_float(hi);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}

__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <bool A_KC, bool B_KC, bool COLSUM, class Epi>
__global__ void __launch_bounds__(128)
gemm_tf32x3_pipelined(Op A, Op B, long long M, int N, long long Kd, long long k_chunk, long long bias_row, Epi epi) {
    constexpr int BM = 32, BN = 64, NT = 128;
    constexpr int STAGES = 6;   // shadows fast::STAGES: this kernel's tiles are small enough for a deeper pipeline
    // mn-contiguous tiles are padded to a row pitch of 8 (mod 32) words so that the fragment reads
    // (4 k-rows x 8 consecutive columns per instruction) hit 32 distinct banks
    constexpr int PMN = 8;
    constexpr int A_LD = A_KC ? BK + 4 : BM + PMN;
    constexpr int B_LD = B_KC ? BK + 4 : BN + PMN;
    constexpr int A_ELEMS = (A_KC ? BM : BK) * A_LD;
    constexpr int B_ELEMS = (B_KC ? BN : BK) * B_LD;
    constexpr int STAGE_ELEMS = A_ELEMS + B_ELEMS;
    constexpr int RED_ELEMS = (BM + 1) * BN;
    constexpr int SMEM_ELEMS = STAGES * STAGE_ELEMS > RED_ELEMS ? STAGES * STAGE_ELEMS : RED_ELEMS;
    __shared__ __align__(16) float smem[SMEM_ELEMS];

    const int tid = threadIdx.x;
    KernelTraceRec* ktr = nullptr;
    if (d_ktrace && tid == 0) {
        const unsigned int slot = atomicAdd(&d_ktrace_n, 1u);
        if (slot < KTRACE_CAP) {
            ktr = d_ktrace + slot;
            ktr->tag = 16 | (A_KC ? 1 : 0) | (B_KC ? 2 : 0) | (COLSUM ? 4 : 0);
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
    const int nk = k_end > k_begin ? (int)((k_end - k_begin + BK - 1) / BK) : 0;

    float big[4][4], small[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int r = 0; r < 4; ++r) { big[j][r] = 0.f; small[j][r] = 0.f; }
    float bsum[4] = {0.f, 0.f, 0.f, 0.f};

    pdl_wait();   // everything above is index arithmetic; the operands may come from the previous launch
    const bool a_fast = A.vec4 != 0;
    const bool b_fast = B.vec4 != 0;
    TileCopier<BM, NT, A_KC, PMN> ca;
    TileCopier<BN, NT, B_KC, PMN> cb;
    if (a_fast) ca.init(A, m0, k_begin, k_end, tid);
    if (b_fast) cb.init(B, n0, k_begin, k_end, tid);
    auto issue = [&](int s) {
        float* sa = smem + (s % STAGES) * STAGE_ELEMS;
        float* sb = sa + A_ELEMS;
        const long long k0 = k_begin + (long long)s * BK;
        if (a_fast) ca.issue(sa, s); else load_tile<BM, NT, A_KC, PMN>(sa, A, m0, k0, k_end, tid);
        if (b_fast) cb.issue(sb, s); else load_tile<BN, NT, B_KC, PMN>(sb, B, n0, k0, k_end, tid);
    };
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) {
        if (s < nk) issue(s);
        cp_commit();
    }
    if (ktr) ktr->t[1] = ktrace_now();
    for (int s = 0; s < nk; ++s) {
        cp_wait<STAGES - 2>();
        __syncthreads();
        if (ktr && s == 0) ktr->t[2] = ktrace_now();
        if (s + STAGES - 1 < nk) issue(s + STAGES - 1);
        cp_commit();
        const float* sa = smem + (s % STAGES) * STAGE_ELEMS;
        const float* sb = sa + A_ELEMS;
#pragma unroll
        for (int ks = 0; ks < BK; ks += 8) {
            // A fragment (16 x 8): rows g, g + 8; k = t, t + 4
            float af[4];
            if (A_KC) {
                af[0] = sa[(mb + g) * A_LD + ks + t];
                af[1] = sa[(mb + g + 8) * A_LD + ks + t];
                af[2] = sa[(mb + g) * A_LD + ks + t + 4];
                af[3] = sa[(mb + g + 8) * A_LD + ks + t + 4];
            } else {
                af[0] = sa[(ks + t) * A_LD + mb + g];
                af[1] = sa[(ks + t) * A_LD + mb + g + 8];
                af[2] = sa[(ks + t + 4) * A_LD + mb + g];
                af[3] = sa[(ks + t + 4) * A_LD + mb + g + 8];
            }
            uint32_t ah[4], al[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) split_tf32(af[r], ah[r], al[r]);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                // B fragment (8 x 8): k = t, t + 4; column g of the j-th 8-column block
                float bf[2];
                if (B_KC) {
                    bf[0] = sb[(nb + 8 * j + g) * B_LD + ks + t];
                    bf[1] = sb[(nb + 8 * j + g) * B_LD + ks + t + 4];
                } else {
                    bf[0] = sb[(ks + t) * B_LD + nb + 8 * j + g];
                    bf[1] = sb[(ks + t + 4) * B_LD + nb + 8 * j + g];
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
    }
    cp_wait<0>();
    pdl_trigger();
    if (ktr) ktr->t[3] = ktrace_now();
    __syncthreads();                           // everyone is done with the pipeline buffers

    // ---- the tile goes through shared memory: the accumulator fragments become rows ----
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int col = nb + 8 * j + 2 * t;
        float2 lo2 = make_float2(big[j][0] + small[j][0], big[j][1] + small[j][1]);
        float2 hi2 = make_float2(big[j][2] + small[j][2], big[j][3] + small[j][3]);
        *reinterpret_cast<float2*>(smem + (mb + g) * BN + col) = lo2;
        *reinterpret_cast<float2*>(smem + (mb + g + 8) * BN + col) = hi2;
    }
    if (COLSUM) {
        // column sums of B: lanes with the same g hold disjoint k's of column g; the two row-block warps
        // of a column half hold the same numbers, so one of them writes
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
        // ---- split-K across the cluster: reduce partial tiles through distributed shared memory ----
        cluster.sync();
        const unsigned rank = cluster.block_rank();
        const int rows_per = BM / (int)csize;   // csize divides BM (host guarantees)
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
        cluster.sync();                        // peers may still be reading this CTA's tile
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
