// The split-bf16 tcgen05 GEMM kernel (see poz_tc3_gemm.cuh for the scheme).
#pragma once
#include "poz_tc3_gemm.cuh"

namespace poz {
namespace tc3 {

// Epilogue of one 128 x BN tile for the 32 TMEM lanes of this warp's quarter and its column half: the two
// accumulators are added, bias / activation (or activation gradient) applied, and the result written as fp32
// and / or as the next GEMM's bf16 planes.
template <int BN, bool HEAD = false, bool ONE_ACC = false>
__device__ __forceinline__ void epilogue_tile(const Epi3& ep, long long M, int N, long long mb, int nb, int sp, bool empty_k,
                                              uint32_t tmem_base, uint32_t ab, int quarter, int chalf, int lane,
                                              const float* head_ws = nullptr, float* head_part = nullptr) {
    const long long m = mb * BM + quarter * 32 + lane;
    // fused head: HEAD_MAX_N running sums as packed fp32 pairs (fma.rn.f32x2: two FMAs per issue slot)
    unsigned long long hacc2[HEAD ? HEAD_MAX_N / 2 : 1];
    if constexpr (HEAD) {
#pragma unroll
        for (int o = 0; o < HEAD_MAX_N / 2; ++o) hacc2[o] = 0ull;
    }
    // ONE_ACC: all six products of a k-step went into one accumulator per buffer (the fp32-input kernel with its
    // X planes in tensor memory); otherwise hi*hi and the five small terms are kept apart and added here
    const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + ab * (ONE_ACC ? BN : 2 * BN) + chalf * (BN / 2);
#pragma unroll 1
    for (int c = 0; c < BN / 2; c += 32) {
        uint32_t v[32], w[32];
        tc::tmem_ld32(taddr + c, v);
        if constexpr (!ONE_ACC) tc::tmem_ld32(taddr + BN + c, w);
        tc::tmem_ld_wait();
        const int n = nb * BN + chalf * (BN / 2) + c;
        if (m < M && n < N) {
            const bool full = (n + 32 <= N);
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j)
                f[j] = empty_k ? 0.f : (ONE_ACC ? __uint_as_float(v[j]) : __uint_as_float(v[j]) + __uint_as_float(w[j]));
            if (ep.mode == EPI_BIAS_ACT) {
                if (ep.bias) {
                    if (full && ((reinterpret_cast<uintptr_t>(ep.bias + n) & 15) == 0)) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            const float4 b4 = __ldg(reinterpret_cast<const float4*>(ep.bias + n + j));
                            f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (n + j < N) f[j] += __ldg(ep.bias + n + j);
                    }
                }
                if (ep.act == ACT_RELU) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = f[j] < 0.f ? 0.f : f[j];     // (keeps a NaN a NaN, as torch.relu does)
                } else if (ep.act != ACT_NONE && ep.act != ACT_SOFTMAX) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) f[j] = act_fwd<float>(ep.act, f[j]);
                }
            } else if (ep.mode == EPI_ACTGRAD) {
                const float* aux = ep.aux + m * ep.ld_aux + n;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (full || n + j < N) f[j] *= act_bwd_from_out<float>(ep.act, __ldg(aux + j));
            }
            if constexpr (HEAD) {
                // this thread's 32 activations times the head's weight rows (staged [128][HEAD_MAX_N], zero beyond N / head_n);
                // every lane of the warp reads the same row: broadcast loads
                const int hq = (ep.head_n + 3) >> 2;
                const uint32_t wrow = tc::smem_u32(head_ws) + (uint32_t)(chalf * (BN / 2) + c) * (HEAD_MAX_N * 4);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    unsigned long long fj;
                    asm("mov.b64 %0, {%1, %1};" : "=l"(fj) : "f"(f[j]));
#pragma unroll
                    for (int q = 0; q < HEAD_MAX_N / 4; ++q)
                        if (q < hq) {
                            unsigned long long w0, w1;
                            asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "r"(wrow + (uint32_t)(j * HEAD_MAX_N + 4 * q) * 4));
                            asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(hacc2[2 * q]) : "l"(fj), "l"(w0));
                            asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(hacc2[2 * q + 1]) : "l"(fj), "l"(w1));
                        }
                }
            }
            if (ep.out_f32) {
                float* out = ep.out_f32 + (long long)sp * ep.slab + m * ep.ld_f32 + n;
                if (full && ((reinterpret_cast<uintptr_t>(out) & 15) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<float4*>(out + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (n + j < N) out[j] = f[j];
                }
            }
            if (ep.out_planes) {
                // pad columns N..ld_planes-1 of the planes must read as zero for the next GEMM
                __nv_bfloat16* o = ep.out_planes + m * ep.ld_planes + n;
                uint32_t ph[16], pm[16], pl[16];
#pragma unroll
                for (int j = 0; j < 32; j += 2)
                    split3_pair((full || n + j < N) ? f[j] : 0.f, (full || n + j + 1 < N) ? f[j + 1] : 0.f, ph[j / 2], pm[j / 2],
                                pl[j / 2]);
                const bool fullp = (n + 32 <= ep.ld_planes);
                if (fullp && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        reinterpret_cast<uint4*>(o)[q] = make_uint4(ph[4 * q], ph[4 * q + 1], ph[4 * q + 2], ph[4 * q + 3]);
                        reinterpret_cast<uint4*>(o + ep.plane_stride)[q] = make_uint4(pm[4 * q], pm[4 * q + 1], pm[4 * q + 2], pm[4 * q + 3]);
                        reinterpret_cast<uint4*>(o + 2 * ep.plane_stride)[q] = make_uint4(pl[4 * q], pl[4 * q + 1], pl[4 * q + 2], pl[4 * q + 3]);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; j += 2)
                        if (n + j < ep.ld_planes) {   // ld_planes is even
                            *reinterpret_cast<uint32_t*>(o + j) = ph[j / 2];
                            *reinterpret_cast<uint32_t*>(o + ep.plane_stride + j) = pm[j / 2];
                            *reinterpret_cast<uint32_t*>(o + 2 * ep.plane_stride + j) = pl[j / 2];
                        }
                }
            }
        }
    }
    if constexpr (HEAD) {
        // the warps of the second column half hand their partial sums over; the first half finishes the rows
        float hacc[HEAD_MAX_N];
#pragma unroll
        for (int o = 0; o < HEAD_MAX_N / 2; ++o) asm("mov.b64 {%0, %1}, %2;" : "=f"(hacc[2 * o]), "=f"(hacc[2 * o + 1]) : "l"(hacc2[o]));
        float4* pr = reinterpret_cast<float4*>(head_part + (quarter * 32 + lane) * HEAD_MAX_N);
        if (chalf == 1) {
#pragma unroll
            for (int q = 0; q < HEAD_MAX_N / 4; ++q) pr[q] = make_float4(hacc[4 * q], hacc[4 * q + 1], hacc[4 * q + 2], hacc[4 * q + 3]);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_WARPS * 32) : "memory");
        if (chalf == 0 && m < M) {
            const int hn = ep.head_n;
#pragma unroll
            for (int q = 0; q < HEAD_MAX_N / 4; ++q) {
                const float4 pv = pr[q];
                hacc[4 * q] += pv.x; hacc[4 * q + 1] += pv.y; hacc[4 * q + 2] += pv.z; hacc[4 * q + 3] += pv.w;
            }
            if (ep.head_b) {
#pragma unroll
                for (int o = 0; o < HEAD_MAX_N; ++o)
                    if (o < hn) hacc[o] += __ldg(ep.head_b + o);
            }
            if (ep.head_act == ACT_SOFTMAX) {
                float mx = -INFINITY;
#pragma unroll
                for (int o = 0; o < HEAD_MAX_N; ++o)
                    if (o < hn) mx = fmaxf(mx, hacc[o]);
                float sum = 0.f;
#pragma unroll
                for (int o = 0; o < HEAD_MAX_N; ++o)
                    if (o < hn) { hacc[o] = expf(hacc[o] - mx); sum += hacc[o]; }
                const float inv = 1.f / sum;
#pragma unroll
                for (int o = 0; o < HEAD_MAX_N; ++o) hacc[o] *= inv;
            } else if (ep.head_act != ACT_NONE) {
#pragma unroll
                for (int o = 0; o < HEAD_MAX_N; ++o) hacc[o] = act_fwd<float>(ep.head_act, hacc[o]);
            }
            float* o_row = ep.head_out + m * hn;
#pragma unroll
            for (int o = 0; o < HEAD_MAX_N; ++o)
                if (o < hn) o_row[o] = hacc[o];
        }
    }
}

template <int A_MAJOR, int B_MAJOR, int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
tc3_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                long long M, int N, long long K, int splits, long long k_per_split, Epi3 ep) {
    using C = Cfg<BN>;
    constexpr int STAGES = C::STAGES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full_bar = empty_bar + STAGES;      // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;      // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long tiles_m = (M + BM - 1) / BM;
    const int tiles_n = (N + BN - 1) / BN;
    const long long num_tiles = tiles_m * tiles_n * splits;

    if (warp == 8 && lane == 0) {
        tc::tma_prefetch_desc(&map_a);
        tc::tma_prefetch_desc(&map_b);
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&tmem_full_bar[b], 1); tc::mbar_init(&tmem_empty_bar[b], NUM_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        // ================= TMA producer =================
        if (tc::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int nb = (int)(t % tiles_n);
                const long long mb = (t / tiles_n) % tiles_m;
                const int sp = (int)(t / (tiles_n * tiles_m));
                const long long k0 = (long long)sp * k_per_split;
                const long long k1 = (k0 + k_per_split < K) ? k0 + k_per_split : K;
                const int nkb = k1 > k0 ? (int)((k1 - k0 + BK - 1) / BK) : 0;
                for (int kb = 0; kb < nkb; ++kb) {
                    tc::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * C::STAGE_BYTES;
                    uint8_t* sb = sa + PLANES * C::A_PLANE;
                    tc::mbar_expect_tx(&full_bar[stage], C::STAGE_BYTES);
                    const int kc = (int)(k0 + (long long)kb * BK);
#pragma unroll
                    for (int p = 0; p < PLANES; ++p) {
                        if (A_MAJOR == MAJOR_K) {
                            tma_load_3d(&map_a, &full_bar[stage], sa + p * C::A_PLANE, kc, (int)(mb * BM), p);
                        } else {
#pragma unroll
                            for (int j = 0; j < BM / 64; ++j)
                                tma_load_3d(&map_a, &full_bar[stage], sa + p * C::A_PLANE + j * 8192,
                                            (int)(mb * BM) + 64 * j, kc, p);
                        }
                        if (B_MAJOR == MAJOR_K) {
                            tma_load_3d(&map_b, &full_bar[stage], sb + p * C::B_PLANE, kc, nb * BN, p);
                        } else {
#pragma unroll
                            for (int j = 0; j < BN / 64; ++j)
                                tma_load_3d(&map_b, &full_bar[stage], sb + p * C::B_PLANE + j * 8192, nb * BN + 64 * j,
                                            kc, p);
                        }
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 9) {
        // ================= MMA issuer =================
        if (tc::elect_one()) {
            constexpr uint32_t idesc = make_idesc3<A_MAJOR, B_MAJOR, BN>();
            int stage = 0;
            uint32_t phase = 0;
            uint32_t it = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
                const int sp = (int)(t / (tiles_n * tiles_m));
                const long long k0 = (long long)sp * k_per_split;
                const long long k1 = (k0 + k_per_split < K) ? k0 + k_per_split : K;
                const int nkb = k1 > k0 ? (int)((k1 - k0 + BK - 1) / BK) : 0;
                const uint32_t ab = it & 1, aphase = (it >> 1) & 1;
                tc::mbar_wait(&tmem_empty_bar[ab], aphase ^ 1);
                tc::tc_fence_after();
                const uint32_t acc = tmem_base + ab * (2 * BN);      // main (hi*hi) accumulator
                const uint32_t acs = acc + BN;                       // small-terms accumulator
                for (int kb = 0; kb < nkb; ++kb) {
                    tc::mbar_wait(&full_bar[stage], phase);
                    tc::tc_fence_after();
                    const uint32_t sa = tc::smem_u32(smem + stage * C::STAGE_BYTES);
                    const uint32_t sb = sa + PLANES * C::A_PLANE;
#pragma unroll
                    for (int kk = 0; kk < BK / UMMA_K; ++kk) {
                        uint64_t ad[PLANES], bd[PLANES];
#pragma unroll
                        for (int p = 0; p < PLANES; ++p) {
                            ad[p] = (A_MAJOR == MAJOR_K) ? tc::make_desc(sa + p * C::A_PLANE + kk * 32, 16, 1024)
                                                         : tc::make_desc(sa + p * C::A_PLANE + kk * 2048, 8192, 1024);
                            bd[p] = (B_MAJOR == MAJOR_K) ? tc::make_desc(sb + p * C::B_PLANE + kk * 32, 16, 1024)
                                                         : tc::make_desc(sb + p * C::B_PLANE + kk * 2048, 8192, 1024);
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
    } else {
        // ================= epilogue =================
        const int quarter = warp & 3, chalf = warp >> 2;
        constexpr bool HEAD_CAPABLE = (A_MAJOR == MAJOR_K && B_MAJOR == MAJOR_MN && BN == 128);
        // fused head: its weights [N][head_n] staged once per CTA as [128][HEAD_MAX_N], zero-filled, behind the barriers
        float* head_ws = reinterpret_cast<float*>(smem + STAGES * C::STAGE_BYTES + 256);
        float* head_part = head_ws + 128 * HEAD_MAX_N;                  // [2][128][HEAD_MAX_N]
        const bool head = HEAD_CAPABLE && ep.head_w != nullptr;
        if (head) {
            for (int i = threadIdx.x; i < 128 * HEAD_MAX_N; i += NUM_EPI_WARPS * 32) {
                const int k = i / HEAD_MAX_N, o = i % HEAD_MAX_N;
                head_ws[i] = (k < N && o < ep.head_n) ? __ldg(ep.head_w + (long long)k * ep.head_n + o) : 0.f;
            }
            asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_WARPS * 32) : "memory");
        }
        uint32_t it = 0;
        for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
            const int nb = (int)(t % tiles_n);
            const long long mb = (t / tiles_n) % tiles_m;
            const int sp = (int)(t / (tiles_n * tiles_m));
            const long long k0 = (long long)sp * k_per_split;
            const bool empty_k = k0 >= K;
            const uint32_t ab = it & 1, aphase = (it >> 1) & 1;
            tc::mbar_wait(&tmem_full_bar[ab], aphase);
            tc::tc_fence_after();
            if (HEAD_CAPABLE && head)
                epilogue_tile<BN, HEAD_CAPABLE>(ep, M, N, mb, nb, sp, empty_k, tmem_base, ab, quarter, chalf, lane, head_ws,
                                                head_part + (it & 1) * 128 * HEAD_MAX_N);
            else
                epilogue_tile<BN>(ep, M, N, mb, nb, sp, empty_k, tmem_base, ab, quarter, chalf, lane);
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
