// "tcs": the deep contractions of a narrow float32 train step at SMALL batch on the tensor cores.
//
// At batch 1024 the two K/N = 784 GEMMs of the 784-256-128-10 step (forward of layer 0, weight gradient of
// layer 0) are 0.4 GFLOP each: half a microsecond of tensor time, 20 us on the FMA pipe as the SIMT kernels ran
// them.  What such a launch needs is not a persistent tile loop but enough CTAs at once and no staging pass:
//
//   * one 128 x 64 output tile per CTA, the contraction split over a thread-block CLUSTER of 2 / 4 / 8 CTAs
//     (grid = tiles_n x tiles_m x splits, ~one CTA per SM); the partial tiles meet through distributed shared
//     memory and every CTA of the cluster finishes 128 / splits rows of the tile -- bias / dropout / activation
//     for the forward, or the OPTIMIZER ITSELF for the weight gradient (poz_tcs.h: UpdateEpi), so the gradient
//     of the biggest layer never exists in memory and the update kernel has nothing left to do for it;
//   * fp32-grade products from bf16 MMAs: three planes per operand, six products per k-step, the five small ones
//     in their own accumulator (the scheme of poz_tc3_gemm.cuh);
//   * the A operand is converted on the fly: sixteen converter warps read the fp32 matrix straight from global
//     memory -- a row per lane for the forward (eight 16-byte loads per 32 k), a COLUMN per lane for the weight
//     gradient (X^T: 32 coalesced 4-byte loads, lanes = consecutive features), with the optional row gather of
//     train()'s sampling folded into the address -- split it into hi / mid / lo and write the planes into TENSOR
//     memory (tcgen05.st), from where tcgen05.mma takes them; nothing of A touches shared memory;
//   * the B operand (W's planes, or dZ's planes) comes by TMA, all k-blocks of the CTA's k range up front
//     (<= MAX_KB x 24 KB): no stage recycling, one mbarrier per k-block.
//
//   warps 0-15 : converters (group = warp / 4 takes half-blocks `group` and `group + 4` of the CTA's <= 8; lane = tile
//                row), then the epilogue     warp 16 : TMA producer of B     warp 17 : tensor memory + MMA issue
// Tensor memory (512 columns): main accumulator 0..63, small-terms accumulator 64..127, then one A-plane slot of
// 3 x 16 columns per half-block of the CTA's k range (<= 2 x MAX_KB): nothing is recycled, so the converters never
// wait for the MMAs.  Split-K exchange: every epilogue warp writes its rows of the partial tile straight into the
// shared memory of the CTA that finishes them (distributed shared memory stores), ONE cluster barrier follows, and
// each CTA sums what it received -- no CTA reads a peer's memory, so none has to wait for its peers before it exits.
#pragma once
#include <cooperative_groups.h>
#include "poz_tc3_x32t_kernel.cuh"
#include "poz_tcs.h"

namespace poz {
namespace tcs {

namespace cg = cooperative_groups;

constexpr int PLANES = 3;
constexpr int CONV_WARPS = 16;                       // four groups of four: every half-block of the CTA's k range is in flight at once
constexpr int THREADS = (CONV_WARPS + 2) * 32;
constexpr int B_PLANE = BN * BK * 2;                 // 8 KB
constexpr int B_KBLOCK = PLANES * B_PLANE;           // 24 KB
constexpr int PART_LD = 68;                          // floats per row of a partial tile: 16-byte rows, conflict-free float4 writes
constexpr int PART_BYTES = BM * PART_LD * 4;         // [splits][128 / splits rows][PART_LD]: the partial rows this CTA finishes, from every peer
constexpr int BAR_OFF = MAX_KB * B_KBLOCK + PART_BYTES;
constexpr int SMEM_BYTES = BAR_OFF + 256 + 1024;
constexpr int TMEM_COLS = 512;
constexpr int COL_MAIN = 0, COL_SMALL = BN, COL_A = 2 * BN, A_SLOT_COLS = PLANES * 16;
static_assert(COL_A + 2 * MAX_KB * A_SLOT_COLS <= TMEM_COLS, "tensor memory: two accumulators + one A-plane slot per half-block");
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

struct Args {
    const float* a;
    long long lda;
    const int* row_idx;
    int a_cols, ones_row;
    long long M;
    int N;
    long long K, k_per_split;
    int a_dep, b_dep;             // A / B's planes are written by an earlier launch of the same step: read only after pdl_wait()
};

// Up to MAX_GROUP problems of one kind in one launch (the weight gradients of ALL layers of a narrow stack: same
// contraction length = batch, so one cluster size fits all): blockIdx.x walks the problems' tiles one after the other.
struct GroupArgs {
    int n;
    int tile_end[MAX_GROUP];      // running sum of tiles_m x tiles_n
    int tiles_n[MAX_GROUP];
    Args a[MAX_GROUP];
};
template <class Epi> struct GroupEpi { Epi e[MAX_GROUP]; };

// 32 consecutive k of this lane's A' row, zero beyond the matrix.  Straight-line code: every load is issued
// unconditionally from a clamped address and the value selected afterwards, so the 8 (or 32) loads of a
// half-block are all in flight together.
template <bool A_T>
__device__ __forceinline__ void load_a(const Args& g, long long m, long long k0, int lane, float (&x)[32]) {
    if (!A_T) {
        const long long mc = m < g.M ? m : g.M - 1;
        const float* rp = g.a + (g.row_idx ? (long long)__ldg(g.row_idx + mc) : mc) * g.lda;
        const bool vec = ((reinterpret_cast<uintptr_t>(rp) & 15) == 0) && (g.K % 4 == 0);
        if (vec) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const long long k = k0 + 4 * c;
                const long long kc = k + 4 <= g.K ? k : g.K - 4;
                const float4 q = __ldg(reinterpret_cast<const float4*>(rp + kc));
                const bool ok = m < g.M && k + 4 <= g.K;
                x[4 * c] = ok ? q.x : 0.f; x[4 * c + 1] = ok ? q.y : 0.f; x[4 * c + 2] = ok ? q.z : 0.f; x[4 * c + 3] = ok ? q.w : 0.f;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const long long k = k0 + j;
                const float v = __ldg(rp + (k < g.K ? k : g.K - 1));
                x[j] = (m < g.M && k < g.K) ? v : 0.f;
            }
        }
    } else {
        const bool real = m < g.a_cols;
        const long long mc = real ? m : g.a_cols - 1;
        const float fill = (g.ones_row && m == g.a_cols) ? 1.f : 0.f;
        // source rows of the 32 k: lane j fetches row_idx[k0 + j] once, the values go round by shuffle
        const long long kl = k0 + lane < g.K ? k0 + lane : g.K - 1;
        const int my_row = g.row_idx ? __ldg(g.row_idx + kl) : (int)kl;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const int r = __shfl_sync(0xffffffffu, my_row, j);
            const float v = __ldg(g.a + (long long)r * g.lda + mc);
            x[j] = (k0 + j < g.K) ? (real ? v : fill) : 0.f;
        }
    }
}

template <bool A_T, class Epi>
__global__ void __launch_bounds__(THREADS, 1)
tcs_gemm_kernel(const __grid_constant__ GroupArgs ga, const __grid_constant__ CUtensorMap map0, const __grid_constant__ CUtensorMap map1,
                const __grid_constant__ CUtensorMap map2, const __grid_constant__ CUtensorMap map3,
                const __grid_constant__ GroupEpi<Epi> ge) {
    // which problem this CTA belongs to, and its tile there
    int prob = 0;
#pragma unroll
    for (int q = 1; q < MAX_GROUP; ++q)
        if (q < ga.n && (int)blockIdx.x >= ga.tile_end[q - 1]) prob = q;
    const int tile = (int)blockIdx.x - (prob > 0 ? ga.tile_end[prob - 1] : 0);
    const Args& g = ga.a[prob];
    const CUtensorMap* map_b_ptr = prob == 0 ? &map0 : prob == 1 ? &map1 : prob == 2 ? &map2 : &map3;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sB = smem;                                                  // [MAX_KB][3][64 k][64 n] bf16, 128-byte swizzle
    float* part = reinterpret_cast<float*>(smem + MAX_KB * B_KBLOCK);    // [splits][128 / splits][PART_LD]
    uint64_t* b_full = reinterpret_cast<uint64_t*>(smem + BAR_OFF);      // [MAX_KB]
    uint64_t* a_full = b_full + MAX_KB;                                  // [2 * MAX_KB] planes of half-block hb are in tensor memory
    uint64_t* acc_full = a_full + 2 * MAX_KB;                            // [1]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nb = tile % ga.tiles_n[prob];
    const long long mb = tile / ga.tiles_n[prob];
    const long long k_beg = (long long)blockIdx.z * g.k_per_split;
    const long long k_end = (k_beg + g.k_per_split < g.K) ? k_beg + g.k_per_split : g.K;
    const int nkb = k_end > k_beg ? (int)((k_end - k_beg + BK - 1) / BK) : 0;
    // POZ_KERNEL_TRACE=1: time stamps of this CTA (tags 200 forward / 201 weight gradient): set-up done, first
    // half-block in tensor memory, accumulators complete, partial tiles exchanged, epilogue done
    fast::KernelTraceRec* ktr = fast::ktrace_open(A_T ? 201 : 200, (int)g.M, g.N, (int)g.K);

    // Programmatic dependent launch (when the launch carries the attribute; no-ops otherwise): the next launch's CTAs may
    // be scheduled as soon as every CTA of this one has got here, and this launch's own reads of what its predecessors
    // wrote -- B's planes, and A where a_dep says so -- come after pdl_wait().
    pdl_trigger();
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned csize = cluster.num_blocks(), rank = cluster.block_rank();
    const int rows_per = BM / (int)csize;                                // rows of the tile this CTA finishes

    if (warp == CONV_WARPS && lane == 0) {
        tc::tma_prefetch_desc(map_b_ptr);
        for (int s = 0; s < MAX_KB; ++s) tc::mbar_init(&b_full[s], 1);
        for (int h = 0; h < 2 * MAX_KB; ++h) tc::mbar_init(&a_full[h], 4);
        tc::mbar_init(acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == CONV_WARPS + 1) tc::tmem_alloc(tmem_slot, TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    if (ktr) ktr->t[1] = fast::ktrace_now();

    const Epi& epi = ge.e[prob];              // stays in the parameter space; only its State lives in registers
    typename Epi::State est;
    if (warp == CONV_WARPS) {
        // ================= TMA producer: every k-block of B's planes, up front =================
        if (tc::elect_one()) {
            if (g.b_dep) pdl_wait();
            for (int kb = 0; kb < nkb; ++kb) {
                tc::mbar_expect_tx(&b_full[kb], B_KBLOCK);
#pragma unroll
                for (int p = 0; p < PLANES; ++p)
                    tc3::tma_load_3d(map_b_ptr, &b_full[kb], sB + kb * B_KBLOCK + p * B_PLANE, nb * BN, (int)(k_beg + (long long)kb * BK), p);
            }
        }
    } else if (warp == CONV_WARPS + 1) {
        // ================= MMA issuer =================
        if (tc::elect_one()) {
            constexpr uint32_t idesc = tc3::make_idesc3<tc::MAJOR_K, tc::MAJOR_MN, BN>();
            const uint32_t acc = tmem_base + COL_MAIN, acs = tmem_base + COL_SMALL;
            for (int kb = 0; kb < nkb; ++kb) {
                tc::mbar_wait(&b_full[kb], 0);
                const uint32_t sb = tc::smem_u32(sB + kb * B_KBLOCK);
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    tc::mbar_wait(&a_full[2 * kb + h], 0);
                    tc::tc_fence_after();
                    const uint32_t ta = tmem_base + COL_A + (2 * kb + h) * A_SLOT_COLS;
#pragma unroll
                    for (int j = 0; j < 2; ++j) {
                        const int kk = 2 * h + j;
                        uint32_t ad[PLANES];
                        uint64_t bd[PLANES];
#pragma unroll
                        for (int p = 0; p < PLANES; ++p) {
                            ad[p] = ta + p * 16 + j * 8;
                            bd[p] = tc::make_desc(sb + p * B_PLANE + kk * 2048, 8192, 1024);
                        }
                        const uint32_t first = (kb | kk) != 0 ? 1u : 0u;
                        tc3::umma_bf16_ts(acs, ad[0], bd[2], idesc, first);
                        tc3::umma_bf16_ts(acs, ad[2], bd[0], idesc, 1u);
                        tc3::umma_bf16_ts(acs, ad[1], bd[1], idesc, 1u);
                        tc3::umma_bf16_ts(acs, ad[0], bd[1], idesc, 1u);
                        tc3::umma_bf16_ts(acs, ad[1], bd[0], idesc, 1u);
                        tc3::umma_bf16_ts(acc, ad[0], bd[0], idesc, first);
                    }
                }
            }
            if (nkb > 0) tc::umma_commit(acc_full);
            else tc::mbar_arrive(acc_full);
        }
    } else {
        // ================= converters, then the partial tile to the CTAs that finish it =================
        // group g (warps 4g..4g+3, one tile row per lane) owns half-blocks g and g + 4 of the CTA's <= 8: both are
        // requested before anything is converted, so the whole A operand of the CTA is one memory round trip
        const int quarter = warp & 3, grp = warp >> 2;
        const int r = quarter * 32 + lane;
        const long long m = mb * BM + r;
        const uint32_t ta0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + COL_A;
        const int nhb = 2 * nkb;
        float xa[32], xb[32];
        if (g.a_dep) pdl_wait();
        if (grp < nhb) load_a<A_T>(g, m, k_beg + 32LL * grp, lane, xa);
        if (grp + 4 < nhb) load_a<A_T>(g, m, k_beg + 32LL * (grp + 4), lane, xb);
        auto convert_store = [&](const float (&x)[32], int hb) {
            const uint32_t ta = ta0 + (uint32_t)hb * A_SLOT_COLS;
#pragma unroll
            for (int half = 0; half < 2; ++half) {            // 16 k at a time: 8 packed words per plane
                uint32_t ph[8], pm[8], pl[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) split3_pair(x[16 * half + 2 * c], x[16 * half + 2 * c + 1], ph[c], pm[c], pl[c]);
                tc3::tmem_st8(ta + 8 * half, ph);
                tc3::tmem_st8(ta + 16 + 8 * half, pm);
                tc3::tmem_st8(ta + 32 + 8 * half, pl);
            }
            tc3::tmem_st_wait();
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&a_full[hb]);
        };
        if (grp < nhb) convert_store(xa, grp);
        if (ktr) ktr->t[2] = fast::ktrace_now();
        if (grp + 4 < nhb) convert_store(xb, grp + 4);
        // what the epilogue will need besides the gradient (the optimizer's w, m, v): requested now, used after the exchange
        epi.prefetch(est, mb * BM + (long long)rank * rows_per, nb * BN, rows_per, g.M, g.N);
        if (warp < 8) {
            // main + small accumulators of this warp's 32 rows and column half -> the shared memory of the CTA that finishes the row
            tc3::mbar_wait_idle(acc_full, 0);
            tc::tc_fence_after();
            if (ktr) ktr->t[3] = fast::ktrace_now();
            const int chalf = grp;
            const uint32_t tacc = tmem_base + ((uint32_t)(quarter * 32) << 16) + chalf * 32;
            uint32_t v[32], w[32];
            tc::tmem_ld32(tacc + COL_MAIN, v);
            tc::tmem_ld32(tacc + COL_SMALL, w);
            tc::tmem_ld_wait();
            const unsigned owner = (unsigned)(r / rows_per);
            float* dst = cluster.map_shared_rank(part, owner) + ((long long)rank * rows_per + (r % rows_per)) * PART_LD + chalf * 32;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                float4 q;
                q.x = nkb > 0 ? __uint_as_float(v[j]) + __uint_as_float(w[j]) : 0.f;
                q.y = nkb > 0 ? __uint_as_float(v[j + 1]) + __uint_as_float(w[j + 1]) : 0.f;
                q.z = nkb > 0 ? __uint_as_float(v[j + 2]) + __uint_as_float(w[j + 2]) : 0.f;
                q.w = nkb > 0 ? __uint_as_float(v[j + 3]) + __uint_as_float(w[j + 3]) : 0.f;
                *reinterpret_cast<float4*>(dst + j) = q;
            }
        }
    }

    // ================= one cluster barrier: every partial row has arrived where it is finished =================
    tc::tc_fence_before();
    cluster.sync();
    pdl_wait();      // the epilogue may read the step state that the launch before wrote (dropout keys)
    if (ktr) ktr->t[4] = fast::ktrace_now();
    if (warp < CONV_WARPS) {
        // 512 threads, rows_per x 16 four-wide pieces: piece i of this thread is its prefetch slot i
        int slot = 0;
        for (int idx = threadIdx.x; idx < rows_per * (BN / 4); idx += CONV_WARPS * 32, ++slot) {
            const int row = idx / (BN / 4);
            const int c4 = (idx % (BN / 4)) * 4;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int q = 0; q < MAX_SPLITS; ++q)
                if (q < (int)csize) {
                    const float4 t = *reinterpret_cast<const float4*>(part + (q * rows_per + row) * PART_LD + c4);
                    s.x += t.x; s.y += t.y; s.z += t.z; s.w += t.w;
                }
            const long long mo = mb * BM + (long long)rank * rows_per + row;
            const int n = nb * BN + c4;
            if (mo < g.M && n < g.N) {
                const float vals[4] = {s.x, s.y, s.z, s.w};
                epi.template row<4>(est, mo, n, vals, (g.N - n < 4) ? (g.N - n) : 4, slot);
            }
        }
    }
    epi.finish(est);
    if (ktr) ktr->t[5] = fast::ktrace_now();
    if (warp == CONV_WARPS + 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

}  // namespace tcs
}  // namespace poz
