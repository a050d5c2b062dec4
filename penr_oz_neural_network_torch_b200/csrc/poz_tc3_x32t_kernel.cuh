// Forward split-bf16 tcgen05 GEMM from fp32 rows, X planes in TENSOR memory: out = epilogue(X[M][K] * W[K][N]).
//
// TMA brings fp32 half-blocks (128 rows x 32 k, 128-byte swizzle) into a five-slot shared-memory ring and eight
// converter warps split them into hi / mid / lo planes.  A first version (round 1, removed) wrote those planes into
// shared memory, and its six MMAs per k-step read both operands from there: 8 KB per 128x128x16 MMA, i.e. 128 B per
// clock at the tensor pipe's full rate -- all the shared-memory bandwidth there is, before the TMA writes and the
// converters' own loads and stores (ncu: tensor pipe 51 % active, LSU data pipe 66 %; 0.60 ms per 262,144-row pass
// against 0.50 ms for this kernel).  Here the converters put the planes into
// tensor memory (`tcgen05.st`, one row per lane, two bf16 per 32-bit column) and the MMAs take A from there
// (`tcgen05.mma` with a TMEM A operand), so shared memory carries only the fp32 half-blocks and W's planes.
// Tensor memory then holds: two 128-column main accumulators (hi*hi; double-buffered tiles), ONE 128-column accumulator
// for the five small products, and two X-plane half-block slots (3 planes x 16 columns each; converter group h owns
// slot h).  The tensor core truncates its fp32 accumulator at every MMA, so the small terms must stay out of the main
// sum (one accumulator for all six: max relative error of the output probabilities 1.8e-5 against 2.3e-6); the small
// accumulator is shared between the two tile buffers by folding it into the main one first thing in the epilogue
// (tcgen05.ld both, add, tcgen05.st) and releasing it to the next tile's MMAs right after.
//   warps 0-7 : epilogue          warp 8 : TMA producer of the fp32 X half-blocks      warp 18 : TMA producer of W's planes
//   warp  9   : TMEM + MMA        warps 10-17 : converters (warp w: rows 32 (w % 4).., half-block (w - 10) / 4 of each k-block)
#pragma once
#include "poz_tc3_kernel.cuh"

namespace poz {
namespace tc3 {

constexpr int X32_CONV_WARPS = 8;                   // two groups of four (default): one tile row per lane
constexpr int X32_THREADS = (NUM_EPI_WARPS + 2 + X32_CONV_WARPS) * 32;
constexpr int x32t_threads(int ncg) { return (NUM_EPI_WARPS + 3 + 4 * ncg) * 32; }
constexpr int X32_RAW_BYTES = BM * 32 * 4;          // one half-block: 128 rows x 32 floats

__device__ __forceinline__ void tma_load_2d_f32(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(tc::smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}

// POZ_X32T_PROBE=1: cycles spent by converter warp 10 of block 0 in the phases of its loop, summed over the launch:
// [0] waiting for the fp32 half-block, [1] loads + conversion, [2] waiting for the tensor-memory slot, [3] tcgen05.st + its
// wait, [4] fences / arrivals, [5] iterations, [6] whole loop
// and of the MMA issuer of block 0: [8] waiting for the accumulators at a tile boundary, [9] for W's planes, [10] for X's planes,
// [11] the rest (issuing), [12] k-blocks
__device__ unsigned long long x32t_probe[16];

constexpr int X32T_SLOTS = 5;                       // fp32 half-blocks in flight in shared memory (default)
constexpr int X32T_SLOT_COLS = PLANES * 16;         // 32 k-values of three planes, two per column
constexpr int X32T_W_WARP = NUM_EPI_WARPS + 2 + X32_CONV_WARPS;   // warp 18: producer of W's planes
constexpr int X32T_THREADS = X32_THREADS + 32;

template <int BN, int STAGES_, bool HEAD, bool F16 = false, int RAW_SLOTS_ = X32T_SLOTS> struct CfgX32T {
    static constexpr int STAGES = STAGES_;          // W-plane stages (one per 64-wide k-block)
    static constexpr int RAW_SLOTS = RAW_SLOTS_;    // fp32 half-blocks of X in flight in shared memory
    static constexpr int NPL = F16 ? 2 : PLANES;    // planes per operand
    static constexpr int B_PLANE = BN * BK * 2;
    static constexpr int B_STAGE = NPL * B_PLANE;
    static constexpr int RAW_OFF = STAGES * B_STAGE;
    static constexpr int BAR_OFF = RAW_OFF + RAW_SLOTS * X32_RAW_BYTES;
    static constexpr int HEAD_OFF = BAR_OFF + 512;  // fused head: staged weights + two partial-sum buffers (poz_tc3.h)
    static constexpr int SMEM_BYTES = HEAD_OFF + (HEAD ? HEAD_SMEM_BYTES : 0) + 1024;
    static constexpr int SMALL_COL0 = 2 * BN;       // after the two main accumulators
    static constexpr int A_COL0 = 3 * BN;
    static constexpr int TMEM_COLS = 512;
    // X-plane slots in tensor memory, each one half-block (32 k) of NPL planes: one per converter group, two in the fp16 form
    // (its planes are smaller; measured neutral: the converters, not the hand-over of a slot, set the pace)
    static constexpr int SLOT_COLS = NPL * 16;
    static constexpr int NSLOT = F16 ? 4 : 2;
    static_assert(A_COL0 + NSLOT * SLOT_COLS <= TMEM_COLS, "tensor memory: accumulators + the X-plane slots");
    static_assert(SMEM_BYTES <= 227 * 1024, "shared memory of the fp32-input kernel");
};

// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// this lane's eight 32-bit columns starting at taddr
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
                 : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
          "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
          "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
          "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t saddr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(saddr) : "memory");
    return v;
}
// Wait of a warp that has nothing else to do for microseconds (the epilogue warps during a tile's main loop, the X
// producer when its ring is full): sleep between polls so that the polling does not take issue slots from the
// converters -- ncu counted 29 % of the kernel's executed instructions in the epilogue warps' poll loop.
__device__ __forceinline__ void mbar_wait_idle(uint64_t* bar, uint32_t parity) {
    for (uint32_t spins = 0; !tc::mbar_try_wait(bar, parity); ++spins) {
        __nanosleep(spins < 64 ? 20 : 100);
        if (spins > 20u * 1000u * 1000u) tc::mbar_timeout();
    }
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// F16: two fp16 planes per operand and three MMAs per product instead of three bf16 planes and six (poz_tc3_gemm.cuh,
// split2h_pair): half the work for the tensor pipe, a third less for the converters and for the W loads.
// NCG: converter groups of four warps each.  Half-blocks are numbered through the whole launch (hbg); group g converts
// hbg = g, g + NCG, ...; half-block hbg uses ring slot hbg % RAW_SLOTS and tensor-memory slot hbg % NSLOT.  A converter
// warp's loop is serial -- wait for data 290, load + convert 510, wait for the slot, publish 380 cycles (POZ_X32T_PROBE=1) --
// so with two groups the MMAs (768 cycles of tensor pipe per k-block in the fp16 form) were fed one k-block per ~1600.
// A third group (736 threads, 80 registers) was measured: 966 instead of 508 cycles for load + convert, the pass slower
// (0.358 against 0.338 ms) -- the warps of a scheduler share the conversion pipe, which is what the time goes to.
template <int BN, int NSTAGES, bool HEAD, bool F16 = false, int RAW_SLOTS = X32T_SLOTS, int NCG = 2>
__global__ void __launch_bounds__((NUM_EPI_WARPS + 3 + 4 * NCG) * 32, 1)
tc3_fwd_x32t_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_b,
                    long long M, int N, long long K, Epi3 ep) {
    using C = CfgX32T<BN, NSTAGES, HEAD, F16, RAW_SLOTS>;
    constexpr int STAGES = C::STAGES;
    constexpr int NPL = C::NPL;
    constexpr int NSLOT = C::NSLOT;
    constexpr int W_WARP = NUM_EPI_WARPS + 2 + 4 * NCG;      // producer of W's planes: the warp after the converters
    if (ep.run_if && *ep.run_if == 0) return;        // the repeat of a pass that needed none (uniform: before any barrier)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* raw = smem + C::RAW_OFF;                                  // [X32T_SLOTS][128 rows][32 floats], swizzled
    uint64_t* b_full = reinterpret_cast<uint64_t*>(smem + C::BAR_OFF); // [STAGES] W planes landed
    uint64_t* b_empty = b_full + STAGES;                               // [STAGES] MMAs of the stage retired
    uint64_t* a_full = b_empty + STAGES;                               // [NSLOT] planes of a half-block are in tensor memory
    uint64_t* a_empty = a_full + NSLOT;                                // [NSLOT] MMAs that read them retired
    uint64_t* raw_full = a_empty + NSLOT;                              // [SLOTS]
    uint64_t* raw_empty = raw_full + RAW_SLOTS;                        // [SLOTS]
    uint64_t* tmem_full_bar = raw_empty + RAW_SLOTS;                   // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;                      // [2]
    uint64_t* small_empty_bar = tmem_empty_bar + 2;                    // [1] small-terms accumulator folded into the main one
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(small_empty_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long tiles_m = (M + BM - 1) / BM;
    const int tiles_n = (N + BN - 1) / BN;
    const long long num_tiles = tiles_m * tiles_n;
    const int nkb = (int)((K + BK - 1) / BK);

    if (warp == 8 && lane == 0) {
        tc::tma_prefetch_desc(&map_x);
        tc::tma_prefetch_desc(&map_b);
        for (int s = 0; s < STAGES; ++s) { tc::mbar_init(&b_full[s], 1); tc::mbar_init(&b_empty[s], 1); }
        for (int h = 0; h < NSLOT; ++h) { tc::mbar_init(&a_full[h], 4); tc::mbar_init(&a_empty[h], 1); }
        for (int q = 0; q < RAW_SLOTS; ++q) { tc::mbar_init(&raw_full[q], 1); tc::mbar_init(&raw_empty[q], 4); }
        tc::mbar_init(small_empty_bar, NUM_EPI_WARPS);
        for (int b = 0; b < 2; ++b) { tc::mbar_init(&tmem_full_bar[b], 1); tc::mbar_init(&tmem_empty_bar[b], NUM_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) tc::tmem_alloc(tmem_slot, C::TMEM_COLS);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 8) {
        // ================= TMA producer, X: fp32 half-blocks into the raw ring =================
        // (W's planes have their own producer warp: with both in one thread the wait for a free W stage also held
        // back the X loads, and the converters of a short-K layer spent their time waiting for data)
        if (tc::elect_one()) {
            uint32_t slot = 0, rphase = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const long long mb = t / tiles_n;
                for (int hb = 0; hb < 2 * nkb; ++hb) {
                    mbar_wait_idle(&raw_empty[slot], rphase ^ 1);
                    tc::mbar_expect_tx(&raw_full[slot], X32_RAW_BYTES);
                    tma_load_2d_f32(&map_x, &raw_full[slot], raw + slot * X32_RAW_BYTES, 32 * hb, (int)(mb * BM));
                    if (++slot == RAW_SLOTS) { slot = 0; rphase ^= 1; }
                }
            }
        }
    } else if (warp == W_WARP) {
        // ================= TMA producer, W: one stage of planes per 64-wide k-block =================
        if (tc::elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int nb = (int)(t % tiles_n);
                for (int kb = 0; kb < nkb; ++kb) {
                    const int kc = kb * BK;
                    tc::mbar_wait(&b_empty[stage], phase ^ 1);
                    uint8_t* sb = smem + stage * C::B_STAGE;
                    tc::mbar_expect_tx(&b_full[stage], C::B_STAGE);
#pragma unroll
                    for (int p = 0; p < NPL; ++p)
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
            constexpr uint32_t idesc = F16 ? make_idesc_f16<MAJOR_K, MAJOR_MN, BN>() : make_idesc3<MAJOR_K, MAJOR_MN, BN>();
            if (F16 && ep.f16_flag && ep.f16_w_amax && f16_range_bad(__uint_as_float(*ep.f16_w_amax))) *ep.f16_flag = 1;
            int stage = 0;
            uint32_t phase = 0;
            uint32_t kbc = 0;                              // k-blocks done: parity of the per-half barriers
            uint32_t it = 0;
            const uint32_t acs = tmem_base + C::SMALL_COL0; // small-terms accumulator (shared by the two tile buffers)
            const bool mprobe = ep.probe && blockIdx.x == 0;
            long long mc[4] = {0, 0, 0, 0}, mt = mprobe ? clock64() : 0, mn = 0;
            for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
                const uint32_t ab = it & 1, aphase = (it >> 1) & 1;
                if (mprobe) { const long long now = clock64(); mc[3] += now - mt; mt = now; }
                tc::mbar_wait(&tmem_empty_bar[ab], aphase ^ 1);
                tc::mbar_wait(small_empty_bar, (it & 1) ^ 1);     // the previous tile's small terms have been folded away
                tc::tc_fence_after();
                if (mprobe) { const long long now = clock64(); mc[0] += now - mt; mt = now; }
                const uint32_t acc = tmem_base + ab * BN;          // main (hi*hi) accumulator
                for (int kb = 0; kb < nkb; ++kb, ++kbc) {
                    if (mprobe) { const long long now = clock64(); mc[3] += now - mt; mt = now; ++mn; }
                    tc::mbar_wait(&b_full[stage], phase);
                    if (mprobe) { const long long now = clock64(); mc[1] += now - mt; mt = now; }
                    const uint32_t sb = tc::smem_u32(smem + stage * C::B_STAGE);
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const uint32_t hbg = 2u * kbc + (uint32_t)h;       // half-block number since the launch began
                        const uint32_t as = hbg % NSLOT, apar = (hbg / NSLOT) & 1u;
                        if (mprobe) { const long long now = clock64(); mc[3] += now - mt; mt = now; }
                        tc::mbar_wait(&a_full[as], apar);
                        tc::tc_fence_after();
                        if (mprobe) { const long long now = clock64(); mc[2] += now - mt; mt = now; }
                        const uint32_t ta = tmem_base + C::A_COL0 + as * C::SLOT_COLS;
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const int kk = 2 * h + j;
                            uint32_t ad[NPL];
                            uint64_t bd[NPL];
#pragma unroll
                            for (int p = 0; p < NPL; ++p) {
                                ad[p] = ta + p * 16 + j * 8;
                                bd[p] = tc::make_desc(sb + p * C::B_PLANE + kk * 2048, 8192, 1024);
                            }
                            const uint32_t first = (kb | kk) != 0 ? 1u : 0u;
                            if constexpr (F16) {
                                umma_bf16_ts(acs, ad[0], bd[1], idesc, first);     // hi * lo'   (both scaled by 2^11)
                                umma_bf16_ts(acs, ad[1], bd[0], idesc, 1u);        // lo' * hi
                                umma_bf16_ts(acc, ad[0], bd[0], idesc, first);     // hi * hi
                            } else {
                                umma_bf16_ts(acs, ad[0], bd[NPL - 1], idesc, first);
                                umma_bf16_ts(acs, ad[NPL - 1], bd[0], idesc, 1u);
                                umma_bf16_ts(acs, ad[1], bd[1], idesc, 1u);
                                umma_bf16_ts(acs, ad[0], bd[1], idesc, 1u);
                                umma_bf16_ts(acs, ad[1], bd[0], idesc, 1u);
                                umma_bf16_ts(acc, ad[0], bd[0], idesc, first);
                            }
                        }
                        tc::umma_commit(&a_empty[as]);       // this half's plane slot may be rewritten
                    }
                    tc::umma_commit(&b_empty[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                tc::umma_commit(&tmem_full_bar[ab]);
            }
            if (mprobe) {
                for (int q = 0; q < 4; ++q) atomicAdd(&x32t_probe[8 + q], (unsigned long long)mc[q]);
                atomicAdd(&x32t_probe[12], (unsigned long long)mn);
            }
        }
    } else if (warp >= NUM_EPI_WARPS + 2 && warp < W_WARP) {
        // ================= converters: one row per lane (the lanes this warp may address in tensor memory) =================
        // NCG groups of four warps; a group's load -> convert -> store chain hides behind the other groups'.
        const int quarter = warp & 3;
        const int grp = (warp - (NUM_EPI_WARPS + 2)) >> 2;
        const int r = quarter * 32 + lane;
        const uint32_t sw = (uint32_t)(r & 7);
        const uint32_t ta0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + C::A_COL0;
        const uint32_t raw_row = tc::smem_u32(raw) + r * 128;
        bool bad = false;                                  // F16: a row of this lane outside the fp16-split range
        const bool probe = ep.probe && blockIdx.x == 0 && warp == NUM_EPI_WARPS + 2;
        long long pc[5] = {0, 0, 0, 0, 0}, pt0 = probe ? clock64() : 0, pt = pt0, pn = 0;
        const long long my_tiles = (long long)blockIdx.x < num_tiles ? (num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
        const uint32_t per_tile = 2u * (uint32_t)nkb;
        const unsigned long long total_hb = (unsigned long long)my_tiles * per_tile;
        uint32_t pos = (uint32_t)grp % per_tile;           // place of the half-block within its tile (a new tile = other rows)
        __half2 amax2 = __floats2half2_rn(0.f, 0.f);       // F16: max |hi| over this lane's share of the current tile's row
        static_assert(NCG <= NSLOT && NCG <= RAW_SLOTS, "a group advances by NCG slots per step");
        uint32_t slot = grp % RAW_SLOTS, rphase = (grp / RAW_SLOTS) & 1u;     // ring slot hbg % RAW_SLOTS and its parity
        uint32_t as = grp % NSLOT, apar = (grp / NSLOT) & 1u;                 // tensor-memory slot hbg % NSLOT and its parity
        for (unsigned long long hbg = grp; hbg < total_hb; hbg += NCG) {
            {
                tc::mbar_wait(&raw_full[slot], rphase);
                if (probe) { const long long now = clock64(); pc[0] += now - pt; pt = now; }
                const uint32_t src = raw_row + slot * X32_RAW_BYTES;
                float4 x[8];
#pragma unroll
                for (int c = 0; c < 8; ++c) x[c] = lds128(src + (((uint32_t)c ^ sw) << 4));
                uint32_t ph[16], pm[16], pl[16];
                if constexpr (F16) {
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        split2h_pair(x[c].x, x[c].y, ph[2 * c], pl[2 * c]);
                        split2h_pair(x[c].z, x[c].w, ph[2 * c + 1], pl[2 * c + 1]);
                        amax2 = __hmax2(amax2, __hmax2(__habs2(*reinterpret_cast<const __half2*>(&ph[2 * c])),
                                                       __habs2(*reinterpret_cast<const __half2*>(&ph[2 * c + 1]))));
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        split3_pair(x[c].x, x[c].y, ph[2 * c], pm[2 * c], pl[2 * c]);
                        split3_pair(x[c].z, x[c].w, ph[2 * c + 1], pm[2 * c + 1], pl[2 * c + 1]);
                    }
                }
                // the MMAs that last read this tensor-memory slot (NSLOT half-blocks ago) must have retired
                const uint32_t ta = ta0 + as * C::SLOT_COLS;
                if (probe) { const long long now = clock64(); pc[1] += now - pt; pt = now; }
                tc::mbar_wait(&a_empty[as], apar ^ 1);
                tc::tc_fence_after();
                if (probe) { const long long now = clock64(); pc[2] += now - pt; pt = now; }
                tmem_st16(ta, ph);
                if constexpr (F16) {
                    tmem_st16(ta + 16, pl);
                } else {
                    tmem_st16(ta + 16, pm);
                    tmem_st16(ta + 32, pl);
                }
                tmem_st_wait();
                if (probe) { const long long now = clock64(); pc[3] += now - pt; pt = now; }
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    tc::mbar_arrive(&raw_empty[slot]);
                    tc::mbar_arrive(&a_full[as]);
                }
                if (probe) { const long long now = clock64(); pc[4] += now - pt; pt = now; ++pn; }
            }
            slot += NCG;
            if (slot >= RAW_SLOTS) { slot -= RAW_SLOTS; rphase ^= 1u; }
            as += NCG;
            if (as >= NSLOT) { as -= NSLOT; apar ^= 1u; }
            pos += NCG;
            if (pos >= per_tile) {                         // this group's next half-block belongs to a later tile
                do pos -= per_tile; while (pos >= per_tile);
                if (F16) bad = bad || f16_range_bad(fmaxf(__low2float(amax2), __high2float(amax2)));
                amax2 = __floats2half2_rn(0.f, 0.f);
            }
        }
        if (F16) bad = bad || f16_range_bad(fmaxf(__low2float(amax2), __high2float(amax2)));
        if (F16 && bad && ep.f16_flag) *ep.f16_flag = 1;
        if (probe && lane == 0) {
            for (int q = 0; q < 5; ++q) atomicAdd(&x32t_probe[q], (unsigned long long)pc[q]);
            atomicAdd(&x32t_probe[5], (unsigned long long)pn);
            atomicAdd(&x32t_probe[6], (unsigned long long)(clock64() - pt0));
        }
    } else {
        // ================= epilogue =================
        const int quarter = warp & 3, chalf = warp >> 2;
        float* head_ws = reinterpret_cast<float*>(smem + C::HEAD_OFF);    // [128][HEAD_MAX_N], zero beyond N / head_n
        float* head_part = head_ws + 128 * HEAD_MAX_N;                     // [2][128][HEAD_MAX_N]
        if (HEAD) {
            for (int i = threadIdx.x; i < 128 * HEAD_MAX_N; i += NUM_EPI_WARPS * 32) {
                const int k = i / HEAD_MAX_N, o = i % HEAD_MAX_N;
                head_ws[i] = (k < N && o < ep.head_n) ? __ldg(ep.head_w + (long long)k * ep.head_n + o) : 0.f;
            }
            asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_WARPS * 32) : "memory");
        }
        uint32_t it = 0;
        for (long long t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
            const int nb = (int)(t % tiles_n);
            const long long mb = t / tiles_n;
            const uint32_t ab = it & 1, aphase = (it >> 1) & 1;
            mbar_wait_idle(&tmem_full_bar[ab], aphase);
            tc::tc_fence_after();
            {   // main += small for this thread's row and column half, then hand the small accumulator to the next tile
                const uint32_t tm = tmem_base + ((uint32_t)(quarter * 32) << 16) + ab * BN + chalf * (BN / 2);
                const uint32_t ts = tmem_base + ((uint32_t)(quarter * 32) << 16) + C::SMALL_COL0 + chalf * (BN / 2);
#pragma unroll
                for (int c = 0; c < BN / 2; c += 32) {
                    uint32_t v[32], w[32];
                    tc::tmem_ld32(tm + c, v);
                    tc::tmem_ld32(ts + c, w);
                    tc::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        v[j] = __float_as_uint(F16 ? fmaf(__uint_as_float(w[j]), F16_LO_UNSCALE, __uint_as_float(v[j]))
                                                   : __uint_as_float(v[j]) + __uint_as_float(w[j]));
                    tmem_st32(tm + c, v);
                }
                tmem_st_wait();
                tc::tc_fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(small_empty_bar);
            }
            epilogue_tile<BN, HEAD, true>(ep, M, N, mb, nb, 0, false, tmem_base, ab, quarter, chalf, lane, head_ws,
                                          head_part + (it & 1) * 128 * HEAD_MAX_N);
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
