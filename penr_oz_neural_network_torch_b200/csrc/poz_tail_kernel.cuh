// The tail of a narrow classifier stack in ONE launch: the last hidden unit's forward, the classifier head
// (last Linear + softmax + cross-entropy + its gradient), the head's backward-x and the hidden unit's
// backward-x -- everything between "the activations of the unit before" and "the gradient handed back to it"
// is row-local, so a block of 8 batch rows can do it without leaving the SM:
//
//   a1 = act1(a0 W1 + b1)                    neural_net_model.py:117-119,172-184        [8 x N1]
//   z2 = a1 W2 + b2;  p = softmax(z2);  cost += -log p[y] / B;  dZ2 = (p - onehot) / B   :188,:400-403, test_torch_backward.py:163-164
//   dZ1 = (dZ2 W2^T) . act1'(a1)             test_torch_backward.py:168-170             [8 x N1]
//   dZ0 = (dZ1 W1^T) . act0'(a0)                                                         [8 x K1]
//
// W1 (up to 256 x 128 fp32, 132 KB) is staged once per block in shared memory with 16-byte cp.async copies into rows
// of pitch N1 + 4 floats: the forward reads it by columns (lanes = consecutive columns), the backward-x by rows with
// 16-byte loads -- eight consecutive rows 528 bytes apart tile the 32 banks, so neither pass has a bank conflict.  At
// batch 1024 this is 128 blocks of 512 threads, one per SM, replacing three launches (7 + 7 + 9 us in the CUDA graph) and the two
// kernel boundaries between them; exact fp32 FMA arithmetic, no tensor cores (0.13 GFLOP in all).
// Optionally the block also leaves what the WEIGHT GRADIENTS of the two units need from these rows: dZ1 as bf16
// planes (B operand of the tensor-core weight-gradient launch, poz_tcs_kernel.cuh), and its eight rows' partial sums
// of [dW2; db2] = [a1 1]^T dZ2 and of db1 = sum dZ1 -- tail_update_kernel adds the blocks' partials in a fixed order and
// applies the optimizer to the head's parameters and to b1, so neither needs a GEMM launch of its own.
// Dropout: the same Philox keys (or recorded masks) as the separate kernels; the hidden unit's two masks are kept in shared
// memory for its backward.  Not here (the step falls back to the separate kernels): diagnostics.
#pragma once
#include "poz_kernels.cuh"
#include "poz_tcs.h"

namespace poz {

constexpr int TAIL_ROWS = 8;
constexpr int TAIL_THREADS = 512;
constexpr int TAIL_CHUNK = 32;            // rows of W1 per asynchronous copy group
constexpr int TAIL_KSPLIT = 8;            // the forward splits K1 over 8 thread groups, the backward-x N1 over 4
constexpr int TAIL_MAX_K1 = 256, TAIL_MAX_N1 = 128, TAIL_MAX_N2 = 32;

struct TailArgs {
    const float* a0;           // recorded output of the unit before the last hidden unit [M][K1]
    const float* w1b;          // [K1 + 1][N1] = [W; b] of the last hidden unit
    const float* w2b;          // [N1 + 1][N2] = [W; b] of the head
    const int* labels;         // int32 labels
    const int* row_idx;        // gather of label rows (nullable)
    float* a1_out;             // recorded output of the last hidden unit [M][N1]
    float* dz2_out;            // gradient of the head's linear output [M][N2]
    float* dz1_out;            // gradient of the last hidden unit's linear output [M][N1]
    float* dz0_out;            // gradient of the linear output of the unit before [M][K1]
    unsigned short* dz0_planes;   // nullable: dz0 also as bf16 planes hi / mid / lo [3][M][ldp] (poz_tcs_kernel.cuh)
    long long dz0_ldp, dz0_pstride;
    unsigned short* dz1_planes;   // nullable: the same for dz1
    long long dz1_ldp, dz1_pstride;
    float* part;                  // nullable: [gridDim.x][part_stride]: [dW2; db2] partial ((N1 + 1) x N2), then db1 partial (N1) at part_b1
    int part_stride, part_b1;
    long long M;
    int K1, N1, N2, has_bias, act1, act0;
    double inv_batch;
    double* cost_base;
    long long cost_cap;
    const DevState* st;
    // dropout points (neural_net_model.py:393-395), as in the separate kernels: on the linear output (z) and on the
    // activation output (a) of the hidden unit, on the head's logits, and -- for the backward-x only -- of the unit before
    DropPoint dz1p, da1p, dz2p, dz0p, da0p;
    int any_drop;
};

// four adjacent values -> their bf16 planes hi / mid / lo, 8-byte stores where the pitch allows
__device__ __forceinline__ void tail_store_planes4(const float4& v, unsigned short* o, long long stride, long long ldp) {
    if ((ldp & 3) == 0 && (stride & 3) == 0) {
        const float x[4] = {v.x, v.y, v.z, v.w};
        unsigned short h[4], md[4], lo[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __nv_bfloat16 bh = __float2bfloat16_rn(x[i]);
            const float r1 = x[i] - __bfloat162float(bh);
            const __nv_bfloat16 bm = __float2bfloat16_rn(r1);
            const __nv_bfloat16 bl = __float2bfloat16_rn(r1 - __bfloat162float(bm));
            h[i] = __bfloat16_as_ushort(bh); md[i] = __bfloat16_as_ushort(bm); lo[i] = __bfloat16_as_ushort(bl);
        }
        *reinterpret_cast<uint2*>(o) = make_uint2(h[0] | ((unsigned)h[1] << 16), h[2] | ((unsigned)h[3] << 16));
        *reinterpret_cast<uint2*>(o + stride) = make_uint2(md[0] | ((unsigned)md[1] << 16), md[2] | ((unsigned)md[3] << 16));
        *reinterpret_cast<uint2*>(o + 2 * stride) = make_uint2(lo[0] | ((unsigned)lo[1] << 16), lo[2] | ((unsigned)lo[3] << 16));
    } else {
        store_planes3<float>(v.x, o, stride); store_planes3<float>(v.y, o + 1, stride);
        store_planes3<float>(v.z, o + 2, stride); store_planes3<float>(v.w, o + 3, stride);
    }
}

// ---- mbarrier / bulk-copy helpers of the W1 staging (the cluster's CTAs share the reads of W1: each fetches a share of
//      the rows and multicasts them into every CTA's shared memory; W1 lies in this GPU's own memory) ----
__device__ __forceinline__ void tail_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void tail_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tail_mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t done = 0;
    long long spins = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (!done && ++spins > (1LL << 26)) __trap();        // a copy that never lands: fail loudly instead of hanging the GPU
    }
}
// one row of W1 (bytes, a multiple of 16) into the same shared-memory offset of every CTA in cta_mask
__device__ __forceinline__ void tail_bulk_row(float* dst_smem, const float* src, uint32_t bytes, uint64_t* bar, uint32_t ncta,
                                              uint16_t cta_mask) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem), b = (uint32_t)__cvta_generic_to_shared(bar);
    if (ncta > 1)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                     ::"r"(d), "l"(src), "r"(bytes), "r"(b), "h"(cta_mask) : "memory");
    else
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(d), "l"(src), "r"(bytes), "r"(b) : "memory");
}
__device__ __forceinline__ void tail_cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void tail_cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

__device__ __forceinline__ void tail_cp_wait(int pending) {      // cp.async.wait_group takes an immediate
    switch (pending) {
        case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
        case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
        case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
        case 3: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
        case 4: asm volatile("cp.async.wait_group 4;" ::: "memory"); break;
        case 5: asm volatile("cp.async.wait_group 5;" ::: "memory"); break;
        case 6: asm volatile("cp.async.wait_group 6;" ::: "memory"); break;
        default: asm volatile("cp.async.wait_group 7;" ::: "memory"); break;
    }
}

__device__ __forceinline__ void tail_cp16(float* dst_smem, const float* src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}

inline size_t tail_smem_bytes(int K1, int N1, int N2) {
    const size_t p1 = N1 + 4, pa = K1 + 4, pn = N1 + 4, p2 = (size_t)N2 | 1;
    const size_t red = (size_t)TAIL_ROWS * (TAIL_KSPLIT * (size_t)N1 > 4 * (size_t)K1 ? TAIL_KSPLIT * (size_t)N1 : 4 * (size_t)K1);
    return (K1 * p1 + N1 + N1 * p2 + 32 + TAIL_ROWS * 32 + TAIL_ROWS * pa + 2 * TAIL_ROWS * pn + red) * sizeof(float) +
           2 * TAIL_ROWS * (size_t)N1;          // keep flags of the hidden unit's two dropout points
}

__global__ void __launch_bounds__(TAIL_THREADS, 1) tail_fused_kernel(const __grid_constant__ TailArgs a) {
    extern __shared__ __align__(16) unsigned char tail_smem[];
    const int K1 = a.K1, N1 = a.N1, N2 = a.N2;
    const int P1 = N1 + 4, PA = K1 + 4, PN = N1 + 4, P2 = N2 | 1;
    float* W1s = reinterpret_cast<float*>(tail_smem);        // [K1][P1]
    float* a0s = W1s + (size_t)K1 * P1;                      // [8][PA]
    float* a1s = a0s + TAIL_ROWS * PA;                       // [8][PN]
    float* dz1s = a1s + TAIL_ROWS * PN;                      // [8][PN]
    float* b1s = dz1s + TAIL_ROWS * PN;                      // [N1]
    float* W2s = b1s + N1;                                   // [N1][P2]
    float* b2s = W2s + (size_t)N1 * P2;                      // [32]
    float* dz2s = b2s + 32;                                  // [8][32]
    float* red = dz2s + TAIL_ROWS * 32;                      // partial sums of the k-split groups: [8 splits][8][N1] / [4][8][K1]
    const int red_elems = TAIL_ROWS * (TAIL_KSPLIT * N1 > 4 * K1 ? TAIL_KSPLIT * N1 : 4 * K1);
    unsigned char* kz1 = reinterpret_cast<unsigned char*>(red + red_elems);   // [8][N1] keep flags: dropout on z1
    unsigned char* ka1 = kz1 + TAIL_ROWS * N1;                                // [8][N1] ... on a1
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const long long m0 = (long long)blockIdx.x * TAIL_ROWS;
    // POZ_KERNEL_TRACE=1, tag 66: operands staged / hidden unit forward done / head done / backward-x done / exit
    fast::KernelTraceRec* ktr = fast::ktrace_open(66, (int)a.M, N1, K1);

    // Index arithmetic below uses shifts and masks over the maxima (K1 <= 256, N1 <= 128, N2 <= 32) with guards, never
    // a division by a run-time size: the first version of this kernel spent 60 % of its issue slots on them.
    // ---- stage W1: bulk copies of whole rows, one mbarrier per chunk of 32 rows, so that the forward below starts on
    //      chunk 0 while the others are still arriving.  The CTAs of a cluster share the reads: each fetches the rows
    //      with (row mod cluster size == its rank) and multicasts them into every CTA's shared memory -- 128 blocks
    //      reading 132 KB each from L2 was 4 us of L2 bandwidth.  Then this block's rows of a0 and the small operands ----
    const int n4 = N1 >> 2, k4 = K1 >> 2;
    const int nchunks = (K1 + TAIL_CHUNK - 1) / TAIL_CHUNK;          // <= 8
    __shared__ __align__(8) uint64_t bars[TAIL_MAX_K1 / TAIL_CHUNK];   // one per chunk, used once (parity 0)
    uint32_t crank, csize;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
    asm("mov.u32 %0, %%cluster_nctarank;" : "=r"(csize));
    if (t == 0) {
        for (int ch = 0; ch < nchunks; ++ch) {
            const int nr = K1 - ch * TAIL_CHUNK < TAIL_CHUNK ? K1 - ch * TAIL_CHUNK : TAIL_CHUNK;
            tail_mbar_init(&bars[ch], 1);
            tail_mbar_expect_tx(&bars[ch], (uint32_t)(nr * N1 * sizeof(float)));   // the whole chunk, whoever fetches it
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    pdl_trigger();                                                    // (programmatic dependent launch: see poz_tcs_kernel.cuh)
    tail_cluster_arrive();                                            // "my barriers are armed"
    tail_cluster_wait();                                              // ... and so are everybody else's: rows may be sent
    if (warp < nchunks) {                                             // warp w sends this CTA's share of chunk w
        const int r = warp * TAIL_CHUNK + lane;
        if (r < K1 && (r & (int)(csize - 1)) == (int)crank)
            tail_bulk_row(W1s + (size_t)r * P1, a.w1b + (size_t)r * N1, (uint32_t)(N1 * sizeof(float)), &bars[warp], csize,
                          (uint16_t)((1u << csize) - 1));
    }
    if (t < N1) b1s[t] = a.has_bias ? a.w1b[(size_t)K1 * N1 + t] : 0.f;
    for (int k = warp; k < N1; k += TAIL_THREADS / 32)
        if (lane < N2) W2s[k * P2 + lane] = a.w2b[k * N2 + lane];
    if (t < 32) b2s[t] = (a.has_bias && t < N2) ? a.w2b[(size_t)N1 * N2 + t] : 0.f;
    const int hrow = warp & (TAIL_ROWS - 1), hhalf = warp >> 3;       // the head's split: row, half of the outputs / of N1
    int y = -1;                                                        // label of this warp's row
    const long long mw = m0 + hrow;
    if (mw < a.M) y = a.labels[a.row_idx ? (long long)a.row_idx[mw] : mw];
    pdl_wait();                                                       // a0 is the launch before's result; nothing above reads it
    {
        const int r = t >> 6, c = t & 63;                            // a0: 8 rows x (up to) 64 16-byte pieces
        if (c < k4) {
            const long long m = m0 + r < a.M ? m0 + r : a.M - 1;
            tail_cp16(a0s + r * PA + c * 4, a.a0 + m * K1 + c * 4);
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");      // a0 rows (cp.async) and the small operands
    __syncthreads();
    if (ktr) ktr->t[1] = fast::ktrace_now();

    // ---- forward of the hidden unit: thread = (four adjacent columns, four of the eight rows, one eighth of each chunk's
    //      k); a warp's 32 lanes read 512 contiguous bytes of a W1 row, the a0 values are broadcasts: 8 loads per 64 FMAs ----
    {
        const int cg = t & 31, rh = (t >> 5) & 1, ks = t >> 6;              // column group, row half, k split
        const int c = cg * 4;
        float acc[4][4];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[r][q] = 0.f;
        const float* ar = a0s + (rh * 4) * PA;
        for (int ch = 0; ch < nchunks; ++ch) {
            tail_mbar_wait(&bars[ch], 0);                                   // chunk ch has landed (all its bytes, from all senders)
            const int k = ch * TAIL_CHUNK + ks * 4;                         // four k of this chunk per split
            if (c < N1 && k < K1) {
                float4 w[4], x[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) w[q] = *reinterpret_cast<const float4*>(W1s + (size_t)(k + q) * P1 + c);
#pragma unroll
                for (int r = 0; r < 4; ++r) x[r] = *reinterpret_cast<const float4*>(ar + r * PA + k);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    acc[r][0] = fmaf(x[r].x, w[0].x, acc[r][0]); acc[r][1] = fmaf(x[r].x, w[0].y, acc[r][1]);
                    acc[r][2] = fmaf(x[r].x, w[0].z, acc[r][2]); acc[r][3] = fmaf(x[r].x, w[0].w, acc[r][3]);
                    acc[r][0] = fmaf(x[r].y, w[1].x, acc[r][0]); acc[r][1] = fmaf(x[r].y, w[1].y, acc[r][1]);
                    acc[r][2] = fmaf(x[r].y, w[1].z, acc[r][2]); acc[r][3] = fmaf(x[r].y, w[1].w, acc[r][3]);
                    acc[r][0] = fmaf(x[r].z, w[2].x, acc[r][0]); acc[r][1] = fmaf(x[r].z, w[2].y, acc[r][1]);
                    acc[r][2] = fmaf(x[r].z, w[2].z, acc[r][2]); acc[r][3] = fmaf(x[r].z, w[2].w, acc[r][3]);
                    acc[r][0] = fmaf(x[r].w, w[3].x, acc[r][0]); acc[r][1] = fmaf(x[r].w, w[3].y, acc[r][1]);
                    acc[r][2] = fmaf(x[r].w, w[3].z, acc[r][2]); acc[r][3] = fmaf(x[r].w, w[3].w, acc[r][3]);
                }
            }
        }
        tail_cluster_arrive();                                              // (waited for at the end: no CTA leaves the cluster early)
        if (c < N1) {
#pragma unroll
            for (int r = 0; r < 4; ++r)
                *reinterpret_cast<float4*>(red + ((size_t)ks * TAIL_ROWS + rh * 4 + r) * N1 + c) =
                    make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
        }
        __syncthreads();
        if (t < TAIL_ROWS * 32) {                                           // 8 rows x (up to) 32 groups of four columns
            const int row = t >> 5, c4 = (t & 31) * 4;
            if (c4 < N1) {
                float4 z4 = *reinterpret_cast<const float4*>(b1s + c4);
#pragma unroll
                for (int sidx = 0; sidx < TAIL_KSPLIT; ++sidx) {
                    const float4 pq = *reinterpret_cast<const float4*>(red + ((size_t)sidx * TAIL_ROWS + row) * N1 + c4);
                    z4.x += pq.x; z4.y += pq.y; z4.z += pq.z; z4.w += pq.w;
                }
                float z[4] = {z4.x, z4.y, z4.z, z4.w}, v[4];
                if (a.dz1p.enabled) {
                    bool keep[4];
                    drop_keep4(a.dz1p, a.st, m0 + row, c4, N1, keep);
#pragma unroll
                    for (int i = 0; i < 4; ++i) { kz1[row * N1 + c4 + i] = keep[i]; z[i] = keep[i] ? z[i] * (float)a.dz1p.scale : 0.f; }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) v[i] = act_fwd<float>(a.act1, z[i]);
                if (a.da1p.enabled) {
                    bool keep[4];
                    drop_keep4(a.da1p, a.st, m0 + row, c4, N1, keep);
#pragma unroll
                    for (int i = 0; i < 4; ++i) { ka1[row * N1 + c4 + i] = keep[i]; v[i] = keep[i] ? v[i] * (float)a.da1p.scale : 0.f; }
                }
                const float4 v4 = make_float4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<float4*>(a1s + row * PN + c4) = v4;
                if (m0 + row < a.M) *reinterpret_cast<float4*>(a.a1_out + (m0 + row) * N1 + c4) = v4;
            }
        }
    }
    __syncthreads();
    if (ktr) ktr->t[2] = fast::ktrace_now();

    // ---- head.  Two warps per row.  Logits: each lane keeps four values of the row's a1, the warp sums one output at a
    //      time (half of the outputs per warp); then both warps of a row do the softmax, and each takes half of
    //      dZ1 = (dZ2 W2^T) . act1'(a1) ----
    double local = 0.0;
    float* zs = red;                                                        // [8][32] logits
    {
        const float* arow = a1s + hrow * PN;
        float av[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) av[j] = lane + 32 * j < N1 ? arow[lane + 32 * j] : 0.f;
        const int nh = (N2 + 1) >> 1;
        const int nb = hhalf * nh, ne = nb + nh < N2 ? nb + nh : N2;
        for (int n = nb; n < ne; ++n) {
            float pz = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (lane + 32 * j < N1) pz = fmaf(av[j], W2s[(lane + 32 * j) * P2 + n], pz);
            pz = warp_sum(pz);
            if (lane == 0) zs[hrow * 32 + n] = pz + b2s[n];
        }
        __syncthreads();
        float g = 0.f;
        if (mw < a.M) {
            float z = lane < N2 ? zs[hrow * 32 + lane] : -INFINITY;
            bool keep2 = true;                                              // dropout on the logits (and back through it)
            if (a.dz2p.enabled && lane < N2) {
                keep2 = drop_keep(a.dz2p, a.st, mw, lane, N2);
                z = keep2 ? z * (float)a.dz2p.scale : 0.f;
            }
            const float mx = warp_max(z);
            const float e = lane < N2 ? expf(z - mx) : 0.f;
            const float ssum = warp_sum(e);
            const float pr = e / ssum;
            g = lane < N2 ? (pr - (lane == y ? 1.f : 0.f)) * (float)a.inv_batch : 0.f;
            if (a.dz2p.enabled) g = keep2 ? g * (float)a.dz2p.scale : 0.f;
            const float zy = __shfl_sync(0xffffffffu, z, (y >= 0 && y < N2) ? y : 0);
            if (hhalf == 0) {
                if (lane < N2) a.dz2_out[mw * N2 + lane] = g;
                if (lane == 0 && y >= 0 && y < N2) local += (double)(mx + logf(ssum) - zy);
            }
        }
        if (hhalf == 0) dz2s[hrow * 32 + lane] = g;
        float dd[2] = {0.f, 0.f};
        const int kb = hhalf * 64 + lane;                                   // this warp's k: kb, kb + 32
        for (int n = 0; n < N2; ++n) {
            const float gn = __shfl_sync(0xffffffffu, g, n);
#pragma unroll
            for (int j = 0; j < 2; ++j)
                if (kb + 32 * j < N1) dd[j] = fmaf(gn, W2s[(kb + 32 * j) * P2 + n], dd[j]);
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int k = kb + 32 * j;
            if (k < N1) {                                                    // (zero for rows past M)
                float ga = dd[j], av = arow[k];
                if (a.da1p.enabled) {
                    ga = ka1[hrow * N1 + k] ? ga * (float)a.da1p.scale : 0.f;
                    av = av * (float)a.da1p.keep_prob;
                }
                float gz = ga * act_bwd_from_out<float>(a.act1, av);
                if (a.dz1p.enabled) gz = kz1[hrow * N1 + k] ? gz * (float)a.dz1p.scale : 0.f;
                dz1s[hrow * PN + k] = gz;
            }
        }
    }
    __syncthreads();
    if (ktr) ktr->t[3] = fast::ktrace_now();

    // dZ1 to global memory, four values per thread: fp32 and (for the weight-gradient launch) the three bf16 planes
    if (t < TAIL_ROWS * 32) {
        const int r = t >> 5, k = lane * 4;
        const long long m = m0 + r;
        if (k < N1 && m < a.M) {
            const float4 v = *reinterpret_cast<const float4*>(dz1s + r * PN + k);
            *reinterpret_cast<float4*>(a.dz1_out + m * N1 + k) = v;
            if (a.dz1_planes) tail_store_planes4(v, a.dz1_planes + m * a.dz1_ldp + k, a.dz1_pstride, a.dz1_ldp);
        }
    }

    // ---- backward-x of the hidden unit: thread = (input features j and j + 128, one quarter of N1), all eight rows: 10
    //      shared-memory loads (8 of them broadcasts) per 64 FMAs; the quarters meet in shared memory ----
    if (a.dz0_out) {
        const int j = t & 127, kq = t >> 7;
        const int nq = ((n4 + 3) >> 2) * 4;
        const int k0 = kq * nq, k1 = k0 + nq < N1 ? k0 + nq : N1;
        float acc[2][TAIL_ROWS];
#pragma unroll
        for (int r = 0; r < TAIL_ROWS; ++r) acc[0][r] = acc[1][r] = 0.f;
        const bool h0 = j < K1, h1 = j + 128 < K1;
        const float* wr0 = W1s + (size_t)(h0 ? j : 0) * P1;
        const float* wr1 = W1s + (size_t)(h1 ? j + 128 : 0) * P1;
#pragma unroll 2
        for (int k = k0; k < k1; k += 4) {
            const float4 w0 = *reinterpret_cast<const float4*>(wr0 + k);
            const float4 w1 = *reinterpret_cast<const float4*>(wr1 + k);
#pragma unroll
            for (int r = 0; r < TAIL_ROWS; ++r) {
                const float4 d = *reinterpret_cast<const float4*>(dz1s + r * PN + k);
                acc[0][r] = fmaf(d.x, w0.x, acc[0][r]); acc[1][r] = fmaf(d.x, w1.x, acc[1][r]);
                acc[0][r] = fmaf(d.y, w0.y, acc[0][r]); acc[1][r] = fmaf(d.y, w1.y, acc[1][r]);
                acc[0][r] = fmaf(d.z, w0.z, acc[0][r]); acc[1][r] = fmaf(d.z, w1.z, acc[1][r]);
                acc[0][r] = fmaf(d.w, w0.w, acc[0][r]); acc[1][r] = fmaf(d.w, w1.w, acc[1][r]);
            }
        }
#pragma unroll
        for (int r = 0; r < TAIL_ROWS; ++r) {
            if (h0) red[((size_t)kq * TAIL_ROWS + r) * K1 + j] = acc[0][r];
            if (h1) red[((size_t)kq * TAIL_ROWS + r) * K1 + j + 128] = acc[1][r];
        }
        __syncthreads();
        const int r = t >> 6, jj = (t & 63) * 4;                            // 8 rows x (up to) 64 groups of four
        const long long m = m0 + r;
        if (jj < K1 && m < a.M) {
            float4 sm = *reinterpret_cast<const float4*>(red + (size_t)r * K1 + jj);
#pragma unroll
            for (int q = 1; q < 4; ++q) {
                const float4 sq = *reinterpret_cast<const float4*>(red + ((size_t)q * TAIL_ROWS + r) * K1 + jj);
                sm.x += sq.x; sm.y += sq.y; sm.z += sq.z; sm.w += sq.w;
            }
            const float4 o = *reinterpret_cast<const float4*>(a0s + r * PA + jj);
            float4 gz;
            if (!a.any_drop) {
                gz.x = sm.x * act_bwd_from_out<float>(a.act0, o.x);
                gz.y = sm.y * act_bwd_from_out<float>(a.act0, o.y);
                gz.z = sm.z * act_bwd_from_out<float>(a.act0, o.z);
                gz.w = sm.w * act_bwd_from_out<float>(a.act0, o.w);
            } else {
                const float dv[4] = {sm.x, sm.y, sm.z, sm.w}, ov[4] = {o.x, o.y, o.z, o.w};
                float gv[4];
                bool ka0[4] = {true, true, true, true}, kz0[4] = {true, true, true, true};
                if (a.da0p.enabled) drop_keep4(a.da0p, a.st, m, jj, K1, ka0);
                if (a.dz0p.enabled) drop_keep4(a.dz0p, a.st, m, jj, K1, kz0);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float ga = dv[i], av = ov[i];
                    if (a.da0p.enabled) {
                        ga = ka0[i] ? ga * (float)a.da0p.scale : 0.f;
                        av = av * (float)a.da0p.keep_prob;
                    }
                    float gq = ga * act_bwd_from_out<float>(a.act0, av);
                    if (a.dz0p.enabled) gq = kz0[i] ? gq * (float)a.dz0p.scale : 0.f;
                    gv[i] = gq;
                }
                gz = make_float4(gv[0], gv[1], gv[2], gv[3]);
            }
            *reinterpret_cast<float4*>(a.dz0_out + m * K1 + jj) = gz;
            if (a.dz0_planes) tail_store_planes4(gz, a.dz0_planes + m * a.dz0_ldp + jj, a.dz0_pstride, a.dz0_ldp);
        }
    }
    // ---- this block's share of [dW2; db2] = [a1 1]^T dZ2 and of db1 = column sums of dZ1 ----
    if (a.part) {
        float* prt = a.part + (size_t)blockIdx.x * a.part_stride;
        if (lane < N2) {
            float g2[TAIL_ROWS];                                            // this lane's column of dZ2
#pragma unroll
            for (int r = 0; r < TAIL_ROWS; ++r) g2[r] = dz2s[r * 32 + lane];
            for (int k = warp; k <= N1; k += TAIL_THREADS / 32) {           // a row of [dW2; db2] per warp
                float sacc = 0.f;
                if (k < N1) {
#pragma unroll
                    for (int r = 0; r < TAIL_ROWS; ++r) sacc = fmaf(a1s[r * PN + k], g2[r], sacc);
                } else {
#pragma unroll
                    for (int r = 0; r < TAIL_ROWS; ++r) sacc = fmaf(1.f, g2[r], sacc);
                }
                prt[k * N2 + lane] = sacc;
            }
        }
        if (t < N1) {
            float sacc = 0.f;
#pragma unroll
            for (int r = 0; r < TAIL_ROWS; ++r) sacc += dz1s[r * PN + t];
            prt[a.part_b1 + t] = sacc;
        }
    }
    if (ktr) ktr->t[4] = fast::ktrace_now();
    tail_cluster_wait();
    if (a.cost_base) block_add_double(local * a.inv_batch, cost_slot(a.cost_base, a.cost_cap, a.st));
    if (ktr) ktr->t[5] = fast::ktrace_now();
}

// The blocks' partial sums of tail_fused_kernel, added in a fixed order, then the optimizer (tcs::UpdateEpi: update_kernel's
// arithmetic) for the head's [W2; b2] and for b1.  A block of 8 warps finishes 32 parameters: lane = parameter, warp w adds
// the partials w, w + 8, ... (four loads in flight per thread), warp 0 adds the eight sums and updates.  (One thread per
// parameter walking all 128 partials took 13.7 us under ncu -- as long as the weight-gradient launch it runs beside.)
constexpr int TAIL_UPDATE_PARAMS = 32;
__global__ void __launch_bounds__(256) tail_update_kernel(const float* __restrict__ part, int nparts, int stride, int part_b1,
                                                          int n_w2, int N2, tcs::UpdateEpi e2, int n_b1, int K1, tcs::UpdateEpi e1) {
    __shared__ float sums[8][TAIL_UPDATE_PARAMS];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i = blockIdx.x * TAIL_UPDATE_PARAMS + lane;
    float g = 0.f;
    if (i < n_w2 + n_b1) {
        const float* src = part + (i < n_w2 ? i : part_b1 + (i - n_w2));
        float g0 = 0.f, g1 = 0.f, g2 = 0.f, g3 = 0.f;
        int q = w;
        for (; q + 24 < nparts; q += 32) {
            g0 += src[(size_t)q * stride]; g1 += src[(size_t)(q + 8) * stride];
            g2 += src[(size_t)(q + 16) * stride]; g3 += src[(size_t)(q + 24) * stride];
        }
        for (; q < nparts; q += 8) g0 += src[(size_t)q * stride];
        g = (g0 + g1) + (g2 + g3);
    }
    sums[w][lane] = g;
    __syncthreads();
    tcs::UpdateEpi::State s2, s1;
    if (w == 0 && i < n_w2 + n_b1) {
        float t = sums[0][lane];
#pragma unroll
        for (int k = 1; k < 8; ++k) t += sums[k][lane];
        const float gv[1] = {t};
        if (i < n_w2) e2.row<1>(s2, i / N2, i % N2, gv, 1, -1);
        else e1.row<1>(s1, K1, i - n_w2, gv, 1, -1);
    }
    e2.finish(s2);
}

}  // namespace poz
