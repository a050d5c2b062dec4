// The last hidden Linear unit and the classifier head in one launch (float32 training step).
//
// In the step's timeline the forward of the last hidden unit is a 6 us cluster GEMM
// followed by a ~2.5 us kernel boundary and the 7 us head kernel, which itself spends its first 2 us staging its
// weights.  Rows are independent through both, so here every warp carries one batch row through
//     a1 = act(drop(a0 . W1 + b1))     (forward of unit p = last - 1, neural_net_model.py:117-119,172-188,393-395)
// and then through head_ce_row (last Linear + softmax / cross-entropy + its backward-x).  W1 is streamed through
// shared memory in 32-row k-tiles by the whole block (cp.async, three stages) while the head's weights are staged
// once; for the product every warp owns an eighth of the columns for all eight rows of the block.
// Limits (host-checked): N1 (width of unit p) a multiple of 32 and <= 256; everything fits in 96 KB of shared memory.
#pragma once
#include "poz_simt_fast.cuh"
#include "poz_kernels.cuh"

namespace poz {

constexpr int FH_KT = 32;          // k-rows of W1 per stage
constexpr int FH_STAGES = 3;
constexpr int FH_WARPS = 8;

struct FwdHeadArgs {
    HeadArgs<float> h;             // h.a_prev is not read (the row comes from this kernel); h.K == N1
    const float* a_in;             // input of unit p: [M][K1] (x when p == 0)
    const int* in_row_idx;         // gather of a_in rows (only when p == 0)
    const float* w1b;              // [K1 (+1)][N1]
    float* a1_out;                 // recorded output of unit p [M][N1]
    float* z1_rec;                 // diagnostics: recorded linear output of unit p (nullable)
    int K1, N1, act1;
    DropPoint dz1, da1;            // dropout on the linear output and on the activation output of unit p
};

__global__ void __launch_bounds__(FH_WARPS * 32) fwd_head_kernel(const __grid_constant__ FwdHeadArgs a) {
    extern __shared__ __align__(16) unsigned char fh_smem[];
    const int K1 = a.K1, N1 = a.N1, N = a.h.N;
    const int Ns = N | 1;
    // carve-up: head weights [N1][Ns] | head bias [32] | rows [8][N1] | inputs [8][K1] | W1 stages [3][32][N1]
    float* W2 = reinterpret_cast<float*>(fh_smem);
    float* bias2 = W2 + (size_t)N1 * Ns;
    float* rows = bias2 + 32;
    size_t off = (size_t)N1 * Ns + 32 + (size_t)FH_WARPS * N1;
    off = (off + 3) & ~(size_t)3;                                      // 16-byte alignment for the 128-bit reads below
    float* xin = reinterpret_cast<float*>(fh_smem) + off;
    const int K1p = ((K1 + FH_KT - 1) / FH_KT) * FH_KT + 1;            // input rows padded with zeros to whole k-tiles; odd pitch:
                                                                       // eight rows are read per k and must sit in different banks
    off += (size_t)FH_WARPS * K1p;
    off = (off + 3) & ~(size_t)3;
    float* w1s = reinterpret_cast<float*>(fh_smem) + off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float* arow = rows + (size_t)warp * N1;
    float* xrow = xin + (size_t)warp * K1p;
    fast::KernelTraceRec* ktr = fast::ktrace_open(65, (int)a.h.M, N, N1);   // trace tag 65: forward + head

    pdl_trigger();
    pdl_wait();
    // head weights, staged once per block
    for (int i = tid; i < N1 * N; i += blockDim.x) W2[(i / N) * Ns + (i % N)] = a.h.wb[i];
    for (int i = tid; i < 32; i += blockDim.x) bias2[i] = (a.h.has_bias && i < N) ? a.h.wb[(long long)N1 * N + i] : 0.f;

    const int nkt = (K1 + FH_KT - 1) / FH_KT;
    const int chunks_per_row = N1 / 4;                                 // 16-byte chunks of a W1 row
    // Every block walks the k-tiles in its own rotation: all blocks read the same W1, and when they all ask for the
    // same tile at the same time the L2 serves those identical lines one requester after the other.
    const int rot = (int)(blockIdx.x % (unsigned)nkt);
    auto issue_w1 = [&](int kt) {
        float* dst = w1s + (size_t)(kt % FH_STAGES) * FH_KT * N1;
        const int k0 = ((kt + rot) % nkt) * FH_KT;
        for (int c = tid; c < FH_KT * chunks_per_row; c += blockDim.x) {
            const int r = c / chunks_per_row, cc = (c - r * chunks_per_row) * 4;
            const bool ok = k0 + r < K1;
            fast::cp_async16(dst + r * N1 + cc, ok ? a.w1b + (long long)(k0 + r) * N1 + cc : a.w1b, ok ? 16 : 0);
        }
    };
    const float invb = (float)a.h.inv_batch;
    double local = 0.0;
    const long long groups = (a.h.M + FH_WARPS - 1) / FH_WARPS;
    bool first = true;
    for (long long grp = blockIdx.x; grp < groups; grp += gridDim.x, first = false) {
        const long long m = grp * FH_WARPS + warp;
        const bool live = m < a.h.M;
        // this warp's input row and label
        int y = -1;
        if (live) {
            const float* src = a.a_in + (a.in_row_idx ? (long long)a.in_row_idx[m] : m) * K1;
            for (int k = lane; k < K1p; k += 32) xrow[k] = k < K1 ? src[k] : 0.f;
            y = a.h.labels[a.h.row_idx ? (long long)a.h.row_idx[m] : m];
        }
        // a1 = x . W1 for the block's eight rows.  Warp w owns columns [w N1/8, (w+1) N1/8); lane = (row r = lane / 4,
        // column group cg = lane % 4): per k one 128-bit read of the W1 tile row (4 addresses per warp, broadcast to the
        // rows) and one read of x[r][k], so every W1 element leaves shared memory once per warp-column-slice, not once
        // per row (a row-per-warp version was shared-memory bound: 13 us for this phase).
        const int r8 = lane >> 2, cg = lane & 3;
        const int ncg = N1 >> 5;                                       // 4-column groups per warp (N1 % 32 == 0)
        const float* xr = xin + (size_t)r8 * K1p;
        float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        for (int s = 0; s < FH_STAGES - 1; ++s) {
            if (s < nkt) issue_w1(s);
            fast::cp_commit();
        }
        for (int kt = 0; kt < nkt; ++kt) {
            fast::cp_wait<FH_STAGES - 2>();
            __syncthreads();                                           // tile kt landed; everyone is done with tile kt - 1 (and the input rows are written)
            if (kt + FH_STAGES - 1 < nkt) issue_w1(kt + FH_STAGES - 1);
            fast::cp_commit();
            const float* wt = w1s + (size_t)(kt % FH_STAGES) * FH_KT * N1 + warp * (N1 >> 3);
            const int k0 = ((kt + rot) % nkt) * FH_KT;
            // (rows of the tile beyond K1 are zero-filled by the copy and the input rows are zero-padded, so the
            // whole tile is always multiplied: a fixed trip count lets the reads of eight k-values run ahead)
#pragma unroll 8
            for (int kk = 0; kk < FH_KT; ++kk) {
                const float xv = xr[k0 + kk];
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int g = cg + 4 * j;
                    if (g < ncg) {
                        const float4 w = *reinterpret_cast<const float4*>(wt + kk * N1 + 4 * g);
                        acc[j][0] = fmaf(xv, w.x, acc[j][0]); acc[j][1] = fmaf(xv, w.y, acc[j][1]);
                        acc[j][2] = fmaf(xv, w.z, acc[j][2]); acc[j][3] = fmaf(xv, w.w, acc[j][3]);
                    }
                }
            }
        }
        fast::cp_wait<0>();
        if (ktr && first) ktr->t[1] = fast::ktrace_now();             // hidden unit's products done
        // bias, dropout, activation, dropout; the recorded output goes to global memory and to the row buffer of the
        // warp that will run the head for row r
        {
            const long long mr = grp * FH_WARPS + r8;
            if (mr < a.h.M) {
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int g = cg + 4 * j;
                    if (g < ncg) {
                        const int c = warp * (N1 >> 3) + 4 * g;
                        float zv[4], av[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            float v = acc[j][q] + (a.h.has_bias ? a.w1b[(long long)K1 * N1 + c + q] : 0.f);
                            if (a.dz1.enabled) v = drop_keep(a.dz1, a.h.st, mr, c + q, N1) ? v * (float)a.dz1.scale : 0.f;
                            zv[q] = v;
                            float w = act_fwd<float>(a.act1, v);
                            if (a.da1.enabled) w = drop_keep(a.da1, a.h.st, mr, c + q, N1) ? w * (float)a.da1.scale : 0.f;
                            av[q] = w;
                        }
                        if (a.z1_rec) *reinterpret_cast<float4*>(a.z1_rec + mr * N1 + c) = make_float4(zv[0], zv[1], zv[2], zv[3]);
                        *reinterpret_cast<float4*>(a.a1_out + mr * N1 + c) = make_float4(av[0], av[1], av[2], av[3]);
                        *reinterpret_cast<float4*>(rows + (size_t)r8 * N1 + c) = make_float4(av[0], av[1], av[2], av[3]);
                    }
                }
            }
        }
        __syncthreads();                                               // head weights staged (first pass); rows complete
        if (live) head_ce_row<float>(a.h, W2, Ns, bias2, arow, m, y, lane, invb, local, ktr, first);
        __syncthreads();                                               // the next group overwrites xrow / arow / W1 stages
    }
    if (ktr) ktr->t[4] = fast::ktrace_now();
    if (a.h.cost_base) block_add_double(local * a.h.inv_batch, cost_slot(a.h.cost_base, a.h.cost_cap, a.h.st));
    if (ktr) ktr->t[5] = fast::ktrace_now();
}

}  // namespace poz
