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
// batch 1024 this is 128 blocks, one per SM, replacing three launches (7 + 7 + 9 us in the CUDA graph) and the two
// kernel boundaries between them; exact fp32 FMA arithmetic, no tensor cores (0.13 GFLOP in all).
// Optionally the block also leaves what the WEIGHT GRADIENTS of the two units need from these rows: dZ1 as bf16
// planes (B operand of the tensor-core weight-gradient launch, poz_tcs_kernel.cuh), and its eight rows' partial sums
// of [dW2; db2] = [a1 1]^T dZ2 and of db1 = sum dZ1 -- tail_update_kernel adds the blocks' partials in a fixed order and
// applies the optimizer to the head's parameters and to b1, so neither needs a GEMM launch of its own.
// Not here (the step falls back to the separate kernels): dropout, recorded masks, diagnostics.
#pragma once
#include "poz_kernels.cuh"
#include "poz_tcs.h"

namespace poz {

constexpr int TAIL_ROWS = 8;
constexpr int TAIL_THREADS = 256;
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
};

__device__ __forceinline__ void tail_cp16(float* dst_smem, const float* src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst_smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}

inline size_t tail_smem_bytes(int K1, int N1, int N2) {
    const size_t p1 = N1 + 4, pa = K1 + 4, pn = N1 + 4, p2 = (size_t)N2 | 1;
    return (K1 * p1 + N1 + N1 * p2 + 32 + TAIL_ROWS * 32 + TAIL_ROWS * pa + 2 * TAIL_ROWS * pn) * sizeof(float);
}

__global__ void __launch_bounds__(TAIL_THREADS, 1) tail_fused_kernel(TailArgs a) {
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
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const long long m0 = (long long)blockIdx.x * TAIL_ROWS;
    // POZ_KERNEL_TRACE=1, tag 66: operands staged / hidden unit forward done / head done / backward-x done / exit
    fast::KernelTraceRec* ktr = fast::ktrace_open(66, (int)a.M, N1, K1);

    // ---- stage W1 and this block's rows of a0 (16-byte asynchronous copies), then the small operands ----
    const int n4 = N1 / 4, k4 = K1 / 4;
    for (int i = t; i < K1 * n4; i += TAIL_THREADS) {
        const int r = i / n4, c = (i - r * n4) * 4;
        tail_cp16(W1s + (size_t)r * P1 + c, a.w1b + (size_t)r * N1 + c);
    }
    for (int i = t; i < TAIL_ROWS * k4; i += TAIL_THREADS) {
        const int r = i / k4, c = (i - r * k4) * 4;
        const long long m = m0 + r < a.M ? m0 + r : a.M - 1;
        tail_cp16(a0s + r * PA + c, a.a0 + m * K1 + c);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (int i = t; i < N1; i += TAIL_THREADS) b1s[i] = a.has_bias ? a.w1b[(size_t)K1 * N1 + i] : 0.f;
    for (int i = t; i < N1 * N2; i += TAIL_THREADS) W2s[(i / N2) * P2 + (i % N2)] = a.w2b[i];
    for (int i = t; i < 32; i += TAIL_THREADS) b2s[i] = (a.has_bias && i < N2) ? a.w2b[(size_t)N1 * N2 + i] : 0.f;
    int y = -1;                                              // label of this warp's row
    const long long mw = m0 + warp;
    if (mw < a.M) y = a.labels[a.row_idx ? (long long)a.row_idx[mw] : mw];
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (ktr) ktr->t[1] = fast::ktrace_now();

    // ---- forward of the hidden unit: thread = (column c, four of the eight rows) ----
    {
        const int c = t & 127, g = t >> 7;
        if (c < N1) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            const float* ar = a0s + (g * 4) * PA;
#pragma unroll 2
            for (int k = 0; k < K1; k += 4) {
                const float w0 = W1s[(size_t)k * P1 + c], w1 = W1s[(size_t)(k + 1) * P1 + c];
                const float w2 = W1s[(size_t)(k + 2) * P1 + c], w3 = W1s[(size_t)(k + 3) * P1 + c];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float4 x = *reinterpret_cast<const float4*>(ar + r * PA + k);
                    acc[r] = fmaf(x.x, w0, acc[r]);
                    acc[r] = fmaf(x.y, w1, acc[r]);
                    acc[r] = fmaf(x.z, w2, acc[r]);
                    acc[r] = fmaf(x.w, w3, acc[r]);
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int row = g * 4 + r;
                const float v = act_fwd<float>(a.act1, acc[r] + b1s[c]);
                a1s[row * PN + c] = v;
                if (m0 + row < a.M) a.a1_out[(m0 + row) * N1 + c] = v;
            }
        }
    }
    __syncthreads();
    if (ktr) ktr->t[2] = fast::ktrace_now();

    // ---- head: one warp per row (logits, softmax, loss, dZ2), then dZ1 = (dZ2 W2^T) . act1'(a1) ----
    double local = 0.0;
    if (mw < a.M) {
        const float* arow = a1s + warp * PN;
        float z = b2s[lane];
        if (lane < N2) {
            float z1 = 0.f, z2 = 0.f, z3 = 0.f;
            for (int k = 0; k < N1; k += 4) {
                z = fmaf(arow[k], W2s[k * P2 + lane], z);
                z1 = fmaf(arow[k + 1], W2s[(k + 1) * P2 + lane], z1);
                z2 = fmaf(arow[k + 2], W2s[(k + 2) * P2 + lane], z2);
                z3 = fmaf(arow[k + 3], W2s[(k + 3) * P2 + lane], z3);
            }
            z = (z + z1) + (z2 + z3);
        }
        const float zm = lane < N2 ? z : -INFINITY;
        const float mx = warp_max(zm);
        const float e = lane < N2 ? expf(z - mx) : 0.f;
        const float s = warp_sum(e);
        const float p = e / s;
        const float g = lane < N2 ? (p - (lane == y ? 1.f : 0.f)) * (float)a.inv_batch : 0.f;
        if (lane < N2) a.dz2_out[mw * N2 + lane] = g;
        dz2s[warp * 32 + lane] = g;
        const float zy = __shfl_sync(0xffffffffu, z, (y >= 0 && y < N2) ? y : 0);
        if (lane == 0 && y >= 0 && y < N2) local += (double)(mx + logf(s) - zy);
        for (int kg = 0; kg < N1; kg += 128) {
            float dd[4] = {0.f, 0.f, 0.f, 0.f};
            for (int n = 0; n < N2; ++n) {
                const float gn = __shfl_sync(0xffffffffu, g, n);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int kk = kg + 32 * j + lane;
                    if (kk < N1) dd[j] = fmaf(gn, W2s[kk * P2 + n], dd[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int k = kg + 32 * j + lane;
                if (k < N1) {
                    const float gz = dd[j] * act_bwd_from_out<float>(a.act1, arow[k]);
                    dz1s[warp * PN + k] = gz;
                    a.dz1_out[mw * N1 + k] = gz;
                    if (a.dz1_planes) store_planes3<float>(gz, a.dz1_planes + mw * a.dz1_ldp + k, a.dz1_pstride);
                }
            }
        }
    } else {
        for (int k = lane; k < N1; k += 32) dz1s[warp * PN + k] = 0.f;
        dz2s[warp * 32 + lane] = 0.f;
    }
    __syncthreads();
    if (ktr) ktr->t[3] = fast::ktrace_now();

    // ---- backward-x of the hidden unit: thread = input feature j, all eight rows ----
    if (a.dz0_out && t < K1) {
        const int j = t;
        float acc[TAIL_ROWS];
#pragma unroll
        for (int r = 0; r < TAIL_ROWS; ++r) acc[r] = 0.f;
        const float* wr = W1s + (size_t)j * P1;
#pragma unroll 2
        for (int k = 0; k < N1; k += 4) {
            const float4 w = *reinterpret_cast<const float4*>(wr + k);
#pragma unroll
            for (int r = 0; r < TAIL_ROWS; ++r) {
                const float4 d = *reinterpret_cast<const float4*>(dz1s + r * PN + k);
                acc[r] = fmaf(d.x, w.x, acc[r]);
                acc[r] = fmaf(d.y, w.y, acc[r]);
                acc[r] = fmaf(d.z, w.z, acc[r]);
                acc[r] = fmaf(d.w, w.w, acc[r]);
            }
        }
#pragma unroll
        for (int r = 0; r < TAIL_ROWS; ++r) {
            const long long m = m0 + r;
            if (m < a.M) {
                const float gz = acc[r] * act_bwd_from_out<float>(a.act0, a0s[r * PA + j]);
                a.dz0_out[m * K1 + j] = gz;
                if (a.dz0_planes) store_planes3<float>(gz, a.dz0_planes + m * a.dz0_ldp + j, a.dz0_pstride);
            }
        }
    }
    // ---- this block's share of [dW2; db2] = [a1 1]^T dZ2 and of db1 = column sums of dZ1 ----
    if (a.part) {
        float* prt = a.part + (size_t)blockIdx.x * a.part_stride;
        for (int i = t; i < (N1 + 1) * N2; i += TAIL_THREADS) {
            const int k = i / N2, n = i - k * N2;
            float sacc = 0.f;
#pragma unroll
            for (int r = 0; r < TAIL_ROWS; ++r) sacc = fmaf(k < N1 ? a1s[r * PN + k] : 1.f, dz2s[r * 32 + n], sacc);
            prt[i] = sacc;
        }
        for (int k = t; k < N1; k += TAIL_THREADS) {
            float sacc = 0.f;
#pragma unroll
            for (int r = 0; r < TAIL_ROWS; ++r) sacc += dz1s[r * PN + k];
            prt[a.part_b1 + k] = sacc;
        }
    }
    if (ktr) ktr->t[4] = fast::ktrace_now();
    if (a.cost_base) block_add_double(local * a.inv_batch, cost_slot(a.cost_base, a.cost_cap, a.st));
    if (ktr) ktr->t[5] = fast::ktrace_now();
}

// The blocks' partial sums of tail_fused_kernel, added in a fixed order, then the optimizer (tcs::UpdateEpi: update_kernel's
// arithmetic) for the head's [W2; b2] and for b1.  One thread per parameter: (N1 + 1) N2 + N1 of them.
__global__ void __launch_bounds__(256) tail_update_kernel(const float* __restrict__ part, int nparts, int stride, int part_b1,
                                                          int n_w2, int N2, tcs::UpdateEpi e2, int n_b1, int K1, tcs::UpdateEpi e1) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    tcs::UpdateEpi::State s2, s1;
    if (i < n_w2 + n_b1) {
        const float* src = part + (i < n_w2 ? i : part_b1 + (i - n_w2));
        float g0 = 0.f, g1 = 0.f, g2 = 0.f, g3 = 0.f;
        int q = 0;
        for (; q + 4 <= nparts; q += 4) {
            g0 += src[(size_t)q * stride]; g1 += src[(size_t)(q + 1) * stride];
            g2 += src[(size_t)(q + 2) * stride]; g3 += src[(size_t)(q + 3) * stride];
        }
        for (; q < nparts; ++q) g0 += src[(size_t)q * stride];
        const float gv[1] = {(g0 + g1) + (g2 + g3)};
        if (i < n_w2) e2.row<1>(s2, i / N2, i % N2, gv, 1, -1);
        else e1.row<1>(s1, K1, i - n_w2, gv, 1, -1);
    }
    e2.finish(s2);
}

}  // namespace poz
