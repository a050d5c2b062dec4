// Element-wise / reduction kernels of the wide (bf16) path: MSE loss + its gradient, bias gradient.
#pragma once
#include <cuda_bf16.h>
#include "poz_kernels.cuh"
#include "poz_tc_gemm.cuh"

namespace poz {

constexpr int COLSUM_ROWS = 32;   // row splits of the bias-gradient reduction
inline int tc_tile_m() { return tc::BM; }
inline int tc_tile_n() { return tc::BN; }

struct WideLossArgs {
    const __nv_bfloat16* y;        // final output [total]
    const __nv_bfloat16* target;   // [total]
    __nv_bfloat16* dz;             // gradient of the last linear's PRE-dropout output (nullable)
    long long total;
    int act;
    double inv_count;              // 1 / (global_batch * out)
    double* cost_base;
    long long cost_cap;
    const DevState* st;
    int N;                         // out features (multiple of 8: the eight elements of a thread share a row)
    DropPoint dzp;                 // dropout point on the last linear's output (only when an activation follows it)
    __nv_bfloat16* da_rec;         // diagnostics (nullable): gradient of the recorded final output
    __nv_bfloat16* dz_rec;         // diagnostics (nullable): gradient of the recorded last linear output
};

// out[b][:] = src[idx[b]][:] -- the sampled rows of train() (neural_net_model.py:460-461) as the dense bf16 batch the
// TMA-fed GEMMs read; 16 bytes per thread and trip (cols is a multiple of 8)
__global__ void wide_gather_rows_kernel(const __nv_bfloat16* __restrict__ src, const int* __restrict__ idx, long long rows,
                                        int cols, __nv_bfloat16* __restrict__ dst) {
    const int vpr = cols / 8;
    const long long total = rows * vpr;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / vpr;
        const int v = (int)(i - r * vpr);
        reinterpret_cast<uint4*>(dst + r * cols)[v] = __ldg(reinterpret_cast<const uint4*>(src + (long long)__ldg(idx + r) * cols) + v);
    }
}

// MSE (neural_net_model.py:405-406) over bf16 tensors, 8 elements (16 bytes) per thread per trip.
__global__ void wide_mse_loss_kernel(WideLossArgs a) {
    double local = 0.0;
    const float two_inv = (float)(2.0 * a.inv_count);
    const long long nvec = a.total / 8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec;
         i += (long long)gridDim.x * blockDim.x) {
        uint4 qy = reinterpret_cast<const uint4*>(a.y)[i];
        uint4 qt = reinterpret_cast<const uint4*>(a.target)[i];
        const __nv_bfloat162* hy = reinterpret_cast<const __nv_bfloat162*>(&qy);
        const __nv_bfloat162* ht = reinterpret_cast<const __nv_bfloat162*>(&qt);
        uint32_t o[4], oa[4];
        float acc = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float2 y2 = __bfloat1622float2(hy[e]), t2 = __bfloat1622float2(ht[e]);
            float d0 = y2.x - t2.x, d1 = y2.y - t2.y;
            acc += d0 * d0 + d1 * d1;
            oa[e] = tc::pack_bf16(two_inv * d0, two_inv * d1);
            o[e] = tc::pack_bf16(two_inv * d0 * act_bwd_from_out<float>(a.act, y2.x),
                                 two_inv * d1 * act_bwd_from_out<float>(a.act, y2.y));
        }
        local += (double)acc;
        if (a.da_rec) reinterpret_cast<uint4*>(a.da_rec)[i] = make_uint4(oa[0], oa[1], oa[2], oa[3]);
        if (a.dz_rec) reinterpret_cast<uint4*>(a.dz_rec)[i] = make_uint4(o[0], o[1], o[2], o[3]);
        if (a.dzp.enabled) {
            // the mask of the forward pass, regenerated (or read back) for the eight elements of this thread
            const long long e0 = i * 8;
            const long long m = e0 / a.N;
            const int n0 = (int)(e0 - m * a.N);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const __nv_bfloat162 h2 = *reinterpret_cast<const __nv_bfloat162*>(&o[e]);
                float2 g = __bfloat1622float2(h2);
                g.x = drop_keep(a.dzp, a.st, m, n0 + 2 * e, a.N) ? g.x * (float)a.dzp.scale : 0.f;
                g.y = drop_keep(a.dzp, a.st, m, n0 + 2 * e + 1, a.N) ? g.y * (float)a.dzp.scale : 0.f;
                o[e] = tc::pack_bf16(g.x, g.y);
            }
        }
        if (a.dz) reinterpret_cast<uint4*>(a.dz)[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
    if (a.cost_base) block_add_double(local * a.inv_count, cost_slot(a.cost_base, a.cost_cap, a.st));
}

// Bias gradient db[n] = sum_b dz[b][n]  (test_torch_backward.py:166), two deterministic passes.
// pass 1: block (x, r) sums rows r, r+R, ... of 64 columns; a warp reads one 128-byte row segment.
__global__ void wide_colsum_partial_kernel(const __nv_bfloat16* dz, long long rows, int N, float* partial) {
    __shared__ float red[8][64];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = blockIdx.x * 64 + lane * 2;
    float s0 = 0.f, s1 = 0.f;
    if (col < N) {
        for (long long r = (long long)blockIdx.y * 8 + warp; r < rows; r += (long long)gridDim.y * 8) {
            __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(dz + r * N + col);
            float2 f = __bfloat1622float2(h);
            s0 += f.x; s1 += f.y;
        }
    }
    red[warp][lane * 2] = s0;
    red[warp][lane * 2 + 1] = s1;
    __syncthreads();
    if (threadIdx.x < 64) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
        int cc = blockIdx.x * 64 + threadIdx.x;
        if (cc < N) partial[(long long)blockIdx.y * N + cc] = t;
    }
}
__global__ void wide_colsum_final_kernel(const float* partial, int N, float* out) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float t = 0.f;
    for (int r = 0; r < COLSUM_ROWS; ++r) t += partial[(long long)r * N + n];
    out[n] = t;
}

}  // namespace poz
