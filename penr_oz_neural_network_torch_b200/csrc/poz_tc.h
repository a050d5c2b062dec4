// Launchers of the tcgen05 GEMM (csrc/poz_tc.cu) used by the bf16 context.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "poz_common.cuh"

namespace poz {
namespace tc {
// What a train step adds to the two bf16-output launches: the unit's dropout points (neural_net_model.py:393-395)
// and, on a diagnostics step (:491-493), the tensors the statistics need.  Everything optional.
struct Extras {
    DropPoint dz{}, da{};
    const DevState* st = nullptr;
    __nv_bfloat16* z_rec = nullptr;    // forward: recorded (post-dropout) linear output
    __nv_bfloat16* da_rec = nullptr;   // backward-x: gradient of the previous unit's recorded activation output
    __nv_bfloat16* dz_rec = nullptr;   // backward-x: gradient of its recorded linear output
};
// out[M][N] = act(x[M][K] * w[K][N] + bias[N])
int gemm_forward(const __nv_bfloat16* x, const __nv_bfloat16* w, long long M, int N, int K, const float* bias,
                 int act, __nv_bfloat16* out, int sms, cudaStream_t s, const Extras* ex = nullptr);
// dz_prev[M][K_in] = (dz[M][N_out] * w[K_in][N_out]^T) .* act_prev'(a_prev[M][K_in])
int gemm_backward_data(const __nv_bfloat16* dz, const __nv_bfloat16* w, long long M, int K_in, int N_out,
                       const __nv_bfloat16* a_prev, int act_prev, __nv_bfloat16* dz_prev, int sms, cudaStream_t s,
                       const Extras* ex = nullptr);
// g[split][K_in][N_out] = partial sums over the batch of x[batch][K_in]^T * dz[batch][N_out]
int gemm_backward_weight(const __nv_bfloat16* x, const __nv_bfloat16* dz, long long batch, int K_in, int N_out,
                         float* g, long long slab, int splits, int sms, cudaStream_t s);
// TMA descriptor of a row-major fp32 matrix [rows][cols] (pitch ld elements) for boxes of 32 columns (one
// 128-byte swizzle row) by box_rows rows; `map_out` points at a CUtensorMap.  Used by the TMA-fed narrow GEMM.
int make_map_f32(void* map_out, const void* ptr, long long rows, long long cols, long long ld, int box_rows);
}  // namespace tc
}  // namespace poz
