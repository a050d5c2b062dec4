// Interface of the split-bf16 tcgen05 GEMM (csrc/poz_tc3.cu): plane stacks, epilogue description,
// launchers.  Host-only header (no device code) so that poz_api.cu can include it.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace poz {
namespace tc3 {

// An fp32 matrix carried as three bf16 planes: [3][rows][ld], plane p at ptr + p * plane_stride.
struct PlaneMat {
    const __nv_bfloat16* ptr;
    long long rows;
    int cols;
    long long ld;            // elements, multiple of 8
    long long plane_stride;  // elements, multiple of 8
    int planes = 3;          // 2: fp16 planes hi, lo' = (x - hi) * 2^11 (the fp16-split inference kernels)
};

enum EpiMode { EPI_BIAS_ACT = 0, EPI_ACTGRAD = 1, EPI_PLAIN = 2 };

struct Epi3 {
    float* out_f32;               // nullable; [M][ld_f32] (+ split * slab)
    long long ld_f32, slab;
    __nv_bfloat16* out_planes;    // nullable; [3][rows][ld_planes]
    long long ld_planes, plane_stride;
    const float* bias;            // EPI_BIAS_ACT: [N] or null
    const float* aux;             // EPI_ACTGRAD: recorded activation of the previous unit [M][ld_aux]
    long long ld_aux;
    int mode, act;
    // Optional fused head (forward, EPI_BIAS_ACT, one 128-wide tile holds whole rows: 64 < N <= 128): the NEXT Linear
    // unit, at most HEAD_MAX_N wide, applied to the tile's activated rows in fp32 by the epilogue warps:
    // head_out[M][head_n] = head_act(row * head_w[N][head_n] + head_b).  out_f32 / out_planes may then be null.
    const float* head_w;          // nullable
    const float* head_b;          // nullable
    float* head_out;
    int head_n, head_act;
    // fp16-split kernels (two planes, three MMAs per product): the kernel stores 1 to *f16_flag when an operand leaves the
    // range in which two fp16 planes carry fp32 precision (|x| >= 2^15, or a whole row / the whole weight matrix below
    // 2^-16); the caller then repeats the pass with the split-bf16 kernels, which are launched with run_if = that flag
    // and return at once while it is 0.
    int* f16_flag;                // nullable
    const unsigned* f16_w_amax;   // bits of max |W| of this layer (device), written when its fp16 planes were made
    const int* run_if;            // nullable: the kernel does nothing unless *run_if != 0
    int probe;                    // POZ_X32T_PROBE=1: one converter warp of block 0 adds up the cycles of its loop's phases
};
constexpr int HEAD_MAX_N = 16;
constexpr int HEAD_SMEM_BYTES = (128 * HEAD_MAX_N + 2 * 128 * HEAD_MAX_N) * 4;   // staged weights + two partial-sum buffers

// row_idx (nullable) gathers source rows: plane row r = source row row_idx[r]
int split_planes(const float* src, const int* row_idx, long long rows, int cols, long long ld, __nv_bfloat16* dst,
                 long long ldp, long long plane_stride, cudaStream_t stream);
// out[M][N] = epilogue(x[M][K] * w[K][N])
int gemm_forward(const PlaneMat& x, const PlaneMat& w, long long M, int N, int K, const Epi3& ep, int sms, cudaStream_t s);
// whether gemm_forward takes ep.head_* for this width
bool gemm_forward_head_ok(int N, int head_n);
// the same with x given as the fp32 matrix itself (pitch ld floats, 16-byte aligned rows): its planes are made
// inside the kernel and never written to global memory (csrc/poz_tc3_x32t_kernel.cuh); N must be > 64
int gemm_forward_x32(const float* x, long long ld, const PlaneMat& w, long long M, int N, int K, const Epi3& ep, int sms,
                     cudaStream_t s);
// fp32 matrix -> two fp16 planes [2][rows][ldp] (hi, (x - hi) * 2^11); max |x| is folded into *amax_bits (atomicMax on the
// float's bits; the caller zeroes it first)
int split_planes_f16(const float* src, long long rows, int cols, long long ld, __nv_bfloat16* dst, long long ldp,
                     long long plane_stride, unsigned* amax_bits, cudaStream_t stream);
bool gemm_forward_x32_ok(const float* x, long long ld, int N);
void x32t_probe_dump();           // prints and clears the sums of POZ_X32T_PROBE=1
// whether gemm_forward_x32 takes ep.head_* (only the tensor-memory variant does)
bool gemm_forward_x32_head_ok();
// out[M][K_in] = epilogue(dz[M][N_out] * w[K_in][N_out]^T)
int gemm_backward_data(const PlaneMat& dz, const PlaneMat& w, long long M, int K_in, int N_out, const Epi3& ep, int sms,
                       cudaStream_t s);
// out[split][K_in][N_out] = partial sums over the batch of x[batch][K_in]^T * dz[batch][N_out]
int gemm_backward_weight(const PlaneMat& x, const PlaneMat& dz, long long batch, int K_in, int N_out, int splits,
                         const Epi3& ep, int sms, cudaStream_t s);

}  // namespace tc3
}  // namespace poz
