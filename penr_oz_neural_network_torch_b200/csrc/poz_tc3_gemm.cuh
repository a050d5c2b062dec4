// Split-bf16 ("bf16x3") tcgen05 GEMM: fp32-grade matrix products on the 5th-gen tensor cores.
//
// Every fp32 operand x is carried as three bf16 planes hi + mid + lo (x = hi + mid + lo to ~2^-24),
// and C = A*B is accumulated in tensor memory from the six plane products that matter
//   hi*hi, hi*mid, mid*hi, mid*mid, hi*lo, lo*hi          (dropped terms are <= 2^-24 relative),
// which keeps the float32 contract of the narrow configurations (rtol 1e-5 vs the float64
// reference) while the contraction runs at tensor-core rate instead of on the fp32 FMA pipe.
//
// The tensor core truncates its fp32 accumulator on every tcgen05.mma, an error proportional to the
// accumulator's magnitude and to the number of accumulation steps, so the five small products go to
// an accumulator of their own (2^-8 of the magnitude) and only hi*hi (K/16 steps) feeds the main
// one; the epilogue adds the two.  Measured: error/max|C| ~1.5e-6 at K = 784 instead of 5e-6.
//
// Tile: 128 x BN output, BK = 64, one stage = 3 A planes + 3 B planes; accumulator pairs
// double-buffered in TMEM (2 x 2 x BN columns) so the epilogue of tile i overlaps the MMAs of
// tile i+1.
//   warps 0-7 : epilogue (TMEM lane quarter w%4, column half w/4)
//   warp  8   : TMA producer       warp 9 : TMEM allocator + MMA issuer
// Planes are stored [3][rows][pitch] and addressed through one 3-D tensor map per operand.
#pragma once
#include <cuda_fp16.h>
#include "poz_tc_gemm.cuh"
#include "poz_tc3.h"

namespace poz {
namespace tc3 {

using tc::MAJOR_K;
using tc::MAJOR_MN;

constexpr int BM = 128, BK = 64, UMMA_K = 16, PLANES = 3;
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = (NUM_EPI_WARPS + 2) * 32;

template <int BN> struct Cfg {
    static constexpr int STAGES = BN <= 64 ? 3 : 2;
    static constexpr int A_PLANE = BM * BK * 2;
    static constexpr int B_PLANE = BN * BK * 2;
    static constexpr int STAGE_BYTES = PLANES * (A_PLANE + B_PLANE);
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
    // two accumulators per buffer (hi*hi, and the five small terms), two buffers
    static constexpr int TMEM_COLS = 4 * BN;                        // 256 or 512: powers of two
};

__device__ __forceinline__ void tma_load_3d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(tc::smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1),
        "r"(c2)
        : "memory");
}

template <int A_MAJOR, int B_MAJOR, int BN> __device__ __forceinline__ constexpr uint32_t make_idesc3() {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)A_MAJOR << 15) | ((uint32_t)B_MAJOR << 16) |
           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

// the same with fp16 operands (a_format = b_format = 0)
template <int A_MAJOR, int B_MAJOR, int BN> __device__ __forceinline__ constexpr uint32_t make_idesc_f16() {
    return (1u << 4) | ((uint32_t)A_MAJOR << 15) | ((uint32_t)B_MAJOR << 16) | ((uint32_t)(BN >> 3) << 17) |
           ((uint32_t)(BM >> 4) << 24);
}

// fp16 split: x = hi + lo with hi = fp16(x) (11 significant bits) and lo = x - hi (exact in fp32, |lo| <= 2^-11 |x|);
// lo is carried as lo' = fp16(lo * 2^11), which keeps it in fp16's normal range wherever hi is, so hi + lo' 2^-11
// has 22 significant bits.  A product needs three MMAs: hi*hi, and hi*lo' + lo'*hi in an accumulator of their own that
// the epilogue folds in with the factor 2^-11 (the dropped lo*lo is below 2^-22 relative).  Valid for
// 2^-16 <= max |x| < 2^15 (poz_tc3.h, f16_flag).
constexpr float F16_LO_SCALE = 2048.f, F16_LO_UNSCALE = 1.f / 2048.f;
constexpr float F16_MAX_OK = 32768.f, F16_MIN_OK = 1.52587890625e-05f;
__device__ __forceinline__ void split2h_pair(float x0, float x1, uint32_t& ph, uint32_t& pl) {
    const __half2 h = __floats2half2_rn(x0, x1);
    ph = *reinterpret_cast<const uint32_t*>(&h);
    const float2 hf = __half22float2(h);
    // (measured: rebuilding hi's fp32 value from its bits with shifts and masks instead of __half22float2 made the
    //  converters' load + convert phase 809 instead of 508 cycles per half-block -- the unpacking is not what costs)
    const __half2 l = __floats2half2_rn(fmaf(x0, F16_LO_SCALE, -F16_LO_SCALE * hf.x), fmaf(x1, F16_LO_SCALE, -F16_LO_SCALE * hf.y));
    pl = *reinterpret_cast<const uint32_t*>(&l);
}
__device__ __forceinline__ bool f16_range_bad(float amax) {
    return amax >= F16_MAX_OK || (amax != 0.f && amax < F16_MIN_OK);
}

// x -> (hi, mid, lo) bf16 with x ~= hi + mid + lo
__device__ __forceinline__ void split3(float x, __nv_bfloat16& h, __nv_bfloat16& m, __nv_bfloat16& l) {
    h = __float2bfloat16_rn(x);
    const float r1 = x - __bfloat162float(h);
    m = __float2bfloat16_rn(r1);
    const float r2 = r1 - __bfloat162float(m);
    l = __float2bfloat16_rn(r2);
}

// The same for two values at once, producing the packed bf16x2 words of the three planes (x0 in the low half).
// One packed conversion per plane (cvt.rn.bf16x2.f32) instead of two scalar ones: the scalar conversion
// issues on the quarter-rate conversion pipe and was what bound the in-kernel converters of
// poz_tc3_x32_kernel.cuh (ncu: F2F.BF16.F32 at the top of their samples).  Same rounding, same bits.
__device__ __forceinline__ void split3_pair(float x0, float x1, uint32_t& ph, uint32_t& pm, uint32_t& pl) {
    __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
    ph = *reinterpret_cast<uint32_t*>(&h);
    const float r0 = x0 - __uint_as_float(ph << 16), r1 = x1 - __uint_as_float(ph & 0xffff0000u);
    __nv_bfloat162 m = __floats2bfloat162_rn(r0, r1);
    pm = *reinterpret_cast<uint32_t*>(&m);
    const float s0 = r0 - __uint_as_float(pm << 16), s1 = r1 - __uint_as_float(pm & 0xffff0000u);
    __nv_bfloat162 l = __floats2bfloat162_rn(s0, s1);
    pl = *reinterpret_cast<uint32_t*>(&l);
}

// fp32 matrix [rows][cols] (pitch ld, optional row gather) -> planes [3][rows][ldp]; pad columns
// cols..ldp-1 are zeroed.  One thread converts 8 consecutive elements: two 16-byte loads when the
// source row allows it, one 16-byte store per plane.
__global__ void split_planes_kernel(const float* __restrict__ src, const int* __restrict__ row_idx, long long rows,
                                    int cols, long long ld, __nv_bfloat16* __restrict__ dst, long long ldp,
                                    long long plane_stride) {
    const int groups = (int)(ldp / 8);
    const long long total = rows * groups;
    const bool vec = ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && (ld % 4 == 0);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / groups;
        const int c = (int)(i - r * groups) * 8;
        const float* s = src + (row_idx ? (long long)row_idx[r] : r) * ld + c;
        float x[8];
        if (vec && c + 8 <= cols) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(s));
            const float4 b = __ldg(reinterpret_cast<const float4*>(s) + 1);
            x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w; x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = (c + j < cols) ? s[j] : 0.f;
        }
        uint32_t ph[4], pm[4], pl[4];
#pragma unroll
        for (int j = 0; j < 8; j += 2) split3_pair(x[j], x[j + 1], ph[j / 2], pm[j / 2], pl[j / 2]);
        __nv_bfloat16* o = dst + r * ldp + c;
        *reinterpret_cast<uint4*>(o) = make_uint4(ph[0], ph[1], ph[2], ph[3]);
        *reinterpret_cast<uint4*>(o + plane_stride) = make_uint4(pm[0], pm[1], pm[2], pm[3]);
        *reinterpret_cast<uint4*>(o + 2 * plane_stride) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
    }
}

// fp32 matrix [rows][cols] (pitch ld) -> fp16 planes [2][rows][ldp] (hi, lo'), pad columns zeroed; max |x| -> *amax_bits
__global__ void split_planes_f16_kernel(const float* __restrict__ src, long long rows, int cols, long long ld,
                                        __nv_bfloat16* __restrict__ dst, long long ldp, long long plane_stride,
                                        unsigned* __restrict__ amax_bits) {
    const int groups = (int)(ldp / 8);
    const long long total = rows * groups;
    float amax = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / groups;
        const int c = (int)(i - r * groups) * 8;
        const float* s = src + r * ld + c;
        float x[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            x[j] = (c + j < cols) ? s[j] : 0.f;
            amax = fmaxf(amax, fabsf(x[j]));
        }
        uint32_t ph[4], pl[4];
#pragma unroll
        for (int j = 0; j < 8; j += 2) split2h_pair(x[j], x[j + 1], ph[j / 2], pl[j / 2]);
        __nv_bfloat16* o = dst + r * ldp + c;
        *reinterpret_cast<uint4*>(o) = make_uint4(ph[0], ph[1], ph[2], ph[3]);
        *reinterpret_cast<uint4*>(o + plane_stride) = make_uint4(pl[0], pl[1], pl[2], pl[3]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    if ((threadIdx.x & 31) == 0 && amax > 0.f) atomicMax(amax_bits, __float_as_uint(amax));   // non-negative floats order as their bits
}

}  // namespace tc3
}  // namespace poz
