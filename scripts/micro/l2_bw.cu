// Microbenchmark: L2 -> SM read bandwidth on this GPU for (a) fully coalesced 128-bit loads and
// (b) the access pattern of the narrow GEMM's operand tiles (cp.async 16 B, 64-byte row segments at
// a 3136-byte row pitch).  The buffers (16 MB / 3.2 MB) stay L2-resident across the repetitions.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void k_ldg(const float4* __restrict__ p, size_t n4, int reps, float* out) {
    float acc = 0.f;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
            float4 v;
            asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p + i));
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 12345.678f) out[0] = acc;
}

// rows x 784 floats; each CTA (128 threads) walks its 32 rows in k-tiles of BK floats: 32 rows x (BK*4) bytes per tile
template <int BK>
__global__ void k_tiles(const float* __restrict__ x, int rows, int reps, float* out) {
    __shared__ __align__(16) float s[4][32 * BK];
    const int tid = threadIdx.x;
    constexpr int CV = BK / 4;                    // 16-byte chunks per row
    float acc = 0.f;
    const int row_tiles = rows / 32;
    for (int r = 0; r < reps; ++r)
        for (int rt = blockIdx.x; rt < row_tiles; rt += gridDim.x) {
            const float* base = x + (size_t)rt * 32 * 784;
            int stage = 0;
            for (int k0 = 0; k0 + BK <= 784; k0 += BK) {
                for (int c = tid; c < 32 * CV; c += 128) {
                    const int rr = c / CV, cc = (c % CV) * 4;
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(&s[stage & 3][rr * BK + cc]);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(base + (size_t)rr * 784 + k0 + cc));
                }
                asm volatile("cp.async.commit_group;");
                asm volatile("cp.async.wait_group 3;");
                ++stage;
            }
            asm volatile("cp.async.wait_group 0;");
            __syncthreads();
            acc += s[0][tid];
        }
    if (acc == 12345.678f) out[0] = acc;
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const size_t bytes = 16u << 20;
    float4* p; cudaMalloc(&p, bytes); cudaMemset(p, 0, bytes);
    float* x; cudaMalloc(&x, 1024 * 784 * 4); cudaMemset(x, 0, 1024 * 784 * 4);
    float* out; cudaMalloc(&out, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float ms;
    for (int per_sm : {1, 2, 4, 8}) {
        const int reps = 50;
        k_ldg<<<sms * per_sm, 256>>>(p, bytes / 16, 2, out);
        cudaEventRecord(e0); k_ldg<<<sms * per_sm, 256>>>(p, bytes / 16, reps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        printf("coalesced ld.global.cg.v4, %d CTAs of 256 per SM: %7.1f GB/s\n", per_sm, (double)bytes * reps / ms * 1e-6);
    }
    for (int per_sm : {1, 2, 4, 8}) {
        const int reps = 400;
        k_tiles<16><<<sms * per_sm, 128>>>(x, 1024, 2, out);
        cudaEventRecord(e0); k_tiles<16><<<sms * per_sm, 128>>>(x, 1024, reps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        printf("cp.async 16 B, 32 x 64-byte row segments per tile, %d CTAs of 128 per SM: %7.1f GB/s\n", per_sm,
               1024.0 * 784 * 4 * reps / ms * 1e-6);
        k_tiles<32><<<sms * per_sm, 128>>>(x, 1024, 2, out);
        cudaEventRecord(e0); k_tiles<32><<<sms * per_sm, 128>>>(x, 1024, reps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        printf("cp.async 16 B, 32 x 128-byte row segments per tile, %d CTAs of 128 per SM: %7.1f GB/s\n", per_sm,
               1024.0 * 768 * 4 * reps / ms * 1e-6);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
