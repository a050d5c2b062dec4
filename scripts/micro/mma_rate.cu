// Microbenchmark: issue rate of the legacy warp-level tensor-core instruction (mma.sync m16n8k8 tf32,
// m16n8k16 bf16) against FFMA on this GPU.  Informs the choice of arithmetic for the narrow-MLP kernels.
#include <cstdio>
#include <cuda_runtime.h>
#include <cstdint>

__global__ void k_tf32(float* out, int iters) {
    float c[8][4];
    for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
    uint32_t a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 5, a3 = 7, b0 = threadIdx.x + 1, b1 = 11;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0.f;
    for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_bf16(float* out, int iters) {
    float c[8][4];
    for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.f;
    uint32_t a0 = threadIdx.x, a1 = threadIdx.x * 3, a2 = 5, a3 = 7, b0 = threadIdx.x + 1, b1 = 11;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3])
                         : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
    float s = 0.f;
    for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma(float* out, int iters, float x, float y) {
    float c[32];
    for (int i = 0; i < 32; ++i) c[i] = i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 32; ++i) c[i] = fmaf(c[i], x, y);
    }
    float s = 0.f;
    for (int i = 0; i < 32; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_ffma2(float* out, int iters, float x, float y) {
    unsigned long long c[16], xx, yy;
    asm("mov.b64 %0, {%1, %1};" : "=l"(xx) : "f"(x));
    asm("mov.b64 %0, {%1, %1};" : "=l"(yy) : "f"(y));
    for (int i = 0; i < 16; ++i) { float v = i; asm("mov.b64 %0, {%1, %1};" : "=l"(c[i]) : "f"(v)); }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(c[i]) : "l"(xx), "l"(yy));
    }
    float s = 0.f;
    for (int i = 0; i < 16; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(c[i])); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* out; cudaMalloc(&out, sizeof(float) * sms * 4 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int warps = 4; warps <= 32; warps *= 2) {
        const int blocks = sms * (warps >= 8 ? 2 : 1), threads = warps * 32 / (warps >= 8 ? 2 : 1);
        float ms;
        k_tf32<<<blocks, threads>>>(out, 100); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_tf32<<<blocks, threads>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        double fl = 2.0 * 16 * 8 * 8 * 8.0 * iters * (double)blocks * threads / 32;
        printf("warps/SM %2d  tf32 m16n8k8 : %8.1f TFLOP/s\n", warps, fl / ms * 1e-9);
        k_bf16<<<blocks, threads>>>(out, 100); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_bf16<<<blocks, threads>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        fl = 2.0 * 16 * 8 * 16 * 8.0 * iters * (double)blocks * threads / 32;
        printf("warps/SM %2d  bf16 m16n8k16: %8.1f TFLOP/s\n", warps, fl / ms * 1e-9);
        k_ffma<<<blocks, threads>>>(out, 100, 1.0001f, 0.5f); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_ffma<<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        fl = 2.0 * 32.0 * iters * (double)blocks * threads;
        printf("warps/SM %2d  ffma         : %8.1f TFLOP/s\n", warps, fl / ms * 1e-9);
        k_ffma2<<<blocks, threads>>>(out, 100, 1.0001f, 0.5f); cudaDeviceSynchronize();
        cudaEventRecord(e0); k_ffma2<<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        printf("warps/SM %2d  ffma2 (f32x2): %8.1f TFLOP/s\n", warps, fl / ms * 1e-9);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
