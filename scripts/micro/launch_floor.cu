// Microbenchmark: what a chain of dependent kernel launches costs inside a CUDA graph on this GPU,
// against grid-wide barriers inside one cooperative kernel.  Informs how the narrow train step is cut.
#include <cstdio>
#include <cuda_runtime.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

__global__ void k_touch(float* p, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = p[i] * 1.0001f + 1.f;
}

__global__ void k_barriers(float* p, int n, int rounds, unsigned* bar) {
    for (int r = 0; r < rounds; ++r) {
        int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n) p[i] = p[i] * 1.0001f + 1.f;
        __syncthreads();
        if (threadIdx.x == 0) {
            volatile unsigned* gen_p = bar + 1;
            const unsigned gen = *gen_p;
            __threadfence();
            if (atomicAdd(bar, 1u) == gridDim.x - 1) { bar[0] = 0u; __threadfence(); atomicAdd(bar + 1, 1u); }
            else while (*gen_p == gen) { }
            __threadfence();
        }
        __syncthreads();
    }
}

__global__ void k_cg_barriers(float* p, int n, int rounds) {
    cg::grid_group grid = cg::this_grid();
    for (int r = 0; r < rounds; ++r) {
        int i = blockIdx.x * blockDim.x + threadIdx.x;
        if (i < n) p[i] = p[i] * 1.0001f + 1.f;
        grid.sync();
    }
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int n = sms * 256;
    float* p; cudaMalloc(&p, n * sizeof(float)); cudaMemset(p, 0, n * sizeof(float));
    unsigned* bar; cudaMalloc(&bar, 8); cudaMemset(bar, 0, 8);
    cudaStream_t s; cudaStreamCreate(&s);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int chain : {1, 2, 4, 9, 18}) {
        for (int blocks : {1, sms, 4 * sms}) {
            cudaGraph_t g; cudaGraphExec_t ge;
            cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
            for (int i = 0; i < chain; ++i) k_touch<<<blocks, 256, 0, s>>>(p, n);
            cudaStreamEndCapture(s, &g);
            cudaGraphInstantiate(&ge, g, 0);
            for (int i = 0; i < 20; ++i) cudaGraphLaunch(ge, s);
            cudaStreamSynchronize(s);
            const int reps = 200;
            cudaEventRecord(e0, s);
            for (int i = 0; i < reps; ++i) cudaGraphLaunch(ge, s);
            cudaEventRecord(e1, s);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            printf("graph of %2d dependent kernels, %4d blocks each: %7.2f us per graph, %5.2f us per kernel\n", chain, blocks,
                   ms * 1000 / reps, ms * 1000 / reps / chain);
            cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
        }
    }
    // the same chain launched as thread-block clusters of 4 (the split-K GEMMs of the step are cluster launches)
    for (int chain : {9, 18}) {
        cudaGraph_t g; cudaGraphExec_t ge;
        cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal);
        for (int i = 0; i < chain; ++i) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(4 * sms); cfg.blockDim = dim3(256); cfg.stream = s;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = 4; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            cudaLaunchKernelEx(&cfg, k_touch, p, n);
        }
        cudaStreamEndCapture(s, &g);
        cudaGraphInstantiate(&ge, g, 0);
        for (int i = 0; i < 20; ++i) cudaGraphLaunch(ge, s);
        cudaStreamSynchronize(s);
        const int reps = 200;
        cudaEventRecord(e0, s);
        for (int i = 0; i < reps; ++i) cudaGraphLaunch(ge, s);
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("graph of %2d dependent CLUSTER(4) kernels, %4d blocks each: %7.2f us per graph, %5.2f us per kernel\n", chain, 4 * sms,
               ms * 1000 / reps, ms * 1000 / reps / chain);
        cudaGraphExecDestroy(ge); cudaGraphDestroy(g);
    }
    for (int rounds : {1, 9, 33}) {
        int nn = n;
        void* args[] = {&p, &nn, &rounds, &bar};
        for (int i = 0; i < 20; ++i) cudaLaunchCooperativeKernel((void*)k_barriers, dim3(sms), dim3(256), args, 0, s);
        cudaStreamSynchronize(s);
        const int reps = 200;
        cudaEventRecord(e0, s);
        for (int i = 0; i < reps; ++i) cudaLaunchCooperativeKernel((void*)k_barriers, dim3(sms), dim3(256), args, 0, s);
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("cooperative kernel, %2d rounds with own barrier : %7.2f us per launch\n", rounds, ms * 1000 / reps);
        void* args2[] = {&p, &nn, &rounds};
        for (int i = 0; i < 20; ++i) cudaLaunchCooperativeKernel((void*)k_cg_barriers, dim3(sms), dim3(256), args2, 0, s);
        cudaStreamSynchronize(s);
        cudaEventRecord(e0, s);
        for (int i = 0; i < reps; ++i) cudaLaunchCooperativeKernel((void*)k_cg_barriers, dim3(sms), dim3(256), args2, 0, s);
        cudaEventRecord(e1, s);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        printf("cooperative kernel, %2d rounds with grid.sync()  : %7.2f us per launch\n", rounds, ms * 1000 / reps);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
