// Inference of a few rows (the REST layer's compute_output on one activation vector, main.py:257 ->
// neural_net_model.py:375-396) in ONE launch.
//
// Per-layer kernels cost ~5 us each at batch 1 whatever they compute (launch, a dependent round trip to
// L2 for the weights, drain), and the layers depend on each other.  Here a single thread-block cluster
// runs the whole stack: every CTA owns a slice of each layer's output columns, reads its slice of the
// weights straight from global memory (for one row this is a matrix-vector product: each weight is used
// once, so there is nothing to stage), reduces the k-partials of its threads through shared memory,
// and pushes its slice of the activations into every peer's input buffer with distributed-shared-memory
// stores; one cluster barrier per layer then separates the layers (against >1 us for a kernel boundary
// and ~2 us for a grid barrier on this part).  Rank 0 applies a trailing softmax and writes the result.
//
// Limits (host-checked): float32, plain Linear + relu/sigmoid/tanh units with an optional trailing
// softmax, batch <= SI_MAX_BATCH, every layer's columns per CTA <= SI_MAX_COLS.
#pragma once
#include <cooperative_groups.h>
#include "poz_common.cuh"
#include "poz_simt_fast.cuh"   // kernel trace records (POZ_KERNEL_TRACE)

namespace poz {
namespace si {

namespace cg = cooperative_groups;

constexpr int SI_THREADS = 512;
constexpr int SI_MAX_BATCH = 4;        // measured: beyond a few rows the per-layer kernels are as fast
constexpr int SI_MAX_COLS = 64;       // output columns of one layer owned by one CTA
constexpr int SI_MAX_LAYERS = 8;

struct Args {
    int nl, has_bias, batch, max_w;   // max_w: widest activation vector (row pitch of the input buffer)
    int sizes[SI_MAX_LAYERS + 1];
    int acts[SI_MAX_LAYERS];
    const float* wb[SI_MAX_LAYERS];   // [K (+1)][N] images of W and b
    const float* x;                   // [batch][sizes[0]]
    float* out;                       // [batch][sizes[nl]]
};

__global__ void __launch_bounds__(SI_THREADS, 1) infer_small_kernel(const __grid_constant__ Args a) {
    extern __shared__ __align__(16) float si_smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int C = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    const int tid = threadIdx.x, B = a.batch;
    const int lane = tid & 31, warp = tid >> 5;
    // two input buffers: while a layer is computed from one, the peers push the next layer's input into the other
    float* inbuf[2] = {si_smem, si_smem + (size_t)B * a.max_w};     // [B][K of the layer]
    float* red = si_smem + (size_t)2 * B * a.max_w;                 // [k-group][B][column]
    __shared__ float sbias[SI_MAX_COLS];

    fast::KernelTraceRec* ktr = nullptr;
    if (fast::d_ktrace && tid == 0) {
        const unsigned int slot = atomicAdd(&fast::d_ktrace_n, 1u);
        if (slot < fast::KTRACE_CAP) {
            ktr = fast::d_ktrace + slot;
            ktr->tag = 64; ktr->m = B; ktr->n = a.sizes[a.nl]; ktr->k = a.sizes[0];
            for (int q = 0; q < 6; ++q) ktr->t[q] = 0;
            ktr->t[0] = fast::ktrace_now();
        }
    }
    pdl_trigger();
    pdl_wait();
    // Layer geometry of this thread, and the first eight weights of its k walk: they are requested before
    // the layer's input is complete (for layer 0 before x is copied in, for the others before the cluster
    // barrier of the layer above), which takes one L2 round trip per layer off the dependent chain.
    struct Geo { int K, N, cols_per, c0, nc, CW, KG, col, kg; };
    auto geometry = [&](int l) {
        Geo g;
        g.K = a.sizes[l]; g.N = a.sizes[l + 1];
        g.cols_per = (g.N + C - 1) / C;
        g.c0 = rank * g.cols_per;
        g.nc = max(0, min(g.cols_per, g.N - g.c0));
        g.CW = SI_THREADS / 32;                            // threads along the columns (power of two); at most
        while (g.CW < g.cols_per) g.CW <<= 1;              // 32 thread groups along k, so that one warp can
        g.KG = SI_THREADS / g.CW;                          // add a column's k-partials with shuffles
        g.col = tid % g.CW; g.kg = tid / g.CW;
        return g;
    };
    float wpre[8];
    auto prefetch = [&](int l, const Geo& g) {
        const float* wp = a.wb[l] + g.c0 + g.col;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int k = g.kg + u * g.KG;
            wpre[u] = (g.col < g.nc && k < g.K) ? __ldg(wp + (size_t)k * g.N) : 0.f;
        }
    };
    Geo g = geometry(0);
    prefetch(0, g);
    {   // the input rows (every CTA needs all of them)
        const int n = B * a.sizes[0];
        for (int i = tid; i < n; i += SI_THREADS) inbuf[0][i] = a.x[i];
    }
    __syncthreads();
    if (ktr) ktr->t[1] = fast::ktrace_now();

    for (int l = 0; l < a.nl; ++l) {
        const float* in = inbuf[l & 1];
        float* nxt = inbuf[(l + 1) & 1];
        const int K = g.K, N = g.N, c0 = g.c0, nc = g.nc, CW = g.CW, KG = g.KG, col = g.col, kg = g.kg;
        const float* W = a.wb[l];
        // bias of this CTA's columns: requested now, needed after the k walk
        float bias_v = 0.f;
        if (a.has_bias && warp * 32 + lane < nc) bias_v = __ldg(W + (size_t)K * N + c0 + tid);
        float acc[SI_MAX_BATCH];
#pragma unroll
        for (int b = 0; b < SI_MAX_BATCH; ++b) acc[b] = 0.f;
        if (col < nc) {
            const float* wp = W + c0 + col;
            // the prefetched first eight k-values of this thread
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int k = kg + u * KG;
                if (k < K) {
#pragma unroll
                    for (int b = 0; b < SI_MAX_BATCH; ++b)
                        if (b < B) acc[b] = fmaf(in[b * K + k], wpre[u], acc[b]);
                }
            }
            int k = kg + 8 * KG;
            // eight weights in flight per thread, then the rows they multiply
            for (; k + 7 * KG < K; k += 8 * KG) {
                float w[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) w[u] = __ldg(wp + (size_t)(k + u * KG) * N);
#pragma unroll
                for (int u = 0; u < 8; ++u)
#pragma unroll
                    for (int b = 0; b < SI_MAX_BATCH; ++b)
                        if (b < B) acc[b] = fmaf(in[b * K + k + u * KG], w[u], acc[b]);
            }
            for (; k < K; k += KG) {
                const float w = __ldg(wp + (size_t)k * N);
#pragma unroll
                for (int b = 0; b < SI_MAX_BATCH; ++b)
                    if (b < B) acc[b] = fmaf(in[b * K + k], w, acc[b]);
            }
        }
#pragma unroll
        for (int b = 0; b < SI_MAX_BATCH; ++b)
            if (b < B) red[((size_t)kg * B + b) * CW + col] = acc[b];
        if (tid < nc) sbias[tid] = bias_v;
        __syncthreads();
        const bool softmax_here = a.acts[l] == ACT_SOFTMAX;   // only the last unit (host-checked)
        const int act = a.acts[l];
        if (l + 1 < a.nl) {                                // next layer's first weights, before this layer's barrier
            g = geometry(l + 1);
            prefetch(l + 1, g);
        }
        // one warp per output: lane q holds k-group q's partial; lane p then stores the result into peer p's
        // next input buffer (remote shared-memory stores are not waited for; the cluster barrier publishes them)
        for (int o = warp; o < B * nc; o += SI_THREADS / 32) {
            const int b = o / nc, cc = o - b * nc;
            float s = lane < KG ? red[((size_t)lane * B + b) * CW + cc] : 0.f;
            s = warp_sum(s);
            if (a.has_bias) s += sbias[cc];
            const float v = softmax_here ? s : act_fwd<float>(act, s);
            if (lane < C) cluster.map_shared_rank(nxt, lane)[b * N + c0 + cc] = v;
        }
        cluster.sync();                                    // every CTA now holds the layer's full output
        if (ktr && l < 3) ktr->t[2 + l] = fast::ktrace_now();
    }
    const float* in = inbuf[a.nl & 1];

    if (rank == 0) {
        const int N = a.sizes[a.nl];
        if (a.acts[a.nl - 1] == ACT_SOFTMAX) {
            // one warp per row (neural_net_model.py:188)
            const int lane = tid & 31, warp = tid >> 5;
            for (int b = warp; b < B; b += SI_THREADS / 32) {
                const float* z = in + b * N;
                float mx = -INFINITY;
                for (int j = lane; j < N; j += 32) mx = fmaxf(mx, z[j]);
                mx = warp_max(mx);
                float s = 0.f;
                for (int j = lane; j < N; j += 32) s += expf(z[j] - mx);
                s = warp_sum(s);
                const float inv = 1.f / s;
                for (int j = lane; j < N; j += 32) a.out[(size_t)b * N + j] = expf(z[j] - mx) * inv;
            }
        } else {
            for (int i = tid; i < B * N; i += SI_THREADS) a.out[i] = in[i];
        }
    }
    if (ktr) ktr->t[5] = fast::ktrace_now();
}

}  // namespace si
}  // namespace poz
