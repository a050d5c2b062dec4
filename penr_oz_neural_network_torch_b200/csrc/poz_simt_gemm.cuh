// Register-blocked SIMT GEMM used by the narrow (fp32 / fp64) path: one template covers
//   forward      Z  = [X 1] @ [W; b]          (A k-contiguous,  B n-contiguous)
//   backward-X   dA = dZ @ W^T                (A k-contiguous,  B k-contiguous)
//   backward-W   [dW; db] = [X 1]^T @ dZ      (A m-contiguous,  B n-contiguous, split over batch)
// The bias rides along as one extra k (forward) or m (backward-W) index whose operand value is 1,
// which works because a layer's bias is stored directly behind its weight matrix, making
// [W; b] one contiguous row-major [in+1][out] matrix.
#pragma once
#include <cuda_bf16.h>
#include "poz_common.cuh"

namespace poz {

template <typename T> struct Vec16;
template <> struct Vec16<float> { using type = float4; static constexpr int V = 4; };
template <> struct Vec16<double> { using type = double2; static constexpr int V = 2; };

// A row-major global matrix with an optional gather on its rows and an optional implicit
// column of ones appended after its last real column.
template <typename T> struct Operand {
    const T* ptr;
    long long ld;
    const int* row_idx;
    long long rows;   // valid rows
    int cols;         // valid (real) columns
    int ones;         // logical column `cols` reads as 1
    int vec_ok;       // base 16B aligned and ld % V == 0
};

// Loads one [MN][BK] logical tile into registers, then stores it k-major into shared memory.
// KC (k contiguous in global): global rows <-> mn index, global cols <-> k index.
// !KC: global rows <-> k index, global cols <-> mn index.
template <typename T, int MN, int BK, int NT, bool KC> struct TileLoader {
    static constexpr int V = Vec16<T>::V;
    static constexpr int ROWS = KC ? MN : BK;
    static constexpr int COLS = KC ? BK : MN;
    static constexpr int CV = COLS / V;
    static constexpr int NVEC = ROWS * CV;
    static constexpr int PER = (NVEC + NT - 1) / NT;
    static_assert(COLS % V == 0, "tile width must be a multiple of the vector width");
    T reg[PER][V];

    __device__ __forceinline__ void load(const Operand<T>& op, long long row0, long long col0,
                                         long long row_limit, long long col_limit, int tid) {
        using VT = typename Vec16<T>::type;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            int v = tid + i * NT;
            if (NVEC % NT != 0 && v >= NVEC) break;
            int r = v / CV, c = (v % CV) * V;
            long long gr = row0 + r, gc = col0 + c;
#pragma unroll
            for (int j = 0; j < V; ++j) reg[i][j] = T(0);
            if (gr < row_limit) {
                long long srow = op.row_idx ? (long long)op.row_idx[gr] : gr;
                const T* p = op.ptr + srow * op.ld + gc;
                if (op.vec_ok && gc + V <= op.cols && gc + V <= col_limit) {
                    VT q = *reinterpret_cast<const VT*>(p);
                    const T* qs = reinterpret_cast<const T*>(&q);
#pragma unroll
                    for (int j = 0; j < V; ++j) reg[i][j] = qs[j];
                } else {
#pragma unroll
                    for (int j = 0; j < V; ++j) {
                        long long cj = gc + j;
                        if (cj < col_limit) {
                            if (cj < op.cols) reg[i][j] = p[j];
                            else if (op.ones && cj == op.cols) reg[i][j] = T(1);
                        }
                    }
                }
            }
        }
    }

    template <int LDS> __device__ __forceinline__ void store(T (*S)[LDS], int tid) const {
        using VT = typename Vec16<T>::type;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            int v = tid + i * NT;
            if (NVEC % NT != 0 && v >= NVEC) break;
            int r = v / CV, c = (v % CV) * V;
            if (KC) {   // r = mn, c = k : transpose on the way in
#pragma unroll
                for (int j = 0; j < V; ++j) S[c + j][r] = reg[i][j];
            } else {    // r = k, c = mn : straight copy
                VT q;
                T* qs = reinterpret_cast<T*>(&q);
#pragma unroll
                for (int j = 0; j < V; ++j) qs[j] = reg[i][j];
                *reinterpret_cast<VT*>(&S[r][c]) = q;
            }
        }
    }
};

// C[M][N] (+)= A[M][Kd] * B[Kd][N] over k in [z*k_chunk, min(Kd, (z+1)*k_chunk)), z = blockIdx.z.
// Kd_a / Kd_b are the logical k extents of the two operands including an implicit ones index.
template <typename T, int BM, int BN, int BK, int TM, int TN, bool A_KC, bool B_KC, class Epi>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_simt_kernel(Operand<T> A, Operand<T> B, long long M, int N, long long Kd, long long k_chunk,
                 Epi epi) {
    constexpr int NT = (BM / TM) * (BN / TN);
    constexpr int V = Vec16<T>::V;
    constexpr int PAD = V;
    static_assert(TM % V == 0 || TM < V, "TM");
    __shared__ __align__(16) T As[2][BK][BM + PAD];
    __shared__ __align__(16) T Bs[2][BK][BN + PAD];

    const int tid = threadIdx.x;
    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    const long long m0 = (long long)blockIdx.y * BM;
    const int n0 = blockIdx.x * BN;
    const long long k_begin = (long long)blockIdx.z * k_chunk;
    const long long k_end = (k_begin + k_chunk < Kd) ? (k_begin + k_chunk) : Kd;
    const int nk = (int)((k_end - k_begin + BK - 1) / BK);

    TileLoader<T, BM, BK, NT, A_KC> la;
    TileLoader<T, BN, BK, NT, B_KC> lb;

    // limits: mn-dimension extents and the k bound of this split
    const long long a_mn_lim = A_KC ? A.rows : (long long)(A.cols + A.ones);
    const long long b_mn_lim = B_KC ? B.rows : (long long)(B.cols + B.ones);
    auto load_tiles = [&](long long k0) {
        if (A_KC) la.load(A, m0, k0, a_mn_lim, k_end, tid);
        else      la.load(A, k0, m0, (k_end < A.rows ? k_end : A.rows), a_mn_lim, tid);
        if (B_KC) lb.load(B, n0, k0, b_mn_lim, k_end, tid);
        else      lb.load(B, k0, n0, (k_end < B.rows ? k_end : B.rows), b_mn_lim, tid);
    };

    T acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = T(0);

    if (nk > 0) {
        load_tiles(k_begin);
        la.store(As[0], tid);
        lb.store(Bs[0], tid);
    }
    __syncthreads();

    for (int t = 0; t < nk; ++t) {
        const int cur = t & 1;
        if (t + 1 < nk) load_tiles(k_begin + (long long)(t + 1) * BK);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            T a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; ++i) a[i] = As[cur][k][ty * TM + i];
#pragma unroll
            for (int j = 0; j < TN; ++j) b[j] = Bs[cur][k][tx * TN + j];
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        if (t + 1 < nk) {
            la.store(As[cur ^ 1], tid);
            lb.store(Bs[cur ^ 1], tid);
        }
        __syncthreads();
    }

#pragma unroll
    for (int i = 0; i < TM; ++i) {
        long long m = m0 + ty * TM + i;
        if (m >= M) continue;
        int n = n0 + tx * TN;
        if (n >= N) continue;
        int nvalid = (N - n < TN) ? (N - n) : TN;
        epi.template row<TN>(m, n, acc[i], nvalid, (int)blockIdx.z);
    }
}

template <typename T, int TN>
__device__ __forceinline__ void store_row(T* p, const T (&v)[TN], int nvalid) {
    using VT = typename Vec16<T>::type;
    constexpr int V = Vec16<T>::V;
    if (TN % V == 0 && nvalid == TN && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
#pragma unroll
        for (int c = 0; c < TN; c += V) {
            VT q;
            T* qs = reinterpret_cast<T*>(&q);
#pragma unroll
            for (int j = 0; j < V; ++j) qs[j] = v[c + j];
            *reinterpret_cast<VT*>(p + c) = q;
        }
    } else {
#pragma unroll
        for (int j = 0; j < TN; ++j)
            if (j < nvalid) p[j] = v[j];
    }
}

// ---- epilogues --------------------------------------------------------------------------------

// keep flags of TN consecutive elements of one dropout point (all true when the point is off): one Philox evaluation
// for an aligned group of four (poz_common.cuh)
template <int TN>
__device__ __forceinline__ void drop_flags(const DropPoint& d, const DevState* st, long long m, int n, int ld, int nvalid,
                                           bool (&keep)[TN]) {
#pragma unroll
    for (int j = 0; j < TN; ++j) keep[j] = true;
    if (!d.enabled) return;
    if constexpr (TN == 4) {
        if (nvalid == 4) {
            drop_keep4(d, st, m, n, ld, keep);
            return;
        }
    }
#pragma unroll
    for (int j = 0; j < TN; ++j)
        if (j < nvalid) keep[j] = drop_keep(d, st, m, n + j, ld);
}

// forward: dropout on z, activation, dropout on a (neural_net_model.py:117-119,172-188,393-395)
template <typename T> struct FwdEpi {
    T* z_out;          // recorded (post-dropout) linear output, nullable
    T* a_out;          // recorded (post-dropout) activation output, nullable
    int ld;
    int act;
    DropPoint dz, da;
    const DevState* st;
    template <int TN>
    __device__ __forceinline__ void row(long long m, int n, const T (&acc)[TN], int nvalid, int) const {
        T z[TN], a[TN];
        bool kz[TN], ka[TN];
        drop_flags<TN>(dz, st, m, n, ld, nvalid, kz);
        drop_flags<TN>(da, st, m, n, ld, nvalid, ka);
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            T v = acc[j];
            if (dz.enabled && j < nvalid) v = kz[j] ? v * T(dz.scale) : T(0);
            z[j] = v;
            T w = act_fwd<T>(act, v);
            if (da.enabled && j < nvalid) w = ka[j] ? w * T(da.scale) : T(0);
            a[j] = w;
        }
        if (z_out) store_row<T, TN>(z_out + m * ld + n, z, nvalid);
        if (a_out) store_row<T, TN>(a_out + m * ld + n, a, nvalid);
    }
};

// backward-X: acc = gradient of the recorded output of the previous unit; push it through that
// unit's activation-dropout, activation and linear-dropout to get the gradient of its
// pre-dropout linear output (test_torch_backward.py:168-170).
template <typename T> struct BwdDataEpi {
    const T* a_prev;   // recorded output of the previous unit [M][ld]
    T* dz_prev;        // gradient w.r.t. the previous unit's pre-dropout linear output
    T* da_rec;         // diagnostics: gradient of the recorded activation output (nullable)
    T* dz_rec;         // diagnostics: gradient of the recorded linear output (nullable)
    int ld;
    int act;
    DropPoint dz, da;
    const DevState* st;
    // optional (float only): dz_prev also as the three bf16 planes hi + mid + lo a tensor-core weight-gradient launch
    // takes as its B operand (poz_tcs_kernel.cuh) -- [3][rows][ldp], pad columns untouched (zeroed at allocation)
    unsigned short* planes = nullptr;
    long long ldp = 0, plane_stride = 0;
    template <int TN>
    __device__ __forceinline__ void row(long long m, int n, const T (&acc)[TN], int nvalid, int) const {
        T g0[TN], g1[TN], g2[TN];
        bool kz[TN], ka[TN];
        drop_flags<TN>(dz, st, m, n, ld, nvalid, kz);
        drop_flags<TN>(da, st, m, n, ld, nvalid, ka);
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            T g = acc[j];
            g0[j] = g;
            T gz = T(0);
            if (j < nvalid) {
                T ga = g;
                T a = a_prev[m * ld + n + j];
                if (da.enabled) {
                    ga = ka[j] ? g * T(da.scale) : T(0);
                    a = a * T(da.keep_prob);
                }
                gz = ga * act_bwd_from_out<T>(act, a);
            }
            g1[j] = gz;
            if (dz.enabled && j < nvalid) gz = kz[j] ? gz * T(dz.scale) : T(0);
            g2[j] = gz;
        }
        if (da_rec) store_row<T, TN>(da_rec + m * ld + n, g0, nvalid);
        if (dz_rec) store_row<T, TN>(dz_rec + m * ld + n, g1, nvalid);
        store_row<T, TN>(dz_prev + m * ld + n, g2, nvalid);
        if (planes && TN == 4 && nvalid == 4) {
            // two packed conversions per plane, one 8-byte store per plane
            uint32_t ph[2], pm[2], pl[2];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const float x0 = (float)g2[2 * q], x1 = (float)g2[2 * q + 1];
                __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
                ph[q] = *reinterpret_cast<uint32_t*>(&h);
                const float r0 = x0 - __uint_as_float(ph[q] << 16), r1 = x1 - __uint_as_float(ph[q] & 0xffff0000u);
                __nv_bfloat162 mm = __floats2bfloat162_rn(r0, r1);
                pm[q] = *reinterpret_cast<uint32_t*>(&mm);
                const float s0 = r0 - __uint_as_float(pm[q] << 16), s1 = r1 - __uint_as_float(pm[q] & 0xffff0000u);
                __nv_bfloat162 ll = __floats2bfloat162_rn(s0, s1);
                pl[q] = *reinterpret_cast<uint32_t*>(&ll);
            }
            unsigned short* o = planes + m * ldp + n;
            *reinterpret_cast<uint2*>(o) = make_uint2(ph[0], ph[1]);
            *reinterpret_cast<uint2*>(o + plane_stride) = make_uint2(pm[0], pm[1]);
            *reinterpret_cast<uint2*>(o + 2 * plane_stride) = make_uint2(pl[0], pl[1]);
        } else if (planes) {
#pragma unroll
            for (int j = 0; j < TN; ++j)
                if (j < nvalid) {
                    const float xv = (float)g2[j];
                    const __nv_bfloat16 h = __float2bfloat16_rn(xv);
                    const float r1 = xv - __bfloat162float(h);
                    const __nv_bfloat16 mm = __float2bfloat16_rn(r1);
                    const __nv_bfloat16 ll = __float2bfloat16_rn(r1 - __bfloat162float(mm));
                    unsigned short* o = planes + m * ldp + n + j;
                    o[0] = __bfloat16_as_ushort(h);
                    o[plane_stride] = __bfloat16_as_ushort(mm);
                    o[2 * plane_stride] = __bfloat16_as_ushort(ll);
                }
        }
    }
};

// backward-W: each batch split writes its partial [in+1][out] image into its own slab.
template <typename T> struct BwdWeightEpi {
    T* g;                  // flat gradient slabs
    long long slab;        // elements per slab (total parameter count)
    long long off;         // offset of this layer's [W; b] image inside a slab
    int ld;                // out features
    template <int TN>
    __device__ __forceinline__ void row(long long m, int n, const T (&acc)[TN], int nvalid, int z) const {
        store_row<T, TN>(g + (long long)z * slab + off + m * ld + n, acc, nvalid);
    }
};

}  // namespace poz
