// fp32 SIMT GEMM with separable-address operand loaders (dense, implicit-GEMM convolution,
// convolution weight-gradient) and a functor epilogue.  This is the exact-precision
// ("fp32 parity", 1e-5 vs the reference) contraction engine; the bf16 tcgen05 path lives in
// tc_gemm.cu.
//
//   C[m,n] = sum_k A(m,k) * B(n,k),   A(m,k) = a.fetch(a.mn_off(m) + a.k_off(k))
//
// A loader provides: mn_off(i) -> element offset of row i (or -1 if out of range),
// k_off(k) -> element offset of reduction index k, fetch(off) -> float, and K_FAST: whether
// consecutive k (true) or consecutive rows (false) are adjacent in memory, which selects the
// thread->element staging order so that global loads coalesce.
#pragma once
#include "common.cuh"

namespace r2d2 {

template <int BM, int BN, int TM, int TN, class AL, class BL, class EP>
__global__ void __launch_bounds__((BM / TM) * (BN / TN))
gemm_kernel(const AL al, const BL bl, const EP ep, const int M, const int N, const int K,
            const int k_per_split) {
    constexpr int BK = 16;
    constexpr int NT = (BM / TM) * (BN / TN);
    constexpr int PA = BM + 4, PB = BN + 4;
    constexpr int A_EL = BM * BK / NT, B_EL = BN * BK / NT;
    static_assert(TM % 4 == 0 && TN % 4 == 0, "thread tile must be built from float4 groups");
    static_assert((BM * BK) % NT == 0 && (BN * BK) % NT == 0, "staging must divide evenly");
    static_assert(NT % BK == 0, "K-fast staging needs NT % 16 == 0");
    static_assert(AL::K_FAST || NT % BM == 0, "row-fast staging of A needs NT % BM == 0");
    static_assert(BL::K_FAST || NT % BN == 0, "row-fast staging of B needs NT % BN == 0");

    __shared__ __align__(16) float As[2][BK][PA];
    __shared__ __align__(16) float Bs[2][BK][PB];
    __shared__ long long rowA[BM];
    __shared__ long long rowB[BN];

    const int tid = threadIdx.x;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int kbeg = blockIdx.z * k_per_split;
    const int kend = min(K, kbeg + k_per_split);

    for (int i = tid; i < BM; i += NT) rowA[i] = (m0 + i < M) ? al.mn_off(m0 + i) : -1;
    for (int i = tid; i < BN; i += NT) rowB[i] = (n0 + i < N) ? bl.mn_off(n0 + i) : -1;
    __syncthreads();

    float ra[A_EL], rb[B_EL];
    // row-fast staging: this thread's row is fixed for the whole kernel
    const long long myrowA = AL::K_FAST ? 0 : rowA[tid % BM];
    const long long myrowB = BL::K_FAST ? 0 : rowB[tid % BN];

    auto gload = [&](int kt) {
        if (AL::K_FAST) {
            const int k = kt + (tid % BK);
            const long long ko = (k < kend) ? al.k_off(k) : -1;
#pragma unroll
            for (int j = 0; j < A_EL; ++j) {
                const long long ro = rowA[tid / BK + j * (NT / BK)];
                ra[j] = (ko >= 0 && ro >= 0) ? al.fetch(ro + ko) : 0.f;
            }
        } else {
#pragma unroll
            for (int j = 0; j < A_EL; ++j) {
                const int k = kt + tid / BM + j * (NT / BM);
                ra[j] = (k < kend && myrowA >= 0) ? al.fetch(myrowA + al.k_off(k)) : 0.f;
            }
        }
        if (BL::K_FAST) {
            const int k = kt + (tid % BK);
            const long long ko = (k < kend) ? bl.k_off(k) : -1;
#pragma unroll
            for (int j = 0; j < B_EL; ++j) {
                const long long ro = rowB[tid / BK + j * (NT / BK)];
                rb[j] = (ko >= 0 && ro >= 0) ? bl.fetch(ro + ko) : 0.f;
            }
        } else {
#pragma unroll
            for (int j = 0; j < B_EL; ++j) {
                const int k = kt + tid / BN + j * (NT / BN);
                rb[j] = (k < kend && myrowB >= 0) ? bl.fetch(myrowB + bl.k_off(k)) : 0.f;
            }
        }
    };
    auto sstore = [&](int buf) {
#pragma unroll
        for (int j = 0; j < A_EL; ++j) {
            if (AL::K_FAST) As[buf][tid % BK][tid / BK + j * (NT / BK)] = ra[j];
            else As[buf][tid / BM + j * (NT / BM)][tid % BM] = ra[j];
        }
#pragma unroll
        for (int j = 0; j < B_EL; ++j) {
            if (BL::K_FAST) Bs[buf][tid % BK][tid / BK + j * (NT / BK)] = rb[j];
            else Bs[buf][tid / BN + j * (NT / BN)][tid % BN] = rb[j];
        }
    };

    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    if (kbeg < kend) {
        gload(kbeg);
        sstore(0);
        __syncthreads();
        int buf = 0;
        for (int kt = kbeg; kt < kend; kt += BK) {
            const bool more = kt + BK < kend;
            if (more) gload(kt + BK);
#pragma unroll
            for (int kk = 0; kk < BK; ++kk) {
                float a[TM], b[TN];
#pragma unroll
                for (int g = 0; g < TM / 4; ++g) {
                    const float4 v = *reinterpret_cast<const float4*>(
                        &As[buf][kk][g * (BM / TM) * 4 + ty * 4]);
                    a[g * 4 + 0] = v.x; a[g * 4 + 1] = v.y; a[g * 4 + 2] = v.z; a[g * 4 + 3] = v.w;
                }
#pragma unroll
                for (int g = 0; g < TN / 4; ++g) {
                    const float4 v = *reinterpret_cast<const float4*>(
                        &Bs[buf][kk][g * (BN / TN) * 4 + tx * 4]);
                    b[g * 4 + 0] = v.x; b[g * 4 + 1] = v.y; b[g * 4 + 2] = v.z; b[g * 4 + 3] = v.w;
                }
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
            if (more) {
                sstore(buf ^ 1);
                __syncthreads();
                buf ^= 1;
            }
        }
    }

#pragma unroll
    for (int gi = 0; gi < TM / 4; ++gi)
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
            const int m = m0 + gi * (BM / TM) * 4 + ty * 4 + ii;
            if (m >= M) continue;
#pragma unroll
            for (int gj = 0; gj < TN / 4; ++gj)
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    const int n = n0 + gj * (BN / TN) * 4 + tx * 4 + jj;
                    if (n < N) ep(m, n, acc[gi * 4 + ii][gj * 4 + jj], (int)blockIdx.z);
                }
        }
}

// ---------------------------------------------------------------- loaders
// dense: element (i,k) at base[i*ld_i + k*ld_k]
template <bool KFAST>
struct DenseLoad {
    static constexpr bool K_FAST = KFAST;
    const float* base;
    long long ld_i, ld_k;
    __device__ __forceinline__ long long mn_off(int i) const { return (long long)i * ld_i; }
    __device__ __forceinline__ long long k_off(int k) const { return (long long)k * ld_k; }
    __device__ __forceinline__ float fetch(long long o) const { return __ldg(base + o); }
};

// Frame addressing shared by the conv loaders: output row r = t*B + b maps to input frame
// (b*S + s0 + t) of the slab (u8) or simply frame r (dense f32 activations: B=0).
struct FrameMap {
    int B, S, s0;
    long long frame_elems;
    // ring mode (seq != null): frame (t, b) is ring row (seq[b] + s0 + t) mod M -- the sampled
    // sequences are read in place from the replay ring, no gathered slab
    const long long* seq = nullptr;
    long long M = 0;
    __device__ __forceinline__ long long base(int r) const {
        if (B == 0) return (long long)r * frame_elems;
        const int t = r / B, b = r - t * B;
        if (seq) {
            long long row = __ldg(seq + b) + s0 + t;
            row %= M;
            return row * frame_elems;
        }
        return ((long long)b * S + s0 + t) * frame_elems;
    }
};

// implicit-GEMM convolution input: row m = (frame r, out pixel p), k = (ci, ky, kx)
template <class T, int CIN, int KH, int HIN, int HOUT, int STRIDE>
struct ConvPatchRows {
    static constexpr bool K_FAST = true;
    const T* base;
    FrameMap fm;
    __device__ __forceinline__ long long mn_off(int m) const {
        const int r = m / (HOUT * HOUT), p = m - r * (HOUT * HOUT);
        const int oy = p / HOUT, ox = p - oy * HOUT;
        return fm.base(r) + (long long)(oy * STRIDE) * HIN + ox * STRIDE;
    }
    __device__ __forceinline__ long long k_off(int k) const {
        const int ci = k / (KH * KH), rem = k - ci * (KH * KH);
        const int ky = rem / KH, kx = rem - ky * KH;
        return (long long)ci * HIN * HIN + ky * HIN + kx;
    }
    __device__ __forceinline__ float fetch(long long o) const;
};
template <class T, int CIN, int KH, int HIN, int HOUT, int STRIDE>
__device__ __forceinline__ float ConvPatchRows<T, CIN, KH, HIN, HOUT, STRIDE>::fetch(long long o) const {
    if constexpr (sizeof(T) == 1) return __fdiv_rn((float)base[o], 255.f);
    else return __ldg(base + o);
}

// weight-gradient operands: reduction index kk = (frame r, out pixel p)
//   A(m=co, kk) = dZ[r, co, p]                    (dense NCHW gradient)
//   B(n=(ci,ky,kx), kk) = in[r, ci, oy*s+ky, ox*s+kx]
template <int COUT, int HOUT>
struct GradRowsT {
    static constexpr bool K_FAST = true;
    const float* base;
    __device__ __forceinline__ long long mn_off(int co) const { return (long long)co * HOUT * HOUT; }
    __device__ __forceinline__ long long k_off(int kk) const {
        const int r = kk / (HOUT * HOUT), p = kk - r * (HOUT * HOUT);
        return (long long)r * COUT * HOUT * HOUT + p;
    }
    __device__ __forceinline__ float fetch(long long o) const { return __ldg(base + o); }
};
template <class T, int CIN, int KH, int HIN, int HOUT, int STRIDE>
struct ConvPatchCols {
    static constexpr bool K_FAST = true;
    const T* base;
    FrameMap fm;
    __device__ __forceinline__ long long mn_off(int n) const {
        const int ci = n / (KH * KH), rem = n - ci * (KH * KH);
        const int ky = rem / KH, kx = rem - ky * KH;
        return (long long)ci * HIN * HIN + ky * HIN + kx;
    }
    __device__ __forceinline__ long long k_off(int kk) const {
        const int r = kk / (HOUT * HOUT), p = kk - r * (HOUT * HOUT);
        const int oy = p / HOUT, ox = p - oy * HOUT;
        return fm.base(r) + (long long)(oy * STRIDE) * HIN + ox * STRIDE;
    }
    __device__ __forceinline__ float fetch(long long o) const {
        if constexpr (sizeof(T) == 1) return __fdiv_rn((float)base[o], 255.f);
        else return __ldg(base + o);
    }
};

// ---------------------------------------------------------------- epilogues
// y[m*ldc + n] = acc (+ y) + b1[n] + b2[n], optional ReLU
struct EpiDense {
    float* c;
    long long ldc;
    const float* b1;
    const float* b2;
    int relu, accumulate;
    __device__ __forceinline__ void operator()(int m, int n, float v, int) const {
        float* p = c + (long long)m * ldc + n;
        if (b1) v += __ldg(b1 + n);
        if (b2) v += __ldg(b2 + n);
        if (accumulate) v += *p;
        if (relu) v = fmaxf(v, 0.f);
        *p = v;
    }
};
// split-K partials: part[z][m*ldc + n]
struct EpiPartial {
    float* part;
    long long ldc, slab;
    __device__ __forceinline__ void operator()(int m, int n, float v, int z) const {
        part[(long long)z * slab + (long long)m * ldc + n] = v;
    }
};
// conv forward: row m = (r,p), n = co -> NCHW out[r, co, p], bias + ReLU
template <int COUT, int PIX>
struct EpiConv {
    float* out;
    const float* bias;
    __device__ __forceinline__ void operator()(int m, int n, float v, int) const {
        const int r = m / PIX, p = m - r * PIX;
        out[((long long)r * COUT + n) * PIX + p] = fmaxf(v + __ldg(bias + n), 0.f);
    }
};

template <int BM, int BN, int TM, int TN, class AL, class BL, class EP>
static inline cudaError_t launch_gemm(const char* name, const AL& al, const BL& bl, const EP& ep,
                                      int M, int N, int K, int splits, cudaStream_t st) {
    constexpr int NT = (BM / TM) * (BN / TN);
    int kps = (K + splits - 1) / splits;
    kps = (kps + 15) / 16 * 16;
    dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, splits);
    {
        ProfScope ps(name, st);
        gemm_kernel<BM, BN, TM, TN, AL, BL, EP><<<grid, NT, 0, st>>>(al, bl, ep, M, N, K, kps);
    }
    return cudaGetLastError();
}

}  // namespace r2d2
