// Persistent LSTM unroll on tcgen05 tensor cores (model.py:52-60), one cooperative launch for all
// T steps of a chain.
//
// Swap-AB: per step  G^T[4H, B] = W_hh[4H, H] . h_{t-1}^T[H, B]  (+ xproj^T).  CTA c owns 16 hidden
// units and all four of their gates = 64 rows of W_hh (UMMA M = 64, N = 64 batch columns); its
// W_hh slice stays resident in shared memory for the whole unroll as bf16 (hi, lo) pairs, the
// cell state c lives in registers, and only h_{t-1} travels: every CTA writes its 16 units of
// h_t (bf16 hi/lo, K-major) to a ping-pong buffer in global memory, a grid-wide arrive/wait
// counter publishes it, and TMA streams it back in as the B operand of the next step.
// Forward contractions are error-compensated bf16x3 (W_hi.h_hi + W_lo.h_hi + W_hi.h_lo, fp32
// accumulation in TMEM) -- see DESIGN.md section 2 for why the forward pass needs this.  The
// (hi, lo) halves of W are stacked along M (UMMA M = 128) and, at NB = 32, those of h along N, so
// one MMA per k-step produces all three terms.
//
// 416 threads: warp 0 = bulk-copy producer, warps 1,10,11,12 = MMA issuers (warp 1 also owns the
// TMEM allocation), warps 2-5 = TMEM readers (lane quarter q = warp % 4 = accumulator rows
// 32q..32q+31: two gates x 16 units of the W_hi rows for q < 2, of the W_lo rows for q >= 2),
// warps 2-9 = pointwise cell update (thread -> batch row, 2-4 units).
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace r2d2 {

constexpr int LS_UNITS = 16, LS_ROWS = 64, LS_NB = 64;   // LS_NB: backward batch tile

__device__ __forceinline__ uint32_t ls_pack(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float ls_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ uint32_t ls_sw128(int row, int chunk) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}
// sigmoid / tanh through ex2.approx + rcp.approx: absolute error ~1e-7 (inside the 1e-5 class),
// a fraction of the instructions of the libm-accurate versions on the per-step critical path
__device__ __forceinline__ float ls_sigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float ls_tanh(float x) { return 1.f - __fdividef(2.f, __expf(2.f * x) + 1.f); }
__device__ __forceinline__ void red_release_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

struct LstmArgs {
    const float* xprojT;   // [4H, ldx]: column col0 + t*B + b holds step t, batch b
    long long ldx, col0;
    const float* w_hh;     // [4H, H] fp32
    float* hs_ext;         // [(T+1), B, H]; slot 0 = h0 (read), slot t+1 written
    float* cs_ext;         // same for c
    float* gates;          // [T, B, 4H] post-activation i,f,g,o or null
    __nv_bfloat16* h16;    // [2, H/64, groups, hi|lo, NB, 64] ping-pong, K-major B operand laid out
                           // so that the (hi, lo) tiles of one k-block and batch group are ONE
                           // contiguous 2*NB*128-byte run = one TMA box = one pipeline stage
    int* counter;          // [gridDim.y] grid barrier counters, zero on entry
    int T, B, H;
    long long* trace;      // optional [T, 8] globaltimer stamps of CTA (0,0) (diagnostics), or null
    __nv_bfloat16* hist_hi;   // optional [T - hist_t0, B, H] bf16 (hi, lo) split of h_t for t >= hist_t0:
    __nv_bfloat16* hist_lo;   //   the operand image the dueling-head GEMM consumes (no split pass)
    int hist_t0;
};

__device__ __forceinline__ long long gtimer() {
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#define LS_STAMP(slot)                                                        \
    do {                                                                      \
        if (a.trace && blockIdx.x == 0 && blockIdx.y == 0) a.trace[t * 32 + (slot)] = gtimer(); \
    } while (0)

template <int N> __device__ __forceinline__ void st_f32(float* p, const float* v) {
    if constexpr (N == 4) *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    else if constexpr (N == 2) *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]);
    else *p = v[0];
}
template <int N> __device__ __forceinline__ void st_bf16_hilo(__nv_bfloat16* hi, __nv_bfloat16* lo,
                                                              const float* v) {
    if constexpr (N == 4) {
        uint2 a, b;
        split_bf16x2(v[0], v[1], a.x, b.x);
        split_bf16x2(v[2], v[3], a.y, b.y);
        *reinterpret_cast<uint2*>(hi) = a;
        *reinterpret_cast<uint2*>(lo) = b;
    } else if constexpr (N == 2) {
        uint32_t a, b;
        split_bf16x2(v[0], v[1], a, b);
        *reinterpret_cast<uint32_t*>(hi) = a;
        *reinterpret_cast<uint32_t*>(lo) = b;
    } else {
        *hi = __float2bfloat16_rn(v[0]);
        *lo = __float2bfloat16_rn(v[0] - ls_round(v[0]));
    }
}

// NB = batch columns per CTA (UMMA N).  Batch rows never interact, so the B/NB batch groups are
// independent chains with their own arrive counter: a smaller NB spreads one chain over more SMs
// (less h traffic, MMA time and pointwise work per CTA per step).
constexpr int LS_ISSUERS = 4;   // MMA-issuing warps: 1, 10, 11, 12 (load groups g % 4 == i)

template <int NB>
__global__ void __launch_bounds__(416, 1)
tc_lstm_fwd_kernel(const LstmArgs a) {
    // h ring: 64 KB = 4 slots of 16 KB.  A slot is a LOAD GROUP of KBG consecutive k-blocks
    // ((hi|lo) tiles, NB rows x 128 B each), contiguous and already swizzled in global memory, so
    // it arrives with ONE 1-D bulk copy on ONE mbarrier: every mbarrier operation of the
    // producer thread costs 30-250 ns on this critical path (tools/ubench/tma_fanin.cu).
    // NB = 32: the ring holds the whole h of a step (H <= 512), so a slot is never reused within
    // a step and the grid barrier already proves it free: no "slot empty" barriers at all.
    constexpr int KBG = NB == 32 ? 2 : 1, STAGES = 4;
    constexpr bool WHOLE = NB == 32;
    constexpr int UPT = LS_UNITS * NB / 256;                    // hidden units per pointwise thread
    // W_hi multiplies both h_hi and h_lo: with NB = 32 they are ONE MMA of N = 64 against the
    // adjacent [h_hi ; h_lo] tiles of a stage (2 MMAs per k-step instead of 3; the tensor pipe and
    // the issuing thread pay per instruction at these sizes).  Slice = SW accumulator columns.
    constexpr bool CAT = NB == 32;
    constexpr int SW = CAT ? 2 * NB : NB;
    static_assert(NB == 64 || NB == 32, "batch tile must be 32 or 64");
    const int H = a.H, B = a.B, NKB = H / 64;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full, acc_empty;
    __shared__ uint32_t tmem_slot;
    __shared__ __align__(16) float G[4][LS_UNITS][NB + 4];      // raw gate pre-activations (no xproj)

    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const uint32_t W_KB = LS_ROWS * 128;                        // 64 rows x 64 bf16
    uint8_t* w_s = base;                                        // [NKB][W_hi 8 KB | W_lo 8 KB]
    uint8_t* h_s = base + 2 * NKB * W_KB;                       // STAGES x (hi | lo)
    constexpr uint32_t H_TILE = NB * 128, STAGE = 2 * H_TILE;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int u0 = blockIdx.x * LS_UNITS, b0 = blockIdx.y * NB;
    const int ncta = gridDim.x;
    int* counter = a.counter + blockIdx.y;
    const long long BH = (long long)B * H;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_full, LS_ISSUERS);
        mbar_init(&acc_empty, 128);
        mbar_fence_init();
    }
    // resident W_hh slice: row = gate*16 + u  <->  W_hh[gate*H + u0 + u, :], split into hi/lo
    for (int i = tid; i < LS_ROWS * (H / 8); i += blockDim.x) {
        const int row = i / (H / 8), ch = i - row * (H / 8);
        const int kb = ch >> 3, c = ch & 7;
        const float* src = a.w_hh + (long long)((row >> 4) * H + u0 + (row & 15)) * H + ch * 8;
        const float4 x0 = __ldg(reinterpret_cast<const float4*>(src));
        const float4 x1 = __ldg(reinterpret_cast<const float4*>(src) + 1);
        const float v[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
        float l[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) l[e] = v[e] - ls_round(v[e]);
        // per k-block ONE 128-row K-major tile: rows 0..63 = W_hi, rows 64..127 = W_lo
        const uint32_t off = (uint32_t)kb * 2 * W_KB + ls_sw128(row, c);
        *reinterpret_cast<uint4*>(w_s + off) =
            make_uint4(ls_pack(v[0], v[1]), ls_pack(v[2], v[3]), ls_pack(v[4], v[5]), ls_pack(v[6], v[7]));
        *reinterpret_cast<uint4*>(w_s + W_KB + off) =
            make_uint4(ls_pack(l[0], l[1]), ls_pack(l[2], l[3]), ls_pack(l[4], l[5]), ls_pack(l[6], l[7]));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 1) tmem_alloc(&tmem_slot, LS_ISSUERS * SW);   // one accumulator slice per issuer
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    const int issuer = warp == 1 ? 0 : (warp >= 10 ? warp - 9 : -1);

    // pointwise mapping (warps 2..9): batch row pb and UPT consecutive hidden units; the
    // LS_UNITS/UPT threads of a batch row are neighbours, so a warp's store covers whole 64-byte
    // runs (16 units of one row) instead of 32 different rows
    const int pt = tid - 64;                      // 0..255 for pointwise threads
    constexpr int TPR = LS_UNITS / UPT;           // threads per batch row
    const int pb = pt / TPR, pu = (pt % TPR) * UPT;
    // offset of (row b0+pb, units u0+pu.., hi) inside one [groups, H/64, hi|lo, NB, 64] bf16
    // image; the 16-byte chunks of a 128-byte row are stored in SWIZZLE_128B order (chunk ^ row%8)
    // so that a linear bulk copy lands the tile exactly as the UMMA descriptor expects it
    const int col16 = (u0 & 63) + pu;
    const long long GRP = (long long)NKB * 2 * NB * 64;          // elements per batch group
    const long long o16r = (long long)blockIdx.y * GRP + ((long long)(u0 >> 6) * 2 * NB + pb) * 64 +
                           ((((col16 >> 3) ^ (pb & 7)) << 3) | (col16 & 7));
    const long long IMG = GRP * gridDim.y;
    __nv_bfloat16* const h16_lo = a.h16 + NB * 64;
    const bool pw = warp >= 2 && warp < 10;
    const bool pvalid = pw && (b0 + pb) < B;
    float c_reg[UPT];
#pragma unroll
    for (int j = 0; j < UPT; ++j) c_reg[j] = 0.f;

    // ---- prologue: publish h0 as bf16 (hi, lo) into ping-pong buffer 0; load c0
    if (pvalid) {
        const long long o = (long long)(b0 + pb) * H + u0 + pu;
        float hv[UPT];
#pragma unroll
        for (int j = 0; j < UPT; ++j) {
            hv[j] = a.hs_ext[o + j];
            c_reg[j] = a.cs_ext[o + j];
        }
        st_bf16_hilo<UPT>(a.h16 + o16r, h16_lo + o16r, hv);
    }
    if (pw) {
        // CTA barrier orders every thread's stores before thread 0's gpu-scope RELEASE, which is
        // cumulative: one releasing arrive per CTA instead of 256 thread fences
        named_bar_sync(1, 256);
        if (pt == 0) red_release_add(counter, 1);
    }

    uint32_t it = 0;   // ring iteration counter (producer and MMA warps keep identical copies)
    for (int t = 0; t < a.T; ++t) {
        const int p = t & 1;
        if (warp == 0) {
            if (lane == 0) {
                // wait until every CTA of this batch group has published h_{t-1}
                const int target = ncta * (t + 1);
                uint32_t spins = 0;
                while (ld_acquire(counter) < target)
                    if (++spins > (1u << 28)) __trap();
                LS_STAMP(0);                                        // barrier passed
                // generic-proxy global writes of the other CTAs -> async-proxy (bulk copy) reads; the
                // .global form is enough (and 0.35 us/step cheaper than the all-spaces fence)
                asm volatile("fence.proxy.async.global;" ::: "memory");
                LS_STAMP(1);
                const uint8_t* img = reinterpret_cast<const uint8_t*>(a.h16 + (long long)p * IMG + (long long)blockIdx.y * GRP);
                for (int kb = 0; kb < NKB; kb += KBG, ++it) {
                    const int s = it % STAGES;
                    const int nk = NKB - kb < KBG ? NKB - kb : KBG;
                    if constexpr (!WHOLE) mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_bar[s], nk * STAGE);
                    bulk_g2s(h_s + (size_t)s * (KBG * STAGE), img + (size_t)kb * STAGE, nk * STAGE, &full_bar[s]);
                }
            }
        } else if (issuer >= 0) {
            if (lane == 0) {
                // Four issuing warps, load groups g % 4 == issuer, each accumulating into its own
                // TMEM slice (the readers add the slices): issuing all the small MMAs of a step
                // from one thread was the critical path (37 ns per MMA, tools/diag_lstm.py).
                // A = [W_hi ; W_lo] stacked along M (128 rows): ONE MMA per k-step gives W_hi.[h_hi|h_lo]
                // in accumulator rows 0..63 and W_lo.h_hi in rows 64..127 (their W_lo.h_lo columns are
                // ~2^-16 of the result and simply added in).  The tensor pipe spends the same ~20 ns
                // on an MMA this small whatever its M, so halving the count halves the MMA phase.
                constexpr uint32_t idesc = umma_idesc_bf16(2 * LS_ROWS, 64);
                static_assert(SW == 64, "accumulator slice = 64 columns");
                mbar_wait(&acc_empty, (t & 1) ^ 1);                // readers drained step t-1
                tc_fence_after();
                const uint64_t dw = umma_desc_k_sw128(smem_u32(w_s));
                const uint64_t dh0 = umma_desc_k_sw128(smem_u32(h_s));
                const uint32_t d = tmem_base + issuer * SW;
                const int ngrp = (NKB + KBG - 1) / KBG;             // load groups per step
                const uint32_t it0 = (uint32_t)t * ngrp;            // ring position of group 0
                for (int g = issuer; g < ngrp; g += LS_ISSUERS) {
                    const uint32_t itk = it0 + g;
                    const int s = itk % STAGES;
                    mbar_wait(&full_bar[s], (itk / STAGES) & 1);
                    tc_fence_after();
#pragma unroll
                    for (int j = 0; j < KBG; ++j) {
                        const int kb = g * KBG + j;
                        if (kb < NKB) {
                            const uint64_t hhi = dh0 + (uint64_t)(((s * KBG + j) * STAGE) >> 4), hlo = hhi + (H_TILE >> 4);
                            const uint64_t wk = dw + (uint64_t)((kb * 2 * W_KB) >> 4);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const uint32_t acc = (g >= LS_ISSUERS) | (j != 0) | (k != 0);
                                if constexpr (CAT) {
                                    umma_f16(d, wk + 2 * k, hhi + 2 * k, idesc, acc);      // B = [h_hi ; h_lo], 64 rows
                                } else {
                                    umma_f16(d, wk + 2 * k, hhi + 2 * k, idesc, acc);
                                    umma_f16(d, wk + 2 * k, hlo + 2 * k, idesc, 1);
                                }
                            }
                        }
                    }
                    if constexpr (!WHOLE) umma_commit(&empty_bar[s]);
                    if (issuer == 0) LS_STAMP(16 + g);
                }
                if (issuer == 0) LS_STAMP(2);
                umma_commit(&acc_full);                             // LS_ISSUERS arrivals complete it
            }
        } else {
            // ---- pointwise threads: prefetch this step's input projection (independent of the MMA)
            float xp[4][UPT];
            if (pvalid) {
                const float* xc = a.xprojT + a.col0 + (long long)t * B + b0 + pb;
#pragma unroll
                for (int g = 0; g < 4; ++g)
#pragma unroll
                    for (int j = 0; j < UPT; ++j)
                        xp[g][j] = __ldg(xc + (long long)(g * H + u0 + pu + j) * a.ldx);
            }
            if (warp < 6) {
                // ---- TMEM readers: warp with lane quarter q = warp % 4 owns accumulator rows
                //      32q..32q+31: q < 2 the W_hi rows of gates 2q, 2q+1 (16 units each), q >= 2 the
                //      W_lo rows of gates 2(q-2), 2(q-2)+1; NB batch columns per row
                const int q = warp & 3;
                const int gate = 2 * (q & 1) + (lane >> 4), unit = lane & 15;
                // one thread polls the accumulator barrier, the other readers sleep on a hardware
                // barrier: pollers take issue slots from the TMA and MMA issuing threads
                if (tid == 64) {
                    mbar_wait_sleep(&acc_full, t & 1, 32);
                    LS_STAMP(3);                                    // accumulator complete
                }
                named_bar_sync(3, 128);
                tc_fence_after();
                const int ngrp_r = (NKB + KBG - 1) / KBG;
                const int nslices = ngrp_r < LS_ISSUERS ? ngrp_r : LS_ISSUERS;   // issuers that ran MMAs
                float acc[NB];
#pragma unroll
                for (int j = 0; j < NB; ++j) acc[j] = 0.f;
#pragma unroll
                for (int i = 0; i < LS_ISSUERS; ++i) {
                    if (i < nslices) {
                        // both halves of the slice in flight before the one wait
                        static_assert(SW == 64, "two 32-column loads per slice");
                        uint32_t r0[32], r1[32];
                        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + i * SW, r0);
                        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + i * SW + 32, r1);
                        tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            acc[j % NB] += __uint_as_float(r0[j]);
                            acc[(32 + j) % NB] += __uint_as_float(r1[j]);
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&acc_empty);
                // W_hi rows store, then the W_lo rows of the same (gate, unit) add on top
                if (q < 2) {
#pragma unroll
                    for (int j = 0; j < NB; j += 4)
                        *reinterpret_cast<float4*>(&G[gate][unit][j]) =
                            make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
                }
                named_bar_sync(5, 128);
                if (q >= 2) {
#pragma unroll
                    for (int j = 0; j < NB; j += 4) {
                        float4 v = *reinterpret_cast<float4*>(&G[gate][unit][j]);
                        v.x += acc[j]; v.y += acc[j + 1]; v.z += acc[j + 2]; v.w += acc[j + 3];
                        *reinterpret_cast<float4*>(&G[gate][unit][j]) = v;
                    }
                }
            }
            named_bar_sync(1, 256);                                 // G complete
            if (pt == 0) LS_STAMP(4);
            float hv[UPT], cv[UPT], gi[UPT], gf[UPT], gg[UPT], go[UPT];
            if (pvalid) {
#pragma unroll
                for (int j = 0; j < UPT; ++j) {
                    gi[j] = ls_sigmoid(G[0][pu + j][pb] + xp[0][j]);
                    gf[j] = ls_sigmoid(G[1][pu + j][pb] + xp[1][j]);
                    gg[j] = ls_tanh(G[2][pu + j][pb] + xp[2][j]);
                    go[j] = ls_sigmoid(G[3][pu + j][pb] + xp[3][j]);
                    cv[j] = gf[j] * c_reg[j] + gi[j] * gg[j];
                    hv[j] = go[j] * ls_tanh(cv[j]);
                    c_reg[j] = cv[j];
                }
                // what the next step waits for goes out first and alone: the releasing arrive
                // below has to wait for every store issued before it
                const long long o16 = (long long)((t + 1) & 1) * IMG + o16r;
                st_bf16_hilo<UPT>(a.h16 + o16, h16_lo + o16, hv);
            }
            if (pt == 0) LS_STAMP(5);                               // pointwise + h stores issued
            named_bar_sync(2, 256);                                 // all of this CTA's h_t is out
            if (pt == 0) {
                LS_STAMP(6);
                red_release_add(counter, 1);
                LS_STAMP(7);
            }
            // fp32 history for the kernels after the unroll (nobody inside it reads these)
            if (pvalid) {
                const long long o = (long long)(b0 + pb) * H + u0 + pu;
                st_f32<UPT>(a.hs_ext + (long long)(t + 1) * BH + o, hv);
                st_f32<UPT>(a.cs_ext + (long long)(t + 1) * BH + o, cv);
                if (a.hist_hi && t >= a.hist_t0) {
                    const long long oh = (long long)(t - a.hist_t0) * BH + o;
                    st_bf16_hilo<UPT>(a.hist_hi + oh, a.hist_lo + oh, hv);
                }
                if (a.gates) {
                    float* gp = a.gates + ((long long)t * B + b0 + pb) * 4 * H + u0 + pu;
                    st_f32<UPT>(gp, gi);
                    st_f32<UPT>(gp + H, gf);
                    st_f32<UPT>(gp + 2 * H, gg);
                    st_f32<UPT>(gp + 3 * H, go);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, LS_ISSUERS * SW);
}

// ------------------------------------------------------------------------------------------
// BPTT over the learn window, persistent, plain bf16 operands (backward contractions do not
// need the compensation).  Per step (t descending)
//   dh[B, u] = dhs_t[B, u] + dG_{t+1}[B, 4H] . W_hh[4H, u]
// with CTA c owning 16 hidden units u: A = dG_{t+1} (M = 64 batch rows, pulled in by four bulk
// copies per step from a bf16 ping-pong image every CTA writes its columns of), B = its W_hh^T
// slice [16, 4H] resident in shared memory, D[64, 16] in TMEM.  In the M=64 accumulator layout
// batch row m sits in lane (m%16) + 32*(m/16): those lanes read the recurrent term and hand it
// over through shared memory, then all 128 reader threads do the gate-gradient math (thread ->
// batch row, 4 or 8 units; dL/dc stays in registers across steps).
struct LstmBwdArgs {
    const float* dhs;      // [T, B, H] dL/dh_t from the head
    const float* w_hh;     // [4H, H]
    const float* gates;    // [T, B, 4H] post-activation i,f,g,o
    const float* cs_ext;   // [(T+1), B, H], slot t = c_{t-1}, slot t+1 = c_t
    float* dgates;         // [T, B, 4H] out: pre-activation gradients
    __nv_bfloat16* dgates16;   // optional [T, B, 4H] bf16 copy (operand of the gradient GEMMs)
    __nv_bfloat16* dg16;   // [2, 4H/64, B, 64] ping-pong, k-block-major (contiguous TMA boxes)
    int* counter;
    int T, B, H;
    long long* trace;      // optional [T, 32] globaltimer stamps of CTA (0,0), or null
};
#define LB_STAMP(slot)                                                        \
    do {                                                                      \
        if (a.trace && blockIdx.x == 0 && blockIdx.y == 0) a.trace[s * 32 + (slot)] = gtimer(); \
    } while (0)
constexpr int LB_RING_BYTES = 128 * 1024;   // A-tile ring: with NB = 32 a whole step's dG (32 tiles of 4 KB)

// NB = batch rows per CTA (32 or 64); the UMMA is always M = 64, with NB = 32 the upper half of
// the A tile is never filled and its accumulator rows are ignored (rows are independent).
template <int NB>
__global__ void __launch_bounds__(288, 1)
tc_lstm_bwd_kernel(const LstmBwdArgs a) {
    // A tile = NB rows x 128 B.  The UMMA reads 64 rows: with NB = 32 the upper half of a tile
    // aliases the next tile (or the pad behind the ring) and only feeds ignored accumulator rows.
    // Ring = 4 slots of 32 KB; a slot is a LOAD GROUP of kpg consecutive k-block tiles, contiguous
    // and pre-swizzled in global memory: one bulk copy and one mbarrier per group (see the forward
    // kernel).  NB = 32: 4 groups = the whole dG of a step (H <= 512), no "slot empty" barriers.
    constexpr int LB_STAGES = 4;
    constexpr uint32_t SLOT = LB_RING_BYTES / LB_STAGES;
    constexpr bool WHOLE = NB == 32;
    const int H = a.H, B = a.B, K4 = 4 * a.H, NKB = K4 / 64;
    const int kpg = WHOLE ? NKB / 4 : (NKB >= 8 ? NKB / 8 : 1);   // tiles per load group (<= 32 KB)
    const int ngrp = NKB / kpg;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[LB_STAGES], empty_bar[LB_STAGES], acc_full, acc_empty;
    __shared__ uint32_t tmem_slot;
    __shared__ __align__(16) float R[64][20];                   // recurrent term, row stride 20: conflict-free float4
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const uint32_t W_KB = LS_UNITS * 128;                       // 16 rows x 64 bf16 = 2 KB
    uint8_t* w_s = base;                                        // [NKB][2 KB]
    constexpr uint32_t A_TILE = NB * 128;                       // NB rows x 64 bf16
    uint8_t* a_s = base + (size_t)NKB * W_KB;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int u0 = blockIdx.x * LS_UNITS, b0 = blockIdx.y * NB;
    const int ncta = gridDim.x;
    int* counter = a.counter + blockIdx.y;
    const long long BH = (long long)B * H;
    const long long GRP = (long long)NKB * NB * 64;             // dg16 image: [groups][NKB][NB][64]
    const long long IMG = GRP * gridDim.y;

    if (tid == 0) {
        for (int s = 0; s < LB_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_full, LS_ISSUERS);
        mbar_init(&acc_empty, 128);
        mbar_fence_init();
    }
    // resident B operand: row = unit u, K = gate index j: W_hh[j, u0+u]
    for (int i = tid; i < LS_UNITS * (K4 / 8); i += blockDim.x) {
        const int u = i % LS_UNITS, ch = i / LS_UNITS;          // consecutive threads -> consecutive u
        const int kb = ch >> 3, c = ch & 7;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __ldg(a.w_hh + (long long)(ch * 8 + e) * H + u0 + u);
        *reinterpret_cast<uint4*>(w_s + (uint32_t)kb * W_KB + ls_sw128(u, c)) =
            make_uint4(ls_pack(v[0], v[1]), ls_pack(v[2], v[3]), ls_pack(v[4], v[5]), ls_pack(v[6], v[7]));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 1) tmem_alloc(&tmem_slot, 64);                   // LS_ISSUERS slices of 16 columns
    const int issuer = warp == 1 ? 0 : (warp >= 6 ? warp - 5 : -1);   // MMA warps 1, 6, 7, 8
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    // warps 2..5: TMEM readers (lane < 16 of warp q = warp % 4 owns accumulator row 16q + lane)
    // hand the recurrent term over through shared memory, then all 128 threads do the gate-gradient
    // math: thread -> batch row mr and EPT consecutive hidden units (dL/dc stays in registers)
    const int q = warp & 3;
    const bool treader = warp >= 2 && warp < 6 && q * 16 < NB;   // warp-uniform: tcgen05.ld is warp-wide
    constexpr int EPT = NB / 8, TPR = LS_UNITS / EPT;           // 4 or 8 units per thread
    const int mi = tid - 64;
    const int mr = mi / TPR, mu = (mi % TPR) * EPT;
    const int rb = b0 + mr;
    const bool rvalid = warp >= 2 && warp < 6 && rb < B;
    float dc_reg[EPT];
#pragma unroll
    for (int j = 0; j < EPT; ++j) dc_reg[j] = 0.f;

    uint32_t it = 0;
    for (int s = 0; s < a.T; ++s) {
        const int t = a.T - 1 - s;
        if (warp == 0) {
            if (lane == 0 && s > 0) {
                const int target = ncta * s;                    // every CTA finished iteration s-1
                uint32_t spins = 0;
                while (ld_acquire(counter) < target)
                    if (++spins > (1u << 28)) __trap();
                LB_STAMP(0);                                    // barrier passed
                asm volatile("fence.proxy.async.global;" ::: "memory");
                const int p = (s - 1) & 1;
                const uint8_t* img = reinterpret_cast<const uint8_t*>(a.dg16 + (long long)p * IMG + (long long)blockIdx.y * GRP);
                for (int g = 0; g < ngrp; ++g, ++it) {
                    const int st = it % LB_STAGES;
                    if constexpr (!WHOLE) mbar_wait(&empty_bar[st], ((it / LB_STAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_bar[st], kpg * A_TILE);
                    bulk_g2s(a_s + (size_t)st * SLOT, img + (size_t)g * kpg * A_TILE, kpg * A_TILE, &full_bar[st]);
                }
                LB_STAMP(1);                                    // all loads issued
            }
        } else if (issuer >= 0) {
            if (lane == 0 && s > 0) {
                // four issuing warps (k-blocks kb % 4 == issuer), private 16-column TMEM slices
                constexpr uint32_t idesc = umma_idesc_bf16(LS_NB, LS_UNITS);
                mbar_wait(&acc_empty, s & 1);   // readers drained iteration s-1 (s=1: passes at once)
                tc_fence_after();
                const uint64_t db0 = umma_desc_k_sw128(smem_u32(w_s));
                const uint64_t da0 = umma_desc_k_sw128(smem_u32(a_s));
                const uint32_t d = tmem_base + issuer * LS_UNITS;
                const uint32_t it0 = (uint32_t)(s - 1) * ngrp;
                for (int g = issuer; g < ngrp; g += LS_ISSUERS) {
                    const uint32_t itk = it0 + g;
                    const int st = itk % LB_STAGES;
                    mbar_wait(&full_bar[st], (itk / LB_STAGES) & 1);
                    tc_fence_after();
                    if (g == 0) LB_STAMP(9);                    // first group landed
                    if (g == ngrp - 1) LB_STAMP(10);            // last group landed
                    for (int j = 0; j < kpg; ++j) {
                        const int kb = g * kpg + j;
                        const uint64_t aa = da0 + (uint64_t)((st * SLOT + j * A_TILE) >> 4);
                        const uint64_t bb = db0 + (uint64_t)((kb * W_KB) >> 4);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16(d, aa + 2 * k, bb + 2 * k, idesc, (g >= LS_ISSUERS) | (j != 0) | (k != 0));
                    }
                    if constexpr (!WHOLE) umma_commit(&empty_bar[st]);
                }
                umma_commit(&acc_full);
                if (issuer == (ngrp - 1) % LS_ISSUERS) LB_STAMP(2);   // last MMA issued
            }
        } else if (warp >= 2 && warp < 6) {
            // ---- everything that does not depend on the recurrent term is loaded BEFORE waiting
            //      for the accumulator
            float gi[EPT], gf[EPT], gg[EPT], go[EPT], ct[EPT], cp[EPT], dh[EPT];
            const long long o = (long long)rb * H + u0 + mu;
            if (rvalid) {
                const float* gp = a.gates + ((long long)t * B + rb) * K4 + u0 + mu;
#pragma unroll
                for (int j = 0; j < EPT; j += 4) {
                    *reinterpret_cast<float4*>(gi + j) = __ldg(reinterpret_cast<const float4*>(gp + j));
                    *reinterpret_cast<float4*>(gf + j) = __ldg(reinterpret_cast<const float4*>(gp + H + j));
                    *reinterpret_cast<float4*>(gg + j) = __ldg(reinterpret_cast<const float4*>(gp + 2 * H + j));
                    *reinterpret_cast<float4*>(go + j) = __ldg(reinterpret_cast<const float4*>(gp + 3 * H + j));
                    *reinterpret_cast<float4*>(ct + j) = __ldg(reinterpret_cast<const float4*>(a.cs_ext + (long long)(t + 1) * BH + o + j));
                    *reinterpret_cast<float4*>(cp + j) = __ldg(reinterpret_cast<const float4*>(a.cs_ext + (long long)t * BH + o + j));
                    *reinterpret_cast<float4*>(dh + j) = __ldg(reinterpret_cast<const float4*>(a.dhs + (long long)t * BH + o + j));
                }
            }
            float rec[EPT];
#pragma unroll
            for (int j = 0; j < EPT; ++j) rec[j] = 0.f;
            if (s > 0) {
                if (tid == 64) {
                    mbar_wait_sleep(&acc_full, (s - 1) & 1, 32);
                    LB_STAMP(3);                                // accumulators complete
                }
                named_bar_sync(3, 128);
                tc_fence_after();
                if (treader) {
                    uint32_t ra[32], rb2[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16), ra);
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + 32, rb2);
                    tmem_ld_wait();
                    float* dst = &R[q * 16 + (lane & 15)][0];
                    if (lane < 16)
#pragma unroll
                    for (int j = 0; j < LS_UNITS; j += 4)
                        *reinterpret_cast<float4*>(dst + j) = make_float4(
                            (__uint_as_float(ra[j]) + __uint_as_float(ra[16 + j])) + (__uint_as_float(rb2[j]) + __uint_as_float(rb2[16 + j])),
                            (__uint_as_float(ra[j + 1]) + __uint_as_float(ra[17 + j])) + (__uint_as_float(rb2[j + 1]) + __uint_as_float(rb2[17 + j])),
                            (__uint_as_float(ra[j + 2]) + __uint_as_float(ra[18 + j])) + (__uint_as_float(rb2[j + 2]) + __uint_as_float(rb2[18 + j])),
                            (__uint_as_float(ra[j + 3]) + __uint_as_float(ra[19 + j])) + (__uint_as_float(rb2[j + 3]) + __uint_as_float(rb2[19 + j])));
                }
                tc_fence_before();
                mbar_arrive(&acc_empty);
                named_bar_sync(4, 128);                          // R complete
                if (tid == 64) LB_STAMP(4);                     // TMEM read + hand-over
#pragma unroll
                for (int j = 0; j < EPT; j += 4)
                    *reinterpret_cast<float4*>(rec + j) = *reinterpret_cast<const float4*>(&R[mr][mu + j]);
            }
            float di[EPT], df[EPT], dg[EPT], dov[EPT];
            if (rvalid) {
#pragma unroll
                for (int j = 0; j < EPT; ++j) {
                    const float d = dh[j] + rec[j];
                    const float tc = ls_tanh(ct[j]);
                    const float dc = dc_reg[j] + d * go[j] * (1.f - tc * tc);
                    di[j] = dc * gg[j] * gi[j] * (1.f - gi[j]);
                    df[j] = dc * cp[j] * gf[j] * (1.f - gf[j]);
                    dg[j] = dc * gi[j] * (1.f - gg[j] * gg[j]);
                    dov[j] = d * tc * go[j] * (1.f - go[j]);
                    dc_reg[j] = dc * gf[j];
                }
                // bf16 image the next step's bulk copies read: gate g, units u0.. = k-block
                // g*H/64 + u0/64; 16-byte chunks in SWIZZLE_128B order (chunk ^ row%8)
                const int col = (u0 & 63) + mu;
                __nv_bfloat16* d16 = a.dg16 + (long long)(s & 1) * IMG + (long long)blockIdx.y * GRP +
                                     ((long long)(u0 >> 6) * NB + mr) * 64 + ((((col >> 3) ^ (mr & 7)) << 3) | (col & 7));
                const long long gstep = (long long)(H >> 6) * NB * 64;
                if constexpr (EPT == 4) {
                    *reinterpret_cast<uint2*>(d16) = make_uint2(ls_pack(di[0], di[1]), ls_pack(di[2], di[3]));
                    *reinterpret_cast<uint2*>(d16 + gstep) = make_uint2(ls_pack(df[0], df[1]), ls_pack(df[2], df[3]));
                    *reinterpret_cast<uint2*>(d16 + 2 * gstep) = make_uint2(ls_pack(dg[0], dg[1]), ls_pack(dg[2], dg[3]));
                    *reinterpret_cast<uint2*>(d16 + 3 * gstep) = make_uint2(ls_pack(dov[0], dov[1]), ls_pack(dov[2], dov[3]));
                } else {
                    *reinterpret_cast<uint4*>(d16) = make_uint4(ls_pack(di[0], di[1]), ls_pack(di[2], di[3]), ls_pack(di[4], di[5]), ls_pack(di[6], di[7]));
                    *reinterpret_cast<uint4*>(d16 + gstep) = make_uint4(ls_pack(df[0], df[1]), ls_pack(df[2], df[3]), ls_pack(df[4], df[5]), ls_pack(df[6], df[7]));
                    *reinterpret_cast<uint4*>(d16 + 2 * gstep) = make_uint4(ls_pack(dg[0], dg[1]), ls_pack(dg[2], dg[3]), ls_pack(dg[4], dg[5]), ls_pack(dg[6], dg[7]));
                    *reinterpret_cast<uint4*>(d16 + 3 * gstep) = make_uint4(ls_pack(dov[0], dov[1]), ls_pack(dov[2], dov[3]), ls_pack(dov[4], dov[5]), ls_pack(dov[6], dov[7]));
                }
            }
            if (tid == 64) LB_STAMP(5);                         // gate math + bf16 stores issued
            named_bar_sync(1, 128);
            if (tid == 64) {
                LB_STAMP(6);
                red_release_add(counter, 1);
                LB_STAMP(7);
            }
            // fp32 gradients for the kernels after the unroll, behind the releasing arrive
            if (rvalid) {
                float* dgp = a.dgates + ((long long)t * B + rb) * K4 + u0 + mu;
#pragma unroll
                for (int j = 0; j < EPT; j += 4) {
                    *reinterpret_cast<float4*>(dgp + j) = make_float4(di[j], di[j + 1], di[j + 2], di[j + 3]);
                    *reinterpret_cast<float4*>(dgp + H + j) = make_float4(df[j], df[j + 1], df[j + 2], df[j + 3]);
                    *reinterpret_cast<float4*>(dgp + 2 * H + j) = make_float4(dg[j], dg[j + 1], dg[j + 2], dg[j + 3]);
                    *reinterpret_cast<float4*>(dgp + 3 * H + j) = make_float4(dov[j], dov[j + 1], dov[j + 2], dov[j + 3]);
                }
                if (a.dgates16) {
                    __nv_bfloat16* d16r = a.dgates16 + ((long long)t * B + rb) * K4 + u0 + mu;
#pragma unroll
                    for (int j = 0; j < EPT; j += 4) {
                        *reinterpret_cast<uint2*>(d16r + j) = make_uint2(ls_pack(di[j], di[j + 1]), ls_pack(di[j + 2], di[j + 3]));
                        *reinterpret_cast<uint2*>(d16r + H + j) = make_uint2(ls_pack(df[j], df[j + 1]), ls_pack(df[j + 2], df[j + 3]));
                        *reinterpret_cast<uint2*>(d16r + 2 * H + j) = make_uint2(ls_pack(dg[j], dg[j + 1]), ls_pack(dg[j + 2], dg[j + 3]));
                        *reinterpret_cast<uint2*>(d16r + 3 * H + j) = make_uint2(ls_pack(dov[j], dov[j + 1]), ls_pack(dov[j + 2], dov[j + 3]));
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 64);
}

// BPTT on CTA PAIRS: the contraction over the 4H gate rows is split between the two CTAs of a
// cluster (same 16 units, same 32 batch rows, half of the k-blocks each): 64 instead of 128 small
// MMAs and 64 instead of 128 KB of dG per CTA and step.  Each CTA finishes 8 of the 16 units: it keeps its
// partial sums of those and PUSHES the partial sums of the other 8 into the peer's shared memory
// (st.async, whose bytes complete a transaction barrier over there); the two partials are
// summed in the order half 0 + half 1 on both sides.
__device__ __forceinline__ uint32_t map_peer(const void* p, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
    return r;
}
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(288, 1)
tc_lstm_bwd_pair_kernel(const LstmBwdArgs a) {
    constexpr int NB = 32;
    // A tile = NB rows x 128 B.  The UMMA reads 64 rows: with NB = 32 the upper half of a tile
    // aliases the next tile (or the pad behind the ring) and only feeds ignored accumulator rows.
    // Ring = 4 slots of 32 KB; a slot is a LOAD GROUP of kpg consecutive k-block tiles, contiguous
    // and pre-swizzled in global memory: one bulk copy and one mbarrier per group (see the forward
    // kernel).  NB = 32: 4 groups = the whole dG of a step (H <= 512), no "slot empty" barriers.
    constexpr int LB_STAGES = 4;
    constexpr uint32_t SLOT = LB_RING_BYTES / LB_STAGES;
    constexpr bool WHOLE = NB == 32;
    const int H = a.H, B = a.B, K4 = 4 * a.H, NKB_ALL = K4 / 64, NKB = NKB_ALL / 2;   // NKB: this CTA's k-blocks
    const uint32_t half = cluster_ctarank();
    const int kb0 = (int)half * NKB;
    const int kpg = WHOLE ? NKB / 4 : (NKB >= 8 ? NKB / 8 : 1);   // tiles per load group (<= 32 KB)
    const int ngrp = NKB / kpg;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[LB_STAGES], empty_bar[LB_STAGES], acc_full, acc_empty, xbar;
    __shared__ uint32_t tmem_slot;
    // partial recurrent terms of step parity s & 1: R = this CTA's k-half, P = the peer's (pushed by the peer),
    // both only for the 8 units this CTA finishes
    __shared__ __align__(16) float R[2][64][20], P[2][64][20];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const uint32_t W_KB = LS_UNITS * 128;                       // 16 rows x 64 bf16 = 2 KB
    uint8_t* w_s = base;                                        // [NKB][2 KB]
    constexpr uint32_t A_TILE = NB * 128;                       // NB rows x 64 bf16
    uint8_t* a_s = base + (size_t)NKB * W_KB;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int u0 = (blockIdx.x >> 1) * LS_UNITS, b0 = blockIdx.y * NB;
    const int ncta = gridDim.x;
    int* counter = a.counter + blockIdx.y;
    const long long BH = (long long)B * H;
    const long long GRP = (long long)NKB_ALL * NB * 64;         // dg16 image: [groups][4H/64][NB][64]
    const long long IMG = GRP * gridDim.y;

    if (tid == 0) {
        for (int s = 0; s < LB_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_full, LS_ISSUERS);
        mbar_init(&acc_empty, 128);
        mbar_init(&xbar, 1);
        mbar_fence_init();
    }
    // resident B operand: row = unit u, K = gate index j: W_hh[j, u0+u]
    for (int i = tid; i < LS_UNITS * NKB * 8; i += blockDim.x) {
        const int u = i % LS_UNITS, chl = i / LS_UNITS;         // consecutive threads -> consecutive u
        const int kb = chl >> 3, c = chl & 7, ch = kb0 * 8 + chl;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __ldg(a.w_hh + (long long)(ch * 8 + e) * H + u0 + u);
        *reinterpret_cast<uint4*>(w_s + (uint32_t)kb * W_KB + ls_sw128(u, c)) =
            make_uint4(ls_pack(v[0], v[1]), ls_pack(v[2], v[3]), ls_pack(v[4], v[5]), ls_pack(v[6], v[7]));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 1) tmem_alloc(&tmem_slot, 64);                   // LS_ISSUERS slices of 16 columns
    const int issuer = warp == 1 ? 0 : (warp >= 6 ? warp - 5 : -1);   // MMA warps 1, 6, 7, 8
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                          // the peer's xbar exists
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    // warps 2..5: TMEM readers (lane < 16 of warp q = warp % 4 owns accumulator row 16q + lane)
    // hand the recurrent term over through shared memory, then all 128 threads do the gate-gradient
    // math: thread -> batch row mr and EPT consecutive hidden units (dL/dc stays in registers)
    const int q = warp & 3;
    const bool treader = warp >= 2 && warp < 6 && q * 16 < NB;   // warp-uniform: tcgen05.ld is warp-wide
    constexpr int EPT = 4, TPR = (LS_UNITS / 2) / EPT;          // this CTA finishes 8 of the 16 units: 64 threads
    const int mi = tid - 64;
    const int mr = (mi / TPR) & 31, mu = (int)half * (LS_UNITS / 2) + (mi % TPR) * EPT;
    const int rb = b0 + mr;
    const bool rvalid = warp >= 2 && warp < 6 && mi < NB * TPR && rb < B;
    float dc_reg[EPT];
#pragma unroll
    for (int j = 0; j < EPT; ++j) dc_reg[j] = 0.f;

    uint32_t it = 0;
    for (int s = 0; s < a.T; ++s) {
        const int t = a.T - 1 - s;
        if (warp == 0) {
            if (lane == 0 && s > 0) {
                const int target = ncta * s;                    // every CTA finished iteration s-1
                uint32_t spins = 0;
                while (ld_acquire(counter) < target)
                    if (++spins > (1u << 28)) __trap();
                LB_STAMP(0);                                    // barrier passed
                asm volatile("fence.proxy.async.global;" ::: "memory");
                const int p = (s - 1) & 1;
                const uint8_t* img = reinterpret_cast<const uint8_t*>(a.dg16 + (long long)p * IMG + (long long)blockIdx.y * GRP);
                for (int g = 0; g < ngrp; ++g, ++it) {
                    const int st = it % LB_STAGES;
                    if constexpr (!WHOLE) mbar_wait(&empty_bar[st], ((it / LB_STAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_bar[st], kpg * A_TILE);
                    bulk_g2s(a_s + (size_t)st * SLOT, img + (size_t)(kb0 + g * kpg) * A_TILE, kpg * A_TILE, &full_bar[st]);
                }
                LB_STAMP(1);                                    // all loads issued
            }
        } else if (issuer >= 0) {
            if (lane == 0 && s > 0) {
                // four issuing warps (k-blocks kb % 4 == issuer), private 16-column TMEM slices
                constexpr uint32_t idesc = umma_idesc_bf16(LS_NB, LS_UNITS);
                mbar_wait(&acc_empty, s & 1);   // readers drained iteration s-1 (s=1: passes at once)
                tc_fence_after();
                const uint64_t db0 = umma_desc_k_sw128(smem_u32(w_s));
                const uint64_t da0 = umma_desc_k_sw128(smem_u32(a_s));
                const uint32_t d = tmem_base + issuer * LS_UNITS;
                const uint32_t it0 = (uint32_t)(s - 1) * ngrp;
                for (int g = issuer; g < ngrp; g += LS_ISSUERS) {
                    const uint32_t itk = it0 + g;
                    const int st = itk % LB_STAGES;
                    mbar_wait(&full_bar[st], (itk / LB_STAGES) & 1);
                    tc_fence_after();
                    if (g == 0) LB_STAMP(9);                    // first group landed
                    if (g == ngrp - 1) LB_STAMP(10);            // last group landed
                    for (int j = 0; j < kpg; ++j) {
                        const int kb = g * kpg + j;
                        const uint64_t aa = da0 + (uint64_t)((st * SLOT + j * A_TILE) >> 4);
                        const uint64_t bb = db0 + (uint64_t)((kb * W_KB) >> 4);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16(d, aa + 2 * k, bb + 2 * k, idesc, (g >= LS_ISSUERS) | (j != 0) | (k != 0));
                    }
                    if constexpr (!WHOLE) umma_commit(&empty_bar[st]);
                }
                umma_commit(&acc_full);
                if (issuer == (ngrp - 1) % LS_ISSUERS) LB_STAMP(2);   // last MMA issued
            }
        } else if (warp >= 2 && warp < 6) {
            // ---- everything that does not depend on the recurrent term is loaded BEFORE waiting
            //      for the accumulator
            float gi[EPT], gf[EPT], gg[EPT], go[EPT], ct[EPT], cp[EPT], dh[EPT];
            const long long o = (long long)rb * H + u0 + mu;
            if (rvalid) {
                const float* gp = a.gates + ((long long)t * B + rb) * K4 + u0 + mu;
#pragma unroll
                for (int j = 0; j < EPT; j += 4) {
                    *reinterpret_cast<float4*>(gi + j) = __ldg(reinterpret_cast<const float4*>(gp + j));
                    *reinterpret_cast<float4*>(gf + j) = __ldg(reinterpret_cast<const float4*>(gp + H + j));
                    *reinterpret_cast<float4*>(gg + j) = __ldg(reinterpret_cast<const float4*>(gp + 2 * H + j));
                    *reinterpret_cast<float4*>(go + j) = __ldg(reinterpret_cast<const float4*>(gp + 3 * H + j));
                    *reinterpret_cast<float4*>(ct + j) = __ldg(reinterpret_cast<const float4*>(a.cs_ext + (long long)(t + 1) * BH + o + j));
                    *reinterpret_cast<float4*>(cp + j) = __ldg(reinterpret_cast<const float4*>(a.cs_ext + (long long)t * BH + o + j));
                    *reinterpret_cast<float4*>(dh + j) = __ldg(reinterpret_cast<const float4*>(a.dhs + (long long)t * BH + o + j));
                }
            }
            float rec[EPT];
#pragma unroll
            for (int j = 0; j < EPT; ++j) rec[j] = 0.f;
            if (s > 0) {
                if (tid == 64) {
                    mbar_expect_tx(&xbar, NB * (LS_UNITS / 2) * 4);  // the peer's share of this step: 32 rows x 8 units
                    mbar_wait_sleep(&acc_full, (s - 1) & 1, 32);
                    LB_STAMP(3);                                // accumulators complete
                }
                named_bar_sync(3, 128);
                tc_fence_after();
                if (treader) {
                    uint32_t ra[32], rb2[32];
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16), ra);
                    tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + 32, rb2);
                    tmem_ld_wait();
                    // units 0-7 are finished by half 0, units 8-15 by half 1: the sums of MY units
                    // stay here, the others are pushed into the peer's shared memory
                    float* dst = &R[s & 1][q * 16 + (lane & 15)][0];
                    const uint32_t rdst = map_peer(&P[s & 1][q * 16 + (lane & 15)][0], half ^ 1u);
                    const uint32_t rbar = map_peer(&xbar, half ^ 1u);
                    if (lane < 16)
#pragma unroll
                    for (int j = 0; j < LS_UNITS; j += 4) {
                        const float4 v = make_float4(
                            (__uint_as_float(ra[j]) + __uint_as_float(ra[16 + j])) + (__uint_as_float(rb2[j]) + __uint_as_float(rb2[16 + j])),
                            (__uint_as_float(ra[j + 1]) + __uint_as_float(ra[17 + j])) + (__uint_as_float(rb2[j + 1]) + __uint_as_float(rb2[17 + j])),
                            (__uint_as_float(ra[j + 2]) + __uint_as_float(ra[18 + j])) + (__uint_as_float(rb2[j + 2]) + __uint_as_float(rb2[18 + j])),
                            (__uint_as_float(ra[j + 3]) + __uint_as_float(ra[19 + j])) + (__uint_as_float(rb2[j + 3]) + __uint_as_float(rb2[19 + j])));
                        if ((uint32_t)(j >> 3) == half)
                            *reinterpret_cast<float4*>(dst + j) = v;
                        else
                            // asynchronous remote store that counts its bytes on the peer's barrier
                            asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(
                                             rdst + j * 4),
                                         "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(rbar)
                                         : "memory");
                    }
                }
                tc_fence_before();
                mbar_arrive(&acc_empty);
                named_bar_sync(4, 128);                          // both partials are written (mine here, the peer's share over there)
                {                                                // the peer's share has landed in P
                    uint32_t ok = 0, spins = 0;
                    while (!ok) {
                        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
                                     "selp.u32 %0, 1, 0, p;\n}"
                                     : "=r"(ok)
                                     : "r"(smem_u32(&xbar)), "r"((uint32_t)((s - 1) & 1))
                                     : "memory");
                        if (!ok && ++spins > (1u << 26)) __trap();
                    }
                }
                if (tid == 64) LB_STAMP(4);                     // TMEM read + hand-over
                {
                    const float4 own = *reinterpret_cast<const float4*>(&R[s & 1][mr][mu]);
                    const float4 peer = *reinterpret_cast<const float4*>(&P[s & 1][mr][mu]);
                    const float4 lo = half == 0 ? own : peer, hi = half == 0 ? peer : own;   // half 0 + half 1
                    rec[0] = lo.x + hi.x, rec[1] = lo.y + hi.y, rec[2] = lo.z + hi.z, rec[3] = lo.w + hi.w;
                }
            }
            float di[EPT], df[EPT], dg[EPT], dov[EPT];
            if (rvalid) {
#pragma unroll
                for (int j = 0; j < EPT; ++j) {
                    const float d = dh[j] + rec[j];
                    const float tc = ls_tanh(ct[j]);
                    const float dc = dc_reg[j] + d * go[j] * (1.f - tc * tc);
                    di[j] = dc * gg[j] * gi[j] * (1.f - gi[j]);
                    df[j] = dc * cp[j] * gf[j] * (1.f - gf[j]);
                    dg[j] = dc * gi[j] * (1.f - gg[j] * gg[j]);
                    dov[j] = d * tc * go[j] * (1.f - go[j]);
                    dc_reg[j] = dc * gf[j];
                }
                // bf16 image the next step's bulk copies read: gate g, units u0.. = k-block
                // g*H/64 + u0/64; 16-byte chunks in SWIZZLE_128B order (chunk ^ row%8)
                const int col = (u0 & 63) + mu;
                __nv_bfloat16* d16 = a.dg16 + (long long)(s & 1) * IMG + (long long)blockIdx.y * GRP +
                                     ((long long)(u0 >> 6) * NB + mr) * 64 + ((((col >> 3) ^ (mr & 7)) << 3) | (col & 7));
                const long long gstep = (long long)(H >> 6) * NB * 64;
                if constexpr (EPT == 4) {
                    *reinterpret_cast<uint2*>(d16) = make_uint2(ls_pack(di[0], di[1]), ls_pack(di[2], di[3]));
                    *reinterpret_cast<uint2*>(d16 + gstep) = make_uint2(ls_pack(df[0], df[1]), ls_pack(df[2], df[3]));
                    *reinterpret_cast<uint2*>(d16 + 2 * gstep) = make_uint2(ls_pack(dg[0], dg[1]), ls_pack(dg[2], dg[3]));
                    *reinterpret_cast<uint2*>(d16 + 3 * gstep) = make_uint2(ls_pack(dov[0], dov[1]), ls_pack(dov[2], dov[3]));
                } else {
                    *reinterpret_cast<uint4*>(d16) = make_uint4(ls_pack(di[0], di[1]), ls_pack(di[2], di[3]), ls_pack(di[4], di[5]), ls_pack(di[6], di[7]));
                    *reinterpret_cast<uint4*>(d16 + gstep) = make_uint4(ls_pack(df[0], df[1]), ls_pack(df[2], df[3]), ls_pack(df[4], df[5]), ls_pack(df[6], df[7]));
                    *reinterpret_cast<uint4*>(d16 + 2 * gstep) = make_uint4(ls_pack(dg[0], dg[1]), ls_pack(dg[2], dg[3]), ls_pack(dg[4], dg[5]), ls_pack(dg[6], dg[7]));
                    *reinterpret_cast<uint4*>(d16 + 3 * gstep) = make_uint4(ls_pack(dov[0], dov[1]), ls_pack(dov[2], dov[3]), ls_pack(dov[4], dov[5]), ls_pack(dov[6], dov[7]));
                }
            }
            if (tid == 64) LB_STAMP(5);                         // gate math + bf16 stores issued
            named_bar_sync(1, 128);
            if (tid == 64) {
                LB_STAMP(6);
                red_release_add(counter, 1);
                LB_STAMP(7);
            }
            // fp32 gradients for the kernels after the unroll, behind the releasing arrive
            if (rvalid) {
                float* dgp = a.dgates + ((long long)t * B + rb) * K4 + u0 + mu;
#pragma unroll
                for (int j = 0; j < EPT; j += 4) {
                    *reinterpret_cast<float4*>(dgp + j) = make_float4(di[j], di[j + 1], di[j + 2], di[j + 3]);
                    *reinterpret_cast<float4*>(dgp + H + j) = make_float4(df[j], df[j + 1], df[j + 2], df[j + 3]);
                    *reinterpret_cast<float4*>(dgp + 2 * H + j) = make_float4(dg[j], dg[j + 1], dg[j + 2], dg[j + 3]);
                    *reinterpret_cast<float4*>(dgp + 3 * H + j) = make_float4(dov[j], dov[j + 1], dov[j + 2], dov[j + 3]);
                }
                if (a.dgates16) {
                    __nv_bfloat16* d16r = a.dgates16 + ((long long)t * B + rb) * K4 + u0 + mu;
#pragma unroll
                    for (int j = 0; j < EPT; j += 4) {
                        *reinterpret_cast<uint2*>(d16r + j) = make_uint2(ls_pack(di[j], di[j + 1]), ls_pack(di[j + 2], di[j + 3]));
                        *reinterpret_cast<uint2*>(d16r + H + j) = make_uint2(ls_pack(df[j], df[j + 1]), ls_pack(df[j + 2], df[j + 3]));
                        *reinterpret_cast<uint2*>(d16r + 2 * H + j) = make_uint2(ls_pack(dg[j], dg[j + 1]), ls_pack(dg[j + 2], dg[j + 3]));
                        *reinterpret_cast<uint2*>(d16r + 3 * H + j) = make_uint2(ls_pack(dov[j], dov[j + 1]), ls_pack(dov[j + 2], dov[j + 3]));
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                          // the peer may still be writing P
    if (warp == 1) tmem_dealloc(tmem_base, 64);
}

}  // namespace r2d2

using namespace r2d2;

static long long* g_lstm_trace = nullptr;

extern "C" {

// diagnostics: device buffer [T,32] of globaltimer stamps written by CTA (0,0) of the next forward
// unrolls (null switches it off)
int r2d2_tc_lstm_set_trace(void* trace) {
    g_lstm_trace = static_cast<long long*>(trace);
    return R2D2_OK;
}

int64_t r2d2_tc_lstm_scratch_bytes(int B, int H) {
    // h ping-pong [2, H/64, groups, hi|lo, NB, 64] bf16 (batch padded to whole 64-row groups),
    // + barrier counters (256 B, kept 256-byte aligned)
    const int64_t Bp = ((int64_t)B + 63) / 64 * 64;
    return 2 * 2 * Bp * H * 2 + 256;
}

int r2d2_tc_lstm_seq_fwd(const float* xprojT, int64_t ldx, int64_t col0, const float* w_hh,
                         float* hs_ext, float* cs_ext, float* gates, void* h16_hi, void* h16_lo,
                         int hist_t0, int T, int B, int H, void* scratch, void* stream) {
    R2D2_REQUIRE((h16_hi == nullptr) == (h16_lo == nullptr) && hist_t0 >= 0 && hist_t0 <= T &&
                     ((uintptr_t)h16_hi % 8) == 0 && ((uintptr_t)h16_lo % 8) == 0,
                 "tc_lstm_seq_fwd: h16_hi/h16_lo go together (8-byte aligned), 0 <= hist_t0 <= T");
    R2D2_REQUIRE(xprojT && w_hh && hs_ext && cs_ext && scratch, "tc_lstm_seq_fwd: null pointer");
    R2D2_REQUIRE(T > 0 && B > 0 && H > 0 && H % 64 == 0 && H <= 512,
                 "tc_lstm_seq_fwd: H must be a multiple of 64 and <= 512 (H=%d)", H);
    R2D2_REQUIRE(((uintptr_t)w_hh % 16) == 0 && ((uintptr_t)hs_ext % 16) == 0 &&
                     ((uintptr_t)cs_ext % 16) == 0 && ((uintptr_t)scratch % 256) == 0 &&
                     (!gates || ((uintptr_t)gates % 16) == 0),
                 "tc_lstm_seq_fwd: misaligned buffer");
    cudaStream_t st = as_stream(stream);
    const int64_t h16 = r2d2_tc_lstm_scratch_bytes(B, H) - 256;
    uint8_t* sc = static_cast<uint8_t*>(scratch);
    LstmArgs a;
    a.xprojT = xprojT; a.ldx = ldx; a.col0 = col0; a.w_hh = w_hh;
    a.hs_ext = hs_ext; a.cs_ext = cs_ext; a.gates = gates;
    a.h16 = reinterpret_cast<__nv_bfloat16*>(sc);
    a.counter = reinterpret_cast<int*>(sc + h16);
    a.T = T; a.B = B; a.H = H;
    a.trace = g_lstm_trace;
    a.hist_hi = static_cast<__nv_bfloat16*>(h16_hi);
    a.hist_lo = static_cast<__nv_bfloat16*>(h16_lo);
    a.hist_t0 = hist_t0;
    // batch tile: 32 columns spreads a chain over twice the SMs; fall back to 64 when that many
    // CTAs (two chains may run side by side) would not be co-resident
    const int NKB = H / 64;
    int nb = 32;
    if ((H / LS_UNITS) * ((B + 31) / 32) * 2 > sm_count()) nb = 64;
    const int groups = (B + nb - 1) / nb;
    R2D2_REQUIRE(groups <= 64, "tc_lstm_seq_fwd: batch too large");
    R2D2_CUDA_OK(cudaMemsetAsync(a.counter, 0, 256, st));
    const int smem = 2 * NKB * LS_ROWS * 128 + (256 / nb) * 2 * nb * 128 + 1024;
    void* kern = nb == 32 ? (void*)tc_lstm_fwd_kernel<32> : (void*)tc_lstm_fwd_kernel<64>;
    R2D2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    dim3 grid(H / LS_UNITS, groups);
    R2D2_REQUIRE((int)(grid.x * grid.y) <= sm_count(),
                 "tc_lstm_seq_fwd: %d CTAs must be co-resident but the device has %d SMs",
                 grid.x * grid.y, sm_count());
    void* params[] = {&a};
    {
        ProfScope ps("tc_lstm_fwd_kernel", st);
        R2D2_CUDA_OK(cudaLaunchCooperativeKernel(kern, grid, dim3(416), params, (size_t)smem, st));
    }
    return R2D2_OK;
}

}  // extern "C"

extern "C" {

int64_t r2d2_tc_lstm_bwd_scratch_bytes(int B, int H) {
    // dG ping-pong [2, groups, 4H/64, NB, 64] bf16 (batch padded to whole 64-row groups) + counters
    const int64_t Bp = ((int64_t)B + 63) / 64 * 64;
    return 2 * Bp * 4 * H * 2 + 256;
}

int r2d2_tc_lstm_seq_bwd(const float* dhs, const float* w_hh, const float* gates,
                         const float* cs_ext, float* dgates, void* dgates_bf16, int T, int B, int H,
                         void* scratch, void* stream) {
    R2D2_REQUIRE(dhs && w_hh && gates && cs_ext && dgates && scratch, "tc_lstm_seq_bwd: null pointer");
    R2D2_REQUIRE(T > 0 && B > 0 && H > 0 && H % 64 == 0 && H <= 512,
                 "tc_lstm_seq_bwd: H must be a multiple of 64 and <= 512 (H=%d)", H);
    R2D2_REQUIRE(((uintptr_t)dhs % 16) == 0 && ((uintptr_t)gates % 16) == 0 && ((uintptr_t)cs_ext % 16) == 0 &&
                     ((uintptr_t)dgates % 16) == 0 && ((uintptr_t)scratch % 256) == 0,
                 "tc_lstm_seq_bwd: misaligned buffer");
    cudaStream_t st = as_stream(stream);
    const int64_t dg16 = r2d2_tc_lstm_bwd_scratch_bytes(B, H) - 256;
    uint8_t* sc = static_cast<uint8_t*>(scratch);
    LstmBwdArgs a;
    a.dhs = dhs; a.w_hh = w_hh; a.gates = gates; a.cs_ext = cs_ext; a.dgates = dgates;
    a.dgates16 = static_cast<__nv_bfloat16*>(dgates_bf16);
    R2D2_REQUIRE(((uintptr_t)dgates_bf16 % 8) == 0, "tc_lstm_seq_bwd: misaligned dgates_bf16");
    a.dg16 = reinterpret_cast<__nv_bfloat16*>(sc);
    a.counter = reinterpret_cast<int*>(sc + dg16);
    a.T = T; a.B = B; a.H = H;
    a.trace = g_lstm_trace;
    int nb = 32;
    if ((H / LS_UNITS) * ((B + 31) / 32) > sm_count()) nb = 64;
    const int groups = (B + nb - 1) / nb;
    R2D2_REQUIRE(groups <= 64, "tc_lstm_seq_bwd: batch too large");
    R2D2_CUDA_OK(cudaMemsetAsync(a.counter, 0, 256, st));
    const int NKB = 4 * H / 64;
    static const bool no_pairs = [] {                    // R2D2_BPTT_PAIRS=0: one CTA per unit block (A/B runs)
        const char* e = getenv("R2D2_BPTT_PAIRS");
        return e && e[0] == '0';
    }();
    if (!no_pairs && nb == 32 && NKB % 8 == 0 && 2 * (H / LS_UNITS) * groups <= sm_count()) {
        const int smem2 = (NKB / 2) * LS_UNITS * 128 + LB_RING_BYTES + 4096 + 1024;
        R2D2_CUDA_OK(cudaFuncSetAttribute(tc_lstm_bwd_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * (H / LS_UNITS), groups);
        cfg.blockDim = dim3(288);
        cfg.dynamicSmemBytes = (size_t)smem2;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeCooperative;     // all CTAs co-resident: they meet at a grid barrier every step
        attr[0].val.cooperative = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        {
            ProfScope ps("tc_lstm_bwd_kernel", st);
            R2D2_CUDA_OK(cudaLaunchKernelEx(&cfg, tc_lstm_bwd_pair_kernel, a));
        }
        return R2D2_OK;
    }
    const int smem = NKB * LS_UNITS * 128 + LB_RING_BYTES + 4096 + 1024;
    void* kern = nb == 32 ? (void*)tc_lstm_bwd_kernel<32> : (void*)tc_lstm_bwd_kernel<64>;
    R2D2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    dim3 grid(H / LS_UNITS, groups);
    R2D2_REQUIRE((int)(grid.x * grid.y) <= sm_count(), "tc_lstm_seq_bwd: %d CTAs must be co-resident", grid.x * grid.y);
    void* params[] = {&a};
    {
        ProfScope ps("tc_lstm_bwd_kernel", st);
        R2D2_CUDA_OK(cudaLaunchCooperativeKernel(kern, grid, dim3(288), params, (size_t)smem, st));
    }
    return R2D2_OK;
}

}  // extern "C"
