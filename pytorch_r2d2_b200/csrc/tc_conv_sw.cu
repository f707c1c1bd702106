// conv2 / conv3 forward (model.py:16-21,46) on tcgen05 tensor cores, "shifted window" form.
//
// The gather form in tc_conv.cu builds every A-tile row (a pixel's receptive field) from global
// memory, so each activation is loaded and stored once per kernel offset that touches it (4x in
// conv2, 9x in conv3) and the L1/shared data pipe, not the tensor pipe, bounds the kernel.  Here a
// frame's activations go to shared memory ONCE, as an image whose row p is one input position
// and whose 128-byte row holds that position's channels (bf16, hi and lo planes):
//   conv2 (4x4, stride 2, 20x20x32): space-to-depth by 2 -> 10x10 positions x 128 channels
//          (dy,dx,ci); the convolution becomes 2x2, stride 1: offsets (ky2,kx2) = row shifts
//          {0, 1, 10, 11}; two planes of 64 channels (dy = 0 | dy = 1);
//   conv3 (3x3, stride 1, 9x9x32): row p = [channels of position p | channels of position p+1];
//          offsets (ky, kx in {0,1}) = row shift 9*ky over the whole row, (ky, 2) = shift 9*ky+2
//          over its first half.
// The A operand of every kernel offset is then THE SAME shared-memory image read through a UMMA
// descriptor whose start address is advanced by `shift` rows (no base-offset field needed: the
// swizzle is a function of the absolute shared-memory address).  One
// M = 128 tile per frame (rows past the last position, and positions whose window leaves the
// frame, produce values nobody stores).
// Arithmetic as in tc_conv.cu: error-compensated bf16x3, A_hi x [W_hi ; W_lo] as one N = 64 MMA
// plus A_lo x W_hi, fp32 accumulation in TMEM (two buffers), bias + ReLU epilogue to NCHW fp32
// (and for conv3 the bf16 (hi, lo) image of the features for the projection GEMM).
#include <cuda_bf16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace r2d2 {

namespace {

// diagnostics: build with -DSWC_TRACE to record globaltimer stamps of CTA 0's first 32 frames
// (slot: 0 stage free seen by producer thread 0, 1 producer arrived, 2 issuer 0 starts, 3 issuer 0
// done issuing, 4 issuer 1 done issuing, 5 epilogue sees the accumulator, 6 epilogue stores done)
#ifdef SWC_TRACE
__device__ long long g_swc_trace[2 * 32 * 8];     // [layer - 2][frame][slot]
#define SWC_STAMP(slot)                                                                      \
    do {                                                                                     \
        if (blockIdx.x == 0 && it < 32 && (threadIdx.x & 31) == 0 &&                          \
            (threadIdx.x == 0 || threadIdx.x >= SWC_PROD)) {                                 \
            long long t__;                                                                   \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t__));                          \
            g_swc_trace[(L - 2) * 256 + it * 8 + (slot)] = t__;                                              \
        }                                                                                    \
    } while (0)
#else
#define SWC_STAMP(slot) do { } while (0)
#endif

constexpr int SWC_PROD = 416;            // producer threads (13 warps)
constexpr int SWC_ISSUERS = 2;           // MMA-issuing warps: warp 0 the A_hi MMAs, warp 1 the A_lo MMAs
constexpr int SWC_THREADS = SWC_PROD + 32 * SWC_ISSUERS + 128;

__device__ __forceinline__ uint32_t swc_pack(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float swc_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ uint32_t swc_off(int row, int chunk) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}
// K-major SWIZZLE_128B descriptor whose start may sit on any 128-byte row of a 1024-byte atom
__device__ __forceinline__ uint64_t swc_desc(uint32_t smem_addr) {
    return umma_desc_k_sw128(smem_addr);
}

template <int L> struct SwCfg;
template <> struct SwCfg<2> {            // conv2
    static constexpr int HIN = 20, GRID = 10, HOUT = 9, NPOS = 100, PLANES = 2, NKB = 8;
    static constexpr int ROWS = 104, STAGES = 2, KTOT = 512, NITEMS = 800;
    static constexpr int PARTS = 2, NSLOT = 2;      // raw frame fetched as two channel halves, one frame of slots
                                                    // (quarters: 0.205 vs 0.187 ms per step -- four half-empty rounds)
    __device__ static constexpr int shift(int kb) { return ((kb >> 1) >> 1) * 10 + ((kb >> 1) & 1); }
    __device__ static constexpr int plane(int kb) { return kb & 1; }
    __device__ static constexpr int nk(int) { return 4; }
    // element e of k-block kb  ->  index into W[co][ci][ky][kx] (row of 512), or -1
    __host__ __device__ static constexpr int widx(int kb, int e) {
        const int o = kb >> 1, dy = kb & 1, dx = e >> 5, ci = e & 31;
        const int ky = 2 * (o >> 1) + dy, kx = 2 * (o & 1) + dx;
        return ci * 16 + ky * 4 + kx;
    }
};
template <> struct SwCfg<3> {            // conv3
    static constexpr int HIN = 9, GRID = 9, HOUT = 7, NPOS = 81, PLANES = 1, NKB = 6;
    static constexpr int ROWS = 88, STAGES = 4, KTOT = 288, NITEMS = 324;
    static constexpr int PARTS = 1, NSLOT = 4;      // whole frames, four in flight
    __device__ static constexpr int shift(int kb) { return (kb >> 1) * 9 + 2 * (kb & 1); }
    __device__ static constexpr int plane(int) { return 0; }
    __device__ static constexpr int nk(int kb) { return (kb & 1) ? 2 : 4; }
    __host__ __device__ static constexpr int widx(int kb, int e) {
        const int ky = kb >> 1, part = kb & 1;
        const int kx = part ? 2 : (e >> 5), ci = e & 31;
        return (part && e >= 32) ? -1 : ci * 9 + ky * 3 + kx;
    }
};

template <int L>
__global__ void __launch_bounds__(SWC_THREADS, 1)
tc_conv_sw_fwd_kernel(const float* __restrict__ in, const float* __restrict__ w,
                      const float* __restrict__ bias, float* __restrict__ out, const int R,
                      __nv_bfloat16* __restrict__ out_hi, __nv_bfloat16* __restrict__ out_lo) {
    using C = SwCfg<L>;
    constexpr int COUT = 32, P = C::HOUT * C::HOUT, NKB = C::NKB, STAGES = C::STAGES;
    // One (hi or lo) plane of 64 channels: ROWS = the positions rounded up to whole 8-row swizzle
    // atoms.  An M = 128 tile read at a row shift runs past the plane's end into whatever follows
    // (the next plane, the next stage, and behind the ring the resident weights): initialised,
    // finite bf16 data that only feed accumulator rows nobody stores -- so the planes carry no
    // padding and a third stage fits.
    constexpr uint32_t PLANE = C::ROWS * 128;
    constexpr uint32_t STAGE_BYTES = C::PLANES * 2 * PLANE;      // [plane][hi|lo][ROWS x 128 B]
    constexpr uint32_t W_KB = 2 * COUT * 128;                    // [W_hi rows ; W_lo rows] x 64 k
    // accumulator columns of one buffer: [A_hi.W_hi | A_hi.W_lo | A_lo.W_hi] (the epilogue adds them)
    constexpr uint32_t SLICE = 2 * COUT, ACC_COLS = 128;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2], raw_full[C::NSLOT], raw_empty[C::NSLOT];
    __shared__ uint32_t tmem_slot;
    constexpr uint32_t FRAME_BYTES = 32 * C::HIN * C::HIN * 4;   // one frame of fp32 NCHW activations

    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t* a_s = base;                                         // ring first, weights behind it (see PLANE)
    uint8_t* w_s = base + STAGES * STAGE_BYTES;
    // raw fp32 activations, brought in by bulk copies into NSLOT slots of 1/PARTS frame each: the
    // producers read shared memory instead of waiting for global loads, and a slot is refilled
    // (with the part NSLOT positions ahead) the moment every producer has read it.  conv2 (51 KB
    // per frame) has room for one frame of slots: two channel halves, so that the fetch of the
    // next frame's first half starts after half a conversion.
    constexpr int PARTS = C::PARTS, NSLOT = C::NSLOT;
    constexpr uint32_t PART_BYTES = FRAME_BYTES / PARTS;
    constexpr int PART_ITEMS = C::NITEMS / PARTS;
    static_assert(PART_ITEMS <= SWC_PROD && PART_BYTES % 16 == 0, "one item per producer thread and part");
    const float* raw_s = reinterpret_cast<const float*>(w_s + NKB * W_KB);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int MMA_WARP = SWC_PROD / 32;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], SWC_PROD);
            mbar_init(&empty_bar[s], SWC_ISSUERS);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], SWC_ISSUERS);
            mbar_init(&acc_empty[b], 128);
        }
        for (int r = 0; r < NSLOT; ++r) {
            mbar_init(&raw_full[r], 1);
            mbar_init(&raw_empty[r], SWC_PROD);
        }
        mbar_fence_init();
    }
    // resident weights, K permuted to the image's channel order, split into (hi, lo), swizzled
    for (int i = tid; i < NKB * COUT * 8; i += blockDim.x) {
        const int kb = i / (COUT * 8), i2 = i - kb * (COUT * 8);
        const int co = i2 >> 3, c = i2 & 7;
        float v[8], l[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int wi = C::widx(kb, c * 8 + e);
            const float x = wi >= 0 ? __ldg(w + (long long)co * C::KTOT + wi) : 0.f;
            v[e] = x;
            l[e] = x - swc_round(x);
        }
        const uint32_t off = (uint32_t)kb * W_KB + swc_off(co, c);
        *reinterpret_cast<uint4*>(w_s + off) =
            make_uint4(swc_pack(v[0], v[1]), swc_pack(v[2], v[3]), swc_pack(v[4], v[5]), swc_pack(v[6], v[7]));
        *reinterpret_cast<uint4*>(w_s + off + COUT * 128) =
            make_uint4(swc_pack(l[0], l[1]), swc_pack(l[2], l[3]), swc_pack(l[4], l[5]), swc_pack(l[6], l[7]));
    }
    // rows the producers never write must not hold NaN patterns that could trip nothing but
    // debuggers; zero the whole ring once
    for (uint32_t i = tid * 16; i < STAGES * STAGE_BYTES; i += blockDim.x * 16)
        *reinterpret_cast<uint4*>(a_s + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 2 * ACC_COLS);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers: raw frame (shared memory, via bulk copy) -> the frame's image;
        // every activation is read and stored once.  Thread 0 also drives the bulk copies: frame
        // f+1 is requested as soon as every producer has finished reading frame f.
        uint32_t it = 0;
        // g = running part index of this CTA (frame it, part g % PARTS) -> slot g % NSLOT
        auto fetch = [&](long long g) {
            const long long f = (long long)blockIdx.x + (g / PARTS) * (long long)gridDim.x;
            if (f >= R) return;
            const int slot = (int)(g % NSLOT), part = (int)(g % PARTS);
            mbar_expect_tx(&raw_full[slot], PART_BYTES);
            bulk_g2s(const_cast<float*>(raw_s) + (size_t)slot * (PART_BYTES / 4),
                     in + f * (FRAME_BYTES / 4) + (size_t)part * (PART_BYTES / 4), PART_BYTES, &raw_full[slot]);
        };
        if (tid == 0)
            for (int g = 0; g < NSLOT; ++g) fetch(g);
        auto chunk = [](const float* v, uint4& h, uint4& l) {
            split_bf16x2(v[0], v[1], h.x, l.x); split_bf16x2(v[2], v[3], h.y, l.y);
            split_bf16x2(v[4], v[5], h.z, l.z); split_bf16x2(v[6], v[7], h.w, l.w);
        };
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int s = it % STAGES;
            mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);           // image stage free
            SWC_STAMP(0);
            uint8_t* st = a_s + (size_t)s * STAGE_BYTES;
#pragma unroll
            for (int part = 0; part < PARTS; ++part) {
                const long long g = (long long)it * PARTS + part;
                const int slot = (int)(g % NSLOT);
                const float* raw_p = raw_s + (size_t)slot * (PART_BYTES / 4);
                mbar_wait(&raw_full[slot], (uint32_t)(g / NSLOT) & 1);   // this part landed in shared memory
                const int i = tid;
                if (i < PART_ITEMS) {
                    uint4 h, l;
                    if constexpr (L == 2) {
                        // item = (ci-chunk, dy, position Y, X) of this channel half: 8 channels x (dx = 0, 1)
                        const int X = i % 10, Y = (i / 10) % 10, dy = (i / 100) & 1, cc2 = i / 200;
                        const int cc = part * (4 / PARTS) + cc2;
                        const float* q = raw_p + (cc2 * 8) * 400 + (2 * Y + dy) * 20 + 2 * X;
                        float v0[8], v1[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const float2 v = *reinterpret_cast<const float2*>(q + e * 400);
                            v0[e] = v.x;
                            v1[e] = v.y;
                        }
                        const int row = Y * 10 + X;
                        uint8_t* pl = st + (size_t)dy * 2 * PLANE;
                        const uint32_t o0 = swc_off(row, cc), o1 = swc_off(row, 4 + cc);   // dx = 0 | dx = 1
                        chunk(v0, h, l);
                        *reinterpret_cast<uint4*>(pl + o0) = h;
                        *reinterpret_cast<uint4*>(pl + PLANE + o0) = l;
                        chunk(v1, h, l);
                        *reinterpret_cast<uint4*>(pl + o1) = h;
                        *reinterpret_cast<uint4*>(pl + PLANE + o1) = l;
                    } else {
                        // item = (ci-chunk cc, position p): own row's first half, previous row's second half
                        const int p = i % 81, cc = i / 81;
                        const float* q = raw_p + (cc * 8) * 81 + p;
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = q[e * 81];
                        chunk(v, h, l);
                        const uint32_t o0 = swc_off(p, cc);
                        *reinterpret_cast<uint4*>(st + o0) = h;
                        *reinterpret_cast<uint4*>(st + PLANE + o0) = l;
                        if (p > 0) {
                            const uint32_t o1 = swc_off(p - 1, 4 + cc);
                            *reinterpret_cast<uint4*>(st + o1) = h;
                            *reinterpret_cast<uint4*>(st + PLANE + o1) = l;
                        }
                    }
                }
                mbar_arrive(&raw_empty[slot]);                           // done reading this slot
                if (tid == 0) {
                    mbar_wait(&raw_empty[slot], (uint32_t)(g / NSLOT) & 1);   // ... and so is everybody else
                    fetch(g + NSLOT);
                }
            }
            SWC_STAMP(7);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
            SWC_STAMP(1);
        }
    } else if (warp < MMA_WARP + SWC_ISSUERS) {
        // One thread issues a tcgen05.mma every ~74 clk whatever its shape and a frame takes 64 of
        // them, so the issue is split between two warps: warp 0 issues the A_hi MMAs (N = 64 against
        // [W_hi ; W_lo]), warp 1 the A_lo MMAs (N = 32 against W_hi), each into its own accumulator
        // columns.  With two issuers the tensor pipe itself is the limit: ~45 clk per M=128 MMA at
        // N <= 64 (tools/ubench/mma_rate.cu), i.e. ~1.5 us of MMA time per frame; the traced phases
        // (tools/diag_conv_sw.py) show the producers' 51 KB of loads per frame taking about as long.
        if (lane == 0) {
            const int issuer = warp - MMA_WARP;
            constexpr uint32_t idesc64 = umma_idesc_bf16(128, SLICE), idesc32 = umma_idesc_bf16(128, COUT);
            const uint32_t wb = smem_u32(w_s);
            uint32_t it = 0;
            for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
                const uint32_t buf = it & 1, use = it >> 1;
                const int s = it % STAGES;
                mbar_wait(&acc_empty[buf], (use & 1) ^ 1);
                mbar_wait(&full_bar[s], (it / STAGES) & 1);
                tc_fence_after();
                if (issuer == 0) SWC_STAMP(2);
                const uint32_t st = smem_u32(a_s + (size_t)s * STAGE_BYTES);
                if (issuer == 0) {
                    const uint32_t d = tmem_base + buf * ACC_COLS;
#pragma unroll
                    for (int kb = 0; kb < NKB; ++kb) {
                        const uint32_t a_hi = st + C::plane(kb) * 2 * PLANE + C::shift(kb) * 128;
                        const uint32_t b_hl = wb + kb * W_KB;
#pragma unroll
                        for (int k = 0; k < C::nk(kb); ++k)
                            umma_f16(d, swc_desc(a_hi + k * 32), umma_desc_k_sw128(b_hl + k * 32), idesc64, (kb | k) != 0);
                    }
                } else {
                    const uint32_t d = tmem_base + buf * ACC_COLS + SLICE;
#pragma unroll
                    for (int kb = 0; kb < NKB; ++kb) {
                        const uint32_t a_lo = st + C::plane(kb) * 2 * PLANE + C::shift(kb) * 128 + PLANE;
                        const uint32_t b_hl = wb + kb * W_KB;
#pragma unroll
                        for (int k = 0; k < C::nk(kb); ++k)
                            umma_f16(d, swc_desc(a_lo + k * 32), umma_desc_k_sw128(b_hl + k * 32), idesc32, (kb | k) != 0);
                    }
                }
                umma_commit(&empty_bar[s]);
                umma_commit(&acc_full[buf]);
                if (issuer == 0) SWC_STAMP(3); else SWC_STAMP(4);
            }
        }
    } else {
        // ================= epilogue: one image row (TMEM lane) per thread, 32 channels
        const int q = warp & 3;
        const int m = q * 32 + lane;
        const int Y = m / C::GRID, X = m - Y * C::GRID;
        const bool valid = m < C::NPOS && Y < C::HOUT && X < C::HOUT;
        const int p = Y * C::HOUT + X;
        float bs[COUT];
#pragma unroll
        for (int co = 0; co < COUT; ++co) bs[co] = __ldg(bias + co);
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const uint32_t buf = it & 1, use = it >> 1;
            mbar_wait(&acc_full[buf], use & 1);
            SWC_STAMP(5);
            tc_fence_after();
            uint32_t r[32], r2[32], r3[32];
            float acc[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * ACC_COLS, r);
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * ACC_COLS + COUT, r2);
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * ACC_COLS + SLICE, r3);
            tmem_ld_wait();
#pragma unroll
            for (int co = 0; co < COUT; ++co)
                acc[co] = (__uint_as_float(r[co]) + __uint_as_float(r2[co])) + __uint_as_float(r3[co]);
            tc_fence_before();
            mbar_arrive(&acc_empty[buf]);
            if (valid) {
                const long long oo = (long long)fr * COUT * P + p;
                if (out_hi) {
#pragma unroll
                    for (int co = 0; co < COUT; ++co) {
                        const float x = fmaxf(acc[co] + bs[co], 0.f);
                        const __nv_bfloat16 h = __float2bfloat16_rn(x);
                        out[oo + (long long)co * P] = x;
                        out_hi[oo + (long long)co * P] = h;
                        out_lo[oo + (long long)co * P] = __float2bfloat16_rn(x - __bfloat162float(h));
                    }
                } else {
#pragma unroll
                    for (int co = 0; co < COUT; ++co)
                        out[oo + (long long)co * P] = fmaxf(acc[co] + bs[co], 0.f);
                }
            }
            if (q == 3) SWC_STAMP(6);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 2 * ACC_COLS);
}

}  // namespace

template <int L>
int launch_tc_conv_sw(const char* name, const float* in, const float* w, const float* b, float* out, int R,
                      cudaStream_t st, __nv_bfloat16* out_hi, __nv_bfloat16* out_lo) {
    using C = SwCfg<L>;
    const int smem = C::NKB * 2 * 32 * 128 + C::STAGES * C::PLANES * 2 * C::ROWS * 128 + C::NSLOT * (32 * C::HIN * C::HIN * 4 / C::PARTS) + 1024;
    static_assert(C::NKB * 2 * 32 * 128 >= 48 * 128, "the weights behind the ring absorb the over-read");
    auto kern = tc_conv_sw_fwd_kernel<L>;
    R2D2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int grid = sm_count();
    if (grid > R) grid = R;
    R2D2_LAUNCH(name, st, kern<<<grid, SWC_THREADS, smem, st>>>(in, w, b, out, R, out_hi, out_lo));
    return R2D2_OK;
}

template int launch_tc_conv_sw<2>(const char*, const float*, const float*, const float*, float*, int, cudaStream_t,
                                  __nv_bfloat16*, __nv_bfloat16*);
template int launch_tc_conv_sw<3>(const char*, const float*, const float*, const float*, float*, int, cudaStream_t,
                                  __nv_bfloat16*, __nv_bfloat16*);

}  // namespace r2d2

#ifdef SWC_TRACE
extern "C" int r2d2_dbg_swc_trace(long long* host) {
    return (int)cudaMemcpyFromSymbol(host, r2d2::g_swc_trace, sizeof(long long) * 2 * 32 * 8);
}
#endif
