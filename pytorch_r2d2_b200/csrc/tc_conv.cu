// Conv torso forward (model.py:13-21,44-46) as an implicit GEMM on tcgen05 tensor cores.
//
//   out[r, co, p] = relu(bias[co] + sum_k patch[(r,p), k] * W[co, k]),  k = (ci, ky, kx)
//
// M = 128 output pixels per tile (one TMEM lane each), N = 32 output channels, K = Cin*Kh*Kw in
// blocks of 64.  Persistent CTAs, 416 threads:
//   warps 0-7  producers: two threads per row of the A tile (the pixel's receptive field), four
//              16-byte chunks each, all patch offsets compile-time constants; they gather
//              straight from the u8 slab / fp32 NCHW activations, splits it into bf16 (hi, lo)
//              and stores it into shared memory in the 128-byte-swizzled K-major layout the UMMA
//              descriptor expects (what TMA would have produced for a dense matrix);
//   warp 8     TMEM allocator + MMA issuer (one lane): error-compensated bf16x3
//              (A_hi.W_hi + A_hi.W_lo + A_lo.W_hi, fp32 accumulate in TMEM); for conv1 the u8
//              pixels are exact in bf16, so A has no lo part and 1/255 is applied in the epilogue;
//   warps 9-12 epilogue: tcgen05.ld -> bias + ReLU -> NCHW fp32 (coalesced along pixels).
// The weights (both halves, all K) stay resident in shared memory; two TMEM accumulator
// buffers let the epilogue of tile i overlap the MMAs of tile i+1.
#include <cuda_bf16.h>
#include <stdlib.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "gemm.cuh"   // FrameMap
#include "ptx.cuh"

namespace r2d2 {

// tc_conv1_sw.cu
int launch_tc_conv1_sw_dual(const uint8_t* in, FrameMap fm, const float* wa, const float* ba, float* outa,
                            const float* wb, const float* bb, float* outb, int R, int skip_b, cudaStream_t st,
                            uint32_t* relu_bits_a);
// tc_wgrad_sw.cu
int launch_tc_conv1_wgrad_sw(const uint8_t* in, FrameMap fm, const float* dz, float* part, int R, float in_scale,
                             int grid, cudaStream_t st, float* bias_part);
int launch_tc_conv2_wgrad_sw(const float* in, const float* dz, float* part, int R, int grid, cudaStream_t st,
                             float* bias_part);
int launch_tc_conv3_wgrad_sw(const float* in, const float* dz, float* part, int R, int grid, cudaStream_t st,
                             float* bias_part);
// tc_dgrad_sw.cu
int launch_tc_conv2_dgrad_sw(const float* dz2, const float* w2, const uint32_t* maskbits, float* dz1, int R,
                             cudaStream_t st);
int launch_tc_conv3_dgrad_sw(const float* dz3, const float* w3, const float* a2, float* dz2, int R, cudaStream_t st);
// tc_conv_sw.cu
template <int L>
int launch_tc_conv_sw(const char* name, const float* in, const float* w, const float* b, float* out, int R,
                      cudaStream_t st, __nv_bfloat16* out_hi, __nv_bfloat16* out_lo);

constexpr int CV_STAGES = 4;

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t pack_f16x2(float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
// two u8 pixels (bytes 2*PAIR, 2*PAIR+1 of w) -> packed fp16x2, exact, two instructions: splice each
// byte under the exponent of 1024.0h (0x64xx = 1024 + x) with one PRMT, subtract (1024, 1024)
template <int PAIR>
__device__ __forceinline__ uint32_t u8x2_to_f16x2(uint32_t w) {
    const uint32_t m = __byte_perm(w, 0x64646464u, PAIR == 0 ? 0x4140u : 0x4342u);
    const __half2 r = __hsub2(*reinterpret_cast<const __half2*>(&m), __half2half2(__ushort_as_half(0x6400)));
    return *reinterpret_cast<const uint32_t*>(&r);
}
__device__ __forceinline__ float bf16_round(float x) {
    return __bfloat162float(__float2bfloat16_rn(x));
}
// byte offset of (row, 16-byte chunk) inside a K-major SWIZZLE_128B tile of 128-byte rows
__device__ __forceinline__ uint32_t sw128_off(int row, int chunk) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}

// NCH 8-element chunks (chunks NCH*PART .. NCH*PART+NCH-1 of k-block kb) of one patch row.
template <class T, int CIN, int KH, int HIN, int PART, int NCH = 4>
__device__ __forceinline__ void gather_chunks(const T* __restrict__ src, const int kb,
                                              const bool valid, float (&v)[NCH][8]) {
    constexpr int K = CIN * KH * KH;
#pragma unroll
    for (int cc = 0; cc < NCH; ++cc) {
        const int c = PART * NCH + cc;
        if constexpr (sizeof(T) == 1 && KH == 8) {
            // conv1 on the u8 slab: k-block = one input channel, chunk = one 8-pixel patch row
            const uint32_t* q = reinterpret_cast<const uint32_t*>(src + kb * (HIN * HIN) + c * HIN);
            const uint32_t lo4 = valid ? __ldg(q) : 0u, hi4 = valid ? __ldg(q + 1) : 0u;
            // u8 -> float without the (quarter-rate) I2F pipe: splice the byte into the mantissa
            // of 2^23 with one PRMT, subtract 2^23 with one FADD (exact for 0..255)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                v[cc][e] = __uint_as_float(__byte_perm(lo4, 0x4B000000u, 0x7650u | e)) - 8388608.f;
                v[cc][4 + e] = __uint_as_float(__byte_perm(hi4, 0x4B000000u, 0x7650u | e)) - 8388608.f;
            }
        } else if constexpr (KH == 4) {
            // conv2: chunk = (ci = 4kb + c/2, ky = 2(c%2) .. +1, kx = 0..3): two rows of 4 floats
            const T* q = src + (kb * 4 + (c >> 1)) * (HIN * HIN) + ((c & 1) * 2) * HIN;
#pragma unroll
            for (int rr = 0; rr < 2; ++rr) {
                const float2* q2 = reinterpret_cast<const float2*>(q + rr * HIN);
                const float2 x0 = valid ? __ldg(q2) : make_float2(0.f, 0.f);
                const float2 x1 = valid ? __ldg(q2 + 1) : make_float2(0.f, 0.f);
                v[cc][rr * 4 + 0] = x0.x; v[cc][rr * 4 + 1] = x0.y;
                v[cc][rr * 4 + 2] = x1.x; v[cc][rr * 4 + 3] = x1.y;
            }
        } else {
            // generic (conv3, f32-frame conv1): scalar loads; kb is unrolled by the caller, so k
            // and the (ci,ky,kx) decode are compile-time constants
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int k = kb * 64 + c * 8 + e;
                const int ci = k / (KH * KH), rem = k - ci * (KH * KH);
                const int ky = rem / KH, kx = rem - ky * KH;
                float x = 0.f;
                if (k < K && valid) x = (float)__ldg(src + ci * HIN * HIN + ky * HIN + kx);
                v[cc][e] = x;
            }
        }
    }
}

// NW = 2 (conv1 on the u8 slab only): TWO networks' first layers in one pass.  The online and the
// target net read the same frames, and the kernel is bound by the producers' gather, so the A
// tile is built once and multiplied against both weight sets stacked along N (one N = 128 MMA per
// k-step); net 1 skips the first `skip1` frames (its window starts later) and writes out1.
// NPW = producer warps (8 or 16): the gather is bound by memory latency, so the fp32 layers run 16
// (two chunks per thread, twice the loads in flight per SM at half the registers per thread).
template <class T, int CIN, int KH, int HIN, int HOUT, int STRIDE, int NW = 1, int NPW = 8>
__global__ void __launch_bounds__(NPW * 32 + 160, 1)
tc_conv_fwd_kernel(const T* __restrict__ in, const FrameMap fm, const float* __restrict__ w,
                   const float* __restrict__ bias, float* __restrict__ out, const int R,
                   const float in_scale, const float* __restrict__ w1 = nullptr,
                   const float* __restrict__ bias1 = nullptr, float* __restrict__ out1 = nullptr,
                   const int skip1 = 0, __nv_bfloat16* __restrict__ out_hi = nullptr,
                   __nv_bfloat16* __restrict__ out_lo = nullptr, uint32_t* __restrict__ relu_bits = nullptr) {
    constexpr int K = CIN * KH * KH, NKB = (K + 63) / 64, KP = NKB * 64;
    constexpr int P = HOUT * HOUT, COUT = 32;
    constexpr bool NEED_LO = sizeof(T) != 1;                 // u8 pixels are exact in bf16
    static_assert(NW == 1 || !NEED_LO, "the two-network pass is for the u8 first layer");
    constexpr uint32_t A_TILE = 128 * 128;                   // 128 rows x 64 bf16
    constexpr uint32_t STAGE_BYTES = NEED_LO ? 2 * A_TILE : A_TILE;
    constexpr uint32_t W_KB = COUT * 128;                    // 32 rows x 64 bf16 per k-block and half
    constexpr uint32_t W_BYTES = NW * 2 * NKB * W_KB;        // per k-block: NW x [hi rows | lo rows]
    constexpr uint32_t ACC_COLS = NW * 2 * COUT;             // accumulator columns per buffer
    // u8 frames: operands in fp16 instead of bf16 -- pixels 0..255 are exact in both, but a byte
    // becomes fp16 with ONE PRMT + half an HSUB2 (bf16 needs PRMT + FADD + half an F2FP), and the
    // weights' (hi, lo) split keeps 22 mantissa bits instead of 16
    constexpr bool U8F16 = sizeof(T) == 1 && KH == 8;
    constexpr int MMA_WARP = NPW;
    constexpr int NCH = 32 / NPW;                            // 16-byte chunks of a row per producer thread
    static_assert(NPW == 8 || NPW == 16, "8 or 16 producer warps");
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[CV_STAGES], empty_bar[CV_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;

    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t* w_s = base;                                     // [NKB][hi|lo][32 x 128 B]
    uint8_t* a_s = base + W_BYTES;                           // CV_STAGES x STAGE_BYTES

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long Mtot = (long long)R * P;
    const int ntiles = (int)((Mtot + 127) / 128);

    // ---- one-time set-up: barriers, k-offset table, resident weights (hi/lo, swizzled)
    if (tid == 0) {
        for (int s = 0; s < CV_STAGES; ++s) {
            mbar_init(&full_bar[s], NPW * 32);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
        }
        mbar_fence_init();
    }
    for (int i = tid; i < NW * COUT * (KP / 8); i += blockDim.x) {
        const int net = i / (COUT * (KP / 8)), i2 = i - net * (COUT * (KP / 8));
        const int co = i2 / (KP / 8), ch = i2 - co * (KP / 8);  // chunk of 8 k's
        const int kb = ch >> 3, c = ch & 7;
        const float* wsrc = net == 0 ? w : w1;
        float v[8], l[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int k = ch * 8 + e;
            const float x = k < K ? __ldg(wsrc + (long long)co * K + k) : 0.f;
            v[e] = x;
            l[e] = x - (U8F16 ? __half2float(__float2half_rn(x)) : bf16_round(x));
        }
        // rows 0..31 of a net's k-block tile = W_hi, rows 32..63 = W_lo: ONE K-major operand of
        // 64*NW rows (1 KB swizzle atoms of 8 rows, so the per-net 8 KB offset keeps the pattern)
        const uint32_t off = (uint32_t)(kb * NW + net) * 2 * W_KB + sw128_off(co, c);
        if constexpr (U8F16) {
            *reinterpret_cast<uint4*>(w_s + off) = make_uint4(pack_f16x2(v[0], v[1]), pack_f16x2(v[2], v[3]),
                                                             pack_f16x2(v[4], v[5]), pack_f16x2(v[6], v[7]));
            *reinterpret_cast<uint4*>(w_s + off + W_KB) =
                make_uint4(pack_f16x2(l[0], l[1]), pack_f16x2(l[2], l[3]), pack_f16x2(l[4], l[5]),
                           pack_f16x2(l[6], l[7]));
        } else {
        *reinterpret_cast<uint4*>(w_s + off) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                                         pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
        *reinterpret_cast<uint4*>(w_s + off + W_KB) =
            make_uint4(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]), pack_bf16x2(l[4], l[5]),
                       pack_bf16x2(l[6], l[7]));
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // weights -> tensor-core proxy
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 2 * ACC_COLS);            // 2 buffers x NW x (32 + 32) columns
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers: thread t -> row (t & 127), chunks 4*(t>>7) .. +3
        const int row = tid & 127, half = tid >> 7;
        uint32_t it = 0;
        if constexpr (sizeof(T) == 1 && KH == 8) {
            // conv1 on the u8 slab.  The gather is memory-latency bound: all 32 four-byte loads of
            // a tile row (4 input channels x this thread's 4 patch rows x 8 pixels) are issued
            // together, and the NEXT tile's loads are issued before the current tile is converted
            // and stored, so that DRAM latency overlaps the ring traffic of the current tile.
            auto issue = [&](int tile, uint32_t (&dst)[NKB][NCH][2]) {
                const long long m = (long long)tile * 128 + row;
                const bool valid = tile < ntiles && m < Mtot;
                const T* src = in;
                if (valid) {
                    const int r = (int)(m / P), p = (int)(m - (long long)r * P);
                    const int oy = p / HOUT, ox = p - oy * HOUT;
                    src = in + fm.base(r) + (long long)(oy * STRIDE) * HIN + ox * STRIDE;
                }
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb)
#pragma unroll
                    for (int cc = 0; cc < NCH; ++cc) {
                        const uint32_t* q = reinterpret_cast<const uint32_t*>(
                            src + kb * (HIN * HIN) + (half * NCH + cc) * HIN);
                        dst[kb][cc][0] = valid ? __ldg(q) : 0u;
                        dst[kb][cc][1] = valid ? __ldg(q + 1) : 0u;
                    }
            };
            uint32_t cur[NKB][NCH][2], nxt[NKB][NCH][2];
            issue(blockIdx.x, cur);
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                issue(tile + gridDim.x, nxt);
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb, ++it) {
                    const int s = it % CV_STAGES;
                    mbar_wait(&empty_bar[s], ((it / CV_STAGES) & 1) ^ 1);
                    uint8_t* a_hi = a_s + (size_t)s * STAGE_BYTES;
#pragma unroll
                    for (int cc = 0; cc < NCH; ++cc) {
                        *reinterpret_cast<uint4*>(a_hi + sw128_off(row, half * NCH + cc)) =
                            make_uint4(u8x2_to_f16x2<0>(cur[kb][cc][0]), u8x2_to_f16x2<1>(cur[kb][cc][0]),
                                       u8x2_to_f16x2<0>(cur[kb][cc][1]), u8x2_to_f16x2<1>(cur[kb][cc][1]));
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    mbar_arrive(&full_bar[s]);
                }
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb)
#pragma unroll
                    for (int cc = 0; cc < NCH; ++cc) {
                        cur[kb][cc][0] = nxt[kb][cc][0];
                        cur[kb][cc][1] = nxt[kb][cc][1];
                    }
            }
        } else {
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const long long m = (long long)tile * 128 + row;
            const bool valid = m < Mtot;
            const T* src = in;                                // any valid address when !valid
            if (valid) {
                const int r = (int)(m / P), p = (int)(m - (long long)r * P);
                const int oy = p / HOUT, ox = p - oy * HOUT;
                src = in + fm.base(r) + (long long)(oy * STRIDE) * HIN + ox * STRIDE;
            }
            {
                // software pipeline: the loads of the next PF k-blocks are in flight while k-block kb
                // is converted and stored (the gather is memory-latency bound)
                constexpr int PF = NPW == 16 ? 2 : 1;             // k-blocks of loads in flight ahead
                float v[PF + 1][NCH][8];
                auto gather = [&](int kb, float (&dst)[NCH][8]) {
                    // `half` (= which group of NCH chunks) selects a compile-time instantiation
                    if (half == 0) gather_chunks<T, CIN, KH, HIN, 0, NCH>(src, kb, valid, dst);
                    else if (half == 1) gather_chunks<T, CIN, KH, HIN, 1, NCH>(src, kb, valid, dst);
                    else if constexpr (NPW == 16) {
                        if (half == 2) gather_chunks<T, CIN, KH, HIN, 2, NCH>(src, kb, valid, dst);
                        else gather_chunks<T, CIN, KH, HIN, 3, NCH>(src, kb, valid, dst);
                    }
                };
#pragma unroll
                for (int kb = 0; kb < PF && kb < NKB; ++kb) gather(kb, v[kb]);
#pragma unroll
                for (int kb = 0; kb < NKB; ++kb, ++it) {
                    if (kb + PF < NKB) gather(kb + PF, v[(kb + PF) % (PF + 1)]);
                    const int s = it % CV_STAGES;
                    mbar_wait(&empty_bar[s], ((it / CV_STAGES) & 1) ^ 1);
                    uint8_t* a_hi = a_s + (size_t)s * STAGE_BYTES;
                    float(&vv)[NCH][8] = v[kb % (PF + 1)];
#pragma unroll
                    for (int cc = 0; cc < NCH; ++cc) {
                        const uint32_t off = sw128_off(row, half * NCH + cc);
                        if constexpr (NEED_LO) {
                            uint4 h, l;
                            split_bf16x2(vv[cc][0], vv[cc][1], h.x, l.x);
                            split_bf16x2(vv[cc][2], vv[cc][3], h.y, l.y);
                            split_bf16x2(vv[cc][4], vv[cc][5], h.z, l.z);
                            split_bf16x2(vv[cc][6], vv[cc][7], h.w, l.w);
                            *reinterpret_cast<uint4*>(a_hi + off) = h;
                            *reinterpret_cast<uint4*>(a_hi + A_TILE + off) = l;
                        } else {
                            *reinterpret_cast<uint4*>(a_hi + off) =
                                make_uint4(pack_bf16x2(vv[cc][0], vv[cc][1]), pack_bf16x2(vv[cc][2], vv[cc][3]),
                                           pack_bf16x2(vv[cc][4], vv[cc][5]), pack_bf16x2(vv[cc][6], vv[cc][7]));
                        }
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    mbar_arrive(&full_bar[s]);
                }
            }
        }
        }
    } else if (warp == MMA_WARP) {
        // ================= MMA issuer
        if (lane == 0) {
            // The tensor pipe spends ~44 cycles on an MMA this small whatever its N, and one thread
            // issues one every ~37 ns, so the count matters: A_hi multiplies both W_hi and W_lo,
            // hence ONE MMA with N = 64 against the stacked [W_hi ; W_lo] rows (columns 0-31 and
            // 32-63 of the accumulator, added in the epilogue) instead of two with N = 32.
            constexpr uint32_t idesc64 = U8F16 ? umma_idesc_f16(128, ACC_COLS) : umma_idesc_bf16(128, ACC_COLS),
                               idesc32 = umma_idesc_bf16(128, COUT);
            const uint32_t wb = smem_u32(w_s);
            uint32_t it = 0, tcount = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tcount) {
                const uint32_t buf = tcount & 1, use = tcount >> 1;
                mbar_wait(&acc_empty[buf], (use & 1) ^ 1);           // epilogue drained this buffer
                tc_fence_after();
                const uint32_t d = tmem_base + buf * ACC_COLS;
                for (int kb = 0; kb < NKB; ++kb, ++it) {
                    const int s = it % CV_STAGES;
                    mbar_wait(&full_bar[s], (it / CV_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(a_s + (size_t)s * STAGE_BYTES), a_lo = a_hi + A_TILE;
                    const uint32_t b_hl = wb + kb * NW * 2 * W_KB;   // [W_hi ; W_lo] (x NW) of this k-block
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        umma_f16(d, umma_desc_k_sw128(a_hi + k * 32), umma_desc_k_sw128(b_hl + k * 32),
                                 idesc64, (kb | k) != 0);
                        if (NEED_LO)
                            umma_f16(d, umma_desc_k_sw128(a_lo + k * 32),
                                     umma_desc_k_sw128(b_hl + k * 32), idesc32, 1);
                    }
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        // ================= epilogue: one pixel (TMEM lane) per thread, 32 channels
        const int q = warp & 3;
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tcount) {
            const uint32_t buf = tcount & 1, use = tcount >> 1;
            mbar_wait(&acc_full[buf], use & 1);
            tc_fence_after();
            const long long m = (long long)tile * 128 + q * 32 + lane;
            const int fr = (int)(m / P), p = (int)(m - (long long)fr * P);
#pragma unroll
            for (int net = 0; net < NW; ++net) {
                uint32_t r[32], r2[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * ACC_COLS + net * 2 * COUT, r);
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * ACC_COLS + net * 2 * COUT + COUT, r2);
                tmem_ld_wait();
#pragma unroll
                for (int c = 0; c < 32; ++c) r[c] = __float_as_uint(__uint_as_float(r[c]) + __uint_as_float(r2[c]));
                if (net == NW - 1) {
                    tc_fence_before();
                    mbar_arrive(&acc_empty[buf]);
                }
                const int fo = net == 0 ? fr : fr - skip1;         // net 1 starts skip1 frames later
                if (m < Mtot && fo >= 0) {
                    const long long oo = (long long)fo * COUT * P + p;
                    float* o = (net == 0 ? out : out1) + oo;
                    const float* bs = net == 0 ? bias : bias1;
                    if (NW == 1 && out_hi) {
                        // also emit the bf16 (hi, lo) split the next contraction consumes
#pragma unroll
                        for (int co = 0; co < COUT; ++co) {
                            const float x = fmaxf(fmaf(__uint_as_float(r[co]), in_scale, __ldg(bs + co)), 0.f);
                            const __nv_bfloat16 h = __float2bfloat16_rn(x);
                            o[(long long)co * P] = x;
                            out_hi[oo + (long long)co * P] = h;
                            out_lo[oo + (long long)co * P] = __float2bfloat16_rn(x - __bfloat162float(h));
                        }
                    } else {
                        uint32_t bits = 0u;
#pragma unroll
                        for (int co = 0; co < COUT; ++co) {
                            const float x = fmaxf(fmaf(__uint_as_float(r[co]), in_scale, __ldg(bs + co)), 0.f);
                            o[(long long)co * P] = x;
                            bits |= (x > 0.f ? 1u : 0u) << co;
                        }
                        // ReLU' of net 0 as one word per frame and pixel (bit co), for the data
                        // gradient that flows back into this layer's output
                        if (net == 0 && relu_bits) relu_bits[(long long)fo * P + p] = bits;
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 2 * ACC_COLS);
}

template <class T, int CIN, int KH, int HIN, int HOUT, int STRIDE>
static int launch_tc_conv(const char* name, const T* in, FrameMap fm, const float* w, const float* b,
                          float* out, int R, float in_scale, cudaStream_t st,
                          __nv_bfloat16* out_hi = nullptr, __nv_bfloat16* out_lo = nullptr) {
    constexpr int K = CIN * KH * KH, NKB = (K + 63) / 64;
    constexpr bool NEED_LO = sizeof(T) != 1;
    constexpr int NPW = (sizeof(T) == 1 && KH == 8) ? 8 : 16;
    const int smem = 2 * NKB * 32 * 128 + CV_STAGES * (NEED_LO ? 2 : 1) * 128 * 128 + 1024;
    auto kern = tc_conv_fwd_kernel<T, CIN, KH, HIN, HOUT, STRIDE, 1, NPW>;
    R2D2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const long long Mtot = (long long)R * HOUT * HOUT;
    const int ntiles = (int)((Mtot + 127) / 128);
    int grid = sm_count();
    if (grid > ntiles) grid = ntiles;
    R2D2_LAUNCH(name, st, kern<<<grid, NPW * 32 + 160, smem, st>>>(in, fm, w, b, out, R, in_scale, nullptr, nullptr, nullptr, 0, out_hi, out_lo, nullptr));
    return R2D2_OK;
}

// first conv layer of two networks over the same u8 frames (see the kernel's NW parameter)
static int launch_tc_conv1_dual(const uint8_t* in, FrameMap fm, const float* wa, const float* ba, float* outa,
                                const float* wb, const float* bb, float* outb, int R, int skip_b,
                                cudaStream_t st, uint32_t* relu_bits_a = nullptr) {
    // shifted-window kernel (tc_conv1_sw.cu) unless R2D2_CONV1=gather (A/B measurements)
    static const bool gather_form = [] {
        const char* e = getenv("R2D2_CONV1");
        return e && e[0] == 'g';
    }();
    if (!gather_form) return launch_tc_conv1_sw_dual(in, fm, wa, ba, outa, wb, bb, outb, R, skip_b, st, relu_bits_a);
    constexpr int NKB = 4;
    const int smem = 2 * 2 * NKB * 32 * 128 + CV_STAGES * 128 * 128 + 1024;
    constexpr int NPW = 8;     // measured: 16 producer warps are no faster here (0.260 vs 0.245 ms)
    auto kern = tc_conv_fwd_kernel<uint8_t, 4, 8, 84, 20, 4, 2, NPW>;
    R2D2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const long long Mtot = (long long)R * 400;
    const int ntiles = (int)((Mtot + 127) / 128);
    int grid = sm_count();
    if (grid > ntiles) grid = ntiles;
    R2D2_LAUNCH("tc_conv1_fwd_dual", st, kern<<<grid, NPW * 32 + 160, smem, st>>>(in, fm, wa, ba, outa, R, 1.f / 255.f, wb, bb, outb, skip_b, nullptr, nullptr, relu_bits_a));
    return R2D2_OK;
}

// ------------------------------------------------------------------------------------------
// Conv weight gradient on tensor cores (plain bf16; backward contractions need no compensation):
//   dW[co, k] = sum_{pixels} dZ[pix, co] * patch[pix, k]
// computed as D[k (M), co (N=32)] = sum_pix A[k, pix] * B[co, pix].  The A tile is the SAME
// shared-memory image the forward producer builds (64 pixel rows x 128-byte rows of 64 patch
// elements, 128B-swizzled) -- consumed here as an MN-major operand (M = patch element, K = pixel);
// B is a K-major [32 co x 64 pixels] tile of dZ.  Each persistent CTA owns a contiguous range of
// 64-pixel blocks, accumulates its partial dW in TMEM (KP/128 accumulators of 32 columns) and
// writes it out once; the partials are then summed in fixed order (reduce_partials_kernel).
template <class T, int CIN, int KH, int HIN, int SUB>
__device__ __forceinline__ void wgrad_gather(const T* __restrict__ src, const bool valid, const int prow,
                                             uint8_t* __restrict__ a_stage) {
    constexpr int K = CIN * KH * KH, NKB = (K + 63) / 64;
    // half-k-blocks SUB, SUB+4, ... of this pixel's patch row (compile-time indices)
#pragma unroll
    for (int hk = SUB; hk < 2 * NKB; hk += 4) {
        constexpr int HALF = SUB & 1;
        const int kb = hk >> 1;
        float v[4][8];
        gather_chunks<T, CIN, KH, HIN, HALF>(src, kb, valid, v);
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)
            *reinterpret_cast<uint4*>(a_stage + kb * 8192 + sw128_off(prow, HALF * 4 + cc)) =
                make_uint4(pack_bf16x2(v[cc][0], v[cc][1]), pack_bf16x2(v[cc][2], v[cc][3]),
                           pack_bf16x2(v[cc][4], v[cc][5]), pack_bf16x2(v[cc][6], v[cc][7]));
    }
}

template <class T, int CIN, int KH, int HIN, int HOUT, int STRIDE, int STAGES>
__global__ void __launch_bounds__(416, 1)
tc_conv_wgrad_kernel(const T* __restrict__ in, const FrameMap fm, const float* __restrict__ dz,
                     float* __restrict__ part, const int R, const float in_scale,
                     float* __restrict__ bias_part) {
    constexpr int K = CIN * KH * KH, NKB = (K + 63) / 64;
    constexpr int MT = (NKB + 1) / 2, KPM = MT * 128;           // M tiles of 128 patch elements
    constexpr int P = HOUT * HOUT, COUT = 32;
    constexpr uint32_t A_STAGE = 2 * MT * 8192, B_TILE = COUT * 128, STAGE_BYTES = A_STAGE + B_TILE;
    constexpr int MMA_WARP = 8;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full;
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long Mtot = (long long)R * P;
    const int nblk = (int)((Mtot + 63) / 64);
    const int per = (nblk + gridDim.x - 1) / gridDim.x;
    const int blk_lo = min(nblk, (int)blockIdx.x * per), blk_hi = min(nblk, blk_lo + per);

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 256);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_full, 1);
        mbar_fence_init();
    }
    if (NKB < 2 * MT) {   // odd number of 64-wide groups: the padding group must hold finite values
        for (int i = tid; i < STAGES * 8192 / 16; i += blockDim.x) {
            const int s = i / (8192 / 16), o = i - s * (8192 / 16);
            *reinterpret_cast<uint4*>(base + (size_t)s * STAGE_BYTES + (2 * MT - 1) * 8192 + o * 16) =
                make_uint4(0u, 0u, 0u, 0u);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers: A: pixel = tid & 63, quarter = tid >> 6; B: (co, chunk) = tid
        const int prow = tid & 63, sub = tid >> 6;
        const int bco = tid >> 3, bch = tid & 7;
        uint32_t it = 0;
        // the producers see every dZ value of their channel anyway: the bias gradient (sum over
        // frames and pixels, in fp32 before the bf16 rounding) falls out of the operand build
        float bsum = 0.f;
        if constexpr (sizeof(T) == 1 && KH == 8 && P % 8 == 0) {
            // first layer on the u8 slab: the gather is memory-latency bound, so the loads of block
            // blk+1 (16 words of patch + 8 gradient values per thread) are issued before block blk
            // is converted and stored.  Thread (pixel prow, sub) owns half-k-blocks sub and sub+4:
            // input channel ci = sub/2 (+2), patch rows 4*(sub%2) .. +3, eight pixels each.
            static_assert(NKB == 4, "u8 first layer: K = 256");
            auto issue = [&](int blk, uint32_t (&raw)[2][4][2], float4 (&g)[2]) {
                const long long m = (long long)blk * 64 + prow;
                const bool valid = blk < blk_hi && m < Mtot;
                const T* src = in;
                if (valid) {
                    const int r = (int)(m / P), p = (int)(m - (long long)r * P);
                    const int oy = p / HOUT, ox = p - oy * HOUT;
                    src = in + fm.base(r) + (long long)(oy * STRIDE) * HIN + ox * STRIDE;
                }
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int cc = 0; cc < 4; ++cc) {
                        const int kb = (sub >> 1) + 2 * h, c = (sub & 1) * 4 + cc;
                        const uint32_t* q = reinterpret_cast<const uint32_t*>(src + kb * (HIN * HIN) + c * HIN);
                        raw[h][cc][0] = valid ? __ldg(q) : 0u;
                        raw[h][cc][1] = valid ? __ldg(q + 1) : 0u;
                    }
                const long long mm = (long long)blk * 64 + bch * 8;
                g[0] = g[1] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (blk < blk_hi && mm < Mtot) {      // P % 8 == 0: the 8 pixels share a frame, 32-byte aligned
                    const int r = (int)(mm / P), p = (int)(mm - (long long)r * P);
                    const float4* gp = reinterpret_cast<const float4*>(dz + ((long long)r * COUT + bco) * P + p);
                    g[0] = __ldg(gp);
                    g[1] = __ldg(gp + 1);
                }
            };
            uint32_t cur[2][4][2], nxt[2][4][2];
            float4 gcur[2], gnxt[2];
            issue(blk_lo, cur, gcur);
            for (int blk = blk_lo; blk < blk_hi; ++blk, ++it) {
                issue(blk + 1, nxt, gnxt);
                const int s = it % STAGES;
                mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
                uint8_t* a_stage = base + (size_t)s * STAGE_BYTES;
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int cc = 0; cc < 4; ++cc) {
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            v[e] = __uint_as_float(__byte_perm(cur[h][cc][0], 0x4B000000u, 0x7650u | e)) - 8388608.f;
                            v[4 + e] = __uint_as_float(__byte_perm(cur[h][cc][1], 0x4B000000u, 0x7650u | e)) - 8388608.f;
                        }
                        const int kb = (sub >> 1) + 2 * h;
                        *reinterpret_cast<uint4*>(a_stage + kb * 8192 + sw128_off(prow, (sub & 1) * 4 + cc)) =
                            make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                       pack_bf16x2(v[6], v[7]));
                    }
                *reinterpret_cast<uint4*>(a_stage + A_STAGE + sw128_off(bco, bch)) =
                    make_uint4(pack_bf16x2(gcur[0].x, gcur[0].y), pack_bf16x2(gcur[0].z, gcur[0].w),
                               pack_bf16x2(gcur[1].x, gcur[1].y), pack_bf16x2(gcur[1].z, gcur[1].w));
                bsum += ((gcur[0].x + gcur[0].y) + (gcur[0].z + gcur[0].w)) +
                        ((gcur[1].x + gcur[1].y) + (gcur[1].z + gcur[1].w));
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(&full_bar[s]);
#pragma unroll
                for (int h = 0; h < 2; ++h)
#pragma unroll
                    for (int cc = 0; cc < 4; ++cc) {
                        cur[h][cc][0] = nxt[h][cc][0];
                        cur[h][cc][1] = nxt[h][cc][1];
                    }
                gcur[0] = gnxt[0];
                gcur[1] = gnxt[1];
            }
        } else {
        for (int blk = blk_lo; blk < blk_hi; ++blk, ++it) {
            const int s = it % STAGES;
            mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
            uint8_t* a_stage = base + (size_t)s * STAGE_BYTES;
            // ---- A: this pixel's patch row
            const long long m = (long long)blk * 64 + prow;
            const bool valid = m < Mtot;
            const T* src = in;
            if (valid) {
                const int r = (int)(m / P), p = (int)(m - (long long)r * P);
                const int oy = p / HOUT, ox = p - oy * HOUT;
                src = in + fm.base(r) + (long long)(oy * STRIDE) * HIN + ox * STRIDE;
            }
            switch (sub) {
                case 0: wgrad_gather<T, CIN, KH, HIN, 0>(src, valid, prow, a_stage); break;
                case 1: wgrad_gather<T, CIN, KH, HIN, 1>(src, valid, prow, a_stage); break;
                case 2: wgrad_gather<T, CIN, KH, HIN, 2>(src, valid, prow, a_stage); break;
                default: wgrad_gather<T, CIN, KH, HIN, 3>(src, valid, prow, a_stage); break;
            }
            // ---- B: dZ[., co, 8 consecutive pixels]
            {
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const long long mm = (long long)blk * 64 + bch * 8 + e;
                    float x = 0.f;
                    if (mm < Mtot) {
                        const int r = (int)(mm / P), p = (int)(mm - (long long)r * P);
                        x = __ldg(dz + ((long long)r * COUT + bco) * P + p);
                    }
                    v[e] = x;
                }
                *reinterpret_cast<uint4*>(a_stage + A_STAGE + sw128_off(bco, bch)) =
                    make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                               pack_bf16x2(v[6], v[7]));
                bsum += ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
        }
        }
        if (bias_part) {   // the 8 chunk-threads of a channel are neighbouring lanes
            bsum += __shfl_xor_sync(FULL, bsum, 1);
            bsum += __shfl_xor_sync(FULL, bsum, 2);
            bsum += __shfl_xor_sync(FULL, bsum, 4);
            if (bch == 0) bias_part[blockIdx.x * COUT + bco] = bsum;
        }
    } else if (warp == MMA_WARP) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_amn(128, COUT);
            uint32_t it = 0;
            for (int blk = blk_lo; blk < blk_hi; ++blk, ++it) {
                const int s = it % STAGES;
                mbar_wait(&full_bar[s], (it / STAGES) & 1);
                tc_fence_after();
                const uint32_t a0 = smem_u32(base + (size_t)s * STAGE_BYTES), b0 = a0 + A_STAGE;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)                      // 16 pixels per MMA
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt)
                        umma_f16(tmem_base + mt * COUT,
                                 umma_desc_mn_sw128(a0 + mt * 16384 + ks * 2048, 8192),
                                 umma_desc_k_sw128(b0 + ks * 32), idesc, (it | ks) != 0);
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_full);
        }
    } else {
        // ================= epilogue (once): lane = patch element, 32 columns = co
        const int q = warp & 3;
        if (blk_hi > blk_lo) {
            mbar_wait(&acc_full, 0);
            tc_fence_after();
        }
        float* out = part + (long long)blockIdx.x * COUT * K;
#pragma unroll 1
        for (int mt = 0; mt < MT; ++mt) {
            uint32_t r[32];
            if (blk_hi > blk_lo) {
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + mt * COUT, r);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int co = 0; co < COUT; ++co) r[co] = 0u;
            }
            const int k = mt * 128 + q * 32 + lane;
            if (k < K) {
#pragma unroll
                for (int co = 0; co < COUT; ++co)
                    out[(long long)co * K + k] = __uint_as_float(r[co]) * in_scale;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 128);
    (void)KPM;
}

template <class T, int CIN, int KH, int HIN, int HOUT, int STRIDE, int STAGES>
static int launch_tc_wgrad(const char* name, const T* in, FrameMap fm, const float* dz, float* part,
                           int R, float in_scale, int* grid_out, cudaStream_t st,
                           float* bias_part = nullptr) {
    constexpr int K = CIN * KH * KH, NKB = (K + 63) / 64, MT = (NKB + 1) / 2;
    const int smem = STAGES * (2 * MT * 8192 + 32 * 128) + 1024;
    auto kern = tc_conv_wgrad_kernel<T, CIN, KH, HIN, HOUT, STRIDE, STAGES>;
    R2D2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const long long Mtot = (long long)R * HOUT * HOUT;
    const int nblk = (int)((Mtot + 63) / 64);
    int grid = sm_count();
    if (grid > nblk) grid = nblk;
    *grid_out = grid;
    R2D2_LAUNCH(name, st, kern<<<grid, 416, smem, st>>>(in, fm, dz, part, R, in_scale, bias_part));
    return R2D2_OK;
}

// dw[i] = sum_z part[z*nw + i]  and  db[c] = sum_z bias_part[z*32 + c]  in one launch.  Fixed order:
// four interleaved chains (z = j, j + 4, ...) per output, each summed in order, then
// (c0 + c1) + (c2 + c3).  64 outputs x 4 chains per block: 148 partials of a 16K-element gradient
// are 256 blocks of short loops instead of 64 blocks of long ones.
__global__ void __launch_bounds__(256)
reduce_wgrad_partials_kernel(const float* __restrict__ part, float* __restrict__ dw, int nw,
                             const float* __restrict__ bias_part, float* __restrict__ db,
                             int splits) {
    __shared__ float sh[4][64];
    const int o = threadIdx.x & 63, chain = threadIdx.x >> 6;
    const int i = blockIdx.x * 64 + o;
    float s = 0.f;
    if (i < nw) {
        for (int z = chain; z < splits; z += 4) s += part[(long long)z * nw + i];
    } else if (i < nw + 32) {
        for (int z = chain; z < splits; z += 4) s += bias_part[z * 32 + (i - nw)];
    }
    sh[chain][o] = s;
    __syncthreads();
    if (chain == 0) {
        const float t = (sh[0][o] + sh[1][o]) + (sh[2][o] + sh[3][o]);
        if (i < nw) dw[i] = t;
        else if (i < nw + 32) db[i - nw] = t;
    }
}
static int reduce_wgrad(const float* part, float* dw, int nw, float* db, int grid, cudaStream_t st) {
    // bias partials live behind the weight partials of all CTAs
    R2D2_LAUNCH("reduce_partials_kernel", st, reduce_wgrad_partials_kernel<<<(nw + 32 + 63) / 64, 256, 0, st>>>(
        part, dw, nw, part + (long long)grid * nw, db, grid));
    return R2D2_OK;
}

// ------------------------------------------------------------------------------------------
// Conv data gradient on tensor cores (plain bf16), ReLU' of the layer input fused:
//   dIn[r,ci,y,x] = relu'(act[r,ci,y,x]) * sum_{co,ky,kx} dZ[r,co,(y-ky)/S,(x-kx)/S] W[co,ci,ky,kx]
// Output pixels are split by parity class (PY,PX) = (y%S, x%S): within a class only the taps
// ky = PY + S*jy, kx = PX + S*jx contribute, so each class is a dense implicit GEMM with
// M = (frame, yy, xx), N = 32 input channels, K = 32 * (KH/S)^2 -- no zero-insertion.
// Same pipeline as the forward kernel (producer warps -> swizzled smem -> UMMA -> TMEM x2).
template <int KH, int HIN, int HOUT, int STRIDE, int PY, int PX>
__global__ void __launch_bounds__(416, 1)
tc_conv_dgrad_kernel(const float* __restrict__ dz, const float* __restrict__ w,
                     const float* __restrict__ act, float* __restrict__ din, const int R) {
    constexpr int TAPS = KH / STRIDE, C = 32;
    constexpr int K = C * TAPS * TAPS, NKB = (K + 63) / 64, KP = NKB * 64;
    constexpr int GY = (HIN - PY + STRIDE - 1) / STRIDE, GX = (HIN - PX + STRIDE - 1) / STRIDE;
    constexpr int P = GY * GX;                                   // output pixels of this class per frame
    constexpr uint32_t A_TILE = 128 * 128, W_KB = C * 128;
    constexpr int MMA_WARP = 8;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[CV_STAGES], empty_bar[CV_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t* w_s = base;                                         // [NKB][32 x 128 B]
    uint8_t* a_s = base + NKB * W_KB;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long Mtot = (long long)R * P;
    const int ntiles = (int)((Mtot + 127) / 128);

    if (tid == 0) {
        for (int s = 0; s < CV_STAGES; ++s) {
            mbar_init(&full_bar[s], 256);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
        }
        mbar_fence_init();
    }
    // resident B: row = ci, k = (co, jy, jx) -> W[co, ci, PY + S*jy, PX + S*jx]
    for (int i = tid; i < C * (KP / 8); i += blockDim.x) {
        const int ci = i / (KP / 8), ch = i - ci * (KP / 8);
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int k = ch * 8 + e;
            float x = 0.f;
            if (k < K) {
                const int co = k / (TAPS * TAPS), rem = k - co * (TAPS * TAPS);
                const int jy = rem / TAPS, jx = rem - jy * TAPS;
                x = __ldg(w + ((long long)(co * C + ci) * KH + PY + STRIDE * jy) * KH + PX + STRIDE * jx);
            }
            v[e] = x;
        }
        *reinterpret_cast<uint4*>(w_s + (uint32_t)(ch >> 3) * W_KB + sw128_off(ci, ch & 7)) =
            make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                       pack_bf16x2(v[6], v[7]));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        const int row = tid & 127, half = tid >> 7;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const long long m = (long long)tile * 128 + row;
            const bool valid = m < Mtot;
            int yy = 0, xx = 0;
            const float* src = dz;
            if (valid) {
                const int r = (int)(m / P), p = (int)(m - (long long)r * P);
                yy = p / GX;
                xx = p - yy * GX;
                src = dz + (long long)r * C * HOUT * HOUT;
            }
#pragma unroll
            for (int kb = 0; kb < NKB; ++kb, ++it) {
                const int s = it % CV_STAGES;
                mbar_wait(&empty_bar[s], ((it / CV_STAGES) & 1) ^ 1);
                uint8_t* a_t = a_s + (size_t)s * A_TILE;
                float v[4][8];
#pragma unroll
                for (int cc = 0; cc < 4; ++cc)
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        // `half` only shifts k by 32: keep the decode compile-time per branch
                        const int kbase = kb * 64 + cc * 8 + e;
                        float x = 0.f;
#pragma unroll
                        for (int hsel = 0; hsel < 2; ++hsel) {
                            if (half == hsel) {
                                const int k = kbase + hsel * 32;
                                const int co = k / (TAPS * TAPS), rem = k - co * (TAPS * TAPS);
                                const int jy = rem / TAPS, jx = rem - jy * TAPS;
                                const int oy = yy - jy, ox = xx - jx;
                                if (k < K && valid && oy >= 0 && oy < HOUT && ox >= 0 && ox < HOUT)
                                    x = __ldg(src + (co * HOUT + oy) * HOUT + ox);
                            }
                        }
                        v[cc][e] = x;
                    }
#pragma unroll
                for (int cc = 0; cc < 4; ++cc)
                    *reinterpret_cast<uint4*>(a_t + sw128_off(row, half * 4 + cc)) =
                        make_uint4(pack_bf16x2(v[cc][0], v[cc][1]), pack_bf16x2(v[cc][2], v[cc][3]),
                                   pack_bf16x2(v[cc][4], v[cc][5]), pack_bf16x2(v[cc][6], v[cc][7]));
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(&full_bar[s]);
            }
        }
    } else if (warp == MMA_WARP) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, C);
            const uint32_t wb = smem_u32(w_s);
            uint32_t it = 0, tcount = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tcount) {
                const uint32_t buf = tcount & 1, use = tcount >> 1;
                mbar_wait(&acc_empty[buf], (use & 1) ^ 1);
                tc_fence_after();
                const uint32_t d = tmem_base + buf * C;
                for (int kb = 0; kb < NKB; ++kb, ++it) {
                    const int s = it % CV_STAGES;
                    mbar_wait(&full_bar[s], (it / CV_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t aa = smem_u32(a_s + (size_t)s * A_TILE), bb = wb + kb * W_KB;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(d, umma_desc_k_sw128(aa + k * 32), umma_desc_k_sw128(bb + k * 32), idesc,
                                 (kb | k) != 0);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        const int q = warp & 3;
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tcount) {
            const uint32_t buf = tcount & 1, use = tcount >> 1;
            mbar_wait(&acc_full[buf], use & 1);
            tc_fence_after();
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * C, r);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&acc_empty[buf]);
            const long long m = (long long)tile * 128 + q * 32 + lane;
            if (m < Mtot) {
                const int fr = (int)(m / P), p = (int)(m - (long long)fr * P);
                const int yy = p / GX, xx = p - yy * GX;
                const long long o = (long long)fr * C * HIN * HIN + (STRIDE * yy + PY) * HIN + STRIDE * xx + PX;
#pragma unroll
                for (int ci = 0; ci < C; ++ci) {
                    const long long oo = o + (long long)ci * HIN * HIN;
                    din[oo] = act[oo] > 0.f ? __uint_as_float(r[ci]) : 0.f;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 64);
}

// Stride-2 layer (conv2: 4x4 kernel, 20x20 <- 9x9), all four output-parity classes in ONE GEMM:
// the A tile (dZ gathered at (yy-jy, xx-jx), k = (co, jy, jx), K = 128) does not depend on the
// class, only the weights do, so the classes are concatenated along N: D[(r,yy,xx), (py,px,ci)],
// N = 128.  One pass over dZ and over the layer's activations (ReLU'), 2x2 output blocks written
// as float2 -- the per-class version re-read the activations four times (ncu: 226 MB per class).
// MASKBITS: ReLU' comes from a bit mask (one u32 per frame and pixel, bit ci = activation > 0,
// written by the first layer's forward epilogue) instead of the fp32 activations themselves:
// 16 bytes per thread instead of 512, and the kernel is HBM-bound.
template <bool MASKBITS>
__global__ void __launch_bounds__(416, 1)
tc_conv_dgrad2_kernel(const float* __restrict__ dz, const float* __restrict__ w,
                      const float* __restrict__ act, const uint32_t* __restrict__ maskbits,
                      float* __restrict__ din, const int R) {
    constexpr int KH = 4, HIN = 20, HOUT = 9, TAPS = 2, C = 32, NCOL = 128;
    constexpr int K = C * TAPS * TAPS, NKB = K / 64;             // 128, 2
    constexpr int G = 10, P = G * G;                              // (yy, xx) grid per frame
    constexpr uint32_t A_TILE = 128 * 128, W_KB = NCOL * 128;     // B k-block: 128 rows x 128 B
    constexpr int MMA_WARP = 8;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[CV_STAGES], empty_bar[CV_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t* w_s = base;
    uint8_t* a_s = base + NKB * W_KB;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const long long Mtot = (long long)R * P;
    const int ntiles = (int)((Mtot + 127) / 128);

    if (tid == 0) {
        for (int s = 0; s < CV_STAGES; ++s) {
            mbar_init(&full_bar[s], 256);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
        }
        mbar_fence_init();
    }
    // resident B: row n = (py*2+px)*32 + ci, k = (co, jy, jx) -> W[co, ci, py + 2 jy, px + 2 jx]
    for (int i = tid; i < NCOL * (K / 8); i += blockDim.x) {
        const int n = i / (K / 8), ch = i - n * (K / 8);
        const int cls = n >> 5, ci = n & 31, py = cls >> 1, px = cls & 1;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int k = ch * 8 + e;
            const int co = k >> 2, jy = (k >> 1) & 1, jx = k & 1;
            v[e] = __ldg(w + ((long long)(co * C + ci) * KH + py + 2 * jy) * KH + px + 2 * jx);
        }
        *reinterpret_cast<uint4*>(w_s + (uint32_t)(ch >> 3) * W_KB + sw128_off(n, ch & 7)) =
            make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                       pack_bf16x2(v[6], v[7]));
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        const int row = tid & 127, half = tid >> 7;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const long long m = (long long)tile * 128 + row;
            const bool valid = m < Mtot;
            int yy = 0, xx = 0;
            const float* src = dz;
            if (valid) {
                const int r = (int)(m / P), p = (int)(m - (long long)r * P);
                yy = p / G;
                xx = p - yy * G;
                src = dz + (long long)r * C * HOUT * HOUT;
            }
            // the four taps of this pixel: validity and offsets once per row
            int off[4];
            bool ok[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int oy = yy - (j >> 1), ox = xx - (j & 1);
                ok[j] = valid && oy >= 0 && oy < HOUT && ox >= 0 && ox < HOUT;
                off[j] = oy * HOUT + ox;
            }
#pragma unroll
            for (int kb = 0; kb < NKB; ++kb, ++it) {
                const int s = it % CV_STAGES;
                mbar_wait(&empty_bar[s], ((it / CV_STAGES) & 1) ^ 1);
                uint8_t* a_t = a_s + (size_t)s * A_TILE;
                float v[4][8];
#pragma unroll
                for (int cc = 0; cc < 4; ++cc)
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const int co = kb * 16 + half * 8 + cc * 2 + (e >> 2);   // k = co*4 + tap
                        v[cc][e] = ok[e & 3] ? __ldg(src + co * (HOUT * HOUT) + off[e & 3]) : 0.f;
                    }
#pragma unroll
                for (int cc = 0; cc < 4; ++cc)
                    *reinterpret_cast<uint4*>(a_t + sw128_off(row, half * 4 + cc)) =
                        make_uint4(pack_bf16x2(v[cc][0], v[cc][1]), pack_bf16x2(v[cc][2], v[cc][3]),
                                   pack_bf16x2(v[cc][4], v[cc][5]), pack_bf16x2(v[cc][6], v[cc][7]));
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(&full_bar[s]);
            }
        }
    } else if (warp == MMA_WARP) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, NCOL);
            const uint32_t wb = smem_u32(w_s);
            uint32_t it = 0, tcount = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tcount) {
                const uint32_t buf = tcount & 1, use = tcount >> 1;
                mbar_wait(&acc_empty[buf], (use & 1) ^ 1);
                tc_fence_after();
                const uint32_t d = tmem_base + buf * NCOL;
                for (int kb = 0; kb < NKB; ++kb, ++it) {
                    const int s = it % CV_STAGES;
                    mbar_wait(&full_bar[s], (it / CV_STAGES) & 1);
                    tc_fence_after();
                    const uint32_t aa = smem_u32(a_s + (size_t)s * A_TILE), bb = wb + kb * W_KB;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(d, umma_desc_k_sw128(aa + k * 32), umma_desc_k_sw128(bb + k * 32), idesc,
                                 (kb | k) != 0);
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        const int q = warp & 3;
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++tcount) {
            const uint32_t buf = tcount & 1, use = tcount >> 1;
            mbar_wait(&acc_full[buf], use & 1);
            tc_fence_after();
            const long long m = (long long)tile * 128 + q * 32 + lane;
            const bool valid = m < Mtot;
            long long o = 0, mo = 0;
            if (valid) {
                const int fr = (int)(m / P), p = (int)(m - (long long)fr * P);
                const int yy = p / G, xx = p - yy * G;
                o = (long long)fr * C * HIN * HIN + (2 * yy) * HIN + 2 * xx;
                mo = (long long)fr * HIN * HIN + (2 * yy) * HIN + 2 * xx;
            }
            (void)mo;
#pragma unroll
            for (int py = 0; py < 2; ++py) {
                uint32_t r0[32], r1[32];                       // px = 0, 1
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * NCOL + (py * 2) * C, r0);
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * NCOL + (py * 2 + 1) * C, r1);
                tmem_ld_wait();
                if (py == 1) {
                    tc_fence_before();
                    mbar_arrive(&acc_empty[buf]);
                }
                if (valid) {
                    if constexpr (MASKBITS) {
                        // the two pixels (2yy+py, 2xx) and (2yy+py, 2xx+1) of this frame
                        const uint2 mk = __ldg(reinterpret_cast<const uint2*>(maskbits + mo + py * HIN));
#pragma unroll
                        for (int ci = 0; ci < C; ++ci) {
                            const long long oo = o + (long long)ci * HIN * HIN + py * HIN;
                            *reinterpret_cast<float2*>(din + oo) =
                                make_float2((mk.x >> ci) & 1u ? __uint_as_float(r0[ci]) : 0.f,
                                            (mk.y >> ci) & 1u ? __uint_as_float(r1[ci]) : 0.f);
                        }
                    } else {
#pragma unroll
                        for (int ci = 0; ci < C; ++ci) {
                            const long long oo = o + (long long)ci * HIN * HIN + py * HIN;
                            const float2 a = *reinterpret_cast<const float2*>(act + oo);
                            *reinterpret_cast<float2*>(din + oo) =
                                make_float2(a.x > 0.f ? __uint_as_float(r0[ci]) : 0.f,
                                            a.y > 0.f ? __uint_as_float(r1[ci]) : 0.f);
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 256);
}

template <int KH, int HIN, int HOUT, int STRIDE, int PY, int PX>
static int launch_tc_dgrad(const char* name, const float* dz, const float* w, const float* act,
                           float* din, int R, cudaStream_t st) {
    constexpr int TAPS = KH / STRIDE, K = 32 * TAPS * TAPS, NKB = (K + 63) / 64;
    constexpr int GY = (HIN - PY + STRIDE - 1) / STRIDE, GX = (HIN - PX + STRIDE - 1) / STRIDE;
    const int smem = NKB * 32 * 128 + CV_STAGES * 128 * 128 + 1024;
    auto kern = tc_conv_dgrad_kernel<KH, HIN, HOUT, STRIDE, PY, PX>;
    R2D2_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const long long Mtot = (long long)R * GY * GX;
    const int ntiles = (int)((Mtot + 127) / 128);
    int grid = sm_count();
    if (grid > ntiles) grid = ntiles;
    R2D2_LAUNCH(name, st, kern<<<grid, 416, smem, st>>>(dz, w, act, din, R));
    return R2D2_OK;
}

template <class T>
static int tc_torso_fwd_impl(const T* frames, FrameMap fm, int R, const float* w1, const float* b1,
                             const float* w2, const float* b2, const float* w3, const float* b3,
                             float* a1, float* a2, float* feat, cudaStream_t st) {
    R2D2_REQUIRE((long long)R * 400 < (1ll << 31), "tc_torso_fwd: too many frames");
    const float s1 = sizeof(T) == 1 ? 1.f / 255.f : 1.f;
    int rc = launch_tc_conv<T, 4, 8, 84, 20, 4>("tc_conv1_fwd", frames, fm, w1, b1, a1, R, s1, st);
    if (rc) return rc;
    rc = launch_tc_conv<float, 32, 4, 20, 9, 2>("tc_conv2_fwd", a1, FrameMap{0, 0, 0, 12800}, w2, b2, a2,
                                                R, 1.f, st);
    if (rc) return rc;
    return launch_tc_conv<float, 32, 3, 9, 7, 1>("tc_conv3_fwd", a2, FrameMap{0, 0, 0, 2592}, w3, b3, feat,
                                                 R, 1.f, st);
}

}  // namespace r2d2

using namespace r2d2;

extern "C" {

int r2d2_tc_torso_fwd_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* w1,
                         const float* b1, const float* w2, const float* b2, const float* w3,
                         const float* b3, float* a1, float* a2, float* feat, void* stream) {
    R2D2_REQUIRE(slab && w1 && b1 && w2 && b2 && w3 && b3 && a1 && a2 && feat, "tc_torso_fwd: null pointer");
    R2D2_REQUIRE(B > 0 && nt > 0 && s0 >= 0 && s0 + nt <= S, "tc_torso_fwd: bad frame window");
    return tc_torso_fwd_impl<uint8_t>(slab, FrameMap{B, S, s0, R2D2_FRAME_BYTES}, B * nt, w1, b1, w2, b2,
                                      w3, b3, a1, a2, feat, as_stream(stream));
}

int r2d2_tc_conv1_fwd_dual_u8(const uint8_t* slab, int B, int S, int nt, const float* w1_a,
                              const float* b1_a, float* a1_a, const float* w1_b, const float* b1_b,
                              float* a1_b, int s0_b, void* stream) {
    R2D2_REQUIRE(slab && w1_a && b1_a && a1_a && w1_b && b1_b && a1_b, "tc_conv1_fwd_dual: null pointer");
    R2D2_REQUIRE(B > 0 && nt > 0 && nt <= S && s0_b >= 0 && s0_b < nt, "tc_conv1_fwd_dual: bad frame window");
    R2D2_REQUIRE((long long)B * nt * 400 < (1ll << 31), "tc_conv1_fwd_dual: too many frames");
    return launch_tc_conv1_dual(slab, FrameMap{B, S, 0, R2D2_FRAME_BYTES}, w1_a, b1_a, a1_a, w1_b, b1_b, a1_b,
                                B * nt, s0_b * B, as_stream(stream));
}

int r2d2_tc_conv1_fwd_dual_ring(const uint8_t* ring_state, int64_t M, const int64_t* seq_index, int B,
                                int nt, const float* w1_a, const float* b1_a, float* a1_a,
                                const float* w1_b, const float* b1_b, float* a1_b, int s0_b,
                                uint32_t* relu_bits_a, void* stream) {
    R2D2_REQUIRE(ring_state && seq_index && w1_a && b1_a && a1_a && w1_b && b1_b && a1_b,
                 "tc_conv1_fwd_dual_ring: null pointer");
    R2D2_REQUIRE(B > 0 && nt > 0 && M > 0 && s0_b >= 0 && s0_b < nt, "tc_conv1_fwd_dual_ring: bad frame window");
    R2D2_REQUIRE((long long)B * nt * 400 < (1ll << 31), "tc_conv1_fwd_dual_ring: too many frames");
    FrameMap fm{B, nt, 0, R2D2_FRAME_BYTES, reinterpret_cast<const long long*>(seq_index), (long long)M};
    return launch_tc_conv1_dual(ring_state, fm, w1_a, b1_a, a1_a, w1_b, b1_b, a1_b, B * nt, s0_b * B,
                                as_stream(stream), relu_bits_a);
}

// grid the weight-gradient kernels run with (one CTA per SM, at most one per 64-pixel block)
static int wgrad_grid(int R, int P) {
    const long long nblk = ((long long)R * P + 63) / 64;
    const int g = sm_count();
    return nblk < g ? (int)nblk : g;
}

static bool wgrad_gather_form() {
    static const bool v = [] {
        const char* e = getenv("R2D2_WGRAD");
        return e && e[0] == 'g';
    }();
    return v;
}
static int launch_conv3_wgrad(const float* a2, const float* dz, float* part, int R, int* grid_out, cudaStream_t st,
                              float* bias_part = nullptr) {
    if (wgrad_gather_form())
        return launch_tc_wgrad<float, 32, 3, 9, 7, 1, 3>("tc_conv3_wgrad", a2, FrameMap{0, 0, 0, 2592}, dz, part, R, 1.f,
                                                         grid_out, st, bias_part);
    *grid_out = wgrad_grid(R, 49);
    return launch_tc_conv3_wgrad_sw(a2, dz, part, R, *grid_out, st, bias_part);
}
static int launch_conv2_wgrad(const float* a1, const float* dz, float* part, int R, int* grid_out, cudaStream_t st,
                              float* bias_part = nullptr) {
    if (wgrad_gather_form())
        return launch_tc_wgrad<float, 32, 4, 20, 9, 2, 3>("tc_conv2_wgrad", a1, FrameMap{0, 0, 0, 12800}, dz, part, R, 1.f,
                                                          grid_out, st, bias_part);
    *grid_out = wgrad_grid(R, 81);
    return launch_tc_conv2_wgrad_sw(a1, dz, part, R, *grid_out, st, bias_part);
}
// first-layer weight gradient: shifted-window kernel (tc_wgrad_sw.cu) unless R2D2_WGRAD=gather
static int launch_conv1_wgrad(const uint8_t* in, FrameMap fm, const float* dz, float* part, int R, int* grid_out,
                              cudaStream_t st, float* bias_part = nullptr) {
    if (wgrad_gather_form())
        return launch_tc_wgrad<uint8_t, 4, 8, 84, 20, 4, 4>("tc_conv1_wgrad", in, fm, dz, part, R, 1.f / 255.f, grid_out,
                                                            st, bias_part);
    *grid_out = wgrad_grid(R, 400);
    return launch_tc_conv1_wgrad_sw(in, fm, dz, part, R, 1.f / 255.f, *grid_out, st, bias_part);
}

int r2d2_tc_conv1_wgrad_ring(const uint8_t* ring_state, int64_t M, const int64_t* seq_index, int B, int s0,
                             int nt, const float* dz, float* dw, float* db, float* part, void* stream) {
    R2D2_REQUIRE(ring_state && seq_index && dz && dw && part && nt > 0 && B > 0 && M > 0 && s0 >= 0,
                 "tc_conv1_wgrad_ring: bad args");
    int grid = wgrad_grid(B * nt, 400);
    FrameMap fm{B, s0 + nt, s0, R2D2_FRAME_BYTES, reinterpret_cast<const long long*>(seq_index), (long long)M};
    int rc = launch_conv1_wgrad(ring_state, fm, dz, part, B * nt, &grid, as_stream(stream),
                                db ? part + (long long)grid * 32 * 256 : nullptr);
    if (rc) return rc;
    if (db) return reduce_wgrad(part, dw, 32 * 256, db, grid, as_stream(stream));
    return r2d2_reduce_partials(part, dw, 32 * 256, grid, stream);
}

int r2d2_tc_torso_fwd_tail(const float* a1, int R, const float* w2, const float* b2, const float* w3,
                           const float* b3, float* a2, float* feat, void* feat_hi, void* feat_lo,
                           void* stream) {
    R2D2_REQUIRE(a1 && w2 && b2 && w3 && b3 && a2 && feat && R > 0, "tc_torso_fwd_tail: bad args");
    R2D2_REQUIRE((feat_hi == nullptr) == (feat_lo == nullptr), "tc_torso_fwd_tail: feat_hi and feat_lo go together");
    // shifted-window kernels (tc_conv_sw.cu): every activation staged once instead of once per
    // kernel offset (per launch at 5440 frames: conv2 0.118 vs 0.149 ms, conv3 0.057 vs 0.065 ms).
    // R2D2_CONV_TAIL=gather selects the gather kernels above (A/B measurements).
    static const bool gather_form = [] {
        const char* e = getenv("R2D2_CONV_TAIL");
        return e && e[0] == 'g';
    }();
    if (!gather_form) {
        int rc = launch_tc_conv_sw<2>("tc_conv2_fwd", a1, w2, b2, a2, R, as_stream(stream), nullptr, nullptr);
        if (rc) return rc;
        return launch_tc_conv_sw<3>("tc_conv3_fwd", a2, w3, b3, feat, R, as_stream(stream),
                                    static_cast<__nv_bfloat16*>(feat_hi), static_cast<__nv_bfloat16*>(feat_lo));
    }
    int rc = launch_tc_conv<float, 32, 4, 20, 9, 2>("tc_conv2_fwd", a1, FrameMap{0, 0, 0, 12800}, w2, b2, a2,
                                                    R, 1.f, as_stream(stream));
    if (rc) return rc;
    return launch_tc_conv<float, 32, 3, 9, 7, 1>("tc_conv3_fwd", a2, FrameMap{0, 0, 0, 2592}, w3, b3, feat,
                                                 R, 1.f, as_stream(stream),
                                                 static_cast<__nv_bfloat16*>(feat_hi),
                                                 static_cast<__nv_bfloat16*>(feat_lo));
}

// data gradients of conv3 (into the 9x9 layer) and conv2 (into the 20x20 layer), ReLU' fused
int r2d2_tc_conv_dgrad3(const float* dz3, const float* w3, const float* a2, float* dz2, int R,
                        void* stream) {
    R2D2_REQUIRE(dz3 && w3 && a2 && dz2 && R > 0, "tc_conv_dgrad3: bad args");
    static const bool gather_form = [] {                 // R2D2_DGRAD=gather: the previous kernel (A/B runs)
        const char* e = getenv("R2D2_DGRAD");
        return e && e[0] == 'g';
    }();
    if (!gather_form) return launch_tc_conv3_dgrad_sw(dz3, w3, a2, dz2, R, as_stream(stream));
    return launch_tc_dgrad<3, 9, 7, 1, 0, 0>("tc_conv3_dgrad", dz3, w3, a2, dz2, R, as_stream(stream));
}
static int dgrad2_launch(const float* dz2, const float* w2, const float* a1, const uint32_t* maskbits,
                         float* dz1, int R, cudaStream_t st) {
    const int smem = 2 * 128 * 128 + CV_STAGES * 128 * 128 + 1024;
    const int ntiles = (int)(((long long)R * 100 + 127) / 128);
    int grid = sm_count();
    if (grid > ntiles) grid = ntiles;
    static const bool gather_form = [] {                 // R2D2_DGRAD=gather: the previous kernels (A/B runs)
        const char* e = getenv("R2D2_DGRAD");
        return e && e[0] == 'g';
    }();
    if (maskbits && !gather_form) return launch_tc_conv2_dgrad_sw(dz2, w2, maskbits, dz1, R, st);
    if (maskbits) {
        R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv_dgrad2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        R2D2_LAUNCH("tc_conv2_dgrad", st, tc_conv_dgrad2_kernel<true><<<grid, 416, smem, st>>>(dz2, w2, nullptr, maskbits, dz1, R));
    } else {
        R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv_dgrad2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        R2D2_LAUNCH("tc_conv2_dgrad", st, tc_conv_dgrad2_kernel<false><<<grid, 416, smem, st>>>(dz2, w2, a1, nullptr, dz1, R));
    }
    return R2D2_OK;
}
int r2d2_tc_conv_dgrad2(const float* dz2, const float* w2, const float* a1, float* dz1, int R,
                        void* stream) {
    R2D2_REQUIRE(dz2 && w2 && a1 && dz1 && R > 0, "tc_conv_dgrad2: bad args");
    return dgrad2_launch(dz2, w2, a1, nullptr, dz1, R, as_stream(stream));
}
int r2d2_tc_conv_dgrad2_masked(const float* dz2, const float* w2, const uint32_t* relu_bits, float* dz1,
                               int R, void* stream) {
    R2D2_REQUIRE(dz2 && w2 && relu_bits && dz1 && R > 0 && ((uintptr_t)relu_bits % 8) == 0,
                 "tc_conv_dgrad2_masked: bad args");
    return dgrad2_launch(dz2, w2, nullptr, relu_bits, dz1, R, as_stream(stream));
}

// partial-sum reduction shared with the fp32 path (nn_kernels.cu)
int r2d2_reduce_partials(const float* part, float* out, int64_t n, int splits, void* stream);

// weight gradients of the three conv layers on tensor cores; dz1/dz2/dz3 are the NCHW
// pre-activation gradients produced by the data-gradient kernels (dz3 = masked dfeat)
int r2d2_tc_torso_wgrad_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* a1,
                           const float* a2, const float* dz1, const float* dz2, const float* dz3,
                           float* dw1, float* dw2, float* dw3, void* scratch, void* stream) {
    R2D2_REQUIRE(slab && a1 && a2 && dz1 && dz2 && dz3 && dw1 && dw2 && dw3 && scratch,
                 "tc_torso_wgrad: null pointer");
    R2D2_REQUIRE(B > 0 && nt > 0 && s0 >= 0 && s0 + nt <= S, "tc_torso_wgrad: bad frame window");
    cudaStream_t st = as_stream(stream);
    const int R = B * nt;
    float* part = static_cast<float*>(scratch);
    int grid = 0;
    int rc = launch_conv3_wgrad(a2, dz3, part, R, &grid, st);
    if (rc) return rc;
    rc = r2d2_reduce_partials(part, dw3, 32 * 288, grid, stream);
    if (rc) return rc;
    rc = launch_conv2_wgrad(a1, dz2, part, R, &grid, st);
    if (rc) return rc;
    rc = r2d2_reduce_partials(part, dw2, 32 * 512, grid, stream);
    if (rc) return rc;
    rc = launch_conv1_wgrad(slab, FrameMap{B, S, s0, R2D2_FRAME_BYTES}, dz1, part, R, &grid, st);
    if (rc) return rc;
    return r2d2_reduce_partials(part, dw1, 32 * 256, grid, stream);
}

// one layer at a time (layer = 1, 2, 3), with its own partial buffer, so that the three weight
// gradients can run on different streams: part needs sm_count * 32 * K floats (K = 256/512/288)
int r2d2_tc_conv_wgrad_layer(int layer, const void* in, int B, int S, int s0, int nt,
                             const float* dz, float* dw, float* db, float* part, void* stream) {
    R2D2_REQUIRE(in && dz && dw && part && nt > 0 && B > 0, "tc_conv_wgrad_layer: bad args");
    cudaStream_t st = as_stream(stream);
    const int R = B * nt;
    int grid = 0, rc = 0, K = 0;
    const int pix = layer == 1 ? 400 : layer == 2 ? 81 : 49, kk = layer == 1 ? 256 : layer == 2 ? 512 : 288;
    float* bias_part = db ? part + (long long)wgrad_grid(R, pix) * 32 * kk : nullptr;
    if (layer == 1) {
        R2D2_REQUIRE(s0 >= 0 && s0 + nt <= S, "tc_conv_wgrad_layer: bad frame window");
        K = 256;
        rc = launch_conv1_wgrad(static_cast<const uint8_t*>(in), FrameMap{B, S, s0, R2D2_FRAME_BYTES}, dz, part, R,
                                &grid, st, bias_part);
    } else if (layer == 2) {
        K = 512;
        rc = launch_conv2_wgrad(static_cast<const float*>(in), dz, part, R, &grid, st, bias_part);
    } else if (layer == 3) {
        K = 288;
        rc = launch_conv3_wgrad(static_cast<const float*>(in), dz, part, R, &grid, st, bias_part);
    } else {
        return set_err(R2D2_EINVAL, "tc_conv_wgrad_layer: layer must be 1, 2 or 3");
    }
    if (rc) return rc;
    if (db) return reduce_wgrad(part, dw, 32 * K, db, grid, st);
    return r2d2_reduce_partials(part, dw, 32 * K, grid, stream);
}

int r2d2_tc_torso_fwd_f32(const float* frames, int R, const float* w1, const float* b1,
                          const float* w2, const float* b2, const float* w3, const float* b3,
                          float* a1, float* a2, float* feat, void* stream) {
    R2D2_REQUIRE(frames && w1 && b1 && w2 && b2 && w3 && b3 && a1 && a2 && feat, "tc_torso_fwd: null pointer");
    R2D2_REQUIRE(R > 0, "tc_torso_fwd: bad frame count");
    return tc_torso_fwd_impl<float>(frames, FrameMap{0, 0, 0, R2D2_FRAME_BYTES}, R, w1, b1, w2, b2, w3, b3,
                                    a1, a2, feat, as_stream(stream));
}

}  // extern "C"
