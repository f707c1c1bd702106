// Conv weight gradients in "shifted window" form (plain bf16; see tc_conv.cu for the gather form).
//
//   dW[co][ci][ky][kx] = sum over frames and output positions of dZ[co][oy][ox] * X[ci][S oy + ky][S ox + kx]
//
// With the input in space-to-depth layout (row = position, 128-byte row = 64 channels (ci, dy, dx),
// exactly the image the forward shifted-window kernels build) the sum over positions is a GEMM whose
// K dimension is the position:   D_shift[chan][co] = sum_m Img[m + shift][chan] * dZ'[co][m]
// -- Img is consumed as an MN-major A operand (M = channel), one kernel offset (ky2, kx2) is a ROW
// SHIFT of the same image, and two shifts that differ by one row are the two 64-element MN groups
// of one M = 128 tile (leading-dimension offset 128 bytes).  dZ' is dZ laid out on the image's
// position grid (the column the s2d grid has and the output grid lacks holds zeros), a K-major
// B operand of N = 32 channels.  Every input pixel and every gradient value is loaded and
// converted once per frame; the gather form builds 64 patch elements per output pixel.
// Each persistent CTA accumulates its frames in TMEM and writes one partial; partials are summed
// in fixed order by reduce_wgrad_partials_kernel (tc_conv.cu).
#include <cuda_bf16.h>

#include "common.cuh"
#include "gemm.cuh"   // FrameMap
#include "ptx.cuh"

namespace r2d2 {

namespace {

__device__ __forceinline__ uint32_t ws_off(int row, int chunk) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}
__device__ __forceinline__ uint32_t ws_pack(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
// four u8 pixels of w -> two packed bf16x2 (exact: 8 significant bits)
__device__ __forceinline__ void ws_u8x4(uint32_t w, uint32_t& lo, uint32_t& hi) {
    const float f0 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7650u)) - 8388608.f;
    const float f1 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7651u)) - 8388608.f;
    const float f2 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7652u)) - 8388608.f;
    const float f3 = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7653u)) - 8388608.f;
    // the upper halves of the fp32 patterns ARE the bf16 values
    lo = __byte_perm(__float_as_uint(f0), __float_as_uint(f1), 0x7632u);
    hi = __byte_perm(__float_as_uint(f2), __float_as_uint(f3), 0x7632u);
}
__device__ __forceinline__ void ws_bar(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------------------------------------
// first layer: 8x8 / stride 4 on u8 84x84x4 -> s2d by 4: 21x21 positions x 64 channels.  The
// position grid is laid out with a row pitch of 24 (three pad columns): every output row of dZ
// then starts on a 16-byte chunk of dZ' and is stored as three whole chunks (20 values + 4 zeros)
// instead of 21 scalar pairs at run-time offsets.  Shifts {0, 1, 24, 25}; dZ [32][20][20] -> dZ' [32][504 (+8)]
constexpr int W1_PROD = 320;                        // producers: 882 image items = 3 rounds; 320 = 32 co x 10 row pairs
constexpr int W1_THREADS = W1_PROD + 64 + 128;       // + two MMA-issuing warps (one thread issues an MMA every ~74 clk)
constexpr int W1_PITCH = 24, W1_ROWS = 512, W1_STAGES = 2, W1_NKB = 8, W1_KSTEPS = 32;
constexpr uint32_t W1_IMG = W1_ROWS * 128;          // 21 x 24 (+8) positions x 64 bf16
constexpr uint32_t W1_DZ = W1_NKB * 32 * 128;       // 8 k-blocks of [32 co x 64 positions]
constexpr int W1_SMEM = W1_STAGES * (W1_IMG + W1_DZ) + 1024;

__global__ void __launch_bounds__(W1_THREADS, 1)
tc_conv1_wgrad_sw_kernel(const uint8_t* __restrict__ in, const FrameMap fm, const float* __restrict__ dz,
                         float* __restrict__ part, const int R, const float in_scale,
                         float* __restrict__ bias_part) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[W1_STAGES], empty_bar[W1_STAGES], acc_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float bias_s[W1_PROD];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    // images first, gradient tiles behind them: a shifted tile's over-read (rows 512..536, always
    // multiplied by zero gradients) must still hold finite values
    uint8_t* a_s = base;
    uint8_t* g_s = base + W1_STAGES * W1_IMG;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int MMA_WARP = W1_PROD / 32;
    const bool has_work = (int)blockIdx.x < R;

    if (tid == 0) {
        for (int s = 0; s < W1_STAGES; ++s) {
            mbar_init(&full_bar[s], W1_PROD);
            mbar_init(&empty_bar[s], 2);
        }
        mbar_init(&acc_full, 2);
        mbar_fence_init();
    }
    for (uint32_t i = tid * 16; i < W1_STAGES * (W1_IMG + W1_DZ); i += blockDim.x * 16)
        *reinterpret_cast<uint4*>(base + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers
        // image item = (input channel, position): 4 raw rows x 4 bytes -> two 16-byte chunks
        constexpr int NIT = 6;                                           // 2 channel pairs x 3 rounds of 320
        int soff[NIT];
        uint32_t doff[NIT];
        bool act[NIT];
#pragma unroll
        for (int k = 0; k < NIT; ++k) {
            const int i = tid + (k % 3) * W1_PROD;
            act[k] = i < 2 * 441;
            const int c2 = i / 441, p = i - c2 * 441;
            const int Y = p / 21, X = p - Y * 21;
            const int ci = (k / 3) * 2 + c2;
            soff[k] = act[k] ? ci * 7056 + (4 * Y) * 84 + 4 * X : 0;
            doff[k] = ws_off(act[k] ? Y * W1_PITCH + X : 0, ci * 2);
        }
        // gradient item = (channel co, output rows 2 yp and 2 yp + 1): 40 consecutive floats ->
        // chunks 6 yp .. 6 yp + 5 of the channel's dZ' row (8 positions each; 20 values + 4 zeros per row)
        const int co = tid / 10, yp = tid - co * 10;
        // register prefetch in two halves: the next frame's pixels are in flight while this
        // frame's gradients are converted, and the other way round
        uint32_t nx[NIT][4];
        float4 ng[10];
        // fm.base() of a ring frame is itself a dependent load (the sequence's start row) plus a
        // 64-bit modulo: it is evaluated one frame ahead of the pixel loads that need it
        auto load_img = [&](long long fbase) {
            const uint8_t* fb = in + fbase;
#pragma unroll
            for (int k = 0; k < NIT; ++k) {
                const uint32_t* q = reinterpret_cast<const uint32_t*>(fb + soff[k]);
#pragma unroll
                for (int j = 0; j < 4; ++j) nx[k][j] = act[k] ? __ldg(q + 21 * j) : 0u;
            }
        };
        auto load_dz = [&](int f) {
            const float4* gp = reinterpret_cast<const float4*>(dz + ((long long)f * 32 + co) * 400 + 40 * yp);
#pragma unroll
            for (int j = 0; j < 10; ++j) ng[j] = __ldg(gp + j);
        };
        // the register prefetch is half a frame deep; an L2 prefetch two frames ahead turns its
        // loads into L2 hits (221 + 400 lines of 128 bytes per frame: two per thread)
        auto prefetch_l2 = [&](int f, long long fbase) {
            if (f < R) {
                const uint8_t* fb = in + fbase;
                const uint8_t* gb = reinterpret_cast<const uint8_t*>(dz + (long long)f * 12800);
                if (tid < 221) asm volatile("prefetch.global.L2 [%0];" ::"l"(fb + tid * 128));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(gb + tid * 128));
                if (tid < 80) asm volatile("prefetch.global.L2 [%0];" ::"l"(gb + (320 + tid) * 128));
            }
        };
        long long b_next = 0;
        if (has_work) {
            load_img(fm.base(blockIdx.x));
            load_dz(blockIdx.x);
            if ((int)(blockIdx.x + gridDim.x) < R) b_next = fm.base(blockIdx.x + gridDim.x);
            prefetch_l2(blockIdx.x + gridDim.x, b_next);
        }
        float bsum = 0.f;
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int s = it % W1_STAGES;
            const bool more = fr + (int)gridDim.x < R;
            const int f2 = fr + 2 * gridDim.x;
            const long long b_pref = f2 < R ? fm.base(f2) : 0;           // needed after this frame's conversions
            mbar_wait(&empty_bar[s], ((it / W1_STAGES) & 1) ^ 1);
            uint8_t* img = a_s + (size_t)s * W1_IMG;
            uint8_t* gt = g_s + (size_t)s * W1_DZ;
#pragma unroll
            for (int k = 0; k < NIT; ++k)
                if (act[k]) {
                    uint4 c0, c1;
                    ws_u8x4(nx[k][0], c0.x, c0.y);
                    ws_u8x4(nx[k][1], c0.z, c0.w);
                    ws_u8x4(nx[k][2], c1.x, c1.y);
                    ws_u8x4(nx[k][3], c1.z, c1.w);
                    *reinterpret_cast<uint4*>(img + doff[k]) = c0;
                    *reinterpret_cast<uint4*>(img + (doff[k] ^ 16u)) = c1;
                }
            if (more) load_img(b_next);
            float g[40];
#pragma unroll
            for (int j = 0; j < 10; ++j) g[4 * j] = ng[j].x, g[4 * j + 1] = ng[j].y, g[4 * j + 2] = ng[j].z, g[4 * j + 3] = ng[j].w;
            float bs = 0.f;
#pragma unroll
            for (int j = 0; j < 40; ++j) bs += g[j];
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                const int C = 6 * yp + c;                                // chunk of the channel's 512-position row
                const float* v = g + (c / 3) * 20 + (c % 3) * 8;
                uint4 w;
                w.x = ws_pack(v[0], v[1]);
                w.y = ws_pack(v[2], v[3]);
                w.z = c % 3 == 2 ? 0u : ws_pack(v[4], v[5]);
                w.w = c % 3 == 2 ? 0u : ws_pack(v[6], v[7]);
                *reinterpret_cast<uint4*>(gt + (C >> 3) * 4096 + ws_off(co, C & 7)) = w;
            }
            bsum += bs;
            if (more) load_dz(fr + gridDim.x);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
            prefetch_l2(f2, b_pref);
            b_next = b_pref;
        }
        if (bias_part) {                                                 // fixed order: thread, then row pairs
            bias_s[tid] = bsum;
            ws_bar(1, W1_PROD);
            if (tid < 32) {
                float t = 0.f;
#pragma unroll
                for (int j = 0; j < 10; ++j) t += bias_s[tid * 10 + j];
                bias_part[blockIdx.x * 32 + tid] = t;
            }
        }
    } else if (warp < MMA_WARP + 2) {
        if (lane == 0) {                                                 // issuer `pair`: shifts (24 pair, 24 pair + 1)
            const int pair = warp - MMA_WARP;
            constexpr uint32_t idesc = umma_idesc_bf16_amn(128, 32);
            uint32_t it = 0;
            for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
                const int s = it % W1_STAGES;
                mbar_wait(&full_bar[s], (it / W1_STAGES) & 1);
                tc_fence_after();
                const uint32_t img = smem_u32(a_s + (size_t)s * W1_IMG), gt = smem_u32(g_s + (size_t)s * W1_DZ);
#pragma unroll 4
                for (int ks = 0; ks < W1_KSTEPS; ++ks)                   // 16 positions per MMA
                    umma_f16(tmem_base + pair * 32, umma_desc_mn_sw128(img + (ks * 16 + pair * W1_PITCH) * 128, 128),
                             umma_desc_k_sw128(gt + (ks >> 2) * 4096 + (ks & 3) * 32), idesc, (it | ks) != 0);
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_full);
        }
    } else {
        // ================= epilogue (once): TMEM lane = (shift kx2, channel), 32 columns = co
        const int q = warp & 3;
        if (has_work) {
            mbar_wait(&acc_full, 0);
            tc_fence_after();
        }
        float* out = part + (long long)blockIdx.x * 32 * 256;
        const int l = q * 32 + lane, kx2 = l >> 6, chan = l & 63;
        const int ci = chan >> 4, dy = (chan >> 2) & 3, dx = chan & 3;
#pragma unroll 1
        for (int ky2 = 0; ky2 < 2; ++ky2) {
            uint32_t r[32];
            if (has_work) {
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + ky2 * 32, r);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int c = 0; c < 32; ++c) r[c] = 0u;
            }
            const int k = ci * 64 + (4 * ky2 + dy) * 8 + 4 * kx2 + dx;
#pragma unroll
            for (int c = 0; c < 32; ++c) out[c * 256 + k] = __uint_as_float(r[c]) * in_scale;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 64);
}

// ---------------------------------------------------------------------------------------------
// second layer: 4x4 / stride 2 on fp32 [32][20][20] -> s2d by 2: 10x10 positions x 128 channels
// (ci, dy, dx) = two 64-channel planes (the two MN groups of an M = 128 tile), shifts
// {0, 1, 10, 11}; dZ [32][9][9] -> dZ' [32][100 (+28)]
constexpr int W2_PROD = 416;                        // 400 image items x 2 + 288 gradient items
constexpr int W2_THREADS = W2_PROD + 64 + 128;
constexpr int W2_STAGES = 4, W2_KSTEPS = 7;
constexpr uint32_t W2_PLANE = 128 * 128;            // 128 position rows x 64 bf16
constexpr uint32_t W2_IMG = 2 * W2_PLANE, W2_DZ = 2 * 32 * 128;
constexpr int W2_SMEM = W2_STAGES * (W2_IMG + W2_DZ) + 1024;

__global__ void __launch_bounds__(W2_THREADS, 1)
tc_conv2_wgrad_sw_kernel(const float* __restrict__ in, const float* __restrict__ dz, float* __restrict__ part,
                         const int R, float* __restrict__ bias_part) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[W2_STAGES], empty_bar[W2_STAGES], acc_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float bias_s[288];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int MMA_WARP = W2_PROD / 32;
    const bool has_work = (int)blockIdx.x < R;

    if (tid == 0) {
        for (int s = 0; s < W2_STAGES; ++s) {
            mbar_init(&full_bar[s], W2_PROD);
            mbar_init(&empty_bar[s], 2);
        }
        mbar_init(&acc_full, 2);
        mbar_fence_init();
    }
    // rows 100..127 of every plane and gradient positions 9, 19, .., 99..127 stay zero for good
    for (uint32_t i = tid * 16; i < W2_STAGES * (W2_IMG + W2_DZ); i += blockDim.x * 16)
        *reinterpret_cast<uint4*>(base + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers
        // image item = (channel pair cp, row Y, column pair xp): 2 ci x 2 dy float4 -> the 16-byte
        // chunks of positions (Y, 2 xp) and (Y, 2 xp + 1); threads 0..399 take items tid, tid + 400
        const bool img_t = tid < 400;
        const int xp = tid % 5, Y = (tid / 5) % 10, cp0 = tid / 50;      // cp0 0..7, second item cp0 + 8
        const int m0 = Y * 10 + 2 * xp;
        const float* src0 = in + (long long)(2 * cp0) * 400 + (2 * Y) * 20 + 4 * xp;
        const uint32_t d0 = ws_off(m0, cp0), d1 = ws_off(m0 + 1, cp0);
        // gradient item = (co, oy): 9 floats -> positions 10 oy .. 10 oy + 8; threads 0..287
        const bool dz_t = tid < 288;
        const int co = tid / 9, oy = tid - co * 9;
        float4 nx[2][4];
        float ng[9];
        auto load_img = [&](int f) {
            if (img_t) {
                const float* fb = src0 + (long long)f * 12800;
#pragma unroll
                for (int it2 = 0; it2 < 2; ++it2)
#pragma unroll
                    for (int j = 0; j < 4; ++j)                          // j = (ci & 1) * 2 + dy
                        nx[it2][j] = __ldg(reinterpret_cast<const float4*>(fb + it2 * 16 * 400 + (j >> 1) * 400 + (j & 1) * 20));
            }
        };
        auto load_dz = [&](int f) {
            if (dz_t) {
                const float* gp = dz + ((long long)f * 32 + co) * 81 + 9 * oy;
#pragma unroll
                for (int j = 0; j < 9; ++j) ng[j] = __ldg(gp + j);
            }
        };
        if (has_work) {
            load_img(blockIdx.x);
            load_dz(blockIdx.x);
        }
        float bsum = 0.f;
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int s = it % W2_STAGES;
            const bool more = fr + (int)gridDim.x < R;
            mbar_wait(&empty_bar[s], ((it / W2_STAGES) & 1) ^ 1);
            uint8_t* img = base + (size_t)s * (W2_IMG + W2_DZ);
            uint8_t* gt = img + W2_IMG;
            if (img_t) {
#pragma unroll
                for (int it2 = 0; it2 < 2; ++it2) {                      // plane it2 (ci 0..15 | 16..31)
                    uint8_t* pl = img + it2 * W2_PLANE;
                    *reinterpret_cast<uint4*>(pl + d0) =
                        make_uint4(ws_pack(nx[it2][0].x, nx[it2][0].y), ws_pack(nx[it2][1].x, nx[it2][1].y),
                                   ws_pack(nx[it2][2].x, nx[it2][2].y), ws_pack(nx[it2][3].x, nx[it2][3].y));
                    *reinterpret_cast<uint4*>(pl + d1) =
                        make_uint4(ws_pack(nx[it2][0].z, nx[it2][0].w), ws_pack(nx[it2][1].z, nx[it2][1].w),
                                   ws_pack(nx[it2][2].z, nx[it2][2].w), ws_pack(nx[it2][3].z, nx[it2][3].w));
                }
            }
            if (more) load_img(fr + gridDim.x);
            if (dz_t) {
                float bs = 0.f;
#pragma unroll
                for (int jp = 0; jp < 5; ++jp) {
                    const float v0 = ng[2 * jp], v1 = jp < 4 ? ng[2 * jp + 1] : 0.f;
                    const int m = 10 * oy + 2 * jp;
                    *reinterpret_cast<uint32_t*>(gt + (m >> 6) * 4096 + ws_off(co, (m & 63) >> 3) + (m & 7) * 2) = ws_pack(v0, v1);
                    bs += v0 + v1;
                }
                bsum += bs;
            }
            if (more) load_dz(fr + gridDim.x);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
        }
        if (bias_part) {                                                 // fixed order: thread, then rows
            if (dz_t) bias_s[tid] = bsum;
            ws_bar(1, W2_PROD);
            if (tid < 32) {
                float t = 0.f;
#pragma unroll
                for (int j = 0; j < 9; ++j) t += bias_s[tid * 9 + j];
                bias_part[blockIdx.x * 32 + tid] = t;
            }
        }
    } else if (warp < MMA_WARP + 2) {
        if (lane == 0) {                                                 // issuer ky2: shifts (10 ky2, 10 ky2 + 1)
            const int ky2 = warp - MMA_WARP;
            constexpr uint32_t idesc = umma_idesc_bf16_amn(128, 32);
            uint32_t it = 0;
            for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
                const int s = it % W2_STAGES;
                mbar_wait(&full_bar[s], (it / W2_STAGES) & 1);
                tc_fence_after();
                const uint32_t img = smem_u32(base + (size_t)s * (W2_IMG + W2_DZ)), gt = img + W2_IMG;
#pragma unroll
                for (int ks = 0; ks < W2_KSTEPS; ++ks) {                 // 16 positions per MMA
                    const uint64_t b = umma_desc_k_sw128(gt + (ks >> 2) * 4096 + (ks & 3) * 32);
#pragma unroll
                    for (int kx2 = 0; kx2 < 2; ++kx2)
                        umma_f16(tmem_base + (ky2 * 2 + kx2) * 32,
                                 umma_desc_mn_sw128(img + (ks * 16 + ky2 * 10 + kx2) * 128, W2_PLANE), b, idesc,
                                 (it | ks) != 0);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_full);
        }
    } else {
        // ================= epilogue (once): TMEM lane = channel (ci, dy, dx), 32 columns = co per shift
        const int q = warp & 3;
        if (has_work) {
            mbar_wait(&acc_full, 0);
            tc_fence_after();
        }
        float* out = part + (long long)blockIdx.x * 32 * 512;
        const int l = q * 32 + lane, ci = l >> 2, dy = (l >> 1) & 1, dx = l & 1;
#pragma unroll 1
        for (int sh = 0; sh < 4; ++sh) {
            uint32_t r[32];
            if (has_work) {
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + sh * 32, r);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int c = 0; c < 32; ++c) r[c] = 0u;
            }
            const int k = ci * 16 + (2 * (sh >> 1) + dy) * 4 + 2 * (sh & 1) + dx;
#pragma unroll
            for (int c = 0; c < 32; ++c) out[c * 512 + k] = __uint_as_float(r[c]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 128);
}

// ---------------------------------------------------------------------------------------------
// third layer: 3x3 / stride 1 on fp32 [32][9][9]: 81 positions x 32 channels.  A 128-byte row holds
// TWO column shifts, row m = [a2[m][0..31] | a2[m+1][0..31]] (kx = 0, 1); the same image two rows
// further down is kx = 2 (and a fourth, unused column) -- the second MN group of the M = 128 tile
// (leading-dimension offset 256 bytes).  One MMA per k-step and ky; dZ [32][7][7] -> dZ' [32][64]
// on the 9-wide grid, K = 64 positions.
constexpr int W3_PROD = 352;                        // 81 positions x 4 channel octets = 324 image items
constexpr int W3_THREADS = W3_PROD + 32 + 128;
constexpr int W3_STAGES = 4;
constexpr uint32_t W3_IMG = 96 * 128, W3_DZ = 32 * 128;
constexpr int W3_SMEM = W3_STAGES * (W3_IMG + W3_DZ) + 1024;

__global__ void __launch_bounds__(W3_THREADS, 1)
tc_conv3_wgrad_sw_kernel(const float* __restrict__ in, const float* __restrict__ dz, float* __restrict__ part,
                         const int R, float* __restrict__ bias_part) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[W3_STAGES], empty_bar[W3_STAGES], acc_full;
    __shared__ uint32_t tmem_slot;
    __shared__ float bias_s[224];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int MMA_WARP = W3_PROD / 32;
    const bool has_work = (int)blockIdx.x < R;

    if (tid == 0) {
        for (int s = 0; s < W3_STAGES; ++s) {
            mbar_init(&full_bar[s], W3_PROD);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_full, 1);
        mbar_fence_init();
    }
    for (uint32_t i = tid * 16; i < W3_STAGES * (W3_IMG + W3_DZ); i += blockDim.x * 16)
        *reinterpret_cast<uint4*>(base + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // image item = (position m, octet of ci): 8 loads -> chunk oc of row m and chunk 4 + oc of row m - 1
        const bool img_t = tid < 324;
        const int m = img_t ? tid % 81 : 0, oc = img_t ? tid / 81 : 0;
        const uint32_t d0 = ws_off(m, oc), d1 = ws_off(m > 0 ? m - 1 : 0, 4 + oc);
        const float* src = in + (long long)(oc * 8) * 81 + m;
        // gradient item = (co, oy): 7 floats -> positions 9 oy .. 9 oy + 6 of row co
        const bool dz_t = tid < 224;
        const int co = tid / 7, oy = tid - co * 7;
        float nx[8], ng[7];
        auto load = [&](int f) {
            if (img_t) {
#pragma unroll
                for (int e = 0; e < 8; ++e) nx[e] = __ldg(src + (long long)f * 2592 + e * 81);
            }
            if (dz_t) {
                const float* gp = dz + ((long long)f * 32 + co) * 49 + 7 * oy;
#pragma unroll
                for (int j = 0; j < 7; ++j) ng[j] = __ldg(gp + j);
            }
        };
        if (has_work) load(blockIdx.x);
        float bsum = 0.f;
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int s = it % W3_STAGES;
            const uint4 c = make_uint4(ws_pack(nx[0], nx[1]), ws_pack(nx[2], nx[3]), ws_pack(nx[4], nx[5]), ws_pack(nx[6], nx[7]));
            __nv_bfloat16 g16[7];
            float bs = 0.f;
#pragma unroll
            for (int j = 0; j < 7; ++j) {
                g16[j] = __float2bfloat16_rn(ng[j]);
                bs += ng[j];
            }
            bsum += bs;
            if (fr + (int)gridDim.x < R) load(fr + gridDim.x);
            mbar_wait(&empty_bar[s], ((it / W3_STAGES) & 1) ^ 1);
            uint8_t* img = base + (size_t)s * (W3_IMG + W3_DZ);
            uint8_t* gt = img + W3_IMG;
            if (img_t) {
                *reinterpret_cast<uint4*>(img + d0) = c;
                if (m > 0) *reinterpret_cast<uint4*>(img + d1) = c;
            }
            if (dz_t) {
#pragma unroll
                for (int j = 0; j < 7; ++j) {
                    const int mm = 9 * oy + j;
                    *reinterpret_cast<__nv_bfloat16*>(gt + ws_off(co, mm >> 3) + (mm & 7) * 2) = g16[j];
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
        }
        if (bias_part) {                                                 // fixed order: thread, then rows
            if (dz_t) bias_s[tid] = bsum;
            ws_bar(1, W3_PROD);
            if (tid < 32) {
                float t = 0.f;
#pragma unroll
                for (int j = 0; j < 7; ++j) t += bias_s[tid * 7 + j];
                bias_part[blockIdx.x * 32 + tid] = t;
            }
        }
    } else if (warp == MMA_WARP) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16_amn(128, 32);
            uint32_t it = 0;
            for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
                const int s = it % W3_STAGES;
                mbar_wait(&full_bar[s], (it / W3_STAGES) & 1);
                tc_fence_after();
                const uint32_t img = smem_u32(base + (size_t)s * (W3_IMG + W3_DZ)), gt = img + W3_IMG;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {                         // 16 positions per MMA
                    const uint64_t b = umma_desc_k_sw128(gt + ks * 32);
#pragma unroll
                    for (int ky = 0; ky < 3; ++ky)
                        umma_f16(tmem_base + ky * 32, umma_desc_mn_sw128(img + (ks * 16 + 9 * ky) * 128, 256), b, idesc,
                                 (it | ks) != 0);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_full);
        }
    } else {
        // ================= epilogue (once): TMEM lane = (kx, ci) (kx = 3 unused), 32 columns = co per ky
        const int q = warp & 3;
        if (has_work) {
            mbar_wait(&acc_full, 0);
            tc_fence_after();
        }
        float* out = part + (long long)blockIdx.x * 32 * 288;
        const int l = q * 32 + lane, kx = l >> 5, ci = l & 31;
#pragma unroll 1
        for (int ky = 0; ky < 3; ++ky) {
            uint32_t r[32];
            if (has_work) {
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + ky * 32, r);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int c = 0; c < 32; ++c) r[c] = 0u;
            }
            if (kx < 3) {
                const int k = ci * 9 + ky * 3 + kx;
#pragma unroll
                for (int c = 0; c < 32; ++c) out[c * 288 + k] = __uint_as_float(r[c]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 128);
}

}  // namespace

// part: [grid][32][256] weight partials; bias_part (may be null): [grid][32]
int launch_tc_conv1_wgrad_sw(const uint8_t* in, FrameMap fm, const float* dz, float* part, int R, float in_scale,
                             int grid, cudaStream_t st, float* bias_part) {
    R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv1_wgrad_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, W1_SMEM));
    R2D2_LAUNCH("tc_conv1_wgrad", st,
                tc_conv1_wgrad_sw_kernel<<<grid, W1_THREADS, W1_SMEM, st>>>(in, fm, dz, part, R, in_scale, bias_part));
    return R2D2_OK;
}

// second layer: `in` = layer-1 activations [R][32][20][20] fp32 (contiguous frames), part: [grid][32][512]
int launch_tc_conv2_wgrad_sw(const float* in, const float* dz, float* part, int R, int grid, cudaStream_t st,
                             float* bias_part) {
    R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv2_wgrad_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, W2_SMEM));
    R2D2_LAUNCH("tc_conv2_wgrad", st,
                tc_conv2_wgrad_sw_kernel<<<grid, W2_THREADS, W2_SMEM, st>>>(in, dz, part, R, bias_part));
    return R2D2_OK;
}

// third layer: `in` = layer-2 activations [R][32][9][9] fp32, part: [grid][32][288]
int launch_tc_conv3_wgrad_sw(const float* in, const float* dz, float* part, int R, int grid, cudaStream_t st,
                             float* bias_part) {
    R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv3_wgrad_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, W3_SMEM));
    R2D2_LAUNCH("tc_conv3_wgrad", st,
                tc_conv3_wgrad_sw_kernel<<<grid, W3_THREADS, W3_SMEM, st>>>(in, dz, part, R, bias_part));
    return R2D2_OK;
}

}  // namespace r2d2
