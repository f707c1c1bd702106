// Conv data gradients in "shifted window" form (plain bf16), ReLU' of the layer input fused.
//
// Second layer (4x4 / stride 2, [32][9][9] gradient -> [32][20][20]): on the space-to-depth grid
// of the layer input (10x10 positions m = 10 Y + X, 128 channels n = (ci, dy, dx)) the transposed
// convolution is
//     dIn[m][n] = sum over (ky2, kx2) and co of  dZ'[m - 10 ky2 - kx2][co] * W[co][ci][2 ky2 + dy][2 kx2 + dx]
// with dZ' = dZ laid out on the same grid (row 9 and column 9 are zero).  One frame = one
// M = 128 tile (100 positions), N = 128, K = 4 shifts x 32 co.  The gradient image is stored with
// TWO column shifts per 128-byte row -- row m = [dZ'[m][0..31] | dZ'[m-1][0..31]] -- so a k-block
// of 64 is one row window and the row shift ky2 is a start address 10 rows further up: every
// gradient value is loaded once and stored twice, instead of being gathered four times.
// The output frame is assembled in shared memory as it lies in global memory ([ci][20][20] fp32,
// 51 200 bytes) and leaves as ONE bulk copy: 2x2 blocks stored straight from the accumulator
// rows are 8-byte pieces of 32-byte sectors shared between warps, which costs more than half the
// write bandwidth (tools/ubench/write_bw.cu).
#include <cuda_bf16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace r2d2 {

namespace {

__device__ __forceinline__ uint32_t dg_off(int row, int chunk) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}
__device__ __forceinline__ uint32_t dg_pack(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void dg_bar(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

constexpr int D2_PROD = 352;                        // 81 positions x 4 channel octets = 324 items
constexpr int D2_EPI = 256;                         // two warps per TMEM lane quarter, half the channels each
constexpr int D2_THREADS = D2_PROD + 32 + D2_EPI;
constexpr int D2_STAGES = 4, D2_MARGIN = 16;        // zero rows in front of the image (shift -10 -1)
constexpr uint32_t D2_IMG = 128 * 128;              // 16 margin + 100 positions (+12) rows x 128 B
constexpr uint32_t D2_WKB = 128 * 128;              // weights: 128 channels x 64 k, two k-blocks
constexpr uint32_t D2_OUT = 32 * 400 * 4;           // one output frame
constexpr int D2_SMEM = 2 * D2_WKB + D2_STAGES * D2_IMG + 2 * D2_OUT + 1024;

__global__ void __launch_bounds__(D2_THREADS, 1)
tc_conv2_dgrad_sw_kernel(const float* __restrict__ dz, const float* __restrict__ w,
                         const uint32_t* __restrict__ maskbits, float* __restrict__ din, const int R) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[D2_STAGES], empty_bar[D2_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t* w_s = base;
    uint8_t* a_s = base + 2 * D2_WKB;
    uint8_t* o_s = a_s + D2_STAGES * D2_IMG;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int MMA_WARP = D2_PROD / 32;

    if (tid == 0) {
        for (int s = 0; s < D2_STAGES; ++s) {
            mbar_init(&full_bar[s], D2_PROD);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], D2_EPI);
        }
        mbar_fence_init();
    }
    // resident B: row n = (ci, dy, dx), k-block ky2, k = kx2 * 32 + co -> W[co][ci][2 ky2 + dy][2 kx2 + dx]
    for (int i = tid; i < 2 * 128 * 8; i += blockDim.x) {
        const int kb = i >> 10, n = (i >> 3) & 127, c = i & 7;
        const int ci = n >> 2, dy = (n >> 1) & 1, dx = n & 1;
        const int kx2 = c >> 2;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int co = (c & 3) * 8 + e;
            v[e] = __ldg(w + ((long long)(co * 32 + ci) * 4 + 2 * kb + dy) * 4 + 2 * kx2 + dx);
        }
        *reinterpret_cast<uint4*>(w_s + (uint32_t)kb * D2_WKB + dg_off(n, c)) =
            make_uint4(dg_pack(v[0], v[1]), dg_pack(v[2], v[3]), dg_pack(v[4], v[5]), dg_pack(v[6], v[7]));
    }
    // images: the margin, row 9 / column 9 of the grid and the second halves nobody writes stay zero
    for (uint32_t i = tid * 16; i < D2_STAGES * D2_IMG; i += blockDim.x * 16)
        *reinterpret_cast<uint4*>(a_s + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers: item = (output position (oy, ox), octet of co): 8 loads
        // (coalesced over the positions), one 16-byte chunk into row m and one into row m + 1
        const bool act = tid < 324;
        const int p = act ? tid % 81 : 0, oc = act ? tid / 81 : 0;
        const int oy = p / 9, ox = p - oy * 9;
        const int m = oy * 10 + ox;
        const uint32_t d0 = dg_off(D2_MARGIN + m, oc), d1 = dg_off(D2_MARGIN + m + 1, 4 + oc);
        const float* src = dz + (long long)(oc * 8) * 81 + p;
        float nx[8];
        auto load = [&](int f) {
            if (act) {
#pragma unroll
                for (int e = 0; e < 8; ++e) nx[e] = __ldg(src + (long long)f * 2592 + e * 81);
            }
        };
        if ((int)blockIdx.x < R) load(blockIdx.x);
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int s = it % D2_STAGES;
            const uint4 c = make_uint4(dg_pack(nx[0], nx[1]), dg_pack(nx[2], nx[3]), dg_pack(nx[4], nx[5]), dg_pack(nx[6], nx[7]));
            if (fr + (int)gridDim.x < R) load(fr + gridDim.x);
            mbar_wait(&empty_bar[s], ((it / D2_STAGES) & 1) ^ 1);
            if (act) {
                uint8_t* img = a_s + (size_t)s * D2_IMG;
                *reinterpret_cast<uint4*>(img + d0) = c;
                *reinterpret_cast<uint4*>(img + d1) = c;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
        }
    } else if (warp == MMA_WARP) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, 128);
            const uint32_t wb = smem_u32(w_s);
            uint32_t it = 0;
            for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
                const int s = it % D2_STAGES;
                const uint32_t buf = it & 1;
                mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
                mbar_wait(&full_bar[s], (it / D2_STAGES) & 1);
                tc_fence_after();
                const uint32_t img = smem_u32(a_s + (size_t)s * D2_IMG);
#pragma unroll
                for (int ky2 = 0; ky2 < 2; ++ky2)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(tmem_base + buf * 128, umma_desc_k_sw128(img + (D2_MARGIN - 10 * ky2) * 128 + k * 32),
                                 umma_desc_k_sw128(wb + ky2 * D2_WKB + k * 32), idesc, (ky2 | k) != 0);
                umma_commit(&empty_bar[s]);
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        // ================= epilogue: TMEM lane = position (Y, X); 128 columns = (ci, dy, dx)
        const int q = warp & 3, et = q * 32 + lane;                      // et = position m
        const int half = (warp - (MMA_WARP + 1)) >> 2;                   // channels 16 half .. 16 half + 15
        const bool leader = et == 0 && half == 0;
        const bool valid = et < 100;
        const int Y = et / 10, X = et - Y * 10;
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const uint32_t buf = it & 1;
            float* stg = reinterpret_cast<float*>(o_s + (size_t)buf * D2_OUT);
            // ReLU' bits of this thread's four pixels (bit ci of word (frame, pixel))
            uint2 mk0 = make_uint2(0u, 0u), mk1 = make_uint2(0u, 0u);
            if (valid) {
                const uint32_t* mb = maskbits + (long long)fr * 400 + (2 * Y) * 20 + 2 * X;
                mk0 = __ldg(reinterpret_cast<const uint2*>(mb));
                mk1 = __ldg(reinterpret_cast<const uint2*>(mb + 20));
            }
            // the staging buffer was handed to the bulk copy two frames ago: it has been read
            if (leader) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            dg_bar(1, D2_EPI);
            mbar_wait(&acc_full[buf], (it >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int cc = 2 * half; cc < 2 * half + 2; ++cc) {           // 8 input channels at a time
                uint32_t r[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * 128 + cc * 32, r);
                tmem_ld_wait();
                if (cc == 2 * half + 1) {
                    tc_fence_before();
                    mbar_arrive(&acc_empty[buf]);
                }
                if (valid) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int ci = cc * 8 + j;
                        float2 top, bot;
                        top.x = (mk0.x >> ci) & 1u ? __uint_as_float(r[4 * j]) : 0.f;
                        top.y = (mk0.y >> ci) & 1u ? __uint_as_float(r[4 * j + 1]) : 0.f;
                        bot.x = (mk1.x >> ci) & 1u ? __uint_as_float(r[4 * j + 2]) : 0.f;
                        bot.y = (mk1.y >> ci) & 1u ? __uint_as_float(r[4 * j + 3]) : 0.f;
                        float* o = stg + ci * 400 + (2 * Y) * 20 + 2 * X;
                        *reinterpret_cast<float2*>(o) = top;
                        *reinterpret_cast<float2*>(o + 20) = bot;
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            dg_bar(1, D2_EPI);
            if (leader) {
                bulk_s2g(din + (long long)fr * 12800, stg, D2_OUT);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
        if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 256);
}

// ---------------------------------------------------------------------------------------------
// Third layer (3x3 / stride 1, [32][7][7] gradient -> [32][9][9]), same scheme on the 9x9 grid of
// the layer input:  dIn[m][ci] = sum over (ky, kx), co of dZ'[m - 9 ky - kx][co] * W[co][ci][ky][kx].
// Row m of the image = [dZ'[m] | dZ'[m-1]]: the k-block at row shift 9 ky holds kx = 0, 1, the one
// at 9 ky + 2 holds kx = 2 (and a fourth column with zero weights).  6 k-blocks, N = 32.
// ReLU' from the layer's forward activations (a2 > 0).
constexpr int D3_PROD = 224;                        // 49 positions x 4 channel octets = 196 items
constexpr int D3_THREADS = D3_PROD + 32 + 128;
constexpr int D3_STAGES = 4, D3_MARGIN = 24;
constexpr uint32_t D3_IMG = 128 * 128;
constexpr uint32_t D3_WKB = 32 * 128;               // 32 input channels x 64 k
constexpr uint32_t D3_OUT = 32 * 81 * 4;            // one output frame (10 368 bytes)
constexpr int D3_SMEM = 6 * D3_WKB + D3_STAGES * D3_IMG + 2 * D3_OUT + 1024;

__global__ void __launch_bounds__(D3_THREADS, 2)
tc_conv3_dgrad_sw_kernel(const float* __restrict__ dz, const float* __restrict__ w, const float* __restrict__ act,
                         float* __restrict__ din, const int R) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[D3_STAGES], empty_bar[D3_STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t* w_s = base;
    uint8_t* a_s = base + 6 * D3_WKB;
    uint8_t* o_s = a_s + D3_STAGES * D3_IMG;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int MMA_WARP = D3_PROD / 32;

    if (tid == 0) {
        for (int s = 0; s < D3_STAGES; ++s) {
            mbar_init(&full_bar[s], D3_PROD);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
        }
        mbar_fence_init();
    }
    // resident B: row ci, k-block kb = 2 ky + (kx >> 1), k = (kx & 1) * 32 + co -> W[co][ci][ky][kx] (kx = 3: 0)
    for (int i = tid; i < 6 * 32 * 8; i += blockDim.x) {
        const int kb = i >> 8, ci = (i >> 3) & 31, c = i & 7;
        const int ky = kb >> 1, kx = 2 * (kb & 1) + (c >> 2);
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int co = (c & 3) * 8 + e;
            v[e] = kx < 3 ? __ldg(w + ((long long)(co * 32 + ci) * 3 + ky) * 3 + kx) : 0.f;
        }
        *reinterpret_cast<uint4*>(w_s + (uint32_t)kb * D3_WKB + dg_off(ci, c)) =
            make_uint4(dg_pack(v[0], v[1]), dg_pack(v[2], v[3]), dg_pack(v[4], v[5]), dg_pack(v[6], v[7]));
    }
    for (uint32_t i = tid * 16; i < D3_STAGES * D3_IMG; i += blockDim.x * 16)
        *reinterpret_cast<uint4*>(a_s + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 64);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers: item = (output position (oy, ox), octet of co)
        const bool actv = tid < 196;
        const int p = actv ? tid % 49 : 0, oc = actv ? tid / 49 : 0;
        const int oy = p / 7, ox = p - oy * 7;
        const int m = oy * 9 + ox;
        const uint32_t d0 = dg_off(D3_MARGIN + m, oc), d1 = dg_off(D3_MARGIN + m + 1, 4 + oc);
        const float* src = dz + (long long)(oc * 8) * 49 + p;
        float nx[8];
        auto load = [&](int f) {
            if (actv) {
#pragma unroll
                for (int e = 0; e < 8; ++e) nx[e] = __ldg(src + (long long)f * 1568 + e * 49);
            }
        };
        if ((int)blockIdx.x < R) load(blockIdx.x);
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int s = it % D3_STAGES;
            const uint4 c = make_uint4(dg_pack(nx[0], nx[1]), dg_pack(nx[2], nx[3]), dg_pack(nx[4], nx[5]), dg_pack(nx[6], nx[7]));
            if (fr + (int)gridDim.x < R) load(fr + gridDim.x);
            mbar_wait(&empty_bar[s], ((it / D3_STAGES) & 1) ^ 1);
            if (actv) {
                uint8_t* img = a_s + (size_t)s * D3_IMG;
                *reinterpret_cast<uint4*>(img + d0) = c;
                *reinterpret_cast<uint4*>(img + d1) = c;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
        }
    } else if (warp == MMA_WARP) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, 32);
            const uint32_t wb = smem_u32(w_s);
            uint32_t it = 0;
            for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
                const int s = it % D3_STAGES;
                const uint32_t buf = it & 1;
                mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
                mbar_wait(&full_bar[s], (it / D3_STAGES) & 1);
                tc_fence_after();
                const uint32_t img = smem_u32(a_s + (size_t)s * D3_IMG);
#pragma unroll
                for (int kb = 0; kb < 6; ++kb)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        umma_f16(tmem_base + buf * 32,
                                 umma_desc_k_sw128(img + (D3_MARGIN - 9 * (kb >> 1) - 2 * (kb & 1)) * 128 + k * 32),
                                 umma_desc_k_sw128(wb + kb * D3_WKB + k * 32), idesc, (kb | k) != 0);
                umma_commit(&empty_bar[s]);
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        // ================= epilogue: TMEM lane = input position m; 32 columns = ci
        const int q = warp & 3, et = q * 32 + lane;
        const bool valid = et < 81;
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const uint32_t buf = it & 1;
            float* stg = reinterpret_cast<float*>(o_s + (size_t)buf * D3_OUT);
            float a[32];                                                 // the layer's forward activations (ReLU')
            if (valid) {
                const float* ap = act + (long long)fr * 2592 + et;
#pragma unroll
                for (int ci = 0; ci < 32; ++ci) a[ci] = __ldg(ap + ci * 81);
            }
            if (et == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            dg_bar(1, 128);
            mbar_wait(&acc_full[buf], (it >> 1) & 1);
            tc_fence_after();
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * 32, r);
            tmem_ld_wait();
            tc_fence_before();
            mbar_arrive(&acc_empty[buf]);
            if (valid) {
#pragma unroll
                for (int ci = 0; ci < 32; ++ci) stg[ci * 81 + et] = a[ci] > 0.f ? __uint_as_float(r[ci]) : 0.f;
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            dg_bar(1, 128);
            if (et == 0) {
                bulk_s2g(din + (long long)fr * 2592, stg, D3_OUT);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
        if (et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 64);
}

}  // namespace

int launch_tc_conv2_dgrad_sw(const float* dz2, const float* w2, const uint32_t* maskbits, float* dz1, int R,
                             cudaStream_t st) {
    R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv2_dgrad_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, D2_SMEM));
    int grid = sm_count();
    if (grid > R) grid = R;
    R2D2_LAUNCH("tc_conv2_dgrad", st, tc_conv2_dgrad_sw_kernel<<<grid, D2_THREADS, D2_SMEM, st>>>(dz2, w2, maskbits, dz1, R));
    return R2D2_OK;
}

int launch_tc_conv3_dgrad_sw(const float* dz3, const float* w3, const float* a2, float* dz2, int R, cudaStream_t st) {
    R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv3_dgrad_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, D3_SMEM));
    int grid = 2 * sm_count();                                           // two CTAs per SM
    if (grid > R) grid = R;
    R2D2_LAUNCH("tc_conv3_dgrad", st, tc_conv3_dgrad_sw_kernel<<<grid, D3_THREADS, D3_SMEM, st>>>(dz3, w3, a2, dz2, R));
    return R2D2_OK;
}

}  // namespace r2d2
