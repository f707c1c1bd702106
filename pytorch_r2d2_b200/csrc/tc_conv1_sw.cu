// conv1 forward (model.py:14-15,46) of TWO networks over the same u8 frames, "shifted window"
// form (see tc_conv_sw.cu for the idea).  8x8 / stride 4 on 84x84x4 = space-to-depth by 4 ->
// 21x21 positions x 64 channels (ci, dy, dx), where the convolution is 2x2 / stride 1: kernel
// offsets (ky2, kx2) are row shifts {0, 1, 21, 22} of ONE shared-memory image per frame.
//
//   raw frame (28 224 B, read in place from the replay ring or a slab) -> two bulk copies (two
//   input channels each) into recycled staging slots -> producers convert every pixel ONCE to fp16
//   (exact for 0..255) and store it into the image: row = position, 128-byte row = 64 channels,
//   SWIZZLE_128B -> per frame 4 M=128 tiles x 4 offsets x 4 k-steps = 64 MMAs of N = 128
//   ([W_hi ; W_lo] of net a, then of net b: one A tile feeds both networks), fp32 accumulation in
//   TMEM (four 128-column buffers, one per tile in flight; two issuing warps take alternate
//   tiles) -> epilogue: (hi + lo part) * 1/255 + bias, ReLU, NCHW fp32 (+ ReLU bit mask of net a).
// The gather form (tc_conv.cu) builds every 8x8 receptive field from global memory, i.e. loads
// and converts each pixel four times.
#include <cuda_fp16.h>

#include "common.cuh"
#include "gemm.cuh"   // FrameMap
#include "ptx.cuh"

namespace r2d2 {

namespace {

constexpr int C1_PROD = 320;                       // producer threads (10 warps: 882 items = 3 rounds)
constexpr int C1_ISSUERS = 2;
constexpr int C1_EPI = 256;                        // epilogue threads: one warp per (lane quarter, network)
constexpr int C1_THREADS = C1_PROD + 32 * C1_ISSUERS + C1_EPI;
constexpr int C1_ROWS = 448;                       // 441 positions rounded up to whole swizzle atoms
constexpr int C1_STAGES = 2, C1_NKB = 4, C1_TILES = 4, C1_BUFS = 4;
constexpr uint32_t C1_STAGE = C1_ROWS * 128;       // one image: 448 rows x 64 fp16
constexpr uint32_t C1_WKB = 128 * 128;             // [a_hi 32 | a_lo 32 | b_hi 32 | b_lo 32] rows x 64 k
constexpr uint32_t C1_HALF = R2D2_FRAME_BYTES / 2; // two input channels
constexpr int C1_SMEM = C1_STAGES * C1_STAGE + C1_NKB * C1_WKB + 2 * C1_HALF + 1024;

__device__ __forceinline__ uint32_t c1_off(int row, int chunk) {
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}
__device__ __forceinline__ uint32_t c1_pack_f16(float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}
// two u8 pixels (bytes 2*PAIR, 2*PAIR+1 of w) -> packed fp16x2, exact: splice each byte under the
// exponent of 1024.0h with one PRMT, subtract (1024, 1024)
template <int PAIR>
__device__ __forceinline__ uint32_t c1_u8x2(uint32_t w) {
    const uint32_t m = __byte_perm(w, 0x64646464u, PAIR == 0 ? 0x4140u : 0x4342u);
    const __half2 r = __hsub2(*reinterpret_cast<const __half2*>(&m), __half2half2(__ushort_as_half(0x6400)));
    return *reinterpret_cast<const uint32_t*>(&r);
}

__global__ void __launch_bounds__(C1_THREADS, 1)
tc_conv1_sw_dual_kernel(const uint8_t* __restrict__ in, const FrameMap fm, const float* __restrict__ wa,
                        const float* __restrict__ ba, float* __restrict__ outa, const float* __restrict__ wb,
                        const float* __restrict__ bb, float* __restrict__ outb, const int R, const int skip_b,
                        uint32_t* __restrict__ relu_bits) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[C1_STAGES], empty_bar[C1_STAGES], acc_full[C1_BUFS], acc_empty[C1_BUFS],
        raw_full[2], raw_empty[2];
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* base = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    uint8_t* a_s = base;                                          // image ring first: a shifted tile's
    uint8_t* w_s = base + C1_STAGES * C1_STAGE;                   // over-read ends in the weights
    uint8_t* raw_s = w_s + C1_NKB * C1_WKB;                       // two half-frame staging slots
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    constexpr int MMA_WARP = C1_PROD / 32;

    if (tid == 0) {
        for (int s = 0; s < C1_STAGES; ++s) {
            mbar_init(&full_bar[s], C1_PROD);
            mbar_init(&empty_bar[s], C1_ISSUERS);
        }
        for (int b = 0; b < C1_BUFS; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], C1_EPI);
        }
        for (int r = 0; r < 2; ++r) {
            mbar_init(&raw_full[r], 1);
            mbar_init(&raw_empty[r], C1_PROD);
        }
        mbar_fence_init();
    }
    // resident weights: k-block o = (ky2, kx2); element e = (ci, dy, dx) -> W[co][ci][4ky2+dy][4kx2+dx]
    for (int i = tid; i < C1_NKB * 2 * 32 * 8; i += blockDim.x) {
        const int kb = i / (2 * 32 * 8), i2 = i - kb * (2 * 32 * 8);
        const int net = i2 / (32 * 8), i3 = i2 - net * (32 * 8);
        const int co = i3 >> 3, c = i3 & 7;
        const float* wsrc = net == 0 ? wa : wb;
        float v[8], l[8];
#pragma unroll
        for (int e8 = 0; e8 < 8; ++e8) {
            const int e = c * 8 + e8, ci = e >> 4, dy = (e >> 2) & 3, dx = e & 3;
            const float x = __ldg(wsrc + (long long)co * 256 + ci * 64 + (4 * (kb >> 1) + dy) * 8 + 4 * (kb & 1) + dx);
            v[e8] = x;
            l[e8] = x - __half2float(__float2half_rn(x));
        }
        const uint32_t off = (uint32_t)kb * C1_WKB + c1_off(net * 64 + co, c);
        *reinterpret_cast<uint4*>(w_s + off) = make_uint4(c1_pack_f16(v[0], v[1]), c1_pack_f16(v[2], v[3]),
                                                         c1_pack_f16(v[4], v[5]), c1_pack_f16(v[6], v[7]));
        *reinterpret_cast<uint4*>(w_s + off + 32 * 128) = make_uint4(c1_pack_f16(l[0], l[1]), c1_pack_f16(l[2], l[3]),
                                                                    c1_pack_f16(l[4], l[5]), c1_pack_f16(l[6], l[7]));
    }
    for (uint32_t i = tid * 16; i < C1_STAGES * C1_STAGE; i += blockDim.x * 16)
        *reinterpret_cast<uint4*>(a_s + i) = make_uint4(0, 0, 0, 0);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == MMA_WARP) tmem_alloc(&tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp < MMA_WARP) {
        // ================= producers
        uint32_t it = 0;
        // g = running half-frame index of this CTA -> slot g % 2; refilled as soon as it has been read
        auto fetch = [&](long long g) {
            const long long f = (long long)blockIdx.x + (g >> 1) * (long long)gridDim.x;
            if (f >= R) return;
            const int slot = (int)(g & 1);
            mbar_expect_tx(&raw_full[slot], C1_HALF);
            bulk_g2s(raw_s + (size_t)slot * C1_HALF, in + fm.base((int)f) + (size_t)slot * C1_HALF, C1_HALF, &raw_full[slot]);
        };
        if (tid == 0) {
            fetch(0);
            fetch(1);
        }
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int s = it % C1_STAGES;
            mbar_wait(&empty_bar[s], ((it / C1_STAGES) & 1) ^ 1);
            uint8_t* st = a_s + (size_t)s * C1_STAGE;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const long long g = (long long)it * 2 + half;
                const uint8_t* rp = raw_s + (size_t)half * C1_HALF;
                mbar_wait(&raw_full[half], (uint32_t)(g >> 1) & 1);
                // item = (input channel of this half, position): 4 rows x 4 bytes -> two 16-byte chunks
                for (int i = tid; i < 2 * 441; i += C1_PROD) {
                    const int c2 = i / 441, p = i - c2 * 441;
                    const int Y = p / 21, X = p - Y * 21;
                    const uint32_t* q = reinterpret_cast<const uint32_t*>(rp + c2 * 7056 + (4 * Y) * 84 + 4 * X);
                    const uint32_t r0 = q[0], r1 = q[21], r2 = q[42], r3 = q[63];
                    const int ci = half * 2 + c2;
                    *reinterpret_cast<uint4*>(st + c1_off(p, ci * 2)) =
                        make_uint4(c1_u8x2<0>(r0), c1_u8x2<1>(r0), c1_u8x2<0>(r1), c1_u8x2<1>(r1));
                    *reinterpret_cast<uint4*>(st + c1_off(p, ci * 2 + 1)) =
                        make_uint4(c1_u8x2<0>(r2), c1_u8x2<1>(r2), c1_u8x2<0>(r3), c1_u8x2<1>(r3));
                }
                mbar_arrive(&raw_empty[half]);
                if (tid == 0) {
                    mbar_wait(&raw_empty[half], (uint32_t)(g >> 1) & 1);
                    fetch(g + 2);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(&full_bar[s]);
        }
    } else if (warp < MMA_WARP + C1_ISSUERS) {
        // ================= MMA issuers: alternate tiles, each tile its own accumulator buffer
        if (lane == 0) {
            const int issuer = warp - MMA_WARP;
            constexpr uint32_t idesc = umma_idesc_f16(128, 128);
            const uint32_t wbase = smem_u32(w_s);
            uint32_t it = 0;
            for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
                const int s = it % C1_STAGES;
                mbar_wait(&full_bar[s], (it / C1_STAGES) & 1);
                tc_fence_after();
                const uint32_t st = smem_u32(a_s + (size_t)s * C1_STAGE);
                for (int tile = issuer; tile < C1_TILES; tile += C1_ISSUERS) {
                    const uint32_t tg = it * C1_TILES + tile;            // running tile index
                    const uint32_t buf = tg % C1_BUFS, use = tg / C1_BUFS;
                    mbar_wait(&acc_empty[buf], (use & 1) ^ 1);
                    tc_fence_after();
                    const uint32_t d = tmem_base + buf * 128;
#pragma unroll
                    for (int kb = 0; kb < C1_NKB; ++kb) {
                        const uint32_t a = st + (tile * 128 + (kb >> 1) * 21 + (kb & 1)) * 128;
                        const uint32_t b = wbase + kb * C1_WKB;
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_f16(d, umma_desc_k_sw128(a + k * 32), umma_desc_k_sw128(b + k * 32), idesc, (kb | k) != 0);
                    }
                    umma_commit(&acc_full[buf]);
                }
                umma_commit(&empty_bar[s]);
            }
        }
    } else {
        // ================= epilogue: one image row (TMEM lane) per thread, one network per warp group
        const int ew = warp - (MMA_WARP + C1_ISSUERS);
        const int q = warp & 3, net = ew >> 2;                           // TMEM lane quarter = warp id % 4
        float* const outn = net == 0 ? outa : outb;
        const float* const bs = net == 0 ? ba : bb;
        const int skip = net == 0 ? 0 : skip_b;                          // net b starts skip_b frames later
        uint32_t it = 0;
        for (int fr = blockIdx.x; fr < R; fr += gridDim.x, ++it) {
            const int fo = fr - skip;
            for (int tile = 0; tile < C1_TILES; ++tile) {
                const uint32_t tg = it * C1_TILES + tile;
                const uint32_t buf = tg % C1_BUFS, use = tg / C1_BUFS;
                mbar_wait(&acc_full[buf], use & 1);
                tc_fence_after();
                const int m = tile * 128 + q * 32 + lane;
                const int Y = m / 21, X = m - Y * 21;
                const bool valid = m < 441 && Y < 20 && X < 20;
                const int p = Y * 20 + X;
                uint32_t r[32], r2[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * 128 + net * 64, r);
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + buf * 128 + net * 64 + 32, r2);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&acc_empty[buf]);
                if (valid && fo >= 0) {
                    float* o = outn + (long long)fo * 32 * 400 + p;
                    uint32_t bits = 0u;
#pragma unroll
                    for (int co = 0; co < 32; ++co) {
                        const float acc = __uint_as_float(r[co]) + __uint_as_float(r2[co]);
                        const float x = fmaxf(fmaf(acc, 1.f / 255.f, __ldg(bs + co)), 0.f);
                        o[(long long)co * 400] = x;
                        bits |= (x > 0.f ? 1u : 0u) << co;
                    }
                    if (net == 0 && relu_bits) relu_bits[(long long)fo * 400 + p] = bits;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 512);
}

}  // namespace

int launch_tc_conv1_sw_dual(const uint8_t* in, FrameMap fm, const float* wa, const float* ba, float* outa,
                            const float* wb, const float* bb, float* outb, int R, int skip_b, cudaStream_t st,
                            uint32_t* relu_bits_a) {
    R2D2_CUDA_OK(cudaFuncSetAttribute(tc_conv1_sw_dual_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C1_SMEM));
    int grid = sm_count();
    if (grid > R) grid = R;
    R2D2_LAUNCH("tc_conv1_fwd_dual", st,
                tc_conv1_sw_dual_kernel<<<grid, C1_THREADS, C1_SMEM, st>>>(in, fm, wa, ba, outa, wb, bb, outb, R, skip_b,
                                                                          relu_bits_a));
    return R2D2_OK;
}

}  // namespace r2d2
