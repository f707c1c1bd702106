// bf16 tensor-core path (north-star "2e-2" precision class): tcgen05.mma with TMEM accumulators,
// operands staged by TMA into 128B-swizzled shared memory through an mbarrier ring.
//
//   C[M,N] = A[M,K] . B[N,K]^T  (+ bias1[n] + bias2[n]) (ReLU)     A, B bf16 K-major; C f32/bf16
//
// One 128 x BN output tile per CTA.  192 threads: warp 0 = TMA producer (one elected lane),
// warp 1 = TMEM allocator + MMA issuer (one elected lane), warps 2-5 = epilogue (each owns the
// TMEM lane quarter warp_idx % 4; tcgen05.ld 32x32b -> registers -> global).
// Used for: LSTM input projection x.W_ih^T (model.py:55), dueling hidden layer (model.py:26-36),
// and, with transposed bf16 copies of the operands, their data/weight gradients.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "ptx.cuh"

namespace r2d2 {

// 3 stages (96 KB): a plain-GEMM CTA then fits on an SM beside a persistent conv-gradient CTA
// (87-99 KB), so the weight-gradient GEMMs of the backward pass fill in instead of queueing
constexpr int TC_BM = 128, TC_BK = 64, TC_STAGES = 3;

struct TcEpilogue {
    float* c32;             // may be null
    __nv_bfloat16* c16;     // may be null
    long long ldc;
    const float* bias1;     // per output column, may be null
    const float* bias2;
    int relu;
    int bias_row;           // biases indexed by output row m instead of column n
    const float* mask;      // optional [M, ldc]: outputs where mask <= 0 are written as 0 (ReLU')
};

__device__ __forceinline__ uint32_t pack2_bf16(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

template <int BN>
__global__ void __launch_bounds__(192, 1)
tc_gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
               const __grid_constant__ CUtensorMap tma_a_lo,
               const __grid_constant__ CUtensorMap tma_b_lo, const int nseg, const TcEpilogue ep,
               const int M, const int N, const int K, const int bn) {
    // bn <= BN (multiple of 16): tile width chosen by the host to fill whole waves of SMs; the
    // stage layout keeps room for BN columns
    constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = BN * TC_BK * 2;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[TC_STAGES], empty_bar[TC_STAGES], acc_bar;
    __shared__ uint32_t tmem_slot;

    // SWIZZLE_128B atoms must start on 1024-byte boundaries
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* tiles = smem_raw + ((1024u - (raw & 1023u)) & 1023u);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * TC_BM, n0 = blockIdx.y * bn;
    // nseg == 1: C = A.B^T in plain bf16.  nseg == 3: error-compensated "bf16x3" -- A = A_hi+A_lo,
    // B = B_hi+B_lo (each the bf16 rounding of the remainder) and C = A_hi.B_hi + A_hi.B_lo +
    // A_lo.B_hi accumulated in fp32 in TMEM: ~2^-16 relative operand error instead of 2^-9.
    const int kb_per_seg = (K + TC_BK - 1) / TC_BK;
    const int num_kb = kb_per_seg * nseg;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_bar, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_slot, BN);   // BN fp32 accumulator columns x 128 lanes
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % TC_STAGES;
                const uint32_t ph = (kb / TC_STAGES) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);            // slot free (first pass: immediate)
                mbar_expect_tx(&full_bar[s], A_BYTES + (uint32_t)bn * TC_BK * 2);
                uint8_t* a_dst = tiles + (size_t)s * STAGE_BYTES;
                const int seg = kb / kb_per_seg, kk = (kb - seg * kb_per_seg) * TC_BK;
                tma_load_2d(a_dst, seg == 2 ? &tma_a_lo : &tma_a, kk, m0, &full_bar[s]);
                tma_load_2d(a_dst + A_BYTES, seg == 1 ? &tma_b_lo : &tma_b, kk, n0, &full_bar[s]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(TC_BM, bn);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % TC_STAGES;
                const uint32_t ph = (kb / TC_STAGES) & 1;
                mbar_wait(&full_bar[s], ph);                 // TMA bytes have landed
                tc_fence_after();
                const uint32_t a_addr = smem_u32(tiles + (size_t)s * STAGE_BYTES);
                const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
                for (int k = 0; k < TC_BK / 16; ++k)         // UMMA_K = 16 bf16 = 32 bytes
                    umma_f16(tmem_base, umma_desc_k_sw128(a_addr + k * 32),
                             umma_desc_k_sw128(b_addr + k * 32), idesc, (kb | k) != 0);
                umma_commit(&empty_bar[s]);                  // frees the slot when the MMAs finish
            }
            umma_commit(&acc_bar);                           // accumulator complete
        }
    } else {
        // ---- epilogue: TMEM -> registers -> global, one output row per thread
        mbar_wait(&acc_bar, 0);
        tc_fence_after();
        const int q = warp & 3;                              // TMEM lane quarter of this warp
        const int m = m0 + q * 32 + lane;
#pragma unroll 1
        for (int c = 0; c < bn; c += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
            tmem_ld_wait();
            if (m < M) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const int n = n0 + c + j;
                    if (n >= N || c + j >= bn) break;
                    float v[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float x = __uint_as_float(r[j + e]);
                        if (ep.bias1) x += __ldg(ep.bias1 + (ep.bias_row ? m : n + e));
                        if (ep.bias2) x += __ldg(ep.bias2 + (ep.bias_row ? m : n + e));
                        if (ep.relu) x = fmaxf(x, 0.f);
                        v[e] = x;
                    }
                    if (ep.mask) {
                        const float4 mk = __ldg(reinterpret_cast<const float4*>(ep.mask + (long long)m * ep.ldc + n));
                        if (!(mk.x > 0.f)) v[0] = 0.f;
                        if (!(mk.y > 0.f)) v[1] = 0.f;
                        if (!(mk.z > 0.f)) v[2] = 0.f;
                        if (!(mk.w > 0.f)) v[3] = 0.f;
                    }
                    if (ep.c32)
                        *reinterpret_cast<float4*>(ep.c32 + (long long)m * ep.ldc + n) =
                            make_float4(v[0], v[1], v[2], v[3]);
                    if (ep.c16) {
                        __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]);
                        __nv_bfloat162 hi = __floats2bfloat162_rn(v[2], v[3]);
                        uint2 pk;
                        pk.x = *reinterpret_cast<uint32_t*>(&lo);
                        pk.y = *reinterpret_cast<uint32_t*>(&hi);
                        *reinterpret_cast<uint2*>(ep.c16 + (long long)m * ep.ldc + n) = pk;
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, BN);
}

// ---------------------------------------------------------------- bf16x3 GEMM
// C = A_hi.B_hi^T + A_hi.B_lo^T + A_lo.B_hi^T with all four operand tiles of a k-block in ONE
// pipeline stage, so every tile crosses L2 -> shared memory once per k-block (the three-pass form
// above re-reads A_hi and B_hi and is bound by exactly that traffic).
//   BN = 128: [B_hi ; B_lo] are adjacent in the stage = one 256-row K-major operand: A_hi multiplies
//             both in ONE N = 256 MMA (accumulator columns 0-127 and 128-255, added in the epilogue),
//             A_lo.B_hi is a second N = 128 MMA into columns 0-127.
//   BN = 256: three N = 256 MMAs into the same 256 columns; twice the output per byte loaded.
template <int BN>
__global__ void __launch_bounds__(192, 1)
tc_gemm_x3_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                  const __grid_constant__ CUtensorMap tma_a_lo,
                  const __grid_constant__ CUtensorMap tma_b_lo, const TcEpilogue ep, const int M,
                  const int N, const int K, const int bn_rt) {
    // wide form: the tile is bn <= 256 columns wide (multiple of 16, chosen by the host so that the
    // number of tiles fills whole waves of SMs); the stage layout keeps room for 256.
    // Persistent: CTA b works through tiles b, b + grid, ... (M-tile fastest, so that the CTAs of a
    // wave share their B columns in L2) with TWO accumulators of 256 TMEM columns -- the epilogue of
    // one tile (TMEM -> registers -> global) runs under the MMAs of the next.
    const int bn = BN == 128 ? BN : bn_rt;
    constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = BN * TC_BK * 2;
    constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
    constexpr int STAGES = BN == 128 ? 3 : 2;
    constexpr bool CAT = BN == 128;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* tiles = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int mt = (M + TC_BM - 1) / TC_BM, ntiles = mt * ((N + bn - 1) / bn);
    const int num_kb = (K + TC_BK - 1) / TC_BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        tma_prefetch_desc(&tma_a_lo);
        tma_prefetch_desc(&tma_b_lo);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;                                             // k-blocks issued so far (all tiles)
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                const int m0 = (t % mt) * TC_BM, n0 = (t / mt) * bn;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_bar[s], 2 * A_BYTES + 2 * (uint32_t)bn * TC_BK * 2);
                    uint8_t* dst = tiles + (size_t)s * STAGE_BYTES;
                    const int kk = kb * TC_BK;
                    tma_load_2d(dst, &tma_a, kk, m0, &full_bar[s]);
                    tma_load_2d(dst + A_BYTES, &tma_a_lo, kk, m0, &full_bar[s]);
                    tma_load_2d(dst + 2 * A_BYTES, &tma_b, kk, n0, &full_bar[s]);
                    tma_load_2d(dst + 2 * A_BYTES + B_BYTES, &tma_b_lo, kk, n0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(TC_BM, bn);
            constexpr uint32_t idesc_cat = umma_idesc_bf16(TC_BM, 256);
            uint32_t it = 0, ti = 0;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++ti) {
                const uint32_t buf = ti & 1, d = tmem_base + buf * 256;
                mbar_wait(&acc_empty[buf], ((ti >> 1) & 1) ^ 1);
                tc_fence_after();
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&full_bar[s], (it / STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(tiles + (size_t)s * STAGE_BYTES), a_lo = a_hi + A_BYTES;
                    const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + B_BYTES;
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; ++k) {
                        const uint32_t acc = (kb | k) != 0;
                        if constexpr (CAT) {
                            umma_f16(d, umma_desc_k_sw128(a_hi + k * 32), umma_desc_k_sw128(b_hi + k * 32), idesc_cat, acc);
                            umma_f16(d, umma_desc_k_sw128(a_lo + k * 32), umma_desc_k_sw128(b_hi + k * 32), idesc, 1);
                        } else {
                            umma_f16(d, umma_desc_k_sw128(a_hi + k * 32), umma_desc_k_sw128(b_hi + k * 32), idesc, acc);
                            umma_f16(d, umma_desc_k_sw128(a_hi + k * 32), umma_desc_k_sw128(b_lo + k * 32), idesc, 1);
                            umma_f16(d, umma_desc_k_sw128(a_lo + k * 32), umma_desc_k_sw128(b_hi + k * 32), idesc, 1);
                        }
                    }
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        const int q = warp & 3;
        uint32_t ti = 0;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++ti) {
            const uint32_t buf = ti & 1, d = tmem_base + buf * 256;
            const int m0 = (t % mt) * TC_BM, n0 = (t / mt) * bn;
            mbar_wait(&acc_full[buf], (ti >> 1) & 1);
            tc_fence_after();
            const int m = m0 + q * 32 + lane;
#pragma unroll 1
            for (int c = 0; c < bn; c += 32) {
                uint32_t r[32];
                tmem_ld32(d + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
                if constexpr (CAT) {
                    uint32_t r2[32];
                    tmem_ld32(d + ((uint32_t)(q * 32) << 16) + (uint32_t)(128 + c), r2);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
                } else {
                    tmem_ld_wait();
                }
                if (c + 32 >= bn) {                                      // last read of this accumulator
                    tc_fence_before();
                    mbar_arrive(&acc_empty[buf]);
                }
                if (m < M) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const int n = n0 + c + j;
                        if (n >= N || c + j >= bn) break;
                        float v[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float x = __uint_as_float(r[j + e]);
                            if (ep.bias1) x += __ldg(ep.bias1 + (ep.bias_row ? m : n + e));
                            if (ep.bias2) x += __ldg(ep.bias2 + (ep.bias_row ? m : n + e));
                            if (ep.relu) x = fmaxf(x, 0.f);
                            v[e] = x;
                        }
                        if (ep.c32)
                            *reinterpret_cast<float4*>(ep.c32 + (long long)m * ep.ldc + n) =
                                make_float4(v[0], v[1], v[2], v[3]);
                        if (ep.c16) {
                            uint2 pk;
                            pk.x = pack2_bf16(v[0], v[1]);
                            pk.y = pack2_bf16(v[2], v[3]);
                            *reinterpret_cast<uint2*>(ep.c16 + (long long)m * ep.ldc + n) = pk;
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------- bf16x3 GEMM, CTA pairs
// The wide bf16x3 kernel moves (128 + bn) x 64 x 2 x 2 bytes from L2 into shared memory per k-block
// and CTA, ~65 GB/s per SM at the measured speed: L2 bandwidth, not the tensor pipe, bounds it.
// Here two CTAs of a cluster work on the two M-tiles 2j, 2j+1 of the SAME n-tile: each loads its
// own A tiles and HALF of the B tiles, multicast into both CTAs' shared memory -- a third less L2
// traffic per CTA.  A stage may be refilled once BOTH CTAs' MMAs have read it: the "stage free"
// arrive (tcgen05.commit) is multicast to both CTAs' barriers.  Everything else as above
// (persistent, two TMEM accumulators).  The tensor maps cover this GPU's memory only.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
tc_gemm_x3_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b_half,
                       const __grid_constant__ CUtensorMap tma_a_lo,
                       const __grid_constant__ CUtensorMap tma_b_lo_half, const TcEpilogue ep, const int M,
                       const int N, const int K, const int bn) {
    constexpr int BN = 256, STAGES = 2;
    constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = BN * TC_BK * 2;
    constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[STAGES], empty_bar[STAGES], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* tiles = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
    const int mt2 = ((M + TC_BM - 1) / TC_BM) >> 1, ntiles = mt2 * ((N + bn - 1) / bn);   // tiles of a pair
    const int num_kb = (K + TC_BK - 1) / TC_BK;
    const uint32_t half_bytes = (uint32_t)(bn >> 1) * TC_BK * 2;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b_half);
        tma_prefetch_desc(&tma_a_lo);
        tma_prefetch_desc(&tma_b_lo_half);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 2);                                 // both CTAs' MMAs have read the stage
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_empty[b], 128);
        }
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                                  // the peer's barriers exist
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = pair; t < ntiles; t += npairs) {
                const int m0 = (2 * (t % mt2) + rank) * TC_BM, n0 = (t / mt2) * bn;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
                    mbar_expect_tx(&full_bar[s], 2 * A_BYTES + 4 * half_bytes);
                    uint8_t* dst = tiles + (size_t)s * STAGE_BYTES;
                    const int kk = kb * TC_BK;
                    tma_load_2d(dst, &tma_a, kk, m0, &full_bar[s]);
                    tma_load_2d(dst + A_BYTES, &tma_a_lo, kk, m0, &full_bar[s]);
                    // my half of the B rows, into both CTAs
                    tma_load_2d_mc(dst + 2 * A_BYTES + rank * half_bytes, &tma_b_half, kk, n0 + rank * (bn >> 1),
                                   &full_bar[s], 3);
                    tma_load_2d_mc(dst + 2 * A_BYTES + B_BYTES + rank * half_bytes, &tma_b_lo_half, kk,
                                   n0 + rank * (bn >> 1), &full_bar[s], 3);
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(TC_BM, bn);
            uint32_t it = 0, ti = 0;
            for (int t = pair; t < ntiles; t += npairs, ++ti) {
                const uint32_t buf = ti & 1, d = tmem_base + buf * 256;
                mbar_wait(&acc_empty[buf], ((ti >> 1) & 1) ^ 1);
                tc_fence_after();
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % STAGES;
                    mbar_wait(&full_bar[s], (it / STAGES) & 1);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(tiles + (size_t)s * STAGE_BYTES), a_lo = a_hi + A_BYTES;
                    const uint32_t b_hi = a_hi + 2 * A_BYTES, b_lo = b_hi + B_BYTES;
#pragma unroll
                    for (int k = 0; k < TC_BK / 16; ++k) {
                        const uint32_t acc = (kb | k) != 0;
                        umma_f16(d, umma_desc_k_sw128(a_hi + k * 32), umma_desc_k_sw128(b_hi + k * 32), idesc, acc);
                        umma_f16(d, umma_desc_k_sw128(a_hi + k * 32), umma_desc_k_sw128(b_lo + k * 32), idesc, 1);
                        umma_f16(d, umma_desc_k_sw128(a_lo + k * 32), umma_desc_k_sw128(b_hi + k * 32), idesc, 1);
                    }
                    umma_commit_mc(&empty_bar[s], 3);
                }
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        const int q = warp & 3;
        uint32_t ti = 0;
        for (int t = pair; t < ntiles; t += npairs, ++ti) {
            const uint32_t buf = ti & 1, d = tmem_base + buf * 256;
            const int m0 = (2 * (t % mt2) + rank) * TC_BM, n0 = (t / mt2) * bn;
            mbar_wait(&acc_full[buf], (ti >> 1) & 1);
            tc_fence_after();
            const int m = m0 + q * 32 + lane;
#pragma unroll 1
            for (int c = 0; c < bn; c += 32) {
                uint32_t r[32];
                tmem_ld32(d + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
                tmem_ld_wait();
                if (c + 32 >= bn) {
                    tc_fence_before();
                    mbar_arrive(&acc_empty[buf]);
                }
                if (m < M) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const int n = n0 + c + j;
                        if (n >= N || c + j >= bn) break;
                        float v[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float x = __uint_as_float(r[j + e]);
                            if (ep.bias1) x += __ldg(ep.bias1 + (ep.bias_row ? m : n + e));
                            if (ep.bias2) x += __ldg(ep.bias2 + (ep.bias_row ? m : n + e));
                            if (ep.relu) x = fmaxf(x, 0.f);
                            v[e] = x;
                        }
                        if (ep.c32)
                            *reinterpret_cast<float4*>(ep.c32 + (long long)m * ep.ldc + n) =
                                make_float4(v[0], v[1], v[2], v[3]);
                        if (ep.c16) {
                            uint2 pk;
                            pk.x = pack2_bf16(v[0], v[1]);
                            pk.y = pack2_bf16(v[2], v[3]);
                            *reinterpret_cast<uint2*>(ep.c16 + (long long)m * ep.ldc + n) = pk;
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();                                                  // nobody leaves while the peer still writes here
    if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------- C = A^T . B, no transposed copies
//   C[P, Q] = sum_k A[k, P] * B[k, Q]        A [Kc, P], B [Kc, Q] row-major bf16; C f32
// The weight gradients dW = dY^T . X contract over the ROWS of two row-major matrices.  Both
// operands are consumed MN-major: a TMA box of 64 elements x 64 rows lands as [k row][128 bytes]
// (SWIZZLE_128B), which is one 64-element MN group of the UMMA operand; a 128-wide tile is two such
// groups 8 KB apart, a k-step of 16 rows is 2 KB further down.  (The K-major form needs A^T and
// B^T materialised first -- five transposing passes per step in the learner.)
template <int BN>
__global__ void __launch_bounds__(192, 1)
tc_gemm_tn_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                  float* __restrict__ c, const int P, const int Q, const int Kc) {
    constexpr uint32_t G_BYTES = 64 * 128;                               // one group: 64 k rows x 64 elements
    constexpr uint32_t A_BYTES = 2 * G_BYTES, B_BYTES = (BN / 64) * G_BYTES;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[TC_STAGES], empty_bar[TC_STAGES], acc_bar;
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* tiles = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * TC_BM, n0 = blockIdx.y * BN;
    const int num_kb = (Kc + 63) / 64;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_bar, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_slot, BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % TC_STAGES;
                mbar_wait(&empty_bar[s], ((kb / TC_STAGES) & 1) ^ 1);
                mbar_expect_tx(&full_bar[s], STAGE_BYTES);               // out-of-range parts arrive as zeros
                uint8_t* dst = tiles + (size_t)s * STAGE_BYTES;
                tma_load_2d(dst, &tma_a, m0, kb * 64, &full_bar[s]);
                tma_load_2d(dst + G_BYTES, &tma_a, m0 + 64, kb * 64, &full_bar[s]);
#pragma unroll
                for (int g = 0; g < BN / 64; ++g)
                    tma_load_2d(dst + A_BYTES + g * G_BYTES, &tma_b, n0 + 64 * g, kb * 64, &full_bar[s]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(TC_BM, BN) | (1u << 15) | (1u << 16);   // A and B MN-major
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % TC_STAGES;
                mbar_wait(&full_bar[s], (kb / TC_STAGES) & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(tiles + (size_t)s * STAGE_BYTES), b_addr = a_addr + A_BYTES;
#pragma unroll
                for (int k = 0; k < 4; ++k)                              // 16 k rows per MMA
                    umma_f16(tmem_base, umma_desc_mn_sw128(a_addr + k * 2048, G_BYTES),
                             umma_desc_mn_sw128(b_addr + k * 2048, G_BYTES), idesc, (kb | k) != 0);
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_bar);
        }
    } else {
        mbar_wait(&acc_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int m = m0 + q * 32 + lane;
#pragma unroll 1
        for (int cc = 0; cc < BN; cc += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)cc, r);
            tmem_ld_wait();
            if (m < P) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const int n = n0 + cc + j;
                    if (n >= Q) break;
                    *reinterpret_cast<float4*>(c + (long long)m * Q + n) =
                        make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                                    __uint_as_float(r[j + 3]));
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, BN);
}

// ---------------------------------------------------------------- C = A . B, B as stored ([K, N] row-major)
//   C[M, N] = sum_k A[m, k] * B[k, n]        A [M, K], B [K, N] row-major bf16; C f32
// The data gradients dX = dY . W multiply by the weight matrix AS STORED ([out, in] = [K, N]):
// A is K-major as in the plain kernel, B is read MN-major like the operands of tc_gemm_tn_kernel,
// so the bf16 "hi" copy of W that the forward pass already made serves and no W^T is built.
template <int BN>
__global__ void __launch_bounds__(192, 1)
tc_gemm_nn_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                  float* __restrict__ c, const int M, const int N, const int K) {
    constexpr uint32_t G_BYTES = 64 * 128;
    constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2, B_BYTES = (BN / 64) * G_BYTES;
    constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full_bar[TC_STAGES], empty_bar[TC_STAGES], acc_bar;
    __shared__ uint32_t tmem_slot;
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* tiles = smem_raw + ((1024u - (raw & 1023u)) & 1023u);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * TC_BM, n0 = blockIdx.y * BN;
    const int num_kb = (K + TC_BK - 1) / TC_BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tma_a);
        tma_prefetch_desc(&tma_b);
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&acc_bar, 1);
        mbar_fence_init();
    }
    if (warp == 1) tmem_alloc(&tmem_slot, BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % TC_STAGES;
                mbar_wait(&empty_bar[s], ((kb / TC_STAGES) & 1) ^ 1);
                mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                uint8_t* dst = tiles + (size_t)s * STAGE_BYTES;
                tma_load_2d(dst, &tma_a, kb * TC_BK, m0, &full_bar[s]);
#pragma unroll
                for (int g = 0; g < BN / 64; ++g)
                    tma_load_2d(dst + A_BYTES + g * G_BYTES, &tma_b, n0 + 64 * g, kb * 64, &full_bar[s]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(TC_BM, BN) | (1u << 16);                // B MN-major
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % TC_STAGES;
                mbar_wait(&full_bar[s], (kb / TC_STAGES) & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(tiles + (size_t)s * STAGE_BYTES), b_addr = a_addr + A_BYTES;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_f16(tmem_base, umma_desc_k_sw128(a_addr + k * 32), umma_desc_mn_sw128(b_addr + k * 2048, G_BYTES),
                             idesc, (kb | k) != 0);
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_bar);
        }
    } else {
        mbar_wait(&acc_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int m = m0 + q * 32 + lane;
#pragma unroll 1
        for (int cc = 0; cc < BN; cc += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)cc, r);
            tmem_ld_wait();
            if (m < M) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const int n = n0 + cc + j;
                    if (n >= N) break;
                    *reinterpret_cast<float4*>(c + (long long)m * N + n) =
                        make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                                    __uint_as_float(r[j + 3]));
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, BN);
}

// ---------------------------------------------------------------- host: tensor maps
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
                cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// row-major bf16 [rows, cols] (cols contiguous); box = [box_rows, 64 cols], 128B swizzle,
// out-of-bounds elements read as zero
static int make_tmap(CUtensorMap* tm, const void* base, long long rows, long long cols,
                     int box_rows) {
    EncodeTiledFn fn = encode_fn();
    R2D2_REQUIRE(fn, "tc_gemm: cuTensorMapEncodeTiled is unavailable in this driver");
    R2D2_REQUIRE(((uintptr_t)base % 16) == 0 && (cols * 2) % 16 == 0,
                 "tc_gemm: operand base must be 16-byte aligned and K a multiple of 8 (K=%lld)", cols);
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                    box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    R2D2_REQUIRE(r == CUDA_SUCCESS, "tc_gemm: cuTensorMapEncodeTiled failed with %d", (int)r);
    return R2D2_OK;
}

// row-major bf16 [rows, cols]; box = 64 cols x 64 rows (one MN group of an MN-major operand)
static int make_tmap_mn(CUtensorMap* tm, const void* base, long long rows, long long cols) {
    EncodeTiledFn fn = encode_fn();
    R2D2_REQUIRE(fn, "tc_gemm: cuTensorMapEncodeTiled is unavailable in this driver");
    R2D2_REQUIRE(((uintptr_t)base % 16) == 0 && (cols * 2) % 16 == 0,
                 "tc_gemm: operand base must be 16-byte aligned and its row a multiple of 8 elements (%lld)", cols);
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {64, 64};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides,
                    box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    R2D2_REQUIRE(r == CUDA_SUCCESS, "tc_gemm: cuTensorMapEncodeTiled failed with %d", (int)r);
    return R2D2_OK;
}

// ---------------------------------------------------------------- casts / transposes
__global__ void cast_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                 int64_t n) {
    const int64_t n4 = n / 4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(in)[i];
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        reinterpret_cast<uint2*>(out)[i] = pk;
    }
    if (blockIdx.x == 0)
        for (int64_t i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) out[i] = __float2bfloat16_rn(in[i]);
}

// hi = bf16(x), lo = bf16(x - hi): x ~= hi + lo to ~2^-17 relative
__global__ void split_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ hi,
                                  __nv_bfloat16* __restrict__ lo, int64_t n) {
    // eight elements per thread: two 16-byte loads, one 16-byte store per output (scalar when the
    // buffers are not 16-byte aligned, and for the tail)
    const bool al = (((uintptr_t)in | (uintptr_t)hi | (uintptr_t)lo) & 15) == 0;
    const int64_t n8 = al ? n / 8 : 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float4 x0 = __ldg(reinterpret_cast<const float4*>(in) + 2 * i);
        const float4 x1 = __ldg(reinterpret_cast<const float4*>(in) + 2 * i + 1);
        uint4 h, l;
        split_bf16x2(x0.x, x0.y, h.x, l.x);
        split_bf16x2(x0.z, x0.w, h.y, l.y);
        split_bf16x2(x1.x, x1.y, h.z, l.z);
        split_bf16x2(x1.z, x1.w, h.w, l.w);
        reinterpret_cast<uint4*>(hi)[i] = h;
        reinterpret_cast<uint4*>(lo)[i] = l;
    }
    for (int64_t i = n8 * 8 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float x = in[i];
        const __nv_bfloat16 h = __float2bfloat16_rn(x);
        hi[i] = h;
        lo[i] = __float2bfloat16_rn(x - __bfloat162float(h));
    }
}

// out[c, r] = bf16(in[r, c]); 32x32 tiles through shared memory
__global__ void transpose_cast_bf16_kernel(const float* __restrict__ in,
                                           __nv_bfloat16* __restrict__ out, int R, int C) {
    __shared__ float t[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        t[i][threadIdx.x] = (r < R && c < C) ? in[(int64_t)r * C + c] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < C && r < R) out[(int64_t)c * R + r] = __float2bfloat16_rn(t[threadIdx.x][i]);
    }
}

}  // namespace r2d2

using namespace r2d2;

extern "C" {

static int tc_linear_impl(const void* x_hi, const void* x_lo, const void* w_hi, const void* w_lo,
                          const float* bias1, const float* bias2, float* y_f32, void* y_bf16, int M,
                          int N, int K, int relu, void* stream, const float* mask = nullptr,
                          int max_ctas = 0) {
    R2D2_REQUIRE(!mask || (!x_lo && ((uintptr_t)mask % 16) == 0), "tc_linear: mask is for the plain bf16 form, 16-byte aligned");
    R2D2_REQUIRE(x_hi && w_hi && (y_f32 || y_bf16) && M > 0 && N > 0 && K > 0, "tc_linear: bad args");
    R2D2_REQUIRE((x_lo == nullptr) == (w_lo == nullptr), "tc_linear: give both lo parts or neither");
    R2D2_REQUIRE(N % 4 == 0, "tc_linear: N must be a multiple of 4 (N=%d)", N);
    R2D2_REQUIRE(!y_f32 || ((uintptr_t)y_f32 % 16) == 0, "tc_linear: y_f32 must be 16-byte aligned");
    R2D2_REQUIRE(!y_bf16 || ((uintptr_t)y_bf16 % 8) == 0, "tc_linear: y_bf16 must be 8-byte aligned");
    if (x_lo) {
        // bf16x3: BN = 256 halves the operand traffic per output; the small problems (heads) keep
        // 128-wide tiles so that they still spread over the SMs
        // Tile width: cost = waves x width, the kernels being persistent (a CTA walks its tiles one
        // after the other).  128-wide tiles ([B_hi;B_lo] as one N = 256 operand, three stages) for
        // small problems; up to 256 wide (half the operand traffic per output, two stages) for
        // large ones and wherever a width between 144 and 256 saves a wave (the dueling heads:
        // 20 x 8 tiles of 128 are two waves on 148 SMs, 20 x 7 tiles of 160 are one).
        const long long mt = (M + TC_BM - 1) / TC_BM;
        const long long sms = max_ctas > 0 && max_ctas < sm_count() ? max_ctas : sm_count();
        const long long cost_cat = (mt * ((N + 127) / 128) + sms - 1) / sms * 128;
        long long best = -1;
        int BNx = 128;
        for (int bn = 256; bn >= 144; bn -= 16) {
            const long long tiles = mt * ((N + bn - 1) / bn), cost = (tiles + sms - 1) / sms * bn;
            if (best < 0 || cost < best) { best = cost; BNx = bn; }
        }
        const bool wide = mt * ((N + 255) / 256) >= 2 * sm_count() || best < cost_cat;
        if (!wide) BNx = 128;
        CUtensorMap ta, tb, ta_lo, tb_lo;
        int rc = make_tmap(&ta, x_hi, M, K, TC_BM);
        if (rc) return rc;
        rc = make_tmap(&ta_lo, x_lo, M, K, TC_BM);
        if (rc) return rc;
        rc = make_tmap(&tb, w_hi, N, K, BNx);
        if (rc) return rc;
        rc = make_tmap(&tb_lo, w_lo, N, K, BNx);
        if (rc) return rc;
        TcEpilogue ep{y_f32, reinterpret_cast<__nv_bfloat16*>(y_bf16), N, bias1, bias2, relu & 1, (relu >> 1) & 1, nullptr};
        const long long tiles_ll = (long long)((M + TC_BM - 1) / TC_BM) * ((N + BNx - 1) / BNx);
        // persistent: at most one CTA per SM; the caller may leave SMs free for a kernel that
        // must start beside this one (a CTA keeps its SM until all its tiles are done)
        const int cta_cap = max_ctas > 0 && max_ctas < sm_count() ? max_ctas : sm_count();
        const int grid = tiles_ll < cta_cap ? (int)tiles_ll : cta_cap;
        static const bool no_pairs = [] {                // R2D2_GEMM_PAIRS=0: the single-CTA kernel (A/B runs)
            const char* e = getenv("R2D2_GEMM_PAIRS");
            return e && e[0] == '0';
        }();
        if (wide && !no_pairs && ((M + TC_BM - 1) / TC_BM) % 2 == 0 && M % TC_BM == 0) {
            CUtensorMap tbh, tbh_lo;
            rc = make_tmap(&tbh, w_hi, N, K, BNx / 2);
            if (rc) return rc;
            rc = make_tmap(&tbh_lo, w_lo, N, K, BNx / 2);
            if (rc) return rc;
            const int smem = 2 * (2 * TC_BM * TC_BK * 2 + 2 * 256 * TC_BK * 2) + 1024;
            R2D2_CUDA_OK(cudaFuncSetAttribute(tc_gemm_x3_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            const long long pair_tiles = tiles_ll / 2;
            int pairs = cta_cap / 2;
            if (pairs > pair_tiles) pairs = (int)pair_tiles;
            R2D2_LAUNCH("tc_gemm_kernel_x3", as_stream(stream),
                        tc_gemm_x3_pair_kernel<<<2 * pairs, 192, smem, as_stream(stream)>>>(ta, tbh, ta_lo, tbh_lo, ep, M, N, K, BNx));
        } else if (wide) {
            const int smem = 2 * (2 * TC_BM * TC_BK * 2 + 2 * 256 * TC_BK * 2) + 1024;
            R2D2_CUDA_OK(cudaFuncSetAttribute(tc_gemm_x3_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            R2D2_LAUNCH("tc_gemm_kernel_x3", as_stream(stream),
                        tc_gemm_x3_kernel<256><<<grid, 192, smem, as_stream(stream)>>>(ta, tb, ta_lo, tb_lo, ep, M, N, K, BNx));
        } else {
            const int smem = 3 * (2 * TC_BM * TC_BK * 2 + 2 * 128 * TC_BK * 2) + 1024;
            R2D2_CUDA_OK(cudaFuncSetAttribute(tc_gemm_x3_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            R2D2_LAUNCH("tc_gemm_kernel_x3", as_stream(stream),
                        tc_gemm_x3_kernel<128><<<grid, 192, smem, as_stream(stream)>>>(ta, tb, ta_lo, tb_lo, ep, M, N, K, 128));
        }
        return R2D2_OK;
    }
    constexpr int BN = 128;
    int bn = BN;
    {   // tile width that wastes the least in the last wave (cost = waves x width); ties -> wider
        const long long mt = (M + TC_BM - 1) / TC_BM, sms = sm_count();
        long long best = -1;
        for (int w = BN; w >= 64; w -= 16) {
            const long long tiles = mt * ((N + w - 1) / w), cost = (tiles + sms - 1) / sms * w;
            if (best < 0 || cost < best) { best = cost; bn = w; }
        }
    }
    CUtensorMap ta, tb, ta_lo, tb_lo;
    int rc = make_tmap(&ta, x_hi, M, K, TC_BM);
    if (rc) return rc;
    rc = make_tmap(&tb, w_hi, N, K, bn);
    if (rc) return rc;
    rc = make_tmap(&ta_lo, x_lo ? x_lo : x_hi, M, K, TC_BM);
    if (rc) return rc;
    rc = make_tmap(&tb_lo, w_lo ? w_lo : w_hi, N, K, bn);
    if (rc) return rc;
    const int smem = TC_STAGES * (TC_BM * TC_BK * 2 + BN * TC_BK * 2) + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        R2D2_CUDA_OK(cudaFuncSetAttribute(tc_gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    TcEpilogue ep{y_f32, reinterpret_cast<__nv_bfloat16*>(y_bf16), N, bias1, bias2, relu & 1,
                  (relu >> 1) & 1, mask};
    dim3 grid((M + TC_BM - 1) / TC_BM, (N + bn - 1) / bn);
    const int nseg = x_lo ? 3 : 1;
    R2D2_LAUNCH(nseg == 3 ? "tc_gemm_kernel_x3" : "tc_gemm_kernel", as_stream(stream),
                tc_gemm_kernel<BN><<<grid, 192, smem, as_stream(stream)>>>(ta, tb, ta_lo, tb_lo, nseg,
                                                                         ep, M, N, K, bn));
    return R2D2_OK;
}

int r2d2_tc_linear(const void* x_bf16, const void* w_bf16, const float* bias1, const float* bias2,
                   float* y_f32, void* y_bf16, int M, int N, int K, int relu, void* stream) {
    return tc_linear_impl(x_bf16, nullptr, w_bf16, nullptr, bias1, bias2, y_f32, y_bf16, M, N, K, relu,
                          stream);
}

int r2d2_tc_linear_tn(const void* a_bf16, const void* b_bf16, float* y_f32, int Kc, int P, int Q, void* stream) {
    R2D2_REQUIRE(a_bf16 && b_bf16 && y_f32 && Kc > 0 && P > 0 && Q > 0, "tc_linear_tn: bad args");
    R2D2_REQUIRE(P % 8 == 0 && Q % 8 == 0 && ((uintptr_t)y_f32 % 16) == 0,
                 "tc_linear_tn: P and Q must be multiples of 8, y 16-byte aligned (P=%d Q=%d)", P, Q);
    CUtensorMap ta, tb;
    int rc = make_tmap_mn(&ta, a_bf16, Kc, P);
    if (rc) return rc;
    rc = make_tmap_mn(&tb, b_bf16, Kc, Q);
    if (rc) return rc;
    constexpr int BN = 128;
    const int smem = TC_STAGES * (2 + BN / 64) * 64 * 128 + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        R2D2_CUDA_OK(cudaFuncSetAttribute(tc_gemm_tn_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    dim3 grid((P + TC_BM - 1) / TC_BM, (Q + BN - 1) / BN);
    R2D2_LAUNCH("tc_gemm_tn_kernel", as_stream(stream),
                tc_gemm_tn_kernel<BN><<<grid, 192, smem, as_stream(stream)>>>(ta, tb, y_f32, P, Q, Kc));
    return R2D2_OK;
}

int r2d2_tc_linear_nn(const void* a_bf16, const void* b_bf16, float* y_f32, int M, int N, int K, void* stream) {
    R2D2_REQUIRE(a_bf16 && b_bf16 && y_f32 && M > 0 && N > 0 && K > 0, "tc_linear_nn: bad args");
    R2D2_REQUIRE(N % 8 == 0 && K % 8 == 0 && ((uintptr_t)y_f32 % 16) == 0,
                 "tc_linear_nn: N and K must be multiples of 8, y 16-byte aligned (N=%d K=%d)", N, K);
    CUtensorMap ta, tb;
    int rc = make_tmap(&ta, a_bf16, M, K, TC_BM);
    if (rc) return rc;
    rc = make_tmap_mn(&tb, b_bf16, K, N);
    if (rc) return rc;
    constexpr int BN = 128;
    const int smem = TC_STAGES * (TC_BM * TC_BK * 2 + (BN / 64) * 64 * 128) + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        R2D2_CUDA_OK(cudaFuncSetAttribute(tc_gemm_nn_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    dim3 grid((M + TC_BM - 1) / TC_BM, (N + BN - 1) / BN);
    R2D2_LAUNCH("tc_gemm_nn_kernel", as_stream(stream),
                tc_gemm_nn_kernel<BN><<<grid, 192, smem, as_stream(stream)>>>(ta, tb, y_f32, M, N, K));
    return R2D2_OK;
}

int r2d2_tc_linear_masked(const void* x_bf16, const void* w_bf16, const float* mask, float* y_f32,
                          int M, int N, int K, void* stream) {
    R2D2_REQUIRE(mask && y_f32, "tc_linear_masked: mask and y_f32 required");
    return tc_linear_impl(x_bf16, nullptr, w_bf16, nullptr, nullptr, nullptr, y_f32, nullptr, M, N, K, 0,
                          stream, mask);
}

int r2d2_tc_linear_x3(const void* x_hi, const void* x_lo, const void* w_hi, const void* w_lo,
                      const float* bias1, const float* bias2, float* y_f32, void* y_bf16, int M,
                      int N, int K, int relu, void* stream) {
    R2D2_REQUIRE(x_lo && w_lo, "tc_linear_x3: lo parts required");
    return tc_linear_impl(x_hi, x_lo, w_hi, w_lo, bias1, bias2, y_f32, y_bf16, M, N, K, relu, stream);
}

int r2d2_tc_linear_x3_ctas(const void* x_hi, const void* x_lo, const void* w_hi, const void* w_lo,
                           const float* bias1, const float* bias2, float* y_f32, void* y_bf16, int M,
                           int N, int K, int relu, int max_ctas, void* stream) {
    R2D2_REQUIRE(x_lo && w_lo && max_ctas >= 0, "tc_linear_x3_ctas: lo parts required, max_ctas >= 0");
    return tc_linear_impl(x_hi, x_lo, w_hi, w_lo, bias1, bias2, y_f32, y_bf16, M, N, K, relu, stream, nullptr,
                          max_ctas);
}

int r2d2_split_bf16(const float* in, void* hi_bf16, void* lo_bf16, int64_t n, void* stream) {
    R2D2_REQUIRE(in && hi_bf16 && lo_bf16 && n > 0, "split_bf16: bad args");
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    R2D2_LAUNCH("split_bf16_kernel", as_stream(stream),
                split_bf16_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(
                    in, reinterpret_cast<__nv_bfloat16*>(hi_bf16),
                    reinterpret_cast<__nv_bfloat16*>(lo_bf16), n));
    return R2D2_OK;
}

int r2d2_cast_bf16(const float* in, void* out_bf16, int64_t n, void* stream) {
    R2D2_REQUIRE(in && out_bf16 && n > 0, "cast_bf16: bad args");
    R2D2_REQUIRE(((uintptr_t)in % 16) == 0 && ((uintptr_t)out_bf16 % 8) == 0, "cast_bf16: misaligned");
    int64_t blocks = (n / 4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks < 1) blocks = 1;
    R2D2_LAUNCH("cast_bf16_kernel", as_stream(stream),
                cast_bf16_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(
                    in, reinterpret_cast<__nv_bfloat16*>(out_bf16), n));
    return R2D2_OK;
}

int r2d2_transpose_cast_bf16(const float* in, void* out_bf16, int R, int C, void* stream) {
    R2D2_REQUIRE(in && out_bf16 && R > 0 && C > 0, "transpose_cast_bf16: bad args");
    dim3 grid((C + 31) / 32, (R + 31) / 32), block(32, 8);
    R2D2_LAUNCH("transpose_cast_bf16_kernel", as_stream(stream),
                transpose_cast_bf16_kernel<<<grid, block, 0, as_stream(stream)>>>(
                    in, reinterpret_cast<__nv_bfloat16*>(out_bf16), R, C));
    return R2D2_OK;
}

}  // extern "C"
