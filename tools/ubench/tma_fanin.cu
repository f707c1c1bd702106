// Microbenchmark behind the LSTM exchange design: G CTAs in lockstep (grid barrier per step) each
// pull the same 128 KB (or their own 128 KB) from L2 into shared memory as 32 x 4 KB pieces.
//   mode 0: 2-D tensor boxes (32 rows x 128 B, SWIZZLE_128B), every CTA reads the SAME tile
//   mode 1: 2-D tensor boxes, every CTA reads its OWN tile
//   mode 2: 1-D bulk copies of 4 KB, same          mode 3: 1-D bulk, own
//   mode 4: 1-D bulk copies of 32 KB x 4, same
//   mode 5/6: as 0/2 but the tile is REWRITTEN by the CTAs every step (1/G each) before the barrier
//   mode 8: as 0 with one mbarrier per piece (expect_tx each); mode 9: as 8 plus a try_wait on an
//           already-complete 'stage free' barrier before every issue (the kernels' producer loop)
//   mode 7: as 6, clusters of 4 with multicast (each CTA issues 8 of the 32 pieces to all four)
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/tma_fanin tools/ubench/tma_fanin.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint64_t* b, uint32_t par) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(smem_u32(b)), "r"(par) : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t par) { while (!mbar_try(b, par)) {} }
__device__ __forceinline__ void tma_2d(void* dst, const void* tm, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_1d_mc(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint16_t mask) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ int ld_acq(const int* p) { int v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ int ld_rlx(const int* p) { int v; asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void red_rel(int* p, int v) { asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }

constexpr int PIECES = 32, PIECE = 4096, TILE = PIECES * PIECE;

__global__ void __launch_bounds__(128, 1)
fanin(const __grid_constant__ CUtensorMap tm, uint8_t* buf, int* counter, long long* out, int steps, int mode) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t bar, pbar[PIECES], ebar[PIECES];
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int G = gridDim.x, c = blockIdx.x;
    const bool own = mode == 1 || mode == 3, rewrite = (mode >= 5 && mode <= 7) || mode >= 10;
    const bool mc = mode == 7;
    uint8_t* src = buf + (own ? (size_t)c * TILE : 0);
    if (threadIdx.x == 0) { for (int p = 0; p < PIECES; ++p) { mbar_init(&pbar[p], 1); mbar_init(&ebar[p], 1); } mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    if (mc) asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    long long tsum = 0;
    for (int s = 0; s < steps; ++s) {
        if (rewrite) {   // every CTA rewrites its 1/G share of the ping-pong image for this step
            uint8_t* img = buf + (size_t)(s & 1) * TILE;
            const int share = TILE / G;
            for (int i = threadIdx.x * 16; i < share; i += blockDim.x * 16)
                *reinterpret_cast<uint4*>(img + (size_t)c * share + i) = make_uint4(s, c, i, 1);
            src = img;
        }
        __syncthreads();
        if (mode >= 15) {
            // several acquire pollers in different warps, staggered in time; the first to see the
            // barrier complete claims the step and issues the loads
            __shared__ int claimed;
            __shared__ long long tshared;
            const int NP = mode == 15 ? 2 : 4;
            const int w = threadIdx.x >> 5;
            if ((threadIdx.x & 31) == 0 && w < NP) {
                if (w == 0) red_rel(counter, 1);
                else __nanosleep(w * (600 / NP));
                while (ld_acq(counter) < G * (s + 1)) {}
                if (atomicExch(&claimed, s + 1) != s + 1) {
                    asm volatile("fence.proxy.async.global;" ::: "memory");
                    const long long t0 = gtimer();
                    mbar_expect_tx(&bar, TILE);
                    for (int p = 0; p < 4; ++p) bulk_1d(base + p * 32768, src + p * 32768, 32768, &bar);
                    mbar_wait(&bar, s & 1);
                    tshared = gtimer() - t0;
                }
            }
            __syncthreads();
            if (threadIdx.x == 0) tsum += tshared;
            continue;
        }
        if (threadIdx.x == 0) {
            red_rel(counter, 1);
            if (mode == 11 || mode == 12) {
                // several relaxed polls in flight, staggered, then one acquire fence
                const int target = G * (s + 1);
                int v[4];
                const int NP = mode == 11 ? 4 : 2;
                for (int i = 0; i < NP; ++i) { v[i] = ld_rlx(counter); if (i + 1 < NP) __nanosleep(600 / NP); }
                bool done = false;
                while (!done) {
                    for (int i = 0; i < NP; ++i) {
                        if (v[i] >= target) { done = true; break; }
                        v[i] = ld_rlx(counter);
                    }
                }
                asm volatile("fence.acq_rel.gpu;" ::: "memory");
            } else {
                while (ld_acq(counter) < G * (s + 1)) {}
            }
            if (mode == 13) asm volatile("fence.proxy.async.global;" ::: "memory");
            else if (mode != 14) asm volatile("fence.proxy.async;" ::: "memory");
            const long long t0 = gtimer();
            if (mode == 8 || mode == 9) {
                for (int p = 0; p < PIECES; ++p) {
                    if (mode == 9) mbar_wait(&ebar[p], 1);
                    mbar_expect_tx(&pbar[p], PIECE);
                    tma_2d(base + p * PIECE, &tm, 0, p * 32, &pbar[p]);
                }
                const long long t1 = gtimer();
                for (int p = 0; p < PIECES; ++p) mbar_wait(&pbar[p], s & 1);
                tsum += gtimer() - t0;
                if (c == 0 && s == steps - 1) out[gridDim.x] = t1 - t0;
            } else {
            mbar_expect_tx(&bar, TILE);
            if (mode == 0 || mode == 1 || mode == 5) {
                const int row0 = (own ? c * PIECES * 32 : 0) + (rewrite ? (s & 1) * PIECES * 32 : 0);
                for (int p = 0; p < PIECES; ++p) tma_2d(base + p * PIECE, &tm, 0, row0 + p * 32, &bar);
            } else if (mode == 4 || mode >= 10) {
                for (int p = 0; p < 4; ++p) bulk_1d(base + p * 32768, src + p * 32768, 32768, &bar);
            } else if (mc) {
                const uint32_t r = cluster_rank();
                for (int p = r; p < PIECES; p += 4) bulk_1d_mc(base + p * PIECE, src + p * PIECE, PIECE, &bar, 0xF);
            } else {
                for (int p = 0; p < PIECES; ++p) bulk_1d(base + p * PIECE, src + p * PIECE, PIECE, &bar);
            }
            mbar_wait(&bar, s & 1);
            tsum += gtimer() - t0;
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) out[c] = tsum;
    if (mc) asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
    const int steps = 200;
    void* fp = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    EncodeFn enc = (EncodeFn)fp;
    uint8_t* buf; int* counter; long long* out;
    const int GMAX = 128;
    CK(cudaMalloc(&buf, (size_t)GMAX * TILE));
    CK(cudaMemset(buf, 1, (size_t)GMAX * TILE));
    CK(cudaMalloc(&counter, 256));
    CK(cudaMalloc(&out, (GMAX + 1) * 8));
    CUtensorMap tm;
    cuuint64_t dims[2] = {64, (cuuint64_t)GMAX * TILE / 128};
    cuuint64_t strides[1] = {128};
    cuuint32_t box[2] = {64, 32}, estr[2] = {1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    const int smem = TILE + 1024;
    CK(cudaFuncSetAttribute(fanin, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const char* names[] = {"2-D boxes, same tile", "2-D boxes, own tile", "1-D 4 KB, same", "1-D 4 KB, own", "1-D 32 KB, same",
                           "2-D boxes, same, rewritten each step", "1-D 4 KB, same, rewritten each step",
                           "1-D 4 KB multicast x4, rewritten each step", "2-D boxes, same, barrier per piece",
                           "2-D boxes, same, barrier per piece + stage-free try_wait",
                           "1-D 32 KB, rewritten, acquire poll", "1-D 32 KB, rewritten, 4 staggered relaxed polls + fence",
                           "1-D 32 KB, rewritten, 2 staggered relaxed polls + fence",
                           "1-D 32 KB, rewritten, acquire poll, fence.proxy.async.global", "1-D 32 KB, rewritten, acquire poll, NO proxy fence (timing only)",
                           "1-D 32 KB, rewritten, .global fence, 2 staggered acquire pollers", "1-D 32 KB, rewritten, .global fence, 4 staggered acquire pollers"};
    for (int G : {64}) {
        for (int mode = 0; mode < 17; ++mode) {
            CK(cudaMemset(counter, 0, 256));
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(G); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute at[2];
            at[0].id = cudaLaunchAttributeCooperative; at[0].val.cooperative = 1;
            int na = 1;
            if (mode == 7) { at[1].id = cudaLaunchAttributeClusterDimension; at[1].val.clusterDim.x = 4; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1; na = 2; }
            cfg.attrs = at; cfg.numAttrs = na;
            cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
            CK(cudaEventRecord(e0));
            cudaError_t le = cudaLaunchKernelEx(&cfg, fanin, tm, buf, counter, out, steps, mode);
            if (le != cudaSuccess) { printf("G=%3d mode %d launch failed: %s\n", G, mode, cudaGetErrorString(le)); cudaGetLastError(); continue; }
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            std::vector<long long> h(G);
            CK(cudaMemcpy(h.data(), out, G * 8, cudaMemcpyDeviceToHost));
            double mean = 0, mx = 0;
            for (int i = 0; i < G; ++i) { mean += h[i]; if (h[i] > mx) mx = h[i]; }
            mean /= G;
            if (mode == 8 || mode == 9) { long long ti; CK(cudaMemcpy(&ti, out + G, 8, cudaMemcpyDeviceToHost)); printf("        (issue loop of the last step: %lld ns)\n", ti); }
            printf("G=%3d  %-44s load %6.0f ns mean / %6.0f ns max per step (%.1f GB/s per SM), step %6.0f ns\n", G, names[mode],
                   mean / steps, mx / steps, TILE / (mean / steps), ms * 1e6 / steps);
        }
    }
    return 0;
}
