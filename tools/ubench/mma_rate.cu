// Microbenchmark behind the conv / LSTM tile-shape choices: how long does the tensor pipe of ONE SM
// spend on a small tcgen05.mma (kind::f16, bf16 operands from 128B-swizzled shared memory, K = 16)
// as a function of its shape, and does a second issuing thread help?
//   for each (M, N): `reps` accumulating MMAs issued back to back by one thread (or split between
//   two threads of different warps, each into its own TMEM columns), one commit, wait; cycles/MMA.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I pytorch_r2d2_b200/csrc -o tools/ubench/mma_rate tools/ubench/mma_rate.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

#include "ptx.cuh"

using namespace r2d2;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(128, 1)
mma_rate(long long* out, int M, int N, int reps, int issuers, int a_rows_shift) {
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t done;
    __shared__ uint32_t tmem_slot;
    uint8_t* base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(base)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(&done, issuers); mbar_fence_init(); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x < 32) tmem_alloc(&tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tm = tmem_slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    long long t0 = 0, t1 = 0;
    if (lane == 0 && warp < issuers) {
        const uint32_t idesc = umma_idesc_bf16(M, N);
        const uint32_t a = smem_u32(base) + a_rows_shift * 128, b = smem_u32(base + 48 * 1024);
        const uint32_t d = tm + warp * 256;
        asm volatile("mov.u64 %0, %%clock64;" : "=l"(t0));
        for (int r = 0; r < reps / issuers; ++r) {
            // walk over 4 k-steps x 4 tiles like the kernels do
            const uint32_t ka = a + (r & 3) * 32 + ((r >> 2) & 1) * 16384;
            umma_f16(d, umma_desc_k_sw128(ka), umma_desc_k_sw128(b + (r & 3) * 32), idesc, r != 0);
        }
        umma_commit(&done);
    }
    if (threadIdx.x == 0) {
        mbar_wait(&done, 0);
        asm volatile("mov.u64 %0, %%clock64;" : "=l"(t1));
        out[0] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

int main() {
    long long* out;
    CK(cudaMalloc(&out, 8));
    const int smem = 97 * 1024;
    CK(cudaFuncSetAttribute(mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int reps = 256;
    const int shapes[][2] = {{128, 16}, {128, 32}, {128, 64}, {128, 128}, {128, 256}, {64, 16}, {64, 32}, {64, 64}, {64, 128}, {64, 256}};
    printf("%6s %6s %8s %14s %14s %14s\n", "M", "N", "issuers", "clk/MMA", "MAC/clk", "shifted-A clk");
    for (auto& s : shapes)
        for (int issuers = 1; issuers <= 2; ++issuers) {
            double clk[2];
            for (int sh = 0; sh < 2; ++sh) {
                long long best = 1ll << 60;
                for (int it = 0; it < 5; ++it) {
                    mma_rate<<<1, 128, smem>>>(out, s[0], s[1], reps, issuers, sh ? 11 : 0);
                    CK(cudaDeviceSynchronize());
                    long long h;
                    CK(cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost));
                    if (h < best) best = h;
                }
                clk[sh] = (double)best / reps;
            }
            printf("%6d %6d %8d %14.1f %14.0f %14.1f\n", s[0], s[1], issuers, clk[0], (double)s[0] * s[1] * 16 / clk[0], clk[1]);
        }
    return 0;
}
