// Microbenchmark behind the conv1 epilogue: how fast can one B200 absorb pure writes, and how much
// does the conv epilogue's pattern (one 128-byte run per warp store, 32 runs 1600 B apart, start
// not line-aligned) lose against dense 16-byte-per-thread stores or bulk (TMA) stores from smem?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I pytorch_r2d2_b200/csrc -o tools/ubench/write_bw tools/ubench/write_bw.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

#include "ptx.cuh"

using namespace r2d2;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int FRAME = 32 * 400;   // floats of one conv1 output frame

// dense: consecutive threads write consecutive float4
__global__ void dense_kernel(float4* out, long long n4) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
        out[i] = make_float4(1.f, 2.f, 3.f, 4.f);
}

// the epilogue's pattern: warp w of a CTA owns image rows [32w', 32w'+32) of tile t; lane = row;
// 32 scalar stores, channel stride 400 floats
__global__ void epilogue_kernel(float* out, int frames) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int f = blockIdx.x; f < frames; f += gridDim.x)
        for (int tile = warp >> 2; tile < 4; tile += blockDim.x >> 7) {
            const int m = tile * 128 + (warp & 3) * 32 + lane;
            const int Y = m / 21, X = m - Y * 21;
            if (m < 441 && Y < 20 && X < 20) {
                float* o = out + (long long)f * FRAME + Y * 20 + X;
#pragma unroll
                for (int co = 0; co < 32; ++co) o[co * 400] = (float)co;
            }
        }
}

// variants of the pattern: warp store = one 128-byte run starting `shift` floats after a line
// boundary (0: whole lines, 16: two half lines, 8: sector-aligned, 3: nothing aligned), channel
// stride 400 floats as before; vec4: each thread stores a float4 (512-byte run per warp store)
__global__ void shifted_kernel(float* out, int frames, int shift) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int f = blockIdx.x; f < frames; f += gridDim.x)
        for (int seg = warp; seg < 12; seg += blockDim.x >> 5) {
            float* o = out + (long long)f * FRAME + seg * 32 + shift + lane;
#pragma unroll
            for (int co = 0; co < 32; co += 2) o[co * 400] = (float)co;     // even channels: line-aligned rows
        }
}
__global__ void vec4_kernel(float* out, int frames) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int f = blockIdx.x; f < frames; f += gridDim.x)
        for (int seg = warp; seg < 3; seg += blockDim.x >> 5) {
            float* o = out + (long long)f * FRAME + seg * 128 + lane * 4;
#pragma unroll
            for (int co = 0; co < 32; co += 2) *reinterpret_cast<float4*>(o + co * 400) = make_float4(1.f, 2.f, 3.f, 4.f);
        }
}

// bulk stores: the frame is assembled in shared memory (here: just left as it is) and leaves as
// `chunk`-byte cp.async.bulk shared -> global copies issued by the lanes of one warp
__global__ void bulk_kernel(float* out, int frames, int chunk) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int nchunk = FRAME * 4 / chunk;
    for (int i = threadIdx.x; i < FRAME; i += blockDim.x) reinterpret_cast<float*>(sm)[i] = (float)i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32)
        for (int f = blockIdx.x; f < frames; f += gridDim.x) {
            uint8_t* dst = reinterpret_cast<uint8_t*>(out + (long long)f * FRAME);
            for (int c = threadIdx.x; c < nchunk; c += 32)
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + (size_t)c * chunk),
                             "r"(smem_u32(sm + (size_t)c * chunk)), "r"(chunk)
                             : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
        }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class F>
static float timeit(F f) {
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    f();
    CK(cudaDeviceSynchronize());
    float best = 1e9f;
    for (int it = 0; it < 5; ++it) {
        CK(cudaEventRecord(a));
        f();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main() {
    const int frames = 2 * 5440;                       // both networks of the bench shape
    const long long bytes = (long long)frames * FRAME * 4;
    float* out;
    CK(cudaMalloc(&out, bytes + 4096));
    printf("%lld MB per pass\n", bytes >> 20);
    float ms = timeit([&] { dense_kernel<<<148 * 8, 256>>>(reinterpret_cast<float4*>(out), bytes / 16); });
    printf("dense float4            %.3f ms  %.0f GB/s\n", ms, bytes / ms * 1e-6);
    for (int threads : {256, 512, 1024}) {
        ms = timeit([&] { epilogue_kernel<<<148, threads>>>(out, frames); });
        printf("epilogue pattern %4d t %.3f ms  %.0f GB/s\n", threads, ms, bytes / ms * 1e-6);
    }
    ms = timeit([&] { epilogue_kernel<<<148 * 2, 512>>>(out, frames); });
    printf("epilogue 2 CTA/SM 512 t %.3f ms  %.0f GB/s\n", ms, bytes / ms * 1e-6);
    for (int shift : {0, 16, 8, 3}) {
        ms = timeit([&] { shifted_kernel<<<148, 384>>>(out, frames, shift); });
        const double b = (double)frames * 12 * 16 * 128;
        printf("128-B runs, shift %2d    %.3f ms  %.0f GB/s\n", shift, ms, b / ms * 1e-6);
    }
    ms = timeit([&] { vec4_kernel<<<148, 96>>>(out, frames); });
    printf("512-B runs (float4)     %.3f ms  %.0f GB/s\n", ms, (double)frames * 3 * 16 * 512 / ms * 1e-6);
    CK(cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FRAME * 4));
    for (int chunk : {64, 128, 256, 320, 384, 448, 480, 512, 640, 800, 1600, 3200, 51200}) {
        ms = timeit([&] { bulk_kernel<<<148, 128, FRAME * 4>>>(out, frames, chunk); });
        printf("bulk store %5d B      %.3f ms  %.0f GB/s\n", chunk, ms, bytes / ms * 1e-6);
    }
    return 0;
}
