// Data-parallel gradient exchange + optimiser step as ONE kernel over NVLink peer memory.
//
// The learner's only exchange is the sum of the flat f32 gradient over the
// ranks, followed by centered RMSprop (learner.py:51,99-100) on every replica.  With NCCL that is
// an all-reduce (reduce-scatter + all-gather of the GRADIENT) and then a full-size optimiser pass
// on every rank.  Here rank r owns shard r of the parameter vector:
//
//   1. every rank announces "my gradient of step s is complete" to its peers (release, .sys);
//   2. rank r waits for all announcements, then streams shard r of EVERY rank's gradient over
//      NVLink (P2P loads from the peers' buffers), sums it in rank order (deterministic), applies
//      RMSprop to its shard of the optimiser state (which exists only on the owner) and
//   3. writes the updated PARAMETER shard into every rank's parameter buffer (P2P stores): the
//      all-gather moves new parameters instead of summed gradients, so the optimiser arithmetic
//      and state traffic are done once per element, not once per element per rank;
//   4. announces "my shard is written", and leaves when every peer has announced the same.
//
// Buffers that peers touch (gradient, parameters, flags) are cudaMalloc'ed by this library and
// exported with CUDA IPC handles (one process per GPU); step numbers only grow, so flags never
// need resetting.  Every wait is bounded: a lost peer traps (after R2D2_PEER_TIMEOUT_S seconds,
// default 600) instead of hanging the GPU.
#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace r2d2 {

constexpr int DP_MAX_WORLD = 16;

struct DpPeers {
    float* grad[DP_MAX_WORLD];     // every rank's flat gradient buffer (mapped locally)
    float* param[DP_MAX_WORLD];    // every rank's flat parameter buffer
    int* flags[DP_MAX_WORLD];      // every rank's flag block: [0..W) gradient ready, [32..32+W) shard written,
                                   // [64] step counter, [65] local CTA-done counter
};

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Bounded wait: `budget` polls of ~0.1-0.2 us each.  The bound exists so that a dead peer ends in
// an error the host sees (the kernel traps) instead of a hung GPU; it is generous by default
// (R2D2_PEER_TIMEOUT_S, default 600 s) because a healthy peer may legitimately be late -- a long
// replay ingest, a checkpoint, a debugger.
__device__ __forceinline__ void wait_flag(const int* p, int want, unsigned long long budget) {
    for (unsigned long long it = 0; it < budget; ++it) {
        if (ld_acquire_sys(p) >= want) return;
        __nanosleep(100);
    }
    __trap();
}

__global__ void __launch_bounds__(256)
dp_reduce_rmsprop_kernel(DpPeers pe, const int rank, const int world, float* __restrict__ v,
                         float* __restrict__ mavg, const long long offset, const int bucket,
                         const long long n, const float lr, const float alpha, const float oma,
                         const float eps, const float gscale, const unsigned long long wait_budget) {
    // a bucket = the slice [offset, offset + n) of the flat vectors with its own flag block, so
    // that slices whose gradients are complete early can be exchanged while the rest of the
    // backward pass is still running
    for (int r = 0; r < world; ++r) {
        pe.grad[r] += offset;
        pe.param[r] += offset;
        pe.flags[r] += bucket * 128;
    }
    int* my = pe.flags[rank];
    const int step = my[64] + 1;               // written only by this kernel's last CTA (below)
    // 1. my gradient is complete (every earlier kernel of this stream has finished)
    if (blockIdx.x == 0 && threadIdx.x < world) {
        __threadfence_system();
        st_release_sys(pe.flags[threadIdx.x] + rank, step);
    }
    // 2. wait for everybody's gradient
    if (threadIdx.x < world) wait_flag(my + threadIdx.x, step, wait_budget);
    __syncthreads();

    // 3. my shard: [lo, hi) in float4 units; the tail (n % 4) belongs to the last rank
    const long long n4 = n / 4, per = (n4 + world - 1) / world;
    const long long lo = per * rank, hi = (lo + per < n4) ? lo + per : n4;
    // U independent positions per thread and iteration: the peer loads cross NVLink (~2 us), so the
    // bytes in flight per thread set the bandwidth
    constexpr int U = 4;
    const long long G = (long long)gridDim.x * blockDim.x;
    for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += U * G) {
        float4 g[U];
#pragma unroll
        for (int u = 0; u < U; ++u) g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
        for (int r = 0; r < world; ++r) {      // fixed order: every run sums identically
            float4 x[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const long long i = i0 + u * G;
                x[u] = i < hi ? __ldcv(reinterpret_cast<const float4*>(pe.grad[r]) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) { g[u].x += x[u].x; g[u].y += x[u].y; g[u].z += x[u].z; g[u].w += x[u].w; }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = i0 + u * G;
            if (i >= hi) break;
            const long long li = i - lo;
            float4 P = reinterpret_cast<const float4*>(pe.param[rank])[i];
            float4 V = reinterpret_cast<float4*>(v)[li], Mv = reinterpret_cast<float4*>(mavg)[li];
            float pp[4] = {P.x, P.y, P.z, P.w}, gg[4] = {g[u].x, g[u].y, g[u].z, g[u].w};
            float vv[4] = {V.x, V.y, V.z, V.w}, mm[4] = {Mv.x, Mv.y, Mv.z, Mv.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float gr = gg[j] * gscale;
                vv[j] = vv[j] * alpha + oma * gr * gr;
                mm[j] = mm[j] + oma * (gr - mm[j]);
                pp[j] = pp[j] - lr * (gr / (sqrtf(vv[j] - mm[j] * mm[j]) + eps));
            }
            reinterpret_cast<float4*>(v)[li] = make_float4(vv[0], vv[1], vv[2], vv[3]);
            reinterpret_cast<float4*>(mavg)[li] = make_float4(mm[0], mm[1], mm[2], mm[3]);
            const float4 Pn = make_float4(pp[0], pp[1], pp[2], pp[3]);
            for (int r = 0; r < world; ++r) reinterpret_cast<float4*>(pe.param[r])[i] = Pn;
        }
    }
    if (rank == world - 1 && blockIdx.x == 0) {    // scalar tail, state stored behind the shard
        for (long long i = n4 * 4 + threadIdx.x; i < n; i += blockDim.x) {
            float g = 0.f;
            for (int r = 0; r < world; ++r) g += __ldcv(pe.grad[r] + i);
            const long long li = (hi - lo) * 4 + (i - n4 * 4);
            const float gr = g * gscale;
            const float vn = v[li] * alpha + oma * gr * gr, mn = mavg[li] + oma * (gr - mavg[li]);
            v[li] = vn; mavg[li] = mn;
            const float pn = pe.param[rank][i] - lr * (gr / (sqrtf(vn - mn * mn) + eps));
            for (int r = 0; r < world; ++r) pe.param[r][i] = pn;
        }
    }

    // 4. last CTA out: publish "shard written", wait for the peers' shards, advance the step
    // one system-scope fence per CTA (cumulative over the CTA barrier), not one per thread: it has
    // to wait for the acknowledgement of every peer store before it
    __syncthreads();
    __shared__ int last;
    if (threadIdx.x == 0) {
        __threadfence_system();
        last = atomicAdd(my + 65, 1) == (int)gridDim.x - 1;
    }
    __syncthreads();
    if (last) {
        if (threadIdx.x < world) {
            __threadfence_system();
            st_release_sys(pe.flags[threadIdx.x] + 32 + rank, step);
            wait_flag(my + 32 + threadIdx.x, step, wait_budget);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            my[65] = 0;
            my[64] = step;
            __threadfence();
        }
    }
}

}  // namespace r2d2

using namespace r2d2;

extern "C" {

// cudaMalloc + IPC export of a buffer peers will map (handle: 64 bytes)
int r2d2_peer_alloc(int64_t bytes, void** ptr, void* handle64) {
    R2D2_REQUIRE(bytes > 0 && ptr && handle64, "peer_alloc: bad args");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    R2D2_CUDA_OK(cudaMalloc(ptr, (size_t)bytes));
    R2D2_CUDA_OK(cudaMemset(*ptr, 0, (size_t)bytes));
    R2D2_CUDA_OK(cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(handle64), *ptr));
    return R2D2_OK;
}
int r2d2_peer_open(const void* handle64, void** ptr) {
    R2D2_REQUIRE(handle64 && ptr, "peer_open: bad args");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, sizeof(h));
    R2D2_CUDA_OK(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return R2D2_OK;
}
int r2d2_peer_close(void* ptr) {
    R2D2_CUDA_OK(cudaIpcCloseMemHandle(ptr));
    return R2D2_OK;
}
int r2d2_peer_free(void* ptr) {
    R2D2_CUDA_OK(cudaFree(ptr));
    return R2D2_OK;
}

int64_t r2d2_dp_shard_elems(int64_t n, int world) {
    const int64_t n4 = n / 4, per = (n4 + world - 1) / world;
    return per * 4 + 4;      // float4 shard + room for the scalar tail (last rank)
}

int r2d2_dp_reduce_rmsprop(int rank, int world, void* const* grads, void* const* params,
                           void* const* flags, float* square_avg_shard, float* grad_avg_shard,
                           int64_t offset, int bucket, int64_t n, double lr, double alpha, double eps,
                           double grad_scale, void* stream) {
    R2D2_REQUIRE(offset >= 0 && offset % 4 == 0 && bucket >= 0 && bucket < 8,
                 "dp_reduce_rmsprop: offset must be a multiple of 4 elements, bucket in 0..7");
    R2D2_REQUIRE(world >= 1 && world <= DP_MAX_WORLD && rank >= 0 && rank < world,
                 "dp_reduce_rmsprop: world must be 1..%d (rank %d of %d)", DP_MAX_WORLD, rank, world);
    R2D2_REQUIRE(grads && params && flags && square_avg_shard && grad_avg_shard && n > 0,
                 "dp_reduce_rmsprop: null pointer");
    DpPeers pe;
    for (int r = 0; r < world; ++r) {
        R2D2_REQUIRE(grads[r] && params[r] && flags[r], "dp_reduce_rmsprop: rank %d is not mapped", r);
        R2D2_REQUIRE(((uintptr_t)grads[r] % 16) == 0 && ((uintptr_t)params[r] % 16) == 0,
                     "dp_reduce_rmsprop: buffers must be 16-byte aligned");
        pe.grad[r] = static_cast<float*>(grads[r]);
        pe.param[r] = static_cast<float*>(params[r]);
        pe.flags[r] = static_cast<int*>(flags[r]);
    }
    const int64_t per = (n / 4 + world - 1) / world;
    int64_t blocks = (per + 255) / 256;
    const int cap = 2 * sm_count();          // every CTA must be resident while it waits for the peers
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    // polls cost ~0.15 us each (100 ns sleep + a system-scope load)
    static const unsigned long long wait_budget = [] {
        const char* e = getenv("R2D2_PEER_TIMEOUT_S");
        double sec = e ? atof(e) : 600.0;
        if (!(sec > 0)) sec = 600.0;
        return (unsigned long long)(sec / 0.15e-6);
    }();
    cudaStream_t st = as_stream(stream);
    R2D2_LAUNCH("dp_reduce_rmsprop_kernel", st,
                dp_reduce_rmsprop_kernel<<<(unsigned)blocks, 256, 0, st>>>(
                    pe, rank, world, square_avg_shard, grad_avg_shard, (long long)offset, bucket,
                    (long long)n, (float)lr,
                    (float)alpha, (float)(1.0 - alpha), (float)eps, (float)grad_scale, wait_budget));
    return R2D2_OK;
}

}  // extern "C"
