// Shared helpers for the r2d2_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/r2d2_b200.h"

namespace r2d2 {

// thread-local error message behind r2d2_last_error()
char* err_buf();
int set_err(int code, const char* fmt, ...);

#define R2D2_REQUIRE(cond, ...)                                  \
    do {                                                         \
        if (!(cond)) return r2d2::set_err(R2D2_EINVAL, __VA_ARGS__); \
    } while (0)

#define R2D2_CUDA_OK(expr)                                                               \
    do {                                                                                 \
        cudaError_t e__ = (expr);                                                        \
        if (e__ != cudaSuccess)                                                          \
            return r2d2::set_err(R2D2_ECUDA, "%s failed: %s (%s:%d)", #expr,             \
                                 cudaGetErrorString(e__), __FILE__, __LINE__);           \
    } while (0)

#define R2D2_LAUNCH_OK() R2D2_CUDA_OK(cudaGetLastError())

// Every kernel launch of the library goes through this: counts the launch and, when the
// profiler is enabled (bench.py's roofline leg), brackets it with CUDA events on its stream.
#define R2D2_LAUNCH(name, st, ...)                \
    do {                                          \
        {                                         \
            r2d2::ProfScope ps__(name, st);       \
            __VA_ARGS__;                          \
        }                                         \
        R2D2_LAUNCH_OK();                         \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ int64_t mod_pos(int64_t a, int64_t m) {
    int64_t r = a % m;
    return r < 0 ? r + m : r;
}

// 16-byte streaming load/store (no L1 allocation): one-pass copy traffic
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(uint4* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

int sm_count();

struct ProfScope {
    ProfScope(const char* name, cudaStream_t st);
    ~ProfScope();
    const char* name_;
    cudaStream_t st_;
    bool on_;
    void* a_ = nullptr;
    void* b_ = nullptr;
};

}  // namespace r2d2
