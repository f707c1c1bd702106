// Sequence-priority arithmetic shared by the replay and the actor kernels (sm_100a only):
// eta*max + (1-eta)*mean over a window of per-step priorities, in numpy's float32 order
// (/root/reference replay_memory.py:187-188; oracle/replay_oracle.py:np_pairwise_sum_f32).
// `pr` is deliberately not __restrict__: the actor kernel reads priorities it wrote itself.
#pragma once
#include "common.cuh"

namespace r2d2 {

// numpy float32 pairwise sum over a window of `n` priorities starting at `start`
// (wrap: element i is priority[(start+i) mod M]); association as numpy's add.reduce.
static __device__ float pw_sum(const float* pr, int64_t start, int n, int64_t M,
                        bool wrap) {
    auto at = [&](int i) -> float {
        int64_t s = start + i;
        if (wrap) s = mod_pos(s, M);
        return pr[s];
    };
    if (n < 8) {
        float res = 0.f;
        for (int i = 0; i < n; ++i) res = __fadd_rn(res, at(i));
        return res;
    }
    if (n <= 128) {
        float r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = at(j);
        int i = 8;
        for (; i < n - (n % 8); i += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = __fadd_rn(r[j], at(i + j));
        }
        float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                              __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
        for (; i < n; ++i) res = __fadd_rn(res, at(i));
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __fadd_rn(pw_sum(pr, start, n2, M, wrap), pw_sum(pr, start + n2, n - n2, M, wrap));
}

static __device__ float seq_priority_value(const float* pr, int64_t start, int n,
                                    int64_t M, bool wrap, float eta, float ometa) {
    float mx = -INFINITY;
    for (int i = 0; i < n; ++i) {
        int64_t s = start + i;
        if (wrap) s = mod_pos(s, M);
        mx = fmaxf(mx, pr[s]);
    }
    const float mean = __fdiv_rn(pw_sum(pr, start, n, M, wrap), (float)n);
    return __fadd_rn(__fmul_rn(eta, mx), __fmul_rn(ometa, mean));
}

}  // namespace r2d2
