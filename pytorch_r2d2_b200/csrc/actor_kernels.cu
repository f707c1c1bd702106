// Actor-side experience construction for N actors stepped together (sm_100a).
// One CTA per actor: n-step ring append, 0..n emissions into the actor's local replay ring
// (with their initial priorities), and at the end of an episode the sequence starts and their
// priorities.  Restates /root/reference actor.py:74-94,144-167 and replay_memory.py:12-55,
// 93-105 (see include/r2d2_b200.h, r2d2_actor_observe); oracle: oracle/actor_oracle.py.
#include "common.cuh"
#include "priority.cuh"

namespace r2d2 {

constexpr int FRAME = R2D2_FRAME_BYTES;
constexpr int NA = R2D2_ACTIONS;

__global__ void __launch_bounds__(256)
actor_observe_kernel(r2d2_actor_buffers b, int n, int H, int64_t C, int T, int overlap,
                     const float* __restrict__ obs, const int64_t* __restrict__ action,
                     const float* __restrict__ q, const float* __restrict__ h,
                     const float* __restrict__ c, const float* __restrict__ tq,
                     const float* __restrict__ th, const float* __restrict__ tc,
                     const double* __restrict__ reward, const uint8_t* __restrict__ done,
                     const double* __restrict__ gamma_pow, float gamma_n, float prio_eps, float alpha,
                     float eta, float ometa) {
    const int a = blockIdx.x, tid = threadIdx.x;
    const int H2 = 2 * H;
    int nsteps = b.counters[2 * a], size = b.counters[2 * a + 1];
    int64_t index = b.index[a];
    const bool dn = done[a] != 0;
    __syncthreads();     // everybody has read the counters before thread 0 rewrites them

    // ---- NStepMemory.add: deques of maxlen n == ring slot (append number mod n)
    {
        const int slot = nsteps % n;
        const size_t e = (size_t)a * n + slot;
        uchar4* dst = reinterpret_cast<uchar4*>(b.ns_state + e * FRAME);
        const float4* src = reinterpret_cast<const float4*>(obs + (size_t)a * FRAME);
        for (int i = tid; i < FRAME / 4; i += blockDim.x) {
            const float4 v = src[i];      // state * 255 in f32, then C truncation (numpy u8 store)
            dst[i] = make_uchar4((unsigned char)__fmul_rn(v.x, 255.f), (unsigned char)__fmul_rn(v.y, 255.f),
                                 (unsigned char)__fmul_rn(v.z, 255.f), (unsigned char)__fmul_rn(v.w, 255.f));
        }
        for (int i = tid; i < H; i += blockDim.x) {
            b.ns_hc[e * H2 + i] = h[(size_t)a * H + i];
            b.ns_hc[e * H2 + H + i] = c[(size_t)a * H + i];
            b.ns_thc[e * H2 + i] = th[(size_t)a * H + i];
            b.ns_thc[e * H2 + H + i] = tc[(size_t)a * H + i];
        }
        if (tid < NA) b.ns_q[e * NA + tid] = q[a * NA + tid];
        if (tid == 0) {
            b.ns_action[e] = (int)action[a];
            b.ns_reward[e] = reward[a];
        }
        ++nsteps;
        size = size + 1 < n ? size + 1 : n;
    }
    __syncthreads();

    // ---- NStepMemory.get + calc_priority + ReplayMemory.add: one per step once n_steps > n,
    //      everything that is left when the episode ends
    const int total = dn ? size : (nsteps > n ? 1 : 0);
    __shared__ float s_reward, s_prio;
    for (int em = 0; em < total; ++em) {
        const int slot = (nsteps - size) % n;            // oldest waiting entry
        const size_t e = (size_t)a * n + slot;
        if (tid == 0) {
            // the reward deque is never popped: the last min(n_steps, n) rewards, oldest first
            const int rsize = nsteps < n ? nsteps : n;
            double sum = 0.0;
            for (int i = 0; i < rsize; ++i)
                sum = __dadd_rn(sum, __dmul_rn(gamma_pow[i], b.ns_reward[(size_t)a * n + (nsteps - rsize + i) % n]));
            const float rf = (float)sum;
            const float pq = b.ns_q[e * NA + b.ns_action[e]];
            float target = rf;
            if (!dn) {
                int na = 0;
                float best = q[a * NA];
                for (int k = 1; k < NA; ++k)
                    if (q[a * NA + k] > best) { best = q[a * NA + k]; na = k; }
                target = __fadd_rn(rf, __fmul_rn(gamma_n, tq[a * NA + na]));
            }
            s_reward = rf;
            s_prio = powf(__fadd_rn(fabsf(__fsub_rn(pq, target)), prio_eps), alpha);
        }
        const int64_t row = (int64_t)a * C + index % C;
        const uint4* fs = reinterpret_cast<const uint4*>(b.ns_state + e * FRAME);
        uint4* fd = reinterpret_cast<uint4*>(b.state + (size_t)row * FRAME);
        for (int i = tid; i < FRAME / 16; i += blockDim.x) fd[i] = fs[i];
        for (int i = tid; i < H2; i += blockDim.x) {
            b.hs_cs[(size_t)row * H2 + i] = b.ns_hc[e * H2 + i];
            b.target_hs_cs[(size_t)row * H2 + i] = b.ns_thc[e * H2 + i];
        }
        __syncthreads();
        if (tid == 0) {
            b.action[row] = (int8_t)b.ns_action[e];
            b.reward[row] = s_reward;
            b.done[row] = dn ? 1.f : 0.f;
            b.stack_count[row] = 1;
            b.priority[row] = s_prio;
        }
        ++index;
        --size;
        __syncthreads();
    }

    // ---- end of episode: Actor.reset -> set_seq_start_index, fresh NStepMemory
    if (dn) {
        const int64_t s0 = b.ep_start[a], last = index;
        const int64_t lim = last - T;
        const int64_t nreg = lim > s0 ? (lim - s0 + overlap - 1) / overlap : 0;
        const float* pr = b.priority + (size_t)a * C;
        for (int64_t k = tid; k <= nreg; k += blockDim.x) {
            const int64_t idx = k < nreg ? s0 + k * overlap : lim;
            if (idx < 0 || idx >= C) continue;
            const int len = (int)(idx + T <= C ? T : C - idx);     // non-wrapping slice
            b.sequence_priority[(size_t)a * C + idx] = seq_priority_value(pr, idx, len, C, false, eta, ometa);
            b.is_seq_start[(size_t)a * C + idx] = 1;
        }
        nsteps = 0;
        size = 0;
        if (tid == 0) b.ep_start[a] = index;
    }
    if (tid == 0) {
        b.counters[2 * a] = nsteps;
        b.counters[2 * a + 1] = size;
        b.index[a] = index;
    }
}

}  // namespace r2d2

using namespace r2d2;

extern "C" int r2d2_actor_observe(const r2d2_actor_buffers* buf, int N, int n, int H, int64_t C, int T,
                                  int overlap, const float* obs, const int64_t* action, const float* q,
                                  const float* h, const float* c, const float* tq, const float* th,
                                  const float* tc, const double* reward, const uint8_t* done,
                                  const double* gamma_pow, float gamma_n, float prio_eps, float alpha,
                                  float eta_f32, float one_minus_eta_f32, void* stream) {
    R2D2_REQUIRE(buf && obs && action && q && h && c && tq && th && tc && reward && done && gamma_pow,
                 "actor_observe: null pointer");
    R2D2_REQUIRE(buf->ns_state && buf->ns_hc && buf->ns_thc && buf->ns_q && buf->ns_action && buf->ns_reward &&
                     buf->counters && buf->state && buf->hs_cs && buf->target_hs_cs && buf->action && buf->reward &&
                     buf->done && buf->stack_count && buf->priority && buf->sequence_priority &&
                     buf->is_seq_start && buf->index && buf->ep_start,
                 "actor_observe: null buffer in r2d2_actor_buffers");
    R2D2_REQUIRE(N > 0 && n > 0 && H > 0 && C > 0 && T > 0 && overlap > 0, "actor_observe: bad sizes");
    R2D2_REQUIRE(C > n, "actor_observe: local ring (C=%lld rows) must hold more than n=%d rows", (long long)C, n);
    R2D2_LAUNCH("actor_observe_kernel", as_stream(stream),
                actor_observe_kernel<<<N, 256, 0, as_stream(stream)>>>(*buf, n, H, C, T, overlap, obs, action, q, h, c, tq,
                                                                       th, tc, reward, done, gamma_pow, gamma_n,
                                                                       prio_eps, alpha, eta_f32, one_minus_eta_f32));
    return R2D2_OK;
}
