// GPU-resident prioritized sequence replay: sum-tree (K1/K8), slab gather (K2), ingest (K9),
// priority scatter and sequence-priority refresh (K8).  Integer/byte work: HBM- or
// latency-bound; no tensor cores.  Reference semantics: replay_memory.py:93-119,175-262.
#include "common.cuh"
#include "priority.cuh"
#include "ptx.cuh"

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

namespace r2d2 {

static thread_local char g_err[512] = "";
char* err_buf() { return g_err; }
int set_err(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
// ---- launch counter and optional per-kernel CUDA-event profiler (bench.py's roofline leg)
static std::atomic<long long> g_launches{0};
static bool g_prof_on = false;
struct ProfRec { const char* name; cudaEvent_t a, b; cudaStream_t st; };
static std::vector<ProfRec> g_recs;
static std::vector<cudaEvent_t> g_pool;
static std::mutex g_prof_mu;

static cudaEvent_t prof_event() {
    if (!g_pool.empty()) { cudaEvent_t e = g_pool.back(); g_pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
}
ProfScope::ProfScope(const char* name, cudaStream_t st) : name_(name), st_(st), on_(g_prof_on) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (on_) {
        std::lock_guard<std::mutex> lk(g_prof_mu);
        a_ = prof_event();
        b_ = prof_event();
        cudaEventRecord((cudaEvent_t)a_, st_);
    }
}
ProfScope::~ProfScope() {
    if (on_) {
        cudaEventRecord((cudaEvent_t)b_, st_);
        std::lock_guard<std::mutex> lk(g_prof_mu);
        g_recs.push_back(ProfRec{name_, (cudaEvent_t)a_, (cudaEvent_t)b_, st_});
    }
}

int sm_count() {
    int dev = 0, n = 148;
    if (cudaGetDevice(&dev) == cudaSuccess)
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    return n;
}

// ======================================================================== sum-tree
constexpr int FAN = R2D2_TREE_FAN;
constexpr int MAX_LEVELS = 8;

struct TreeLayout {
    int n_levels;
    int64_t size[MAX_LEVELS];
    int64_t off[MAX_LEVELS];
    int64_t total;
};

static TreeLayout make_layout(int64_t M) {
    TreeLayout L{};
    int64_t n = M, off = 0;
    int l = 0;
    while (l < MAX_LEVELS) {
        n = (n + FAN - 1) / FAN;
        L.size[l] = n;
        L.off[l] = off;
        off += n;
        ++l;
        if (n <= FAN) break;
    }
    L.n_levels = l;
    L.total = off;
    return L;
}

// Inclusive Hillis-Steele scan across the warp, f64.  The association (own + lower, d =
// 1,2,4,8,16) is part of the parity contract: oracle/replay_oracle.py:_hs_scan mirrors it.
__device__ __forceinline__ double hs_scan(double v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        double o = __shfl_up_sync(FULL, v, d);
        if (lane >= d) v = __dadd_rn(v, o);
    }
    return v;
}

__device__ __forceinline__ double leaf_value(const float* __restrict__ sp,
                                             const uint8_t* __restrict__ st, int64_t slot,
                                             int64_t M) {
    if (slot >= M) return 0.0;
    return st[slot] == 1 ? (double)sp[slot] : 0.0;
}

// one warp per parent node of level `lvl` (0 = parents of leaves)
__global__ void tree_build_level(const float* __restrict__ sp, const uint8_t* __restrict__ st,
                                 int64_t M, double* __restrict__ nodes, int lvl,
                                 int64_t child_off, int64_t child_size, int64_t out_off,
                                 int64_t out_size) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t j = warp; j < out_size; j += nwarps) {
        const int64_t c = j * FAN + lane;
        double v;
        if (lvl == 0) v = leaf_value(sp, st, c, M);
        else v = c < child_size ? nodes[child_off + c] : 0.0;
        v = hs_scan(v, lane);
        if (lane == 31) nodes[out_off + j] = v;
    }
}

// Descent rule (shared with oracle SumTree._pick): first lane with value>0 and inclusive
// prefix >= target; else the last lane with value>0; lane 0 if the node is empty.
__device__ __forceinline__ int pick_child(double v, double pre, double& target, int lane) {
    unsigned hit = __ballot_sync(FULL, v > 0.0 && pre >= target);
    int sel;
    if (hit) sel = __ffs(hit) - 1;
    else {
        unsigned pos = __ballot_sync(FULL, v > 0.0);
        sel = pos ? 31 - __clz(pos) : 0;
    }
    double excl = __shfl_sync(FULL, pre, sel > 0 ? sel - 1 : 0);
    if (sel == 0) excl = 0.0;
    target = __dsub_rn(target, excl);
    return sel;
}

__global__ void tree_sample_kernel(const float* __restrict__ sp, const uint8_t* __restrict__ st,
                                   int64_t M, const double* __restrict__ nodes, TreeLayout L,
                                   const double* __restrict__ u, int B,
                                   int64_t* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int draw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (draw >= B) return;
    const int top = L.n_levels - 1;
    double v = lane < L.size[top] ? nodes[L.off[top] + lane] : 0.0;
    double pre = hs_scan(v, lane);
    const double root = __shfl_sync(FULL, pre, 31);
    double target = __dmul_rn(u[draw], root);
    int64_t node = pick_child(v, pre, target, lane);
    for (int lvl = top; lvl >= 0; --lvl) {
        const int64_t c = node * FAN + lane;
        if (lvl == 0) v = leaf_value(sp, st, c, M);
        else v = c < L.size[lvl - 1] ? nodes[L.off[lvl - 1] + c] : 0.0;
        pre = hs_scan(v, lane);
        node = node * FAN + pick_child(v, pre, target, lane);
    }
    if (lane == 0) out[draw] = node;
}

// single CTA; level-synchronous recompute of the ancestors of `n` leaves
__global__ void tree_update_kernel(const float* __restrict__ sp, const uint8_t* __restrict__ st,
                                   int64_t M, double* __restrict__ nodes, TreeLayout L,
                                   const int64_t* __restrict__ slots, int n) {
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    for (int lvl = 0; lvl < L.n_levels; ++lvl) {
        for (int i = warp; i < n; i += nwarps) {
            int64_t s = slots[i];
            if (s < 0 || s >= M) continue;  // warp-uniform
            int64_t node = s;
            for (int k = 0; k <= lvl; ++k) node /= FAN;
            const int64_t c = node * FAN + lane;
            double v;
            if (lvl == 0) v = leaf_value(sp, st, c, M);
            else v = c < L.size[lvl - 1] ? nodes[L.off[lvl - 1] + c] : 0.0;
            v = hs_scan(v, lane);
            if (lane == 31) nodes[L.off[lvl] + node] = v;
        }
        __threadfence_block();
        __syncthreads();
    }
}

// one level of the same refresh, grid-wide: one warp per touched leaf recomputes that leaf's
// ancestor at level `lvl` from its 32 children (duplicates recompute the same value).  Levels
// are separate launches (level l+1 reads what level l wrote).  Used for large batches of
// touched leaves, where a single CTA would serialise thousands of warp-scans.
__global__ void tree_update_level_kernel(const float* __restrict__ sp, const uint8_t* __restrict__ st,
                                         int64_t M, double* __restrict__ nodes, TreeLayout L,
                                         const int64_t* __restrict__ slots, int n, int lvl) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int i = warp; i < n; i += nwarps) {
        int64_t s = slots[i];
        if (s < 0 || s >= M) continue;  // warp-uniform
        int64_t node = s;
        for (int k = 0; k <= lvl; ++k) node /= FAN;
        const int64_t c = node * FAN + lane;
        double v;
        if (lvl == 0) v = leaf_value(sp, st, c, M);
        else v = c < L.size[lvl - 1] ? nodes[L.off[lvl - 1] + c] : 0.0;
        v = hs_scan(v, lane);
        if (lane == 31) nodes[L.off[lvl] + node] = v;
    }
}

// ======================================================================== gather
constexpr int FRAME_U4 = R2D2_FRAME_BYTES / 16;  // 1764 x 16 B per transition row

// one CTA per (sequence, row): a 28 224-byte contiguous row copy, 16 B per lane per access
__global__ void __launch_bounds__(256)
gather_rows_kernel(const uint4* __restrict__ state, int64_t M,
                   const int64_t* __restrict__ seq_index, int S, uint4* __restrict__ slab) {
    const int b = blockIdx.y, s = blockIdx.x;
    const int64_t src = (seq_index[b] + s) % M;
    const uint4* in = state + src * FRAME_U4;
    uint4* out = slab + ((int64_t)b * S + s) * FRAME_U4;
    uint4 v[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        int i = threadIdx.x + j * 256;
        if (i < FRAME_U4) v[j] = ld_stream(in + i);
    }
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        int i = threadIdx.x + j * 256;
        if (i < FRAME_U4) st_stream(out + i, v[j]);
    }
}

// ---- TMA (bulk async copy) variant: global -> smem -> global, one elected thread drives a
// 4-deep ring of 28 224-byte row buffers per CTA; persistent grid.
constexpr int TMA_STAGES = 4;

__global__ void __launch_bounds__(32)
gather_rows_tma_kernel(const uint8_t* __restrict__ state, int64_t M,
                       const int64_t* __restrict__ seq_index, int B, int S,
                       uint8_t* __restrict__ slab) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t full[TMA_STAGES];
    if (threadIdx.x != 0) return;
    for (int i = 0; i < TMA_STAGES; ++i) mbar_init(&full[i], 1);
    mbar_fence_init();
    const int64_t total = (int64_t)B * S;
    const int64_t per = (total + gridDim.x - 1) / gridDim.x;
    const int64_t lo = blockIdx.x * per;
    const int64_t hi = lo + per < total ? lo + per : total;
    const int64_t cnt = hi > lo ? hi - lo : 0;
    // prologue: fill the ring
    for (int64_t i = 0; i < cnt && i < TMA_STAGES; ++i) {
        const int64_t r = lo + i;
        const int64_t src = (seq_index[r / S] + (r % S)) % M;
        mbar_expect_tx(&full[i], R2D2_FRAME_BYTES);
        bulk_g2s(smem + i * R2D2_FRAME_BYTES, state + src * R2D2_FRAME_BYTES, R2D2_FRAME_BYTES,
                 &full[i]);
    }
    for (int64_t i = 0; i < cnt; ++i) {
        const int st = (int)(i % TMA_STAGES);
        const uint32_t parity = (uint32_t)((i / TMA_STAGES) & 1);
        mbar_wait(&full[st], parity);
        bulk_s2g(slab + (lo + i) * R2D2_FRAME_BYTES, smem + st * R2D2_FRAME_BYTES,
                 R2D2_FRAME_BYTES);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        const int64_t nx = i + TMA_STAGES;
        if (nx < cnt) {
            // the buffer may be refilled once its store has finished READING smem
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            const int64_t r = lo + nx;
            const int64_t src = (seq_index[r / S] + (r % S)) % M;
            mbar_expect_tx(&full[st], R2D2_FRAME_BYTES);
            bulk_g2s(smem + st * R2D2_FRAME_BYTES, state + src * R2D2_FRAME_BYTES,
                     R2D2_FRAME_BYTES, &full[st]);
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// hidden rows, action/reward/done over the learn window, index matrix
__global__ void gather_small_kernel(const float* __restrict__ hs_cs,
                                    const float* __restrict__ ths_cs,
                                    const int8_t* __restrict__ action,
                                    const float* __restrict__ reward,
                                    const float* __restrict__ done, int64_t M, int cell,
                                    const int64_t* __restrict__ seq_index, int B, int T,
                                    int n_step, int burn_in, float* __restrict__ hs,
                                    float* __restrict__ cs, float* __restrict__ ths,
                                    float* __restrict__ tcs, int64_t* __restrict__ action_out,
                                    float* __restrict__ reward_out, float* __restrict__ done_out,
                                    int64_t* __restrict__ index_out) {
    const int L = T - burn_in;
    const int64_t n_hidden = (int64_t)B * cell;
    const int64_t n_learn = (int64_t)L * B;
    const int64_t n_index = (int64_t)T * B;
    const int64_t total = n_hidden + n_learn + n_index;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        if (i < n_hidden) {
            const int b = (int)(i / cell), j = (int)(i % cell);
            const int64_t s0 = seq_index[b], s1 = (s0 + n_step) % M;
            hs[i] = hs_cs[s0 * 2 * cell + j];
            cs[i] = hs_cs[s0 * 2 * cell + cell + j];
            ths[i] = ths_cs[s1 * 2 * cell + j];
            tcs[i] = ths_cs[s1 * 2 * cell + cell + j];
        } else if (i < n_hidden + n_learn) {
            const int64_t k = i - n_hidden;
            const int t = (int)(k / B), b = (int)(k % B);
            const int64_t slot = (seq_index[b] + burn_in + t) % M;
            action_out[k] = (int64_t)action[slot];
            reward_out[k] = reward[slot];
            done_out[k] = done[slot];
        } else {
            const int64_t k = i - n_hidden - n_learn;
            const int s = (int)(k / B), b = (int)(k % B);
            index_out[k] = (seq_index[b] + s) % M;
        }
    }
}

// slab u8 [B,S,frame] -> f32/255 [T,B,frame] starting at frame offset s0
__global__ void slab_to_f32_kernel(const uint8_t* __restrict__ slab, int B, int S, int s0,
                                   int T, float* __restrict__ out) {
    const int64_t row = blockIdx.x;  // t*B + b
    const int t = (int)(row / B), b = (int)(row % B);
    const uchar4* in =
        reinterpret_cast<const uchar4*>(slab + ((int64_t)b * S + s0 + t) * R2D2_FRAME_BYTES);
    float4* o = reinterpret_cast<float4*>(out + row * R2D2_FRAME_BYTES);
    for (int i = threadIdx.x; i < R2D2_FRAME_BYTES / 4; i += blockDim.x) {
        uchar4 v = in[i];
        o[i] = make_float4(__fdiv_rn((float)v.x, 255.f), __fdiv_rn((float)v.y, 255.f),
                           __fdiv_rn((float)v.z, 255.f), __fdiv_rn((float)v.w, 255.f));
    }
}

// ======================================================================== ingest
__global__ void __launch_bounds__(256)
ingest_rows_kernel(const uint4* __restrict__ src, uint4* __restrict__ state, int64_t M,
                   int64_t write_index) {
    const int64_t i = blockIdx.x;
    const int64_t slot = (write_index + i) % M;
    const uint4* in = src + i * FRAME_U4;
    uint4* out = state + slot * FRAME_U4;
    uint4 v[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        int k = threadIdx.x + j * 256;
        if (k < FRAME_U4) v[j] = ld_stream(in + k);
    }
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        int k = threadIdx.x + j * 256;
        if (k < FRAME_U4) st_stream(out + k, v[j]);
    }
}

__global__ void ingest_small_kernel(const float* __restrict__ s_hs, const float* __restrict__ s_ths,
                                    const int8_t* __restrict__ s_action,
                                    const float* __restrict__ s_reward,
                                    const float* __restrict__ s_done,
                                    const int8_t* __restrict__ s_stack,
                                    const float* __restrict__ s_prio,
                                    const float* __restrict__ s_sp,
                                    const uint8_t* __restrict__ s_start, int64_t rows,
                                    float* hs_cs, float* ths_cs, int8_t* action, float* reward,
                                    float* done, int8_t* stack, float* prio, float* sp,
                                    uint8_t* start, int64_t M, int cell, int64_t write_index) {
    const int w = 2 * cell;
    const int64_t total = rows * w;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / w;
        const int j = (int)(i % w);
        const int64_t slot = (write_index + r) % M;
        hs_cs[slot * w + j] = s_hs[i];
        ths_cs[slot * w + j] = s_ths[i];
        if (j == 0) {
            action[slot] = s_action[r];
            reward[slot] = s_reward[r];
            done[slot] = s_done[r];
            stack[slot] = s_stack[r];
            prio[slot] = s_prio[r];
            sp[slot] = s_sp[r];
            start[slot] = s_start[r];
        }
    }
}

// ======================================================================== priorities
// generic scatter with numpy's last-occurrence-wins on duplicates
__global__ void update_priority_kernel(float* __restrict__ priority, int64_t M,
                                       const int64_t* __restrict__ index,
                                       const float* __restrict__ p, int64_t n) {
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x) {
        int64_t s = index[k];
        if (s < 0) s += M;  // numpy negative index
        if (s < 0 || s >= M) continue;
        bool later = false;
        for (int64_t k2 = k + 1; k2 < n && !later; ++k2) {
            int64_t s2 = index[k2];
            if (s2 < 0) s2 += M;
            later = (s2 == s);
        }
        if (!later) priority[s] = p[k];
    }
}

// learner form: flat position k = t*B+b writes slot (seq_index[b]+burn_in+t)%M
__global__ void update_priority_seq_kernel(float* __restrict__ priority, int64_t M,
                                           const int64_t* __restrict__ seq_index, int B,
                                           int burn_in, int L, const float* __restrict__ p) {
    extern __shared__ int64_t s_seq[];
    for (int i = threadIdx.x; i < B; i += blockDim.x) s_seq[i] = seq_index[i];
    __syncthreads();
    const int64_t n = (int64_t)L * B;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x) {
        const int t = (int)(k / B), b = (int)(k % B);
        const int64_t slot = (s_seq[b] + burn_in + t) % M;
        bool later = false;
        const int64_t base = slot - burn_in % M;        // in (-M, M)
        for (int b2 = 0; b2 < B && !later; ++b2) {
            // smallest t2 >= 0 with (seq[b2]+burn_in+t2) % M == slot; further solutions are
            // t2 + j*M, all checked while < L (only matters when M < L).  All terms are in
            // [0, M), so the modulo is two conditional adds (a 64-bit % per pair made this
            // kernel take milliseconds at B = 4096).
            int64_t t2 = base - s_seq[b2];               // in (-2M, M)
            if (t2 < 0) t2 += M;
            if (t2 < 0) t2 += M;
            for (; t2 < L; t2 += M)
                if (t2 * B + b2 > k) { later = true; break; }
        }
        if (!later) priority[slot] = p[k];
    }
}

// Last-writer-wins in O(n) without scratch, for large batches (the scans above are O(n*B)):
//   claim : atomicMax(bits(priority[slot]), OWNER | k)   OWNER = 0xFFC00000 (quiet-NaN space:
//           above the bit pattern of every finite float and of +-inf, so the largest flat
//           position k among the writers of a slot survives, whatever the slot held before)
//   commit: atomicCAS(bits(priority[slot]), OWNER | k, bits(p[k]))  -- only the owner's succeeds
// Two launches; max is order-independent, so the result is deterministic.  k < 2^22.
constexpr unsigned OWNER_TAG = 0xFFC00000u;
constexpr int64_t OWNER_MAX_N = 1 << 22;

__device__ __forceinline__ int64_t seq_slot(const int64_t* __restrict__ seq_index, int B, int burn_in,
                                            int64_t M, int64_t k) {
    const int t = (int)(k / B), b = (int)(k % B);
    return (seq_index[b] + burn_in + t) % M;
}

// seq_index != nullptr: learner form (slot from the sampled starts); else generic index list
__global__ void priority_claim_kernel(float* __restrict__ priority, int64_t M,
                                      const int64_t* __restrict__ seq_index, int B, int burn_in,
                                      const int64_t* __restrict__ index, int64_t n, int64_t k0) {
    unsigned* bits = reinterpret_cast<unsigned*>(priority);
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x) {
        int64_t s;
        if (seq_index) s = seq_slot(seq_index, B, burn_in, M, k0 + k);
        else { s = index[k0 + k]; if (s < 0) s += M; if (s < 0 || s >= M) continue; }
        atomicMax(bits + s, OWNER_TAG | (unsigned)k);
    }
}

__global__ void priority_commit_kernel(float* __restrict__ priority, int64_t M,
                                       const int64_t* __restrict__ seq_index, int B, int burn_in,
                                       const int64_t* __restrict__ index, const float* __restrict__ p,
                                       int64_t n, int64_t k0) {
    unsigned* bits = reinterpret_cast<unsigned*>(priority);
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n;
         k += (int64_t)gridDim.x * blockDim.x) {
        int64_t s;
        if (seq_index) s = seq_slot(seq_index, B, burn_in, M, k0 + k);
        else { s = index[k0 + k]; if (s < 0) s += M; if (s < 0 || s >= M) continue; }
        atomicCAS(bits + s, OWNER_TAG | (unsigned)k, __float_as_uint(p[k0 + k]));
    }
}

// Single CTA.  op = 3*i + j: j=0 own slot (non-wrapping, truncated window), j=1 "previous",
// j=2 "next" (both write slot idx-i from a wrapping window) -- replay_memory.py:184-213.
// All values depend only on `priority`/`is_start`, which this kernel does not modify, so the
// sequential loop of the reference reduces to "last op targeting a slot wins"; the two
// formulas only differ for slots within T of the array end, so only those need the check.
__global__ void update_seq_priority_kernel(const float* __restrict__ pr, float* __restrict__ sp,
                                           const uint8_t* __restrict__ st, int64_t M,
                                           const int64_t* __restrict__ seq_index, int n, int T,
                                           float eta, float ometa, int flag,
                                           int64_t* __restrict__ touched) {
    const int nops = 3 * n;
    // phase 1: target slot (hit may be negative: numpy wraps it) or -1
    for (int op = threadIdx.x; op < nops; op += blockDim.x) {
        const int i = op / 3, j = op % 3;
        const int64_t idx = seq_index[i];
        int64_t hit = -1;
        bool valid = false;
        if (j == 0) { hit = idx; valid = true; }
        else if (flag) {
            const int dir = j == 1 ? -1 : 1;
            for (int k = 1; k <= T; ++k) {
                if (st[mod_pos(idx + (int64_t)dir * k, M)] == 1) { hit = idx - k; valid = true; break; }
            }
        }
        touched[op] = valid ? mod_pos(hit, M) : -1;
    }
    __threadfence_block();
    __syncthreads();
    // phase 2: value + write unless a later op owns the slot
    for (int op = threadIdx.x; op < nops; op += blockDim.x) {
        const int64_t slot = touched[op];
        if (slot < 0) continue;
        const int i = op / 3, j = op % 3;
        if (slot + T > M) {
            bool later = false;
            for (int o2 = op + 1; o2 < nops && !later; ++o2) later = (touched[o2] == slot);
            if (later) continue;
        }
        float v;
        if (j == 0) {
            const int64_t idx = seq_index[i];
            const int len = (int)((idx + T <= M) ? T : (M - idx));
            v = seq_priority_value(pr, idx, len, M, false, eta, ometa);
        } else {
            // wrapping window starting at the (possibly negative) hit == slot mod M
            v = seq_priority_value(pr, slot, T, M, true, eta, ometa);
        }
        sp[slot] = v;
    }
}

// ---- the same refresh, grid-wide (large batches): phase 1 = one thread per op, phase 2 = one
// warp per op.  Lanes 0..7 carry numpy's eight strided partial sums, all lanes share the max.
__device__ float seq_priority_value_warp(const float* __restrict__ pr, int64_t start, int n,
                                         int64_t M, bool wrap, float eta, float ometa, int lane) {
    if (n < 8 || n > 128) {
        float v = 0.f;
        if (lane == 0) v = seq_priority_value(pr, start, n, M, wrap, eta, ometa);
        return __shfl_sync(FULL, v, 0);
    }
    auto at = [&](int i) -> float {
        int64_t s = start + i;
        if (wrap) s = mod_pos(s, M);
        return pr[s];
    };
    float mx = -INFINITY;
    for (int i = lane; i < n; i += 32) mx = fmaxf(mx, at(i));
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, d));
    const int nfull = n - (n % 8);
    float r = 0.f;
    if (lane < 8) {
        r = at(lane);
        for (int i = 8 + lane; i < nfull; i += 8) r = __fadd_rn(r, at(i));
    }
    float s1 = __fadd_rn(r, __shfl_down_sync(FULL, r, 1));      // lanes 0,2,4,6: r0+r1, r2+r3, ...
    float s2 = __fadd_rn(s1, __shfl_down_sync(FULL, s1, 2));    // lanes 0,4
    float res = __fadd_rn(s2, __shfl_down_sync(FULL, s2, 4));   // lane 0
    float v = 0.f;
    if (lane == 0) {
        for (int i = nfull; i < n; ++i) res = __fadd_rn(res, at(i));
        const float mean = __fdiv_rn(res, (float)n);
        v = __fadd_rn(__fmul_rn(eta, mx), __fmul_rn(ometa, mean));
    }
    return __shfl_sync(FULL, v, 0);
}

__global__ void seq_priority_targets_kernel(const uint8_t* __restrict__ st, int64_t M,
                                            const int64_t* __restrict__ seq_index, int n, int T,
                                            int flag, int64_t* __restrict__ touched) {
    const int nops = 3 * n;
    for (int op = blockIdx.x * blockDim.x + threadIdx.x; op < nops; op += gridDim.x * blockDim.x) {
        const int i = op / 3, j = op % 3;
        const int64_t idx = seq_index[i];
        int64_t hit = -1;
        bool valid = false;
        if (j == 0) { hit = idx; valid = true; }
        else if (flag) {
            const int dir = j == 1 ? -1 : 1;
            for (int k = 1; k <= T; ++k) {
                if (st[mod_pos(idx + (int64_t)dir * k, M)] == 1) { hit = idx - k; valid = true; break; }
            }
        }
        touched[op] = valid ? mod_pos(hit, M) : -1;
    }
}

__global__ void seq_priority_values_kernel(const float* __restrict__ pr, float* __restrict__ sp,
                                           int64_t M, const int64_t* __restrict__ seq_index, int n,
                                           int T, float eta, float ometa,
                                           const int64_t* __restrict__ touched) {
    const int nops = 3 * n;
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int op = warp; op < nops; op += nwarps) {
        const int64_t slot = touched[op];
        if (slot < 0) continue;                       // warp-uniform
        const int i = op / 3, j = op % 3;
        if (slot + T > M) {                           // formulas differ only near the array end
            bool later = false;
            for (int o2 = op + 1 + lane; o2 < nops; o2 += 32) later |= (touched[o2] == slot);
            if (__any_sync(FULL, later)) continue;
        }
        float v;
        if (j == 0) {
            const int64_t idx = seq_index[i];
            const int len = (int)((idx + T <= M) ? T : (M - idx));
            v = seq_priority_value_warp(pr, idx, len, M, false, eta, ometa, lane);
        } else {
            v = seq_priority_value_warp(pr, slot, T, M, true, eta, ometa, lane);
        }
        if (lane == 0) sp[slot] = v;
    }
}

}  // namespace r2d2

using namespace r2d2;

// ======================================================================== C ABI
extern "C" {

int r2d2_version(void) { return 100; }
const char* r2d2_last_error(void) { return r2d2::err_buf(); }
int r2d2_sm_count(void) { return r2d2::sm_count(); }
long long r2d2_launch_count(void) { return r2d2::g_launches.load(); }
int r2d2_prof_enable(int on) {
    r2d2::g_prof_on = on != 0;
    return R2D2_OK;
}
int r2d2_prof_report(char* buf, int buflen, int reset) {
    // synchronises, then writes one "name count total_ms" line per kernel name
    R2D2_CUDA_OK(cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lk(r2d2::g_prof_mu);
    std::map<std::string, std::pair<long long, double>> agg;
    for (auto& r : r2d2::g_recs) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
            auto& e = agg[r.name];
            e.first += 1;
            e.second += ms;
        }
    }
    int pos = 0;
    if (buf && buflen > 0) buf[0] = 0;
    for (auto& kv : agg) {
        int w = snprintf(buf + pos, pos < buflen ? buflen - pos : 0, "%s %lld %.6f\n",
                         kv.first.c_str(), kv.second.first, kv.second.second);
        if (w < 0 || pos + w >= buflen) break;
        pos += w;
    }
    if (reset) {
        for (auto& r : r2d2::g_recs) { r2d2::g_pool.push_back(r.a); r2d2::g_pool.push_back(r.b); }
        r2d2::g_recs.clear();
    }
    return R2D2_OK;
}

int r2d2_prof_trace(char* buf, int buflen) {
    // synchronises, then writes one "name stream_id start_ms end_ms" line per recorded launch, in
    // launch order, times relative to the first recorded launch (events of different streams
    // share one device clock)
    R2D2_CUDA_OK(cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lk(r2d2::g_prof_mu);
    int pos = 0;
    if (buf && buflen > 0) buf[0] = 0;
    std::map<cudaStream_t, int> ids;
    for (auto& r : r2d2::g_recs) {
        float t0 = 0.f, t1 = 0.f;
        if (cudaEventElapsedTime(&t0, r2d2::g_recs[0].a, r.a) != cudaSuccess ||
            cudaEventElapsedTime(&t1, r2d2::g_recs[0].a, r.b) != cudaSuccess)
            continue;
        int id = ids.emplace(r.st, (int)ids.size()).first->second;
        int w = snprintf(buf + pos, pos < buflen ? buflen - pos : 0, "%s %d %.6f %.6f\n", r.name, id,
                         t0, t1);
        if (w < 0 || pos + w >= buflen) break;
        pos += w;
    }
    return R2D2_OK;
}

int r2d2_tree_layout(int64_t M, int max_levels, int64_t* level_sizes, int64_t* level_offsets,
                     int* n_levels, int64_t* total_nodes) {
    R2D2_REQUIRE(M > 0, "tree_layout: M must be positive");
    TreeLayout L = make_layout(M);
    R2D2_REQUIRE(L.size[L.n_levels - 1] <= FAN, "tree_layout: M=%lld needs more than %d levels",
                 (long long)M, MAX_LEVELS);
    R2D2_REQUIRE(max_levels >= L.n_levels, "tree_layout: need room for %d levels", L.n_levels);
    for (int i = 0; i < L.n_levels; ++i) {
        if (level_sizes) level_sizes[i] = L.size[i];
        if (level_offsets) level_offsets[i] = L.off[i];
    }
    if (n_levels) *n_levels = L.n_levels;
    if (total_nodes) *total_nodes = L.total;
    return R2D2_OK;
}

int r2d2_tree_build(const float* seq_prio, const uint8_t* is_start, int64_t M, double* nodes,
                    void* stream) {
    R2D2_REQUIRE(seq_prio && is_start && nodes && M > 0, "tree_build: null pointer or M<=0");
    TreeLayout L = make_layout(M);
    R2D2_REQUIRE(L.size[L.n_levels - 1] <= FAN, "tree_build: M too large");
    const int sms = sm_count();
    for (int l = 0; l < L.n_levels; ++l) {
        const int64_t warps = L.size[l];
        int64_t blocks = (warps + 7) / 8;  // 8 warps per CTA
        if (blocks > (int64_t)sms * 16) blocks = (int64_t)sms * 16;
        R2D2_LAUNCH("tree_build_level", as_stream(stream), tree_build_level<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(
            seq_prio, is_start, M, nodes, l, l ? L.off[l - 1] : 0, l ? L.size[l - 1] : M, L.off[l],
            L.size[l]));
    }
    return R2D2_OK;
}

int r2d2_tree_sample(const float* seq_prio, const uint8_t* is_start, int64_t M,
                     const double* nodes, const double* u, int B, int64_t* out_slot,
                     void* stream) {
    R2D2_REQUIRE(seq_prio && is_start && nodes && u && out_slot, "tree_sample: null pointer");
    R2D2_REQUIRE(M > 0 && B >= 0, "tree_sample: bad sizes");
    if (B == 0) return R2D2_OK;
    TreeLayout L = make_layout(M);
    const int warps_per_cta = 4;
    R2D2_LAUNCH("tree_sample_kernel", as_stream(stream), tree_sample_kernel<<<(B + warps_per_cta - 1) / warps_per_cta, warps_per_cta * 32, 0,
                         as_stream(stream)>>>(seq_prio, is_start, M, nodes, L, u, B, out_slot));
    return R2D2_OK;
}

int r2d2_tree_update(const float* seq_prio, const uint8_t* is_start, int64_t M, double* nodes,
                     const int64_t* slots, int n, void* stream) {
    R2D2_REQUIRE(seq_prio && is_start && nodes && slots, "tree_update: null pointer");
    if (n <= 0) return R2D2_OK;
    TreeLayout L = make_layout(M);
    if (n <= 96) {      // one launch, levels synchronised inside the CTA
        R2D2_LAUNCH("tree_update_kernel", as_stream(stream), tree_update_kernel<<<1, 1024, 0, as_stream(stream)>>>(seq_prio, is_start, M, nodes, L, slots,
                                                               n));
        return R2D2_OK;
    }
    int blocks = (n + 7) / 8;                          // 8 warps per CTA, one warp per touched leaf
    blocks = blocks > sm_count() * 8 ? sm_count() * 8 : blocks;
    for (int l = 0; l < L.n_levels; ++l)
        R2D2_LAUNCH("tree_update_level_kernel", as_stream(stream),
                    tree_update_level_kernel<<<blocks, 256, 0, as_stream(stream)>>>(seq_prio, is_start, M, nodes, L, slots, n, l));
    return R2D2_OK;
}

int r2d2_replay_gather(const uint8_t* state, const float* hs_cs, const float* target_hs_cs,
                       const int8_t* action, const float* reward, const float* done, int64_t M,
                       int cell, const int64_t* seq_index, int B, int T, int n_step,
                       int burn_in, uint8_t* slab, float* hs, float* cs, float* ths, float* tcs,
                       int64_t* action_out, float* reward_out, float* done_out,
                       int64_t* index_out, int use_tma, void* stream) {
    R2D2_REQUIRE(state && hs_cs && target_hs_cs && action && reward && done && seq_index &&
                     hs && cs && ths && tcs && action_out && reward_out && done_out && index_out,
                 "replay_gather: null pointer");
    R2D2_REQUIRE(B > 0 && T > 0 && n_step >= 0 && burn_in >= 0 && burn_in < T && cell > 0 && M > 0,
                 "replay_gather: bad sizes B=%d T=%d n=%d burn_in=%d", B, T, n_step, burn_in);
    R2D2_REQUIRE(((uintptr_t)state % 16) == 0 && ((uintptr_t)slab % 16) == 0,
                 "replay_gather: state/slab must be 16-byte aligned");
    const int S = T + n_step;
    if (!slab) {
        // frames stay in the ring: the consumer reads them in place through seq_index
        // (r2d2_tc_conv1_fwd_dual_ring / r2d2_tc_conv1_wgrad_ring)
    } else if (use_tma) {
        const int smem = TMA_STAGES * R2D2_FRAME_BYTES;
        R2D2_CUDA_OK(cudaFuncSetAttribute(gather_rows_tma_kernel,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        int64_t total = (int64_t)B * S;
        int grid = sm_count();
        if (grid > total) grid = (int)total;
        R2D2_LAUNCH("gather_rows_tma_kernel", as_stream(stream),
                    gather_rows_tma_kernel<<<grid, 32, smem, as_stream(stream)>>>(state, M, seq_index,
                                                                                 B, S, slab));
    } else {
        R2D2_LAUNCH("gather_rows_kernel", as_stream(stream),
                    gather_rows_kernel<<<dim3(S, B), 256, 0, as_stream(stream)>>>(
                        reinterpret_cast<const uint4*>(state), M, seq_index, S,
                        reinterpret_cast<uint4*>(slab)));
    }
    const int64_t total = (int64_t)B * cell + (int64_t)(T - burn_in) * B + (int64_t)T * B;
    R2D2_LAUNCH("gather_small_kernel", as_stream(stream),
                gather_small_kernel<<<(unsigned)((total + 255) / 256), 256, 0, as_stream(stream)>>>(
                    hs_cs, target_hs_cs, action, reward, done, M, cell, seq_index, B, T, n_step,
                    burn_in, hs, cs, ths, tcs, action_out, reward_out, done_out, index_out));
    return R2D2_OK;
}

int r2d2_slab_to_f32(const uint8_t* slab, int B, int S, int s0, int T, float* out, void* stream) {
    R2D2_REQUIRE(slab && out && B > 0 && T > 0 && s0 >= 0 && s0 + T <= S, "slab_to_f32: bad args");
    R2D2_LAUNCH("slab_to_f32_kernel", as_stream(stream), slab_to_f32_kernel<<<(unsigned)((int64_t)T * B), 256, 0, as_stream(stream)>>>(slab, B, S, s0,
                                                                                   T, out));
    return R2D2_OK;
}

int r2d2_replay_ingest(const uint8_t* src_state, const float* src_hs_cs,
                       const float* src_target_hs_cs, const int8_t* src_action,
                       const float* src_reward, const float* src_done,
                       const int8_t* src_stack_count, const float* src_priority,
                       const float* src_seq_prio, const uint8_t* src_is_start, int64_t rows,
                       uint8_t* state, float* hs_cs, float* target_hs_cs, int8_t* action,
                       float* reward, float* done, int8_t* stack_count, float* priority,
                       float* seq_prio, uint8_t* is_start, int64_t M, int cell,
                       int64_t write_index, void* stream) {
    R2D2_REQUIRE(rows >= 0 && rows <= M, "replay_ingest: rows=%lld exceeds ring size %lld",
                 (long long)rows, (long long)M);
    if (rows == 0) return R2D2_OK;
    R2D2_REQUIRE(src_state && src_hs_cs && src_target_hs_cs && src_action && src_reward &&
                     src_done && src_stack_count && src_priority && src_seq_prio && src_is_start &&
                     state && hs_cs && target_hs_cs && action && reward && done && stack_count &&
                     priority && seq_prio && is_start,
                 "replay_ingest: null pointer");
    R2D2_REQUIRE(((uintptr_t)state % 16) == 0 && ((uintptr_t)src_state % 16) == 0,
                 "replay_ingest: state buffers must be 16-byte aligned");
    R2D2_LAUNCH("ingest_rows_kernel", as_stream(stream), ingest_rows_kernel<<<(unsigned)rows, 256, 0, as_stream(stream)>>>(
        reinterpret_cast<const uint4*>(src_state), reinterpret_cast<uint4*>(state), M,
        write_index % M));
    const int64_t total = rows * 2 * cell;
    int64_t blocks = (total + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    R2D2_LAUNCH("ingest_small_kernel", as_stream(stream), ingest_small_kernel<<<(unsigned)blocks, 256, 0, as_stream(stream)>>>(
        src_hs_cs, src_target_hs_cs, src_action, src_reward, src_done, src_stack_count,
        src_priority, src_seq_prio, src_is_start, rows, hs_cs, target_hs_cs, action, reward, done,
        stack_count, priority, seq_prio, is_start, M, cell, write_index % M));
    return R2D2_OK;
}

int r2d2_update_priority(float* priority, int64_t M, const int64_t* index, const float* p,
                         int64_t n, void* stream) {
    R2D2_REQUIRE(priority && index && p && M > 0, "update_priority: bad args");
    if (n <= 0) return R2D2_OK;
    int64_t blocks = (n + 127) / 128;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (n <= 2048) {     // few indices: one launch, each index scans the ones behind it
        R2D2_LAUNCH("update_priority_kernel", as_stream(stream), update_priority_kernel<<<(unsigned)blocks, 128, 0, as_stream(stream)>>>(priority, M, index, p,
                                                                                n));
        return R2D2_OK;
    }
    // many: claim/commit in O(n), in chunks the owner tag can number (later chunks overwrite earlier ones)
    for (int64_t k0 = 0; k0 < n; k0 += OWNER_MAX_N) {
        const int64_t nn = n - k0 < OWNER_MAX_N ? n - k0 : OWNER_MAX_N;
        int64_t bl = (nn + 255) / 256;
        if (bl > 148 * 16) bl = 148 * 16;
        R2D2_LAUNCH("priority_claim_kernel", as_stream(stream),
                    priority_claim_kernel<<<(unsigned)bl, 256, 0, as_stream(stream)>>>(priority, M, nullptr, 0, 0, index, nn, k0));
        R2D2_LAUNCH("priority_commit_kernel", as_stream(stream),
                    priority_commit_kernel<<<(unsigned)bl, 256, 0, as_stream(stream)>>>(priority, M, nullptr, 0, 0, index, p, nn, k0));
    }
    return R2D2_OK;
}

int r2d2_update_priority_seq(float* priority, int64_t M, const int64_t* seq_index, int B,
                             int burn_in, int L, const float* p, void* stream) {
    R2D2_REQUIRE(priority && seq_index && p && M > 0 && B > 0 && L > 0 && burn_in >= 0,
                 "update_priority_seq: bad args");
    const int64_t n = (int64_t)L * B;
    if (B > 32) {       // larger batches: claim/commit in O(n) instead of an O(B) scan per element
        R2D2_REQUIRE(n <= OWNER_MAX_N && M >= L, "update_priority_seq: L*B=%lld exceeds %lld (or M < L)",
                     (long long)n, (long long)OWNER_MAX_N);
        int64_t bl = (n + 255) / 256;
        if (bl > 148 * 16) bl = 148 * 16;
        R2D2_LAUNCH("priority_claim_kernel", as_stream(stream),
                    priority_claim_kernel<<<(unsigned)bl, 256, 0, as_stream(stream)>>>(priority, M, seq_index, B, burn_in, nullptr, n, 0));
        R2D2_LAUNCH("priority_commit_kernel", as_stream(stream),
                    priority_commit_kernel<<<(unsigned)bl, 256, 0, as_stream(stream)>>>(priority, M, seq_index, B, burn_in, nullptr, p, n, 0));
        return R2D2_OK;
    }
    int64_t blocks = (n + 127) / 128;
    if (blocks > 148 * 16) blocks = 148 * 16;
    R2D2_LAUNCH("update_priority_seq_kernel", as_stream(stream), update_priority_seq_kernel<<<(unsigned)blocks, 128, B * sizeof(int64_t), as_stream(stream)>>>(
        priority, M, seq_index, B, burn_in, L, p));
    return R2D2_OK;
}

int r2d2_update_sequence_priority(const float* priority, float* seq_prio,
                                  const uint8_t* is_start, int64_t M, const int64_t* seq_index,
                                  int n, int T, float eta_f32, float one_minus_eta_f32,
                                  int update_pre_next, int64_t* touched, void* stream) {
    R2D2_REQUIRE(priority && seq_prio && is_start && seq_index && touched,
                 "update_sequence_priority: null pointer");
    R2D2_REQUIRE(M > 0 && T > 0 && T <= M, "update_sequence_priority: need 0 < T <= M");
    if (n <= 0) return R2D2_OK;
    if (n <= 32) {      // small batches: one launch, phases synchronised inside the CTA
        R2D2_LAUNCH("update_seq_priority_kernel", as_stream(stream), update_seq_priority_kernel<<<1, 1024, 0, as_stream(stream)>>>(
            priority, seq_prio, is_start, M, seq_index, n, T, eta_f32, one_minus_eta_f32,
            update_pre_next, touched));
        return R2D2_OK;
    }
    const int nops = 3 * n;
    int b1 = (nops + 127) / 128, b2 = (nops + 7) / 8;
    const int cap = sm_count() * 8;
    b1 = b1 > cap ? cap : b1;
    b2 = b2 > cap ? cap : b2;
    R2D2_LAUNCH("seq_priority_targets_kernel", as_stream(stream),
                seq_priority_targets_kernel<<<b1, 128, 0, as_stream(stream)>>>(is_start, M, seq_index, n, T, update_pre_next, touched));
    R2D2_LAUNCH("seq_priority_values_kernel", as_stream(stream),
                seq_priority_values_kernel<<<b2, 256, 0, as_stream(stream)>>>(priority, seq_prio, M, seq_index, n, T, eta_f32,
                                                                              one_minus_eta_f32, touched));
    return R2D2_OK;
}

}  // extern "C"
