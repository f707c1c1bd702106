/* r2d2_b200 -- C ABI of the B200-native R2D2 learner hot path.
 *
 * Drop-in boundary for the hot path of LiXirong/pytorch-r2d2 (reference citations are
 * file:line into that repo).  The reference has no FFI of its own (it is six Python files);
 * the entry points below are what a ctypes binding of its hot path binds, one per step of
 * `Learner.train()` (learner.py:68-104) and of the `ReplayMemory` methods it calls.
 *
 * Conventions
 *   - every function returns 0 on success, a negative R2D2_E* code on failure, never throws;
 *     `r2d2_last_error()` returns a thread-local message for the last failure;
 *   - all tensor arguments are raw DEVICE pointers, borrowed for the duration of the enqueue;
 *     the caller (PyTorch) owns them;
 *   - `stream` is a `cudaStream_t` passed as `void*`; calls are asynchronous on it;
 *   - no entry point allocates device memory; scratch is passed in by the caller and sized
 *     by the matching `*_scratch_bytes` query;
 *   - layouts are row-major; "time-major" means row index t*B + b (model.py:59,64-66).
 */
#ifndef R2D2_B200_H_
#define R2D2_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define R2D2_OK 0
#define R2D2_EINVAL (-1)
#define R2D2_ECUDA (-2)

#define R2D2_TREE_FAN 32
#define R2D2_FRAME_BYTES 28224 /* 4*84*84 u8, replay_memory.py:65,77 */
#define R2D2_FEAT 1568        /* 7*7*32, model.py:24,46 */
#define R2D2_ACTIONS 6        /* model.py:35 */

int r2d2_version(void);
const char* r2d2_last_error(void);
/* number of SMs of the current device (grid sizing for callers / bench) */
int r2d2_sm_count(void);
/* kernels launched by this library since it was loaded (bench.py's gpu_launches) */
long long r2d2_launch_count(void);
/* optional per-kernel CUDA-event timing on the launch stream (bench.py's roofline leg; not for
 * use under stream capture).  r2d2_prof_report synchronises the device and writes one
 * "name count total_ms" line per kernel into buf. */
int r2d2_prof_enable(int on);
int r2d2_prof_report(char* buf, int buflen, int reset);
/* one "name stream_id start_ms end_ms" line per recorded launch, in launch order, relative to the
 * first recorded launch: the timeline of a multi-stream step (tools/timeline.py).  Call before
 * r2d2_prof_report(reset=1). */
int r2d2_prof_trace(char* buf, int buflen);

/* ------------------------------------------------------------------ sum-tree (K1, K8)
 * Replaces the O(M) mask + normalise + torch.multinomial of replay_memory.py:225-232.
 * Leaves are implicit: leaf(slot) = is_seq_start[slot]==1 ? sequence_priority[slot] : 0 (f32).
 * Internal nodes are f64, fan-out 32; level k+1 node j = last element of a Hillis-Steele
 * warp scan over its 32 children.  `nodes` holds all internal levels bottom-up, back to back;
 * r2d2_tree_layout gives sizes/offsets. */
int r2d2_tree_layout(int64_t M, int max_levels, int64_t* level_sizes, int64_t* level_offsets,
                     int* n_levels, int64_t* total_nodes);
int r2d2_tree_build(const float* seq_prio, const uint8_t* is_start, int64_t M, double* nodes,
                    void* stream);
/* out_slot[i] = slot drawn for the f64 uniform u[i] in [0,1): target = u*root, descend by
 * "first child with value>0 and inclusive prefix >= target".  Same draws as multinomial
 * (replay_memory.py:227-232) given the same uniforms; see DESIGN.md for the contract. */
int r2d2_tree_sample(const float* seq_prio, const uint8_t* is_start, int64_t M,
                     const double* nodes, const double* u, int B, int64_t* out_slot,
                     void* stream);
/* recompute the ancestors of `n` leaves (slots < 0 are skipped) */
int r2d2_tree_update(const float* seq_prio, const uint8_t* is_start, int64_t M, double* nodes,
                     const int64_t* slots, int n, void* stream);

/* ------------------------------------------------------------------ slab gather (K2)
 * Replaces the Python gather loop of replay_memory.py:233-262 (get_stacked_state :215-222 in
 * its only used mode, stack_count==1).  One contiguous (T+n)-row u8 slab per sequence:
 *   slab[b, s]      = state[(seq_index[b]+s) % M]             s < T+n   ([B,T+n,4,84,84] u8)
 *   state[s]        = slab[:, s], next_state[s] = slab[:, s+n] (replay_memory.py:236-237)
 *   hs,cs           = hs_cs[seq_index] split at `cell`         (:238-239)
 *   ths,tcs         = target_hs_cs[(seq_index+n)%M] split      (:240-241)
 *   action/reward/done[t,b] at (seq_index[b]+burn_in+t) % M    (:242-244) time-major [L,B]
 *   index[s,b]      = (seq_index[b]+s) % M                     (:245) time-major [T,B]   */
int r2d2_replay_gather(const uint8_t* state, const float* hs_cs, const float* target_hs_cs,
                       const int8_t* action, const float* reward, const float* done, int64_t M,
                       int cell, const int64_t* seq_index, int B, int T, int n_step,
                       int burn_in, uint8_t* slab, float* hs, float* cs, float* ths, float* tcs,
                       int64_t* action_out, float* reward_out, float* done_out,
                       int64_t* index_out, int use_tma, void* stream);
/* compat path for ReplayMemory.sample(): u8 slab -> f32/255 [T,B,4,84,84] at frame offset s0
 * (replay_memory.py:255-256) */
int r2d2_slab_to_f32(const uint8_t* slab, int B, int S, int s0, int T, float* out, void* stream);

/* ------------------------------------------------------------------ ingest (K9)
 * ReplayMemory.extend (replay_memory.py:107-119): copy `rows` staged rows of all ten keys
 * into ring slots (write_index + i) % M.  src_* are device staging buffers. */
int r2d2_replay_ingest(const uint8_t* src_state, const float* src_hs_cs,
                       const float* src_target_hs_cs, const int8_t* src_action,
                       const float* src_reward, const float* src_done,
                       const int8_t* src_stack_count, const float* src_priority,
                       const float* src_seq_prio, const uint8_t* src_is_start, int64_t rows,
                       uint8_t* state, float* hs_cs, float* target_hs_cs, int8_t* action,
                       float* reward, float* done, int8_t* stack_count, float* priority,
                       float* seq_prio, uint8_t* is_start, int64_t M, int cell,
                       int64_t write_index, void* stream);

/* ------------------------------------------------------------------ priorities (K8)
 * update_priority (replay_memory.py:175-176): priority[index[k]] = p[k], last occurrence
 * wins on duplicate slots (numpy fancy assignment order). */
int r2d2_update_priority(float* priority, int64_t M, const int64_t* index, const float* p,
                         int64_t n, void* stream);
/* The learner's fused form (learner.py:103): index = (seq_index[b]+burn_in+t)%M at flat
 * position t*B+b; duplicates resolved to the last flat position without scratch. */
int r2d2_update_priority_seq(float* priority, int64_t M, const int64_t* seq_index, int B,
                             int burn_in, int L, const float* p, void* stream);
/* update_sequence_priority (replay_memory.py:184-213) incl. its neighbour quirks; numpy f32
 * pairwise-sum order for the mean; eta_f32/one_minus_eta_f32 are the f32 roundings of the
 * Python floats.  touched[3*n] receives the slots written (or -1), for r2d2_tree_update. */
int r2d2_update_sequence_priority(const float* priority, float* seq_prio,
                                  const uint8_t* is_start, int64_t M, const int64_t* seq_index,
                                  int n, int T, float eta_f32, float one_minus_eta_f32,
                                  int update_pre_next, int64_t* touched, void* stream);

/* ------------------------------------------------------------------ network, forward
 * conv torso (model.py:13-21,44-46).  Frames come from the u8 slab: output row r = t*B + b
 * for t < nt reads slab[b, s0 + t]; /255 fused (replay_memory.py:256).  a1 [R,32,20,20],
 * a2 [R,32,9,9], feat [R,1568] (C-major flatten). */
int r2d2_torso_fwd_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* w1,
                      const float* b1, const float* w2, const float* b2, const float* w3,
                      const float* b3, float* a1, float* a2, float* feat, void* stream);
/* same for caller-provided f32 frames [R,4,84,84] (QNet.forward's input, model.py:38-44) */
int r2d2_torso_fwd_f32(const float* frames, int R, const float* w1, const float* b1,
                       const float* w2, const float* b2, const float* w3, const float* b3,
                       float* a1, float* a2, float* feat, void* stream);
/* y[M,N] = x[M,K] . w[N,K]^T + bias1 + bias2 (either may be NULL), optional ReLU */
int r2d2_linear_fwd(const float* x, const float* w, const float* bias1, const float* bias2,
                    float* y, int M, int N, int K, int relu, void* stream);
/* LSTM unroll (model.py:52-60): xproj [T,B,4H] already holds x.W_ih^T + b_ih + b_hh.
 * hs_ext/cs_ext are [(T+1),B,H]; slot 0 is the initial state (written by the caller), slot
 * t+1 receives step t.  gates [T,B,4H] (post-activation i,f,g,o) may be NULL (inference). */
int r2d2_lstm_seq_fwd(const float* xproj, const float* w_hh, float* hs_ext, float* cs_ext,
                      float* gates, int T, int B, int H, void* stream);
/* dueling head (model.py:26-36,62-64): w1 [2H,H] = (val.0.weight; adv.0.weight),
 * b1 [2H], w2 [7,H] = (val.2.weight; adv.2.weight), b2 [7].  hid [M,2H] scratch/stash. */
int r2d2_head_fwd(const float* hs, const float* w1, const float* b1, const float* w2,
                  const float* b2, float* hid, float* q, int M, int H, void* stream);

/* second half of r2d2_head_fwd alone (val/adv output layers + dueling combine) for callers that
 * computed hid = relu(hs.w1^T + b1) themselves (bf16 tensor-core path) */
int r2d2_head_out(const float* hid, const float* w2, const float* b2, float* q, int M, int H,
                  void* stream);

/* ------------------------------------------------------------------ target / TD (K7)
 * learner.py:86-102: a* = argmax q_next_online (first max), y = r + gamma_n*Qt[a*]*(1-done)
 * (optionally h(r + gamma_n*h^-1(Qt)*(1-done)), NOT in the reference, default off),
 * delta = q[action]-y, loss = mean(0.5 delta^2), dq = dloss/dq_all, prio=(|delta|+eps)^alpha */
int r2d2_td_target(const float* q_all, const float* q_next_online, const float* q_next_target,
                   const int64_t* action, const float* reward, const float* done, int n,
                   float gamma_n, int rescale, float rescale_eps, float prio_eps, float alpha,
                   const int64_t* next_action_override, int64_t* next_action, float* y,
                   float* delta, float* dq, float* prio, float* loss, void* stream);

/* ------------------------------------------------------------------ network, backward
 * (autograd of learner.py:97-99 through model.py:44-66).  head_bwd: gradient of the dueling
 * output layers w.r.t. the hidden layer (dhid, ReLU' applied) and w.r.t. their weights (dw2, db2);
 * either half may be skipped by passing NULL (dhid, or dw2 and db2 together). */
int r2d2_head_bwd(const float* dq, const float* hid, const float* w2, float* dhid, float* dw2,
                  float* db2, int M, int H, void* stream);
/* dx[M,K] (+)= dy[M,N] . w[N,K] */
int r2d2_linear_dgrad(const float* dy, const float* w, float* dx, int M, int N, int K,
                      int accumulate, void* stream);
/* dw[N,K] = dy[M,N]^T . x[M,K]; db[N] = column sums of dy (db may be NULL).
 * scratch: r2d2_linear_wgrad_scratch_bytes(M,N,K) bytes (split-K partials). */
int64_t r2d2_linear_wgrad_scratch_bytes(int M, int N, int K);
int r2d2_linear_wgrad(const float* dy, const float* x, float* dw, float* db, int M, int N,
                      int K, void* scratch, void* stream);
/* No counterpart in the reference.  Same result as r2d2_linear_wgrad for callers that cannot
 * supply (or allocate) a workspace: a C/C++ actor or serving host that was given exactly its
 * input/output buffers by another component, a process whose whole HBM budget was handed out
 * at start-up, or a latency-critical loop that must never reach cudaMalloc because it stalls
 * every stream on the device.  The split-K partial sums go to an array built into the library
 * (R2D2_STATIC_WGRAD_SCRATCH_BYTES per GPU, resident until exit).  Problems whose scratch
 * need exceeds it are refused with R2D2_EINVAL.  The first call on each device resolves the
 * array's address (do it at set-up, before any stream capture: a first call on a capturing
 * stream is refused); `stream` must belong to the current device (checked); all calls share the
 * array, so calls that may overlap on the GPU must be serialised by the caller.  Never allocates. */
#define R2D2_STATIC_WGRAD_SCRATCH_BYTES (32 << 20)
int r2d2_linear_wgrad_static(const float* dy, const float* x, float* dw, float* db, int M, int N,
                             int K, void* stream);
/* db[N] = column sums of dy[M,N] (bias gradients when the weight gradient runs on tensor cores) */
int r2d2_colsum(const float* dy, float* db, int M, int N, void* stream);
/* same, the sums also written to db_copy (LSTMCell's bias_ih and bias_hh receive one gradient) */
int r2d2_colsum2(const float* dy, float* db, float* db_copy, int M, int N, void* stream);
/* BPTT over the learn window: dhs [T,B,H] is dL/dh_t from the head; gates/cs_ext from
 * r2d2_lstm_seq_fwd.  Writes dgates [T,B,4H] (pre-activation grads).  dc_work [B,H] carries
 * dL/dc between steps (caller-owned, contents undefined on entry). */
int r2d2_lstm_seq_bwd(const float* dhs, const float* w_hh, const float* gates,
                      const float* cs_ext, float* dgates, float* dc_work, int T, int B, int H,
                      void* stream);
/* conv torso backward for R frames (slab rows as in r2d2_torso_fwd_u8): dfeat [R,1568] is
 * overwritten; dw*,db* written.  scratch: r2d2_torso_bwd_scratch_bytes(R). */
int64_t r2d2_torso_bwd_scratch_bytes(int R);
int r2d2_torso_bwd_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* a1,
                      const float* a2, const float* feat, float* dfeat, const float* w2,
                      const float* w3, float* dw1, float* db1, float* dw2, float* db2,
                      float* dw3, float* db3, void* scratch, void* stream);
/* same for caller-provided f32 frames [R,4,84,84] (autograd of QNet.forward) */
int r2d2_torso_bwd_f32(const float* frames, int R, const float* a1, const float* a2,
                       const float* feat, float* dfeat, const float* w2, const float* w3,
                       float* dw1, float* db1, float* dw2, float* db2, float* dw3, float* db3,
                       void* scratch, void* stream);

/* ------------------------------------------------------------------ bf16 tensor-core path
 * (north-star 2e-2 precision class; tcgen05.mma + TMEM accumulators + TMA operand staging)
 * y[M,N] = x[M,K] . w[N,K]^T + bias1 + bias2 (ReLU optional); x, w bf16 row-major with K
 * contiguous (K % 8 == 0, N % 4 == 0); either or both outputs (f32 / bf16) may be given.
 * `relu` is a flag word: bit 0 = ReLU, bit 1 = biases are per output ROW (used to produce the
 * transposed input projection W_ih.x^T that the persistent LSTM kernel consumes).
 * The same entry point serves model.py:55 (x.W_ih^T), the dueling hidden layer (model.py:26-36)
 * and - on transposed bf16 copies - the data and weight gradients of both. */
int r2d2_tc_linear(const void* x_bf16, const void* w_bf16, const float* bias1, const float* bias2,
                   float* y_f32, void* y_bf16, int M, int N, int K, int relu, void* stream);
/* y[M,N] = mask[M,N] > 0 ? x.w^T : 0 -- a data gradient with the ReLU' of the layer it flows
 * into fused (mask = that layer's forward output), e.g. dL/dfeat masked by feat. */
int r2d2_tc_linear_masked(const void* x_bf16, const void* w_bf16, const float* mask, float* y_f32,
                          int M, int N, int K, void* stream);
/* conv torso forward on tensor cores (same contract as r2d2_torso_fwd_{u8,f32}): implicit GEMM,
 * patches gathered by producer warps into swizzled shared memory, bf16x3 compensation, weights
 * resident, two TMEM accumulators */
int r2d2_tc_torso_fwd_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* w1,
                         const float* b1, const float* w2, const float* b2, const float* w3,
                         const float* b3, float* a1, float* a2, float* feat, void* stream);
/* The first conv layer of TWO networks over the same u8 frames in one pass (the learner's online
 * and target nets, learner.py:71-72,79: both read the sampled frames; the pass is bound by
 * gathering the receptive fields, which is done once).  Net a covers frames 0..nt of every
 * sequence -> a1_a [nt*B,32,20,20]; net b covers frames s0_b..nt -> a1_b [(nt-s0_b)*B,32,20,20].
 * r2d2_tc_torso_fwd_tail runs the remaining two layers of one network from its a1; with feat_hi /
 * feat_lo (bf16 [R,1568], optional, together) the last layer also writes the r2d2_split_bf16 image
 * of feat that the input projection consumes. */
int r2d2_tc_conv1_fwd_dual_u8(const uint8_t* slab, int B, int S, int nt, const float* w1_a,
                              const float* b1_a, float* a1_a, const float* w1_b, const float* b1_b,
                              float* a1_b, int s0_b, void* stream);
/* Ring forms: the sampled sequences are read IN PLACE from the replay ring (frame (t,b) = ring row
 * (seq_index[b] + s0 + t) mod M, seq_index on the device = r2d2_tree_sample's output), so the
 * frames are never copied into a slab (r2d2_replay_gather with slab = NULL still fills the small
 * per-sequence fields).  Same results as the slab forms. */
int r2d2_tc_conv1_fwd_dual_ring(const uint8_t* ring_state, int64_t M, const int64_t* seq_index, int B,
                                int nt, const float* w1_a, const float* b1_a, float* a1_a,
                                const float* w1_b, const float* b1_b, float* a1_b, int s0_b,
                                uint32_t* relu_bits_a, void* stream);
/* relu_bits_a (optional, u32 [nt*B, 20*20]): bit co of word (frame, pixel) = (a1_a[frame, co, pixel] > 0);
 * r2d2_tc_conv_dgrad2_masked takes ReLU' from it instead of reading the fp32 activations back. */
int r2d2_tc_conv_dgrad2_masked(const float* dz2, const float* w2, const uint32_t* relu_bits, float* dz1,
                               int R, void* stream);
int r2d2_tc_conv1_wgrad_ring(const uint8_t* ring_state, int64_t M, const int64_t* seq_index, int B, int s0,
                             int nt, const float* dz, float* dw, float* db, float* part, void* stream);
int r2d2_tc_torso_fwd_tail(const float* a1, int R, const float* w2, const float* b2, const float* w3,
                           const float* b3, float* a2, float* feat, void* feat_hi, void* feat_lo,
                           void* stream);
int r2d2_tc_torso_fwd_f32(const float* frames, int R, const float* w1, const float* b1,
                          const float* w2, const float* b2, const float* w3, const float* b3,
                          float* a1, float* a2, float* feat, void* stream);
/* Persistent LSTM unroll on tensor cores (model.py:52-60): one cooperative launch runs all T
 * steps; W_hh stays resident in shared memory (bf16 hi/lo), the cell state in registers, h_t is
 * exchanged between the CTAs through a ping-pong buffer + grid-wide counter in `scratch`
 * (r2d2_tc_lstm_scratch_bytes, 256-byte aligned).  xprojT is the TRANSPOSED input projection
 * [4H, ldx] (column col0 + t*B + b), biases included.  Outputs as r2d2_lstm_seq_fwd.
 * H % 64 == 0, H <= 512; H/16 * ceil(B/64) CTAs must be co-resident. */
int64_t r2d2_tc_lstm_scratch_bytes(int B, int H);
/* diagnostics: [T,32] int64 device buffer receiving globaltimer stamps of CTA (0,0) per step
 * (barrier passed, proxy fence, MMAs issued, accumulator ready, gates staged, stores issued,
 * fence done, arrive done); null switches it off */
int r2d2_tc_lstm_set_trace(void* trace);
/* h16_hi / h16_lo (optional, together; bf16 [T - hist_t0, B, H]): the r2d2_split_bf16 image of
 * h_t for t >= hist_t0, written by the kernel itself for the dueling-head GEMM. */
int r2d2_tc_lstm_seq_fwd(const float* xprojT, int64_t ldx, int64_t col0, const float* w_hh,
                         float* hs_ext, float* cs_ext, float* gates, void* h16_hi, void* h16_lo,
                         int hist_t0, int T, int B, int H, void* scratch, void* stream);
/* Persistent BPTT on tensor cores (same contract as r2d2_lstm_seq_bwd, plain bf16 operands):
 * dG_{t+1} pulled in by bulk copies, W_hh^T slice resident, dL/dc in registers, grid-wide counter.
 * dgates_bf16 (optional, [T,B,4H] bf16): a rounded copy of dgates for the gradient GEMMs. */
int64_t r2d2_tc_lstm_bwd_scratch_bytes(int B, int H);
int r2d2_tc_lstm_seq_bwd(const float* dhs, const float* w_hh, const float* gates,
                         const float* cs_ext, float* dgates, void* dgates_bf16, int T, int B, int H,
                         void* scratch, void* stream);
/* conv torso backward split for the tensor-core path: r2d2_torso_bwd_dgrad_u8 runs the ReLU
 * masks, data gradients and bias gradients (scratch as r2d2_torso_bwd_u8: [dz2 | dz1 | partials])
 * and r2d2_tc_torso_wgrad_u8 the three weight gradients on tcgen05 (plain bf16, the forward's
 * patch tile reused as an MN-major operand; per-CTA partials summed in fixed order). */
int r2d2_torso_bwd_dgrad_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* a1,
                            const float* a2, const float* feat, float* dfeat, const float* w2,
                            const float* w3, float* db1, float* db2, float* db3, void* scratch,
                            void* stream);
int r2d2_tc_torso_wgrad_u8(const uint8_t* slab, int B, int S, int s0, int nt, const float* a1,
                           const float* a2, const float* dz1, const float* dz2, const float* dz3,
                           float* dw1, float* dw2, float* dw3, void* scratch, void* stream);
int r2d2_reduce_partials(const float* part, float* out, int64_t n, int splits, void* stream);
/* one conv layer's weight gradient (layer 1: `in` = u8 slab, B/S/s0/nt as r2d2_torso_fwd_u8;
 * layers 2/3: `in` = that layer's fp32 NCHW input activations, S/s0 ignored) and, when db is given,
 * its bias gradient db[co] = sum over frames and pixels of dz (accumulated by the producer threads
 * while they build the dZ operand).  Private partial buffer of r2d2_sm_count()*(32*K + 32) floats
 * (K = 256/512/288) so layers can run on separate streams. */
int r2d2_tc_conv_wgrad_layer(int layer, const void* in, int B, int S, int s0, int nt,
                             const float* dz, float* dw, float* db, float* part, void* stream);
/* conv data gradients on tcgen05 (plain bf16, ReLU' of the layer input fused): dz3 [R,32,7,7]
 * -> dz2 [R,32,9,9] (a2 = that layer's activations), dz2 -> dz1 [R,32,20,20]; the stride-2 layer
 * runs as four output-parity classes.  Plus the two small pieces of the conv backward that stay
 * elementwise: d[i] = a[i] > 0 ? d[i] : 0, and per-channel sums of a [R,32,PIX] gradient, PIX in {49,81,400}
 * (part: >= R2D2_CHANSUM_PART_FLOATS floats of scratch). */
#define R2D2_CHANSUM_PART_FLOATS (296 * 32)
int r2d2_tc_conv_dgrad3(const float* dz3, const float* w3, const float* a2, float* dz2, int R,
                        void* stream);
int r2d2_tc_conv_dgrad2(const float* dz2, const float* w2, const float* a1, float* dz1, int R,
                        void* stream);
int r2d2_relu_mask(float* d, const float* a, int64_t n, void* stream);
int r2d2_chansum(const float* dz, float* db, int R, int PIX, float* part, void* stream);
/* Error-compensated variant ("bf16x3"): x = x_hi + x_lo, w = w_hi + w_lo (r2d2_split_bf16) and
 * y = x_hi.w_hi^T + x_hi.w_lo^T + x_lo.w_hi^T, fp32 accumulation in TMEM -- operand error 2^-16
 * instead of 2^-9.  Used for FORWARD contractions: a 1e-3 forward perturbation flips ReLU masks
 * and moves the (heavily cancelling) gradient sums by several percent, see DESIGN.md section 2. */
int r2d2_tc_linear_x3(const void* x_hi, const void* x_lo, const void* w_hi, const void* w_lo,
                      const float* bias1, const float* bias2, float* y_f32, void* y_bf16, int M,
                      int N, int K, int relu, void* stream);
/* y[P, Q] = a^T . b for row-major bf16 a [Kc, P] and b [Kc, Q] (P, Q multiples of 8): the weight
 * gradients dW = dY^T . X (learner.py:98, autograd of model.py:26-36,55) straight from the
 * activations and gradients as they lie in memory -- both operands are read MN-major by the
 * tensor cores, no transposed copies. */
int r2d2_tc_linear_tn(const void* a_bf16, const void* b_bf16, float* y_f32, int Kc, int P, int Q, void* stream);
/* y[M, N] = a . b for row-major bf16 a [M, K] and b [K, N] (N, K multiples of 8): the data gradients
 * dX = dY . W with the weight matrix as stored ([out, in]) -- b is read MN-major, no W^T copy. */
int r2d2_tc_linear_nn(const void* a_bf16, const void* b_bf16, float* y_f32, int M, int N, int K, void* stream);
/* Same, on at most max_ctas SMs (0 = all).  The kernel is persistent -- a CTA keeps its SM until all
 * of its tiles are done -- so a caller that wants another kernel to START beside it (the learner:
 * the online network's recurrence beside the target network's input projection) leaves SMs free. */
int r2d2_tc_linear_x3_ctas(const void* x_hi, const void* x_lo, const void* w_hi, const void* w_lo,
                           const float* bias1, const float* bias2, float* y_f32, void* y_bf16, int M,
                           int N, int K, int relu, int max_ctas, void* stream);
int r2d2_split_bf16(const float* in, void* hi_bf16, void* lo_bf16, int64_t n, void* stream);
/* f32 -> bf16 copies feeding r2d2_tc_linear: flat cast, and out[c,r] = in[r,c] */
int r2d2_cast_bf16(const float* in, void* out_bf16, int64_t n, void* stream);
int r2d2_transpose_cast_bf16(const float* in, void* out_bf16, int R, int C, void* stream);

/* ------------------------------------------------------------------ optimiser (K10)
 * torch.optim.RMSprop(centered=True) of learner.py:51 over a flat f32 buffer:
 * g *= grad_scale; v = a v + (1-a) g^2; m = a m + (1-a) g; p -= lr*g/(sqrt(v-m^2)+eps) */
int r2d2_rmsprop_centered(float* param, const float* grad, float* square_avg, float* grad_avg,
                          int64_t n, double lr, double alpha, double eps, double grad_scale,
                          void* stream);

/* ------------------------------------------------------------------ data-parallel exchange (section 8e)
 * No reference counterpart: the reference has ONE learner process (learner.py:68-104).  With one
 * process per GPU the only exchange is the sum of the flat gradient; these entry points do it and
 * the optimiser step (learner.py:51,99-100) in one kernel over NVLink peer memory.
 * r2d2_peer_alloc: cudaMalloc (zero-filled) + CUDA IPC export (handle64: 64 bytes) of a buffer the
 * other ranks map with r2d2_peer_open.  r2d2_dp_reduce_rmsprop: grads/params/flags are arrays of
 * `world` device pointers (index = rank; the local rank's own buffers included; flags: 4 KiB,
 * zero at start).  One call handles the slice [offset, offset+n) of the flat vectors ("bucket",
 * 0..7, each with its own flags and its own state shards), so that slices whose gradients are
 * complete early can be exchanged while the rest of the backward pass still runs.  Rank r reduces shard r of all gradients in rank order, applies centered RMSprop
 * (state only for its shard: r2d2_dp_shard_elems floats each) and writes the new parameters of its
 * shard into every rank's parameter buffer; the call returns (on the stream) when every rank's
 * shard has arrived.  All ranks must call it once per step; a missing rank traps after
 * R2D2_PEER_TIMEOUT_S seconds (environment, default 600). */
int r2d2_peer_alloc(int64_t bytes, void** ptr, void* handle64);
int r2d2_peer_open(const void* handle64, void** ptr);
int r2d2_peer_close(void* ptr);
int r2d2_peer_free(void* ptr);
int64_t r2d2_dp_shard_elems(int64_t n, int world);
int r2d2_dp_reduce_rmsprop(int rank, int world, void* const* grads, void* const* params,
                           void* const* flags, float* square_avg_shard, float* grad_avg_shard,
                           int64_t offset, int bucket, int64_t n, double lr, double alpha, double eps,
                           double grad_scale, void* stream);

/* ------------------------------------------------------------------ actor-side experience (section 8f)
 * Replaces, for N actors stepped together, what /root/reference actor.py:74-94 does per actor
 * after env.step returned: NStepMemory.add / get (replay_memory.py:27-52, deques of maxlen n --
 * so the first transition of an episode is pushed out before the first get, and the reward
 * deque, never popped, gives every transition flushed at the end of an episode the same
 * discounted sum), calc_priority (actor.py:144-157: the CURRENT step's Q-values stand in for
 * the next state's, `done` is the current step's), ReplayMemory.add into the actor's local ring
 * (replay_memory.py:93-105: frames stored as state*255 truncated to u8), and at the end of an
 * episode set_seq_start_index (actor.py:159-167: a start every `overlap` rows from the
 * episode's first row while < last-T, plus a forced start at last-T; their sequence
 * priorities eta*max + (1-eta)*mean in numpy's order, replay_memory.py:187-188).  If fewer than
 * T rows are in the ring at the end of an episode the reference indexes a negative slot (and
 * raises on the empty slice); here that forced start is skipped.
 *
 * All pointers are device memory owned by the caller; leading dimension N (actor) everywhere.
 *   n-step ring : ns_state u8[N,n,28224], ns_hc / ns_thc f32[N,n,2H] (h|c of the online / target
 *                 net), ns_q f32[N,n,6], ns_action i32[N,n], ns_reward f64[N,n],
 *                 counters i32[N,2] = {steps in this episode, entries waiting}
 *   local rings : the ten replay keys with shapes [N,C,...] (same dtypes as the replay ring),
 *                 index i64[N] (rows written), ep_start i64[N]; the caller hands the rows to
 *                 the learner's replay (r2d2_replay_ingest) before C rows are reached.
 *   per call    : obs f32[N,28224] in [0,1], action i64[N], q / tq f32[N,6], h c th tc
 *                 f32[N,H], reward f64[N], done u8[N]; gamma_pow f64[n] = gamma^0..gamma^(n-1)
 *                 as the host computes them (the reference sums gamma**i * r in Python floats),
 *                 gamma_n = gamma^n rounded to f32. */
typedef struct {
    uint8_t* ns_state; float* ns_hc; float* ns_thc; float* ns_q; int32_t* ns_action;
    double* ns_reward; int32_t* counters;
    uint8_t* state; float* hs_cs; float* target_hs_cs; int8_t* action; float* reward; float* done;
    int8_t* stack_count; float* priority; float* sequence_priority; uint8_t* is_seq_start;
    int64_t* index; int64_t* ep_start;
} r2d2_actor_buffers;
int r2d2_actor_observe(const r2d2_actor_buffers* buf, int N, int n, int H, int64_t C, int T,
                       int overlap, const float* obs, const int64_t* action, const float* q,
                       const float* h, const float* c, const float* tq, const float* th,
                       const float* tc, const double* reward, const uint8_t* done,
                       const double* gamma_pow, float gamma_n, float prio_eps, float alpha,
                       float eta_f32, float one_minus_eta_f32, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* R2D2_B200_H_ */
