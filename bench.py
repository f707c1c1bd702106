#!/usr/bin/env python
"""Benchmark of the R2D2 learner hot path (BASELINE.json metric: learner sequences/sec at
B=64, T=80 = burn-in 40 + learn 40, n=5, LSTM 512, synthetic 84x84 frames).

    python bench.py --gpus N --steps K --warmup W            # this repo, one rank per GPU
    python bench.py --impl reference --steps K --warmup W    # the reference path on host cores

One "step" = one full ``Learner.train()``: PER draw -> slab gather -> conv torso + LSTM burn-in /
learn / double-Q unrolls of both nets -> n-step double-Q target + TD + loss -> backward ->
centered RMSprop -> per-step + per-sequence priority update + tree refresh.  Nothing is
skipped or cached.  Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FRAME = 4 * 84 * 84
FEAT = 1568
LATENCY_FLOOR_US = 2.5   # measured floor of the recurrence's per-step exchange protocol


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--burn-in", type=int, default=40)
    ap.add_argument("--learn", type=int, default=40)
    ap.add_argument("--n-step", type=int, default=5)
    ap.add_argument("--cell", type=int, default=512)
    ap.add_argument("--memory-size", type=int, default=500000)
    ap.add_argument("--ingest-rows", type=int, default=640,
                    help="e2e leg: fresh transitions uploaded from pinned host memory per step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-sub", action="store_true",
                    help="skip the N=1 side records (fp32 class, configs 3 and 5, eager-torch reference)")
    ap.add_argument("--tma-gather", action="store_true")
    ap.add_argument("--dp-exchange", default="peer", choices=["peer", "nccl"],
                    help="N > 1: gradient exchange + optimiser as one kernel over NVLink peer memory "
                         "(csrc/dp_peer.cu), or NCCL all-reduce followed by the per-rank optimiser kernel")
    ap.add_argument("--no-cuda-graph", action="store_true",
                    help="enqueue every kernel from Python instead of replaying the captured step")
    ap.add_argument("--precision", default="bf16", choices=["fp32", "bf16"],
                    help="bf16: tcgen05 tensor-core contractions (2e-2 class); fp32: FFMA (1e-5 class)")
    return ap.parse_args()


# ------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------ work model
def work_model(B, BI, L, n, H, nparams):
    """Algorithmic work per STEP of every named kernel: (kind, amount) with kind 'flop' (tensor
    roofline) or 'byte' (HBM roofline).  FLOPs are the algorithm's: the 3x
    MMA issue of the compensated bf16x3 kernels is NOT counted.  Bytes are compulsory traffic."""
    T, S = BI + L, BI + L + n
    M = L * B
    c1, c2, c3 = 6553600, 2654208, 903168
    fr_fwd = (S + T) * B
    proj = 2 * 4 * H * FEAT
    lin_fwd = (S + T) * B * proj + 3 * M * 2 * (2 * H * H)
    lin_bwd = 2 * M * proj + 2 * M * 2 * (2 * H * H) + M * 2 * 4 * H * H
    steps_fwd = T + L + T
    lstm = 2 * B * 4 * H * H
    w = {
        # fp32 path
        "conv1_fwd": ("flop", c1 * fr_fwd), "conv2_fwd": ("flop", c2 * fr_fwd), "conv3_fwd": ("flop", c3 * fr_fwd),
        "linear_fwd": ("flop", lin_fwd), "linear_dgrad": ("flop", 2 * M * (2 * H * H) + M * proj),
        "linear_wgrad": ("flop", 2 * M * (2 * H * H) + M * 2 * 4 * H * H + M * proj),
        "lstm_step_fwd_kernel": ("flop", lstm * steps_fwd), "lstm_step_bwd_kernel": ("flop", lstm * (L - 1)),
        "conv1_wgrad": ("flop", c1 * M), "conv2_wgrad": ("flop", c2 * M), "conv3_wgrad": ("flop", c3 * M),
        "conv_dgrad_kernel<4,20,9,2>": ("flop", c2 * M), "conv_dgrad_kernel<3,9,7,1>": ("flop", c3 * M),
        # tensor-core path
        "tc_conv1_fwd": ("flop", c1 * fr_fwd), "tc_conv1_fwd_dual": ("flop", c1 * fr_fwd),
        "tc_conv2_fwd": ("flop", c2 * fr_fwd),
        "tc_conv3_fwd": ("flop", c3 * fr_fwd),
        "tc_gemm_kernel_x3": ("flop", lin_fwd), "tc_gemm_kernel": ("flop", lin_bwd),
        # weight gradients dW = dY^T.X (heads, W_hh, W_ih) and data gradients dX = dY.W (heads, W_ih)
        "tc_gemm_tn_kernel": ("flop", M * 2 * (2 * H * H) + M * 2 * 4 * H * H + M * proj),
        "tc_gemm_nn_kernel": ("flop", M * 2 * (2 * H * H) + M * proj),
        "tc_lstm_fwd_kernel": ("flop", lstm * steps_fwd), "tc_lstm_bwd_kernel": ("flop", lstm * (L - 1)),
        "tc_conv1_wgrad": ("flop", c1 * M), "tc_conv2_wgrad": ("flop", c2 * M), "tc_conv3_wgrad": ("flop", c3 * M),
        "tc_conv2_dgrad": ("flop", c2 * M), "tc_conv3_dgrad": ("flop", c3 * M),
        # HBM-bound
        "gather_rows_kernel": ("byte", 2 * B * S * FRAME), "gather_rows_tma_kernel": ("byte", 2 * B * S * FRAME),
        "rmsprop_kernel": ("byte", 28 * nparams),
        # (the feature split is fused into the conv3 epilogue; what is left: both nets' W_ih and
        #  head weights, and the three head inputs)
        "split_bf16_kernel": ("byte", 8 * (2 * (4 * H * FEAT + 2 * H * H) + 3 * M * H)),
        "dp_reduce_rmsprop_kernel": ("byte", 4 * nparams * 3),   # per rank: grads in, params out, shard state
        "chansum_partial_kernel": ("byte", 4 * M * (1568 + 2592 + 12800)),
    }
    total = fr_fwd * (c1 + c2 + c3 + proj) + steps_fwd * lstm + 3 * M * 2 * (2 * H * H + 7 * H) \
        + M * (2 * (c1 + c2 + c3 + proj + 4 * H * H * 2 + 2 * (2 * H * H + 7 * H)) - c1)
    return w, total


# ------------------------------------------------------------------------------ reference arm
def workload_config(args, world):
    """The workload, in the same words from both arms (the driver compares the two dicts)."""
    return {"workload": "configs[1]: synthetic Atari 84x84 R2D2 learner step "
                        "(PER sample -> LSTM unroll -> n-step TD -> priority update)",
            "batch_per_gpu": args.batch, "global_batch": world * args.batch,
            "seq_len": args.burn_in + args.learn, "burn_in": args.burn_in,
            "n_step": args.n_step, "lstm": args.cell, "frame": "4x84x84 u8",
            "replay_slots_per_gpu": args.memory_size, "parallelism": f"dp{world}",
            "l2_policy": "inputs larger than L2 (replay ring >> 126 MB; every step reads a fresh "
                         "154 MB of sampled frames)"}


def cpu_reference_step_factory(args, M):
    """One learner step of the reference on the host cores, at the bench's full shapes.

    Preferred: the UNMODIFIED reference, byte-compiled from /root/reference into oracle/_ref by
    oracle/make_ref.py in the build container (kind "reference"; /root/reference itself does not
    exist on the GPU box).  Fallback when oracle/_ref did not travel: the oracle port of
    ReplayMemory.sample + Learner.train (pinned to the reference by oracle/make_golden.py),
    kind "port".  Returns (step_fn, kind, actual_slots)."""
    import numpy as np
    import torch
    from oracle import ref_harness as RH

    torch.set_num_threads(os.cpu_count() or 1)
    B, BI, L, n, H = args.batch, args.burn_in, args.learn, args.n_step, args.cell
    # the ring costs ~36 KB of host RAM per slot: shrink it when the box cannot hold it
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 32 << 30
    while M > 20000 and M * (28224 + 8 * H + 64) > 0.5 * avail:
        M //= 2
    state = {"i": 0}
    if RH.available():
        Lr = RH.make_reference_learner(B, BI, L, n, H, M, device="cpu")

        def step():
            torch.manual_seed(1000 + state["i"])
            state["i"] += 1
            Lr.train()
        return step, "reference", M

    from oracle import net_oracle as NO
    from oracle import replay_oracle as RO
    rep = RO.ReplayOracle(M, H, n, BI, L)
    N = int(0.8 * M)
    RH.fill_tiled(rep.memory, N, H)
    rep.index = N
    starts = np.arange(0, max(N - rep.T - n, 1), 40)
    rep.update_sequence_priority(starts)
    rep.memory["is_seq_start"][starts] = 1
    p = NO.init_params(0, H)
    pt = {k: v.clone() for k, v in p.items()}
    sq = {k: torch.zeros_like(v) for k, v in p.items()}
    ga = {k: torch.zeros_like(v) for k, v in p.items()}

    def step():
        torch.manual_seed(1000 + state["i"])
        state["i"] += 1
        u = torch.rand(B, dtype=torch.float64).numpy()
        si = rep.sample_indices(u)
        g = rep.gather(si)
        res = NO.train_step(p, pt, NO.to_torch_batch(g["batch"]), burn_in=BI, n_step=n)
        for k in NO.PARAM_KEYS:
            NO.rmsprop_centered(p[k], res["grads"][k], sq[k], ga[k])
        rep.update_priority(g["index"][BI:].reshape(-1), res["priority"])
        rep.update_sequence_priority(si, True)

    return step, "port", M


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    step, kind, slots = cpu_reference_step_factory(args, args.memory_size)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    B = args.batch
    val = B / dt
    cores = os.cpu_count() or 1
    what = ("the unmodified reference (oracle/_ref: Learner.train on device='cpu', LSTM/head widened to "
            f"{args.cell})" if kind == "reference" else "oracle port of ReplayMemory.sample + Learner.train")
    line = {
        "impl": "reference", "metric": "learner_sequences_per_sec", "value": val, "unit": "seq/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": val, "unit": "seq/s", "cores": cores, "kind": kind,
                         "sample": f"{args.steps} full steps (B={B}) after {args.warmup} warm-up: {what}, "
                                   f"torch CPU fp32 with {cores} threads, {slots}-slot replay on rank 0"},
        "e2e": {"value": val, "unit": "seq/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


# ------------------------------------------------------------------------------ our arm
def fill_synthetic(L, args, rank):
    """SYN-R2D2 generated on the device: M slots, 80 % filled, sequence
    starts every 40 slots, done every 997th row."""
    import torch
    rm = L.replay_memory
    M = rm.memory_size
    N = int(0.8 * M)
    gen = torch.Generator(device=rm.device).manual_seed(1 + rank)
    raw = lambda k: dict.__getitem__(rm.memory, k)  # noqa: E731
    raw("state")[:N].random_(0, 256, generator=gen)
    raw("hs_cs")[:N].normal_(0, 0.1, generator=gen)
    raw("target_hs_cs")[:N].normal_(0, 0.1, generator=gen)
    raw("action")[:N].random_(0, 6, generator=gen)
    raw("reward")[:N].normal_(0, 1, generator=gen)
    raw("done")[:N, 0] = (torch.arange(N, device=rm.device) % 997 == 996).float()
    raw("stack_count")[:N] = 1
    raw("priority")[:N].uniform_(0.01, 1.01, generator=gen)
    rm.index = N
    rm.burn_in_length, rm.learning_length = args.burn_in, args.learn
    rm.sequence_length = args.burn_in + args.learn
    starts = torch.arange(0, N - rm.sequence_length - args.n_step, 40, device=rm.device)
    for i in range(0, starts.numel(), 4096):
        rm.update_sequence_priority(starts[i:i + 4096])
    rm.memory["is_seq_start"][starts] = 1
    rm.rebuild_tree()


def make_host_dump(rows, H):
    import torch
    g = torch.Generator().manual_seed(3)
    pin = lambda t: t.pin_memory()  # noqa: E731
    return {
        "state": pin(torch.randint(0, 256, (rows, 4, 84, 84), dtype=torch.uint8, generator=g)),
        "hs_cs": pin(torch.randn(rows, 2 * H, generator=g) * 0.1),
        "target_hs_cs": pin(torch.randn(rows, 2 * H, generator=g) * 0.1),
        "action": pin(torch.randint(0, 6, (rows, 1), generator=g).to(torch.int8)),
        "reward": pin(torch.randn(rows, 1, generator=g)),
        "done": pin(torch.zeros(rows, 1)),
        "stack_count": pin(torch.ones(rows, dtype=torch.int8)),
        "priority": pin(torch.rand(rows, generator=g) + 0.01),
        "sequence_priority": pin(torch.rand(rows, generator=g) + 0.01),
        "is_seq_start": pin((torch.arange(rows) % 40 == 0).to(torch.uint8)),
    }


def time_device(L, steps, barrier):
    """K un-instrumented steps, device time (CUDA events on the current stream)."""
    import torch
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(steps):
        L.train()
    ev1.record()
    barrier()
    return ev0.elapsed_time(ev1)


def time_e2e(L, steps, dump, barrier, warmup=5):
    """The public API with host buffers, wall clock: every step uploads `dump` (fresh experience)
    from pinned host memory through ReplayMemory.extend, runs Learner.train() (uniforms H2D
    inside) and copies the loss to pinned host memory; the host READS loss k-1 while step k runs,
    the way a training loop logs.  `warmup` untimed iterations of the same loop body come first
    (pinned staging allocation, the copy stream's first use).  Returns (seconds, last loss)."""
    import torch
    loss_pin = [torch.empty(1).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]
    losses = []

    def loop(n):
        for i in range(n):
            loss = L.train()                                  # lands the dump staged during step i-1
            L.replay_memory.extend(dump, overlap=True)        # H2D of the next dump overlaps step i
            loss_pin[i & 1].copy_(loss, non_blocking=True)
            loss_ev[i & 1].record()
            if i > 0:
                loss_ev[(i - 1) & 1].synchronize()
                losses.append(float(loss_pin[(i - 1) & 1]))
        loss_ev[(n - 1) & 1].synchronize()
        losses.append(float(loss_pin[(n - 1) & 1]))
        torch.cuda.synchronize()

    loop(max(2, warmup))
    barrier()
    t0 = time.perf_counter()
    loop(steps)
    return time.perf_counter() - t0, losses[-1]


def ev_time(fn, iters=50, warm=5):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def sub_fp32_class(args, L, dump):
    """The same workload in the exact-precision mode (every contraction fp32 FFMA: the 1e-5
    parity class of BASELINE.json's north_star), a short run: device-timed value and e2e."""
    import torch
    from pytorch_r2d2_b200 import Learner
    torch.manual_seed(0)
    L32 = Learner(1, {}, device=L.device, cell_size=args.cell, memory_size=64, precision="fp32")
    L32.batch_size, L32.burn_in_length, L32.learning_length, L32.bootstrap_steps = (
        args.batch, args.burn_in, args.learn, args.n_step)
    L32.sequence_length = args.burn_in + args.learn
    L32.replay_memory = L.replay_memory            # same ring, same tree
    L32.use_cuda_graph = not args.no_cuda_graph
    steps = max(3, min(args.steps, 10))
    sync = torch.cuda.synchronize
    for _ in range(3):
        L32.train()
    ms = time_device(L32, steps, sync) / steps
    sec, _ = time_e2e(L32, steps, dump, sync)
    L.replay_memory._wait_ingest()
    sync()
    return {"dtype": "f32", "tolerance_class": "1e-5 (tests/test_gpu_fullshape.py, fp32)",
            "value": args.batch / (ms * 1e-3), "unit": "seq/s", "ms_per_step": ms, "steps": steps,
            "warmup": 3, "e2e": {"value": args.batch * steps / sec, "unit": "seq/s"}}


def sub_config3(args):
    """BASELINE.json configs[2]: 4 M sequence starts, sum-tree sample + priority-update sweep,
    B = 64..4096, GPU-resident; the CPU restatement of the reference's selection
    (replay_memory.py:225-232) and self-refresh loop (:185-188) timed beside it."""
    import numpy as np
    import torch
    from pytorch_r2d2_b200 import _lib
    from pytorch_r2d2_b200.replay_memory import tree_layout
    N, T = 4_000_000, 80
    p = ((np.random.RandomState(0).rand(N).astype(np.float32) + np.float32(1e-3)) ** np.float32(0.6)).astype(np.float32)
    sp = torch.from_numpy(p).cuda()
    st = torch.ones(N, dtype=torch.uint8, device="cuda")
    prio = torch.rand(N, device="cuda") + 0.01
    _, _, total = tree_layout(N)
    nodes = torch.zeros(total, dtype=torch.float64, device="cuda")
    S, P = _lib.stream_ptr, _lib.ptr
    out = {"leaves": N, "seq_len": T, "cores": os.cpu_count() or 1}
    out["tree_build_ms"] = ev_time(lambda: _lib.call("r2d2_tree_build", P(sp), P(st), N, P(nodes), S()), 20) * 1e3
    prio_h, sp_h = prio.cpu().numpy(), p.copy()
    eta, ometa = np.float32(0.9), np.float32(1 - 0.9)
    for B in (64, 256, 1024, 4096):
        u = torch.rand(B, dtype=torch.float64, generator=torch.Generator().manual_seed(123)).cuda()
        idx = torch.empty(B, dtype=torch.int64, device="cuda")
        ts = ev_time(lambda: _lib.call("r2d2_tree_sample", P(sp), P(st), N, P(nodes), P(u), B, P(idx), S()))
        touched = torch.empty(3 * B, dtype=torch.int64, device="cuda")
        pvals = torch.rand(T, B, device="cuda")

        def update():
            _lib.call("r2d2_update_priority_seq", P(prio), N, P(idx), B, 0, T, P(pvals), S())
            _lib.call("r2d2_update_sequence_priority", P(prio), P(sp), P(st), N, P(idx), B, T,
                      float(eta), float(ometa), 1, P(touched), S())
            _lib.call("r2d2_tree_update", P(sp), P(st), N, P(nodes), P(touched), 3 * B, S())
        tu = ev_time(update, 20)
        # CPU: selection = mask + f32 normalise + multinomial (f64 weights, with replacement)
        t0 = time.perf_counter()
        starts = np.arange(N)[np.ones(N, np.uint8) == 1]
        w = torch.as_tensor(p[starts] / np.sum(p[starts]), dtype=torch.double)
        torch.multinomial(w, B, True)
        tc = time.perf_counter() - t0
        # CPU: per-step priority scatter + the per-sequence eta*max + (1-eta)*mean refresh
        idx_h, pv_h = idx.cpu().numpy(), pvals.cpu().numpy()
        t0 = time.perf_counter()
        prio_h[(idx_h[None, :] + np.arange(T)[:, None]) % N] = pv_h
        for i in idx_h:
            pr = prio_h[i:i + T]
            sp_h[i] = eta * pr.max() + ometa * pr.mean()
        tcu = time.perf_counter() - t0
        out[f"B{B}"] = {"gpu_sample_us": ts * 1e6, "gpu_samples_per_s": B / ts,
                        "gpu_update_us": tu * 1e6, "gpu_seq_updates_per_s": B / tu,
                        "cpu_sample_ms": tc * 1e3, "cpu_samples_per_s": B / tc,
                        "cpu_update_ms": tcu * 1e3, "cpu_seq_updates_per_s": B / tcu}
    return out


def sub_config5(args):
    """BASELINE.json configs[4] on one GPU: 256 actors stepped together (T=1), online AND target
    net every step with per-actor recurrent state (actor.py:96-106), next to the same two
    forwards of the oracle's QNet restatement on the host cores."""
    import numpy as np
    import torch
    from oracle import net_oracle as NO
    from pytorch_r2d2_b200 import BatchedActor
    N, H = 256, args.cell
    out = {"actors": N, "lstm": H}
    obs = torch.from_numpy(np.random.RandomState(2).randint(0, 256, (N, 4, 84, 84)).astype(np.float32)
                           / np.float32(255)).cuda()
    for prec in ("bf16", "fp32"):
        torch.manual_seed(0)
        act = BatchedActor(N, "cuda:0", H, precision=prec)
        act.forward_both(obs)
        t = ev_time(lambda: act.forward_both(obs), 30)
        out[prec] = {"ms_per_step": t * 1e3, "actor_steps_per_s": N / t}
    pcpu = NO.init_params(0, H)
    xc = obs.cpu()
    torch.set_num_threads(os.cpu_count() or 1)
    with torch.no_grad():
        NO.qnet_forward(pcpu, xc, None, None)
        t0 = time.perf_counter()
        for _ in range(3):
            NO.qnet_forward(pcpu, xc, None, None)
            NO.qnet_forward(pcpu, xc, None, None)   # both nets
        tc = (time.perf_counter() - t0) / 3
    out["cpu_port"] = {"ms_per_step": tc * 1e3, "actor_steps_per_s": N / tc, "cores": os.cpu_count() or 1}
    return out


def sub_reference_cuda_eager(args):
    """Informational: the unmodified reference with its own default device 'cuda:0' (learner.py:19)
    -- PyTorch eager / cuDNN / cuBLAS on this B200, host-side sample() and all."""
    import torch
    from oracle import ref_harness as RH
    if not RH.available():
        return {"unavailable": "oracle/_ref did not travel"}
    Lr = RH.make_reference_learner(args.batch, args.burn_in, args.learn, args.n_step, args.cell, 20000,
                                   device="cuda:0")
    for i in range(2):
        Lr.train()
    torch.cuda.synchronize()
    n = 3
    t0 = time.perf_counter()
    for _ in range(n):
        Lr.train()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n
    return {"value": args.batch / dt, "unit": "seq/s", "ms_per_step": dt * 1e3, "steps": n,
            "what": "unmodified reference Learner.train(), device='cuda:0', fp32 eager (cuDNN/cuBLAS), "
                    "20000-slot host replay, wall clock incl. its host-side sample() and H2D upload"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    from pytorch_r2d2_b200 import Learner, _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    _lib.load()
    B, BI, Ln, n, H = args.batch, args.burn_in, args.learn, args.n_step, args.cell

    torch.manual_seed(0)
    L = Learner(1, {}, device=f"cuda:{local}", cell_size=H, memory_size=args.memory_size,
                precision=args.precision)
    L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = B, BI, Ln, n
    L.sequence_length = BI + Ln
    L.replay_memory.use_tma_gather = args.tma_gather
    # world > 1: gradient exchange + optimiser in one NVLink peer-memory kernel; with it the whole
    # step is again a CUDA graph (capturing the NCCL all-reduce together with the cooperative LSTM
    # launches hung in testing, so the NCCL variant is enqueued eagerly)
    peer = world > 1 and args.dp_exchange == "peer" and L.enable_peer_exchange()
    L.use_cuda_graph = not args.no_cuda_graph and (world == 1 or peer)
    fill_synthetic(L, args, rank)
    torch.manual_seed(1234 + rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    barrier()      # ranks enter the first (peer-synchronised) step together
    for _ in range(max(args.warmup, 3)):
        L.train()
    barrier()

    # ---- timed region: K steps, device time, inputs resident in HBM ----------------------
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    launches0, replays0 = _lib.launch_count(), L.graph_replays
    ms = time_device(L, args.steps, barrier)
    # kernels of this library launched inside the timed region: the ones that went through the
    # C entry points plus, for CUDA-graph replays, the kernel nodes counted at capture
    launches = (_lib.launch_count() - launches0) + (L.graph_replays - replays0) * L.graph_kernel_nodes
    # ---- same K steps again with every kernel launch bracketed by CUDA events on its stream
    #      (per-kernel durations for the roofline; kept out of `value` because the ~600 event
    #      records per step cost a few percent)
    L.overlap_streams = False            # one stream: per-kernel durations free of co-runners
    graphed, L.use_cuda_graph = L.use_cuda_graph, False   # eager, so that launches can be bracketed
    L.train()
    torch.cuda.synchronize()
    _lib.prof_report(reset=True)
    _lib.prof_enable(True)
    for _ in range(args.steps):
        L.train()
    _lib.prof_enable(False)
    prof = _lib.prof_report(reset=True)
    L.overlap_streams, L.use_cuda_graph = True, graphed
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    ms_per_step = ms / args.steps
    value = world * B / (ms_per_step * 1e-3)

    # ---- e2e: the public API with host buffers ---------------------------------------------
    rows = args.ingest_rows
    dump = make_host_dump(rows, H)
    h2d = sum(v.numel() * v.element_size() for v in dump.values()) + B * 8
    e2e_s, loss_host = time_e2e(L, args.steps, dump, barrier, args.warmup)
    t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * B * args.steps / float(t.item())

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel -----------------------------------------------------
    peaks = {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        d = json.load(open(pk))
        peaks = {"hbm_gbs": d["hbm_gbs"], "bf16_tflops_sustained": d["bf16_tflops_sustained"],
                 "src": "measured"}
    work, total_flops = work_model(B, BI, Ln, n, H, L.net.layout.total)
    top = sorted(prof.items(), key=lambda kv: -kv[1][1])
    kernel_ms_per_step = sum(v[1] for v in prof.values()) / args.steps

    def roof_of(name, cnt, tot_ms):
        """achieved = algorithmic work per launch / average launch duration (CUDA events)."""
        avg_ms = tot_ms / cnt
        r = {"kernel": name, "launches_per_step": cnt / args.steps, "avg_ms": avg_ms,
             "ms_per_step": tot_ms / args.steps}
        if name in work:
            kind, amount = work[name]
            per_launch = amount / (cnt / args.steps)
            if kind == "flop":
                ach = per_launch / (avg_ms * 1e-3) / 1e12
                r.update({"bound": "tensor", "achieved": ach, "peak": peaks["bf16_tflops_sustained"],
                          "unit": "TFLOP/s", "frac": ach / peaks["bf16_tflops_sustained"],
                          "algorithmic_flops_per_launch": per_launch})
            else:
                ach = per_launch / (avg_ms * 1e-3) / 1e9
                r.update({"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                          "frac": ach / peaks["hbm_gbs"], "algorithmic_bytes_per_launch": per_launch})
        return r

    # DRAM traffic per launch from the committed `ncu --set full` capture of the same workload
    # (tools/ncu_traffic.py stamps the capture with the commit and a hash of csrc/; a capture
    # taken before the kernels last changed is NOT reported)
    traffic, tmeta, tnote = {}, {}, None
    tfiles = sorted(f for f in os.listdir(os.path.join(ROOT, "profiles")) if f.endswith("_ncu_traffic.json")) \
        if os.path.isdir(os.path.join(ROOT, "profiles")) else []
    if tfiles:
        traffic = json.load(open(os.path.join(ROOT, "profiles", tfiles[-1])))
        tmeta = traffic.pop("_meta", {})
        if tmeta.get("csrc_sha16") != csrc_sha16():
            tnote = (f"{tfiles[-1]} was captured at csrc hash {tmeta.get('csrc_sha16')} (commit "
                     f"{tmeta.get('commit')}); csrc is now {csrc_sha16()}: stale, not reported")
            print("bench.py: " + tnote, file=sys.stderr)
            traffic = {}
    dom_name, (dom_cnt, dom_ms) = top[0]
    roof = roof_of(dom_name, dom_cnt, dom_ms)
    tkey = dom_name if dom_name in traffic else ("tc_gemm_kernel*" if dom_name.startswith("tc_gemm") and
                                                 "tc_gemm_kernel*" in traffic else None)
    roof.update({"share_of_serial_step": dom_ms / args.steps / kernel_ms_per_step,
                 "traffic": traffic[tkey]["dram_bytes_per_launch"] if tkey else None,
                 "traffic_source": ({"file": tfiles[-1], **tmeta} if tkey else tnote),
                 "peak_source": peaks["src"],
                 "timing": "CUDA events around every launch, single-stream pass of the same K steps"})
    kernels = {k: roof_of(k, c, m) for k, (c, m) in top[:16]}
    for k in ("gather_rows_kernel", "gather_rows_tma_kernel", "rmsprop_kernel"):
        if k in prof and k not in kernels:
            kernels[k] = roof_of(k, *prof[k])
    # the recurrences are chains of dependent 134-MFLOP steps: tensor-peak fraction says little,
    # microseconds per dependent step against the exchange protocol's own floor says more
    dep = {"tc_lstm_fwd_kernel": 2 * (BI + Ln) + Ln, "tc_lstm_bwd_kernel": Ln - 1}
    for k, nsteps in dep.items():
        for r in ([roof] if roof["kernel"] == k else []) + ([kernels[k]] if k in kernels else []):
            r["dependent_steps_per_step"] = nsteps
            r["us_per_dependent_step"] = r["ms_per_step"] * 1e3 / nsteps
            r["latency_floor_us"] = LATENCY_FLOOR_US
            r["latency_floor_source"] = ("tools/ubench/tma_fanin.cu: one step of the exchange protocol "
                                         "with no compute (write 1/64 of 128 KB, release, poll, bulk-load)")

    line = {
        "metric": "learner_sequences_per_sec", "value": value, "unit": "seq/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16" if args.precision == "bf16" else "f32",
        "data": "synthetic",
        "config": workload_config(args, world),
        "detail": {
            "precision": args.precision, "cuda_graph": bool(L.use_cuda_graph),
            "dp_exchange": ("none" if world == 1 else
                            "fused reduce+rmsprop kernel over NVLink peer memory" if peer else
                            "nccl all-reduce + per-rank rmsprop"),
            "algorithmic_gflop_per_step_per_gpu": total_flops / 1e9,
            "e2e_ingest_rows_per_step": rows},
        "e2e": {"value": e2e_value, "unit": "seq/s", "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 4, "last_loss": loss_host},
        "gpu_launches": launches, "clocks": clk, "roofline": roof,
        "kernel_ms_per_step": kernel_ms_per_step, "kernels": kernels,
        "achieved_tflops_whole_step": total_flops / (ms_per_step * 1e-3) / 1e12,
    }

    # ---- N=1 only: the other named workloads and the CPU baseline, each bounded to seconds ----
    if world == 1 and not args.no_sub:
        for name, fn in (("fp32_class", lambda: sub_fp32_class(args, L, dump)),
                         ("config3", lambda: sub_config3(args)),
                         ("config5", lambda: sub_config5(args)),
                         ("reference_cuda_eager", lambda: sub_reference_cuda_eager(args))):
            try:
                line[name] = fn()
            except Exception as e:     # a side record must not take the headline line down
                line[name] = {"error": f"{type(e).__name__}: {e}"}
            torch.cuda.synchronize()
    if world == 1 and not args.no_cpu_baseline:
        # bounded sample of the same workload on the host cores
        step, kind, slots = cpu_reference_step_factory(args, 20000)
        step()
        t0 = time.perf_counter()
        ncpu = 3
        for _ in range(ncpu):
            step()
        dt = (time.perf_counter() - t0) / ncpu
        cores = os.cpu_count() or 1
        line["cpu_baseline"] = {
            "value": B / dt, "unit": "seq/s", "cores": cores, "kind": kind,
            "sample": f"{ncpu} full B={B} steps after 1 warm-up ("
                      + ("the unmodified reference from oracle/_ref" if kind == "reference" else "oracle port")
                      + f"), torch CPU fp32, {cores} threads, {slots}-slot replay"}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def csrc_sha16():
    """Hash of the CUDA sources (ties a committed ncu capture to the kernels it measured)."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "pytorch_r2d2_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


class _StdoutToStderr:
    """Route everything written to fd 1 (NCCL banners, library chatter) to stderr while the
    benchmark runs, so that stdout carries exactly ONE line: the JSON result."""

    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


_RESULT = []


def emit(line: dict):
    _RESULT.append(json.dumps(line))


def main():
    args = parse()
    with _StdoutToStderr():
        if args.impl == "reference":
            run_reference(args)
        else:
            run_ours(args)
    for r in _RESULT:
        print(r, flush=True)


if __name__ == "__main__":
    main()
