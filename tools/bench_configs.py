"""Side benchmarks for BASELINE.json configs 3 and 5 (not the headline bench line).

  config 3: large replay (4M sequence starts) sum-tree sample / priority-update sweep, B = 64..4096
  config 5: 256 concurrent actors' batched inference (both nets, T=1) on the learner's kernels

Prints one JSON object; CPU legs use the oracle restatement of the reference (numpy / torch)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from pytorch_r2d2_b200 import _lib, BatchedActor
from pytorch_r2d2_b200.replay_memory import tree_layout

def ev_time(fn, iters=50, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3

out = {"config3": {}, "config5": {}}
# ---------------------------------------------------------------- config 3 (SYN-TREE)
N = 4_000_000
p = ((np.random.RandomState(0).rand(N).astype(np.float32) + np.float32(1e-3)) ** np.float32(0.6)).astype(np.float32)
sp = torch.from_numpy(p).cuda(); st = torch.ones(N, dtype=torch.uint8, device="cuda")
prio = torch.rand(N, device="cuda") + 0.01
sizes, offs, total = tree_layout(N)
nodes = torch.zeros(total, dtype=torch.float64, device="cuda")
S = _lib.stream_ptr; P = _lib.ptr
build = lambda: _lib.call("r2d2_tree_build", P(sp), P(st), N, P(nodes), S())
out["config3"]["tree_build_4M_ms"] = ev_time(build, 20) * 1e3
T = 80
for B in (64, 256, 1024, 4096):
    u = torch.rand(B, dtype=torch.float64, generator=torch.Generator().manual_seed(123)).cuda()
    idx = torch.empty(B, dtype=torch.int64, device="cuda")
    sample = lambda: _lib.call("r2d2_tree_sample", P(sp), P(st), N, P(nodes), P(u), B, P(idx), S())
    ts = ev_time(sample)
    touched = torch.empty(3 * B, dtype=torch.int64, device="cuda")
    pvals = torch.rand(T, B, device="cuda")
    def update():
        _lib.call("r2d2_update_priority_seq", P(prio), N, P(idx), B, 0, T, P(pvals), S())
        _lib.call("r2d2_update_sequence_priority", P(prio), P(sp), P(st), N, P(idx), B, T,
                  float(np.float32(0.9)), float(np.float32(1 - 0.9)), 1, P(touched), S())
        _lib.call("r2d2_tree_update", P(sp), P(st), N, P(nodes), P(touched), 3 * B, S())
    tu = ev_time(update, 20)
    # CPU: the reference's selection (mask + f32 normalise + multinomial) restated
    t0 = time.perf_counter()
    starts = np.nonzero(np.ones(N, np.uint8) == 1)[0]
    w = torch.as_tensor(p[starts] / np.sum(p[starts]), dtype=torch.double)
    torch.multinomial(w, B, True)
    tc = time.perf_counter() - t0
    out["config3"][f"B{B}"] = {"gpu_sample_us": ts * 1e6, "gpu_samples_per_s": B / ts,
                                "gpu_update_us": tu * 1e6, "gpu_seq_updates_per_s": B / tu,
                                "cpu_sample_ms": tc * 1e3, "cpu_samples_per_s": B / tc}
# ---------------------------------------------------------------- config 5 (SYN-ACT)
from oracle import net_oracle as NO
for prec in ("bf16", "fp32"):
    torch.manual_seed(0)
    act = BatchedActor(256, "cuda:0", 512, precision=prec)
    obs = torch.from_numpy(np.random.RandomState(2).randint(0, 256, (256, 4, 84, 84)).astype(np.float32) / np.float32(255)).cuda()
    act.forward_both(obs)
    t = ev_time(lambda: act.forward_both(obs), 30)
    out["config5"][prec] = {"ms_per_step_256_actors": t * 1e3, "actor_steps_per_s": 256 / t}
pcpu = NO.init_params(0, 512)
xc = obs.cpu()
torch.set_num_threads(os.cpu_count())
h = c = None
with torch.no_grad():
    NO.qnet_forward(pcpu, xc, None, None)
    t0 = time.perf_counter()
    for _ in range(3):
        NO.qnet_forward(pcpu, xc, None, None); NO.qnet_forward(pcpu, xc, None, None)   # both nets
    tc = (time.perf_counter() - t0) / 3
out["config5"]["cpu_port"] = {"ms_per_step_256_actors": tc * 1e3, "actor_steps_per_s": 256 / tc, "cores": os.cpu_count()}
print(json.dumps(out, indent=1))
