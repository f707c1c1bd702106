"""Timeline of one multi-stream learner step at the benchmark configuration: every kernel launch
with its stream, start and end (CUDA events on the launch streams, eager launches), to see what is
on the critical path.  python tools/timeline.py [--steps 3] > gpurun_out/timeline.txt"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from pytorch_r2d2_b200 import Learner, _lib

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--memory-size", type=int, default=100_000)
a = ap.parse_args()
torch.manual_seed(0)
L = Learner(1, {}, device="cuda:0", cell_size=512, memory_size=a.memory_size, precision="bf16")
L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = 64, 40, 40, 5
L.sequence_length = 80
class A: pass
fa = A(); fa.memory_size = a.memory_size; fa.cell = 512; fa.batch = 64; fa.burn_in = 40; fa.learn = 40; fa.n_step = 5
bench.fill_synthetic(L, fa, 0)
for _ in range(3):
    L.train()
torch.cuda.synchronize()
_lib.prof_report(reset=True)
_lib.prof_enable(True)
for _ in range(a.steps):
    L.train()
_lib.prof_enable(False)
tr = _lib.prof_trace()
_lib.prof_report(reset=True)
# last step only: launches after the last tree/sample kernel
starts = [i for i, r in enumerate(tr) if r[0].startswith("tree_build") or r[0].startswith("tree_sample")]
first = [i for i in starts if i == 0 or i - 1 not in starts]
lo = first[-1]
t0 = tr[lo][2]
print(f"# step = launches {lo}..{len(tr)-1}; total {tr[-1][3]-t0:.3f} ms (eager, event-instrumented)")
end_prev = {}
for name, sid, s, e in tr[lo:]:
    gap = s - end_prev.get(sid, s)
    end_prev[sid] = e
    print(f"s{sid} {s-t0:8.3f} {e-t0:8.3f} {e-s:7.3f} gap {gap:6.3f}  {name}")
