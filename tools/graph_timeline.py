"""Timeline of one CUDA-graph-replayed learner step at the benchmark configuration, from
torch.profiler (CUPTI): kernel name, stream, start, end.  python tools/graph_timeline.py > out.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from pytorch_r2d2_b200 import Learner

torch.manual_seed(0)
L = Learner(1, {}, device="cuda:0", cell_size=512, memory_size=100_000, precision="bf16")
L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = 64, 40, 40, 5
L.sequence_length = 80
L.use_cuda_graph = True
class A: pass
fa = A(); fa.memory_size = 100_000; fa.cell = 512; fa.batch = 64; fa.burn_in = 40; fa.learn = 40; fa.n_step = 5
bench.fill_synthetic(L, fa, 0)
for _ in range(5):
    L.train()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        L.train()
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
# last step = events after the last tree_sample kernel
idx = [i for i, e in enumerate(evs) if "tree_sample" in e.name]
lo = idx[-1] if idx else 0
t0 = evs[lo].time_range.start
print(f"# graph replay, step total {(evs[-1].time_range.end - t0) / 1e3:.3f} ms")
for e in evs[lo:]:
    name = e.name.replace("r2d2::", "").split("(")[0][:60]
    print(f"{(e.time_range.start - t0) / 1e3:8.3f} {(e.time_range.end - t0) / 1e3:8.3f} {(e.time_range.end - e.time_range.start) / 1e3:7.3f}  {name}")
