"""Host-side cost of the e2e loop's calls (bench.time_e2e): perf_counter around Learner.train(),
ReplayMemory.extend(overlap=True) and the loss read-back, plus the device time of the same steps."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
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
dump = bench.make_host_dump(640, 512)
for _ in range(5):
    L.train(); L.replay_memory.extend(dump, overlap=True)
torch.cuda.synchronize()
N = 30
tt = te = 0.0
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter(); e0.record()
for i in range(N):
    a = time.perf_counter(); loss = L.train(); b = time.perf_counter()
    L.replay_memory.extend(dump, overlap=True); c = time.perf_counter()
    tt += b - a; te += c - b
e1.record(); torch.cuda.synchronize()
wall = time.perf_counter() - t0
print(f"per step: wall {wall/N*1e3:.3f} ms  device {e0.elapsed_time(e1)/N:.3f} ms  host train() {tt/N*1e3:.3f} ms  host extend() {te/N*1e3:.3f} ms")
t0 = time.perf_counter(); e0.record()
for i in range(N):
    L.train()
e1.record(); torch.cuda.synchronize()
print(f"train only: wall {(time.perf_counter()-t0)/N*1e3:.3f} ms device {e0.elapsed_time(e1)/N:.3f} ms")
# ---- bench.time_e2e's loop (loss read back one step behind), with host timers per statement
loss_pin = [torch.empty(1).pin_memory() for _ in range(2)]
loss_ev = [torch.cuda.Event() for _ in range(2)]
tt = te = tc = ts = 0.0
torch.cuda.synchronize()
t0 = time.perf_counter(); e0.record()
for i in range(N):
    a = time.perf_counter(); loss = L.train(); b = time.perf_counter()
    L.replay_memory.extend(dump, overlap=True); c = time.perf_counter()
    loss_pin[i & 1].copy_(loss, non_blocking=True); loss_ev[i & 1].record(); d = time.perf_counter()
    if i > 0:
        loss_ev[(i - 1) & 1].synchronize(); float(loss_pin[(i - 1) & 1])
    e = time.perf_counter()
    tt += b - a; te += c - b; tc += d - c; ts += e - d
e1.record(); torch.cuda.synchronize()
wall = time.perf_counter() - t0
print(f"bench loop: wall {wall/N*1e3:.3f} ms device {e0.elapsed_time(e1)/N:.3f} ms | host train {tt/N*1e3:.3f} extend {te/N*1e3:.3f} copy {tc/N*1e3:.3f} sync {ts/N*1e3:.3f}")
