"""A few eager single-stream learner steps at the benchmark configuration, for ncu:
python tools/prof_step.py [steps]   (no CUDA graph, no side streams: kernels appear in program order)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from pytorch_r2d2_b200 import Learner

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
torch.manual_seed(0)
L = Learner(1, {}, device="cuda:0", cell_size=512, memory_size=100_000, precision="bf16")
L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = 64, 40, 40, 5
L.sequence_length = 80
L.overlap_streams = False
class A: pass
fa = A(); fa.memory_size = 100_000; fa.cell = 512; fa.batch = 64; fa.burn_in = 40; fa.learn = 40; fa.n_step = 5
bench.fill_synthetic(L, fa, 0)
for _ in range(steps):
    loss = L.train()
torch.cuda.synchronize()
print("loss", float(loss))
