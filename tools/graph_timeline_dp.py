"""CUPTI timeline of one CUDA-graph-replayed DATA-PARALLEL learner step (rank 0's view), launched
with torchrun:  python -m torch.distributed.run --nproc-per-node N tools/graph_timeline_dp.py > out.txt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch.profiler import profile, ProfilerActivity
import bench
from pytorch_r2d2_b200 import Learner

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
torch.manual_seed(0)
L = Learner(1, {}, device=f"cuda:{local}", cell_size=512, memory_size=100_000, precision="bf16")
L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = 64, 40, 40, 5
L.sequence_length = 80
assert L.enable_peer_exchange()
L.use_cuda_graph = True
class A: pass
fa = A(); fa.memory_size = 100_000; fa.cell = 512; fa.batch = 64; fa.burn_in = 40; fa.learn = 40; fa.n_step = 5
bench.fill_synthetic(L, fa, rank)
torch.manual_seed(1234 + rank)
dist.barrier(); torch.cuda.synchronize()
for _ in range(8):
    L.train()
torch.cuda.synchronize(); dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(4):
        L.train()
    torch.cuda.synchronize()
dist.barrier()
if rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    idx = [i for i, e in enumerate(evs) if "tree_sample" in e.name]
    lo, prev = idx[-1], idx[-2]
    t0 = evs[lo].time_range.start
    print(f"# world {world}: graph replay, step period {(t0 - evs[prev].time_range.start) / 1e3:.3f} ms, "
          f"last step's kernels span {(evs[-1].time_range.end - t0) / 1e3:.3f} ms")
    for e in evs[lo:]:
        name = e.name.replace("r2d2::", "").split("(")[0][:60]
        print(f"{(e.time_range.start - t0) / 1e3:8.3f} {(e.time_range.end - t0) / 1e3:8.3f} {(e.time_range.end - e.time_range.start) / 1e3:7.3f}  {name}")
dist.destroy_process_group()
