"""Stress: the 256-actor step must be bit-reproducible (any difference between repeated identical
steps = a race).  python tools/stress_actor.py [precision] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pytorch_r2d2_b200 import BatchedActor

prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
N, H = 256, 512
torch.manual_seed(11)
for graph in (True, False):
    act = BatchedActor(N, "cuda:0", H, precision=prec)
    act.use_cuda_graph = graph
    x = torch.from_numpy(np.random.RandomState(4).randint(0, 256, (N, 4, 84, 84)).astype(np.float32) / np.float32(255)).cuda()
    ref = None
    bad = 0
    for r in range(reps):
        act.reset()
        out = act.select_action(x, rng=np.random.RandomState(0))
        cur = [t.clone() for t in out[1:]]
        torch.cuda.synchronize()
        if ref is None:
            ref = cur
        else:
            d = [float((a - b).abs().max()) for a, b in zip(cur, ref)]
            if max(d) > 0:
                bad += 1
                if bad <= 5:
                    print(f"graph={graph} rep {r}: max diffs {d}")
    print(f"precision {prec} graph={graph}: {bad} of {reps - 1} repeats differ")
