"""Diagnostic: per-step phase timing of the persistent tensor-core BPTT kernel (alone on the GPU)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from pytorch_r2d2_b200 import _lib, engine

H, B, T = 512, 64, 40
dhs = torch.randn(T, B, H, device="cuda") * 0.01
w_hh = torch.randn(4 * H, H, device="cuda") * 0.05
gates = torch.rand(T, B, 4 * H, device="cuda")
cs = torch.randn(T + 1, B, H, device="cuda")
dgates = torch.empty(T, B, 4 * H, device="cuda")
scratch = torch.empty(engine.tc_lstm_bwd_scratch_bytes(B, H), dtype=torch.uint8, device="cuda")
trace = torch.zeros(T, 32, dtype=torch.int64, device="cuda")
run = lambda: engine.tc_lstm_seq_bwd(dhs, w_hh, gates, cs, dgates, T, B, H, scratch)
for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    run()
e1.record(); torch.cuda.synchronize()
print(f"T={T}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per launch, {e0.elapsed_time(e1) / 10 / T * 1e3:.2f} us per step")
_lib.call("r2d2_tc_lstm_set_trace", _lib.ptr(trace))
run()
torch.cuda.synchronize()
_lib.call("r2d2_tc_lstm_set_trace", None)
tr = trace.cpu().double()[5:35]
print(f"steady-state step period {float((tr[1:, 0] - tr[:-1, 0]).mean()):.0f} ns")
names = {0: "barrier passed", 1: "all loads issued", 9: "first group landed",
         10: "last group landed", 2: "last MMA issued", 3: "accumulators complete", 4: "TMEM read",
         5: "gate math + stores issued", 6: "CTA barrier", 7: "arrive issued"}
# stamps 3..7 of iteration s belong to the recurrent term loaded in iteration s (same s)
for k in (1, 9, 10, 2, 3, 4, 5, 6, 7):
    print(f"  {names[k]:>28s}  +{float((tr[:, k] - tr[:, 0]).mean()):8.0f} ns after the barrier")
print(f"  {'next barrier passed':>28s}  +{float((tr[1:, 0] - tr[:-1, 0]).mean()):8.0f} ns")
