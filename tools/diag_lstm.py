"""Diagnostic: per-step phase timing of the persistent tensor-core LSTM kernel (alone on the GPU)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from pytorch_r2d2_b200 import _lib, engine

H, B, T = 512, 64, 80
xprojT = (torch.randn(4 * H, T * B, device="cuda") * 0.5).contiguous()
w_hh = torch.randn(4 * H, H, device="cuda") * 0.05
hs = torch.zeros(T + 1, B, H, device="cuda"); cs = torch.zeros(T + 1, B, H, device="cuda")
gates = torch.empty(T, B, 4 * H, device="cuda")
scratch = torch.empty(engine.tc_lstm_scratch_bytes(B, H), dtype=torch.uint8, device="cuda")
trace = torch.zeros(T, 32, dtype=torch.int64, device="cuda")
for _ in range(3):
    engine.tc_lstm_seq_fwd(xprojT, 0, w_hh, hs, cs, gates, T, B, H, scratch)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    engine.tc_lstm_seq_fwd(xprojT, 0, w_hh, hs, cs, gates, T, B, H, scratch)
e1.record(); torch.cuda.synchronize()
print(f"T={T}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per launch, {e0.elapsed_time(e1) / 10 / T * 1e3:.2f} us per step")
_lib.call("r2d2_tc_lstm_set_trace", _lib.ptr(trace))
engine.tc_lstm_seq_fwd(xprojT, 0, w_hh, hs, cs, gates, T, B, H, scratch)
torch.cuda.synchronize()
_lib.call("r2d2_tc_lstm_set_trace", None)
tr = trace.cpu().double()[10:70]
names = ["barrier passed", "proxy fence", "MMAs issued", "acc ready", "gates staged", "stores issued", "threadfence+bar", "arrive issued"]
step = (tr[1:, 0] - tr[:-1, 0]).mean()
print(f"steady-state step period {step:.0f} ns")
for i in range(1, 8):
    print(f"  {names[i-1]:>18s} -> {names[i]:<18s} {float((tr[:, i] - tr[:, i-1]).mean()):8.0f} ns")
print(f"  {'arrive issued':>18s} -> {'next barrier passed':<18s} {float((tr[1:, 0] - tr[:-1, 7]).mean()):8.0f} ns")

print("issuer 0, relative to 'proxy fence' (ns): MMAs of k-block issued")
for kb in (0, 4):
    print(f"  kb {kb}: issued {float((tr[:, 16 + kb] - tr[:, 1]).mean()):7.0f}")
