"""Diagnostic: per-frame phase timing of the shifted-window conv2 kernel (CTA 0, first 32 frames).
Needs a library built with R2D2_NVCC_EXTRA=-DSWC_TRACE (python -m pytorch_r2d2_b200.build)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pytorch_r2d2_b200 import QNet, _lib, engine

torch.manual_seed(0)
net = QNet("cuda:0", 512)
w = net._weights()
R = 5440
a1 = torch.rand(R, 32, 20, 20, device="cuda")
a2 = torch.empty(R, 32, 9, 9, device="cuda"); feat = torch.empty(R, 1568, device="cuda")
for _ in range(3):
    engine.torso_fwd_tail(a1, w, a2, feat)
torch.cuda.synchronize()
lib = _lib.load()
buf = (ctypes.c_longlong * 512)()
rc = lib.r2d2_dbg_swc_trace(buf)
names = ["stage free (producer)", "producer arrived", "issuer0 start", "issuer0 issued", "issuer1 issued", "epilogue sees acc", "epilogue done", "producer stores issued"]
for layer in (2, 3):
    t = torch.tensor(list(buf), dtype=torch.float64).reshape(2, 32, 8)[layer - 2]
    print(f"conv{layer}: frame period {float((t[5:28,0]-t[4:27,0]).mean()):.0f} ns (CTA 0, frames 4..27)")
    for i, n in enumerate(names):
        # stamps of frame f relative to the moment ITS stage was seen free by the producer
        print(f"  {n:>24s}  +{float((t[4:28, i] - t[4:28, 0]).mean()):8.0f} ns")
    print(f"  issuer0: start -> issued {float((t[4:28,3]-t[4:28,2]).mean()):.0f} ns; acc seen -> epilogue done {float((t[4:28,6]-t[4:28,5]).mean()):.0f} ns")
