import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import net_oracle as NO
N,H=256,512
warm = len(sys.argv) > 1
if warm:
    with torch.no_grad():
        pw = NO.init_params(1, 64 if sys.argv[1]=="small" else H)
        xw = torch.rand(8 if sys.argv[1]=="small" else N, 4, 84, 84)
        NO.qnet_forward(pw, xw, None, None)
p = NO.init_params(11, H)
x = torch.from_numpy(np.random.RandomState(4).randint(0,256,(N,4,84,84)).astype(np.float32)/np.float32(255))
with torch.no_grad():
    a = NO.qnet_forward(p, x, None, None)[3][0]
    b = NO.qnet_forward(p, x, None, None)[3][0]
print(f"{float((a-b).abs().max()/b.abs().max()):.2e}")
