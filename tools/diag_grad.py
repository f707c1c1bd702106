"""Diagnostic: per-key / per-channel gradient errors of one Learner step vs the oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from oracle import net_oracle as NO
from util import make_oracle_replay, mirror_to_gpu, rel_err
from pytorch_r2d2_b200 import Learner

H, B, BI, LN, NS = [int(x) for x in sys.argv[1:6]] if len(sys.argv) > 5 else (256, 5, 2, 3, 1)
M = 80
rep, _ = make_oracle_replay(M, H, NS, BI, LN, seed=40 + H)
torch.manual_seed(7)
L = Learner(1, {}, device="cuda:0", cell_size=H, memory_size=M)
L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = B, BI, LN, NS
L.replay_memory = mirror_to_gpu(rep, B)
p = {k: v.cpu().clone() for k, v in L.net.state_dict().items()}
pt = {k: v.cpu().clone() for k, v in L.target_net.state_dict().items()}
u = np.random.RandomState(H).rand(B)
si = rep.sample_indices(u)
batch = NO.to_torch_batch(rep.gather(si)["batch"])
res = NO.train_step(p, pt, batch, burn_in=BI, n_step=NS)
L.train(u=torch.from_numpy(u), next_action_override=res["next_action"].cuda())
for k in NO.PARAM_KEYS:
    g = L.net.layout.view(L.optim.grad, k).cpu()
    r = res["grads"][k]
    print(f"{k:24s} rel_err {rel_err(g, r):.3e}")
    if k.startswith("vis_layers") and k.endswith("weight"):
        e = (g - r).abs().reshape(32, -1).max(1).values / r.abs().max()
        bad = [(i, float(x)) for i, x in enumerate(e) if x > 1e-5]
        print("   bad out-channels:", bad)
# pre-activation magnitudes near zero at conv1 for the learn frames
frames = batch["state"][BI:].reshape(-1, 4, 84, 84)
z1 = torch.nn.functional.conv2d(frames, p["vis_layers.0.weight"], p["vis_layers.0.bias"], stride=4)
a1 = L._ws["a1_on"][BI * B:(BI + LN) * B].cpu()
flip = ((z1 > 0) != (a1 > 0))
print("conv1 relu sign mismatches:", int(flip.sum()), "min |z1|:", float(z1.abs().min()))
if flip.any():
    idx = flip.nonzero()
    print(idx[:10], z1[flip][:10], a1[flip][:10])
z2 = torch.nn.functional.conv2d(torch.relu(z1), p["vis_layers.2.weight"], p["vis_layers.2.bias"], stride=2)
a2 = L._ws["a2_on"][BI * B:(BI + LN) * B].cpu()
print("conv2 relu sign mismatches:", int(((z2 > 0) != (a2 > 0)).sum()))
