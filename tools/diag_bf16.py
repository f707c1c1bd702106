"""Diagnostic: bf16-mode gradients vs (a) the fp32 oracle, (b) the bf16-emulating oracle."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from oracle import net_oracle as NO
from util import make_oracle_replay, mirror_to_gpu, rel_err
from pytorch_r2d2_b200 import Learner

def l2(a, b):
    a = a.detach().double().cpu().reshape(-1); b = b.detach().double().cpu().reshape(-1)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))

for (H, B, BI, LN, NS, M) in [(256, 4, 3, 4, 2, 80), (512, 8, 4, 4, 5, 80), (512, 32, 8, 8, 5, 400)]:
    rep, _ = make_oracle_replay(M, H, NS, BI, LN, seed=40 + H)
    torch.manual_seed(7)
    L = Learner(1, {}, device="cuda:0", cell_size=H, memory_size=M, precision="bf16")
    L.batch_size, L.burn_in_length, L.learning_length, L.bootstrap_steps = B, BI, LN, NS
    L.replay_memory = mirror_to_gpu(rep, B)
    p = {k: v.cpu().clone() for k, v in L.net.state_dict().items()}
    pt = {k: v.cpu().clone() for k, v in L.target_net.state_dict().items()}
    u = np.random.RandomState(H).rand(B)
    si = rep.sample_indices(u)
    tb = NO.to_torch_batch(rep.gather(si)["batch"])
    r32 = NO.train_step(p, pt, tb, burn_in=BI, n_step=NS)
    r16 = NO.train_step(p, pt, tb, burn_in=BI, n_step=NS, tc_emulation=("proj", "head"),
                        next_action_override=r32["next_action"])
    L.train(u=torch.from_numpy(u), next_action_override=r32["next_action"].cuda())
    print(f"== H={H} B={B} L={LN}: q_all vs fp32 {rel_err(L.last['q_all'], r32['q_all']):.2e}  vs emu {rel_err(L.last['q_all'], r16['q_all']):.2e}")
    for k in NO.PARAM_KEYS:
        g = L.net.layout.view(L.optim.grad, k)
        print(f"  {k:22s} max-rel vs fp32 {rel_err(g, r32['grads'][k]):.2e} (l2 {l2(g, r32['grads'][k]):.2e}) | vs emu {rel_err(g, r16['grads'][k]):.2e} (l2 {l2(g, r16['grads'][k]):.2e}) | emu vs fp32 {rel_err(r16['grads'][k], r32['grads'][k]):.2e}")
