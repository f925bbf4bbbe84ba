"""Error anatomy of the tcgen05 policy kernel against fp32 / tf32-rounded PyTorch references (GPU)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity_common as pc
from quadruped_rl_locomotion_b200.lib import Policy, Sim, Go1Lib

w = dict(np.load(os.path.join(pc.GOLDEN, "policy_pos_0428.npz")))
n = 4096
sim = Sim(pc.model_blob("position"), 64, lib=Go1Lib(), seed=4)
lo, hi = sim.action_bounds()
dev = [torch.from_numpy(np.ascontiguousarray(w[k], dtype=np.float32)).cuda() for k in ("w0", "b0", "w1", "b1", "w2", "b2", "log_std")]
rng = np.random.default_rng(0)
obs = torch.from_numpy(rng.normal(size=(n, 48)).astype(np.float32)).cuda()
act = torch.zeros(n, 12, device="cuda")
det = Policy(*[d.data_ptr() for d in dev[:6]], None)
sim.policy_act(det, obs, act, n=n); torch.cuda.synchronize()
torch.backends.cuda.matmul.allow_tf32 = False
def fwd(o, rnd):
    r = (lambda x: ((x.view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)) if rnd else (lambda x: x)
    h = torch.tanh(r(o).double() @ r(dev[0]).double().T + dev[1]).float()
    h = torch.tanh(r(h).double() @ r(dev[2]).double().T + dev[3]).float()
    return (r(h).double() @ r(dev[4]).double().T + dev[5]).float()
mu32, mutf = fwd(obs, False), fwd(obs, True)
lo_t, hi_t = torch.from_numpy(lo).cuda(), torch.from_numpy(hi).cuda()
clip = lambda x: torch.minimum(torch.maximum(x, lo_t), hi_t)
e32 = (act - clip(mu32)).abs(); etf = (act - clip(mutf)).abs()
print("vs fp32 ref: max %.3e mean %.3e | vs tf32-rounded-operand ref: max %.3e mean %.3e" % (e32.max(), e32.mean(), etf.max(), etf.mean()))
i = torch.argmax(e32).item(); r, c = divmod(i, 12)
print("worst element row %d (row %% 128 = %d) col %d: kernel %.6f fp32 %.6f tf32ref %.6f |mu| scale %.3f" % (r, r % 128, c, act[r, c], clip(mu32)[r, c], clip(mutf)[r, c], mu32.abs().mean()))
print("weight norms: |W0| row max %.2f |W1| row max %.2f |W2| row max %.2f" % (dev[0].norm(dim=1).max(), dev[2].norm(dim=1).max(), dev[4].norm(dim=1).max()))
rows = e32.max(dim=1).values
print("error by row%%128 (first 16):", [float("%.2e" % rows[torch.arange(n, device='cuda') % 128 == k].max().item()) for k in range(16)])
