"""Device time of the policy forward (row a21) at n rows: tcgen05 kernel (actor only / actor + critic), the fp32 scalar kernel,
and plain PyTorch fp32 (two cuBLAS GEMM chains + elementwise kernels).  CUDA events on a non-default torch stream that is also
passed to the library (a NULL stream means "the handle's own stream", which torch events do not see); 200 launches each after 20
warm-ups; inputs are re-read from HBM / L2 every launch.

    python tools/policy_bench.py [n=16384]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from quadruped_rl_locomotion_b200.lib import Critic, Go1Lib, Policy, Sim  # noqa: E402
from quadruped_rl_locomotion_b200.vec_env import load_model_blob  # noqa: E402

STREAM = None


def timed(f, k=200):
    with torch.cuda.stream(STREAM):
        for _ in range(20):
            f()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(STREAM)
        for _ in range(k):
            f()
        b.record(STREAM); torch.cuda.synchronize()
    return a.elapsed_time(b) / k * 1e3


def main():
    global STREAM
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    sim = Sim(load_model_blob("torque"), 64, lib=Go1Lib(os.environ.get("GO1_LIB")) if os.environ.get("GO1_LIB") else Go1Lib(), seed=1)
    STREAM = torch.cuda.Stream()
    s = STREAM.cuda_stream
    assert s != 0
    g = torch.Generator().manual_seed(0)
    mk = lambda *sh: (torch.randn(*sh, generator=g) * 0.2).cuda().contiguous()
    aw = [mk(64, 48), mk(64), mk(64, 64), mk(64), mk(12, 64), mk(12), torch.full((12,), -1.0).cuda()]
    cw = [mk(64, 48), mk(64), mk(64, 64), mk(64), mk(1, 64), mk(1)]
    pol, cri = Policy(*[t.data_ptr() for t in aw]), Critic(*[t.data_ptr() for t in cw])
    obs = torch.randn(n, 48, generator=g).cuda(); act = torch.zeros(n, 12, device="cuda")
    val = torch.zeros(n, device="cuda"); lp = torch.zeros(n, device="cuda")
    torch.backends.cuda.matmul.allow_tf32 = False

    def torch_fwd():
        h = torch.tanh(torch.addmm(aw[1], obs, aw[0].T)); h = torch.tanh(torch.addmm(aw[3], h, aw[2].T)); mu = torch.addmm(aw[5], h, aw[4].T)
        hc = torch.tanh(torch.addmm(cw[1], obs, cw[0].T)); hc = torch.tanh(torch.addmm(cw[3], hc, cw[2].T)); v = torch.addmm(cw[5], hc, cw[4].T)
        a = mu + torch.exp(aw[6]) * torch.randn_like(mu)
        return a.clamp(-1, 1), v
    res = {
        "tcgen05 actor + critic (go1_policy_forward)": timed(lambda: sim.policy_forward(pol, cri, obs, act, val, lp, n=n, stream=s)),
        "tcgen05 actor only (go1_policy_act)": timed(lambda: sim.policy_act(pol, obs, act, n=n, stream=s)),
        "fp32 scalar kernel, actor only (go1_policy_act_scalar)": timed(lambda: sim.policy_act_scalar(pol, obs, act, n=n, stream=s)),
        "PyTorch fp32 eager, actor + critic": timed(torch_fwd),
    }
    flop = 2 * n * (48 * 64 + 64 * 64 + 64 * 12 + 48 * 64 + 64 * 64 + 64)
    byt = n * (48 + 12 + 1 + 1) * 4
    for k, v in res.items():
        print(f"{k:58s} {v:8.1f} us")
    t = res["tcgen05 actor + critic (go1_policy_forward)"] * 1e-6
    print(f"n = {n}: {flop / 1e6:.1f} MFLOP, {byt / 1e6:.2f} MB in + out -> {flop / t / 1e12:.2f} TFLOP/s, {byt / t / 1e9:.0f} GB/s")
    sim.close()


if __name__ == "__main__":
    main()
