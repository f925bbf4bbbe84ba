"""bench.py -- env-steps/s of the Go1 hot path (BASELINE.json metric) on N GPUs of one node.

    python bench.py --gpus 1 --steps 200 --warmup 20            # our CUDA path
    python bench.py --impl reference --gpus 1 --steps 3 --warmup 1   # the reference algorithm on host cores

A "step" is one control step (frame_skip=10 mj_steps + obs/reward/termination/auto-reset) of EVERY
environment of the slab: 16384 envs per GPU, torque control, random actions U[-1,1] generated on the
device (BASELINE config 3(i)); `value` = env-steps/s summed over ranks with state resident in HBM;
`e2e` = the same through Go1CudaVecEnv.step (host numpy actions in, host obs/reward/done out).
One process per GPU (torchrun), no data-path collective: the max-over-ranks timing and, every 100 control steps, the
12 rollout statistics (BASELINE config 4) are all-reduced over NCCL.  The timed window is STEADY STATE whatever --warmup
says: an untimed burn-in (burn_in) brings the slab to the stationary mix of episode phases first.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "env-steps/sec (Go1, 16k envs/GPU)"
UNIT = "env-steps/s"
ENVS_PER_GPU = 16384
ALG_BYTES_PER_ENV_STEP = 904.0     # 82 state words r+w, 12 action floats in, obs/reward/flags out
ALG_FLOP_PER_ENV_STEP = 0.67e6     # SURVEY 8(d) nominal (10 substeps x (25k + 3 Newton iterations x 14k))
# FP32 operations the kernel actually EXECUTES per env-step (fadd + fmul + 2 ffma, predicated-on threads) in the committed
# ncu capture (profiles/r1_ncu_k_step_final_raw_selected.txt): 1.885e10 per launch of 16384 envs; 3.7 iterations per substep
# instead of the nominal 3 and the redundancy of lane-parallel evaluation make up the difference to the nominal figure
NCU_EXECUTED_FLOP_PER_ENV_STEP = 1.32e6     # profiles/r2_ncu_k_step_final_raw_selected.txt (round 1, floor contacts only: 1.15e6)
FP32_FMA_MEASURED_TFLOPS = 72.1    # tools/ubench/fma.cu on this pool's B200 (148 SMs, 1965 MHz): 97 % of 148 x 128 x 2 x 1.965 GHz
BURNIN_FALL, BURNIN_MIX = 300, 100   # untimed control steps before / after the episode phases are staggered (see burn_in)
STATS_EVERY = 100                    # BASELINE config 4: the rollout statistics are all-reduced every 100 control steps


def kernel_source_sha():
    """sha256 over the kernel sources: ties an ncu capture (profiles/*traffic*.json) to the code it was taken from."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "quadruped_rl_locomotion_b200", "csrc")
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".h")) and f not in ("go1_policy_tc.cuh", "go1_render.cuh"):     # k_step does not include these two
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def ncu_traffic(n_envs, ctrl):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE k_step launch from the committed `ncu --set full` capture of THIS
    kernel source (profiles/r2_ncu_k_step_traffic.json, written by tools/ncu_traffic.py); None when the capture is stale."""
    p = os.path.join(ROOT, "profiles", "r2_ncu_k_step_traffic.json")
    try:
        t = json.load(open(p))
        if t.get("kernel_source_sha") == kernel_source_sha() and t.get("n_envs") == n_envs and t.get("ctrl") == ctrl:
            return float(t["dram_bytes_per_launch"])
    except Exception:  # noqa: BLE001
        pass
    return None


def burn_in(sim, mode, policy):
    """Brings a freshly reset slab to the stationary mix of episode phases WITHOUT timing it: the robots fall under random torques
    (BURNIN_FALL steps), the consecutive-unhealthy counters are spread over [0, 750] (go1_stagger_phases), BURNIN_MIX more steps.
    From then on ~N/750 episodes end per control step, so every timed window holds terminations, auto-resets, fresh upright
    starts and long-fallen many-contact states in their steady-state proportions."""
    sim.rollout(BURNIN_FALL, mode, policy)
    sim.stagger_phases()
    sim.rollout(BURNIN_MIX, mode, policy)


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)), "measured"
        except Exception:  # noqa: BLE001
            pass
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu: int):
        self.gpu, self.rows, self.proc = gpu, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.2] or [r for _, r in self.rows]
        sm, reasons, smax = [], set(), None
        for r in rows:
            try:
                sm.append(float(r[1])); smax = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:  # noqa: BLE001
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm)}


def _dist_env():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ------------------------------------------------------------------------------------------------
def cpu_port_rate(seconds: float, cores: int, ctrl: str = "torque"):
    """The oracle (fp64 C restatement of the reference's mj_step path) on `cores` host threads: each thread
    owns one environment and runs do_simulation (10 x mj_step + mj_rnePostConstraint) with random torques."""
    from oracle.oracle import OracleSim
    from quadruped_rl_locomotion_b200.mjcf import Model
    model = Model.load(os.path.join(ROOT, "quadruped_rl_locomotion_b200", "models", f"go1_{ctrl}.go1m"))
    sims = [OracleSim(model) for _ in range(cores)]
    lo, hi = model.actuator_ctrlrange[:, 0], model.actuator_ctrlrange[:, 1]
    counts = [0] * cores
    stop = [False]

    def work(i):
        rng = np.random.default_rng(1000 + i)
        s = sims[i]
        s.data.qpos[:] = model.key_qpos[0]
        while not stop[0]:
            s.data.ctrl[:] = rng.uniform(lo, hi)
            s.step(10)                   # ctypes releases the GIL: threads run on separate cores
            s.rne_post_constraint()
            counts[i] += 1
            if counts[i] % 750 == 0:     # episode boundary like the env's unhealthy cap
                s.reset_data(); s.data.qpos[:] = model.key_qpos[0]

    th = [threading.Thread(target=work, args=(i,)) for i in range(cores)]
    t0 = time.time()
    for t in th:
        t.start()
    time.sleep(seconds)
    stop[0] = True
    for t in th:
        t.join()
    dt = time.time() - t0
    return sum(counts) / dt, sum(counts), dt


# ------------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """--impl reference: the reference algorithm (oracle port; MuJoCo itself is not installable here) on all
    host threads.  Rank 0 only; each 'step' is a bounded sample of the workload."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = float(os.environ.get("GO1_REF_STEP_SECONDS", "4"))
    for _ in range(args.warmup):
        cpu_port_rate(min(1.0, per_step), cores)
    tot_steps, tot_t = 0, 0.0
    for _ in range(args.steps):
        _, n, dt = cpu_port_rate(per_step, cores)
        tot_steps += n; tot_t += dt
    v = tot_steps / tot_t
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / max(1, args.steps), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "Go1 torque control, random actions U[-1,1], do_simulation = 10 x mj_step + mj_rnePostConstraint",
                   "envs": cores, "note": "mujoco==2.3.0 is not installable in this image: fp64 C restatement (oracle/), physics only"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{cores} envs (one per host thread), {tot_steps} control steps in {tot_t:.1f} s"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_ours(args, rank, world, local):
    import torch
    import torch.distributed as dist
    from quadruped_rl_locomotion_b200.lib import ACTION_RANDOM, ACTION_POLICY, Go1Lib, Policy, Sim
    from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv, load_model_blob

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = args.envs_per_gpu
    lib = Go1Lib()
    cfg = lib.default_config()
    cfg.biped = 1 if args.biped else 0      # go1_mujoco_env.py:43 (train.py --biped); default off = BASELINE's configs
    sim = Sim(load_model_blob(args.ctrl), n, lib=lib, cfg=cfg, device=local, seed=1234, env_id_offset=rank * n)
    policy = None
    mode = ACTION_RANDOM
    keep = []
    if args.actions == "policy":
        # SB3-default initialised 48-64-64-12 tanh actor (no 48-d torque policy is shipped, SURVEY finding 4)
        g = torch.Generator().manual_seed(0)
        def ortho(o, i, gain):
            w = torch.empty(o, i); torch.nn.init.orthogonal_(w, gain=gain, generator=g); return w
        ws = [ortho(64, 48, 2 ** 0.5), torch.zeros(64), ortho(64, 64, 2 ** 0.5), torch.zeros(64), ortho(12, 64, 0.01), torch.zeros(12), torch.zeros(12)]
        keep = [w.cuda().contiguous() for w in ws]
        policy = Policy(*[k.data_ptr() for k in keep])
        mode = ACTION_POLICY
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    burn_in(sim, mode, policy)
    for _ in range(args.warmup):
        sim.rollout(1, mode, policy)
    barrier()
    clocks = ClockSampler(local); clocks.start()
    t0 = time.time()
    dev_ms, launches = 0.0, 0
    stats = None
    STAT_KEYS = ("sum_reward", "n_episodes", "sum_episode_length", "sum_episode_return", "n_terminated", "n_truncated", "n_bad_state",
                 "env_steps", "n_contact_overflow", "sum_contacts", "sum_solver_iterations", "sum_self_contacts")
    gstats = torch.zeros(len(STAT_KEYS), dtype=torch.float64, device="cuda")     # global (all ranks) statistics
    pending = torch.zeros_like(gstats)
    n_allreduce = 0

    def reduce_stats():
        nonlocal n_allreduce
        if world > 1:
            dist.all_reduce(pending, op=dist.ReduceOp.SUM)     # NCCL over NVLink: 12 doubles (config 4)
        gstats.add_(pending); pending.zero_()
        n_allreduce += 1

    for it in range(args.steps):
        flush.zero_()                      # L2 flush between timed iterations (not timed)
        torch.cuda.synchronize()
        st = sim.rollout(1, mode, policy)  # CUDA events on the launch stream inside go1_rollout
        dev_ms += st["device_ms"]; launches += st["launches"]
        stats = st if stats is None else {k: stats[k] + st[k] for k in st}
        pending += torch.tensor([st[k] for k in STAT_KEYS], dtype=torch.float64, device="cuda")
        if (it + 1) % STATS_EVERY == 0:
            reduce_stats()
    reduce_stats()
    barrier()
    t1 = time.time()
    clk = clocks.stop(t0, t1)
    ms = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_max = float(ms.item())
    value = world * n * args.steps / (ms_max * 1e-3)
    g = dict(zip(STAT_KEYS, gstats.tolist()))
    if args.steps >= 200:
        assert g["n_episodes"] > 0, "a steady-state window of >= 200 control steps must contain episode ends"

    # ---- e2e through the public API: host actions -> VecEnv.step -> host obs/reward/done
    sim.close()
    env = Go1CudaVecEnv(n, ctrl_type=args.ctrl, biped=bool(args.biped), device=local, seed=1234, env_id_offset=rank * n, library=lib)
    env.reset()
    burn_in(env.sim, ACTION_RANDOM, None)       # same stationary mix as the device-timed run (untimed)
    lo, hi = env.action_space.low, env.action_space.high
    rng = np.random.default_rng(rank)
    e2e_steps = max(5, min(args.steps, 50))
    # one fresh i.i.d. action batch per step, drawn BEFORE the timed region (a short cycle of repeated batches is a different,
    # slower workload: the same torques every few steps drive the robots into contact-rich limit cycles)
    acts = rng.uniform(lo, hi, size=(e2e_steps + 3, n, 12)).astype(np.float32)
    for i in range(3):
        env.step(acts[e2e_steps + i])
    barrier()
    te = time.perf_counter()
    for i in range(e2e_steps):
        obs, rew, dones, infos = env.step(acts[i])
    barrier()
    e2e_t = torch.tensor([time.perf_counter() - te], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_val = world * n * e2e_steps / float(e2e_t.item())
    env.close()

    if rank == 0:
        peaks, pk = _peaks()
        per_launch_ms = ms_max / max(1, args.steps)
        ach_gbs = ALG_BYTES_PER_ENV_STEP * n / (per_launch_ms * 1e-3) / 1e9
        sm_mhz = clk.get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        nsm = torch.cuda.get_device_properties(local).multi_processor_count
        fp32_nominal = nsm * 128 * 2 * peaks.get("sm_max_mhz", 1965.0) * 1e6 / 1e12
        fp32_peak = FP32_FMA_MEASURED_TFLOPS if nsm == 148 else fp32_nominal
        ach_tf = ALG_FLOP_PER_ENV_STEP * n / (per_launch_ms * 1e-3) / 1e12
        cores = os.cpu_count() or 1
        cpu_v, cpu_n, cpu_dt = cpu_port_rate(float(os.environ.get("GO1_CPU_BASELINE_SECONDS", "12")), cores, args.ctrl)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": per_launch_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"Go1 {args.ctrl} control{' (biped mode)' if args.biped else ''}, {n} envs/GPU, {args.actions} actions, frame_skip 10, fused obs/reward/auto-reset",
                       "envs_per_gpu": n, "parallelism": f"env-slab x{world} (no data-path collective)",
                       "l2": "flushed between timed steps (256 MiB write, untimed)", "timing": "CUDA events on the launch stream, max over ranks",
                       "state": "steady state (staggered episode phases, see steady_state)", "physics": "floor + robot-robot contacts"},
            "roofline": {"bound": "hbm", "achieved": ach_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach_gbs / peaks["hbm_gbs"],
                         "traffic": ncu_traffic(n, args.ctrl), "traffic_source": "profiles/r2_ncu_k_step_traffic.json (ncu --set full of this kernel source; null = no capture of this source)",
                         "algorithmic_bytes_per_launch": ALG_BYTES_PER_ENV_STEP * n, "peak_source": pk,
                         "note": "neither HBM nor FP32 binds: the GPC instruction cache does (gcc instruction requests at 80 % of "
                                 "peak, DESIGN.md section 3); see fp32 block",
                         "fp32": {"achieved_tflops": ach_tf, "peak_tflops": fp32_peak, "frac": ach_tf / fp32_peak,
                                  "flop_per_env_step": ALG_FLOP_PER_ENV_STEP, "executed_flop_per_env_step_ncu": NCU_EXECUTED_FLOP_PER_ENV_STEP,
                                  "executed_tflops": NCU_EXECUTED_FLOP_PER_ENV_STEP * n / (per_launch_ms * 1e-3) / 1e12,
                                  "peak": "measured FMA throughput (tools/ubench/fma.cu, profiles/r1_fma_ubench.json)" if nsm == 148 else "n_SM*128*2*f_max (nominal)",
                                  "nominal_peak_tflops": fp32_nominal, "sm_mhz_under_load": sm_mhz}},
            "cpu_baseline": {"value": cpu_v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"fp64 oracle (restatement, not MuJoCo): {cores} envs on {cores} threads, {cpu_n} control steps in {cpu_dt:.1f} s, physics only"},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": n * 12 * 4,
                    "d2h_bytes_per_step": n * (48 * 8 + 4 + 1 + 1 + 48 * 4 + 20 * 4), "steps": e2e_steps,
                    "api": "Go1CudaVecEnv.step(np.ndarray[N,12]) -> (obs[N,48] f64, rew, dones, infos)"},
            "gpu_launches": launches, "clocks": clk,
            "rollout_stats": dict(g, scope=f"all {world} ranks, NCCL all-reduce every {STATS_EVERY} control steps ({n_allreduce} all-reduces)"),
            "steady_state": {"burn_in": f"{BURNIN_FALL} untimed control steps, go1_stagger_phases, {BURNIN_MIX} more (+ --warmup)",
                             "episodes_per_step": g["n_episodes"] / max(1, args.steps),
                             "newton_iterations_per_env_step": g["sum_solver_iterations"] / max(1.0, g["env_steps"]),
                             "contacts_per_env": g["sum_contacts"] / max(1.0, g["env_steps"]),
                             "robot_robot_contacts_per_env": g["sum_self_contacts"] / max(1.0, g["env_steps"]),
                             "contacts_dropped": g["n_contact_overflow"]},
            "kernel_source_sha": kernel_source_sha(),
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_RESULT_FD = None


def own_stdout():
    """stdout must carry exactly ONE JSON line, but libraries write banners to fd 1 from C (NCCL prints
    "NCCL version ..." at communicator creation): keep a private copy of the real stdout for the result and point
    fd 1 at stderr for everything else."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    own_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs-per-gpu", type=int, default=ENVS_PER_GPU)
    ap.add_argument("--ctrl", default="torque", choices=["torque", "position"])
    ap.add_argument("--actions", default="random", choices=["random", "policy"])
    ap.add_argument("--biped", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank, world, local = _dist_env()
    if args.impl == "reference":
        if args.steps > 20:
            args.steps = 20   # bounded: each step is a multi-second CPU sample
        run_reference(args, rank, world)
        return
    run_ours(args, rank, world, local)


if __name__ == "__main__":
    main()
