"""N > 1 host logic on CPU: two ranks over ``gloo``, one slab each (emulation build of the CUDA
source), Philox stream = global env id, and the ONLY collectives of the path: max-over-ranks timing and the
8-element rollout-statistics all-reduce.  The sharded result must equal the single-slab result."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity_common as pc
from conftest import build_emu
from quadruped_rl_locomotion_b200.lib import Go1Lib, Sim

STAT_KEYS = ("sum_reward", "n_episodes", "sum_episode_length", "sum_episode_return", "n_terminated", "n_truncated",
             "n_bad_state", "env_steps")
N_PER_RANK, STEPS = 6, 4


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, libpath, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sim = Sim(pc.model_blob("torque"), N_PER_RANK, lib=Go1Lib(libpath), seed=77, env_id_offset=rank * N_PER_RANK)
    sim.reset_host()
    tot = np.zeros(len(STAT_KEYS))
    ms = 0.0
    for _ in range(STEPS):
        st = sim.rollout(1)
        tot += [st[k] for k in STAT_KEYS]
        ms += st["device_ms"]
    t = torch.tensor(tot, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)                   # rollout statistics (config 4)
    tm = torch.tensor([ms + rank], dtype=torch.float64)        # + rank: make the max identifiable
    dist.all_reduce(tm, op=dist.ReduceOp.MAX)                  # bench.py: max over ranks
    q = np.array([sim.get_state(e)["qpos"] for e in range(N_PER_RANK)])
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), stats=t.numpy(), tmax=tm.numpy(), qpos=q, ms=ms)
    sim.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_slabs_equal_one_slab(tmp_path):
    libpath = build_emu("float")
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), libpath, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(tmp_path / f"rank{k}.npz") for k in range(world)]
    assert np.array_equal(r[0]["stats"], r[1]["stats"])        # every rank holds the reduced vector
    assert r[0]["tmax"][0] == r[1]["tmax"][0] == pytest.approx(max(r[0]["ms"], r[1]["ms"] + 1))
    # single process, one slab of 2 x N envs
    sim = Sim(pc.model_blob("torque"), world * N_PER_RANK, lib=Go1Lib(libpath), seed=77)
    sim.reset_host()
    tot = np.zeros(len(STAT_KEYS))
    for _ in range(STEPS):
        st = sim.rollout(1)
        tot += [st[k] for k in STAT_KEYS]
    q = np.array([sim.get_state(e)["qpos"] for e in range(world * N_PER_RANK)])
    sim.close()
    assert np.array_equal(q[:N_PER_RANK], r[0]["qpos"]) and np.array_equal(q[N_PER_RANK:], r[1]["qpos"])
    assert r[0]["stats"][7] == world * N_PER_RANK * STEPS
    assert np.allclose(r[0]["stats"], tot, rtol=1e-6)


def test_bench_reference_arm_skips_nonzero_ranks():
    """bench.py --impl reference under torchrun: rank 0 alone works and prints; other ranks exit 0 silently."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""
