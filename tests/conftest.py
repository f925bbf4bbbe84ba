import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


EMU_DIR = os.path.join(ROOT, "tests", "emu")
CSRC = os.path.join(ROOT, "quadruped_rl_locomotion_b200", "csrc")


def build_emu(prec: str) -> str:
    """g++ build of the CUDA source against tests/emu/cuda_emu.h (CPU SIMT emulation; test-only)."""
    out = os.path.join(EMU_DIR, f"libgo1sim_emu_{prec}.so")
    deps = [os.path.join(CSRC, f) for f in ("go1sim.cu", "go1_warp.cuh", "go1_pair.cuh", "go1_model.h")] + \
           [os.path.join(EMU_DIR, "cuda_emu.h"), os.path.join(ROOT, "include", "go1sim.h")]
    if not os.path.exists(out) or any(os.path.getmtime(d) > os.path.getmtime(out) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DGO1_EMU", "-DGO1_WARPS=1",
                               f"-DGO1_REAL={prec}", "-I", EMU_DIR, "-x", "c++", os.path.join(CSRC, "go1sim.cu"),
                               "-o", out, "-lpthread"])
    return out


@pytest.fixture(scope="session")
def emu_lib_f32():
    from quadruped_rl_locomotion_b200.lib import Go1Lib
    return Go1Lib(build_emu("float"))


@pytest.fixture(scope="session")
def emu_lib_f64():
    from quadruped_rl_locomotion_b200.lib import Go1Lib
    return Go1Lib(build_emu("double"))


@pytest.fixture(scope="session")
def cuda_lib_f32():
    from quadruped_rl_locomotion_b200.lib import Go1Lib
    return Go1Lib(precision="f32")


@pytest.fixture(scope="session")
def cuda_lib_f64():
    from quadruped_rl_locomotion_b200.lib import Go1Lib
    return Go1Lib(precision="f64")
