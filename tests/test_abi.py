"""The C-ABI library loads and exports every symbol include/go1sim.h declares (no compute calls: CPU box)."""
import ctypes
import os
import re

from quadruped_rl_locomotion_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "go1sim.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(go1_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_bound():
    assert set(_declared()) == set(L.EXPORTS)


def test_libraries_export_all_symbols():
    import __graft_entry__ as g
    g.build()
    for prec in ("f32", "f64"):
        path = L.library_path(prec)
        assert os.path.exists(path)
        dll = ctypes.CDLL(path)
        for name in _declared():
            assert hasattr(dll, name), (path, name)
        dll.go1_precision_bits.restype = ctypes.c_int
        assert dll.go1_precision_bits() == (32 if prec == "f32" else 64)


def test_struct_sizes_match_header():
    # sizes implied by the header's field lists (catches drift between go1sim.h and lib.py)
    assert ctypes.sizeof(L.State) == 8 * (19 + 18 + 18 + 12 + 4 + 3 + 3 + 1) + 4 * 8
    assert ctypes.sizeof(L.Diag) == 8 * (18 * 3 + 14 * 6 + 1 + 3) + 4 * 4
    assert ctypes.sizeof(L.EnvConfig) == 4 * (4 + 2 + 3 + 3 + 6 + 11 + 2 + 2 + 2 + 3 + 1 + 8)
    assert ctypes.sizeof(L.Stats) == 64 and ctypes.sizeof(L.Policy) == 56


def test_missing_library_fails_loudly(tmp_path):
    import pytest
    with pytest.raises(L.Go1Error):
        L.Go1Lib(str(tmp_path / "libgo1sim.so"))


def test_create_rejects_garbage_without_gpu():
    """Model validation happens before any CUDA call, so it is testable on the CPU box."""
    lib = L.Go1Lib()
    import pytest
    with pytest.raises(L.Go1Error, match="model"):
        L.Sim(b"GO1MDL01" + b"\0" * 64, 4, lib=lib)
    cfg = lib.default_config(); cfg.biped = 1
    blob = open(os.path.join(ROOT, "quadruped_rl_locomotion_b200", "models", "go1_torque.go1m"), "rb").read()
    with pytest.raises(L.Go1Error, match="biped"):
        L.Sim(blob, 4, lib=lib, cfg=cfg)
