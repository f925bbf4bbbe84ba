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


def test_struct_sizes_match_header(tmp_path):
    """sizeof of every struct as gcc sees include/go1sim.h == the ctypes mirror in lib.py (catches drift)."""
    import subprocess
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "go1sim.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %d %d\\n",'
                   'sizeof(Go1State),sizeof(Go1Diag),sizeof(Go1EnvConfig),sizeof(Go1Stats),sizeof(Go1Policy),sizeof(Go1Contact),'
                   'GO1_INFO_DIM,GO1_BIPED_INFO_DIM);return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    assert got == [ctypes.sizeof(L.State), ctypes.sizeof(L.Diag), ctypes.sizeof(L.EnvConfig), ctypes.sizeof(L.Stats),
                   ctypes.sizeof(L.Policy), ctypes.sizeof(L.Contact), L.INFO_DIM, L.BIPED_INFO_DIM]
    assert ctypes.sizeof(L.State) == 8 * (19 + 18 + 18 + 12 + 4 + 3 + 3 + 1) + 4 * 10
    assert ctypes.sizeof(L.EnvConfig) == 4 * (4 + 2 + 3 + 3 + 6 + 11 + 2 + 2 + 2 + 3 + 1 + 10 + 12 + 2)


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
    cfg = lib.default_config()
    assert list(cfg.biped_w) == pytest.approx([15, 50, 5, 8, 5, 9, 6, 0.7, 150, 25])          # go1_mujoco_env.py:92-101
    assert list(cfg.biped_ready_joints) == pytest.approx([0, 4, -2, 0, 4, -2, 0, 2.8, -1.2, 0, 2.8, -1.2])   # :32-39
    assert cfg.biped_reset_z == pytest.approx(0.65) and cfg.biped_reset_pitch_deg == -95.0
