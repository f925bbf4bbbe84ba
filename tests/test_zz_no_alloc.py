"""``go1_step_noalloc`` (include/go1sim.h): the control step with the kernel's scratch in a file-scope ``static __device__`` array, for
callers that cannot reach cudaMalloc.  No counterpart in the reference, so the checker
is the existing entry point ``go1_step``: two simulators created from the same seed, one stepped through each entry point with the
same actions, must return the same bytes (the tolerance of ``test_determinism_and_env_independence``: a step's results do not depend
on the queue order, hence not on which scratch holds it), on two calls in a row (the second call finds the first call's contents in
the shared array).  This file is collected last on purpose."""
import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.lib import Go1Error, Sim

NOALLOC_MAX_ENVS = 16384      # GO1_NOALLOC_MAX_ENVS
OUT = (("obs", 48, np.float32), ("rew", 0, np.float32), ("done", 0, np.uint8), ("trunc", 0, np.uint8), ("tobs", 48, np.float32), ("info", 20, np.float32))


def _twin_steps(lib, n, make, fetch, put, stream=None, sync=lambda: None, n_calls=2, ctrl="torque", seed=21):
    """Steps twin simulators n_calls times: A through go1_step, B through go1_step_noalloc.  `make(shape, dtype)` creates a caller
    buffer, `put(buf, array)` fills it, `fetch(buf)` returns it as numpy; `stream` / `sync`: the CUDA stream of the calls."""
    a = Sim(pc.model_blob(ctrl), n, lib=lib, seed=seed)
    b = Sim(pc.model_blob(ctrl), n, lib=lib, seed=seed)
    lo, hi = a.action_bounds()
    rng = np.random.default_rng(5)
    act = make((n, 12), np.float32)
    bufs = [{k: make((n, w) if w else (n,), dt) for k, w, dt in OUT} for _ in range(2)]
    res = []
    for _ in range(n_calls):
        act_np = rng.uniform(lo, hi, size=(n, 12)).astype(np.float32)
        put(act, act_np)
        a.step(act, *[bufs[0][k] for k, _, _ in OUT], stream)
        b.step_noalloc(act, *[bufs[1][k] for k, _, _ in OUT], stream)       # same stream: the calls do not overlap on the GPU
        sync()
        res.append(({k: fetch(v).copy() for k, v in bufs[0].items()}, {k: fetch(v).copy() for k, v in bufs[1].items()}))
    sa, sb = a.get_state(n - 1), b.get_state(n - 1)
    a.close(); b.close()
    return res, sa, sb


def _assert_same(res, sa, sb):
    assert len(res) == 2
    for ra, rb in res:
        for k in ra:
            assert np.array_equal(ra[k], rb[k]), k
        assert np.isfinite(ra["obs"]).all() and np.abs(ra["obs"]).max() > 0
    assert not np.array_equal(res[0][0]["obs"], res[1][0]["obs"])          # the two calls are two different steps
    assert np.array_equal(sa["qpos"], sb["qpos"]) and np.array_equal(sa["qvel"], sb["qvel"])


def test_noalloc_entry_point_matches_go1_step_emu(emu_lib_f32):
    """CPU (SIMT emulation build of the same source): the host logic of the entry point and the shared scratch."""
    def put(buf, arr):
        buf[:] = arr
    res, sa, sb = _twin_steps(emu_lib_f32, 7, lambda s, d: np.zeros(s, d), lambda v: v, put)
    _assert_same(res, sa, sb)


@pytest.mark.gpu
def test_noalloc_entry_point_matches_go1_step(cuda_lib_f32):
    import torch
    tdt = {np.float32: torch.float32, np.uint8: torch.uint8}
    make = lambda s, d: torch.zeros(*s, dtype=tdt[d], device="cuda")
    fetch = lambda v: v.cpu().numpy()
    st = torch.cuda.Stream()                       # a stream of the caller's: NULL would mean "the handle's own stream"

    def put(buf, arr):
        with torch.cuda.stream(st):
            buf.copy_(torch.from_numpy(arr).cuda(), non_blocking=False)
        st.synchronize()
    # a ragged size, then the documented maximum (the second pair of calls finds the first pair's work keys in the array)
    for n in (1000 + 3, NOALLOC_MAX_ENVS):
        res, sa, sb = _twin_steps(cuda_lib_f32, n, make, fetch, put, stream=st.cuda_stream, sync=st.synchronize)
        _assert_same(res, sa, sb)
    # beyond the maximum: a clear error, no launch
    big = Sim(pc.model_blob("torque"), NOALLOC_MAX_ENVS + 1, lib=cuda_lib_f32, seed=1)
    n = NOALLOC_MAX_ENVS + 1
    bufs = [make((n, 12), np.float32)] + [make((n, w) if w else (n,), dt) for _, w, dt in OUT]
    with pytest.raises(Go1Error, match="GO1_NOALLOC_MAX_ENVS"):
        big.step_noalloc(*bufs, st.cuda_stream)
    big.step(*bufs, st.cuda_stream); st.synchronize()                       # the handle is still usable through go1_step
    big.close()
