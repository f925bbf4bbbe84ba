"""Headless CUDA ray-caster for render_mode="rgb_array" (reference: train.py:53-61 creates the record env with
camera_name="tracking", width=1024, height=768; training_callback.py:686-738 collects ``render()`` frames).  The reference's OpenGL
renderer is unavailable on this box, so there is no reference image: pixel parity with MuJoCo's renderer is UNPINNED.  The kernel
is checked against the numpy restatement of the same ray-caster (oracle/render_ref.py) and through properties of the image."""
import numpy as np
import pytest

import parity_common as pc
from quadruped_rl_locomotion_b200.mjcf import Model
from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv


def test_render_refused_without_gpu_kernel(emu_lib_f32):
    """The CPU emulation build has no renderer: the call fails loudly instead of returning a blank frame."""
    from quadruped_rl_locomotion_b200.lib import Go1Error
    env = Go1CudaVecEnv(2, library=emu_lib_f32, render_mode="rgb_array", width=64, height=48)
    env.reset()
    with pytest.raises(Go1Error):
        env.render()
    env.close()
    with pytest.raises(ValueError):
        Go1CudaVecEnv(2, library=emu_lib_f32, render_mode="human")
    env = Go1CudaVecEnv(2, library=emu_lib_f32)
    assert env.render() is None and env.get_images() == [None, None]      # render_mode=None: gymnasium returns None too
    env.close()


@pytest.mark.gpu
def test_frame_matches_numpy_restatement(cuda_lib_f32):
    from oracle import render_ref
    W, H = 256, 192
    env = Go1CudaVecEnv(4, ctrl_type="torque", library=cuda_lib_f32, seed=3, render_mode="rgb_array", width=W, height=H)
    env.reset()
    rng = np.random.default_rng(0)
    for _ in range(15):
        env.step(rng.uniform(-1, 1, size=(4, 12)).astype(np.float32))
    model = Model.from_blob(pc.model_blob("torque"))
    frames = env.get_images()
    assert frames[0].shape == (H, W, 3) and frames[0].dtype == np.uint8
    bad = total = robot = 0
    for e in range(2):
        qpos = env.sim.get_state(e)["qpos"]
        for _ in range(150):
            px, py = int(rng.integers(W)), int(rng.integers(H))
            ref = render_ref.pixel(model, qpos, px, py, W, H)
            got = frames[e][py, px]
            total += 1
            bad += int(np.abs(ref.astype(int) - got.astype(int)).max() > 2)       # silhouette / checker edges may land on either side
            robot += int(abs(int(ref[0]) - int(ref[1])) <= 2 and ref[0] > 20)
    assert bad <= total * 0.04, (bad, total)
    assert robot > 5                                                              # the sample did hit the robot
    env.close()


@pytest.mark.gpu
def test_camera_tracks_the_robot_and_scene_structure(cuda_lib_f32):
    env = Go1CudaVecEnv(2, ctrl_type="torque", library=cuda_lib_f32, seed=1, render_mode="rgb_array")
    env.reset()
    for e in range(2):
        env.sim.set_state(e, qpos=Model.from_blob(pc.model_blob("torque")).key_qpos[0], qvel=np.zeros(18))
    f0 = env.render()
    assert f0.shape == (768, 1024, 3)
    # sky above, checker floor below, the robot (grey: r == g) around the image centre
    assert f0[:60].astype(int)[..., 2].mean() > f0[:60].astype(int)[..., 0].mean() + 20
    floor = f0[700:, :, :].reshape(-1, 3)
    assert len(np.unique(floor, axis=0)) >= 2 and (floor[:, 2] >= floor[:, 0]).all()
    grey = (np.abs(f0[..., 0].astype(int) - f0[..., 1].astype(int)) <= 1) & (f0[..., 0] > 25)
    ys, xs = np.nonzero(grey)
    assert len(xs) > 5000 and abs(xs.mean() - 512) < 60 and 250 < ys.mean() < 600
    # trackcom: moving the robot 3 m along x leaves it at the same place in the image (the floor pattern moves instead)
    s = env.sim.get_state(0)
    q = s["qpos"].copy(); q[0] += 3.0; q[1] -= 1.25
    env.sim.set_state(0, qpos=q)
    f1 = env.render()
    grey1 = (np.abs(f1[..., 0].astype(int) - f1[..., 1].astype(int)) <= 1) & (f1[..., 0] > 25)
    assert (grey ^ grey1).sum() < 0.02 * grey.sum()
    # a different pose gives a different silhouette
    q2 = s["qpos"].copy(); q2[7:] += 0.4
    env.sim.set_state(0, qpos=q2)
    f2 = env.render()
    grey2 = (np.abs(f2[..., 0].astype(int) - f2[..., 1].astype(int)) <= 1) & (f2[..., 0] > 25)
    assert (grey ^ grey2).sum() > 0.05 * grey.sum()
    # device output buffer + several environments in one launch
    import torch
    out = torch.zeros(2, 96, 128, 3, dtype=torch.uint8, device="cuda"); torch.cuda.synchronize()
    env.sim.render((0, 1), 128, 96, out=out)
    assert out.float().std().item() > 5
    env.close()
