"""Data formats and schedules of the reference's callbacks: training_callback.py:47-69 (curriculum),
:300-331 / :472-520 (training_history.json + best_model.zip), :452-466 (failure analysis)."""
import json
import os

import numpy as np
import pytest

from quadruped_rl_locomotion_b200 import callbacks as cb


def test_curriculum_schedule_matches_reference_formula():
    total, init = 5_000_000, 1.0
    end = int(total * 0.7)
    assert cb.curriculum_rand_power(0, total, init) == 1.0
    assert cb.curriculum_rand_power(end // 2, total, init) == pytest.approx(0.5, abs=1e-6)
    assert cb.curriculum_rand_power(end - 1, total, init) == pytest.approx(1.0 / end, rel=1e-6)
    assert cb.curriculum_rand_power(end, total, init) == 0.0 and cb.curriculum_rand_power(total, total, init) == 0.0

    class Env:                                   # what CurriculumCallback talks to (training_env.set_attr, :64)
        def __init__(self): self.calls = []
        def set_attr(self, name, value, indices=None): self.calls.append((name, value))
    e = Env()
    s = cb.CurriculumSchedule(e, total, 0.8)
    assert s(end // 4) == pytest.approx(0.6) and e.calls[-1] == ("_rand_power", pytest.approx(0.6))


def test_training_history_file_format_and_best_model(tmp_path):
    d = str(tmp_path / "run")
    saved = []
    h = cb.TrainingHistory(best_model_save_path=d)
    new = h.update(300_000, rewards=[10.0, 14.0], lengths=[100, 140], components=[{"upright": 0.5}, {"upright": 0.7}],
                   stability=[{"angular_stability": 0.1, "cop_stability": 0.2, "zmp_stability": 0.3}], failures=["too_low"], successes=1,
                   save_model=lambda p: saved.append(p) or open(p, "wb").write(b"zip"))
    assert new and saved == [os.path.join(d, "best_model.zip")]
    assert not h.update(600_000, rewards=[5.0], lengths=[50])            # worse: no new best model, history still grows
    raw = json.load(open(os.path.join(d, "training_history.json")))
    assert set(raw) == set(cb.HISTORY_KEYS)                              # the eight keys of training_callback.py:488-497
    assert raw["best_mean_reward"] == 12.0 and raw["rewards_history"] == [12.0] and raw["timesteps_history"] == [300_000]
    assert raw["success_rates"] == [0.5] and raw["lengths_history"] == [120.0]
    # the file is written at the moment of the new best, BEFORE that evaluation's components are appended (:499-520)
    assert raw["reward_components_history"] == [] and raw["failure_reasons"] == []
    assert open(os.path.join(d, "training_history.json")).read().startswith('{\n    "best_mean_reward"')      # indent=4
    # reader (:300-331): resume continues counting evaluations
    h2 = cb.TrainingHistory(best_model_save_path=d, load_history_from=os.path.join(d, "training_history.json"))
    assert h2.eval_count == 1 and h2.best_mean_reward == 12.0 and h2.timesteps_history == [300_000]
    h2.update(900_000, rewards=[20.0], lengths=[10], save_model=lambda p: None)
    raw = json.load(open(os.path.join(d, "training_history.json")))
    assert raw["rewards_history"] == [12.0, 20.0] and raw["best_mean_reward"] == 20.0
    # a missing / broken file starts a fresh history like the reference
    assert cb.TrainingHistory(load_history_from=str(tmp_path / "nope.json")).eval_count == 0
    (tmp_path / "bad.json").write_text("{not json")
    assert cb.TrainingHistory(load_history_from=str(tmp_path / "bad.json")).rewards_history == []


def test_reference_written_history_loads(tmp_path):
    """A file exactly as the reference writes it (numpy scalars through default=float, -inf as -Infinity)."""
    data = {"best_mean_reward": np.float32(3.5), "rewards_history": [np.float64(1.0), 3.5], "lengths_history": [10.0, 20.0],
            "success_rates": [0.0, 0.5], "timesteps_history": [300000, 600000], "reward_components_history": [{}, {"upright": np.float32(0.25)}],
            "stability_metrics": [], "failure_reasons": ["unknown"]}
    p = tmp_path / "training_history.json"
    with open(p, "w") as f:
        json.dump(data, f, indent=4, default=float)
    h = cb.TrainingHistory(load_history_from=str(p))
    assert h.eval_count == 2 and h.best_mean_reward == 3.5 and h.reward_components_history[1]["upright"] == 0.25


def test_analyze_failure_rules():
    assert cb.analyze_failure({"upright": 0.2}, np.array([1.0])) == "fall_forward"
    assert cb.analyze_failure({"upright": 0.2}, np.array([-1.0])) == "fall_backward"
    assert cb.analyze_failure({"height": 0.1}, np.zeros(1)) == "too_low"
    assert cb.analyze_failure({"lateral_vel": 0.1}, np.zeros(1)) == "lateral_instability"
    assert cb.analyze_failure({"joint_limit": -1.0}, np.zeros(1)) == "joint_limit_violation"
    assert cb.analyze_failure({}, np.zeros(1)) == "unknown"


def test_evaluate_policy_on_the_vec_env(emu_lib_f32, tmp_path):
    """The batched evaluation loop + best-model bookkeeping on the (emulated) CUDA VecEnv."""
    from quadruped_rl_locomotion_b200.vec_env import Go1CudaVecEnv
    env = Go1CudaVecEnv(4, ctrl_type="torque", library=emu_lib_f32, seed=2)
    env.sim.set_state(0, step=5995)                         # this environment hits the time limit during the evaluation
    predict = lambda obs, deterministic=True: (np.zeros((4, 12), np.float32), None)
    env.reset = lambda: env._obs.astype(np.float64)                      # keep the state edited above (reset would overwrite it)
    res = cb.evaluate_policy(env, predict, max_steps=8)
    assert res["lengths"][0] == 5 and np.all(res["lengths"][1:] == 8) and len(res["failures"]) == 1
    h = cb.TrainingHistory(best_model_save_path=str(tmp_path))
    assert h.update(1000, res["rewards"], res["lengths"], failures=res["failures"], successes=res["successes"], save_model=lambda p: None)
    env.close()
