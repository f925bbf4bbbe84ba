"""The data formats and schedules of the reference's training callbacks, so that resume, dashboards and
checkpoints written by either side stay interchangeable with the CUDA VecEnv / :class:`~quadruped_rl_locomotion_b200.ppo.DevicePPO`.

* :func:`curriculum_rand_power` -- ``CurriculumCallback._on_step`` (training_callback.py:47-69): ``rand_power`` falls linearly
  from its initial value to 0 at 70 % of the run, then stays 0; :class:`CurriculumSchedule` applies it through
  ``env.set_attr("_rand_power", v)`` exactly like the reference (:64).
* :class:`TrainingHistory` -- ``training_history.json`` (writer :483-511, reader ``_load_history`` :300-331): the eight keys
  ``best_mean_reward``, ``rewards_history``, ``lengths_history``, ``success_rates``, ``timesteps_history``,
  ``reward_components_history``, ``stability_metrics``, ``failure_reasons``; ``json.dump(..., indent=4, default=float)``.
  ``update`` reproduces ``_update_history``: append the evaluation, and on a new best mean reward save ``best_model.zip`` next
  to ``training_history.json``.
* :func:`analyze_failure` -- ``_analyze_failure`` (:452-466).
* :func:`evaluate_policy` -- the evaluation loop (:349-408) without its wall-clock cap and ``time.sleep``: deterministic actions,
  first episode of every environment of a VecEnv, so 1000 evaluation episodes are one batched rollout.

The video / matplotlib parts of the callbacks (OpenGL renderer, plots) are not reproduced here; frames for
``VideoRecordingCallback`` come from the CUDA ray-caster (``quadruped_rl_locomotion_b200.render``)."""
from __future__ import annotations

import json
import os
from typing import Callable, Optional

import numpy as np

HISTORY_KEYS = ("best_mean_reward", "rewards_history", "lengths_history", "success_rates", "timesteps_history",
                "reward_components_history", "stability_metrics", "failure_reasons")
COMPONENT_KEYS = ("upright", "height", "feet", "forward_vel", "lateral_vel", "cop_stab", "zmp_stab")      # :372-373


def curriculum_rand_power(num_timesteps: int, total_timesteps: int, initial_rand_power: float) -> float:
    """training_callback.py:47-69."""
    end = int(total_timesteps * 0.7)
    if num_timesteps < end:
        return initial_rand_power * (1.0 - num_timesteps / end)
    return 0.0


class CurriculumSchedule:
    """``CurriculumCallback``: call once per step / rollout with the current ``num_timesteps``."""

    def __init__(self, env, total_timesteps: int, initial_rand_power: float):
        self.env, self.total, self.initial = env, int(total_timesteps), float(initial_rand_power)
        self.value = float(initial_rand_power)

    def __call__(self, num_timesteps: int) -> float:
        self.value = curriculum_rand_power(int(num_timesteps), self.total, self.initial)
        self.env.set_attr("_rand_power", self.value)                         # :64
        return self.value


def analyze_failure(info: dict, obs) -> str:
    """training_callback.py:452-466."""
    reasons = []
    if info.get("upright", 1) < 0.5:
        reasons.append("fall_forward" if obs[0] > 0 else "fall_backward")
    if info.get("height", 1) < 0.3:
        reasons.append("too_low")
    if info.get("lateral_vel", 1) < 0.5:
        reasons.append("lateral_instability")
    if info.get("joint_limit", 0) < -0.5:
        reasons.append("joint_limit_violation")
    return reasons[0] if reasons else "unknown"


class TrainingHistory:
    """``training_history.json`` + ``best_model.zip`` of ``EnhancedVisualCallback`` (training_callback.py:270-331, :472-520)."""

    def __init__(self, best_model_save_path: Optional[str] = None, load_history_from: Optional[str] = None):
        self.best_model_save_path = best_model_save_path
        self.best_mean_reward = -np.inf
        self.rewards_history, self.lengths_history, self.success_rates, self.timesteps_history = [], [], [], []
        self.reward_components_history, self.stability_metrics, self.failure_reasons = [], [], []
        self.eval_count = 0
        if load_history_from:
            self.load(load_history_from)

    # -- reader (:300-331): a missing or broken file starts a fresh history, like the reference
    def load(self, path: str) -> bool:
        if not os.path.exists(path):
            return False
        try:
            with open(path, "r") as f:
                h = json.load(f)
        except Exception:  # noqa: BLE001
            return False
        self.best_mean_reward = h.get("best_mean_reward", -np.inf)
        self.rewards_history = h.get("rewards_history", [])
        self.lengths_history = h.get("lengths_history", [])
        self.success_rates = h.get("success_rates", [])
        self.timesteps_history = h.get("timesteps_history", [])
        self.reward_components_history = h.get("reward_components_history", [])
        self.stability_metrics = h.get("stability_metrics", [])
        self.failure_reasons = h.get("failure_reasons", [])
        self.eval_count = len(self.rewards_history)
        return True

    def as_dict(self) -> dict:
        return {"best_mean_reward": self.best_mean_reward, "rewards_history": self.rewards_history, "lengths_history": self.lengths_history,
                "success_rates": self.success_rates, "timesteps_history": self.timesteps_history,
                "reward_components_history": self.reward_components_history, "stability_metrics": self.stability_metrics,
                "failure_reasons": self.failure_reasons}

    # -- writer (:499-511)
    def save(self, directory: Optional[str] = None) -> str:
        d = directory or self.best_model_save_path or "."
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "training_history.json")
        with open(path, "w") as f:
            json.dump(self.as_dict(), f, indent=4, default=float)
        return path

    def append(self, num_timesteps: int, mean_reward: float, mean_length: float = float("nan"), success_rate: float = 0.0):
        self.rewards_history.append(float(mean_reward)); self.lengths_history.append(float(mean_length))
        self.success_rates.append(float(success_rate)); self.timesteps_history.append(int(num_timesteps))
        self.eval_count = len(self.rewards_history)

    def update(self, num_timesteps: int, rewards, lengths, components=(), stability=(), failures=(), successes: int = 0,
               save_model: Optional[Callable[[str], object]] = None) -> bool:
        """``_update_history`` (:472-520).  ``save_model(path)`` is ``self.model.save`` of the reference; returns True when a new
        best model was written."""
        n = max(1, len(rewards))
        mean_reward = float(np.mean(rewards))
        self.append(num_timesteps, mean_reward, float(np.mean(lengths)), successes / n)
        new_best = False
        if self.best_model_save_path is not None and mean_reward > self.best_mean_reward:
            self.best_mean_reward = mean_reward
            new_best = True
            os.makedirs(self.best_model_save_path, exist_ok=True)
            if save_model is not None:
                save_model(os.path.join(self.best_model_save_path, "best_model.zip"))
            self.save(self.best_model_save_path)
        avg = {}
        if components:
            for key in components[0].keys():
                avg[key] = float(np.mean([c.get(key, 0) for c in components]))
        self.reward_components_history.append(avg)                                                  # :514-520
        if stability:
            self.stability_metrics.append({k: float(np.mean([s[k] for s in stability])) for k in stability[0].keys()})
        self.failure_reasons.extend(list(failures))
        return new_best


def evaluate_policy(env, predict: Callable, max_steps: int = 6000):
    """The evaluation loop of training_callback.py:349-408 for a VecEnv: deterministic actions, FIRST episode of every
    environment.  Returns dict(rewards, lengths, successes, failures) -- the arguments of ``TrainingHistory.update``."""
    obs = env.reset()
    n = env.num_envs
    ret = np.zeros(n); length = np.zeros(n, int); alive = np.ones(n, bool)
    successes, failures = 0, []
    for _ in range(max_steps):
        act, _ = predict(obs, deterministic=True)
        obs, rew, dones, infos = env.step(act)
        ret += np.where(alive, rew, 0.0); length += alive
        for i in np.flatnonzero(dones & alive):
            info = infos[i]
            if info.get("bipedal_success", False):
                successes += 1
            else:
                failures.append(analyze_failure(info, info.get("terminal_observation", obs[i])))
        alive &= ~dones
        if not alive.any():
            break
    return dict(rewards=ret, lengths=length, successes=successes, failures=failures)
