"""numpy (fp64) restatement of the reference's environment logic -- TEST INFRASTRUCTURE ONLY.

Follows ``/root/reference/go1_mujoco_env.py`` (non-biped path) line by line:

* observation            ``_get_obs``              go1_mujoco_env.py:915-947
* projected gravity      ``projected_gravity``     :596-608, ``euler_from_quaternion`` :1017-1038
* health                 ``_get_health_status``    :433-471
* reward / cost terms    ``_calc_reward``          :791-913 and the properties it calls
* step bookkeeping       ``step``                  :473-532
* reset                  ``reset_model``           :949-1001 (deterministic part; sampling is the caller's)

It is validated against the reference's literal code in ``tests/test_env_logic.py`` (through
``oracle/ref_env.py``, only where ``/root/reference`` exists) and against the committed golden
vectors everywhere.  The CUDA path is then compared with THIS restatement on the GPU box,
where the reference sources are not available.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

REWARD_WEIGHTS = {  # go1_mujoco_env.py:69-76
    "linear_vel_tracking": 2.0, "angular_vel_tracking": 1.0, "healthy": 1.0, "feet_airtime": 5.0,
    "recovery": 10.0, "get_up": 20.0,
}
COST_WEIGHTS = {  # go1_mujoco_env.py:77-89
    "torque": 0.0002, "vertical_vel": 2.0, "xy_angular_vel": 0.05, "action_rate": 0.01, "joint_limit": 10.0,
    "joint_velocity": 0.01, "joint_acceleration": 2.0e-4, "orientation": 1.0, "collision": 1.0,
    "default_joint_position": 0.1, "unhealthy_state": 5.0,
}
HEALTHY_Z = (0.22, 1.8)                                # :118
HEALTHY_PITCH = (-np.pi, 0.0)                          # :119  (applied to state[5] == qy)
HEALTHY_ROLL = (-np.deg2rad(80), np.deg2rad(80))       # :120  (applied to state[4] == qx)
FEET_BODIES = [4, 7, 10, 13]                           # :124
CONTACT_BODIES = [2, 3, 5, 6, 8, 9, 11, 12]            # :126
DT = 0.02                                              # timestep 0.002 * frame_skip 10
MAX_UNHEALTHY_TIME = 15.0                              # :130
MAX_EPISODE_STEPS = 120.0 / DT                         # :64, :498
BIPED_REWARD_WEIGHT = 15.0                              # :92  biped_upright
BIPED_COST_WEIGHTS = {  # :93-101
    "front_contact": 50.0, "rear_feet_airborne": 5.0, "front_foot_height": 8.0, "crossed_legs": 5.0, "low_rear_hips": 9.0,
    "front_feet_below_hips": 6.0, "abduction_joints": 0.7, "unwanted_contact": 150.0, "self_collision": 25.0,
}
BIPEDAL_READY_JOINTS = np.array([0.0, 4.0, -2.0, 0.0, 4.0, -2.0, 0.0, 2.8, -1.2, 0.0, 2.8, -1.2])   # :32-39
UNWANTED_CONTACT_BODIES = [1, 3, 6, 9, 12, 4, 7]       # :192-200 trunk, thighs, front calves (MuJoCo body ids)
BIPED_INFO_KEYS = ("biped_upright_reward", "biped_front_contact_cost", "biped_rear_feet_airborne_cost",
                   "biped_front_foot_height_cost", "biped_crossed_legs_cost", "biped_low_rear_hips_cost",
                   "biped_front_feet_below_hips_cost", "biped_abduction_joints_cost", "biped_unwanted_contact_cost",
                   "self_collision_cost")
INFO_KEYS = ("linear_vel_tracking_reward", "reward_ctrl", "reward_survive", "recovery_reward", "get_up_reward",
             "unhealthy_state_cost", "orientation_cost", "default_joint_position_cost")


def euler_from_quaternion(w, x, y, z):  # :1017-1038
    t0 = 2.0 * (w * x + y * z)
    t1 = 1.0 - 2.0 * (x * x + y * y)
    roll = np.arctan2(t0, t1)
    t2 = 2.0 * (w * y - z * x)
    t2 = 1.0 if t2 > 1.0 else t2
    t2 = -1.0 if t2 < -1.0 else t2
    pitch = np.arcsin(t2)
    t3 = 2.0 * (w * z + x * y)
    t4 = 1.0 - 2.0 * (y * y + z * z)
    yaw = np.arctan2(t3, t4)
    return roll, pitch, yaw


def projected_gravity(quat, gravity=(0.0, 0.0, -9.81)):  # :596-608
    e = np.array(euler_from_quaternion(*quat))
    v = np.dot(np.asarray(gravity), e) * e
    n = np.linalg.norm(v)
    return v if n == 0 else v / n


def health(qpos, qvel):  # :433-471 -> (ok, reason_code) 0 ok,1 not finite,2 z,3 roll,4 pitch
    s = np.concatenate([qpos, qvel])
    if not np.isfinite(s).all():
        return False, 1
    if not (HEALTHY_Z[0] <= s[2] <= HEALTHY_Z[1]):
        return False, 2
    if not (HEALTHY_ROLL[0] <= s[4] <= HEALTHY_ROLL[1]):
        return False, 3
    if not (HEALTHY_PITCH[0] <= s[5] <= HEALTHY_PITCH[1]):
        return False, 4
    return True, 0


def soft_joint_range(ctrlrange):  # :134-145
    off = 0.5 * (1 - 0.9) * (ctrlrange[:, 1] - ctrlrange[:, 0])
    r = ctrlrange.copy()
    r[:, 0] += off
    r[:, 1] -= off
    return r


@dataclass
class EpisodeState:
    step: int = 0
    last_action: np.ndarray = field(default_factory=lambda: np.zeros(12))
    feet_air_time: np.ndarray = field(default_factory=lambda: np.zeros(4))
    last_contacts: np.ndarray = field(default_factory=lambda: np.zeros(4, bool))
    desired_velocity: np.ndarray = field(default_factory=lambda: np.zeros(3))
    time_unhealthy: float = 0.0
    last_dev: np.ndarray = field(default_factory=lambda: np.zeros(3))  # z, roll, pitch
    front_feet_touched: bool = False                                   # biped only (:476-479)


def count_self_collisions(contacts, geom_bodyid):
    """``self_collision_cost`` before its weight (go1_mujoco_env.py:269-312): contacts of ``data.contact`` between the
    front-right and the front-left limb (bodies FR_hip/thigh/calf = 2..4, FL = 5..7) or between the rear pair (RR = 8..10,
    RL = 11..13), in either geom order."""
    fr, fl, rr, rl = {2, 3, 4}, {5, 6, 7}, {8, 9, 10}, {11, 12, 13}
    n = 0
    for c in contacts:
        b1, b2 = int(geom_bodyid[c.geom1]), int(geom_bodyid[c.geom2])
        if (b1 in fr and b2 in fl) or (b2 in fr and b1 in fl) or (b1 in rr and b2 in rl) or (b2 in rr and b1 in rl):
            n += 1
    return n


class EnvLogic:
    """Per-step env logic given the physics results of ``do_simulation``."""

    def __init__(self, model, biped=False):
        self.biped = bool(biped)
        self.key_qpos = np.array(model.key_qpos[0])
        self.default_joint_position = np.array(model.key_ctrl[0])      # :105
        self.soft_range = soft_joint_range(np.array(model.actuator_ctrlrange))
        self.gravity = np.array(model.opt_gravity)

    def obs(self, qpos, qvel, ep: EpisodeState):  # :915-947
        o = np.concatenate([
            qvel[:3] * 2.0, qvel[3:6] * 0.25, projected_gravity(qpos[3:7], self.gravity),
            ep.desired_velocity * 2.0, (qpos[7:] - self.key_qpos[7:]) * 1.0, qvel[6:] * 0.05, ep.last_action,
        ])
        return np.clip(o, -100.0, 100.0)

    def step(self, ep: EpisodeState, action, qpos, qvel, qacc, qfrc_actuator, cfrc_ext, trunk_z_xpos, kin=None,
             prev_front_feet_forces=None, n_self_collisions=0):
        """Everything ``Go1MujocoEnv.step`` does after ``do_simulation`` (:483-532).

        ``qpos/qvel`` are post-step; ``qacc/qfrc_actuator/cfrc_ext/trunk_z_xpos`` are the values MuJoCo
        leaves in ``mjData`` (start of the last substep / ``mj_rnePostConstraint``).
        Returns (obs, reward, terminated, truncated, info dict); mutates ``ep``.
        """
        action = np.asarray(action, dtype=np.float64)
        ep.step += 1                                                    # :474
        if self.biped and prev_front_feet_forces is not None and np.any(np.asarray(prev_front_feet_forces) > 1.0):
            ep.front_feet_touched = True                                # :477-479: cfrc_ext of the PREVIOUS step
        obs = self.obs(qpos, qvel, ep)                                  # :483 (last_action = a_{t-1})
        ok, reason = health(qpos, qvel)
        pg = projected_gravity(qpos[3:7], self.gravity)
        v_des = ep.desired_velocity
        # ---- rewards (:793-818)
        lin = np.exp(-np.sum(np.square(v_des[:2] - qvel[:2])) / 0.25) * REWARD_WEIGHTS["linear_vel_tracking"]
        ang = np.exp(-np.square(v_des[2] - qvel[5]) / 0.25) * REWARD_WEIGHTS["angular_vel_tracking"]
        healthy_r = (1.0 if ok else 0.0) * REWARD_WEIGHTS["healthy"]
        # feet air time (:633-668)
        feet_f = np.linalg.norm(cfrc_ext[FEET_BODIES], axis=1)
        curr = feet_f > 1.0
        if self.biped:                                                  # :646-649: alternating rear feet, no bookkeeping
            air = float(curr[2] != curr[3])
        else:
            filt = np.logical_or(curr, ep.last_contacts)
            ep.last_contacts = curr.copy()
            first = (ep.feet_air_time > 0.0) * filt
            ep.feet_air_time = ep.feet_air_time + DT
            air = np.sum(np.square((ep.feet_air_time - 0.2).clip(min=0.0)) * first)
            air *= np.linalg.norm(v_des[:2]) > 0.1
            ep.feet_air_time = ep.feet_air_time * ~filt
        air_r = air * REWARD_WEIGHTS["feet_airtime"]
        # recovery (:534-580)
        z, roll, pitch = qpos[2], qpos[4], qpos[5]
        dev = np.zeros(3)
        if not (HEALTHY_Z[0] <= z <= HEALTHY_Z[1]):
            dev[0] = min(abs(z - HEALTHY_Z[0]), abs(z - HEALTHY_Z[1]))
        if not (HEALTHY_ROLL[0] <= roll <= HEALTHY_ROLL[1]):
            dev[1] = min(abs(roll - HEALTHY_ROLL[0]), abs(roll - HEALTHY_ROLL[1]))
        if not (HEALTHY_PITCH[0] <= pitch <= HEALTHY_PITCH[1]):
            dev[2] = min(abs(pitch - HEALTHY_PITCH[0]), abs(pitch - HEALTHY_PITCH[1]))
        if ok:
            ep.last_dev = np.zeros(3)
            rec = 0.0
        else:
            rec = float(np.sum(ep.last_dev - dev))
            ep.last_dev = dev
        rec_r = rec * REWARD_WEIGHTS["recovery"]
        # get up (:219-243)
        if ok:
            gu = 0.0
        else:
            gu = float(np.clip(trunk_z_xpos / HEALTHY_Z[0], 0.0, 1.0)) * 1.5 + (1.0 - np.sum(np.square(pg[:2]))) * 0.5
        gu_r = gu * REWARD_WEIGHTS["get_up"]
        rewards = lin + ang + healthy_r + air_r + rec_r + gu_r
        # ---- costs (:821-858)
        disc = 1.0 if ok else 0.1
        ctrl_c = np.sum(np.square(qfrc_actuator[-12:])) * disc * COST_WEIGHTS["torque"]
        rate_c = np.sum(np.square(ep.last_action - action)) * disc * COST_WEIGHTS["action_rate"]
        vert_c = np.square(qvel[2]) * COST_WEIGHTS["vertical_vel"]
        xyang_c = np.sum(np.square(qvel[3:5])) * COST_WEIGHTS["xy_angular_vel"]
        jl = (self.soft_range[:, 0] - qpos[7:]).clip(min=0.0) + (qpos[7:] - self.soft_range[:, 1]).clip(min=0.0)
        jl_c = np.sum(jl) * COST_WEIGHTS["joint_limit"]
        jv_c = np.sum(np.square(qvel[6:])) * COST_WEIGHTS["joint_velocity"]
        acc_c = np.sum(np.square(qacc[6:]) / (np.abs(qvel[6:]) + 1e-6)) * disc * COST_WEIGHTS["joint_acceleration"]
        coll_c = (1.0 * (np.linalg.norm(cfrc_ext[CONTACT_BODIES]) > 0.1)) * COST_WEIGHTS["collision"]
        unh_c = (0.0 if ok else 1.0) * COST_WEIGHTS["unhealthy_state"]
        orient_c = np.sum(np.square(pg[:2])) * COST_WEIGHTS["orientation"]
        djp_c = np.sum(np.square(qpos[7:] - self.default_joint_position)) * COST_WEIGHTS["default_joint_position"]
        costs = ctrl_c + rate_c + vert_c + xyang_c + jl_c + jv_c + acc_c + coll_c + unh_c
        binfo = {}
        if self.biped:                                                  # :869-904
            w = BIPED_COST_WEIGHTS
            front_f = feet_f[:2]
            up = max(0.0, float(kin["trunk_xaxis_z"])) * BIPED_REWARD_WEIGHT                      # :388-401
            fc = float(np.sum(np.square(front_f))) * w["front_contact"]                              # :426-431
            ra = w["rear_feet_airborne"] if np.all(feet_f[2:] < 1.0) else 0.0                        # :880-882
            fh = float(np.sum((kin["front_calf_z"] - kin["front_feet_z"]).clip(min=0.0))) * w["front_foot_height"]   # :403-423
            cl = max(0.0, float(kin["rear_hip_y"][0] - kin["rear_hip_y"][1])) * w["crossed_legs"]    # :315-331
            lh = float(np.sum((0.2 - kin["rear_hip_z"]).clip(min=0.0))) * 10.0 * w["low_rear_hips"]  # :334-351
            fb = float(np.sum(np.square((kin["front_hip_z"] - kin["front_feet_z"]).clip(min=0.0)))) * w["front_feet_below_hips"]  # :354-376
            ab = float(np.sum(np.square(qpos[7:][[0, 3, 6, 9]]))) * w["abduction_joints"]            # :759-769
            uc = float(np.sum(np.square(np.linalg.norm(cfrc_ext[UNWANTED_CONTACT_BODIES], axis=1)))) * w["unwanted_contact"]  # :772-781
            sc = float(n_self_collisions) * w["self_collision"]                                      # :269-312
            rewards = rewards + up
            costs = costs + fc + ra + fh + cl + lh + fb + ab + uc + sc
            binfo = dict(zip(BIPED_INFO_KEYS, (up, -fc, -ra, -fh, -cl, -lh, -fb, -ab, -uc, -sc)))
        else:
            costs = costs + orient_c + djp_c
        reward = max(0.0, rewards - costs)                              # :911
        # ---- termination (:487-498)
        if not ok:
            ep.time_unhealthy += DT
        else:
            ep.time_unhealthy = 0.0
        terminated = ep.time_unhealthy > MAX_UNHEALTHY_TIME
        truncated = ep.step >= MAX_EPISODE_STEPS
        info = {
            "x_position": qpos[0], "y_position": qpos[1], "healthy": ok, "health_reason": reason,
            "linear_vel_tracking_reward": lin, "reward_ctrl": -ctrl_c, "reward_survive": healthy_r,
            "recovery_reward": rec_r, "get_up_reward": gu_r, "unhealthy_state_cost": -unh_c,
            "orientation_cost": -orient_c, "default_joint_position_cost": -djp_c,
            "feet_contact_forces": feet_f, **binfo,
            "bipedal_success": bool(truncated and self.biped and not ep.front_feet_touched),         # :520-521
        }
        ep.last_action = action.copy()                                  # :529
        return obs, reward, terminated, truncated, info
