/*
 * go1sim.h -- C ABI of the B200-native batched Go1 simulator (libgo1sim.so / libgo1sim_f64.so).
 *
 * This is the drop-in boundary for the reference's environment hot path.  Every entry point
 * names the reference interface it replaces (paths relative to the reference repo):
 *
 *   go1_create      <- Go1MujocoEnv.__init__ -> MujocoEnv.__init__ -> MjModel.from_xml_path
 *                      (go1_mujoco_env.py:41-53) for N environments at once, and the worker
 *                      construction of SubprocVecEnv (train.py:44-50)
 *   go1_reset       <- VecEnv.reset() -> MujocoEnv.reset() -> Go1MujocoEnv.reset_model()
 *                      (go1_mujoco_env.py:949-1001)
 *   go1_step        <- SubprocVecEnv.step_async/step_wait -> Monitor.step -> Go1MujocoEnv.step
 *                      (go1_mujoco_env.py:473-532: do_simulation = mj_step x frame_skip +
 *                      mj_rnePostConstraint, _get_obs :915-947, _calc_reward :791-913,
 *                      termination :487-498) + the worker's auto-reset / terminal_observation
 *   go1_step_host   <- the same call with HOST numpy buffers (what SB3 hands to VecEnv.step); go1_step_host_f64 returns
 *                      the observations in the env's float64 observation dtype
 *   go1_rollout     <- test_max_envs.py:129-138 (random-action stepping loop, steps/s)
 *   go1_set_state / go1_get_state / go1_get_diag / go1_get_contacts
 *                   <- direct reads/writes of mjData the reference does
 *                      (data.qpos/qvel/qacc/qfrc_actuator/cfrc_ext/xpos/contact)
 *   go1_set_attr_f32<- VecEnv.set_attr("_rand_power", v) (training_callback.py:64) and the
 *                      constructor attributes of go1_mujoco_env.py:107-108
 *
 * Conventions: every function returns 0 on success or a negative GO1_E* code; the message is
 * available from go1_last_error().  No exceptions cross the ABI.  One handle = one GPU = one host
 * thread at a time.  Device-pointer arguments are raw CUDA device pointers (e.g. torch
 * tensor.data_ptr()); `stream` is a cudaStream_t cast to void*.  NULL = the handle's own stream, which is NON-BLOCKING: it is not
 * the legacy default stream and does not synchronise with it.  A caller whose own GPU work produces the inputs or consumes the
 * outputs passes the stream that work runs on (from PyTorch: a torch.cuda.Stream(); the default stream's handle is 0 = NULL) or
 * synchronises around the call.
 * Per-environment numerical failure is not an error: the env is flagged, auto-reset and counted.
 */
#ifndef GO1SIM_H_
#define GO1SIM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GO1_OBS_DIM 48
#define GO1_ACT_DIM 12
#define GO1_NQ 19
#define GO1_NV 18
#define GO1_NBODY 14
#define GO1_INFO_DIM 20
#define GO1_BIPED_INFO_DIM 12
#define GO1_MAX_CONTACTS 40        /* contacts kept per environment and substep (floor + robot-robot); more are dropped and counted */
#define GO1_MAX_PAIR_CONTACTS 16   /* robot-robot contacts among them */

enum {
    GO1_OK = 0,
    GO1_EINVAL = -1,      /* bad argument */
    GO1_EMODEL = -2,      /* model blob rejected (format / unsupported topology or feature) */
    GO1_ECUDA = -3,       /* CUDA runtime error */
    GO1_ENOMEM = -4,
    GO1_EUNSUPPORTED = -5 /* e.g. a model the kernel layout does not cover, biped info without biped mode */
};

/* Environment constants; zero-initialise then call go1_default_config() to obtain the
 * reference's literals (go1_mujoco_env.py:64-150). */
typedef struct Go1EnvConfig {
    int32_t biped;                 /* go1_mujoco_env.py:41,:43: 1 = bipedal mode (reset pose :967-972, terms :869-904) */
    int32_t frame_skip;            /* :49  -> 10 */
    int32_t max_episode_steps;     /* :64,:498 -> 6000 = 120 s / dt.  Both limits are STEP COUNTS for dt = 10 x 0.002 s; a caller that
                                    * changes frame_skip rescales them (the air-time bookkeeping :656 uses the model's dt itself) */
    int32_t max_unhealthy_steps;   /* :130,:497 -> terminate at the 751st consecutive unhealthy step: 750 */
    float rand_power;              /* :44,:974 */
    float fallen_start_prob;       /* :953 -> 0.2 */
    float desired_vel_min[3];      /* :107 */
    float desired_vel_max[3];      /* :108 */
    float reward_w[6];             /* :69-76  lin_vel, ang_vel, healthy, feet_airtime, recovery, get_up */
    float cost_w[11];              /* :77-89  torque, vertical_vel, xy_angular_vel, action_rate, joint_limit,
                                             joint_velocity, joint_acceleration, orientation, collision,
                                             default_joint_position, unhealthy_state */
    float healthy_z[2];            /* :118 */
    float healthy_roll[2];         /* :120 (applied to qpos[4], as the reference does) */
    float healthy_pitch[2];        /* :119 (applied to qpos[5]) */
    float tracking_sigma;          /* :116 */
    float obs_clip;                /* :150 */
    float solver_tolerance;        /* <0: use the model's (MJCF <option tolerance>) */
    int32_t solver_iterations;     /* <=0: use the model's */
    float biped_w[10];             /* :92-101 biped_upright (reward); front_contact, rear_feet_airborne, front_foot_height,
                                             crossed_legs, low_rear_hips, front_feet_below_hips, abduction_joints,
                                             unwanted_contact, self_collision (costs).  Read only when biped != 0. */
    float biped_ready_joints[12];  /* :32-39 BIPEDAL_READY_JOINTS (FR, FL, RR, RL) */
    float biped_reset_z;           /* :969 -> 0.65 */
    float biped_reset_pitch_deg;   /* :970 -> -95 */
} Go1EnvConfig;

/* Full per-environment state for teacher-forced parity tests (doubles in both builds). */
typedef struct Go1State {
    double qpos[GO1_NQ];
    double qvel[GO1_NV];
    double qacc_warmstart[GO1_NV];
    double last_action[GO1_ACT_DIM];
    double feet_air_time[4];
    double last_health_deviation[3];
    double desired_velocity[3];
    double episode_return;
    int32_t last_contacts[4];
    int32_t step;
    int32_t unhealthy_steps;
    int32_t reset_count;
    int32_t bad_state;
    int32_t front_feet_touched;    /* :66,:477-479 (biped only) */
    int32_t pad_;
} Go1State;

/* What mjData holds after do_simulation, as the reference reads it. */
typedef struct Go1Diag {
    double qacc[GO1_NV];
    double qfrc_actuator[GO1_NV];
    double qfrc_constraint[GO1_NV];
    double cfrc_ext[GO1_NBODY][6];   /* [torque; force] about the robot COM, world axes (mj_rnePostConstraint) */
    double trunk_xpos_z;             /* data.xpos[trunk][2] (start of the last substep) */
    double subtree_com[3];
    int32_t ncon;                    /* active contacts in the last substep */
    int32_t solver_iter;             /* Newton iterations of the last substep */
    int32_t solver_iter_total;       /* summed over the frame_skip substeps */
    int32_t contact_overflow;        /* contacts dropped because a per-leg list was full */
} Go1Diag;

typedef struct Go1Contact {
    int32_t geom;        /* mjContact.geom2: MuJoCo geom id of the robot geom */
    int32_t dim;
    double dist;
    double pos[3];       /* world */
    double force[6];     /* contact-frame force: normal, 2 tangents, torsion, 2 rolling */
    int32_t geom1;       /* mjContact.geom1: 0 = the floor plane, otherwise the robot geom of the lower geom type (robot-robot contact) */
    int32_t pad_;
} Go1Contact;

typedef struct Go1Stats {
    double sum_reward;
    double n_episodes;
    double sum_episode_length;
    double sum_episode_return;
    double n_terminated;
    double n_truncated;
    double n_bad_state;
    double env_steps;
    double n_contact_overflow;     /* contacts dropped because an environment had more than GO1_MAX_CONTACTS (or GO1_MAX_PAIR_CONTACTS robot-robot) */
    double sum_contacts;           /* active contacts in the last substep, summed over env-steps */
    double sum_solver_iterations;  /* Newton iterations, summed over substeps and env-steps */
    double sum_self_contacts;      /* robot-robot contacts in the last substep, summed over env-steps */
} Go1Stats;
#define GO1_NSTATS 12

/* Policy for go1_rollout(action_mode=GO1_ACTION_POLICY): SB3 MlpPolicy actor 48-64-64-12 (tanh),
 * a = clip(mu + exp(log_std) * eps, low, high). Row-major [out][in] weights (torch layout), device ptrs. */
typedef struct Go1Policy {
    const float* w0; const float* b0;    /* [64][48], [64] */
    const float* w1; const float* b1;    /* [64][64], [64] */
    const float* w2; const float* b2;    /* [12][64], [12] */
    const float* log_std;                /* [12] or NULL for deterministic */
} Go1Policy;

enum { GO1_ACTION_RANDOM = 0, GO1_ACTION_POLICY = 1, GO1_ACTION_ZERO = 2 };

typedef struct Go1Sim* go1_handle;

void go1_default_config(Go1EnvConfig* cfg);
int go1_precision_bits(void);                        /* 32 (product) or 64 (check build) */
const char* go1_version(void);

int go1_create(const void* model_blob, size_t nbytes, const Go1EnvConfig* cfg, int n_envs, int cuda_device,
               uint64_t seed, uint64_t env_id_offset, go1_handle* out);
int go1_destroy(go1_handle h);
const char* go1_last_error(go1_handle h);            /* h may be NULL: last create error */

int go1_num_envs(go1_handle h);
int go1_action_bounds(go1_handle h, float* low12, float* high12);   /* actuator_ctrlrange */

/* reset all envs (mask_dev NULL) or those with mask_dev[i] != 0; writes obs_dev[N][48] (may be NULL) */
int go1_reset(go1_handle h, const uint8_t* mask_dev, float* obs_dev, void* stream);

/* one control step for all envs, device buffers; asynchronous on `stream`.
 * obs_dev[N][48] f32 (reset obs for done envs), rew_dev[N] f32, done_dev[N] u8, trunc_dev[N] u8
 * (TimeLimit.truncated), terminal_obs_dev[N][48] (rows valid where done), info_dev[N][GO1_INFO_DIM] (nullable):
 * [0] x, [1] y, [2..9] 8 reward_info terms (go1_mujoco_env.py:860-867,908-909), [10] episode return, [11] episode length,
 * [12] health reason code (0 healthy, 1 not finite, 2 z, 3 roll, 4 pitch), [13] healthy flag, [14] Newton iterations of the
 * control step, [15] active contacts, [16] the offending value of termination_details (:444-459), [17] robot-robot contacts,
 * [18] contacts dropped for lack of slots, [19] spare. */
int go1_step(go1_handle h, const float* actions_dev, float* obs_dev, float* rew_dev, uint8_t* done_dev,
             uint8_t* trunc_dev, float* terminal_obs_dev, float* info_dev, void* stream);

/* go1_step without any allocation on the call path.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot reach cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments,
 * same results as go1_step, bit for bit; the step kernel's scratch (claim counter of the work queue, queue order, per-environment
 * work keys) lives in one file-scope `static __device__` array of the library instead of the handle's buffers, 128 KB that every
 * GPU which loads the library holds until exit.  Limit: handles of at most GO1_NOALLOC_MAX_ENVS environments (GO1_EINVAL with a
 * message otherwise).  The call performs no cudaMalloc / cudaFree / allocating library routine (the handle itself is created
 * beforehand by go1_create, which does allocate).  ALL handles on one device share the array: calls that overlap on the GPU
 * (different streams, or another host thread's handle) race on it -- serialise them.  The array keeps the previous call's work
 * keys on purpose (a sorting hint only: results do not depend on it); the claim counter is cleared by every call. */
#define GO1_NOALLOC_MAX_ENVS 16384
int go1_step_noalloc(go1_handle h, const float* actions_dev, float* obs_dev, float* rew_dev, uint8_t* done_dev,
                     uint8_t* trunc_dev, float* terminal_obs_dev, float* info_dev, void* stream);

/* same with HOST buffers: H2D of actions, kernel, D2H of results, synchronous (the VecEnv path). */
int go1_step_host(go1_handle h, const float* actions, float* obs, float* rew, uint8_t* done, uint8_t* trunc,
                  float* terminal_obs, float* info);
/* go1_step_host with the observation batch delivered as float64, obs64[N][48]: the dtype of the reference env's
 * observation_space (gymnasium Box float64, go1_mujoco_env.py:151-153) that SB3's rollout buffer receives from VecEnv.step.  The
 * widening runs on the device after the step kernel; every other buffer is as in go1_step_host. */
int go1_step_host_f64(go1_handle h, const float* actions, double* obs64, float* rew, uint8_t* done, uint8_t* trunc,
                      float* terminal_obs, float* info);
int go1_reset_host(go1_handle h, const uint8_t* mask, float* obs);
/* Optional: page-lock a caller-owned host buffer (cudaHostRegister).  go1_step_host copies to / from buffers that lie
 * inside a registered range directly (no staging through the library's own pinned buffers).  What SubprocVecEnv's
 * pipes + np.stack were in the reference (SB3 subproc_vec_env.py step_wait) becomes one DMA per array.
 * Unregister before freeing the memory; go1_destroy unregisters what is left. */
int go1_host_register(go1_handle h, void* ptr, size_t bytes);
int go1_host_unregister(go1_handle h, void* ptr);

/* device-resident stepping loop (throughput path): n_steps control steps with actions produced on the
 * device; accumulates stats on the device and copies them to *out (nullable) at the end (synchronous). */
int go1_rollout(go1_handle h, int n_steps, int action_mode, const Go1Policy* policy, Go1Stats* out, void* stream);
/* Throughput runs only (no counterpart in the reference: SubprocVecEnv workers de-synchronise by themselves, test_max_envs.py:129-138):
 * a slab that was reset all at once terminates all at once ~800 control steps later.  Sets the consecutive-unhealthy-step counter
 * (go1_mujoco_env.py:487-493) of every currently unhealthy environment to a Philox draw in [0, max_unhealthy_steps], so that episode ends
 * (and the auto-reset path) arrive at a uniform rate.  Physics state untouched. */
int go1_stagger_phases(go1_handle h);
/* elapsed device time (ms, CUDA events on the launch stream) and kernel launches of the last go1_rollout (per control step:
 * k_order + k_step, + the policy kernel in GO1_ACTION_POLICY mode) */
int go1_last_rollout_timing(go1_handle h, float* ms, int* launches);

/* Actor forward + action sampling for n rows of observations (device pointers): replaces the SB3
 * ActorCriticPolicy forward / predict the PPO loop runs between env steps (train.py:122-128,
 * OnPolicyAlgorithm.collect_rollouts).  log_std == NULL -> deterministic (predict(deterministic=True)).
 * The Philox stream is (seed, env_id_offset + row, launch counter); every call advances the counter. */
int go1_policy_act(go1_handle h, const Go1Policy* policy, const float* obs_dev, float* act_dev, int n, void* stream);

/* Critic of the SB3 MlpPolicy (train.py:122-128, net_arch pi=[64,64], vf=[64,64]): mlp_extractor.value_net + value_net.
 * Row-major [out][in] weights (torch layout), device pointers. */
typedef struct Go1Critic {
    const float* w0; const float* b0;    /* [64][48], [64] */
    const float* w1; const float* b1;    /* [64][64], [64] */
    const float* w2; const float* b2;    /* [1][64], [1]   */
} Go1Critic;

/* fp32 scalar reference of the actor forward (one thread per row, CUDA cores): what the tensor-core kernel is tested against */
int go1_policy_act_scalar(go1_handle h, const Go1Policy* policy, const float* obs_dev, float* act_dev, int n, void* stream);

/* ActorCriticPolicy.forward for n rows of observations on the tensor cores (one tcgen05 CTA per 128 rows, tf32 operands,
 * fp32 accumulation in tensor memory): actions = clip(mu + exp(log_std) * eps) (deterministic mean when log_std == NULL),
 * values[n] (critic may be NULL -> values untouched), log_probs[n] of the UNCLIPPED Gaussian sample as SB3 stores them in its
 * RolloutBuffer, raw_act[n][12] = that unclipped sample (any of the three may be NULL).  Same Philox stream as go1_policy_act. */
int go1_policy_forward(go1_handle h, const Go1Policy* actor, const Go1Critic* critic, const float* obs_dev, float* act_dev,
                       float* values_dev, float* log_probs_dev, float* raw_act_dev, int n, void* stream);

/* render_mode="rgb_array" (train.py:53-61: camera_name="tracking", width=1024, height=768; frames consumed by
 * VideoRecordingCallback, training_callback.py:686-738).  The reference renders with MuJoCo's OpenGL renderer, which is unavailable
 * on this box; this entry ray-casts the model's collision primitives and the checker floor on the GPU (no graphics API) into
 * rgb[n_img][height][width][3] u8 (HOST buffer when rgb_is_host != 0, else device).  Camera: mode trackcom (follows the robot's
 * centre of mass with a fixed world orientation); cam == NULL = the reference's `tracking` camera (go1_torque.xml:84). */
typedef struct Go1Camera {
    float pos[3];        /* camera position in the trunk frame at the reference pose qpos0 ("0 -1 0.8") */
    float xaxis[3];      /* xyaxes "1 0 0 0 1 1": image x and y directions in the world (orthonormalised here) */
    float yaxis[3];
    float fovy_deg;      /* MuJoCo default 45 */
} Go1Camera;
int go1_render(go1_handle h, const Go1Camera* cam, const int32_t* env_ids_host, int n_img, int width, int height, uint8_t* rgb,
               int rgb_is_host);

/* Biped mode only (cfg.biped != 0; GO1_EUNSUPPORTED otherwise): the extra reward_info entries of the LAST control
 * step, out[N][GO1_BIPED_INFO_DIM] f32 HOST buffer (synchronous D2H): biped_upright_reward, biped_front_contact_cost,
 * biped_rear_feet_airborne_cost, biped_front_foot_height_cost, biped_crossed_legs_cost, biped_low_rear_hips_cost,
 * biped_front_feet_below_hips_cost, biped_abduction_joints_cost, biped_unwanted_contact_cost, self_collision_cost
 * (go1_mujoco_env.py:893-902, costs negated as there), bipedal_success (:520-521), front_feet_touched (:477-479).
 * go1_biped_info_dev returns the device pointer of the same rows for device-resident callers.
 * go1_biped_info_host synchronises the handle's own stream (go1_step_host, go1_rollout with stream NULL); after
 * go1_step on a caller-owned stream, synchronise that stream first. */
int go1_biped_info_host(go1_handle h, float* out);
int go1_biped_info_dev(go1_handle h, const float** ptr);

int go1_set_state(go1_handle h, int env, const Go1State* s);
int go1_get_state(go1_handle h, int env, Go1State* s);
int go1_enable_diag(go1_handle h, int enable);       /* diag/contact capture costs HBM traffic: off by default */
int go1_get_diag(go1_handle h, int env, Go1Diag* d);
int go1_get_contacts(go1_handle h, int env, Go1Contact* buf, int cap, int* n);

/* names: "rand_power", "fallen_start_prob", "desired_vel_min_x|y|z", "desired_vel_max_x|y|z" */
int go1_set_attr_f32(go1_handle h, const char* name, float v);
int go1_get_attr_f32(go1_handle h, const char* name, float* v);

#ifdef __cplusplus
}
#endif
#endif /* GO1SIM_H_ */
