/*
 * b200lab.h -- C ABI of libb200lab.so: the B200-native (sm_100a) non-physics step of Isaac Lab
 * (ManagerBasedRLEnv.step MDP pipeline, actuator models, contact/ray-cast sensors) and the
 * rsl_rl PPO hot path (GAE, minibatch gather, actor-critic MLP fwd/bwd, loss, Adam).
 *
 * The reference (godk1122/IsaacLab, Isaac Lab 2.1.0 + rsl_rl 3.0.0) is 100 % Python and has no FFI for
 * this path, so every entry point below names the reference Python symbol it replaces
 * (path:line relative to the reference root; LAB = source/isaaclab/isaaclab, RSL = source/rsl_rl/rsl_rl,
 * TASKS = source/isaaclab_tasks/isaaclab_tasks).  INTEGRATION.md shows the ctypes binding a maintainer
 * of the reference would add.
 *
 * Conventions (all entry points):
 *   - plain C types only; every pointer inside an *Args struct is a DEVICE pointer unless the field
 *     name starts with `h_` (host pointer);
 *   - tensors are contiguous, env-major, fp32 unless stated (i64 = int64_t, u8 = uint8_t/bool);
 *   - quaternions are (w, x, y, z) as in ArticulationData (LAB/assets/articulation/articulation_data.py:366);
 *   - work is enqueued on `stream` (a cudaStream_t passed as void*) and the call returns without
 *     synchronising; nothing is allocated on the hot path;
 *   - return value 0 = ok, negative = error; b200lab_last_error() returns a static message;
 *   - not thread-safe per device.  No CPU fallback exists: without a CUDA device every compute
 *     entry point returns B200LAB_ERR_CUDA.
 */
#ifndef B200LAB_H_
#define B200LAB_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200LAB_VERSION 100

#define B200LAB_OK 0
#define B200LAB_ERR_ARG (-1)
#define B200LAB_ERR_CUDA (-2)
#define B200LAB_ERR_UNSUPPORTED (-3)

#define B200LAB_MAX_TERMS 32
#define B200LAB_MAX_ACT_GROUPS 8
#define B200LAB_MAX_INTERVAL_EVENTS 4
#define B200LAB_MAX_RESET_JOINT_EVENTS 4

const char* b200lab_last_error(void);
int b200lab_version(void);
/* number of kernels this library has launched since load (bench.py's `gpu_launches`) */
int64_t b200lab_launch_count(void);
/* sizeof() of the named *Args struct as compiled (binding self-check); -1 if unknown */
int64_t b200lab_sizeof(const char* name);

/* ------------------------------------------------------------------------------------------------
 * Term programs: the compiled form of the manager term cfgs (`func(env, **params) -> Tensor`,
 * LAB/managers/manager_term_cfg.py:28-344).  Host-side Python maps recognised `func` identities to
 * the opcodes below; unknown user functions are evaluated in Python and enter as *_EXTERNAL columns.
 * ---------------------------------------------------------------------------------------------- */
enum B200LabRewardOp {
  B200LAB_REW_IS_ALIVE = 1,                    /* LAB/envs/mdp/rewards.py:31 */
  B200LAB_REW_IS_TERMINATED = 2,               /* rewards.py:36 */
  B200LAB_REW_LIN_VEL_Z_L2 = 3,                /* rewards.py:76 */
  B200LAB_REW_ANG_VEL_XY_L2 = 4,               /* rewards.py:83 */
  B200LAB_REW_FLAT_ORIENTATION_L2 = 5,         /* rewards.py:90 */
  B200LAB_REW_JOINT_TORQUES_L2 = 6,            /* rewards.py:136 */
  B200LAB_REW_JOINT_VEL_L1 = 7,                /* rewards.py:146 */
  B200LAB_REW_JOINT_VEL_L2 = 8,                /* rewards.py:153 */
  B200LAB_REW_JOINT_ACC_L2 = 9,                /* rewards.py:163 */
  B200LAB_REW_JOINT_DEVIATION_L1 = 10,         /* rewards.py:173 */
  B200LAB_REW_JOINT_POS_LIMITS = 11,           /* rewards.py:182 */
  B200LAB_REW_ACTION_RATE_L2 = 12,             /* rewards.py:245 */
  B200LAB_REW_ACTION_L2 = 13,                  /* rewards.py:250 */
  B200LAB_REW_UNDESIRED_CONTACTS = 14,         /* rewards.py:260  p[0]=threshold, body_mask */
  B200LAB_REW_CONTACT_FORCES = 15,             /* rewards.py:271  p[0]=threshold, body_mask */
  B200LAB_REW_TRACK_LIN_VEL_XY_EXP = 16,       /* rewards.py:287  p[0]=std**2 (fp32) */
  B200LAB_REW_TRACK_ANG_VEL_Z_EXP = 17,        /* rewards.py:301  p[0]=std**2 */
  B200LAB_REW_FEET_AIR_TIME = 18,              /* TASKS/manager_based/locomotion/velocity/mdp/rewards.py:25  p[0]=threshold, p[1]=step_dt+1e-8 */
  B200LAB_REW_FEET_AIR_TIME_POSITIVE_BIPED = 19, /* same file :47  p[0]=threshold */
  B200LAB_REW_FEET_SLIDE = 20,                 /* same file :69  body_mask=contact bodies, body_mask2 in i64[0] = asset bodies */
  B200LAB_REW_TRACK_LIN_VEL_XY_YAW_FRAME_EXP = 21, /* same file :86 p[0]=std**2 */
  B200LAB_REW_TRACK_ANG_VEL_Z_WORLD_EXP = 22,  /* same file :99 p[0]=std**2 */
  B200LAB_REW_JOINT_POS_TARGET_L2 = 23,        /* TASKS/manager_based/classic/cartpole/mdp/rewards.py:19 p[0]=target */
  B200LAB_REW_EXTERNAL = 100                   /* value read from ext_values[col=i0][N] (user Python term) */
};

enum B200LabTerminationOp {
  B200LAB_TERM_TIME_OUT = 1,                   /* LAB/envs/mdp/terminations.py:30 */
  B200LAB_TERM_ILLEGAL_CONTACT = 2,            /* terminations.py:150  p[0]=threshold, body_mask */
  B200LAB_TERM_JOINT_POS_OUT_OF_MANUAL_LIMIT = 3, /* terminations.py:90  p[0]=lo p[1]=hi, joint_mask */
  B200LAB_TERM_ROOT_HEIGHT_BELOW_MINIMUM = 4,  /* terminations.py:74  p[0]=minimum_height */
  B200LAB_TERM_BAD_ORIENTATION = 5,            /* terminations.py:60  p[0]=limit_angle */
  B200LAB_TERM_EXTERNAL = 100
};

enum B200LabObsOp {
  B200LAB_OBS_BASE_LIN_VEL = 1,                /* LAB/envs/mdp/observations.py:40 */
  B200LAB_OBS_BASE_ANG_VEL = 2,                /* observations.py:47 */
  B200LAB_OBS_PROJECTED_GRAVITY = 3,           /* observations.py:54 */
  B200LAB_OBS_GENERATED_COMMANDS = 4,          /* observations.py:529 */
  B200LAB_OBS_JOINT_POS_REL = 5,               /* observations.py:114 */
  B200LAB_OBS_JOINT_VEL_REL = 6,               /* observations.py:150 */
  B200LAB_OBS_JOINT_POS = 7,                   /* observations.py:104 */
  B200LAB_OBS_JOINT_VEL = 8,                   /* observations.py:140 */
  B200LAB_OBS_LAST_ACTION = 9,                 /* observations.py:512 */
  B200LAB_OBS_HEIGHT_SCAN = 10,                /* observations.py:165  p[0]=offset */
  B200LAB_OBS_BASE_POS_Z = 11,                 /* observations.py:33 */
  B200LAB_OBS_EXTERNAL = 100
};

enum B200LabNoiseKind {                        /* LAB/utils/noise/noise_model.py:20-98 */
  B200LAB_NOISE_NONE = 0,
  B200LAB_NOISE_UNIFORM_ADD = 1,               /* x + u*(hi-lo) + lo */
  B200LAB_NOISE_UNIFORM_SCALE = 2,
  B200LAB_NOISE_UNIFORM_ABS = 3,
  B200LAB_NOISE_CONSTANT_ADD = 4,              /* x + bias (n_lo) */
  B200LAB_NOISE_CONSTANT_SCALE = 5,            /* x * bias */
  B200LAB_NOISE_CONSTANT_ABS = 6,              /* bias */
  /* gaussian_noise (noise_model.py:73-98): z ~ N(0,1); noise_lo = mean, noise_span = std.  In parity mode the
   * element's u_obs slot carries the recorded N(0,1) draw instead of a uniform. */
  B200LAB_NOISE_GAUSSIAN_ADD = 7,              /* (x + mean) + std*z */
  B200LAB_NOISE_GAUSSIAN_SCALE = 8,            /* x * (mean + std*z) */
  B200LAB_NOISE_GAUSSIAN_ABS = 9               /* mean + std*z */
};

typedef struct B200LabTerm {
  int32_t op;
  int32_t time_out;          /* termination terms: 1 -> ORed into truncated, else terminated */
  float weight;              /* reward terms */
  float p[5];
  uint64_t joint_mask;       /* bit j set -> joint j selected (SceneEntityCfg.joint_ids) */
  uint64_t body_mask;        /* bit b set -> body b selected (SceneEntityCfg.body_ids) */
  uint64_t body_mask2;
  int32_t i0;                /* EXTERNAL: column */
  int32_t dim;               /* obs terms: number of columns */
  int32_t noise_kind;
  int32_t clip_on;
  float noise_lo, noise_span; /* n_min and fp32(n_max - n_min) */
  float clip_lo, clip_hi;
  float scale;               /* obs terms; 1.0 when cfg.scale is None */
  int32_t scale_on;
  /* obs terms, LAB/managers/observation_manager.py:305-329.  history_length > 0: the term occupies dim * history_length
   * columns, oldest row first (CircularBuffer.buffer flattened), kept in the ring at float offset `hist_offset` of the
   * env's obs_hist row.  Function modifiers (LAB/utils/modifiers/modifier.py:23-66) run before the noise, in order:
   * kind 1 scale (x * p), 2 bias (x + p), 3 clip (p, p+1); parameters consecutive in mod_p. */
  int32_t history_length;
  int32_t hist_offset;
  int32_t num_mod;
  int32_t mod_kind[3];
  float mod_p[4];
} B200LabTerm;

enum B200LabActuatorKind {
  B200LAB_ACT_IMPLICIT = 1,   /* LAB/actuators/actuator_pd.py:115-140 */
  B200LAB_ACT_IDEAL_PD = 2,   /* actuator_pd.py:184-198 */
  B200LAB_ACT_DC_MOTOR = 3,   /* actuator_pd.py:264-286 */
  B200LAB_ACT_NET_LSTM = 4,   /* LAB/actuators/actuator_net.py:76-99 */
  B200LAB_ACT_NET_MLP = 5,    /* actuator_net.py:147-189 */
  B200LAB_ACT_DELAYED_PD = 6, /* actuator_pd.py:289-347: IdealPD behind a per-env command delay (DelayBuffer) */
  B200LAB_ACT_REMOTIZED_PD = 7 /* actuator_pd.py:349-416: DelayedPD + torque clamp to an angle-dependent limit (lookup table) */
};

/* One actuator group (Articulation.actuators[name], LAB/assets/articulation/articulation.py:1432-1469).
 * Per-joint parameter arrays have length J (articulation joint order); only joints in joint_mask are touched. */
typedef struct B200LabActuatorGroup {
  int32_t kind;
  int32_t _pad;
  uint64_t joint_mask;
  const float* stiffness;      /* [J] */
  const float* damping;        /* [J] */
  const float* effort_limit;   /* [J] */
  const float* velocity_limit; /* [J] */
  float saturation_effort;     /* DCMotor; inf if None */
  /* ActuatorNetLSTM: weights packed per layer l: W_ih[4H,in_l] W_hh[4H,H] b_ih[4H] b_hh[4H]; then head W[H] b[1].
   * State h,c: [L, N*Jg, H] with row = env*Jg + local joint index (reference layout, actuator_net.py:56-64). */
  int32_t lstm_layers, lstm_hidden;
  const float* net_weights;
  float* lstm_h;
  float* lstm_c;
  /* ActuatorNetMLP: layers n_layers; dims[0..n_layers]; weights packed W0[d1,d0] b0[d1] W1 ... ;
   * activation 0=softsign 1=elu 2=relu 3=tanh.  History [N,Hist,Jg] x2 (reference layout, actuator_net.py:134-140). */
  int32_t mlp_layers;
  int32_t mlp_dims[6];
  int32_t mlp_act;
  int32_t mlp_hist;            /* history_length = max(input_idx)+1 */
  int32_t mlp_num_idx;
  int32_t mlp_input_idx[4];
  int32_t mlp_vel_pos_order;   /* 0 = "pos_vel", 1 = "vel_pos" */
  float mlp_pos_scale, mlp_vel_scale, mlp_torque_scale;
  float* mlp_pos_hist;
  float* mlp_vel_hist;
  int32_t group_joints;        /* Jg */
  int32_t _pad2;
  /* DelayedPDActuator (LAB/utils/buffers/delay_buffer.py, circular_buffer.py): per (env, local joint) a ring of
   * max_delay+1 rows {pos, vel, effort set-point} and its push count since the last reset (the first push after a
   * reset fills the ring); per env the lag in physics steps, re-drawn at reset as
   * min_delay + floor(u * (max_delay - min_delay + 1)) from reset-uniform slot `delay_u_slot`. */
  int32_t min_delay, max_delay;
  int32_t delay_u_slot;
  int32_t _pad3;
  float* delay_ring;           /* [N, Jg, max_delay+1, 3] */
  int32_t* delay_count;        /* [N, Jg] */
  int32_t* delay_lag;          /* [N] */
  /* RemotizedPD: |applied torque| <= y(joint_pos), y piecewise linear through (lookup_x[i], lookup_y[i]), i < lookup_n,
   * x ascending, zero-order hold outside (LAB/utils/interpolation/linear_interpolation.py:53-77) */
  int32_t lookup_n;
  int32_t _pad4;
  const float* lookup_x;
  const float* lookup_y;
} B200LabActuatorGroup;

/* ------------------------------------------------------------------------------------------------
 * Environment state: raw views of the tensors the reference keeps in ArticulationData,
 * ContactSensorData, RayCasterData and the managers.  Owned by the caller (torch tensors).
 * ---------------------------------------------------------------------------------------------- */
/* Views of the simulator's state tensors for the current physics step (PhysX tensor views in the reference:
 * root_physx_view.get_root_transforms/velocities, get_dof_positions/velocities, contact_physx_view
 * .get_net_contact_forces).  Passed per call so a recorded slab can be replayed without copies.
 * root_state_w / joint_pos / joint_vel are in/out: reset events write the sampled reset state into them
 * (the reference writes it to PhysX, LAB/assets/articulation/articulation.py:317-556). */
typedef struct B200LabSimView {
  float* root_state_w;         /* [N,13] pos, quat wxyz, lin vel, ang vel */
  float* joint_pos;            /* [N,J] */
  float* joint_vel;            /* [N,J] */
  const float* net_forces_w;   /* [N,B,3] net contact forces of the current physics step (or NULL) */
  const float* body_lin_vel_w; /* [N,B,3] or NULL */
} B200LabSimView;

typedef struct B200LabEnvState {
  int32_t N, J, A, B, R;       /* envs, joints, action dim, bodies, rays */
  int32_t hist;                /* contact history length (>=1) */
  /* --- articulation constants --- */
  const float* default_root_state; /* [13] (identical for all envs, articulation.py:1279) */
  const float* default_joint_pos;  /* [J] */
  const float* default_joint_vel;  /* [J] */
  const float* soft_joint_pos_limits; /* [J,2] */
  float* soft_joint_vel_limits;    /* [J] (written by the actuator pass, articulation.py:1466) */
  /* --- articulation buffers --- */
  float* joint_pos_target;     /* [N,J] */
  float* joint_vel_target;     /* [N,J] */
  float* joint_effort_target;  /* [N,J] */
  float* computed_torque;      /* [N,J] */
  float* applied_torque;       /* [N,J] */
  float* joint_pos_target_sim; /* [N,J] implicit actuators: pos target handed to the simulator */
  float* joint_effort_target_sim; /* [N,J] explicit actuators: effort handed to the simulator */
  float* prev_joint_vel;       /* [N,J] */
  float* joint_acc;            /* [N,J] */
  /* --- action manager (LAB/managers/action_manager.py:318-339) --- */
  float* action;               /* [N,A] */
  float* prev_action;          /* [N,A] */
  float* raw_actions;          /* [N,A] term raw actions (joint_actions.py:131) */
  float* processed_actions;    /* [N,A] */
  const float* action_scale;   /* [A] */
  const float* action_offset;  /* [A] */
  const float* action_clip;    /* [A,2] or NULL */
  const int32_t* action_joint_ids; /* [A] joint index each action column drives */
  int32_t action_kind;         /* 0 = joint position target, 1 = joint effort target */
  int32_t num_act_groups;
  int8_t joint_action_col[64]; /* joint -> action column driving it, -1 if none */
  int8_t joint_group[64];      /* joint -> actuator group index, -1 if un-actuated */
  int8_t joint_local[64];      /* joint -> index of the joint inside its group (row of the net state) */
  B200LabActuatorGroup act[B200LAB_MAX_ACT_GROUPS];
  /* --- contact sensor (LAB/sensors/contact_sensor/contact_sensor.py:320-379) --- */
  float* cs_net_forces_w;      /* [N,B,3] */
  float* cs_forces_hist;       /* [N,hist,B,3] */
  float* cs_last_air_time;     /* [N,B] */
  float* cs_current_air_time;  /* [N,B] */
  float* cs_last_contact_time; /* [N,B] */
  float* cs_current_contact_time; /* [N,B] */
  float* cs_timestamp;         /* [N] */
  float* cs_timestamp_last_update; /* [N] */
  uint8_t* cs_is_outdated;     /* [N] */
  float cs_force_threshold;
  float cs_update_period;
  /* --- ray caster (LAB/sensors/ray_caster/ray_caster.py:220-260) --- */
  float* rc_timestamp;         /* [N] */
  float* rc_timestamp_last_update; /* [N] */
  uint8_t* rc_is_outdated;     /* [N] */
  float* rc_drift;             /* [N,3] */
  float* rc_pos_w;             /* [N,3] */
  float* rc_quat_w;            /* [N,4] */
  float* rc_ray_hits_w;        /* [N,R,3] */
  const float* rc_ray_starts;  /* [R,3] local pattern incl. offset (identical per env) */
  const float* rc_ray_dirs;    /* [R,3] */
  float rc_update_period;
  float rc_max_distance;
  int32_t rc_attach_yaw_only;
  float rc_drift_lo, rc_drift_span;
  const void* rc_mesh;         /* device-side mesh accel handle from b200lab_mesh_create (or NULL) */
  /* --- env counters --- */
  int64_t* episode_length_buf; /* [N] */
  int64_t max_episode_length;
  float step_dt;
  float physics_dt;
  float max_episode_length_s;
  /* --- terminations (LAB/managers/termination_manager.py:151-174) --- */
  int32_t num_term;
  B200LabTerm term[B200LAB_MAX_TERMS];
  uint8_t* term_dones;         /* [num_term, N] */
  uint8_t* terminated;         /* [N] */
  uint8_t* truncated;          /* [N] */
  uint8_t* reset_buf;          /* [N] */
  /* --- rewards (LAB/managers/reward_manager.py:128-157) --- */
  int32_t num_rew;
  B200LabTerm rew[B200LAB_MAX_TERMS];
  float* reward_buf;           /* [N] */
  float* episode_sums;         /* [num_rew, N] */
  float* step_reward;          /* [N, num_rew] */
  const float* ext_values;     /* [n_ext, N] external term columns or NULL */
  /* --- command: UniformVelocityCommand (LAB/envs/mdp/commands/velocity_command.py:111-160) --- */
  int32_t has_command;
  int32_t cmd_heading_command;
  float* cmd_vel_b;            /* [N,3] */
  float* cmd_heading_target;   /* [N] */
  uint8_t* cmd_is_heading;     /* [N] */
  uint8_t* cmd_is_standing;    /* [N] */
  float* cmd_time_left;        /* [N] */
  int64_t* cmd_counter;        /* [N] */
  float* cmd_metric_xy;        /* [N] */
  float* cmd_metric_yaw;       /* [N] */
  float cmd_resample_lo, cmd_resample_span;
  float cmd_range[8];          /* lin_x lo,hi, lin_y lo,hi, ang_z lo,hi, heading lo,hi */
  float cmd_span[4];           /* fp32(hi) - fp32(lo) of the four ranges (Tensor.uniform_) */
  float cmd_rel_standing, cmd_rel_heading, cmd_heading_stiffness;
  float cmd_max_command_step;  /* resampling_time_range[1] / step_dt (fp32 of the double quotient) */
  /* --- events --- */
  int32_t ev_reset_root_state; /* reset_root_state_uniform present (LAB/envs/mdp/events.py:823) */
  /* every sampled range is passed as (lo, span): value = u * span + lo, with `span` computed by the host exactly
   * as the reference computes `upper - lower` (fp32 tensor difference or python-double difference cast to fp32) */
  float ev_pose_range[12];     /* x y z roll pitch yaw  (lo, span) */
  float ev_vel_range[12];
  /* reset_joints_by_scale (events.py:987) / reset_joints_by_offset (events.py:1020) terms in cfg order; each one
   * rewrites ALL joints (the reference ignores asset_cfg.joint_ids here), so the last one determines the state.
   * Term k reads u_reset columns [B200LAB_U_RESET_JOINT + 2*J*k, +J) for positions and the next J for velocities. */
  int32_t ev_num_reset_joints;
  int32_t ev_reset_joints_mode[B200LAB_MAX_RESET_JOINT_EVENTS];  /* 1 by_scale, 2 by_offset */
  float ev_joint_pos_range[B200LAB_MAX_RESET_JOINT_EVENTS][2];   /* (lo, span) */
  float ev_joint_vel_range[B200LAB_MAX_RESET_JOINT_EVENTS][2];
  int32_t num_interval;        /* push_by_setting_velocity terms (events.py:795) */
  float* ev_interval_time_left[B200LAB_MAX_INTERVAL_EVENTS]; /* [N] each */
  float ev_interval_range[B200LAB_MAX_INTERVAL_EVENTS][2];  /* (lo, span) */
  float ev_push_range[B200LAB_MAX_INTERVAL_EVENTS][12];     /* (lo, span) x 6 */
  /* EventManager gating (LAB/managers/event_manager.py:204-270).
   * reset mode, min_step_count_between_reset > 0: the term is applied to an env being reset only if
   * global_step - last_triggered >= min_steps, or it has never been triggered; state per term: last_step i32 [N], once u8 [N]
   * (NULL pointers / 0 = apply on every reset).  global_step = env._sim_step_counter // decimation = step_counter + 1.
   * interval mode, is_global_time: ONE timer for all envs, kept as two floats indexed by step parity (read [step & 1], env 0
   * writes [(step + 1) & 1]); when it fires every env is pushed and the new interval is env 0's draw. */
  int32_t ev_root_min_steps;
  int32_t ev_reset_joints_min_steps[B200LAB_MAX_RESET_JOINT_EVENTS];
  int32_t* ev_root_last_step;
  uint8_t* ev_root_once;
  int32_t* ev_reset_joints_last_step[B200LAB_MAX_RESET_JOINT_EVENTS];
  uint8_t* ev_reset_joints_once[B200LAB_MAX_RESET_JOINT_EVENTS];
  int32_t ev_interval_global[B200LAB_MAX_INTERVAL_EVENTS];
  float* ev_interval_global_time[B200LAB_MAX_INTERVAL_EVENTS];   /* [2] each, or NULL */
  /* --- curriculum: terrain_levels_vel (TASKS/.../velocity/mdp/curriculums.py:26-55) --- */
  int32_t has_terrain_curriculum;
  int64_t* terrain_levels;     /* [N] */
  const int64_t* terrain_types;/* [N] */
  float* env_origins;          /* [N,3] */
  const float* terrain_origins;/* [rows, cols, 3] */
  int32_t terrain_rows, terrain_cols;
  float terrain_size_x;        /* terrain_generator.size[0] */
  /* --- observations (LAB/managers/observation_manager.py:260-335), one concatenated group --- */
  int32_t num_obs;
  int32_t obs_dim;
  B200LabTerm obs[B200LAB_MAX_TERMS];
  float* obs_buf;              /* [N, obs_dim] */
  int32_t obs_enable_corruption;
  /* observation history (ObservationTermCfg.history_length, LAB/utils/buffers/circular_buffer.py): obs_hist [N,
   * obs_hist_total] holds every history term's ring, row-major [history_length, dim] at the term's hist_offset.
   * obs_hist_count[0] = appends so far (b200lab_post_step advances it once per step, so the observation computed at
   * reset time and the one of the first step use consecutive slots).  obs_hist_reset_step[n] = env step at which env n
   * was last reset (written by post_step): a compute_obs of that same step fills every slot of the env's rings with the
   * fresh row (the first push after a reset).  The host initialises it to the current step counter, so the first
   * compute fills every env, and sets it to -1 after a compute_obs issued outside a step (the observation returned by
   * env.reset()).  All NULL / 0 when no term keeps history. */
  int32_t obs_hist_total;
  int32_t _pad_hist;
  float* obs_hist;
  int32_t* obs_hist_count;
  int64_t* obs_hist_reset_step;
  /* --- randomness (SURVEY.md 7 hard part 3) --- */
  /* parity mode: pre-drawn uniforms in [0,1): u_reset [N, n_u_reset], u_step [N, n_u_step], u_obs [N, obs_dim];
   * perf mode (pointers NULL): Philox4x32-10 keyed by (seed, step_counter, env, slot). */
  const float* u_reset;
  const float* u_step;
  const float* u_obs;
  int32_t n_u_reset, n_u_step;
  uint64_t seed;
  /* optional device-side step counter (CUDA-graph replay): when non-NULL, process_action increments it and the
   * kernels key their Philox streams on it instead of the host-passed `step_counter` argument */
  int64_t* step_counter_dev;
  /* device flag "some env was reset in this step": cleared by process_action / decimation_fused, set by post_step,
   * read by post_reset_write (required by that call, else optional) */
  int32_t* any_reset_dev;
  /* --- episode log accumulators (extras["log"], LAB/envs/manager_based_rl_env.py:365-389) --- */
  double* log_accum;           /* [B200LAB_LOG_SLOTS] per-step sums over the envs reset in this step (post_step adds,
                                  b200lab_log_fold consumes and zeroes) */
  double* log_running;         /* [2 * B200LAB_LOG_SLOTS] or NULL: running sums of the per-step log VALUES (slot 0 = number of
                                  steps that had resets), then the values of the last folded step (slot 0 = its reset count) */
} B200LabEnvState;

/* Layout of u_reset columns (per reset env): */
#define B200LAB_U_RESET_CURRICULUM 0   /* randint_like(level, max_level) as floor(u*max_level) */
#define B200LAB_U_RESET_DRIFT 1        /* 3: RayCaster.reset drift */
#define B200LAB_U_RESET_POSE 4         /* 6 */
#define B200LAB_U_RESET_VEL 10         /* 6 */
#define B200LAB_U_RESET_CMD 16         /* 7: time_left, x, y, z, heading, is_heading, is_standing */
#define B200LAB_U_RESET_JOINT 23       /* per reset-joints term: 2*J (joint pos scale/offset [J], joint vel [J]) */
/* Layout of u_step columns (per env): */
#define B200LAB_U_STEP_CMD 0           /* 7 (same order as above) */
#define B200LAB_U_STEP_INTERVAL 7      /* per interval term k: 7 at 7+7k: time, x y z roll pitch yaw */

/* log_accum layout: [0] = number of reset envs; [1 .. 1+num_rew) = sum over reset envs of episode_sums;
 * then num_term termination counts; then 2 command metric sums; then 1 terrain level sum (all envs). */
#define B200LAB_LOG_SLOTS 128
/* extras["log"] with the reference's semantics (manager_based_rl_env.py:365-389: one dict per step that had resets, each
 * value a statistic over THAT step's reset envs; the runner averages the dicts).  Per-step values, in log_accum's slot
 * order: Episode_Reward/<term> = mean over the step's reset envs of the episodic sum / max_episode_length_s
 * (reward_manager.py:100-126); Episode_Termination/<term> = number of those envs the term fired for
 * (termination_manager.py:127-149); Metrics/<command>/error_vel_{xy,yaw} = mean metric (command_manager.py:120-149);
 * slot 3+num_rew+num_term: Curriculum/terrain_levels = mean level over ALL envs (curriculums.py:55).
 * b200lab_log_fold turns the sums post_step left in log_accum into these values, adds them to log_running, keeps them as
 * "last step", and zeroes log_accum.  b200lab_step_tail[_begin] issues it on the library's side stream; callers that use
 * post_step / compute_obs separately call it once per step after post_step. */
int b200lab_log_fold(void* env, void* stream);

/* The env descriptor is uploaded to the device once; kernels read it through a pointer (it is constant per
 * environment instance).  b200lab_env_update re-uploads after a cfg change (e.g. RewardManager.set_term_cfg). */
void* b200lab_env_create(const B200LabEnvState* st);
int b200lab_env_update(void* env, const B200LabEnvState* st);
void b200lab_env_destroy(void* env);

/* ActionManager.process_action (LAB/managers/action_manager.py:318-339) + JointAction.process_actions
 * (LAB/envs/mdp/actions/joint_actions.py:130-139).  `action_in` [N,A]. */
int b200lab_process_action(void* env, const float* action_in, void* stream);

/* The two halves of one physics sub-step the reference has around sim.step() (manager_based_rl_env.py:182-196):
 *   pre : action_manager.apply_action + scene.write_data_to_sim -> Articulation._apply_actuator_model
 *         (reads the state BEFORE the physics step, writes the efforts/targets handed to the simulator)
 *   post: scene.update(dt): ArticulationData.update (joint_acc), ContactSensor update, RayCaster timestamps
 *         (reads the state AFTER the physics step) */
int b200lab_substep_pre(void* env, const B200LabSimView* sim, void* stream);
int b200lab_substep_post(void* env, const B200LabSimView* sim, void* stream);

/* Speculative height scan.  Replaces nothing in the reference; it re-schedules RayCaster._update_buffers_impl
 * (LAB/sensors/ray_caster/ray_caster.py:220-260): the sensor pose after the last physics sub-step only changes
 * afterwards for envs that are reset, so the rays of the height scanner (vertical pattern, attach_yaw_only) can be cast
 * from `sim` on an internal stream while the caller's stream runs the actuator / contact / termination / reward
 * kernels.  b200lab_compute_obs joins the scan, uses its hits for every env whose sensor pose (root pose + drift) is
 * bitwise the pose the scan was cast from, and casts the others (reset envs) in place -- results are identical with
 * and without the call.  No-op for other sensor configurations.  Opt-in: on B200 at 4096 envs the step kernels already
 * fill the register file, so the overlap only re-distributes time (measurement). */
int b200lab_height_scan_prefetch(void* env, const B200LabSimView* sim, void* stream);

/* REPLAY fast path: the whole physics loop of one env step in one launch, when the simulator states of all
 * `num_substeps` (= decimation, <= 8) physics steps are already known (recorded Isaac Sim run or synthetic state):
 * process_action, then for s = 0..S-1 { substep_pre on state s-1 (`prev` for s = 0); substep_post on `views[s]` }.
 * Identical arithmetic and ordering to the three calls above; per-thread state stays in registers across sub-steps.
 * `views` is a HOST array of num_substeps views. */
int b200lab_decimation_fused(void* env, const float* action_in, const B200LabSimView* prev,
                             const B200LabSimView* views, int32_t num_substeps, void* stream);

/* Everything in ManagerBasedRLEnv.step after the physics loop and before the observations
 * (manager_based_rl_env.py:198-236): episode_length_buf += 1, TerminationManager.compute,
 * RewardManager.compute, _reset_idx for the envs in reset_buf (masked, no host sync),
 * CommandManager.compute, interval events.  `step_counter` keys the Philox streams (perf mode) and is
 * env._sim_step_counter // decimation for the reset-event bookkeeping.  * Stream ordering: b200lab_post_step and b200lab_compute_obs[_range] are launched with programmatic stream serialization;
 * before the dependency on the preceding kernel of the stream is resolved they read only library-owned buffers that
 * kernel cannot write (command state, episode sums, event timers, term/column tables) -- the caller's simulator tensors
 * are read after it, so any kernel of the caller may directly precede them. */
int b200lab_post_step(void* env, const B200LabSimView* sim, int64_t step_counter, void* stream);

/* The `scene.write_data_to_sim()` that step() issues right after `_reset_idx` when ANY env was reset
 * (manager_based_rl_env.py:217-223): Articulation._apply_actuator_model runs once more for ALL envs -- stateful
 * actuator nets advance, computed/applied torque are rewritten -- on the joint state that holds the reset writes and
 * the targets of the last apply_action.  The condition is evaluated on the device (`any_reset_dev`), so the call is
 * unconditional on the host and needs no synchronisation; `sim` is the view post_step was given. */
int b200lab_post_reset_write(void* env, const B200LabSimView* sim, void* stream);

/* ObservationManager.compute (observation_manager.py:238-335) including RayCaster lazy update +
 * raycast_mesh for outdated envs (ray_caster.py:220-260, LAB/utils/warp/ops.py:24-127).  * The per-env quantities shared by the columns (base-frame velocities, projected gravity, sensor pose; the ray-caster's
 * outdated/timestamp bookkeeping of sensor_base.py:287-296) are derived by b200lab_post_step from the state it leaves and
 * stay valid until the next call that advances the env (process_action, decimation_fused, substep_pre/post,
 * env_update); called without a preceding post_step on the same view, compute_obs derives them itself.  Writing the
 * simulator tensors between post_step and compute_obs therefore requires b200lab_env_update. */
int b200lab_compute_obs(void* env, const B200LabSimView* sim, int64_t step_counter, void* stream);

/* post_reset_write and compute_obs issued together: the actuator pass runs on a library-owned side stream forked from
 * and joined back into `stream` (both are latency-bound and touch disjoint buffers), everything `stream` enqueues
 * afterwards sees both results.  Same results as the two calls in sequence. */
int b200lab_step_tail(void* env, const B200LabSimView* sim, int64_t step_counter, void* stream);

/* The same work split so that the caller can pipeline its own consumers behind the observation kernels: `_begin` forks
 * the conditional post-reset actuator pass onto the library's side stream, `b200lab_compute_obs_range` computes the
 * observation rows of envs [env_begin, env_end) (any partition of [0, N) gives the rows b200lab_compute_obs gives),
 * `_end` joins the side stream.  bench.py's end-to-end arm copies each finished range to the host while the next one
 * is computed. */
int b200lab_step_tail_begin(void* env, const B200LabSimView* sim, void* stream);
int b200lab_compute_obs_range(void* env, const B200LabSimView* sim, int64_t step_counter, int32_t env_begin,
                              int32_t env_end, void* stream);
int b200lab_step_tail_end(void* env, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Ray casting.  Replaces convert_to_warp_mesh (LAB/utils/warp/ops.py:130-145) and raycast_mesh
 * (ops.py:24-127 / kernels.py:13-75).
 * ---------------------------------------------------------------------------------------------- */
/* Builds the device-side acceleration structure from HOST vertex/face arrays.  Returns an opaque handle
 * (host pointer to a descriptor; `b200lab_mesh_device_ptr` gives the device-side struct for rc_mesh). */
void* b200lab_mesh_create(const float* h_vertices, int64_t num_vertices, const int32_t* h_faces,
                          int64_t num_faces, float cell_size);
const void* b200lab_mesh_device_ptr(void* mesh);
int64_t b200lab_mesh_bytes(void* mesh);
void b200lab_mesh_destroy(void* mesh);
/* ray_starts/ray_dirs [n,3]; ray_hits [n,3] (inf on miss); optional ray_distance [n], ray_normal [n,3],
 * ray_face_id [n] (int32, -1 on miss) may be NULL. */
int b200lab_raycast_mesh(void* mesh, const float* ray_starts, const float* ray_dirs, int64_t n, float max_dist,
                         float* ray_hits, float* ray_distance, float* ray_normal, int32_t* ray_face_id,
                         void* stream);

/* ------------------------------------------------------------------------------------------------
 * rsl_rl hot path.
 * ---------------------------------------------------------------------------------------------- */
/* RolloutStorage.compute_returns (RSL/storage/rollout_storage.py:127-149).
 * rewards, values [T,N]; dones u8 [T,N]; last_values [N]; returns, advantages [T,N] (out);
 * scratch: >= 2*(number of CTAs)+2 doubles, see b200lab_gae_scratch_bytes. */
int64_t b200lab_gae_scratch_bytes(int32_t T, int32_t N);
int b200lab_gae(const float* rewards, const float* values, const uint8_t* dones, const float* last_values,
                float* returns, float* advantages, int32_t T, int32_t N, float gamma, float lam,
                int32_t normalize_advantage, void* scratch, void* stream);

/* b200lab_gae WITHOUT caller scratch and without any allocation.  No counterpart in the reference: added
 * for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  Same kernels, launch geometry and results (bitwise) as b200lab_gae; the per-CTA partial sums
 * live in a file-scope `static __device__` array inside the library: 2 * (B200LAB_GAE_STATIC_MAX_ENVS / 64) + 2 doubles
 * = 64 KB, held by every GPU that loads the library until the process exits.  N > B200LAB_GAE_STATIC_MAX_ENVS is rejected
 * with B200LAB_ERR_ARG.  The call makes no cudaMalloc / cudaFree / allocating library call; the first call on a device
 * resolves the array's device address (cudaGetSymbolAddress, with that device current) -- make it once before capturing.
 * EVERY call on a device shares the one array and it keeps the previous call's contents (nothing needs clearing: each
 * call overwrites the slots it reads), so two calls that overlap on the GPU -- different streams, or graphs replayed
 * concurrently -- RACE on it: serialise them, or use b200lab_gae with separate scratch buffers. */
#define B200LAB_GAE_STATIC_MAX_ENVS 262144
int b200lab_gae_static(const float* rewards, const float* values, const uint8_t* dones, const float* last_values,
                       float* returns, float* advantages, int32_t T, int32_t N, float gamma, float lam,
                       int32_t normalize_advantage, void* stream);

/* PPO.process_env_step time-out bootstrap + RolloutStorage.add_transitions (RSL/algorithms/ppo.py:142-169,
 * rollout_storage.py:77-103) for one step `t` of the [T,N,.] slabs. */
typedef struct B200LabTransition {
  int32_t T, N, obs_dim, act_dim, t;
  float gamma;
  const float* obs;          /* [N,obs_dim], or NULL when the caller has already stored the row of step t */
  const float* actions;      /* [N,A] */
  const float* rewards;      /* [N] */
  const uint8_t* dones;      /* [N] (u8) */
  const uint8_t* time_outs;  /* [N] or NULL */
  const float* values;       /* [N] */
  const float* log_prob;     /* [N] */
  const float* mu;           /* [N,A] */
  const float* sigma;        /* [N,A] */
  float* s_obs; float* s_actions; float* s_rewards; uint8_t* s_dones; float* s_values;
  float* s_log_prob; float* s_mu; float* s_sigma;
} B200LabTransition;
int b200lab_add_transition(const B200LabTransition* tr, void* stream);

/* RolloutStorage.mini_batch_generator gathers (rollout_storage.py:188-197): one fused gather of the eight
 * flattened [T*N,.] slabs by `idx` (int64 [M]). */
typedef struct B200LabGather {
  int64_t M; int32_t obs_dim, act_dim;
  const int64_t* idx;
  const float* obs; const float* actions; const float* values; const float* returns;
  const float* log_prob; const float* advantages; const float* mu; const float* sigma;
  float* o_obs; float* o_actions; float* o_values; float* o_returns;
  float* o_log_prob; float* o_advantages; float* o_mu; float* o_sigma;
  int64_t o_obs_ld;   /* row stride of o_obs in elements; 0 = obs_dim.  A caller that wants 16-byte aligned observation rows
                       * (the first MLP layer's TMA operand) passes a zero-initialised [M, ld] buffer: pad columns are not written */
} B200LabGather;
int b200lab_gather_minibatch(const B200LabGather* g, void* stream);

/* OnPolicyRunner.learn's finished-episode bookkeeping (RSL/runners/on_policy_runner.py:113-132): cur_reward_sum +=
 * rewards, cur_episode_length += 1, push both for envs with dones > 0 (ascending env id, as nonzero()) into rings of
 * `ring_size` (the reference's deque(maxlen=100)), zero them.  `pushed` (device int64) counts pushes since creation;
 * slot = push index % ring_size.  dones_elem_size: 1 (u8/bool) or 8 (int64, the VecEnv contract's dtype). */
int b200lab_episode_stats(const float* rewards, const void* dones, int32_t dones_elem_size, int32_t N,
                          float* cur_reward_sum, float* cur_episode_length, float* ring_reward, float* ring_length,
                          int32_t ring_size, int64_t* pushed, void* stream);

/* Recurrent policy cells (RSL/networks/memory.py:21-38 -> torch.nn.LSTM / torch.nn.GRU; used by
 * RSL/modules/actor_critic_recurrent.py:141-160).  The matrix products of a step are b200lab_linear_fwd calls
 * (gi = x W_ih^T + b_ih for all time steps at once, gh = h W_hh^T + b_hh per step); these entry points are the gate
 * arithmetic and its derivative.  All tensors row-major fp32; B = trajectories in the batch, H = hidden size.
 * LSTM: gi, gh, gates, dgates [B,4H] in torch's gate order i,f,g,o; `gates` (post-activation, may be NULL in inference)
 * is what backward reads.  `dc` is read as dL/dc_t and overwritten with dL/dc_{t-1}; dh_out / dh_rec (either may be NULL)
 * are the gradient of h_t from the layer above / from step t+1.
 * GRU: gi, gh, dgi, dgh [B,3H] in order r,z,n; `saved` [B,4H] = r|z|n|(W_hn h + b_hn).  dh_dir_in / dh_dir_out carry
 * the z * h_{t-1} path of dL/dh between steps (they may be the same buffer); dh_rec is dgh . W_hh of step t+1. */
int b200lab_lstm_cell_fwd(const float* gi, const float* gh, const float* c_prev, float* h, float* c, float* gates,
                          int64_t B, int32_t H, void* stream);
int b200lab_lstm_cell_bwd(const float* gates, const float* c_prev, const float* c, const float* dh_out,
                          const float* dh_rec, float* dc, float* dgates, int64_t B, int32_t H, void* stream);
int b200lab_gru_cell_fwd(const float* gi, const float* gh, const float* h_prev, float* h, float* saved, int64_t B,
                         int32_t H, void* stream);
int b200lab_gru_cell_bwd(const float* saved, const float* h_prev, const float* dh_out, const float* dh_rec,
                         const float* dh_dir_in, float* dh_dir_out, float* dgi, float* dgh, int64_t B, int32_t H,
                         void* stream);

/* Actor-critic MLP (RSL/networks/mlp.py:60-80: Linear+ELU ... Linear).  Y = act(X W^T + b).
 * X [M,K] row-major, W [Nout,K] row-major (nn.Linear.weight), b [Nout], Y [M,Nout].
 * precision: 0 = fp32 FFMA (exact-order reference path), 1 = tcgen05 kind::tf32, 2 = tcgen05 3xTF32 split. */
int b200lab_linear_fwd(const float* X, const float* W, const float* b, float* Y, int64_t M, int32_t K,
                       int32_t Nout, int32_t elu, int32_t precision, void* stream);
/* Backward of the same layer.  dY [M,Nout] is the gradient w.r.t. the POST-activation output; Y is the saved
 * post-activation output (ELU'(x) recovered from Y).  dX may be NULL (first layer).  dW [Nout,K], db [Nout]
 * are OVERWRITTEN.  dZ is a [M,Nout] scratch that receives dY*ELU'(.); `workspace` holds the split
 * partial sums of the dW reduction (b200lab_linear_bwd_workspace_bytes). */
int64_t b200lab_linear_bwd_workspace_bytes(int64_t M, int32_t K, int32_t Nout);
int b200lab_linear_bwd(const float* X, const float* W, const float* Y, const float* dY, float* dZ, float* dX,
                       float* dW, float* db, int64_t M, int32_t K, int32_t Nout, int32_t elu, int32_t precision,
                       void* workspace, void* stream);
/* The same backward for a layer inside a Linear+ELU chain (autograd of RSL/networks/mlp.py:60-80 walks the layers from the
 * output back; each ELU's backward is an elementwise pass between two matrix products).  flags:
 *   B200LAB_BWD_ELU_OUT  as elu = 1 above: dZ = dY * ELU'(Y), db = column sums of dZ
 *   B200LAB_BWD_ELU_IN   X is itself an ELU output (the previous layer's activation): dX receives (dZ W) * ELU'(X), i.e. the
 *                        previous layer's dZ (in place), and db_in [K] its column sums (that layer's bias gradient); the
 *                        reductions of dW, db and db_in then share one launch
 *   B200LAB_BWD_DZ_GIVEN dY already is this layer's dZ and db is already filled (the previous call ran with ELU_IN)
 * Everything else as b200lab_linear_bwd (same workspace size). */
#define B200LAB_BWD_ELU_OUT 1
#define B200LAB_BWD_ELU_IN 2
#define B200LAB_BWD_DZ_GIVEN 4
int b200lab_linear_bwd_chain(const float* X, const float* W, const float* Y, const float* dY, float* dZ, float* dX,
                             float* dW, float* db, float* db_in, int64_t M, int32_t K, int32_t Nout, int32_t flags,
                             int32_t precision, void* workspace, void* stream);

/* Rollout-side sampling (ActorCritic.act / get_actions_log_prob, RSL/modules/actor_critic.py:115-143):
 * actions = mu + std*eps; log_prob = sum_a Normal(mu,std).log_prob(actions). eps [M,A] standard normal draws. */
int b200lab_sample_actions(const float* mu, const float* std, const float* eps, float* actions, float* log_prob,
                           float* sigma_out, int64_t M, int32_t A, void* stream);

/* PPO loss forward+backward for one minibatch (RSL/algorithms/ppo.py:249-315).
 * Outputs d_mu [M,A], d_value [M], d_std [A] (+=atomic, caller zeroes), and stats[8] (device):
 * [0]=surrogate loss, [1]=value loss, [2]=mean entropy, [3]=mean KL, [4..7] scratch. */
typedef struct B200LabPpoLoss {
  int64_t M; int32_t A;
  const float* mu; const float* std; const float* value;   /* current policy outputs: [M,A], [A], [M] */
  const float* actions; const float* old_log_prob; const float* old_mu; const float* old_sigma;
  const float* advantages; const float* returns; const float* target_values;
  float clip_param, value_loss_coef, entropy_coef;
  int32_t use_clipped_value_loss;
  float* d_mu; float* d_value; float* d_std; float* stats;
} B200LabPpoLoss;
int b200lab_ppo_loss(const B200LabPpoLoss* a, void* stream);

/* clip_grad_norm_ + Adam.step over flat parameter/gradient buffers (ppo.py:380-381).
 * lr is read from device memory (lr_dev[0]) so that the KL-adaptive schedule needs no host sync;
 * b200lab_adapt_lr updates it from stats[3] (ppo.py:260-294). */
int b200lab_adapt_lr(float* lr_dev, const float* stats, float desired_kl, float inv_world_size, void* stream);
int b200lab_grad_sumsq(const float* grad, int64_t n, float* out_sumsq, void* stream);
int b200lab_adam_clip_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n,
                           const float* lr_dev, const float* grad_sumsq, float grad_scale, float max_grad_norm,
                           float beta1, float beta2, float eps, int64_t step, void* stream);

/* EmpiricalNormalization (RSL/networks/normalization.py:14-72).  update: learn the batch x[M,D] into the running
 * mean / var / std ([D] each); count_new = samples seen INCLUDING this batch (the host owns the counter: M is known
 * there, and `until` is checked before the call).  apply: out = (x - mean) / (std + eps). */
int b200lab_empirical_norm_update(const float* x, int64_t M, int32_t D, int64_t count_new, float* mean, float* var,
                                  float* std, void* stream);
int b200lab_empirical_norm_apply(const float* x, const float* mean, const float* std, float eps, float* out, int64_t M,
                                 int32_t D, void* stream);
/* The same update with GLOBAL statistics over the ranks of a multi-GPU run (normalization.py:50-66 applied to the union of
 * the ranks' batches; BASELINE north_star: "NCCL ... for global observation-normaliser statistics").  `partial` reduces the
 * rank's batch to sums[0..D) = sum x, sums[D..2D) = sum x^2 (fp64), sums[2D] = M; the caller all-reduces (SUM) the 2D+1
 * doubles; `merge` applies the running update with rate = M_total / (count_old + M_total), count_old = samples seen by ALL
 * ranks before this batch.  With one rank the result equals b200lab_empirical_norm_update up to fp32 rounding of the
 * batch variance (one-pass fp64 instead of two-pass fp32). */
int b200lab_empirical_norm_partial(const float* x, int64_t M, int32_t D, double* sums, void* stream);
int b200lab_empirical_norm_merge(const double* sums, int32_t D, int64_t count_old, float* mean, float* var, float* std,
                                 void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200LAB_H_ */
