/*
 * dexstep.h — C ABI of libdexstep.so: the B200-native per-step tensor pipeline of dexhand_env.
 *
 * The reference (CloudEngineHub/dexrobot_isaac) has NO FFI for this path: it is eager PyTorch
 * behind Python component classes.  This ABI is new and sits *below* those classes; each entry
 * point names the reference method(s) it replaces (paths relative to dexhand_env/):
 *
 *   dex_pre_physics      ActionProcessor.process_actions        components/action/action_processor.py:284-352
 *                        (+ DefaultActionRules                  components/action/default_rules.py:33-112,
 *                           ActionRules post filters            components/action/rules.py:113-190,
 *                           ActionProcessor.apply_coupling      components/action/action_processor.py:571-614,
 *                           ObservationEncoder.update_prev_actions  components/observation/observation_encoder.py:287-296)
 *   dex_post_physics     StepProcessor.process_physics_step     components/step_processor.py:37-131
 *                        (= TensorManager.refresh_tensors gather components/physics/tensor_manager.py:441-446,
 *                           ObservationEncoder.compute_observations/concatenate_observations
 *                                                               components/observation/observation_encoder.py:442-494,576-986,1483-1521,
 *                           BlindGraspingTask.get_task_observations/check_*_criteria/compute_task_reward_terms
 *                                                               tasks/blind_grasping_task.py:549-1364,
 *                           TerminationManager.evaluate         components/termination/termination_manager.py:120-321,
 *                           RewardCalculator.compute_common_reward_terms/compute_total_reward
 *                                                               components/reward/reward_calculator.py:66-275,
 *                           and, for envs whose reset flag fires, DexHandBase.reset_idx -> ResetManager.reset_idx
 *                                                               tasks/dexhand_base.py:743-756, components/reset/reset_manager.py:92-190)
 *   dex_reset_idx        DexHandBase.reset_idx (masked)         tasks/dexhand_base.py:743-756 (+ reset_manager.py:92-190,
 *                                                               blind_grasping_task.py:433-547, action_processor.py:524-568,616-666,
 *                                                               observation_encoder.py:363-383, termination_manager.py:341-357)
 *   dex_step_epilogue    torch.nonzero(reset_buf) + termination rates + update_consecutive_successes
 *                                                               step_processor.py:109-111, termination_manager.py:259-266,323-339
 *   dex_apply_coupling / dex_extract_active_targets
 *                        ActionProcessor.apply_coupling / extract_active_targets_from_full_dof
 *                                                               action_processor.py:571-614 / 616-666
 *   dex_reset_actor_indices  the index tensors of the indexed gym setters   components/reset/reset_manager.py:139-168,
 *                                                               components/physics/physics_manager.py:121-160 (fixed capacity, no host read)
 *   dex_post_physics_noalloc  (no counterpart in the reference: dex_post_physics + fused epilogue with library-owned scratch)
 *   dex_point_in_hand_frame  utils/coordinate_transforms.py:17-35 (known-answer tests)
 *   (DexOutBuffers.episode_sums/.stat_totals: RewardComponentObserver accumulation, rl/reward_observer.py:160-293,
 *    fused into dex_post_physics; the totals vector is the payload of the single NCCL all-reduce)
 *
 * Conventions
 *   - extern "C", POD structs only, raw device pointers, element counts as int32.  No torch types.
 *   - All pointers are DEVICE pointers unless a parameter is documented as host.  The caller
 *     (PyTorch / Isaac Gym) owns every buffer; the library never allocates or frees device memory.
 *   - Every function is asynchronous w.r.t. the host, ordered on the passed cudaStream_t
 *     (passed as void*), performs no allocation and no synchronisation and reads no device result on
 *     the host => safe to capture in a CUDA graph.  Pointers must stay valid/fixed while a captured
 *     graph that contains the call is alive.
 *   - Return value: 0 on success, a negative DEX_ERR_* for argument errors detected before launch,
 *     or a positive cudaError_t from the launch.  dex_strerror() maps both.
 *   - Layouts.  Buffers dictated by Isaac Gym / the policy are row-major AoS exactly as the reference
 *     has them: dof_state [N,26,2], rigid_body_state [N,NB,13], net_contact_force [N,NB,3],
 *     actor_root_state [N,NA,13], actions [N,NACT], obs_buf [N,obs_dim], full_dof_targets [N,26].
 *     Path-internal persistent state is stored TRANSPOSED (SoA, [K][N], env index fastest) so that a
 *     thread-per-env kernel reads and writes it fully coalesced; the Python layer exposes it as
 *     [N,K] strided views.
 */
#ifndef DEXSTEP_H_
#define DEXSTEP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DEX_ABI_VERSION 11

#define DEX_NUM_DOF 26          /* constants.py:8-10: 6 base + 20 finger DOFs */
#define DEX_NUM_BASE 6
#define DEX_NUM_ACTIVE 18       /* 6 base + 12 finger controls */
#define DEX_NUM_ACTIVE_FINGER 12
#define DEX_NUM_FINGER_DOF 20
#define DEX_NUM_FINGERS 5
#define DEX_NUM_UNIFORMS 29     /* box x,y,yaw | hand translation 3 | hand rotation 3 | finger DOFs 20 */

/* error codes (negative); positive values are cudaError_t */
#define DEX_OK 0
#define DEX_ERR_NULL -1
#define DEX_ERR_SIZE -2
#define DEX_ERR_ALIGN -3
#define DEX_ERR_PARAM -4
#define DEX_ERR_VERSION -5
#define DEX_ERR_CAPACITY -6   /* dex_post_physics_noalloc: n_envs exceeds DEX_NOALLOC_MAX_ENVS */

/* task / mode enums */
#define DEX_TASK_BASE 0
#define DEX_TASK_BLIND_GRASPING 1
#define DEX_CONTROL_POSITION 0
#define DEX_CONTROL_POSITION_DELTA 1
#define DEX_MODE_FULL 0      /* StepProcessor.process_physics_step */
#define DEX_MODE_OBS_ONLY 1  /* ObservationEncoder.compute_observations + concatenate (first pass of reset()) */

/* flags for dex_post_physics */
#define DEX_FLAG_FIRST_REWARD_CALL 1u  /* RewardCalculator prev-state is None (reward_calculator.py:149-153,189-190) */
#define DEX_FLAG_STALE_CONTACTS 2u     /* read monitored contact forces from state.contact_forces instead of gathering
                                          (reset() computes its first obs pass without refresh_tensors, dexhand_base.py:813) */
#define DEX_FLAG_NO_FUSED_RESET 4u     /* compute reset_buf but leave the reset to a later dex_reset_idx call */
#define DEX_FLAG_EARLY_SIM_LOADS 8u    /* the kernel directly ahead in the stream does not write the four sim tensors (it is
                                          dex_pre_physics of the same step, or any kernel that has completed): their rows are
                                          loaded before the programmatic-launch wait, overlapping that kernel.  Never set it when
                                          the previous launch is a dex_post_physics of the same envs (its fused reset writes
                                          dof_state / actor_root_state) or a physics kernel that signals early */

#define DEX_FLAG_TILE_KERNEL 16u       /* run the table-driven one-tile-per-CTA kernel (post_physics.cuh) even where the persistent
                                          warp-specialised kernel (post_physics_v2.cuh) covers the launch: A/B runs and the
                                          bit-for-bit cross-check of the two implementations (tests/test_cuda_v2.py) */
#define DEX_FLAG_SINGLE_PASS 64u      /* prefer the single-pass kernel (one 32-env tile per thread block, four blocks per SM, no ring)
                                         where it applies: BlindGrasping with more than 2 and at most 4 tiles per SM (16 384 envs on a
                                         B200).  It is 3-5 % slower than the persistent kernel when every input comes from HBM and
                                         20 % faster when the sim tensors are still L2-resident, i.e. when dex_post_physics runs right
                                         behind the physics step that wrote them.  Ignored elsewhere; same results bit for bit */
#define DEX_FLAG_TMA_STAGING 32u      /* persistent kernel with TMA producers: every gathered Isaac Gym row / state buffer is one
                                         2-D tensor copy (cp.async.bulk.tensor, SASS UTMALDG) instead of 4-byte asynchronous
                                         copies.  Bit-identical results; MEASURED SLOWER on B200 (20.98 vs 18.87 us per launch at
                                         16 384 envs: the TMA unit serves these 28..64-byte box rows at ~1 row per 6.5 clk per
                                         SM, profiles/r02_variants.md), kept selectable so the measurement can be repeated.
                                         BlindGrasping, N % 32 == 0, num_bodies % 4 == 0, num_actors % 2 == 0; ignored otherwise */
#define DEX_FLAG_PAIR_GEOMETRY 128u   /* persistent kernel with TWO tiles per pass: one thread block of 16 consumer + 8 producer
                                         warps per SM, the two groups of 8 consumer warps run the per-tile program in lockstep
                                         on two tiles (all barriers shared).  Bit-identical results; MEASURED SLOWER on B200
                                         (18.9 vs 18.3 us per launch at 16 384 envs, 194 vs 183 us at 262 144: the per-SM rate of
                                         ~one tile per 3.3 us is not an instruction-fetch limit that lockstep groups could
                                         share, profiles/r02_variants.md), kept selectable so the measurement can be repeated.
                                         BlindGrasping with an even number of 32-env tiles; ignored otherwise and with
                                         DEX_FLAG_TMA_STAGING / DEX_FLAG_SINGLE_PASS where those apply */

/* observation keys (obs_dict entries of the reference, observation_encoder.py:613-756,
 * step_processor.py:73, blind_grasping_task.py:569-652,702-746,895-899) */
enum DexObsKey {
  DEX_K_BASE_DOF_POS = 0, DEX_K_BASE_DOF_VEL, DEX_K_ACTIVE_FINGER_DOF_POS, DEX_K_ACTIVE_FINGER_DOF_VEL,
  DEX_K_ALL_FINGER_DOF_POS, DEX_K_ALL_FINGER_DOF_VEL, DEX_K_HAND_POSE, DEX_K_HAND_POSE_ARR_ALIGNED,
  DEX_K_CONTACT_FORCES, DEX_K_PREV_ACTIONS, DEX_K_ACTIVE_PREV_TARGETS, DEX_K_BASE_DOF_TARGET,
  DEX_K_ACTIVE_FINGER_DOF_TARGET, DEX_K_ALL_FINGER_DOF_TARGET, DEX_K_CONTACT_FORCE_MAGNITUDE,
  DEX_K_CONTACT_BINARY, DEX_K_CONTACT_DURATION, DEX_K_FINGERTIP_POSES_WORLD, DEX_K_FINGERTIP_POSES_HAND,
  DEX_K_FINGERPAD_POSES_WORLD, DEX_K_FINGERPAD_POSES_HAND, DEX_K_EPISODE_TIME, DEX_K_ACTIVE_RULE_TARGETS,
  /* BlindGrasping task observations */
  DEX_K_OBJECT_POS, DEX_K_OBJECT_VEL, DEX_K_FINGER_TO_OBJECT_DISTANCES, DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE,
  DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF, DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF, DEX_K_HAND_TO_OBJECT_DISTANCE,
  DEX_K_FINGERPAD_DISTANCES, DEX_K_FIRST_THREE_FINGERPAD_CENTROID, DEX_K_THUMB_CONTACT,
  DEX_K_OTHER_FINGERS_CONTACT, DEX_K_GRASP_STATE, DEX_K_GRASP_DURATION, DEX_K_CURRENT_STAGE,
  DEX_K_TIME_IN_STAGE, DEX_K_STAGE_PROGRESS,
  DEX_NUM_OBS_KEYS
};

/* reward terms, in the reference's dict order: common (reward_calculator.py:66-221) then task
 * (blind_grasping_task.py:980-1053); termination rewards (termination_manager.py:280-321) last */
enum DexRewardTerm {
  DEX_R_ALIVE = 0, DEX_R_HEIGHT_SAFETY, DEX_R_FINGER_VELOCITY, DEX_R_HAND_VELOCITY, DEX_R_HAND_ANGULAR_VELOCITY,
  DEX_R_JOINT_LIMIT, DEX_R_FINGER_ACCELERATION, DEX_R_HAND_ACCELERATION, DEX_R_HAND_ANGULAR_ACCELERATION,
  DEX_R_CONTACT_STABILITY,
  DEX_R_S1_HEIGHT_ALIGNMENT, DEX_R_S1_CENTROID_POSITIONING, DEX_R_S1_OBJECT_STABILITY,
  DEX_R_S1_FINGER_HEIGHT_CONSISTENCY, DEX_R_S1_THUMB_ROTATION,
  DEX_R_S2_THUMB_CONTACT, DEX_R_S2_OTHER_FINGERS_CONTACT, DEX_R_S2_GRASP_ACHIEVEMENT, DEX_R_S2_FINGERPAD_PROXIMITY,
  DEX_R_S2_BASE_STABILITY,
  DEX_R_S3_OBJECT_HEIGHT, DEX_R_S3_GRASP_MAINTENANCE, DEX_R_S3_GRASP_DURATION,
  DEX_R_S1_COMPLETION, DEX_R_S2_COMPLETION, DEX_R_PENETRATION_PENALTY,
  DEX_NUM_REWARD_TERMS
};
#define DEX_NUM_COMMON_TERMS 10

/* termination criteria (termination_manager.py:160-227; dict order: builtin first, then task) */
enum DexCriterion {
  DEX_C_GRASP_LIFT_SUCCESS = 0,  /* success */
  DEX_C_HITTING_GROUND,          /* failures ... */
  DEX_C_BOX_TOO_FAR, DEX_C_STAGE1_PREGRASP_FAILED, DEX_C_STAGE2_CONTACT_FAILED, DEX_C_STAGE3_GRASP_LOST,
  DEX_NUM_CRITERIA
};

/* int32 counters accumulated by dex_post_physics (zeroed by dex_step_epilogue after it consumed them) */
enum DexCounter {
  DEX_CNT_SUCCESS = 0, DEX_CNT_FAILURE, DEX_CNT_TIMEOUT, DEX_CNT_RESET,
  DEX_CNT_CRIT0, /* .. DEX_CNT_CRIT0 + DEX_NUM_CRITERIA - 1: raw criterion hit counts */
  DEX_CNT_EPISODE_LENGTH = 10, /* sum of episode_step_count over the envs that reset this step */
  DEX_NUM_COUNTERS = 16
};

/* floats written by dex_step_epilogue */
enum DexStat {
  DEX_STAT_SUCCESS_RATE = 0, DEX_STAT_FAILURE_RATE, DEX_STAT_TIMEOUT_RATE,
  DEX_STAT_CRIT0, /* .. + DEX_NUM_CRITERIA - 1: mean of the raw criterion */
  DEX_NUM_STATS = 16
};

/* ------------------------------------------------------------------------------------------------
 * Parameter block: everything the reference reads from cfg / derives once at init.  Filled on the
 * host (double arithmetic as Python does, then narrowed to float exactly where torch narrows).
 * ---------------------------------------------------------------------------------------------- */
typedef struct DexParams {
  int32_t abi_version;            /* must be DEX_ABI_VERSION */
  int32_t task;                   /* DEX_TASK_* */
  /* --- layout of the sim buffers (hand_initializer.py:439-560, blind_grasping_task.py:388-419) --- */
  int32_t num_bodies;             /* NB: rigid bodies per env; one env of rigid_body_state is NB * rigid_row_floats floats */
  int32_t num_contact_bodies;     /* bodies per env in net_contact_force (its row pitch; contact_body[] and box_body index it).
                                     Equal to num_bodies for the Isaac Gym tensors; smaller for a compact host-staging layout
                                     that ships only the monitored rows (DexStepEnv.alloc_host_buffers(compact=True)) */
  int32_t num_actors;             /* NA: actors per env */
  int32_t hand_body;              /* right_hand_base */
  int32_t tip_body[DEX_NUM_FINGERS];      /* r_f_link{1..5}_tip */
  int32_t pad_body[DEX_NUM_FINGERS];      /* r_f_link{1..5}_pad */
  int32_t contact_body[DEX_NUM_FINGERS];  /* contact_force_bodies */
  int32_t box_body;               /* -1 without box */
  int32_t box_actor;              /* -1 without box */
  /* --- action pipeline (action_processor.py:181-247,380-503) --- */
  int32_t num_actions;            /* NACT = 18 / 12 / 6 */
  int32_t control_mode;           /* DEX_CONTROL_* */
  int32_t policy_controls_base;
  int32_t policy_controls_fingers;
  int32_t filter_velocity_clamp;  /* post_action_filters contains "velocity_clamp" */
  int32_t filter_position_clamp;
  int32_t dof_to_active[DEX_NUM_DOF];   /* per DOF: index into the 18 active targets, -1 = forced 0 (r_f_joint3_1) */
  float dof_coupling_scale[DEX_NUM_DOF];
  int32_t inverse_src_dof[DEX_NUM_ACTIVE]; /* per active target: DOF whose position resets it (last DOF of the group) */
  float inverse_scale[DEX_NUM_ACTIVE];
  float active_lower[DEX_NUM_ACTIVE];
  float active_upper[DEX_NUM_ACTIVE];
  float max_deltas[DEX_NUM_ACTIVE];
  /* --- observations --- */
  int32_t active_finger_dof[DEX_NUM_ACTIVE_FINGER]; /* observation_encoder.py:496-574 */
  int32_t obs_dim;
  int32_t obs_offset[DEX_NUM_OBS_KEYS];   /* column of the key in obs_buf, -1 = not a policy key */
  int32_t dict_dim;
  int32_t dict_offset[DEX_NUM_OBS_KEYS];  /* column of the key in the optional obs_dict buffer */
  float control_dt;
  float contact_binary_threshold;
  /* --- termination --- */
  int32_t max_episode_length;
  int32_t has_height_safety;
  float handbase_threshold;
  float fingertip_threshold;
  int32_t criterion_active[DEX_NUM_CRITERIA];
  float success_reward;           /* +termination_success */
  float failure_reward;           /* -termination_failure_penalty */
  float timeout_reward;           /* -termination_timeout_penalty */
  int32_t max_consecutive_successes;
  /* --- rewards --- */
  float reward_weight[DEX_NUM_REWARD_TERMS];
  int32_t component_row[DEX_NUM_REWARD_TERMS]; /* row of the raw term in reward_components (weighted = row+1), -1 if weight == 0 */
  int32_t total_row;                           /* row of "total" */
  int32_t termination_row;                     /* first of 6 rows: success, success_w, failure_penalty, .._w, timeout_penalty, .._w */
  int32_t num_component_rows;
  int32_t need_common_rewards;                 /* any common weight != 0 */
  float dof_lower[DEX_NUM_DOF];
  float dof_upper[DEX_NUM_DOF];
  /* --- BlindGrasping (blind_grasping_task.py:114-205) --- */
  float box_size, box_z, box_xy_range;
  float success_height_threshold;
  float contact_duration_threshold_s;
  int32_t contact_duration_threshold_steps;
  int32_t min_fingers_for_grasp;
  float max_box_distance;
  float stage1_duration, stage2_duration, stage2_contact_success_threshold;
  float hand_translation_range, hand_rotation_range, thumb_rotation_range, other_finger_range;
  float height_alignment_decay, centroid_positioning_decay, object_stability_decay;
  float height_consistency_decay, fingerpad_proximity_decay, base_stability_decay;
  float height_tolerance, centroid_tolerance, position_drift_tolerance, velocity_tolerance;
  float proximity_threshold;       /* sqrt(3)*size/2*1.2 */
  float penetration_threshold;     /* size/2*geometricPenetrationFactor */
  float min_reward_distance;       /* size/2*proximityMinDistanceFactor */
  float penetration_depth_scale;
  float lift_height_range;         /* success_height_threshold - box_z (double, then narrowed) */
  float half_pi;                   /* (float)(pi/2) */
  float pi;                        /* (float)pi */
  /* --- packed host staging (no counterpart in the reference; 0 or 13 = the Isaac Gym tensor) --- */
  int32_t rigid_row_floats;        /* floats between consecutive body rows of rigid_body_state: body b starts at float
                                      b * rigid_row_floats of its env.  13 (or 0) for Isaac Gym's [N,NB,13].  7..12 for a
                                      host-staging buffer that ships poses only (pos3 + quat4 per tip / pad row,
                                      DexStepEnv.alloc_host_buffers(compact="packed")): the path reads 7 floats of every tip /
                                      pad row and 13 of the hand row (7 when no common reward term is weighted), so with 7 the
                                      hand takes two consecutive slots.  hand_body * rigid_row_floats + 13 must stay inside
                                      the env's NB * rigid_row_floats floats (checked).  With a pitch other than 13
                                      DEX_FLAG_TMA_STAGING is ignored (the tensor maps assume Isaac Gym's rows). */
} DexParams;

/* Isaac Gym tensor-API buffers, zero-copy (tensor_manager.py:173-328). */
typedef struct DexSimViews {
  float* dof_state;               /* [N,26,2]  read; rows of reset envs are written (reset_manager.py:121-124) */
  const float* rigid_body_state;  /* [N,NB,13] read only */
  const float* net_contact_force; /* [N,NB,3]  read only */
  float* actor_root_state;        /* [N,NA,13] box row read; written on reset (blind_grasping_task.py:475-490) */
} DexSimViews;

/* Persistent path state.  [K][N] = transposed/SoA.  Pointers marked (BG) may be NULL
 * for DEX_TASK_BASE; (common) may be NULL when need_common_rewards == 0. */
typedef struct DexStateBuffers {
  float* active_prev_targets;     /* [18][N]  ActionProcessor.active_prev_targets */
  float* active_rule_targets;     /* [18][N]  obs_dict["active_rule_targets"] */
  float* full_dof_targets;        /* [N,26]   AoS: handed to gym.set_dof_position_target_tensor */
  float* prev_actions;            /* [NACT][N] ObservationEncoder.prev_actions */
  float* prev_dof_pos;            /* [26][N] */
  float* contact_duration_steps;  /* [5][N] */
  float* prev_contact_binary;     /* [5][N] */
  float* contact_forces;          /* [15][N] gathered monitored forces (tensor_manager.contact_forces); may be NULL */
  int64_t* episode_step_count;    /* [N] */
  float* prev_finger_dof_vel;     /* [20][N] (common) */
  float* prev_hand_vel;           /* [3][N]  (common) */
  float* prev_hand_ang_vel;       /* [3][N]  (common) */
  uint8_t* prev_contacts;         /* [5][N]  (common) */
  int64_t* success_duration_steps;        /* [N] (BG) */
  uint8_t* success_conditions_met;        /* [N] (BG) */
  int64_t* current_stage;                 /* [N] (BG) */
  float* time_in_stage;                   /* [N] (BG) */
  float* stage_contact_duration;          /* [N] (BG) */
  uint8_t* just_transitioned_to_stage2;   /* [N] (BG) */
  uint8_t* just_transitioned_to_stage3;   /* [N] (BG) */
  float* initial_box_positions;           /* [3][N] (BG) */
  uint8_t* failure_tensors;               /* [3][N] (BG) stage1_pregrasp_failed, stage2_contact_failed, stage3_grasp_lost */
  uint64_t* rng_step;                     /* [1] optional device-resident Philox step counter: when non-NULL it replaces the
                                             by-value step_index (which a captured CUDA graph would freeze);
                                             dex_step_epilogue increments it */
} DexStateBuffers;

typedef struct DexOutBuffers {
  float* obs_buf;                 /* [N,obs_dim] */
  float* obs_dict;                /* [N,dict_dim] optional full obs_dict materialisation (NULL = skip) */
  float* rew_buf;                 /* [N] */
  uint8_t* reset_buf;             /* [N] bool */
  uint8_t* success;               /* [N] termination_info["success"] */
  uint8_t* failure;               /* [N] */
  uint8_t* timeout;               /* [N] */
  float* reward_components;       /* [num_component_rows][N] */
  uint32_t* reset_bits;           /* [ceil(N/32)] ballot words of reset_buf */
  int32_t* counters;              /* [DEX_NUM_COUNTERS] */
  /* Optional device-side RewardComponentObserver (rl/reward_observer.py:160-293); all NULL = off.
   * Fused into the post-physics pass so the per-step component arrays need not be re-read. */
  float* episode_sums;            /* [num_component_rows][N] running per-episode sum of each component */
  int32_t* episode_lengths;       /* [N] observer-side episode length (reward_observer.py:166) */
  double* stat_totals;            /* [S][4][2 + num_component_rows], classes {all, success, failure, timeout}:
                                     n_episodes, sum(episode_length), sum of each component's episode sum.
                                     S = stat_slices blocks; the totals are the SUM OVER THE S BLOCKS (the host adds them
                                     when it reads).  Summed, this vector is the payload of the single NCCL all-reduce
                                     (SURVEY 8e). */
  int32_t stat_slices;            /* S >= 1 (0 is read as 1).  Every finished episode adds 2 x (2 + rows) fp64 atomics; with
                                     one block they all land on the same ~80 addresses and serialise in L2 (measured:
                                     +4.3 us per launch at 16 384 envs with 1.5 % of the envs finishing per step).  The
                                     32-env tile t adds to block t % S, which divides the contention by S */
  /* Optional policy-input stage: the input normalisation rl_games applies to obs_buf before the
   * network (cfg/train/BaseTaskPPO.yaml:44 normalize_input -> rl_games RunningMeanStd.forward with frozen statistics:
   * y = clamp((obs - running_mean.float()) / sqrt(running_var.float() + epsilon), -5, 5)), written from the same
   * shared-memory slab as obs_buf so the policy never re-reads obs_buf.  obs_norm == NULL = off.  The division is a
   * multiplication by the caller's fp32 reciprocal: within 2 ulp of the torch expression (tests: rtol 1e-6).
   * The running statistics themselves stay with the caller (they change only in training, from the raw obs_buf). */
  float* obs_norm;                /* [N,obs_dim] normalised + clipped observations */
  const float* norm_mean;         /* [obs_dim] running_mean.float() */
  const float* norm_scale;        /* [obs_dim] 1 / sqrt(running_var.float() + epsilon), computed by the caller in fp32 */
  float norm_clip;                /* 5.0 in rl_games; +inf = no clipping */
} DexOutBuffers;

typedef struct DexEpilogueBuffers {
  const uint32_t* reset_bits;     /* [ceil(N/32)] */
  int32_t* counters;              /* [DEX_NUM_COUNTERS] consumed and zeroed */
  int32_t* env_ids;               /* [N] ascending ids of the reset envs (== torch.nonzero(reset_buf)) */
  int32_t* num_reset;             /* [1] */
  float* stats;                   /* [DEX_NUM_STATS] */
  int32_t* consecutive_successes; /* [1] persistent */
  uint64_t* rng_step;             /* [1] optional: incremented once per call (see DexStateBuffers.rng_step) */
  int64_t* totals;                /* [DEX_NUM_COUNTERS + 1] optional running totals since the caller last zeroed them:
                                     totals[k] += counters[k], totals[DEX_NUM_COUNTERS] += 1 (steps).  Together with
                                     DexOutBuffers.stat_totals this is the payload of the NCCL all-reduce */
  uint64_t* tile_records;         /* [4 * ceil(N/32)] (three planes of ceil(N/32) words are used) zero-initialised once by the caller; required only when the epilogue is
                                     fused into dex_post_physics (and then rng_step is required too).  Every 32-env tile
                                     publishes its reset ballot word and its counter contributions as three self-validating
                                     8-byte records (top 16 bits = step stamp derived from *rng_step); one extra thread block
                                     behind the tiles polls them - no fence, no atomic, no counter contention on the step path.
                                     SCHEDULING: the polling block has the highest block index of the launch.  The persistent
                                     and single-pass kernels launch at most (SMs x occupancy) blocks, polling block included,
                                     so every tile's block is resident by construction; the tile kernel relies on blocks being
                                     dispatched in index order (every tile's block has started before the polling block does),
                                     which holds on every CUDA GPU to date but is not a documented guarantee.  All waits are
                                     therefore BOUNDED: a record that does not appear within ~2 s traps the kernel and the call
                                     that next synchronises the stream returns a CUDA error (never a hang); pass ep == NULL to
                                     dex_post_physics and call dex_step_epilogue for a configuration without that dependence.
                                     STAMPS: a record is current when its stamp == (*rng_step + 1) & 0xffff, so *rng_step must
                                     only move forward by dex_step_epilogue / the fused epilogue; a caller that rewinds or
                                     restores *rng_step, or shares tile_records between pipelines, must re-zero tile_records */
} DexEpilogueBuffers;

const char* dex_strerror(int code);
int dex_abi_version(void);
/* sizeof(DexParams) as compiled, so a binding can verify its struct mirror */
int dex_sizeof_params(void);
/* number of dex_post_physics calls of this process that launched the persistent warp-specialised kernel
 * (csrc/post_physics_v2.cuh) rather than the tile kernel: lets a test assert which kernel the lean path runs */
int dex_debug_v2_launches(void);
/* which kernel the last dex_post_physics call of this process launched: 0 = table-driven tile kernel (post_physics_kernel),
 * 1 = persistent warp-specialised kernel (post_physics_v2_kernel, LDGSTS producers), 2 = single-pass kernel
 * (post_physics_v2s_kernel), 3 = persistent kernel with TMA producers (tensor copies), 4 = persistent kernel with two tiles
 * per pass (DEX_FLAG_PAIR_GEOMETRY) */
int dex_debug_last_post_kernel(void);
/* measurement aid (bench.py): launches an EMPTY kernel with dex_pre_physics' grid, block and programmatic-launch
 * attribute for n_envs - the per-launch floor of a kernel in a stream / graph of step kernels on this GPU */
int dex_debug_launch_floor(int32_t n_envs, void* stream);

/* Pre-physics: actions -> 26 DOF position targets.  `actions` [N,NACT] row-major.  Optional debug
 * outputs (may be NULL): active_raw_targets [18][N] (ActionProcessor.active_raw_targets). */
int dex_pre_physics(const DexParams* params, const DexStateBuffers* state, const float* actions,
                    float* active_raw_targets, int32_t n_envs, void* stream);

/* Post-physics: one pass per env over the sim buffers; see header comment.  `uniforms` [N,29]
 * (row-major) are the reset randoms in [0,1); if NULL they are generated in-kernel (Philox4x32-10
 * keyed by (seed, step_index, env)).  `epilogue` (may be NULL): when given and mode == DEX_MODE_FULL,
 * the work of dex_step_epilogue is done by one extra thread block of the same launch (no second
 * kernel); the caller must then NOT call dex_step_epilogue for this step. */
int dex_post_physics(const DexParams* params, const DexSimViews* sim, const DexStateBuffers* state,
                     const DexOutBuffers* out, const DexEpilogueBuffers* epilogue, int32_t n_envs, int32_t mode,
                     uint32_t flags, const float* uniforms, uint64_t seed, uint64_t step_index, void* stream);

/* dex_post_physics with the fused step epilogue, for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, same results and the same kernels as dex_post_physics called with
 * `epilogue != NULL`, with one difference: the tile-record scratch the fused epilogue polls (DexEpilogueBuffers.tile_records,
 * otherwise allocated and zero-initialised by the caller) is a file-scope `static __device__` array inside the library,
 * 4 * DEX_NOALLOC_MAX_ENVS / 32 words = 256 KB, resident on every GPU that loads the library until the process exits.
 * `epilogue->tile_records` is ignored (may be NULL); `epilogue` itself and mode == DEX_MODE_FULL are required.
 *   - n_envs > DEX_NOALLOC_MAX_ENVS returns DEX_ERR_CAPACITY before anything is launched.
 *   - The call allocates nothing: no cudaMalloc / cudaFree / allocating library routine, on the first call or later (the
 *     array's device address is looked up once per device with cudaGetSymbolAddress).
 *   - ALL calls of a process on one GPU share the array and it keeps the last call's contents.  Each call therefore first
 *     launches a small kernel on `stream` that overwrites the records it will use with a stamp that cannot be this step's
 *     (so records of an earlier call - of this or of any other pipeline - never pass as current); two launches are issued.
 *   - Calls that OVERLAP on the GPU (different streams, or graphs replayed concurrently) race on the array: serialise them. */
#define DEX_NOALLOC_MAX_ENVS 262144
int dex_post_physics_noalloc(const DexParams* params, const DexSimViews* sim, const DexStateBuffers* state,
                             const DexOutBuffers* out, const DexEpilogueBuffers* epilogue, int32_t n_envs, int32_t mode,
                             uint32_t flags, const float* uniforms, uint64_t seed, uint64_t step_index, void* stream);

/* Fixed-capacity actor-index lists for Isaac Gym's indexed setters (replaces the host-sized index tensors of
 * components/reset/reset_manager.py:139-168 and components/physics/physics_manager.py:121-160).  From the epilogue's
 * env_ids / num_reset (device resident) it writes, without any host read:
 *   dof_actor_ids  [capacity]               global id (env * num_actors + dof_actor) of the hand actor of the i-th reset env
 *   root_actor_ids [capacity * num_actors]  global ids of ALL actors of the i-th reset env (reset_manager.py:139-156)
 * Entries past num_reset are PADDED with the first reset env (its rows hold the state the reset just wrote: applying them
 * twice changes nothing) or, when no env was reset, with env 0 (whose rows hold its current simulation state).  The
 * caller then always issues set_actor_root_state_tensor_indexed(..., capacity * num_actors) and
 * set_dof_state_tensor_indexed(..., capacity): constant launch arguments, so the whole control step - setters included -
 * can live in one CUDA graph.  capacity >= n_envs can never overflow; with a smaller capacity *overflow (optional, device,
 * int32, accumulates) counts the resets that did not fit and were therefore NOT applied in that step. */
int dex_reset_actor_indices(const DexEpilogueBuffers* epilogue, int32_t capacity, int32_t num_actors, int32_t dof_actor,
                            int32_t* root_actor_ids, int32_t* dof_actor_ids, int32_t* overflow, void* stream);

/* Policy-input stage, TRAINING half (rl_games RunningMeanStd.forward with self.training, the model the reference
 * trains with: cfg/train/BaseTaskPPO.yaml:44 normalize_input).  In this order, on `stream`, without a host read:
 *   1. batch moments of obs_buf over the n_envs rows (mean, unbiased variance, rounded to fp32 as rl_games' input.mean(0) /
 *      input.var(0) are),
 *   2. Chan merge into the float64 running statistics: running_mean / running_var [obs_dim], count [1] (rl_games initialises
 *      them to 0 / 1 / 1), and norm_mean = (float)running_mean, norm_scale = 1 / sqrtf((float)running_var + epsilon) - the two
 *      vectors DexOutBuffers.norm_mean / norm_scale point at, so the fused frozen-statistics stage of the NEXT dex_post_physics
 *      calls uses the new statistics,
 *   3. obs_norm (optional, [n_envs, obs_dim]) = clamp((obs_buf - norm_mean) * norm_scale, -clip, clip) with the UPDATED statistics,
 *      which is what the network sees in training mode.
 * A reduction over all envs sits between obs_buf and its normalised form, so this half is three small launches behind
 * dex_post_physics rather than part of it; evaluation / play use the fused stage alone.  scratch: [2 * obs_dim] doubles. */
int dex_policy_input_update(const float* obs_buf, int32_t n_envs, int32_t obs_dim, double* running_mean, double* running_var,
                            double* count, float* norm_mean, float* norm_scale, float* obs_norm, float epsilon, float clip,
                            double* scratch, void* stream);

/* Masked reset of envs with reset_mask[env] != 0 (NULL = all envs). */
int dex_reset_idx(const DexParams* params, const DexSimViews* sim, const DexStateBuffers* state,
                  const uint8_t* reset_mask, int32_t n_envs, const float* uniforms, uint64_t seed,
                  uint64_t step_index, void* stream);

/* Ordered compaction of the reset mask + termination statistics + consecutive-success counter. */
int dex_step_epilogue(const DexParams* params, const DexEpilogueBuffers* ep, int32_t n_envs, void* stream);

/* Stand-alone coupling helpers; active [18][N] SoA, full [N,26] AoS. */
int dex_apply_coupling(const DexParams* params, const float* active, float* full, int32_t n_envs, void* stream);
int dex_extract_active_targets(const DexParams* params, const float* full, float* active, int32_t n_envs,
                               void* stream);

/* Known-answer entry: p_hand = R(q)^-1 (p_world - hand_pos); all [n,3]/[n,4] row-major. */
int dex_point_in_hand_frame(const float* p_world, const float* hand_pos, const float* hand_quat, float* out,
                            int32_t n, void* stream);
/* q_out = conj(a) (x) b  and  q_out = a (x) conj([0,s,0,s]) helpers used by the same tests. */
int dex_quat_mul(const float* a, const float* b, float* out, int32_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DEXSTEP_H_ */
