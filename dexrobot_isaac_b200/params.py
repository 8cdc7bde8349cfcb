"""Fills the ``DexParams`` block of the C ABI from the task config dict + the body/DOF layout.

Everything the reference derives once at init lives here (SURVEY §8a "Derived constants"):
coupling tables (action_processor.py:437-503), active limits (:380-413), max deltas (:416-434),
active_finger_dof_indices (observation_encoder.py:496-574), observation column layout
(:783-829), reward rows (reward_calculator.py:223-275), termination parameters
(termination_manager.py:27-77) and the BlindGrasping thresholds (blind_grasping_task.py:114-205).
Python-double arithmetic first, narrowing to fp32 exactly where torch narrows a Python scalar.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Sequence

import numpy as np

from . import _lib as L

NUM_DOF = 26
NUM_BASE = 6
NUM_ACTIVE_FINGER = 12
NUM_FINGER_DOF = 20
NUM_ACTIVE = 18
NUM_FINGERS = 5

# constants.py:14-37
BASE_JOINT_NAMES = ["ARTx", "ARTy", "ARTz", "ARRx", "ARRy", "ARRz"]
FINGER_JOINT_NAMES = [f"r_f_joint{f}_{j}" for f in range(1, 6) for j in range(1, 5)]
DOF_NAMES = BASE_JOINT_NAMES + FINGER_JOINT_NAMES
FINGERTIP_BODY_NAMES = [f"r_f_link{f}_tip" for f in range(1, 6)]   # constants.py:40-46
FINGERPAD_BODY_NAMES = [f"r_f_link{f}_pad" for f in range(1, 6)]   # constants.py:48-54

# constants.py:71-88 — action-order finger control -> coupled joints (joint, scale)
FINGER_COUPLING_MAP: Dict[int, List] = {
    0: [("r_f_joint1_1", 1.0)],
    1: [("r_f_joint1_2", 1.0)],
    2: [("r_f_joint1_3", 1.0), ("r_f_joint1_4", 1.0)],
    3: [("r_f_joint2_1", 1.0), ("r_f_joint4_1", 1.0), ("r_f_joint5_1", 2.0)],
    4: [("r_f_joint2_2", 1.0)],
    5: [("r_f_joint2_3", 1.0), ("r_f_joint2_4", 1.0)],
    6: [("r_f_joint3_2", 1.0)],
    7: [("r_f_joint3_3", 1.0), ("r_f_joint3_4", 1.0)],
    8: [("r_f_joint4_2", 1.0)],
    9: [("r_f_joint4_3", 1.0), ("r_f_joint4_4", 1.0)],
    10: [("r_f_joint5_2", 1.0)],
    11: [("r_f_joint5_3", 1.0), ("r_f_joint5_4", 1.0)],
}
SPREAD_FIXED_JOINT = "r_f_joint3_1"  # action_processor.py:612: forced to 0

# hand_initializer.py:20-34 HardwareMapping, enum order: control name -> joints (first = primary)
HARDWARE_MAPPING = [
    ("th_dip", ["r_f_joint1_3", "r_f_joint1_4"]), ("th_mcp", ["r_f_joint1_2"]), ("th_rot", ["r_f_joint1_1"]),
    ("ff_spr", ["r_f_joint2_1", "r_f_joint4_1", "r_f_joint5_1"]), ("ff_dip", ["r_f_joint2_3", "r_f_joint2_4"]),
    ("ff_mcp", ["r_f_joint2_2"]), ("mf_dip", ["r_f_joint3_3", "r_f_joint3_4"]), ("mf_mcp", ["r_f_joint3_2"]),
    ("rf_dip", ["r_f_joint4_3", "r_f_joint4_4"]), ("rf_mcp", ["r_f_joint4_2"]),
    ("lf_dip", ["r_f_joint5_3", "r_f_joint5_4"]), ("lf_mcp", ["r_f_joint5_2"]),
]

# obs_dict key widths (observation_encoder.py:613-756, blind_grasping_task.py:569-652); prev_actions = NACT
OBS_KEY_DIMS = {
    "base_dof_pos": 6, "base_dof_vel": 6, "active_finger_dof_pos": 12, "active_finger_dof_vel": 12,
    "all_finger_dof_pos": 20, "all_finger_dof_vel": 20, "hand_pose": 7, "hand_pose_arr_aligned": 7,
    "contact_forces": 15, "prev_actions": None, "active_prev_targets": 18, "base_dof_target": 6,
    "active_finger_dof_target": 12, "all_finger_dof_target": 20, "contact_force_magnitude": 5,
    "contact_binary": 5, "contact_duration": 5, "fingertip_poses_world": 35, "fingertip_poses_hand": 35,
    "fingerpad_poses_world": 35, "fingerpad_poses_hand": 35, "episode_time": 1, "active_rule_targets": 18,
    "object_pos": 3, "object_vel": 3, "finger_to_object_distances": 5, "avg_finger_to_object_distance": 1,
    "finger_to_object_height_diff": 5, "avg_finger_to_object_height_diff": 1, "hand_to_object_distance": 1,
    "fingerpad_distances": 10, "first_three_fingerpad_centroid": 3, "thumb_contact": 1,
    "other_fingers_contact": 1, "grasp_state": 1, "grasp_duration": 1, "current_stage": 1, "time_in_stage": 1,
    "stage_progress": 1,
}
# keys whose reference tensors have shape (N,) rather than (N, 1) (blind_grasping_task.py:702-724)
OBS_KEYS_1D = ("avg_finger_to_object_distance", "avg_finger_to_object_height_diff", "hand_to_object_distance")
NUM_DEFAULT_KEYS = L.CONSTS["DEX_K_OBJECT_POS"]  # keys below this enum value exist for every task

FAILURE_CRITERIA = ["hitting_ground", "box_too_far", "stage1_pregrasp_failed", "stage2_contact_failed",
                    "stage3_grasp_lost"]
SUCCESS_CRITERIA = ["grasp_lift_success"]


@dataclass
class BodyLayout:
    """Rigid-body / actor indices of one env: what HandInitializer.initialize_rigid_body_indices
    (hand_initializer.py:439-560) and BlindGraspingTask (blind_grasping_task.py:388-419) discover
    from the asset.  The hand MJCF is absent from the reference checkout, so these are runtime
    parameters; ``default`` gives the layout the benchmarks assume."""

    num_bodies: int
    num_actors: int
    hand_body: int
    tip_bodies: Sequence[int]
    pad_bodies: Sequence[int]
    contact_bodies: Sequence[int]
    box_body: int = -1
    box_actor: int = -1
    body_names: Dict[int, str] = field(default_factory=dict)
    num_contact_bodies: int = 0  # bodies per env in net_contact_force (0 = num_bodies, the Isaac Gym tensors)
    rigid_row_floats: int = 13   # floats per body row of rigid_body_state (13 = Isaac Gym; 7 = packed host staging)

    @property
    def contact_rows(self) -> int:
        return self.num_contact_bodies or self.num_bodies

    def compact(self, packed: bool = False) -> "BodyLayout":
        """The host-staging layout that ships only the rows the path reads (DexStepEnv.alloc_host_buffers(compact=True)):
        rigid_body_state rows [hand, 5 tips, 5 pads], net_contact_force rows [5 monitored bodies (, box)],
        actor_root_state row [box] (or the hand actor when there is no box).

        ``packed``: the rigid rows additionally shrink to what the path reads of them - 7-float slots (DexParams.
        rigid_row_floats = 7): the 13-float hand row takes slots 0-1, the tip / pad poses (pos3 + quat4) slots 2-6 / 7-11,
        [N,12,7] = 336 B per env instead of 572."""
        has_box = self.box_body >= 0
        first = 2 if packed else 1
        return BodyLayout(num_bodies=first + 10, num_actors=1, hand_body=0,
                          tip_bodies=[first + f for f in range(NUM_FINGERS)],
                          pad_bodies=[first + 5 + f for f in range(NUM_FINGERS)], contact_bodies=list(range(NUM_FINGERS)),
                          box_body=5 if has_box else -1, box_actor=0 if has_box else -1,
                          num_contact_bodies=6 if has_box else 5, rigid_row_floats=7 if packed else 13)

    def compact_gather_indices(self):
        """-> (rigid rows, contact rows, actor rows) of THIS layout that make up the compact layout, in its order."""
        rigid = [self.hand_body] + list(self.tip_bodies) + list(self.pad_bodies)
        contact = list(self.contact_bodies) + ([self.box_body] if self.box_body >= 0 else [])
        actors = [self.box_actor if self.box_actor >= 0 else 0]
        return rigid, contact, actors

    @staticmethod
    def default(with_box: bool = False, nb_hand: int = 31) -> "BodyLayout":
        per = (nb_hand - 1) // NUM_FINGERS
        if per < 6:
            raise ValueError("nb_hand too small: need link1..4 + tip + pad per finger")
        names = {0: "right_hand_base"}
        for f in range(NUM_FINGERS):
            for j in range(4):
                names[1 + per * f + j] = f"r_f_link{f + 1}_{j + 1}"
            names[1 + per * f + 4] = f"r_f_link{f + 1}_tip"
            names[1 + per * f + 5] = f"r_f_link{f + 1}_pad"
        if with_box:
            names[nb_hand] = "box"
        return BodyLayout(
            num_bodies=nb_hand + (1 if with_box else 0), num_actors=2 if with_box else 1, hand_body=0,
            tip_bodies=[1 + per * f + 4 for f in range(NUM_FINGERS)],
            pad_bodies=[1 + per * f + 5 for f in range(NUM_FINGERS)],
            contact_bodies=[1 + per * f + 3 for f in range(NUM_FINGERS)],
            box_body=nb_hand if with_box else -1, box_actor=1 if with_box else -1, body_names=names)

    @staticmethod
    def from_synth(layout) -> "BodyLayout":
        """Adapter for the test generator's layout object (oracle/synth.py) — duck-typed, no import."""
        return BodyLayout(num_bodies=layout.nb, num_actors=layout.na, hand_body=layout.hand,
                          tip_bodies=list(layout.tips), pad_bodies=list(layout.pads),
                          contact_bodies=list(layout.contacts), box_body=layout.box_body,
                          box_actor=layout.box_actor, body_names=dict(layout.body_names()))


def f32(x) -> float:
    """Narrow a Python double the way torch narrows a scalar that meets an fp32 tensor."""
    return float(np.float32(x))


def coupling_tables():
    """-> dof_to_active[26], dof_scale[26], inverse_src_dof[18], inverse_scale[18]
    (action_processor.py:437-503; duplicates in the inverse scatter: last DOF in index order wins)."""
    name_to_idx = {n: i for i, n in enumerate(DOF_NAMES)}
    dof_to_active = [-1] * NUM_DOF
    dof_scale = [1.0] * NUM_DOF
    for d in range(NUM_BASE):
        dof_to_active[d] = d
    for f, joints in FINGER_COUPLING_MAP.items():
        for joint, scale in joints:
            d = name_to_idx[joint]
            dof_to_active[d] = NUM_BASE + f
            dof_scale[d] = scale
    dof_to_active[name_to_idx[SPREAD_FIXED_JOINT]] = -1
    inv_src = list(range(NUM_BASE)) + [0] * NUM_ACTIVE_FINGER
    inv_scale = [1.0] * NUM_ACTIVE
    for d in range(NUM_BASE, NUM_DOF):  # ascending: the last DOF of a group wins
        a = dof_to_active[d]
        if a >= 0:
            inv_src[a] = d
            inv_scale[a] = dof_scale[d]
    return dof_to_active, dof_scale, inv_src, inv_scale


def first_joint_of_group() -> List[int]:
    name_to_idx = {n: i for i, n in enumerate(DOF_NAMES)}
    return [name_to_idx[FINGER_COUPLING_MAP[f][0][0]] for f in range(NUM_ACTIVE_FINGER)]


def active_finger_dof_indices() -> List[int]:
    name_to_idx = {n: i for i, n in enumerate(DOF_NAMES)}
    return [name_to_idx[joints[0]] for _, joints in HARDWARE_MAPPING]


@dataclass
class PipelineSpec:
    """Host-side description derived alongside DexParams (names <-> rows/columns)."""

    task: str
    is_blind_grasping: bool
    num_actions: int
    control_dt: float
    obs_keys: List[str]
    obs_dim: int
    obs_slices: Dict[str, tuple]        # key -> (start, end) in obs_buf == component_slice_indices
    dict_keys: List[str]
    dict_dim: int
    dict_slices: Dict[str, tuple]       # key -> (start, end) in the obs_dict buffer
    component_names: List[str]          # row -> reward_components key
    active_success: List[str]
    active_failure: List[str]
    reward_weights: Dict[str, float]
    max_episode_length: int


def make_params(cfg: dict, layout: BodyLayout, dof_lower: Sequence[float], dof_upper: Sequence[float],
                post_action_filters: Sequence[str] = ("velocity_clamp", "position_clamp"),
                physics_steps_per_control_step: int = 2):
    """-> (DexParams, PipelineSpec).  ``control_dt = physics_steps_per_control_step * sim.dt``: the
    reference measures 2 (one step + one reset step, dexhand_base.py:270-320, physics_manager.py:243-259)."""
    C = L.CONSTS
    tc, env_cfg, sim_cfg = cfg["task"], cfg["env"], cfg["sim"]
    task_name = tc.get("name", "BaseTask")
    is_bg = task_name == "BlindGrasping"
    if task_name not in ("BaseTask", "BlindGrasping"):
        raise ValueError(f"fused kernels cover BaseTask and BlindGrasping, not '{task_name}'")
    P = L.DexParams()
    P.abi_version = C["DEX_ABI_VERSION"]
    P.task = C["DEX_TASK_BLIND_GRASPING"] if is_bg else C["DEX_TASK_BASE"]
    P.num_bodies, P.num_actors, P.hand_body = layout.num_bodies, layout.num_actors, layout.hand_body
    P.num_contact_bodies = layout.contact_rows
    P.rigid_row_floats = layout.rigid_row_floats
    for f in range(NUM_FINGERS):
        P.tip_body[f], P.pad_body[f], P.contact_body[f] = (layout.tip_bodies[f], layout.pad_bodies[f],
                                                           layout.contact_bodies[f])
    P.box_body, P.box_actor = (layout.box_body, layout.box_actor) if is_bg else (-1, -1)

    # ---- action pipeline ----
    mode = tc["controlMode"]
    if mode not in ("position", "position_delta"):
        raise ValueError(f"Invalid control mode: {mode}")  # action_processor.py:208-215
    ctrl_base, ctrl_fingers = bool(tc["policy_controls_hand_base"]), bool(tc["policy_controls_fingers"])
    nact = (NUM_BASE if ctrl_base else 0) + (NUM_ACTIVE_FINGER if ctrl_fingers else 0)
    if nact == 0:
        raise ValueError("policy must control the hand base, the fingers or both")
    P.num_actions = nact
    P.control_mode = C["DEX_CONTROL_POSITION"] if mode == "position" else C["DEX_CONTROL_POSITION_DELTA"]
    P.policy_controls_base, P.policy_controls_fingers = int(ctrl_base), int(ctrl_fingers)
    unknown = set(post_action_filters) - {"velocity_clamp", "position_clamp"}
    if unknown:
        raise ValueError(f"unknown post action filters: {sorted(unknown)}")
    P.filter_velocity_clamp = int("velocity_clamp" in post_action_filters)
    P.filter_position_clamp = int("position_clamp" in post_action_filters)
    d2a, dscale, inv_src, inv_scale = coupling_tables()
    for d in range(NUM_DOF):
        P.dof_to_active[d], P.dof_coupling_scale[d] = d2a[d], dscale[d]
        P.dof_lower[d], P.dof_upper[d] = f32(dof_lower[d]), f32(dof_upper[d])
    first = first_joint_of_group()
    control_dt = float(sim_cfg["dt"]) * physics_steps_per_control_step
    lin, ang, fin = (tc["max_base_linear_velocity"], tc["max_base_angular_velocity"], tc["max_finger_joint_velocity"])
    for a in range(NUM_ACTIVE):
        src = a if a < NUM_BASE else first[a - NUM_BASE]
        P.active_lower[a], P.active_upper[a] = f32(dof_lower[src]), f32(dof_upper[src])
        P.inverse_src_dof[a], P.inverse_scale[a] = inv_src[a], inv_scale[a]
        P.max_deltas[a] = f32(control_dt * (lin if a < 3 else ang if a < 6 else fin))
    P.control_dt = f32(control_dt)
    for j, d in enumerate(active_finger_dof_indices()):
        P.active_finger_dof[j] = d

    # ---- observation layout ----
    key_id = {name: i for i, name in enumerate(L.OBS_KEY_NAMES)}
    dims = dict(OBS_KEY_DIMS)
    dims["prev_actions"] = nact
    obs_keys = list(tc["policy_observation_keys"])
    for i in range(C["DEX_NUM_OBS_KEYS"]):
        P.obs_offset[i] = -1
        P.dict_offset[i] = 0
    col, obs_slices = 0, {}
    for k in obs_keys:
        if k not in key_id:
            raise ValueError(f"observation key '{k}' is not produced by the fused kernels")
        if not is_bg and key_id[k] >= NUM_DEFAULT_KEYS:
            raise ValueError(f"observation key '{k}' needs the BlindGrasping task")
        if k in obs_slices:
            raise ValueError(f"duplicate observation key '{k}'")
        P.obs_offset[key_id[k]] = col
        obs_slices[k] = (col, col + dims[k])
        col += dims[k]
    P.obs_dim = col
    dcol, dict_slices, dict_keys = 0, {}, []
    for i, k in enumerate(L.OBS_KEY_NAMES):
        if not is_bg and i >= NUM_DEFAULT_KEYS:
            break
        P.dict_offset[i] = dcol
        dict_slices[k] = (dcol, dcol + dims[k])
        dict_keys.append(k)
        dcol += dims[k]
    P.dict_dim = dcol
    P.contact_binary_threshold = f32(tc["contact_binary_threshold"])

    # ---- termination ----
    term = tc.get("termination", {}) or {}
    P.max_episode_length = int(env_cfg["episodeLength"])
    hs = term.get("height_safety")
    P.has_height_safety = int(hs is not None)
    if hs is not None:
        P.handbase_threshold, P.fingertip_threshold = f32(hs["handbase_threshold"]), f32(hs["fingertip_threshold"])
    act_s, act_f = list(term.get("active_success_criteria", []) or []), list(term.get("active_failure_criteria", []) or [])
    avail_s = SUCCESS_CRITERIA if is_bg else []
    avail_f = (["hitting_ground"] if hs is not None else []) + (FAILURE_CRITERIA[1:] if is_bg else [])
    for name in act_s:
        if name not in avail_s:  # termination_manager.py:98-118 validate_criteria
            raise RuntimeError(f"active success criterion '{name}' is not provided by task {task_name}")
    for name in act_f:
        if name not in avail_f:
            raise RuntimeError(f"active failure criterion '{name}' is not provided by task {task_name}")
    crit_names = SUCCESS_CRITERIA + FAILURE_CRITERIA
    for c, name in enumerate(crit_names):
        if name in SUCCESS_CRITERIA:
            on = name in avail_s and (not act_s or name in act_s)
        else:
            on = name in avail_f and (not act_f or name in act_f)
        P.criterion_active[c] = int(on)
    rw = tc["reward_weights"]
    P.success_reward = f32(rw["termination_success"])
    P.failure_reward = f32(-rw["termination_failure_penalty"])
    P.timeout_reward = f32(-rw["termination_timeout_penalty"])
    P.max_consecutive_successes = int(tc["max_consecutive_successes"])

    # ---- reward rows (dict order: common, task; weight != 0 only) ----
    names: List[str] = []
    n_terms = C["DEX_NUM_REWARD_TERMS"] if is_bg else C["DEX_NUM_COMMON_TERMS"]
    any_common = False
    for i in range(C["DEX_NUM_REWARD_TERMS"]):
        term_name = L.REWARD_TERM_NAMES[i]
        w = float(rw.get(term_name, 0.0)) if i < n_terms else 0.0
        P.reward_weight[i] = f32(w)
        if w != 0:
            P.component_row[i] = len(names)
            names += [term_name, f"{term_name}_weighted"]
            any_common = any_common or i < C["DEX_NUM_COMMON_TERMS"]
        else:
            P.component_row[i] = -1
    P.total_row = len(names)
    names.append("total")
    P.termination_row = len(names)
    for key in ("success", "failure_penalty", "timeout_penalty"):
        names += [f"termination_{key}", f"termination_{key}_weighted"]
    P.num_component_rows = len(names)
    P.need_common_rewards = int(any_common)

    # ---- BlindGrasping ----
    if is_bg:
        box = env_cfg["box"]
        size, box_z = box["size"], box["initial_position"]["z"]
        pp, rc, q = tc["penetration_prevention"], tc["reward_calculation"], tc["quality_thresholds"]
        P.box_size, P.box_z, P.box_xy_range = f32(size), f32(box_z), f32(box["initial_position"]["xy_range"])
        P.success_height_threshold = f32(tc["success_height_threshold"])
        P.contact_duration_threshold_s = f32(tc["contact_duration_threshold"])
        P.contact_duration_threshold_steps = int(tc["contact_duration_threshold"] / control_dt)
        P.min_fingers_for_grasp = int(tc["min_fingers_for_grasp"])
        P.max_box_distance = f32(tc["max_box_distance"])
        P.stage1_duration, P.stage2_duration = f32(tc["stage1_duration"]), f32(tc["stage2_duration"])
        P.stage2_contact_success_threshold = f32(tc["stage_evaluation"]["stage2_contact_success_threshold"])
        P.hand_translation_range, P.hand_rotation_range = f32(tc["hand_translation_range"]), f32(tc["hand_rotation_range"])
        P.thumb_rotation_range = f32(tc["finger_randomization"]["thumb_rotation_range"])
        P.other_finger_range = f32(tc["finger_randomization"]["other_finger_range"])
        P.height_alignment_decay = f32(rc["height_alignment_decay"])
        P.centroid_positioning_decay = f32(rc["centroid_positioning_decay"])
        P.object_stability_decay = f32(rc["object_stability_decay"])
        P.height_consistency_decay = f32(rc["first_three_height_consistency_decay"])
        P.fingerpad_proximity_decay = f32(rc["fingerpad_proximity_decay"])
        P.base_stability_decay = f32(rc["base_stability_decay"])
        P.height_tolerance, P.centroid_tolerance = f32(q["height_tolerance"]), f32(q["centroid_tolerance"])
        P.position_drift_tolerance, P.velocity_tolerance = f32(q["position_drift_tolerance"]), f32(q["velocity_tolerance"])
        P.proximity_threshold = f32(math.sqrt(3) * size / 2 * 1.2)
        P.penetration_threshold = f32(size / 2.0 * pp["geometricPenetrationFactor"])
        P.min_reward_distance = f32(size / 2.0 * pp["proximityMinDistanceFactor"])
        P.penetration_depth_scale = f32(pp["penetrationDepthScale"])
        P.lift_height_range = f32(tc["success_height_threshold"] - box_z)
    P.half_pi, P.pi = f32(math.pi / 2), f32(math.pi)

    spec = PipelineSpec(task=task_name, is_blind_grasping=is_bg, num_actions=nact, control_dt=control_dt,
                        obs_keys=obs_keys, obs_dim=col, obs_slices=obs_slices, dict_keys=dict_keys, dict_dim=dcol,
                        dict_slices=dict_slices, component_names=names,
                        active_success=[n for c, n in enumerate(crit_names) if n in SUCCESS_CRITERIA and P.criterion_active[c]],
                        active_failure=[n for c, n in enumerate(crit_names) if n in FAILURE_CRITERIA and P.criterion_active[c]],
                        reward_weights=dict(rw), max_episode_length=int(env_cfg["episodeLength"]))
    return P, spec
