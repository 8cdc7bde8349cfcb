"""BlindGraspingTask facade (reference: dexhand_env/tasks/blind_grasping_task.py).

Everything this task computes per control step in the reference - task observations (:549-653), finger-box contacts
(:773-833), the three-stage FSM (:835-978), the 16 reward terms (:980-1208), the success / failure criteria (:1210-1364)
and ``reset_task_state`` (:433-547) - runs inside ``dex_post_physics`` / ``dex_reset_idx``.  The hooks below keep the
reference's names and signatures and hand out the kernel's results of the CURRENT step (views of device buffers, valid
until the next launch); nothing is recomputed on the host.  The task states the reference registers with the observation
encoder (:226-257) alias the pipeline's state buffers, so ``task.task_states["current_stage"]`` is the kernel's own tensor.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from .. import _lib as L
from ..params import OBS_KEY_DIMS
from .task_interface import DexTask

_TASK_OBS_KEYS = [k for k in L.OBS_KEY_NAMES[L.CONSTS["DEX_K_OBJECT_POS"]:]]
_TASK_STATES = (("success_duration_steps", torch.long), ("success_conditions_met", torch.bool), ("current_stage", torch.long),
                ("time_in_stage", torch.float32), ("stage_contact_duration", torch.float32),
                ("just_transitioned_to_stage2", torch.bool), ("just_transitioned_to_stage3", torch.bool))
_TASK_FAILURES = ("stage1_pregrasp_failed", "stage2_contact_failed", "stage3_grasp_lost")


class BlindGraspingTask(DexTask):
    def __init__(self, sim, gym, device, num_envs, full_config):
        self.sim, self.gym, self.device, self.num_envs, self.cfg = sim, gym, device, num_envs, full_config
        self.parent_env = None
        env_config, task_config = full_config["env"], full_config["task"]
        box = env_config["box"]  # :137-146
        self.box_size, self.box_mass = box["size"], box["mass"]
        self.box_friction, self.box_restitution = box["friction"], box["restitution"]
        self.box_xy_range, self.box_z = box["initial_position"]["xy_range"], box["initial_position"]["z"]
        self.height_threshold = task_config["success_height_threshold"]  # :147-170
        self.contact_duration_threshold_seconds = task_config["contact_duration_threshold"]
        self.min_fingers_for_grasp = task_config["min_fingers_for_grasp"]
        self.max_box_distance = task_config["max_box_distance"]
        self.stage1_duration, self.stage2_duration = task_config["stage1_duration"], task_config["stage2_duration"]
        self.hand_translation_range = task_config["hand_translation_range"]
        self.hand_rotation_range = task_config["hand_rotation_range"]
        self.thumb_rotation_range = task_config["finger_randomization"]["thumb_rotation_range"]
        self.other_finger_range = task_config["finger_randomization"]["other_finger_range"]
        self.stage2_contact_success_threshold = task_config["stage_evaluation"]["stage2_contact_success_threshold"]
        rc, q, pp = task_config["reward_calculation"], task_config["quality_thresholds"], task_config["penetration_prevention"]
        self.height_alignment_decay = rc["height_alignment_decay"]  # :175-200
        self.centroid_positioning_decay = rc["centroid_positioning_decay"]
        self.object_stability_decay = rc["object_stability_decay"]
        self.first_three_height_consistency_decay = rc["first_three_height_consistency_decay"]
        self.fingerpad_proximity_decay, self.base_stability_decay = rc["fingerpad_proximity_decay"], rc["base_stability_decay"]
        self.geometric_penetration_factor = pp["geometricPenetrationFactor"]
        self.proximity_min_distance_factor = pp["proximityMinDistanceFactor"]
        self.penetration_depth_scale = pp["penetrationDepthScale"]
        self.height_tolerance, self.centroid_tolerance = q["height_tolerance"], q["centroid_tolerance"]
        self.position_drift_tolerance, self.velocity_tolerance = q["position_drift_tolerance"], q["velocity_tolerance"]
        self.contact_duration_threshold_steps = None  # set in finalize_setup (:259-298)
        self.box_local_actor_index = None
        self.box_local_rigid_body_index = None

    # ------------------------------------------------------------------ wiring
    @property
    def _p(self):
        if self.parent_env is None:
            raise RuntimeError("parent_env is None - initialization failed")
        return self.parent_env.pipeline

    def initialize_task_states(self):
        """:226-257: the reference registers seven task states; here they alias the pipeline's kernel-owned buffers."""
        for name, dtype in _TASK_STATES:
            if name not in self.task_states:
                self.register_task_state(name, (self.num_envs,), dtype=dtype)

    def finalize_setup(self):
        """:259-298: needs control_dt (measured by the parent)."""
        p = self._p
        self.contact_duration_threshold_steps = int(p.params.contact_duration_threshold_steps)
        self.box_local_actor_index = int(p.params.box_actor)
        self.box_local_rigid_body_index = int(p.params.box_body)
        self._failure_tensors = {k: p.state["failure_tensors"][i] for i, k in enumerate(_TASK_FAILURES)}

    def set_tensor_references(self, root_state_tensor: torch.Tensor):
        """:368-431: views of the box rows of actor_root_state."""
        self.root_state_tensor = root_state_tensor
        p = self._p
        box = root_state_tensor.view(self.num_envs, p.layout.num_actors, 13)[:, int(p.params.box_actor)]
        self.box_states, self.box_positions, self.box_velocities = box, box[:, 0:3], box[:, 7:10]

    @property
    def initial_box_positions(self) -> torch.Tensor:
        return self._p.view("initial_box_positions")

    def load_task_assets(self):
        """:300-319 creates the Isaac Gym box asset: simulator set-up, outside the step path (done by the host
        application that owns the gym handles)."""

    def create_task_objects(self, gym, sim, env_ptr, env_id: int):
        """:321-366 creates the box actor of one env: simulator set-up, outside the step path."""

    # ------------------------------------------------------------------ per-step hooks: views of the fused results
    def get_task_observations(self, obs_dict: Dict[str, torch.Tensor]) -> Optional[Dict[str, torch.Tensor]]:
        """:549-653 -> the task's obs_dict entries of the step the kernel just processed.  With
        ``materialize_obs_dict=True`` all 16 keys exist; otherwise the ones that are policy observation keys."""
        p = self._p
        if p.out["obs_dict"] is not None:
            d = p.obs_dict()
            return {k: d[k] for k in _TASK_OBS_KEYS}
        return {k: p.out["obs_buf"][:, a:b] for k, (a, b) in p.spec.obs_slices.items() if k in _TASK_OBS_KEYS}

    def compute_task_reward_terms(self, obs_dict: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        """:980-1053 -> the raw (unweighted) task terms of the current step."""
        return self.parent_env.reward_calculator.compute_task_reward_terms()

    def check_task_success_criteria(self, obs_dict: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
        """:1210-1236.  The kernel evaluated grasp_lift_success before it reset the finished envs; the mask that
        survives the step is the exclusive success class (== the criterion whenever it is the active success criterion)."""
        return {"grasp_lift_success": self._p.out["success"]}

    def check_task_failure_criteria(self, obs_dict: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
        """:1238-1332.  stage1_pregrasp_failed / stage2_contact_failed / stage3_grasp_lost are the kernel's own per-env
        masks of this step (DexStateBuffers.failure_tensors).  box_too_far is consumed inside the kernel and only its
        per-step mean is published (extras["failure_box_too_far"]); the mask returned here is re-derived on the device
        from the current sim tensors, i.e. for envs that reset in this step it describes the post-reset pose."""
        p = self._p
        out = {k: p.state["failure_tensors"][i] for i, k in enumerate(_TASK_FAILURES)}
        lay, P = p.layout, p.params
        hand = p.sim_tensors["rigid_body_state"].view(self.num_envs, lay.num_bodies, 13)[:, int(P.hand_body), 0:3]
        box = p.sim_tensors["actor_root_state"].view(self.num_envs, lay.num_actors, 13)[:, int(P.box_actor), 0:3]
        out["box_too_far"] = torch.norm(hand - box, dim=1) > float(P.max_box_distance)
        return out

    def reset_task_state(self, env_ids: torch.Tensor):
        """:433-547.  The task-state reset is part of the one masked reset launch (ResetManager.reset_idx ->
        dex_reset_idx; fused into dex_post_physics during a step), which the facade's ResetManager has already issued
        when the reference would call this hook - a second call must not randomise the envs again: no-op."""

    @property
    def task_observation_dims(self) -> Dict[str, int]:
        return {k: OBS_KEY_DIMS[k] for k in _TASK_OBS_KEYS}
