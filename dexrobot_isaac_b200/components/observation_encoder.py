"""ObservationEncoder facade (reference: components/observation/observation_encoder.py).

In the fused design the whole encoder (`_compute_default_observations` :576-758, task observations
via `BlindGraspingTask.get_task_observations` blind_grasping_task.py:549-653, hand-frame transforms
:915-986, contact-duration tracking :323-361, `_concat_selected_observations` :783-829) runs inside
``dex_post_physics``.  ``compute_observations`` therefore launches that kernel in
``DEX_MODE_OBS_ONLY`` (what ``DexHandBase.reset`` needs, dexhand_base.py:813); during a control
step ``StepProcessor.process_physics_step`` does the single full launch and this class only hands
out views of its results.

``obs_dict`` entries are views into one ``[N, dict_dim]`` buffer written by the kernel when the
pipeline was built with ``materialize_obs_dict=True`` (the policy needs ``obs_buf`` only; the
privileged dictionary is a debug / viewer / custom-task facility, SURVEY §8d excludes it from the
algorithmic bytes).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

from ..params import BASE_JOINT_NAMES, FINGERPAD_BODY_NAMES, FINGERTIP_BODY_NAMES, HARDWARE_MAPPING
from ..pipeline import FLAG_STALE_CONTACTS, MODE_OBS_ONLY


class ObservationEncoder:
    def __init__(self, parent):
        self.parent = parent
        self.task_states: Dict[str, torch.Tensor] = {}
        self.observation_keys: List[str] = []
        self.component_slice_indices: Dict[str, Tuple[int, int]] = {}
        self.action_processor = None
        self._initialized = False

    @property
    def _p(self):
        return self.parent.pipeline

    @property
    def num_envs(self):
        return self.parent.num_envs

    @property
    def device(self):
        return self.parent.device

    @property
    def control_dt(self):
        return self.parent.control_dt

    @property
    def num_observations(self) -> int:
        return self._p.spec.obs_dim

    @property
    def obs_buf(self) -> torch.Tensor:
        return self._p.out["obs_buf"]

    @property
    def states_buf(self) -> torch.Tensor:
        """observation_encoder.py:474-476: a copy of obs_buf for asymmetric actor-critic.  The copy is made on
        access (the fused path does not pay 2x the observation bytes every step for it)."""
        return self._p.out["obs_buf"].clone()

    @property
    def prev_actions(self) -> torch.Tensor:
        return self._p.view("prev_actions")

    @property
    def prev_dof_pos(self) -> torch.Tensor:
        return self._p.view("prev_dof_pos")

    @property
    def contact_duration_steps(self) -> torch.Tensor:
        return self._p.view("contact_duration_steps")

    @property
    def prev_contact_binary(self) -> torch.Tensor:
        return self._p.view("prev_contact_binary")

    @property
    def active_finger_dof_indices(self) -> torch.Tensor:
        return torch.tensor(list(self._p.params.active_finger_dof), dtype=torch.long, device=self.device)

    # ------------------------------------------------------------------ setup (:119-243)
    def initialize(self, observation_keys: List[str], joint_to_control: Dict[str, str] = None,
                   active_joint_names: List[str] = None, num_actions: int = None, action_processor=None,
                   index_mappings: Dict = None):
        spec = self._p.spec
        if list(observation_keys) != list(spec.obs_keys):
            raise RuntimeError("observation_keys differ from the policy_observation_keys the pipeline was built with: "
                               f"{list(observation_keys)} vs {spec.obs_keys}")
        if num_actions is not None and int(num_actions) != spec.num_actions:
            raise RuntimeError(f"num_actions {num_actions} != {spec.num_actions}")
        self.observation_keys = list(observation_keys)
        self.joint_to_control = dict(joint_to_control or {j: c for c, js in HARDWARE_MAPPING for j in js})
        self.active_joint_names = list(active_joint_names or [js[0] for _, js in HARDWARE_MAPPING])
        self.num_actions = spec.num_actions
        self.action_processor = action_processor
        self.index_mappings = index_mappings or {}
        self.component_slice_indices = dict(spec.obs_slices)
        self._initialized = True

    # ------------------------------------------------------------------ state the encoder owns
    def update_prev_actions(self, actions: torch.Tensor):
        """observation_encoder.py:287-296.  ``ActionProcessor.process_actions`` (dex_pre_physics) already stored
        the actions it processed; only a different tensor needs a copy."""
        buf = self.action_processor.action_buffer if self.action_processor is not None else None
        if buf is not None and actions.data_ptr() == buf.data_ptr():
            return
        a = actions.to(self.device, torch.float32)
        if a.shape != self.prev_actions.shape:
            raise RuntimeError(f"actions shape {tuple(a.shape)} != {tuple(self.prev_actions.shape)}")
        self.prev_actions.copy_(a)  # a strided device copy; no comparison (that would be a host sync per step)

    def reset_observer_state(self, env_ids: torch.Tensor):
        """observation_encoder.py:363-383: contact duration / previous contacts / previous DOF positions (zero,
        not the new pose) / previous actions of the given envs."""
        ids = torch.as_tensor(env_ids, device=self.device, dtype=torch.long)
        if ids.numel() == 0:
            return
        st = self._p.state
        for name in ("contact_duration_steps", "prev_contact_binary", "prev_dof_pos", "prev_actions"):
            st[name][:, ids] = 0

    def register_task_state(self, name: str, shape: Tuple[int, ...], dtype: torch.dtype = torch.float32):
        """:385-408.  The BlindGrasping states live in the pipeline (kernel-owned); their names map onto those
        buffers, anything else gets a plain tensor that the fused kernels never touch."""
        if name in self.task_states:
            raise ValueError(f"Task state '{name}' already registered")
        if shape[0] != self.num_envs:
            raise ValueError(f"First dimension of task state must be num_envs ({self.num_envs}), got {shape[0]}")
        st = self._p.state
        if name in st and tuple(self._p.view(name).shape) == tuple(shape):
            self.task_states[name] = self._p.view(name)
        else:
            self.task_states[name] = torch.zeros(shape, dtype=dtype, device=self.device)

    def get_task_state(self, name: str) -> torch.Tensor:
        if name not in self.task_states:
            raise KeyError(f"Task state '{name}' not registered")
        return self.task_states[name]

    def set_task_state(self, name: str, value: torch.Tensor):
        if name not in self.task_states:
            raise KeyError(f"Task state '{name}' not registered")
        if value.shape != self.task_states[name].shape:
            raise ValueError(f"Shape mismatch for task state '{name}': expected {self.task_states[name].shape}, "
                             f"got {value.shape}")
        self.task_states[name][:] = value

    # ------------------------------------------------------------------ observations
    def compute_observations(self, exclude_components: Optional[List[str]] = None, stale_contacts: bool = False):
        """:442-478 -> one dex_post_physics launch in DEX_MODE_OBS_ONLY (updates the finite-difference and
        contact-duration state exactly like the reference's call does).  Returns ``obs_dict`` when
        ``exclude_components`` is given, else ``(obs_buf, obs_dict)``."""
        p = self._p
        p.post_physics(MODE_OBS_ONLY, FLAG_STALE_CONTACTS if stale_contacts else 0)
        d = self.obs_dict()
        if exclude_components:
            for c in exclude_components:
                d.pop(c, None)
            return d
        return self.obs_buf, d

    def obs_dict(self) -> Dict[str, torch.Tensor]:
        """The obs_dict of the last launch (requires ``materialize_obs_dict=True``), else the policy keys as
        slices of obs_buf."""
        p = self._p
        if p.out["obs_dict"] is not None:
            return p.obs_dict()
        return {k: self.obs_buf[:, a:b] for k, (a, b) in p.spec.obs_slices.items()}

    def concatenate_observations(self, obs_dict: Dict[str, torch.Tensor]) -> torch.Tensor:
        """:480-494.  The kernel already concatenated in ``policy_observation_keys`` order; entries of
        ``obs_dict`` that are not the kernel's own views are copied over the corresponding columns."""
        own = {self.obs_buf.untyped_storage().data_ptr()}
        if self._p.out["obs_dict"] is not None:
            own.add(self._p.out["obs_dict"].untyped_storage().data_ptr())
        for k, (a, b) in self._p.spec.obs_slices.items():
            if k not in obs_dict:
                raise KeyError(f"observation key '{k}' missing from obs_dict")
            v = obs_dict[k]
            if v.untyped_storage().data_ptr() not in own:  # a tensor the caller replaced: honour it
                self.obs_buf[:, a:b].copy_(v.reshape(self.num_envs, b - a))
        return self.obs_buf

    def get_observation_space(self):
        import numpy as np
        return {"shape": (self.num_observations,), "low": -np.inf, "high": np.inf, "dtype": np.float32}

    # ------------------------------------------------------------------ index helpers (:999-1481)
    def _slice(self, key: str) -> Tuple[int, int]:
        if key not in self.component_slice_indices:
            raise KeyError(f"'{key}' is not part of the policy observation")
        return self.component_slice_indices[key]

    def get_obs_index_for_base_joint(self, joint_name: str, obs_type: str = "pos") -> int:
        key = {"pos": "base_dof_pos", "vel": "base_dof_vel", "target": "base_dof_target"}[obs_type]
        return self._slice(key)[0] + BASE_JOINT_NAMES.index(joint_name)

    def get_obs_index_for_finger_control(self, control_name: str, obs_type: str = "pos") -> int:
        key = {"pos": "active_finger_dof_pos", "vel": "active_finger_dof_vel",
               "target": "active_finger_dof_target"}[obs_type]
        names = [c for c, _ in HARDWARE_MAPPING]
        return self._slice(key)[0] + names.index(control_name)

    def get_obs_range_for_finger_pose(self, finger_body_name: str, frame: str = "world") -> Tuple[int, int]:
        if finger_body_name in FINGERTIP_BODY_NAMES:
            key, i = f"fingertip_poses_{frame}", FINGERTIP_BODY_NAMES.index(finger_body_name)
        else:
            key, i = f"fingerpad_poses_{frame}", FINGERPAD_BODY_NAMES.index(finger_body_name)
        a = self._slice(key)[0] + 7 * i
        return a, a + 7

    def get_obs_range_for_contact_force(self, finger_idx: int) -> Tuple[int, int]:
        a = self._slice("contact_forces")[0] + 3 * finger_idx
        return a, a + 3

    def get_base_dof_value(self, joint_name: str, obs_type: str = "pos", obs_dict: Dict = None, env_idx=None):
        key = {"pos": "base_dof_pos", "vel": "base_dof_vel", "target": "base_dof_target"}[obs_type]
        d = obs_dict if obs_dict is not None else self.obs_dict()
        v = d[key][:, BASE_JOINT_NAMES.index(joint_name)]
        return v if env_idx is None else v[env_idx]

    def get_active_finger_dof_value(self, control_name: str, obs_type: str = "pos", obs_dict: Dict = None,
                                    env_idx=None):
        key = {"pos": "active_finger_dof_pos", "vel": "active_finger_dof_vel",
               "target": "active_finger_dof_target"}[obs_type]
        d = obs_dict if obs_dict is not None else self.obs_dict()
        v = d[key][:, [c for c, _ in HARDWARE_MAPPING].index(control_name)]
        return v if env_idx is None else v[env_idx]

    def parse_obs_tensor(self, obs_tensor: torch.Tensor = None, obs_dict: Dict[str, torch.Tensor] = None,
                         env_idx: int = 0):
        """:1452-1481 -> ObservationParser over one env's observation row."""
        if obs_tensor is None:
            obs_tensor = self.obs_buf
        if obs_tensor is None:
            raise ValueError("No observation tensor available")
        row = obs_tensor if obs_tensor.dim() == 1 else obs_tensor[env_idx]
        return ObservationParser(row, self, obs_dict)


class ObservationParser:
    """observation_encoder.py:1523-1665: named access into one observation row."""

    def __init__(self, obs_data: torch.Tensor, encoder: ObservationEncoder, obs_dict: Dict = None):
        self.obs_data, self.encoder, self.obs_dict = obs_data, encoder, obs_dict

    def get_base_dof(self, joint_name: str, obs_type: str = "pos"):
        return self.obs_data[self.encoder.get_obs_index_for_base_joint(joint_name, obs_type)]

    def get_finger_dof(self, control_name: str, obs_type: str = "pos"):
        return self.obs_data[self.encoder.get_obs_index_for_finger_control(control_name, obs_type)]

    def get_finger_pose(self, finger_body_name: str, frame: str = "world"):
        a, b = self.encoder.get_obs_range_for_finger_pose(finger_body_name, frame)
        pose = self.obs_data[a:b]
        return {"position": pose[:3], "orientation": pose[3:7]}

    def get_contact_force(self, finger_idx: int):
        a, b = self.encoder.get_obs_range_for_contact_force(finger_idx)
        return self.obs_data[a:b]
