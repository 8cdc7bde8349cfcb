"""DexTask - the task plugin interface (reference: dexhand_env/tasks/task_interface.py:14-405).

Same hook names, argument names and defaults as the reference.  A task whose hooks are all the defaults below (or one of
the two shipped tasks) runs entirely inside the fused kernels; a task that overrides ``pre_action_rule`` / ``action_rule`` /
``post_action_filters`` + ``register_custom_filters``, or adds reward terms in ``compute_task_reward_terms``, is routed to
the un-fused GPU path at construction time (``DexStepEnv``: the DICT kernel instance materialises ``obs_dict`` on the
device, the Python hooks run on those tensors, ``dex_apply_coupling`` finishes the action pipeline) - SURVEY 8f N4."""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Callable, Dict, List, Optional

import torch


class DexTask(ABC):
    parent_env = None  # set by the environment that owns the task (dexhand_base.py:430-452)

    # ------------------------------------------------------------------ task_interface.py:23-84
    @property
    def task_states(self):
        if self.parent_env is None:
            raise RuntimeError("parent_env is None - initialization failed")
        return self.parent_env.observation_encoder.task_states

    def register_task_state(self, name: str, shape: tuple, dtype=torch.float32):
        if self.parent_env is None:
            raise RuntimeError("parent_env is None - initialization failed")
        return self.parent_env.observation_encoder.register_task_state(name, shape, dtype=dtype)

    @property
    def observation_encoder(self):
        if self.parent_env is None:
            raise RuntimeError("parent_env is None - initialization failed")
        return self.parent_env.observation_encoder

    # ------------------------------------------------------------------ abstract hooks :86-169
    @abstractmethod
    def compute_task_reward_terms(self, obs_dict: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        ...

    @abstractmethod
    def check_task_success_criteria(self, obs_dict: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
        ...

    @abstractmethod
    def check_task_failure_criteria(self, obs_dict: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
        ...

    @abstractmethod
    def reset_task_state(self, env_ids: torch.Tensor):
        ...

    @abstractmethod
    def create_task_objects(self, gym, sim, env_ptr, env_id: int):
        ...

    @abstractmethod
    def load_task_assets(self):
        ...

    # ------------------------------------------------------------------ optional hooks :170-405 (reference defaults)
    def get_task_observations(self, obs_dict: Dict[str, torch.Tensor]) -> Optional[Dict[str, torch.Tensor]]:
        return None

    def get_task_dof_targets(self, num_envs: int, device: str, base_controlled: bool,
                             fingers_controlled: bool) -> Optional[Dict[str, torch.Tensor]]:
        return None

    def set_tensor_references(self, root_state_tensor: torch.Tensor):
        pass

    def initialize_task_states(self):
        pass

    def finalize_setup(self):
        pass

    @property
    def pre_action_rule(self) -> Optional[Callable]:
        """(active_prev_targets, state) -> active_rule_targets, or None for the identity rule (:290-317)."""
        return None

    @property
    def action_rule(self) -> Optional[Callable]:
        """(active_prev_targets, active_rule_targets, actions, config) -> raw targets, or None for the default
        position / position_delta rule (:319-356)."""
        return None

    @property
    def post_action_filters(self) -> List[str]:
        """Names of additional post-action filters, registered through register_custom_filters (:358-377)."""
        return []

    def register_custom_filters(self, action_processor) -> None:
        pass

    # ------------------------------------------------------------------ routing helper (not in the reference)
    def uses_custom_action_hooks(self) -> bool:
        """True when the task overrides any hook the fused pre-physics kernel cannot honour."""
        return self.pre_action_rule is not None or self.action_rule is not None or len(self.post_action_filters) > 0
