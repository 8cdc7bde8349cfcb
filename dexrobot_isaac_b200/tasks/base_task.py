"""BaseTask facade (reference: dexhand_env/tasks/base_task.py:14-160): the minimal task - no task rewards, no task
criteria, no task observations, no objects."""
from __future__ import annotations

from typing import Dict, Optional

import torch

from .task_interface import DexTask


class BaseTask(DexTask):
    def __init__(self, sim, gym, device, num_envs, cfg):
        self.sim, self.gym, self.device, self.num_envs, self.cfg = sim, gym, device, num_envs, cfg
        self.parent_env = None

    def compute_task_reward_terms(self, obs_dict: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        return {}  # base_task.py:48-63

    def check_task_success_criteria(self, obs_dict: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
        return {}  # :65-80

    def check_task_failure_criteria(self, obs_dict: Optional[Dict[str, torch.Tensor]] = None) -> Dict[str, torch.Tensor]:
        return {}  # :82-97

    def reset_task_state(self, env_ids: torch.Tensor):
        pass  # :99-108

    def create_task_objects(self, gym, sim, env_ptr, env_id: int):
        pass  # :110-122

    def load_task_assets(self):
        pass  # :124-130

    def get_task_observations(self, obs_dict: Dict[str, torch.Tensor]) -> Optional[Dict[str, torch.Tensor]]:
        return None  # :132-146

    def set_tensor_references(self, root_state_tensor: torch.Tensor):
        self.root_state_tensor = root_state_tensor  # :148-160
