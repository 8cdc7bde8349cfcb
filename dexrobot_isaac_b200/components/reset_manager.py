"""ResetManager facade (reference: components/reset/reset_manager.py:16-200).

``reset_idx`` of the reference (:92-190) resets the episode counter, DOF state, action targets,
observer state, termination tracking and calls ``task.reset_task_state``; then applies the DOF /
root states to the simulator through the indexed gym setters.  Here the whole per-env state reset
(incl. BlindGraspingTask.reset_task_state, blind_grasping_task.py:433-547) is one masked launch of
``dex_reset_idx``; during a control step the same reset is fused into ``dex_post_physics``.  With
Isaac Gym present the caller still has to push ``dof_state`` / ``actor_root_state`` with the indexed
setters - ``parent.pipeline.env_ids()`` is the ascending id list (== ``torch.nonzero(reset_buf)``).
"""
from __future__ import annotations

import torch


class ResetManager:
    def __init__(self, parent, dof_state=None, root_state_tensor=None, hand_local_actor_index: int = 0,
                 default_dof_pos=None, task=None):
        self.parent = parent
        self.dof_state = dof_state if dof_state is not None else parent.dof_state
        self.root_state_tensor = root_state_tensor if root_state_tensor is not None else parent.actor_root_state_tensor
        self.hand_local_actor_index = hand_local_actor_index
        self.task = task
        if default_dof_pos is None:
            raise RuntimeError("default_dof_pos is None - ResetManager requires default DOF positions")
        if float(torch.as_tensor(default_dof_pos).abs().max()) != 0.0:
            # reset_manager.py:107-124 writes default_dof_pos; the shipped configs use zeros, which is what the
            # fused reset writes before the task randomisation.
            raise NotImplementedError("non-zero default_dof_pos is not part of the fused reset kernel")
        self.default_dof_pos = default_dof_pos

    @property
    def num_envs(self):
        return self.parent.num_envs

    @property
    def device(self):
        return self.parent.device

    @property
    def action_processor(self):
        return self.parent.action_processor

    @property
    def episode_step_count(self):
        return self.parent.episode_step_count

    def reset_idx(self, env_ids) -> bool:
        """reset_manager.py:92-190 (masked form of the id list).  Returns True like the reference."""
        ids = torch.as_tensor(env_ids, device=self.device, dtype=torch.long)
        if ids.numel() == 0:
            return True
        if int(ids.min()) < 0 or int(ids.max()) >= self.num_envs:
            raise RuntimeError(f"env_ids out of range [0, {self.num_envs})")
        mask = torch.zeros(self.num_envs, dtype=torch.bool, device=self.device)
        mask[ids] = True
        self.parent.pipeline.reset_idx(mask)
        tm = getattr(self.parent, "termination_manager", None)
        if tm is not None:
            tm.reset_tracking(ids)
        return True

    def reset_mask(self, mask: torch.Tensor) -> bool:
        """reset_idx for a boolean mask that lives on the device (no nonzero(), no host sync)."""
        self.parent.pipeline.reset_idx(mask)
        tm = getattr(self.parent, "termination_manager", None)
        if tm is not None:
            for t in [tm.episode_success, tm.episode_failure, *tm.success_reasons.values(), *tm.failure_reasons.values()]:
                t.logical_and_(~mask.bool())
        return True

    def reset_all(self) -> bool:
        """reset_manager.py:192-200."""
        self.parent.pipeline.reset_idx(None)
        tm = getattr(self.parent, "termination_manager", None)
        if tm is not None:
            tm.reset_tracking(torch.arange(self.num_envs, device=self.device))
        return True
