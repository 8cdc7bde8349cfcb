"""RewardCalculator facade (reference: components/reward/reward_calculator.py:15-275).

All reward terms (10 common terms :66-221, 16 BlindGrasping terms blind_grasping_task.py:980-1208),
their weighting and the ordered sum (:223-275) are computed by ``dex_post_physics`` in the same launch
as the observations; the four prev-state tensors of the common terms live in the pipeline.  The two
public methods hand out the rows the kernel wrote for the CURRENT step (they are views, valid until
the next launch); they do not recompute.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import torch

from .. import _lib as L

_COMMON = [L.REWARD_TERM_NAMES[i] for i in range(L.CONSTS["DEX_NUM_COMMON_TERMS"])]


class RewardCalculator:
    def __init__(self, parent, task_cfg: dict):
        self.parent = parent
        self.task_cfg = task_cfg
        rw = task_cfg.get("reward_weights")
        if rw is None:
            raise RuntimeError("task_cfg['reward_weights'] is required")  # reward_calculator.py:41-45 (fail fast)
        self.reward_weights: Dict[str, float] = {k: v for k, v in rw.items() if k != "_delete_"}

    @property
    def _p(self):
        return self.parent.pipeline

    @property
    def num_envs(self):
        return self.parent.num_envs

    @property
    def device(self):
        return self.parent.device

    # prev-state of the common terms (reward_calculator.py:47-55); None until the first call, like the reference
    def _prev(self, name):
        p = self._p
        return p.view(name) if p._reward_initialized and p.params.need_common_rewards else None

    @property
    def prev_finger_dof_vel(self):
        return self._prev("prev_finger_dof_vel")

    @property
    def prev_hand_vel(self):
        return self._prev("prev_hand_vel")

    @property
    def prev_hand_ang_vel(self):
        return self._prev("prev_hand_ang_vel")

    @property
    def prev_contacts(self):
        return self._prev("prev_contacts")

    def compute_common_reward_terms(self, obs_dict=None, parent_env=None) -> Dict[str, torch.Tensor]:
        """:66-221 -> the raw common terms of the current step (only the non-zero-weight ones are
        materialised by the kernel; the total is unaffected, :247-251)."""
        comps = self._p.reward_components()
        return {k: comps[k] for k in _COMMON if k in comps}

    def compute_task_reward_terms(self) -> Dict[str, torch.Tensor]:
        comps = self._p.reward_components()
        return {k: v for k, v in comps.items()
                if k not in _COMMON and not k.endswith("_weighted") and k != "total" and not k.startswith("termination_")}

    def compute_total_reward(self, common_rewards=None, task_rewards=None,
                             success_failure_rewards: Optional[dict] = None) -> Tuple[torch.Tensor, Dict[str, torch.Tensor]]:
        """:223-275 -> (total reward WITHOUT termination rewards, components {name, name_weighted, total})."""
        comps = self._p.reward_components()
        # terms the kernel does not know (a custom task's extra reward) or values that are not the kernel's own rows
        # cannot be folded into the fused total: refuse instead of returning something else than the caller asked for
        for d in (common_rewards, task_rewards):
            for name, v in (d or {}).items():
                if name not in comps:
                    raise NotImplementedError(f"reward term '{name}' is not computed by the fused kernel (SURVEY 8f N4)")
                if isinstance(v, torch.Tensor) and v.data_ptr() != comps[name].data_ptr() and v.shape == comps[name].shape \
                        and not torch.equal(v, comps[name]):
                    raise NotImplementedError(f"reward term '{name}': the values handed in differ from the kernel's")
        out = {k: v for k, v in comps.items() if not k.startswith("termination_")}
        return comps["total"], out
