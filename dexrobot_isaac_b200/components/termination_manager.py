"""TerminationManager facade (reference: components/termination/termination_manager.py:11-359).

``evaluate`` of the reference (:120-278) OR-reduces the active success / failure criteria, applies
the timeout, makes the three classes exclusive and builds per-criterion statistics.  The fused
post-physics kernel does this per env (masks, bit-exact) and its epilogue finishes the cross-env
reductions (rates, criterion means, consecutive successes) on the device: no host sync.  This class
exposes the results under the reference's names.
"""
from __future__ import annotations

from typing import Dict, Tuple

import torch

from .. import _lib as L
from ..params import FAILURE_CRITERIA, SUCCESS_CRITERIA

C = L.CONSTS


class TerminationManager:
    def __init__(self, parent, task_cfg: dict):
        self.parent = parent
        self.task_cfg = task_cfg
        p = parent.pipeline
        term = task_cfg.get("termination", {}) or {}
        self.active_success_criteria = list(term.get("active_success_criteria", []) or [])
        self.active_failure_criteria = list(term.get("active_failure_criteria", []) or [])
        self.max_episode_length = int(p.params.max_episode_length)
        self.max_consecutive_successes = int(p.params.max_consecutive_successes)
        rw = task_cfg["reward_weights"]
        self.success_reward = float(rw["termination_success"])
        self.failure_penalty = float(rw["termination_failure_penalty"])
        self.timeout_penalty = float(rw["termination_timeout_penalty"])
        n, dev = parent.num_envs, parent.device
        # sticky per-episode flags (:64-76, :87-96).  Every env whose criterion fires is reset in the same step,
        # and reset_tracking clears them before extras are built, so outside of a step they
        # are all-False; kept as real tensors for API compatibility.
        self.episode_success = torch.zeros(n, dtype=torch.bool, device=dev)
        self.episode_failure = torch.zeros(n, dtype=torch.bool, device=dev)
        self.success_reasons: Dict[str, torch.Tensor] = {k: torch.zeros(n, dtype=torch.bool, device=dev)
                                                          for k in p.spec.active_success}
        self.failure_reasons: Dict[str, torch.Tensor] = {k: torch.zeros(n, dtype=torch.bool, device=dev)
                                                          for k in p.spec.active_failure}

    @property
    def num_envs(self):
        return self.parent.num_envs

    @property
    def device(self):
        return self.parent.device

    @property
    def consecutive_successes(self) -> torch.Tensor:
        """Device-resident 0-d int32 (the reference keeps a host int and pays a sync per step, :331)."""
        return self.parent.pipeline.epi["consecutive_successes"][0]

    def validate_criteria(self, builtin_dict, task_dict, active_list, criteria_type):
        """:98-118: every active criterion must be provided by the env or the task."""
        have = set(builtin_dict) | set(task_dict)
        for name in active_list:
            if name not in have:
                raise RuntimeError(f"{criteria_type} criterion '{name}' is active but not implemented")

    def evaluate(self, episode_step_count=None, builtin_success=None, task_success=None, builtin_failure=None,
                 task_failure=None) -> Tuple[torch.Tensor, dict, dict, dict]:
        """:120-278 -> (should_reset, termination_info, termination_rewards, raw_termination_rewards) of the
        step the fused kernel just processed.  The criterion dictionaries of the reference signature are
        accepted for validation only: the criteria themselves were evaluated in the kernel."""
        if builtin_success is not None:
            self.validate_criteria(builtin_success or {}, task_success or {}, self.active_success_criteria, "Success")
            self.validate_criteria(builtin_failure or {}, task_failure or {}, self.active_failure_criteria, "Failure")
        p = self.parent.pipeline
        # The criteria were evaluated inside the kernel: dictionaries that carry anything else than the kernel's own
        # criteria (a custom task's extra criterion, or a mask the caller computed differently) cannot be honoured here
        # and must not be ignored silently - such tasks belong on the generic-task route (DexStepEnv(task=...)).
        known = set(SUCCESS_CRITERIA) | set(FAILURE_CRITERIA)
        for d in (builtin_success, task_success, builtin_failure, task_failure):
            for name, mask in (d or {}).items():
                if name not in known:
                    raise NotImplementedError(f"criterion '{name}' is not evaluated by the fused kernel (SURVEY 8f N4)")
                own = self._own_mask(name)
                if own is not None and isinstance(mask, torch.Tensor) and mask.data_ptr() != own.data_ptr() and \
                        mask.shape == own.shape and not torch.equal(mask.to(own.dtype), own):
                    raise NotImplementedError(f"criterion '{name}': the mask handed in differs from the one the fused "
                                              "kernel evaluated for this step")
        stats, out = p.epi["stats"], p.out
        info = {"success": out["success"], "failure": out["failure"], "timeout": out["timeout"],
                "success_rate": stats[C["DEX_STAT_SUCCESS_RATE"]], "failure_rate": stats[C["DEX_STAT_FAILURE_RATE"]],
                "timeout_rate": stats[C["DEX_STAT_TIMEOUT_RATE"]]}
        for c, name in enumerate(SUCCESS_CRITERIA + FAILURE_CRITERIA):
            kind = "success" if name in SUCCESS_CRITERIA else "failure"
            if name in (p.spec.active_success if kind == "success" else p.spec.active_failure):
                info[f"{kind}_{name}"] = stats[C["DEX_STAT_CRIT0"] + c]
                info[f"{kind}_reason_{name}"] = (self.success_reasons if kind == "success" else self.failure_reasons)[name]
        comps = p.reward_components()
        rewards = {"success": comps["termination_success_weighted"],
                   "failure_penalty": comps["termination_failure_penalty_weighted"],
                   "timeout_penalty": comps["termination_timeout_penalty_weighted"]}
        raw = {"success": comps["termination_success"], "failure_penalty": comps["termination_failure_penalty"],
               "timeout_penalty": comps["termination_timeout_penalty"]}
        return out["reset_buf"], info, rewards, raw

    def _own_mask(self, name):
        """The kernel's per-env mask of a criterion, where it is materialised."""
        p = self.parent.pipeline
        ft = ("stage1_pregrasp_failed", "stage2_contact_failed", "stage3_grasp_lost")
        if name in ft and "failure_tensors" in p.state:
            return p.state["failure_tensors"][ft.index(name)]
        return None

    def update_consecutive_successes(self, success_tensor=None):
        """:323-339.  Done by the step epilogue on the device (``min(c + 1, max)`` if any env succeeded, else
        0); calling it again would double-count, so this is a no-op kept for call-site compatibility."""
        return None

    def reset_tracking(self, env_ids):
        """:341-357."""
        ids = torch.as_tensor(env_ids, device=self.device, dtype=torch.long)
        self.episode_success[ids] = False
        self.episode_failure[ids] = False
        for t in list(self.success_reasons.values()) + list(self.failure_reasons.values()):
            t[ids] = False
