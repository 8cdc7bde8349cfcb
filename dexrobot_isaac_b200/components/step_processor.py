"""StepProcessor facade (reference: components/step_processor.py:12-232).

``process_physics_step`` of the reference (:37-131) orchestrates tensor refresh, observation
encoding, the pre-action rule, the step counter, termination, rewards, resets and extras through
~300 eager kernel launches and 6+ host syncs.  Here it is ONE launch (``dex_post_physics`` with the
reset of finished envs and the step epilogue fused in) and zero host syncs; the returned tensors
are the pipeline's output buffers.
"""
from __future__ import annotations


class StepProcessor:
    def __init__(self, parent):
        self.parent = parent

    @property
    def device(self):
        return self.parent.device

    def process_physics_step(self):
        """-> (obs_buf, rew_buf, reset_buf, extras), step_processor.py:37-131."""
        env = self.parent
        if getattr(env, "generic_task", False):
            return self._process_physics_step_generic()
        env.pipeline.post_step()
        self._update_extras()
        env.obs_dict = env.observation_encoder.obs_dict() if env.pipeline.out["obs_dict"] is not None else \
            {"active_rule_targets": env.action_processor.active_rule_targets}
        env.last_reward_components = env.extras["reward_components"]
        return env.obs_buf, env.rew_buf, env.reset_buf, env.extras

    def _update_extras(self, termination_info=None):
        """step_processor.py:221-232.  Lazy: the entries are views / slices of device buffers, no host sync happens
        here.  Works for any parent that owns a `pipeline` (it does not have to be DexStepEnv)."""
        from .. import _lib as L
        from ..params import FAILURE_CRITERIA, SUCCESS_CRITERIA

        C = L.CONSTS
        env = self.parent
        p, sp = env.pipeline, env.pipeline.spec
        stats = p.epi["stats"]
        ex = {"consecutive_successes": p.epi["consecutive_successes"][0],
              "episode_length": p.state["episode_step_count"],
              "success": p.out["success"], "failure": p.out["failure"], "timeout": p.out["timeout"],
              "success_rate": stats[C["DEX_STAT_SUCCESS_RATE"]], "failure_rate": stats[C["DEX_STAT_FAILURE_RATE"]],
              "timeout_rate": stats[C["DEX_STAT_TIMEOUT_RATE"]]}
        tm = getattr(env, "termination_manager", None)
        for c, name in enumerate(SUCCESS_CRITERIA + FAILURE_CRITERIA):
            kind = "success" if name in SUCCESS_CRITERIA else "failure"
            if name in (sp.active_success if kind == "success" else sp.active_failure):
                ex[f"{kind}_{name}"] = stats[C["DEX_STAT_CRIT0"] + c]
                # termination_manager.py:353-357 + step_processor.py:109-111: every env whose criterion fired was
                # reset in the same step, which clears its sticky reason bit before extras are built -> the masks
                # that reach extras are all-False.
                if tm is not None:
                    ex[f"{kind}_reason_{name}"] = (tm.success_reasons if kind == "success" else tm.failure_reasons)[name]
        if termination_info:
            ex.update(termination_info)
        ex["reward_components"] = p.reward_components()
        env.extras = ex

    # the reference's private helpers, kept as thin accessors of the fused results
    def _evaluate_termination(self):
        return self.parent.termination_manager.evaluate()

    def _compute_rewards(self):
        return self.parent.reward_calculator.compute_total_reward()[1]

    # ------------------------------------------------------------------ generic-task route
    def _process_physics_step_generic(self):
        """A task with Python hooks: ONE launch of the obs_dict-materialising kernel instance with the reset held back
        (DEX_FLAG_NO_FUSED_RESET), then the hooks on the device tensors in the reference's order (step_processor.py:66-131):
        pre-action rule -> extra task reward terms -> reset of the finished envs (one dex_reset_idx launch)."""
        import torch

        from ..pipeline import FLAG_NO_FUSED_RESET

        env = self.parent
        p, ap, task = env.pipeline, env.action_processor, env.task
        p.post_step(flags=FLAG_NO_FUSED_RESET)
        self._update_extras()
        obs_dict = env.observation_encoder.obs_dict()
        # pre-action rule (step_processor.py:66-73): the kernel wrote the identity rule; a custom rule replaces it
        if ap.pre_action_rule is not None:
            art = ap.pre_action_rule(ap.active_prev_targets, {"obs_dict": obs_dict, "env": env})
            ap.active_rule_targets.copy_(art)                        # the state the next process_actions reads
            obs_dict["active_rule_targets"].copy_(art)               # the materialised obs_dict entry (a buffer view)
            sl = p.spec.obs_slices.get("active_rule_targets")
            if sl is not None:
                env.obs_buf[:, sl[0]:sl[1]].copy_(art)
        env.obs_dict = obs_dict
        # task reward terms the kernel does not know (step_processor.py:183-202, reward_calculator.py:223-275)
        comps = dict(env.extras["reward_components"])
        weights = env.reward_calculator.reward_weights
        for name, val in (task.compute_task_reward_terms(obs_dict) or {}).items():
            if name in comps:
                continue  # a kernel term handed back by a facade
            w = float(weights.get(name, 0.0))
            if w != 0:
                weighted = val * w
                env.rew_buf.add_(weighted)
                comps[name], comps[f"{name}_weighted"] = val, weighted
                comps["total"] = comps["total"] + weighted
        env.extras["reward_components"] = comps
        env.last_reward_components = comps
        # reset of the finished envs (step_processor.py:109-111), after the hooks have seen the terminal state
        env.reset_manager.reset_mask(env.reset_buf)
        return env.obs_buf, env.rew_buf, env.reset_buf, env.extras
