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
        env.pipeline.post_step()
        env._update_extras()
        env.obs_dict = env.observation_encoder.obs_dict() if env.pipeline.out["obs_dict"] is not None else \
            {"active_rule_targets": env.action_processor.active_rule_targets}
        env.last_reward_components = env.extras["reward_components"]
        return env.obs_buf, env.rew_buf, env.reset_buf, env.extras

    # the reference's private helpers, kept as thin accessors of the fused results
    def _evaluate_termination(self):
        return self.parent.termination_manager.evaluate()

    def _compute_rewards(self):
        return self.parent.reward_calculator.compute_total_reward()[1]
