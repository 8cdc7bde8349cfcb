"""A toy third task for the generic-task route: one custom action rule, one custom pre-action rule, one
custom post-action filter and one extra reward term, written as device-agnostic torch functions with the reference's hook
signatures (components/action/rules.py:196-244, tasks/task_interface.py:286-405).

TEST INFRASTRUCTURE ONLY.  The same callables are plugged into
  * the UNMODIFIED reference components (oracle/ref_harness.RefEnv(hooks=...)),
  * the oracle port (OracleEnv.hooks) - pinned against the former by tests/test_oracle_vs_reference.py,
  * the CUDA product through a DexTask subclass (tests/test_cuda_generic_task.py).
"""
import torch

EXTRA_REWARD_NAME = "toy_hand_height"
EXTRA_REWARD_WEIGHT = 0.7
FILTER_NAME = "toy_deadband"


def action_rule(active_prev_targets, active_rule_targets, actions, config):
    """Leaky position-delta rule: the rule targets pull 10 % towards themselves, the policy adds a small delta."""
    t = 0.9 * active_prev_targets + 0.1 * active_rule_targets
    fs = 6 if config["policy_controls_base"] else 0
    if config["policy_controls_base"]:
        t[:, :6] = t[:, :6] + 0.004 * actions[:, :6]
    if config["policy_controls_fingers"]:
        t[:, 6:] = t[:, 6:] + 0.015 * actions[:, fs:fs + 12]
    return t


def pre_action_rule(active_prev_targets, state):
    """Rule targets: previous targets with the vertical base target steered against the observed hand height."""
    rule = active_prev_targets.clone()
    rule[:, 2] = active_prev_targets[:, 2] - 0.01 * state["obs_dict"]["hand_pose"][:, 2]
    return rule


def deadband_filter(active_prev_targets, active_rule_targets, active_targets):
    """Changes below 2e-4 are dropped."""
    return torch.where((active_targets - active_prev_targets).abs() < 2e-4, active_prev_targets, active_targets)


def extra_reward_terms(obs_dict):
    return {EXTRA_REWARD_NAME: torch.exp(-obs_dict["hand_pose"][:, 2].abs())}


HOOKS = {"action_rule": action_rule, "pre_action_rule": pre_action_rule, "filters": [(FILTER_NAME, deadband_filter)],
         "extra_reward_terms": extra_reward_terms, "extra_reward_weights": {EXTRA_REWARD_NAME: EXTRA_REWARD_WEIGHT}}
