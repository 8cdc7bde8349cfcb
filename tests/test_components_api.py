"""The reference-named component facades keep the reference's public surface: every public
method / constructor of the reference classes on the hot path exists with the same parameter names.  The
reference itself is only importable in the build container (/root/reference); elsewhere the pinned list is used."""
import inspect
import os
import sys

import pytest

from dexrobot_isaac_b200 import components as C

# (class, method) -> parameter names (without self), pinned from the reference @ dc85a0b
PINNED = {
    ("ActionProcessor", "__init__"): ["parent"],
    ("ActionProcessor", "initialize_from_config"): ["config"],
    ("ActionProcessor", "finalize_setup"): [],
    ("ActionProcessor", "apply_pre_action_rule"): ["active_prev_targets", "state"],
    ("ActionProcessor", "process_actions"): ["actions", "active_rule_targets"],
    ("ActionProcessor", "reset_targets"): ["env_ids", "dof_positions"],
    ("ActionProcessor", "apply_coupling"): ["active_targets"],
    ("ActionProcessor", "extract_active_targets_from_full_dof"): ["full_dof_positions"],
    ("ActionProcessor", "set_pre_action_rule"): ["rule"],
    ("ActionProcessor", "set_action_rule"): ["rule"],
    ("ActionProcessor", "set_coupling_rule"): ["rule"],
    ("ActionProcessor", "register_post_action_filter"): ["name", "filter_fn"],
    ("ActionProcessor", "unscale_actions"): ["actions"],
    ("ObservationEncoder", "__init__"): ["parent"],
    ("ObservationEncoder", "initialize"): ["observation_keys", "joint_to_control", "active_joint_names", "num_actions",
                                           "action_processor", "index_mappings"],
    ("ObservationEncoder", "update_prev_actions"): ["actions"],
    ("ObservationEncoder", "reset_observer_state"): ["env_ids"],
    ("ObservationEncoder", "register_task_state"): ["name", "shape", "dtype"],
    ("ObservationEncoder", "get_task_state"): ["name"],
    ("ObservationEncoder", "set_task_state"): ["name", "value"],
    ("ObservationEncoder", "compute_observations"): ["exclude_components"],
    ("ObservationEncoder", "concatenate_observations"): ["obs_dict"],
    ("ObservationEncoder", "parse_obs_tensor"): ["obs_tensor", "obs_dict", "env_idx"],
    ("RewardCalculator", "__init__"): ["parent", "task_cfg"],
    ("RewardCalculator", "compute_common_reward_terms"): ["obs_dict", "parent_env"],
    ("RewardCalculator", "compute_total_reward"): ["common_rewards", "task_rewards", "success_failure_rewards"],
    ("TerminationManager", "__init__"): ["parent", "task_cfg"],
    ("TerminationManager", "evaluate"): ["episode_step_count", "builtin_success", "task_success", "builtin_failure",
                                         "task_failure"],
    ("TerminationManager", "update_consecutive_successes"): ["success_tensor"],
    ("TerminationManager", "reset_tracking"): ["env_ids"],
    ("TerminationManager", "validate_criteria"): ["builtin_dict", "task_dict", "active_list", "criteria_type"],
    ("ResetManager", "__init__"): ["parent", "dof_state", "root_state_tensor", "hand_local_actor_index",
                                   "default_dof_pos", "task"],
    ("ResetManager", "reset_idx"): ["env_ids"],
    ("ResetManager", "reset_all"): [],
    ("StepProcessor", "__init__"): ["parent"],
    ("StepProcessor", "process_physics_step"): [],
}


def _params(fn):
    return [p for p in inspect.signature(fn).parameters if p != "self"]


@pytest.mark.parametrize("cls_name,meth", sorted(PINNED))
def test_signature_matches_pinned(cls_name, meth):
    fn = getattr(getattr(C, cls_name), meth)
    have = _params(fn)
    want = PINNED[(cls_name, meth)]
    assert have[:len(want)] == want, f"{cls_name}.{meth}: {have} vs reference {want}"
    # extra parameters of the facade must be optional
    sig = inspect.signature(fn)
    for name in have[len(want):]:
        assert sig.parameters[name].default is not inspect.Parameter.empty


REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present (GPU box)")
def test_pinned_list_matches_reference_source():
    """The pinned table above is what the reference really declares (checked where the reference exists)."""
    stubs = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "stubs")
    for p in (stubs, REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    from dexhand_env.components.action.action_processor import ActionProcessor
    from dexhand_env.components.observation.observation_encoder import ObservationEncoder
    from dexhand_env.components.reset.reset_manager import ResetManager
    from dexhand_env.components.reward.reward_calculator import RewardCalculator
    from dexhand_env.components.step_processor import StepProcessor
    from dexhand_env.components.termination.termination_manager import TerminationManager

    ref = {c.__name__: c for c in (ActionProcessor, ObservationEncoder, ResetManager, RewardCalculator, StepProcessor,
                                   TerminationManager)}
    for (cls_name, meth), want in PINNED.items():
        fn = getattr(ref[cls_name], meth)
        fn = getattr(fn, "__wrapped__", fn)
        assert _params(fn) == want, f"reference {cls_name}.{meth}: {_params(fn)}"


def test_no_oracle_import_in_product():
    """The product package never imports the oracle (parity infrastructure)."""
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dexrobot_isaac_b200")
    for dp, _, files in os.walk(root):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
