"""The task plugin interface keeps the reference's surface: DexTask / BaseTask /
BlindGraspingTask expose every hook of dexhand_env/tasks/{task_interface,base_task,blind_grasping_task}.py with the same
parameter names, the same abstract set and the same property-vs-method kinds.  The pinned table is checked against the
reference's source where the reference exists (build container); on the GPU box only the table is used.  CPU only."""
import inspect
import os
import sys

import pytest

from dexrobot_isaac_b200 import tasks as T

# method -> parameter names (without self), pinned from the reference @ dc85a0b (tasks/task_interface.py)
PINNED_METHODS = {
    "register_task_state": ["name", "shape", "dtype"],
    "compute_task_reward_terms": ["obs_dict"],
    "check_task_success_criteria": ["obs_dict"],
    "check_task_failure_criteria": ["obs_dict"],
    "reset_task_state": ["env_ids"],
    "create_task_objects": ["gym", "sim", "env_ptr", "env_id"],
    "load_task_assets": [],
    "get_task_observations": ["obs_dict"],
    "get_task_dof_targets": ["num_envs", "device", "base_controlled", "fingers_controlled"],
    "set_tensor_references": ["root_state_tensor"],
    "initialize_task_states": [],
    "finalize_setup": [],
    "register_custom_filters": ["action_processor"],
}
PINNED_PROPERTIES = ["task_states", "observation_encoder", "pre_action_rule", "action_rule", "post_action_filters"]
PINNED_ABSTRACT = {"compute_task_reward_terms", "check_task_success_criteria", "check_task_failure_criteria",
                   "reset_task_state", "create_task_objects", "load_task_assets"}
PINNED_CTORS = {"BaseTask": ["sim", "gym", "device", "num_envs", "cfg"],
                "BlindGraspingTask": ["sim", "gym", "device", "num_envs", "full_config"]}


def _params(fn):
    return [p for p in inspect.signature(fn).parameters if p != "self"]


@pytest.mark.parametrize("cls", ["DexTask", "BaseTask", "BlindGraspingTask"])
def test_task_hooks_match_pinned(cls):
    c = getattr(T, cls)
    for meth, want in PINNED_METHODS.items():
        fn = getattr(c, meth)
        assert not isinstance(inspect.getattr_static(c, meth), property), f"{cls}.{meth} must be a method"
        assert _params(fn) == want, f"{cls}.{meth}: {_params(fn)} vs reference {want}"
    for prop in PINNED_PROPERTIES:
        assert isinstance(inspect.getattr_static(c, prop), property), f"{cls}.{prop} must be a property"
    if cls == "DexTask":
        assert set(c.__abstractmethods__) == PINNED_ABSTRACT
    else:
        assert not c.__abstractmethods__
        assert _params(c.__init__) == PINNED_CTORS[cls]


def test_default_hooks_are_the_reference_defaults():
    cfg = {"env": {}, "task": {}}
    t = T.BaseTask(None, None, "cpu", 4, cfg)
    assert t.pre_action_rule is None and t.action_rule is None and t.post_action_filters == []
    assert t.compute_task_reward_terms({}) == {} and t.check_task_success_criteria({}) == {}
    assert t.check_task_failure_criteria({}) == {} and t.get_task_observations({}) is None
    assert t.get_task_dof_targets(4, "cpu", True, True) is None
    assert not t.uses_custom_action_hooks()
    with pytest.raises(RuntimeError):  # task_interface.py:23-29: fail fast without a parent
        t.task_states
    with pytest.raises(TypeError):
        T.DexTask()  # abstract


REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present (GPU box)")
def test_pinned_table_matches_reference_source():
    stubs = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "stubs")
    for p in (stubs, REF):
        if p not in sys.path:
            sys.path.insert(0, p)
    from dexhand_env.tasks.base_task import BaseTask
    from dexhand_env.tasks.blind_grasping_task import BlindGraspingTask
    from dexhand_env.tasks.task_interface import DexTask

    for meth, want in PINNED_METHODS.items():
        assert _params(getattr(DexTask, meth)) == want, meth
    public = {n for n, v in vars(DexTask).items() if not n.startswith("_")}
    assert public == set(PINNED_METHODS) | set(PINNED_PROPERTIES), public ^ (set(PINNED_METHODS) | set(PINNED_PROPERTIES))
    for prop in PINNED_PROPERTIES:
        assert isinstance(inspect.getattr_static(DexTask, prop), property), prop
    assert set(DexTask.__abstractmethods__) == PINNED_ABSTRACT
    assert _params(BaseTask.__init__) == PINNED_CTORS["BaseTask"]
    assert _params(BlindGraspingTask.__init__) == PINNED_CTORS["BlindGraspingTask"]
    # every public hook the reference's shipped tasks define exists on the facades
    for ref_cls, own in ((BaseTask, T.BaseTask), (BlindGraspingTask, T.BlindGraspingTask)):
        for n, v in vars(ref_cls).items():
            if n.startswith("_") or not callable(v) and not isinstance(v, property):
                continue
            assert hasattr(own, n), f"{own.__name__} lacks {n}"
