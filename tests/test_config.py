"""Host logic: the shipped task configs equal the reference's on every key the path reads."""
import pytest

from dexrobot_isaac_b200.config import deep_merge, load_task_config


def _subset_equal(mine, ref, path=""):
    bad = []
    for k, v in mine.items():
        if k == "numEnvs":
            continue
        if k not in ref:
            bad.append(f"{path}{k}: absent in reference")
        elif isinstance(v, dict):
            bad += _subset_equal(v, ref[k], f"{path}{k}.")
        elif v != ref[k]:
            bad.append(f"{path}{k}: {v!r} != {ref[k]!r}")
    return bad


@pytest.mark.parametrize("task", ["BaseTask", "BlindGrasping"])
def test_shipped_config_matches_reference(task):
    from oracle.ref_harness import load_reference_cfg, reference_available

    if not reference_available():
        pytest.skip("/root/reference not present")
    mine, ref = load_task_config(task, 64), load_reference_cfg(task, 64)
    assert not _subset_equal(mine, ref)
    assert mine["task"]["reward_weights"] == ref["task"]["reward_weights"]
    assert list(mine["task"]["reward_weights"]) == list(ref["task"]["reward_weights"])  # dict order matters
    assert mine["task"]["policy_observation_keys"] == ref["task"]["policy_observation_keys"]


def test_delete_semantics():
    out = deep_merge({"a": {"x": 1, "y": 2}, "b": 1}, {"a": {"_delete_": True, "z": 3}})
    assert out == {"a": {"z": 3}, "b": 1}
    out = deep_merge({"a": {"x": 1, "y": 2}}, {"a": {"y": 5}})
    assert out == {"a": {"x": 1, "y": 5}}


def test_overrides_and_errors():
    cfg = load_task_config("BlindGrasping", 8, control_mode="position", policy_controls_hand_base=False)
    assert cfg["task"]["controlMode"] == "position" and cfg["task"]["policy_controls_hand_base"] is False
    assert cfg["env"]["episodeLength"] == 500 and cfg["sim"]["dt"] == 0.01
    assert "alive" in cfg["task"]["reward_weights"] and cfg["task"]["reward_weights"]["alive"] == 0.0
    with pytest.raises(ValueError):
        load_task_config("BaseTask", 8, control_mode="torque")
    with pytest.raises(ValueError):
        load_task_config("NoSuchTask", 8)
