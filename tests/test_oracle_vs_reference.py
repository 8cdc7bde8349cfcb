"""PINS the CPU oracle: the restatement in oracle/dexstep_oracle.py is run in lock-step with the
reference's own components (imported unmodified from /root/reference by oracle/ref_harness.py)
over scripted 520-step trajectories.  Everything the parity gates name must be BIT-EXACT here
(same torch CPU ops in the same order), floats included.  Skipped where the reference is absent
(the GPU box); there the committed golden fixtures take over (test_oracle_golden.py)."""
import pytest
import torch

from oracle import synth
from oracle.driver import compare_snapshots, run_lockstep
from oracle.ref_harness import RefEnv, load_reference_cfg, reference_available

pytestmark = pytest.mark.skipif(not reference_available(), reason="/root/reference not present")

CASES = [
    ("BaseTask", 48, 300, {}),
    ("BlindGrasping", 48, 520, {}),
    ("BaseTask", 16, 40, dict(control_mode="position")),
    ("BlindGrasping", 16, 240, dict(control_mode="position")),
    ("BlindGrasping", 16, 240, dict(policy_controls_hand_base=False)),
    ("BaseTask", 16, 40, dict(policy_controls_fingers=False)),
]


def make_pair(task, n, kw, nb_hand=31):
    from oracle.dexstep_oracle import OracleEnv

    layout = synth.BodyLayout(nb_hand=nb_hand, with_box=task == "BlindGrasping")
    ref = RefEnv(task, n, layout=layout, **kw)
    cfg = load_reference_cfg(task, n)
    if "control_mode" in kw:
        cfg["task"]["controlMode"] = kw["control_mode"]
    cfg["task"]["policy_controls_hand_base"] = kw.get("policy_controls_hand_base", True)
    cfg["task"]["policy_controls_fingers"] = kw.get("policy_controls_fingers", True)
    ora = OracleEnv(cfg, n, layout, synth.dof_props_tensor())
    return ref, ora


@pytest.mark.parametrize("task,n,steps,kw", CASES)
def test_oracle_bit_exact_vs_reference(task, n, steps, kw):
    ref, ora = make_pair(task, n, kw)
    assert ref.num_observations == ora.obs_buf.shape[1] if hasattr(ora, "obs_buf") else True
    world = synth.ScriptedWorld(n, ref.layout, seed=101, nact=ref.num_actions)
    bad = []
    seen = {"reset": 0, "success": 0, "failure": 0, "timeout": 0}

    def on_step(t, envs):
        a, b = envs[0].snapshot(), envs[1].snapshot()
        for m in compare_snapshots(a, b, rtol=0.0, atol=0.0):
            if len(bad) < 10:
                bad.append(f"step {t}: {m}")
        seen["reset"] += int(a["reset_buf"].sum())
        for k in ("success", "failure", "timeout"):
            seen[k] += int(a[f"extras/{k}"].sum())

    run_lockstep([ref, ora], world, steps, on_step)
    assert not bad, "\n".join(bad)
    assert seen["reset"] > 0
    if task == "BlindGrasping" and steps >= 520 and not kw:
        assert seen["success"] > 0 and seen["failure"] > 0 and seen["timeout"] > 0, seen


def test_oracle_other_body_count():
    """NB is a runtime parameter (hand MJCF missing): same parity at NB_hand = 36."""
    ref, ora = make_pair("BlindGrasping", 8, {}, nb_hand=36)
    world = synth.ScriptedWorld(8, ref.layout, seed=5, nact=18)
    bad = []

    def on_step(t, envs):
        bad.extend(compare_snapshots(envs[0].snapshot(), envs[1].snapshot(), 0.0, 0.0))

    run_lockstep([ref, ora], world, 30, on_step)
    assert not bad, "\n".join(bad[:10])


def test_reference_dims_and_constants():
    """Probe-confirmed facts of SURVEY §8(a)/App. B."""
    ref, ora = make_pair("BaseTask", 4, {})
    assert ref.num_observations == 224 and abs(ref.control_dt - 0.01) < 1e-12
    assert ref.observation_encoder.active_finger_dof_indices.tolist() == [8, 7, 6, 10, 12, 11, 16, 15, 20, 19, 24, 23]
    assert ora.act_idx.tolist() == [8, 7, 6, 10, 12, 11, 16, 15, 20, 19, 24, 23]
    assert torch.equal(ref.action_processor.coupling_indices[:, 1], ora.cpl_d)
    assert torch.equal(ref.action_processor.max_deltas, ora.max_deltas)
    ref, ora = make_pair("BlindGrasping", 4, {})
    assert ref.num_observations == 158 and abs(ref.control_dt - 0.02) < 1e-12
    assert ref.task.contact_duration_threshold_steps == 100 == ora.contact_duration_threshold_steps


def test_oracle_with_custom_task_hooks_vs_reference():
    """the toy third task (custom action rule, pre-action rule, post-action filter, extra reward term;
    oracle/toy_hooks.py) through the UNMODIFIED reference components (ActionRules / StepProcessor / RewardCalculator run
    the callables) against the oracle port running the same callables - bit-exact over 220 steps."""
    from oracle import toy_hooks
    from oracle.dexstep_oracle import OracleEnv

    n, steps = 24, 220
    layout = synth.BodyLayout(with_box=False)
    ref = RefEnv("BaseTask", n, layout=layout, hooks=toy_hooks.HOOKS)
    cfg = load_reference_cfg("BaseTask", n)
    cfg["task"]["reward_weights"][toy_hooks.EXTRA_REWARD_NAME] = toy_hooks.EXTRA_REWARD_WEIGHT
    ora = OracleEnv(cfg, n, layout, synth.dof_props_tensor())
    ora.hooks = toy_hooks.HOOKS
    world = synth.ScriptedWorld(n, layout, seed=131, nact=ref.num_actions)
    bad = []

    def on_step(t, envs):
        a, b = envs[0].snapshot(), envs[1].snapshot()
        for m in compare_snapshots(a, b, rtol=0.0, atol=0.0):
            if len(bad) < 10:
                bad.append(f"step {t}: {m}")
        if t > 2:
            assert f"rc/{toy_hooks.EXTRA_REWARD_NAME}_weighted" in a

    run_lockstep([ref, ora], world, steps, on_step)
    assert not bad, "\n".join(bad)


@pytest.mark.parametrize("task,n,steps", [("BaseTask", 48, 300), ("BlindGrasping", 48, 520)])
def test_oracle_bit_exact_vs_reference_on_threshold_ties(task, n, steps):
    """thresholds hit EXACTLY (synth.TieWorld: forces of exactly f32(0.1) / f32(contact threshold),
    a pad exactly at the proximity threshold, box_z == f32(0.2), tip / hand height on the height-safety thresholds, hand at
    exactly max_box_distance, box speed == velocity tolerance; synth.tie_config: a dyadic control_dt so time_in_stage /
    stage_contact_duration / grasp duration land exactly on 200 / 150 / 25 / 100 steps' worth of seconds).  The unmodified
    reference components and the oracle port must agree bit for bit - strict vs non-strict comparisons and f32-vs-double
    thresholds are exactly what this distinguishes."""
    from oracle.dexstep_oracle import OracleEnv

    layout = synth.BodyLayout(with_box=task == "BlindGrasping")
    cfg_r, cfg_o = (synth.tie_config(load_reference_cfg(task, n)) for _ in range(2))
    ref = RefEnv(task, n, layout=layout, cfg=cfg_r)
    ora = OracleEnv(cfg_o, n, layout, synth.dof_props_tensor())
    assert ref.control_dt == 2.0 ** -6 == ora.control_dt
    world = synth.TieWorld(n, layout, cfg_o, seed=141, nact=ref.num_actions)
    bad, seen = [], {"reset": 0, "success": 0, "failure": 0, "timeout": 0, "tis_tie": 0}

    def on_step(t, envs):
        a, b = envs[0].snapshot(), envs[1].snapshot()
        for m in compare_snapshots(a, b, rtol=0.0, atol=0.0):
            if len(bad) < 10:
                bad.append(f"step {t}: {m}")
        seen["reset"] += int(a["reset_buf"].sum())
        for k in ("success", "failure", "timeout"):
            seen[k] += int(a[f"extras/{k}"].sum())
        if task == "BlindGrasping":  # a stage timer that sits exactly on its duration was seen (and acted upon)
            tis = a["task/time_in_stage"]
            seen["tis_tie"] += int(((tis == 200 / 64) | (tis == 150 / 64)).sum())

    run_lockstep([ref, ora], world, steps, on_step)
    assert not bad, "\n".join(bad)
    assert all(v > 0 for v in world.hits.values()), world.hits
    assert seen["reset"] > 0 and seen["failure"] > 0
    if task == "BlindGrasping":
        assert seen["timeout"] + seen["success"] > 0 and seen["tis_tie"] > 0, seen


def test_oracle_vs_reference_other_physics_step_count():
    """control_dt is MEASURED by the reference (physics steps per control step x sim.dt,
    physics_manager.py:243-259) and feeds max_deltas, the stage timers, contact_duration_threshold_steps and the finite
    differences.  A count of 3 instead of the stock 2: the port follows."""
    from oracle.dexstep_oracle import OracleEnv

    n, layout = 24, synth.BodyLayout(with_box=True)
    ref = RefEnv("BlindGrasping", n, layout=layout, physics_steps_per_control_step=3)
    ora = OracleEnv(load_reference_cfg("BlindGrasping", n), n, layout, synth.dof_props_tensor(),
                    physics_steps_per_control_step=3)
    assert abs(ref.control_dt - 0.03) < 1e-12 and ref.control_dt == ora.control_dt
    assert ref.task.contact_duration_threshold_steps == ora.contact_duration_threshold_steps == 66
    world = synth.ScriptedWorld(n, layout, seed=161, nact=ref.num_actions)
    bad = []

    def on_step(t, envs):
        for m in compare_snapshots(envs[0].snapshot(), envs[1].snapshot(), rtol=0.0, atol=0.0):
            if len(bad) < 10:
                bad.append(f"step {t}: {m}")

    run_lockstep([ref, ora], world, 380, on_step)
    assert not bad, "\n".join(bad)
