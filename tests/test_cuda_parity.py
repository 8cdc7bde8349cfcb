"""GPU parity tests proper: the CUDA path (through the C ABI) against
  (1) the golden fixtures the REFERENCE produced (tests/golden/*.npz),
  (2) the CPU oracle in lockstep on the same seeded scripted world (N=64 over 520 steps, N=4096 shorter),
  (3) size-independent properties at BASELINE.json's full sizes (16 384 envs).
Tolerances (BASELINE.json north_star): bit-exact for masks / counters / indices, 1e-5 relative
(+1e-7 absolute near zero; tools/debug/atol_sweep.py shows what each key needs) for observations, rewards and DOF targets.
"""
import glob
import os

import pytest
import torch

from oracle import synth
from oracle.driver import check_against_golden, compare_snapshots, golden_cfg, penetration_floors, run_lockstep
from dexrobot_isaac_b200.config import load_task_config

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-5, 1e-7
GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def make_cuda_env(task, n, layout=None, materialize=True, **kw):
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    cfg = load_task_config(task, num_envs=n, **kw)
    layout = layout or synth.BodyLayout(with_box=task == "BlindGrasping")
    return DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device="cuda:0",
                      materialize_obs_dict=materialize)


def make_oracle_env(task, n, layout=None, **kw):
    from oracle.dexstep_oracle import OracleEnv

    cfg = load_task_config(task, num_envs=n, **kw)
    layout = layout or synth.BodyLayout(with_box=task == "BlindGrasping")
    return OracleEnv(cfg, n, layout, synth.dof_props_tensor())


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_cuda_matches_reference_golden(path):
    def make(meta):
        from dexrobot_isaac_b200.env import DexStepEnv
        from dexrobot_isaac_b200.params import BodyLayout

        cfg = golden_cfg(meta, load_task_config)  # incl. the dyadic-dt tie configuration of the *_ties_* fixtures
        layout = synth.BodyLayout(with_box=meta["task"] == "BlindGrasping")
        return DexStepEnv(cfg, meta["n"], BodyLayout.from_synth(layout), synth.dof_props_tensor(), device="cuda:0",
                          materialize_obs_dict=True)

    bad = check_against_golden(path, make, rtol=RTOL, atol=ATOL)
    assert not bad, "\n".join(bad)


def _lockstep(task, n, steps, seed, layout=None, check_every=1, **kw):
    layout = layout or synth.BodyLayout(with_box=task == "BlindGrasping")
    ora = make_oracle_env(task, n, layout, **kw)
    cud = make_cuda_env(task, n, layout, **kw)
    world = synth.ScriptedWorld(n, layout, seed=seed, nact=ora.num_actions)
    bad, seen = [], {"reset": 0, "success": 0, "failure": 0, "timeout": 0}
    floors = penetration_floors(load_task_config(task, num_envs=n, **kw))

    def on_step(t, envs):
        if len(bad) > 20 or (t % check_every and t > 3):
            return
        a, b = envs[0].snapshot(), envs[1].snapshot()
        seen["reset"] += int(a["reset_buf"].sum())
        for k in ("success", "failure", "timeout"):
            seen[k] += int(a[f"extras/{k}"].sum())
        for m in compare_snapshots(a, b, RTOL, ATOL, skip_rows=a["reset_buf"].bool(), floors=floors):
            bad.append(f"step {t} {m}")
        # the ordered id list of the envs that reset == torch.nonzero(reset_buf)
        ids = envs[1].pipeline.env_ids().cpu()
        if not torch.equal(ids, torch.nonzero(a["reset_buf"]).flatten()):
            bad.append(f"step {t} env_ids differ")

    run_lockstep([ora, cud], world, steps, on_step)
    assert not bad, "\n".join(bad)
    return seen


def test_lockstep_basetask_n64():
    seen = _lockstep("BaseTask", 64, 520, seed=101)
    assert seen["timeout"] > 0 and seen["failure"] > 0


def test_lockstep_blindgrasping_n64():
    seen = _lockstep("BlindGrasping", 64, 520, seed=102)
    assert seen["success"] > 0 and seen["failure"] > 0 and seen["timeout"] > 0


@pytest.mark.parametrize("kw", [dict(control_mode="position"), dict(policy_controls_hand_base=False),
                                dict(policy_controls_fingers=False)],
                         ids=["position", "nobase", "nofingers"])
def test_lockstep_action_variants(kw):
    _lockstep("BlindGrasping", 48, 230, seed=103, **kw)
    _lockstep("BaseTask", 33, 40, seed=104, **kw)


def test_lockstep_ragged_and_other_body_layout():
    # N not a multiple of the env tile / warp, and a 37-body hand (SURVEY §8d sensitivity layout)
    _lockstep("BaseTask", 1, 12, seed=105)
    _lockstep("BlindGrasping", 77, 215, seed=106, layout=synth.BodyLayout(nb_hand=37, with_box=True))


def test_lockstep_basetask_n4096():
    _lockstep("BaseTask", 4096, 40, seed=107, check_every=8)


def test_lockstep_blindgrasping_n4096():
    seen = _lockstep("BlindGrasping", 4096, 240, seed=108, check_every=40)
    assert seen["reset"] > 0


@pytest.mark.parametrize("materialize", [False, True], ids=["persistent", "tile+obs_dict"])
def test_lockstep_other_physics_step_count(materialize):
    """physics_steps_per_control_step (the reference measures it, physics_manager.py:243-259) is a constructor
    argument of DexStepEnv; 3 instead of the stock 2 changes control_dt -> max_deltas, stage timers, the contact-duration
    step threshold and every finite difference.  Oracle pinned for the same value by tests/test_oracle_vs_reference.py."""
    from oracle.dexstep_oracle import OracleEnv
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    n, layout = 96, synth.BodyLayout(with_box=True)
    cfg = load_task_config("BlindGrasping", num_envs=n)
    ora = OracleEnv(cfg, n, layout, synth.dof_props_tensor(), physics_steps_per_control_step=3)
    cud = DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device="cuda:0",
                     materialize_obs_dict=materialize, physics_steps_per_control_step=3)
    assert abs(cud.control_dt - 0.03) < 1e-12
    assert int(cud.pipeline.params.contact_duration_threshold_steps) == ora.contact_duration_threshold_steps == 66
    world = synth.ScriptedWorld(n, layout, seed=171, nact=ora.num_actions)
    floors = penetration_floors(cfg)
    bad = []

    def on_step(t, envs):
        a, b = envs[0].snapshot(), envs[1].snapshot()
        ign = () if materialize else ("obs/",)
        for m in compare_snapshots(a, b, RTOL, ATOL, skip_rows=a["reset_buf"].bool(), ignore=ign, floors=floors):
            if len(bad) < 20:
                bad.append(f"step {t} {m}")

    run_lockstep([ora, cud], world, 380, on_step)
    assert not bad, "\n".join(bad)
