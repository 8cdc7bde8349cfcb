"""GPU lockstep at the thresholds: oracle/synth.TieWorld puts forces, distances, heights and speeds
EXACTLY on the float32 thresholds of the path and synth.tie_config makes the stage timers land exactly on their durations
(dyadic control_dt).  Masks, counters and FSM state must stay bit-exact against the oracle - which is itself pinned
bit-exact against the unmodified reference on the same world (tests/test_oracle_vs_reference.py::*threshold_ties).
Both kernel families are covered: the persistent kernel (lean configuration) and the tile kernel (obs_dict materialised)."""
import pytest
import torch

from oracle import synth
from oracle.driver import compare_snapshots, penetration_floors, run_lockstep
from dexrobot_isaac_b200.config import load_task_config

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL, ATOL = 1e-5, 1e-7


@pytest.mark.parametrize("obs_dict", [False, True], ids=["persistent", "tile+obs_dict"])
@pytest.mark.parametrize("task,n,steps", [("BaseTask", 192, 300), ("BlindGrasping", 192, 520)])
def test_threshold_ties_lockstep(task, n, steps, obs_dict):
    from oracle.dexstep_oracle import OracleEnv
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    layout = synth.BodyLayout(with_box=task == "BlindGrasping")
    cfg = synth.tie_config(load_task_config(task, num_envs=n))
    ora = OracleEnv(cfg, n, layout, synth.dof_props_tensor())
    cud = DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device=DEV,
                     materialize_obs_dict=obs_dict)
    assert cud.control_dt == 2.0 ** -6 == ora.control_dt
    world = synth.TieWorld(n, layout, cfg, seed=151, nact=ora.num_actions)
    floors = penetration_floors(cfg)
    bad, seen = [], {"reset": 0, "success": 0, "failure": 0, "timeout": 0, "tis_tie": 0}

    def on_step(t, envs):
        a, b = envs[0].snapshot(), envs[1].snapshot()
        seen["reset"] += int(a["reset_buf"].sum())
        for k in ("success", "failure", "timeout"):
            seen[k] += int(a[f"extras/{k}"].sum())
        if task == "BlindGrasping":
            tis = a["task/time_in_stage"]
            seen["tis_tie"] += int(((tis == 200 / 64) | (tis == 150 / 64)).sum())
        ign = () if obs_dict else ("obs/",)
        for m in compare_snapshots(a, b, RTOL, ATOL, skip_rows=a["reset_buf"].bool(), ignore=ign, floors=floors):
            if len(bad) < 20:
                bad.append(f"step {t} {m}")

    run_lockstep([ora, cud], world, steps, on_step)
    assert not bad, "\n".join(bad)
    assert all(v > 0 for v in world.hits.values()), world.hits
    assert seen["reset"] > 0 and seen["failure"] > 0
    if task == "BlindGrasping":
        assert seen["tis_tie"] > 0 and seen["timeout"] + seen["success"] > 0, seen
