"""Which absolute tolerance does each snapshot key need?  Runs the CUDA path against the oracle on the scripted world and
prints, per float key, max over steps/elements of (|cuda - oracle| - rtol * |oracle|), i.e. the smallest atol that would
pass, plus the magnitude range of the key.  Used to set tests' per-key absolute tolerances.
usage: python tools/debug/atol_sweep.py [task] [n] [steps]"""
import sys

import torch

sys.path.insert(0, ".")
from oracle import synth  # noqa: E402
from oracle.driver import run_lockstep  # noqa: E402
from oracle.dexstep_oracle import OracleEnv  # noqa: E402
from dexrobot_isaac_b200.config import load_task_config  # noqa: E402
from dexrobot_isaac_b200.env import DexStepEnv  # noqa: E402
from dexrobot_isaac_b200.params import BodyLayout  # noqa: E402

task = sys.argv[1] if len(sys.argv) > 1 else "BlindGrasping"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 256
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 520
RTOL = 1e-5
layout = synth.BodyLayout(with_box=task == "BlindGrasping")
cfg = load_task_config(task, num_envs=n)
ora = OracleEnv(cfg, n, layout, synth.dof_props_tensor())
cud = DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device="cuda:0",
                 materialize_obs_dict=True)
world = synth.ScriptedWorld(n, layout, seed=777, nact=ora.num_actions)
need, mags = {}, {}


def on_step(t, envs):
    a, b = envs[0].snapshot(), envs[1].snapshot()
    keep = ~a["reset_buf"].bool()
    for k, x in a.items():
        if k not in b or not x.dtype.is_floating_point:
            continue
        y = b[k].cpu().to(x.dtype).reshape(x.shape)
        if k.startswith("obs/") and x.dim() >= 1 and x.shape[0] == n:
            x, y = x[keep], y[keep]
        if x.numel() == 0:
            continue
        excess = ((x - y).abs() - RTOL * x.abs()).max().item()
        need[k] = max(need.get(k, -1.0), excess)
        mags[k] = max(mags.get(k, 0.0), x.abs().max().item())


run_lockstep([ora, cud], world, steps, on_step)
for k in sorted(need, key=lambda k: -need[k]):
    print(f"{k:45s} needs atol {max(need[k], 0.0):.3e}   max|ref| {mags[k]:.3e}")
