"""Long lockstep soak of the lean CUDA path against the oracle (not part of the test suite: minutes, not seconds).
    python tools/debug/soak_lockstep.py [task] [n] [steps] [single]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from oracle import synth  # noqa: E402
from oracle.dexstep_oracle import OracleEnv  # noqa: E402
from oracle.driver import compare_snapshots, penetration_floors, run_lockstep  # noqa: E402
from dexrobot_isaac_b200.config import load_task_config  # noqa: E402
from dexrobot_isaac_b200.env import DexStepEnv  # noqa: E402
from dexrobot_isaac_b200.params import BodyLayout  # noqa: E402

task = sys.argv[1] if len(sys.argv) > 1 else "BlindGrasping"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 256
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3000
single = len(sys.argv) > 4 and sys.argv[4] == "single"
lay = synth.BodyLayout(with_box=task == "BlindGrasping")
cfg = load_task_config(task, num_envs=n)
ora = OracleEnv(cfg, n, lay, synth.dof_props_tensor())
cud = DexStepEnv(cfg, n, BodyLayout.from_synth(lay), synth.dof_props_tensor(), device="cuda:0", materialize_obs_dict=False)
cud.pipeline.single_pass = single
world = synth.ScriptedWorld(n, lay, seed=4242, nact=ora.num_actions)
floors = penetration_floors(cfg)
bad, seen = [], {"reset": 0, "success": 0, "failure": 0, "timeout": 0}
t0 = time.time()


def on_step(t, envs):
    a, b = envs[0].snapshot(), envs[1].snapshot()
    seen["reset"] += int(a["reset_buf"].sum())
    for k in ("success", "failure", "timeout"):
        seen[k] += int(a[f"extras/{k}"].sum())
    for m in compare_snapshots(a, b, 1e-5, 1e-7, skip_rows=a["reset_buf"].bool(), ignore=("obs/",), floors=floors):
        if len(bad) < 20:
            bad.append(f"step {t} {m}")


run_lockstep([ora, cud], world, steps, on_step)
print(f"{task} n={n} steps={steps} kernel={cud.pipeline.lib.dex_debug_last_post_kernel()} {seen} "
      f"mismatches={len(bad)} ({time.time() - t0:.0f} s)")
for m in bad:
    print("  ", m)
sys.exit(1 if bad else 0)
