"""Debug: lockstep of a CUDA env (v2 or tile kernel) against the oracle; print the first mismatching step in detail."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from oracle import synth
from oracle.driver import compare_snapshots, run_lockstep
from tests.test_cuda_v2 import lean_env, oracle_env

task, n, steps, seed, tile = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
lay = synth.BodyLayout(with_box=task == "BlindGrasping")
ora, cud = oracle_env(task, n, lay), lean_env(task, n, lay)
cud.pipeline.tile_kernel = bool(tile)
world = synth.ScriptedWorld(n, lay, seed=seed, nact=ora.num_actions)
found = []
prev = {}
def on_step(t, envs):
    if found: return
    a, b = envs[0].snapshot(), envs[1].snapshot()
    bad = compare_snapshots(a, b, 1e-5, 1e-6, skip_rows=a["reset_buf"].bool(), ignore=("obs/", "rew_buf", "rc/"))
    if bad:
        found.append(t)
        print("step", t, bad[:6])
        for k in ("active_prev_targets", "full_dof_targets"):
            x, y = a[k].cpu(), b[k].cpu()
            d = ((x - y).abs() > 1e-5 + 1e-5 * x.abs())
            envs_bad = torch.nonzero(d.any(dim=1)).flatten()
            print(k, "bad envs", envs_bad[:10].tolist(), "of", int(envs_bad.numel()))
            for e in envs_bad[:3].tolist():
                print("  env", e, "cols", torch.nonzero(d[e]).flatten().tolist())
                print("   ref", x[e].tolist()); print("   got", y[e].tolist())
                print("   reset now", bool(a["reset_buf"][e]), "prev reset", prev.get("reset", torch.zeros(n, dtype=torch.bool))[e].item(),
                      "step count", int(a["episode_step_count"][e]), "ref prev apt", prev["apt"][e].tolist() if "apt" in prev else None)
    prev["reset"] = a["reset_buf"].clone().cpu(); prev["apt"] = a["active_prev_targets"].clone().cpu()
run_lockstep([ora, cud], world, steps, on_step)
print("done; first mismatch:", found)
