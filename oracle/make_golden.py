"""Generates tests/golden/*.npz by running the REFERENCE's own components (ref_harness.RefEnv)
on the seeded scripted world.  Run in the build container only:

    python -m oracle.make_golden

Each fixture holds, for N=32 envs over 520 steps: per-step compact records (reset mask, step
count, stage, termination classes, rewards, float64 row-sums of obs_buf and full_dof_targets)
and full snapshots at selected steps.  Inputs are not stored: they are regenerated from the seed
by oracle/synth.py (same image on both machines); a per-step input checksum is stored to detect
generator drift.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

from . import synth
from .driver import run_lockstep
from .ref_harness import RefEnv, load_reference_cfg

FULL_STEPS = (0, 1, 2, 3, 150, 201, 202, 203, 228, 232, 300, 352, 400, 499, 500, 520)
GOLDEN = {
    "basetask_n32": dict(task="BaseTask", n=32, seed=11, steps=520, kw={}),
    "blindgrasping_n32": dict(task="BlindGrasping", n=32, seed=12, steps=520, kw={}),
    "basetask_position_n32": dict(task="BaseTask", n=32, seed=13, steps=64, kw=dict(control_mode="position")),
    "blindgrasping_nobase_n32": dict(task="BlindGrasping", n=32, seed=14, steps=260,
                                     kw=dict(policy_controls_hand_base=False)),
    # every element of obs_buf / full_dof_targets / active_prev_targets of EVERY step (not row sums)
    "blindgrasping_elementwise_n12": dict(task="BlindGrasping", n=12, seed=15, steps=520, kw={}, elementwise=True),
    "basetask_elementwise_n8": dict(task="BaseTask", n=8, seed=16, steps=310, kw={}, elementwise=True),
    # thresholds hit exactly (synth.TieWorld + synth.tie_config), the reference's own answers at the ties
    "blindgrasping_ties_n32": dict(task="BlindGrasping", n=32, seed=17, steps=520, kw={}, ties=True),
    "basetask_ties_n32": dict(task="BaseTask", n=32, seed=18, steps=310, kw={}, ties=True),
}
ELEMENTWISE_KEYS = ("obs_buf", "full_dof_targets", "active_prev_targets")


def input_checksum(world, t, ep_step, box0):
    st = world.state(t, ep_step, box0)
    return float(sum(v.double().sum() for v in st.values()) + world.actions(t).double().sum()
                 + world.uniforms(t).double().sum())


def record(name, spec, out_dir):
    ties, elementwise = bool(spec.get("ties")), bool(spec.get("elementwise"))
    if ties:
        cfg = synth.tie_config(load_reference_cfg(spec["task"], spec["n"]))
        env = RefEnv(spec["task"], spec["n"], cfg=cfg, **spec["kw"])
        world = synth.TieWorld(spec["n"], env.layout, cfg, seed=spec["seed"], nact=env.num_actions)
    else:
        env = RefEnv(spec["task"], spec["n"], **spec["kw"])
        world = synth.ScriptedWorld(spec["n"], env.layout, seed=spec["seed"], nact=env.num_actions)
    rec = {}
    elem = {k: [] for k in ELEMENTWISE_KEYS} if elementwise else {}
    per_step = {k: [] for k in ("reset_buf", "episode_step_count", "rew_buf", "obs_rowsum", "tgt_rowsum",
                                "success", "failure", "timeout", "stage", "success_duration_steps")}

    def on_step(t, envs):
        s = envs[0].snapshot()
        per_step["reset_buf"].append(s["reset_buf"].numpy().copy())
        per_step["episode_step_count"].append(s["episode_step_count"].numpy().astype(np.int16))
        per_step["rew_buf"].append(s["rew_buf"].numpy().copy())
        per_step["obs_rowsum"].append(s["obs_buf"].double().sum(dim=1).numpy())
        per_step["tgt_rowsum"].append(s["full_dof_targets"].double().sum(dim=1).numpy())
        for k in ("success", "failure", "timeout"):
            per_step[k].append(s[f"extras/{k}"].numpy().copy())
        if "task/current_stage" in s:
            per_step["stage"].append(s["task/current_stage"].numpy().astype(np.int8))
            per_step["success_duration_steps"].append(s["task/success_duration_steps"].numpy().astype(np.int16))
        for k in elem:
            elem[k].append(s[k].numpy().copy())
        if t in FULL_STEPS and t <= spec["steps"]:
            for k, v in s.items():
                if v is not None:
                    rec[f"full/{t}/{k}"] = v.numpy().copy()

    run_lockstep([env], world, spec["steps"], on_step)
    for k, v in per_step.items():
        if v:
            rec[f"steps/{k}"] = np.stack(v)
    for k, v in elem.items():
        rec[f"elem/{k}"] = np.stack(v)
    rec["meta/ties"] = np.array(ties)
    rec["meta/task"] = np.array(spec["task"])
    rec["meta/n"] = np.array(spec["n"])
    rec["meta/seed"] = np.array(spec["seed"])
    rec["meta/steps"] = np.array(spec["steps"])
    rec["meta/control_mode"] = np.array(env.task_cfg["controlMode"])
    rec["meta/ctrl_base"] = np.array(bool(env.task_cfg["policy_controls_hand_base"]))
    rec["meta/ctrl_fingers"] = np.array(bool(env.task_cfg["policy_controls_fingers"]))
    path = os.path.join(out_dir, f"{name}.npz")
    np.savez_compressed(path, **rec)
    print(f"wrote {path}: {os.path.getsize(path) / 1e6:.2f} MB, "
          f"resets={int(rec['steps/reset_buf'].sum())} success={int(rec['steps/success'].sum())} "
          f"failure={int(rec['steps/failure'].sum())} timeout={int(rec['steps/timeout'].sum())}")


def main():
    """python -m oracle.make_golden [name ...]   (no names = all fixtures)"""
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, spec in GOLDEN.items():
        if len(sys.argv) > 1 and name not in sys.argv[1:]:
            continue
        record(name, spec, out_dir)


if __name__ == "__main__":
    sys.exit(main())
