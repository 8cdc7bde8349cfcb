"""Drives any implementation of the step pipeline (reference harness, CPU oracle, CUDA
product) through the same scripted trajectory and compares their snapshots.

TEST INFRASTRUCTURE ONLY.  The three implementations expose the same driver API:
``load_state(state)``, ``reset(uniforms)``, ``pre_physics_step(actions)``,
``post_physics_step(uniforms)``, ``snapshot()`` and the attributes ``episode_step_count``,
``num_actions``.
"""
from __future__ import annotations

from typing import Callable, Dict, Iterable, List, Optional, Tuple

import torch

from . import synth

# bit-exact keys (SURVEY §8d parity gates): masks, counters, indices
INT_KEYS_PREFIX = ("reset_buf", "episode_step_count", "extras/success", "extras/failure", "extras/timeout",
                   "extras/episode_length", "extras/consecutive_successes", "task/current_stage",
                   "task/success_duration_steps", "task/success_conditions_met",
                   "task/just_transitioned", "obs/contact_binary", "extras/success_reason",
                   "extras/failure_reason")


def box0_of(env):
    for name in ("initial_box_positions",):
        if hasattr(env, name):
            return getattr(env, name)
    task = getattr(env, "task", None)
    if task is not None and hasattr(task, "initial_box_positions"):
        return task.initial_box_positions
    return None


def run_lockstep(envs: List, world: synth.ScriptedWorld, steps: int,
                 on_step: Callable[[int, List], None], leader: int = 0):
    """Step all ``envs`` on the same inputs; the scripted world is driven by ``envs[leader]``'s
    episode_step_count / initial_box_positions (bit-exact state, identical on all sides when
    parity holds)."""
    lead = envs[leader]
    s0 = world.state(0, lead.episode_step_count, box0_of(lead))
    u0 = world.uniforms(0)
    for e in envs:
        e.load_state(s0)
        e.reset(u0)
    on_step(0, envs)
    for t in range(1, steps + 1):
        a = world.actions(t)
        for e in envs:
            e.pre_physics_step(a)
        st = world.state(t, lead.episode_step_count, box0_of(lead))
        u = world.uniforms(t)
        for e in envs:
            e.load_state(st)
            e.post_physics_step(u)
        on_step(t, envs)


def penetration_floors(cfg: dict) -> Dict[str, float]:
    """Magnitude floors for the one reward term that is a scaled DIFFERENCE of nearly equal numbers:
    penetration_penalty = max(threshold - min_tip_distance, 0) * depth_scale (blind_grasping_task.py:1194-1208), weighted by
    -100.  A 1-ulp difference in the distance (torch's CPU norm vs sqrtf of a plain sum of squares) is amplified by
    distance / (threshold - distance); the 1e-5 relative bound is therefore taken relative to threshold * scale (* |w|)."""
    try:
        box, pp = cfg["env"]["box"], cfg["task"]["penetration_prevention"]
        w = abs(float(cfg["task"]["reward_weights"].get("penetration_penalty", 0.0)))
        raw = float(box["size"]) / 2.0 * float(pp["geometricPenetrationFactor"]) * float(pp["penetrationDepthScale"])
        return {"rc/penetration_penalty": raw, "rc/penetration_penalty_weighted": raw * w}
    except (KeyError, TypeError):
        return {}


def compare_snapshots(ref: Dict[str, torch.Tensor], got: Dict[str, torch.Tensor], rtol: float,
                      atol: float, skip_rows: Optional[torch.Tensor] = None,
                      ignore: Iterable[str] = (), floors: Optional[Dict[str, float]] = None) -> List[str]:
    """Return a list of human-readable mismatches (empty = parity).  Integer/bool tensors must be
    bit-exact; float tensors must satisfy |a-b| <= atol + rtol*|ref|.  ``skip_rows`` masks out envs
    (used for obs_dict entries that alias live buffers mutated by the in-step reset)."""
    bad = []
    # rew_buf and rc/total are SUMS of weighted terms of either sign (e.g. +10 s3_object_height - 11 penetration): the
    # 1e-5 relative bound of the north star holds per term; for the sums it is taken relative to the sum of the terms'
    # magnitudes (cancellation otherwise turns a 1-ulp expf difference in one term into > 1e-5 of a small total)
    mag = None
    floors = floors or {}
    for k, x in ref.items():
        if k.startswith("rc/") and k.endswith("_weighted") and x is not None:
            m = torch.clamp(x.detach().cpu().abs().double(), min=float(floors.get(k, 0.0)))
            mag = m if mag is None else mag + m
    for k, x in ref.items():
        if x is None or any(k.startswith(p) for p in ignore):
            continue
        if k not in got:
            bad.append(f"{k}: missing")
            continue
        y = got[k]
        if y is None:
            continue
        x, y = x.detach().cpu(), y.detach().cpu()
        if x.shape != y.shape:
            if x.numel() == y.numel():
                y = y.reshape(x.shape)
            else:
                bad.append(f"{k}: shape {tuple(x.shape)} vs {tuple(y.shape)}")
                continue
        if skip_rows is not None and k.startswith("obs/") and x.dim() >= 1 and x.shape[0] == skip_rows.shape[0]:
            keep = ~skip_rows
            x, y = x[keep], y[keep]
        if x.dtype in (torch.bool, torch.int64, torch.int32, torch.uint8, torch.int16, torch.int8):
            if not torch.equal(x.to(torch.int64), y.to(torch.int64)):
                bad.append(f"{k}: {int((x.to(torch.int64) != y.to(torch.int64)).sum())} integer mismatches")
        else:
            y = y.to(x.dtype)
            err = (x - y).abs()
            tol = atol + rtol * x.abs()
            if k in ("rew_buf", "rc/total") and mag is not None and mag.shape == x.shape:
                tol = atol + rtol * torch.maximum(x.abs().double(), mag).to(x.dtype)
            elif k in floors:
                tol = atol + rtol * torch.clamp(x.abs(), min=float(floors[k]))
            if not bool((err <= tol).all()):
                i = int(torch.argmax(err - tol))
                bad.append(f"{k}: max err {float(err.max()):.3e} (ref {float(x.flatten()[i]):.6e} "
                           f"got {float(y.flatten()[i]):.6e})")
    return bad


# ------------------------------------------------------------------ golden fixtures
def golden_cfg(meta, loader):
    """The task config a fixture was recorded with: ``loader`` is load_task_config (or load_reference_cfg-compatible)."""
    cfg = loader(meta["task"], num_envs=meta["n"], control_mode=meta["control_mode"],
                 policy_controls_hand_base=meta["ctrl_base"], policy_controls_fingers=meta["ctrl_fingers"])
    if meta.get("ties"):
        from .synth import tie_config
        tie_config(cfg)
    return cfg


def load_golden(path):
    import numpy as np

    z = np.load(path, allow_pickle=False)
    meta = {k.split("/", 1)[1]: z[k].item() for k in z.files if k.startswith("meta/")}
    return z, meta


def check_against_golden(path, make_env, rtol=1e-5, atol=1e-6, max_report=12, ignore=()):
    """Run ``make_env(meta)`` through the fixture's trajectory; return a list of mismatches.

    The per-step records (every step) are compared bit-exactly for masks/counters and to
    ``rtol``/``atol`` for rewards and row checksums; the full snapshots at the stored steps are
    compared key by key (obs/* entries only on envs that did not reset in that step, because
    several of them alias live buffers in the reference)."""
    import numpy as np

    z, meta = load_golden(path)
    env = make_env(meta)
    from .synth import BodyLayout, ScriptedWorld, TieWorld

    layout = BodyLayout(with_box=meta["task"] == "BlindGrasping")
    if meta.get("ties"):
        world = TieWorld(meta["n"], layout, env.cfg, seed=meta["seed"], nact=env.num_actions)
    else:
        world = ScriptedWorld(meta["n"], layout, seed=meta["seed"], nact=env.num_actions)
    E = {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith("elem/")}
    steps = meta["steps"]
    full_steps = sorted({int(k.split("/")[1]) for k in z.files if k.startswith("full/")})
    bad: List[str] = []
    S = {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith("steps/")}

    def on_step(t, envs):
        if len(bad) >= max_report:
            return
        s = envs[0].snapshot()
        s = {k: (v.detach().cpu() if v is not None else None) for k, v in s.items()}
        exact = {"reset_buf": s["reset_buf"], "episode_step_count": s["episode_step_count"],
                 "success": s["extras/success"], "failure": s["extras/failure"], "timeout": s["extras/timeout"]}
        if "stage" in S:
            exact["stage"] = s["task/current_stage"]
            exact["success_duration_steps"] = s["task/success_duration_steps"]
        for k, v in exact.items():
            if not np.array_equal(S[k][t].astype(np.int64), v.numpy().astype(np.int64)):
                bad.append(f"step {t} {k}: integer mismatch")
        approx = {"rew_buf": s["rew_buf"].double(), "obs_rowsum": s["obs_buf"].double().sum(dim=1),
                  "tgt_rowsum": s["full_dof_targets"].double().sum(dim=1)}
        for k, v in approx.items():
            ref = torch.from_numpy(S[k][t]).double()
            scale = 1.0 if k == "rew_buf" else 64.0  # row sums accumulate per-element tolerance
            if not bool(((ref - v).abs() <= scale * atol + rtol * ref.abs() * (1.0 if k == "rew_buf" else 8.0)).all()):
                bad.append(f"step {t} {k}: max err {float((ref - v).abs().max()):.3e}")
        if E:  # element-wise, every step (same bound as the full snapshots)
            ref_e = {k: torch.from_numpy(v[t]) for k, v in E.items()}
            for m in compare_snapshots(ref_e, s, rtol, atol):
                bad.append(f"step {t} {m}")
        if t in full_steps:
            ref_full = {k.split("/", 2)[2]: torch.from_numpy(z[k]) for k in z.files if k.startswith(f"full/{t}/")}
            for m in compare_snapshots(ref_full, s, rtol, atol, skip_rows=ref_full["reset_buf"].bool(),
                                       ignore=ignore):
                bad.append(f"step {t} {m}")

    run_lockstep([env], world, steps, on_step)
    return bad
