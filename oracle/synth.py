"""Synthetic sim-state generator for the step-pipeline oracle, the parity tests and bench.py.

TEST / MEASUREMENT INFRASTRUCTURE ONLY — never imported by the product package.

It stands in for PhysX: it produces the four Isaac Gym tensor-API buffers the path reads
(layout contract: reference tensor_manager.py:173-328)

    dof_state          [N, 26, 2]   (pos, vel) interleaved
    rigid_body_states  [N, NB, 13]  pos3, quat4 (x,y,z,w), linvel3, angvel3
    net_contact_force  [N, NB, 3]
    actor_root_state   [N, NA, 13]

with the value distributions of SURVEY.md §8(d).  The hand MJCF is missing from the
reference checkout (un-vendored submodule), so the body count is an assumption: NB_hand=31
by default (base=0; finger f: links 1+6f..4+6f, tip 5+6f, pad 6+6f), box last.

Two modes:
  * ``random_state``   — i.i.d. per step (bench ring, quick parity cases);
  * ``ScriptedWorld``  — per-env scripted episodes that are a function of the env's own
    ``episode_step_count`` and ``initial_box_positions`` (both bit-exact integer / exactly
    reproducible state of the path under test), so that every FSM stage, every termination
    criterion, success and timeout occur within ~520 steps.
All quantities that feed a thresholded mask are guard-banded (pushed >=1e-4 relative away
from the threshold) unless ``dyadic=True``, in which case every component is a multiple of
2^-6 so that squares, sums and IEEE sqrt are exact on every implementation.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

NUM_DOF = 26
NUM_FINGERS = 5
N_UNIFORMS = 29  # box x, y, yaw | hand translation 3 | hand rotation 3 | finger DOFs 20

BASE_JOINT_NAMES = ["ARTx", "ARTy", "ARTz", "ARRx", "ARRy", "ARRz"]
FINGER_JOINT_NAMES = [f"r_f_joint{f}_{j}" for f in range(1, 6) for j in range(1, 5)]
DOF_NAMES = BASE_JOINT_NAMES + FINGER_JOINT_NAMES


@dataclass
class BodyLayout:
    """Rigid-body / actor index layout of one env (runtime parameter of the path)."""

    nb_hand: int = 31
    with_box: bool = False
    hand: int = 0
    tips: List[int] = field(default_factory=list)
    pads: List[int] = field(default_factory=list)
    contacts: List[int] = field(default_factory=list)

    def __post_init__(self):
        per = (self.nb_hand - 1) // NUM_FINGERS  # 6 for 31, 7 for 36/37
        if per < 6:
            raise ValueError("nb_hand too small for link1..4 + tip + pad per finger")
        if not self.tips:
            self.tips = [1 + per * f + 4 for f in range(NUM_FINGERS)]
            self.pads = [1 + per * f + 5 for f in range(NUM_FINGERS)]
            self.contacts = [1 + per * f + 3 for f in range(NUM_FINGERS)]

    @property
    def nb(self) -> int:
        return self.nb_hand + (1 if self.with_box else 0)

    @property
    def na(self) -> int:
        return 2 if self.with_box else 1

    @property
    def box_body(self) -> int:
        return self.nb_hand if self.with_box else -1

    @property
    def box_actor(self) -> int:
        return 1 if self.with_box else -1

    def body_names(self) -> Dict[int, str]:
        names = {self.hand: "right_hand_base"}
        per = (self.nb_hand - 1) // NUM_FINGERS
        for f in range(NUM_FINGERS):
            for j in range(4):
                names[1 + per * f + j] = f"r_f_link{f + 1}_{j + 1}"
            names[self.tips[f]] = f"r_f_link{f + 1}_tip"
            names[self.pads[f]] = f"r_f_link{f + 1}_pad"
        if self.with_box:
            names[self.box_body] = "box"
        return names


def dof_limits() -> torch.Tensor:
    """Synthetic DOF limits [26, 2]: no finger DOF has a zero range."""
    lim = torch.zeros(NUM_DOF, 2)
    lim[0:3, 0], lim[0:3, 1] = -1.0, 1.0
    lim[3:6, 0], lim[3:6, 1] = -math.pi, math.pi
    for f in range(NUM_FINGERS):
        for j in range(4):
            d = 6 + 4 * f + j
            if j == 0:
                hi = 2.2 if f == 0 else 0.3  # thumb rotation / spread joints
            else:
                hi = 1.3
            lim[d, 0], lim[d, 1] = 0.0, hi
    return lim


def dof_props_tensor() -> torch.Tensor:
    """[26, 6] = stiffness, damping, friction, armature, lower, upper (tensor_manager.py:242)."""
    props = torch.zeros(NUM_DOF, 6)
    props[:, 4:6] = dof_limits()
    return props


# ----------------------------------------------------------------------------------------
# helpers
# ----------------------------------------------------------------------------------------
def _guard(mag: torch.Tensor, thresholds, rel=1e-4) -> torch.Tensor:
    """Return a per-element scale (1.0 or 1.01) that moves ``mag`` out of every guard band."""
    scale = torch.ones_like(mag)
    for thr in thresholds:
        near = (mag - thr).abs() <= rel * 4 * thr
        scale = torch.where(near, torch.full_like(scale, 1.01), scale)
    return scale


def _snap(x: torch.Tensor, bits: int = 6) -> torch.Tensor:
    q = float(2**bits)
    return torch.round(x * q) / q


def _rand_quat(shape, gen) -> torch.Tensor:
    q = torch.randn(*shape, 4, generator=gen)
    return q / q.norm(dim=-1, keepdim=True).clamp_min(1e-6)


def _contact_forces(n, k, gen, dyadic=False) -> torch.Tensor:
    """0 w.p. 0.5 else N(0,1)^3 * U{0.05, 0.5, 5}: straddles thresholds 0.1 and 1.0."""
    on = torch.rand(n, k, generator=gen) < 0.5
    scale = torch.tensor([0.05, 0.5, 5.0])[torch.randint(0, 3, (n, k), generator=gen)]
    f = torch.randn(n, k, 3, generator=gen) * scale.unsqueeze(-1) * on.unsqueeze(-1)
    if dyadic:
        return _snap(f, 6)
    mag = f.double().norm(dim=-1)
    return f * _guard(mag, (0.1, 1.0)).float().unsqueeze(-1)


def random_uniforms(n: int, gen: torch.Generator) -> torch.Tensor:
    """Injected reset randoms [N, 29] in [0, 1) (stand-in for the reference's torch.rand draws,
    blind_grasping_task.py:453-533)."""
    return torch.rand(n, N_UNIFORMS, generator=gen)


def random_actions(n: int, nact: int, gen: torch.Generator) -> torch.Tensor:
    return torch.rand(n, nact, generator=gen) * 2.0 - 1.0


def random_state(n: int, layout: BodyLayout, gen: torch.Generator, dyadic: bool = False,
                 ground_frac: float = 0.01) -> Dict[str, torch.Tensor]:
    """One i.i.d. synthetic sim state (SURVEY §8d distributions)."""
    lim = dof_limits()
    nb, na = layout.nb, layout.na
    u = torch.rand(n, NUM_DOF, generator=gen)
    dof_pos = lim[:, 0] + u * (lim[:, 1] - lim[:, 0])
    dof_vel = torch.randn(n, NUM_DOF, generator=gen)
    dof_state = torch.stack([dof_pos, dof_vel], dim=-1)

    rb = torch.zeros(n, nb, 13)
    base = torch.empty(n, 3)
    base[:, 0:2] = torch.rand(n, 2, generator=gen) * 0.3 - 0.15
    base[:, 2] = 0.05 + torch.rand(n, generator=gen) * 0.45
    rb[:, :, 0:3] = base.unsqueeze(1) + (torch.rand(n, nb, 3, generator=gen) * 0.24 - 0.12)
    rb[:, layout.hand, 0:3] = base
    # keep bodies above ground unless the env is one of the "hitting ground" ones
    rb[:, :, 2] = rb[:, :, 2].abs() + 1e-3
    hit = torch.rand(n, generator=gen) < ground_frac
    which = torch.randint(0, NUM_FINGERS, (n,), generator=gen)
    tip_idx = torch.tensor(layout.tips)[which]
    rows = torch.nonzero(hit).flatten()
    rb[rows, tip_idx[rows], 2] = -0.01 - 0.02 * torch.rand(rows.numel(), generator=gen)
    rb[:, :, 3:7] = _rand_quat((n, nb), gen)
    rb[:, :, 7:13] = torch.randn(n, nb, 6, generator=gen) * 0.1

    cf = torch.zeros(n, nb, 3)
    mon = list(layout.contacts) + ([layout.box_body] if layout.with_box else [])
    cf[:, mon, :] = _contact_forces(n, len(mon), gen, dyadic)

    root = torch.zeros(n, na, 13)
    root[:, 0, 0:7] = rb[:, layout.hand, 0:7]
    if layout.with_box:
        box = torch.empty(n, 3)
        box[:, 0:2] = torch.rand(n, 2, generator=gen) * 0.1 - 0.05
        box[:, 2] = 0.02 + torch.rand(n, generator=gen) * 0.23
        root[:, layout.box_actor, 0:3] = box
        root[:, layout.box_actor, 6] = 1.0
        root[:, layout.box_actor, 7:10] = torch.randn(n, 3, generator=gen) * 0.003
        rb[:, layout.box_body, 0:3] = box
        rb[:, layout.box_body, 3:7] = torch.tensor([0.0, 0.0, 0.0, 1.0])
    if dyadic:
        rb[:, :, 0:3] = _snap(rb[:, :, 0:3], 6)
        root[:, :, 0:3] = _snap(root[:, :, 0:3], 6)
        root[:, :, 7:10] = _snap(root[:, :, 7:10], 9)
    else:
        _guard_geometry(rb, root, layout)
    return {"dof_state": dof_state, "rigid_body_states": rb, "net_contact_force": cf,
            "actor_root_state": root}


def _guard_geometry(rb, root, layout, box0: Optional[torch.Tensor] = None):
    """Push distance-type quantities out of the guard bands of their thresholds, in place."""
    if not layout.with_box:
        return
    box = root[:, layout.box_actor, 0:3]
    prox = math.sqrt(3) * 0.05 / 2 * 1.2
    for idx, thresholds in ((layout.pads, (prox,)), (layout.tips, (0.025,)), ([layout.hand], (0.8,))):
        p = rb[:, idx, 0:3]
        d = (p - box.unsqueeze(1)).double().norm(dim=-1)
        s = _guard(d, thresholds).float().unsqueeze(-1)
        rb[:, idx, 0:3] = box.unsqueeze(1) + (p - box.unsqueeze(1)) * s
    # centroid tolerance 0.08 (first three pads): shift pad 0 slightly if in the band
    cen = rb[:, layout.pads[:3], 0:3].mean(dim=1)
    d = (cen - box).double().norm(dim=-1)
    near = (d - 0.08).abs() <= 4e-4 * 0.08
    rb[near, layout.pads[0], 0] += 0.003
    v = root[:, layout.box_actor, 7:10]
    vm = v.double().norm(dim=-1)
    root[:, layout.box_actor, 7:10] = v * _guard(vm, (0.005,)).float().unsqueeze(-1)
    if box0 is not None:
        drift = (box - box0).double().norm(dim=-1)
        near = (drift - 0.01).abs() <= 4e-4 * 0.01
        root[near, layout.box_actor, 0] += 1e-4
        rb[near, layout.box_body, 0] = root[near, layout.box_actor, 0]


# ----------------------------------------------------------------------------------------
# scripted episodes (BlindGrasping coverage)
# ----------------------------------------------------------------------------------------
SCENARIOS = ("noise", "success", "no_contact", "grasp_lost", "never_lifts", "box_too_far",
             "hit_ground", "bad_pregrasp", "flicker_then_grasp")


class ScriptedWorld:
    """Per-env scripted fake physics.  ``state(t, ep_step, box0)`` is a pure function of the
    global step ``t``, the env's episode step count (int64 state of the path under test) and
    its ``initial_box_positions`` (written by the path's reset from injected uniforms)."""

    def __init__(self, n: int, layout: BodyLayout, seed: int = 1234, nact: int = 18,
                 scenario_weights: Optional[List[float]] = None):
        self.n, self.layout, self.seed, self.nact = n, layout, seed, nact
        g = torch.Generator().manual_seed(seed)
        if not layout.with_box:
            self.scenario = torch.zeros(n, dtype=torch.long)
        else:
            w = torch.tensor(scenario_weights or [3, 3, 2, 2, 1, 1, 1, 2, 1], dtype=torch.float)
            self.scenario = torch.multinomial(w, n, replacement=True, generator=g)
            # make sure every scenario is present when n is large enough
            k = min(n, len(SCENARIOS))
            self.scenario[:k] = torch.arange(k)
        self.phase = torch.randint(0, 40, (n,), generator=g)  # per-env event jitter
        self.pad_dirs = torch.tensor([[1.0, 0.0, 0.0], [-0.5, 0.8, 0.0], [-0.5, -0.8, 0.0],
                                      [0.0, 1.0, 0.3], [0.0, -1.0, 0.3]])
        self.pad_dirs = self.pad_dirs / self.pad_dirs.norm(dim=1, keepdim=True)

    def _gen(self, t: int, salt: int = 0) -> torch.Generator:
        return torch.Generator().manual_seed(self.seed * 1000003 + 7919 * t + salt)

    def actions(self, t: int) -> torch.Tensor:
        return random_actions(self.n, self.nact, self._gen(t, 1))

    def uniforms(self, t: int) -> torch.Tensor:
        return random_uniforms(self.n, self._gen(t, 2))

    def state(self, t: int, ep_step: torch.Tensor, box0: Optional[torch.Tensor] = None
              ) -> Dict[str, torch.Tensor]:
        lay, n = self.layout, self.n
        gen = self._gen(t, 3)
        st = random_state(n, lay, gen, ground_frac=0.01 if not lay.with_box else 0.0)
        if not lay.with_box:
            return st
        s = ep_step.to(torch.long).cpu()
        sc = self.scenario
        scripted = sc != 0
        # noise envs keep a 1% ground-hit rate
        rb, cf, root = st["rigid_body_states"], st["net_contact_force"], st["actor_root_state"]
        noise_rows = torch.nonzero(~scripted).flatten()
        hit = torch.rand(noise_rows.numel(), generator=gen) < 0.01
        rb[noise_rows[hit], lay.tips[2], 2] = -0.02
        if box0 is None:
            box0 = torch.zeros(n, 3)
            box0[:, 2] = 0.027
        box0 = box0.detach().cpu().float()
        jit = torch.randn(n, 3, generator=gen) * 5e-4  # small numeric variety

        # ---- box trajectory -------------------------------------------------------------
        box = box0.clone()
        box[:, 0:2] += jit[:, 0:2] * 2  # drift ~1 mm << 0.01 tolerance
        lift_start = 232 + self.phase // 4
        lifting = (sc == 1) | (sc == 3) | (sc == 8)
        rise = torch.clamp((s - lift_start).float() * 0.004, 0.0, 0.23)
        box[:, 2] = torch.where(lifting, box0[:, 2] + rise, box0[:, 2])
        boxvel = jit * 2.0  # ~1e-3 << 0.005
        boxvel[:, 2] = torch.where(lifting & (rise > 0) & (rise < 0.23),
                                   torch.full((n,), 0.2), boxvel[:, 2])

        # ---- hand / pads / tips ----------------------------------------------------------
        hand = box.clone()
        hand[:, 2] = box[:, 2] + 0.12
        hand[:, 0] += 0.02
        far = (sc == 5) & (s >= 150 + self.phase)
        hand[far, 0] += 1.0
        pad_r = torch.full((n, 1, 1), 0.036)
        pads = box.unsqueeze(1) + self.pad_dirs.unsqueeze(0) * pad_r
        pads[:, :, 2] += jit[:, 2:3] * 4  # within 0.03 of box height
        bad = sc == 7
        pads[bad, 0:3, 2] += 0.06  # pregrasp height check fails at the 1->2 transition
        tips = box.unsqueeze(1) + self.pad_dirs.unsqueeze(0) * 0.05
        tips[:, :, 2] += 0.02
        pen = (self.phase % 5 == 0) & (s > 210)  # some envs penetrate (tip < 0.025 from centre)
        tips[pen, 1] = box[pen] + torch.tensor([0.012, 0.0, 0.008])
        ground = (sc == 6) & (s >= 60 + 3 * self.phase)
        tips[ground, 3, 2] = -0.015

        rows = torch.nonzero(scripted).flatten()
        rb[rows, lay.hand, 0:3] = hand[rows]
        for f in range(NUM_FINGERS):
            rb[rows, lay.pads[f], 0:3] = pads[rows, f]
            rb[rows, lay.tips[f], 0:3] = tips[rows, f]
        root[rows, lay.box_actor, 0:3] = box[rows]
        root[rows, lay.box_actor, 7:10] = boxvel[rows]
        rb[rows, lay.box_body, 0:3] = box[rows]

        # ---- contacts ---------------------------------------------------------------------
        contact_on = torch.zeros(n, dtype=torch.bool)
        st2 = s >= 203
        contact_on |= ((sc == 1) | (sc == 4)) & st2
        contact_on |= (sc == 3) & st2 & (s < 300 + self.phase)
        flick = (sc == 8) & st2
        # flicker: thumb contact drops every 10th step so the 0.5 s timer never completes,
        # then steady contact from the stage-2 timeout onwards
        contact_on |= flick & ((s % 10 != 0) | (s >= 345))
        f_on = torch.zeros(n, NUM_FINGERS, 3)
        f_on[:, 0] = torch.tensor([0.0, 1.5, 1.0])
        f_on[:, 1] = torch.tensor([0.5, -1.0, 1.2])
        f_on[:, 2] = torch.tensor([0.4, 0.2, 0.9])
        f_on = f_on + torch.randn(n, NUM_FINGERS, 3, generator=gen) * 0.05
        f_on[:, 3:] = 0.0
        forces = f_on * contact_on.view(n, 1, 1)
        boxf = torch.tensor([0.0, 0.3, 3.0]).expand(n, 3) * contact_on.view(n, 1)
        cf[rows.unsqueeze(1), torch.tensor(lay.contacts).unsqueeze(0)] = forces[rows]
        cf[rows, lay.box_body] = boxf[rows]
        _guard_geometry(rb, root, lay, box0)
        return st


# ---------------------------------------------------------------------------------------------------------------------
# tie cases: quantities that feed a thresholded mask placed EXACTLY on the threshold
def tie_config(cfg: dict) -> dict:
    """In place: a dyadic control_dt (sim.dt = 2^-7 -> control_dt = 2^-6) and stage durations that are exact multiples of it
    with the SAME step counts as the shipped config (200 / 150 / 25 / 100 control steps), so that ``time_in_stage``,
    ``stage_contact_duration`` and the grasp duration land exactly on their thresholds (blind_grasping_task.py:835-978,
    1210-1236).  Everything else is the shipped configuration."""
    cfg["sim"]["dt"] = 2.0 ** -7
    tc = cfg["task"]
    if "stage1_duration" in tc:
        tc["stage1_duration"] = 200 / 64
        tc["stage2_duration"] = 150 / 64
        tc["stage_evaluation"]["stage2_contact_success_threshold"] = 25 / 64
        tc["contact_duration_threshold"] = 100 / 64
    return cfg


class TieWorld(ScriptedWorld):
    """ScriptedWorld whose states are post-processed so that, in a random ~1/6 of the envs per step and quantity, a
    quantity that feeds a strict / non-strict comparison of the path sits exactly on its float32 threshold:

      * monitored contact forces of magnitude exactly f32(contact_binary_threshold) and exactly f32(0.1)
        (observation_encoder.py:712 ``>``; reward_calculator.py:186 ``> 0.1``), box force likewise (blind_grasping_task.py:797),
      * a fingerpad at exactly f32(proximity_threshold) from the box centre (:797-812 ``<``),
      * box height exactly f32(success_height_threshold) (:885 ``>``),
      * fingertip / hand-base height exactly the height-safety thresholds (step_processor.py:150-160 ``<``),
      * hand at exactly f32(max_box_distance) from the box (:1290 ``>``),
      * box speed exactly f32(velocity_tolerance) (:1305-1330 ``<=``).

    The vectors are axis aligned and the other coordinate of the difference is zero, so squares, sums, differences and
    IEEE sqrt are exact whatever the summation order or FMA contraction of an implementation.  ``hits`` counts the injected
    ties per kind (tests assert they happened)."""

    def __init__(self, n, layout, cfg: dict, seed: int = 1234, nact: int = 18, rate: float = 1.0 / 6.0,
                 scripted_rate: float = 0.002, **kw):
        kw.setdefault("scenario_weights", [8, 3, 2, 2, 1, 1, 1, 2, 1])
        super().__init__(n, layout, seed=seed, nact=nact, **kw)
        self.scripted_rate = scripted_rate
        tc = cfg["task"]
        f32 = lambda v: float(torch.tensor(float(v), dtype=torch.float32))  # noqa: E731
        self.rate = rate
        self.thr_contact = f32(tc["contact_binary_threshold"])
        self.thr_stability = f32(0.1)
        hs = (tc.get("termination", {}) or {}).get("height_safety")
        self.thr_hand_z = f32(hs["handbase_threshold"]) if hs else None
        self.thr_tip_z = f32(hs["fingertip_threshold"]) if hs else None
        if layout.with_box:
            self.thr_prox = f32(math.sqrt(3) * cfg["env"]["box"]["size"] / 2 * 1.2)
            self.thr_box_z = f32(tc["success_height_threshold"])
            self.thr_far = f32(tc["max_box_distance"])
            self.thr_vel = f32(tc["quality_thresholds"]["velocity_tolerance"])
        self.hits: Dict[str, int] = {}

    def _pick(self, gen, kind: str) -> torch.Tensor:
        # scripted envs (scenario != 0) need undisturbed streaks of 25 / 100 steps to walk the FSM (and so reach the
        # timer ties), so they get ties rarely; the noise envs get them at the full rate
        rate = torch.where(self.scenario == 0, self.rate, self.scripted_rate)
        rows = torch.nonzero(torch.rand(self.n, generator=gen) < rate).flatten()
        self.hits[kind] = self.hits.get(kind, 0) + int(rows.numel())
        return rows

    @staticmethod
    def _axis(rows, gen):
        """Per-row signed unit vector along a random axis."""
        ax = torch.randint(0, 3, (rows.numel(),), generator=gen)
        sg = torch.randint(0, 2, (rows.numel(),), generator=gen).float() * 2 - 1
        v = torch.zeros(rows.numel(), 3)
        v[torch.arange(rows.numel()), ax] = sg
        return v, ax

    def state(self, t, ep_step, box0=None):
        st = super().state(t, ep_step, box0)
        lay, n = self.layout, self.n
        gen = self._gen(t, 9)
        rb, cf, root = st["rigid_body_states"], st["net_contact_force"], st["actor_root_state"]
        mon = torch.tensor(lay.contacts)
        for kind, thr in (("force==contact_thr", self.thr_contact), ("force==0.1", self.thr_stability)):
            rows = self._pick(gen, kind)
            body = mon[torch.randint(0, len(lay.contacts), (rows.numel(),), generator=gen)]
            cf[rows, body] = self._axis(rows, gen)[0] * thr
        if self.thr_tip_z is not None:
            rows = self._pick(gen, "tip_z==thr")
            tip = torch.tensor(lay.tips)[torch.randint(0, NUM_FINGERS, (rows.numel(),), generator=gen)]
            rb[rows, tip, 2] = self.thr_tip_z
            rows = self._pick(gen, "hand_z==thr")
            rb[rows, lay.hand, 2] = self.thr_hand_z
            root[rows, 0, 2] = self.thr_hand_z
        if not lay.with_box:
            return st
        ba, bb = lay.box_actor, lay.box_body

        def zero_box_axis(rows, ax):
            root[rows, ba, ax] = 0.0
            rb[rows, bb, ax] = 0.0

        rows = self._pick(gen, "box_force==contact_thr")
        cf[rows, bb] = self._axis(rows, gen)[0] * self.thr_contact
        rows = self._pick(gen, "box_z==height_thr")
        root[rows, ba, 2] = self.thr_box_z
        rb[rows, bb, 2] = self.thr_box_z
        rows = self._pick(gen, "box_speed==tol")
        root[rows, ba, 7:10] = self._axis(rows, gen)[0] * self.thr_vel
        # geometry ties: zero the box coordinate along the chosen axis so that (box + d) - box == d exactly
        rows = self._pick(gen, "pad_dist==prox")
        v, ax = self._axis(rows, gen)
        zero_box_axis(rows, ax)
        pad = torch.tensor(lay.pads)[torch.randint(0, NUM_FINGERS, (rows.numel(),), generator=gen)]
        rb[rows, pad, 0:3] = root[rows, ba, 0:3] + v * self.thr_prox
        rows = self._pick(gen, "hand_dist==max")
        v, ax = self._axis(rows, gen)
        v[:, 2] = v[:, 2].abs()  # never below the ground plane (that would be a plain hitting_ground failure)
        zero_box_axis(rows, ax)
        # (a pad tie of the same env stays valid only if it used the same axis; re-place pads of these rows is not needed:
        # every kind is counted on its own and the comparison is on whatever the state finally holds)
        rb[rows, lay.hand, 0:3] = root[rows, ba, 0:3] + v * self.thr_far
        root[rows, 0, 0:3] = rb[rows, lay.hand, 0:3]
        return st

    def count_exact(self, st: Dict[str, torch.Tensor]) -> Dict[str, int]:
        """How many quantities of ``st`` sit exactly on a threshold (computed like the path computes them)."""
        lay = self.layout
        rb, cf, root = st["rigid_body_states"], st["net_contact_force"], st["actor_root_state"]
        mag = torch.norm(cf[:, list(lay.contacts)], dim=2)
        out = {"force==contact_thr": int((mag == self.thr_contact).sum()), "force==0.1": int((mag == self.thr_stability).sum())}
        if self.thr_tip_z is not None:
            out["tip_z==thr"] = int((rb[:, list(lay.tips), 2] == self.thr_tip_z).sum())
            out["hand_z==thr"] = int((rb[:, lay.hand, 2] == self.thr_hand_z).sum())
        if lay.with_box:
            box = root[:, lay.box_actor, 0:3]
            out["box_force==contact_thr"] = int((torch.norm(cf[:, lay.box_body], dim=1) == self.thr_contact).sum())
            out["box_z==height_thr"] = int((box[:, 2] == self.thr_box_z).sum())
            out["box_speed==tol"] = int((torch.norm(root[:, lay.box_actor, 7:10], dim=1) == self.thr_vel).sum())
            out["pad_dist==prox"] = int((torch.norm(rb[:, list(lay.pads), 0:3] - box.unsqueeze(1), dim=2) == self.thr_prox).sum())
            out["hand_dist==max"] = int((torch.norm(rb[:, lay.hand, 0:3] - box, dim=1) == self.thr_far).sum())
        return out
