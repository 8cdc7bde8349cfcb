"""Seeded synthetic sim-state tensors of the shapes Isaac Gym hands the path.

Used by bench.py and ``__graft_entry__.smoke()`` as the stand-in for PhysX (which is out of scope):
the generator runs on whatever device it is given, with plain torch ops — it is input plumbing, not
part of the hot path, and is never inside a timed region.

Shapes: dof_state [N,26,2], rigid_body_state [N,NB,13] (pos3, quat4 xyzw, linvel3, angvel3),
net_contact_force [N,NB,3], actor_root_state [N,NA,13]; actions [N,NACT] in [-1,1]; reset
uniforms [N,29] in [0,1).  Distributions follow SURVEY §8d: DOF positions uniform inside synthetic
limits (no finger DOF has a zero range), DOF velocities N(0,1), hand base in a 0.3 m square
0.05-0.5 m above ground, bodies within +-0.12 m of the base, ~1 % of envs with a fingertip below
ground, unit quaternions, body velocities N(0,0.1), monitored contact forces zero w.p. 0.5 else
N(0,1)^3 * {0.05,0.5,5} (straddling the 0.1 N and 1.0 N thresholds), box near the origin.
"""
from __future__ import annotations

import math
from typing import Dict

import torch

from .params import NUM_DOF, NUM_FINGERS, BodyLayout


def dof_limits() -> torch.Tensor:
    """[26,2] lower/upper: base ART +-1 m, ARR +-pi, thumb rotation [0,2.2], spread joints [0,0.3],
    flexion joints [0,1.3] rad (docs/reference-dof-control-api.md:142-156, docs/DESIGN_DECISIONS.md:83-87)."""
    lim = torch.zeros(NUM_DOF, 2)
    lim[0:3, 0], lim[0:3, 1] = -1.0, 1.0
    lim[3:6, 0], lim[3:6, 1] = -math.pi, math.pi
    for f in range(NUM_FINGERS):
        for j in range(4):
            hi = (2.2 if f == 0 else 0.3) if j == 0 else 1.3
            lim[6 + 4 * f + j, 1] = hi
    return lim


def dof_props() -> torch.Tensor:
    """[26,6] = stiffness, damping, friction, armature, lower, upper (tensor_manager.py:242)."""
    props = torch.zeros(NUM_DOF, 6)
    props[:, 4:6] = dof_limits()
    return props


def random_state(n: int, layout: BodyLayout, gen: torch.Generator, device="cpu",
                 ground_frac: float = 0.01) -> Dict[str, torch.Tensor]:
    dev = torch.device(device)

    def rand(*s):
        return torch.rand(*s, generator=gen, device=dev)

    def randn(*s):
        return torch.randn(*s, generator=gen, device=dev)

    nb, na = layout.num_bodies, layout.num_actors
    lim = dof_limits().to(dev)
    dof_state = torch.stack([lim[:, 0] + rand(n, NUM_DOF) * (lim[:, 1] - lim[:, 0]), randn(n, NUM_DOF)], dim=-1)

    rb = torch.zeros(n, nb, 13, device=dev)
    base = torch.cat([rand(n, 2) * 0.3 - 0.15, 0.05 + rand(n, 1) * 0.45], dim=1)
    rb[:, :, 0:3] = base.unsqueeze(1) + (rand(n, nb, 3) * 0.24 - 0.12)
    rb[:, layout.hand_body, 0:3] = base
    rb[:, :, 2] = rb[:, :, 2].abs() + 1e-3
    hit = rand(n) < ground_frac
    tip = torch.tensor(list(layout.tip_bodies), device=dev)[torch.randint(0, NUM_FINGERS, (n,), generator=gen, device=dev)]
    rows = torch.nonzero(hit).flatten()
    rb[rows, tip[rows], 2] = -0.01 - 0.02 * rand(rows.numel())
    q = randn(n, nb, 4)
    rb[:, :, 3:7] = q / q.norm(dim=-1, keepdim=True)
    rb[:, :, 7:13] = randn(n, nb, 6) * 0.1

    cf = torch.zeros(n, nb, 3, device=dev)
    mon = list(layout.contact_bodies) + ([layout.box_body] if layout.box_body >= 0 else [])
    k = len(mon)
    scale = torch.tensor([0.05, 0.5, 5.0], device=dev)[torch.randint(0, 3, (n, k), generator=gen, device=dev)]
    on = (rand(n, k) >= 0.5).float()
    cf[:, mon, :] = randn(n, k, 3) * (scale * on).unsqueeze(-1)

    root = torch.zeros(n, na, 13, device=dev)
    root[:, 0, 0:7] = rb[:, layout.hand_body, 0:7]
    if layout.box_actor >= 0:
        box = torch.cat([rand(n, 2) * 0.1 - 0.05, 0.02 + rand(n, 1) * 0.23], dim=1)
        root[:, layout.box_actor, 0:3] = box
        root[:, layout.box_actor, 6] = 1.0
        root[:, layout.box_actor, 7:10] = randn(n, 3) * 0.003
        rb[:, layout.box_body, 0:3] = box
        rb[:, layout.box_body, 3:7] = torch.tensor([0.0, 0.0, 0.0, 1.0], device=dev)
    return {"dof_state": dof_state.contiguous(), "rigid_body_state": rb, "net_contact_force": cf,
            "actor_root_state": root}


def random_actions(n: int, nact: int, gen: torch.Generator, device="cpu") -> torch.Tensor:
    return torch.rand(n, nact, generator=gen, device=torch.device(device)) * 2.0 - 1.0


def random_uniforms(n: int, gen: torch.Generator, device="cpu") -> torch.Tensor:
    return torch.rand(n, 29, generator=gen, device=torch.device(device))
