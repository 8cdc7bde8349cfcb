"""Multi-GPU plumbing of the path: contiguous env sharding and the ONE collective — a sum
all-reduce of the packed episode / reward statistics vector, once per logging interval, off the step path.

Payload (float64, < 2 KB):
    [ counters totals (DEX_NUM_COUNTERS) | steps | stat_totals [4][2 + rows] (optional observer) ]
It replaces the per-step ``.item()`` / ``.any()`` host reductions of rl/reward_observer.py:223-293 and
termination_manager.py:164-185,259-266,331.  Works with any torch.distributed backend (nccl on the GPUs,
gloo in the CPU tests).
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence, Tuple

import torch

from . import _lib as L

C = L.CONSTS
N_COUNTERS = C["DEX_NUM_COUNTERS"]


def shard_range(num_envs_total: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous env range of ``rank``: (first env, count).  Remainder envs go to the lowest ranks."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, rem = divmod(int(num_envs_total), world)
    start = rank * base + min(rank, rem)
    return start, base + (1 if rank < rem else 0)


def pack(totals: torch.Tensor, stat_totals: Optional[torch.Tensor] = None) -> torch.Tensor:
    """totals: int64 [N_COUNTERS + 1] (DexEpilogueBuffers.totals); stat_totals: float64 [4, 2 + rows] or None."""
    if totals.numel() != N_COUNTERS + 1:
        raise ValueError(f"totals must have {N_COUNTERS + 1} elements")
    parts = [totals.to(torch.float64).flatten()]
    if stat_totals is not None:
        parts.append(stat_totals.to(torch.float64).flatten())
    return torch.cat(parts)


def pack_pipelines(pipelines: Sequence) -> torch.Tensor:
    """Sum of the payloads of several pipelines that live on this rank (the bench's ring of replicas)."""
    tot = torch.stack([p.epi["totals"] for p in pipelines]).sum(dim=0)
    st = None
    if pipelines[0].out["stat_totals"] is not None:
        st = torch.stack([p.stat_totals() for p in pipelines]).sum(dim=0)
    return pack(tot, st)


def all_reduce(payload: torch.Tensor, group=None, async_op: bool = False):
    """The single collective of the path.  No-op without an initialised process group (1 GPU)."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None
    return dist.all_reduce(payload, op=dist.ReduceOp.SUM, group=group, async_op=async_op)


def unpack(payload: torch.Tensor, component_names: Optional[Sequence[str]] = None) -> Dict:
    """-> {"steps", "resets", "success", "failure", "timeout", "episode_length_sum", "criteria": {...},
    and, with the observer payload, "episodes": {class: {"n", "mean_length", "components": {name: mean}}}}."""
    v = payload.detach().to("cpu", torch.float64).tolist()
    out = {"steps": v[N_COUNTERS], "resets": v[C["DEX_CNT_RESET"]], "success": v[C["DEX_CNT_SUCCESS"]],
           "failure": v[C["DEX_CNT_FAILURE"]], "timeout": v[C["DEX_CNT_TIMEOUT"]],
           "episode_length_sum": v[C["DEX_CNT_EPISODE_LENGTH"]],
           "criteria": {L.CRITERION_NAMES[c]: v[C["DEX_CNT_CRIT0"] + c] for c in range(C["DEX_NUM_CRITERIA"])}}
    rest = v[N_COUNTERS + 1:]
    if rest:
        if len(rest) % 4:
            raise ValueError("observer payload must be [4][2 + rows]")
        stride = len(rest) // 4
        rows = stride - 2
        names = list(component_names) if component_names is not None else [f"row{i}" for i in range(rows)]
        if len(names) != rows:
            raise ValueError(f"{len(names)} component names for {rows} rows")
        eps = {}
        for k, cls in enumerate(("all", "success", "failure", "timeout")):
            blk = rest[k * stride:(k + 1) * stride]
            n = blk[0]
            eps[cls] = {"n": n, "mean_length": blk[1] / n if n else 0.0,
                        "components": {nm: (blk[2 + i] / n if n else 0.0) for i, nm in enumerate(names)}}
        out["episodes"] = eps
    return out
