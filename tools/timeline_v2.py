#!/usr/bin/env python
"""Per-CTA, per-tile phase timeline of the persistent post-physics kernel (debug build with -DDEXSTEP_TL2).

    python tools/timeline_v2.py [envs] [task]      (on the GPU box)
Builds dexrobot_isaac_b200/libdexstep_tl2.so, runs L2-cold launches over a ring of replicas and prints, per tile
iteration of a CTA, how long each consumer phase and each producer step took (SM clocks -> us at 1.965 GHz)."""
import ctypes
import os
import statistics as st_
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from dexrobot_isaac_b200 import build as B  # noqa: E402
from dexrobot_isaac_b200 import _lib as L  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
task = sys.argv[2] if len(sys.argv) > 2 else "BlindGrasping"
tl_path = os.path.join(B.PKG_DIR, "libdexstep_tl2.so")
cmd = [B.find_nvcc()] + B.NVCC_FLAGS + ["-DDEXSTEP_TL2"] + os.environ.get("TL_FLAGS", "").split() + \
      ["-I", os.path.join(ROOT, "include")] + B.SOURCES + ["-o", tl_path]
subprocess.run(cmd, check=True)
L._lib = L.load(tl_path)
lib = L._lib
lib.dex_debug_set_timeline2.argtypes = [ctypes.c_void_p]

from dexrobot_isaac_b200 import synthetic  # noqa: E402
from dexrobot_isaac_b200.config import load_task_config  # noqa: E402
from dexrobot_isaac_b200.env import DexStepEnv  # noqa: E402
from dexrobot_isaac_b200.params import BodyLayout  # noqa: E402

dev = torch.device("cuda:0")
torch.zeros(1, device=dev)
lay = BodyLayout.default(with_box=task == "BlindGrasping")
cfg = load_task_config(task, num_envs=n)
gen = torch.Generator(device=dev).manual_seed(1)
envs = []
for i in range(6):
    e = DexStepEnv(cfg, n, lay, synthetic.dof_props(), device=dev, observer=False, seed=i,
                   sim=synthetic.random_state(n, lay, gen, dev))
    e.pipeline.uniforms = None
    e.reset()
    envs.append((e, synthetic.random_actions(n, e.num_actions, gen, dev).contiguous()))
ITERS, GRID = 8, 2 * 148 + 8
tl = torch.zeros(GRID, ITERS, 32, dtype=torch.int64, device=dev)
for _ in range(3):
    for e, a in envs:
        e.pipeline.step(a)
torch.cuda.synchronize()
assert lib.dex_debug_set_timeline2(tl.data_ptr()) == 0
runs = []
for e, a in envs:  # one L2-cold launch per replica
    tl.zero_()
    e.pipeline.post_step()
    torch.cuda.synchronize()
    runs.append(tl.cpu().double())
assert lib.dex_debug_set_timeline2(None) == 0
GHZ = 1.965e3  # clocks per us


def med(f):
    vals = [f(t) for t in runs]
    vals = [v for v in vals if v is not None]
    return st_.median(vals) if vals else float("nan")


t = runs[0]
used = (t[:, 0, 0] > 0)
print(f"envs {n} task {task}: {int(used.sum())} worker CTAs; tiles per CTA: "
      f"{ {k: int(((t[:, :, 1] > 0).sum(dim=1) == k).sum()) for k in range(1, ITERS + 1) if int(((t[:, :, 1] > 0).sum(dim=1) == k).sum())} }")
cons = [((0, 1), "wait full (load latency seen by consumers)"), ((1, 2), "items (thread 0's warp)"), ((2, 3), "  barrier A"),
        ((3, 4), "scalar jobs (warp 0)"), ((4, 5), "  barrier B"), ((5, 6), "combine + record (warp 0)"), ((6, 7), "  barrier C"),
        ((7, 8), "total (warp 0)"), ((8, 9), "obs tile out"), ((9, 10), "reset + store read wait"), ((10, 11), "  barrier D"),
        ((0, 11), "whole tile (consumer)")]
prod = [((16, 17), "producer: wait empty"), ((17, 18), "producer: dof tile + gather issue"), ((18, 19), "producer: state rows issue")]
for it in range(ITERS):
    if not any(bool((r[:, it, 1] > 0).any()) for r in runs):
        break
    print(f"--- tile iteration {it} ---")

    def dur(r, a_, b_):
        m = (r[:, it, a_] > 0) & (r[:, it, b_] > 0)
        if not bool(m.any()):
            return None
        return (r[m, it, b_] - r[m, it, a_]) / GHZ

    for (a_, b_), nm in cons + prod:
        d_med = med(lambda r: None if dur(r, a_, b_) is None else float(dur(r, a_, b_).median()))
        d_max = med(lambda r: None if dur(r, a_, b_) is None else float(dur(r, a_, b_).max()))
        print(f"  {nm:52s} median {d_med:7.2f}  max {d_max:7.2f} us")
# absolute: when does each CTA finish its last tile, relative to the earliest CTA start (globaltimer, 32 ns ticks or coarser)


def ends(r):
    start = r[:, 0, 30]
    m = start > 0
    t0 = start[m].min()
    last_it = (r[:, :, 1] > 0).sum(dim=1) - 1
    # globaltimer of the last iteration's start + its duration in clocks
    out = []
    for cta in torch.nonzero(m).flatten().tolist():
        li = int(last_it[cta])
        out.append(float(r[cta, li, 30] - t0) / 1e3 + float(r[cta, li, 11] - r[cta, li, 0]) / GHZ)
    return out


e_all = [ends(r) for r in runs]
print(f"CTA end after first CTA start: median {st_.median(st_.median(e) for e in e_all):.2f} us, max {st_.median(max(e) for e in e_all):.2f} us")
starts = [((r[:, 0, 30][r[:, 0, 30] > 0]) - (r[:, 0, 30][r[:, 0, 30] > 0]).min()) / 1e3 for r in runs]
print(f"CTA start skew: median {st_.median(float(s.median()) for s in starts):.2f} us, max {st_.median(float(s.max()) for s in starts):.2f} us")
e0_, e1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 40
e0_.record()
for _ in range(reps):
    for e, a in envs:
        e.pipeline.post_step()
e1_.record()
torch.cuda.synchronize()
print(f"post-physics launch avg (stream-ordered eager launches, instrumented build): {1e3 * e0_.elapsed_time(e1_) / (reps * len(envs)):.2f} us")
