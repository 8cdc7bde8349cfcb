#!/usr/bin/env python
"""Per-CTA phase timeline of the post-physics kernel (debug build with -DDEXSTEP_TIMELINE).

    python tools/timeline.py [envs] [task]      (on the GPU box)
Builds dexrobot_isaac_b200/libdexstep_tl.so, runs a few L2-cold steps and prints when each phase of each CTA
started/ended relative to the first CTA start: stage-in, items, scalar, stage-out, reset, epilogue."""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from dexrobot_isaac_b200 import build as B  # noqa: E402
from dexrobot_isaac_b200 import _lib as L  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
task = sys.argv[2] if len(sys.argv) > 2 else "BlindGrasping"
tl_path = os.path.join(B.PKG_DIR, "libdexstep_tl.so")
cmd = [B.find_nvcc()] + B.NVCC_FLAGS + ["-DDEXSTEP_TIMELINE"] + os.environ.get("TL_FLAGS", "").split() + ["-I", os.path.join(ROOT, "include")] + B.SOURCES + ["-o", tl_path]
subprocess.run(cmd, check=True)
L._lib = L.load(tl_path)  # every pipeline created from here on uses the instrumented library
lib = L._lib
lib.dex_debug_set_timeline.argtypes = [ctypes.c_void_p]

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
grid = (n + 31) // 32
tl = torch.zeros(grid + 1, 32, dtype=torch.int64, device=dev)  # + the epilogue CTA (marks 0 and 6 only)
for _ in range(3):
    for e, a in envs:
        e.pipeline.step(a)
torch.cuda.synchronize()
assert lib.dex_debug_set_timeline(tl.data_ptr()) == 0
runs, epi_end = [], []
for e, a in envs:  # one L2-cold launch per replica (6 x ~63 MB > L2)
    tl.zero_()
    e.pipeline.post_step()
    torch.cuda.synchronize()
    full = tl.cpu().double()
    epi_end.append(float(full[grid, 6] - full[:grid, 0].min()) / 1e3)
    runs.append(full[:grid])
assert lib.dex_debug_set_timeline(None) == 0
LIGHT = "-DDEXSTEP_TIMELINE_LIGHT" in os.environ.get("TL_FLAGS", "")
import statistics as st_
def med(f):
    return st_.median(f(t) for t in runs)
print(f"envs {n} task {task} grid {grid}; median over {len(runs)} launches")
print(f"epilogue CTA end   median {st_.median(epi_end):7.2f} us after the first CTA start")
if LIGHT:
    # light marks: 32-bit SM clocks per mark (differences mod 2^32), globaltimer only at CTA start / end
    ghz = 1.965
    def d(t, a_, b_):
        return ((t[:, 16 + b_] - t[:, 16 + a_]) % 2 ** 32) / ghz / 1e3
    print(f"CTA start skew     median {med(lambda t: float(((t[:, 0] - t[:, 0].min()) / 1e3).median())):6.2f}  "
          f"max {med(lambda t: float(((t[:, 0] - t[:, 0].min()) / 1e3).max())):6.2f} us")
    print(f"CTA end            median {med(lambda t: float(((t[:, 6] - t[:, 0].min()) / 1e3).median())):6.2f}  "
          f"max {med(lambda t: float(((t[:, 6] - t[:, 0].min()) / 1e3).max())):6.2f} us")
    for (a_, b_), nm in [((0, 1), "stage-in"), ((1, 2), "items"), ((2, 3), "scalar || early stage-out"),
                         ((2, 7), "  scalar: task obs + FSM"), ((7, 8), "  scalar: termination"),
                         ((8, 9), "  scalar: ballots"), ((9, 10), "  scalar: record"), ((10, 3), "  scalar: vote barrier"),
                         ((11, 12), "  warp 1: early f32 rows"), ((12, 13), "  warp 1: early obs chunks"),
                         ((3, 4), "terms + late stage-out"), ((4, 5), "reset"), ((0, 5), "whole CTA")]:
        print(f"phase {nm:28s} median {med(lambda t: float(d(t, a_, b_).median())):6.2f}  "
              f"max {med(lambda t: float(d(t, a_, b_).max())):6.2f} us")
    sys.exit(0)
names = ["start", "stage-in done", "items done", "scalar done", "stage-out done", "reset done", "end"]
print(f"envs {n} task {task} grid {grid}; times in us relative to the first CTA start; median over {len(runs)} launches")
for k, nm in enumerate(names):
    print(f"{nm:16s} median {med(lambda t: float(((t[:, k] - t[:, 0].min()) / 1e3).median())):7.2f}  "
          f"max {med(lambda t: float(((t[:, k] - t[:, 0].min()) / 1e3).max())):7.2f}")
for k, nm in enumerate(["stage-in", "items", "scalar", "stage-out", "reset", "epilogue"]):
    print(f"phase {nm:10s} median {med(lambda t: float(((t[:, k + 1] - t[:, k]) / 1e3).median())):6.2f}  "
          f"max {med(lambda t: float(((t[:, k + 1] - t[:, k]) / 1e3).max())):6.2f}")
# SM-clock deltas inside the scalar phase (marks 2 -> 7 FSM done -> 8 termination done -> 9 counters -> 10 ticket -> 3)
seq = [2, 7, 8, 9, 10, 3]
lab = ["task obs + FSM", "termination + masks", "ballots + counters", "len-sum + ticket", "barrier (wait for warps 1-7)"]
for (a_, b_), nm in zip(zip(seq[:-1], seq[1:]), lab):
    print(f"scalar: {nm:30s} median {med(lambda t: float((t[:, 16 + b_] - t[:, 16 + a_]).median())):8.0f} clk  "
          f"max {med(lambda t: float((t[:, 16 + b_] - t[:, 16 + a_]).max())):8.0f} clk")
for (a_, b_), nm in (((11, 12), "warp 1: early f32 rows"), ((12, 13), "warp 1: early obs chunks")):
    print(f"early:  {nm:30s} median {med(lambda t: float((t[:, 16 + b_] - t[:, 16 + a_]).median())):8.0f} clk  "
          f"max {med(lambda t: float((t[:, 16 + b_] - t[:, 16 + a_]).max())):8.0f} clk")
# event-timed average of the (instrumentation-free path of the) same library: ring over the replicas
e0_, e1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 40
e0_.record()
for _ in range(reps):
    for e, a in envs:
        e.pipeline.post_step()
e1_.record()
torch.cuda.synchronize()
print(f"post-physics launch avg (stream-ordered, eager launches): {1e3 * e0_.elapsed_time(e1_) / (reps * len(envs)):.2f} us")
