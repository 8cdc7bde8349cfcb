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
cmd = [B.find_nvcc()] + B.NVCC_FLAGS + ["-DDEXSTEP_TIMELINE", "-I", os.path.join(ROOT, "include")] + B.SOURCES + ["-o", tl_path]
subprocess.run(cmd, check=True)
L._lib = L.load(tl_path)  # every pipeline created from here on uses the instrumented library
lib = L._lib
lib.dex_debug_set_timeline.argtypes = [ctypes.c_void_p]

from dexrobot_isaac_b200 import synthetic  # noqa: E402
from dexrobot_isaac_b200.config import load_task_config  # noqa: E402
from dexrobot_isaac_b200.env import DexStepEnv  # noqa: E402
from dexrobot_isaac_b200.params import BodyLayout  # noqa: E402

dev = torch.device("cuda:0")
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
tl = torch.zeros(grid, 8, dtype=torch.int64, device=dev)
for _ in range(3):
    for e, a in envs:
        e.pipeline.step(a)
torch.cuda.synchronize()
assert lib.dex_debug_set_timeline(tl.data_ptr()) == 0
e, a = envs[0]
e.pipeline.post_step()
torch.cuda.synchronize()
assert lib.dex_debug_set_timeline(None) == 0
t = tl.cpu().double()
t0 = t[:, 0].min()
names = ["start", "stage-in done", "items done", "scalar done", "stage-out done", "reset done", "end"]
print(f"envs {n} task {task} grid {grid}; times in us relative to the first CTA start")
for k, nm in enumerate(names):
    c = (t[:, k] - t0) / 1e3
    print(f"{nm:16s} min {c.min():7.2f}  median {c.median():7.2f}  p90 {c.quantile(0.9):7.2f}  max {c.max():7.2f}")
d = (t[:, 1:7] - t[:, 0:6]) / 1e3
for k, nm in enumerate(["stage-in", "items", "scalar", "stage-out", "reset", "epilogue"]):
    print(f"phase {nm:10s} median {d[:, k].median():6.2f}  p90 {d[:, k].quantile(0.9):6.2f}  max {d[:, k].max():6.2f}")
sm = t[:, 7].long()
print("CTAs per SM: min", int(torch.bincount(sm).min()), "max", int(torch.bincount(sm).max()), "SMs used", int((torch.bincount(sm) > 0).sum()))
