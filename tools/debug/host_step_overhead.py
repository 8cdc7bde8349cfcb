"""Host-side cost of one DexStepEnv.step_host call (no device sync inside the loop) next to the device-side step time, and
how a SHORT timed region (20 steps, the driver's protocol) compares with a long one, region after region."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from dexrobot_isaac_b200 import synthetic  # noqa: E402
from dexrobot_isaac_b200.config import load_task_config  # noqa: E402
from dexrobot_isaac_b200.env import DexStepEnv  # noqa: E402
from dexrobot_isaac_b200.params import BodyLayout  # noqa: E402

n, dev = 16384, "cuda:0"
cfg = load_task_config("BlindGrasping", num_envs=n)
lay = BodyLayout.default(with_box=True)
gen = torch.Generator(device=dev).manual_seed(1)
envs, actions = [], []
for i in range(2):
    e = DexStepEnv(cfg, n, lay, synthetic.dof_props(), device=dev, materialize_obs_dict=False, observer=False, seed=1 + i,
                   sim=synthetic.random_state(n, lay, gen, dev))
    e.pipeline.uniforms = None  # resets draw from the in-kernel Philox stream
    e.reset()
    a = synthetic.random_actions(n, e.num_actions, gen, dev).contiguous()
    e.pipeline.step(a)
    envs.append(e)
    actions.append(a)


def region(compact, k, warm=20):
    bufs = []
    for e, a in zip(envs, actions):
        hin, hout = e.alloc_host_buffers(compact=compact)
        e.fill_host_inputs(hin, a)
        bufs.append((hin, hout))
    for i in range(warm):
        envs[i % 2].step_host(*bufs[i % 2])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main = torch.cuda.current_stream()
    t0 = time.perf_counter()
    e0.record(main)
    for i in range(k):
        envs[i % 2].step_host(*bufs[i % 2])
    t1 = time.perf_counter()
    for e in envs:
        e.host_sync()
    e1.record(main)
    torch.cuda.synchronize()
    for e in envs:
        e.end_host_mode()
    return e0.elapsed_time(e1) * 1e3 / k, 1e6 * (t1 - t0) / k


for compact, k in (("packed", 20), ("packed", 20), ("packed", 20), ("packed", 200), (True, 20), (True, 200), ("packed", 20)):
    dev_us, host_us = region(compact, k)
    print(f"compact={compact!s:7s} {k:4d} steps: {dev_us:7.1f} us per step on the device, host issue {host_us:5.1f} us per step")
