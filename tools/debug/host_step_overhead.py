"""Host-side cost of one DexStepEnv.step_host call (no device sync inside the loop) next to the device-side step time, and
how a SHORT timed region (20 steps, the driver's protocol) compares with a long one, region after region."""
import time
import torch
from oracle import synth
from dexrobot_isaac_b200.config import load_task_config
from dexrobot_isaac_b200.env import DexStepEnv
from dexrobot_isaac_b200.params import BodyLayout

n, dev = 16384, "cuda:0"
cfg = load_task_config("BlindGrasping", num_envs=n)
lay = synth.BodyLayout(with_box=True)
envs = [DexStepEnv(cfg, n, BodyLayout.from_synth(lay), synth.dof_props_tensor(), device=dev, materialize_obs_dict=False,
                   observer=False) for _ in range(2)]
world = synth.ScriptedWorld(n, lay, seed=1, nact=envs[0].num_actions)
for e in envs:
    e.load_state(world.state(0, e.episode_step_count, None))
    e.reset(world.uniforms(0))


def region(compact, k, warm=20):
    bufs = []
    for e in envs:
        hin, hout = e.alloc_host_buffers(compact=compact)
        e.fill_host_inputs(hin, world.actions(1).to(dev))
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
