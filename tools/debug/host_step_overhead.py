"""Host-side cost of one DexStepEnv.step_host call (no device sync inside the loop) next to the device-side step time."""
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
bufs = []
for e in envs:
    e.load_state(world.state(0, e.episode_step_count, None))
    e.reset(world.uniforms(0))
    hin, hout = e.alloc_host_buffers(compact="packed")
    e.fill_host_inputs(hin, world.actions(1).to(dev))
    bufs.append((hin, hout))
for i in range(20):
    envs[i % 2].step_host(*bufs[i % 2])
torch.cuda.synchronize()
for k in (200, 200):
    t0 = time.perf_counter()
    for i in range(k):
        envs[i % 2].step_host(*bufs[i % 2])
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"host issue {1e6 * (t1 - t0) / k:.1f} us per step_host; with the device drained {1e6 * (t2 - t0) / k:.1f} us per step")
