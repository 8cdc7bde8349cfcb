"""Debug: drive the v2 kernel and the tile kernel through the same scripted world; report the first divergence."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from oracle import synth
from tests.test_cuda_v2 import lean_env

task, n, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
lay = synth.BodyLayout(with_box=task == "BlindGrasping")
e1, e2 = lean_env(task, n), lean_env(task, n)
e1.pipeline.tile_kernel = True
world = synth.ScriptedWorld(n, lay, seed=209, nact=e1.num_actions)
s0 = world.state(0, e1.episode_step_count, getattr(e1, "initial_box_positions", None))
u0 = world.uniforms(0)
for e in (e1, e2):
    e.load_state(s0); e.reset(u0)
last_reset = torch.zeros(n, dtype=torch.long)
for t in range(1, steps + 1):
    a = world.actions(t)
    for e in (e1, e2): e.pre_physics_step(a.to("cuda:0"))
    st = world.state(t, e1.episode_step_count, getattr(e1, "initial_box_positions", None))
    u = world.uniforms(t)
    for e in (e1, e2):
        e.load_state(st); e.post_physics_step(u)
    torch.cuda.synchronize()
    p1, p2 = e1.pipeline, e2.pipeline
    rb = p1.out["reset_buf"].cpu()
    bad = []
    for k, v in list(p1.state.items()) + [("out/" + k, v) for k, v in p1.out.items() if v is not None and k not in ("counters",)]:
        w = p2.state[k] if not k.startswith("out/") else p2.out[k[4:]]
        if k == "rng_step": continue
        if not torch.equal(v, w):
            d = (v != w)
            idx = torch.nonzero(d.reshape(-1, d.shape[-1]).any(dim=0) if d.dim() == 2 and d.shape[-1] == n else d.reshape(n, -1).any(dim=1)).flatten().cpu()
            bad.append((k, idx[:8].tolist(), int(idx.numel())))
    if bad:
        print("step", t, "first divergence:")
        for k, idx, cnt in bad:
            print("  ", k, "envs", idx, "count", cnt, "last reset at", [int(last_reset[i]) for i in idx[:8]], "reset now", [bool(rb[i]) for i in idx[:8]])
        break
    last_reset[rb] = t
else:
    print("no divergence in", steps, "steps; resets total", int(e1.pipeline.epi["totals"][3]))
