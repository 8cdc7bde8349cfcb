"""CPU time per control step of the UNMODIFIED reference components (oracle/ref_harness.RefEnv) against the oracle port
(oracle/dexstep_oracle.OracleEnv) on the same scripted inputs - build container only (needs /root/reference).
    python tools/debug/port_vs_reference_cpu.py [envs] [steps]"""
import sys
import time

import torch

sys.path.insert(0, ".")
from oracle import synth  # noqa: E402
from oracle.dexstep_oracle import OracleEnv  # noqa: E402
from oracle.ref_harness import RefEnv, load_reference_cfg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 12
layout = synth.BodyLayout(with_box=True)
world = synth.ScriptedWorld(n, layout, seed=1, nact=18)
states = [world.state(t, torch.full((n,), t), None) for t in range(steps + 3)]
actions = [world.actions(t) for t in range(steps + 3)]
unis = [world.uniforms(t) for t in range(steps + 3)]
print(f"torch threads {torch.get_num_threads()}, envs {n}, steps {steps}")
for name, make in (("reference", lambda: RefEnv("BlindGrasping", n, layout=layout)),
                   ("port", lambda: OracleEnv(load_reference_cfg("BlindGrasping", n), n, layout, synth.dof_props_tensor()))):
    env = make()
    env.load_state(states[0])
    env.reset(unis[0])
    ts = []
    for t in range(1, steps + 3):
        t0 = time.perf_counter()
        env.pre_physics_step(actions[t])
        t1 = time.perf_counter()
        env.load_state(states[t])   # fake physics: not timed
        t2 = time.perf_counter()
        env.post_physics_step(unis[t])
        t3 = time.perf_counter()
        if t > 2:
            ts.append((t1 - t0) + (t3 - t2))
    ts.sort()
    print(f"{name:10s} median {1e3 * ts[len(ts) // 2]:.1f} ms/step  (min {1e3 * ts[0]:.1f})")
