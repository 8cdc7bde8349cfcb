"""Helper of tests/test_cuda_epilogue.py, run as a subprocess (the environment variables it is run under -
CUDA_LAUNCH_BLOCKING, DEXSTEP_PDL, DEXSTEP_POST_V1, DEXSTEP_POST_GEOMETRY - are read at CUDA / library initialisation).

    python -m tests.epilogue_check TASK N STEPS

Steps two identical environments on the same seeded random states with in-kernel Philox resets: one with the epilogue fused
into dex_post_physics (one extra CTA that polls the tile records), one with the stand-alone dex_step_epilogue launch (the safe
configuration).  Every output, state and epilogue buffer must be identical bit for bit; prints OK <resets> on success."""
import sys

import torch

from oracle import synth
from dexrobot_isaac_b200.config import load_task_config
from dexrobot_isaac_b200.env import DexStepEnv
from dexrobot_isaac_b200.params import BodyLayout


def main(task, n, steps):
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    cfg = load_task_config(task, num_envs=n)
    envs = [DexStepEnv(cfg, n, BodyLayout.from_synth(lay), synth.dof_props_tensor(), device="cuda:0",
                       materialize_obs_dict=False) for _ in range(2)]
    envs[1].pipeline.fuse_epilogue = False
    gen = torch.Generator().manual_seed(77)
    u = synth.random_uniforms(n, gen)
    for t in range(steps):
        st = synth.random_state(n, lay, gen, ground_frac=0.02)
        a = synth.random_actions(n, envs[0].num_actions, gen).to("cuda:0").contiguous()
        for e in envs:
            e.load_state(st)
            if t == 0:
                e.reset(u)
                e.pipeline.uniforms = None
            else:
                e.pipeline.pre_physics(a)
                e.pipeline.post_step()
        torch.cuda.synchronize()
        p1, p2 = envs[0].pipeline, envs[1].pipeline
        for name, d1, d2 in (("out", p1.out, p2.out), ("state", p1.state, p2.state), ("sim", p1.sim_tensors, p2.sim_tensors)):
            for k, v in d1.items():
                if k == "stat_totals" and v is not None:
                    # observer sums: float64 atomics over all CTAs, order-dependent in the last bits by construction
                    if not torch.allclose(v, d2[k], rtol=1e-9, atol=1e-9):
                        print(f"MISMATCH step {t} {name}/{k} (beyond summation order)")
                        return 1
                    continue
                if v is not None and not torch.equal(v, d2[k]):
                    print(f"MISMATCH step {t} {name}/{k}")
                    return 1
        if not torch.equal(p1.env_ids(), p2.env_ids()):
            print(f"MISMATCH step {t} env_ids")
            return 1
        for k in ("num_reset", "stats", "consecutive_successes", "totals"):
            if not torch.equal(p1.epi[k], p2.epi[k]):
                print(f"MISMATCH step {t} epi/{k}")
                return 1
        ids = p1.env_ids()
        if not torch.equal(ids, torch.nonzero(p1.out["reset_buf"]).flatten()):
            print(f"MISMATCH step {t} env_ids vs nonzero(reset_buf)")
            return 1
    print("OK", int(envs[0].pipeline.epi["totals"][3]))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1], int(sys.argv[2]), int(sys.argv[3])))
