"""The fused step epilogue (one extra CTA behind the tiles that polls their records) under the conditions that would
expose a scheduling dependence (VERDICT r01 item 8 / ADVICE r01 medium): serialised launches (CUDA_LAUNCH_BLOCKING=1),
no programmatic dependent launch (DEXSTEP_PDL=0), many waves of CTAs (262 144 envs = 8 192 tiles + 1), every kernel
family (persistent, single-pass, tile kernel) - each compared bit for bit with the stand-alone dex_step_epilogue launch
(``fuse_epilogue=False``, the safe configuration).  The waits are bounded: a tile that never publishes traps the kernel
(a CUDA error on the stream) instead of hanging it - see include/dexstep.h, DexEpilogueBuffers.tile_records."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(task, n, steps, **env):
    e = dict(os.environ)
    e.update({k: str(v) for k, v in env.items()})
    r = subprocess.run([sys.executable, "-m", "tests.epilogue_check", task, str(n), str(steps)], cwd=ROOT, env=e,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and r.stdout.strip().startswith("OK"), f"{env}\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}"
    assert int(r.stdout.strip().split()[1]) > 0  # resets happened


@pytest.mark.parametrize("env", [dict(), dict(CUDA_LAUNCH_BLOCKING=1), dict(DEXSTEP_PDL=0),
                                 dict(CUDA_LAUNCH_BLOCKING=1, DEXSTEP_PDL=0), dict(DEXSTEP_POST_V1=1, CUDA_LAUNCH_BLOCKING=1)],
                         ids=["default", "launch_blocking", "no_pdl", "blocking+no_pdl", "tile_kernel+blocking"])
def test_fused_epilogue_262144_envs(env):
    _run("BlindGrasping", 262144, 4, **env)


@pytest.mark.parametrize("n", [16384, 12000, 4096, 100003, 33])
def test_fused_epilogue_every_geometry_launch_blocking(n):
    """persistent kernel from under one tile per SM to many, ragged sizes (the single-pass kernel: next test)."""
    _run("BlindGrasping", n, 5, CUDA_LAUNCH_BLOCKING=1)
    _run("BaseTask", n, 5, CUDA_LAUNCH_BLOCKING=1, DEXSTEP_PDL=0)


def test_fused_epilogue_single_pass_geometry():
    _run("BlindGrasping", 16384, 5, DEXSTEP_POST_GEOMETRY="single", CUDA_LAUNCH_BLOCKING=1)
    _run("BlindGrasping", 12000, 5, DEXSTEP_POST_GEOMETRY="single", DEXSTEP_PDL=0)


def test_rewinding_the_step_counter_rezeroes_the_tile_records():
    """a record counts as current when its 16-bit stamp matches the step counter; StepPipeline.set_rng_step
    re-zeroes the records, so the steps after a rewind still publish exact env_ids / counters."""
    import torch

    from oracle import synth
    from dexrobot_isaac_b200.config import load_task_config
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    n, lay = 4096, synth.BodyLayout(with_box=True)
    env = DexStepEnv(load_task_config("BlindGrasping", num_envs=n), n, BodyLayout.from_synth(lay),
                     synth.dof_props_tensor(), device="cuda:0")
    gen = torch.Generator().manual_seed(5)
    env.load_state(synth.random_state(n, lay, gen))
    env.reset(synth.random_uniforms(n, gen))
    env.pipeline.uniforms = None
    p = env.pipeline
    for t in range(6):
        if t == 3:
            p.set_rng_step(1)  # the stamps of steps 1..3 are in the records already
            assert int(p.epi["tile_records"].abs().sum()) == 0
        env.load_state(synth.random_state(n, lay, gen, ground_frac=0.05))
        p.pre_physics(synth.random_actions(n, env.num_actions, gen).to("cuda:0").contiguous())
        p.post_step()
        ids = p.env_ids()
        assert ids.numel() > 0 and torch.equal(ids, torch.nonzero(p.out["reset_buf"]).flatten())
        assert int(p.epi["num_reset"]) == int(p.out["reset_buf"].sum())
