"""Small runs of every kernel family and option combination (tile kernel with obs_dict, persistent kernel lean / with the
observer / with TMA producers, single-pass kernel, policy-input update), a few steps each with in-kernel resets:
    python tools/debug/all_kernels_run.py [steps]
(written for compute-sanitizer, which is closed on this pool; still a quick "does every combination launch" check)."""
import sys

import torch

sys.path.insert(0, ".")
from oracle import synth  # noqa: E402
from dexrobot_isaac_b200.config import load_task_config  # noqa: E402
from dexrobot_isaac_b200.env import DexStepEnv  # noqa: E402
from dexrobot_isaac_b200.params import BodyLayout  # noqa: E402

DEV = "cuda:0"
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
for task, n, dict_, observer, tma in (("BlindGrasping", 77, True, True, False), ("BaseTask", 33, True, True, False),
                                      ("BlindGrasping", 77, False, False, False), ("BlindGrasping", 96, False, True, False),
                                      ("BaseTask", 70, False, True, False), ("BlindGrasping", 128, False, False, True),
                                      ("BlindGrasping", 9472, False, False, False), ("BlindGrasping", 9472, False, True, False)):
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    env = DexStepEnv(load_task_config(task, num_envs=n), n, BodyLayout.from_synth(lay), synth.dof_props_tensor(), device=DEV,
                     materialize_obs_dict=dict_, observer=observer)
    env.pipeline.tma_staging = tma
    env.pipeline.single_pass = n == 9472
    gen = torch.Generator().manual_seed(1)
    env.load_state(synth.random_state(n, lay, gen))
    env.reset(synth.random_uniforms(n, gen))
    env.pipeline.uniforms = None
    env.enable_obs_norm_training()
    for t in range(steps):
        env.load_state(synth.random_state(n, lay, gen, ground_frac=0.1))
        env.pre_physics_step(synth.random_actions(n, env.num_actions, gen))
        env.post_physics_step()
        env.update_obs_normalization()
    torch.cuda.synchronize()
    print(task, n, "obs_dict" if dict_ else "lean", "observer" if observer else "", "tma" if tma else "",
          "kernel", env.pipeline.lib.dex_debug_last_post_kernel(), "resets", int(env.pipeline.epi["totals"][3]), flush=True)
print("done")
