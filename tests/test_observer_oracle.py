"""The oracle restatement of the reference's RewardComponentObserver accumulation (oracle/reward_observer.py) against
the UNMODIFIED reference observer (dexhand_env/rl/reward_observer.py) run behind rl_games stubs.  Build container only
(needs /root/reference); pins the checker that the GPU tests use for the fused device-side accumulator."""
import pytest
import torch

from oracle import ref_harness, synth
from oracle.reward_observer import OracleObserver

pytestmark = pytest.mark.skipif(not ref_harness.reference_available(), reason="/root/reference not present")


@pytest.mark.parametrize("task,steps", [("BlindGrasping", 330), ("BaseTask", 320)])
def test_oracle_observer_matches_reference_observer(task, steps):
    n = 48
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    env = ref_harness.RefEnv(task, n, lay)
    ref = ref_harness.RefObserver(env)
    world = synth.ScriptedWorld(n, lay, seed=77, nact=env.num_actions)
    env.load_state(world.state(0, env.episode_step_count, getattr(env, "initial_box_positions", None)))
    env.reset(world.uniforms(0))
    names, ora = None, None
    done_total = 0
    for t in range(1, steps + 1):
        env.pre_physics_step(world.actions(t))
        env.load_state(world.state(t, env.episode_step_count, getattr(env, "initial_box_positions", None)))
        env.post_physics_step(world.uniforms(t))
        rc = env.extras["reward_components"]
        if ora is None:
            names = list(rc.keys())
            ora = OracleObserver(n, names)
        ora.after_steps(rc)
        ora.process_infos(env.extras, torch.nonzero(env.reset_buf).flatten())
        ref.step()
        done_total += int(env.reset_buf.sum())
    assert done_total > n  # every env finished at least one episode
    want = ref.totals(names)
    got = ora.totals.clone()
    got[:, 2 + names.index("total")] = 0  # the reference skips "total" (reward_observer.py:254)
    assert torch.equal(got[:, 0], want[:, 0])          # episode counts per class
    assert torch.equal(got[:, 1], want[:, 1])          # summed lengths per class
    assert torch.allclose(got[:, 2:], want[:, 2:], rtol=1e-6, atol=1e-6)
    assert float(want[0, 0]) == done_total and float(want[1:, 0].sum()) == done_total
    # the live per-env accumulators agree too
    for k in names:
        assert torch.allclose(ora.episode_sums[k], ref.obs.episode_reward_sums[k], rtol=0, atol=0), k
    assert torch.equal(ora.episode_lengths, ref.obs.episode_lengths)
