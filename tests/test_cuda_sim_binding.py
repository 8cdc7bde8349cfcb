"""the fixed-capacity indexed-setter recipe against a FAKE gym.

The fake implements the documented semantics of Isaac Gym's indexed setters (the simulator copies, from the FULL tensor it
is handed, the rows of the listed global actor ids into its internal state).  Two fakes are driven side by side for many
steps: one by the reference's recipe (host-side ``nonzero(reset_buf)``, index tensors whose length changes every step -
components/reset/reset_manager.py:139-168, physics_manager.py:121-160), one by ``IndexedResetApplier`` (fixed capacity,
padded, no host sync).  Their internal states must stay identical, and ``apply()`` must not synchronise with the host."""
import pytest
import torch

from oracle import synth
from dexrobot_isaac_b200.config import load_task_config

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


class FakeGym:
    """Internal simulator state + the two indexed setters (what PhysX would hold)."""

    def __init__(self, n, num_actors, dof_per_actor0=26):
        self.root = torch.full((n * num_actors, 13), -7.0, device=DEV)   # internal state, != anything a reset writes
        self.dof = torch.full((n, dof_per_actor0, 2), -7.0, device=DEV)
        self.num_actors = num_actors
        self.calls = []

    def set_actor_root_state_tensor_indexed(self, sim, root_tensor, ids, count):
        ids = ids[:count].long()
        self.root[ids] = root_tensor.view(-1, 13)[ids]
        self.calls.append(("root", int(count)))

    def set_dof_state_tensor_indexed(self, sim, dof_tensor, ids, count):
        ids = ids[:count].long()                      # global actor ids of the hand actors
        env = ids // self.num_actors
        self.dof[env] = dof_tensor.view(-1, 26, 2)[env]
        self.calls.append(("dof", int(count)))


@pytest.mark.parametrize("task,capacity", [("BlindGrasping", None), ("BaseTask", None), ("BlindGrasping", 256)])
def test_fixed_capacity_setters_equal_the_reference_recipe(task, capacity):
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout
    from dexrobot_isaac_b200.sim_binding import IndexedResetApplier

    n, steps = 2048, 40
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    env = DexStepEnv(load_task_config(task, num_envs=n), n, BodyLayout.from_synth(lay), synth.dof_props_tensor(), device=DEV)
    p = env.pipeline
    na = lay.na
    ref_gym, new_gym = FakeGym(n, na), FakeGym(n, na)
    applier = IndexedResetApplier(p, new_gym, sim=None, unwrap_tensor=lambda t: t, dof_actor=0, capacity=capacity)
    gen = torch.Generator().manual_seed(3)
    env.load_state(synth.random_state(n, lay, gen))
    env.reset(synth.random_uniforms(n, gen))
    p.uniforms = None
    seen, most = 0, 0
    for t in range(steps):
        env.load_state(synth.random_state(n, lay, gen, ground_frac=0.0 if t % 5 == 4 else 0.04))
        if t % 5 == 4:  # a step without any reset (padding with env 0) - BlindGrasping's random boxes may still fail some
            pass
        env.pre_physics_step(synth.random_actions(n, env.num_actions, gen))
        env.post_physics_step()
        # ---- the new recipe: no host sync allowed inside apply() ----
        torch.cuda.synchronize()
        torch.cuda.set_sync_debug_mode("error")
        try:
            applier.apply()
        finally:
            torch.cuda.set_sync_debug_mode("default")
        # ---- the reference's recipe (reset_manager.py:139-168, physics_manager.py:121-160) ----
        env_ids = torch.nonzero(env.reset_buf).flatten()
        if len(env_ids) > 0:
            all_actor = torch.arange(n * na, device=DEV, dtype=torch.int32).reshape(n, na)
            ids = all_actor[env_ids].flatten()
            ref_gym.set_actor_root_state_tensor_indexed(None, env.actor_root_state_tensor, ids, len(ids))
            gids = (env_ids * na + 0).to(torch.int32)
            ref_gym.set_dof_state_tensor_indexed(None, env.dof_state.reshape(-1, 2), gids, len(env_ids))
        k = int(env_ids.numel())
        seen, most = seen + k, max(most, k)
        cap = applier.capacity
        assert new_gym.calls[-2:] == [("root", cap * na), ("dof", cap)]  # constant arguments, every step
        if k <= cap:
            # rows of the envs that were reset are identical in both simulators; the padded id changed nothing that the
            # reference's simulator would not also hold after a refresh (it rewrote a row with the tensor's own content)
            assert torch.equal(new_gym.root.view(n, na, 13)[env_ids], ref_gym.root.view(n, na, 13)[env_ids]), f"step {t}"
            assert torch.equal(new_gym.dof[env_ids], ref_gym.dof[env_ids]), f"step {t}"
            touched = torch.zeros(n, dtype=torch.bool, device=DEV)
            touched[env_ids] = True
            pad = int(env_ids[0]) if k > 0 else 0
            touched[pad] = True
            # nothing but the reset envs (and the one padding env) was written in this step
            if t == 0:
                assert bool((new_gym.dof[~touched] == -7.0).all()) and bool((new_gym.root.view(n, na, 13)[~touched] == -7.0).all())
            if k == 0:  # padding with env 0: its rows now hold env 0's tensor rows, i.e. its current simulation state
                assert torch.equal(new_gym.dof[0], env.dof_state[0])
    assert seen > 0 and most > 0
    if capacity is None:
        assert int(applier.overflow) == 0
    else:
        # the small capacity must have overflowed at least once in this run for the counter to be exercised
        assert (most > capacity) == (int(applier.overflow) > 0)


def test_reset_actor_indices_layout_and_padding():
    """The kernel's output, element by element, for a hand-made env_ids list."""
    import ctypes

    from dexrobot_isaac_b200 import _lib as L

    lib = L.load()
    ep = L.DexEpilogueBuffers()
    env_ids = torch.tensor([5, 9, 11, 0, 0, 0, 0, 0], dtype=torch.int32, device=DEV)
    for k, want_first in ((3, [5, 9, 11, 5, 5, 5]), (0, [0, 0, 0, 0, 0, 0]), (8, [5, 9, 11, 0, 0, 0])):
        num = torch.tensor([k], dtype=torch.int32, device=DEV)
        ep.env_ids, ep.num_reset = env_ids.data_ptr(), num.data_ptr()
        cap, na = 6, 3
        root = torch.full((cap * na,), -1, dtype=torch.int32, device=DEV)
        dof = torch.full((cap,), -1, dtype=torch.int32, device=DEV)
        ovf = torch.zeros(1, dtype=torch.int32, device=DEV)
        rc = lib.dex_reset_actor_indices(ctypes.byref(ep), cap, na, 1, root.data_ptr(), dof.data_ptr(), ovf.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream)
        assert rc == 0
        e = torch.tensor(want_first, dtype=torch.int32, device=DEV)
        assert torch.equal(dof, e * na + 1)
        assert torch.equal(root.view(cap, na), (e * na).unsqueeze(1) + torch.arange(na, dtype=torch.int32, device=DEV))
        assert int(ovf) == max(k - cap, 0)
    assert lib.dex_reset_actor_indices(ctypes.byref(ep), 0, 3, 1, root.data_ptr(), dof.data_ptr(), None, None) == L.CONSTS["DEX_ERR_SIZE"]
    assert lib.dex_reset_actor_indices(ctypes.byref(ep), 6, 3, 3, root.data_ptr(), dof.data_ptr(), None, None) == L.CONSTS["DEX_ERR_SIZE"]
