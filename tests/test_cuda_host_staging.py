"""GPU tests of the host-buffer entry (DexStepEnv.alloc_host_buffers / fill_host_inputs / step_host, the call bench.py's
``e2e`` times): one env steps on its own device tensors, a second one is fed the same states through pinned host buffers in
each staging layout - the Isaac Gym tensors as they are, the compact layout (monitored rows only) and the packed one
(7-float pose slots, DexParams.rigid_row_floats = 7).  The kernels and their arithmetic are the same, only the addressing of
the sim rows differs, so every output and every state buffer must agree BIT FOR BIT.  Also: the library rejects a row pitch
it cannot address."""
import ctypes

import pytest
import torch

from oracle import synth
from oracle.driver import box0_of
from dexrobot_isaac_b200.config import load_task_config

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

STATE_KEYS = ("active_prev_targets", "full_dof_targets", "prev_actions", "prev_dof_pos", "contact_duration_steps",
              "prev_contact_binary", "active_rule_targets")
TASK_KEYS = ("success_duration_steps", "current_stage", "time_in_stage", "stage_contact_duration", "initial_box_positions")


def _env(task, n, **kw):
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    cfg = load_task_config(task, num_envs=n)
    layout = synth.BodyLayout(with_box=task == "BlindGrasping")
    return DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device=DEV,
                      materialize_obs_dict=False, **kw), layout


def _put(env, state):
    """The fake PhysX step of the host-fed env: its FULL-layout device tensors stand in for the simulator's memory."""
    full = env._full_sim()
    for k, v in full.items():
        src = state[k] if k in state else state["rigid_body_states"]
        v.copy_(src.reshape(v.shape))


@pytest.mark.parametrize("task", ["BaseTask", "BlindGrasping"])
@pytest.mark.parametrize("compact", [False, True, "packed"], ids=["full", "compact", "packed"])
@pytest.mark.parametrize("n,blocks", [(100, True), (100, False), (4096, True)], ids=["n100", "n100_per_tensor", "n4096"])
def test_step_host_equals_device_step(task, compact, n, blocks):
    steps = 150 if n == 100 else 40
    a, layout = _env(task, n)
    b, _ = _env(task, n)
    world = synth.ScriptedWorld(n, layout, seed=77, nact=a.num_actions)
    hin, hout = b.alloc_host_buffers(compact=compact)
    if compact == "packed":
        assert tuple(hin["rigid_body_state"].shape) == (n, 12, 7)
    assert all(v.is_pinned() for v in list(hin.values()) + list(hout.values()))
    if not blocks:  # a caller's own pinned tensors: one copy per tensor
        hin = {k: v for k, v in hin.items() if k != "_block"}
        hout = {k: v for k, v in hout.items() if k != "_block"}
    s0 = world.state(0, a.episode_step_count, box0_of(a))
    a.load_state(s0)
    b.load_state(s0)
    u0 = world.uniforms(0)
    a.reset(u0)
    b.reset(u0)
    resets = 0
    for t in range(1, steps + 1):
        act = world.actions(t).to(DEV)
        st = world.state(t, a.episode_step_count, box0_of(a))
        a.load_state(st)
        a.pipeline.step(act)
        _put(b, st)
        b.fill_host_inputs(hin, act)
        h2d, d2h = b.step_host(hin, hout)
        torch.cuda.synchronize()
        assert torch.equal(hout["obs_buf"], a.obs_buf.cpu()), f"step {t} obs_buf"
        assert torch.equal(hout["rew_buf"], a.rew_buf.cpu()), f"step {t} rew_buf"
        assert torch.equal(hout["reset_buf"], a.reset_buf.cpu()), f"step {t} reset_buf"
        assert torch.equal(b.pipeline.env_ids(), a.pipeline.env_ids()), f"step {t} env_ids"
        for k in STATE_KEYS + (TASK_KEYS if task == "BlindGrasping" else ()):
            assert torch.equal(b.pipeline.view(k), a.pipeline.view(k)), f"step {t} {k}"
        resets += int(a.reset_buf.sum())
    assert resets > 0
    per_env = {False: None, True: (11 * 13 + 52 + (6 if task == "BlindGrasping" else 5) * 3 + 13) * 4,
               "packed": (12 * 7 + 52 + (6 if task == "BlindGrasping" else 5) * 3 + 13) * 4}[compact]
    if per_env is not None:  # as copied: the single-block path rounds each of its 5 (3) segments up to 256 bytes
        payload = n * (per_env + 4 * a.num_actions)
        assert h2d == payload if not blocks else payload <= h2d < payload + 5 * 256
    payload = n * (4 * a.obs_buf.shape[1] + 4 + 1)
    assert d2h == payload if not blocks else payload <= d2h < payload + 3 * 256
    b.end_host_mode()
    assert b.pipeline.layout.rigid_row_floats == 13 and int(b.pipeline.params.rigid_row_floats) == 13
    # back on its own tensors the env keeps stepping in lockstep with the device-fed one
    st = world.state(steps + 1, a.episode_step_count, box0_of(a))
    act = world.actions(steps + 1).to(DEV)
    for e in (a, b):
        e.load_state(st)
        e.pipeline.step(act)
    assert torch.equal(a.obs_buf, b.obs_buf) and torch.equal(a.reset_buf, b.reset_buf)


@pytest.mark.parametrize("kernel", ["tile", "persistent", "single"])
def test_packed_rows_every_kernel_family(kernel):
    """The packed row pitch through each post-physics kernel family (16 384 envs: the single-pass geometry is eligible)."""
    n = 16384
    a, layout = _env("BlindGrasping", n)
    b, _ = _env("BlindGrasping", n)
    for e in (a, b):
        e.pipeline.tile_kernel = kernel == "tile"
        e.pipeline.single_pass = kernel == "single"
    world = synth.ScriptedWorld(n, layout, seed=5, nact=a.num_actions)
    hin, hout = b.alloc_host_buffers(compact="packed")
    s0 = world.state(0, a.episode_step_count, box0_of(a))
    u0 = world.uniforms(0)
    for e in (a, b):
        e.load_state(s0)
        e.reset(u0)
    for t in range(1, 7):
        act = world.actions(t).to(DEV)
        st = world.state(t, a.episode_step_count, box0_of(a))
        a.load_state(st)
        a.pipeline.step(act)
        _put(b, st)
        b.fill_host_inputs(hin, act)
        b.step_host(hin, hout)
        torch.cuda.synchronize()
        assert torch.equal(hout["obs_buf"], a.obs_buf.cpu()) and torch.equal(hout["rew_buf"], a.rew_buf.cpu())
        assert torch.equal(hout["reset_buf"], a.reset_buf.cpu())
    from dexrobot_isaac_b200 import _lib as L
    assert L.load().dex_debug_last_post_kernel() in {"tile": (0,), "persistent": (1, 4), "single": (2,)}[kernel]


def test_row_pitch_is_validated():
    from dexrobot_isaac_b200 import _lib as L

    env, _ = _env("BlindGrasping", 64)
    p = env.pipeline
    act = torch.zeros(64, env.num_actions, device=DEV)
    p.step(act)
    torch.cuda.synchronize()
    saved = int(p.params.rigid_row_floats)
    for bad in (6, 14, -1):
        p.params.rigid_row_floats = bad
        with pytest.raises(L.DexStepError):
            p.step(act)
    # hand row of 13 floats in the LAST 7-float slot would run past the env's block
    p.params.rigid_row_floats = 7
    hand = int(p.params.hand_body)
    p.params.hand_body = int(p.params.num_bodies) - 1
    with pytest.raises(L.DexStepError):
        p.step(act)
    p.params.hand_body, p.params.rigid_row_floats = hand, saved
    p.step(act)
    torch.cuda.synchronize()
    assert ctypes.sizeof(L.DexParams) % 4 == 0
