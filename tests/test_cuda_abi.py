"""GPU tests of the C ABI itself: known-answer transforms, argument validation / error codes,
CUDA-graph replay == eager launches, in-kernel Philox reset ranges, and size-independent
properties at BASELINE.json's full sizes (env-permutation equivariance, env_ids == nonzero,
counters == mask sums, device-side episode statistics == host recomputation)."""
import ctypes

import pytest
import torch

from oracle import dexstep_oracle as O
from oracle import synth
from tests.test_oracle_transforms import S, known_answer_tensors

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _env(task, n, **kw):
    from tests.test_cuda_parity import make_cuda_env

    return make_cuda_env(task, n, **kw)


def test_known_answer_transforms_cuda():
    from dexrobot_isaac_b200 import _lib as L

    lib = L.load()
    hp, hq, pw, ex = [t.to(DEV).contiguous() for t in known_answer_tensors()]
    out = torch.empty_like(pw)
    L.check(lib.dex_point_in_hand_frame(pw.data_ptr(), hp.data_ptr(), hq.data_ptr(), out.data_ptr(), pw.shape[0],
                                        torch.cuda.current_stream().cuda_stream))
    assert torch.allclose(out.cpu(), ex.cpu(), atol=1e-5)
    ref = O.point_in_hand_frame(pw.cpu(), hp.cpu(), hq.cpu())
    assert torch.allclose(out.cpu(), ref, rtol=1e-6, atol=1e-7)
    # ARR alignment identity: [0,s,0,s] (x) conj([0,s,0,s]) = [0,0,0,1]
    a = torch.tensor([[0.0, S, 0.0, S]], device=DEV)
    b = torch.tensor([[-0.0, -S, -0.0, S]], device=DEV)
    q = torch.empty_like(a)
    L.check(lib.dex_quat_mul(a.data_ptr(), b.data_ptr(), q.data_ptr(), 1, torch.cuda.current_stream().cuda_stream))
    assert torch.allclose(q.cpu(), torch.tensor([[0.0, 0.0, 0.0, 1.0]]), atol=1e-6)
    g = torch.Generator().manual_seed(3)
    a, b = torch.randn(257, 4, generator=g), torch.randn(257, 4, generator=g)
    q = torch.empty(257, 4, device=DEV)
    ad, bd = a.to(DEV), b.to(DEV)
    L.check(lib.dex_quat_mul(ad.data_ptr(), bd.data_ptr(), q.data_ptr(), 257, torch.cuda.current_stream().cuda_stream))
    assert torch.allclose(q.cpu(), O.quat_mul(a, b), rtol=1e-6, atol=1e-6)


def test_error_codes():
    from dexrobot_isaac_b200 import _lib as L

    env = _env("BaseTask", 8)
    p = env.pipeline
    lib, C = p.lib, L.CONSTS
    stream = torch.cuda.current_stream().cuda_stream
    a = torch.zeros(8, 18, device=DEV)
    assert lib.dex_pre_physics(ctypes.byref(p.params), ctypes.byref(p._state), None, None, 8, stream) == C["DEX_ERR_NULL"]
    assert lib.dex_pre_physics(ctypes.byref(p.params), ctypes.byref(p._state), a.data_ptr(), None, 0, stream) == C["DEX_ERR_SIZE"]
    bad = L.DexParams.from_buffer_copy(p.params)
    bad.abi_version = 999
    assert lib.dex_pre_physics(ctypes.byref(bad), ctypes.byref(p._state), a.data_ptr(), None, 8, stream) == C["DEX_ERR_VERSION"]
    bad = L.DexParams.from_buffer_copy(p.params)
    bad.hand_body = 1000
    assert lib.dex_pre_physics(ctypes.byref(bad), ctypes.byref(p._state), a.data_ptr(), None, 8, stream) == C["DEX_ERR_PARAM"]
    sim = L.DexSimViews.from_buffer_copy(p._sim)
    sim.dof_state = p._sim.dof_state + 4
    rc = lib.dex_post_physics(ctypes.byref(p.params), ctypes.byref(sim), ctypes.byref(p._state), ctypes.byref(p._out), None,
                              8, 0, 0, None, 0, 0, stream)
    assert rc == C["DEX_ERR_ALIGN"]
    assert b"aligned" in lib.dex_strerror(rc)
    with pytest.raises(RuntimeError):
        p.pre_physics(torch.zeros(8, 17, device=DEV))
    with pytest.raises(RuntimeError):
        L.check(C["DEX_ERR_NULL"], "x")
    torch.cuda.synchronize()


def test_no_cpu_fallback():
    from dexrobot_isaac_b200 import _lib as L
    from dexrobot_isaac_b200.config import load_task_config
    from dexrobot_isaac_b200.params import BodyLayout
    from dexrobot_isaac_b200.pipeline import StepPipeline

    lim = synth.dof_limits()
    with pytest.raises(L.DexStepError):
        StepPipeline(load_task_config("BaseTask", 4), 4, BodyLayout.default(), lim[:, 0].tolist(), lim[:, 1].tolist(),
                     device="cpu")


@pytest.mark.parametrize("task", ["BaseTask", "BlindGrasping"])
def test_graph_replay_equals_eager(task):
    n, steps = 1000, 12
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    e1, e2 = _env(task, n, materialize=False), _env(task, n, materialize=False)
    world = synth.ScriptedWorld(n, lay, seed=7, nact=e1.num_actions)
    u = world.uniforms(0).to(DEV)
    st0 = world.state(0, e1.episode_step_count, None)
    for e in (e1, e2):
        e.load_state(st0)
        e.reset(u)
    # static input buffers for the graph: the graph's "physics" is a device copy of the next state
    nxt = {k: torch.empty_like(v) for k, v in e2.pipeline.sim_tensors.items()}
    act = torch.zeros(n, e2.num_actions, device=DEV)
    names = {"dof_state": "dof_state", "rigid_body_state": "rigid_body_states",
             "net_contact_force": "net_contact_force", "actor_root_state": "actor_root_state"}

    def physics():
        for k, t in e2.pipeline.sim_tensors.items():
            t.copy_(nxt[k])

    e2.pipeline.uniforms = u
    e1.pipeline.uniforms = u
    g = None
    for t in range(1, steps + 1):
        st = world.state(t, e1.episode_step_count, None)
        a = world.actions(t).to(DEV)
        e1.pre_physics_step(a)
        e1.load_state(st)
        e1.post_physics_step()
        for k, src in names.items():
            nxt[k].copy_(st[src].reshape(nxt[k].shape))
        act.copy_(a)
        if t < 3:
            e2.pipeline.step(act, physics)
        else:
            if g is None:
                # capture replays the step once as warm-up and once inside the capture: snapshot/restore state
                saved = {k: v.clone() for k, v in e2.pipeline.state.items()}
                saved_sim = {k: v.clone() for k, v in e2.pipeline.sim_tensors.items()}
                saved_out = {k: v.clone() for k, v in e2.pipeline.out.items() if v is not None}
                saved_epi = {k: v.clone() for k, v in e2.pipeline.epi.items()}
                g = e2.pipeline.capture_step_graph(act, physics)
                for k, v in saved.items():
                    e2.pipeline.state[k].copy_(v)
                for k, v in saved_sim.items():
                    e2.pipeline.sim_tensors[k].copy_(v)
                for k, v in saved_out.items():
                    e2.pipeline.out[k].copy_(v)
                for k, v in saved_epi.items():
                    e2.pipeline.epi[k].copy_(v)
            e2.pipeline.replay()
        torch.cuda.synchronize()
        for k in ("obs_buf", "rew_buf", "reset_buf", "reward_components"):
            assert torch.equal(e1.pipeline.out[k], e2.pipeline.out[k]), f"step {t} {k}"
        for k, v in e1.pipeline.state.items():
            if k != "rng_step":
                assert torch.equal(v, e2.pipeline.state[k]), f"step {t} state {k}"
        assert torch.equal(e1.pipeline.env_ids(), e2.pipeline.env_ids())


def test_philox_reset_ranges_and_freshness():
    n = 4096
    env = _env("BlindGrasping", n, materialize=False)
    p = env.pipeline
    p.uniforms = None
    draws = []
    for _ in range(2):
        p.reset_idx(None)
        p.epilogue()  # advances the device-resident Philox step counter
        torch.cuda.synchronize()
        pos = env.dof_pos.clone()
        box = env.actor_root_state_tensor[:, 1].clone()
        P = p.params
        assert float(pos[:, 0:3].abs().max()) <= P.hand_translation_range
        assert float(pos[:, 3:6].abs().max()) <= P.hand_rotation_range
        assert 0.0 <= float(pos[:, 6].min()) and float(pos[:, 6].max()) < P.thumb_rotation_range
        assert 0.0 <= float(pos[:, 7:].min()) and float(pos[:, 7:].max()) < P.other_finger_range
        assert float(box[:, 0:2].abs().max()) <= P.box_xy_range and torch.all(box[:, 2] == P.box_z)
        assert torch.allclose(box[:, 5] ** 2 + box[:, 6] ** 2, torch.ones(n, device=DEV), atol=1e-5)
        assert torch.all(env.dof_vel == 0) and torch.all(box[:, 7:] == 0)
        # roughly uniform: mean of U(-r, r) ~ 0, every column distinct across envs
        assert abs(float(pos[:, 0].mean())) < 0.02 and pos[:, 0].unique().numel() > n // 2
        draws.append(pos)
    assert not torch.equal(draws[0], draws[1])  # a new step index gives new randoms
    # active targets after reset are the inverse coupling of the new pose (last DOF of the group wins)
    apt = p.view("active_prev_targets")
    assert torch.equal(apt[:, :6], draws[1][:, :6])
    assert torch.equal(apt[:, 6], draws[1][:, 6]) and torch.equal(apt[:, 8], draws[1][:, 9])
    assert torch.allclose(apt[:, 9], draws[1][:, 22] / 2.0)
    # the stand-alone reset and the reset fused into dex_post_physics draw from different Philox streams: an env reset by
    # both within one step (same step counter) gets two different poses
    step_before = int(p.state["rng_step"])
    p.reset_idx(None)
    a = env.dof_pos.clone()
    env.episode_step_count.fill_(int(p.params.max_episode_length) - 1)  # every env times out -> fused reset of all
    p.pre_physics(torch.zeros(n, env.num_actions, device=DEV))
    p.post_step()
    torch.cuda.synchronize()
    assert int(p.out["reset_buf"].sum()) == n and int(p.state["rng_step"]) == step_before + 1
    assert not torch.equal(a[:, 0:6], env.dof_pos[:, 0:6])


@pytest.mark.parametrize("task", ["BaseTask", "BlindGrasping"])
def test_full_size_properties(task):
    """N = 16 384 (BASELINE configs[3]): permuting the envs permutes every output; the compacted
    env_ids equal nonzero(reset_buf); counters equal the mask sums; the fused episode statistics
    equal a host-side recomputation from the per-step component arrays."""
    n, steps = 16384, 6
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    gen = torch.Generator().manual_seed(5)
    perm = torch.randperm(n, generator=gen)
    e1, e2 = _env(task, n, materialize=False), _env(task, n, materialize=False)
    u = synth.random_uniforms(n, gen)
    ep_sum = torch.zeros(e1.pipeline.params.num_component_rows, n, dtype=torch.float64)
    tot = torch.zeros(4, 2 + ep_sum.shape[0], dtype=torch.float64)
    ep_len = torch.zeros(n, dtype=torch.int64)
    for t in range(steps):
        st = synth.random_state(n, lay, gen, ground_frac=0.01)
        if t == 3:  # force some timeouts
            for e, idx in ((e1, torch.arange(n)), (e2, perm)):
                cnt = torch.where(torch.arange(n) % 7 == 0, e.pipeline.params.max_episode_length - 2, 5)
                e.episode_step_count.copy_(cnt[idx].to(DEV))
        a = synth.random_actions(n, e1.num_actions, gen)
        for e, idx in ((e1, torch.arange(n)), (e2, perm)):
            e.load_state({k: v[idx] for k, v in st.items()})
            if t == 0:
                e.reset(u[idx])
            else:
                e.pre_physics_step(a[idx])
                e.post_physics_step(u[idx])
        torch.cuda.synchronize()
        p1, p2 = e1.pipeline, e2.pipeline
        pd = perm.to(DEV)
        assert torch.equal(p1.out["obs_buf"][pd], p2.out["obs_buf"]), f"step {t} obs"
        assert torch.equal(p1.out["rew_buf"][pd], p2.out["rew_buf"])
        assert torch.equal(p1.out["reset_buf"][pd], p2.out["reset_buf"])
        assert torch.equal(p1.state["full_dof_targets"][pd], p2.state["full_dof_targets"])
        assert torch.equal(p1.state["prev_dof_pos"][:, pd], p2.state["prev_dof_pos"])
        rb = p1.out["reset_buf"]
        assert torch.equal(p1.env_ids(), torch.nonzero(rb).flatten())
        assert torch.equal(rb, p1.out["success"] | p1.out["failure"] | p1.out["timeout"])
        assert int(p1.out["success"].sum() + p1.out["failure"].sum() + p1.out["timeout"].sum()) == int(rb.sum())
        stats = p1.epi["stats"].cpu()
        C = __import__("dexrobot_isaac_b200._lib", fromlist=["CONSTS"]).CONSTS
        assert abs(float(stats[C["DEX_STAT_TIMEOUT_RATE"]]) - float(p1.out["timeout"].float().mean())) < 1e-7
        assert abs(float(stats[C["DEX_STAT_FAILURE_RATE"]]) - float(p1.out["failure"].float().mean())) < 1e-7
        # host recomputation of the RewardComponentObserver accumulation (rl/reward_observer.py:160-293)
        comp = p1.out["reward_components"].double().cpu()
        ep_sum += comp
        ep_len += 1
        done = rb.cpu()
        cls = torch.where(p1.out["success"].cpu(), 1, torch.where(p1.out["failure"].cpu(), 2, 3))
        for c in (0, 1, 2, 3):
            m = done if c == 0 else done & (cls == c)
            tot[c, 0] += m.sum()
            tot[c, 1] += ep_len[m].sum()
            tot[c, 2:] += ep_sum[:, m].sum(dim=1)
        ep_sum[:, done] = 0
        ep_len[done] = 0
        got = p1.stat_totals().cpu()
        assert torch.allclose(got, tot, rtol=1e-6, atol=1e-4), f"step {t} stat_totals"
        if t == 3:
            assert int(p1.out["timeout"].sum()) > 0


@pytest.mark.parametrize("n", [33, 4097, 100003])
def test_fused_epilogue_records(n):
    """The fused step epilogue (tile records polled by the last tile's CTA, DexEpilogueBuffers.tile_records) on a
    single tile + ragged tail, a ragged few-wave grid and a ragged many-wave grid (100 003 envs = 3 126 tiles, more
    than five waves): over several graph-free steps env_ids == nonzero(reset_buf) in ascending order, the rates
    equal the mask means, the running totals equal the summed masks and episode lengths, the Philox / stamp
    counter advances by one per step, and the stand-alone dex_step_epilogue path (REDs + reset words) agrees."""
    task = "BlindGrasping"
    lay = synth.BodyLayout(with_box=True)
    gen = torch.Generator().manual_seed(11)
    e1 = _env(task, n, materialize=False)                       # fused epilogue
    e2 = _env(task, n, materialize=False)                       # post-physics + stand-alone dex_step_epilogue
    e2.pipeline.fuse_epilogue = False
    u = synth.random_uniforms(n, gen)
    C = __import__("dexrobot_isaac_b200._lib", fromlist=["CONSTS"]).CONSTS
    tot = {k: 0 for k in ("success", "failure", "timeout", "reset", "len")}
    for t in range(5):
        st = synth.random_state(n, lay, gen, ground_frac=0.02)
        a = synth.random_actions(n, e1.num_actions, gen)
        for e in (e1, e2):
            e.load_state(st)
            if t == 0:
                e.reset(u)
            else:
                if t == 3:  # force some timeouts
                    cnt = torch.where(torch.arange(n) % 5 == 0, e.pipeline.params.max_episode_length - 2, 7)
                    e.episode_step_count.copy_(cnt.to(DEV))
                e.pre_physics_step(a)
                e.post_physics_step(u)
        torch.cuda.synchronize()
        if t == 0:
            base = {k: int(e1.pipeline.epi["totals"][C[f"DEX_CNT_{k.upper()}"]]) for k in ("success", "failure", "timeout", "reset")}
            base["len"] = int(e1.pipeline.epi["totals"][C["DEX_CNT_EPISODE_LENGTH"]])
            base["steps"] = int(e1.pipeline.epi["totals"][C["DEX_NUM_COUNTERS"]])
            step0 = int(e1.pipeline.state["rng_step"][0])
            continue
        p1, p2 = e1.pipeline, e2.pipeline
        rb = p1.out["reset_buf"]
        assert torch.equal(rb, p2.out["reset_buf"])
        ids = p1.env_ids()
        assert torch.equal(ids, torch.nonzero(rb).flatten()), f"step {t}"
        assert torch.equal(ids, p2.env_ids())
        assert torch.equal(p1.epi["stats"], p2.epi["stats"])
        stats = p1.epi["stats"].cpu()
        for key, name in (("SUCCESS", "success"), ("FAILURE", "failure"), ("TIMEOUT", "timeout")):
            assert abs(float(stats[C[f"DEX_STAT_{key}_RATE"]]) - float(p1.out[name].float().mean())) < 1e-7
            tot[name] += int(p1.out[name].sum())
        tot["reset"] += int(rb.sum())
        # the episode length of a finished env is its step count before the reset zeroed it: recompute from e2's
        # buffers is not possible after the fused reset, so compare the two implementations instead
        assert torch.equal(p1.epi["totals"], p2.epi["totals"]), f"step {t} totals"
        assert int(p1.state["rng_step"][0]) == step0 + t
        if t == 3:
            assert int(p1.out["timeout"].sum()) > 0
    T = e1.pipeline.epi["totals"].cpu()
    for k in ("success", "failure", "timeout", "reset"):
        assert int(T[C[f"DEX_CNT_{k.upper()}"]]) - base[k] == tot[k], k
    assert int(T[C["DEX_NUM_COUNTERS"]]) - base["steps"] == 4


@pytest.mark.parametrize("task", ["BaseTask", "BlindGrasping"])
def test_early_sim_loads_equal_plain(task):
    """StepPipeline.step() without a physics launch sets DEX_FLAG_EARLY_SIM_LOADS (the sim rows are loaded before the
    programmatic-launch wait, overlapping dex_pre_physics): eager and graph-replayed it must give exactly what the
    separately launched pre-/post-physics calls give, with the sim state rewritten before every step."""
    n, steps = 5000, 9
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    e1, e2 = _env(task, n, materialize=False), _env(task, n, materialize=False)
    world = synth.ScriptedWorld(n, lay, seed=21, nact=e1.num_actions)
    u = world.uniforms(0).to(DEV)
    st0 = world.state(0, e1.episode_step_count, None)
    for e in (e1, e2):
        e.load_state(st0)
        e.reset(u)
        e.pipeline.uniforms = u
    act = torch.zeros(n, e2.num_actions, device=DEV)
    g = None
    for t in range(1, steps + 1):
        st = world.state(t, e1.episode_step_count, None)
        a = world.actions(t).to(DEV)
        e1.load_state(st)
        e1.pre_physics_step(a)
        e1.post_physics_step()
        e2.load_state(st)
        act.copy_(a)
        if t < 4:
            e2.pipeline.step(act)
        else:
            if g is None:
                saved = [{k: v.clone() for k, v in d.items() if v is not None}
                         for d in (e2.pipeline.state, e2.pipeline.sim_tensors, e2.pipeline.out, e2.pipeline.epi)]
                g = e2.pipeline.capture_step_graph(act)
                for d, sv in zip((e2.pipeline.state, e2.pipeline.sim_tensors, e2.pipeline.out, e2.pipeline.epi), saved):
                    for k, v in sv.items():
                        d[k].copy_(v)
            e2.pipeline.replay()
        torch.cuda.synchronize()
        for k in ("obs_buf", "rew_buf", "reset_buf", "reward_components", "success", "failure", "timeout"):
            assert torch.equal(e1.pipeline.out[k], e2.pipeline.out[k]), f"step {t} {k}"
        for k, v in e1.pipeline.state.items():
            if k != "rng_step":
                assert torch.equal(v, e2.pipeline.state[k]), f"step {t} state {k}"
        for k, v in e1.pipeline.sim_tensors.items():
            assert torch.equal(v, e2.pipeline.sim_tensors[k]), f"step {t} sim {k}"
        assert torch.equal(e1.pipeline.env_ids(), e2.pipeline.env_ids())


@pytest.mark.parametrize("task,materialize", [("BaseTask", False), ("BlindGrasping", False), ("BlindGrasping", True)])
def test_fused_obs_normalization(task, materialize):
    """DexOutBuffers.obs_norm == rl_games' RunningMeanStd.forward (frozen statistics) of obs_buf,
    to rtol 1e-6 (fp32 subtract, reciprocal multiply, clamp), eager and graph-replayed, with the statistics replaced mid-run
    and obs_buf itself unchanged by the extra output."""
    from oracle.policy_input import RunningMeanStd

    n, steps = 4097, 8
    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    e1, e2 = _env(task, n, materialize=materialize), _env(task, n, materialize=materialize)
    world = synth.ScriptedWorld(n, lay, seed=5, nact=e1.num_actions)
    d = e2.pipeline.spec.obs_dim
    g = torch.Generator().manual_seed(11)
    rms = RunningMeanStd(d)
    rms.running_mean = torch.randn(d, generator=g, dtype=torch.float64) * 0.3
    rms.running_var = torch.rand(d, generator=g, dtype=torch.float64) ** 4 * 2.0  # some columns far below 1: clamp hits
    e2.set_obs_normalization(rms.running_mean, rms.running_var, rms.epsilon, rms.clip)
    u = world.uniforms(0).to(DEV)
    st0 = world.state(0, e1.episode_step_count, None)
    for e in (e1, e2):
        e.load_state(st0)
        e.reset(u)
        e.pipeline.uniforms = u
    act = torch.zeros(n, e2.num_actions, device=DEV)
    captured = False
    clamped = 0
    for t in range(1, steps + 1):
        st = world.state(t, e1.episode_step_count, None)
        a = world.actions(t).to(DEV)
        e1.load_state(st)
        e1.pre_physics_step(a)
        e1.post_physics_step()
        e2.load_state(st)
        act.copy_(a)
        if t == 4:  # new statistics (a training-time update on the host): same buffers, so the graph below sees them
            rms.training = True
            rms.forward(e2.obs_buf.cpu())
            rms.training = False
            e2.set_obs_normalization(rms.running_mean, rms.running_var, rms.epsilon, rms.clip)
        if t < 6:
            e2.pipeline.step(act)
        else:
            if not captured:
                dicts = (e2.pipeline.state, e2.pipeline.sim_tensors, e2.pipeline.out, e2.pipeline.epi)
                saved = [{k: v.clone() for k, v in dd.items() if v is not None} for dd in dicts]
                e2.pipeline.capture_step_graph(act)
                for dd, sv in zip(dicts, saved):
                    for k, v in sv.items():
                        dd[k].copy_(v)
                captured = True
            e2.pipeline.replay()
        torch.cuda.synchronize()
        assert torch.equal(e1.obs_buf, e2.obs_buf), f"step {t}: obs_buf changed by the fused stage"
        assert torch.equal(e1.rew_buf, e2.rew_buf)
        if materialize:  # the privileged-observation instance of the kernel carries the stage too
            assert torch.equal(e1.pipeline.out["obs_dict"], e2.pipeline.out["obs_dict"])
        want = rms.forward(e2.obs_buf.cpu())
        got = e2.obs_norm.cpu()
        # fp32: the kernel multiplies by the reciprocal of sqrt(var + eps) instead of dividing (<= 2 ulp)
        assert torch.allclose(got, want, rtol=1e-6, atol=1e-7), f"step {t}: max diff {(got - want).abs().max()}"
        clamped += int((want.abs() == rms.clip).sum())
    assert clamped > 0  # the clip branch was exercised
    with pytest.raises(ValueError):
        e2.set_obs_normalization(torch.zeros(d + 1), torch.ones(d + 1))


@pytest.mark.parametrize("task,n", [("BlindGrasping", 4096), ("BaseTask", 1000)])
def test_policy_input_training_update(task, n):
    """SURVEY 8f N3, training half (dex_policy_input_update): the running mean / variance / count after every step equal
    rl_games' RunningMeanStd in training mode (oracle/policy_input.py) fed with the same observations, obs_norm equals the
    forward pass with the UPDATED statistics, a CUDA-graph replay of step + update equals the eager twin, and after
    freeze_obs_normalization() the fused stage of dex_post_physics carries on with the statistics reached."""
    from oracle.policy_input import RunningMeanStd

    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    e1, e2 = _env(task, n, materialize=False), _env(task, n, materialize=False)
    d = e1.pipeline.spec.obs_dim
    world = synth.ScriptedWorld(n, lay, seed=9, nact=e1.num_actions)
    u = world.uniforms(0).to(DEV)
    st0 = world.state(0, e1.episode_step_count, None)
    for e in (e1, e2):
        e.load_state(st0)
        e.reset(u)
        e.pipeline.uniforms = u
        e.enable_obs_norm_training(epsilon=1e-5, clip=5.0)
    rms = RunningMeanStd(d)
    rms.training = True
    act = torch.zeros(n, e2.num_actions, device=DEV)
    graph = None
    for t in range(1, 9):
        st = world.state(t, e1.episode_step_count, None)
        a = world.actions(t).to(DEV)
        e1.load_state(st)
        e1.pipeline.step(a.contiguous())
        e1.update_obs_normalization()
        e2.load_state(st)
        act.copy_(a)
        if t < 5:
            e2.pipeline.step(act)
            e2.update_obs_normalization()
        else:
            if graph is None:
                torch.cuda.synchronize()
                dicts = (e2.pipeline.state, e2.pipeline.sim_tensors, e2.pipeline.out, e2.pipeline.epi, e2.pipeline.rms)
                saved = [{k: v.clone() for k, v in dd.items() if v is not None} for dd in dicts]
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    e2.pipeline.step(act)
                    e2.update_obs_normalization()
                for dd, sv in zip(dicts, saved):
                    for k, v in sv.items():
                        dd[k].copy_(v)
            graph.replay()
        torch.cuda.synchronize()
        obs = e1.obs_buf.cpu()
        rms.forward(obs)  # training: update, then normalise
        r = e1.pipeline.rms
        assert float(r["count"]) == float(rms.count) == 1.0 + t * n
        gm, gv = r["running_mean"].cpu(), r["running_var"].cpu()
        # the batch moments enter in fp32 on both sides (torch's fp32 reduction vs float64 sums rounded to fp32): 1 fp32 ulp of
        # the column's magnitude for the mean, of its squared magnitude for the variance
        mag = obs.abs().max(dim=0).values.double()
        assert bool(((gm - rms.running_mean).abs() <= 1e-6 * rms.running_mean.abs() + 2e-7 * mag + 1e-9).all()), f"step {t} mean"
        assert bool(((gv - rms.running_var).abs() <= 1e-5 * rms.running_var.abs() + 4e-7 * mag * mag + 1e-9).all()), f"step {t} var"
        # obs_norm: forward with the statistics the device holds (isolates the element-wise part from the moment rounding)
        chk = RunningMeanStd(d)
        chk.running_mean, chk.running_var = gm, gv
        want = chk.forward(obs)
        got = e1.obs_norm.cpu()
        assert torch.allclose(got, want, rtol=1e-6, atol=1e-6), f"step {t}: max diff {(got - want).abs().max()}"
        # the graph-replayed twin is bit-identical (same kernels, same order; the fp64 atomics of 64 row blocks commute
        # only approximately, so the statistics are compared to 1e-12 relative)
        r2 = e2.pipeline.rms
        assert torch.allclose(r2["running_mean"], r["running_mean"], rtol=1e-12, atol=1e-15)
        assert torch.allclose(r2["running_var"], r["running_var"], rtol=1e-12, atol=1e-15)
        assert torch.allclose(e2.obs_norm, e1.obs_norm, rtol=1e-6, atol=1e-6)
        assert torch.equal(e1.obs_buf, e2.obs_buf)
    # play: frozen statistics, fused stage
    e1.pipeline.freeze_obs_normalization()
    st = world.state(9, e1.episode_step_count, None)
    e1.load_state(st)
    e1.pipeline.step(world.actions(9).to(DEV).contiguous())
    torch.cuda.synchronize()
    chk = RunningMeanStd(d)
    chk.running_mean, chk.running_var = e1.pipeline.rms["running_mean"].cpu(), e1.pipeline.rms["running_var"].cpu()
    assert torch.allclose(e1.obs_norm.cpu(), chk.forward(e1.obs_buf.cpu()), rtol=1e-6, atol=1e-6)
    assert float(e1.pipeline.rms["count"]) == 1.0 + 8 * n  # no update in play
