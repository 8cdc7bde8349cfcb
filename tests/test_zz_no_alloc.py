"""dex_post_physics_noalloc (include/dexstep.h): the fused-epilogue post-physics launch with its tile-record scratch in a
static __device__ array of the library instead of a caller-allocated buffer.  No counterpart in the reference.
Collected LAST on purpose (file name).

Called through the C ABI (ctypes) with DexEpilogueBuffers.tile_records = NULL and compared with the existing entry point
(dex_post_physics with caller-provided, zero-initialised records) on identical twins, several calls in a row: every
output, state and epilogue buffer bit for bit (the observer's float64 atomics to summation-order tolerance, as in
tests/epilogue_check.py)."""
import ctypes

import pytest
import torch

from oracle import synth
from dexrobot_isaac_b200.config import load_task_config

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _twins(task, n, obs_dict):
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    lay = synth.BodyLayout(with_box=task == "BlindGrasping")
    cfg = load_task_config(task, num_envs=n)
    envs = [DexStepEnv(cfg, n, BodyLayout.from_synth(lay), synth.dof_props_tensor(), device=DEV,
                       materialize_obs_dict=obs_dict) for _ in range(2)]
    return lay, envs


def _noalloc_post(p, n_envs=None):  # (honours p.single_pass like StepPipeline.post_physics does)
    """One dex_post_physics_noalloc call with the pipeline's own buffers; tile_records is NOT passed."""
    from dexrobot_isaac_b200 import _lib as L

    ep = L.DexEpilogueBuffers()
    ctypes.memmove(ctypes.byref(ep), ctypes.byref(p._epi), ctypes.sizeof(ep))
    ep.tile_records = None
    rc = p.lib.dex_post_physics_noalloc(ctypes.byref(p.params), ctypes.byref(p._sim), ctypes.byref(p._state),
                                        ctypes.byref(p._out), ctypes.byref(ep), p.n if n_envs is None else n_envs,
                                        L.CONSTS["DEX_MODE_FULL"], L.CONSTS["DEX_FLAG_SINGLE_PASS"] if p.single_pass else 0, None,
                                        p.seed, p.host_step, p._stream())
    if rc == 0:
        p.host_step += 1
    return rc


def _same(p1, p2, tag):
    for name, d1, d2 in (("out", p1.out, p2.out), ("state", p1.state, p2.state), ("sim", p1.sim_tensors, p2.sim_tensors)):
        for k, v in d1.items():
            if v is None:
                continue
            if k == "stat_totals":  # float64 atomics over all CTAs: order-dependent in the last bits by construction
                assert torch.allclose(v, d2[k], rtol=1e-9, atol=1e-9), f"{tag} {name}/{k}"
            else:
                assert torch.equal(v, d2[k]), f"{tag} {name}/{k}"
    for k in ("num_reset", "stats", "consecutive_successes", "totals"):
        assert torch.equal(p1.epi[k], p2.epi[k]), f"{tag} epi/{k}"
    ids = p1.env_ids()
    assert torch.equal(ids, p2.env_ids()), f"{tag} env_ids"
    assert torch.equal(ids, torch.nonzero(p1.out["reset_buf"]).flatten()), f"{tag} env_ids vs reset_buf"


def _compare(task, n, steps, obs_dict, seed, single_pass=False):
    lay, (e1, e2) = _twins(task, n, obs_dict)
    e1.pipeline.single_pass = e2.pipeline.single_pass = single_pass
    gen = torch.Generator().manual_seed(seed)
    u = synth.random_uniforms(n, gen)
    st = synth.random_state(n, lay, gen, ground_frac=0.03)
    for e in (e1, e2):
        e.load_state(st)
        e.reset(u)
        e.pipeline.uniforms = None  # in-kernel Philox from here on
    torch.cuda.synchronize()
    e2.pipeline.epi["tile_records"].zero_()  # (reset() above went through the existing entry point)
    resets = 0
    for t in range(steps):
        st = synth.random_state(n, lay, gen, ground_frac=0.03)
        a = synth.random_actions(n, e1.num_actions, gen).to(DEV).contiguous()
        for e in (e1, e2):
            e.load_state(st)
            e.pipeline.pre_physics(a)
        e1.pipeline.post_step()                      # existing entry point, caller-allocated records
        assert _noalloc_post(e2.pipeline) == 0       # new entry point, records inside the library
        torch.cuda.synchronize()
        _same(e1.pipeline, e2.pipeline, f"{task} n={n} step {t}")
        resets += int(e1.pipeline.out["reset_buf"].sum())
    assert resets > 0
    # the caller-side record buffer of the twin was never touched: the scratch really is the library's own array
    assert int(e2.pipeline.epi["tile_records"].ne(0).sum()) == 0
    return e2


@pytest.mark.parametrize("task,n,obs_dict,single", [("BlindGrasping", 16384, False, True), ("BlindGrasping", 16384, False, False),
                                                    ("BaseTask", 5000, False, False), ("BlindGrasping", 1000, True, False),
                                                    ("BlindGrasping", 262144, False, False)],
                         ids=["single_pass_16384", "persistent_16384", "persistent_ragged_5000", "tile_kernel_obs_dict_1000",
                              "capacity_262144"])
def test_noalloc_equals_existing_entry_point(task, n, obs_dict, single):
    """Every kernel family behind dex_post_physics, 4 calls in a row (the capacity case 2), bit for bit."""
    _compare(task, n, 2 if n > 100000 else 4, obs_dict, seed=41, single_pass=single)


def test_noalloc_shared_scratch_keeps_no_stale_records():
    """All calls share one array that keeps the last call's contents.  Two pipelines of the same size run one after the
    other from the same step count (so their record stamps coincide) on different states, then interleaved call by call:
    a record left by the other pipeline must never pass as current."""
    _compare("BlindGrasping", 4096, 3, False, seed=51)
    _compare("BlindGrasping", 4096, 3, False, seed=52)
    # interleaved: A and B alternate on the shared array, each against its own twin on the existing entry point
    pairs = []
    for seed in (61, 62):
        lay, (e1, e2) = _twins("BlindGrasping", 2048, False)
        gen = torch.Generator().manual_seed(seed)
        u, st = synth.random_uniforms(2048, gen), synth.random_state(2048, lay, gen, ground_frac=0.05)
        for e in (e1, e2):
            e.load_state(st)
            e.reset(u)
            e.pipeline.uniforms = None
        pairs.append((lay, gen, e1, e2))
    for t in range(4):
        for lay, gen, e1, e2 in pairs:
            st = synth.random_state(2048, lay, gen, ground_frac=0.05)
            a = synth.random_actions(2048, e1.num_actions, gen).to(DEV).contiguous()
            for e in (e1, e2):
                e.load_state(st)
                e.pipeline.pre_physics(a)
            e1.pipeline.post_step()
            assert _noalloc_post(e2.pipeline) == 0
        torch.cuda.synchronize()
        for i, (_, _, e1, e2) in enumerate(pairs):
            _same(e1.pipeline, e2.pipeline, f"interleaved pipeline {i} step {t}")


def test_noalloc_rejects_sizes_beyond_the_maximum_and_bad_arguments():
    from dexrobot_isaac_b200 import _lib as L

    _, (e1, _) = _twins("BlindGrasping", 64, False)
    p = e1.pipeline
    cap = L.CONSTS["DEX_NOALLOC_MAX_ENVS"]
    assert cap == 262144
    launches = p.lib.dex_debug_v2_launches()
    rc = _noalloc_post(p, n_envs=cap + 1)  # rejected before any launch (the buffers here hold 64 envs only)
    assert rc == L.CONSTS["DEX_ERR_CAPACITY"] and b"DEX_NOALLOC_MAX_ENVS" in p.lib.dex_strerror(rc)
    assert p.lib.dex_debug_v2_launches() == launches
    # epilogue is required (it is what the scratch is for); OBS_ONLY mode has no epilogue
    a = (ctypes.byref(p.params), ctypes.byref(p._sim), ctypes.byref(p._state), ctypes.byref(p._out))
    assert p.lib.dex_post_physics_noalloc(*a, None, p.n, L.CONSTS["DEX_MODE_FULL"], 0, None, 1, 0, p._stream()) == L.CONSTS["DEX_ERR_NULL"]
    assert p.lib.dex_post_physics_noalloc(*a, ctypes.byref(p._epi), p.n, L.CONSTS["DEX_MODE_OBS_ONLY"], 0, None, 1, 0,
                                          p._stream()) == L.CONSTS["DEX_ERR_PARAM"]
    torch.cuda.synchronize()


def test_noalloc_allocates_nothing_and_runs_under_graph_capture():
    """Free device memory (cudaMemGetInfo) does not move across calls, and the call is legal inside stream capture (an
    allocating or synchronising CUDA call would invalidate the capture): pre-physics + the no-alloc post-physics captured
    in one CUDA graph and replayed == the existing entry point launched eagerly."""
    n = 4096
    lay, (e1, e2) = _twins("BlindGrasping", n, False)
    gen = torch.Generator().manual_seed(71)
    u, st = synth.random_uniforms(n, gen), synth.random_state(n, lay, gen, ground_frac=0.03)
    for e in (e1, e2):
        e.load_state(st)
        e.reset(u)
        e.pipeline.uniforms = None
    act = torch.zeros(n, e1.num_actions, device=DEV)
    states = [synth.random_state(n, lay, gen, ground_frac=0.03) for _ in range(4)]
    actions = [synth.random_actions(n, e1.num_actions, gen).to(DEV) for _ in range(4)]

    def eager_existing(t):
        e1.load_state(states[t])
        e1.pipeline.pre_physics(actions[t].contiguous())
        e1.pipeline.post_step()

    # step 0 eagerly on both sides (first call: looks the array's address up)
    eager_existing(0)
    e2.load_state(states[0])
    act.copy_(actions[0])
    e2.pipeline.pre_physics(act)
    assert _noalloc_post(e2.pipeline) == 0
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    # step 1 eagerly: no allocation
    eager_existing(1)
    e2.load_state(states[1])
    act.copy_(actions[1])
    e2.pipeline.pre_physics(act)
    assert _noalloc_post(e2.pipeline) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0  # free memory did not shrink: nothing was allocated
    _same(e1.pipeline, e2.pipeline, "eager step 1")
    # steps 2, 3: captured once, replayed twice
    e2.load_state(states[2])
    act.copy_(actions[2])
    g = torch.cuda.CUDAGraph()
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        e2.pipeline.pre_physics(act)
        assert _noalloc_post(e2.pipeline) == 0
    for t in (2, 3):
        eager_existing(t)
        e2.load_state(states[t])
        act.copy_(actions[t])
        g.replay()
        torch.cuda.synchronize()
        _same(e1.pipeline, e2.pipeline, f"graph replay step {t}")
