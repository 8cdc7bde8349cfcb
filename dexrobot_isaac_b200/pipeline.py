"""StepPipeline — owns the persistent path state in HBM and drives libdexstep.so (C ABI, raw pointers).

This is the single object the reference-named component classes (components/*.py) delegate to.
PyTorch is used for device memory, streams and CUDA-graph capture only; every arithmetic op of
the path runs in the hand-written kernels of csrc/dexstep.cu.  There is no eager/CPU fallback:
constructing a StepPipeline on a non-CUDA device raises.

HBM layout:
  * buffers dictated from outside stay row-major AoS exactly as Isaac Gym / the policy hand them
    over: dof_state [N,26,2], rigid_body_state [N,NB,13], net_contact_force [N,NB,3],
    actor_root_state [N,NA,13], actions [N,NACT], obs_buf [N,obs_dim], full_dof_targets [N,26];
  * path-internal persistent state is SoA [K,N] (env index fastest); it is exposed to callers as
    [N,K] strided *views* (``.t()``), so reference-style indexing ``state[env_ids, k]`` works.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, Optional, Sequence

import torch

from . import _lib as L
from .params import NUM_ACTIVE, NUM_DOF, NUM_FINGERS, BodyLayout, PipelineSpec, make_params

C = L.CONSTS
MODE_FULL, MODE_OBS_ONLY = C["DEX_MODE_FULL"], C["DEX_MODE_OBS_ONLY"]
FLAG_FIRST_REWARD_CALL = C["DEX_FLAG_FIRST_REWARD_CALL"]
FLAG_STALE_CONTACTS = C["DEX_FLAG_STALE_CONTACTS"]
FLAG_NO_FUSED_RESET = C["DEX_FLAG_NO_FUSED_RESET"]
FLAG_EARLY_SIM_LOADS = C["DEX_FLAG_EARLY_SIM_LOADS"]
FLAG_TILE_KERNEL = C["DEX_FLAG_TILE_KERNEL"]
FLAG_TMA_STAGING = C["DEX_FLAG_TMA_STAGING"]
FLAG_SINGLE_PASS = C["DEX_FLAG_SINGLE_PASS"]
FLAG_PAIR_GEOMETRY = C["DEX_FLAG_PAIR_GEOMETRY"]
_EARLY_SIM = os.environ.get("DEXSTEP_EARLY_SIM_LOADS", "1") != "0"  # A/B switch for StepPipeline.step()
NUM_UNIFORMS = C["DEX_NUM_UNIFORMS"]


STAT_SLICES = 32  # blocks of DexOutBuffers.stat_totals


def result_segments(n: int, obs_dim: int):
    return [("obs_buf", (n, obs_dim), torch.float32), ("rew_buf", (n,), torch.float32), ("reset_buf", (n,), torch.bool)]


def carve_block(segments, device=None, pin_memory: bool = False):
    """One zeroed uint8 allocation holding every (name, shape, dtype) of ``segments`` at 256-byte aligned offsets
    -> (block, {name: view}).  Host and device blocks carved from the same list have the same offsets, so one copy of the
    block moves all of them."""
    offs, total = [], 0
    for _, shape, dtype in segments:
        offs.append(total)
        nbytes = torch.empty((), dtype=dtype).element_size()
        for d in shape:
            nbytes *= d
        total += (nbytes + 255) // 256 * 256
    block = (torch.zeros(total, dtype=torch.uint8, pin_memory=True) if pin_memory
             else torch.zeros(total, dtype=torch.uint8, device=device))
    views = {}
    for (name, shape, dtype), off in zip(segments, offs):
        nbytes = torch.empty((), dtype=dtype).element_size()
        for d in shape:
            nbytes *= d
        views[name] = block[off:off + nbytes].view(dtype).view(*shape)
    return block, views


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class StepPipeline:
    def __init__(self, cfg: dict, num_envs: int, layout: BodyLayout, dof_lower: Sequence[float],
                 dof_upper: Sequence[float], device="cuda:0", sim: Optional[Dict[str, torch.Tensor]] = None,
                 materialize_obs_dict: bool = False, observer: bool = True, seed: int = 1234,
                 post_action_filters: Sequence[str] = ("velocity_clamp", "position_clamp"),
                 physics_steps_per_control_step: int = 2, fuse_epilogue: bool = True):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise L.DexStepError("StepPipeline needs a CUDA device: this path has no CPU or eager fallback")
        self.lib = L.load()
        self.n = self.num_envs = int(num_envs)
        if self.n <= 0:
            raise ValueError("num_envs must be positive")
        self.layout = layout
        self.cfg = cfg
        self.params, self.spec = make_params(cfg, layout, dof_lower, dof_upper, post_action_filters,
                                             physics_steps_per_control_step)
        self.spec: PipelineSpec
        self.seed = int(seed)
        self.host_step = 0
        self._reward_initialized = False
        self._graph = None
        self.fuse_epilogue = bool(fuse_epilogue)  # step epilogue done by one extra CTA of the post-physics launch
        self.tile_kernel = False  # True: DEX_FLAG_TILE_KERNEL on every post-physics launch (A/B against the v2 kernel)
        self.tma_staging = False  # True: DEX_FLAG_TMA_STAGING (persistent kernel with tensor-copy producers; measured slower)
        # True: DEX_FLAG_SINGLE_PASS - the single-pass kernel where it applies (2-4 tiles per SM): 20 % faster than the
        # persistent one when the sim tensors are L2-resident (post-physics right behind the simulator), 3-5 % slower from HBM
        self.single_pass = False
        self.pair_geometry = False  # True: DEX_FLAG_PAIR_GEOMETRY (two tiles per pass and thread block; measured slower)
        n, dev, sp = self.n, self.device, self.spec
        f32, i64, u8 = torch.float32, torch.int64, torch.uint8

        def z(*shape, dtype=f32):
            return torch.zeros(*shape, dtype=dtype, device=dev)

        # ---- sim views (own storage unless bound to external, e.g. gymtorch-wrapped, tensors) ----
        self.sim_tensors: Dict[str, torch.Tensor] = {}
        self._sim = L.DexSimViews()
        self.bind_sim(sim or {
            "dof_state": z(n, NUM_DOF, 2), "rigid_body_state": z(n, layout.num_bodies, layout.rigid_row_floats),
            "net_contact_force": z(n, layout.contact_rows, 3), "actor_root_state": z(n, layout.num_actors, 13)})

        # ---- persistent state, SoA [K,N] ----
        nact = sp.num_actions
        s = self.state = {
            "active_prev_targets": z(NUM_ACTIVE, n), "active_rule_targets": z(NUM_ACTIVE, n),
            "full_dof_targets": z(n, NUM_DOF), "prev_actions": z(nact, n), "prev_dof_pos": z(NUM_DOF, n),
            "contact_duration_steps": z(NUM_FINGERS, n), "prev_contact_binary": z(NUM_FINGERS, n),
            "contact_forces": z(3 * NUM_FINGERS, n), "episode_step_count": z(n, dtype=i64),
            "prev_finger_dof_vel": z(20, n), "prev_hand_vel": z(3, n), "prev_hand_ang_vel": z(3, n),
            "prev_contacts": z(NUM_FINGERS, n, dtype=torch.bool),
            "rng_step": z(1, dtype=i64),
        }
        if sp.is_blind_grasping:
            s.update({
                "success_duration_steps": z(n, dtype=i64), "success_conditions_met": z(n, dtype=torch.bool),
                "current_stage": z(n, dtype=i64), "time_in_stage": z(n), "stage_contact_duration": z(n),
                "just_transitioned_to_stage2": z(n, dtype=torch.bool),
                "just_transitioned_to_stage3": z(n, dtype=torch.bool),
                "initial_box_positions": z(3, n), "failure_tensors": z(3, n, dtype=torch.bool)})
            s["initial_box_positions"][2].fill_(float(self.params.box_z))  # blind_grasping_task.py:424-428
        self._state = L.DexStateBuffers()
        for name, _ in L.DexStateBuffers._fields_:
            setattr(self._state, name, _ptr(s.get(name)))

        # ---- outputs ----
        rows = self.params.num_component_rows
        # what a step hands back to its caller - obs_buf, rew_buf, reset_buf - lives in ONE allocation, so that a host-side
        # caller fetches it with one device-to-host copy (DexStepEnv.step_host)
        self.result_block, res = carve_block(result_segments(n, sp.obs_dim), device=dev)
        o = self.out = {
            "obs_buf": res["obs_buf"], "obs_dict": z(n, sp.dict_dim) if materialize_obs_dict else None,
            "rew_buf": res["rew_buf"], "reset_buf": res["reset_buf"], "success": z(n, dtype=torch.bool),
            "failure": z(n, dtype=torch.bool), "timeout": z(n, dtype=torch.bool),
            "reward_components": z(rows, n), "reset_bits": z((n + 31) // 32, dtype=torch.int32),
            "counters": z(C["DEX_NUM_COUNTERS"], dtype=torch.int32),
            "episode_sums": z(rows, n) if observer else None,
            "episode_lengths": z(n, dtype=torch.int32) if observer else None,
            # [S][4][2 + rows]: S blocks spread the fp64 atomics of finished episodes (include/dexstep.h, stat_slices);
            # read through stat_totals(), which sums them
            "stat_totals": z(STAT_SLICES, 4, 2 + rows, dtype=torch.float64) if observer else None,
        }
        self._out = L.DexOutBuffers()
        for name, ctype in L.DexOutBuffers._fields_:
            if ctype is ctypes.c_void_p:
                setattr(self._out, name, _ptr(o.get(name)))
        self._out.stat_slices = STAT_SLICES
        e = self.epi = {"env_ids": z(n, dtype=torch.int32), "num_reset": z(1, dtype=torch.int32),
                        "stats": z(C["DEX_NUM_STATS"]), "consecutive_successes": z(1, dtype=torch.int32),
                        "totals": z(C["DEX_NUM_COUNTERS"] + 1, dtype=i64), "tile_records": z((n + 31) // 32, 4, dtype=i64)}
        self._epi = L.DexEpilogueBuffers()
        self._epi.reset_bits, self._epi.counters = _ptr(o["reset_bits"]), _ptr(o["counters"])
        self._epi.env_ids, self._epi.num_reset = _ptr(e["env_ids"]), _ptr(e["num_reset"])
        self._epi.stats, self._epi.consecutive_successes = _ptr(e["stats"]), _ptr(e["consecutive_successes"])
        self._epi.rng_step, self._epi.totals = _ptr(s["rng_step"]), _ptr(e["totals"])
        self._epi.tile_records = _ptr(e["tile_records"])
        self.uniforms: Optional[torch.Tensor] = None  # [N,29] injected reset randoms (None = in-kernel Philox)
        self.launches = 0  # kernels launched through the C ABI (bench.py's gpu_launches)

    # ------------------------------------------------------------------ policy-input stage
    def set_obs_normalization(self, running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor] = None,
                              epsilon: float = 1e-5, clip: float = 5.0):
        """Have dex_post_physics also write ``out["obs_norm"]``: the network input rl_games derives from obs_buf when
        ``normalize_input`` is on (cfg/train/BaseTaskPPO.yaml:44), i.e. RunningMeanStd.forward with frozen statistics:
        ``clamp((obs - running_mean.float()) / sqrt(running_var.float() + epsilon), -clip, clip)``.
        ``running_mean`` / ``running_var`` are rl_games' float64 buffers [obs_dim]; call again whenever they change
        (the values are copied into pipeline-owned buffers, so a captured graph stays valid).  ``None`` switches the
        stage off - only before a graph is captured."""
        o = self.out
        if running_mean is None:
            o["obs_norm"] = None
            self._out.obs_norm = None
            return
        d = self.spec.obs_dim
        if running_mean.numel() != d or running_var is None or running_var.numel() != d:
            raise ValueError(f"running_mean / running_var must have obs_dim = {d} elements")
        # same expressions as RunningMeanStd.forward, evaluated on the device the statistics live on (torch's CPU and
        # CUDA sqrt differ in the last bit for some inputs), then moved
        mean = running_mean.detach().float().reshape(d).to(self.device)
        scale = (1.0 / torch.sqrt(running_var.detach().float().reshape(d) + epsilon)).to(self.device)
        if o.get("obs_norm") is None:
            o["obs_norm"] = torch.zeros(self.n, d, dtype=torch.float32, device=self.device)
            o["norm_mean"], o["norm_scale"] = mean.clone(), scale.clone()
            self._out.obs_norm, self._out.norm_mean = _ptr(o["obs_norm"]), _ptr(o["norm_mean"])
            self._out.norm_scale = _ptr(o["norm_scale"])
        else:
            o["norm_mean"].copy_(mean)
            o["norm_scale"].copy_(scale)
        if self._out.norm_clip != float(clip):
            if getattr(self, "_graph", None) is not None:
                raise RuntimeError("norm_clip is a launch parameter: set it before capture_step_graph()")
            self._out.norm_clip = float(clip)

    def enable_obs_norm_training(self, epsilon: float = 1e-5, clip: float = 5.0, running_mean: Optional[torch.Tensor] = None,
                                 running_var: Optional[torch.Tensor] = None, count: float = 1.0, fused_forward: bool = False):
        """Training mode of the policy-input stage (rl_games RunningMeanStd with ``self.training``): device-resident float64
        running statistics (``self.rms``: running_mean, running_var, count - rl_games' initial values 0 / 1 / 1 unless given)
        that ``update_obs_normalization()`` advances with the batch of observations of the step just computed, after which
        ``out["obs_norm"]`` holds that batch normalised with the UPDATED statistics (dex_policy_input_update).  The fused
        frozen-statistics stage of dex_post_physics stays off in this mode unless ``fused_forward`` (it would normalise with
        the statistics of the previous step and be overwritten); ``freeze_obs_normalization()`` switches to it for play."""
        d, dev, o = self.spec.obs_dim, self.device, self.out
        f64 = torch.float64
        self.rms = {
            "running_mean": (torch.zeros(d, dtype=f64) if running_mean is None else running_mean.detach().to(f64).reshape(d).clone()).to(dev),
            "running_var": (torch.ones(d, dtype=f64) if running_var is None else running_var.detach().to(f64).reshape(d).clone()).to(dev),
            "count": torch.full((1,), float(count), dtype=f64, device=dev),
            "scratch": torch.zeros(2 * d, dtype=f64, device=dev)}
        self._rms_eps, self._rms_clip = float(epsilon), float(clip)
        if o.get("obs_norm") is None:
            o["obs_norm"] = torch.zeros(self.n, d, dtype=torch.float32, device=dev)
            o["norm_mean"] = torch.zeros(d, dtype=torch.float32, device=dev)
            o["norm_scale"] = torch.ones(d, dtype=torch.float32, device=dev)
        o["norm_mean"].copy_(self.rms["running_mean"].float())
        o["norm_scale"].copy_(1.0 / torch.sqrt(self.rms["running_var"].float() + epsilon))
        self._out.norm_mean, self._out.norm_scale = _ptr(o["norm_mean"]), _ptr(o["norm_scale"])
        self._out.norm_clip = float(clip)
        self._out.obs_norm = _ptr(o["obs_norm"]) if fused_forward else None

    def update_obs_normalization(self):
        """dex_policy_input_update on the current obs_buf: batch moments -> merge into the running statistics -> obs_norm
        with the updated statistics.  Three small launches + one memset on the current stream, no host read (capturable)."""
        if getattr(self, "rms", None) is None:
            raise RuntimeError("enable_obs_norm_training() first")
        o, r = self.out, self.rms
        L.check(self.lib.dex_policy_input_update(o["obs_buf"].data_ptr(), self.n, self.spec.obs_dim, r["running_mean"].data_ptr(),
                                                 r["running_var"].data_ptr(), r["count"].data_ptr(), o["norm_mean"].data_ptr(),
                                                 o["norm_scale"].data_ptr(), o["obs_norm"].data_ptr(), self._rms_eps,
                                                 self._rms_clip, r["scratch"].data_ptr(), self._stream()),
                "dex_policy_input_update")
        self.launches += 3

    def freeze_obs_normalization(self):
        """Evaluation / play: stop updating, let dex_post_physics write obs_norm with the statistics reached so far."""
        if getattr(self, "_graph", None) is not None:
            raise RuntimeError("switch before capture_step_graph()")
        self._out.obs_norm = _ptr(self.out["obs_norm"])

    # ------------------------------------------------------------------ sim binding (zero-copy)
    def bind_sim(self, sim: Dict[str, torch.Tensor]):
        """Point the kernels at the caller's sim tensors.  With Isaac Gym these are the
        gymtorch.wrap_tensor views (tensor_manager.py:173-328): no copy is ever made."""
        n, lay = self.n, self.layout
        want = {"dof_state": n * NUM_DOF * 2, "rigid_body_state": n * lay.num_bodies * lay.rigid_row_floats,
                "net_contact_force": n * lay.contact_rows * 3, "actor_root_state": n * lay.num_actors * 13}
        for k, numel in want.items():
            t = sim.get(k)
            if t is None:
                raise L.DexStepError(f"sim tensor '{k}' missing")
            if t.device != self.device or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != numel:
                raise L.DexStepError(f"sim tensor '{k}': need contiguous float32 on {self.device} with {numel} "
                                     f"elements, got {t.dtype} {tuple(t.shape)} on {t.device}")
            self.sim_tensors[k] = t
            setattr(self._sim, k, t.data_ptr())
        self._graph = None

    def use_layout(self, layout: BodyLayout, sim: Dict[str, torch.Tensor]):
        """Switch the body / actor indexing of the sim tensors (they are run-time parameters of the kernels) and bind
        tensors of that layout: used to run the path on a compact host-staging layout (``BodyLayout.compact()``) and to
        switch back.  Everything else in the parameter block is unchanged; a captured graph is invalidated."""
        P = self.params
        P.num_bodies, P.num_contact_bodies, P.num_actors = layout.num_bodies, layout.contact_rows, layout.num_actors
        P.hand_body, P.rigid_row_floats = layout.hand_body, layout.rigid_row_floats
        for f in range(NUM_FINGERS):
            P.tip_body[f], P.pad_body[f], P.contact_body[f] = (layout.tip_bodies[f], layout.pad_bodies[f],
                                                               layout.contact_bodies[f])
        if self.spec.is_blind_grasping:
            P.box_body, P.box_actor = layout.box_body, layout.box_actor
        self.layout = layout
        self.bind_sim(sim)

    # ------------------------------------------------------------------ [N,K] views of SoA state
    def view(self, name: str) -> torch.Tensor:
        t = self.state[name]
        return t.t() if t.dim() == 2 and name != "full_dof_targets" else t

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    # ------------------------------------------------------------------ kernel launches
    def pre_physics(self, actions: torch.Tensor, raw_out: Optional[torch.Tensor] = None):
        """dex_pre_physics: ActionProcessor.process_actions + update_prev_actions (1 launch)."""
        if actions.shape != (self.n, self.spec.num_actions):
            raise RuntimeError(f"actions shape {tuple(actions.shape)} != {(self.n, self.spec.num_actions)}")
        if actions.dtype != torch.float32 or actions.device != self.device or not actions.is_contiguous():
            raise RuntimeError("actions must be a contiguous float32 tensor on the pipeline's device")
        L.check(self.lib.dex_pre_physics(ctypes.byref(self.params), ctypes.byref(self._state), actions.data_ptr(),
                                         _ptr(raw_out), self.n, self._stream()), "dex_pre_physics")
        self.launches += 1

    def post_physics(self, mode: int = MODE_FULL, flags: int = 0, uniforms: Optional[torch.Tensor] = None,
                     fused_epilogue: bool = False):
        """dex_post_physics: StepProcessor.process_physics_step incl. the fused reset (1 launch).
        ``fused_epilogue``: the last CTA also does dex_step_epilogue's work (do not call epilogue() then)."""
        u = uniforms if uniforms is not None else self.uniforms
        if u is not None and (u.shape != (self.n, NUM_UNIFORMS) or u.dtype != torch.float32 or not u.is_contiguous()
                              or u.device != self.device):
            raise RuntimeError(f"uniforms must be contiguous float32 [{self.n},{NUM_UNIFORMS}] on {self.device}")
        if mode == MODE_FULL and not self._reward_initialized:
            flags |= FLAG_FIRST_REWARD_CALL
            self._reward_initialized = True
        if self.tile_kernel:
            flags |= FLAG_TILE_KERNEL
        if self.tma_staging:
            flags |= FLAG_TMA_STAGING
        if self.single_pass:
            flags |= FLAG_SINGLE_PASS
        if self.pair_geometry:
            flags |= FLAG_PAIR_GEOMETRY
        L.check(self.lib.dex_post_physics(ctypes.byref(self.params), ctypes.byref(self._sim), ctypes.byref(self._state),
                                          ctypes.byref(self._out),
                                          ctypes.byref(self._epi) if fused_epilogue and mode == MODE_FULL else None,
                                          self.n, mode, flags, _ptr(u), self.seed, self.host_step, self._stream()),
                "dex_post_physics")
        self.launches += 1
        if fused_epilogue and mode == MODE_FULL:
            self.host_step += 1

    def reset_idx(self, mask: Optional[torch.Tensor] = None, uniforms: Optional[torch.Tensor] = None):
        """dex_reset_idx: masked reset (mask None = all envs)."""
        u = uniforms if uniforms is not None else self.uniforms
        if mask is not None and (mask.shape != (self.n,) or mask.dtype not in (torch.bool, torch.uint8)
                                 or mask.device != self.device):
            raise RuntimeError("reset mask must be a bool/uint8 [N] tensor on the pipeline's device")
        L.check(self.lib.dex_reset_idx(ctypes.byref(self.params), ctypes.byref(self._sim), ctypes.byref(self._state),
                                       _ptr(mask), self.n, _ptr(u), self.seed, self.host_step, self._stream()),
                "dex_reset_idx")
        self.launches += 1

    def set_rng_step(self, step: int):
        """Rewind / restore the device-resident Philox step counter (e.g. when resuming from a checkpoint).  The tile
        records are stamped with it (include/dexstep.h, DexEpilogueBuffers.tile_records): they are re-zeroed so that no
        stale record can pass as current."""
        self.state["rng_step"].fill_(int(step))
        self.epi["tile_records"].zero_()
        self.host_step = int(step)

    def epilogue(self):
        """dex_step_epilogue: env_ids = nonzero(reset_buf), rates, consecutive_successes (1 launch)."""
        L.check(self.lib.dex_step_epilogue(ctypes.byref(self.params), ctypes.byref(self._epi), self.n, self._stream()),
                "dex_step_epilogue")
        self.host_step += 1
        self.launches += 1

    def apply_coupling(self, active_soa: torch.Tensor, full_out: torch.Tensor):
        L.check(self.lib.dex_apply_coupling(ctypes.byref(self.params), active_soa.data_ptr(), full_out.data_ptr(),
                                            active_soa.shape[1], self._stream()), "dex_apply_coupling")
        self.launches += 1

    def extract_active_targets(self, full: torch.Tensor, active_soa_out: torch.Tensor):
        L.check(self.lib.dex_extract_active_targets(ctypes.byref(self.params), full.data_ptr(), active_soa_out.data_ptr(),
                                                    full.shape[0], self._stream()), "dex_extract_active_targets")
        self.launches += 1

    # ------------------------------------------------------------------ whole control step
    def step(self, actions: torch.Tensor, physics=None):
        """pre-physics -> physics(callable, may be None) -> post-physics (+fused reset) -> epilogue."""
        self.pre_physics(actions)
        if physics is not None:
            physics()
        # without a physics launch in between, the kernel ahead of post-physics is pre-physics, which does not touch
        # the sim tensors: their rows may be loaded before the programmatic-launch wait (DEX_FLAG_EARLY_SIM_LOADS)
        self.post_step(flags=FLAG_EARLY_SIM_LOADS if (physics is None and _EARLY_SIM) else 0)

    def post_step(self, uniforms: Optional[torch.Tensor] = None, flags: int = 0):
        """post-physics + step epilogue: one launch when fuse_epilogue, else two."""
        if self.fuse_epilogue:
            self.post_physics(uniforms=uniforms, fused_epilogue=True, flags=flags)
        else:
            self.post_physics(uniforms=uniforms, flags=flags)
            self.epilogue()

    def capture_step_graph(self, actions: torch.Tensor, physics=None):
        """Capture one control step (2-3 launches + the physics callable's work) in a CUDA graph.
        ``actions`` and every bound buffer must keep their addresses.  The first reward call must
        already have happened (its flag differs), and in-kernel Philox uses the device-resident
        step counter so replays draw fresh randoms."""
        if not self._reward_initialized:
            raise RuntimeError("run at least one eager step before capturing (first reward call initialises prev-state)")
        side = torch.cuda.Stream(self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side):
            self.step(actions, physics)  # warm-up on the side stream
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        before = self.launches
        with torch.cuda.graph(g):
            self.step(actions, physics)
        self.launches_per_graph = self.launches - before
        self._graph = g
        return g

    def replay(self):
        if self._graph is None:
            raise RuntimeError("capture_step_graph() first")
        self._graph.replay()
        self.host_step += 1
        self.launches += self.launches_per_graph

    # ------------------------------------------------------------------ results
    def env_ids(self) -> torch.Tensor:
        """== torch.nonzero(reset_buf).flatten() of the last step (host sync: reads the count)."""
        k = int(self.epi["num_reset"].item())
        return self.epi["env_ids"][:k].to(torch.int64)

    def stat_totals(self) -> Optional[torch.Tensor]:
        """The observer's episode totals [4][2 + rows] (classes all / success / failure / timeout): sum of the blocks."""
        t = self.out["stat_totals"]
        return None if t is None else t.sum(dim=0)

    def reward_components(self) -> Dict[str, torch.Tensor]:
        rc = self.out["reward_components"]
        return {name: rc[i] for i, name in enumerate(self.spec.component_names)}

    def obs_dict(self) -> Dict[str, torch.Tensor]:
        d = self.out["obs_dict"]
        if d is None:
            raise RuntimeError("construct the pipeline with materialize_obs_dict=True to read obs_dict")
        from .params import OBS_KEYS_1D
        out = {}
        for k, (a, b) in self.spec.dict_slices.items():
            out[k] = d[:, a] if k in OBS_KEYS_1D else d[:, a:b]
        return out
