"""DexStepEnv — the caller-side wiring DexHandBase does around the step pipeline, without Isaac Gym.

It plays the role of the reference's ``parent`` object (tasks/dexhand_base.py): it owns (or is
bound to) the four sim-state tensors, constructs the reference-named components around one
``StepPipeline`` and exposes the call sequence of dexhand_base.py:

    reset()                 :805-838   reset_idx(all) -> obs pass (stale contact gather) -> full pass
    pre_physics_step(a)     :840-887   ActionProcessor.process_actions + update_prev_actions
    post_physics_step()     :921       StepProcessor.process_physics_step (+ fused reset_idx)
    step(a)                 :893-942   pre -> physics callable -> post

With Isaac Gym present, pass the gymtorch-wrapped tensors as ``sim`` (zero-copy) and a ``physics``
callable that runs gym.simulate/fetch_results/refresh_*; here tests and the bench pass a callable
that writes the next synthetic state.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional

import torch

from . import _lib as L
from .params import BodyLayout
from .pipeline import NUM_UNIFORMS, StepPipeline, carve_block, result_segments

C = L.CONSTS


_COPY_STREAMS: Dict[int, tuple] = {}


def _copy_streams(device: torch.device):
    """One upload and one download stream per device, shared by every env on it (DexStepEnv.step_host)."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _COPY_STREAMS:
        _COPY_STREAMS[idx] = (torch.cuda.Stream(device), torch.cuda.Stream(device))
    return _COPY_STREAMS[idx]


class DexStepEnv:
    def __init__(self, cfg: dict, num_envs: int, layout: Optional[BodyLayout] = None,
                 dof_props: Optional[torch.Tensor] = None, device="cuda:0",
                 sim: Optional[Dict[str, torch.Tensor]] = None, materialize_obs_dict: bool = False,
                 observer: bool = True, seed: int = 1234, init_uniforms: Optional[torch.Tensor] = None,
                 physics: Optional[Callable[[], None]] = None, task=None, physics_steps_per_control_step: int = 2):
        self.cfg, self.env_cfg, self.task_cfg, self.sim_cfg = cfg, cfg["env"], cfg["task"], cfg["sim"]
        self.task_name = self.task_cfg.get("name", "BaseTask")
        self.num_envs = int(num_envs)
        dev = torch.device(device)
        if dev.type == "cuda" and dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        self.layout = layout or BodyLayout.default(with_box=self.task_name == "BlindGrasping")
        if dof_props is None:
            raise ValueError("dof_props [26,6] (lower/upper limits in columns 4,5; tensor_manager.py:242) is required")
        self.dof_props = dof_props.detach().to("cpu", torch.float32)
        self.physics = physics
        # ---- the task (dexhand_base.py:430-452).  The two shipped tasks are facades over the fused kernels; any other
        # DexTask is a GENERIC task: its Python hooks run un-fused on the device tensors, which needs the
        # obs_dict materialised by the kernel and the reset applied after the hooks have seen the step
        from .tasks import BaseTask as _BaseTask, BlindGraspingTask as _BGTask
        if task is None:
            task = (_BGTask if self.task_name == "BlindGrasping" else _BaseTask)(None, None, dev, self.num_envs, cfg)
        self.task = task
        self.generic_task = type(task) not in (_BaseTask, _BGTask)
        if self.generic_task:
            materialize_obs_dict = True
        self.pipeline = StepPipeline(cfg, num_envs, self.layout, self.dof_props[:, 4].tolist(),
                                     self.dof_props[:, 5].tolist(), device=dev, sim=sim,
                                     materialize_obs_dict=materialize_obs_dict, observer=observer, seed=seed,
                                     physics_steps_per_control_step=physics_steps_per_control_step)
        p = self.pipeline
        self.spec = p.spec
        self.num_actions = p.spec.num_actions
        self.num_observations = p.spec.obs_dim
        self.control_dt = p.spec.control_dt
        # buffers the reference's parent exposes (initialization_manager.py:34-117)
        self.dof_state = p.sim_tensors["dof_state"].view(self.num_envs, 26, 2)
        self.dof_pos, self.dof_vel = self.dof_state[..., 0], self.dof_state[..., 1]
        self.rigid_body_states = p.sim_tensors["rigid_body_state"].view(self.num_envs, self.layout.num_bodies, 13)
        self.contact_forces_all = p.sim_tensors["net_contact_force"].view(self.num_envs, self.layout.num_bodies, 3)
        self.actor_root_state_tensor = p.sim_tensors["actor_root_state"].view(self.num_envs, self.layout.num_actors, 13)
        self.obs_buf, self.rew_buf, self.reset_buf = p.out["obs_buf"], p.out["rew_buf"], p.out["reset_buf"]
        self.episode_step_count = p.state["episode_step_count"]
        self.extras: Dict = {}
        self._zero_mask = torch.zeros(self.num_envs, dtype=torch.bool, device=dev)
        self._actions = torch.zeros(self.num_envs, self.num_actions, device=dev)
        # control-cycle measurement of the reference (dexhand_base.py:270-320) does one full reset with
        # whatever the global RNG yields; here: uniforms 0.5 unless given.
        u = init_uniforms if init_uniforms is not None else torch.full((self.num_envs, NUM_UNIFORMS), 0.5)
        p.reset_idx(None, u.to(dev, torch.float32).contiguous())
        # ---- the reference's components (dexhand_base.py:453-537), facades over the pipeline ----
        from .components import (ActionProcessor, ObservationEncoder, ResetManager, RewardCalculator, StepProcessor,
                                 TerminationManager)
        self.obs_dict: Dict = {}
        self.last_reward_components: Dict = {}
        self.action_processor = ActionProcessor(self)
        self.action_processor.initialize_from_config({
            "control_mode": self.task_cfg["controlMode"], "num_dof": 26,
            "policy_controls_hand_base": self.task_cfg["policy_controls_hand_base"],
            "policy_controls_fingers": self.task_cfg["policy_controls_fingers"],
            "finger_vel_limit": self.task_cfg["max_finger_joint_velocity"],
            "base_lin_vel_limit": self.task_cfg["max_base_linear_velocity"],
            "base_ang_vel_limit": self.task_cfg["max_base_angular_velocity"],
            "post_action_filters": ["velocity_clamp", "position_clamp"]})
        self.observation_encoder = ObservationEncoder(self)
        self.observation_encoder.initialize(observation_keys=self.task_cfg["policy_observation_keys"],
                                            num_actions=self.num_actions, action_processor=self.action_processor)
        self.termination_manager = TerminationManager(self, self.task_cfg)
        self.reward_calculator = RewardCalculator(self, self.task_cfg)
        self.reset_manager = ResetManager(self, self.dof_state, self.actor_root_state_tensor, 0,
                                          torch.zeros(26, device=dev), task=self.task)
        self.step_processor = StepProcessor(self)
        # task wiring, in the reference's order (dexhand_base.py:238-262, 430-452, 322-340)
        self.task.parent_env = self
        self.task.device = dev
        self.task.initialize_task_states()
        self.task.set_tensor_references(self.actor_root_state_tensor)
        ap = self.action_processor
        if self.task.pre_action_rule is not None:
            ap.set_pre_action_rule(self.task.pre_action_rule)
        if self.task.action_rule is not None:
            ap.set_action_rule(self.task.action_rule)
        self.task.register_custom_filters(ap)
        if self.task.post_action_filters:
            ap._enabled_post_action_filters = list(ap._enabled_post_action_filters) + list(self.task.post_action_filters)
        self.action_processor.finalize_setup()
        self.task.finalize_setup()

    # ------------------------------------------------------------------ reference call sequence
    def reset_idx(self, env_ids):
        """dexhand_base.py:743-756 (masked form of the index list)."""
        if len(env_ids) == 0:
            return
        self.reset_manager.reset_idx(env_ids)

    def reset(self, uniforms: Optional[torch.Tensor] = None):
        p = self.pipeline
        if uniforms is not None:
            p.uniforms = uniforms.to(self.device, torch.float32).contiguous()
        self.reset_manager.reset_all()                                      # dexhand_base.py:805-811
        self.observation_encoder.compute_observations(stale_contacts=True)  # :813, no refresh_tensors before it
        self.step_processor.process_physics_step()                          # :815-838
        return self.obs_buf

    def pre_physics_step(self, actions: torch.Tensor):
        """dexhand_base.py:840-887."""
        self.action_processor.process_actions(actions=actions,
                                              active_rule_targets=self.action_processor.active_rule_targets)
        self.observation_encoder.update_prev_actions(self._actions)

    def post_physics_step(self, uniforms: Optional[torch.Tensor] = None):
        p = self.pipeline
        if uniforms is not None:
            p.uniforms = uniforms.to(self.device, torch.float32).contiguous()
        return self.step_processor.process_physics_step()

    def step(self, actions: torch.Tensor):
        self.pre_physics_step(actions)
        if self.physics is not None:
            self.physics()
        return self.post_physics_step()

    # ------------------------------------------------------------------ policy-input stage
    def set_obs_normalization(self, running_mean, running_var=None, epsilon: float = 1e-5, clip: float = 5.0):
        """Fuse rl_games' input normalisation (normalize_input, cfg/train/BaseTaskPPO.yaml:44) into the post-physics
        pass: ``self.obs_norm`` then holds clamp((obs_buf - mean) / sqrt(var + epsilon), -clip, clip) after every
        step, computed from the same on-chip copy of the observation as obs_buf.  See StepPipeline.set_obs_normalization."""
        self.pipeline.set_obs_normalization(running_mean, running_var, epsilon, clip)

    def enable_obs_norm_training(self, epsilon: float = 1e-5, clip: float = 5.0, **kw):
        """Training mode of the policy-input stage (rl_games updates its running mean / variance with every batch before it
        normalises it): see StepPipeline.enable_obs_norm_training / update_obs_normalization / freeze_obs_normalization."""
        self.pipeline.enable_obs_norm_training(epsilon, clip, **kw)

    def update_obs_normalization(self):
        self.pipeline.update_obs_normalization()

    @property
    def obs_norm(self) -> Optional[torch.Tensor]:
        return self.pipeline.out.get("obs_norm")

    # ------------------------------------------------------------------ host-buffer entry (end-to-end path)
    def alloc_host_buffers(self, compact=False):
        """Pinned host mirrors of one step's inputs (sim tensors + actions) and results (obs, rew, reset).

        ``compact=False``: the four Isaac Gym tensors as they are (2 432 B/env for BlindGrasping with 32 bodies).
        ``compact=True``: a host staging layout that holds only the rows the path reads - rigid_body_state
        [N,11,13] (hand, 5 tips, 5 pads), net_contact_force [N,6,3] (5 monitored bodies, box), actor_root_state
        [N,1,13] (box), dof_state [N,26,2] - 904 B/env.  The body indices are run-time parameters of the kernels, so the
        same kernels run on the compact device mirrors (``BodyLayout.compact()``); a host-side simulator fills the
        compact buffers directly, ``fill_host_inputs`` gathers them from full tensors.
        ``compact="packed"``: as above with the rigid rows cut to the floats the path reads (``DexParams.rigid_row_floats``
        = 7): rigid_body_state [N,12,7] - the 13-float hand row in slots 0-1, tip and pad poses (pos3 + quat4) in slots
        2-11 - 668 B/env.

        Every tensor of ``inputs`` is a view of ONE pinned allocation (``inputs["_block"]``), likewise ``outputs``:
        ``step_host`` then moves a step with one copy per direction (on the bench box 20 % faster than one copy per
        tensor when both directions are busy, tools/microbench/pcie_duplex.py)."""
        if compact not in (False, True, "packed"):
            raise ValueError(f"compact must be False, True or 'packed', not {compact!r}")
        n = self.num_envs
        lay = self._full_layout().compact(packed=compact == "packed") if compact else self._full_layout()
        block, inputs = carve_block(self._host_segments(lay), pin_memory=True)
        inputs["_block"] = block
        oblock, outputs = carve_block(result_segments(n, self.obs_buf.shape[1]), pin_memory=True)
        outputs["_block"] = oblock
        return inputs, outputs

    def _host_segments(self, lay):
        n, f32 = self.num_envs, torch.float32
        return [("dof_state", (n, 26, 2), f32), ("rigid_body_state", (n, lay.num_bodies, lay.rigid_row_floats), f32),
                ("net_contact_force", (n, lay.contact_rows, 3), f32), ("actor_root_state", (n, lay.num_actors, 13), f32),
                ("actions", (n, self.num_actions), f32)]

    def _full_layout(self):
        saved = getattr(self, "_host_saved", None)
        return saved[0] if saved is not None else self.layout

    def _full_sim(self):
        saved = getattr(self, "_host_saved", None)
        return saved[1] if saved is not None else self.pipeline.sim_tensors

    def fill_host_inputs(self, host_inputs: Dict[str, torch.Tensor], actions: torch.Tensor):
        """Copy the current device sim state (+ ``actions``) into host buffers of either layout (outside any timed
        region: this stands in for the host-side simulator that would write them)."""
        full, lay = self._full_sim(), self._full_layout()
        if tuple(host_inputs["rigid_body_state"].shape[1:]) == (lay.num_bodies, 13):
            for k, v in full.items():
                host_inputs[k].copy_(v.view(host_inputs[k].shape))
        else:
            rigid, contact, actors = lay.compact_gather_indices()
            n = self.num_envs
            host_inputs["dof_state"].copy_(full["dof_state"].view(n, 26, 2))
            rows = full["rigid_body_state"].view(n, lay.num_bodies, 13)[:, rigid]
            dst = host_inputs["rigid_body_state"]
            if dst.shape[2] == 13:
                dst.copy_(rows)
            else:  # packed: hand row (13 floats) in slots 0-1, the 10 poses behind it
                dst[:, 0:2].view(n, 14)[:, :13].copy_(rows[:, 0])
                dst[:, 2:].copy_(rows[:, 1:, :7])
            host_inputs["net_contact_force"].copy_(full["net_contact_force"].view(n, lay.contact_rows, 3)[:, contact])
            host_inputs["actor_root_state"].copy_(full["actor_root_state"].view(n, lay.num_actors, 13)[:, actors])
        host_inputs["actions"].copy_(actions)

    def step_host(self, host_inputs: Dict[str, torch.Tensor], host_outputs: Dict[str, torch.Tensor]):
        """One control step fed from HOST buffers: H2D of the sim state + actions, the two kernel launches, D2H of
        obs/rew/reset - all asynchronous.  The layout (full, compact or packed) is taken from the buffers' shapes; the first
        call switches the pipeline to device mirrors of that layout (``end_host_mode`` switches back to the env's own sim
        tensors).  Buffers from ``alloc_host_buffers`` move with one copy per direction; plain dicts of pinned tensors
        (no ``"_block"``) with one copy per tensor.

        The three stages run on three streams - uploads on one per-device upload stream, the kernels on the CALLER's
        current stream, downloads on one per-device download stream - chained by events, so that the upload of the next step
        (of this env or of another one) overlaps this step's kernels and download and the link carries at most one copy per
        direction at any time (two uploads at once only share the link and push the downloads out of phase: measured, 36-52 M
        instead of 60 M env-steps/s at 16 384 envs).  ``host_inputs`` must stay untouched until the upload is done and
        ``host_outputs`` are valid once the download is: call ``host_sync()`` (makes the current stream wait for both) and
        synchronise the stream, or synchronise the device.  Returns (h2d_bytes, d2h_bytes) as copied."""
        p = self.pipeline
        shape = tuple(host_inputs["rigid_body_state"].shape[1:])
        mode = (shape, "_block" in host_inputs)
        if mode != getattr(self, "_host_mode", None):
            self.end_host_mode()
            full = self.layout
            lay = full if shape == (full.num_bodies, 13) else full.compact(packed=shape[1] == 7)
            self._host_saved = (full, dict(p.sim_tensors))
            self._host_mode = mode
            self._host_dev_block, dev = carve_block(self._host_segments(lay), device=self.device)
            self._host_dev_actions = dev.pop("actions")
            p.use_layout(lay, dev)
            self._host_ev = {k: torch.cuda.Event() for k in ("in_ready", "in_free", "out_ready", "out_free")}
        up, down = _copy_streams(self.device)
        cur, ev = torch.cuda.current_stream(self.device), self._host_ev
        h2d = d2h = 0
        up.wait_event(ev["in_free"])  # the kernels of this env's previous step have read the device mirrors
        with torch.cuda.stream(up):
            if "_block" in host_inputs:
                self._host_dev_block.copy_(host_inputs["_block"], non_blocking=True)
                h2d = self._host_dev_block.numel()
            else:
                for k, dst in list(p.sim_tensors.items()) + [("actions", self._host_dev_actions)]:
                    dst.copy_(host_inputs[k].view(dst.shape), non_blocking=True)
                    h2d += dst.numel() * dst.element_size()
            ev["in_ready"].record(up)
        cur.wait_event(ev["in_ready"])
        cur.wait_event(ev["out_free"])  # the previous step's results have left the device
        p.step(self._host_dev_actions)
        ev["in_free"].record(cur)
        ev["out_ready"].record(cur)
        down.wait_event(ev["out_ready"])
        with torch.cuda.stream(down):
            if "_block" in host_outputs:
                host_outputs["_block"].copy_(p.result_block, non_blocking=True)
                d2h = p.result_block.numel()
            else:
                for k, src in (("obs_buf", self.obs_buf), ("rew_buf", self.rew_buf), ("reset_buf", self.reset_buf)):
                    host_outputs[k].copy_(src, non_blocking=True)
                    d2h += src.numel() * src.element_size()
            ev["out_free"].record(down)
        return h2d, d2h

    def host_sync(self):
        """Make the current stream wait for the last ``step_host``'s download (and therefore for its upload and kernels):
        after a synchronise of that stream the host outputs are valid and the host inputs may be rewritten."""
        ev = getattr(self, "_host_ev", None)
        if ev is not None:
            torch.cuda.current_stream(self.device).wait_event(ev["out_free"])

    def end_host_mode(self):
        """Back to the env's own (full-layout) sim tensors after ``step_host`` calls."""
        saved = getattr(self, "_host_saved", None)
        if saved is not None:
            torch.cuda.synchronize(self.device)
            self.pipeline.use_layout(*saved)
            self._host_saved = None
            self._host_mode = None
            self._host_dev_block = self._host_dev_actions = self._host_ev = None

    # ------------------------------------------------------------------ extras (step_processor.py:221-232)
    def _update_extras(self):
        self.step_processor._update_extras()

    # ------------------------------------------------------------------ test driver API (oracle/driver.py)
    def load_state(self, state: Dict[str, torch.Tensor]):
        """Fake PhysX step: overwrite the sim buffers in place."""
        t = self.pipeline.sim_tensors
        t["dof_state"].copy_(state["dof_state"].reshape(t["dof_state"].shape))
        rb = state["rigid_body_state"] if "rigid_body_state" in state else state["rigid_body_states"]
        t["rigid_body_state"].copy_(rb.reshape(t["rigid_body_state"].shape))
        t["net_contact_force"].copy_(state["net_contact_force"].reshape(t["net_contact_force"].shape))
        t["actor_root_state"].copy_(state["actor_root_state"].reshape(t["actor_root_state"].shape))

    @property
    def initial_box_positions(self):
        if not self.spec.is_blind_grasping:
            raise AttributeError("initial_box_positions exists only for BlindGrasping")
        return self.pipeline.view("initial_box_positions")

    def snapshot(self) -> Dict[str, torch.Tensor]:
        p = self.pipeline
        out = {"obs_buf": self.obs_buf, "rew_buf": self.rew_buf, "reset_buf": self.reset_buf,
               "episode_step_count": self.episode_step_count, "dof_state": self.dof_state,
               "actor_root_state": self.actor_root_state_tensor}
        for k in ("active_prev_targets", "full_dof_targets", "prev_actions", "prev_dof_pos", "contact_duration_steps",
                  "prev_contact_binary", "active_rule_targets"):
            out[k] = p.view(k)
        for k, v in self.extras.items():
            if k == "reward_components":
                for ck, cv in v.items():
                    out[f"rc/{ck}"] = cv
            else:
                out[f"extras/{k}"] = v
        if self.spec.is_blind_grasping:
            for k in ("success_duration_steps", "success_conditions_met", "current_stage", "time_in_stage",
                      "stage_contact_duration", "just_transitioned_to_stage2", "just_transitioned_to_stage3",
                      "initial_box_positions"):
                out[f"task/{k}"] = p.view(k)
        if p.out["obs_dict"] is not None:
            for k, v in p.obs_dict().items():
                out[f"obs/{k}"] = v
        return {k: v.detach().clone() for k, v in out.items()}
