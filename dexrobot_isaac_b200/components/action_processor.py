"""ActionProcessor facade (reference: components/action/action_processor.py:64-760).

``process_actions`` is ONE launch of ``dex_pre_physics``: action rule (default_rules.py:33-112),
post-action filters (rules.py:113-190), coupling 18 -> 26 (action_processor.py:571-614),
``active_prev_targets <- next`` (:342) and ``ObservationEncoder.update_prev_actions``
(observation_encoder.py:287-296) are fused.  The open extension points of the reference
(``set_action_rule`` / ``set_pre_action_rule`` / ``set_coupling_rule`` / custom post filters) take
arbitrary Python callables.  The fused kernel covers the default rules; as soon as a callable is
registered the processor switches to the GENERIC route: the registered rule / filters
run as torch ops on the pipeline's device tensors (same call signatures as rules.py:113-190), the
built-in filters are the two clamps of rules.py:141-190, and the 18 -> 26 coupling is
``dex_apply_coupling``.  Still no host round trip and no CPU fallback - just un-fused.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..params import NUM_ACTIVE, NUM_ACTIVE_FINGER, NUM_BASE, NUM_DOF, coupling_tables


class ActionProcessor:
    def __init__(self, parent):
        self.parent = parent
        self._initialized = False
        self._finalized = False
        self._raw = None  # [18, N] optional debug output (active_raw_targets)
        self.actions: Optional[torch.Tensor] = None
        self.pre_action_rule = None
        self.action_rule = None
        self.coupling_rule = None
        self._enabled_post_action_filters = []
        self._custom_filters = {}

    # ------------------------------------------------------------------ properties of the reference
    @property
    def _p(self):
        return self.parent.pipeline

    @property
    def num_envs(self):
        return self.parent.num_envs

    @property
    def device(self):
        return self.parent.device

    @property
    def control_dt(self):
        return self.parent.control_dt

    @property
    def dof_names(self):
        from ..params import DOF_NAMES
        return list(DOF_NAMES)

    @property
    def active_prev_targets(self) -> torch.Tensor:
        return self._p.view("active_prev_targets")

    @property
    def active_rule_targets(self) -> torch.Tensor:
        return self._p.view("active_rule_targets")

    @property
    def active_next_targets(self) -> torch.Tensor:  # == active_prev_targets after process_actions (:342)
        return self._p.view("active_prev_targets")

    @property
    def active_raw_targets(self) -> Optional[torch.Tensor]:
        return None if self._raw is None else self._raw.t()

    @property
    def full_dof_targets(self) -> torch.Tensor:
        return self._p.state["full_dof_targets"]

    @property
    def current_targets(self) -> torch.Tensor:  # legacy alias used by docs snippets
        return self.full_dof_targets

    # ------------------------------------------------------------------ setup (action_processor.py:181-266)
    def initialize_from_config(self, config: dict):
        """Validates ``config`` (keys of action_processor.py:181-247) against the parameter block the
        pipeline was built with; the kernels' constants are fixed at pipeline construction."""
        if self._finalized:
            raise RuntimeError("initialize_from_config() cannot be called after finalize_setup()")
        P, dev = self._p.params, self.device
        mode = config.get("control_mode", self.parent.task_cfg["controlMode"])
        if mode not in ("position", "position_delta"):
            raise ValueError(f"Invalid control mode: {mode}. Must be 'position' or 'position_delta'")
        want = {"position": 0, "position_delta": 1}[mode]
        if want != P.control_mode:
            raise RuntimeError(f"control_mode '{mode}' differs from the pipeline's parameter block")
        if int(config.get("num_dof", NUM_DOF)) != NUM_DOF:
            raise RuntimeError(f"num_dof must be {NUM_DOF}")
        for key, have in (("policy_controls_hand_base", P.policy_controls_base),
                          ("policy_controls_fingers", P.policy_controls_fingers)):
            if key in config and bool(config[key]) != bool(have):
                raise RuntimeError(f"{key}={config[key]} differs from the pipeline's parameter block")
        filters = list(config.get("post_action_filters", ["velocity_clamp", "position_clamp"]))
        unknown = set(filters) - {"velocity_clamp", "position_clamp"}
        if unknown:
            raise NotImplementedError(f"post action filters {sorted(unknown)} are not part of the fused kernel")
        if ("velocity_clamp" in filters) != bool(P.filter_velocity_clamp) or \
                ("position_clamp" in filters) != bool(P.filter_position_clamp):
            raise RuntimeError("post_action_filters differ from the pipeline's parameter block")
        self._enabled_post_action_filters = filters
        self.action_control_mode = mode
        self.policy_controls_hand_base = bool(P.policy_controls_base)
        self.policy_controls_fingers = bool(P.policy_controls_fingers)
        self.num_actions = int(P.num_actions)
        f32 = dict(dtype=torch.float32, device=dev)
        self.active_lower_limits = torch.tensor(list(P.active_lower), **f32)
        self.active_upper_limits = torch.tensor(list(P.active_upper), **f32)
        self.max_deltas = torch.tensor(list(P.max_deltas), **f32)
        mask = [bool(P.policy_controls_base)] * NUM_BASE + [bool(P.policy_controls_fingers)] * NUM_ACTIVE_FINGER
        self.active_target_mask = torch.tensor(mask, dtype=torch.bool, device=dev)
        # coupling table as (finger_idx, dof_idx, scale) triples (action_processor.py:437-503)
        d2a, scale, _, _ = coupling_tables()
        pairs = sorted((d2a[d] - NUM_BASE, d) for d in range(NUM_BASE, NUM_DOF) if d2a[d] >= 0)
        self.coupling_indices = torch.tensor(pairs, dtype=torch.long, device=dev)
        self.coupling_scales = torch.tensor([scale[d] for _, d in pairs], **f32)
        self._initialized = True

    def finalize_setup(self):
        if not self._initialized:
            raise RuntimeError("initialize_from_config() must be called before finalize_setup()")
        self._finalized = True

    # ------------------------------------------------------------------ hot path
    def apply_pre_action_rule(self, active_prev_targets: torch.Tensor, state: dict) -> torch.Tensor:
        """Identity rule (rules.py:78-95).  The fused post-physics kernel already wrote
        ``active_rule_targets = active_prev_targets``; this returns that buffer."""
        return self.active_rule_targets

    def process_actions(self, actions: torch.Tensor, active_rule_targets: Optional[torch.Tensor] = None) -> bool:
        """action_processor.py:284-352 -> one dex_pre_physics launch."""
        if not self._finalized:
            raise RuntimeError("process_actions() called before finalize_setup()")
        p = self._p
        if actions is None:
            raise RuntimeError("actions is None")
        a = actions.to(self.device, torch.float32)
        if a.shape != (self.num_envs, p.spec.num_actions):
            raise RuntimeError(f"actions shape {tuple(a.shape)} != {(self.num_envs, p.spec.num_actions)}")
        if active_rule_targets is not None:
            own = self.active_rule_targets
            if active_rule_targets.data_ptr() != own.data_ptr():
                if active_rule_targets.shape != own.shape:
                    raise RuntimeError(f"active_rule_targets shape {tuple(active_rule_targets.shape)} != {tuple(own.shape)}")
                own.copy_(active_rule_targets)
        buf = self.action_buffer
        if a.data_ptr() != buf.data_ptr():
            buf.copy_(a)
        self.actions = buf
        if self.uses_generic_route:
            self._process_actions_generic(buf)
        else:
            p.pre_physics(buf, self._raw)
        return True

    @property
    def action_buffer(self) -> torch.Tensor:
        """The fixed-address [N, num_actions] device buffer dex_pre_physics reads (graph-capturable).  The parent may
        own it (``parent._actions``, DexStepEnv does, to copy host actions straight into it); any other parent - e.g. the
        reference's DexHandBase - gets one allocated here."""
        buf = getattr(self.parent, "_actions", None)
        if buf is None:
            buf = getattr(self, "_own_actions", None)
            if buf is None:
                buf = self._own_actions = torch.zeros(self.num_envs, self._p.spec.num_actions, device=self.device)
        return buf

    # ------------------------------------------------------------------ generic route (custom callables)
    @property
    def uses_generic_route(self) -> bool:
        return self.action_rule is not None or self.coupling_rule is not None or bool(self._custom_filters) or \
            any(f not in ("velocity_clamp", "position_clamp") for f in self._enabled_post_action_filters)

    def _default_action_rule(self, prev, rule, actions, config):
        """default_rules.py:33-112 as torch ops (only used when a custom filter / coupling rule forces the generic
        route while the action rule itself is the default)."""
        t = rule.clone()
        lo, hi, md = self.active_lower_limits, self.active_upper_limits, self.max_deltas
        fs = NUM_BASE if config["policy_controls_base"] else 0
        if config["control_mode"] == "position":
            if config["policy_controls_base"]:
                t[:, :NUM_BASE] = (actions[:, :NUM_BASE] + 1.0) * 0.5 * (hi[:NUM_BASE] - lo[:NUM_BASE]) + lo[:NUM_BASE]
            if config["policy_controls_fingers"]:
                a = actions[:, fs:fs + NUM_ACTIVE_FINGER]
                t[:, NUM_BASE:] = (a + 1.0) * 0.5 * (hi[NUM_BASE:] - lo[NUM_BASE:]) + lo[NUM_BASE:]
            return t
        if config["policy_controls_base"]:
            t[:, :NUM_BASE] = prev[:, :NUM_BASE] + actions[:, :NUM_BASE] * md[:NUM_BASE]
        if config["policy_controls_fingers"]:
            t[:, NUM_BASE:] = prev[:, NUM_BASE:] + actions[:, fs:fs + NUM_ACTIVE_FINGER] * md[NUM_BASE:]
        return torch.clamp(t, lo, hi)  # default_rules.py:107-110

    def _process_actions_generic(self, actions: torch.Tensor):
        """action_processor.py:320-343 with the registered callables, on device tensors."""
        p = self._p
        prev, rule = self.active_prev_targets, self.active_rule_targets  # [N,18] views of the SoA state
        config = {"control_mode": self.action_control_mode, "policy_controls_base": self.policy_controls_hand_base,
                  "policy_controls_fingers": self.policy_controls_fingers}
        fn = self.action_rule if self.action_rule is not None else self._default_action_rule
        raw = fn(prev, rule, actions, config)
        if self._raw is not None:
            self._raw.copy_(raw.t())
        nxt = raw
        for name in self._enabled_post_action_filters:  # rules.py:113-139
            if name == "velocity_clamp":
                nxt = prev + torch.clamp(nxt - prev, -self.max_deltas, self.max_deltas)
            elif name == "position_clamp":
                nxt = torch.clamp(nxt, self.active_lower_limits, self.active_upper_limits)
            elif name in self._custom_filters:
                nxt = self._custom_filters[name](prev, rule, nxt)
        nxt_soa = nxt.t().contiguous()
        if self.coupling_rule is not None:
            p.state["full_dof_targets"].copy_(self.coupling_rule(nxt))
        else:
            p.apply_coupling(nxt_soa, p.state["full_dof_targets"])
        p.state["active_prev_targets"].copy_(nxt_soa)                       # :342
        p.state["prev_actions"].copy_(actions.t())                          # observation_encoder.py:287-296

    def enable_raw_targets(self, on: bool = True):
        """Materialise ``active_raw_targets`` (debug attribute of the reference) on the next calls."""
        self._raw = torch.zeros(NUM_ACTIVE, self.num_envs, device=self.device) if on else None

    def reset_targets(self, env_ids=None, dof_positions: Optional[torch.Tensor] = None):
        """action_processor.py:524-568: targets of the given envs from the current (or given) DOF positions:
        ``full_dof_targets[ids] = dof_pos[ids]`` and the inverse coupling into the 18 active targets."""
        p = self._p
        pos = self.parent.dof_pos if dof_positions is None else dof_positions.to(self.device, torch.float32)
        if env_ids is None:
            full = pos.contiguous()
            p.state["full_dof_targets"].copy_(full)
            p.extract_active_targets(full, p.state["active_prev_targets"])
            return
        ids = torch.as_tensor(env_ids, device=self.device, dtype=torch.long)
        if ids.numel() == 0:
            return
        rows = pos[ids] if pos.shape[0] == self.num_envs else pos
        p.state["full_dof_targets"][ids] = rows
        act = torch.empty(NUM_ACTIVE, ids.numel(), device=self.device)
        p.extract_active_targets(rows.contiguous(), act)
        p.state["active_prev_targets"][:, ids] = act

    def apply_coupling(self, active_targets: torch.Tensor) -> torch.Tensor:
        """action_processor.py:571-614 ([M,18] -> [M,26]) through dex_apply_coupling."""
        t = active_targets.to(self.device, torch.float32)
        out = torch.empty(t.shape[0], NUM_DOF, device=self.device)
        self._p.apply_coupling(t.t().contiguous(), out)
        return out

    def extract_active_targets_from_full_dof(self, full_dof_positions: torch.Tensor) -> torch.Tensor:
        """action_processor.py:616-666 ([M,26] -> [M,18]); last DOF of a coupling group wins."""
        f = full_dof_positions.to(self.device, torch.float32).contiguous()
        out = torch.empty(NUM_ACTIVE, f.shape[0], device=self.device)
        self._p.extract_active_targets(f, out)
        return out.t()

    def unscale_actions(self, actions: torch.Tensor) -> torch.Tensor:
        """action_processor.py:722-760: [-1,1] actions -> physical units of the controlled targets."""
        P = self._p.params
        lo, hi, md = self.active_lower_limits, self.active_upper_limits, self.max_deltas
        cols = ([*range(NUM_BASE)] if P.policy_controls_base else []) + \
               ([*range(NUM_BASE, NUM_ACTIVE)] if P.policy_controls_fingers else [])
        idx = torch.tensor(cols, device=self.device)
        a = actions.to(self.device, torch.float32)
        if P.control_mode == 0:
            return (a + 1.0) * 0.5 * (hi[idx] - lo[idx]) + lo[idx]
        return a * md[idx]

    # ------------------------------------------------------------------ extension points (rules.py:196-244)
    def set_pre_action_rule(self, rule):
        """rule(active_prev_targets, state) -> active_rule_targets; None restores the identity rule.  Applied by
        StepProcessor after the post-physics launch (step_processor.py:66-73), on device tensors."""
        self.pre_action_rule = rule

    def set_action_rule(self, rule):
        self.action_rule = rule

    def set_coupling_rule(self, rule):
        self.coupling_rule = rule

    def register_post_action_filter(self, name, filter_fn):
        self._custom_filters[name] = filter_fn
