"""Runs the REFERENCE's own step-pipeline components (imported unmodified from
/root/reference) on CPU, behind stub ``isaacgym`` / ``gym`` modules and a fake parent that
owns synthetic sim-state tensors.

TEST INFRASTRUCTURE ONLY.  Works only where /root/reference exists (the build container);
it is how the CPU restatement ``oracle/dexstep_oracle.py`` is pinned and how the golden
fixtures under tests/golden/ are generated (``oracle/make_golden.py``).  Nothing here is
reference code: only the wiring that DexHandBase does around the components is re-created
(dexhand_base.py:453-537 component construction, initialization_manager.py:34-172,
dexhand_base.py:238-320 action rules + control-cycle measurement, :743-756 reset_idx,
:805-838 reset, :840-887 pre_physics_step).
"""
from __future__ import annotations

import copy
import os
import sys
from typing import Dict, Optional

import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = os.environ.get("DEXSTEP_REFERENCE_ROOT", "/root/reference")

from . import synth  # noqa: E402


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "dexhand_env"))


def _ensure_paths():
    stubs = os.path.join(HERE, "stubs")
    for p in (REFERENCE_ROOT, stubs):
        if p not in sys.path:
            sys.path.insert(0, p)
    # stubs must win over any real package of the same name
    sys.path.remove(stubs)
    sys.path.insert(0, stubs)
    try:
        from loguru import logger

        logger.remove()
    except Exception:
        pass


def _deep_merge(base: dict, over: dict) -> dict:
    """Hydra-style merge: dicts merge recursively; a dict carrying ``_delete_: true``
    replaces the parent's dict instead of merging (BlindGrasping.yaml:129)."""
    out = copy.deepcopy(base)
    for k, v in over.items():
        if isinstance(v, dict):
            if v.get("_delete_", False):
                out[k] = {kk: copy.deepcopy(vv) for kk, vv in v.items() if kk != "_delete_"}
            elif isinstance(out.get(k), dict):
                out[k] = _deep_merge(out[k], v)
            else:
                out[k] = copy.deepcopy(v)
        else:
            out[k] = copy.deepcopy(v)
    return out


def load_reference_cfg(task_name: str, num_envs: int) -> dict:
    """yaml.safe_load of the reference's own task YAMLs (Hydra is absent here)."""
    cfg_dir = os.path.join(REFERENCE_ROOT, "dexhand_env", "cfg", "task")
    with open(os.path.join(cfg_dir, "BaseTask.yaml")) as f:
        cfg = yaml.safe_load(f)
    cfg.pop("defaults", None)
    if task_name != "BaseTask":
        with open(os.path.join(cfg_dir, f"{task_name}.yaml")) as f:
            over = yaml.safe_load(f)
        over.pop("defaults", None)
        cfg = _deep_merge(cfg, over)
    cfg["env"]["numEnvs"] = num_envs
    return cfg


class _FakeGym:
    """No-op gym; records the three device->PhysX setter calls."""

    def __init__(self):
        self.calls = []
        self.last = {}

    def set_dof_position_target_tensor(self, sim, t):
        self.last["dof_position_target"] = t

    def set_actor_root_state_tensor_indexed(self, sim, root, idx, n):
        self.last["root_state"] = root.clone()
        self.last["root_indices"] = idx.clone()

    def set_dof_state_tensor_indexed(self, sim, dof, idx, n):
        self.last["dof_state"] = dof.clone()
        self.last["dof_indices"] = idx.clone()

    def __getattr__(self, name):
        def _noop(*a, **k):
            return None

        return _noop


class _HandInit:
    pass


class RefEnv:
    """The reference's step pipeline on synthetic state.  Public driver methods:
    ``load_state`` (fake physics), ``reset``, ``pre_physics_step``, ``post_physics_step``."""

    def __init__(self, task_name: str, num_envs: int, layout: Optional[synth.BodyLayout] = None,
                 control_mode: Optional[str] = None, policy_controls_hand_base: bool = True,
                 policy_controls_fingers: bool = True, init_uniforms: Optional[torch.Tensor] = None,
                 cfg: Optional[dict] = None, hooks: Optional[dict] = None, physics_steps_per_control_step: int = 2):
        if not reference_available():
            raise RuntimeError(f"reference not found at {REFERENCE_ROOT}")
        _ensure_paths()
        from dexhand_env.components.action import DefaultActionRules
        from dexhand_env.components.action.action_processor import ActionProcessor
        from dexhand_env.components.initialization.hand_initializer import HardwareMapping
        from dexhand_env.components.initialization.initialization_manager import InitializationManager
        from dexhand_env.components.observation.observation_encoder import ObservationEncoder
        from dexhand_env.components.physics.physics_manager import PhysicsManager
        from dexhand_env.components.physics.tensor_manager import TensorManager
        from dexhand_env.components.reset.reset_manager import ResetManager
        from dexhand_env.components.reward.reward_calculator import RewardCalculator
        from dexhand_env.components.step_processor import StepProcessor
        from dexhand_env.components.termination.termination_manager import TerminationManager
        from dexhand_env.constants import (BASE_JOINT_NAMES, FINGERPAD_BODY_NAMES,
                                           FINGERTIP_BODY_NAMES)
        from dexhand_env.tasks.base_task import BaseTask
        from dexhand_env.tasks.blind_grasping_task import BlindGraspingTask

        self.task_name = task_name
        with_box = task_name == "BlindGrasping"
        self.layout = layout or synth.BodyLayout(with_box=with_box)
        assert self.layout.with_box == with_box
        self.num_envs = num_envs
        self.device = torch.device("cpu")
        self.cfg = cfg if cfg is not None else load_reference_cfg(task_name, num_envs)
        self.env_cfg, self.task_cfg, self.sim_cfg = self.cfg["env"], self.cfg["task"], self.cfg["sim"]
        if control_mode is not None:
            self.task_cfg["controlMode"] = control_mode
        self.task_cfg["policy_controls_hand_base"] = policy_controls_hand_base
        self.task_cfg["policy_controls_fingers"] = policy_controls_fingers
        self.gym, self.sim = _FakeGym(), object()
        self._uniforms = None
        n, lay = num_envs, self.layout

        # --- sim tensors (persistent storage; fake physics copies into them in place) ---
        self.dof_state = torch.zeros(n, synth.NUM_DOF, 2)
        self.rigid_body_states = torch.zeros(n, lay.nb, 13)
        self.net_contact_force = torch.zeros(n * lay.nb, 3)
        self.actor_root_state_tensor = torch.zeros(n, lay.na, 13)
        self.dof_pos = self.dof_state[..., 0]
        self.dof_vel = self.dof_state[..., 1]
        self.num_dof = synth.NUM_DOF
        self.dof_props = synth.dof_props_tensor()

        # --- hand initializer stand-in (hand_initializer.py:105-117, 499-560) ---
        hi = _HandInit()
        hi.dof_names = list(synth.DOF_NAMES)
        hi.fingertip_local_indices = list(lay.tips)
        hi.fingerpad_local_indices = list(lay.pads)
        hi.contact_force_body_names = list(self.task_cfg["contact_force_bodies"])
        mapping = {m.control_name: m.joint_names for m in HardwareMapping}
        hi.active_joint_names = list(mapping.keys())
        hi.joint_to_control = {j: c for c, js in mapping.items() for j in js}
        hi.local_rigid_body_index_to_name = lay.body_names()
        self.hand_initializer = hi
        self.hand_local_rigid_body_index = lay.hand
        self.hand_local_actor_index = 0
        self.contact_force_local_body_indices = list(lay.contacts)
        self.base_joint_names = BASE_JOINT_NAMES
        self.fingertip_body_names = FINGERTIP_BODY_NAMES
        self.fingerpad_body_names = FINGERPAD_BODY_NAMES

        # --- tensor manager assembled by hand (acquisition is Isaac-Gym-only) ---
        tm = TensorManager.__new__(TensorManager)
        tm.parent, tm.gym, tm.sim = self, self.gym, self.sim
        tm.tensors_initialized = True
        tm.dof_state, tm.dof_pos, tm.dof_vel = self.dof_state, self.dof_pos, self.dof_vel
        tm.actor_root_state_tensor = self.actor_root_state_tensor
        tm.num_dof, tm.dof_props = self.num_dof, self.dof_props
        tm.rigid_body_states = self.rigid_body_states
        tm._contact_force_tensor_handle = self.net_contact_force
        tm.contact_forces_all = self.net_contact_force.view(n, lay.nb, 3)
        tm.contact_forces = torch.zeros(n, len(lay.contacts), 3)
        self.tensor_manager = tm
        self.contact_forces = tm.contact_forces

        # --- task (blind_grasping_task.py:368-431 wiring done by hand) ---
        if with_box:
            task = BlindGraspingTask(self.sim, self.gym, self.device, n, self.cfg)
            task.root_state_tensor = self.actor_root_state_tensor
            task.box_actor_indices = torch.arange(n) * lay.na + lay.box_actor
            task.box_local_actor_index = lay.box_actor
            task.box_local_rigid_body_index = lay.box_body
            task.box_states = self.actor_root_state_tensor[:, lay.box_actor, :]
            task.box_positions = task.box_states[:, :3]
            task.box_velocities = task.box_states[:, 7:10]
            task.initial_box_positions = torch.zeros(n, 3)
            task.initial_box_positions[:, 2] = task.box_z
        else:
            task = BaseTask(self.sim, self.gym, self.device, n, self.cfg)
        task.parent_env = self
        task.device = self.device
        self.task = task
        self._patch_task_rng()

        self.physics_dt = self.sim_cfg["dt"]
        self.physics_manager = PhysicsManager(parent=self, physics_dt=self.physics_dt)
        self.action_processor = ActionProcessor(parent=self)
        self.action_processor.initialize_from_config({
            "control_mode": self.task_cfg["controlMode"],
            "num_dof": self.num_dof,
            "policy_controls_hand_base": self.task_cfg["policy_controls_hand_base"],
            "policy_controls_fingers": self.task_cfg["policy_controls_fingers"],
            "finger_vel_limit": self.task_cfg["max_finger_joint_velocity"],
            "base_lin_vel_limit": self.task_cfg["max_base_linear_velocity"],
            "base_ang_vel_limit": self.task_cfg["max_base_angular_velocity"],
            "post_action_filters": ["velocity_clamp", "position_clamp"],
        })
        self.observation_encoder = ObservationEncoder(parent=self)
        self.reset_manager = ResetManager(
            parent=self, dof_state=self.dof_state, root_state_tensor=self.actor_root_state_tensor,
            hand_local_actor_index=self.hand_local_actor_index,
            default_dof_pos=torch.zeros(self.num_dof), task=self.task)
        self.termination_manager = TerminationManager(parent=self, task_cfg=self.task_cfg)
        self.reward_calculator = RewardCalculator(parent=self, task_cfg=self.task_cfg)
        self.initialization_manager = InitializationManager(parent=self)
        self.step_processor = StepProcessor(parent=self)
        self.obs_dict = {}
        self.initialization_manager.create_index_mappings()
        self.initialization_manager.setup_additional_tensors()
        DefaultActionRules.setup_default_action_rule(self.action_processor, self.task_cfg["controlMode"])
        if hooks:  # a custom task's hooks, registered the way DexHandBase._setup_action_rules does (dexhand_base.py:238-262)
            if hooks.get("pre_action_rule") is not None:
                self.action_processor.set_pre_action_rule(hooks["pre_action_rule"])
            if hooks.get("action_rule") is not None:
                self.action_processor.set_action_rule(hooks["action_rule"])
            names = [name for name, _ in hooks.get("filters", [])]
            if names:
                self.action_processor._enabled_post_action_filters = \
                    list(self.action_processor._enabled_post_action_filters) + names
            if hooks.get("extra_reward_terms") is not None:
                base_terms, extra = self.task.compute_task_reward_terms, hooks["extra_reward_terms"]
                self.task.compute_task_reward_terms = lambda od: {**base_terms(od), **extra(od)}
                self.reward_calculator.reward_weights.update(hooks.get("extra_reward_weights", {}))

        # control-cycle measurement (dexhand_base.py:270-320)
        self._uniforms = init_uniforms if init_uniforms is not None else torch.full((n, synth.N_UNIFORMS), 0.5)
        self.physics_manager.start_control_cycle_measurement()
        self.action_processor.process_actions(torch.zeros(n, self.num_actions), torch.zeros(n, 18))
        self.physics_manager.step_physics(refresh_tensors=True)
        # the reference MEASURES the number of physics steps per control step here (physics_manager.py:243-259): one
        # step above + the one reset_idx takes = 2 with the stock integration; an integration whose reset needs more
        # steps measures more - emulated by stepping the (fake) physics again
        for _ in range(physics_steps_per_control_step - 2):
            self.physics_manager.step_physics(refresh_tensors=True)
        self.reset_idx(torch.arange(n))
        self.physics_manager.finish_control_cycle_measurement()
        self.action_processor.finalize_setup()
        if hooks:
            # NOTE (reference quirk): DexHandBase registers the task's custom filters BEFORE finalize_setup()
            # (dexhand_base.py:216/253 vs :307), and finalize_setup() re-creates the filter registry with the two built-ins
            # only (action_processor.py:267 -> rules.py:58-62): a task filter registered through the documented hook is
            # dropped with a "enabled but not registered" warning.  The harness registers it again here, so that the
            # filter the task API documents actually runs; parity of the generic route is defined on that behaviour.
            for name, fn in hooks.get("filters", []):
                self.action_processor.register_post_action_filter(name, fn)
        if hasattr(self.task, "finalize_setup"):
            self.task.finalize_setup()
        self.control_dt = self.physics_manager.control_dt

    # ------------------------------------------------------------------ plumbing
    def _patch_task_rng(self):
        """Route the task's len(env_ids)-sized torch.rand draws (blind_grasping_task.py:453-533)
        to the injected uniforms tensor so that both sides of a parity test see the same randoms."""
        if self.task_name != "BlindGrasping":
            return
        orig = self.task.reset_task_state
        env = self

        def patched(env_ids):
            if len(env_ids) == 0:
                return orig(env_ids)
            u = env._uniforms[env_ids]
            cols = [u[:, 0], u[:, 1], u[:, 2], u[:, 3:6], u[:, 6:9], u[:, 9:29]]
            it = iter(cols)
            real_rand = torch.rand

            def fake_rand(*shape, **kw):
                out = next(it).clone()
                assert tuple(out.shape) == tuple(shape), (out.shape, shape)
                return out

            torch.rand = fake_rand
            try:
                return orig(env_ids)
            finally:
                torch.rand = real_rand

        self.task.reset_task_state = patched

    @property
    def episode_time(self):
        if self.physics_manager.control_dt is None:
            return torch.zeros(self.num_envs)
        return self.episode_step_count.float() * self.physics_manager.control_dt

    def reset_idx(self, env_ids):
        """dexhand_base.py:743-756 minus video recording."""
        if len(env_ids) == 0:
            return
        self.termination_manager.reset_tracking(env_ids)
        self.reset_manager.reset_idx(env_ids)
        self.observation_encoder.reset_observer_state(env_ids)

    # ------------------------------------------------------------------ driver API
    def load_state(self, state: Dict[str, torch.Tensor]):
        """Fake PhysX step: overwrite the sim buffers in place (views stay valid)."""
        self.dof_state.copy_(state["dof_state"])
        self.rigid_body_states.copy_(state["rigid_body_states"])
        self.net_contact_force.copy_(state["net_contact_force"].reshape(-1, 3))
        self.actor_root_state_tensor.copy_(state["actor_root_state"])

    def reset(self, uniforms: Optional[torch.Tensor] = None):
        """dexhand_base.py:805-838."""
        if uniforms is not None:
            self._uniforms = uniforms
        self.reset_idx(torch.arange(self.num_envs))
        obs_dict = self.observation_encoder.compute_observations(exclude_components=["active_rule_targets"])
        state = {"obs_dict": obs_dict, "env": self}
        rule = self.action_processor.apply_pre_action_rule(self.action_processor.active_prev_targets, state)
        obs_dict["active_rule_targets"] = rule
        self.obs_buf = self.observation_encoder.concatenate_observations(obs_dict)
        self.obs_dict = obs_dict
        obs, rew, done, info = self.step_processor.process_physics_step()
        return obs

    def pre_physics_step(self, actions: torch.Tensor):
        """dexhand_base.py:840-887 minus viewer random-actions mode."""
        self.actions = actions.clone()
        self.action_processor.process_actions(actions=self.actions,
                                              active_rule_targets=self.obs_dict["active_rule_targets"])
        self.observation_encoder.update_prev_actions(self.actions)

    def post_physics_step(self, uniforms: Optional[torch.Tensor] = None):
        if uniforms is not None:
            self._uniforms = uniforms
        return self.step_processor.process_physics_step()

    # ------------------------------------------------------------------ snapshots for parity
    def snapshot(self) -> Dict[str, torch.Tensor]:
        """Everything the parity gates compare (SURVEY §8d 'Parity gates')."""
        ap, oe, tmn = self.action_processor, self.observation_encoder, self.termination_manager
        out = {
            "obs_buf": self.obs_buf.clone(), "rew_buf": self.rew_buf.clone(),
            "reset_buf": self.reset_buf.clone(),
            "episode_step_count": self.episode_step_count.clone(),
            "active_prev_targets": ap.active_prev_targets.clone(),
            "full_dof_targets": ap.full_dof_targets.clone(),
            "prev_actions": oe.prev_actions.clone(),
            "prev_dof_pos": oe.prev_dof_pos.clone() if oe.prev_dof_pos is not None else None,
            "contact_duration_steps": oe.contact_duration_steps.clone(),
            "prev_contact_binary": oe.prev_contact_binary.clone(),
            "dof_state": self.dof_state.clone(),
            "actor_root_state": self.actor_root_state_tensor.clone(),
            "active_rule_targets": self.obs_dict["active_rule_targets"].clone(),
        }
        for k, v in self.extras.items():
            if k == "reward_components":
                for ck, cv in v.items():
                    out[f"rc/{ck}"] = cv.clone()
            elif torch.is_tensor(v):
                out[f"extras/{k}"] = v.clone()
            else:
                out[f"extras/{k}"] = torch.tensor(v)
        for k, v in oe.task_states.items():
            out[f"task/{k}"] = v.clone()
        if self.task_name == "BlindGrasping":
            out["task/initial_box_positions"] = self.task.initial_box_positions.clone()
        for k, v in self.obs_dict.items():
            out[f"obs/{k}"] = v.clone()
        return out


class RefObserver:
    """Runs the UNMODIFIED reference RewardComponentObserver (dexhand_env/rl/reward_observer.py) on a RefEnv, behind
    the rl_games stubs of oracle/stubs/rl_games.  `step()` is called once per control step, after
    RefEnv.post_physics_step: after_steps() then process_infos(extras, done ids) - the order the observer's own
    docstrings describe."""

    def __init__(self, env: "RefEnv"):
        _ensure_paths()
        # the module file itself, not the dexhand_env.rl package (whose __init__ pulls in the rl_games training stack)
        import importlib.util

        spec = importlib.util.spec_from_file_location(
            "_ref_reward_observer", os.path.join(REFERENCE_ROOT, "dexhand_env", "rl", "reward_observer.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        RewardComponentObserver = mod.RewardComponentObserver

        class _Vec:
            pass

        class _Algo:
            pass

        self.env = env
        algo = _Algo()
        algo.vec_env = _Vec()
        algo.vec_env.env = env          # needs .num_envs and .extras
        algo.device = "cpu"
        self.obs = RewardComponentObserver(log_interval=10 ** 9)
        self.obs.after_init(algo)

    def step(self):
        self.obs.after_steps()
        done = torch.nonzero(self.env.reset_buf).flatten()
        self.obs.process_infos(self.env.extras, done)

    def totals(self, names):
        """-> [4][2 + len(names)] float64 in the layout of DexOutBuffers.stat_totals ('total' stays 0: the reference skips it)."""
        o = self.obs
        out = torch.zeros(4, 2 + len(names), dtype=torch.float64)
        out[0, 0] = o.total_episodes
        for c, t in enumerate(("success", "failure", "timeout"), start=1):
            out[c, 0] = o.episodes_by_type[t]
        for c, t in enumerate(("all", "success", "failure", "timeout")):
            for r, k in enumerate(names):
                if k in o.cumulative_sums[t]:
                    out[c, 2 + r] = o.cumulative_sums[t][k]["rewards"]
                    out[c, 1] = o.cumulative_sums[t][k]["steps"]
        return out
