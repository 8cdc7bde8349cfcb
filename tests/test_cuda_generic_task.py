"""GPU tests of the task-hook facades and of the generic-task route.

  (1) A toy third task - a ``DexTask`` subclass with a custom action rule, pre-action rule, post-action filter and an
      extra reward term (oracle/toy_hooks.py, pinned through the UNMODIFIED reference components by
      tests/test_oracle_vs_reference.py) - runs through ``DexStepEnv`` un-fused on the device and is compared in
      lockstep with the oracle running the same callables.
  (2) The reference's ``DexHandBase`` call sequence replayed by a parent that is NOT DexStepEnv (it owns a StepPipeline
      and builds the component facades itself, like dexhand_base.py:_init_components) gives bit-identical results to
      DexStepEnv, and the task hooks are called in the reference's order.
  (3) The two shipped task facades report the kernel's terms / criteria through the reference's hook signatures.
"""
import pytest
import torch

from oracle import synth, toy_hooks
from oracle.driver import compare_snapshots, penetration_floors, run_lockstep
from dexrobot_isaac_b200.config import load_task_config

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL, ATOL = 1e-5, 1e-7


def _toy_task_class():
    from dexrobot_isaac_b200.tasks import BaseTask

    class ToyTask(BaseTask):
        """The hook surface of task_interface.py:286-405, overridden the way a third-party task would."""
        calls = None

        def _log(self, what):
            if self.calls is not None:
                self.calls.append(what)

        @property
        def pre_action_rule(self):
            return toy_hooks.pre_action_rule

        @property
        def action_rule(self):
            return toy_hooks.action_rule

        @property
        def post_action_filters(self):
            return [toy_hooks.FILTER_NAME]

        def register_custom_filters(self, action_processor):
            self._log("register_custom_filters")
            action_processor.register_post_action_filter(toy_hooks.FILTER_NAME, toy_hooks.deadband_filter)

        def compute_task_reward_terms(self, obs_dict):
            self._log("compute_task_reward_terms")
            return toy_hooks.extra_reward_terms(obs_dict)

        def initialize_task_states(self):
            self._log("initialize_task_states")

        def set_tensor_references(self, root_state_tensor):
            self._log("set_tensor_references")
            super().set_tensor_references(root_state_tensor)

        def finalize_setup(self):
            self._log("finalize_setup")

    return ToyTask


def _toy_cfg(n):
    cfg = load_task_config("BaseTask", num_envs=n)
    cfg["task"]["reward_weights"][toy_hooks.EXTRA_REWARD_NAME] = toy_hooks.EXTRA_REWARD_WEIGHT
    return cfg


def _toy_env(n, layout, calls=None, **kw):
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    cfg = _toy_cfg(n)
    task = _toy_task_class()(None, None, DEV, n, cfg)
    task.calls = calls
    return DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device=DEV, task=task, **kw)


@pytest.mark.parametrize("n,steps,seed", [(64, 330, 301), (1000, 60, 302)])
def test_generic_task_lockstep_vs_oracle_with_hooks(n, steps, seed):
    """The toy task through DexStepEnv (generic route: obs_dict kernel instance + torch hooks + dex_apply_coupling +
    dex_reset_idx) == the oracle with the same hooks, every step, every buffer the snapshot holds."""
    from oracle.dexstep_oracle import OracleEnv

    layout = synth.BodyLayout(with_box=False)
    cfg = _toy_cfg(n)
    ora = OracleEnv(cfg, n, layout, synth.dof_props_tensor())
    ora.hooks = toy_hooks.HOOKS
    cud = _toy_env(n, layout)
    assert cud.generic_task and cud.action_processor.uses_generic_route
    world = synth.ScriptedWorld(n, layout, seed=seed, nact=ora.num_actions)
    bad, seen = [], {"reset": 0, "deadband": 0}
    floors = penetration_floors(cfg)
    prev = {}

    def on_step(t, envs):
        a, b = envs[0].snapshot(), envs[1].snapshot()
        seen["reset"] += int(a["reset_buf"].sum())
        if t > 1 and "apt" in prev:
            seen["deadband"] += int((a["active_prev_targets"] == prev["apt"]).sum())
        prev["apt"] = a["active_prev_targets"].clone()
        if t > 2:
            assert f"rc/{toy_hooks.EXTRA_REWARD_NAME}_weighted" in b, sorted(b)
        for m in compare_snapshots(a, b, RTOL, ATOL, skip_rows=a["reset_buf"].bool(), floors=floors):
            if len(bad) < 20:
                bad.append(f"step {t} {m}")

    run_lockstep([ora, cud], world, steps, on_step)
    assert not bad, "\n".join(bad)
    assert seen["reset"] > 0
    assert seen["deadband"] > 0  # the custom filter did hold some targets


def test_generic_task_hooks_change_the_result():
    """Guard against a route that silently ignores the hooks: the toy task and the plain BaseTask facade give
    different targets and rewards on the same inputs."""
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    n, layout = 32, synth.BodyLayout(with_box=False)
    toy = _toy_env(n, layout)
    plain = DexStepEnv(_toy_cfg(n), n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device=DEV,
                       materialize_obs_dict=True)
    assert not plain.generic_task and not plain.action_processor.uses_generic_route
    world = synth.ScriptedWorld(n, layout, seed=303, nact=toy.num_actions)
    out = []
    run_lockstep([toy, plain], world, 6, lambda t, envs: out.append([e.snapshot() for e in envs]))
    a, b = out[-1]
    assert not torch.equal(a["active_prev_targets"], b["active_prev_targets"])
    assert not torch.equal(a["rew_buf"], b["rew_buf"])
    assert f"rc/{toy_hooks.EXTRA_REWARD_NAME}" in a and f"rc/{toy_hooks.EXTRA_REWARD_NAME}" not in b


class ForeignParent:
    """NOT DexStepEnv: a parent in the shape of the reference's DexHandBase (dexhand_base.py:453-537 builds the
    components with parent=self; :840-887 / :889-960 call them).  It owns a StepPipeline the way the reference's
    TensorManager owns the gym tensors, and nothing of DexStepEnv's private surface (no ``_actions`` buffer, no
    ``_update_extras``)."""

    def __init__(self, cfg, n, layout, task, calls):
        from dexrobot_isaac_b200.components import (ActionProcessor, ObservationEncoder, ResetManager, RewardCalculator,
                                                    StepProcessor, TerminationManager)
        from dexrobot_isaac_b200.params import BodyLayout
        from dexrobot_isaac_b200.pipeline import NUM_UNIFORMS, StepPipeline
        from dexrobot_isaac_b200.tasks import BaseTask, BlindGraspingTask

        self.calls = calls
        self.cfg, self.env_cfg, self.task_cfg = cfg, cfg["env"], cfg["task"]
        self.num_envs, self.device = n, torch.device(DEV)
        self.task = task
        self.generic_task = type(task) not in (BaseTask, BlindGraspingTask)
        lay = BodyLayout.from_synth(layout)
        props = synth.dof_props_tensor()
        self.pipeline = p = StepPipeline(cfg, n, lay, props[:, 4].tolist(), props[:, 5].tolist(), device=self.device,
                                         materialize_obs_dict=True)
        self.layout = lay
        self.control_dt = p.spec.control_dt
        self.num_actions = p.spec.num_actions
        self.dof_state = p.sim_tensors["dof_state"].view(n, 26, 2)
        self.dof_pos, self.dof_vel = self.dof_state[..., 0], self.dof_state[..., 1]
        self.actor_root_state_tensor = p.sim_tensors["actor_root_state"].view(n, lay.num_actors, 13)
        self.obs_buf, self.rew_buf, self.reset_buf = p.out["obs_buf"], p.out["rew_buf"], p.out["reset_buf"]
        self.episode_step_count = p.state["episode_step_count"]
        self.extras, self.obs_dict, self.last_reward_components = {}, {}, {}
        p.reset_idx(None, torch.full((n, NUM_UNIFORMS), 0.5, device=self.device))
        # dexhand_base.py:_init_components order
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
                                          torch.zeros(26, device=self.device), task=task)
        self.step_processor = StepProcessor(self)
        task.parent_env = self
        task.initialize_task_states()
        task.set_tensor_references(self.actor_root_state_tensor)
        ap = self.action_processor
        if task.pre_action_rule is not None:
            ap.set_pre_action_rule(task.pre_action_rule)
        if task.action_rule is not None:
            ap.set_action_rule(task.action_rule)
        task.register_custom_filters(ap)
        if task.post_action_filters:
            ap._enabled_post_action_filters = list(ap._enabled_post_action_filters) + list(task.post_action_filters)
        ap.finalize_setup()
        task.finalize_setup()

    # the driver protocol of oracle/driver.run_lockstep, written with the reference's component calls only
    def load_state(self, state):
        from dexrobot_isaac_b200.env import DexStepEnv
        DexStepEnv.load_state(self, state)

    def reset(self, uniforms=None):
        self.pipeline.uniforms = uniforms.to(self.device, torch.float32).contiguous()
        self.calls.append("reset_manager.reset_all")
        self.reset_manager.reset_all()                                      # dexhand_base.py:805-811
        self.calls.append("observation_encoder.compute_observations")
        self.observation_encoder.compute_observations(stale_contacts=True)  # :813
        self.calls.append("step_processor.process_physics_step")
        self.step_processor.process_physics_step()                          # :815-838

    def pre_physics_step(self, actions):
        self.actions = actions.to(self.device)                              # dexhand_base.py:846: the parent's own tensor
        self.calls.append("action_processor.process_actions")
        self.action_processor.process_actions(actions=self.actions,
                                              active_rule_targets=self.action_processor.active_rule_targets)
        self.calls.append("observation_encoder.update_prev_actions")
        self.observation_encoder.update_prev_actions(self.actions)

    def post_physics_step(self, uniforms=None):
        self.pipeline.uniforms = uniforms.to(self.device, torch.float32).contiguous()
        self.calls.append("step_processor.process_physics_step")
        return self.step_processor.process_physics_step()

    def snapshot(self):
        from dexrobot_isaac_b200.env import DexStepEnv
        self.spec = self.pipeline.spec
        return DexStepEnv.snapshot(self)


def _make_pair(kind, n):
    """(DexStepEnv, ForeignParent, calls of the foreign side) for the same task and config."""
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout
    from dexrobot_isaac_b200.tasks import BaseTask, BlindGraspingTask

    calls = []
    if kind == "toy":
        layout, cfg = synth.BodyLayout(with_box=False), _toy_cfg(n)
        t1, t2 = (_toy_task_class()(None, None, DEV, n, cfg) for _ in range(2))
        t2.calls = calls
    else:
        layout = synth.BodyLayout(with_box=kind == "BlindGrasping")
        cfg = load_task_config(kind, num_envs=n)
        cls = BlindGraspingTask if kind == "BlindGrasping" else BaseTask
        t1, t2 = (cls(None, None, DEV, n, cfg) for _ in range(2))
    own = DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device=DEV, task=t1,
                     materialize_obs_dict=True)
    foreign = ForeignParent(cfg, n, layout, t2, calls)
    return own, foreign, calls, layout


@pytest.mark.parametrize("kind,steps", [("BaseTask", 330), ("BlindGrasping", 330), ("toy", 120)])
def test_reference_call_sequence_with_foreign_parent(kind, steps):
    """The component facades + task facades driven by a parent that is not DexStepEnv, through the reference's
    DexHandBase call sequence: bit-identical to DexStepEnv on every snapshot key, every step."""
    n = 96
    own, foreign, calls, layout = _make_pair(kind, n)
    assert not isinstance(foreign, type(own))
    world = synth.ScriptedWorld(n, layout, seed=311, nact=own.num_actions)
    bad, seen = [], {"reset": 0}

    def on_step(t, envs):
        a, b = envs[0].snapshot(), envs[1].snapshot()
        seen["reset"] += int(a["reset_buf"].sum())
        if set(a) != set(b):
            bad.append(f"step {t} keys differ: {sorted(set(a) ^ set(b))}")
        for k in a:
            if k in b and not torch.equal(a[k], b[k]) and len(bad) < 20:
                bad.append(f"step {t} {k}")

    run_lockstep([own, foreign], world, steps, on_step)
    assert not bad, "\n".join(bad)
    assert seen["reset"] > 0
    # the sequence the foreign parent issued is the reference's (dexhand_base.py:805-838, :840-887, :889-960)
    per_step = ["action_processor.process_actions", "observation_encoder.update_prev_actions",
                "step_processor.process_physics_step"]
    seq = [c for c in calls if "." in c]
    assert seq[:3] == ["reset_manager.reset_all", "observation_encoder.compute_observations",
                       "step_processor.process_physics_step"]
    assert seq[3:6] == per_step and len(seq) == 3 + 3 * steps
    if kind == "toy":
        hooks = [c for c in calls if "." not in c]
        # dexhand_base.py:238-262, 430-452, 322-340: states -> tensor references -> filters -> finalize, then one
        # compute_task_reward_terms per process_physics_step (step_processor.py:183-202)
        assert hooks[:4] == ["initialize_task_states", "set_tensor_references", "register_custom_filters",
                             "finalize_setup"]
        assert hooks[4:] == ["compute_task_reward_terms"] * (steps + 1)
        i = calls.index("compute_task_reward_terms")
        assert calls[i - 1] == "step_processor.process_physics_step"


def test_blind_grasping_facade_hooks_hand_out_the_kernel_results():
    """BlindGraspingTask's hooks (reference signatures) == the oracle's task observations / reward terms / failure
    masks of the same step; task_states alias the kernel-owned state."""
    from oracle.dexstep_oracle import OracleEnv
    from dexrobot_isaac_b200.env import DexStepEnv
    from dexrobot_isaac_b200.params import BodyLayout

    n, steps = 64, 260
    layout = synth.BodyLayout(with_box=True)
    cfg = load_task_config("BlindGrasping", num_envs=n)
    ora = OracleEnv(cfg, n, layout, synth.dof_props_tensor())
    cud = DexStepEnv(cfg, n, BodyLayout.from_synth(layout), synth.dof_props_tensor(), device=DEV,
                     materialize_obs_dict=True)
    task = cud.task
    assert type(task).__name__ == "BlindGraspingTask" and task.parent_env is cud
    assert task.contact_duration_threshold_steps == ora.contact_duration_threshold_steps
    assert task.task_states["current_stage"].data_ptr() == cud.pipeline.view("current_stage").data_ptr()
    world = synth.ScriptedWorld(n, layout, seed=321, nact=ora.num_actions)
    floors = penetration_floors(cfg)
    bad, seen = [], {"fail": 0, "terms": set()}

    def on_step(t, envs):
        snap = envs[0].snapshot()
        keep = ~snap["reset_buf"].bool()
        obs = task.get_task_observations(cud.obs_dict)
        for k, v in obs.items():
            a, b = snap[f"obs/{k}"][keep], v.cpu()[keep]
            if not torch.allclose(a.float(), b.float(), rtol=RTOL, atol=ATOL):
                bad.append(f"step {t} task obs {k}")
        terms = task.compute_task_reward_terms(cud.obs_dict)
        for k, v in terms.items():
            seen["terms"].add(k)
            # the reference (and the oracle) only materialise a stage's terms while some env is in that stage
            a, b = snap.get(f"rc/{k}", torch.zeros(n)), v.cpu()
            tol = ATOL + RTOL * max(floors.get(f"rc/{k}", 0.0), 0.0)
            if not torch.allclose(a, b, rtol=RTOL, atol=tol):
                bad.append(f"step {t} reward term {k}")
        fails = task.check_task_failure_criteria(cud.obs_dict)
        assert set(fails) == {"stage1_pregrasp_failed", "stage2_contact_failed", "stage3_grasp_lost", "box_too_far"}
        any_task_failure = torch.zeros(n, dtype=torch.bool)
        for k in ("stage1_pregrasp_failed", "stage2_contact_failed", "stage3_grasp_lost"):
            any_task_failure |= fails[k].bool().cpu()
        seen["fail"] += int(any_task_failure.sum())
        # a task failure mask implies the env is in the failure class of this step (or succeeded in the same step)
        cls = (snap["extras/failure"].bool() | snap["extras/success"].bool())
        if bool((any_task_failure & ~cls).any()):
            bad.append(f"step {t} failure mask outside the failure class")
        succ = task.check_task_success_criteria(cud.obs_dict)["grasp_lift_success"].cpu().bool()
        if not torch.equal(succ, snap["extras/success"].bool()):
            bad.append(f"step {t} success mask")

    run_lockstep([ora, cud], world, steps, on_step)
    assert not bad, "\n".join(bad[:20])
    assert seen["fail"] > 0 and len(seen["terms"]) >= 10


def test_facades_refuse_what_the_kernel_did_not_compute():
    """a facade never accepts foreign dictionaries silently."""
    cud = _make_pair("BaseTask", 16)[0]
    world = synth.ScriptedWorld(16, synth.BodyLayout(with_box=False), seed=5, nact=cud.num_actions)
    run_lockstep([cud], world, 2, lambda t, envs: None)
    with pytest.raises(NotImplementedError):
        cud.termination_manager.evaluate(cud.episode_step_count, {}, {"my_criterion": cud.reset_buf}, {}, {})
    with pytest.raises(NotImplementedError):
        cud.reward_calculator.compute_total_reward({}, {"my_term": cud.rew_buf})
    common = cud.reward_calculator.compute_common_reward_terms(cud.obs_dict, cud)
    name = next(iter(common))
    with pytest.raises(NotImplementedError):
        cud.reward_calculator.compute_total_reward({name: common[name] + 1.0}, {})
    total, comps = cud.reward_calculator.compute_total_reward(common, {})  # the kernel's own rows are accepted
    assert "total" in comps
