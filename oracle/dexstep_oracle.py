"""CPU restatement (torch, fp32, eager) of the reference's per-step tensor pipeline.

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  This is the *checker*, never the product: only
tests/, ``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference``
legs may import it.  The product path (dexrobot_isaac_b200) never does and fails loudly when
its CUDA library is missing.

Why it exists: the reference is pure Python/torch and can be imported in the build container
(``oracle/ref_harness.py``) but /root/reference does not exist on the GPU box, so the parity
tests there need a self-contained statement of the same algorithm.  PINNING: this file is
checked, step by step over 520-step scripted trajectories of both tasks, against the
reference's own components executed by ``ref_harness.RefEnv`` (tests/test_oracle_vs_reference.py,
bit-exact masks/counters, <=1e-6 floats) and against the golden fixtures under tests/golden/
that were produced by the reference (oracle/make_golden.py).  The quaternion helpers are pinned
by the 7 known-answer cases of the reference's utils/test_coordinate_transforms.py.

Each function cites the reference file:line it follows (paths relative to
/root/reference/dexhand_env/).  torch ops are used on purpose (``torch.norm``, true division,
``torch.var``) so that CPU numerics are the reference's numerics.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch

NB_BASE = 6           # constants.py:8
N_ACTIVE_FINGER = 12  # constants.py:9
N_FINGER_DOF = 20     # constants.py:10
N_ACTIVE = NB_BASE + N_ACTIVE_FINGER
N_UNIFORMS = 29

BASE_JOINTS = ["ARTx", "ARTy", "ARTz", "ARRx", "ARRy", "ARRz"]              # constants.py:14
FINGER_JOINTS = [f"r_f_joint{f}_{j}" for f in range(1, 6) for j in range(1, 5)]  # constants.py:16-37
DOF_NAMES = BASE_JOINTS + FINGER_JOINTS

# constants.py:71-88  (finger control -> [(joint, scale)]), action order
COUPLING = {
    0: [("r_f_joint1_1", 1.0)], 1: [("r_f_joint1_2", 1.0)],
    2: [("r_f_joint1_3", 1.0), ("r_f_joint1_4", 1.0)],
    3: [("r_f_joint2_1", 1.0), ("r_f_joint4_1", 1.0), ("r_f_joint5_1", 2.0)],
    4: [("r_f_joint2_2", 1.0)], 5: [("r_f_joint2_3", 1.0), ("r_f_joint2_4", 1.0)],
    6: [("r_f_joint3_2", 1.0)], 7: [("r_f_joint3_3", 1.0), ("r_f_joint3_4", 1.0)],
    8: [("r_f_joint4_2", 1.0)], 9: [("r_f_joint4_3", 1.0), ("r_f_joint4_4", 1.0)],
    10: [("r_f_joint5_2", 1.0)], 11: [("r_f_joint5_3", 1.0), ("r_f_joint5_4", 1.0)],
}
# hand_initializer.py:20-34 (HardwareMapping, enum order) -> primary joint per control
HARDWARE_PRIMARY = ["r_f_joint1_3", "r_f_joint1_2", "r_f_joint1_1", "r_f_joint2_1", "r_f_joint2_3",
                    "r_f_joint2_2", "r_f_joint3_3", "r_f_joint3_2", "r_f_joint4_3", "r_f_joint4_2",
                    "r_f_joint5_3", "r_f_joint5_2"]


# ------------------------------------------------------------------ quaternion helpers
# isaacgym.torch_utils (third party, absent): same formulas/op order as oracle/stubs.
def quat_mul(a, b):
    x1, y1, z1, w1 = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    x2, y2, z2, w2 = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    ww = (z1 + x1) * (x2 + y2)
    yy = (w1 - y1) * (w2 + z2)
    zz = (w1 + y1) * (w2 - z2)
    xx = ww + yy + zz
    qq = 0.5 * (xx + (z1 - x1) * (x2 - y2))
    w = qq - ww + (z1 - y1) * (y2 - z2)
    x = qq - xx + (x1 + w1) * (x2 + w2)
    y = qq - yy + (w1 - x1) * (y2 + z2)
    z = qq - zz + (z1 + y1) * (w2 - x2)
    return torch.stack([x, y, z, w], dim=-1)


def quat_conjugate(q):
    return torch.cat((-q[..., :3], q[..., 3:]), dim=-1)


def quat_rotate_inverse(q, v):
    """v(2w^2-1) - 2w(q_v x v) + 2 q_v (q_v . v)   (utils/coordinate_transforms.py:33)."""
    n = q.shape[0]
    q_w, q_vec = q[:, -1], q[:, :3]
    a = v * (2.0 * q_w**2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.reshape(n, 1, 3), v.reshape(n, 3, 1)).squeeze(-1) * 2.0
    return a - b + c


def quat_rotate(q, v):
    n = q.shape[0]
    q_w, q_vec = q[:, -1], q[:, :3]
    a = v * (2.0 * q_w**2 - 1.0).unsqueeze(-1)
    b = torch.cross(q_vec, v, dim=-1) * q_w.unsqueeze(-1) * 2.0
    c = q_vec * torch.bmm(q_vec.reshape(n, 1, 3), v.reshape(n, 3, 1)).squeeze(-1) * 2.0
    return a + b + c


def point_in_hand_frame(p_world, hand_pos, hand_quat):
    """utils/coordinate_transforms.py:17-35."""
    return quat_rotate_inverse(hand_quat, p_world - hand_pos)


def point_in_world_frame(p_local, hand_pos, hand_quat):
    """utils/coordinate_transforms.py:38-56."""
    return quat_rotate(hand_quat, p_local) + hand_pos


# ------------------------------------------------------------------ derived constants
def coupling_table():
    """action_processor.py:437-503 -> (finger_idx[19], dof_idx[19], scale[19]), spread DOF idx."""
    name_to_idx = {n: i for i, n in enumerate(DOF_NAMES)}
    f_idx, d_idx, sc = [], [], []
    for f in range(N_ACTIVE_FINGER):
        for joint, s in COUPLING[f]:
            f_idx.append(f)
            d_idx.append(name_to_idx[joint])
            sc.append(s)
    return (torch.tensor(f_idx), torch.tensor(d_idx), torch.tensor(sc, dtype=torch.float32),
            name_to_idx["r_f_joint3_1"])


def active_finger_dof_indices():
    """observation_encoder.py:496-574: primary DOF of each control in HardwareMapping order."""
    name_to_idx = {n: i for i, n in enumerate(DOF_NAMES)}
    return torch.tensor([name_to_idx[j] for j in HARDWARE_PRIMARY])


class OracleEnv:
    """Same driver API as ref_harness.RefEnv: load_state / reset / pre_physics_step /
    post_physics_step / snapshot."""

    def __init__(self, cfg: dict, num_envs: int, layout, dof_props: torch.Tensor,
                 init_uniforms: Optional[torch.Tensor] = None, physics_steps_per_control_step: int = 2):
        self.cfg = cfg
        self.env_cfg, self.task_cfg, self.sim_cfg = cfg["env"], cfg["task"], cfg["sim"]
        self.task_name = self.task_cfg["name"]
        self.is_bg = self.task_name == "BlindGrasping"
        self.n = self.num_envs = num_envs
        self.layout = layout
        n, tc = num_envs, self.task_cfg
        self.dof_props = dof_props.clone().float()
        self.control_dt = None

        # sim buffers
        self.dof_state = torch.zeros(n, 26, 2)
        self.rigid_body_states = torch.zeros(n, layout.nb, 13)
        self.contact_forces_all = torch.zeros(n, layout.nb, 3)
        self.actor_root_state = torch.zeros(n, layout.na, 13)
        self.dof_pos = self.dof_state[..., 0]
        self.dof_vel = self.dof_state[..., 1]
        self.contact_forces = torch.zeros(n, 5, 3)

        # ---- action processor state (action_processor.py:181-247) ----
        self.control_mode = tc["controlMode"]
        self.ctrl_base = bool(tc["policy_controls_hand_base"])
        self.ctrl_fingers = bool(tc["policy_controls_fingers"])
        self.num_actions = (NB_BASE if self.ctrl_base else 0) + (N_ACTIVE_FINGER if self.ctrl_fingers else 0)
        self.lower = self.dof_props[:, 4].clone()
        self.upper = self.dof_props[:, 5].clone()
        self.active_prev_targets = torch.zeros(n, N_ACTIVE)
        self.full_dof_targets = torch.zeros(n, 26)
        self.cpl_f, self.cpl_d, self.cpl_s, self.spread_idx = coupling_table()
        # action_processor.py:380-413: limits of the first joint of each coupling group
        name_to_idx = {nm: i for i, nm in enumerate(DOF_NAMES)}
        first = [name_to_idx[COUPLING[f][0][0]] for f in range(N_ACTIVE_FINGER)]
        self.active_lower = torch.cat([self.lower[:NB_BASE], self.lower[first]])
        self.active_upper = torch.cat([self.upper[:NB_BASE], self.upper[first]])
        # action_processor.py:491-503 inverse map
        self.inv_finger = torch.full((26,), -1, dtype=torch.long)
        self.inv_scale = torch.ones(26)
        for i in range(len(self.cpl_f)):
            self.inv_finger[self.cpl_d[i]] = self.cpl_f[i]
            self.inv_scale[self.cpl_d[i]] = self.cpl_s[i]
        self.max_deltas = None
        self.finalized = False
        self.active_rule_targets = torch.zeros(n, N_ACTIVE)
        self.active_raw_targets = None
        self.active_next_targets = None

        # ---- observation encoder state (observation_encoder.py:119-242) ----
        self.obs_keys: List[str] = list(tc["policy_observation_keys"])
        self.act_idx = active_finger_dof_indices()
        self.prev_actions = torch.zeros(n, self.num_actions)
        self.prev_dof_pos = torch.zeros(n, 26)  # set by the init dry-run on an all-zero sim
        self.contact_duration_steps = torch.zeros(n, 5)
        self.prev_contact_binary = torch.zeros(n, 5)
        self.contact_thr = tc["contact_binary_threshold"]

        self.rew_buf = torch.zeros(n)
        self.reset_buf = torch.zeros(n, dtype=torch.bool)
        self.episode_step_count = torch.zeros(n, dtype=torch.long)
        self.obs_dict: Dict[str, torch.Tensor] = {}
        self.extras: Dict = {}

        # ---- termination manager (termination_manager.py:27-77) ----
        term = tc.get("termination", {})
        self.active_success = term.get("active_success_criteria", [])
        self.active_failure = term.get("active_failure_criteria", [])
        self.success_reasons: Dict[str, torch.Tensor] = {}
        self.failure_reasons: Dict[str, torch.Tensor] = {}
        self.max_episode_length = self.env_cfg["episodeLength"]
        rw = tc["reward_weights"]
        self.reward_weights = rw
        self.success_reward = rw["termination_success"]
        self.failure_penalty = rw["termination_failure_penalty"]
        self.timeout_penalty = rw["termination_timeout_penalty"]
        self.consecutive_successes = 0
        self.max_consecutive_successes = tc["max_consecutive_successes"]

        # ---- reward calculator state (reward_calculator.py:49-53) ----
        self.prev_finger_dof_vel = None
        self.prev_hand_vel = None
        self.prev_hand_ang_vel = None
        self.prev_contacts = None

        # ---- BlindGrasping task (blind_grasping_task.py:114-257) ----
        if self.is_bg:
            box = self.env_cfg["box"]
            self.box_size, self.box_z = box["size"], box["initial_position"]["z"]
            self.box_xy_range = box["initial_position"]["xy_range"]
            self.box_pos = self.actor_root_state[:, layout.box_actor, 0:3]
            self.box_vel = self.actor_root_state[:, layout.box_actor, 7:10]
            self.initial_box_positions = torch.zeros(n, 3)
            self.initial_box_positions[:, 2] = self.box_z
            self.ts = {
                "success_duration_steps": torch.zeros(n, dtype=torch.long),
                "success_conditions_met": torch.zeros(n, dtype=torch.bool),
                "current_stage": torch.zeros(n, dtype=torch.long),
                "time_in_stage": torch.zeros(n),
                "stage_contact_duration": torch.zeros(n),
                "just_transitioned_to_stage2": torch.zeros(n, dtype=torch.bool),
                "just_transitioned_to_stage3": torch.zeros(n, dtype=torch.bool),
            }
            self.failure_tensors = None
            self.contact_duration_threshold_steps = None
        else:
            self.ts = {}
        self._uniforms = init_uniforms if init_uniforms is not None else torch.full((n, N_UNIFORMS), 0.5)

        # ---- control-cycle measurement (dexhand_base.py:270-320): dummy process_actions before
        # finalize zeroes the targets (action_processor.py:305-318); one full reset; 2 physics
        # steps were taken -> control_dt = 2 * sim.dt (physics_manager.py:243-259)
        self.full_dof_targets.zero_()
        self.reset_idx(torch.arange(n))
        self.control_dt = self.sim_cfg["dt"] * physics_steps_per_control_step
        # action_processor.py:416-434
        self.max_deltas = torch.zeros(N_ACTIVE)
        self.max_deltas[0:3] = self.control_dt * tc["max_base_linear_velocity"]
        self.max_deltas[3:6] = self.control_dt * tc["max_base_angular_velocity"]
        self.max_deltas[6:] = self.control_dt * tc["max_finger_joint_velocity"]
        self.finalized = True
        if self.is_bg:  # blind_grasping_task.py:259-296
            self.contact_duration_threshold_steps = int(tc["contact_duration_threshold"] / self.control_dt)
            self._reset_task_state(torch.arange(n))
            self.failure_tensors = {k: torch.zeros(n, dtype=torch.bool) for k in
                                    ("stage1_pregrasp_failed", "stage2_contact_failed", "stage3_grasp_lost")}

    # ================================================================== fake physics
    def load_state(self, state):
        self.dof_state.copy_(state["dof_state"])
        self.rigid_body_states.copy_(state["rigid_body_states"])
        self.contact_forces_all.copy_(state["net_contact_force"].reshape(self.n, -1, 3))
        self.actor_root_state.copy_(state["actor_root_state"])

    # ================================================================== pre-physics (P1-P7)
    def apply_coupling(self, active):
        """action_processor.py:571-614: 18 -> 26 gather*scale scatter, joint3_1 := 0."""
        full = torch.zeros(active.shape[0], 26)
        full[:, :NB_BASE] = active[:, :NB_BASE]
        full[:, self.cpl_d] = active[:, NB_BASE:][:, self.cpl_f] * self.cpl_s.unsqueeze(0)
        full[:, self.spread_idx] = 0.0
        return full

    def extract_active_targets(self, full):
        """action_processor.py:616-666 (inverse coupling).  The reference scatters with DUPLICATE column indices
        (`active[:, 6 + finger_indices] = scaled`, :657-659: three spread DOFs map to one control, two DOFs to each
        dip control), which torch documents as undefined order: single-threaded (small batches) the last DOF in index
        order wins, but a large reset batch is split over threads and other DOFs of the group can win for some rows
        (seen at 4096 envs when ~100 envs time out in one step).  The oracle states the deterministic rule - the LAST
        DOF in index order wins - which is what the reference does whenever its scatter runs sequentially and what
        tests/test_oracle_vs_reference.py pins; the CUDA kernels implement the same rule (DexParams.inverse_src_dof)."""
        out = torch.zeros(full.shape[0], N_ACTIVE)
        out[:, :NB_BASE] = full[:, :NB_BASE]
        fi, sc = self.inv_finger[NB_BASE:], self.inv_scale[NB_BASE:]
        for j in range(fi.numel()):  # ascending DOF order: later DOFs overwrite earlier ones
            if int(fi[j]) >= 0:
                out[:, NB_BASE + int(fi[j])] = full[:, NB_BASE + j] / sc[j]
        return out

    def _action_rule(self, prev, rule, actions):
        t = rule.clone()
        fs = NB_BASE if self.ctrl_base else 0
        if self.control_mode == "position":
            # default_rules.py:33-64 + scaling.py:28-45
            if self.ctrl_base:
                lo, hi = self.active_lower[:6], self.active_upper[:6]
                t[:, :6] = (actions[:, :6] + 1.0) * 0.5 * (hi - lo) + lo
            if self.ctrl_fingers:
                lo, hi = self.active_lower[6:], self.active_upper[6:]
                t[:, 6:] = (actions[:, fs:fs + 12] + 1.0) * 0.5 * (hi - lo) + lo
            return t
        # default_rules.py:80-112 + scaling.py:48-80
        if self.ctrl_base:
            t[:, :6] = prev[:, :6] + actions[:, :6] * self.max_deltas[:6]
        if self.ctrl_fingers:
            t[:, 6:] = prev[:, 6:] + actions[:, fs:fs + 12] * self.max_deltas[6:]
        return torch.clamp(t, self.active_lower, self.active_upper)

    def process_actions(self, actions, rule_targets):
        """action_processor.py:284-352 (post-finalize branch) + rules.py:113-190.  `self.hooks` (optional, tests of the
        generic-task route) carries custom callables with the reference's hook signatures (rules.py:196-244)."""
        hooks = getattr(self, "hooks", None) or {}
        self.actions = actions.clone()
        self.active_rule_targets = rule_targets
        if hooks.get("action_rule") is not None:
            config = {"control_mode": self.control_mode, "policy_controls_base": self.ctrl_base,
                      "policy_controls_fingers": self.ctrl_fingers}                # action_processor.py:322-326
            raw = hooks["action_rule"](self.active_prev_targets, rule_targets, actions, config)
        else:
            raw = self._action_rule(self.active_prev_targets, rule_targets, actions)
        self.active_raw_targets = raw
        delta = raw - self.active_prev_targets                                    # rules.py:141-166
        nxt = self.active_prev_targets + torch.clamp(delta, -self.max_deltas, self.max_deltas)
        nxt = torch.clamp(nxt, self.active_lower, self.active_upper)              # rules.py:168-190
        for _, fn in hooks.get("filters", []):                                    # rules.py:113-139, task filters last
            nxt = fn(self.active_prev_targets, rule_targets, nxt)
        self.active_next_targets = nxt
        self.full_dof_targets = self.apply_coupling(nxt)
        self.active_prev_targets = nxt.clone()

    def pre_physics_step(self, actions):
        """dexhand_base.py:840-887."""
        self.process_actions(actions.clone(), self.obs_dict["active_rule_targets"])
        self.prev_actions = actions.clone()                                       # observation_encoder.py:287-296

    # ================================================================== observations (O1-O8, T1-T3)
    def _hand_frame(self, poses_world):
        """observation_encoder.py:915-986."""
        n = self.n
        hand_pos = self.rigid_body_states[:, self.layout.hand, :3]
        hand_quat = self.rigid_body_states[:, self.layout.hand, 3:7]
        p = poses_world.view(n, 5, 7)
        pos = p[:, :, :3].reshape(-1, 3)
        quat = p[:, :, 3:7].reshape(-1, 4)
        hp = hand_pos.unsqueeze(1).expand(-1, 5, -1).reshape(-1, 3)
        hq = hand_quat.unsqueeze(1).expand(-1, 5, -1).reshape(-1, 4)
        pos_h = point_in_hand_frame(pos, hp, hq)
        quat_h = quat_mul(quat_conjugate(hq), quat)
        return torch.cat([pos_h.reshape(n, 5, 3), quat_h.reshape(n, 5, 4)], dim=2).reshape(n, -1)

    def _default_observations(self):
        """observation_encoder.py:576-758."""
        n, lay = self.n, self.layout
        od = {}
        dof_pos = self.dof_pos
        # O2  observation_encoder.py:298-321
        if self.control_dt is None:
            self.prev_dof_pos = dof_pos.clone()
            dof_vel = torch.zeros_like(dof_pos)
        else:
            dof_vel = (dof_pos - self.prev_dof_pos) / self.control_dt
            self.prev_dof_pos = dof_pos.clone()
        od["base_dof_pos"] = dof_pos[:, :NB_BASE]
        od["base_dof_vel"] = dof_vel[:, :NB_BASE]
        od["active_finger_dof_pos"] = dof_pos[:, self.act_idx]
        od["active_finger_dof_vel"] = dof_vel[:, self.act_idx]
        od["all_finger_dof_pos"] = dof_pos[:, NB_BASE:].clone()
        od["all_finger_dof_vel"] = dof_vel[:, NB_BASE:].clone()
        hand_pose = self.rigid_body_states[:, lay.hand, :7]
        od["hand_pose"] = hand_pose
        # O6  observation_encoder.py:1483-1521
        s = 0.7071067811865476
        inv = quat_conjugate(torch.tensor([0.0, s, 0.0, s]))
        q_al = quat_mul(hand_pose[:, 3:7], inv.unsqueeze(0).expand(n, -1))
        od["hand_pose_arr_aligned"] = torch.cat([hand_pose[:, :3], q_al], dim=1)
        od["contact_forces"] = self.contact_forces.reshape(n, -1)
        od["prev_actions"] = self.prev_actions
        od["active_prev_targets"] = self.active_prev_targets
        od["base_dof_target"] = self.full_dof_targets[:, :NB_BASE]
        od["active_finger_dof_target"] = self.full_dof_targets[:, self.act_idx]
        od["all_finger_dof_target"] = self.full_dof_targets[:, NB_BASE:].clone()
        mag = torch.norm(self.contact_forces, dim=2)
        od["contact_force_magnitude"] = mag
        binary = mag > self.contact_thr
        od["contact_binary"] = binary.float()
        # O4  observation_encoder.py:323-361
        bf = binary.float()
        started = (bf == 1.0) & (self.prev_contact_binary == 0.0)
        active = bf == 1.0
        self.contact_duration_steps = torch.where(
            started, torch.ones_like(self.contact_duration_steps),
            torch.where(active, self.contact_duration_steps + 1, torch.zeros_like(self.contact_duration_steps)))
        self.prev_contact_binary = bf.clone()
        if self.control_dt is not None:
            od["contact_duration"] = self.contact_duration_steps * self.control_dt
        else:
            od["contact_duration"] = torch.zeros_like(bf)
        tips_w = self.rigid_body_states[:, lay.tips, :7].reshape(n, 35)
        pads_w = self.rigid_body_states[:, lay.pads, :7].reshape(n, 35)
        od["fingertip_poses_world"] = tips_w
        od["fingertip_poses_hand"] = self._hand_frame(tips_w)
        od["fingerpad_poses_world"] = pads_w
        od["fingerpad_poses_hand"] = self._hand_frame(pads_w)
        if self.control_dt is None:
            od["episode_time"] = torch.zeros(n, 1)
        else:  # dexhand_base.py:703-712
            od["episode_time"] = (self.episode_step_count.float() * self.control_dt).unsqueeze(1)
        return od

    def _finger_box_contacts(self, od):
        """blind_grasping_task.py:773-833."""
        lay = self.layout
        box_mag = torch.norm(self.contact_forces_all[:, lay.box_body, :], dim=1)
        box_has = box_mag > self.contact_thr
        pads = self.rigid_body_states[:, lay.pads, :3]
        dist = torch.norm(pads - self.box_pos.unsqueeze(1), dim=2)
        near = dist < (math.sqrt(3) * self.box_size / 2 * 1.2)
        return od["contact_binary"].bool() & box_has.unsqueeze(1) & near

    def _task_observations(self, od):
        """blind_grasping_task.py:549-653 (+ :684-771, :835-978)."""
        if not self.is_bg:
            return {}
        n, tc = self.n, self.task_cfg
        to = {"object_pos": self.box_pos, "object_vel": self.box_vel}
        pads = od["fingerpad_poses_world"].view(n, 5, 7)[:, :, :3]
        # spatial relationships :684-724
        d = torch.norm(pads - self.box_pos.unsqueeze(1).expand(-1, 5, -1), dim=2)
        to["finger_to_object_distances"] = d
        to["avg_finger_to_object_distance"] = torch.mean(d, dim=1)
        hd = torch.abs(pads[:, :, 2] - self.box_pos[:, 2].unsqueeze(1))
        to["finger_to_object_height_diff"] = hd
        to["avg_finger_to_object_height_diff"] = torch.mean(hd, dim=1)
        to["hand_to_object_distance"] = torch.norm(od["hand_pose"][:, :3] - self.box_pos, dim=1)
        # geometry :726-771
        iu = torch.triu_indices(5, 5, offset=1)
        to["fingerpad_distances"] = torch.norm(pads[:, iu[0]] - pads[:, iu[1]], dim=2)
        to["first_three_fingerpad_centroid"] = torch.mean(pads[:, :3, :], dim=1)
        # contact state :668-682
        fbc = self._finger_box_contacts(od)
        thumb = fbc[:, 0]
        others = fbc[:, 1:].any(dim=1)
        grasp = thumb & others
        to["thumb_contact"] = thumb.unsqueeze(1).float()
        to["other_fingers_contact"] = others.unsqueeze(1).float()
        to["grasp_state"] = grasp.unsqueeze(1).float()
        if self.control_dt is None:  # init dry-run :609-623
            to["grasp_duration"] = torch.zeros(n, 1)
            to["current_stage"] = torch.ones(n, 1)
            to["time_in_stage"] = torch.zeros(n, 1)
            to["stage_progress"] = torch.zeros(n, 1)
            return to
        # success duration :626-648
        fbc = self._finger_box_contacts(od)
        cond = (self.box_pos[:, 2] > tc["success_height_threshold"]) & (fbc.sum(dim=1) >= tc["min_fingers_for_grasp"])
        sds = self.ts["success_duration_steps"]
        sds[cond] += 1
        sds[~cond] = 0
        self.ts["success_conditions_met"][:] = cond
        to["grasp_duration"] = (sds.float() * self.control_dt).unsqueeze(1)
        # stage FSM :835-899
        stage, tis, scd = self.ts["current_stage"], self.ts["time_in_stage"], self.ts["stage_contact_duration"]
        j2, j3 = self.ts["just_transitioned_to_stage2"], self.ts["just_transitioned_to_stage3"]
        tis += self.control_dt
        cb = od["contact_binary"].bool()
        thumb_other = cb[:, 0] & cb[:, 1:].any(dim=1)
        m2 = stage == 2
        scd[m2] = torch.where(thumb_other[m2], scd[m2] + self.control_dt, torch.zeros_like(scd[m2]))
        # transitions :901-956
        j2.fill_(False)
        j3.fill_(False)
        s1_done = (stage == 1) & (tis >= tc["stage1_duration"])
        s2_done = ((stage == 2) & (scd >= tc["stage_evaluation"]["stage2_contact_success_threshold"])) | \
                  ((stage == 2) & (tis >= tc["stage2_duration"]))
        j2[s1_done] = True
        j3[s2_done] = True
        stage[s1_done] = 2
        stage[s2_done] = 3
        tis[s1_done | s2_done] = 0.0
        scd[s1_done | s2_done] = 0.0
        to["current_stage"] = stage.unsqueeze(1).float()
        to["time_in_stage"] = tis.unsqueeze(1)
        # progress :958-978
        prog = torch.zeros_like(tis)
        m1, m2, m3 = stage == 1, stage == 2, stage == 3
        prog[m1] = torch.clamp(tis[m1] / tc["stage1_duration"], 0, 1)
        prog[m2] = torch.clamp(tis[m2] / tc["stage2_duration"], 0, 1)
        prog[m3] = 1.0
        to["stage_progress"] = prog.unsqueeze(1)
        return to

    def compute_observations(self):
        """observation_encoder.py:442-478 with exclude_components=['active_rule_targets'].
        NB: does NOT re-gather the monitored contact forces; only process_physics_step does
        (step_processor.py:46-48), so the first obs pass of reset() sees the stale gather."""
        od = self._default_observations()
        od.update(self._task_observations(od))
        return od

    def concatenate(self, od):
        """observation_encoder.py:783-829."""
        parts = []
        for k in self.obs_keys:
            t = od[k]
            parts.append(t.reshape(self.n, -1) if t.dim() > 2 else t)
        return torch.cat(parts, dim=1)

    # ================================================================== termination (M1-M6)
    def _evaluate_termination(self):
        """step_processor.py:133-181 + termination_manager.py:120-321."""
        n, tc, od = self.n, self.task_cfg, self.obs_dict
        builtin_f, task_s, task_f = {}, {}, {}
        if "height_safety" in tc["termination"]:
            hs = tc["termination"]["height_safety"]
            hand_z = self.rigid_body_states[:, self.layout.hand, 2]
            tip_z = od["fingertip_poses_world"].view(n, 5, 7)[:, :, 2]
            builtin_f["hitting_ground"] = (hand_z < hs["handbase_threshold"]) | \
                                          (torch.min(tip_z, dim=1)[0] < hs["fingertip_threshold"])
        if self.is_bg:
            # blind_grasping_task.py:1210-1236
            task_s["grasp_lift_success"] = self.ts["success_duration_steps"] >= self.contact_duration_threshold_steps
            # blind_grasping_task.py:1238-1364
            task_f["box_too_far"] = od["hand_to_object_distance"] > tc["max_box_distance"]
            stage = self.ts["current_stage"]
            j2, j3 = self.ts["just_transitioned_to_stage2"], self.ts["just_transitioned_to_stage3"]
            for t in self.failure_tensors.values():
                t.fill_(False)
            grasp = od["grasp_state"].squeeze(-1).bool()
            if j2.any():
                q = tc["quality_thresholds"]
                pads = od["fingerpad_poses_world"].view(n, 5, 7)[:, :, :3]
                h_ok = torch.all(torch.abs(pads[:, :3, 2] - self.box_pos[:, 2].unsqueeze(1)) <= q["height_tolerance"], dim=1)
                c_ok = torch.norm(od["first_three_fingerpad_centroid"] - self.box_pos, dim=1) <= q["centroid_tolerance"]
                drift = torch.norm(self.box_pos - self.initial_box_positions[:, :3], dim=1)
                vmag = torch.norm(self.box_vel, dim=1)
                s_ok = (drift <= q["position_drift_tolerance"]) & (vmag <= q["velocity_tolerance"])
                self.failure_tensors["stage1_pregrasp_failed"][:] = j2 & ~(h_ok & c_ok & s_ok)
            if j3.any():
                self.failure_tensors["stage2_contact_failed"][:] = j3 & ~grasp
            m3 = stage == 3
            if m3.any():
                self.failure_tensors["stage3_grasp_lost"][:] = m3 & ~grasp
            task_f.update(self.failure_tensors)

        info = {}
        act_s, act_f = {}, {}
        for name, crit in task_s.items():
            if name in self.active_success or not self.active_success:
                act_s[name] = crit
                info[f"success_{name}"] = crit.float().mean()
        for src in (builtin_f, task_f):
            for name, crit in src.items():
                if name in self.active_failure or not self.active_failure:
                    act_f[name] = crit
                    info[f"failure_{name}"] = crit.float().mean()
        ep_s = torch.zeros(n, dtype=torch.bool)
        ep_f = torch.zeros(n, dtype=torch.bool)
        for name, crit in act_s.items():
            prev = self.success_reasons.get(name, torch.zeros(n, dtype=torch.bool))
            self.success_reasons[name] = (~ep_s & crit) | prev
            ep_s = ep_s | crit
        for name, crit in act_f.items():
            prev = self.failure_reasons.get(name, torch.zeros(n, dtype=torch.bool))
            self.failure_reasons[name] = (~ep_f & crit) | prev
            ep_f = ep_f | crit
        timeout = self.episode_step_count >= self.max_episode_length - 1
        should_reset = ep_s | ep_f | timeout
        s_t = ep_s & should_reset
        f_t = ep_f & ~ep_s & should_reset
        t_t = timeout & ~ep_s & ~ep_f & should_reset
        info["success"], info["failure"], info["timeout"] = s_t, f_t, t_t
        for name, m in self.success_reasons.items():
            info[f"success_reason_{name}"] = m
        for name, m in self.failure_reasons.items():
            info[f"failure_reason_{name}"] = m
        info["success_rate"] = s_t.sum().float() / n
        info["failure_rate"] = f_t.sum().float() / n
        info["timeout_rate"] = t_t.sum().float() / n
        # termination_manager.py:280-321
        rew, raw = {}, {}
        for key, mask, val in (("success", s_t, self.success_reward),
                               ("failure_penalty", f_t, -self.failure_penalty),
                               ("timeout_penalty", t_t, -self.timeout_penalty)):
            r, w = torch.zeros(n), torch.zeros(n)
            r[mask] = val
            w[mask] = 1.0
            rew[key], raw[key] = r, w
        return should_reset, info, rew, raw

    # ================================================================== rewards (R1-R3)
    def _common_rewards(self):
        """reward_calculator.py:66-221."""
        n, od, lay = self.n, self.obs_dict, self.layout
        r = {}
        hand_vel = self.rigid_body_states[:, lay.hand, 7:10]
        hand_ang = self.rigid_body_states[:, lay.hand, 10:13]
        f_vel = self.dof_vel[:, NB_BASE:]
        f_pos = self.dof_pos[:, NB_BASE:]
        lo, hi = self.dof_props[NB_BASE:, 4], self.dof_props[NB_BASE:, 5]
        r["alive"] = torch.ones(n)
        tip_z = od["fingertip_poses_world"].view(n, 5, 7)[:, :, 2]
        min_z = torch.min(tip_z, dim=1)[0]
        r["height_safety"] = torch.clamp(1.0 - torch.exp(-(min_z - 0.02) * 20), 0.0, 1.0)
        r["finger_velocity"] = torch.exp(-0.1 * torch.norm(f_vel, dim=1))
        r["hand_velocity"] = torch.exp(-0.2 * torch.norm(hand_vel, dim=1))
        r["hand_angular_velocity"] = torch.exp(-0.2 * torch.norm(hand_ang, dim=1))
        norm_j = 2.0 * (f_pos - lo) / (hi - lo) - 1.0
        pen = torch.clamp(torch.abs(norm_j) - 0.8, 0.0, 1.0)
        r["joint_limit"] = 1.0 - torch.sum(pen, dim=1) / f_pos.shape[1]
        if self.prev_finger_dof_vel is None:
            self.prev_finger_dof_vel = f_vel.clone()
            self.prev_hand_vel = hand_vel.clone()
            self.prev_hand_ang_vel = hand_ang.clone()
        r["finger_acceleration"] = torch.exp(-2.0 * torch.norm(f_vel - self.prev_finger_dof_vel, dim=1))
        r["hand_acceleration"] = torch.exp(-0.5 * torch.norm(hand_vel - self.prev_hand_vel, dim=1))
        r["hand_angular_acceleration"] = torch.exp(-0.5 * torch.norm(hand_ang - self.prev_hand_ang_vel, dim=1))
        cf = od["contact_forces"]
        contacts = torch.norm(cf.view(n, cf.shape[1] // 3, 3), dim=2) > 0.1
        if self.prev_contacts is None:
            self.prev_contacts = contacts.clone()
        changes = torch.sum(torch.logical_xor(contacts, self.prev_contacts).float(), dim=1)
        r["contact_stability"] = torch.exp(-changes)
        self.prev_finger_dof_vel.copy_(f_vel)
        self.prev_hand_vel.copy_(hand_vel)
        self.prev_hand_ang_vel.copy_(hand_ang)
        self.prev_contacts.copy_(contacts)
        return r

    def _tip_box_min_distance(self):
        """blind_grasping_task.py:1178-1192."""
        tips = self.obs_dict["fingertip_poses_world"].view(-1, 5, 7)[:, :, :3]
        return torch.norm(tips - self.box_pos.unsqueeze(1), dim=-1).min(dim=1)[0]

    def _task_rewards(self):
        """blind_grasping_task.py:980-1208."""
        if not self.is_bg:
            return {}
        n, od, tc = self.n, self.obs_dict, self.task_cfg
        rc = tc["reward_calculation"]
        pp = tc["penetration_prevention"]
        r = {}
        stage = self.ts["current_stage"]
        m1, m2, m3 = stage == 1, stage == 2, stage == 3
        if m1.any():  # :1055-1112
            f = m1.float()
            r["s1_height_alignment"] = torch.exp(-rc["height_alignment_decay"] * od["avg_finger_to_object_height_diff"]) * f
            cd = torch.norm(od["first_three_fingerpad_centroid"] - self.box_pos, dim=1)
            r["s1_centroid_positioning"] = torch.exp(-rc["centroid_positioning_decay"] * cd) * f
            drift = torch.norm(self.box_pos - self.initial_box_positions[:, :3], dim=1)
            vmag = torch.norm(self.box_vel, dim=1)
            r["s1_object_stability"] = torch.exp(-rc["object_stability_decay"] * (drift + vmag)) * f
            pads = od["fingerpad_poses_world"].view(n, 5, 7)[:, :, :3]
            var = torch.var(pads[:, :3, 2], dim=1)
            r["s1_finger_height_consistency"] = torch.exp(-rc["first_three_height_consistency_decay"] * var) * f
            thumb_rot = od["all_finger_dof_pos"][:, 0]
            r["s1_thumb_rotation"] = torch.exp(-5.0 * torch.abs(thumb_rot - math.pi / 2)) * f
        if m2.any():  # :1114-1150
            f = m2.float()
            r["s2_thumb_contact"] = od["thumb_contact"].squeeze(-1).bool().float() * f
            r["s2_other_fingers_contact"] = od["other_fingers_contact"].squeeze(-1).bool().float() * f
            r["s2_grasp_achievement"] = od["grasp_state"].squeeze(-1).bool().float() * f
            min_d = self._tip_box_min_distance()
            safe = torch.clamp(min_d, min=self.box_size / 2.0 * pp["proximityMinDistanceFactor"])
            r["s2_fingerpad_proximity"] = torch.exp(-rc["fingerpad_proximity_decay"] * safe) * f
            r["s2_base_stability"] = torch.exp(-rc["base_stability_decay"] * torch.norm(od["base_dof_vel"], dim=1)) * f
        if m3.any():  # :1152-1176
            f = m3.float()
            h = self.box_pos[:, 2] - self.box_z
            r["s3_object_height"] = torch.clamp(h / (tc["success_height_threshold"] - self.box_z), 0, 1) * f
            r["s3_grasp_maintenance"] = od["grasp_state"].squeeze(-1).bool().float() * f
            gd = od["grasp_duration"].squeeze(-1)
            r["s3_grasp_duration"] = torch.clamp(gd / tc["contact_duration_threshold"], 0, 1) * f
        j2, j3 = self.ts["just_transitioned_to_stage2"], self.ts["just_transitioned_to_stage3"]
        r["s1_completion"] = (j2 & ~self.failure_tensors["stage1_pregrasp_failed"]).float()
        r["s2_completion"] = (j3 & ~self.failure_tensors["stage2_contact_failed"]).float()
        thr = self.box_size / 2.0 * pp["geometricPenetrationFactor"]
        depth = torch.clamp(thr - self._tip_box_min_distance(), min=0.0)
        r["penetration_penalty"] = depth * pp["penetrationDepthScale"]
        return r

    def _total_reward(self, common, task):
        """reward_calculator.py:223-275."""
        total = torch.zeros(self.n)
        comps = {}
        for name, val in {**common, **task}.items():
            w = self.reward_weights.get(name, 0.0)
            if w != 0:
                weighted = val * w
                total += weighted
                comps[name] = val
                comps[f"{name}_weighted"] = weighted
        comps["total"] = total
        return total, comps

    # ================================================================== reset (Z1, P8)
    def _reset_task_state(self, ids):
        """blind_grasping_task.py:433-547 with the torch.rand draws replaced by injected uniforms."""
        if len(ids) == 0 or not self.is_bg:
            return
        u = self._uniforms[ids]
        tc, a = self.task_cfg, self.layout.box_actor
        x = (u[:, 0] * 2 - 1) * self.box_xy_range
        y = (u[:, 1] * 2 - 1) * self.box_xy_range
        ang = (u[:, 2] * 2 - 1) * math.pi
        self.initial_box_positions[ids, 0] = x
        self.initial_box_positions[ids, 1] = y
        self.initial_box_positions[ids, 2] = self.box_z
        root = self.actor_root_state
        root[ids, a, 0], root[ids, a, 1], root[ids, a, 2] = x, y, self.box_z
        root[ids, a, 3] = 0
        root[ids, a, 4] = 0
        root[ids, a, 5] = torch.sin(ang / 2)
        root[ids, a, 6] = torch.cos(ang / 2)
        root[ids, a, 7:13] = 0
        self.ts["success_duration_steps"][ids] = 0
        self.ts["success_conditions_met"][ids] = False
        self.ts["current_stage"][ids] = 1
        self.ts["time_in_stage"][ids] = 0.0
        self.ts["stage_contact_duration"][ids] = 0.0
        self.ts["just_transitioned_to_stage2"][ids] = False
        self.ts["just_transitioned_to_stage3"][ids] = False
        self.dof_state[ids, 0:3, 0] = (u[:, 3:6] * 2 - 1) * tc["hand_translation_range"]
        self.dof_state[ids, 3:6, 0] = (u[:, 6:9] * 2 - 1) * tc["hand_rotation_range"]
        ranges = torch.full((N_FINGER_DOF,), tc["finger_randomization"]["other_finger_range"])
        ranges[0] = tc["finger_randomization"]["thumb_rotation_range"]
        self.dof_state[ids, 6:26, 0] = u[:, 9:29].clone() * ranges.unsqueeze(0)

    def reset_idx(self, ids):
        """dexhand_base.py:743-756 -> termination_manager.py:341-357, reset_manager.py:92-190,
        action_processor.py:524-568, observation_encoder.py:363-383."""
        if len(ids) == 0:
            return
        for m in list(self.success_reasons.values()) + list(self.failure_reasons.values()):
            m[ids] = False
        self.episode_step_count[ids] = 0
        self.dof_state[ids, :, 0] = torch.zeros(26)
        self.dof_state[ids, :, 1] = 0
        self._reset_task_state(ids)
        pos = self.dof_state[ids, :, 0]
        self.active_prev_targets[ids, :] = self.extract_active_targets(pos)
        self.full_dof_targets[ids, :] = pos
        self.contact_duration_steps[ids] = 0
        self.prev_contact_binary[ids] = 0
        self.prev_dof_pos[ids] = 0
        self.prev_actions[ids] = 0

    # ================================================================== step orchestration
    def reset(self, uniforms=None):
        """dexhand_base.py:805-838."""
        if uniforms is not None:
            self._uniforms = uniforms
        self.reset_idx(torch.arange(self.n))
        od = self.compute_observations()
        od["active_rule_targets"] = self.active_prev_targets.clone()
        self.obs_buf = self.concatenate(od)
        self.obs_dict = od
        return self.post_physics_step()[0]

    def post_physics_step(self, uniforms=None):
        """step_processor.py:37-131."""
        if uniforms is not None:
            self._uniforms = uniforms
        # O1 tensor_manager.py:441-446: gather of the monitored contact bodies
        self.contact_forces = self.contact_forces_all[:, torch.tensor(self.layout.contacts), :]
        od = self.compute_observations()
        hooks = getattr(self, "hooks", None) or {}
        if hooks.get("pre_action_rule") is not None:                   # step_processor.py:66-73
            od["active_rule_targets"] = hooks["pre_action_rule"](self.active_prev_targets, {"obs_dict": od, "env": self})
        else:
            od["active_rule_targets"] = self.active_prev_targets.clone()   # rules.py:78-95 identity rule
        self.obs_buf = self.concatenate(od)
        self.states_buf = self.obs_buf.clone()
        self.obs_dict = od
        self.episode_step_count += 1
        should_reset, info, term_rew, raw_rew = self._evaluate_termination()
        self.reset_buf = should_reset
        task_rewards = dict(self._task_rewards())
        if hooks.get("extra_reward_terms") is not None:                # a custom task's compute_task_reward_terms
            task_rewards.update(hooks["extra_reward_terms"](od))
        self.rew_buf[:], comps = self._total_reward(self._common_rewards(), task_rewards)
        # termination_manager.py:323-339
        if torch.any(info["success"]):
            self.consecutive_successes += 1
        else:
            self.consecutive_successes = 0
        self.consecutive_successes = min(self.consecutive_successes, self.max_consecutive_successes)
        for key, val in term_rew.items():   # step_processor.py:204-219
            self.rew_buf += val
            comps[f"termination_{key}"] = raw_rew[key]
            comps[f"termination_{key}_weighted"] = val
        self.env_ids = torch.nonzero(self.reset_buf).flatten()
        if torch.any(self.reset_buf):
            self.reset_idx(self.env_ids)
        self.extras = {"consecutive_successes": self.consecutive_successes,
                       "episode_length": self.episode_step_count.clone()}
        self.extras.update(info)
        self.extras["reward_components"] = comps
        return self.obs_buf, self.rew_buf, self.reset_buf, self.extras

    # ================================================================== snapshot (same keys as RefEnv)
    def snapshot(self):
        out = {
            "obs_buf": self.obs_buf.clone(), "rew_buf": self.rew_buf.clone(),
            "reset_buf": self.reset_buf.clone(),
            "episode_step_count": self.episode_step_count.clone(),
            "active_prev_targets": self.active_prev_targets.clone(),
            "full_dof_targets": self.full_dof_targets.clone(),
            "prev_actions": self.prev_actions.clone(),
            "prev_dof_pos": self.prev_dof_pos.clone(),
            "contact_duration_steps": self.contact_duration_steps.clone(),
            "prev_contact_binary": self.prev_contact_binary.clone(),
            "dof_state": self.dof_state.clone(),
            "actor_root_state": self.actor_root_state.clone(),
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
        for k, v in self.ts.items():
            out[f"task/{k}"] = v.clone()
        if self.is_bg:
            out["task/initial_box_positions"] = self.initial_box_positions.clone()
        for k, v in self.obs_dict.items():
            out[f"obs/{k}"] = v.clone()
        return out
