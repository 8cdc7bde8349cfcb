// post_physics.cuh — the table-driven TILE kernel of the post-physics pass (StepProcessor.process_physics_step,
// step_processor.py:37-131).  Round 1's production kernel; since round 2 the lean path runs the kernels of
// post_physics_v2.cuh and this one serves the instances those have no stage for - obs_dict materialisation
// (DexOutBuffers.obs_dict != NULL), DEX_MODE_OBS_ONLY, DEX_FLAG_STALE_CONTACTS, non-standard layouts - and A/B runs
// (DEX_FLAG_TILE_KERNEL / DEXSTEP_POST_V1=1).  It is the kernel pinned first against the oracle; the v2 kernels are
// compared with it bit for bit (tests/test_cuda_v2.py).
//
// Design: "gather table -> compute on a shared-memory slab -> scatter table".
//   One CTA of 256 threads (8 warps) owns a tile of 32 consecutive envs.  Every env of the tile has a SLAB of
//   floats in shared memory (odd stride, so `slab[lane * stride + const]` never bank-conflicts when lane = env).
//   phase 0  stage-in    plain global loads staged through registers, ALL issued before the first shared store (one
//                        DRAM round trip per tile):
//                        * the Isaac Gym AoS rows (dof_state row, 11 rigid-body rows, 5(+1) contact rows, box root
//                          row, full_dof_targets row) row-per-warp: consecutive lanes = consecutive floats of a row
//                          (with DEX_FLAG_EARLY_SIM_LOADS they are issued before griddepcontrol.wait);
//                        * the path's SoA state rows through a host-built IN TABLE (row pointer, slab slot, type):
//                          lane = env, one 128 B line per request.
//                        Buffers the fused reset writes (dof_state, actor_root_state, full_dof_targets, state rows)
//                        are read with coherent loads (ld.global.cg), only rigid bodies / contact forces with .nc.
//   phase 1  items       the per-env arithmetic is cut into independent work items (one fingertip/pad pose, one
//                        contact sensor, one pad pair, one DOF velocity, hand, box).  Warps take items round-robin,
//                        lane = env: item-type branches are warp-uniform, all shared accesses conflict-free.
//                        Items read slab slots and write slab slots - no global memory, no layout logic.
//   phase 2  scalar      warp 0, lane = env: contact logic, stage FSM, termination, reward terms, masks -> slab;
//                        cross-env counters via warp ballots, published as the tile's three 8-byte records.
//   phase 3  stage-out   * obs_buf rows (and the optional obs_dict rows) are GATHERED from the slab through a
//                          host-built column map (obs column -> slab slot), row-per-warp, coalesced;
//                        * every updated state / output row leaves through the OUT TABLE (f32 / u8 / i64 rows),
//                          lane = env, 128 B per request.
//   phase 4  reset       tiles that contain finished envs (CTA-uniform vote) apply the fused reset.
//   epilogue             optional (ep != NULL): ONE extra CTA behind the tiles polls the tile records (bounded wait,
//                        include/dexstep.h DexEpilogueBuffers.tile_records), compacts the reset words into the ordered
//                        env-id list and finalises the statistics - no extra launch, no atomics on the step path.
#pragma once
#include "dexstep_common.cuh"

namespace {

constexpr int kTile = 32;          // envs per CTA == warp width (lane = env)
constexpr int kPostThreads = 256;  // 8 warps
constexpr int kPostWarps = kPostThreads / 32;
constexpr int kRbW = 13 + 10 * 7;  // hand row (13) + 5 tips + 5 pads (7 each) = 83
constexpr int kMaxIn = 160, kMaxOut = 192, kMaxObs = 256, kMaxDict = 512;
constexpr int kRowU = 4;                          // table rows per warp per batch (loads issued before stores)
constexpr int kRowBatch = kPostWarps * kRowU;     // f32 table segments are padded to a multiple of this

// ---- slab slots (floats).  Fixed part; the optional sections follow at run-time offsets (PostArgs.o_*). ----
enum : int {
  I_DOF = 0,                 // 52  dof_state row (pos, vel interleaved)
  I_RB = I_DOF + 52,         // 83  hand row (13) | 5 tip poses | 5 pad poses
  I_CF = I_RB + kRbW,        // 18  5 monitored contact forces | box contact force
  I_BOX = I_CF + 18,         // 6   box pos | box linvel
  S_PDP = I_BOX + 6,         // 26  prev_dof_pos in  -> finite-difference velocity out (in place)
  S_APT = S_PDP + 26,        // 18  active_prev_targets
  S_PACT = S_APT + 18,       // 18  prev_actions
  S_CDUR = S_PACT + 18,      // 5   contact_duration_steps (updated in place)
  S_PBIN = S_CDUR + 5,       // 5   prev_contact_binary in -> contact_binary out (in place)
  S_BOX0 = S_PBIN + 5,       // 3   initial_box_positions
  S_STEP = S_BOX0 + 3,       // int episode_step_count (updated in place)
  S_SDS = S_STEP + 1,        // int success_duration_steps
  S_STAGE = S_SDS + 1,       // int current_stage
  S_TIS = S_STAGE + 1,       // time_in_stage
  S_SCD = S_TIS + 1,         // stage_contact_duration
  C_CMAG = S_SCD + 1,        // 5   contact force magnitude
  C_CDURS = C_CMAG + 5,      // 5   contact duration in seconds
  C_PADBOX = C_CDURS + 5,    // 5   pad -> box distance
  C_HDIFF = C_PADBOX + 5,    // 5   |pad_z - box_z|
  C_TIPBOX = C_HDIFF + 5,    // 5   tip -> box distance
  C_PAIR = C_TIPBOX + 5,     // 10  pad pair distances
  C_CEN = C_PAIR + 10,       // 3   centroid of the first three pads
  C_HANDBOX = C_CEN + 3,
  C_CDIST,
  C_DRIFT,
  C_VMAG,
  C_BOXHAS,
  C_AVGDIST,
  C_AVGHDIFF,
  C_THUMB,
  C_OTHERS,
  C_GRASP,
  C_GRASPDUR,
  C_STAGEF,
  C_PROG,
  C_EPTIME,
  C_MINTIPZ,
  C_MINTIPBOX,
  M_RESET,  // masks as float 0/1 (they leave as u8 rows)
  M_SUCC,
  M_FAIL,
  M_TIMEOUT,
  M_SCM,
  M_J2,
  M_J3,
  M_FT0,
  M_FT1,
  M_FT2,
  C_REW,
  C_COMP  // num_component_rows reward components; the optional sections follow
};
// offsets inside the optional sections
enum : int { HF_TIP = 0, HF_PAD = 35, HF_ARR = 70, HF_SIZE = 74 };  // hand-frame poses, ARR-aligned quaternion
enum : int {                                                         // common-reward state and partial terms
  CP_PFDV = 0, CP_PHV = 20, CP_PHA = 23, CP_PCON = 26, CP_CNOW = 31, CP_CHG = 36, CP_RHAND = 41,
  CP_FV2 = 45, CP_FA2 = 65, CP_JL = 85, CP_SIZE = 105
};
enum : int { ROW_F32 = 0, ROW_U8 = 1, ROW_I64 = 2 };

struct PostArgs {
  DexParams p;
  DexSimViews sim;
  DexStateBuffers st;
  DexOutBuffers out;
  DexEpilogueBuffers ep;  // used when fused_epilogue != 0
  int32_t n;
  int32_t mode;
  uint32_t flags;
  int32_t fused_epilogue;
  const float* uniforms;
  uint64_t seed;
  uint64_t step;
  // ---- derived on the host at launch (build_post_tables) ----
  int32_t stride;                       // slab stride (odd)
  int32_t o_tgt, o_hf, o_cpv;           // optional sections (-1 = absent)
  int32_t want_hf_tip, want_hf_pad, want_arr, common, dict_dim;
  int32_t n_in_f32, n_in_i64, n_in;     // IN table, sorted by type: [0,n_in_f32) f32 | [..,n_in_i64) i64 | [..,n_in) u8
  // OUT table: [0,n_out_early) f32 rows that are final after the items phase | [..,n_out_f32) f32 rows that the scalar
  // phase / the reward terms write | [..,n_out_i64) int64 rows | [..,n_out) bool rows
  int32_t n_out_early, n_out_f32, n_out_i64, n_out;
  uint32_t obs_late, dict_late;         // bit c: columns [32c, 32c+32) read a slot that the scalar phase writes
  int32_t rb_off[11];                   // float offset of the hand / 5 tip / 5 pad rows inside one env's [NB,13] slab
  int32_t rb_pitch;                     // floats of one env in rigid_body_state (NB * DexParams.rigid_row_floats)
  int32_t cf_off[6];                    // float offset of the 5 monitored contact rows (+ box) inside [NB,3]
  uint64_t in_ptr[kMaxIn];              // row base pointers
  uint16_t in_meta[kMaxIn];             // slab slot
  uint64_t out_ptr[kMaxOut];
  uint16_t out_meta[kMaxOut];
  int16_t obs_src[kMaxObs];             // obs column -> slab slot
  int16_t dict_src[kMaxDict];           // obs_dict column -> slab slot
};

// ------------------------------------------------------------------------------------------------
// host: slab layout, IN / OUT tables, observation column maps
// ------------------------------------------------------------------------------------------------
inline int obs_key_dim(const DexParams& P, const int key) {
  switch (key) {
    case DEX_K_BASE_DOF_POS: case DEX_K_BASE_DOF_VEL: case DEX_K_BASE_DOF_TARGET: return 6;
    case DEX_K_ACTIVE_FINGER_DOF_POS: case DEX_K_ACTIVE_FINGER_DOF_VEL: case DEX_K_ACTIVE_FINGER_DOF_TARGET: return 12;
    case DEX_K_ALL_FINGER_DOF_POS: case DEX_K_ALL_FINGER_DOF_VEL: case DEX_K_ALL_FINGER_DOF_TARGET: return 20;
    case DEX_K_HAND_POSE: case DEX_K_HAND_POSE_ARR_ALIGNED: return 7;
    case DEX_K_CONTACT_FORCES: return 15;
    case DEX_K_PREV_ACTIONS: return P.num_actions;
    case DEX_K_ACTIVE_PREV_TARGETS: case DEX_K_ACTIVE_RULE_TARGETS: return 18;
    case DEX_K_CONTACT_FORCE_MAGNITUDE: case DEX_K_CONTACT_BINARY: case DEX_K_CONTACT_DURATION: return 5;
    case DEX_K_FINGERTIP_POSES_WORLD: case DEX_K_FINGERTIP_POSES_HAND: case DEX_K_FINGERPAD_POSES_WORLD:
    case DEX_K_FINGERPAD_POSES_HAND: return 35;
    case DEX_K_OBJECT_POS: case DEX_K_OBJECT_VEL: case DEX_K_FIRST_THREE_FINGERPAD_CENTROID: return 3;
    case DEX_K_FINGER_TO_OBJECT_DISTANCES: case DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF: return 5;
    case DEX_K_FINGERPAD_DISTANCES: return 10;
    default: return 1;
  }
}

// slab slot that holds element j of observation key `key`
inline int obs_key_slot(const PostArgs& a, const int key, const int j) {
  const DexParams& P = a.p;
  switch (key) {
    case DEX_K_BASE_DOF_POS: return I_DOF + 2 * j;
    case DEX_K_BASE_DOF_VEL: return S_PDP + j;
    case DEX_K_ACTIVE_FINGER_DOF_POS: return I_DOF + 2 * P.active_finger_dof[j];
    case DEX_K_ACTIVE_FINGER_DOF_VEL: return S_PDP + P.active_finger_dof[j];
    case DEX_K_ALL_FINGER_DOF_POS: return I_DOF + 2 * (DEX_NUM_BASE + j);
    case DEX_K_ALL_FINGER_DOF_VEL: return S_PDP + DEX_NUM_BASE + j;
    case DEX_K_HAND_POSE: return I_RB + j;
    case DEX_K_HAND_POSE_ARR_ALIGNED: return j < 3 ? I_RB + j : a.o_hf + HF_ARR + (j - 3);
    case DEX_K_CONTACT_FORCES: return I_CF + j;
    case DEX_K_PREV_ACTIONS: return S_PACT + j;
    case DEX_K_ACTIVE_PREV_TARGETS: case DEX_K_ACTIVE_RULE_TARGETS: return S_APT + j;
    case DEX_K_BASE_DOF_TARGET: return a.o_tgt + j;
    case DEX_K_ACTIVE_FINGER_DOF_TARGET: return a.o_tgt + P.active_finger_dof[j];
    case DEX_K_ALL_FINGER_DOF_TARGET: return a.o_tgt + DEX_NUM_BASE + j;
    case DEX_K_CONTACT_FORCE_MAGNITUDE: return C_CMAG + j;
    case DEX_K_CONTACT_BINARY: return S_PBIN + j;
    case DEX_K_CONTACT_DURATION: return C_CDURS + j;
    case DEX_K_FINGERTIP_POSES_WORLD: return I_RB + 13 + j;
    case DEX_K_FINGERTIP_POSES_HAND: return a.o_hf + HF_TIP + j;
    case DEX_K_FINGERPAD_POSES_WORLD: return I_RB + 13 + 35 + j;
    case DEX_K_FINGERPAD_POSES_HAND: return a.o_hf + HF_PAD + j;
    case DEX_K_EPISODE_TIME: return C_EPTIME;
    case DEX_K_OBJECT_POS: return I_BOX + j;
    case DEX_K_OBJECT_VEL: return I_BOX + 3 + j;
    case DEX_K_FINGER_TO_OBJECT_DISTANCES: return C_PADBOX + j;
    case DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE: return C_AVGDIST;
    case DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF: return C_HDIFF + j;
    case DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF: return C_AVGHDIFF;
    case DEX_K_HAND_TO_OBJECT_DISTANCE: return C_HANDBOX;
    case DEX_K_FINGERPAD_DISTANCES: return C_PAIR + j;
    case DEX_K_FIRST_THREE_FINGERPAD_CENTROID: return C_CEN + j;
    case DEX_K_THUMB_CONTACT: return C_THUMB;
    case DEX_K_OTHER_FINGERS_CONTACT: return C_OTHERS;
    case DEX_K_GRASP_STATE: return C_GRASP;
    case DEX_K_GRASP_DURATION: return C_GRASPDUR;
    case DEX_K_CURRENT_STAGE: return C_STAGEF;
    case DEX_K_TIME_IN_STAGE: return S_TIS;
    case DEX_K_STAGE_PROGRESS: return C_PROG;
    default: return -1;
  }
}

// Fills the derived part of PostArgs.  Returns DEX_OK or DEX_ERR_SIZE when a table overflows.
inline int build_post_tables(PostArgs& a) {
  const DexParams& P = a.p;
  const DexStateBuffers& st = a.st;
  const bool bg = P.task == DEX_TASK_BLIND_GRASPING, dict = a.out.obs_dict != nullptr;
  const bool full = a.mode == DEX_MODE_FULL;
  const bool stale = (a.flags & DEX_FLAG_STALE_CONTACTS) != 0;
  const size_t N = (size_t)a.n;
  auto wanted = [&](int key) { return dict || P.obs_offset[key] >= 0; };
  a.common = (full && P.need_common_rewards) ? 1 : 0;
  a.want_hf_tip = wanted(DEX_K_FINGERTIP_POSES_HAND);
  a.want_hf_pad = wanted(DEX_K_FINGERPAD_POSES_HAND);
  a.want_arr = wanted(DEX_K_HAND_POSE_ARR_ALIGNED);
  const bool want_tgt = wanted(DEX_K_BASE_DOF_TARGET) || wanted(DEX_K_ACTIVE_FINGER_DOF_TARGET) ||
                        wanted(DEX_K_ALL_FINGER_DOF_TARGET);
  int o = C_COMP + P.num_component_rows;
  a.o_tgt = want_tgt ? o : -1;
  o += want_tgt ? DEX_NUM_DOF : 0;
  const bool want_hf = a.want_hf_tip || a.want_hf_pad || a.want_arr;
  a.o_hf = want_hf ? o : -1;
  o += want_hf ? HF_SIZE : 0;
  a.o_cpv = P.need_common_rewards ? o : -1;
  o += P.need_common_rewards ? CP_SIZE : 0;
  a.stride = o | 1;
  if (a.stride >= 4096) return DEX_ERR_SIZE;

  const int rw = P.rigid_row_floats > 0 ? P.rigid_row_floats : 13;  // 13 = Isaac Gym rows, 7..12 = packed host staging
  a.rb_pitch = P.num_bodies * rw;
  a.rb_off[0] = P.hand_body * rw;
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
    a.rb_off[1 + f] = P.tip_body[f] * rw;
    a.rb_off[6 + f] = P.pad_body[f] * rw;
    a.cf_off[f] = P.contact_body[f] * 3;
  }
  a.cf_off[5] = bg ? P.box_body * 3 : 0;

  // ---- IN table (filled per type, then concatenated f32 | i64 | u8) ----
  struct Row { uint64_t ptr; uint16_t slot; };
  Row rows[3][kMaxOut];
  int cnt[3] = {0, 0, 0};
  bool ok = true;
  auto add = [&](const void* base, size_t row, int elem, int slot, int type) {
    if (cnt[type] >= kMaxOut) { ok = false; return; }
    rows[type][cnt[type]++] = Row{(uint64_t)(uintptr_t)((const char*)base + row * N * elem), (uint16_t)slot};
  };
  auto in = add;
  for (int d = 0; d < DEX_NUM_DOF; ++d) in(st.prev_dof_pos, d, 4, S_PDP + d, ROW_F32);
  for (int k = 0; k < DEX_NUM_ACTIVE; ++k) in(st.active_prev_targets, k, 4, S_APT + k, ROW_F32);
  for (int k = 0; k < P.num_actions; ++k) in(st.prev_actions, k, 4, S_PACT + k, ROW_F32);
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) in(st.contact_duration_steps, f, 4, S_CDUR + f, ROW_F32);
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) in(st.prev_contact_binary, f, 4, S_PBIN + f, ROW_F32);
  in(st.episode_step_count, 0, 8, S_STEP, ROW_I64);
  if (bg) {
    for (int c = 0; c < 3; ++c) in(st.initial_box_positions, c, 4, S_BOX0 + c, ROW_F32);
    in(st.success_duration_steps, 0, 8, S_SDS, ROW_I64);
    in(st.current_stage, 0, 8, S_STAGE, ROW_I64);
    in(st.time_in_stage, 0, 4, S_TIS, ROW_F32);
    in(st.stage_contact_duration, 0, 4, S_SCD, ROW_F32);
  }
  if (a.common && !(a.flags & DEX_FLAG_FIRST_REWARD_CALL)) {
    for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) in(st.prev_finger_dof_vel, j, 4, a.o_cpv + CP_PFDV + j, ROW_F32);
    for (int c = 0; c < 3; ++c) in(st.prev_hand_vel, c, 4, a.o_cpv + CP_PHV + c, ROW_F32);
    for (int c = 0; c < 3; ++c) in(st.prev_hand_ang_vel, c, 4, a.o_cpv + CP_PHA + c, ROW_F32);
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) in(st.prev_contacts, f, 1, a.o_cpv + CP_PCON + f, ROW_U8);
  }
  if (stale)
    for (int i = 0; i < 15; ++i) in(st.contact_forces, i, 4, I_CF + i, ROW_F32);
  {
    int k = 0;
    const int order[3] = {ROW_F32, ROW_I64, ROW_U8};
    if (cnt[0] + cnt[1] + cnt[2] + kRowBatch > kMaxIn) return DEX_ERR_SIZE;
    for (int t = 0; t < 3; ++t) {
      for (int i = 0; i < cnt[order[t]]; ++i, ++k) {
        a.in_ptr[k] = rows[order[t]][i].ptr;
        a.in_meta[k] = rows[order[t]][i].slot;
      }
      if (t == 0) {  // pad the f32 segment to a multiple of kRowBatch rows (repeat the last row: idempotent)
        while (k > 0 && k % kRowBatch != 0 && k < kMaxIn) {
          a.in_ptr[k] = a.in_ptr[k - 1];
          a.in_meta[k] = a.in_meta[k - 1];
          ++k;
        }
        a.n_in_f32 = k;
      }
      if (t == 1) a.n_in_i64 = k;
    }
    a.n_in = k;
    cnt[0] = cnt[1] = cnt[2] = 0;
  }

  // ---- OUT table ----
  auto out = add;
  for (int d = 0; d < DEX_NUM_DOF; ++d) out(st.prev_dof_pos, d, 4, I_DOF + 2 * d, ROW_F32);  // prev_dof_pos <- pos
  for (int k = 0; k < DEX_NUM_ACTIVE; ++k) out(st.active_rule_targets, k, 4, S_APT + k, ROW_F32);  // rules.py:78-95
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) out(st.contact_duration_steps, f, 4, S_CDUR + f, ROW_F32);
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) out(st.prev_contact_binary, f, 4, S_PBIN + f, ROW_F32);
  if (st.contact_forces != nullptr && !stale)
    for (int i = 0; i < 15; ++i) out(st.contact_forces, i, 4, I_CF + i, ROW_F32);
  if (bg) {
    out(st.success_duration_steps, 0, 8, S_SDS, ROW_I64);
    out(st.success_conditions_met, 0, 1, M_SCM, ROW_U8);
    out(st.current_stage, 0, 8, S_STAGE, ROW_I64);
    out(st.time_in_stage, 0, 4, S_TIS, ROW_F32);
    out(st.stage_contact_duration, 0, 4, S_SCD, ROW_F32);
    out(st.just_transitioned_to_stage2, 0, 1, M_J2, ROW_U8);
    out(st.just_transitioned_to_stage3, 0, 1, M_J3, ROW_U8);
  }
  if (full) {
    out(st.episode_step_count, 0, 8, S_STEP, ROW_I64);
    if (bg)
      for (int c = 0; c < 3; ++c) out(st.failure_tensors, c, 1, M_FT0 + c, ROW_U8);
    for (int i = 0; i < DEX_NUM_REWARD_TERMS; ++i)  // raw + weighted rows of the non-zero-weight terms; the total,
      if (P.component_row[i] >= 0) {                // termination rows and rew_buf leave from post_total()
        out(a.out.reward_components, P.component_row[i], 4, C_COMP + P.component_row[i], ROW_F32);
        out(a.out.reward_components, P.component_row[i] + 1, 4, C_COMP + P.component_row[i] + 1, ROW_F32);
      }
    out(a.out.reset_buf, 0, 1, M_RESET, ROW_U8);
    out(a.out.success, 0, 1, M_SUCC, ROW_U8);
    out(a.out.failure, 0, 1, M_FAIL, ROW_U8);
    out(a.out.timeout, 0, 1, M_TIMEOUT, ROW_U8);
    if (a.common) {
      for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j)
        out(st.prev_finger_dof_vel, j, 4, I_DOF + 2 * (DEX_NUM_BASE + j) + 1, ROW_F32);  // Isaac dof_vel
      for (int c = 0; c < 3; ++c) out(st.prev_hand_vel, c, 4, I_RB + 7 + c, ROW_F32);
      for (int c = 0; c < 3; ++c) out(st.prev_hand_ang_vel, c, 4, I_RB + 10 + c, ROW_F32);
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) out(st.prev_contacts, f, 1, a.o_cpv + CP_CNOW + f, ROW_U8);
    }
  }
  {
    int k = 0;
    const int order[3] = {ROW_F32, ROW_I64, ROW_U8};
    if (cnt[0] + cnt[1] + cnt[2] + kRowBatch > kMaxOut) return DEX_ERR_SIZE;
    auto late = [&](int slot) {
      return slot == S_TIS || slot == S_SCD || (slot >= C_COMP && slot < C_COMP + P.num_component_rows);
    };
    for (int pass = 0; pass < 2; ++pass) {  // f32 rows: early ones first
      for (int i = 0; i < cnt[ROW_F32]; ++i)
        if (late(rows[ROW_F32][i].slot) == (pass == 1)) {
          a.out_ptr[k] = rows[ROW_F32][i].ptr;
          a.out_meta[k++] = rows[ROW_F32][i].slot;
        }
      if (pass == 0) a.n_out_early = k;
    }
    a.n_out_f32 = k;
    for (int t = 1; t < 3; ++t) {
      for (int i = 0; i < cnt[order[t]]; ++i, ++k) {
        a.out_ptr[k] = rows[order[t]][i].ptr;
        a.out_meta[k] = rows[order[t]][i].slot;
      }
      if (t == 1) a.n_out_i64 = k;
    }
    a.n_out = k;
  }

  // ---- observation column maps ----
  if (P.obs_dim > kMaxObs) return DEX_ERR_SIZE;
  for (int c = 0; c < kMaxObs; ++c) a.obs_src[c] = 0;
  const int nkeys = bg ? DEX_NUM_OBS_KEYS : DEX_K_OBJECT_POS;
  for (int key = 0; key < DEX_NUM_OBS_KEYS; ++key) {
    const int off = P.obs_offset[key];
    if (off < 0) continue;
    const int dim = obs_key_dim(P, key);
    if (key >= nkeys || off + dim > P.obs_dim) return DEX_ERR_PARAM;
    for (int j = 0; j < dim; ++j) a.obs_src[off + j] = (int16_t)obs_key_slot(a, key, j);
  }
  auto slot_is_late = [](int s) {  // slots that the scalar phase (warp 0) writes
    return s == C_AVGDIST || s == C_AVGHDIFF || s == C_THUMB || s == C_OTHERS || s == C_GRASP || s == C_GRASPDUR ||
           s == C_STAGEF || s == S_TIS || s == C_PROG || s == C_EPTIME || s == S_STEP || s == S_SDS || s == S_STAGE ||
           s == S_SCD;
  };
  a.obs_late = a.dict_late = 0;
  for (int c = 0; c < P.obs_dim; ++c)
    if (slot_is_late(a.obs_src[c])) a.obs_late |= 1u << (c >> 5);
  a.dict_dim = 0;
  if (dict) {
    if (P.dict_dim > kMaxDict) return DEX_ERR_SIZE;
    a.dict_dim = P.dict_dim;
    for (int c = 0; c < kMaxDict; ++c) a.dict_src[c] = 0;
    for (int key = 0; key < nkeys; ++key) {
      const int off = P.dict_offset[key], dim = obs_key_dim(P, key);
      if (off < 0 || off + dim > P.dict_dim) return DEX_ERR_PARAM;
      for (int j = 0; j < dim; ++j) a.dict_src[off + j] = (int16_t)obs_key_slot(a, key, j);
    }
    for (int c = 0; c < P.dict_dim; ++c)
      if (slot_is_late(a.dict_src[c])) a.dict_late |= 1u << (c >> 5);
  }
  return ok ? DEX_OK : DEX_ERR_SIZE;
}

#ifdef DEXSTEP_TIMELINE  // debug build only (tools/timeline.py): per-CTA phase timestamps, [grid][32]:
// [k] = globaltimer ns (256 ns ticks on B200) and [16 + k] = SM clock at mark k (k < 15), [15] = smid
__device__ unsigned long long* g_timeline = nullptr;
#ifdef DEXSTEP_TIMELINE_LIGHT
// light marks: one CS2R + one shared store per mark (the heavy marks below put an L2 round trip - the load of
// g_timeline - and two global stores on the critical path of warp 0, ~0.3 us each); dumped once at the end:
// [16 + k] = 32-bit SM clock at mark k, [0] / [6] = globaltimer at the start / end of the CTA, [15] = smid
__device__ __forceinline__ unsigned* tl_buf() {
  __shared__ unsigned s_tl[20];  // [0..14] clocks, [16..17] globaltimer at mark 0
  return s_tl;
}
__device__ __forceinline__ void tl_mark(const int k) {
  if (threadIdx.x == 0) {
    unsigned* b = tl_buf();
    b[k] = (unsigned)clock();
    if (k == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      b[16] = (unsigned)t;
      b[17] = (unsigned)(t >> 32);
    }
    if (k == 6) {
      unsigned long long* g = g_timeline;
      if (g != nullptr) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        unsigned sm;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
        g += (size_t)blockIdx.x * 32;
        g[0] = ((unsigned long long)b[17] << 32) | b[16];
        g[6] = t;
        g[15] = sm;
#pragma unroll
        for (int i = 0; i < 15; ++i) g[16 + i] = b[i];
      }
    }
  }
}
__device__ __forceinline__ void tl_mark_w1(const int k) {
  if (threadIdx.x == 32) tl_buf()[k] = (unsigned)clock();
}
#else
__device__ __forceinline__ void tl_mark(const int k) {
  if (threadIdx.x == 0 && g_timeline != nullptr) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    g_timeline[(size_t)blockIdx.x * 32 + k] = t;
    g_timeline[(size_t)blockIdx.x * 32 + 16 + k] = (unsigned long long)clock64();
    if (k == 0) {
      unsigned sm;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
      g_timeline[(size_t)blockIdx.x * 32 + 15] = sm;
    }
  }
}
__device__ __forceinline__ void tl_mark_w1(const int k) {  // SM clock only, taken by warp 1 (first stage-out warp)
  if (threadIdx.x == 32 && g_timeline != nullptr) g_timeline[(size_t)blockIdx.x * 32 + 16 + k] = (unsigned long long)clock64();
}
#endif
#define TL_MARK(k) tl_mark(k)
#define TL_MARK_W1(k) tl_mark_w1(k)
#else
#define TL_MARK(k)
#define TL_MARK_W1(k)
#endif
// ------------------------------------------------------------------------------------------------
// phase 0: stage-in.  Plain LDG -> STS batches (all loads of a batch issued before the first store), no
// cp.async: on sm_100a every 4-byte LDGSTS costs four issue slots, and this kernel is issue-bound.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float lds_f32(const unsigned a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f32(const unsigned a, const float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
}

// FULL = the tile holds 32 envs (every tile but a ragged last one): no per-row predicates
template <bool BG, bool FULL>
__device__ __forceinline__ void post_stage_in(const PostArgs& A, float* smem, const int e0, const int ne) {
  const DexParams& P = A.p;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int S = A.stride;
  const unsigned sbase = (unsigned)__cvta_generic_to_shared(smem);

  // ---- AoS sim rows: warps take env rows round-robin (4 each), consecutive lanes = consecutive floats ----
  {
    auto rbo = [&](int s) {  // rigid slot (0..82) -> float offset inside the env's [NB,13] slab
      s = min(s, kRbW - 1);
      const int row = s < 13 ? 0 : 1 + (s - 13) / 7;
      return A.rb_off[row] + (s < 13 ? s : (s - 13) - (row - 1) * 7);
    };
    const int rbo0 = rbo(lane), rbo1 = rbo(lane + 32), rbo2 = rbo(lane + 64);
    const bool stale = (A.flags & DEX_FLAG_STALE_CONTACTS) != 0;
    const bool p_rb0 = lane < 7 || lane >= 13 || A.common;  // hand velocities only feed the common reward terms
    const bool p_rb2 = lane + 64 < kRbW;
    const bool p_dof1 = lane + 32 < 2 * DEX_NUM_DOF;
    const bool p_tgt = A.o_tgt >= 0 && lane < DEX_NUM_DOF;
    // lanes 0..17: contact rows (5 monitored + box), lanes 18..23: box root state (pos, linvel); slots I_CF + lane
    const bool is_box = BG && lane >= 18 && lane < 24;
    const bool p_x = is_box || (lane < (BG ? 18 : 15) && (lane >= 15 || !stale));
    const int xrow = is_box ? P.num_actors * 13 : P.num_contact_bodies * 3;
    int xoff;
    if (is_box) {
      const int k = lane - 18;
      xoff = P.box_actor * 13 + (k < 3 ? k : k + 4);
    } else {
      const int f = min(lane / 3, 5);
      xoff = A.cf_off[f] + (lane - 3 * f);
    }
    const float* gx = (is_box ? A.sim.actor_root_state : A.sim.net_contact_force) + (size_t)(e0 + warp) * xrow + xoff;
    const float* g_dof = A.sim.dof_state + (size_t)(e0 + warp) * (2 * DEX_NUM_DOF) + lane;
    const float* g_rb = A.sim.rigid_body_state + (size_t)(e0 + warp) * A.rb_pitch;
    const float* g_tgt = A.st.full_dof_targets + (size_t)(e0 + warp) * DEX_NUM_DOF + lane;
    const unsigned rb_step = kPostWarps * A.rb_pitch, x_step = kPostWarps * xrow;
    constexpr int R = kTile / kPostWarps;  // env rows per warp
    // ---- the Isaac Gym sim rows (80 % of the tile's bytes) go first.  With DEX_FLAG_EARLY_SIM_LOADS they are issued
    // BEFORE griddepcontrol.wait: the kernel ahead in the stream (dex_pre_physics of the same step) does not write
    // them, so they overlap its execution and the launch gap.  Without the flag the wait sits in front (a second
    // griddepcontrol.wait is a no-op); AoS rows ahead of the SoA state rows is also the faster order there
    // (-1.2 us at 16 K envs: the long strided rows have their DRAM pages opened first).
    if ((A.flags & DEX_FLAG_EARLY_SIM_LOADS) == 0) pdl_wait();
    float v[R][7];
#pragma unroll
    for (int i = 0; i < R; ++i) {
      if (FULL || warp + kPostWarps * i < ne) {
        v[i][0] = __ldcg(g_dof + i * (kPostWarps * 2 * DEX_NUM_DOF));
        if (p_dof1) v[i][1] = __ldcg(g_dof + i * (kPostWarps * 2 * DEX_NUM_DOF) + 32);
        const float* rb = g_rb + i * rb_step;
        if (p_rb0) v[i][2] = __ldg(rb + rbo0);
        v[i][3] = __ldg(rb + rbo1);
        if (p_rb2) v[i][4] = __ldg(rb + rbo2);
        if (p_x) v[i][5] = __ldcg(gx + i * x_step);
      }
    }
    pdl_wait();  // everything above touched only kernel parameters and the sim tensors
    // ---- SoA state rows (IN table, lane = env, one 128 B line per request); every load of the tile is in flight
    // before the first store: one DRAM round trip for the whole stage-in instead of one per batch.
    const bool act = FULL || lane < ne;
    const size_t col = (size_t)e0 + (act ? lane : 0);
    const size_t colb = col * 4;
    constexpr int MAXB = kMaxIn / kRowBatch;  // 5 batches of 4 rows per warp
    const int nbatch = A.n_in_f32 / kRowBatch;
    float w[MAXB][kRowU];
#pragma unroll
    for (int b = 0; b < MAXB; ++b) {
      if (b < nbatch && act) {
#pragma unroll
        for (int u = 0; u < kRowU; ++u)
          w[b][u] = __ldcg(reinterpret_cast<const float*>(A.in_ptr[b * kRowBatch + warp + kPostWarps * u] + colb));
      }
    }
    // typed rows (int64 counters, bool flags): at most a few per warp, loaded as 32-bit values
    const int r_typed = A.n_in_f32 + warp;
    float wt = 0.0f;
    if (act && r_typed < A.n_in) {
      if (r_typed < A.n_in_i64)  // int64 state is small (counts, stage): keep the low word
        wt = __int_as_float((int)__ldcg(reinterpret_cast<const long long*>(A.in_ptr[r_typed]) + col));
      else
        wt = __ldcg(reinterpret_cast<const unsigned char*>(A.in_ptr[r_typed]) + col) ? 1.0f : 0.0f;
    }
    if (p_tgt) {  // full_dof_targets is written by dex_pre_physics: always after the wait
#pragma unroll
      for (int i = 0; i < R; ++i)
        if (FULL || warp + kPostWarps * i < ne) v[i][6] = __ldcg(g_tgt + i * (kPostWarps * DEX_NUM_DOF));
    }
    // ---- stores ----
    const unsigned Eb = sbase + (unsigned)(lane * S) * 4u;
#pragma unroll
    for (int b = 0; b < MAXB; ++b) {
      if (b < nbatch && act) {
#pragma unroll
        for (int u = 0; u < kRowU; ++u)
          sts_f32(Eb + (unsigned)A.in_meta[b * kRowBatch + warp + kPostWarps * u] * 4u, w[b][u]);
      }
    }
    if (act && r_typed < A.n_in) sts_f32(Eb + (unsigned)A.in_meta[r_typed] * 4u, wt);
    const unsigned E0 = sbase + (unsigned)(warp * S + lane) * 4u;
    const unsigned estep = (unsigned)(kPostWarps * S) * 4u;
    const unsigned otgt = (unsigned)A.o_tgt * 4u;
#pragma unroll
    for (int i = 0; i < R; ++i) {
      if (FULL || warp + kPostWarps * i < ne) {
        const unsigned E = E0 + i * estep;
        sts_f32(E + I_DOF * 4, v[i][0]);
        if (p_dof1) sts_f32(E + (I_DOF + 32) * 4, v[i][1]);
        if (p_rb0) sts_f32(E + I_RB * 4, v[i][2]);
        sts_f32(E + (I_RB + 32) * 4, v[i][3]);
        if (p_rb2) sts_f32(E + (I_RB + 64) * 4, v[i][4]);
        if (p_x) sts_f32(E + I_CF * 4, v[i][5]);
        if (p_tgt) sts_f32(E + otgt, v[i][6]);
      }
    }
    // more than one typed row per warp (only with the common-reward state): the rest, after the burst
    for (int r = r_typed + kPostWarps; act && r < A.n_in; r += kPostWarps) {
      float t;
      if (r < A.n_in_i64)
        t = __int_as_float((int)__ldcg(reinterpret_cast<const long long*>(A.in_ptr[r]) + col));
      else
        t = __ldcg(reinterpret_cast<const unsigned char*>(A.in_ptr[r]) + col) ? 1.0f : 0.0f;
      sts_f32(Eb + (unsigned)A.in_meta[r] * 4u, t);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// phase 1: items (lane = env, item type warp-uniform; slab in, slab out)
// ------------------------------------------------------------------------------------------------
template <bool BG>
__device__ __forceinline__ void post_items(const PostArgs& A, float* smem, const int ne) {
  const DexParams& P = A.p;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane >= ne) return;
  float* E = smem + lane * A.stride;
  const float* rb = E + I_RB;
  const float dt = P.control_dt;
  const bool common = A.common != 0;
  const bool first = (A.flags & DEX_FLAG_FIRST_REWARD_CALL) != 0;

  // item ranges (heavy items first so the round-robin spreads them over the warps)
  const int i_contact = 10, i_pair = 15, i_hand = BG ? 25 : 15, i_box = i_hand + 1;
  const int i_dof = BG ? i_box + 1 : i_box, n_items = i_dof + DEX_NUM_DOF;

  for (int it = warp; it < n_items; it += kPostWarps) {
    if (it < i_contact) {
      // ---- tip / pad pose f: hand-frame pose (observation_encoder.py:915-986), distances to the box ----
      const int f = it;
      const bool is_pad = f >= DEX_NUM_FINGERS;
      const int g = is_pad ? f - DEX_NUM_FINGERS : f;
      const float* p = rb + 13 + 7 * f;
      const float px = p[0], py = p[1], pz = p[2];
      if (is_pad ? A.want_hf_pad : A.want_hf_tip) {
        const Quat hq{rb[3], rb[4], rb[5], rb[6]};
        const Vec3 ph = quat_rotate_inverse(hq, Vec3{px - rb[0], py - rb[1], pz - rb[2]});
        const Quat qh = quat_mul(quat_conj(hq), Quat{p[3], p[4], p[5], p[6]});
        float* o = E + A.o_hf + (is_pad ? HF_PAD : HF_TIP) + 7 * g;
        o[0] = ph.x;
        o[1] = ph.y;
        o[2] = ph.z;
        o[3] = qh.x;
        o[4] = qh.y;
        o[5] = qh.z;
        o[6] = qh.w;
      }
      if (BG) {  // blind_grasping_task.py:684-724, 1178-1192
        const float* box = E + I_BOX;
        const float d = norm3(px - box[0], py - box[1], pz - box[2]);
        if (is_pad) {
          E[C_PADBOX + g] = d;
          E[C_HDIFF + g] = fabsf(pz - box[2]);
        } else {
          E[C_TIPBOX + g] = d;
        }
      }
    } else if (it < i_pair) {
      // ---- contact sensor f (observation_encoder.py:658-735, 323-361; reward_calculator.py:183-199) ----
      const int f = it - i_contact;
      const float* cf = E + I_CF + 3 * f;
      const float mag = norm3(cf[0], cf[1], cf[2]);
      const bool bin = mag > P.contact_binary_threshold;
      const float dur = E[S_CDUR + f], prevb = E[S_PBIN + f];
      const float ndur = (bin && prevb == 0.0f) ? 1.0f : (bin ? dur + 1.0f : 0.0f);
      E[S_CDUR + f] = ndur;
      E[S_PBIN + f] = bin ? 1.0f : 0.0f;
      E[C_CMAG + f] = mag;
      E[C_CDURS + f] = ndur * dt;
      if (common) {
        float* cp = E + A.o_cpv;
        const bool c_now = mag > 0.1f;  // hard-coded 0.1 N (reward_calculator.py:186)
        const bool c_prev = first ? c_now : (cp[CP_PCON + f] != 0.0f);
        cp[CP_CNOW + f] = c_now ? 1.0f : 0.0f;
        cp[CP_CHG + f] = (c_now != c_prev) ? 1.0f : 0.0f;
      }
    } else if (BG && it < i_hand) {
      // ---- pad pair distance, triu order (blind_grasping_task.py:726-771) ----
      const int k = it - i_pair;
      // k -> (a, b): rows of the strict upper triangle of a 5x5 matrix start at 0, 4, 7, 9
      const int a = (k >= 4) + (k >= 7) + (k >= 9);
      const int b = k - (a == 0 ? 0 : a == 1 ? 4 : a == 2 ? 7 : 9) + a + 1;
      const float* pa = rb + 13 + 35 + 7 * a;
      const float* pb = rb + 13 + 35 + 7 * b;
      E[C_PAIR + k] = norm3(pa[0] - pb[0], pa[1] - pb[1], pa[2] - pb[2]);
    } else if (it == i_hand) {
      // ---- ARR-aligned pose (:1483-1521), hand -> box distance, hand velocity reward terms ----
      if (A.want_arr) {
        const float sh = 0.7071067811865476f;
        const Quat qa = quat_mul(Quat{rb[3], rb[4], rb[5], rb[6]}, Quat{-0.0f, -sh, -0.0f, sh});
        float* o = E + A.o_hf + HF_ARR;
        o[0] = qa.x;
        o[1] = qa.y;
        o[2] = qa.z;
        o[3] = qa.w;
      }
      if (BG) {
        const float* box = E + I_BOX;
        E[C_HANDBOX] = norm3(rb[0] - box[0], rb[1] - box[1], rb[2] - box[2]);
      }
      if (common) {  // reward_calculator.py:100-181
        float* cp = E + A.o_cpv;
        float hv[3], ha[3], phv[3], pha[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          hv[c] = rb[7 + c];
          ha[c] = rb[10 + c];
          phv[c] = first ? hv[c] : cp[CP_PHV + c];
          pha[c] = first ? ha[c] : cp[CP_PHA + c];
        }
        cp[CP_RHAND + 0] = expf(-0.2f * norm3(hv[0], hv[1], hv[2]));
        cp[CP_RHAND + 1] = expf(-0.2f * norm3(ha[0], ha[1], ha[2]));
        cp[CP_RHAND + 2] = expf(-0.5f * norm3(hv[0] - phv[0], hv[1] - phv[1], hv[2] - phv[2]));
        cp[CP_RHAND + 3] = expf(-0.5f * norm3(ha[0] - pha[0], ha[1] - pha[1], ha[2] - pha[2]));
      }
    } else if (BG && it == i_box) {
      // ---- pad centroid, box stability inputs (blind_grasping_task.py:746-771, 1334-1364) ----
      const float* box = E + I_BOX;
      float cen[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        cen[c] = (rb[13 + 35 + c] + rb[13 + 35 + 7 + c] + rb[13 + 35 + 14 + c]) / 3.0f;
        E[C_CEN + c] = cen[c];
      }
      E[C_CDIST] = norm3(cen[0] - box[0], cen[1] - box[1], cen[2] - box[2]);
      const float* b0 = E + S_BOX0;
      E[C_DRIFT] = norm3(box[0] - b0[0], box[1] - b0[1], box[2] - b0[2]);
      E[C_VMAG] = norm3(box[3], box[4], box[5]);
      const float* cfb = E + I_CF + 15;
      E[C_BOXHAS] = (norm3(cfb[0], cfb[1], cfb[2]) > P.contact_binary_threshold) ? 1.0f : 0.0f;
    } else {
      // ---- DOF d: finite-difference velocity (observation_encoder.py:298-321), common-reward partial terms ----
      const int d = it - i_dof;
      const float pos = E[I_DOF + 2 * d];
      const float dp = pos - E[S_PDP + d];
      E[S_PDP + d] = div_rn(dp, dt);  // joints at rest (dp == 0) are the common case: see div_rn
      if (common && d >= DEX_NUM_BASE) {  // reward_calculator.py:94-98,138-160: Isaac dof_vel, not the finite difference
        float* cp = E + A.o_cpv;
        const int j = d - DEX_NUM_BASE;
        const float iv = E[I_DOF + 2 * d + 1];
        const float pv = first ? iv : cp[CP_PFDV + j];
        cp[CP_FV2 + j] = iv * iv;
        const float acc = iv - pv;
        cp[CP_FA2 + j] = acc * acc;
        const float nj = div_rn(2.0f * (pos - P.dof_lower[d]), P.dof_upper[d] - P.dof_lower[d]) - 1.0f;
        cp[CP_JL + j] = clampf(fabsf(nj) - 0.8f, 0.0f, 1.0f);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// phase 2: per-env scalar part (warp 0, lane = env).  Returns should_reset of this lane.
// ------------------------------------------------------------------------------------------------
template <bool BG>
__device__ __forceinline__ bool post_scalar(const PostArgs& A, float* smem, const int e0, const int ne,
                                            const unsigned stamp, unsigned* s_rbits) {
  const DexParams& P = A.p;
  const int lane = threadIdx.x & 31;
  const bool active = lane < ne;
  float* E = smem + lane * A.stride;
  const float* rb = E + I_RB;
  const float dt = P.control_dt;
  const bool full = A.mode == DEX_MODE_FULL;

  bool should_reset = false, t_success = false, t_failure = false, t_timeout = false;
  int ep_len_done = 0;
  bool crit[DEX_NUM_CRITERIA];
#pragma unroll
  for (int c = 0; c < DEX_NUM_CRITERIA; ++c) crit[c] = false;

  if (active) {
    // episode_time from the PRE-increment step count (dexhand_base.py:703-712)
    const int step_before = __float_as_int(E[S_STEP]);
    E[C_EPTIME] = (float)step_before * dt;
    float min_tip_z = rb[13 + 2];
#pragma unroll
    for (int f = 1; f < DEX_NUM_FINGERS; ++f) min_tip_z = fminf(min_tip_z, rb[13 + 7 * f + 2]);

    // ---------------- BlindGrasping task observations + FSM (blind_grasping_task.py:549-978) ----
    float avg_hdiff = 0.0f, grasp_duration = 0.0f, min_tip_box = 0.0f, box_z = 0.0f;
    bool thumb_c = false, others_c = false, grasp = false, j2 = false, j3 = false;
    int stage = 0, sds = 0;
    if (BG) {
      box_z = E[I_BOX + 2];
      float dsum = 0.0f, hsum = 0.0f;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        dsum += E[C_PADBOX + f];
        hsum += E[C_HDIFF + f];
      }
      avg_hdiff = div_rn(hsum, 5.0f);
      E[C_AVGDIST] = div_rn(dsum, 5.0f);
      E[C_AVGHDIFF] = avg_hdiff;
      // finger-box contacts :773-833
      const bool box_has = E[C_BOXHAS] != 0.0f;
      bool cbin[DEX_NUM_FINGERS], fbc[DEX_NUM_FINGERS];
      int n_on_box = 0;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        cbin[f] = E[S_PBIN + f] != 0.0f;
        fbc[f] = cbin[f] && box_has && (E[C_PADBOX + f] < P.proximity_threshold);
        n_on_box += fbc[f] ? 1 : 0;
      }
      thumb_c = fbc[0];
      others_c = fbc[1] || fbc[2] || fbc[3] || fbc[4];
      grasp = thumb_c && others_c;
      E[C_THUMB] = thumb_c ? 1.0f : 0.0f;
      E[C_OTHERS] = others_c ? 1.0f : 0.0f;
      E[C_GRASP] = grasp ? 1.0f : 0.0f;
      // success duration :626-648
      const bool cond = (box_z > P.success_height_threshold) && (n_on_box >= P.min_fingers_for_grasp);
      sds = cond ? __float_as_int(E[S_SDS]) + 1 : 0;
      E[S_SDS] = __int_as_float(sds);
      E[M_SCM] = cond ? 1.0f : 0.0f;
      grasp_duration = (float)sds * dt;
      E[C_GRASPDUR] = grasp_duration;
      // stage FSM :835-956 (plain fp32 accumulation of (float)control_dt decides the transition step)
      stage = __float_as_int(E[S_STAGE]);
      float tis = E[S_TIS] + dt;
      float scd = E[S_SCD];
      const bool thumb_other = cbin[0] && (cbin[1] || cbin[2] || cbin[3] || cbin[4]);
      if (stage == 2) scd = thumb_other ? scd + dt : 0.0f;
      const bool s1_done = (stage == 1) && (tis >= P.stage1_duration);
      const bool s2_done = (stage == 2) && ((scd >= P.stage2_contact_success_threshold) || (tis >= P.stage2_duration));
      j2 = s1_done;
      j3 = s2_done;
      if (s1_done) stage = 2;
      if (s2_done) stage = 3;
      if (s1_done || s2_done) {
        tis = 0.0f;
        scd = 0.0f;
      }
      E[S_STAGE] = __int_as_float(stage);
      E[S_TIS] = tis;
      E[S_SCD] = scd;
      E[M_J2] = j2 ? 1.0f : 0.0f;
      E[M_J3] = j3 ? 1.0f : 0.0f;
      float prog = 0.0f;  // :958-978
      if (stage == 1) prog = clampf(div_rn(tis, P.stage1_duration), 0.0f, 1.0f);
      if (stage == 2) prog = clampf(div_rn(tis, P.stage2_duration), 0.0f, 1.0f);
      if (stage == 3) prog = 1.0f;
      E[C_STAGEF] = (float)stage;
      E[C_PROG] = prog;
      TL_MARK(7);
      min_tip_box = E[C_TIPBOX];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) min_tip_box = fminf(min_tip_box, E[C_TIPBOX + f]);
    }

    if (full) {
      // step_processor.py:80
      const int step_now = step_before + 1;
      E[S_STEP] = __int_as_float(step_now);
      ep_len_done = step_now;

      // ---------------- termination (step_processor.py:133-181, termination_manager.py:120-278) ----
      if (P.has_height_safety)
        crit[DEX_C_HITTING_GROUND] = (rb[2] < P.handbase_threshold) || (min_tip_z < P.fingertip_threshold);
      bool pregrasp_failed = false, contact_failed = false, grasp_lost = false;
      if (BG) {
        crit[DEX_C_GRASP_LIFT_SUCCESS] = sds >= P.contact_duration_threshold_steps;
        crit[DEX_C_BOX_TOO_FAR] = E[C_HANDBOX] > P.max_box_distance;
        if (j2) {  // _evaluate_pregrasp_quality :1334-1364
          const bool h_ok = (E[C_HDIFF + 0] <= P.height_tolerance) && (E[C_HDIFF + 1] <= P.height_tolerance) &&
                            (E[C_HDIFF + 2] <= P.height_tolerance);
          const bool c_ok = E[C_CDIST] <= P.centroid_tolerance;
          const bool s_ok = (E[C_DRIFT] <= P.position_drift_tolerance) && (E[C_VMAG] <= P.velocity_tolerance);
          pregrasp_failed = !(h_ok && c_ok && s_ok);
        }
        if (j3) contact_failed = !grasp;
        if (stage == 3) grasp_lost = !grasp;
        crit[DEX_C_STAGE1_PREGRASP_FAILED] = pregrasp_failed;
        crit[DEX_C_STAGE2_CONTACT_FAILED] = contact_failed;
        crit[DEX_C_STAGE3_GRASP_LOST] = grasp_lost;
        E[M_FT0] = pregrasp_failed ? 1.0f : 0.0f;
        E[M_FT1] = contact_failed ? 1.0f : 0.0f;
        E[M_FT2] = grasp_lost ? 1.0f : 0.0f;
      }
      bool ep_s = false, ep_f = false;
#pragma unroll
      for (int c = 0; c < DEX_NUM_CRITERIA; ++c) {
        crit[c] = crit[c] && (P.criterion_active[c] != 0);
        if (c == DEX_C_GRASP_LIFT_SUCCESS)
          ep_s = ep_s || crit[c];
        else
          ep_f = ep_f || crit[c];
      }
      const bool timeout = step_now >= (P.max_episode_length - 1);
      should_reset = ep_s || ep_f || timeout;
      t_success = ep_s;
      t_failure = ep_f && !ep_s;
      t_timeout = timeout && !ep_s && !ep_f;

      E[C_MINTIPZ] = min_tip_z;
      if (BG) E[C_MINTIPBOX] = min_tip_box;
      E[M_RESET] = should_reset ? 1.0f : 0.0f;
      E[M_SUCC] = t_success ? 1.0f : 0.0f;
      E[M_FAIL] = t_failure ? 1.0f : 0.0f;
      E[M_TIMEOUT] = t_timeout ? 1.0f : 0.0f;
    }
  }

  TL_MARK(8);
  if (full) {
    // reset ballot word + cross-env counters (termination_manager.py:164-185,259-266)
    const unsigned fullmask = 0xffffffffu;
    const unsigned rbits = __ballot_sync(fullmask, should_reset);
    if (lane == 0) {
      A.out.reset_bits[e0 >> 5] = rbits;
      *s_rbits = rbits;  // the CTA's own copy for the fused reset (no L2 round trip)
    }
    const unsigned bs = __ballot_sync(fullmask, t_success), bf = __ballot_sync(fullmask, t_failure),
                   bt = __ballot_sync(fullmask, t_timeout);
    unsigned bc[DEX_NUM_CRITERIA];
#pragma unroll
    for (int c = 0; c < DEX_NUM_CRITERIA; ++c) bc[c] = __ballot_sync(fullmask, crit[c]);
    int len = 0;
    if (rbits) {  // sum of the finished episodes' lengths (warp-uniform branch)
      len = should_reset ? ep_len_done : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) len += __shfl_xor_sync(fullmask, len, o);
    }
    TL_MARK(9);
    if (lane == 0) {
      if (A.fused_epilogue) {
        // Fused epilogue: the tile PUBLISHES its reset word and its counter contributions as three self-validating
        // 8-byte records (aligned 8-byte stores are single-copy atomic; bits 48..63 = stamp of this step).  The
        // CTA of the last tile polls the stamps: no fence, no atomic, no contention on shared counters here.
        //   rec0: [0,32) reset word | [32,38) #success | [38,44) #failure     (#timeout = popc(word) - both)
        //   rec1: [0,36) six 6-bit criterion hit counts
        //   rec2: [0,48) sum of the finished episodes' lengths
        const unsigned long long st = (unsigned long long)stamp << 48;
        unsigned long long r1 = 0;
#pragma unroll
        for (int c = 0; c < DEX_NUM_CRITERIA; ++c) r1 |= (unsigned long long)__popc(bc[c]) << (6 * c);
        // three planes of [tiles] words (not [tiles][3]): the epilogue CTA reads them lane-contiguously
        const size_t tiles = (size_t)((A.n + kTile - 1) / kTile);
        unsigned long long* rec = reinterpret_cast<unsigned long long*>(A.ep.tile_records) + (size_t)(e0 >> 5);
        __stcg(rec, st | rbits | ((unsigned long long)__popc(bs) << 32) | ((unsigned long long)__popc(bf) << 38));
        __stcg(rec + tiles, st | r1);
        __stcg(rec + 2 * tiles, st | (unsigned long long)(unsigned)len);
      } else {
#ifndef EXP_NO_COUNTERS
        if (bs) atomicAdd(A.out.counters + DEX_CNT_SUCCESS, __popc(bs));
        if (bf) atomicAdd(A.out.counters + DEX_CNT_FAILURE, __popc(bf));
        if (bt) atomicAdd(A.out.counters + DEX_CNT_TIMEOUT, __popc(bt));
        if (rbits) atomicAdd(A.out.counters + DEX_CNT_RESET, __popc(rbits));
#pragma unroll
        for (int c = 0; c < DEX_NUM_CRITERIA; ++c)
          if (bc[c]) atomicAdd(A.out.counters + DEX_CNT_CRIT0 + c, __popc(bc[c]));
        if (rbits) atomicAdd(A.out.counters + DEX_CNT_EPISODE_LENGTH, len);
#endif
      }
    }
    TL_MARK(10);
  }
  return should_reset;
}

// ------------------------------------------------------------------------------------------------
// phase 2b: reward terms as work items (all warps; lane = env, term index warp-uniform).
// reward_calculator.py:66-221 (common terms), blind_grasping_task.py:980-1208 (task terms).
// Writes the raw term and the weighted term into the component rows of the slab.
// ------------------------------------------------------------------------------------------------
template <bool BG>
__device__ __forceinline__ void post_reward_terms(const PostArgs& A, float* smem, const int ne, const int nw) {
  const DexParams& P = A.p;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane >= ne) return;
  float* E = smem + lane * A.stride;
  const float* rb = E + I_RB;
  float* comp = E + C_COMP;
  const int nterms = BG ? DEX_NUM_REWARD_TERMS : DEX_NUM_COMMON_TERMS;
  for (int i = warp; i < nterms; i += nw) {  // nw warps share the terms
    const float w = P.reward_weight[i];
    if (w == 0.0f) continue;  // reward_calculator.py:247-251: zero-weight terms are not materialised
    float r = 0.0f;
    if (i < DEX_NUM_COMMON_TERMS) {
      const float* cp = E + A.o_cpv;
      switch (i) {
        case DEX_R_ALIVE: r = 1.0f; break;
        case DEX_R_HEIGHT_SAFETY: r = clampf(1.0f - expf(-(E[C_MINTIPZ] - 0.02f) * 20.0f), 0.0f, 1.0f); break;
        case DEX_R_FINGER_VELOCITY: {
          float a = 0.0f;
#pragma unroll
          for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) a += cp[CP_FV2 + j];
          r = expf(-0.1f * sqrtf(a));
        } break;
        case DEX_R_HAND_VELOCITY: r = cp[CP_RHAND + 0]; break;
        case DEX_R_HAND_ANGULAR_VELOCITY: r = cp[CP_RHAND + 1]; break;
        case DEX_R_JOINT_LIMIT: {
          float a = 0.0f;
#pragma unroll
          for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) a += cp[CP_JL + j];
          r = 1.0f - div_rn(a, 20.0f);
        } break;
        case DEX_R_FINGER_ACCELERATION: {
          float a = 0.0f;
#pragma unroll
          for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) a += cp[CP_FA2 + j];
          r = expf(-2.0f * sqrtf(a));
        } break;
        case DEX_R_HAND_ACCELERATION: r = cp[CP_RHAND + 2]; break;
        case DEX_R_HAND_ANGULAR_ACCELERATION: r = cp[CP_RHAND + 3]; break;
        default: {  // DEX_R_CONTACT_STABILITY
          float a = 0.0f;
#pragma unroll
          for (int f = 0; f < DEX_NUM_FINGERS; ++f) a += cp[CP_CHG + f];
          r = expf(-a);
        } break;
      }
    } else if (BG) {
      const int stage = __float_as_int(E[S_STAGE]);
      const float m1 = (stage == 1) ? 1.0f : 0.0f, m2 = (stage == 2) ? 1.0f : 0.0f, m3 = (stage == 3) ? 1.0f : 0.0f;
      switch (i) {
        // stage 1 :1055-1112
        case DEX_R_S1_HEIGHT_ALIGNMENT: r = expf(-P.height_alignment_decay * E[C_AVGHDIFF]) * m1; break;
        case DEX_R_S1_CENTROID_POSITIONING: r = expf(-P.centroid_positioning_decay * E[C_CDIST]) * m1; break;
        case DEX_R_S1_OBJECT_STABILITY: r = expf(-P.object_stability_decay * (E[C_DRIFT] + E[C_VMAG])) * m1; break;
        case DEX_R_S1_FINGER_HEIGHT_CONSISTENCY: {  // torch.var(dim=1), unbiased, z of the first three pads
          const float z0 = rb[13 + 35 + 2], z1 = rb[13 + 35 + 7 + 2], z2 = rb[13 + 35 + 14 + 2];
          const float mean = (z0 + z1 + z2) / 3.0f;
          const float d0 = z0 - mean, d1 = z1 - mean, d2 = z2 - mean;
          const float var = (d0 * d0 + d1 * d1 + d2 * d2) / 2.0f;
          r = expf(-P.height_consistency_decay * var) * m1;
        } break;
        case DEX_R_S1_THUMB_ROTATION: r = expf(-5.0f * fabsf(E[I_DOF + 2 * DEX_NUM_BASE] - P.half_pi)) * m1; break;
        // stage 2 :1114-1150
        case DEX_R_S2_THUMB_CONTACT: r = E[C_THUMB] * m2; break;
        case DEX_R_S2_OTHER_FINGERS_CONTACT: r = E[C_OTHERS] * m2; break;
        case DEX_R_S2_GRASP_ACHIEVEMENT: r = E[C_GRASP] * m2; break;
        case DEX_R_S2_FINGERPAD_PROXIMITY:
          r = expf(-P.fingerpad_proximity_decay * fmaxf(E[C_MINTIPBOX], P.min_reward_distance)) * m2;
          break;
        case DEX_R_S2_BASE_STABILITY: {
          float bv2 = 0.0f;
#pragma unroll
          for (int d = 0; d < DEX_NUM_BASE; ++d) bv2 += E[S_PDP + d] * E[S_PDP + d];  // finite-difference base velocity
          r = expf(-P.base_stability_decay * sqrtf(bv2)) * m2;
        } break;
        // stage 3 :1152-1176
        case DEX_R_S3_OBJECT_HEIGHT: r = clampf(div_rn(E[I_BOX + 2] - P.box_z, P.lift_height_range), 0.0f, 1.0f) * m3; break;
        case DEX_R_S3_GRASP_MAINTENANCE: r = E[C_GRASP] * m3; break;
        case DEX_R_S3_GRASP_DURATION:
          r = clampf(div_rn(E[C_GRASPDUR], P.contact_duration_threshold_s), 0.0f, 1.0f) * m3;
          break;
        // completion bonuses :1025-1048, penetration :1194-1208
        case DEX_R_S1_COMPLETION: r = (E[M_J2] != 0.0f && E[M_FT0] == 0.0f) ? 1.0f : 0.0f; break;
        case DEX_R_S2_COMPLETION: r = (E[M_J3] != 0.0f && E[M_FT1] == 0.0f) ? 1.0f : 0.0f; break;
        default:  // DEX_R_PENETRATION_PENALTY
          r = fmaxf(P.penetration_threshold - E[C_MINTIPBOX], 0.0f) * P.penetration_depth_scale;
          break;
      }
    }
    const int row = P.component_row[i];
    comp[row] = r;
    comp[row + 1] = r * w;
  }
}

// ------------------------------------------------------------------------------------------------
// phase 2c: ordered weighted sum (reward_calculator.py:223-275), termination rewards
// (termination_manager.py:280-321, step_processor.py:204-219), optional fused RewardComponentObserver.
// Warp 0, lane = env.  The rows it produces leave straight from registers (coalesced).
// ------------------------------------------------------------------------------------------------
template <bool BG>
__device__ __forceinline__ void post_total(const PostArgs& A, float* smem, const int e0, const int ne) {
  const DexParams& P = A.p;
  const int lane = threadIdx.x & 31;
  if (lane >= ne) return;
  const size_t N = (size_t)A.n;
  const size_t env = (size_t)e0 + lane;
  float* E = smem + lane * A.stride;
  float* comp = E + C_COMP;
  float total = 0.0f;
#pragma unroll
  for (int i = 0; i < DEX_NUM_REWARD_TERMS; ++i) {
    if (!BG && i >= DEX_NUM_COMMON_TERMS) break;
    if (P.reward_weight[i] != 0.0f) total += comp[P.component_row[i] + 1];  // dict order: common, then task
  }
  const bool t_success = E[M_SUCC] != 0.0f, t_failure = E[M_FAIL] != 0.0f, t_timeout = E[M_TIMEOUT] != 0.0f;
  const float rs = t_success ? P.success_reward : 0.0f;
  const float rf = t_failure ? P.failure_reward : 0.0f;
  const float rt = t_timeout ? P.timeout_reward : 0.0f;
  float rew = total;
  rew += rs;
  rew += rf;
  rew += rt;
  float* g = A.out.reward_components + env;
  const float tv[7] = {total, t_success ? 1.0f : 0.0f, rs, t_failure ? 1.0f : 0.0f, rf, t_timeout ? 1.0f : 0.0f, rt};
  g[(size_t)P.total_row * N] = tv[0];
#pragma unroll
  for (int k = 0; k < 6; ++k) g[(size_t)(P.termination_row + k) * N] = tv[1 + k];
  A.out.rew_buf[env] = rew;

  // ---------------- fused RewardComponentObserver (rl/reward_observer.py:160-293) -----------
  if (A.out.episode_sums != nullptr) {
    comp[P.total_row] = tv[0];
#pragma unroll
    for (int k = 0; k < 6; ++k) comp[P.termination_row + k] = tv[1 + k];
    const bool should_reset = E[M_RESET] != 0.0f;
    const int rows = P.num_component_rows;
    const int len = A.out.episode_lengths[env] + 1;
    const int stride = 2 + rows;
    double* tot = A.out.stat_totals + (size_t)((env >> 5) % max(1, A.out.stat_slices)) * 4 * stride;  // block of this tile
    const int cls = t_success ? 1 : (t_failure ? 2 : 3);
    if (should_reset) {
      atomicAdd(tot + 0, 1.0);
      atomicAdd(tot + 1, (double)len);
      atomicAdd(tot + cls * stride + 0, 1.0);
      atomicAdd(tot + cls * stride + 1, (double)len);
    }
    for (int c = 0; c < rows; ++c) {
      const float s = A.out.episode_sums[(size_t)c * N + env] + comp[c];
      if (should_reset) {
        atomicAdd(tot + 2 + c, (double)s);
        atomicAdd(tot + cls * stride + 2 + c, (double)s);
      }
      A.out.episode_sums[(size_t)c * N + env] = should_reset ? 0.0f : s;
    }
    A.out.episode_lengths[env] = should_reset ? 0 : len;
  }
}

// ------------------------------------------------------------------------------------------------
// phase 3: stage-out
// ------------------------------------------------------------------------------------------------
// rows of a [N, W] row-major buffer gathered from the slab through a column map (shared copy of the map):
// lane = column (mod 32), so every global store is a full coalesced 128 B request.  Only the 32-column chunks
// whose bit is set in `chunks` are written; the env rows are spread over NW warps, this warp is number `w`.
// NORM: the optional policy-input stage rides on the same registers:
//   norm.dst[env, k] = clamp((obs[k] - mean[k]) * scale[k], -clip, clip),  scale = 1 / sqrt(var + eps)
// i.e. the frozen-statistics form of rl_games' RunningMeanStd.forward (normalize_input, cfg/train/BaseTaskPPO.yaml:44)
// (fp32 subtract, multiply by the caller's fp32 reciprocal - within 2 ulp of the torch division -, clamp that lets
// NaN through like torch.clamp).
struct NormOut {
  float* dst;
  const float* mean;  // shared-memory copies of DexOutBuffers.norm_mean / norm_scale
  const float* scale;
  float clip;
};
template <int NW, bool NORM = false>
__device__ __forceinline__ void gather_rows_out(float* __restrict__ dst, const int W, const short* s_map,
                                                const float* smem, const int S, const int e0, const int ne,
                                                unsigned chunks, const int w, const NormOut norm = NormOut{}) {
  const int lane = threadIdx.x & 31;
  constexpr int R = (kTile + NW - 1) / NW;  // env rows per warp
  const unsigned s0 = (unsigned)__cvta_generic_to_shared(smem) + (unsigned)(w * S) * 4u;
  const unsigned sstep = (unsigned)(NW * S) * 4u;
  const size_t off0 = ((size_t)e0 + w) * W + lane;
  float* d0 = dst + off0;
  const size_t dstep = (size_t)NW * W;
  const int nchunk = (W + 31) >> 5;
  if (nchunk < 32) chunks &= (1u << nchunk) - 1u;
  while (chunks) {
    const int c = __ffs(chunks) - 1;
    chunks &= chunks - 1;
    const int k = 32 * c + lane;
    if (k < W) {
      const unsigned a = s0 + (unsigned)s_map[k] * 4u;
      float v[R];
#pragma unroll
      for (int i = 0; i < R; ++i) {
        if (NORM) v[i] = 0.0f;
        if (w + NW * i < ne) v[i] = lds_f32(a + i * sstep);
      }
      float* d = d0 + 32 * c;
#pragma unroll
      for (int i = 0; i < R; ++i) {
#ifdef EXP_NO_OBS
        if (v[i] == 123.456f)
#endif
        if (w + NW * i < ne) *d = v[i];
        d += dstep;
      }
      if (NORM) {
        const float m = norm.mean[k], sc = norm.scale[k];  // shared copies
        float* d2 = norm.dst + off0 + 32 * c;
#pragma unroll
        for (int i = 0; i < R; ++i) {
          const float y = (v[i] - m) * sc;
          if (w + NW * i < ne) *d2 = (y != y) ? y : fminf(fmaxf(y, -norm.clip), norm.clip);
          d2 += dstep;
        }
      }
    }
  }
}

// f32 rows [r_begin, r_end) of the OUT table: lane = env, one 128 B request per row; rows spread over NW warps
template <int NW>
__device__ __forceinline__ void post_rows_f32(const PostArgs& A, const float* smem, const int e0, const int ne,
                                              const int r_begin, const int r_end, const int w) {
  const int lane = threadIdx.x & 31;
  if (lane >= ne) return;
  const unsigned Eb = (unsigned)__cvta_generic_to_shared(smem) + (unsigned)(lane * A.stride) * 4u;
  const size_t colb = ((size_t)e0 + lane) * 4;
  for (int r0 = r_begin + w; r0 < r_end; r0 += NW * kRowU) {
    float v[kRowU];
#pragma unroll
    for (int u = 0; u < kRowU; ++u)
      if (r0 + NW * u < r_end) v[u] = lds_f32(Eb + (unsigned)A.out_meta[r0 + NW * u] * 4u);
#pragma unroll
    for (int u = 0; u < kRowU; ++u)
#ifdef EXP_NO_ROWS
      if (v[u] == 123.456f)
#endif
      if (r0 + NW * u < r_end) *reinterpret_cast<float*>(A.out_ptr[r0 + NW * u] + colb) = v[u];
  }
}

// int64 / bool rows of the OUT table (counters, stage, masks)
__device__ __forceinline__ void post_rows_typed(const PostArgs& A, const float* smem, const int e0, const int ne,
                                                const int nw) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane >= ne) return;
  const unsigned Eb = (unsigned)__cvta_generic_to_shared(smem) + (unsigned)(lane * A.stride) * 4u;
  const size_t col = (size_t)e0 + lane;
  for (int r = A.n_out_f32 + warp; r < A.n_out; r += nw) {
    const float v = lds_f32(Eb + (unsigned)A.out_meta[r] * 4u);
    if (r < A.n_out_i64)
      reinterpret_cast<long long*>(A.out_ptr[r])[col] = (long long)__float_as_int(v);
    else
      reinterpret_cast<unsigned char*>(A.out_ptr[r])[col] = (v != 0.0f) ? 1 : 0;
  }
}

// ------------------------------------------------------------------------------------------------
// phase 4: fused reset of ONE finished env by ONE warp (ResetManager.reset_idx reset_manager.py:92-190,
// BlindGraspingTask.reset_task_state blind_grasping_task.py:433-547, ActionProcessor.reset_targets
// action_processor.py:524-568,616-666, ObservationEncoder.reset_observer_state observation_encoder.py:363-383).
// Same arithmetic as reset_env() (dexstep_common.cuh), lane = element; the warps of the CTA reset different
// envs of the tile concurrently, no CTA barrier inside.  s_u: this warp's 32-float scratch.
// ------------------------------------------------------------------------------------------------
template <bool BG, typename ArgsT>
__device__ __forceinline__ void reset_env_warp(const ArgsT& A, const int env, const uint64_t step, float* s_u) {
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const int lane = threadIdx.x & 31;
  const size_t N = (size_t)A.n;
  const unsigned full = 0xffffffffu;
  if (BG) {
    if (A.uniforms != nullptr) {
      if (lane < DEX_NUM_UNIFORMS) s_u[lane] = A.uniforms[(size_t)env * DEX_NUM_UNIFORMS + lane];
    } else if (lane < 8) {
      uint32_t r[4];
      philox4x32_10((uint32_t)env, (uint32_t)lane, (uint32_t)step, (uint32_t)(step >> 32), (uint32_t)A.seed,
                    (uint32_t)(A.seed >> 32), r);
#pragma unroll
      for (int k = 0; k < 4; ++k) s_u[4 * lane + k] = (float)(r[k] >> 8) * (1.0f / 16777216.0f);
    }
    __syncwarp();
  }
  float p = 0.0f;  // new position of DOF `lane`
  if (lane < DEX_NUM_DOF) {
    if (BG) {  // blind_grasping_task.py:511-547 (all 20 finger DOFs, including r_f_joint3_1)
      const float u = s_u[3 + lane];
      if (lane < 3)
        p = (u * 2.0f - 1.0f) * P.hand_translation_range;
      else if (lane < 6)
        p = (u * 2.0f - 1.0f) * P.hand_rotation_range;
      else if (lane == 6)
        p = u * P.thumb_rotation_range;
      else
        p = u * P.other_finger_range;
    }
    float* dof = A.sim.dof_state + (size_t)env * (DEX_NUM_DOF * 2);
    reinterpret_cast<float2*>(dof)[lane] = make_float2(p, 0.0f);
    st.full_dof_targets[(size_t)env * DEX_NUM_DOF + lane] = p;  // action_processor.py:553 (verbatim, DOF 14 included)
    st.prev_dof_pos[lane * N + env] = 0.0f;                     // observation_encoder.py:375-376 (zero, not the new pose)
  }
  // inverse coupling: the last DOF of a coupling group wins (action_processor.py:616-666)
  const int src = (lane < DEX_NUM_ACTIVE) ? P.inverse_src_dof[lane] : 0;
  const float v = __shfl_sync(full, p, src);
  if (lane < DEX_NUM_ACTIVE) st.active_prev_targets[lane * N + env] = (lane < DEX_NUM_BASE) ? v : v / P.inverse_scale[lane];
  if (lane < DEX_NUM_FINGERS) {
    st.contact_duration_steps[lane * N + env] = 0.0f;
    st.prev_contact_binary[lane * N + env] = 0.0f;
  }
  if (lane < P.num_actions) st.prev_actions[lane * N + env] = 0.0f;
  if (lane == 26) {
    st.episode_step_count[env] = 0;
    if (BG) {  // blind_grasping_task.py:492-509
      st.success_duration_steps[env] = 0;
      st.success_conditions_met[env] = 0;
      st.current_stage[env] = 1;
      st.time_in_stage[env] = 0.0f;
      st.stage_contact_duration[env] = 0.0f;
      st.just_transitioned_to_stage2[env] = 0;
      st.just_transitioned_to_stage3[env] = 0;
    }
  } else if (BG && lane == 27) {  // blind_grasping_task.py:449-490: box pose
    const float bx = (s_u[0] * 2.0f - 1.0f) * P.box_xy_range;
    const float by = (s_u[1] * 2.0f - 1.0f) * P.box_xy_range;
    const float ang = (s_u[2] * 2.0f - 1.0f) * P.pi;
    st.initial_box_positions[0 * N + env] = bx;
    st.initial_box_positions[1 * N + env] = by;
    st.initial_box_positions[2 * N + env] = P.box_z;
    float* root = A.sim.actor_root_state + ((size_t)env * P.num_actors + P.box_actor) * 13;
    root[0] = bx;
    root[1] = by;
    root[2] = P.box_z;
    root[3] = 0.0f;
    root[4] = 0.0f;
    root[5] = sinf(ang / 2.0f);
    root[6] = cosf(ang / 2.0f);
#pragma unroll
    for (int i = 7; i < 13; ++i) root[i] = 0.0f;
  }
  __syncwarp();  // s_u is reused by this warp's next env
}

// ------------------------------------------------------------------------------------------------
// step epilogue body: ordered compaction of the reset mask (== torch.nonzero(reset_buf),
// step_processor.py:109-111), termination rates (termination_manager.py:259-266), consecutive
// successes (:323-339), running totals.  Runs in ONE CTA of T threads: either the stand-alone
// epilogue kernel or the last CTA of the post-physics kernel.
// ------------------------------------------------------------------------------------------------
// T threads take part (tid = 0..T-1 within the group); BAR = 0: the whole CTA (__syncthreads), else the id of a
// named barrier for a T-thread warp group of a larger CTA.
template <int T, int BAR>
__device__ __forceinline__ void epilogue_sync() {
  if (BAR == 0)
    __syncthreads();
  else
    asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(T) : "memory");
}
template <int T, int BAR>
__device__ void epilogue_body(const DexEpilogueBuffers& ep, const int n, const int max_consecutive, int* s_warp,
                              int* s_base, const int tid) {
  const int lane = tid & 31, warp = tid >> 5;
  constexpr int NW = T / 32;
  const int words = (n + 31) >> 5;
  // thread t owns the contiguous word range [t*K, (t+1)*K): one block scan for any n
  const int K = (words + T - 1) / T;
  const int w_begin = min(tid * K, words), w_end = min(w_begin + K, words);
  // all L2 reads of the epilogue are issued up front (one round trip): the counters and, when a thread owns at
  // most kCache words (n <= 32768 with 256 threads), its reset words
  constexpr int kCache = 4;
  const int c = (warp == 0 && lane < DEX_NUM_COUNTERS) ? __ldcg(ep.counters + lane) : 0;
  const bool has_tot = warp == 0 && lane <= DEX_NUM_COUNTERS && ep.totals != nullptr;
  const long long tot0 = has_tot ? __ldcg(reinterpret_cast<const long long*>(ep.totals) + lane) : 0;
  const bool l0 = warp == 0 && lane == 0;
  const int cs0 = l0 ? __ldcg(ep.consecutive_successes) : 0;
  const unsigned long long rs0 = (l0 && ep.rng_step != nullptr) ? __ldcg(reinterpret_cast<const unsigned long long*>(ep.rng_step)) : 0ull;
  unsigned cache[kCache];
#pragma unroll
  for (int i = 0; i < kCache; ++i) cache[i] = (w_begin + i < w_end) ? __ldcg(ep.reset_bits + w_begin + i) : 0u;
  int cnt = 0;
#pragma unroll
  for (int i = 0; i < kCache; ++i) cnt += __popc(cache[i]);
  for (int w = w_begin + kCache; w < w_end; ++w) cnt += __popc(__ldcg(ep.reset_bits + w));
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  epilogue_sync<T, BAR>();
  if (warp == 0) {
    int v = (lane < NW) ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += u;
    }
    if (lane < NW) s_warp[lane] = v;
    if (lane == NW - 1) *s_base = v;
  }
  epilogue_sync<T, BAR>();
  int pos = (warp > 0 ? s_warp[warp - 1] : 0) + incl - cnt;
  for (int w = w_begin; w < w_end; ++w) {
    const int i = w - w_begin;
    unsigned bits = (i < kCache) ? (i == 0 ? cache[0] : i == 1 ? cache[1] : i == 2 ? cache[2] : cache[3])
                                 : __ldcg(ep.reset_bits + w);
    while (bits) {
      const int b = __ffs(bits) - 1;
      ep.env_ids[pos++] = w * 32 + b;
      bits &= bits - 1;
    }
  }
  // counters -> rates (termination_manager.py:259-266); thread k owns counter k
  if (warp == 0) {
    const float fn = (float)n;
    if (lane == DEX_CNT_SUCCESS) ep.stats[DEX_STAT_SUCCESS_RATE] = (float)c / fn;
    if (lane == DEX_CNT_FAILURE) ep.stats[DEX_STAT_FAILURE_RATE] = (float)c / fn;
    if (lane == DEX_CNT_TIMEOUT) ep.stats[DEX_STAT_TIMEOUT_RATE] = (float)c / fn;
    if (lane >= DEX_CNT_CRIT0 && lane < DEX_CNT_CRIT0 + DEX_NUM_CRITERIA)
      ep.stats[DEX_STAT_CRIT0 + (lane - DEX_CNT_CRIT0)] = (float)c / fn;
    if (has_tot) ep.totals[lane] = tot0 + (lane < DEX_NUM_COUNTERS ? c : 1);  // [DEX_NUM_COUNTERS] counts steps
    if (lane < DEX_NUM_COUNTERS) ep.counters[lane] = 0;
    const int n_success = __shfl_sync(0xffffffffu, c, DEX_CNT_SUCCESS);
    if (lane == 0) {
      *ep.num_reset = *s_base;
      const int cs = (n_success > 0) ? cs0 + 1 : 0;  // termination_manager.py:323-339
      *ep.consecutive_successes = min(cs, max_consecutive);
      if (ep.rng_step != nullptr) *ep.rng_step = rs0 + 1;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// fused step epilogue over the tile records (see post_scalar): T threads of the LAST tile's CTA (named barrier
// BAR) wait until every tile has published this step's stamp - tiles with a lower block index were scheduled
// before this CTA, so they are running or done - then do the work of epilogue_body() without any counter array.
// ------------------------------------------------------------------------------------------------
template <int T, int BAR>
__device__ void epilogue_records(const DexEpilogueBuffers& ep, const int n, const int max_consecutive, const unsigned stamp,
                                 int* s_warp, int* s_base, int* s_cnt, unsigned* s_words, const int cap, const int tid) {
  const int lane = tid & 31, warp = tid >> 5;
  constexpr int NW = T / 32;
  const int tiles = (n + 31) >> 5;
  const volatile unsigned long long* rec0 = reinterpret_cast<const volatile unsigned long long*>(ep.tile_records);
  const volatile unsigned long long* rec1 = rec0 + tiles;
  const volatile unsigned long long* rec2 = rec1 + tiles;
  // loads that do not depend on the other tiles: issued first
  const bool has_tot = warp == 0 && lane <= DEX_NUM_COUNTERS && ep.totals != nullptr;
  const long long tot0 = has_tot ? __ldcg(reinterpret_cast<const long long*>(ep.totals) + lane) : 0;
  const bool l0 = tid == 0;
  const int cs0 = l0 ? __ldcg(ep.consecutive_successes) : 0;
  const unsigned long long rs0 = l0 ? __ldcg(reinterpret_cast<const unsigned long long*>(ep.rng_step)) : 0ull;
  if (tid < DEX_NUM_COUNTERS) s_cnt[tid] = 0;
  // ---- pass 1, lane = tile (consecutive lanes read consecutive 8-byte records of a plane: 8 sectors per request
  // instead of 32): wait for the stamps, sum the counters, park the reset words in shared memory (s_words[cap])
  constexpr int kBatch = 2;
  int n_s = 0, n_f = 0, crit[DEX_NUM_CRITERIA];
  long long len = 0;
#pragma unroll
  for (int c = 0; c < DEX_NUM_CRITERIA; ++c) crit[c] = 0;
  // thread t owns the contiguous tile range [t*K, (t+1)*K) in pass 2.  Up to kBatch tiles per thread (<= 32 768 envs)
  // it also reads exactly those in pass 1 - no barrier between the passes; beyond, pass 1 is lane-contiguous.
  const int K = (tiles + T - 1) / T;
  const int w_begin = min(tid * K, tiles), w_end = min(w_begin + K, tiles);
  const bool direct = K <= kBatch;  // CTA-uniform
  for (int i0 = tid; i0 < (direct ? T : tiles); i0 += T * kBatch) {  // kBatch tiles per round trip (volatile: L2 every time)
    unsigned long long r0[kBatch], r1[kBatch], r2[kBatch];
    int idx[kBatch];
#pragma unroll
    for (int j = 0; j < kBatch; ++j) {
      idx[j] = direct ? w_begin + j : i0 + j * T;
      if (direct && j >= K) idx[j] = tiles;  // nothing to do
    }
    bool ok;
    unsigned polls = 0;
    do {
#pragma unroll
      for (int j = 0; j < kBatch; ++j) {
        const int i = min(idx[j], tiles - 1);
        r0[j] = rec0[i];
        r1[j] = rec1[i];
        r2[j] = rec2[i];
      }
      ok = true;
#pragma unroll
      for (int j = 0; j < kBatch; ++j)
        ok = ok && (unsigned)(r0[j] >> 48) == stamp && (unsigned)(r1[j] >> 48) == stamp && (unsigned)(r2[j] >> 48) == stamp;
      // bounded (each poll is an L2 round trip, ~0.5 us): a tile that never publishes - a fault in its CTA, or a record
      // buffer shared with another pipeline - ends the launch with a trap after a few seconds instead of hanging the stream
      if (!ok && ++polls > (1u << 23)) __trap();
    } while (!ok);
#pragma unroll
    for (int j = 0; j < kBatch; ++j) {
      const int i = idx[j];
      if (i < tiles) {
        if (i < cap) s_words[i] = (unsigned)r0[j];
        n_s += (int)((r0[j] >> 32) & 63u);
        n_f += (int)((r0[j] >> 38) & 63u);
#pragma unroll
        for (int c = 0; c < DEX_NUM_CRITERIA; ++c) crit[c] += (int)((r1[j] >> (6 * c)) & 63u);
        len += (long long)(r2[j] & 0xffffffffffffull);
      }
    }
  }
  if (!direct) epilogue_sync<T, BAR>();  // s_words complete (direct: every thread reads back its own words only)
  auto word = [&](const int w) { return w < cap ? s_words[w] : (unsigned)rec0[w]; };
  int cnt = 0;
  for (int w = w_begin; w < w_end; ++w) cnt += __popc(word(w));
  // block scan of the reset counts
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) s_warp[warp] = incl;
  // counter totals: warp reduce, then one shared atomic per warp and counter
  int red[4 + DEX_NUM_CRITERIA] = {n_s, n_f, cnt, (int)len};
#pragma unroll
  for (int c = 0; c < DEX_NUM_CRITERIA; ++c) red[4 + c] = crit[c];
#pragma unroll
  for (int k = 0; k < 4 + DEX_NUM_CRITERIA; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) red[k] += __shfl_xor_sync(0xffffffffu, red[k], o);
  }
  epilogue_sync<T, BAR>();  // s_cnt zeroed, s_warp written
  if (lane == 0) {
    atomicAdd(s_cnt + DEX_CNT_SUCCESS, red[0]);
    atomicAdd(s_cnt + DEX_CNT_FAILURE, red[1]);
    atomicAdd(s_cnt + DEX_CNT_RESET, red[2]);
    atomicAdd(s_cnt + DEX_CNT_EPISODE_LENGTH, red[3]);
#pragma unroll
    for (int c = 0; c < DEX_NUM_CRITERIA; ++c) atomicAdd(s_cnt + DEX_CNT_CRIT0 + c, red[4 + c]);
  }
  if (warp == 0) {
    int v = (lane < NW) ? s_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += u;
    }
    if (lane < NW) s_warp[lane] = v;
    if (lane == NW - 1) *s_base = v;
  }
  epilogue_sync<T, BAR>();
  int pos = (warp > 0 ? s_warp[warp - 1] : 0) + incl - cnt;
  for (int w = w_begin; w < w_end; ++w) {
    unsigned bits = word(w);
    while (bits) {
      const int b = __ffs(bits) - 1;
      ep.env_ids[pos++] = w * 32 + b;
      bits &= bits - 1;
    }
  }
  if (warp == 0) {  // counters -> rates (termination_manager.py:259-266); lane k owns counter k
    int c = (lane < DEX_NUM_COUNTERS) ? s_cnt[lane] : 0;
    const int n_reset = __shfl_sync(0xffffffffu, c, DEX_CNT_RESET);
    const int n_success = __shfl_sync(0xffffffffu, c, DEX_CNT_SUCCESS);
    const int n_failure = __shfl_sync(0xffffffffu, c, DEX_CNT_FAILURE);
    if (lane == DEX_CNT_TIMEOUT) c = n_reset - n_success - n_failure;  // the three classes partition the resets
    const float fn = (float)n;
    if (lane == DEX_CNT_SUCCESS) ep.stats[DEX_STAT_SUCCESS_RATE] = (float)c / fn;
    if (lane == DEX_CNT_FAILURE) ep.stats[DEX_STAT_FAILURE_RATE] = (float)c / fn;
    if (lane == DEX_CNT_TIMEOUT) ep.stats[DEX_STAT_TIMEOUT_RATE] = (float)c / fn;
    if (lane >= DEX_CNT_CRIT0 && lane < DEX_CNT_CRIT0 + DEX_NUM_CRITERIA)
      ep.stats[DEX_STAT_CRIT0 + (lane - DEX_CNT_CRIT0)] = (float)c / fn;
    if (has_tot) ep.totals[lane] = tot0 + (lane < DEX_NUM_COUNTERS ? c : 1);  // [DEX_NUM_COUNTERS] counts steps
    if (lane == 0) {
      *ep.num_reset = *s_base;
      const int cs = (n_success > 0) ? cs0 + 1 : 0;  // termination_manager.py:323-339
      *ep.consecutive_successes = min(cs, max_consecutive);
      *ep.rng_step = rs0 + 1;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------------
template <bool BG, bool DICT, bool NORM>
__global__ void __launch_bounds__(kPostThreads, DICT ? 1 : 4) post_physics_kernel(const __grid_constant__ PostArgs A) {
  extern __shared__ __align__(16) float smem[];
  __shared__ short s_obs_map[kMaxObs];
  __shared__ float s_norm[NORM ? 2 * kMaxObs : 2];  // mean | denominator of the policy-input stage
  __shared__ short s_dict_map[DICT ? kMaxDict : 1];
  __shared__ int s_epi_warp[kPostWarps];
  __shared__ int s_epi_base;
  __shared__ int s_epi_cnt[DEX_NUM_COUNTERS];
  __shared__ float s_reset_u[32 * kPostWarps];
  __shared__ unsigned long long s_rng_step;
  __shared__ unsigned s_rbits;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int e0 = blockIdx.x * kTile;
  const int ne = min(kTile, A.n - e0);

  TL_MARK(0);
  // Fused step epilogue: the launch carries ONE extra CTA behind the tiles (launch_post_t).  It owns no tile: it polls
  // the tile records - every tile publishes its own at the end of its scalar phase - and finishes the step
  // (env_ids, rates, running totals, stamp counter) while the tiles are still staging out.  Tiles have lower block
  // indices, so they were scheduled before this CTA: the wait cannot deadlock.
  if (ne <= 0) {
    pdl_wait();
    pdl_launch_dependents();
    if (tid == 0) {
      // one thread waits for the record of the LAST tile (scheduled last, so among the last to publish) before the
      // CTA starts the full poll: 256 threads polling from the start of the launch would load the LSU of an SM
      // that also runs tiles
      const unsigned long long rs = *A.st.rng_step;
      const unsigned stamp = ((unsigned)rs + 1u) & 0xffffu;
      const volatile unsigned long long* last =
          reinterpret_cast<const volatile unsigned long long*>(A.ep.tile_records) + (size_t)(blockIdx.x - 1);  // plane 0
      unsigned polls = 0;
      while ((unsigned)(last[0] >> 48) != stamp) {
        if (++polls > (1u << 23)) __trap();  // bounded wait, see epilogue_records
      }
      s_rng_step = rs;
    }
    __syncthreads();
    epilogue_records<kPostThreads, 0>(A.ep, A.n, A.p.max_consecutive_successes, ((unsigned)s_rng_step + 1u) & 0xffffu,
                                      s_epi_warp, &s_epi_base, s_epi_cnt, reinterpret_cast<unsigned*>(smem),
                                      A.stride * kTile, tid);  // the slab of this CTA is free: it holds the reset words
    TL_MARK(5);
    TL_MARK(6);
    return;
  }
  s_obs_map[tid] = A.obs_src[tid];  // kMaxObs == kPostThreads
  float nm = 0.0f, ns = 1.0f;  // NORM: issued here, parked in shared memory after the stage-in loads are in flight
  if (NORM && tid < A.p.obs_dim) {  // written by the caller's copies, which precede dex_pre_physics in the stream
    nm = __ldg(A.out.norm_mean + tid);
    ns = __ldg(A.out.norm_scale + tid);
  }
  if (DICT) {
    s_dict_map[tid] = A.dict_src[tid];
    s_dict_map[tid + kPostThreads] = A.dict_src[tid + kPostThreads];
  }
  const bool full_tile = ne == kTile;  // CTA-uniform
  const NormOut norm{A.out.obs_norm, s_norm, s_norm + (NORM ? kMaxObs : 1), A.out.norm_clip};  // NORM instances only
  if (full_tile)
    post_stage_in<BG, true>(A, smem, e0, ne);
  else
    post_stage_in<BG, false>(A, smem, e0, ne);
  if (NORM) {
    s_norm[tid] = nm;
    s_norm[kMaxObs + tid] = ns;
  }
  // Philox step counter (also the stamp of this step's tile records): read before this CTA publishes its record
  // (scalar phase), i.e. before the fused epilogue can advance it - by ONE thread and after the stage-in loads
  // are in flight (it is an HBM miss of its own)
  if (tid == 64) s_rng_step = A.st.rng_step ? *A.st.rng_step : A.step;
  __syncthreads();
  pdl_launch_dependents();
  TL_MARK(1);
  post_items<BG>(A, smem, ne);
  __syncthreads();
  TL_MARK(2);
  bool sr = false;
  if (warp == 0) {
    sr = post_scalar<BG>(A, smem, e0, ne, ((unsigned)s_rng_step + 1u) & 0xffffu, &s_rbits);
  } else {  // meanwhile everything that is final after the items phase leaves through warps 1..7
    constexpr int NW = kPostWarps - 1;
    TL_MARK_W1(11);
    post_rows_f32<NW>(A, smem, e0, ne, 0, A.n_out_early, warp - 1);
    TL_MARK_W1(12);
    gather_rows_out<NW, NORM>(A.out.obs_buf, A.p.obs_dim, s_obs_map, smem, A.stride, e0, ne, ~A.obs_late, warp - 1, norm);
    TL_MARK_W1(13);
    if (DICT) gather_rows_out<NW>(A.out.obs_dict, A.dict_dim, s_dict_map, smem, A.stride, e0, ne, ~A.dict_late, warp - 1);
  }
  const int any_reset = __syncthreads_or(sr ? 1 : 0);
  TL_MARK(3);
  const bool full_mode = A.mode == DEX_MODE_FULL;
  {
    if (full_mode) post_reward_terms<BG>(A, smem, ne, kPostWarps);
    gather_rows_out<kPostWarps, NORM>(A.out.obs_buf, A.p.obs_dim, s_obs_map, smem, A.stride, e0, ne, A.obs_late, warp, norm);
    if (DICT) gather_rows_out<kPostWarps>(A.out.obs_dict, A.dict_dim, s_dict_map, smem, A.stride, e0, ne, A.dict_late, warp);
    post_rows_typed(A, smem, e0, ne, kPostWarps);
    if (full_mode) {
      __syncthreads();  // the reward-term rows of the slab are complete
      if (warp == 0) post_total<BG>(A, smem, e0, ne);
    }
    post_rows_f32<kPostWarps>(A, smem, e0, ne, A.n_out_early, A.n_out_f32, warp);
  }
  TL_MARK(4);
#ifdef EXP_NO_RESET
  if (false) {
#else
  if (any_reset && !(A.flags & DEX_FLAG_NO_FUSED_RESET)) {  // fused reset (dexhand_base.py:743-756), CTA-uniform branch
#endif
    __syncthreads();  // the OUT-table rows of the tile are written before the reset overwrites some of them
    unsigned mask = s_rbits;  // written by warp 0 before the vote barrier
    const uint64_t rng_step = s_rng_step;
    // warp w resets the (w)th, (w+8)th, ... finished env of the tile
    for (int k = 0; mask; ++k) {
      const int e = __ffs(mask) - 1;
      mask &= mask - 1;
      if ((k & (kPostWarps - 1)) == warp) reset_env_warp<BG>(A, e0 + e, rng_step, s_reset_u + 32 * warp);
    }
  }
  TL_MARK(5);
  TL_MARK(6);
}

__global__ void __launch_bounds__(1024) epilogue_kernel(const DexEpilogueBuffers ep, const int n, const int max_consecutive) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  epilogue_body<1024, 0>(ep, n, max_consecutive, s_warp, &s_base, threadIdx.x);
}

template <bool BG, bool DICT, bool NORM>
int launch_post_t(const PostArgs& a, cudaStream_t s) {
  const size_t smem = (size_t)a.stride * kTile * sizeof(float);
  if (smem > 200 * 1024) return DEX_ERR_SIZE;
  static bool configured = false;  // per template instance
  if (!configured) {
    const cudaError_t e = cudaFuncSetAttribute(post_physics_kernel<BG, DICT, NORM>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               200 * 1024);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int grid = (a.n + kTile - 1) / kTile + ((a.fused_epilogue && a.mode == DEX_MODE_FULL) ? 1 : 0);  // + epilogue CTA
  const cudaError_t e = launch_step_kernel(post_physics_kernel<BG, DICT, NORM>, grid, kPostThreads, smem, s, a);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

inline int launch_post(PostArgs& a, cudaStream_t s) {
  const int rc = build_post_tables(a);
  if (rc) return rc;
  const bool bg = a.p.task == DEX_TASK_BLIND_GRASPING, dict = a.out.obs_dict != nullptr;
  const bool norm = a.out.obs_norm != nullptr;
  if (norm) {
    if (bg) return dict ? launch_post_t<true, true, true>(a, s) : launch_post_t<true, false, true>(a, s);
    return dict ? launch_post_t<false, true, true>(a, s) : launch_post_t<false, false, true>(a, s);
  }
  if (bg) return dict ? launch_post_t<true, true, false>(a, s) : launch_post_t<true, false, false>(a, s);
  return dict ? launch_post_t<false, true, false>(a, s) : launch_post_t<false, false, false>(a, s);
}

}  // namespace
