// post_physics_v2.cuh — the post-physics kernels the lean production path runs (round 2).
//
// Same arithmetic as post_physics.cuh (StepProcessor.process_physics_step, step_processor.py:37-131), op for op (that
// file is pinned to the oracle and kept for the obs_dict / OBS_ONLY / stale-contact instances); different machine
// mapping.  Two geometries (launch_post2_t):
//
//   post_physics_v2_kernel  PERSISTENT, warp-specialised, the default: <= SMs x CTAs/SM worker CTAs (+ one epilogue CTA) loop
//     over tiles; 4 producer warps stream tile t+1 into a 2-stage shared-memory ring (mbarrier full/empty pairs) while 8
//     consumer warps compute tile t.  All CTAs are co-resident, so the epilogue CTA's poll of the tile records needs no
//     scheduling-order assumption.
//   post_physics_v2s_kernel SINGLE PASS, on request (DEX_FLAG_SINGLE_PASS) for 2 < tiles/SM <= 4 (16 384 envs on 148 SMs =
//     3.46 tiles per SM): one tile per CTA, 8 warps, 4 CTAs per SM, loads staged through registers, the work block aliased
//     into the stage (Lay<BG, true>) so that four CTAs fit.  3-5 % slower than the persistent kernel when the inputs come
//     from HBM, 20 % faster when the sim tensors are still in L2.
//   Template flags: STD = compile-time standard observation layout, TMA = tensor-copy producers (DEX_FLAG_TMA_STAGING, measured
//     slower), OBS = fused RewardComponentObserver (a separate instance: even dormant observer code costs the lean path 1-4 %),
//     G = tiles per pass and CTA (2 = the pair geometry, DEX_FLAG_PAIR_GEOMETRY: two groups of 8 consumer warps in lockstep on
//     two tiles, one 24-warp CTA per SM; measured 3-6 % slower, kept selectable).
//
// What shaped them (profiles/r02_variants.md has the measurements of the variants that did NOT work):
//   * The path is bound by INSTRUCTION SUPPLY for cold code, not by bytes or issue slots: every CTA runs ~10 K distinct
//     SASS instructions once per tile.  Cold straight-line code costs ~3 cycles/instruction/SM, code that is resident in
//     the 32 KB L1.5 ~1.  Hence: all consumer warps run the SAME item code at the same time (round-robin work items,
//     phase_items), loops are not unrolled, compile-time observation layouts (STD instances, std_obs_off) remove the
//     table walks, and the scalar phase is split over 4 warps only where the code is short.
//   * Producer warps gather the scattered Isaac Gym rows and the SoA state rows with 4/8-byte cp.async (LDGSTS, completion
//     counted on the stage's mbarrier with cp.async.mbarrier.arrive.noinc); cp.async.bulk (UBLKCP) is used only for the
//     two tile-contiguous arrays (dof_state 32 x 208 B, full_dof_targets) - a bulk copy costs ~50 issue cycles, one per
//     128-byte row starved the SM sub-partition the producers share with consumers.
//   * Waiting warps block in hardware: ONE consumer warp polls the stage mbarrier (try_wait.parity with a suspend hint),
//     the other seven wait behind a named barrier (bar.sync 1, 256).  All polls are bounded (kSpinLimitClocks -> trap).
//   * NO TABLES on the way out: every state / reward row leaves straight from the registers of the lane (= env) that
//     computed it (one 128 B request per row); observation values are written into an obs-shaped shared tile by the
//     work item that produces them and the tile leaves as one bulk store (flat stride) or coalesced row copies.
//   * Fused extras (all optional, all in the same pass): reset of finished envs, step epilogue (tile records + one
//     polling CTA), RewardComponentObserver sums, rl_games input normalisation of obs_buf (obs_norm).
// Numerics: lane = env for all arithmetic; ordered reductions stay single-lane per env (parity with torch's order).
#pragma once
#include <string.h>

#include <cuda.h>  // CUtensorMap (types only: the encoder is fetched with cudaGetDriverEntryPoint, libcuda is not linked)

#include "post_physics.cuh"

namespace {
namespace pp2 {

constexpr int kConsWarps = 8, kProdWarps = 4;
constexpr int kConsThreads = kConsWarps * 32, kProdThreads = kProdWarps * 32;
constexpr int kThreads = kConsThreads + kProdThreads;
constexpr int kStages = 2;
constexpr int kMaxBulk = 128;  // rows / tile arrays of one tile (the plain-load fallback walks this table)

enum : int { BK_F32 = 0, BK_I64 = 1, BK_U8 = 2, BK_DOF = 3, BK_TGT = 4 };

// ---- shared-memory layout of one stage (bytes) and of the work block (fields of 32 floats) ----
// AL: the single-pass geometry (one tile per CTA, four CTAs per SM) folds the work block into the stage: fields that are
// written after phase A alias state rows that are dead by then, the phase-A fields sit right behind the stage
template <bool BG, bool AL = false>
struct Lay {
  // SoA state rows, [row][32] floats
  static constexpr int R_PDP = 0, R_APT = 26, R_PACT = 44, R_CDUR = 62, R_PBIN = 67;
  static constexpr int R_BOX0 = 72, R_TIS = 75, R_SCD = 76;  // BlindGrasping
  static constexpr int R_PFDV = 72, R_PHV = 92, R_PHA = 95;  // BaseTask: common-reward previous state
  static constexpr int NF32 = BG ? 77 : 98;
  static constexpr int I_STEP = 0, I_SDS = 1, I_STAGE = 2;   // int64 rows, [row][32] x 8 B
  static constexpr int NI64 = BG ? 3 : 1;
  static constexpr int NU8 = BG ? 0 : 5;                     // prev_contacts, [row][32] x 1 B
  static constexpr int OFF_I64 = NF32 * 128;
  static constexpr int OFF_U8 = OFF_I64 + NI64 * 256;
  static constexpr int OFF_DOF = OFF_U8 + NU8 * 32;          // dof_state tile, [32][52] floats as in HBM
  static constexpr int OFF_TGT = OFF_DOF + 32 * 52 * 4;      // full_dof_targets tile, [32][26] (BaseTask)
  static constexpr int OFF_AOS = OFF_TGT + (BG ? 0 : 32 * 26 * 4);
  // gathered Isaac rows, [32][SA] floats, SA odd (lane = env reads are conflict-free)
  static constexpr int A_RB = 0, A_CF = 83, A_BOX = 101;     // hand row 13 | 5 tips x 7 | 5 pads x 7 ; contacts ; box
  static constexpr int SA = BG ? 107 : 99;
  static constexpr int STAGE_BYTES = (OFF_AOS + 32 * SA * 4 + 127) / 128 * 128;
  // work block: values that cross a warp or a phase, [field][32] floats (lane = env: conflict-free), indexed from WK0:
  // !AL: a block of its own (WK0 = 0).  AL: indexed from the stage base in 128-byte rows - the phase-A fields behind the stage
  // (row STAGE_BYTES / 128 on), the phase-B fields over the active_prev_targets rows and the weighted terms over the
  // prev_dof_pos rows (both only read in phase A)
  static constexpr int WKA = AL ? STAGE_BYTES / 128 : 0;        // phase-A fields
  static constexpr int W_BIN = WKA + 0, W_PADBOX = WKA + 5, W_HDIFF = WKA + 10, W_TIPBOX = WKA + 15;
  static constexpr int W_CDIST = WKA + 20, W_DRIFT = WKA + 21, W_VMAG = WKA + 22, W_BOXHAS = WKA + 23, W_HANDBOX = WKA + 24;
  static constexpr int W_BVEL = WKA + 25;  // 6
  static constexpr int NWKA = 31;
  static constexpr int WKB = AL ? R_APT : 31;                   // phase-B fields (13)
  static constexpr int W_GRASP = WKB + 0, W_THUMB = WKB + 1, W_OTHERS = WKB + 2, W_GRASPDUR = WKB + 3, W_SDS = WKB + 4,
                       W_STAGE = WKB + 5, W_J2 = WKB + 6, W_J3 = WKB + 7;
  static constexpr int W_HITG = WKB + 8, W_BOXFAR = WKB + 9, W_QUALOK = WKB + 10, W_MINTIPBOX = WKB + 11, W_AVGHDIFF = WKB + 12;
  static constexpr int W_CHG = 44, W_RHAND = 49, W_FV2 = 53, W_FA2 = 73, W_JL = 93;  // BaseTask (common rewards; !AL only)
  static constexpr int W_WGT = AL ? R_PDP : (BG ? 44 : 113);    // 26 weighted reward terms
  static constexpr int W_COUNT = AL ? NWKA : W_WGT + DEX_NUM_REWARD_TERMS;  // rows of the separate block
};


// ------------------------------------------------------------------------------------------------
// The reference's two shipped observation layouts (cfg/task/BaseTask.yaml:103-116, BlindGrasping.yaml:50-66; SURVEY
// App. B) as compile-time constants.  A launch whose parameter block matches them (std_layout_matches) runs the STD
// instance of the kernel: every obs column offset is an immediate, no key test and no zero-weight test survives, so
// the per-env program is (almost) free of taken branches - with one compute warp per scheduler a taken branch costs a
// front-end refetch that nothing hides (profiles/r02c_*).  Any other layout runs the generic instance.
// ------------------------------------------------------------------------------------------------
template <bool BG>
__host__ __device__ constexpr int std_obs_off(const int k) {
  if (BG) {
    switch (k) {
      case DEX_K_BASE_DOF_POS: return 0;
      case DEX_K_ACTIVE_FINGER_DOF_POS: return 6;
      case DEX_K_BASE_DOF_VEL: return 18;
      case DEX_K_ACTIVE_FINGER_DOF_VEL: return 24;
      case DEX_K_CONTACT_BINARY: return 36;
      case DEX_K_CONTACT_DURATION: return 41;
      case DEX_K_HAND_POSE: return 46;
      case DEX_K_PREV_ACTIONS: return 53;
      case DEX_K_FINGERTIP_POSES_WORLD: return 71;
      case DEX_K_FINGERPAD_POSES_WORLD: return 106;
      case DEX_K_FIRST_THREE_FINGERPAD_CENTROID: return 141;
      case DEX_K_EPISODE_TIME: return 144;
      case DEX_K_FINGERPAD_DISTANCES: return 145;
      case DEX_K_CURRENT_STAGE: return 155;
      case DEX_K_TIME_IN_STAGE: return 156;
      case DEX_K_STAGE_PROGRESS: return 157;
      default: return -1;
    }
  } else {
    switch (k) {
      case DEX_K_BASE_DOF_POS: return 0;
      case DEX_K_BASE_DOF_VEL: return 6;
      case DEX_K_ACTIVE_FINGER_DOF_POS: return 12;
      case DEX_K_ACTIVE_FINGER_DOF_VEL: return 24;
      case DEX_K_HAND_POSE: return 36;
      case DEX_K_CONTACT_BINARY: return 43;
      case DEX_K_PREV_ACTIONS: return 48;
      case DEX_K_BASE_DOF_TARGET: return 66;
      case DEX_K_ACTIVE_FINGER_DOF_TARGET: return 72;
      case DEX_K_FINGERTIP_POSES_WORLD: return 84;
      case DEX_K_FINGERTIP_POSES_HAND: return 119;
      case DEX_K_FINGERPAD_POSES_WORLD: return 154;
      case DEX_K_FINGERPAD_POSES_HAND: return 189;
      default: return -1;
    }
  }
}
constexpr int kStdObsDim[2] = {224, 158};
// HardwareMapping order of the 12 active finger DOFs (hand_initializer.py:20-34): dof -> column, -1 = not active
__host__ __device__ constexpr int std_dof2afj(const int d) {
  switch (d) {
    case 8: return 0; case 7: return 1; case 6: return 2; case 10: return 3; case 12: return 4; case 11: return 5;
    case 16: return 6; case 15: return 7; case 20: return 8; case 19: return 9; case 24: return 10; case 23: return 11;
    default: return -1;
  }
}

struct Post2Args {
  DexParams p;
  DexSimViews sim;
  DexStateBuffers st;
  DexOutBuffers out;
  DexEpilogueBuffers ep;
  int32_t n, mode;
  uint32_t flags;
  int32_t fused_epilogue;
  const float* uniforms;
  uint64_t seed, step;
  // ---- derived on the host (build_post2) ----
  int32_t tiles, workers;
  int32_t common, first, want_tgt;
  int32_t bulk_ok, n_bulk, n_bulk_early;  // entries [0, n_bulk_early) touch sim tensors only
  int32_t n_tile_entries;                 // entries [0, n_tile_entries) are whole AoS tiles (bulk copies), the rest SoA rows
  uint32_t tx_bytes;                      // bytes of the bulk-copied AoS tiles of a full tile
  int32_t wp;                             // obs tile row stride in floats
  int32_t obs_bulk;                       // 1: the obs tile of a full tile is flat ([32][obs_dim]) and leaves with ONE bulk store
  uint32_t obs_late;                      // bit c: obs columns [32c, 32c+32) hold a value that phase B writes
  uint32_t smem_bytes;
  int32_t rb_off[11], cf_off[6];
  int32_t rb_pitch;                       // floats of one env in rigid_body_state (NB * DexParams.rigid_row_floats)
  int32_t dof2afj[DEX_NUM_DOF];           // dof -> index in active_finger_dof, -1 = not an active finger DOF
  // ---- TMA staging of the persistent kernel (build_tma): one copy per lane of producer warp 0 and tile ----
  int32_t tma;                            // 1: tensor copies instead of LDGSTS gathers
  uint32_t stage_bytes;                   // bytes of one ring stage
  uint32_t tma_tx;                        // bytes that land per tile (expect_tx)
  int32_t t_ncopy;                        // copies per tile (<= 32)
  int32_t t_root_g, t_root_sh;            // actor_root_state is viewed as [N / g][g * NA * 13] (16-byte pitch); sh = log2(g)
  uint16_t t_hand, t_pose[10], t_cf[6], t_box[2];  // consumer view: float offset of element 0 of lane 0 inside the raw block
  uint8_t t_pose_w[10], t_cf_w[6];                 //                box width in floats (lane stride)
  uint8_t t_kind[32];                     // 0: sim box {c0, e0}   1: root box {c0, e0 / g}   2: state box {e0, 0}   3: plain bulk copy
  uint8_t t_early[32];                    // 1: reads a sim tensor only (may be issued ahead of griddepcontrol.wait)
  uint8_t t_map[32];                      // index into tmap
  uint32_t t_dst[32];                     // byte offset inside the stage
  int32_t t_c0[32];                       // inner coordinate (kinds 0, 1)
  uint32_t t_bytes[32];                   // kind 3: bytes per tile (== source stride per tile)
  uint64_t t_ptr[32];                     // kind 3: global base
  alignas(64) CUtensorMap tmap[14];
  uint64_t bk_ptr[kMaxBulk];              // global base of the row / tile array
  uint16_t bk_dst16[kMaxBulk];            // byte offset inside the stage / 16
  uint16_t bk_len16[kMaxBulk];            // bytes per full tile / 16 (== source stride per tile)
  uint8_t bk_kind[kMaxBulk];
};

// ------------------------------------------------------------------------------------------------
// host: eligibility + derived arguments
// ------------------------------------------------------------------------------------------------
inline bool post2_eligible(const DexParams& P, const DexOutBuffers& out, const int mode, const uint32_t flags) {
  if (mode != DEX_MODE_FULL || out.obs_dict != nullptr) return false;
  if (flags & DEX_FLAG_STALE_CONTACTS) return false;
  if (P.obs_dim > 256) return false;
  const bool bg = P.task == DEX_TASK_BLIND_GRASPING;
  auto has = [&](int k) { return P.obs_offset[k] >= 0; };
  const bool want_tgt = has(DEX_K_BASE_DOF_TARGET) || has(DEX_K_ACTIVE_FINGER_DOF_TARGET) || has(DEX_K_ALL_FINGER_DOF_TARGET);
  if (bg && (P.need_common_rewards || want_tgt)) return false;  // the BlindGrasping stage layout carries neither
  if (!bg)
    for (int k = DEX_K_OBJECT_POS; k < DEX_NUM_OBS_KEYS; ++k)
      if (has(k)) return false;
  return true;
}

template <bool BG>
inline int build_post2(Post2Args& a) {
  using LY = Lay<BG>;
  const DexParams& P = a.p;
  const DexStateBuffers& st = a.st;
  const size_t N = (size_t)a.n;
  auto has = [&](int k) { return P.obs_offset[k] >= 0; };
  a.common = P.need_common_rewards ? 1 : 0;
  a.first = (a.flags & DEX_FLAG_FIRST_REWARD_CALL) ? 1 : 0;
  a.want_tgt = (has(DEX_K_BASE_DOF_TARGET) || has(DEX_K_ACTIVE_FINGER_DOF_TARGET) || has(DEX_K_ALL_FINGER_DOF_TARGET)) ? 1 : 0;
  a.tiles = (a.n + 31) / 32;
  // obs tile: flat rows (stride obs_dim) whenever lane = env stores stay <= 4-way bank conflicted, so that a full tile is
  // one contiguous 16-byte-aligned block and leaves with a single cp.async.bulk store; else odd stride + row copies
  a.obs_bulk = (P.obs_dim % 8 != 0 && ((uintptr_t)a.out.obs_buf & 15u) == 0) ? 1 : 0;
  a.wp = a.obs_bulk ? P.obs_dim : (P.obs_dim | 1);
  a.obs_late = 0;
  {
    const int late_keys[10] = {DEX_K_THUMB_CONTACT, DEX_K_OTHER_FINGERS_CONTACT, DEX_K_GRASP_STATE, DEX_K_GRASP_DURATION,
                               DEX_K_CURRENT_STAGE, DEX_K_TIME_IN_STAGE, DEX_K_STAGE_PROGRESS, DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE,
                               DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF, DEX_K_EPISODE_TIME};
    for (int i = 0; i < 10; ++i)
      if (BG && has(late_keys[i])) a.obs_late |= 1u << (P.obs_offset[late_keys[i]] >> 5);
  }
  const int rw = P.rigid_row_floats > 0 ? P.rigid_row_floats : 13;  // 13 = Isaac Gym rows, 7..12 = packed host staging
  a.rb_pitch = P.num_bodies * rw;
  a.rb_off[0] = P.hand_body * rw;
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
    a.rb_off[1 + f] = P.tip_body[f] * rw;
    a.rb_off[6 + f] = P.pad_body[f] * rw;
    a.cf_off[f] = P.contact_body[f] * 3;
  }
  a.cf_off[5] = BG ? P.box_body * 3 : 0;
  for (int d = 0; d < DEX_NUM_DOF; ++d) a.dof2afj[d] = -1;
  for (int j = 0; j < DEX_NUM_ACTIVE_FINGER; ++j) a.dof2afj[P.active_finger_dof[j]] = j;

  int k = 0;
  bool aligned = true;  // the AoS tiles: base 16-byte aligned (tile offsets are multiples of 16 for any N)
  uint32_t tx = 0;
  auto add = [&](const void* base, size_t row, int elem, int dst, int kind) {
    if (k >= kMaxBulk) return;
    const uintptr_t ptr = (uintptr_t)((const char*)base + row * N * elem);
    const int len = kind == BK_DOF ? 32 * 52 * 4 : kind == BK_TGT ? 32 * 26 * 4 : 32 * elem;
    if ((kind == BK_DOF || kind == BK_TGT) && (ptr & 15u)) aligned = false;
    a.bk_ptr[k] = (uint64_t)ptr;
    a.bk_dst16[k] = (uint16_t)(dst / 16);
    a.bk_len16[k] = (uint16_t)(len / 16);
    a.bk_kind[k] = (uint8_t)kind;
    if (kind == BK_DOF || kind == BK_TGT) tx += (uint32_t)len;
    ++k;
  };
  add(a.sim.dof_state, 0, 4, LY::OFF_DOF, BK_DOF);
  a.n_bulk_early = k;
  if (!BG && a.want_tgt) add(st.full_dof_targets, 0, 4, LY::OFF_TGT, BK_TGT);
  a.n_tile_entries = k;
  if (BG) {
  } else if (a.common && !a.first) {
  }
  for (int d = 0; d < DEX_NUM_DOF; ++d) add(st.prev_dof_pos, d, 4, (LY::R_PDP + d) * 128, BK_F32);
  for (int i = 0; i < DEX_NUM_ACTIVE; ++i) add(st.active_prev_targets, i, 4, (LY::R_APT + i) * 128, BK_F32);
  for (int i = 0; i < P.num_actions; ++i) add(st.prev_actions, i, 4, (LY::R_PACT + i) * 128, BK_F32);
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) add(st.contact_duration_steps, f, 4, (LY::R_CDUR + f) * 128, BK_F32);
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) add(st.prev_contact_binary, f, 4, (LY::R_PBIN + f) * 128, BK_F32);
  add(st.episode_step_count, 0, 8, LY::OFF_I64 + LY::I_STEP * 256, BK_I64);
  if (BG) {
    for (int c = 0; c < 3; ++c) add(st.initial_box_positions, c, 4, (LY::R_BOX0 + c) * 128, BK_F32);
    add(st.time_in_stage, 0, 4, LY::R_TIS * 128, BK_F32);
    add(st.stage_contact_duration, 0, 4, LY::R_SCD * 128, BK_F32);
    add(st.success_duration_steps, 0, 8, LY::OFF_I64 + LY::I_SDS * 256, BK_I64);
    add(st.current_stage, 0, 8, LY::OFF_I64 + LY::I_STAGE * 256, BK_I64);
  } else if (a.common && !a.first) {
    for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) add(st.prev_finger_dof_vel, j, 4, (LY::R_PFDV + j) * 128, BK_F32);
    for (int c = 0; c < 3; ++c) add(st.prev_hand_vel, c, 4, (LY::R_PHV + c) * 128, BK_F32);
    for (int c = 0; c < 3; ++c) add(st.prev_hand_ang_vel, c, 4, (LY::R_PHA + c) * 128, BK_F32);
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) add(st.prev_contacts, f, 1, LY::OFF_U8 + f * 32, BK_U8);
  }
  if (k >= kMaxBulk) return DEX_ERR_SIZE;
  a.n_bulk = k;
  a.tx_bytes = tx;
  a.bulk_ok = aligned ? 1 : 0;
  a.smem_bytes = (uint32_t)(kStages * LY::STAGE_BYTES + LY::W_COUNT * 128 + (32 * a.wp * 4 + 127) / 128 * 128);
  return DEX_OK;
}


// does the launch match the compile-time layout of the STD kernel instance?
template <bool BG>
inline bool std_layout_matches(const Post2Args& a) {
  const DexParams& P = a.p;
  if (P.obs_dim != kStdObsDim[BG ? 1 : 0] || P.num_actions != DEX_NUM_ACTIVE) return false;
  for (int k = 0; k < DEX_NUM_OBS_KEYS; ++k)
    if (P.obs_offset[k] != std_obs_off<BG>(k)) return false;
  for (int d = 0; d < DEX_NUM_DOF; ++d)
    if (a.dof2afj[d] != std_dof2afj(d)) return false;
  for (int i = 0; i < DEX_NUM_REWARD_TERMS; ++i) {  // every term of the task is materialised, no other one
    const bool want = BG ? i >= DEX_NUM_COMMON_TERMS : i < DEX_NUM_COMMON_TERMS;
    if ((P.reward_weight[i] != 0.0f) != want) return false;
  }
  if (a.st.contact_forces == nullptr || a.first) return false;
  if (!BG && (!a.common || !a.want_tgt)) return false;
  return true;
}

// ------------------------------------------------------------------------------------------------
// host: TMA staging of the persistent kernel (BlindGrasping).  Every gathered Isaac Gym row becomes one 2-D tensor copy
// (box = [w floats] x [32 envs] out of rigid_body_state / net_contact_force viewed as [N][NB * 13 | NB * 3]); the TMA
// unit wants 16-byte aligned box origins and row pitches, so a row that starts at float offset r inside an env's record is
// fetched from (r & ~3) with the smallest box width that still covers it (hand: 16; pose: 8 or 12; contact: 4 or 8) and
// the consumers read from element (r & 3) on.  actor_root_state (pitch NA * 52 B) is viewed as [N / g][g * NA * 13] with
// g = 1 or 2 so that the pitch is a multiple of 16.  The [K][N] state rows are one box [32 envs] x [K rows] per buffer and
// land exactly in the stage's row layout; the dof tile and the single 128 / 256-byte rows are plain bulk copies.  One copy
// per lane of producer warp 0; all of them complete on the stage's mbarrier (expect_tx).  Needs: N % 32 == 0, NB % 4 == 0,
// NA % 2 == 0 (or % 4), 16-byte aligned bases; anything else keeps the LDGSTS producers.
// ------------------------------------------------------------------------------------------------
typedef CUresult (*TmapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline TmapEncodeFn tmap_encoder() {
  static TmapEncodeFn fn = []() -> TmapEncodeFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    return (TmapEncodeFn)p;
  }();
  return fn;
}

inline bool tmap_2d(CUtensorMap* m, const void* base, const uint64_t inner, const uint64_t outer, const uint64_t pitch_bytes,
                    const uint32_t box0, const uint32_t box1) {
  TmapEncodeFn enc = tmap_encoder();
  if (enc == nullptr) return false;
  const cuuint64_t dims[2] = {inner, outer}, strides[1] = {pitch_bytes};
  const cuuint32_t box[2] = {box0, box1}, es[2] = {1, 1};
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// fills a.tma / a.t_* / a.tmap / a.stage_bytes / a.smem_bytes; leaves a.tma = 0 when the launch does not qualify
inline void build_tma(Post2Args& a) {
  using LY = Lay<true>;
  a.tma = 0;
  a.stage_bytes = LY::STAGE_BYTES;
  static const bool env_on = []() {  // DEXSTEP_POST_TMA=1: as if every call carried DEX_FLAG_TMA_STAGING
    const char* e = getenv("DEXSTEP_POST_TMA");
    return e != nullptr && e[0] == '1';
  }();
  const bool allow = env_on || (a.flags & DEX_FLAG_TMA_STAGING) != 0;
  const DexParams& P = a.p;
  const DexStateBuffers& st = a.st;
  const int N = a.n, NB = P.num_bodies, NC = P.num_contact_bodies, NA = P.num_actors;
  if (!allow || (P.rigid_row_floats != 0 && P.rigid_row_floats != 13) || N % 32 != 0 || NB % 4 != 0 || NC % 4 != 0 || NA % 2 != 0 || !a.bulk_ok) return;
  const void* bases[] = {a.sim.rigid_body_state, a.sim.net_contact_force, a.sim.actor_root_state, a.sim.dof_state,
                         st.prev_dof_pos, st.active_prev_targets, st.prev_actions, st.contact_duration_steps,
                         st.prev_contact_binary, st.initial_box_positions, st.time_in_stage, st.stage_contact_duration,
                         st.success_duration_steps, st.current_stage, st.episode_step_count};
  for (const void* b : bases)
    if (b == nullptr || ((uintptr_t)b & 15u)) return;
  const int g = (NA % 4 == 0) ? 1 : 2;
  // ---- the maps are cached: a pipeline calls with the same buffers every step ----
  struct Key {
    const void* p[9];
    int n, nb, nc, na, nact;
  };
  static thread_local Key key = {};
  static thread_local CUtensorMap maps[14];
  static thread_local bool maps_ok = false;
  const Key now = {{a.sim.rigid_body_state, a.sim.net_contact_force, a.sim.actor_root_state, st.prev_dof_pos, st.active_prev_targets,
                    st.prev_actions, st.contact_duration_steps, st.prev_contact_binary, st.initial_box_positions},
                   N, NB, NC, NA, P.num_actions};
  enum { M_RB16, M_RB12, M_RB8, M_CF8, M_CF4, M_ROOT, M_PDP, M_APT, M_PACT, M_CDUR, M_PBIN, M_BOX0 };
  if (!maps_ok || memcmp(&key, &now, sizeof(Key)) != 0) {
    maps_ok = false;
    const uint64_t rbp = (uint64_t)NB * 52, cfp = (uint64_t)NC * 12, rtp = (uint64_t)g * NA * 52, sp = (uint64_t)N * 4;
    bool ok = tmap_2d(&maps[M_RB16], a.sim.rigid_body_state, (uint64_t)NB * 13, N, rbp, 16, 32) &&
              tmap_2d(&maps[M_RB12], a.sim.rigid_body_state, (uint64_t)NB * 13, N, rbp, 12, 32) &&
              tmap_2d(&maps[M_RB8], a.sim.rigid_body_state, (uint64_t)NB * 13, N, rbp, 8, 32) &&
              tmap_2d(&maps[M_CF8], a.sim.net_contact_force, (uint64_t)NC * 3, N, cfp, 8, 32) &&
              tmap_2d(&maps[M_CF4], a.sim.net_contact_force, (uint64_t)NC * 3, N, cfp, 4, 32) &&
              tmap_2d(&maps[M_ROOT], a.sim.actor_root_state, (uint64_t)g * NA * 13, N / g, rtp, 16, 32 / g) &&
              tmap_2d(&maps[M_PDP], st.prev_dof_pos, N, DEX_NUM_DOF, sp, 32, DEX_NUM_DOF) &&
              tmap_2d(&maps[M_APT], st.active_prev_targets, N, DEX_NUM_ACTIVE, sp, 32, DEX_NUM_ACTIVE) &&
              tmap_2d(&maps[M_PACT], st.prev_actions, N, P.num_actions, sp, 32, P.num_actions) &&
              tmap_2d(&maps[M_CDUR], st.contact_duration_steps, N, DEX_NUM_FINGERS, sp, 32, DEX_NUM_FINGERS) &&
              tmap_2d(&maps[M_PBIN], st.prev_contact_binary, N, DEX_NUM_FINGERS, sp, 32, DEX_NUM_FINGERS) &&
              tmap_2d(&maps[M_BOX0], st.initial_box_positions, N, 3, sp, 32, 3);
    if (!ok) return;
    key = now;
    maps_ok = true;
  }
  memcpy(a.tmap, maps, sizeof(maps));
  // ---- copy table + consumer view ----
  int k = 0;
  uint32_t raw = 0, tx = 0;  // raw: byte offset inside the raw block (behind OFF_AOS)
  auto box_copy = [&](int kind, int map, uint32_t dst, int c0, uint32_t bytes) {
    a.t_kind[k] = (uint8_t)kind;
    a.t_early[k] = kind < 2 ? 1 : 0;
    a.t_map[k] = (uint8_t)map;
    a.t_dst[k] = dst;
    a.t_c0[k] = c0;
    a.t_bytes[k] = bytes;
    a.t_ptr[k] = 0;
    tx += bytes;
    ++k;
  };
  auto bulk_copy = [&](const void* base, uint32_t dst, uint32_t bytes) {
    a.t_kind[k] = 3;
    a.t_early[k] = base == (const void*)a.sim.dof_state ? 1 : 0;
    a.t_map[k] = 0;
    a.t_dst[k] = dst;
    a.t_c0[k] = 0;
    a.t_bytes[k] = bytes;
    a.t_ptr[k] = (uint64_t)(uintptr_t)base;
    tx += bytes;
    ++k;
  };
  bulk_copy(a.sim.dof_state, LY::OFF_DOF, 32 * 52 * 4);
  {  // hand row: 13 floats
    const int r = a.rb_off[0];
    a.t_hand = (uint16_t)(raw / 4 + (r & 3));
    box_copy(0, M_RB16, LY::OFF_AOS + raw, r & ~3, 32 * 16 * 4);
    raw += 32 * 16 * 4;
  }
  for (int f = 0; f < 10; ++f) {  // tip / pad poses: 7 floats
    const int r = a.rb_off[1 + f], w = ((r & 3) + 7 <= 8) ? 8 : 12;
    a.t_pose[f] = (uint16_t)(raw / 4 + (r & 3));
    a.t_pose_w[f] = (uint8_t)w;
    box_copy(0, w == 8 ? M_RB8 : M_RB12, LY::OFF_AOS + raw, r & ~3, 32u * w * 4);
    raw += (32u * w * 4 + 127) / 128 * 128;
  }
  for (int f = 0; f < 6; ++f) {  // contact forces (5 monitored bodies + the box): 3 floats
    const int r = a.cf_off[f], w = ((r & 3) + 3 <= 4) ? 4 : 8;
    a.t_cf[f] = (uint16_t)(raw / 4 + (r & 3));
    a.t_cf_w[f] = (uint8_t)w;
    box_copy(0, w == 4 ? M_CF4 : M_CF8, LY::OFF_AOS + raw, r & ~3, 32u * w * 4);
    raw += (32u * w * 4 + 127) / 128 * 128;
  }
  a.t_root_g = g;
  a.t_root_sh = g == 1 ? 0 : 1;
  for (int j = 0; j < g; ++j) {  // box root row of the envs e = g * row + j: position [0..2], velocity [7..9]
    const int r = (j * NA + P.box_actor) * 13;
    a.t_box[j] = (uint16_t)(raw / 4 + (r & 3));
    box_copy(1, M_ROOT, LY::OFF_AOS + raw, r & ~3, (32u / g) * 16 * 4);
    raw += (32u / g) * 16 * 4;
  }
  if (g == 1) a.t_box[1] = a.t_box[0];
  box_copy(2, M_PDP, LY::R_PDP * 128, 0, 32 * 4 * DEX_NUM_DOF);
  box_copy(2, M_APT, LY::R_APT * 128, 0, 32 * 4 * DEX_NUM_ACTIVE);
  box_copy(2, M_PACT, LY::R_PACT * 128, 0, 32u * 4 * P.num_actions);
  box_copy(2, M_CDUR, LY::R_CDUR * 128, 0, 32 * 4 * DEX_NUM_FINGERS);
  box_copy(2, M_PBIN, LY::R_PBIN * 128, 0, 32 * 4 * DEX_NUM_FINGERS);
  box_copy(2, M_BOX0, LY::R_BOX0 * 128, 0, 32 * 4 * 3);
  bulk_copy(st.time_in_stage, LY::R_TIS * 128, 128);
  bulk_copy(st.stage_contact_duration, LY::R_SCD * 128, 128);
  bulk_copy(st.episode_step_count, LY::OFF_I64 + LY::I_STEP * 256, 256);
  bulk_copy(st.success_duration_steps, LY::OFF_I64 + LY::I_SDS * 256, 256);
  bulk_copy(st.current_stage, LY::OFF_I64 + LY::I_STAGE * 256, 256);
  if (k > 32) return;
  const uint32_t stage_bytes = (LY::OFF_AOS + raw + 127) / 128 * 128;
  const uint32_t smem = (uint32_t)(kStages * stage_bytes + LY::W_COUNT * 128 + (32 * a.wp * 4 + 127) / 128 * 128);
  if (smem > 113 * 1024) return;  // two CTAs per SM must still fit
  a.t_ncopy = k;
  a.tma_tx = tx;
  a.stage_bytes = stage_bytes;
  a.smem_bytes = smem;
  a.tma = 1;
}

#ifdef DEXSTEP_TL2  // debug build only (tools/timeline_v2.py): SM-clock marks per (CTA, tile iteration), [cta][kTlIters][32] u64:
// [0..15] consumer thread 0, [16..23] producer thread 0, [30] globaltimer ns at the mark 0 of the iteration, [31] smid
constexpr int kTlIters = 8;
__device__ unsigned long long* g_tl2 = nullptr;
#define TL2_DECL unsigned long long* const tl2 = g_tl2
#define TL2_PARAM , unsigned long long* const tl2
#define TL2_ARG , tl2
#define TL2(it, k)                                                                                                 \
  do {                                                                                                             \
    if (tl2 != nullptr && (it) < kTlIters && (threadIdx.x == 0 || threadIdx.x == kConsThreads)) {                 \
      unsigned long long* q_ = tl2 + ((size_t)blockIdx.x * kTlIters + (it)) * 32;                                  \
      q_[(k)] = (unsigned long long)clock64();                                                                     \
      if ((k) == 0) {                                                                                              \
        unsigned long long t_;                                                                                     \
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                                     \
        unsigned sm_;                                                                                              \
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm_));                                                           \
        q_[30] = t_;                                                                                               \
        q_[31] = sm_;                                                                                              \
      }                                                                                                            \
    }                                                                                                              \
  } while (0)
#else
#define TL2_DECL
#define TL2_PARAM
#define TL2_ARG
#define TL2(it, k)
#endif

// ------------------------------------------------------------------------------------------------
// device: mbarrier / bulk-copy / named-barrier primitives
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(const uint32_t bar, const uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(const uint32_t bar) {
  asm volatile("{\n.reg .b64 t;\nmbarrier.arrive.shared::cta.b64 t, [%0];\n}" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_tx(const uint32_t bar, const uint32_t bytes) {
  asm volatile("{\n.reg .b64 t;\nmbarrier.arrive.expect_tx.shared::cta.b64 t, [%0], %1;\n}" ::"r"(bar), "r"(bytes) : "memory");
}
// try_wait with a suspend-time hint: the warp sleeps in hardware until the phase completes (or ~the hint elapses) instead
// of spinning through the issue slots of the warps that do the work
__device__ __forceinline__ bool mbar_try(const uint32_t bar, const uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\nselp.u32 %0, 1, 0, p;\n}"
               : "=r"(ok)
               : "r"(bar), "r"(parity), "r"(1000000u)
               : "memory");
  return ok != 0;
}
// Bounded wait: a stage that is never filled / released (a bug, or a faulted peer warp) traps after ~2 s of SM clocks
// instead of hanging the stream - the host sees a launch failure on its next CUDA call.
constexpr long long kSpinLimitClocks = 4000000000ll;
__device__ __forceinline__ void mbar_wait(const uint32_t bar, const uint32_t parity) {
  if (mbar_try(bar, parity)) return;
  const long long t0 = clock64();
  unsigned n = 0;
  while (!mbar_try(bar, parity)) {
    if ((++n & 255u) == 0u && clock64() - t0 > kSpinLimitClocks) __trap();
  }
}
// global -> shared bulk copy (TMA engine, SASS UBLKCP); src, dst 16-byte aligned, bytes a multiple of 16
__device__ __forceinline__ void bulk_g2s(const uint32_t dst, const void* src, const uint32_t bytes, const uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}
// Ampere-style asynchronous copies (SASS LDGSTS): global -> shared without a register in between; their completion is
// tied to the stage's mbarrier by cp.async.mbarrier.arrive.noinc (one counted arrival per producer thread and tile)
__device__ __forceinline__ void cpa4(const uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cpa8(const uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cpa_arrive(const uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void obs_put(float* obs, const int off, const int j, const float v) {
  if (off >= 0) obs[off + j] = v;
}

// ------------------------------------------------------------------------------------------------
// producer side: one producer warp fills the slot of its compute warp
// ------------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------------
// consumer phases (lane = env)
// ------------------------------------------------------------------------------------------------
template <int LM>
constexpr int kBoxVel = LM == 2 ? 7 : 3;  // where the box's linear velocity sits behind Cx::box()

template <bool BG, bool STD>
struct Cx {
  const unsigned char* stage;
  float* wk;   // work block + lane
  float* obs;  // obs tile row of this lane's env
  const int32_t* obs_offset;  // generic instance: DexParams.obs_offset
  const int32_t* dof2afj;
  int lane, cw, e0, ne;
  bool act;
  size_t N, env;
  __device__ __forceinline__ int off(const int k) const {
    if constexpr (STD) return std_obs_off<BG>(k);
    else return obs_offset[k];
  }
  __device__ __forceinline__ int afj(const int d) const { return dof2afj[d]; }  // run-time d: one constant-bank load
  __device__ __forceinline__ const float* soa() const { return reinterpret_cast<const float*>(stage) + lane; }
  __device__ __forceinline__ const float* aos() const {
    return reinterpret_cast<const float*>(stage + Lay<BG>::OFF_AOS) + lane * Lay<BG>::SA;
  }
  // ---- the gathered Isaac Gym rows of this lane's env.  LM (layout mode of the stage): 0 = [32][SA] records filled by
  // LDGSTS / register-staged loads, 1 = the same inside the aliased single-pass stage, 2 = the dense boxes the TMA unit
  // wrote (one [32 envs][w floats] box per gathered row, w = 4 .. 16; element offsets from build_tma) ----
  __device__ __forceinline__ const float* rawf() const { return reinterpret_cast<const float*>(stage + Lay<BG>::OFF_AOS); }
  template <int LM>
  __device__ __forceinline__ const float* hand(const Post2Args& A) const {  // 13 floats: pos, quat, lin vel, ang vel
    if constexpr (LM == 2) return rawf() + A.t_hand + lane * 16;
    else return aos() + Lay<BG>::A_RB;
  }
  template <int LM>
  __device__ __forceinline__ const float* pose(const Post2Args& A, const int f) const {  // 7 floats; f: 5 tips, 5 pads
    if constexpr (LM == 2) return rawf() + A.t_pose[f] + lane * (int)A.t_pose_w[f];
    else return aos() + Lay<BG>::A_RB + 13 + 7 * f;
  }
  template <int LM>
  __device__ __forceinline__ const float* contact(const Post2Args& A, const int f) const {  // 3 floats; f = 5: the box
    if constexpr (LM == 2) return rawf() + A.t_cf[f] + lane * (int)A.t_cf_w[f];
    else return aos() + Lay<BG>::A_CF + 3 * f;
  }
  template <int LM>
  __device__ __forceinline__ const float* box(const Post2Args& A) const {  // position at [0..2], velocity at [kBoxVel<LM> ..]
    if constexpr (LM == 2) return rawf() + ((lane & (A.t_root_g - 1)) ? A.t_box[1] : A.t_box[0]) + (lane >> A.t_root_sh) * 16;
    else return aos() + Lay<BG>::A_BOX;
  }
  __device__ __forceinline__ int i64lo(const int r) const {
    return reinterpret_cast<const int*>(stage + Lay<BG>::OFF_I64 + r * 256)[2 * lane];
  }
};

// phase A: per-env work items, taken round-robin by the 8 consumer warps (item type is warp-uniform)
template <bool BG, bool STD, int LM>
__device__ __forceinline__ void phase_items(const Post2Args& A, const Cx<BG, STD>& c) {
  using LY = Lay<BG, LM == 1>;
  if (!c.act) return;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float dt = P.control_dt;
  const float* soa = c.soa();
  const float* rb = c.template hand<LM>(A);  // the hand row (13 floats)
  float* wk = c.wk;
  float* obs = c.obs;
  const size_t N = c.N, env = c.env;
  const bool common = !BG && (STD || A.common != 0), first = !STD && A.first != 0;
  constexpr int I_CONTACT = 10, I_DOF = 15, I_PACT = 28, I_APT = 31, I_HAND = 34, I_BOX = 35, I_PAIR = 36;
  constexpr int N_ITEMS = BG ? I_PAIR + 5 : I_HAND + 1;

  for (int it = c.cw; it < N_ITEMS; it += kConsWarps) {
    if (it < I_CONTACT) {
      // ---- tip / pad pose f (observation_encoder.py:915-986; blind_grasping_task.py:684-724, 1178-1192) ----
      const int f = it;
      const bool is_pad = f >= DEX_NUM_FINGERS;
      const int g = is_pad ? f - DEX_NUM_FINGERS : f;
      const float* p = c.template pose<LM>(A, f);
      const float px = p[0], py = p[1], pz = p[2], qx = p[3], qy = p[4], qz = p[5], qw = p[6];
      const int ow = c.off(is_pad ? DEX_K_FINGERPAD_POSES_WORLD : DEX_K_FINGERTIP_POSES_WORLD);
      if (ow >= 0) {
        float* o = obs + ow + 7 * g;
        o[0] = px;
        o[1] = py;
        o[2] = pz;
        o[3] = qx;
        o[4] = qy;
        o[5] = qz;
        o[6] = qw;
      }
      const int oh = c.off(is_pad ? DEX_K_FINGERPAD_POSES_HAND : DEX_K_FINGERTIP_POSES_HAND);
      if (oh >= 0) {
        const Quat hq{rb[3], rb[4], rb[5], rb[6]};
        const Vec3 ph = quat_rotate_inverse(hq, Vec3{px - rb[0], py - rb[1], pz - rb[2]});
        const Quat qh = quat_mul(quat_conj(hq), Quat{qx, qy, qz, qw});
        float* o = obs + oh + 7 * g;
        o[0] = ph.x;
        o[1] = ph.y;
        o[2] = ph.z;
        o[3] = qh.x;
        o[4] = qh.y;
        o[5] = qh.z;
        o[6] = qh.w;
      }
      if (BG) {
        const float* box = c.template box<LM>(A);
        const float d = norm3(px - box[0], py - box[1], pz - box[2]);
        if (is_pad) {
          const float hd = fabsf(pz - box[2]);
          wk[(LY::W_PADBOX + g) * 32] = d;
          wk[(LY::W_HDIFF + g) * 32] = hd;
          obs_put(obs, c.off(DEX_K_FINGER_TO_OBJECT_DISTANCES), g, d);
          obs_put(obs, c.off(DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF), g, hd);
        } else {
          wk[(LY::W_TIPBOX + g) * 32] = d;
        }
      }
    } else if (it < I_DOF) {
      // ---- contact sensor f (observation_encoder.py:658-735, 323-361; reward_calculator.py:183-199) ----
      const int f = it - I_CONTACT;
      const float* cf = c.template contact<LM>(A, f);
      const float fx = cf[0], fy = cf[1], fz = cf[2];
      const float mag = norm3(fx, fy, fz);
      const bool bin = mag > P.contact_binary_threshold;
      const float dur = soa[(LY::R_CDUR + f) * 32], prevb = soa[(LY::R_PBIN + f) * 32];
      const float ndur = (bin && prevb == 0.0f) ? 1.0f : (bin ? dur + 1.0f : 0.0f);
      const float fb = bin ? 1.0f : 0.0f;
      st.contact_duration_steps[f * N + env] = ndur;
      st.prev_contact_binary[f * N + env] = fb;
      if (STD || st.contact_forces != nullptr) {
        st.contact_forces[(3 * f + 0) * N + env] = fx;
        st.contact_forces[(3 * f + 1) * N + env] = fy;
        st.contact_forces[(3 * f + 2) * N + env] = fz;
      }
      wk[(LY::W_BIN + f) * 32] = fb;
      const int ocf = c.off(DEX_K_CONTACT_FORCES);
      if (ocf >= 0) {
        obs[ocf + 3 * f] = fx;
        obs[ocf + 3 * f + 1] = fy;
        obs[ocf + 3 * f + 2] = fz;
      }
      obs_put(obs, c.off(DEX_K_CONTACT_FORCE_MAGNITUDE), f, mag);
      obs_put(obs, c.off(DEX_K_CONTACT_BINARY), f, fb);
      obs_put(obs, c.off(DEX_K_CONTACT_DURATION), f, ndur * dt);
      if (!BG && common) {
        const bool c_now = mag > 0.1f;  // hard-coded 0.1 N (reward_calculator.py:186)
        const bool c_prev = first ? c_now : (c.stage[LY::OFF_U8 + f * 32 + c.lane] != 0);
        wk[(LY::W_CHG + f) * 32] = (c_now != c_prev) ? 1.0f : 0.0f;
        st.prev_contacts[f * N + env] = c_now ? 1 : 0;
      }
    } else if (it < I_PACT) {
      // ---- DOF pair (2q, 2q+1): finite-difference velocity (observation_encoder.py:298-321), obs columns,
      //      common-reward partial terms (reward_calculator.py:94-98,138-160: Isaac dof_vel, not the finite difference)
      const int q = it - I_DOF;
      const float4 pv4 = reinterpret_cast<const float4*>(c.stage + LY::OFF_DOF)[c.lane * 13 + q];
      const float pos2[2] = {pv4.x, pv4.z}, vel2[2] = {pv4.y, pv4.w};
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int d = 2 * q + h;
        const float pos = pos2[h];
        const float dp = pos - soa[(LY::R_PDP + d) * 32];
        const float fd = div_rn(dp, dt);
        st.prev_dof_pos[d * N + env] = pos;
        float tgt = 0.0f;
        if (!BG && (STD || A.want_tgt)) tgt = reinterpret_cast<const float*>(c.stage + LY::OFF_TGT)[c.lane * 26 + d];
        if (d < DEX_NUM_BASE) {
          obs_put(obs, c.off(DEX_K_BASE_DOF_POS), d, pos);
          obs_put(obs, c.off(DEX_K_BASE_DOF_VEL), d, fd);
          if (!BG) obs_put(obs, c.off(DEX_K_BASE_DOF_TARGET), d, tgt);
          if (BG) wk[(LY::W_BVEL + d) * 32] = fd;
        } else {
          const int j = d - DEX_NUM_BASE;
          obs_put(obs, c.off(DEX_K_ALL_FINGER_DOF_POS), j, pos);
          obs_put(obs, c.off(DEX_K_ALL_FINGER_DOF_VEL), j, fd);
          if (!BG) obs_put(obs, c.off(DEX_K_ALL_FINGER_DOF_TARGET), j, tgt);
          const int aj = c.afj(d);
          if (aj >= 0) {
            obs_put(obs, c.off(DEX_K_ACTIVE_FINGER_DOF_POS), aj, pos);
            obs_put(obs, c.off(DEX_K_ACTIVE_FINGER_DOF_VEL), aj, fd);
            if (!BG) obs_put(obs, c.off(DEX_K_ACTIVE_FINGER_DOF_TARGET), aj, tgt);
          }
          if (!BG && common) {
            const float iv = vel2[h];
            const float pvv = first ? iv : soa[(LY::R_PFDV + j) * 32];
            wk[(LY::W_FV2 + j) * 32] = iv * iv;
            const float acc = iv - pvv;
            wk[(LY::W_FA2 + j) * 32] = acc * acc;
            const float nj = div_rn(2.0f * (pos - P.dof_lower[d]), P.dof_upper[d] - P.dof_lower[d]) - 1.0f;
            wk[(LY::W_JL + j) * 32] = clampf(fabsf(nj) - 0.8f, 0.0f, 1.0f);
            st.prev_finger_dof_vel[j * N + env] = iv;
          }
        }
      }
    } else if (it < I_APT) {
      // ---- prev_actions columns (observation_encoder.py:287-296) ----
      const int k0 = 6 * (it - I_PACT);
      const int op = c.off(DEX_K_PREV_ACTIONS);
      if (op >= 0 && k0 < P.num_actions) {
#pragma unroll
        for (int k = 0; k < 6; ++k) obs[op + k0 + k] = soa[(LY::R_PACT + k0 + k) * 32];  // num_actions is 6, 12 or 18
      }
    } else if (it < I_HAND) {
      // ---- active_rule_targets <- active_prev_targets (rules.py:78-95, step_processor.py:73) ----
      const int k0 = 6 * (it - I_APT);
      const int o1 = c.off(DEX_K_ACTIVE_PREV_TARGETS), o2 = c.off(DEX_K_ACTIVE_RULE_TARGETS);
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        const float v = soa[(LY::R_APT + k0 + k) * 32];
        st.active_rule_targets[(k0 + k) * N + env] = v;
        obs_put(obs, o1, k0 + k, v);
        obs_put(obs, o2, k0 + k, v);
      }
    } else if (it == I_HAND) {
      // ---- hand pose, ARR-aligned pose (observation_encoder.py:1483-1521), hand -> box distance,
      //      hand velocity reward terms (reward_calculator.py:100-181) ----
      const int oh = c.off(DEX_K_HAND_POSE);
      if (oh >= 0) {
#pragma unroll
        for (int j = 0; j < 7; ++j) obs[oh + j] = rb[j];
      }
      const int oa = c.off(DEX_K_HAND_POSE_ARR_ALIGNED);
      if (oa >= 0) {
        const float sh = 0.7071067811865476f;
        const Quat qa = quat_mul(Quat{rb[3], rb[4], rb[5], rb[6]}, Quat{-0.0f, -sh, -0.0f, sh});
        obs[oa + 0] = rb[0];
        obs[oa + 1] = rb[1];
        obs[oa + 2] = rb[2];
        obs[oa + 3] = qa.x;
        obs[oa + 4] = qa.y;
        obs[oa + 5] = qa.z;
        obs[oa + 6] = qa.w;
      }
      if (BG) {
        const float* box = c.template box<LM>(A);
        const float hb = norm3(rb[0] - box[0], rb[1] - box[1], rb[2] - box[2]);
        wk[LY::W_HANDBOX * 32] = hb;
        obs_put(obs, c.off(DEX_K_HAND_TO_OBJECT_DISTANCE), 0, hb);
      }
      if (!BG && common) {
        float hv[3], ha[3], phv[3], pha[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          hv[k] = rb[7 + k];
          ha[k] = rb[10 + k];
          phv[k] = first ? hv[k] : soa[(LY::R_PHV + k) * 32];
          pha[k] = first ? ha[k] : soa[(LY::R_PHA + k) * 32];
          st.prev_hand_vel[k * N + env] = hv[k];
          st.prev_hand_ang_vel[k * N + env] = ha[k];
        }
        wk[(LY::W_RHAND + 0) * 32] = expf(-0.2f * norm3(hv[0], hv[1], hv[2]));
        wk[(LY::W_RHAND + 1) * 32] = expf(-0.2f * norm3(ha[0], ha[1], ha[2]));
        wk[(LY::W_RHAND + 2) * 32] = expf(-0.5f * norm3(hv[0] - phv[0], hv[1] - phv[1], hv[2] - phv[2]));
        wk[(LY::W_RHAND + 3) * 32] = expf(-0.5f * norm3(ha[0] - pha[0], ha[1] - pha[1], ha[2] - pha[2]));
      }
    } else if (BG && it == I_BOX) {
      // ---- pad centroid, box stability inputs (blind_grasping_task.py:746-771, 1334-1364) ----
      const float* box = c.template box<LM>(A);
      constexpr int BV = kBoxVel<LM>;
      float cen[3];
#pragma unroll
      for (int k = 0; k < 3; ++k)
        cen[k] = (c.template pose<LM>(A, 5)[k] + c.template pose<LM>(A, 6)[k] + c.template pose<LM>(A, 7)[k]) / 3.0f;
      const int oc = c.off(DEX_K_FIRST_THREE_FINGERPAD_CENTROID);
      const int opos = c.off(DEX_K_OBJECT_POS), ovel = c.off(DEX_K_OBJECT_VEL);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        obs_put(obs, oc, k, cen[k]);
        obs_put(obs, opos, k, box[k]);
        obs_put(obs, ovel, k, box[BV + k]);
      }
      wk[LY::W_CDIST * 32] = norm3(cen[0] - box[0], cen[1] - box[1], cen[2] - box[2]);
      wk[LY::W_DRIFT * 32] = norm3(box[0] - soa[(LY::R_BOX0 + 0) * 32], box[1] - soa[(LY::R_BOX0 + 1) * 32],
                                   box[2] - soa[(LY::R_BOX0 + 2) * 32]);
      wk[LY::W_VMAG * 32] = norm3(box[BV], box[BV + 1], box[BV + 2]);
      const float* cfb = c.template contact<LM>(A, 5);
      wk[LY::W_BOXHAS * 32] = (norm3(cfb[0], cfb[1], cfb[2]) > P.contact_binary_threshold) ? 1.0f : 0.0f;
    } else if (BG) {
      // ---- two pad pair distances, triu order (blind_grasping_task.py:726-771) ----
      const int op = c.off(DEX_K_FINGERPAD_DISTANCES);
      if (op >= 0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int k = 2 * (it - I_PAIR) + h;
          // k -> (a, b): rows of the strict upper triangle of a 5x5 matrix start at 0, 4, 7, 9
          const int a = (k >= 4) + (k >= 7) + (k >= 9);
          const int b = k - (a == 0 ? 0 : a == 1 ? 4 : a == 2 ? 7 : 9) + a + 1;
          const float* pa = c.template pose<LM>(A, 5 + a);
          const float* pb = c.template pose<LM>(A, 5 + b);
          obs[op + k] = norm3(pa[0] - pb[0], pa[1] - pb[1], pa[2] - pb[2]);
        }
      }
    }
  }
}

// one common reward term (reward_calculator.py:66-221); BaseTask, phase B
template <bool BG, bool STD, int LM>
__device__ __forceinline__ float common_term(const Post2Args& A, const Cx<BG, STD>& c, const int i) {
  using LY = Lay<BG, LM == 1>;
  const float* wk = c.wk;
  switch (i) {
    case DEX_R_ALIVE: return 1.0f;
    case DEX_R_HEIGHT_SAFETY: {
      float mz = c.template pose<LM>(A, 0)[2];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) mz = fminf(mz, c.template pose<LM>(A, f)[2]);
      return clampf(1.0f - expf(-(mz - 0.02f) * 20.0f), 0.0f, 1.0f);
    }
    case DEX_R_FINGER_VELOCITY: {
      float a = 0.0f;
#pragma unroll
      for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) a += wk[(LY::W_FV2 + j) * 32];
      return expf(-0.1f * sqrtf(a));
    }
    case DEX_R_HAND_VELOCITY: return wk[(LY::W_RHAND + 0) * 32];
    case DEX_R_HAND_ANGULAR_VELOCITY: return wk[(LY::W_RHAND + 1) * 32];
    case DEX_R_JOINT_LIMIT: {
      float a = 0.0f;
#pragma unroll
      for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) a += wk[(LY::W_JL + j) * 32];
      return 1.0f - div_rn(a, 20.0f);
    }
    case DEX_R_FINGER_ACCELERATION: {
      float a = 0.0f;
#pragma unroll
      for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) a += wk[(LY::W_FA2 + j) * 32];
      return expf(-2.0f * sqrtf(a));
    }
    case DEX_R_HAND_ACCELERATION: return wk[(LY::W_RHAND + 2) * 32];
    case DEX_R_HAND_ANGULAR_ACCELERATION: return wk[(LY::W_RHAND + 3) * 32];
    default: {  // DEX_R_CONTACT_STABILITY
      float a = 0.0f;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) a += wk[(LY::W_CHG + f) * 32];
      return expf(-a);
    }
  }
}

// one BlindGrasping task term (blind_grasping_task.py:980-1208); phase C (after the FSM)
template <bool BG, bool STD, int LM>
__device__ __forceinline__ float task_term(const Post2Args& A, const Cx<BG, STD>& c, const int i) {
  using LY = Lay<BG, LM == 1>;
  const DexParams& P = A.p;
  const float* wk = c.wk;
  const int stage = __float_as_int(wk[LY::W_STAGE * 32]);
  const float m1 = (stage == 1) ? 1.0f : 0.0f, m2 = (stage == 2) ? 1.0f : 0.0f, m3 = (stage == 3) ? 1.0f : 0.0f;
  switch (i) {
    case DEX_R_S1_HEIGHT_ALIGNMENT: return expf(-P.height_alignment_decay * wk[LY::W_AVGHDIFF * 32]) * m1;
    case DEX_R_S1_CENTROID_POSITIONING: return expf(-P.centroid_positioning_decay * wk[LY::W_CDIST * 32]) * m1;
    case DEX_R_S1_OBJECT_STABILITY:
      return expf(-P.object_stability_decay * (wk[LY::W_DRIFT * 32] + wk[LY::W_VMAG * 32])) * m1;
    case DEX_R_S1_FINGER_HEIGHT_CONSISTENCY: {  // torch.var(dim=1), unbiased, z of the first three pads
      const float z0 = c.template pose<LM>(A, 5)[2], z1 = c.template pose<LM>(A, 6)[2], z2 = c.template pose<LM>(A, 7)[2];
      const float mean = (z0 + z1 + z2) / 3.0f;
      const float d0 = z0 - mean, d1 = z1 - mean, d2 = z2 - mean;
      const float var = (d0 * d0 + d1 * d1 + d2 * d2) / 2.0f;
      return expf(-P.height_consistency_decay * var) * m1;
    }
    case DEX_R_S1_THUMB_ROTATION: {
      const float th = reinterpret_cast<const float*>(c.stage + LY::OFF_DOF)[c.lane * 52 + 2 * DEX_NUM_BASE];
      return expf(-5.0f * fabsf(th - P.half_pi)) * m1;
    }
    case DEX_R_S2_THUMB_CONTACT: return wk[LY::W_THUMB * 32] * m2;
    case DEX_R_S2_OTHER_FINGERS_CONTACT: return wk[LY::W_OTHERS * 32] * m2;
    case DEX_R_S2_GRASP_ACHIEVEMENT: return wk[LY::W_GRASP * 32] * m2;
    case DEX_R_S2_FINGERPAD_PROXIMITY:
      return expf(-P.fingerpad_proximity_decay * fmaxf(wk[LY::W_MINTIPBOX * 32], P.min_reward_distance)) * m2;
    case DEX_R_S2_BASE_STABILITY: {
      float bv2 = 0.0f;
#pragma unroll
      for (int d = 0; d < DEX_NUM_BASE; ++d) {
        const float v = wk[(LY::W_BVEL + d) * 32];  // finite-difference base velocity
        bv2 += v * v;
      }
      return expf(-P.base_stability_decay * sqrtf(bv2)) * m2;
    }
    case DEX_R_S3_OBJECT_HEIGHT:
      return clampf(div_rn(c.template box<LM>(A)[2] - P.box_z, P.lift_height_range), 0.0f, 1.0f) * m3;
    case DEX_R_S3_GRASP_MAINTENANCE: return wk[LY::W_GRASP * 32] * m3;
    case DEX_R_S3_GRASP_DURATION:
      return clampf(div_rn(wk[LY::W_GRASPDUR * 32], P.contact_duration_threshold_s), 0.0f, 1.0f) * m3;
    // completion bonuses :1025-1048 (transition and not failed), penetration :1194-1208
    case DEX_R_S1_COMPLETION: return (wk[LY::W_J2 * 32] != 0.0f && wk[LY::W_QUALOK * 32] != 0.0f) ? 1.0f : 0.0f;
    case DEX_R_S2_COMPLETION: return (wk[LY::W_J3 * 32] != 0.0f && wk[LY::W_GRASP * 32] != 0.0f) ? 1.0f : 0.0f;
    default:  // DEX_R_PENETRATION_PENALTY
      return fmaxf(P.penetration_threshold - wk[LY::W_MINTIPBOX * 32], 0.0f) * P.penetration_depth_scale;
  }
}

// raw + weighted rows of term i leave from registers; the weighted value is parked for the ordered total
template <bool BG, bool STD, int LM>
__device__ __forceinline__ void emit_term(const Post2Args& A, const Cx<BG, STD>& c, const int i, const float r) {
  const float w = A.p.reward_weight[i];
  if (!STD && w == 0.0f) return;  // reward_calculator.py:247-251: zero-weight terms are not materialised (STD: none is)
  const int row = A.p.component_row[i];
  const float rw = r * w;
  float* g = A.out.reward_components + c.env;
  g[(size_t)row * c.N] = r;
  g[(size_t)(row + 1) * c.N] = rw;
  c.wk[(Lay<BG, LM == 1>::W_WGT + i) * 32] = rw;  // (the observer's running sums of these rows: observer_term_rows)
}

// phase B: the per-env scalar part, split over the consumer warps (warp-uniform job index)
template <bool BG, bool STD, int LM>
__device__ __forceinline__ void phase_scalar(const Post2Args& A, const Cx<BG, STD>& c) {
  using LY = Lay<BG, LM == 1>;
  if (!c.act) return;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float dt = P.control_dt;
  float* wk = c.wk;
  float* obs = c.obs;
  const float* soa = c.soa();
  const float* rb = c.template hand<LM>(A);
  const size_t env = c.env;
  if (BG) {
    if (c.cw == 0) {
      // ---- finger-box contacts (blind_grasping_task.py:773-833), success duration (:626-648) ----
      const bool box_has = wk[LY::W_BOXHAS * 32] != 0.0f;
      bool fbc[DEX_NUM_FINGERS];
      int n_on_box = 0;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        fbc[f] = (wk[(LY::W_BIN + f) * 32] != 0.0f) && box_has && (wk[(LY::W_PADBOX + f) * 32] < P.proximity_threshold);
        n_on_box += fbc[f] ? 1 : 0;
      }
      const bool thumb_c = fbc[0], others_c = fbc[1] || fbc[2] || fbc[3] || fbc[4];
      const bool grasp = thumb_c && others_c;
      const float ft = thumb_c ? 1.0f : 0.0f, fo = others_c ? 1.0f : 0.0f, fg = grasp ? 1.0f : 0.0f;
      wk[LY::W_THUMB * 32] = ft;
      wk[LY::W_OTHERS * 32] = fo;
      wk[LY::W_GRASP * 32] = fg;
      obs_put(obs, c.off(DEX_K_THUMB_CONTACT), 0, ft);
      obs_put(obs, c.off(DEX_K_OTHER_FINGERS_CONTACT), 0, fo);
      obs_put(obs, c.off(DEX_K_GRASP_STATE), 0, fg);
      const float box_z = c.template box<LM>(A)[2];
      const bool cond = (box_z > P.success_height_threshold) && (n_on_box >= P.min_fingers_for_grasp);
      const int sds = cond ? c.i64lo(LY::I_SDS) + 1 : 0;
      st.success_duration_steps[env] = (long long)sds;
      st.success_conditions_met[env] = cond ? 1 : 0;
      const float gd = (float)sds * dt;
      wk[LY::W_SDS * 32] = __int_as_float(sds);
      wk[LY::W_GRASPDUR * 32] = gd;
      obs_put(obs, c.off(DEX_K_GRASP_DURATION), 0, gd);
    } else if (c.cw == 1) {
      // ---- stage FSM :835-956 (plain fp32 accumulation of (float)control_dt decides the transition step) ----
      int stage = c.i64lo(LY::I_STAGE);
      float tis = soa[LY::R_TIS * 32] + dt;
      float scd = soa[LY::R_SCD * 32];
      const bool b0 = wk[(LY::W_BIN + 0) * 32] != 0.0f;
      const bool bo = (wk[(LY::W_BIN + 1) * 32] != 0.0f) || (wk[(LY::W_BIN + 2) * 32] != 0.0f) ||
                      (wk[(LY::W_BIN + 3) * 32] != 0.0f) || (wk[(LY::W_BIN + 4) * 32] != 0.0f);
      const bool thumb_other = b0 && bo;
      if (stage == 2) scd = thumb_other ? scd + dt : 0.0f;
      const bool s1_done = (stage == 1) && (tis >= P.stage1_duration);
      const bool s2_done = (stage == 2) && ((scd >= P.stage2_contact_success_threshold) || (tis >= P.stage2_duration));
      if (s1_done) stage = 2;
      if (s2_done) stage = 3;
      if (s1_done || s2_done) {
        tis = 0.0f;
        scd = 0.0f;
      }
      st.current_stage[env] = (long long)stage;
      st.time_in_stage[env] = tis;
      st.stage_contact_duration[env] = scd;
      st.just_transitioned_to_stage2[env] = s1_done ? 1 : 0;
      st.just_transitioned_to_stage3[env] = s2_done ? 1 : 0;
      float prog = 0.0f;  // :958-978
      if (stage == 1) prog = clampf(div_rn(tis, P.stage1_duration), 0.0f, 1.0f);
      if (stage == 2) prog = clampf(div_rn(tis, P.stage2_duration), 0.0f, 1.0f);
      if (stage == 3) prog = 1.0f;
      wk[LY::W_STAGE * 32] = __int_as_float(stage);
      wk[LY::W_J2 * 32] = s1_done ? 1.0f : 0.0f;
      wk[LY::W_J3 * 32] = s2_done ? 1.0f : 0.0f;
      obs_put(obs, c.off(DEX_K_CURRENT_STAGE), 0, (float)stage);
      obs_put(obs, c.off(DEX_K_TIME_IN_STAGE), 0, tis);
      obs_put(obs, c.off(DEX_K_STAGE_PROGRESS), 0, prog);
    } else if (c.cw == 2) {
      // ---- minima, criteria inputs (step_processor.py:133-181, blind_grasping_task.py:1290-1364) ----
      float min_tip_z = c.template pose<LM>(A, 0)[2], min_tip_box = wk[LY::W_TIPBOX * 32];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) {
        min_tip_z = fminf(min_tip_z, c.template pose<LM>(A, f)[2]);
        min_tip_box = fminf(min_tip_box, wk[(LY::W_TIPBOX + f) * 32]);
      }
      const bool hitg = P.has_height_safety && ((rb[2] < P.handbase_threshold) || (min_tip_z < P.fingertip_threshold));
      const bool h_ok = (wk[(LY::W_HDIFF + 0) * 32] <= P.height_tolerance) && (wk[(LY::W_HDIFF + 1) * 32] <= P.height_tolerance) &&
                        (wk[(LY::W_HDIFF + 2) * 32] <= P.height_tolerance);
      const bool c_ok = wk[LY::W_CDIST * 32] <= P.centroid_tolerance;
      const bool s_ok = (wk[LY::W_DRIFT * 32] <= P.position_drift_tolerance) && (wk[LY::W_VMAG * 32] <= P.velocity_tolerance);
      wk[LY::W_MINTIPBOX * 32] = min_tip_box;
      wk[LY::W_HITG * 32] = hitg ? 1.0f : 0.0f;
      wk[LY::W_BOXFAR * 32] = (wk[LY::W_HANDBOX * 32] > P.max_box_distance) ? 1.0f : 0.0f;
      wk[LY::W_QUALOK * 32] = (h_ok && c_ok && s_ok) ? 1.0f : 0.0f;
    } else if (c.cw == 3) {
      // ---- averages (blind_grasping_task.py:702-724), episode time (dexhand_base.py:703-712: PRE-increment count) ----
      float dsum = 0.0f, hsum = 0.0f;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        dsum += wk[(LY::W_PADBOX + f) * 32];
        hsum += wk[(LY::W_HDIFF + f) * 32];
      }
      const float avg_h = div_rn(hsum, 5.0f);
      wk[LY::W_AVGHDIFF * 32] = avg_h;
      obs_put(obs, c.off(DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE), 0, div_rn(dsum, 5.0f));
      obs_put(obs, c.off(DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF), 0, avg_h);
      obs_put(obs, c.off(DEX_K_EPISODE_TIME), 0, (float)c.i64lo(LY::I_STEP) * dt);
    }
  } else {
    if (c.cw == 2) {
      float min_tip_z = c.template pose<LM>(A, 0)[2];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) min_tip_z = fminf(min_tip_z, c.template pose<LM>(A, f)[2]);
      const bool hitg = P.has_height_safety && ((rb[2] < P.handbase_threshold) || (min_tip_z < P.fingertip_threshold));
      wk[LY::W_HITG * 32] = hitg ? 1.0f : 0.0f;
      obs_put(obs, c.off(DEX_K_EPISODE_TIME), 0, (float)c.i64lo(LY::I_STEP) * dt);
    } else {
      // common reward terms over the seven other warps (0, 1, 3..7)
      for (int i = (c.cw < 2 ? c.cw : c.cw - 1); i < DEX_NUM_COMMON_TERMS; i += kConsWarps - 1) {
        emit_term<BG, STD, LM>(A, c, i, common_term<BG, STD, LM>(A, c, i));
      }
    }
  }
}

// phase C, warp 0: termination (step_processor.py:133-181, termination_manager.py:120-278), masks, tile record.
// Returns the exclusive termination class of this lane's env: bit 0 success, 1 failure, 2 timeout.
template <bool BG, bool STD, int LM>
__device__ __forceinline__ unsigned phase_combine(const Post2Args& A, const Cx<BG, STD>& c, const unsigned stamp, unsigned* s_rbits) {
  using LY = Lay<BG, LM == 1>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float* wk = c.wk;
  bool should_reset = false, t_success = false, t_failure = false, t_timeout = false;
  int ep_len_done = 0;
  bool crit[DEX_NUM_CRITERIA];
#pragma unroll
  for (int k = 0; k < DEX_NUM_CRITERIA; ++k) crit[k] = false;
  if (c.act) {
    const int step_now = c.i64lo(LY::I_STEP) + 1;  // step_processor.py:80
    st.episode_step_count[c.env] = (long long)step_now;
    ep_len_done = step_now;
    crit[DEX_C_HITTING_GROUND] = wk[LY::W_HITG * 32] != 0.0f;
    if (BG) {
      const bool grasp = wk[LY::W_GRASP * 32] != 0.0f;
      const bool j2 = wk[LY::W_J2 * 32] != 0.0f, j3 = wk[LY::W_J3 * 32] != 0.0f;
      const int stage = __float_as_int(wk[LY::W_STAGE * 32]);
      const int sds = __float_as_int(wk[LY::W_SDS * 32]);
      crit[DEX_C_GRASP_LIFT_SUCCESS] = sds >= P.contact_duration_threshold_steps;
      crit[DEX_C_BOX_TOO_FAR] = wk[LY::W_BOXFAR * 32] != 0.0f;
      const bool pregrasp_failed = j2 && (wk[LY::W_QUALOK * 32] == 0.0f);  // _evaluate_pregrasp_quality :1334-1364
      const bool contact_failed = j3 && !grasp;
      const bool grasp_lost = (stage == 3) && !grasp;
      crit[DEX_C_STAGE1_PREGRASP_FAILED] = pregrasp_failed;
      crit[DEX_C_STAGE2_CONTACT_FAILED] = contact_failed;
      crit[DEX_C_STAGE3_GRASP_LOST] = grasp_lost;
      st.failure_tensors[0 * c.N + c.env] = pregrasp_failed ? 1 : 0;
      st.failure_tensors[1 * c.N + c.env] = contact_failed ? 1 : 0;
      st.failure_tensors[2 * c.N + c.env] = grasp_lost ? 1 : 0;
    }
    bool ep_s = false, ep_f = false;
#pragma unroll
    for (int k = 0; k < DEX_NUM_CRITERIA; ++k) {
      crit[k] = crit[k] && (P.criterion_active[k] != 0);
      if (k == DEX_C_GRASP_LIFT_SUCCESS)
        ep_s = ep_s || crit[k];
      else
        ep_f = ep_f || crit[k];
    }
    const bool timeout = step_now >= (P.max_episode_length - 1);
    should_reset = ep_s || ep_f || timeout;
    t_success = ep_s;
    t_failure = ep_f && !ep_s;
    t_timeout = timeout && !ep_s && !ep_f;
    A.out.reset_buf[c.env] = should_reset ? 1 : 0;
    A.out.success[c.env] = t_success ? 1 : 0;
    A.out.failure[c.env] = t_failure ? 1 : 0;
    A.out.timeout[c.env] = t_timeout ? 1 : 0;
  }
  // reset ballot word + cross-env counters (termination_manager.py:164-185,259-266)
  const unsigned fullmask = 0xffffffffu;
  const unsigned rbits = __ballot_sync(fullmask, should_reset);
  const int lane = c.lane;
  if (lane == 0) {
    A.out.reset_bits[c.e0 >> 5] = rbits;
    s_rbits[0] = rbits;
  }
  const unsigned bs = __ballot_sync(fullmask, t_success), bf = __ballot_sync(fullmask, t_failure),
                 bt = __ballot_sync(fullmask, t_timeout);
  if (lane == 0) {  // termination class of every env of the tile, for the observer flush of the warps that reset them
    s_rbits[1] = bs;
    s_rbits[2] = bf;
  }
  unsigned bc[DEX_NUM_CRITERIA];
#pragma unroll
  for (int k = 0; k < DEX_NUM_CRITERIA; ++k) bc[k] = __ballot_sync(fullmask, crit[k]);
  int len = 0;
  if (rbits) {  // sum of the finished episodes' lengths (warp-uniform branch)
    len = should_reset ? ep_len_done : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) len += __shfl_xor_sync(fullmask, len, o);
  }
  if (lane == 0) {
    if (A.fused_epilogue) {
      // three self-validating 8-byte records per tile (see post_physics.cuh post_scalar): bits 48..63 = step stamp
      const unsigned long long stp = (unsigned long long)stamp << 48;
      unsigned long long r1 = 0;
#pragma unroll
      for (int k = 0; k < DEX_NUM_CRITERIA; ++k) r1 |= (unsigned long long)__popc(bc[k]) << (6 * k);
      const size_t tiles = (size_t)A.tiles;
      unsigned long long* rec = reinterpret_cast<unsigned long long*>(A.ep.tile_records) + (size_t)(c.e0 >> 5);
      __stcg(rec, stp | rbits | ((unsigned long long)__popc(bs) << 32) | ((unsigned long long)__popc(bf) << 38));
      __stcg(rec + tiles, stp | r1);
      __stcg(rec + 2 * tiles, stp | (unsigned long long)(unsigned)len);
    } else {
      if (bs) atomicAdd(A.out.counters + DEX_CNT_SUCCESS, __popc(bs));
      if (bf) atomicAdd(A.out.counters + DEX_CNT_FAILURE, __popc(bf));
      if (bt) atomicAdd(A.out.counters + DEX_CNT_TIMEOUT, __popc(bt));
      if (rbits) atomicAdd(A.out.counters + DEX_CNT_RESET, __popc(rbits));
#pragma unroll
      for (int k = 0; k < DEX_NUM_CRITERIA; ++k)
        if (bc[k]) atomicAdd(A.out.counters + DEX_CNT_CRIT0 + k, __popc(bc[k]));
      if (rbits) atomicAdd(A.out.counters + DEX_CNT_EPISODE_LENGTH, len);
    }
  }
  return (t_success ? 1u : 0u) | (t_failure ? 2u : 0u) | (t_timeout ? 4u : 0u);
}

// phase D, warp 0: ordered weighted sum (reward_calculator.py:223-275) + termination rewards
// (termination_manager.py:280-321, step_processor.py:204-219)
template <bool BG, bool STD, int LM, bool OBS>
__device__ __forceinline__ void phase_total(const Post2Args& A, const Cx<BG, STD>& c, const unsigned cls) {
  using LY = Lay<BG, LM == 1>;
  if (!c.act) return;
  const DexParams& P = A.p;
  const float* wk = c.wk;
  // observer on: the running sums of the seven rows this warp produces and the episode length are requested first, so
  // that their (L2) latency runs under the ordered sum below instead of behind it, one round trip after the other
  constexpr bool observer = OBS;  // a separate kernel instance: even dormant observer code costs the lean path 1-4 %
  float es7[7];
  int ep_len = 0;
  if constexpr (observer) {
    const float* es = A.out.episode_sums + c.env;
    es7[0] = __ldcg(es + (size_t)P.total_row * c.N);
#pragma unroll
    for (int k = 0; k < 6; ++k) es7[1 + k] = __ldcg(es + (size_t)(P.termination_row + k) * c.N);
    ep_len = __ldcg(A.out.episode_lengths + c.env);
  }
  // STD: exactly the task's own terms are materialised (BlindGrasping: the 16 task terms, BaseTask: the 10 common ones)
  constexpr int I0 = (STD && BG) ? DEX_NUM_COMMON_TERMS : 0, I1 = BG ? DEX_NUM_REWARD_TERMS : DEX_NUM_COMMON_TERMS;
  float total = 0.0f;
#pragma unroll
  for (int i = I0; i < I1; ++i) {
    if (STD || P.reward_weight[i] != 0.0f) total += wk[(LY::W_WGT + i) * 32];  // dict order: common, then task
  }
  const bool t_success = cls & 1u, t_failure = cls & 2u, t_timeout = cls & 4u;
  const float rs = t_success ? P.success_reward : 0.0f;
  const float rf = t_failure ? P.failure_reward : 0.0f;
  const float rt = t_timeout ? P.timeout_reward : 0.0f;
  float rew = total;
  rew += rs;
  rew += rf;
  rew += rt;
  float* g = A.out.reward_components + c.env;
  const float tv[7] = {total, t_success ? 1.0f : 0.0f, rs, t_failure ? 1.0f : 0.0f, rf, t_timeout ? 1.0f : 0.0f, rt};
  g[(size_t)P.total_row * c.N] = tv[0];
#pragma unroll
  for (int k = 0; k < 6; ++k) g[(size_t)(P.termination_row + k) * c.N] = tv[1 + k];
  A.out.rew_buf[c.env] = rew;
  // ---- fused RewardComponentObserver (rl/reward_observer.py:160-293): the rows this warp produces (total + the six
  //      termination rows), the episode length and the episode counters; the term rows are warps 1-7's (observer_term_rows)
  if constexpr (observer) {
    const bool done = cls != 0u;
    const int stride = 2 + P.num_component_rows;
    const int kc = t_success ? 1 : (t_failure ? 2 : 3);
    double* tot = A.out.stat_totals + (size_t)((c.e0 >> 5) % max(1, A.out.stat_slices)) * 4 * stride;  // block of this tile
    const int len = ep_len + 1;
    A.out.episode_lengths[c.env] = done ? 0 : len;
    if (done) {
      atomicAdd(tot + 0, 1.0);
      atomicAdd(tot + 1, (double)len);
      atomicAdd(tot + kc * stride + 0, 1.0);
      atomicAdd(tot + kc * stride + 1, (double)len);
    }
    float* es = A.out.episode_sums + c.env;
#pragma unroll
    for (int k = 0; k < 7; ++k) {
      const int row = k == 0 ? P.total_row : P.termination_row + (k - 1);
      const float sum = es7[k] + tv[k];
      if (done) {
        atomicAdd(tot + 2 + row, (double)sum);
        atomicAdd(tot + kc * stride + 2 + row, (double)sum);
      }
      es[(size_t)row * c.N] = done ? 0.0f : sum;
    }
  }
}

// The observer's running sums (episode_sums [rows][N], one read-modify-write per row, env and step) are touched in the
// middle of the per-tile program, where a DRAM miss cannot hide behind anything: one warp pulls the tile's 128-byte row
// segments into L2 while the tile's inputs are still on their way (prefetch.global.L2 changes no state: always safe).
__device__ __forceinline__ void observer_prefetch(const Post2Args& A, const int e0, const int lane) {
  if (A.out.episode_sums == nullptr) return;
  const int rows = A.p.num_component_rows;
  for (int r = lane; r < rows; r += 32)
    asm volatile("prefetch.global.L2 [%0];" ::"l"(A.out.episode_sums + (size_t)r * A.n + e0));
  if (lane == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(A.out.episode_lengths + e0));
}

// The observer's running sums of the TERM rows (rows below total_row: raw / weighted value of every materialised reward
// term; rl/reward_observer.py:176-188 `sums += component`, :190-230 flush at episode end), lane = env, rows dealt over
// warps w = 0 .. nw-1.  Runs behind barrier C: every term row of the tile is in reward_components by then (read back
// with coherent loads - other warps of this CTA wrote them) and the tile's reset / class ballots are in s_rbits.  All
// loads of a batch of four rows are in flight together: one L2 round trip per batch instead of one per row.
template <bool BG, bool STD>
__device__ __forceinline__ void observer_term_rows(const Post2Args& A, const Cx<BG, STD>& c, const unsigned* s_rbits, const int w,
                                                   const int nw) {
  if (!c.act) return;
  const DexParams& P = A.p;
  const int nrow = P.total_row;
  const bool done = (s_rbits[0] >> c.lane) & 1u;
  const int kc = ((s_rbits[1] >> c.lane) & 1u) ? 1 : (((s_rbits[2] >> c.lane) & 1u) ? 2 : 3);
  const int stride = 2 + P.num_component_rows;
  double* tot = A.out.stat_totals + (size_t)((c.e0 >> 5) % max(1, A.out.stat_slices)) * 4 * stride;  // block of this tile
  const float* rc = A.out.reward_components + c.env;
  float* es = A.out.episode_sums + c.env;
#pragma unroll 1
  for (int r0 = w; r0 < nrow; r0 += 4 * nw) {
    float a[4], b[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = r0 + j * nw;
      if (r < nrow) {
        a[j] = __ldcg(rc + (size_t)r * c.N);
        b[j] = __ldcg(es + (size_t)r * c.N);
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = r0 + j * nw;
      if (r < nrow) {
        const float sum = b[j] + a[j];
        if (done) {
          atomicAdd(tot + 2 + r, (double)sum);
          atomicAdd(tot + kc * stride + 2 + r, (double)sum);
        }
        es[(size_t)r * c.N] = done ? 0.0f : sum;
      }
    }
  }
}


// ------------------------------------------------------------------------------------------------
// tile loads, shared by both geometries: NW warps issue asynchronous copies of one tile into `stage`; every thread
// ends with one counted cp.async arrival on the stage's mbarrier, every warp with one plain arrival, and warp 0 / lane 0
// carries the byte count of the bulk copies.
// ------------------------------------------------------------------------------------------------
template <bool BG>
struct LoadCtx {
  int rbo0, rbo1, rbo2, xrow;
  bool p_rb0, p_rb2, p_x;
  const float* x_base;
  __device__ __forceinline__ void setup(const Post2Args& A, const int* s_rbo, const int lane) {
    rbo0 = s_rbo[lane];
    rbo1 = s_rbo[lane + 32];
    rbo2 = s_rbo[min(lane + 64, kRbW - 1)];
    p_rb0 = lane < 7 || lane >= 13 || A.common;  // hand velocities feed only the common reward terms
    p_rb2 = lane + 64 < kRbW;
    const bool is_box = BG && lane >= 18 && lane < 24;
    p_x = is_box || lane < (BG ? 18 : 15);
    xrow = is_box ? A.p.num_actors * 13 : A.p.num_contact_bodies * 3;
    int xoff;
    if (is_box) {
      const int k = lane - 18;
      xoff = A.p.box_actor * 13 + (k < 3 ? k : k + 4);
    } else {
      const int f = min(lane / 3, 5);
      xoff = A.cf_off[f] + (lane - 3 * f);
    }
    x_base = (is_box ? (const float*)A.sim.actor_root_state : A.sim.net_contact_force) + xoff;
  }
};

template <bool BG, int NW>
__device__ __forceinline__ void issue_tile_loads(const Post2Args& A, const LoadCtx<BG>& lc, unsigned char* stage, const uint32_t fb,
                                                 const int tile, const int pw, const int lane, const bool first) {
  using LY = Lay<BG>;
  const DexStateBuffers& st = A.st;
  const size_t N = (size_t)A.n;
  const int e0 = tile * 32, ne = min(32, A.n - e0);
  const uint32_t sst = s32(stage);
  const bool full = ne == 32;           // CTA-uniform
  const bool bulk = A.bulk_ok && full;  // the AoS tiles (dof_state, full_dof_targets) as ONE bulk copy each
  const bool early = (A.flags & DEX_FLAG_EARLY_SIM_LOADS) != 0;
  const bool split = first && early;    // first tile: the state rows wait for griddepcontrol.wait, the sim rows do not
  const bool act = lane < ne;
  if (first && !early) pdl_wait();
  // a contiguous AoS tile: one bulk copy, or (ragged tile / unaligned base) 4-byte asynchronous copies
  auto aos_tile = [&](const float* base, const int per, const int off) {
    const float* src = base + (size_t)e0 * per;
    if (bulk) {
      if (pw == 0 && lane == 0) bulk_g2s(sst + off, src, (uint32_t)(32 * per * 4), fb);
    } else {
#pragma unroll 1
      for (int i = pw * 32 + lane; i < ne * per; i += NW * 32) cpa4(sst + off + 4u * i, src + i);
    }
  };
  aos_tile(A.sim.dof_state, 52, LY::OFF_DOF);
  // ---- scattered Isaac Gym rows: warp pw takes env rows pw, pw+NW, ..; consecutive lanes = consecutive floats ----
  {
    const size_t nb13 = (size_t)A.rb_pitch;
    const float* q0 = A.sim.rigid_body_state + (size_t)(e0 + pw) * nb13;
    const float* qx = lc.x_base + (size_t)(e0 + pw) * lc.xrow;
    uint32_t a = sst + LY::OFF_AOS + (uint32_t)(pw * LY::SA + lane) * 4u;
#pragma unroll 1
    for (int i = 0; i < 32 / NW; ++i) {
      if (full || pw + NW * i < ne) {
        if (lc.p_rb0) cpa4(a + LY::A_RB * 4, q0 + lc.rbo0);
        cpa4(a + (LY::A_RB + 32) * 4, q0 + lc.rbo1);
        if (lc.p_rb2) cpa4(a + (LY::A_RB + 64) * 4, q0 + lc.rbo2);
        if (lc.p_x) cpa4(a + LY::A_CF * 4, qx);
      }
      q0 += NW * nb13;
      qx += (size_t)NW * lc.xrow;
      a += NW * LY::SA * 4;
    }
  }
  if (split) pdl_wait();  // everything above touched only kernel parameters and the sim tensors
  if (!BG && A.want_tgt) aos_tile(st.full_dof_targets, 26, LY::OFF_TGT);
  // ---- SoA state rows: lane = env, one 128-byte line per row; row r of a buffer goes to warp r % NW ----
  auto rows = [&](const float* base, const int nrows, const int r0) {
    const float* src = base + (size_t)pw * N + e0 + lane;
    uint32_t dst = sst + (uint32_t)(r0 + pw) * 128u + 4u * lane;
#pragma unroll 1
    for (int r = pw; r < nrows; r += NW) {
      if (act) cpa4(dst, src);
      src += NW * N;
      dst += NW * 128u;
    }
  };
  auto row64 = [&](const int64_t* base, const int slot) {
    if (act) cpa8(sst + LY::OFF_I64 + slot * 256 + 8u * lane, base + e0 + lane);
  };
  rows(st.prev_dof_pos, DEX_NUM_DOF, LY::R_PDP);
  rows(st.active_prev_targets, DEX_NUM_ACTIVE, LY::R_APT);
  rows(st.prev_actions, A.p.num_actions, LY::R_PACT);
  rows(st.contact_duration_steps, DEX_NUM_FINGERS, LY::R_CDUR);
  rows(st.prev_contact_binary, DEX_NUM_FINGERS, LY::R_PBIN);
  if (BG) {
    rows(st.initial_box_positions, 3, LY::R_BOX0);
    if (pw == 3 && act) cpa4(sst + LY::R_TIS * 128u + 4u * lane, st.time_in_stage + e0 + lane);  // single rows
    if (pw == 0 && act) cpa4(sst + LY::R_SCD * 128u + 4u * lane, st.stage_contact_duration + e0 + lane);
    if (pw == 1) row64(st.success_duration_steps, LY::I_SDS);
    if (pw == 2) row64(st.current_stage, LY::I_STAGE);
  } else if (A.common && !A.first) {
    rows(st.prev_finger_dof_vel, DEX_NUM_FINGER_DOF, LY::R_PFDV);
    rows(st.prev_hand_vel, 3, LY::R_PHV);
    rows(st.prev_hand_ang_vel, 3, LY::R_PHA);
    for (int f = pw; f < DEX_NUM_FINGERS; f += NW) {  // prev_contacts: 5 bool rows of 32 bytes
      const unsigned char* src = st.prev_contacts + (size_t)f * N + e0;
      if (full && (A.n & 3) == 0) {
        if (lane < 8) cpa4(sst + LY::OFF_U8 + f * 32 + 4u * lane, src + 4 * lane);
      } else if (act) {
        stage[LY::OFF_U8 + f * 32 + lane] = src[lane];
      }
    }
  }
  if (pw == 3) row64(st.episode_step_count, LY::I_STEP);
  cpa_arrive(fb);  // counted arrival when every asynchronous copy of this thread has landed
  __syncwarp();
  if (lane == 0) {
    if (pw == 0 && bulk)
      mbar_arrive_tx(fb, A.tx_bytes);
    else
      mbar_arrive(fb);
  }
}

// Register-staged tile load for the single-pass geometry (8 warps, nothing to overlap with: plain LDG costs the LSU a
// quarter of an LDGSTS).  Every load of the tile is in flight before the first shared store; the dof_state tile still
// comes as one bulk copy on `fb` (thread 0 carries its byte count).  Ends with the stores issued, not yet synchronised.
template <bool BG>
__device__ __forceinline__ void load_tile_regs(const Post2Args& A, const LoadCtx<BG>& lc, unsigned char* stage, const uint32_t fb,
                                               const int tile, const int w, const int lane) {
  using LY = Lay<BG>;
  static_assert(BG, "single-pass geometry: BlindGrasping stage layout");
  constexpr int NW = kConsWarps;
  const DexStateBuffers& st = A.st;
  const size_t N = (size_t)A.n;
  const int e0 = tile * 32, ne = min(32, A.n - e0);
  const uint32_t sst = s32(stage);
  const bool full = ne == 32;
  const bool bulk = A.bulk_ok && full;
  const bool early = (A.flags & DEX_FLAG_EARLY_SIM_LOADS) != 0;
  const bool act = lane < ne;
  if (!early) pdl_wait();
  // dof_state tile: contiguous 32 x 208 B -> 416 float4, two coalesced 128-bit loads per thread (16-byte aligned: check_sim)
  float4 d4[2];
  const float4* dsrc = reinterpret_cast<const float4*>(A.sim.dof_state + (size_t)e0 * 52);
  const int tq = w * 32 + lane, nq = ne * 13;
#pragma unroll
  for (int j = 0; j < 2; ++j)
    if (tq + 256 * j < nq) d4[j] = dsrc[tq + 256 * j];
  (void)bulk;
  (void)sst;
  (void)fb;
  // scattered Isaac Gym rows: warp w takes env rows w, w+8, w+16, w+24
  float g[4][4];
  {
    const size_t nb13 = (size_t)A.rb_pitch;
    const float* q0 = A.sim.rigid_body_state + (size_t)(e0 + w) * nb13;
    const float* qx = lc.x_base + (size_t)(e0 + w) * lc.xrow;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (full || w + NW * i < ne) {
        if (lc.p_rb0) g[i][0] = __ldg(q0 + lc.rbo0);
        g[i][1] = __ldg(q0 + lc.rbo1);
        if (lc.p_rb2) g[i][2] = __ldg(q0 + lc.rbo2);
        if (lc.p_x) g[i][3] = *qx;
      }
      q0 += NW * nb13;
      qx += (size_t)NW * lc.xrow;
    }
  }
  if (early) pdl_wait();  // everything above touched only kernel parameters and the sim tensors
  // SoA state rows: lane = env; row r of a buffer goes to warp r % 8
  const size_t col = (size_t)e0 + (act ? lane : 0);
  float r_pdp[4], r_apt[3], r_pact[3], r_one[4];
  long long r_i64 = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (w + NW * j < DEX_NUM_DOF) r_pdp[j] = st.prev_dof_pos[(size_t)(w + NW * j) * N + col];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    if (w + NW * j < DEX_NUM_ACTIVE) r_apt[j] = st.active_prev_targets[(size_t)(w + NW * j) * N + col];
    if (w + NW * j < A.p.num_actions) r_pact[j] = st.prev_actions[(size_t)(w + NW * j) * N + col];
  }
  if (w < DEX_NUM_FINGERS) {
    r_one[0] = st.contact_duration_steps[(size_t)w * N + col];
    r_one[1] = st.prev_contact_binary[(size_t)w * N + col];
  }
  if (w < 3) r_one[2] = st.initial_box_positions[(size_t)w * N + col];
  if (w == 3) r_one[3] = st.time_in_stage[col];
  if (w == 4) r_one[3] = st.stage_contact_duration[col];
  if (w == 5) r_i64 = st.success_duration_steps[col];
  if (w == 6) r_i64 = st.current_stage[col];
  if (w == 7) r_i64 = st.episode_step_count[col];
  // ---- stores ----
  {
    float4* ddst = reinterpret_cast<float4*>(stage + LY::OFF_DOF);
#pragma unroll
    for (int j = 0; j < 2; ++j)
      if (tq + 256 * j < nq) ddst[tq + 256 * j] = d4[j];
  }
  {
    float* a = reinterpret_cast<float*>(stage + LY::OFF_AOS) + w * LY::SA + lane;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (full || w + NW * i < ne) {
        if (lc.p_rb0) a[LY::A_RB] = g[i][0];
        a[LY::A_RB + 32] = g[i][1];
        if (lc.p_rb2) a[LY::A_RB + 64] = g[i][2];
        if (lc.p_x) a[LY::A_CF] = g[i][3];
      }
      a += NW * LY::SA;
    }
  }
  float* rows = reinterpret_cast<float*>(stage) + lane;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    if (w + NW * j < DEX_NUM_DOF) rows[(LY::R_PDP + w + NW * j) * 32] = r_pdp[j];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    if (w + NW * j < DEX_NUM_ACTIVE) rows[(LY::R_APT + w + NW * j) * 32] = r_apt[j];
    if (w + NW * j < A.p.num_actions) rows[(LY::R_PACT + w + NW * j) * 32] = r_pact[j];
  }
  if (w < DEX_NUM_FINGERS) {
    rows[(LY::R_CDUR + w) * 32] = r_one[0];
    rows[(LY::R_PBIN + w) * 32] = r_one[1];
  }
  if (w < 3) rows[(LY::R_BOX0 + w) * 32] = r_one[2];
  if (w == 3) rows[LY::R_TIS * 32] = r_one[3];
  if (w == 4) rows[LY::R_SCD * 32] = r_one[3];
  long long* i64 = reinterpret_cast<long long*>(stage + LY::OFF_I64) + lane;
  if (w == 5) i64[LY::I_SDS * 32] = r_i64;
  if (w == 6) i64[LY::I_STAGE * 32] = r_i64;
  if (w == 7) i64[LY::I_STEP * 32] = r_i64;
}

template <int NT = kConsThreads>
__device__ __forceinline__ void cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory"); }

// shared -> global bulk store of the flat obs tile (TMA engine): one instruction for 20 KB
__device__ __forceinline__ void bulk_s2g(void* dst, const uint32_t src, const uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// obs tile -> obs_buf with row copies (ragged tiles, layouts whose flat tile would bank-conflict): the env rows are spread
// over the 8 consumer warps, lane = column: every global store is one full 128-byte request
__device__ __forceinline__ void copy_obs_rows(const float* s_obs, float* g_tile, const int W, const int WP, const int ne,
                                              const int w, const int lane) {
  const int nchunk = (W + 31) >> 5;
#pragma unroll 1
  for (int ch = 0; ch < nchunk; ++ch) {
    const int col = 32 * ch + lane;
    if (col < W) {
      float v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = w + kConsWarps * i;
        if (row < ne) v[i] = s_obs[row * WP + col];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = w + kConsWarps * i;
        if (row < ne) g_tile[(size_t)row * W + col] = v[i];
      }
    }
  }
}

// optional policy-input stage, from the same shared obs tile: obs_norm = clamp((obs - mean) * scale, +-clip),
// the frozen-statistics form of rl_games' RunningMeanStd.forward (normalize_input, cfg/train/BaseTaskPPO.yaml:44); fp32
// subtract, multiply by the caller's fp32 reciprocal, clamp that lets NaN through like torch.clamp.  Row-per-warp,
// lane = column, 128-byte stores.
__device__ __forceinline__ void copy_obs_norm_rows(const float* s_obs, const DexOutBuffers& out, const int e0, const int W, const int WP,
                                                   const int ne, const int w, const int lane) {
  const int nchunk = (W + 31) >> 5;
  const float clip = out.norm_clip;
  float* g_tile = out.obs_norm + (size_t)e0 * W;
#pragma unroll 1
  for (int ch = 0; ch < nchunk; ++ch) {
    const int col = 32 * ch + lane;
    if (col < W) {
      const float m = __ldg(out.norm_mean + col), sc = __ldg(out.norm_scale + col);
      float v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = w + kConsWarps * i;
        if (row < ne) v[i] = s_obs[row * WP + col];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int row = w + kConsWarps * i;
        if (row < ne) {
          const float y = (v[i] - m) * sc;
          g_tile[(size_t)row * W + col] = (y != y) ? y : fminf(fmaxf(y, -clip), clip);
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// the kernel: persistent, 8 consumer + 4 producer warps per CTA, 2-stage ring, two CTAs per SM (BlindGrasping).
// Instruction supply is the scarce resource of this path (profiles/r02_variants.md): the consumer warps run the SAME
// item code on different items at the same time (one fetched line serves all), loops are not unrolled, and a warp that
// waits blocks in hardware (bar.sync behind one polling warp) instead of spinning through fetch slots.
// ------------------------------------------------------------------------------------------------
template <bool BG, bool STD, bool TMA = false, bool OBS = false, int G = 1>
__global__ void __launch_bounds__(G * kThreads, (BG && G == 1) ? 2 : 1) post_physics_v2_kernel(const __grid_constant__ Post2Args A) {
  using LY = Lay<BG>;
  // G tiles per pass: G groups of 8 consumer warps run the per-tile program on G tiles IN LOCKSTEP (they share every barrier),
  // so one pass of the instruction stream through the SM's front end serves G tiles; G groups of 4 producer warps fill the
  // G sub-stages of a ring stage.  G = 2 is the "pair" geometry (one CTA per SM), G = 1 the plain persistent kernel.
  constexpr int CW = G * kConsWarps, CT = CW * 32;
  static_assert(G == 1 || (BG && !TMA), "pair geometry: BlindGrasping, LDGSTS producers");
  static_assert(BG || !TMA, "TMA staging: BlindGrasping stage layout");
  constexpr int LM = TMA ? 2 : 0;
  const uint32_t stage_bytes = TMA ? A.stage_bytes : (uint32_t)LY::STAGE_BYTES;
  extern __shared__ __align__(128) unsigned char smem2[];
  __shared__ __align__(8) unsigned long long s_bar[2 * kStages];  // full[0..], empty[0..]
  __shared__ float s_reset_u[32 * CW];
  __shared__ int s_rbo[kRbW + 1];  // gathered rigid slot -> float offset inside one env's [NB,13] block
  __shared__ unsigned long long s_rng_step;
  __shared__ unsigned s_rbits_all[G][4];  // ballots of this tile: reset | success | failure
  __shared__ int s_epi_warp[kConsWarps];
  __shared__ int s_epi_base;
  __shared__ int s_epi_cnt[DEX_NUM_COUNTERS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- the epilogue CTA (last block): polls the tile records of the workers, all of which are co-resident ----
  if (A.fused_epilogue && (int)blockIdx.x == A.workers) {
    pdl_wait();
    pdl_launch_dependents();
    if (tid == 0) {
      const unsigned long long rs = *A.st.rng_step;
      const unsigned stamp = ((unsigned)rs + 1u) & 0xffffu;
      const volatile unsigned long long* last =
          reinterpret_cast<const volatile unsigned long long*>(A.ep.tile_records) + (size_t)(A.tiles - 1);  // plane 0
      const long long t0 = clock64();
      while ((unsigned)(last[0] >> 48) != stamp) {
        __nanosleep(100);
        if (clock64() - t0 > kSpinLimitClocks) __trap();  // bounded: see mbar_wait
      }
      s_rng_step = rs;
    }
    __syncthreads();
    if (tid < kConsThreads)
      epilogue_records<kConsThreads, 2>(A.ep, A.n, A.p.max_consecutive_successes, ((unsigned)s_rng_step + 1u) & 0xffffu,
                                        s_epi_warp, &s_epi_base, s_epi_cnt, reinterpret_cast<unsigned*>(smem2),
                                        (int)(A.smem_bytes / 4), tid);
    return;
  }

  // ---- setup ----
  const uint32_t bar_full = s32(&s_bar[0]), bar_empty = s32(&s_bar[kStages]);
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) {
      // LDGSTS producers: one cp.async-completion arrival per producer thread + one per warp; TMA: one arrival (expect_tx)
      mbar_init(bar_full + 8 * s, TMA ? 1 : G * (kProdThreads + kProdWarps));
      mbar_init(bar_empty + 8 * s, CW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (tid < kRbW) {  // one slot per thread: the integer division happens once per CTA, not per producer lane
    const int s = tid;
    const int row = s < 13 ? 0 : 1 + (s - 13) / 7;
    s_rbo[s] = A.rb_off[row] + (s < 13 ? s : (s - 13) - (row - 1) * 7);
  }
  __syncthreads();
  pdl_launch_dependents();
  TL2_DECL;
  const int tiles = A.tiles, workers = A.workers;

  if (warp >= CW) {
    // =========================== producer warps ===========================
    const int pw = (warp - CW) % kProdWarps, pg = G == 1 ? 0 : (warp - CW) / kProdWarps;
    if constexpr (TMA) {
      // ONE warp, one copy per lane and tile (build_tma): 2-D tensor boxes for the gathered Isaac Gym rows and the [K][N]
      // state rows, plain bulk copies for the dof tile and the single rows.  The TMA unit serves a CTA's copies in issue
      // order, so the first tile lands while the second is still being fetched - and no load instruction competes with
      // the consumers for issue slots.
      if (pw != 0) return;
      const bool mine = lane < A.t_ncopy;
      const int kind = mine ? A.t_kind[lane] : 3;
      const bool early_copy = mine && A.t_early[lane] != 0 && (A.flags & DEX_FLAG_EARLY_SIM_LOADS) != 0;
      const uint32_t dst_off = mine ? A.t_dst[lane] : 0u, bytes = mine ? A.t_bytes[lane] : 0u;
      const int c0 = mine ? A.t_c0[lane] : 0;
      const CUtensorMap* map = &A.tmap[mine ? A.t_map[lane] : 0];
      const char* gptr = reinterpret_cast<const char*>(mine ? A.t_ptr[lane] : 0ull);
      auto issue = [&](const uint32_t sst, const uint32_t fb, const int e0) {
        if (kind == 3) {
          bulk_g2s(sst + dst_off, gptr + (size_t)(e0 >> 5) * bytes, bytes, fb);
        } else {
          const int x = kind == 2 ? e0 : c0, y = kind == 0 ? e0 : (kind == 1 ? (e0 >> A.t_root_sh) : 0);
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                           sst + dst_off),
                       "l"(map), "r"(x), "r"(y), "r"(fb)
                       : "memory");
        }
      };
      int it = 0;
#pragma unroll 1
      for (int tile = blockIdx.x; tile < tiles; tile += workers, ++it) {
        const int s = it & 1, use = it >> 1;
        const uint32_t sst = s32(smem2 + s * stage_bytes), fb = bar_full + 8 * s;
        TL2(it, 16);
        if (use > 0) mbar_wait(bar_empty + 8 * s, (unsigned)(use - 1) & 1u);
        TL2(it, 17);
        if (lane == 0) mbar_arrive_tx(fb, A.tma_tx);
        __syncwarp();
        if (it == 0) {  // the kernel ahead in the stream may still be writing the path state (and, without the flag, the sim tensors)
          if (early_copy) issue(sst, fb, tile * 32);
          pdl_wait();
          if (mine && !early_copy) issue(sst, fb, tile * 32);
        } else if (mine) {
          issue(sst, fb, tile * 32);
        }
        TL2(it, 19);
      }
      return;
    }
    LoadCtx<BG> lc;
    lc.setup(A, s_rbo, lane);
    int it = 0;
#pragma unroll 1
    for (int tile = blockIdx.x * G + pg; tile < tiles; tile += workers * G, ++it) {
      const int s = it & 1, use = it >> 1;
      unsigned char* stage = smem2 + (s * G + pg) * stage_bytes;
      const uint32_t fb = bar_full + 8 * s;
      TL2(it, 16);
      if (use > 0) mbar_wait(bar_empty + 8 * s, (unsigned)(use - 1) & 1u);
      TL2(it, 17);
      issue_tile_loads<BG, kProdWarps>(A, lc, stage, fb, tile, pw, lane, it == 0);
      TL2(it, 19);
    }
    return;
  }

  // =========================== consumer warps ===========================
  // consumer group g works on tile (pass * workers + block) * G + g; inside a group the warps are numbered 0..7 as before
  const int cg = G == 1 ? 0 : warp / kConsWarps;
  const int wg = G == 1 ? warp : warp % kConsWarps;
  float* const wk_base = reinterpret_cast<float*>(smem2 + kStages * G * stage_bytes) +
                         cg * (LY::W_COUNT * 32 + (32 * A.wp * 4 + 127) / 128 * 32);
  float* const s_obs = wk_base + LY::W_COUNT * 32;
  unsigned* const s_rbits = s_rbits_all[cg];
  Cx<BG, STD> c;
  c.lane = lane;
  c.cw = wg;
  c.N = (size_t)A.n;
  c.wk = wk_base + lane;
  c.obs = s_obs + lane * A.wp;
  c.obs_offset = A.p.obs_offset;
  c.dof2afj = A.dof2afj;
  unsigned stamp = 0;
  uint64_t rng_step = 0;
  bool store_pending = false;
  const int W = A.p.obs_dim;
  int it = 0;
#pragma unroll 1
  for (int tile = blockIdx.x * G + cg; tile < tiles; tile += workers * G, ++it) {  // tiles % G == 0: every group has a tile
    const int s = it & 1, use = it >> 1;
    c.e0 = tile * 32;
    c.ne = min(32, A.n - c.e0);
    c.act = lane < c.ne;
    c.env = (size_t)c.e0 + (c.act ? lane : 0);
    c.stage = smem2 + (s * G + cg) * stage_bytes;
    float* const g_obs = A.out.obs_buf + (size_t)c.e0 * W;
    const bool bulk_out = A.obs_bulk && c.ne == 32;  // CTA-uniform
    TL2(it, 0);
    if constexpr (OBS) {
      if (wg == kConsWarps - 1) {  // observer rows of this tile (first iteration) and of the next one
        if (it == 0) observer_prefetch(A, c.e0, lane);
        if (tile + workers * G < tiles) observer_prefetch(A, (tile + workers * G) * 32, lane);
      }
    }
    // ONE warp polls the stage's mbarrier, the other seven block in hardware behind it
    if (warp == 0) mbar_wait(bar_full + 8 * s, (unsigned)use & 1u);
    // Philox step counter == stamp of this step's tile records: read once per CTA, behind the first full barrier (i.e.
    // behind the producers' griddepcontrol.wait; the epilogue CTA advances it only after every record of this launch is out)
    if (it == 0 && tid == 0) s_rng_step = A.st.rng_step ? *reinterpret_cast<const volatile unsigned long long*>(A.st.rng_step) : A.step;
    cons_sync<CT>();
    TL2(it, 1);
    if (it == 0) {
      rng_step = s_rng_step;
      stamp = ((unsigned)rng_step + 1u) & 0xffffu;
    }
    phase_items<BG, STD, LM>(A, c);
    TL2(it, 2);
    cons_sync<CT>();
    TL2(it, 3);
    phase_scalar<BG, STD, LM>(A, c);
    TL2(it, 4);
    cons_sync<CT>();
    TL2(it, 5);
    unsigned cls = 0;
    if (wg == 0) {
      cls = phase_combine<BG, STD, LM>(A, c, stamp, s_rbits);
    } else if (BG) {
      if (c.act) {
#pragma unroll 1
        for (int i = DEX_NUM_COMMON_TERMS + (wg - 1); i < DEX_NUM_REWARD_TERMS; i += kConsWarps - 1)
          emit_term<BG, STD, LM>(A, c, i, task_term<BG, STD, LM>(A, c, i));
      }
    }
    if (bulk_out) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // obs tile writes -> visible to the bulk store
    __syncwarp();
    TL2(it, 6);
    cons_sync<CT>();
    TL2(it, 7);
    // the raw stage is not read after this point: hand it back to the producers
    if (lane == 0) mbar_arrive(bar_empty + 8 * s);
    if (wg == 0) phase_total<BG, STD, LM, OBS>(A, c, cls);
    else if constexpr (OBS) observer_term_rows<BG, STD>(A, c, s_rbits, wg - 1, kConsWarps - 1);
    TL2(it, 8);
    if (bulk_out) {
      if (wg == 1 && lane == 0) {
        bulk_s2g(g_obs, s32(s_obs), (uint32_t)(32 * W * 4));
        store_pending = true;
      }
    } else {
      copy_obs_rows(s_obs, g_obs, W, A.wp, c.ne, wg, lane);
    }
    if (A.out.obs_norm != nullptr) copy_obs_norm_rows(s_obs, A.out, c.e0, W, A.wp, c.ne, wg, lane);
    TL2(it, 9);
    {
      unsigned mask = s_rbits[0];  // written by warp 0 before the barrier above
      // finished envs: one warp per env flushes the observer's term rows and applies the fused reset (dexhand_base.py:743-756)
      const bool do_reset = !(A.flags & DEX_FLAG_NO_FUSED_RESET);
      if (mask && do_reset) {
        for (int k = 0; mask; ++k) {
          const int e = __ffs(mask) - 1;
          mask &= mask - 1;
          if ((k & (kConsWarps - 1)) == wg) {
            reset_env_warp<BG>(A, c.e0 + e, rng_step, s_reset_u + 32 * warp);
          }
        }
      }
    }
    if (bulk_out && wg == 1 && lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // obs tile may be rewritten
    TL2(it, 10);
    cons_sync<CT>();  // work block, obs tile and s_rbits are reused by the next tile
    TL2(it, 11);
  }
  if (store_pending) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the last obs tile has left shared memory
}

// ------------------------------------------------------------------------------------------------
// single-pass geometry: ONE tile per CTA of 8 warps, four CTAs per SM (64 registers), no producer warps and no ring - every
// warp issues its share of the tile's asynchronous copies, then all run the same four phases.  Chosen when the launch
// has more tiles than the persistent grid has CTAs but no more than 4 x SMs (e.g. 16 384 envs = 512 tiles on 148 SMs):
// there the persistent kernel would run two full tile rounds on most CTAs, this one runs a single round.
// The work block is folded into the stage (Lay<BG, true>) to fit four CTAs into the SM's shared memory.
// ------------------------------------------------------------------------------------------------
constexpr int kOneThreads = kConsThreads;

template <bool BG, bool STD, bool OBS = false>
__global__ void __launch_bounds__(kOneThreads, 4) post_physics_v2s_kernel(const __grid_constant__ Post2Args A) {
  using LY = Lay<BG, true>;
  static_assert(BG, "the aliased work block covers the BlindGrasping fields only");
  extern __shared__ __align__(128) unsigned char smem2[];
  __shared__ __align__(8) unsigned long long s_bar[1];
  __shared__ float s_reset_u[32 * kConsWarps];
  __shared__ int s_rbo[kRbW + 1];
  __shared__ unsigned long long s_rng_step;
  __shared__ unsigned s_rbits[4];  // ballots of this tile: reset | success | failure
  __shared__ int s_epi_warp[kConsWarps];
  __shared__ int s_epi_base;
  __shared__ int s_epi_cnt[DEX_NUM_COUNTERS];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tile = blockIdx.x;

  if (tile >= A.tiles) {  // the epilogue CTA (fused launches carry one extra block)
    pdl_wait();
    pdl_launch_dependents();
    if (tid == 0) {
      const unsigned long long rs = *A.st.rng_step;
      const unsigned stamp = ((unsigned)rs + 1u) & 0xffffu;
      const volatile unsigned long long* last =
          reinterpret_cast<const volatile unsigned long long*>(A.ep.tile_records) + (size_t)(A.tiles - 1);  // plane 0
      const long long t0 = clock64();
      while ((unsigned)(last[0] >> 48) != stamp) {
        __nanosleep(100);
        if (clock64() - t0 > kSpinLimitClocks) __trap();
      }
      s_rng_step = rs;
    }
    __syncthreads();
    epilogue_records<kOneThreads, 0>(A.ep, A.n, A.p.max_consecutive_successes, ((unsigned)s_rng_step + 1u) & 0xffffu, s_epi_warp,
                                     &s_epi_base, s_epi_cnt, reinterpret_cast<unsigned*>(smem2), (int)(A.smem_bytes / 4), tid);
    return;
  }

  const uint32_t fb = s32(&s_bar[0]);
  if (tid == 0) {
    mbar_init(fb, 1);  // tracks the bulk copy of the dof_state tile only; everything else is register-staged
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (tid < kRbW) {
    const int s = tid;
    const int row = s < 13 ? 0 : 1 + (s - 13) / 7;
    s_rbo[s] = A.rb_off[row] + (s < 13 ? s : (s - 13) - (row - 1) * 7);
  }
  __syncthreads();
  pdl_launch_dependents();
  unsigned char* const stage = smem2;
  float* const s_obs = reinterpret_cast<float*>(smem2 + LY::STAGE_BYTES + LY::NWKA * 128);
  if constexpr (OBS) {
    if (warp == kConsWarps - 1) observer_prefetch(A, tile * 32, lane);
  }
  {
    LoadCtx<BG> lc;
    lc.setup(A, s_rbo, lane);
    load_tile_regs<BG>(A, lc, stage, fb, tile, warp, lane);
  }
  Cx<BG, STD> c;
  c.lane = lane;
  c.cw = warp;
  c.N = (size_t)A.n;
  c.wk = reinterpret_cast<float*>(stage) + lane;  // aliased work block: indexed from the stage base
  c.obs = s_obs + lane * A.wp;
  c.obs_offset = A.p.obs_offset;
  c.dof2afj = A.dof2afj;
  c.e0 = tile * 32;
  c.ne = min(32, A.n - c.e0);
  c.act = lane < c.ne;
  c.env = (size_t)c.e0 + (c.act ? lane : 0);
  c.stage = stage;
  const int W = A.p.obs_dim;
  float* const g_obs = A.out.obs_buf + (size_t)c.e0 * W;
  const bool bulk_out = A.obs_bulk && c.ne == 32;
  if (tid == 0) s_rng_step = A.st.rng_step ? *reinterpret_cast<const volatile unsigned long long*>(A.st.rng_step) : A.step;
  cons_sync();
  const uint64_t rng_step = s_rng_step;
  const unsigned stamp = ((unsigned)rng_step + 1u) & 0xffffu;
  phase_items<BG, STD, true>(A, c);
  cons_sync();
  phase_scalar<BG, STD, true>(A, c);
  cons_sync();
  unsigned cls = 0;
  if (warp == 0) {
    cls = phase_combine<BG, STD, true>(A, c, stamp, s_rbits);
  } else if (c.act) {
#pragma unroll 1
    for (int i = DEX_NUM_COMMON_TERMS + (warp - 1); i < DEX_NUM_REWARD_TERMS; i += kConsWarps - 1)
      emit_term<BG, STD, true>(A, c, i, task_term<BG, STD, true>(A, c, i));
  }
  if (bulk_out) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  cons_sync();
  if (warp == 0) phase_total<BG, STD, true, OBS>(A, c, cls);
  else if constexpr (OBS) observer_term_rows<BG, STD>(A, c, s_rbits, warp - 1, kConsWarps - 1);
  if (bulk_out) {
    if (tid == 32) bulk_s2g(g_obs, s32(s_obs), (uint32_t)(32 * W * 4));
  } else {
    copy_obs_rows(s_obs, g_obs, W, A.wp, c.ne, warp, lane);
  }
  if (A.out.obs_norm != nullptr) copy_obs_norm_rows(s_obs, A.out, c.e0, W, A.wp, c.ne, warp, lane);
  {
    unsigned mask = s_rbits[0];
    const bool do_reset = !(A.flags & DEX_FLAG_NO_FUSED_RESET);
    if (mask && do_reset) {
      for (int k = 0; mask; ++k) {
        const int e = __ffs(mask) - 1;
        mask &= mask - 1;
        if ((k & (kConsWarps - 1)) == warp) {
          reset_env_warp<BG>(A, c.e0 + e, rng_step, s_reset_u + 32 * warp);
        }
      }
    }
  }
  if (bulk_out && tid == 32) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the obs tile has left shared memory
}

static int g_last_geometry = 0;  // 0 tile kernel, 1 persistent, 2 single pass, 3 persistent + TMA, 4 pair (dex_debug_last_post_kernel)

template <bool STD, bool OBS>
int launch_post2_single(Post2Args& a, cudaStream_t s, int& ok) {
  using LY = Lay<true, true>;
  static int ctas_per_sm = -1, num_sms = 0;
  const uint32_t smem = (uint32_t)(LY::STAGE_BYTES + LY::NWKA * 128 + (32 * a.wp * 4 + 127) / 128 * 128);
  static uint32_t smem_cfg = 0;
  if (smem != smem_cfg) {
    cudaError_t e = cudaFuncSetAttribute(post_physics_v2s_kernel<true, STD, OBS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    int dev = 0, occ = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, post_physics_v2s_kernel<true, STD, OBS>, kOneThreads, smem);
    if (e != cudaSuccess) return (int)e;
    ctas_per_sm = occ;
    smem_cfg = smem;
  }
  const int fused = a.fused_epilogue ? 1 : 0;
  // every CTA (tiles + epilogue) must be resident at once: the epilogue CTA polls the tiles' records
  ok = a.tiles + fused <= num_sms * ctas_per_sm && a.tiles + fused > num_sms * 2;
  if (!ok) return DEX_OK;
  a.smem_bytes = smem;
  a.workers = a.tiles;
  const cudaError_t e = launch_step_kernel(post_physics_v2s_kernel<true, STD, OBS>, a.tiles + fused, kOneThreads, smem, s, a);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

template <bool BG, bool STD, bool TMA, bool OBS, int G = 1>
int launch_post2_k(Post2Args& a, cudaStream_t s) {
  static int ctas_per_sm = -1, num_sms = 0;
  static size_t smem_cfg = 0, smem_occ = 0;
  if (a.smem_bytes > smem_cfg) {
    const size_t want = a.smem_bytes > 100 * 1024 ? a.smem_bytes : 100 * 1024;
    const cudaError_t e = cudaFuncSetAttribute(post_physics_v2_kernel<BG, STD, TMA, OBS, G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want);
    if (e != cudaSuccess) return (int)e;
    smem_cfg = want;
    ctas_per_sm = -1;
  }
  if (ctas_per_sm < 0 || smem_occ != a.smem_bytes) {
    int dev = 0, occ = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, post_physics_v2_kernel<BG, STD, TMA, OBS, G>, G * kThreads, a.smem_bytes);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return DEX_ERR_SIZE;
    ctas_per_sm = occ;
    smem_occ = a.smem_bytes;
  }
  const int fused = a.fused_epilogue ? 1 : 0;
  int cap = num_sms * ctas_per_sm - fused;  // every CTA of the launch (workers + epilogue) is resident at once
  if (cap < 1) cap = 1;
  const int units = a.tiles / G;  // G tiles per CTA and pass (the pair geometry requires tiles % G == 0)
  a.workers = units < cap ? units : cap;
  const cudaError_t e = launch_step_kernel(post_physics_v2_kernel<BG, STD, TMA, OBS, G>, a.workers + fused, G * kThreads, a.smem_bytes, s, a);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

template <bool BG>
int launch_post2_t(Post2Args& a, cudaStream_t s) {
  const int rc = build_post2<BG>(a);
  if (rc) return rc;
  if (a.smem_bytes > 227 * 1024 - 2048) return DEX_ERR_SIZE;
  const bool std_layout = std_layout_matches<BG>(a);
  const bool obs = a.out.episode_sums != nullptr;  // the fused observer is compiled into separate kernel instances
  if constexpr (BG) {
    // geometry: the persistent kernel everywhere, unless the caller asks for the single-pass kernel (DEX_FLAG_SINGLE_PASS, or
    // DEXSTEP_POST_GEOMETRY=single for every call of the process) and the launch has between 2 x SMs and 4 x SMs tiles.
    // Measured at 16 384 envs (3.46 tiles per SM): inputs from HBM 18.3 us persistent / 19.1 us single pass per launch; sim
    // tensors still in L2 (a physics step just wrote them) 15.1 / 12.0 us.  The default follows the first, which is how the
    // benchmark of this path is defined; a deployment whose post-physics pass runs right behind the simulator sets the flag.
    static const bool env_single = []() {
      const char* e = getenv("DEXSTEP_POST_GEOMETRY");
      return e != nullptr && e[0] == 's';
    }();
    const bool allow_single = env_single || (a.flags & DEX_FLAG_SINGLE_PASS) != 0;
    if (allow_single && !(a.flags & DEX_FLAG_TMA_STAGING)) {
      int ok = 0;
      const int rc1 = obs ? (std_layout ? launch_post2_single<true, true>(a, s, ok) : launch_post2_single<false, true>(a, s, ok))
                          : (std_layout ? launch_post2_single<true, false>(a, s, ok) : launch_post2_single<false, false>(a, s, ok));
      if (ok) g_last_geometry = 2;
      if (ok || rc1) return rc1;
    }
  }
  if constexpr (BG) {
    // pair geometry: two tiles per pass and CTA (one CTA of 24 warps per SM), on request only (DEX_FLAG_PAIR_GEOMETRY, or
    // DEXSTEP_POST_GEOMETRY=pair for every call of the process; =nopair overrides the flag): measured slower, see the header.
    static const int env_pair = []() {
      const char* e = getenv("DEXSTEP_POST_GEOMETRY");
      return e == nullptr ? 0 : (strcmp(e, "pair") == 0 ? 1 : (strcmp(e, "nopair") == 0 ? -1 : 0));
    }();
    const bool eligible = a.tiles % 2 == 0 && !(a.flags & DEX_FLAG_TMA_STAGING) && 2u * a.smem_bytes <= 227u * 1024u - 4096u;
    if (eligible && env_pair >= 0 && (env_pair > 0 || (a.flags & DEX_FLAG_PAIR_GEOMETRY) != 0)) {
      a.tma = 0;
      a.smem_bytes *= 2;
      g_last_geometry = 4;
      return obs ? (std_layout ? launch_post2_k<true, true, false, true, 2>(a, s) : launch_post2_k<true, false, false, true, 2>(a, s))
                 : (std_layout ? launch_post2_k<true, true, false, false, 2>(a, s) : launch_post2_k<true, false, false, false, 2>(a, s));
    }
  }
  g_last_geometry = 1;
  if constexpr (BG) {
    build_tma(a);
    if (a.tma) {
      g_last_geometry = 3;
      return obs ? (std_layout ? launch_post2_k<true, true, true, true>(a, s) : launch_post2_k<true, false, true, true>(a, s))
                 : (std_layout ? launch_post2_k<true, true, true, false>(a, s) : launch_post2_k<true, false, true, false>(a, s));
    }
  } else {
    a.tma = 0;
  }
  return obs ? (std_layout ? launch_post2_k<BG, true, false, true>(a, s) : launch_post2_k<BG, false, false, true>(a, s))
             : (std_layout ? launch_post2_k<BG, true, false, false>(a, s) : launch_post2_k<BG, false, false, false>(a, s));
}

}  // namespace pp2
}  // namespace
