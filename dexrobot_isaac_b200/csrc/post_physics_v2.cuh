// post_physics_v2.cuh — the persistent, warp-specialised post-physics kernel (round 2).
//
// Same arithmetic as post_physics.cuh (StepProcessor.process_physics_step, step_processor.py:37-131), different
// machine mapping:
//   * PERSISTENT grid: <= SMs x CTAs/SM worker CTAs (+ one epilogue CTA), each looping over 32-env tiles, so all
//     CTAs are co-resident (the epilogue CTA's poll of the tile records needs no scheduling-order assumption) and
//     every per-CTA setup cost is paid once per launch instead of once per tile.
//   * WARP SPECIALISATION: 4 producer warps stream tile t+1 into a 2-stage shared-memory ring while 8 consumer
//     warps compute tile t (mbarrier full/empty pairs).  Memory time and compute latency overlap instead of adding.
//   * BULK ASYNC COPIES (cp.async.bulk -> SASS UBLKCP, completion on the stage's mbarrier) for everything that is
//     contiguous in HBM at tile granularity: the dof_state tile (32 x 208 B), the full_dof_targets tile and every
//     128-byte row segment of the SoA path state; only the scattered Isaac Gym rows (rigid bodies, contact forces,
//     box root) are gathered with LDG by the producer warps (row-per-warp, sector-exact).
//   * NO TABLES on the way out: every state row / reward row leaves straight from the registers of the lane (= env)
//     that computed it (one 128 B request per row), observation values are written into an obs-shaped shared tile
//     by the work item that produces them and the tile leaves with coalesced row stores.
//   * The one-warp scalar phase of v1 is split over four warps (finger-box contacts + success counter | stage FSM |
//     minima + criteria inputs | averages), the ordered reductions stay single-lane per env (parity).
// Numerics are op-for-op those of v1 (which is pinned to the oracle); lane = env for all arithmetic.
#pragma once
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
template <bool BG>
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
  // work block: values that cross a warp, [field][32] floats (lane = env: conflict-free)
  static constexpr int W_BIN = 0, W_PADBOX = 5;  // 5 each
  static constexpr int W_CDIST = 10, W_DRIFT = 11, W_VMAG = 12, W_AVGHDIFF = 13, W_MINTIPBOX = 14, W_MINTIPZ = 15;
  static constexpr int W_BVEL = 16;              // 6: finite-difference base velocity
  static constexpr int W_BOXHAS = 22, W_SOK = 23, W_BOXFAR = 24, W_HITG = 25, W_HCOK = 26;
  static constexpr int W_WGT = 27;               // 26 weighted reward terms
  static constexpr int W_CHG = 53, W_RHAND = 54, W_FV2 = 58, W_FA2 = 78, W_JL = 98;  // BaseTask (common rewards)
  static constexpr int W_COUNT = BG ? 53 : 118;
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
  int32_t dof2afj[DEX_NUM_DOF];           // dof -> index in active_finger_dof, -1 = not an active finger DOF
  uint64_t bk_ptr[kMaxBulk];              // global base of the row / tile array
  uint16_t bk_dst16[kMaxBulk];            // byte offset inside the stage / 16
  uint16_t bk_len16[kMaxBulk];            // bytes per full tile / 16 (== source stride per tile)
  uint8_t bk_kind[kMaxBulk];
};

// ------------------------------------------------------------------------------------------------
// host: eligibility + derived arguments
// ------------------------------------------------------------------------------------------------
inline bool post2_eligible(const DexParams& P, const DexOutBuffers& out, const int mode, const uint32_t flags) {
  if (mode != DEX_MODE_FULL || out.obs_dict != nullptr || out.obs_norm != nullptr || out.episode_sums != nullptr) return false;
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
    const int late_keys[7] = {DEX_K_THUMB_CONTACT, DEX_K_OTHER_FINGERS_CONTACT, DEX_K_GRASP_STATE, DEX_K_GRASP_DURATION,
                              DEX_K_CURRENT_STAGE, DEX_K_TIME_IN_STAGE, DEX_K_STAGE_PROGRESS};
    for (int i = 0; i < 7; ++i)
      if (BG && has(late_keys[i])) a.obs_late |= 1u << (P.obs_offset[late_keys[i]] >> 5);
  }
  a.rb_off[0] = P.hand_body * 13;
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
    a.rb_off[1 + f] = P.tip_body[f] * 13;
    a.rb_off[6 + f] = P.pad_body[f] * 13;
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
// compute side: ONE warp runs the whole per-env program for its tile, lane = env.  Every compute warp of the SM
// executes the same straight-line instruction stream (instruction fetch is shared, no intra-tile barrier exists),
// intermediates stay in registers, and the compiler schedules the independent chains (5 fingers, 26 DOFs, 10 pad
// pairs, 16 reward terms) of one env side by side.
// ------------------------------------------------------------------------------------------------
template <bool BG, bool STD>
struct Cx {
  const unsigned char* stage;
  float* obs;  // obs tile row of this lane's env
  float* wk;   // work block + lane
  const int32_t* obs_offset;  // generic instance: DexParams.obs_offset
  const int32_t* dof2afj;
  int lane, e0, ne;
  bool act;
  size_t N, env;
  __device__ __forceinline__ int off(const int k) const {
    if constexpr (STD) return std_obs_off<BG>(k);
    else return obs_offset[k];
  }
  __device__ __forceinline__ int afj(const int d) const {
    if constexpr (STD) return std_dof2afj(d);
    else return dof2afj[d];
  }
  __device__ __forceinline__ const float* soa() const { return reinterpret_cast<const float*>(stage) + lane; }
  __device__ __forceinline__ const float* aos() const {
    return reinterpret_cast<const float*>(stage + Lay<BG>::OFF_AOS) + lane * Lay<BG>::SA;
  }
  __device__ __forceinline__ int i64lo(const int r) const {
    return reinterpret_cast<const int*>(stage + Lay<BG>::OFF_I64 + r * 256)[2 * lane];
  }
};

// values that flow between the parts of the per-env program (registers)
struct Mid {
  float bin[DEX_NUM_FINGERS], padbox[DEX_NUM_FINGERS];
  float cdist, drift, vmag, min_tip_box, avg_hdiff, graspdur;
  float bvel[DEX_NUM_BASE];
  bool boxhas, boxfar, hitg, hcok, sok, grasp, thumb, others, j2, j3;
  int sds, stage;
  float fv2, fa2, jl, chg, rhand[4], min_tip_z;  // BaseTask: ordered partial sums of the common reward terms
  float wgt[DEX_NUM_REWARD_TERMS];
};

// ---- the five tip (or pad) poses: world columns, hand-frame columns (observation_encoder.py:915-986) and, for
//      BlindGrasping, the box distances (blind_grasping_task.py:684-771, 1178-1192, 1334-1364)
template <bool BG, bool STD, bool PAD>
__device__ __forceinline__ void part_poses(const Post2Args& A, const Cx<BG, STD>& c, Mid& m) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const float* aos = c.aos();
  const float* rb = aos + LY::A_RB;
  float* obs = c.obs;
  float p[DEX_NUM_FINGERS][7];
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f)
#pragma unroll
    for (int j = 0; j < 7; ++j) p[f][j] = rb[13 + (PAD ? 35 : 0) + 7 * f + j];
  const int ow = c.off(PAD ? DEX_K_FINGERPAD_POSES_WORLD : DEX_K_FINGERTIP_POSES_WORLD);
  if (ow >= 0) {
#pragma unroll
    for (int f = 0; f < DEX_NUM_FINGERS; ++f)
#pragma unroll
      for (int j = 0; j < 7; ++j) obs[ow + 7 * f + j] = p[f][j];
  }
  const int oh = c.off(PAD ? DEX_K_FINGERPAD_POSES_HAND : DEX_K_FINGERTIP_POSES_HAND);
  if (oh >= 0) {
    const Quat hq{rb[3], rb[4], rb[5], rb[6]};
    const float hx = rb[0], hy = rb[1], hz = rb[2];
#pragma unroll
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
      const Vec3 ph = quat_rotate_inverse(hq, Vec3{p[f][0] - hx, p[f][1] - hy, p[f][2] - hz});
      const Quat qh = quat_mul(quat_conj(hq), Quat{p[f][3], p[f][4], p[f][5], p[f][6]});
      float* o = obs + oh + 7 * f;
      o[0] = ph.x;
      o[1] = ph.y;
      o[2] = ph.z;
      o[3] = qh.x;
      o[4] = qh.y;
      o[5] = qh.z;
      o[6] = qh.w;
    }
  }
  if (!PAD) {
    // hitting_ground inputs (step_processor.py:133-181): min over the tips in finger order
    float mz = p[0][2];
#pragma unroll
    for (int f = 1; f < DEX_NUM_FINGERS; ++f) mz = fminf(mz, p[f][2]);
    m.min_tip_z = mz;
    m.hitg = P.has_height_safety && ((rb[2] < P.handbase_threshold) || (mz < P.fingertip_threshold));
  }
  if (BG) {
    const float* box = aos + LY::A_BOX;
    const float bx = box[0], by = box[1], bz = box[2];
    float d[DEX_NUM_FINGERS];
#pragma unroll
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) d[f] = norm3(p[f][0] - bx, p[f][1] - by, p[f][2] - bz);
    if (!PAD) {
      float mb = d[0];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) mb = fminf(mb, d[f]);
      m.min_tip_box = mb;
    } else {
      const int od = c.off(DEX_K_FINGER_TO_OBJECT_DISTANCES), ohd = c.off(DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF);
      float hd[DEX_NUM_FINGERS], dsum = 0.0f, hsum = 0.0f;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        hd[f] = fabsf(p[f][2] - bz);
        m.padbox[f] = d[f];
        obs_put(obs, od, f, d[f]);
        obs_put(obs, ohd, f, hd[f]);
        dsum += d[f];
        hsum += hd[f];
      }
      m.avg_hdiff = div_rn(hsum, 5.0f);  // blind_grasping_task.py:702-724
      obs_put(obs, c.off(DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE), 0, div_rn(dsum, 5.0f));
      obs_put(obs, c.off(DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF), 0, m.avg_hdiff);
      // pad centroid + the height / centroid part of _evaluate_pregrasp_quality (:746-771, :1334-1364)
      float cen[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) cen[k] = (p[0][k] + p[1][k] + p[2][k]) / 3.0f;
      const int oc = c.off(DEX_K_FIRST_THREE_FINGERPAD_CENTROID);
#pragma unroll
      for (int k = 0; k < 3; ++k) obs_put(obs, oc, k, cen[k]);
      m.cdist = norm3(cen[0] - bx, cen[1] - by, cen[2] - bz);
      const bool h_ok = (hd[0] <= P.height_tolerance) && (hd[1] <= P.height_tolerance) && (hd[2] <= P.height_tolerance);
      m.hcok = h_ok && (m.cdist <= P.centroid_tolerance);
      // the ten pad pair distances, triu order (:726-771)
      const int op = c.off(DEX_K_FINGERPAD_DISTANCES);
      if (op >= 0) {
        int k = 0;
#pragma unroll
        for (int a = 0; a < DEX_NUM_FINGERS; ++a)
#pragma unroll
          for (int b = a + 1; b < DEX_NUM_FINGERS; ++b, ++k)
            obs[op + k] = norm3(p[a][0] - p[b][0], p[a][1] - p[b][1], p[a][2] - p[b][2]);
      }
    }
  }
}

// ---- the five contact sensors (observation_encoder.py:658-735, 323-361; reward_calculator.py:183-199) and, for
//      BlindGrasping, the box contact flag and the box stability inputs
template <bool BG, bool STD>
__device__ __forceinline__ void part_contacts(const Post2Args& A, const Cx<BG, STD>& c, Mid& m) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float* aos = c.aos();
  const float* soa = c.soa();
  float* obs = c.obs;
  const size_t N = c.N, env = c.env;
  const float dt = P.control_dt;
  float cf[DEX_NUM_FINGERS][3], dur[DEX_NUM_FINGERS], prevb[DEX_NUM_FINGERS];
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
#pragma unroll
    for (int k = 0; k < 3; ++k) cf[f][k] = aos[LY::A_CF + 3 * f + k];
    dur[f] = soa[(LY::R_CDUR + f) * 32];
    prevb[f] = soa[(LY::R_PBIN + f) * 32];
  }
  const int ocf = c.off(DEX_K_CONTACT_FORCES), om = c.off(DEX_K_CONTACT_FORCE_MAGNITUDE);
  const int ob = c.off(DEX_K_CONTACT_BINARY), od = c.off(DEX_K_CONTACT_DURATION);
  const bool common = !BG && (STD || A.common != 0), first = !STD && A.first != 0;
  float chg = 0.0f;
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
    const float mag = norm3(cf[f][0], cf[f][1], cf[f][2]);
    const bool bin = mag > P.contact_binary_threshold;
    const float ndur = (bin && prevb[f] == 0.0f) ? 1.0f : (bin ? dur[f] + 1.0f : 0.0f);
    const float fb = bin ? 1.0f : 0.0f;
    st.contact_duration_steps[f * N + env] = ndur;
    st.prev_contact_binary[f * N + env] = fb;
    if (STD || st.contact_forces != nullptr) {
#pragma unroll
      for (int k = 0; k < 3; ++k) st.contact_forces[(3 * f + k) * N + env] = cf[f][k];
    }
    m.bin[f] = fb;
    if (ocf >= 0) {
#pragma unroll
      for (int k = 0; k < 3; ++k) obs[ocf + 3 * f + k] = cf[f][k];
    }
    obs_put(obs, om, f, mag);
    obs_put(obs, ob, f, fb);
    obs_put(obs, od, f, ndur * dt);
    if (common) {
      const bool c_now = mag > 0.1f;  // hard-coded 0.1 N (reward_calculator.py:186)
      const bool c_prev = first ? c_now : (c.stage[LY::OFF_U8 + f * 32 + c.lane] != 0);
      chg += (c_now != c_prev) ? 1.0f : 0.0f;  // summed in finger order as the term does (reward_calculator.py:183-199)
      st.prev_contacts[f * N + env] = c_now ? 1 : 0;
    }
  }
  m.chg = chg;
  if (BG) {
    const float* cfb = aos + LY::A_CF + 15;
    const float* box = aos + LY::A_BOX;
    const float b[6] = {box[0], box[1], box[2], box[3], box[4], box[5]};
    m.boxhas = norm3(cfb[0], cfb[1], cfb[2]) > P.contact_binary_threshold;
    m.drift = norm3(b[0] - soa[(LY::R_BOX0 + 0) * 32], b[1] - soa[(LY::R_BOX0 + 1) * 32], b[2] - soa[(LY::R_BOX0 + 2) * 32]);
    m.vmag = norm3(b[3], b[4], b[5]);
    m.sok = (m.drift <= P.position_drift_tolerance) && (m.vmag <= P.velocity_tolerance);
    const int opos = c.off(DEX_K_OBJECT_POS), ovel = c.off(DEX_K_OBJECT_VEL);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      obs_put(obs, opos, k, b[k]);
      obs_put(obs, ovel, k, b[3 + k]);
    }
  }
}

// ---- hand pose, ARR-aligned pose (observation_encoder.py:1483-1521), hand -> box distance, hand velocity reward
//      terms (reward_calculator.py:100-181), episode time (dexhand_base.py:703-712: PRE-increment count)
template <bool BG, bool STD>
__device__ __forceinline__ void part_hand(const Post2Args& A, const Cx<BG, STD>& c, Mid& m) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float* aos = c.aos();
  const float* soa = c.soa();
  const float* rb = aos + LY::A_RB;
  float* obs = c.obs;
  float h[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) h[j] = rb[j];
  obs_put(obs, c.off(DEX_K_EPISODE_TIME), 0, (float)c.i64lo(LY::I_STEP) * P.control_dt);
  const int oh = c.off(DEX_K_HAND_POSE);
  if (oh >= 0) {
#pragma unroll
    for (int j = 0; j < 7; ++j) obs[oh + j] = h[j];
  }
  const int oa = c.off(DEX_K_HAND_POSE_ARR_ALIGNED);
  if (oa >= 0) {
    const float sh = 0.7071067811865476f;
    const Quat qa = quat_mul(Quat{h[3], h[4], h[5], h[6]}, Quat{-0.0f, -sh, -0.0f, sh});
    obs[oa + 0] = h[0];
    obs[oa + 1] = h[1];
    obs[oa + 2] = h[2];
    obs[oa + 3] = qa.x;
    obs[oa + 4] = qa.y;
    obs[oa + 5] = qa.z;
    obs[oa + 6] = qa.w;
  }
  if (BG) {
    const float* box = aos + LY::A_BOX;
    const float hb = norm3(h[0] - box[0], h[1] - box[1], h[2] - box[2]);
    m.boxfar = hb > P.max_box_distance;
    obs_put(obs, c.off(DEX_K_HAND_TO_OBJECT_DISTANCE), 0, hb);
  }
  if (!BG && (STD || A.common)) {
    const bool first = !STD && A.first != 0;
    float hv[3], ha[3], phv[3], pha[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      hv[k] = rb[7 + k];
      ha[k] = rb[10 + k];
      phv[k] = first ? hv[k] : soa[(LY::R_PHV + k) * 32];
      pha[k] = first ? ha[k] : soa[(LY::R_PHA + k) * 32];
      st.prev_hand_vel[k * c.N + c.env] = hv[k];
      st.prev_hand_ang_vel[k * c.N + c.env] = ha[k];
    }
    m.rhand[0] = expf(-0.2f * norm3(hv[0], hv[1], hv[2]));
    m.rhand[1] = expf(-0.2f * norm3(ha[0], ha[1], ha[2]));
    m.rhand[2] = expf(-0.5f * norm3(hv[0] - phv[0], hv[1] - phv[1], hv[2] - phv[2]));
    m.rhand[3] = expf(-0.5f * norm3(ha[0] - pha[0], ha[1] - pha[1], ha[2] - pha[2]));
  }
}

// ---- the 26 DOFs: finite-difference velocity (observation_encoder.py:298-321), obs columns, common-reward partial
//      sums (reward_calculator.py:94-98,138-160: Isaac dof_vel, not the finite difference; summed in DOF order)
template <bool BG, bool STD, int Q0, int Q1>  // DOF pairs [Q0, Q1)
__device__ __forceinline__ void part_dofs(const Post2Args& A, const Cx<BG, STD>& c, Mid& m) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float* soa = c.soa();
  float* obs = c.obs;
  const size_t N = c.N, env = c.env;
  const float dt = P.control_dt;
  const float4* d4 = reinterpret_cast<const float4*>(c.stage + LY::OFF_DOF) + c.lane * 13;
  const bool want_tgt = !BG && (STD || A.want_tgt != 0);
  const bool common = !BG && (STD || A.common != 0), first = !STD && A.first != 0;
  const int o_bp = c.off(DEX_K_BASE_DOF_POS), o_bv = c.off(DEX_K_BASE_DOF_VEL), o_bt = c.off(DEX_K_BASE_DOF_TARGET);
  const int o_ap = c.off(DEX_K_ALL_FINGER_DOF_POS), o_av = c.off(DEX_K_ALL_FINGER_DOF_VEL),
            o_at = c.off(DEX_K_ALL_FINGER_DOF_TARGET);
  const int o_fp = c.off(DEX_K_ACTIVE_FINGER_DOF_POS), o_fv = c.off(DEX_K_ACTIVE_FINGER_DOF_VEL),
            o_ft = c.off(DEX_K_ACTIVE_FINGER_DOF_TARGET);
#pragma unroll
  for (int q = Q0; q < Q1; ++q) {
    const float4 v4 = d4[q];
    const float pos2[2] = {v4.x, v4.z}, vel2[2] = {v4.y, v4.w};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int d = 2 * q + h;
      const float pos = pos2[h];
      const float fd = div_rn(pos - soa[(LY::R_PDP + d) * 32], dt);
      st.prev_dof_pos[d * N + env] = pos;
      const float tgt = want_tgt ? reinterpret_cast<const float*>(c.stage + LY::OFF_TGT)[c.lane * 26 + d] : 0.0f;
      if (d < DEX_NUM_BASE) {
        obs_put(obs, o_bp, d, pos);
        obs_put(obs, o_bv, d, fd);
        if (!BG) obs_put(obs, o_bt, d, tgt);
        if (BG) c.wk[(LY::W_BVEL + (d < DEX_NUM_BASE ? d : 0)) * 32] = fd;
      } else {
        const int j = d - DEX_NUM_BASE;
        obs_put(obs, o_ap, j, pos);
        obs_put(obs, o_av, j, fd);
        if (!BG) obs_put(obs, o_at, j, tgt);
        const int aj = c.afj(d);
        if (aj >= 0) {
          obs_put(obs, o_fp, aj, pos);
          obs_put(obs, o_fv, aj, fd);
          if (!BG) obs_put(obs, o_ft, aj, tgt);
        }
        if (common) {
          const float iv = vel2[h];
          const float pvv = first ? iv : soa[(LY::R_PFDV + j) * 32];
          // partial terms: summed in DOF order by the term itself (reward_calculator.py:138-160)
          c.wk[(LY::W_FV2 + j) * 32] = iv * iv;
          const float acc = iv - pvv;
          c.wk[(LY::W_FA2 + j) * 32] = acc * acc;
          const float nj = div_rn(2.0f * (pos - P.dof_lower[d]), P.dof_upper[d] - P.dof_lower[d]) - 1.0f;
          c.wk[(LY::W_JL + j) * 32] = clampf(fabsf(nj) - 0.8f, 0.0f, 1.0f);
          st.prev_finger_dof_vel[j * N + env] = iv;
        }
      }
    }
  }
}

// ---- prev_actions columns (observation_encoder.py:287-296); active_rule_targets <- active_prev_targets
//      (rules.py:78-95, step_processor.py:73)
template <bool BG, bool STD>
__device__ __forceinline__ void part_copies(const Post2Args& A, const Cx<BG, STD>& c) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const float* soa = c.soa();
  float* obs = c.obs;
  float apt[DEX_NUM_ACTIVE];
#pragma unroll
  for (int k = 0; k < DEX_NUM_ACTIVE; ++k) apt[k] = soa[(LY::R_APT + k) * 32];
  const int op = c.off(DEX_K_PREV_ACTIONS);
  if (op >= 0) {
#pragma unroll
    for (int k0 = 0; k0 < DEX_NUM_ACTIVE; k0 += 6) {  // num_actions is 6, 12 or 18
      if (k0 < P.num_actions) {
        float v[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) v[k] = soa[(LY::R_PACT + k0 + k) * 32];
#pragma unroll
        for (int k = 0; k < 6; ++k) obs[op + k0 + k] = v[k];
      }
    }
  }
  const int o1 = c.off(DEX_K_ACTIVE_PREV_TARGETS), o2 = c.off(DEX_K_ACTIVE_RULE_TARGETS);
  float* art = A.st.active_rule_targets + c.env;
#pragma unroll
  for (int k = 0; k < DEX_NUM_ACTIVE; ++k) {
    art[(size_t)k * c.N] = apt[k];
    obs_put(obs, o1, k, apt[k]);
    obs_put(obs, o2, k, apt[k]);
  }
}

// ---- BlindGrasping: finger-box contacts (blind_grasping_task.py:773-833), success duration (:626-648), stage FSM
//      (:835-978; plain fp32 accumulation of (float)control_dt decides the transition step)
// WRITE = false: the reward-term warps recompute the (cheap) chain for themselves instead of waiting for warp 0
template <bool STD, bool WRITE>
__device__ __forceinline__ void part_task_state(const Post2Args& A, const Cx<true, STD>& c, Mid& m) {
  using LY = Lay<true>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float dt = P.control_dt;
  float* obs = c.obs;
  const float* soa = c.soa();
  const size_t env = c.env;
  const float box_z = c.aos()[LY::A_BOX + 2];
  const int sds_old = c.i64lo(LY::I_SDS);
  int stage = c.i64lo(LY::I_STAGE);
  float tis = soa[LY::R_TIS * 32] + dt;
  float scd = soa[LY::R_SCD * 32];
  bool fbc[DEX_NUM_FINGERS];
  int n_on_box = 0;
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
    fbc[f] = (m.bin[f] != 0.0f) && m.boxhas && (m.padbox[f] < P.proximity_threshold);
    n_on_box += fbc[f] ? 1 : 0;
  }
  m.thumb = fbc[0];
  m.others = fbc[1] || fbc[2] || fbc[3] || fbc[4];
  m.grasp = m.thumb && m.others;
  if (WRITE) obs_put(obs, c.off(DEX_K_THUMB_CONTACT), 0, m.thumb ? 1.0f : 0.0f);
  if (WRITE) obs_put(obs, c.off(DEX_K_OTHER_FINGERS_CONTACT), 0, m.others ? 1.0f : 0.0f);
  if (WRITE) obs_put(obs, c.off(DEX_K_GRASP_STATE), 0, m.grasp ? 1.0f : 0.0f);
  const bool cond = (box_z > P.success_height_threshold) && (n_on_box >= P.min_fingers_for_grasp);
  m.sds = cond ? sds_old + 1 : 0;
  if (WRITE) st.success_duration_steps[env] = (long long)m.sds;
  if (WRITE) st.success_conditions_met[env] = cond ? 1 : 0;
  m.graspdur = (float)m.sds * dt;
  if (WRITE) obs_put(obs, c.off(DEX_K_GRASP_DURATION), 0, m.graspdur);
  // FSM
  const bool thumb_other = (m.bin[0] != 0.0f) && ((m.bin[1] != 0.0f) || (m.bin[2] != 0.0f) || (m.bin[3] != 0.0f) || (m.bin[4] != 0.0f));
  if (stage == 2) scd = thumb_other ? scd + dt : 0.0f;
  const bool s1_done = (stage == 1) && (tis >= P.stage1_duration);
  const bool s2_done = (stage == 2) && ((scd >= P.stage2_contact_success_threshold) || (tis >= P.stage2_duration));
  if (s1_done) stage = 2;
  if (s2_done) stage = 3;
  if (s1_done || s2_done) {
    tis = 0.0f;
    scd = 0.0f;
  }
  if (WRITE) st.current_stage[env] = (long long)stage;
  if (WRITE) st.time_in_stage[env] = tis;
  if (WRITE) st.stage_contact_duration[env] = scd;
  if (WRITE) st.just_transitioned_to_stage2[env] = s1_done ? 1 : 0;
  if (WRITE) st.just_transitioned_to_stage3[env] = s2_done ? 1 : 0;
  float prog = 0.0f;  // :958-978
  if (stage == 1) prog = clampf(div_rn(tis, P.stage1_duration), 0.0f, 1.0f);
  if (stage == 2) prog = clampf(div_rn(tis, P.stage2_duration), 0.0f, 1.0f);
  if (stage == 3) prog = 1.0f;
  m.stage = stage;
  m.j2 = s1_done;
  m.j3 = s2_done;
  if (WRITE) obs_put(obs, c.off(DEX_K_CURRENT_STAGE), 0, (float)stage);
  if (WRITE) obs_put(obs, c.off(DEX_K_TIME_IN_STAGE), 0, tis);
  if (WRITE) obs_put(obs, c.off(DEX_K_STAGE_PROGRESS), 0, prog);
}

// raw + weighted rows of term i leave from registers; the weighted value is parked in the work block for the ordered total
template <bool BG, bool STD>
__device__ __forceinline__ void emit_term(const Post2Args& A, const Cx<BG, STD>& c, const int i, const float r) {
  const float w = A.p.reward_weight[i];
  if (!STD && w == 0.0f) return;  // reward_calculator.py:247-251: zero-weight terms are not materialised (STD: none is)
  const int row = A.p.component_row[i];
  const float rw = r * w;
  float* g = A.out.reward_components + c.env;
  g[(size_t)row * c.N] = r;
  g[(size_t)(row + 1) * c.N] = rw;
  c.wk[(Lay<BG>::W_WGT + i) * 32] = rw;
}

// the reward terms are dealt round-robin to the consumer warps 1..7; warp CW runs the terms with (term % 7) == CW - 1 as
// one straight-line block (TERM's value expression is only instantiated for the owner)
#define DEX_TERM(BGF, i, expr)                                   \
  if constexpr (((i) % (kConsWarps - 1)) == CW - 1) emit_term<BGF, STD>(A, c, (i), (expr))

// ---- common reward terms (reward_calculator.py:66-221); BaseTask.  The ordered partial sums are formed here.
template <bool STD, int CW>
__device__ __forceinline__ void part_common_terms(const Post2Args& A, const Cx<false, STD>& c) {
  using LY = Lay<false>;
  const float* wk = c.wk;
  auto sum20 = [&](const int f0) {
    float a = 0.0f;
#pragma unroll
    for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) a += wk[(f0 + j) * 32];
    return a;
  };
  DEX_TERM(false, DEX_R_ALIVE, 1.0f);
  DEX_TERM(false, DEX_R_HEIGHT_SAFETY, clampf(1.0f - expf(-(wk[LY::W_MINTIPZ * 32] - 0.02f) * 20.0f), 0.0f, 1.0f));
  DEX_TERM(false, DEX_R_FINGER_VELOCITY, expf(-0.1f * sqrtf(sum20(LY::W_FV2))));
  DEX_TERM(false, DEX_R_HAND_VELOCITY, wk[(LY::W_RHAND + 0) * 32]);
  DEX_TERM(false, DEX_R_HAND_ANGULAR_VELOCITY, wk[(LY::W_RHAND + 1) * 32]);
  DEX_TERM(false, DEX_R_JOINT_LIMIT, 1.0f - div_rn(sum20(LY::W_JL), 20.0f));
  DEX_TERM(false, DEX_R_FINGER_ACCELERATION, expf(-2.0f * sqrtf(sum20(LY::W_FA2))));
  DEX_TERM(false, DEX_R_HAND_ACCELERATION, wk[(LY::W_RHAND + 2) * 32]);
  DEX_TERM(false, DEX_R_HAND_ANGULAR_ACCELERATION, wk[(LY::W_RHAND + 3) * 32]);
  DEX_TERM(false, DEX_R_CONTACT_STABILITY, expf(-wk[LY::W_CHG * 32]));  // summed in finger order by part_contacts
}

// ---- BlindGrasping task terms (blind_grasping_task.py:980-1208), after the FSM (m: part_task_state of this warp)
template <bool STD, int CW>
__device__ __forceinline__ void part_task_terms(const Post2Args& A, const Cx<true, STD>& c, const Mid& m) {
  using LY = Lay<true>;
  const DexParams& P = A.p;
  const float* wk = c.wk;
  const float* aos = c.aos();
  const float* rb = aos + LY::A_RB;
  const float m1 = (m.stage == 1) ? 1.0f : 0.0f, m2 = (m.stage == 2) ? 1.0f : 0.0f, m3 = (m.stage == 3) ? 1.0f : 0.0f;
  const float fg = m.grasp ? 1.0f : 0.0f;
  DEX_TERM(true, DEX_R_S1_HEIGHT_ALIGNMENT, expf(-P.height_alignment_decay * wk[LY::W_AVGHDIFF * 32]) * m1);
  DEX_TERM(true, DEX_R_S1_CENTROID_POSITIONING, expf(-P.centroid_positioning_decay * wk[LY::W_CDIST * 32]) * m1);
  DEX_TERM(true, DEX_R_S1_OBJECT_STABILITY, expf(-P.object_stability_decay * (wk[LY::W_DRIFT * 32] + wk[LY::W_VMAG * 32])) * m1);
  if constexpr ((DEX_R_S1_FINGER_HEIGHT_CONSISTENCY % (kConsWarps - 1)) == CW - 1) {  // torch.var(dim=1), unbiased, z of the first three pads
    const float z0 = rb[13 + 35 + 2], z1 = rb[13 + 35 + 7 + 2], z2 = rb[13 + 35 + 14 + 2];
    const float mean = (z0 + z1 + z2) / 3.0f;
    const float d0 = z0 - mean, d1 = z1 - mean, d2 = z2 - mean;
    const float var = (d0 * d0 + d1 * d1 + d2 * d2) / 2.0f;
    emit_term<true, STD>(A, c, DEX_R_S1_FINGER_HEIGHT_CONSISTENCY, expf(-P.height_consistency_decay * var) * m1);
  }
  DEX_TERM(true, DEX_R_S1_THUMB_ROTATION,
           expf(-5.0f * fabsf(reinterpret_cast<const float*>(c.stage + LY::OFF_DOF)[c.lane * 52 + 2 * DEX_NUM_BASE] - P.half_pi)) * m1);
  DEX_TERM(true, DEX_R_S2_THUMB_CONTACT, (m.thumb ? 1.0f : 0.0f) * m2);
  DEX_TERM(true, DEX_R_S2_OTHER_FINGERS_CONTACT, (m.others ? 1.0f : 0.0f) * m2);
  DEX_TERM(true, DEX_R_S2_GRASP_ACHIEVEMENT, fg * m2);
  DEX_TERM(true, DEX_R_S2_FINGERPAD_PROXIMITY,
           expf(-P.fingerpad_proximity_decay * fmaxf(wk[LY::W_MINTIPBOX * 32], P.min_reward_distance)) * m2);
  if constexpr ((DEX_R_S2_BASE_STABILITY % (kConsWarps - 1)) == CW - 1) {
    float bv2 = 0.0f;
#pragma unroll
    for (int d = 0; d < DEX_NUM_BASE; ++d) {
      const float v = wk[(LY::W_BVEL + d) * 32];  // finite-difference base velocity
      bv2 += v * v;
    }
    emit_term<true, STD>(A, c, DEX_R_S2_BASE_STABILITY, expf(-P.base_stability_decay * sqrtf(bv2)) * m2);
  }
  DEX_TERM(true, DEX_R_S3_OBJECT_HEIGHT, clampf(div_rn(aos[LY::A_BOX + 2] - P.box_z, P.lift_height_range), 0.0f, 1.0f) * m3);
  DEX_TERM(true, DEX_R_S3_GRASP_MAINTENANCE, fg * m3);
  DEX_TERM(true, DEX_R_S3_GRASP_DURATION, clampf(div_rn(m.graspdur, P.contact_duration_threshold_s), 0.0f, 1.0f) * m3);
  // completion bonuses :1025-1048 (transition and not failed), penetration :1194-1208
  DEX_TERM(true, DEX_R_S1_COMPLETION, (m.j2 && m.hcok && m.sok) ? 1.0f : 0.0f);
  DEX_TERM(true, DEX_R_S2_COMPLETION, (m.j3 && m.grasp) ? 1.0f : 0.0f);
  DEX_TERM(true, DEX_R_PENETRATION_PENALTY, fmaxf(P.penetration_threshold - wk[LY::W_MINTIPBOX * 32], 0.0f) * P.penetration_depth_scale);
}

// ---- termination (step_processor.py:133-181, termination_manager.py:120-278), masks, tile record.
// Returns the exclusive termination class of this lane's env (bit 0 success, 1 failure, 2 timeout) and the reset ballot.
template <bool BG, bool STD, bool FULL>
__device__ __forceinline__ unsigned part_terminate(const Post2Args& A, const Cx<BG, STD>& c, const Mid& m, const unsigned stamp,
                                                   unsigned& rbits_out) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  bool should_reset = false, t_success = false, t_failure = false, t_timeout = false;
  int ep_len_done = 0;
  bool crit[DEX_NUM_CRITERIA];
#pragma unroll
  for (int k = 0; k < DEX_NUM_CRITERIA; ++k) crit[k] = false;
  if (c.act) {
    const int step_now = c.i64lo(LY::I_STEP) + 1;  // step_processor.py:80
    st.episode_step_count[c.env] = (long long)step_now;
    ep_len_done = step_now;
    crit[DEX_C_HITTING_GROUND] = m.hitg;
    if (BG) {
      crit[DEX_C_GRASP_LIFT_SUCCESS] = m.sds >= P.contact_duration_threshold_steps;
      crit[DEX_C_BOX_TOO_FAR] = m.boxfar;
      const bool pregrasp_failed = m.j2 && !(m.hcok && m.sok);  // _evaluate_pregrasp_quality :1334-1364
      const bool contact_failed = m.j3 && !m.grasp;
      const bool grasp_lost = (m.stage == 3) && !m.grasp;
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
  rbits_out = rbits;
  const int lane = c.lane;
  if (lane == 0) A.out.reset_bits[c.e0 >> 5] = rbits;
  const unsigned bs = __ballot_sync(fullmask, t_success), bf = __ballot_sync(fullmask, t_failure),
                 bt = __ballot_sync(fullmask, t_timeout);
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

// ---- ordered weighted sum (reward_calculator.py:223-275) + termination rewards
//      (termination_manager.py:280-321, step_processor.py:204-219)
template <bool BG, bool STD>
__device__ __forceinline__ void part_total(const Post2Args& A, const Cx<BG, STD>& c, const unsigned cls) {
  const DexParams& P = A.p;
  constexpr int I0 = BG ? DEX_NUM_COMMON_TERMS : 0, I1 = BG ? DEX_NUM_REWARD_TERMS : DEX_NUM_COMMON_TERMS;
  float wv[I1 - I0];
#pragma unroll
  for (int i = I0; i < I1; ++i) wv[i - I0] = c.wk[(Lay<BG>::W_WGT + i) * 32];  // zero-weight terms hold stale values: skipped below
  float total = 0.0f;
#pragma unroll
  for (int i = I0; i < I1; ++i)
    if (STD || P.reward_weight[i] != 0.0f) total += wv[i - I0];  // dict order: common, then task
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

// phase A: everything that depends only on the staged tile.  One fixed job per consumer warp, straight-line code; the
// values another warp needs later go to the work block.
template <bool BG, bool STD>
__device__ __forceinline__ void phase_a(const Post2Args& A, const Cx<BG, STD>& c, const int cw) {
  using LY = Lay<BG>;
  float* wk = c.wk;
  Mid m;
  switch (cw) {
    case 0:  // tips
      part_poses<BG, STD, false>(A, c, m);
      wk[LY::W_HITG * 32] = m.hitg ? 1.0f : 0.0f;
      if (BG) wk[LY::W_MINTIPBOX * 32] = m.min_tip_box;
      else wk[LY::W_MINTIPZ * 32] = m.min_tip_z;
      break;
    case 1:  // pads
      part_poses<BG, STD, true>(A, c, m);
      if (BG) {
#pragma unroll
        for (int f = 0; f < DEX_NUM_FINGERS; ++f) wk[(LY::W_PADBOX + f) * 32] = m.padbox[f];
        wk[LY::W_AVGHDIFF * 32] = m.avg_hdiff;
        wk[LY::W_CDIST * 32] = m.cdist;
        wk[LY::W_HCOK * 32] = m.hcok ? 1.0f : 0.0f;
      }
      break;
    case 2:  // contact sensors, box stability
      part_contacts<BG, STD>(A, c, m);
      if (BG) {
#pragma unroll
        for (int f = 0; f < DEX_NUM_FINGERS; ++f) wk[(LY::W_BIN + f) * 32] = m.bin[f];
        wk[LY::W_BOXHAS * 32] = m.boxhas ? 1.0f : 0.0f;
        wk[LY::W_DRIFT * 32] = m.drift;
        wk[LY::W_VMAG * 32] = m.vmag;
        wk[LY::W_SOK * 32] = m.sok ? 1.0f : 0.0f;
      } else {
        wk[LY::W_CHG * 32] = m.chg;
      }
      break;
    case 3:  // hand, copies
      part_hand<BG, STD>(A, c, m);
      if (BG) {
        wk[LY::W_BOXFAR * 32] = m.boxfar ? 1.0f : 0.0f;
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) wk[(LY::W_RHAND + k) * 32] = m.rhand[k];
      }
      part_copies<BG, STD>(A, c);
      break;
    case 4: part_dofs<BG, STD, 0, 4>(A, c, m); break;
    case 5: part_dofs<BG, STD, 4, 7>(A, c, m); break;
    case 6: part_dofs<BG, STD, 7, 10>(A, c, m); break;
    default: part_dofs<BG, STD, 10, 13>(A, c, m); break;
  }
}

// work block -> the inputs of part_task_state / part_task_terms / part_terminate
template <bool BG>
__device__ __forceinline__ void import_mid(const float* wk, Mid& m) {
  using LY = Lay<BG>;
  m.hitg = wk[LY::W_HITG * 32] != 0.0f;
  if (BG) {
#pragma unroll
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
      m.bin[f] = wk[(LY::W_BIN + f) * 32];
      m.padbox[f] = wk[(LY::W_PADBOX + f) * 32];
    }
    m.boxhas = wk[LY::W_BOXHAS * 32] != 0.0f;
    m.boxfar = wk[LY::W_BOXFAR * 32] != 0.0f;
    m.hcok = wk[LY::W_HCOK * 32] != 0.0f;
    m.sok = wk[LY::W_SOK * 32] != 0.0f;
  }
}

// phase C, warps 1..7: the reward terms (BlindGrasping: behind a private copy of the task-state chain)
template <bool BG, bool STD, int CW>
__device__ __forceinline__ void phase_c_terms(const Post2Args& A, const Cx<BG, STD>& c) {
  if constexpr (BG) {
    Mid m;
    import_mid<true>(c.wk, m);
    part_task_state<STD, false>(A, c, m);
    part_task_terms<STD, CW>(A, c, m);
  } else {
    part_common_terms<STD, CW>(A, c);
  }
}

__device__ __forceinline__ void cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsThreads) : "memory"); }

// shared -> global bulk store of the flat obs tile (TMA engine): one instruction for 20 KB
__device__ __forceinline__ void bulk_s2g(void* dst, const uint32_t src, const uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

// one tile on the 8 consumer warps: phase A | barrier | phase C | barrier | phase D | barrier
template <bool BG, bool STD, bool FULL>
__device__ __forceinline__ void consume_tile(const Post2Args& A, Cx<BG, STD>& c, const int warp, const int it, const unsigned stamp_in,
                                             unsigned long long* s_rng_step, unsigned* s_rbits, unsigned& stamp, uint64_t& rng_step,
                                             const uint32_t bar_empty_s, float* s_obs, float* s_reset_u, bool& store_pending TL2_PARAM) {
  const int lane = c.lane, tid = warp * 32 + lane;
  // Philox step counter == stamp of this step's tile records: read once per CTA, behind the first full barrier (i.e.
  // behind the producers' griddepcontrol.wait; the epilogue CTA advances it only after every record of this launch is out)
  unsigned long long rs_tmp = 0ull;
  if (it == 0 && tid == 0) rs_tmp = A.st.rng_step ? *reinterpret_cast<const volatile unsigned long long*>(A.st.rng_step) : A.step;
  if (FULL || c.act) phase_a<BG, STD>(A, c, warp);
  if (it == 0 && tid == 0) *s_rng_step = rs_tmp;
  TL2(it, 2);
  cons_sync();
  TL2(it, 3);
  if (it == 0) {
    rng_step = *s_rng_step;
    stamp = ((unsigned)rng_step + 1u) & 0xffffu;
  }
  (void)stamp_in;
  // ---- phase C: warp 0 = task state + termination + tile record; warps 1..7 = reward terms ----
  unsigned cls = 0;
  if (warp == 0) {
    Mid m;
    if (FULL || c.act) {
      import_mid<BG>(c.wk, m);
      if constexpr (BG) part_task_state<STD, true>(A, c, m);
    }
    unsigned rbits;
    cls = part_terminate<BG, STD, FULL>(A, c, m, stamp, rbits);
    if (lane == 0) *s_rbits = rbits;
  } else if (FULL || c.act) {
    switch (warp) {
      case 1: phase_c_terms<BG, STD, 1>(A, c); break;
      case 2: phase_c_terms<BG, STD, 2>(A, c); break;
      case 3: phase_c_terms<BG, STD, 3>(A, c); break;
      case 4: phase_c_terms<BG, STD, 4>(A, c); break;
      case 5: phase_c_terms<BG, STD, 5>(A, c); break;
      case 6: phase_c_terms<BG, STD, 6>(A, c); break;
      default: phase_c_terms<BG, STD, 7>(A, c); break;
    }
  }
  const bool bulk_out = FULL && A.obs_bulk;
  if (bulk_out) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // obs tile writes -> visible to the bulk store
  __syncwarp();
  TL2(it, 4);
  cons_sync();
  TL2(it, 5);
  // the raw stage is not read after this point: hand it back to the producers
  if (lane == 0) mbar_arrive(bar_empty_s);
  // ---- phase D: total, obs tile out, fused reset ----
  if (warp == 0 && (FULL || c.act)) part_total<BG, STD>(A, c, cls);
  const int W = A.p.obs_dim;
  float* const g_obs = A.out.obs_buf + (size_t)c.e0 * W;
  if (bulk_out) {
    if (tid == 32) {
      bulk_s2g(g_obs, s32(s_obs), (uint32_t)(32 * W * 4));
      store_pending = true;
    }
  } else {
    copy_obs_rows(s_obs, g_obs, W, A.wp, c.ne, warp, lane);
  }
  {
    unsigned mask = *s_rbits;  // written by warp 0 before the barrier above
    if (mask && !(A.flags & DEX_FLAG_NO_FUSED_RESET)) {  // fused reset (dexhand_base.py:743-756): one warp per env
      for (int k = 0; mask; ++k) {
        const int e = __ffs(mask) - 1;
        mask &= mask - 1;
        if ((k & (kConsWarps - 1)) == warp) reset_env_warp<BG>(A, c.e0 + e, rng_step, s_reset_u + 32 * warp);
      }
    }
  }
  if (bulk_out && tid == 32) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // obs tile may be rewritten
  TL2(it, 7);
  cons_sync();  // work block, obs tile and s_rbits are reused by the next tile
}

// ------------------------------------------------------------------------------------------------
// the kernel: persistent, 8 consumer + 4 producer warps per CTA, 2-stage ring, two CTAs per SM (BlindGrasping)
// ------------------------------------------------------------------------------------------------
template <bool BG, bool STD>
__global__ void __launch_bounds__(kThreads, BG ? 2 : 1) post_physics_v2_kernel(const __grid_constant__ Post2Args A) {
  using LY = Lay<BG>;
  extern __shared__ __align__(128) unsigned char smem2[];
  __shared__ __align__(8) unsigned long long s_bar[2 * kStages];  // full[0..], empty[0..]
  __shared__ float s_reset_u[32 * kConsWarps];
  __shared__ int s_rbo[kRbW + 1];  // gathered rigid slot -> float offset inside one env's [NB,13] block
  __shared__ unsigned long long s_rng_step;
  __shared__ unsigned s_rbits;
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
      mbar_init(bar_full + 8 * s, kProdThreads + kProdWarps);  // one cp.async-completion arrival per producer thread + one per warp
      mbar_init(bar_empty + 8 * s, kConsWarps);
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
  float* const wk_base = reinterpret_cast<float*>(smem2 + kStages * LY::STAGE_BYTES);
  float* const s_obs = wk_base + LY::W_COUNT * 32;
  const int tiles = A.tiles, workers = A.workers;

  if (warp >= kConsWarps) {
    // =========================== producer warps ===========================
    const int pw = warp - kConsWarps;
    const int rbo0 = s_rbo[lane], rbo1 = s_rbo[lane + 32], rbo2 = s_rbo[min(lane + 64, kRbW - 1)];
    const bool p_rb0 = lane < 7 || lane >= 13 || A.common;  // hand velocities feed only the common reward terms
    const bool p_rb2 = lane + 64 < kRbW;
    const bool is_box = BG && lane >= 18 && lane < 24;
    const bool p_x = is_box || lane < (BG ? 18 : 15);
    const int xrow = is_box ? A.p.num_actors * 13 : A.p.num_contact_bodies * 3;
    int xoff;
    if (is_box) {
      const int k = lane - 18;
      xoff = A.p.box_actor * 13 + (k < 3 ? k : k + 4);
    } else {
      const int f = min(lane / 3, 5);
      xoff = A.cf_off[f] + (lane - 3 * f);
    }
    const float* const x_base = (is_box ? (const float*)A.sim.actor_root_state : A.sim.net_contact_force) + xoff;
    const size_t nb13 = (size_t)A.p.num_bodies * 13;
    const size_t rb_step = kProdWarps * nb13, x_step = (size_t)kProdWarps * xrow;
    const bool early = (A.flags & DEX_FLAG_EARLY_SIM_LOADS) != 0;
    const int n_rows = A.n_bulk;
    constexpr int R = 32 / kProdWarps;  // env rows per producer warp
    int it = 0;
    for (int tile = blockIdx.x; tile < tiles; tile += workers, ++it) {
      const int s = it & 1, use = it >> 1;
      const int e0 = tile * 32, ne = min(32, A.n - e0);
      unsigned char* stage = smem2 + s * LY::STAGE_BYTES;
      const uint32_t sst = s32(stage), fb = bar_full + 8 * s;
      TL2(it, 16);
      if (use > 0) mbar_wait(bar_empty + 8 * s, (unsigned)(use - 1) & 1u);
      TL2(it, 17);
      const bool full = ne == 32;               // CTA-uniform
      const bool bulk = A.bulk_ok && full;      // the two AoS tiles (dof_state, full_dof_targets) as bulk copies
      const bool split = it == 0 && early;      // first tile: the state rows wait for griddepcontrol.wait, the sim rows do not
      const bool act = lane < ne;
      if (it == 0 && !early) pdl_wait();
      // table entry k: [0, n_bulk_early) sim-tensor tiles (dof_state) | state tiles (full_dof_targets) | state rows
      auto issue_tiles = [&](const int k0, const int k1) {
        for (int k = k0; k < k1; ++k) {
          const int kind = A.bk_kind[k];
          if (kind != BK_DOF && kind != BK_TGT) break;
          const uint32_t len = (uint32_t)A.bk_len16[k] * 16u;
          const char* src = reinterpret_cast<const char*>(A.bk_ptr[k]) + (size_t)tile * len;
          const uint32_t dst = sst + (uint32_t)A.bk_dst16[k] * 16u;
          if (bulk) {
            if (pw == 0 && lane == 0) bulk_g2s(dst, src, len, fb);
          } else {  // ragged tile / unaligned base: 4-byte asynchronous copies of the contiguous block
            const int nfl = ne * (kind == BK_DOF ? 52 : 26);
            for (int i = pw * 32 + lane; i < nfl; i += kProdThreads) cpa4(dst + 4u * i, src + 4 * i);
          }
        }
      };
      issue_tiles(0, A.n_bulk_early);
      // ---- scattered Isaac Gym rows: warp pw takes env rows pw, pw+4, ..; consecutive lanes = consecutive floats ----
      {
        const float* q0 = A.sim.rigid_body_state + (size_t)(e0 + pw) * nb13;
        const float* qx = x_base + (size_t)(e0 + pw) * xrow;
        uint32_t a = sst + LY::OFF_AOS + (uint32_t)(pw * LY::SA + lane) * 4u;
#pragma unroll
        for (int i = 0; i < R; ++i) {
          if (full || pw + kProdWarps * i < ne) {
            if (p_rb0) cpa4(a + LY::A_RB * 4, q0 + rbo0);
            cpa4(a + (LY::A_RB + 32) * 4, q0 + rbo1);
            if (p_rb2) cpa4(a + (LY::A_RB + 64) * 4, q0 + rbo2);
            if (p_x) cpa4(a + LY::A_CF * 4, qx);
          }
          q0 += rb_step;
          qx += x_step;
          a += kProdWarps * LY::SA * 4;
        }
      }
      TL2(it, 18);
      if (split) pdl_wait();  // everything above touched only kernel parameters and the sim tensors
      issue_tiles(A.n_bulk_early, n_rows);
      // ---- SoA state rows: lane = env, one 128-byte line per row; row k of the table goes to warp k % 4 ----
      for (int k = A.n_tile_entries + pw; k < n_rows; k += kProdWarps) {
        const int kind = A.bk_kind[k];
        const char* src = reinterpret_cast<const char*>(A.bk_ptr[k]);
        const uint32_t dst = sst + (uint32_t)A.bk_dst16[k] * 16u;
        if (kind == BK_F32) {
          if (act) cpa4(dst + 4u * lane, src + ((size_t)e0 + lane) * 4);
        } else if (kind == BK_I64) {
          if (act) cpa8(dst + 8u * lane, src + ((size_t)e0 + lane) * 8);
        } else {  // BK_U8: 32 bytes of a bool row
          if (full && (A.n & 3) == 0) {
            if (lane < 8) cpa4(dst + 4u * lane, src + (size_t)e0 + 4 * lane);
          } else if (act) {
            stage[(size_t)A.bk_dst16[k] * 16 + lane] = *reinterpret_cast<const unsigned char*>(src + (size_t)e0 + lane);
          }
        }
      }
      cpa_arrive(fb);  // counted arrival when every asynchronous copy of this thread has landed
      __syncwarp();
      TL2(it, 19);
      if (lane == 0) {
        if (pw == 0 && bulk)
          mbar_arrive_tx(fb, A.tx_bytes);
        else
          mbar_arrive(fb);
      }
    }
    return;
  }

  // =========================== consumer warps ===========================
  Cx<BG, STD> c;
  c.lane = lane;
  c.N = (size_t)A.n;
  c.wk = wk_base + lane;
  c.obs = s_obs + lane * A.wp;
  c.obs_offset = A.p.obs_offset;
  c.dof2afj = A.dof2afj;
  unsigned stamp = 0;
  uint64_t rng_step = 0;
  bool store_pending = false;
  int it = 0;
  for (int tile = blockIdx.x; tile < tiles; tile += workers, ++it) {
    const int s = it & 1, use = it >> 1;
    c.e0 = tile * 32;
    c.ne = min(32, A.n - c.e0);
    c.act = lane < c.ne;
    c.env = (size_t)c.e0 + (c.act ? lane : 0);
    c.stage = smem2 + s * LY::STAGE_BYTES;
    TL2(it, 0);
    mbar_wait(bar_full + 8 * s, (unsigned)use & 1u);
    TL2(it, 1);
    if (c.ne == 32)  // CTA-uniform
      consume_tile<BG, STD, true>(A, c, warp, it, 0u, &s_rng_step, &s_rbits, stamp, rng_step, bar_empty + 8 * s, s_obs, s_reset_u,
                                  store_pending TL2_ARG);
    else
      consume_tile<BG, false, false>(A, reinterpret_cast<Cx<BG, false>&>(c), warp, it, 0u, &s_rng_step, &s_rbits, stamp, rng_step,
                                     bar_empty + 8 * s, s_obs, s_reset_u, store_pending TL2_ARG);
    TL2(it, 6);
  }
  if (store_pending) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the last obs tile has left shared memory
}

template <bool BG, bool STD>
int launch_post2_k(Post2Args& a, cudaStream_t s) {
  static int ctas_per_sm = -1, num_sms = 0;
  static size_t smem_cfg = 0, smem_occ = 0;
  if (a.smem_bytes > smem_cfg) {
    const size_t want = a.smem_bytes > 100 * 1024 ? a.smem_bytes : 100 * 1024;
    const cudaError_t e = cudaFuncSetAttribute(post_physics_v2_kernel<BG, STD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want);
    if (e != cudaSuccess) return (int)e;
    smem_cfg = want;
    ctas_per_sm = -1;
  }
  if (ctas_per_sm < 0 || smem_occ != a.smem_bytes) {
    int dev = 0, occ = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, post_physics_v2_kernel<BG, STD>, kThreads, a.smem_bytes);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return DEX_ERR_SIZE;
    ctas_per_sm = occ;
    smem_occ = a.smem_bytes;
  }
  const int fused = a.fused_epilogue ? 1 : 0;
  int cap = num_sms * ctas_per_sm - fused;  // every CTA of the launch (workers + epilogue) is resident at once
  if (cap < 1) cap = 1;
  a.workers = a.tiles < cap ? a.tiles : cap;
  const cudaError_t e = launch_step_kernel(post_physics_v2_kernel<BG, STD>, a.workers + fused, kThreads, a.smem_bytes, s, a);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

template <bool BG>
int launch_post2_t(Post2Args& a, cudaStream_t s) {
  const int rc = build_post2<BG>(a);
  if (rc) return rc;
  if (a.smem_bytes > 227 * 1024 - 2048) return DEX_ERR_SIZE;
  return std_layout_matches<BG>(a) ? launch_post2_k<BG, true>(a, s) : launch_post2_k<BG, false>(a, s);
}

}  // namespace pp2
}  // namespace
