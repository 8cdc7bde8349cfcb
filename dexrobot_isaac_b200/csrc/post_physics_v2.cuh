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
constexpr int kMaxBulk = kProdThreads;  // one bulk entry per producer thread

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
  // work block: values that cross a warp or a phase, [field][32] floats
  static constexpr int W_BIN = 0, W_PADBOX = 5;           // 5 each
  static constexpr int W_CDIST = 10, W_DRIFT = 11, W_VMAG = 12, W_BOXHAS = 13, W_BOXFAR = 14;
  static constexpr int W_BVEL = 15;                       // 6: finite-difference base velocity
  static constexpr int W_GRASP = 21, W_THUMB = 22, W_OTHERS = 23, W_GRASPDUR = 24, W_SDS = 25, W_STAGE = 26, W_J2 = 27,
                       W_J3 = 28;
  static constexpr int W_HITG = 29, W_HCOK = 30, W_SOK = 31, W_MINTIPBOX = 32, W_AVGHDIFF = 33;
  static constexpr int W_CHG = 34, W_RHAND = 39, W_FV2 = 43, W_FA2 = 63, W_JL = 83;  // BaseTask (common rewards)
  static constexpr int W_WGT = BG ? 34 : 103;  // 26 weighted reward terms
  static constexpr int W_COUNT = W_WGT + DEX_NUM_REWARD_TERMS;
};

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
  uint32_t tx_bytes;                      // bytes of all bulk entries of a full tile
  int32_t wp;                             // obs tile row stride (odd)
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
  a.wp = P.obs_dim | 1;
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
  bool aligned = (a.n % 16) == 0;
  uint32_t tx = 0;
  auto add = [&](const void* base, size_t row, int elem, int dst, int kind) {
    if (k >= kMaxBulk) return;
    const uintptr_t ptr = (uintptr_t)((const char*)base + row * N * elem);
    const int len = kind == BK_DOF ? 32 * 52 * 4 : kind == BK_TGT ? 32 * 26 * 4 : 32 * elem;
    if (ptr & 15u) aligned = false;
    a.bk_ptr[k] = (uint64_t)ptr;
    a.bk_dst16[k] = (uint16_t)(dst / 16);
    a.bk_len16[k] = (uint16_t)(len / 16);
    a.bk_kind[k] = (uint8_t)kind;
    tx += (uint32_t)len;
    ++k;
  };
  add(a.sim.dof_state, 0, 4, LY::OFF_DOF, BK_DOF);
  a.n_bulk_early = k;
  if (!BG && a.want_tgt) add(st.full_dof_targets, 0, 4, LY::OFF_TGT, BK_TGT);
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
  a.smem_bytes = (uint32_t)(kStages * LY::STAGE_BYTES + LY::W_COUNT * 128 + 32 * a.wp * 4);
  return DEX_OK;
}

#ifdef DEXSTEP_TL2  // debug build only (tools/timeline_v2.py): SM-clock marks per (CTA, tile iteration), [cta][kTlIters][32] u64:
// [0..15] consumer thread 0, [16..23] producer thread 0, [30] globaltimer ns at the mark 0 of the iteration, [31] smid
constexpr int kTlIters = 8;
__device__ unsigned long long* g_tl2 = nullptr;
#define TL2_DECL unsigned long long* const tl2 = g_tl2
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
__device__ __forceinline__ void cons_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsThreads) : "memory"); }

__device__ __forceinline__ void obs_put(float* obs, const int off, const int j, const float v) {
  if (off >= 0) obs[off + j] = v;
}

// ------------------------------------------------------------------------------------------------
// producer side
// ------------------------------------------------------------------------------------------------
// generic path (ragged tile, or row bases / N that are not 16-byte friendly): every row with plain loads
template <bool BG>
__device__ __forceinline__ void prod_rows_ldg(const Post2Args& A, unsigned char* stage, const int e0, const int ne, const int pw,
                                              const int lane) {
  const bool act = lane < ne;
  for (int k0 = pw; k0 < A.n_bulk; k0 += kProdWarps * 4) {
    unsigned long long v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + kProdWarps * u;
      v[u] = 0ull;
      if (k < A.n_bulk) {
        const int kind = A.bk_kind[k];
        const char* src = reinterpret_cast<const char*>(A.bk_ptr[k]);
        if (kind == BK_F32) {
          if (act) v[u] = __float_as_uint(*reinterpret_cast<const float*>(src + ((size_t)e0 + lane) * 4));
        } else if (kind == BK_I64) {
          if (act) v[u] = *reinterpret_cast<const unsigned long long*>(src + ((size_t)e0 + lane) * 8);
        } else if (kind == BK_U8) {
          if (act) v[u] = *reinterpret_cast<const unsigned char*>(src + ((size_t)e0 + lane));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = k0 + kProdWarps * u;
      if (k < A.n_bulk && act) {
        const int kind = A.bk_kind[k];
        unsigned char* dst = stage + (size_t)A.bk_dst16[k] * 16;
        if (kind == BK_F32)
          reinterpret_cast<unsigned*>(dst)[lane] = (unsigned)v[u];
        else if (kind == BK_I64)
          reinterpret_cast<unsigned long long*>(dst)[lane] = v[u];
        else if (kind == BK_U8)
          dst[lane] = (unsigned char)v[u];
      }
    }
  }
  // the two AoS tiles are contiguous: flat copies
  for (int k = 0; k < A.n_bulk; ++k) {
    const int kind = A.bk_kind[k];
    if (kind != BK_DOF && kind != BK_TGT) continue;
    const int per = kind == BK_DOF ? 52 : 26;
    const float* src = reinterpret_cast<const float*>(A.bk_ptr[k]) + (size_t)e0 * per;
    float* dst = reinterpret_cast<float*>(stage + (size_t)A.bk_dst16[k] * 16);
    for (int i = pw * 32 + lane; i < ne * per; i += kProdThreads) dst[i] = src[i];
  }
}

// ------------------------------------------------------------------------------------------------
// consumer side (lane = env).  Each of the 8 consumer warps owns a fixed JOB per phase: its code is straight-line
// (loads first, then the independent per-finger / per-DOF chains side by side), so one warp covers its own latencies
// and keeps re-running the same few hundred instructions tile after tile.
// ------------------------------------------------------------------------------------------------
template <bool BG>
struct Cx {
  const unsigned char* stage;
  float* wk;   // work block + lane
  float* obs;  // obs tile row of this lane's env
  int lane, cw, e0, ne;
  bool act;
  size_t N, env;
  __device__ __forceinline__ const float* soa() const { return reinterpret_cast<const float*>(stage) + lane; }
  __device__ __forceinline__ const float* aos() const {
    return reinterpret_cast<const float*>(stage + Lay<BG>::OFF_AOS) + lane * Lay<BG>::SA;
  }
  __device__ __forceinline__ int i64lo(const int r) const {
    return reinterpret_cast<const int*>(stage + Lay<BG>::OFF_I64 + r * 256)[2 * lane];
  }
};

// ---- job: the five tip (or pad) poses: world columns, hand-frame columns (observation_encoder.py:915-986) and, for
//      BlindGrasping, the box distances (blind_grasping_task.py:684-771, 1178-1192, 1334-1364)
template <bool BG, bool PAD, bool WORLD, bool HANDF>
__device__ __forceinline__ void job_poses(const Post2Args& A, const Cx<BG>& c) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const float* aos = c.aos();
  const float* rb = aos + LY::A_RB;
  float* wk = c.wk;
  float* obs = c.obs;
  float p[DEX_NUM_FINGERS][7];
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f)
#pragma unroll
    for (int j = 0; j < 7; ++j) p[f][j] = rb[13 + (PAD ? 35 : 0) + 7 * f + j];
  if (WORLD) {
    const int ow = P.obs_offset[PAD ? DEX_K_FINGERPAD_POSES_WORLD : DEX_K_FINGERTIP_POSES_WORLD];
    if (ow >= 0) {
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f)
#pragma unroll
        for (int j = 0; j < 7; ++j) obs[ow + 7 * f + j] = p[f][j];
    }
  }
  if (HANDF) {
    const int oh = P.obs_offset[PAD ? DEX_K_FINGERPAD_POSES_HAND : DEX_K_FINGERTIP_POSES_HAND];
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
  }
  if (WORLD && !PAD) {
    // hitting_ground inputs (step_processor.py:133-181): min over the tips in finger order
    float min_tip_z = p[0][2];
#pragma unroll
    for (int f = 1; f < DEX_NUM_FINGERS; ++f) min_tip_z = fminf(min_tip_z, p[f][2]);
    const bool hitg = P.has_height_safety && ((rb[2] < P.handbase_threshold) || (min_tip_z < P.fingertip_threshold));
    wk[LY::W_HITG * 32] = hitg ? 1.0f : 0.0f;
  }
  if (BG && WORLD) {
    const float* box = aos + LY::A_BOX;
    const float bx = box[0], by = box[1], bz = box[2];
    float d[DEX_NUM_FINGERS];
#pragma unroll
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) d[f] = norm3(p[f][0] - bx, p[f][1] - by, p[f][2] - bz);
    if (!PAD) {
      float m = d[0];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) m = fminf(m, d[f]);
      wk[LY::W_MINTIPBOX * 32] = m;
    } else {
      const int od = P.obs_offset[DEX_K_FINGER_TO_OBJECT_DISTANCES], ohd = P.obs_offset[DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF];
      float hd[DEX_NUM_FINGERS], dsum = 0.0f, hsum = 0.0f;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        hd[f] = fabsf(p[f][2] - bz);
        wk[(LY::W_PADBOX + f) * 32] = d[f];
        obs_put(obs, od, f, d[f]);
        obs_put(obs, ohd, f, hd[f]);
        dsum += d[f];
        hsum += hd[f];
      }
      const float avg_h = div_rn(hsum, 5.0f);  // blind_grasping_task.py:702-724
      wk[LY::W_AVGHDIFF * 32] = avg_h;
      obs_put(obs, P.obs_offset[DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE], 0, div_rn(dsum, 5.0f));
      obs_put(obs, P.obs_offset[DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF], 0, avg_h);
      // pad centroid + the height / centroid part of _evaluate_pregrasp_quality (:746-771, :1334-1364)
      float cen[3];
#pragma unroll
      for (int k = 0; k < 3; ++k) cen[k] = (p[0][k] + p[1][k] + p[2][k]) / 3.0f;
      const int oc = P.obs_offset[DEX_K_FIRST_THREE_FINGERPAD_CENTROID];
#pragma unroll
      for (int k = 0; k < 3; ++k) obs_put(obs, oc, k, cen[k]);
      const float cdist = norm3(cen[0] - bx, cen[1] - by, cen[2] - bz);
      wk[LY::W_CDIST * 32] = cdist;
      const bool h_ok = (hd[0] <= P.height_tolerance) && (hd[1] <= P.height_tolerance) && (hd[2] <= P.height_tolerance);
      wk[LY::W_HCOK * 32] = (h_ok && (cdist <= P.centroid_tolerance)) ? 1.0f : 0.0f;
    }
  }
}

// ---- job: the five contact sensors (observation_encoder.py:658-735, 323-361; reward_calculator.py:183-199) and, for
//      BlindGrasping, the box contact flag and the box stability inputs
template <bool BG>
__device__ __forceinline__ void job_contacts(const Post2Args& A, const Cx<BG>& c) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float* aos = c.aos();
  const float* soa = c.soa();
  float* wk = c.wk;
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
  const int ocf = P.obs_offset[DEX_K_CONTACT_FORCES], om = P.obs_offset[DEX_K_CONTACT_FORCE_MAGNITUDE];
  const int ob = P.obs_offset[DEX_K_CONTACT_BINARY], od = P.obs_offset[DEX_K_CONTACT_DURATION];
  const bool common = !BG && A.common != 0, first = A.first != 0;
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
    const float mag = norm3(cf[f][0], cf[f][1], cf[f][2]);
    const bool bin = mag > P.contact_binary_threshold;
    const float ndur = (bin && prevb[f] == 0.0f) ? 1.0f : (bin ? dur[f] + 1.0f : 0.0f);
    const float fb = bin ? 1.0f : 0.0f;
    st.contact_duration_steps[f * N + env] = ndur;
    st.prev_contact_binary[f * N + env] = fb;
    if (st.contact_forces != nullptr) {
#pragma unroll
      for (int k = 0; k < 3; ++k) st.contact_forces[(3 * f + k) * N + env] = cf[f][k];
    }
    wk[(LY::W_BIN + f) * 32] = fb;
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
      wk[(LY::W_CHG + f) * 32] = (c_now != c_prev) ? 1.0f : 0.0f;
      st.prev_contacts[f * N + env] = c_now ? 1 : 0;
    }
  }
  if (BG) {
    const float* cfb = aos + LY::A_CF + 15;
    const float* box = aos + LY::A_BOX;
    const float b[6] = {box[0], box[1], box[2], box[3], box[4], box[5]};
    wk[LY::W_BOXHAS * 32] = (norm3(cfb[0], cfb[1], cfb[2]) > P.contact_binary_threshold) ? 1.0f : 0.0f;
    const float drift = norm3(b[0] - soa[(LY::R_BOX0 + 0) * 32], b[1] - soa[(LY::R_BOX0 + 1) * 32], b[2] - soa[(LY::R_BOX0 + 2) * 32]);
    const float vmag = norm3(b[3], b[4], b[5]);
    wk[LY::W_DRIFT * 32] = drift;
    wk[LY::W_VMAG * 32] = vmag;
    wk[LY::W_SOK * 32] = ((drift <= P.position_drift_tolerance) && (vmag <= P.velocity_tolerance)) ? 1.0f : 0.0f;
    const int opos = P.obs_offset[DEX_K_OBJECT_POS], ovel = P.obs_offset[DEX_K_OBJECT_VEL];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      obs_put(obs, opos, k, b[k]);
      obs_put(obs, ovel, k, b[3 + k]);
    }
  }
}

// ---- job: hand pose, ARR-aligned pose (observation_encoder.py:1483-1521), hand -> box distance, hand velocity
//      reward terms (reward_calculator.py:100-181), episode time (dexhand_base.py:703-712: PRE-increment count)
template <bool BG>
__device__ __forceinline__ void job_hand(const Post2Args& A, const Cx<BG>& c) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float* aos = c.aos();
  const float* soa = c.soa();
  const float* rb = aos + LY::A_RB;
  float* wk = c.wk;
  float* obs = c.obs;
  float h[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) h[j] = rb[j];
  obs_put(obs, P.obs_offset[DEX_K_EPISODE_TIME], 0, (float)c.i64lo(LY::I_STEP) * P.control_dt);
  const int oh = P.obs_offset[DEX_K_HAND_POSE];
  if (oh >= 0) {
#pragma unroll
    for (int j = 0; j < 7; ++j) obs[oh + j] = h[j];
  }
  const int oa = P.obs_offset[DEX_K_HAND_POSE_ARR_ALIGNED];
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
    wk[LY::W_BOXFAR * 32] = (hb > P.max_box_distance) ? 1.0f : 0.0f;
    obs_put(obs, P.obs_offset[DEX_K_HAND_TO_OBJECT_DISTANCE], 0, hb);
  }
  if (!BG && A.common) {
    const bool first = A.first != 0;
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
    wk[(LY::W_RHAND + 0) * 32] = expf(-0.2f * norm3(hv[0], hv[1], hv[2]));
    wk[(LY::W_RHAND + 1) * 32] = expf(-0.2f * norm3(ha[0], ha[1], ha[2]));
    wk[(LY::W_RHAND + 2) * 32] = expf(-0.5f * norm3(hv[0] - phv[0], hv[1] - phv[1], hv[2] - phv[2]));
    wk[(LY::W_RHAND + 3) * 32] = expf(-0.5f * norm3(ha[0] - pha[0], ha[1] - pha[1], ha[2] - pha[2]));
  }
}

// ---- job: the ten pad pair distances, triu order (blind_grasping_task.py:726-771)
template <bool BG>
__device__ __forceinline__ void job_pairs(const Post2Args& A, const Cx<BG>& c) {
  using LY = Lay<BG>;
  const int op = A.p.obs_offset[DEX_K_FINGERPAD_DISTANCES];
  if (op < 0) return;
  const float* rb = c.aos() + LY::A_RB;
  float q[DEX_NUM_FINGERS][3];
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f)
#pragma unroll
    for (int k = 0; k < 3; ++k) q[f][k] = rb[13 + 35 + 7 * f + k];
  int k = 0;
#pragma unroll
  for (int a = 0; a < DEX_NUM_FINGERS; ++a)
#pragma unroll
    for (int b = a + 1; b < DEX_NUM_FINGERS; ++b, ++k)
      c.obs[op + k] = norm3(q[a][0] - q[b][0], q[a][1] - q[b][1], q[a][2] - q[b][2]);
}

// ---- job: DOFs [D0, D1): finite-difference velocity (observation_encoder.py:298-321), obs columns, common-reward
//      partial terms (reward_calculator.py:94-98,138-160: Isaac dof_vel, not the finite difference)
template <bool BG, int D0, int D1>
__device__ __forceinline__ void job_dofs(const Post2Args& A, const Cx<BG>& c) {
  using LY = Lay<BG>;
  static_assert(D0 % 2 == 0 && D1 % 2 == 0, "whole (pos, vel) pairs of DOFs");
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float* soa = c.soa();
  float* wk = c.wk;
  float* obs = c.obs;
  const size_t N = c.N, env = c.env;
  const float dt = P.control_dt;
  constexpr int ND = D1 - D0;
  float pos[ND], vel[ND], prev[ND], tgt[ND];
  const float4* d4 = reinterpret_cast<const float4*>(c.stage + LY::OFF_DOF) + c.lane * 13;
#pragma unroll
  for (int q = 0; q < ND / 2; ++q) {
    const float4 v = d4[D0 / 2 + q];
    pos[2 * q] = v.x;
    vel[2 * q] = v.y;
    pos[2 * q + 1] = v.z;
    vel[2 * q + 1] = v.w;
  }
  const bool want_tgt = !BG && A.want_tgt != 0;
#pragma unroll
  for (int i = 0; i < ND; ++i) {
    prev[i] = soa[(LY::R_PDP + D0 + i) * 32];
    tgt[i] = want_tgt ? reinterpret_cast<const float*>(c.stage + LY::OFF_TGT)[c.lane * 26 + D0 + i] : 0.0f;
  }
  const bool common = !BG && A.common != 0, first = A.first != 0;
  const int o_bp = P.obs_offset[DEX_K_BASE_DOF_POS], o_bv = P.obs_offset[DEX_K_BASE_DOF_VEL], o_bt = P.obs_offset[DEX_K_BASE_DOF_TARGET];
  const int o_ap = P.obs_offset[DEX_K_ALL_FINGER_DOF_POS], o_av = P.obs_offset[DEX_K_ALL_FINGER_DOF_VEL],
            o_at = P.obs_offset[DEX_K_ALL_FINGER_DOF_TARGET];
  const int o_fp = P.obs_offset[DEX_K_ACTIVE_FINGER_DOF_POS], o_fv = P.obs_offset[DEX_K_ACTIVE_FINGER_DOF_VEL],
            o_ft = P.obs_offset[DEX_K_ACTIVE_FINGER_DOF_TARGET];
#pragma unroll
  for (int i = 0; i < ND; ++i) {
    const int d = D0 + i;
    const float fd = div_rn(pos[i] - prev[i], dt);
    st.prev_dof_pos[d * N + env] = pos[i];
    if (d < DEX_NUM_BASE) {
      obs_put(obs, o_bp, d, pos[i]);
      obs_put(obs, o_bv, d, fd);
      if (!BG) obs_put(obs, o_bt, d, tgt[i]);
      if (BG) wk[(LY::W_BVEL + d) * 32] = fd;
    } else {
      const int j = d - DEX_NUM_BASE;
      obs_put(obs, o_ap, j, pos[i]);
      obs_put(obs, o_av, j, fd);
      if (!BG) obs_put(obs, o_at, j, tgt[i]);
      const int aj = A.dof2afj[d];
      if (aj >= 0) {
        obs_put(obs, o_fp, aj, pos[i]);
        obs_put(obs, o_fv, aj, fd);
        if (!BG) obs_put(obs, o_ft, aj, tgt[i]);
      }
      if (common) {
        const float iv = vel[i];
        const float pvv = first ? iv : soa[(LY::R_PFDV + j) * 32];
        wk[(LY::W_FV2 + j) * 32] = iv * iv;
        const float acc = iv - pvv;
        wk[(LY::W_FA2 + j) * 32] = acc * acc;
        const float nj = div_rn(2.0f * (pos[i] - P.dof_lower[d]), P.dof_upper[d] - P.dof_lower[d]) - 1.0f;
        wk[(LY::W_JL + j) * 32] = clampf(fabsf(nj) - 0.8f, 0.0f, 1.0f);
        st.prev_finger_dof_vel[j * N + env] = iv;
      }
    }
  }
}

// ---- job: prev_actions columns (observation_encoder.py:287-296); active_rule_targets <- active_prev_targets
//      (rules.py:78-95, step_processor.py:73)
template <bool BG>
__device__ __forceinline__ void job_copies(const Post2Args& A, const Cx<BG>& c) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const float* soa = c.soa();
  float* obs = c.obs;
  float apt[DEX_NUM_ACTIVE];
#pragma unroll
  for (int k = 0; k < DEX_NUM_ACTIVE; ++k) apt[k] = soa[(LY::R_APT + k) * 32];
  const int op = P.obs_offset[DEX_K_PREV_ACTIONS];
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
  const int o1 = P.obs_offset[DEX_K_ACTIVE_PREV_TARGETS], o2 = P.obs_offset[DEX_K_ACTIVE_RULE_TARGETS];
  float* art = A.st.active_rule_targets + c.env;
#pragma unroll
  for (int k = 0; k < DEX_NUM_ACTIVE; ++k) {
    art[(size_t)k * c.N] = apt[k];
    obs_put(obs, o1, k, apt[k]);
    obs_put(obs, o2, k, apt[k]);
  }
}

// phase A: everything that depends only on the staged tile
template <bool BG>
__device__ __forceinline__ void phase_a(const Post2Args& A, const Cx<BG>& c) {
  if (!c.act) return;
  switch (c.cw) {
    case 0: job_poses<BG, false, true, BG>(A, c); break;   // tips: world (+ hand frame for BlindGrasping configs that ask)
    case 1: job_poses<BG, true, true, BG>(A, c); break;    // pads
    case 2: job_contacts<BG>(A, c); break;
    case 3: job_hand<BG>(A, c); break;
    case 4:
      if (BG) job_pairs<BG>(A, c);
      else job_poses<BG, false, false, true>(A, c);         // BaseTask: tip poses in the hand frame
      break;
    case 5: job_dofs<BG, 0, 14>(A, c); break;
    case 6: job_dofs<BG, 14, 26>(A, c); break;
    default:
      job_copies<BG>(A, c);
      if (!BG) job_poses<BG, true, false, true>(A, c);      // BaseTask: pad poses in the hand frame
      break;
  }
}

// one common reward term (reward_calculator.py:66-221); BaseTask
template <bool BG>
__device__ __forceinline__ float common_term(const Post2Args& A, const Cx<BG>& c, const int i) {
  using LY = Lay<BG>;
  const float* wk = c.wk;
  const float* rb = c.aos() + LY::A_RB;
  switch (i) {
    case DEX_R_ALIVE: return 1.0f;
    case DEX_R_HEIGHT_SAFETY: {
      float mz = rb[13 + 2];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) mz = fminf(mz, rb[13 + 7 * f + 2]);
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
template <bool BG>
__device__ __forceinline__ float task_term(const Post2Args& A, const Cx<BG>& c, const int i, const int stage) {
  using LY = Lay<BG>;
  const DexParams& P = A.p;
  const float* wk = c.wk;
  const float* aos = c.aos();
  const float* rb = aos + LY::A_RB;
  const float m1 = (stage == 1) ? 1.0f : 0.0f, m2 = (stage == 2) ? 1.0f : 0.0f, m3 = (stage == 3) ? 1.0f : 0.0f;
  switch (i) {
    case DEX_R_S1_HEIGHT_ALIGNMENT: return expf(-P.height_alignment_decay * wk[LY::W_AVGHDIFF * 32]) * m1;
    case DEX_R_S1_CENTROID_POSITIONING: return expf(-P.centroid_positioning_decay * wk[LY::W_CDIST * 32]) * m1;
    case DEX_R_S1_OBJECT_STABILITY:
      return expf(-P.object_stability_decay * (wk[LY::W_DRIFT * 32] + wk[LY::W_VMAG * 32])) * m1;
    case DEX_R_S1_FINGER_HEIGHT_CONSISTENCY: {  // torch.var(dim=1), unbiased, z of the first three pads
      const float z0 = rb[13 + 35 + 2], z1 = rb[13 + 35 + 7 + 2], z2 = rb[13 + 35 + 14 + 2];
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
      return clampf(div_rn(aos[LY::A_BOX + 2] - P.box_z, P.lift_height_range), 0.0f, 1.0f) * m3;
    case DEX_R_S3_GRASP_MAINTENANCE: return wk[LY::W_GRASP * 32] * m3;
    case DEX_R_S3_GRASP_DURATION:
      return clampf(div_rn(wk[LY::W_GRASPDUR * 32], P.contact_duration_threshold_s), 0.0f, 1.0f) * m3;
    // completion bonuses :1025-1048 (transition and not failed), penetration :1194-1208
    case DEX_R_S1_COMPLETION:
      return (wk[LY::W_J2 * 32] != 0.0f && wk[LY::W_HCOK * 32] != 0.0f && wk[LY::W_SOK * 32] != 0.0f) ? 1.0f : 0.0f;
    case DEX_R_S2_COMPLETION: return (wk[LY::W_J3 * 32] != 0.0f && wk[LY::W_GRASP * 32] != 0.0f) ? 1.0f : 0.0f;
    default:  // DEX_R_PENETRATION_PENALTY
      return fmaxf(P.penetration_threshold - wk[LY::W_MINTIPBOX * 32], 0.0f) * P.penetration_depth_scale;
  }
}

// raw + weighted rows of term i leave from registers; the weighted value is parked for the ordered total
template <bool BG>
__device__ __forceinline__ void emit_term(const Post2Args& A, const Cx<BG>& c, const int i, const float r) {
  const float w = A.p.reward_weight[i];
  const int row = A.p.component_row[i];
  const float rw = r * w;
  float* g = A.out.reward_components + c.env;
  g[(size_t)row * c.N] = r;
  g[(size_t)(row + 1) * c.N] = rw;
  c.wk[(Lay<BG>::W_WGT + i) * 32] = rw;
}

// phase B (BlindGrasping): the two scalar chains that the reward terms and the termination wait for
template <bool BG>
__device__ __forceinline__ void phase_b(const Post2Args& A, const Cx<BG>& c) {
  using LY = Lay<BG>;
  if (!c.act) return;
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const float dt = P.control_dt;
  float* wk = c.wk;
  float* obs = c.obs;
  const float* soa = c.soa();
  const size_t env = c.env;
  if (c.cw == 0) {
    // ---- finger-box contacts (blind_grasping_task.py:773-833), success duration (:626-648) ----
    const bool box_has = wk[LY::W_BOXHAS * 32] != 0.0f;
    const float box_z = c.aos()[LY::A_BOX + 2];
    const int sds_old = c.i64lo(LY::I_SDS);
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
    obs_put(obs, P.obs_offset[DEX_K_THUMB_CONTACT], 0, ft);
    obs_put(obs, P.obs_offset[DEX_K_OTHER_FINGERS_CONTACT], 0, fo);
    obs_put(obs, P.obs_offset[DEX_K_GRASP_STATE], 0, fg);
    const bool cond = (box_z > P.success_height_threshold) && (n_on_box >= P.min_fingers_for_grasp);
    const int sds = cond ? sds_old + 1 : 0;
    st.success_duration_steps[env] = (long long)sds;
    st.success_conditions_met[env] = cond ? 1 : 0;
    const float gd = (float)sds * dt;
    wk[LY::W_SDS * 32] = __int_as_float(sds);
    wk[LY::W_GRASPDUR * 32] = gd;
    obs_put(obs, P.obs_offset[DEX_K_GRASP_DURATION], 0, gd);
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
    obs_put(obs, P.obs_offset[DEX_K_CURRENT_STAGE], 0, (float)stage);
    obs_put(obs, P.obs_offset[DEX_K_TIME_IN_STAGE], 0, tis);
    obs_put(obs, P.obs_offset[DEX_K_STAGE_PROGRESS], 0, prog);
  }
}

// phase C, warp 0: termination (step_processor.py:133-181, termination_manager.py:120-278), masks, tile record.
// Returns the exclusive termination class of this lane's env: bit 0 success, 1 failure, 2 timeout.
template <bool BG>
__device__ __forceinline__ unsigned phase_combine(const Post2Args& A, const Cx<BG>& c, const unsigned stamp, unsigned* s_rbits) {
  using LY = Lay<BG>;
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
      const bool qual_ok = (wk[LY::W_HCOK * 32] != 0.0f) && (wk[LY::W_SOK * 32] != 0.0f);
      const int stage = __float_as_int(wk[LY::W_STAGE * 32]);
      const int sds = __float_as_int(wk[LY::W_SDS * 32]);
      crit[DEX_C_GRASP_LIFT_SUCCESS] = sds >= P.contact_duration_threshold_steps;
      crit[DEX_C_BOX_TOO_FAR] = wk[LY::W_BOXFAR * 32] != 0.0f;
      const bool pregrasp_failed = j2 && !qual_ok;  // _evaluate_pregrasp_quality :1334-1364
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
    *s_rbits = rbits;
  }
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

// phase D, warp 0: ordered weighted sum (reward_calculator.py:223-275) + termination rewards
// (termination_manager.py:280-321, step_processor.py:204-219)
template <bool BG>
__device__ __forceinline__ void phase_total(const Post2Args& A, const Cx<BG>& c, const unsigned cls) {
  using LY = Lay<BG>;
  if (!c.act) return;
  const DexParams& P = A.p;
  const float* wk = c.wk;
  constexpr int NT = BG ? DEX_NUM_REWARD_TERMS : DEX_NUM_COMMON_TERMS;
  float wv[NT];
#pragma unroll
  for (int i = 0; i < NT; ++i) wv[i] = wk[(LY::W_WGT + i) * 32];  // terms with weight 0 hold stale values: skipped below
  float total = 0.0f;
#pragma unroll
  for (int i = 0; i < NT; ++i)
    if (P.reward_weight[i] != 0.0f) total += wv[i];  // dict order: common, then task
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

// obs tile -> obs_buf for the 32-column chunks in `chunks`: the env rows are spread over NW warps (this one is number w),
// lane = column: every global store is one full 128-byte request.  Row pointers are formed once; the chunk offset is
// an immediate.
template <int NW>
__device__ __forceinline__ void copy_obs(const float* s_obs, float* g_tile, const int W, const int WP, const int ne,
                                         const unsigned chunks, const int w, const int lane) {
  constexpr int R = (32 + NW - 1) / NW;
  const float* sp[R];
  float* gp[R];
  bool ok[R];
#pragma unroll
  for (int i = 0; i < R; ++i) {
    const int row = w + NW * i;
    ok[i] = row < ne;
    sp[i] = s_obs + row * WP + lane;
    gp[i] = g_tile + (size_t)row * W + lane;
  }
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) {  // obs_dim <= 256
    if (!((chunks >> ch) & 1u)) continue;
    if (32 * ch + lane < W) {
      float v[R];
#pragma unroll
      for (int i = 0; i < R; ++i)
        if (ok[i]) v[i] = sp[i][32 * ch];
#pragma unroll
      for (int i = 0; i < R; ++i)
        if (ok[i]) gp[i][32 * ch] = v[i];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------------
template <bool BG>
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
      mbar_init(bar_full + 8 * s, kProdWarps);
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
    const int pw = warp - kConsWarps, ptid = tid - kConsThreads;
    // this thread's bulk entry (one row / tile array per producer thread)
    const bool e_has = ptid < A.n_bulk, e_early = ptid < A.n_bulk_early;
    const char* const e_ptr = e_has ? reinterpret_cast<const char*>(A.bk_ptr[ptid]) : nullptr;
    const uint32_t e_dst = e_has ? (uint32_t)A.bk_dst16[ptid] * 16u : 0u;
    const uint32_t e_len = e_has ? (uint32_t)A.bk_len16[ptid] * 16u : 0u;
    // scattered Isaac Gym rows: lane = gathered float (rigid slots lane, lane + 32, lane + 64; contact / box slot lane)
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
      const bool bulk = A.bulk_ok && ne == 32;  // CTA-uniform
      if (it == 0 && !early) pdl_wait();
      if (bulk && e_has && e_early) bulk_g2s(sst + e_dst, e_ptr + (size_t)tile * e_len, e_len, fb);
      if (bulk && e_has && !e_early && !(it == 0 && early)) bulk_g2s(sst + e_dst, e_ptr + (size_t)tile * e_len, e_len, fb);
      // ---- scattered rows: warp pw takes env rows pw, pw+4, ..; consecutive lanes = consecutive floats ----
      float v[R][4];
      {
        const float* q0 = A.sim.rigid_body_state + (size_t)(e0 + pw) * nb13;
        const float* qx = x_base + (size_t)(e0 + pw) * xrow;
        if (ne == 32) {
#pragma unroll
          for (int i = 0; i < R; ++i) {
            if (p_rb0) v[i][0] = __ldg(q0 + rbo0);
            v[i][1] = __ldg(q0 + rbo1);
            if (p_rb2) v[i][2] = __ldg(q0 + rbo2);
            if (p_x) v[i][3] = *qx;  // actor_root_state rows are rewritten by the fused reset of this kernel: plain load
            q0 += rb_step;
            qx += x_step;
          }
        } else {
#pragma unroll
          for (int i = 0; i < R; ++i) {
            if (pw + kProdWarps * i < ne) {
              if (p_rb0) v[i][0] = __ldg(q0 + rbo0);
              v[i][1] = __ldg(q0 + rbo1);
              if (p_rb2) v[i][2] = __ldg(q0 + rbo2);
              if (p_x) v[i][3] = *qx;
            }
            q0 += rb_step;
            qx += x_step;
          }
        }
      }
      TL2(it, 18);
      if (it == 0 && early) {
        pdl_wait();  // everything above touched only kernel parameters and the sim tensors
        if (bulk && e_has && !e_early) bulk_g2s(sst + e_dst, e_ptr + (size_t)tile * e_len, e_len, fb);
      }
      if (!bulk) prod_rows_ldg<BG>(A, stage, e0, ne, pw, lane);
      {
        float* a = reinterpret_cast<float*>(stage + LY::OFF_AOS) + pw * LY::SA + lane;
#pragma unroll
        for (int i = 0; i < R; ++i) {
          if (ne == 32 || pw + kProdWarps * i < ne) {
            if (p_rb0) a[LY::A_RB] = v[i][0];
            a[LY::A_RB + 32] = v[i][1];
            if (p_rb2) a[LY::A_RB + 64] = v[i][2];
            if (p_x) a[LY::A_CF] = v[i][3];
          }
          a += kProdWarps * LY::SA;
        }
      }
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
  Cx<BG> c;
  c.lane = lane;
  c.cw = warp;
  c.N = (size_t)A.n;
  c.wk = wk_base + lane;
  c.obs = s_obs + lane * A.wp;
  unsigned stamp = 0;
  uint64_t rng_step = 0;
  const int W = A.p.obs_dim, WP = A.wp;
  const unsigned all_chunks = (W >= 256) ? 0xffu : ((1u << ((W + 31) >> 5)) - 1u);
  const unsigned late_chunks = BG ? (A.obs_late & all_chunks) : all_chunks;  // BaseTask has no phase B: one copy at the end
  const unsigned early_chunks = all_chunks & ~late_chunks;
  int it = 0;
  for (int tile = blockIdx.x; tile < tiles; tile += workers, ++it) {
    const int s = it & 1, use = it >> 1;
    c.e0 = tile * 32;
    c.ne = min(32, A.n - c.e0);
    c.act = lane < c.ne;
    c.env = (size_t)c.e0 + (c.act ? lane : 0);
    c.stage = smem2 + s * LY::STAGE_BYTES;
    float* const g_obs = A.out.obs_buf + (size_t)c.e0 * W;
    TL2(it, 0);
    mbar_wait(bar_full + 8 * s, (unsigned)use & 1u);
    TL2(it, 1);
    // Philox step counter == stamp of this step's tile records: read once per CTA, behind the first full barrier
    // (i.e. behind the producers' griddepcontrol.wait; the epilogue CTA advances it only after every record of this
    // launch has been published), by one thread, and parked while phase A runs
    unsigned long long rs_tmp = 0ull;
    if (it == 0 && tid == 0) rs_tmp = A.st.rng_step ? *reinterpret_cast<const volatile unsigned long long*>(A.st.rng_step) : A.step;

    phase_a<BG>(A, c);
    if (it == 0 && tid == 0) s_rng_step = rs_tmp;
    TL2(it, 2);
    cons_sync();
    TL2(it, 3);
    if (it == 0) {
      rng_step = s_rng_step;
      stamp = ((unsigned)rng_step + 1u) & 0xffffu;
    }
    if (BG) {
      // phase B: warps 0 / 1 run the two scalar chains, the other six send off the obs chunks that are already final
      if (warp < 2)
        phase_b<BG>(A, c);
      else
        copy_obs<kConsWarps - 2>(s_obs, g_obs, W, WP, c.ne, early_chunks, warp - 2, lane);
      TL2(it, 4);
      cons_sync();
      TL2(it, 5);
    }
    // phase C: warp 0 termination + tile record; warps 1..7 the reward terms
    unsigned cls = 0;
    if (warp == 0) {
      cls = phase_combine<BG>(A, c, stamp, &s_rbits);
    } else if (c.act) {
      if (BG) {
        const int stage_now = __float_as_int(c.wk[LY::W_STAGE * 32]);
        for (int i = DEX_NUM_COMMON_TERMS + (warp - 1); i < DEX_NUM_REWARD_TERMS; i += kConsWarps - 1) {
          if (A.p.reward_weight[i] == 0.0f) continue;  // reward_calculator.py:247-251: zero-weight terms are not materialised
          emit_term<BG>(A, c, i, task_term<BG>(A, c, i, stage_now));
        }
      } else {
        for (int i = warp - 1; i < DEX_NUM_COMMON_TERMS; i += kConsWarps - 1) {
          if (A.p.reward_weight[i] == 0.0f) continue;
          emit_term<BG>(A, c, i, common_term<BG>(A, c, i));
        }
      }
    }
    // the raw stage is not read after this point: hand it back to the producers
    __syncwarp();
    TL2(it, 6);
    cons_sync();
    TL2(it, 7);
    if (lane == 0) mbar_arrive(bar_empty + 8 * s);
    if (warp == 0) phase_total<BG>(A, c, cls);
    TL2(it, 8);
    copy_obs<kConsWarps>(s_obs, g_obs, W, WP, c.ne, late_chunks, warp, lane);
    TL2(it, 9);
    // ---- fused reset of the finished envs (dexhand_base.py:743-756): one warp per env ----
    {
      unsigned mask = s_rbits;  // written by warp 0 before the barrier above
      if (mask && !(A.flags & DEX_FLAG_NO_FUSED_RESET)) {
        for (int k = 0; mask; ++k) {
          const int e = __ffs(mask) - 1;
          mask &= mask - 1;
          if ((k & (kConsWarps - 1)) == warp) reset_env_warp<BG>(A, c.e0 + e, rng_step, s_reset_u + 32 * warp);
        }
      }
    }
    TL2(it, 10);
    cons_sync();  // work block, obs tile and s_rbits are reused by the next tile
    TL2(it, 11);
  }
}

template <bool BG>
int launch_post2_t(Post2Args& a, cudaStream_t s) {
  const int rc = build_post2<BG>(a);
  if (rc) return rc;
  static int ctas_per_sm = -1, num_sms = 0;
  static size_t smem_cfg = 0;
  if (a.smem_bytes > smem_cfg) {
    const size_t want = a.smem_bytes > 120 * 1024 ? a.smem_bytes : 120 * 1024;
    const cudaError_t e = cudaFuncSetAttribute(post_physics_v2_kernel<BG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want);
    if (e != cudaSuccess) return (int)e;
    smem_cfg = want;
    ctas_per_sm = -1;
  }
  static size_t smem_occ = 0;
  if (ctas_per_sm < 0 || smem_occ != a.smem_bytes) {
    int dev = 0, occ = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, post_physics_v2_kernel<BG>, kThreads, a.smem_bytes);
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return DEX_ERR_SIZE;
    ctas_per_sm = occ;
    smem_occ = a.smem_bytes;
  }
  const int fused = a.fused_epilogue ? 1 : 0;
  int cap = num_sms * ctas_per_sm - fused;  // every CTA of the launch (workers + epilogue) is resident at once
  if (cap < 1) cap = 1;
  a.workers = a.tiles < cap ? a.tiles : cap;
  const cudaError_t e = launch_step_kernel(post_physics_v2_kernel<BG>, a.workers + fused, kThreads, a.smem_bytes, s, a);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

}  // namespace pp2
}  // namespace
