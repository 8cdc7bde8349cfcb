// post_physics.cuh — the fused post-physics kernel (StepProcessor.process_physics_step, step_processor.py:37-131).
//
// Mapping: one CTA of kPostThreads (8 warps) owns a tile of 32 consecutive envs.
//   phase 0  stage-in   every input of the tile is copied global -> shared with cp.async (LDGSTS): AoS sim
//                       buffers row-per-warp (consecutive lanes = consecutive floats of one env row, so each
//                       request is one or two sectors), SoA state rows lane-per-env (128 B per request).
//                       All copies are in flight at once; one wait.
//   phase 1  items      the per-env work is cut into independent *items* (one DOF, one fingertip/pad pose, one
//                       contact sensor, one pad pair, ...).  Warps take items round-robin; inside a warp
//                       lane = env, so every branch on the item type is warp-uniform and every shared-memory
//                       access `slab[lane * stride + const]` (stride odd) is bank-conflict free.  Items write
//                       observation columns into the shared obs tile, state rows straight to global (128 B
//                       coalesced) and a few per-env scalars into a scratch area.
//   phase 2  scalar     warp 0 (lane = env) runs the short serial part: finger-box contact logic, stage FSM,
//                       termination, reward terms, masks, counters (warp ballots), fused reset of finished envs.
//   phase 3  stage-out  obs tile shared -> global, row-per-warp, fully coalesced.
//   phase 4  epilogue   optional: the last CTA to finish (atomic ticket) compacts the reset mask into the
//                       ordered env-id list and finalises the termination statistics (no extra launch).
#pragma once
#include "dexstep_common.cuh"

namespace {

constexpr int kTile = 32;          // envs per CTA == warp width (lane = env in phases 1-2)
constexpr int kPostThreads = 256;  // 8 warps
constexpr int kPostWarps = kPostThreads / 32;
constexpr int kRbW = 13 + 10 * 7;  // hand row (13) + 5 tips + 5 pads (7 each) = 83

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
};

// Per-env slab in shared memory (float offsets).  Computed identically on host (smem size) and device.
struct Slab {
  int dof;   // 52  dof_state row (pos, vel interleaved)
  int rb;    // 83  hand row + 5 tip + 5 pad poses
  int cf;    // 18  5 monitored contact forces + box contact force
  int box;   // 6   box pos + linvel
  int vel;   // 26  prev_dof_pos in, finite-difference velocity out (in place)
  int st;    // 49  active_prev_targets 18 | prev_actions 18 | contact_duration 5 | prev_contact_binary 5 | box0 3
  int tgt;   // 26  full_dof_targets row (only when a *_dof_target key is wanted)
  int cpv;   // 26  common-reward prev state: finger dof vel 20 | hand vel 3 | hand ang vel 3
  int obs;   // obs_dim
  int scr;   // scratch scalars
  int stride;
};

// scratch slots
enum {
  SC_CBIN = 0,      // 5 contact_binary (0/1)
  SC_CMAG = 5,      // 5 contact force magnitude
  SC_CHG = 10,      // 5 contact change flags (common rewards)
  SC_PADBOX = 15,   // 5 pad -> box distance
  SC_HDIFF = 20,    // 5 |pad_z - box_z|
  SC_TIPBOX = 25,   // 5 tip -> box distance
  SC_HANDBOX = 30,  // hand -> box distance
  SC_CDIST = 31,    // |centroid - box|
  SC_DRIFT = 32,    // |box - box0|
  SC_VMAG = 33,     // |v_box|
  SC_BOXHAS = 34,   // box contact force above threshold (0/1)
  SC_RHAND = 35,    // 4 hand velocity / acceleration reward terms (common)
  SC_BASE = 39,
  SC_FV2 = 39,      // 20 squared finger dof velocities (common)
  SC_FA2 = 59,      // 20 squared finger dof accelerations
  SC_JL = 79,       // 20 joint-limit terms
  SC_COMMON_END = 99
};

__host__ __device__ inline bool post_want_targets(const DexParams& P, bool dict) {
  return dict || P.obs_offset[DEX_K_BASE_DOF_TARGET] >= 0 || P.obs_offset[DEX_K_ACTIVE_FINGER_DOF_TARGET] >= 0 ||
         P.obs_offset[DEX_K_ALL_FINGER_DOF_TARGET] >= 0;
}

__host__ __device__ inline Slab make_slab(const DexParams& P, bool dict) {
  Slab s;
  int o = 0;
  s.dof = o; o += 2 * DEX_NUM_DOF;
  s.rb = o; o += kRbW;
  s.cf = o; o += 18;
  s.box = o; o += 6;
  s.vel = o; o += DEX_NUM_DOF;
  s.st = o; o += 49;
  s.tgt = o; o += post_want_targets(P, dict) ? DEX_NUM_DOF : 0;
  s.cpv = o; o += P.need_common_rewards ? 26 : 0;
  s.obs = o; o += P.obs_dim;
  s.scr = o; o += P.need_common_rewards ? SC_COMMON_END : SC_BASE;
  s.stride = o | 1;  // odd stride: lane-per-env accesses hit 32 different banks
  return s;
}

__host__ __device__ inline size_t post_smem_bytes(const DexParams& P, bool dict) {
  return (size_t)make_slab(P, dict).stride * kTile * sizeof(float);
}

__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

// rigid-body slot (0..82) -> float offset inside one env's [NB,13] slab
__device__ __forceinline__ int rb_slot_offset(const DexParams& P, const int s) {
  if (s < 13) return P.hand_body * 13 + s;
  const int t = s - 13, f = t / 7, c = t - f * 7;
  int body = 0;
#pragma unroll
  for (int k = 0; k < DEX_NUM_FINGERS; ++k) {
    if (f == k) body = P.tip_body[k];
    if (f == k + DEX_NUM_FINGERS) body = P.pad_body[k];
  }
  return body * 13 + c;
}

// ------------------------------------------------------------------------------------------------
// phase 0: stage-in
// ------------------------------------------------------------------------------------------------
template <bool BG>
__device__ __forceinline__ void post_stage_in(const PostArgs& A, const Slab& L, float* smem, const int e0, const int ne,
                                              const bool want_tgt) {
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t N = (size_t)A.n;
  const int S = L.stride;
  const bool full = A.mode == DEX_MODE_FULL;
  const bool common = full && P.need_common_rewards;

  // ---- SoA state rows: lane = env, warps take rows round-robin (each request = one 128 B line) ----
  if (lane < ne) {
    float* E = smem + lane * S;
    const size_t col = (size_t)e0 + lane;
    int r = warp;
    for (; r < DEX_NUM_DOF; r += kPostWarps) cp_async4(E + L.vel + r, st.prev_dof_pos + r * N + col);
    r -= DEX_NUM_DOF;
    for (; r < DEX_NUM_ACTIVE; r += kPostWarps) cp_async4(E + L.st + r, st.active_prev_targets + r * N + col);
    r -= DEX_NUM_ACTIVE;
    for (; r < P.num_actions; r += kPostWarps) cp_async4(E + L.st + 18 + r, st.prev_actions + r * N + col);
    r -= P.num_actions;
    for (; r < DEX_NUM_FINGERS; r += kPostWarps) cp_async4(E + L.st + 36 + r, st.contact_duration_steps + r * N + col);
    r -= DEX_NUM_FINGERS;
    for (; r < DEX_NUM_FINGERS; r += kPostWarps) cp_async4(E + L.st + 41 + r, st.prev_contact_binary + r * N + col);
    r -= DEX_NUM_FINGERS;
    if (BG) {
      for (; r < 3; r += kPostWarps) cp_async4(E + L.st + 46 + r, st.initial_box_positions + r * N + col);
      r -= 3;
    }
    if (common) {
      for (; r < DEX_NUM_FINGER_DOF; r += kPostWarps) cp_async4(E + L.cpv + r, st.prev_finger_dof_vel + r * N + col);
      r -= DEX_NUM_FINGER_DOF;
      for (; r < 3; r += kPostWarps) cp_async4(E + L.cpv + 20 + r, st.prev_hand_vel + r * N + col);
      r -= 3;
      for (; r < 3; r += kPostWarps) cp_async4(E + L.cpv + 23 + r, st.prev_hand_ang_vel + r * N + col);
      r -= 3;
    }
    if (A.flags & DEX_FLAG_STALE_CONTACTS) {
      for (; r < 15; r += kPostWarps) cp_async4(E + L.cf + r, st.contact_forces + r * N + col);
    }
  }

  // ---- AoS sim rows: warps take env rows round-robin, consecutive lanes = consecutive floats ----
  const int nb = P.num_bodies;
  // per-lane gather offsets (depend on the lane only): rigid slots lane, lane+32, lane+64; contact slot lane
  const int rbo0 = rb_slot_offset(P, lane), rbo1 = rb_slot_offset(P, lane + 32);
  const int rbo2 = (lane + 64 < kRbW) ? rb_slot_offset(P, lane + 64) : 0;
  const bool need_hand_vel = common;  // hand velocities only feed the common reward terms
  int cfo = 0;
  {
    const int f = lane / 3, c = lane - f * 3;
    int body = 0;
#pragma unroll
    for (int k = 0; k < DEX_NUM_FINGERS; ++k)
      if (f == k) body = P.contact_body[k];
    if (BG && f == DEX_NUM_FINGERS) body = P.box_body;
    cfo = body * 3 + c;
  }
  const int ncf = BG ? 18 : 15;
  for (int e = warp; e < ne; e += kPostWarps) {
    float* E = smem + e * S;
    const size_t env = (size_t)e0 + e;
    {  // dof_state row: 52 contiguous floats
      const float* src = A.sim.dof_state + env * (2 * DEX_NUM_DOF);
      cp_async4(E + L.dof + lane, src + lane);
      if (lane + 32 < 2 * DEX_NUM_DOF) cp_async4(E + L.dof + lane + 32, src + lane + 32);
    }
    {  // rigid-body rows
      const float* src = A.sim.rigid_body_state + env * (size_t)nb * 13;
      if (lane < 7 || lane >= 13 || need_hand_vel) cp_async4(E + L.rb + lane, src + rbo0);
      cp_async4(E + L.rb + lane + 32, src + rbo1);
      if (lane + 64 < kRbW) cp_async4(E + L.rb + lane + 64, src + rbo2);
    }
    if (lane < ncf && (lane >= 15 || !(A.flags & DEX_FLAG_STALE_CONTACTS)))
      cp_async4(E + L.cf + lane, A.sim.net_contact_force + env * (size_t)nb * 3 + cfo);
    if (BG && lane < 6)
      cp_async4(E + L.box + lane,
                A.sim.actor_root_state + (env * P.num_actors + P.box_actor) * 13 + (lane < 3 ? lane : lane + 4));
    if (want_tgt && lane < DEX_NUM_DOF) cp_async4(E + L.tgt + lane, st.full_dof_targets + env * DEX_NUM_DOF + lane);
  }
  cp_async_wait_all();
}

// ------------------------------------------------------------------------------------------------
// phase 1: items (lane = env, item type warp-uniform)
// ------------------------------------------------------------------------------------------------
template <bool DICT>
struct Emitter {
  const DexParams& P;
  float* obs;   // shared obs row of this env
  float* dict;  // global obs_dict row of this env (DICT only)
  __device__ __forceinline__ void operator()(const int key, const int j, const float v) const {
    const int o = P.obs_offset[key];
    if (o >= 0) obs[o + j] = v;
    if (DICT) dict[P.dict_offset[key] + j] = v;
  }
  __device__ __forceinline__ bool want(const int key) const { return DICT || P.obs_offset[key] >= 0; }
};

template <bool BG, bool DICT>
__device__ __forceinline__ void post_items(const PostArgs& A, const Slab& L, float* smem, const int* s_dof2fin,
                                           const int e0, const int ne, const bool want_tgt) {
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane >= ne) return;
  const size_t N = (size_t)A.n;
  const size_t env = (size_t)e0 + lane;
  float* E = smem + lane * L.stride;
  float* scr = E + L.scr;
  const Emitter<DICT> emit{P, E + L.obs, DICT ? A.out.obs_dict + env * P.dict_dim : nullptr};
  const float dt = P.control_dt;
  const bool full = A.mode == DEX_MODE_FULL;
  const bool common = full && P.need_common_rewards;
  const bool first = (A.flags & DEX_FLAG_FIRST_REWARD_CALL) != 0;
  const float* rb = E + L.rb;

  // item ranges (heavy items first so the round-robin spreads them over the warps)
  const int i_pose = 0, i_contact = 10, i_pair = 15, i_hand = BG ? 25 : 15, i_box = i_hand + 1;
  const int i_dof = BG ? i_box + 1 : i_box, i_apt = i_dof + DEX_NUM_DOF, i_pact = i_apt + DEX_NUM_ACTIVE;
  const int i_tgt = i_pact + P.num_actions, n_items = i_tgt + (want_tgt ? DEX_NUM_DOF : 0);

  for (int it = warp; it < n_items; it += kPostWarps) {
    if (it < i_contact) {
      // ---- tip / pad pose: world + hand frame (observation_encoder.py:737-751, 915-986) ----
      const int f = it - i_pose;
      const bool is_pad = f >= DEX_NUM_FINGERS;
      const int g = is_pad ? f - DEX_NUM_FINGERS : f;
      const float* p = rb + 13 + 7 * f;
      const float px = p[0], py = p[1], pz = p[2];
      const Quat q{p[3], p[4], p[5], p[6]};
      const int kw = is_pad ? DEX_K_FINGERPAD_POSES_WORLD : DEX_K_FINGERTIP_POSES_WORLD;
      const int kh = is_pad ? DEX_K_FINGERPAD_POSES_HAND : DEX_K_FINGERTIP_POSES_HAND;
      emit(kw, 7 * g + 0, px);
      emit(kw, 7 * g + 1, py);
      emit(kw, 7 * g + 2, pz);
      emit(kw, 7 * g + 3, q.x);
      emit(kw, 7 * g + 4, q.y);
      emit(kw, 7 * g + 5, q.z);
      emit(kw, 7 * g + 6, q.w);
      if (emit.want(kh)) {
        const Quat hq{rb[3], rb[4], rb[5], rb[6]};
        const Vec3 ph = quat_rotate_inverse(hq, Vec3{px - rb[0], py - rb[1], pz - rb[2]});
        const Quat qh = quat_mul(quat_conj(hq), q);
        emit(kh, 7 * g + 0, ph.x);
        emit(kh, 7 * g + 1, ph.y);
        emit(kh, 7 * g + 2, ph.z);
        emit(kh, 7 * g + 3, qh.x);
        emit(kh, 7 * g + 4, qh.y);
        emit(kh, 7 * g + 5, qh.z);
        emit(kh, 7 * g + 6, qh.w);
      }
      if (BG) {  // blind_grasping_task.py:684-724, 1178-1192
        const float* box = E + L.box;
        const float d = norm3(px - box[0], py - box[1], pz - box[2]);
        if (is_pad) {
          const float hd = fabsf(pz - box[2]);
          scr[SC_PADBOX + g] = d;
          scr[SC_HDIFF + g] = hd;
          emit(DEX_K_FINGER_TO_OBJECT_DISTANCES, g, d);
          emit(DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF, g, hd);
        } else {
          scr[SC_TIPBOX + g] = d;
        }
      }
    } else if (it < i_pair) {
      // ---- contact sensor f (observation_encoder.py:658-735, 323-361; reward_calculator.py:183-199) ----
      const int f = it - i_contact;
      const float* cf = E + L.cf + 3 * f;
      const float fx = cf[0], fy = cf[1], fz = cf[2];
      emit(DEX_K_CONTACT_FORCES, 3 * f, fx);
      emit(DEX_K_CONTACT_FORCES, 3 * f + 1, fy);
      emit(DEX_K_CONTACT_FORCES, 3 * f + 2, fz);
      if (st.contact_forces != nullptr && !(A.flags & DEX_FLAG_STALE_CONTACTS)) {
        st.contact_forces[(size_t)(3 * f) * N + env] = fx;
        st.contact_forces[(size_t)(3 * f + 1) * N + env] = fy;
        st.contact_forces[(size_t)(3 * f + 2) * N + env] = fz;
      }
      const float mag = norm3(fx, fy, fz);
      const bool bin = mag > P.contact_binary_threshold;
      emit(DEX_K_CONTACT_FORCE_MAGNITUDE, f, mag);
      emit(DEX_K_CONTACT_BINARY, f, bin ? 1.0f : 0.0f);
      const float dur = E[L.st + 36 + f], prevb = E[L.st + 41 + f];
      const float ndur = (bin && prevb == 0.0f) ? 1.0f : (bin ? dur + 1.0f : 0.0f);
      st.contact_duration_steps[f * N + env] = ndur;
      st.prev_contact_binary[f * N + env] = bin ? 1.0f : 0.0f;
      emit(DEX_K_CONTACT_DURATION, f, ndur * dt);
      scr[SC_CBIN + f] = bin ? 1.0f : 0.0f;
      scr[SC_CMAG + f] = mag;
      if (common) {
        const bool c_now = mag > 0.1f;  // hard-coded 0.1 N (reward_calculator.py:186)
        const bool c_prev = first ? c_now : (st.prev_contacts[f * N + env] != 0);
        st.prev_contacts[f * N + env] = c_now ? 1 : 0;
        scr[SC_CHG + f] = (c_now != c_prev) ? 1.0f : 0.0f;
      }
    } else if (BG && it < i_hand) {
      // ---- pad pair distance, triu order (blind_grasping_task.py:726-771) ----
      const int k = it - i_pair;
      int a = 0, b = 0, c = 0;
#pragma unroll
      for (int x = 0; x < DEX_NUM_FINGERS; ++x)
#pragma unroll
        for (int y = x + 1; y < DEX_NUM_FINGERS; ++y) {
          if (c == k) {
            a = x;
            b = y;
          }
          ++c;
        }
      const float* pa = rb + 13 + 35 + 7 * a;
      const float* pb = rb + 13 + 35 + 7 * b;
      emit(DEX_K_FINGERPAD_DISTANCES, k, norm3(pa[0] - pb[0], pa[1] - pb[1], pa[2] - pb[2]));
    } else if (it == i_hand) {
      // ---- hand pose (:632-656), ARR-aligned pose (:1483-1521), hand velocity reward terms ----
#pragma unroll
      for (int c = 0; c < 7; ++c) emit(DEX_K_HAND_POSE, c, rb[c]);
      if (emit.want(DEX_K_HAND_POSE_ARR_ALIGNED)) {
        const float sh = 0.7071067811865476f;
        const Quat qa = quat_mul(Quat{rb[3], rb[4], rb[5], rb[6]}, Quat{-0.0f, -sh, -0.0f, sh});
        emit(DEX_K_HAND_POSE_ARR_ALIGNED, 0, rb[0]);
        emit(DEX_K_HAND_POSE_ARR_ALIGNED, 1, rb[1]);
        emit(DEX_K_HAND_POSE_ARR_ALIGNED, 2, rb[2]);
        emit(DEX_K_HAND_POSE_ARR_ALIGNED, 3, qa.x);
        emit(DEX_K_HAND_POSE_ARR_ALIGNED, 4, qa.y);
        emit(DEX_K_HAND_POSE_ARR_ALIGNED, 5, qa.z);
        emit(DEX_K_HAND_POSE_ARR_ALIGNED, 6, qa.w);
      }
      if (BG) {
        const float* box = E + L.box;
        const float d = norm3(rb[0] - box[0], rb[1] - box[1], rb[2] - box[2]);
        scr[SC_HANDBOX] = d;
        emit(DEX_K_HAND_TO_OBJECT_DISTANCE, 0, d);
      }
      if (common) {  // reward_calculator.py:100-181
        float hv[3], ha[3], phv[3], pha[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          hv[c] = rb[7 + c];
          ha[c] = rb[10 + c];
          phv[c] = first ? hv[c] : E[L.cpv + 20 + c];
          pha[c] = first ? ha[c] : E[L.cpv + 23 + c];
          st.prev_hand_vel[c * N + env] = hv[c];
          st.prev_hand_ang_vel[c * N + env] = ha[c];
        }
        scr[SC_RHAND + 0] = expf(-0.2f * norm3(hv[0], hv[1], hv[2]));
        scr[SC_RHAND + 1] = expf(-0.2f * norm3(ha[0], ha[1], ha[2]));
        scr[SC_RHAND + 2] = expf(-0.5f * norm3(hv[0] - phv[0], hv[1] - phv[1], hv[2] - phv[2]));
        scr[SC_RHAND + 3] = expf(-0.5f * norm3(ha[0] - pha[0], ha[1] - pha[1], ha[2] - pha[2]));
      }
    } else if (BG && it == i_box) {
      // ---- box state, pad centroid, stability inputs (blind_grasping_task.py:569-600, 746-771, 1334-1364) ----
      const float* box = E + L.box;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        emit(DEX_K_OBJECT_POS, c, box[c]);
        emit(DEX_K_OBJECT_VEL, c, box[3 + c]);
      }
      float cen[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        cen[c] = (rb[13 + 35 + c] + rb[13 + 35 + 7 + c] + rb[13 + 35 + 14 + c]) / 3.0f;
        emit(DEX_K_FIRST_THREE_FINGERPAD_CENTROID, c, cen[c]);
      }
      scr[SC_CDIST] = norm3(cen[0] - box[0], cen[1] - box[1], cen[2] - box[2]);
      const float* b0 = E + L.st + 46;
      scr[SC_DRIFT] = norm3(box[0] - b0[0], box[1] - b0[1], box[2] - b0[2]);
      scr[SC_VMAG] = norm3(box[3], box[4], box[5]);
      const float* cfb = E + L.cf + 15;
      scr[SC_BOXHAS] = (norm3(cfb[0], cfb[1], cfb[2]) > P.contact_binary_threshold) ? 1.0f : 0.0f;
    } else if (it < i_apt) {
      // ---- DOF d: finite-difference velocity (:298-321), position / velocity slices (:612-630) ----
      const int d = it - i_dof;
      const float pos = E[L.dof + 2 * d];
      const float v = (pos - E[L.vel + d]) / dt;
      E[L.vel + d] = v;
      st.prev_dof_pos[d * N + env] = pos;
      if (d < DEX_NUM_BASE) {
        emit(DEX_K_BASE_DOF_POS, d, pos);
        emit(DEX_K_BASE_DOF_VEL, d, v);
      } else {
        const int j = d - DEX_NUM_BASE;
        emit(DEX_K_ALL_FINGER_DOF_POS, j, pos);
        emit(DEX_K_ALL_FINGER_DOF_VEL, j, v);
        if (common) {  // reward_calculator.py:94-98,138-160: Isaac dof_vel, not the finite difference
          const float iv = E[L.dof + 2 * d + 1];
          const float pv = first ? iv : E[L.cpv + j];
          st.prev_finger_dof_vel[j * N + env] = iv;
          scr[SC_FV2 + j] = iv * iv;
          const float acc = iv - pv;
          scr[SC_FA2 + j] = acc * acc;
          const float nj = 2.0f * (pos - P.dof_lower[d]) / (P.dof_upper[d] - P.dof_lower[d]) - 1.0f;
          scr[SC_JL + j] = clampf(fabsf(nj) - 0.8f, 0.0f, 1.0f);
        }
      }
      const int af = s_dof2fin[d];
      if (af >= 0) {
        emit(DEX_K_ACTIVE_FINGER_DOF_POS, af, pos);
        emit(DEX_K_ACTIVE_FINGER_DOF_VEL, af, v);
      }
    } else if (it < i_pact) {
      // ---- active target a: identity pre-action rule (rules.py:78-95) ----
      const int a = it - i_apt;
      const float t = E[L.st + a];
      st.active_rule_targets[a * N + env] = t;
      emit(DEX_K_ACTIVE_PREV_TARGETS, a, t);
      emit(DEX_K_ACTIVE_RULE_TARGETS, a, t);
    } else if (it < i_tgt) {
      const int k = it - i_pact;
      emit(DEX_K_PREV_ACTIONS, k, E[L.st + 18 + k]);
    } else {
      // ---- DOF target d (:664-701) ----
      const int d = it - i_tgt;
      const float t = E[L.tgt + d];
      if (d < DEX_NUM_BASE)
        emit(DEX_K_BASE_DOF_TARGET, d, t);
      else
        emit(DEX_K_ALL_FINGER_DOF_TARGET, d - DEX_NUM_BASE, t);
      const int af = s_dof2fin[d];
      if (af >= 0) emit(DEX_K_ACTIVE_FINGER_DOF_TARGET, af, t);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// phase 2: per-env scalar part (warp 0, lane = env)
// ------------------------------------------------------------------------------------------------
template <bool BG, bool DICT>
__device__ __forceinline__ void post_scalar(const PostArgs& A, const Slab& L, float* smem, const int e0, const int ne) {
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const int lane = threadIdx.x & 31;
  const bool active = lane < ne;
  const size_t N = (size_t)A.n;
  const size_t env = (size_t)e0 + lane;
  float* E = smem + lane * L.stride;
  const float* scr = E + L.scr;
  const float* rb = E + L.rb;
  const float dt = P.control_dt;
  const bool full = A.mode == DEX_MODE_FULL;

  bool should_reset = false, t_success = false, t_failure = false, t_timeout = false;
  int ep_len_done = 0;
  bool crit[DEX_NUM_CRITERIA];
#pragma unroll
  for (int c = 0; c < DEX_NUM_CRITERIA; ++c) crit[c] = false;

  if (active) {
    const Emitter<DICT> emit{P, E + L.obs, DICT ? A.out.obs_dict + env * P.dict_dim : nullptr};
    // episode_time from the PRE-increment step count (dexhand_base.py:703-712)
    const int64_t step_before = st.episode_step_count[env];
    emit(DEX_K_EPISODE_TIME, 0, (float)step_before * dt);
    float min_tip_z = rb[13 + 2];
#pragma unroll
    for (int f = 1; f < DEX_NUM_FINGERS; ++f) min_tip_z = fminf(min_tip_z, rb[13 + 7 * f + 2]);

    // ---------------- BlindGrasping task observations + FSM (blind_grasping_task.py:549-978) ----
    float avg_hdiff = 0.0f, grasp_duration = 0.0f, min_tip_box = 0.0f, box_z = 0.0f;
    bool thumb_c = false, others_c = false, grasp = false, j2 = false, j3 = false;
    int64_t stage = 0, sds = 0;
    if (BG) {
      box_z = E[L.box + 2];
      float dsum = 0.0f, hsum = 0.0f;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        dsum += scr[SC_PADBOX + f];
        hsum += scr[SC_HDIFF + f];
      }
      avg_hdiff = hsum / 5.0f;
      emit(DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE, 0, dsum / 5.0f);
      emit(DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF, 0, avg_hdiff);
      // finger-box contacts :773-833
      const bool box_has = scr[SC_BOXHAS] != 0.0f;
      bool cbin[DEX_NUM_FINGERS], fbc[DEX_NUM_FINGERS];
      int n_on_box = 0;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        cbin[f] = scr[SC_CBIN + f] != 0.0f;
        fbc[f] = cbin[f] && box_has && (scr[SC_PADBOX + f] < P.proximity_threshold);
        n_on_box += fbc[f] ? 1 : 0;
      }
      thumb_c = fbc[0];
      others_c = fbc[1] || fbc[2] || fbc[3] || fbc[4];
      grasp = thumb_c && others_c;
      emit(DEX_K_THUMB_CONTACT, 0, thumb_c ? 1.0f : 0.0f);
      emit(DEX_K_OTHER_FINGERS_CONTACT, 0, others_c ? 1.0f : 0.0f);
      emit(DEX_K_GRASP_STATE, 0, grasp ? 1.0f : 0.0f);
      // success duration :626-648
      const bool cond = (box_z > P.success_height_threshold) && (n_on_box >= P.min_fingers_for_grasp);
      sds = cond ? st.success_duration_steps[env] + 1 : 0;
      st.success_duration_steps[env] = sds;
      st.success_conditions_met[env] = cond ? 1 : 0;
      grasp_duration = (float)sds * dt;
      emit(DEX_K_GRASP_DURATION, 0, grasp_duration);
      // stage FSM :835-956 (plain fp32 accumulation of (float)control_dt decides the transition step)
      stage = st.current_stage[env];
      float tis = st.time_in_stage[env] + dt;
      float scd = st.stage_contact_duration[env];
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
      st.current_stage[env] = stage;
      st.time_in_stage[env] = tis;
      st.stage_contact_duration[env] = scd;
      st.just_transitioned_to_stage2[env] = j2 ? 1 : 0;
      st.just_transitioned_to_stage3[env] = j3 ? 1 : 0;
      float prog = 0.0f;  // :958-978
      if (stage == 1) prog = clampf(tis / P.stage1_duration, 0.0f, 1.0f);
      if (stage == 2) prog = clampf(tis / P.stage2_duration, 0.0f, 1.0f);
      if (stage == 3) prog = 1.0f;
      emit(DEX_K_CURRENT_STAGE, 0, (float)stage);
      emit(DEX_K_TIME_IN_STAGE, 0, tis);
      emit(DEX_K_STAGE_PROGRESS, 0, prog);
      min_tip_box = scr[SC_TIPBOX];
#pragma unroll
      for (int f = 1; f < DEX_NUM_FINGERS; ++f) min_tip_box = fminf(min_tip_box, scr[SC_TIPBOX + f]);
    }

    if (full) {
      // step_processor.py:80
      const int64_t step_now = step_before + 1;
      st.episode_step_count[env] = step_now;
      ep_len_done = (int)step_now;

      // ---------------- termination (step_processor.py:133-181, termination_manager.py:120-278) ----
      if (P.has_height_safety)
        crit[DEX_C_HITTING_GROUND] = (rb[2] < P.handbase_threshold) || (min_tip_z < P.fingertip_threshold);
      bool pregrasp_failed = false, contact_failed = false, grasp_lost = false;
      if (BG) {
        crit[DEX_C_GRASP_LIFT_SUCCESS] = sds >= (int64_t)P.contact_duration_threshold_steps;
        crit[DEX_C_BOX_TOO_FAR] = scr[SC_HANDBOX] > P.max_box_distance;
        if (j2) {  // _evaluate_pregrasp_quality :1334-1364
          const bool h_ok = (scr[SC_HDIFF + 0] <= P.height_tolerance) && (scr[SC_HDIFF + 1] <= P.height_tolerance) &&
                            (scr[SC_HDIFF + 2] <= P.height_tolerance);
          const bool c_ok = scr[SC_CDIST] <= P.centroid_tolerance;
          const bool s_ok = (scr[SC_DRIFT] <= P.position_drift_tolerance) && (scr[SC_VMAG] <= P.velocity_tolerance);
          pregrasp_failed = !(h_ok && c_ok && s_ok);
        }
        if (j3) contact_failed = !grasp;
        if (stage == 3) grasp_lost = !grasp;
        crit[DEX_C_STAGE1_PREGRASP_FAILED] = pregrasp_failed;
        crit[DEX_C_STAGE2_CONTACT_FAILED] = contact_failed;
        crit[DEX_C_STAGE3_GRASP_LOST] = grasp_lost;
        st.failure_tensors[0 * N + env] = pregrasp_failed ? 1 : 0;
        st.failure_tensors[1 * N + env] = contact_failed ? 1 : 0;
        st.failure_tensors[2 * N + env] = grasp_lost ? 1 : 0;
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
      const bool timeout = step_now >= (int64_t)(P.max_episode_length - 1);
      should_reset = ep_s || ep_f || timeout;
      t_success = ep_s;
      t_failure = ep_f && !ep_s;
      t_timeout = timeout && !ep_s && !ep_f;

      // ---------------- rewards (reward_calculator.py:66-275, blind_grasping_task.py:980-1208) ----
      float r[DEX_NUM_REWARD_TERMS];
#pragma unroll
      for (int i = 0; i < DEX_NUM_REWARD_TERMS; ++i) r[i] = 0.0f;
      if (P.need_common_rewards) {
        r[DEX_R_ALIVE] = 1.0f;
        r[DEX_R_HEIGHT_SAFETY] = clampf(1.0f - expf(-(min_tip_z - 0.02f) * 20.0f), 0.0f, 1.0f);
        float fv2 = 0.0f, fa2 = 0.0f, jl = 0.0f;
#pragma unroll
        for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) {
          fv2 += scr[SC_FV2 + j];
          fa2 += scr[SC_FA2 + j];
          jl += scr[SC_JL + j];
        }
        r[DEX_R_FINGER_VELOCITY] = expf(-0.1f * sqrtf(fv2));
        r[DEX_R_JOINT_LIMIT] = 1.0f - jl / 20.0f;
        r[DEX_R_FINGER_ACCELERATION] = expf(-2.0f * sqrtf(fa2));
        r[DEX_R_HAND_VELOCITY] = scr[SC_RHAND + 0];
        r[DEX_R_HAND_ANGULAR_VELOCITY] = scr[SC_RHAND + 1];
        r[DEX_R_HAND_ACCELERATION] = scr[SC_RHAND + 2];
        r[DEX_R_HAND_ANGULAR_ACCELERATION] = scr[SC_RHAND + 3];
        float changes = 0.0f;
#pragma unroll
        for (int f = 0; f < DEX_NUM_FINGERS; ++f) changes += scr[SC_CHG + f];
        r[DEX_R_CONTACT_STABILITY] = expf(-changes);
      }
      if (BG) {
        const float m1 = (stage == 1) ? 1.0f : 0.0f, m2 = (stage == 2) ? 1.0f : 0.0f, m3 = (stage == 3) ? 1.0f : 0.0f;
        // stage 1 :1055-1112
        r[DEX_R_S1_HEIGHT_ALIGNMENT] = expf(-P.height_alignment_decay * avg_hdiff) * m1;
        r[DEX_R_S1_CENTROID_POSITIONING] = expf(-P.centroid_positioning_decay * scr[SC_CDIST]) * m1;
        r[DEX_R_S1_OBJECT_STABILITY] = expf(-P.object_stability_decay * (scr[SC_DRIFT] + scr[SC_VMAG])) * m1;
        {  // torch.var(dim=1), unbiased, over the z of the first three pads
          const float z0 = rb[13 + 35 + 2], z1 = rb[13 + 35 + 7 + 2], z2 = rb[13 + 35 + 14 + 2];
          const float mean = (z0 + z1 + z2) / 3.0f;
          const float d0 = z0 - mean, d1 = z1 - mean, d2 = z2 - mean;
          const float var = (d0 * d0 + d1 * d1 + d2 * d2) / 2.0f;
          r[DEX_R_S1_FINGER_HEIGHT_CONSISTENCY] = expf(-P.height_consistency_decay * var) * m1;
        }
        r[DEX_R_S1_THUMB_ROTATION] = expf(-5.0f * fabsf(E[L.dof + 2 * DEX_NUM_BASE] - P.half_pi)) * m1;
        // stage 2 :1114-1150
        r[DEX_R_S2_THUMB_CONTACT] = (thumb_c ? 1.0f : 0.0f) * m2;
        r[DEX_R_S2_OTHER_FINGERS_CONTACT] = (others_c ? 1.0f : 0.0f) * m2;
        r[DEX_R_S2_GRASP_ACHIEVEMENT] = (grasp ? 1.0f : 0.0f) * m2;
        r[DEX_R_S2_FINGERPAD_PROXIMITY] = expf(-P.fingerpad_proximity_decay * fmaxf(min_tip_box, P.min_reward_distance)) * m2;
        {
          float bv2 = 0.0f;
#pragma unroll
          for (int d = 0; d < DEX_NUM_BASE; ++d) bv2 += E[L.vel + d] * E[L.vel + d];  // finite-difference base velocity
          r[DEX_R_S2_BASE_STABILITY] = expf(-P.base_stability_decay * sqrtf(bv2)) * m2;
        }
        // stage 3 :1152-1176
        r[DEX_R_S3_OBJECT_HEIGHT] = clampf((box_z - P.box_z) / P.lift_height_range, 0.0f, 1.0f) * m3;
        r[DEX_R_S3_GRASP_MAINTENANCE] = (grasp ? 1.0f : 0.0f) * m3;
        r[DEX_R_S3_GRASP_DURATION] = clampf(grasp_duration / P.contact_duration_threshold_s, 0.0f, 1.0f) * m3;
        // completion bonuses :1025-1048, penetration :1194-1208
        r[DEX_R_S1_COMPLETION] = (j2 && !pregrasp_failed) ? 1.0f : 0.0f;
        r[DEX_R_S2_COMPLETION] = (j3 && !contact_failed) ? 1.0f : 0.0f;
        r[DEX_R_PENETRATION_PENALTY] = fmaxf(P.penetration_threshold - min_tip_box, 0.0f) * P.penetration_depth_scale;
      }
      // weighted sum in dict order (reward_calculator.py:223-275)
      float* comp = A.out.reward_components;
      float total = 0.0f;
#pragma unroll
      for (int i = 0; i < DEX_NUM_REWARD_TERMS; ++i) {
        if (!BG && i >= DEX_NUM_COMMON_TERMS) break;
        const float w = P.reward_weight[i];
        if (w != 0.0f) {
          const float weighted = r[i] * w;
          total += weighted;
          comp[(size_t)P.component_row[i] * N + env] = r[i];
          comp[(size_t)(P.component_row[i] + 1) * N + env] = weighted;
        }
      }
      comp[(size_t)P.total_row * N + env] = total;
      // termination rewards (termination_manager.py:280-321, step_processor.py:204-219)
      const float rs = t_success ? P.success_reward : 0.0f;
      const float rf = t_failure ? P.failure_reward : 0.0f;
      const float rt = t_timeout ? P.timeout_reward : 0.0f;
      float rew = total;
      rew += rs;
      rew += rf;
      rew += rt;
      {
        const size_t tr = (size_t)P.termination_row;
        comp[(tr + 0) * N + env] = t_success ? 1.0f : 0.0f;
        comp[(tr + 1) * N + env] = rs;
        comp[(tr + 2) * N + env] = t_failure ? 1.0f : 0.0f;
        comp[(tr + 3) * N + env] = rf;
        comp[(tr + 4) * N + env] = t_timeout ? 1.0f : 0.0f;
        comp[(tr + 5) * N + env] = rt;
      }
      A.out.rew_buf[env] = rew;
      A.out.reset_buf[env] = should_reset ? 1 : 0;
      A.out.success[env] = t_success ? 1 : 0;
      A.out.failure[env] = t_failure ? 1 : 0;
      A.out.timeout[env] = t_timeout ? 1 : 0;

      // ---------------- fused RewardComponentObserver (rl/reward_observer.py:160-293) -----------
      if (A.out.episode_sums != nullptr) {
        const int rows = P.num_component_rows;
        const int len = A.out.episode_lengths[env] + 1;
        const int stride = 2 + rows;
        double* tot = A.out.stat_totals;
        const int cls = t_success ? 1 : (t_failure ? 2 : 3);
        if (should_reset) {
          atomicAdd(tot + 0, 1.0);
          atomicAdd(tot + 1, (double)len);
          atomicAdd(tot + cls * stride + 0, 1.0);
          atomicAdd(tot + cls * stride + 1, (double)len);
        }
        for (int c = 0; c < rows; ++c) {
          const float s = A.out.episode_sums[(size_t)c * N + env] + comp[(size_t)c * N + env];
          if (should_reset) {
            atomicAdd(tot + 2 + c, (double)s);
            atomicAdd(tot + cls * stride + 2 + c, (double)s);
          }
          A.out.episode_sums[(size_t)c * N + env] = should_reset ? 0.0f : s;
        }
        A.out.episode_lengths[env] = should_reset ? 0 : len;
      }

      // ---------------- fused reset (dexhand_base.py:743-756) -----------------------------------
      if (should_reset && !(A.flags & DEX_FLAG_NO_FUSED_RESET)) {
        const ResetCtx rc{A.uniforms, A.seed, st.rng_step ? *st.rng_step : A.step};
        reset_env<BG>(P, A.sim, st, rc, (int)env, N);
      }
    }
  }

  if (full) {
    // reset ballot word + cross-env counters (termination_manager.py:164-185,259-266)
    const unsigned fullmask = 0xffffffffu;
    const unsigned rbits = __ballot_sync(fullmask, should_reset);
    if (lane == 0) A.out.reset_bits[e0 >> 5] = rbits;
    const unsigned bs = __ballot_sync(fullmask, t_success), bf = __ballot_sync(fullmask, t_failure),
                   bt = __ballot_sync(fullmask, t_timeout);
    unsigned bc[DEX_NUM_CRITERIA];
#pragma unroll
    for (int c = 0; c < DEX_NUM_CRITERIA; ++c) bc[c] = __ballot_sync(fullmask, crit[c]);
    if (lane == 0) {
      if (bs) atomicAdd(A.out.counters + DEX_CNT_SUCCESS, __popc(bs));
      if (bf) atomicAdd(A.out.counters + DEX_CNT_FAILURE, __popc(bf));
      if (bt) atomicAdd(A.out.counters + DEX_CNT_TIMEOUT, __popc(bt));
      if (rbits) atomicAdd(A.out.counters + DEX_CNT_RESET, __popc(rbits));
#pragma unroll
      for (int c = 0; c < DEX_NUM_CRITERIA; ++c)
        if (bc[c]) atomicAdd(A.out.counters + DEX_CNT_CRIT0 + c, __popc(bc[c]));
    }
    if (rbits) {  // sum of the finished episodes' lengths (warp-uniform branch)
      int len = should_reset ? ep_len_done : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) len += __shfl_xor_sync(fullmask, len, o);
      if (lane == 0) atomicAdd(A.out.counters + DEX_CNT_EPISODE_LENGTH, len);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// step epilogue body: ordered compaction of the reset mask (== torch.nonzero(reset_buf),
// step_processor.py:109-111), termination rates (termination_manager.py:259-266), consecutive
// successes (:323-339), running totals.  Runs in ONE CTA of T threads: either the stand-alone
// epilogue kernel or the last CTA of the post-physics kernel.
// ------------------------------------------------------------------------------------------------
template <int T>
__device__ void epilogue_body(const DexEpilogueBuffers& ep, const int n, const int max_consecutive, int* s_warp,
                              int* s_base) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = T / 32;
  const int words = (n + 31) >> 5;
  if (tid == 0) *s_base = 0;
  __syncthreads();
  for (int w0 = 0; w0 < words; w0 += T) {
    const int w = w0 + tid;
    unsigned bits = (w < words) ? __ldcg(ep.reset_bits + w) : 0u;
    const int cnt = __popc(bits);
    int incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int v = (lane < NW) ? s_warp[lane] : 0;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
      }
      if (lane < NW) s_warp[lane] = v;
    }
    __syncthreads();
    int pos = *s_base + (warp > 0 ? s_warp[warp - 1] : 0) + incl - cnt;
    while (bits) {
      const int b = __ffs(bits) - 1;
      ep.env_ids[pos++] = w * 32 + b;
      bits &= bits - 1;
    }
    __syncthreads();
    if (tid == 0) *s_base += s_warp[NW - 1];
    __syncthreads();
  }
  if (tid == 0) {
    *ep.num_reset = *s_base;
    int32_t c[DEX_NUM_COUNTERS];
#pragma unroll
    for (int k = 0; k < DEX_NUM_COUNTERS; ++k) c[k] = __ldcg(ep.counters + k);
    const float fn = (float)n;
    ep.stats[DEX_STAT_SUCCESS_RATE] = (float)c[DEX_CNT_SUCCESS] / fn;
    ep.stats[DEX_STAT_FAILURE_RATE] = (float)c[DEX_CNT_FAILURE] / fn;
    ep.stats[DEX_STAT_TIMEOUT_RATE] = (float)c[DEX_CNT_TIMEOUT] / fn;
#pragma unroll
    for (int k = 0; k < DEX_NUM_CRITERIA; ++k) ep.stats[DEX_STAT_CRIT0 + k] = (float)c[DEX_CNT_CRIT0 + k] / fn;
    int cs = *ep.consecutive_successes;
    cs = (c[DEX_CNT_SUCCESS] > 0) ? cs + 1 : 0;
    *ep.consecutive_successes = min(cs, max_consecutive);
    if (ep.totals != nullptr) {
#pragma unroll
      for (int k = 0; k < DEX_NUM_COUNTERS; ++k) ep.totals[k] += c[k];
      ep.totals[DEX_NUM_COUNTERS] += 1;
    }
#pragma unroll
    for (int k = 0; k < DEX_NUM_COUNTERS; ++k) ep.counters[k] = 0;
    if (ep.rng_step != nullptr) *ep.rng_step += 1;
  }
}

// ------------------------------------------------------------------------------------------------
// the kernel
// ------------------------------------------------------------------------------------------------
template <bool BG, bool DICT>
__global__ void __launch_bounds__(kPostThreads) post_physics_kernel(const __grid_constant__ PostArgs A) {
  extern __shared__ __align__(16) float smem[];
  __shared__ int s_dof2fin[DEX_NUM_DOF];
  __shared__ int s_epi_warp[kPostWarps];
  __shared__ int s_epi_base;
  __shared__ int s_last;
  const DexParams& P = A.p;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int e0 = blockIdx.x * kTile;
  const int ne = min(kTile, A.n - e0);
  const Slab L = make_slab(P, DICT);
  const bool want_tgt = post_want_targets(P, DICT);

  if (tid < DEX_NUM_DOF) {  // DOF -> column inside active_finger_dof_* (observation_encoder.py:496-574), -1 = none
    int j = -1;
#pragma unroll
    for (int k = 0; k < DEX_NUM_ACTIVE_FINGER; ++k)
      if (P.active_finger_dof[k] == tid) j = k;
    s_dof2fin[tid] = j;
  }
  post_stage_in<BG>(A, L, smem, e0, ne, want_tgt);
  __syncthreads();
  post_items<BG, DICT>(A, L, smem, s_dof2fin, e0, ne, want_tgt);
  __syncthreads();
  if (warp == 0) post_scalar<BG, DICT>(A, L, smem, e0, ne);
  __syncthreads();
  {  // stage-out: obs rows, consecutive lanes = consecutive columns
    const int W = P.obs_dim;
    for (int e = warp; e < ne; e += kPostWarps) {
      const float* src = smem + e * L.stride + L.obs;
      float* dst = A.out.obs_buf + ((size_t)e0 + e) * W;
      for (int k = lane; k < W; k += 32) dst[k] = src[k];
    }
  }
  if (A.fused_epilogue) {
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      const unsigned t = atomicAdd(A.ep.ticket, 1u);
      s_last = (t == gridDim.x - 1) ? 1 : 0;
    }
    __syncthreads();
    if (s_last) {
      __threadfence();
      epilogue_body<kPostThreads>(A.ep, A.n, P.max_consecutive_successes, s_epi_warp, &s_epi_base);
      if (tid == 0) *A.ep.ticket = 0u;
    }
  }
}

__global__ void __launch_bounds__(1024) epilogue_kernel(const DexEpilogueBuffers ep, const int n, const int max_consecutive) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  epilogue_body<1024>(ep, n, max_consecutive, s_warp, &s_base);
}

template <bool BG, bool DICT>
int launch_post_t(const PostArgs& a, cudaStream_t s) {
  const size_t smem = post_smem_bytes(a.p, DICT);
  if (smem > 220 * 1024) return DEX_ERR_SIZE;
  static bool configured = false;  // per template instance
  if (!configured) {
    const cudaError_t e = cudaFuncSetAttribute(post_physics_kernel<BG, DICT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               220 * 1024);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  const int grid = (a.n + kTile - 1) / kTile;
  post_physics_kernel<BG, DICT><<<grid, kPostThreads, smem, s>>>(a);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

inline int launch_post(const PostArgs& a, cudaStream_t s) {
  const bool bg = a.p.task == DEX_TASK_BLIND_GRASPING, dict = a.out.obs_dict != nullptr;
  if (bg) return dict ? launch_post_t<true, true>(a, s) : launch_post_t<true, false>(a, s);
  return dict ? launch_post_t<false, true>(a, s) : launch_post_t<false, false>(a, s);
}

}  // namespace
