// dexstep.cu — hand-written sm_100a kernels + C ABI for the per-step tensor pipeline of dexhand_env.
//
// Compile: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -shared -Xcompiler -fPIC
//
// -fmad=false is deliberate: the reference computes every term with separately rounded fp32 torch ops,
// and thresholded masks sit downstream of several of them.  The path is HBM/latency bound (~0.3 flop/B),
// so un-fused multiplies cost nothing measurable.  No --use_fast_math: sqrtf/expf/division are the IEEE
// (<=1-2 ulp) versions.
//
// Layout / mapping:
//   * one thread per env for all arithmetic (envs are independent; no cross-env term on the path);
//   * path-internal state is SoA [K][N] -> thread-per-env loads/stores are naturally 128 B-coalesced;
//   * buffers whose layout is dictated from outside (Isaac Gym AoS sim tensors, actions, obs_buf,
//     full_dof_targets) are staged through shared memory per tile of envs: coalesced 128-bit / sector
//     gathers global->smem, conflict-free odd-stride rows for the thread-per-env side, coalesced
//     smem->global copy-out.
#include <cuda_runtime.h>
#include <stdint.h>

#include "dexstep.h"

namespace {

// ------------------------------------------------------------------------------------------------
// small math, in the operation order of the reference's torch ops (see oracle/dexstep_oracle.py)
// ------------------------------------------------------------------------------------------------
struct Quat {
  float x, y, z, w;
};
struct Vec3 {
  float x, y, z;
};

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

__device__ __forceinline__ float norm3(float x, float y, float z) { return sqrtf(x * x + y * y + z * z); }

// isaacgym.torch_utils.quat_mul (factored Hamilton product on x,y,z,w)
__device__ __forceinline__ Quat quat_mul(const Quat a, const Quat b) {
  const float ww = (a.z + a.x) * (b.x + b.y);
  const float yy = (a.w - a.y) * (b.w + b.z);
  const float zz = (a.w + a.y) * (b.w - b.z);
  const float xx = ww + yy + zz;
  const float qq = 0.5f * (xx + (a.z - a.x) * (b.x - b.y));
  Quat r;
  r.w = qq - ww + (a.z - a.y) * (b.y - b.z);
  r.x = qq - xx + (a.x + a.w) * (b.x + b.w);
  r.y = qq - yy + (a.w - a.x) * (b.y + b.z);
  r.z = qq - zz + (a.z + a.y) * (b.w - b.x);
  return r;
}

__device__ __forceinline__ Quat quat_conj(const Quat q) { return Quat{-q.x, -q.y, -q.z, q.w}; }

// isaacgym.torch_utils.quat_rotate_inverse: v(2w^2-1) - 2w(q x v) + 2 q (q.v)
__device__ __forceinline__ Vec3 quat_rotate_inverse(const Quat q, const Vec3 v) {
  const float s = 2.0f * (q.w * q.w) - 1.0f;
  const float cx = q.y * v.z - q.z * v.y;
  const float cy = q.z * v.x - q.x * v.z;
  const float cz = q.x * v.y - q.y * v.x;
  const float d = q.x * v.x + q.y * v.y + q.z * v.z;
  Vec3 r;
  r.x = v.x * s - cx * q.w * 2.0f + q.x * d * 2.0f;
  r.y = v.y * s - cy * q.w * 2.0f + q.y * d * 2.0f;
  r.z = v.z * s - cz * q.w * 2.0f + q.z * d * 2.0f;
  return r;
}

// Philox4x32-10 (counter-based; no RNG state in memory, no extra launches)
__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0;
  out[1] = c1;
  out[2] = c2;
  out[3] = c3;
}

// ------------------------------------------------------------------------------------------------
// reset of one env (ResetManager.reset_idx + BlindGraspingTask.reset_task_state +
// ActionProcessor.reset_targets + ObservationEncoder.reset_observer_state)
// ------------------------------------------------------------------------------------------------
struct ResetCtx {
  const float* uniforms;  // [N,29] or nullptr
  uint64_t seed, step;
};

template <bool BG>
__device__ void reset_env(const DexParams& P, const DexSimViews& sim, const DexStateBuffers& st, const ResetCtx& rc,
                          const int env, const size_t N) {
  // reset_manager.py:107-124
  st.episode_step_count[env] = 0;
  float pos[DEX_NUM_DOF];
#pragma unroll
  for (int d = 0; d < DEX_NUM_DOF; ++d) pos[d] = 0.0f;

  if (BG) {
    float u[32];
    if (rc.uniforms != nullptr) {
      const float* src = rc.uniforms + (size_t)env * DEX_NUM_UNIFORMS;
#pragma unroll
      for (int i = 0; i < DEX_NUM_UNIFORMS; ++i) u[i] = src[i];
    } else {
#pragma unroll
      for (int b = 0; b < 8; ++b) {
        uint32_t r[4];
        philox4x32_10((uint32_t)env, (uint32_t)b, (uint32_t)rc.step, (uint32_t)(rc.step >> 32), (uint32_t)rc.seed,
                      (uint32_t)(rc.seed >> 32), r);
#pragma unroll
        for (int k = 0; k < 4; ++k) u[4 * b + k] = (float)(r[k] >> 8) * (1.0f / 16777216.0f);
      }
    }
    // blind_grasping_task.py:449-490: box pose
    const float bx = (u[0] * 2.0f - 1.0f) * P.box_xy_range;
    const float by = (u[1] * 2.0f - 1.0f) * P.box_xy_range;
    const float ang = (u[2] * 2.0f - 1.0f) * P.pi;
    st.initial_box_positions[0 * N + env] = bx;
    st.initial_box_positions[1 * N + env] = by;
    st.initial_box_positions[2 * N + env] = P.box_z;
    float* root = sim.actor_root_state + ((size_t)env * P.num_actors + P.box_actor) * 13;
    root[0] = bx;
    root[1] = by;
    root[2] = P.box_z;
    root[3] = 0.0f;
    root[4] = 0.0f;
    root[5] = sinf(ang / 2.0f);
    root[6] = cosf(ang / 2.0f);
#pragma unroll
    for (int i = 7; i < 13; ++i) root[i] = 0.0f;
    // :492-509 task states
    st.success_duration_steps[env] = 0;
    st.success_conditions_met[env] = 0;
    st.current_stage[env] = 1;
    st.time_in_stage[env] = 0.0f;
    st.stage_contact_duration[env] = 0.0f;
    st.just_transitioned_to_stage2[env] = 0;
    st.just_transitioned_to_stage3[env] = 0;
    // :511-547 hand DOF randomisation (all 20 finger DOFs, including r_f_joint3_1)
#pragma unroll
    for (int i = 0; i < 3; ++i) pos[i] = (u[3 + i] * 2.0f - 1.0f) * P.hand_translation_range;
#pragma unroll
    for (int i = 0; i < 3; ++i) pos[3 + i] = (u[6 + i] * 2.0f - 1.0f) * P.hand_rotation_range;
    pos[6] = u[9] * P.thumb_rotation_range;
#pragma unroll
    for (int i = 1; i < DEX_NUM_FINGER_DOF; ++i) pos[6 + i] = u[9 + i] * P.other_finger_range;
  }

  float* dof = sim.dof_state + (size_t)env * (DEX_NUM_DOF * 2);
  float* full = st.full_dof_targets + (size_t)env * DEX_NUM_DOF;
#pragma unroll
  for (int d = 0; d < DEX_NUM_DOF; ++d) {
    dof[2 * d] = pos[d];
    dof[2 * d + 1] = 0.0f;
    full[d] = pos[d];  // action_processor.py:553 (verbatim copy, DOF 14 included)
    st.prev_dof_pos[d * N + env] = 0.0f;  // observation_encoder.py:375-376 (zero, not the new pose)
  }
  // action_processor.py:616-666 inverse coupling: the last DOF of a coupling group wins
#pragma unroll
  for (int a = 0; a < DEX_NUM_ACTIVE; ++a) {
    float v = 0.0f;
#pragma unroll
    for (int d = 0; d < DEX_NUM_DOF; ++d)
      if (P.inverse_src_dof[a] == d) v = pos[d];
    st.active_prev_targets[a * N + env] = (a < DEX_NUM_BASE) ? v : v / P.inverse_scale[a];
  }
  // observation_encoder.py:363-380
#pragma unroll
  for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
    st.contact_duration_steps[f * N + env] = 0.0f;
    st.prev_contact_binary[f * N + env] = 0.0f;
  }
  for (int k = 0; k < P.num_actions; ++k) st.prev_actions[k * N + env] = 0.0f;
}

// ------------------------------------------------------------------------------------------------
// post-physics kernel
// ------------------------------------------------------------------------------------------------
struct PostArgs {
  DexParams p;
  DexSimViews sim;
  DexStateBuffers st;
  DexOutBuffers out;
  int32_t n;
  int32_t mode;
  uint32_t flags;
  const float* uniforms;
  uint64_t seed;
  uint64_t step;
};

constexpr int kDofW = DEX_NUM_DOF * 2;  // 52 floats per env, 13 float4
constexpr int kRbW = 13 + 10 * 7;       // hand row (13) + 5 tips + 5 pads (7 each) = 83 (odd stride)
constexpr int kCfW = 19;                // 5*3 monitored + 3 box (+1 pad -> odd stride)
constexpr int kBoxW = 7;                // pos3 + linvel3 (+1 pad)
constexpr int kVelW = 27;               // 26 finite-difference velocities (+1 pad)
constexpr int kTgtW = 27;               // 26 DOF targets (+1 pad)

__host__ __device__ inline int post_smem_floats(int tile, int obs_dim) {
  return tile * (kDofW + kRbW + kCfW + kBoxW + kVelW + kTgtW + (obs_dim | 1)) + 128;
}

template <bool BG, int TILE>
__global__ void __launch_bounds__(TILE) post_physics_kernel(const __grid_constant__ PostArgs A) {
  extern __shared__ __align__(16) float smem[];
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const int tid = threadIdx.x;
  const int e0 = blockIdx.x * TILE;
  const int ne = min(TILE, A.n - e0);
  const size_t N = (size_t)A.n;
  const int OS = P.obs_dim | 1;

  float* s_dof = smem;
  float* s_rb = s_dof + TILE * kDofW;
  float* s_cf = s_rb + TILE * kRbW;
  float* s_box = s_cf + TILE * kCfW;
  float* s_vel = s_box + TILE * kBoxW;
  float* s_tgt = s_vel + TILE * kVelW;
  float* s_obs = s_tgt + TILE * kTgtW;
  int* s_lut = reinterpret_cast<int*>(s_obs + TILE * OS);  // [83] rigid slot -> offset inside the env slab

  // ---- stage in -----------------------------------------------------------------------------
  for (int s = tid; s < kRbW; s += TILE) {
    int body, comp;
    if (s < 13) {
      body = P.hand_body;
      comp = s;
    } else {
      const int t = s - 13, f = t / 7;
      comp = t - f * 7;
      body = 0;
#pragma unroll
      for (int k = 0; k < DEX_NUM_FINGERS; ++k) {
        if (f == k) body = P.tip_body[k];
        if (f == k + DEX_NUM_FINGERS) body = P.pad_body[k];
      }
    }
    s_lut[s] = body * 13 + comp;
  }
  {  // dof_state tile: contiguous, 16 B aligned (52 floats = 13 float4 per env)
    const float4* src = reinterpret_cast<const float4*>(A.sim.dof_state + (size_t)e0 * kDofW);
    float4* dst = reinterpret_cast<float4*>(s_dof);
    for (int i = tid; i < ne * (kDofW / 4); i += TILE) dst[i] = __ldg(src + i);
  }
  const bool want_targets = A.out.obs_dict != nullptr || P.obs_offset[DEX_K_BASE_DOF_TARGET] >= 0 ||
                            P.obs_offset[DEX_K_ACTIVE_FINGER_DOF_TARGET] >= 0 ||
                            P.obs_offset[DEX_K_ALL_FINGER_DOF_TARGET] >= 0;
  if (want_targets) {  // full_dof_targets tile [ne,26]: contiguous, coalesced
    const float* src = st.full_dof_targets + (size_t)e0 * DEX_NUM_DOF;
    for (int i = tid; i < ne * DEX_NUM_DOF; i += TILE) {
      const int e = i / DEX_NUM_DOF, k = i - e * DEX_NUM_DOF;
      s_tgt[e * kTgtW + k] = src[i];
    }
  }
  __syncthreads();
  {  // rigid-body rows: sector gathers, consecutive lanes read consecutive floats of a row
    const float* rb = A.sim.rigid_body_state + (size_t)e0 * P.num_bodies * 13;
    const int slab = P.num_bodies * 13;
    const int nrb = (A.mode == DEX_MODE_FULL && P.need_common_rewards) ? 13 : 7;  // hand velocities only for R1
    for (int i = tid; i < ne * kRbW; i += TILE) {
      const int e = i / kRbW, s = i - e * kRbW;
      if (s < 13 && s >= nrb) continue;
      s_rb[i] = __ldg(rb + (size_t)e * slab + s_lut[s]);
    }
    if (A.flags & DEX_FLAG_STALE_CONTACTS) {
      for (int i = tid; i < ne * 15; i += TILE) {
        const int e = i / 15, s = i - e * 15;
        s_cf[e * kCfW + s] = st.contact_forces[(size_t)s * N + e0 + e];
      }
    } else {
      const float* cf = A.sim.net_contact_force + (size_t)e0 * P.num_bodies * 3;
      for (int i = tid; i < ne * 15; i += TILE) {
        const int e = i / 15, s = i - e * 15, f = s / 3, c = s - f * 3;
        int body = 0;
#pragma unroll
        for (int k = 0; k < DEX_NUM_FINGERS; ++k)
          if (f == k) body = P.contact_body[k];
        s_cf[e * kCfW + s] = __ldg(cf + ((size_t)e * P.num_bodies + body) * 3 + c);
      }
    }
    if (BG) {
      const float* cf = A.sim.net_contact_force + (size_t)e0 * P.num_bodies * 3;
      for (int i = tid; i < ne * 3; i += TILE) {
        const int e = i / 3, c = i - e * 3;
        s_cf[e * kCfW + 15 + c] = __ldg(cf + ((size_t)e * P.num_bodies + P.box_body) * 3 + c);
      }
      const float* root = A.sim.actor_root_state + (size_t)e0 * P.num_actors * 13;
      for (int i = tid; i < ne * 6; i += TILE) {
        const int e = i / 6, c = i - e * 6;
        s_box[e * kBoxW + c] = root[((size_t)e * P.num_actors + P.box_actor) * 13 + (c < 3 ? c : c + 4)];
      }
    }
  }
  __syncthreads();

  // ---- per-env compute ----------------------------------------------------------------------
  const bool active = tid < ne;
  const int env = e0 + tid;
  const bool full = (A.mode == DEX_MODE_FULL);
  bool should_reset = false, t_success = false, t_failure = false, t_timeout = false;
  int ep_len_done = 0;
  bool crit[DEX_NUM_CRITERIA];
#pragma unroll
  for (int c = 0; c < DEX_NUM_CRITERIA; ++c) crit[c] = false;

  if (active) {
    float* obs = s_obs + tid * OS;
    float* dict = A.out.obs_dict ? A.out.obs_dict + (size_t)env * P.dict_dim : nullptr;
    auto emit = [&](const int key, const int j, const float v) {
      const int o = P.obs_offset[key];
      if (o >= 0) obs[o + j] = v;
      if (dict) dict[P.dict_offset[key] + j] = v;
    };
    auto want = [&](const int key) { return P.obs_offset[key] >= 0 || dict != nullptr; };

    const float* dof = s_dof + tid * kDofW;
    const float* rb = s_rb + tid * kRbW;
    const float* cf = s_cf + tid * kCfW;
    float* vel = s_vel + tid * kVelW;
    const float dt = P.control_dt;

    // O2 finite-difference velocity (observation_encoder.py:298-321), prev_dof_pos <- pos
#pragma unroll
    for (int d = 0; d < DEX_NUM_DOF; ++d) {
      const float pos = dof[2 * d];
      const float prev = st.prev_dof_pos[d * N + env];
      vel[d] = (pos - prev) / dt;
      st.prev_dof_pos[d * N + env] = pos;
    }
    // O3 slices / gathers (observation_encoder.py:612-630)
#pragma unroll
    for (int d = 0; d < DEX_NUM_BASE; ++d) {
      emit(DEX_K_BASE_DOF_POS, d, dof[2 * d]);
      emit(DEX_K_BASE_DOF_VEL, d, vel[d]);
    }
#pragma unroll
    for (int j = 0; j < DEX_NUM_ACTIVE_FINGER; ++j) {
      const int d = P.active_finger_dof[j];
      emit(DEX_K_ACTIVE_FINGER_DOF_POS, j, dof[2 * d]);
      emit(DEX_K_ACTIVE_FINGER_DOF_VEL, j, vel[d]);
    }
    if (want(DEX_K_ALL_FINGER_DOF_POS) || want(DEX_K_ALL_FINGER_DOF_VEL)) {
#pragma unroll
      for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) {
        emit(DEX_K_ALL_FINGER_DOF_POS, j, dof[2 * (DEX_NUM_BASE + j)]);
        emit(DEX_K_ALL_FINGER_DOF_VEL, j, vel[DEX_NUM_BASE + j]);
      }
    }
    // hand pose (:632-656) and ARR-aligned pose (:1483-1521)
    const Vec3 hp{rb[0], rb[1], rb[2]};
    const Quat hq{rb[3], rb[4], rb[5], rb[6]};
#pragma unroll
    for (int c = 0; c < 7; ++c) emit(DEX_K_HAND_POSE, c, rb[c]);
    if (want(DEX_K_HAND_POSE_ARR_ALIGNED)) {
      const float sh = 0.7071067811865476f;
      const Quat inv{-0.0f, -sh, -0.0f, sh};
      const Quat qa = quat_mul(hq, inv);
      emit(DEX_K_HAND_POSE_ARR_ALIGNED, 0, hp.x);
      emit(DEX_K_HAND_POSE_ARR_ALIGNED, 1, hp.y);
      emit(DEX_K_HAND_POSE_ARR_ALIGNED, 2, hp.z);
      emit(DEX_K_HAND_POSE_ARR_ALIGNED, 3, qa.x);
      emit(DEX_K_HAND_POSE_ARR_ALIGNED, 4, qa.y);
      emit(DEX_K_HAND_POSE_ARR_ALIGNED, 5, qa.z);
      emit(DEX_K_HAND_POSE_ARR_ALIGNED, 6, qa.w);
    }
    // targets / previous actions (:664-701)
    if (want(DEX_K_PREV_ACTIONS))
      for (int k = 0; k < P.num_actions; ++k) emit(DEX_K_PREV_ACTIONS, k, st.prev_actions[k * N + env]);
    {
      const bool w_apt = want(DEX_K_ACTIVE_PREV_TARGETS), w_art = want(DEX_K_ACTIVE_RULE_TARGETS);
#pragma unroll
      for (int a = 0; a < DEX_NUM_ACTIVE; ++a) {
        const float t = st.active_prev_targets[a * N + env];
        // O8 identity pre-action rule (rules.py:78-95): rule targets for the next process_actions
        st.active_rule_targets[a * N + env] = t;
        if (w_apt) emit(DEX_K_ACTIVE_PREV_TARGETS, a, t);
        if (w_art) emit(DEX_K_ACTIVE_RULE_TARGETS, a, t);
      }
    }
    if (want_targets) {
      const float* fdt = s_tgt + tid * kTgtW;
#pragma unroll
      for (int d = 0; d < DEX_NUM_BASE; ++d) emit(DEX_K_BASE_DOF_TARGET, d, fdt[d]);
#pragma unroll
      for (int j = 0; j < DEX_NUM_ACTIVE_FINGER; ++j) emit(DEX_K_ACTIVE_FINGER_DOF_TARGET, j, fdt[P.active_finger_dof[j]]);
      if (want(DEX_K_ALL_FINGER_DOF_TARGET))
#pragma unroll
        for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) emit(DEX_K_ALL_FINGER_DOF_TARGET, j, fdt[DEX_NUM_BASE + j]);
    }
    // contacts (:658-735) + O4 duration tracking (:323-361)
    float cmag[DEX_NUM_FINGERS];
    bool cbin[DEX_NUM_FINGERS];
#pragma unroll
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
      const float fx = cf[3 * f], fy = cf[3 * f + 1], fz = cf[3 * f + 2];
      emit(DEX_K_CONTACT_FORCES, 3 * f, fx);
      emit(DEX_K_CONTACT_FORCES, 3 * f + 1, fy);
      emit(DEX_K_CONTACT_FORCES, 3 * f + 2, fz);
      if (st.contact_forces != nullptr && !(A.flags & DEX_FLAG_STALE_CONTACTS)) {
        st.contact_forces[(size_t)(3 * f) * N + env] = fx;
        st.contact_forces[(size_t)(3 * f + 1) * N + env] = fy;
        st.contact_forces[(size_t)(3 * f + 2) * N + env] = fz;
      }
      cmag[f] = norm3(fx, fy, fz);
      cbin[f] = cmag[f] > P.contact_binary_threshold;
      emit(DEX_K_CONTACT_FORCE_MAGNITUDE, f, cmag[f]);
      emit(DEX_K_CONTACT_BINARY, f, cbin[f] ? 1.0f : 0.0f);
      const float prevb = st.prev_contact_binary[f * N + env];
      const float dur = st.contact_duration_steps[f * N + env];
      const float ndur = (cbin[f] && prevb == 0.0f) ? 1.0f : (cbin[f] ? dur + 1.0f : 0.0f);
      st.contact_duration_steps[f * N + env] = ndur;
      st.prev_contact_binary[f * N + env] = cbin[f] ? 1.0f : 0.0f;
      emit(DEX_K_CONTACT_DURATION, f, ndur * dt);
    }
    // tip / pad poses, world and hand frame (:737-751, :915-986)
    const bool w_tip_h = want(DEX_K_FINGERTIP_POSES_HAND), w_pad_h = want(DEX_K_FINGERPAD_POSES_HAND);
    const Quat hqc = quat_conj(hq);
    float min_tip_z = 0.0f;
#pragma unroll
    for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
      const float* tp = rb + 13 + 7 * f;
      const float* pp = rb + 13 + 35 + 7 * f;
#pragma unroll
      for (int c = 0; c < 7; ++c) {
        emit(DEX_K_FINGERTIP_POSES_WORLD, 7 * f + c, tp[c]);
        emit(DEX_K_FINGERPAD_POSES_WORLD, 7 * f + c, pp[c]);
      }
      min_tip_z = (f == 0) ? tp[2] : fminf(min_tip_z, tp[2]);
      if (w_tip_h) {
        const Vec3 ph = quat_rotate_inverse(hq, Vec3{tp[0] - hp.x, tp[1] - hp.y, tp[2] - hp.z});
        const Quat qh = quat_mul(hqc, Quat{tp[3], tp[4], tp[5], tp[6]});
        emit(DEX_K_FINGERTIP_POSES_HAND, 7 * f + 0, ph.x);
        emit(DEX_K_FINGERTIP_POSES_HAND, 7 * f + 1, ph.y);
        emit(DEX_K_FINGERTIP_POSES_HAND, 7 * f + 2, ph.z);
        emit(DEX_K_FINGERTIP_POSES_HAND, 7 * f + 3, qh.x);
        emit(DEX_K_FINGERTIP_POSES_HAND, 7 * f + 4, qh.y);
        emit(DEX_K_FINGERTIP_POSES_HAND, 7 * f + 5, qh.z);
        emit(DEX_K_FINGERTIP_POSES_HAND, 7 * f + 6, qh.w);
      }
      if (w_pad_h) {
        const Vec3 ph = quat_rotate_inverse(hq, Vec3{pp[0] - hp.x, pp[1] - hp.y, pp[2] - hp.z});
        const Quat qh = quat_mul(hqc, Quat{pp[3], pp[4], pp[5], pp[6]});
        emit(DEX_K_FINGERPAD_POSES_HAND, 7 * f + 0, ph.x);
        emit(DEX_K_FINGERPAD_POSES_HAND, 7 * f + 1, ph.y);
        emit(DEX_K_FINGERPAD_POSES_HAND, 7 * f + 2, ph.z);
        emit(DEX_K_FINGERPAD_POSES_HAND, 7 * f + 3, qh.x);
        emit(DEX_K_FINGERPAD_POSES_HAND, 7 * f + 4, qh.y);
        emit(DEX_K_FINGERPAD_POSES_HAND, 7 * f + 5, qh.z);
        emit(DEX_K_FINGERPAD_POSES_HAND, 7 * f + 6, qh.w);
      }
    }
    // episode_time from the PRE-increment step count (dexhand_base.py:703-712)
    const int64_t step_before = st.episode_step_count[env];
    emit(DEX_K_EPISODE_TIME, 0, (float)step_before * dt);

    // ---------------- BlindGrasping task observations + FSM (blind_grasping_task.py:549-978) ----
    float avg_hdiff = 0.0f, hand_box_dist = 0.0f, grasp_duration = 0.0f, min_tip_box = 0.0f;
    float cen[3] = {0.0f, 0.0f, 0.0f};
    float box[3] = {0.0f, 0.0f, 0.0f}, bvel[3] = {0.0f, 0.0f, 0.0f};
    bool thumb_c = false, others_c = false, grasp = false, j2 = false, j3 = false;
    int64_t stage = 0, sds = 0;
    if (BG) {
      const float* bx = s_box + tid * kBoxW;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        box[c] = bx[c];
        bvel[c] = bx[3 + c];
        emit(DEX_K_OBJECT_POS, c, box[c]);
        emit(DEX_K_OBJECT_VEL, c, bvel[c]);
      }
      float pad[DEX_NUM_FINGERS][3];
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f)
#pragma unroll
        for (int c = 0; c < 3; ++c) pad[f][c] = rb[13 + 35 + 7 * f + c];
      // spatial relationships :684-724
      float dsum = 0.0f, hsum = 0.0f, pdist[DEX_NUM_FINGERS];
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        pdist[f] = norm3(pad[f][0] - box[0], pad[f][1] - box[1], pad[f][2] - box[2]);
        const float hd = fabsf(pad[f][2] - box[2]);
        dsum += pdist[f];
        hsum += hd;
        emit(DEX_K_FINGER_TO_OBJECT_DISTANCES, f, pdist[f]);
        emit(DEX_K_FINGER_TO_OBJECT_HEIGHT_DIFF, f, hd);
      }
      avg_hdiff = hsum / 5.0f;
      emit(DEX_K_AVG_FINGER_TO_OBJECT_DISTANCE, 0, dsum / 5.0f);
      emit(DEX_K_AVG_FINGER_TO_OBJECT_HEIGHT_DIFF, 0, avg_hdiff);
      hand_box_dist = norm3(hp.x - box[0], hp.y - box[1], hp.z - box[2]);
      emit(DEX_K_HAND_TO_OBJECT_DISTANCE, 0, hand_box_dist);
      // geometry :726-771 (triu order (0,1),(0,2),(0,3),(0,4),(1,2),...)
      {
        int k = 0;
#pragma unroll
        for (int a = 0; a < DEX_NUM_FINGERS; ++a)
#pragma unroll
          for (int b = a + 1; b < DEX_NUM_FINGERS; ++b) {
            emit(DEX_K_FINGERPAD_DISTANCES, k,
                 norm3(pad[a][0] - pad[b][0], pad[a][1] - pad[b][1], pad[a][2] - pad[b][2]));
            ++k;
          }
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        cen[c] = (pad[0][c] + pad[1][c] + pad[2][c]) / 3.0f;
        emit(DEX_K_FIRST_THREE_FINGERPAD_CENTROID, c, cen[c]);
      }
      // finger-box contacts :773-833
      const float box_mag = norm3(cf[15], cf[16], cf[17]);
      const bool box_has = box_mag > P.contact_binary_threshold;
      bool fbc[DEX_NUM_FINGERS];
      int n_on_box = 0;
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        fbc[f] = cbin[f] && box_has && (pdist[f] < P.proximity_threshold);
        n_on_box += fbc[f] ? 1 : 0;
      }
      thumb_c = fbc[0];
      others_c = fbc[1] || fbc[2] || fbc[3] || fbc[4];
      grasp = thumb_c && others_c;
      emit(DEX_K_THUMB_CONTACT, 0, thumb_c ? 1.0f : 0.0f);
      emit(DEX_K_OTHER_FINGERS_CONTACT, 0, others_c ? 1.0f : 0.0f);
      emit(DEX_K_GRASP_STATE, 0, grasp ? 1.0f : 0.0f);
      // success duration :626-648
      const bool cond = (box[2] > P.success_height_threshold) && (n_on_box >= P.min_fingers_for_grasp);
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
      // closest fingertip to the box centre :1178-1192
#pragma unroll
      for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
        const float* tp = rb + 13 + 7 * f;
        const float d = norm3(tp[0] - box[0], tp[1] - box[1], tp[2] - box[2]);
        min_tip_box = (f == 0) ? d : fminf(min_tip_box, d);
      }
    }

    if (full) {
      // step_processor.py:80
      const int64_t step_now = step_before + 1;
      st.episode_step_count[env] = step_now;
      ep_len_done = (int)step_now;

      // ---------------- termination (step_processor.py:133-181, termination_manager.py:120-278) ----
      if (P.has_height_safety)
        crit[DEX_C_HITTING_GROUND] = (hp.z < P.handbase_threshold) || (min_tip_z < P.fingertip_threshold);
      bool pregrasp_failed = false, contact_failed = false, grasp_lost = false;
      if (BG) {
        crit[DEX_C_GRASP_LIFT_SUCCESS] = sds >= (int64_t)P.contact_duration_threshold_steps;
        crit[DEX_C_BOX_TOO_FAR] = hand_box_dist > P.max_box_distance;
        if (j2) {  // _evaluate_pregrasp_quality :1334-1364
          bool h_ok = true;
#pragma unroll
          for (int f = 0; f < 3; ++f) h_ok = h_ok && (fabsf(rb[13 + 35 + 7 * f + 2] - box[2]) <= P.height_tolerance);
          const bool c_ok = norm3(cen[0] - box[0], cen[1] - box[1], cen[2] - box[2]) <= P.centroid_tolerance;
          const float drift = norm3(box[0] - st.initial_box_positions[0 * N + env], box[1] - st.initial_box_positions[1 * N + env],
                                    box[2] - st.initial_box_positions[2 * N + env]);
          const float vmag = norm3(bvel[0], bvel[1], bvel[2]);
          const bool s_ok = (drift <= P.position_drift_tolerance) && (vmag <= P.velocity_tolerance);
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
        const bool first = (A.flags & DEX_FLAG_FIRST_REWARD_CALL) != 0;
        r[DEX_R_ALIVE] = 1.0f;
        r[DEX_R_HEIGHT_SAFETY] = clampf(1.0f - expf(-(min_tip_z - 0.02f) * 20.0f), 0.0f, 1.0f);
        float fv2 = 0.0f, fa2 = 0.0f, jl = 0.0f;
#pragma unroll
        for (int j = 0; j < DEX_NUM_FINGER_DOF; ++j) {
          const int d = DEX_NUM_BASE + j;
          const float v = dof[2 * d + 1];  // Isaac dof_vel, not the finite difference (reward_calculator.py:94-98)
          const float pv = first ? v : st.prev_finger_dof_vel[j * N + env];
          st.prev_finger_dof_vel[j * N + env] = v;
          fv2 += v * v;
          const float a = v - pv;
          fa2 += a * a;
          const float nj = 2.0f * (dof[2 * d] - P.dof_lower[d]) / (P.dof_upper[d] - P.dof_lower[d]) - 1.0f;
          jl += clampf(fabsf(nj) - 0.8f, 0.0f, 1.0f);
        }
        r[DEX_R_FINGER_VELOCITY] = expf(-0.1f * sqrtf(fv2));
        r[DEX_R_JOINT_LIMIT] = 1.0f - jl / 20.0f;
        r[DEX_R_FINGER_ACCELERATION] = expf(-2.0f * sqrtf(fa2));
        float hv[3], ha[3], phv[3], pha[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          hv[c] = rb[7 + c];
          ha[c] = rb[10 + c];
          phv[c] = first ? hv[c] : st.prev_hand_vel[c * N + env];
          pha[c] = first ? ha[c] : st.prev_hand_ang_vel[c * N + env];
          st.prev_hand_vel[c * N + env] = hv[c];
          st.prev_hand_ang_vel[c * N + env] = ha[c];
        }
        r[DEX_R_HAND_VELOCITY] = expf(-0.2f * norm3(hv[0], hv[1], hv[2]));
        r[DEX_R_HAND_ANGULAR_VELOCITY] = expf(-0.2f * norm3(ha[0], ha[1], ha[2]));
        r[DEX_R_HAND_ACCELERATION] = expf(-0.5f * norm3(hv[0] - phv[0], hv[1] - phv[1], hv[2] - phv[2]));
        r[DEX_R_HAND_ANGULAR_ACCELERATION] = expf(-0.5f * norm3(ha[0] - pha[0], ha[1] - pha[1], ha[2] - pha[2]));
        float changes = 0.0f;
#pragma unroll
        for (int f = 0; f < DEX_NUM_FINGERS; ++f) {
          const bool c_now = cmag[f] > 0.1f;  // hard-coded 0.1 N (reward_calculator.py:186)
          const bool c_prev = first ? c_now : (st.prev_contacts[f * N + env] != 0);
          st.prev_contacts[f * N + env] = c_now ? 1 : 0;
          changes += (c_now != c_prev) ? 1.0f : 0.0f;
        }
        r[DEX_R_CONTACT_STABILITY] = expf(-changes);
      }
      if (BG) {
        const float m1 = (stage == 1) ? 1.0f : 0.0f, m2 = (stage == 2) ? 1.0f : 0.0f, m3 = (stage == 3) ? 1.0f : 0.0f;
        // stage 1 :1055-1112
        r[DEX_R_S1_HEIGHT_ALIGNMENT] = expf(-P.height_alignment_decay * avg_hdiff) * m1;
        const float cdist = norm3(cen[0] - box[0], cen[1] - box[1], cen[2] - box[2]);
        r[DEX_R_S1_CENTROID_POSITIONING] = expf(-P.centroid_positioning_decay * cdist) * m1;
        const float drift = norm3(box[0] - st.initial_box_positions[0 * N + env], box[1] - st.initial_box_positions[1 * N + env],
                                  box[2] - st.initial_box_positions[2 * N + env]);
        const float vmag = norm3(bvel[0], bvel[1], bvel[2]);
        r[DEX_R_S1_OBJECT_STABILITY] = expf(-P.object_stability_decay * (drift + vmag)) * m1;
        {  // torch.var(dim=1), unbiased, over the z of the first three pads
          const float z0 = rb[13 + 35 + 2], z1 = rb[13 + 35 + 7 + 2], z2 = rb[13 + 35 + 14 + 2];
          const float mean = (z0 + z1 + z2) / 3.0f;
          const float d0 = z0 - mean, d1 = z1 - mean, d2 = z2 - mean;
          const float var = (d0 * d0 + d1 * d1 + d2 * d2) / 2.0f;
          r[DEX_R_S1_FINGER_HEIGHT_CONSISTENCY] = expf(-P.height_consistency_decay * var) * m1;
        }
        r[DEX_R_S1_THUMB_ROTATION] = expf(-5.0f * fabsf(dof[2 * DEX_NUM_BASE] - P.half_pi)) * m1;
        // stage 2 :1114-1150
        r[DEX_R_S2_THUMB_CONTACT] = (thumb_c ? 1.0f : 0.0f) * m2;
        r[DEX_R_S2_OTHER_FINGERS_CONTACT] = (others_c ? 1.0f : 0.0f) * m2;
        r[DEX_R_S2_GRASP_ACHIEVEMENT] = (grasp ? 1.0f : 0.0f) * m2;
        r[DEX_R_S2_FINGERPAD_PROXIMITY] = expf(-P.fingerpad_proximity_decay * fmaxf(min_tip_box, P.min_reward_distance)) * m2;
        {
          float bv2 = 0.0f;
#pragma unroll
          for (int d = 0; d < DEX_NUM_BASE; ++d) bv2 += vel[d] * vel[d];  // finite-difference base velocity
          r[DEX_R_S2_BASE_STABILITY] = expf(-P.base_stability_decay * sqrtf(bv2)) * m2;
        }
        // stage 3 :1152-1176
        r[DEX_R_S3_OBJECT_HEIGHT] = clampf((box[2] - P.box_z) / P.lift_height_range, 0.0f, 1.0f) * m3;
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
        reset_env<BG>(P, A.sim, st, rc, env, N);
      }
    }
  }

  if (full) {
    // reset ballot word + cross-env counters (termination_manager.py:164-185,259-266)
    const unsigned fullmask = 0xffffffffu;
    const unsigned rbits = __ballot_sync(fullmask, should_reset);
    const int lane = tid & 31;
    if (lane == 0) A.out.reset_bits[(e0 + tid) >> 5] = rbits;
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

  // ---- stage out: obs tile, coalesced -------------------------------------------------------
  __syncthreads();
  {
    const int W = P.obs_dim;
    float* dst = A.out.obs_buf + (size_t)e0 * W;
    for (int i = tid; i < ne * W; i += TILE) {
      const int e = i / W, k = i - e * W;
      dst[i] = s_obs[e * OS + k];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// stand-alone masked reset
// ------------------------------------------------------------------------------------------------
struct ResetArgs {
  DexParams p;
  DexSimViews sim;
  DexStateBuffers st;
  const uint8_t* mask;
  int32_t n;
  const float* uniforms;
  uint64_t seed;
  uint64_t step;
};

template <bool BG>
__global__ void __launch_bounds__(128) reset_kernel(const __grid_constant__ ResetArgs A) {
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= A.n) return;
  if (A.mask != nullptr && A.mask[env] == 0) return;
  const ResetCtx rc{A.uniforms, A.seed, A.st.rng_step ? *A.st.rng_step : A.step};
  reset_env<BG>(A.p, A.sim, A.st, rc, env, (size_t)A.n);
}

// ------------------------------------------------------------------------------------------------
// pre-physics kernel: actions -> 18 active targets -> 26 DOF targets
// ------------------------------------------------------------------------------------------------
struct PreArgs {
  DexParams p;
  DexStateBuffers st;
  const float* actions;
  float* raw_out;
  int32_t n;
};

constexpr int kPreTile = 128;
constexpr int kActS = 19;   // odd smem row stride for <=18 actions
constexpr int kFullS = 27;  // odd smem row stride for 26 DOF targets

__global__ void __launch_bounds__(kPreTile) pre_physics_kernel(const __grid_constant__ PreArgs A) {
  __shared__ float s_act[kPreTile * kActS];
  __shared__ float s_next[kPreTile * kActS];
  __shared__ float s_full[kPreTile * kFullS];
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const int tid = threadIdx.x;
  const int e0 = blockIdx.x * kPreTile;
  const int ne = min(kPreTile, A.n - e0);
  const size_t N = (size_t)A.n;
  const int NA = P.num_actions;
  {  // stage the action tile (row-major [N,NACT]) with coalesced loads
    const float* src = A.actions + (size_t)e0 * NA;
    for (int i = tid; i < ne * NA; i += kPreTile) {
      const int e = i / NA, k = i - e * NA;
      s_act[e * kActS + k] = __ldg(src + i);
    }
  }
  __syncthreads();
  if (tid < ne) {
    const int env = e0 + tid;
    const float* act = s_act + tid * kActS;
    float* nxt = s_next + tid * kActS;
    const int fs = P.policy_controls_base ? DEX_NUM_BASE : 0;  // default_rules.py:53,97
#pragma unroll
    for (int a = 0; a < DEX_NUM_ACTIVE; ++a) {
      const float prev = st.active_prev_targets[a * N + env];
      const float rule = st.active_rule_targets[a * N + env];
      const float lo = P.active_lower[a], hi = P.active_upper[a], md = P.max_deltas[a];
      const bool controlled = (a < DEX_NUM_BASE) ? (P.policy_controls_base != 0) : (P.policy_controls_fingers != 0);
      float t = rule;
      if (controlled) {
        const float x = act[(a < DEX_NUM_BASE) ? a : fs + (a - DEX_NUM_BASE)];
        if (P.control_mode == DEX_CONTROL_POSITION)
          t = (x + 1.0f) * 0.5f * (hi - lo) + lo;  // scaling.py:28-45
        else
          t = prev + x * md;  // scaling.py:48-63
      }
      if (P.control_mode == DEX_CONTROL_POSITION_DELTA) t = clampf(t, lo, hi);  // default_rules.py:107-110
      if (A.raw_out != nullptr) A.raw_out[a * N + env] = t;
      if (P.filter_velocity_clamp) t = prev + clampf(t - prev, -md, md);  // rules.py:141-166
      if (P.filter_position_clamp) t = clampf(t, lo, hi);                  // rules.py:168-190
      nxt[a] = t;
      st.active_prev_targets[a * N + env] = t;  // action_processor.py:342
    }
    for (int k = 0; k < NA; ++k) st.prev_actions[k * N + env] = act[k];  // observation_encoder.py:287-296
    // coupling 18 -> 26 (action_processor.py:571-614)
    float* fullrow = s_full + tid * kFullS;
#pragma unroll
    for (int d = 0; d < DEX_NUM_DOF; ++d) {
      const int a = P.dof_to_active[d];
      float v = 0.0f;
      if (a >= 0) v = (d < DEX_NUM_BASE) ? nxt[a] : nxt[a] * P.dof_coupling_scale[d];
      fullrow[d] = v;
    }
  }
  __syncthreads();
  {
    float* dst = st.full_dof_targets + (size_t)e0 * DEX_NUM_DOF;
    for (int i = tid; i < ne * DEX_NUM_DOF; i += kPreTile) {
      const int e = i / DEX_NUM_DOF, k = i - e * DEX_NUM_DOF;
      dst[i] = s_full[e * kFullS + k];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// step epilogue: ordered compaction of reset ids + rates + consecutive successes (single CTA)
// ------------------------------------------------------------------------------------------------
struct EpiArgs {
  DexEpilogueBuffers ep;
  int32_t n;
  int32_t max_consecutive;
};

__global__ void __launch_bounds__(1024) epilogue_kernel(const __grid_constant__ EpiArgs A) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int words = (A.n + 31) >> 5;
  if (tid == 0) s_base = 0;
  __syncthreads();
  for (int w0 = 0; w0 < words; w0 += 1024) {
    const int w = w0 + tid;
    unsigned bits = (w < words) ? A.ep.reset_bits[w] : 0u;
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
      int v = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int u = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += u;
      }
      s_warp[lane] = v;
    }
    __syncthreads();
    int pos = s_base + (warp > 0 ? s_warp[warp - 1] : 0) + incl - cnt;
    while (bits) {
      const int b = __ffs(bits) - 1;
      A.ep.env_ids[pos++] = w * 32 + b;
      bits &= bits - 1;
    }
    __syncthreads();
    if (tid == 0) s_base += s_warp[31];
    __syncthreads();
  }
  if (tid == 0) {
    *A.ep.num_reset = s_base;
    int32_t* c = A.ep.counters;
    const float n = (float)A.n;
    A.ep.stats[DEX_STAT_SUCCESS_RATE] = (float)c[DEX_CNT_SUCCESS] / n;
    A.ep.stats[DEX_STAT_FAILURE_RATE] = (float)c[DEX_CNT_FAILURE] / n;
    A.ep.stats[DEX_STAT_TIMEOUT_RATE] = (float)c[DEX_CNT_TIMEOUT] / n;
    for (int k = 0; k < DEX_NUM_CRITERIA; ++k) A.ep.stats[DEX_STAT_CRIT0 + k] = (float)c[DEX_CNT_CRIT0 + k] / n;
    // termination_manager.py:323-339
    int cs = *A.ep.consecutive_successes;
    cs = (c[DEX_CNT_SUCCESS] > 0) ? cs + 1 : 0;
    *A.ep.consecutive_successes = min(cs, A.max_consecutive);
    if (A.ep.totals != nullptr) {
      for (int k = 0; k < DEX_NUM_COUNTERS; ++k) A.ep.totals[k] += c[k];
      A.ep.totals[DEX_NUM_COUNTERS] += 1;
    }
    for (int k = 0; k < DEX_NUM_COUNTERS; ++k) c[k] = 0;
    if (A.ep.rng_step != nullptr) *A.ep.rng_step += 1;
  }
}

// ------------------------------------------------------------------------------------------------
// small stand-alone helpers
// ------------------------------------------------------------------------------------------------
struct CoupleArgs {
  DexParams p;
  const float* in;
  float* out;
  int32_t n;
};

__global__ void apply_coupling_kernel(const __grid_constant__ CoupleArgs A) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (env, dof)
  if (i >= A.n * DEX_NUM_DOF) return;
  const int env = i / DEX_NUM_DOF, d = i - env * DEX_NUM_DOF;
  const int a = A.p.dof_to_active[d];
  float v = 0.0f;
  if (a >= 0) {
    v = A.in[(size_t)a * A.n + env];
    if (d >= DEX_NUM_BASE) v = v * A.p.dof_coupling_scale[d];
  }
  A.out[i] = v;
}

__global__ void extract_active_kernel(const __grid_constant__ CoupleArgs A) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (active, env)
  if (i >= A.n * DEX_NUM_ACTIVE) return;
  const int a = i / A.n, env = i - a * A.n;
  const float v = A.in[(size_t)env * DEX_NUM_DOF + A.p.inverse_src_dof[a]];
  A.out[i] = (a < DEX_NUM_BASE) ? v : v / A.p.inverse_scale[a];
}

__global__ void point_in_hand_frame_kernel(const float* p, const float* hp, const float* hq, float* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Quat q{hq[4 * i], hq[4 * i + 1], hq[4 * i + 2], hq[4 * i + 3]};
  const Vec3 r = quat_rotate_inverse(q, Vec3{p[3 * i] - hp[3 * i], p[3 * i + 1] - hp[3 * i + 1], p[3 * i + 2] - hp[3 * i + 2]});
  out[3 * i] = r.x;
  out[3 * i + 1] = r.y;
  out[3 * i + 2] = r.z;
}

__global__ void quat_mul_kernel(const float* a, const float* b, float* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Quat r = quat_mul(Quat{a[4 * i], a[4 * i + 1], a[4 * i + 2], a[4 * i + 3]},
                          Quat{b[4 * i], b[4 * i + 1], b[4 * i + 2], b[4 * i + 3]});
  out[4 * i] = r.x;
  out[4 * i + 1] = r.y;
  out[4 * i + 2] = r.z;
  out[4 * i + 3] = r.w;
}

// ------------------------------------------------------------------------------------------------
// host-side validation
// ------------------------------------------------------------------------------------------------
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

int check_params(const DexParams* p) {
  if (p == nullptr) return DEX_ERR_NULL;
  if (p->abi_version != DEX_ABI_VERSION) return DEX_ERR_VERSION;
  if (p->task != DEX_TASK_BASE && p->task != DEX_TASK_BLIND_GRASPING) return DEX_ERR_PARAM;
  if (p->num_bodies <= 0 || p->num_actors <= 0) return DEX_ERR_SIZE;
  if (p->num_actions != 18 && p->num_actions != 12 && p->num_actions != 6) return DEX_ERR_SIZE;
  if (p->obs_dim <= 0 || p->obs_dim > 1024) return DEX_ERR_SIZE;
  if (!(p->control_dt > 0.0f)) return DEX_ERR_PARAM;
  auto body_ok = [&](int b) { return b >= 0 && b < p->num_bodies; };
  if (!body_ok(p->hand_body)) return DEX_ERR_PARAM;
  for (int f = 0; f < DEX_NUM_FINGERS; ++f)
    if (!body_ok(p->tip_body[f]) || !body_ok(p->pad_body[f]) || !body_ok(p->contact_body[f])) return DEX_ERR_PARAM;
  if (p->task == DEX_TASK_BLIND_GRASPING) {
    if (!body_ok(p->box_body) || p->box_actor < 0 || p->box_actor >= p->num_actors) return DEX_ERR_PARAM;
  }
  for (int a = 0; a < DEX_NUM_ACTIVE; ++a)
    if (p->inverse_src_dof[a] < 0 || p->inverse_src_dof[a] >= DEX_NUM_DOF) return DEX_ERR_PARAM;
  for (int d = 0; d < DEX_NUM_DOF; ++d)
    if (p->dof_to_active[d] < -1 || p->dof_to_active[d] >= DEX_NUM_ACTIVE) return DEX_ERR_PARAM;
  for (int j = 0; j < DEX_NUM_ACTIVE_FINGER; ++j)
    if (p->active_finger_dof[j] < 0 || p->active_finger_dof[j] >= DEX_NUM_DOF) return DEX_ERR_PARAM;
  if (p->num_component_rows <= 0 || p->num_component_rows > 2 * DEX_NUM_REWARD_TERMS + 7) return DEX_ERR_PARAM;
  return DEX_OK;
}

int check_state(const DexParams* p, const DexStateBuffers* st) {
  if (st == nullptr) return DEX_ERR_NULL;
  if (!st->active_prev_targets || !st->active_rule_targets || !st->full_dof_targets || !st->prev_actions ||
      !st->prev_dof_pos || !st->contact_duration_steps || !st->prev_contact_binary || !st->episode_step_count)
    return DEX_ERR_NULL;
  if (p->need_common_rewards &&
      (!st->prev_finger_dof_vel || !st->prev_hand_vel || !st->prev_hand_ang_vel || !st->prev_contacts))
    return DEX_ERR_NULL;
  if (p->task == DEX_TASK_BLIND_GRASPING &&
      (!st->success_duration_steps || !st->success_conditions_met || !st->current_stage || !st->time_in_stage ||
       !st->stage_contact_duration || !st->just_transitioned_to_stage2 || !st->just_transitioned_to_stage3 ||
       !st->initial_box_positions || !st->failure_tensors))
    return DEX_ERR_NULL;
  return DEX_OK;
}

int check_sim(const DexParams* p, const DexSimViews* sim) {
  if (sim == nullptr) return DEX_ERR_NULL;
  if (!sim->dof_state || !sim->rigid_body_state || !sim->net_contact_force) return DEX_ERR_NULL;
  if (p->task == DEX_TASK_BLIND_GRASPING && !sim->actor_root_state) return DEX_ERR_NULL;
  if (!aligned16(sim->dof_state)) return DEX_ERR_ALIGN;
  return DEX_OK;
}

int launch_status() {
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

constexpr int kPostTile = 32;

template <bool BG>
int launch_post(const PostArgs& a, cudaStream_t s) {
  const size_t smem = (size_t)post_smem_floats(kPostTile, a.p.obs_dim) * sizeof(float);
  static bool configured = false;  // per template instance
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(post_physics_kernel<BG, kPostTile>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         200 * 1024);
    if (e != cudaSuccess) return (int)e;
    configured = true;
  }
  if (smem > 200 * 1024) return DEX_ERR_SIZE;
  const int grid = (a.n + kPostTile - 1) / kPostTile;
  post_physics_kernel<BG, kPostTile><<<grid, kPostTile, smem, s>>>(a);
  return launch_status();
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

const char* dex_strerror(int code) {
  switch (code) {
    case DEX_OK: return "ok";
    case DEX_ERR_NULL: return "dexstep: required pointer is NULL";
    case DEX_ERR_SIZE: return "dexstep: invalid size (n_envs, body/actor/action/obs counts)";
    case DEX_ERR_ALIGN: return "dexstep: dof_state must be 16-byte aligned";
    case DEX_ERR_PARAM: return "dexstep: invalid parameter block (index out of range, bad enum or control_dt)";
    case DEX_ERR_VERSION: return "dexstep: DexParams.abi_version does not match the library";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "dexstep: unknown error";
  }
}

int dex_abi_version(void) { return DEX_ABI_VERSION; }
int dex_sizeof_params(void) { return (int)sizeof(DexParams); }

int dex_pre_physics(const DexParams* params, const DexStateBuffers* state, const float* actions,
                    float* active_raw_targets, int32_t n_envs, void* stream) {
  int rc = check_params(params);
  if (rc) return rc;
  rc = check_state(params, state);
  if (rc) return rc;
  if (actions == nullptr) return DEX_ERR_NULL;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  PreArgs a{*params, *state, actions, active_raw_targets, n_envs};
  const int grid = (n_envs + kPreTile - 1) / kPreTile;
  pre_physics_kernel<<<grid, kPreTile, 0, (cudaStream_t)stream>>>(a);
  return launch_status();
}

int dex_post_physics(const DexParams* params, const DexSimViews* sim, const DexStateBuffers* state,
                     const DexOutBuffers* out, int32_t n_envs, int32_t mode, uint32_t flags, const float* uniforms,
                     uint64_t seed, uint64_t step_index, void* stream) {
  int rc = check_params(params);
  if (rc) return rc;
  rc = check_sim(params, sim);
  if (rc) return rc;
  rc = check_state(params, state);
  if (rc) return rc;
  if (out == nullptr || out->obs_buf == nullptr) return DEX_ERR_NULL;
  if (mode != DEX_MODE_FULL && mode != DEX_MODE_OBS_ONLY) return DEX_ERR_PARAM;
  if (mode == DEX_MODE_FULL &&
      (!out->rew_buf || !out->reset_buf || !out->success || !out->failure || !out->timeout || !out->reward_components ||
       !out->reset_bits || !out->counters))
    return DEX_ERR_NULL;
  if ((flags & DEX_FLAG_STALE_CONTACTS) && state->contact_forces == nullptr) return DEX_ERR_NULL;
  if (out->episode_sums != nullptr && (out->episode_lengths == nullptr || out->stat_totals == nullptr)) return DEX_ERR_NULL;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  PostArgs a{*params, *sim, *state, *out, n_envs, mode, flags, uniforms, seed, step_index};
  if (params->task == DEX_TASK_BLIND_GRASPING) return launch_post<true>(a, (cudaStream_t)stream);
  return launch_post<false>(a, (cudaStream_t)stream);
}

int dex_reset_idx(const DexParams* params, const DexSimViews* sim, const DexStateBuffers* state,
                  const uint8_t* reset_mask, int32_t n_envs, const float* uniforms, uint64_t seed, uint64_t step_index,
                  void* stream) {
  int rc = check_params(params);
  if (rc) return rc;
  rc = check_sim(params, sim);
  if (rc) return rc;
  rc = check_state(params, state);
  if (rc) return rc;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  ResetArgs a{*params, *sim, *state, reset_mask, n_envs, uniforms, seed, step_index};
  const int grid = (n_envs + 127) / 128;
  if (params->task == DEX_TASK_BLIND_GRASPING)
    reset_kernel<true><<<grid, 128, 0, (cudaStream_t)stream>>>(a);
  else
    reset_kernel<false><<<grid, 128, 0, (cudaStream_t)stream>>>(a);
  return launch_status();
}

int dex_step_epilogue(const DexParams* params, const DexEpilogueBuffers* ep, int32_t n_envs, void* stream) {
  int rc = check_params(params);
  if (rc) return rc;
  if (ep == nullptr || !ep->reset_bits || !ep->counters || !ep->env_ids || !ep->num_reset || !ep->stats ||
      !ep->consecutive_successes)
    return DEX_ERR_NULL;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  EpiArgs a{*ep, n_envs, params->max_consecutive_successes};
  epilogue_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(a);
  return launch_status();
}

int dex_apply_coupling(const DexParams* params, const float* active, float* full, int32_t n_envs, void* stream) {
  int rc = check_params(params);
  if (rc) return rc;
  if (!active || !full) return DEX_ERR_NULL;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  CoupleArgs a{*params, active, full, n_envs};
  const int total = n_envs * DEX_NUM_DOF;
  apply_coupling_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a);
  return launch_status();
}

int dex_extract_active_targets(const DexParams* params, const float* full, float* active, int32_t n_envs, void* stream) {
  int rc = check_params(params);
  if (rc) return rc;
  if (!active || !full) return DEX_ERR_NULL;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  CoupleArgs a{*params, full, active, n_envs};
  const int total = n_envs * DEX_NUM_ACTIVE;
  extract_active_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a);
  return launch_status();
}

int dex_point_in_hand_frame(const float* p_world, const float* hand_pos, const float* hand_quat, float* out, int32_t n,
                            void* stream) {
  if (!p_world || !hand_pos || !hand_quat || !out) return DEX_ERR_NULL;
  if (n <= 0) return DEX_ERR_SIZE;
  point_in_hand_frame_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(p_world, hand_pos, hand_quat, out, n);
  return launch_status();
}

int dex_quat_mul(const float* a, const float* b, float* out, int32_t n, void* stream) {
  if (!a || !b || !out) return DEX_ERR_NULL;
  if (n <= 0) return DEX_ERR_SIZE;
  quat_mul_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(a, b, out, n);
  return launch_status();
}

}  // extern "C"
