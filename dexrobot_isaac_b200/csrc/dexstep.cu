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
//   * lane = env for all arithmetic (envs are independent; no cross-env term on the path except the step statistics);
//   * path-internal state is SoA [K][N]: a warp's access to one row is one 128 B request;
//   * buffers whose layout is dictated from outside (Isaac Gym AoS sim tensors, actions, obs_buf, full_dof_targets)
//     are staged through shared memory per tile of 32 envs;
//   * dex_post_physics dispatches between three kernels over the same arithmetic: the persistent warp-specialised kernel
//     and the single-pass kernel of post_physics_v2.cuh (lean production path, chosen by tiles per SM) and the
//     table-driven tile kernel of post_physics.cuh (obs_dict materialisation, OBS_ONLY, stale contacts, A/B runs);
//   * dex_pre_physics: one kernel, 8 warps x 32 envs per CTA (this file).
#include <cuda_runtime.h>
#include <stdint.h>

#include "dexstep.h"

#include "dexstep_common.cuh"
#include "post_physics.cuh"
#include "post_physics_v2.cuh"

namespace {

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

// One CTA of 8 warps owns 32 consecutive envs.  The 18 active targets are work items taken round-robin by the
// warps with lane = env: the SoA state rows (prev / rule targets in, new targets out, prev_actions out) move as
// full 128 B lines, the branch on the target index is warp-uniform.  The AoS tiles (actions in,
// full_dof_targets out) are contiguous in global memory and cross a small shared tile with odd row stride.
constexpr int kPreTile = 32;
constexpr int kPreThreads = 256;
constexpr int kPreWarps = kPreThreads / 32;
constexpr int kActS = 19;  // odd smem row stride for <=18 actions / 18 active targets

__global__ void __launch_bounds__(kPreThreads) pre_physics_kernel(const __grid_constant__ PreArgs A) {
  __shared__ float s_act[kPreTile * kActS];
  __shared__ float s_next[kPreTile * kActS];  // column 18 of every row stays 0: the source of the DOFs no active target drives
  __shared__ float2 s_map[DEX_NUM_DOF];       // DOF -> (column of s_next as int bits, coupling scale)
  const DexParams& P = A.p;
  const DexStateBuffers& st = A.st;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int e0 = blockIdx.x * kPreTile;
  const int ne = min(kPreTile, A.n - e0);
  const size_t N = (size_t)A.n;
  const int NA = P.num_actions;
  const size_t env = (size_t)e0 + lane;
  const bool act_lane = lane < ne;

  // coupling tables -> shared (a per-lane index into the constant bank would serialise)
  if (tid < DEX_NUM_DOF) {
    const int a = P.dof_to_active[tid];
    s_map[tid] = make_float2(__int_as_float(a >= 0 ? a : DEX_NUM_ACTIVE), (tid < DEX_NUM_BASE) ? 1.0f : P.dof_coupling_scale[tid]);
  }
  if (tid < kPreTile) s_next[tid * kActS + DEX_NUM_ACTIVE] = 0.0f;
  pdl_wait();               // the previous kernel of the stream (post-physics of the last step) has completed
  pdl_launch_dependents();  // the next kernel's CTAs may be scheduled as this grid drains
  // issue every global load of the tile first: the action rows (warp w: env rows w, w+8, ..; lane = column) ...
  constexpr int R = kPreTile / kPreWarps;
  const float* asrc = A.actions + ((size_t)e0 + warp) * NA + lane;
  float av[R];
#pragma unroll
  for (int j = 0; j < R; ++j)
    av[j] = (lane < NA && warp + j * kPreWarps < ne) ? __ldg(asrc + j * kPreWarps * NA) : 0.0f;
  // ... and the prev / rule target rows of this warp's items (a = warp, warp + 8, warp + 16)
  float prev[3], rule[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int a = warp + j * kPreWarps;
    if (a < DEX_NUM_ACTIVE && act_lane) {
      prev[j] = st.active_prev_targets[a * N + env];
      rule[j] = st.active_rule_targets[a * N + env];
    }
  }
#pragma unroll
  for (int j = 0; j < R; ++j)
    if (lane < NA && warp + j * kPreWarps < ne) s_act[(warp + j * kPreWarps) * kActS + lane] = av[j];
  __syncthreads();

  const int fs = P.policy_controls_base ? DEX_NUM_BASE : 0;  // default_rules.py:53,97
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int a = warp + j * kPreWarps;
    if (a < DEX_NUM_ACTIVE && act_lane) {
      const float lo = P.active_lower[a], hi = P.active_upper[a], md = P.max_deltas[a];
      const bool controlled = (a < DEX_NUM_BASE) ? (P.policy_controls_base != 0) : (P.policy_controls_fingers != 0);
      float t = rule[j];
      if (controlled) {
        const float x = s_act[lane * kActS + ((a < DEX_NUM_BASE) ? a : fs + (a - DEX_NUM_BASE))];
        if (P.control_mode == DEX_CONTROL_POSITION)
          t = (x + 1.0f) * 0.5f * (hi - lo) + lo;  // scaling.py:28-45
        else
          t = prev[j] + x * md;  // scaling.py:48-63
      }
      if (P.control_mode == DEX_CONTROL_POSITION_DELTA) t = clampf(t, lo, hi);  // default_rules.py:107-110
      if (A.raw_out != nullptr) A.raw_out[a * N + env] = t;
      if (P.filter_velocity_clamp) t = prev[j] + clampf(t - prev[j], -md, md);  // rules.py:141-166
      if (P.filter_position_clamp) t = clampf(t, lo, hi);                      // rules.py:168-190
      s_next[lane * kActS + a] = t;
      st.active_prev_targets[a * N + env] = t;  // action_processor.py:342
    }
  }
  // prev_actions rows (observation_encoder.py:287-296): row k for 32 envs = one 128 B line
  if (act_lane) {
    float* pa = st.prev_actions + (size_t)warp * N + env;
    const float* sa = s_act + lane * kActS + warp;
#pragma unroll
    for (int j = 0; j < 3; ++j)  // NA <= 18: rows warp, warp + 8, warp + 16
      if (warp + j * kPreWarps < NA) pa[(size_t)j * kPreWarps * N] = sa[j * kPreWarps];
  }
  __syncthreads();
  // coupling 18 -> 26 (action_processor.py:571-614) straight into the contiguous full_dof_targets tile: element i of the
  // tile is DOF d = i % 26 of env e = i / 26; one division per thread, then (e, d) advance by (9, 22) per 256 elements
  float* dst = st.full_dof_targets + (size_t)e0 * DEX_NUM_DOF + tid;
  const int n_out = ne * DEX_NUM_DOF;
  int e = tid / DEX_NUM_DOF, d = tid - e * DEX_NUM_DOF;
#pragma unroll
  for (int j = 0; j < (kPreTile * DEX_NUM_DOF + kPreThreads - 1) / kPreThreads; ++j) {
    if (tid + j * kPreThreads < n_out) {
      const float2 m = s_map[d];
      dst[j * kPreThreads] = s_next[e * kActS + __float_as_int(m.x)] * m.y;  // (an undriven DOF reads the zero column: 0 * scale)
    }
    e += kPreThreads / DEX_NUM_DOF;
    d += kPreThreads % DEX_NUM_DOF;
    if (d >= DEX_NUM_DOF) {
      d -= DEX_NUM_DOF;
      ++e;
    }
  }
}

// The cost of A launch in a stream of step kernels, for bench.py: same grid / block / programmatic-launch attribute as
// dex_pre_physics, no work.  What dex_pre_physics costs beyond this is what its loads, arithmetic and stores cost.
struct NullArgs {
  int32_t n;
};
__global__ void __launch_bounds__(kPreThreads) launch_floor_kernel(const __grid_constant__ NullArgs A) {
  pdl_wait();
  pdl_launch_dependents();
  (void)A;
}

// ------------------------------------------------------------------------------------------------
// fixed-capacity actor-index lists for Isaac Gym's indexed setters: see include/dexstep.h
// ------------------------------------------------------------------------------------------------
__global__ void reset_actor_indices_kernel(const int32_t* env_ids, const int32_t* num_reset, const int capacity,
                                           const int num_actors, const int dof_actor, int32_t* root_ids, int32_t* dof_ids,
                                           int32_t* overflow) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= capacity) return;
  const int k = *num_reset;
  // padding: the first reset env again (its rows hold the state just written, so applying them twice changes
  // nothing); without any reset, env 0 - whose rows hold its current simulation state
  const int pad = k > 0 ? env_ids[0] : 0;
  const int e = i < k ? env_ids[i] : pad;
  dof_ids[i] = e * num_actors + dof_actor;
  for (int a = 0; a < num_actors; ++a) root_ids[i * num_actors + a] = e * num_actors + a;
  if (i == 0 && overflow != nullptr) *overflow += max(k - capacity, 0);
}

// ------------------------------------------------------------------------------------------------
// policy-input stage, training half: rl_games' RunningMeanStd in training mode first merges the moments of
// the current batch of observations into its running statistics and only then normalises that batch with the UPDATED
// statistics - a reduction over all envs sits between obs_buf and the network input, so this half cannot ride in the
// post-physics pass.  Three small launches on the caller's stream, no host read: column sums -> merge -> normalise.
// ------------------------------------------------------------------------------------------------
constexpr int kMomThreads = 256;
constexpr int kMomRowBlocks = 64;

// grid (ceil(W / 32), kMomRowBlocks): lane = column, the 8 warps of a block stride over the block's rows; fp64 partial
// sums (the batch moments rl_games takes in fp32 are reproduced to well below 1e-6 relative), one atomic per column and block
__global__ void __launch_bounds__(kMomThreads) obs_moments_kernel(const float* __restrict__ obs, const int n, const int W, double* acc) {
  __shared__ double s_sum[kMomThreads / 32][32], s_sq[kMomThreads / 32][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + lane;
  const int rows_per = (n + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per, r1 = min(n, r0 + rows_per);
  double sum = 0.0, sq = 0.0;
  if (col < W)
    for (int r = r0 + warp; r < r1; r += kMomThreads / 32) {
      const double v = (double)obs[(size_t)r * W + col];
      sum += v;
      sq += v * v;
    }
  s_sum[warp][lane] = sum;
  s_sq[warp][lane] = sq;
  __syncthreads();
  if (warp == 0 && col < W) {
#pragma unroll
    for (int w = 1; w < kMomThreads / 32; ++w) {
      sum += s_sum[w][lane];
      sq += s_sq[w][lane];
    }
    atomicAdd(acc + col, sum);
    atomicAdd(acc + W + col, sq);
  }
}

// one block: rl_games/algos_torch/running_mean_std.py _update_mean_var_count_from_moments (Chan et al.), float64 state,
// fed with the batch's fp32 mean and UNBIASED fp32 variance like `input.mean(0)`, `input.var(0)`
__global__ void obs_moments_merge_kernel(const double* acc, const int n, const int W, double* running_mean, double* running_var,
                                         double* count, float* norm_mean, float* norm_scale, const float epsilon) {
  const double cnt = *count;
  __syncthreads();  // every thread has read the old count before thread 0 replaces it
  for (int c = threadIdx.x; c < W; c += blockDim.x) {
    const double s = acc[c], q = acc[W + c];
    const double mean_d = s / n;
    const double var_d = n > 1 ? fmax((q - s * mean_d) / (n - 1), 0.0) : 0.0;
    const double bm = (double)(float)mean_d, bv = (double)(float)var_d;  // the fp32 tensors rl_games hands to the merge
    const double rm = running_mean[c], rv = running_var[c];
    const double delta = bm - rm, tot = cnt + (double)n;
    const double new_mean = rm + delta * (double)n / tot;
    const double m2 = rv * cnt + bv * (double)n + delta * delta * cnt * (double)n / tot;
    const double new_var = m2 / tot;
    running_mean[c] = new_mean;
    running_var[c] = new_var;
    norm_mean[c] = (float)new_mean;
    norm_scale[c] = 1.0f / sqrtf((float)new_var + epsilon);
  }
  if (threadIdx.x == 0) *count = cnt + (double)n;
}

// y = clamp((x - mean) * scale, -clip, clip), NaN passes through: the arithmetic of the fused stage (copy_obs_norm_rows)
__global__ void obs_normalize_kernel(const float* __restrict__ obs, const float* __restrict__ mean, const float* __restrict__ scale,
                                     const float clip, const size_t total, const int W, float* __restrict__ out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % (size_t)W);
    const float y = (obs[i] - mean[c]) * scale[c];
    out[i] = (y != y) ? y : fminf(fmaxf(y, -clip), clip);
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
  if (p->num_bodies <= 0 || p->num_actors <= 0 || p->num_contact_bodies <= 0) return DEX_ERR_SIZE;
  if (p->num_actions != 18 && p->num_actions != 12 && p->num_actions != 6) return DEX_ERR_SIZE;
  if (p->obs_dim <= 0 || p->obs_dim > 256) return DEX_ERR_SIZE;  // kMaxObs of the kernels
  if (!(p->control_dt > 0.0f)) return DEX_ERR_PARAM;
  auto body_ok = [&](int b) { return b >= 0 && b < p->num_bodies; };
  auto cbody_ok = [&](int b) { return b >= 0 && b < p->num_contact_bodies; };
  if (!body_ok(p->hand_body)) return DEX_ERR_PARAM;
  if (p->rigid_row_floats != 0) {  // packed host staging: rows of 7..13 floats, the 13-float hand row must end inside its env
    const int rw = p->rigid_row_floats;
    if (rw < 7 || rw > 13 || p->hand_body * rw + 13 > p->num_bodies * rw) return DEX_ERR_PARAM;
  }
  for (int f = 0; f < DEX_NUM_FINGERS; ++f)
    if (!body_ok(p->tip_body[f]) || !body_ok(p->pad_body[f]) || !cbody_ok(p->contact_body[f])) return DEX_ERR_PARAM;
  if (p->task == DEX_TASK_BLIND_GRASPING) {
    if (!cbody_ok(p->box_body) || p->box_actor < 0 || p->box_actor >= p->num_actors) return DEX_ERR_PARAM;
  }
  for (int a = 0; a < DEX_NUM_ACTIVE; ++a)
    if (p->inverse_src_dof[a] < 0 || p->inverse_src_dof[a] >= DEX_NUM_DOF) return DEX_ERR_PARAM;
  for (int d = 0; d < DEX_NUM_DOF; ++d)
    if (p->dof_to_active[d] < -1 || p->dof_to_active[d] >= DEX_NUM_ACTIVE) return DEX_ERR_PARAM;
  for (int j = 0; j < DEX_NUM_ACTIVE_FINGER; ++j)
    if (p->active_finger_dof[j] < 0 || p->active_finger_dof[j] >= DEX_NUM_DOF) return DEX_ERR_PARAM;
  if (p->num_component_rows <= 0 || p->num_component_rows > 2 * DEX_NUM_REWARD_TERMS + 7) return DEX_ERR_PARAM;
  for (int i = 0; i < DEX_NUM_REWARD_TERMS; ++i)  // rows index the reward_components buffer and the shared slab
    if (p->component_row[i] < -1 || p->component_row[i] + 1 >= p->num_component_rows) return DEX_ERR_PARAM;
  if (p->total_row < 0 || p->total_row >= p->num_component_rows) return DEX_ERR_PARAM;
  if (p->termination_row < 0 || p->termination_row + 5 >= p->num_component_rows) return DEX_ERR_PARAM;
  for (int k = 0; k < DEX_NUM_OBS_KEYS; ++k)
    if (p->obs_offset[k] < -1 || p->obs_offset[k] >= p->obs_dim) return DEX_ERR_PARAM;
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

}  // namespace

// ------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------------
// scratch of dex_post_physics_noalloc: tile records for up to DEX_NOALLOC_MAX_ENVS envs (4 words per 32-env tile, 256 KB).
// Shared by every call on this GPU, keeps the last call's contents: each call invalidates the part it uses first.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned long long g_noalloc_tile_records[4 * (DEX_NOALLOC_MAX_ENVS / 32)];

// A record is current when bits 48..63 == (*rng_step + 1) & 0xffff.  Writing that stamp with bit 15 flipped makes every
// record of the coming step invalid whatever an earlier call (of any pipeline, at any step count) left there - a plain
// clear to zero would pass as current once every 65 536 steps.
__global__ void noalloc_invalidate_records_kernel(unsigned long long* rec, const int words, const uint64_t* rng_step) {
  const unsigned stamp = ((unsigned)*rng_step + 1u) & 0xffffu;
  const unsigned long long v = (unsigned long long)(stamp ^ 0x8000u) << 48;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < words) rec[i] = v;
}

extern "C" {

const char* dex_strerror(int code) {
  switch (code) {
    case DEX_OK: return "ok";
    case DEX_ERR_NULL: return "dexstep: required pointer is NULL";
    case DEX_ERR_SIZE: return "dexstep: invalid size (n_envs, body/actor/action/obs counts)";
    case DEX_ERR_ALIGN: return "dexstep: dof_state must be 16-byte aligned";
    case DEX_ERR_PARAM: return "dexstep: invalid parameter block (index out of range, bad enum or control_dt)";
    case DEX_ERR_VERSION: return "dexstep: DexParams.abi_version does not match the library";
    case DEX_ERR_CAPACITY: return "dexstep: n_envs exceeds DEX_NOALLOC_MAX_ENVS (262144), the capacity of the built-in scratch of dex_post_physics_noalloc";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "dexstep: unknown error";
  }
}

static int g_v2_launches = 0;
int dex_debug_v2_launches(void) { return g_v2_launches; }
int dex_debug_last_post_kernel(void) { return pp2::g_last_geometry; }
int dex_debug_launch_floor(int32_t n_envs, void* stream) {
  if (n_envs <= 0) return DEX_ERR_SIZE;
  const cudaError_t e = launch_step_kernel(launch_floor_kernel, (n_envs + kPreTile - 1) / kPreTile, kPreThreads, 0,
                                           (cudaStream_t)stream, NullArgs{n_envs});
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
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
  const cudaError_t e = launch_step_kernel(pre_physics_kernel, grid, kPreThreads, 0, (cudaStream_t)stream, a);
  if (e != cudaSuccess) (void)cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

static int check_epilogue(const DexEpilogueBuffers* ep) {
  if (ep == nullptr || !ep->reset_bits || !ep->counters || !ep->env_ids || !ep->num_reset || !ep->stats ||
      !ep->consecutive_successes)
    return DEX_ERR_NULL;
  return DEX_OK;
}

int dex_post_physics(const DexParams* params, const DexSimViews* sim, const DexStateBuffers* state,
                     const DexOutBuffers* out, const DexEpilogueBuffers* epilogue, int32_t n_envs, int32_t mode,
                     uint32_t flags, const float* uniforms, uint64_t seed, uint64_t step_index, void* stream) {
  int rc = check_params(params);
  if (rc) return rc;
  rc = check_sim(params, sim);
  if (rc) return rc;
  rc = check_state(params, state);
  if (rc) return rc;
  if (out == nullptr || out->obs_buf == nullptr) return DEX_ERR_NULL;
  if (out->obs_norm != nullptr && (out->norm_mean == nullptr || out->norm_scale == nullptr)) return DEX_ERR_NULL;
  if (mode != DEX_MODE_FULL && mode != DEX_MODE_OBS_ONLY) return DEX_ERR_PARAM;
  if (mode == DEX_MODE_FULL &&
      (!out->rew_buf || !out->reset_buf || !out->success || !out->failure || !out->timeout || !out->reward_components ||
       !out->reset_bits || !out->counters))
    return DEX_ERR_NULL;
  if ((flags & DEX_FLAG_STALE_CONTACTS) && state->contact_forces == nullptr) return DEX_ERR_NULL;
  if (out->episode_sums != nullptr && (out->episode_lengths == nullptr || out->stat_totals == nullptr)) return DEX_ERR_NULL;
  if (out->stat_slices < 0 || out->stat_slices > 4096) return DEX_ERR_SIZE;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  const bool fused = epilogue != nullptr && mode == DEX_MODE_FULL;
  if (fused) {
    rc = check_epilogue(epilogue);
    if (rc) return rc;
    if (epilogue->tile_records == nullptr || epilogue->rng_step == nullptr || epilogue->rng_step != state->rng_step ||
        epilogue->reset_bits != out->reset_bits || epilogue->counters != out->counters)
      return DEX_ERR_NULL;
  }
  // The persistent warp-specialised kernel (post_physics_v2.cuh) takes every launch it covers; the table-driven tile
  // kernel (post_physics.cuh) keeps the rest: obs_dict materialisation, the observation-only pass and the stale-contact
  // pass of reset(), and layouts v2 has no stage for.  DEXSTEP_POST_V1=1 forces the tile kernel (A/B runs).
  static const bool force_v1 = []() {
    const char* e = getenv("DEXSTEP_POST_V1");
    return e != nullptr && e[0] == '1';
  }();
  if (!force_v1 && !(flags & DEX_FLAG_TILE_KERNEL) && pp2::post2_eligible(*params, *out, mode, flags)) {
    static thread_local pp2::Post2Args b;
    b.p = *params;
    b.sim = *sim;
    b.st = *state;
    b.out = *out;
    b.ep = fused ? *epilogue : DexEpilogueBuffers{};
    b.n = n_envs;
    b.mode = mode;
    b.flags = flags;
    b.fused_epilogue = fused ? 1 : 0;
    b.uniforms = uniforms;
    b.seed = seed;
    b.step = step_index;
    ++g_v2_launches;
    return params->task == DEX_TASK_BLIND_GRASPING ? pp2::launch_post2_t<true>(b, (cudaStream_t)stream)
                                                   : pp2::launch_post2_t<false>(b, (cudaStream_t)stream);
  }
  static thread_local PostArgs a;  // ~7 KB of launch arguments: keep it off the stack
  a.p = *params;
  a.sim = *sim;
  a.st = *state;
  a.out = *out;
  a.ep = fused ? *epilogue : DexEpilogueBuffers{};
  a.n = n_envs;
  a.mode = mode;
  a.flags = flags;
  a.fused_epilogue = fused ? 1 : 0;
  a.uniforms = uniforms;
  a.seed = seed;
  a.step = step_index;
  pp2::g_last_geometry = 0;
  return launch_post(a, (cudaStream_t)stream);
}

// ---- dex_post_physics_noalloc: the fused-epilogue launch with its tile-record scratch in a static device array ----
// (no counterpart in the reference; see include/dexstep.h).  In HOST code the array's name is only a stand-in for the
// symbol: its device address comes from cudaGetSymbolAddress, once per device, with that device current.
int dex_post_physics_noalloc(const DexParams* params, const DexSimViews* sim, const DexStateBuffers* state,
                             const DexOutBuffers* out, const DexEpilogueBuffers* epilogue, int32_t n_envs, int32_t mode,
                             uint32_t flags, const float* uniforms, uint64_t seed, uint64_t step_index, void* stream) {
  if (params == nullptr || sim == nullptr || state == nullptr || out == nullptr || epilogue == nullptr) return DEX_ERR_NULL;
  if (mode != DEX_MODE_FULL) return DEX_ERR_PARAM;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  if (n_envs > DEX_NOALLOC_MAX_ENVS) return DEX_ERR_CAPACITY;
  if (state->rng_step == nullptr) return DEX_ERR_NULL;  // the records are stamped with the device-resident step counter
  int rc = check_params(params);  // everything that can be rejected is rejected before the first launch
  if (rc) return rc;
  rc = check_sim(params, sim);
  if (rc) return rc;
  rc = check_state(params, state);
  if (rc) return rc;
  rc = check_epilogue(epilogue);
  if (rc) return rc;
  constexpr int kMaxDevices = 64;
  static unsigned long long* s_records[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return DEX_ERR_SIZE;
  if (s_records[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_noalloc_tile_records);
    if (e != cudaSuccess) {
      (void)cudaGetLastError();
      return (int)e;
    }
    s_records[dev] = static_cast<unsigned long long*>(ptr);
  }
  unsigned long long* const rec = s_records[dev];
  const int words = 4 * ((n_envs + 31) / 32);
  noalloc_invalidate_records_kernel<<<(words + 255) / 256, 256, 0, (cudaStream_t)stream>>>(rec, words, state->rng_step);
  e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  DexEpilogueBuffers ep = *epilogue;
  ep.tile_records = reinterpret_cast<uint64_t*>(rec);
  return dex_post_physics(params, sim, state, out, &ep, n_envs, mode, flags, uniforms, seed, step_index, stream);
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
  rc = check_epilogue(ep);
  if (rc) return rc;
  if (n_envs <= 0) return DEX_ERR_SIZE;
  epilogue_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(*ep, n_envs, params->max_consecutive_successes);
  return launch_status();
}

int dex_reset_actor_indices(const DexEpilogueBuffers* ep, int32_t capacity, int32_t num_actors, int32_t dof_actor,
                            int32_t* root_actor_ids, int32_t* dof_actor_ids, int32_t* overflow, void* stream) {
  if (ep == nullptr || ep->env_ids == nullptr || ep->num_reset == nullptr || root_actor_ids == nullptr || dof_actor_ids == nullptr)
    return DEX_ERR_NULL;
  if (capacity <= 0 || num_actors <= 0 || dof_actor < 0 || dof_actor >= num_actors) return DEX_ERR_SIZE;
  reset_actor_indices_kernel<<<(capacity + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      ep->env_ids, ep->num_reset, capacity, num_actors, dof_actor, root_actor_ids, dof_actor_ids, overflow);
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
}

int dex_policy_input_update(const float* obs_buf, int32_t n_envs, int32_t obs_dim, double* running_mean, double* running_var,
                            double* count, float* norm_mean, float* norm_scale, float* obs_norm, float epsilon, float clip,
                            double* scratch, void* stream) {
  if (!obs_buf || !running_mean || !running_var || !count || !norm_mean || !norm_scale || !scratch) return DEX_ERR_NULL;
  if (n_envs <= 0 || obs_dim <= 0 || obs_dim > 1024) return DEX_ERR_SIZE;
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(double) * 2 * (size_t)obs_dim, s);
  if (e != cudaSuccess) return (int)e;
  const int row_blocks = n_envs < kMomRowBlocks * 8 ? (n_envs + 7) / 8 : kMomRowBlocks;
  obs_moments_kernel<<<dim3((obs_dim + 31) / 32, row_blocks), kMomThreads, 0, s>>>(obs_buf, n_envs, obs_dim, scratch);
  obs_moments_merge_kernel<<<1, 256, 0, s>>>(scratch, n_envs, obs_dim, running_mean, running_var, count, norm_mean, norm_scale,
                                            epsilon);
  if (obs_norm != nullptr) {
    const size_t total = (size_t)n_envs * obs_dim;
    const int blocks = (int)((total + 4 * 256 - 1) / (4 * 256));
    obs_normalize_kernel<<<blocks > 4 * 148 ? 4 * 148 : blocks, 256, 0, s>>>(obs_buf, norm_mean, norm_scale, clip, total, obs_dim, obs_norm);
  }
  e = cudaGetLastError();
  return e == cudaSuccess ? DEX_OK : (int)e;
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

#ifdef DEXSTEP_TL2
/* debug builds only (tools/timeline_v2.py): device buffer [grid][8][32] of uint64 for the v2 kernel's phase marks */
int dex_debug_set_timeline2(void* dev_ptr) { return (int)cudaMemcpyToSymbol(pp2::g_tl2, &dev_ptr, sizeof(dev_ptr)); }
#endif

#ifdef DEXSTEP_TIMELINE
/* debug builds only: device buffer [grid][8] of uint64 that receives per-CTA phase timestamps */
int dex_debug_set_timeline(void* dev_ptr) {
  return (int)cudaMemcpyToSymbol(g_timeline, &dev_ptr, sizeof(dev_ptr));
}
#endif

}  // extern "C"
