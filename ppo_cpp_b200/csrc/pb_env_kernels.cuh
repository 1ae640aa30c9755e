// pb_env_kernels.cuh — the fused vector-env step: action clamp -> (autoreset | frame_skip x mj_step -> cfrc_ext ->
// reward/termination/obs) -> wrapper chain, one env per warp, everything between the state load and the output
// store in shared memory.  Mirrors ParVectorEnv::step / reset (libs/gymcpp/gym.h:310-365), the four task headers
// (libs/gymcpp/mujoco/*_v?.h) and the wrapper chain of src/ppo_continuous_action.cpp:41-49.
#pragma once
#include "pb_sim_step.cuh"

#include "pb_env_types.h"

// ------------------------------------------------------------------ std::mt19937 + libstdc++ distributions (lane 0)
__device__ inline void mt_seed(uint32_t *mt, int *idx, uint32_t seed) {
  uint32_t x = seed;
  mt[0] = x;
  for (int i = 1; i < 624; i++) { x = 1812433253u * (x ^ (x >> 30)) + (uint32_t)i; mt[i] = x; }
  *idx = 624;
}
__device__ inline uint32_t mt_u32(uint32_t *mt, int *idx) {
  if (*idx >= 624) {
    for (int i = 0; i < 624; i++) {
      uint32_t y = (mt[i] & 0x80000000u) | (mt[(i + 1) % 624] & 0x7fffffffu);
      mt[i] = mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    *idx = 0;
  }
  uint32_t y = mt[(*idx)++];
  y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
  return y;
}
struct Rng { uint32_t *mt; int idx; int ok; double saved; };
__device__ inline double rng_canonical(Rng &r) {
  double sum = (double)mt_u32(r.mt, &r.idx);
  sum += (double)mt_u32(r.mt, &r.idx) * 4294967296.0;
  double ret = sum / 18446744073709551616.0;
  if (ret >= 1.0) ret = 0.99999999999999988898;
  return ret;
}
__device__ inline double rng_uniform(Rng &r, double a, double b) { return __dadd_rn(__dmul_rn(b - a, rng_canonical(r)), a); }
__device__ inline double rng_normal(Rng &r, double mean, double stddev) {
  double ret;
  if (r.ok) { r.ok = 0; ret = r.saved; }
  else {
    double x, y, r2;
    do {
      x = __dadd_rn(__dmul_rn(2.0, rng_canonical(r)), -1.0);
      y = __dadd_rn(__dmul_rn(2.0, rng_canonical(r)), -1.0);
      r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
    } while (r2 > 1.0 || r2 == 0.0);
    double mult = sqrt(-2.0 * log(r2) / r2);
    r.saved = __dmul_rn(x, mult); r.ok = 1;
    ret = __dmul_rn(y, mult);
  }
  return __dadd_rn(__dmul_rn(ret, stddev), mean);
}

// reset noise per task (half_cheetah_v5.h:62-69, hopper_v5.h:69-76, ant_v5.h:71-78, humanoid_v4.h:58-65)
template <class C>
__device__ void task_reset_noise(EnvS<C> &s, const DevModel &m, Rng &r) {
  if (C::TASK == PB_TASK_HALFCHEETAH) {
    for (int i = 0; i < C::NQ; i++) s.qpos[i] = (float)((double)m.qpos0[i] + rng_uniform(r, -0.1, 0.1));
    for (int i = 0; i < C::NV; i++) s.qvel[i] = (float)(0.0 + rng_normal(r, 0.0, 0.1));
  } else if (C::TASK == PB_TASK_HOPPER) {
    for (int i = 0; i < C::NQ; i++) s.qpos[i] = (float)((double)m.qpos0[i] + rng_uniform(r, -5e-3, 5e-3));
    for (int i = 0; i < C::NV; i++) s.qvel[i] = (float)(0.0 + rng_uniform(r, -5e-3, 5e-3));
  } else if (C::TASK == PB_TASK_ANT) {
    for (int i = 0; i < C::NQ; i++) s.qpos[i] = (float)((double)m.qpos0[i] + rng_uniform(r, -0.1, 0.1));
    for (int i = 0; i < C::NV; i++) s.qvel[i] = (float)(0.0 + 0.1 * rng_normal(r, 0.0, 1.0));
  } else {
    for (int i = 0; i < C::NQ; i++) s.qpos[i] = (float)((double)m.qpos0[i] + rng_uniform(r, -1e-2, 1e-2));
    for (int i = 0; i < C::NV; i++) s.qvel[i] = (float)(0.0 + rng_uniform(r, -1e-2, 1e-2));
  }
}

// raw observation of the task from the last forward evaluation (writeState of each task header)
template <class C>
__device__ void task_write_obs(const EnvS<C> &s, float *obs, int lane) {
  constexpr int SKIP = (C::TASK == PB_TASK_HALFCHEETAH || C::TASK == PB_TASK_HOPPER) ? 1 : 2;
  constexpr int NP = C::NQ - SKIP;
  for (int i = lane; i < NP; i += 32) obs[i] = s.qpos[SKIP + i];
  for (int i = lane; i < C::NV; i += 32) {
    float v = s.qvel[i];
    if (C::TASK == PB_TASK_HOPPER) v = clampf(v, -10.f, 10.f);
    obs[NP + i] = v;
  }
  constexpr int O = NP + C::NV;
  if (C::TASK == PB_TASK_ANT) {
    for (int i = lane; i < (C::NB - 1) * 6; i += 32) obs[O + i] = clampf((&s.cfrc[1][0])[i], -1.f, 1.f);
  } else if (C::TASK == PB_TASK_HUMANOID) {
    for (int i = lane; i < C::NB * 10; i += 32) obs[O + i] = (&s.cinert[0][0])[i];
    for (int i = lane; i < C::NB * 6; i += 32) obs[O + C::NB * 10 + i] = (&s.cvel[0][0])[i];
    for (int i = lane; i < C::NV; i += 32) obs[O + C::NB * 16 + i] = s.qfrc_actuator[i];
    for (int i = lane; i < C::NB * 6; i += 32) obs[O + C::NB * 16 + C::NV + i] = (&s.cfrc[0][0])[i];
  }
}

// Welford update with batch_count = 1 in the reference's float32 operation order (stateful_observation.h:64-84);
// intrinsics keep nvcc from contracting the mul/add pairs into FMAs.
__device__ __forceinline__ void welford_f32(float x, float &mean, float &var, float count, float tot) {
  float delta = __fsub_rn(x, mean);
  float new_mean = __fadd_rn(mean, __fdiv_rn(__fmul_rn(delta, 1.0f), tot));
  float m_a = __fmul_rn(var, count);
  float M2 = __fadd_rn(__fadd_rn(m_a, 0.0f), __fdiv_rn(__fmul_rn(__fmul_rn(__fmul_rn(delta, delta), count), 1.0f), tot));
  var = __fdiv_rn(M2, tot);
  mean = new_mean;
}

// Per-env scalars that live in registers between the prologue and the epilogue of one env step (warp-uniform).
struct EnvCtx {
  int env, elapsed;
  bool do_reset, fwd_only;
  float xref, x_before;
};

// ---- prologue: load the persistent state of env `env`, clamp the action, apply the reset noise if the env resets
template <class C>
__device__ void env_begin(EnvS<C> &s, const DevModel &m, const EnvBuffers &eb, const StepArgs &a, int env, int lane, EnvCtx &c) {
  constexpr int NXY = (C::TASK == PB_TASK_HALFCHEETAH || C::TASK == PB_TASK_HOPPER) ? 1 : 2;
  for (int i = lane; i < C::NQ; i += 32) s.qpos[i] = eb.qpos[(size_t)env * C::NQ + i];
  for (int i = lane; i < C::NV; i += 32) { s.qvel[i] = eb.qvel[(size_t)env * C::NV + i]; s.warm[i] = eb.warm[(size_t)env * C::NV + i]; }
  for (int i = lane; i < C::NU; i += 32) {
    float u = a.actions ? a.actions[(size_t)env * C::NU + i] : 0.f;
    if (a.clip_actions) u = clampf(u, a.act_min, a.act_max);
    s.ctrl[i] = u;
  }
  if (lane == 0) { s.overflow = 0; s.ncon = 0; s.nefc = 0; s.pgs_sweeps = 0; s.pgs_rowsteps = 0;
#ifdef PB_IMBALANCE
    s.dbg_t[0] = s.dbg_t[1] = 0;
#endif
  }
  for (int it = lane; it < C::NB * 6; it += 32) (&s.cfrc[0][0])[it] = 0.f;
  __syncwarp();
  c.env = env;
  c.fwd_only = a.mode == 2;
  c.do_reset = c.fwd_only || a.mode == 1 || eb.autoreset[env] != 0;
  c.xref = eb.xref[env];
  c.elapsed = eb.elapsed[env];
  if (c.do_reset) {
    if (lane == 0 && !c.fwd_only) {
      Rng r = {eb.mt + (size_t)env * 624, eb.mt_idx[env], eb.nrm_ok[env], eb.nrm_saved[env]};
      if (a.mode == 1 && a.seed + env > 0) mt_seed(r.mt, &r.idx, (uint32_t)(a.seed + env));
      task_reset_noise<C>(s, m, r);
      eb.mt_idx[env] = r.idx; eb.nrm_ok[env] = r.ok; eb.nrm_saved[env] = r.saved;
      // absolute root translation goes to the double-precision base, qpos keeps the relative part (zero)
      for (int k = 0; k < NXY; k++) { eb.xy_base[(size_t)env * 2 + k] = (double)s.qpos[k]; s.qpos[k] = 0.f; }
      for (int k = NXY; k < 2; k++) eb.xy_base[(size_t)env * 2 + k] = 0.0;
    }
    if (lane < C::NV && !c.fwd_only) s.warm[lane] = 0.f;
    __syncwarp();
    if (!c.fwd_only) c.elapsed = 0;
  }
  c.x_before = (C::TASK == PB_TASK_ANT || C::TASK == PB_TASK_HUMANOID) ? c.xref : s.qpos[0];
}

// ---- epilogue: cfrc_ext, reward / termination, observation, wrapper chain, stores
template <class C>
__device__ void env_finish(EnvS<C> &s, const DevModel &m, const EnvBuffers &eb, const StepArgs &a, int lane, EnvCtx &c) {
  constexpr int NXY = (C::TASK == PB_TASK_HALFCHEETAH || C::TASK == PB_TASK_HOPPER) ? 1 : 2;
  constexpr int NSTAGE = C::INTEGRATOR == PB_INT_RK4 ? 4 : 1;
  float *obs = &s.J[0][0];  // the constraint Jacobian is dead once the last forward is done
  static_assert(sizeof(s.J) / sizeof(float) >= C::OBS, "obs staging does not fit the Jacobian storage");
  const int env = c.env;
  const bool do_reset = c.do_reset, fwd_only = c.fwd_only;
  float reward = 0.f, xref = c.xref;
  int terminated = 0, truncated = 0, elapsed = c.elapsed;
  double dxy[2] = {0.0, 0.0};
  unsigned nforward = 0;
  if (do_reset) {
    nforward = 1;
    // cfrc_ext stays zero after a reset (mj_forward does not run rnePostConstraint)
    for (int it = lane; it < C::NB * 6; it += 32) (&s.cfrc[0][0])[it] = 0.f;
    __syncwarp();
  } else {
    const float x_before = c.x_before;
    nforward = C::FRAME_SKIP * NSTAGE;
    if (C::NEED_CFRC) sim_cfrc_ext<C>(s, m, lane);
    const float dt = m.timestep * C::FRAME_SKIP;
    float x_after = C::TASK == PB_TASK_ANT ? s.xpos[1][0] : (C::TASK == PB_TASK_HUMANOID ? s.com[0] : s.qpos[0]);
    float xv = (x_after - x_before) / dt;
    float c2 = 0.f;
    for (int i = 0; i < C::NU; i++) c2 += s.ctrl[i] * s.ctrl[i];
    elapsed += 1;
    truncated = elapsed >= 1000;
    if (C::TASK == PB_TASK_HALFCHEETAH) {
      reward = xv - 0.1f * c2;
    } else if (C::TASK == PB_TASK_HOPPER) {
      bool healthy = s.qpos[1] >= 0.7f && s.qpos[2] >= -0.2f && s.qpos[2] <= 0.2f;
      for (int i = 2; i < C::NQ; i++) healthy &= s.qpos[i] >= -100.f && s.qpos[i] <= 100.f;
      for (int i = 0; i < C::NV; i++) healthy &= s.qvel[i] >= -100.f && s.qvel[i] <= 100.f;
      reward = xv + (healthy ? 1.f : 0.f) - 1e-3f * c2;
      terminated = !healthy;
    } else if (C::TASK == PB_TASK_ANT) {
      bool healthy = s.qpos[2] >= 0.2f && s.qpos[2] <= 1.0f;
      for (int i = 0; i < C::NQ; i++) healthy &= isfinite(s.qpos[i]);
      for (int i = 0; i < C::NV; i++) healthy &= isfinite(s.qvel[i]);
      float cc = 0.f;
      for (int i = 6; i < C::NB * 6; i++) { float f = clampf((&s.cfrc[0][0])[i], -1.f, 1.f); cc += f * f; }
      reward = xv + (healthy ? 1.f : 0.f) - (0.5f * c2 + 5e-4f * cc);
      terminated = !healthy;
    } else {
      reward = 1.25f * xv + 5.0f - 0.1f * c2;
      terminated = !(1.0f < s.qpos[2] && s.qpos[2] < 2.0f);
    }
  }

  // ---- re-centre the root translation; keep the stale x reference relative to the new origin
  {
    float stale = C::TASK == PB_TASK_ANT ? s.xpos[1][0] : (C::TASK == PB_TASK_HUMANOID ? s.com[0] : 0.f);
    __syncwarp();
    if (lane == 0) {
      for (int k = 0; k < NXY; k++) dxy[k] = (double)s.qpos[k];
      xref = stale - s.qpos[0];
      for (int k = 0; k < NXY; k++) s.qpos[k] = 0.f;
    }
    __syncwarp();
  }
  task_write_obs<C>(s, obs, lane);
  __syncwarp();

  // ---- wrapper chain
  float ep_return = eb.ep_return[env];
  int ep_length = eb.ep_length[env], has_info = 0;
  if (fwd_only) {}
  else if (do_reset) { ep_return = 0.f; ep_length = 0; }
  else { ep_return = __fadd_rn(ep_return, reward); ep_length += 1; has_info = terminated || truncated; }
  float *out = a.obs + (size_t)env * C::OBS;
  if (a.chain == 1 && !fwd_only) {
    float count = eb.ob_count[env], tot = __fadd_rn(count, 1.0f);
    for (int i = lane; i < C::OBS; i += 32) {
      size_t g = (size_t)env * C::OBS + i;
      float mean = eb.ob_mean[g], var = eb.ob_var[g], x = obs[i];
      welford_f32(x, mean, var, count, tot);
      eb.ob_mean[g] = mean; eb.ob_var[g] = var;
      float y = __fdiv_rn(__fsub_rn(x, mean), __fsqrt_rn(__fadd_rn(var, 1e-4f)));
      out[i] = clampf(y, -10.f, 10.f);
    }
    if (lane == 0) {
      eb.ob_count[env] = tot;
      if (!do_reset) {
        float acc = __fadd_rn(__fmul_rn(__fmul_rn(eb.rw_acc[env], a.gamma), __fsub_rn(1.0f, (float)terminated)), reward);
        float rc = eb.rw_count[env], rt = __fadd_rn(rc, 1.0f), rm = eb.rw_mean[env], rv = eb.rw_var[env];
        welford_f32(acc, rm, rv, rc, rt);
        eb.rw_acc[env] = acc; eb.rw_count[env] = rt; eb.rw_mean[env] = rm; eb.rw_var[env] = rv;
        reward = clampf(__fdiv_rn(reward, __fsqrt_rn(__fadd_rn(rv, 1e-8f))), -10.f, 10.f);
      }
    }
  } else if (a.obs) {
    for (int i = lane; i < C::OBS; i += 32) out[i] = obs[i];
  }

  // ---- store persistent state and step outputs
  for (int i = lane; i < C::NQ; i += 32) eb.qpos[(size_t)env * C::NQ + i] = s.qpos[i];
  for (int i = lane; i < C::NV; i += 32) { eb.qvel[(size_t)env * C::NV + i] = s.qvel[i]; eb.warm[(size_t)env * C::NV + i] = s.warm[i]; }
  if (lane == 0) {
    if (!do_reset) for (int k = 0; k < NXY; k++) eb.xy_base[(size_t)env * 2 + k] += dxy[k];
    eb.xref[env] = xref;
    eb.elapsed[env] = elapsed;
    if (!fwd_only) eb.autoreset[env] = (a.mode == 1) ? 0 : (do_reset ? 0 : (terminated || truncated));
    eb.ep_return[env] = ep_return; eb.ep_length[env] = ep_length;
    if (a.rew) a.rew[env] = do_reset ? 0.f : reward;
    if (a.term) a.term[env] = do_reset ? 0.f : (float)terminated;
    if (a.trunc) a.trunc[env] = do_reset ? 0.f : (float)truncated;
    if (a.ep_stat && has_info) { atomicAdd(&a.ep_stat[0], ep_return); atomicAdd(&a.ep_stat[1], (float)ep_length); atomicAdd(&a.ep_stat[2], 1.f); }
    if (a.has_info) { a.has_info[env] = has_info; a.info_r[env] = has_info ? ep_return : 0.f; a.info_l[env] = has_info ? ep_length : 0; }
    atomicAdd(&eb.counters[0], (unsigned long long)nforward);
    if (s.overflow) atomicAdd(&eb.counters[1], 1ull);
#ifdef PB_IMBALANCE
    atomicAdd(&eb.counters[24], (unsigned long long)s.dbg_t[0]); atomicAdd(&eb.counters[25], (unsigned long long)s.dbg_t[1]);
#endif
    if (s.pgs_sweeps) { atomicAdd(&eb.counters[2], (unsigned long long)s.pgs_sweeps); atomicAdd(&eb.counters[3], (unsigned long long)s.pgs_rowsteps); }
  }
  __syncwarp();
}

// Order of the next step's work queue: envs sorted by the time their warp spent in the previous step's forward evaluations
// (descending; measured with clock(), so whatever makes an env expensive - contacts, solver iterations - is captured),
// envs that are about to autoreset (one forward evaluation, no physics) last.  Lockstep CTAs take WPB consecutive queue
// entries at a time, so the warps that meet at the per-evaluation barrier carry similar contact loads and the barrier
// costs little; heavy envs go first so the tail of the launch is made of cheap ones.  Counting sort, one CTA.
#define PB_ORDER_BUCKETS 256
__global__ void __launch_bounds__(1024) env_order_kernel(EnvBuffers eb) {
  __shared__ int hist[PB_ORDER_BUCKETS];
  __shared__ int sh_lo, sh_hi;
  const int n = eb.num_envs;
  for (int i = threadIdx.x; i < PB_ORDER_BUCKETS; i += blockDim.x) hist[i] = 0;
  if (threadIdx.x == 0) { sh_lo = 0x7fffffff; sh_hi = 0; }
  __syncthreads();
  int lo = 0x7fffffff, hi = 0;
  for (int e = threadIdx.x; e < n; e += blockDim.x) if (!eb.autoreset[e]) { int c = eb.cost[e]; lo = min(lo, c); hi = max(hi, c); }
  for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(PB_FULL, lo, o)); hi = max(hi, __shfl_xor_sync(PB_FULL, hi, o)); }
  if ((threadIdx.x & 31) == 0) { atomicMin(&sh_lo, lo); atomicMax(&sh_hi, hi); }
  __syncthreads();
  lo = sh_lo; hi = sh_hi;
  const float scale = hi > lo ? (float)(PB_ORDER_BUCKETS - 2) / (float)(hi - lo) : 0.f;
  // bucket 0 = heaviest ... bucket 254 = lightest stepping env, bucket 255 = autoresets
  auto key = [&](int e) {
    if (eb.autoreset[e]) return PB_ORDER_BUCKETS - 1;
    const int k = (int)((float)(eb.cost[e] - lo) * scale);
    return (PB_ORDER_BUCKETS - 2) - max(0, min(PB_ORDER_BUCKETS - 2, k));
  };
  for (int e = threadIdx.x; e < n; e += blockDim.x) atomicAdd(&hist[key(e)], 1);
  __syncthreads();
  if (threadIdx.x < 32) {  // exclusive prefix over the buckets
    int carry = 0;
    for (int base = 0; base < PB_ORDER_BUCKETS; base += 32) {
      int v = hist[base + threadIdx.x], x = v;
      for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(PB_FULL, x, o); if ((int)threadIdx.x >= o) x += y; }
      hist[base + threadIdx.x] = carry + x - v;
      carry += __shfl_sync(PB_FULL, x, 31);
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < n; e += blockDim.x) eb.order[atomicAdd(&hist[key(e)], 1)] = e;
}

// One warp per env; the WPB warps of a CTA fetch envs from a device-side work counter (so a resetting env, which
// needs one forward evaluation instead of frame_skip x stages, frees its warp early and the last wave has no tail).
// LOCKSTEP: the warps of a CTA meet at a barrier before every forward evaluation, so they stream the (large, mostly
// straight-line) dynamics code through the instruction cache together instead of each at its own place; the warps that
// need work at a round take consecutive entries of the (optionally cost-ordered) queue with one atomic per CTA.
// CTAs per SM the shared-memory footprint allows (228 KB per SM, 1 KB reserved per CTA); the register allocation is bounded
// to match (the kernel is latency-bound: resident warps are worth more than a few spilled registers).
template <class C, int WPB, bool MSM>
constexpr int env_smem_bytes() { return ((int)sizeof(EnvS<C>) * WPB + 15) / 16 * 16 + (MSM ? (int)sizeof(DevModel) : 0); }
template <class C, int WPB, bool MSM>
constexpr int env_ctas_per_sm() {
  int n = (228 * 1024) / (env_smem_bytes<C, WPB, MSM>() + 1024);
  while (n > 1 && 65536 / (n * WPB * 32) < 80) n--;  // never squeeze below 80 registers per thread
  // the register-resident PGS (a row of A per lane) needs the full 128: below that it spills inside the sweeps
  while (n > 1 && C::SOLVER == PB_SOL_PGS && 65536 / (n * WPB * 32) < 128) n--;
  return n < 1 ? 1 : n;
}
// MSM: the CTA keeps its own copy of the model tables in shared memory (with ~216 KB of the SM configured as shared memory
// the L1 is a few KB, so table reads from global memory are mostly L2 round trips).
template <class C, int WPB, bool LOCKSTEP, int SYNCMASK = 0, bool MSM = false>
__global__ void __launch_bounds__(WPB * 32, env_ctas_per_sm<C, WPB, MSM>()) env_step_kernel(const DevModel *__restrict__ mp, EnvBuffers eb, StepArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int sh_status[2][WPB], sh_base[2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  EnvS<C> &s = reinterpret_cast<EnvS<C> *>(smem_raw)[warp];
  if constexpr (MSM) {
    static_assert(sizeof(DevModel) % 4 == 0, "DevModel is copied word by word");
    unsigned *dst = reinterpret_cast<unsigned *>(smem_raw + env_smem_bytes<C, WPB, false>());
    const unsigned *src = reinterpret_cast<const unsigned *>(mp);
    for (int i = threadIdx.x; i < (int)(sizeof(DevModel) / 4); i += WPB * 32) dst[i] = src[i];
    __syncthreads();
  }
  const DevModel &m = MSM ? *reinterpret_cast<const DevModel *>(smem_raw + env_smem_bytes<C, WPB, false>()) : *mp;
  constexpr int NSTAGE = C::INTEGRATOR == PB_INT_RK4 ? 4 : 1;
  if (blockIdx.x == 0 && threadIdx.x == 0) eb.work[a.parity ^ 1] = 0;  // the next launch's counter
  EnvCtx c;
  bool have = false, exhausted = false;
  int ev = 0, nev = 0;
  unsigned cost = 0;
  // SYNCMASK's upper bits: barrier period in forward evaluations (0/1: every evaluation).  With a period P > 1 the warps
  // of a CTA may drift apart by up to P evaluations' worth of load difference between two barriers.
  constexpr int PERIOD = (SYNCMASK >> 8) > 1 ? (SYNCMASK >> 8) : 1;
  for (int round = 0;; round++) {
    if constexpr (LOCKSTEP) {
      if (round % PERIOD == 0) {
        const int par = (round / PERIOD) & 1;
        const bool need = !have && !exhausted;
        if (lane == 0) sh_status[par][warp] = need ? 1 : (have ? 2 : 0);
        __syncthreads();
        int n = 0, rank = 0, anyhave = 0;
#pragma unroll
        for (int w = 0; w < WPB; w++) {
          const int st = sh_status[par][w];
          n += st == 1; rank += (st == 1 && w < warp); anyhave |= st == 2;
        }
        if (n > 0) {
          if (threadIdx.x == 0) sh_base[par] = atomicAdd(eb.work + a.parity, n);
          __syncthreads();
          const int base = sh_base[par];
          anyhave |= base < eb.num_envs;
          if (need) {
            const int pos = base + rank;
            if (pos < eb.num_envs) {
              env_begin<C>(s, m, eb, a, a.ordered ? eb.order[pos] : pos, lane, c);
              have = true; ev = 0; cost = 0; nev = c.do_reset ? 1 : C::FRAME_SKIP * NSTAGE;
            } else exhausted = true;
          }
        }
        if (!anyhave) break;
      }
      if (!have) { if constexpr ((SYNCMASK & 0xff) != 0) sim_forward_idle<SYNCMASK & 0xff>(); continue; }
    } else {
      if (!have) {
        int e = 0;
        if (lane == 0) e = atomicAdd(eb.work + a.parity, 1);
        e = __shfl_sync(PB_FULL, e, 0);
        if (e >= eb.num_envs) break;
        env_begin<C>(s, m, eb, a, e, lane, c);
        have = true; ev = 0; cost = 0; nev = c.do_reset ? 1 : C::FRAME_SKIP * NSTAGE;
      }
    }
    const unsigned t_begin = (unsigned)clock();
    // frame_skip x mj_step as a flat sequence of forward evaluations; a resetting env runs only one
#ifdef PB_IMBALANCE
    unsigned stage_t[6];
    sim_forward<C, LOCKSTEP ? (SYNCMASK & 0xff) : 0>(s, m, lane, stage_t);
#else
    sim_forward<C, LOCKSTEP ? (SYNCMASK & 0xff) : 0>(s, m, lane);
#endif
    if (!c.do_reset) {
      if constexpr (C::INTEGRATOR == PB_INT_RK4) sim_rk4_post<C>(s, m, ev % NSTAGE, lane); else sim_euler<C>(s, m, lane);
    }
    const unsigned t_eval = (unsigned)clock() - t_begin;
    cost += t_eval >> 6;  // this warp's time in the evaluation, in 64-cycle units
#ifdef PB_IMBALANCE   // debug build: per stage, sum over CTA rounds of the slowest warp's and of all warps' cycles -> counters[8..]
    if constexpr (LOCKSTEP) {
      __shared__ unsigned sh_te[8][WPB];
      if (lane == 0) { for (int k = 0; k < 6; k++) sh_te[k][warp] = stage_t[k]; sh_te[6][warp] = t_eval - (stage_t[0] + stage_t[1] + stage_t[2] + stage_t[3] + stage_t[4] + stage_t[5]); sh_te[7][warp] = t_eval; }
      __syncthreads();
      if (threadIdx.x < 8) {
        unsigned mx = 0, sm = 0;
        for (int w = 0; w < WPB; w++) { mx = max(mx, sh_te[threadIdx.x][w]); sm += sh_te[threadIdx.x][w]; }
        atomicAdd(&eb.counters[8 + 2 * threadIdx.x], (unsigned long long)mx); atomicAdd(&eb.counters[9 + 2 * threadIdx.x], (unsigned long long)sm);
        if (threadIdx.x == 0) atomicAdd(&eb.counters[6], 1ull);
      }
      __syncthreads();
    }
#endif
    if (++ev == nev) {
      env_finish<C>(s, m, eb, a, lane, c);
      if (lane == 0 && !c.fwd_only) eb.cost[c.env] = c.do_reset ? 0 : (int)cost;
      have = false;
    }
  }
}
