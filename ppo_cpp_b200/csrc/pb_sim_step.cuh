// pb_sim_step.cuh — mj_forward / mj_step (Euler with implicit joint damping, RK4) and the cfrc_ext part of
// mj_rnePostConstraint for one env per warp (reference call site mujoco_env.h:102-108).
#pragma once
#include "pb_sim_vel.cuh"

// SYNCMASK: CTA-wide barriers between the stages (bit k = barrier after stage k) used by the lockstep kernel to keep
// the warps of a CTA on the same stretch of code; a warp that skips the evaluation calls sim_forward_idle instead.
template <int SYNCMASK, int BIT>
__device__ __forceinline__ void stage_sync() { if constexpr ((SYNCMASK >> BIT) & 1) __syncthreads(); }
template <int SYNCMASK>
__device__ __forceinline__ void sim_forward_idle() {
#pragma unroll
  for (int k = 0; k < 8; k++) if ((SYNCMASK >> k) & 1) __syncthreads();
}

#ifdef PB_IMBALANCE
#define PB_STAGE_T(k) do { unsigned t_ = (unsigned)clock(); if (stage_t) stage_t[k] = t_ - t_prev; t_prev = t_; } while (0)
#else
#define PB_STAGE_T(k) do { } while (0)
#endif
template <class C, int SYNCMASK = 0>
__device__ void sim_forward(EnvS<C> &s, const DevModel &m, int lane, unsigned *stage_t = nullptr) {
#ifdef PB_IMBALANCE
  unsigned t_prev = (unsigned)clock();
#endif
  sim_kinematics<C>(s, m, lane);
  sim_com_pos<C>(s, m, lane);
  PB_STAGE_T(0);
  stage_sync<SYNCMASK, 0>();
  sim_collision<C>(s, m, lane);   // before crb_factor: the geom frames share storage with its scratch
  PB_STAGE_T(1);
  stage_sync<SYNCMASK, 1>();
  sim_crb_factor<C>(s, m, lane);
  PB_STAGE_T(2);
  stage_sync<SYNCMASK, 2>();
  sim_make_constraint<C>(s, m, lane);
  PB_STAGE_T(3);
  stage_sync<SYNCMASK, 3>();
  sim_velocity_smooth<C>(s, m, lane);
  PB_STAGE_T(4);
  stage_sync<SYNCMASK, 4>();
  if (s.nefc == 0) {
    if (lane < C::NV) { s.qacc[lane] = s.qacc_smooth[lane]; if (!EnvS<C>::PGS) s.qfrc_constraint[lane] = 0.f; }
  } else {
    // the PGS path forms aref / b on the row it keeps in registers (sim_solve_pgs)
    if constexpr (C::SOLVER == PB_SOL_PGS) sim_solve_pgs<C>(s, m, lane);
    else { sim_reference<C>(s, lane); sim_solve_newton<C>(s, m, lane); }
  }
  __syncwarp();
  if (lane < C::NV) s.warm[lane] = s.qacc[lane];
  __syncwarp();
  PB_STAGE_T(5);
}

// qpos <- qpos (+) h*vel, quaternion-aware for the free joint (mj_integratePos); lane = joint
template <class C>
__device__ void sim_integrate_pos(EnvS<C> &s, const DevModel &m, const float *vel, float h, int lane) {
  if (lane < C::NJ) {
    int qa = m.jnt_qposadr[lane], da = m.jnt_dofadr[lane];
    if (m.jnt_type[lane] == PB_JNT_FREE) {
      for (int k = 0; k < 3; k++) s.qpos[qa + k] += h * vel[da + k];
      float w[3] = {vel[da + 3], vel[da + 4], vel[da + 5]};
      float ang = h * normalize3(w), sn, cs, q[4], ql[4], qn[4];
      sincosf(0.5f * ang, &sn, &cs);
      ql[0] = cs; ql[1] = w[0] * sn; ql[2] = w[1] * sn; ql[3] = w[2] * sn;
      for (int k = 0; k < 4; k++) q[k] = s.qpos[qa + 3 + k];
      quat_normalize(q);
      quat_mul(qn, q, ql);
      quat_normalize(qn);
      for (int k = 0; k < 4; k++) s.qpos[qa + 3 + k] = qn[k];
    } else {
      s.qpos[qa] += h * vel[da];
    }
  }
  __syncwarp();
}

// semi-implicit Euler; joint damping handled implicitly: (M + h B) qacc = qfrc_smooth + qfrc_constraint
template <class C>
__device__ void sim_euler(EnvS<C> &s, const DevModel &m, int lane) {
  constexpr int NV = C::NV, LDH = NV + 1;
  float h = m.timestep;
  bool damped = false;
  for (int i = 0; i < NV; i++) damped |= m.dof_damping[i] > 0.f;
  float *x = s.vecA;
  if (damped) {
    static_assert(C::INTEGRATOR != PB_INT_EULER || C::SOLVER == PB_SOL_NEWTON, "Euler path reuses the Newton scratch");
    float *Md = s.solv, *H = s.solv + NV * LDH;
    if (s.nefc == 0) {  // the Newton solver did not run: build the dense M copy here
      for (int e = lane; e < NV * LDH; e += 32) Md[e] = 0.f;
      __syncwarp();
      for (int e = lane; e < C::NM; e += 32) { int i = m.M_i[e], j = m.M_j[e]; Md[i * LDH + j] = s.qM[e]; Md[j * LDH + i] = s.qM[e]; }
      __syncwarp();
    }
    for (int e = lane; e < NV * NV; e += 32) {
      int i = e / NV, j = e - i * NV;
      if (j <= i) H[i * LDH + j] = Md[i * LDH + j] + (i == j ? h * m.dof_damping[i] : 0.f);
    }
    if (lane < NV) x[lane] = s.qfrc_smooth[lane] + s.qfrc_constraint[lane];
    __syncwarp();
    for (int j = 0; j < NV; j++) {
      float d = sqrtf(fmaxf(H[j * LDH + j], PB_MINVAL));
      __syncwarp();
      if (lane > j && lane < NV) H[lane * LDH + j] /= d;
      if (lane == j) H[j * LDH + j] = d;
      __syncwarp();
      for (int e = lane; e < (NV - j - 1) * (NV - j - 1); e += 32) {
        int a = j + 1 + e / (NV - j - 1), b = j + 1 + e % (NV - j - 1);
        if (b <= a) H[a * LDH + b] -= H[a * LDH + j] * H[b * LDH + j];
      }
      __syncwarp();
    }
    for (int k = 0; k < NV; k++) {
      if (lane == k) x[k] /= H[k * LDH + k];
      __syncwarp();
      if (lane > k && lane < NV) x[lane] -= H[lane * LDH + k] * x[k];
      __syncwarp();
    }
    for (int k = NV - 1; k >= 0; k--) {
      if (lane == k) x[k] /= H[k * LDH + k];
      __syncwarp();
      if (lane < k) x[lane] -= H[k * LDH + lane] * x[k];
      __syncwarp();
    }
  } else {
    if (lane < NV) x[lane] = s.qacc[lane];
    __syncwarp();
  }
  if (lane < NV) s.qvel[lane] += h * x[lane];
  __syncwarp();
  sim_integrate_pos<C>(s, m, s.qvel, h, lane);
}

// classic RK4 written as four (forward, post-stage update) pairs: stage i evaluates the dynamics at X[i], the post
// step accumulates the weighted sums and moves the state to X[i+1] (stages 1-3 each run a full forward, collision and
// constraint solve included).  Splitting it this way lets the env kernel keep the warps of a CTA in lockstep.
template <class C>
__device__ void sim_rk4_post(EnvS<C> &s, const DevModel &m, int i, int lane) {
  const float h = m.timestep;
  const float A[3] = {0.5f, 0.5f, 1.f}, B[4] = {1.f / 6.f, 1.f / 3.f, 1.f / 3.f, 1.f / 6.f};
  if (i == 0) {
    if (lane < C::NQ) s.q0[lane] = s.qpos[lane];
    if (lane < C::NV) {
      s.v0[lane] = s.qvel[lane];
      s.accv[lane] = B[0] * s.qvel[lane];
      s.acca[lane] = B[0] * s.qacc[lane];
    }
  } else {
    if (lane < C::NV) s.acca[lane] += B[i] * s.qacc[lane];
  }
  __syncwarp();
  if (i < 3) {
    // X[i+1] = X[0] (+) h*A*(stage i velocity, acceleration)
    if (lane < C::NV) s.vecA[lane] = A[i] * s.qvel[lane];  // qvel still holds the stage-i velocity
    if (lane < C::NQ) s.qpos[lane] = s.q0[lane];
    __syncwarp();
    sim_integrate_pos<C>(s, m, s.vecA, h, lane);
    if (lane < C::NV) {
      float v = s.v0[lane] + h * A[i] * s.qacc[lane];
      s.qvel[lane] = v;
      s.accv[lane] += B[i + 1] * v;
    }
    __syncwarp();
  } else {
    if (lane < C::NQ) s.qpos[lane] = s.q0[lane];
    if (lane < C::NV) s.qvel[lane] = s.v0[lane] + h * s.acca[lane];
    __syncwarp();
    sim_integrate_pos<C>(s, m, s.accv, h, lane);
  }
}

// cfrc_ext from the contact forces of the last forward evaluation (stale frames, as the library leaves them)
template <class C>
__device__ void sim_cfrc_ext(EnvS<C> &s, const DevModel &m, int lane) {
  for (int it = lane; it < C::NB * 6; it += 32) (&s.cfrc[0][0])[it] = 0.f;
  __syncwarp();
  // lane = body: gather over contacts (no atomics, fixed summation order)
  if (lane >= 1 && lane < C::NB) {
    float acc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < s.ncon; c++) {
      const Contact &con = s.con[c];
      if (con.efc < 0) continue;
      int p = con.pair, b1 = m.geom_body[m.pair_g1[p]], b2 = m.geom_body[m.pair_g2[p]];
      if (b1 != lane && b2 != lane) continue;
      const float *f = s.efc_force + con.efc;
      float l0, l1 = 0.f, l2 = 0.f, mu = m.pair_mu[p];
      if (m.pair_condim[p] == 1) l0 = f[0];
      else { l0 = f[0] + f[1] + f[2] + f[3]; l1 = (f[0] - f[1]) * mu; l2 = (f[2] - f[3]) * mu; }
      float t2[3], wf[3], off[3], tq[3];
      cross3(t2, con.n, con.t1);
      for (int k = 0; k < 3; k++) { wf[k] = con.n[k] * l0 + con.t1[k] * l1 + t2[k] * l2; off[k] = con.pos[k] - s.com[k]; }
      cross3(tq, off, wf);
      float sg = (b2 == lane ? 1.f : 0.f) - (b1 == lane ? 1.f : 0.f);
      for (int k = 0; k < 3; k++) { acc[k] += sg * tq[k]; acc[3 + k] += sg * wf[k]; }
    }
    for (int k = 0; k < 6; k++) s.cfrc[lane][k] = acc[k];
  }
  __syncwarp();
}
