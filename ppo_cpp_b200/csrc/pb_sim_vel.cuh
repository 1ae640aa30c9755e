// pb_sim_vel.cuh — velocity stage, smooth acceleration and the two constraint solvers for one env per warp.
// Replaces the velocity/acceleration/constraint half of mj_forward.
#pragma once
#include "pb_sim_pos.cuh"

// Sum of v over the dof chain of every lane (itself and its ancestors) by pointer jumping over dof_parent: `rounds` shuffle
// rounds instead of a walk down the tree levels.
__device__ __forceinline__ void chain_sum6(float *v, int ptr, int rounds, int lane) {
  for (int r = 0; r < rounds; r++) {
    const int src = ptr < 0 ? lane : ptr;
    float t[6];
#pragma unroll
    for (int c = 0; c < 6; c++) t[c] = __shfl_sync(PB_FULL, v[c], src);
    const int nptr = __shfl_sync(PB_FULL, ptr, src);
    if (ptr >= 0) {
#pragma unroll
      for (int c = 0; c < 6; c++) v[c] += t[c];
      ptr = nptr;
    }
  }
}

// cvel, cdof_dot, RNE bias (gravity + Coriolis), passive and actuator forces -> qfrc_smooth.
// mj_comVel / mj_rne accumulate cvel and cacc joint by joint down the tree; both are sums over the dof chain, so dof lanes
// form them with chain_sum6 (registers + shuffles) and body lanes pick the value at their last dof.
template <class C>
__device__ void sim_velocity_smooth(EnvS<C> &s, const DevModel &m, int lane) {
  float cd[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, vel[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, qv = 0.f;
  int ptr = -1, vs = -1;
  bool cdd_zero = true;
  if (lane < C::NV) {
    qv = s.qvel[lane];
#pragma unroll
    for (int c = 0; c < 6; c++) { cd[c] = s.cdof[lane][c]; vel[c] = cd[c] * qv; }
    ptr = m.dof_parent[lane]; vs = m.dof_velsrc[lane]; cdd_zero = m.dof_cdd_zero[lane] != 0;
  }
  chain_sum6(vel, ptr, m.dof_rounds, lane);
  float cv[6], acc[6];
#pragma unroll
  for (int c = 0; c < 6; c++) { float t = __shfl_sync(PB_FULL, vel[c], vs < 0 ? lane : vs); cv[c] = vs < 0 ? 0.f : t; }
  cross_motion(acc, cv, cd);  // cdof_dot
#pragma unroll
  for (int c = 0; c < 6; c++) acc[c] = cdd_zero ? 0.f : acc[c] * qv;
  chain_sum6(acc, ptr, m.dof_rounds, lane);
  {
    const int ld = lane < C::NB ? m.body_lastdof[lane] : -1;
    float bv[6], ba[6];
#pragma unroll
    for (int c = 0; c < 6; c++) {
      float t0 = __shfl_sync(PB_FULL, vel[c], ld < 0 ? lane : ld), t1 = __shfl_sync(PB_FULL, acc[c], ld < 0 ? lane : ld);
      bv[c] = ld < 0 ? 0.f : t0; ba[c] = ld < 0 ? 0.f : t1;
    }
    ba[3] -= m.gravity[0]; ba[4] -= m.gravity[1]; ba[5] -= m.gravity[2];
    if (lane < C::NB) {
      float f[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (lane > 0) {
        float t[6], t1[6];
        mul_inert_vec(f, s.cinert[lane], ba);
        mul_inert_vec(t, s.cinert[lane], bv);
        cross_force(t1, bv, t);
#pragma unroll
        for (int c = 0; c < 6; c++) f[c] += t1[c];
      }
#pragma unroll
      for (int c = 0; c < 6; c++) { s.cvel[lane][c] = bv[c]; s.cfrc[lane][c] = f[c]; }
    }
  }
  __syncwarp();
  // subtree force sums, then project on the dof axes
  constexpr int NT6 = (C::NB * 6 + 31) / 32;
  float sub[NT6];
#pragma unroll
  for (int t = 0; t < NT6; t++) {
    int it = lane + 32 * t;
    float acc = 0.f;
    if (it < C::NB * 6) {
      int b = it / 6, c = it - 6 * b, n = m.body_subtree[b];
      if (b > 0) for (int k = 0; k < n; k++) acc += s.cfrc[b + k][c];
    }
    sub[t] = acc;
  }
  __syncwarp();
#pragma unroll
  for (int t = 0; t < NT6; t++) { int it = lane + 32 * t; if (it < C::NB * 6) (&s.cfrc[0][0])[it] = sub[t]; }
  __syncwarp();
  if (lane < C::NV) {
    int i = lane;
    const float *a = s.cdof[i], *f = s.cfrc[m.dof_body[i]];
    float bias = a[0] * f[0] + a[1] * f[1] + a[2] * f[2] + a[3] * f[3] + a[4] * f[4] + a[5] * f[5];
    float passive = -m.dof_damping[i] * s.qvel[i];
    float act = 0.f;
    int u = m.dof_act[i];
    if (u >= 0) act = m.dof_gear[i] * clampf(s.ctrl[u], m.dof_ctrl_lo[i], m.dof_ctrl_hi[i]);
    s.qfrc_actuator[i] = act;
    s.qfrc_smooth[i] = passive + (act - bias);  // spring term added below by the joint lanes
  }
  __syncwarp();
  if (lane < C::NJ && m.jnt_type[lane] != PB_JNT_FREE && m.jnt_stiffness[lane] != 0.f)
    s.qfrc_smooth[m.jnt_dofadr[lane]] -= m.jnt_stiffness[lane] * (s.qpos[m.jnt_qposadr[lane]] - m.jnt_springref[lane]);
  __syncwarp();
  if (lane < C::NV) s.qacc_smooth[lane] = s.qfrc_smooth[lane];
  __syncwarp();
  sim_solve_LD<C>(s, m, s.qacc_smooth, lane);
}

// efc_vel, aref, b = J qacc_smooth - aref (mj_referenceConstraint + start of mj_fwdConstraint); lane = row
template <class C>
__device__ void sim_reference(EnvS<C> &s, int lane) {
  int nefc = s.nefc;
  for (int r = lane; r < nefc; r += 32) {
    float vel = 0.f, ja = 0.f;
    for (int i = 0; i < C::NV; i++) { float j = s.J[r][i]; vel += j * s.qvel[i]; ja += j * s.qacc_smooth[i]; }
    float aref = -s.efc_kip[r] - s.efc_bd[r] * vel;
    s.efc_aref[r] = aref;
    s.efc_b[r] = ja - aref;
  }
  __syncwarp();
}

// x <- L^-T x for one vector held in registers, fully unrolled over the compile-time dof tree: dof K pushes x[K] to each
// of its ancestors J with the factor entry LD[madr(K) + 1 + a] (broadcast shared-memory reads).
template <class C, int K, int J, int E>
__device__ __forceinline__ void lt_regs_anc(float (&x)[C::NV], const float *__restrict__ qLD) {
  if constexpr (J >= 0) {
    x[J] -= qLD[E] * x[K];
    lt_regs_anc<C, K, C::DOF_PARENT[J], E + 1>(x, qLD);
  }
}
template <class C, int K>
__device__ __forceinline__ void lt_regs(float (&x)[C::NV], const float *__restrict__ qLD) {
  if constexpr (K > 0) {
    lt_regs_anc<C, K, C::DOF_PARENT[K], Tree<C>::madr(K) + 1>(x, qLD);
    lt_regs<C, K - 1>(x, qLD);
  }
}

// ------------------------------------------------------------------ PGS (dual), Humanoid
// At most 32 rows, row r in lane r.  A Gauss-Seidel sweep is an unrolled sequence of row steps left through a warp-uniform
// exit after step nefc-1 (an entry jump into the sequence - Duff's device with the rows in the last lanes - was measured
// first: its compare tree + indirect branch cost ~80 cycles at the head of every sweep).  Everything a sweep touches sits in the owning lane's registers: its force f, its unconstrained minimiser
// g = f - res / A_rr (res = (A f + b)_r) and ROW r of -A / A_rr (ah[0..31], compile-time indices, zero on the diagonal: a
// row's own update does not move its own minimiser).  One row step i is
//   FMNMX, FADD   every lane forms the update of its own row, dl = max(0, g) - f (only lane i's is used)
//   SHFL          broadcast of lane i's step
//   FFMA          every lane moves its minimiser by ah[i] * delta
//   ISETP, FSEL   the owning lane records the minimiser its step used; the new force max(0, gs) and the cost decrease
//                 -dl (0.5 A_ii dl + res) = dl A_ii (gs - f - 0.5 dl) are formed once per sweep, all rows in parallel
// against 21 instructions per row step for the shared-memory version of round 1 (packed-triangle index arithmetic, a load
// and a warp-uniform branch per row step), which was 28 % of the kernel's instructions at the walking state
// (profiles/r02_envstep_walk_v0_lines.txt).  Same iterates as mj_solPGS on scalar rows (limits, pyramid edges):
// f_i <- max(0, f_i - res_i / A_ii) in row order, cost decrease summed per sweep, stop when decrease * scale < tolerance.
// The set-up (mj_referenceConstraint, warm start, Y = L^-T J^T, A) runs on the row held in registers as well.
template <class C, int I>
__device__ __forceinline__ void pgs_row_step(float f, float &g, float &gs, const float (&ah)[32], int lane) {
  const float dl = fmaxf(0.f, g) - f;
  const float delta = __shfl_sync(PB_FULL, dl, I);
  if (lane == I) gs = g;   // the owner keeps the minimiser its step used (its own step does not move it: ah[I] = 0 there)
  g += ah[I] * delta;
}
template <class C>
__device__ void sim_solve_pgs(EnvS<C> &s, const DevModel &m, int lane) {
  static_assert(C::MAXEFC <= 32, "lane-per-row PGS needs MAXEFC <= 32");
  constexpr int NV = C::NV;
  // row count through a warp vote: warp-uniform by construction
  const int nefc = __popc(__ballot_sync(PB_FULL, lane < s.nefc));
#ifdef PB_IMBALANCE
  const unsigned dbg_t0 = (unsigned)clock();
#endif
  float *AR = s.solv;   // packed lower triangle, element (r, c), r >= c, at r(r+1)/2 + c
  const int row = lane;
  const bool mine = lane < nefc;
  const int tri = mine ? row * (row + 1) / 2 : 0;
  float f = 0.f, b = 0.f, R = 1.f;
  {
    float x[NV];
#pragma unroll
    for (int i = 0; i < NV; i++) x[i] = 0.f;
    if (mine) {
#pragma unroll
      for (int i = 0; i < NV; i++) x[i] = s.J[row][i];
      // efc_vel, aref, b = J qacc_smooth - aref and the warm-start residual: three dot products over the row
      float vel = 0.f, ja = 0.f, jw = 0.f;
#pragma unroll
      for (int i = 0; i < NV; i++) { vel += x[i] * s.qvel[i]; ja += x[i] * s.qacc_smooth[i]; jw += x[i] * s.warm[i]; }
      const float aref = -s.efc_kip[row] - s.efc_bd[row] * vel;
      b = ja - aref;
      R = s.efc_R[row];
      const float jar = jw - aref;
      f = jar < 0.f ? -jar / R : 0.f;   // warm-start force from qacc_warmstart
      // Y = L^-T J^T in place (rows are lane-private, L entries are broadcast reads)
      lt_regs<C, NV - 1>(x, s.qLD);
#pragma unroll
      for (int i = 0; i < NV; i++) { s.J[row][i] = x[i]; x[i] *= s.qLDinv[i]; }
    }
    __syncwarp();
    // A = Y D^-1 Y^T + diag(R): the lane of row c forms element (r, c) of row r from its own (scaled) row and a broadcast of row r
    for (int r = 0; r < nefc; r++) {
      if (mine && row <= r) {
        float acc = row == r ? R : 0.f;
#pragma unroll
        for (int i = 0; i < NV; i++) acc += x[i] * s.J[r][i];
        AR[r * (r + 1) / 2 + row] = acc;
      }
    }
  }
  __syncwarp();
  const float aii = mine ? AR[tri + row] : 1.f, ainv = mine ? 1.f / aii : 0.f;
  // ah[c] = -A[row][c] / A_rr for every row c, 0 on the diagonal and outside the rows
  float ah[32];
#pragma unroll
  for (int c = 0; c < 32; c++) {
    ah[c] = 0.f;
    if (mine && c < nefc && c != row) ah[c] = -ainv * AR[c <= row ? tri + c : c * (c + 1) / 2 + row];
  }
  // res = A f + b for the warm start; the dual cost decides between the warm start and zero.  In terms of g:
  // g = f - res / A_rr = -b / A_rr + sum_{c != r} ah[c] f_c
  float g = -b * ainv;
  {
    float gw = g;
#pragma unroll
    for (int p = 0; p < 32; p++) gw += ah[p] * __shfl_sync(PB_FULL, f, p);
    const float res = aii * (f - gw);
    const float cost = warp_sum(f * (0.5f * (res - b) + b));
    if (cost > 0.f) f = 0.f; else g = gw;
  }
#ifdef PB_IMBALANCE
  const unsigned dbg_t1 = (unsigned)clock();
#endif
  int sweeps = 0;
  const unsigned tol_q = (unsigned)(m.tolerance * 0x1p36f);
  for (int iter = 0; iter < m.iterations; iter++) {
    float gs = 0.f;
    sweeps++;
    do {   // row steps 0 .. nefc-1: straight-line code with a warp-uniform exit after every step (nefc >= 1 here)
#define PB_PGS_STEP(I) pgs_row_step<C, I>(f, g, gs, ah, lane); if (nefc == I + 1) break;
      PB_PGS_STEP(0) PB_PGS_STEP(1) PB_PGS_STEP(2) PB_PGS_STEP(3) PB_PGS_STEP(4) PB_PGS_STEP(5) PB_PGS_STEP(6) PB_PGS_STEP(7)
      PB_PGS_STEP(8) PB_PGS_STEP(9) PB_PGS_STEP(10) PB_PGS_STEP(11) PB_PGS_STEP(12) PB_PGS_STEP(13) PB_PGS_STEP(14) PB_PGS_STEP(15)
      PB_PGS_STEP(16) PB_PGS_STEP(17) PB_PGS_STEP(18) PB_PGS_STEP(19) PB_PGS_STEP(20) PB_PGS_STEP(21) PB_PGS_STEP(22) PB_PGS_STEP(23)
      PB_PGS_STEP(24) PB_PGS_STEP(25) PB_PGS_STEP(26) PB_PGS_STEP(27) PB_PGS_STEP(28) PB_PGS_STEP(29) PB_PGS_STEP(30) PB_PGS_STEP(31)
#undef PB_PGS_STEP
    } while (0);
    // end of the sweep, all rows at once: new force and the row's cost decrease
    const float fn = fmaxf(0.f, gs), dl = fn - f;
    float imp = dl * aii * (gs - f - 0.5f * dl);
    f = fn;
    // sum of the rows' cost decreases against the tolerance, as ONE integer warp reduction (redux.sync) instead of five
    // shuffle + add rounds on the sweep's critical path: each lane's scaled decrease in fixed point, 2^-36 resolution (the
    // tolerance 1e-8 is 687 units; 32 lanes round by at most 32), clamped to 2^-10 per lane so the sum cannot wrap - a lane at
    // the clamp is five orders of magnitude above the tolerance, so the decision is unchanged
    const unsigned q = (unsigned)(fminf(fmaxf(imp * m.solver_scale, 0.f), 0x1p-10f) * 0x1p36f);
    if (__reduce_add_sync(PB_FULL, q) < tol_q) break;
  }
#ifdef PB_IMBALANCE
  const unsigned dbg_t2 = (unsigned)clock();
  if (lane == 0) { s.dbg_t[0] += dbg_t1 - dbg_t0; s.dbg_t[1] += dbg_t2 - dbg_t1; }
#endif
  if (mine) s.efc_force[row] = f;
  if (lane == 0) { s.pgs_sweeps += sweeps; s.pgs_rowsteps += sweeps * nefc; }
  __syncwarp();
  // qacc = qacc_smooth + L^-1 D^-1 (Y^T f); qfrc_constraint is not needed downstream for RK4 models
  {
    const int myend = lane < C::NV ? m.dof_Madr[lane] + m.dof_depth[lane] : 0;
    float xl = 0.f;
    if (lane < C::NV) {
      float acc = 0.f;
      for (int r = 0; r < nefc; r++) acc += s.J[r][lane] * s.efc_force[r];
      xl = acc * s.qLDinv[lane];
    }
    solve_down<C, 0>(xl, s.qLD, lane, myend);
    if (lane < C::NV) s.qacc[lane] = s.qacc_smooth[lane] + xl;
  }
  __syncwarp();
}

// ------------------------------------------------------------------ Newton (primal), HalfCheetah / Hopper / Ant
// cost(a) = 1/2 (a-a0)^T M (a-a0) + sum_r 1/2 D_r min(0, J_r a - aref_r)^2 ; exact line search on the convex
// piecewise-quadratic by safeguarded Newton iterations on its derivative.
template <class C>
__device__ void sim_solve_newton(EnvS<C> &s, const DevModel &m, int lane) {
  constexpr int NV = C::NV, LDH = NV + 1;
  const int nefc = s.nefc;
  float *Md = s.solv, *H = s.solv + NV * LDH;
  float *qacc = s.qacc, *Ma = s.vecA, *grad = s.vecB, *srch = s.vecC, *Mv = s.vecD;
  // dense copy of M (NV <= 14 for the Newton models)
  for (int e = lane; e < NV * LDH; e += 32) Md[e] = 0.f;
  __syncwarp();
  for (int e = lane; e < C::NM; e += 32) { int i = m.M_i[e], j = m.M_j[e]; Md[i * LDH + j] = s.qM[e]; Md[j * LDH + i] = s.qM[e]; }
  __syncwarp();
  // warm start: best of qacc_warmstart and qacc_smooth
  float cost_ws, cost_sm;
  {
    float part = 0.f;
    for (int r = lane; r < nefc; r += 32) {
      float jar = -s.efc_aref[r];
      for (int i = 0; i < NV; i++) jar += s.J[r][i] * s.warm[i];
      if (jar < 0.f) part += 0.5f * jar * jar / s.efc_R[r];
    }
    if (lane < NV) {
      float ma = 0.f;
      for (int j = 0; j < NV; j++) ma += Md[lane * LDH + j] * s.warm[j];
      part += 0.5f * (ma - s.qfrc_smooth[lane]) * (s.warm[lane] - s.qacc_smooth[lane]);
    }
    cost_ws = warp_sum(part);
    part = 0.f;
    for (int r = lane; r < nefc; r += 32) { float b = s.efc_b[r]; if (b < 0.f) part += 0.5f * b * b / s.efc_R[r]; }
    cost_sm = warp_sum(part);
  }
  if (lane < NV) qacc[lane] = (cost_ws > cost_sm) ? s.qacc_smooth[lane] : s.warm[lane];
  __syncwarp();
  float cost = 0.f;
  const int maxiter = min(m.iterations, 20);  // fp32: the improvement test fires long before this
  for (int iter = 0;; iter++) {
    // Ma, jar, force, cost, gradient
    if (lane < NV) {
      float ma = 0.f;
      for (int j = 0; j < NV; j++) ma += Md[lane * LDH + j] * qacc[j];
      Ma[lane] = ma;
    }
    float part = 0.f;
    for (int r = lane; r < nefc; r += 32) {
      float jar = -s.efc_aref[r];
      for (int i = 0; i < NV; i++) jar += s.J[r][i] * qacc[i];
      s.efc_jar[r] = jar;
      float f = jar < 0.f ? -jar / s.efc_R[r] : 0.f;
      s.efc_force[r] = f;
      part += 0.5f * f * -jar;
    }
    __syncwarp();
    float gpart = 0.f;
    if (lane < NV) {
      float g = Ma[lane] - s.qfrc_smooth[lane];
      part += 0.5f * g * (qacc[lane] - s.qacc_smooth[lane]);
      float fc = 0.f;
      for (int r = 0; r < nefc; r++) fc += s.J[r][lane] * s.efc_force[r];
      s.qfrc_constraint[lane] = fc;
      g -= fc;
      grad[lane] = g;
      gpart = g * g;
    }
    float newcost = warp_sum(part), gnorm = sqrtf(warp_sum(gpart));
    bool done = iter > 0 && (m.solver_scale * (cost - newcost) < m.tolerance || m.solver_scale * gnorm < m.tolerance);
    cost = newcost;
    if (done || iter >= maxiter) break;
    // Hessian H = M + J^T diag(D active) J, Cholesky H = L L^T and search = -H^-1 grad, all in registers: lane i keeps
    // row i of H / L (h[j], j <= i is the part that matters) and, as it goes by, row i of L^T (ht[b] = L[b][i]); column
    // steps and substitutions exchange values with shuffles over compile-time register indices — no shared-memory
    // traffic, no warp barriers, no index arithmetic (the shared-memory version spent ~40 % of the Ant kernel here).
    for (int r = lane; r < nefc; r += 32) s.efc_jv[r] = s.efc_jar[r] < 0.f ? 1.f / s.efc_R[r] : 0.f;  // D, until jv is formed
    __syncwarp();
    const int li = lane < NV ? lane : NV - 1;   // lanes >= NV shadow the last row; nothing reads them
    float h[NV], ht[NV], hinv = 1.f;
#pragma unroll
    for (int j = 0; j < NV; j++) { h[j] = Md[li * LDH + j]; ht[j] = 0.f; }
    for (int r = 0; r < nefc; r++) {
      const float d = s.efc_jv[r];
      if (d != 0.f) {   // warp-uniform
        const float w = s.J[r][li] * d;
#pragma unroll
        for (int j = 0; j < NV; j++) h[j] += w * s.J[r][j];
      }
    }
#pragma unroll
    for (int j = 0; j < NV; j++) {  // right-looking Cholesky, column j
      // pivot: 1 / sqrt by MUFU.RSQ (2 ulp; the direction only has to be a descent direction, the line search and the
      // next iteration absorb the rounding) — no sqrt + divide on the serial column chain
      const float pv = fmaxf(__shfl_sync(PB_FULL, h[j], j), PB_MINVAL), dinv = rsqrtf(pv);
      if (lane == j) hinv = dinv;          // 1 / L[j][j], used by both substitutions
      h[j] = lane == j ? pv * dinv : h[j] * dinv;
#pragma unroll
      for (int b = 0; b < NV; b++) {   // constant trip count (b > j resolved after unrolling): h / ht must stay in registers
        if (b > j) {
          const float hb = __shfl_sync(PB_FULL, h[j], b);   // L[b][j]
          if (lane == j) ht[b] = hb;
          h[b] -= h[j] * hb;                                // L[a][b] -= L[a][j] L[b][j] for the rows a >= b that need it
        }
      }
    }
    float x = lane < NV ? -grad[lane] : 0.f;
    x *= hinv;   // lane k carries x[k] / L[k][k] throughout: the value it has to broadcast at its own step
#pragma unroll
    for (int k = 0; k < NV; k++) {  // forward substitution
      const float xk = __shfl_sync(PB_FULL, x, k);
      if (lane > k) x -= h[k] * hinv * xk;
    }
    x *= hinv;
#pragma unroll
    for (int k = NV - 1; k >= 0; k--) {  // back substitution with L^T
      const float xk = __shfl_sync(PB_FULL, x, k);
      if (lane < k) x -= ht[k] * hinv * xk;
    }
    __syncwarp();
    if (lane < NV) srch[lane] = x;
    __syncwarp();
    // line search: phi'(alpha) = c0 + c1 alpha + sum_{active(alpha)} D (jar + alpha jv) jv
    float c0p = 0.f, c1p = 0.f;
    if (lane < NV) {
      float mv = 0.f;
      for (int j = 0; j < NV; j++) mv += Md[lane * LDH + j] * srch[j];
      Mv[lane] = mv;
      c0p = srch[lane] * (Ma[lane] - s.qfrc_smooth[lane]);
      c1p = srch[lane] * mv;
    }
    // the rows of this lane stay in registers for the whole search (jar, jv, D = 1 / R)
    constexpr int RPL = (C::MAXEFC + 31) / 32;
    float ls_jar[RPL], ls_jv[RPL], ls_D[RPL];
#pragma unroll
    for (int q = 0; q < RPL; q++) {
      const int r = lane + 32 * q;
      ls_jar[q] = 0.f; ls_jv[q] = 0.f; ls_D[q] = 0.f;
      if (r < nefc) {
        float jv = 0.f;
        for (int i = 0; i < NV; i++) jv += s.J[r][i] * srch[i];
        s.efc_jv[r] = jv;
        ls_jv[q] = jv; ls_jar[q] = s.efc_jar[r]; ls_D[q] = 1.f / s.efc_R[r];
      }
    }
    float c0 = warp_sum(c0p), c1 = warp_sum(c1p);
    __syncwarp();
    float alpha = 0.f, lo = 0.f, hi = INFINITY;
    for (int ls = 0; ls < 24; ls++) {
      float d1p = 0.f, d2p = 0.f;
#pragma unroll
      for (int q = 0; q < RPL; q++) {
        const float jv = ls_jv[q], xx = ls_jar[q] + alpha * jv;
        if (ls_D[q] != 0.f && (xx < 0.f || (xx == 0.f && jv < 0.f))) { d1p += ls_D[q] * xx * jv; d2p += ls_D[q] * jv * jv; }
      }
      float d1 = c0 + c1 * alpha + warp_sum(d1p), d2 = c1 + warp_sum(d2p);
      if (d1 < 0.f) lo = alpha; else hi = alpha;
      float next = alpha - d1 / d2;
      if (!(next > lo && next < hi)) next = (hi < INFINITY) ? 0.5f * (lo + hi) : 2.f * alpha + 1e-6f;
      bool conv = fabsf(next - alpha) <= 1e-6f * fabsf(next);   // 1e-7 is below one fp32 ulp: the iterate can bounce forever
      alpha = next;
      if (conv) break;
    }
    if (lane < NV) qacc[lane] += alpha * srch[lane];
    __syncwarp();
  }
}
