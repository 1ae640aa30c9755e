// pb_sim_pos.cuh — position stage of the forward dynamics for one env per warp:
// kinematics -> tree com -> cinert/cdof -> composite inertia -> sparse M -> L^T D L -> collision -> constraint rows.
// Replaces the position half of mj_forward reached through libs/gymcpp/mujoco/mujoco_env.h:103.
#pragma once
#include "pb_sim_common.cuh"

// Kinematics without a level-by-level walk (a level of these trees holds 1-3 bodies, so a level loop leaves ~2 of 32 lanes
// busy): (A) every body lane builds its transform relative to its PARENT frame - joint rotations, off-centre anchors - in
// registers; (B) pointer jumping over the parent links composes them into world transforms in ceil(log2(depth)) shuffle
// rounds; (C) joint lanes move their parent-frame anchors / axes to the world frame.  Same quantities as mj_kinematics; the
// products are associated differently, i.e. equal up to fp32 rounding.
template <class C>
__device__ void sim_kinematics(EnvS<C> &s, const DevModel &m, int lane) {
  float R[9] = {1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f}, pos[3] = {0.f, 0.f, 0.f};
  int ptr = -1;
  if (lane >= 1 && lane < C::NB) {
    const int b = lane, ja = m.body_jntadr[b], jn = m.body_jntnum[b];
    const int par = m.body_parent[b];
    ptr = par > 0 ? par : -1;
    float quat[4], t[3];
    if (jn == 1 && m.jnt_type[ja] == PB_JNT_FREE) {
      int a = m.jnt_qposadr[ja];
      quat[0] = s.qpos[a + 3]; quat[1] = s.qpos[a + 4]; quat[2] = s.qpos[a + 5]; quat[3] = s.qpos[a + 6];
      quat_normalize(quat);
      for (int k = 0; k < 4; k++) s.qpos[a + 3 + k] = quat[k];  // normalised in place, as the library does
      for (int k = 0; k < 3; k++) { pos[k] = s.qpos[a + k]; s.xanchor[ja][k] = pos[k]; }
      s.xaxis[ja][0] = 0.f; s.xaxis[ja][1] = 0.f; s.xaxis[ja][2] = 1.f;
    } else {
      for (int k = 0; k < 3; k++) pos[k] = m.body_pos[b][k];
      for (int k = 0; k < 4; k++) quat[k] = m.body_quat[b][k];
      for (int j = ja; j < ja + jn; j++) {
        float q = s.qpos[m.jnt_qposadr[j]] - m.jnt_qpos0[j];
        quat2mat(R, quat);
        mat3vec(t, R, m.jnt_pos[j]);
        float anchor[3], axis[3];
        for (int k = 0; k < 3; k++) anchor[k] = t[k] + pos[k];
        mat3vec(axis, R, m.jnt_axis[j]);
        for (int k = 0; k < 3; k++) { s.xanchor[j][k] = anchor[k]; s.xaxis[j][k] = axis[k]; }  // parent frame until (C)
        if (m.jnt_type[j] == PB_JNT_SLIDE) {
          for (int k = 0; k < 3; k++) pos[k] += axis[k] * q;
        } else {
          float sn, cs, ql[4], qn[4];
          sincosf(0.5f * q, &sn, &cs);
          ql[0] = cs; ql[1] = m.jnt_axis[j][0] * sn; ql[2] = m.jnt_axis[j][1] * sn; ql[3] = m.jnt_axis[j][2] * sn;
          quat_mul(qn, quat, ql);
          for (int k = 0; k < 4; k++) quat[k] = qn[k];
          quat2mat(R, quat);
          mat3vec(t, R, m.jnt_pos[j]);  // off-centre rotation: the anchor stays where it was
          for (int k = 0; k < 3; k++) pos[k] = anchor[k] - t[k];
        }
      }
    }
    quat_normalize(quat);
    quat2mat(R, quat);
  }
  for (int r = 0; r < m.body_rounds; r++) {  // (B) T_b <- T_ptr(b) o T_b ; ptr(b) <- ptr(ptr(b))
    const int src = ptr < 0 ? lane : ptr;
    float Rp[9], pp[3];
#pragma unroll
    for (int k = 0; k < 9; k++) Rp[k] = __shfl_sync(PB_FULL, R[k], src);
#pragma unroll
    for (int k = 0; k < 3; k++) pp[k] = __shfl_sync(PB_FULL, pos[k], src);
    const int nptr = __shfl_sync(PB_FULL, ptr, src);
    if (ptr >= 0) {
      float np[3], nR[9];
      mat3vec(np, Rp, pos);
#pragma unroll
      for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) nR[3 * i + j] = Rp[3 * i] * R[j] + Rp[3 * i + 1] * R[3 + j] + Rp[3 * i + 2] * R[6 + j];
#pragma unroll
      for (int k = 0; k < 3; k++) pos[k] = pp[k] + np[k];
#pragma unroll
      for (int k = 0; k < 9; k++) R[k] = nR[k];
      ptr = nptr;
    }
  }
  if (lane < C::NB) {  // lane 0 is the world frame (identity); its slots are shared with later stages, so rewrite them every time
    float t[3];
    mat3vec(t, R, m.body_ipos[lane]);
    for (int k = 0; k < 3; k++) { s.xpos[lane][k] = pos[k]; s.xipos[lane][k] = pos[k] + t[k]; }
    for (int k = 0; k < 9; k++) s.xmat[lane][k] = R[k];
  }
  __syncwarp();
  if (lane < C::NJ) {  // (C)
    const int j = lane, p = m.body_parent[m.jnt_body[j]];
    float a[3], x[3];
    mat3vec(a, s.xmat[p], s.xanchor[j]);
    mat3vec(x, s.xmat[p], s.xaxis[j]);
    for (int k = 0; k < 3; k++) { s.xanchor[j][k] = s.xpos[p][k] + a[k]; s.xaxis[j][k] = x[k]; }
  }
  for (int g = lane; g < C::NG; g += 32) {
    int b = m.geom_body[g];
    float t[3];
    mat3vec(t, s.xmat[b], m.geom_pos[g]);
    for (int k = 0; k < 3; k++) s.gxpos[g][k] = s.xpos[b][k] + t[k];
    mat3vec(s.gz[g], s.xmat[b], m.geom_zaxis[g]);
  }
  __syncwarp();
}

// tree com, cinert, cdof (mj_comPos); all bodies share the root body 1
template <class C>
__device__ void sim_com_pos(EnvS<C> &s, const DevModel &m, int lane) {
  float cx = 0.f, cy = 0.f, cz = 0.f;
  if (lane >= 1 && lane < C::NB) {
    float ms = m.body_mass[lane];
    cx = ms * s.xipos[lane][0]; cy = ms * s.xipos[lane][1]; cz = ms * s.xipos[lane][2];
  }
  float inv = 1.f / m.total_mass;
  cx = warp_sum(cx) * inv; cy = warp_sum(cy) * inv; cz = warp_sum(cz) * inv;
  if (lane == 0) { s.com[0] = cx; s.com[1] = cy; s.com[2] = cz; }
  const float com[3] = {cx, cy, cz};
  if (lane < C::NB) {
    float *ci = s.cinert[lane];
    if (lane == 0) {
      for (int k = 0; k < 10; k++) ci[k] = 0.f;
    } else {
      const float *R = s.xmat[lane], *I = m.body_inertia[lane];
      // W = R * Isym * R^T
      float RI[9];
      for (int r = 0; r < 3; r++) {
        float a = R[3 * r], b = R[3 * r + 1], c = R[3 * r + 2];
        RI[3 * r] = a * I[0] + b * I[3] + c * I[4];
        RI[3 * r + 1] = a * I[3] + b * I[1] + c * I[5];
        RI[3 * r + 2] = a * I[4] + b * I[5] + c * I[2];
      }
      float W00 = RI[0] * R[0] + RI[1] * R[1] + RI[2] * R[2], W11 = RI[3] * R[3] + RI[4] * R[4] + RI[5] * R[5];
      float W22 = RI[6] * R[6] + RI[7] * R[7] + RI[8] * R[8], W01 = RI[0] * R[3] + RI[1] * R[4] + RI[2] * R[5];
      float W02 = RI[0] * R[6] + RI[1] * R[7] + RI[2] * R[8], W12 = RI[3] * R[6] + RI[4] * R[7] + RI[5] * R[8];
      float d0 = s.xipos[lane][0] - com[0], d1 = s.xipos[lane][1] - com[1], d2 = s.xipos[lane][2] - com[2];
      float ms = m.body_mass[lane], dd = d0 * d0 + d1 * d1 + d2 * d2;
      ci[0] = W00 + ms * (dd - d0 * d0); ci[1] = W11 + ms * (dd - d1 * d1); ci[2] = W22 + ms * (dd - d2 * d2);
      ci[3] = W01 - ms * d0 * d1; ci[4] = W02 - ms * d0 * d2; ci[5] = W12 - ms * d1 * d2;
      ci[6] = ms * d0; ci[7] = ms * d1; ci[8] = ms * d2; ci[9] = ms;
    }
  }
  if (lane < C::NJ) {
    int j = lane, da = m.jnt_dofadr[j], type = m.jnt_type[j];
    float off[3] = {com[0] - s.xanchor[j][0], com[1] - s.xanchor[j][1], com[2] - s.xanchor[j][2]};
    if (type == PB_JNT_FREE) {
      int b = 1;  // the free joint belongs to the root body
      for (int k = 0; k < 3; k++) {
        for (int c = 0; c < 6; c++) s.cdof[da + k][c] = (c == 3 + k) ? 1.f : 0.f;
        float ax[3] = {s.xmat[b][k], s.xmat[b][3 + k], s.xmat[b][6 + k]}, cr[3];
        cross3(cr, ax, off);
        for (int c = 0; c < 3; c++) { s.cdof[da + 3 + k][c] = ax[c]; s.cdof[da + 3 + k][3 + c] = cr[c]; }
      }
    } else if (type == PB_JNT_SLIDE) {
      for (int c = 0; c < 3; c++) { s.cdof[da][c] = 0.f; s.cdof[da][3 + c] = s.xaxis[j][c]; }
    } else {
      float cr[3];
      cross3(cr, s.xaxis[j], off);
      for (int c = 0; c < 3; c++) { s.cdof[da][c] = s.xaxis[j][c]; s.cdof[da][3 + c] = cr[c]; }
    }
  }
  __syncwarp();
}

// One elimination step of the in-place L^T D L (dof K = NV-1-STEP), unrolled over the compile-time dof tree: the update
// list of the step is [fac_start, fac_start + c(c+1)/2) in the precomputed schedule, c = number of ancestors of K.  The
// schedule words of step STEP+1 are fetched before step STEP's arithmetic, so the (L1-hit, ~40 cycle) loads are off the
// serial load -> shared-memory read-modify-write -> __syncwarp chain.
template <class C, int STEP>
struct FacStep {
  static constexpr int K = C::NV - 1 - STEP, KK = Tree<C>::madr(K < 0 ? 0 : K), NC = K < 0 ? 0 : Tree<C>::depth(K) - 1;
  static constexpr int E0 = Tree<C>::fac_start(STEP > C::NV ? C::NV : STEP), NE = K < 0 ? 0 : Tree<C>::nanc_pairs(K), NBATCH = (NE + 31) / 32;
};
constexpr int PB_FAC_MAXBATCH = 3;  // 12 ancestors -> 78 updates
template <class C, int STEP>
__device__ __forceinline__ void factor_fetch(const DevModel &m, int lane, unsigned (&u)[PB_FAC_MAXBATCH]) {
  if constexpr (STEP < C::NV) {
    using F = FacStep<C, STEP>;
    static_assert(F::NBATCH <= PB_FAC_MAXBATCH, "raise PB_FAC_MAXBATCH");
#pragma unroll
    for (int bt = 0; bt < F::NBATCH; bt++) if (32 * bt + lane < F::NE) u[bt] = m.fac_tab[F::E0 + 32 * bt + lane];
  }
}
template <class C, int STEP>
__device__ __forceinline__ void factor_steps(EnvS<C> &s, const DevModel &m, int lane, const unsigned (&u)[PB_FAC_MAXBATCH]) {
  if constexpr (STEP < C::NV) {
    using F = FacStep<C, STEP>;
    unsigned un[PB_FAC_MAXBATCH];
    factor_fetch<C, STEP + 1>(m, lane, un);
    // pivot reciprocal by MUFU.RCP alone (1 ulp): the IEEE-rounded form spends ~6 more instructions per step on the
    // serial chain for half an ulp that the fp32 factor does not keep anyway
    float inv;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(s.qLD[F::KK]));
    if constexpr (F::NC > 0) {
#pragma unroll
      for (int bt = 0; bt < F::NBATCH; bt++) {
        if (32 * bt + lane < F::NE) {
          const unsigned w = u[bt];
          s.qLD[w & 255u] -= s.qLD[(w >> 8) & 255u] * s.qLD[(w >> 16) & 255u] * inv;
        }
      }
      __syncwarp();
      if (lane < F::NC) s.qLD[F::KK + 1 + lane] *= inv;
    }
    if (lane == 0) s.qLDinv[F::K] = inv;
    __syncwarp();
    factor_steps<C, STEP + 1>(s, m, lane, un);
  }
}

// In-place L^T D L in REGISTERS.  Lane i holds row i of the sparse M indexed by the depth of the column dof on i's chain
// (r[d] = M[i][ancestor of i at depth d], r[depth(i)-1] the diagonal); all ancestors of a dof lie on one chain, so the rows of K
// and of any ancestor of K use the same index for the same column.  Elimination step K (NV-1 .. 0, compile-time): row K is
// broadcast with depth(K) shuffles, every strict ancestor i of K picks L_Ki = M[K][i] / M[K][K] out of the broadcast (a select
// chain on its own depth) and subtracts L_Ki * M[K][j] from its entries j <= i.  No shared-memory round trips and no warp
// barriers on the 23-step serial chain, which cost ~300 cycles per step in the shared-memory version (7 k of the 52 k cycles of
// a forward evaluation, profiles/r02_stage_cycles.txt).  Rows are scaled to L (and 1/D kept) when they are written back.
template <class C, int K>
__device__ __forceinline__ void factor_reg_steps(float (&r)[Tree<C>::maxdepth()], float &myinv, int lane, int mydiag) {
  if constexpr (K >= 0) {
    constexpr int DK = Tree<C>::depth(K);
    float b[DK];
#pragma unroll
    for (int d = 0; d < DK; d++) b[d] = __shfl_sync(PB_FULL, r[d], K);
    float inv;   // MUFU.RCP alone (1 ulp), as before
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(b[DK - 1]));
    if (lane == K) myinv = inv;
    if constexpr (DK > 1) {
      constexpr unsigned ANC = Tree<C>::ancmask(K);
      float l = 0.f;
#pragma unroll
      for (int d = 0; d < DK - 1; d++) if (mydiag == d) l = b[d];
      l = ((ANC >> lane) & 1u) ? l * inv : 0.f;
      // entries beyond the lane's own depth are scratch: updated along, never stored or broadcast
#pragma unroll
      for (int d = 0; d < DK - 1; d++) r[d] -= l * b[d];
    }
    factor_reg_steps<C, K - 1>(r, myinv, lane, mydiag);
  }
}

// composite inertia by subtree range sums, sparse M, in-place L^T D L (mj_crb + mj_factorM)
template <class C>
__device__ void sim_crb_factor(EnvS<C> &s, const DevModel &m, int lane) {
  const float my_armature = lane < C::NV ? m.dof_armature[lane] : 0.f;
  const int my_diag = lane < C::NV ? m.dof_Madr[lane] : 0;
  for (int it = lane; it < C::NB * 10; it += 32) {
    int b = it / 10, c = it - 10 * b;
    float acc = 0.f;
    if (b > 0) {
      int n = m.body_subtree[b];
      for (int k = 0; k < n; k++) acc += s.cinert[b + k][c];
    }
    s.tmpA[b][c] = acc;
  }
  __syncwarp();
  if (lane < C::NV) mul_inert_vec(s.tmpB[lane], s.tmpA[m.dof_body[lane]], s.cdof[lane]);
  __syncwarp();
  for (int e = lane; e < C::NM; e += 32) {
    int i = m.M_i[e], j = m.M_j[e];
    const float *a = s.cdof[j], *b = s.tmpB[i];
    float v = a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] + a[5] * b[5];
    if constexpr (!EnvS<C>::PGS) s.qM[e] = v;
    s.qLD[e] = v;
  }
  __syncwarp();
  if (lane < C::NV) {  // armature on the diagonal
    const float v = s.qLD[my_diag] + my_armature;
    if constexpr (!EnvS<C>::PGS) s.qM[my_diag] = v;
    s.qLD[my_diag] = v;
  }
  __syncwarp();
  {
    constexpr int MD = Tree<C>::maxdepth();
    const int mydepth = lane < C::NV ? m.dof_depth[lane] : 0, mydiag = mydepth - 1;
    float r[MD], myinv = 0.f;
#pragma unroll
    for (int d = 0; d < MD; d++) r[d] = d < mydepth ? s.qLD[my_diag + mydiag - d] : 0.f;   // row k: diagonal first, then the ancestors going up
    factor_reg_steps<C, C::NV - 1>(r, myinv, lane, mydiag);
#pragma unroll
    for (int d = 0; d < MD; d++) if (d < mydiag) s.qLD[my_diag + mydiag - d] = r[d] * myinv;
#pragma unroll
    for (int d = 0; d < MD; d++) if (d == mydiag) s.qLD[my_diag] = r[d];
    if (lane < C::NV) s.qLDinv[lane] = myinv;
  }
  __syncwarp();
}

// mj_solveLD for one right-hand side with element i in lane i: both triangular sweeps run in registers over the
// compile-time dof tree - the finished x[K] is broadcast with a shuffle and the lanes that are ancestors (L^-T sweep) or
// descendants (L^-1 sweep) of K apply it; the factor is read from shared memory.  No barriers inside.
template <class C, int K>
__device__ __forceinline__ void solve_up(float &xl, const float *qLD, int lane, int mydepth) {  // x <- L^-T x, K = NV-1 .. 1
  if constexpr (K >= 1) {
    constexpr unsigned ANC = Tree<C>::ancmask(K);
    if constexpr (ANC != 0u) {
      const float xk = __shfl_sync(PB_FULL, xl, K);
      if ((ANC >> lane) & 1u) xl -= qLD[Tree<C>::madr(K) + Tree<C>::depth(K) - mydepth] * xk;
    }
    solve_up<C, K - 1>(xl, qLD, lane, mydepth);
  }
}
template <class C, int K>
__device__ __forceinline__ void solve_down(float &xl, const float *qLD, int lane, int myend) {  // x <- L^-1 x, K = 0 .. NV-2
  if constexpr (K < C::NV - 1) {
    constexpr unsigned DESC = Tree<C>::descmask(K);
    if constexpr (DESC != 0u) {
      const float xk = __shfl_sync(PB_FULL, xl, K);
      if ((DESC >> lane) & 1u) xl -= qLD[myend - Tree<C>::depth(K)] * xk;
    }
    solve_down<C, K + 1>(xl, qLD, lane, myend);
  }
}
// x <- M^-1 x, x in shared memory
template <class C>
__device__ void sim_solve_LD(EnvS<C> &s, const DevModel &m, float *x, int lane) {
  const int mydepth = lane < C::NV ? m.dof_depth[lane] : 0, myend = lane < C::NV ? m.dof_Madr[lane] + mydepth : 0;
  float xl = lane < C::NV ? x[lane] : 0.f;
  solve_up<C, C::NV - 1>(xl, s.qLD, lane, mydepth);
  if (lane < C::NV) xl *= s.qLDinv[lane];
  solve_down<C, 0>(xl, s.qLD, lane, myend);
  if (lane < C::NV) x[lane] = xl;
  __syncwarp();
}

// ------------------------------------------------------------------ collision
struct Cand { float dist, pos[3], n[3]; };

__device__ __forceinline__ bool sphere_sphere(Cand &c, float margin, const float *p1, float r1, const float *p2, float r2) {
  float d[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  float cd = sqrtf(dot3(d, d)), dist = cd - r1 - r2;
  if (dist > margin) return false;
  if (cd >= PB_MINVAL) { float i = 1.f / cd; c.n[0] = d[0] * i; c.n[1] = d[1] * i; c.n[2] = d[2] * i; }
  else { c.n[0] = 1.f; c.n[1] = 0.f; c.n[2] = 0.f; }
  c.dist = dist;
  for (int k = 0; k < 3; k++) c.pos[k] = p1[k] + c.n[k] * (r1 + 0.5f * dist);
  return true;
}
__device__ __forceinline__ bool plane_sphere(Cand &c, float margin, const float *pp, const float *pn, const float *sp, float r) {
  float d[3] = {sp[0] - pp[0], sp[1] - pp[1], sp[2] - pp[2]};
  float dist = dot3(d, pn) - r;
  if (dist > margin) return false;
  c.dist = dist;
  for (int k = 0; k < 3; k++) { c.n[k] = pn[k]; c.pos[k] = sp[k] - pn[k] * (r + 0.5f * dist); }
  return true;
}
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

template <class C>
__device__ void sim_collision(EnvS<C> &s, const DevModel &m, int lane) {
  int ncon = 0;
  for (int base = 0; base < C::NPAIR; base += 32) {
    int p = base + lane;
    Cand c0, c1;
    bool v0 = false, v1 = false, cap_axis = false;
    float axis2[3] = {0.f, 0.f, 0.f};
    if (p < C::NPAIR) {
      const int g1 = m.pair_g1[p], g2 = m.pair_g2[p];
      const float *p1 = s.gxpos[g1], *p2 = s.gxpos[g2];
      const float d[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
      // bounding-sphere reject first (most pairs, most of the time); plane pairs have an infinite reach
      if (dot3(d, d) <= m.pair_reach2[p]) {
        const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
        const float margin = m.pair_margin[p];
        const float *a1 = s.gz[g1], *a2 = s.gz[g2];
        const float r1 = m.geom_size[g1][0], h1 = m.geom_size[g1][1], r2 = m.geom_size[g2][0], h2 = m.geom_size[g2][1];
        if (t1 == PB_GEOM_PLANE) {
          if (t2 == PB_GEOM_SPHERE) {
            v0 = plane_sphere(c0, margin, p1, a1, p2, r2);
          } else {
            float e[3];
            for (int k = 0; k < 3; k++) { axis2[k] = a2[k]; e[k] = p2[k] + a2[k] * h2; }
            v0 = plane_sphere(c0, margin, p1, a1, e, r2);
            for (int k = 0; k < 3; k++) e[k] = p2[k] - a2[k] * h2;
            v1 = plane_sphere(c1, margin, p1, a1, e, r2);
            cap_axis = true;
          }
        } else if (t1 == PB_GEOM_SPHERE && t2 == PB_GEOM_SPHERE) {
          v0 = sphere_sphere(c0, margin, p1, r1, p2, r2);
        } else if (t1 == PB_GEOM_SPHERE) {  // sphere - capsule
          float v[3];
          float x = clampf(dot3(a2, d), -h2, h2);
          for (int k = 0; k < 3; k++) v[k] = p2[k] + a2[k] * x;
          v0 = sphere_sphere(c0, margin, p1, r1, v, r2);
        } else {  // capsule - capsule: closest points of the axis segments
          float mb = -dot3(a1, a2), u = -dot3(a1, d), v = dot3(a2, d), det = 1.f - mb * mb, x1, x2;
          if (fabsf(det) >= 1e-12f) { x1 = (u - mb * v) / det; x2 = (v - mb * u) / det; } else { x1 = 0.f; x2 = v; }
          if (x1 > h1) { x1 = h1; x2 = v - mb * x1; } else if (x1 < -h1) { x1 = -h1; x2 = v - mb * x1; }
          if (x2 > h2) { x2 = h2; x1 = clampf(u - mb * x2, -h1, h1); }
          else if (x2 < -h2) { x2 = -h2; x1 = clampf(u - mb * x2, -h1, h1); }
          float v1p[3], v2p[3];
          for (int k = 0; k < 3; k++) { v1p[k] = p1[k] + a1[k] * x1; v2p[k] = p2[k] + a2[k] * x2; }
          v0 = sphere_sphere(c0, margin, v1p, r1, v2p, r2);
        }
      }
    }
    if (!__any_sync(PB_FULL, v0 || v1)) continue;  // nothing found in this batch of 32 pairs
    int cnt = (v0 ? 1 : 0) + (v1 ? 1 : 0), total;
    int off = ncon + warp_excl_scan(cnt, lane, &total);
    for (int w = 0; w < 2; w++) {
      bool v = w ? v1 : v0;
      if (!v) continue;
      if (off >= C::MAXCON) { s.overflow = 1; continue; }
      const Cand &c = w ? c1 : c0;
      Contact &o = s.con[off++];
      o.dist = c.dist; o.pair = p; o.efc = -1;
      float f[6] = {c.n[0], c.n[1], c.n[2], 0.f, 0.f, 0.f};
      if (cap_axis) { f[3] = axis2[0]; f[4] = axis2[1]; f[5] = axis2[2]; }
      // mju_makeFrame: default y axis, orthogonalise against the normal
      if (dot3(f + 3, f + 3) < 0.25f) { f[3] = 0.f; f[4] = 0.f; f[5] = 0.f; if (f[1] < 0.5f && f[1] > -0.5f) f[4] = 1.f; else f[5] = 1.f; }
      float t = dot3(f, f + 3);
      for (int k = 0; k < 3; k++) f[3 + k] -= t * f[k];
      normalize3(f + 3);
      for (int k = 0; k < 3; k++) { o.pos[k] = c.pos[k]; o.n[k] = f[k]; o.t1[k] = f[3 + k]; }
    }
    ncon = min(ncon + total, C::MAXCON);
  }
  if (lane == 0) s.ncon = ncon;
  __syncwarp();
}

// impedance d(r) of one row (MuJoCo solimp sigmoid); solimp already clamped on the host
__device__ __forceinline__ float impedance(const float *si, float pos, float margin) {
  if (si[0] == si[1] || si[2] <= PB_MINVAL) return 0.5f * (si[0] + si[1]);
  float x = fabsf((pos - margin) / si[2]);
  if (x >= 1.f) return si[1];
  if (x <= 0.f) return si[0];
  float y;
  if (si[4] == 1.f) y = x;
  else if (si[4] == 2.f) y = (x <= si[3]) ? x * x / si[3] : 1.f - (1.f - x) * (1.f - x) / (1.f - si[3]);
  else y = (x <= si[3]) ? powf(x, si[4]) / powf(si[3], si[4] - 1.f) : 1.f - powf(1.f - x, si[4]) / powf(1.f - si[3], si[4] - 1.f);
  return si[0] + y * (si[1] - si[0]);
}

// constraint rows: active joint limits (lower side first) then contacts in pair order (mj_makeConstraint)
template <class C>
__device__ void sim_make_constraint(EnvS<C> &s, const DevModel &m, int lane) {
  // --- limits: lane = joint
  bool lo = false, hi = false;
  float dlo = 0.f, dhi = 0.f;
  if (lane < C::NJ && m.jnt_limited[lane]) {
    float v = s.qpos[m.jnt_qposadr[lane]], mg = m.jnt_margin[lane];
    dlo = v - m.jnt_range[lane][0]; dhi = m.jnt_range[lane][1] - v;
    lo = dlo < mg; hi = dhi < mg;
  }
  int nlim, row = warp_excl_scan((lo ? 1 : 0) + (hi ? 1 : 0), lane, &nlim);
  // --- contacts: lane = contact, rows per included contact
  int ncon = s.ncon, crow = 0, nrows = 0;
  if (lane < ncon) {
    const Contact &c = s.con[lane];
    if (c.dist < m.pair_incmargin[c.pair]) nrows = (m.pair_condim[c.pair] == 1) ? 1 : 4;
  }
  int ncrow;
  crow = nlim + warp_excl_scan(nrows, lane, &ncrow);
  // rows that do not fit are dropped together with every later contact (keeps the row block contiguous)
  unsigned dropped = __ballot_sync(PB_FULL, nrows > 0 && crow + nrows > C::MAXEFC);
  int first_drop = dropped ? (__ffs(dropped) - 1) : 32;
  bool included = nrows > 0 && lane < first_drop;
  if (lane < ncon) s.con[lane].efc = included ? crow : -1;
  if (dropped && lane == 0) s.overflow = 1;
  int nefc = nlim + ncrow;
  if (dropped) nefc = __shfl_sync(PB_FULL, crow, first_drop);
  // zero the limit rows, then drop the single entries in
  for (int i = lane; i < nlim * EnvS<C>::LDJ; i += 32) (&s.J[0][0])[i] = 0.f;
  __syncwarp();
  for (int side = 0; side < 2; side++) {
    bool act = side ? hi : lo;
    if (!act) continue;
    int j = lane;
    float dist = side ? dhi : dlo, mg = m.jnt_margin[j];
    float imp = impedance(m.jnt_solimp[j], dist, mg);
    s.J[row][m.jnt_dofadr[j]] = side ? -1.f : 1.f;
    s.efc_R[row] = fmaxf(PB_MINVAL, (1.f - imp) * m.jnt_diag[j] / imp);
    s.efc_kip[row] = m.jnt_k[j] * imp * (dist - mg);
    s.efc_bd[row] = m.jnt_b[j];
    row++;
  }
  // --- contact rows: lane = dof column
  for (int c = 0; c < ncon; c++) {
    const Contact &con = s.con[c];
    int r0 = con.efc;
    if (r0 < 0) continue;
    int p = con.pair, dim = m.pair_condim[p];
    float mu = m.pair_mu[p];
    if (lane < C::NV) {
      unsigned m1 = m.body_dofmask[m.geom_body[m.pair_g1[p]]], m2 = m.body_dofmask[m.geom_body[m.pair_g2[p]]];
      float sg = (float)((int)((m2 >> lane) & 1u) - (int)((m1 >> lane) & 1u));
      float off[3] = {con.pos[0] - s.com[0], con.pos[1] - s.com[1], con.pos[2] - s.com[2]}, jd[3];
      cross3(jd, s.cdof[lane], off);  // point Jacobian column = lin + ang x (p - com)
      for (int k = 0; k < 3; k++) jd[k] = sg * (jd[k] + s.cdof[lane][3 + k]);
      float jn = dot3(con.n, jd);
      if (dim == 1) {
        s.J[r0][lane] = jn;
      } else {
        float t2[3];
        cross3(t2, con.n, con.t1);
        float j1 = mu * dot3(con.t1, jd), j2 = mu * dot3(t2, jd);
        s.J[r0][lane] = jn + j1; s.J[r0 + 1][lane] = jn - j1; s.J[r0 + 2][lane] = jn + j2; s.J[r0 + 3][lane] = jn - j2;
      }
    }
    int nr = dim == 1 ? 1 : 4;
    if (lane < nr) {
      float mg = m.pair_incmargin[p], imp = impedance(m.pair_solimp[p], con.dist, mg);
      float R = fmaxf(PB_MINVAL, (1.f - imp) * m.pair_diag[p] / imp);
      if (dim > 1) R = fmaxf(PB_MINVAL, 2.f * mu * mu * R);
      s.efc_R[r0 + lane] = R;
      s.efc_kip[r0 + lane] = m.pair_k[p] * imp * (con.dist - mg);
      s.efc_bd[r0 + lane] = m.pair_b[p];
    }
  }
  if (lane == 0) s.nefc = nefc;
  __syncwarp();
}
