// pb_devmodel.cpp — host-side packing of pb_model into the float DevModel the kernels read (see pb_devmodel.h).
#include "pb_devmodel.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace {
const double kMinImp = 0.0001, kMaxImp = 0.9999;

void quat2mat(const double *q, double *m) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}

// constraint stiffness/damping from (solref, solimp) with refsafe; solimp is clamped in place
void kb_from_sol(const double *solref, const double *solimp_in, double timestep, float *solimp_out, float *k, float *b) {
  double s[5];
  s[0] = std::min(kMaxImp, std::max(kMinImp, solimp_in[0]));
  s[1] = std::min(kMaxImp, std::max(kMinImp, solimp_in[1]));
  s[2] = std::max(0.0, solimp_in[2]);
  s[3] = std::min(kMaxImp, std::max(kMinImp, solimp_in[3]));
  s[4] = std::max(1.0, solimp_in[4]);
  for (int i = 0; i < 5; i++) solimp_out[i] = (float)s[i];
  double r0 = solref[0], r1 = solref[1];
  if (r0 > 0 && r0 < 2 * timestep) r0 = 2 * timestep;
  if (r0 > 0) {
    *k = (float)(1.0 / (s[1] * s[1] * r0 * r0 * r1 * r1));
    *b = (float)(2.0 / (s[1] * r0));
  } else {
    *k = (float)(-r0 / (s[1] * s[1]));
    *b = (float)(-r1 / s[1]);
  }
}
}  // namespace

extern "C" int pb_build_devmodel(const pb_model *m, DevModel *d) {
  std::memset(d, 0, sizeof(*d));
  d->nq = m->nq; d->nv = m->nv; d->nu = m->nu; d->nbody = m->nbody; d->njnt = m->njnt; d->ngeom = m->ngeom;
  d->npair = m->npair; d->iterations = m->iterations;
  d->timestep = (float)m->timestep; d->tolerance = (float)m->tolerance; d->meaninertia = (float)m->meaninertia;
  d->solver_scale = (float)(1.0 / (m->meaninertia * std::max(1, m->nv)));
  for (int k = 0; k < 3; k++) d->gravity[k] = (float)m->gravity[k];
  double tm = 0;
  for (int b = 0; b < m->nbody; b++) {
    tm += m->body_mass[b];
    d->body_parent[b] = m->body_parent[b]; d->body_jntadr[b] = m->body_jntadr[b]; d->body_jntnum[b] = m->body_jntnum[b];
    d->body_dofadr[b] = m->body_dofadr[b]; d->body_dofnum[b] = m->body_dofnum[b]; d->body_subtree[b] = m->body_subtree[b];
    for (int k = 0; k < 3; k++) { d->body_pos[b][k] = (float)m->body_pos[b][k]; d->body_ipos[b][k] = (float)m->body_ipos[b][k]; }
    for (int k = 0; k < 4; k++) d->body_quat[b][k] = (float)m->body_quat[b][k];
    d->body_mass[b] = (float)m->body_mass[b];
    const double *I = m->body_inertia[b];
    const double sym[6] = {I[0], I[4], I[8], I[1], I[2], I[5]};
    for (int k = 0; k < 6; k++) d->body_inertia[b][k] = (float)sym[k];
    if (m->body_root[b] != 1 && b > 0) return -2;  // kernels assume a single kinematic tree rooted at body 1
  }
  d->total_mass = (float)tm;
  // dof ancestor masks
  for (int b = 1; b < m->nbody; b++) {
    unsigned mask = d->body_dofmask[m->body_parent[b]];
    for (int i = 0; i < m->body_dofnum[b]; i++) mask |= 1u << (m->body_dofadr[b] + i);
    d->body_dofmask[b] = mask;
  }
  // levels by depth (world excluded)
  int maxdepth = 0;
  for (int b = 1; b < m->nbody; b++) maxdepth = std::max(maxdepth, m->body_depth[b]);
  d->nlevel = maxdepth;
  int n = 0;
  for (int dep = 1; dep <= maxdepth; dep++) {
    d->level_start[dep - 1] = n;
    for (int b = 1; b < m->nbody; b++) if (m->body_depth[b] == dep) d->level_body[n++] = b;
  }
  d->level_start[maxdepth] = n;
  d->body_rounds = 0;
  while ((1 << d->body_rounds) < maxdepth) d->body_rounds++;
  for (int b = 0; b < m->nbody; b++) {
    int a = b;
    while (a > 0 && m->body_dofnum[a] == 0) a = m->body_parent[a];
    d->body_lastdof[b] = a > 0 ? m->body_dofadr[a] + m->body_dofnum[a] - 1 : -1;
    for (int j = 0; j < m->body_jntnum[b]; j++) d->jnt_body[m->body_jntadr[b] + j] = b;
  }
  // joints
  for (int j = 0; j < m->njnt; j++) {
    d->jnt_type[j] = m->jnt_type[j]; d->jnt_qposadr[j] = m->jnt_qposadr[j]; d->jnt_dofadr[j] = m->jnt_dofadr[j];
    d->jnt_limited[j] = m->jnt_limited[j] && m->jnt_type[j] != PB_JNT_FREE;
    for (int k = 0; k < 3; k++) { d->jnt_pos[j][k] = (float)m->jnt_pos[j][k]; d->jnt_axis[j][k] = (float)m->jnt_axis[j][k]; }
    d->jnt_range[j][0] = (float)m->jnt_range[j][0]; d->jnt_range[j][1] = (float)m->jnt_range[j][1];
    d->jnt_qpos0[j] = (float)m->qpos0[m->jnt_qposadr[j]];
    d->jnt_stiffness[j] = (float)m->jnt_stiffness[j]; d->jnt_springref[j] = (float)m->jnt_springref[j];
    d->jnt_margin[j] = (float)m->jnt_margin[j];
    kb_from_sol(m->jnt_solref[j], m->jnt_solimp[j], m->timestep, d->jnt_solimp[j], &d->jnt_k[j], &d->jnt_b[j]);
    d->jnt_diag[j] = (float)m->dof_invweight0[m->jnt_dofadr[j]];
  }
  // dofs, sparse M layout
  int nM = 0;
  for (int i = 0; i < m->nv; i++) {
    d->dof_body[i] = m->dof_body[i]; d->dof_parent[i] = m->dof_parent[i];
    d->dof_armature[i] = (float)m->dof_armature[i]; d->dof_damping[i] = (float)m->dof_damping[i];
    d->dof_act[i] = -1;
    d->dof_Madr[i] = nM;
    int depth = 0;
    d->dof_ancmask[i] = 0;
    for (int j = i; j >= 0; j = m->dof_parent[j]) {
      d->dof_ancmask[i] |= 1u << j;
      if (nM >= PB_MAXM) return -3;
      d->M_i[nM] = (unsigned char)i; d->M_j[nM] = (unsigned char)j; nM++; depth++;
    }
    d->dof_depth[i] = depth;  // chain length including self
  }
  d->nM = nM;
  int maxchain = 1;
  for (int i = 0; i < m->nv; i++) { maxchain = std::max(maxchain, d->dof_depth[i]); d->dof_velsrc[i] = m->dof_parent[i]; }
  d->dof_rounds = 0;
  while ((1 << d->dof_rounds) < maxchain) d->dof_rounds++;
  for (int j = 0; j < m->njnt; j++) {
    if (m->jnt_type[j] != PB_JNT_FREE) continue;
    const int da = m->jnt_dofadr[j];
    for (int k = 0; k < 3; k++) d->dof_cdd_zero[da + k] = 1;
    for (int k = 3; k < 6; k++) d->dof_velsrc[da + k] = da + 2;
  }
  for (int u = 0; u < m->nu; u++) {
    int i = m->act_dof[u];
    if (d->dof_act[i] >= 0) return -4;  // one motor per dof in these models
    d->dof_act[i] = u;
    d->dof_gear[i] = (float)m->act_gear[u];
    d->dof_ctrl_lo[i] = m->act_ctrllimited[u] ? (float)m->act_ctrlrange[u][0] : -INFINITY;
    d->dof_ctrl_hi[i] = m->act_ctrllimited[u] ? (float)m->act_ctrlrange[u][1] : INFINITY;
  }
  for (int i = 0; i < m->nq; i++) d->qpos0[i] = (float)m->qpos0[i];
  // factorisation schedule, steps k = nv-1 .. 0 stored in that order
  int nf = 0;
  for (int k = m->nv - 1, step = 0; k >= 0; k--, step++) {
    d->fac_start[step] = nf;
    int c = d->dof_depth[k] - 1;  // number of ancestors
    std::vector<int> anc;
    for (int j = m->dof_parent[k]; j >= 0; j = m->dof_parent[j]) anc.push_back(j);
    for (int a = 0; a < c; a++)
      for (int b = a; b < c; b++) {
        if (nf >= PB_MAXFAC) return -5;
        unsigned tgt = d->dof_Madr[anc[a]] + (b - a), sa = d->dof_Madr[k] + 1 + a, sb = d->dof_Madr[k] + 1 + b;
        d->fac_tab[nf++] = tgt | (sa << 8) | (sb << 16);
      }
  }
  d->fac_start[m->nv] = nf;
  // geoms
  for (int g = 0; g < m->ngeom; g++) {
    d->geom_type[g] = m->geom_type[g]; d->geom_body[g] = m->geom_body[g];
    d->geom_size[g][0] = (float)m->geom_size[g][0]; d->geom_size[g][1] = (float)m->geom_size[g][1];
    double mat[9];
    quat2mat(m->geom_quat[g], mat);
    for (int k = 0; k < 3; k++) { d->geom_pos[g][k] = (float)m->geom_pos[g][k]; d->geom_zaxis[g][k] = (float)mat[3 * k + 2]; }
  }
  for (int p = 0; p < m->npair; p++) {
    d->pair_g1[p] = (unsigned char)m->pair_g1[p]; d->pair_g2[p] = (unsigned char)m->pair_g2[p];
    d->pair_condim[p] = (unsigned char)m->pair_condim[p];
    d->pair_margin[p] = (float)m->pair_margin[p];
    d->pair_incmargin[p] = (float)(m->pair_margin[p] - m->pair_gap[p]);
    d->pair_mu[p] = (float)m->pair_friction[p][0];
    {
      int g1 = m->pair_g1[p], g2 = m->pair_g2[p];
      auto reach = [&](int g) { return m->geom_size[g][0] + (m->geom_type[g] == PB_GEOM_CAPSULE ? m->geom_size[g][1] : 0.0); };
      double r = reach(g1) + reach(g2) + m->pair_margin[p];
      d->pair_reach2[p] = (m->geom_type[g1] == PB_GEOM_PLANE || m->geom_type[g2] == PB_GEOM_PLANE) ? INFINITY : (float)(r * r * (1.0 + 1e-6));
    }
    kb_from_sol(m->pair_solref[p], m->pair_solimp[p], m->timestep, d->pair_solimp[p], &d->pair_k[p], &d->pair_b[p]);
    int b1 = m->geom_body[m->pair_g1[p]], b2 = m->geom_body[m->pair_g2[p]];
    double tran = m->body_invweight0[b1][0] + m->body_invweight0[b2][0], mu = m->pair_friction[p][0];
    d->pair_diag[p] = (float)(m->pair_condim[p] == 1 ? tran : tran + mu * mu * tran);
  }
  return 0;
}
