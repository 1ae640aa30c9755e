/* pbo_physics.c — CPU ORACLE (test infrastructure): FP64 restatement of the MuJoCo 3.2.0 forward dynamics
 * subset the reference reaches through mujoco_env.h:97-108.  See pbo.h for the parity statement (UNPINNED).
 * Stage names follow SURVEY.md §8a rows A4.1-A4.9. */
#include "pbo.h"
#include "../ppo_cpp_b200/csrc/model_tables.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* Test-only switches (oracle/ASSUMPTIONS.md, tools/oracle_sensitivity.py): each bit replaces one behaviour that this file
 * restates from memory by its plausible alternative, so its effect on one env step can be MEASURED.  0 = the restatement. */
int pbo_variant = 0;
enum { PBO_V_WARM_FIXED = 1,      /* qacc_warmstart is not overwritten by the forward evaluations of RK4 stages 1-3 */
       PBO_V_REVERSE_CONTACTS = 2,/* contacts (hence constraint rows, hence the PGS visiting order) in reverse pair order */
       PBO_V_EXACT_DIAG = 4,      /* R from the exact diagonal of J M^-1 J^T instead of the invweight0 approximation */
       PBO_V_DEFAULT_FRAME = 8,   /* plane-capsule contact frames from the default y axis instead of the capsule axis */
       PBO_V_TOL100 = 16,         /* solver tolerance x 100 (stands for "a line search / stopping rule that ends earlier") */
       PBO_V_ITER2X = 32,         /* twice the solver iterations (how far from converged the capped PGS is) */
       PBO_V_NO_EULERDAMP = 64 }; /* explicit joint damping in the Euler step */
static __thread int g_rk_stage = 0;

#define MINVAL 1e-15
#define MINIMP 0.0001
#define MAXIMP 0.9999

/* ------------------------------------------------------------------ small vector helpers */
static double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static void cross3(double *r, const double *a, const double *b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static double norm3(const double *a) { return sqrt(dot3(a, a)); }
static double normalize3(double *a) {
  double n = norm3(a);
  if (n < MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; } else { a[0] /= n; a[1] /= n; a[2] /= n; }
  return n;
}
static void mulmat3vec(double *r, const double *m, const double *v) {
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void quat_mul(double *r, const double *a, const double *b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
static void quat_normalize(double *q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; } else { q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n; }
}
static void quat2mat(double *m, const double *q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
static void axisangle2quat(double *q, const double *axis, double angle) {
  double s = sin(0.5 * angle);
  q[0] = cos(0.5 * angle); q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}

/* spatial 6-vectors are [angular(3), linear(3)] about the kinematic tree's centre of mass (MuJoCo "c-frame") */
static void mul_inert_vec(double *r, const double *i, const double *v) {
  r[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  r[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  r[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  r[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  r[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  r[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
static void cross_motion(double *r, const double *vel, const double *v) {
  double a[3], b[3], c[3];
  cross3(a, vel, v);
  cross3(b, vel, v + 3);
  cross3(c, vel + 3, v);
  for (int k = 0; k < 3; k++) { r[k] = a[k]; r[3 + k] = b[k] + c[k]; }
}
static void cross_force(double *r, const double *vel, const double *f) {
  double a[3], b[3], c[3];
  cross3(a, vel, f);
  cross3(b, vel + 3, f + 3);
  cross3(c, vel, f + 3);
  for (int k = 0; k < 3; k++) { r[k] = a[k] + b[k]; r[3 + k] = c[k]; }
}

const pb_model *pbo_model_by_name(const char *env_id) {
  if (!strcmp(env_id, "HalfCheetah-v5")) return &PB_MODEL_HALFCHEETAH;
  if (!strcmp(env_id, "Hopper-v5")) return &PB_MODEL_HOPPER;
  if (!strcmp(env_id, "Ant-v5")) return &PB_MODEL_ANT;
  if (!strcmp(env_id, "Humanoid-v4")) return &PB_MODEL_HUMANOID;
  return 0;
}

void pbo_reset_data(const pb_model *m, pbo_data *d) {
  int cap_con = d->cap_con, cap_efc = d->cap_efc; /* engine caps are a property of the data object, not of the state */
  memset(d, 0, sizeof(*d));
  d->m = m;
  d->cap_con = cap_con; d->cap_efc = cap_efc;
  memcpy(d->qpos, m->qpos0, sizeof(double) * m->nq);
  d->xquat[0][0] = 1;
  d->xmat[0][0] = d->xmat[0][4] = d->xmat[0][8] = 1;
}

/* ------------------------------------------------------------------ A4.1 kinematics (mj_kinematics) */
static void kinematics(pbo_data *d) {
  const pb_model *m = d->m;
  for (int b = 1; b < m->nbody; b++) {
    int p = m->body_parent[b], ja = m->body_jntadr[b], jn = m->body_jntnum[b];
    double pos[3], quat[4];
    if (jn == 1 && m->jnt_type[ja] == PB_JNT_FREE) {
      int a = m->jnt_qposadr[ja];
      quat_normalize(d->qpos + a + 3); /* in place, as the library does */
      memcpy(pos, d->qpos + a, sizeof(pos));
      memcpy(quat, d->qpos + a + 3, sizeof(quat));
      memcpy(d->xanchor[ja], pos, sizeof(pos));
      d->xaxis[ja][0] = 0; d->xaxis[ja][1] = 0; d->xaxis[ja][2] = 1;
    } else {
      double t[3];
      mulmat3vec(t, d->xmat[p], m->body_pos[b]);
      for (int k = 0; k < 3; k++) pos[k] = d->xpos[p][k] + t[k];
      quat_mul(quat, d->xquat[p], m->body_quat[b]);
      for (int j = ja; j < ja + jn; j++) {
        double mat[9], q = d->qpos[m->jnt_qposadr[j]] - m->qpos0[m->jnt_qposadr[j]];
        quat2mat(mat, quat);
        mulmat3vec(t, mat, m->jnt_pos[j]);
        for (int k = 0; k < 3; k++) d->xanchor[j][k] = t[k] + pos[k];
        mulmat3vec(d->xaxis[j], mat, m->jnt_axis[j]);
        if (m->jnt_type[j] == PB_JNT_SLIDE) {
          for (int k = 0; k < 3; k++) pos[k] += d->xaxis[j][k] * q;
        } else {
          double ql[4], qn[4];
          axisangle2quat(ql, m->jnt_axis[j], q);
          quat_mul(qn, quat, ql);
          memcpy(quat, qn, sizeof(quat));
          quat2mat(mat, quat); /* off-centre rotation: keep the anchor where it was */
          mulmat3vec(t, mat, m->jnt_pos[j]);
          for (int k = 0; k < 3; k++) pos[k] = d->xanchor[j][k] - t[k];
        }
      }
    }
    quat_normalize(quat);
    memcpy(d->xpos[b], pos, sizeof(pos));
    memcpy(d->xquat[b], quat, sizeof(quat));
    quat2mat(d->xmat[b], quat);
    double t[3];
    mulmat3vec(t, d->xmat[b], m->body_ipos[b]);
    for (int k = 0; k < 3; k++) d->xipos[b][k] = d->xpos[b][k] + t[k];
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_body[g];
    double t[3], q[4];
    mulmat3vec(t, d->xmat[b], m->geom_pos[g]);
    for (int k = 0; k < 3; k++) d->gxpos[g][k] = d->xpos[b][k] + t[k];
    quat_mul(q, d->xquat[b], m->geom_quat[g]);
    quat_normalize(q);
    quat2mat(d->gxmat[g], q);
  }
}

/* ------------------------------------------------------------------ A4.1 comPos (subtree_com, cinert, cdof) */
static void com_pos(pbo_data *d) {
  const pb_model *m = d->m;
  double smass[PB_MAXBODY];
  for (int b = 0; b < m->nbody; b++) {
    smass[b] = m->body_mass[b];
    for (int k = 0; k < 3; k++) d->subtree_com[b][k] = m->body_mass[b] * d->xipos[b][k];
  }
  for (int b = m->nbody - 1; b > 0; b--) {
    int p = m->body_parent[b];
    smass[p] += smass[b];
    for (int k = 0; k < 3; k++) d->subtree_com[p][k] += d->subtree_com[b][k];
  }
  for (int b = 0; b < m->nbody; b++)
    for (int k = 0; k < 3; k++)
      d->subtree_com[b][k] = smass[b] < MINVAL ? d->xipos[b][k] : d->subtree_com[b][k] / smass[b];

  memset(d->cinert[0], 0, sizeof(d->cinert[0]));
  for (int b = 1; b < m->nbody; b++) {
    const double *R = d->xmat[b], *I = m->body_inertia[b], *com = d->subtree_com[m->body_root[b]];
    double RI[9], W[9], dif[3], mass = m->body_mass[b];
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) RI[3 * r + c] = R[3 * r] * I[c] + R[3 * r + 1] * I[3 + c] + R[3 * r + 2] * I[6 + c];
    for (int r = 0; r < 3; r++)
      for (int c = 0; c < 3; c++) W[3 * r + c] = RI[3 * r] * R[3 * c] + RI[3 * r + 1] * R[3 * c + 1] + RI[3 * r + 2] * R[3 * c + 2];
    for (int k = 0; k < 3; k++) dif[k] = d->xipos[b][k] - com[k];
    double dd = dot3(dif, dif), *ci = d->cinert[b];
    ci[0] = W[0] + mass * (dd - dif[0] * dif[0]);
    ci[1] = W[4] + mass * (dd - dif[1] * dif[1]);
    ci[2] = W[8] + mass * (dd - dif[2] * dif[2]);
    ci[3] = W[1] - mass * dif[0] * dif[1];
    ci[4] = W[2] - mass * dif[0] * dif[2];
    ci[5] = W[5] - mass * dif[1] * dif[2];
    ci[6] = mass * dif[0]; ci[7] = mass * dif[1]; ci[8] = mass * dif[2];
    ci[9] = mass;
  }
  for (int j = 0; j < m->njnt; j++) {
    int b = m->jnt_body[j], da = m->jnt_dofadr[j];
    double off[3];
    for (int k = 0; k < 3; k++) off[k] = d->subtree_com[m->body_root[b]][k] - d->xanchor[j][k];
    if (m->jnt_type[j] == PB_JNT_FREE) {
      for (int k = 0; k < 3; k++) {
        memset(d->cdof[da + k], 0, sizeof(d->cdof[0]));
        d->cdof[da + k][3 + k] = 1;
        double ax[3] = {d->xmat[b][k], d->xmat[b][3 + k], d->xmat[b][6 + k]};
        memcpy(d->cdof[da + 3 + k], ax, sizeof(ax));
        cross3(d->cdof[da + 3 + k] + 3, ax, off);
      }
    } else if (m->jnt_type[j] == PB_JNT_SLIDE) {
      memset(d->cdof[da], 0, sizeof(d->cdof[0]));
      memcpy(d->cdof[da] + 3, d->xaxis[j], sizeof(double) * 3);
    } else {
      memcpy(d->cdof[da], d->xaxis[j], sizeof(double) * 3);
      cross3(d->cdof[da] + 3, d->xaxis[j], off);
    }
  }
}

/* ------------------------------------------------------------------ A4.1 CRB mass matrix + factorisation */
static int chol_factor(int n, double A[][PB_MAXDOF], double L[][PB_MAXDOF]) {
  for (int i = 0; i < n; i++)
    for (int j = 0; j <= i; j++) {
      double s = A[i][j];
      for (int k = 0; k < j; k++) s -= L[i][k] * L[j][k];
      if (i == j) { if (s < MINVAL) s = MINVAL; L[i][i] = sqrt(s); } else L[i][j] = s / L[j][j];
    }
  return 0;
}
static void chol_solve(int n, double L[][PB_MAXDOF], double *x) {
  for (int i = 0; i < n; i++) {
    double s = x[i];
    for (int k = 0; k < i; k++) s -= L[i][k] * x[k];
    x[i] = s / L[i][i];
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = x[i];
    for (int k = i + 1; k < n; k++) s -= L[k][i] * x[k];
    x[i] = s / L[i][i];
  }
}

static void crb(pbo_data *d) {
  const pb_model *m = d->m;
  double c[PB_MAXBODY][10];
  memcpy(c, d->cinert, sizeof(c));
  for (int b = m->nbody - 1; b > 0; b--) {
    int p = m->body_parent[b];
    if (p > 0) for (int k = 0; k < 10; k++) c[p][k] += c[b][k];
  }
  memset(d->M, 0, sizeof(d->M));
  for (int i = 0; i < m->nv; i++) {
    double buf[6];
    mul_inert_vec(buf, c[m->dof_body[i]], d->cdof[i]);
    for (int j = i; j >= 0; j = m->dof_parent[j]) {
      double s = 0;
      for (int k = 0; k < 6; k++) s += d->cdof[j][k] * buf[k];
      d->M[i][j] = d->M[j][i] = s;
    }
    d->M[i][i] += m->dof_armature[i];
  }
  chol_factor(m->nv, d->M, d->L);
}

/* ------------------------------------------------------------------ A4.2 collision (static pair list, narrowphase) */
static void make_frame(double *f) {
  normalize3(f);
  if (norm3(f + 3) < 0.5) {
    f[3] = f[4] = f[5] = 0;
    if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1;
  }
  double t = dot3(f, f + 3);
  for (int k = 0; k < 3; k++) f[3 + k] -= t * f[k];
  normalize3(f + 3);
  cross3(f + 6, f, f + 3);
}

static int add_contact(pbo_data *d, int pair, double dist, const double *pos, const double *normal,
                       const double *yaxis) {
  const pb_model *m = d->m;
  if (d->ncon >= (d->cap_con > 0 && d->cap_con < PBO_MAXCON ? d->cap_con : PBO_MAXCON)) { d->overflow = 1; return 0; }
  pbo_contact *c = &d->contact[d->ncon++];
  memset(c, 0, sizeof(*c));
  c->dist = dist;
  memcpy(c->pos, pos, sizeof(c->pos));
  memcpy(c->frame, normal, sizeof(double) * 3);
  if (yaxis) memcpy(c->frame + 3, yaxis, sizeof(double) * 3);
  make_frame(c->frame);
  c->includemargin = m->pair_margin[pair] - m->pair_gap[pair];
  memcpy(c->friction, m->pair_friction[pair], sizeof(c->friction));
  memcpy(c->solref, m->pair_solref[pair], sizeof(c->solref));
  memcpy(c->solimp, m->pair_solimp[pair], sizeof(c->solimp));
  c->dim = m->pair_condim[pair];
  c->g1 = m->pair_g1[pair];
  c->g2 = m->pair_g2[pair];
  c->efc_address = -1;
  return 1;
}

/* sphere (p1,r1) vs sphere (p2,r2); normal from 1 to 2 */
static void sphere_sphere(pbo_data *d, int pair, double margin, const double *p1, double r1, const double *p2,
                          double r2, const double *yaxis) {
  double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  double cd = norm3(dif), dist = cd - r1 - r2;
  if (dist > margin) return;
  double n[3] = {1, 0, 0}, pos[3];
  if (cd >= MINVAL) for (int k = 0; k < 3; k++) n[k] = dif[k] / cd;
  for (int k = 0; k < 3; k++) pos[k] = p1[k] + n[k] * (r1 + 0.5 * dist);
  add_contact(d, pair, dist, pos, n, yaxis);
}

static void plane_sphere(pbo_data *d, int pair, double margin, int gp, const double *sp, double r,
                         const double *yaxis) {
  double n[3] = {d->gxmat[gp][2], d->gxmat[gp][5], d->gxmat[gp][8]};
  double dif[3] = {sp[0] - d->gxpos[gp][0], sp[1] - d->gxpos[gp][1], sp[2] - d->gxpos[gp][2]};
  double dist = dot3(dif, n) - r;
  if (dist > margin) return;
  double pos[3];
  for (int k = 0; k < 3; k++) pos[k] = sp[k] - n[k] * (r + 0.5 * dist);
  add_contact(d, pair, dist, pos, n, yaxis);
}

static double clampd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }

static void collide_pair(pbo_data *d, int pair) {
  const pb_model *m = d->m;
  int g1 = m->pair_g1[pair], g2 = m->pair_g2[pair], t1 = m->geom_type[g1], t2 = m->geom_type[g2];
  double margin = m->pair_margin[pair];
  const double *p1 = d->gxpos[g1], *p2 = d->gxpos[g2], *s1 = m->geom_size[g1], *s2 = m->geom_size[g2];
  double ax1[3] = {d->gxmat[g1][2], d->gxmat[g1][5], d->gxmat[g1][8]};
  double ax2[3] = {d->gxmat[g2][2], d->gxmat[g2][5], d->gxmat[g2][8]};
  if (t1 == PB_GEOM_PLANE && t2 == PB_GEOM_SPHERE) {
    plane_sphere(d, pair, margin, g1, p2, s2[0], 0);
  } else if (t1 == PB_GEOM_PLANE && t2 == PB_GEOM_CAPSULE) {
    double e[3];
    for (int k = 0; k < 3; k++) e[k] = p2[k] + ax2[k] * s2[1];
    const double *yax = (pbo_variant & PBO_V_DEFAULT_FRAME) ? 0 : ax2;
    plane_sphere(d, pair, margin, g1, e, s2[0], yax); /* frame y-axis aligned with the capsule axis */
    for (int k = 0; k < 3; k++) e[k] = p2[k] - ax2[k] * s2[1];
    plane_sphere(d, pair, margin, g1, e, s2[0], yax);
  } else if (t1 == PB_GEOM_SPHERE && t2 == PB_GEOM_SPHERE) {
    sphere_sphere(d, pair, margin, p1, s1[0], p2, s2[0], 0);
  } else if (t1 == PB_GEOM_SPHERE && t2 == PB_GEOM_CAPSULE) {
    double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]}, v[3];
    double x = clampd(dot3(ax2, dif), -s2[1], s2[1]);
    for (int k = 0; k < 3; k++) v[k] = p2[k] + ax2[k] * x;
    sphere_sphere(d, pair, margin, p1, s1[0], v, s2[0], 0);
  } else if (t1 == PB_GEOM_CAPSULE && t2 == PB_GEOM_CAPSULE) {
    /* closest points of the two axis segments (non-parallel case; near-parallel falls back to a clamped
       single contact — MuJoCo emits up to two contacts for exactly parallel axes, a measure-zero case) */
    double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
    double mb = -dot3(ax1, ax2), u = -dot3(ax1, dif), v = dot3(ax2, dif), det = 1.0 - mb * mb;
    double x1, x2;
    if (fabs(det) >= MINVAL) {
      x1 = (u - mb * v) / det;
      x2 = (v - mb * u) / det;
    } else {
      x1 = 0; x2 = v;
    }
    if (x1 > s1[1]) { x1 = s1[1]; x2 = v - mb * x1; }
    else if (x1 < -s1[1]) { x1 = -s1[1]; x2 = v - mb * x1; }
    if (x2 > s2[1]) { x2 = s2[1]; x1 = clampd(u - mb * x2, -s1[1], s1[1]); }
    else if (x2 < -s2[1]) { x2 = -s2[1]; x1 = clampd(u - mb * x2, -s1[1], s1[1]); }
    double v1[3], v2[3];
    for (int k = 0; k < 3; k++) { v1[k] = p1[k] + ax1[k] * x1; v2[k] = p2[k] + ax2[k] * x2; }
    sphere_sphere(d, pair, margin, v1, s1[0], v2, s2[0], 0);
  }
}

static void collision(pbo_data *d) {
  d->ncon = 0;
  d->overflow = 0;
  for (int p = 0; p < d->m->npair; p++) collide_pair(d, p);
  if (pbo_variant & PBO_V_REVERSE_CONTACTS)
    for (int i = 0, j = d->ncon - 1; i < j; i++, j--) { pbo_contact t = d->contact[i]; d->contact[i] = d->contact[j]; d->contact[j] = t; }
}

/* ------------------------------------------------------------------ A4.3 constraint rows */
static void jac_point(const pbo_data *d, int body, const double *point, double jp[][3]) {
  const pb_model *m = d->m;
  for (int i = 0; i < m->nv; i++) jp[i][0] = jp[i][1] = jp[i][2] = 0;
  for (int b = body; b > 0; b = m->body_parent[b]) {
    for (int j = m->body_jntadr[b]; j < m->body_jntadr[b] + m->body_jntnum[b]; j++) {
      int da = m->jnt_dofadr[j];
      double off[3] = {point[0] - d->xanchor[j][0], point[1] - d->xanchor[j][1], point[2] - d->xanchor[j][2]};
      if (m->jnt_type[j] == PB_JNT_FREE) {
        for (int k = 0; k < 3; k++) {
          jp[da + k][k] = 1;
          double ax[3] = {d->xmat[b][k], d->xmat[b][3 + k], d->xmat[b][6 + k]};
          cross3(jp[da + 3 + k], ax, off);
        }
      } else if (m->jnt_type[j] == PB_JNT_SLIDE) {
        memcpy(jp[da], d->xaxis[j], sizeof(double) * 3);
      } else {
        cross3(jp[da], d->xaxis[j], off);
      }
    }
  }
}

static int add_row(pbo_data *d, double pos, double margin, double diag) {
  int i = d->nefc;
  if (i >= PBO_MAXEFC) return -1;
  memset(d->efc_J[i], 0, sizeof(d->efc_J[i]));
  d->efc_pos[i] = pos;
  d->efc_margin[i] = margin;
  d->efc_diagApprox[i] = diag;
  d->nefc++;
  return i;
}

static void impedance(const double *solimp_in, const double *solref_in, double timestep, double pos, double margin,
                      double *imp, double *k, double *b) {
  double s[5], ref0 = solref_in[0], ref1 = solref_in[1];
  s[0] = clampd(solimp_in[0], MINIMP, MAXIMP);
  s[1] = clampd(solimp_in[1], MINIMP, MAXIMP);
  s[2] = solimp_in[2] < 0 ? 0 : solimp_in[2];
  s[3] = clampd(solimp_in[3], MINIMP, MAXIMP);
  s[4] = solimp_in[4] < 1 ? 1 : solimp_in[4];
  if (ref0 > 0 && ref0 < 2 * timestep) ref0 = 2 * timestep; /* refsafe */
  if (s[0] == s[1] || s[2] <= MINVAL) {
    *imp = 0.5 * (s[0] + s[1]);
  } else {
    double x = fabs((pos - margin) / s[2]);
    if (x >= 1) *imp = s[1];
    else if (x <= 0) *imp = s[0];
    else {
      double y;
      if (s[4] == 1) y = x;
      else if (x <= s[3]) y = pow(x, s[4]) / pow(s[3], s[4] - 1);
      else y = 1 - pow(1 - x, s[4]) / pow(1 - s[3], s[4] - 1);
      *imp = s[0] + y * (s[1] - s[0]);
    }
  }
  if (ref0 > 0) {
    *k = 1.0 / (s[1] * s[1] * ref0 * ref0 * ref1 * ref1);
    *b = 2.0 / (s[1] * ref0);
  } else {
    *k = -ref0 / (s[1] * s[1]);
    *b = -ref1 / s[1];
  }
}

static void make_constraint(pbo_data *d) {
  const pb_model *m = d->m;
  d->nefc = 0;
  /* joint limits, lower side first (hinge/slide only in these models) */
  for (int j = 0; j < m->njnt; j++) {
    if (!m->jnt_limited[j] || m->jnt_type[j] == PB_JNT_FREE) continue;
    double value = d->qpos[m->jnt_qposadr[j]];
    for (int side = -1; side <= 1; side += 2) {
      double dist = side * (m->jnt_range[j][(side + 1) / 2] - value);
      if (dist < m->jnt_margin[j]) {
        int r = add_row(d, dist, m->jnt_margin[j], m->dof_invweight0[m->jnt_dofadr[j]]);
        if (r < 0) continue;
        d->efc_J[r][m->jnt_dofadr[j]] = -side;
        impedance(m->jnt_solimp[j], m->jnt_solref[j], m->timestep, dist, m->jnt_margin[j], &d->efc_imp[r],
                  &d->efc_k[r], &d->efc_bdamp[r]);
        d->efc_R[r] = fmax(MINVAL, (1 - d->efc_imp[r]) * d->efc_diagApprox[r] / d->efc_imp[r]);
      }
    }
  }
  /* contacts: frictionless (condim 1) -> 1 row, pyramidal condim 3 -> 4 rows */
  for (int c = 0; c < d->ncon; c++) {
    pbo_contact *con = &d->contact[c];
    if (con->dist >= con->includemargin) continue;
    int b1 = m->geom_body[con->g1], b2 = m->geom_body[con->g2];
    double j1[PB_MAXDOF][3], j2[PB_MAXDOF][3], jd[3][PB_MAXDOF];
    jac_point(d, b1, con->pos, j1);
    jac_point(d, b2, con->pos, j2);
    for (int i = 0; i < m->nv; i++) {
      double df[3] = {j2[i][0] - j1[i][0], j2[i][1] - j1[i][1], j2[i][2] - j1[i][2]};
      for (int k = 0; k < 3; k++) jd[k][i] = dot3(con->frame + 3 * k, df);
    }
    double tran = m->body_invweight0[b1][0] + m->body_invweight0[b2][0];
    int nrow = con->dim == 1 ? 1 : 2 * (con->dim - 1), first = d->nefc;
    if (first + nrow > (d->cap_efc > 0 && d->cap_efc < PBO_MAXEFC ? d->cap_efc : PBO_MAXEFC)) { d->overflow = 1; break; }
    con->efc_address = first;
    for (int r = 0; r < nrow; r++) {
      double fri = con->dim == 1 ? 0.0 : con->friction[r / 2 < 2 ? 0 : 1];
      int row = add_row(d, con->dist, con->includemargin, tran + fri * fri * tran);
      for (int i = 0; i < m->nv; i++)
        d->efc_J[row][i] = con->dim == 1 ? jd[0][i] : jd[0][i] + ((r & 1) ? -fri : fri) * jd[1 + r / 2][i];
      if (pbo_variant & PBO_V_EXACT_DIAG) {
        double x[PB_MAXDOF], a = 0;
        memcpy(x, d->efc_J[row], sizeof(double) * m->nv);
        chol_solve(m->nv, d->L, x);
        for (int i = 0; i < m->nv; i++) a += d->efc_J[row][i] * x[i];
        d->efc_diagApprox[row] = a;
      }
      impedance(con->solimp, con->solref, m->timestep, con->dist, con->includemargin, &d->efc_imp[row],
                &d->efc_k[row], &d->efc_bdamp[row]);
      d->efc_R[row] = fmax(MINVAL, (1 - d->efc_imp[row]) * d->efc_diagApprox[row] / d->efc_imp[row]);
    }
    if (con->dim > 1) { /* pyramidal: all rows share Rpy = 2 mu^2 R(first row) */
      double rpy = 2 * con->friction[0] * con->friction[0] * d->efc_R[first];
      for (int r = 0; r < nrow; r++) d->efc_R[first + r] = fmax(MINVAL, rpy);
    }
  }
  for (int i = 0; i < d->nefc; i++) d->efc_D[i] = 1.0 / d->efc_R[i];
}

/* ------------------------------------------------------------------ A4.4 velocity stage */
static void com_vel(pbo_data *d) {
  const pb_model *m = d->m;
  memset(d->cvel[0], 0, sizeof(d->cvel[0]));
  for (int b = 1; b < m->nbody; b++) {
    double cv[6];
    memcpy(cv, d->cvel[m->body_parent[b]], sizeof(cv));
    for (int j = m->body_jntadr[b]; j < m->body_jntadr[b] + m->body_jntnum[b]; j++) {
      int da = m->jnt_dofadr[j];
      if (m->jnt_type[j] == PB_JNT_FREE) {
        for (int k = 0; k < 3; k++) memset(d->cdof_dot[da + k], 0, sizeof(d->cdof_dot[0]));
        for (int k = 0; k < 3; k++)
          for (int c = 0; c < 6; c++) cv[c] += d->cdof[da + k][c] * d->qvel[da + k];
        for (int k = 3; k < 6; k++) cross_motion(d->cdof_dot[da + k], cv, d->cdof[da + k]);
        for (int k = 3; k < 6; k++)
          for (int c = 0; c < 6; c++) cv[c] += d->cdof[da + k][c] * d->qvel[da + k];
      } else {
        cross_motion(d->cdof_dot[da], cv, d->cdof[da]);
        for (int c = 0; c < 6; c++) cv[c] += d->cdof[da][c] * d->qvel[da];
      }
    }
    memcpy(d->cvel[b], cv, sizeof(cv));
  }
}

static void passive(pbo_data *d) {
  const pb_model *m = d->m;
  for (int i = 0; i < m->nv; i++) d->qfrc_passive[i] = -m->dof_damping[i] * d->qvel[i];
  for (int j = 0; j < m->njnt; j++) {
    if (m->jnt_type[j] == PB_JNT_FREE || m->jnt_stiffness[j] == 0) continue;
    d->qfrc_passive[m->jnt_dofadr[j]] -= m->jnt_stiffness[j] * (d->qpos[m->jnt_qposadr[j]] - m->jnt_springref[j]);
  }
}

/* RNE with optional acceleration; out[nv] = cdof . (accumulated body forces) */
static void rne(pbo_data *d, int flg_acc, double *out) {
  const pb_model *m = d->m;
  double cacc[PB_MAXBODY][6], cfrc[PB_MAXBODY][6];
  memset(cacc, 0, sizeof(cacc));
  for (int k = 0; k < 3; k++) cacc[0][3 + k] = -m->gravity[k];
  memset(cfrc[0], 0, sizeof(cfrc[0]));
  for (int b = 1; b < m->nbody; b++) {
    memcpy(cacc[b], cacc[m->body_parent[b]], sizeof(cacc[b]));
    for (int i = m->body_dofadr[b]; i >= 0 && i < m->body_dofadr[b] + m->body_dofnum[b]; i++)
      for (int c = 0; c < 6; c++)
        cacc[b][c] += d->cdof_dot[i][c] * d->qvel[i] + (flg_acc ? d->cdof[i][c] * d->qacc[i] : 0.0);
    double t[6], t1[6];
    mul_inert_vec(cfrc[b], d->cinert[b], cacc[b]);
    mul_inert_vec(t, d->cinert[b], d->cvel[b]);
    cross_force(t1, d->cvel[b], t);
    for (int c = 0; c < 6; c++) cfrc[b][c] += t1[c];
  }
  for (int b = m->nbody - 1; b > 0; b--)
    for (int c = 0; c < 6; c++) cfrc[m->body_parent[b]][c] += cfrc[b][c];
  for (int i = 0; i < m->nv; i++) {
    double s = 0;
    for (int c = 0; c < 6; c++) s += d->cdof[i][c] * cfrc[m->dof_body[i]][c];
    out[i] = s;
  }
}

static void reference_constraint(pbo_data *d) {
  const pb_model *m = d->m;
  for (int r = 0; r < d->nefc; r++) {
    double v = 0;
    for (int i = 0; i < m->nv; i++) v += d->efc_J[r][i] * d->qvel[i];
    d->efc_vel[r] = v;
    d->efc_aref[r] = -d->efc_k[r] * d->efc_imp[r] * (d->efc_pos[r] - d->efc_margin[r]) - d->efc_bdamp[r] * v;
  }
}

/* ------------------------------------------------------------------ A4.5 actuation + smooth acceleration */
static void actuation(pbo_data *d) {
  const pb_model *m = d->m;
  memset(d->qfrc_actuator, 0, sizeof(d->qfrc_actuator));
  for (int u = 0; u < m->nu; u++) {
    double c = d->ctrl[u];
    if (m->act_ctrllimited[u]) c = clampd(c, m->act_ctrlrange[u][0], m->act_ctrlrange[u][1]);
    d->qfrc_actuator[m->act_dof[u]] += m->act_gear[u] * c;
  }
}

static void acceleration(pbo_data *d) {
  const pb_model *m = d->m;
  for (int i = 0; i < m->nv; i++) {
    d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_actuator[i];
    d->qacc_smooth[i] = d->qfrc_smooth[i];
  }
  chol_solve(m->nv, d->L, d->qacc_smooth);
}

/* ------------------------------------------------------------------ A4.6 constraint solvers */
/* all rows in these models are scalar inequalities: s(jar) = 0.5 D jar^2 if jar < 0 else 0 */
static double constraint_update(pbo_data *d, const double *jar, double *force) {
  double cost = 0;
  for (int r = 0; r < d->nefc; r++) {
    if (jar[r] < 0) { force[r] = -d->efc_D[r] * jar[r]; cost += 0.5 * d->efc_D[r] * jar[r] * jar[r]; }
    else force[r] = 0;
  }
  return cost;
}

static void mul_M(const pbo_data *d, double *r, const double *v) {
  int nv = d->m->nv;
  for (int i = 0; i < nv; i++) {
    double s = 0;
    for (int j = 0; j < nv; j++) s += d->M[i][j] * v[j];
    r[i] = s;
  }
}

static void solve_pgs(pbo_data *d) {
  const pb_model *m = d->m;
  int nv = m->nv, ne = d->nefc;
  static __thread double AR[PBO_MAXEFC][PBO_MAXEFC], MinvJT[PBO_MAXEFC][PB_MAXDOF];
  double jar[PBO_MAXEFC], *f = d->efc_force;
  for (int r = 0; r < ne; r++) {
    memcpy(MinvJT[r], d->efc_J[r], sizeof(double) * nv);
    chol_solve(nv, d->L, MinvJT[r]);
  }
  for (int r = 0; r < ne; r++)
    for (int c = 0; c < ne; c++) {
      double s = 0;
      for (int i = 0; i < nv; i++) s += d->efc_J[r][i] * MinvJT[c][i];
      AR[r][c] = s + (r == c ? d->efc_R[r] : 0.0);
    }
  /* warm start: forces implied by qacc_warmstart; fall back to zero if the dual cost is positive */
  for (int r = 0; r < ne; r++) {
    double s = -d->efc_aref[r];
    for (int i = 0; i < nv; i++) s += d->efc_J[r][i] * d->qacc_warmstart[i];
    jar[r] = s;
  }
  constraint_update(d, jar, f);
  double cost = 0;
  for (int r = 0; r < ne; r++) {
    double s = 0;
    for (int c = 0; c < ne; c++) s += AR[r][c] * f[c];
    cost += f[r] * (0.5 * s + d->efc_b[r]);
  }
  if (cost > 0) memset(f, 0, sizeof(double) * ne);
  double scale = 1.0 / (m->meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  const int maxit = m->iterations * ((pbo_variant & PBO_V_ITER2X) ? 2 : 1);
  const double tol = m->tolerance * ((pbo_variant & PBO_V_TOL100) ? 100.0 : 1.0);
  while (iter < maxit) {
    double improvement = 0;
    for (int r = 0; r < ne; r++) {
      double res = d->efc_b[r];
      for (int c = 0; c < ne; c++) res += AR[r][c] * f[c];
      double old = f[r];
      f[r] -= res / AR[r][r];
      if (f[r] < 0) f[r] = 0;
      double delta = f[r] - old;
      improvement -= 0.5 * delta * delta * AR[r][r] + delta * res;
    }
    iter++;
    if (improvement * scale < tol) break;
  }
  d->solver_niter = iter;
  for (int i = 0; i < nv; i++) {
    double s = 0;
    for (int r = 0; r < ne; r++) s += d->efc_J[r][i] * f[r];
    d->qfrc_constraint[i] = s;
    d->qacc[i] = s;
  }
  chol_solve(nv, d->L, d->qacc);
  for (int i = 0; i < nv; i++) d->qacc[i] += d->qacc_smooth[i];
}

typedef struct { double t; int row; } brk_t;
static int brk_cmp(const void *a, const void *b) {
  double x = ((const brk_t *)a)->t, y = ((const brk_t *)b)->t;
  return x < y ? -1 : (x > y ? 1 : 0);
}

/* primal Newton with an exact one-dimensional line search (the objective is a convex piecewise quadratic, so
 * the minimiser along a direction is found by walking its sorted breakpoints).  MuJoCo's own line search is
 * iterative with tolerance ls_tolerance; both converge to the unique minimiser within opt.tolerance. */
static void solve_newton(pbo_data *d) {
  const pb_model *m = d->m;
  int nv = m->nv, ne = d->nefc;
  double qacc[PB_MAXDOF], Ma[PB_MAXDOF], jar[PBO_MAXEFC], grad[PB_MAXDOF], search[PB_MAXDOF];
  double Mv[PB_MAXDOF], jv[PBO_MAXEFC], *f = d->efc_force, tmp[PB_MAXDOF];
  double H[PB_MAXDOF][PB_MAXDOF], HL[PB_MAXDOF][PB_MAXDOF];
  double scale = 1.0 / (m->meaninertia * (nv > 1 ? nv : 1));
  /* warm start: best of qacc_warmstart and qacc_smooth */
  double cost_ws, cost_sm;
  {
    for (int r = 0; r < ne; r++) {
      double s = -d->efc_aref[r];
      for (int i = 0; i < nv; i++) s += d->efc_J[r][i] * d->qacc_warmstart[i];
      jar[r] = s;
    }
    cost_ws = constraint_update(d, jar, f);
    mul_M(d, Ma, d->qacc_warmstart);
    for (int i = 0; i < nv; i++) cost_ws += 0.5 * (Ma[i] - d->qfrc_smooth[i]) * (d->qacc_warmstart[i] - d->qacc_smooth[i]);
    cost_sm = constraint_update(d, d->efc_b, f);
    memcpy(qacc, cost_ws > cost_sm ? d->qacc_smooth : d->qacc_warmstart, sizeof(double) * nv);
  }
  double cost = 0;
  int iter = 0;
  for (;;) {
    mul_M(d, Ma, qacc);
    for (int r = 0; r < ne; r++) {
      double s = -d->efc_aref[r];
      for (int i = 0; i < nv; i++) s += d->efc_J[r][i] * qacc[i];
      jar[r] = s;
    }
    double newcost = constraint_update(d, jar, f);
    for (int i = 0; i < nv; i++) newcost += 0.5 * (Ma[i] - d->qfrc_smooth[i]) * (qacc[i] - d->qacc_smooth[i]);
    double gnorm = 0;
    for (int i = 0; i < nv; i++) {
      double s = Ma[i] - d->qfrc_smooth[i];
      for (int r = 0; r < ne; r++) s -= d->efc_J[r][i] * f[r];
      grad[i] = s;
      gnorm += s * s;
    }
    gnorm = sqrt(gnorm);
    const double tol = m->tolerance * ((pbo_variant & PBO_V_TOL100) ? 100.0 : 1.0);
    if (iter > 0 && (scale * (cost - newcost) < tol || scale * gnorm < tol)) { cost = newcost; break; }
    cost = newcost;
    if (iter >= m->iterations * ((pbo_variant & PBO_V_ITER2X) ? 2 : 1) || scale * gnorm < 1e-300) break;
    /* Hessian of the active set and Newton direction */
    for (int i = 0; i < nv; i++)
      for (int j = 0; j < nv; j++) {
        double s = d->M[i][j];
        for (int r = 0; r < ne; r++) if (jar[r] < 0) s += d->efc_D[r] * d->efc_J[r][i] * d->efc_J[r][j];
        H[i][j] = s;
      }
    chol_factor(nv, H, HL);
    for (int i = 0; i < nv; i++) search[i] = -grad[i];
    chol_solve(nv, HL, search);
    /* exact line search along `search` */
    mul_M(d, Mv, search);
    for (int r = 0; r < ne; r++) {
      double s = 0;
      for (int i = 0; i < nv; i++) s += d->efc_J[r][i] * search[i];
      jv[r] = s;
    }
    double c0 = 0, c1 = 0; /* derivative = c0 + c1*alpha on the current piece */
    for (int i = 0; i < nv; i++) { c0 += search[i] * grad[i]; c1 += search[i] * Mv[i]; }
    /* grad already contains the active rows' contribution at alpha=0; split it out to track activity */
    brk_t brk[PBO_MAXEFC];
    int nb = 0;
    for (int r = 0; r < ne; r++) {
      int active = jar[r] < 0 || (jar[r] == 0 && jv[r] < 0);
      if (active) c1 += d->efc_D[r] * jv[r] * jv[r];
      if (jv[r] != 0) {
        double t = -jar[r] / jv[r];
        if (t > 0) { brk[nb].t = t; brk[nb].row = r; nb++; }
      }
    }
    qsort(brk, nb, sizeof(brk_t), brk_cmp);
    double alpha = 0;
    {
      double a = 0; /* current position; derivative at a is c0 + c1*a with piece-local (c0,c1) */
      int k = 0;
      for (;;) {
        double root = c1 > 0 ? -c0 / c1 : 1e300;
        double next = k < nb ? brk[k].t : 1e300;
        if (root <= next) { alpha = root < a ? a : root; break; }
        if (k >= nb) { alpha = next; break; }
        int r = brk[k].row;
        double D = d->efc_D[r], sgn = jv[r] > 0 ? -1.0 : 1.0; /* jv>0: row leaves the active set at t */
        c0 += sgn * D * jar[r] * jv[r];
        c1 += sgn * D * jv[r] * jv[r];
        a = next;
        k++;
      }
    }
    for (int i = 0; i < nv; i++) qacc[i] += alpha * search[i];
    (void)tmp;
    iter++;
  }
  d->solver_niter = iter;
  memcpy(d->qacc, qacc, sizeof(double) * nv);
  for (int i = 0; i < nv; i++) {
    double s = 0;
    for (int r = 0; r < ne; r++) s += d->efc_J[r][i] * f[r];
    d->qfrc_constraint[i] = s;
  }
}

static void fwd_constraint(pbo_data *d) {
  const pb_model *m = d->m;
  int nv = m->nv;
  if (d->nefc == 0) {
    memcpy(d->qacc, d->qacc_smooth, sizeof(double) * nv);
    memcpy(d->qacc_warmstart, d->qacc_smooth, sizeof(double) * nv);
    memset(d->qfrc_constraint, 0, sizeof(d->qfrc_constraint));
    d->solver_niter = 0;
    return;
  }
  for (int r = 0; r < d->nefc; r++) {
    double s = -d->efc_aref[r];
    for (int i = 0; i < nv; i++) s += d->efc_J[r][i] * d->qacc_smooth[i];
    d->efc_b[r] = s;
  }
  if (m->solver == PB_SOL_PGS) solve_pgs(d); else solve_newton(d);
  if (!((pbo_variant & PBO_V_WARM_FIXED) && g_rk_stage)) memcpy(d->qacc_warmstart, d->qacc, sizeof(double) * nv);
}

/* ------------------------------------------------------------------ mj_forward */
void pbo_forward(pbo_data *d) {
  kinematics(d);
  com_pos(d);
  crb(d);
  collision(d);
  make_constraint(d);
  com_vel(d);
  passive(d);
  reference_constraint(d);
  rne(d, 0, d->qfrc_bias);
  actuation(d);
  acceleration(d);
  fwd_constraint(d);
  if (d->trace_n < 64) { d->trace_nefc[d->trace_n] = (short)d->nefc; d->trace_niter[d->trace_n] = (short)d->solver_niter; d->trace_n++; }
  d->n_forward++;
}

/* ------------------------------------------------------------------ A4.7 integrators */
static void integrate_pos(const pb_model *m, double *qpos, const double *qvel, double h) {
  for (int j = 0; j < m->njnt; j++) {
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    if (m->jnt_type[j] == PB_JNT_FREE) {
      for (int k = 0; k < 3; k++) qpos[qa + k] += h * qvel[da + k];
      double w[3] = {qvel[da + 3], qvel[da + 4], qvel[da + 5]};
      double ang = h * normalize3(w); /* normalize3 returns the original norm */
      double ql[4], qn[4];
      axisangle2quat(ql, w, ang);
      quat_normalize(qpos + qa + 3);
      quat_mul(qn, qpos + qa + 3, ql);
      quat_normalize(qn);
      memcpy(qpos + qa + 3, qn, sizeof(qn));
    } else {
      qpos[qa] += h * qvel[da];
    }
  }
}

static void euler(pbo_data *d) {
  const pb_model *m = d->m;
  int nv = m->nv, damped = 0;
  double qacc[PB_MAXDOF];
  memcpy(qacc, d->qacc, sizeof(qacc));
  for (int i = 0; i < nv; i++) if (m->dof_damping[i] > 0) damped = 1;
  if (damped && !(pbo_variant & PBO_V_NO_EULERDAMP)) { /* eulerdamp: (M + h diag(B)) qacc = qfrc_smooth + qfrc_constraint */
    double A[PB_MAXDOF][PB_MAXDOF], AL[PB_MAXDOF][PB_MAXDOF];
    memcpy(A, d->M, sizeof(A));
    for (int i = 0; i < nv; i++) {
      A[i][i] += m->timestep * m->dof_damping[i];
      qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i];
    }
    chol_factor(nv, A, AL);
    chol_solve(nv, AL, qacc);
  }
  for (int i = 0; i < nv; i++) d->qvel[i] += m->timestep * qacc[i];
  integrate_pos(m, d->qpos, d->qvel, m->timestep);
}

static void rk4(pbo_data *d) {
  static const double A[3][3] = {{0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 1}}, B[4] = {1.0 / 6, 1.0 / 3, 1.0 / 3, 1.0 / 6};
  const pb_model *m = d->m;
  int nv = m->nv, nq = m->nq;
  double h = m->timestep, q0[PB_MAXQ], v0[PB_MAXDOF], Xv[4][PB_MAXDOF], F[4][PB_MAXDOF], dv[PB_MAXDOF], da[PB_MAXDOF];
  memcpy(q0, d->qpos, sizeof(double) * nq);
  memcpy(v0, d->qvel, sizeof(double) * nv);
  memcpy(Xv[0], d->qvel, sizeof(double) * nv);
  memcpy(F[0], d->qacc, sizeof(double) * nv);
  for (int i = 1; i < 4; i++) {
    for (int k = 0; k < nv; k++) {
      dv[k] = da[k] = 0;
      for (int j = 0; j < i; j++) { dv[k] += A[i - 1][j] * Xv[j][k]; da[k] += A[i - 1][j] * F[j][k]; }
    }
    memcpy(d->qpos, q0, sizeof(double) * nq);
    integrate_pos(m, d->qpos, dv, h);
    for (int k = 0; k < nv; k++) d->qvel[k] = v0[k] + h * da[k];
    memcpy(Xv[i], d->qvel, sizeof(double) * nv);
    g_rk_stage = i;
    pbo_forward(d);
    g_rk_stage = 0;
    memcpy(F[i], d->qacc, sizeof(double) * nv);
  }
  for (int k = 0; k < nv; k++) {
    dv[k] = da[k] = 0;
    for (int j = 0; j < 4; j++) { dv[k] += B[j] * Xv[j][k]; da[k] += B[j] * F[j][k]; }
  }
  memcpy(d->qpos, q0, sizeof(double) * nq);
  for (int k = 0; k < nv; k++) d->qvel[k] = v0[k] + h * da[k];
  integrate_pos(m, d->qpos, dv, h);
}

void pbo_step(pbo_data *d) {
  pbo_forward(d);
  if (d->m->integrator == PB_INT_RK4) rk4(d); else euler(d);
}

/* ------------------------------------------------------------------ A4.8 mj_rnePostConstraint (cfrc_ext) */
void pbo_rne_post_constraint(pbo_data *d) {
  const pb_model *m = d->m;
  memset(d->cfrc_ext, 0, sizeof(d->cfrc_ext));
  for (int c = 0; c < d->ncon; c++) {
    const pbo_contact *con = &d->contact[c];
    if (con->efc_address < 0) continue;
    const double *f = d->efc_force + con->efc_address;
    double lf[3] = {0, 0, 0}, wf[3];
    if (con->dim == 1) lf[0] = f[0];
    else {
      lf[0] = f[0] + f[1] + f[2] + f[3];
      lf[1] = (f[0] - f[1]) * con->friction[0];
      lf[2] = (f[2] - f[3]) * con->friction[0];
    }
    for (int k = 0; k < 3; k++) wf[k] = con->frame[k] * lf[0] + con->frame[3 + k] * lf[1] + con->frame[6 + k] * lf[2];
    int bs[2] = {m->geom_body[con->g1], m->geom_body[con->g2]};
    for (int s = 0; s < 2; s++) {
      int b = bs[s];
      if (!b) continue;
      const double *com = d->subtree_com[m->body_root[b]];
      double off[3] = {con->pos[0] - com[0], con->pos[1] - com[1], con->pos[2] - com[2]}, tq[3];
      cross3(tq, off, wf);
      double sg = s == 0 ? -1.0 : 1.0;
      for (int k = 0; k < 3; k++) { d->cfrc_ext[b][k] += sg * tq[k]; d->cfrc_ext[b][3 + k] += sg * wf[k]; }
    }
  }
}

double pbo_energy(const pbo_data *d, double *kinetic, double *potential) {
  const pb_model *m = d->m;
  double ke = 0, pe = 0, Mv[PB_MAXDOF];
  mul_M(d, Mv, d->qvel);
  for (int i = 0; i < m->nv; i++) ke += 0.5 * d->qvel[i] * Mv[i];
  for (int b = 1; b < m->nbody; b++) pe -= m->body_mass[b] * dot3(m->gravity, d->xipos[b]);
  for (int j = 0; j < m->njnt; j++)
    if (m->jnt_type[j] != PB_JNT_FREE) {
      double q = d->qpos[m->jnt_qposadr[j]] - m->jnt_springref[j];
      pe += 0.5 * m->jnt_stiffness[j] * q * q;
    }
  if (kinetic) *kinetic = ke;
  if (potential) *potential = pe;
  return ke + pe;
}
