/* pb_devmodel.h — float32, GPU-side packing of a compiled MJCF model (pb_model) plus the static schedules the
 * warp-per-env kernels are driven by (tree levels, sparse mass-matrix entry list, L^T D L update schedule,
 * per-pair / per-limit constraint constants).  Built once on the host by pb_build_devmodel(). */
#ifndef PB_DEVMODEL_H
#define PB_DEVMODEL_H

#include "pb_model.h"

#define PB_MAXM 192   /* sparse mass-matrix entries (Humanoid: 185) */
#define PB_MAXFAC 800 /* factorisation update schedule entries (Humanoid: 778) */

struct DevModel {
  int nq, nv, nu, nbody, njnt, ngeom, npair, nM, nlevel, iterations;
  float timestep, tolerance, meaninertia, solver_scale;
  float gravity[3];
  float total_mass;
  /* bodies */
  int body_parent[PB_MAXBODY], body_jntadr[PB_MAXBODY], body_jntnum[PB_MAXBODY];
  int body_dofadr[PB_MAXBODY], body_dofnum[PB_MAXBODY], body_subtree[PB_MAXBODY];
  unsigned body_dofmask[PB_MAXBODY]; /* bit i set: dof i moves this body */
  int body_lastdof[PB_MAXBODY];      /* last dof of the nearest body on the path to the root that has dofs (itself included), -1: none */
  int body_rounds, dof_rounds;       /* pointer-jumping rounds: ceil(log2(deepest body chain / dof chain)) */
  float body_pos[PB_MAXBODY][3], body_quat[PB_MAXBODY][4], body_mass[PB_MAXBODY], body_ipos[PB_MAXBODY][3];
  float body_inertia[PB_MAXBODY][6]; /* xx yy zz xy xz yz about the body com, body axes */
  int level_start[PB_MAXBODY + 2], level_body[PB_MAXBODY];
  /* joints */
  int jnt_type[PB_MAXJNT], jnt_qposadr[PB_MAXJNT], jnt_dofadr[PB_MAXJNT], jnt_limited[PB_MAXJNT], jnt_body[PB_MAXJNT];
  float jnt_pos[PB_MAXJNT][3], jnt_axis[PB_MAXJNT][3], jnt_range[PB_MAXJNT][2], jnt_qpos0[PB_MAXJNT];
  float jnt_stiffness[PB_MAXJNT], jnt_springref[PB_MAXJNT], jnt_margin[PB_MAXJNT];
  float jnt_solimp[PB_MAXJNT][5], jnt_k[PB_MAXJNT], jnt_b[PB_MAXJNT], jnt_diag[PB_MAXJNT];
  /* dofs */
  int dof_body[PB_MAXDOF], dof_parent[PB_MAXDOF], dof_Madr[PB_MAXDOF], dof_depth[PB_MAXDOF];
  unsigned dof_ancmask[PB_MAXDOF]; /* bit j set: dof j is dof i or one of its ancestors */
  int dof_velsrc[PB_MAXDOF];       /* dof whose chain-inclusive velocity is the running cvel when cdof_dot of dof i is formed
                                      (mj_comVel): the parent dof, except that the three rotational dofs of a free joint all
                                      see the velocity after its translational dofs; -1: zero */
  int dof_cdd_zero[PB_MAXDOF];     /* 1: cdof_dot is zero (translational dofs of a free joint) */
  float dof_armature[PB_MAXDOF], dof_damping[PB_MAXDOF];
  float dof_gear[PB_MAXDOF], dof_ctrl_lo[PB_MAXDOF], dof_ctrl_hi[PB_MAXDOF];
  int dof_act[PB_MAXDOF]; /* actuator driving this dof, -1 if none */
  float qpos0[PB_MAXQ];
  /* sparse M: entry e is M(M_i[e], M_j[e]), row-major by dof with the diagonal first then ancestors */
  unsigned char M_i[PB_MAXM], M_j[PB_MAXM];
  /* L^T D L schedule: for step k (nv-1..0) entries [fac_start[k], fac_start[k+1]) hold tgt | a<<8 | b<<16 :
     LD[tgt] -= LD[a]*LD[b]/LD[kk] */
  int fac_start[PB_MAXDOF + 1];
  unsigned fac_tab[PB_MAXFAC];
  /* geoms */
  int geom_type[PB_MAXGEOM], geom_body[PB_MAXGEOM];
  float geom_size[PB_MAXGEOM][2], geom_pos[PB_MAXGEOM][3], geom_zaxis[PB_MAXGEOM][3];
  /* collision pairs with pre-mixed contact constants */
  unsigned char pair_g1[PB_MAXPAIR], pair_g2[PB_MAXPAIR], pair_condim[PB_MAXPAIR];
  float pair_margin[PB_MAXPAIR], pair_incmargin[PB_MAXPAIR], pair_mu[PB_MAXPAIR];
  float pair_reach2[PB_MAXPAIR]; /* squared bounding-sphere reach (r+h of both geoms + margin); INF for plane pairs */
  float pair_solimp[PB_MAXPAIR][5], pair_k[PB_MAXPAIR], pair_b[PB_MAXPAIR], pair_diag[PB_MAXPAIR];
};

#ifdef __cplusplus
extern "C" {
#endif
/* returns 0 on success, <0 if a table overflows */
int pb_build_devmodel(const pb_model *src, struct DevModel *dst);
#ifdef __cplusplus
}
#endif

#endif
