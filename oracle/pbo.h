/* pbo.h — CPU ORACLE (test infrastructure, NOT the product path).
 *
 * FP64 plain-C restatement of the reference hot path: the MuJoCo 3.2.0 forward dynamics the reference calls
 * through mj_step / mj_forward / mj_resetData / mj_rnePostConstraint (libs/gymcpp/mujoco/mujoco_env.h:97-108),
 * the four task definitions (libs/gymcpp/mujoco/{half_cheetah_v5,hopper_v5,ant_v5,humanoid_v4}.h), the wrapper
 * chain (libs/gymcpp/wrappers/*.h) and the vector env (libs/gymcpp/gym.h:310-365).
 *
 * PARITY UNPINNED for the physics layer: MuJoCo 3.2.0 is a prebuilt third-party binary that is not vendored
 * in the reference (CMakeLists.txt:33-34,65) and cannot be installed here, and the reference has no tests.
 * The physics below restates MuJoCo's published algorithm (computation pipeline chapter + engine_* structure);
 * it is pinned only by (i) the body-mass constants in src/ac_ppo_continuous_action.cpp:496-497 and (ii) the
 * physical invariants in tests/test_oracle_physics.py.  Task/wrapper/vector-env logic follows the reference
 * sources line by line and is pinned by construction.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use this. */
#ifndef PBO_H
#define PBO_H

#include "../ppo_cpp_b200/csrc/pb_model.h"

#define PBO_MAXCON 64
#define PBO_MAXEFC 200

typedef struct pbo_contact {
  double dist, pos[3], frame[9], includemargin, friction[3], solref[2], solimp[5];
  int dim, g1, g2, efc_address;
} pbo_contact;

typedef struct pbo_data {
  const pb_model *m;
  /* state */
  double qpos[PB_MAXQ], qvel[PB_MAXDOF], ctrl[PB_MAXU], qacc[PB_MAXDOF], qacc_warmstart[PB_MAXDOF];
  /* position-dependent */
  double xpos[PB_MAXBODY][3], xquat[PB_MAXBODY][4], xmat[PB_MAXBODY][9], xipos[PB_MAXBODY][3];
  double xanchor[PB_MAXJNT][3], xaxis[PB_MAXJNT][3];
  double gxpos[PB_MAXGEOM][3], gxmat[PB_MAXGEOM][9];
  double subtree_com[PB_MAXBODY][3];
  double cinert[PB_MAXBODY][10], cdof[PB_MAXDOF][6];
  double M[PB_MAXDOF][PB_MAXDOF], L[PB_MAXDOF][PB_MAXDOF]; /* dense mass matrix and its Cholesky factor */
  int ncon;
  pbo_contact contact[PBO_MAXCON];
  int nefc;
  double efc_J[PBO_MAXEFC][PB_MAXDOF], efc_pos[PBO_MAXEFC], efc_margin[PBO_MAXEFC], efc_diagApprox[PBO_MAXEFC];
  double efc_R[PBO_MAXEFC], efc_D[PBO_MAXEFC], efc_k[PBO_MAXEFC], efc_bdamp[PBO_MAXEFC], efc_imp[PBO_MAXEFC];
  double efc_vel[PBO_MAXEFC], efc_aref[PBO_MAXEFC], efc_b[PBO_MAXEFC], efc_force[PBO_MAXEFC];
  /* velocity-dependent */
  double cvel[PB_MAXBODY][6], cdof_dot[PB_MAXDOF][6];
  double qfrc_passive[PB_MAXDOF], qfrc_bias[PB_MAXDOF], qfrc_actuator[PB_MAXDOF], qfrc_smooth[PB_MAXDOF];
  double qacc_smooth[PB_MAXDOF], qfrc_constraint[PB_MAXDOF];
  double cfrc_ext[PB_MAXBODY][6];
  int solver_niter;
  /* storage caps of the engine under test (0 = the oracle's own PBO_MAXCON / PBO_MAXEFC): contacts beyond cap_con are
   * dropped in pair order; the first contact whose rows do not fit cap_efc is dropped together with every later contact
   * (limit rows always fit).  Same rule as csrc/pb_sim_pos.cuh (sim_collision / sim_make_constraint); overflow is set
   * whenever something was dropped.  Survives pbo_reset_data. */
  int cap_con, cap_efc, overflow;
  /* per-evaluation trace of (nefc, solver iterations) since the last pbo_trace_reset, for cost-model studies */
  int trace_n;
  short trace_nefc[64], trace_niter[64];
  long n_forward; /* forward-dynamics evaluations performed (for FLOP/throughput accounting) */
} pbo_data;

void pbo_reset_data(const pb_model *m, pbo_data *d);     /* mj_resetData */
void pbo_forward(pbo_data *d);                           /* mj_forward (sensors skipped) */
void pbo_step(pbo_data *d);                              /* mj_step */
void pbo_rne_post_constraint(pbo_data *d);               /* mj_rnePostConstraint (cfrc_ext only) */
double pbo_energy(const pbo_data *d, double *kinetic, double *potential); /* test helper */

const pb_model *pbo_model_by_name(const char *env_id);

#endif
