/* pbo_api.c — CPU ORACLE (test infrastructure): flat accessors so pytest can drive the physics through ctypes
 * without mirroring the C structs.  Not part of the product path. */
#include "pbo.h"

#include <stdlib.h>
#include <string.h>

pbo_data *pbo_data_new(const pb_model *m) {
  pbo_data *d = (pbo_data *)malloc(sizeof(pbo_data));
  d->cap_con = d->cap_efc = 0;
  pbo_reset_data(m, d);
  return d;
}
void pbo_data_free(pbo_data *d) { free(d); }
void pbo_data_reset(pbo_data *d) { pbo_reset_data(d->m, d); }

/* model variants for invariant tests: bit flags switch features off on a private copy */
enum { PBO_NO_GRAVITY = 1, PBO_NO_CONTACT = 2, PBO_NO_LIMIT = 4, PBO_NO_PASSIVE = 8, PBO_NO_ARMATURE = 16 };
pb_model *pbo_model_variant(const char *env_id, int flags, int integrator, double timestep) {
  const pb_model *src = pbo_model_by_name(env_id);
  if (!src) return 0;
  pb_model *m = (pb_model *)malloc(sizeof(pb_model));
  memcpy(m, src, sizeof(pb_model));
  if (flags & PBO_NO_GRAVITY) m->gravity[0] = m->gravity[1] = m->gravity[2] = 0;
  if (flags & PBO_NO_CONTACT) m->npair = 0;
  if (flags & PBO_NO_LIMIT) for (int j = 0; j < m->njnt; j++) m->jnt_limited[j] = 0;
  if (flags & PBO_NO_PASSIVE) {
    for (int i = 0; i < m->nv; i++) m->dof_damping[i] = 0;
    for (int j = 0; j < m->njnt; j++) m->jnt_stiffness[j] = 0;
  }
  if (flags & PBO_NO_ARMATURE) for (int i = 0; i < m->nv; i++) m->dof_armature[i] = 0;
  if (integrator >= 0) m->integrator = integrator;
  if (timestep > 0) m->timestep = timestep;
  return m;
}
void pbo_model_free(pb_model *m) { free(m); }
int pbo_model_dims(const pb_model *m, int *out) {
  out[0] = m->nq; out[1] = m->nv; out[2] = m->nu; out[3] = m->nbody; out[4] = m->njnt; out[5] = m->ngeom;
  out[6] = m->npair;
  return 0;
}

/* returns element count and a pointer into d (double fields only) */
int pbo_field(pbo_data *d, const char *name, double **ptr) {
  const pb_model *m = d->m;
#define F(n, p, c) if (!strcmp(name, n)) { *ptr = (double *)(p); return (c); }
  F("qpos", d->qpos, m->nq) F("qvel", d->qvel, m->nv) F("ctrl", d->ctrl, m->nu) F("qacc", d->qacc, m->nv)
  F("qacc_warmstart", d->qacc_warmstart, m->nv) F("xpos", d->xpos, 3 * m->nbody) F("xquat", d->xquat, 4 * m->nbody)
  F("xmat", d->xmat, 9 * m->nbody) F("xipos", d->xipos, 3 * m->nbody) F("subtree_com", d->subtree_com, 3 * m->nbody)
  F("cinert", d->cinert, 10 * m->nbody) F("cdof", d->cdof, 6 * m->nv) F("cvel", d->cvel, 6 * m->nbody)
  F("M", d->M, PB_MAXDOF * PB_MAXDOF) F("qfrc_bias", d->qfrc_bias, m->nv) F("qfrc_passive", d->qfrc_passive, m->nv)
  F("qfrc_actuator", d->qfrc_actuator, m->nv) F("qfrc_smooth", d->qfrc_smooth, m->nv)
  F("qacc_smooth", d->qacc_smooth, m->nv) F("qfrc_constraint", d->qfrc_constraint, m->nv)
  F("cfrc_ext", d->cfrc_ext, 6 * m->nbody) F("efc_force", d->efc_force, d->nefc) F("efc_pos", d->efc_pos, d->nefc)
  F("efc_aref", d->efc_aref, d->nefc) F("efc_R", d->efc_R, d->nefc) F("efc_J", d->efc_J, PB_MAXDOF * d->nefc)
  F("gxpos", d->gxpos, 3 * m->ngeom) F("gxmat", d->gxmat, 9 * m->ngeom)
#undef F
  return -1;
}
void pbo_set_caps(pbo_data *d, int cap_con, int cap_efc) { d->cap_con = cap_con; d->cap_efc = cap_efc; }
int pbo_overflow(const pbo_data *d) { return d->overflow; }
void pbo_trace_reset(pbo_data *d) { d->trace_n = 0; }
int pbo_trace_get(const pbo_data *d, short *nefc, short *niter) {
  memcpy(nefc, d->trace_nefc, sizeof(short) * d->trace_n);
  memcpy(niter, d->trace_niter, sizeof(short) * d->trace_n);
  return d->trace_n;
}
int pbo_counts(const pbo_data *d, int *out) {
  out[0] = d->ncon; out[1] = d->nefc; out[2] = d->solver_niter; out[3] = (int)d->n_forward;
  return 0;
}
int pbo_contact_get(const pbo_data *d, int i, double *out /* dist,pos3,frame9 */, int *iout /* dim,g1,g2,efc */) {
  if (i < 0 || i >= d->ncon) return -1;
  const pbo_contact *c = &d->contact[i];
  out[0] = c->dist;
  memcpy(out + 1, c->pos, sizeof(c->pos));
  memcpy(out + 4, c->frame, sizeof(c->frame));
  iout[0] = c->dim; iout[1] = c->g1; iout[2] = c->g2; iout[3] = c->efc_address;
  return 0;
}

/* model constants for the closed-form / independent-model tests: element count and a pointer into m (doubles or ints) */
int pbo_model_field(const pb_model *m, const char *name, const double **ptr) {
#define F(n, p, c) if (!strcmp(name, n)) { *ptr = (const double *)(p); return (c); }
  F("meaninertia", &m->meaninertia, 1) F("timestep", &m->timestep, 1) F("tolerance", &m->tolerance, 1) F("gravity", m->gravity, 3)
  F("body_mass", m->body_mass, m->nbody) F("body_invweight0", m->body_invweight0, 2 * m->nbody) F("body_ipos", m->body_ipos, 3 * m->nbody)
  F("dof_invweight0", m->dof_invweight0, m->nv) F("dof_damping", m->dof_damping, m->nv) F("dof_armature", m->dof_armature, m->nv)
  F("jnt_stiffness", m->jnt_stiffness, m->njnt) F("jnt_springref", m->jnt_springref, m->njnt) F("jnt_range", m->jnt_range, 2 * m->njnt)
  F("jnt_solref", m->jnt_solref, 2 * m->njnt) F("jnt_solimp", m->jnt_solimp, 5 * m->njnt) F("jnt_margin", m->jnt_margin, m->njnt)
  F("qpos0", m->qpos0, m->nq) F("act_gear", m->act_gear, m->nu)
#undef F
  return -1;
}
int pbo_model_ifield(const pb_model *m, const char *name, const int **ptr) {
#define F(n, p, c) if (!strcmp(name, n)) { *ptr = (const int *)(p); return (c); }
  F("jnt_type", m->jnt_type, m->njnt) F("jnt_dofadr", m->jnt_dofadr, m->njnt) F("jnt_qposadr", m->jnt_qposadr, m->njnt) F("jnt_body", m->jnt_body, m->njnt)
  F("jnt_limited", m->jnt_limited, m->njnt) F("geom_body", m->geom_body, m->ngeom) F("act_dof", m->act_dof, m->nu) F("body_parent", m->body_parent, m->nbody)
  F("dof_body", m->dof_body, m->nv)
#undef F
  return -1;
}
/* parameters of contact i that the row build used: includemargin, friction[3], solref[2], solimp[5] */
int pbo_contact_params(const pbo_data *d, int i, double *out11) {
  if (i < 0 || i >= d->ncon) return -1;
  const pbo_contact *c = &d->contact[i];
  out11[0] = c->includemargin;
  memcpy(out11 + 1, c->friction, sizeof(c->friction));
  memcpy(out11 + 4, c->solref, sizeof(c->solref));
  memcpy(out11 + 6, c->solimp, sizeof(c->solimp));
  return 0;
}

extern int pbo_variant;
void pbo_set_variant(int bits) { pbo_variant = bits; }
