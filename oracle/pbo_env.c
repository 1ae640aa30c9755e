/* pbo_env.c — CPU ORACLE (test infrastructure): task definitions, wrapper chain and vector env of the
 * reference, restated line by line in plain C on top of pbo_physics.c.
 *   tasks     libs/gymcpp/mujoco/half_cheetah_v5.h:62-130, hopper_v5.h:69-183, ant_v5.h:71-198,
 *             humanoid_v4.h:58-158; mujocoStep libs/gymcpp/mujoco/mujoco_env.h:97-108
 *   wrappers  libs/gymcpp/wrappers/common.h:25-65, stateful_observation.h:44-84, transform_observation.h:33-44,
 *             stateful_reward.h:49-91, vectorize_reward.h:31-42; chain src/ppo_continuous_action.cpp:41-49
 *   vector    libs/gymcpp/gym.h:310-365 (ParVectorEnv reset/step, next-step autoreset)
 *   rng       std::mt19937 + libstdc++ uniform_real_distribution<double>/normal_distribution<double>
 * Wrapper arithmetic is float32 exactly as the reference (kFloat32 tensors, float scalars); build with
 * -ffp-contract=off so no FMA contraction changes it. */
#include "pbo.h"

#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ std::mt19937 + libstdc++ distributions */
typedef struct { uint32_t mt[624]; int idx; int saved_ok; double saved; } pbo_rng;

void pbo_rng_seed(pbo_rng *r, uint32_t seed) {
  r->mt[0] = seed;
  for (int i = 1; i < 624; i++) r->mt[i] = 1812433253u * (r->mt[i - 1] ^ (r->mt[i - 1] >> 30)) + (uint32_t)i;
  r->idx = 624;
}
uint32_t pbo_rng_u32(pbo_rng *r) {
  if (r->idx >= 624) {
    for (int i = 0; i < 624; i++) {
      uint32_t y = (r->mt[i] & 0x80000000u) | (r->mt[(i + 1) % 624] & 0x7fffffffu);
      r->mt[i] = r->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    r->idx = 0;
  }
  uint32_t y = r->mt[r->idx++];
  y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
  return y;
}
static double rng_canonical(pbo_rng *r) { /* generate_canonical<double,53> over a 32-bit engine: 2 draws */
  double sum = (double)pbo_rng_u32(r);
  sum += (double)pbo_rng_u32(r) * 4294967296.0;
  double ret = sum / 18446744073709551616.0;
  if (ret >= 1.0) ret = nextafter(1.0, 0.0);
  return ret;
}
static double rng_uniform(pbo_rng *r, double a, double b) { return (b - a) * rng_canonical(r) + a; }
static double rng_normal(pbo_rng *r, double mean, double stddev) { /* Marsaglia polar, second value cached */
  double ret;
  if (r->saved_ok) { r->saved_ok = 0; ret = r->saved; }
  else {
    double x, y, r2;
    do { x = 2.0 * rng_canonical(r) - 1.0; y = 2.0 * rng_canonical(r) - 1.0; r2 = x * x + y * y; } while (r2 > 1.0 || r2 == 0.0);
    double mult = sqrt(-2 * log(r2) / r2);
    r->saved = x * mult; r->saved_ok = 1;
    ret = y * mult;
  }
  return ret * stddev + mean;
}

/* ------------------------------------------------------------------ one env = task + wrapper chain */
enum { TASK_HALFCHEETAH, TASK_HOPPER, TASK_ANT, TASK_HUMANOID };
enum { CHAIN_STATS_ONLY = 0, CHAIN_PPO = 1 }; /* ac_ppo chain (:50-53) vs ppo chain (:41-49) */
#define MAXOBS 376

typedef struct {
  int task, frame_skip, obs_dim, elapsed;
  pbo_data d;
  pbo_rng rng;
  /* RecordEpisodeStatistics */
  float ep_return; int ep_length;
  /* NormalizeObservation (float32, per env) */
  float ob_mean[MAXOBS], ob_var[MAXOBS], ob_count;
  /* NormalizeReward */
  float rw_mean, rw_var, rw_acc, rw_count;
  int autoreset;
} pbo_env;

typedef struct pbo_vec {
  int num_envs, chain, clip_actions, obs_dim, act_dim, task, nthreads;
  float gamma, act_min, act_max;
  pbo_env *envs;
} pbo_vec;

static const struct { const char *id; int task, frame_skip, obs, act; float lo, hi; } TASKS[] = {
    {"HalfCheetah-v5", TASK_HALFCHEETAH, 5, 17, 6, -1.0f, 1.0f},
    {"Hopper-v5", TASK_HOPPER, 4, 11, 3, -1.0f, 1.0f},
    {"Ant-v5", TASK_ANT, 5, 105, 8, -1.0f, 1.0f},
    {"Humanoid-v4", TASK_HUMANOID, 5, 376, 17, -0.4f, 0.4f}};

static void task_write_state(pbo_env *e, float *obs) {
  const pbo_data *d = &e->d;
  const pb_model *m = d->m;
  int k = 0;
  if (e->task == TASK_HALFCHEETAH || e->task == TASK_HOPPER) {
    for (int i = 1; i < m->nq; i++) obs[k++] = (float)d->qpos[i];
    for (int i = 0; i < m->nv; i++) {
      double v = d->qvel[i];
      if (e->task == TASK_HOPPER) v = v < -10.0 ? -10.0 : (v > 10.0 ? 10.0 : v);
      obs[k++] = (float)v;
    }
  } else {
    for (int i = 2; i < m->nq; i++) obs[k++] = (float)d->qpos[i];
    for (int i = 0; i < m->nv; i++) obs[k++] = (float)d->qvel[i];
    if (e->task == TASK_ANT) {
      const double *c = &d->cfrc_ext[0][0];
      for (int i = 6; i < 6 * m->nbody; i++) obs[k++] = (float)(c[i] < -1.0 ? -1.0 : (c[i] > 1.0 ? 1.0 : c[i]));
    } else {
      const double *ci = &d->cinert[0][0], *cv = &d->cvel[0][0], *cf = &d->cfrc_ext[0][0];
      for (int i = 0; i < 10 * m->nbody; i++) obs[k++] = (float)ci[i];
      for (int i = 0; i < 6 * m->nbody; i++) obs[k++] = (float)cv[i];
      for (int i = 0; i < m->nv; i++) obs[k++] = (float)d->qfrc_actuator[i];
      for (int i = 0; i < 6 * m->nbody; i++) obs[k++] = (float)cf[i];
    }
  }
}

static void task_reset(pbo_env *e, int seed, float *obs) {
  const pb_model *m = e->d.m;
  if (seed > 0) pbo_rng_seed(&e->rng, (uint32_t)seed);
  e->elapsed = 0;
  long nfwd = e->d.n_forward;
  pbo_reset_data(m, &e->d);
  pbo_data *d = &e->d;
  d->n_forward = nfwd;
  switch (e->task) {
    case TASK_HALFCHEETAH:
      for (int i = 0; i < m->nq; i++) d->qpos[i] = m->qpos0[i] + rng_uniform(&e->rng, -0.1, 0.1);
      for (int i = 0; i < m->nv; i++) d->qvel[i] = 0.0 + rng_normal(&e->rng, 0.0, 0.1);
      break;
    case TASK_HOPPER:
      for (int i = 0; i < m->nq; i++) d->qpos[i] = m->qpos0[i] + rng_uniform(&e->rng, -5e-3, 5e-3);
      for (int i = 0; i < m->nv; i++) d->qvel[i] = 0.0 + rng_uniform(&e->rng, -5e-3, 5e-3);
      break;
    case TASK_ANT:
      for (int i = 0; i < m->nq; i++) d->qpos[i] = m->qpos0[i] + rng_uniform(&e->rng, -0.1, 0.1);
      for (int i = 0; i < m->nv; i++) d->qvel[i] = 0.0 + 0.1 * rng_normal(&e->rng, 0.0, 1.0);
      break;
    default:
      for (int i = 0; i < m->nq; i++) d->qpos[i] = m->qpos0[i] + rng_uniform(&e->rng, -1e-2, 1e-2);
      for (int i = 0; i < m->nv; i++) d->qvel[i] = 0.0 + rng_uniform(&e->rng, -1e-2, 1e-2);
  }
  pbo_forward(d);
  task_write_state(e, obs);
}

static void humanoid_mass_center(const pbo_data *d, double *xy) {
  const pb_model *m = d->m;
  double ms = 0, mx = 0, my = 0;
  for (int b = 0; b < m->nbody; b++) {
    ms += m->body_mass[b];
    mx += m->body_mass[b] * d->xipos[b][0];
    my += m->body_mass[b] * d->xipos[b][1];
  }
  xy[0] = mx / ms; xy[1] = my / ms;
}

static void mujoco_step(pbo_env *e, const float *action) {
  pbo_data *d = &e->d;
  for (int i = 0; i < d->m->nu; i++) d->ctrl[i] = (double)action[i];
  for (int i = 0; i < e->frame_skip; i++) pbo_step(d);
  pbo_rne_post_constraint(d);
}

static void task_step(pbo_env *e, const float *action, float *obs, float *reward, int *term, int *trunc) {
  pbo_data *d = &e->d;
  const pb_model *m = d->m;
  double dt = m->timestep * e->frame_skip, ctrl2 = 0;
  if (e->task == TASK_HALFCHEETAH) {
    double xb = d->qpos[0];
    mujoco_step(e, action);
    double xv = (d->qpos[0] - xb) / dt, cost = 0;
    for (int i = 0; i < m->nu; i++) cost += 0.1 * d->ctrl[i] * d->ctrl[i];
    *reward = (float)(1.0 * xv - cost);
    *term = 0;
    *trunc = (++e->elapsed >= 1000);
    task_write_state(e, obs);
  } else if (e->task == TASK_HOPPER) {
    double xb = d->qpos[0];
    mujoco_step(e, action);
    double xv = (d->qpos[0] - xb) / dt;
    task_write_state(e, obs);
    int healthy = 1;
    double z = d->qpos[1], ang = d->qpos[2];
    if (z < 0.7 || z > INFINITY) healthy = 0;
    if (ang < -0.2 || ang > 0.2) healthy = 0;
    for (int i = 2; i < m->nq; i++) if (d->qpos[i] < -100.0 || d->qpos[i] > 100.0) healthy = 0;
    for (int i = 0; i < m->nv; i++) if (d->qvel[i] < -100.0 || d->qvel[i] > 100.0) healthy = 0;
    double rewards = xv * 1.0 + (double)healthy * 1.0, cost = 0;
    for (int i = 0; i < m->nu; i++) cost += 1e-3 * d->ctrl[i] * d->ctrl[i];
    *reward = (float)(rewards - cost);
    *term = !healthy;
    *trunc = (++e->elapsed >= 1000);
  } else if (e->task == TASK_ANT) {
    double xb = d->xpos[1][0];
    mujoco_step(e, action);
    double xv = (d->xpos[1][0] - xb) / dt;
    task_write_state(e, obs);
    int healthy = d->qpos[2] >= 0.2 && d->qpos[2] <= 1.0;
    for (int i = 0; i < m->nq; i++) if (!isfinite(d->qpos[i])) healthy = 0;
    for (int i = 0; i < m->nv; i++) if (!isfinite(d->qvel[i])) healthy = 0;
    double rewards = xv * 1.0 + (double)healthy * 1.0, cost = 0, ccost = 0;
    for (int i = 0; i < m->nu; i++) cost += 0.5 * d->ctrl[i] * d->ctrl[i];
    const double *c = &d->cfrc_ext[0][0];
    for (int i = 6; i < 6 * m->nbody; i++) {
      double f = c[i] < -1.0 ? -1.0 : (c[i] > 1.0 ? 1.0 : c[i]);
      ccost += f * f;
    }
    ccost = 5e-4 * ccost;
    *reward = (float)(rewards - (cost + ccost));
    *term = !healthy;
    *trunc = (++e->elapsed >= 1000);
  } else {
    double before[2], after[2];
    humanoid_mass_center(d, before);
    mujoco_step(e, action);
    humanoid_mass_center(d, after);
    for (int i = 0; i < m->nu; i++) ctrl2 += 0.1 * d->ctrl[i] * d->ctrl[i];
    double xv = (after[0] - before[0]) / dt;
    *reward = (float)(xv * 1.25 + 5.0 - ctrl2);
    ++e->elapsed;
    *term = !(1.0 < d->qpos[2] && d->qpos[2] < 2.0);
    *trunc = (e->elapsed >= 1000);
    task_write_state(e, obs);
  }
}

/* Welford update with batch_count = 1, float32, operation order of stateful_observation.h:64-84 */
static void welford_f32(float x, float *mean, float *var, float count, float tot) {
  float delta = x - *mean;
  float new_mean = *mean + delta * 1.0f / tot;
  float m_a = *var * count;
  float M2 = m_a + 0.0f + (delta * delta) * count * 1.0f / tot;
  *var = M2 / tot;
  *mean = new_mean;
}

static void chain_observation(pbo_vec *v, pbo_env *e, float *obs) {
  if (v->chain != CHAIN_PPO) return;
  float tot = e->ob_count + 1.0f;
  for (int i = 0; i < e->obs_dim; i++) {
    welford_f32(obs[i], &e->ob_mean[i], &e->ob_var[i], e->ob_count, tot);
    float y = (obs[i] - e->ob_mean[i]) / sqrtf(e->ob_var[i] + 1e-4f);
    obs[i] = y < -10.0f ? -10.0f : (y > 10.0f ? 10.0f : y);
  }
  e->ob_count = tot;
}

static void env_reset(pbo_vec *v, pbo_env *e, int seed, float *obs) {
  task_reset(e, seed, obs);
  e->ep_return = 0.0f;
  e->ep_length = 0;
  chain_observation(v, e, obs);
}

static void env_step(pbo_vec *v, pbo_env *e, const float *action, float *obs, float *rew, float *term, float *trunc,
                     int *has_info, float *info_r, int *info_l) {
  float a[PB_MAXU];
  for (int i = 0; i < v->act_dim; i++) {
    a[i] = action[i];
    if (v->clip_actions) a[i] = a[i] < v->act_min ? v->act_min : (a[i] > v->act_max ? v->act_max : a[i]);
  }
  if (e->autoreset) {
    env_reset(v, e, -1, obs);
    *rew = 0.0f; *term = 0.0f; *trunc = 0.0f; *has_info = 0;
    e->autoreset = 0;
    return;
  }
  float r; int te, tr;
  task_step(e, a, obs, &r, &te, &tr);
  e->ep_return += r;
  e->ep_length += 1;
  *has_info = te || tr;
  if (*has_info) { *info_r = e->ep_return; *info_l = e->ep_length; }
  chain_observation(v, e, obs);
  if (v->chain == CHAIN_PPO) {
    e->rw_acc = e->rw_acc * v->gamma * (1.0f - (float)te) + r;
    float tot = e->rw_count + 1.0f;
    welford_f32(e->rw_acc, &e->rw_mean, &e->rw_var, e->rw_count, tot);
    e->rw_count = tot;
    r = r / sqrtf(e->rw_var + 1e-8f);
    r = r < -10.0f ? -10.0f : (r > 10.0f ? 10.0f : r);
  }
  *rew = r; *term = (float)te; *trunc = (float)tr;
  e->autoreset = te || tr;
}

/* ------------------------------------------------------------------ vector env (gym.h:276-366) */
pbo_vec *pbo_vec_create(const char *env_id, int num_envs, int chain, float gamma, int clip_actions, int nthreads) {
  int t = -1;
  for (int i = 0; i < 4; i++) if (!strcmp(env_id, TASKS[i].id)) t = i;
  if (t < 0) return 0;
  pbo_vec *v = (pbo_vec *)calloc(1, sizeof(pbo_vec));
  v->num_envs = num_envs; v->chain = chain; v->gamma = gamma; v->clip_actions = clip_actions;
  v->obs_dim = TASKS[t].obs; v->act_dim = TASKS[t].act; v->act_min = TASKS[t].lo; v->act_max = TASKS[t].hi;
  v->task = TASKS[t].task; v->nthreads = nthreads < 1 ? 1 : nthreads;
  v->envs = (pbo_env *)calloc(num_envs, sizeof(pbo_env));
  const pb_model *m = pbo_model_by_name(env_id);
  for (int i = 0; i < num_envs; i++) {
    pbo_env *e = &v->envs[i];
    e->task = TASKS[t].task; e->frame_skip = TASKS[t].frame_skip; e->obs_dim = TASKS[t].obs;
    e->elapsed = 1001;
    pbo_reset_data(m, &e->d);
    pbo_rng_seed(&e->rng, 5489u);
    for (int k = 0; k < MAXOBS; k++) e->ob_var[k] = 1.0f;
    e->ob_count = 1e-4f;
    e->rw_var = 1.0f; e->rw_count = 1e-8f;
  }
  return v;
}
void pbo_vec_destroy(pbo_vec *v) { if (v) { free(v->envs); free(v); } }
int pbo_vec_obs_dim(const pbo_vec *v) { return v->obs_dim; }
int pbo_vec_act_dim(const pbo_vec *v) { return v->act_dim; }

void pbo_vec_reset(pbo_vec *v, int seed, float *obs) {
  for (int i = 0; i < v->num_envs; i++) {
    env_reset(v, &v->envs[i], seed + i, obs + (size_t)i * v->obs_dim);
    v->envs[i].autoreset = 0;
  }
}

typedef struct { pbo_vec *v; int lo, hi; const float *act; float *obs, *rew, *term, *trunc, *info_r; int *has, *info_l; } job_t;
static void *step_job(void *p) {
  job_t *j = (job_t *)p;
  pbo_vec *v = j->v;
  for (int i = j->lo; i < j->hi; i++)
    env_step(v, &v->envs[i], j->act + (size_t)i * v->act_dim, j->obs + (size_t)i * v->obs_dim, j->rew + i, j->term + i,
             j->trunc + i, j->has + i, j->info_r + i, j->info_l + i);
  return 0;
}
void pbo_vec_step(pbo_vec *v, const float *actions, float *obs, float *rew, float *term, float *trunc, int *has_info,
                  float *info_r, int *info_l) {
  int nt = v->nthreads > v->num_envs ? v->num_envs : v->nthreads;
  job_t jobs[256];
  pthread_t th[256];
  if (nt > 256) nt = 256;
  for (int t = 0; t < nt; t++) {
    job_t j = {v, (int)((long)v->num_envs * t / nt), (int)((long)v->num_envs * (t + 1) / nt), actions, obs, rew, term,
               trunc, info_r, has_info, info_l};
    jobs[t] = j;
  }
  if (nt == 1) { step_job(&jobs[0]); return; }
  for (int t = 0; t < nt; t++) pthread_create(&th[t], 0, step_job, &jobs[t]);
  for (int t = 0; t < nt; t++) pthread_join(th[t], 0);
}

/* state access for parity tests: copy physics state of env i out / in */
void pbo_vec_get_state(pbo_vec *v, int i, double *qpos, double *qvel, double *qacc_warmstart) {
  const pbo_data *d = &v->envs[i].d;
  memcpy(qpos, d->qpos, sizeof(double) * d->m->nq);
  memcpy(qvel, d->qvel, sizeof(double) * d->m->nv);
  memcpy(qacc_warmstart, d->qacc_warmstart, sizeof(double) * d->m->nv);
}
pbo_data *pbo_vec_data(pbo_vec *v, int i) { return &v->envs[i].d; }
/* give every env the same normaliser statistics (a trained policy's normaliser; mirrors pb_env_set_norm_stats) */
void pbo_vec_set_norm_stats(pbo_vec *v, const float *ob_mean, const float *ob_var, float ob_count, const float *rw4) {
  for (int i = 0; i < v->num_envs; i++) {
    pbo_env *e = &v->envs[i];
    memcpy(e->ob_mean, ob_mean, sizeof(float) * v->obs_dim);
    memcpy(e->ob_var, ob_var, sizeof(float) * v->obs_dim);
    e->ob_count = ob_count;
    if (rw4) { e->rw_mean = rw4[0]; e->rw_var = rw4[1]; e->rw_acc = rw4[2]; e->rw_count = rw4[3]; }
  }
}
long pbo_vec_total_forwards(const pbo_vec *v) {
  long n = 0;
  for (int i = 0; i < v->num_envs; i++) n += v->envs[i].d.n_forward;
  return n;
}
