// pb_capi.cu — C ABI of the vector env (include/ppo_b200.h) over the warp-per-env kernels.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/ppo_b200.h"
#include "model_tables.h"
#include "pb_env_kernels.cuh"

#include "pb_internal.h"

static thread_local std::string g_err;
extern "C" const char *pb_last_error(void) { return g_err.c_str(); }
int pb_fail(int code, const std::string &msg) { g_err = msg; return code; }

namespace {
struct TaskInfo { const char *id; int task; const pb_model *model; int obs, act; float lo, hi; };
const TaskInfo kTasks[] = {
    {"HalfCheetah-v5", PB_TASK_HALFCHEETAH, &PB_MODEL_HALFCHEETAH, 17, 6, -1.f, 1.f},
    {"Hopper-v5", PB_TASK_HOPPER, &PB_MODEL_HOPPER, 11, 3, -1.f, 1.f},
    {"Ant-v5", PB_TASK_ANT, &PB_MODEL_ANT, 105, 8, -1.f, 1.f},
    {"Humanoid-v4", PB_TASK_HUMANOID, &PB_MODEL_HUMANOID, 376, 17, -0.4f, 0.4f},
};

template <class C>
bool dims_match(const DevModel &d) {
  bool ok = d.nq == C::NQ && d.nv == C::NV && d.nu == C::NU && d.nbody == C::NB && d.njnt == C::NJ && d.ngeom == C::NG &&
            d.npair == C::NPAIR && d.nM == C::NM;
  // the kernels unroll over a compile-time copy of the dof tree
  for (int i = 0; ok && i < C::NV; i++) ok = d.dof_parent[i] == C::DOF_PARENT[i] && d.dof_Madr[i] == Tree<C>::madr(i);
  return ok;
}

template <class C, int WPB, bool LOCKSTEP, int SYNCMASK = 0, bool MSM = false>
cudaError_t launch(const pb_env *e, const StepArgs &a_in) {
  size_t smem = env_smem_bytes<C, WPB, MSM>();
  static int resident[64] = {};  // CTAs the device holds at once (persistent grid), per device
  if (!resident[e->device]) {
    cudaError_t err = cudaFuncSetAttribute(env_step_kernel<C, WPB, LOCKSTEP, SYNCMASK, MSM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    int per_sm = 0, sms = 0;
    err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, env_step_kernel<C, WPB, LOCKSTEP, SYNCMASK, MSM>, WPB * 32, smem);
    if (err != cudaSuccess) return err;
    err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device);
    if (err != cudaSuccess) return err;
    resident[e->device] = std::max(1, per_sm * sms);
  }
  StepArgs a = a_in;
  a.parity = e->launch_parity;
  e->launch_parity ^= 1;
  int grid = std::min((e->num_envs + WPB - 1) / WPB, resident[e->device]);
  // cost-ordered queue: pays when the persistent grid takes several chunks per CTA and the envs carry last step's costs
  static const int no_order = getenv("PB_ENV_ORDER") ? atoi(getenv("PB_ENV_ORDER")) == 0 : 0;
  a.ordered = LOCKSTEP && !no_order && a.mode == 0 && e->num_envs >= 2 * grid * WPB;
  // The single-CTA counting sort (16 us) is launched on a side stream right after the env step whose costs it reads, so it runs
  // beside the policy forward instead of in front of the next env step; every env launch first waits for a pending sort (it reads
  // the cost / autoreset words this launch writes).  PB_ENV_ORDER_ASYNC=0: in line, as before.
  static const int order_async = getenv("PB_ENV_ORDER_ASYNC") ? atoi(getenv("PB_ENV_ORDER_ASYNC")) : 1;
  const bool had_order = e->order_pending;
  if (e->order_pending) { cudaStreamWaitEvent(e->stream, e->order_done, 0); e->order_pending = false; }
  if (a.ordered && !had_order) env_order_kernel<<<1, 1024, 0, e->stream>>>(e->eb);
  env_step_kernel<C, WPB, LOCKSTEP, SYNCMASK, MSM><<<grid, WPB * 32, smem, e->stream>>>(e->d_model, e->eb, a);
  if (a.ordered && order_async && e->order_stream) {
    cudaEventRecord(e->step_done, e->stream);
    cudaStreamWaitEvent(e->order_stream, e->step_done, 0);
    env_order_kernel<<<1, 1024, 0, e->order_stream>>>(e->eb);
    cudaEventRecord(e->order_done, e->order_stream);
    e->order_pending = true;
  }
  return cudaGetLastError();
}

cudaError_t dispatch_impl(const pb_env *e, const StepArgs &a) {
  // CTA shapes measured in round 1 (profiles/r01_cta_variants.txt): the warps of a CTA run in lockstep (one instruction
  // stream), so the shape trades barrier imbalance against instruction-fetch sharing.  PB_ENV_TUNING=1 selects the
  // free-running 4-warp CTAs (no lockstep) for A/B measurements.
  static const int tuning = getenv("PB_ENV_TUNING") ? atoi(getenv("PB_ENV_TUNING")) : 0;
  switch (e->task) {
    case PB_TASK_HALFCHEETAH:
      if (tuning == 3) return launch<CfgHalfCheetah, 12, true>(e, a);   // A/B: two lockstep CTAs of 12 warps
      if (tuning == 4) return launch<CfgHalfCheetah, 8, true>(e, a);    // A/B: three of 8
      // two lockstep CTAs of 12 warps per SM: measured against 1 x 24, 3 x 8 and free-running 4-warp CTAs at the ac_ppo workload
      // (rollout 108.2-108.8 ms against 111.1-112.0 / 113.7 / 133.1 ms, profiles/r02_env_balance_ab.txt)
      return tuning == 1 ? launch<CfgHalfCheetah, 4, false>(e, a) : launch<CfgHalfCheetah, 12, true>(e, a);
    case PB_TASK_HOPPER:
      return tuning == 1 ? launch<CfgHopper, 4, false>(e, a) : launch<CfgHopper, 8, true>(e, a);
    case PB_TASK_ANT:
      return tuning == 1 ? launch<CfgAnt, 4, false>(e, a) : launch<CfgAnt, 16, true>(e, a);   // one lockstep CTA of 16 warps per SM
    default: {
      // CTA shape (PB_HUM_WPB selects it for A/B measurements).  Measured at the walking state (profiles/r02_wpb_ab.txt): 2 x 8 warps
      // 3.63 ms, 2 x 9 4.01, 2 x 10 3.83 (96 registers), 1 x 16 3.72, free-running 4-warp CTAs 7.09 (instruction-cache thrash)
      static const int wpb = getenv("PB_HUM_WPB") ? atoi(getenv("PB_HUM_WPB")) : 8;
      if (tuning == 1) return launch<CfgHumanoid, 4, false>(e, a);
      if (tuning == 2) return launch<CfgHumanoid, 16, true>(e, a);
      switch (wpb) {
        case 4: return launch<CfgHumanoid, 4, true>(e, a);    // 4 CTAs x 4 warps: smaller lockstep groups, four instruction streams
        default: return launch<CfgHumanoid, 8, true>(e, a);   // 2 CTAs x 8 warps per SM, 128 registers, no spills
      }
    }
  }
}

template <class T>
int dev_alloc(pb_env *e, T **p, size_t n, bool zero = true) {
  CK(cudaMalloc((void **)p, n * sizeof(T)));
  e->allocs.push_back(*p);
  if (zero) CK(cudaMemsetAsync(*p, 0, n * sizeof(T), e->stream));
  return 0;
}
}  // namespace

cudaError_t pb_env_dispatch(const pb_env *e, const StepArgs &a) { return dispatch_impl(e, a); }

extern "C" int pb_env_create(const char *env_id, int num_envs, int device, int chain, float gamma, int clip_actions,
                             pb_env **out) {
  if (!env_id || !out || num_envs <= 0) return fail(PB_ERR_ARG, "pb_env_create: bad argument");
  const TaskInfo *ti = nullptr;
  for (const auto &t : kTasks) if (!strcmp(t.id, env_id)) ti = &t;
  if (!ti) return fail(PB_ERR_BAD_ENV, std::string("Env not implemented: ") + env_id);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return fail(PB_ERR_BAD_DEVICE, "pb_env_create: CUDA device " + std::to_string(device) + " not available");
  CK(cudaSetDevice(device));
  DevModel hm;
  if (pb_build_devmodel(ti->model, &hm) != 0) return fail(PB_ERR_MODEL, "pb_build_devmodel failed");
  bool ok = ti->task == PB_TASK_HALFCHEETAH ? dims_match<CfgHalfCheetah>(hm)
            : ti->task == PB_TASK_HOPPER    ? dims_match<CfgHopper>(hm)
            : ti->task == PB_TASK_ANT       ? dims_match<CfgAnt>(hm)
                                            : dims_match<CfgHumanoid>(hm);
  if (!ok) return fail(PB_ERR_MODEL, std::string("An incorrect model is compiled in for ") + env_id);
  pb_env *e = new pb_env();
  e->task = ti->task; e->num_envs = num_envs; e->device = device; e->chain = chain; e->clip_actions = clip_actions;
  e->nq = hm.nq; e->nv = hm.nv; e->nu = hm.nu; e->obs_dim = ti->obs; e->gamma = gamma; e->act_min = ti->lo; e->act_max = ti->hi;
  CK(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&e->order_stream, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&e->step_done, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&e->order_done, cudaEventDisableTiming));
  CK(cudaMalloc((void **)&e->d_model, sizeof(DevModel)));
  e->allocs.push_back(e->d_model);
  CK(cudaMemcpyAsync(e->d_model, &hm, sizeof(DevModel), cudaMemcpyHostToDevice, e->stream));
  CK(cudaStreamSynchronize(e->stream));  // hm is a stack object
  const size_t N = num_envs;
  EnvBuffers &b = e->eb;
  b.num_envs = num_envs;
  int rc = 0;
  rc |= dev_alloc(e, &b.qpos, N * hm.nq); rc |= dev_alloc(e, &b.qvel, N * hm.nv); rc |= dev_alloc(e, &b.warm, N * hm.nv);
  rc |= dev_alloc(e, &b.xy_base, N * 2); rc |= dev_alloc(e, &b.xref, N);
  rc |= dev_alloc(e, &b.elapsed, N); rc |= dev_alloc(e, &b.autoreset, N); rc |= dev_alloc(e, &b.ep_length, N);
  rc |= dev_alloc(e, &b.ep_return, N);
  rc |= dev_alloc(e, &b.ob_mean, N * ti->obs); rc |= dev_alloc(e, &b.ob_var, N * ti->obs); rc |= dev_alloc(e, &b.ob_count, N);
  rc |= dev_alloc(e, &b.rw_mean, N); rc |= dev_alloc(e, &b.rw_var, N); rc |= dev_alloc(e, &b.rw_acc, N); rc |= dev_alloc(e, &b.rw_count, N);
  rc |= dev_alloc(e, &b.mt, N * 624); rc |= dev_alloc(e, &b.mt_idx, N); rc |= dev_alloc(e, &b.nrm_ok, N); rc |= dev_alloc(e, &b.nrm_saved, N);
  rc |= dev_alloc(e, &b.counters, 32); rc |= dev_alloc(e, &b.work, 2);
  rc |= dev_alloc(e, &b.cost, N); rc |= dev_alloc(e, &b.order, N);
  rc |= dev_alloc(e, &e->d_act, N * hm.nu); rc |= dev_alloc(e, &e->d_obs, N * ti->obs); rc |= dev_alloc(e, &e->d_rew, N);
  rc |= dev_alloc(e, &e->d_term, N); rc |= dev_alloc(e, &e->d_trunc, N); rc |= dev_alloc(e, &e->d_info_r, N);
  rc |= dev_alloc(e, &e->d_has, N); rc |= dev_alloc(e, &e->d_info_l, N);
  if (rc) { pb_env_destroy(e); return PB_ERR_CUDA; }
  // initial values of the reference objects: elapsed_step_ = 1001 (half_cheetah_v5.h:29), var = 1, count = eps
  // (stateful_observation.h:25-28, stateful_reward.h:27-33), default-constructed mt19937 (seed 5489)
  {
    std::vector<float> ones(N * ti->obs, 1.f), c1(N, 1e-4f), c2(N, 1e-8f), v1(N, 1.f);
    std::vector<int> el(N, 1001), idx(N, 624);
    std::vector<uint32_t> mt(N * 624);
    uint32_t x = 5489u;
    mt[0] = x;
    for (int i = 1; i < 624; i++) { x = 1812433253u * (x ^ (x >> 30)) + (uint32_t)i; mt[i] = x; }
    for (size_t n = 1; n < N; n++) memcpy(&mt[n * 624], &mt[0], 624 * sizeof(uint32_t));
    std::vector<float> q0(N * hm.nq);
    for (size_t n = 0; n < N; n++) for (int i = 0; i < hm.nq; i++) q0[n * hm.nq + i] = hm.qpos0[i];
    CK(cudaMemcpy(b.ob_var, ones.data(), ones.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b.ob_count, c1.data(), N * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b.rw_count, c2.data(), N * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b.rw_var, v1.data(), N * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b.elapsed, el.data(), N * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b.mt_idx, idx.data(), N * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b.mt, mt.data(), mt.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b.qpos, q0.data(), q0.size() * 4, cudaMemcpyHostToDevice));
  }
  size_t pin = N * (hm.nu + ti->obs + 6) * sizeof(float);
  CK(cudaMallocHost((void **)&e->h_pin, pin));
  CK(cudaStreamSynchronize(e->stream));
  *out = e;
  return PB_OK;
}

extern "C" int pb_env_destroy(pb_env *e) {
  if (!e) return PB_OK;
  cudaSetDevice(e->device);
  cudaStreamSynchronize(e->stream);
  if (e->order_stream) { cudaStreamSynchronize(e->order_stream); cudaStreamDestroy(e->order_stream); }
  if (e->step_done) cudaEventDestroy(e->step_done);
  if (e->order_done) cudaEventDestroy(e->order_done);
  for (void *p : e->allocs) cudaFree(p);
  if (e->h_pin) cudaFreeHost(e->h_pin);
  cudaStreamDestroy(e->stream);
  delete e;
  return PB_OK;
}

extern "C" int pb_env_num_envs(const pb_env *e) { return e->num_envs; }
extern "C" int pb_env_obs_dim(const pb_env *e) { return e->obs_dim; }
extern "C" int pb_env_act_dim(const pb_env *e) { return e->nu; }
extern "C" float pb_env_act_min(const pb_env *e) { return e->act_min; }
extern "C" float pb_env_act_max(const pb_env *e) { return e->act_max; }
extern "C" void *pb_env_stream(const pb_env *e) { return (void *)e->stream; }
extern "C" int pb_env_model_dims(const pb_env *e, int *nq, int *nv, int *nu) {
  if (nq) *nq = e->nq;
  if (nv) *nv = e->nv;
  if (nu) *nu = e->nu;
  return PB_OK;
}

StepArgs pb_env_make_args(const pb_env *e, int mode, int seed) {
  StepArgs a;
  memset(&a, 0, sizeof(a));
  a.chain = e->chain; a.clip_actions = e->clip_actions; a.act_min = e->act_min; a.act_max = e->act_max;
  a.gamma = e->gamma; a.mode = mode; a.seed = seed;
  return a;
}

extern "C" int pb_env_reset(pb_env *e, int seed, float *obs_out) {
  if (!e || !obs_out) return fail(PB_ERR_ARG, "pb_env_reset: null argument");
  CK(cudaSetDevice(e->device));
  StepArgs a = pb_env_make_args(e, 1, seed);
  a.obs = obs_out;
  CK(pb_env_dispatch(e, a));
  return PB_OK;
}

extern "C" int pb_env_step(pb_env *e, const float *actions, float *obs_out, float *rew_out, float *term_out,
                           float *trunc_out, int *has_info, float *info_r, int *info_l) {
  if (!e || !actions || !obs_out || !rew_out || !term_out || !trunc_out) return fail(PB_ERR_ARG, "pb_env_step: null argument");
  CK(cudaSetDevice(e->device));
  StepArgs a = pb_env_make_args(e, 0, 0);
  a.actions = actions; a.obs = obs_out; a.rew = rew_out; a.term = term_out; a.trunc = trunc_out;
  a.has_info = has_info ? has_info : e->d_has; a.info_r = info_r ? info_r : e->d_info_r; a.info_l = info_l ? info_l : e->d_info_l;
  CK(pb_env_dispatch(e, a));
  return PB_OK;
}

extern "C" int pb_env_forward(pb_env *e, float *raw_obs_out) {
  if (!e) return fail(PB_ERR_ARG, "pb_env_forward: null argument");
  CK(cudaSetDevice(e->device));
  StepArgs a = pb_env_make_args(e, 2, 0);
  a.obs = raw_obs_out;
  CK(pb_env_dispatch(e, a));
  return PB_OK;
}

extern "C" int pb_env_sync(pb_env *e) {
  CK(cudaSetDevice(e->device));
  CK(cudaStreamSynchronize(e->stream));
  return PB_OK;
}

extern "C" int pb_env_reset_host(pb_env *e, int seed, float *obs_out) {
  int rc = pb_env_reset(e, seed, e->d_obs);
  if (rc) return rc;
  size_t nb = (size_t)e->num_envs * e->obs_dim * sizeof(float);
  CK(cudaMemcpyAsync(e->h_pin, e->d_obs, nb, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  memcpy(obs_out, e->h_pin, nb);
  return PB_OK;
}

extern "C" int pb_env_step_host(pb_env *e, const float *actions, float *obs_out, float *rew_out, float *term_out,
                                float *trunc_out, int *has_info, float *info_r, int *info_l) {
  if (!e || !actions) return fail(PB_ERR_ARG, "pb_env_step_host: null argument");
  CK(cudaSetDevice(e->device));
  const size_t N = e->num_envs, na = N * e->nu, no = N * e->obs_dim;
  float *h_act = e->h_pin, *h_out = e->h_pin + na;  // out: obs | rew | term | trunc | info_r | (int) has | info_l
  memcpy(h_act, actions, na * sizeof(float));
  CK(cudaMemcpyAsync(e->d_act, h_act, na * sizeof(float), cudaMemcpyHostToDevice, e->stream));
  int rc = pb_env_step(e, e->d_act, e->d_obs, e->d_rew, e->d_term, e->d_trunc, e->d_has, e->d_info_r, e->d_info_l);
  if (rc) return rc;
  CK(cudaMemcpyAsync(h_out, e->d_obs, no * 4, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaMemcpyAsync(h_out + no, e->d_rew, N * 4, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaMemcpyAsync(h_out + no + N, e->d_term, N * 4, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaMemcpyAsync(h_out + no + 2 * N, e->d_trunc, N * 4, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaMemcpyAsync(h_out + no + 3 * N, e->d_info_r, N * 4, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaMemcpyAsync(h_out + no + 4 * N, e->d_has, N * 4, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaMemcpyAsync(h_out + no + 5 * N, e->d_info_l, N * 4, cudaMemcpyDeviceToHost, e->stream));
  CK(cudaStreamSynchronize(e->stream));
  if (obs_out) memcpy(obs_out, h_out, no * 4);
  if (rew_out) memcpy(rew_out, h_out + no, N * 4);
  if (term_out) memcpy(term_out, h_out + no + N, N * 4);
  if (trunc_out) memcpy(trunc_out, h_out + no + 2 * N, N * 4);
  if (info_r) memcpy(info_r, h_out + no + 3 * N, N * 4);
  if (has_info) memcpy(has_info, h_out + no + 4 * N, N * 4);
  if (info_l) memcpy(info_l, h_out + no + 5 * N, N * 4);
  return PB_OK;
}

extern "C" int pb_env_get_state(pb_env *e, int i, double *qpos, double *qvel, double *warm) {
  if (!e || i < 0 || i >= e->num_envs) return fail(PB_ERR_ARG, "pb_env_get_state: bad index");
  CK(cudaSetDevice(e->device));
  CK(cudaStreamSynchronize(e->stream));
  std::vector<float> q(e->nq), v(e->nv), w(e->nv);
  double base[2];
  CK(cudaMemcpy(q.data(), e->eb.qpos + (size_t)i * e->nq, e->nq * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(v.data(), e->eb.qvel + (size_t)i * e->nv, e->nv * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(w.data(), e->eb.warm + (size_t)i * e->nv, e->nv * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(base, e->eb.xy_base + (size_t)i * 2, 16, cudaMemcpyDeviceToHost));
  int nxy = (e->task == PB_TASK_HALFCHEETAH || e->task == PB_TASK_HOPPER) ? 1 : 2;
  for (int k = 0; k < e->nq; k++) if (qpos) qpos[k] = (double)q[k] + (k < nxy ? base[k] : 0.0);
  for (int k = 0; k < e->nv; k++) { if (qvel) qvel[k] = v[k]; if (warm) warm[k] = w[k]; }
  return PB_OK;
}

extern "C" int pb_env_set_state(pb_env *e, int i, const double *qpos, const double *qvel, const double *warm) {
  if (!e || i < 0 || i >= e->num_envs || !qpos || !qvel) return fail(PB_ERR_ARG, "pb_env_set_state: bad argument");
  CK(cudaSetDevice(e->device));
  CK(cudaStreamSynchronize(e->stream));
  std::vector<float> q(e->nq), v(e->nv), w(e->nv, 0.f);
  int nxy = (e->task == PB_TASK_HALFCHEETAH || e->task == PB_TASK_HOPPER) ? 1 : 2;
  double base[2] = {0.0, 0.0};
  for (int k = 0; k < e->nq; k++) { q[k] = (float)qpos[k]; if (k < nxy) { base[k] = qpos[k]; q[k] = 0.f; } }
  for (int k = 0; k < e->nv; k++) { v[k] = (float)qvel[k]; if (warm) w[k] = (float)warm[k]; }
  CK(cudaMemcpy(e->eb.qpos + (size_t)i * e->nq, q.data(), e->nq * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(e->eb.qvel + (size_t)i * e->nv, v.data(), e->nv * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(e->eb.warm + (size_t)i * e->nv, w.data(), e->nv * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(e->eb.xy_base + (size_t)i * 2, base, 16, cudaMemcpyHostToDevice));
  return PB_OK;
}

extern "C" int pb_env_solver_counters(pb_env *e, unsigned long long *c) {
  CK(cudaSetDevice(e->device));
  CK(cudaStreamSynchronize(e->stream));
  CK(cudaMemcpy(c, e->eb.counters + 2, 16, cudaMemcpyDeviceToHost));
  if (getenv("PB_IMBALANCE_PRINT")) {   // debug build with -DPB_IMBALANCE: lockstep imbalance of the env kernel
    unsigned long long d[32];
    CK(cudaMemcpy(d, e->eb.counters, sizeof(d), cudaMemcpyDeviceToHost));
    const char *names[8] = {"kinematics+com", "collision", "crb+factor", "make_constraint", "velocity+smooth", "solver", "post stage", "whole evaluation"};
    if (d[6]) for (int k = 0; k < 8; k++)
      fprintf(stderr, "imbalance %-18s mean warp %7.0f cycles, slowest of the CTA %7.0f (+%5.0f)\n", names[k], (double)d[9 + 2 * k] / d[6] / 8.0,
              (double)d[8 + 2 * k] / d[6], (double)d[8 + 2 * k] / d[6] - (double)d[9 + 2 * k] / d[6] / 8.0);
    if (d[0]) fprintf(stderr, "solver per evaluation: set-up %.0f cycles, sweeps %.0f cycles (%.1f sweeps, %.1f row steps -> %.0f cycles per sweep)\n",
                      (double)d[24] / d[0], (double)d[25] / d[0], (double)d[2] / d[0], (double)d[3] / d[0], (double)d[25] / (double)d[2]);
  }
  return PB_OK;
}

extern "C" int pb_env_counters(pb_env *e, unsigned long long *c) {
  CK(cudaSetDevice(e->device));
  CK(cudaStreamSynchronize(e->stream));
  CK(cudaMemcpy(c, e->eb.counters, 16, cudaMemcpyDeviceToHost));
  return PB_OK;
}

// NormalizeObservation / NormalizeReward statistics of env i (stateful_observation.h:25-28, stateful_reward.h:27-33):
// read one env's, or overwrite one env's (i >= 0) / every env's (i < 0) with the given values.  rw4 = mean, var, acc, count.
extern "C" int pb_env_get_norm_stats(pb_env *e, int i, float *ob_mean, float *ob_var, float *ob_count, float *rw4) {
  if (!e || i < 0 || i >= e->num_envs) return fail(PB_ERR_ARG, "pb_env_get_norm_stats: bad index");
  CK(cudaSetDevice(e->device));
  CK(cudaStreamSynchronize(e->stream));
  const size_t O = e->obs_dim;
  if (ob_mean) CK(cudaMemcpy(ob_mean, e->eb.ob_mean + i * O, O * 4, cudaMemcpyDeviceToHost));
  if (ob_var) CK(cudaMemcpy(ob_var, e->eb.ob_var + i * O, O * 4, cudaMemcpyDeviceToHost));
  if (ob_count) CK(cudaMemcpy(ob_count, e->eb.ob_count + i, 4, cudaMemcpyDeviceToHost));
  if (rw4) {
    CK(cudaMemcpy(rw4 + 0, e->eb.rw_mean + i, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(rw4 + 1, e->eb.rw_var + i, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(rw4 + 2, e->eb.rw_acc + i, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(rw4 + 3, e->eb.rw_count + i, 4, cudaMemcpyDeviceToHost));
  }
  return PB_OK;
}

extern "C" int pb_env_set_norm_stats(pb_env *e, int i, const float *ob_mean, const float *ob_var, float ob_count, const float *rw4) {
  if (!e || i >= e->num_envs || !ob_mean || !ob_var) return fail(PB_ERR_ARG, "pb_env_set_norm_stats: bad argument");
  CK(cudaSetDevice(e->device));
  CK(cudaStreamSynchronize(e->stream));
  const size_t O = e->obs_dim, lo = i < 0 ? 0 : i, hi = i < 0 ? e->num_envs : i + 1, n = hi - lo;
  std::vector<float> m(n * O), v(n * O), c(n, ob_count);
  for (size_t k = 0; k < n; k++) { memcpy(&m[k * O], ob_mean, O * 4); memcpy(&v[k * O], ob_var, O * 4); }
  CK(cudaMemcpy(e->eb.ob_mean + lo * O, m.data(), n * O * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(e->eb.ob_var + lo * O, v.data(), n * O * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(e->eb.ob_count + lo, c.data(), n * 4, cudaMemcpyHostToDevice));
  if (rw4) {
    float *dst[4] = {e->eb.rw_mean, e->eb.rw_var, e->eb.rw_acc, e->eb.rw_count};
    for (int f = 0; f < 4; f++) {
      std::fill(c.begin(), c.end(), rw4[f]);
      CK(cudaMemcpy(dst[f] + lo, c.data(), n * 4, cudaMemcpyHostToDevice));
    }
  }
  return PB_OK;
}
