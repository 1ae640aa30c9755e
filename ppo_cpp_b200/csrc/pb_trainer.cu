// pb_trainer.cu — fused rollout-and-update engine behind the pb_trainer_* C ABI (include/ppo_b200.h): the iteration
// loop of src/ppo_continuous_action.cpp:375-586 on one GPU, with an optional NCCL gradient average across replicas.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <chrono>
#include <vector>

#include "pb_internal.h"
#include "pb_ppo_fused.cuh"
#include "pb_ppo_tc.cuh"
#include "pb_ppo_rollout.cuh"

#define PB_SPLITS_DEFAULT 32

#include "pb_nccl.h"
#define PB_PEER_HOST
#include "pb_peer.cuh"

struct pb_rng {
  std::mt19937 gen;
};

struct pb_trainer {
  pb_ppo_config cfg;
  std::string env_id;
  pb_env *env = nullptr;
  cudaStream_t stream;
  AgentLayout L;
  PpoHyper hp;
  int N, T, O, A, B, mb, P, nblk_loss, npart, splits, tm;
  // parameters and optimiser state
  float *params, *adam_m, *adam_v, *grad, *W1T, *W2T, *Wc, *Hp2;
  float *Hr1 = nullptr, *Hr2 = nullptr;   // first-layer pre-activations of the rollout forward from the tensor-core kernel, [N][128] per k-range
  // rollout storage [T][N][.]
  float *obs, *actions, *logprobs, *rewards, *dones, *values, *advantages, *returns;
  float *next_obs, *term, *trunc, *next_value;
  // update scratch
  float *H1, *H2, *D1, *D2, *D3, *slabs, *part, *partials, *adv_stats, *scalars, *mb_stats, *ep_stat, *colsum;
  float *Xp, *actp; float4 *scp;  // epoch-ordered gathered rows [B][O], [B][A], [B]
  int *idx_all, *cursor, *h_idx = nullptr;
  float *clip_acc;  // running sum of clipfrac over the minibatches of one update
  std::vector<void *> allocs;
  pb_rng perm_rng;
  uint32_t step_index = 0;
  cudaEvent_t ev[4];
  cudaStream_t side = nullptr;          // dw1_tc_kernel runs here, beside wgrad_kernel on the main stream
  cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
  std::vector<cudaEvent_t> idx_ev;   // per-epoch: upload of that epoch's permutation slot has completed
  std::vector<char> idx_ev_used;
  float ms_rollout = 0, ms_gae = 0, ms_update = 0;
  int clip_count = 0;  // minibatches whose clipfrac clip_acc holds (zeroed at the start of every update)
  ncclComm_t comm = nullptr;
  int world = 1, rank = 0;
  cudaGraphExec_t mb_graph = nullptr;   // the num_minibatches steps of one epoch, captured once (launch arguments never change)
  PeerExchange peer;   // fused peer-memory gradient average (pb_peer.cuh); falls back to ncclAllReduce when unavailable
};

namespace {
template <class T>
int talloc(pb_trainer *t, T **p, size_t n) {
  CK(cudaMalloc((void **)p, n * sizeof(T)));
  t->allocs.push_back(*p);
  CK(cudaMemsetAsync(*p, 0, n * sizeof(T), t->stream));
  return 0;
}

// nn::init::orthogonal_(w[rows x cols], gain): QR of a Gaussian matrix with the sign fix, rows >= cols after the
// transpose rule torch applies (src/ppo_continuous_action.cpp:159-164).  Distribution-equivalent to torch, not
// stream-equivalent (torch draws from its global generator).
void orthogonal_init(float *w, int rows, int cols, double gain, std::mt19937 &gen) {
  std::normal_distribution<double> nd(0.0, 1.0);
  bool transpose = rows < cols;
  int r = transpose ? cols : rows, c = transpose ? rows : cols;  // r >= c
  std::vector<double> a((size_t)r * c);
  for (auto &x : a) x = nd(gen);
  // modified Gram-Schmidt on the columns; R's diagonal is positive, which is torch's q *= sign(diag(r))
  for (int j = 0; j < c; j++) {
    for (int k = 0; k < j; k++) {
      double d = 0;
      for (int i = 0; i < r; i++) d += a[(size_t)i * c + k] * a[(size_t)i * c + j];
      for (int i = 0; i < r; i++) a[(size_t)i * c + j] -= d * a[(size_t)i * c + k];
    }
    double n = 0;
    for (int i = 0; i < r; i++) n += a[(size_t)i * c + j] * a[(size_t)i * c + j];
    n = std::sqrt(n);
    for (int i = 0; i < r; i++) a[(size_t)i * c + j] /= n;
  }
  for (int i = 0; i < rows; i++)
    for (int j = 0; j < cols; j++) w[(size_t)i * cols + j] = (float)(gain * (transpose ? a[(size_t)j * c + i] : a[(size_t)i * c + j]));
}
}  // namespace

extern "C" void pb_ppo_config_default(pb_ppo_config *c) {
  memset(c, 0, sizeof(*c));
  c->env_id = "Humanoid-v4"; c->num_envs = 1; c->num_steps = 2048; c->device = 0; c->seed = 1;
  c->gamma = 0.99f; c->gae_lambda = 0.95f; c->num_minibatches = 32; c->update_epochs = 10; c->norm_adv = 1;
  c->clip_coef = 0.2f; c->clip_vloss = 1; c->ent_coef = 0.0f; c->vf_coef = 0.5f; c->max_grad_norm = 0.5f;
  c->adam_eps = 1e-5f; c->clip_actions = 1; c->rank = 0; c->world_size = 1;
}

extern "C" pb_rng *pb_rng_create(unsigned long long seed) {
  pb_rng *r = new pb_rng();
  r->gen.seed((uint32_t)seed);
  return r;
}
extern "C" void pb_rng_destroy(pb_rng *r) { delete r; }
void pb_randperm_i64(std::mt19937 &gen, int n, long long *out) {
  for (int i = 0; i < n; i++) out[i] = i;
  for (int i = 0; i < n - 1; i++) {
    long long z = (long long)(gen() % (uint32_t)(n - i));
    long long tmp = out[i]; out[i] = out[z + i]; out[z + i] = tmp;
  }
}
// DD-PPO index rule (src/ac_ppo_continuous_action.cpp:804-810): b_inds = randperm(num_collected); when fewer samples than
// batch were collected, b_inds.repeat(ceil(batch / num_collected)) truncated to batch.  One permutation is drawn either way.
void pb_dd_ppo_indices(std::mt19937 &gen, int num_collected, int batch, long long *out) {
  if (num_collected >= batch) { pb_randperm_i64(gen, batch, out); return; }   // not pre-empted: num_collected == batch
  std::vector<long long> p(num_collected);
  pb_randperm_i64(gen, num_collected, p.data());
  for (int i = 0; i < batch; i++) out[i] = p[i % num_collected];
}
extern "C" int pb_dd_ppo_batch_indices(pb_rng *r, int num_collected, int batch, long long *out) {
  if (!r || num_collected < 1 || batch < 1 || num_collected > batch || !out) return pb_fail(PB_ERR_ARG, "pb_dd_ppo_batch_indices: bad argument");
  pb_dd_ppo_indices(r->gen, num_collected, batch, out);
  return PB_OK;
}
// the pre-emption test of a collector after `step` steps (:680-688)
extern "C" int pb_dd_ppo_should_preempt(int num_done, int num_envs, float threshold, int step, float min_perc, int num_steps) {
  const long min_steps = lroundf(min_perc * (float)num_steps);
  return ((float)num_done / (float)num_envs > threshold && step > min_steps) ? 1 : 0;
}
extern "C" int pb_rng_randperm(pb_rng *r, int n, long long *out) {
  if (!r || n < 0 || !out) return pb_fail(PB_ERR_ARG, "pb_rng_randperm: bad argument");
  pb_randperm_i64(r->gen, n, out);
  return PB_OK;
}

extern "C" int pb_trainer_create(const pb_ppo_config *cfg, pb_trainer **out) {
  if (!cfg || !out || !cfg->env_id) return pb_fail(PB_ERR_ARG, "pb_trainer_create: null argument");
  if (cfg->num_steps <= 0 || cfg->num_minibatches <= 0 || cfg->update_epochs < 0)
    return pb_fail(PB_ERR_ARG, "pb_trainer_create: bad sizes");
  long long batch = (long long)cfg->num_envs * cfg->num_steps;
  if (batch % cfg->num_minibatches) return pb_fail(PB_ERR_ARG, "batch_size must be divisible by num_minibatches");
  pb_trainer *t = new pb_trainer();
  t->cfg = *cfg; t->env_id = cfg->env_id; t->cfg.env_id = t->env_id.c_str();
  int rc = pb_env_create(cfg->env_id, cfg->num_envs, cfg->device, PB_CHAIN_PPO, cfg->gamma, cfg->clip_actions, &t->env);
  if (rc) { delete t; return rc; }
  t->stream = t->env->stream;
  t->N = cfg->num_envs; t->T = cfg->num_steps; t->O = t->env->obs_dim; t->A = t->env->nu;
  t->B = (int)batch; t->mb = t->B / cfg->num_minibatches;
  t->L = AgentLayout::make(t->O, t->A); t->P = t->L.total;
  t->hp = {cfg->clip_coef, cfg->ent_coef, cfg->vf_coef, cfg->max_grad_norm, cfg->adam_eps, 0.9f, 0.999f, cfg->norm_adv, cfg->clip_vloss};
  t->tm = getenv("PB_FB_TM") ? atoi(getenv("PB_FB_TM")) : 64;
  if (t->tm != 32) t->tm = 64;
  t->nblk_loss = (t->mb + t->tm - 1) / t->tm;
  t->splits = getenv("PB_SPLITS") ? std::max(1, atoi(getenv("PB_SPLITS"))) : PB_SPLITS_DEFAULT;
  if (cfg->num_envs < 1 || t->env->nu + 1 > FB_MAXA + 1) { pb_trainer_destroy(t); return pb_fail(PB_ERR_ARG, "action dimension too large"); }
  t->npart = (t->P + 255) / 256;
  t->rank = cfg->rank; t->world = cfg->world_size > 0 ? cfg->world_size : 1;
  const size_t N = t->N, T = t->T, O = t->O, A = t->A, P = t->P, mb = t->mb;
  int e = 0;
  e |= talloc(t, &t->params, P); e |= talloc(t, &t->adam_m, P); e |= talloc(t, &t->adam_v, P); e |= talloc(t, &t->grad, P);
  e |= talloc(t, &t->W1T, O * 128); e |= talloc(t, &t->W2T, 2 * 64 * 64);
  e |= talloc(t, &t->Wc, (size_t)2 * ((O + L1_BK - 1) / L1_BK) * L1_B_FLOATS); e |= talloc(t, &t->Hp2, (size_t)t->mb * 128);
  e |= talloc(t, &t->Hr1, (size_t)t->N * 128); e |= talloc(t, &t->Hr2, (size_t)t->N * 128);
  e |= talloc(t, &t->obs, T * N * O); e |= talloc(t, &t->actions, T * N * A); e |= talloc(t, &t->logprobs, T * N);
  e |= talloc(t, &t->rewards, T * N); e |= talloc(t, &t->dones, T * N); e |= talloc(t, &t->values, T * N);
  e |= talloc(t, &t->advantages, T * N); e |= talloc(t, &t->returns, T * N);
  e |= talloc(t, &t->next_obs, N * O); e |= talloc(t, &t->term, N); e |= talloc(t, &t->trunc, N); e |= talloc(t, &t->next_value, N);
  e |= talloc(t, &t->H1, mb * 128); e |= talloc(t, &t->H2, mb * 128); e |= talloc(t, &t->D1, mb * 128); e |= talloc(t, &t->D2, mb * 128);
  e |= talloc(t, &t->D3, mb * 64); e |= talloc(t, &t->slabs, (size_t)t->splits * P);
  e |= talloc(t, &t->Xp, T * N * O); e |= talloc(t, &t->actp, T * N * A); e |= talloc(t, &t->scp, T * N);
  e |= talloc(t, &t->part, (size_t)t->nblk_loss * (PB_NSTAT + A)); e |= talloc(t, &t->partials, t->npart);
  e |= talloc(t, &t->adv_stats, 2 * (size_t)cfg->num_minibatches); e |= talloc(t, &t->scalars, 4); e |= talloc(t, &t->mb_stats, PB_NSTAT);
  e |= talloc(t, &t->colsum, 64); e |= talloc(t, &t->ep_stat, 4); e |= talloc(t, &t->clip_acc, 1);
  e |= talloc(t, &t->idx_all, (size_t)std::max(1, cfg->update_epochs) * t->B); e |= talloc(t, &t->cursor, 1);
  if (e) { pb_trainer_destroy(t); return PB_ERR_CUDA; }
  CK(cudaMallocHost((void **)&t->h_idx, (size_t)std::max(1, cfg->update_epochs) * t->B * sizeof(int)));
  for (auto &ev : t->ev) CK(cudaEventCreate(&ev));
  CK(cudaStreamCreateWithFlags(&t->side, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&t->fork_ev, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&t->join_ev, cudaEventDisableTiming));
  t->idx_ev.resize(std::max(1, cfg->update_epochs)); t->idx_ev_used.assign(t->idx_ev.size(), 0);
  for (auto &ev : t->idx_ev) CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  // agent init: orthogonal(sqrt 2) hidden, 1.0 critic head, 0.01 actor head, zero biases and logstd (:124-137)
  {
    std::vector<float> h(P, 0.f);
    std::mt19937 gen((uint32_t)cfg->seed);
    const double g2 = std::sqrt(2.0);
    orthogonal_init(&h[t->L.cW1], 64, t->O, g2, gen); orthogonal_init(&h[t->L.cW2], 64, 64, g2, gen);
    orthogonal_init(&h[t->L.cW3], 1, 64, 1.0, gen);
    orthogonal_init(&h[t->L.aW1], 64, t->O, g2, gen); orthogonal_init(&h[t->L.aW2], 64, 64, g2, gen);
    orthogonal_init(&h[t->L.aW3], t->A, 64, 0.01, gen);
    CK(cudaMemcpy(t->params, h.data(), P * sizeof(float), cudaMemcpyHostToDevice));
  }
  t->perm_rng.gen.seed((uint32_t)cfg->seed);
  CK(cudaStreamSynchronize(t->stream));
  *out = t;
  return PB_OK;
}

extern "C" int pb_trainer_destroy(pb_trainer *t) {
  if (!t) return PB_OK;
  if (t->env) { cudaSetDevice(t->env->device); cudaStreamSynchronize(t->stream); if (t->side) cudaStreamSynchronize(t->side); }
  // Data-parallel teardown: no rank may unmap / free its exported exchange buffer while a peer can still write to it (a rank
  // that evaluates longer than the peer barrier's own 5 s bound would otherwise find freed mappings - ADVICE r1).  A one-float
  // ncclAllReduce is the rendezvous: it returns only when every rank has reached its destroy call.
  if (t->env && t->comm && t->peer.on && g_nccl.AllReduce && t->scalars) {
    if (g_nccl.AllReduce(t->scalars, t->scalars, 1, ncclFloat, ncclSum, t->comm, t->stream) == ncclSuccess) cudaStreamSynchronize(t->stream);
  }
  if (t->env) t->peer.teardown(t->stream);
  if (t->mb_graph) cudaGraphExecDestroy(t->mb_graph);
  if (t->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(t->comm);
  for (void *p : t->allocs) cudaFree(p);
  if (t->h_idx) cudaFreeHost(t->h_idx);
  for (auto &ev : t->ev) if (ev) cudaEventDestroy(ev);
  if (t->fork_ev) cudaEventDestroy(t->fork_ev);
  if (t->join_ev) cudaEventDestroy(t->join_ev);
  if (t->side) cudaStreamDestroy(t->side);
  for (auto &ev : t->idx_ev) if (ev) cudaEventDestroy(ev);
  if (t->env) pb_env_destroy(t->env);
  delete t;
  return PB_OK;
}

extern "C" int pb_trainer_num_params(const pb_trainer *t) { return t->P; }
extern "C" void *pb_trainer_stream(const pb_trainer *t) { return (void *)t->stream; }
extern "C" pb_env *pb_trainer_env(pb_trainer *t) { return t->env; }
extern "C" int pb_trainer_sync(pb_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  return PB_OK;
}
extern "C" int pb_trainer_get_params(pb_trainer *t, float *h) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(h, t->params, t->P * sizeof(float), cudaMemcpyDeviceToHost));
  return PB_OK;
}
extern "C" int pb_trainer_set_params(pb_trainer *t, const float *h) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(t->params, h, t->P * sizeof(float), cudaMemcpyHostToDevice));
  return PB_OK;
}

// ------------------------------------------------------------------ rollout / GAE
static int launch_policy(pb_trainer *t, int row, int sample) {
  PolicyArgs a;
  memset(&a, 0, sizeof(a));
  a.params = t->params; a.W1T = t->W1T; a.W2T = t->W2T; a.L = t->L; a.num_envs = t->N;
  a.obs_in = t->next_obs; a.term_in = t->term; a.trunc_in = t->trunc;
  const size_t N = t->N;
  if (sample) {
    a.obs_row = t->obs + row * N * t->O; a.act_row = t->actions + row * N * t->A; a.logp_row = t->logprobs + row * N;
    a.val_row = t->values + row * N; a.done_row = t->dones + row * N;
  } else {
    a.val_row = t->next_value;
  }
  a.seed_lo = (uint32_t)t->cfg.seed; a.seed_hi = 0x9E3779B9u ^ (uint32_t)t->rank; a.step_index = t->step_index; a.sample = sample;
  size_t smem = ((size_t)t->O * PB_POL_TILE + 2 * 128 * PB_POL_TILE + PB_POL_TILE * (t->A + 1)) * sizeof(float);
  static bool cfgd[64] = {};   // function attributes are per device
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (!cfgd[dev & 63]) {
    CK(cudaFuncSetAttribute(policy_forward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    CK(cudaFuncSetAttribute(tc::l1_forward_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L1_SMEM_BYTES));
    cfgd[dev & 63] = true;
  }
  // First layer (72 % of the forward MACs) on tcgen05, fp32-accurate 3xTF32, with the kernel the update uses: t->Wc is current
  // (refresh_transposed at the start of the rollout, adam2_kernel after every optimiser step).  PB_POLICY_TC=0 or fewer than 128
  // envs: the FFMA layer inside policy_forward_kernel.
  static const int pol_tc = (getenv("PB_POLICY_TC") ? atoi(getenv("PB_POLICY_TC")) : 1) && (getenv("PB_UPDATE_MMA") ? atoi(getenv("PB_UPDATE_MMA")) == 2 : 1);
  if (pol_tc && t->N >= 128) {
    const int nch = (t->O + L1_BK - 1) / L1_BK, ksplit = std::min(L1_KSPLIT, nch);
    tc::L1Args l1 = {t->next_obs, nullptr, t->N, t->O, t->Wc, {t->Hr1, t->Hr2}, 0};
    tc::l1_forward_tc_kernel<<<dim3((t->N + 127) / 128, ksplit), L1_THREADS, L1_SMEM_BYTES, t->stream>>>(l1);
    a.hpre = t->Hr1; a.hpre2 = ksplit > 1 ? t->Hr2 : nullptr;
  }
  policy_forward_kernel<<<(t->N + PB_POL_TILE - 1) / PB_POL_TILE, PB_POL_THREADS, smem, t->stream>>>(a);
  CK(cudaGetLastError());
  return PB_OK;
}

static int refresh_transposed(pb_trainer *t) {
  int n = t->O * 128 + 2 * 64 * 64;
  transpose_weights_kernel<<<(n + 255) / 256, 256, 0, t->stream>>>(t->params, t->L, t->W1T, t->W2T);
  const int nch = (t->O + L1_BK - 1) / L1_BK;
  tc::w1_split_kernel<<<(nch * L1_B_FLOATS + 255) / 256, 256, 0, t->stream>>>(t->params, t->L.cW1, t->L.aW1, t->O, nch, t->Wc);
  CK(cudaGetLastError());
  return PB_OK;
}

extern "C" int pb_trainer_reset(pb_trainer *t, int seed) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaMemsetAsync(t->term, 0, t->N * sizeof(float), t->stream));
  CK(cudaMemsetAsync(t->trunc, 0, t->N * sizeof(float), t->stream));
  return pb_env_reset(t->env, seed, t->next_obs);
}

extern "C" int pb_trainer_rollout(pb_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaEventRecord(t->ev[0], t->stream));
  int rc = refresh_transposed(t);
  if (rc) return rc;
  const size_t N = t->N;
  for (int s = 0; s < t->T; s++) {
    if ((rc = launch_policy(t, s, 1))) return rc;
    StepArgs a = pb_env_make_args(t->env, 0, 0);
    a.actions = t->actions + s * N * t->A; a.obs = t->next_obs; a.rew = t->rewards + s * N; a.term = t->term; a.trunc = t->trunc;
    a.has_info = nullptr; a.ep_stat = t->ep_stat;
    CK(pb_env_dispatch(t->env, a));
    t->step_index++;
  }
  CK(cudaEventRecord(t->ev[1], t->stream));
  return PB_OK;
}

extern "C" int pb_trainer_gae(pb_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  int rc = launch_policy(t, 0, 0);  // next_value = V(next_obs) with the rollout weights
  if (rc) return rc;
  gae_kernel<<<(t->N + 127) / 128, 128, 0, t->stream>>>(t->T, t->N, t->rewards, t->values, t->dones, t->next_value, t->term,
                                                         t->trunc, t->cfg.gamma, t->cfg.gae_lambda, t->advantages, t->returns);
  CK(cudaGetLastError());
  CK(cudaEventRecord(t->ev[2], t->stream));
  return PB_OK;
}

// ------------------------------------------------------------------ epoch prologue + one minibatch (four launches)
static int epoch_prologue(pb_trainer *t, const int *idx_dev, int rows, int nmb) {
  GatherArgs g;
  g.idx = idx_dev; g.rows = rows; g.O = t->O; g.A = t->A;
  g.obs = t->obs; g.actions = t->actions; g.logprobs = t->logprobs; g.advantages = t->advantages; g.returns = t->returns;
  g.values = t->values; g.Xp = t->Xp; g.actp = t->actp; g.scp = t->scp;
  epoch_gather_kernel<<<(rows + 7) / 8, 256, 0, t->stream>>>(g);
  adv_stats_all_kernel<<<nmb, 256, 0, t->stream>>>(t->scp, t->mb, t->adv_stats);
  CK(cudaMemsetAsync(t->cursor, 0, sizeof(int), t->stream));
  CK(cudaGetLastError());
  return PB_OK;
}

// PB_UPDATE_TIMING=1 (debug): events at every kernel boundary of minibatches [64, 128) of an update plus host clocks
// around the permutation generator and the launch loop; pb_trainer_update prints the averages once the stream drains.
namespace {
constexpr int UT_SEG = 6, UT_FIRST = 64, UT_COUNT = 64;
struct UpdateTiming {
  int on = -1, mbi = 0;
  std::vector<cudaEvent_t> ev;      // [UT_COUNT][UT_SEG + 1]
  double host_perm_ms = 0, host_loop_ms = 0;
} g_ut;
inline void ut_tick(cudaStream_t st, int seg) {
  if (g_ut.on != 1) return;
  const int i = g_ut.mbi - UT_FIRST;
  if (i < 0 || i >= UT_COUNT) return;
  if (g_ut.ev.empty()) { g_ut.ev.resize((size_t)UT_COUNT * (UT_SEG + 1)); for (auto &e : g_ut.ev) cudaEventCreate(&e); }
  cudaEventRecord(g_ut.ev[(size_t)i * (UT_SEG + 1) + seg], st);
}
}  // namespace

// one-time function attributes of the update kernels (outside any stream capture)
static int configure_update_kernels() {
  static bool cfgd_dev[64] = {};   // function attributes are per device: a second trainer on another device needs them too
  int dev = 0;
  CK(cudaGetDevice(&dev));
  bool &cfgd = cfgd_dev[dev & 63];
  if (!cfgd) {
    CK(cudaFuncSetAttribute(fwd_bwd_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(FB_SMEM_FLOATS(32) * sizeof(float))));
    CK(cudaFuncSetAttribute(fwd_bwd_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(FB_SMEM_FLOATS(64) * sizeof(float))));
    CK(cudaFuncSetAttribute(fwd_bwd_kernel<64, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(FB_SMEM_FLOATS_MMA * sizeof(float))));
    CK(cudaFuncSetAttribute(wgrad_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(WG_SMEM_FLOATS_MMA * sizeof(float))));
    CK(cudaFuncSetAttribute(tc::l1_forward_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L1_SMEM_BYTES));
    CK(cudaFuncSetAttribute(tc::dw1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DW_SMEM_BYTES));
    // wgrad_small_kernel shares SMs with dw1_tc_kernel (192 KB of shared memory): both must want the same carve-out,
    // or the SM would have to drain before switching configuration
    CK(cudaFuncSetAttribute(wgrad_small_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    // PB_WG_CARVEOUT=1 (experiment): give wgrad_kernel the same maximum shared-memory carve-out as dw1_tc_kernel so that
    // CTAs of both can be resident on one SM without a carve-out change
    if (getenv("PB_WG_CARVEOUT") && atoi(getenv("PB_WG_CARVEOUT")))
      CK(cudaFuncSetAttribute(wgrad_kernel<2>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    cfgd = true;
  }
  return PB_OK;
}

static int minibatch_step(pb_trainer *t) {
  const int mb = t->mb, A = t->A;
  cudaStream_t st = t->stream;
  ut_tick(st, 0);
  { int rc = configure_update_kernels(); if (rc) return rc; }
  FwdBwdArgs f;
  f.params = t->params; f.W1T = t->W1T; f.W2T = t->W2T; f.L = t->L; f.hp = t->hp; f.Xp = t->Xp; f.actp = t->actp; f.scp = t->scp;
  f.adv_stats = t->adv_stats; f.cursor = t->cursor; f.mb = mb; f.H1 = t->H1; f.H2 = t->H2; f.D1 = t->D1; f.D2 = t->D2; f.D3 = t->D3;
  f.part = t->part;
  static int dbg_left = getenv("PB_FB_DEBUG") ? atoi(getenv("PB_FB_DEBUG")) : 0;
  f.debug = dbg_left > 0 ? (dbg_left--, 1) : 0;
  // PB_UPDATE_MMA: 0 fp32 FFMA first-layer products, 1 mma.sync 3xTF32 (slower than FFMA on B200: kept as a reference
  // point), 2 tcgen05 3xTF32 (first-layer forward on the tensor cores, accumulators in TMEM)
  static const int mma_mode = getenv("PB_UPDATE_MMA") ? atoi(getenv("PB_UPDATE_MMA")) : 2;
  const int use_mma = mma_mode == 1;
  f.Hpre = nullptr; f.Hpre2 = nullptr;
  if (mma_mode == 2) {
    const int nch = (t->O + L1_BK - 1) / L1_BK;   // t->Wc is kept current by refresh_transposed and adam2_kernel
    static int l1dbg = getenv("PB_L1_DEBUG") ? atoi(getenv("PB_L1_DEBUG")) : 0;
    const int ksplit = std::min(L1_KSPLIT, nch);
    tc::L1Args l1 = {t->Xp, t->cursor, mb, t->O, t->Wc, {t->H1, t->Hp2}, l1dbg > 0 ? (l1dbg--, 1) : 0};
    tc::l1_forward_tc_kernel<<<dim3((mb + 127) / 128, ksplit), L1_THREADS, L1_SMEM_BYTES, st>>>(l1);
    f.Hpre2 = ksplit > 1 ? t->Hp2 : nullptr;
    f.Hpre = t->H1;
  }
  ut_tick(st, 1);
  if (t->tm == 32) fwd_bwd_kernel<32><<<t->nblk_loss, FB_THREADS, FB_SMEM_FLOATS(32) * sizeof(float), st>>>(f);
  else if (use_mma) fwd_bwd_kernel<64, true><<<t->nblk_loss, FB_THREADS, FB_SMEM_FLOATS_MMA * sizeof(float), st>>>(f);
  else fwd_bwd_kernel<64><<<t->nblk_loss, FB_THREADS, FB_SMEM_FLOATS(64) * sizeof(float), st>>>(f);
  ut_tick(st, 2);
  WgradArgs w;
  w.Xp = t->Xp; w.cursor = t->cursor; w.mb = mb; w.splits = t->splits; w.L = t->L; w.H1 = t->H1; w.H2 = t->H2; w.D1 = t->D1;
  w.D2 = t->D2; w.D3 = t->D3; w.slabs = t->slabs; w.slab_stride = t->P;
  // PB_DW1_TC=0 keeps the first-layer weight gradient on the FFMA tiles
  static const int dw_tc = getenv("PB_DW1_TC") ? atoi(getenv("PB_DW1_TC")) : 1;
  const bool dw1_tc = dw_tc && mma_mode == 2;
  w.skip_w1 = dw1_tc ? 1 : 0;
  // PB_WG_OVERLAP=0 keeps both weight-gradient kernels on one stream
  static const int wg_overlap = getenv("PB_WG_OVERLAP") ? atoi(getenv("PB_WG_OVERLAP")) : 1;
  const bool forked = dw1_tc && wg_overlap;
  if (dw1_tc) {
    // the two kernels read D1 / X / H1 / H2 and write disjoint slab rows; dw1_tc_kernel fills 96 SMs, so they overlap
    cudaStream_t ds = forked ? t->side : st;
    if (forked) { CK(cudaEventRecord(t->fork_ev, st)); CK(cudaStreamWaitEvent(ds, t->fork_ev, 0)); }
    tc::DwArgs d = {t->Xp, t->D1, t->cursor, mb, t->O, t->splits, t->L.cW1, t->L.aW1, t->slabs, (size_t)t->P};
    tc::dw1_tc_kernel<<<dim3((t->O + 127) / 128, t->splits), DW_THREADS, DW_SMEM_BYTES, ds>>>(d);
    if (forked) CK(cudaEventRecord(t->join_ev, ds));
  }
  const int njobs = dw1_tc ? 5 : (t->O + 63) / 64 + 4;
  static const int wg_occ = getenv("PB_WG_OCC") ? atoi(getenv("PB_WG_OCC")) : 2;
  constexpr size_t wg_smem = 2 * WG_BK * (128 + 64) * sizeof(float);
  // PB_WG_SMALL=0 keeps the five small jobs in wgrad_kernel (same slabs bit for bit; slower, and it cannot share an SM
  // with dw1_tc_kernel)
  static const int wg_small = getenv("PB_WG_SMALL") ? atoi(getenv("PB_WG_SMALL")) : 1;
  if (dw1_tc && wg_small && !use_mma) wgrad_small_kernel<<<5 * t->splits, 256, WGA_SMEM_BYTES, st>>>(w);
  else if (use_mma) wgrad_kernel<2, true><<<njobs * t->splits, 256, WG_SMEM_FLOATS_MMA * sizeof(float), st>>>(w);
  else if (wg_occ == 3) wgrad_kernel<3><<<njobs * t->splits, 256, wg_smem, st>>>(w);
  else wgrad_kernel<2><<<njobs * t->splits, 256, wg_smem, st>>>(w);
  ut_tick(st, 3);   // end of wgrad_kernel on the main stream
  if (forked) CK(cudaStreamWaitEvent(st, t->join_ev, 0));
  ut_tick(st, 4);   // both weight-gradient kernels done
  // More than one rank: the flat gradient goes into the exported exchange buffer and peer_avg_kernel reads every
  // rank's copy over NVLink (average + squared-norm partials in one launch); ncclAllReduce only as the fallback.
  const bool peer = t->peer.on;
  if (peer && t->peer.ll) {
    // slab reduction, push of {value, seq} words to every peer, poll of the own receive area, average, squared-norm partials
    grad_reduce_peer_ll_kernel<<<t->npart, 256, 0, st>>>(t->slabs, t->P, t->splits, t->part, t->nblk_loss, t->L, t->scalars + 1, t->colsum,
                                                         t->peer.args);
  } else if (peer && t->peer.fused) {
    // slab reduction, exchange, average and squared-norm partials in one launch, slice by slice
    grad_reduce_peer_kernel<<<t->npart, 256, 0, st>>>(t->slabs, t->P, t->splits, t->part, t->nblk_loss, t->L, t->scalars + 1, t->colsum,
                                                      t->peer.args);
  } else {
    grad_reduce_kernel<<<t->npart, 256, 0, st>>>(t->slabs, t->P, t->splits, t->part, t->nblk_loss, t->L, peer ? t->peer.gbuf() : t->grad,
                                                 t->partials, t->scalars + 1, t->colsum, peer ? t->peer.seq : nullptr, t->peer.Ppad);
  }
  if (peer && !t->peer.fused && !t->peer.ll) {
    peer_avg_kernel<<<t->npart, 256, 0, st>>>(t->peer.args);
  } else if (!peer && t->comm) {
    ncclResult_t r = g_nccl.AllReduce(t->grad, t->grad, t->P, ncclFloat, ncclAvg, t->comm, st);
    if (r != ncclSuccess) return pb_fail(PB_ERR_COMM, std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?"));
    sqnorm_partials_kernel<<<t->npart, 256, 0, st>>>(t->grad, t->P, t->partials);
  }
  ut_tick(st, 5);
  AdamState as = {t->params, t->adam_m, t->adam_v, t->grad, t->P};
  adam2_kernel<<<t->npart, 256, 0, st>>>(as, t->partials, t->npart, t->hp, t->scalars, t->cursor, t->mb_stats, t->clip_acc, t->colsum,
                                         mb, t->L, t->W1T, t->W2T, t->Wc, peer ? t->peer.seq : nullptr);
  ut_tick(st, 6);
  g_ut.mbi++;
  CK(cudaGetLastError());
  return PB_OK;
}

static int set_lr(pb_trainer *t, float lr) {
  CK(cudaMemcpyAsync(t->scalars, &lr, sizeof(float), cudaMemcpyHostToDevice, t->stream));
  return PB_OK;
}

extern "C" int pb_trainer_update(pb_trainer *t, float lr) {
  CK(cudaSetDevice(t->env->device));
  const int E = t->cfg.update_epochs, B = t->B, nmb = t->cfg.num_minibatches;
  int rc = set_lr(t, lr);
  if (rc) return rc;
  if ((rc = refresh_transposed(t))) return rc;
  // losses/clipfrac is the mean over all update_epochs x num_minibatches minibatches of the iteration (:505,581)
  CK(cudaMemsetAsync(t->clip_acc, 0, sizeof(float), t->stream));
  t->clip_count = E * nmb;
  // Per epoch: host mt19937 Fisher-Yates permutation (bit-exact with torch::randperm) into its own pinned slot, one
  // upload, then the epoch's launches.  The host draws epoch e while the device is still busy with the rollout / epoch
  // e-1, so the stream never waits on the generator; a slot is rewritten only after its previous upload has finished.
  std::vector<long long> perm(B);
  if (g_ut.on < 0) g_ut.on = getenv("PB_UPDATE_TIMING") ? atoi(getenv("PB_UPDATE_TIMING")) : 0;
  g_ut.mbi = 0; g_ut.host_perm_ms = 0;
  const auto ut_t0 = std::chrono::steady_clock::now();
  for (int e = 0; e < E; e++) {
    int *slot = t->h_idx + (size_t)e * B;
    if (t->idx_ev_used[e]) CK(cudaEventSynchronize(t->idx_ev[e]));
    const auto ut_p0 = std::chrono::steady_clock::now();
    pb_randperm_i64(t->perm_rng.gen, B, perm.data());
    for (int i = 0; i < B; i++) slot[i] = (int)perm[i];
    g_ut.host_perm_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - ut_p0).count();
    CK(cudaMemcpyAsync(t->idx_all + (size_t)e * B, slot, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, t->stream));
    CK(cudaEventRecord(t->idx_ev[e], t->stream));
    t->idx_ev_used[e] = 1;
    if ((rc = epoch_prologue(t, t->idx_all + (size_t)e * B, B, nmb))) return rc;
    // The nmb minibatch steps of an epoch are one CUDA graph: every launch argument is constant (the minibatch cursor,
    // the optimiser step and the exchange sequence number live in device memory), so the graph is captured on first use
    // and replayed — one submission instead of 6-7 launches x nmb, and no launch gaps between the short kernels.
    // PB_UPDATE_GRAPH=0 (and the per-kernel timing mode) launch directly.
    static const int use_graph = getenv("PB_UPDATE_GRAPH") ? atoi(getenv("PB_UPDATE_GRAPH")) : 1;
    const bool nccl_exchange = t->comm && !t->peer.on;   // NCCL's lazy setup inside a capture is not worth the risk
    if (use_graph && g_ut.on != 1 && !nccl_exchange) {
      if (!t->mb_graph) {
        if ((rc = configure_update_kernels())) return rc;
        cudaGraph_t g = nullptr;
        CK(cudaStreamBeginCapture(t->stream, cudaStreamCaptureModeThreadLocal));
        for (int k = 0; k < nmb && !rc; k++) rc = minibatch_step(t);
        cudaError_t ce = cudaStreamEndCapture(t->stream, &g);
        if (rc) { if (g) cudaGraphDestroy(g); return rc; }
        CK(ce);
        ce = cudaGraphInstantiate(&t->mb_graph, g, 0);
        cudaGraphDestroy(g);
        CK(ce);
      }
      CK(cudaGraphLaunch(t->mb_graph, t->stream));
    } else {
      for (int k = 0; k < nmb; k++) if ((rc = minibatch_step(t))) return rc;
    }
  }
  CK(cudaEventRecord(t->ev[3], t->stream));
  if (g_ut.on == 1) {
    g_ut.host_loop_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - ut_t0).count();
    CK(cudaStreamSynchronize(t->stream));
    const double wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - ut_t0).count();
    double seg[UT_SEG] = {0}, span = 0;
    int n = 0;
    if (!g_ut.ev.empty() && g_ut.mbi >= UT_FIRST + UT_COUNT) {
      for (int i = 0; i < UT_COUNT; i++) {
        for (int k = 0; k < UT_SEG; k++) {
          float ms = 0; cudaEventElapsedTime(&ms, g_ut.ev[(size_t)i * (UT_SEG + 1) + k], g_ut.ev[(size_t)i * (UT_SEG + 1) + k + 1]);
          seg[k] += ms * 1e3;
        }
        n++;
      }
      float ms = 0; cudaEventElapsedTime(&ms, g_ut.ev[0], g_ut.ev[(size_t)(UT_COUNT - 1) * (UT_SEG + 1) + UT_SEG]);
      span = ms * 1e3 / UT_COUNT;
    }
    fprintf(stderr, "[update timing] host: perm %.2f ms of launch loop %.2f ms (returns before the device drains), wall to drain %.2f ms, %d minibatches\n",
            g_ut.host_perm_ms, g_ut.host_loop_ms, wall_ms, g_ut.mbi);
    if (n) fprintf(stderr, "[update timing] device us/minibatch over %d: l1 %.1f | fwd_bwd %.1f | wgrad(main) %.1f | wait dw1 join %.1f | grad_reduce(+allreduce) %.1f | adam %.1f | sum %.1f | span incl. gaps %.1f\n",
                   n, seg[0] / n, seg[1] / n, seg[2] / n, seg[3] / n, seg[4] / n, seg[5] / n,
                   (seg[0] + seg[1] + seg[2] + seg[3] + seg[4] + seg[5]) / n, span);
  }
  return PB_OK;
}

extern "C" int pb_trainer_update_minibatch(pb_trainer *t, const int *idx_host, int n, float lr) {
  if (n != t->mb) return pb_fail(PB_ERR_ARG, "pb_trainer_update_minibatch: n must equal the minibatch size");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(t->idx_all, idx_host, (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemset(t->clip_acc, 0, sizeof(float)));
  t->clip_count = 1;
  int rc = set_lr(t, lr);
  if (rc) return rc;
  if ((rc = refresh_transposed(t))) return rc;
  if ((rc = epoch_prologue(t, t->idx_all, n, 1))) return rc;
  if ((rc = minibatch_step(t))) return rc;
  CK(cudaStreamSynchronize(t->stream));
  return PB_OK;
}

extern "C" int pb_trainer_iteration(pb_trainer *t, float lr) {
  int rc;
  if ((rc = pb_trainer_rollout(t))) return rc;
  if ((rc = pb_trainer_gae(t))) return rc;
  return pb_trainer_update(t, lr);
}

extern "C" int pb_trainer_stats(pb_trainer *t, pb_ppo_stats *out) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  memset(out, 0, sizeof(*out));
  float s[PB_NSTAT], ep[4];
  CK(cudaMemcpy(s, t->mb_stats, sizeof(s), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(ep, t->ep_stat, sizeof(ep), cudaMemcpyDeviceToHost));
  CK(cudaMemset(t->ep_stat, 0, sizeof(ep)));
  out->policy_loss = s[0]; out->value_loss = s[1]; out->entropy = s[2]; out->old_approx_kl = s[3]; out->approx_kl = s[4]; out->clipfrac = s[5];
  if (t->clip_count > 0) {   // mean of the per-minibatch clipfracs of the last update, as the reference logs it (:505,581)
    float acc = 0.f;
    CK(cudaMemcpy(&acc, t->clip_acc, sizeof(float), cudaMemcpyDeviceToHost));
    out->clipfrac = acc / (float)t->clip_count;
  }
  out->episodic_return_sum = ep[0]; out->episodic_length_sum = ep[1]; out->episodes = (int)ep[2];
  cudaEventElapsedTime(&out->rollout_ms, t->ev[0], t->ev[1]);
  cudaEventElapsedTime(&out->gae_ms, t->ev[1], t->ev[2]);
  cudaEventElapsedTime(&out->update_ms, t->ev[2], t->ev[3]);
  cudaGetLastError();  // events may not have been recorded yet
  unsigned long long c[2];
  pb_env_counters(t->env, c);
  out->forward_evals = c[0]; out->overflow_envs = c[1];
  return PB_OK;
}

// ------------------------------------------------------------------ final evaluation and checkpoints
extern "C" int pb_trainer_eval(pb_trainer *t, int eval_seed, int num_eval_runs, float *returns_out, int cap, int *n_out) {
  if (!t || !returns_out || !n_out) return pb_fail(PB_ERR_ARG, "pb_trainer_eval: null argument");
  CK(cudaSetDevice(t->env->device));
  int rc = refresh_transposed(t);
  if (rc) return rc;
  if ((rc = pb_trainer_reset(t, eval_seed))) return rc;
  const size_t N = t->N;
  std::vector<int> has(N);
  std::vector<float> ret(N);
  int n = 0;
  for (long step = 0; n < num_eval_runs && n < cap && step < 1000000; step++) {
    if ((rc = launch_policy(t, 0, 1))) return rc;   // samples into row 0 of the rollout buffers
    StepArgs a = pb_env_make_args(t->env, 0, 0);
    a.actions = t->actions; a.obs = t->next_obs; a.rew = t->rewards; a.term = t->term; a.trunc = t->trunc;
    a.has_info = t->env->d_has; a.info_r = t->env->d_info_r; a.info_l = t->env->d_info_l;
    CK(pb_env_dispatch(t->env, a));
    t->step_index++;
    CK(cudaMemcpyAsync(has.data(), t->env->d_has, N * sizeof(int), cudaMemcpyDeviceToHost, t->stream));
    CK(cudaMemcpyAsync(ret.data(), t->env->d_info_r, N * sizeof(float), cudaMemcpyDeviceToHost, t->stream));
    CK(cudaStreamSynchronize(t->stream));
    for (size_t i = 0; i < N && n < cap; i++) if (has[i]) returns_out[n++] = ret[i];
  }
  *n_out = n;
  return PB_OK;
}

namespace {
struct CkptHeader { char magic[8]; int version, P, N, O; float step; int reserved[3]; };
struct CkptField { void *dev; size_t bytes; };
std::vector<CkptField> ckpt_fields(pb_trainer *t) {
  const size_t N = t->N, O = t->O, P = t->P;
  EnvBuffers &b = t->env->eb;
  return {{t->params, P * 4}, {t->adam_m, P * 4}, {t->adam_v, P * 4}, {b.ob_mean, N * O * 4}, {b.ob_var, N * O * 4},
          {b.ob_count, N * 4}, {b.rw_mean, N * 4}, {b.rw_var, N * 4}, {b.rw_acc, N * 4}, {b.rw_count, N * 4}};
}
}  // namespace

extern "C" int pb_trainer_save(pb_trainer *t, const char *path) {
  if (!t || !path) return pb_fail(PB_ERR_ARG, "pb_trainer_save: null argument");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  FILE *f = fopen(path, "wb");
  if (!f) return pb_fail(PB_ERR_ARG, std::string("pb_trainer_save: cannot open ") + path);
  CkptHeader h;
  memset(&h, 0, sizeof(h));
  memcpy(h.magic, "PB200CK", 8); h.version = 1; h.P = t->P; h.N = t->N; h.O = t->O;
  CK(cudaMemcpy(&h.step, t->scalars + 1, sizeof(float), cudaMemcpyDeviceToHost));
  bool ok = fwrite(&h, sizeof(h), 1, f) == 1;
  std::vector<char> buf;
  for (auto &fd : ckpt_fields(t)) {
    buf.resize(fd.bytes);
    CK(cudaMemcpy(buf.data(), fd.dev, fd.bytes, cudaMemcpyDeviceToHost));
    ok = ok && fwrite(buf.data(), 1, fd.bytes, f) == fd.bytes;
  }
  ok = (fclose(f) == 0) && ok;
  return ok ? PB_OK : pb_fail(PB_ERR_ARG, std::string("pb_trainer_save: short write to ") + path);
}

extern "C" int pb_trainer_load(pb_trainer *t, const char *path) {
  if (!t || !path) return pb_fail(PB_ERR_ARG, "pb_trainer_load: null argument");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  FILE *f = fopen(path, "rb");
  if (!f) return pb_fail(PB_ERR_ARG, std::string("pb_trainer_load: cannot open ") + path);
  CkptHeader h;
  bool ok = fread(&h, sizeof(h), 1, f) == 1 && !memcmp(h.magic, "PB200CK", 8) && h.version == 1;
  if (ok && (h.P != t->P || h.N != t->N || h.O != t->O)) { fclose(f); return pb_fail(PB_ERR_ARG, "pb_trainer_load: checkpoint is for a different env id / num_envs"); }
  std::vector<char> buf;
  for (auto &fd : ckpt_fields(t)) {
    if (!ok) break;
    buf.resize(fd.bytes);
    ok = fread(buf.data(), 1, fd.bytes, f) == fd.bytes;
    if (ok) CK(cudaMemcpy(fd.dev, buf.data(), fd.bytes, cudaMemcpyHostToDevice));
  }
  fclose(f);
  if (!ok) return pb_fail(PB_ERR_ARG, std::string("pb_trainer_load: bad or truncated checkpoint ") + path);
  CK(cudaMemcpy(t->scalars + 1, &h.step, sizeof(float), cudaMemcpyHostToDevice));
  return PB_OK;
}

// ------------------------------------------------------------------ named buffers
static int find_buffer(pb_trainer *t, const char *name, float **p, long long *n) {
  const long long TN = (long long)t->T * t->N;
  struct { const char *n; float *p; long long c; } tab[] = {
      {"obs", t->obs, TN * t->O}, {"actions", t->actions, TN * t->A}, {"logprobs", t->logprobs, TN}, {"rewards", t->rewards, TN},
      {"dones", t->dones, TN}, {"values", t->values, TN}, {"advantages", t->advantages, TN}, {"returns", t->returns, TN},
      {"next_obs", t->next_obs, (long long)t->N * t->O}, {"grad", t->grad, t->P}, {"term", t->term, t->N}, {"trunc", t->trunc, t->N},
      {"next_value", t->next_value, t->N}, {"adam_m", t->adam_m, t->P}, {"adam_v", t->adam_v, t->P}, {"scalars", t->scalars, 2}, {"mb_stats", t->mb_stats, PB_NSTAT}};
  for (auto &e : tab) if (!strcmp(e.n, name)) { *p = e.p; *n = e.c; return PB_OK; }
  return pb_fail(PB_ERR_ARG, std::string("unknown buffer: ") + name);
}
extern "C" int pb_trainer_buffer(pb_trainer *t, const char *name, void **dev_ptr, long long *count) {
  float *p; long long n;
  int rc = find_buffer(t, name, &p, &n);
  if (rc) return rc;
  if (dev_ptr) *dev_ptr = p;
  if (count) *count = n;
  return PB_OK;
}
extern "C" int pb_trainer_write_buffer(pb_trainer *t, const char *name, const float *host, long long count) {
  float *p; long long n;
  int rc = find_buffer(t, name, &p, &n);
  if (rc) return rc;
  if (count > n) return pb_fail(PB_ERR_ARG, "pb_trainer_write_buffer: too many elements");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(p, host, count * sizeof(float), cudaMemcpyHostToDevice));
  return PB_OK;
}
extern "C" int pb_trainer_read_buffer(pb_trainer *t, const char *name, float *host, long long count) {
  float *p; long long n;
  int rc = find_buffer(t, name, &p, &n);
  if (rc) return rc;
  if (count > n) return pb_fail(PB_ERR_ARG, "pb_trainer_read_buffer: too many elements");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(host, p, count * sizeof(float), cudaMemcpyDeviceToHost));
  return PB_OK;
}

// ------------------------------------------------------------------ NCCL data-parallel plumbing
extern "C" int pb_comm_unique_id(void *out128) {
  if (!g_nccl.load()) return pb_fail(PB_ERR_COMM, "libnccl.so.2 not found");
  ncclUniqueId id;
  ncclResult_t r = g_nccl.GetUniqueId(&id);
  if (r != ncclSuccess) return pb_fail(PB_ERR_COMM, "ncclGetUniqueId failed");
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  memcpy(out128, &id, 128);
  return PB_OK;
}
extern "C" int pb_trainer_comm_init(pb_trainer *t, const void *unique_id128, int rank, int world_size) {
  if (world_size <= 1) return PB_OK;
  if (!g_nccl.load()) return pb_fail(PB_ERR_COMM, "libnccl.so.2 not found");
  CK(cudaSetDevice(t->env->device));
  ncclUniqueId id;
  memcpy(&id, unique_id128, 128);
  ncclResult_t r = g_nccl.CommInitRank(&t->comm, world_size, id, rank);
  if (r != ncclSuccess) return pb_fail(PB_ERR_COMM, std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?"));
  t->rank = rank; t->world = world_size;
  if (t->mb_graph) { cudaGraphExecDestroy(t->mb_graph); t->mb_graph = nullptr; }   // captured without the exchange step
  if (t->peer.setup(t->comm, t->stream, rank, world_size, t->P, t->grad, t->partials))
    return pb_fail(PB_ERR_COMM, "peer exchange setup: NCCL failed");
  return PB_OK;
}
extern "C" int pb_trainer_broadcast_params(pb_trainer *t) {
  if (!t->comm) return PB_OK;
  CK(cudaSetDevice(t->env->device));
  ncclResult_t r = g_nccl.Broadcast(t->params, t->params, t->P, ncclFloat, 0, t->comm, t->stream);
  if (r != ncclSuccess) return pb_fail(PB_ERR_COMM, "ncclBroadcast failed");
  return PB_OK;
}
extern "C" int pb_trainer_exchange_mode(const pb_trainer *t) { return !t || !t->comm ? 0 : (t->peer.on ? 2 : 1); }

// ------------------------------------------------------------------ FP32 FMA peak (roofline denominator for the physics)
// 8 independent FFMA chains per thread, 1024 threads x 2 CTAs per SM: enough ILP x TLP to keep all four FMA pipes of
// every SM sub-partition issuing each cycle; nothing touches memory until the final store.
static __global__ void __launch_bounds__(1024, 2) fma_peak_kernel(float *out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int k = 0; k < 16; k++) {
      x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
      x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}
extern "C" int pb_fp32_fma_peak(int device, float ms_target, float *tflops_out) {
  if (!tflops_out) return pb_fail(PB_ERR_ARG, "pb_fp32_fma_peak: null argument");
  CK(cudaSetDevice(device));
  cudaDeviceProp pr;
  CK(cudaGetDeviceProperties(&pr, device));
  const int blocks = pr.multiProcessorCount * 2, threads = 1024;
  float *out = nullptr;
  CK(cudaMalloc((void **)&out, (size_t)blocks * threads * sizeof(float)));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  auto run = [&](int iters, float *ms) {
    cudaEventRecord(e0);
    fma_peak_kernel<<<blocks, threads>>>(out, iters, 0.999f, 1e-3f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(ms, e0, e1);
  };
  float ms = 0.f, best = 0.f;
  run(256, &ms);                                   // warm-up and a first rate estimate
  run(4096, &ms);
  int iters = (int)fminf(4.0e6f, fmaxf(4096.f, 4096.f * (ms_target > 0.f ? ms_target : 20.f) / fmaxf(ms, 1e-3f)));
  for (int rep = 0; rep < 3; rep++) {
    run(iters, &ms);
    const double flops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
    best = fmaxf(best, (float)(flops / (ms * 1e-3) / 1e12));
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
  CK(cudaGetLastError());
  *tflops_out = best;
  return PB_OK;
}
