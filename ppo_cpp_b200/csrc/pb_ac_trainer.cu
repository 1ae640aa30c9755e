// pb_ac_trainer.cu — rollout-and-update engine of ac_ppo_continuous_action behind the pb_ac_* C ABI (include/ppo_b200.h):
// the iteration loop of src/ac_ppo_continuous_action.cpp:624-950 on one GPU per process.  The reference collects with one
// host thread + CUDA stream + generator per env (:641-698); with the simulator on the device there is no straggler, so
// collection is one batched step for all envs of the rank (same data distribution: every env's trajectory depends only on
// its own RNG stream and the frozen policy).  Gradients and the advantage moments are averaged across ranks with NCCL
// (:830-849, :878-885).
#include <cuda_runtime.h>
#include <cstdint>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "pb_internal.h"
#include "pb_nccl.h"
#include "pb_ppo_fused.cuh"
#include "pb_ppo_rollout.cuh"
#include "pb_ppo_tc.cuh"
#include "pb_ac_fused.cuh"
#include "pb_ac_tc.cuh"
#define PB_PEER_HOST
#include "pb_peer.cuh"
#include "pb_ac_kernels.cuh"

void pb_randperm_i64(std::mt19937 &gen, int n, long long *out);  // pb_trainer.cu

struct AcNet {  // activations of one 2-layer encoder for up to R rows
  float *xh1, *h1, *xh2, *h2, *d2, *d1, *rstd1, *rstd2;
};

struct pb_ac_trainer {
  pb_ac_config cfg;
  std::string env_id;
  pb_env *env = nullptr;
  cudaStream_t stream;
  AcLayout L;
  PpoHyper hp;
  int N, T, O, A, B, mb, P, R, npart, splits, nloss;
  float *params, *adam_m, *adam_v, *grad, *obs_mean, *obs_std;
  float *obs, *actions, *logprobs, *rewards, *dones, *values, *advantages, *returns, *next_obs, *term, *trunc, *next_value;
  float *Xp, *actp; float4 *scp; float *Xn;
  AcNet c, a;
  int clip_count = 0;
  float *dwpart_a = nullptr, *dwpart_c = nullptr; // per-CTA partial weight gradients of actc::dw256_tc_kernel
  float *fpart_a = nullptr, *fpart_c = nullptr;   // per-block partial sums of the fused row kernels (pb_ac_fused.cuh), actor / critic
  int fblocks = 148, fused = 1;                    // persistent blocks of those kernels (one per SM); PB_AC_FUSED=0 keeps the one-op kernels
  cudaStream_t side = nullptr;       // critic chain of small forwards (rollout, GAE bootstrap) runs here beside the actor chain
  cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
  float *tc_wc2 = nullptr;           // split weights of the chain on the side stream
  double *adv_part = nullptr;   // block partials + ticket of the cross-rank advantage-moment kernels
  float *tc_wc = nullptr;   // TF32 hi / lo halves of the weight operand of the current tensor-core GEMM (stage layout, pb_ppo_tc.cuh)
  int train_steps = 0;   // steps per env the last collection gathered (num_train_steps_per_env, :722); num_steps unless pre-empted
  float *zab, *vbuf, *dzab, *dvw, *lnpart, *slabs, *partials, *losspart, *adv_stats, *moments, *scalars, *mb_stats, *ep_stat, *clip_acc;
  int *idx_all, *cursor, *h_idx = nullptr;
  std::vector<void *> allocs;
  std::mt19937 perm_gen;
  uint32_t step_index = 0;
  int host_cursor = 0;   // host mirror of *cursor (minibatch index inside the epoch buffers)
  cudaEvent_t ev[4];
  std::vector<cudaEvent_t> idx_ev; std::vector<char> idx_ev_used;
  ncclComm_t comm = nullptr;
  int world = 1, rank = 0;
  PeerExchange peer;   // LL exchange over NVLink peer memory for the gradient and the advantage moments (pb_peer.cuh); else ncclAllReduce
};

static bool ac_fused_mode(const pb_ac_trainer *t);
namespace {
template <class T>
int talloc(pb_ac_trainer *t, T **p, size_t n) {
  CK(cudaMalloc((void **)p, std::max<size_t>(n, 1) * sizeof(T)));
  t->allocs.push_back(*p);
  CK(cudaMemsetAsync(*p, 0, std::max<size_t>(n, 1) * sizeof(T), t->stream));
  return 0;
}
int alloc_net(pb_ac_trainer *t, AcNet &n) {
  const size_t R = t->R;
  int e = 0;
  e |= talloc(t, &n.xh1, R * AC_HID); e |= talloc(t, &n.h1, R * AC_HID); e |= talloc(t, &n.xh2, R * AC_HID); e |= talloc(t, &n.h2, R * AC_HID);
  e |= talloc(t, &n.d2, R * AC_HID); e |= talloc(t, &n.d1, R * AC_HID); e |= talloc(t, &n.rstd1, R); e |= talloc(t, &n.rstd2, R);
  return e;
}

// ---- small kernels private to this engine
// step[0] = lr stays; scalars[1] += 1 (optimiser step); sum of squares partials of the flat gradient
__global__ void ac_sqnorm_kernel(const float *__restrict__ g, int n, float *__restrict__ partials, float *step_count) {
  __shared__ float sm[256];
  int i = blockIdx.x * 256 + threadIdx.x;
  float v = i < n ? g[i] : 0.f;
  sm[threadIdx.x] = v * v;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) partials[blockIdx.x] = sm[0];
  if (step_count && blockIdx.x == 0 && threadIdx.x == 0) *step_count += 1.f;
}
__global__ void ac_slab_reduce_kernel(const float *__restrict__ slabs, size_t stride, int splits, int n, float *__restrict__ grad) {
  int p = blockIdx.x * 256 + threadIdx.x;
  if (p >= n) return;
  float s = 0.f;
  for (int z = 0; z < splits; z++) s += slabs[(size_t)z * stride + p];
  grad[p] = s;
}
// clip_grad_norm_ + torch Adam; last block publishes the minibatch statistics and advances the cursor
__global__ void __launch_bounds__(256) ac_adam_kernel(AdamState s, const float *__restrict__ partials, int npart, PpoHyper hp, float *scalars,
                                                      int *cursor, float *stats_out, float *clip_acc, const float *__restrict__ losspart,
                                                      int nloss, int mb, unsigned *peer_seq = nullptr) {
  __shared__ float sm[256];
  float v = 0.f;
  for (int i = threadIdx.x; i < npart; i += 256) v += partials[i];
  sm[threadIdx.x] = v;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  const float cn = hp.max_grad_norm / (sqrtf(sm[0]) + 1e-6f), clip_s = cn < 1.f ? cn : 1.f;
  const float lr = scalars[0], t = scalars[1];
  const float bc1 = 1.f - powf(hp.beta1, t), bc2 = 1.f - powf(hp.beta2, t);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < s.n) {
    float g = s.g[i] * clip_s;
    float m = hp.beta1 * s.m[i] + (1.f - hp.beta1) * g;
    float vv = hp.beta2 * s.v[i] + (1.f - hp.beta2) * g * g;
    s.m[i] = m; s.v[i] = vv;
    s.p[i] -= (lr / bc1) * (m / (sqrtf(vv) / sqrtf(bc2) + hp.adam_eps));
  }
  if (blockIdx.x == gridDim.x - 1) {   // minibatch statistics: per-block loss partials summed by all 256 threads, fixed order
    __shared__ float red[AC_NSTAT][256];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < AC_NSTAT; k++) {
      float acc = 0.f;
      for (int b = threadIdx.x; b < nloss; b += 256) acc += losspart[(size_t)b * AC_NSTAT + k];
      red[k][threadIdx.x] = acc;
    }
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) {
#pragma unroll
        for (int k = 0; k < AC_NSTAT; k++) red[k][threadIdx.x] += red[k][threadIdx.x + o];
      }
      __syncthreads();
    }
    if (threadIdx.x < AC_NSTAT) {
      const float acc = red[threadIdx.x][0] / (float)mb;
      stats_out[threadIdx.x] = acc;
      if (threadIdx.x == 5 && clip_acc) *clip_acc += acc;
    }
    if (threadIdx.x == 32) { *cursor += 1; if (peer_seq) *peer_seq += 1u; }   // the exchanges of this step are over on this rank
  }
}
// cross-rank advantage moments, the reference's two passes (:830-849): mean of the local means, then sum of squares
// Both passes run on AC_ADV_BLOCKS blocks (a minibatch is 4 MB of float4 rows - one block took ~50 us); every block leaves its fp64
// partial in `part`, the last block to finish (ticket counter behind the partials) adds them up in block order, so the result does not
// depend on the schedule.
#define AC_ADV_BLOCKS 64
__device__ __forceinline__ bool ac_adv_block_reduce(double a, double *part, double &total) {
  __shared__ double s1[256];
  __shared__ bool is_last;
  s1[threadIdx.x] = a;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) s1[threadIdx.x] += s1[threadIdx.x + o]; __syncthreads(); }
  unsigned *ticket = reinterpret_cast<unsigned *>(part + AC_ADV_BLOCKS);
  if (threadIdx.x == 0) {
    part[blockIdx.x] = s1[0];
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return false;
  if (threadIdx.x == 0) {
    __threadfence();
    double t = 0;
    for (unsigned b = 0; b < gridDim.x; b++) t += reinterpret_cast<volatile double *>(part)[b];
    total = t;
    *ticket = 0;   // ready for the next launch
  }
  return threadIdx.x == 0;
}
__global__ void __launch_bounds__(256) ac_adv_mean_kernel(const float4 *__restrict__ scp, const int *cursor, int mb, float *out, double *part) {
  const float4 *p = scp + (size_t)(*cursor) * mb;
  double a = 0, total = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < mb; i += gridDim.x * 256) a += p[i].y;
  if (ac_adv_block_reduce(a, part, total)) out[0] = (float)(total / mb);
}
__global__ void __launch_bounds__(256) ac_adv_sumsq_kernel(const float4 *__restrict__ scp, const int *cursor, int mb, const float *mean, float *out,
                                                            double *part) {
  const float4 *p = scp + (size_t)(*cursor) * mb;
  const double mu = mean[0];
  double a = 0, total = 0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < mb; i += gridDim.x * 256) { double d = p[i].y - mu; a += d * d; }
  if (ac_adv_block_reduce(a, part, total)) out[0] = (float)total;
}
__global__ void ac_adv_finish_kernel(const float *mean, const float *sumsq, int total, float *moments) {
  moments[0] = mean[0];
  moments[1] = 1.f / (sqrtf(sumsq[0] / (float)(total - 1)) + 1e-8f);
}
__global__ void ac_pick_moments_kernel(const float *adv_stats, const int *cursor, float *moments) {
  moments[0] = adv_stats[2 * (*cursor)]; moments[1] = adv_stats[2 * (*cursor) + 1];
}

// generic split-K weight gradient: C[j][k] = sum_m Aop[m][j] * Bop[m][k]; grid = a-tiles x b-tiles x splits
struct AcWgrad {
  const float *Aop; int lda, arows;     // arows valid columns of Aop (output rows)
  const float *Bop; int ldb, bcols;     // bcols valid columns of Bop (output columns)
  int M, splits, atiles, btiles;
  float *slabs; size_t stride; int w_off, ldc, b_off;   // weight block at w_off (row stride ldc), bias at b_off (-1: none)
  int row_split, w_off2, b_off2;                        // output rows >= row_split go to a second block (alpha | beta heads)
};
__global__ void __launch_bounds__(256, 2) ac_wgrad_kernel(AcWgrad g) {
  __shared__ __align__(16) float sA[2 * WG_BK * 64];
  __shared__ __align__(16) float sB[2 * WG_BK * 64];
  int id = blockIdx.x;
  const int z = id % g.splits; id /= g.splits;
  const int bt = id % g.btiles, at = id / g.btiles;
  const int chunk = (g.M + g.splits - 1) / g.splits, m0 = z * chunk, m1 = min(g.M, m0 + chunk);
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  float acc[4][4], cs[4];
#pragma unroll
  for (int i = 0; i < 4; i++) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f; cs[i] = 0.f; }
  const int j0 = at * 64, k0 = bt * 64;
  wg_tile<64, true>(g.Aop + j0, g.lda, g.Bop + k0, g.ldb, g.bcols - k0, (g.ldb & 3) == 0, m0, m1, sA, sB, acc, cs);
  float *slab = g.slabs + (size_t)z * g.stride;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int j = j0 + ty * 4 + i;
    if (j >= g.arows) continue;
    const bool second = j >= g.row_split;
    const int jr = second ? j - g.row_split : j, wo = second ? g.w_off2 : g.w_off, bo = second ? g.b_off2 : g.b_off;
#pragma unroll
    for (int c = 0; c < 4; c++) { const int k = k0 + tx * 4 + c; if (k < g.bcols) slab[wo + (size_t)jr * g.ldc + k] = acc[i][c]; }
    if (bt == 0 && tx == 0 && bo >= 0) slab[bo + jr] = cs[i];
  }
}

// 128 x 128 tile of the same split-K weight gradient for the hidden layers (256 x 256 per net): 8 x 8 accumulators per
// thread, 4 LDS.128 per 64 FFMA instead of 2 per 16, both operands streamed as they lie (rows of dY and of the layer
// input are already "k-major" for A^T B).  Same summation order over the rows as ac_wgrad_kernel -> identical slabs.
__global__ void __launch_bounds__(256, 1) ac_wgrad128_kernel(AcWgrad g) {
  constexpr int BK = 16;
  __shared__ __align__(16) float sA[2][BK][128];
  __shared__ __align__(16) float sB[2][BK][128];
  int id = blockIdx.x;
  const int z = id % g.splits; id /= g.splits;
  const int bt = id % g.btiles, at = id / g.btiles;
  const int chunk = (g.M + g.splits - 1) / g.splits, m0 = z * chunk, m1 = min(g.M, m0 + chunk);
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int j0 = at * 128, k0 = bt * 128;
  const float *Aop = g.Aop + j0, *Bop = g.Bop + k0;
  float acc[8][8], cs[8];
#pragma unroll
  for (int i = 0; i < 8; i++) {
    cs[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
  }
  float4 ar[2], br[2];
  auto gload = [&](int ms) {
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const int e = tid + 256 * i, kk = e >> 5, c4 = e & 31, m = ms + kk;
      ar[i] = m < m1 ? *reinterpret_cast<const float4 *>(Aop + (size_t)m * g.lda + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      br[i] = m < m1 ? *reinterpret_cast<const float4 *>(Bop + (size_t)m * g.ldb + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < 2; i++) {
      const int e = tid + 256 * i, kk = e >> 5, c4 = e & 31;
      *reinterpret_cast<float4 *>(&sA[buf][kk][c4 * 4]) = ar[i];
      *reinterpret_cast<float4 *>(&sB[buf][kk][c4 * 4]) = br[i];
    }
  };
  const int nst = (m1 - m0 + BK - 1) / BK;
  if (nst > 0) {
    gload(m0); sstore(0);
    __syncthreads();
    for (int s = 0; s < nst; s++) {
      if (s + 1 < nst) gload(m0 + (s + 1) * BK);
      const int buf = s & 1;
#pragma unroll
      for (int kk = 0; kk < BK; kk++) {
        const float4 a0 = *reinterpret_cast<const float4 *>(&sA[buf][kk][ty * 4]);
        const float4 a1 = *reinterpret_cast<const float4 *>(&sA[buf][kk][64 + ty * 4]);
        const float4 b0 = *reinterpret_cast<const float4 *>(&sB[buf][kk][tx * 4]);
        const float4 b1 = *reinterpret_cast<const float4 *>(&sB[buf][kk][64 + tx * 4]);
        const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int i = 0; i < 8; i++) {
          cs[i] += av[i];
#pragma unroll
          for (int j = 0; j < 8; j++) acc[i][j] += av[i] * bv[j];
        }
      }
      if (s + 1 < nst) sstore((s + 1) & 1);
      __syncthreads();
    }
  }
  float *slab = g.slabs + (size_t)z * g.stride;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const int j = j0 + (i >> 2) * 64 + ty * 4 + (i & 3);
    if (j >= g.arows) continue;
    const bool second = j >= g.row_split;
    const int jr = second ? j - g.row_split : j, wo = second ? g.w_off2 : g.w_off, bo = second ? g.b_off2 : g.b_off;
#pragma unroll
    for (int jh = 0; jh < 2; jh++)
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const int k = k0 + jh * 64 + tx * 4 + c;
        if (k < g.bcols) slab[wo + (size_t)jr * g.ldc + k] = acc[i][jh * 4 + c];
      }
    if (bt == 0 && tx == 0 && bo >= 0) slab[bo + jr] = cs[i];
  }
}
}  // namespace

extern "C" void pb_ac_config_default(pb_ac_config *c) {  // GlobalConfig of src/ac_ppo_continuous_action.cpp:55-100
  memset(c, 0, sizeof(*c));
  c->env_id = "Ant-v5"; c->num_envs = 8; c->num_steps = 128; c->device = 0; c->seed = 1;
  c->gamma = 0.99f; c->gae_lambda = 0.95f; c->num_minibatches = 4; c->update_epochs = 4; c->norm_adv = 1;
  c->clip_coef = 0.1f; c->clip_vloss = 1; c->ent_coef = 0.01f; c->vf_coef = 0.5f; c->max_grad_norm = 0.5f;
  c->adam_eps = 1e-5f; c->clip_actions = 1; c->rank = 0; c->world_size = 1; c->obs_mean = nullptr; c->obs_std = nullptr;
}

extern "C" int pb_ac_trainer_destroy(pb_ac_trainer *t) {
  if (!t) return PB_OK;
  if (t->env) { cudaSetDevice(t->env->device); cudaStreamSynchronize(t->stream); }
  // no rank may unmap its exchange buffer while a peer can still write to it: a one-float ncclAllReduce is the rendezvous (pb_trainer.cu)
  if (t->env && t->comm && t->peer.on && g_nccl.AllReduce && t->vbuf) {
    if (g_nccl.AllReduce(t->vbuf, t->vbuf, 1, ncclFloat, ncclSum, t->comm, t->stream) == ncclSuccess) cudaStreamSynchronize(t->stream);
  }
  if (t->env) t->peer.teardown(t->stream);
  if (t->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(t->comm);
  for (void *p : t->allocs) cudaFree(p);
  if (t->h_idx) cudaFreeHost(t->h_idx);
  for (auto &ev : t->ev) if (ev) cudaEventDestroy(ev);
  if (t->side) { cudaStreamSynchronize(t->side); cudaStreamDestroy(t->side); }
  if (t->fork_ev) cudaEventDestroy(t->fork_ev);
  if (t->join_ev) cudaEventDestroy(t->join_ev);
  for (auto &ev : t->idx_ev) if (ev) cudaEventDestroy(ev);
  if (t->env) pb_env_destroy(t->env);
  delete t;
  return PB_OK;
}

extern "C" int pb_ac_trainer_create(const pb_ac_config *cfg, pb_ac_trainer **out) {
  if (!cfg || !out || !cfg->env_id) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_create: null argument");
  if (cfg->num_steps <= 0 || cfg->num_minibatches <= 0 || cfg->update_epochs < 0) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_create: bad sizes");
  long long batch = (long long)cfg->num_envs * cfg->num_steps;
  if (batch % cfg->num_minibatches) return pb_fail(PB_ERR_ARG, "batch_size must be divisible by num_minibatches");
  pb_ac_trainer *t = new pb_ac_trainer();
  memset(t->ev, 0, sizeof(t->ev));
  t->cfg = *cfg; t->env_id = cfg->env_id; t->cfg.env_id = t->env_id.c_str();
  // make_env of :50-53: RecordEpisodeStatistics only, no normalisation wrappers
  int rc = pb_env_create(cfg->env_id, cfg->num_envs, cfg->device, PB_CHAIN_STATS_ONLY, cfg->gamma, cfg->clip_actions, &t->env);
  if (rc) { delete t; return rc; }
  t->stream = t->env->stream;
  t->N = cfg->num_envs; t->T = cfg->num_steps; t->O = t->env->obs_dim; t->A = t->env->nu;
  t->B = (int)batch; t->mb = t->B / cfg->num_minibatches; t->R = std::max(t->N, t->mb);
  if (2 * t->A > 64 || t->A >= AC_MAXA) { pb_ac_trainer_destroy(t); return pb_fail(PB_ERR_ARG, "action dimension too large"); }
  t->L = AcLayout::make(t->O, t->A); t->P = t->L.total;
  t->hp = {cfg->clip_coef, cfg->ent_coef, cfg->vf_coef, cfg->max_grad_norm, cfg->adam_eps, 0.9f, 0.999f, cfg->norm_adv, cfg->clip_vloss};
  t->npart = (t->P + 255) / 256;
  // 64 row splits for the large minibatches: the 128 x 128 weight-gradient tiles of a 256 x 256 layer are only 4 per
  // split, and 128 CTAs leave 20 SMs idle (update 171.9 -> 166.7 ms at 262 144 rows)
  t->splits = getenv("PB_AC_SPLITS") ? std::max(1, atoi(getenv("PB_AC_SPLITS"))) : (t->mb >= 131072 ? 64 : 32);
  if (t->mb < 4096) t->splits = std::max(1, std::min(t->splits, t->mb / 64 + 1));
  t->nloss = (t->mb + 127) / 128;
  t->rank = cfg->rank; t->world = cfg->world_size > 0 ? cfg->world_size : 1;
  const size_t N = t->N, T = t->T, O = t->O, A = t->A, P = t->P, R = t->R;
  int e = 0;
  e |= talloc(t, &t->params, P); e |= talloc(t, &t->adam_m, P); e |= talloc(t, &t->adam_v, P); e |= talloc(t, &t->grad, P);
  e |= talloc(t, &t->obs_mean, O); e |= talloc(t, &t->obs_std, O);
  e |= talloc(t, &t->obs, T * N * O); e |= talloc(t, &t->actions, T * N * A); e |= talloc(t, &t->logprobs, T * N);
  e |= talloc(t, &t->rewards, T * N); e |= talloc(t, &t->dones, T * N); e |= talloc(t, &t->values, T * N);
  e |= talloc(t, &t->advantages, T * N); e |= talloc(t, &t->returns, T * N);
  e |= talloc(t, &t->next_obs, N * O); e |= talloc(t, &t->term, N); e |= talloc(t, &t->trunc, N); e |= talloc(t, &t->next_value, N);
  e |= talloc(t, &t->Xp, T * N * O); e |= talloc(t, &t->actp, T * N * A); e |= talloc(t, &t->scp, T * N); e |= talloc(t, &t->Xn, R * O);
  e |= alloc_net(t, t->c); e |= alloc_net(t, t->a);
  e |= talloc(t, &t->zab, R * 2 * A); e |= talloc(t, &t->vbuf, R); e |= talloc(t, &t->dzab, R * 64); e |= talloc(t, &t->dvw, R * 64);
  e |= talloc(t, &t->lnpart, (size_t)4 * AC_LN_BLOCKS * 2 * AC_HID); e |= talloc(t, &t->slabs, (size_t)t->splits * P);
  e |= talloc(t, &t->partials, t->npart); e |= talloc(t, &t->losspart, (size_t)t->nloss * AC_NSTAT);
  e |= talloc(t, &t->adv_stats, 2 * (size_t)cfg->num_minibatches); e |= talloc(t, &t->moments, 4); e |= talloc(t, &t->scalars, 4);
  e |= talloc(t, &t->mb_stats, 8); e |= talloc(t, &t->ep_stat, 4); e |= talloc(t, &t->clip_acc, 1);
  e |= talloc(t, &t->tc_wc, (size_t)2 * AC_HID * AC_HID);
  e |= talloc(t, &t->adv_part, (size_t)AC_ADV_BLOCKS + 1);
  e |= talloc(t, &t->tc_wc2, (size_t)2 * AC_HID * AC_HID);
  {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg->device));
    t->fblocks = prop.multiProcessorCount;
    t->fused = getenv("PB_AC_FUSED") ? atoi(getenv("PB_AC_FUSED")) : 1;
    const size_t per_block = (size_t)acf::heads_part_ld<16>() + 512 + (size_t)AC_HID * AC_L1_MAXO + AC_HID;   // heads partials | LN1 + W1 partials
    e |= talloc(t, &t->fpart_a, per_block * t->fblocks); e |= talloc(t, &t->fpart_c, per_block * t->fblocks);
    if (ac_fused_mode(t)) { e |= talloc(t, &t->dwpart_a, (size_t)W_PART_LD * t->fblocks); e |= talloc(t, &t->dwpart_c, (size_t)W_PART_LD * t->fblocks); }
  }
  e |= talloc(t, &t->idx_all, (size_t)std::max(1, cfg->update_epochs) * t->B); e |= talloc(t, &t->cursor, 1);
  if (e) { pb_ac_trainer_destroy(t); return PB_ERR_CUDA; }
  CK(cudaMallocHost((void **)&t->h_idx, (size_t)std::max(1, cfg->update_epochs) * t->B * sizeof(int)));
  for (auto &ev : t->ev) CK(cudaEventCreate(&ev));
  CK(cudaStreamCreateWithFlags(&t->side, cudaStreamNonBlocking));
  CK(cudaEventCreateWithFlags(&t->fork_ev, cudaEventDisableTiming));
  CK(cudaEventCreateWithFlags(&t->join_ev, cudaEventDisableTiming));
  t->idx_ev.resize(std::max(1, cfg->update_epochs)); t->idx_ev_used.assign(t->idx_ev.size(), 0);
  for (auto &ev : t->idx_ev) CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  {  // fixed observation statistics (:496-534); default mean 0 / std 1 (the reference's HalfCheetah setting)
    std::vector<float> mu(O, 0.f), sd(O, 1.f);
    if (cfg->obs_mean) mu.assign(cfg->obs_mean, cfg->obs_mean + O);
    if (cfg->obs_std) sd.assign(cfg->obs_std, cfg->obs_std + O);
    CK(cudaMemcpy(t->obs_mean, mu.data(), O * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(t->obs_std, sd.data(), O * 4, cudaMemcpyHostToDevice));
  }
  {  // torch default init (no orthogonal call in this agent's constructor): Linear weight and bias ~ U(+-1/sqrt(fan_in)),
     // LayerNorm weight 1 / bias 0.  Distribution-equivalent to torch, not stream-equivalent.
    std::vector<float> h(P, 0.f);
    std::mt19937 gen((uint32_t)cfg->seed);
    auto lin = [&](int w, int b, int rows, int fan_in) {
      std::uniform_real_distribution<float> u(-1.f / std::sqrt((float)fan_in), 1.f / std::sqrt((float)fan_in));
      for (int i = 0; i < rows * fan_in; i++) h[w + i] = u(gen);
      for (int i = 0; i < rows; i++) h[b + i] = u(gen);
    };
    const AcLayout &L = t->L;
    lin(L.cW1, L.cb1, AC_HID, t->O); lin(L.cW2, L.cb2, AC_HID, AC_HID); lin(L.cW3, L.cb3, 1, AC_HID);
    lin(L.aW1, L.ab1, AC_HID, t->O); lin(L.aW2, L.ab2, AC_HID, AC_HID); lin(L.Wal, L.bal, t->A, AC_HID); lin(L.Wbe, L.bbe, t->A, AC_HID);
    for (int i = 0; i < AC_HID; i++) h[L.cg1 + i] = h[L.cg2 + i] = h[L.ag1 + i] = h[L.ag2 + i] = 1.f;
    CK(cudaMemcpy(t->params, h.data(), P * sizeof(float), cudaMemcpyHostToDevice));
  }
  t->perm_gen.seed((uint32_t)(cfg->seed * 1500 + cfg->rank));  // train_generator, :618
  CK(cudaStreamSynchronize(t->stream));
  *out = t;
  return PB_OK;
}

// ------------------------------------------------------------------ forward of both encoders + heads for M rows of Xn
static int ac_gemm(pb_ac_trainer *t, int mode, const float *A, int lda, const float *B, int ldb, float *C, int ldc, const float *bias,
                   int M, int N, int K) {
  AcGemm g = {A, lda, B, ldb, C, ldc, bias, M, N, K};
  // The 256-wide products (hidden-layer forward and input-gradient GEMMs of both encoders: 2/3 of the update's FLOPs together with
  // the weight gradients) run on the tcgen05 tensor cores, fp32-accurate as 3xTF32: tc::l1_forward_tc_kernel with a 128-column
  // output tile per blockIdx.z, the weight operand (W for C = A W^T, W^T for C = A W) split into TF32 hi / lo halves in the stage
  // layout right before the launch (65 k elements - a few microseconds; the weights change every optimiser step anyway), bias
  // added in the epilogue.  PB_AC_TC=0 keeps the FFMA kernels (A/B, parity tests run both).
  static const int tc_ok = getenv("PB_AC_TC") ? atoi(getenv("PB_AC_TC")) : 1;
  if (tc_ok && K == AC_HID && N == AC_HID && lda == K && (((uintptr_t)A) % 16) == 0 && M >= 128) {
    static bool cfgd[64] = {};
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!cfgd[dev & 63]) {
      CK(cudaFuncSetAttribute(tc::l1_forward_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L1_SMEM_BYTES));
      cfgd[dev & 63] = true;
    }
    const int nchunk = K / L1_BK, ntiles = N / 128, total = ntiles * nchunk * L1_B_FLOATS;
    tc::w_split_generic_kernel<<<(total + 255) / 256, 256, 0, t->stream>>>(B, ldb, mode == AC_NN ? 1 : 0, N, K, nchunk, ntiles, t->tc_wc);
    tc::L1Args a = {A, nullptr, M, K, t->tc_wc, {C, nullptr}, 0, ldc, (size_t)2 * nchunk * L1_B_FLOATS, bias};
    tc::l1_forward_tc_kernel<<<dim3((M + 127) / 128, 1, ntiles), L1_THREADS, L1_SMEM_BYTES, t->stream>>>(a);
    CK(cudaGetLastError());
    return PB_OK;
  }
  // 128 x 128 tiles when they still give every SM two CTAs (the update's minibatch GEMMs), else 64 x 64 (rollout forward).
  // PB_AC_GEMM128=0 keeps the small tile everywhere.
  static const int big_ok = getenv("PB_AC_GEMM128") ? atoi(getenv("PB_AC_GEMM128")) : 1;
  const long tiles128 = (long)((N + 127) / 128) * ((M + 127) / 128);
  if (big_ok && N >= 128 && (tiles128 >= 296 || big_ok == 2)) {   // 2: force the big tile (parity tests at small sizes)
    dim3 grid((N + 127) / 128, (M + 127) / 128);
    const bool vec = (K % 4 == 0) && (lda % 4 == 0) && (ldb % 4 == 0) && (((uintptr_t)A | (uintptr_t)B) % 16 == 0) &&
                     (mode == AC_NT || N % 4 == 0);
    if (mode == AC_NT) { if (vec) ac_gemm128_kernel<AC_NT, true><<<grid, 256, 0, t->stream>>>(g); else ac_gemm128_kernel<AC_NT, false><<<grid, 256, 0, t->stream>>>(g); }
    else { if (vec) ac_gemm128_kernel<AC_NN, true><<<grid, 256, 0, t->stream>>>(g); else ac_gemm128_kernel<AC_NN, false><<<grid, 256, 0, t->stream>>>(g); }
    return PB_OK;
  }
  dim3 grid((N + 63) / 64, (M + 63) / 64);
  if (mode == AC_NT) ac_gemm_kernel<AC_NT><<<grid, 256, 0, t->stream>>>(g); else ac_gemm_kernel<AC_NN><<<grid, 256, 0, t->stream>>>(g);
  return PB_OK;
}
// Fused mode (pb_ac_fused.cuh / pb_ac_tc.cuh): every consumer of a post-activation matrix recomputes it from x_hat, the LayerNorms
// ride in the GEMM epilogues.  Needs the fused heads kernel (2 A <= 16 outputs); PB_AC_FUSED=0 keeps one kernel per op.
static bool ac_fused_mode(const pb_ac_trainer *t) { return t->fused && 2 * t->A <= 16; }
static bool ac_tc_ln(const pb_ac_trainer *t, int M) {
  static const int tc_ok = getenv("PB_AC_TC") ? atoi(getenv("PB_AC_TC")) : 1;
  return ac_fused_mode(t) && tc_ok && M >= 64;
}
template <int EPI>
static int ac_gemm256(pb_ac_trainer *t, const float *W, int transposed, actc::GemmArgs g, cudaStream_t st = nullptr, float *wc = nullptr) {
  if (!st) st = t->stream;
  if (!wc) wc = t->tc_wc;
  static bool cfgd[64] = {};
  int dev = 0;
  CK(cudaGetDevice(&dev));
  if (!cfgd[dev & 63]) {
    CK(cudaFuncSetAttribute(actc::gemm256_tc_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM_BYTES));
    cfgd[dev & 63] = true;
  }
  actc::w_split256_kernel<<<AC_HID * AC_HID / 256, 256, 0, st>>>(W, transposed, wc);
  g.Wc = wc;
  const int ntiles = (g.M + 127) / 128;
  static const int debug = getenv("PB_AC_GEMM_DEBUG") ? atoi(getenv("PB_AC_GEMM_DEBUG")) : 0;
  static unsigned long long *dbg = nullptr;
  if (debug && !dbg) CK(cudaMalloc((void **)&dbg, 16 * sizeof(unsigned long long)));
  if (debug && ntiles >= 148) g.dbg = dbg;
  actc::gemm256_tc_kernel<EPI><<<std::min(t->fblocks, ntiles), G_THREADS, G_SMEM_BYTES, st>>>(g);
  CK(cudaGetLastError());
  if (g.dbg) {   // debugging aid: where CTA 0's issuer / first producer warp / first epilogue warp spent their cycles
    unsigned long long h[16];
    CK(cudaStreamSynchronize(st));
    CK(cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost));
    fprintf(stderr, "[gemm256 epi=%d] tiles %llu | issuer %llu cyc: wait A %llu, wait B %llu, wait acc-empty %llu | producer %llu: wait cp.async %llu, wait A-empty %llu | epilogue %llu: wait acc-full %llu\n",
            EPI, h[9], h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7], h[8]);
  }
  return PB_OK;
}
// heads of one net in one pass over h2 (acf::ac_heads_fwd_kernel); W1 == nullptr: a single weight block
static void ac_heads_fwd(pb_ac_trainer *t, const float *h2, const float *gamma, const float *beta, const float *W0, const float *b0, int n0,
                         const float *W1, const float *b1, int n1, float *out, int ldo, int M, cudaStream_t st = nullptr) {
  if (!st) st = t->stream;
  acf::AcHeadsFwd a = {h2, gamma, beta, W0, b0, n0, W1, b1, n1, out, ldo, M};
  const int NO = n0 + n1, grid = std::min(2 * t->fblocks, (M + 31) / 32);
  if (NO == 1) acf::ac_heads_fwd_kernel<1><<<grid, 256, NO * AC_HID * sizeof(float), st>>>(a);
  else acf::ac_heads_fwd_kernel<8><<<grid, 256, NO * AC_HID * sizeof(float), st>>>(a);
}
static int ac_forward(pb_ac_trainer *t, const float *X, int M, bool critic, bool actor) {
  const AcLayout &L = t->L;
  const float *P = t->params;
  cudaStream_t st = t->stream;
  const size_t n = (size_t)M * t->O;
  ac_standardise_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(X, t->obs_mean, t->obs_std, t->Xn, n, t->O);
  const int rb = (M + 7) / 8;
  const bool fused = ac_fused_mode(t), tcln = ac_tc_ln(t, M);
  // Small forwards (rollout step, GAE bootstrap: 8192 rows = 64 row tiles for 148 SMs): the critic chain runs on a side stream beside
  // the actor chain - the two nets share nothing but the standardised input.  Only on the all-fused path (own split-weight buffer
  // per chain, every launch takes its stream); PB_AC_FORK=0 keeps one stream.
  static const int fork_ok = getenv("PB_AC_FORK") ? atoi(getenv("PB_AC_FORK")) : 1;
  const bool fork = fork_ok && critic && actor && fused && tcln && t->O <= AC_L1_MAXO && 2 * ((M + 127) / 128) <= t->fblocks;
  auto net = [&](AcNet &nn, int W1, int b1, int g1, int B1, int W2, int b2, int g2, int B2, cudaStream_t s, float *wc) {
    if (t->fused && t->O <= AC_L1_MAXO) {   // layer 1 + LayerNorm + ReLU in one kernel, no pre-activation round trip
      acf::AcL1Fwd f = {t->Xn, t->O, P + W1, P + b1, P + g1, P + B1, nn.xh1, tcln ? nullptr : nn.h1, nn.rstd1, M};
      acf::ac_l1_ln_fwd_kernel<<<std::min(2 * t->fblocks, (M + 31) / 32), 256, 0, s>>>(f);
    } else {
      ac_gemm(t, AC_NT, t->Xn, t->O, P + W1, t->O, nn.xh1, AC_HID, P + b1, M, AC_HID, t->O);
      ac_ln_relu_fwd_kernel<<<rb, 256, 0, s>>>(nn.xh1, P + g1, P + B1, nn.rstd1, nn.h1, M);
    }
    if (tcln) {   // x_hat2 = LN(relu(gamma1 x_hat1 + beta1) W2^T + b2) in one kernel; h2 is never stored
      actc::GemmArgs g = {};
      g.A = nn.xh1; g.M = M; g.pg = P + g1; g.pb = P + B1; g.bias = P + b2; g.out = nn.xh2; g.rstd = nn.rstd2;
      ac_gemm256<actc::G_EPI_LN_FWD>(t, P + W2, 0, g, s, wc);
    } else {
      ac_gemm(t, AC_NT, nn.h1, AC_HID, P + W2, AC_HID, nn.xh2, AC_HID, P + b2, M, AC_HID, AC_HID);
      ac_ln_relu_fwd_kernel<<<rb, 256, 0, s>>>(nn.xh2, P + g2, P + B2, nn.rstd2, nn.h2, M);
    }
  };
  if (critic) {
    cudaStream_t s = fork ? t->side : st;
    if (fork) { CK(cudaEventRecord(t->fork_ev, st)); CK(cudaStreamWaitEvent(t->side, t->fork_ev, 0)); }
    net(t->c, L.cW1, L.cb1, L.cg1, L.cB1, L.cW2, L.cb2, L.cg2, L.cB2, s, fork ? t->tc_wc2 : t->tc_wc);
    if (fused) ac_heads_fwd(t, t->c.xh2, P + L.cg2, P + L.cB2, P + L.cW3, P + L.cb3, 1, nullptr, nullptr, 0, t->vbuf, 1, M, s);
    else ac_rowdot_kernel<<<rb, 256, 0, st>>>(t->c.h2, P + L.cW3, P + L.cb3, 1, t->vbuf, 1, 0, M);
    if (fork) CK(cudaEventRecord(t->join_ev, t->side));
  }
  if (actor) {
    net(t->a, L.aW1, L.ab1, L.ag1, L.aB1, L.aW2, L.ab2, L.ag2, L.aB2, st, t->tc_wc);
    if (fused) ac_heads_fwd(t, t->a.xh2, P + L.ag2, P + L.aB2, P + L.Wal, P + L.bal, t->A, P + L.Wbe, P + L.bbe, t->A, t->zab, 2 * t->A, M, st);
    else {
      ac_rowdot_kernel<<<rb, 256, 0, st>>>(t->a.h2, P + L.Wal, P + L.bal, t->A, t->zab, 2 * t->A, 0, M);
      ac_rowdot_kernel<<<rb, 256, 0, st>>>(t->a.h2, P + L.Wbe, P + L.bbe, t->A, t->zab, 2 * t->A, t->A, M);
    }
  }
  if (fork) CK(cudaStreamWaitEvent(st, t->join_ev, 0));
  CK(cudaGetLastError());
  return PB_OK;
}

static int ac_policy_step(pb_ac_trainer *t, int row, int mode, bool store) {
  int rc = ac_forward(t, t->next_obs, t->N, true, true);
  if (rc) return rc;
  const size_t N = t->N;
  AcSampleArgs a;
  memset(&a, 0, sizeof(a));
  a.zab = t->zab; a.ldz = 2 * t->A; a.v = t->vbuf; a.N = t->N; a.A = t->A; a.act_low = t->env->act_min; a.act_high = t->env->act_max;
  a.term_in = t->term; a.trunc_in = t->trunc; a.O = t->O;
  a.act_row = t->actions + row * N * t->A;
  if (store) { a.logp_row = t->logprobs + row * N; a.val_row = t->values + row * N; a.done_row = t->dones + row * N; }
  a.seed_lo = (uint32_t)t->cfg.seed; a.seed_hi = 0xAC99u ^ ((uint32_t)t->rank * 0x9E3779B9u); a.step_index = t->step_index; a.mode = mode;
  {
    const int epb = 256 / AC_SAMPLE_SLOT(t->A);
    ac_beta_sample_kernel<<<(t->N + epb - 1) / epb, 256, 0, t->stream>>>(a);
  }
  if (store) CK(cudaMemcpyAsync(t->obs + row * N * t->O, t->next_obs, N * t->O * sizeof(float), cudaMemcpyDeviceToDevice, t->stream));
  CK(cudaGetLastError());
  return PB_OK;
}

extern "C" int pb_ac_trainer_reset(pb_ac_trainer *t, int seed) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaMemsetAsync(t->term, 0, t->N * sizeof(float), t->stream));
  CK(cudaMemsetAsync(t->trunc, 0, t->N * sizeof(float), t->stream));
  return pb_env_reset(t->env, seed, t->next_obs);
}

// Collection (:641-715) as begin / step / end so that a host loop can stop it early (DD-PPO pre-emption, :680-693): with batched
// collection every env of the rank has gathered the same number of steps, which is min_collected_steps (:703-722).
extern "C" int pb_ac_trainer_collect_begin(pb_ac_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaEventRecord(t->ev[0], t->stream));
  t->train_steps = 0;
  return PB_OK;
}
extern "C" int pb_ac_trainer_collect_step(pb_ac_trainer *t, int s) {
  if (s < 0 || s >= t->T || s != t->train_steps) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_collect_step: steps must be collected in order, below num_steps");
  CK(cudaSetDevice(t->env->device));
  const size_t N = t->N;
  int rc = ac_policy_step(t, s, 0, true);
  if (rc) return rc;
  StepArgs a = pb_env_make_args(t->env, 0, 0);
  a.actions = t->actions + s * N * t->A; a.obs = t->next_obs; a.rew = t->rewards + s * N; a.term = t->term; a.trunc = t->trunc;
  a.has_info = nullptr; a.ep_stat = t->ep_stat;
  CK(pb_env_dispatch(t->env, a));
  t->step_index++;
  t->train_steps = s + 1;
  return PB_OK;
}
// preempted != 0: the collector broke out after the step it had just taken - the reference's loop then returns that step's INDEX
// (:649,688,693), so the last collected step is in the buffers (it is what GAE bootstraps from) but is not trained on.
extern "C" int pb_ac_trainer_collect_end(pb_ac_trainer *t, int preempted) {
  if (preempted) t->train_steps -= 1;
  if (t->train_steps < 1) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_collect_end: nothing to train on");
  CK(cudaSetDevice(t->env->device));
  CK(cudaEventRecord(t->ev[1], t->stream));
  return PB_OK;
}
// all ranks rendezvous (comm->allreduce(sychronize), :632): a one-float all-reduce, then a stream sync
extern "C" int pb_ac_trainer_barrier(pb_ac_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  if (t->comm) {
    if (g_nccl.AllReduce(t->vbuf, t->vbuf, 1, ncclFloat, ncclSum, t->comm, t->stream) != ncclSuccess) return pb_fail(PB_ERR_COMM, "ncclAllReduce failed");
  }
  CK(cudaStreamSynchronize(t->stream));
  return PB_OK;
}
extern "C" int pb_ac_trainer_train_steps(const pb_ac_trainer *t) { return t->train_steps > 0 ? t->train_steps : t->T; }

extern "C" int pb_ac_trainer_rollout(pb_ac_trainer *t) {
  int rc = pb_ac_trainer_collect_begin(t);
  for (int s = 0; s < t->T && !rc; s++) rc = pb_ac_trainer_collect_step(t, s);
  return rc ? rc : pb_ac_trainer_collect_end(t, 0);
}

extern "C" int pb_ac_trainer_gae(pb_ac_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  int rc = ac_forward(t, t->next_obs, t->N, true, false);   // next_value = V(next_obs), :760
  if (rc) return rc;
  CK(cudaMemcpyAsync(t->next_value, t->vbuf, t->N * sizeof(float), cudaMemcpyDeviceToDevice, t->stream));
  const int k = pb_ac_trainer_train_steps(t);
  if (k == t->T) {
    gae_kernel<<<(t->N + 127) / 128, 128, 0, t->stream>>>(t->T, t->N, t->rewards, t->values, t->dones, t->next_value, t->term, t->trunc,
                                                           t->cfg.gamma, t->cfg.gae_lambda, t->advantages, t->returns);
  } else {
    // pre-empted, training on k < num_steps steps: the reference's loop bootstraps from next_value only at t == num_steps - 1
    // (:766); at t = k - 1 it reads values[k] / dones[k], which the collector wrote for the step it took before breaking.
    // Slot k serves as "next value" / "next done" (one done array here: passed as the termination flag, zero truncation flag).
    CK(cudaMemsetAsync(t->vbuf, 0, t->N * sizeof(float), t->stream));
    gae_kernel<<<(t->N + 127) / 128, 128, 0, t->stream>>>(k, t->N, t->rewards, t->values, t->dones, t->values + (size_t)k * t->N,
                                                           t->dones + (size_t)k * t->N, t->vbuf, t->cfg.gamma, t->cfg.gae_lambda,
                                                           t->advantages, t->returns);
  }
  CK(cudaGetLastError());
  CK(cudaEventRecord(t->ev[2], t->stream));
  return PB_OK;
}

// ------------------------------------------------------------------ one minibatch (:816-888)
static int ac_wgrad(pb_ac_trainer *t, const float *Aop, int lda, int arows, const float *Bop, int ldb, int bcols, int w_off, int ldc, int b_off,
                    int row_split = 1 << 30, int w_off2 = 0, int b_off2 = -1, const float *xg = nullptr, const float *xb = nullptr) {
  AcWgrad g;
  g.Aop = Aop; g.lda = lda; g.arows = arows; g.Bop = Bop; g.ldb = ldb; g.bcols = bcols; g.M = t->mb; g.splits = t->splits;
  g.atiles = (arows + 63) / 64; g.btiles = (bcols + 63) / 64; g.slabs = t->slabs; g.stride = t->P; g.w_off = w_off; g.ldc = ldc; g.b_off = b_off;
  g.row_split = row_split; g.w_off2 = w_off2; g.b_off2 = b_off2;
  // hidden layers (both dimensions multiples of 128, 16-byte aligned rows): the 128 x 128 tile; PB_AC_GEMM128=0 disables
  static const int big_ok = getenv("PB_AC_GEMM128") ? atoi(getenv("PB_AC_GEMM128")) : 1;
  // hidden-layer weight gradients dW = dY^T H (K = the 262144 minibatch rows): tcgen05 3xTF32, tc::dw1_tc_kernel with a 128 x 128 output
  // tile per (blockIdx.z, blockIdx.x) and split, bias gradient = column sums of dY formed by the producers.  PB_AC_TC=0: FFMA tiles.
  static const int tc_ok = getenv("PB_AC_TC") ? atoi(getenv("PB_AC_TC")) : 1;
  if (tc_ok && arows % 128 == 0 && bcols % 128 == 0 && ldb == bcols && lda % 4 == 0 && (((uintptr_t)Aop | (uintptr_t)Bop) % 16) == 0 &&
      row_split >= arows) {
    static bool cfgd[64] = {};
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!cfgd[dev & 63]) {
      CK(cudaFuncSetAttribute(tc::dw1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DW_SMEM_BYTES));
      cfgd[dev & 63] = true;
    }
    tc::DwArgs d = {Bop, Aop, nullptr, t->mb, bcols, t->splits, 0, 0, t->slabs, (size_t)t->P, 1, lda, w_off, ldc, b_off, xg, xb};
    tc::dw1_tc_kernel<<<dim3(bcols / 128, t->splits, arows / 128), DW_THREADS, DW_SMEM_BYTES, t->stream>>>(d);
    CK(cudaGetLastError());
    return PB_OK;
  }
  if (xg) return pb_fail(PB_ERR_ARG, "ac_wgrad: the input transform needs the tensor-core kernel");
  if (big_ok && arows % 128 == 0 && bcols % 128 == 0 && lda % 4 == 0 && ldb % 4 == 0 && (((uintptr_t)Aop | (uintptr_t)Bop) % 16) == 0) {
    g.atiles = arows / 128; g.btiles = bcols / 128;
    ac_wgrad128_kernel<<<g.atiles * g.btiles * g.splits, 256, 0, t->stream>>>(g);
    return PB_OK;
  }
  ac_wgrad_kernel<<<g.atiles * g.btiles * g.splits, 256, 0, t->stream>>>(g);
  return PB_OK;
}
// LayerNorm gradients are written to t->grad by ac_colpart_reduce_kernel *after* the slab reduction has produced every
// other entry, so the order below is: forward, loss, all slab-producing launches, slab reduce, then the LN reductions.
static int ac_minibatch_step(pb_ac_trainer *t) {
  const int M = t->mb, A = t->A;
  const AcLayout &L = t->L;
  const float *P = t->params;
  cudaStream_t st = t->stream;
  int rc;
  // rows of this minibatch in the epoch buffers: kernels that take `cursor` offset themselves on the device; the forward
  // takes a pointer, so the cursor is mirrored on the host (it only ever counts launches of this function)
  const float *X = t->Xp + (size_t)t->host_cursor * M * t->O;
  if ((rc = ac_forward(t, X, M, true, true))) return rc;
  // advantage moments
  if (t->comm && t->peer.on) {   // the two scalar all-reduces of :830-849 as LL exchanges (words P and P + 1 of the receive rows)
    ac_adv_mean_kernel<<<AC_ADV_BLOCKS, 256, 0, st>>>(t->scp, t->cursor, M, t->moments + 2, t->adv_part);
    peer_ll_exchange_kernel<<<1, 32, 0, st>>>(t->moments + 2, 1, t->P, 1, nullptr, nullptr, t->peer.args);
    ac_adv_sumsq_kernel<<<AC_ADV_BLOCKS, 256, 0, st>>>(t->scp, t->cursor, M, t->moments + 2, t->moments + 3, t->adv_part);
    peer_ll_exchange_kernel<<<1, 32, 0, st>>>(t->moments + 3, 1, t->P + 1, 0, nullptr, nullptr, t->peer.args);
    ac_adv_finish_kernel<<<1, 1, 0, st>>>(t->moments + 2, t->moments + 3, t->world * M, t->moments);
  } else if (t->comm) {
    ac_adv_mean_kernel<<<AC_ADV_BLOCKS, 256, 0, st>>>(t->scp, t->cursor, M, t->moments + 2, t->adv_part);
    if (g_nccl.AllReduce(t->moments + 2, t->moments + 2, 1, ncclFloat, ncclAvg, t->comm, st) != ncclSuccess) return pb_fail(PB_ERR_COMM, "ncclAllReduce(adv mean)");
    ac_adv_sumsq_kernel<<<AC_ADV_BLOCKS, 256, 0, st>>>(t->scp, t->cursor, M, t->moments + 2, t->moments + 3, t->adv_part);
    if (g_nccl.AllReduce(t->moments + 3, t->moments + 3, 1, ncclFloat, ncclSum, t->comm, st) != ncclSuccess) return pb_fail(PB_ERR_COMM, "ncclAllReduce(adv sumsq)");
    ac_adv_finish_kernel<<<1, 1, 0, st>>>(t->moments + 2, t->moments + 3, t->world * M, t->moments);
  } else {
    ac_pick_moments_kernel<<<1, 1, 0, st>>>(t->adv_stats, t->cursor, t->moments);
  }
  AcLossArgs la;
  memset(&la, 0, sizeof(la));
  la.zab = t->zab; la.v = t->vbuf; la.ldz = 2 * A; la.M = M; la.A = A; la.actp = t->actp; la.scp = t->scp; la.cursor = t->cursor; la.mb = M;
  la.adv_moments = t->moments; la.hp = t->hp; la.act_low = t->env->act_min; la.act_high = t->env->act_max;
  la.dzab = t->dzab; la.dvw = t->dvw; la.part = t->losspart;
  ac_beta_loss_kernel<<<t->nloss, 128, 0, st>>>(la);
  const int rb = (M + 7) / 8;
  acf::AcReduceSegs segs;
  int nseg = 0, seg_cols = 0;
  auto seg = [&](const float *part, size_t ld, int nblk, int ncol, float *out0, float *out1, int split) {
    segs.s[nseg++] = {part, ld, nblk, ncol, out0, out1, split};
    seg_cols = std::max(seg_cols, ncol);
  };
  const size_t LNP = (size_t)AC_LN_BLOCKS * 2 * AC_HID;
  // fused row kernels (pb_ac_fused.cuh): heads backward + ReLU/LayerNorm-2 backward + head gradients in one pass over x_hat2
  const int NOa = 2 * A, noc = NOa <= 8 ? 8 : (NOa <= 12 ? 12 : 16);
  const bool fuse_heads = ac_fused_mode(t), fuse_w1 = t->fused && (t->O == 11 || t->O == 17), tcln = ac_tc_ln(t, M);
  float *w1part_a = t->fpart_a + (size_t)acf::heads_part_ld<16>() * t->fblocks, *w1part_c = t->fpart_c + (size_t)acf::heads_part_ld<16>() * t->fblocks;
  if (fuse_heads) {
    static bool cfgd[64] = {};
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (!cfgd[dev & 63]) {
      CK(cudaFuncSetAttribute(acf::ac_heads_ln_bwd_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, acf::heads_smem_bytes<16>()));
      CK(cudaFuncSetAttribute(acf::ac_heads_ln_bwd_kernel<12>, cudaFuncAttributeMaxDynamicSharedMemorySize, acf::heads_smem_bytes<12>()));
      CK(cudaFuncSetAttribute(acf::ac_heads_ln_bwd_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, acf::heads_smem_bytes<8>()));
      CK(cudaFuncSetAttribute(acf::ac_heads_ln_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, acf::heads_smem_bytes<1>()));
      CK(cudaFuncSetAttribute(acf::ac_w1_grad_kernel<17>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((AC_HID * 17 + AC_HID) * sizeof(float))));
      cfgd[dev & 63] = true;
    }
    acf::AcHeadsLnBwd ha = {t->dzab, 64, P + L.Wal, A, P + L.Wbe, A, t->a.xh2, t->a.rstd2, P + L.ag2, P + L.aB2, t->a.d2, t->fpart_a, M};
    acf::AcHeadsLnBwd hc = {t->dvw, 64, P + L.cW3, 1, nullptr, 0, t->c.xh2, t->c.rstd2, P + L.cg2, P + L.cB2, t->c.d2, t->fpart_c, M};
    const int nb = t->fblocks;
    int lda_ = 0;
    if (noc == 8) { lda_ = acf::heads_part_ld<8>(); acf::ac_heads_ln_bwd_kernel<8><<<nb, 256, acf::heads_smem_bytes<8>(), st>>>(ha); }
    else if (noc == 12) { lda_ = acf::heads_part_ld<12>(); acf::ac_heads_ln_bwd_kernel<12><<<nb, 256, acf::heads_smem_bytes<12>(), st>>>(ha); }
    else { lda_ = acf::heads_part_ld<16>(); acf::ac_heads_ln_bwd_kernel<16><<<nb, 256, acf::heads_smem_bytes<16>(), st>>>(ha); }
    const int ldc_ = acf::heads_part_ld<1>();
    acf::ac_heads_ln_bwd_kernel<1><<<nb, 256, acf::heads_smem_bytes<1>(), st>>>(hc);
    seg(t->fpart_a, lda_, nb, 2 * AC_HID, t->grad + L.ag2, t->grad + L.aB2, AC_HID);
    seg(t->fpart_a + 512, lda_, nb, NOa * AC_HID, t->grad + L.Wal, t->grad + L.Wbe, A * AC_HID);
    seg(t->fpart_a + 512 + noc * AC_HID, lda_, nb, NOa, t->grad + L.bal, t->grad + L.bbe, A);
    seg(t->fpart_c, ldc_, nb, 2 * AC_HID, t->grad + L.cg2, t->grad + L.cB2, AC_HID);
    seg(t->fpart_c + 512, ldc_, nb, AC_HID, t->grad + L.cW3, t->grad + L.cW3, AC_HID);
    seg(t->fpart_c + 512 + AC_HID, ldc_, nb, 1, t->grad + L.cb3, t->grad + L.cb3, 1);
  } else {
    // heads: dh2 of both nets, head weight / bias gradients
    ac_heads_bwd_kernel<<<rb, 256, 0, st>>>(t->dzab, 64, P + L.Wal, A, P + L.Wbe, A, t->a.d2, M);
    ac_heads_bwd_kernel<<<rb, 256, 0, st>>>(t->dvw, 64, P + L.cW3, 1, nullptr, 0, t->c.d2, M);
    ac_wgrad(t, t->dzab, 64, 2 * A, t->a.h2, AC_HID, AC_HID, L.Wal, AC_HID, L.bal, A, L.Wbe, L.bbe);
    ac_wgrad(t, t->dvw, 64, 1, t->c.h2, AC_HID, AC_HID, L.cW3, AC_HID, L.cb3);
  }
  // slab-producing part of both encoders, then the slab reduction, then the per-block partial sums (LayerNorm and fused-kernel
  // gradients), which overwrite their entries of t->grad
  auto enc = [&](AcNet &n, int W1, int b1, int g1, int B1, int W2, int b2, int g2, int B2, float *lnp2, float *lnp1, float *w1part, float *dwpart) {
    if (!fuse_heads) {
      ac_ln_relu_bwd_kernel<<<AC_LN_BLOCKS, 256, 0, st>>>(n.d2, n.h2, n.xh2, n.rstd2, P + g2, lnp2, M);
      seg(lnp2, 2 * AC_HID, AC_LN_BLOCKS, 2 * AC_HID, t->grad + g2, t->grad + B2, AC_HID);
    }
    if (tcln) {   // dW2 = d2^T relu(gamma1 x_hat1 + beta1), both operands read once (actc::dw256_tc_kernel)
      static bool cfgd[64] = {};
      int dev = 0;
      cudaGetDevice(&dev);
      if (!cfgd[dev & 63]) {
        if (cudaFuncSetAttribute(actc::dw256_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)W_SMEM_BYTES) != cudaSuccess) { rc = PB_ERR_CUDA; return; }
        cfgd[dev & 63] = true;
      }
      actc::Dw256Args w = {n.d2, n.xh1, M, P + g1, P + B1, dwpart};
      actc::dw256_tc_kernel<<<t->fblocks, W_THREADS, W_SMEM_BYTES, st>>>(w);
      seg(dwpart, W_PART_LD, t->fblocks, W_PART_LD, t->grad + W2, t->grad + b2, AC_HID * AC_HID);
    } else ac_wgrad(t, n.d2, AC_HID, AC_HID, n.h1, AC_HID, AC_HID, W2, AC_HID, b2);
    if (tcln) {   // dh1 = d2 W2 on tcgen05, then ReLU/LayerNorm-1 backward + dW1 / db1 in one row pass (d1 stays in registers)
      actc::GemmArgs g = {};
      g.A = n.d2; g.M = M; g.out = n.d1;
      ac_gemm256<actc::G_EPI_PLAIN>(t, P + W2, 1, g);
      acf::AcLnBwdW1 q = {n.d1, n.xh1, n.rstd1, P + g1, P + B1, t->Xn, n.d1, w1part, M};
      static bool cfgd[64] = {};
      int dev = 0;
      cudaGetDevice(&dev);
      if (!cfgd[dev & 63]) {
        cudaFuncSetAttribute(acf::ac_ln_bwd_w1_kernel<17>, cudaFuncAttributeMaxDynamicSharedMemorySize, acf::lnb_smem_bytes<17>());
        cudaFuncSetAttribute(acf::ac_ln_bwd_w1_kernel<11>, cudaFuncAttributeMaxDynamicSharedMemorySize, acf::lnb_smem_bytes<11>());
        cudaFuncSetAttribute(acf::ac_ln_bwd_w1_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, acf::lnb_smem_bytes<0>());
        cfgd[dev & 63] = true;
      }
      const int nb = t->fblocks;
      if (fuse_w1 && t->O == 17) acf::ac_ln_bwd_w1_kernel<17><<<nb, 256, acf::lnb_smem_bytes<17>(), st>>>(q);
      else if (fuse_w1) acf::ac_ln_bwd_w1_kernel<11><<<nb, 256, acf::lnb_smem_bytes<11>(), st>>>(q);
      else acf::ac_ln_bwd_w1_kernel<0><<<nb, 256, acf::lnb_smem_bytes<0>(), st>>>(q);
      const int ld = fuse_w1 ? 512 + AC_HID * t->O + AC_HID : 512;
      seg(w1part, ld, nb, 2 * AC_HID, t->grad + g1, t->grad + B1, AC_HID);
      if (fuse_w1) seg(w1part + 512, ld, nb, AC_HID * t->O + AC_HID, t->grad + W1, t->grad + b1, AC_HID * t->O);
      else ac_wgrad(t, n.d1, AC_HID, AC_HID, t->Xn, t->O, t->O, W1, t->O, b1);
      return;
    }
    ac_gemm(t, AC_NN, n.d2, AC_HID, P + W2, AC_HID, n.d1, AC_HID, nullptr, M, AC_HID, AC_HID);
    ac_ln_relu_bwd_kernel<<<AC_LN_BLOCKS, 256, 0, st>>>(n.d1, n.h1, n.xh1, n.rstd1, P + g1, lnp1, M);
    seg(lnp1, 2 * AC_HID, AC_LN_BLOCKS, 2 * AC_HID, t->grad + g1, t->grad + B1, AC_HID);
    if (fuse_w1) {
      acf::AcW1Grad w = {n.d1, t->Xn, w1part, M};
      const int ld = AC_HID * t->O + AC_HID;
      if (t->O == 17) acf::ac_w1_grad_kernel<17><<<t->fblocks, 256, ld * sizeof(float), st>>>(w);
      else acf::ac_w1_grad_kernel<11><<<t->fblocks, 256, ld * sizeof(float), st>>>(w);
      seg(w1part, ld, t->fblocks, ld, t->grad + W1, t->grad + b1, AC_HID * t->O);
    } else {
      ac_wgrad(t, n.d1, AC_HID, AC_HID, t->Xn, t->O, t->O, W1, t->O, b1);
    }
  };
  enc(t->a, L.aW1, L.ab1, L.ag1, L.aB1, L.aW2, L.ab2, L.ag2, L.aB2, t->lnpart, t->lnpart + LNP, w1part_a, t->dwpart_a);
  enc(t->c, L.cW1, L.cb1, L.cg1, L.cB1, L.cW2, L.cb2, L.cg2, L.cB2, t->lnpart + 2 * LNP, t->lnpart + 3 * LNP, w1part_c, t->dwpart_c);
  if (rc) return rc;
  ac_slab_reduce_kernel<<<t->npart, 256, 0, st>>>(t->slabs, t->P, t->splits, t->P, t->grad);
  acf::ac_reduce_segs_kernel<<<dim3((seg_cols + 31) / 32, nseg), 1024, 0, st>>>(segs);
  if (t->comm && t->peer.on) {   // average of every gradient tensor across ranks (:878-885): one LL exchange, squared-norm partials included
    peer_ll_exchange_kernel<<<t->npart, 256, 0, st>>>(t->grad, t->P, 0, 1, t->partials, t->scalars + 1, t->peer.args);
  } else {
    if (t->comm) {  // ... or as one flat ncclAllReduce
      if (g_nccl.AllReduce(t->grad, t->grad, t->P, ncclFloat, ncclAvg, t->comm, st) != ncclSuccess) return pb_fail(PB_ERR_COMM, "ncclAllReduce(grad)");
    }
    ac_sqnorm_kernel<<<t->npart, 256, 0, st>>>(t->grad, t->P, t->partials, t->scalars + 1);
  }
  AdamState as = {t->params, t->adam_m, t->adam_v, t->grad, t->P};
  ac_adam_kernel<<<t->npart, 256, 0, st>>>(as, t->partials, t->npart, t->hp, t->scalars, t->cursor, t->mb_stats, t->clip_acc, t->losspart,
                                           t->nloss, M, t->peer.on ? t->peer.seq : nullptr);
  CK(cudaGetLastError());
  t->host_cursor++;
  return PB_OK;
}

static int ac_epoch_prologue(pb_ac_trainer *t, const int *idx_dev, int rows, int nmb) {
  GatherArgs g;
  g.idx = idx_dev; g.rows = rows; g.O = t->O; g.A = t->A;
  g.obs = t->obs; g.actions = t->actions; g.logprobs = t->logprobs; g.advantages = t->advantages; g.returns = t->returns;
  g.values = t->values; g.Xp = t->Xp; g.actp = t->actp; g.scp = t->scp;
  epoch_gather_kernel<<<(rows + 7) / 8, 256, 0, t->stream>>>(g);
  adv_stats_all_kernel<<<nmb, 256, 0, t->stream>>>(t->scp, t->mb, t->adv_stats);
  CK(cudaMemsetAsync(t->cursor, 0, sizeof(int), t->stream));
  t->host_cursor = 0;
  CK(cudaGetLastError());
  return PB_OK;
}
static int ac_set_lr(pb_ac_trainer *t, float lr) {
  CK(cudaMemcpyAsync(t->scalars, &lr, sizeof(float), cudaMemcpyHostToDevice, t->stream));
  return PB_OK;
}

extern "C" int pb_ac_trainer_update(pb_ac_trainer *t, float lr) {
  CK(cudaSetDevice(t->env->device));
  const int E = t->cfg.update_epochs, B = t->B, nmb = t->cfg.num_minibatches;
  int rc = ac_set_lr(t, lr);
  if (rc) return rc;
  CK(cudaMemsetAsync(t->clip_acc, 0, sizeof(float), t->stream));   // losses/clipfrac = mean over the iteration's minibatches (:873,899)
  t->clip_count = E * nmb;
  std::vector<long long> perm(B);
  for (int e = 0; e < E; e++) {  // torch::randperm(num_samples, train_generator), :804
    int *slot = t->h_idx + (size_t)e * B;
    if (t->idx_ev_used[e]) CK(cudaEventSynchronize(t->idx_ev[e]));
    // b_inds = randperm(num_samples_collected); pre-empted collections repeat it up to batch_size_per_device (:804-810)
    pb_dd_ppo_indices(t->perm_gen, pb_ac_trainer_train_steps(t) * t->N, B, perm.data());
    for (int i = 0; i < B; i++) slot[i] = (int)perm[i];
    CK(cudaMemcpyAsync(t->idx_all + (size_t)e * B, slot, (size_t)B * sizeof(int), cudaMemcpyHostToDevice, t->stream));
    CK(cudaEventRecord(t->idx_ev[e], t->stream));
    t->idx_ev_used[e] = 1;
    if ((rc = ac_epoch_prologue(t, t->idx_all + (size_t)e * B, B, nmb))) return rc;
    for (int k = 0; k < nmb; k++) if ((rc = ac_minibatch_step(t))) return rc;
  }
  CK(cudaEventRecord(t->ev[3], t->stream));
  return PB_OK;
}

extern "C" int pb_ac_trainer_update_minibatch(pb_ac_trainer *t, const int *idx_host, int n, float lr) {
  if (n != t->mb) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_update_minibatch: n must equal the minibatch size");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(t->idx_all, idx_host, (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemset(t->clip_acc, 0, sizeof(float)));
  t->clip_count = 1;
  int rc = ac_set_lr(t, lr);
  if (rc) return rc;
  if ((rc = ac_epoch_prologue(t, t->idx_all, n, 1))) return rc;
  if ((rc = ac_minibatch_step(t))) return rc;
  CK(cudaStreamSynchronize(t->stream));
  return PB_OK;
}

extern "C" int pb_ac_trainer_iteration(pb_ac_trainer *t, float lr) {
  int rc;
  if ((rc = pb_ac_trainer_rollout(t))) return rc;
  if ((rc = pb_ac_trainer_gae(t))) return rc;
  return pb_ac_trainer_update(t, lr);
}

extern "C" int pb_ac_trainer_stats(pb_ac_trainer *t, pb_ppo_stats *out) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  memset(out, 0, sizeof(*out));
  float s[AC_NSTAT], ep[4];
  CK(cudaMemcpy(s, t->mb_stats, sizeof(s), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(ep, t->ep_stat, sizeof(ep), cudaMemcpyDeviceToHost));
  CK(cudaMemset(t->ep_stat, 0, sizeof(ep)));
  out->policy_loss = s[0]; out->value_loss = s[1]; out->entropy = s[2]; out->old_approx_kl = s[3]; out->approx_kl = s[4]; out->clipfrac = s[5];
  if (t->clip_count > 0) {   // mean of the per-minibatch clipfracs of the last update
    float acc = 0.f;
    CK(cudaMemcpy(&acc, t->clip_acc, sizeof(float), cudaMemcpyDeviceToHost));
    out->clipfrac = acc / (float)t->clip_count;
  }
  out->episodic_return_sum = ep[0]; out->episodic_length_sum = ep[1]; out->episodes = (int)ep[2];
  cudaEventElapsedTime(&out->rollout_ms, t->ev[0], t->ev[1]);
  cudaEventElapsedTime(&out->gae_ms, t->ev[1], t->ev[2]);
  cudaEventElapsedTime(&out->update_ms, t->ev[2], t->ev[3]);
  cudaGetLastError();
  unsigned long long c[2];
  pb_env_counters(t->env, c);
  out->forward_evals = c[0]; out->overflow_envs = c[1];
  return PB_OK;
}

extern "C" int pb_ac_trainer_num_params(const pb_ac_trainer *t) { return t->P; }
extern "C" void *pb_ac_trainer_stream(const pb_ac_trainer *t) { return (void *)t->stream; }
extern "C" pb_env *pb_ac_trainer_env(pb_ac_trainer *t) { return t->env; }
extern "C" int pb_ac_trainer_sync(pb_ac_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  return PB_OK;
}
extern "C" int pb_ac_trainer_get_params(pb_ac_trainer *t, float *h) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(h, t->params, t->P * sizeof(float), cudaMemcpyDeviceToHost));
  return PB_OK;
}
extern "C" int pb_ac_trainer_set_params(pb_ac_trainer *t, const float *h) {
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(t->params, h, t->P * sizeof(float), cudaMemcpyHostToDevice));
  return PB_OK;
}

static int ac_find_buffer(pb_ac_trainer *t, const char *name, float **p, long long *n) {
  const long long TN = (long long)t->T * t->N;
  struct { const char *n; float *p; long long c; } tab[] = {
      {"obs", t->obs, TN * t->O}, {"actions", t->actions, TN * t->A}, {"logprobs", t->logprobs, TN}, {"rewards", t->rewards, TN},
      {"dones", t->dones, TN}, {"values", t->values, TN}, {"advantages", t->advantages, TN}, {"returns", t->returns, TN},
      {"next_obs", t->next_obs, (long long)t->N * t->O}, {"grad", t->grad, t->P}, {"term", t->term, t->N}, {"trunc", t->trunc, t->N},
      {"next_value", t->next_value, t->N}, {"adam_m", t->adam_m, t->P}, {"adam_v", t->adam_v, t->P}, {"scalars", t->scalars, 2}, {"mb_stats", t->mb_stats, AC_NSTAT},
      {"zab", t->zab, (long long)t->R * 2 * t->A}, {"vbuf", t->vbuf, t->R}};
  for (auto &e : tab) if (!strcmp(e.n, name)) { *p = e.p; *n = e.c; return PB_OK; }
  return pb_fail(PB_ERR_ARG, std::string("unknown buffer: ") + name);
}
extern "C" int pb_ac_trainer_write_buffer(pb_ac_trainer *t, const char *name, const float *host, long long count) {
  float *p; long long n;
  int rc = ac_find_buffer(t, name, &p, &n);
  if (rc) return rc;
  if (count > n) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_write_buffer: too many elements");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(p, host, count * sizeof(float), cudaMemcpyHostToDevice));
  return PB_OK;
}
extern "C" int pb_ac_trainer_read_buffer(pb_ac_trainer *t, const char *name, float *host, long long count) {
  float *p; long long n;
  int rc = ac_find_buffer(t, name, &p, &n);
  if (rc) return rc;
  if (count > n) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_read_buffer: too many elements");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(host, p, count * sizeof(float), cudaMemcpyDeviceToHost));
  return PB_OK;
}
extern "C" long long pb_ac_trainer_buffer_count(pb_ac_trainer *t, const char *name) {
  float *p; long long n;
  return ac_find_buffer(t, name, &p, &n) ? -1 : n;
}

// forward only, for parity tests: alpha/beta logits and values of the current next_obs rows -> "zab" / "vbuf"
extern "C" int pb_ac_trainer_forward(pb_ac_trainer *t) {
  CK(cudaSetDevice(t->env->device));
  return ac_forward(t, t->next_obs, t->N, true, true);
}

// final evaluation (:965-1001): envs reset with eval_seed, action = Beta mean, episodic returns of finished episodes
extern "C" int pb_ac_trainer_eval(pb_ac_trainer *t, int eval_seed, int num_eval_runs, float *returns_out, int cap, int *n_out) {
  if (!t || !returns_out || !n_out) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_eval: null argument");
  CK(cudaSetDevice(t->env->device));
  int rc = pb_ac_trainer_reset(t, eval_seed);
  if (rc) return rc;
  const size_t N = t->N;
  std::vector<int> has(N);
  std::vector<float> ret(N);
  int n = 0;
  for (long step = 0; n < num_eval_runs && n < cap && step < 1000000; step++) {
    if ((rc = ac_policy_step(t, 0, 1, false))) return rc;
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

extern "C" int pb_ac_trainer_comm_init(pb_ac_trainer *t, const void *unique_id128, int rank, int world_size) {
  if (world_size <= 1) return PB_OK;
  if (!g_nccl.load()) return pb_fail(PB_ERR_COMM, "libnccl.so.2 not found");
  CK(cudaSetDevice(t->env->device));
  ncclUniqueId id;
  memcpy(&id, unique_id128, 128);
  ncclResult_t r = g_nccl.CommInitRank(&t->comm, world_size, id, rank);
  if (r != ncclSuccess) return pb_fail(PB_ERR_COMM, std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?"));
  t->rank = rank; t->world = world_size;
  // receive rows of P + 2 words: the flat gradient, then the advantage mean and sum-of-squares scalars
  if (t->peer.setup(t->comm, t->stream, rank, world_size, t->P + 2, t->grad, t->partials)) return pb_fail(PB_ERR_COMM, "peer exchange setup: NCCL failed");
  if (t->peer.on && !t->peer.ll) { t->peer.teardown(t->stream); }   // only the LL form is wired into this trainer
  return PB_OK;
}
extern "C" int pb_ac_trainer_exchange_mode(const pb_ac_trainer *t) { return !t || !t->comm ? 0 : (t->peer.on ? 2 : 1); }
extern "C" int pb_ac_trainer_broadcast_params(pb_ac_trainer *t) {  // :551-553
  if (!t->comm) return PB_OK;
  CK(cudaSetDevice(t->env->device));
  if (g_nccl.Broadcast(t->params, t->params, t->P, ncclFloat, 0, t->comm, t->stream) != ncclSuccess) return pb_fail(PB_ERR_COMM, "ncclBroadcast failed");
  return PB_OK;
}

// ---- observation statistics (--estimate_mean_std, :663-665, 956-963): population mean / std per column of the rollout's
// observations, all envs of the rank (the reference samples env 0 only).  One block per column, fp64 accumulation.
namespace {
__global__ void __launch_bounds__(256) ac_obs_moments_kernel(const float *__restrict__ X, long long rows, int O, double *__restrict__ out) {
  __shared__ double sh[2][256];
  const int o = blockIdx.x;
  double s = 0.0, q = 0.0;
  for (long long r = threadIdx.x; r < rows; r += 256) {
    const double v = (double)X[r * O + o];
    s += v; q += v * v;
  }
  sh[0][threadIdx.x] = s; sh[1][threadIdx.x] = q;
  __syncthreads();
  for (int k = 128; k > 0; k >>= 1) {
    if ((int)threadIdx.x < k) { sh[0][threadIdx.x] += sh[0][threadIdx.x + k]; sh[1][threadIdx.x] += sh[1][threadIdx.x + k]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out[2 * o] = sh[0][0]; out[2 * o + 1] = sh[1][0]; }
}
}  // namespace

// sums[2*o] = sum x, sums[2*o+1] = sum x^2 over the T*N rows of the rollout buffer; the caller merges iterations / ranks
extern "C" int pb_ac_trainer_obs_sums(pb_ac_trainer *t, double *sums_host, long long *rows_out) {
  if (!t || !sums_host || !rows_out) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_obs_sums: null argument");
  CK(cudaSetDevice(t->env->device));
  double *d = nullptr;
  CK(cudaMalloc((void **)&d, 2 * (size_t)t->O * sizeof(double)));
  const long long rows = (long long)t->T * t->N;
  ac_obs_moments_kernel<<<t->O, 256, 0, t->stream>>>(t->obs, rows, t->O, d);
  bool comm_ok = true;  // data parallel: every rank ends up with the same sums, hence the same standardisation
  if (t->comm) comm_ok = g_nccl.AllReduce(d, d, 2 * (size_t)t->O, ncclDouble, ncclSum, t->comm, t->stream) == ncclSuccess;
  cudaError_t e = cudaMemcpyAsync(sums_host, d, 2 * (size_t)t->O * sizeof(double), cudaMemcpyDeviceToHost, t->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
  cudaFree(d);
  CK(e);
  if (!comm_ok) return pb_fail(PB_ERR_COMM, "ncclAllReduce failed");
  *rows_out = rows * (t->comm ? t->world : 1);
  return PB_OK;
}
extern "C" int pb_ac_trainer_set_obs_stats(pb_ac_trainer *t, const float *mean_host, const float *std_host) {
  if (!t || !mean_host || !std_host) return pb_fail(PB_ERR_ARG, "pb_ac_trainer_set_obs_stats: null argument");
  CK(cudaSetDevice(t->env->device));
  CK(cudaStreamSynchronize(t->stream));
  CK(cudaMemcpy(t->obs_mean, mean_host, (size_t)t->O * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(t->obs_std, std_host, (size_t)t->O * 4, cudaMemcpyHostToDevice));
  return PB_OK;
}
extern "C" int pb_ac_trainer_obs_dim(const pb_ac_trainer *t) { return t->O; }
extern "C" int pb_ac_trainer_act_dim(const pb_ac_trainer *t) { return t->A; }

// ---- pb_obs_moments_noalloc: the observation-moment reduction of pb_ac_trainer_obs_sums for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// pb_ac_trainer_obs_sums cudaMallocs / cudaFrees its 2 x O doubles per call; this
// variant reduces in two launches through block partials kept in a file-scope __device__ array (below), so it allocates nothing:
//   stage 1  grid (O, PB_NOALLOC_SPLITS): block (o, k) sums x and x^2 of column o over its slice of the rows (fp64) -> partial
//   stage 2  one thread per column adds the PB_NOALLOC_SPLITS partials in index order (fixed order -> bit-reproducible)
// Limits: obs_dim <= PB_NOALLOC_MAX_OBS (512; Humanoid-v4 has 376).  The array is 2 * 512 * 64 doubles = 512 KiB and every GPU
// that loads the library holds it until exit.  Every call on a device shares it and it keeps the last call's contents: each call
// overwrites every slot it later reads (nothing has to start at zero, so nothing is cleared), and calls that overlap on the
// GPU - two streams, two host threads - race on it; serialise them.
#define PB_NOALLOC_MAX_OBS 512
#define PB_NOALLOC_SPLITS 64
static __device__ __align__(256) double pb_noalloc_scratch[2 * PB_NOALLOC_MAX_OBS * PB_NOALLOC_SPLITS];

namespace {
__global__ void __launch_bounds__(256) obs_moments_partial_kernel(const float *__restrict__ X, long long rows, int O, double *__restrict__ part) {
  __shared__ double sh[2][256];
  const int o = blockIdx.x, k = blockIdx.y;
  const long long per = (rows + PB_NOALLOC_SPLITS - 1) / PB_NOALLOC_SPLITS, r0 = k * per, r1 = r0 + per < rows ? r0 + per : rows;
  double s = 0.0, q = 0.0;
  for (long long r = r0 + threadIdx.x; r < r1; r += 256) {
    const double v = (double)X[r * O + o];
    s += v; q += v * v;
  }
  sh[0][threadIdx.x] = s; sh[1][threadIdx.x] = q;
  __syncthreads();
  for (int h = 128; h > 0; h >>= 1) {
    if ((int)threadIdx.x < h) { sh[0][threadIdx.x] += sh[0][threadIdx.x + h]; sh[1][threadIdx.x] += sh[1][threadIdx.x + h]; }
    __syncthreads();
  }
  // written unconditionally (an empty slice writes zeros): stage 2 never reads a slot of an earlier call
  if (threadIdx.x == 0) { part[((size_t)k * O + o) * 2] = sh[0][0]; part[((size_t)k * O + o) * 2 + 1] = sh[1][0]; }
}
__global__ void obs_moments_final_kernel(const double *__restrict__ part, int O, double *__restrict__ out) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= O) return;
  double s = 0.0, q = 0.0;
  for (int k = 0; k < PB_NOALLOC_SPLITS; k++) { s += part[((size_t)k * O + o) * 2]; q += part[((size_t)k * O + o) * 2 + 1]; }
  out[2 * o] = s; out[2 * o + 1] = q;
}
}  // namespace

extern "C" int pb_obs_moments_noalloc(int device, void *stream, const float *obs_dev, long long rows, int obs_dim, double *sums_out_dev) {
  if (!obs_dev || !sums_out_dev || rows <= 0 || obs_dim <= 0) return pb_fail(PB_ERR_ARG, "pb_obs_moments_noalloc: bad argument");
  if (obs_dim > PB_NOALLOC_MAX_OBS)
    return pb_fail(PB_ERR_ARG, "pb_obs_moments_noalloc: obs_dim " + std::to_string(obs_dim) + " exceeds the static scratch limit of " +
                                   std::to_string(PB_NOALLOC_MAX_OBS) + " columns");
  if (device < 0 || device >= 64) return pb_fail(PB_ERR_BAD_DEVICE, "pb_obs_moments_noalloc: bad device");
  CK(cudaSetDevice(device));
  // In host code the array's name is a host-side stand-in, not a device address: resolve it once per device, with that device
  // current, and hand the kernels the resolved pointer.
  static double *scratch[64] = {};
  if (!scratch[device]) {
    double *p = nullptr;
    CK(cudaGetSymbolAddress((void **)&p, pb_noalloc_scratch));
    scratch[device] = p;
  }
  cudaStream_t st = (cudaStream_t)stream;
  obs_moments_partial_kernel<<<dim3(obs_dim, PB_NOALLOC_SPLITS), 256, 0, st>>>(obs_dev, rows, obs_dim, scratch[device]);
  obs_moments_final_kernel<<<(obs_dim + 127) / 128, 128, 0, st>>>(scratch[device], obs_dim, sums_out_dev);
  CK(cudaGetLastError());
  return PB_OK;
}
extern "C" int pb_obs_moments_noalloc_max_obs(void) { return PB_NOALLOC_MAX_OBS; }
