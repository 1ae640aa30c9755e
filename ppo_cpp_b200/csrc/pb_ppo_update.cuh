// pb_ppo_update.cuh — PPO clipped loss with its hand-derived gradients, bias/logstd gradient reductions, deterministic
// slab reduction, global-norm clip and Adam (src/ppo_continuous_action.cpp:496-541; torch optim::Adam semantics).
#pragma once
#include "pb_ppo_common.cuh"

#define PB_LOSS_THREADS 64
#define PB_NSTAT 6  // pg_loss, v_loss, entropy, old_approx_kl, approx_kl, clipfrac (sums over the minibatch)

struct PpoHyper {
  float clip_coef, ent_coef, vf_coef, max_grad_norm, adam_eps, beta1, beta2;
  int norm_adv, clip_vloss;
};

// mean and 1/(unbiased std + 1e-8) of the gathered advantages (one block; fp64 accumulation)
// also advances the optimiser step count (scalars[1]) once per minibatch, before any kernel reads it
static __global__ void adv_stats_kernel(const float *__restrict__ adv, const int *__restrict__ idx_base, const int *cursor, int mb,
                                 float *__restrict__ out2, float *scalars) {
  __shared__ double s1[256], s2[256];
  const int *idx = idx_base + (size_t)(*cursor) * mb;
  double a = 0, b = 0;
  for (int i = threadIdx.x; i < mb; i += 256) { double v = adv[idx[i]]; a += v; b += v * v; }
  s1[threadIdx.x] = a; s2[threadIdx.x] = b;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    scalars[1] += 1.f;
    double mean = s1[0] / mb, var = (s2[0] - mb * mean * mean) / (mb - 1);
    out2[0] = (float)mean;
    out2[1] = (float)(1.0 / (sqrt(var > 0 ? var : 0) + 1e-8));
  }
}

struct LossArgs {
  const float *params; AgentLayout L; PpoHyper hp;
  const float *H2; int ldh;                 // [mb][128]
  const float *actions, *logp_old, *adv, *returns, *values;  // flat rollout buffers [T*N]
  const int *idx_base; const int *cursor; int mb;
  const float *adv_stats;                   // mean, inv_std
  float *D3; int ldd3;                      // [mb][A+1]  (dL/dmean..., dL/dv)
  float *D2; int ldd2;                      // [mb][128]  delta at the second tanh (critic | actor)
  float *part;                              // [nblk][PB_NSTAT + A] per-block sums (stats, dlogstd)
};

static __global__ void __launch_bounds__(PB_LOSS_THREADS) ppo_loss_kernel(LossArgs a) {
  const int A = a.L.act, b = blockIdx.x * PB_LOSS_THREADS + threadIdx.x;
  extern __shared__ float red[];  // [PB_LOSS_THREADS/32][PB_NSTAT + A]
  const int NP = PB_NSTAT + A;
  float stat[PB_NSTAT] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float dls[20];
  for (int i = 0; i < A; i++) dls[i] = 0.f;
  if (b < a.mb) {
    const int s = a.idx_base[(size_t)(*a.cursor) * a.mb + b];
    const float *h = a.H2 + (size_t)b * a.ldh;
    const float invB = 1.f / (float)a.mb;
    // heads
    float v = a.params[a.L.cb3];
    for (int k = 0; k < 64; k++) v += a.params[a.L.cW3 + k] * h[k];
    float logp = 0.f, ent = 0.f, dmean_unit[20];
    for (int i = 0; i < A; i++) {
      float mean = a.params[a.L.ab3 + i];
      const float *w = a.params + a.L.aW3 + i * 64;
      for (int k = 0; k < 64; k++) mean += w[k] * h[64 + k];
      float ls = a.params[a.L.logstd + i], sd = expf(ls), d = a.actions[(size_t)s * A + i] - mean, var = sd * sd;
      logp += -(d * d) / (2.f * var) - ls - 0.91893853320467274178f;
      ent += 0.5f + 0.91893853320467274178f + ls;
      dmean_unit[i] = d / var;          // dlogp/dmean
      dls[i] = d * d / var - 1.f;       // dlogp/dlogstd
    }
    float logratio = logp - a.logp_old[s], ratio = expf(logratio);
    float adv = a.adv[s];
    if (a.hp.norm_adv) adv = (adv - a.adv_stats[0]) * a.adv_stats[1];
    float c = a.hp.clip_coef;
    float pg1 = -adv * ratio, pg2 = -adv * fminf(fmaxf(ratio, 1.f - c), 1.f + c);
    // d max(pg1,pg2)/d ratio with torch's tie rule (halves) folded: inside the clip range both branches carry -adv
    float dpg = 0.f;
    bool inside = ratio >= 1.f - c && ratio <= 1.f + c;
    if (pg1 > pg2) dpg = -adv; else if (pg1 == pg2) dpg = inside ? -adv : -0.5f * adv; else dpg = inside ? -adv : 0.f;
    float g_logp = invB * dpg * ratio;
    // value loss
    float R = a.returns[s], vold = a.values[s], dv;
    float vu = (v - R) * (v - R), vloss = vu;
    if (a.hp.clip_vloss) {
      float dvo = v - vold, vc = vold + fminf(fmaxf(dvo, -c), c), vcl = (vc - R) * (vc - R);
      bool in = dvo >= -c && dvo <= c;
      float gu = 2.f * (v - R), gc = in ? 2.f * (vc - R) : 0.f;
      if (vu > vcl) dv = gu; else if (vu == vcl) dv = 0.5f * (gu + gc); else dv = gc;
      vloss = fmaxf(vu, vcl);
    } else dv = 2.f * (v - R);
    dv *= 0.5f * a.hp.vf_coef * invB;
    stat[0] = fmaxf(pg1, pg2); stat[1] = 0.5f * vloss; stat[2] = ent; stat[3] = -logratio;
    stat[4] = (ratio - 1.f) - logratio; stat[5] = fabsf(ratio - 1.f) > c ? 1.f : 0.f;
    // deltas at the heads
    float *d3 = a.D3 + (size_t)b * a.ldd3;
    for (int i = 0; i < A; i++) { dmean_unit[i] *= g_logp; d3[i] = dmean_unit[i]; dls[i] = g_logp * dls[i] - a.hp.ent_coef * invB; }
    d3[A] = dv;
    // delta through the second tanh of both nets
    float *d2 = a.D2 + (size_t)b * a.ldd2;
    for (int k = 0; k < 64; k++) {
      float hc = h[k];
      d2[k] = dv * a.params[a.L.cW3 + k] * (1.f - hc * hc);
      float sacc = 0.f;
      for (int i = 0; i < A; i++) sacc += dmean_unit[i] * a.params[a.L.aW3 + i * 64 + k];
      float ha = h[64 + k];
      d2[64 + k] = sacc * (1.f - ha * ha);
    }
  }
  // block reduction of stats and dlogstd, fixed order
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = 0; i < NP; i++) {
    float v = i < PB_NSTAT ? stat[i] : dls[i - PB_NSTAT];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp * NP + i] = v;
  }
  __syncthreads();
  if (threadIdx.x < NP) {
    float v = 0.f;
    for (int w = 0; w < PB_LOSS_THREADS / 32; w++) v += red[w * NP + threadIdx.x];
    a.part[(size_t)blockIdx.x * NP + threadIdx.x] = v;
  }
}

// bias gradients: column sums of D1 (128), D2 (128), D3 (A+1) over row chunk z -> slab z
static __global__ void bias_grad_kernel(const float *__restrict__ D1, const float *__restrict__ D2, const float *__restrict__ D3,
                                 int ldd3, int mb, int splits, AgentLayout L, float *__restrict__ slabs, size_t slab_stride) {
  int z = blockIdx.x, col = threadIdx.x, A = L.act;
  int chunk = (mb + splits - 1) / splits, r0 = z * chunk, r1 = min(mb, r0 + chunk);
  if (col >= 256 + A + 1) return;
  const float *src; int ld, c, dst;
  if (col < 128) { src = D1; ld = 128; c = col; dst = col < 64 ? L.cb1 + col : L.ab1 + col - 64; }
  else if (col < 256) { src = D2; ld = 128; c = col - 128; dst = c < 64 ? L.cb2 + c : L.ab2 + c - 64; }
  else { src = D3; ld = ldd3; c = col - 256; dst = c < A ? L.ab3 + c : L.cb3; }
  float s = 0.f;
  for (int r = r0; r < r1; r++) s += src[(size_t)r * ld + c];
  slabs[(size_t)z * slab_stride + dst] = s;
}

// grad[p] = sum_z slab[z][p] in fixed order; logstd entries come from the loss kernel's per-block partials
static __global__ void slab_reduce_kernel(const float *__restrict__ slabs, size_t slab_stride, int splits, const float *__restrict__ part,
                                   int nblk, AgentLayout L, float *__restrict__ grad) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= L.total) return;
  float s = 0.f;
  if (p >= L.logstd) {
    int i = PB_NSTAT + (p - L.logstd), NP = PB_NSTAT + L.act;
    for (int b = 0; b < nblk; b++) s += part[(size_t)b * NP + i];
  } else {
    for (int z = 0; z < splits; z++) s += slabs[(size_t)z * slab_stride + p];
  }
  grad[p] = s;
}

// per-block partial sums of g^2 (after the optional allreduce)
static __global__ void sqnorm_partials_kernel(const float *__restrict__ grad, int n, float *__restrict__ partials) {
  __shared__ float sm[256];
  int i = blockIdx.x * 256 + threadIdx.x;
  float v = i < n ? grad[i] : 0.f;
  sm[threadIdx.x] = v * v;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) partials[blockIdx.x] = sm[0];
}

struct AdamState { float *p, *m, *v; const float *g; int n; };
// clip_grad_norm_ (scale by max_norm/(norm+1e-6) when < 1) + torch Adam step; `scalars`: [0]=lr, [1]=step count (float)
static __global__ void adam_kernel(AdamState s, const float *__restrict__ partials, int npart, PpoHyper hp, float *scalars,
                            int *cursor, float *stats_out, const float *__restrict__ part, int nblk, int NP, int mb) {
  __shared__ float clip_s;
  if (threadIdx.x == 0) {
    float tot = 0.f;
    for (int i = 0; i < npart; i++) tot += partials[i];
    float c = hp.max_grad_norm / (sqrtf(tot) + 1e-6f);
    clip_s = c < 1.f ? c : 1.f;
  }
  __syncthreads();
  const float lr = scalars[0], t = scalars[1];
  const float bc1 = 1.f - powf(hp.beta1, t), bc2 = 1.f - powf(hp.beta2, t);
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < s.n) {
    float g = s.g[i] * clip_s;
    float m = hp.beta1 * s.m[i] + (1.f - hp.beta1) * g;
    float v = hp.beta2 * s.v[i] + (1.f - hp.beta2) * g * g;
    s.m[i] = m; s.v[i] = v;
    float denom = sqrtf(v) / sqrtf(bc2) + hp.adam_eps;
    s.p[i] -= (lr / bc1) * (m / denom);
  }
  // last block bookkeeping: publish minibatch stats, advance the step count and the minibatch cursor
  if (blockIdx.x == gridDim.x - 1) {
    __syncthreads();
    if (threadIdx.x < PB_NSTAT && stats_out) {
      float v = 0.f;
      for (int b = 0; b < nblk; b++) v += part[(size_t)b * NP + threadIdx.x];
      stats_out[threadIdx.x] = v / (float)mb;
    }
    if (threadIdx.x == 32) *cursor += 1;
  }
}
