// pb_ppo_rollout.cuh — rollout-side agent kernels: fused actor+critic forward with Normal sampling and log-prob
// (get_action_and_value, src/ppo_continuous_action.cpp:145-157 + include/rl_utils.h:21-46), storage of one rollout
// row (src/ppo_continuous_action.cpp:391-399), and the GAE reverse scan (src/ppo_continuous_action.cpp:447-467).
#pragma once
#include "pb_ppo_common.cuh"

#define PB_POL_TILE 16     // envs per CTA
#define PB_POL_THREADS 128 // 64 critic + 64 actor hidden units

// Transposed first/second layer weights for coalesced reads in the rollout kernel:
// W1T[k][j] (j<64 critic, j>=64 actor), W2T[net][k][u].
static __global__ void transpose_weights_kernel(const float *__restrict__ p, AgentLayout L, float *__restrict__ W1T, float *__restrict__ W2T) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int n1 = L.obs * 128;
  if (i < n1) {
    int k = i / 128, j = i - k * 128;
    W1T[i] = j < 64 ? p[L.cW1 + j * L.obs + k] : p[L.aW1 + (j - 64) * L.obs + k];
  } else if (i < n1 + 2 * 64 * 64) {
    int r = i - n1, net = r / 4096, k = (r % 4096) / 64, u = r % 64;
    W2T[r] = p[(net ? L.aW2 : L.cW2) + u * 64 + k];
  }
}

struct PolicyArgs {
  const float *params, *W1T, *W2T;
  AgentLayout L;
  int num_envs;
  const float *obs_in;            // [N][O] current observation (next_obs of the loop)
  const float *term_in, *trunc_in;  // [N] flags of the previous transition (next_done = term || trunc)
  // rollout row t (any may be NULL when only the value is wanted)
  float *obs_row, *act_row, *logp_row, *val_row, *done_row;
  uint32_t seed_lo, seed_hi, step_index;  // Philox key / counter word
  int sample;                     // 0: value only (GAE bootstrap)
  // first-layer pre-activations (without bias) from tc::l1_forward_tc_kernel, [N][128] per k-range (hpre2 may be NULL);
  // NULL: layer 1 is computed here on the FFMA pipe
  const float *hpre, *hpre2;
};

static __global__ void __launch_bounds__(PB_POL_THREADS) policy_forward_kernel(PolicyArgs a) {
  extern __shared__ __align__(16) float sm[];
  const int O = a.L.obs, A = a.L.act, tid = threadIdx.x, e0 = blockIdx.x * PB_POL_TILE;
  float *xT = sm;                            // [O][16]
  float *h1 = xT + O * PB_POL_TILE;          // [128][16]
  float *h2 = h1 + 128 * PB_POL_TILE;        // [128][16]
  float *outv = h2 + 128 * PB_POL_TILE;      // [16][A+1] : mean..., value
  float acc[PB_POL_TILE];
  if (a.hpre) {
    // layer 1 came from the tensor cores: copy the observation rows into the rollout buffer, add the k-range partials and the bias
    if (a.obs_row) {
      const size_t base = (size_t)e0 * O;
      const int nfl = min(PB_POL_TILE, a.num_envs - e0) * O;   // floats of this tile's rows (contiguous in both buffers)
      if ((((uintptr_t)(a.obs_in + base) | (uintptr_t)(a.obs_row + base)) & 15) == 0) {
        // 16-byte copies, four in flight per thread (a scalar copy loop was 31 % of this kernel's stall samples)
        const float4 *src = reinterpret_cast<const float4 *>(a.obs_in + base);
        float4 *dst = reinterpret_cast<float4 *>(a.obs_row + base);
        const int n4 = nfl >> 2;
        for (int i = tid; i < n4; i += 4 * PB_POL_THREADS) {
          float4 v[4];
#pragma unroll
          for (int u = 0; u < 4; u++) if (i + u * PB_POL_THREADS < n4) v[u] = src[i + u * PB_POL_THREADS];
#pragma unroll
          for (int u = 0; u < 4; u++) if (i + u * PB_POL_THREADS < n4) dst[i + u * PB_POL_THREADS] = v[u];
        }
        for (int i = (n4 << 2) + tid; i < nfl; i += PB_POL_THREADS) a.obs_row[base + i] = a.obs_in[base + i];
      } else {
        for (int i = tid; i < nfl; i += PB_POL_THREADS) a.obs_row[base + i] = a.obs_in[base + i];
      }
    }
    const float b = tid < 64 ? a.params[a.L.cb1 + tid] : a.params[a.L.ab1 + tid - 64];
#pragma unroll
    for (int e = 0; e < PB_POL_TILE; e++) {
      const int env = e0 + e;
      float v = b;
      if (env < a.num_envs) { v += a.hpre[(size_t)env * 128 + tid]; if (a.hpre2) v += a.hpre2[(size_t)env * 128 + tid]; }
      h1[tid * PB_POL_TILE + e] = tanhf(v);
    }
  } else {
    // stage the observation tile transposed (and copy it into the rollout row)
    for (int i = tid; i < PB_POL_TILE * O; i += PB_POL_THREADS) {
      int e = i / O, k = i - e * O, env = e0 + e;
      float v = env < a.num_envs ? a.obs_in[(size_t)env * O + k] : 0.f;
      xT[k * PB_POL_TILE + e] = v;
      if (a.obs_row && env < a.num_envs) a.obs_row[(size_t)env * O + k] = v;
    }
    __syncthreads();
    // layer 1: unit tid of the concatenated [critic | actor] hidden layer
    float b = tid < 64 ? a.params[a.L.cb1 + tid] : a.params[a.L.ab1 + tid - 64];
#pragma unroll
    for (int e = 0; e < PB_POL_TILE; e++) acc[e] = b;
    for (int k = 0; k < O; k++) {
      float w = a.W1T[k * 128 + tid];
      const float4 *x4 = reinterpret_cast<const float4 *>(xT + k * PB_POL_TILE);
#pragma unroll
      for (int q = 0; q < PB_POL_TILE / 4; q++) {
        float4 x = x4[q];
        acc[4 * q] += w * x.x; acc[4 * q + 1] += w * x.y; acc[4 * q + 2] += w * x.z; acc[4 * q + 3] += w * x.w;
      }
    }
#pragma unroll
    for (int e = 0; e < PB_POL_TILE; e++) h1[tid * PB_POL_TILE + e] = tanhf(acc[e]);
  }
  __syncthreads();
  {  // layer 2 of each net
    int net = tid >> 6, u = tid & 63;
    float b = a.params[(net ? a.L.ab2 : a.L.cb2) + u];
#pragma unroll
    for (int e = 0; e < PB_POL_TILE; e++) acc[e] = b;
    const float *W = a.W2T + net * 4096;
    for (int k = 0; k < 64; k++) {
      float w = W[k * 64 + u];
      const float4 *x4 = reinterpret_cast<const float4 *>(h1 + (net * 64 + k) * PB_POL_TILE);
#pragma unroll
      for (int q = 0; q < PB_POL_TILE / 4; q++) {
        float4 x = x4[q];
        acc[4 * q] += w * x.x; acc[4 * q + 1] += w * x.y; acc[4 * q + 2] += w * x.z; acc[4 * q + 3] += w * x.w;
      }
    }
#pragma unroll
    for (int e = 0; e < PB_POL_TILE; e++) h2[tid * PB_POL_TILE + e] = tanhf(acc[e]);
  }
  __syncthreads();
  // layer 3: (env, output) pairs; output A is the value.  The head rows ((A + 1) x 64 floats) are staged in shared memory (the
  // observation tile's slot when it is free and large enough): read from global memory the 64-step dot products were a chain of
  // dependent L1 / L2 round trips (28 % of the stall samples).
  const bool w3_smem = a.hpre && O * PB_POL_TILE >= (A + 1) * 64;
  if (w3_smem) {
    for (int i = tid; i < (A + 1) * 64; i += PB_POL_THREADS) {
      const int o = i >> 6, k = i & 63;
      xT[i] = o < A ? a.params[a.L.aW3 + o * 64 + k] : a.params[a.L.cW3 + k];
    }
    __syncthreads();
  }
  for (int i = tid; i < PB_POL_TILE * (A + 1); i += PB_POL_THREADS) {
    int e = i / (A + 1), o = i - e * (A + 1);
    const float *w = w3_smem ? xT + o * 64 : (o < A ? a.params + a.L.aW3 + o * 64 : a.params + a.L.cW3);
    const float *h = h2 + (o < A ? 64 : 0) * PB_POL_TILE + e;
    float s = o < A ? a.params[a.L.ab3 + o] : a.params[a.L.cb3];
    for (int k = 0; k < 64; k++) s += w[k] * h[k * PB_POL_TILE];
    outv[e * (A + 1) + o] = s;
  }
  __syncthreads();
  if (tid < PB_POL_TILE) {
    int env = e0 + tid;
    if (env < a.num_envs) {
      const float *o = outv + tid * (A + 1);
      if (a.val_row) a.val_row[env] = o[A];
      if (a.done_row) a.done_row[env] = (a.term_in[env] != 0.f || a.trunc_in[env] != 0.f) ? 1.f : 0.f;
      if (a.sample) {
        float logp = 0.f;
        for (int i = 0; i < A; i += 4) {
          uint32_t c[4] = {(uint32_t)env, a.step_index, (uint32_t)(i >> 2), 0x5eedu};
          philox4x32_10(c, a.seed_lo, a.seed_hi);
          float z[4];
          box_muller(c[0], c[1], &z[0], &z[1]);
          box_muller(c[2], c[3], &z[2], &z[3]);
          for (int j = 0; j < 4 && i + j < A; j++) {
            float ls = a.params[a.L.logstd + i + j], sd = expf(ls), mean = o[i + j];
            float act = mean + sd * z[j];
            a.act_row[(size_t)env * A + i + j] = act;
            float d = act - mean;
            logp += -(d * d) / (2.f * sd * sd) - ls - 0.91893853320467274178f;
          }
        }
        a.logp_row[env] = logp;
      }
    }
  }
}

// GAE reverse scan, one thread per env; rows are [T][N] so a warp reads consecutive addresses at every t.
static __global__ void gae_kernel(int T, int N, const float *__restrict__ rewards, const float *__restrict__ values,
                           const float *__restrict__ dones, const float *__restrict__ next_value,
                           const float *__restrict__ term, const float *__restrict__ trunc, float gamma, float lam,
                           float *__restrict__ advantages, float *__restrict__ returns) {
  int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  float last = 0.f, nextval = next_value[n];
  float nextnonterm = 1.0f - ((term[n] != 0.f || trunc[n] != 0.f) ? 1.f : 0.f);
  for (int t = T - 1; t >= 0; t--) {
    size_t i = (size_t)t * N + n;
    float v = values[i];
    float delta = rewards[i] + gamma * nextval * nextnonterm - v;
    last = delta + gamma * lam * nextnonterm * last;
    advantages[i] = last;
    returns[i] = last + v;
    nextnonterm = 1.0f - dones[i];
    nextval = v;
  }
}
