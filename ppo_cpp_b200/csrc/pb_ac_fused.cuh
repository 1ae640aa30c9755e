// pb_ac_fused.cuh — row kernels of the ac_ppo update that stream a [rows x 256] activation matrix exactly once.
//
// A 262 144-row minibatch makes every 256-wide activation matrix 268 MB, twice the L2: the update is bound by how many
// times those matrices cross HBM, not by arithmetic (src/ac_ppo_continuous_action.cpp:816-888 through libtorch makes one
// pass per op).  The kernels here fuse the ops that share a row:
//   ac_l1_ln_fwd_kernel      Linear(obs -> 256) + LayerNorm (+ ReLU)        for obs_dim <= 24          (:150-172 encoder, layer 1)
//   ac_heads_fwd_kernel      all head products of one net in one pass over h2                          (dist_alpha | dist_beta, critic.6)
//   ac_heads_ln_bwd_kernel   heads backward + ReLU/LayerNorm-2 backward + head weight / bias gradients + dgamma / dbeta partials
//   ac_w1_grad_kernel        first-layer weight / bias gradient dW1 = d1^T Xn                          for obs_dim <= 24
//   ac_ln_bwd_w1_kernel      ReLU/LayerNorm-1 backward + that gradient while d1 is in registers (d1 is never stored)
//   ac_reduce_segs_kernel    fixed-order sums of all per-block partial vectors of a minibatch in one launch
// One warp owns a row; lane l owns columns [4l, 4l + 4) and [128 + 4l, 128 + 4l + 4), so every 16-byte access of a warp
// covers 512 contiguous bytes in global memory and is conflict-free in shared memory.  Blocks are persistent (one or two per
// SM), each owning a contiguous range of rows; rows are taken several at a time so that enough loads are in flight to cover
// the HBM latency.  Per-block partial sums are reduced in a fixed order: results do not depend on the launch.
#pragma once
#include "pb_ac_kernels.cuh"

namespace acf {

__device__ __forceinline__ void ld8(const float *row, int lane, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4 *>(row + 4 * lane), b = *reinterpret_cast<const float4 *>(row + 128 + 4 * lane);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void st8(float *row, int lane, const float (&v)[8]) {
  *reinterpret_cast<float4 *>(row + 4 * lane) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4 *>(row + 128 + 4 * lane) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ int col8(int lane, int i) { return i < 4 ? 4 * lane + i : 128 + 4 * lane + (i - 4); }
__device__ __forceinline__ float warp_sum(float s) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  return s;
}

// ------------------------------------------------------------------ layer 1: pre = Xn W1^T + b1, LayerNorm, ReLU
// K = obs_dim <= AC_L1_MAXO.  W1^T lives in shared memory ([k][256]); a warp takes 4 rows at a time, whose observations it
// stages in its own shared-memory slot and reads back as broadcasts.  Writes x_hat (and h = relu(gamma x_hat + beta) when
// h != nullptr) and rstd, exactly what ac_gemm + ac_ln_relu_fwd_kernel produced in three passes.
#define AC_L1_MAXO 24
struct AcL1Fwd {
  const float *Xn; int O; const float *W, *bias, *gamma, *beta;
  float *xhat, *h, *rstd; int M;
};
static __global__ void __launch_bounds__(256, 2) ac_l1_ln_fwd_kernel(AcL1Fwd a) {
  __shared__ __align__(16) float Wt[AC_L1_MAXO][AC_HID];
  __shared__ __align__(16) float xs[8][4][AC_L1_MAXO];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, O = a.O;
  for (int i = tid; i < O * AC_HID; i += 256) { const int c = i / O, k = i - c * O; Wt[k][c] = a.W[i]; }
  float g[8], be[8], bi[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { const int c = col8(lane, i); g[i] = a.gamma[c]; be[i] = a.beta[c]; bi[i] = a.bias[c]; }
  __syncthreads();
  const int ngroups = (a.M + 31) / 32;   // 8 warps x 4 rows
  for (int grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
    const int r0 = grp * 32 + warp * 4;
    __syncwarp();
    for (int i = lane; i < 4 * O; i += 32) {
      const int r = i / O, k = i - r * O;
      xs[warp][r][k] = (r0 + r < a.M) ? a.Xn[(size_t)(r0 + r) * O + k] : 0.f;
    }
    __syncwarp();
    float acc[4][8];
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
      for (int i = 0; i < 8; i++) acc[r][i] = 0.f;
    for (int k = 0; k < O; k++) {
      const float4 wa = *reinterpret_cast<const float4 *>(&Wt[k][4 * lane]), wb = *reinterpret_cast<const float4 *>(&Wt[k][128 + 4 * lane]);
      const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
      for (int r = 0; r < 4; r++) {
        const float x = xs[warp][r][k];
#pragma unroll
        for (int i = 0; i < 8; i++) acc[r][i] += x * w[i];
      }
    }
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int row = r0 + r;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < 8; i++) { acc[r][i] += bi[i]; s += acc[r][i]; }
      const float mean = warp_sum(s) * (1.f / AC_HID);
      float v = 0.f;
#pragma unroll
      for (int i = 0; i < 8; i++) { acc[r][i] -= mean; v += acc[r][i] * acc[r][i]; }
      const float rs = rsqrtf(warp_sum(v) * (1.f / AC_HID) + 1e-5f);
      if (row < a.M) {
        float y[8];
#pragma unroll
        for (int i = 0; i < 8; i++) { acc[r][i] *= rs; y[i] = fmaxf(0.f, g[i] * acc[r][i] + be[i]); }
        st8(a.xhat + (size_t)row * AC_HID, lane, acc[r]);
        if (a.h) st8(a.h + (size_t)row * AC_HID, lane, y);
        if (lane == 0) a.rstd[row] = rs;
      }
    }
  }
}

// ------------------------------------------------------------------ heads forward: out[m][o] = b[o] + sum_k h[m][k] W[o][k]
// Two weight blocks (n0 + n1 <= 32 outputs: alpha | beta, or the critic's single row), one pass over h.  A warp takes 4
// rows; the 4 x NO partial dot products of a lane are summed across the warp with the transposing butterfly (31 shuffles
// per 32 values instead of 5 each).  When xhat != nullptr the input is recomputed as relu(gamma xhat + beta).
struct AcHeadsFwd {
  const float *h, *gamma, *beta;   // gamma == nullptr: h is the activation itself
  const float *W0, *b0; int n0; const float *W1, *b1; int n1;
  float *out; int ldo, M;
};
// Transposing butterfly: every lane brings 32 values; on return lane l holds the warp-wide total of value l.  Step s keeps, in
// lanes with bit s set, the upper half of the remaining values and hands the other half to the partner lane: 31 shuffles.
__device__ __forceinline__ float butterfly32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16, n = 16; s > 0; s >>= 1, n >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int i = 0; i < n; i++) {
      const float send = up ? v[i] : v[i + n], keep = up ? v[i + n] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}
template <int NOC>   // outputs per butterfly pass and row: 1 (critic) or 8; 4 rows x NOC values <= 32
static __global__ void __launch_bounds__(256, 2) ac_heads_fwd_kernel(AcHeadsFwd a) {
  extern __shared__ __align__(16) float hw[];   // [NO][256]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NO = a.n0 + a.n1;
  for (int i = tid; i < NO * AC_HID; i += 256) { const int o = i >> 8; hw[i] = o < a.n0 ? a.W0[i] : a.W1[i - a.n0 * AC_HID]; }
  float g[8], be[8];
  if (a.gamma) {
#pragma unroll
    for (int i = 0; i < 8; i++) { const int c = col8(lane, i); g[i] = a.gamma[c]; be[i] = a.beta[c]; }
  }
  __syncthreads();
  constexpr int RPW = 32 / NOC > 4 ? 4 : 32 / NOC;   // rows per warp pass
  const int ngroups = (a.M + 8 * RPW - 1) / (8 * RPW);
  for (int grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
    const int r0 = grp * 8 * RPW + warp * RPW;
    float x[RPW][8];
#pragma unroll
    for (int r = 0; r < RPW; r++) {
      if (r0 + r < a.M) ld8(a.h + (size_t)(r0 + r) * AC_HID, lane, x[r]);
      else {
#pragma unroll
        for (int i = 0; i < 8; i++) x[r][i] = 0.f;
      }
      if (a.gamma) {
#pragma unroll
        for (int i = 0; i < 8; i++) x[r][i] = fmaxf(0.f, g[i] * x[r][i] + be[i]);
      }
    }
    for (int o0 = 0; o0 < NO; o0 += NOC) {
      float v[32];
#pragma unroll
      for (int i = 0; i < 32; i++) v[i] = 0.f;
#pragma unroll
      for (int oo = 0; oo < NOC; oo++) {
        const int o = o0 + oo;
        if (o < NO) {
          const float4 wa = *reinterpret_cast<const float4 *>(hw + o * AC_HID + 4 * lane), wb = *reinterpret_cast<const float4 *>(hw + o * AC_HID + 128 + 4 * lane);
#pragma unroll
          for (int r = 0; r < RPW; r++)
            v[r * NOC + oo] = x[r][0] * wa.x + x[r][1] * wa.y + x[r][2] * wa.z + x[r][3] * wa.w + x[r][4] * wb.x + x[r][5] * wb.y + x[r][6] * wb.z + x[r][7] * wb.w;
        }
      }
      const float tot = butterfly32(v, lane);
      const int j = lane, r = j / NOC, oo = j - r * NOC, o = o0 + oo, row = r0 + r;
      if (r < RPW && o < NO && row < a.M) a.out[(size_t)row * a.ldo + o] = tot + (o < a.n0 ? a.b0[o] : a.b1[o - a.n0]);
    }
  }
}

// ------------------------------------------------------------------ heads backward + ReLU / LayerNorm-2 backward + head gradients
// Per row: dh = dz W_heads, dy = dh masked by relu, (dgamma, dbeta) += (dy x_hat, dy), dpre = rstd (dy gamma - mean - x_hat mean'),
// dW_head[o] += dz[o] h, db_head[o] += dz[o]; h = relu(gamma x_hat + beta) is recomputed, so the layer's post-activation matrix is
// never read.  One pass: reads x_hat, writes dpre.  Per-block partials: [dgamma 256 | dbeta 256 | dW_head NOC x 256 | db_head 32].
struct AcHeadsLnBwd {
  const float *dz; int ldz;
  const float *W0; int n0; const float *W1; int n1;
  const float *xhat, *rstd, *gamma, *beta;
  float *d;
  float *part;
  int M;
};
template <int NOC> __host__ __device__ constexpr int heads_part_ld() { return 512 + NOC * AC_HID + 32; }

#define HLB_DEPTH 6
template <int NOC> __host__ __device__ constexpr int heads_smem_bytes() { return 8 * HLB_DEPTH * (2 * 32 * 16 + 32 * 4) + (NOC * AC_HID + heads_part_ld<NOC>()) * 4; }
__device__ __forceinline__ void acf_cp16(void *dst, const void *src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void acf_cp4(void *dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}

// The rows of a warp stream through a private cp.async ring (HLB_DEPTH rows: x_hat, the dz row, rstd) that every lane fills and
// reads back for its own columns - enough bytes in flight without holding them in registers (see ac_ln_bwd_w1_kernel).
template <int NOC>
static __global__ void __launch_bounds__(256, 1) ac_heads_ln_bwd_kernel(AcHeadsLnBwd a) {
  extern __shared__ __align__(16) unsigned char hlb_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NO = a.n0 + a.n1;
  float4 *ring4 = reinterpret_cast<float4 *>(hlb_smem) + warp * HLB_DEPTH * 2 * 32;                          // [depth][2][lane]
  float *ring1 = reinterpret_cast<float *>(hlb_smem + 8 * HLB_DEPTH * 2 * 32 * 16) + warp * HLB_DEPTH * 32;  // [depth][lane]: dz | rstd
  float *hw = reinterpret_cast<float *>(hlb_smem + 8 * HLB_DEPTH * (2 * 32 * 16 + 32 * 4));                  // [NOC][256] head weights
  float *accs = hw + NOC * AC_HID;                                                                           // block accumulator
  for (int i = tid; i < NOC * AC_HID; i += 256) {
    const int o = i >> 8;
    hw[i] = o < a.n0 ? a.W0[i] : (o < NO ? a.W1[i - a.n0 * AC_HID] : 0.f);
  }
  float g[8], be[8], dg[8], db[8], dW[NOC][8], hb = 0.f;
#pragma unroll
  for (int i = 0; i < 8; i++) { const int c = col8(lane, i); g[i] = a.gamma[c]; be[i] = a.beta[c]; dg[i] = 0.f; db[i] = 0.f; }
#pragma unroll
  for (int o = 0; o < NOC; o++)
#pragma unroll
    for (int i = 0; i < 8; i++) dW[o][i] = 0.f;
  __syncthreads();
  const int chunk = (a.M + gridDim.x - 1) / gridDim.x, rb = blockIdx.x * chunk, re = min(a.M, rb + chunk);
  const int nrows = re > rb + warp ? (re - rb - warp + 7) / 8 : 0;
  auto issue = [&](int j) {
    if (j < nrows) {
      const size_t row = (size_t)(rb + warp + 8 * j);
      float4 *slot = ring4 + (j % HLB_DEPTH) * 2 * 32;
      acf_cp16(slot + lane, a.xhat + row * AC_HID + 4 * lane);
      acf_cp16(slot + 32 + lane, a.xhat + row * AC_HID + 128 + 4 * lane);
      if (lane < NO) acf_cp4(ring1 + (j % HLB_DEPTH) * 32 + lane, a.dz + row * a.ldz + lane);
      if (lane == 31) acf_cp4(ring1 + (j % HLB_DEPTH) * 32 + 31, a.rstd + row);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int p = 0; p < HLB_DEPTH - 1; p++) issue(p);
  for (int j = 0; j < nrows; j++) {
    issue(j + HLB_DEPTH - 1);
    asm volatile("cp.async.wait_group %0;" ::"n"(HLB_DEPTH - 1) : "memory");
    __syncwarp();
    const float4 *slot = ring4 + (j % HLB_DEPTH) * 2 * 32;
    const float4 x0 = slot[lane], x1 = slot[32 + lane];
    const float rs = ring1[(j % HLB_DEPTH) * 32 + 31];
    const float z = lane < NO ? ring1[(j % HLB_DEPTH) * 32 + lane] : 0.f;
    const float x[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
    float dh[8], h[8], zo[NOC];
#pragma unroll
    for (int i = 0; i < 8; i++) dh[i] = 0.f;
#pragma unroll
    for (int o = 0; o < NOC; o++) {
      const float4 wa = *reinterpret_cast<const float4 *>(hw + o * AC_HID + 4 * lane), wb = *reinterpret_cast<const float4 *>(hw + o * AC_HID + 128 + 4 * lane);
      const float w[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
      zo[o] = __shfl_sync(0xffffffffu, z, o);
#pragma unroll
      for (int i = 0; i < 8; i++) dh[i] += zo[o] * w[i];
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const float hv = g[i] * x[i] + be[i];
      const bool on = hv > 0.f;
      h[i] = on ? hv : 0.f;
      const float dy = on ? dh[i] : 0.f;
      dg[i] += dy * x[i]; db[i] += dy;
      const float dxh = dy * g[i];
      dh[i] = dxh;
      s1 += dxh; s2 += dxh * x[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
    const float m1 = s1 * (1.f / AC_HID), m2 = s2 * (1.f / AC_HID);
    float out[8];
#pragma unroll
    for (int i = 0; i < 8; i++) out[i] = rs * (dh[i] - m1 - x[i] * m2);
    st8(a.d + (size_t)(rb + warp + 8 * j) * AC_HID, lane, out);
#pragma unroll
    for (int o = 0; o < NOC; o++)
#pragma unroll
      for (int i = 0; i < 8; i++) dW[o][i] += zo[o] * h[i];
    hb += z;
    __syncwarp();   // every lane has read the slot's shared words before a later copy may overwrite them
  }
  // block reduction, warps in ascending order
  for (int w = 0; w < 8; w++) {
    if (warp == w) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int c = col8(lane, i);
        if (w == 0) { accs[c] = dg[i]; accs[256 + c] = db[i]; } else { accs[c] += dg[i]; accs[256 + c] += db[i]; }
#pragma unroll
        for (int o = 0; o < NOC; o++) { if (w == 0) accs[512 + o * AC_HID + c] = dW[o][i]; else accs[512 + o * AC_HID + c] += dW[o][i]; }
      }
      if (w == 0) accs[512 + NOC * AC_HID + lane] = hb; else accs[512 + NOC * AC_HID + lane] += hb;
    }
    __syncthreads();
  }
  constexpr int LD = heads_part_ld<NOC>();
  for (int i = tid; i < LD; i += 256) a.part[(size_t)blockIdx.x * LD + i] = accs[i];
}

// ------------------------------------------------------------------ first-layer weight gradient dW1[c][k] = sum_m d1[m][c] Xn[m][k]
// obs_dim = O is a template parameter (O x 8 accumulators per lane).  Per-block partials in the parameter layout: [c * O + k | bias 256].
struct AcW1Grad { const float *d1, *Xn; float *part; int M; };
template <int O>
static __global__ void __launch_bounds__(256, 1) ac_w1_grad_kernel(AcW1Grad a) {
  extern __shared__ __align__(16) float accs[];   // [256 * O + 256]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float acc[O][8], bacc[8];
#pragma unroll
  for (int i = 0; i < 8; i++) bacc[i] = 0.f;
#pragma unroll
  for (int k = 0; k < O; k++)
#pragma unroll
    for (int i = 0; i < 8; i++) acc[k][i] = 0.f;
  const int chunk = (a.M + gridDim.x - 1) / gridDim.x, rb = blockIdx.x * chunk, re = min(a.M, rb + chunk);
  float nd[2][8], nx[2];
  auto fetch = [&](int row, float (&d)[8], float &x) {
    if (row < re) { ld8(a.d1 + (size_t)row * AC_HID, lane, d); x = lane < O ? a.Xn[(size_t)row * O + lane] : 0.f; }
    else {
#pragma unroll
      for (int i = 0; i < 8; i++) d[i] = 0.f;
      x = 0.f;
    }
  };
  fetch(rb + warp, nd[0], nx[0]);
  fetch(rb + warp + 8, nd[1], nx[1]);
  for (int row0 = rb + warp; row0 < re; row0 += 16) {
    float d[2][8], x[2];
#pragma unroll
    for (int r = 0; r < 2; r++) {
#pragma unroll
      for (int i = 0; i < 8; i++) d[r][i] = nd[r][i];
      x[r] = nx[r];
    }
    fetch(row0 + 16, nd[0], nx[0]);
    fetch(row0 + 24, nd[1], nx[1]);
#pragma unroll
    for (int r = 0; r < 2; r++) {
#pragma unroll
      for (int i = 0; i < 8; i++) bacc[i] += d[r][i];
#pragma unroll
      for (int k = 0; k < O; k++) {
        const float xk = __shfl_sync(0xffffffffu, x[r], k);
#pragma unroll
        for (int i = 0; i < 8; i++) acc[k][i] += xk * d[r][i];
      }
    }
  }
  for (int w = 0; w < 8; w++) {
    if (warp == w) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int c = col8(lane, i);
        if (w == 0) accs[AC_HID * O + c] = bacc[i]; else accs[AC_HID * O + c] += bacc[i];
#pragma unroll
        for (int k = 0; k < O; k++) { if (w == 0) accs[c * O + k] = acc[k][i]; else accs[c * O + k] += acc[k][i]; }
      }
    }
    __syncthreads();
  }
  constexpr int LD = AC_HID * O + AC_HID;
  for (int i = tid; i < LD; i += 256) a.part[(size_t)blockIdx.x * LD + i] = accs[i];
}

// ------------------------------------------------------------------ ReLU / LayerNorm-1 backward + first-layer gradient, d1 never stored
// Input: dh = d2 W2 (the plain product of actc::gemm256_tc_kernel), x_hat1, rstd1.  Per row: dy = dh masked by gamma x_hat + beta > 0,
// (dgamma, dbeta) += (dy x_hat, dy), d1 = rstd (dy gamma - mean - x_hat mean'), and then dW1[c][k] += d1[c] Xn[k], db1 += d1 while
// d1 is still in registers (O > 0; obs_dim is a template parameter), or d1 is stored for a separate weight-gradient kernel (O == 0).
// The rows of a warp stream through a private cp.async ring (LNB_DEPTH rows of 2 KB + the observation row) that every lane fills
// and reads back for its own columns only - no barrier, and enough bytes in flight without holding them in registers.
// Per-block partials: [dgamma 256 | dbeta 256 | dW1 in the parameter layout c * O + k | db1 256].
#define LNB_DEPTH 6
struct AcLnBwdW1 {
  const float *dh, *xhat, *rstd, *gamma, *beta, *Xn;
  float *dout;     // O == 0 only
  float *part;
  int M;
};
template <int O> __host__ __device__ constexpr int lnb_part_ld() { return 512 + (O > 0 ? AC_HID * O + AC_HID : 0); }
template <int O> __host__ __device__ constexpr int lnb_smem_bytes() { return 8 * LNB_DEPTH * (4 * 32 * 16 + 32 * 4) + lnb_part_ld<O>() * 4; }


template <int O>
static __global__ void __launch_bounds__(256, 1) ac_ln_bwd_w1_kernel(AcLnBwdW1 a) {
  extern __shared__ __align__(16) unsigned char lnb_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float4 *ring4 = reinterpret_cast<float4 *>(lnb_smem) + warp * LNB_DEPTH * 4 * 32;                    // [depth][piece 4][lane]
  float *ring1 = reinterpret_cast<float *>(lnb_smem + 8 * LNB_DEPTH * 4 * 32 * 16) + warp * LNB_DEPTH * 32;   // [depth][lane]: Xn | rstd
  float *accs = reinterpret_cast<float *>(lnb_smem + 8 * LNB_DEPTH * (4 * 32 * 16 + 32 * 4));
  constexpr int OK = O > 0 ? O : 1;
  float g[8], be[8], dg[8], db[8], acc[OK][8], bacc[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { const int c = col8(lane, i); g[i] = a.gamma[c]; be[i] = a.beta[c]; dg[i] = 0.f; db[i] = 0.f; bacc[i] = 0.f; }
#pragma unroll
  for (int k = 0; k < OK; k++)
#pragma unroll
    for (int i = 0; i < 8; i++) acc[k][i] = 0.f;
  const int chunk = (a.M + gridDim.x - 1) / gridDim.x, rb = blockIdx.x * chunk, re = min(a.M, rb + chunk);
  const int nrows = re > rb + warp ? (re - rb - warp + 7) / 8 : 0;
  auto issue = [&](int j) {
    if (j < nrows) {
      const size_t row = (size_t)(rb + warp + 8 * j);
      float4 *slot = ring4 + (j % LNB_DEPTH) * 4 * 32;
      acf_cp16(slot + lane, a.dh + row * AC_HID + 4 * lane);
      acf_cp16(slot + 32 + lane, a.dh + row * AC_HID + 128 + 4 * lane);
      acf_cp16(slot + 64 + lane, a.xhat + row * AC_HID + 4 * lane);
      acf_cp16(slot + 96 + lane, a.xhat + row * AC_HID + 128 + 4 * lane);
      if (O > 0 && lane < O) acf_cp4(ring1 + (j % LNB_DEPTH) * 32 + lane, a.Xn + row * O + lane);
      if (lane == 31) acf_cp4(ring1 + (j % LNB_DEPTH) * 32 + 31, a.rstd + row);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
#pragma unroll
  for (int p = 0; p < LNB_DEPTH - 1; p++) issue(p);
  for (int j = 0; j < nrows; j++) {
    issue(j + LNB_DEPTH - 1);
    asm volatile("cp.async.wait_group %0;" ::"n"(LNB_DEPTH - 1) : "memory");
    __syncwarp();   // the rstd / Xn words were copied by other lanes
    const float4 *slot = ring4 + (j % LNB_DEPTH) * 4 * 32;
    const float4 d0 = slot[lane], d1 = slot[32 + lane], x0 = slot[64 + lane], x1 = slot[96 + lane];
    const float rs = ring1[(j % LNB_DEPTH) * 32 + 31];
    const float xo = (O > 0 && lane < O) ? ring1[(j % LNB_DEPTH) * 32 + lane] : 0.f;
    const float dh[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w}, x[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
    float dxh[8], s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const float dy = (g[i] * x[i] + be[i] > 0.f) ? dh[i] : 0.f;
      dg[i] += dy * x[i]; db[i] += dy;
      dxh[i] = dy * g[i];
      s1 += dxh[i]; s2 += dxh[i] * x[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
    const float m1 = s1 * (1.f / AC_HID), m2 = s2 * (1.f / AC_HID);
    float d[8];
#pragma unroll
    for (int i = 0; i < 8; i++) d[i] = rs * (dxh[i] - m1 - x[i] * m2);
    if (O > 0) {
#pragma unroll
      for (int i = 0; i < 8; i++) bacc[i] += d[i];
#pragma unroll
      for (int k = 0; k < OK; k++) {
        const float xk = __shfl_sync(0xffffffffu, xo, k);
#pragma unroll
        for (int i = 0; i < 8; i++) acc[k][i] += xk * d[i];
      }
    } else {
      st8(a.dout + (size_t)(rb + warp + 8 * j) * AC_HID, lane, d);
    }
    __syncwarp();   // every lane has read the slot's shared words before a later copy may overwrite them
  }
  for (int w = 0; w < 8; w++) {
    if (warp == w) {
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const int c = col8(lane, i);
        if (w == 0) { accs[c] = dg[i]; accs[256 + c] = db[i]; } else { accs[c] += dg[i]; accs[256 + c] += db[i]; }
        if (O > 0) {
          if (w == 0) accs[512 + AC_HID * O + c] = bacc[i]; else accs[512 + AC_HID * O + c] += bacc[i];
#pragma unroll
          for (int k = 0; k < OK; k++) { if (w == 0) accs[512 + c * O + k] = acc[k][i]; else accs[512 + c * O + k] += acc[k][i]; }
        }
      }
    }
    __syncthreads();
  }
  constexpr int LD = lnb_part_ld<O>();
  for (int i = tid; i < LD; i += 256) a.part[(size_t)blockIdx.x * LD + i] = accs[i];
}

// ------------------------------------------------------------------ all partial-sum reductions of one minibatch in one launch
#define AC_MAXSEG 16
struct AcReduceSeg { const float *part; size_t ld; int nblk, ncol; float *out0, *out1; int split; };
struct AcReduceSegs { AcReduceSeg s[AC_MAXSEG]; };
static __global__ void __launch_bounds__(1024) ac_reduce_segs_kernel(AcReduceSegs S) {
  __shared__ float red[32][33];
  const AcReduceSeg &q = S.s[blockIdx.y];
  if ((int)blockIdx.x * 32 >= q.ncol) return;
  const int cx = threadIdx.x & 31, g = threadIdx.x >> 5, c = blockIdx.x * 32 + cx;
  float s = 0.f;
  if (c < q.ncol) for (int b = g; b < q.nblk; b += 32) s += q.part[(size_t)b * q.ld + c];
  red[g][cx] = s;
  __syncthreads();
  if (g == 0 && c < q.ncol) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 32; w++) t += red[w][cx];
    if (c < q.split) q.out0[c] = t; else q.out1[c - q.split] = t;
  }
}

}  // namespace acf
