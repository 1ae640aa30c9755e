// pb_ac_kernels.cuh — kernels of the ac_ppo_continuous_action agent (src/ac_ppo_continuous_action.cpp:150-266):
// observation standardisation with fixed mean/std, Linear -> LayerNorm -> ReLU (x2, 256 wide) critic and actor encoder,
// Beta(alpha, beta) policy with alpha, beta = softplus(Linear(256, A)) + 1 (include/rl_utils.h:48-132), forward and
// hand-derived backward.  fp32 on the FFMA pipe; the per-sample special functions (lgamma, digamma, trigamma) run in fp64.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <stdint.h>

#include "pb_ppo_common.cuh"
#include "pb_ppo_update.cuh"

#define AC_HID 256
#define AC_MAXA 20

struct AcLayout {
  int obs, act;
  // torch registration order: critic.{0,1,3,4,6}.{weight,bias}, actor_mean.{0,1,3,4}, dist_alpha.0, dist_beta.0
  int cW1, cb1, cg1, cB1, cW2, cb2, cg2, cB2, cW3, cb3, aW1, ab1, ag1, aB1, aW2, ab2, ag2, aB2, Wal, bal, Wbe, bbe, total;
  __host__ __device__ static AcLayout make(int O, int A) {
    AcLayout l;
    l.obs = O; l.act = A;
    int p = 0;
    l.cW1 = p; p += AC_HID * O; l.cb1 = p; p += AC_HID; l.cg1 = p; p += AC_HID; l.cB1 = p; p += AC_HID;
    l.cW2 = p; p += AC_HID * AC_HID; l.cb2 = p; p += AC_HID; l.cg2 = p; p += AC_HID; l.cB2 = p; p += AC_HID;
    l.cW3 = p; p += AC_HID; l.cb3 = p; p += 1;
    l.aW1 = p; p += AC_HID * O; l.ab1 = p; p += AC_HID; l.ag1 = p; p += AC_HID; l.aB1 = p; p += AC_HID;
    l.aW2 = p; p += AC_HID * AC_HID; l.ab2 = p; p += AC_HID; l.ag2 = p; p += AC_HID; l.aB2 = p; p += AC_HID;
    l.Wal = p; p += A * AC_HID; l.bal = p; p += A; l.Wbe = p; p += A * AC_HID; l.bbe = p; p += A;
    l.total = p;
    return l;
  }
};

// ------------------------------------------------------------------ dense products, 64x64 tiles, 256 threads, 4x4 per thread
// MODE_NT: C[m][n] = sum_k A[m][k] * B[n][k] (+ bias[n])      x @ W^T   (B = weight [N][K])
// MODE_NN: C[m][n] = sum_k A[m][k] * B[k][n]                  dY @ W    (B = weight [K][N])
enum { AC_NT = 0, AC_NN = 1 };
struct AcGemm { const float *A; int lda; const float *B; int ldb; float *C; int ldc; const float *bias; int M, N, K; };

template <int MODE>
__global__ void __launch_bounds__(256) ac_gemm_kernel(AcGemm g) {
  constexpr int BK = 16, LD = 68;
  __shared__ __align__(16) float As[2][BK][LD], Bs[2][BK][LD];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f;
  float ar[4], br[4];
  const int lr = tid >> 2, lk = (tid & 3) * 4;   // row-major "4 consecutive k of one row" loader
  const int nk = tid >> 4, nc = (tid & 15) * 4;  // "4 consecutive columns of one k" loader (MODE_NN B operand)
  auto gload = [&](int k0) {
    {
      const int m = m0 + lr;
      const float *p = g.A + (size_t)m * g.lda + k0 + lk;
#pragma unroll
      for (int j = 0; j < 4; j++) ar[j] = (m < g.M && k0 + lk + j < g.K) ? p[j] : 0.f;
    }
    if (MODE == AC_NT) {
      const int n = n0 + lr;
      const float *p = g.B + (size_t)n * g.ldb + k0 + lk;
#pragma unroll
      for (int j = 0; j < 4; j++) br[j] = (n < g.N && k0 + lk + j < g.K) ? p[j] : 0.f;
    } else {
      const int k = k0 + nk;
      const float *p = g.B + (size_t)k * g.ldb + n0 + nc;
#pragma unroll
      for (int j = 0; j < 4; j++) br[j] = (k < g.K && n0 + nc + j < g.N) ? p[j] : 0.f;
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int j = 0; j < 4; j++) As[buf][lk + j][lr] = ar[j];
    if (MODE == AC_NT) {
#pragma unroll
      for (int j = 0; j < 4; j++) Bs[buf][lk + j][lr] = br[j];
    } else {
      *reinterpret_cast<float4 *>(&Bs[buf][nk][nc]) = make_float4(br[0], br[1], br[2], br[3]);
    }
  };
  const int nst = (g.K + BK - 1) / BK;
  gload(0); sstore(0);
  __syncthreads();
  for (int s = 0; s < nst; s++) {
    if (s + 1 < nst) gload((s + 1) * BK);
    const int buf = s & 1;
#pragma unroll
    for (int kk = 0; kk < BK; kk++) {
      float4 a = *reinterpret_cast<const float4 *>(&As[buf][kk][ty * 4]);
      float4 b = *reinterpret_cast<const float4 *>(&Bs[buf][kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int i = 0; i < 4; i++) { acc[i][0] += av[i] * b.x; acc[i][1] += av[i] * b.y; acc[i][2] += av[i] * b.z; acc[i][3] += av[i] * b.w; }
    }
    if (s + 1 < nst) sstore((s + 1) & 1);
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int m = m0 + ty * 4 + i;
    if (m >= g.M) continue;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int n = n0 + tx * 4 + j;
      if (n < g.N) g.C[(size_t)m * g.ldc + n] = acc[i][j] + (g.bias ? g.bias[n] : 0.f);
    }
  }
}

// 128 x 128 output tile, 8 x 8 accumulators per thread (rows ty*4+i and 64+ty*4+i, columns tx*4+j and 64+tx*4+j), BK = 16,
// register-staged double buffering.  Per k step a thread issues 4 LDS.128 for 64 FFMA (the 64 x 64 tile above: 2 for 16),
// so the kernel is bound by the FMA pipe rather than by shared-memory bandwidth.  Used when the grid still fills the GPU
// (the minibatch GEMMs of the update: M = 65 536 ... 262 144 rows); same arithmetic per output (k ascending).
// VEC: rows are 16-byte aligned and K % 4 == 0 -> float4 global loads.
template <int MODE, bool VEC>
__global__ void __launch_bounds__(256, 2) ac_gemm128_kernel(AcGemm g) {
  constexpr int BK = 16, LD = 132;
  __shared__ __align__(16) float As[2][BK][LD], Bs[2][BK][LD];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * 128, n0 = blockIdx.x * 128;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
  float4 ar[2], br[2];
  const int lr = tid >> 2, lk = (tid & 3) * 4;   // "4 consecutive k of one row" loader: rows lr and lr + 64
  const int nk = tid >> 5, nc = (tid & 31) * 4;  // "4 consecutive columns of one k" loader (NN B operand): k rows nk and nk + 8
  auto ld4 = [&](const float *p, int valid) -> float4 {   // valid = number of in-range elements (0..4) from p
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid >= 4 && VEC) return *reinterpret_cast<const float4 *>(p);
    if (valid > 0) v.x = p[0];
    if (valid > 1) v.y = p[1];
    if (valid > 2) v.z = p[2];
    if (valid > 3) v.w = p[3];
    return v;
  };
  auto gload = [&](int k0) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int m = m0 + lr + 64 * h;
      ar[h] = ld4(g.A + (size_t)m * g.lda + k0 + lk, m < g.M ? g.K - (k0 + lk) : 0);
      if (MODE == AC_NT) {
        const int n = n0 + lr + 64 * h;
        br[h] = ld4(g.B + (size_t)n * g.ldb + k0 + lk, n < g.N ? g.K - (k0 + lk) : 0);
      } else {
        const int k = k0 + nk + 8 * h;
        br[h] = ld4(g.B + (size_t)k * g.ldb + n0 + nc, k < g.K ? g.N - (n0 + nc) : 0);
      }
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      As[buf][lk + 0][lr + 64 * h] = ar[h].x; As[buf][lk + 1][lr + 64 * h] = ar[h].y;
      As[buf][lk + 2][lr + 64 * h] = ar[h].z; As[buf][lk + 3][lr + 64 * h] = ar[h].w;
      if (MODE == AC_NT) {
        Bs[buf][lk + 0][lr + 64 * h] = br[h].x; Bs[buf][lk + 1][lr + 64 * h] = br[h].y;
        Bs[buf][lk + 2][lr + 64 * h] = br[h].z; Bs[buf][lk + 3][lr + 64 * h] = br[h].w;
      } else {
        *reinterpret_cast<float4 *>(&Bs[buf][nk + 8 * h][nc]) = br[h];
      }
    }
  };
  const int nst = (g.K + BK - 1) / BK;
  gload(0); sstore(0);
  __syncthreads();
  for (int s = 0; s < nst; s++) {
    if (s + 1 < nst) gload((s + 1) * BK);
    const int buf = s & 1, kend = g.K - s * BK;   // a partial last stage (K = 17: one row of sixteen) skips its zero rows
#pragma unroll
    for (int kk = 0; kk < BK; kk++) {
      if (kk >= kend) break;
      const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][kk][64 + tx * 4]);
      const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] += av[i] * bv[j];
    }
    if (s + 1 < nst) sstore((s + 1) & 1);
    __syncthreads();
  }
  // epilogue: bias once per column group, 16-byte stores when the output rows allow it (scalar stores from 16 lanes
  // at a 16-byte stride touch every sector four times)
  const bool vecc = (g.ldc % 4 == 0) && ((reinterpret_cast<uintptr_t>(g.C) & 15) == 0);
  float bz[8];
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const int n = n0 + (j >> 2) * 64 + tx * 4 + (j & 3);
    bz[j] = (g.bias && n < g.N) ? g.bias[n] : 0.f;
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const int m = m0 + (i >> 2) * 64 + ty * 4 + (i & 3);
    if (m >= g.M) continue;
#pragma unroll
    for (int jh = 0; jh < 2; jh++) {
      const int n = n0 + jh * 64 + tx * 4;
      float *c = g.C + (size_t)m * g.ldc + n;
      if (vecc && n + 3 < g.N) {
        *reinterpret_cast<float4 *>(c) = make_float4(acc[i][jh * 4] + bz[jh * 4], acc[i][jh * 4 + 1] + bz[jh * 4 + 1],
                                                     acc[i][jh * 4 + 2] + bz[jh * 4 + 2], acc[i][jh * 4 + 3] + bz[jh * 4 + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; j++)
          if (n + j < g.N) c[j] = acc[i][jh * 4 + j] + bz[jh * 4 + j];
      }
    }
  }
}

// ------------------------------------------------------------------ elementwise / row kernels
// Xn[m][k] = (X[m][k] - mean[k]) / std[k]   (get_value / get_action_and_value, :189,215)
__global__ void ac_standardise_kernel(const float *__restrict__ X, const float *__restrict__ mean, const float *__restrict__ std_,
                                      float *__restrict__ Xn, size_t n, int O) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { int k = (int)(i % O); Xn[i] = (X[i] - mean[k]) / std_[k]; }
}

// LayerNorm(256, eps 1e-5, biased variance) + ReLU, one warp per row, 8 columns per lane.
// pre is overwritten with x_hat; rstd[m] and h = relu(gamma * x_hat + beta) are written for the backward pass.
__global__ void __launch_bounds__(256) ac_ln_relu_fwd_kernel(float *__restrict__ pre, const float *__restrict__ gamma,
                                                             const float *__restrict__ beta, float *__restrict__ rstd,
                                                             float *__restrict__ h, int M) {
  const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  float *p = pre + (size_t)row * AC_HID + lane * 8;
  float4 a = *reinterpret_cast<const float4 *>(p), b = *reinterpret_cast<const float4 *>(p + 4);
  float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; i++) s += x[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s * (1.f / AC_HID);
  float v = 0.f;
#pragma unroll
  for (int i = 0; i < 8; i++) { x[i] -= mean; v += x[i] * x[i]; }
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const float rs = rsqrtf(v * (1.f / AC_HID) + 1e-5f);
  float y[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { x[i] *= rs; y[i] = fmaxf(0.f, gamma[lane * 8 + i] * x[i] + beta[lane * 8 + i]); }
  *reinterpret_cast<float4 *>(p) = make_float4(x[0], x[1], x[2], x[3]);
  *reinterpret_cast<float4 *>(p + 4) = make_float4(x[4], x[5], x[6], x[7]);
  float *q = h + (size_t)row * AC_HID + lane * 8;
  *reinterpret_cast<float4 *>(q) = make_float4(y[0], y[1], y[2], y[3]);
  *reinterpret_cast<float4 *>(q + 4) = make_float4(y[4], y[5], y[6], y[7]);
  if (lane == 0) rstd[row] = rs;
}

// Backward of ReLU + LayerNorm: d holds dL/dh on entry and dL/dpre on exit.  Each block owns a contiguous chunk of rows and
// writes its partial dgamma / dbeta sums to part[block][512] (fixed order; reduced by ac_colpart_reduce_kernel).
#define AC_LN_BLOCKS 592
__global__ void __launch_bounds__(256) ac_ln_relu_bwd_kernel(float *__restrict__ d, const float *__restrict__ h,
                                                             const float *__restrict__ xhat, const float *__restrict__ rstd,
                                                             const float *__restrict__ gamma, float *__restrict__ part, int M) {
  __shared__ float red[8][2 * AC_HID];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int chunk = (M + gridDim.x - 1) / gridDim.x, r0 = blockIdx.x * chunk, r1 = min(M, r0 + chunk);
  float g[8], dg[8], db[8];
#pragma unroll
  for (int i = 0; i < 8; i++) { g[i] = gamma[lane * 8 + i]; dg[i] = 0.f; db[i] = 0.f; }
  for (int row = r0 + warp; row < r1; row += 8) {
    const size_t o = (size_t)row * AC_HID + lane * 8;
    float4 a = *reinterpret_cast<const float4 *>(d + o), b = *reinterpret_cast<const float4 *>(d + o + 4);
    float4 ha = *reinterpret_cast<const float4 *>(h + o), hb = *reinterpret_cast<const float4 *>(h + o + 4);
    float4 xa = *reinterpret_cast<const float4 *>(xhat + o), xb = *reinterpret_cast<const float4 *>(xhat + o + 4);
    float dy[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    const float hv[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
    const float xh[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
    float s1 = 0.f, s2 = 0.f, dx[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (!(hv[i] > 0.f)) dy[i] = 0.f;     // ReLU mask
      dg[i] += dy[i] * xh[i]; db[i] += dy[i];
      dx[i] = dy[i] * g[i];                // dL/dx_hat
      s1 += dx[i]; s2 += dx[i] * xh[i];
    }
    for (int o2 = 16; o2 > 0; o2 >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o2); s2 += __shfl_xor_sync(0xffffffffu, s2, o2); }
    const float rs = rstd[row], m1 = s1 * (1.f / AC_HID), m2 = s2 * (1.f / AC_HID);
#pragma unroll
    for (int i = 0; i < 8; i++) dx[i] = rs * (dx[i] - m1 - xh[i] * m2);
    *reinterpret_cast<float4 *>(d + o) = make_float4(dx[0], dx[1], dx[2], dx[3]);
    *reinterpret_cast<float4 *>(d + o + 4) = make_float4(dx[4], dx[5], dx[6], dx[7]);
  }
#pragma unroll
  for (int i = 0; i < 8; i++) { red[warp][lane * 8 + i] = dg[i]; red[warp][AC_HID + lane * 8 + i] = db[i]; }
  __syncthreads();
  for (int c = threadIdx.x; c < 2 * AC_HID; c += 256) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; w++) s += red[w][c];
    part[(size_t)blockIdx.x * 2 * AC_HID + c] = s;
  }
}

// out[c] = sum_b part[b][c], c < ncol (fixed order)
__global__ void ac_colpart_reduce_kernel(const float *__restrict__ part, int nblk, int ncol, float *__restrict__ out0,
                                         float *__restrict__ out1, int split) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncol) return;
  float s = 0.f;
  for (int b = 0; b < nblk; b++) s += part[(size_t)b * ncol + c];
  if (c < split) out0[c] = s; else out1[c - split] = s;
}

// heads forward: out[m][o] = bias[o] + sum_k h[m][k] * W[o][k], o < NO (<= 32), one warp per row
__global__ void __launch_bounds__(256) ac_rowdot_kernel(const float *__restrict__ h, const float *__restrict__ W,
                                                        const float *__restrict__ bias, int NO, float *__restrict__ out, int ldo,
                                                        int col0, int M) {
  const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  const float *p = h + (size_t)row * AC_HID + lane * 8;
  float4 a = *reinterpret_cast<const float4 *>(p), b = *reinterpret_cast<const float4 *>(p + 4);
  const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  for (int o = 0; o < NO; o++) {
    const float *w = W + (size_t)o * AC_HID + lane * 8;
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i] * w[i];
    for (int q = 16; q > 0; q >>= 1) s += __shfl_xor_sync(0xffffffffu, s, q);
    if (lane == 0) out[(size_t)row * ldo + col0 + o] = s + bias[o];
  }
}

// heads backward: dh[m][k] = sum_o dz[m][col0 + o] * W[o][k], one warp per row (dz rows are ldz wide)
__global__ void __launch_bounds__(256) ac_heads_bwd_kernel(const float *__restrict__ dz, int ldz, const float *__restrict__ W0, int n0,
                                                           const float *__restrict__ W1, int n1, float *__restrict__ dh, int M) {
  const int lane = threadIdx.x & 31, row = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const float *z = dz + (size_t)row * ldz;
  for (int o = 0; o < n0 + n1; o++) {
    const float g = z[o];
    const float *w = (o < n0 ? W0 + (size_t)o * AC_HID : W1 + (size_t)(o - n0) * AC_HID) + lane * 8;
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] += g * w[i];
  }
  float *q = dh + (size_t)row * AC_HID + lane * 8;
  *reinterpret_cast<float4 *>(q) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  *reinterpret_cast<float4 *>(q + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
}

// ------------------------------------------------------------------ special functions (fp64; arguments >= 1 here)
__device__ inline double ac_digamma(double x) {
  double r = 0.0;
  while (x < 6.0) { r -= 1.0 / x; x += 1.0; }
  const double f = 1.0 / (x * x);
  return r + log(x) - 0.5 / x - f * (1.0 / 12.0 - f * (1.0 / 120.0 - f * (1.0 / 252.0 - f * (1.0 / 240.0 - f * (1.0 / 132.0)))));
}
__device__ inline double ac_trigamma(double x) {
  double r = 0.0;
  while (x < 6.0) { r += 1.0 / (x * x); x += 1.0; }
  const double f = 1.0 / (x * x);
  return r + 1.0 / x + 0.5 * f + (1.0 / x) * f * (1.0 / 6.0 - f * (1.0 / 30.0 - f * (1.0 / 42.0 - f * (1.0 / 30.0))));
}
__device__ inline float ac_softplus(float z) { return z > 20.f ? z : log1pf(expf(z)); }  // torch: beta 1, threshold 20
__device__ inline float ac_sigmoid(float z) { return 1.f / (1.f + expf(-z)); }
__device__ inline double ac_xlogy(double x, double y) { return x == 0.0 ? 0.0 : x * log(y); }

// ------------------------------------------------------------------ Beta heads: log-prob, entropy, clipped PPO loss, gradients
#define AC_NSTAT 6
struct AcLossArgs {
  const float *zab;     // [M][ldz]: alpha logits in [0, A), beta logits in [A, 2A)
  const float *v;       // [M] critic output
  int ldz, M, A;
  const float *actp; const float4 *scp;  // epoch-ordered gathered rows: env-scale action [.][A]; (logp_old, adv, ret, val)
  const int *cursor; int mb;             // rows of this minibatch start at cursor * mb
  const float *adv_moments;              // [2]: mean, 1/(std + 1e-8) (cross-rank when data parallel)
  PpoHyper hp; float act_low, act_high;
  float *dzab;          // [M][64] zero padded: dL/dalpha_logit, dL/dbeta_logit
  float *dvw;           // [M][64] zero padded: column 0 = dL/dv
  float *part;          // [blocks][AC_NSTAT] per-block sums
};

__global__ void __launch_bounds__(128) ac_beta_loss_kernel(AcLossArgs a) {
  __shared__ float red[4][AC_NSTAT];
  const int m = blockIdx.x * 128 + threadIdx.x, A = a.A;
  float stat[AC_NSTAT] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (m < a.M) {
    const size_t r = (size_t)(*a.cursor) * a.mb + m;
    const float *z = a.zab + (size_t)m * a.ldz, *act = a.actp + r * A;
    const float4 sc = a.scp[r];
    const double invB = 1.0 / (double)a.mb;
    double logp = 0.0, ent = 0.0;
    float dlp_a[AC_MAXA], dlp_b[AC_MAXA], dH_a[AC_MAXA], dH_b[AC_MAXA], sg_a[AC_MAXA], sg_b[AC_MAXA];
    for (int i = 0; i < A; i++) {
      const float al = ac_softplus(z[i]) + 1.0f, be = ac_softplus(z[A + i]) + 1.0f;
      // scale_action (:194-203): to [0,1], clamp to [1e-7, 1 + 1e-7] in float32
      float s = (act[i] - a.act_low) / (a.act_high - a.act_low);
      s = fminf(fmaxf(s, 0.0f + 1e-7f), 1.0f + 1e-7f);
      const double x = (double)s, y = (double)(1.0f - s), da = (double)al, dbt = (double)be, a0 = (double)(al + be);
      const double psi0 = ac_digamma(a0), psia = ac_digamma(da), psib = ac_digamma(dbt);
      logp += ac_xlogy(da - 1.0, x) + ac_xlogy(dbt - 1.0, y) + lgamma(a0) - lgamma(da) - lgamma(dbt);
      ent += lgamma(da) + lgamma(dbt) - lgamma(a0) - (2.0 - a0) * psi0 - (da - 1.0) * psia - (dbt - 1.0) * psib;
      dlp_a[i] = (float)((da - 1.0 == 0.0 ? 0.0 : log(x)) + psi0 - psia);
      dlp_b[i] = (float)((dbt - 1.0 == 0.0 ? 0.0 : log(y)) + psi0 - psib);
      const double t0 = -(2.0 - a0) * ac_trigamma(a0);
      dH_a[i] = (float)(t0 - (da - 1.0) * ac_trigamma(da));
      dH_b[i] = (float)(t0 - (dbt - 1.0) * ac_trigamma(dbt));
      sg_a[i] = z[i] > 20.f ? 1.f : ac_sigmoid(z[i]);
      sg_b[i] = z[A + i] > 20.f ? 1.f : ac_sigmoid(z[A + i]);
    }
    const float logratio = (float)logp - sc.x, ratio = expf(logratio);
    float adv = sc.y;
    if (a.hp.norm_adv) adv = (adv - a.adv_moments[0]) * a.adv_moments[1];
    const float c = a.hp.clip_coef;
    const float pg1 = -adv * ratio, pg2 = -adv * fminf(fmaxf(ratio, 1.f - c), 1.f + c);
    float dpg;
    const bool inside = ratio >= 1.f - c && ratio <= 1.f + c;
    if (pg1 > pg2) dpg = -adv; else if (pg1 == pg2) dpg = inside ? -adv : -0.5f * adv; else dpg = inside ? -adv : 0.f;
    const float g_logp = (float)invB * dpg * ratio, g_ent = -a.hp.ent_coef * (float)invB;
    const float v = a.v[m], R = sc.z, vold = sc.w;
    float dv, vu = (v - R) * (v - R), vloss = vu;
    if (a.hp.clip_vloss) {
      float dvo = v - vold, vc = vold + fminf(fmaxf(dvo, -c), c), vcl = (vc - R) * (vc - R);
      const bool in = dvo >= -c && dvo <= c;
      float gu = 2.f * (v - R), gc = in ? 2.f * (vc - R) : 0.f;
      if (vu > vcl) dv = gu; else if (vu == vcl) dv = 0.5f * (gu + gc); else dv = gc;
      vloss = fmaxf(vu, vcl);
    } else dv = 2.f * (v - R);
    dv *= 0.5f * a.hp.vf_coef * (float)invB;
    float *dz = a.dzab + (size_t)m * 64;
    for (int i = 0; i < A; i++) {
      dz[i] = (g_logp * dlp_a[i] + g_ent * dH_a[i]) * sg_a[i];
      dz[A + i] = (g_logp * dlp_b[i] + g_ent * dH_b[i]) * sg_b[i];
    }
    a.dvw[(size_t)m * 64] = dv;
    stat[0] = fmaxf(pg1, pg2); stat[1] = 0.5f * vloss; stat[2] = (float)ent; stat[3] = -logratio;
    stat[4] = (ratio - 1.f) - logratio; stat[5] = fabsf(ratio - 1.f) > c ? 1.f : 0.f;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < AC_NSTAT; i++) {
    float v = stat[i];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < AC_NSTAT) a.part[(size_t)blockIdx.x * AC_NSTAT + threadIdx.x] = red[0][threadIdx.x] + red[1][threadIdx.x] + red[2][threadIdx.x] + red[3][threadIdx.x];
}

// ------------------------------------------------------------------ rollout: sample Beta actions, log-prob, store the row
// Gamma(shape >= 1) by Marsaglia-Tsang from Philox normals/uniforms; Beta = X / (X + Y) like at::_sample_dirichlet followed
// by select(-1, 0) (rl_utils.h:95-98).  Distribution-exact, not stream-exact with torch's generator (documented deviation).
struct AcSampleArgs {
  const float *zab; int ldz; const float *v; int N, A;
  float act_low, act_high;
  const float *obs_in, *term_in, *trunc_in; int O;
  float *obs_row, *act_row, *logp_row, *val_row, *done_row;
  uint32_t seed_lo, seed_hi, step_index;
  int mode;   // 0: sample, 1: mean (final evaluation, :229-231,981)
};

__device__ inline float ac_gamma_sample(float shape, uint32_t env, uint32_t step, uint32_t slot, uint32_t k0, uint32_t k1) {
  const float d = shape - 1.f / 3.f, c = rsqrtf(9.f * d);
  for (uint32_t it = 0; it < 64; it++) {
    uint32_t ctr[4] = {env, step, slot, 0xBE7Au + it};
    philox4x32_10(ctr, k0, k1);
    float z0, z1;
    box_muller(ctr[0], ctr[1], &z0, &z1);
    const float u = ((float)ctr[2] + 1.0f) * 2.3283064365386963e-10f;
    const float t = 1.f + c * z0;
    if (t <= 0.f) continue;
    const float vv = t * t * t;
    if (logf(u) < 0.5f * z0 * z0 + d - d * vv + d * logf(vv)) return d * vv;
  }
  return d;
}

// One thread per (env, action dimension): the two gamma draws and the three fp64 lgamma of a dimension are independent of the other
// dimensions, and a thread per env left 64 CTAs of 4 warps for the whole GPU (21 us per rollout step).  The per-dimension log-prob
// terms go through shared memory and are added by the env's first thread in dimension order, so the sum is the same as before bit
// for bit.  blockDim = 256 = AC_SAMPLE_EPB(A) envs x dims-per-env slot (8 for A <= 8, else 32).
#define AC_SAMPLE_SLOT(A) ((A) <= 8 ? 8 : 32)
__global__ void __launch_bounds__(256) ac_beta_sample_kernel(AcSampleArgs a) {
  __shared__ double terms[256];
  const int A = a.A, slot = AC_SAMPLE_SLOT(A), epb = 256 / slot;
  const int le = threadIdx.x / slot, i = threadIdx.x - le * slot, env = blockIdx.x * epb + le;
  const bool live = env < a.N && i < A;
  double term = 0.0;
  if (live) {
    const float *z = a.zab + (size_t)env * a.ldz;
    const float al = ac_softplus(z[i]) + 1.0f, be = ac_softplus(z[A + i]) + 1.0f;
    float s;
    if (a.mode == 0) {
      const float x = ac_gamma_sample(al, env, a.step_index, 2 * i, a.seed_lo, a.seed_hi);
      const float y = ac_gamma_sample(be, env, a.step_index, 2 * i + 1, a.seed_lo, a.seed_hi);
      s = x / (x + y);
      s = fminf(fmaxf(s, 1.17549435e-38f), 1.f - 1.1920929e-7f);  // _sample_dirichlet clamps to [tiny, 1 - eps]
    } else {
      s = al / (al + be);
    }
    const double da = al, dbt = be, a0 = (double)(al + be);
    term = ac_xlogy(da - 1.0, (double)s) + ac_xlogy(dbt - 1.0, (double)(1.0f - s)) + lgamma(a0) - lgamma(da) - lgamma(dbt);
    if (a.act_row) a.act_row[(size_t)env * A + i] = s * (a.act_high - a.act_low) + a.act_low;   // unscale_action (:205-210)
  }
  terms[threadIdx.x] = term;
  __syncthreads();
  if (i == 0 && env < a.N) {
    double logp = 0.0;
    for (int k = 0; k < A; k++) logp += terms[le * slot + k];
    if (a.logp_row) a.logp_row[env] = (float)logp;
    if (a.val_row) a.val_row[env] = a.v[env];
    if (a.done_row) a.done_row[env] = (a.term_in[env] != 0.f || a.trunc_in[env] != 0.f) ? 1.f : 0.f;
  }
}
