// pb_ppo_gemm.cuh — fp32 tiled GEMM used by the PPO minibatch forward/backward (src/ppo_continuous_action.cpp:496-538).
// One kernel, three operand layouts, fused epilogues; fp32 accumulate on the FMA pipe so losses/gradients match the
// reference's fp32 CPU arithmetic to rounding (the tensor-core variant is a later step).
//   MODE_FWD : C[m][n] = tanh(bias[n] + sum_k A[row(m)][k] * W[n][k])          (x @ W^T, rows optionally gathered)
//   MODE_BWD : C[m][n] = (sum_k D[m][k] * W[k][n]) * (1 - H[m][n]^2)           (delta @ W, through tanh)
//   MODE_WGT : C_z[m][n] = sum_{k in chunk z} D[k][m] * X[row(k)][n]           (delta^T @ x, split-K into slabs)
// blockIdx.z selects the network (0 critic, 1 actor) in FWD/BWD, and (net, split) in WGT.
#pragma once
#include <cuda_runtime.h>

enum { MODE_FWD = 0, MODE_BWD = 1, MODE_WGT = 2 };

struct GemmArgs {
  int M, N, K;
  const float *A[2]; int lda;       // FWD: activations; BWD/WGT: deltas
  const float *B[2]; int ldb;       // FWD/BWD: weights; WGT: activations
  float *C[2]; int ldc;
  const float *bias[2];             // FWD
  const float *H[2]; int ldh;       // BWD: tanh outputs for the derivative
  const int *rowidx;                // gather index for A rows (FWD) / X rows (WGT), NULL = identity
  const int *cursor; int mb;        // minibatch cursor: rowidx += (*cursor) * mb
  int gatherB;                      // WGT: apply rowidx to B (X) rows
  int splits; size_t slab_stride;   // WGT
  int nets;
};

template <int MODE>
__global__ void __launch_bounds__(256) gemm_kernel(GemmArgs g) {
  constexpr int BM = 64, BN = 64, BK = 16;
  __shared__ float As[BK][BM + 4], Bs[BK][BN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  int net = 0, split = 0;
  if (MODE == MODE_WGT) { net = blockIdx.z / g.splits; split = blockIdx.z - net * g.splits; } else net = blockIdx.z;
  const float *__restrict__ A = g.A[net], *__restrict__ B = g.B[net];
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int *idx = g.rowidx ? g.rowidx + (g.cursor ? (size_t)(*g.cursor) * g.mb : 0) : nullptr;
  int k_begin = 0, k_end = g.K;
  if (MODE == MODE_WGT) {
    int chunk = (g.K + g.splits - 1) / g.splits;
    k_begin = split * chunk; k_end = min(g.K, k_begin + chunk);
  }
  float acc[4][4] = {};
  for (int k0 = k_begin; k0 < k_end; k0 += BK) {
    // ---- stage tiles (each thread 4 elements of each)
#pragma unroll
    for (int i = 0; i < 4; i++) {
      if (MODE == MODE_WGT) {
        // A' = D^T: element (m, k) = D[k][m]  -> threads run along m (contiguous)
        int e = tid + 256 * i, kk = e / BM, mm = e % BM, k = k0 + kk, m = m0 + mm;
        As[kk][mm] = (k < k_end && m < g.M) ? A[(size_t)k * g.lda + m] : 0.f;
        // B' = X: element (k, n) = X[row(k)][n] -> threads run along n
        int kb = e / BN, nn = e % BN, k2 = k0 + kb, n = n0 + nn;
        float v = 0.f;
        if (k2 < k_end && n < g.N) { int r = (g.gatherB && idx) ? idx[k2] : k2; v = B[(size_t)r * g.ldb + n]; }
        Bs[kb][nn] = v;
      } else {
        // A element (m, k) = A[row(m)][k] -> threads run along k
        int e = tid + 256 * i, mm = e / BK, kk = e % BK, m = m0 + mm, k = k0 + kk;
        float v = 0.f;
        if (m < g.M && k < g.K) { int r = (MODE == MODE_FWD && idx) ? idx[m] : m; v = A[(size_t)r * g.lda + k]; }
        As[kk][mm] = v;
        if (MODE == MODE_FWD) {  // B' (k, n) = W[n][k] -> threads along k
          int nn = e / BK, kb = e % BK, n = n0 + nn, k2 = k0 + kb;
          Bs[kb][nn] = (n < g.N && k2 < g.K) ? B[(size_t)n * g.ldb + k2] : 0.f;
        } else {                 // B' (k, n) = W[k][n] -> threads along n
          int kb = e / BN, nn = e % BN, k2 = k0 + kb, n = n0 + nn;
          Bs[kb][nn] = (k2 < g.K && n < g.N) ? B[(size_t)k2 * g.ldb + n] : 0.f;
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; kk++) {
      float4 a = *reinterpret_cast<const float4 *>(&As[kk][ty * 4]);
      float4 b = *reinterpret_cast<const float4 *>(&Bs[kk][tx * 4]);
      float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] += av[i] * bv[j];
    }
    __syncthreads();
  }
  float *C = g.C[net] + (MODE == MODE_WGT ? (size_t)split * g.slab_stride : 0);
#pragma unroll
  for (int i = 0; i < 4; i++) {
    int m = m0 + ty * 4 + i;
    if (m >= g.M) continue;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      int n = n0 + tx * 4 + j;
      if (n >= g.N) continue;
      float v = acc[i][j];
      if (MODE == MODE_FWD) v = tanhf(v + g.bias[net][n]);
      if (MODE == MODE_BWD) { float h = g.H[net][(size_t)m * g.ldh + n]; v *= (1.f - h * h); }
      C[(size_t)m * g.ldc + n] = v;
    }
  }
}
