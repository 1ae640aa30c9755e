// pb_ppo_fused.cuh — the PPO minibatch step as four launches (src/ppo_continuous_action.cpp:489-541):
//   epoch prologue : gather the epoch's permuted rows into contiguous buffers + per-minibatch advantage moments
//   fwd_bwd_kernel : per 64-sample tile: both MLPs forward, heads, clipped loss, and every activation delta
//   wgrad_kernel   : all weight / bias gradients as split-K "A^T B" tiles into per-split slabs (fixed summation order)
//   grad_reduce    : slabs -> flat gradient (+ sum of squares partials)      adam2_kernel : clip + Adam (+ transposed copies)
// fp32 throughout (FFMA pipe) so gradients match the reference's fp32 CPU arithmetic to rounding.
#pragma once
#include "pb_ppo_common.cuh"
#include "pb_ppo_update.cuh"

#define FB_THREADS 256
#define FB_BK 16
#define FB_LDH 132      // activations [m][n] row stride (16 B aligned rows, conflict-free float4 row stores)
#define FB_LDW3 68
#define FB_MAXA 19      // A + 1 <= 20

// ------------------------------------------------------------------ epoch prologue
struct GatherArgs {
  const int *idx;                 // [rows] permuted sample ids of this epoch (or one explicit minibatch)
  int rows, O, A;
  const float *obs, *actions, *logprobs, *advantages, *returns, *values;  // flat rollout buffers [T*N][.]
  float *Xp, *actp;               // [rows][O], [rows][A]
  float4 *scp;                    // [rows] (logp_old, adv, ret, val)
};

// one warp per row: coalesced 4-byte copies of the observation row, lanes < A copy the action
static __global__ void __launch_bounds__(256) epoch_gather_kernel(GatherArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r = blockIdx.x * 8 + warp;
  if (r >= a.rows) return;
  const int s = a.idx[r];
  const float *src = a.obs + (size_t)s * a.O;
  float *dst = a.Xp + (size_t)r * a.O;
  if ((a.O & 3) == 0) {
    const float4 *s4 = reinterpret_cast<const float4 *>(src);
    float4 *d4 = reinterpret_cast<float4 *>(dst);
    for (int i = lane; i < a.O / 4; i += 32) d4[i] = s4[i];
  } else {
    for (int i = lane; i < a.O; i += 32) dst[i] = src[i];
  }
  if (lane < a.A) a.actp[(size_t)r * a.A + lane] = a.actions[(size_t)s * a.A + lane];
  if (lane == 0) a.scp[r] = make_float4(a.logprobs[s], a.advantages[s], a.returns[s], a.values[s]);
}

// mean and 1/(unbiased std + 1e-8) of the advantages of every minibatch of the epoch (block = minibatch; fp64 sums)
static __global__ void __launch_bounds__(256) adv_stats_all_kernel(const float4 *__restrict__ scp, int mb, float *__restrict__ out) {
  __shared__ double s1[256], s2[256];
  const float4 *p = scp + (size_t)blockIdx.x * mb;
  double a = 0, b = 0;
  for (int i = threadIdx.x; i < mb; i += 256) { double v = p[i].y; a += v; b += v * v; }
  s1[threadIdx.x] = a; s2[threadIdx.x] = b;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { s1[threadIdx.x] += s1[threadIdx.x + o]; s2[threadIdx.x] += s2[threadIdx.x + o]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double mean = s1[0] / mb, var = (s2[0] - mb * mean * mean) / (mb - 1);
    out[2 * blockIdx.x] = (float)mean;
    out[2 * blockIdx.x + 1] = (float)(1.0 / (sqrt(var > 0 ? var : 0) + 1e-8));
  }
}

// ------------------------------------------------------------------ forward + loss + activation deltas, one 64-sample tile per CTA
struct FwdBwdArgs {
  const float *params, *W1T, *W2T;   // flat torch-layout parameters; W1T[k][128] (critic | actor), W2T[net][k][u]
  AgentLayout L; PpoHyper hp;
  const float *Xp, *actp; const float4 *scp;   // epoch-ordered gathered rows
  const float *adv_stats;            // [minibatch][2] mean, 1/(std+1e-8)
  const int *cursor;                 // minibatch index inside the epoch buffers
  int mb;
  float *H1, *H2, *D1, *D2;          // [mb][128] activations and deltas of the two hidden layers (critic | actor)
  float *D3;                         // [mb][64]   dL/dmean..., dL/dvalue, zero padding (never written)
  float *part;                       // [tiles][PB_NSTAT + A] per-tile sums (loss statistics, dL/dlogstd)
  int debug;                         // 1: block 0 prints the cycle count of each phase (tuning aid)
  const float *Hpre, *Hpre2;         // optional [mb][128] (+ second partial or null): first-layer products X W1^T from l1_forward_tc_kernel
};
#define FB_TICK(i) do { if (a.debug && blockIdx.x == 0 && threadIdx.x == 0) tick[i] = clock64(); } while (0)

// TM samples per CTA tile (32 or 64); As[kk][m] row stride TM + 4; TM / 16 rows per thread
#define FB_SMEM_FLOATS(TM) (2 * FB_BK * ((TM) + 4) + 2 * FB_BK * 128 + 2 * (TM) * FB_LDH + 2 * 64 * 64 + 20 * FB_LDW3 + 2 * (TM) * 20 + 320)
// tensor-core variant of the first layer (TM = 64): hi / lo TF32 halves of the X chunk [k][m] (stride 72) and of the W1T chunk
// [k][n] (stride 136); the strides are 8 mod 32 so the m16n8k8 fragment loads (4 k x 8 m|n per instruction) hit 32 banks
#define FB_LDA_M 20      // X chunk [m][k], 16 k per row + pad: (g*20 + t) mod 32 is a permutation for g < 8, t < 4
#define FB_LDB_M 136     // W1T chunk [k][n]: (t*136 + g) mod 32 = 8t + g
#define FB_NST 6         // cp.async stages in flight (the X rows come from HBM: ~1.5-2k cycles of latency to cover)
#define FB_SMEM_FLOATS_MMA (FB_NST * 64 * FB_LDA_M + FB_NST * FB_BK * FB_LDB_M + 2 * 64 * FB_LDH + 2 * 64 * 64 + 20 * FB_LDW3 + 2 * 64 * 20 + 320)

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem, bool valid) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  const int n = valid ? 16 : 0;  // src-size 0: the 16 destination bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sa), "l"(gmem), "r"(n));
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem, bool valid) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  const int n = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(sa), "l"(gmem), "r"(n));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }

// ---- 3xTF32: x = hi + lo with both halves TF32; a*b ~ a_lo*b_hi + a_hi*b_lo + a_hi*b_hi keeps fp32-level accuracy on the
// tensor cores (the dropped a_lo*b_lo term is 2^-22 relative), accumulation in fp32.
__device__ __forceinline__ void tf32_split(float x, float &hi, float &lo) {
  unsigned h, l;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  const float r = x - __uint_as_float(h);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(r));
  hi = __uint_as_float(h); lo = __uint_as_float(l);
}
// D(16x8) += A(16x8, row) * B(8x8, col); lane = 4*g + t: a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4); b0 (k=t, n=g)
// b1 (k=t+4, n=g); c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1)
__device__ __forceinline__ void mma_tf32(float (&c)[4], const float (&a)[4], float b0, float b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
                 "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}
__device__ __forceinline__ void mma_3xtf32(float (&c)[4], const float (&ah)[4], const float (&al)[4], float bh0, float bh1, float bl0,
                                           float bl1) {
#ifndef PB_TF32_1X
  mma_tf32(c, al, bh0, bh1);
  mma_tf32(c, ah, bl0, bl1);
#endif
  mma_tf32(c, ah, bh0, bh1);
}

// acc[i][jh*4+j] += sum_k act[ty*4+i][jh*64+k] * W[jh][k][tx*4+j]   (two independent 64x64x64 products, one per net)
template <int FB_RPT>
__device__ __forceinline__ void fb_layer64(float (&acc)[FB_RPT][8], const float *act, const float *W, int tx, int ty) {
#pragma unroll
  for (int jh = 0; jh < 2; jh++) {
    const float *ap = act + ty * FB_RPT * FB_LDH + jh * 64;
    const float *wp = W + jh * 4096 + tx * 4;
#pragma unroll 2
    for (int k4 = 0; k4 < 64; k4 += 4) {
      float av[FB_RPT][4];
#pragma unroll
      for (int i = 0; i < FB_RPT; i++) {
        float4 a = *reinterpret_cast<const float4 *>(ap + i * FB_LDH + k4);
        av[i][0] = a.x; av[i][1] = a.y; av[i][2] = a.z; av[i][3] = a.w;
      }
#pragma unroll
      for (int kk = 0; kk < 4; kk++) {
        float4 b = *reinterpret_cast<const float4 *>(wp + (k4 + kk) * 64);
#pragma unroll
        for (int i = 0; i < FB_RPT; i++) {
          acc[i][jh * 4 + 0] += av[i][kk] * b.x; acc[i][jh * 4 + 1] += av[i][kk] * b.y;
          acc[i][jh * 4 + 2] += av[i][kk] * b.z; acc[i][jh * 4 + 3] += av[i][kk] * b.w;
        }
      }
    }
  }
}

template <int FB_TM, bool MMA = false>
static __global__ void __launch_bounds__(FB_THREADS, FB_TM == 32 ? 2 : 1) fwd_bwd_kernel(FwdBwdArgs a) {
  static_assert(!MMA || FB_TM == 64, "the tensor-core first layer is written for 64-sample tiles");
  constexpr int FB_LDA = MMA ? FB_LDA_M : FB_TM + 4, FB_RPT = FB_TM / 16;
  extern __shared__ __align__(16) float sm[];
  float *As = sm;                               // [2][BK][LDA]   X chunk, transposed      (MMA: [NST][64][20] rows of X)
  float *Bs = As + (MMA ? FB_NST * 64 * FB_LDA_M : 2 * FB_BK * FB_LDA);   // [2][BK][128] W1T chunk  (MMA: [NST][BK][136])
  float *H1s = Bs + (MMA ? FB_NST * FB_BK * FB_LDB_M : 2 * FB_BK * 128);  // [64][LDH]
  float *H2s = H1s + FB_TM * FB_LDH;            // [64][LDH]      h2, later the delta D2 in place
  float *W2s = H2s + FB_TM * FB_LDH;            // [2][64][64]    W2T for the forward, W2 for the backward
  float *W3s = W2s + 2 * 64 * 64;               // [A+1][LDW3]    actor head rows, then the critic head row
  float *outs = W3s + 20 * FB_LDW3;             // [64][20]       mean..., value
  float *d3s = outs + FB_TM * 20;               // [64][20]
  float *red = d3s + FB_TM * 20;                // [8 warps][PB_NSTAT + A]
  float *lsS = red + 8 * 32, *ivS = lsS + 32;   // logstd and 1 / var per action dimension
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int O = a.L.obs, A = a.L.act, NP = PB_NSTAT + A;
  const int cur = *a.cursor;
  const size_t row0 = (size_t)cur * a.mb + (size_t)blockIdx.x * FB_TM;   // row in the epoch buffers
  const size_t out0 = (size_t)blockIdx.x * FB_TM;                        // row in the per-minibatch scratch
  const int rows = min(FB_TM, a.mb - blockIdx.x * FB_TM);
  const bool aligned = (O & 3) == 0;
  const float *P = a.params;
  long long tick[8];
  FB_TICK(0);

  // ---- small weights into shared memory
  if (tid < a.L.act) { const float ls = P[a.L.logstd + tid], sd = expf(ls); lsS[tid] = ls; ivS[tid] = 1.f / (sd * sd); }
  for (int e = tid; e < 2048; e += FB_THREADS) reinterpret_cast<float4 *>(W2s)[e] = reinterpret_cast<const float4 *>(a.W2T)[e];
  for (int e = tid; e < (A + 1) * 64; e += FB_THREADS) {
    int o = e >> 6, k = e & 63;
    W3s[o * FB_LDW3 + k] = o < A ? P[a.L.aW3 + o * 64 + k] : P[a.L.cW3 + k];
  }

  // ---- layer 1: acc[64 x 128] = X[64 x O] W1T[O x 128], register-staged double buffering over k
  float acc[FB_RPT][8];
  float4 xr, wr[2];
  const int xm = tid >> 2, xk = (tid & 3) * 4;
  auto gload = [&](int c) {
    const int k0 = c * FB_BK, k = k0 + xk;
    xr = make_float4(0.f, 0.f, 0.f, 0.f);
    if (xm < rows && xm < FB_TM) {
      const float *p = a.Xp + (row0 + xm) * O + k;
      if (aligned) { if (k < O) xr = *reinterpret_cast<const float4 *>(p); }
      else { if (k < O) xr.x = p[0]; if (k + 1 < O) xr.y = p[1]; if (k + 2 < O) xr.z = p[2]; if (k + 3 < O) xr.w = p[3]; }
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
      int e = tid + FB_THREADS * i, kk = e >> 5, n4 = e & 31, kg = k0 + kk;
      wr[i] = kg < O ? *reinterpret_cast<const float4 *>(a.W1T + (size_t)kg * 128 + n4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  const int nchunk = (O + FB_BK - 1) / FB_BK;
  if constexpr (MMA) {
    // 8 warps = 4 (16-row blocks) x 2 (nets): each warp owns a 16 x 64 block = 8 m16n8 accumulators
    const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int mrow = (warp & 3) * 16, ncol0 = (warp >> 2) * 64;
    float c1[8][4];
#pragma unroll
    for (int j = 0; j < 8; j++) c1[j][0] = c1[j][1] = c1[j][2] = c1[j][3] = 0.f;
    // cp.async pipeline, FB_NST chunks of 16 k in flight: thread -> (row xm, 4 k) of X and two (k, 4 n) pieces of W1T
    auto issue = [&](int c) {
      if (c < nchunk) {
        const int st = c % FB_NST, k0 = c * FB_BK, k = k0 + xk;
        float *dx = As + (st * 64 + xm) * FB_LDA_M + xk;
        const float *px = a.Xp + (row0 + xm) * O + k;
        const bool rv = xm < rows;
        if (aligned) cp_async16(dx, rv && k < O ? px : a.Xp, rv && k < O);
        else {
#pragma unroll
          for (int i = 0; i < 4; i++) cp_async4(dx + i, rv && k + i < O ? px + i : a.Xp, rv && k + i < O);
        }
#pragma unroll
        for (int i = 0; i < 2; i++) {
          const int e = tid + FB_THREADS * i, kk = e >> 5, n4 = e & 31, kg = k0 + kk;
          cp_async16(Bs + (st * FB_BK + kk) * FB_LDB_M + n4 * 4, a.W1T + (size_t)(kg < O ? kg : 0) * 128 + n4 * 4, kg < O);
        }
      }
      cp_async_commit();
    };
#pragma unroll
    for (int c = 0; c < FB_NST - 1; c++) issue(c);
    for (int c = 0; c < nchunk; c++) {
      cp_async_wait<FB_NST - 2>();
      __syncthreads();
      issue(c + FB_NST - 1);
      const float *Ax = As + (c % FB_NST) * 64 * FB_LDA_M + (mrow + g) * FB_LDA_M + t;
      const float *Bx = Bs + (c % FB_NST) * FB_BK * FB_LDB_M + t * FB_LDB_M + ncol0 + g;
#pragma unroll
      for (int ks = 0; ks < FB_BK / 8; ks++) {
        const float ax[4] = {Ax[ks * 8], Ax[8 * FB_LDA_M + ks * 8], Ax[ks * 8 + 4], Ax[8 * FB_LDA_M + ks * 8 + 4]};
        float ah[4], al[4];
#pragma unroll
        for (int i = 0; i < 4; i++) tf32_split(ax[i], ah[i], al[i]);
#pragma unroll
        for (int j = 0; j < 8; j++) {
          float bh0, bl0, bh1, bl1;
          tf32_split(Bx[(ks * 8) * FB_LDB_M + 8 * j], bh0, bl0);
          tf32_split(Bx[(ks * 8 + 4) * FB_LDB_M + 8 * j], bh1, bl1);
          mma_3xtf32(c1[j], ah, al, bh0, bh1, bl0, bl1);
        }
      }
    }
    cp_async_wait<0>();
    // h1 = tanh(acc + b1) -> shared [m][n] and global
    const float *bp = P + (ncol0 ? a.L.ab1 : a.L.cb1);
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int nl = 8 * j + 2 * t, col = ncol0 + nl;
      const float b0 = bp[nl], b1 = bp[nl + 1];
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int m = mrow + g + 8 * h;
        const float2 v = make_float2(tanhf(c1[j][2 * h] + b0), tanhf(c1[j][2 * h + 1] + b1));
        *reinterpret_cast<float2 *>(H1s + m * FB_LDH + col) = v;
        if (m < rows) *reinterpret_cast<float2 *>(a.H1 + (out0 + m) * 128 + col) = v;
      }
    }
    __syncthreads();
  } else {
#pragma unroll
  for (int i = 0; i < FB_RPT; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
  auto sstore = [&](int buf) {
    float *A_ = As + buf * FB_BK * FB_LDA;
    if (xm < FB_TM) {
      A_[(xk + 0) * FB_LDA + xm] = xr.x; A_[(xk + 1) * FB_LDA + xm] = xr.y;
      A_[(xk + 2) * FB_LDA + xm] = xr.z; A_[(xk + 3) * FB_LDA + xm] = xr.w;
    }
    float *B_ = Bs + buf * FB_BK * 128;
#pragma unroll
    for (int i = 0; i < 2; i++) {
      int e = tid + FB_THREADS * i, kk = e >> 5, n4 = e & 31;
      *reinterpret_cast<float4 *>(B_ + kk * 128 + n4 * 4) = wr[i];
    }
  };
  if (a.Hpre) {  // tensor-core path: pre-activations of this tile come from global memory
#pragma unroll
    for (int i = 0; i < FB_RPT; i++) {
      const int m = ty * FB_RPT + i;
#pragma unroll
      for (int jh = 0; jh < 2; jh++) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m < rows) {
          v = *reinterpret_cast<const float4 *>(a.Hpre + (out0 + m) * 128 + jh * 64 + tx * 4);
          if (a.Hpre2) {
            const float4 w = *reinterpret_cast<const float4 *>(a.Hpre2 + (out0 + m) * 128 + jh * 64 + tx * 4);
            v.x += w.x; v.y += w.y; v.z += w.z; v.w += w.w;
          }
        }
        acc[i][jh * 4] = v.x; acc[i][jh * 4 + 1] = v.y; acc[i][jh * 4 + 2] = v.z; acc[i][jh * 4 + 3] = v.w;
      }
    }
  } else {
  gload(0); sstore(0);
  __syncthreads();
  for (int c = 0; c < nchunk; c++) {
    if (c + 1 < nchunk) gload(c + 1);
    const float *A_ = As + (c & 1) * FB_BK * FB_LDA + ty * FB_RPT, *B_ = Bs + (c & 1) * FB_BK * 128 + tx * 4;
#pragma unroll
    for (int kk = 0; kk < FB_BK; kk++) {
      float ar[FB_RPT];
      if constexpr (FB_RPT == 4) {
        float4 av = *reinterpret_cast<const float4 *>(A_ + kk * FB_LDA);
        ar[0] = av.x; ar[1] = av.y; ar[2] = av.z; ar[3] = av.w;
      } else {
        float2 av = *reinterpret_cast<const float2 *>(A_ + kk * FB_LDA);
        ar[0] = av.x; ar[1] = av.y;
      }
      float4 b0 = *reinterpret_cast<const float4 *>(B_ + kk * 128), b1 = *reinterpret_cast<const float4 *>(B_ + kk * 128 + 64);
#pragma unroll
      for (int i = 0; i < FB_RPT; i++) {
        acc[i][0] += ar[i] * b0.x; acc[i][1] += ar[i] * b0.y; acc[i][2] += ar[i] * b0.z; acc[i][3] += ar[i] * b0.w;
        acc[i][4] += ar[i] * b1.x; acc[i][5] += ar[i] * b1.y; acc[i][6] += ar[i] * b1.z; acc[i][7] += ar[i] * b1.w;
      }
    }
    if (c + 1 < nchunk) sstore((c + 1) & 1);
    __syncthreads();
  }
  }
  // h1 = tanh(acc + b1) -> shared [m][n] and global (the weight-gradient kernel reads it)
#pragma unroll
  for (int jh = 0; jh < 2; jh++) {
    const float *bp = P + (jh ? a.L.ab1 : a.L.cb1) + tx * 4;
    const float b0 = bp[0], b1 = bp[1], b2 = bp[2], b3 = bp[3];
#pragma unroll
    for (int i = 0; i < FB_RPT; i++) {
      const int m = ty * FB_RPT + i;
      float4 h = make_float4(tanhf(acc[i][jh * 4] + b0), tanhf(acc[i][jh * 4 + 1] + b1), tanhf(acc[i][jh * 4 + 2] + b2),
                             tanhf(acc[i][jh * 4 + 3] + b3));
      *reinterpret_cast<float4 *>(H1s + m * FB_LDH + jh * 64 + tx * 4) = h;
      if (m < rows) *reinterpret_cast<float4 *>(a.H1 + (out0 + m) * 128 + jh * 64 + tx * 4) = h;
    }
  }
  __syncthreads();
  }
  FB_TICK(1);
  // ---- layer 2
#pragma unroll
  for (int i = 0; i < FB_RPT; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
  fb_layer64<FB_RPT>(acc, H1s, W2s, tx, ty);
#pragma unroll
  for (int jh = 0; jh < 2; jh++) {
    const float *bp = P + (jh ? a.L.ab2 : a.L.cb2) + tx * 4;
    const float b0 = bp[0], b1 = bp[1], b2 = bp[2], b3 = bp[3];
#pragma unroll
    for (int i = 0; i < FB_RPT; i++) {
      const int m = ty * FB_RPT + i;
      float4 h = make_float4(tanhf(acc[i][jh * 4] + b0), tanhf(acc[i][jh * 4 + 1] + b1), tanhf(acc[i][jh * 4 + 2] + b2),
                             tanhf(acc[i][jh * 4 + 3] + b3));
      *reinterpret_cast<float4 *>(H2s + m * FB_LDH + jh * 64 + tx * 4) = h;
      if (m < rows) *reinterpret_cast<float4 *>(a.H2 + (out0 + m) * 128 + jh * 64 + tx * 4) = h;
    }
  }
  __syncthreads();
  FB_TICK(2);
  // ---- heads (thread = sample tid/4, outputs tid%4 + 4i); the second-layer weights are swapped for the untransposed
  // copy the backward product needs while this runs
  for (int e = tid; e < 8192; e += FB_THREADS) W2s[e] = e < 4096 ? P[a.L.cW2 + e] : P[a.L.aW2 + e - 4096];
  {
    constexpr int TPR = FB_THREADS / FB_TM;  // threads per sample
    const int m = tid / TPR;
    for (int o = tid % TPR; o <= A; o += TPR) {
      const float *h = H2s + m * FB_LDH + (o < A ? 64 : 0), *w = W3s + o * FB_LDW3;
      float s = o < A ? P[a.L.ab3 + o] : P[a.L.cb3];
#pragma unroll 4
      for (int k = 0; k < 64; k += 4) {
        float4 hv = *reinterpret_cast<const float4 *>(h + k), wv = *reinterpret_cast<const float4 *>(w + k);
        s += wv.x * hv.x; s += wv.y * hv.y; s += wv.z * hv.z; s += wv.w * hv.w;
      }
      outs[m * 20 + o] = s;
    }
  }
  __syncthreads();
  FB_TICK(3);
  // ---- clipped PPO loss and its gradient at the heads (same math as ppo_loss_kernel): TPS threads per sample split the
  // action dimensions, the scalar part is computed redundantly by the TPS threads of a sample
  {
    constexpr int TPS = FB_THREADS / FB_TM, DPT = (FB_MAXA + TPS) / TPS;
    const int m = tid / TPS, part = tid % TPS, lane = tid & 31, warp = tid >> 5;
    const bool valid = m < rows;
    float stat[PB_NSTAT] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float dl[DPT], dmean[DPT];
#pragma unroll
    for (int j = 0; j < DPT; j++) { dl[j] = 0.f; dmean[j] = 0.f; }
    float *d3 = d3s + m * 20;
    const float *o = outs + m * 20;
    const float invB = 1.f / (float)a.mb;
    float logp = 0.f, ent = 0.f;
    if (valid) {
      const float *act = a.actp + (row0 + m) * A;
#pragma unroll
      for (int j = 0; j < DPT; j++) {
        const int i = part + TPS * j;
        if (i < A) {
          const float ls = lsS[i], iv = ivS[i], d = act[i] - o[i], q = d * d * iv;
          logp += -0.5f * q - ls - 0.91893853320467274178f;
          ent += 0.5f + 0.91893853320467274178f + ls;
          dmean[j] = d * iv;   // dlogp/dmean
          dl[j] = q - 1.f;     // dlogp/dlogstd
        }
      }
    }
#pragma unroll
    for (int sft = TPS / 2; sft > 0; sft >>= 1) { logp += __shfl_xor_sync(0xffffffffu, logp, sft); ent += __shfl_xor_sync(0xffffffffu, ent, sft); }
    if (valid) {
      const float4 sc = a.scp[row0 + m];   // logp_old, adv, ret, val
      const float v = o[A];
      float logratio = logp - sc.x, ratio = expf(logratio);
      float adv = sc.y;
      if (a.hp.norm_adv) adv = (adv - a.adv_stats[2 * cur]) * a.adv_stats[2 * cur + 1];
      const float c = a.hp.clip_coef;
      float pg1 = -adv * ratio, pg2 = -adv * fminf(fmaxf(ratio, 1.f - c), 1.f + c);
      // d max(pg1,pg2)/d ratio with torch's tie rule (halves) folded: inside the clip range both branches carry -adv
      float dpg;
      const bool inside = ratio >= 1.f - c && ratio <= 1.f + c;
      if (pg1 > pg2) dpg = -adv; else if (pg1 == pg2) dpg = inside ? -adv : -0.5f * adv; else dpg = inside ? -adv : 0.f;
      const float g_logp = invB * dpg * ratio;
      float *g3 = a.D3 + (out0 + m) * 64;
      if (part == 0) {
        const float R = sc.z, vold = sc.w;
        float dv, vu = (v - R) * (v - R), vloss = vu;
        if (a.hp.clip_vloss) {
          float dvo = v - vold, vc = vold + fminf(fmaxf(dvo, -c), c), vcl = (vc - R) * (vc - R);
          const bool in = dvo >= -c && dvo <= c;
          float gu = 2.f * (v - R), gc = in ? 2.f * (vc - R) : 0.f;
          if (vu > vcl) dv = gu; else if (vu == vcl) dv = 0.5f * (gu + gc); else dv = gc;
          vloss = fmaxf(vu, vcl);
        } else dv = 2.f * (v - R);
        dv *= 0.5f * a.hp.vf_coef * invB;
        stat[0] = fmaxf(pg1, pg2); stat[1] = 0.5f * vloss; stat[2] = ent; stat[3] = -logratio;
        stat[4] = (ratio - 1.f) - logratio; stat[5] = fabsf(ratio - 1.f) > c ? 1.f : 0.f;
        d3[A] = dv; g3[A] = dv;
      }
#pragma unroll
      for (int j = 0; j < DPT; j++) {
        const int i = part + TPS * j;
        if (i < A) { const float g = dmean[j] * g_logp; d3[i] = g; g3[i] = g; dl[j] = g_logp * dl[j] - a.hp.ent_coef * invB; }
      }
    } else {
#pragma unroll
      for (int j = 0; j < DPT; j++) { const int i = part + TPS * j; if (i <= A) d3[i] = 0.f; }
    }
    // fixed-order reduction over the tile: shuffle tree inside each warp, the warps summed below in order
    for (int i = 0; i < NP; i++) {
      float v = 0.f;
      if (i < PB_NSTAT) {
#pragma unroll
        for (int q = 0; q < PB_NSTAT; q++) if (q == i) v = stat[q];
      } else if (((i - PB_NSTAT) % TPS) == part) {
#pragma unroll
        for (int q = 0; q < DPT; q++) if (q == (i - PB_NSTAT) / TPS) v = dl[q];
      }
      for (int sft = 16; sft > 0; sft >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sft);
      if (lane == 0) red[warp * NP + i] = v;
    }
  }
  __syncthreads();
  if (tid < NP) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < FB_THREADS / 32; w++) v += red[w * NP + tid];
    a.part[(size_t)blockIdx.x * NP + tid] = v;
  }
  FB_TICK(4);
  // ---- delta through the second tanh: D2 = (D3 W3) * (1 - h2^2), in place over h2
#pragma unroll
  for (int i = 0; i < FB_RPT; i++) {
    const int m = ty * FB_RPT + i;
    const float *d3 = d3s + m * 20;
    float *hp = H2s + m * FB_LDH + tx * 4;
    float4 hc = *reinterpret_cast<const float4 *>(hp), ha = *reinterpret_cast<const float4 *>(hp + 64);
    const float dv = d3[A];
    float4 wc = *reinterpret_cast<const float4 *>(W3s + A * FB_LDW3 + tx * 4);
    float4 dc = make_float4(dv * wc.x * (1.f - hc.x * hc.x), dv * wc.y * (1.f - hc.y * hc.y), dv * wc.z * (1.f - hc.z * hc.z),
                            dv * wc.w * (1.f - hc.w * hc.w));
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (int o = 0; o < A; o++) {
      float4 w = *reinterpret_cast<const float4 *>(W3s + o * FB_LDW3 + tx * 4);
      const float g = d3[o];
      s0 += g * w.x; s1 += g * w.y; s2 += g * w.z; s3 += g * w.w;
    }
    float4 da = make_float4(s0 * (1.f - ha.x * ha.x), s1 * (1.f - ha.y * ha.y), s2 * (1.f - ha.z * ha.z), s3 * (1.f - ha.w * ha.w));
    *reinterpret_cast<float4 *>(hp) = dc;
    *reinterpret_cast<float4 *>(hp + 64) = da;
    if (m < rows) {
      *reinterpret_cast<float4 *>(a.D2 + (out0 + m) * 128 + tx * 4) = dc;
      *reinterpret_cast<float4 *>(a.D2 + (out0 + m) * 128 + 64 + tx * 4) = da;
    }
  }
  __syncthreads();
  FB_TICK(5);
  // ---- delta through the first tanh: D1 = (D2 W2) * (1 - h1^2)
#pragma unroll
  for (int i = 0; i < FB_RPT; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.f;
  fb_layer64<FB_RPT>(acc, H2s, W2s, tx, ty);
#pragma unroll
  for (int i = 0; i < FB_RPT; i++) {
    const int m = ty * FB_RPT + i;
    if (m >= rows) continue;
#pragma unroll
    for (int jh = 0; jh < 2; jh++) {
      float4 h = *reinterpret_cast<const float4 *>(H1s + m * FB_LDH + jh * 64 + tx * 4);
      float4 d = make_float4(acc[i][jh * 4] * (1.f - h.x * h.x), acc[i][jh * 4 + 1] * (1.f - h.y * h.y),
                             acc[i][jh * 4 + 2] * (1.f - h.z * h.z), acc[i][jh * 4 + 3] * (1.f - h.w * h.w));
      *reinterpret_cast<float4 *>(a.D1 + (out0 + m) * 128 + jh * 64 + tx * 4) = d;
    }
  }
  FB_TICK(6);
  if (a.debug && blockIdx.x == 0 && threadIdx.x == 0)
    printf("fwd_bwd phases (cycles): layer1 %lld layer2 %lld heads %lld loss %lld d2 %lld d1 %lld total %lld\n", tick[1] - tick[0],
           tick[2] - tick[1], tick[3] - tick[2], tick[4] - tick[3], tick[5] - tick[4], tick[6] - tick[5], tick[6] - tick[0]);
}

// ------------------------------------------------------------------ weight gradients: C[j][k] = sum_m Aop[m][j] * Bop[m][k]
// One launch, jobs laid out job-major so the heavy first-layer tiles start first:
//   jobs [0, nt1)   : dW1 column tile (128 hidden units of both nets x 64 observation columns)
//   jobs nt1, nt1+1 : dW2 of the critic / actor (64 x 64)
//   jobs nt1+2, +3  : head weights (dW3) against the critic / actor half of h2
// bias gradients are the column sums of the A operands, accumulated inside the same tiles
// blockIdx.x = job * splits + z; split z reduces its own chunk of the minibatch rows into slab z (fixed order).
struct WgradArgs {
  const float *Xp; const int *cursor; int mb, splits;
  AgentLayout L;
  const float *H1, *H2, *D1, *D2, *D3;
  float *slabs; size_t slab_stride;
  int skip_w1;   // 1: dW1 comes from dw1_tc_kernel; the job list is dW2 x2, dW3 x2 and the first-layer bias column sums
};

#define WG_BK 16

// BM x 64 output tile over rows [m0, m1): Aop rows have lda floats (BM consecutive columns from a0), Bop rows ldb (64 columns
// from b0, bcols of them valid).  256 threads, (BM/16) x 4 outputs each, register-staged double buffering.
// COLSUM additionally accumulates csum[r] = sum_m Aop[m][row r of this thread] (the bias gradient of that layer).
template <int BM, bool COLSUM>
__device__ __forceinline__ void wg_tile(const float *__restrict__ Aop, int lda, const float *__restrict__ Bop, int ldb, int bcols,
                                        bool baligned, int m0, int m1, float *sA, float *sB, float (&acc)[BM / 16][4],
                                        float (&csum)[BM / 16]) {
  constexpr int RA = BM / 64;  // float4 loads of A per thread per stage
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  float4 ar[RA], br;
  auto gload = [&](int ms) {
#pragma unroll
    for (int i = 0; i < RA; i++) {
      int e = tid + 256 * i, kk = e / (BM / 4), j4 = e % (BM / 4), m = ms + kk;
      ar[i] = m < m1 ? *reinterpret_cast<const float4 *>(Aop + (size_t)m * lda + j4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    int kk = tid >> 4, c = (tid & 15) * 4, m = ms + kk;
    br = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m < m1) {
      const float *p = Bop + (size_t)m * ldb + c;
      if (baligned) { if (c < bcols) br = *reinterpret_cast<const float4 *>(p); }
      else { if (c < bcols) br.x = p[0]; if (c + 1 < bcols) br.y = p[1]; if (c + 2 < bcols) br.z = p[2]; if (c + 3 < bcols) br.w = p[3]; }
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < RA; i++) {
      int e = tid + 256 * i, kk = e / (BM / 4), j4 = e % (BM / 4);
      *reinterpret_cast<float4 *>(sA + (buf * WG_BK + kk) * BM + j4 * 4) = ar[i];
    }
    *reinterpret_cast<float4 *>(sB + (buf * WG_BK + (tid >> 4)) * 64 + (tid & 15) * 4) = br;
  };
  const int nst = (m1 - m0 + WG_BK - 1) / WG_BK;
  if (nst <= 0) return;
  gload(m0); sstore(0);
  __syncthreads();
  for (int s = 0; s < nst; s++) {
    if (s + 1 < nst) gload(m0 + (s + 1) * WG_BK);
    const float *A_ = sA + (s & 1) * WG_BK * BM + ty * 4, *B_ = sB + (s & 1) * WG_BK * 64 + tx * 4;
#pragma unroll
    for (int kk = 0; kk < WG_BK; kk++) {
      float4 b = *reinterpret_cast<const float4 *>(B_ + kk * 64);
#pragma unroll
      for (int g = 0; g < BM / 64; g++) {
        float4 av = *reinterpret_cast<const float4 *>(A_ + kk * BM + g * 64);
        const float a4[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
        for (int i = 0; i < 4; i++) {
          acc[g * 4 + i][0] += a4[i] * b.x; acc[g * 4 + i][1] += a4[i] * b.y; acc[g * 4 + i][2] += a4[i] * b.z; acc[g * 4 + i][3] += a4[i] * b.w;
          if (COLSUM) csum[g * 4 + i] += a4[i];
        }
      }
    }
    if (s + 1 < nst) sstore((s + 1) & 1);
    __syncthreads();
  }
}

// Tensor-core form of the 128 x 64 first-layer tile: C[j][k] = sum_m D1[m][j] X[m][k] as m16n8k8 3xTF32 products with the
// minibatch rows as the MMA k dimension.  hi / lo halves of the D1 chunk [m][j] (stride 136) and of the X chunk [m][k]
// (stride 72) in shared memory; 8 warps = 4 (32 hidden units) x 2 (32 observation columns), 2 x 4 accumulators each.
#define WG_LDA_M 136
#define WG_LDB_M 72
#define WG_SMEM_FLOATS_MMA (4 * WG_BK * WG_LDA_M + 4 * WG_BK * WG_LDB_M)
template <bool COLSUM>
__device__ __forceinline__ void wg_tile_mma(const float *__restrict__ Aop, const float *__restrict__ Bop, int ldb, int bcols, bool baligned,
                                            int m0, int m1, float *sA, float *sB, float (&acc)[2][4][4], float &csum) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int j0 = (warp & 3) * 32, n0 = (warp >> 2) * 32;
  float4 ar[2], br;
  auto gload = [&](int ms) {
#pragma unroll
    for (int i = 0; i < 2; i++) {
      int e = tid + 256 * i, kk = e >> 5, j4 = e & 31, m = ms + kk;
      ar[i] = m < m1 ? *reinterpret_cast<const float4 *>(Aop + (size_t)m * 128 + j4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    int kk = tid >> 4, c = (tid & 15) * 4, m = ms + kk;
    br = make_float4(0.f, 0.f, 0.f, 0.f);
    if (m < m1) {
      const float *p = Bop + (size_t)m * ldb + c;
      if (baligned) { if (c < bcols) br = *reinterpret_cast<const float4 *>(p); }
      else { if (c < bcols) br.x = p[0]; if (c + 1 < bcols) br.y = p[1]; if (c + 2 < bcols) br.z = p[2]; if (c + 3 < bcols) br.w = p[3]; }
    }
  };
  auto sstore = [&](int buf) {
    float *Ah = sA + (buf * 2) * WG_BK * WG_LDA_M, *Al = Ah + WG_BK * WG_LDA_M;
#pragma unroll
    for (int i = 0; i < 2; i++) {
      int e = tid + 256 * i, kk = e >> 5, j4 = e & 31;
      float4 h, l;
      tf32_split(ar[i].x, h.x, l.x); tf32_split(ar[i].y, h.y, l.y); tf32_split(ar[i].z, h.z, l.z); tf32_split(ar[i].w, h.w, l.w);
      *reinterpret_cast<float4 *>(Ah + kk * WG_LDA_M + j4 * 4) = h;
      *reinterpret_cast<float4 *>(Al + kk * WG_LDA_M + j4 * 4) = l;
    }
    float *Bh = sB + (buf * 2) * WG_BK * WG_LDB_M, *Bl = Bh + WG_BK * WG_LDB_M;
    float4 h, l;
    tf32_split(br.x, h.x, l.x); tf32_split(br.y, h.y, l.y); tf32_split(br.z, h.z, l.z); tf32_split(br.w, h.w, l.w);
    *reinterpret_cast<float4 *>(Bh + (tid >> 4) * WG_LDB_M + (tid & 15) * 4) = h;
    *reinterpret_cast<float4 *>(Bl + (tid >> 4) * WG_LDB_M + (tid & 15) * 4) = l;
  };
  const int nst = (m1 - m0 + WG_BK - 1) / WG_BK;
  if (nst <= 0) return;
  gload(m0); sstore(0);
  __syncthreads();
  for (int s = 0; s < nst; s++) {
    if (s + 1 < nst) gload(m0 + (s + 1) * WG_BK);
    const float *Ah = sA + ((s & 1) * 2) * WG_BK * WG_LDA_M, *Al = Ah + WG_BK * WG_LDA_M;
    const float *Bh = sB + ((s & 1) * 2) * WG_BK * WG_LDB_M, *Bl = Bh + WG_BK * WG_LDB_M;
#pragma unroll
    for (int ks = 0; ks < WG_BK / 8; ks++) {
      float ah[2][4], al[2][4];
#pragma unroll
      for (int mt = 0; mt < 2; mt++) {
        const int ia = (ks * 8 + t) * WG_LDA_M + j0 + mt * 16 + g;
        ah[mt][0] = Ah[ia]; ah[mt][1] = Ah[ia + 8]; ah[mt][2] = Ah[ia + 4 * WG_LDA_M]; ah[mt][3] = Ah[ia + 4 * WG_LDA_M + 8];
        al[mt][0] = Al[ia]; al[mt][1] = Al[ia + 8]; al[mt][2] = Al[ia + 4 * WG_LDA_M]; al[mt][3] = Al[ia + 4 * WG_LDA_M + 8];
      }
#pragma unroll
      for (int nt = 0; nt < 4; nt++) {
        const int ib = (ks * 8 + t) * WG_LDB_M + n0 + nt * 8 + g;
        const float bh0 = Bh[ib], bh1 = Bh[ib + 4 * WG_LDB_M], bl0 = Bl[ib], bl1 = Bl[ib + 4 * WG_LDB_M];
#pragma unroll
        for (int mt = 0; mt < 2; mt++) mma_3xtf32(acc[mt][nt], ah[mt], al[mt], bh0, bh1, bl0, bl1);
      }
    }
    if (COLSUM && tid < 128) {
#pragma unroll
      for (int kk = 0; kk < WG_BK; kk++) csum += Ah[kk * WG_LDA_M + tid] + Al[kk * WG_LDA_M + tid];
    }
    if (s + 1 < nst) sstore((s + 1) & 1);
    __syncthreads();
  }
}

template <int MINB, bool MMA = false>
static __global__ void __launch_bounds__(256, MINB) wgrad_kernel(WgradArgs a) {
  extern __shared__ __align__(16) float wg_sm[];   // MMA: WG_SMEM_FLOATS_MMA floats, else 2*WG_BK*(128+64)
  float *sA = wg_sm, *sB = wg_sm + (MMA ? 4 * WG_BK * WG_LDA_M : 2 * WG_BK * 128);
  const int O = a.L.obs, A = a.L.act, nt1 = a.skip_w1 ? 0 : (O + 63) / 64;
  const int job = blockIdx.x / a.splits, z = blockIdx.x - job * a.splits;
  const int chunk = (a.mb + a.splits - 1) / a.splits, m0 = z * chunk, m1 = min(a.mb, m0 + chunk);
  float *slab = a.slabs + (size_t)z * a.slab_stride;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  if (MMA && job < nt1) {
    // tensor-core tile; accumulator (mt, nt, c): hidden unit j0 + 16 mt + g (+8 for c >= 2), column n0 + 8 nt + 2 t + (c & 1)
    float acc[2][4][4], cs = 0.f;
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
      for (int nt = 0; nt < 4; nt++) acc[mt][nt][0] = acc[mt][nt][1] = acc[mt][nt][2] = acc[mt][nt][3] = 0.f;
    const int k0 = job * 64;
    const float *X = a.Xp + (size_t)(*a.cursor) * a.mb * O + k0;
    if (job == 0) wg_tile_mma<true>(a.D1, X, O, O - k0, (O & 3) == 0, m0, m1, sA, sB, acc, cs);
    else wg_tile_mma<false>(a.D1, X, O, O - k0, (O & 3) == 0, m0, m1, sA, sB, acc, cs);
    const int lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int j0 = (warp & 3) * 32, n0 = (warp >> 2) * 32;
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int j = j0 + mt * 16 + g + 8 * h, net = j >> 6, jl = j & 63;
        float *dst = slab + (net ? a.L.aW1 : a.L.cW1) + (size_t)jl * O + k0;
#pragma unroll
        for (int nt = 0; nt < 4; nt++) {
          const int k = n0 + nt * 8 + 2 * t;
          if (k0 + k < O) dst[k] = acc[mt][nt][2 * h];
          if (k0 + k + 1 < O) dst[k + 1] = acc[mt][nt][2 * h + 1];
        }
      }
    if (job == 0 && tid < 128) slab[(tid >> 6 ? a.L.ab1 : a.L.cb1) + (tid & 63)] = cs;
  } else if (job < nt1) {
    // dW1[j][k0 + 0..63] for the 128 hidden units (critic rows, then actor rows); tile 0 also sums the first-layer biases
    float acc[8][4], cs[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f; cs[i] = 0.f; }
    const int k0 = job * 64;
    const float *X = a.Xp + (size_t)(*a.cursor) * a.mb * O + k0;
    if (job == 0) wg_tile<128, true>(a.D1, 128, X, O, O - k0, (O & 3) == 0, m0, m1, sA, sB, acc, cs);
    else wg_tile<128, false>(a.D1, 128, X, O, O - k0, (O & 3) == 0, m0, m1, sA, sB, acc, cs);
#pragma unroll
    for (int g = 0; g < 2; g++)
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int j = ty * 4 + i;  // hidden unit inside net g
        float *dst = slab + (g ? a.L.aW1 : a.L.cW1) + (size_t)j * O + k0 + tx * 4;
#pragma unroll
        for (int c = 0; c < 4; c++) if (k0 + tx * 4 + c < O) dst[c] = acc[g * 4 + i][c];
        if (job == 0 && tx == 0) slab[(g ? a.L.ab1 : a.L.cb1) + j] = cs[g * 4 + i];
      }
  } else if (job == nt1 + 4) {
    // first-layer bias gradients: column sums of D1 over the split's rows; 8 row groups x 32 lanes of 16-byte loads,
    // partial sums combined in fixed order
    const int c4 = tid & 31, rg = tid >> 5;
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int m = m0 + rg; m < m1; m += 8) {
      const float4 d = *reinterpret_cast<const float4 *>(a.D1 + (size_t)m * 128 + 4 * c4);
      sum.x += d.x; sum.y += d.y; sum.z += d.z; sum.w += d.w;
    }
    *reinterpret_cast<float4 *>(sA + rg * 128 + 4 * c4) = sum;
    __syncthreads();
    if (tid < 128) {
      float r = 0.f;
#pragma unroll
      for (int g = 0; g < 8; g++) r += sA[g * 128 + tid];
      slab[(tid >> 6 ? a.L.ab1 : a.L.cb1) + (tid & 63)] = r;
    }
  } else if (job < nt1 + 2) {
    const int net = job - nt1;
    float acc[4][4], cs[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f; cs[i] = 0.f; }
    wg_tile<64, true>(a.D2 + net * 64, 128, a.H1 + net * 64, 128, 64, true, m0, m1, sA, sB, acc, cs);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float *dst = slab + (net ? a.L.aW2 : a.L.cW2) + (ty * 4 + i) * 64 + tx * 4;
      dst[0] = acc[i][0]; dst[1] = acc[i][1]; dst[2] = acc[i][2]; dst[3] = acc[i][3];
      if (tx == 0) slab[(net ? a.L.ab2 : a.L.cb2) + ty * 4 + i] = cs[i];
    }
  } else {
    // head weights: D3w[m][o] (64 zero-padded columns: o < A actor means, o == A value) against one half of H2
    const int half = job - nt1 - 2;   // 0: critic half (row A only), 1: actor half (rows < A)
    float acc[4][4], cs[4];
#pragma unroll
    for (int i = 0; i < 4; i++) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f; cs[i] = 0.f; }
    wg_tile<64, true>(a.D3, 64, a.H2 + half * 64, 128, 64, true, m0, m1, sA, sB, acc, cs);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int o = ty * 4 + i;
      if (half == 1 && o < A) {
        float *dst = slab + a.L.aW3 + o * 64 + tx * 4;
        dst[0] = acc[i][0]; dst[1] = acc[i][1]; dst[2] = acc[i][2]; dst[3] = acc[i][3];
        if (tx == 0) slab[a.L.ab3 + o] = cs[i];
      } else if (half == 0 && o == A) {
        float *dst = slab + a.L.cW3 + tx * 4;
        dst[0] = acc[i][0]; dst[1] = acc[i][1]; dst[2] = acc[i][2]; dst[3] = acc[i][3];
        if (tx == 0) slab[a.L.cb3] = cs[i];
      }
    }
  }
}

// ------------------------------------------------------------------ small weight-gradient jobs beside dw1_tc_kernel
// On the tcgen05 path dW1 comes from dw1_tc_kernel; what remains is dW2 (2 nets), dW3 (2 halves of h2) and the first-layer
// bias column sums: five jobs x splits CTAs of 64 x 64 tiles over a few hundred rows each.  wgrad_small_kernel does the
// same arithmetic in the same summation order as wgrad_kernel's branches for those jobs (bit-identical slabs), with two
// differences that matter for time:
//  * operands are streamed with cp.async through a ring of WGA_ST stages of WGA_BK rows, three stages in flight — the
//    one-stage register prefetch of wg_tile left every 16-row step exposed to an L2 round trip, so the 160-CTA launch was
//    latency-bound at ~22 us;
//  * it is sized to share an SM with a dw1_tc_kernel CTA (192 KB shared memory, 288 threads x 141 registers): 32 KB of
//    shared memory and <= 80 registers, so the side-stream fork really overlaps (wgrad_kernel's 112 registers x 256
//    threads plus dw1_tc_kernel's 144 x 288 exceed the 64 K register file, which is why the two never co-resided).
#define WGA_BK 16
#define WGA_ST 4
#define WGA_SMEM_BYTES (2 * WGA_ST * WGA_BK * 64 * 4)
__device__ __forceinline__ void wg_cp_async16(float *dst_smem, const float *src, bool valid) {
  const unsigned d = (unsigned)__cvta_generic_to_shared(dst_smem);
  const int n = valid ? 16 : 0;   // src-size 0: the 16 bytes are zero-filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void wg_tile_async(const float *__restrict__ Aop, int lda, const float *__restrict__ Bop, int ldb,
                                              int m0, int m1, float *sm, float (&acc)[4][4], float (&csum)[4]) {
  float *sA = sm, *sB = sm + WGA_ST * WGA_BK * 64;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int nst = (m1 - m0 + WGA_BK - 1) / WGA_BK;
  if (nst <= 0) return;
  auto load = [&](int st) {   // 16 rows x 16 float4 per operand: one A and one B copy per thread
    const int buf = st % WGA_ST, kk = tid >> 4, j4 = tid & 15, m = m0 + st * WGA_BK + kk;
    const bool ok = m < m1;
    const size_t row = ok ? m : m1 - 1;
    wg_cp_async16(sA + (buf * WGA_BK + kk) * 64 + j4 * 4, Aop + row * lda + j4 * 4, ok);
    wg_cp_async16(sB + (buf * WGA_BK + kk) * 64 + j4 * 4, Bop + row * ldb + j4 * 4, ok);
  };
#pragma unroll
  for (int p = 0; p < WGA_ST - 1; p++) {
    if (p < nst) load(p);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (int s = 0; s < nst; s++) {
    asm volatile("cp.async.wait_group %0;" ::"n"(WGA_ST - 2) : "memory");
    __syncthreads();   // stage s has landed for every thread; buffer (s - 1) % WGA_ST is no longer being read
    if (s + WGA_ST - 1 < nst) load(s + WGA_ST - 1);
    asm volatile("cp.async.commit_group;" ::: "memory");
    const float *A_ = sA + (s % WGA_ST) * WGA_BK * 64 + ty * 4, *B_ = sB + (s % WGA_ST) * WGA_BK * 64 + tx * 4;
#pragma unroll
    for (int kk = 0; kk < WGA_BK; kk++) {
      const float4 b = *reinterpret_cast<const float4 *>(B_ + kk * 64);
      const float4 av = *reinterpret_cast<const float4 *>(A_ + kk * 64);
      const float a4[4] = {av.x, av.y, av.z, av.w};
#pragma unroll
      for (int i = 0; i < 4; i++) {
        acc[i][0] += a4[i] * b.x; acc[i][1] += a4[i] * b.y; acc[i][2] += a4[i] * b.z; acc[i][3] += a4[i] * b.w;
        csum[i] += a4[i];
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// grid = 5 * splits: jobs 0,1 dW2 (critic, actor) | 2,3 dW3 against the critic / actor half of h2 | 4 first-layer bias sums
static __global__ void __launch_bounds__(256, 3) wgrad_small_kernel(WgradArgs a) {
  extern __shared__ __align__(16) float wgs_sm[];   // WGA_SMEM_BYTES
  const int A = a.L.act;
  const int job = blockIdx.x / a.splits, z = blockIdx.x - job * a.splits;
  const int chunk = (a.mb + a.splits - 1) / a.splits, m0 = z * chunk, m1 = min(a.mb, m0 + chunk);
  float *slab = a.slabs + (size_t)z * a.slab_stride;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  if (job == 4) {
    // column sums of D1 over the split's rows; 8 row groups x 32 lanes of 16-byte loads, partial sums combined in fixed order
    const int c4 = tid & 31, rg = tid >> 5;
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int m = m0 + rg; m < m1; m += 8) {
      const float4 d = *reinterpret_cast<const float4 *>(a.D1 + (size_t)m * 128 + 4 * c4);
      sum.x += d.x; sum.y += d.y; sum.z += d.z; sum.w += d.w;
    }
    *reinterpret_cast<float4 *>(wgs_sm + rg * 128 + 4 * c4) = sum;
    __syncthreads();
    if (tid < 128) {
      float r = 0.f;
#pragma unroll
      for (int g = 0; g < 8; g++) r += wgs_sm[g * 128 + tid];
      slab[(tid >> 6 ? a.L.ab1 : a.L.cb1) + (tid & 63)] = r;
    }
    return;
  }
  float acc[4][4], cs[4];
#pragma unroll
  for (int i = 0; i < 4; i++) { acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.f; cs[i] = 0.f; }
  if (job < 2) {
    const int net = job;
    wg_tile_async(a.D2 + net * 64, 128, a.H1 + net * 64, 128, m0, m1, wgs_sm, acc, cs);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float *dst = slab + (net ? a.L.aW2 : a.L.cW2) + (ty * 4 + i) * 64 + tx * 4;
      dst[0] = acc[i][0]; dst[1] = acc[i][1]; dst[2] = acc[i][2]; dst[3] = acc[i][3];
      if (tx == 0) slab[(net ? a.L.ab2 : a.L.cb2) + ty * 4 + i] = cs[i];
    }
  } else {
    // head weights: D3w[m][o] (64 zero-padded columns: o < A actor means, o == A value) against one half of H2
    const int half = job - 2;   // 0: critic half (row A only), 1: actor half (rows < A)
    wg_tile_async(a.D3, 64, a.H2 + half * 64, 128, m0, m1, wgs_sm, acc, cs);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int o = ty * 4 + i;
      if (half == 1 && o < A) {
        float *dst = slab + a.L.aW3 + o * 64 + tx * 4;
        dst[0] = acc[i][0]; dst[1] = acc[i][1]; dst[2] = acc[i][2]; dst[3] = acc[i][3];
        if (tx == 0) slab[a.L.ab3 + o] = cs[i];
      } else if (half == 0 && o == A) {
        float *dst = slab + a.L.cW3 + tx * 4;
        dst[0] = acc[i][0]; dst[1] = acc[i][1]; dst[2] = acc[i][2]; dst[3] = acc[i][3];
        if (tx == 0) slab[a.L.cb3] = cs[i];
      }
    }
  }
}

// grad[p] = sum_z slab[z][p] in fixed order (logstd entries from the per-tile partials) + per-block sum of squares.
// The block(s) holding the logstd entries first reduce the per-tile partial columns (statistics | dL/dlogstd) with one
// warp per column; the last block publishes those column sums for adam2_kernel's statistics.
static __global__ void __launch_bounds__(256) grad_reduce_kernel(const float *__restrict__ slabs, size_t slab_stride, int splits,
                                                          const float *__restrict__ part, int ntile, AgentLayout L,
                                                          float *__restrict__ grad, float *__restrict__ partials,
                                                          float *step_count, float *__restrict__ colsum,
                                                          const unsigned *__restrict__ peer_seq = nullptr, int peer_stride = 0) {
  __shared__ float sm[256];
  __shared__ float cols[PB_NSTAT + FB_MAXA + 1];
  // data-parallel peer path (pb_peer.cuh): `grad` is the double-buffered exchange buffer, slot = completed steps & 1
  if (peer_seq) grad += (size_t)(*peer_seq & 1u) * peer_stride;
  const int p = blockIdx.x * 256 + threadIdx.x, NP = PB_NSTAT + L.act;
  const bool last = blockIdx.x == gridDim.x - 1;
  if (last || (blockIdx.x + 1) * 256 > L.logstd) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int c = warp; c < NP; c += 8) {
      float v = 0.f;
      for (int b = lane; b < ntile; b += 32) v += part[(size_t)b * NP + c];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) { cols[c] = v; if (last) colsum[c] = v; }
    }
    __syncthreads();
  }
  float s = 0.f;
  if (p < L.total) {
    if (p >= L.logstd) s = cols[PB_NSTAT + (p - L.logstd)];
    else for (int z = 0; z < splits; z++) s += slabs[(size_t)z * slab_stride + p];
    grad[p] = s;
  }
  if (step_count && blockIdx.x == 0 && threadIdx.x == 0) *step_count += 1.f;  // optimiser step, read by adam2_kernel
  sm[threadIdx.x] = s * s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) partials[blockIdx.x] = sm[0];
}

// clip_grad_norm_ + torch Adam step; also refreshes the transposed weight copies the forward kernels read, publishes the
// minibatch statistics and advances the minibatch cursor (last block).
static __global__ void __launch_bounds__(256) adam2_kernel(AdamState s, const float *__restrict__ partials, int npart, PpoHyper hp,
                                                    float *scalars, int *cursor, float *stats_out, float *clip_acc,
                                                    const float *__restrict__ colsum, int mb, AgentLayout L,
                                                    float *__restrict__ W1T, float *__restrict__ W2T, float *__restrict__ Wc,
                                                    unsigned *peer_seq = nullptr) {
  __shared__ float sm[256];
  {
    float v = 0.f;
    for (int i = threadIdx.x; i < npart; i += 256) v += partials[i];
    sm[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  }
  const float cn = hp.max_grad_norm / (sqrtf(sm[0]) + 1e-6f), clip_s = cn < 1.f ? cn : 1.f;
  const float lr = scalars[0], t = scalars[1];   // step count was advanced by grad_reduce_kernel
  const float bc1 = 1.f - powf(hp.beta1, t), bc2 = 1.f - powf(hp.beta2, t);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < s.n) {
    float g = s.g[i] * clip_s;
    float m = hp.beta1 * s.m[i] + (1.f - hp.beta1) * g;
    float v = hp.beta2 * s.v[i] + (1.f - hp.beta2) * g * g;
    s.m[i] = m; s.v[i] = v;
    float denom = sqrtf(v) / sqrtf(bc2) + hp.adam_eps;
    float pn = s.p[i] - (lr / bc1) * (m / denom);
    s.p[i] = pn;
    const int O = L.obs;
    // Wc: TF32 hi / lo halves of W1 in the stage layout of l1_forward_tc_kernel: [hi|lo][k/32][(k%32)/4][n 128][k%4]
    int wn = -1, wk = 0;
    if (i < L.cb1) { int j = i / O, k = i - j * O; W1T[k * 128 + j] = pn; wn = j; wk = k; }
    else if (i >= L.cW2 && i < L.cb2) { int r = i - L.cW2, u = r >> 6, k = r & 63; W2T[k * 64 + u] = pn; }
    else if (i >= L.aW1 && i < L.ab1) { int r = i - L.aW1, j = r / O, k = r - j * O; W1T[k * 128 + 64 + j] = pn; wn = 64 + j; wk = k; }
    else if (i >= L.aW2 && i < L.ab2) { int r = i - L.aW2, u = r >> 6, k = r & 63; W2T[4096 + k * 64 + u] = pn; }
    if (Wc && wn >= 0) {
      float hi, lo;
      tf32_split(pn, hi, lo);
      const int idx = (((wk >> 5) * 8 + ((wk & 31) >> 2)) * 128 + wn) * 4 + (wk & 3);
      Wc[idx] = hi; Wc[(size_t)((O + 31) >> 5) * 4096 + idx] = lo;
    }
  }
  if (blockIdx.x == gridDim.x - 1) {
    if (threadIdx.x < PB_NSTAT && stats_out) {
      float v = colsum[threadIdx.x] / (float)mb;
      stats_out[threadIdx.x] = v;
      if (threadIdx.x == 5 && clip_acc) *clip_acc += v;
    }
    if (threadIdx.x == 32) *cursor += 1;
    if (threadIdx.x == 64 && peer_seq) *peer_seq += 1u;   // exchange step completed (pb_peer.cuh)
  }
}
