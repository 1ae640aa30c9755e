// pb_ac_tc.cuh — the 256 x 256 products of the ac_ppo encoders on tcgen05 with the LayerNorm that follows them fused in.
//
//   forward   x_hat2, rstd2 = LayerNorm( relu(gamma1 x_hat1 + beta1) W2^T + b2 )          (src/ac_ppo_continuous_action.cpp:150-172)
//   backward  d1 = LayerNormReLU'( d2 W2 ; x_hat1, rstd1, gamma1, beta1 ), dgamma1 / dbeta1 partial sums
//
// Both are C[M x 256] = A[M x 256] B^T with M = 262 144 minibatch rows: 268 MB in, 268 MB out, 103 GFLOP as 3xTF32 — the HBM time
// (82 us) and the tensor-pipe time (94 us) are about equal, so the kernel is persistent (one CTA per SM, 128-row tiles taken
// round-robin) and every stage of a tile overlaps the neighbouring tiles:
//   warps 4-11  producers: load the A chunk (128 rows x 16 floats) G_PF chunks ahead into registers, apply the input transform
//               (forward: relu(gamma x_hat + beta), so the post-activation matrix is never stored), split into TF32 hi / lo and
//               store both halves in the K-major canonical layout of pb_ppo_tc.cuh
//   warp 12     one lane: TMA bulk copies of the pre-split weight chunk (hi / lo, 16 KB each, L2 resident) two chunks ahead and
//               the 6 tcgen05.mma (128 x 256 x 8, a_lo b_hi + a_hi b_lo + a_hi b_hi) of each chunk; the accumulator is one of
//               two 256-column TMEM buffers, committed to acc_full when the tile's 16 chunks are issued
//   warps 0-3   epilogue, one thread per row: reads the finished accumulator from TMEM (the other buffer is being filled),
//               does the whole LayerNorm row computation in registers and writes the result; arrives on acc_empty
// The chunk pipeline (G_NST shared-memory stages) runs across tile boundaries, so the next tile's A is in flight while the
// current one finishes.
#pragma once
#include "pb_ppo_tc.cuh"
#include "pb_ac_fused.cuh"

namespace actc {
using namespace tc;

#define G_BK 16
#define G_NST 4
#define G_NCHUNK (AC_HID / G_BK)                              // 16 chunks per tile
#define G_A_HALF (128 * G_BK)                                 // floats: A hi or lo of one stage, 8 KB
#define G_B_HALF (AC_HID * G_BK)                              // floats: B hi or lo of one stage, 16 KB
#define G_STAGE_FLOATS (2 * G_A_HALF + 2 * G_B_HALF)          // 48 KB
#define G_PF 8                                                // register prefetch distance of the producers, in chunks
#define G_THREADS 416
#define G_VEC_FLOATS (4 * AC_HID)                             // gamma | beta (input transform or LN backward) | bias | spare
#define G_SMEM_BYTES (G_NST * G_STAGE_FLOATS * 4 + G_VEC_FLOATS * 4 + 256)

enum { G_EPI_LN_FWD = 0, G_EPI_LN_BWD = 1 };

struct GemmArgs {
  const float *A; int M;               // [M][256]
  const float *Wc;                     // pre-split B: [hi | lo][chunk 16][c16 4][n 256][4] (w_split256_kernel)
  const float *pg, *pb;                // input transform relu(pg a + pb); nullptr: none
  // G_EPI_LN_FWD: x_hat = LN(acc + bias), rstd
  const float *bias; float *xhat, *rstd;
  // G_EPI_LN_BWD: dout = rstd (dy gamma - mean(dy gamma) - x_hat mean(dy gamma x_hat)), dy = acc masked by gamma x_hat + beta > 0;
  // part[cta][0..256) = sum dy x_hat, part[cta][256..512) = sum dy over the CTA's rows
  const float *xh_in, *rstd_in, *gamma, *beta; float *dout, *part;
};

// B(n, k) = transposed ? W[k][n] : W[n][k] -> hi / lo halves, chunk-major stage layout (a chunk's half is 16 KB contiguous)
static __global__ void w_split256_kernel(const float *__restrict__ W, int transposed, float *__restrict__ Wc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= AC_HID * AC_HID) return;
  const int e = i & 3, n = (i >> 2) & 255, c16 = (i >> 10) & 3, chunk = i >> 12, k = chunk * G_BK + c16 * 4 + e;
  float hi, lo;
  split_tf32(transposed ? W[(size_t)k * AC_HID + n] : W[(size_t)n * AC_HID + k], hi, lo);
  Wc[i] = hi; Wc[AC_HID * AC_HID + i] = lo;
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

template <int EPI>
static __global__ void __launch_bounds__(G_THREADS, 1) gemm256_tc_kernel(GemmArgs a) {
  extern __shared__ __align__(128) unsigned char g_smem[];
  float *tiles = reinterpret_cast<float *>(g_smem);
  float *vec = tiles + G_NST * G_STAGE_FLOATS;                    // [0,256) gamma | [256,512) beta | [512,768) bias
  uint64_t *empty_bar = reinterpret_cast<uint64_t *>(vec + G_VEC_FLOATS);   // [G_NST] the MMAs that read the stage have completed
  uint64_t *fulla_bar = empty_bar + G_NST;                                  // [G_NST] A halves stored
  uint64_t *fullb_bar = fulla_bar + G_NST;                                  // [G_NST] B halves landed
  uint64_t *accf_bar = fullb_bar + G_NST;                                   // [2] accumulator buffer complete
  uint64_t *acce_bar = accf_bar + 2;                                        // [2] accumulator buffer drained by the epilogue
  __shared__ uint32_t tmem_base_sh;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = (a.M + 127) / 128;
  const int my_tiles = (int)blockIdx.x < ntiles ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == 0) tmem_alloc(&tmem_base_sh, 512);
  if (tid == 32) {
    for (int s = 0; s < G_NST; s++) { mbar_init(empty_bar + s, 1); mbar_init(fulla_bar + s, 1); mbar_init(fullb_bar + s, 1); }
    for (int b = 0; b < 2; b++) { mbar_init(accf_bar + b, 1); mbar_init(acce_bar + b, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < AC_HID; i += G_THREADS) {
    if (EPI == G_EPI_LN_FWD) {
      vec[i] = a.pg ? a.pg[i] : 1.f; vec[256 + i] = a.pb ? a.pb[i] : 0.f; vec[512 + i] = a.bias ? a.bias[i] : 0.f;
    } else {
      vec[i] = a.gamma[i]; vec[256 + i] = a.beta[i]; vec[512 + i] = 0.f;
    }
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_d = tmem_base_sh;
  auto tile = [&](int s, int which) {  // 0 A hi, 1 A lo, 2 B hi, 3 B lo
    return tiles + s * G_STAGE_FLOATS + (which < 2 ? which * G_A_HALF : 2 * G_A_HALF + (which - 2) * G_B_HALF);
  };
  const int total_chunks = my_tiles * G_NCHUNK;

  if (warp == 12) {
    // ------------------------------------------------ issuer: weight TMA two chunks ahead, 6 MMAs per chunk
    if (lane == 0 && my_tiles > 0) {
      const float *Wh = a.Wc, *Wl = a.Wc + AC_HID * AC_HID;
      auto issue_b = [&](int g) {
        const int s = g % G_NST, use = g / G_NST, c = g % G_NCHUNK;
        if (use > 0) mbar_wait(empty_bar + s, (uint32_t)((use - 1) & 1));
        mbar_expect_tx(fullb_bar + s, 2 * G_B_HALF * 4);
        bulk_g2s(tile(s, 2), Wh + (size_t)c * G_B_HALF, G_B_HALF * 4, fullb_bar + s);
        bulk_g2s(tile(s, 3), Wl + (size_t)c * G_B_HALF, G_B_HALF * 4, fullb_bar + s);
      };
      constexpr uint32_t IDESC = idesc_tf32(128, 256);
      issue_b(0);
      if (total_chunks > 1) issue_b(1);
      for (int g = 0; g < total_chunks; g++) {
        const int s = g % G_NST, use = g / G_NST, c = g % G_NCHUNK, tl = g / G_NCHUNK, buf = tl & 1;
        if (g + 2 < total_chunks) issue_b(g + 2);
        if (c == 0 && tl >= 2) { mbar_wait(acce_bar + buf, (uint32_t)(((tl >> 1) - 1) & 1)); fence_after_sync(); }
        mbar_wait(fulla_bar + s, (uint32_t)(use & 1));
        mbar_wait(fullb_bar + s, (uint32_t)(use & 1));
        fence_after_sync();
        const uint32_t ah = smem_u32(tile(s, 0)), al = smem_u32(tile(s, 1)), bh = smem_u32(tile(s, 2)), bl = smem_u32(tile(s, 3));
        const uint32_t dcol = tmem_d + (uint32_t)(buf * 256);
#pragma unroll
        for (int ks = 0; ks < G_BK / 8; ks++) {
          const uint64_t dah = smem_desc(ah + ks * 2 * (128 * 16), 128 * 16, 128), dal = smem_desc(al + ks * 2 * (128 * 16), 128 * 16, 128);
          const uint64_t dbh = smem_desc(bh + ks * 2 * (256 * 16), 256 * 16, 128), dbl = smem_desc(bl + ks * 2 * (256 * 16), 256 * 16, 128);
          mma_tf32_ss(dcol, dal, dbh, IDESC, (c > 0 || ks > 0) ? 1u : 0u);
          mma_tf32_ss(dcol, dah, dbl, IDESC, 1u);
          mma_tf32_ss(dcol, dah, dbh, IDESC, 1u);
        }
        // a commit is an asynchronous arrive on this CTA's shared memory: issue one only if somebody waits for it before exit
        if (g + G_NST < total_chunks) mma_commit(empty_bar + s);
        if (c == G_NCHUNK - 1) mma_commit(accf_bar + buf);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ------------------------------------------------ producers
    const int pt = tid - 128, pw = pt >> 5;
    const int rl = lane & 7, c16 = lane >> 3;               // 8 rows x 4 sixteen-byte pieces per warp and item
    const bool xform = a.pg != nullptr;
    auto gload = [&](int g, float4 (&x)[2]) {
      const int tl = g / G_NCHUNK, c = g % G_NCHUNK;
      const int row0 = ((int)blockIdx.x + tl * (int)gridDim.x) * 128;
#pragma unroll
      for (int i = 0; i < 2; i++) {
        const int r = (pw + 8 * i) * 8 + rl, row = row0 + r;
        x[i] = (g < total_chunks && row < a.M) ? *reinterpret_cast<const float4 *>(a.A + (size_t)row * AC_HID + c * G_BK + c16 * 4)
                                                 : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    auto stage = [&](int g, const float4 (&x)[2]) {
      const int s = g % G_NST, use = g / G_NST, c = g % G_NCHUNK;
      const int tl = g / G_NCHUNK, row0 = ((int)blockIdx.x + tl * (int)gridDim.x) * 128;
      float4 y[2] = {x[0], x[1]};
      if (xform) {
        const float4 gg = *reinterpret_cast<const float4 *>(vec + c * G_BK + c16 * 4), bb = *reinterpret_cast<const float4 *>(vec + 256 + c * G_BK + c16 * 4);
#pragma unroll
        for (int i = 0; i < 2; i++) {
          const bool live = row0 + (pw + 8 * i) * 8 + rl < a.M;   // rows beyond M stay zero (relu(beta) is not)
          y[i].x = live ? fmaxf(0.f, gg.x * y[i].x + bb.x) : 0.f; y[i].y = live ? fmaxf(0.f, gg.y * y[i].y + bb.y) : 0.f;
          y[i].z = live ? fmaxf(0.f, gg.z * y[i].z + bb.z) : 0.f; y[i].w = live ? fmaxf(0.f, gg.w * y[i].w + bb.w) : 0.f;
        }
      }
      if (use > 0) mbar_wait(empty_bar + s, (uint32_t)((use - 1) & 1));
      float *Ah = tile(s, 0), *Al = tile(s, 1);
#pragma unroll
      for (int i = 0; i < 2; i++) store_split(Ah, Al, c16 * (128 * 4) + ((pw + 8 * i) * 8 + rl) * 4, y[i]);
      fence_proxy_async();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (pt == 0) mbar_arrive(fulla_bar + s);
    };
    float4 xr[G_PF][2];
#pragma unroll
    for (int p = 0; p < G_PF - 1; p++) gload(p, xr[p]);
    for (int g0 = 0; g0 < total_chunks; g0 += G_PF) {
#pragma unroll
      for (int p = 0; p < G_PF; p++) {
        const int g = g0 + p;
        if (g < total_chunks) {
          gload(g + G_PF - 1, xr[(p + G_PF - 1) % G_PF]);
          stage(g, xr[p]);
        }
      }
    }
  } else {
    // ------------------------------------------------ epilogue: thread = accumulator row (TMEM lane)
    const int m = warp * 32 + lane;
    float colacc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) colacc[i] = 0.f;
    for (int tl = 0; tl < my_tiles; tl++) {
      const int buf = tl & 1, row = ((int)blockIdx.x + tl * (int)gridDim.x) * 128 + m;
      const bool live = row < a.M;
      mbar_wait(accf_bar + buf, (uint32_t)((tl >> 1) & 1));
      fence_after_sync();
      const uint32_t tbase = tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)(buf * 256);
      if (EPI == G_EPI_LN_FWD) {
        float s = 0.f;
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32];
          tmem_ld32(tbase + cb * 32, v);
#pragma unroll
          for (int i = 0; i < 32; i++) s += v[i] + vec[512 + cb * 32 + i];
        }
        const float mean = s * (1.f / AC_HID);
        float q = 0.f;
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32];
          tmem_ld32(tbase + cb * 32, v);
#pragma unroll
          for (int i = 0; i < 32; i++) { const float d = v[i] + vec[512 + cb * 32 + i] - mean; q += d * d; }
        }
        const float rs = rsqrtf(q * (1.f / AC_HID) + 1e-5f);
        float *dst = a.xhat + (size_t)row * AC_HID;
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32];
          tmem_ld32(tbase + cb * 32, v);
          if (live) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 bi = *reinterpret_cast<const float4 *>(vec + 512 + cb * 32 + i);
              *reinterpret_cast<float4 *>(dst + cb * 32 + i) = make_float4((v[i] + bi.x - mean) * rs, (v[i + 1] + bi.y - mean) * rs,
                                                                          (v[i + 2] + bi.z - mean) * rs, (v[i + 3] + bi.w - mean) * rs);
            }
          }
        }
        if (live) a.rstd[row] = rs;
      } else {
        const float *xr = a.xh_in + (size_t)row * AC_HID;
        const float rs = live ? a.rstd_in[row] : 0.f;
        float s1 = 0.f, s2 = 0.f;
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32], w[32];
          tmem_ld32(tbase + cb * 32, v);
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 x = live ? *reinterpret_cast<const float4 *>(xr + cb * 32 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 gg = *reinterpret_cast<const float4 *>(vec + cb * 32 + i), bb = *reinterpret_cast<const float4 *>(vec + 256 + cb * 32 + i);
            const float xs[4] = {x.x, x.y, x.z, x.w}, gs[4] = {gg.x, gg.y, gg.z, gg.w}, bs[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
            for (int e = 0; e < 4; e++) {
              const float dy = (live && gs[e] * xs[e] + bs[e] > 0.f) ? v[i + e] : 0.f;
              const float dxh = dy * gs[e];
              s1 += dxh; s2 += dxh * xs[e];
              v[i + e] = dy * xs[e];      // dgamma term
              w[i + e] = dy;              // dbeta term
            }
          }
          colacc[cb] += acf::butterfly32(v, lane);
          colacc[8 + cb] += acf::butterfly32(w, lane);
        }
        const float m1 = s1 * (1.f / AC_HID), m2 = s2 * (1.f / AC_HID);
        float *dst = a.dout + (size_t)row * AC_HID;
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32];
          tmem_ld32(tbase + cb * 32, v);
          if (live) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              const float4 x = *reinterpret_cast<const float4 *>(xr + cb * 32 + i);
              const float4 gg = *reinterpret_cast<const float4 *>(vec + cb * 32 + i), bb = *reinterpret_cast<const float4 *>(vec + 256 + cb * 32 + i);
              const float xs[4] = {x.x, x.y, x.z, x.w}, gs[4] = {gg.x, gg.y, gg.z, gg.w}, bs[4] = {bb.x, bb.y, bb.z, bb.w};
              float o[4];
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const float dxh = (gs[e] * xs[e] + bs[e] > 0.f) ? v[i + e] * gs[e] : 0.f;
                o[e] = rs * (dxh - m1 - xs[e] * m2);
              }
              *reinterpret_cast<float4 *>(dst + cb * 32 + i) = make_float4(o[0], o[1], o[2], o[3]);
            }
          }
        }
      }
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(acce_bar + buf);
    }
    if (EPI == G_EPI_LN_BWD) {
      // per-CTA partial sums: the four epilogue warps in ascending order through the (now idle) first operand stage
      float *red = tiles;   // [4][512]
      asm volatile("bar.sync 2, 128;" ::: "memory");   // all epilogue warps are past their last TMEM read; MMAs are complete
#pragma unroll
      for (int cb = 0; cb < 8; cb++) { red[warp * 512 + cb * 32 + lane] = colacc[cb]; red[warp * 512 + 256 + cb * 32 + lane] = colacc[8 + cb]; }
      asm volatile("bar.sync 2, 128;" ::: "memory");
      for (int c = tid; c < 512; c += 128) a.part[(size_t)blockIdx.x * 512 + c] = red[c] + red[512 + c] + red[1024 + c] + red[1536 + c];
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, 512);
}

// ------------------------------------------------------------------ weight gradient of a 256 x 256 layer in one pass over both operands
//   dW[j][n] = sum_m D[m][j] X'[m][n],  db[j] = sum_m D[m][j],  X' = relu(xg x_hat + xb) recomputed while staging (or X itself)
// One CTA per SM owns a contiguous range of minibatch rows and the WHOLE 256 x 256 output: two 128 x 256 accumulators fill the
// 512 TMEM columns, so D and X are each read exactly once (the 128 x 128-tile kernel of pb_ppo_tc.cuh reads both twice).  The
// rows are the MMA K dimension; producers transpose 4 x 4 blocks on the fly into the K-major canonical layout as in
// tc::dw1_tc_kernel.  Chunks of 16 rows, W_NST stages of 64 KB.  Per-CTA partial results go to part[cta][256 * 256 | 256]
// and are summed in a fixed order by acf::ac_reduce_segs_kernel.
#define W_BK 16
#define W_NST 3
#define W_HALF (AC_HID * W_BK)                               // floats: one operand half (hi or lo) of one stage, 16 KB
#define W_STAGE_FLOATS (4 * W_HALF)                          // D hi, D lo, X hi, X lo: 64 KB
#define W_SMEM_BYTES (W_NST * W_STAGE_FLOATS * 4 + 256)
#define W_THREADS 288
#define W_PART_LD (AC_HID * AC_HID + AC_HID)
struct Dw256Args {
  const float *D, *X; int M;
  const float *xg, *xb;      // nullptr: X is used as it is
  float *part;               // [gridDim.x][W_PART_LD]
};

static __global__ void __launch_bounds__(W_THREADS, 1) dw256_tc_kernel(Dw256Args a) {
  extern __shared__ __align__(128) unsigned char w_smem[];
  float *tiles = reinterpret_cast<float *>(w_smem);
  uint64_t *empty_bar = reinterpret_cast<uint64_t *>(w_smem + W_NST * W_STAGE_FLOATS * 4);
  uint64_t *full_bar = empty_bar + W_NST;
  uint64_t *done_bar = full_bar + W_NST;
  __shared__ uint32_t tmem_base_sh;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int per = (((a.M + (int)gridDim.x - 1) / (int)gridDim.x) + W_BK - 1) / W_BK * W_BK;   // rows per CTA, whole chunks
  const int m0 = min(a.M, (int)blockIdx.x * per), m1 = min(a.M, m0 + per);
  const int nchunk = (m1 - m0 + W_BK - 1) / W_BK;
  float *part = a.part + (size_t)blockIdx.x * W_PART_LD;
  if (nchunk == 0) {   // an idle CTA still owns its partial
    for (int i = tid; i < W_PART_LD; i += W_THREADS) part[i] = 0.f;
    return;
  }
  if (warp == 0) tmem_alloc(&tmem_base_sh, 512);
  if (tid == 32) {
    for (int s = 0; s < W_NST; s++) { mbar_init(empty_bar + s, 1); mbar_init(full_bar + s, 1); }
    mbar_init(done_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_d = tmem_base_sh;
  auto tile = [&](int s, int which) { return tiles + s * W_STAGE_FLOATS + which * W_HALF; };   // 0 D hi, 1 D lo, 2 X hi, 3 X lo

  if (warp == 8) {
    if (lane == 0) {
      constexpr uint32_t IDESC = idesc_tf32(128, 256);
      for (int ch = 0; ch < nchunk; ch++) {
        const int s = ch % W_NST, use = ch / W_NST;
        mbar_wait(full_bar + s, (uint32_t)(use & 1));
        fence_after_sync();
        const uint32_t dh = smem_u32(tile(s, 0)), dl = smem_u32(tile(s, 1)), xh = smem_u32(tile(s, 2)), xl = smem_u32(tile(s, 3));
#pragma unroll
        for (int ks = 0; ks < W_BK / 8; ks++) {
          const uint64_t dbh = smem_desc(xh + ks * 2 * (256 * 16), 256 * 16, 128), dbl = smem_desc(xl + ks * 2 * (256 * 16), 256 * 16, 128);
#pragma unroll
          for (int jt = 0; jt < 2; jt++) {
            // D half-stage: [jt][c16 4][j 128][4 floats]
            const uint32_t o = jt * (4 * 128 * 16) + ks * 2 * (128 * 16);
            const uint64_t dah = smem_desc(dh + o, 128 * 16, 128), dal = smem_desc(dl + o, 128 * 16, 128);
            const uint32_t dcol = tmem_d + (uint32_t)(jt * 256);
            mma_tf32_ss(dcol, dal, dbh, IDESC, (ch > 0 || ks > 0) ? 1u : 0u);
            mma_tf32_ss(dcol, dah, dbl, IDESC, 1u);
            mma_tf32_ss(dcol, dah, dbh, IDESC, 1u);
          }
        }
        if (ch + W_NST < nchunk) mma_commit(empty_bar + s);
        if (ch == nchunk - 1) mma_commit(done_bar);
      }
    }
    __syncwarp();
  } else {
    // producers: thread = (k-unit u of 4 rows, column block cb of 4 columns) of both operands
    const int cb = tid & 63, u = tid >> 6;
    float4 tg = make_float4(1.f, 1.f, 1.f, 1.f), tb = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool xform = a.xg != nullptr;
    if (xform) {   // scalar loads: parameter blocks are not 16-byte aligned
      const float *pg = a.xg + 4 * cb, *pb = a.xb + 4 * cb;
      tg = make_float4(pg[0], pg[1], pg[2], pg[3]); tb = make_float4(pb[0], pb[1], pb[2], pb[3]);
    }
    auto gload = [&](int ch, float4 (&da)[4], float4 (&xa)[4]) {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int m = m0 + ch * W_BK + 4 * u + i;
        float4 d = make_float4(0.f, 0.f, 0.f, 0.f), x = d;
        if (ch < nchunk && m < m1) {
          d = *reinterpret_cast<const float4 *>(a.D + (size_t)m * AC_HID + 4 * cb);
          x = *reinterpret_cast<const float4 *>(a.X + (size_t)m * AC_HID + 4 * cb);
          if (xform) {
            x.x = fmaxf(0.f, tg.x * x.x + tb.x); x.y = fmaxf(0.f, tg.y * x.y + tb.y);
            x.z = fmaxf(0.f, tg.z * x.z + tb.z); x.w = fmaxf(0.f, tg.w * x.w + tb.w);
          }
        }
        da[i] = d; xa[i] = x;
      }
    };
    // 4 x 4 transpose; column 4 cb + c of the operand becomes the 16-byte k-unit (4 consecutive rows) at `base + col * 4` floats,
    // stores rotated so that a quarter warp writes eight different 16-byte slots
    auto put = [&](float *hi, float *lo, const float4 (&r)[4], bool split_tiles) {
      const float4 v0 = make_float4(r[0].x, r[1].x, r[2].x, r[3].x), v1 = make_float4(r[0].y, r[1].y, r[2].y, r[3].y);
      const float4 v2 = make_float4(r[0].z, r[1].z, r[2].z, r[3].z), v3 = make_float4(r[0].w, r[1].w, r[2].w, r[3].w);
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const int c = (t + (cb >> 1)) & 3, col = 4 * cb + c;
        // D (A operand): two 128-row tiles, each [c16 = u][j][4]; X (B operand): one 256-row tile [c16 = u][n][4]
        const int off = split_tiles ? (col >> 7) * (4 * 128 * 4) + u * (128 * 4) + (col & 127) * 4 : u * (256 * 4) + col * 4;
        store_split(hi, lo, off, sel4(v0, v1, v2, v3, c));
      }
    };
    float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);
    auto stage = [&](int ch, const float4 (&da)[4], const float4 (&xa)[4]) {
      const int s = ch % W_NST, use = ch / W_NST;
#pragma unroll
      for (int i = 0; i < 4; i++) { bsum.x += da[i].x; bsum.y += da[i].y; bsum.z += da[i].z; bsum.w += da[i].w; }
      if (use > 0) mbar_wait(empty_bar + s, (uint32_t)((use - 1) & 1));
      put(tile(s, 0), tile(s, 1), da, true);
      put(tile(s, 2), tile(s, 3), xa, false);
      fence_proxy_async();
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (tid == 0) mbar_arrive(full_bar + s);
    };
    float4 d0[4], x0[4], d1[4], x1[4], d2[4], x2[4];
    gload(0, d0, x0);
    gload(1, d1, x1);
    for (int ch = 0; ch < nchunk; ch += 3) {
      gload(ch + 2, d2, x2);
      stage(ch, d0, x0);
      if (ch + 1 < nchunk) { gload(ch + 3, d0, x0); stage(ch + 1, d1, x1); }
      if (ch + 2 < nchunk) { gload(ch + 4, d1, x1); stage(ch + 2, d2, x2); }
    }
    // epilogue: warp w reads TMEM lane quadrant w & 3 of accumulator jt = w >> 2; thread = output row j, 256 columns
    mbar_wait(done_bar, 0);
    fence_after_sync();
    {
      const int q = warp & 3, jt = warp >> 2, j = jt * 128 + q * 32 + lane;
      float *dst = part + (size_t)j * AC_HID;
#pragma unroll 1
      for (int cbk = 0; cbk < 8; cbk++) {
        float v[32];
        tmem_ld32(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(jt * 256 + cbk * 32), v);
#pragma unroll
        for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4 *>(dst + cbk * 32 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
      }
    }
    float *bs = tiles;   // [4 k-units][256] partial column sums (the operand stages are idle once done_bar has fired)
    *reinterpret_cast<float4 *>(bs + u * AC_HID + 4 * cb) = bsum;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    part[AC_HID * AC_HID + tid] = (bs[tid] + bs[AC_HID + tid]) + (bs[2 * AC_HID + tid] + bs[3 * AC_HID + tid]);
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, 512);
}

}  // namespace actc
