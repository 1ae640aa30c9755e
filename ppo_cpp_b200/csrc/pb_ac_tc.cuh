// pb_ac_tc.cuh — the 256 x 256 products of the ac_ppo encoders on tcgen05 with the LayerNorm that follows them fused in.
//
//   forward   x_hat2, rstd2 = LayerNorm( relu(gamma1 x_hat1 + beta1) W2^T + b2 )          (src/ac_ppo_continuous_action.cpp:150-172)
//   backward  dh1 = d2 W2 (plain epilogue; the LayerNorm backward and the first-layer gradient follow in acf::ac_ln_bwd_w1_kernel)
//
// Both are C[M x 256] = A[M x 256] B^T with M = 262 144 minibatch rows: 268 MB in, 268 MB out, 103 GFLOP as 3xTF32 — the HBM time
// (82 us) and the tensor-pipe time (94 us) are about equal, so the kernel is persistent (one CTA per SM, 128-row tiles taken
// round-robin) and every stage of a tile overlaps the neighbouring tiles:
//   warps 4-11  producers: cp.async the A chunk (128 rows x 16 floats) into a private raw ring G_RAW chunks deep (register prefetch
//               cannot go that deep: a wait on one of the 6 scoreboards also waits for the newest loads that share it), read the
//               own pieces back, apply the input transform (forward: relu(gamma x_hat + beta), so the post-activation matrix is
//               never stored), split into TF32 hi / lo and store both halves in the K-major canonical layout of pb_ppo_tc.cuh;
//               the warps run independently and each arrives on the stage barrier itself
//   warp 13     one lane: TMA bulk copies of the pre-split weight chunk (hi / lo, 16 KB each, L2 resident) into a ring of 4 stages
//   warp 12     one lane: the 6 tcgen05.mma (128 x 256 x 8, a_lo b_hi + a_hi b_lo + a_hi b_hi) of each chunk; the accumulator is
//               one of two 256-column TMEM buffers, committed to acc_full when the tile's 16 chunks are issued
//   warps 0-3   epilogue, one thread per row: reads the finished accumulator from TMEM (the other buffer is being filled),
//               does the whole LayerNorm row computation in registers and writes the result; arrives on acc_empty
// The chunk pipeline (G_NSA / G_NSB shared-memory stages) runs across tile boundaries, so the next tile's A is in flight while the
// current one finishes.
#pragma once
#include "pb_ppo_tc.cuh"
#include "pb_ac_fused.cuh"

namespace actc {
using namespace tc;

#define G_BK 16
#define G_NSA 3                                               // A stages (converted hi / lo halves)
#define G_NSB 4                                               // B stages (TMA)
#define G_RAW 5                                               // raw fp32 ring of the producers (cp.async), in chunks
#define G_NCHUNK (AC_HID / G_BK)                              // 16 chunks per tile
#define G_A_HALF (128 * G_BK)                                 // floats: A hi or lo of one stage, 8 KB
#define G_B_HALF (AC_HID * G_BK)                              // floats: B hi or lo of one stage, 16 KB
#define G_RAW_FLOATS (128 * G_BK)                             // one raw chunk, 8 KB
#define G_THREADS 448
#define G_VEC_FLOATS (3 * AC_HID)                             // gamma | beta of the input transform | bias
#define G_SMEM_BYTES ((G_NSA * 2 * G_A_HALF + G_NSB * 2 * G_B_HALF + G_RAW * G_RAW_FLOATS + G_VEC_FLOATS) * 4 + 256)

enum { G_EPI_LN_FWD = 0, G_EPI_PLAIN = 1 };

struct GemmArgs {
  const float *A; int M;               // [M][256]
  const float *Wc;                     // pre-split B: [hi | lo][chunk 16][c16 4][n 256][4] (w_split256_kernel)
  const float *pg, *pb;                // input transform relu(pg a + pb); nullptr: none
  const float *bias;                   // G_EPI_LN_FWD: x_hat = LN(acc + bias), rstd
  float *out, *rstd;                   // G_EPI_LN_FWD: x_hat; G_EPI_PLAIN: the product itself
  unsigned long long *dbg;             // PB_AC_GEMM_DEBUG: CTA 0 writes where its roles spent their cycles (nullptr: off)
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
// 16-byte asynchronous copy global -> shared (L2 only); bytes = 0 zero-fills.  Completion is tracked per thread in commit groups,
// not in the register scoreboards, so any number of chunks can be in flight.
__device__ __forceinline__ void cp_async16(void *dst_smem, const void *src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int EPI>
static __global__ void __launch_bounds__(G_THREADS, 1) gemm256_tc_kernel(GemmArgs a) {
  extern __shared__ __align__(128) unsigned char g_smem[];
  float *a_tiles = reinterpret_cast<float *>(g_smem);                       // [G_NSA][hi | lo][G_A_HALF]
  float *b_tiles = a_tiles + G_NSA * 2 * G_A_HALF;                          // [G_NSB][hi | lo][G_B_HALF]
  float *raw = b_tiles + G_NSB * 2 * G_B_HALF;                              // [G_RAW][2][256] float4: slot of (piece i, producer thread)
  float *vec = raw + G_RAW * G_RAW_FLOATS;                                  // [0,256) gamma | [256,512) beta | [512,768) bias
  uint64_t *emptya_bar = reinterpret_cast<uint64_t *>(vec + G_VEC_FLOATS);  // [G_NSA] the MMAs that read the A stage have completed
  uint64_t *fulla_bar = emptya_bar + G_NSA;                                 // [G_NSA] A halves stored (one arrive per producer warp)
  uint64_t *emptyb_bar = fulla_bar + G_NSA;                                 // [G_NSB]
  uint64_t *fullb_bar = emptyb_bar + G_NSB;                                 // [G_NSB] B halves landed
  uint64_t *accf_bar = fullb_bar + G_NSB;                                   // [2] accumulator buffer complete
  uint64_t *acce_bar = accf_bar + 2;                                        // [2] accumulator buffer drained by the epilogue
  __shared__ uint32_t tmem_base_sh;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = (a.M + 127) / 128;
  const int my_tiles = (int)blockIdx.x < ntiles ? (ntiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  if (warp == 0) tmem_alloc(&tmem_base_sh, 512);
  if (tid == 32) {
    for (int s = 0; s < G_NSA; s++) { mbar_init(emptya_bar + s, 1); mbar_init(fulla_bar + s, 8); }
    for (int s = 0; s < G_NSB; s++) { mbar_init(emptyb_bar + s, 1); mbar_init(fullb_bar + s, 1); }
    for (int b = 0; b < 2; b++) { mbar_init(accf_bar + b, 1); mbar_init(acce_bar + b, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  for (int i = tid; i < AC_HID; i += G_THREADS) {
    vec[i] = a.pg ? a.pg[i] : 1.f; vec[256 + i] = a.pb ? a.pb[i] : 0.f; vec[512 + i] = a.bias ? a.bias[i] : 0.f;
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_d = tmem_base_sh;
  const int total_chunks = my_tiles * G_NCHUNK;

  if (warp == 13) {
    // ------------------------------------------------ weight loader: a B stage is refilled as soon as the MMAs that read it are done
    if (lane == 0) {
      const float *Wh = a.Wc, *Wl = a.Wc + AC_HID * AC_HID;
      for (int g = 0; g < total_chunks; g++) {
        const int s = g % G_NSB, use = g / G_NSB, c = g % G_NCHUNK;
        if (use > 0) mbar_wait(emptyb_bar + s, (uint32_t)((use - 1) & 1));
        float *dst = b_tiles + s * 2 * G_B_HALF;
        mbar_expect_tx(fullb_bar + s, 2 * G_B_HALF * 4);
        bulk_g2s(dst, Wh + (size_t)c * G_B_HALF, G_B_HALF * 4, fullb_bar + s);
        bulk_g2s(dst + G_B_HALF, Wl + (size_t)c * G_B_HALF, G_B_HALF * 4, fullb_bar + s);
      }
    }
    __syncwarp();
  } else if (warp == 12) {
    // ------------------------------------------------ issuer: 6 MMAs per chunk
    if (lane == 0) {
      constexpr uint32_t IDESC = idesc_tf32(128, 256);
      const bool dbg = a.dbg && blockIdx.x == 0;
      long long t_acce = 0, t_fa = 0, t_fb = 0, t00 = clock64();
      for (int g = 0; g < total_chunks; g++) {
        const int sa = g % G_NSA, usea = g / G_NSA, sb = g % G_NSB, useb = g / G_NSB, c = g % G_NCHUNK, tl = g / G_NCHUNK, buf = tl & 1;
        long long t0 = dbg ? clock64() : 0;
        if (c == 0 && tl >= 2) { mbar_wait(acce_bar + buf, (uint32_t)(((tl >> 1) - 1) & 1)); fence_after_sync(); }
        if (dbg) { const long long t1 = clock64(); t_acce += t1 - t0; t0 = t1; }
        mbar_wait(fulla_bar + sa, (uint32_t)(usea & 1));
        if (dbg) { const long long t1 = clock64(); t_fa += t1 - t0; t0 = t1; }
        mbar_wait(fullb_bar + sb, (uint32_t)(useb & 1));
        if (dbg) { const long long t1 = clock64(); t_fb += t1 - t0; t0 = t1; }
        fence_after_sync();
        const uint32_t ah = smem_u32(a_tiles + sa * 2 * G_A_HALF), al = ah + G_A_HALF * 4;
        const uint32_t bh = smem_u32(b_tiles + sb * 2 * G_B_HALF), bl = bh + G_B_HALF * 4;
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
        if (g + G_NSA < total_chunks) mma_commit(emptya_bar + sa);
        if (g + G_NSB < total_chunks) mma_commit(emptyb_bar + sb);
        if (c == G_NCHUNK - 1) mma_commit(accf_bar + buf);
      }
      if (dbg) { a.dbg[0] = clock64() - t00; a.dbg[1] = t_fa; a.dbg[2] = t_fb; a.dbg[3] = t_acce; }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ------------------------------------------------ producers: cp.async raw ring -> transform -> TF32 split -> A stage.
    // Every thread reads back exactly the two 16-byte pieces it copied itself, so the raw ring needs no barrier; the warps run
    // independently of each other and each arrives on the stage's barrier on its own.
    const int pt = tid - 128, pw = pt >> 5;
    const int rl = lane & 7, c16 = lane >> 3;               // 8 rows x 4 sixteen-byte pieces per warp and item
    const bool xform = a.pg != nullptr;
    float4 *rawv = reinterpret_cast<float4 *>(raw);
    auto issue = [&](int g) {
      if (g < total_chunks) {
        const int tl = g / G_NCHUNK, c = g % G_NCHUNK, row0 = ((int)blockIdx.x + tl * (int)gridDim.x) * 128;
#pragma unroll
        for (int i = 0; i < 2; i++) {
          const int row = row0 + (pw + 8 * i) * 8 + rl;
          const bool live = row < a.M;
          cp_async16(rawv + ((g % G_RAW) * 2 + i) * 256 + pt, a.A + (size_t)(live ? row : 0) * AC_HID + c * G_BK + c16 * 4, live ? 16 : 0);
        }
      }
      cp_async_commit();
    };
    const bool dbg = a.dbg && blockIdx.x == 0 && pt == 0;
    long long t_cp = 0, t_ea = 0, t00 = clock64();
#pragma unroll
    for (int p = 0; p < G_RAW - 1; p++) issue(p);
    for (int g = 0; g < total_chunks; g++) {
      issue(g + G_RAW - 1);
      long long t0 = dbg ? clock64() : 0;
      cp_async_wait<G_RAW - 1>();
      if (dbg) t_cp += clock64() - t0;
      const int s = g % G_NSA, use = g / G_NSA, c = g % G_NCHUNK;
      const int tl = g / G_NCHUNK, row0 = ((int)blockIdx.x + tl * (int)gridDim.x) * 128;
      float4 y[2];
#pragma unroll
      for (int i = 0; i < 2; i++) y[i] = rawv[((g % G_RAW) * 2 + i) * 256 + pt];
      if (xform) {
        const float4 gg = *reinterpret_cast<const float4 *>(vec + c * G_BK + c16 * 4), bb = *reinterpret_cast<const float4 *>(vec + 256 + c * G_BK + c16 * 4);
#pragma unroll
        for (int i = 0; i < 2; i++) {
          const bool live = row0 + (pw + 8 * i) * 8 + rl < a.M;   // rows beyond M stay zero (relu(beta) is not)
          y[i].x = live ? fmaxf(0.f, gg.x * y[i].x + bb.x) : 0.f; y[i].y = live ? fmaxf(0.f, gg.y * y[i].y + bb.y) : 0.f;
          y[i].z = live ? fmaxf(0.f, gg.z * y[i].z + bb.z) : 0.f; y[i].w = live ? fmaxf(0.f, gg.w * y[i].w + bb.w) : 0.f;
        }
      }
      t0 = dbg ? clock64() : 0;
      if (use > 0) mbar_wait(emptya_bar + s, (uint32_t)((use - 1) & 1));
      if (dbg) t_ea += clock64() - t0;
      float *Ah = a_tiles + s * 2 * G_A_HALF, *Al = Ah + G_A_HALF;
#pragma unroll
      for (int i = 0; i < 2; i++) store_split(Ah, Al, c16 * (128 * 4) + ((pw + 8 * i) * 8 + rl) * 4, y[i]);
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(fulla_bar + s);
    }
    if (dbg) { a.dbg[4] = clock64() - t00; a.dbg[5] = t_cp; a.dbg[6] = t_ea; }
  } else {
    // ------------------------------------------------ epilogue: thread = accumulator row (TMEM lane)
    const int m = warp * 32 + lane;
    const bool dbg = a.dbg && blockIdx.x == 0 && tid == 0;
    long long t_af = 0, t00 = clock64();
    for (int tl = 0; tl < my_tiles; tl++) {
      const int buf = tl & 1, row = ((int)blockIdx.x + tl * (int)gridDim.x) * 128 + m;
      const bool live = row < a.M;
      const long long t0 = dbg ? clock64() : 0;
      mbar_wait(accf_bar + buf, (uint32_t)((tl >> 1) & 1));
      if (dbg) t_af += clock64() - t0;
      fence_after_sync();
      const uint32_t tbase = tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)(buf * 256);
      float *dst = a.out + (size_t)row * AC_HID;
      if (EPI == G_EPI_LN_FWD) {
        float s = 0.f;
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32];
          tmem_ld32(tbase + cb * 32, v);
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 bi = *reinterpret_cast<const float4 *>(vec + 512 + cb * 32 + i);
            s += (v[i] + bi.x) + (v[i + 1] + bi.y) + (v[i + 2] + bi.z) + (v[i + 3] + bi.w);
          }
        }
        const float mean = s * (1.f / AC_HID);
        float q = 0.f;
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32];
          tmem_ld32(tbase + cb * 32, v);
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const float4 bi = *reinterpret_cast<const float4 *>(vec + 512 + cb * 32 + i);
            const float d0 = v[i] + bi.x - mean, d1 = v[i + 1] + bi.y - mean, d2 = v[i + 2] + bi.z - mean, d3 = v[i + 3] + bi.w - mean;
            q += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
          }
        }
        const float rs = rsqrtf(q * (1.f / AC_HID) + 1e-5f);
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
#pragma unroll 1
        for (int cb = 0; cb < 8; cb++) {
          float v[32];
          tmem_ld32(tbase + cb * 32, v);
          if (live) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4 *>(dst + cb * 32 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
        }
      }
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(acce_bar + buf);
    }
    if (dbg) { a.dbg[7] = clock64() - t00; a.dbg[8] = t_af; a.dbg[9] = (unsigned long long)my_tiles; }
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
#define W_NST 2
#define W_RAW 3                                              // raw fp32 ring of the producers (cp.async), in chunks of 32 KB
#define W_HALF (AC_HID * W_BK)                               // floats: one operand half (hi or lo) of one stage, 16 KB
#define W_STAGE_FLOATS (4 * W_HALF)                          // D hi, D lo, X hi, X lo: 64 KB
#define W_RAW_FLOATS (2 * AC_HID * W_BK)                     // D and X rows of one chunk
#define W_SMEM_BYTES ((W_NST * W_STAGE_FLOATS + W_RAW * W_RAW_FLOATS) * 4 + 256)
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
  float4 *rawv = reinterpret_cast<float4 *>(tiles + W_NST * W_STAGE_FLOATS);   // [W_RAW][8 pieces][256 threads]
  uint64_t *empty_bar = reinterpret_cast<uint64_t *>(w_smem + (W_NST * W_STAGE_FLOATS + W_RAW * W_RAW_FLOATS) * 4);
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
    for (int s = 0; s < W_NST; s++) { mbar_init(empty_bar + s, 1); mbar_init(full_bar + s, 8); }
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
    // cp.async raw ring: pieces 0-3 = this thread's 4 rows of D, 4-7 = of X; a thread reads back only what it copied itself
    auto issue = [&](int ch) {
      if (ch < nchunk) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const int m = m0 + ch * W_BK + 4 * u + i;
          const bool live = m < m1;
          const size_t o = (size_t)(live ? m : m0) * AC_HID + 4 * cb;
          cp_async16(rawv + ((ch % W_RAW) * 8 + i) * 256 + tid, a.D + o, live ? 16 : 0);
          cp_async16(rawv + ((ch % W_RAW) * 8 + 4 + i) * 256 + tid, a.X + o, live ? 16 : 0);
        }
      }
      cp_async_commit();
    };
    auto fetch = [&](int ch, float4 (&da)[4], float4 (&xa)[4]) {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        da[i] = rawv[((ch % W_RAW) * 8 + i) * 256 + tid];
        float4 x = rawv[((ch % W_RAW) * 8 + 4 + i) * 256 + tid];
        if (xform && m0 + ch * W_BK + 4 * u + i < m1) {   // rows beyond the range stay zero (relu(beta) is not)
          x.x = fmaxf(0.f, tg.x * x.x + tb.x); x.y = fmaxf(0.f, tg.y * x.y + tb.y);
          x.z = fmaxf(0.f, tg.z * x.z + tb.z); x.w = fmaxf(0.f, tg.w * x.w + tb.w);
        }
        xa[i] = x;
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
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar + s);
    };
#pragma unroll
    for (int p = 0; p < W_RAW - 1; p++) issue(p);
    for (int ch = 0; ch < nchunk; ch++) {
      issue(ch + W_RAW - 1);
      cp_async_wait<W_RAW - 1>();
      float4 da[4], xa[4];
      fetch(ch, da, xa);
      stage(ch, da, xa);
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
