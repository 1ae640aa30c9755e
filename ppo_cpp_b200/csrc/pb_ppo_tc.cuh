// pb_ppo_tc.cuh — first-layer products of the PPO update on the 5th-generation tensor cores (tcgen05 + TMEM), fp32-accurate.
//
// The update's two large GEMMs are the first-layer forward  Hpre[mb x 128] = X[mb x O] W1^T  (both nets side by side) and
// the first-layer weight gradient dW1[128 x O] = D1^T X.  Both are computed as 3xTF32: every fp32 operand is split into
// hi = tf32(x) and lo = tf32(x - hi) while it is staged into shared memory, and a_lo*b_hi + a_hi*b_lo + a_hi*b_hi is
// accumulated in fp32 in TMEM by three tcgen05.mma.kind::tf32 instructions per k-step (the dropped a_lo*b_lo term is
// 2^-22 relative), so the result matches the reference's fp32 arithmetic to rounding (tests/test_gpu_ppo_parity.py).
//
// Operand staging: K-major, no-swizzle canonical layout — core matrices of 8 rows x 16 bytes stored contiguously; a tile of
// R rows keeps, for every 16-byte k-chunk c, its R rows back to back: byte(r, c) = c * (R*16) + r*16, i.e. the stride
// between 8-row groups (SBO) is 128 B and between the two k-chunks of one MMA (LBO) is R*16 B.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "pb_ppo_common.cuh"

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
// Bounded spin: a tensor-core pipeline that never signals must not hang the GPU box — trap instead.
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  for (uint32_t spins = 0;; spins++) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    if (ok) return;
    if (spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t cols) {  // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(cols));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {  // the same warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols));
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the tensor core's async-proxy reads
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// K-major, SWIZZLE_NONE shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address, LBO, SBO in 16-byte
// units, descriptor version 1 (Blackwell) in bits [46,48).
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): fp32 accumulate, TF32 x TF32, both operands K-major, M x N tile.
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]^T, one k-step of 8 TF32; issued by ONE thread
__device__ __forceinline__ void mma_tf32_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(0u) : "memory");
}
// all previously issued MMAs of this thread arrive on the mbarrier when they have completed
__device__ __forceinline__ void mma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 TMEM lanes (this warp's quadrant) x 16 consecutive fp32 columns -> 16 registers per thread (thread t = lane t)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                 "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void split_tf32(float x, float &hi, float &lo) {
  uint32_t h, l;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  const float r = x - __uint_as_float(h);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(r));
  hi = __uint_as_float(h); lo = __uint_as_float(l);
}
__device__ __forceinline__ void store_split(float *hi_base, float *lo_base, int off_floats, float4 v) {
  float4 h, l;
  split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y); split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
  *reinterpret_cast<float4 *>(hi_base + off_floats) = h;
  *reinterpret_cast<float4 *>(lo_base + off_floats) = l;
}

// TMA bulk copy (1-D) global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ------------------------------------------------------------------ forward: Hpre[ks][m][n] = sum_{k in range ks} X[m][k] W1[n][k]
// One CTA = 128 minibatch rows x 128 hidden units (critic | actor) x one of L1_KSPLIT contiguous ranges of k-chunks
// (blockIdx.y), so a minibatch of 8192 rows gives 128 CTAs for the 148 SMs, no operand is read twice and the serial chain
// of tcgen05.mma issues (~80 cycles each, 36 per 96 columns of k) is halved; the consumer adds the partial products.
// K is streamed in chunks of 32 floats (4 MMA k-steps) through a ring of L1_NST shared-memory stages, warp-specialised:
//   warps 0-7 (producers): every thread loads its 4 pieces (row, 16-byte k-chunk) of the X chunk L1_PF chunks ahead into
//     registers, splits them into TF32 hi / lo and stores both halves in the canonical layout; after a producer-only named
//     barrier one thread arrives on the stage's "A full" barrier.
//   warp 8, one lane (issuer): requests the stage's weight halves with two TMA bulk copies one chunk ahead (w1_split_kernel
//     keeps hi / lo of W1 in global memory in exactly the stage layout, refreshed once per optimiser step), waits for A and
//     B, issues the 12 tcgen05.mma of the chunk (accumulator: 128 TMEM columns) and commits them to the stage's "empty"
//     barrier, which producers and the next TMA of that stage wait on.
#define L1_BK 32
#define L1_NST 3
#define L1_PF 4                                               // register prefetch distance for X, in chunks
#define L1_A_FLOATS (128 * L1_BK)                             // X half (hi or lo) of one stage: 16 KB
#define L1_B_FLOATS (128 * L1_BK)                             // W half of one stage: 16 KB
#define L1_STAGE_FLOATS (2 * L1_A_FLOATS + 2 * L1_B_FLOATS)   // 64 KB
#define L1_KSPLIT 2                                           // CTAs per row tile, each owning a contiguous range of k-chunks
#define L1_SMEM_BYTES (L1_NST * L1_STAGE_FLOATS * 4 + 128)
#define L1_THREADS 288
struct L1Args {
  const float *Xp; const int *cursor; int mb, O;
  const float *Wc;                     // [hi|lo 2][nchunk][8][128][4]: pre-split W1 (critic rows, actor rows) in stage layout
  float *Hpre[L1_KSPLIT];              // [mb][ldc] partial products of each k range
  int debug;                           // 1: block 0 prints where the issuer spends its cycles
  // generalisation for the ac_ppo 256-wide layers (pb_ac_trainer.cu): blockIdx.z selects a 128-column tile of the output,
  // whose pre-split weights start wc_tile_stride floats further on; the output has leading dimension ldc (0 = 128) and an
  // optional bias per column.  The ppo path leaves all of them zero / null and launches with gridDim.z = 1.
  int ldc; size_t wc_tile_stride; const float *bias;
};

// W[n][k] (torch layout, leading dimension ldw; transposed != 0: the operand is W^T, i.e. element (n, k) = W[k][n]) -> hi / lo
// halves in the stage layout, one block of 2 x nchunk x 4096 floats per 128-row tile of n; k >= K and n >= N are zero
static __global__ void w_split_generic_kernel(const float *__restrict__ W, int ldw, int transposed, int N, int K, int nchunk, int ntiles,
                                              float *__restrict__ Wc) {
  const int per_tile = nchunk * L1_B_FLOATS;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ntiles * per_tile) return;
  const int z = i / per_tile, rem = i - z * per_tile;
  const int e = rem & 3, n = (rem >> 2) & 127, c16 = (rem >> 9) & 7, chunk = rem >> 12, k = chunk * L1_BK + c16 * 4 + e, nn = z * 128 + n;
  float hi = 0.f, lo = 0.f;
  if (k < K && nn < N) split_tf32(transposed ? W[(size_t)k * ldw + nn] : W[(size_t)nn * ldw + k], hi, lo);
  Wc[(size_t)z * 2 * per_tile + rem] = hi; Wc[(size_t)z * 2 * per_tile + per_tile + rem] = lo;
}

// W1 (torch layout, one [64][O] block per net) -> hi / lo halves in the stage layout; k >= O is zero
static __global__ void w1_split_kernel(const float *__restrict__ params, int cW1, int aW1, int O, int nchunk, float *__restrict__ Wc) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, total = nchunk * L1_B_FLOATS;
  if (i >= total) return;
  // i = ((chunk * 8 + c16) * 128 + n) * 4 + e
  const int e = i & 3, n = (i >> 2) & 127, c16 = (i >> 9) & 7, chunk = i >> 12, k = chunk * L1_BK + c16 * 4 + e;
  float hi = 0.f, lo = 0.f;
  if (k < O) split_tf32(params[(n < 64 ? cW1 + n * O : aW1 + (n - 64) * O) + k], hi, lo);
  Wc[i] = hi; Wc[(size_t)total + i] = lo;
}

__device__ __forceinline__ void l1_item(int q, int &r, int &c) {  // 8 rows x 4 chunks per warp: full 32 B sectors from global
  r = (q & 7) | ((q >> 6) << 3);
  c = ((q >> 3) & 3) | (((q >> 5) & 1) << 2);
}

static __global__ void __launch_bounds__(L1_THREADS, 1) l1_forward_tc_kernel(L1Args a) {
  extern __shared__ __align__(128) unsigned char l1_smem[];
  float *tiles = reinterpret_cast<float *>(l1_smem);
  uint64_t *empty_bar = reinterpret_cast<uint64_t *>(l1_smem + L1_NST * L1_STAGE_FLOATS * 4);  // [L1_NST] MMAs of the stage done
  uint64_t *fulla_bar = empty_bar + L1_NST;                                                    // [L1_NST] X halves stored
  uint64_t *fullb_bar = fulla_bar + L1_NST;                                                    // [L1_NST] W halves landed
  uint64_t *done_bar = fullb_bar + L1_NST;
  __shared__ uint32_t tmem_base_sh;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int O = a.O, nchunk_all = (O + L1_BK - 1) / L1_BK, cpk = (nchunk_all + gridDim.y - 1) / gridDim.y;
  const int chunk0 = blockIdx.y * cpk, nchunk = max(0, min(nchunk_all, chunk0 + cpk) - chunk0);   // host guarantees nchunk > 0
  const size_t row0 = (size_t)(a.cursor ? *a.cursor : 0) * a.mb + (size_t)blockIdx.x * 128;
  const int rows = min(128, a.mb - blockIdx.x * 128);
  const bool aligned = (O & 3) == 0;

  if (warp == 0) tmem_alloc(&tmem_base_sh, 128);
  if (tid == 32) {
    for (int s = 0; s < L1_NST; s++) { mbar_init(empty_bar + s, 1); mbar_init(fulla_bar + s, 1); mbar_init(fullb_bar + s, 1); }
    mbar_init(done_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_d = tmem_base_sh;
  auto tile = [&](int s, int which) {  // 0 A hi, 1 A lo, 2 B hi, 3 B lo
    return tiles + s * L1_STAGE_FLOATS + (which < 2 ? which * L1_A_FLOATS : 2 * L1_A_FLOATS + (which - 2) * L1_B_FLOATS);
  };

  if (warp == 8) {
    // ---------------- issuer
    if (lane == 0) {
      const float *Wh = a.Wc + (size_t)blockIdx.z * a.wc_tile_stride + (size_t)chunk0 * L1_B_FLOATS, *Wl = Wh + (size_t)nchunk_all * L1_B_FLOATS;
      auto issue_b = [&](int ch) {
        const int s = ch % L1_NST;
        mbar_expect_tx(fullb_bar + s, 2 * L1_B_FLOATS * 4);
        bulk_g2s(tile(s, 2), Wh + (size_t)ch * L1_B_FLOATS, L1_B_FLOATS * 4, fullb_bar + s);
        bulk_g2s(tile(s, 3), Wl + (size_t)ch * L1_B_FLOATS, L1_B_FLOATS * 4, fullb_bar + s);
      };
      constexpr uint32_t IDESC = idesc_tf32(128, 128);
      long long t_wait = 0, t_issue = 0;
      const bool dbg = a.debug && blockIdx.x == 0 && blockIdx.y == 0;
      const long long t_start = clock64();
      issue_b(0);
      for (int ch = 0; ch < nchunk; ch++) {
        const int s = ch % L1_NST, use = ch / L1_NST;
        long long t0 = clock64();
        if (ch + 1 < nchunk) {  // weights of the next chunk: its stage was last read by chunk ch + 1 - L1_NST
          const int s1 = (ch + 1) % L1_NST, use1 = (ch + 1) / L1_NST;
          if (use1 > 0) mbar_wait(empty_bar + s1, (uint32_t)((use1 - 1) & 1));
          issue_b(ch + 1);
        }
        mbar_wait(fulla_bar + s, (uint32_t)(use & 1));
        mbar_wait(fullb_bar + s, (uint32_t)(use & 1));
        fence_after_sync();
        if (dbg) { long long t1 = clock64(); t_wait += t1 - t0; t0 = t1; }
        const uint32_t ah = smem_u32(tile(s, 0)), al = smem_u32(tile(s, 1)), bh = smem_u32(tile(s, 2)), bl = smem_u32(tile(s, 3));
#pragma unroll
        for (int ks = 0; ks < L1_BK / 8; ks++) {   // two 16-byte k-chunks per k-step
          const uint64_t dah = smem_desc(ah + ks * 2 * (128 * 16), 128 * 16, 128), dal = smem_desc(al + ks * 2 * (128 * 16), 128 * 16, 128);
          const uint64_t dbh = smem_desc(bh + ks * 2 * (128 * 16), 128 * 16, 128), dbl = smem_desc(bl + ks * 2 * (128 * 16), 128 * 16, 128);
          mma_tf32_ss(tmem_d, dal, dbh, IDESC, (ch > 0 || ks > 0) ? 1u : 0u);
          mma_tf32_ss(tmem_d, dah, dbl, IDESC, 1u);
          mma_tf32_ss(tmem_d, dah, dbh, IDESC, 1u);
        }
        // every commit must be waited for before the CTA exits (the arrive is an asynchronous write to this CTA's shared
        // memory): the stage's barrier only when a later chunk reuses the stage, the final barrier for the epilogue
        if (ch + L1_NST < nchunk) mma_commit(empty_bar + s);
        if (ch == nchunk - 1) mma_commit(done_bar);
        if (dbg) { long long t1 = clock64(); t_issue += t1 - t0; }
      }
      if (dbg) printf("l1 tc issuer (cycles): loop %lld [waits %lld issue %lld]\n", clock64() - t_start, t_wait, t_issue);
    }
    __syncwarp();
  } else {
    // ---------------- producers: this thread's 4 pieces of a chunk of X
    int ir[4], ic[4];
#pragma unroll
    for (int i = 0; i < 4; i++) l1_item(tid + 256 * i, ir[i], ic[i]);
    auto gload = [&](int ch, float4 (&xa)[4]) {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int k = (chunk0 + ch) * L1_BK + ic[i] * 4, r = ir[i];
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ch < nchunk && r < rows) {
          const float *p = a.Xp + (row0 + r) * O + k;
          if (aligned) { if (k < O) v = *reinterpret_cast<const float4 *>(p); }
          else { if (k < O) v.x = p[0]; if (k + 1 < O) v.y = p[1]; if (k + 2 < O) v.z = p[2]; if (k + 3 < O) v.w = p[3]; }
        }
        xa[i] = v;
      }
    };
    long long p_empty = 0, p_store = 0, p_bar = 0;
    const bool pdbg = a.debug && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0;
    const long long p_start = clock64();
    auto stage = [&](int ch, const float4 (&xa)[4]) {
      const int s = ch % L1_NST, use = ch / L1_NST;
      long long t0 = clock64();
      if (use > 0) mbar_wait(empty_bar + s, (uint32_t)((use - 1) & 1));  // the MMAs that read this stage have completed
      if (pdbg) { long long t1 = clock64(); p_empty += t1 - t0; t0 = t1; }
      float *Ah = tile(s, 0), *Al = tile(s, 1);
#pragma unroll
      for (int i = 0; i < 4; i++) store_split(Ah, Al, ic[i] * (128 * 4) + ir[i] * 4, xa[i]);   // chunk c -> 128 rows x 16 B
      fence_proxy_async();
      if (pdbg) { long long t1 = clock64(); p_store += t1 - t0; t0 = t1; }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (tid == 0) mbar_arrive(fulla_bar + s);
      if (pdbg) { long long t1 = clock64(); p_bar += t1 - t0; }
    };
    // register prefetch L1_PF chunks ahead; the chunk loop is unrolled by L1_PF so the register sets are statically named
    float4 xr[L1_PF][4];
#pragma unroll
    for (int p = 0; p < L1_PF - 1; p++) gload(p, xr[p]);
    for (int ch0 = 0; ch0 < nchunk; ch0 += L1_PF) {
#pragma unroll
      for (int p = 0; p < L1_PF; p++) {
        const int ch = ch0 + p;
        if (ch < nchunk) {
          gload(ch + L1_PF - 1, xr[(p + L1_PF - 1) % L1_PF]);
          stage(ch, xr[p]);
        }
      }
    }
    // epilogue: accumulator rows = TMEM lanes; warp w reads lane quadrant w % 4, column half w / 4 (32 columns)
    if (pdbg) printf("l1 tc producer (cycles): loop %lld [empty-wait %lld load-wait+split+store %lld bar %lld]\n", clock64() - p_start, p_empty, p_store, p_bar);
    mbar_wait(done_bar, 0);
    fence_after_sync();
    const int q = warp & 3, half = warp >> 2, m = q * 32 + lane;
    const int ldc = a.ldc ? a.ldc : 128, col0 = blockIdx.z * 128 + half * 64;
    float *dst = a.Hpre[blockIdx.y] + ((size_t)blockIdx.x * 128 + m) * ldc + col0;
#pragma unroll
    for (int cb = 0; cb < 4; cb++) {
      float v[16];
      tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * 64 + cb * 16), v);
      if (a.bias) {
#pragma unroll
        for (int i = 0; i < 16; i++) v[i] += a.bias[col0 + cb * 16 + i];
      }
      if (m < rows) {
#pragma unroll
        for (int i = 0; i < 4; i++) *reinterpret_cast<float4 *>(dst + cb * 16 + i * 4) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, 128);
}

// ------------------------------------------------------------------ weight gradient: dW1[j][n] = sum_m D1[m][j] X[m][n]
// One CTA = the 128 first-layer units of both nets x 128 observation columns (blockIdx.x) x one split of the minibatch rows
// (blockIdx.y, the same split -> slab scheme as wgrad_kernel, so grad_reduce_kernel is unchanged).  The minibatch rows are
// the MMA K dimension; both operands are stored row-major over m in global memory, i.e. MN-major for this product, so the
// producers transpose on the fly: a thread loads a 4 (rows m) x 4 (columns) block of D1 and of X with four 16-byte loads
// each, and the four column vectors (4 consecutive m of one unit / one observation column) are exactly the 16-byte
// k-units of the K-major canonical layout.  The four stores of a thread are rotated by its block index so that a quarter
// warp always writes eight different 16-byte slots.  Register prefetch two chunks (of 32 rows) ahead; issuer and barriers
// as in the forward kernel.  Epilogue: TMEM -> shared memory tile -> coalesced rows of the split's slab.
#define DW_BK 32
#define DW_NST 3
#define DW_HALF_FLOATS (128 * DW_BK)                          // one operand half of one stage: 16 KB
#define DW_STAGE_FLOATS (4 * DW_HALF_FLOATS)                  // A hi, A lo, B hi, B lo: 64 KB
#define DW_SMEM_BYTES (DW_NST * DW_STAGE_FLOATS * 4 + 128)
#define DW_THREADS 288
#define DW_LDT 132                                            // epilogue tile row stride (floats)
struct DwArgs {
  const float *Xp, *D1; const int *cursor; int mb, O, splits;
  int cW1, aW1;                        // offsets of the two [64][O] blocks inside a slab
  float *slabs; size_t slab_stride;
  // generalisation for the ac_ppo hidden layers (pb_ac_trainer.cu), all zero on the ppo path: generic != 0 -> D1 has row stride
  // ldd and blockIdx.z selects its 128-column tile, the output tile goes to slab + w_off + (j0 + j) * ldw + n0 + n, and the
  // column sums of the D tile (the bias gradient) to slab + b_off + j0 + j (b_off < 0: none; written by the blockIdx.x == 0 CTA)
  int generic, ldd, w_off, ldw, b_off;
  // optional transform of the X operand while it is staged: x := relu(xg[n] x + xb[n]) (ac_ppo fused mode: the post-activation
  // matrix is recomputed from x_hat instead of being stored); nullptr: none
  const float *xg, *xb;
};

__device__ __forceinline__ float4 sel4(const float4 &v0, const float4 &v1, const float4 &v2, const float4 &v3, int c) {
  float4 r = v0;
  if (c == 1) r = v1;
  if (c == 2) r = v2;
  if (c == 3) r = v3;
  return r;
}

static __global__ void __launch_bounds__(DW_THREADS, 1) dw1_tc_kernel(DwArgs a) {
  extern __shared__ __align__(128) unsigned char dw_smem[];
  float *tiles = reinterpret_cast<float *>(dw_smem);
  uint64_t *empty_bar = reinterpret_cast<uint64_t *>(dw_smem + DW_NST * DW_STAGE_FLOATS * 4);  // [DW_NST]
  uint64_t *full_bar = empty_bar + DW_NST;                                                     // [DW_NST]
  uint64_t *done_bar = full_bar + DW_NST;
  __shared__ uint32_t tmem_base_sh;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int O = a.O, n0 = blockIdx.x * 128, z = blockIdx.y, j0 = a.generic ? blockIdx.z * 128 : 0, ldd = a.generic ? a.ldd : 128;
  const bool want_bias = a.generic && a.b_off >= 0 && blockIdx.x == 0;
  const int per = (a.mb + a.splits - 1) / a.splits, m0 = z * per, m1 = min(a.mb, m0 + per);
  const int nchunk = m1 > m0 ? (m1 - m0 + DW_BK - 1) / DW_BK : 0;
  float *slab = a.slabs + (size_t)z * a.slab_stride;
  auto out_row = [&](int j) {
    if (a.generic) return slab + a.w_off + (size_t)(j0 + j) * a.ldw + n0;
    return slab + (j < 64 ? a.cW1 + (size_t)j * O : a.aW1 + (size_t)(j - 64) * O) + n0;
  };
  if (nchunk == 0) {  // an empty split still owns its slab entries
    if (warp < 8) for (int j = warp; j < 128; j += 8) for (int n = lane; n < 128; n += 32) if (n0 + n < O) out_row(j)[n] = 0.f;
    if (want_bias && tid < 128) slab[a.b_off + j0 + tid] = 0.f;
    return;
  }
  const size_t xrow0 = (size_t)(a.cursor ? *a.cursor : 0) * a.mb;
  const bool aligned = (O & 3) == 0;

  if (warp == 0) tmem_alloc(&tmem_base_sh, 128);
  if (tid == 32) {
    for (int s = 0; s < DW_NST; s++) { mbar_init(empty_bar + s, 1); mbar_init(full_bar + s, 1); }
    mbar_init(done_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_d = tmem_base_sh;
  auto tile = [&](int s, int which) { return tiles + s * DW_STAGE_FLOATS + which * DW_HALF_FLOATS; };  // 0 A hi, 1 A lo, 2 B hi, 3 B lo

  if (warp == 8) {
    if (lane == 0) {
      constexpr uint32_t IDESC = idesc_tf32(128, 128);
      for (int ch = 0; ch < nchunk; ch++) {
        const int s = ch % DW_NST, use = ch / DW_NST;
        mbar_wait(full_bar + s, (uint32_t)(use & 1));
        fence_after_sync();
        const uint32_t ah = smem_u32(tile(s, 0)), al = smem_u32(tile(s, 1)), bh = smem_u32(tile(s, 2)), bl = smem_u32(tile(s, 3));
#pragma unroll
        for (int ks = 0; ks < DW_BK / 8; ks++) {
          const uint32_t o = ks * 2 * (128 * 16);
          const uint64_t dah = smem_desc(ah + o, 128 * 16, 128), dal = smem_desc(al + o, 128 * 16, 128);
          const uint64_t dbh = smem_desc(bh + o, 128 * 16, 128), dbl = smem_desc(bl + o, 128 * 16, 128);
          mma_tf32_ss(tmem_d, dal, dbh, IDESC, (ch > 0 || ks > 0) ? 1u : 0u);
          mma_tf32_ss(tmem_d, dah, dbl, IDESC, 1u);
          mma_tf32_ss(tmem_d, dah, dbh, IDESC, 1u);
        }
        if (ch + DW_NST < nchunk) mma_commit(empty_bar + s);   // waited for by the producers of chunk ch + DW_NST
        if (ch == nchunk - 1) mma_commit(done_bar);            // waited for by the epilogue
      }
    }
    __syncwarp();
  } else {
    // producers: block (k-unit u = 4 rows, column block jb = 4 columns) of both operands
    const int jb = tid & 31, u = tid >> 5;
    float4 tg = make_float4(1.f, 1.f, 1.f, 1.f), tb = make_float4(0.f, 0.f, 0.f, 0.f);
    const bool xform = a.xg != nullptr;
    if (xform && n0 + 4 * jb + 3 < O) {   // scalar loads: parameter blocks are not 16-byte aligned
      const float *pg = a.xg + n0 + 4 * jb, *pb = a.xb + n0 + 4 * jb;
      tg = make_float4(pg[0], pg[1], pg[2], pg[3]); tb = make_float4(pb[0], pb[1], pb[2], pb[3]);
    }
    auto gload = [&](int ch, float4 (&da)[4], float4 (&xb)[4]) {
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int m = m0 + ch * DW_BK + 4 * u + i;
        float4 d = make_float4(0.f, 0.f, 0.f, 0.f), x = d;
        if (ch < nchunk && m < m1) {
          d = *reinterpret_cast<const float4 *>(a.D1 + (size_t)m * ldd + j0 + 4 * jb);
          const int n = n0 + 4 * jb;
          const float *p = a.Xp + (xrow0 + m) * O + n;
          if (aligned) { if (n < O) x = *reinterpret_cast<const float4 *>(p); }
          else { if (n < O) x.x = p[0]; if (n + 1 < O) x.y = p[1]; if (n + 2 < O) x.z = p[2]; if (n + 3 < O) x.w = p[3]; }
          if (xform) {
            x.x = fmaxf(0.f, tg.x * x.x + tb.x); x.y = fmaxf(0.f, tg.y * x.y + tb.y);
            x.z = fmaxf(0.f, tg.z * x.z + tb.z); x.w = fmaxf(0.f, tg.w * x.w + tb.w);
          }
        }
        da[i] = d; xb[i] = x;
      }
    };
    auto put = [&](float *hi, float *lo, const float4 (&r)[4]) {   // 4 x 4 transpose, rotated 16-byte stores
      const float4 v0 = make_float4(r[0].x, r[1].x, r[2].x, r[3].x), v1 = make_float4(r[0].y, r[1].y, r[2].y, r[3].y);
      const float4 v2 = make_float4(r[0].z, r[1].z, r[2].z, r[3].z), v3 = make_float4(r[0].w, r[1].w, r[2].w, r[3].w);
#pragma unroll
      for (int t = 0; t < 4; t++) {
        const int c = (t + (jb >> 1)) & 3;
        store_split(hi, lo, u * (128 * 4) + (4 * jb + c) * 4, sel4(v0, v1, v2, v3, c));   // unit u -> 128 rows x 16 B
      }
    };
    float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);   // column sums of this thread's 4 D columns over its rows (bias gradient)
    auto stage = [&](int ch, const float4 (&da)[4], const float4 (&xb)[4]) {
      const int s = ch % DW_NST, use = ch / DW_NST;
      if (want_bias) {
#pragma unroll
        for (int i = 0; i < 4; i++) { bsum.x += da[i].x; bsum.y += da[i].y; bsum.z += da[i].z; bsum.w += da[i].w; }
      }
      if (use > 0) mbar_wait(empty_bar + s, (uint32_t)((use - 1) & 1));
      put(tile(s, 0), tile(s, 1), da);
      put(tile(s, 2), tile(s, 3), xb);
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
    // epilogue: TMEM (lane = unit j) -> shared tile [j][n] (the operand stages are free once done_bar has fired) -> slab
    mbar_wait(done_bar, 0);
    fence_after_sync();
    float *ts = tiles;
    {
      const int q = warp & 3, half = warp >> 2, j = q * 32 + lane;
#pragma unroll
      for (int cb = 0; cb < 4; cb++) {
        float v[16];
        tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * 64 + cb * 16), v);
#pragma unroll
        for (int i = 0; i < 4; i++)
          *reinterpret_cast<float4 *>(ts + j * DW_LDT + half * 64 + cb * 16 + i * 4) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
      }
    }
    float *bs = ts + 128 * DW_LDT;   // [8 row groups][128 columns] partial column sums, behind the output tile
    if (want_bias) *reinterpret_cast<float4 *>(bs + u * 128 + 4 * jb) = bsum;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    for (int j = warp; j < 128; j += 8) {
      float *dst = out_row(j);
#pragma unroll
      for (int t = 0; t < 4; t++) { const int n = lane + 32 * t; if (n0 + n < O) dst[n] = ts[j * DW_LDT + n]; }
    }
    if (want_bias && tid < 128) {   // fixed summation order over the 8 row groups
      float acc = 0.f;
#pragma unroll
      for (int g = 0; g < 8; g++) acc += bs[g * 128 + tid];
      slab[a.b_off + j0 + tid] = acc;
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_d, 128);
}

}  // namespace tc
