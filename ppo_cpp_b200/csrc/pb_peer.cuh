// pb_peer.cuh — data-parallel gradient average over NVLink peer memory, fused with the squared-norm partials that
// clip_grad_norm_ needs.  Replaces "ncclAllReduce(avg) + sqnorm_partials_kernel" between grad_reduce_kernel and
// adam2_kernel (the exchange step of src/ac_ppo_continuous_action.cpp:878-885 / src/distributed.cpp:117-133): the flat
// gradient is 0.04-0.23 MB, so the collective is pure latency and one kernel that reads every peer's buffer directly
// (one-shot all-reduce: world x P floats over NVSwitch) beats a ring/tree that has to be launched, synchronised and
// followed by a second kernel.
//
// Three forms exist, bit-identical in their results: the LL form (grad_reduce_peer_ll_kernel / peer_ll_exchange_kernel, the
// default since round 2: {value, seq} words pushed into every peer's receive area, no flags, fences or remote reads), the
// two-kernel form described next (PB_PEER_LL=0) and the one-kernel-with-flags form (PB_PEER_FUSED=1).
//
// Memory (one cudaMalloc per rank, exported with cudaIpcGetMemHandle, opened by every peer):
//   float gbuf[2][Ppad]      double-buffered flat gradient written by grad_reduce_kernel (slot = seq & 1)
//   uint  flags[PB_PEER_MAX][NB] flags[p][b] = highest sequence number rank p has published for slice b (256 floats) of
//                                its gradient (written by rank p, remotely); the two-kernel variant uses b = 0 only
//   uint  bye[PB_PEER_MAX]       the same for the teardown barrier
//   u64   ll[2][world][Ppad]     LL receive area: ll[slot][source rank][i] = {sequence number : value bits} of element i
// Protocol for optimiser step `seq` (1, 2, ...; kept in device memory so the launch arguments never change):
//   1. grad_reduce_kernel has written gbuf[(seq-1)&1] on every rank (stream order).
//   2. peer_avg_kernel, block 0: thread p stores `seq` into rank p's flags[my rank] with release.sys (push: waiting
//      ranks poll their own memory, nothing spins over NVLink).
//   3. every block: threads p < world spin (acquire.sys) until flags[p] >= seq, then the block sums its 256-float slice
//      over the peers in rank order (identical summation order on every rank -> bit-identical parameters), scales by
//      1/world, writes the local averaged gradient and its partial sum of squares.
//   4. adam2_kernel advances seq.
// Variant (PB_PEER_FUSED=1): grad_reduce_peer_kernel does steps 1-3 in ONE launch, slice by slice — block b reduces its
// 256-float slice of the split slabs into the exchange buffer, publishes flags[.][b] to every peer and waits only for the
// peers' slice b, so there is no grid-wide dependency.  Blocks wait only on same-index blocks of the peers and every GPU
// dispatches blocks in index order, so the waits cannot deadlock even if a grid were not fully resident.  Parity-green
// (bit-identical to the two-kernel path), but MEASURED SLOWER at 2 GPUs — update 35.0-35.8 ms against 31.2-32.5 ms
// (profiles/r01_bench_g2_fused_ab.txt): 226 blocks each pay a system-scope release (MEMBAR.SYS) and a cross-GPU round
// trip, where the two-kernel path pays them once per step in block 0 while the launch boundary does the grid-wide join
// for free.  Kept as an option; the default is the two-kernel path.
// A slot is overwritten at step seq+2, which a rank reaches only after every peer has published seq+1, i.e. has left the
// kernel that read the slot at step seq — two slots are enough.  Waits are bounded (globaltimer) and trap rather than
// hang the GPU.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

#define PB_PEER_MAX 16
#define PB_PEER_TIMEOUT_NS 60000000000ull   // 60 s: a rank that has not arrived by then is dead

struct PeerArgs {
  const float *gbuf[PB_PEER_MAX];   // peers' gbuf bases (own entry = local pointer)
  unsigned *flags[PB_PEER_MAX];     // peers' flag arrays
  unsigned long long *ll[PB_PEER_MAX];   // peers' LL receive areas: ll[slot 2][source rank][Ppad] words of {value, seq}
  const unsigned *seq;              // local: number of completed steps
  float *own;                       // this rank's gbuf (writable)
  float *grad_out;                  // local averaged gradient [P]
  float *partials;                  // [gridDim.x] sum of squares of the averaged slice
  int world, rank, P, Ppad, NB;     // NB = blocks of 256 floats = flag slots per source rank
};

__device__ __forceinline__ void peer_st_release(unsigned *p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned peer_ld_acquire(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float peer_ld_f32(const float *p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long peer_timer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// returns false on timeout
__device__ __forceinline__ bool peer_wait(const unsigned *flag, unsigned seq, unsigned long long timeout_ns) {
  if ((int)(peer_ld_acquire(flag) - seq) >= 0) return true;
  const unsigned long long t0 = peer_timer();
  for (;;) {
    for (int k = 0; k < 64; k++)
      if ((int)(peer_ld_acquire(flag) - seq) >= 0) return true;
    if (peer_timer() - t0 > timeout_ns) return false;
    __nanosleep(64);
  }
}

static __global__ void __launch_bounds__(256) peer_avg_kernel(PeerArgs a) {
  __shared__ float sm[256];
  const unsigned seq = *a.seq + 1u;
  if (blockIdx.x == 0 && threadIdx.x < a.world) {
    __threadfence_system();
    peer_st_release(a.flags[threadIdx.x] + (size_t)a.rank * a.NB, seq);
  }
  if (threadIdx.x < a.world) {
    if (!peer_wait(a.flags[a.rank] + (size_t)threadIdx.x * a.NB, seq, PB_PEER_TIMEOUT_NS)) {
      printf("peer_avg_kernel: rank %d timed out waiting for rank %d at step %u\n", a.rank, (int)threadIdx.x, seq);
      __trap();
    }
  }
  __syncthreads();
  const int i = blockIdx.x * 256 + threadIdx.x;
  const size_t off = (size_t)((seq - 1u) & 1u) * a.Ppad + i;
  float s = 0.f;
  if (i < a.P) {
    float v[PB_PEER_MAX];
#pragma unroll
    for (int p = 0; p < PB_PEER_MAX; p++) if (p < a.world) v[p] = peer_ld_f32(a.gbuf[p] + off);
#pragma unroll
    for (int p = 0; p < PB_PEER_MAX; p++) if (p < a.world) s += v[p];
    s *= 1.f / (float)a.world;
    a.grad_out[i] = s;
  }
  sm[threadIdx.x] = s * s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) a.partials[blockIdx.x] = sm[0];
}

// Teardown barrier: publish `seq`, wait (bounded, non-fatal) until every peer has published it too.  After it no peer
// reads this rank's buffers any more.  result[0] = 1 when every peer arrived.
static __global__ void peer_barrier_kernel(PeerArgs a, unsigned seq, unsigned long long timeout_ns, int *result) {
  __shared__ int ok;
  if (threadIdx.x == 0) ok = 1;
  __syncthreads();
  if (threadIdx.x < a.world) {
    __threadfence_system();
    peer_st_release(a.flags[threadIdx.x] + (size_t)PB_PEER_MAX * a.NB + a.rank, seq);
    if (!peer_wait(a.flags[a.rank] + (size_t)PB_PEER_MAX * a.NB + threadIdx.x, seq, timeout_ns)) ok = 0;
  }
  __syncthreads();
  if (threadIdx.x == 0) *result = ok;
}

__device__ __forceinline__ void peer_st_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long peer_ld_u64(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Generic LL exchange of n floats that already lie in `data` (the ac_ppo trainer: its flat gradient, and the two advantage-moment
// scalars of src/ac_ppo_continuous_action.cpp:830-849).  Words go to offset `word_off` of this rank's row in every peer's receive
// area, so several exchanges of one optimiser step use disjoint ranges and share the step's sequence number.  Result in place:
// the rank-ordered sum, divided by the world size when `average`; `partials` (optional) gets each block's sum of squares of the
// result (clip_grad_norm_), `step_count` (optional) is advanced by one.  See grad_reduce_peer_ll_kernel for the protocol.
static __global__ void __launch_bounds__(256) peer_ll_exchange_kernel(float *__restrict__ data, int n, int word_off, int average,
                                                               float *__restrict__ partials, float *step_count, PeerArgs a) {
  __shared__ float sm[256];
  const unsigned seq = *a.seq + 1u;
  const size_t slot = (size_t)((seq - 1u) & 1u);
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (step_count && blockIdx.x == 0 && threadIdx.x == 0) *step_count += 1.f;
  float s = 0.f;
  if (i < n) {
    const float mine = data[i];
    const unsigned long long word = ((unsigned long long)seq << 32) | (unsigned long long)__float_as_uint(mine);
    const size_t out = (slot * a.world + a.rank) * a.Ppad + word_off + i;
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) if (r < a.world && r != a.rank) peer_st_u64(a.ll[r] + out, word);
    float v[PB_PEER_MAX];
    unsigned pending = 0;
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) {
      v[r] = 0.f;
      if (r < a.world && r != a.rank) pending |= 1u << r;
    }
    const unsigned long long *in = a.ll[a.rank] + slot * a.world * a.Ppad + word_off + i;
    unsigned long long t0 = 0;
    for (unsigned spins = 0; pending; spins++) {
#pragma unroll
      for (int r = 0; r < PB_PEER_MAX; r++) {
        if (pending & (1u << r)) {
          const unsigned long long w = peer_ld_u64(in + (size_t)r * a.Ppad);
          if ((unsigned)(w >> 32) == seq) { v[r] = __uint_as_float((unsigned)w); pending &= ~(1u << r); }
        }
      }
      if (pending && (spins & 1023u) == 1023u) {
        const unsigned long long now = peer_timer();
        if (!t0) t0 = now;
        else if (now - t0 > PB_PEER_TIMEOUT_NS) {
          printf("peer_ll_exchange_kernel: rank %d timed out at step %u (offset %d, pending mask %x)\n", a.rank, seq, word_off + i, pending);
          __trap();
        }
      }
    }
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) if (r < a.world) s += (r == a.rank) ? mine : v[r];
    if (average) s *= 1.f / (float)a.world;
    data[i] = s;
  }
  if (partials) {
    sm[threadIdx.x] = s * s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
    if (threadIdx.x == 0) partials[blockIdx.x] = sm[0];
  }
}

#ifdef PB_NSTAT   // pb_ppo_fused.cuh was included first: the slab reduction fused with the exchange
// grad_reduce_kernel + peer_avg_kernel in one launch (see the header comment).  Same slab summation order as
// grad_reduce_kernel, same rank-ordered average as peer_avg_kernel: results are bit-identical to the two-kernel path.
static __global__ void __launch_bounds__(256) grad_reduce_peer_kernel(const float *__restrict__ slabs, size_t slab_stride, int splits,
                                                               const float *__restrict__ part, int ntile, AgentLayout L,
                                                               float *step_count, float *__restrict__ colsum, PeerArgs a) {
  __shared__ float sm[256];
  __shared__ float cols[PB_NSTAT + FB_MAXA + 1];
  const unsigned seq = *a.seq + 1u;
  const size_t slot = (size_t)((seq - 1u) & 1u) * a.Ppad;
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
  if (p < L.total) {
    float s = 0.f;
    if (p >= L.logstd) s = cols[PB_NSTAT + (p - L.logstd)];
    else for (int z = 0; z < splits; z++) s += slabs[(size_t)z * slab_stride + p];
    a.own[slot + p] = s;
  }
  if (step_count && blockIdx.x == 0 && threadIdx.x == 0) *step_count += 1.f;  // optimiser step, read by adam2_kernel
  __syncthreads();   // the slice is written; thread t publishes it to rank t and waits for rank t's slice
  if (threadIdx.x < a.world) {
    __threadfence_system();
    peer_st_release(a.flags[threadIdx.x] + (size_t)a.rank * a.NB + blockIdx.x, seq);
    if (!peer_wait(a.flags[a.rank] + (size_t)threadIdx.x * a.NB + blockIdx.x, seq, PB_PEER_TIMEOUT_NS)) {
      printf("grad_reduce_peer_kernel: rank %d timed out waiting for rank %d (slice %d, step %u)\n", a.rank, (int)threadIdx.x,
             (int)blockIdx.x, seq);
      __trap();
    }
  }
  __syncthreads();
  float s = 0.f;
  if (p < a.P) {
    float v[PB_PEER_MAX];
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) if (r < a.world) v[r] = peer_ld_f32(a.gbuf[r] + slot + p);
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) if (r < a.world) s += v[r];
    s *= 1.f / (float)a.world;
    a.grad_out[p] = s;
  }
  sm[threadIdx.x] = s * s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) a.partials[blockIdx.x] = sm[0];
}
// "LL" variant (PB_PEER_LL, the default): value and sequence number travel in ONE 8-byte store, so there is no flag, no fence and
// no read over NVLink at all.  Block b reduces its 256-float slice of the slabs, pushes {value, seq} words straight into every
// peer's receive area (one-way NVLink stores, 256 B per warp), then polls its OWN receive area until the words of all peers carry
// `seq` - a peer's word is complete the moment it is visible, because an aligned 8-byte store is single-copy atomic.  Cost over the
// single-GPU grad_reduce_kernel: the one-way flight of the stores plus rank skew; the two-kernel form pays a launch, a flag flight
// and a remote-read round trip (update at 8 GPUs 35.0 ms with flags + remote reads; see DESIGN.md section 5 for the LL figure).
// Same slab summation order as grad_reduce_kernel and the same rank-ordered average as peer_avg_kernel: bit-identical results.
// Two slots suffice for the same reason as in the header comment (a peer pushes seq + 2 only after it has seen my seq + 1, which I
// send after I have finished polling seq).
static __global__ void __launch_bounds__(256) grad_reduce_peer_ll_kernel(const float *__restrict__ slabs, size_t slab_stride, int splits,
                                                                  const float *__restrict__ part, int ntile, AgentLayout L,
                                                                  float *step_count, float *__restrict__ colsum, PeerArgs a) {
  __shared__ float sm[256];
  __shared__ float cols[PB_NSTAT + FB_MAXA + 1];
  const unsigned seq = *a.seq + 1u;
  const size_t slot = (size_t)((seq - 1u) & 1u);
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
  if (step_count && blockIdx.x == 0 && threadIdx.x == 0) *step_count += 1.f;  // optimiser step, read by adam2_kernel
  float s = 0.f;
  if (p < a.P) {
    float mine = 0.f;
    if (p >= L.logstd) mine = cols[PB_NSTAT + (p - L.logstd)];
    else for (int z = 0; z < splits; z++) mine += slabs[(size_t)z * slab_stride + p];
    const unsigned long long word = ((unsigned long long)seq << 32) | (unsigned long long)__float_as_uint(mine);
    const size_t out = (slot * a.world + a.rank) * a.Ppad + p;      // my row in every peer's receive area
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) if (r < a.world && r != a.rank) peer_st_u64(a.ll[r] + out, word);
    float v[PB_PEER_MAX];
    unsigned pending = 0;
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) {
      v[r] = 0.f;
      if (r < a.world && r != a.rank) pending |= 1u << r;
    }
    const unsigned long long *in = a.ll[a.rank] + slot * a.world * a.Ppad + p;
    unsigned long long t0 = 0;
    for (unsigned spins = 0; pending; spins++) {
#pragma unroll
      for (int r = 0; r < PB_PEER_MAX; r++) {
        if (pending & (1u << r)) {
          const unsigned long long w = peer_ld_u64(in + (size_t)r * a.Ppad);
          if ((unsigned)(w >> 32) == seq) { v[r] = __uint_as_float((unsigned)w); pending &= ~(1u << r); }
        }
      }
      if (pending && (spins & 1023u) == 1023u) {
        const unsigned long long now = peer_timer();
        if (!t0) t0 = now;
        else if (now - t0 > PB_PEER_TIMEOUT_NS) {
          printf("grad_reduce_peer_ll_kernel: rank %d timed out at step %u (slice %d, pending mask %x)\n", a.rank, seq, (int)blockIdx.x, pending);
          __trap();
        }
      }
    }
#pragma unroll
    for (int r = 0; r < PB_PEER_MAX; r++) if (r < a.world) s += (r == a.rank) ? mine : v[r];
    s *= 1.f / (float)a.world;
    a.grad_out[p] = s;
  }
  sm[threadIdx.x] = s * s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) a.partials[blockIdx.x] = sm[0];
}
#endif

// ------------------------------------------------------------------ host side
#ifdef PB_PEER_HOST
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "pb_nccl.h"

// Owns the exported buffer and the peers' mappings.  setup() is collective over the NCCL communicator (the IPC handles
// travel through one ncclAllReduce, the go / no-go decision through another), so either every rank uses the peer path
// or every rank stays on ncclAllReduce.  PB_PEER_AR=0 forces the NCCL path.
struct PeerExchange {
  bool on = false, fused = false, ll = false;
  int world = 1, rank = 0, P = 0, Ppad = 0, NB = 0;
  void *base = nullptr;
  void *peer_base[PB_PEER_MAX] = {};
  unsigned *seq = nullptr;
  int *result = nullptr;
  PeerArgs args;

  float *gbuf() const { return (float *)base; }

  // returns 0 on success (on == true or a clean collective fallback), non-zero only when NCCL itself failed
  int setup(ncclComm_t comm, cudaStream_t st, int rank_, int world_, int P_, float *grad_out, float *partials) {
    rank = rank_; world = world_; P = P_; Ppad = (P + 63) & ~63; NB = (P + 255) / 256;
    fused = getenv("PB_PEER_FUSED") && atoi(getenv("PB_PEER_FUSED"));   // measured slower than two kernels, see the header
    int ok = world > 1 && world <= PB_PEER_MAX;
    if (getenv("PB_PEER_AR") && !atoi(getenv("PB_PEER_AR"))) ok = 0;
    ll = !(getenv("PB_PEER_LL") && !atoi(getenv("PB_PEER_LL"))) && !fused;   // default: the LL one-kernel exchange
    // [gbuf 2 x Ppad floats][flags PB_PEER_MAX x NB + PB_PEER_MAX words][pad to 16][LL words 2 x world x Ppad x 8 B]
    const size_t ll_off = ((size_t)2 * Ppad * sizeof(float) + ((size_t)PB_PEER_MAX * NB + PB_PEER_MAX) * sizeof(unsigned) + 15) / 16 * 16;
    const size_t bytes = ll_off + (size_t)2 * (world > 0 ? world : 1) * Ppad * sizeof(unsigned long long);
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    if (ok) {
      if (cudaMalloc(&base, bytes) != cudaSuccess || cudaMemset(base, 0, bytes) != cudaSuccess ||
          cudaMalloc((void **)&seq, 2 * sizeof(unsigned)) != cudaSuccess || cudaMemset(seq, 0, 2 * sizeof(unsigned)) != cudaSuccess ||
          cudaIpcGetMemHandle(&mine, base) != cudaSuccess) { ok = 0; cudaGetLastError(); }
      result = (int *)(seq + 1);
    }
    // all-gather of the handles as a sum over zero-initialised slots (the loader resolves only AllReduce / Broadcast)
    const int W = world > 0 ? world : 1;
    int *x = nullptr;
    if (cudaMalloc((void **)&x, (size_t)(W * 16 + 1) * sizeof(int)) != cudaSuccess) return 1;
    cudaMemset(x, 0, (size_t)(W * 16 + 1) * sizeof(int));
    if (ok) cudaMemcpy(x + rank * 16, &mine, 64, cudaMemcpyHostToDevice);
    if (g_nccl.AllReduce(x, x, (size_t)W * 16, ncclInt32, ncclSum, comm, st) != ncclSuccess) { cudaFree(x); return 1; }
    cudaStreamSynchronize(st);
    cudaIpcMemHandle_t all[PB_PEER_MAX];
    if (ok) {
      cudaMemcpy(all, x, (size_t)W * 64, cudaMemcpyDeviceToHost);
      cudaIpcMemHandle_t zero;
      memset(&zero, 0, sizeof(zero));
      for (int p = 0; p < W && ok; p++) {
        if (p == rank) { peer_base[p] = base; continue; }
        if (!memcmp(&all[p], &zero, 64)) { ok = 0; break; }   // that rank could not export
        if (cudaIpcOpenMemHandle(&peer_base[p], all[p], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          peer_base[p] = nullptr; ok = 0; cudaGetLastError();
        }
      }
    }
    // unanimous decision: sum of the per-rank verdicts must equal world
    int h = ok;
    cudaMemcpy(x + W * 16, &h, sizeof(int), cudaMemcpyHostToDevice);
    if (g_nccl.AllReduce(x + W * 16, x + W * 16, 1, ncclInt32, ncclSum, comm, st) != ncclSuccess) { cudaFree(x); return 1; }
    cudaStreamSynchronize(st);
    cudaMemcpy(&h, x + W * 16, sizeof(int), cudaMemcpyDeviceToHost);
    cudaFree(x);
    on = ok && h == W;
    if (!on) {
      release_mappings();
      if (rank == 0) fprintf(stderr, "[ppo_b200] gradient exchange: ncclAllReduce over %d ranks (peer-memory path %s)\n", W,
                             getenv("PB_PEER_AR") && !atoi(getenv("PB_PEER_AR")) ? "disabled by PB_PEER_AR=0" : "unavailable");
      return 0;
    }
    memset(&args, 0, sizeof(args));
    for (int p = 0; p < W; p++) {
      args.gbuf[p] = (const float *)peer_base[p];
      args.flags[p] = (unsigned *)((float *)peer_base[p] + (size_t)2 * Ppad);
      args.ll[p] = (unsigned long long *)((char *)peer_base[p] + ll_off);
    }
    args.seq = seq; args.own = (float *)base; args.grad_out = grad_out; args.partials = partials;
    args.world = W; args.rank = rank; args.P = P; args.Ppad = Ppad; args.NB = NB;
    if (rank == 0) fprintf(stderr, "[ppo_b200] gradient exchange: fused peer-memory all-reduce over %d ranks (CUDA IPC, %d floats, %s)\n", W, P,
                           ll ? "one kernel, {value, seq} words pushed to the peers" : (fused ? "slab reduction + exchange in one kernel" : "two kernels"));
    return 0;
  }

  void release_mappings() {
    for (int p = 0; p < PB_PEER_MAX; p++) {
      if (peer_base[p] && peer_base[p] != base) cudaIpcCloseMemHandle(peer_base[p]);
      peer_base[p] = nullptr;
    }
    if (base) cudaFree(base);
    if (seq) cudaFree(seq);
    base = nullptr; seq = nullptr; result = nullptr;
    cudaGetLastError();
  }

  // after the stream has drained: wait (bounded) until no peer can still be reading this rank's buffers, then unmap / free
  void teardown(cudaStream_t st) {
    if (!on) return;
    on = false;
    if (cudaStreamQuery(st) == cudaSuccess) {
      peer_barrier_kernel<<<1, 32, 0, st>>>(args, 1u, 5000000000ull, result);
      cudaStreamSynchronize(st);
    }
    release_mappings();
  }
};
#endif
