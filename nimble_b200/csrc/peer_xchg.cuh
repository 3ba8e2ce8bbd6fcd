// peer_xchg.cuh -- device-side all-reduce (sum, FP64) over NVLink peer memory for the data-sharded mode.
//
// The per-proposal exchange of a data-sharded run is `chains` doubles (64 bytes for the LMM's 8 chains): pure latency.
// A host-issued ncclAllReduce costs a proxy hand-off per call and cannot sit inside a captured CUDA graph of ours without
// taking the whole iteration out of graph replay.  Here every rank owns one exchange buffer in its own HBM that its
// peers map (CUDA IPC between processes, plain peer access inside one process); one small kernel per all-reduce
//   1. PUSHES this rank's partial vector into slot [parity][this rank] of every rank's buffer (plain st.global over NVLink),
//   2. fences and stores an epoch flag next to it on every rank,
//   3. spins on the flags in its OWN memory until every rank's epoch has arrived,
//   4. adds the R vectors up in rank order 0..R-1 -- the same order on every rank, hence bit-identical sums and identical
//      accept / reject decisions everywhere.
// Only plain loads and stores touch peer memory (no TMA, no multicast).  The epoch is the sequence number of the all-reduce
// (one per kernel, the same on every rank because every rank runs the same schedule); its low bit selects one of two
// copies of the buffer.  Two suffice: a rank can only start all-reduce e+2 after it has seen every peer's flags of e+1,
// which a peer writes after its own all-reduce e -- reads included -- has finished.
// The sequence number lives in device memory and is advanced by the kernel itself (by its last block), so the kernel can
// be captured into the iteration's CUDA graph and replayed.  A rank whose peers never arrive gives up after XCHG_TIMEOUT_CYCLES, raises the
// error word and lets the stream drain (the host reports it) -- a kernel must never spin forever; once the error word is set
// every later all-reduce returns at once (NaN), so a broken group costs one time-out, not one per exchange.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace nimb {

constexpr int XCHG_MAX_RANKS = 8;          // one NVSwitch box
constexpr int XCHG_MAX_BLOCKS = 64;        // blocks of one all-reduce kernel (each owns a slice of the vector and its own flags)
constexpr int XCHG_THREADS = 256;
constexpr size_t XCHG_CAP = 1u << 18;      // doubles per (parity, source rank): 2 MiB; longer vectors go in pieces
constexpr long long XCHG_TIMEOUT_CYCLES = 1ll << 33;  // ~4 s at 1.9 GHz

// layout of one rank's exchange allocation: data[2][R][XCHG_CAP] doubles, then flags[2][R][XCHG_MAX_BLOCKS] (u64),
// then seq (u64), ticket (u32), err (u32) -- the last three for local use only
struct PeerXchg {
  int nranks = 0, rank = 0;
  double* data[XCHG_MAX_RANKS] = {nullptr};                 // data[p]: rank p's buffer as mapped into this process
  unsigned long long* flags[XCHG_MAX_RANKS] = {nullptr};
  unsigned long long* seq = nullptr;                        // all-reduces completed so far on this rank
  unsigned int* ticket = nullptr;                           // block ticket: the last block of a kernel advances seq
  unsigned int* err = nullptr;                              // this rank's error word (1 = a wait timed out)
};

static inline size_t xchg_data_doubles(int R) { return (size_t)2 * R * XCHG_CAP; }
static inline size_t xchg_bytes(int R) {
  return xchg_data_doubles(R) * sizeof(double) + sizeof(unsigned long long) * ((size_t)2 * R * XCHG_MAX_BLOCKS + 1) + 64;
}

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(double* p, double v) {
  asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}

// buf[0..count) <- sum over ranks of buf, in place, identical bits on every rank.  count <= XCHG_CAP.  Every rank launches
// the same grid for the same all-reduce (the grid only depends on count).
__global__ void __launch_bounds__(XCHG_THREADS) k_peer_allreduce(PeerXchg X, double* __restrict__ buf, int count) {
  __shared__ unsigned long long s_epoch;
  __shared__ int s_bad;
  const int b = blockIdx.x, R = X.nranks, tid = threadIdx.x;
  if (tid == 0) {
    s_epoch = *X.seq + 1;   // nobody advances seq before every block of this kernel has passed this read (see the end)
    s_bad = *(volatile unsigned int*)X.err != 0;   // a wait has timed out before: the group is broken, do not wait again
  }
  __syncthreads();
  const unsigned long long e = s_epoch;
  const int par = (int)(e & 1ull);
  const int per = (count + gridDim.x - 1) / gridDim.x;
  const int lo = b * per, hi = min(count, lo + per);
  // 1. push my slice to every rank (my own copy included), starting with my right-hand neighbour to spread the traffic
  for (int k = 1; k <= R; k++) {
    const int p = (X.rank + k) % R;
    double* dst = X.data[p] + ((size_t)par * R + X.rank) * XCHG_CAP;
    for (int i = lo + tid; i < hi; i += XCHG_THREADS) st_relaxed_sys(dst + i, buf[i]);
  }
  __syncthreads();
  // 2. one thread per destination rank publishes the epoch.  st.release at system scope is cumulative over everything that
  // happens-before it, and the CTA barrier orders every thread's data stores before it: no per-thread fence is needed
  if (tid < R) st_release_sys(X.flags[tid] + ((size_t)par * R + X.rank) * XCHG_MAX_BLOCKS + b, e);
  // 3. one thread per source rank waits for that rank's epoch in MY memory
  if (tid < R && !s_bad) {
    const unsigned long long* f = X.flags[X.rank] + ((size_t)par * R + tid) * XCHG_MAX_BLOCKS + b;
    const long long t0 = clock64();
    while (ld_acquire_sys(f) < e) {
      if (clock64() - t0 > XCHG_TIMEOUT_CYCLES) { s_bad = 1; break; }
    }
  }
  __syncthreads();
  if (s_bad) {
    if (tid == 0) atomicExch(X.err, 1u);
    for (int i = lo + tid; i < hi; i += XCHG_THREADS) buf[i] = __longlong_as_double(0x7ff8000000000000ll);
  } else {
    // 4. fixed-order sum of the R contributions that now sit in my own buffer
    const double* mine = X.data[X.rank] + (size_t)par * R * XCHG_CAP;
    for (int i = lo + tid; i < hi; i += XCHG_THREADS) {
      double s = ld_relaxed_sys(mine + i);
      for (int q = 1; q < R; q++) s += ld_relaxed_sys(mine + (size_t)q * XCHG_CAP + i);
      buf[i] = s;
    }
  }
  // the last block to get here advances the sequence number for the next all-reduce (ordinary stream order: that kernel
  // starts after this one has ended)
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(X.ticket, 1u) == gridDim.x - 1) { *X.ticket = 0; *X.seq = e; }
  }
}

static inline int xchg_blocks(size_t count) {
  size_t nb = (count + 4 * XCHG_THREADS - 1) / (4 * XCHG_THREADS);
  return (int)(nb < 1 ? 1 : nb > XCHG_MAX_BLOCKS ? XCHG_MAX_BLOCKS : nb);
}

}  // namespace nimb
