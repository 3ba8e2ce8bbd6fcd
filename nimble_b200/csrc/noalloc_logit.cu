// noalloc_logit.cu -- nimb200_bern_logit_block_logprob_noalloc(): the per-chain block log-probability of
//   y[i] ~ dbern(ilogit(inprod(X[i,1:p], beta[1:p])))
// for callers that are not allowed to allocate device memory.
//
// Who that is: a host program that gives the whole GPU to its own memory manager (an arena sized at start-up, a
// fixed per-worker budget under MPS, a framework that treats any foreign cudaMalloc as a leak) and so cannot let a
// library call cudaMalloc behind its back; and a steady-state sampling loop in which cudaMalloc / cudaFree are
// unacceptable because they synchronise the device and have no latency bound.  Every other entry point of this
// library works on a nimb200_model whose creation allocates; this one does not need a model at all.
//
// It is a variant of the fused block-sum kernel (fused_logit.cu): same FP64 mma.sync contraction, same Bernoulli
// epilogue, but X is read in place from the caller's row-major array, and the tile partial sums -- the kernel's only
// scratch -- live in a file-scope __device__ array of fixed size instead of a cudaMalloc'ed buffer.  Consequences the
// caller must know:
//   * limits: at most NIMB200_NOALLOC_MAX_CHAINS chains and p <= 64; larger requests are rejected with an error;
//   * every GPU that loads this library holds the 2.3 MiB array until the process exits;
//   * the FIRST call on each device resolves the array's device address (cudaGetSymbolAddress; in host code the
//     array's name is only a host-side handle, not something a kernel may dereference).  Make that first call before
//     any stream capture begins;
//   * all calls on a device share the one array.  A launch overwrites every element it later reads, so nothing has to
//     be cleared between calls, but two calls whose kernels overlap on the GPU (different streams, or different host
//     threads) race on it: serialise them.
//   * nothing is allocated or freed here: no cudaMalloc / cudaFree / cudaMallocAsync, no library that allocates.
#include <cuda_runtime.h>
#include <cmath>
#include <cstdint>
#include <mutex>
#include <string>
#include "../../include/nimb200.h"

namespace nimb { void set_error(const std::string& s); }

namespace {

constexpr int NA_CHUNKS = 148;      // node chunks = rows of the partial-sum scratch
constexpr int NA_NC = 64;           // chains per CTA
constexpr int NA_THREADS = 256;
constexpr int NA_MAXK = 16;         // p <= 64

// the scratch: partial sums [NA_CHUNKS][C], C <= NIMB200_NOALLOC_MAX_CHAINS
static __device__ __align__(256) double g_noalloc_scratch[(size_t)NA_CHUNKS * NIMB200_NOALLOC_MAX_CHAINS];

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// same arithmetic as bern_logit_lp in fused_logit.cu (see the derivation there)
__device__ __forceinline__ double bern_lp(double eta, bool y1) {
  double e = exp(-eta);
  double p = 1.0 / (1.0 + e);
  double q = 1.0 - p;
  double t = y1 ? p : q;
  double delta = (1.0 - q) - p;
  double corr = (y1 || q == 0.0) ? 0.0 : delta * (double)__frcp_rn((float)q);
  return log(t) + corr;
}

// grid (NA_CHUNKS, ceil(C/64)); scratch[blockIdx.x][c] receives the chunk's partial sum for chain c
__global__ void __launch_bounds__(NA_THREADS)
k_bern_logit_direct(const double* __restrict__ X, long long ld, const double* __restrict__ y, int n, int p,
                    const double* __restrict__ beta, int C, int rows_per_chunk, double* __restrict__ scratch) {
  __shared__ double red[2 * NA_NC];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wn = warp & 3, wg = warp >> 2;
  const int chain0 = blockIdx.y * NA_NC + wn * 16;
  const int nk = (p + 3) >> 2;
  double bf[NA_MAXK][2];
#pragma unroll
  for (int ks = 0; ks < NA_MAXK; ks++)
#pragma unroll
    for (int nt = 0; nt < 2; nt++) {
      int k = ks * 4 + (lane & 3), c = chain0 + nt * 8 + (lane >> 2);
      bf[ks][nt] = (k < p && c < C) ? beta[(size_t)k * C + c] : 0.0;
    }
  double csum[4] = {0.0, 0.0, 0.0, 0.0};
  const int row_lo = blockIdx.x * rows_per_chunk, row_hi = min(n, row_lo + rows_per_chunk);
  // each warp group walks its half of the chunk in steps of 8 rows
  for (int r0 = row_lo + wg * 8; r0 < row_hi; r0 += 16) {
    int row = r0 + (lane >> 2);
    bool live = row < row_hi;
    const double* xr = X + (size_t)(live ? row : row_lo) * ld;
    double acc[2][2] = {{0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
    for (int ks = 0; ks < NA_MAXK; ks++) {
      if (ks < nk) {
        int k = ks * 4 + (lane & 3);
        double a = (live && k < p) ? __ldg(xr + k) : 0.0;
        dmma(acc[0][0], acc[0][1], a, bf[ks][0]);
        dmma(acc[1][0], acc[1][1], a, bf[ks][1]);
      }
    }
    bool y1 = live ? (__ldg(y + row) != 0.0) : false;
#pragma unroll
    for (int nt = 0; nt < 2; nt++) {
      double l0 = bern_lp(acc[nt][0], y1), l1 = bern_lp(acc[nt][1], y1);
      csum[nt * 2 + 0] += live ? l0 : 0.0;
      csum[nt * 2 + 1] += live ? l1 : 0.0;
    }
  }
#pragma unroll
  for (int j = 0; j < 4; j++) {
    double v = csum[j];
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    csum[j] = v;
  }
  if ((lane >> 2) == 0) {
#pragma unroll
    for (int nt = 0; nt < 2; nt++) {
      red[wg * NA_NC + wn * 16 + nt * 8 + 2 * (lane & 3) + 0] = csum[nt * 2 + 0];
      red[wg * NA_NC + wn * 16 + nt * 8 + 2 * (lane & 3) + 1] = csum[nt * 2 + 1];
    }
  }
  __syncthreads();
  if (tid < NA_NC) {
    int c = blockIdx.y * NA_NC + tid;
    if (c < C) scratch[(size_t)blockIdx.x * C + c] = red[tid] + red[NA_NC + tid];  // written even for empty chunks
  }
}

__global__ void k_noalloc_reduce(const double* __restrict__ scratch, int C, double* __restrict__ out) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s = 0.0;
  for (int b = 0; b < NA_CHUNKS; b++) s += scratch[(size_t)b * C + c];
  out[c] = s;
}

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
double* g_scratch_addr[MAX_DEVICES] = {nullptr};

int fail(const std::string& s) { nimb::set_error(s); return 1; }

}  // namespace

extern "C" int nimb200_bern_logit_block_logprob_noalloc(const double* d_X, int64_t ld, const double* d_y, int32_t n, int32_t p,
                                                        const double* d_beta, int32_t nchains, double* d_out,
                                                        void* cuda_stream) {
  if (!d_X || !d_y || !d_beta || !d_out) return fail("nimb200_bern_logit_block_logprob_noalloc: null pointer");
  if (n < 0) return fail("nimb200_bern_logit_block_logprob_noalloc: n < 0");
  // the kernel's row counters are 32-bit: keep chunk starts and the 16-row stride clear of INT32_MAX
  if (n > INT32_MAX - 32 * NA_CHUNKS) return fail("nimb200_bern_logit_block_logprob_noalloc: n exceeds " + std::to_string(INT32_MAX - 32 * NA_CHUNKS));
  if (p < 1 || p > 4 * NA_MAXK) return fail("nimb200_bern_logit_block_logprob_noalloc: p must be between 1 and 64");
  if (ld < p) return fail("nimb200_bern_logit_block_logprob_noalloc: row stride ld smaller than p");
  if (nchains < 1 || nchains > NIMB200_NOALLOC_MAX_CHAINS)
    return fail("nimb200_bern_logit_block_logprob_noalloc: nchains exceeds NIMB200_NOALLOC_MAX_CHAINS (" +
                std::to_string(NIMB200_NOALLOC_MAX_CHAINS) + "), the fixed size of the built-in scratch array");
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(std::string("nimb200_bern_logit_block_logprob_noalloc: ") + cudaGetErrorString(e));
  if (dev < 0 || dev >= MAX_DEVICES) return fail("nimb200_bern_logit_block_logprob_noalloc: device index out of range");
  double* scratch = nullptr;
  {
    // set-up, once per device, with that device current: the name g_noalloc_scratch is only a host-side handle here
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_addr[dev]) {
      void* ptr = nullptr;
      e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
      if (e != cudaSuccess || !ptr)
        return fail(std::string("nimb200_bern_logit_block_logprob_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
      g_scratch_addr[dev] = (double*)ptr;
    }
    scratch = g_scratch_addr[dev];
  }
  cudaStream_t st = (cudaStream_t)cuda_stream;
  int rows_per_chunk = (int)(((long long)n + NA_CHUNKS - 1) / NA_CHUNKS);   // 64-bit: n may be within NA_CHUNKS of INT32_MAX
  rows_per_chunk = (rows_per_chunk + 15) & ~15;
  if (rows_per_chunk == 0) rows_per_chunk = 16;
  dim3 grid(NA_CHUNKS, (nchains + NA_NC - 1) / NA_NC);
  k_bern_logit_direct<<<grid, NA_THREADS, 0, st>>>(d_X, (long long)ld, d_y, n, p, d_beta, nchains, rows_per_chunk, scratch);
  k_noalloc_reduce<<<(nchains + 127) / 128, 128, 0, st>>>(scratch, nchains, d_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail(std::string("nimb200_bern_logit_block_logprob_noalloc: launch failed: ") + cudaGetErrorString(e));
  return 0;
}
