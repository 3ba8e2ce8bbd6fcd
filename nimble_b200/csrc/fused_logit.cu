// fused_logit.cu -- fused block log-probability of  y[i] ~ dbern(ilogit(inprod(X[i,1:p], beta[1:p])))
// for ALL chains at once (kernels K1+K2).
//
// The reference walks, per chain and per node, a lifted deterministic node (Eigen inprod + ilogit,
// P/R/BUGS_mathCompatibility.R:3-6,38-39; P/inst/include/nimble/Utils.h:187) and then a stochastic node
// (dbinom(y, 1, pr, log=1) from Rmath) through virtual calls (P/inst/CppCode/accessorClasses.cpp:49-57).
// Here the linear predictor of a (128 node x 64 chain) tile is one FP64 tensor-core contraction
// (mma.sync m8n8k4 f64: X tile [128 x p] times the chains' beta block [p x 64]), and the link, the Bernoulli
// log-density (same branches as nmath's dbinom_raw) and the per-chain reduction run on the accumulators in
// registers: X and y are read from HBM once per proposal for all chains, nothing per-node is ever stored.
//
// Data movement: the X tile is a contiguous 128 x P4 block (rows padded to a multiple of 4 doubles), fetched
// with one 1-D TMA bulk copy (cp.async.bulk, completion on an mbarrier); beta is staged in shared memory once
// per block; per-chain partial sums leave through registers -> warp shuffles -> shared -> partials[tile][C].
#include <cfloat>
#include <cmath>
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include "engine.cuh"

namespace nimb {

static const int FT_M = 128;      // nodes per tile
static const int FT_NC = 64;      // chains per block
static const int FT_THREADS = 256;
static const int FT_MAXP4 = 64;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void dmma_m8n8k4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// bd0(x, np) with x = 1 (nmath/bd0.c), as reached from dbinom_raw with n = 1
__device__ __forceinline__ double bd0_one(double np) {
  if (!isfinite(np) || np == 0.0) return NAN;
  if (fabs(1.0 - np) < 0.1 * (1.0 + np)) {
    double v = (1.0 - np) / (1.0 + np);
    double s = (1.0 - np) * v;
    if (fabs(s) < DBL_MIN) return s;
    double ej = 2 * v;
    v = v * v;
    for (int j = 1; j < 1000; j++) {
      ej *= v;
      double s1 = s + ej / (double)((j << 1) + 1);
      if (s1 == s) return s1;
      s = s1;
    }
  }
  return log(1.0 / np) + np - 1.0;
}

// dbinom(y, 1, ilogit(eta), log = 1) for y in {0, 1}: nmath dbinom -> dbinom_raw(x, 1, p, 1-p)
__device__ __forceinline__ double bern_logit_lp(double eta, double y) {
  double p = 1. / (1. + exp(-eta));
  if (isnan(p)) return p;
  double q = 1 - p;
  if (y == 0.0) {
    if (p == 0) return 0.;
    if (q == 0) return -INFINITY;
    return (p < 0.1) ? -bd0_one(q) - p : log(q);
  }
  if (p == 0) return -INFINITY;
  if (q == 0) return 0.;
  return (q < 0.1) ? -bd0_one(p) - q : log(p);
}

// grid: (node chunks, chain tiles).  partials[blockIdx.x][C]
__global__ void __launch_bounds__(FT_THREADS, 2)
k_fused_bern_logit(const double* __restrict__ Xpad, const double* __restrict__ y, int n, int P4, int p,
                   const double* __restrict__ state, int beta_slot, int C, int tiles_per_chunk,
                   double* __restrict__ partials) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int LDB = FT_NC + 4;  // (k*LDB*2 + n*2) mod 32 distinct over a half warp -> conflict-free B fragments
  double* Xs = (double*)smem_raw;                       // [FT_M][P4]
  double* Bs = Xs + FT_M * P4;                          // [P4][LDB]
  double* ys = Bs + P4 * LDB;                           // [FT_M]
  double* red = ys + FT_M;                              // [4][FT_NC]
  uint64_t* bar = (uint64_t*)(red + 4 * FT_NC);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp & 3, wn = warp >> 2;              // warp tile: nodes wm*32.., chains wn*32..
  const int chain0 = blockIdx.y * FT_NC;

  for (int e = tid; e < P4 * FT_NC; e += FT_THREADS) {  // stage this block's beta columns (zero padded)
    int k = e / FT_NC, cc = e % FT_NC;
    int c = chain0 + cc;
    Bs[k * LDB + cc] = (k < p && c < C) ? state[(size_t)(beta_slot + k) * C + c] : 0.0;
  }
  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double csum[8];
#pragma unroll
  for (int j = 0; j < 8; j++) csum[j] = 0.0;

  const int ntiles_total = (n + FT_M - 1) / FT_M;
  const int tile_lo = blockIdx.x * tiles_per_chunk;
  const int tile_hi = min(ntiles_total, tile_lo + tiles_per_chunk);
  const uint32_t tile_bytes = (uint32_t)(FT_M * P4 * sizeof(double));
  uint32_t phase = 0;

  for (int tile = tile_lo; tile < tile_hi; tile++) {
    const int node0 = tile * FT_M;
    if (tid == 0) {
      mbar_expect_tx(bar, tile_bytes);
      bulk_g2s(Xs, Xpad + (size_t)node0 * P4, tile_bytes, bar);
    }
    if (tid < FT_M) ys[tid] = (node0 + tid < n) ? y[node0 + tid] : -1.0;  // -1 marks padding rows
    mbar_wait(bar, phase);
    phase ^= 1;
    __syncthreads();

    double acc[4][4][2];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
      for (int b = 0; b < 4; b++) acc[a][b][0] = acc[a][b][1] = 0.0;

    const double* xa = Xs + (wm * 32 + (lane >> 2)) * P4 + (lane & 3);
    const double* bb = Bs + (lane & 3) * LDB + wn * 32 + (lane >> 2);
    for (int k0 = 0; k0 < P4; k0 += 4) {
      double af[4], bf[4];
#pragma unroll
      for (int a = 0; a < 4; a++) af[a] = xa[a * 8 * P4 + k0];
#pragma unroll
      for (int b = 0; b < 4; b++) bf[b] = bb[k0 * LDB + b * 8];
#pragma unroll
      for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 4; b++) dmma_m8n8k4(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
    // epilogue on the accumulators: lane holds rows (lane>>2) of each m-tile, columns 2*(lane&3)+{0,1} of each n-tile
#pragma unroll
    for (int a = 0; a < 4; a++) {
      double yv = ys[wm * 32 + a * 8 + (lane >> 2)];
      if (yv >= 0.0) {
#pragma unroll
        for (int b = 0; b < 4; b++) {
          csum[b * 2 + 0] += bern_logit_lp(acc[a][b][0], yv);
          csum[b * 2 + 1] += bern_logit_lp(acc[a][b][1], yv);
        }
      }
    }
    __syncthreads();  // everyone is done with Xs / ys before the next bulk copy lands
  }

  // reduce over the 8 row-lanes of the warp (lanes with equal lane&3), then over the 4 node-warps
#pragma unroll
  for (int j = 0; j < 8; j++) {
    double v = csum[j];
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    csum[j] = v;
  }
  if ((lane >> 2) == 0) {
#pragma unroll
    for (int b = 0; b < 4; b++) {
      red[wm * FT_NC + wn * 32 + b * 8 + 2 * (lane & 3) + 0] = csum[b * 2 + 0];
      red[wm * FT_NC + wn * 32 + b * 8 + 2 * (lane & 3) + 1] = csum[b * 2 + 1];
    }
  }
  __syncthreads();
  if (tid < FT_NC) {
    int c = chain0 + tid;
    if (c < C) partials[(size_t)blockIdx.x * C + c] = ((red[tid] + red[FT_NC + tid]) + red[2 * FT_NC + tid]) + red[3 * FT_NC + tid];
  }
}

// X [n][ld] (first p columns used) -> Xpad [ntiles*FT_M][P4], zero padded
__global__ void k_pad_x(const double* __restrict__ X, int n, int p, int ld, int P4, size_t total, double* __restrict__ Xpad) {
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  size_t i = gid / P4;
  int k = (int)(gid % P4);
  Xpad[gid] = (i < (size_t)n && k < p) ? X[i * ld + k] : 0.0;
}

static int p4_of(int p) { return (p + 3) & ~3; }

static int g_chunks_cache_n = -1, g_chunks_cache = 0;
static int chunks_for(int n) {
  // node chunks: aim at ~2 waves of 2 resident blocks per SM when there are 16 chain tiles; the count only
  // depends on n so that the scratch size is known up front
  if (n == g_chunks_cache_n) return g_chunks_cache;
  int ntiles = (n + FT_M - 1) / FT_M;
  int chunks = std::min(ntiles, 148);
  g_chunks_cache_n = n; g_chunks_cache = chunks;
  return chunks;
}

int fused_bern_logit_tiles(int n) { return chunks_for(n); }

size_t fused_bern_logit_xpad_len(int n, int p) {
  if (p < 1 || p4_of(p) > FT_MAXP4) return 0;
  if (getenv("NIMB200_NO_FUSED")) return 0;
  size_t ntiles = (size_t)(n + FT_M - 1) / FT_M;
  return ntiles * FT_M * p4_of(p);
}

void fused_bern_logit_prepare_x(const double* X, int n, int p, int ld, double* Xpad, cudaStream_t st) {
  int P4 = p4_of(p);
  size_t total = fused_bern_logit_xpad_len(n, p);
  k_pad_x<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(X, n, p, ld, P4, total, Xpad);
}

void fused_bern_logit_launch(const double* Xpad, const double* y, int n, int p, const double* state, int beta_slot, int C,
                             double* partials, cudaStream_t st) {
  int P4 = p4_of(p);
  int LDB = FT_NC + 4;
  size_t smem = sizeof(double) * ((size_t)FT_M * P4 + (size_t)P4 * LDB + FT_M + 4 * FT_NC) + 16;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(k_fused_bern_logit, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024);
    attr_set = true;
  }
  int chunks = chunks_for(n);
  int ntiles = (n + FT_M - 1) / FT_M;
  int tpc = (ntiles + chunks - 1) / chunks;
  dim3 grid(chunks, (C + FT_NC - 1) / FT_NC);
  k_fused_bern_logit<<<grid, FT_THREADS, smem, st>>>(Xpad, y, n, P4, p, state, beta_slot, C, tpc, partials);
}

}  // namespace nimb
