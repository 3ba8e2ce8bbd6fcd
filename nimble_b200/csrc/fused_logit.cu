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
#include "fastmath.cuh"
#include "bern_logit.cuh"

namespace nimb {

static const int FT_M = 128;      // nodes per tile (one TMA bulk copy)
static const int FT_NC = 64;      // chains per block: 4 warps x 16 chains, two such warp groups split the rows
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

// grid: (chain tiles, node chunks): the chain tiles of one node chunk are scheduled together and share its X tiles
// through L2.  partials[chunk][C]
// smem: 2 stages x { X tile [FT_M][P4], y tile [FT_M] }, 2 mbarriers.   NK = k-steps of 4 (P4 = 4*NK).
template <int NK>
__global__ void __launch_bounds__(FT_THREADS, 2)
k_fused_bern_logit(const double* __restrict__ Xpad, const double* __restrict__ ypad, int n, int p,
                   const double* __restrict__ state, int beta_slot, int C, int tiles_per_chunk, int tile_begin, int tile_end,
                   double* __restrict__ partials) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int P4 = 4 * NK;
  constexpr int stage_doubles = FT_M * P4 + FT_M;
  double* stage0 = (double*)smem_raw;
  double* red = stage0 + 2 * stage_doubles;             // [2][FT_NC]
  uint64_t* bar = (uint64_t*)(red + 2 * FT_NC);          // [2]
  MathTab* tab = (MathTab*)(bar + 2);                    // exp / log tables
  for (int i = threadIdx.x; i < (int)(sizeof(MathTab) / sizeof(double)); i += FT_THREADS)
    ((double*)tab)[i] = ((const double*)&g_mathtab)[i];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wn = warp & 3, wg = warp >> 2;               // warp: chains wn*16 .. +15, rows wg*64 .. +63 of each tile
  const int chain0 = blockIdx.x * FT_NC + wn * 16;
  const int chunk = blockIdx.y;

  // B fragments stay in registers for the whole kernel: beta[k0 + lane%4][chain0 + nt*8 + lane/4]
  double bf[NK][2];
#pragma unroll
  for (int ks = 0; ks < NK; ks++) {
#pragma unroll
    for (int nt = 0; nt < 2; nt++) {
      int k = ks * 4 + (lane & 3), c = chain0 + nt * 8 + (lane >> 2);
      bf[ks][nt] = (k < p && c < C) ? state[(size_t)(beta_slot + k) * C + c] : 0.0;
    }
  }
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  double csum[4] = {0.0, 0.0, 0.0, 0.0};
  const int tile_lo = tile_begin + chunk * tiles_per_chunk;
  const int tile_hi = min(tile_end, tile_lo + tiles_per_chunk);
  const uint32_t x_bytes = (uint32_t)(FT_M * P4 * sizeof(double)), y_bytes = (uint32_t)(FT_M * sizeof(double));

  if (tid == 0 && tile_lo < tile_hi) {
    mbar_expect_tx(&bar[0], x_bytes + y_bytes);
    bulk_g2s(stage0, Xpad + (size_t)tile_lo * FT_M * P4, x_bytes, &bar[0]);
    bulk_g2s(stage0 + FT_M * P4, ypad + (size_t)tile_lo * FT_M, y_bytes, &bar[0]);
  }
  for (int tile = tile_lo; tile < tile_hi; tile++) {
    const int it = tile - tile_lo, b = it & 1;
    if (tid == 0 && tile + 1 < tile_hi) {  // prefetch the next tile into the other stage (freed by the last __syncthreads)
      double* nxt = stage0 + (b ^ 1) * stage_doubles;
      mbar_expect_tx(&bar[b ^ 1], x_bytes + y_bytes);
      bulk_g2s(nxt, Xpad + (size_t)(tile + 1) * FT_M * P4, x_bytes, &bar[b ^ 1]);
      bulk_g2s(nxt + FT_M * P4, ypad + (size_t)(tile + 1) * FT_M, y_bytes, &bar[b ^ 1]);
    }
    mbar_wait(&bar[b], (uint32_t)((it >> 1) & 1));
    const double* Xs = stage0 + b * stage_doubles;
    const double* ys = Xs + FT_M * P4;
    const double* xa = Xs + (wg * 64 + (lane >> 2)) * P4 + (lane & 3);
#pragma unroll 1
    for (int mp = 0; mp < 4; mp++) {  // two 8-row m-tiles at a time -> 4 independent accumulator chains
      double acc[2][4] = {{0.0, 0.0, 0.0, 0.0}, {0.0, 0.0, 0.0, 0.0}};  // [m-tile][n-tile*2 + column]
      const double* x0 = xa + (mp * 16) * P4;
#pragma unroll
      for (int ks = 0; ks < NK; ks++) {
        double a0 = x0[ks * 4], a1 = x0[8 * P4 + ks * 4];
        dmma_m8n8k4(acc[0][0], acc[0][1], a0, bf[ks][0]);
        dmma_m8n8k4(acc[0][2], acc[0][3], a0, bf[ks][1]);
        dmma_m8n8k4(acc[1][0], acc[1][1], a1, bf[ks][0]);
        dmma_m8n8k4(acc[1][2], acc[1][3], a1, bf[ks][1]);
      }
      // epilogue on the accumulators: lane holds row (lane>>2) of each m-tile, columns 2*(lane&3)+{0,1} of each n-tile
#pragma unroll
      for (int mt = 0; mt < 2; mt++) {
        double yv = ys[wg * 64 + mp * 16 + mt * 8 + (lane >> 2)];
        bool live = yv >= 0.0, y1 = yv != 0.0;          // -1 marks the zero-padded rows past n
        double lp[4];
        bern_logit_lp_vec<4>(acc[mt], y1, lp, nullptr, tab);
#pragma unroll
        for (int j = 0; j < 4; j++) csum[j] += live ? lp[j] : 0.0;
      }
    }
    __syncthreads();  // everyone is done with stage b before it is refilled
  }

  // reduce over the 8 row-lanes of the warp (lanes with equal lane&3), then over the 2 row groups
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
      red[wg * FT_NC + wn * 16 + nt * 8 + 2 * (lane & 3) + 0] = csum[nt * 2 + 0];
      red[wg * FT_NC + wn * 16 + nt * 8 + 2 * (lane & 3) + 1] = csum[nt * 2 + 1];
    }
  }
  __syncthreads();
  if (tid < FT_NC) {
    int c = blockIdx.x * FT_NC + tid;
    if (c < C) partials[(size_t)chunk * C + c] = red[tid] + red[FT_NC + tid];
  }
}

// X [n][ld] (first p columns used) -> Xpad [ntiles*FT_M][P4], zero padded
__global__ void k_pad_x(const double* __restrict__ X, const double* __restrict__ y, int n, int p, int ld, int P4, size_t total,
                        double* __restrict__ Xpad, double* __restrict__ ypad) {
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  size_t i = gid / P4;
  int k = (int)(gid % P4);
  Xpad[gid] = (i < (size_t)n && k < p) ? X[i * ld + k] : 0.0;
  if (k == 0) ypad[i] = i < (size_t)n ? y[i] : -1.0;  // -1 marks padding rows
}

// k-steps are specialised at compile time: NK in {4, 8, 13, 16}; rows are padded to 4*NK doubles
static int nk_of(int p) { int nk = (p + 3) / 4; return nk <= 4 ? 4 : nk <= 8 ? 8 : nk <= 13 ? 13 : 16; }
static int p4_of(int p) { return 4 * nk_of(p); }

static int chunks_for(int n) {
  // node chunks: aim at ~2 waves of 2 resident blocks per SM when there are 16 chain tiles; the count only
  // depends on n so that the scratch size is known up front (stateless: models on several devices / threads share it)
  int ntiles = (n + FT_M - 1) / FT_M;
  return std::min(ntiles, 148);
}

int fused_bern_logit_tiles(int n) { return chunks_for(n); }

// doubles needed for the padded copy: X tiles [ntiles*FT_M][P4] followed by y tiles [ntiles*FT_M]
size_t fused_bern_logit_xpad_len(int n, int p) {
  if (p < 1 || p > FT_MAXP4) return 0;
  if (getenv("NIMB200_NO_FUSED")) return 0;
  size_t ntiles = (size_t)(n + FT_M - 1) / FT_M;
  return ntiles * FT_M * p4_of(p) + ntiles * FT_M;
}
static size_t xpad_x_len(int n, int p) { return (size_t)((n + FT_M - 1) / FT_M) * FT_M * p4_of(p); }

void fused_bern_logit_prepare_x(const double* X, const double* y, int n, int p, int ld, double* Xpad, cudaStream_t st) {
  int P4 = p4_of(p);
  size_t total = xpad_x_len(n, p);
  k_pad_x<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(X, y, n, p, ld, P4, total, Xpad, Xpad + total);
}

template <int NK>
static void launch_nk(const double* Xpad, const double* ypad, int n, int p, const double* state, int beta_slot, int C,
                      int chunks, int tpc, int tb, int te, double* partials, cudaStream_t st) {
  size_t smem = sizeof(double) * (2 * ((size_t)FT_M * 4 * NK + FT_M) + 2 * FT_NC) + 16 + sizeof(MathTab);
  static std::atomic<unsigned long long> attr_set{0};
  ensure_dyn_smem(k_fused_bern_logit<NK>, smem, attr_set);
  dim3 grid((C + FT_NC - 1) / FT_NC, chunks);
  k_fused_bern_logit<NK><<<grid, FT_THREADS, smem, st>>>(Xpad, ypad, n, p, state, beta_slot, C, tpc, tb, te, partials);
}

void fused_bern_logit_launch(const double* Xpad, int n, int p, const double* state, int beta_slot, int C, double* partials,
                             int tile_begin, int tile_end, cudaStream_t st) {
  int chunks = chunks_for(n);
  int ntiles = std::max(0, tile_end - tile_begin);
  int tpc = std::max(1, (ntiles + chunks - 1) / chunks);
  const double* ypad = Xpad + xpad_x_len(n, p);
  switch (nk_of(p)) {
    case 4: launch_nk<4>(Xpad, ypad, n, p, state, beta_slot, C, chunks, tpc, tile_begin, tile_end, partials, st); break;
    case 8: launch_nk<8>(Xpad, ypad, n, p, state, beta_slot, C, chunks, tpc, tile_begin, tile_end, partials, st); break;
    case 13: launch_nk<13>(Xpad, ypad, n, p, state, beta_slot, C, chunks, tpc, tile_begin, tile_end, partials, st); break;
    default: launch_nk<16>(Xpad, ypad, n, p, state, beta_slot, C, chunks, tpc, tile_begin, tile_end, partials, st); break;
  }
}

__global__ void k_debug_math(int which, const double* __restrict__ x, int n, double* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double v = x[i];
  out[i] = which == 0 ? exp_f64(v) : which == 1 ? log_f64(v) : which == 2 ? rcp_f64(v)
         : which == 3 ? bern_logit_lp(v, false) : bern_logit_lp(v, true);
}

}  // namespace nimb

// diagnostic: evaluates the kernels' own exp (0) / log (1) / reciprocal (2) / Bernoulli-logit log-density for y = 0 (3)
// and y = 1 (4) on a host array, so tests can measure their accuracy against libm.
extern "C" int nimb200_debug_math(int32_t which, const double* x, int32_t n, double* out) {
  if (!x || !out || n < 0 || which < 0 || which > 4) return 1;
  if (n == 0) return 0;
  double *dx = nullptr, *dout = nullptr;
  if (cudaMalloc(&dx, sizeof(double) * n) != cudaSuccess) return 1;
  if (cudaMalloc(&dout, sizeof(double) * n) != cudaSuccess) { cudaFree(dx); return 1; }
  cudaMemcpy(dx, x, sizeof(double) * n, cudaMemcpyHostToDevice);
  nimb::k_debug_math<<<(n + 255) / 256, 256>>>(which, dx, n, dout);
  cudaError_t e = cudaMemcpy(out, dout, sizeof(double) * n, cudaMemcpyDeviceToHost);
  cudaFree(dx);
  cudaFree(dout);
  return e == cudaSuccess ? 0 : 1;
}
