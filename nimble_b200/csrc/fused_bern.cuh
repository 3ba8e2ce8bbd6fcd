// fused_bern.cuh -- fused kernels for  v[i] ~ dbern(gate[G(i)] * ilogit(c0 + c1 * x[i]))  (BASELINE config 4: occupancy model,
// z[i] ~ dbern(ilogit(b0 + b1 x[i])),  y[i,j] ~ dbern(z[i] * ilogit(a0 + a1 w[i,j])) ).  Included by engine.cu.
//
// The value v is data (y) or a latent 0/1 variable (z); the optional gate is a latent 0/1 variable gathered through an
// index map.  gate = 1: the product z*p is p itself, the log-density is the table-driven bern_logit_lp of bern_logit.cuh
// (same derivation from nmath's dbinom_raw as the logistic kernel); gate = 0: dbinom(v, 1, 0) = (v == 0 ? 0 : -Inf).
// Anything else (a value or gate that is not exactly 0 or 1) takes the exact generic path dbinom(v, 1, gate * ilogit(eta)).
// The binary Gibbs sweep (MCMC_samplers.R:127-151) over the sites is one thread per (site, chain): both states' prior and
// visit log-densities from one pass over the site's visits.
#pragma once
#include "bern_logit.cuh"

namespace nimb {

struct BernArgs {
  const double* x;
  const double* vdata;   // value if data, else null
  int v_slot;            // value slot base if latent
  int c0_slot, c1_slot;
  int gate_slot;         // -1: no gate
  const int32_t* gidx;
  int gmode, garg;
  int n;
};

__device__ __forceinline__ int bern_gate_index(const BernArgs& a, int i) {
  switch (a.gmode) {
    case NIMB_IDX_I: return i;
    case NIMB_IDX_DIV: return i / a.garg;
    case NIMB_IDX_MOD: return i % a.garg;
    default: return a.gidx[i];
  }
}

__device__ __forceinline__ double bernlin_lp(double v, double gate, double eta, const double* tab) {
  bool v01 = v == 0.0 || v == 1.0;
  if (v01 && gate == 1.0) return bern_logit_lp1(eta, v != 0.0, tab);
  if (v01 && gate == 0.0 && eta == eta) return v == 0.0 ? 0.0 : -INFINITY;
  return dbinom(v, 1.0, gate * (1. / (1. + exp(-eta))));  // exact generic path
}

static const int BF_THREADS = 256;
static const int BF_TILE = 256;

// block sums: grid (chain tiles of 32, node chunks); lane = chain, the 8 warps of a CTA share the staged node tile
// (warp w takes the w-th run of 32 consecutive nodes).  partials[chunk][C]
__global__ void __launch_bounds__(BF_THREADS)
k_bernlin_sum(DevTab T, BernArgs a, int lo, int hi, int nodes_per_chunk, double* __restrict__ partials) {
  __shared__ double sx[BF_TILE], sv[BF_TILE];
  __shared__ int sg[BF_TILE];
  __shared__ double red[BF_THREADS / 32][32];
  __shared__ __align__(16) double tab[SOFTPLUS_DOUBLES];
  load_softplus(tab, threadIdx.x, BF_THREADS);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = BF_THREADS / 32;
  const int c = blockIdx.x * 32 + lane;
  const bool live = c < T.C;
  const int cc = live ? c : T.C - 1;
  const double c0 = T.state[(size_t)a.c0_slot * T.C + cc], c1 = T.state[(size_t)a.c1_slot * T.C + cc];
  const int c_lo = lo + blockIdx.y * nodes_per_chunk, c_hi = min(hi, c_lo + nodes_per_chunk);
  double acc = 0.0;
  int gprev = -1;
  double gate = 1.0;
  for (int t0 = c_lo; t0 < c_hi; t0 += BF_TILE) {
    __syncthreads();
    int i = t0 + threadIdx.x;
    if (i < c_hi) {
      sx[threadIdx.x] = a.x[i];
      if (a.vdata) sv[threadIdx.x] = a.vdata[i];
      sg[threadIdx.x] = a.gate_slot >= 0 ? bern_gate_index(a, i) : 0;
    }
    __syncthreads();
    int cnt = min(BF_TILE, c_hi - t0);
    const int per = BF_TILE / nw, k_hi = min(cnt, (warp + 1) * per);
    for (int k = warp * per; k < k_hi; k++) {  // a contiguous run per warp: consecutive nodes mostly share the gate
      if (a.gate_slot >= 0 && sg[k] != gprev) { gprev = sg[k]; gate = T.state[(size_t)(a.gate_slot + gprev) * T.C + cc]; }
      double v = a.vdata ? sv[k] : T.state[(size_t)(a.v_slot + t0 + k) * T.C + cc];
      acc += bernlin_lp(v, gate, fma(c1, sx[k], c0), tab);
    }
  }
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && live) {
    double s = 0.0;
    for (int w = 0; w < nw; w++) s += red[w][lane];
    partials[(size_t)blockIdx.y * T.C + c] = s;
  }
}

// binary Gibbs sweep over latent 0/1 targets z[t] whose prior is  z[t] ~ dbern(ilogit(b0 + b1 xz[t]))  (pz) and whose dependents
// are the data nodes of  y ~ dbern(z[G] * ilogit(a0 + a1 w))  (py) selected by the sampler's stage-1 entry.
__global__ void __launch_bounds__(BF_THREADS)
k_bernlin_binary_sweep(DevTab T, int sid, uint64_t seed, uint64_t iter, BernArgs pz, BernArgs py, double* __restrict__ contrib_z,
                       double* __restrict__ contrib_y) {
  __shared__ __align__(16) double tab[SOFTPLUS_DOUBLES];
  load_softplus(tab, threadIdx.x, BF_THREADS);
  __syncthreads();
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c = (int)(gid % T.C);
  size_t tq = gid / T.C;
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq;
  double* sv = T.state + (size_t)(sp.target_slot + t) * T.C + c;
  const double cur = *sv, other = 1 - cur;
  const nimb200_dep d0 = T.deps[sp.dep_off], d1 = T.deps[sp.dep_off + 1];
  const int znode = (d0.sel == NIMB_SEL_SAME ? d0.a + t : d0.a);
  const double b0 = T.state[(size_t)pz.c0_slot * T.C + c], b1 = T.state[(size_t)pz.c1_slot * T.C + c];
  const double a0 = T.state[(size_t)py.c0_slot * T.C + c], a1 = T.state[(size_t)py.c1_slot * T.C + c];
  const double eta_z = fma(b1, pz.x[znode], b0);
  double curPrior = bernlin_lp(cur, 1.0, eta_z, tab), otherPrior = bernlin_lp(other, 1.0, eta_z, tab);
  double curDeps = 0.0, otherDeps = 0.0;
  Sel s = sel_of(T, d1, t);
  for (int k = s.lo; k < s.hi; k++) {
    int node = s.idx ? s.idx[k] : k;
    double eta = fma(a1, py.x[node], a0), yv = py.vdata[node];
    curDeps += bernlin_lp(yv, cur, eta, tab);
    otherDeps += bernlin_lp(yv, other, eta, tab);
  }
  double currentLogProb = curPrior + curDeps;
  double otherLogProb = (otherPrior == -INFINITY) ? otherPrior : otherPrior + otherDeps;
  double acceptanceProb = 1 / (exp(currentLogProb - otherLogProb) + 1);
  DrawKey key = make_key(T, seed, iter, sid, t, c);
  double u, u2;
  unif2(key, 0, u, u2);
  bool jump = (!isnan(acceptanceProb)) & (u < acceptanceProb);
  if (jump) *sv = other;
  // by-product: under the state it leaves behind, this site's share of the block log-density of its own declaration and of the
  // visits' declaration -- their fixed-order sums over the sites refresh the cached current-state block sums, so the RW steps on
  // the detection / occupancy coefficients that follow need no extra pass
  if (contrib_z) {
    contrib_z[(size_t)t * T.C + c] = jump ? otherPrior : curPrior;
    contrib_y[(size_t)t * T.C + c] = jump ? otherDeps : curDeps;
  }
}

// column sums of contrib[n][C] over chunks of rows, fixed order: partials[chunk][C] (lane = chain, the 8 warps of a block take
// consecutive eighths of the chunk in turn and meet in shared memory)
static const int CS_ROWS = 256;
__global__ void __launch_bounds__(256)
k_colsum(const double* __restrict__ contrib, int n, int C, double* __restrict__ partials) {
  __shared__ double red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  const int r0 = blockIdx.y * CS_ROWS, r1 = min(n, r0 + CS_ROWS);
  const int per = CS_ROWS / 8, lo = r0 + warp * per, hi = min(r1, lo + per);
  double a0 = 0.0, a1 = 0.0;
  if (c < C) {
    int r = lo;
    for (; r + 1 < hi; r += 2) { a0 += contrib[(size_t)r * C + c]; a1 += contrib[(size_t)(r + 1) * C + c]; }
    if (r < hi) a0 += contrib[(size_t)r * C + c];
  }
  red[warp][lane] = a0 + a1;
  __syncthreads();
  if (warp == 0 && c < C) {
    double s = 0.0;
    for (int w = 0; w < 8; w++) s += red[w][lane];
    partials[(size_t)blockIdx.y * C + c] = s;
  }
}

}  // namespace nimb
