// fused_pois.cuh -- fused kernels for  y[i] ~ dpois(lam[G(i)] * exp(b0 + b1 * x[i]))  (BASELINE config 5: Poisson GLMM).
// Included by engine.cu after the interpreter helpers (sel_of, get_param, make_key, rgamma1).
//
// The generic interpreter spends > 1000 instructions per evaluation on this declaration: the postfix program, CUDA's
// exp, and nmath's dpois_raw with its stirlerr table look-up (divergent constant-memory index), log(2 pi y) and the
// bd0 Taylor loop -- most of which depends on the data only.  Here
//   A[i] = -0.5 log(2 pi y[i]) - stirlerr(y[i])  is computed once per model (data only, same nmath functions),
//   mu   = lam * exp(b0 + b1 x)                   table-driven exp (fastmath.cuh),
//   lp   = y == 0 ? -mu : A - (y log(y / mu) + mu - y)      bd0 in its closed form (no series, no divergence).
// nmath evaluates bd0 by a Taylor series when y ~ mu to keep its RELATIVE accuracy; the closed form has absolute error
// ~ 2e-16 y, which relative to lp (|lp| >= |A| ~ 0.5 log(2 pi y) for y >= 1) is < 1e-13 for counts up to ~1e3:
// inside the 1e-12 parity bar (tests/test_gpu_parity.py::test_pois_family_*).  Anything unusual (lam <= 0, NaN,
// overflowing linear predictor, mu outside [1e-290, 1e290]) takes the exact generic path per element.
#pragma once

namespace nimb {

struct PoisArgs {
  const double* y;
  const double* x;
  const double* A;       // node constants
  const int32_t* gidx;   // index array (mode NIMB_IDX_IARR) or null
  int mode, arg;         // index map i -> lam element
  int lam_slot, b0_slot, b1_slot;
  int n;
};

__device__ __forceinline__ int pois_lam_index(const PoisArgs& a, int i) {
  switch (a.mode) {
    case NIMB_IDX_I: return i;
    case NIMB_IDX_DIV: return i / a.arg;
    case NIMB_IDX_MOD: return i % a.arg;
    default: return a.gidx[i];
  }
}

__device__ __forceinline__ double pois_lp(double y, double A, double lam, double eta, const MathTab* tab) {
  bool ok = lam > 0.0 && lam < 1e300 && eta > -700.0 && eta < 700.0;
  double mu = 0.0;
  if (ok) {
    mu = lam * exp_tab(eta, tab);
    ok = mu > 1e-290 && mu < 1e290;
  }
  if (!ok) return dpois(y, lam * exp(eta));  // exact generic path (rare)
  if (y == 0.0) return -mu;
  double q = y * rcp_f64(mu);
  return A - fma(y, log_tab(q, tab), mu - y);
}

// node constants A[i] = -0.5 log(2 pi y) - stirlerr(y), the data-only part of nmath's dpois_raw
__global__ void k_pois_nodeconst(const double* __restrict__ y, int n, double* __restrict__ A) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double v = y[i];
  A[i] = v > 0.0 ? -0.5 * log(NIMB_2PI * v) + (-stirlerr(v)) : 0.0;
}

static const int PF_THREADS = 256;  // chains per block
static const int PF_TILE = 1024;    // nodes staged per step

// block sums: grid (chain tiles of 32, node chunks); lane = chain (lam gathers coalesced over the warp), the 8 warps of a
// CTA share the staged node tile (warp w takes the w-th run of 32 consecutive nodes).  partials[chunk][C].
__global__ void __launch_bounds__(PF_THREADS)
k_pois_sum(DevTab T, PoisArgs a, int lo, int hi, int nodes_per_chunk, double* __restrict__ partials) {
  __shared__ double sy[PF_TILE], sx[PF_TILE], sA[PF_TILE];
  __shared__ int sg[PF_TILE];
  __shared__ double red[PF_THREADS / 32][32];
  __shared__ MathTab tab;
  load_mathtab(&tab, threadIdx.x, PF_THREADS);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = PF_THREADS / 32;
  const int c = blockIdx.x * 32 + lane;
  const bool live = c < T.C;
  const int cc = live ? c : T.C - 1;
  const double b0 = T.state[(size_t)a.b0_slot * T.C + cc], b1 = T.state[(size_t)a.b1_slot * T.C + cc];
  const int c_lo = lo + blockIdx.y * nodes_per_chunk, c_hi = min(hi, c_lo + nodes_per_chunk);
  double acc0 = 0.0, acc1 = 0.0;
  int gprev = -1;
  double lam = 0.0;
  for (int t0 = c_lo; t0 < c_hi; t0 += PF_TILE) {
    __syncthreads();
    for (int q = threadIdx.x; q < PF_TILE && t0 + q < c_hi; q += PF_THREADS) {
      int i = t0 + q;
      sy[q] = a.y[i]; sx[q] = a.x[i]; sA[q] = a.A[i]; sg[q] = pois_lam_index(a, i);
    }
    __syncthreads();
    int cnt = min(PF_TILE, c_hi - t0);
    const int per = PF_TILE / nw, k_hi = min(cnt, (warp + 1) * per);
    for (int k = warp * per; k < k_hi; k += 2) {  // a contiguous run per warp (consecutive nodes mostly share lam), two nodes per trip
      int g0 = sg[k];
      if (g0 != gprev) { lam = T.state[(size_t)(a.lam_slot + g0) * T.C + cc]; gprev = g0; }
      acc0 += pois_lp(sy[k], sA[k], lam, fma(b1, sx[k], b0), &tab);
      if (k + 1 < k_hi) {
        int g1 = sg[k + 1];
        if (g1 != gprev) { lam = T.state[(size_t)(a.lam_slot + g1) * T.C + cc]; gprev = g1; }
        acc1 += pois_lp(sy[k + 1], sA[k + 1], lam, fma(b1, sx[k + 1], b0), &tab);
      }
    }
  }
  red[warp][lane] = acc0 + acc1;
  __syncthreads();
  if (warp == 0 && live) {
    double s = 0.0;
    for (int w = 0; w < nw; w++) s += red[w][lane];
    partials[(size_t)blockIdx.y * T.C + c] = s;
  }
}

// conjugate gamma-Poisson sweep over the lam[g] targets (MCMC_conjugacy.R:80-93: contribution_shape = value,
// contribution_rate = coeff, coeff = the dependent's lambda with the target set to 1 = exp(b0 + b1 x)).
// thread = (target g, chain c), chains fastest.
__global__ void __launch_bounds__(PF_THREADS)
k_pois_conj_sweep(DevTab T, int sid, uint64_t seed, uint64_t iter, PoisArgs a) {
  __shared__ MathTab tab;
  load_mathtab(&tab, threadIdx.x, PF_THREADS);
  __syncthreads();
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c = (int)(gid % T.C);
  size_t tq = gid / T.C;
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq;
  const double b0 = T.state[(size_t)a.b0_slot * T.C + c], b1 = T.state[(size_t)a.b1_slot * T.C + c];
  const nimb200_dep dp = T.deps[sp.dep_off + 1];
  Sel s = sel_of(T, dp, t);
  double c0 = 0.0, c1 = 0.0;
  bool exact = !(b0 == b0 && b1 == b1);
  for (int k = s.lo; k < s.hi; k++) {
    int node = s.idx ? s.idx[k] : k;
    double eta = fma(b1, a.x[node], b0);
    exact |= !(eta > -700.0 && eta < 700.0);
    c0 += a.y[node];
    c1 += exp_tab(fmin(fmax(eta, -700.0), 700.0), &tab);
  }
  if (exact) {  // overflowing / NaN linear predictor somewhere: redo with the library exp
    c1 = 0.0;
    for (int k = s.lo; k < s.hi; k++) {
      int node = s.idx ? s.idx[k] : k;
      c1 += 1.0 * exp(b0 + b1 * a.x[node]);
    }
  }
  const nimb200_dep tp = T.deps[sp.dep_off];
  int tnode = (tp.sel == NIMB_SEL_SAME ? tp.a + t : tp.a);
  double prior_shape = get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
  double prior_rate = get_param(T, tp.decl, tnode, c, NIMB_PARAM_RATE, -1, 0.0);
  DrawKey key = make_key(T, seed, iter, sid, t, c);
  double shape = prior_shape + c0, scale = 1 / (prior_rate + c1);
  T.state[(size_t)(sp.target_slot + t) * T.C + c] = rgamma1(key, shape) * scale;
}

}  // namespace nimb
