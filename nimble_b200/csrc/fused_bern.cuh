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

// scalar form: the exact generic path for anything irregular, the table-driven form otherwise (per lane, no warp vote)
__device__ __forceinline__ double bernlin_lp(double v, double gate, double eta, const MathTab* tab) {
  bool v01 = v == 0.0 || v == 1.0;
  if (v01 && gate == 1.0) {
    if (!(eta >= -690.0)) return bern_logit_lp_slow(eta, v != 0.0);
    double in[1] = {eta}, o[1];
    bern_logit_lp_vec<1, false, false>(in, v != 0.0, o, nullptr, tab);
    return o[0];
  }
  if (v01 && gate == 0.0 && eta == eta) return v == 0.0 ? 0.0 : -INFINITY;
  return dbinom(v, 1.0, gate * (1. / (1. + exp(-eta))));  // exact generic path
}

// out of line: the rare branches of the kernels below call it, and the hot loops stay small in the instruction cache
static __device__ __noinline__ double bernlin_lp_cold(double v, double gate, double eta, const MathTab* tab) {
  return bernlin_lp(v, gate, eta, tab);
}

static const int BF_THREADS = 256;
static const int BF_TILE = 1024;  // 16 KiB of staged nodes (with the 41 KiB of tables three CTAs still fit an SM)
static const size_t BF_SUM_SMEM = sizeof(double) * 2 * BF_TILE + BL_REP_BYTES;   // dynamic shared memory of k_bernlin_sum (BernNode = 16 bytes)
static const int BF_W = 4;   // nodes in flight per lane: four independent exp / log chains keep the FP64 pipe fed

// one staged node: 16 bytes (code: the data value 0 / 1, 2 = anything else)
struct __align__(16) BernNode {
  double x;
  int code, g;
};

// is the double exactly 0.0 or 1.0?  On its bits (integer pipe): nonzero when it is neither
__device__ __forceinline__ unsigned not01_bits(double v) {
  const unsigned h = (unsigned)__double2hiint(v);
  return (unsigned)__double2loint(v) | min(h, h ^ 0x3ff00000u);
}

// block sums: grid (chain tiles of 32, node chunks); lane = chain, the 8 warps of a CTA share the staged node tile
// (warp w takes the w-th run of 256 consecutive nodes, BF_W at a time).  partials[chunk][C].
// Round 2: (1) the softplus table cost 80 bytes of divergent shared-memory reads per evaluation and made this kernel
// shared-memory-bound (ncu round 1: short_scoreboard 7.5 per issue, 87 instructions per evaluation); the exp / log-table form
// of the fused logistic kernel reads 24 bytes, and BF_W evaluations per lane travel through it together.  (2) The per-chain
// operands (gate = the latent site state, or the latent value itself) are loaded one trip ahead: their global-memory latency
// was the top stall (long_scoreboard on the first use, 34 % of the samples).  (3) One instantiation per (latent, gated)
// combination; the value / gate / predictor checks are integer tests on the bits, ORed into one flag per trip.
template <bool LATENT, bool GATED>
__global__ void __launch_bounds__(BF_THREADS, 3)
k_bernlin_sum(DevTab T, BernArgs a, int lo, int hi, int nodes_per_chunk, double* __restrict__ partials) {
  extern __shared__ __align__(16) unsigned char bf_smem[];   // staged nodes | replicated tables (BL_REP_BYTES)
  BernNode* sn = reinterpret_cast<BernNode*>(bf_smem);
  double* rep = reinterpret_cast<double*>(bf_smem + sizeof(BernNode) * BF_TILE);
  __shared__ double red[BF_THREADS / 32][32];
  __shared__ __align__(16) MathTab tab;
  load_mathtab(&tab, threadIdx.x, BF_THREADS);
  load_bern_rep_tables(rep, threadIdx.x, BF_THREADS);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = BF_THREADS / 32;
  const double* __restrict__ expL = rep + (lane & 15);
  const int c = blockIdx.x * 32 + lane;
  const bool live = c < T.C;
  const int cc = live ? c : T.C - 1;
  const size_t C = (size_t)T.C;
  const double c0 = T.state[(size_t)a.c0_slot * C + cc], c1 = T.state[(size_t)a.c1_slot * C + cc];
  const int c_lo = lo + blockIdx.y * nodes_per_chunk, c_hi = min(hi, c_lo + nodes_per_chunk);
  const double* __restrict__ gatep = T.state + (size_t)(GATED ? a.gate_slot : 0) * C + cc;
  const double* __restrict__ valp = T.state + (size_t)(LATENT ? a.v_slot : 0) * C + cc;
  double acc = 0.0;
  for (int t0 = c_lo; t0 < c_hi; t0 += BF_TILE) {
    __syncthreads();
    for (int q = threadIdx.x; q < BF_TILE; q += BF_THREADS) {
      const int i = t0 + q;
      if (i < c_hi) {
        BernNode nd;
        nd.x = a.x[i];
        const double v = LATENT ? 0.0 : a.vdata[i];
        nd.code = v == 0.0 ? 0 : v == 1.0 ? 1 : 2;
        nd.g = GATED ? bern_gate_index(a, i) : 0;
        sn[q] = nd;
      }
    }
    __syncthreads();
    const int cnt = min(BF_TILE, c_hi - t0);
    const int per = ((cnt + nw - 1) / nw + BF_W - 1) & ~(BF_W - 1);   // a partial tile is still shared by all warps
    const int k_lo = min(cnt, warp * per), k_hi = min(cnt, (warp + 1) * per);
    const int k_full = k_lo + (max(0, k_hi - k_lo) & ~(BF_W - 1));
    double gn[BF_W], vn[BF_W];   // the next trip's per-chain operands
#pragma unroll
    for (int j = 0; j < BF_W; j++) { gn[j] = 1.0; vn[j] = 0.0; }
    auto fetch = [&](int k) {
#pragma unroll
      for (int j = 0; j < BF_W; j++) {
        if (GATED) gn[j] = gatep[(size_t)sn[k + j].g * C];
        if (LATENT) vn[j] = valp[(size_t)(t0 + k + j) * C];
      }
    };
    if (k_lo < k_full) fetch(k_lo);
    for (int k = k_lo; k < k_full; k += BF_W) {
      double eta[BF_W], gt[BF_W], vv[BF_W], lp[BF_W];
      bool y1[BF_W];
      unsigned irr = 0;
#pragma unroll
      for (int j = 0; j < BF_W; j++) { gt[j] = gn[j]; vv[j] = vn[j]; }
      if (k + BF_W < k_full) fetch(k + BF_W);
#pragma unroll
      for (int j = 0; j < BF_W; j++) {
        const BernNode nd = sn[k + j];
        eta[j] = fma(c1, nd.x, c0);
        // irregular: a value or gate that is not exactly 0 or 1, a linear predictor that is NaN or infinite
        irr |= (unsigned)((__double2hiint(eta[j]) & 0x7ff00000) == 0x7ff00000);
        if (LATENT) { y1[j] = __double2hiint(vv[j]) != 0; irr |= not01_bits(vv[j]); }
        else { y1[j] = nd.code == 1; irr |= (unsigned)(nd.code & 2); }
        if (GATED) irr |= not01_bits(gt[j]);
      }
      if (__any_sync(0xffffffffu, irr != 0)) {  // the scalar form sorts every case out (warp-uniform branch, practically never taken)
#pragma unroll
        for (int j = 0; j < BF_W; j++) acc += bernlin_lp_cold(LATENT ? vv[j] : a.vdata[t0 + k + j], gt[j], eta[j], &tab);
        continue;
      }
      bern_logit_lp_rep<BF_W, true>(eta, y1, lp, expL, &tab);
#pragma unroll
      for (int j = 0; j < BF_W; j++) {
        if (GATED) acc += __double2hiint(gt[j]) != 0 ? lp[j] : (y1[j] ? -INFINITY : 0.0);   // gate 0: dbinom(v, 1, 0)
        else acc += lp[j];
      }
    }
    for (int k = k_full; k < k_hi; k++) {   // at most BF_W - 1 nodes at the end of a chunk
      const BernNode nd = sn[k];
      const double v = LATENT ? valp[(size_t)(t0 + k) * C] : a.vdata[t0 + k];
      acc += bernlin_lp_cold(v, GATED ? gatep[(size_t)nd.g * C] : 1.0, fma(c1, nd.x, c0), &tab);
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

// The same block sum for a DATA declaration (values staged as codes), written for the issue slots: per evaluation the loop
// holds one 16-byte node read, the gate's address and load (one trip ahead), four integer instructions of gate validation, the
// moderate path's arithmetic and one select.  What the general kernel above checks per evaluation is established per tile
// (values that are not 0 / 1: the whole tile goes through the scalar form) or per trip by the vote the moderate path needs anyway
// (a predictor that is NaN, infinite or beyond 8 sends the trip through the scalar form, which also knows dbinom(v, 1, 0 * NaN)).
template <bool GATED>
__global__ void __launch_bounds__(BF_THREADS, 3)
k_bernlin_sum_data(DevTab T, BernArgs a, int lo, int hi, int nodes_per_chunk, double* __restrict__ partials) {
  extern __shared__ __align__(16) unsigned char bf_smem[];   // staged nodes | replicated tables (BL_REP_BYTES)
  BernNode* sn = reinterpret_cast<BernNode*>(bf_smem);
  double* rep = reinterpret_cast<double*>(bf_smem + sizeof(BernNode) * BF_TILE);
  __shared__ double red[BF_THREADS / 32][32];
  __shared__ __align__(16) MathTab tab;
  load_mathtab(&tab, threadIdx.x, BF_THREADS);
  load_bern_rep_tables(rep, threadIdx.x, BF_THREADS);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = BF_THREADS / 32;
  const double* __restrict__ expL = rep + (lane & 15);
  const int c = blockIdx.x * 32 + lane;
  const bool live = c < T.C;
  const int cc = live ? c : T.C - 1;
  const size_t C = (size_t)T.C;
  const double c0 = T.state[(size_t)a.c0_slot * C + cc], c1 = T.state[(size_t)a.c1_slot * C + cc];
  const int c_lo = lo + blockIdx.y * nodes_per_chunk, c_hi = min(hi, c_lo + nodes_per_chunk);
  const double* __restrict__ gatep = T.state + (size_t)(GATED ? a.gate_slot : 0) * C + cc;
  double acc = 0.0;
  for (int t0 = c_lo; t0 < c_hi; t0 += BF_TILE) {
    __syncthreads();
    int odd = 0;
    for (int q = threadIdx.x; q < BF_TILE; q += BF_THREADS) {
      const int i = t0 + q;
      if (i < c_hi) {
        BernNode nd;
        nd.x = a.x[i];
        const double v = a.vdata[i];
        nd.code = v == 0.0 ? 0 : v == 1.0 ? 1 : 2;
        nd.g = GATED ? bern_gate_index(a, i) : 0;
        odd |= nd.code & 2;
        sn[q] = nd;
      }
    }
    const int tile_odd = __syncthreads_or(odd);   // a value that is neither 0 nor 1 somewhere in the tile
    const int cnt = min(BF_TILE, c_hi - t0);
    const int per = ((cnt + nw - 1) / nw + BF_W - 1) & ~(BF_W - 1);
    const int k_lo = min(cnt, warp * per), k_hi = min(cnt, (warp + 1) * per);
    const int k_full = tile_odd ? k_lo : k_lo + (max(0, k_hi - k_lo) & ~(BF_W - 1));
    double gn[BF_W];   // the next trip's gates
#pragma unroll
    for (int j = 0; j < BF_W; j++) gn[j] = 1.0;
    auto fetch = [&](int k) {
#pragma unroll
      for (int j = 0; j < BF_W; j++) gn[j] = gatep[(size_t)sn[k + j].g * C];
    };
    if (GATED && k_lo < k_full) fetch(k_lo);
    for (int k = k_lo; k < k_full; k += BF_W) {
      double eta[BF_W], gt[BF_W], lp[BF_W];
      bool y1[BF_W];
      unsigned irr = 0;
#pragma unroll
      for (int j = 0; j < BF_W; j++) {
        const BernNode nd = sn[k + j];
        eta[j] = fma(c1, nd.x, c0);
        y1[j] = nd.code != 0;
        gt[j] = gn[j];
        if (GATED) irr |= not01_bits(gt[j]);
      }
      if (GATED && k + BF_W < k_full) fetch(k + BF_W);
      if (!__all_sync(0xffffffffu, irr == 0 && bern_logit_moderate<BF_W>(eta))) {   // practically never: the scalar form sorts it out
#pragma unroll
        for (int j = 0; j < BF_W; j++) acc += bernlin_lp_cold(y1[j] ? 1.0 : 0.0, gt[j], eta[j], &tab);
        continue;
      }
      bern_logit_lp_rep_mod<BF_W, 16>(eta, y1, lp, expL);
#pragma unroll
      for (int j = 0; j < BF_W; j++) {
        if (GATED) acc += __double2hiint(gt[j]) != 0 ? lp[j] : (y1[j] ? -INFINITY : 0.0);   // gate 0: dbinom(v, 1, 0)
        else acc += lp[j];
      }
    }
    for (int k = k_full; k < k_hi; k++) {   // the end of a chunk, or a tile that holds an odd value
      const BernNode nd = sn[k];
      acc += bernlin_lp_cold(a.vdata[t0 + k], GATED ? gatep[(size_t)nd.g * C] : 1.0, fma(c1, nd.x, c0), &tab);
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
// With z = 0 a visit contributes 0 (y = 0) or -Inf (y = 1): only the z = 1 state needs the link, two visits at a time.
__global__ void __launch_bounds__(BF_THREADS)
k_bernlin_binary_sweep(DevTab T, int sid, uint64_t seed, uint64_t iter, BernArgs pz, BernArgs py, double* __restrict__ contrib_z,
                       double* __restrict__ contrib_y) {
  __shared__ __align__(16) MathTab tab;   // plain tables: a CTA serves 256 (site, chain) pairs, too few to pay for 41 KiB of copies
  load_mathtab(&tab, threadIdx.x, BF_THREADS);
  __syncthreads();
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c;
  size_t tq;
  split_index(gid, T.C, tq, c);
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq;
  double* sv = T.state + (size_t)(sp.target_slot + t) * T.C + c;
  const double cur = *sv, other = 1 - cur;
  const nimb200_dep d0 = T.deps[sp.dep_off], d1 = T.deps[sp.dep_off + 1];
  const int znode = (d0.sel == NIMB_SEL_SAME ? d0.a + t : d0.a);
  const double b0 = T.state[(size_t)pz.c0_slot * T.C + c], b1 = T.state[(size_t)pz.c1_slot * T.C + c];
  const double a0 = T.state[(size_t)py.c0_slot * T.C + c], a1 = T.state[(size_t)py.c1_slot * T.C + c];
  const double eta_z = fma(b1, pz.x[znode], b0);
  const Sel s = sel_of(T, d1, t);
  double curPrior, otherPrior, curDeps = 0.0, otherDeps = 0.0;
  bool fast = not01_bits(cur) == 0 && (__double2hiint(eta_z) & 0x7ff00000) != 0x7ff00000;
  if (fast) {
    const bool z1 = __double2hiint(cur) != 0;
    {  // the prior under z = 1 and z = 0: one predictor, both values
      const double e2[2] = {eta_z, eta_z};
      const bool y2[2] = {true, false};
      double l2[2];
      bern_logit_lp_vecy<2, false, false>(e2, y2, l2, nullptr, &tab);
      curPrior = z1 ? l2[0] : l2[1];
      otherPrior = z1 ? l2[1] : l2[0];
    }
    // the visits' log-density with z = 1 (sum1, in visit order); with z = 0 it is 0, or -Inf once a visit detected the species.
    // Two visits per trip, the next trip's (x, y) loaded before this trip's arithmetic (their latency was the top stall).
    double sum1 = 0.0, xn[2], yn[2];
    bool seen = false;
    auto fetch = [&](int k) {
#pragma unroll
      for (int j = 0; j < 2; j++) {
        const int kk = min(k + j, s.hi - 1);
        const int node = s.idx ? s.idx[kk] : kk;
        xn[j] = __ldg(py.x + node);
        yn[j] = __ldg(py.vdata + node);
      }
    };
    if (s.lo < s.hi) fetch(s.lo);
    for (int k = s.lo; k < s.hi; k += 2) {
      double eta[2], lp[2];
      bool y1[2];
      unsigned irr = 0;
#pragma unroll
      for (int j = 0; j < 2; j++) {
        eta[j] = fma(a1, xn[j], a0);
        y1[j] = __double2hiint(yn[j]) != 0;
        irr |= not01_bits(yn[j]) | (unsigned)((__double2hiint(eta[j]) & 0x7ff00000) == 0x7ff00000);
      }
      if (k + 2 < s.hi) fetch(k + 2);
      if (irr) { fast = false; break; }
      bern_logit_lp_vecy<2, false, false>(eta, y1, lp, nullptr, &tab);
      sum1 += lp[0];
      seen |= y1[0];
      if (k + 1 < s.hi) { sum1 += lp[1]; seen |= y1[1]; }
    }
    const double sum0 = seen ? -INFINITY : 0.0;
    curDeps = z1 ? sum1 : sum0;
    otherDeps = z1 ? sum0 : sum1;
  }
  if (!fast) {  // anything irregular: visit by visit through the scalar form
    curPrior = bernlin_lp_cold(cur, 1.0, eta_z, &tab);
    otherPrior = bernlin_lp_cold(other, 1.0, eta_z, &tab);
    curDeps = 0.0; otherDeps = 0.0;
    for (int k = s.lo; k < s.hi; k++) {
      int node = s.idx ? s.idx[k] : k;
      double eta = fma(a1, py.x[node], a0), yv = py.vdata[node];
      curDeps += bernlin_lp_cold(yv, cur, eta, &tab);
      otherDeps += bernlin_lp_cold(yv, other, eta, &tab);
    }
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
