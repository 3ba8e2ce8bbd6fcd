// fused_norm.cuh -- streaming kernels for  y[i] ~ dnorm(A + v[G(i)], sd = S)  with A, S chain-level expressions and v a
// latent vector gathered through an index map (BASELINE config 3: linear mixed model, y[i] ~ dnorm(mu + u[grp[i]], sd = sig_e)).
// Included by engine.cu after the interpreter helpers.
//
// No transcendental per node: log(sd) and 1/sd^2 are chain-level, so per (node, chain) only the residual and a fused
// multiply-add remain and the pass is bound by reading y and the index array from HBM (12 bytes per node for up to 8
// chains per thread).  Block sum:  sum_i dnorm(y_i | m_i, s) = -n (log sqrt(2 pi) + log s) - 0.5 (sum_i d_i^2) / s^2,
// d_i = y_i - m_i, which differs from nmath's per-node -(C + 0.5 x^2 + log s), x = d/s, only in where the division by s
// happens (relative 1e-16 per term).  nmath's special cases are reproduced per chain from the reduced quantities:
// s < 0 -> NaN, s = Inf -> -Inf, s = 0 -> +Inf / -Inf / NaN by the count of exact zeros among the d_i, non-finite d -> -Inf
// or NaN through the arithmetic.  Conjugate contributions (MCMC_conjugacy.R:109-115, additive link, coeff = 1):
// contribution_mean = tau sum_i (y_i - offset_i), contribution_tau = n tau with offset_i = A(target = 0) + v[G(i)].
#pragma once
#include "tma_util.cuh"

namespace nimb {

struct NormArgs {
  const double* y;
  const int32_t* gidx;  // index array or null
  int mode, arg;        // index map (NIMB_IDX_*), mode < 0: no gathered term
  int v_slot;           // slot base of the gathered vector
  int a_off, a_len;     // chain-level program A
  int decl;
};

__device__ __forceinline__ int norm_index(const NormArgs& a, int i) {
  switch (a.mode) {
    case NIMB_IDX_I: return i;
    case NIMB_IDX_DIV: return i / a.arg;
    case NIMB_IDX_MOD: return i % a.arg;
    default: return a.gidx[i];
  }
}

static const int NF_THREADS = 256;
static const int NF_CT = 8;  // chains per thread

// grid (node blocks, chain groups of NF_CT).  MODE 0: out[blk][c] = sum d^2;  MODE 1: out[blk][c] = sum d with A evaluated
// under ov_slot := 0 (the conjugate sampler's offset probing).  Per (node, chain): one gathered load, two adds, one fma.
template <int MODE>
__global__ void __launch_bounds__(NF_THREADS)
k_norm_stream(DevTab T, NormArgs a, int lo, int hi, int nodes_per_block, const int32_t* idx, int ov_slot, double* __restrict__ partials) {
  __shared__ double red[NF_CT][NF_THREADS / 32];
  __shared__ double sA[NF_CT];
  const int c0 = blockIdx.y * NF_CT;
  if (threadIdx.x < NF_CT)  // the chain-level term once per block, not once per thread
    sA[threadIdx.x] = eval_prog(T, min(c0 + (int)threadIdx.x, T.C - 1), a.a_off, a.a_len, 0, 0, MODE == 1 ? ov_slot : -1, 0.0);
  __syncthreads();
  double A[NF_CT], s0[NF_CT];
  int co[NF_CT];
#pragma unroll
  for (int k = 0; k < NF_CT; k++) { A[k] = sA[k]; s0[k] = 0.0; co[k] = min(c0 + k, T.C - 1); }
  const int b_lo = lo + blockIdx.x * nodes_per_block, b_hi = min(hi, b_lo + nodes_per_block);
  const int slo = T.shard_lo[a.decl], shi = T.shard_hi[a.decl];
  const double* __restrict__ vbase = T.state + (size_t)a.v_slot * T.C;
  // four nodes per trip, their loads issued together: what matters for an HBM-bound pass is bytes in flight per thread
  // (prefetching the next trip as well was tried and lost to register pressure)
  for (int k0 = b_lo + threadIdx.x; k0 < b_hi; k0 += 4 * NF_THREADS) {
    double yv[4];
    int gi[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      int kk = k0 + u * NF_THREADS;
      bool ok = kk < b_hi;
      int node = ok ? (idx ? __ldg(idx + kk) : kk) : -1;
      if (idx && ok && (node < slo || node >= shi)) node = -1;
      yv[u] = node >= 0 ? __ldg(a.y + node) : 0.0;
      gi[u] = node < 0 ? -1 : (a.mode < 0 ? 0 : (a.mode == NIMB_IDX_IARR ? __ldg(a.gidx + node) : norm_index(a, node)));
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (gi[u] < 0) continue;
      if (a.mode >= 0) {
        const double* __restrict__ vp = vbase + (size_t)gi[u] * T.C;
#pragma unroll
        for (int k = 0; k < NF_CT; k++) {
          double d = yv[u] - (A[k] + vp[co[k]]);
          if (MODE == 0) s0[k] = fma(d, d, s0[k]); else s0[k] += d;
        }
      } else {
#pragma unroll
        for (int k = 0; k < NF_CT; k++) {
          double d = yv[u] - A[k];
          if (MODE == 0) s0[k] = fma(d, d, s0[k]); else s0[k] += d;
        }
      }
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NF_CT; k++) {
    double v0 = s0[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v0 += __shfl_xor_sync(0xffffffffu, v0, o);
    if (lane == 0) red[k][warp] = v0;
  }
  __syncthreads();
  if (threadIdx.x < NF_CT) {
    int k = threadIdx.x, c = c0 + k;
    if (c < T.C) {
      double s = 0.0;
      for (int w = 0; w < NF_THREADS / 32; w++) s += red[k][w];
      partials[(size_t)blockIdx.x * T.C + c] = s;
    }
  }
}

// TMA-staged form of k_norm_stream for contiguous node ranges: y (and the group-index array) travel HBM -> shared memory as
// 1-D bulk copies through a 4-stage ring of full / empty mbarriers (producer: thread 0; consumers: the 8 warps, each arriving on
// the stage's empty barrier when it is done, so there is no CTA-wide barrier per tile), so each CTA keeps 48 KB of loads in flight without spending registers on
// them; the threads only touch shared memory and the (L2-resident) gathered vector.  Same arithmetic as k_norm_stream.
static const int NT_TILE = 1024;   // nodes per stage
static const int NT_STAGES = 4;
template <int MODE>
__global__ void __launch_bounds__(NF_THREADS)
k_norm_stream_tma(DevTab T, NormArgs a, int lo, int hi, int nodes_per_block, int ov_slot, double* __restrict__ partials) {
  extern __shared__ __align__(128) unsigned char nsm[];
  double* ys = (double*)nsm;                                         // [NT_STAGES][NT_TILE]
  int32_t* gs = (int32_t*)(ys + NT_STAGES * NT_TILE);                // [NT_STAGES][NT_TILE]
  uint64_t* bar = (uint64_t*)(gs + NT_STAGES * NT_TILE);             // full[NT_STAGES] (TMA landed), empty[NT_STAGES] (all warps done)
  uint64_t* ebar = bar + NT_STAGES;
  double* red = (double*)(bar + 2 * NT_STAGES);                          // [NF_CT][NF_THREADS/32]
  double* sA = red + NF_CT * (NF_THREADS / 32);                      // [NF_CT]
  const bool use_idx = a.mode == NIMB_IDX_IARR;
  const int c0 = blockIdx.y * NF_CT;
  if (threadIdx.x < NF_CT)
    sA[threadIdx.x] = eval_prog(T, min(c0 + (int)threadIdx.x, T.C - 1), a.a_off, a.a_len, 0, 0, MODE == 1 ? ov_slot : -1, 0.0);
  if (threadIdx.x == 0) {
    for (int s = 0; s < NT_STAGES; s++) { tma::mbar_init(&bar[s], 1); tma::mbar_init(&ebar[s], NF_THREADS / 32); }
    tma::fence_init();
  }
  __syncthreads();
  double A[NF_CT], s0[NF_CT];
  int co[NF_CT];
#pragma unroll
  for (int k = 0; k < NF_CT; k++) { A[k] = sA[k]; s0[k] = 0.0; co[k] = min(c0 + k, T.C - 1); }
  const int b_lo = lo + blockIdx.x * nodes_per_block, b_hi = min(hi, b_lo + nodes_per_block);
  const int ntile = b_hi > b_lo ? (b_hi - b_lo + NT_TILE - 1) / NT_TILE : 0;
  const double* __restrict__ vbase = T.state + (size_t)a.v_slot * T.C;
  const bool vec = (T.C % 2) == 0 && c0 + NF_CT <= T.C && ((size_t)a.v_slot * T.C + c0) % 2 == 0;
  double mk[NF_CT];
  int gprev = -1;
#pragma unroll
  for (int k = 0; k < NF_CT; k++) mk[k] = 0.0;
  auto issue = [&](int tile) {  // thread 0 only: full tiles by TMA (the ragged last tile is loaded by the threads themselves)
    int n0 = b_lo + tile * NT_TILE;
    if (n0 + NT_TILE > b_hi) return;
    int s = tile % NT_STAGES;
    // the stage is free once every warp has arrived on its empty barrier for the previous use (no CTA-wide barrier per tile)
    if (tile >= NT_STAGES) tma::wait(&ebar[s], (uint32_t)((tile / NT_STAGES - 1) & 1));
    uint32_t bytes = NT_TILE * 8 + (use_idx ? NT_TILE * 4 : 0);
    tma::expect_tx(&bar[s], bytes);
    tma::bulk_g2s(ys + s * NT_TILE, a.y + n0, NT_TILE * 8, &bar[s]);
    if (use_idx) tma::bulk_g2s(gs + s * NT_TILE, a.gidx + n0, NT_TILE * 4, &bar[s]);
  };
  if (threadIdx.x == 0)
    for (int tl = 0; tl < NT_STAGES - 1 && tl < ntile; tl++) issue(tl);
  for (int tile = 0; tile < ntile; tile++) {
    const int s = tile % NT_STAGES, n0 = b_lo + tile * NT_TILE;
    const bool full = n0 + NT_TILE <= b_hi;
    if (threadIdx.x == 0 && tile + NT_STAGES - 1 < ntile) issue(tile + NT_STAGES - 1);  // refill the stage freed last trip
    if (full) tma::wait(&bar[s], (uint32_t)((tile / NT_STAGES) & 1));
    else {  // ragged last tile: every warp has consumed all earlier tiles once it is past this barrier
      __syncthreads();
      for (int q = threadIdx.x; q < b_hi - n0; q += NF_THREADS) { ys[s * NT_TILE + q] = a.y[n0 + q]; if (use_idx) gs[s * NT_TILE + q] = a.gidx[n0 + q]; }
      __syncthreads();
    }
    const int cnt = min(NT_TILE, b_hi - n0);
#pragma unroll
    for (int u = 0; u < NT_TILE / NF_THREADS; u++) {
      int q = u * NF_THREADS + threadIdx.x;
      if (q < cnt) {
        double yv = ys[s * NT_TILE + q];
        if (a.mode >= 0) {
          int g = use_idx ? gs[s * NT_TILE + q] : norm_index(a, n0 + q);
          if (g != gprev) {  // the per-chain means A + v[g] are kept while the thread stays in one group (sorted data: most trips)
            gprev = g;
            const double* __restrict__ vp = vbase + (size_t)g * T.C;
            if (vec) {  // the CTA's 8 chain copies of v[g] are 64 contiguous, 16-byte aligned bytes: four 128-bit loads
              const double2* __restrict__ v2 = reinterpret_cast<const double2*>(vp + c0);
              double2 w0 = v2[0], w1 = v2[1], w2 = v2[2], w3 = v2[3];
              mk[0] = A[0] + w0.x; mk[1] = A[1] + w0.y; mk[2] = A[2] + w1.x; mk[3] = A[3] + w1.y;
              mk[4] = A[4] + w2.x; mk[5] = A[5] + w2.y; mk[6] = A[6] + w3.x; mk[7] = A[7] + w3.y;
            } else {
#pragma unroll
              for (int k = 0; k < NF_CT; k++) mk[k] = A[k] + vp[co[k]];
            }
          }
#pragma unroll
          for (int k = 0; k < NF_CT; k++) {
            double d = yv - mk[k];
            if (MODE == 0) s0[k] = fma(d, d, s0[k]); else s0[k] += d;
          }
        } else {
#pragma unroll
          for (int k = 0; k < NF_CT; k++) {
            double d = yv - A[k];
            if (MODE == 0) s0[k] = fma(d, d, s0[k]); else s0[k] += d;
          }
        }
      }
    }
    if (full) {  // this warp is done with stage s
      __syncwarp();
      if ((threadIdx.x & 31) == 0) tma::arrive(&ebar[s]);
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NF_CT; k++) {
    double v0 = s0[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v0 += __shfl_xor_sync(0xffffffffu, v0, o);
    if (lane == 0) red[k * (NF_THREADS / 32) + warp] = v0;
  }
  __syncthreads();
  if (threadIdx.x < NF_CT) {
    int k = threadIdx.x, c = c0 + k;
    if (c < T.C) {
      double sm = 0.0;
      for (int w = 0; w < NF_THREADS / 32; w++) sm += red[k * (NF_THREADS / 32) + w];
      partials[(size_t)blockIdx.x * T.C + c] = sm;
    }
  }
}
static const size_t NT_SMEM = (size_t)NT_STAGES * NT_TILE * 12 + 2 * NT_STAGES * 8 + (NF_CT * (NF_THREADS / 32) + NF_CT) * 8;

// sd == 0 (degenerate): dnorm is +Inf where y == mean and -Inf elsewhere, so the sum depends on how many residuals are
// exactly zero.  Counted here by a plain loop, only for the chain that needs it.
__device__ double norm_sd0_sum(const DevTab& T, const NormArgs& a, int c, int lo, int hi, const int32_t* idx) {
  double A = eval_prog(T, c, a.a_off, a.a_len, 0, 0, -1, 0.0);
  long zeros = 0, n = 0;
  for (int k = lo; k < hi; k++) {
    int node = idx ? idx[k] : k;
    if (idx && (node < T.shard_lo[a.decl] || node >= T.shard_hi[a.decl])) continue;
    double m = a.mode >= 0 ? A + T.state[(size_t)(a.v_slot + norm_index(a, node)) * T.C + c] : A;
    zeros += (a.y[node] - m == 0.0);
    n++;
  }
  return n == 0 ? 0.0 : (zeros == n ? INFINITY : (zeros == 0 ? -INFINITY : NAN));
}

__device__ __forceinline__ double norm_sum_from_ss(double sigma, double S, int count) {
  double n = (double)count;
  if (count == 0) return 0.0;
  if (isnan(sigma) || isnan(S)) return sigma + S;
  if (sigma < 0) return NAN;
  if (!isfinite(sigma)) return -INFINITY;
  double q = S / (sigma * sigma);
  return !isfinite(q) ? -INFINITY : -(n * (NIMB_LN_SQRT_2PI + log(sigma)) + 0.5 * q);
}

// all-reduced path: per chain sum d^2 -> sum of log-densities (sd == 0 gives NaN here: the zero count is not exchanged)
__global__ void k_norm_finish(DevTab T, NormArgs a, int count, const double* __restrict__ SS, double* __restrict__ dst) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= T.C) return;
  const nimb200_decl& d = T.decl[a.decl];
  double sigma = eval_prog(T, c, d.prog_off[1], d.prog_len[1], 0, 0, -1, 0.0);
  dst[c] = (sigma == 0 && count > 0 && !isnan(SS[c])) ? NAN : norm_sum_from_ss(sigma, SS[c], count);
}

// single-rank path: reduce the block partials and finish in one launch (one warp per chain, fixed order)
template <int MODE>
__global__ void k_norm_reduce_finish(DevTab T, NormArgs a, int lo, int hi, const int32_t* idx, const double* __restrict__ partials,
                                     int nb, double* __restrict__ dst0, double* __restrict__ dst1) {
  int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (c >= T.C) return;
  double S = 0.0;
  for (int b = lane; b < nb; b += 32) S += partials[(size_t)b * T.C + c];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
  if (lane != 0) return;
  int count = hi - lo;
  if (MODE == 1) {
    double tau = get_param(T, a.decl, 0, c, NIMB_PARAM_TAU, -1, 0.0);
    dst0[c] = S * tau;
    dst1[c] = (double)count * tau;
    return;
  }
  const nimb200_decl& d = T.decl[a.decl];
  double sigma = eval_prog(T, c, d.prog_off[1], d.prog_len[1], 0, 0, -1, 0.0);
  dst0[c] = (sigma == 0 && count > 0 && !isnan(S)) ? norm_sd0_sum(T, a, c, lo, hi, idx) : norm_sum_from_ss(sigma, S, count);
}

// per chain: sum of residuals -> (contribution_mean, contribution_tau)
__global__ void k_norm_conj_finish(DevTab T, NormArgs a, int count, const double* __restrict__ R, double* __restrict__ dst0,
                                   double* __restrict__ dst1) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= T.C) return;
  double tau = get_param(T, a.decl, 0, c, NIMB_PARAM_TAU, -1, 0.0);
  dst0[c] = R[c] * tau;
  dst1[c] = (double)count * tau;
}

// conjugate normal-normal sweep over the gathered vector's elements (targets v[t]): one block per target.
// Phase 1: the block adds up y over the target's nodes (chain independent, fixed order); phase 2: one thread per chain
// forms contribution_mean = tau (sum y - n A), contribution_tau = n tau, draws and writes.
__global__ void __launch_bounds__(128)
k_norm_conj_sweep(DevTab T, int sid, uint64_t seed, uint64_t iter, NormArgs a) {
  __shared__ double red[4];
  __shared__ double s_sum;
  const nimb200_sampler sp = T.samplers[sid];
  const int t = blockIdx.x;
  const nimb200_dep dp = T.deps[sp.dep_off + 1];
  Sel s = sel_of(T, dp, t);
  double acc = 0.0;
  int cnt = 0;
  for (int k = s.lo + threadIdx.x; k < s.hi; k += 128) {
    int node = s.idx ? s.idx[k] : k;
    if (s.idx && (node < s.nlo || node >= s.nhi)) continue;
    acc += a.y[node];
    cnt++;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { acc += __shfl_xor_sync(0xffffffffu, acc, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o); }
  __shared__ int cred[4];
  if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = acc; cred[threadIdx.x >> 5] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) { s_sum = ((red[0] + red[1]) + red[2]) + red[3]; cred[0] = cred[0] + cred[1] + cred[2] + cred[3]; }
  __syncthreads();
  const double ysum = s_sum;
  const double n = (double)cred[0];
  const nimb200_dep tp = T.deps[sp.dep_off];
  const int tnode = (tp.sel == NIMB_SEL_SAME ? tp.a + t : tp.a);
  const int slot = sp.target_slot + t;
  for (int c = threadIdx.x; c < T.C; c += 128) {
    double A = eval_prog(T, c, a.a_off, a.a_len, 0, 0, -1, 0.0);
    double tau = get_param(T, a.decl, 0, c, NIMB_PARAM_TAU, -1, 0.0);
    double cm = (ysum - n * A) * tau, ct = n * tau;
    double prior_mean = get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
    double prior_tau = get_param(T, tp.decl, tnode, c, NIMB_PARAM_TAU, -1, 0.0);
    DrawKey key = make_key(T, seed, iter, sid, t, c);
    double mean = (prior_mean * prior_tau + cm) / (prior_tau + ct), sd = pow(prior_tau + ct, -0.5), z, z1;
    normal2(key, 0, z, z1);
    T.state[(size_t)slot * T.C + c] = mean + sd * z;
  }
}

// data-sharded form of the sweep: phase 1 leaves this rank's (sum y, count) per target in gs[t][2] (chain independent, so the
// all-reduce between the phases carries 2 doubles per target); phase 2 is one thread per (target, chain).
__global__ void __launch_bounds__(128)
k_norm_group_sums(DevTab T, int sid, NormArgs a, double* __restrict__ gs) {
  __shared__ double red[4];
  __shared__ int cred[4];
  const nimb200_sampler sp = T.samplers[sid];
  const int t = blockIdx.x;
  const nimb200_dep dp = T.deps[sp.dep_off + 1];
  Sel s = sel_of(T, dp, t);
  double acc = 0.0;
  int cnt = 0;
  for (int k = s.lo + threadIdx.x; k < s.hi; k += 128) {
    int node = s.idx ? s.idx[k] : k;
    if (s.idx && (node < s.nlo || node >= s.nhi)) continue;
    acc += a.y[node];
    cnt++;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { acc += __shfl_xor_sync(0xffffffffu, acc, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o); }
  if ((threadIdx.x & 31) == 0) { red[threadIdx.x >> 5] = acc; cred[threadIdx.x >> 5] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) {
    gs[2 * (size_t)t] = ((red[0] + red[1]) + red[2]) + red[3];
    gs[2 * (size_t)t + 1] = (double)(cred[0] + cred[1] + cred[2] + cred[3]);
  }
}

__global__ void k_norm_conj_draw(DevTab T, int sid, uint64_t seed, uint64_t iter, NormArgs a, const double* __restrict__ gs) {
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c = (int)(gid % T.C);
  size_t tq = gid / T.C;
  if (tq >= (size_t)sp.n_targets) return;
  int t = (int)tq;
  const double ysum = gs[2 * (size_t)t], n = gs[2 * (size_t)t + 1];
  const nimb200_dep tp = T.deps[sp.dep_off];
  const int tnode = (tp.sel == NIMB_SEL_SAME ? tp.a + t : tp.a);
  double A = eval_prog(T, c, a.a_off, a.a_len, 0, 0, -1, 0.0);
  double tau = get_param(T, a.decl, 0, c, NIMB_PARAM_TAU, -1, 0.0);
  double cm = (ysum - n * A) * tau, ct = n * tau;
  double prior_mean = get_param(T, tp.decl, tnode, c, 0, -1, 0.0);
  double prior_tau = get_param(T, tp.decl, tnode, c, NIMB_PARAM_TAU, -1, 0.0);
  DrawKey key = make_key(T, seed, iter, sid, t, c);
  double mean = (prior_mean * prior_tau + cm) / (prior_tau + ct), sd = pow(prior_tau + ct, -0.5), z, z1;
  normal2(key, 0, z, z1);
  T.state[(size_t)(sp.target_slot + t) * T.C + c] = mean + sd * z;
}

}  // namespace nimb
