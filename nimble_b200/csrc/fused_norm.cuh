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
  int a_slot;           // >= 0: A is just this state slot (read directly, no interpreter)
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

// TMA-staged form of k_norm_stream for contiguous node ranges.  Every warp is autonomous: it owns a contiguous run of nodes
// and a private 3-stage ring of 256-node stages (y and the group-index array arrive as 1-D bulk copies, completion on the
// stage's mbarrier); lane 0 refills a stage right after the warp's __syncwarp, so the main loop has no cross-warp
// synchronisation at all and a CTA keeps 48 KB of loads in flight without spending registers on them (3 CTAs per SM).
// Because a warp walks consecutive nodes, the per-chain means A + v[g] stay in registers while it remains inside one
// group (sorted data: a gather every ~4 stages).  IDX: 0 = no gathered term, 1 = index array (staged), 2 = computed index.  Same arithmetic as
// k_norm_stream.
static const int NW_TILE = 256;                        // nodes per stage
static const int NW_STAGES = 3;
static const int NW_STAGE_BYTES = NW_TILE * 12;        // y [256] doubles, then g [256] ints
static const int NW_WARPS = NF_THREADS / 32;
static const size_t NW_SMEM = (size_t)NW_WARPS * NW_STAGES * NW_STAGE_BYTES + NW_WARPS * NW_STAGES * 8 + (NF_CT * NW_WARPS + NF_CT) * 8;
template <int MODE, int IDX>
__global__ void __launch_bounds__(NF_THREADS, 3)
k_norm_stream_tma(DevTab T, NormArgs a, int lo, int hi, int nodes_per_warp, int ov_slot, double* __restrict__ partials) {
  extern __shared__ __align__(128) unsigned char nsm[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned char* ring = nsm + warp * (NW_STAGES * NW_STAGE_BYTES);
  uint64_t* bar = (uint64_t*)(nsm + NW_WARPS * NW_STAGES * NW_STAGE_BYTES) + warp * NW_STAGES;
  double* red = (double*)(nsm + NW_WARPS * NW_STAGES * NW_STAGE_BYTES + NW_WARPS * NW_STAGES * 8);  // [NF_CT][NW_WARPS]
  double* sA = red + NF_CT * NW_WARPS;                                                              // [NF_CT]
  const int c0 = blockIdx.y * NF_CT;
  const long long w0 = (long long)lo + ((long long)blockIdx.x * NW_WARPS + warp) * nodes_per_warp;
  const int w_lo = (int)min((long long)hi, w0), w_hi = (int)min((long long)hi, w0 + nodes_per_warp);
  const int ntile = (w_hi - w_lo + NW_TILE - 1) / NW_TILE;
  auto issue = [&](int tile) {  // lane 0: whole and 4-node-multiple stages by TMA (a ragged global tail is read directly)
    const int n0 = w_lo + tile * NW_TILE, cnt = min(NW_TILE, w_hi - n0);
    if (cnt & 3) return;
    unsigned char* st = ring + (tile % NW_STAGES) * NW_STAGE_BYTES;
    uint64_t* b = &bar[tile % NW_STAGES];
    tma::expect_tx(b, (uint32_t)(cnt * 8 + (IDX == 1 ? cnt * 4 : 0)));
    tma::bulk_g2s(st, a.y + n0, (uint32_t)(cnt * 8), b);
    if (IDX == 1) tma::bulk_g2s(st + NW_TILE * 8, a.gidx + n0, (uint32_t)(cnt * 4), b);
  };
  cudaTriggerProgrammaticLaunchCompletion();  // lets the reduce / finish launch behind this kernel be scheduled as CTAs retire
  // the first copies start before anything else; warp 1 meanwhile evaluates the chain-level term (dependent global loads)
  if (lane == 0) {
    for (int s = 0; s < NW_STAGES; s++) tma::mbar_init(&bar[s], 1);
    tma::fence_init();
    for (int tl = 0; tl < NW_STAGES - 1 && tl < ntile; tl++) issue(tl);
  }
  if (warp == 1 && lane < NF_CT) {
    const int c = min(c0 + lane, T.C - 1);
    if (a.a_slot >= 0) sA[lane] = (MODE == 1 && a.a_slot == ov_slot) ? 0.0 : T.state[(size_t)a.a_slot * T.C + c];
    else sA[lane] = eval_prog(T, c, a.a_off, a.a_len, 0, 0, MODE == 1 ? ov_slot : -1, 0.0);
  }
  __syncthreads();
  double s0[NF_CT], mk[NF_CT];  // the chain-level terms stay in shared memory: they are only needed when the group changes
#pragma unroll
  for (int k = 0; k < NF_CT; k++) { mk[k] = sA[k]; s0[k] = 0.0; }
  const double* __restrict__ vb = T.state + (size_t)a.v_slot * T.C + c0;
  const bool vec = (T.C % 2) == 0 && c0 + NF_CT <= T.C && ((size_t)a.v_slot * T.C + c0) % 2 == 0;
  int gprev = -1;
  auto step = [&](double yv, int g) {
    if (IDX != 0 && g != gprev) {  // new group: the CTA's 8 chain copies of v[g] are 64 contiguous bytes
      gprev = g;
      const double* __restrict__ vp = vb + (long long)g * (long long)T.C;
      if (vec) {
        const double2* __restrict__ v2 = reinterpret_cast<const double2*>(vp);
        double2 q0 = v2[0], q1 = v2[1], q2 = v2[2], q3 = v2[3];
        mk[0] = sA[0] + q0.x; mk[1] = sA[1] + q0.y; mk[2] = sA[2] + q1.x; mk[3] = sA[3] + q1.y;
        mk[4] = sA[4] + q2.x; mk[5] = sA[5] + q2.y; mk[6] = sA[6] + q3.x; mk[7] = sA[7] + q3.y;
      } else {
#pragma unroll
        for (int k = 0; k < NF_CT; k++) mk[k] = sA[k] + vp[min(k, T.C - 1 - c0)];
      }
    }
#pragma unroll
    for (int k = 0; k < NF_CT; k++) {
      double d = yv - mk[k];
      if (MODE == 0) s0[k] = fma(d, d, s0[k]); else s0[k] += d;
    }
  };
  for (int tile = 0; tile < ntile; tile++) {
    const int s = tile % NW_STAGES, n0 = w_lo + tile * NW_TILE, cnt = min(NW_TILE, w_hi - n0);
    // stage (tile - 1) % NW_STAGES was released by the __syncwarp that ended the previous trip
    if (lane == 0 && tile + NW_STAGES - 1 < ntile) issue(tile + NW_STAGES - 1);
    if (!(cnt & 3)) {
      tma::wait(&bar[s], (uint32_t)((tile / NW_STAGES) & 1));
      const unsigned char* st = ring + s * NW_STAGE_BYTES;
      if (cnt == NW_TILE) {  // lane: nodes 64 j + 2 lane + {0, 1}, j = 0..3 (conflict-free 128- / 64-bit shared loads)
#pragma unroll
        for (int j = 0; j < NW_TILE / 64; j++) {
          const double2 yy = reinterpret_cast<const double2*>(st)[j * 32 + lane];
          int2 gg = make_int2(0, 0);
          if (IDX == 1) gg = reinterpret_cast<const int2*>(st + NW_TILE * 8)[j * 32 + lane];
          else if (IDX == 2) gg = make_int2(norm_index(a, n0 + j * 64 + 2 * lane), norm_index(a, n0 + j * 64 + 2 * lane + 1));
          step(yy.x, gg.x);
          step(yy.y, gg.y);
        }
      } else {
        for (int q = lane; q < cnt; q += 32)
          step(reinterpret_cast<const double*>(st)[q],
               IDX == 1 ? reinterpret_cast<const int32_t*>(st + NW_TILE * 8)[q] : (IDX == 2 ? norm_index(a, n0 + q) : 0));
      }
    } else {
      for (int q = lane; q < cnt; q += 32)
        step(a.y[n0 + q], IDX == 1 ? a.gidx[n0 + q] : (IDX == 2 ? norm_index(a, n0 + q) : 0));
    }
    __syncwarp();
  }
#pragma unroll
  for (int k = 0; k < NF_CT; k++) {
    double v0 = s0[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v0 += __shfl_xor_sync(0xffffffffu, v0, o);
    if (lane == 0) red[k * NW_WARPS + warp] = v0;
  }
  __syncthreads();
  if (threadIdx.x < NF_CT) {
    int k = threadIdx.x, c = c0 + k;
    if (c < T.C) {
      double sm = 0.0;
      for (int w = 0; w < NW_WARPS; w++) sm += red[k * NW_WARPS + w];
      partials[(size_t)blockIdx.x * T.C + c] = sm;
    }
  }
}
// launch geometry of the TMA form: at most 444 CTAs (3 per SM), every warp a 4-node multiple of at least one stage
static inline int norm_tma_nodes_per_warp(int count) {
  long long warps = std::min<long long>(444LL * NW_WARPS, ((long long)count + NW_TILE - 1) / NW_TILE);
  if (warps < 1) warps = 1;
  long long npw = ((long long)count + warps - 1) / warps;
  return (int)((npw + 3) / 4 * 4);
}
static inline int norm_tma_blocks(int count) {
  long long per_block = (long long)norm_tma_nodes_per_warp(count) * NW_WARPS;
  return (int)std::max<long long>(1, ((long long)count + per_block - 1) / per_block);
}

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
  const bool live = c < T.C;
  // Launched with programmatic stream serialization right behind the streaming kernel: everything that does not read the
  // partials (the chain-level sd / tau through the interpreter, its logarithm) runs while that kernel drains; the grid
  // dependency wait below returns once it has completed and its writes are visible.
  const int count = hi - lo;
  double par = 0.0;
  if (live && lane == 0) {
    if (MODE == 1) par = get_param(T, a.decl, 0, c, NIMB_PARAM_TAU, -1, 0.0);
    else { const nimb200_decl& d = T.decl[a.decl]; par = eval_prog(T, c, d.prog_off[1], d.prog_len[1], 0, 0, -1, 0.0); }
  }
  cudaGridDependencySynchronize();
  if (!live) return;
  double S = 0.0;
  for (int b = lane; b < nb; b += 32) S += partials[(size_t)b * T.C + c];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
  if (lane != 0) return;
  if (MODE == 1) {
    dst0[c] = S * par;
    dst1[c] = (double)count * par;
    return;
  }
  const double sigma = par;
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

// Conjugate contributions of the whole declaration to a target inside the chain-level term A (the LMM's mu), from the per-group
// data-only sums gs[g] = (sum y, count) the sweep over v keeps:  sum_i (y_i - A(target := 0) - v[G(i)]) = sum_g (Sy_g - n_g (A0 + v_g)).
// One block per chain, threads stride over the groups (four independent partial sums each, so the gathers overlap), fixed-order
// butterfly and warp order: G terms per chain instead of a pass over the nodes.
__global__ void __launch_bounds__(1024)
k_norm_conj_from_groups(DevTab T, NormArgs a, const double* __restrict__ gs, int G, int ov_slot,
                        double* __restrict__ dst0, double* __restrict__ dst1) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  __shared__ double sR[32], sN[32];
  const int c = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, NT = blockDim.x;
  const double A0 = eval_prog(T, c, a.a_off, a.a_len, 0, 0, ov_slot, 0.0);
  const double* __restrict__ vb = T.state + (size_t)a.v_slot * T.C + c;
  double R[4] = {0.0, 0.0, 0.0, 0.0}, cnt = 0.0;
  for (int g0 = threadIdx.x; g0 < G; g0 += 4 * NT) {
    double sy[4], n[4], v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {  // all loads of a trip first (clamped index, no branch in between), then the arithmetic
      const int g = min(g0 + u * NT, G - 1);
      const double2 q = reinterpret_cast<const double2*>(gs)[g];
      sy[u] = q.x; n[u] = g0 + u * NT < G ? q.y : 0.0;
      v[u] = vb[(size_t)g * T.C];
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      R[u] += n[u] > 0.0 ? sy[u] - n[u] * (A0 + v[u]) : 0.0;   // an empty group contributes nothing, whatever its v holds
      cnt += n[u];
    }
  }
  double r = (R[0] + R[1]) + (R[2] + R[3]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { r += __shfl_xor_sync(0xffffffffu, r, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o); }
  if (lane == 0) { sR[warp] = r; sN[warp] = cnt; }
  __syncthreads();
  if (threadIdx.x != 0) return;
  double rr = 0.0, nn = 0.0;
  for (int w = 0; w < (NT >> 5); w++) { rr += sR[w]; nn += sN[w]; }
  const double tau = get_param(T, a.decl, 0, c, NIMB_PARAM_TAU, -1, 0.0);
  dst0[c] = rr * tau;
  dst1[c] = nn * tau;
}

// Conjugate normal-normal sweep over the gathered vector's elements (targets v[t]).  (sum y, count) over a target's nodes
// depend on the data only: k_norm_group_sums leaves this rank's share in gs[t][2] once (one block per target, fixed order;
// data-sharded runs all-reduce the 2 doubles per target), k_norm_conj_draw is one thread per (target, chain) forming
// contribution_mean = tau (sum y - n A), contribution_tau = n tau, drawing and writing.
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

// Second data-only statistic per target: the centred sum of squares  ss[t] = sum_i (y_i - ybar_t)^2  over the target's nodes,
// with ybar_t from the (already complete, i.e. all-reduced when data-sharded) group sums.  Same node walk as above.
__global__ void __launch_bounds__(128)
k_norm_group_ss(DevTab T, int sid, NormArgs a, const double* __restrict__ gs, double* __restrict__ ss) {
  __shared__ double red[4];
  const nimb200_sampler sp = T.samplers[sid];
  const int t = blockIdx.x;
  const nimb200_dep dp = T.deps[sp.dep_off + 1];
  Sel s = sel_of(T, dp, t);
  const double n = gs[2 * (size_t)t + 1];
  const double ybar = n > 0.0 ? gs[2 * (size_t)t] / n : 0.0;
  double acc = 0.0;
  for (int k = s.lo + threadIdx.x; k < s.hi; k += 128) {
    int node = s.idx ? s.idx[k] : k;
    if (s.idx && (node < s.nlo || node >= s.nhi)) continue;
    const double d = a.y[node] - ybar;
    acc += d * d;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) ss[t] = ((red[0] + red[1]) + red[2]) + red[3];
}

// Block sum of the whole declaration from the per-group data-only statistics (sum y, n, centred sum of squares):
//   sum_i (y_i - A - v[G(i)])^2 = sum_g [ ss_g + n_g (ybar_g - A - v_g)^2 ]
// -- every term non-negative, so nothing cancels -- then the same finish as the streaming kernel.  G terms per chain instead of a
// pass over the N nodes: with the group statistics in place an MCMC iteration of the linear mixed model reads no data node at all.
// One block per chain, fixed-order butterfly and warp order.
__global__ void __launch_bounds__(1024)
k_norm_sum_from_groups(DevTab T, NormArgs a, const double* __restrict__ gs, const double* __restrict__ ss, int G, int n_nodes,
                       double* __restrict__ dst) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  __shared__ double sS[32];
  const int c = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5, NT = blockDim.x;
  const double A = eval_prog(T, c, a.a_off, a.a_len, 0, 0, -1, 0.0);
  const double* __restrict__ vb = T.state + (size_t)a.v_slot * T.C + c;
  double S[4] = {0.0, 0.0, 0.0, 0.0};
  for (int g0 = threadIdx.x; g0 < G; g0 += 4 * NT) {
    double sy[4], n[4], v[4], q2[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int g = min(g0 + u * NT, G - 1);
      const double2 q = reinterpret_cast<const double2*>(gs)[g];
      sy[u] = q.x; n[u] = g0 + u * NT < G ? q.y : 0.0;
      q2[u] = ss[g];
      v[u] = vb[(size_t)g * T.C];
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const double e = sy[u] / fmax(n[u], 1.0) - (A + v[u]);
      S[u] += n[u] > 0.0 ? q2[u] + n[u] * (e * e) : 0.0;
    }
  }
  double r = (S[0] + S[1]) + (S[2] + S[3]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  if (lane == 0) sS[warp] = r;
  __syncthreads();
  if (threadIdx.x != 0) return;
  double tot = 0.0;
  for (int w = 0; w < (NT >> 5); w++) tot += sS[w];
  const nimb200_decl& d = T.decl[a.decl];
  const double sigma = eval_prog(T, c, d.prog_off[1], d.prog_len[1], 0, 0, -1, 0.0);
  // sd == 0: the result depends on how many residuals are exactly zero, which the group statistics cannot tell -- walk the nodes
  // (data-sharded ranks only hold their slice, and report NaN like the all-reduced streaming path does)
  if (sigma == 0 && n_nodes > 0 && !isnan(tot)) dst[c] = T.nranks > 1 ? NAN : norm_sd0_sum(T, a, c, 0, n_nodes, nullptr);
  else dst[c] = norm_sum_from_ss(sigma, tot, n_nodes);
}

__global__ void k_norm_conj_draw(DevTab T, int sid, uint64_t seed, uint64_t iter, NormArgs a, const double* __restrict__ gs) {
  cudaTriggerProgrammaticLaunchCompletion();
  cudaGridDependencySynchronize();
  const nimb200_sampler sp = T.samplers[sid];
  size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c;
  size_t tq;
  split_index(gid, T.C, tq, c);
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
