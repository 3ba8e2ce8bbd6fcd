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

// grid (node blocks, chain groups of NF_CT).  MODE 0: out[blk][0][c] = sum d^2, out[blk][1][c] = #(d == 0)
//                                              MODE 1: out[blk][0][c] = sum d (A evaluated with ov_slot := 0), out[blk][1][c] = 0
template <int MODE>
__global__ void __launch_bounds__(NF_THREADS)
k_norm_stream(DevTab T, NormArgs a, int lo, int hi, int nodes_per_block, const int32_t* idx, int ov_slot, double* __restrict__ partials) {
  __shared__ double red[2][NF_CT][NF_THREADS / 32];
  const int c0 = blockIdx.y * NF_CT;
  double A[NF_CT], s0[NF_CT], s1[NF_CT];
#pragma unroll
  for (int k = 0; k < NF_CT; k++) {
    int c = min(c0 + k, T.C - 1);
    A[k] = eval_prog(T, c, a.a_off, a.a_len, 0, 0, MODE == 1 ? ov_slot : -1, 0.0);
    s0[k] = 0.0; s1[k] = 0.0;
  }
  const int b_lo = lo + blockIdx.x * nodes_per_block, b_hi = min(hi, b_lo + nodes_per_block);
  for (int k0 = b_lo + threadIdx.x; k0 < b_hi; k0 += NF_THREADS) {
    int node = idx ? idx[k0] : k0;
    if (idx && (node < T.shard_lo[a.decl] || node >= T.shard_hi[a.decl])) continue;
    double yv = a.y[node];
    const double* vp = a.mode >= 0 ? T.state + (size_t)(a.v_slot + norm_index(a, node)) * T.C : nullptr;
#pragma unroll
    for (int k = 0; k < NF_CT; k++) {
      int c = min(c0 + k, T.C - 1);
      double m = vp ? A[k] + vp[c] : A[k];
      double d = yv - m;
      if (MODE == 0) { s0[k] = fma(d, d, s0[k]); s1[k] += d == 0.0 ? 1.0 : 0.0; }
      else s0[k] += d;
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NF_CT; k++) {
    double v0 = s0[k], v1 = s1[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { v0 += __shfl_xor_sync(0xffffffffu, v0, o); v1 += __shfl_xor_sync(0xffffffffu, v1, o); }
    if (lane == 0) { red[0][k][warp] = v0; red[1][k][warp] = v1; }
  }
  __syncthreads();
  if (threadIdx.x < 2 * NF_CT) {
    int which = threadIdx.x / NF_CT, k = threadIdx.x % NF_CT, c = c0 + k;
    if (c < T.C) {
      double s = 0.0;
      for (int w = 0; w < NF_THREADS / 32; w++) s += red[which][k][w];
      partials[((size_t)blockIdx.x * 2 + which) * T.C + c] = s;
    }
  }
}

// per chain: (sum d^2, #zeros) -> sum of log-densities, with nmath's dnorm special cases
__global__ void k_norm_finish(DevTab T, NormArgs a, int count, const double* __restrict__ SZ /*[2][C]*/, double* __restrict__ dst) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= T.C) return;
  const nimb200_decl& d = T.decl[a.decl];
  double sigma = eval_prog(T, c, d.prog_off[1], d.prog_len[1], 0, 0, -1, 0.0);
  double S = SZ[c], Z = SZ[(size_t)T.C + c], n = (double)count, r;
  if (count == 0) r = 0.0;
  else if (isnan(sigma) || isnan(S)) r = sigma + S;
  else if (sigma < 0) r = NAN;
  else if (!isfinite(sigma)) r = -INFINITY;
  else if (sigma == 0) r = (Z == n) ? INFINITY : (Z == 0.0 ? -INFINITY : NAN);
  else {
    double q = S / (sigma * sigma);
    r = !isfinite(q) ? -INFINITY : -(n * (NIMB_LN_SQRT_2PI + log(sigma)) + 0.5 * q);
  }
  dst[c] = r;
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

}  // namespace nimb
