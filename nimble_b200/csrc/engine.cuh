// engine.cuh -- shared declarations of the sm_100a library (host model object + device table view).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/nimb200.h"

namespace nimb {

// cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE attribute of a kernel: set it once for every device a
// model is created on (`done` is one call site's bit mask of devices already configured; concurrent callers at worst
// set the attribute twice).
template <typename K>
static inline void ensure_dyn_smem(K kernel, size_t bytes, std::atomic<unsigned long long>& done) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (done.load(std::memory_order_acquire) & bit) return;
  cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  done.fetch_or(bit, std::memory_order_release);
}

// Device view of the node table + chain-major state.  Passed to kernels by value.
struct DevTab {
  const nimb200_op* ops;
  const nimb200_decl* decl;
  const nimb200_dep* deps;
  const nimb200_sampler* samplers;
  const double* const* darr;
  const int32_t* const* iarr;
  double* state;  // [n_slots][C]: chain index fastest, so a warp reads 32 chains' copies of one node coalesced
  int n_slots, C, first_chain;
  // data-node sharding: this rank owns instances [shard_lo[d], shard_hi[d]) of declaration d (whole range if not sharded)
  const int32_t* shard_lo;
  const int32_t* shard_hi;
  const int32_t* sharded;
  int rank, nranks;
  // pconst[3*d + p] != 0: parameter program p of declaration d does not depend on the node index (hoisted out of node loops)
  const int32_t* pconst;
  // ctr[0]: global iteration index of the iteration being executed (device resident so that one captured CUDA graph can be
  // replayed for every iteration); kernels read it instead of their `iter` argument
  const unsigned long long* ctr;
};

// Per-sampler adaptive state, all [T][C] (or [d*d][C], [AI*d][C]) chain-fastest.
struct SamplerDev {
  double* scale;
  int32_t* timesRan;
  int32_t* timesAccepted;
  int32_t* timesAdapted;
  double* propCov;
  double* chol;
  double* cholScale;
  double* empirSamp;
  // MCMCsaveHistory (P/R/MCMC_samplers.R:394-399, 726-739): hist[adaptation][k][T*C], k = 0 scale before the update, 1 acceptance
  // rate, 2.. the d*d entries of propCov before the update (RW_block with d <= 10); null when histories are off
  double* hist;
  int hist_cap, hist_K;
};

// fused y[i] ~ dbern(ilogit(X[i,] . beta)) block sum (fused_logit.cu).  partials: [n_tiles][C].
// Returns the number of node tiles written (0 if the shape is unsupported and the generic kernel must run).
int fused_bern_logit_tiles(int n);
size_t fused_bern_logit_xpad_len(int n, int p);
void fused_bern_logit_prepare_x(const double* X_rowmajor_dev, const double* y_dev, int n, int p, int ld, double* Xpad_dev,
                                cudaStream_t st);
// tiles [tile_begin, tile_end) of 128 nodes (data sharding gives each rank whole tiles)
void fused_bern_logit_launch(const double* Xpad, int n, int p, const double* state, int beta_slot, int C, double* partials,
                             int tile_begin, int tile_end, cudaStream_t st);

}  // namespace nimb
