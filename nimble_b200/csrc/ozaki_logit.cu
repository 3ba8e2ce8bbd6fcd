// ozaki_logit.cu -- the fused dbern-logit-inprod block sum with its contraction on the 5th-generation tensor cores.
//
// Why: in k_fused_bern_logit (fused_logit.cu) the FP64 multipliers are the saturated unit -- the FP64 mma.sync contraction
// takes 52 of the ~77 multiplier slots per evaluation, the link + density epilogue the rest (ncu: DMMA sub-pipe 59 % + FP64
// pipe 26 %).  tcgen05.mma has no f64 kind, but X is DATA: it is split once, at model-create time, into eight signed 7-bit
// slices per element (x = 2^e_row * sum_s d_s 2^(-7(s+1)), |d_s| <= 64, exact to 2^-57 of the row maximum); beta is split the
// same way per chain at every launch; the 36 slice products with s + t <= 7 are exact int8 x int8 -> int32 tensor-core
// products (tcgen05.mma kind::i8, accumulators in TMEM, one 32-column accumulator per weight s + t), recombined exactly in
// 64-bit integer arithmetic and scaled by the two exponents: the linear predictor carries an absolute error below
// 2^-50 * 2^(e_row + f_chain), i.e. ~1e-14 for the benchmark's magnitudes, far inside the 1e-12 block-sum bar.  The FP64 pipe
// is left with two conversions, one fma, one multiply and the 24.8-instruction epilogue per evaluation.
//
// Structure (one CTA per SM, 576 threads): warp 0 = producer (one 1-D TMA bulk copy per 128-row tile: 64 KiB of slices laid
// out in global memory exactly as the UMMA canonical K-major no-swizzle layout wants them in shared memory, plus y and the row
// exponents), warp 1 = MMA issuer, warps 2-17 = epilogue (tcgen05.ld 32x32b.x8, row = TMEM lane = thread).  A CTA owns 64
// chains as two halves of 32: while the epilogue works on one half's accumulators the tensor cores fill the other's -- the two
// halves are the two TMEM stages (2 x 8 x 32 = 512 columns).
// MMA shape (round 2, second version): the beta slices of a half are stacked along N into ONE 256-row operand, so A slice s
// against slices t = 0 .. 7 - s is one tcgen05.mma of N = (8 - s) 32 per K step whose accumulator columns are the weights
// s .. 7 -- 16 MMAs per half tile instead of 72, 136 KiB of shared-memory operand reads instead of 360 (tensor + TMEM-load
// pipeline alone: 2.97 -> 1.47 ms).  Epilogue: tables held 16 times (bank-conflict free), scale as two multiplies by
// precomputed powers of two, 77 instructions per evaluation instead of 88.
// Measured (profiles/r2q_ozaki_summary.md): 4.6 - 4.7 ms per launch against 4.85 ms for the FP64 tensor-core kernel, epilogue
// alone 3.3 - 3.5 ms.  The clock64 timeline (nimb200_debug_ozaki_trace) shows why the two do not overlap: per half and warp
// 1000 cycles of waiting for the tensor cores and 570 of TMEM loads against 3300 of arithmetic, the waits of the four warps
// of a scheduler coincide, and the arithmetic itself runs 16 % slower while the MMAs stream their operands from shared memory.
// Still a diagnostic (nimb200_debug_ozaki), not the sampling path: a 5 % gain does not pay for 524 bytes per row of sliced X.
#include <cfloat>
#include <cmath>
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include "engine.cuh"
#include "fastmath.cuh"
#include "bern_logit.cuh"
#include <string>
#include <vector>

namespace nimb { void set_error(const std::string& s); }

namespace nimb {

static const int OZ_M = 128;                 // rows (nodes) per tile
static const int OZ_K = 64;                  // inner dimension padded to two K = 32 steps
static const int OZ_S = 8;                   // slices per operand
static const int OZ_NC = 64;                 // chains per CTA
static const int OZ_NH = 32;                 // chains per half (UMMA N)
static const int OZ_SLICE_A = OZ_M * OZ_K;   // bytes of one A slice of a tile (8192)
static const int OZ_SLICE_B = OZ_NC * OZ_K;  // bytes of one B slice (4096)
static const int OZ_TILE_BYTES = OZ_S * OZ_SLICE_A + OZ_M * 8 + OZ_M * 4;  // slices + y[128] + e[128] = 67072
static const int OZ_EPI_WARPS = 16;                        // four per TMEM lane quarter, each a quarter (8) of a half's columns
static const int OZ_THREADS = 64 + 32 * OZ_EPI_WARPS;      // producer warp + MMA warp + epilogue warps

__device__ __forceinline__ uint32_t oz_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void oz_mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(oz_smem(bar)), "r"(count));
}
__device__ __forceinline__ void oz_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(oz_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void oz_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(oz_smem(bar)) : "memory");
}
// bounded wait: a barrier that never completes (a protocol error) traps instead of hanging the GPU
template <unsigned SLEEP = 0>
__device__ __forceinline__ void oz_mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = oz_smem(bar);
  for (unsigned spins = 0;; spins++) {
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(a), "r"(parity)
        : "memory");
    if (done) return;
    if (SLEEP) __nanosleep(SLEEP);   // producer / MMA warps: a spinning warp must not take issue slots from the epilogue warps of its scheduler
    if (spins > (1u << 26)) __trap();
  }
}
__device__ __forceinline__ void oz_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(oz_smem(dst)),
               "l"(src), "r"(bytes), "r"(oz_smem(bar))
               : "memory");
}
// shared-memory matrix descriptor, K-major, no swizzle: core matrix = 8 rows x 16 bytes, contiguous (128 B);
// lbo = byte distance between the two core matrices of one K = 32 step, sbo = byte distance between 8-row groups
__device__ __forceinline__ uint64_t oz_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         (1ull << 46);
}
// instruction descriptor: D = S32, A = B = signed int8, both K-major, M = 128, N given (a multiple of 32 up to 256)
__host__ __device__ constexpr uint32_t oz_idesc(int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(OZ_M >> 4) << 24);
}

// The whole MMA warp runs the issue loop (warp-uniform control flow keeps the descriptors in uniform registers, which is what
// the instruction reads); only the elected lane's instruction takes effect.
template <int ACC>
__device__ __forceinline__ void oz_mma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t leader) {
  asm volatile(
      "{\n"
      ".reg .pred p, q;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "setp.ne.b32 q, %6, 0;\n"
      "@q tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n"
      "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "n"(ACC), "r"(0u), "r"(leader)
      : "memory");
}
__device__ __forceinline__ void oz_commit(uint64_t* bar, uint32_t leader) {
  asm volatile(
      "{\n"
      ".reg .pred q;\n"
      "setp.ne.b32 q, %1, 0;\n"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
      "}\n" ::"r"(oz_smem(bar)), "r"(leader)
      : "memory");
}
__device__ __forceinline__ uint32_t oz_elect() {
  uint32_t pred;
  asm volatile(
      "{\n"
      ".reg .pred P;\n"
      "elect.sync _|P, 0xffffffff;\n"
      "selp.u32 %0, 1, 0, P;\n"
      "}\n"
      : "=r"(pred));
  return pred;
}
__device__ __forceinline__ void oz_ld4(uint32_t taddr, int (&v)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3])
               : "r"(taddr));
}

__device__ __forceinline__ void oz_ld8(uint32_t taddr, int (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr));
}

// ---------------------------------------------------------------------------------------------------------------- slicing
// digit s of r (|r| < 1/2): d = rint(128 r), remainder 128 r - d (exact)
__device__ __forceinline__ int oz_digit(double& r) {
  const double t = r * 128.0;
  const double d = rint(t);
  r = t - d;
  return (int)d;
}

// X [n][ld] -> per 128-row tile: 8 slices in the UMMA canonical layout [slice][k16 chunk 0..3][row group 0..15][row 0..7][16 B],
// then y[128] (-1 marks rows past n), then the row exponents e[128].  One thread per row.  flag[0] is raised when a row holds a
// non-finite value or an exponent outside the supported range (the caller then keeps the FP64 tensor-core kernel).
__global__ void k_oz_slice_x(const double* __restrict__ X, const double* __restrict__ y, int n, int p, int ld,
                             unsigned char* __restrict__ out, int* __restrict__ flag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int ntile_rows = ((n + OZ_M - 1) / OZ_M) * OZ_M;
  if (i >= ntile_rows) return;
  unsigned char* tile = out + (size_t)(i / OZ_M) * OZ_TILE_BYTES;
  const int r = i % OZ_M, rg = r >> 3, rr = r & 7;
  double mx = 0.0;
  bool bad = false;
  if (i < n)
    for (int k = 0; k < p; k++) {
      const double v = X[(size_t)i * ld + k];
      bad |= !isfinite(v);
      mx = fmax(mx, fabs(v));
    }
  int e = 0;
  if (mx > 0.0 && !bad) e = ilogb(mx) + 2;   // |x| 2^-e < 1/2
  if (e < -400 || e > 400) { bad = true; e = 0; }
  if (bad) atomicExch(flag, 1);
  for (int k = 0; k < OZ_K; k++) {
    double rem = (i < n && k < p && !bad) ? scalbn(X[(size_t)i * ld + k], -e) : 0.0;
    for (int s = 0; s < OZ_S; s++) {
      const int d = oz_digit(rem);
      tile[(size_t)s * OZ_SLICE_A + (k >> 4) * 2048 + rg * 128 + rr * 16 + (k & 15)] = (unsigned char)(signed char)d;
    }
  }
  double* yt = (double*)(tile + OZ_S * OZ_SLICE_A);
  int* et = (int*)(tile + OZ_S * OZ_SLICE_A + OZ_M * 8);
  yt[r] = i < n ? y[i] : -1.0;
  et[r] = e;
}

// timeline diagnostic (nimb200_debug_ozaki_trace): clock64 stamps of CTA (0, 0), [warp 0..17][tile 0..63][half][4 events]
__device__ long long* g_oz_trace = nullptr;
#define OZ_STAMP(it, h, k)                                                                                           \
  do {                                                                                                               \
    if (trace && lane == 0 && (it) < 64) trace[(((size_t)warp * 64 + (it)) * 2 + (h)) * 4 + (k)] = clock64();        \
  } while (0)

// ---------------------------------------------------------------------------------------------------------------- main kernel
// grid (chain tiles of 64, node chunks).  partials[chunk][C].  eta_out (diagnostic, may be null): [n][C] linear predictors.
__global__ void __launch_bounds__(OZ_THREADS, 1)
k_oz_bern_logit(const unsigned char* __restrict__ Xq, int n, int p, const double* __restrict__ state, int beta_slot, int C,
                int tiles_per_chunk, int tile_begin, int tile_end, double* __restrict__ partials, double* __restrict__ eta_out,
                int dbg) {
  extern __shared__ __align__(1024) unsigned char smem[];
  // carve-up: A stages | B slices | barriers, TMEM address, chain exponents | reduction scratch | math tables
  unsigned char* sA = smem;                                         // 2 x OZ_TILE_BYTES (67072 each, 128-byte multiple)
  unsigned char* sB = sA + 2 * OZ_TILE_BYTES;                       // 8 x 4096
  uint64_t* bars = (uint64_t*)(sB + OZ_S * OZ_SLICE_B);             // a_full[2] a_empty[2] t_full[2] t_empty[2]
  uint32_t* s_tmem = (uint32_t*)(bars + 8);
  int* s_f = (int*)(s_tmem + 2);                                    // chain exponents [64]
  double* s_red = (double*)(s_f + OZ_NC);                           // [4 quarters][64 chains]
  double* s_cs = s_red + 4 * OZ_NC;                                 // chain scales 2^f [64] (NaN: poisoned chain)
  MathTab* tab = (MathTab*)(s_cs + OZ_NC);
  double* s_rep = (double*)(tab + 1);                               // the replicated tables (BL_REP_BYTES)
  uint64_t *a_full = bars, *a_empty = bars + 2, *t_full = bars + 4, *t_empty = bars + 6;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  long long* trace = (blockIdx.x == 0 && blockIdx.y == 0) ? g_oz_trace : nullptr;
  const int chain0 = blockIdx.x * OZ_NC, chunk = blockIdx.y;
  const int tile_lo = tile_begin + chunk * tiles_per_chunk, tile_hi = min(tile_end, tile_lo + tiles_per_chunk);
  const int ntiles = max(0, tile_hi - tile_lo);

  for (int i = tid; i < (int)(sizeof(MathTab) / sizeof(double)); i += OZ_THREADS) ((double*)tab)[i] = ((const double*)&g_mathtab)[i];
  load_bern_rep_tables(s_rep, tid, OZ_THREADS);
  // beta of this CTA's 64 chains -> int8 slices in the canonical layout [half][k16 chunk][slice][chain group 0..3][chain 0..7][16 B]:
  // per half, the eight slices of its 32 chains stacked along N make ONE 256-row K-major operand
  if (tid < OZ_NC) {
    const int c = chain0 + tid;
    double mx = 0.0;
    bool bad = false;
    if (c < C)
      for (int k = 0; k < p; k++) {
        const double v = state[(size_t)(beta_slot + k) * C + c];
        bad |= !isfinite(v);
        mx = fmax(mx, fabs(v));
      }
    int f = (mx > 0.0 && !bad) ? ilogb(mx) + 2 : 0;
    if (f < -400 || f > 400) { bad = true; f = 0; }
    s_f[tid] = bad ? (1 << 20) : f;     // a non-finite or out-of-range beta poisons the chain: its sum is recomputed as NaN below
    s_cs[tid] = bad ? __longlong_as_double(0x7ff8000000000000ll) : __hiloint2double((1023 + f) << 20, 0);
    for (int k = 0; k < OZ_K; k++) {
      double rem = (c < C && k < p && !bad) ? scalbn(state[(size_t)(beta_slot + k) * C + c], -f) : 0.0;
      for (int s = 0; s < OZ_S; s++) {
        const int d = oz_digit(rem);
        // half h = tid / 32 holds one 256-row operand: row = slice * 32 + chain of the half (see the MMA loop)
        sB[(size_t)(tid >> 5) * (OZ_S * OZ_NH * OZ_K) + (k >> 4) * (OZ_S * OZ_NH * 16) + (s * 4 + ((tid & 31) >> 3)) * 128 + (tid & 7) * 16 +
           (k & 15)] = (unsigned char)(signed char)d;
      }
    }
  }
  if (tid == 0) {
    oz_mbar_init(&a_full[0], 1); oz_mbar_init(&a_full[1], 1);
    oz_mbar_init(&a_empty[0], 1 + OZ_EPI_WARPS); oz_mbar_init(&a_empty[1], 1 + OZ_EPI_WARPS);
    oz_mbar_init(&t_full[0], 1); oz_mbar_init(&t_full[1], 1);
    oz_mbar_init(&t_empty[0], OZ_EPI_WARPS); oz_mbar_init(&t_empty[1], OZ_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // the MMA warp owns the TMEM allocation: all 512 columns (one CTA per SM)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(oz_smem(s_tmem)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // the generic-proxy writes of the beta slices must be visible to the tensor core's (async proxy) reads
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;

  if (warp == 0) {
    // ===================================================== producer: one bulk copy per tile into a 2-stage ring
    if (lane == 0) {
      for (int it = 0; it < ntiles; it++) {
        const int st = it & 1;
        if (it >= 2) oz_mbar_wait<200>(&a_empty[st], (uint32_t)(((it >> 1) - 1) & 1));
        oz_mbar_expect_tx(&a_full[st], (uint32_t)OZ_TILE_BYTES);
        oz_bulk_g2s(sA + (size_t)st * OZ_TILE_BYTES, Xq + (size_t)(tile_lo + it) * OZ_TILE_BYTES, (uint32_t)OZ_TILE_BYTES, &a_full[st]);
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer: 36 slice pairs x 2 K-steps per half tile
    {
      const uint32_t aB = oz_smem(sB);
      const uint32_t leader = oz_elect();
      for (int it = 0; it < ntiles; it++) {
        const int st = it & 1;
        oz_mbar_wait(&a_full[st], (uint32_t)((it >> 1) & 1));
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t aA = oz_smem(sA + (size_t)st * OZ_TILE_BYTES);
        // descriptors differ from the base ones only in the 14-bit start-address field (16-byte units): one 64-bit add per
        // operand and MMA, everything else is a compile-time constant of the fully unrolled 72-instruction sequence (a single
        // thread issues them all, so every instruction of this loop is on the critical path of the tensor pipe's occupancy)
        const uint64_t da0 = oz_desc(aA, 2048, 128);
#pragma unroll
        for (int h = 0; h < 2; h++) {
          OZ_STAMP(it, h, 0);
          if (it >= 1) {  // the epilogue has drained this half's accumulators of the previous tile
            oz_mbar_wait(&t_empty[h], (uint32_t)((it - 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          }
          OZ_STAMP(it, h, 1);
          if ((dbg & 3) == 2) { oz_commit(&t_full[h], leader); continue; }   // diagnostic: no tensor work
          // A slice s against the stacked beta slices t = 0 .. 7 - s (rows 0 .. (8 - s) 32 of the half's operand) lands in the
          // accumulator columns of the weights s .. 7: one MMA of N = (8 - s) 32 per K step instead of 8 - s MMAs of N = 32 --
          // the A tile (4 KiB per K step) is read once per slice, not once per slice pair (round 2: 136 KiB of shared-memory
          // operand reads per half tile instead of 360; the tensor work is the same 1152 cycles)
          const uint64_t db0 = oz_desc(aB + h * (OZ_S * OZ_NH * OZ_K), OZ_S * OZ_NH * 16, 128);
          const uint32_t d0 = tmem + (uint32_t)(h * 256);
#pragma unroll
          for (int s = 0; s < OZ_S; s++) {
#pragma unroll
            for (int ks = 0; ks < 2; ks++) {
              const uint64_t da = da0 + (uint64_t)((s * OZ_SLICE_A + ks * 4096) >> 4);
              const uint64_t db = db0 + (uint64_t)((ks * 2 * OZ_S * OZ_NH * 16) >> 4);
              if ((s | ks) != 0) oz_mma<1>(d0 + (uint32_t)(s * OZ_NH), da, db, oz_idesc((OZ_S - s) * OZ_NH), leader);
              else oz_mma<0>(d0, da, db, oz_idesc(OZ_S * OZ_NH), leader);
            }
          }
          oz_commit(&t_full[h], leader);         // accumulators of (tile, half) complete -> epilogue
          OZ_STAMP(it, h, 2);
        }
        oz_commit(&a_empty[st], leader);         // every MMA reading this A stage has completed
      }
    }
  } else {
    // ===================================================== epilogue: row = TMEM lane = thread, 8 columns per half
    const int ew = warp - 2;
    const int q = warp & 3;                      // TMEM lane quarter this warp may touch
    const int cq = ew >> 2;                      // which 8 of a half's 32 columns
    const int row = q * 32 + lane;
    const double* __restrict__ expL = s_rep + (lane & 15);
    double csum[2][8];
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
      for (int j = 0; j < 8; j++) csum[h][j] = 0.0;
    for (int it = 0; it < ntiles; it++) {
      const int st = it & 1;
      oz_mbar_wait(&a_full[st], (uint32_t)((it >> 1) & 1));   // y and the row exponents travel with the tile
      const unsigned char* tile = sA + (size_t)st * OZ_TILE_BYTES;
      const double yv = ((const double*)(tile + OZ_S * OZ_SLICE_A))[row];
      const int er = ((const int*)(tile + OZ_S * OZ_SLICE_A + OZ_M * 8))[row];
      __syncwarp();
      if (lane == 0) oz_mbar_arrive(&a_empty[st]);            // this warp no longer needs the stage
      const bool live = __double2hiint(yv) >= 0, y1 = (__double2hiint(yv) | __double2loint(yv)) != 0;   // y is 0, 1 or -1 (past n)
      const double rs = __hiloint2double((1023 + er - 35) << 20, 0);   // 2^(er - 35): the row's share of the scale
      const long long grow = (long long)(tile_lo + it) * OZ_M + row;
#pragma unroll
      for (int h = 0; h < 2; h++) {
        if (dbg & 4) asm volatile("bar.sync 1, %0;" ::"n"(32 * OZ_EPI_WARPS) : "memory");   // experiment: epilogue warps in lockstep
        OZ_STAMP(it, h, 0);
        oz_mbar_wait(&t_full[h], (uint32_t)(it & 1));
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        OZ_STAMP(it, h, 1);
        // the thread's 8 columns of the half, all eight weights, two weights (16 accumulator registers) per wait; exact
        // recombination in 64-bit integers, then TMEM goes back to the tensor cores before any floating-point arithmetic
        const uint32_t base = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(h * 256 + cq * 8);
        long long hi[8], lo[8];
        {
          int a[2][8], b[8];
          oz_ld8(base, a[0]); oz_ld8(base + (uint32_t)OZ_NH, a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int j = 0; j < 8; j++) b[j] = a[0][j] * 128 + a[1][j];
          oz_ld8(base + (uint32_t)(2 * OZ_NH), a[0]); oz_ld8(base + (uint32_t)(3 * OZ_NH), a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int j = 0; j < 8; j++) hi[j] = (long long)b[j] * 16384 + (a[0][j] * 128 + a[1][j]);
          oz_ld8(base + (uint32_t)(4 * OZ_NH), a[0]); oz_ld8(base + (uint32_t)(5 * OZ_NH), a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int j = 0; j < 8; j++) b[j] = a[0][j] * 128 + a[1][j];
          oz_ld8(base + (uint32_t)(6 * OZ_NH), a[0]); oz_ld8(base + (uint32_t)(7 * OZ_NH), a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) oz_mbar_arrive(&t_empty[h]);
#pragma unroll
          for (int j = 0; j < 8; j++) lo[j] = (long long)b[j] * 16384 + (a[0][j] * 128 + a[1][j]);
        }
        OZ_STAMP(it, h, 2);
        if ((dbg & 3) == 1) { csum[h][0] += (double)(hi[0] + lo[7]); continue; }   // diagnostic: no epilogue arithmetic
#pragma unroll
        for (int cb = 0; cb < 2; cb++) {
          double eta[4], lp[4];
#pragma unroll
          for (int j = 0; j < 4; j++) {
            // exact int64 -> double on the integer pipe (|hi|, |lo| < 2^51): 2^52 + 2^51 + x has x in its low mantissa bits;
            // for lo the same trick 28 binary places down yields lo 2^-28 directly.  Two exact subtractions, one rounding add.
            const double dh = __longlong_as_double(hi[cb * 4 + j] + 0x4338000000000000ll) - 6755399441055744.0;
            const double dl = __longlong_as_double(lo[cb * 4 + j] + 0x4178000000000000ll) - 25165824.0;
            eta[j] = ((dh + dl) * rs) * s_cs[h * OZ_NH + cq * 8 + cb * 4 + j];
          }
          { const bool yv4[4] = {y1, y1, y1, y1}; bern_logit_lp_rep<4, true>(eta, yv4, lp, expL, tab); }
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (live) csum[h][cb * 4 + j] += lp[j];
          if (eta_out) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
              const int c = chain0 + h * OZ_NH + cq * 8 + cb * 4 + j;
              if (live && grow < n && c < C) eta_out[(size_t)grow * C + c] = eta[j];
            }
          }
        }
        OZ_STAMP(it, h, 3);
      }
    }
    // per-chain sums over the 128 rows: butterfly inside the warp, then the four lane quarters through shared memory
#pragma unroll
    for (int h = 0; h < 2; h++)
#pragma unroll
      for (int j = 0; j < 8; j++) {
        double v = csum[h][j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[q * OZ_NC + h * OZ_NH + cq * 8 + j] = v;
      }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid < OZ_NC) {
    const int c = chain0 + tid;
    if (c < C) partials[(size_t)chunk * C + c] = ((s_red[tid] + s_red[OZ_NC + tid]) + s_red[2 * OZ_NC + tid]) + s_red[3 * OZ_NC + tid];
  }
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------- two CTAs per SM
// Third structure (round 2): a CTA owns 32 chains and ONE accumulator stage (8 weights x 32 columns = 256 TMEM columns), one
// A stage, 8 epilogue warps (lane quarter x 16 columns) -- and two such CTAs share an SM.  Inside a CTA the chain
// MMA -> TMEM loads -> arithmetic is serial apart from the second half of the arithmetic (the stage is handed back as soon
// as the second batch of 8 columns is in registers); the other CTA's arithmetic fills the gap.  Why: with 16 epilogue warps on
// two stages of one CTA the warps drift a whole tile apart and the four warps of a scheduler end up waiting for the tensor
// cores together (timeline in profiles/r2q_ozaki_summary.md); two independent CTAs cannot wait together for long.
static const int OZ2_NC = 32;                              // chains per CTA = UMMA rows of the stacked operand / 8
static const int OZ2_EPI_WARPS = 8;
static const int OZ2_THREADS = 64 + 32 * OZ2_EPI_WARPS;    // 320
static const int OZ2_REP = 8;                              // copies of the look-up tables (two CTAs' shared memory must fit an SM)
static const int OZ2_B_BYTES = OZ_S * OZ2_NC * OZ_K;       // 16384

__global__ void __launch_bounds__(OZ2_THREADS, 2)
k_oz2_bern_logit(const unsigned char* __restrict__ Xq, int n, int p, const double* __restrict__ state, int beta_slot, int C,
                 int tiles_per_chunk, int tile_begin, int tile_end, double* __restrict__ partials, double* __restrict__ eta_out,
                 int dbg) {
  extern __shared__ __align__(1024) unsigned char smem[];
  unsigned char* sA = smem;                                         // one stage: OZ_TILE_BYTES
  unsigned char* sB = sA + OZ_TILE_BYTES;                           // OZ2_B_BYTES
  uint64_t* bars = (uint64_t*)(sB + OZ2_B_BYTES);                   // a_full a_empty t_full t_empty
  uint32_t* s_tmem = (uint32_t*)(bars + 4);
  double* s_red = (double*)(s_tmem + 2);                            // [4 quarters][32 chains]
  double* s_cs = s_red + 4 * OZ2_NC;                                // chain scales 2^f [32] (NaN: poisoned chain)
  MathTab* tab = (MathTab*)(s_cs + OZ2_NC);
  double* s_rep = (double*)(tab + 1);                               // tables, OZ2_REP copies
  uint64_t *a_full = bars, *a_empty = bars + 1, *t_full = bars + 2, *t_empty = bars + 3;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  long long* trace = (blockIdx.x == 0 && blockIdx.y == 0) ? g_oz_trace : nullptr;
  const int chain0 = blockIdx.x * OZ2_NC, chunk = blockIdx.y;
  const int tile_lo = tile_begin + chunk * tiles_per_chunk, tile_hi = min(tile_end, tile_lo + tiles_per_chunk);
  const int ntiles = max(0, tile_hi - tile_lo);

  for (int i = tid; i < (int)(sizeof(MathTab) / sizeof(double)); i += OZ2_THREADS) ((double*)tab)[i] = ((const double*)&g_mathtab)[i];
  for (int i = tid; i < BL_REP_ENTRIES * OZ2_REP; i += OZ2_THREADS) {
    const int e = i / OZ2_REP;
    s_rep[i] = e < 64 ? g_mathtab.expT[e] : e < 64 + 129 ? g_mathtab.invc[e - 64] : (e - 193 == 128 ? c_tab_misc[7] : g_mathtab.logA[e - 193]);
  }
  if (tid < OZ2_NC) {   // beta of this CTA's 32 chains -> [k16 chunk][slice][chain group 0..3][chain 0..7][16 B]: one 256-row operand
    const int c = chain0 + tid;
    double mx = 0.0;
    bool bad = false;
    if (c < C)
      for (int k = 0; k < p; k++) {
        const double v = state[(size_t)(beta_slot + k) * C + c];
        bad |= !isfinite(v);
        mx = fmax(mx, fabs(v));
      }
    int f = (mx > 0.0 && !bad) ? ilogb(mx) + 2 : 0;
    if (f < -400 || f > 400) { bad = true; f = 0; }
    s_cs[tid] = bad ? __longlong_as_double(0x7ff8000000000000ll) : __hiloint2double((1023 + f) << 20, 0);
    for (int k = 0; k < OZ_K; k++) {
      double rem = (c < C && k < p && !bad) ? scalbn(state[(size_t)(beta_slot + k) * C + c], -f) : 0.0;
      for (int sl = 0; sl < OZ_S; sl++) {
        const int d = oz_digit(rem);
        sB[(k >> 4) * (OZ_S * OZ2_NC * 16) + (sl * 4 + (tid >> 3)) * 128 + (tid & 7) * 16 + (k & 15)] = (unsigned char)(signed char)d;
      }
    }
  }
  if (tid == 0) {
    oz_mbar_init(a_full, 1);
    oz_mbar_init(a_empty, 1 + OZ2_EPI_WARPS);
    oz_mbar_init(t_full, 1);
    oz_mbar_init(t_empty, OZ2_EPI_WARPS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // half of the SM's TMEM: the co-resident CTA takes the other half
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(oz_smem(s_tmem)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *s_tmem;

  if (warp == 0) {
    // ===================================================== producer: one bulk copy per tile into the single stage
    if (lane == 0) {
      for (int it = 0; it < ntiles; it++) {
        if (it >= 1) oz_mbar_wait<100>(a_empty, (uint32_t)((it - 1) & 1));
        oz_mbar_expect_tx(a_full, (uint32_t)OZ_TILE_BYTES);
        oz_bulk_g2s(sA, Xq + (size_t)(tile_lo + it) * OZ_TILE_BYTES, (uint32_t)OZ_TILE_BYTES, a_full);
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer: 8 slices x 2 K steps per tile, N = (8 - s) 32
    const uint32_t aB = oz_smem(sB), aA = oz_smem(sA);
    const uint32_t leader = oz_elect();
    const uint64_t da0 = oz_desc(aA, 2048, 128);
    const uint64_t db0 = oz_desc(aB, OZ_S * OZ2_NC * 16, 128);
    for (int it = 0; it < ntiles; it++) {
      OZ_STAMP(it, 0, 0);
      oz_mbar_wait(a_full, (uint32_t)(it & 1));
      if (it >= 1) oz_mbar_wait(t_empty, (uint32_t)((it - 1) & 1));
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      OZ_STAMP(it, 0, 1);
      if ((dbg & 3) != 2) {
#pragma unroll
        for (int sl = 0; sl < OZ_S; sl++) {
#pragma unroll
          for (int ks = 0; ks < 2; ks++) {
            const uint64_t da = da0 + (uint64_t)((sl * OZ_SLICE_A + ks * 4096) >> 4);
            const uint64_t db = db0 + (uint64_t)((ks * 2 * OZ_S * OZ2_NC * 16) >> 4);
            if ((sl | ks) != 0) oz_mma<1>(tmem + (uint32_t)(sl * OZ2_NC), da, db, oz_idesc((OZ_S - sl) * OZ2_NC), leader);
            else oz_mma<0>(tmem, da, db, oz_idesc(OZ_S * OZ2_NC), leader);
          }
        }
      }
      oz_commit(t_full, leader);      // accumulators of the tile complete -> epilogue
      oz_commit(a_empty, leader);     // every MMA reading the A stage has completed -> producer
      OZ_STAMP(it, 0, 2);
    }
  } else {
    // ===================================================== epilogue: row = TMEM lane = thread, 16 columns in two batches of 8
    const int ew = warp - 2;
    const int q = warp & 3;                      // TMEM lane quarter this warp may touch
    const int c16 = (ew >> 2) * 16;              // which 16 of the 32 columns
    const int row = q * 32 + lane;
    const double* __restrict__ expL = s_rep + (lane & (OZ2_REP - 1));
    double csum[2][8];
#pragma unroll
    for (int b8 = 0; b8 < 2; b8++)
#pragma unroll
      for (int j = 0; j < 8; j++) csum[b8][j] = 0.0;
    for (int it = 0; it < ntiles; it++) {
      oz_mbar_wait(a_full, (uint32_t)(it & 1));   // y and the row exponents travel with the tile
      const double yv = ((const double*)(sA + OZ_S * OZ_SLICE_A))[row];
      const int er = ((const int*)(sA + OZ_S * OZ_SLICE_A + OZ_M * 8))[row];
      __syncwarp();
      if (lane == 0) oz_mbar_arrive(a_empty);     // this warp no longer needs the stage
      const bool live = __double2hiint(yv) >= 0, y1 = (__double2hiint(yv) | __double2loint(yv)) != 0;   // y is 0, 1 or -1 (past n)
      const double rs = __hiloint2double((1023 + er - 35) << 20, 0);   // 2^(er - 35): the row's share of the scale
      const long long grow = (long long)(tile_lo + it) * OZ_M + row;
      OZ_STAMP(it, 0, 0);
      oz_mbar_wait(t_full, (uint32_t)(it & 1));
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      OZ_STAMP(it, 0, 1);
#pragma unroll
      for (int b8 = 0; b8 < 2; b8++) {
        const uint32_t base = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(c16 + b8 * 8);
        long long hi[8], lo[8];
        {
          int a[2][8], b[8];
          oz_ld8(base, a[0]); oz_ld8(base + (uint32_t)OZ2_NC, a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int j = 0; j < 8; j++) b[j] = a[0][j] * 128 + a[1][j];
          oz_ld8(base + (uint32_t)(2 * OZ2_NC), a[0]); oz_ld8(base + (uint32_t)(3 * OZ2_NC), a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int j = 0; j < 8; j++) hi[j] = (long long)b[j] * 16384 + (a[0][j] * 128 + a[1][j]);
          oz_ld8(base + (uint32_t)(4 * OZ2_NC), a[0]); oz_ld8(base + (uint32_t)(5 * OZ2_NC), a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int j = 0; j < 8; j++) b[j] = a[0][j] * 128 + a[1][j];
          oz_ld8(base + (uint32_t)(6 * OZ2_NC), a[0]); oz_ld8(base + (uint32_t)(7 * OZ2_NC), a[1]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
          if (b8 == 1) {   // everything of the stage is in registers: hand it back before the arithmetic
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) oz_mbar_arrive(t_empty);
          }
#pragma unroll
          for (int j = 0; j < 8; j++) lo[j] = (long long)b[j] * 16384 + (a[0][j] * 128 + a[1][j]);
        }
        OZ_STAMP(it, b8, 2);
        if ((dbg & 3) == 1) { csum[b8][0] += (double)(hi[0] + lo[7]); continue; }   // diagnostic: no epilogue arithmetic
#pragma unroll
        for (int cb = 0; cb < 2; cb++) {
          double eta[4], lp[4];
#pragma unroll
          for (int j = 0; j < 4; j++) {
            const double dh = __longlong_as_double(hi[cb * 4 + j] + 0x4338000000000000ll) - 6755399441055744.0;
            const double dl = __longlong_as_double(lo[cb * 4 + j] + 0x4178000000000000ll) - 25165824.0;
            eta[j] = ((dh + dl) * rs) * s_cs[c16 + b8 * 8 + cb * 4 + j];
          }
          { const bool yv4[4] = {y1, y1, y1, y1}; bern_logit_lp_rep<4, true, OZ2_REP>(eta, yv4, lp, expL, tab); }
#pragma unroll
          for (int j = 0; j < 4; j++)
            if (live) csum[b8][cb * 4 + j] += lp[j];
          if (eta_out) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
              const int c = chain0 + c16 + b8 * 8 + cb * 4 + j;
              if (live && grow < n && c < C) eta_out[(size_t)grow * C + c] = eta[j];
            }
          }
        }
        OZ_STAMP(it, b8, 3);
      }
    }
#pragma unroll
    for (int b8 = 0; b8 < 2; b8++)
#pragma unroll
      for (int j = 0; j < 8; j++) {
        double v = csum[b8][j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[q * OZ2_NC + c16 + b8 * 8 + j] = v;
      }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid < OZ2_NC) {
    const int c = chain0 + tid;
    if (c < C) partials[(size_t)chunk * C + c] = ((s_red[tid] + s_red[OZ2_NC + tid]) + s_red[2 * OZ2_NC + tid]) + s_red[3 * OZ2_NC + tid];
  }
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------- host side
static size_t oz_smem_bytes() {
  return 2 * (size_t)OZ_TILE_BYTES + OZ_S * OZ_SLICE_B + 8 * 8 + 8 + OZ_NC * 4 + 4 * OZ_NC * 8 + OZ_NC * 8 + sizeof(MathTab) + BL_REP_BYTES + 64;
}

size_t ozaki_xq_bytes(int n, int p) {
  if (p < 1 || p > OZ_K) return 0;
  return (size_t)((n + OZ_M - 1) / OZ_M) * OZ_TILE_BYTES;
}

// returns 0 when the sliced copy is usable, 1 when X holds values the slicing does not represent (non-finite, extreme exponents)
int ozaki_prepare_x(const double* X, const double* y, int n, int p, int ld, unsigned char* Xq, int* d_flag, cudaStream_t st) {
  cudaMemsetAsync(d_flag, 0, sizeof(int), st);
  const int rows = ((n + OZ_M - 1) / OZ_M) * OZ_M;
  k_oz_slice_x<<<(rows + 127) / 128, 128, 0, st>>>(X, y, n, p, ld, Xq, d_flag);
  int h = 0;
  cudaMemcpyAsync(&h, d_flag, sizeof(int), cudaMemcpyDeviceToHost, st);
  cudaStreamSynchronize(st);
  return h;
}

int ozaki_chunks(int n) {  // 16 chain tiles x 37 chunks = 4 waves of 148 one-CTA-per-SM blocks at 1024 chains
  return std::min((n + OZ_M - 1) / OZ_M, 37);
}

static size_t oz2_smem_bytes() {
  return (size_t)OZ_TILE_BYTES + OZ2_B_BYTES + 4 * 8 + 8 + 4 * OZ2_NC * 8 + OZ2_NC * 8 + sizeof(MathTab) + BL_REP_ENTRIES * OZ2_REP * 8 + 64;
}

void ozaki_bern_logit_launch(const unsigned char* Xq, int n, int p, const double* state, int beta_slot, int C, double* partials,
                             int tile_begin, int tile_end, double* eta_out, cudaStream_t st) {
  static std::atomic<unsigned long long> attr{0}, attr2{0};
  const int chunks = ozaki_chunks(n);
  const int ntiles = std::max(0, tile_end - tile_begin);
  const int tpc = (ntiles + chunks - 1) / chunks;
  const int dbg = getenv("NIMB200_OZ_DBG") ? atoi(getenv("NIMB200_OZ_DBG")) : 0;
  if (!getenv("NIMB200_OZ_ONE_CTA")) {   // two CTAs of 32 chains per SM
    ensure_dyn_smem(k_oz2_bern_logit, oz2_smem_bytes(), attr2);
    dim3 grid2((C + OZ2_NC - 1) / OZ2_NC, chunks);
    k_oz2_bern_logit<<<grid2, OZ2_THREADS, oz2_smem_bytes(), st>>>(Xq, n, p, state, beta_slot, C, std::max(1, tpc), tile_begin, tile_end,
                                                                 partials, eta_out, dbg);
    return;
  }
  ensure_dyn_smem(k_oz_bern_logit, oz_smem_bytes(), attr);
  dim3 grid((C + OZ_NC - 1) / OZ_NC, chunks);
  k_oz_bern_logit<<<grid, OZ_THREADS, oz_smem_bytes(), st>>>(Xq, n, p, state, beta_slot, C, std::max(1, tpc), tile_begin, tile_end,
                                                             partials, eta_out, getenv("NIMB200_OZ_DBG") ? atoi(getenv("NIMB200_OZ_DBG")) : 0);
}

}  // namespace nimb

// diagnostic: run nimb200_debug_ozaki-style launches with the timeline stamps of CTA (0, 0) switched on; out[18][64][2][4]
extern "C" int nimb200_debug_ozaki_trace(long long* out_or_null_to_stop) {
  using namespace nimb;
  static long long* dbuf = nullptr;
  const size_t n = (size_t)18 * 64 * 2 * 4;
  if (!dbuf) {
    if (cudaMalloc(&dbuf, n * sizeof(long long)) != cudaSuccess) return 1;
    cudaMemset(dbuf, 0, n * sizeof(long long));
    if (cudaMemcpyToSymbol(g_oz_trace, &dbuf, sizeof(dbuf)) != cudaSuccess) return 1;
    return 0;
  }
  if (out_or_null_to_stop && cudaMemcpy(out_or_null_to_stop, dbuf, n * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess) return 1;
  long long* z = nullptr;
  cudaMemcpyToSymbol(g_oz_trace, &z, sizeof(z));
  cudaFree(dbuf);
  dbuf = nullptr;
  return 0;
}

// diagnostic (tests, measurements): linear predictors and per-chain block sums of the tensor-core path for host arrays
// X [n][p] row-major, y [n], beta [p][C] (chain fastest); eta_out [n][C] and sums_out [C] may each be null.
extern "C" int nimb200_debug_ozaki(const double* X, const double* y, int32_t n, int32_t p, const double* beta, int32_t C,
                                   double* eta_out, double* sums_out, double* ms_out) {
  using namespace nimb;
  if (!X || !y || !beta || n < 1 || p < 1 || p > OZ_K || C < 1) { set_error("nimb200_debug_ozaki: bad argument"); return 1; }
  double *dX = nullptr, *dy = nullptr, *db = nullptr, *dpart = nullptr, *deta = nullptr;
  unsigned char* dq = nullptr;
  int* dflag = nullptr;
  const int chunks = ozaki_chunks(n), ntiles = (n + OZ_M - 1) / OZ_M;
  cudaError_t e = cudaSuccess;
  auto ck = [&](cudaError_t x) { if (e == cudaSuccess) e = x; };
  ck(cudaMalloc(&dX, sizeof(double) * (size_t)n * p));
  ck(cudaMalloc(&dy, sizeof(double) * n));
  ck(cudaMalloc(&db, sizeof(double) * (size_t)p * C));
  ck(cudaMalloc(&dpart, sizeof(double) * (size_t)chunks * C));
  ck(cudaMalloc(&dq, ozaki_xq_bytes(n, p)));
  ck(cudaMalloc(&dflag, sizeof(int)));
  if (eta_out) ck(cudaMalloc(&deta, sizeof(double) * (size_t)n * C));
  int rc = 0;
  if (e == cudaSuccess) {
    ck(cudaMemcpy(dX, X, sizeof(double) * (size_t)n * p, cudaMemcpyHostToDevice));
    ck(cudaMemcpy(dy, y, sizeof(double) * n, cudaMemcpyHostToDevice));
    ck(cudaMemcpy(db, beta, sizeof(double) * (size_t)p * C, cudaMemcpyHostToDevice));
    if (ozaki_prepare_x(dX, dy, n, p, p, dq, dflag, 0)) { set_error("nimb200_debug_ozaki: X is not representable by the slicing"); rc = 1; }
    if (!rc) {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      ozaki_bern_logit_launch(dq, n, p, db, 0, C, dpart, 0, ntiles, deta, 0);   // warm
      const int reps = ms_out ? 5 : 1;
      cudaEventRecord(e0, 0);
      for (int r = 0; r < reps; r++) ozaki_bern_logit_launch(dq, n, p, db, 0, C, dpart, 0, ntiles, deta, 0);
      cudaEventRecord(e1, 0);
      ck(cudaDeviceSynchronize());
      ck(cudaGetLastError());
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms_out) *ms_out = ms / reps;
      cudaEventDestroy(e0); cudaEventDestroy(e1);
      if (e == cudaSuccess && sums_out) {
        std::vector<double> hp((size_t)chunks * C);
        ck(cudaMemcpy(hp.data(), dpart, sizeof(double) * hp.size(), cudaMemcpyDeviceToHost));
        for (int c = 0; c < C; c++) { double t = 0; for (int k = 0; k < chunks; k++) t += hp[(size_t)k * C + c]; sums_out[c] = t; }
      }
      if (e == cudaSuccess && eta_out) ck(cudaMemcpy(eta_out, deta, sizeof(double) * (size_t)n * C, cudaMemcpyDeviceToHost));
    }
  }
  cudaFree(dX); cudaFree(dy); cudaFree(db); cudaFree(dpart); cudaFree(dq); cudaFree(dflag); cudaFree(deta);
  if (e != cudaSuccess) { set_error(std::string("nimb200_debug_ozaki: CUDA error ") + cudaGetErrorString(e)); return 1; }
  return rc;
}
