// Non-template kernels of the planning iteration (compiled once, in ps_engine.cu): K1 knot-noise
// perturbation, K3 grid-level min-loc + winner record, and the SoA read-back helpers.
#pragma once
#include "ps_kernels.cuh"

namespace ps {

// ---------------------------------------------------------------- counter-based RNG (Philox4x32-10)
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
  c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
}
__device__ __forceinline__ void philox4x32(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int i = 0; i < 10; i++) { philox_round(c, k0, k1); k0 += 0x9E3779B9u; k1 += 0xBB67AE85u; }
}
__device__ __forceinline__ double u01_open(uint32_t hi, uint32_t lo) {   // (0, 1), 53 bits
  const unsigned long long x = ((unsigned long long)hi << 32) | lo;
  return ((double)(x >> 11) + 0.5) * (1.0 / 9007199254740992.0);
}

// ---------------------------------------------------------------- K1
// grid: (ceil(n / block), P*nu); knots[(e) * stride + i] for local sample i, element e = p*nu + a
__global__ void perturb_kernel(const double* __restrict__ nominal, const double* __restrict__ ctrlrange, int n, int nelem, int nu,
                               double sigma0, double sigma1, unsigned long long seed, unsigned long long iteration,
                               long long sample_offset, const double* __restrict__ z_injected,
                               const unsigned char* __restrict__ std2_injected, double* __restrict__ knots, int stride,
                               const double* __restrict__ param_std) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int e = blockIdx.y;
  if (i >= n) return;
  const long long gi = sample_offset + i;          // global sample index (rank-independent noise)
  const int a = e % nu;
  const double lo = ctrlrange[2 * a], hi = ctrlrange[2 * a + 1];
  double v = nominal[e];
  // Predictive Sampling leaves sample 0 un-noised (sampling/planner.cc:361); the cross-entropy planner
  // perturbs every candidate (cross_entropy/planner.cc:395-409) and rolls the nominal out separately:
  // here that is the LAST column of the batch (param_std != nullptr selects this mode)
  const bool noisy = param_std ? (i != n - 1) : (gi != 0);
  if (noisy) {
    double sd = sigma0, z;
    if (z_injected) {
      z = z_injected[(size_t)i * nelem + e];
      if (sigma1 > 0 && std2_injected && std2_injected[i]) sd = sigma1;
    } else {
      const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
      if (sigma1 > 0) {
        uint32_t c[4] = {(uint32_t)gi, (uint32_t)(gi >> 32), 0xFFFFFFFFu, (uint32_t)iteration};
        philox4x32(c, k0, k1);
        if (u01_open(c[0], c[1]) < 0.2) sd = sigma1;     // absl::Bernoulli(gen, 0.2), planner.cc:323
      }
      uint32_t c[4] = {(uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)e, (uint32_t)iteration};
      philox4x32(c, k0, k1);
      const double u1 = u01_open(c[0], c[1]), u2 = u01_open(c[2], c[3]);
      z = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);        // Box-Muller
    }
    if (param_std) v += param_std[e] * z;                  // N(0, max(sqrt(variance_k), std_min)), cross_entropy/planner.cc:360-363
    else { const double scale = 0.5 * (hi - lo); v += (scale * sd) * z; }
    v = clampd(v, lo, hi);
  }
  knots[(size_t)e * stride + i] = v;
}

// standard-normal fill (Philox + Box-Muller, one draw per element): the force-perturbation tensor of the
// robust rollouts when no noise is injected
__global__ void normal_fill_kernel(double* __restrict__ out, size_t n, unsigned long long seed, unsigned long long iteration) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t c[4] = {(uint32_t)i, (uint32_t)(i >> 32), 0xF0F0F0F0u, (uint32_t)iteration};
  philox4x32(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  out[i] = sqrt(-2.0 * log(u01_open(c[0], c[1]))) * cospi(2.0 * u01_open(c[2], c[3]));
}

// ---------------------------------------------------------------- capacity-overflow fix-up support
// compact list of the samples whose rollout overflowed the scoring kernel's contact / row capacity (flag bit 1)
__global__ void collect_flagged_kernel(const unsigned char* __restrict__ failure, int first, int n, int* __restrict__ list, int* __restrict__ count) {
  const int i = first + blockIdx.x * blockDim.x + threadIdx.x;
  if (i < first + n && (failure[i] & 2)) list[atomicAdd(count, 1)] = i;
}
// block-level (return, global index) min straight from the per-sample returns: the input of K3 whenever returns may have
// been rewritten after K2 (fix-up pass).  One entry per 1024 samples.
__global__ void __launch_bounds__(256)
minloc_returns_kernel(const double* __restrict__ returns, int n, long long sample_offset, MinLoc* __restrict__ block_min) {
  __shared__ MinLoc wm[8];
  double c = 1.0e300; long long gi = 0x7fffffffffffffffLL;
  for (int k = 0; k < 4; k++) {
    const int i = blockIdx.x * 1024 + k * 256 + threadIdx.x;
    if (i < n) { double v = returns[i]; if (!(v == v)) v = 1.0e300; if (better(v, sample_offset + i, c, gi)) { c = v; gi = sample_offset + i; } }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_down_sync(0xffffffffu, c, o); const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
    if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { wm[threadIdx.x >> 5].cost = c; wm[threadIdx.x >> 5].index = gi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; w++) if (better(wm[w].cost, wm[w].index, c, gi)) { c = wm[w].cost; gi = wm[w].index; }
    block_min[blockIdx.x].cost = c; block_min[blockIdx.x].index = gi;
  }
}

// ---------------------------------------------------------------- K3
// record = [winner_cost, winner_global_index, nominal_cost (local sample 0), knots[nelem]...]
__global__ void argmin_kernel(const MinLoc* __restrict__ block_min, int nblocks, const double* __restrict__ returns,
                              const double* __restrict__ knots, int stride, int nelem, long long sample_offset,
                              double* __restrict__ record) {
  __shared__ MinLoc wm[32];
  __shared__ long long win;
  double c = 1.0e300; long long gi = 0x7fffffffffffffffLL;
  for (int b = threadIdx.x; b < nblocks; b += blockDim.x) {
    const MinLoc v = block_min[b];
    if (better(v.cost, v.index, c, gi)) { c = v.cost; gi = v.index; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_down_sync(0xffffffffu, c, o);
    const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
    if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { wm[threadIdx.x >> 5].cost = c; wm[threadIdx.x >> 5].index = gi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = (blockDim.x + 31) / 32;
    c = threadIdx.x < nw ? wm[threadIdx.x].cost : 1.0e300;
    gi = threadIdx.x < nw ? wm[threadIdx.x].index : 0x7fffffffffffffffLL;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double oc = __shfl_down_sync(0xffffffffu, c, o);
      const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
      if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
    }
    if (threadIdx.x == 0) {
      if (gi == 0x7fffffffffffffffLL) gi = sample_offset;     // all NaN: fall back to the first sample
      record[0] = c; record[1] = (double)gi; record[2] = returns[0]; win = gi - sample_offset;
    }
  }
  __syncthreads();
  const long long w = win;
  for (int e = threadIdx.x; e < nelem; e += blockDim.x) record[3 + e] = knots[(size_t)e * stride + w];
}

// ---------------------------------------------------------------- K3x: cross-rank min-loc over gathered winner records
// all[world][count] (one ncclAllGather of every rank's record) -> out[count] = the record with the lowest return,
// ties to the lowest global sample index; out[2] = the nominal return, i.e. rank 0's (global sample 0 lives there).
__global__ void select_winner_kernel(const double* __restrict__ all, int world, int count, double* __restrict__ out) {
  __shared__ int best_s;
  if (threadIdx.x == 0) {
    int best = 0;
    for (int r = 1; r < world; r++) {
      double c = all[(size_t)r * count], bc = all[(size_t)best * count];
      if (!(c == c)) c = 1.0e300;
      if (!(bc == bc)) bc = 1.0e300;
      if (c < bc || (c == bc && all[(size_t)r * count + 1] < all[(size_t)best * count + 1])) best = r;
    }
    best_s = best;
  }
  __syncthreads();
  const double* src = all + (size_t)best_s * count;
  for (int k = threadIdx.x; k < count; k += blockDim.x) out[k] = k == 2 ? all[2] : src[k];
}

// ---------------------------------------------------------------- K4 / K5: cross-entropy elite update
// Rank of every candidate under (return, index) order -- the total order of the partial_sort in
// cross_entropy/planner.cc:201-207 with ties broken by index; NaN sorts last.  O(n^2 / tile) compares
// through a shared-memory tile of the returns: 65,536 candidates = 4.3e9 compares, well under a
// millisecond of the iteration.  Candidates ranked below n_elite write their index to elite[rank].
__global__ void rank_kernel(const double* __restrict__ returns, int n, int n_elite, int* __restrict__ elite) {
  __shared__ double tile[1024];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double mine = i < n ? returns[i] : 0.0;
  if (!(mine == mine)) mine = 1.0e300;
  int rank = 0;
  for (int base = 0; base < n; base += 1024) {
    const int cnt = n - base < 1024 ? n - base : 1024;
    __syncthreads();
    for (int k = threadIdx.x; k < cnt; k += blockDim.x) { double v = returns[base + k]; tile[k] = (v == v) ? v : 1.0e300; }
    __syncthreads();
    for (int k = 0; k < cnt; k++) { const double v = tile[k]; rank += (v < mine) || (v == mine && base + k < i); }
  }
  if (i < n && rank < n_elite) elite[rank] = i;
}

// out = [avg_return, best_return, mean[nelem], variance[nelem]].  One thread per spline parameter; the
// sums run over the elites in rank order, sequentially, exactly as cross_entropy/planner.cc:224-262
// (including its variance, which measures the BEST candidate's distance from the elite mean n_elite
// times -- `n` is fetched once from trajectory_order[0] at :246 and not per elite).
__global__ void elite_stats_kernel(const double* __restrict__ returns, const double* __restrict__ knots, int stride, int nelem,
                                   const int* __restrict__ elite, int n_elite, double* __restrict__ out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e == 0) {
    double s = 0;
    for (int r = 0; r < n_elite; r++) s += returns[elite[r]];
    out[0] = s / n_elite; out[1] = returns[elite[0]];
  }
  if (e >= nelem) return;
  const double* row = knots + (size_t)e * stride;
  double s = 0;
  for (int r = 0; r < n_elite; r++) s += row[elite[r]];
  const double avg = s * (1.0 / n_elite);                 // mju_scl(.., 1.0 / n_elite), :239
  const double diff = row[elite[0]] - avg;
  double var = 0;
  for (int r = 0; r < n_elite; r++) var += diff * diff / (n_elite - 1);
  out[2 + e] = avg; out[2 + nelem + e] = var;
}

// gather one sample's SoA column into a dense host-order array: out[r] = src[r * stride + i]
__global__ void gather_column_kernel(const double* __restrict__ src, int rows, int stride, int i, double* __restrict__ out) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r < rows) out[r] = src[(size_t)r * stride + i];
}

// transpose SoA [rows][stride] -> [n][rows] (parity read-back of all samples)
__global__ void transpose_kernel(const double* __restrict__ src, int rows, int stride, int n, double* __restrict__ out) {
  __shared__ double tile[32][33];
  const int x = blockIdx.x * 32 + threadIdx.x;   // sample
  for (int ry = threadIdx.y; ry < 32; ry += blockDim.y) {
    const int r = blockIdx.y * 32 + ry;
    tile[ry][threadIdx.x] = (r < rows && x < n) ? src[(size_t)r * stride + x] : 0.0;
  }
  __syncthreads();
  const int r2 = blockIdx.y * 32 + threadIdx.x;  // row
  for (int sy = threadIdx.y; sy < 32; sy += blockDim.y) {
    const int s = blockIdx.x * 32 + sy;
    if (s < n && r2 < rows) out[(size_t)s * rows + r2] = tile[threadIdx.x][sy];
  }
}

// FP64 FMA throughput probe: the roofline denominator for the rollout kernel.
// 8 independent dependency chains per thread, 2 flops per DFMA.
__global__ void dfma_probe_kernel(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

}  // namespace ps
