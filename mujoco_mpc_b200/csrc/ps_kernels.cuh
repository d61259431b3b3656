// sm_100a kernels of one Predictive Sampling planning iteration:
//   K1 perturb_kernel  : candidate knots = clamp(nominal + sigma * scale * z)      (AddNoiseToPolicy,
//                        mjpc/planners/sampling/planner.cc:313-339; sample 0 stays un-noised, :361)
//   K2 rollout_kernel  : per-sample time recursion  policy -> forward pass -> residual -> cost ->
//                        integrate  (Trajectory::NoisyRollout + UpdateReturn, mjpc/trajectory.cc:100-210,
//                        312-326), block-level (cost, index) min
//   K3 argmin_kernel   : grid-level min-loc with lowest-index tie break, packs the winner record
//                        (OptimizePolicyCandidates tail, planner.cc:168-187)
// Layout: everything per-sample is structure-of-arrays with the sample index fastest, so the 32
// rollouts of a warp read/write 256 contiguous bytes per field.
#pragma once
#include <cuda_runtime.h>
#include "ps_rollout.cuh"

namespace ps {

// ---------------------------------------------------------------- TMA bulk copy helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

struct MinLoc { double cost; long long index; };
__device__ __forceinline__ bool better(double c, long long i, double bc, long long bi) {
  return (c < bc) || (c == bc && i < bi);
}

// ---------------------------------------------------------------- K2
// model-specialised instantiations keep their working set in registers: cap the count so that PS_SPEC_MINBLOCKS blocks stay
// resident per SM (the generic instantiations keep the plain bound they were tuned with)
#ifndef PS_SPEC_MINBLOCKS
#define PS_SPEC_MINBLOCKS 2
#endif
template <class T, int BLOCK, bool WB>
__global__ void __launch_bounds__(BLOCK, (T::SPEC ? PS_SPEC_MINBLOCKS : 1))
rollout_kernel(const Constants<T>* __restrict__ gconst, const double* __restrict__ knots, int stride, int first, int n,
               long long sample_offset, double* __restrict__ returns, unsigned char* __restrict__ failure,
               MinLoc* __restrict__ block_min, Writeback wb) {
  __shared__ Constants<T> sc;
  __shared__ alignas(8) uint64_t bar;
  __shared__ MinLoc warp_min[BLOCK / 32];

  // stage the model / task / plan constants with one TMA bulk copy
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  if (threadIdx.x == 0) {
    constexpr uint32_t total = (uint32_t)sizeof(Constants<T>);
    mbar_expect_tx(&bar, total);
    const char* src = reinterpret_cast<const char*>(gconst); char* dst = reinterpret_cast<char*>(&sc);
    for (uint32_t off = 0; off < total; off += 32768u) {
      const uint32_t sz = total - off < 32768u ? total - off : 32768u;
      tma_bulk_g2s(dst + off, src + off, sz, &bar);
    }
  }
  mbar_wait(&bar, 0);

  const int i = first + blockIdx.x * blockDim.x + threadIdx.x;  // local sample index (blockDim.x <= BLOCK, chosen per batch)
  const bool active = i < first + n;
  double total = 1.0e300; bool failed = false;

  if (active) {
    int flags = 0;
    if constexpr (T::SPEC) {      // model-specialised: the cost specification's integers are compile-time too
      const SpecTaskView<typename T::Topo> tv(sc.t);
      total = rollout_sample<T, WB, false, SpecTaskView<typename T::Topo>>(sc.m, tv, sc.p, knots, stride, i, wb, &flags);
    } else total = rollout_sample<T, WB>(sc.m, sc.t, sc.p, knots, stride, i, wb, &flags);
    failed = flags & 1;
    returns[i] = total;
    failure[i] = (unsigned char)flags;
  }

  // block-level min-loc: warp shuffles, then one warp over the per-warp results
  double c = active ? total : 1.0e300; long long gi = active ? sample_offset + i : 0x7fffffffffffffffLL;
  if (!(c == c)) c = 1.0e300;                        // NaN never wins
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_down_sync(0xffffffffu, c, o);
    const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
    if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { warp_min[threadIdx.x >> 5].cost = c; warp_min[threadIdx.x >> 5].index = gi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    c = threadIdx.x < (blockDim.x >> 5) ? warp_min[threadIdx.x].cost : 1.0e300;
    gi = threadIdx.x < (blockDim.x >> 5) ? warp_min[threadIdx.x].index : 0x7fffffffffffffffLL;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double oc = __shfl_down_sync(0xffffffffu, c, o);
      const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
      if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
    }
    if (threadIdx.x == 0) { block_min[blockIdx.x].cost = c; block_min[blockIdx.x].index = gi; }
  }
}

// ---------------------------------------------------------------- robust re-rolls
// RobustPlanner's NoisyRollout (mjpc/trajectory.cc:100-210 with xfrc_std > 0): the same thread-per-rollout
// recursion with full writeback plus Ornstein-Uhlenbeck forces on every body.  Its own kernel so that the
// planning kernels above carry none of this; batches are candidates x repetitions, i.e. small.
template <class T, int BLOCK>
__global__ void __launch_bounds__(BLOCK)
robust_rollout_kernel(const Constants<T>* __restrict__ gconst, const double* __restrict__ knots, int stride, int n,
                      double* __restrict__ returns, unsigned char* __restrict__ failure, Writeback wb, XfrcNoise xf) {
  __shared__ Constants<T> sc;
  __shared__ alignas(8) uint64_t bar;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  if (threadIdx.x == 0) {
    constexpr uint32_t total = (uint32_t)sizeof(Constants<T>);
    mbar_expect_tx(&bar, total);
    const char* src = reinterpret_cast<const char*>(gconst); char* dst = reinterpret_cast<char*>(&sc);
    for (uint32_t off = 0; off < total; off += 32768u) {
      const uint32_t sz = total - off < 32768u ? total - off : 32768u;
      tma_bulk_g2s(dst + off, src + off, sz, &bar);
    }
  }
  mbar_wait(&bar, 0);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int flags = 0;
  const double total = rollout_sample<T, true, true>(sc.m, sc.t, sc.p, knots, stride, i, wb, &flags, &xf);
  returns[i] = total;
  failure[i] = (unsigned char)flags;
}

}  // namespace ps
