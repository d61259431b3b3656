// One engine instantiation ("variant") = the rollout kernel + host staging for one Traits capacity
// class.  Each variant is its own translation unit so the three kernels compile in parallel.
#pragma once
#include "ps_kernels.cuh"
#include "ps_coop_rollout.cuh"
#include <cstddef>
#include <cstdlib>
#include "ps_host_model.h"

namespace ps {

// Threads per block for a thread-per-rollout launch of n rollouts: small batches are latency bound
// (one warp per scheduler at best), so they are spread over as many SMs as possible -- each warp then
// has an SM's instruction cache and L1 to itself (measured: acrobot N=4096 5.87 -> 4.60 ms); large
// batches use the full block so one Constants copy serves more threads.
inline int pick_block(int n, int max_block) {
  static int sms = 0;
  if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
  int b = 32;
  while (b < max_block && (n + b - 1) / b > sms) b <<= 1;
  return b;
}

// thread-per-rollout launch; returns the number of blocks (= entries written to block_min)
template <class T, int BLOCK>
inline int launch_rollout(cudaStream_t st, const void* dconst, const double* knots, int stride, int first, int n,
                          long long offset, double* returns, unsigned char* failure, MinLoc* block_min, Writeback wb, bool writeback,
                          int* counter) {
  (void)counter;
  const int block = pick_block(n, BLOCK), grid = (n + block - 1) / block;
  const Constants<T>* c = reinterpret_cast<const Constants<T>*>(dconst);
  if (writeback) rollout_kernel<T, BLOCK, true><<<grid, block, 0, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb);
  else rollout_kernel<T, BLOCK, false><<<grid, block, 0, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb);
  return grid;
}

// force-perturbed re-rolls of candidate policies (dense thread-per-rollout instantiations only)
template <class T, int BLOCK>
inline int launch_robust(cudaStream_t st, const void* dconst, const double* knots, int stride, int n, double* returns,
                         unsigned char* failure, Writeback wb, XfrcNoise xf) {
  const int block = pick_block(n, BLOCK), grid = (n + block - 1) / block;
  robust_rollout_kernel<T, BLOCK><<<grid, block, 0, st>>>(reinterpret_cast<const Constants<T>*>(dconst), knots, stride, n, returns, failure, wb, xf);
  return grid;
}

// force-perturbed re-rolls on the warp-per-rollout coding
template <class T, int WARPS>
inline int launch_coop_robust(cudaStream_t st, const void* dconst, const double* knots, int stride, int n, double* returns,
                              unsigned char* failure, Writeback wb, XfrcNoise xf) {
  constexpr size_t kConst = (sizeof(Constants<T>) + 127) / 128 * 128, kData = (sizeof(CoopData<T>) + 127) / 128 * 128;
  constexpr size_t kXf = (sizeof(double) * 6 * T::NBODY + 127) / 128 * 128;
  constexpr size_t smem = kConst + WARPS * (kData + kXf);
  static_assert(smem <= 227 * 1024, "robust warp-per-rollout kernel: working sets exceed the shared memory of an SM");
  static bool configured = false;
  if (!configured) { cudaFuncSetAttribute(coop_robust_kernel<T, WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); configured = true; }
  const int grid = (n + WARPS - 1) / WARPS;
  coop_robust_kernel<T, WARPS><<<grid, WARPS * 32, smem, st>>>(reinterpret_cast<const Constants<T>*>(dconst), knots, stride, n, returns, failure, wb, xf);
  return grid;
}

// warp-per-rollout persistent launch: as many blocks as stay resident, WARPS rollouts in flight per block
template <class T, int WARPS>
inline int launch_coop(cudaStream_t st, const void* dconst, const double* knots, int stride, int first, int n,
                       long long offset, double* returns, unsigned char* failure, MinLoc* block_min, Writeback wb, bool writeback,
                       int* counter) {
  constexpr size_t kConst = (sizeof(Constants<T>) + 127) / 128 * 128, kData = (sizeof(CoopData<T>) + 127) / 128 * 128;
  constexpr size_t smem = kConst + WARPS * kData;
  static int resident[2] = {0, 0};                 // blocks the device keeps resident, per kernel flavour
  static int sm_count = 0;
  const int f = writeback ? 1 : 0;
  if (!resident[f]) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    sm_count = sms > 0 ? sms : 148;
    if (writeback) {
      cudaFuncSetAttribute(coop_rollout_kernel<T, WARPS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, coop_rollout_kernel<T, WARPS, true>, WARPS * 32, smem);
    } else {
      cudaFuncSetAttribute(coop_rollout_kernel<T, WARPS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, coop_rollout_kernel<T, WARPS, false>, WARPS * 32, smem);
    }
    resident[f] = sms * (per_sm > 0 ? per_sm : 1);
  }
  // small batches are spread over the SMs: only as many warps per block pull rollouts as the batch needs (coop_rollout_kernel)
  int active = (n + sm_count - 1) / sm_count;
  if (active < 1) active = 1;
  if (active > WARPS) active = WARPS;
  int grid = (n + active - 1) / active;
  if (grid > resident[f]) grid = resident[f];
  const Constants<T>* c = reinterpret_cast<const Constants<T>*>(dconst);
  cudaMemsetAsync(counter, 0, sizeof(int), st);
  if (writeback) coop_rollout_kernel<T, WARPS, true><<<grid, WARPS * 32, smem, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb, counter, active);
  else coop_rollout_kernel<T, WARPS, false><<<grid, WARPS * 32, smem, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb, counter, active);
  return grid;
}

#ifndef PS_SPARSE_BLOCKS_PER_SM_DEFAULT
#define PS_SPARSE_BLOCKS_PER_SM_DEFAULT 0       // 0: one block per group of rollouts (no cap)
#endif
// thread-per-rollout launch of the compressed-Jacobian phases (sparse_rollout_kernel)
template <class T, int BLOCK>
inline int launch_sparse(cudaStream_t st, const void* dconst, const double* knots, int stride, int first, int n,
                         long long offset, double* returns, unsigned char* failure, MinLoc* block_min, Writeback wb, bool writeback,
                         int* counter) {
  (void)counter;
  // the block walks its rollouts in lockstep (Lanes1Block): one block per SM, as large as the batch allows -- the wider
  // the block, the more warps share each instruction-cache line (measured: 512-thread blocks +22 % over 128-thread ones)
  static int sms_ = 0;
  if (!sms_) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms_, cudaDevAttrMultiProcessorCount, dev); if (sms_ <= 0) sms_ = 148; }
  int block = pick_block(n, BLOCK);
  if (n >= sms_ * 64) { block = ((n + sms_ - 1) / sms_ + 31) & ~31; if (block > BLOCK) block = BLOCK; }
  static const int block_rt = std::getenv("PS_SPARSE_BLOCK_RT") ? std::atoi(std::getenv("PS_SPARSE_BLOCK_RT")) : 0;   // diagnostics / tuning
  if (block_rt >= 32 && block_rt <= BLOCK && block > block_rt) block = block_rt & ~31;
  int grid = (n + block - 1) / block;
  // PS_SPARSE_BLOCKS_PER_SM (diagnostics / tuning): cap the grid at that many blocks per SM; the kernel strides over the batch
  static const int per_sm = std::getenv("PS_SPARSE_BLOCKS_PER_SM") ? std::atoi(std::getenv("PS_SPARSE_BLOCKS_PER_SM")) : PS_SPARSE_BLOCKS_PER_SM_DEFAULT;
  if (per_sm > 0) {
    static int sms = 0;
    if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
    if (grid > sms * per_sm) grid = sms * per_sm;
  }
  const Constants<T>* c = reinterpret_cast<const Constants<T>*>(dconst);
  if (writeback) sparse_rollout_kernel<T, BLOCK, true><<<grid, block, 0, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb);
  else sparse_rollout_kernel<T, BLOCK, false><<<grid, block, 0, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb);
  return grid;
}

struct Variant {
  const char* name; size_t const_bytes; size_t plan_offset, plan_bytes; int block; bool coop; bool euler_only;
  int (*fill)(const PsModel*, const PairList&, void*, std::string*);
  void (*options)(void*, double, int);
  void (*debug)(void*, int);
  void (*task)(void*, const DevTask&, int);
  int (*plan)(void*, const double*, int, const double*, int, double, const double*, int, int, int);
  int (*launch)(cudaStream_t, const void*, const double*, int, int, int, long long, double*, unsigned char*, MinLoc*, Writeback, bool, int*);
  int (*robust)(cudaStream_t, const void*, const double*, int, int, double*, unsigned char*, Writeback, XfrcNoise);   // null: no force perturbations
  int (*task_ok)(const DevTask&, int) = nullptr;
  // re-roll of the samples that overflowed this kernel's contact / row capacity (null: the class has no contacts or no larger class)
  void (*fixup)(cudaStream_t, const void*, const double*, int, double*, unsigned char*, const int*, const int*, int*, Writeback, bool) = nullptr;   // model-specialised instantiations: was the kernel compiled for this cost specification?
  // qacc_warmstart of the next launches' rollouts (null pointer: zeros) -- DevPlan::warmstart, ps_set_warmstart
  void (*warmstart)(void*, const double*, int) = nullptr;
};
// the force-perturbed kernels live in their own translation unit (ps_variant_robust.cu), so that the code
// generated for the planning kernels of a variant's unit does not depend on them
using RobustFn = int (*)(cudaStream_t, const void*, const double*, int, int, double*, unsigned char*, Writeback, XfrcNoise);
int launch_robust_tiny(cudaStream_t, const void*, const double*, int, int, double*, unsigned char*, Writeback, XfrcNoise);
int launch_robust_planar(cudaStream_t, const void*, const double*, int, int, double*, unsigned char*, Writeback, XfrcNoise);
int launch_robust_legged(cudaStream_t, const void*, const double*, int, int, double*, unsigned char*, Writeback, XfrcNoise);
int launch_robust_planar_coop(cudaStream_t, const void*, const double*, int, int, double*, unsigned char*, Writeback, XfrcNoise);
int launch_robust_legged_coop(cudaStream_t, const void*, const double*, int, int, double*, unsigned char*, Writeback, XfrcNoise);
template <class T, int BLOCK>
inline Variant make_variant(const char* name, RobustFn robust) {
  Variant v{name, sizeof(Constants<T>), offsetof(Constants<T>, p), sizeof(DevPlan<T>), BLOCK, false, T::LEAN, fill_model<T>, set_options<T>, set_debug<T>, set_task<T>, set_plan<T>, launch_rollout<T, BLOCK>, robust};
  v.warmstart = set_warmstart<T>;
  return v;
}
template <class T, int WARPS>
inline Variant make_coop_variant(const char* name, RobustFn robust = nullptr) {
  Variant v{name, sizeof(Constants<T>), offsetof(Constants<T>, p), sizeof(DevPlan<T>), WARPS * 32, true, true, fill_model<T>, set_options<T>, set_debug<T>, set_task<T>, set_plan<T>, launch_coop<T, WARPS>, robust};
  v.warmstart = set_warmstart<T>;
  return v;
}
// model-specialised lean instantiation (SpecTraits<Spec>, ps_types.cuh): compile-time topology and cost layout
template <class T, int BLOCK>
inline Variant make_spec_variant(const char* name) {
  Variant v{name, sizeof(Constants<T>), offsetof(Constants<T>, p), sizeof(DevPlan<T>), BLOCK, false, true, fill_model_spec<T>, set_options_spec<T>,
            set_debug_spec<T>, set_task_spec<T>, set_plan<T>, launch_rollout<T, BLOCK>, nullptr};
  v.task_ok = task_matches_spec<T>;
  v.warmstart = set_warmstart<T>;
  return v;
}
#define PS_DECL_SPEC_VARIANT(N) const Variant* variant_spec_##N();
PS_SPEC_LIST(PS_DECL_SPEC_VARIANT)
#undef PS_DECL_SPEC_VARIANT
const Variant* variant_tiny();
const Variant* variant_planar();
const Variant* variant_legged();
template <class T, int BLOCK>
inline Variant make_sparse_variant(const char* name) {
  Variant v{name, sizeof(Constants<T>), offsetof(Constants<T>, p), sizeof(DevPlan<T>), BLOCK, true, true, fill_model<T>, set_options<T>, set_debug<T>, set_task<T>, set_plan<T>, launch_sparse<T, BLOCK>, nullptr};
  v.warmstart = set_warmstart<T>;
  return v;
}
// capacity-overflow re-rolls of the legged class (ps_variant_legged_xl.cu)
void launch_fixup_legged(cudaStream_t, const void*, const double*, int, double*, unsigned char*, const int*, const int*, int*, Writeback, bool);
const Variant* variant_planar_sparse();
const Variant* variant_legged_sparse();
const Variant* variant_tiny_coop();
const Variant* variant_planar_coop();
const Variant* variant_legged_coop();

}  // namespace ps
