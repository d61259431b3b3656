// Capacity-overflow fix-up of the legged (quadruped) class: coop_fixup_kernel (one warp per flagged rollout) at the LeggedXLT
// capacities (64 contacts / 256 constraint rows), shared by the three legged scoring kernels.
#include "ps_variant.h"
namespace ps {
void launch_fixup_legged(cudaStream_t st, const void* dconst, const double* knots, int stride, double* returns, unsigned char* failure,
                         const int* list, const int* count, int* counter, Writeback wb, bool writeback) {
  constexpr int WARPS = 2;
  constexpr size_t kConst = (sizeof(Constants<LeggedXLT>) + 127) / 128 * 128, kData = (sizeof(CoopData<LeggedXLT>) + 127) / 128 * 128;
  constexpr size_t smem = kConst + WARPS * kData;
  static_assert(smem <= 227 * 1024, "fix-up kernel: two large-capacity working sets must fit the shared memory of an SM");
  static int sms = 0;
  if (!sms) {
    int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148;
    cudaFuncSetAttribute(coop_fixup_kernel<LeggedXLT, WARPS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(coop_fixup_kernel<LeggedXLT, WARPS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  const Constants<LeggedXLT>* c = reinterpret_cast<const Constants<LeggedXLT>*>(dconst);
  cudaMemsetAsync(counter, 0, sizeof(int), st);
  if (writeback) coop_fixup_kernel<LeggedXLT, WARPS, true><<<sms, WARPS * 32, smem, st>>>(c, knots, stride, returns, failure, list, count, counter, wb);
  else coop_fixup_kernel<LeggedXLT, WARPS, false><<<sms, WARPS * 32, smem, st>>>(c, knots, stride, returns, failure, list, count, counter, wb);
}
}  // namespace ps
