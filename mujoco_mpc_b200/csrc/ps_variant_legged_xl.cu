// Capacity-overflow fix-up of the legged (quadruped) class: sparse_fixup_kernel at the LeggedXLT capacities
// (64 contacts / 256 constraint rows), shared by the three legged scoring kernels.
#include "ps_variant.h"
namespace ps {
void launch_fixup_legged(cudaStream_t st, const void* dconst, const double* knots, int stride, double* returns, unsigned char* failure,
                         const int* list, const int* count, Writeback wb, bool writeback) {
  static int sms = 0;
  if (!sms) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); if (sms <= 0) sms = 148; }
  const Constants<LeggedXLT>* c = reinterpret_cast<const Constants<LeggedXLT>*>(dconst);
  // one warp per block, spread over the SMs: flagged rollouts are few, and each is latency bound
  if (writeback) sparse_fixup_kernel<LeggedXLT, 32, true><<<sms * 4, 32, 0, st>>>(c, knots, stride, returns, failure, list, count, wb);
  else sparse_fixup_kernel<LeggedXLT, 32, false><<<sms * 4, 32, 0, st>>>(c, knots, stride, returns, failure, list, count, wb);
}
}  // namespace ps
