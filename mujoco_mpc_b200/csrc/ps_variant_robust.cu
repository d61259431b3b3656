// Force-perturbed re-rolls (RobustPlanner, mjpc/planners/robust/robust_planner.cc:108-130): the
// robust_rollout_kernel instantiations of the three dense thread-per-rollout capacity classes.  Their own
// translation unit: batches are candidates x repetitions (small), so nothing here is tuned for throughput, and
// the planning kernels of the other units are compiled exactly as they were before this file existed.
#include "ps_variant.h"
#ifndef PS_TINY_BLOCK
#define PS_TINY_BLOCK 128
#endif
namespace ps {
int launch_robust_tiny(cudaStream_t st, const void* dconst, const double* knots, int stride, int n, double* returns,
                       unsigned char* failure, Writeback wb, XfrcNoise xf) {
  return launch_robust<TinyT, PS_TINY_BLOCK>(st, dconst, knots, stride, n, returns, failure, wb, xf);
}
int launch_robust_planar(cudaStream_t st, const void* dconst, const double* knots, int stride, int n, double* returns,
                         unsigned char* failure, Writeback wb, XfrcNoise xf) {
  return launch_robust<PlanarT, 64>(st, dconst, knots, stride, n, returns, failure, wb, xf);
}
int launch_robust_legged(cudaStream_t st, const void* dconst, const double* knots, int stride, int n, double* returns,
                         unsigned char* failure, Writeback wb, XfrcNoise xf) {
  return launch_robust<LeggedT, 64>(st, dconst, knots, stride, n, returns, failure, wb, xf);
}
// warp-per-rollout coding (walker / quadruped class models on the Euler integrator): the latency engine, which is
// what a batch of a few dozen re-rolls wants
int launch_robust_planar_coop(cudaStream_t st, const void* dconst, const double* knots, int stride, int n, double* returns,
                              unsigned char* failure, Writeback wb, XfrcNoise xf) {
  return launch_coop_robust<PlanarCoopT, 4>(st, dconst, knots, stride, n, returns, failure, wb, xf);
}
int launch_robust_legged_coop(cudaStream_t st, const void* dconst, const double* knots, int stride, int n, double* returns,
                              unsigned char* failure, Writeback wb, XfrcNoise xf) {
  return launch_coop_robust<LeggedCoopT, 4>(st, dconst, knots, stride, n, returns, failure, wb, xf);
}
}  // namespace ps
