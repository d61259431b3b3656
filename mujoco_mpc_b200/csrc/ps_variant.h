// One engine instantiation ("variant") = the rollout kernel + host staging for one Traits capacity
// class.  Each variant is its own translation unit so the three kernels compile in parallel.
#pragma once
#include "ps_kernels.cuh"
#include "ps_host_model.h"

namespace ps {

template <class T, int BLOCK>
inline void launch_rollout(cudaStream_t st, const void* dconst, const double* knots, int stride, int first, int n,
                           long long offset, double* returns, unsigned char* failure, MinLoc* block_min, Writeback wb, bool writeback) {
  const int grid = (n + BLOCK - 1) / BLOCK;
  const Constants<T>* c = reinterpret_cast<const Constants<T>*>(dconst);
  if (writeback) rollout_kernel<T, BLOCK, true><<<grid, BLOCK, 0, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb);
  else rollout_kernel<T, BLOCK, false><<<grid, BLOCK, 0, st>>>(c, knots, stride, first, n, offset, returns, failure, block_min, wb);
}

struct Variant {
  const char* name; size_t const_bytes; int block;
  int (*fill)(const PsModel*, const PairList&, void*, std::string*);
  void (*options)(void*, double, int);
  void (*task)(void*, const DevTask&, int);
  int (*plan)(void*, const double*, int, const double*, int, double, const double*, int, int, int);
  void (*launch)(cudaStream_t, const void*, const double*, int, int, int, long long, double*, unsigned char*, MinLoc*, Writeback, bool);
};
template <class T, int BLOCK>
inline Variant make_variant(const char* name) {
  return Variant{name, sizeof(Constants<T>), BLOCK, fill_model<T>, set_options<T>, set_task<T>, set_plan<T>, launch_rollout<T, BLOCK>};
}
const Variant* variant_tiny();
const Variant* variant_planar();
const Variant* variant_legged();

}  // namespace ps
