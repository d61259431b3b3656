// Allocation-free variant of the winner selection (K2's block-level min-loc + K3's grid-level min-loc, the argmin
// of SamplingPlanner::OptimizePolicyCandidates).  No counterpart in the reference: added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
// that forbid it).  The engine's own path keeps its per-block partial results in a cudaMalloc'ed buffer
// (PsEngine_::d_blockmin); here that scratch is a file-scope __device__ array, so the entry point performs no
// cudaMalloc / cudaFree / allocating library call.  Every GPU that loads the library holds the array
// (kScratchBytes below) until the process exits.
//
// The array's name is only a host-side stand-in in host code: its device address is fetched once per device with
// cudaGetSymbolAddress (that device current) and handed to the kernel.
#include <cstdio>
#include <mutex>
#include <string>
#include <cuda_runtime.h>
#include "../../include/ps_engine.h"

namespace {

struct MinLocRec { double cost; long long index; };

constexpr int kThreads = 256, kPerThread = 4, kPerBlock = kThreads * kPerThread;       // returns reduced by one block
constexpr int kMaxBlocks = PS_NOALLOC_MAX_RETURNS / kPerBlock;                           // 1024 partial results
constexpr int kMaxDevices = 64;
constexpr long long kNoIndex = 0x7fffffffffffffffLL;

// scratch: [0, 256) bytes hold the blocks-done counter (must be zero when a launch starts), then one partial
// result per block.  16 KB + 256 B per device.
struct alignas(256) Scratch { unsigned int done; unsigned int pad[63]; MinLocRec part[kMaxBlocks]; };
static __device__ __align__(256) unsigned char g_scratch[sizeof(Scratch)];
constexpr size_t kScratchBytes = sizeof(Scratch);

__device__ __forceinline__ bool better(double c, long long i, double bc, long long bi) { return (c < bc) || (c == bc && i < bi); }

__device__ __forceinline__ void warp_min(double& c, long long& gi) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_down_sync(0xffffffffu, c, o);
    const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
    if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
  }
}

// block-level (return, global index) min of this block's slice -> scratch; the block that finishes last reduces
// the partial results and writes out = [min return, global index].  Same order as the engine's K2 + K3: NaN never
// wins, ties go to the lowest index, an all-NaN batch selects the first sample.
__global__ void __launch_bounds__(kThreads)
argmin_noalloc_kernel(const double* __restrict__ returns, int n, long long sample_offset, Scratch* __restrict__ s, double* __restrict__ out) {
  __shared__ MinLocRec wm[kThreads / 32];
  __shared__ bool last;
  double c = 1.0e300; long long gi = kNoIndex;
  const int base = blockIdx.x * kPerBlock;
  for (int k = 0; k < kPerThread; k++) {
    const int i = base + k * kThreads + threadIdx.x;
    if (i < n) {
      double v = returns[i];
      if (!(v == v)) v = 1.0e300;
      if (better(v, sample_offset + i, c, gi)) { c = v; gi = sample_offset + i; }
    }
  }
  warp_min(c, gi);
  if ((threadIdx.x & 31) == 0) { wm[threadIdx.x >> 5].cost = c; wm[threadIdx.x >> 5].index = gi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    c = threadIdx.x < kThreads / 32 ? wm[threadIdx.x].cost : 1.0e300;
    gi = threadIdx.x < kThreads / 32 ? wm[threadIdx.x].index : kNoIndex;
    warp_min(c, gi);
    if (threadIdx.x == 0) {
      s->part[blockIdx.x].cost = c; s->part[blockIdx.x].index = gi;
      __threadfence();
      last = atomicAdd(&s->done, 1u) == gridDim.x - 1;
    }
  }
  __syncthreads();
  if (!last) return;
  __threadfence();
  c = 1.0e300; gi = kNoIndex;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += kThreads) {
    const MinLocRec v = s->part[b];
    if (better(v.cost, v.index, c, gi)) { c = v.cost; gi = v.index; }
  }
  warp_min(c, gi);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) { wm[threadIdx.x >> 5].cost = c; wm[threadIdx.x >> 5].index = gi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    c = threadIdx.x < kThreads / 32 ? wm[threadIdx.x].cost : 1.0e300;
    gi = threadIdx.x < kThreads / 32 ? wm[threadIdx.x].index : kNoIndex;
    warp_min(c, gi);
    if (threadIdx.x == 0) {
      if (gi == kNoIndex) gi = sample_offset;
      out[0] = c; out[1] = (double)gi;
      s->done = 0;                                  // leave the counter ready for the next launch
    }
  }
}

// device address of g_scratch, resolved once per device
std::mutex g_mu;
Scratch* g_ptr[kMaxDevices] = {nullptr};

}  // namespace

// ps_engine.cu owns ps_last_error(); this unit reports through the same text
extern "C" void ps_set_last_error_(const char* s);

static int fail_na(const std::string& s) { ps_set_last_error_(s.c_str()); return -1; }

extern "C" int ps_argmin_noalloc(int device_ordinal, void* stream, const double* d_returns, int n, long long sample_offset, double* d_out) {
  if (!d_returns || !d_out) return fail_na("ps_argmin_noalloc: null device pointer");
  if (n < 1) return fail_na("ps_argmin_noalloc: n must be positive");
  if (n > PS_NOALLOC_MAX_RETURNS)
    return fail_na("ps_argmin_noalloc: n = " + std::to_string(n) + " exceeds PS_NOALLOC_MAX_RETURNS = " + std::to_string(PS_NOALLOC_MAX_RETURNS) +
                   " (the static scratch holds one partial result per 1024 returns)");
  if (device_ordinal < 0 || device_ordinal >= kMaxDevices) return fail_na("ps_argmin_noalloc: device ordinal out of range");
  cudaError_t err = cudaSetDevice(device_ordinal);
  if (err != cudaSuccess) return fail_na(std::string("ps_argmin_noalloc: cudaSetDevice: ") + cudaGetErrorString(err));
  Scratch* s = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ptr[device_ordinal]) {
      void* p = nullptr;
      err = cudaGetSymbolAddress(&p, g_scratch);                  // with device_ordinal current: that device's copy
      if (err != cudaSuccess || !p) return fail_na(std::string("ps_argmin_noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(err));
      g_ptr[device_ordinal] = reinterpret_cast<Scratch*>(p);
    }
    s = g_ptr[device_ordinal];
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // the array keeps the last call's contents: the counter must start at zero even if an earlier launch was cut short
  err = cudaMemsetAsync(s, 0, 256, st);
  if (err != cudaSuccess) return fail_na(std::string("ps_argmin_noalloc: cudaMemsetAsync: ") + cudaGetErrorString(err));
  const int grid = (n + kPerBlock - 1) / kPerBlock;
  argmin_noalloc_kernel<<<grid, kThreads, 0, st>>>(d_returns, n, sample_offset, s, d_out);
  err = cudaGetLastError();
  if (err != cudaSuccess) return fail_na(std::string("ps_argmin_noalloc: launch: ") + cudaGetErrorString(err));
  return 0;
}

extern "C" int ps_noalloc_scratch_bytes(void) { return (int)kScratchBytes; }
