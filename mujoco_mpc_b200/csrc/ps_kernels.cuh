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
#include "ps_cost.cuh"

namespace ps {

template <class T>
struct alignas(128) Constants { DevModel<T> m; DevTask t; DevPlan<T> p; };

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
                               const unsigned char* __restrict__ std2_injected, double* __restrict__ knots, int stride) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int e = blockIdx.y;
  if (i >= n) return;
  const long long gi = sample_offset + i;          // global sample index (rank-independent noise)
  const int a = e % nu;
  const double lo = ctrlrange[2 * a], hi = ctrlrange[2 * a + 1];
  double v = nominal[e];
  if (gi != 0) {
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
    const double scale = 0.5 * (hi - lo);
    v += (scale * sd) * z;
    v = clampd(v, lo, hi);
  }
  knots[(size_t)e * stride + i] = v;
}

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

// optional full-trajectory outputs, SoA [t][field][stride]
struct Writeback {
  double* states; double* actions; double* residual; double* costs; double* times; double* trace;
};

// ---------------------------------------------------------------- K2
template <class T, int BLOCK, bool WB>
__global__ void __launch_bounds__(BLOCK)
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

  const DevModel<T>& m = sc.m; const DevTask& task = sc.t; const DevPlan<T>& plan = sc.p;
  const int i = first + blockIdx.x * BLOCK + threadIdx.x;       // local sample index
  const bool active = i < first + n;
  double total = 1.0e300; bool failed = false;

  if (active) {
    RolloutData<T> d;
    const int nq = m.nq, nv = m.nv, na = m.na, nu = m.nu, ns = nq + nv + na, nr = task.num_residual;
    const int H = plan.horizon, P = plan.num_spline_points;
    for (int k = 0; k < nq; k++) d.qpos[k] = plan.state[k];
    for (int k = 0; k < nv; k++) { d.qvel[k] = plan.state[nq + k]; d.qacc_warmstart[k] = 0; d.qacc[k] = 0; }
    for (int k = 0; k < na; k++) d.act[k] = plan.state[nq + nv + k];
    for (int k = 0; k < m.nmocap; k++) { vcopy(d.mocap_pos + 3 * k, plan.mocap + 7 * k, 3); vcopy(d.mocap_quat + 4 * k, plan.mocap + 7 * k + 3, 4); }
    d.time = plan.time; d.warning = 0; d.ncon = 0; d.nefc = 0;
    const double* kcol = knots + i;
    auto kn = [&](int p, int a) { return kcol[(size_t)(p * nu + a) * stride]; };
    double residual[kMaxResidual];
    double sum = 0;
    if (WB) { for (int k = 0; k < ns; k++) wb.states[(size_t)k * stride + i] = plan.state[k]; wb.times[i] = d.time; }

    for (int t = 0; t < H - 1 && !failed; t++) {
      spline_sample(plan.spline_times, P, nu, plan.interpolation, d.time, kn, d.ctrl);
      for (int a = 0; a < nu; a++) d.ctrl[a] = clampd(d.ctrl[a], m.act_ctrlrange[2 * a], m.act_ctrlrange[2 * a + 1]);
      if (WB) for (int a = 0; a < nu; a++) wb.actions[((size_t)t * nu + a) * stride + i] = d.ctrl[a];
      if (state_is_bad(m, d)) { failed = true; break; }
      forward_pass(m, d);
      task_residual(m, d, task, residual);
      const double c = cost_value(task, residual);
      if (WB) {
        for (int k = 0; k < nr; k++) wb.residual[((size_t)t * nr + k) * stride + i] = residual[k];
        wb.costs[(size_t)t * stride + i] = c;
        for (int k = 0; k < task.num_trace; k++) {
          double tr[3]; read_sensor(m, d, task.trace_sensor[k], tr);
          for (int x = 0; x < 3; x++) wb.trace[((size_t)t * 3 * task.num_trace + 3 * k + x) * stride + i] = tr[x];
        }
      }
      if (acc_is_bad(m, d) || (d.warning & (WARN_BADQPOS | WARN_BADQVEL | WARN_BADQACC | WARN_BADCTRL))) { failed = true; break; }
      sum += c;
      integrate(m, d);
      if (WB) {
        double* s = wb.states + ((size_t)(t + 1) * ns) * stride + i;
        for (int k = 0; k < nq; k++) s[(size_t)k * stride] = d.qpos[k];
        for (int k = 0; k < nv; k++) s[(size_t)(nq + k) * stride] = d.qvel[k];
        for (int k = 0; k < na; k++) s[(size_t)(nq + nv + k) * stride] = d.act[k];
        wb.times[(size_t)(t + 1) * stride + i] = d.time;
      }
    }
    if (!failed) {
      // terminal: ctrl is NOT rewritten (trajectory.cc:190-198): r(x_{H-1}, u_{H-2})
      if (WB && H > 1) for (int a = 0; a < nu; a++) wb.actions[((size_t)(H - 1) * nu + a) * stride + i] = d.ctrl[a];
      if (H <= 1) for (int a = 0; a < nu; a++) d.ctrl[a] = 0;
      forward_pass(m, d);
      task_residual(m, d, task, residual);
      const double c = cost_value(task, residual);
      if (WB) {
        for (int k = 0; k < nr; k++) wb.residual[((size_t)(H - 1) * nr + k) * stride + i] = residual[k];
        wb.costs[(size_t)(H - 1) * stride + i] = c;
        for (int k = 0; k < task.num_trace; k++) {
          double tr[3]; read_sensor(m, d, task.trace_sensor[k], tr);
          for (int x = 0; x < 3; x++) wb.trace[((size_t)(H - 1) * 3 * task.num_trace + 3 * k + x) * stride + i] = tr[x];
        }
      }
      sum += c;
      total = sum / (H > 1 ? H : 1);
    } else total = 1.0e6;
    returns[i] = total;
    failure[i] = (unsigned char)((failed ? 1 : 0) | ((d.warning & WARN_CAPACITY) ? 2 : 0));
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
    c = threadIdx.x < BLOCK / 32 ? warp_min[threadIdx.x].cost : 1.0e300;
    gi = threadIdx.x < BLOCK / 32 ? warp_min[threadIdx.x].index : 0x7fffffffffffffffLL;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double oc = __shfl_down_sync(0xffffffffu, c, o);
      const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
      if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
    }
    if (threadIdx.x == 0) { block_min[blockIdx.x].cost = c; block_min[blockIdx.x].index = gi; }
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

}  // namespace ps
