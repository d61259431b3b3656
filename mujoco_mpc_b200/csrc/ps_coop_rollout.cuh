// Warp-cooperative rollout: forward pass, Euler integration (implicit joint damping), the
// Trajectory::NoisyRollout time recursion (mjpc/trajectory.cc:100-210,312-326) and, on the device,
// the persistent kernel that hands rollouts to warps.
#pragma once
#include "ps_coop_solver.cuh"
#include "ps_cost.cuh"
#include "ps_rollout.cuh"
#ifdef __CUDACC__
#include "ps_kernels.cuh"
#endif

namespace ps {

// lanes a stage runs on: all of them, or (diagnostics) a single working lane
PS_HD Lanes1 stage_lanes(const Lanes1& L, int, int) { return L; }
#ifdef __CUDACC__
__device__ __forceinline__ Lanes1Block stage_lanes(const Lanes1Block& L, int, int) { return L; }
__device__ __forceinline__ Lanes stage_lanes(const Lanes& L, int debug_serial, int stage) {
  if ((debug_serial >> stage) & 1) { Lanes S; S.lane = L.lane == 0 ? 0 : (1 << 28); S.width = 1; return S; }
  return L;
}
#endif

// `active` = the rollout is still alive.  Under block lockstep (LT::kBlock) a dead rollout keeps walking the stage
// sequence -- doing nothing -- so that the block's barriers stay matched; for every other lane type the caller only gets
// here with active == true or skips the stages altogether.
template <class T, class LT>
PS_FN void co_forward_pass(const LT& L, const DevModel<T>& m, CoopData<T>& d, bool active = true) {
  const int dbg = m.debug_serial;
  if (active) co_position_smooth(stage_lanes(L, dbg, 0), m, d);
  L.block_sync();
  if (active) co_make_constraints(stage_lanes(L, dbg, 1), m, d);
  L.block_sync();
  if (active) {
    co_velocity_stage(stage_lanes(L, dbg, 2), m, d);
    co_reference_acceleration(stage_lanes(L, dbg, 3), m, d);
    if (m.need_subtreevel) co_subtree_velocity(stage_lanes(L, dbg, 4), m, d);
    co_actuation_stage(stage_lanes(L, dbg, 5), m, d);
    co_acceleration_stage(stage_lanes(L, dbg, 6), m, d);
  }
  L.block_sync();
  co_constraint_stage(stage_lanes(L, dbg, 7), m, d, active);
}

// forward pass under Cartesian perturbation forces (robust re-rolls only)
template <class T, class LT>
PS_FN void co_forward_pass_xfrc(const LT& L, const DevModel<T>& m, CoopData<T>& d, const double* xfrc_applied) {
  const int dbg = m.debug_serial;
  co_position_smooth(stage_lanes(L, dbg, 0), m, d);
  co_make_constraints(stage_lanes(L, dbg, 1), m, d);
  co_velocity_stage(stage_lanes(L, dbg, 2), m, d);
  co_reference_acceleration(stage_lanes(L, dbg, 3), m, d);
  if (m.need_subtreevel) co_subtree_velocity(stage_lanes(L, dbg, 4), m, d);
  co_actuation_stage(stage_lanes(L, dbg, 5), m, d);
  co_acceleration_stage_xfrc(stage_lanes(L, dbg, 6), m, d, xfrc_applied);
  co_constraint_stage(stage_lanes(L, dbg, 7), m, d);
}

// Per-rollout perturbation state of the phase-parallel coding (see Perturbation<T, XF> in ps_rollout.cuh): the
// planning kernels use the empty one; the robust kernels hand in storage for xfrc_applied (shared memory for a
// warp, a local array for a single lane) and the Ornstein-Uhlenbeck driving noise.
template <class T, bool XF> struct CoPerturbation;
template <class T> struct CoPerturbation<T, false> { static constexpr bool kActive = false; };   // never touched by the rollout
template <class T> struct CoPerturbation<T, true> {
  static constexpr bool kActive = true;
  double* applied; const XfrcNoise* xf;          // applied: [6 * nbody], visible to every lane of the rollout
  template <class LT> PS_HD void reset(const LT& L, const DevModel<T>& m) const { CO_FOR(k, 6 * m.nbody) applied[k] = 0; L.sync(); }
  template <class LT> PS_HD void advance(const LT& L, const DevModel<T>& m, int t, int stride, int i) const {
    if (!xf || !xf->z) return;
    const int n6 = 6 * m.nbody; const double* zc = xf->z + (size_t)t * n6 * stride + i;
    CO_FOR(k, n6) applied[k] = xf->rate * applied[k] + xf->scale * zc[(size_t)k * stride];
    L.sync();
  }
  template <class LT> PS_HD void forward(const LT& L, const DevModel<T>& m, CoopData<T>& d) const { co_forward_pass_xfrc(L, m, d, applied); }
};

// semi-implicit Euler with implicit joint damping (mj_Euler); other integrators stay with the
// thread-per-rollout engine
template <class T, class LT>
PS_FN void co_integrate(const LT& L, const DevModel<T>& m, CoopData<T>& d) {
  constexpr int S = T::NV; const int nv = m.nv; const double dt = m.timestep;
  double* qacc = d.sv_x;
  if (!m.any_damping || (m.disableflags & DSBL_EULERDAMP)) { CO_FOR(i, nv) qacc[i] = d.qacc[i]; L.sync(); }
  else {
    double* A = d.qH;
    CO_FOR(w, nv * S) A[w] = d.qM[w];
    L.sync();
    CO_FOR(i, nv) { A[i * S + i] += dt * m.dof_damping[i]; qacc[i] = d.qfrc_smooth[i] + d.qfrc_constraint[i]; }
    L.sync();
    co_factor_tree(L, m, A, d.qLDiagInv);
    co_solve_tree(L, m, qacc, A, d.qLDiagInv);
  }
  if (m.na && !(m.disableflags & DSBL_ACTUATION)) {
    CO_FOR(i, m.nu) {
      if (m.act_dyntype[i] == DYN_NONE) continue;
      const int a = m.act_actadr[i];
      double act = d.act[a];
      if (m.act_dyntype[i] == DYN_FILTEREXACT) {
        const double tau = fmax(kMinVal, m.act_dynprm[3 * i]);
        act = act + d.act_dot[a] * tau * (1 - exp(-dt / tau));
      } else act = act + d.act_dot[a] * dt;
      if (m.act_actlimited[i]) act = clampd(act, m.act_actrange[2 * i], m.act_actrange[2 * i + 1]);
      d.act[a] = act;
    }
  }
  CO_FOR(i, nv) d.qvel[i] += qacc[i] * dt;
  L.sync();
  CO_FOR(j, m.njnt) {
    const int qa = m.jnt_qposadr[j], da = m.jnt_dofadr[j], jt = m.jnt_type[j];
    if (jt == JNT_FREE) {
      PS_LOOP2 for (int k = 0; k < 3; k++) d.qpos[qa + k] += dt * d.qvel[da + k];
      quat_integrate(d.qpos + qa + 3, d.qvel + da + 3, dt);
    } else if (jt == JNT_BALL) quat_integrate(d.qpos + qa, d.qvel + da, dt);
    else d.qpos[qa] += dt * d.qvel[da];
  }
  if (L.lane == 0) d.time += dt;
  L.sync();
}

// What a rollout carries from one time step to the next besides its CoopData state: the running cost sum and its fate.
struct CoCarry { double sum; bool failed, dead; };

// co_rollout = co_rollout_begin + co_rollout_steps(0, H) + co_rollout_end.  The split exists for the step-wise launches of
// the thread-per-rollout kernel (sparse_step_kernel), which save / restore the state between launches so that rollouts
// can be re-binned by workload; the arithmetic of a rollout does not depend on how its steps are grouped.
template <class T, bool WB, class LT, class PT = CoPerturbation<T, false>>
PS_FN CoCarry co_rollout_begin(const LT& L, const DevModel<T>& m, const DevPlan<T>& plan, int stride, int i, const Writeback& wb,
                               CoopData<T>& d, const PT pert = PT(), bool live = true) {
  // `live` = false: a padding thread of a block that walks in lockstep (LT::kBlock) -- it takes part in the block's
  // barriers and touches nothing else.  A rollout that fails stops working but, under lockstep, keeps pace the same way;
  // for every other lane type the remaining time steps of a dead rollout are an empty loop.
  const int nq = m.nq, nv = m.nv, na = m.na, nu = m.nu, ns = nq + nv + na;
  if (live) {
    CO_FOR(k, nq) d.qpos[k] = plan.state[k];
    CO_FOR(k, nv) { d.qvel[k] = plan.state[nq + k]; d.qacc_warmstart[k] = plan.warmstart[k]; d.qacc[k] = 0; }
    CO_FOR(k, na) d.act[k] = plan.state[nq + nv + k];
    CO_FOR(k, nu) d.ctrl[k] = 0;
    CO_FOR(k, m.nmocap) { vcopy(d.mocap_pos + 3 * k, plan.mocap + 7 * k, 3); vcopy(d.mocap_quat + 4 * k, plan.mocap + 7 * k + 3, 4); }
    if (L.lane == 0) { d.time = plan.time; d.warning = 0; d.ncon = 0; d.nefc = 0; d.nlimit = 0; d.solver_iter = 0; }
    if (WB) { CO_FOR(k, ns) wb.states[(size_t)k * stride + i] = plan.state[k]; if (L.lane == 0) wb.times[i] = plan.time; }
    L.sync();
    if constexpr (PT::kActive) pert.reset(L, m);
  }
  CoCarry c; c.sum = 0; c.failed = false; c.dead = !live;
  return c;
}

// time steps [t_begin, t_end) of the recursion (t = H - 1 is the terminal forward pass)
template <class T, bool WB, class LT, class PT = CoPerturbation<T, false>>
PS_FN void co_rollout_steps(const LT& L, const DevModel<T>& m, const DevTask& task, const DevPlan<T>& plan, const double* knots,
                            int stride, int i, const Writeback& wb, CoopData<T>& d, CoCarry& carry, int t_begin, int t_end,
                            const PT pert = PT()) {
  const int nq = m.nq, nv = m.nv, na = m.na, nu = m.nu, ns = nq + nv + na, nr = task.num_residual;
  const int H = plan.horizon, P = plan.num_spline_points, ntr = 3 * task.num_trace;
  const double* kcol = knots + i;
  double sum = carry.sum; bool failed = carry.failed, dead = carry.dead;
  PS_LOOP2 for (int t = t_begin; t < t_end; t++) {
    const bool terminal = t == H - 1;
    L.block_sync();
    if (!terminal && !dead) {
      // policy: spline sample + clamp, one lane per actuator
      const double time = d.time;
      CO_FOR(a, nu) {
        auto kn = [&](int p, int) { return kcol[(size_t)(p * nu + a) * stride]; };
        double u; spline_sample(plan.spline_times, P, 1, plan.interpolation, time, kn, &u);
        d.ctrl[a] = clampd(u, m.act_ctrlrange[2 * a], m.act_ctrlrange[2 * a + 1]);
      }
      if constexpr (PT::kActive) pert.advance(L, m, t, stride, i);
      bool bad = false;
      CO_FOR(k, nq) bad = bad || bad_number(d.qpos[k]);
      CO_FOR(k, nv) bad = bad || bad_number(d.qvel[k]);
      L.sync();
      if (L.any(bad)) { failed = true; dead = true; if (WB) { CO_FOR(a, nu) wb.actions[((size_t)t * nu + a) * stride + i] = d.ctrl[a]; } }
    }
    if (!LT::kBlock && dead) break;
    if (WB && !dead && (!terminal || H > 1)) CO_FOR(a, nu) wb.actions[((size_t)t * nu + a) * stride + i] = d.ctrl[a];
    // what an mjData stepped H - 1 times carries into its next mj_step (the terminal forward pass below is the planner's, not the plant's)
    if (WB && terminal && !dead) CO_FOR(k, nv) wb.warmstart[(size_t)k * stride + i] = d.qacc_warmstart[k];
    if constexpr (PT::kActive) { if (!dead) pert.forward(L, m, d); } else co_forward_pass(L, m, d, !dead);
    if (!dead) {
      if (L.lane == 0) {
        task_residual(m, d, task, d.residual);
        d.sc_cost = cost_value(task, d.residual);
      }
      L.sync();
      const double c = d.sc_cost;
      if (WB) {
        CO_FOR(k, nr) wb.residual[((size_t)t * nr + k) * stride + i] = d.residual[k];
        if (L.lane == 0) {
          wb.costs[(size_t)t * stride + i] = c;
          PS_LOOP2 for (int k = 0; k < task.num_trace; k++) {
            double tr[3]; read_sensor(m, d, task.trace_sensor[k], tr);
            PS_LOOP2 for (int x = 0; x < 3; x++) wb.trace[((size_t)t * ntr + 3 * k + x) * stride + i] = tr[x];
          }
        }
      }
      if (terminal) sum += c;
      else {
        bool badacc = false;
        CO_FOR(k, nv) badacc = badacc || bad_number(d.qacc[k]);
        if (L.any(badacc) || (d.warning & (WARN_BADQPOS | WARN_BADQVEL | WARN_BADQACC | WARN_BADCTRL))) { failed = true; dead = true; }
        else {
          sum += c;
          co_integrate(stage_lanes(L, m.debug_serial, 8), m, d);
          if (WB) {
            double* s = wb.states + ((size_t)(t + 1) * ns) * stride + i;
            CO_FOR(k, nq) s[(size_t)k * stride] = d.qpos[k];
            CO_FOR(k, nv) s[(size_t)(nq + k) * stride] = d.qvel[k];
            CO_FOR(k, na) s[(size_t)(nq + nv + k) * stride] = d.act[k];
            if (L.lane == 0) wb.times[(size_t)(t + 1) * stride + i] = d.time;
          }
        }
      }
    }
    if (!LT::kBlock && dead) break;
  }
  carry.sum = sum; carry.failed = failed; carry.dead = dead;
}

template <class T, class LT>
PS_FN double co_rollout_end(const LT& L, const DevPlan<T>& plan, const CoopData<T>& d, const CoCarry& carry, bool live, int* flags) {
  const int H = plan.horizon;
  const double total = carry.failed ? 1.0e6 : carry.sum / (H > 1 ? H : 1);
  *flags = !live ? 0 : ((carry.failed ? 1 : 0) | ((d.warning & WARN_CAPACITY) ? 2 : 0));
  L.sync();
  return total;
}

// knots: SoA [P*nu][stride]; i = this rollout's column.  flags: bit0 failure, bit1 capacity overflow
template <class T, bool WB, class LT, class PT = CoPerturbation<T, false>>
PS_FN double co_rollout(const LT& L, const DevModel<T>& m, const DevTask& task, const DevPlan<T>& plan, const double* knots,
                        int stride, int i, const Writeback& wb, CoopData<T>& d, int* flags, const PT pert = PT(), bool live = true) {
  CoCarry carry = co_rollout_begin<T, WB>(L, m, plan, stride, i, wb, d, pert, live);
  co_rollout_steps<T, WB>(L, m, task, plan, knots, stride, i, wb, d, carry, 0, plan.horizon, pert);
  return co_rollout_end(L, plan, d, carry, live, flags);
}

#ifdef __CUDACC__
// ---------------------------------------------------------------- persistent kernel: one warp per rollout
// Each warp owns a CoopData slot in dynamic shared memory and pulls rollout indices from a global
// counter until the batch is exhausted, so long (contact-rich) and short rollouts balance across
// the grid.  Constants arrive once per block by TMA bulk copy.
template <class T, int WARPS, bool WB>
__global__ void __launch_bounds__(WARPS * 32)
coop_rollout_kernel(const Constants<T>* __restrict__ gconst, const double* __restrict__ knots, int stride, int first, int n,
                    long long sample_offset, double* __restrict__ returns, unsigned char* __restrict__ failure,
                    MinLoc* __restrict__ block_min, Writeback wb, int* __restrict__ counter, int active_warps) {
  // active_warps: how many of the block's WARPS pull rollouts (launch_coop).  A batch smaller than WARPS x the SM count is
  // spread over the SMs instead of packed WARPS to an SM: a rollout alone on its SM runs about twice as fast as one that
  // shares the SM's instruction cache and issue slots with three others at other places in the code (ncu, N = 128:
  // `no_instruction` 1.8 cycles per issue on 32 of 148 SMs).  The other warps go straight to the closing barrier.
  extern __shared__ __align__(128) unsigned char smem[];
  Constants<T>& sc = *reinterpret_cast<Constants<T>*>(smem);
  constexpr size_t kConst = (sizeof(Constants<T>) + 127) / 128 * 128;
  constexpr size_t kData = (sizeof(CoopData<T>) + 127) / 128 * 128;
  __shared__ alignas(8) uint64_t bar;
  __shared__ MinLoc warp_min[WARPS];

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

  const int warp = threadIdx.x >> 5;
  Lanes L; L.lane = threadIdx.x & 31; L.width = 32;
  CoopData<T>& d = *reinterpret_cast<CoopData<T>*>(smem + kConst + warp * kData);
  double best = 1.0e300; long long best_i = 0x7fffffffffffffffLL;
  if (warp < active_warps) PS_LOOP2 for (;;) {
    int k = 0;
    if (L.lane == 0) k = atomicAdd(counter, 1);
    k = __shfl_sync(0xffffffffu, k, 0);
    if (k >= n) break;
    const int i = first + k;
    int flags = 0;
    double total = co_rollout<T, WB>(L, sc.m, sc.t, sc.p, knots, stride, i, wb, d, &flags);
    if (L.lane == 0) { returns[i] = total; failure[i] = (unsigned char)flags; }
    if (!(total == total)) total = 1.0e300;              // NaN never wins
    if (better(total, sample_offset + i, best, best_i)) { best = total; best_i = sample_offset + i; }
  }
  if (L.lane == 0) { warp_min[warp].cost = best; warp_min[warp].index = best_i; }
  __syncthreads();
  if (threadIdx.x == 0) {
    PS_LOOP2 for (int w = 1; w < WARPS; w++) if (better(warp_min[w].cost, warp_min[w].index, best, best_i)) { best = warp_min[w].cost; best_i = warp_min[w].index; }
    block_min[blockIdx.x].cost = best; block_min[blockIdx.x].index = best_i;
  }
}

// ---------------------------------------------------------------- robust re-rolls, one warp per rollout
// RobustPlanner's candidates x repetitions under force perturbations (robust_planner.cc:108-130): the batch is
// small, so each rollout gets a warp and its working set (+ xfrc_applied) in shared memory; always full writeback.
template <class T, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
coop_robust_kernel(const Constants<T>* __restrict__ gconst, const double* __restrict__ knots, int stride, int n,
                   double* __restrict__ returns, unsigned char* __restrict__ failure, Writeback wb, XfrcNoise xf) {
  extern __shared__ __align__(128) unsigned char smem[];
  Constants<T>& sc = *reinterpret_cast<Constants<T>*>(smem);
  constexpr size_t kConst = (sizeof(Constants<T>) + 127) / 128 * 128;
  constexpr size_t kData = (sizeof(CoopData<T>) + 127) / 128 * 128;
  constexpr size_t kXf = (sizeof(double) * 6 * T::NBODY + 127) / 128 * 128;
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
  const int warp = threadIdx.x >> 5;
  const int i = blockIdx.x * WARPS + warp;
  if (i >= n) return;                                  // whole warps leave together; no block-level sync below
  Lanes L; L.lane = threadIdx.x & 31; L.width = 32;
  CoopData<T>& d = *reinterpret_cast<CoopData<T>*>(smem + kConst + warp * (kData + kXf));
  CoPerturbation<T, true> pert{reinterpret_cast<double*>(smem + kConst + warp * (kData + kXf) + kData), &xf};
  int flags = 0;
  const double total = co_rollout<T, true, Lanes, CoPerturbation<T, true>>(L, sc.m, sc.t, sc.p, knots, stride, i, wb, d, &flags, pert);
  if (L.lane == 0) { returns[i] = total; failure[i] = (unsigned char)flags; }
}

// ---------------------------------------------------------------- thread-per-rollout flavour of the same phases
// One THREAD per rollout with a single working lane (Lanes1): the compressed-Jacobian algorithms of
// the cooperative engine, SIMD across rollouts.  Its working set sits in local memory, so it needs
// tens of thousands of rollouts in flight to hide that latency -- the large-batch engine, where the
// warp-per-rollout kernel above is the small-batch (latency) engine.
#ifndef PS_SPARSE_LOCKSTEP
#define PS_SPARSE_LOCKSTEP 1
#endif
#ifndef PS_SPARSE_RESIDENT
#define PS_SPARSE_RESIDENT 512          // threads the register cap keeps resident per SM
#endif
template <class T, int BLOCK, bool WB>
__global__ void __launch_bounds__(BLOCK, PS_SPARSE_RESIDENT / BLOCK)
sparse_rollout_kernel(const Constants<T>* __restrict__ gconst, const double* __restrict__ knots, int stride, int first, int n,
                      long long sample_offset, double* __restrict__ returns, unsigned char* __restrict__ failure,
                      MinLoc* __restrict__ block_min, Writeback wb) {
  __shared__ Constants<T> sc;
  __shared__ alignas(8) uint64_t bar;
  __shared__ MinLoc warp_min[BLOCK / 32];
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
  // grid-stride over the batch: the launch may use fewer blocks than the batch needs (launch_sparse: a bounded number of
  // resident rollouts per SM keeps their thread-private working sets inside L1 / L2), each block then rolls several groups
  double c = 1.0e300; long long gi = 0x7fffffffffffffffLL;
  for (int base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {      // block-uniform trip count
    const int i = first + base + threadIdx.x;
    const bool live = i < first + n;
    int flags = 0;
    CoopData<T> d;
#if PS_SPARSE_LOCKSTEP
    Lanes1Block L;         // the block walks steps / stages / Newton iterations together (ps_coop.cuh); padding threads idle along
    double total = co_rollout<T, WB, Lanes1Block>(L, sc.m, sc.t, sc.p, knots, stride, live ? i : first, wb, d, &flags, CoPerturbation<T, false>(), live);
#else
    Lanes1 L;
    double total = live ? co_rollout<T, WB>(L, sc.m, sc.t, sc.p, knots, stride, i, wb, d, &flags) : 0.0;
#endif
    if (live) {
      returns[i] = total;
      failure[i] = (unsigned char)flags;
      if (!(total == total)) total = 1.0e300;
      if (better(total, sample_offset + i, c, gi)) { c = total; gi = sample_offset + i; }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_down_sync(0xffffffffu, c, o);
    const long long oi = __shfl_down_sync(0xffffffffu, gi, o);
    if (better(oc, oi, c, gi)) { c = oc; gi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { warp_min[threadIdx.x >> 5].cost = c; warp_min[threadIdx.x >> 5].index = gi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); w++) if (better(warp_min[w].cost, warp_min[w].index, c, gi)) { c = warp_min[w].cost; gi = warp_min[w].index; }
    block_min[blockIdx.x].cost = c; block_min[blockIdx.x].index = gi;
  }
}

// ---------------------------------------------------------------- capacity-overflow fix-up
// A rollout that ran out of the scoring kernel's contact / constraint-row capacity carries flag bit 1 and was scored with the
// rows that fit.  The reference never truncates (mjpc/trajectory.cc:169-173: a rollout is valid or 1e6), so every flagged
// sample is rolled again here at capacities (TXL) large enough for a fallen robot, and its return / flags / trajectory
// replace the truncated ones before the argmin.  Flagged samples are few and each one is latency bound, so this is the
// warp-per-rollout coding (working set in shared memory, ~0.8 ms per step instead of the ~25 ms a lone thread needs):
// persistent blocks pull entries of `list` (collect_flagged_kernel) through `counter` until `*count` are done; with
// count == 0 -- the usual case -- every block exits before touching shared memory.
template <class TXL, int WARPS, bool WB>
__global__ void __launch_bounds__(WARPS * 32)
coop_fixup_kernel(const Constants<TXL>* __restrict__ gconst, const double* __restrict__ knots, int stride,
                  double* __restrict__ returns, unsigned char* __restrict__ failure, const int* __restrict__ list,
                  const int* __restrict__ count, int* __restrict__ counter, Writeback wb) {
  const int n = *count;
  if ((int)(blockIdx.x * WARPS) >= n) return;                // nothing for this block (uniform: before any barrier)
  extern __shared__ __align__(128) unsigned char smem[];
  Constants<TXL>& sc = *reinterpret_cast<Constants<TXL>*>(smem);
  constexpr size_t kConst = (sizeof(Constants<TXL>) + 127) / 128 * 128;
  constexpr size_t kData = (sizeof(CoopData<TXL>) + 127) / 128 * 128;
  __shared__ alignas(8) uint64_t bar;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  if (threadIdx.x == 0) {
    constexpr uint32_t total = (uint32_t)sizeof(Constants<TXL>);
    mbar_expect_tx(&bar, total);
    const char* src = reinterpret_cast<const char*>(gconst); char* dst = reinterpret_cast<char*>(&sc);
    for (uint32_t off = 0; off < total; off += 32768u) {
      const uint32_t sz = total - off < 32768u ? total - off : 32768u;
      tma_bulk_g2s(dst + off, src + off, sz, &bar);
    }
  }
  mbar_wait(&bar, 0);
  const int warp = threadIdx.x >> 5;
  Lanes L; L.lane = threadIdx.x & 31; L.width = 32;
  CoopData<TXL>& d = *reinterpret_cast<CoopData<TXL>*>(smem + kConst + warp * kData);
  for (;;) {
    int k = 0;
    if (L.lane == 0) k = atomicAdd(counter, 1);
    k = __shfl_sync(0xffffffffu, k, 0);
    if (k >= n) break;
    const int i = list[k];
    int flags = 0;
    const double total = co_rollout<TXL, WB>(L, sc.m, sc.t, sc.p, knots, stride, i, wb, d, &flags);
    if (L.lane == 0) { returns[i] = total; failure[i] = (unsigned char)flags; }
  }
}
#endif

}  // namespace ps
