// The time recursion of ONE candidate rollout: policy -> forward pass -> residual -> cost ->
// integrate, H-1 times, then the terminal forward pass.  Mirrors Trajectory::NoisyRollout +
// UpdateReturn (mjpc/trajectory.cc:100-210,312-326): residual[t] = r(x_t, u_t), states[t+1] is the
// post-integration state, the terminal residual is r(x_{H-1}, u_{H-2}), the return is the mean of
// the per-step costs, and a diverged rollout scores 1e6 (trajectory.cc:169-173).
#pragma once
#include "ps_cost.cuh"

namespace ps {

template <class T>
struct alignas(128) Constants { DevModel<T> m; DevTask t; DevPlan<T> p; };

// optional full-trajectory outputs, SoA [t][field][stride] with the sample index fastest
struct Writeback {
  double* states; double* actions; double* residual; double* costs; double* times; double* trace;
  double* warmstart;      // [nv][stride]: qacc_warmstart after the last integrated step (before the terminal forward pass)
};

// Force perturbation of the robust re-rolls: standard-normal tensor [(t * 6 nbody + k)][stride] and the discrete
// Ornstein-Uhlenbeck coefficients rate = exp(-dt / xfrc_rate), scale = xfrc_std sqrt(1 - rate^2) (trajectory.cc:149-151)
struct XfrcNoise { const double* z; double rate, scale; };

#ifndef PS_ROLLOUT_SINGLE_SITE
#define PS_ROLLOUT_SINGLE_SITE 0
#endif

// Per-rollout perturbation state.  The planning kernels instantiate the empty one (their forward pass is the plain
// forward_pass they were verified with); robust_rollout_kernel instantiates the one that carries
// mjData.xfrc_applied ([force; torque] per body) through the Ornstein-Uhlenbeck recursion of trajectory.cc:147-155.
template <class T, bool XF> struct Perturbation;
template <class T> struct Perturbation<T, false> {
  PS_HD Perturbation(const DevModel<T>&, const XfrcNoise*) {}
  PS_HD void advance(const DevModel<T>&, int, int, int) {}
  PS_HD void forward(const DevModel<T>& m, RolloutData<T>& d) { forward_pass(m, d); }
};
template <class T> struct Perturbation<T, true> {
  const XfrcNoise* xf; double applied[6 * T::NBODY];
  PS_HD Perturbation(const DevModel<T>& m, const XfrcNoise* x) : xf(x) { for (int k = 0; k < 6 * m.nbody; k++) applied[k] = 0; }
  PS_HD void advance(const DevModel<T>& m, int t, int stride, int i) {
    if (!xf || !xf->z) return;
    const int n6 = 6 * m.nbody; const double* zc = xf->z + (size_t)t * n6 * stride + i;
    for (int k = 0; k < n6; k++) applied[k] = xf->rate * applied[k] + xf->scale * zc[(size_t)k * stride];
  }
  PS_HD void forward(const DevModel<T>& m, RolloutData<T>& d) { forward_pass_xfrc(m, d, applied); }
};

// knots: SoA [P*nu][stride]; i = this rollout's column.  flags: bit0 failure, bit1 capacity overflow
// XF: apply Cartesian perturbation forces (own kernel instantiation: robust_rollout_kernel)
template <class T, bool WB, bool XF = false, class TK = DevTask>
PS_FN double rollout_sample(const DevModel<T>& m, const TK& task, const DevPlan<T>& plan, const double* knots,
                            int stride, int i, const Writeback& wb, int* flags, const XfrcNoise* xf = nullptr) {
  typename RolloutData<T>::Rows rows; RolloutData<T> d; d.bind(rows);
  const int nq = nq_of(m), nv = nv_of(m), na = m.na, nu = m.nu, ns = nq + nv + na, nr = task.num_residual;
  const int H = plan.horizon, P = plan.num_spline_points, ntr = 3 * task.num_trace;
  PS_SPEC_UNROLL
  for (int k = 0; k < nq; k++) d.qpos[k] = plan.state[k];
  PS_SPEC_UNROLL
  for (int k = 0; k < nv; k++) { d.qvel[k] = plan.state[nq + k]; d.qacc_warmstart[k] = plan.warmstart[k]; d.qacc[k] = 0; }
  PS_SPEC_UNROLL
  for (int k = 0; k < na; k++) d.act[k] = plan.state[nq + nv + k];
  PS_SPEC_UNROLL
  for (int k = 0; k < nu; k++) d.ctrl[k] = 0;
  PS_SPEC_UNROLL
  for (int k = 0; k < m.nmocap; k++) { vcopy(d.mocap_pos + 3 * k, plan.mocap + 7 * k, 3); vcopy(d.mocap_quat + 4 * k, plan.mocap + 7 * k + 3, 4); }
  d.time = plan.time; d.warning = 0; d.ncon = 0; d.nefc = 0; d.solver_iter = 0;
  const double* kcol = knots + i;
  auto kn = [&](int p, int a) { return kcol[(size_t)(p * nu + a) * stride]; };
  double residual[kMaxResidual];
  double sum = 0; bool failed = false;
  if (WB) { PS_SPEC_UNROLL for (int k = 0; k < ns; k++) wb.states[(size_t)k * stride + i] = plan.state[k]; wb.times[i] = d.time; }
  Perturbation<T, XF> pert(m, xf);

#if PS_ROLLOUT_SINGLE_SITE
  // Experiment (off by default): ONE call site of the forward pass / residual / cost for the running steps and the
  // terminal step, instead of a loop plus a terminal copy.  Where the stage functions are inlined (tiny, planar)
  // the terminal copy duplicates the whole step in the instruction stream, and the tiny kernel is bound by
  // instruction fetch.  Same arithmetic in the same order.
  for (int t = 0; t < H; t++) {
    const bool terminal = t == H - 1;
    if (!terminal) {
      spline_sample(plan.spline_times, P, nu, plan.interpolation, d.time, kn, d.ctrl);
      PS_SPEC_UNROLL
      for (int a = 0; a < nu; a++) d.ctrl[a] = clampd(d.ctrl[a], m.act_ctrlrange[2 * a], m.act_ctrlrange[2 * a + 1]);
      pert.advance(m, t, stride, i);
    }
    // terminal: ctrl is NOT rewritten before the final forward pass (trajectory.cc:190-198)
    if (WB && (!terminal || H > 1)) PS_SPEC_UNROLL for (int a = 0; a < nu; a++) wb.actions[((size_t)t * nu + a) * stride + i] = d.ctrl[a];
    if (!terminal && state_is_bad(m, d)) { failed = true; break; }
    // what an mjData stepped H - 1 times carries into its next mj_step (the terminal forward pass below is the planner's, not the plant's)
    if (WB && terminal) { PS_SPEC_UNROLL for (int k = 0; k < nv; k++) wb.warmstart[(size_t)k * stride + i] = d.qacc_warmstart[k]; }
    pert.forward(m, d);
    task_residual(m, d, task, residual);
    const double c = cost_value(task, residual);
    if (WB) {
      PS_SPEC_UNROLL
      for (int k = 0; k < nr; k++) wb.residual[((size_t)t * nr + k) * stride + i] = residual[k];
      wb.costs[(size_t)t * stride + i] = c;
      PS_SPEC_UNROLL
      for (int k = 0; k < task.num_trace; k++) {
        double tr[3]; read_sensor(m, d, task.trace_sensor[k], tr);
        PS_SPEC_UNROLL for (int x = 0; x < 3; x++) wb.trace[((size_t)t * ntr + 3 * k + x) * stride + i] = tr[x];
      }
    }
    if (terminal) { sum += c; break; }
    if (acc_is_bad(m, d) || (d.warning & (WARN_BADQPOS | WARN_BADQVEL | WARN_BADQACC | WARN_BADCTRL))) { failed = true; break; }
    sum += c;
    integrate(m, d);
    if (WB) {
      double* s = wb.states + ((size_t)(t + 1) * ns) * stride + i;
      PS_SPEC_UNROLL
      for (int k = 0; k < nq; k++) s[(size_t)k * stride] = d.qpos[k];
      PS_SPEC_UNROLL
      for (int k = 0; k < nv; k++) s[(size_t)(nq + k) * stride] = d.qvel[k];
      PS_SPEC_UNROLL
      for (int k = 0; k < na; k++) s[(size_t)(nq + nv + k) * stride] = d.act[k];
      wb.times[(size_t)(t + 1) * stride + i] = d.time;
    }
  }
  const double total = failed ? 1.0e6 : sum / (H > 1 ? H : 1);
#else
  for (int t = 0; t < H - 1; t++) {
    spline_sample(plan.spline_times, P, nu, plan.interpolation, d.time, kn, d.ctrl);
    PS_SPEC_UNROLL
    for (int a = 0; a < nu; a++) d.ctrl[a] = clampd(d.ctrl[a], m.act_ctrlrange[2 * a], m.act_ctrlrange[2 * a + 1]);
    if (WB) PS_SPEC_UNROLL for (int a = 0; a < nu; a++) wb.actions[((size_t)t * nu + a) * stride + i] = d.ctrl[a];
    pert.advance(m, t, stride, i);
    if (state_is_bad(m, d)) { failed = true; break; }
    pert.forward(m, d);
    task_residual(m, d, task, residual);
    const double c = cost_value(task, residual);
    if (WB) {
      PS_SPEC_UNROLL
      for (int k = 0; k < nr; k++) wb.residual[((size_t)t * nr + k) * stride + i] = residual[k];
      wb.costs[(size_t)t * stride + i] = c;
      PS_SPEC_UNROLL
      for (int k = 0; k < task.num_trace; k++) {
        double tr[3]; read_sensor(m, d, task.trace_sensor[k], tr);
        PS_SPEC_UNROLL for (int x = 0; x < 3; x++) wb.trace[((size_t)t * ntr + 3 * k + x) * stride + i] = tr[x];
      }
    }
    if (acc_is_bad(m, d) || (d.warning & (WARN_BADQPOS | WARN_BADQVEL | WARN_BADQACC | WARN_BADCTRL))) { failed = true; break; }
    sum += c;
    integrate(m, d);
    if (WB) {
      double* s = wb.states + ((size_t)(t + 1) * ns) * stride + i;
      PS_SPEC_UNROLL
      for (int k = 0; k < nq; k++) s[(size_t)k * stride] = d.qpos[k];
      PS_SPEC_UNROLL
      for (int k = 0; k < nv; k++) s[(size_t)(nq + k) * stride] = d.qvel[k];
      PS_SPEC_UNROLL
      for (int k = 0; k < na; k++) s[(size_t)(nq + nv + k) * stride] = d.act[k];
      wb.times[(size_t)(t + 1) * stride + i] = d.time;
    }
  }
  double total;
  if (!failed) {
    // terminal: ctrl is NOT rewritten before the final forward pass (trajectory.cc:190-198)
    if (WB && H > 1) PS_SPEC_UNROLL for (int a = 0; a < nu; a++) wb.actions[((size_t)(H - 1) * nu + a) * stride + i] = d.ctrl[a];
    // what an mjData stepped H - 1 times carries into its next mj_step (the terminal forward pass below is the planner's, not the plant's)
    if (WB) { PS_SPEC_UNROLL for (int k = 0; k < nv; k++) wb.warmstart[(size_t)k * stride + i] = d.qacc_warmstart[k]; }
    pert.forward(m, d);
    task_residual(m, d, task, residual);
    const double c = cost_value(task, residual);
    if (WB) {
      PS_SPEC_UNROLL
      for (int k = 0; k < nr; k++) wb.residual[((size_t)(H - 1) * nr + k) * stride + i] = residual[k];
      wb.costs[(size_t)(H - 1) * stride + i] = c;
      PS_SPEC_UNROLL
      for (int k = 0; k < task.num_trace; k++) {
        double tr[3]; read_sensor(m, d, task.trace_sensor[k], tr);
        PS_SPEC_UNROLL for (int x = 0; x < 3; x++) wb.trace[((size_t)(H - 1) * ntr + 3 * k + x) * stride + i] = tr[x];
      }
    }
    sum += c;
    total = sum / (H > 1 ? H : 1);
  } else total = 1.0e6;
#endif
  *flags = (failed ? 1 : 0) | ((d.warning & WARN_CAPACITY) ? 2 : 0);
  return total;
}

}  // namespace ps
