// Policy evaluation and cost side of one rollout: TimeSpline sampling (zero / linear / cubic
// Hermite with finite-difference slopes), the MJPC norms, the weighted cost with risk transform,
// and the per-task residual device functions.
//
// Reference semantics: mjpc/spline/spline.cc:103-156,240-277 (Sample / Slope),
// mjpc/planners/sampling/policy.cc:52-59 (Action = Sample + Clamp), mjpc/norm.cc:50-210 (value
// path), mjpc/task.cc:71-110 (CostTerms / CostValue), mjpc/tasks/<task>/<task>.cc (Residual).
#pragma once
#include "ps_step.cuh"
#include "ps_quadruped.cuh"

namespace ps {

// ---------------------------------------------------------------- spline
// Knots(p, a) returns knot value of spline point p, action dimension a.
template <class Knots>
PS_FN void spline_sample(const double* times, int P, int dim, int interp, double t, const Knots& knots, double* out) {
  if (P == 0) { for (int a = 0; a < dim; a++) out[a] = 0; return; }
  int up = 0;
  while (up < P && !(t < times[up])) up++;          // first node with time > t
  if (up == P) { for (int a = 0; a < dim; a++) out[a] = knots(P - 1, a); return; }
  if (up == 0) { for (int a = 0; a < dim; a++) out[a] = knots(0, a); return; }
  const int lo = up - 1;
  if (interp == SPLINE_ZERO) { for (int a = 0; a < dim; a++) out[a] = knots(lo, a); return; }
  const double dt = times[up] - times[lo];
  const double tt = (t - times[lo]) / dt;
  if (interp == SPLINE_LINEAR) { for (int a = 0; a < dim; a++) out[a] = knots(lo, a) * (1 - tt) + knots(up, a) * tt; return; }
  const double c0 = 2.0 * tt * tt * tt - 3.0 * tt * tt + 1.0;
  const double c1 = (tt * tt * tt - 2.0 * tt * tt + tt) * dt;
  const double c2 = -2.0 * tt * tt * tt + 3 * tt * tt;
  const double c3 = (tt * tt * tt - tt * tt) * dt;
  for (int a = 0; a < dim; a++) {
    const double vl = knots(lo, a), vu = knots(up, a);
    // slopes: one-sided at the ends, mean of the adjacent secants inside
    double sl, su;
    const double sec = (vu - vl) / dt;
    if (lo == 0) sl = sec;
    else sl = 0.5 * sec + 0.5 * (vl - knots(lo - 1, a)) / (times[lo] - times[lo - 1]);
    if (up == P - 1) su = sec;
    else su = 0.5 * (knots(up + 1, a) - vu) / (times[up + 1] - times[up]) + 0.5 * sec;
    out[a] = c0 * vl + c1 * sl + c2 * vu + c3 * su;
  }
}

// ---------------------------------------------------------------- norms (NormType, mjpc/norm.h:24-35)
PS_NT double norm_value(const double* x, const double* params, int n, int type) {
  double y = 0; const double p = params[0], q = params[1];
  switch (type) {
    case -1: y = x[0]; break;                                                           // kNull
    case 0: for (int i = 0; i < n; i++) y += x[i] * x[i]; y *= 0.5; break;              // kQuadratic
    case 1: { double c = 0; for (int i = 0; i < n; i++) c += x[i] * x[i];               // kL22
              const double a = pow(c, q / 2) + pow(p, q); y = pow(a, 1 / q) - p; } break;
    case 2: { double c = 0; for (int i = 0; i < n; i++) c += x[i] * x[i]; y = sqrt(c + p * p) - p; } break;  // kL2
    case 3: for (int i = 0; i < n; i++) y += p * p * (cosh(x[i] / p) - 1.0); break;     // kCosh
    case 5: for (int i = 0; i < n; i++) y += pow(fabs(x[i]), p); break;                 // kPowerLoss
    case 6: for (int i = 0; i < n; i++) y += sqrt(x[i] * x[i] + p * p) - p; break;      // kSmoothAbsLoss
    case 7: for (int i = 0; i < n; i++) { const double e = pow(fabs(x[i]), q) + pow(p, q); y += pow(e, 1 / q) - p; } break;  // kSmoothAbs2Loss
    case 8: for (int i = 0; i < n; i++) { if (p > 0) y += p * log(1 + exp(x[i] / p)); else y += x[i] > 0 ? x[i] : 0; } break;  // kRectifyLoss
    default: break;
  }
  return y;
}

template <class TK>
PS_NT double cost_value(const TK& t, const double* residual) {
  double cost = 0; int fs = 0, ps_ = 0;
  PS_SPEC_UNROLL
  for (int k = 0; k < t.num_term; k++) {
    double prm[2] = {0, 0};
    PS_SPEC_UNROLL
    for (int j = 0; j < t.num_norm_parameter[k]; j++) prm[j] = t.norm_parameter[ps_ + j];
    cost += t.weight[k] * norm_value(residual + fs, prm, t.dim_norm_residual[k], t.norm[k]);
    fs += t.dim_norm_residual[k]; ps_ += t.num_norm_parameter[k];
  }
  if (fabs(t.risk) < 1.0e-6) return cost;
  return (exp(t.risk * cost) - 1.0) / t.risk;
}

// ---------------------------------------------------------------- residuals
enum { TASK_PARTICLE = 0, TASK_PARTICLE_FIXED, TASK_PARTICLE_TEST, TASK_CARTPOLE, TASK_ACROBOT, TASK_SWIMMER,
       TASK_WALKER, TASK_QUADRUPED };

template <class T, class D, class TK>
PS_NT void task_residual(const DevModel<T>& m, const D& d, const TK& t, double* r) {
  double s0[3], s1[3];
  switch (t.task_id) {
    case TASK_PARTICLE: case TASK_PARTICLE_FIXED: {
      // particle.cc:36-56 (time-varying goal) / :73-78 (mocap goal); sensor_id = {position, velocity}
      double goal[2];
      if (t.task_id == TASK_PARTICLE) { goal[0] = 0.25 * sin_det(d.time); goal[1] = 0.25 * cos_det(d.time / kPi); }
      else { goal[0] = d.mocap_pos[0]; goal[1] = d.mocap_pos[1]; }
      read_sensor(m, d, t.sensor_id[0], s0); read_sensor(m, d, t.sensor_id[1], s1);
      r[0] = s0[0] - goal[0]; r[1] = s0[1] - goal[1]; r[2] = s1[0]; r[3] = s1[1];
      r[4] = d.ctrl[0]; r[5] = d.ctrl[1];
    } break;
    case TASK_PARTICLE_TEST:
      // test/testdata/particle_residual.h:33-43
      r[0] = d.qpos[0] - d.mocap_pos[0]; r[1] = d.qpos[1] - d.mocap_pos[1]; r[2] = d.qvel[0]; r[3] = d.qvel[1];
      break;
    case TASK_CARTPOLE:
      // cartpole.cc:36-49
      r[0] = cos_det(d.qpos[1]) - 1; r[1] = d.qpos[0] - t.parameters[0]; r[2] = d.qvel[1]; r[3] = d.ctrl[0];
      break;
    case TASK_ACROBOT:
      // acrobot.cc:36-49: tip (site 0) to target (site 1), z then x
      r[0] = d.site_xpos[2] - d.site_xpos[5]; r[1] = d.site_xpos[0] - d.site_xpos[3];
      r[2] = d.qvel[0]; r[3] = d.qvel[1]; r[4] = d.ctrl[0];
      break;
    case TASK_SWIMMER: if (!T::LEAN) {
      // swimmer.cc:33-45; sensor_id = {nose, target}
      for (int i = 0; i < m.nu; i++) r[i] = d.ctrl[i];
      read_sensor(m, d, t.sensor_id[0], s0); read_sensor(m, d, t.sensor_id[1], s1);
      r[m.nu] = s0[0] - s1[0]; r[m.nu + 1] = s0[1] - s1[1];
    } break;
    case TASK_WALKER: if (!T::LEAN) {
      // walker.cc:38-69; sensor_id = {torso_position, torso_zaxis, torso_subtreelinvel}
      int c = 0; for (int i = 0; i < m.nu; i++) r[c++] = d.ctrl[i];
      read_sensor(m, d, t.sensor_id[0], s0); r[c++] = s0[2] - t.parameters[0];
      read_sensor(m, d, t.sensor_id[1], s0); r[c++] = s0[2] - 1.0;
      read_sensor(m, d, t.sensor_id[2], s0); r[c++] = s0[0] - t.parameters[1];
    } break;
    case TASK_QUADRUPED: if constexpr (T::NQ != T::NV) quadruped_residual(m, d, t, r); break;   // needs the free joint
    default: break;
  }
}

}  // namespace ps
