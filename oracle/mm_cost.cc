// ORACLE (test infrastructure).  Restatement of the reference's cost side: TimeSpline::Sample
// (mjpc/spline/spline.cc:103-156,240-277), Norm value path (mjpc/norm.cc:50-210),
// BaseResidualFn::CostTerms/CostValue (mjpc/task.cc:71-110) and the task residuals
// (mjpc/tasks/<task>/<task>.cc).  Pinned against the reference's known answers in
// tests/test_oracle_known_answers.py (spline_test.cc:115-160, task_test.cc:76-95).
#include <algorithm>
#include <cmath>
#include "mm_planner.h"
#include "mm_math.h"

namespace mm {

// TimeSpline::Sample: times[P] ascending, knots[P][dim]
void splineSample(const double* times, const double* knots, int P, int dim, int interp, double t, double* out) {
  if (P == 0) { zero(out, dim); return; }
  int up = (int)(std::upper_bound(times, times + P, t) - times);
  if (up == P) { copy(out, knots + (P - 1) * dim, dim); return; }
  if (up == 0) { copy(out, knots, dim); return; }
  int lo = up - 1;
  double tt = (t - times[lo]) / (times[up] - times[lo]);
  const double* pl = knots + lo * dim; const double* pu = knots + up * dim;
  if (interp == PS_SPLINE_ZERO) { copy(out, pl, dim); return; }
  if (interp == PS_SPLINE_LINEAR) { for (int i = 0; i < dim; i++) out[i] = pl[i] * (1 - tt) + pu[i] * tt; return; }
  double dt = times[up] - times[lo];
  double c0 = 2.0 * tt * tt * tt - 3.0 * tt * tt + 1.0;
  double c1 = (tt * tt * tt - 2.0 * tt * tt + tt) * dt;
  double c2 = -2.0 * tt * tt * tt + 3 * tt * tt;
  double c3 = (tt * tt * tt - tt * tt) * dt;
  auto slope = [&](int n, int i) {
    const double* v = knots + n * dim;
    if (n == 0) return (knots[(n + 1) * dim + i] - v[i]) / (times[n + 1] - times[n]);
    const double* pv = knots + (n - 1) * dim;
    if (n == P - 1) return (v[i] - pv[i]) / (times[n] - times[n - 1]);
    const double* nx = knots + (n + 1) * dim;
    return 0.5 * (nx[i] - v[i]) / (times[n + 1] - times[n]) + 0.5 * (v[i] - pv[i]) / (times[n] - times[n - 1]);
  };
  for (int i = 0; i < dim; i++) out[i] = c0 * pl[i] + c1 * slope(lo, i) + c2 * pu[i] + c3 * slope(up, i);
}

double normValue(const double* x, const double* params, int n, int type) {
  double y = 0, p = params ? params[0] : double(0), q = params ? params[1] : double(0);
  switch (type) {
    case -1: y = x[0]; break;
    case 0: for (int i = 0; i < n; i++) y += x[i] * x[i]; y *= 0.5; break;
    case 1: { double c = 0; for (int i = 0; i < n; i++) c += x[i] * x[i];
              double a = std::pow(c, q / 2) + std::pow(p, q); y = std::pow(a, 1 / q) - p; } break;
    case 2: y = std::sqrt(dot(x, x, n) + p * p) - p; break;
    case 3: for (int i = 0; i < n; i++) y += p * p * (std::cosh(x[i] / p) - 1.0); break;
    case 5: for (int i = 0; i < n; i++) y += std::pow(std::fabs(x[i]), p); break;
    case 6: for (int i = 0; i < n; i++) y += std::sqrt(x[i] * x[i] + p * p) - p; break;
    case 7: for (int i = 0; i < n; i++) { double e = std::pow(std::fabs(x[i]), q) + std::pow(p, q); y += std::pow(e, 1 / q) - p; } break;
    case 8: for (int i = 0; i < n; i++) { if (p > 0) y += p * std::log(1 + std::exp(x[i] / p)); else y += x[i] > 0 ? x[i] : double(0); } break;
    default: break;
  }
  return y;
}

void costTerms(const PsTask* t, const double* residual, double* terms, bool weighted) {
  int fs = 0, ps = 0;
  for (int k = 0; k < t->num_term; k++) {
    terms[k] = (weighted ? t->weight[k] : double(1)) * normValue(residual + fs, t->norm_parameter + ps, t->dim_norm_residual[k], t->norm[k]);
    fs += t->dim_norm_residual[k]; ps += t->num_norm_parameter[k];
  }
}

double costValue(const PsTask* t, const double* residual) {
  double terms[PS_MAX_COST_TERMS]; costTerms(t, residual, terms, true);
  double cost = 0; for (int i = 0; i < t->num_term; i++) cost += terms[i];
  if (std::fabs(t->risk) < 1.0e-6) return cost;
  return (std::exp(t->risk * cost) - 1.0) / t->risk;
}

static const double* sensorByName(const Model* m, const Data* d, const char* name) {
  int id = name2id(m, PS_OBJ_SENSOR, name);
  return id < 0 ? nullptr : &d->sensordata[m->sensor_adr[id]];
}

void quadrupedResidual(const Model* m, const Data* d, const PsTask* t, double* r);

void residual(const Model* m, const Data* d, const PsTask* t, double* r) {
  switch (t->task_id) {
    case PS_TASK_PARTICLE: case PS_TASK_PARTICLE_FIXED: {
      double goal[2];
      if (t->task_id == PS_TASK_PARTICLE) { goal[0] = 0.25 * sinDet(d->time); goal[1] = 0.25 * cosDet(d->time / kPi); }
      else { goal[0] = d->mocap_pos[0]; goal[1] = d->mocap_pos[1]; }
      const double* pos = sensorByName(m, d, "position"); const double* vel = sensorByName(m, d, "velocity");
      for (int i = 0; i < m->nq; i++) r[i] = pos[i] - goal[i];
      for (int i = 0; i < m->nv; i++) r[2 + i] = vel[i];
      for (int i = 0; i < m->nu; i++) r[4 + i] = d->ctrl[i];
    } break;
    case PS_TASK_PARTICLE_TEST:
      for (int i = 0; i < m->nq; i++) r[i] = d->qpos[i];
      r[0] -= d->mocap_pos[0]; r[1] -= d->mocap_pos[1];
      for (int i = 0; i < m->nv; i++) r[2 + i] = d->qvel[i];
      break;
    case PS_TASK_CARTPOLE:
      r[0] = cosDet(d->qpos[1]) - 1; r[1] = d->qpos[0] - t->parameters[0]; r[2] = d->qvel[1]; r[3] = d->ctrl[0];
      break;
    case PS_TASK_ACROBOT:
      r[0] = d->site_xpos[2] - d->site_xpos[5]; r[1] = d->site_xpos[0] - d->site_xpos[3];
      r[2] = d->qvel[0]; r[3] = d->qvel[1]; r[4] = d->ctrl[0];
      break;
    case PS_TASK_SWIMMER: {
      for (int i = 0; i < m->nu; i++) r[i] = d->ctrl[i];
      const double* target = sensorByName(m, d, "target"); const double* nose = sensorByName(m, d, "nose");
      r[m->nu] = nose[0] - target[0]; r[m->nu + 1] = nose[1] - target[1];
    } break;
    case PS_TASK_WALKER: {
      int c = 0; for (int i = 0; i < m->nu; i++) r[c++] = d->ctrl[i];
      r[c++] = sensorByName(m, d, "torso_position")[2] - t->parameters[0];
      r[c++] = sensorByName(m, d, "torso_zaxis")[2] - 1.0;
      r[c++] = sensorByName(m, d, "torso_subtreelinvel")[0] - t->parameters[1];
    } break;
    case PS_TASK_QUADRUPED: quadrupedResidual(m, d, t, r); break;
  }
}

}  // namespace mm
