// CudaCrossEntropyPlanner implementation: the host-side control flow of mjpc::CrossEntropyPlanner
// (mjpc/planners/cross_entropy/planner.cc) with Rollouts() and the elite statistics on the B200
// engine.  The host keeps what is O(P * nu): resampling the nominal spline, the policies served
// by ActionFromPolicy, the variance vector.
#include "cuda_cross_entropy_planner.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace mjpc_b200 {

namespace {
double Micros(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
}
int FindNumeric(const PsModel* m, const char* name) {
  for (int i = 0; i < m->nnumeric; i++) if (!std::strcmp(m->names + m->name_numericadr[i], name)) return i;
  return -1;
}
double NumberOrDefault(double dflt, const PsModel* m, const char* name) {   // mjpc/utilities.h:48-64
  const int i = FindNumeric(m, name);
  return i < 0 ? dflt : m->numeric_data[m->numeric_adr[i]];
}
}  // namespace

void CudaCrossEntropyPlanner::Fail(const char* what) const {
  std::fprintf(stderr, "CudaCrossEntropyPlanner: %s: %s\n", what, ps_last_error());
  std::abort();
}

CudaCrossEntropyPlanner::~CudaCrossEntropyPlanner() { if (engine_) ps_destroy(engine_); }

// cross_entropy/planner.cc:40-57
void CudaCrossEntropyPlanner::Initialize(const PsModel* model, const PsTask* task, int device_ordinal) {
  model_ = model; task_ = *task;
  std_initial_ = NumberOrDefault(0.1, model, "sampling_exploration");
  std_min_ = NumberOrDefault(0.1, model, "std_min");
  num_trajectory_ = (int)NumberOrDefault(10, model, "sampling_trajectories");
  n_elite_ = (int)NumberOrDefault(std::max(num_trajectory_ / 10, 2), model, "n_elite");
  interpolation_ = (int)NumberOrDefault(kCubicSpline, model, "sampling_representation");
  // the reference aborts above kMaxTrajectory = 128 (:52-55); this engine has no such cap
  if (num_trajectory_ < 1) { std::fprintf(stderr, "CudaCrossEntropyPlanner: sampling_trajectories must be positive\n"); std::abort(); }
  if (engine_) { ps_destroy(engine_); engine_ = nullptr; }
  engine_ = ps_create(device_ordinal);
  if (!engine_) Fail("ps_create");
  if (ps_set_model(engine_, model)) Fail("ps_set_model");
  if (ps_set_task(engine_, task)) Fail("ps_set_task");
}

void CudaCrossEntropyPlanner::Allocate() {
  state.assign(model_->nq + model_->nv + model_->na, 0.0);
  mocap.assign(7 * model_->nmocap, 0.0);
  userdata.assign(model_->nuserdata, 0.0);
  const int P = std::min((int)NumberOrDefault(kMaxTrajectoryHorizon, model_, "sampling_spline_points"), kMaxSplinePoints);
  for (SamplingPolicy* p : {&policy, &previous_policy, &resampled_policy}) {
    p->model = model_; p->num_spline_points = P; p->plan = TimeSpline(model_->nu);
  }
  variance.assign((size_t)model_->nu * kMaxSplinePoints, 0.0);
}

void CudaCrossEntropyPlanner::Reset(int /*horizon*/, const double* initial_repeated_action) {
  std::fill(state.begin(), state.end(), 0.0); std::fill(mocap.begin(), mocap.end(), 0.0);
  std::fill(userdata.begin(), userdata.end(), 0.0); time = 0.0;
  for (SamplingPolicy* p : {&policy, &previous_policy, &resampled_policy}) {
    p->plan.Clear();
    if (initial_repeated_action) p->plan.AddNode(0, initial_repeated_action);
  }
  std::fill(variance.begin(), variance.end(), std_initial_ * std_initial_);   // :109-111
  trajectory_order.clear();
  improvement = 0.0; nominal_valid_ = false;
}

void CudaCrossEntropyPlanner::SetState(const double* s, const double* mo, const double* ud, double t) {
  std::copy(s, s + state.size(), state.begin());
  if (mo) std::copy(mo, mo + mocap.size(), mocap.begin());
  if (ud) std::copy(ud, ud + userdata.size(), userdata.begin());
  time = t;
}

void CudaCrossEntropyPlanner::SetTask(const PsTask* task) {
  task_ = *task;
  if (ps_set_task(engine_, task)) Fail("ps_set_task");
}

void CudaCrossEntropyPlanner::SetNoise(const double* z, int N, int P) {
  noise_pending_ = true; noise_n_ = z ? N : 0; noise_z_.clear();
  if (z) noise_z_.assign(z, z + (size_t)N * (P > 0 ? P : policy.num_spline_points) * model_->nu);
}
void CudaCrossEntropyPlanner::SetSeed(unsigned long long seed) { ps_set_seed(engine_, seed); }

// cross_entropy/planner.cc:316-341 (no zero-order special case in the knot spacing)
void CudaCrossEntropyPlanner::ResamplePolicy(int horizon) {
  const int num_spline_points = resampled_policy.num_spline_points;
  double nominal_time = time;
  const double time_shift = std::max((horizon - 1) * model_->opt_timestep / (num_spline_points - 1), 1.0e-5);
  TimeSpline scratch(model_->nu, resampled_policy.plan.Interpolation());
  std::vector<double> v(model_->nu);
  for (int t = 0; t < num_spline_points; t++) {
    resampled_policy.Action(v.data(), nullptr, nominal_time);
    scratch.AddNode(nominal_time, v.data());
    nominal_time += time_shift;
  }
  scratch.SetInterpolation(policy.plan.Interpolation());
  resampled_policy.plan = scratch;
}

// cross_entropy/planner.cc:165-284
void CudaCrossEntropyPlanner::OptimizePolicy(int horizon) {
  const int num_trajectory = num_trajectory_;          // read once (:170)
  n_elite_ = std::min(n_elite_, num_trajectory);
  const int n_elite = n_elite_;
  {
    std::shared_lock<std::shared_mutex> lock(mtx_);
    resampled_policy.plan = policy.plan;
    resampled_policy.num_spline_points = policy.num_spline_points;
  }
  resampled_policy.plan.SetInterpolation((SplineInterpolation)interpolation_);
  ResamplePolicy(horizon);

  const auto rollouts_start = std::chrono::steady_clock::now();
  const int P = resampled_policy.plan.Size(), nu = model_->nu, np = P * nu;
  if (P < 2 || P > kMaxSplinePoints) { std::fprintf(stderr, "CudaCrossEntropyPlanner: %d spline points (supported: 2..%d)\n", P, kMaxSplinePoints); std::abort(); }
  if (ps_set_options(engine_, model_->opt_timestep, model_->opt_integrator)) Fail("ps_set_options");
  if (ps_set_state(engine_, state.data(), mocap.data(), userdata.data(), time)) Fail("ps_set_state");
  if (ps_set_nominal(engine_, resampled_policy.plan.times().data(), resampled_policy.plan.values().data(), P, interpolation_)) Fail("ps_set_nominal");
  if (noise_n_ && (noise_n_ < num_trajectory || noise_z_.size() != (size_t)noise_n_ * np)) {
    std::fprintf(stderr, "CudaCrossEntropyPlanner: injected noise is [%d][%zu] but the iteration needs [>= %d][%d]\n", noise_n_,
                 noise_n_ ? noise_z_.size() / noise_n_ : 0, num_trajectory, np);
    std::abort();
  }
  if (noise_pending_) {
    // the injected tensor covers the num_trajectory candidates; the nominal column is never noised
    std::vector<double> z(noise_z_);
    if (noise_n_) z.resize((size_t)(noise_n_ + 1) * np, 0.0);
    if (ps_set_noise(engine_, noise_n_ ? z.data() : nullptr, nullptr, noise_n_ ? noise_n_ + 1 : 0)) Fail("ps_set_noise");
    noise_pending_ = false;
  }
  std::vector<double> sd(np);
  for (int k = 0; k < np; k++) sd[k] = std::max(std::sqrt(variance[k]), std_min_);   // :360-363
  if (ps_set_param_std(engine_, sd.data())) Fail("ps_set_param_std");
  // candidates 0 .. N-1 plus the nominal rollout (:416) as column N of the same launch
  int w = 0; double wret = 0, first = 0;
  if (ps_rollouts(engine_, num_trajectory + 1, horizon, 0.0, 0.0, 0, &w, &wret, &first)) Fail("ps_rollouts");
  if (ps_set_param_std(engine_, nullptr)) Fail("ps_set_param_std");
  last_horizon_ = horizon; last_n_ = num_trajectory; nominal_valid_ = false;
  rollouts_compute_time = Micros(rollouts_start);

  const auto policy_update_start = std::chrono::steady_clock::now();
  std::vector<double> mean(np), var(np);
  trajectory_order.assign(n_elite, 0);
  if (ps_elite_update(engine_, num_trajectory, n_elite, mean.data(), var.data(), &avg_return, &best_return, trajectory_order.data())) Fail("ps_elite_update");
  std::fill(variance.begin(), variance.end(), 0.0);     // :243
  std::copy(var.begin(), var.end(), variance.begin());
  {
    std::unique_lock<std::shared_mutex> lock(mtx_);
    TimeSpline plan(nu, (SplineInterpolation)interpolation_);
    for (int t = 0; t < P; t++) plan.AddNode(resampled_policy.plan.NodeTime(t), mean.data() + (size_t)t * nu);
    policy.plan = plan;
  }
  improvement = std::max(avg_return - best_return, 0.0);   // :277-278
  policy_update_compute_time = Micros(policy_update_start);
  float pm = 0, rm = 0, am = 0;
  if (!ps_last_kernel_ms(engine_, &pm, &rm, &am)) noise_compute_time = pm * 1.0e3;
}

// cross_entropy/planner.cc:287-301: rolls out the resampled nominal with full writeback
void CudaCrossEntropyPlanner::NominalTrajectory(int horizon) {
  const SamplingPolicy& src = resampled_policy;
  if (src.plan.Size() < 1) return;
  if (ps_set_options(engine_, model_->opt_timestep, model_->opt_integrator)) Fail("ps_set_options");
  if (ps_set_state(engine_, state.data(), mocap.data(), userdata.data(), time)) Fail("ps_set_state");
  if (ps_set_nominal(engine_, src.plan.times().data(), src.plan.values().data(), src.plan.Size(), src.plan.Interpolation())) Fail("ps_set_nominal");
  int w; double a, b;
  if (ps_rollouts(engine_, 1, horizon, 0.0, 0.0, 0, &w, &a, &b)) Fail("ps_rollouts");
  last_horizon_ = horizon; last_n_ = 0; nominal_valid_ = false;
}

void CudaCrossEntropyPlanner::ActionFromPolicy(double* action, const double* s, double t, bool use_previous) {
  std::shared_lock<std::shared_mutex> lock(mtx_);
  if (use_previous) previous_policy.Action(action, s, t); else policy.Action(action, s, t);
}

// cross_entropy/planner.cc:426-428: the NOMINAL trajectory (column last_n_ of the last launch)
const Trajectory* CudaCrossEntropyPlanner::BestTrajectory() {
  if (last_horizon_ <= 0) return nullptr;
  if (!nominal_valid_) {
    const int H = last_horizon_, ns = (int)state.size(), nu = model_->nu, nr = task_.num_residual, nt = 3 * task_.num_trace;
    nominal_.horizon = H; nominal_.dim_state = ns; nominal_.dim_action = nu; nominal_.dim_residual = nr; nominal_.dim_trace = nt;
    nominal_.states.assign((size_t)H * ns, 0); nominal_.actions.assign((size_t)H * nu, 0); nominal_.times.assign(H, 0);
    nominal_.residual.assign((size_t)H * nr, 0); nominal_.costs.assign(H, 0); nominal_.trace.assign((size_t)H * std::max(nt, 1), 0);
    int failure = 0;
    if (ps_get_trajectory(engine_, last_n_, nominal_.states.data(), nominal_.actions.data(), nominal_.times.data(), nominal_.residual.data(),
                          nominal_.costs.data(), nominal_.trace.data(), &nominal_.total_return, &failure)) Fail("ps_get_trajectory");
    nominal_.failure = failure != 0; nominal_valid_ = true; nominal_return = nominal_.total_return;
  }
  return &nominal_;
}

}  // namespace mjpc_b200
