// CudaSamplingPlanner implementation: the host-side control flow of
// mjpc::SamplingPlanner (mjpc/planners/sampling/planner.cc) with Rollouts() replaced by the B200
// engine behind the C ABI.  What stays on the host is exactly what is O(P * nu): resampling the
// nominal spline, keeping policy / previous_policy for ActionFromPolicy, and bookkeeping.
#include "cuda_sampling_planner.h"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <numeric>

namespace mjpc_b200 {

namespace {
double Micros(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count();
}
int FindNumeric(const PsModel* m, const char* name) {
  for (int i = 0; i < m->nnumeric; i++) if (!std::strcmp(m->names + m->name_numericadr[i], name)) return i;
  return -1;
}
// GetNumberOrDefault (mjpc/utilities.h:48-64)
double NumberOrDefault(double dflt, const PsModel* m, const char* name) {
  const int i = FindNumeric(m, name);
  return i < 0 ? dflt : m->numeric_data[m->numeric_adr[i]];
}
void Clamp(double* x, const double* range, int n) {   // mjpc/utilities.cc:94-98
  for (int i = 0; i < n; i++) x[i] = std::min(std::max(x[i], range[2 * i]), range[2 * i + 1]);
}
}  // namespace

void SamplingPolicy::Action(double* action, const double* /*state*/, double time) const {
  plan.Sample(time, action);
  Clamp(action, model->actuator_ctrlrange, model->nu);
}

// Fatal configuration / device errors abort like mju_error does in the reference
// (e.g. planner.cc:69-72); rollout divergence is NOT an error (it scores 1e6).
void CudaSamplingPlanner::Fail(const char* what) const {
  std::fprintf(stderr, "CudaSamplingPlanner: %s: %s\n", what, ps_last_error());
  std::abort();
}

CudaSamplingPlanner::~CudaSamplingPlanner() { if (engine_) ps_destroy(engine_); }

void CudaSamplingPlanner::Initialize(const PsModel* model, const PsTask* task, int device_ordinal) {
  model_ = model; task_ = *task; device_ = device_ordinal;
  noise_exploration[0] = NumberOrDefault(0.1, model, "sampling_exploration");
  noise_exploration[1] = 0.0;
  const int se = FindNumeric(model, "sampling_exploration");
  if (se >= 0 && model->numeric_size[se] > 1) noise_exploration[1] = model->numeric_data[model->numeric_adr[se] + 1];
  num_trajectory_ = (int)NumberOrDefault(10, model, "sampling_trajectories");
  interpolation_ = (int)NumberOrDefault(kCubicSpline, model, "sampling_representation");
  sliding_plan_ = (int)NumberOrDefault(0, model, "sampling_sliding_plan");
  // the reference aborts above kMaxTrajectory = 128 (planner.cc:69-72); this engine has no such cap
  if (num_trajectory_ < 1) { std::fprintf(stderr, "CudaSamplingPlanner: sampling_trajectories must be positive\n"); std::abort(); }
  for (int i = 0; i < model->nu; i++)
    if (!model->actuator_ctrllimited[i]) { std::fprintf(stderr, "Ctrl limits required for all actuators.\n"); std::abort(); }  // agent.cc
  if (engine_) { ps_destroy(engine_); engine_ = nullptr; }
  engine_ = ps_create(device_ordinal);
  if (!engine_) Fail("ps_create");
  if (ps_set_model(engine_, model)) Fail("ps_set_model");
  if (ps_set_task(engine_, task)) Fail("ps_set_task");
  winner = 0;
}

void CudaSamplingPlanner::Allocate() {
  state.assign(model_->nq + model_->nv + model_->na, 0.0);
  mocap.assign(7 * model_->nmocap, 0.0);
  userdata.assign(model_->nuserdata, 0.0);
  const int P = std::min((int)NumberOrDefault(kMaxTrajectoryHorizon, model_, "sampling_spline_points"), kMaxSplinePoints);
  for (SamplingPolicy* p : {&policy, &previous_policy, &candidate_winner_}) {
    p->model = model_; p->num_spline_points = P; p->plan = TimeSpline(model_->nu);
  }
  winner = -1;
}

void CudaSamplingPlanner::Reset(int /*horizon*/, const double* initial_repeated_action) {
  std::fill(state.begin(), state.end(), 0.0); std::fill(mocap.begin(), mocap.end(), 0.0);
  std::fill(userdata.begin(), userdata.end(), 0.0); time = 0.0;
  for (SamplingPolicy* p : {&policy, &previous_policy, &candidate_winner_}) {
    p->plan.Clear();
    if (initial_repeated_action) p->plan.AddNode(0, initial_repeated_action);   // policy.cc:44-49
  }
  ranked_.clear(); ranked_score_.clear(); trajectory_order.clear(); ranked_local_.clear();
  improvement = 0.0; winner = 0; best_valid_ = false;
}

void CudaSamplingPlanner::SetState(const double* s, const double* mo, const double* ud, double t) {
  std::copy(s, s + state.size(), state.begin());
  if (mo) std::copy(mo, mo + mocap.size(), mocap.begin());
  if (ud) std::copy(ud, ud + userdata.size(), userdata.begin());
  time = t;
}

void CudaSamplingPlanner::SetTask(const PsTask* task) {
  task_ = *task;
  if (ps_set_task(engine_, task)) Fail("ps_set_task");
}

// The tensor is [N][P][nu]; P is only fixed once the nominal has been resampled, so the upload is
// deferred to the next OptimizePolicyCandidates (after ps_set_nominal).
void CudaSamplingPlanner::SetNoise(const double* z, const unsigned char* use_std2, int N, int P) {
  noise_pending_ = true; noise_n_ = z ? N : 0;
  noise_z_.clear(); noise_std2_.clear();
  if (z) {
    noise_z_.assign(z, z + (size_t)N * (P > 0 ? P : policy.num_spline_points) * model_->nu);
    if (use_std2) noise_std2_.assign(use_std2, use_std2 + N);
  }
}
void CudaSamplingPlanner::SetSeed(unsigned long long seed) { ps_set_seed(engine_, seed); }
// Replaces the nominal AND the policy the next UpdateNominalPolicy resamples from (candidate_policy[winner]).
void CudaSamplingPlanner::SetPolicy(const double* times, const double* knots, int P, int interpolation) {
  std::unique_lock<std::shared_mutex> lock(mtx_);
  interpolation_ = interpolation;
  policy.plan = TimeSpline(model_->nu, (SplineInterpolation)interpolation);
  for (int k = 0; k < P; k++) policy.plan.AddNode(times[k], knots + (size_t)k * model_->nu);
  candidate_winner_.plan = policy.plan;
}
void CudaSamplingPlanner::SetSharding(int rank, int world, AllGatherFn fn, void* ctx) {
  rank_ = rank; world_ = world; gather_ = fn; gather_ctx_ = ctx; native_comm_ = false;
}

// the engine's own NCCL exchange (ps_comm_init): the record ps_rollouts fetches is already the global one
void CudaSamplingPlanner::SetShardingNative(int rank, int world, const void* nccl_unique_id) {
  if (ps_comm_init(engine_, rank, world, nccl_unique_id)) Fail("ps_comm_init");
  rank_ = rank; world_ = world; gather_ = nullptr; gather_ctx_ = nullptr; native_comm_ = world > 1;
}

// planner.cc:236-310
void CudaSamplingPlanner::UpdateNominalPolicy(int horizon) {
  const int num_spline_points = candidate_winner_.num_spline_points;
  double nominal_time = time;
  const double time_horizon = (horizon - 1) * model_->opt_timestep;
  if (sliding_plan_) {
    const int extra_points = interpolation_ == kZeroSpline ? 1 : interpolation_ == kLinearSpline ? 2 : 4;
    const double time_shift = num_spline_points > extra_points
                                  ? std::max(time_horizon / (num_spline_points - extra_points), 1.0e-5) : time_horizon;
    std::unique_lock<std::shared_mutex> lock(mtx_);
    policy.plan.SetInterpolation((SplineInterpolation)interpolation_);
    policy.plan.DiscardBefore(nominal_time);
    if (policy.plan.Size() == 0) policy.plan.AddNode(time);
    while (policy.plan.Size() < num_spline_points) {
      const int last = policy.plan.Size() - 1;
      std::vector<double> v(policy.plan.NodeValues(last), policy.plan.NodeValues(last) + model_->nu);
      policy.plan.AddNode(policy.plan.NodeTime(last) + time_shift, v.data());
    }
  } else {
    const double time_shift = interpolation_ == kZeroSpline ? std::max(time_horizon / num_spline_points, 1.0e-5)
                                                            : std::max(time_horizon / (num_spline_points - 1), 1.0e-5);
    TimeSpline scratch(model_->nu, (SplineInterpolation)interpolation_);
    std::vector<double> v(model_->nu);
    for (int t = 0; t < num_spline_points; t++) {
      candidate_winner_.Action(v.data(), nullptr, nominal_time);
      scratch.AddNode(nominal_time, v.data());
      nominal_time += time_shift;
    }
    std::unique_lock<std::shared_mutex> lock(mtx_);
    policy.plan = scratch;
  }
}

// planner.cc:151-187 with Rollouts() = K1 + K2 + K3 on the device
int CudaSamplingPlanner::OptimizePolicyCandidates(int ncandidates, int horizon) {
  const int num_trajectory = num_trajectory_;          // read once: the GUI thread may change it (planner.cc:153-156)
  ncandidates = std::min(ncandidates, num_trajectory);
  const auto rollouts_start = std::chrono::steady_clock::now();
  {
    std::unique_lock<std::shared_mutex> lock(mtx_);
    policy.plan.SetInterpolation((SplineInterpolation)interpolation_);
  }
  const int P = policy.plan.Size(), nu = model_->nu;
  if (P < 1 || P > kMaxSplinePoints) { std::fprintf(stderr, "CudaSamplingPlanner: %d spline points (supported: 1..%d)\n", P, kMaxSplinePoints); std::abort(); }
  if (ps_set_options(engine_, model_->opt_timestep, model_->opt_integrator)) Fail("ps_set_options");   // agent.cc:275-276
  if (ps_set_state(engine_, state.data(), mocap.data(), userdata.data(), time)) Fail("ps_set_state");
  if (ps_set_nominal(engine_, policy.plan.times().data(), policy.plan.values().data(), P, interpolation_)) Fail("ps_set_nominal");
  if (noise_n_ && (noise_n_ < num_trajectory || noise_z_.size() != (size_t)noise_n_ * P * nu)) {
    // the injected tensor is indexed [sample][knot][actuator] with THIS iteration's P: a mismatch would misalign it
    std::fprintf(stderr, "CudaSamplingPlanner: injected noise is [%d][%zu] but the iteration needs [>= %d][%d]\n", noise_n_,
                 noise_n_ ? noise_z_.size() / noise_n_ : 0, num_trajectory, P * nu);
    std::abort();
  }
  if (noise_pending_) {
    if (ps_set_noise(engine_, noise_n_ ? noise_z_.data() : nullptr, noise_std2_.empty() ? nullptr : noise_std2_.data(), noise_n_)) Fail("ps_set_noise");
    noise_pending_ = false;
  }
  int local_winner = 0; double wret = 0, nret = 0;
  if (ps_rollouts(engine_, num_trajectory, horizon, noise_exploration[0], noise_exploration[1], rank_ * num_trajectory,
                  &local_winner, &wret, &nret)) Fail("ps_rollouts");
  last_horizon_ = horizon; best_valid_ = false;
  const int count = 3 + P * nu;
  std::vector<double> rec(count);
  if (ps_get_winner_record(engine_, rec.data(), count)) Fail("ps_get_winner_record");
  long long global_winner = (long long)rec[1];
  winner_is_local_ = local_winner >= 0; local_winner_ = local_winner >= 0 ? local_winner : 0;
  if (native_comm_) { wret = rec[0]; nret = rec[2]; }   // the all-gather min-loc already ran on the engine's stream
  if (world_ > 1 && gather_) {
    // min-loc over ranks: every rank contributes (return, global index, nominal return, knots)
    std::vector<double> all((size_t)world_ * count);
    if (gather_(gather_ctx_, rec.data(), all.data(), count)) { std::fprintf(stderr, "CudaSamplingPlanner: winner exchange failed\n"); std::abort(); }
    int best = 0;
    for (int r = 1; r < world_; r++) {
      const double c = all[(size_t)r * count], bc = all[(size_t)best * count];
      if (c < bc || (c == bc && all[(size_t)r * count + 1] < all[(size_t)best * count + 1])) best = r;
    }
    nret = all[2];                                        // global sample 0 (the un-noised nominal) lives on rank 0
    std::copy(all.begin() + (size_t)best * count, all.begin() + (size_t)(best + 1) * count, rec.begin());
    global_winner = (long long)rec[1]; wret = rec[0];
    winner_is_local_ = best == rank_;
  }
  nominal_return_ = nret; winner_return_ = wret;
  ranked_.clear(); ranked_score_.clear(); trajectory_order.clear(); ranked_local_.clear();
  auto make_policy = [&](const double* knots) {
    SamplingPolicy p; p.model = model_; p.num_spline_points = policy.num_spline_points;
    p.plan = TimeSpline(nu, (SplineInterpolation)interpolation_);
    for (int k = 0; k < P; k++) p.plan.AddNode(policy.plan.NodeTime(k), knots + (size_t)k * nu);
    return p;
  };
  trajectory_order.push_back((int)global_winner); ranked_.push_back(make_policy(rec.data() + 3)); ranked_score_.push_back(wret);
  ranked_local_.push_back(winner_is_local_ ? local_winner : -1);
  if (ncandidates > 1) {
    // top-k beyond the argmin: rank the per-sample returns on the host (single-rank only; RobustPlanner path)
    std::vector<double> ret(num_trajectory);
    if (ps_get_returns(engine_, ret.data(), nullptr)) Fail("ps_get_returns");
    std::vector<int> order(num_trajectory); std::iota(order.begin(), order.end(), 0);
    std::partial_sort(order.begin(), order.begin() + ncandidates, order.end(),
                      [&](int a, int b) { return ret[a] < ret[b] || (ret[a] == ret[b] && a < b); });
    std::vector<double> knots((size_t)P * nu);
    for (int c = 1; c < ncandidates; c++) {
      if (ps_get_knots(engine_, order[c], knots.data())) Fail("ps_get_knots");
      trajectory_order.push_back(order[c] + rank_ * num_trajectory); ranked_.push_back(make_policy(knots.data())); ranked_score_.push_back(ret[order[c]]);
      ranked_local_.push_back(order[c]);
    }
  }
  rollouts_compute_time = Micros(rollouts_start);
  float pm = 0, rm = 0, am = 0;
  if (!ps_last_kernel_ms(engine_, &pm, &rm, &am)) noise_compute_time = pm * 1.0e3;
  return ncandidates;
}

// planner.cc:190-208
void CudaSamplingPlanner::OptimizePolicy(int horizon) {
  UpdateNominalPolicy(horizon);
  OptimizePolicyCandidates(1, horizon);
  const auto policy_update_start = std::chrono::steady_clock::now();
  CopyCandidateToPolicy(0);
  improvement = std::max(nominal_return_ - winner_return_, 0.0);
  policy_update_compute_time = Micros(policy_update_start);
}

double CudaSamplingPlanner::CandidateScore(int candidate) const { return ranked_score_[candidate]; }

void CudaSamplingPlanner::ActionFromCandidatePolicy(double* action, int candidate, const double* s, double t) {
  ranked_[candidate].Action(action, s, t);
}

// planner.cc:525-534
void CudaSamplingPlanner::CopyCandidateToPolicy(int candidate) {
  winner = trajectory_order[candidate];
  // BestTrajectory() is trajectory[winner] (planner.cc:383-385): follow the candidate that was picked
  if (candidate < (int)ranked_local_.size() && ranked_local_[candidate] != local_winner_) {
    winner_is_local_ = ranked_local_[candidate] >= 0;
    if (winner_is_local_) local_winner_ = ranked_local_[candidate];
    best_valid_ = false;
  }
  std::unique_lock<std::shared_mutex> lock(mtx_);
  previous_policy = policy;
  policy = ranked_[candidate];
  candidate_winner_ = ranked_[candidate];
}

// planner.cc:225-233
void CudaSamplingPlanner::ActionFromPolicy(double* action, const double* s, double t, bool use_previous) {
  std::shared_lock<std::shared_mutex> lock(mtx_);
  if (use_previous) previous_policy.Action(action, s, t); else policy.Action(action, s, t);
}

// planner.cc:211-223: rolls out the current nominal (un-noised sample 0) with full writeback
void CudaSamplingPlanner::NominalTrajectory(int horizon) {
  const SamplingPolicy& src = candidate_winner_;
  if (src.plan.Size() < 1) return;
  if (ps_set_options(engine_, model_->opt_timestep, model_->opt_integrator)) Fail("ps_set_options");
  if (ps_set_state(engine_, state.data(), mocap.data(), userdata.data(), time)) Fail("ps_set_state");
  if (ps_set_nominal(engine_, src.plan.times().data(), src.plan.values().data(), src.plan.Size(), src.plan.Interpolation())) Fail("ps_set_nominal");
  int w; double a, b;
  if (ps_rollouts(engine_, 1, horizon, 0.0, 0.0, 0, &w, &a, &b)) Fail("ps_rollouts");
  // the engine batch is now this single rollout: candidates of the previous iteration are gone
  ranked_.clear(); ranked_score_.clear(); trajectory_order.clear(); ranked_local_.clear();
  last_horizon_ = horizon; local_winner_ = 0; winner_is_local_ = true; best_valid_ = false; winner = 0;
}

// planner.cc:383-385.  The winner's full trajectory is re-rolled with writeback on demand (the
// cost-only production kernel writes 8 bytes per sample); on a rank that does not own the global
// winner this returns nullptr.
const Trajectory* CudaSamplingPlanner::BestTrajectory() {
  if (winner < 0 || !winner_is_local_ || last_horizon_ <= 0) return nullptr;
  if (!best_valid_) {
    const int H = last_horizon_, ns = (int)state.size(), nu = model_->nu, nr = task_.num_residual, nt = 3 * task_.num_trace;
    best_.horizon = H; best_.dim_state = ns; best_.dim_action = nu; best_.dim_residual = nr; best_.dim_trace = nt;
    best_.states.assign((size_t)H * ns, 0); best_.actions.assign((size_t)H * nu, 0); best_.times.assign(H, 0);
    best_.residual.assign((size_t)H * nr, 0); best_.costs.assign(H, 0); best_.trace.assign((size_t)H * std::max(nt, 1), 0);
    int failure = 0;
    if (ps_get_trajectory(engine_, local_winner_, best_.states.data(), best_.actions.data(), best_.times.data(), best_.residual.data(),
                          best_.costs.data(), best_.trace.data(), &best_.total_return, &failure)) Fail("ps_get_trajectory");
    best_.failure = failure != 0; best_valid_ = true;
  }
  return &best_;
}

}  // namespace mjpc_b200
