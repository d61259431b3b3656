// CudaRobustPlanner implementation: the control flow of mjpc::RobustPlanner
// (mjpc/planners/robust/robust_planner.cc) with both rollout stages on the B200 engine.
#include "cuda_robust_planner.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace mjpc_b200 {

namespace {
// GetNumberOrDefault (mjpc/utilities.h:48-64)
double Numeric(double dflt, const PsModel* m, const char* name) {
  for (int i = 0; i < m->nnumeric; i++)
    if (!std::strcmp(m->names + m->name_numericadr[i], name)) return m->numeric_data[m->numeric_adr[i]];
  return dflt;
}
}  // namespace

void CudaRobustPlanner::Fail(const char* what) const {
  std::fprintf(stderr, "CudaRobustPlanner: %s: %s\n", what, ps_last_error());
  std::abort();
}

CudaRobustPlanner::~CudaRobustPlanner() { if (engine_) ps_destroy(engine_); }

// robust_planner.cc:31-58
void CudaRobustPlanner::Initialize(const PsModel* model, const PsTask* task, int device_ordinal) {
  delegate_->Initialize(model, task, device_ordinal);
  model_ = model; task_ = *task;
  nrepetitions_ = (int)Numeric(5, model, "robust_repetitions");
  // if robust_candidates is not defined, it follows from the number of rollouts of the sampling configuration
  ncandidates_ = (int)Numeric(-1, model, "robust_candidates");
  if (ncandidates_ == -1) ncandidates_ = (int)Numeric(10, model, "sampling_trajectories") / std::max(nrepetitions_, 1);
  xfrc_std_ = Numeric(0.1, model, "robust_xfrc");
  xfrc_rate_ = Numeric(0.1, model, "robust_xfrc_rate");
  if (engine_) { ps_destroy(engine_); engine_ = nullptr; }
  engine_ = ps_create(device_ordinal);
  if (!engine_) Fail("ps_create");
  if (ps_set_engine_flavour(engine_, engine_flavour_)) Fail("ps_set_engine_flavour");
  if (ps_set_model(engine_, model)) Fail("ps_set_model");
  if (ps_set_task(engine_, task)) Fail("ps_set_task");
}

// robust_planner.cc:60-72 (trajectory storage lives on the device: ps_rollout_candidates sizes it)
void CudaRobustPlanner::Allocate() {
  delegate_->Allocate();
  state_.assign(model_->nq + model_->nv + model_->na, 0.0);
  mocap_.assign(7 * model_->nmocap, 0.0);
  userdata_.assign(model_->nuserdata, 0.0);
}

// robust_planner.cc:74-85
void CudaRobustPlanner::Reset(int horizon, const double* initial_repeated_action) {
  delegate_->Reset(horizon, initial_repeated_action);
  std::fill(state_.begin(), state_.end(), 0.0); std::fill(mocap_.begin(), mocap_.end(), 0.0);
  std::fill(userdata_.begin(), userdata_.end(), 0.0); time_ = 0.0;
  best_candidate = -1; candidate_score.clear(); noisy_return.clear(); noisy_failure.clear();
}

// robust_planner.cc:87-90
void CudaRobustPlanner::SetState(const double* s, const double* mo, const double* ud, double t) {
  delegate_->SetState(s, mo, ud, t);
  std::copy(s, s + state_.size(), state_.begin());
  if (mo) std::copy(mo, mo + mocap_.size(), mocap_.begin());
  if (ud) std::copy(ud, ud + userdata_.size(), userdata_.begin());
  time_ = t;
}

void CudaRobustPlanner::SetTask(const PsTask* task) {
  task_ = *task;
  delegate_->SetTask(task);
  if (ps_set_task(engine_, task)) Fail("ps_set_task");
}

void CudaRobustPlanner::SetForceNoise(const double* z, int rollouts, int horizon) {
  z_.clear(); z_rollouts_ = z_horizon_ = 0;
  if (!z) return;
  z_.assign(z, z + (size_t)rollouts * (horizon - 1) * 6 * model_->nbody);
  z_rollouts_ = rollouts; z_horizon_ = horizon;
}

void CudaRobustPlanner::SetSeed(unsigned long long seed) {
  delegate_->SetSeed(seed);
  ps_set_seed(engine_, seed ^ 0x9e3779b97f4a7c15ull);      // force noise must not replay the knot noise stream
}

// robust_planner.cc:92-157
void CudaRobustPlanner::OptimizePolicy(int horizon) {
  // the best candidates of the delegate's rollouts (no resampling of the nominal: the reference calls
  // OptimizePolicyCandidates, not OptimizePolicy)
  const int ncandidates = delegate_->OptimizePolicyCandidates(ncandidates_, horizon);
  best_candidate = -1; candidate_score.clear(); noisy_return.clear(); noisy_failure.clear();
  if (!ncandidates) return;
  if (ncandidates == 1) {        // a single candidate is taken as it is
    delegate_->CopyCandidateToPolicy(0);
    best_candidate = 0;
    return;
  }
  const auto start = std::chrono::steady_clock::now();
  const int repetitions = nrepetitions_;
  const int nu = model_->nu;
  const TimeSpline& first = delegate_->CandidatePolicy(0).plan;
  const int P = first.Size();
  std::vector<double> knots((size_t)ncandidates * P * nu);
  for (int c = 0; c < ncandidates; c++) {
    const TimeSpline& plan = delegate_->CandidatePolicy(c).plan;
    std::memcpy(knots.data() + (size_t)c * P * nu, plan.values().data(), sizeof(double) * P * nu);
  }
  if (ps_set_options(engine_, model_->opt_timestep, model_->opt_integrator)) Fail("ps_set_options");
  if (ps_set_state(engine_, state_.data(), mocap_.data(), userdata_.data(), time_)) Fail("ps_set_state");
  // knot times and interpolation of the candidates (they all share the nominal's)
  if (ps_set_nominal(engine_, first.times().data(), knots.data(), P, (int)first.Interpolation())) Fail("ps_set_nominal");
  const int total = ncandidates * repetitions;
  noisy_return.assign(total, 0.0); noisy_failure.assign(total, 0);
  const double* z = nullptr;
  if (!z_.empty()) {
    if (z_rollouts_ < total || z_horizon_ != horizon) { std::fprintf(stderr, "CudaRobustPlanner: injected force noise is [%d][%d-1], need [%d][%d-1]\n", z_rollouts_, z_horizon_, total, horizon); std::abort(); }
    z = z_.data();
  }
  if (ps_rollout_candidates(engine_, knots.data(), ncandidates, repetitions, horizon, xfrc_std_, xfrc_rate_, z,
                            noisy_return.data(), noisy_failure.data())) Fail("ps_rollout_candidates");

  // every candidate's score is the running mean of its noise-free return and its valid noisy returns; the
  // lowest mean wins (strict <: ties keep the better-ranked candidate)
  double best_score = 0;
  candidate_score.assign(ncandidates, 0.0);
  for (int candidate = 0; candidate < ncandidates; candidate++) {
    double mean_return = delegate_->CandidateScore(candidate);
    int valid_rollouts = 0;
    for (int j = 0; j < repetitions; j++) {
      if (noisy_failure[repetitions * candidate + j] & 1) continue;   // a failed rollout does not touch the score (bit 1 = capacity note)
      const double total_return = noisy_return[repetitions * candidate + j];
      mean_return = (valid_rollouts * mean_return + total_return) / (valid_rollouts + 1);
      valid_rollouts++;
    }
    candidate_score[candidate] = mean_return;
    if (best_candidate == -1 || mean_return < best_score) { best_candidate = candidate; best_score = mean_return; }
  }
  noisy_rollouts_compute_time = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - start).count();
  delegate_->CopyCandidateToPolicy(best_candidate);
}

}  // namespace mjpc_b200
