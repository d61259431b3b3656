// extern "C" surface over CudaSamplingPlanner for language bindings (tests/, bench.py use it via
// ctypes).  House style follows the reference's own C plugin ABI, mjpc/interface.cc:77-115.
#include <algorithm>
#include <cstring>
#include <memory>
#include "cuda_sampling_planner.h"

using mjpc_b200::CudaSamplingPlanner;

extern "C" {

// pure host logic of the cross-rank min-loc (lowest return, ties -> lowest global sample index);
// records are [world][count] with record = [return, global index, nominal return, knots...]
int psp_select_winner(const double* all, int world, int count) {
  int best = 0;
  for (int r = 1; r < world; r++) {
    const double c = all[(size_t)r * count], bc = all[(size_t)best * count];
    if (c < bc || (c == bc && all[(size_t)r * count + 1] < all[(size_t)best * count + 1])) best = r;
  }
  return best;
}

void* psp_create(const PsModel* model, const PsTask* task, int device) {
  CudaSamplingPlanner* p = new CudaSamplingPlanner;
  p->Initialize(model, task, device); p->Allocate(); p->Reset(mjpc_b200::kMaxTrajectoryHorizon);
  return p;
}
void psp_destroy(void* h) { delete (CudaSamplingPlanner*)h; }
void psp_reset(void* h, int horizon, const double* initial_repeated_action) { ((CudaSamplingPlanner*)h)->Reset(horizon, initial_repeated_action); }
void psp_set_state(void* h, const double* state, const double* mocap, const double* userdata, double time) {
  ((CudaSamplingPlanner*)h)->SetState(state, mocap, userdata, time);
}
void psp_set_task(void* h, const PsTask* t) { ((CudaSamplingPlanner*)h)->SetTask(t); }
void psp_configure(void* h, int num_trajectory, int interpolation, int num_spline_points, double sigma0, double sigma1, int sliding) {
  CudaSamplingPlanner* p = (CudaSamplingPlanner*)h;
  if (num_trajectory > 0) p->num_trajectory_ = num_trajectory;
  if (interpolation >= 0) p->interpolation_ = interpolation;
  if (num_spline_points > 0) p->SetNumSplinePoints(num_spline_points);
  if (sigma0 >= 0) p->noise_exploration[0] = sigma0;
  if (sigma1 >= 0) p->noise_exploration[1] = sigma1;
  if (sliding >= 0) p->sliding_plan_ = sliding;
}
void psp_set_noise(void* h, const double* z, const unsigned char* use_std2, int N, int P) { ((CudaSamplingPlanner*)h)->SetNoise(z, use_std2, N, P); }
void psp_set_seed(void* h, unsigned long long seed) { ((CudaSamplingPlanner*)h)->SetSeed(seed); }
void psp_set_sharding(void* h, int rank, int world, mjpc_b200::AllGatherFn fn, void* ctx) { ((CudaSamplingPlanner*)h)->SetSharding(rank, world, fn, ctx); }
void psp_set_sharding_native(void* h, int rank, int world, const void* nccl_unique_id) { ((CudaSamplingPlanner*)h)->SetShardingNative(rank, world, nccl_unique_id); }
void psp_optimize_policy(void* h, int horizon) { ((CudaSamplingPlanner*)h)->OptimizePolicy(horizon); }
int psp_optimize_policy_candidates(void* h, int ncandidates, int horizon) { return ((CudaSamplingPlanner*)h)->OptimizePolicyCandidates(ncandidates, horizon); }
double psp_candidate_score(void* h, int c) { return ((CudaSamplingPlanner*)h)->CandidateScore(c); }
void psp_action_from_candidate_policy(void* h, double* action, int c, double time) { ((CudaSamplingPlanner*)h)->ActionFromCandidatePolicy(action, c, nullptr, time); }
void psp_copy_candidate_to_policy(void* h, int c) { ((CudaSamplingPlanner*)h)->CopyCandidateToPolicy(c); }
void psp_nominal_trajectory(void* h, int horizon) { ((CudaSamplingPlanner*)h)->NominalTrajectory(horizon); }
void psp_action_from_policy(void* h, double* action, double time, int use_previous) {
  ((CudaSamplingPlanner*)h)->ActionFromPolicy(action, nullptr, time, use_previous != 0);
}
int psp_winner(void* h) { return ((CudaSamplingPlanner*)h)->winner; }
double psp_improvement(void* h) { return ((CudaSamplingPlanner*)h)->improvement; }
int psp_num_parameters(void* h) { return ((CudaSamplingPlanner*)h)->NumParameters(); }
// which: 0 policy, 1 previous_policy.  Returns the number of spline points written.
int psp_get_policy(void* h, int which, double* times, double* knots) {
  CudaSamplingPlanner* p = (CudaSamplingPlanner*)h;
  const mjpc_b200::TimeSpline& s = which ? p->previous_policy.plan : p->policy.plan;
  std::memcpy(times, s.times().data(), sizeof(double) * s.times().size());
  std::memcpy(knots, s.values().data(), sizeof(double) * s.values().size());
  return s.Size();
}
void psp_timers(void* h, double* noise_us, double* rollouts_us, double* policy_update_us) {
  CudaSamplingPlanner* p = (CudaSamplingPlanner*)h;
  *noise_us = p->noise_compute_time; *rollouts_us = p->rollouts_compute_time; *policy_update_us = p->policy_update_compute_time;
}
// BestTrajectory(): returns horizon (0 if none); any pointer may be NULL
int psp_best_trajectory(void* h, double* states, double* actions, double* times, double* residual, double* costs, double* trace,
                        double* total_return, int* failure) {
  const mjpc_b200::Trajectory* t = ((CudaSamplingPlanner*)h)->BestTrajectory();
  if (!t) return 0;
  if (states) std::memcpy(states, t->states.data(), sizeof(double) * t->states.size());
  if (actions) std::memcpy(actions, t->actions.data(), sizeof(double) * t->actions.size());
  if (times) std::memcpy(times, t->times.data(), sizeof(double) * t->times.size());
  if (residual) std::memcpy(residual, t->residual.data(), sizeof(double) * t->residual.size());
  if (costs) std::memcpy(costs, t->costs.data(), sizeof(double) * t->costs.size());
  if (trace && t->dim_trace) std::memcpy(trace, t->trace.data(), sizeof(double) * (size_t)t->horizon * t->dim_trace);
  if (total_return) *total_return = t->total_return;
  if (failure) *failure = t->failure;
  return t->horizon;
}
void* psp_engine(void* h) { return ((CudaSamplingPlanner*)h)->engine(); }

}  // extern "C"

// ---- CudaCrossEntropyPlanner
#include "cuda_cross_entropy_planner.h"
using mjpc_b200::CudaCrossEntropyPlanner;

extern "C" {

void* pce_create(const PsModel* model, const PsTask* task, int device) {
  CudaCrossEntropyPlanner* p = new CudaCrossEntropyPlanner;
  p->Initialize(model, task, device); p->Allocate(); p->Reset(mjpc_b200::kMaxTrajectoryHorizon);
  return p;
}
void pce_destroy(void* h) { delete (CudaCrossEntropyPlanner*)h; }
void pce_reset(void* h, int horizon, const double* initial_repeated_action) { ((CudaCrossEntropyPlanner*)h)->Reset(horizon, initial_repeated_action); }
void pce_set_state(void* h, const double* state, const double* mocap, const double* userdata, double time) {
  ((CudaCrossEntropyPlanner*)h)->SetState(state, mocap, userdata, time);
}
void pce_set_task(void* h, const PsTask* t) { ((CudaCrossEntropyPlanner*)h)->SetTask(t); }
// negative / non-positive arguments keep the value read from the model's numerics at Initialize
void pce_configure(void* h, int num_trajectory, int n_elite, int interpolation, int num_spline_points, double std_initial, double std_min) {
  CudaCrossEntropyPlanner* p = (CudaCrossEntropyPlanner*)h;
  if (num_trajectory > 0) p->num_trajectory_ = num_trajectory;
  if (n_elite > 0) p->n_elite_ = n_elite;
  if (interpolation >= 0) p->interpolation_ = interpolation;
  if (num_spline_points > 0) for (mjpc_b200::SamplingPolicy* q : {&p->policy, &p->previous_policy, &p->resampled_policy}) q->num_spline_points = num_spline_points;
  if (std_initial >= 0) p->std_initial_ = std_initial;
  if (std_min >= 0) p->std_min_ = std_min;
}
void pce_set_policy(void* h, const double* times, const double* knots, int P, int interpolation) {
  CudaCrossEntropyPlanner* p = (CudaCrossEntropyPlanner*)h;
  mjpc_b200::TimeSpline plan(p->NumParameters() / p->policy.num_spline_points, (mjpc_b200::SplineInterpolation)interpolation);
  const int nu = p->NumParameters() / p->policy.num_spline_points;
  for (int k = 0; k < P; k++) plan.AddNode(times[k], knots + (size_t)k * nu);
  p->policy.plan = plan;
}
void pce_set_variance(void* h, const double* v, int n) { CudaCrossEntropyPlanner* p = (CudaCrossEntropyPlanner*)h; for (int i = 0; i < n && i < (int)p->variance.size(); i++) p->variance[i] = v[i]; }
int pce_get_variance(void* h, double* v, int cap) { CudaCrossEntropyPlanner* p = (CudaCrossEntropyPlanner*)h; int n = cap < (int)p->variance.size() ? cap : (int)p->variance.size(); std::memcpy(v, p->variance.data(), sizeof(double) * n); return n; }
void pce_set_noise(void* h, const double* z, int N, int P) { ((CudaCrossEntropyPlanner*)h)->SetNoise(z, N, P); }
void pce_set_seed(void* h, unsigned long long seed) { ((CudaCrossEntropyPlanner*)h)->SetSeed(seed); }
void pce_optimize_policy(void* h, int horizon) { ((CudaCrossEntropyPlanner*)h)->OptimizePolicy(horizon); }
void pce_nominal_trajectory(void* h, int horizon) { ((CudaCrossEntropyPlanner*)h)->NominalTrajectory(horizon); }
void pce_action_from_policy(void* h, double* action, double time, int use_previous) {
  ((CudaCrossEntropyPlanner*)h)->ActionFromPolicy(action, nullptr, time, use_previous != 0);
}
double pce_improvement(void* h) { return ((CudaCrossEntropyPlanner*)h)->improvement; }
double pce_avg_return(void* h) { return ((CudaCrossEntropyPlanner*)h)->avg_return; }
double pce_best_return(void* h) { return ((CudaCrossEntropyPlanner*)h)->best_return; }
int pce_order(void* h, int* order, int cap) {
  CudaCrossEntropyPlanner* p = (CudaCrossEntropyPlanner*)h;
  int n = cap < (int)p->trajectory_order.size() ? cap : (int)p->trajectory_order.size();
  for (int i = 0; i < n; i++) order[i] = p->trajectory_order[i];
  return n;
}
int pce_get_policy(void* h, int which, double* times, double* knots) {
  CudaCrossEntropyPlanner* p = (CudaCrossEntropyPlanner*)h;
  const mjpc_b200::TimeSpline& s = which == 1 ? p->previous_policy.plan : which == 2 ? p->resampled_policy.plan : p->policy.plan;
  std::memcpy(times, s.times().data(), sizeof(double) * s.times().size());
  std::memcpy(knots, s.values().data(), sizeof(double) * s.values().size());
  return s.Size();
}
int pce_best_trajectory(void* h, double* states, double* actions, double* times, double* residual, double* costs, double* total_return, int* failure) {
  const mjpc_b200::Trajectory* t = ((CudaCrossEntropyPlanner*)h)->BestTrajectory();
  if (!t) return 0;
  if (states) std::memcpy(states, t->states.data(), sizeof(double) * t->states.size());
  if (actions) std::memcpy(actions, t->actions.data(), sizeof(double) * t->actions.size());
  if (times) std::memcpy(times, t->times.data(), sizeof(double) * t->times.size());
  if (residual) std::memcpy(residual, t->residual.data(), sizeof(double) * t->residual.size());
  if (costs) std::memcpy(costs, t->costs.data(), sizeof(double) * t->costs.size());
  if (total_return) *total_return = t->total_return;
  if (failure) *failure = t->failure;
  return t->horizon;
}
void pce_timers(void* h, double* noise_us, double* rollouts_us, double* policy_update_us) {
  CudaCrossEntropyPlanner* p = (CudaCrossEntropyPlanner*)h;
  *noise_us = p->noise_compute_time; *rollouts_us = p->rollouts_compute_time; *policy_update_us = p->policy_update_compute_time;
}
void* pce_engine(void* h) { return ((CudaCrossEntropyPlanner*)h)->engine(); }

}  // extern "C"

// ---- CudaRobustPlanner
#include "cuda_robust_planner.h"
using mjpc_b200::CudaRobustPlanner;

extern "C" {

// parity runs start from a given nominal spline instead of the Reset() one
void psp_set_policy(void* h, const double* times, const double* knots, int P, int interpolation) {
  ((CudaSamplingPlanner*)h)->SetPolicy(times, knots, P, interpolation);
}

void* prb_create(const PsModel* model, const PsTask* task, int device) {
  CudaRobustPlanner* p = new CudaRobustPlanner(std::make_unique<CudaSamplingPlanner>());
  p->Initialize(model, task, device); p->Allocate(); p->Reset(mjpc_b200::kMaxTrajectoryHorizon);
  return p;
}
// flavour of the re-roll kernel: 0 auto (warp-per-rollout where available), 3 dense thread-per-rollout
void* prb_create_flavour(const PsModel* model, const PsTask* task, int device, int flavour) {
  CudaRobustPlanner* p = new CudaRobustPlanner(std::make_unique<CudaSamplingPlanner>());
  p->engine_flavour_ = flavour;
  p->Initialize(model, task, device); p->Allocate(); p->Reset(mjpc_b200::kMaxTrajectoryHorizon);
  return p;
}
void prb_destroy(void* h) { delete (CudaRobustPlanner*)h; }
void* prb_delegate(void* h) { return ((CudaRobustPlanner*)h)->delegate(); }      // a psp_* handle owned by the robust planner
void* prb_engine(void* h) { return ((CudaRobustPlanner*)h)->engine(); }
void prb_reset(void* h, int horizon, const double* initial_repeated_action) { ((CudaRobustPlanner*)h)->Reset(horizon, initial_repeated_action); }
void prb_set_state(void* h, const double* state, const double* mocap, const double* userdata, double time) {
  ((CudaRobustPlanner*)h)->SetState(state, mocap, userdata, time);
}
void prb_set_task(void* h, const PsTask* t) { ((CudaRobustPlanner*)h)->SetTask(t); }
// negative / zero arguments keep the value read from the model's numerics
void prb_configure(void* h, int ncandidates, int nrepetitions, double xfrc_std, double xfrc_rate) {
  CudaRobustPlanner* p = (CudaRobustPlanner*)h;
  if (ncandidates > 0) p->ncandidates_ = ncandidates;
  if (nrepetitions > 0) p->nrepetitions_ = nrepetitions;
  if (xfrc_std >= 0) p->xfrc_std_ = xfrc_std;
  if (xfrc_rate > 0) p->xfrc_rate_ = xfrc_rate;
}
void prb_get_settings(void* h, int* ncandidates, int* nrepetitions, double* xfrc_std, double* xfrc_rate) {
  CudaRobustPlanner* p = (CudaRobustPlanner*)h;
  *ncandidates = p->ncandidates_; *nrepetitions = p->nrepetitions_; *xfrc_std = p->xfrc_std_; *xfrc_rate = p->xfrc_rate_;
}
void prb_set_force_noise(void* h, const double* z, int rollouts, int horizon) { ((CudaRobustPlanner*)h)->SetForceNoise(z, rollouts, horizon); }
void prb_set_seed(void* h, unsigned long long seed) { ((CudaRobustPlanner*)h)->SetSeed(seed); }
void prb_optimize_policy(void* h, int horizon) { ((CudaRobustPlanner*)h)->OptimizePolicy(horizon); }
void prb_nominal_trajectory(void* h, int horizon) { ((CudaRobustPlanner*)h)->NominalTrajectory(horizon); }
void prb_action_from_policy(void* h, double* action, double time, int use_previous) {
  ((CudaRobustPlanner*)h)->ActionFromPolicy(action, nullptr, time, use_previous != 0);
}
int prb_best_candidate(void* h) { return ((CudaRobustPlanner*)h)->best_candidate; }
// scores[ncandidates], noisy returns / failure [ncandidates][nrepetitions]; returns the number of candidates scored
int prb_results(void* h, double* scores, double* noisy_return, unsigned char* noisy_failure, int cap_candidates) {
  CudaRobustPlanner* p = (CudaRobustPlanner*)h;
  const int n = std::min((int)p->candidate_score.size(), cap_candidates);
  const int rep = n ? (int)(p->noisy_return.size() / p->candidate_score.size()) : 0;
  if (scores) std::memcpy(scores, p->candidate_score.data(), sizeof(double) * n);
  if (noisy_return) std::memcpy(noisy_return, p->noisy_return.data(), sizeof(double) * n * rep);
  if (noisy_failure) std::memcpy(noisy_failure, p->noisy_failure.data(), (size_t)n * rep);
  return n;
}
double prb_noisy_rollouts_us(void* h) { return ((CudaRobustPlanner*)h)->noisy_rollouts_compute_time; }

}  // extern "C"
