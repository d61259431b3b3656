// CudaCrossEntropyPlanner: host-side drop-in for mjpc::CrossEntropyPlanner
// (mjpc/planners/cross_entropy/planner.h:36-165) on the same B200 rollout engine as
// CudaSamplingPlanner.  Rollouts() is the identical kernel pipeline (K1 perturb in its
// per-parameter-std mode, K2 rollouts); what differs is the update: the candidates are ranked on the
// device (K4 rank_kernel) and the elites reduced to the new nominal knots and variance (K5
// elite_stats_kernel) instead of the argmin.  Same method names, public members and error behaviour
// as the reference; SURVEY.md section 8(f) row 1.
#ifndef MJPC_B200_CUDA_CROSS_ENTROPY_PLANNER_H_
#define MJPC_B200_CUDA_CROSS_ENTROPY_PLANNER_H_
#include <shared_mutex>
#include <vector>
#include "cuda_sampling_planner.h"

namespace mjpc_b200 {

class CudaCrossEntropyPlanner {
 public:
  CudaCrossEntropyPlanner() = default;
  ~CudaCrossEntropyPlanner();
  CudaCrossEntropyPlanner(const CudaCrossEntropyPlanner&) = delete;

  // ---- mjpc::Planner interface
  void Initialize(const PsModel* model, const PsTask* task, int device_ordinal = 0);   // cross_entropy/planner.cc:40-57
  void Allocate();                                                                    // :60-116
  void Reset(int horizon, const double* initial_repeated_action = nullptr);           // :119-156
  void SetState(const double* state, const double* mocap, const double* userdata, double time);   // :159-162
  void SetTask(const PsTask* task);
  void OptimizePolicy(int horizon);                                                   // :165-284
  void NominalTrajectory(int horizon);                                                // :287-301
  void ActionFromPolicy(double* action, const double* state, double time, bool use_previous = false);   // :304-313
  const Trajectory* BestTrajectory();                                                 // :426-428 (the NOMINAL trajectory)
  int NumParameters() const { return policy.num_spline_points * model_->nu; }

  // ---- engine-specific controls
  void SetNoise(const double* z, int N, int P = -1);   // injected N(0,1) tensor [N][P][nu] (parity runs); P < 0: num_spline_points
  void SetSeed(unsigned long long seed);
  PsEngine* engine() { return engine_; }

  // ---- public members the reference exposes (cross_entropy/planner.h:101-160)
  SamplingPolicy policy, previous_policy, resampled_policy;
  std::vector<double> variance;               // per spline parameter, [P][nu]
  std::vector<int> trajectory_order;          // elite candidates, best first
  double std_initial_ = 0.1, std_min_ = 0.1;
  int num_trajectory_ = 10, n_elite_ = 2;
  int interpolation_ = kCubicSpline;
  double improvement = 0, avg_return = 0, best_return = 0, nominal_return = 0;
  double noise_compute_time = 0, rollouts_compute_time = 0, policy_update_compute_time = 0;   // microseconds
  std::vector<double> state, mocap, userdata; double time = 0;

 private:
  void ResamplePolicy(int horizon);                                                   // :316-341
  void Fail(const char* what) const;
  const PsModel* model_ = nullptr; PsTask task_{}; PsEngine* engine_ = nullptr;
  Trajectory nominal_; bool nominal_valid_ = false; int last_horizon_ = 0, last_n_ = 0;
  std::vector<double> noise_z_; int noise_n_ = 0; bool noise_pending_ = false;
  mutable std::shared_mutex mtx_;
};

}  // namespace mjpc_b200
#endif
