// CudaRobustPlanner: host-side drop-in for mjpc::RobustPlanner
// (mjpc/planners/robust/robust_planner.h:36-92) over the B200 rollout engine.  The delegate is a
// CudaSamplingPlanner (the reference wraps any RankedPlanner; SamplingPlanner is the one it ships,
// planners/include.cc); its best `ncandidates_` policies are re-rolled `nrepetitions_` times each under
// Ornstein-Uhlenbeck force perturbations on every body (Trajectory::NoisyRollout,
// mjpc/trajectory.cc:100-210) by ONE launch of the force-perturbed rollout kernel (coop_robust_kernel, one warp per
// rollout, or robust_rollout_kernel, one thread per rollout) on a second engine handle, and the candidate with the lowest running-mean return is
// copied into the policy (robust_planner.cc:132-156).  SURVEY.md section 8(f) row 2.
#ifndef MJPC_B200_CUDA_ROBUST_PLANNER_H_
#define MJPC_B200_CUDA_ROBUST_PLANNER_H_
#include <memory>
#include <vector>
#include "cuda_sampling_planner.h"

namespace mjpc_b200 {

class CudaRobustPlanner {
 public:
  explicit CudaRobustPlanner(std::unique_ptr<CudaSamplingPlanner> delegate) : delegate_(std::move(delegate)) {}
  ~CudaRobustPlanner();
  CudaRobustPlanner(const CudaRobustPlanner&) = delete;

  // ---- mjpc::Planner interface
  void Initialize(const PsModel* model, const PsTask* task, int device_ordinal = 0);   // robust_planner.cc:31-58
  void Allocate();                                                                    // :60-72
  void Reset(int horizon, const double* initial_repeated_action = nullptr);           // :74-85
  void SetState(const double* state, const double* mocap, const double* userdata, double time);   // :87-90
  void SetTask(const PsTask* task);
  void OptimizePolicy(int horizon);                                                   // :92-157
  void NominalTrajectory(int horizon) { delegate_->NominalTrajectory(horizon); }      // :159-161
  void ActionFromPolicy(double* action, const double* state, double time, bool use_previous = false) {
    delegate_->ActionFromPolicy(action, state, time, use_previous);                   // :162-165
  }
  const Trajectory* BestTrajectory() { return delegate_->BestTrajectory(); }          // :166-168
  int NumParameters() const { return delegate_->NumParameters(); }

  // ---- engine-specific controls
  void SetForceNoise(const double* z, int rollouts, int horizon);   // injected N(0,1) [rollouts][horizon-1][6 nbody] (parity runs)
  void SetSeed(unsigned long long seed);
  CudaSamplingPlanner* delegate() { return delegate_.get(); }
  PsEngine* engine() { return engine_; }

  // ---- settings (private in the reference, set from the model's numerics / its GUI sliders)
  int ncandidates_ = 12, nrepetitions_ = 5;
  // kernel of the re-rolls, read by Initialize(): 0 = warp-per-rollout where the model class has one (walker /
  // quadruped on the Euler integrator), else dense thread-per-rollout; 3 = always dense (ps_set_engine_flavour)
  int engine_flavour_ = 0;
  double xfrc_std_ = 0.1, xfrc_rate_ = 0.1;

  // ---- results of the last OptimizePolicy
  int best_candidate = -1;                    // index into the delegate's ranking
  std::vector<double> candidate_score;        // running-mean score per candidate
  std::vector<double> noisy_return;           // [ncandidates][nrepetitions]
  std::vector<unsigned char> noisy_failure;
  double noisy_rollouts_compute_time = 0;     // microseconds

 private:
  void Fail(const char* what) const;
  const PsModel* model_ = nullptr; PsTask task_{};
  std::unique_ptr<CudaSamplingPlanner> delegate_;
  PsEngine* engine_ = nullptr;                // second handle: dense flavour, carries xfrc_applied
  std::vector<double> state_, mocap_, userdata_; double time_ = 0;
  std::vector<double> z_; int z_rollouts_ = 0, z_horizon_ = 0;
};

}  // namespace mjpc_b200
#endif
