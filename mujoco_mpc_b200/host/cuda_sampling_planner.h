// CudaSamplingPlanner: host-side drop-in for mjpc::SamplingPlanner whose planning iteration runs on
// the B200 engine behind the C ABI (include/ps_engine.h).
//
// Mirrors the reference's public surface (mjpc/planners/sampling/planner.h:41-165 and the
// mjpc::Planner / RankedPlanner virtuals, mjpc/planners/planner.h:32-102): same method names,
// argument meaning, public data members and error behaviour.  MuJoCo and the rest of MJPC are not
// available in this build environment, so the model / task / state arrive as the C-ABI PODs
// (PsModel mirrors mjModel field for field; PsTask is the Task::Residual() snapshot); the adapter a
// maintainer adds inside MJPC to derive this from mjpc::RankedPlanner is shown in INTEGRATION.md.
// The ThreadPool& argument of the reference signatures is accepted nowhere: rollouts are a kernel
// launch, not pool jobs.
#ifndef MJPC_B200_CUDA_SAMPLING_PLANNER_H_
#define MJPC_B200_CUDA_SAMPLING_PLANNER_H_
#include <shared_mutex>
#include <vector>
#include "../../include/ps_engine.h"
#include "time_spline.h"

namespace mjpc_b200 {

inline constexpr int kMaxTrajectoryHorizon = 512;   // mjpc/trajectory.h:27
inline constexpr int kMaxSplinePoints = 36;         // engine cap == GUI cap (sampling/planner.h:36)

// mjpc::SamplingPolicy (planners/sampling/policy.h): a TimeSpline plan + clamp to ctrlrange
struct SamplingPolicy {
  TimeSpline plan; int num_spline_points = 0; const PsModel* model = nullptr;
  void Action(double* action, const double* state, double time) const;   // policy.cc:52-59
};

// mjpc::Trajectory buffers of ONE rollout (mjpc/trajectory.h), row-major by time step
struct Trajectory {
  int horizon = 0, dim_state = 0, dim_action = 0, dim_residual = 0, dim_trace = 0;
  bool failure = false; double total_return = 0;
  std::vector<double> states, actions, times, residual, costs, trace;
};

// cross-rank winner exchange for multi-GPU runs: gathers `count` doubles from every rank into
// all[world][count] (e.g. ncclAllGather + stream sync, or MPI_Allgather); returns 0 on success
typedef int (*AllGatherFn)(void* ctx, const double* mine, double* all, int count);

class CudaSamplingPlanner {
 public:
  CudaSamplingPlanner() = default;
  ~CudaSamplingPlanner();
  CudaSamplingPlanner(const CudaSamplingPlanner&) = delete;

  // ---- mjpc::Planner interface
  void Initialize(const PsModel* model, const PsTask* task, int device_ordinal = 0);   // planner.cc:40-75
  void Allocate();                                                                    // :78-103
  void Reset(int horizon, const double* initial_repeated_action = nullptr);           // :106-143
  void SetState(const double* state, const double* mocap, const double* userdata, double time);   // :146-149
  void SetTask(const PsTask* task);           // per-iteration Task::Residual() snapshot (agent.cc:290)
  void OptimizePolicy(int horizon);                                                   // :190-208
  void NominalTrajectory(int horizon);                                                // :211-223
  void ActionFromPolicy(double* action, const double* state, double time, bool use_previous = false);   // :225-233
  const Trajectory* BestTrajectory();                                                 // :383-385
  int NumParameters() const { return policy.num_spline_points * model_->nu; }
  // ---- mjpc::RankedPlanner interface
  int OptimizePolicyCandidates(int ncandidates, int horizon);                         // :151-187
  double CandidateScore(int candidate) const;
  void ActionFromCandidatePolicy(double* action, int candidate, const double* state, double time);
  void CopyCandidateToPolicy(int candidate);                                          // :525-534

  // ---- engine-specific controls
  // injected N(0,1) tensor z[N][P][nu] (parity runs); P = knots per sample in z (< 0: policy.num_spline_points).
  // The next iteration aborts unless N >= num_trajectory and P equals the number of nodes of its plan.
  void SetNoise(const double* z, const unsigned char* use_std2, int N, int P = -1);
  void SetSeed(unsigned long long seed);
  // the GUI's spline-point slider (planner.cc:439-446): policy, previous_policy and the winner the next
  // UpdateNominalPolicy resamples from all take the new count at once
  void SetNumSplinePoints(int n) { policy.num_spline_points = n; previous_policy.num_spline_points = n; candidate_winner_.num_spline_points = n; }
  void SetPolicy(const double* times, const double* knots, int P, int interpolation);   // start from a given nominal (parity runs)
  void SetSharding(int rank, int world, AllGatherFn fn, void* ctx);       // samples [rank*N, (rank+1)*N) of world*N
  void SetShardingNative(int rank, int world, const void* nccl_unique_id);   // same, exchange = the engine's ncclAllGather (ps_comm_init)
  PsEngine* engine() { return engine_; }
  const SamplingPolicy& CandidatePolicy(int candidate) const { return ranked_[candidate]; }   // candidate_policy[trajectory_order[c]]
  const PsModel* model() const { return model_; }
  const PsTask& task() const { return task_; }
  int device() const { return device_; }

  // ---- public members the reference exposes (sampling/planner.h:115-162)
  SamplingPolicy policy, previous_policy;
  double noise_exploration[2] = {0.1, 0.0};
  int winner = -1;
  int num_trajectory_ = 10;
  int interpolation_ = kCubicSpline;
  int sliding_plan_ = 0;
  double improvement = 0;
  double noise_compute_time = 0, rollouts_compute_time = 0, policy_update_compute_time = 0;   // microseconds
  std::vector<double> state, mocap, userdata; double time = 0;

 private:
  void UpdateNominalPolicy(int horizon);                                              // :236-310
  void Fail(const char* what) const;
  const PsModel* model_ = nullptr; PsTask task_{}; PsEngine* engine_ = nullptr; int device_ = 0;
  SamplingPolicy candidate_winner_;          // candidate_policy[winner] of the last iteration
  std::vector<SamplingPolicy> ranked_;       // candidate policies of the last OptimizePolicyCandidates, best first
  std::vector<double> ranked_score_; std::vector<int> trajectory_order;
  std::vector<int> ranked_local_;            // local sample index of each candidate (-1: lives on another rank)
  Trajectory best_; bool best_valid_ = false; int last_horizon_ = 0;
  double nominal_return_ = 0, winner_return_ = 0;
  int rank_ = 0, world_ = 1; AllGatherFn gather_ = nullptr; void* gather_ctx_ = nullptr; bool native_comm_ = false;
  bool winner_is_local_ = true; int local_winner_ = 0;
  std::vector<double> noise_z_; std::vector<unsigned char> noise_std2_; int noise_n_ = 0; bool noise_pending_ = false;
  mutable std::shared_mutex mtx_;
};

}  // namespace mjpc_b200
#endif
