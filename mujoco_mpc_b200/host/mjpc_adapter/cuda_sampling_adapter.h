// mjpc::CudaSamplingAdapter -- the reference-side binding of the B200 Predictive Sampling engine.
//
// This is the file a maintainer adds INSIDE the MJPC tree (e.g. mjpc/planners/cuda_sampling/planner.h): it derives from
// mjpc::RankedPlanner (mjpc/planners/planner.h:84-102), so Agent, the GUI and the gRPC services keep calling the
// mjpc::Planner virtuals (planner.h:32-82) and get the CUDA engine behind them.  It compiles against the REAL
// <mujoco/mujoco.h> + MJPC headers; in this repository it is compiled and driven through the mjpc::Planner* interface
// against the reference's own headers with the oracle's mujoco.h build shim (tests/test_mjpc_adapter.py), on the host
// stand-in of the engine ABI (CPU) and on the CUDA library (GPU).
//
// Registration (mjpc/planners/include.cc:41-54):
//     - planners.emplace_back(new mjpc::SamplingPlanner);
//     + planners.emplace_back(new mjpc::CudaSamplingAdapter);
#ifndef MJPC_PLANNERS_CUDA_SAMPLING_ADAPTER_H_
#define MJPC_PLANNERS_CUDA_SAMPLING_ADAPTER_H_
#include <functional>
#include <vector>
#include <mujoco/mujoco.h>
#include "mjpc/planners/planner.h"
#include "mjpc/states/state.h"
#include "mjpc/task.h"
#include "mjpc/trajectory.h"
#include "../cuda_sampling_planner.h"

namespace mjpc {

// mjModel -> PsModel (field for field; mjtByte arrays widened into `scratch`)
void FillPsModel(const mjModel* m, PsModel* p, std::vector<std::vector<int>>* scratch);
// Task cost specification + parameters -> PsTask (the Task::Residual() snapshot, mjpc/task.cc:112-128); resolves the
// SensorByName lookups of the task's residual once (mjpc/utilities.cc:166-175).  Returns false for a task whose residual
// the engine does not carry.
bool FillPsTask(const mjModel* m, const Task& task, PsTask* p);

class CudaSamplingAdapter : public RankedPlanner {
 public:
  CudaSamplingAdapter() = default;
  ~CudaSamplingAdapter() override = default;

  // ---- mjpc::Planner
  void Initialize(mjModel* model, const Task& task) override;
  void Allocate() override;
  void Reset(int horizon, const double* initial_repeated_action = nullptr) override;
  void SetState(const State& state) override;
  void OptimizePolicy(int horizon, ThreadPool& pool) override;       // the pool is not used: rollouts are one kernel launch
  void NominalTrajectory(int horizon, ThreadPool& pool) override;
  void ActionFromPolicy(double* action, const double* state, double time, bool use_previous = false) override;
  const Trajectory* BestTrajectory() override;
  void Traces(mjvScene* scn) override {}                             // drawing: out of scope (no graphics stack)
  void GUI(mjUI& ui) override {}
  void Plots(mjvFigure*, mjvFigure*, int, int, int, int*) override {}
  int NumParameters() override { return impl_.NumParameters(); }
  // ---- mjpc::RankedPlanner
  int OptimizePolicyCandidates(int ncandidates, int horizon, ThreadPool& pool) override;
  double CandidateScore(int candidate) const override { return impl_.CandidateScore(candidate); }
  void ActionFromCandidatePolicy(double* action, int candidate, const double* state, double time) override {
    impl_.ActionFromCandidatePolicy(action, candidate, state, time);
  }
  void CopyCandidateToPolicy(int candidate) override { impl_.CopyCandidateToPolicy(candidate); }

  // task state that lives in private members of a task's ResidualFn (the quadruped's gait / mode clock) reaches the engine
  // as PsTask.aux / iaux: the integrator supplies the accessor (a friend of that task class)
  std::function<void(const mjModel*, const Task&, PsTask*)> task_state_hook;
  int device_ordinal = 0;
  mjpc_b200::CudaSamplingPlanner& impl() { return impl_; }

 private:
  void Snapshot();                       // model options + task parameters as of this iteration (agent.cc:275-290)
  mjpc_b200::CudaSamplingPlanner impl_;
  PsModel pm_{}; PsTask pt_{}; std::vector<std::vector<int>> widened_;
  mjModel* model_ = nullptr; const Task* task_ = nullptr;
  std::vector<double> state_, mocap_, userdata_; double time_ = 0;
  Trajectory best_;
};

}  // namespace mjpc
#endif  // MJPC_PLANNERS_CUDA_SAMPLING_ADAPTER_H_
