// ORACLE (test infrastructure).  Planner-side restatement of mjpc::SamplingPlanner /
// Trajectory / SamplingPolicy with noise taken as an INPUT tensor (the reference draws it from a
// fresh absl::BitGen per call, mjpc/planners/sampling/planner.cc:318, so it cannot be replayed).
#ifndef ORACLE_MM_PLANNER_H_
#define ORACLE_MM_PLANNER_H_
#include <vector>
#include "../include/ps_engine.h"
#include "mm_engine.h"

namespace mm {

void splineSample(const double* times, const double* knots, int P, int dim, int interp, double t, double* out);
double normValue(const double* x, const double* params, int n, int type);
void costTerms(const PsTask* t, const double* residual, double* terms, bool weighted);
double costValue(const PsTask* t, const double* residual);
void residual(const Model* m, const Data* d, const PsTask* t, double* r);

struct Trajectory {           // mjpc/trajectory.h
  int horizon = 0; bool failure = false; double total_return = 0;
  std::vector<double> states, actions, times, residual, costs, trace;
};

struct Policy {               // SamplingPolicy: TimeSpline plan (times + knots[P][nu])
  std::vector<double> times, knots; int interp = PS_SPLINE_ZERO;
  int P() const { return (int)times.size(); }
};

struct Planner {
  const Model* model = nullptr; PsTask task;
  std::vector<double> state, mocap, userdata; double time = 0;
  Policy policy, previous_policy;
  std::vector<Policy> candidate;
  std::vector<Trajectory> trajectory;
  int winner = 0; double improvement = 0;
  int num_spline_points = 0, interpolation = PS_SPLINE_CUBIC;
  int sliding_plan = 0;                      // sampling_sliding_plan (planner.cc:65, :244-281)
  double noise_exploration[2] = {0.1, 0};
  std::vector<Data> data;     // one per worker thread
  // CrossEntropyPlanner state (mjpc/planners/cross_entropy/planner.h): per-parameter variance, elites
  std::vector<double> variance; double std_min = 0.1; int n_elite = 2;
  std::vector<int> trajectory_order; Trajectory nominal_trajectory; double avg_return = 0;
  // RobustPlanner (mjpc/planners/robust/robust_planner.h): re-rolled candidates, their mean scores, the pick
  std::vector<Trajectory> robust_trajectory; std::vector<double> robust_score; int robust_best = 0;
};

// SamplingPolicy::Action (policy.cc:52-59): spline sample then clamp to ctrlrange
void policyAction(const Model* m, const Policy& p, double t, double* action);
// Trajectory::NoisyRollout + UpdateReturn (trajectory.cc:100-210,312-326), xfrc noise off
// xfrc_w (optional): standard-normal tensor [steps-1][6*nbody] driving the Ornstein-Uhlenbeck force
// perturbation of trajectory.cc:147-155 (the reference draws it from absl::BitGen)
void rollout(const Model* m, const PsTask* task, Data* d, const Policy& p, const double* state, double time,
             const double* mocap, const double* userdata, int steps, Trajectory* out,
             const double* xfrc_w = nullptr, double xfrc_std = 0, double xfrc_rate = 1);

// SamplingPlanner::OptimizePolicy (resample = true) / OptimizePolicyCandidates + CopyCandidateToPolicy(0) (false)
void optimizePolicy(Planner* pl, int N, int horizon, const double* z, const unsigned char* use_std2, int nthread, bool resample = true);

}  // namespace mm
#endif
