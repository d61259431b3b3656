// TEST INFRASTRUCTURE: drives a planner exclusively through the reference's mjpc::Planner* virtuals
// (mjpc/planners/planner.h:32-82) -- Initialize / Allocate / Reset / SetState / OptimizePolicy / ActionFromPolicy /
// BestTrajectory -- with the reference's own Task, State and ThreadPool objects.  kind 0 = mjpc::CudaSamplingAdapter (this
// repo's binding, mujoco_mpc_b200/host/mjpc_adapter/), kind 1 = the reference's mjpc::SamplingPlanner.  The same call
// sequence on both must give the same policy / winner / best trajectory under the same injected knot noise.  Built against
// the reference headers with the oracle's mujoco.h shim; linked with oracle/_ref/libmjpc_ref.so (the reference's objects) and
// either the host stand-in of the engine ABI (CPU) or libps_engine.so (GPU).
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <absl/random/random.h>
#include <mujoco/mujoco.h>
#include "mjpc/planners/planner.h"
#include "mjpc/planners/sampling/planner.h"
#include "mjpc/states/state.h"
#include "mjpc/task.h"
#include "mjpc/tasks/acrobot/acrobot.h"
#include "mjpc/tasks/cartpole/cartpole.h"
#include "mjpc/tasks/particle/particle.h"
#include "mjpc/tasks/swimmer/swimmer.h"
#include "mjpc/tasks/walker/walker.h"
#include "mjpc/threadpool.h"
#include "../../mujoco_mpc_b200/host/mjpc_adapter/cuda_sampling_adapter.h"

extern "C" mjModel* ref_make_model(const PsModel* p);

namespace {
struct Handle {
  mjModel* model = nullptr; std::unique_ptr<mjpc::Task> task; std::unique_ptr<mjpc::Planner> planner;   // base-class pointer only
  mjpc::State state; std::unique_ptr<mjpc::ThreadPool> pool; int kind = 0;
};
std::unique_ptr<mjpc::ResidualFn> g_residual;
void residual_callback(const mjModel* model, mjData* data, int stage) {
  if (stage == mjSTAGE_ACC && g_residual) g_residual->Residual(model, data, data->sensordata);
}
mjpc::Task* make_task(const std::string& n) {
  if (n == "Cartpole") return new mjpc::Cartpole();
  if (n == "Acrobot") return new mjpc::Acrobot();
  if (n == "Particle") return new mjpc::Particle();
  if (n == "Swimmer") return new mjpc::Swimmer();
  if (n == "Walker") return new mjpc::Walker();
  return nullptr;
}
}  // namespace

extern "C" {
void* adp_create(const PsModel* m, const char* task_name, int kind, int device) {
  std::unique_ptr<Handle> h(new Handle());
  h->task.reset(make_task(task_name));
  if (!h->task) return nullptr;
  h->kind = kind;
  h->model = ref_make_model(m);
  h->task->Reset(h->model);
  if (kind == 0) { auto* a = new mjpc::CudaSamplingAdapter(); a->device_ordinal = device; h->planner.reset(a); }
  else h->planner.reset(new mjpc::SamplingPlanner());
  h->planner->Initialize(h->model, *h->task);
  h->planner->Allocate();
  h->planner->Reset(mjpc::kMaxTrajectoryHorizon);
  h->state.Initialize(h->model); h->state.Allocate(h->model);
  h->pool.reset(new mjpc::ThreadPool(1));
  return h.release();
}
void adp_destroy(void* hp) { Handle* h = (Handle*)hp; if (!h) return; h->planner.reset(); h->pool.reset(); mj_deleteModel(h->model); delete h; }

/* the GUI's knobs: public members of either concrete class */
void adp_configure(void* hp, int P, int interp, double sigma0, double sigma1, int N) {
  Handle* h = (Handle*)hp;
  if (auto* a = dynamic_cast<mjpc::CudaSamplingAdapter*>(h->planner.get())) {
    auto& p = a->impl(); p.SetNumSplinePoints(P); p.interpolation_ = interp; p.noise_exploration[0] = sigma0; p.noise_exploration[1] = sigma1; p.num_trajectory_ = N;
  } else if (auto* s = dynamic_cast<mjpc::SamplingPlanner*>(h->planner.get())) {
    s->policy.num_spline_points = P; s->previous_policy.num_spline_points = P; s->interpolation_ = (mjpc::spline::SplineInterpolation)interp;
    s->noise_exploration[0] = sigma0; s->noise_exploration[1] = sigma1; s->num_trajectory_ = N;
  }
}
/* one planning iteration through the base-class interface */
void adp_plan(void* hp, const double* state, const double* mocap, double time, const double* z, const unsigned char* std2, int N, int P, int H) {
  Handle* h = (Handle*)hp; const mjModel* m = h->model;
  std::vector<double> mp(3 * m->nmocap + 1, 0.0), mq(4 * m->nmocap + 1, 0.0), ud(m->nuserdata + 1, 0.0);
  for (int i = 0; i < m->nmocap && mocap; i++) { std::memcpy(&mp[3 * i], mocap + 7 * i, 24); std::memcpy(&mq[4 * i], mocap + 7 * i + 3, 32); }
  h->state.Set(m, state, state + m->nq, state + m->nq + m->nv, mp.data(), mq.data(), ud.data(), time);
  h->planner->SetState(h->state);
  if (h->kind == 0) dynamic_cast<mjpc::CudaSamplingAdapter*>(h->planner.get())->impl().SetNoise(z, std2, N, P);
  else { g_residual = h->task->Residual(); mjcb_sensor = residual_callback; ref_shim_random::begin_replay(z, std2, N, (long long)P * m->nu); }
  h->planner->OptimizePolicy(H, *h->pool);
  ref_shim_random::end_replay();
}
void adp_action(void* hp, double t, int use_previous, double* out) { ((Handle*)hp)->planner->ActionFromPolicy(out, nullptr, t, use_previous != 0); }
int adp_num_parameters(void* hp) { return ((Handle*)hp)->planner->NumParameters(); }
/* BestTrajectory() through the base class; returns the horizon, or -1 when null */
int adp_best(void* hp, double* states, double* actions, double* times, double* residual, double* costs, double* total_return) {
  const mjpc::Trajectory* t = ((Handle*)hp)->planner->BestTrajectory();
  if (!t) return -1;
  const int H = t->horizon;
  std::memcpy(states, t->states.data(), sizeof(double) * H * t->dim_state); std::memcpy(actions, t->actions.data(), sizeof(double) * H * t->dim_action);
  std::memcpy(times, t->times.data(), sizeof(double) * H); std::memcpy(residual, t->residual.data(), sizeof(double) * H * t->dim_residual);
  std::memcpy(costs, t->costs.data(), sizeof(double) * H); *total_return = t->total_return;
  return H;
}
/* RankedPlanner: OptimizePolicyCandidates + CandidateScore through the interface */
int adp_candidates(void* hp, int ncand, int H, double* scores) {
  Handle* h = (Handle*)hp; auto* r = dynamic_cast<mjpc::RankedPlanner*>(h->planner.get());
  if (!r) return -1;
  if (h->kind != 0) { g_residual = h->task->Residual(); mjcb_sensor = residual_callback; }
  const int n = r->OptimizePolicyCandidates(ncand, H, *h->pool);
  for (int i = 0; i < n; i++) scores[i] = r->CandidateScore(i);
  return n;
}
}  // extern "C"
