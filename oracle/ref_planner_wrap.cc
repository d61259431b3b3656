// ORACLE (test infrastructure): C entry points over the reference's OWN Predictive Sampling planner -- mjpc::SamplingPlanner
// (planners/sampling/planner.cc), SamplingPolicy, Trajectory::NoisyRollout, Task / BaseResidualFn cost code, State, ThreadPool,
// utilities.cc and the six task residuals -- all compiled UNMODIFIED from /root/reference (recipe: oracle/Makefile target
// `ref`, output oracle/_ref/libmjpc_ref.so, git-ignored).  Below mj_step the physics is the oracle's restated MuJoCo step
// (ref_glue.cc).  tools/gen_golden_planner.py drives this to write tests/golden/planner_reference_*.npz, against which
// the restatement (mm_planner.cc / mm_cost.cc / mm_quadruped.cc) and the CUDA engine are checked; bench.py's CPU arms
// time it (cpu_baseline.kind = "reference").
#include <atomic>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include <absl/random/random.h>
#include <mujoco/mujoco.h>
#include "../include/ps_model.h"
#include "mjpc/planners/sampling/planner.h"
#include "mjpc/states/state.h"
#include "mjpc/task.h"
#include "mjpc/tasks/acrobot/acrobot.h"
#include "mjpc/tasks/cartpole/cartpole.h"
#include "mjpc/tasks/particle/particle.h"
#include "mjpc/tasks/swimmer/swimmer.h"
#include "mjpc/tasks/walker/walker.h"
#include "mjpc/threadpool.h"
#include "mjpc/trajectory.h"
// The quadruped's ResidualFn keeps task state (mode, gait phase clock, walk / flip anchors) in private members that
// Transition() manages.  The engine under test receives that state as PsTask.aux, so the harness must be able to READ it
// from the reference object: this one header is included with its access specifiers opened (every header it includes is
// already in; class layout does not depend on access control).  The reference's own translation units are untouched.
#define private public
#define protected public
#include "mjpc/tasks/quadruped/quadruped.h"
#undef private
#undef protected

extern "C" mjModel* ref_make_model(const PsModel* p);

namespace {
struct Handle {
  mjModel* model = nullptr;
  std::unique_ptr<mjpc::Task> task;
  std::unique_ptr<mjpc::SamplingPlanner> planner;
  std::unique_ptr<mjpc::ThreadPool> pool; int pool_threads = 0;
  mjpc::State state;
  mjData* scratch = nullptr;       // Transition / residual probes
};
// MJPC routes the residual through MuJoCo's user-sensor callback.  As in the application (mjpc/app.cc:109-130) the rollout
// threads evaluate a lock-free ResidualFn snapshot taken from the task (Agent::PlanningResidual, Task::Residual() ->
// ResidualLocked); mjpc/testspeed.cc:36-40 goes through the locking Task::Residual instead, which serialises the worker
// threads on the task mutex -- the faster of the reference's two routes is the one timed and compared here.
const mjpc::Task* g_task = nullptr;
std::unique_ptr<mjpc::ResidualFn> g_residual;
void residual_callback(const mjModel* model, mjData* data, int stage) {
  if (stage == mjSTAGE_ACC && g_residual) g_residual->Residual(model, data, data->sensordata);
}
mjpc::Task* make_task(const std::string& n) {
  if (n == "Cartpole") return new mjpc::Cartpole();
  if (n == "Acrobot") return new mjpc::Acrobot();
  if (n == "Particle") return new mjpc::Particle();
  if (n == "ParticleFixed") return new mjpc::ParticleFixed();
  if (n == "Swimmer") return new mjpc::Swimmer();
  if (n == "Walker") return new mjpc::Walker();
  if (n == "Quadruped Flat" || n == "Quadruped") return new mjpc::QuadrupedFlat();
  return nullptr;
}
void activate(Handle* h) { g_task = h->task.get(); g_residual = h->task->Residual(); mjcb_sensor = residual_callback; }
}  // namespace

extern "C" {

void* refp_create(const PsModel* m, const char* task_name) {
  std::unique_ptr<Handle> h(new Handle());
  h->task.reset(make_task(task_name));
  if (!h->task) return nullptr;
  h->model = ref_make_model(m);
  h->task->Reset(h->model);                       // cost specification from the model's user sensors / numerics (task.cc:147-245)
  activate(h.get());
  h->planner.reset(new mjpc::SamplingPlanner());
  h->planner->Initialize(h->model, *h->task);
  h->planner->Allocate();
  h->planner->Reset(mjpc::kMaxTrajectoryHorizon);
  h->state.Initialize(h->model); h->state.Allocate(h->model);
  h->scratch = mj_makeData(h->model);
  return h.release();
}
void refp_destroy(void* hp) {
  Handle* h = (Handle*)hp; if (!h) return;
  if (g_task == h->task.get()) { g_task = nullptr; g_residual.reset(); mjcb_sensor = nullptr; }
  mj_deleteData(h->scratch); h->planner.reset(); h->pool.reset(); mj_deleteModel(h->model); delete h;
}
int refp_max_trajectory(void) { return mjpc::kMaxTrajectory; }

/* mju_error handling of the timed reference arm (ref_shim/mujoco/mujoco.h: mju_shim_error_hook).  on = 1: an error raised
 * inside the reference's code is counted and execution continues, as MuJoCo does when mju_user_error is installed;
 * on = 0 (default, every test): abort.  The count is reported next to the timing so that the reader sees it. */
static std::atomic<long> g_mju_errors{0};
static void count_mju_error(const char*) { g_mju_errors.fetch_add(1, std::memory_order_relaxed); }
void refp_tolerate_errors(int on) { mju_shim_error_hook() = on ? count_mju_error : nullptr; }
long refp_error_count(void) { return g_mju_errors.load(); }
void refp_error_reset(void) { g_mju_errors.store(0); }

/* planner settings the GUI / numerics expose (public members of SamplingPlanner) */
void refp_configure(void* hp, int num_spline_points, int interpolation, double sigma0, double sigma1, int num_trajectory, int sliding_plan) {
  Handle* h = (Handle*)hp; mjpc::SamplingPlanner& p = *h->planner;
  p.policy.num_spline_points = num_spline_points; p.previous_policy.num_spline_points = num_spline_points;
  p.interpolation_ = (mjpc::spline::SplineInterpolation)interpolation;
  p.noise_exploration[0] = sigma0; p.noise_exploration[1] = sigma1;
  p.num_trajectory_ = num_trajectory; p.sliding_plan_ = sliding_plan;
}
void refp_set_state(void* hp, const double* state, const double* mocap, const double* userdata, double time) {
  Handle* h = (Handle*)hp; const mjModel* m = h->model;
  std::vector<double> mp(3 * m->nmocap + 1, 0.0), mq(4 * m->nmocap + 1, 0.0), ud(m->nuserdata + 1, 0.0);
  for (int i = 0; i < m->nmocap; i++) {
    if (mocap) { for (int k = 0; k < 3; k++) mp[3 * i + k] = mocap[7 * i + k]; for (int k = 0; k < 4; k++) mq[4 * i + k] = mocap[7 * i + 3 + k]; }
    else { mq[4 * i] = 1; }
  }
  if (userdata) for (int i = 0; i < m->nuserdata; i++) ud[i] = userdata[i];
  h->state.Set(m, state, state + m->nq, state + m->nq + m->nv, mp.data(), mq.data(), ud.data(), time);
  h->planner->SetState(h->state);
}
/* nominal plan: policy.plan = the given knots (times ascending) */
void refp_set_policy(void* hp, const double* times, const double* knots, int P, int interpolation) {
  Handle* h = (Handle*)hp; mjpc::SamplingPlanner& p = *h->planner; const int nu = h->model->nu;
  mjpc::spline::TimeSpline plan(nu, (mjpc::spline::SplineInterpolation)interpolation);
  plan.Reserve(P);
  for (int i = 0; i < P; i++) plan.AddNode(times[i], absl::MakeConstSpan(knots + (size_t)i * nu, nu));
  p.policy.SetPlan(plan);
}
/* one SamplingPlanner::OptimizePolicy.  z != NULL: replay the injected standard-normal knot noise z[N][P*nu] (and second-std
 * flags) through the absl shim on a ONE-thread pool (deterministic candidate order); z == NULL: the reference's own random
 * draws on `nthread` workers.  Returns the winner. */
int refp_optimize(void* hp, int N, int H, const double* z, const unsigned char* std2, int nthread) {
  Handle* h = (Handle*)hp; mjpc::SamplingPlanner& p = *h->planner;
  if (N > mjpc::kMaxTrajectory) return -1;
  activate(h);
  if (z) nthread = 1;
  if (!h->pool || h->pool_threads != nthread) { h->pool.reset(new mjpc::ThreadPool(nthread)); h->pool_threads = nthread; }
  p.num_trajectory_ = N;
  const int P = p.policy.num_spline_points;
  if (z) ref_shim_random::begin_replay(z, std2, N, (long long)P * h->model->nu);
  p.OptimizePolicy(H, *h->pool);
  ref_shim_random::end_replay();
  return p.winner;
}
void refp_get_returns(void* hp, int N, double* returns, unsigned char* failure) {
  Handle* h = (Handle*)hp;
  for (int i = 0; i < N; i++) { returns[i] = h->planner->trajectory[i].total_return; if (failure) failure[i] = h->planner->trajectory[i].failure; }
}
void refp_get_trajectory(void* hp, int i, double* states, double* actions, double* times, double* residual, double* costs, double* trace,
                         double* total_return, int* failure) {
  Handle* h = (Handle*)hp; const mjpc::Trajectory& t = h->planner->trajectory[i];
  const int H = t.horizon;
  if (states) std::memcpy(states, t.states.data(), sizeof(double) * H * t.dim_state);
  if (actions) std::memcpy(actions, t.actions.data(), sizeof(double) * H * t.dim_action);
  if (times) std::memcpy(times, t.times.data(), sizeof(double) * H);
  if (residual) std::memcpy(residual, t.residual.data(), sizeof(double) * H * t.dim_residual);
  if (costs) std::memcpy(costs, t.costs.data(), sizeof(double) * H);
  if (trace) std::memcpy(trace, t.trace.data(), sizeof(double) * H * 3 * h->task->num_trace);
  if (total_return) *total_return = t.total_return;
  if (failure) *failure = t.failure;
}
/* which: 0 policy, 1 previous_policy, 2 + i candidate_policy[i].  Returns the number of spline nodes. */
int refp_get_policy(void* hp, int which, double* times, double* knots) {
  Handle* h = (Handle*)hp; mjpc::SamplingPlanner& p = *h->planner; const int nu = h->model->nu;
  mjpc::SamplingPolicy& pol = which == 0 ? p.policy : (which == 1 ? p.previous_policy : p.candidate_policy[which - 2]);
  int n = 0;
  for (const auto& node : pol.plan) { times[n] = node.time(); for (int k = 0; k < nu; k++) knots[(size_t)n * nu + k] = node.values()[k]; n++; }
  return n;
}
void refp_action(void* hp, double t, int use_previous, double* out) { ((Handle*)hp)->planner->ActionFromPolicy(out, nullptr, t, use_previous != 0); }
double refp_improvement(void* hp) { return ((Handle*)hp)->planner->improvement; }

/* ---- task side: cost specification, residual and cost on an arbitrary state */
int refp_task_dims(void* hp, int* num_residual, int* num_term, int* num_trace, int* num_parameters) {
  mjpc::Task& t = *((Handle*)hp)->task;
  if (num_residual) *num_residual = t.num_residual; if (num_term) *num_term = t.num_term; if (num_trace) *num_trace = t.num_trace;
  if (num_parameters) *num_parameters = (int)t.parameters.size();
  return 0;
}
void refp_task_spec(void* hp, int* dim_norm_residual, int* norm, int* num_norm_parameter, double* weight, double* norm_parameter, double* risk,
                    double* parameters) {
  mjpc::Task& t = *((Handle*)hp)->task;
  for (int k = 0; k < t.num_term; k++) { dim_norm_residual[k] = t.dim_norm_residual[k]; norm[k] = (int)t.norm[k]; num_norm_parameter[k] = t.num_norm_parameter[k]; weight[k] = t.weight[k]; }
  int np = 0; for (int k = 0; k < t.num_term; k++) np += t.num_norm_parameter[k];
  for (int k = 0; k < np; k++) norm_parameter[k] = t.norm_parameter[k];
  *risk = t.risk;
  for (size_t k = 0; k < t.parameters.size(); k++) parameters[k] = t.parameters[k];
}
void refp_task_set(void* hp, int mode, const double* parameters, int nparam, const double* weight, int nweight) {
  Handle* h = (Handle*)hp; mjpc::Task& t = *h->task;
  if (mode >= 0) t.mode = mode;
  for (int k = 0; k < nparam && k < (int)t.parameters.size(); k++) t.parameters[k] = parameters[k];
  for (int k = 0; k < nweight && k < (int)t.weight.size(); k++) t.weight[k] = weight[k];
  t.UpdateResidual();
}
/* QuadrupedFlat: the ResidualFn state Transition() manages, in the PsTask.aux layout of mujoco_mpc_b200/task.py (A_MODE ... A_ORIENTATION),
 * 17 doubles.  Returns 0, or -1 when the task is not the quadruped. */
int refp_quadruped_state(void* hp, double* aux) {
  Handle* h = (Handle*)hp;
  mjpc::QuadrupedFlat* q = dynamic_cast<mjpc::QuadrupedFlat*>(h->task.get());
  if (!q) return -1;
  const mjpc::QuadrupedFlat::ResidualFn& r = q->residual_;
  aux[0] = (double)r.current_mode_; aux[1] = r.mode_start_time_; aux[2] = r.position_[0]; aux[3] = r.position_[1];
  aux[4] = r.heading_[0]; aux[5] = r.heading_[1]; aux[6] = r.speed_; aux[7] = r.angvel_;
  aux[8] = (double)mjpc::ReinterpretAsInt(r.current_gait_); aux[9] = r.phase_start_; aux[10] = r.phase_start_time_; aux[11] = r.phase_velocity_;
  aux[12] = r.ground_; for (int k = 0; k < 4; k++) aux[13 + k] = r.orientation_[k];
  return 0;
}
/* Task::Transition on (state, mocap, time); writes back the mocap poses it may have moved.  Returns task.mode. */
int refp_task_transition(void* hp, const double* state, double* mocap, double time) {
  Handle* h = (Handle*)hp; const mjModel* m = h->model; mjData* d = h->scratch;
  activate(h);
  mju_copy(d->qpos, state, m->nq); mju_copy(d->qvel, state + m->nq, m->nv); mju_copy(d->act, state + m->nq + m->nv, m->na);
  for (int i = 0; i < m->nmocap && mocap; i++) { mju_copy(d->mocap_pos + 3 * i, mocap + 7 * i, 3); mju_copy(d->mocap_quat + 4 * i, mocap + 7 * i + 3, 4); }
  d->time = time;
  mj_forward(m, d);
  h->task->Transition(h->model, d);
  for (int i = 0; i < m->nmocap && mocap; i++) { mju_copy(mocap + 7 * i, d->mocap_pos + 3 * i, 3); mju_copy(mocap + 7 * i + 3, d->mocap_quat + 4 * i, 4); }
  return h->task->mode;
}
/* residual + cost terms + cost of the task at (state, ctrl, mocap, time): mj_forward with the sensor callback */
double refp_task_residual(void* hp, const double* state, const double* ctrl, const double* mocap, double time, double* residual, double* terms) {
  Handle* h = (Handle*)hp; const mjModel* m = h->model; mjData* d = h->scratch;
  activate(h);
  mju_copy(d->qpos, state, m->nq); mju_copy(d->qvel, state + m->nq, m->nv); mju_copy(d->act, state + m->nq + m->nv, m->na);
  if (ctrl) mju_copy(d->ctrl, ctrl, m->nu);
  for (int i = 0; i < m->nmocap && mocap; i++) { mju_copy(d->mocap_pos + 3 * i, mocap + 7 * i, 3); mju_copy(d->mocap_quat + 4 * i, mocap + 7 * i + 3, 4); }
  d->time = time;
  mj_forward(m, d);
  if (residual) mju_copy(residual, d->sensordata, h->task->num_residual);
  if (terms) h->task->CostTerms(terms, d->sensordata);
  return h->task->CostValue(d->sensordata);
}

}  // extern "C"
