// mjpc::CudaSamplingAdapter: see cuda_sampling_adapter.h.
#include "cuda_sampling_adapter.h"
#include <cstring>
#include <string>

namespace mjpc {

void FillPsModel(const mjModel* m, PsModel* p, std::vector<std::vector<int>>* scratch) {
  scratch->clear();
  auto widen = [scratch](const mjtByte* v, size_t n) {
    scratch->emplace_back(n ? n : 1);
    for (size_t i = 0; i < n; i++) scratch->back()[i] = v[i];
    return (const int*)scratch->back().data();
  };
  std::memset(p, 0, sizeof(*p));
#include "fill_model.inc"
}

namespace {
struct TaskEntry { const char* name; int id; const char* sensors[4]; };
// Task::Name() -> residual implemented on the device + the sensors that residual looks up by name
const TaskEntry kTasks[] = {
    {"Particle", PS_TASK_PARTICLE, {"position", "velocity", nullptr, nullptr}},
    {"ParticleFixed", PS_TASK_PARTICLE_FIXED, {"position", "velocity", nullptr, nullptr}},
    {"Cartpole", PS_TASK_CARTPOLE, {nullptr, nullptr, nullptr, nullptr}},
    {"Acrobot", PS_TASK_ACROBOT, {nullptr, nullptr, nullptr, nullptr}},
    {"Swimmer", PS_TASK_SWIMMER, {"nose", "target", nullptr, nullptr}},
    {"Walker", PS_TASK_WALKER, {"torso_position", "torso_zaxis", "torso_subtreelinvel", nullptr}},
    {"Quadruped Flat", PS_TASK_QUADRUPED, {"torso_subtreecom", "torso_subtreelinvel", "torso_angmom", nullptr}},
};
}  // namespace

bool FillPsTask(const mjModel* m, const Task& t, PsTask* p) {
  const std::string name = t.Name();
  const TaskEntry* e = nullptr;
  for (const TaskEntry& k : kTasks) if (name == k.name) e = &k;
  if (!e) return false;
  // aux / iaux (task state) are owned by task_state_hook: keep them across snapshots
  p->task_id = e->id;
  p->num_residual = t.num_residual; p->num_term = t.num_term; p->num_trace = t.num_trace; p->risk = t.risk;
  if (t.num_term > PS_MAX_COST_TERMS || (int)t.parameters.size() > PS_MAX_TASK_PARAMS) return false;
  int np = 0;
  for (int k = 0; k < t.num_term; k++) {
    p->dim_norm_residual[k] = t.dim_norm_residual[k]; p->norm[k] = (int)t.norm[k]; p->weight[k] = t.weight[k];
    p->num_norm_parameter[k] = t.num_norm_parameter[k]; np += t.num_norm_parameter[k];
  }
  for (int k = 0; k < np; k++) p->norm_parameter[k] = t.norm_parameter[k];
  p->num_parameters = (int)t.parameters.size();
  for (int k = 0; k < p->num_parameters; k++) p->parameters[k] = t.parameters[k];
  for (int k = 0; k < 16; k++) p->sensor_id[k] = -1;
  for (int k = 0; k < 4 && e->sensors[k]; k++) {
    p->sensor_id[k] = mj_name2id(m, mjOBJ_SENSOR, e->sensors[k]);
    if (p->sensor_id[k] < 0) return false;
  }
  for (int k = 0; k < 4; k++) p->trace_sensor[k] = 0;
  for (int k = 0; k < t.num_trace && k < 4; k++) p->trace_sensor[k] = mj_name2id(m, mjOBJ_SENSOR, ("trace" + std::to_string(k)).c_str());
  return true;
}

void CudaSamplingAdapter::Snapshot() {
  // Agent::PlanIteration overrides the planning model's time step and integrator every iteration (agent.cc:275-276) and
  // snapshots the task's residual parameters (agent.cc:290)
  pm_.opt_timestep = model_->opt.timestep; pm_.opt_integrator = model_->opt.integrator;
  if (!FillPsTask(model_, *task_, &pt_)) mju_error("CudaSamplingAdapter: task '%s' has no device residual", task_->Name().c_str());
  if (task_state_hook) task_state_hook(model_, *task_, &pt_);
}

void CudaSamplingAdapter::Initialize(mjModel* model, const Task& task) {
  model_ = model; task_ = &task;
  FillPsModel(model, &pm_, &widened_);
  std::memset(&pt_, 0, sizeof(pt_));
  for (int k = 0; k < 16; k++) pt_.iaux[k] = -1;
  Snapshot();
  impl_.Initialize(&pm_, &pt_, device_ordinal);       // reads sampling_* numerics like SamplingPlanner::Initialize
}

void CudaSamplingAdapter::Allocate() {
  state_.resize(model_->nq + model_->nv + model_->na); mocap_.resize(7 * model_->nmocap); userdata_.resize(model_->nuserdata);
  impl_.Allocate();
}

void CudaSamplingAdapter::Reset(int horizon, const double* initial_repeated_action) { impl_.Reset(horizon, initial_repeated_action); }

void CudaSamplingAdapter::SetState(const State& state) {
  state.CopyTo(state_.data(), mocap_.data(), userdata_.data(), &time_);          // mjpc/states/state.cc:128-135
  impl_.SetState(state_.data(), mocap_.data(), userdata_.data(), time_);
}

void CudaSamplingAdapter::OptimizePolicy(int horizon, ThreadPool&) {
  Snapshot(); impl_.SetTask(&pt_);
  impl_.OptimizePolicy(horizon);
}

int CudaSamplingAdapter::OptimizePolicyCandidates(int ncandidates, int horizon, ThreadPool&) {
  Snapshot(); impl_.SetTask(&pt_);
  return impl_.OptimizePolicyCandidates(ncandidates, horizon);
}

void CudaSamplingAdapter::NominalTrajectory(int horizon, ThreadPool&) { impl_.NominalTrajectory(horizon); }

void CudaSamplingAdapter::ActionFromPolicy(double* action, const double* state, double time, bool use_previous) {
  impl_.ActionFromPolicy(action, state, time, use_previous);
}

const Trajectory* CudaSamplingAdapter::BestTrajectory() {
  const mjpc_b200::Trajectory* b = impl_.BestTrajectory();
  if (!b) return nullptr;
  best_.horizon = b->horizon; best_.dim_state = b->dim_state; best_.dim_action = b->dim_action; best_.dim_residual = b->dim_residual;
  best_.dim_trace = b->dim_trace; best_.states = b->states; best_.actions = b->actions; best_.times = b->times; best_.residual = b->residual;
  best_.costs = b->costs; best_.trace = b->trace; best_.total_return = b->total_return; best_.failure = b->failure;
  return &best_;
}

}  // namespace mjpc
