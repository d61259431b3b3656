// ORACLE (test infrastructure).  Restates Trajectory::NoisyRollout (mjpc/trajectory.cc:100-210),
// Trajectory::UpdateReturn (:312-326), SamplingPolicy::Action (planners/sampling/policy.cc:52-59)
// and SamplingPlanner::{UpdateNominalPolicy,AddNoiseToPolicy,Rollouts,OptimizePolicy}
// (planners/sampling/planner.cc:190-208,236-380,525-534) with injected noise.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include "mm_planner.h"
#include "mm_math.h"

namespace mm {

// Persistent worker pool of the CPU arm (stands in for mjpc::ThreadPool, mjpc/threadpool.cc:30-85): threads are
// created once, every parallelFor hands out indices one at a time from an atomic counter (dynamic scheduling, so a
// slow contact-rich rollout does not hold a whole static stripe back), worker k always passes tid = k.
class WorkerPool {
 public:
  static WorkerPool& get() { static WorkerPool p; return p; }
  void parallelFor(int n, int nthread, const std::function<void(int, int)>& fn) {
    if (nthread <= 1 || n <= 1) { for (int i = 0; i < n; i++) fn(0, i); return; }
    std::unique_lock<std::mutex> run(run_mtx_);              // one parallelFor at a time
    grow(nthread - 1);
    {
      std::lock_guard<std::mutex> lk(mtx_);
      fn_ = &fn; n_ = n; next_.store(0); active_ = nthread - 1; pending_ = nthread - 1; generation_++;
    }
    cv_.notify_all();
    for (int i; (i = next_.fetch_add(1)) < n;) fn(0, i);     // the caller is worker 0
    std::unique_lock<std::mutex> lk(mtx_);
    done_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }
  ~WorkerPool() {
    { std::lock_guard<std::mutex> lk(mtx_); stop_ = true; generation_++; }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
 private:
  void grow(int want) {
    while ((int)threads_.size() < want) {
      const int id = (int)threads_.size() + 1;
      threads_.emplace_back([this, id] {
        unsigned long seen = 0;
        for (;;) {
          const std::function<void(int, int)>* fn; int n;
          {
            std::unique_lock<std::mutex> lk(mtx_);
            cv_.wait(lk, [&] { return stop_ || (generation_ != seen && id <= active_); });
            if (stop_) return;
            seen = generation_; fn = fn_; n = n_;
          }
          for (int i; (i = next_.fetch_add(1)) < n;) (*fn)(id, i);
          { std::lock_guard<std::mutex> lk(mtx_); if (--pending_ == 0) done_.notify_one(); }
        }
      });
    }
  }
  std::vector<std::thread> threads_; std::mutex mtx_, run_mtx_; std::condition_variable cv_, done_;
  const std::function<void(int, int)>* fn_ = nullptr; int n_ = 0, active_ = 0, pending_ = 0; std::atomic<int> next_{0};
  unsigned long generation_ = 0; bool stop_ = false;
};

void policyAction(const Model* m, const Policy& p, double t, double* action) {
  splineSample(p.times.data(), p.knots.data(), p.P(), m->nu, p.interp, t, action);
  for (int i = 0; i < m->nu; i++) action[i] = clip(action[i], m->actuator_ctrlrange[2 * i], m->actuator_ctrlrange[2 * i + 1]);
}

struct CbUser { const PsTask* task; };
static void residualCallback(const Model* m, Data* d, void* user) {
  residual(m, d, ((CbUser*)user)->task, d->sensordata.data());
}

static bool checkWarnings(Data* d) {
  bool f = false;
  for (int i = 0; i < N_WARN; i++) if (d->warning[i] > 0) { d->warning[i] = 0; f = true; }
  return f;
}

static void getTraces(const Model* m, const Data* d, const PsTask* t, double* out) {
  for (int i = 0; i < t->num_trace; i++) copy(out + 3 * i, &d->sensordata[m->sensor_adr[t->trace_sensor[i]]], 3);
}

void rollout(const Model* m, const PsTask* task, Data* d, const Policy& p, const double* state, double time,
             const double* mocap, const double* userdata, int steps, Trajectory* tr,
             const double* xfrc_w, double xfrc_std, double xfrc_rate) {
  int nq = m->nq, nv = m->nv, na = m->na, nu = m->nu, ns = nq + nv + na, nr = task->num_residual, ntr = 3 * task->num_trace;
  tr->failure = false; tr->horizon = steps;
  tr->states.assign(ns * steps, 0); tr->actions.assign(nu * steps, 0); tr->times.assign(steps, 0);
  tr->residual.assign(nr * steps, 0); tr->costs.assign(steps, 0); tr->trace.assign(ntr * steps, 0);
  for (int i = 0; i < m->nmocap; i++) { copy(&d->mocap_pos[3 * i], mocap + 7 * i, 3); copy(&d->mocap_quat[4 * i], mocap + 7 * i + 3, 4); }
  copy(d->userdata.data(), userdata, m->nuserdata);
  copy(tr->states.data(), state, ns);
  copy(d->qpos.data(), state, nq); copy(d->qvel.data(), state + nq, nv); copy(d->act.data(), state + nq + nv, na);
  tr->times[0] = time; d->time = time;
  // deterministic warm start (the reference inherits qacc_warmstart from whatever rollout last
  // used this worker's mjData; see DESIGN.md "deviations")
  zero(d->qacc_warmstart.data(), nv);
  // same for xfrc_applied: the reference's OU state carries over between rollouts of one worker
  zero(d->xfrc_applied.data(), 6 * m->nbody);
  for (int i = 0; i < N_WARN; i++) d->warning[i] = 0;
  CbUser cu{task};
  for (int t = 0; t < steps - 1; t++) {
    policyAction(m, p, d->time, &tr->actions[t * nu]);
    copy(d->ctrl.data(), &tr->actions[t * nu], nu);
    if (xfrc_w && xfrc_std > 0) {     // trajectory.cc:147-155
      const double rate = std::exp(-m->opt_timestep / xfrc_rate), scale = xfrc_std * std::sqrt(1 - rate * rate);
      for (int i = 0; i < 6 * m->nbody; i++) d->xfrc_applied[i] = rate * d->xfrc_applied[i] + scale * xfrc_w[(size_t)t * 6 * m->nbody + i];
    }
    step(m, d, residualCallback, &cu);
    copy(&tr->residual[t * nr], d->sensordata.data(), nr);
    getTraces(m, d, task, &tr->trace[t * ntr]);
    if ((tr->failure |= checkWarnings(d))) { tr->total_return = 1.0e6; return; }
    copy(&tr->states[(t + 1) * ns], d->qpos.data(), nq); copy(&tr->states[(t + 1) * ns + nq], d->qvel.data(), nv);
    copy(&tr->states[(t + 1) * ns + nq + nv], d->act.data(), na);
    tr->times[t + 1] = d->time;
  }
  if (steps > 1) copy(&tr->actions[(steps - 1) * nu], &tr->actions[(steps - 2) * nu], nu);
  forward(m, d, residualCallback, &cu);
  copy(&tr->residual[(steps - 1) * nr], d->sensordata.data(), nr);
  getTraces(m, d, task, &tr->trace[(steps - 1) * ntr]);
  tr->total_return = 0;
  for (int t = 0; t < steps; t++) { tr->costs[t] = costValue(task, &tr->residual[t * nr]); tr->total_return += tr->costs[t]; }
  tr->total_return /= std::max(steps, 1);
}

// TimeSpline::DiscardBefore (mjpc/spline/spline.cc:164-188): drop the nodes the interpolation at `time` no
// longer needs -- everything before the last node with t <= time, one more kept for the cubic's slope
static void discardBefore(Policy* p, double time, int nu) {
  int first = (int)(std::upper_bound(p->times.begin(), p->times.end(), time) - p->times.begin());
  if (first == 0) return;
  int last = first - 1;
  if (p->interp == PS_SPLINE_CUBIC && last > 0) last--;
  p->times.erase(p->times.begin(), p->times.begin() + last);
  p->knots.erase(p->knots.begin(), p->knots.begin() + (size_t)last * nu);
}

// SamplingPlanner::UpdateNominalPolicy, sliding branch (planner.cc:244-281): the plan keeps its nodes, the
// ones behind `time` are dropped and copies of the last node are appended further in the future
static void updateNominalSliding(Planner* pl, int horizon) {
  const Model* m = pl->model;
  int P = pl->num_spline_points, nu = m->nu;
  double th = (horizon - 1) * m->opt_timestep;
  int extra = pl->interpolation == PS_SPLINE_ZERO ? 1 : pl->interpolation == PS_SPLINE_LINEAR ? 2 : 4;
  double shift = P > extra ? std::fmax(th / (P - extra), 1.0e-5) : th;
  Policy& p = pl->policy;
  discardBefore(&p, pl->time, nu);
  if (p.times.empty()) { p.times.push_back(pl->time); p.knots.insert(p.knots.end(), nu, 0.0); }
  while ((int)p.times.size() < P) {
    p.times.push_back(p.times.back() + shift);
    std::vector<double> last(p.knots.end() - nu, p.knots.end());
    p.knots.insert(p.knots.end(), last.begin(), last.end());
  }
}

// SamplingPlanner::UpdateNominalPolicy, non-sliding branch (planner.cc:282-309)
static void updateNominal(Planner* pl, int horizon) {
  if (pl->sliding_plan) { updateNominalSliding(pl, horizon); return; }
  const Model* m = pl->model;
  const Policy& src = pl->candidate.empty() ? pl->policy : pl->candidate[pl->winner];
  int P = pl->num_spline_points, nu = m->nu;
  double t = pl->time, th = (horizon - 1) * m->opt_timestep;
  double shift = pl->interpolation == PS_SPLINE_ZERO ? std::fmax(th / P, 1.0e-5) : std::fmax(th / (P - 1), 1.0e-5);
  Policy np; np.interp = pl->interpolation; np.times.resize(P); np.knots.resize(P * nu);
  for (int k = 0; k < P; k++) { np.times[k] = t; policyAction(m, src, t, &np.knots[k * nu]); t += shift; }
  pl->policy = np;
}

// OptimizePolicy with injected noise z[N][P][nu] (sample-major, node, actuator: the draw order of
// planner.cc:327-333) and per-sample use_std2 flags (the Bernoulli draw of :323)
void optimizePolicy(Planner* pl, int N, int horizon, const double* z, const unsigned char* use_std2, int nthread, bool resample) {
  const Model* m = pl->model; int nu = m->nu;
  if (resample) updateNominal(pl, horizon);     // OptimizePolicy; OptimizePolicyCandidates alone (:151-187) does not re-time the nominal
  pl->policy.interp = pl->interpolation;
  int P = pl->policy.P();
  pl->candidate.assign(N, pl->policy);
  pl->trajectory.resize(N);
  for (int i = 1; i < N; i++) {
    double sd = pl->noise_exploration[0];
    if (pl->noise_exploration[1] > 0 && use_std2 && use_std2[i]) sd = pl->noise_exploration[1];
    Policy& c = pl->candidate[i];
    for (int k = 0; k < P; k++) {
      for (int a = 0; a < nu; a++) {
        double scale = 0.5 * (m->actuator_ctrlrange[2 * a + 1] - m->actuator_ctrlrange[2 * a]);
        c.knots[k * nu + a] += (scale * sd) * z[((size_t)i * P + k) * nu + a];
      }
      for (int a = 0; a < nu; a++) c.knots[k * nu + a] = clip(c.knots[k * nu + a], m->actuator_ctrlrange[2 * a], m->actuator_ctrlrange[2 * a + 1]);
    }
  }
  nthread = std::max(1, nthread);
  if ((int)pl->data.size() < nthread) { size_t old = pl->data.size(); pl->data.resize(nthread); for (size_t k = old; k < pl->data.size(); k++) makeData(m, &pl->data[k]); }
  WorkerPool::get().parallelFor(N, nthread, [&](int tid, int i) {
    rollout(m, &pl->task, &pl->data[tid], pl->candidate[i], pl->state.data(), pl->time, pl->mocap.data(), pl->userdata.data(), horizon, &pl->trajectory[i]);
  });
  int best = 0;
  for (int i = 1; i < N; i++) if (pl->trajectory[i].total_return < pl->trajectory[best].total_return) best = i;
  pl->winner = best;
  pl->previous_policy = pl->policy; pl->policy = pl->candidate[best];
  pl->improvement = std::fmax(pl->trajectory[0].total_return - pl->trajectory[best].total_return, 0.0);
}

// CrossEntropyPlanner::OptimizePolicy (mjpc/planners/cross_entropy/planner.cc:164-284) with injected noise
// z[N][P][nu]: ResamplePolicy (:317-341, no zero-order special case), AddNoiseToPolicy for EVERY candidate
// with std max(sqrt(variance_k), std_min) un-scaled by the control range (:343-375), N rollouts plus the
// nominal one (:378-423), elite mean -> new policy, the reference's variance formula (:243-256: `n` is the
// BEST candidate's node for every elite i), improvement = max(avg_return - best, 0).  partial_sort ties
// are resolved towards the lower index.
void optimizePolicyCE(Planner* pl, int N, int horizon, const double* z, int nthread) {
  const Model* m = pl->model; int nu = m->nu;
  int P = pl->num_spline_points, np = P * nu;
  int n_elite = std::min(pl->n_elite, N);
  if ((int)pl->variance.size() != np) pl->variance.assign(np, pl->noise_exploration[0] * pl->noise_exploration[0]);   // Reset: std_initial^2 (:109-111)
  // resample the nominal to the current time
  Policy rp; rp.interp = pl->interpolation; rp.times.resize(P); rp.knots.resize(np);
  {
    Policy src = pl->policy; src.interp = pl->interpolation;
    double t = pl->time, shift = std::fmax((horizon - 1) * m->opt_timestep / (P - 1), 1.0e-5);
    for (int k = 0; k < P; k++) { rp.times[k] = t; policyAction(m, src, t, &rp.knots[k * nu]); t += shift; }
  }
  pl->candidate.assign(N, rp);
  pl->trajectory.resize(N);
  for (int i = 0; i < N; i++) {
    Policy& c = pl->candidate[i];
    for (int k = 0; k < P; k++) {
      for (int a = 0; a < nu; a++) c.knots[k * nu + a] += std::max(std::sqrt(pl->variance[k * nu + a]), pl->std_min) * z[((size_t)i * P + k) * nu + a];
      for (int a = 0; a < nu; a++) c.knots[k * nu + a] = clip(c.knots[k * nu + a], m->actuator_ctrlrange[2 * a], m->actuator_ctrlrange[2 * a + 1]);
    }
  }
  nthread = std::max(1, nthread);
  if ((int)pl->data.size() < nthread) { size_t old = pl->data.size(); pl->data.resize(nthread); for (size_t k = old; k < pl->data.size(); k++) makeData(m, &pl->data[k]); }
  WorkerPool::get().parallelFor(N + 1, nthread, [&](int tid, int i) {
    rollout(m, &pl->task, &pl->data[tid], i < N ? pl->candidate[i] : rp, pl->state.data(), pl->time, pl->mocap.data(), pl->userdata.data(), horizon,
            i < N ? &pl->trajectory[i] : &pl->nominal_trajectory);
  });
  pl->trajectory_order.resize(N);
  for (int i = 0; i < N; i++) pl->trajectory_order[i] = i;
  std::stable_sort(pl->trajectory_order.begin(), pl->trajectory_order.end(), [&](int a, int b) {
    double ra = pl->trajectory[a].total_return, rb = pl->trajectory[b].total_return;
    if (!(ra == ra)) ra = 1.0e300;
    if (!(rb == rb)) rb = 1.0e300;
    return ra < rb; });
  std::vector<double> mean(np, 0.0); double avg_return = 0;
  for (int e = 0; e < n_elite; e++) {
    int idx = pl->trajectory_order[e];
    for (int k = 0; k < np; k++) mean[k] += pl->candidate[idx].knots[k];
    avg_return += pl->trajectory[idx].total_return;
  }
  for (int k = 0; k < np; k++) mean[k] = mean[k] * (1.0 / n_elite);
  avg_return /= n_elite;
  std::fill(pl->variance.begin(), pl->variance.end(), 0.0);
  const Policy& best = pl->candidate[pl->trajectory_order[0]];
  for (int k = 0; k < np; k++) {
    for (int i = 0; i < n_elite; i++) { double diff = best.knots[k] - mean[k]; pl->variance[k] += diff * diff / (n_elite - 1); }
  }
  // previous_policy is only ever set by Reset in the reference (cross_entropy/planner.cc:129)
  pl->policy.times = rp.times; pl->policy.knots = mean; pl->policy.interp = pl->interpolation;
  pl->winner = pl->trajectory_order[0]; pl->avg_return = avg_return;
  pl->improvement = std::fmax(avg_return - pl->trajectory[pl->trajectory_order[0]].total_return, 0.0);
}

// RobustPlanner::OptimizePolicy over a SamplingPlanner delegate (mjpc/planners/robust/robust_planner.cc:91-157):
// OptimizePolicyCandidates(ncandidates) = one sampling iteration ranked by (return, index); every candidate is
// re-rolled `repetitions` times under OU force noise w[(c*repetitions+j)][H-1][6*nbody]; a candidate's score is
// the running mean of its own return and its non-failed noisy returns (:136-149); the lowest mean wins
// (first on ties) and becomes the policy.
void optimizePolicyRobust(Planner* pl, int N, int horizon, const double* z, int ncandidates, int repetitions,
                          const double* w, double xfrc_std, double xfrc_rate) {
  const Model* m = pl->model;
  // RobustPlanner calls OptimizePolicyCandidates directly, so the nominal spline is NOT resampled (:91-95)
  optimizePolicy(pl, N, horizon, z, nullptr, 1, false);   // leaves candidate[], trajectory[]; policy = argmin
  ncandidates = std::min(ncandidates, N);
  std::vector<int> order(N);
  for (int i = 0; i < N; i++) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return pl->trajectory[a].total_return < pl->trajectory[b].total_return; });
  pl->trajectory_order.assign(order.begin(), order.begin() + ncandidates);
  pl->robust_best = 0; pl->robust_score.assign(ncandidates, 0.0);
  if (ncandidates <= 1) return;
  pl->robust_trajectory.assign((size_t)ncandidates * repetitions, Trajectory());
  const size_t wn = (size_t)(horizon - 1) * 6 * m->nbody;
  for (int c = 0; c < ncandidates; c++)
    for (int j = 0; j < repetitions; j++)
      rollout(m, &pl->task, &pl->data[0], pl->candidate[order[c]], pl->state.data(), pl->time, pl->mocap.data(), pl->userdata.data(), horizon,
              &pl->robust_trajectory[(size_t)c * repetitions + j], w + ((size_t)c * repetitions + j) * wn, xfrc_std, xfrc_rate);
  int best = -1; double best_score = 0;
  for (int c = 0; c < ncandidates; c++) {
    double mean = pl->trajectory[order[c]].total_return; int valid = 0;
    for (int j = 0; j < repetitions; j++) {
      const Trajectory& t = pl->robust_trajectory[(size_t)c * repetitions + j];
      if (t.failure) continue;
      mean = (valid * mean + t.total_return) / (valid + 1); valid++;
    }
    pl->robust_score[c] = mean;
    if (best == -1 || mean < best_score) { best = c; best_score = mean; }
  }
  pl->robust_best = best;
  // CopyCandidateToPolicy(best): previous_policy = the resampled nominal of this iteration (set by optimizePolicy)
  pl->winner = order[best]; pl->policy = pl->candidate[order[best]];
}

}  // namespace mm
