// TEST INFRASTRUCTURE (not product): a host-compiled stand-in for the subset of the C ABI
// (include/ps_engine.h) that the host planner classes call, so that CudaSamplingPlanner /
// CudaCrossEntropyPlanner / CudaRobustPlanner (mujoco_mpc_b200/host/*.cc, compiled unchanged) can be exercised against the
// oracle in this GPU-less container.  Rollouts run the engine's own kernel-source headers
// (csrc/ps_*.cuh: rollout_sample<T, true> and <T, true, true>) through the host compiler, one sample
// after the other; the knot perturbation restates K1 for INJECTED noise only (no generator here).
// tests/hostsim/driver.py links this file with the planner sources into libps_planner_host.so with
// -Bsymbolic; nothing in the product links, loads or falls back to it.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>
#include "../../include/ps_engine.h"
#include "../../mujoco_mpc_b200/csrc/ps_host_model.h"
#include "../../mujoco_mpc_b200/csrc/ps_coop_rollout.cuh"

using namespace ps;

namespace {
std::string g_err;
int fail(const std::string& s) { g_err = s; return -1; }

struct Runner {
  virtual ~Runner() {}
  virtual int fill(const PsModel* m, const PairList& pl, std::string* why) = 0;
  virtual void options(double timestep, int integrator) = 0;
  virtual void task(const DevTask& t, int need) = 0;
  virtual void plan(const double* state, int ns, const double* mocap, int nmocap, double time, const double* times, int P, int interp, int H) = 0;
  virtual double roll(const double* knots, const Writeback& wb, int* flags, const XfrcNoise* xf) = 0;
  virtual double timestep() const = 0;
  virtual void warmstart(const double* w, int nv) = 0;
};
template <class T>
struct RunnerT : Runner {
  std::vector<char> buf; Constants<T>* c;
  RunnerT() : buf(sizeof(Constants<T>) + 256) {
    c = (Constants<T>*)(((uintptr_t)buf.data() + 127) & ~(uintptr_t)127);
    std::memset((void*)c, 0, sizeof(Constants<T>));
  }
  int fill(const PsModel* m, const PairList& pl, std::string* why) override { return fill_model<T>(m, pl, c, why); }
  void options(double timestep, int integrator) override { set_options<T>(c, timestep, integrator); }
  void task(const DevTask& t, int need) override { set_task<T>(c, t, need); }
  void plan(const double* state, int ns, const double* mocap, int nmocap, double time, const double* times, int P, int interp, int H) override {
    set_plan<T>(c, state, ns, mocap, nmocap, time, times, P, interp, H);
  }
  double roll(const double* knots, const Writeback& wb, int* flags, const XfrcNoise* xf) override {
    if (xf && xf->z) return rollout_sample<T, true, true>(c->m, c->t, c->p, knots, 1, 0, wb, flags, xf);
    return rollout_sample<T, true>(c->m, c->t, c->p, knots, 1, 0, wb, flags);
  }
  double timestep() const override { return c->m.timestep; }
  void warmstart(const double* w, int nv) override { set_warmstart<T>(c, w, nv); }
};

struct Sample {
  std::vector<double> states, actions, times, residual, costs, trace, warm;
  void size(int H, int ns, int nu, int nr, int nt) {
    warm.assign(64, 0);       // >= nv of every capacity class
    states.assign((size_t)H * ns, 0); actions.assign((size_t)H * (nu ? nu : 1), 0); times.assign(H, 0);
    residual.assign((size_t)H * (nr ? nr : 1), 0); costs.assign(H, 0); trace.assign((size_t)H * (nt ? nt : 1), 0);
  }
  Writeback wb() { return Writeback{states.data(), actions.data(), residual.data(), costs.data(), times.data(), trace.data(), warm.data()}; }
};
}  // namespace

struct PsEngine_ {
  const PsModel* m = nullptr; PsTask task{}; bool have_task = false;
  std::unique_ptr<Runner> run;
  std::vector<double> state, mocap; double time = 0;
  std::vector<double> times, nominal; int P = 0, interp = 0;
  std::vector<double> z; std::vector<unsigned char> std2; int zn = 0;
  std::vector<double> param_std;
  int N = 0, H = 0; long long offset = 0;
  std::vector<double> knots, returns; std::vector<unsigned char> failure;
  std::vector<Sample> samples;
  std::vector<double> record;
  long long launches = 0;
};

extern "C" {

const char* ps_last_error(void) { return g_err.c_str(); }
PsEngine* ps_create(int) { return new PsEngine_; }
void ps_destroy(PsEngine* e) { delete e; }
int ps_set_engine_flavour(PsEngine*, int) { return 0; }
int ps_set_seed(PsEngine*, unsigned long long) { return 0; }
int ps_last_kernel_ms(PsEngine*, float* a, float* b, float* c) { if (a) *a = 0; if (b) *b = 0; if (c) *c = 0; return 0; }

int ps_set_model(PsEngine* e, const PsModel* m) {
  PairList pl = static_pairs(m);
  std::string why;
  e->m = m; e->run.reset();
  { auto r = std::make_unique<RunnerT<TinyT>>(); if (!r->fill(m, pl, &why)) e->run = std::move(r); }
  if (!e->run) { auto r = std::make_unique<RunnerT<PlanarT>>(); if (!r->fill(m, pl, &why)) e->run = std::move(r); }
  if (!e->run) { auto r = std::make_unique<RunnerT<LeggedT>>(); if (!r->fill(m, pl, &why)) e->run = std::move(r); }
  if (!e->run) return fail("ps_set_model: " + why);
  // mocap bodies start at their model pose until ps_set_state says otherwise
  e->mocap.assign(7 * (m->nmocap ? m->nmocap : 1), 0.0);
  for (int b = 0; b < m->nbody; b++) if (m->body_mocapid[b] >= 0) {
    const int id = m->body_mocapid[b];
    for (int k = 0; k < 3; k++) e->mocap[7 * id + k] = m->body_pos[3 * b + k];
    for (int k = 0; k < 4; k++) e->mocap[7 * id + 3 + k] = m->body_quat[4 * b + k];
  }
  e->state.assign(m->nq + m->nv + m->na, 0.0);
  return 0;
}
int ps_set_options(PsEngine* e, double timestep, int integrator) {
  if (!e->run) return fail("ps_set_options: no model");
  e->run->options(timestep, integrator); return 0;
}
int ps_set_task(PsEngine* e, const PsTask* t) {
  if (!e->run) return fail("ps_set_task: no model");
  DevTask dt; int need = 0; std::string why;
  if (convert_task(t, e->m->sensor_type, e->m->nsensor, &dt, &need, &why)) return fail("ps_set_task: " + why);
  e->run->task(dt, need); e->task = *t; e->have_task = true; return 0;
}
int ps_set_state(PsEngine* e, const double* state, const double* mocap, const double*, double time) {
  if (!e->run) return fail("ps_set_state: no model");
  std::copy(state, state + e->state.size(), e->state.begin());
  if (mocap && e->m->nmocap) std::copy(mocap, mocap + 7 * e->m->nmocap, e->mocap.begin());
  e->time = time; return 0;
}
int ps_set_warmstart(PsEngine* e, const double* w, int nv) {
  if (!e->run) return fail("ps_set_warmstart: no model");
  if (w && nv != e->m->nv) return fail("ps_set_warmstart: nv does not match the model");
  e->run->warmstart(w, e->m->nv); return 0;
}
int ps_get_warmstart(PsEngine* e, int sample, double* w) {
  if (sample < 0 || sample >= (int)e->samples.size() || !w) return fail("ps_get_warmstart: bad sample");
  std::copy(e->samples[sample].warm.begin(), e->samples[sample].warm.begin() + e->m->nv, w); return 0;
}
int ps_set_nominal(PsEngine* e, const double* times, const double* knots, int P, int interp) {
  if (!e->run) return fail("ps_set_nominal: no model");
  e->times.assign(times, times + P); e->nominal.assign(knots, knots + (size_t)P * e->m->nu); e->P = P; e->interp = interp; return 0;
}
int ps_set_noise(PsEngine* e, const double* z, const unsigned char* use_std2, int N) {
  e->z.clear(); e->std2.clear(); e->zn = 0;
  if (z) { e->z.assign(z, z + (size_t)N * e->P * e->m->nu); e->zn = N; if (use_std2) e->std2.assign(use_std2, use_std2 + N); }
  return 0;
}
int ps_set_param_std(PsEngine* e, const double* s) {
  e->param_std.clear();
  if (s) e->param_std.assign(s, s + (size_t)e->P * e->m->nu);
  return 0;
}

static void roll_one(PsEngine* e, int i, const double* knots, const XfrcNoise* xf) {
  const PsModel* m = e->m;
  const int ns = m->nq + m->nv + m->na, nu = m->nu, nr = e->task.num_residual, nt = 3 * e->task.num_trace;
  Sample& s = e->samples[i];
  s.size(e->H, ns, nu, nr, nt);
  int flags = 0;
  e->returns[i] = e->run->roll(knots, s.wb(), &flags, xf);
  e->failure[i] = (unsigned char)flags;
}

static int begin_batch(PsEngine* e, int N, int H) {
  if (!e->run || !e->have_task) return fail("model and task must be set");
  if (e->P <= 0) return fail("set the nominal plan first");
  if (N < 1 || H < 2 || H > 512) return fail("bad sizes");
  e->N = N; e->H = H;
  const int nelem = e->P * e->m->nu;
  e->knots.assign((size_t)N * nelem, 0.0); e->returns.assign(N, 0.0); e->failure.assign(N, 0);
  e->samples.resize(N);
  e->run->plan(e->state.data(), (int)e->state.size(), e->mocap.data(), e->m->nmocap, e->time, e->times.data(), e->P, e->interp, H);
  return 0;
}

// K1 (injected noise) + K2 + K3, restating ps_kernels_aux.cuh: perturb_kernel / argmin_kernel
int ps_rollouts(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset, int* winner, double* winner_return,
                double* nominal_return) {
  if (begin_batch(e, N, H)) return -1;
  if (e->zn < N && (sigma0 > 0 || !e->param_std.empty())) return fail("the host stand-in has no noise generator: inject z with ps_set_noise");
  const PsModel* m = e->m;
  const int nu = m->nu, nelem = e->P * nu;
  e->offset = sample_offset;
  const bool ce = !e->param_std.empty();
  for (int i = 0; i < N; i++) {
    const long long gi = (long long)sample_offset + i;
    const bool noisy = ce ? (i != N - 1) : (gi != 0);
    double* k = e->knots.data() + (size_t)i * nelem;
    for (int el = 0; el < nelem; el++) {
      const int a = el % nu;
      const double lo = m->actuator_ctrlrange[2 * a], hi = m->actuator_ctrlrange[2 * a + 1];
      double v = e->nominal[el];
      if (noisy && e->zn >= N) {
        double sd = sigma0; const double z = e->z[(size_t)i * nelem + el];
        if (sigma1 > 0 && !e->std2.empty() && e->std2[i]) sd = sigma1;
        if (ce) v += e->param_std[el] * z;
        else { const double scale = 0.5 * (hi - lo); v += (scale * sd) * z; }
        v = v < lo ? lo : (v > hi ? hi : v);
      }
      k[el] = v;
    }
    roll_one(e, i, k, nullptr);
  }
  int best = 0;
  for (int i = 1; i < N; i++) if (e->returns[i] < e->returns[best]) best = i;     // lowest return, ties -> lowest index
  e->record.assign(3 + nelem, 0.0);
  e->record[0] = e->returns[best]; e->record[1] = (double)(sample_offset + best); e->record[2] = e->returns[0];
  std::copy(e->knots.begin() + (size_t)best * nelem, e->knots.begin() + (size_t)(best + 1) * nelem, e->record.begin() + 3);
  if (winner) *winner = best;
  if (winner_return) *winner_return = e->record[0];
  if (nominal_return) *nominal_return = e->record[2];
  e->launches += 3;
  return 0;
}

// the stand-in always keeps every sample's trajectory
int ps_rollouts_writeback(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset, int* winner,
                          double* winner_return, double* nominal_return) {
  return ps_rollouts(e, N, H, sigma0, sigma1, sample_offset, winner, winner_return, nominal_return);
}
int ps_variant(PsEngine*, char* out, int cap) { std::snprintf(out, cap, "host stand-in"); return 0; }
int ps_kernel_launch_count(PsEngine* e, long long* launches) { if (launches) *launches = e->launches; return 0; }

int ps_rollout_candidates(PsEngine* e, const double* knots, int n, int repetitions, int H, double xfrc_std, double xfrc_rate,
                          const double* z, double* returns, unsigned char* failure) {
  if (n < 1 || repetitions < 1) return fail("ps_rollout_candidates: bad sizes");
  if (begin_batch(e, n * repetitions, H)) return -1;
  if (xfrc_std > 0 && !z) return fail("the host stand-in has no noise generator: inject the force noise");
  const int nelem = e->P * e->m->nu, n6 = 6 * e->m->nbody;
  e->offset = 0;
  XfrcNoise xf{nullptr, 0.0, 0.0};
  if (xfrc_std > 0) { xf.rate = std::exp(-e->run->timestep() / xfrc_rate); xf.scale = xfrc_std * std::sqrt(1 - xf.rate * xf.rate); }
  for (int c = 0; c < n; c++) for (int j = 0; j < repetitions; j++) {
    const int i = c * repetitions + j;
    std::copy(knots + (size_t)c * nelem, knots + (size_t)(c + 1) * nelem, e->knots.begin() + (size_t)i * nelem);
    if (xfrc_std > 0) xf.z = z + (size_t)i * (H - 1) * n6;
    roll_one(e, i, e->knots.data() + (size_t)i * nelem, xfrc_std > 0 ? &xf : nullptr);
  }
  if (returns) std::copy(e->returns.begin(), e->returns.end(), returns);
  if (failure) std::copy(e->failure.begin(), e->failure.end(), failure);
  e->launches += 1;
  return 0;
}

int ps_get_returns(PsEngine* e, double* total_return, unsigned char* failure) {
  if (!e->N) return fail("ps_get_returns: no batch");
  if (total_return) std::copy(e->returns.begin(), e->returns.end(), total_return);
  if (failure) std::copy(e->failure.begin(), e->failure.end(), failure);
  return 0;
}
int ps_get_knots(PsEngine* e, int sample, double* knots) {
  if (sample < 0 || sample >= e->N) return fail("ps_get_knots: sample out of range");
  const int nelem = e->P * e->m->nu;
  std::copy(e->knots.begin() + (size_t)sample * nelem, e->knots.begin() + (size_t)(sample + 1) * nelem, knots);
  return 0;
}
// no multi-GPU exchange in the host stand-in (the sharded-planner CPU tests use the AllGatherFn callback)
int ps_comm_init(PsEngine*, int, int world, const void*) { return world == 1 ? 0 : fail("ps_comm_init: host stand-in has no NCCL"); }
int ps_get_winner_record(PsEngine* e, double* out, int cap) {
  if (e->record.empty() || cap < (int)e->record.size()) return fail("ps_get_winner_record: no record / buffer too small");
  std::copy(e->record.begin(), e->record.end(), out); return 0;
}
int ps_get_trajectory(PsEngine* e, int sample, double* states, double* actions, double* times, double* residual, double* costs,
                      double* trace, double* total_return, int* failure) {
  if (sample < 0 || sample >= e->N) return fail("ps_get_trajectory: sample out of range");
  const Sample& s = e->samples[sample];
  const int ns = (int)e->state.size(), nu = e->m->nu, nr = e->task.num_residual, nt = 3 * e->task.num_trace, H = e->H;
  if (states) std::copy(s.states.begin(), s.states.begin() + (size_t)H * ns, states);
  if (actions) std::copy(s.actions.begin(), s.actions.begin() + (size_t)H * nu, actions);
  if (times) std::copy(s.times.begin(), s.times.begin() + H, times);
  if (residual) std::copy(s.residual.begin(), s.residual.begin() + (size_t)H * nr, residual);
  if (costs) std::copy(s.costs.begin(), s.costs.begin() + H, costs);
  if (trace && nt) std::copy(s.trace.begin(), s.trace.begin() + (size_t)H * nt, trace);
  if (total_return) *total_return = e->returns[sample];
  if (failure) *failure = e->failure[sample] & 1;
  return 0;
}
// K4 + K5 restated (csrc/ps_kernels_aux.cuh: rank_kernel, elite_stats_kernel): rank by (return, index), NaN last;
// elite mean; the reference's variance formula (cross_entropy/planner.cc:224-262); average and best return
int ps_elite_update(PsEngine* e, int n_candidates, int n_elite, double* mean_knots, double* variance, double* avg_return,
                    double* best_return, int* order) {
  if (e->N <= 0) return fail("ps_elite_update: no rollouts yet");
  if (n_candidates < 1 || n_candidates > e->N || n_elite < 1 || n_elite > n_candidates) return fail("ps_elite_update: sizes out of range");
  const int nelem = e->P * e->m->nu;
  std::vector<double> r(n_candidates);
  for (int i = 0; i < n_candidates; i++) r[i] = e->returns[i] == e->returns[i] ? e->returns[i] : 1.0e300;
  std::vector<int> elite(n_elite, 0);
  for (int i = 0; i < n_candidates; i++) {
    int rank = 0;
    for (int k = 0; k < n_candidates; k++) rank += (r[k] < r[i]) || (r[k] == r[i] && k < i);
    if (rank < n_elite) elite[rank] = i;
  }
  double s = 0;
  for (int k = 0; k < n_elite; k++) s += e->returns[elite[k]];
  if (avg_return) *avg_return = s / n_elite;
  if (best_return) *best_return = e->returns[elite[0]];
  for (int el = 0; el < nelem; el++) {
    double sum = 0;
    for (int k = 0; k < n_elite; k++) sum += e->knots[(size_t)elite[k] * nelem + el];
    const double avg = sum * (1.0 / n_elite);
    const double diff = e->knots[(size_t)elite[0] * nelem + el] - avg;
    double var = 0;
    for (int k = 0; k < n_elite; k++) var += diff * diff / (n_elite - 1);
    if (mean_knots) mean_knots[el] = avg;
    if (variance) variance[el] = var;
  }
  if (order) for (int k = 0; k < n_elite; k++) order[k] = elite[k];
  e->launches += 2;
  return 0;
}

}  // extern "C"
