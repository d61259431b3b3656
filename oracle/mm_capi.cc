// ORACLE (test infrastructure).  ctypes-facing C entry points over the CPU restatement.  Used only
// by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
#include <cstring>
#include <string>
#include "mm_planner.h"
#include "mm_math.h"

namespace mm {
void optimizePolicyCE(Planner* pl, int N, int horizon, const double* z, int nthread);
void optimizePolicyRobust(Planner* pl, int N, int horizon, const double* z, int ncandidates, int repetitions, const double* w, double xfrc_std, double xfrc_rate);
}
using namespace mm;

struct Handle { Planner pl; Data d; };

static void cb(const Model* m, Data* d, void* user) { residual(m, d, (const PsTask*)user, d->sensordata.data()); }

extern "C" {

void* mmo_create(const PsModel* m, const PsTask* t) {
  Handle* h = new Handle;
  h->pl.model = m; h->pl.task = *t;
  makeData(m, &h->d);
  h->pl.state.assign(m->nq + m->nv + m->na, 0); h->pl.mocap.assign(7 * m->nmocap, 0); h->pl.userdata.assign(m->nuserdata, 0);
  for (int i = 0; i < m->nmocap; i++) { copy(&h->pl.mocap[7 * i], &h->d.mocap_pos[3 * i], 3); copy(&h->pl.mocap[7 * i + 3], &h->d.mocap_quat[4 * i], 4); }
  return h;
}
void mmo_destroy(void* hp) { delete (Handle*)hp; }
void mmo_set_task(void* hp, const PsTask* t) { ((Handle*)hp)->pl.task = *t; }

static std::vector<double>* field(Data* d, const std::string& f) {
#define F(n) if (f == #n) return &d->n;
  F(qpos) F(qvel) F(act) F(ctrl) F(qacc) F(qacc_warmstart) F(mocap_pos) F(mocap_quat) F(sensordata)
  F(xpos) F(xquat) F(xmat) F(xipos) F(ximat) F(geom_xpos) F(geom_xmat) F(site_xpos) F(subtree_com) F(cdof) F(cinert)
  F(qM) F(qLD) F(cvel) F(qfrc_bias) F(qfrc_passive) F(qfrc_fluid) F(qfrc_actuator) F(qfrc_smooth) F(qacc_smooth)
  F(qfrc_constraint) F(actuator_force) F(efc_J) F(efc_pos) F(efc_D) F(efc_R) F(efc_aref) F(efc_force) F(efc_vel)
  F(qDeriv) F(subtree_linvel) F(subtree_angmom) F(act_dot) F(xfrc_applied)
#undef F
  return nullptr;
}
int mmo_set(void* hp, const char* f, const double* v, int n) {
  Handle* h = (Handle*)hp;
  if (!std::strcmp(f, "time")) { h->d.time = v[0]; return 1; }
  auto* x = field(&h->d, f); if (!x || (int)x->size() < n) return -1;
  copy(x->data(), v, n); return n;
}
int mmo_get(void* hp, const char* f, double* out, int cap) {
  Handle* h = (Handle*)hp;
  if (!std::strcmp(f, "time")) { out[0] = h->d.time; return 1; }
  if (!std::strcmp(f, "nefc")) { out[0] = h->d.nefc; return 1; }
  if (!std::strcmp(f, "ncon")) { out[0] = h->d.ncon; return 1; }
  if (!std::strcmp(f, "solver_iter")) { out[0] = h->d.solver_iter; return 1; }
  if (!std::strcmp(f, "dbg_line_evals")) { out[0] = (double)mm::g_dbg_line_evals; return 1; }
  if (!std::strcmp(f, "warnings")) { double s = 0; for (int i = 0; i < N_WARN; i++) s += h->d.warning[i]; out[0] = s; return 1; }
  if (!std::strcmp(f, "contact")) {  // [dist, pos3, frame9, dim, geom1, geom2] per contact
    int n = 0; for (auto& c : h->d.contact) { if (n + 16 > cap) break; out[n] = c.dist; copy(out + n + 1, c.pos, 3); copy(out + n + 4, c.frame, 9); out[n + 13] = c.dim; out[n + 14] = c.geom1; out[n + 15] = c.geom2; n += 16; }
    return n;
  }
  auto* x = field(&h->d, f); if (!x) return -1;
  int n = std::min<int>(cap, (int)x->size()); copy(out, x->data(), n); return (int)x->size();
}
void mmo_reset_data(void* hp) { Handle* h = (Handle*)hp; resetData(h->pl.model, &h->d); }
void mmo_forward(void* hp) { Handle* h = (Handle*)hp; forward(h->pl.model, &h->d, cb, &h->pl.task); }
void mmo_step(void* hp) { Handle* h = (Handle*)hp; step(h->pl.model, &h->d, cb, &h->pl.task); }
double mmo_cost(void* hp, const double* res) { return costValue(&((Handle*)hp)->pl.task, res); }
void mmo_cost_terms(void* hp, const double* res, double* terms) { costTerms(&((Handle*)hp)->pl.task, res, terms, true); }
double mmo_norm(const double* x, const double* params, int n, int type) { return normValue(x, params, n, type); }
void mmo_spline_sample(const double* times, const double* knots, int P, int dim, int interp, double t, double* out) { splineSample(times, knots, P, dim, interp, t, out); }

// ---- planner level
void mmo_planner_config(void* hp, int num_spline_points, int interpolation, double s0, double s1) {
  Planner& p = ((Handle*)hp)->pl; p.num_spline_points = num_spline_points; p.interpolation = interpolation;
  p.noise_exploration[0] = s0; p.noise_exploration[1] = s1;
}
void mmo_set_sliding_plan(void* hp, int on) { ((Handle*)hp)->pl.sliding_plan = on; }
void mmo_set_state(void* hp, const double* state, const double* mocap, const double* userdata, double time) {
  Planner& p = ((Handle*)hp)->pl;
  copy(p.state.data(), state, (int)p.state.size());
  if (mocap) copy(p.mocap.data(), mocap, (int)p.mocap.size());
  if (userdata) copy(p.userdata.data(), userdata, (int)p.userdata.size());
  p.time = time;
}
void mmo_set_policy(void* hp, const double* times, const double* knots, int P, int interp) {
  Planner& p = ((Handle*)hp)->pl; int nu = p.model->nu;
  p.policy.times.assign(times, times + P); p.policy.knots.assign(knots, knots + P * nu); p.policy.interp = interp;
  p.previous_policy = p.policy; p.candidate.clear(); p.winner = 0;
}
void mmo_optimize(void* hp, int N, int H, const double* z, const unsigned char* use_std2, int nthread) {
  optimizePolicy(&((Handle*)hp)->pl, N, H, z, use_std2, nthread);
}
// cross-entropy planner: std_min / n_elite, variance in/out, one iteration, elite order and nominal return
void mmo_ce_config(void* hp, double std_min, int n_elite) { Planner& p = ((Handle*)hp)->pl; p.std_min = std_min; p.n_elite = n_elite; }
void mmo_ce_set_variance(void* hp, const double* v, int n) { ((Handle*)hp)->pl.variance.assign(v, v + n); }
int mmo_ce_get_variance(void* hp, double* v, int cap) { Planner& p = ((Handle*)hp)->pl; int n = std::min<int>(cap, (int)p.variance.size()); copy(v, p.variance.data(), n); return (int)p.variance.size(); }
void mmo_optimize_ce(void* hp, int N, int H, const double* z, int nthread) { optimizePolicyCE(&((Handle*)hp)->pl, N, H, z, nthread); }
int mmo_ce_order(void* hp, int* order, int cap) { Planner& p = ((Handle*)hp)->pl; int n = std::min<int>(cap, (int)p.trajectory_order.size()); for (int i = 0; i < n; i++) order[i] = p.trajectory_order[i]; return n; }
double mmo_ce_avg_return(void* hp) { return ((Handle*)hp)->pl.avg_return; }
double mmo_ce_nominal_return(void* hp) { return ((Handle*)hp)->pl.nominal_trajectory.total_return; }
// robust planner: one iteration; scores[ncandidates] and order[ncandidates] out; returns the picked candidate
int mmo_optimize_robust(void* hp, int N, int H, const double* z, int ncandidates, int repetitions, const double* w, double xfrc_std,
                        double xfrc_rate, double* scores, int* order, double* noisy_returns, unsigned char* noisy_failure) {
  Planner& p = ((Handle*)hp)->pl;
  optimizePolicyRobust(&p, N, H, z, ncandidates, repetitions, w, xfrc_std, xfrc_rate);
  for (size_t c = 0; c < p.robust_score.size(); c++) { if (scores) scores[c] = p.robust_score[c]; if (order) order[c] = p.trajectory_order[c]; }
  for (size_t k = 0; k < p.robust_trajectory.size(); k++) { if (noisy_returns) noisy_returns[k] = p.robust_trajectory[k].total_return; if (noisy_failure) noisy_failure[k] = p.robust_trajectory[k].failure; }
  return p.robust_best;
}
int mmo_winner(void* hp) { return ((Handle*)hp)->pl.winner; }
double mmo_improvement(void* hp) { return ((Handle*)hp)->pl.improvement; }
int mmo_get_policy(void* hp, int which /*0 policy,1 previous,2+i candidate i*/, double* times, double* knots) {
  Planner& p = ((Handle*)hp)->pl; const Policy& q = which == 0 ? p.policy : which == 1 ? p.previous_policy : p.candidate[which - 2];
  copy(times, q.times.data(), q.P()); copy(knots, q.knots.data(), (int)q.knots.size()); return q.P();
}
void mmo_action(void* hp, double t, int use_previous, double* action) {
  Planner& p = ((Handle*)hp)->pl; policyAction(p.model, use_previous ? p.previous_policy : p.policy, t, action);
}
void mmo_get_returns(void* hp, double* ret, unsigned char* fail) {
  Planner& p = ((Handle*)hp)->pl;
  for (size_t i = 0; i < p.trajectory.size(); i++) { ret[i] = p.trajectory[i].total_return; if (fail) fail[i] = p.trajectory[i].failure; }
}
void mmo_get_trajectory(void* hp, int i, double* states, double* actions, double* times, double* res, double* costs,
                        double* trace, double* total_return, int* failure) {
  const Trajectory& t = ((Handle*)hp)->pl.trajectory[i];
  if (states) copy(states, t.states.data(), (int)t.states.size());
  if (actions) copy(actions, t.actions.data(), (int)t.actions.size());
  if (times) copy(times, t.times.data(), (int)t.times.size());
  if (res) copy(res, t.residual.data(), (int)t.residual.size());
  if (costs) copy(costs, t.costs.data(), (int)t.costs.size());
  if (trace) copy(trace, t.trace.data(), (int)t.trace.size());
  if (total_return) *total_return = t.total_return;
  if (failure) *failure = t.failure;
}

}  // extern "C"

#ifdef MM_COUNT_FLOPS
// libmm_oracle_flops.so only (oracle/flopcount.h): the tally of FP64 operations since the last reset
FlopCounter g_flops;
extern "C" void mmo_flops_reset(void) { g_flops = FlopCounter(); }
extern "C" void mmo_flops(long long* out /*[5]: add, mul, div, sqrt, special*/) {
  out[0] = g_flops.add; out[1] = g_flops.mul; out[2] = g_flops.div; out[3] = g_flops.sqrt_; out[4] = g_flops.special;
}
#endif
