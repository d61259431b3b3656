// TEST INFRASTRUCTURE (not product): compiles the engine's kernel-source headers
// (mujoco_mpc_b200/csrc/ps_*.cuh) with the HOST compiler so the exact device code path can be
// diffed against the oracle in this GPU-less container before a kernel ever runs on a B200.
// Nothing in the product links or loads this; the shipped engine is CUDA-only.
#include <cstring>
#include <string>
#include <vector>
#include "../../mujoco_mpc_b200/csrc/ps_host_model.h"
#include "../../mujoco_mpc_b200/csrc/ps_coop_rollout.cuh"

using namespace ps;

// optional force perturbation of the NEXT hs_rollout call (thread-per-rollout path): standard-normal tensor
// [H-1][6*nbody] + OU parameters (mjpc/trajectory.cc:147-155)
static const double* g_xfrc_z = nullptr; static double g_xfrc_std = 0, g_xfrc_rate = 1;
extern "C" void hs_set_xfrc(const double* z, double xfrc_std, double xfrc_rate) { g_xfrc_z = z; g_xfrc_std = xfrc_std; g_xfrc_rate = xfrc_rate; }

template <class T>
static int run(const PsModel* m, const PsTask* t, const PairList& pl, const double* state, const double* mocap, double time,
               const double* times, const double* knots, int P, int interp, int H, double timestep, int integrator,
               double* states, double* actions, double* residual, double* costs, double* total, int* flags, std::string* why) {
  std::vector<char> buf(sizeof(Constants<T>) + 256);
  void* c = (void*)(((uintptr_t)buf.data() + 127) & ~(uintptr_t)127);
  std::memset(c, 0, sizeof(Constants<T>));
  if (fill_model<T>(m, pl, c, why)) return -1;
  if (timestep > 0) set_options<T>(c, timestep, integrator);
  DevTask dt; int need = 0;
  if (convert_task(t, m->sensor_type, m->nsensor, &dt, &need, why)) return -2;
  set_task<T>(c, dt, need);
  set_plan<T>(c, state, m->nq + m->nv + m->na, mocap, m->nmocap, time, times, P, interp, H);
  Constants<T>* k = (Constants<T>*)c;
  const int ns = m->nq + m->nv + m->na, nr = t->num_residual, nu = m->nu;
  std::vector<double> tm(H), tr((size_t)H * 3 * (t->num_trace ? t->num_trace : 1));
  std::vector<double> warm(64, 0.0);
  Writeback wb{states, actions, residual, costs, tm.data(), tr.data(), warm.data()};
  (void)ns; (void)nr; (void)nu;
  if (g_xfrc_z && g_xfrc_std > 0) {      // the instantiation robust_rollout_kernel runs
    XfrcNoise xf{g_xfrc_z, std::exp(-k->m.timestep / g_xfrc_rate), 0.0};
    xf.scale = g_xfrc_std * std::sqrt(1 - xf.rate * xf.rate);
    *total = rollout_sample<T, true, true>(k->m, k->t, k->p, knots, 1, 0, wb, flags, &xf);
    return 0;
  }
  *total = rollout_sample<T, true>(k->m, k->t, k->p, knots, 1, 0, wb, flags);
  return 0;
}

// the warp-per-rollout engine with a lane count of 1 (ps_coop.cuh): same phases, same arithmetic
template <class T>
static int run_coop(const PsModel* m, const PsTask* t, const PairList& pl, const double* state, const double* mocap, double time,
                    const double* times, const double* knots, int P, int interp, int H, double timestep, int integrator,
                    double* states, double* actions, double* residual, double* costs, double* total, int* flags, std::string* why) {
  if (integrator != PS_INT_EULER || m->opt_density > 0 || m->opt_viscosity > 0) { *why = "not a cooperative-engine model"; return -1; }
  std::vector<char> buf(sizeof(Constants<T>) + 256);
  void* c = (void*)(((uintptr_t)buf.data() + 127) & ~(uintptr_t)127);
  std::memset(c, 0, sizeof(Constants<T>));
  if (fill_model<T>(m, pl, c, why)) return -1;
  if (timestep > 0) set_options<T>(c, timestep, integrator);
  DevTask dt; int need = 0;
  if (convert_task(t, m->sensor_type, m->nsensor, &dt, &need, why)) return -2;
  set_task<T>(c, dt, need);
  set_plan<T>(c, state, m->nq + m->nv + m->na, mocap, m->nmocap, time, times, P, interp, H);
  Constants<T>* k = (Constants<T>*)c;
  std::vector<double> tm(H), tr((size_t)H * 3 * (t->num_trace ? t->num_trace : 1));
  std::vector<double> warm(64, 0.0);
  Writeback wb{states, actions, residual, costs, tm.data(), tr.data(), warm.data()};
  std::vector<char> dbuf(sizeof(CoopData<T>) + 256);
  CoopData<T>* d = (CoopData<T>*)(((uintptr_t)dbuf.data() + 127) & ~(uintptr_t)127);
  std::memset((void*)d, 0xFF, sizeof(CoopData<T>));   // poison: a read-before-write shows up as NaN
  Lanes L;
  if (g_xfrc_z && g_xfrc_std > 0) {      // the instantiation coop_robust_kernel runs (with one lane here)
    XfrcNoise xf{g_xfrc_z, std::exp(-k->m.timestep / g_xfrc_rate), 0.0};
    xf.scale = g_xfrc_std * std::sqrt(1 - xf.rate * xf.rate);
    std::vector<double> applied(6 * T::NBODY, -1.0e300);     // poison: reset() must clear it
    CoPerturbation<T, true> pert{applied.data(), &xf};
    *total = co_rollout<T, true, Lanes, CoPerturbation<T, true>>(L, k->m, k->t, k->p, knots, 1, 0, wb, *d, flags, pert);
    return 0;
  }
  *total = co_rollout<T, true>(L, k->m, k->t, k->p, knots, 1, 0, wb, *d, flags);
  return 0;
}

extern "C" int hs_rollout_coop(const PsModel* m, const PsTask* t, const double* state, const double* mocap, double time,
                               const double* times, const double* knots, int P, int interp, int H, double timestep,
                               int integrator, double* states, double* actions, double* residual, double* costs, double* total,
                               int* flags, char* variant, int cap) {
  PairList pl = static_pairs(m);
  std::string why;
  std::vector<double> mo(7 * (m->nmocap ? m->nmocap : 1), 0.0);
  if (!mocap) {
    for (int b = 0; b < m->nbody; b++) if (m->body_mocapid[b] >= 0) {
      int id = m->body_mocapid[b];
      for (int k = 0; k < 3; k++) mo[7 * id + k] = m->body_pos[3 * b + k];
      for (int k = 0; k < 4; k++) mo[7 * id + 3 + k] = m->body_quat[4 * b + k];
    }
    mocap = mo.data();
  }
  int r;
#define TRYC(T, name) r = run_coop<T>(m, t, pl, state, mocap, time, times, knots, P, interp, H, timestep, integrator, states, actions, residual, costs, total, flags, &why); \
  if (r == 0) { std::snprintf(variant, cap, "%s (sizeof CoopData %zu, Constants %zu)", name, sizeof(CoopData<T>), sizeof(Constants<T>)); return 0; } if (r == -2) { std::snprintf(variant, cap, "%s", why.c_str()); return -2; }
  TRYC(PlanarCoopT, "planar_coop")
  // legged class: as the engine does, a rollout that overflows the scoring capacities (flag bit 1) is rolled again at the
  // large-capacity instantiation of the fix-up pass (coop_fixup_kernel<LeggedXLT>)
  r = run_coop<LeggedCoopT>(m, t, pl, state, mocap, time, times, knots, P, interp, H, timestep, integrator, states, actions, residual, costs, total, flags, &why);
  if (r == 0 && (*flags & 2)) {
    r = run_coop<LeggedXLT>(m, t, pl, state, mocap, time, times, knots, P, interp, H, timestep, integrator, states, actions, residual, costs, total, flags, &why);
    if (r == 0) { std::snprintf(variant, cap, "legged_coop + legged_xl fix-up (sizeof CoopData %zu)", sizeof(CoopData<LeggedXLT>)); return 0; }
  }
  if (r == 0) { std::snprintf(variant, cap, "legged_coop (sizeof CoopData %zu, Constants %zu)", sizeof(CoopData<LeggedCoopT>), sizeof(Constants<LeggedCoopT>)); return 0; }
  if (r == -2) { std::snprintf(variant, cap, "%s", why.c_str()); return -2; }
  std::snprintf(variant, cap, "%s", why.c_str());
  return -1;
}

extern "C" int hs_rollout(const PsModel* m, const PsTask* t, const double* state, const double* mocap, double time,
                          const double* times, const double* knots /*[P][nu]*/, int P, int interp, int H, double timestep,
                          int integrator, double* states /*[H][ns]*/, double* actions /*[H][nu]*/, double* residual /*[H][nr]*/,
                          double* costs /*[H]*/, double* total, int* flags, char* variant, int cap) {
  PairList pl = static_pairs(m);
  std::string why;
  std::vector<double> mo(7 * (m->nmocap ? m->nmocap : 1), 0.0);
  if (!mocap) {
    for (int b = 0; b < m->nbody; b++) if (m->body_mocapid[b] >= 0) {
      int id = m->body_mocapid[b];
      for (int k = 0; k < 3; k++) mo[7 * id + k] = m->body_pos[3 * b + k];
      for (int k = 0; k < 4; k++) mo[7 * id + 3 + k] = m->body_quat[4 * b + k];
    }
    mocap = mo.data();
  }
  int r;
#define TRY(T, name) r = run<T>(m, t, pl, state, mocap, time, times, knots, P, interp, H, timestep, integrator, states, actions, residual, costs, total, flags, &why); \
  if (r == 0) { std::snprintf(variant, cap, "%s", name); return 0; } if (r == -2) { std::snprintf(variant, cap, "%s", why.c_str()); return -2; }
  TRY(TinyT, "tiny") TRY(PlanarT, "planar") TRY(LeggedT, "legged")
  std::snprintf(variant, cap, "%s", why.c_str());
  return -1;
}
