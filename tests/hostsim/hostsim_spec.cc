// TEST INFRASTRUCTURE (not product): the model-specialised lean instantiations (csrc/ps_variant_tiny_spec.cu) compiled with
// the HOST compiler under the same defines, so that their code path -- compile-time tables, bounded dof-tree walks, the
// split row storage -- is diffed against the oracle in the GPU-less container.
#define PS_THREAD_INLINE 1
#define PS_SPEC_BUILD 1
#define PS_ROLLOUT_SINGLE_SITE 1
#include <cstring>
#include <string>
#include <vector>
#include "../../mujoco_mpc_b200/csrc/ps_host_model.h"

using namespace ps;

template <class T>
static int run_spec(const PsModel* m, const PsTask* t, const PairList& pl, const double* state, const double* mocap, double time,
                    const double* times, const double* knots, int P, int interp, int H, double timestep,
                    double* states, double* actions, double* residual, double* costs, double* total, int* flags, std::string* why) {
  std::vector<char> buf(sizeof(Constants<T>) + 256);
  void* c = (void*)(((uintptr_t)buf.data() + 127) & ~(uintptr_t)127);
  std::memset(c, 0, sizeof(Constants<T>));
  if (fill_model_spec<T>(m, pl, c, why)) return -1;
  if (timestep > 0) set_options_spec<T>(c, timestep, PS_INT_EULER);
  DevTask dt; int need = 0;
  if (convert_task(t, m->sensor_type, m->nsensor, &dt, &need, why)) return -2;
  if (!task_matches_spec<T>(dt, need)) { *why = "cost specification differs from the specialisation"; return -1; }
  set_task_spec<T>(c, dt, need);
  set_plan<T>(c, state, m->nq + m->nv + m->na, mocap, m->nmocap, time, times, P, interp, H);
  Constants<T>* k = (Constants<T>*)c;
  std::vector<double> tm(H), tr((size_t)H * 3 * (t->num_trace ? t->num_trace : 1));
  std::vector<double> warm(64, 0.0);
  Writeback wb{states, actions, residual, costs, tm.data(), tr.data(), warm.data()};
  const SpecTaskView<typename T::Topo> tv(k->t);
  *total = rollout_sample<T, true, false, SpecTaskView<typename T::Topo>>(k->m, tv, k->p, knots, 1, 0, wb, flags);
  return 0;
}

extern "C" int hs_rollout_spec(const PsModel* m, const PsTask* t, const double* state, const double* mocap, double time,
                               const double* times, const double* knots, int P, int interp, int H, double timestep,
                               double* states, double* actions, double* residual, double* costs, double* total, int* flags,
                               char* variant, int cap) {
  PairList pl = static_pairs(m);
  std::string why, all;
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
#define TRY(N) r = run_spec<SpecTraits<Spec##N>>(m, t, pl, state, mocap, time, times, knots, P, interp, H, timestep, states, actions, residual, costs, total, flags, &why); \
  if (r == 0) { std::snprintf(variant, cap, "tiny_%s", #N); return 0; } all += why + "; ";
  PS_SPEC_LIST(TRY)
  std::snprintf(variant, cap, "%s", all.c_str());
  return -1;
}
