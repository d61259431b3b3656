// Build-time tool (not product): flattens a PsModel + PsTask with the engine's own host code (fill_model / convert_task,
// csrc/ps_host_model.h) and prints the compile-time integer description of a model-specialised kernel instantiation
// (one `struct Spec<Name>` of csrc/ps_spec_generated.h).  Driven by tools/gen_spec_header.py.
#include <cstdio>
#include <string>
#include <vector>
#include "../mujoco_mpc_b200/csrc/ps_host_model.h"

using namespace ps;

static void tab(std::string& o, const char* name, const int* v, int n) {
  // pure expression in i_: i_ == 0 ? v0 : i_ == 1 ? v1 : ... : v[n-1]
  std::string e; char b[64];
  if (n <= 0) e = "0";
  for (int i = 0; i < n; i++) {
    if (i + 1 < n) std::snprintf(b, sizeof b, "i_ == %d ? %d : ", i, v[i]); else std::snprintf(b, sizeof b, "%d", v[i]);
    e += b;
  }
  o += std::string("    PS_TAB(") + name + ", " + e + ");\n";
}

extern "C" int spec_dump(const PsModel* s, const PsTask* t, const char* ident, const char* title, char* out, int cap) {
  std::string why;
  std::vector<char> buf(sizeof(Constants<TinyT>));
  PairList pl = static_pairs(s);
  if (fill_model<TinyT>(s, pl, buf.data(), &why)) { std::snprintf(out, cap, "fill_model: %s", why.c_str()); return -1; }
  const DevModel<TinyT>& m = reinterpret_cast<Constants<TinyT>*>(buf.data())->m;
  std::vector<int> st(s->sensor_type, s->sensor_type + s->nsensor);
  DevTask d; int need_sub = 0;
  if (convert_task(t, st.data(), s->nsensor, &d, &need_sub, &why)) { std::snprintf(out, cap, "convert_task: %s", why.c_str()); return -1; }
  std::string o; char b[512];
  std::snprintf(b, sizeof b, "// %s\nstruct Spec%s {\n  static constexpr const char* kName = \"%s\";\n", title, ident, title); o += b;
  std::snprintf(b, sizeof b, "  static constexpr int NQ = %d, NV = %d, NU = %d, NBODY = %d, NJNT = %d, NGEOM = %d, NSITE = %d, NMOCAP = %d, NSENSOR = %d;\n",
                m.nq, m.nv, m.nu, m.nbody, m.njnt, m.ngeom, m.nsite, m.nmocap, m.nsensor); o += b;
  o += "  struct ModelInts {\n";
  std::snprintf(b, sizeof b, "    static constexpr int integrator = %d, cone = %d, iterations = %d, ls_iterations = %d, disableflags = %d;\n",
                m.integrator, m.cone, m.iterations, m.ls_iterations, m.disableflags); o += b;
  std::snprintf(b, sizeof b, "    static constexpr int nq = %d, nv = %d, nu = %d, na = %d, nbody = %d, njnt = %d, ngeom = %d, nsite = %d, nmocap = %d, nsensor = %d, npair = %d;\n",
                m.nq, m.nv, m.nu, m.na, m.nbody, m.njnt, m.ngeom, m.nsite, m.nmocap, m.nsensor, m.npair); o += b;
  std::snprintf(b, sizeof b, "    static constexpr int any_damping = %d, need_subtreevel = %d, nray = %d, ray_complete = %d, debug_serial = 0;\n",
                m.any_damping, need_sub, m.nray, m.ray_complete); o += b;
  tab(o, "ray_geom", m.ray_geom, m.nray);
#define T1(f, n) tab(o, #f, m.f, n);
  T1(body_parentid, m.nbody) T1(body_rootid, m.nbody) T1(body_jntnum, m.nbody) T1(body_jntadr, m.nbody) T1(body_dofnum, m.nbody)
  T1(body_dofadr, m.nbody) T1(body_mocapid, m.nbody) T1(body_weldid, m.nbody)
  T1(jnt_type, m.njnt) T1(jnt_qposadr, m.njnt) T1(jnt_dofadr, m.njnt) T1(jnt_bodyid, m.njnt) T1(jnt_limited, m.njnt)
  T1(dof_bodyid, m.nv) T1(dof_jntid, m.nv) T1(dof_parentid, m.nv)
  T1(geom_type, m.ngeom) T1(geom_condim, m.ngeom) T1(geom_bodyid, m.ngeom) T1(geom_priority, m.ngeom) T1(geom_group, m.ngeom) T1(geom_used, m.ngeom)
  T1(pair_geom1, m.npair) T1(pair_geom2, m.npair) T1(site_bodyid, m.nsite)
  T1(act_dyntype, m.nu) T1(act_gaintype, m.nu) T1(act_biastype, m.nu) T1(act_jntid, m.nu) T1(act_actadr, m.nu) T1(act_ctrllimited, m.nu)
  T1(act_forcelimited, m.nu) T1(act_actlimited, m.nu)
  T1(sensor_type, m.nsensor) T1(sensor_objtype, m.nsensor) T1(sensor_objid, m.nsensor)
#undef T1
  o += "  };\n  struct TaskInts {\n";
  std::snprintf(b, sizeof b, "    static constexpr int task_id = %d, num_residual = %d, num_term = %d, num_trace = %d;\n", d.task_id, d.num_residual, d.num_term, d.num_trace); o += b;
  tab(o, "dim_norm_residual", d.dim_norm_residual, d.num_term); tab(o, "num_norm_parameter", d.num_norm_parameter, d.num_term);
  tab(o, "norm", d.norm, d.num_term); tab(o, "sensor_id", d.sensor_id, 16); tab(o, "trace_sensor", d.trace_sensor, 4);
  o += "  };\n};\n";
  if ((int)o.size() + 1 > cap) return -2;
  std::snprintf(out, cap, "%s", o.c_str());
  return (int)o.size();
}
