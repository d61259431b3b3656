// Host-side flattening of the C-ABI model / task descriptions (include/ps_model.h, ps_engine.h) into
// the fixed-capacity device PODs of ps_types.cuh, plus the static part of the collision filter.
// Plain C++ (no CUDA): shared by the engine (ps_engine.cu) and by the host-compiled kernel-source
// checker in tests/hostsim.
#pragma once
#include <cmath>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/ps_engine.h"
#include <cstdlib>
#include "ps_rollout.cuh"
#ifndef PS_NO_SPEC_HEADER      // tools/spec_dump.cc (the generator of that header) compiles without it
#include "ps_spec_generated.h"
#endif

namespace ps {

// ---------------------------------------------------------------- engine instantiations
//                        NQ NV NU NA NBODY NJNT NGEOM NSITE NMOCAP NPAIR MAXCON MAXEFC NSENSOR ELLIPTIC
typedef Traits<2, 2, 2, 0, 4, 2, 8, 2, 1, 0, 0, 4, 12, false> TinyT;          // particle, cartpole, acrobot
typedef Traits<9, 9, 6, 5, 9, 9, 18, 6, 1, 8, 14, 64, 20, false> PlanarT;    // swimmer, walker
typedef Traits<19, 18, 12, 0, 16, 13, 44, 8, 2, 40, 16, 96, 16, true> LeggedT;  // quadruped (free joint, elliptic cones)

// warp-per-rollout instantiations: + NCH (dofs one contact touches), NJROW (compressed Jacobian rows)
typedef Traits<2, 2, 2, 0, 4, 2, 8, 2, 1, 0, 0, 4, 12, false, 2, 4> TinyCoopT;
typedef Traits<9, 9, 6, 5, 9, 9, 18, 6, 1, 8, 14, 64, 20, false, 9, 52> PlanarCoopT;
typedef Traits<19, 18, 12, 0, 16, 13, 44, 8, 2, 40, 24, 120, 16, true, 12, 84> LeggedCoopT;
// capacity-overflow re-rolls (csrc/ps_coop_rollout.cuh: coop_fixup_kernel): the same model capacities -- so the same
// Constants layout -- with room for 64 contacts / 256 constraint rows (a fallen robot: 39 / 141 in the capacity survey)
typedef Traits<19, 18, 12, 0, 16, 13, 44, 8, 2, 40, 64, 256, 16, true, 12, 192> LeggedXLT;
static_assert(sizeof(Constants<LeggedXLT>) == sizeof(Constants<LeggedCoopT>) && sizeof(Constants<LeggedXLT>) == sizeof(Constants<LeggedT>),
              "the legged instantiations share one Constants layout (the fix-up pass reads whichever copy is resident)");

struct PairList { std::vector<int> g1, g2; int unsupported = 0; };

inline int weld_body(const PsModel* m, int b) { while (b > 0 && m->body_dofnum[b] == 0) b = m->body_parentid[b]; return b; }

inline bool pair_supported(int t1, int t2) {
  if (t1 == PS_GEOM_PLANE) return t2 == PS_GEOM_SPHERE || t2 == PS_GEOM_CAPSULE || t2 == PS_GEOM_CYLINDER || t2 == PS_GEOM_BOX;
  if (t1 == PS_GEOM_SPHERE) return t2 == PS_GEOM_SPHERE || t2 == PS_GEOM_CAPSULE;
  return t1 == PS_GEOM_CAPSULE && t2 == PS_GEOM_CAPSULE;
}

// inline part of the collision filter (contype/conaffinity, same/welded body, parent-child)
inline PairList static_pairs(const PsModel* m) {
  PairList pl;
  if (m->opt_disableflags & (PS_DSBL_CONTACT | PS_DSBL_CONSTRAINT)) return pl;
  for (int ga = 0; ga < m->ngeom; ga++)
    for (int gb = ga + 1; gb < m->ngeom; gb++) {
      const int b1 = m->geom_bodyid[ga], b2 = m->geom_bodyid[gb];
      if (b1 == b2) continue;
      if (!((m->geom_contype[ga] & m->geom_conaffinity[gb]) || (m->geom_contype[gb] & m->geom_conaffinity[ga]))) continue;
      const int w1 = weld_body(m, b1), w2 = weld_body(m, b2);
      if (w1 == w2) continue;
      if (!(m->opt_disableflags & PS_DSBL_FILTERPARENT) && w1 != 0 && w2 != 0)
        if (weld_body(m, m->body_parentid[w1]) == w2 || weld_body(m, m->body_parentid[w2]) == w1) continue;
      int g1 = ga, g2 = gb;
      if (m->geom_type[g1] > m->geom_type[g2]) { int t = g1; g1 = g2; g2 = t; }
      if (m->geom_type[g1] == PS_GEOM_PLANE && m->geom_type[g2] == PS_GEOM_PLANE) continue;
      if (!pair_supported(m->geom_type[g1], m->geom_type[g2])) { pl.unsupported++; continue; }
      pl.g1.push_back(g1); pl.g2.push_back(g2);
    }
  return pl;
}

// Ground() (utilities.cc:538-556) ray-casts against every group-0 geom.  The device carries static
// plane / sphere / box scenery; this reports whether that covers the model's whole group 0.
inline bool ray_scenery_ok(const PsModel* s) {
  int n = 0;
  for (int g = 0; g < s->ngeom; g++) {
    if (s->geom_group[g] != 0) continue;
    const int tp = s->geom_type[g];
    if (weld_body(s, s->geom_bodyid[g]) != 0 || !(tp == PS_GEOM_PLANE || tp == PS_GEOM_SPHERE || tp == PS_GEOM_BOX)) return false;
    if (++n > kMaxRayGeom) return false;
  }
  return n > 0;
}

template <class T>
inline bool track_geom(DevModel<T>& m, int g) {
  if (g < 0 || g >= m.ngeom) return false;
  if (m.geom_trackslot[g] >= 0) return true;
  if (m.ntrack >= kMaxTrackGeom) return false;
  m.geom_trackslot[g] = (short)m.ntrack; m.track_geom[m.ntrack++] = (short)g;
  return true;
}

// Schedules of the warp-cooperative engine: every table lists independent work items of one phase in
// the order that reproduces the sequential recursions of the thread-per-rollout code bit for bit.
template <class T>
inline int build_schedules(const PsModel* s, DevModel<T>& m, std::string* why) {
  const int nb = s->nbody, nv = s->nv;
  std::vector<int> depth(nb, 0);
  int maxd = 0;
  for (int b = 1; b < nb; b++) { depth[b] = depth[s->body_parentid[b]] + 1; if (depth[b] > maxd) maxd = depth[b]; }
  m.nlevel = maxd + 1;
  int at = 0;
  for (int l = 0; l <= maxd; l++) { m.level_adr[l] = (short)at; for (int b = 0; b < nb; b++) if (depth[b] == l) m.level_body[at++] = (short)b; }
  m.level_adr[maxd + 1] = (short)at;
  at = 0;
  for (int p = 0; p < nb; p++) { m.child_adr[p] = (short)at; for (int c = nb - 1; c > 0; c--) if (s->body_parentid[c] == p) m.child_list[at++] = (short)c; }
  m.child_adr[nb] = (short)at;
  std::vector<int> dd(nv > 0 ? nv : 1, 0); int maxdd = -1;
  for (int i = 0; i < nv; i++) { const int p = s->dof_parentid[i]; dd[i] = p < 0 ? 0 : dd[p] + 1; if (dd[i] > maxdd) maxdd = dd[i]; }
  m.ndoflevel = maxdd + 1; at = 0;
  for (int l = 0; l <= maxdd; l++) { m.doflevel_adr[l] = (short)at; for (int i = 0; i < nv; i++) if (dd[i] == l) m.doflevel_dof[at++] = (short)i; }
  m.doflevel_adr[maxdd + 1] = (short)at;
  auto is_anc = [&](int a, int i) { for (int j = s->dof_parentid[i]; j >= 0; j = s->dof_parentid[j]) if (j == a) return true; return false; };
  int ad = 0, aa = 0, fp = 0, mp = 0;
  for (int i = 0; i < nv; i++) {
    m.dofdesc_adr[i] = (short)ad;
    for (int k = nv - 1; k > i; k--) if (is_anc(i, k)) { if (ad >= T::NTRI) { *why = "dof tree too dense"; return -1; } m.dofdesc[ad++] = (short)k; }
    m.dofanc_adr[i] = (short)aa;
    for (int j = s->dof_parentid[i]; j >= 0; j = s->dof_parentid[j]) { if (aa >= T::NTRI) { *why = "dof tree too deep"; return -1; } m.dofanc[aa++] = (short)j; }
    m.fpair_adr[i] = (short)fp;
    for (int a = s->dof_parentid[i]; a >= 0; a = s->dof_parentid[a])
      for (int j = a; j >= 0; j = s->dof_parentid[j]) {
        if (fp >= T::NFPAIR) { *why = "elimination schedule overflow"; return -1; }
        m.fpair_i[fp] = (unsigned char)a; m.fpair_j[fp] = (unsigned char)j; fp++;
      }
    for (int j = i; j >= 0; j = s->dof_parentid[j]) { m.mpair_i[mp] = (unsigned char)i; m.mpair_j[mp] = (unsigned char)j; mp++; }
  }
  m.dofdesc_adr[nv] = (short)ad; m.dofanc_adr[nv] = (short)aa; m.fpair_adr[nv] = (short)fp; m.nmpair = mp;
  for (int a = 0, w = 0; a < T::NV; a++) for (int b = 0; b <= a; b++, w++) { m.tri_a[w] = (unsigned char)a; m.tri_b[w] = (unsigned char)b; }
  m.nfrictionrow = 0;
  for (int i = 0; i < nv; i++) {
    m.dof_frictionrow[i] = -1;
    if (s->dof_frictionloss[i] > 0) { m.dof_frictionrow[i] = (short)m.nfrictionrow; m.frictionrow_dof[m.nfrictionrow++] = (short)i; }
  }
  return 0;
}

template <class T>
inline int fill_model(const PsModel* s, const PairList& pl, void* out, std::string* why) {
  DevModel<T>& m = reinterpret_cast<Constants<T>*>(out)->m;
  std::memset(&m, 0, sizeof(m));
#define CAP(field, cap) if (s->field > cap) { *why = std::string(#field) + " exceeds capacity"; return -1; }
  CAP(nq, T::NQ) CAP(nv, T::NV) CAP(nu, T::NU) CAP(na, T::NA) CAP(nbody, T::NBODY) CAP(njnt, T::NJNT) CAP(ngeom, T::NGEOM)
  CAP(nsite, T::NSITE) CAP(nmocap, T::NMOCAP) CAP(nsensor, T::NSENSOR)
#undef CAP
  if ((int)pl.g1.size() > T::NPAIR) { *why = "collision pairs exceed capacity"; return -1; }
  if (s->opt_integrator == PS_INT_RK4) { *why = "RK4 integrator is not supported (Euler / implicit / implicitfast)"; return -1; }
  if (s->opt_solver != PS_SOL_NEWTON && !(s->opt_disableflags & PS_DSBL_CONSTRAINT)) { *why = "only the Newton solver is supported"; return -1; }
  if (pl.unsupported) { *why = "model has colliding geom pairs outside the supported narrow-phase set"; return -1; }
  if (s->opt_cone == PS_CONE_ELLIPTIC && !T::ELLIPTIC && !pl.g1.empty()) { *why = "elliptic cones not in this instantiation"; return -1; }
  if (!T::CONTACTS && !pl.g1.empty()) { *why = "contacts not in this instantiation"; return -1; }
  if (!T::QUAT) for (int j = 0; j < s->njnt; j++) if (s->jnt_type[j] != PS_JNT_HINGE && s->jnt_type[j] != PS_JNT_SLIDE) { *why = "free / ball joints not in this instantiation"; return -1; }
  if (T::LEAN) {     // features compiled out of the lean instantiation (ps_types.cuh: Traits::LEAN)
    if (s->nv != T::NV || s->nq != T::NQ) { *why = "lean instantiation needs nq == nv == its capacity"; return -1; }
    if (s->na) { *why = "activations not in the lean instantiation"; return -1; }
    if (s->opt_integrator != PS_INT_EULER) { *why = "lean instantiation is Euler only"; return -1; }
    if (s->opt_density > 0 || s->opt_viscosity > 0) { *why = "fluid forces not in the lean instantiation"; return -1; }
    for (int j = 0; j < s->njnt; j++) if (s->jnt_type[j] != PS_JNT_HINGE && s->jnt_type[j] != PS_JNT_SLIDE) { *why = "lean instantiation carries hinge / slide joints only"; return -1; }
    for (int i = 0; i < s->nu; i++) if (s->actuator_dyntype[i] != 0) { *why = "actuator dynamics not in the lean instantiation"; return -1; }
  }
  m.timestep = s->opt_timestep; m.impratio = s->opt_impratio; m.tolerance = s->opt_tolerance; m.ls_tolerance = s->opt_ls_tolerance;
  m.density = s->opt_density; m.viscosity = s->opt_viscosity; m.meaninertia = s->stat_meaninertia;
  for (int k = 0; k < 3; k++) { m.gravity[k] = s->opt_gravity[k]; m.wind[k] = s->opt_wind[k]; }
  m.integrator = s->opt_integrator; m.cone = s->opt_cone; m.iterations = s->opt_iterations; m.ls_iterations = s->opt_ls_iterations;
  m.disableflags = s->opt_disableflags;
  m.nq = s->nq; m.nv = s->nv; m.nu = s->nu; m.na = s->na; m.nbody = s->nbody; m.njnt = s->njnt; m.ngeom = s->ngeom;
  m.nsite = s->nsite; m.nmocap = s->nmocap; m.nsensor = s->nsensor; m.npair = (int)pl.g1.size();
#define CPY(dst, src, n) for (int i_ = 0; i_ < (n); i_++) m.dst[i_] = s->src[i_];
  CPY(qpos0, qpos0, s->nq) CPY(qpos_spring, qpos_spring, s->nq)
  CPY(body_parentid, body_parentid, s->nbody) CPY(body_rootid, body_rootid, s->nbody) CPY(body_jntnum, body_jntnum, s->nbody)
  CPY(body_jntadr, body_jntadr, s->nbody) CPY(body_dofnum, body_dofnum, s->nbody) CPY(body_dofadr, body_dofadr, s->nbody)
  CPY(body_mocapid, body_mocapid, s->nbody)
  for (int b = 0; b < s->nbody; b++) m.body_weldid[b] = weld_body(s, b);
  CPY(body_pos, body_pos, 3 * s->nbody) CPY(body_quat, body_quat, 4 * s->nbody) CPY(body_ipos, body_ipos, 3 * s->nbody)
  CPY(body_iquat, body_iquat, 4 * s->nbody) CPY(body_mass, body_mass, s->nbody) CPY(body_subtreemass, body_subtreemass, s->nbody)
  CPY(body_inertia, body_inertia, 3 * s->nbody) CPY(body_invweight0, body_invweight0, 2 * s->nbody)
  CPY(jnt_type, jnt_type, s->njnt) CPY(jnt_qposadr, jnt_qposadr, s->njnt) CPY(jnt_dofadr, jnt_dofadr, s->njnt)
  CPY(jnt_bodyid, jnt_bodyid, s->njnt) CPY(jnt_limited, jnt_limited, s->njnt) CPY(jnt_pos, jnt_pos, 3 * s->njnt)
  CPY(jnt_axis, jnt_axis, 3 * s->njnt) CPY(jnt_stiffness, jnt_stiffness, s->njnt) CPY(jnt_range, jnt_range, 2 * s->njnt)
  CPY(jnt_margin, jnt_margin, s->njnt) CPY(jnt_solref, jnt_solref, 2 * s->njnt) CPY(jnt_solimp, jnt_solimp, 5 * s->njnt)
  CPY(dof_bodyid, dof_bodyid, s->nv) CPY(dof_jntid, dof_jntid, s->nv) CPY(dof_parentid, dof_parentid, s->nv)
  CPY(dof_armature, dof_armature, s->nv) CPY(dof_damping, dof_damping, s->nv) CPY(dof_frictionloss, dof_frictionloss, s->nv)
  CPY(dof_solref, dof_solref, 2 * s->nv) CPY(dof_solimp, dof_solimp, 5 * s->nv) CPY(dof_invweight0, dof_invweight0, s->nv)
  CPY(geom_type, geom_type, s->ngeom) CPY(geom_condim, geom_condim, s->ngeom) CPY(geom_bodyid, geom_bodyid, s->ngeom)
  CPY(geom_priority, geom_priority, s->ngeom) CPY(geom_group, geom_group, s->ngeom)
  CPY(geom_size, geom_size, 3 * s->ngeom) CPY(geom_pos, geom_pos, 3 * s->ngeom) CPY(geom_quat, geom_quat, 4 * s->ngeom)
  CPY(geom_friction, geom_friction, 3 * s->ngeom) CPY(geom_solmix, geom_solmix, s->ngeom) CPY(geom_solref, geom_solref, 2 * s->ngeom)
  CPY(geom_solimp, geom_solimp, 5 * s->ngeom) CPY(geom_margin, geom_margin, s->ngeom) CPY(geom_gap, geom_gap, s->ngeom)
  CPY(geom_rbound, geom_rbound, s->ngeom)
  CPY(site_bodyid, site_bodyid, s->nsite) CPY(site_pos, site_pos, 3 * s->nsite) CPY(site_quat, site_quat, 4 * s->nsite)
  CPY(act_dyntype, actuator_dyntype, s->nu) CPY(act_gaintype, actuator_gaintype, s->nu) CPY(act_biastype, actuator_biastype, s->nu)
  CPY(act_actadr, actuator_actadr, s->nu) CPY(act_ctrllimited, actuator_ctrllimited, s->nu)
  CPY(act_forcelimited, actuator_forcelimited, s->nu) CPY(act_actlimited, actuator_actlimited, s->nu)
  CPY(act_dynprm, actuator_dynprm, 3 * s->nu) CPY(act_gainprm, actuator_gainprm, 3 * s->nu) CPY(act_biasprm, actuator_biasprm, 3 * s->nu)
  CPY(act_ctrlrange, actuator_ctrlrange, 2 * s->nu) CPY(act_forcerange, actuator_forcerange, 2 * s->nu)
  CPY(act_actrange, actuator_actrange, 2 * s->nu) CPY(act_gear, actuator_gear, s->nu)
  CPY(sensor_type, sensor_type, s->nsensor) CPY(sensor_objtype, sensor_objtype, s->nsensor) CPY(sensor_objid, sensor_objid, s->nsensor)
#undef CPY
  for (int i = 0; i < s->nu; i++) {
    if (s->actuator_trntype[i] != PS_TRN_JOINT) { *why = "only joint transmissions are supported"; return -1; }
    m.act_jntid[i] = s->actuator_trnid[2 * i];
  }
  for (int i = 0; i < s->nv; i++) if (s->dof_damping[i] > 0) m.any_damping = 1;
  for (int p = 0; p < m.npair; p++) { m.pair_geom1[p] = pl.g1[p]; m.pair_geom2[p] = pl.g2[p]; m.geom_used[pl.g1[p]] = 1; m.geom_used[pl.g2[p]] = 1; }
  for (int i = 0; i < s->nsensor; i++) if (s->sensor_objtype[i] == PS_OBJ_GEOM && s->sensor_objid[i] >= 0) m.geom_used[s->sensor_objid[i]] = 1;
  if (build_schedules<T>(s, m, why)) return -1;
  // Ground() ray candidates: group-0 geoms; the device carries static plane / sphere / box scenery only
  m.ray_complete = 1;
  for (int g = 0; g < s->ngeom; g++) {
    if (s->geom_group[g] != 0) continue;
    const int tp = s->geom_type[g];
    const bool ok = weld_body(s, s->geom_bodyid[g]) == 0 && (tp == PS_GEOM_PLANE || tp == PS_GEOM_SPHERE || tp == PS_GEOM_BOX);
    if (ok && m.nray < kMaxRayGeom) { m.ray_geom[m.nray++] = g; m.geom_used[g] = 1; } else m.ray_complete = 0;
  }
  // geoms whose world frame the cooperative engine keeps: ray scenery + sensor geoms (+ task geoms, set_task)
  for (int g = 0; g < s->ngeom; g++) m.geom_trackslot[g] = -1;
  m.ntrack = 0;
  for (int k = 0; k < m.nray; k++) track_geom(m, m.ray_geom[k]);
  for (int i = 0; i < s->nsensor; i++) if (s->sensor_objtype[i] == PS_OBJ_GEOM && s->sensor_objid[i] >= 0) track_geom(m, s->sensor_objid[i]);
  m.ntrack_model = m.ntrack;
  return 0;
}

#ifndef PS_NO_SPEC_HEADER
// ---------------------------------------------------------------- model-specialised instantiations (SpecTraits, ps_types.cuh)
// The model is flattened by the generic fill_model into a scratch DevModel of the same capacities; every integer of it
// must equal the Spec's compile-time value (else this instantiation is not for this model and the caller falls through
// to the generic lean kernel), the doubles are copied.
template <class T>
inline int fill_model_spec(const PsModel* s, const PairList& pl, void* out, std::string* why) {
  using B = typename T::Base; using Spec = typename T::Topo;
  if (s->nq != Spec::NQ || s->nv != Spec::NV || s->nu != Spec::NU || s->na != 0 || s->nbody != Spec::NBODY || s->njnt != Spec::NJNT ||
      s->ngeom != Spec::NGEOM || s->nsite != Spec::NSITE || s->nmocap != Spec::NMOCAP || s->nsensor != Spec::NSENSOR) {
    *why = std::string("sizes differ from the ") + Spec::kName + " specialisation"; return -1;
  }
  void* raw = std::aligned_alloc(128, (sizeof(Constants<B>) + 127) / 128 * 128);
  if (!raw) { *why = "out of memory"; return -1; }
  struct Free { void* p; ~Free() { std::free(p); } } guard{raw};
  if (fill_model<B>(s, pl, raw, why)) return -1;
  const DevModel<B>& g = reinterpret_cast<Constants<B>*>(raw)->m;
  DevModel<T>& m = reinterpret_cast<Constants<T>*>(out)->m;
  const typename Spec::ModelInts si{};
  const char* bad = nullptr;
#define CK(f) if (!bad && g.f != si.f) bad = #f;
#define CKT(f, n) for (int i_ = 0; i_ < (n) && !bad; i_++) if (g.f[i_] != si.f[i_]) bad = #f;
  CK(integrator) CK(cone) CK(iterations) CK(ls_iterations) CK(disableflags) CK(npair) CK(any_damping) CK(nray) CK(ray_complete)
  CKT(ray_geom, g.nray)
  CKT(body_parentid, g.nbody) CKT(body_rootid, g.nbody) CKT(body_jntnum, g.nbody) CKT(body_jntadr, g.nbody) CKT(body_dofnum, g.nbody)
  CKT(body_dofadr, g.nbody) CKT(body_mocapid, g.nbody) CKT(body_weldid, g.nbody)
  CKT(jnt_type, g.njnt) CKT(jnt_qposadr, g.njnt) CKT(jnt_dofadr, g.njnt) CKT(jnt_bodyid, g.njnt) CKT(jnt_limited, g.njnt)
  CKT(dof_bodyid, g.nv) CKT(dof_jntid, g.nv) CKT(dof_parentid, g.nv)
  CKT(geom_type, g.ngeom) CKT(geom_condim, g.ngeom) CKT(geom_bodyid, g.ngeom) CKT(geom_priority, g.ngeom) CKT(geom_group, g.ngeom) CKT(geom_used, g.ngeom)
  CKT(site_bodyid, g.nsite)
  CKT(act_dyntype, g.nu) CKT(act_gaintype, g.nu) CKT(act_biastype, g.nu) CKT(act_jntid, g.nu) CKT(act_actadr, g.nu) CKT(act_ctrllimited, g.nu)
  CKT(act_forcelimited, g.nu) CKT(act_actlimited, g.nu)
  CKT(sensor_type, g.nsensor) CKT(sensor_objtype, g.nsensor) CKT(sensor_objid, g.nsensor)
#undef CK
#undef CKT
  if (bad) { *why = std::string(bad) + " differs from the " + Spec::kName + " specialisation"; return -1; }
  std::memset(static_cast<void*>(&m), 0, sizeof(m));
  m.timestep = g.timestep; m.impratio = g.impratio; m.tolerance = g.tolerance; m.ls_tolerance = g.ls_tolerance;
  m.density = g.density; m.viscosity = g.viscosity; m.meaninertia = g.meaninertia;
#define CPD(f) static_assert(sizeof(m.f) == sizeof(g.f), #f); std::memcpy(m.f, g.f, sizeof(m.f));
  CPD(gravity) CPD(wind) CPD(qpos0) CPD(qpos_spring)
  CPD(body_pos) CPD(body_quat) CPD(body_ipos) CPD(body_iquat) CPD(body_mass) CPD(body_subtreemass) CPD(body_inertia) CPD(body_invweight0)
  CPD(jnt_pos) CPD(jnt_axis) CPD(jnt_stiffness) CPD(jnt_range) CPD(jnt_margin) CPD(jnt_solref) CPD(jnt_solimp)
  CPD(dof_armature) CPD(dof_damping) CPD(dof_frictionloss) CPD(dof_solref) CPD(dof_solimp) CPD(dof_invweight0)
  CPD(geom_size) CPD(geom_pos) CPD(geom_quat) CPD(geom_friction) CPD(geom_solmix) CPD(geom_solref) CPD(geom_solimp) CPD(geom_margin)
  CPD(geom_gap) CPD(geom_rbound) CPD(site_pos) CPD(site_quat)
  CPD(act_dynprm) CPD(act_gainprm) CPD(act_biasprm) CPD(act_ctrlrange) CPD(act_forcerange) CPD(act_actrange) CPD(act_gear)
#undef CPD
  return 0;
}

template <class T>
inline void set_options_spec(void* c, double timestep, int integrator) {
  (void)integrator;      // Euler only (Variant::euler_only): ps_set_options refuses anything else before it gets here
  reinterpret_cast<Constants<T>*>(c)->m.timestep = timestep;
}
template <class T> inline void set_debug_spec(void*, int) {}
template <class T>
inline void set_task_spec(void* c, const DevTask& t, int) { reinterpret_cast<Constants<T>*>(c)->t = t; }

// does the cost specification carry the integers this instantiation was compiled for?
template <class T>
inline int task_matches_spec(const DevTask& d, int need_subtreevel) {
  using Spec = typename T::Topo;
  const typename Spec::TaskInts ti{};
  if (d.task_id != ti.task_id || d.num_residual != ti.num_residual || d.num_term != ti.num_term || d.num_trace != ti.num_trace) return 0;
  if (need_subtreevel != Spec::ModelInts::need_subtreevel) return 0;
  for (int k = 0; k < d.num_term; k++)
    if (d.dim_norm_residual[k] != ti.dim_norm_residual[k] || d.num_norm_parameter[k] != ti.num_norm_parameter[k] || d.norm[k] != ti.norm[k]) return 0;
  for (int k = 0; k < 16; k++) if (d.sensor_id[k] != ti.sensor_id[k]) return 0;
  for (int k = 0; k < d.num_trace; k++) if (d.trace_sensor[k] != ti.trace_sensor[k]) return 0;
  return 1;
}

#endif  // PS_NO_SPEC_HEADER

template <class T>
inline void set_options(void* c, double timestep, int integrator) {
  DevModel<T>& m = reinterpret_cast<Constants<T>*>(c)->m; m.timestep = timestep; m.integrator = integrator;
}

template <class T>
inline void set_debug(void* c, int mask) { reinterpret_cast<Constants<T>*>(c)->m.debug_serial = mask; }

template <class T>
inline void set_task(void* c, const DevTask& t, int need_subtreevel) {
  Constants<T>* k = reinterpret_cast<Constants<T>*>(c); k->t = t; k->m.need_subtreevel = need_subtreevel;
  DevModel<T>& m = k->m;
  for (int i = m.ntrack_model; i < m.ntrack; i++) m.geom_trackslot[m.track_geom[i]] = -1;
  m.ntrack = m.ntrack_model;
  if (t.task_id == TASK_QUADRUPED) for (int f = 0; f < 4; f++) track_geom(m, t.iaux[3 + f]);
}

template <class T>
inline int set_plan(void* c, const double* state, int ns, const double* mocap, int nmocap, double time, const double* times,
                    int P, int interp, int H) {
  DevPlan<T>& p = reinterpret_cast<Constants<T>*>(c)->p;
  if (P > 36) return -1;
  if (state) for (int i = 0; i < ns; i++) p.state[i] = state[i];
  if (mocap) for (int i = 0; i < 7 * nmocap; i++) p.mocap[i] = mocap[i];
  if (time == time) p.time = time;
  if (times) { for (int i = 0; i < P; i++) p.spline_times[i] = times[i]; p.num_spline_points = P; p.interpolation = interp; }
  if (H > 0) p.horizon = H;
  return 0;
}

// DevPlan::warmstart: w (nv entries) or zeros
template <class T>
inline void set_warmstart(void* c, const double* w, int nv) {
  DevPlan<T>& p = reinterpret_cast<Constants<T>*>(c)->p;
  for (int i = 0; i < T::NV; i++) p.warmstart[i] = (w && i < nv) ? w[i] : 0.0;
}


// PsTask (C ABI) -> DevTask; also reports whether the task reads subtree-velocity sensors
inline int convert_task(const PsTask* t, const int* sensor_type, int nsensor, DevTask* out, int* need_subtreevel, std::string* why) {
  if (t->num_term > kMaxTerms) { *why = "more cost terms than the device carries"; return -1; }
  if (t->num_residual > kMaxResidual) { *why = "residual too long"; return -1; }
  if (t->num_parameters > kMaxTaskParam) { *why = "too many task parameters"; return -1; }
  DevTask d; std::memset(&d, 0, sizeof(d));
  d.task_id = t->task_id; d.num_residual = t->num_residual; d.num_term = t->num_term; d.num_trace = t->num_trace > 4 ? 4 : t->num_trace;
  int np = 0;
  for (int k = 0; k < t->num_term; k++) {
    d.dim_norm_residual[k] = t->dim_norm_residual[k]; d.num_norm_parameter[k] = t->num_norm_parameter[k]; d.norm[k] = t->norm[k];
    d.weight[k] = t->weight[k]; np += t->num_norm_parameter[k];
  }
  for (int k = 0; k < np && k < 2 * kMaxTerms; k++) d.norm_parameter[k] = t->norm_parameter[k];
  d.risk = t->risk;
  for (int k = 0; k < t->num_parameters; k++) d.parameters[k] = t->parameters[k];
  for (int k = 0; k < kMaxTaskAux; k++) d.aux[k] = t->aux[k];
  int need_sub = 0;
  for (int k = 0; k < 16; k++) {
    d.sensor_id[k] = t->sensor_id[k]; d.iaux[k] = t->iaux[k];
    const int s = t->sensor_id[k];
    if (s >= nsensor) { *why = "sensor id out of range"; return -1; }
    if (s < 0) continue;
    if (sensor_type[s] == PS_SENS_SUBTREELINVEL || sensor_type[s] == PS_SENS_SUBTREEANGMOM) need_sub = 1;
  }
  for (int k = 0; k < 4; k++) {
    d.trace_sensor[k] = t->trace_sensor[k];
    if (k < d.num_trace && (t->trace_sensor[k] < 0 || t->trace_sensor[k] >= nsensor)) { *why = "trace sensor id out of range"; return -1; }
  }
  *out = d; *need_subtreevel = need_sub;
  return 0;
}

}  // namespace ps
