// ORACLE (test infrastructure): the engine side of oracle/ref_shim/mujoco/mujoco.h.  mjModel is a view filled from a
// PsModel, mjData's arrays alias an mm::Data, and mj_step / mj_forward / mj_ray / mj_name2id run the oracle's restated
// MuJoCo 3.1.4 step (mm_*.cc) -- with the user-sensor callback (mjcb_sensor, which MJPC points at Task::Residual,
// mjpc/testspeed.cc:36-40) invoked at the end of the acceleration stage, as engine_sensor.c does.  Linked with the
// reference's own sources into oracle/_ref/libmjpc_ref.so; the physics stays the restatement (UNPINNED vs real MuJoCo).
#include <cstring>
#include <new>
#include <type_traits>
#include <vector>
#include <atomic>
#include <absl/random/random.h>
#include <mujoco/mujoco.h>
#include "../include/ps_model.h"
#include "mm_engine.h"

mjfSensor mjcb_sensor = nullptr;
mjfGeneric mjcb_control = nullptr;

namespace {
struct DataBox { mjData view; mm::Data d; const mjModel* m; };
struct ModelBox { mjModel view; std::vector<double> key_mpos, key_mquat; std::vector<std::vector<mjtByte>> bytes; };

// Test-harness switch: the reference's worker mjData keeps qacc_warmstart from whatever rollout it ran last
// (planners/sampling/planner.cc:374), the engine under test starts every rollout from a zero warm start (DESIGN.md,
// deviation 2).  When set, a step whose time is EARLIER than the previous step's on the same mjData -- i.e. the first
// step of a new rollout -- clears qacc_warmstart first, which is what a fresh mjData per rollout would do.
bool g_zero_warmstart_on_rewind = false;
thread_local double t_last_time = -1e300;

void bind(DataBox* b) {
  mjData& v = b->view; mm::Data& d = b->d;
  v.qpos = d.qpos.data(); v.qvel = d.qvel.data(); v.act = d.act.data(); v.ctrl = d.ctrl.data(); v.qacc = d.qacc.data();
  v.qacc_warmstart = d.qacc_warmstart.data(); v.mocap_pos = d.mocap_pos.data(); v.mocap_quat = d.mocap_quat.data();
  v.userdata = d.userdata.data(); v.sensordata = d.sensordata.data(); v.xfrc_applied = d.xfrc_applied.data();
  v.xpos = d.xpos.data(); v.xquat = d.xquat.data(); v.xmat = d.xmat.data(); v.xipos = d.xipos.data(); v.ximat = d.ximat.data();
  v.geom_xpos = d.geom_xpos.data(); v.geom_xmat = d.geom_xmat.data(); v.site_xpos = d.site_xpos.data(); v.site_xmat = d.site_xmat.data();
  v.subtree_com = d.subtree_com.data(); v.actuator_force = d.actuator_force.data();
  v.shim = b;
}
DataBox* box(const mjData* d) { return static_cast<DataBox*>(d->shim); }

void sensor_cb(const mm::Model*, mm::Data*, void* user) {
  DataBox* b = static_cast<DataBox*>(user);
  if (mjcb_sensor) mjcb_sensor(b->m, &b->view, mjSTAGE_ACC);
}
void sync_in(DataBox* b) { b->d.time = b->view.time; }
void sync_out(DataBox* b, const int* w0) {
  mjData& v = b->view; const mm::Data& d = b->d;
  v.time = d.time; v.nefc = d.nefc; v.ncon = d.ncon;
  static const int map[mm::N_WARN] = {mjWARN_BADQPOS, mjWARN_BADQVEL, mjWARN_BADQACC, mjWARN_BADCTRL};
  for (int i = 0; i < mm::N_WARN; i++) v.warning[map[i]].number += d.warning[i] - w0[i];
}
}  // namespace

extern "C" void ref_set_zero_warmstart_on_rewind(int on) { g_zero_warmstart_on_rewind = on != 0; }

// mjModel view over a PsModel (the caller keeps the PsModel and its arrays alive)
extern "C" mjModel* ref_make_model(const PsModel* p) {
  ModelBox* b = new ModelBox();
  mjModel& m = b->view; std::memset(&m, 0, sizeof(m));
  auto bytes = [b](const int* v, size_t n) { b->bytes.emplace_back(n ? n : 1); for (size_t i = 0; i < n; i++) b->bytes.back()[i] = (mjtByte)v[i]; return b->bytes.back().data(); };
#include "ref_model_fill.inc"
  // keyframe mocap poses are not part of PsModel: the model's own body poses (what the compiler stores by default)
  b->key_mpos.assign((size_t)3 * p->nmocap * (p->nkey > 0 ? p->nkey : 1), 0.0); b->key_mquat.assign((size_t)4 * p->nmocap * (p->nkey > 0 ? p->nkey : 1), 0.0);
  for (int k = 0; k < (p->nkey > 0 ? p->nkey : 1); k++)
    for (int i = 0; i < p->nbody; i++) {
      const int id = p->body_mocapid[i]; if (id < 0) continue;
      for (int c = 0; c < 3; c++) b->key_mpos[(size_t)3 * (k * p->nmocap + id) + c] = p->body_pos[3 * i + c];
      for (int c = 0; c < 4; c++) b->key_mquat[(size_t)4 * (k * p->nmocap + id) + c] = p->body_quat[4 * i + c];
    }
  m.key_mpos = b->key_mpos.data(); m.key_mquat = b->key_mquat.data();
  m.ps = p;
  return &b->view;
}
void mj_deleteModel(mjModel* m) { if (m) delete reinterpret_cast<ModelBox*>(m); }

mjData* mj_makeData(const mjModel* m) {
  DataBox* b = new DataBox();
  std::memset(&b->view, 0, sizeof(b->view));
  b->m = m;
  mm::makeData(m->ps, &b->d);
  bind(b);
  b->view.time = 0;
  return &b->view;
}
void mj_deleteData(mjData* d) { if (d) delete box(d); }
void mj_resetData(const mjModel* m, mjData* d) {
  DataBox* b = box(d); mm::resetData(m->ps, &b->d);
  d->time = 0; d->nefc = 0; d->ncon = 0; for (auto& w : d->warning) { w.lastinfo = 0; w.number = 0; }
  std::fill(b->d.sensordata.begin(), b->d.sensordata.end(), 0.0);
}
void mj_resetDataKeyframe(const mjModel* m, mjData* d, int key) {
  mj_resetData(m, d);
  if (key < 0 || key >= m->nkey) return;
  mju_copy(d->qpos, m->key_qpos + (size_t)key * m->nq, m->nq); mju_copy(d->qvel, m->key_qvel + (size_t)key * m->nv, m->nv);
  mju_copy(d->act, m->key_act + (size_t)key * m->na, m->na); mju_copy(d->ctrl, m->key_ctrl + (size_t)key * m->nu, m->nu);
}
mjData* mj_copyData(mjData* dest, const mjModel* m, const mjData* src) {
  if (!dest) dest = mj_makeData(m);
  DataBox* a = box(dest); const DataBox* s = box(src);
  a->d = s->d; bind(a);
  a->view.time = src->time; a->view.nefc = src->nefc; a->view.ncon = src->ncon;
  std::memcpy(a->view.warning, src->warning, sizeof(src->warning));
  return dest;
}

void mj_step(const mjModel* m, mjData* d) {
  DataBox* b = box(d);
  if (g_zero_warmstart_on_rewind && d->time < t_last_time) std::fill(b->d.qacc_warmstart.begin(), b->d.qacc_warmstart.end(), 0.0);
  t_last_time = d->time;
  int w0[mm::N_WARN]; for (int i = 0; i < mm::N_WARN; i++) w0[i] = b->d.warning[i];
  sync_in(b);
  mm::step(m->ps, &b->d, sensor_cb, b);
  sync_out(b, w0);
}
void mj_forward(const mjModel* m, mjData* d) {
  DataBox* b = box(d);
  int w0[mm::N_WARN]; for (int i = 0; i < mm::N_WARN; i++) w0[i] = b->d.warning[i];
  sync_in(b);
  mm::forward(m->ps, &b->d, sensor_cb, b);
  sync_out(b, w0);
}
void mj_kinematics(const mjModel* m, mjData* d) { mm::fwdPosition(m->ps, &box(d)->d); }

int mj_name2id(const mjModel* m, int type, const char* name) {
  // the PS_OBJ_* constants are MuJoCo's for every type the oracle resolves, except keyframes
  return mm::name2id(m->ps, type == mjOBJ_KEY ? PS_OBJ_KEY : type, name);
}
const char* mj_id2name(const mjModel* m, int type, int id) {
  const int* adr = nullptr; int n = 0;
  switch (type) {
    case mjOBJ_BODY: case mjOBJ_XBODY: adr = m->name_bodyadr; n = m->nbody; break;
    case mjOBJ_JOINT: adr = m->name_jntadr; n = m->njnt; break;
    case mjOBJ_GEOM: adr = m->name_geomadr; n = m->ngeom; break;
    case mjOBJ_SITE: adr = m->name_siteadr; n = m->nsite; break;
    case mjOBJ_ACTUATOR: adr = m->name_actuatoradr; n = m->nu; break;
    case mjOBJ_SENSOR: adr = m->name_sensoradr; n = m->nsensor; break;
    case mjOBJ_NUMERIC: adr = m->name_numericadr; n = m->nnumeric; break;
    case mjOBJ_TEXT: adr = m->name_textadr; n = m->ntext; break;
    case mjOBJ_KEY: adr = m->name_keyadr; n = m->nkey; break;
    default: return nullptr;
  }
  if (id < 0 || id >= n) return nullptr;
  const char* s = m->names + adr[id];
  return *s ? s : nullptr;
}
mjtNum mj_ray(const mjModel* m, const mjData* d, const mjtNum pnt[3], const mjtNum vec[3], const mjtByte* geomgroup, mjtByte flg_static,
              int bodyexclude, int geomid[1]) {
  return mm::ray(m->ps, &box(d)->d, pnt, vec, geomgroup, flg_static, bodyexclude, geomid);
}
void mj_jac(const mjModel* m, const mjData* d, mjtNum* jacp, mjtNum* jacr, const mjtNum point[3], int body) {
  mm::jac(m->ps, &box(d)->d, jacp, jacr, point, body);
}
void mj_differentiatePos(const mjModel* m, mjtNum* qvel, mjtNum dt, const mjtNum* qpos1, const mjtNum* qpos2) {
  for (int j = 0; j < m->njnt; j++) {
    int pa = m->jnt_qposadr[j], va = m->jnt_dofadr[j];
    switch (m->jnt_type[j]) {
      case mjJNT_FREE:
        for (int i = 0; i < 3; i++) qvel[va + i] = (qpos2[pa + i] - qpos1[pa + i]) / dt;
        va += 3; pa += 3;
        [[fallthrough]];
      case mjJNT_BALL: {
        mjtNum neg[4], dif[4]; mju_negQuat(neg, qpos1 + pa); mju_mulQuat(dif, neg, qpos2 + pa); mju_quat2Vel(qvel + va, dif, dt);
      } break;
      default: qvel[va] = (qpos2[pa] - qpos1[pa]) / dt;
    }
  }
}
void mj_integratePos(const mjModel* m, mjtNum* qpos, const mjtNum* qvel, mjtNum dt) {
  for (int j = 0; j < m->njnt; j++) {
    int pa = m->jnt_qposadr[j], va = m->jnt_dofadr[j];
    switch (m->jnt_type[j]) {
      case mjJNT_FREE:
        for (int i = 0; i < 3; i++) qpos[pa + i] += dt * qvel[va + i];
        pa += 3; va += 3;
        [[fallthrough]];
      case mjJNT_BALL: mju_quatIntegrate(qpos + pa, qvel + va, dt); break;
      default: qpos[pa] += dt * qvel[va];
    }
  }
}
// visualisation / UI: never reached on the planning path
mjModel* mj_loadXML(const char*, const void*, char* error, int error_sz) { if (error && error_sz > 0) std::snprintf(error, error_sz, "shim: no XML compiler"); return nullptr; }
mjModel* mj_copyModel(mjModel*, const mjModel*) { return nullptr; }
void mjv_initGeom(mjvGeom* g, int type, const mjtNum*, const mjtNum*, const mjtNum*, const float*) { if (g) { std::memset(g, 0, sizeof(*g)); g->type = type; } }
void mjv_makeConnector(mjvGeom*, int, mjtNum, mjtNum, mjtNum, mjtNum, mjtNum, mjtNum, mjtNum) {}
void mjui_add(mjUI*, const mjuiDef*) {}
void mjr_figure(mjrRect, mjvFigure*, const mjrContext*) {}
mjtNum mju_cholFactorBand(mjtNum*, int, int, int, mjtNum, mjtNum) { mju_error("shim: mju_cholFactorBand is outside the sampling path"); return 0; }
void mju_cholSolveBand(mjtNum*, const mjtNum*, const mjtNum*, int, int, int) { mju_error("shim: mju_cholSolveBand is outside the sampling path"); }
void mju_band2Dense(mjtNum*, const mjtNum*, int, int, int, mjtByte) { mju_error("shim: band matrices are outside the sampling path"); }
void mju_dense2Band(mjtNum*, const mjtNum*, int, int, int) { mju_error("shim: band matrices are outside the sampling path"); }
int mju_bandDiag(int, int, int, int) { return 0; }
int mju_eig3(mjtNum*, mjtNum*, mjtNum*, const mjtNum*) { mju_error("shim: mju_eig3 is outside the sampling path"); return 0; }
void mju_mat2Quat(mjtNum*, const mjtNum*) { mju_error("shim: mju_mat2Quat is outside the sampling path"); }

namespace ref_shim_random { Replay& replay() { static Replay r; return r; } }
namespace absl {
BitGen::BitGen() { static std::atomic<unsigned long long> ctr{0x9E3779B97F4A7C15ull}; eng_.seed(ctr.fetch_add(0xD1B54A32D192ED03ull)); }
}  // namespace absl
