// ORACLE (test infrastructure).  Top level of the restated MuJoCo 3.1.4 step: mj_forward stage
// order, sensors, state checks and the Euler / implicit / implicitfast integrators (engine_forward,
// engine_sensor; third-party, not in /root/reference; call sites mjpc/trajectory.cc:158,198).
// Parity vs MuJoCo: UNPINNED.
#include <cstring>
#include "mm_engine.h"
#include "mm_math.h"

namespace mm {

void fwdPositionSmooth(const Model* m, Data* d);
void makeConstraint(const Model* m, Data* d);
void referenceConstraint(const Model* m, Data* d);
void factorI(const Model* m, double* L, double* dinv);
void solveI(const Model* m, double* x, const double* L, const double* dinv);
void smoothVelDerivative(const Model* m, Data* d, bool flg_rne);

void fwdPosition(const Model* m, Data* d) { fwdPositionSmooth(m, d); makeConstraint(m, d); }

// 6D velocity [ang; lin] of a frame at `pos` rigidly attached to `body`, world orientation
static void objectVelocity(const Model* m, const Data* d, int body, const double* pos, double* res) {
  const double* cv = &d->cvel[6 * body];
  double dif[3], cr[3];
  for (int k = 0; k < 3; k++) dif[k] = pos[k] - d->subtree_com[3 * m->body_rootid[body] + k];
  cross(cr, dif, cv);
  for (int k = 0; k < 3; k++) { res[k] = cv[k]; res[3 + k] = cv[3 + k] - cr[k]; }
}

static void objectFrame(const Model* m, const Data* d, int objtype, int id, const double** pos, const double** mat, int* body) {
  switch (objtype) {
    case PS_OBJ_BODY: *pos = &d->xipos[3 * id]; *mat = &d->ximat[9 * id]; *body = id; break;
    case PS_OBJ_XBODY: *pos = &d->xpos[3 * id]; *mat = &d->xmat[9 * id]; *body = id; break;
    case PS_OBJ_GEOM: *pos = &d->geom_xpos[3 * id]; *mat = &d->geom_xmat[9 * id]; *body = m->geom_bodyid[id]; break;
    default: *pos = &d->site_xpos[3 * id]; *mat = &d->site_xmat[9 * id]; *body = m->site_bodyid[id]; break;
  }
}

void sensorPos(const Model* m, Data* d) {
  for (int s = 0; s < m->nsensor; s++) {
    double* out = &d->sensordata[m->sensor_adr[s]];
    int id = m->sensor_objid[s], ot = m->sensor_objtype[s];
    const double *pos, *mat; int body;
    switch (m->sensor_type[s]) {
      case PS_SENS_JOINTPOS: out[0] = d->qpos[m->jnt_qposadr[id]]; break;
      case PS_SENS_FRAMEPOS: objectFrame(m, d, ot, id, &pos, &mat, &body); copy(out, pos, 3); break;
      case PS_SENS_FRAMEXAXIS: case PS_SENS_FRAMEYAXIS: case PS_SENS_FRAMEZAXIS: {
        objectFrame(m, d, ot, id, &pos, &mat, &body);
        int c = m->sensor_type[s] - PS_SENS_FRAMEXAXIS;
        out[0] = mat[c]; out[1] = mat[3 + c]; out[2] = mat[6 + c];
      } break;
      case PS_SENS_SUBTREECOM: copy(out, &d->subtree_com[3 * id], 3); break;
      default: break;
    }
  }
}

static void subtreeVel(const Model* m, Data* d) {
  int nb = m->nbody;
  std::vector<double> bodyvel(6 * nb, 0.0);
  double* lv = d->subtree_linvel.data(); double* am = d->subtree_angmom.data();
  for (int i = 0; i < nb; i++) {
    objectVelocity(m, d, i, &d->xipos[3 * i], &bodyvel[6 * i]);
    for (int k = 0; k < 3; k++) lv[3 * i + k] = m->body_mass[i] * bodyvel[6 * i + 3 + k];
    double t[3], t2[3]; mulMatTVec3(t, &d->ximat[9 * i], &bodyvel[6 * i]);
    for (int k = 0; k < 3; k++) t[k] *= m->body_inertia[3 * i + k];
    mulMatVec3(t2, &d->ximat[9 * i], t); copy(am + 3 * i, t2, 3);
  }
  for (int i = nb - 1; i >= 0; i--) {
    if (i) for (int k = 0; k < 3; k++) lv[3 * m->body_parentid[i] + k] += lv[3 * i + k];
    for (int k = 0; k < 3; k++) lv[3 * i + k] /= std::fmax(kMinVal, m->body_subtreemass[i]);
  }
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i]; double dx[3], dv[3], dL[3];
    for (int k = 0; k < 3; k++) { dx[k] = d->xipos[3 * i + k] - d->subtree_com[3 * i + k]; dv[k] = (bodyvel[6 * i + 3 + k] - lv[3 * i + k]) * m->body_mass[i]; }
    cross(dL, dx, dv);
    for (int k = 0; k < 3; k++) { am[3 * i + k] += dL[k]; am[3 * p + k] += am[3 * i + k]; }
    for (int k = 0; k < 3; k++) { dx[k] = d->subtree_com[3 * i + k] - d->subtree_com[3 * p + k]; dv[k] = (lv[3 * i + k] - lv[3 * p + k]) * m->body_subtreemass[i]; }
    cross(dL, dx, dv);
    for (int k = 0; k < 3; k++) am[3 * p + k] += dL[k];
  }
}

void sensorVel(const Model* m, Data* d) {
  bool sub = false;
  for (int s = 0; s < m->nsensor; s++) {
    int t = m->sensor_type[s];
    if (t == PS_SENS_SUBTREELINVEL || t == PS_SENS_SUBTREEANGMOM) sub = true;
  }
  if (sub) subtreeVel(m, d);
  for (int s = 0; s < m->nsensor; s++) {
    double* out = &d->sensordata[m->sensor_adr[s]];
    int id = m->sensor_objid[s], ot = m->sensor_objtype[s];
    const double *pos, *mat; int body; double v[6];
    switch (m->sensor_type[s]) {
      case PS_SENS_JOINTVEL: out[0] = d->qvel[m->jnt_dofadr[id]]; break;
      case PS_SENS_FRAMELINVEL: objectFrame(m, d, ot, id, &pos, &mat, &body); objectVelocity(m, d, body, pos, v); copy(out, v + 3, 3); break;
      case PS_SENS_FRAMEANGVEL: objectFrame(m, d, ot, id, &pos, &mat, &body); objectVelocity(m, d, body, pos, v); copy(out, v, 3); break;
      case PS_SENS_SUBTREELINVEL: copy(out, &d->subtree_linvel[3 * id], 3); break;
      case PS_SENS_SUBTREEANGMOM: copy(out, &d->subtree_angmom[3 * id], 3); break;
      default: break;
    }
  }
}

// acceleration-stage sensors (touch, accelerometer) are only consumed by the reference's state
// estimators, never by a residual: they are left at zero (documented in DESIGN.md).
void sensorAcc(const Model* m, Data* d) { (void)m; (void)d; }

static bool bad(double x) { return std::isnan(x) || x > kMaxVal || x < -kMaxVal; }

void forward(const Model* m, Data* d, SensorCallback cb, void* user) {
  fwdPosition(m, d);
  sensorPos(m, d);
  fwdVelocity(m, d);
  referenceConstraint(m, d);
  sensorVel(m, d);
  fwdActuation(m, d);
  fwdAcceleration(m, d);
  fwdConstraint(m, d);
  sensorAcc(m, d);
  if (cb) cb(m, d, user);
}

static void integratePos(const Model* m, double* qpos, const double* qvel, double dt) {
  for (int j = 0; j < m->njnt; j++) {
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    switch (m->jnt_type[j]) {
      case PS_JNT_FREE:
        for (int k = 0; k < 3; k++) qpos[qa + k] += dt * qvel[da + k];
        quatIntegrate(qpos + qa + 3, qvel + da + 3, dt);
        break;
      case PS_JNT_BALL: quatIntegrate(qpos + qa, qvel + da, dt); break;
      default: qpos[qa] += dt * qvel[da];
    }
  }
}

static void advance(const Model* m, Data* d, const double* qacc) {
  double dt = m->opt_timestep;
  if (m->na && !(m->opt_disableflags & PS_DSBL_ACTUATION)) {
    for (int i = 0; i < m->nu; i++) {
      int a = m->actuator_actadr[i];
      if (a < 0) continue;
      double act = d->act[a];
      if (m->actuator_dyntype[i] == PS_DYN_FILTEREXACT) {
        double tau = std::fmax(kMinVal, m->actuator_dynprm[3 * i]);
        act = act + d->act_dot[a] * tau * (1 - std::exp(-dt / tau));
      } else act = act + d->act_dot[a] * dt;
      if (m->actuator_actlimited[i]) act = clip(act, m->actuator_actrange[2 * i], m->actuator_actrange[2 * i + 1]);
      d->act[a] = act;
    }
  }
  for (int i = 0; i < m->nv; i++) d->qvel[i] += qacc[i] * dt;
  integratePos(m, d->qpos.data(), d->qvel.data(), dt);
  d->time += dt;
}

void step(const Model* m, Data* d, SensorCallback cb, void* user) {
  int nv = m->nv;
  for (int i = 0; i < m->nq; i++) if (bad(d->qpos[i])) { d->warning[WARN_BADQPOS]++; return; }
  for (int i = 0; i < nv; i++) if (bad(d->qvel[i])) { d->warning[WARN_BADQVEL]++; return; }
  forward(m, d, cb, user);
  for (int i = 0; i < nv; i++) if (bad(d->qacc[i])) { d->warning[WARN_BADQACC]++; return; }
  double dt = m->opt_timestep;
  std::vector<double> qacc(nv), L(nv * nv), dinv(nv);
  if (m->opt_integrator == PS_INT_EULER || m->opt_integrator == PS_INT_RK4) {
    bool damp = false;
    if (!(m->opt_disableflags & PS_DSBL_EULERDAMP)) for (int i = 0; i < nv; i++) if (m->dof_damping[i] > 0) damp = true;
    if (!damp) copy(qacc.data(), d->qacc.data(), nv);
    else {
      copy(L.data(), d->qM.data(), nv * nv);
      for (int i = 0; i < nv; i++) L[i * nv + i] += dt * m->dof_damping[i];
      factorI(m, L.data(), dinv.data());
      for (int i = 0; i < nv; i++) qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i];
      solveI(m, qacc.data(), L.data(), dinv.data());
    }
  } else {
    bool full = m->opt_integrator == PS_INT_IMPLICIT;
    smoothVelDerivative(m, d, full);
    for (int i = 0; i < nv; i++) qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i];
    if (full) {
      // dense LU (no pivoting, as MuJoCo's mju_factorLUSparse) of M - dt*qDeriv
      for (int i = 0; i < nv * nv; i++) L[i] = d->qM[i] - dt * d->qDeriv[i];
      for (int k = 0; k < nv; k++)
        for (int i = k + 1; i < nv; i++) {
          double f = L[i * nv + k] / L[k * nv + k]; L[i * nv + k] = f;
          for (int j = k + 1; j < nv; j++) L[i * nv + j] -= f * L[k * nv + j];
        }
      for (int i = 0; i < nv; i++) for (int k = 0; k < i; k++) qacc[i] -= L[i * nv + k] * qacc[k];
      for (int i = nv - 1; i >= 0; i--) { for (int k = i + 1; k < nv; k++) qacc[i] -= L[i * nv + k] * qacc[k]; qacc[i] /= L[i * nv + i]; }
    } else {
      // implicitfast: symmetrized derivative restricted to M's sparsity, L'DL
      for (int i = 0; i < nv; i++) for (int j = 0; j < nv; j++)
        L[i * nv + j] = d->qM[i * nv + j] - dt * 0.5 * (d->qDeriv[i * nv + j] + d->qDeriv[j * nv + i]);
      factorI(m, L.data(), dinv.data());
      solveI(m, qacc.data(), L.data(), dinv.data());
    }
  }
  advance(m, d, qacc.data());
}

int name2id(const Model* m, int objtype, const char* name) {
  const int* adr = nullptr; int n = 0;
  switch (objtype) {
    case PS_OBJ_BODY: case PS_OBJ_XBODY: adr = m->name_bodyadr; n = m->nbody; break;
    case PS_OBJ_JOINT: adr = m->name_jntadr; n = m->njnt; break;
    case PS_OBJ_GEOM: adr = m->name_geomadr; n = m->ngeom; break;
    case PS_OBJ_SITE: adr = m->name_siteadr; n = m->nsite; break;
    case PS_OBJ_ACTUATOR: adr = m->name_actuatoradr; n = m->nu; break;
    case PS_OBJ_SENSOR: adr = m->name_sensoradr; n = m->nsensor; break;
    case PS_OBJ_NUMERIC: adr = m->name_numericadr; n = m->nnumeric; break;
    case PS_OBJ_TEXT: adr = m->name_textadr; n = m->ntext; break;
    case PS_OBJ_KEY: adr = m->name_keyadr; n = m->nkey; break;
    default: return -1;
  }
  for (int i = 0; i < n; i++) if (!std::strcmp(m->names + adr[i], name)) return i;
  return -1;
}

}  // namespace mm
