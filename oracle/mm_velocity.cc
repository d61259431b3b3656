// ORACLE (test infrastructure).  Velocity / force stages: mj_comVel, mj_passive (spring, damper,
// inertia-box fluid), mj_rne, mj_fwdActuation, mj_fwdAcceleration.  Restates MuJoCo 3.1.4
// engine_core_smooth / engine_passive / engine_forward (third-party, not in /root/reference;
// call sites /root/reference/mjpc/trajectory.cc:158,198).  Parity vs MuJoCo: UNPINNED.
#include "mm_engine.h"
#include "mm_math.h"

namespace mm {

// ------------------------------------------------------------------ mj_comVel
static void comVel(const Model* m, Data* d) {
  zero(d->cvel.data(), 6);
  for (int i = 1; i < m->nbody; i++) {
    double cvel[6]; copy(cvel, &d->cvel[6 * m->body_parentid[i]], 6);
    for (int j = m->body_jntadr[i]; j >= 0 && j < m->body_jntadr[i] + m->body_jntnum[i]; j++) {
      int da = m->jnt_dofadr[j];
      int type = m->jnt_type[j];
      if (type == PS_JNT_FREE) {
        for (int k = 0; k < 3; k++) {
          zero(&d->cdof_dot[6 * (da + k)], 6);
          for (int r = 0; r < 6; r++) cvel[r] += d->cdof[6 * (da + k) + r] * d->qvel[da + k];
        }
        da += 3;
      }
      if (type == PS_JNT_FREE || type == PS_JNT_BALL) {
        for (int k = 0; k < 3; k++) crossMotion(&d->cdof_dot[6 * (da + k)], cvel, &d->cdof[6 * (da + k)]);
        for (int k = 0; k < 3; k++)
          for (int r = 0; r < 6; r++) cvel[r] += d->cdof[6 * (da + k) + r] * d->qvel[da + k];
      } else {
        crossMotion(&d->cdof_dot[6 * da], cvel, &d->cdof[6 * da]);
        for (int r = 0; r < 6; r++) cvel[r] += d->cdof[6 * da + r] * d->qvel[da];
      }
    }
    copy(&d->cvel[6 * i], cvel, 6);
  }
}

// ------------------------------------------------------------------ mj_rne (flg_acc = 0)
// result = C(q, v) v + g(q).  If qvel is given, cvel/cdof_dot are recomputed for that velocity
// into scratch (used by the implicit integrator's bias derivative).
void rne(const Model* m, Data* d, const double* qvel, double* result) {
  int nb = m->nbody, nv = m->nv;
  std::vector<double> save_qvel, save_cvel, save_cdd;
  if (qvel) {
    save_qvel = d->qvel; save_cvel = d->cvel; save_cdd = d->cdof_dot;
    copy(d->qvel.data(), qvel, nv); comVel(m, d);
  }
  std::vector<double> cacc(6 * nb, 0.0), cfrc(6 * nb, 0.0);
  bool grav = !(m->opt_disableflags & PS_DSBL_GRAVITY);
  for (int k = 0; k < 3; k++) cacc[3 + k] = grav ? -m->opt_gravity[k] : double(0.0);
  for (int i = 1; i < nb; i++) {
    int bda = m->body_dofadr[i];
    double* ca = &cacc[6 * i]; copy(ca, &cacc[6 * m->body_parentid[i]], 6);
    for (int j = 0; j < m->body_dofnum[i]; j++)
      for (int r = 0; r < 6; r++) ca[r] += d->cdof_dot[6 * (bda + j) + r] * d->qvel[bda + j];
    double t1[6], t2[6], t3[6];
    mulInertVec(t1, &d->cinert[10 * i], ca);
    mulInertVec(t2, &d->cinert[10 * i], &d->cvel[6 * i]);
    crossForce(t3, &d->cvel[6 * i], t2);
    for (int r = 0; r < 6; r++) cfrc[6 * i + r] = t1[r] + t3[r];
  }
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    if (p > 0) for (int r = 0; r < 6; r++) cfrc[6 * p + r] += cfrc[6 * i + r];
  }
  for (int i = 0; i < nv; i++) result[i] = dot(&d->cdof[6 * i], &cfrc[6 * m->dof_bodyid[i]], 6);
  if (qvel) { d->qvel = save_qvel; d->cvel = save_cvel; d->cdof_dot = save_cdd; }
}

// velocity of body inertial frame, local orientation (mj_objectVelocity(mjOBJ_BODY, flg_local=1))
static void bodyLocalVelocity(const Model* m, const Data* d, int i, double* lvel) {
  const double* cv = &d->cvel[6 * i];
  double dif[3], cr[3], lin[3];
  for (int k = 0; k < 3; k++) dif[k] = d->xipos[3 * i + k] - d->subtree_com[3 * m->body_rootid[i] + k];
  cross(cr, dif, cv);
  for (int k = 0; k < 3; k++) lin[k] = cv[3 + k] - cr[k];
  mulMatTVec3(lvel, &d->ximat[9 * i], cv);
  mulMatTVec3(lvel + 3, &d->ximat[9 * i], lin);
}

// qfrc += J(point, body)' * [force; torque]   (mj_applyFT)
static void applyFT(const Model* m, const Data* d, const double* force, const double* torque,
                    const double* point, int body, double* qfrc) {
  int nv = m->nv;
  std::vector<double> jp(3 * nv), jr(3 * nv);
  jac(m, d, jp.data(), jr.data(), point, body);
  for (int i = 0; i < nv; i++)
    for (int k = 0; k < 3; k++) qfrc[i] += jp[k * nv + i] * force[k] + jr[k * nv + i] * torque[k];
}

// equivalent inertia box half... (full) dimensions used by the inertia-box fluid model
void inertiaBox(const Model* m, int i, double box[3]) {
  const double* in = m->body_inertia + 3 * i; double mass = m->body_mass[i];
  box[0] = std::sqrt(std::fmax(kMinVal, in[1] + in[2] - in[0]) / mass * 6.0);
  box[1] = std::sqrt(std::fmax(kMinVal, in[0] + in[2] - in[1]) / mass * 6.0);
  box[2] = std::sqrt(std::fmax(kMinVal, in[0] + in[1] - in[2]) / mass * 6.0);
}

// local fluid force/torque for a local 6D velocity: lfrc = [torque; force]
void inertiaBoxForce(const Model* m, const double box[3], const double lvel[6], double lfrc[6]) {
  zero(lfrc, 6);
  double visc = m->opt_viscosity, dens = m->opt_density;
  if (visc > 0) {
    double diam = (box[0] + box[1] + box[2]) / 3.0;
    for (int k = 0; k < 3; k++) {
      lfrc[k] = -kPi * diam * diam * diam * visc * lvel[k];
      lfrc[3 + k] = -3.0 * kPi * diam * visc * lvel[3 + k];
    }
  }
  if (dens > 0) {
    lfrc[3] -= 0.5 * dens * box[1] * box[2] * std::fabs(lvel[3]) * lvel[3];
    lfrc[4] -= 0.5 * dens * box[0] * box[2] * std::fabs(lvel[4]) * lvel[4];
    lfrc[5] -= 0.5 * dens * box[0] * box[1] * std::fabs(lvel[5]) * lvel[5];
    double b0 = box[0], b1 = box[1], b2 = box[2];
    lfrc[0] -= dens * b0 * (b1 * b1 * b1 * b1 + b2 * b2 * b2 * b2) * std::fabs(lvel[0]) * lvel[0] / 64.0;
    lfrc[1] -= dens * b1 * (b0 * b0 * b0 * b0 + b2 * b2 * b2 * b2) * std::fabs(lvel[1]) * lvel[1] / 64.0;
    lfrc[2] -= dens * b2 * (b0 * b0 * b0 * b0 + b1 * b1 * b1 * b1) * std::fabs(lvel[2]) * lvel[2] / 64.0;
  }
}

// ------------------------------------------------------------------ mj_passive
static void passive(const Model* m, Data* d) {
  int nv = m->nv;
  zero(d->qfrc_spring.data(), nv); zero(d->qfrc_damper.data(), nv); zero(d->qfrc_fluid.data(), nv);
  zero(d->qfrc_passive.data(), nv);
  if (m->opt_disableflags & PS_DSBL_PASSIVE) return;
  for (int j = 0; j < m->njnt; j++) {
    double k = m->jnt_stiffness[j];
    if (k == 0) continue;
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    if (m->jnt_type[j] == PS_JNT_HINGE || m->jnt_type[j] == PS_JNT_SLIDE)
      d->qfrc_spring[da] = -k * (d->qpos[qa] - m->qpos_spring[qa]);
    // free / ball springs are not used by the supported model class
  }
  for (int i = 0; i < nv; i++) d->qfrc_damper[i] = -m->dof_damping[i] * d->qvel[i];
  if (m->opt_viscosity > 0 || m->opt_density > 0) {
    for (int i = 1; i < m->nbody; i++) {
      if (m->body_mass[i] < kMinVal) continue;
      double box[3], lvel[6], lfrc[6], bfrc[6];
      inertiaBox(m, i, box);
      bodyLocalVelocity(m, d, i, lvel);
      // wind: subtract local wind from translational velocity
      double lw[3]; mulMatTVec3(lw, &d->ximat[9 * i], m->opt_wind);
      for (int k = 0; k < 3; k++) lvel[3 + k] -= lw[k];
      inertiaBoxForce(m, box, lvel, lfrc);
      mulMatVec3(bfrc, &d->ximat[9 * i], lfrc);
      mulMatVec3(bfrc + 3, &d->ximat[9 * i], lfrc + 3);
      applyFT(m, d, bfrc + 3, bfrc, &d->xipos[3 * i], i, d->qfrc_fluid.data());
    }
  }
  for (int i = 0; i < nv; i++) d->qfrc_passive[i] = d->qfrc_spring[i] + d->qfrc_damper[i] + d->qfrc_fluid[i];
}

void fwdVelocity(const Model* m, Data* d) {
  for (int i = 0; i < m->nu; i++) {
    int j = m->actuator_trnid[2 * i];
    d->actuator_velocity[i] = m->actuator_gear[i] * d->qvel[m->jnt_dofadr[j]];
  }
  comVel(m, d);
  passive(m, d);
  rne(m, d, nullptr, d->qfrc_bias.data());
}

// ------------------------------------------------------------------ mj_fwdActuation
void fwdActuation(const Model* m, Data* d) {
  int nv = m->nv, nu = m->nu;
  zero(d->qfrc_actuator.data(), nv); zero(d->actuator_force.data(), nu);
  if (m->na) zero(d->act_dot.data(), m->na);
  if (!nu || (m->opt_disableflags & PS_DSBL_ACTUATION)) return;
  for (int i = 0; i < nu; i++) {
    double ctrl = d->ctrl[i];
    if (std::isnan(ctrl) || std::fabs(ctrl) > kMaxVal) { d->warning[WARN_BADCTRL]++; ctrl = 0; for (int k = 0; k < nu; k++) d->ctrl[k] = 0; }
    if (m->actuator_ctrllimited[i] && !(m->opt_disableflags & PS_DSBL_CLAMPCTRL))
      ctrl = clip(ctrl, m->actuator_ctrlrange[2 * i], m->actuator_ctrlrange[2 * i + 1]);
    int aa = m->actuator_actadr[i];
    double in = ctrl;
    if (m->actuator_dyntype[i] != PS_DYN_NONE) {
      const double* prm = m->actuator_dynprm + 3 * i;
      switch (m->actuator_dyntype[i]) {
        case PS_DYN_INTEGRATOR: d->act_dot[aa] = ctrl; break;
        case PS_DYN_FILTER:
        case PS_DYN_FILTEREXACT: d->act_dot[aa] = (ctrl - d->act[aa]) / std::fmax(kMinVal, prm[0]); break;
      }
      in = d->act[aa];
    }
    const double* gp = m->actuator_gainprm + 3 * i; const double* bp = m->actuator_biasprm + 3 * i;
    double gain = gp[0];
    if (m->actuator_gaintype[i] == PS_GAIN_AFFINE) gain = gp[0] + gp[1] * d->actuator_length[i] + gp[2] * d->actuator_velocity[i];
    double bias = 0;
    if (m->actuator_biastype[i] == PS_BIAS_AFFINE) bias = bp[0] + bp[1] * d->actuator_length[i] + bp[2] * d->actuator_velocity[i];
    double f = gain * in + bias;
    if (m->actuator_forcelimited[i]) f = clip(f, m->actuator_forcerange[2 * i], m->actuator_forcerange[2 * i + 1]);
    d->actuator_force[i] = f;
    int j = m->actuator_trnid[2 * i];
    d->qfrc_actuator[m->jnt_dofadr[j]] += m->actuator_gear[i] * f;
  }
}

// ------------------------------------------------------------------ mj_fwdAcceleration
void fwdAcceleration(const Model* m, Data* d) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++) d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_actuator[i];
  // mj_xfrcAccumulate: Cartesian forces / torques applied at the body inertial frames (RobustPlanner's
  // perturbation, mjpc/trajectory.cc:147-155); bodies with an all-zero entry are skipped
  for (int b = 1; b < m->nbody; b++) {
    const double* x = &d->xfrc_applied[6 * b];
    if (x[0] == 0 && x[1] == 0 && x[2] == 0 && x[3] == 0 && x[4] == 0 && x[5] == 0) continue;
    applyFT(m, d, x, x + 3, &d->xipos[3 * b], b, d->qfrc_smooth.data());
  }
  copy(d->qacc_smooth.data(), d->qfrc_smooth.data(), nv);
  solveM(m, d, d->qacc_smooth.data());
}

}  // namespace mm
