// ORACLE (test infrastructure).  Smooth dynamics: kinematics, CoM frame, CRB, L'DL, RNE, passive
// forces, actuation.  Restates MuJoCo 3.1.4 engine_core_smooth / engine_passive / engine_forward
// (third-party, not in /root/reference; see mm_engine.h).  Parity vs MuJoCo: UNPINNED.
#include <cstring>
#include "mm_engine.h"
#include "mm_math.h"

namespace mm {

void makeData(const Model* m, Data* d) {
  int nq = m->nq, nv = m->nv, nb = m->nbody;
  d->qpos.assign(nq, 0); d->qvel.assign(nv, 0); d->act.assign(m->na, 0); d->ctrl.assign(m->nu, 0);
  d->qacc.assign(nv, 0); d->qacc_warmstart.assign(nv, 0); d->act_dot.assign(m->na, 0);
  d->mocap_pos.assign(3 * m->nmocap, 0); d->mocap_quat.assign(4 * m->nmocap, 0);
  d->userdata.assign(m->nuserdata, 0); d->sensordata.assign(m->nsensordata, 0); d->xfrc_applied.assign(6 * nb, 0);
  d->xpos.assign(3 * nb, 0); d->xquat.assign(4 * nb, 0); d->xmat.assign(9 * nb, 0);
  d->xipos.assign(3 * nb, 0); d->ximat.assign(9 * nb, 0);
  d->xanchor.assign(3 * m->njnt, 0); d->xaxis.assign(3 * m->njnt, 0);
  d->geom_xpos.assign(3 * m->ngeom, 0); d->geom_xmat.assign(9 * m->ngeom, 0);
  d->site_xpos.assign(3 * m->nsite, 0); d->site_xmat.assign(9 * m->nsite, 0);
  d->subtree_com.assign(3 * nb, 0); d->cinert.assign(10 * nb, 0); d->crb.assign(10 * nb, 0);
  d->cdof.assign(6 * nv, 0); d->qM.assign(nv * nv, 0); d->qLD.assign(nv * nv, 0);
  d->qLDiagInv.assign(nv, 0); d->actuator_length.assign(m->nu, 0);
  d->cvel.assign(6 * nb, 0); d->cdof_dot.assign(6 * nv, 0); d->qfrc_bias.assign(nv, 0);
  d->qfrc_spring.assign(nv, 0); d->qfrc_damper.assign(nv, 0); d->qfrc_fluid.assign(nv, 0);
  d->qfrc_passive.assign(nv, 0); d->actuator_velocity.assign(m->nu, 0);
  d->subtree_linvel.assign(3 * nb, 0); d->subtree_angmom.assign(3 * nb, 0);
  d->actuator_force.assign(m->nu, 0); d->qfrc_actuator.assign(nv, 0); d->qfrc_smooth.assign(nv, 0);
  d->qacc_smooth.assign(nv, 0); d->qfrc_constraint.assign(nv, 0);
  d->cacc.assign(6 * nb, 0); d->cfrc_body.assign(6 * nb, 0); d->qDeriv.assign(nv * nv, 0);
  resetData(m, d);
}

void resetData(const Model* m, Data* d) {
  copy(d->qpos.data(), m->qpos0, m->nq);
  zero(d->qvel.data(), m->nv); zero(d->act.data(), m->na); zero(d->ctrl.data(), m->nu);
  zero(d->qacc.data(), m->nv); zero(d->qacc_warmstart.data(), m->nv); zero(d->xfrc_applied.data(), 6 * m->nbody);
  for (int i = 0; i < m->nbody; i++) {
    int id = m->body_mocapid[i];
    if (id >= 0) { copy(&d->mocap_pos[3 * id], m->body_pos + 3 * i, 3); copy(&d->mocap_quat[4 * id], m->body_quat + 4 * i, 4); }
  }
  d->time = 0; d->nefc = 0; d->ncon = 0;
  for (int i = 0; i < N_WARN; i++) d->warning[i] = 0;
}

// ------------------------------------------------------------------ mj_kinematics
static void kinematics(const Model* m, Data* d) {
  double* xpos = d->xpos.data(); double* xquat = d->xquat.data(); double* xmat = d->xmat.data();
  zero(xpos, 3); xquat[0] = 1; xquat[1] = xquat[2] = xquat[3] = 0;
  zero(xmat, 9); xmat[0] = xmat[4] = xmat[8] = 1;
  zero(d->xipos.data(), 3); copy(d->ximat.data(), xmat, 9);
  // normalize all quaternions in qpos and mocap
  for (int j = 0; j < m->njnt; j++)
    if (m->jnt_type[j] == PS_JNT_FREE) normalize4(&d->qpos[m->jnt_qposadr[j] + 3]);
    else if (m->jnt_type[j] == PS_JNT_BALL) normalize4(&d->qpos[m->jnt_qposadr[j]]);
  for (int i = 0; i < m->nmocap; i++) normalize4(&d->mocap_quat[4 * i]);

  for (int i = 1; i < m->nbody; i++) {
    double pos[3], quat[4];
    int pid = m->body_parentid[i];
    int jn = m->body_jntnum[i], ja = m->body_jntadr[i];
    if (m->body_mocapid[i] >= 0) {
      copy(pos, &d->mocap_pos[3 * m->body_mocapid[i]], 3);
      copy(quat, &d->mocap_quat[4 * m->body_mocapid[i]], 4);
    } else if (jn == 1 && m->jnt_type[ja] == PS_JNT_FREE) {
      int qa = m->jnt_qposadr[ja];
      copy(pos, &d->qpos[qa], 3); copy(quat, &d->qpos[qa + 3], 4);
      copy(&d->xanchor[3 * ja], pos, 3); copy(&d->xaxis[3 * ja], m->jnt_axis + 3 * ja, 3);
    } else {
      double t[3]; mulMatVec3(t, xmat + 9 * pid, m->body_pos + 3 * i);
      for (int k = 0; k < 3; k++) pos[k] = xpos[3 * pid + k] + t[k];
      mulQuat(quat, xquat + 4 * pid, m->body_quat + 4 * i);
      for (int j = ja; j < ja + jn; j++) {
        int qa = m->jnt_qposadr[j];
        double* anchor = &d->xanchor[3 * j]; double* axis = &d->xaxis[3 * j];
        double v[3]; rotVecQuat(v, m->jnt_pos + 3 * j, quat);
        for (int k = 0; k < 3; k++) anchor[k] = v[k] + pos[k];
        rotVecQuat(axis, m->jnt_axis + 3 * j, quat);
        switch (m->jnt_type[j]) {
          case PS_JNT_SLIDE:
            for (int k = 0; k < 3; k++) pos[k] += axis[k] * (d->qpos[qa] - m->qpos0[qa]);
            break;
          case PS_JNT_BALL:
          case PS_JNT_HINGE: {
            double qloc[4];
            if (m->jnt_type[j] == PS_JNT_BALL) copy(qloc, &d->qpos[qa], 4);
            else axisAngle2Quat(qloc, m->jnt_axis + 3 * j, d->qpos[qa] - m->qpos0[qa]);
            mulQuat(quat, quat, qloc);
            // correct for off-center rotation
            rotVecQuat(v, m->jnt_pos + 3 * j, quat);
            for (int k = 0; k < 3; k++) pos[k] = anchor[k] - v[k];
          } break;
          default: break;
        }
      }
    }
    normalize4(quat);
    copy(xquat + 4 * i, quat, 4); copy(xpos + 3 * i, pos, 3);
    quat2Mat(xmat + 9 * i, quat);
    double t[3]; mulMatVec3(t, xmat + 9 * i, m->body_ipos + 3 * i);
    for (int k = 0; k < 3; k++) d->xipos[3 * i + k] = pos[k] + t[k];
    double qi[4]; mulQuat(qi, quat, m->body_iquat + 4 * i); quat2Mat(&d->ximat[9 * i], qi);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_bodyid[g]; double t[3], q[4];
    mulMatVec3(t, xmat + 9 * b, m->geom_pos + 3 * g);
    for (int k = 0; k < 3; k++) d->geom_xpos[3 * g + k] = xpos[3 * b + k] + t[k];
    mulQuat(q, xquat + 4 * b, m->geom_quat + 4 * g); quat2Mat(&d->geom_xmat[9 * g], q);
  }
  for (int s = 0; s < m->nsite; s++) {
    int b = m->site_bodyid[s]; double t[3], q[4];
    mulMatVec3(t, xmat + 9 * b, m->site_pos + 3 * s);
    for (int k = 0; k < 3; k++) d->site_xpos[3 * s + k] = xpos[3 * b + k] + t[k];
    mulQuat(q, xquat + 4 * b, m->site_quat + 4 * s); quat2Mat(&d->site_xmat[9 * s], q);
  }
}

// ------------------------------------------------------------------ mj_comPos
static void dofCom(double* res, const double* axis, const double* offset) {
  if (offset) { copy(res, axis, 3); cross(res + 3, axis, offset); }
  else { zero(res, 3); copy(res + 3, axis, 3); }
}

static void comPos(const Model* m, Data* d) {
  int nb = m->nbody;
  double* sc = d->subtree_com.data();
  for (int i = 0; i < nb; i++) for (int k = 0; k < 3; k++) sc[3 * i + k] = m->body_mass[i] * d->xipos[3 * i + k];
  for (int i = nb - 1; i > 0; i--) { int p = m->body_parentid[i]; for (int k = 0; k < 3; k++) sc[3 * p + k] += sc[3 * i + k]; }
  for (int i = 0; i < nb; i++) {
    if (m->body_subtreemass[i] < kMinVal) copy(sc + 3 * i, &d->xipos[3 * i], 3);
    else for (int k = 0; k < 3; k++) sc[3 * i + k] /= m->body_subtreemass[i];
  }
  zero(d->cinert.data(), 10);
  for (int i = 1; i < nb; i++) {
    double off[3]; const double* root = sc + 3 * m->body_rootid[i];
    for (int k = 0; k < 3; k++) off[k] = d->xipos[3 * i + k] - root[k];
    // mju_inertCom
    const double* mat = &d->ximat[9 * i]; const double* in = m->body_inertia + 3 * i; double mass = m->body_mass[i];
    double tmp[9];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++)
      tmp[3 * r + c] = mat[3 * r] * in[0] * mat[3 * c] + mat[3 * r + 1] * in[1] * mat[3 * c + 1] + mat[3 * r + 2] * in[2] * mat[3 * c + 2];
    double* res = &d->cinert[10 * i];
    res[0] = tmp[0] + mass * (off[1] * off[1] + off[2] * off[2]);
    res[1] = tmp[4] + mass * (off[0] * off[0] + off[2] * off[2]);
    res[2] = tmp[8] + mass * (off[0] * off[0] + off[1] * off[1]);
    res[3] = tmp[1] - mass * off[0] * off[1];
    res[4] = tmp[2] - mass * off[0] * off[2];
    res[5] = tmp[5] - mass * off[1] * off[2];
    res[6] = mass * off[0]; res[7] = mass * off[1]; res[8] = mass * off[2]; res[9] = mass;
  }
  for (int j = 0; j < m->njnt; j++) {
    int b = m->jnt_bodyid[j], da = m->jnt_dofadr[j];
    double off[3]; const double* root = sc + 3 * m->body_rootid[b];
    for (int k = 0; k < 3; k++) off[k] = root[k] - d->xanchor[3 * j + k];
    switch (m->jnt_type[j]) {
      case PS_JNT_FREE: {
        for (int k = 0; k < 3; k++) { double ax[3] = {0, 0, 0}; ax[k] = 1; dofCom(&d->cdof[6 * (da + k)], ax, nullptr); }
        da += 3;
      }  // fallthrough to ball part
      case PS_JNT_BALL:
        for (int k = 0; k < 3; k++) {
          double ax[3] = {d->xmat[9 * b + k], d->xmat[9 * b + 3 + k], d->xmat[9 * b + 6 + k]};
          dofCom(&d->cdof[6 * (da + k)], ax, off);
        }
        break;
      case PS_JNT_SLIDE: dofCom(&d->cdof[6 * da], &d->xaxis[3 * j], nullptr); break;
      case PS_JNT_HINGE: dofCom(&d->cdof[6 * da], &d->xaxis[3 * j], off); break;
    }
  }
}

// ------------------------------------------------------------------ mj_crb + mj_factorM
static void crbAndFactor(const Model* m, Data* d) {
  int nb = m->nbody, nv = m->nv;
  copy(d->crb.data(), d->cinert.data(), 10 * nb);
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    if (p > 0) for (int k = 0; k < 10; k++) d->crb[10 * p + k] += d->crb[10 * i + k];
  }
  zero(d->qM.data(), nv * nv);
  for (int i = 0; i < nv; i++) {
    double buf[6]; mulInertVec(buf, &d->crb[10 * m->dof_bodyid[i]], &d->cdof[6 * i]);
    d->qM[i * nv + i] = m->dof_armature[i];
    for (int j = i; j >= 0; j = m->dof_parentid[j]) {
      double v = dot(&d->cdof[6 * j], buf, 6);
      if (j == i) d->qM[i * nv + i] += v; else { d->qM[i * nv + j] = v; d->qM[j * nv + i] = v; }
    }
  }
  // L'DL in dof-tree order; qLD(i,j), j ancestor of i, stored at [i*nv+j]
  double* L = d->qLD.data(); copy(L, d->qM.data(), nv * nv);
  for (int k = nv - 1; k >= 0; k--) {
    if (L[k * nv + k] < kMinVal) L[k * nv + k] = kMinVal;
    for (int i = m->dof_parentid[k]; i >= 0; i = m->dof_parentid[i]) {
      double tmp = L[k * nv + i] / L[k * nv + k];
      for (int j = i; j >= 0; j = m->dof_parentid[j]) L[i * nv + j] -= L[k * nv + j] * tmp;
      L[k * nv + i] = tmp;
    }
  }
  for (int i = 0; i < nv; i++) d->qLDiagInv[i] = 1.0 / L[i * nv + i];
}

// solve with a given L'DL factor (mj_solveLD)
static void solveLD(const Model* m, double* x, const double* L, const double* dinv) {
  int nv = m->nv;
  for (int i = nv - 1; i >= 0; i--)
    if (x[i] != 0) for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[j] -= L[i * nv + j] * x[i];
  for (int i = 0; i < nv; i++) x[i] *= dinv[i];
  for (int i = 0; i < nv; i++)
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[i] -= L[i * nv + j] * x[j];
}

void solveM(const Model* m, const Data* d, double* x) { solveLD(m, x, d->qLD.data(), d->qLDiagInv.data()); }

// factor an arbitrary matrix with M's sparsity in place (mj_factorI); used by Euler damping / implicitfast
void factorI(const Model* m, double* L, double* dinv) {
  int nv = m->nv;
  for (int k = nv - 1; k >= 0; k--) {
    if (L[k * nv + k] < kMinVal) L[k * nv + k] = kMinVal;
    for (int i = m->dof_parentid[k]; i >= 0; i = m->dof_parentid[i]) {
      double tmp = L[k * nv + i] / L[k * nv + k];
      for (int j = i; j >= 0; j = m->dof_parentid[j]) L[i * nv + j] -= L[k * nv + j] * tmp;
      L[k * nv + i] = tmp;
    }
  }
  for (int i = 0; i < nv; i++) dinv[i] = 1.0 / L[i * nv + i];
}
void solveI(const Model* m, double* x, const double* L, const double* dinv) { solveLD(m, x, L, dinv); }

void mulM(const Model* m, const Data* d, double* res, const double* vec) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++) res[i] = dot(&d->qM[i * nv], vec, nv);
}

// mj_jac: translational / rotational Jacobian of a world point attached to body
void jac(const Model* m, const Data* d, double* jacp, double* jacr, const double point[3], int body) {
  int nv = m->nv;
  if (jacp) zero(jacp, 3 * nv);
  if (jacr) zero(jacr, 3 * nv);
  while (body > 0 && m->body_dofnum[body] == 0) body = m->body_parentid[body];
  if (body <= 0) return;
  double off[3]; const double* root = &d->subtree_com[3 * m->body_rootid[body]];
  for (int k = 0; k < 3; k++) off[k] = point[k] - root[k];
  for (int i = m->body_dofadr[body] + m->body_dofnum[body] - 1; i >= 0; i = m->dof_parentid[i]) {
    const double* c = &d->cdof[6 * i];
    if (jacr) for (int k = 0; k < 3; k++) jacr[k * nv + i] = c[k];
    if (jacp) { double t[3]; cross(t, c, off); for (int k = 0; k < 3; k++) jacp[k * nv + i] = c[3 + k] + t[k]; }
  }
}

void fwdPositionSmooth(const Model* m, Data* d) {
  kinematics(m, d); comPos(m, d);
  // mj_transmission (joint transmissions only)
  for (int i = 0; i < m->nu; i++) {
    int j = m->actuator_trnid[2 * i];
    d->actuator_length[i] = m->actuator_gear[i] * d->qpos[m->jnt_qposadr[j]];
  }
  crbAndFactor(m, d);
}

}  // namespace mm
