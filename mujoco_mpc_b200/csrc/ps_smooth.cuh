// Unconstrained ("smooth") rigid-body dynamics of one rollout: forward kinematics, subtree CoM
// frame, composite-rigid-body mass matrix, tree-ordered L'DL, velocity propagation, Coriolis /
// gravity bias, passive forces (spring, damper, inertia-box fluid) and actuation.
//
// This is the device replacement for what the reference obtains from MuJoCo 3.1.4 inside
// mj_step / mj_forward (call sites mjpc/trajectory.cc:158,198): mj_kinematics, mj_comPos, mj_crb,
// mj_factorM, mj_comVel, mj_passive, mj_rne, mj_fwdActuation, mj_fwdAcceleration.
#pragma once
#include "ps_math.cuh"
#include "ps_coop.cuh"

namespace ps {

// ---------------------------------------------------------------- forward kinematics
template <class T>
PS_NT void kinematics(const DevModel<T>& m, RolloutData<T>& d) {
  vzero(d.xpos, 3); d.xquat[0] = 1; d.xquat[1] = d.xquat[2] = d.xquat[3] = 0;
  vzero(d.xmat, 9); d.xmat[0] = d.xmat[4] = d.xmat[8] = 1;
  vzero(d.xipos, 3); vcopy(d.ximat, d.xmat, 9);
  if (T::QUAT) PS_SPEC_UNROLL for (int j = 0; j < m.njnt; j++) {
    if (m.jnt_type[j] == JNT_FREE) unit4(d.qpos + m.jnt_qposadr[j] + 3);
    else if (m.jnt_type[j] == JNT_BALL) unit4(d.qpos + m.jnt_qposadr[j]);
  }
  PS_SPEC_UNROLL for (int i = 0; i < m.nmocap; i++) unit4(d.mocap_quat + 4 * i);

  PS_SPEC_UNROLL for (int b = 1; b < m.nbody; b++) {
    double pos[3], quat[4];
    const int par = m.body_parentid[b], jn = m.body_jntnum[b], ja = m.body_jntadr[b];
    if (m.body_mocapid[b] >= 0) {
      vcopy(pos, d.mocap_pos + 3 * m.body_mocapid[b], 3);
      vcopy(quat, d.mocap_quat + 4 * m.body_mocapid[b], 4);
    } else if (T::QUAT && jn == 1 && m.jnt_type[ja] == JNT_FREE) {
      const int qa = m.jnt_qposadr[ja];
      vcopy(pos, d.qpos + qa, 3); vcopy(quat, d.qpos + qa + 3, 4);
      vcopy(d.xanchor + 3 * ja, pos, 3); vcopy(d.xaxis + 3 * ja, m.jnt_axis + 3 * ja, 3);
    } else {
      double t[3]; mat_vec3(t, d.xmat + 9 * par, m.body_pos + 3 * b);
      PS_SPEC_UNROLL for (int k = 0; k < 3; k++) pos[k] = d.xpos[3 * par + k] + t[k];
      quat_mul(quat, d.xquat + 4 * par, m.body_quat + 4 * b);
      PS_SPEC_UNROLL for (int j = ja; j < ja + jn; j++) {
        const int qa = m.jnt_qposadr[j];
        double* anchor = d.xanchor + 3 * j; double* axis = d.xaxis + 3 * j;
        double v[3]; quat_rotate(v, m.jnt_pos + 3 * j, quat);
        PS_SPEC_UNROLL for (int k = 0; k < 3; k++) anchor[k] = v[k] + pos[k];
        quat_rotate(axis, m.jnt_axis + 3 * j, quat);
        const int jt = m.jnt_type[j];
        if (jt == JNT_SLIDE) {
          const double dq = d.qpos[qa] - m.qpos0[qa];
          PS_SPEC_UNROLL for (int k = 0; k < 3; k++) pos[k] += axis[k] * dq;
        } else if (jt == JNT_HINGE || (T::QUAT && jt == JNT_BALL)) {
          double ql[4];
          if (T::QUAT && jt == JNT_BALL) vcopy(ql, d.qpos + qa, 4);
          else axis_angle_quat(ql, m.jnt_axis + 3 * j, d.qpos[qa] - m.qpos0[qa]);
          quat_mul(quat, quat, ql);
          quat_rotate(v, m.jnt_pos + 3 * j, quat);   // keep the anchor fixed under the rotation
          PS_SPEC_UNROLL for (int k = 0; k < 3; k++) pos[k] = anchor[k] - v[k];
        }
      }
    }
    unit4(quat);
    vcopy(d.xquat + 4 * b, quat, 4); vcopy(d.xpos + 3 * b, pos, 3);
    quat_to_mat(d.xmat + 9 * b, quat);
    double t[3]; mat_vec3(t, d.xmat + 9 * b, m.body_ipos + 3 * b);
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) d.xipos[3 * b + k] = pos[k] + t[k];
    double qi[4]; quat_mul(qi, quat, m.body_iquat + 4 * b); quat_to_mat(d.ximat + 9 * b, qi);
  }
  PS_SPEC_UNROLL for (int g = 0; g < m.ngeom; g++) {
    if (!m.geom_used[g]) continue;     // geoms that neither collide nor are read by the task
    const int b = m.geom_bodyid[g]; double t[3], q[4];
    mat_vec3(t, d.xmat + 9 * b, m.geom_pos + 3 * g);
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) d.geom_xpos[3 * g + k] = d.xpos[3 * b + k] + t[k];
    quat_mul(q, d.xquat + 4 * b, m.geom_quat + 4 * g); quat_to_mat(d.geom_xmat + 9 * g, q);
  }
  PS_SPEC_UNROLL for (int s = 0; s < m.nsite; s++) {
    const int b = m.site_bodyid[s]; double t[3], q[4];
    mat_vec3(t, d.xmat + 9 * b, m.site_pos + 3 * s);
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) d.site_xpos[3 * s + k] = d.xpos[3 * b + k] + t[k];
    quat_mul(q, d.xquat + 4 * b, m.site_quat + 4 * s); quat_to_mat(d.site_xmat + 9 * s, q);
  }
}

// ---------------------------------------------------------------- CoM frame quantities
PS_HD void dof_motion(double* res, const double* axis, const double* offset_or_null) {
  if (offset_or_null) { vcopy(res, axis, 3); cross3(res + 3, axis, offset_or_null); }
  else { vzero(res, 3); vcopy(res + 3, axis, 3); }
}

template <class T>
PS_NT void com_position(const DevModel<T>& m, RolloutData<T>& d) {
  const int nb = m.nbody;
  double* sc = d.subtree_com;
  PS_SPEC_UNROLL for (int i = 0; i < nb; i++) PS_SPEC_UNROLL for (int k = 0; k < 3; k++) sc[3 * i + k] = m.body_mass[i] * d.xipos[3 * i + k];
  PS_SPEC_UNROLL for (int i = nb - 1; i > 0; i--) { const int p = m.body_parentid[i]; PS_SPEC_UNROLL for (int k = 0; k < 3; k++) sc[3 * p + k] += sc[3 * i + k]; }
  PS_SPEC_UNROLL for (int i = 0; i < nb; i++) {
    if (m.body_subtreemass[i] < kMinVal) vcopy(sc + 3 * i, d.xipos + 3 * i, 3);
    else PS_SPEC_UNROLL for (int k = 0; k < 3; k++) sc[3 * i + k] /= m.body_subtreemass[i];
  }
  vzero(d.cinert, 10);
  PS_SPEC_UNROLL for (int i = 1; i < nb; i++) {
    double off[3]; const double* root = sc + 3 * m.body_rootid[i];
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) off[k] = d.xipos[3 * i + k] - root[k];
    const double* R = d.ximat + 9 * i; const double* I = m.body_inertia + 3 * i; const double mass = m.body_mass[i];
    // world-aligned rotational inertia R diag(I) R'
    double w00 = R[0] * I[0] * R[0] + R[1] * I[1] * R[1] + R[2] * I[2] * R[2];
    double w11 = R[3] * I[0] * R[3] + R[4] * I[1] * R[4] + R[5] * I[2] * R[5];
    double w22 = R[6] * I[0] * R[6] + R[7] * I[1] * R[7] + R[8] * I[2] * R[8];
    double w01 = R[0] * I[0] * R[3] + R[1] * I[1] * R[4] + R[2] * I[2] * R[5];
    double w02 = R[0] * I[0] * R[6] + R[1] * I[1] * R[7] + R[2] * I[2] * R[8];
    double w12 = R[3] * I[0] * R[6] + R[4] * I[1] * R[7] + R[5] * I[2] * R[8];
    double* c = d.cinert + 10 * i;
    c[0] = w00 + mass * (off[1] * off[1] + off[2] * off[2]);
    c[1] = w11 + mass * (off[0] * off[0] + off[2] * off[2]);
    c[2] = w22 + mass * (off[0] * off[0] + off[1] * off[1]);
    c[3] = w01 - mass * off[0] * off[1];
    c[4] = w02 - mass * off[0] * off[2];
    c[5] = w12 - mass * off[1] * off[2];
    c[6] = mass * off[0]; c[7] = mass * off[1]; c[8] = mass * off[2]; c[9] = mass;
  }
  PS_SPEC_UNROLL for (int j = 0; j < m.njnt; j++) {
    const int b = m.jnt_bodyid[j]; int da = m.jnt_dofadr[j];
    double off[3]; const double* root = sc + 3 * m.body_rootid[b];
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) off[k] = root[k] - d.xanchor[3 * j + k];
    const int jt = m.jnt_type[j];
    if (T::QUAT && jt == JNT_FREE) {
      PS_SPEC_UNROLL for (int k = 0; k < 3; k++) { double ax[3] = {0, 0, 0}; ax[k] = 1; dof_motion(d.cdof + 6 * (da + k), ax, nullptr); }
      da += 3;
    }
    if (T::QUAT && (jt == JNT_FREE || jt == JNT_BALL)) {
      PS_SPEC_UNROLL for (int k = 0; k < 3; k++) {
        double ax[3] = {d.xmat[9 * b + k], d.xmat[9 * b + 3 + k], d.xmat[9 * b + 6 + k]};
        dof_motion(d.cdof + 6 * (da + k), ax, off);
      }
    } else if (jt == JNT_SLIDE) dof_motion(d.cdof + 6 * da, d.xaxis + 3 * j, nullptr);
    else dof_motion(d.cdof + 6 * da, d.xaxis + 3 * j, off);
  }
}

// ---------------------------------------------------------------- mass matrix + factorisation
// factor a symmetric matrix with the dof tree's sparsity in place: L'DL, reverse dof order
template <class T>
PS_NT void factor_tree(const DevModel<T>& m, double* L, double* dinv) {
  constexpr int S = T::NV; const int nv = nv_of(m);
  PS_SPEC_UNROLL for (int k = nv - 1; k >= 0; k--) {
    if (L[k * S + k] < kMinVal) L[k * S + k] = kMinVal;
    const double dk = L[k * S + k];
    PS_CHAIN(i, m.dof_parentid[k], k, m.dof_parentid) {
      const double tmp = L[k * S + i] / dk;
      PS_CHAIN(j, i, i + 1, m.dof_parentid) L[i * S + j] -= L[k * S + j] * tmp;
      L[k * S + i] = tmp;
    }
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) dinv[i] = 1.0 / L[i * S + i];
}

template <class T>
PS_NT void solve_tree(const DevModel<T>& m, double* x, const double* L, const double* dinv) {
  constexpr int S = T::NV; const int nv = nv_of(m);
  PS_SPEC_UNROLL for (int i = nv - 1; i >= 0; i--) {
    const double xi = x[i];
    if (xi != 0) PS_CHAIN(j, m.dof_parentid[i], i, m.dof_parentid) x[j] -= L[i * S + j] * xi;
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) x[i] *= dinv[i];
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++)
    PS_CHAIN(j, m.dof_parentid[i], i, m.dof_parentid) x[i] -= L[i * S + j] * x[j];
}

template <class T>
PS_NT void mass_matrix(const DevModel<T>& m, RolloutData<T>& d) {
  constexpr int S = T::NV; const int nb = m.nbody, nv = nv_of(m);
  vcopy(d.crb, d.cinert, 10 * nb);
  PS_SPEC_UNROLL for (int i = nb - 1; i > 0; i--) {
    const int p = m.body_parentid[i];
    if (p > 0) PS_SPEC_UNROLL for (int k = 0; k < 10; k++) d.crb[10 * p + k] += d.crb[10 * i + k];
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++) d.qM[i * S + j] = 0;
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) {
    double f[6]; inert_vec(f, d.crb + 10 * m.dof_bodyid[i], d.cdof + 6 * i);
    d.qM[i * S + i] = m.dof_armature[i];
    PS_CHAIN(j, i, i + 1, m.dof_parentid) {
      const double v = vdot(d.cdof + 6 * j, f, 6);
      if (j == i) d.qM[i * S + i] += v; else { d.qM[i * S + j] = v; d.qM[j * S + i] = v; }
    }
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++) d.qLD[i * S + j] = d.qM[i * S + j];
  factor_tree(m, d.qLD, d.qLDiagInv);
}

template <class T>
PS_HD void mul_M(const DevModel<T>& m, const RolloutData<T>& d, double* res, const double* vec) {
  PS_SPEC_UNROLL for (int i = 0; i < nv_of(m); i++) res[i] = vdot(d.qM + i * T::NV, vec, nv_of(m));
}

// translational / rotational Jacobian rows of a world point attached to `body`;
// jacp / jacr are [3][T::NV] (may be null)
template <class T>
PS_NT void point_jacobian(const DevModel<T>& m, const RolloutData<T>& d, double* jacp, double* jacr,
                          const double* point, int body) {
  constexpr int S = T::NV;
  if (jacp) vzero(jacp, 3 * S);
  if (jacr) vzero(jacr, 3 * S);
  body = m.body_weldid[body];
  if (body <= 0) return;
  double off[3]; const double* root = d.subtree_com + 3 * m.body_rootid[body];
  PS_SPEC_UNROLL for (int k = 0; k < 3; k++) off[k] = point[k] - root[k];
  PS_SPEC_UNROLL for (int i = m.body_dofadr[body] + m.body_dofnum[body] - 1; i >= 0; i = m.dof_parentid[i]) {
    const double* c = d.cdof + 6 * i;
    if (jacr) PS_SPEC_UNROLL for (int k = 0; k < 3; k++) jacr[k * S + i] = c[k];
    if (jacp) { double t[3]; cross3(t, c, off); PS_SPEC_UNROLL for (int k = 0; k < 3; k++) jacp[k * S + i] = c[3 + k] + t[k]; }
  }
}

template <class T>
PS_FN void position_smooth(const DevModel<T>& m, RolloutData<T>& d) {
  kinematics(m, d);
  com_position(m, d);
  PS_SPEC_UNROLL for (int i = 0; i < m.nu; i++) d.actuator_length[i] = m.act_gear[i] * d.qpos[m.jnt_qposadr[m.act_jntid[i]]];
  mass_matrix(m, d);
}

// ---------------------------------------------------------------- velocity propagation
// cvel / cdof_dot for generalized velocity qvel (outputs may alias d's own arrays)
template <class T>
PS_NT void com_velocity(const DevModel<T>& m, const RolloutData<T>& d, const double* qvel, double* cvel, double* cdof_dot) {
  vzero(cvel, 6);
  PS_SPEC_UNROLL for (int i = 1; i < m.nbody; i++) {
    double v[6]; vcopy(v, cvel + 6 * m.body_parentid[i], 6);
    const int ja = m.body_jntadr[i];
    PS_SPEC_UNROLL for (int j = ja; j < ja + m.body_jntnum[i]; j++) {
      int da = m.jnt_dofadr[j]; const int jt = m.jnt_type[j];
      if (T::QUAT && jt == JNT_FREE) {
        PS_SPEC_UNROLL for (int k = 0; k < 3; k++) {
          vzero(cdof_dot + 6 * (da + k), 6);
          PS_SPEC_UNROLL for (int r = 0; r < 6; r++) v[r] += d.cdof[6 * (da + k) + r] * qvel[da + k];
        }
        da += 3;
      }
      if (T::QUAT && (jt == JNT_FREE || jt == JNT_BALL)) {
        PS_SPEC_UNROLL for (int k = 0; k < 3; k++) cross_motion(cdof_dot + 6 * (da + k), v, d.cdof + 6 * (da + k));
        PS_SPEC_UNROLL for (int k = 0; k < 3; k++) PS_SPEC_UNROLL for (int r = 0; r < 6; r++) v[r] += d.cdof[6 * (da + k) + r] * qvel[da + k];
      } else {
        cross_motion(cdof_dot + 6 * da, v, d.cdof + 6 * da);
        PS_SPEC_UNROLL for (int r = 0; r < 6; r++) v[r] += d.cdof[6 * da + r] * qvel[da];
      }
    }
    vcopy(cvel + 6 * i, v, 6);
  }
}

// bias = C(q, v) v + g(q) by recursive Newton-Euler with zero acceleration
template <class T>
PS_NT void rne_bias(const DevModel<T>& m, const RolloutData<T>& d, const double* qvel, const double* cvel,
                    const double* cdof_dot, double* result) {
  const int nb = m.nbody, nv = nv_of(m);
  double cacc[6 * T::NBODY], cfrc[6 * T::NBODY];
  const bool grav = !(m.disableflags & DSBL_GRAVITY);
  PS_SPEC_UNROLL for (int k = 0; k < 3; k++) { cacc[k] = 0; cacc[3 + k] = grav ? -m.gravity[k] : 0.0; }
  vzero(cfrc, 6);
  PS_SPEC_UNROLL for (int i = 1; i < nb; i++) {
    const int bda = m.body_dofadr[i];
    double* ca = cacc + 6 * i; vcopy(ca, cacc + 6 * m.body_parentid[i], 6);
    PS_SPEC_UNROLL for (int j = 0; j < m.body_dofnum[i]; j++)
      PS_SPEC_UNROLL for (int r = 0; r < 6; r++) ca[r] += cdof_dot[6 * (bda + j) + r] * qvel[bda + j];
    double t1[6], t2[6], t3[6];
    inert_vec(t1, d.cinert + 10 * i, ca);
    inert_vec(t2, d.cinert + 10 * i, cvel + 6 * i);
    cross_force(t3, cvel + 6 * i, t2);
    PS_SPEC_UNROLL for (int r = 0; r < 6; r++) cfrc[6 * i + r] = t1[r] + t3[r];
  }
  PS_SPEC_UNROLL for (int i = nb - 1; i > 0; i--) {
    const int p = m.body_parentid[i];
    if (p > 0) PS_SPEC_UNROLL for (int r = 0; r < 6; r++) cfrc[6 * p + r] += cfrc[6 * i + r];
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) result[i] = vdot(d.cdof + 6 * i, cfrc + 6 * m.dof_bodyid[i], 6);
}

// 6D velocity [ang; lin] of a frame at `pos` fixed to `body`, world orientation
template <class T, class D>
PS_HD void frame_velocity(const DevModel<T>& m, const D& d, int body, const double* pos, double* res) {
  const double* cv = body_cvel(d) + 6 * body;
  double dif[3], cr[3];
  PS_SPEC_UNROLL for (int k = 0; k < 3; k++) dif[k] = pos[k] - d.subtree_com[3 * m.body_rootid[body] + k];
  cross3(cr, dif, cv);
  PS_SPEC_UNROLL for (int k = 0; k < 3; k++) { res[k] = cv[k]; res[3 + k] = cv[3 + k] - cr[k]; }
}

// ---------------------------------------------------------------- inertia-box fluid model
template <class T>
PS_HD void inertia_box(const DevModel<T>& m, int i, double* box) {
  const double* I = m.body_inertia + 3 * i; const double mass = m.body_mass[i];
  box[0] = sqrt(fmax(kMinVal, I[1] + I[2] - I[0]) / mass * 6.0);
  box[1] = sqrt(fmax(kMinVal, I[0] + I[2] - I[1]) / mass * 6.0);
  box[2] = sqrt(fmax(kMinVal, I[0] + I[1] - I[2]) / mass * 6.0);
}

// local-frame velocity [ang; lin] of body i's inertial frame, wind removed
template <class T>
PS_HD void body_local_velocity(const DevModel<T>& m, const RolloutData<T>& d, int i, double* lvel) {
  double v[6]; frame_velocity(m, d, i, d.xipos + 3 * i, v);
  matT_vec3(lvel, d.ximat + 9 * i, v);
  matT_vec3(lvel + 3, d.ximat + 9 * i, v + 3);
  double lw[3]; matT_vec3(lw, d.ximat + 9 * i, m.wind);
  PS_SPEC_UNROLL for (int k = 0; k < 3; k++) lvel[3 + k] -= lw[k];
}

template <class T>
PS_NT void fluid_forces(const DevModel<T>& m, const RolloutData<T>& d, double* qfrc) {
  constexpr int S = T::NV; const int nv = nv_of(m);
  double jp[3 * S], jr[3 * S];
  PS_SPEC_UNROLL for (int i = 1; i < m.nbody; i++) {
    if (m.body_mass[i] < kMinVal) continue;
    double box[3], lvel[6], lf[6], wf[6];
    inertia_box(m, i, box);
    body_local_velocity(m, d, i, lvel);
    vzero(lf, 6);
    const double visc = m.viscosity, dens = m.density;
    if (visc > 0) {
      const double diam = (box[0] + box[1] + box[2]) / 3.0;
      PS_SPEC_UNROLL for (int k = 0; k < 3; k++) {
        lf[k] = -kPi * diam * diam * diam * visc * lvel[k];
        lf[3 + k] = -3.0 * kPi * diam * visc * lvel[3 + k];
      }
    }
    if (dens > 0) {
      lf[3] -= 0.5 * dens * box[1] * box[2] * fabs(lvel[3]) * lvel[3];
      lf[4] -= 0.5 * dens * box[0] * box[2] * fabs(lvel[4]) * lvel[4];
      lf[5] -= 0.5 * dens * box[0] * box[1] * fabs(lvel[5]) * lvel[5];
      const double b0 = box[0], b1 = box[1], b2 = box[2];
      lf[0] -= dens * b0 * (b1 * b1 * b1 * b1 + b2 * b2 * b2 * b2) * fabs(lvel[0]) * lvel[0] / 64.0;
      lf[1] -= dens * b1 * (b0 * b0 * b0 * b0 + b2 * b2 * b2 * b2) * fabs(lvel[1]) * lvel[1] / 64.0;
      lf[2] -= dens * b2 * (b0 * b0 * b0 * b0 + b1 * b1 * b1 * b1) * fabs(lvel[2]) * lvel[2] / 64.0;
    }
    mat_vec3(wf, d.ximat + 9 * i, lf);           // torque, world
    mat_vec3(wf + 3, d.ximat + 9 * i, lf + 3);   // force, world
    point_jacobian(m, d, jp, jr, d.xipos + 3 * i, i);
    PS_SPEC_UNROLL for (int c = 0; c < nv; c++)
      PS_SPEC_UNROLL for (int k = 0; k < 3; k++) qfrc[c] += jp[k * S + c] * wf[3 + k] + jr[k * S + c] * wf[k];
  }
}

template <class T>
PS_NT void velocity_stage(const DevModel<T>& m, RolloutData<T>& d) {
  const int nv = nv_of(m);
  PS_SPEC_UNROLL for (int i = 0; i < m.nu; i++) d.actuator_velocity[i] = m.act_gear[i] * d.qvel[m.jnt_dofadr[m.act_jntid[i]]];
  com_velocity(m, d, d.qvel, d.cvel, d.cdof_dot);
  // passive forces: spring + damper + fluid
  if (m.disableflags & DSBL_PASSIVE) vzero(d.qfrc_passive, nv);
  else {
    double spring[T::NV], fluid[T::NV];
    vzero(spring, nv); vzero(fluid, nv);
    PS_SPEC_UNROLL for (int j = 0; j < m.njnt; j++) {
      const double k = m.jnt_stiffness[j];
      if (k == 0) continue;
      if (m.jnt_type[j] == JNT_HINGE || m.jnt_type[j] == JNT_SLIDE)
        spring[m.jnt_dofadr[j]] = -k * (d.qpos[m.jnt_qposadr[j]] - m.qpos_spring[m.jnt_qposadr[j]]);
    }
    if (!T::LEAN && (m.viscosity > 0 || m.density > 0)) fluid_forces(m, d, fluid);
    PS_SPEC_UNROLL for (int i = 0; i < nv; i++) d.qfrc_passive[i] = spring[i] + (-m.dof_damping[i] * d.qvel[i]) + fluid[i];
  }
  rne_bias(m, d, d.qvel, d.cvel, d.cdof_dot, d.qfrc_bias);
}

// ---------------------------------------------------------------- actuation + smooth acceleration
template <class T>
PS_NT void actuation_stage(const DevModel<T>& m, RolloutData<T>& d) {
  const int nv = nv_of(m), nu = m.nu;
  vzero(d.qfrc_actuator, nv); vzero(d.actuator_force, nu);
  PS_SPEC_UNROLL for (int i = 0; i < m.na; i++) d.act_dot[i] = 0;
  if (!nu || (m.disableflags & DSBL_ACTUATION)) return;
  PS_SPEC_UNROLL for (int i = 0; i < nu; i++) {
    double ctrl = d.ctrl[i];
    if (bad_number(ctrl)) { d.warning |= WARN_BADCTRL; ctrl = 0; PS_SPEC_UNROLL for (int k = 0; k < nu; k++) d.ctrl[k] = 0; }
    if (m.act_ctrllimited[i] && !(m.disableflags & DSBL_CLAMPCTRL))
      ctrl = clampd(ctrl, m.act_ctrlrange[2 * i], m.act_ctrlrange[2 * i + 1]);
    double in = ctrl;
    const int dyn = m.act_dyntype[i];
    if (!T::LEAN && dyn != DYN_NONE) {
      const int aa = m.act_actadr[i];
      if (dyn == DYN_INTEGRATOR) d.act_dot[aa] = ctrl;
      else d.act_dot[aa] = (ctrl - d.act[aa]) / fmax(kMinVal, m.act_dynprm[3 * i]);
      in = d.act[aa];
    }
    const double* gp = m.act_gainprm + 3 * i; const double* bp = m.act_biasprm + 3 * i;
    double gain = gp[0];
    if (m.act_gaintype[i] == 1) gain = gp[0] + gp[1] * d.actuator_length[i] + gp[2] * d.actuator_velocity[i];
    double bias = 0;
    if (m.act_biastype[i] == 1) bias = bp[0] + bp[1] * d.actuator_length[i] + bp[2] * d.actuator_velocity[i];
    double f = gain * in + bias;
    if (m.act_forcelimited[i]) f = clampd(f, m.act_forcerange[2 * i], m.act_forcerange[2 * i + 1]);
    d.actuator_force[i] = f;
    d.qfrc_actuator[m.jnt_dofadr[m.act_jntid[i]]] += m.act_gear[i] * f;
  }
}

template <class T>
PS_NT void acceleration_stage(const DevModel<T>& m, RolloutData<T>& d) {
  const int nv = nv_of(m);
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) {
    d.qfrc_smooth[i] = d.qfrc_passive[i] - d.qfrc_bias[i] + d.qfrc_actuator[i];
    d.qacc_smooth[i] = d.qfrc_smooth[i];
  }
  solve_tree(m, d.qacc_smooth, d.qLD, d.qLDiagInv);
}

// the same stage with Cartesian perturbation forces (mjData.xfrc_applied, [force; torque] per body) added to
// qfrc_smooth by mj_xfrcAccumulate: J(xipos_b, b)' x for every body with a non-zero entry.  A separate function so
// that the planning kernels keep the exact code they were verified with.
template <class T>
PS_NT void acceleration_stage_xfrc(const DevModel<T>& m, RolloutData<T>& d, const double* xfrc_applied) {
  constexpr int S = T::NV; const int nv = nv_of(m);
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) d.qfrc_smooth[i] = d.qfrc_passive[i] - d.qfrc_bias[i] + d.qfrc_actuator[i];
  double jp[3 * S], jr[3 * S];
  PS_SPEC_UNROLL for (int b = 1; b < m.nbody; b++) {
    const double* x = xfrc_applied + 6 * b;
    if (x[0] == 0 && x[1] == 0 && x[2] == 0 && x[3] == 0 && x[4] == 0 && x[5] == 0) continue;
    point_jacobian(m, d, jp, jr, d.xipos + 3 * b, b);
    PS_SPEC_UNROLL for (int c = 0; c < nv; c++)
      PS_SPEC_UNROLL for (int k = 0; k < 3; k++) d.qfrc_smooth[c] += jp[k * S + c] * x[k] + jr[k * S + c] * x[3 + k];
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) d.qacc_smooth[i] = d.qfrc_smooth[i];
  solve_tree(m, d.qacc_smooth, d.qLD, d.qLDiagInv);
}

}  // namespace ps
