// One physics step of one rollout: stage order of the forward pass, on-demand sensors, state
// checks and the Euler (implicit joint damping) / implicit / implicitfast integrators.
//
// Device replacement for mj_step / mj_forward as called by Trajectory::NoisyRollout
// (mjpc/trajectory.cc:158,198).  The residual is evaluated by the caller between forward_pass()
// and integrate(), i.e. at the pre-step state with the control just written -- the timing the
// reference gets from MuJoCo's mjcb_sensor callback at mjSTAGE_ACC (mjpc/testspeed.cc:36-40).
#pragma once
#include "ps_constraint.cuh"

namespace ps {

// ---------------------------------------------------------------- subtree momentum (sensor support)
template <class T>
PS_NT void subtree_velocity(const DevModel<T>& m, RolloutData<T>& d) {
  const int nb = m.nbody;
  double bodyvel[6 * T::NBODY];
  double* lv = d.subtree_linvel; double* am = d.subtree_angmom;
  PS_SPEC_UNROLL for (int i = 0; i < nb; i++) {
    frame_velocity(m, d, i, d.xipos + 3 * i, bodyvel + 6 * i);
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) lv[3 * i + k] = m.body_mass[i] * bodyvel[6 * i + 3 + k];
    double t[3], t2[3]; matT_vec3(t, d.ximat + 9 * i, bodyvel + 6 * i);
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) t[k] *= m.body_inertia[3 * i + k];
    mat_vec3(t2, d.ximat + 9 * i, t); vcopy(am + 3 * i, t2, 3);
  }
  PS_SPEC_UNROLL for (int i = nb - 1; i >= 0; i--) {
    if (i) PS_SPEC_UNROLL for (int k = 0; k < 3; k++) lv[3 * m.body_parentid[i] + k] += lv[3 * i + k];
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) lv[3 * i + k] /= fmax(kMinVal, m.body_subtreemass[i]);
  }
  PS_SPEC_UNROLL for (int i = nb - 1; i > 0; i--) {
    const int p = m.body_parentid[i]; double dx[3], dv[3], dL[3];
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) { dx[k] = d.xipos[3 * i + k] - d.subtree_com[3 * i + k]; dv[k] = (bodyvel[6 * i + 3 + k] - lv[3 * i + k]) * m.body_mass[i]; }
    cross3(dL, dx, dv);
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) { am[3 * i + k] += dL[k]; am[3 * p + k] += am[3 * i + k]; }
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) { dx[k] = d.subtree_com[3 * i + k] - d.subtree_com[3 * p + k]; dv[k] = (lv[3 * i + k] - lv[3 * p + k]) * m.body_subtreemass[i]; }
    cross3(dL, dx, dv);
    PS_SPEC_UNROLL for (int k = 0; k < 3; k++) am[3 * p + k] += dL[k];
  }
}

// evaluate model sensor `s` from the current rollout data (replaces SensorByName + sensordata,
// mjpc/utilities.cc:166-175); returns the number of values written
template <class T, class D>
PS_NT int read_sensor(const DevModel<T>& m, const D& d, int s, double* out) {
  const int type = m.sensor_type[s], ot = m.sensor_objtype[s], id = m.sensor_objid[s];
  const double* pos; const double* mat; int body;
  if (ot == OBJ_BODY) { pos = d.xipos + 3 * id; mat = d.ximat + 9 * id; body = id; }
  else if (ot == OBJ_XBODY) { pos = d.xpos + 3 * id; mat = d.xmat + 9 * id; body = id; }
  else if (ot == OBJ_GEOM) { pos = geom_world_pos(m, d, id); mat = geom_world_mat(m, d, id); body = m.geom_bodyid[id]; }
  else { pos = d.site_xpos + 3 * id; mat = d.site_xmat + 9 * id; body = m.site_bodyid[id]; }
  switch (type) {
    case SENS_JOINTPOS: out[0] = d.qpos[m.jnt_qposadr[id]]; return 1;
    case SENS_JOINTVEL: out[0] = d.qvel[m.jnt_dofadr[id]]; return 1;
    case SENS_FRAMEPOS: vcopy(out, pos, 3); return 3;
    case SENS_FRAMEXAXIS: case SENS_FRAMEYAXIS: case SENS_FRAMEZAXIS: {
      const int c = type - SENS_FRAMEXAXIS; out[0] = mat[c]; out[1] = mat[3 + c]; out[2] = mat[6 + c]; return 3; }
    case SENS_FRAMELINVEL: { double v[6]; frame_velocity(m, d, body, pos, v); vcopy(out, v + 3, 3); return 3; }
    case SENS_FRAMEANGVEL: { double v[6]; frame_velocity(m, d, body, pos, v); vcopy(out, v, 3); return 3; }
    case SENS_SUBTREECOM: vcopy(out, d.subtree_com + 3 * id, 3); return 3;
    case SENS_SUBTREELINVEL: vcopy(out, d.subtree_linvel + 3 * id, 3); return 3;
    case SENS_SUBTREEANGMOM: vcopy(out, d.subtree_angmom + 3 * id, 3); return 3;
    default: out[0] = out[1] = out[2] = 0; return 3;
  }
}

// ---------------------------------------------------------------- forward pass
template <class T>
PS_FN void forward_pass(const DevModel<T>& m, RolloutData<T>& d) {
  position_smooth(m, d);
  make_constraints(m, d);
  velocity_stage(m, d);
  reference_acceleration(m, d);
  if (m.need_subtreevel) subtree_velocity(m, d);
  actuation_stage(m, d);
  acceleration_stage(m, d);
  constraint_stage(m, d);
}

// forward pass under Cartesian perturbation forces (robust_rollout_kernel only)
template <class T>
PS_FN void forward_pass_xfrc(const DevModel<T>& m, RolloutData<T>& d, const double* xfrc_applied) {
  position_smooth(m, d);
  make_constraints(m, d);
  velocity_stage(m, d);
  reference_acceleration(m, d);
  if (m.need_subtreevel) subtree_velocity(m, d);
  actuation_stage(m, d);
  acceleration_stage_xfrc(m, d, xfrc_applied);
  constraint_stage(m, d);
}

// ---------------------------------------------------------------- d(qfrc_smooth)/d(qvel) for implicit integrators
template <class T>
PS_NT void smooth_vel_derivative(const DevModel<T>& m, RolloutData<T>& d, double* D, bool with_bias) {
  constexpr int S = T::NV; const int nv = nv_of(m);
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++) D[i * S + j] = 0;
  PS_SPEC_UNROLL for (int i = 0; i < m.nu; i++) {
    double bv = 0;
    if (m.act_biastype[i] == 1) bv += m.act_biasprm[3 * i + 2];
    if (m.act_gaintype[i] == 1) {
      const double in = m.act_dyntype[i] == DYN_NONE ? d.ctrl[i] : d.act[m.act_actadr[i]];
      bv += m.act_gainprm[3 * i + 2] * in;
    }
    if (bv != 0) { const int dof = m.jnt_dofadr[m.act_jntid[i]]; const double g = m.act_gear[i]; D[dof * S + dof] += g * g * bv; }
  }
  if (!(m.disableflags & DSBL_PASSIVE)) {
    PS_SPEC_UNROLL for (int i = 0; i < nv; i++) D[i * S + i] -= m.dof_damping[i];
    if (m.viscosity > 0 || m.density > 0) {
      double jp[3 * S], jr[3 * S], Jl[6 * S];
      PS_SPEC_UNROLL for (int b = 1; b < m.nbody; b++) {
        if (m.body_mass[b] < kMinVal) continue;
        double box[3], lvel[6]; inertia_box(m, b, box); body_local_velocity(m, d, b, lvel);
        double B[6] = {0, 0, 0, 0, 0, 0}; const double dens = m.density, visc = m.viscosity;
        if (visc > 0) { const double diam = (box[0] + box[1] + box[2]) / 3.0; PS_SPEC_UNROLL for (int k = 0; k < 3; k++) { B[k] -= kPi * diam * diam * diam * visc; B[3 + k] -= 3.0 * kPi * diam * visc; } }
        if (dens > 0) {
          B[3] -= dens * box[1] * box[2] * fabs(lvel[3]); B[4] -= dens * box[0] * box[2] * fabs(lvel[4]); B[5] -= dens * box[0] * box[1] * fabs(lvel[5]);
          const double b0 = box[0], b1 = box[1], b2 = box[2];
          B[0] -= dens * b0 * (b1 * b1 * b1 * b1 + b2 * b2 * b2 * b2) * 2 * fabs(lvel[0]) / 64.0;
          B[1] -= dens * b1 * (b0 * b0 * b0 * b0 + b2 * b2 * b2 * b2) * 2 * fabs(lvel[1]) / 64.0;
          B[2] -= dens * b2 * (b0 * b0 * b0 * b0 + b1 * b1 * b1 * b1) * 2 * fabs(lvel[2]) / 64.0;
        }
        point_jacobian(m, d, jp, jr, d.xipos + 3 * b, b);
        const double* R = d.ximat + 9 * b;
        PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int r = 0; r < 3; r++) {
          Jl[r * S + i] = R[r] * jr[i] + R[3 + r] * jr[S + i] + R[6 + r] * jr[2 * S + i];
          Jl[(3 + r) * S + i] = R[r] * jp[i] + R[3 + r] * jp[S + i] + R[6 + r] * jp[2 * S + i];
        }
        PS_SPEC_UNROLL for (int r = 0; r < 6; r++) PS_SPEC_UNROLL for (int i = 0; i < nv; i++) {
          const double s = B[r] * Jl[r * S + i]; if (s == 0) continue;
          PS_SPEC_UNROLL for (int j = 0; j < nv; j++) D[i * S + j] += s * Jl[r * S + j];
        }
      }
    }
  }
  if (with_bias) {
    // bias(v) = c(v) + g with c a homogeneous quadratic form: column j of dc/dv is exactly
    // c(v + e_j) - c(v) - c(e_j)
    double g[S], cv[S], ce[S], cve[S], v[S], cvel[6 * T::NBODY], cdd[6 * S];
    vzero(v, nv); com_velocity(m, d, v, cvel, cdd); rne_bias(m, d, v, cvel, cdd, g);
    com_velocity(m, d, d.qvel, cvel, cdd); rne_bias(m, d, d.qvel, cvel, cdd, cv);
    PS_SPEC_UNROLL for (int j = 0; j < nv; j++) {
      vzero(v, nv); v[j] = 1; com_velocity(m, d, v, cvel, cdd); rne_bias(m, d, v, cvel, cdd, ce);
      vcopy(v, d.qvel, nv); v[j] += 1; com_velocity(m, d, v, cvel, cdd); rne_bias(m, d, v, cvel, cdd, cve);
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) D[i * S + j] -= (cve[i] - g[i]) - (cv[i] - g[i]) - (ce[i] - g[i]);
    }
  }
}

// ---------------------------------------------------------------- integration
template <class T>
PS_NT void integrate(const DevModel<T>& m, RolloutData<T>& d) {
  constexpr int S = T::NV; const int nv = nv_of(m); const double dt = m.timestep;
  double qacc[S];
  if (T::LEAN || m.integrator == INT_EULER || m.integrator == INT_RK4) {
    if (!m.any_damping || (m.disableflags & DSBL_EULERDAMP)) vcopy(qacc, d.qacc, nv);
    else {
      double* L = d.sv_H; double dinv[S];
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++) L[i * S + j] = d.qM[i * S + j];
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) L[i * S + i] += dt * m.dof_damping[i];
      factor_tree(m, L, dinv);
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) qacc[i] = d.qfrc_smooth[i] + d.qfrc_constraint[i];
      solve_tree(m, qacc, L, dinv);
    }
  } else {
    const bool full = m.integrator == INT_IMPLICIT;
    double D[S * S]; double* L = d.sv_H;
    smooth_vel_derivative(m, d, D, full);
    PS_SPEC_UNROLL for (int i = 0; i < nv; i++) qacc[i] = d.qfrc_smooth[i] + d.qfrc_constraint[i];
    if (full) {
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++) L[i * S + j] = d.qM[i * S + j] - dt * D[i * S + j];
      PS_SPEC_UNROLL for (int k = 0; k < nv; k++)
        PS_SPEC_UNROLL for (int i = k + 1; i < nv; i++) {
          const double f = L[i * S + k] / L[k * S + k]; L[i * S + k] = f;
          PS_SPEC_UNROLL for (int j = k + 1; j < nv; j++) L[i * S + j] -= f * L[k * S + j];
        }
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int k = 0; k < i; k++) qacc[i] -= L[i * S + k] * qacc[k];
      PS_SPEC_UNROLL for (int i = nv - 1; i >= 0; i--) { PS_SPEC_UNROLL for (int k = i + 1; k < nv; k++) qacc[i] -= L[i * S + k] * qacc[k]; qacc[i] /= L[i * S + i]; }
    } else {
      double dinv[S];
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++)
        L[i * S + j] = d.qM[i * S + j] - dt * 0.5 * (D[i * S + j] + D[j * S + i]);
      factor_tree(m, L, dinv);
      solve_tree(m, qacc, L, dinv);
    }
  }
  // advance activations, velocities, positions, time
  if (!T::LEAN && m.na && !(m.disableflags & DSBL_ACTUATION)) {
    PS_SPEC_UNROLL for (int i = 0; i < m.nu; i++) {
      if (m.act_dyntype[i] == DYN_NONE) continue;
      const int a = m.act_actadr[i];
      double act = d.act[a];
      if (m.act_dyntype[i] == DYN_FILTEREXACT) {
        const double tau = fmax(kMinVal, m.act_dynprm[3 * i]);
        act = act + d.act_dot[a] * tau * (1 - exp(-dt / tau));
      } else act = act + d.act_dot[a] * dt;
      if (m.act_actlimited[i]) act = clampd(act, m.act_actrange[2 * i], m.act_actrange[2 * i + 1]);
      d.act[a] = act;
    }
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) d.qvel[i] += qacc[i] * dt;
  PS_SPEC_UNROLL for (int j = 0; j < m.njnt; j++) {
    const int qa = m.jnt_qposadr[j], da = m.jnt_dofadr[j], jt = m.jnt_type[j];
    if (T::QUAT && jt == JNT_FREE) {
      PS_SPEC_UNROLL for (int k = 0; k < 3; k++) d.qpos[qa + k] += dt * d.qvel[da + k];
      quat_integrate(d.qpos + qa + 3, d.qvel + da + 3, dt);
    } else if (T::QUAT && jt == JNT_BALL) quat_integrate(d.qpos + qa, d.qvel + da, dt);
    else d.qpos[qa] += dt * d.qvel[da];
  }
  d.time += dt;
}

// state sanity checks that bracket the forward pass in mj_step; a flagged rollout is scored 1e6
// by the caller (CheckWarnings, mjpc/utilities.cc:787-800; mjpc/trajectory.cc:169-173)
template <class T>
PS_HD bool state_is_bad(const DevModel<T>& m, RolloutData<T>& d) {
  PS_SPEC_UNROLL for (int i = 0; i < nq_of(m); i++) if (bad_number(d.qpos[i])) { d.warning |= WARN_BADQPOS; return true; }
  PS_SPEC_UNROLL for (int i = 0; i < nv_of(m); i++) if (bad_number(d.qvel[i])) { d.warning |= WARN_BADQVEL; return true; }
  return false;
}
template <class T>
PS_HD bool acc_is_bad(const DevModel<T>& m, RolloutData<T>& d) {
  PS_SPEC_UNROLL for (int i = 0; i < nv_of(m); i++) if (bad_number(d.qacc[i])) { d.warning |= WARN_BADQACC; return true; }
  return false;
}

}  // namespace ps
