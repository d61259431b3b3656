// Warp-cooperative smooth dynamics: the phases of mj_kinematics / mj_comPos / mj_crb / mj_factorM /
// mj_comVel / mj_rne / mj_passive / mj_fwdActuation / mj_fwdAcceleration for ONE rollout whose
// working set (CoopData) lives in shared memory.  Per-item arithmetic is that of ps_smooth.cuh.
#pragma once
#include "ps_smooth.cuh"

namespace ps {

// ---------------------------------------------------------------- kinematics, one tree level per phase
template <class T, class LT>
PS_NI void co_body_frame(const DevModel<T>& m, CoopData<T>& d, int b) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  double pos[3], quat[4];
  const int par = m.body_parentid[b], jn = m.body_jntnum[b], ja = m.body_jntadr[b];
  if (m.body_mocapid[b] >= 0) {
    vcopy(pos, d.mocap_pos + 3 * m.body_mocapid[b], 3);
    vcopy(quat, d.mocap_quat + 4 * m.body_mocapid[b], 4);
  } else if (jn == 1 && m.jnt_type[ja] == JNT_FREE) {
    const int qa = m.jnt_qposadr[ja];
    vcopy(pos, d.qpos + qa, 3); vcopy(quat, d.qpos + qa + 3, 4);
    vcopy(d.xanchor + 3 * ja, pos, 3); vcopy(d.xaxis + 3 * ja, m.jnt_axis + 3 * ja, 3);
  } else {
    double t[3]; mat_vec3(t, d.xmat + 9 * par, m.body_pos + 3 * b);
    PS_LOOP2 for (int k = 0; k < 3; k++) pos[k] = d.xpos[3 * par + k] + t[k];
    quat_mul(quat, d.xquat + 4 * par, m.body_quat + 4 * b);
    PS_LOOP2 for (int j = ja; j < ja + jn; j++) {
      const int qa = m.jnt_qposadr[j];
      double* anchor = d.xanchor + 3 * j; double* axis = d.xaxis + 3 * j;
      double v[3]; quat_rotate(v, m.jnt_pos + 3 * j, quat);
      PS_LOOP2 for (int k = 0; k < 3; k++) anchor[k] = v[k] + pos[k];
      quat_rotate(axis, m.jnt_axis + 3 * j, quat);
      const int jt = m.jnt_type[j];
      if (jt == JNT_SLIDE) {
        const double dq = d.qpos[qa] - m.qpos0[qa];
        PS_LOOP2 for (int k = 0; k < 3; k++) pos[k] += axis[k] * dq;
      } else if (jt == JNT_HINGE || jt == JNT_BALL) {
        double ql[4];
        if (jt == JNT_BALL) vcopy(ql, d.qpos + qa, 4);
        else axis_angle_quat(ql, m.jnt_axis + 3 * j, d.qpos[qa] - m.qpos0[qa]);
        quat_mul(quat, quat, ql);
        quat_rotate(v, m.jnt_pos + 3 * j, quat);
        PS_LOOP2 for (int k = 0; k < 3; k++) pos[k] = anchor[k] - v[k];
      }
    }
  }
  unit4(quat);
  vcopy(d.xquat + 4 * b, quat, 4); vcopy(d.xpos + 3 * b, pos, 3);
  quat_to_mat(d.xmat + 9 * b, quat);
  double t[3]; mat_vec3(t, d.xmat + 9 * b, m.body_ipos + 3 * b);
  PS_LOOP2 for (int k = 0; k < 3; k++) d.xipos[3 * b + k] = pos[k] + t[k];
  double qi[4]; quat_mul(qi, quat, m.body_iquat + 4 * b); quat_to_mat(d.ximat + 9 * b, qi);
}

template <class T, class LT>
PS_NI void co_kinematics(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  // quaternion normalisation of the state (mj_kinematics does it in place) + world body
  CO_FOR(j, m.njnt) {
    if (m.jnt_type[j] == JNT_FREE) unit4(d.qpos + m.jnt_qposadr[j] + 3);
    else if (m.jnt_type[j] == JNT_BALL) unit4(d.qpos + m.jnt_qposadr[j]);
  }
  CO_FOR(i, m.nmocap) unit4(d.mocap_quat + 4 * i);
  if (L.lane == 0) {
    vzero(d.xpos, 3); d.xquat[0] = 1; d.xquat[1] = d.xquat[2] = d.xquat[3] = 0;
    vzero(d.xmat, 9); d.xmat[0] = d.xmat[4] = d.xmat[8] = 1;
    vzero(d.xipos, 3); vzero(d.ximat, 9); d.ximat[0] = d.ximat[4] = d.ximat[8] = 1;
  }
  L.sync();
  PS_LOOP2 for (int l = 1; l < m.nlevel; l++) {
    const int a = m.level_adr[l], n = m.level_adr[l + 1] - a;
    CO_FOR(k, n) co_body_frame<T, LT>(m, d, m.level_body[a + k]);
    L.sync();
  }
  CO_FOR(s, m.nsite) {
    const int b = m.site_bodyid[s]; double t[3], q[4];
    mat_vec3(t, d.xmat + 9 * b, m.site_pos + 3 * s);
    PS_LOOP2 for (int k = 0; k < 3; k++) d.site_xpos[3 * s + k] = d.xpos[3 * b + k] + t[k];
    quat_mul(q, d.xquat + 4 * b, m.site_quat + 4 * s); quat_to_mat(d.site_xmat + 9 * s, q);
  }
  CO_FOR(k, m.ntrack) geom_pose(m, d, m.track_geom[k], d.track_xpos + 3 * k, d.track_xmat + 9 * k);
}

// fold children into parents, deepest level first: x[p] += x[c] for c in children(p), decreasing c,
// `n` values per body.  include_world: whether body 0 accumulates too.
// the same for ONE working lane and N values per body known at compile time: a parent's N sums sit in registers while its
// children go by (each element still receives its children in list order), a child's N loads are independent of each other
template <int N, class T>
PS_FN void co_fold_children_single(const DevModel<T>& m, double* x, bool include_world) {
  PS_LOOP2 for (int l = m.nlevel - 2; l >= (include_world ? 0 : 1); l--) {
    const int a = m.level_adr[l], nb = m.level_adr[l + 1] - a;
    PS_LOOP2 for (int bi = 0; bi < nb; bi++) {
      const int p = m.level_body[a + bi];
      const int c0 = m.child_adr[p], c1 = m.child_adr[p + 1];
      if (c0 == c1) continue;                        // a leaf keeps its values
      double acc[N];
#pragma unroll
      for (int k = 0; k < N; k++) acc[k] = x[N * p + k];
      PS_LOOP2 for (int c = c0; c < c1; c++) {
        const double* xc = x + N * m.child_list[c];
        double v[N];
#pragma unroll
        for (int k = 0; k < N; k++) v[k] = xc[k];
#pragma unroll
        for (int k = 0; k < N; k++) acc[k] += v[k];
      }
#pragma unroll
      for (int k = 0; k < N; k++) x[N * p + k] = acc[k];
    }
  }
}

template <class T, class LT>
PS_NI void co_fold_children(const LT& L, const DevModel<T>& m, double* x, int n, bool include_world) { PS_ASSUME_SHARED(&m);
  if constexpr (LT::single) {
    if (n == 3) { co_fold_children_single<3>(m, x, include_world); return; }
    if (n == 6) { co_fold_children_single<6>(m, x, include_world); return; }
    if (n == 10) { co_fold_children_single<10>(m, x, include_world); return; }
  }
  PS_LOOP2 for (int l = m.nlevel - 2; l >= (include_world ? 0 : 1); l--) {
    const int a = m.level_adr[l], nb = m.level_adr[l + 1] - a;
    CO_FOR2(w, nb, bi, n, k) {
      const int p = m.level_body[a + bi];
      double s = x[n * p + k];
      PS_LOOP2 for (int c = m.child_adr[p]; c < m.child_adr[p + 1]; c++) s += x[n * m.child_list[c] + k];
      x[n * p + k] = s;
    }
    L.sync();
  }
}

// ---------------------------------------------------------------- CoM frame quantities
template <class T, class LT>
PS_NI void co_com_position(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nb = m.nbody;
  double* sc = d.subtree_com;
  CO_FOR(w, 3 * nb) sc[w] = m.body_mass[w / 3] * d.xipos[w];
  L.sync();
  co_fold_children(L, m, sc, 3, true);
  CO_FOR(i, nb) {
    if (m.body_subtreemass[i] < kMinVal) vcopy(sc + 3 * i, d.xipos + 3 * i, 3);
    else PS_LOOP2 for (int k = 0; k < 3; k++) sc[3 * i + k] /= m.body_subtreemass[i];
  }
  L.sync();
  CO_FOR(i, nb) {
    double* c = d.cinert + 10 * i;
    if (i == 0) { vzero(c, 10); continue; }
    double off[3]; const double* root = sc + 3 * m.body_rootid[i];
    PS_LOOP2 for (int k = 0; k < 3; k++) off[k] = d.xipos[3 * i + k] - root[k];
    const double* R = d.ximat + 9 * i; const double* I = m.body_inertia + 3 * i; const double mass = m.body_mass[i];
    const double w00 = R[0] * I[0] * R[0] + R[1] * I[1] * R[1] + R[2] * I[2] * R[2];
    const double w11 = R[3] * I[0] * R[3] + R[4] * I[1] * R[4] + R[5] * I[2] * R[5];
    const double w22 = R[6] * I[0] * R[6] + R[7] * I[1] * R[7] + R[8] * I[2] * R[8];
    const double w01 = R[0] * I[0] * R[3] + R[1] * I[1] * R[4] + R[2] * I[2] * R[5];
    const double w02 = R[0] * I[0] * R[6] + R[1] * I[1] * R[7] + R[2] * I[2] * R[8];
    const double w12 = R[3] * I[0] * R[6] + R[4] * I[1] * R[7] + R[5] * I[2] * R[8];
    c[0] = w00 + mass * (off[1] * off[1] + off[2] * off[2]);
    c[1] = w11 + mass * (off[0] * off[0] + off[2] * off[2]);
    c[2] = w22 + mass * (off[0] * off[0] + off[1] * off[1]);
    c[3] = w01 - mass * off[0] * off[1];
    c[4] = w02 - mass * off[0] * off[2];
    c[5] = w12 - mass * off[1] * off[2];
    c[6] = mass * off[0]; c[7] = mass * off[1]; c[8] = mass * off[2]; c[9] = mass;
  }
  CO_FOR(j, m.njnt) {
    const int b = m.jnt_bodyid[j]; int da = m.jnt_dofadr[j];
    double off[3]; const double* root = sc + 3 * m.body_rootid[b];
    PS_LOOP2 for (int k = 0; k < 3; k++) off[k] = root[k] - d.xanchor[3 * j + k];
    const int jt = m.jnt_type[j];
    if (jt == JNT_FREE) {
      PS_LOOP2 for (int k = 0; k < 3; k++) { double ax[3] = {0, 0, 0}; ax[k] = 1; dof_motion(d.cdof + 6 * (da + k), ax, nullptr); }
      da += 3;
    }
    if (jt == JNT_FREE || jt == JNT_BALL) {
      PS_LOOP2 for (int k = 0; k < 3; k++) {
        double ax[3] = {d.xmat[9 * b + k], d.xmat[9 * b + 3 + k], d.xmat[9 * b + 6 + k]};
        dof_motion(d.cdof + 6 * (da + k), ax, off);
      }
    } else if (jt == JNT_SLIDE) dof_motion(d.cdof + 6 * da, d.xaxis + 3 * j, nullptr);
    else dof_motion(d.cdof + 6 * da, d.xaxis + 3 * j, off);
  }
  L.sync();
}

// ---------------------------------------------------------------- tree-sparse L'DL
// in-place factorisation of the symmetric matrix in A (lower triangle, tree sparsity)
template <class T, class LT>
PS_NI void co_factor_tree(const LT& L, const DevModel<T>& m, double* A, double* dinv) { PS_ASSUME_SHARED(&m);
  constexpr int S = T::NV; const int nv = m.nv;
  PS_LOOP2 for (int k = nv - 1; k >= 0; k--) {
    double dk = A[k * S + k];
    if (dk < kMinVal) { dk = kMinVal; if (L.lane == 0) A[k * S + k] = kMinVal; }
    const int pa = m.fpair_adr[k], np = m.fpair_adr[k + 1] - pa;
    if (np == 0) continue;                    // dof without ancestors: nothing to eliminate
    if constexpr (LT::single) {
      // ONE working lane: the quotient A[k][i] / dk depends on i alone and the pair list is grouped by i (ps_host_model.h),
      // so it is formed once per ancestor instead of once per pair, and it IS the normalised entry the second loop below
      // would divide out again (row k is not touched by the pair updates): 99 FP64 divisions per factorisation of the
      // quadruped's tree instead of 474, the same values in the same places
      int last = -1; double tmp = 0;
      PS_LOOP2 for (int w = 0; w < np; w++) {
        const int i = m.fpair_i[pa + w], j = m.fpair_j[pa + w];
        if (i != last) { if (last >= 0) A[k * S + last] = tmp; tmp = A[k * S + i] / dk; last = i; }
        A[i * S + j] -= A[k * S + j] * tmp;
      }
      if (last >= 0) A[k * S + last] = tmp;
      continue;
    }
    CO_FOR(w, np) {
      const int i = m.fpair_i[pa + w], j = m.fpair_j[pa + w];
      const double tmp = A[k * S + i] / dk;
      A[i * S + j] -= A[k * S + j] * tmp;
    }
    L.sync();
    const int aa = m.dofanc_adr[k], na = m.dofanc_adr[k + 1] - aa;
    CO_FOR(w, na) { const int i = m.dofanc[aa + w]; A[k * S + i] = A[k * S + i] / dk; }
    L.sync();
  }
  L.sync();
  CO_FOR(i, nv) dinv[i] = 1.0 / A[i * S + i];
  L.sync();
}

// x <- (L' D L)^-1 x
template <class T, class LT>
PS_NI void co_solve_tree(const LT& L, const DevModel<T>& m, double* x, const double* A, const double* dinv) { PS_ASSUME_SHARED(&m);
  constexpr int S = T::NV; const int nv = m.nv;
  // x <- L^-T x : a dof collects from its descendants (deepest first, decreasing index)
  PS_LOOP2 for (int l = m.ndoflevel - 2; l >= 0; l--) {
    const int a = m.doflevel_adr[l], n = m.doflevel_adr[l + 1] - a;
    CO_FOR(w, n) {
      const int j = m.doflevel_dof[a + w];
      double s = x[j];
      PS_LOOP2 for (int c = m.dofdesc_adr[j]; c < m.dofdesc_adr[j + 1]; c++) { const int i = m.dofdesc[c]; s -= A[i * S + j] * x[i]; }
      x[j] = s;
    }
    L.sync();
  }
  CO_FOR(i, nv) x[i] *= dinv[i];
  L.sync();
  // x <- L^-1 x : a dof subtracts its ancestors (shallowest first)
  PS_LOOP2 for (int l = 1; l < m.ndoflevel; l++) {
    const int a = m.doflevel_adr[l], n = m.doflevel_adr[l + 1] - a;
    CO_FOR(w, n) {
      const int i = m.doflevel_dof[a + w];
      double s = x[i];
      PS_LOOP2 for (int c = m.dofanc_adr[i]; c < m.dofanc_adr[i + 1]; c++) { const int j = m.dofanc[c]; s -= A[i * S + j] * x[j]; }
      x[i] = s;
    }
    L.sync();
  }
}

// ---------------------------------------------------------------- composite inertia, mass matrix
template <class T, class LT>
PS_NI void co_mass_matrix(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  constexpr int S = T::NV; const int nb = m.nbody, nv = m.nv;
  double* crb = d.crb();
  CO_FOR(w, 10 * nb) crb[w] = d.cinert[w];
  CO_FOR(w, nv * S) { d.qM[w] = 0; }
  L.sync();
  co_fold_children(L, m, crb, 10, false);
  // f_i = crb(body_i) * cdof_i, kept in the scratch area
  double* f = d.scratch;
  CO_FOR(i, nv) inert_vec(f + 6 * i, crb + 10 * m.dof_bodyid[i], d.cdof + 6 * i);
  L.sync();
  CO_FOR(w, m.nmpair) {
    const int i = m.mpair_i[w], j = m.mpair_j[w];
    const double v = vdot(d.cdof + 6 * j, f + 6 * i, 6);
    if (j == i) d.qM[i * S + i] = m.dof_armature[i] + v; else { d.qM[i * S + j] = v; d.qM[j * S + i] = v; }
  }
  L.sync();
  CO_FOR(w, nv * S) d.qH[w] = d.qM[w];
  L.sync();
  co_factor_tree(L, m, d.qH, d.qLDiagInv);
}

// One working lane at the class capacity: co_mul_M can leave a copy of M's lower triangle in `Hcopy` while the rows are in
// registers, so that the Newton direction that follows does not load M again (co_newton_direction; every direction is
// preceded by an M v whose H is dead: the product before the solver's loop, or the line search's)
template <class T, class LT> PS_HD bool co_H_prefilled(const DevModel<T>& m) { return LT::single && m.nv == T::NV; }

template <class T, class LT>
PS_NI void co_mul_M(const LT& L, const DevModel<T>& m, const CoopData<T>& d, double* res, const double* vec, double* Hcopy = nullptr) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  if constexpr (LT::single) {
    // one working lane: the vector is loaded once and stays in registers (the dots are those of vdot, j ascending)
    constexpr int S = T::NV; const int nv = m.nv;
    if (nv == S) {          // (a model below the class capacity takes the plain loop)
      double v[S];
#pragma unroll
      for (int j = 0; j < S; j++) v[j] = vec[j];
      // two rows at a time: their 2 S loads are independent of each other and all in flight before the first product
      // (this kernel waits on its thread-private memory: ncu put 11 % of the stall samples on the row loads of one row at a time)
      PS_LOOP2 for (int i = 0; i + 1 < S; i += 2) {
        const double* row = d.qM + i * S;
        double r0[S], r1[S];
#pragma unroll
        for (int j = 0; j < S; j++) { r0[j] = row[j]; r1[j] = row[S + j]; }
        if (Hcopy) {
#pragma unroll
          for (int j = 0; j < S; j++) { if (j <= i) Hcopy[i * S + j] = r0[j]; if (j <= i + 1) Hcopy[(i + 1) * S + j] = r1[j]; }
        }
        double s0 = 0, s1 = 0;
#pragma unroll
        for (int j = 0; j < S; j++) { s0 += r0[j] * v[j]; s1 += r1[j] * v[j]; }
        res[i] = s0; res[i + 1] = s1;
      }
      if (S & 1) {
        const double* row = d.qM + (S - 1) * S;
        double s = 0;
#pragma unroll
        for (int j = 0; j < S; j++) { const double x = row[j]; if (Hcopy) Hcopy[(S - 1) * S + j] = x; s += x * v[j]; }
        res[S - 1] = s;
      }
      return;
    }
  }
  CO_FOR(i, m.nv) res[i] = vdot(d.qM + i * T::NV, vec, m.nv);
}

template <class T, class LT>
PS_NI void co_position_smooth(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  co_kinematics(L, m, d);
  CO_FOR(i, m.nu) d.actuator_length[i] = m.act_gear[i] * d.qpos[m.jnt_qposadr[m.act_jntid[i]]];
  L.sync();
  co_com_position(L, m, d);
  co_mass_matrix(L, m, d);
}

// ---------------------------------------------------------------- velocity stage
template <class T, class LT>
PS_NI void co_body_velocity(const DevModel<T>& m, const CoopData<T>& d, const double* qvel, double* cvel, double* cdof_dot, int i) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  double v[6]; vcopy(v, cvel + 6 * m.body_parentid[i], 6);
  const int ja = m.body_jntadr[i];
  PS_LOOP2 for (int j = ja; j < ja + m.body_jntnum[i]; j++) {
    int da = m.jnt_dofadr[j]; const int jt = m.jnt_type[j];
    if (jt == JNT_FREE) {
      PS_LOOP2 for (int k = 0; k < 3; k++) {
        vzero(cdof_dot + 6 * (da + k), 6);
        PS_LOOP2 for (int r = 0; r < 6; r++) v[r] += d.cdof[6 * (da + k) + r] * qvel[da + k];
      }
      da += 3;
    }
    if (jt == JNT_FREE || jt == JNT_BALL) {
      PS_LOOP2 for (int k = 0; k < 3; k++) cross_motion(cdof_dot + 6 * (da + k), v, d.cdof + 6 * (da + k));
      PS_LOOP2 for (int k = 0; k < 3; k++) PS_LOOP2 for (int r = 0; r < 6; r++) v[r] += d.cdof[6 * (da + k) + r] * qvel[da + k];
    } else {
      cross_motion(cdof_dot + 6 * da, v, d.cdof + 6 * da);
      PS_LOOP2 for (int r = 0; r < 6; r++) v[r] += d.cdof[6 * da + r] * qvel[da];
    }
  }
  vcopy(cvel + 6 * i, v, 6);
}

template <class T, class LT>
PS_NI void co_velocity_stage(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nv = m.nv, nb = m.nbody;
  double* cvel = d.cvel(); double* cdd = d.cdof_dot();
  CO_FOR(i, m.nu) d.actuator_velocity[i] = m.act_gear[i] * d.qvel[m.jnt_dofadr[m.act_jntid[i]]];
  if (L.lane == 0) vzero(cvel, 6);
  L.sync();
  PS_LOOP2 for (int l = 1; l < m.nlevel; l++) {
    const int a = m.level_adr[l], n = m.level_adr[l + 1] - a;
    CO_FOR(k, n) co_body_velocity<T, LT>(m, d, d.qvel, cvel, cdd, m.level_body[a + k]);
    L.sync();
  }
  // passive: joint springs + dof damping (the fluid model stays with the thread-per-rollout engine)
  CO_FOR(i, nv) {
    if (m.disableflags & DSBL_PASSIVE) { d.qfrc_passive[i] = 0; continue; }
    double spring = 0;
    const int j = m.dof_jntid[i]; const double k = m.jnt_stiffness[j];
    if (k != 0 && (m.jnt_type[j] == JNT_HINGE || m.jnt_type[j] == JNT_SLIDE))
      spring = -k * (d.qpos[m.jnt_qposadr[j]] - m.qpos_spring[m.jnt_qposadr[j]]);
    d.qfrc_passive[i] = spring + (-m.dof_damping[i] * d.qvel[i]) + 0.0;
  }
  // bias forces by RNE with zero acceleration
  double* cacc = d.scratch; double* cfrc = d.scratch + 6 * nb;
  if (L.lane == 0) {
    const bool grav = !(m.disableflags & DSBL_GRAVITY);
    PS_LOOP2 for (int k = 0; k < 3; k++) { cacc[k] = 0; cacc[3 + k] = grav ? -m.gravity[k] : 0.0; }
    vzero(cfrc, 6);
  }
  L.sync();
  PS_LOOP2 for (int l = 1; l < m.nlevel; l++) {
    const int a = m.level_adr[l], n = m.level_adr[l + 1] - a;
    CO_FOR(k, n) {
      const int i = m.level_body[a + k], bda = m.body_dofadr[i];
      double ca[6]; vcopy(ca, cacc + 6 * m.body_parentid[i], 6);
      PS_LOOP2 for (int j = 0; j < m.body_dofnum[i]; j++)
        PS_LOOP2 for (int r = 0; r < 6; r++) ca[r] += cdd[6 * (bda + j) + r] * d.qvel[bda + j];
      vcopy(cacc + 6 * i, ca, 6);
      double t1[6], t2[6], t3[6];
      inert_vec(t1, d.cinert + 10 * i, ca);
      inert_vec(t2, d.cinert + 10 * i, cvel + 6 * i);
      cross_force(t3, cvel + 6 * i, t2);
      PS_LOOP2 for (int r = 0; r < 6; r++) cfrc[6 * i + r] = t1[r] + t3[r];
    }
    L.sync();
  }
  co_fold_children(L, m, cfrc, 6, false);
  CO_FOR(i, nv) d.qfrc_bias[i] = vdot(d.cdof + 6 * i, cfrc + 6 * m.dof_bodyid[i], 6);
  L.sync();
}

// ---------------------------------------------------------------- subtree momentum (sensor support)
template <class T, class LT>
PS_NI void co_subtree_velocity(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nb = m.nbody;
  double* bodyvel = d.scratch; double* lv = d.subtree_linvel; double* am = d.subtree_angmom;
  CO_FOR(i, nb) {
    frame_velocity(m, d, i, d.xipos + 3 * i, bodyvel + 6 * i);
    PS_LOOP2 for (int k = 0; k < 3; k++) lv[3 * i + k] = m.body_mass[i] * bodyvel[6 * i + 3 + k];
    double t[3], t2[3]; matT_vec3(t, d.ximat + 9 * i, bodyvel + 6 * i);
    PS_LOOP2 for (int k = 0; k < 3; k++) t[k] *= m.body_inertia[3 * i + k];
    mat_vec3(t2, d.ximat + 9 * i, t); vcopy(am + 3 * i, t2, 3);
  }
  L.sync();
  co_fold_children(L, m, lv, 3, true);
  CO_FOR(i, nb) PS_LOOP2 for (int k = 0; k < 3; k++) lv[3 * i + k] /= fmax(kMinVal, m.body_subtreemass[i]);
  L.sync();
  // angular momentum: own offset term of a body, then fold it (and its transport term) into the parent
  PS_LOOP2 for (int l = m.nlevel - 1; l >= 1; l--) {
    const int a = m.level_adr[l], n = m.level_adr[l + 1] - a;
    CO_FOR(k, n) {
      const int i = m.level_body[a + k]; double dx[3], dv[3], dL[3];
      PS_LOOP2 for (int c = 0; c < 3; c++) { dx[c] = d.xipos[3 * i + c] - d.subtree_com[3 * i + c]; dv[c] = (bodyvel[6 * i + 3 + c] - lv[3 * i + c]) * m.body_mass[i]; }
      cross3(dL, dx, dv);
      PS_LOOP2 for (int c = 0; c < 3; c++) am[3 * i + c] += dL[c];
    }
    L.sync();
    const int pa = m.level_adr[l - 1], pn = m.level_adr[l] - pa;
    CO_FOR(k, pn) {
      const int p = m.level_body[pa + k];
      PS_LOOP2 for (int cc = m.child_adr[p]; cc < m.child_adr[p + 1]; cc++) {
        const int i = m.child_list[cc]; double dx[3], dv[3], dL[3];
        PS_LOOP2 for (int c = 0; c < 3; c++) am[3 * p + c] += am[3 * i + c];
        PS_LOOP2 for (int c = 0; c < 3; c++) { dx[c] = d.subtree_com[3 * i + c] - d.subtree_com[3 * p + c]; dv[c] = (lv[3 * i + c] - lv[3 * p + c]) * m.body_subtreemass[i]; }
        cross3(dL, dx, dv);
        PS_LOOP2 for (int c = 0; c < 3; c++) am[3 * p + c] += dL[c];
      }
    }
    L.sync();
  }
}

// ---------------------------------------------------------------- actuation + smooth acceleration
template <class T, class LT>
PS_NI void co_actuation_stage(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nv = m.nv, nu = m.nu;
  CO_FOR(i, nv) d.qfrc_actuator[i] = 0;
  CO_FOR(i, m.na) d.act_dot[i] = 0;
  bool bad = false;
  if (nu && !(m.disableflags & DSBL_ACTUATION)) CO_FOR(i, nu) bad = bad || bad_number(d.ctrl[i]);
  bad = L.any(bad);
  L.sync();
  if (bad) { d.warning |= WARN_BADCTRL; CO_FOR(i, nu) d.ctrl[i] = 0; L.sync(); }
  CO_FOR(i, nu) {
    if (m.disableflags & DSBL_ACTUATION) { d.actuator_force[i] = 0; continue; }
    double ctrl = d.ctrl[i];
    if (m.act_ctrllimited[i] && !(m.disableflags & DSBL_CLAMPCTRL))
      ctrl = clampd(ctrl, m.act_ctrlrange[2 * i], m.act_ctrlrange[2 * i + 1]);
    double in = ctrl;
    const int dyn = m.act_dyntype[i];
    if (dyn != DYN_NONE) {
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
  }
  L.sync();
  // several actuators may drive one dof: accumulate in actuator order, one lane per dof
  if (nu && !(m.disableflags & DSBL_ACTUATION)) CO_FOR(i, nv) {
    double s = 0;
    PS_LOOP2 for (int a = 0; a < nu; a++) if (m.jnt_dofadr[m.act_jntid[a]] == i) s += m.act_gear[a] * d.actuator_force[a];
    d.qfrc_actuator[i] = s;
  }
  L.sync();
}

template <class T, class LT>
PS_NI void co_acceleration_stage(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  CO_FOR(i, m.nv) {
    d.qfrc_smooth[i] = d.qfrc_passive[i] - d.qfrc_bias[i] + d.qfrc_actuator[i];
    d.qacc_smooth[i] = d.qfrc_smooth[i];
  }
  L.sync();
  co_solve_tree(L, m, d.qacc_smooth, d.qH, d.qLDiagInv);
}

// The same stage under Cartesian perturbation forces (mjData.xfrc_applied, [force; torque] per body; robust
// re-rolls only): one dof per lane, each adding J(xipos_b, b)' x of the bodies whose chain holds that dof, in
// body order -- the summation order of mj_xfrcAccumulate / the dense acceleration_stage_xfrc.
template <class T, class LT>
PS_NI void co_acceleration_stage_xfrc(const LT& L, const DevModel<T>& m, CoopData<T>& d, const double* xfrc_applied) {
  PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  CO_FOR(c, m.nv) {
    double acc = d.qfrc_passive[c] - d.qfrc_bias[c] + d.qfrc_actuator[c];
    const double* cd = d.cdof + 6 * c;
    PS_LOOP2 for (int b = 1; b < m.nbody; b++) {
      const double* x = xfrc_applied + 6 * b;
      if (x[0] == 0 && x[1] == 0 && x[2] == 0 && x[3] == 0 && x[4] == 0 && x[5] == 0) continue;
      const int w = m.body_weldid[b];
      if (w <= 0) continue;
      int i = m.body_dofadr[w] + m.body_dofnum[w] - 1;
      while (i > c) i = m.dof_parentid[i];          // a dof's parent has a lower index
      if (i != c) continue;                         // dof c does not move body b
      double off[3], t[3];
      const double* root = d.subtree_com + 3 * m.body_rootid[w];
      PS_LOOP2 for (int k = 0; k < 3; k++) off[k] = d.xipos[3 * b + k] - root[k];
      cross3(t, cd, off);
      PS_LOOP2 for (int k = 0; k < 3; k++) acc += (cd[3 + k] + t[k]) * x[k] + cd[k] * x[3 + k];
    }
    d.qfrc_smooth[c] = acc;
    d.qacc_smooth[c] = acc;
  }
  L.sync();
  co_solve_tree(L, m, d.qacc_smooth, d.qH, d.qLDiagInv);
}

}  // namespace ps
