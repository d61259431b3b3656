// Warp-cooperative contacts and constraint rows for one rollout (working set in shared memory).
// Rows keep mj_makeConstraint's order -- friction-loss dofs, joint limits, contacts -- but their
// Jacobians are never stored densely: friction / limit rows are unit rows (a dof and a sign), a
// contact's rows are compressed to the dofs of the two kinematic chains it touches.
#pragma once
#include "ps_coop_smooth.cuh"
#include "ps_constraint.cuh"

namespace ps {

// J_r . vec for constraint row r (summation in ascending dof order, as a dense dot would)
template <class T>
PS_HD double co_row_dot(const DevModel<T>& m, const CoopData<T>& d, int r, const double* vec) {
  const int tp = d.efc_type[r], id = d.efc_id[r];
  if (tp == EFC_FRICTION_DOF) return vec[id];
  if (tp == EFC_LIMIT_JOINT) return (double)d.efc_jrow[r] * vec[m.jnt_dofadr[id]];
  const double* J = d.jpool + d.efc_jrow[r] * T::NCH; const signed char* dof = d.con_dof + id * T::NCH;
  const int n = d.contact[id].ndof;
  double s = 0;
  PS_LOOP2 for (int k = 0; k < n; k++) s += J[k] * vec[dof[k]];
  return s;
}

// ---------------------------------------------------------------- narrow phase over the static pair list
template <class T, class LT>
PS_NI void co_detect_contacts(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  int ncon = 0;
  if (!(m.disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT))) {
    CO_ROUNDS(base, m.npair) {
      const int p = base + L.lane;
      RawContact rc[4]; int n = 0, g1 = 0, g2 = 0; double margin = 0, gap = 0;
      if (p < m.npair) {
        g1 = m.pair_geom1[p]; g2 = m.pair_geom2[p];                   // type(g1) <= type(g2)
        const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
        margin = fmax(m.geom_margin[g1], m.geom_margin[g2]); gap = fmax(m.geom_gap[g1], m.geom_gap[g2]);
        double p1[3], m1[9], p2[3], m2[9];
        geom_pose(m, d, g1, p1, m1); geom_pose(m, d, g2, p2, (double*)nullptr);
        const double* s1 = m.geom_size + 3 * g1; const double* s2 = m.geom_size + 3 * g2;
        const double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
        bool near;
        if (t1 == GEOM_PLANE) { const double nrm[3] = {m1[2], m1[5], m1[8]}; near = !(dot3(dif, nrm) > margin + m.geom_rbound[g2]); }
        else { const double bound = m.geom_rbound[g1] + m.geom_rbound[g2] + margin; near = !(dot3(dif, dif) > bound * bound); }
        if (near) {
          geom_pose(m, d, g2, p2, m2);
          if (t1 == GEOM_PLANE) {
            if (t2 == GEOM_SPHERE) n = hit_plane_sphere(rc, p1, m1, p2, s2[0], margin);
            else if (t2 == GEOM_CAPSULE) n = hit_plane_capsule(rc, p1, m1, p2, m2, s2, margin);
            else if (t2 == GEOM_CYLINDER) n = hit_plane_cylinder(rc, p1, m1, p2, m2, s2, margin);
            else if (t2 == GEOM_BOX) n = hit_plane_box(rc, p1, m1, p2, m2, s2, margin);
          } else if (t1 == GEOM_SPHERE) {
            if (t2 == GEOM_SPHERE) n = hit_sphere_sphere(rc, p1, s1[0], p2, s2[0], margin);
            else if (t2 == GEOM_CAPSULE) n = hit_sphere_capsule(rc, p1, s1[0], p2, m2, s2, margin);
          } else if (t1 == GEOM_CAPSULE && t2 == GEOM_CAPSULE) n = hit_capsule_capsule(rc, p1, m1, s1, p2, m2, s2, margin);
        }
      }
      int total; const int off = L.scan(n, &total);
      PS_LOOP2 for (int k = 0; k < n; k++) {
        const int idx = ncon + off + k;
        if (idx >= T::MAXCON) break;
        CoopContact& con = d.contact[idx];
        con.dist = rc[k].dist; vcopy(con.pos, rc[k].pos, 3); vcopy(con.frame, rc[k].frame, 6);
        complete_frame(con.frame);
        con.includemargin = margin - gap; con.geom1 = (short)g1; con.geom2 = (short)g2; con.efc_address = -1; con.mu = 0;
        con.ndof = 0; con.jadr = 0; con.nrow = 0;
        mix_contact_params(m, g1, g2, &con);
      }
      if (ncon + total > T::MAXCON) { ncon = T::MAXCON; if (L.lane == 0) d.warning |= WARN_CAPACITY; break; }
      ncon += total;
    }
  }
  if (L.lane == 0) d.ncon = ncon;
  L.sync();
}

// impedance / reference parameters of one row; R is left un-adjusted (cone scaling follows)
PS_NI void co_row_params(double timestep, int disableflags, const double* sr, const double* si, double pos, double margin,
                         double diag, bool zero_stiffness, double* R, double* B, double* c) {
  double solref[2], solimp[5];
  sanitize_sol(timestep, disableflags, sr, si, solref, solimp);
  const double imp = impedance_at(solimp, pos, margin);
  const double dmax = solimp[1]; double K, Bv;
  if (solref[0] > 0) {
    K = 1 / fmax(kMinVal, dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]);
    Bv = 2 / fmax(kMinVal, dmax * solref[0]);
  } else { K = -solref[0] / fmax(kMinVal, dmax * dmax); Bv = -solref[1] / fmax(kMinVal, dmax); }
  if (zero_stiffness) K = 0;
  *B = Bv; *c = K * imp * (pos - margin);
  *R = fmax(kMinVal, (1 - imp) * diag / imp);
}

// rows of contact ci: compressed Jacobian, impedance, cone scaling.  Runs on ONE lane.
template <class T, class LT>
PS_NI void co_contact_rows(const DevModel<T>& m, CoopData<T>& d, int ci) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  constexpr int NCH = T::NCH;
  CoopContact& con = d.contact[ci];
  const int dim = con.dim, adr = con.efc_address;
  const int b1 = m.geom_bodyid[con.geom1], b2 = m.geom_bodyid[con.geom2];
  const int w1 = m.body_weldid[b1], w2 = m.body_weldid[b2];
  signed char* dof = d.con_dof + ci * NCH; signed char* slot = d.con_slot + ci * T::NV;
  // union of the two kinematic chains, ascending
  int n = 0;
  {
    int i1 = w1 > 0 ? m.body_dofadr[w1] + m.body_dofnum[w1] - 1 : -1;
    int i2 = w2 > 0 ? m.body_dofadr[w2] + m.body_dofnum[w2] - 1 : -1;
    signed char tmp[T::NV];
    while (i1 >= 0 || i2 >= 0) {
      const int i = i1 > i2 ? i1 : i2;
      tmp[n++] = (signed char)i;
      if (i1 == i) i1 = m.dof_parentid[i1];
      if (i2 == i) i2 = m.dof_parentid[i2];
    }
    if (n > NCH) { d.warning |= WARN_CAPACITY; n = NCH; }       // cannot happen when NCH covers two chains
    PS_LOOP2 for (int i = 0; i < m.nv; i++) slot[i] = -1;
    PS_LOOP2 for (int k = 0; k < n; k++) { dof[k] = tmp[n - 1 - k]; slot[dof[k]] = (signed char)k; }
  }
  con.ndof = (short)n;
  double off1[3], off2[3];
  PS_LOOP2 for (int k = 0; k < 3; k++) {
    off1[k] = w1 > 0 ? con.pos[k] - d.subtree_com[3 * m.body_rootid[w1] + k] : 0.0;
    off2[k] = w2 > 0 ? con.pos[k] - d.subtree_com[3 * m.body_rootid[w2] + k] : 0.0;
  }
  // membership of each dof in the two chains
  double Jc[6 * NCH];
  {
    int i1 = w1 > 0 ? m.body_dofadr[w1] + m.body_dofnum[w1] - 1 : -1;
    int i2 = w2 > 0 ? m.body_dofadr[w2] + m.body_dofnum[w2] - 1 : -1;
    PS_LOOP2 for (int k = n - 1; k >= 0; k--) {
      const int i = dof[k]; const double* c = d.cdof + 6 * i;
      double a[3] = {0, 0, 0}, b[3] = {0, 0, 0}, ar[3] = {0, 0, 0}, br[3] = {0, 0, 0};
      if (i1 == i) { double t[3]; cross3(t, c, off1); PS_LOOP2 for (int x = 0; x < 3; x++) { a[x] = c[3 + x] + t[x]; ar[x] = c[x]; } i1 = m.dof_parentid[i1]; }
      if (i2 == i) { double t[3]; cross3(t, c, off2); PS_LOOP2 for (int x = 0; x < 3; x++) { b[x] = c[3 + x] + t[x]; br[x] = c[x]; } i2 = m.dof_parentid[i2]; }
      PS_LOOP2 for (int r = 0; r < dim; r++) {
        const double* f = con.frame + 3 * (r % 3);
        Jc[r * NCH + k] = r < 3 ? f[0] * (b[0] - a[0]) + f[1] * (b[1] - a[1]) + f[2] * (b[2] - a[2])
                                : f[0] * (br[0] - ar[0]) + f[1] * (br[1] - ar[1]) + f[2] * (br[2] - ar[2]);
      }
    }
  }
  const double tran = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
  const double rot = m.body_invweight0[2 * b1 + 1] + m.body_invweight0[2 * b2 + 1];
  double* R = d.efc_R + adr;
  const int nrow = con.nrow;
  PS_LOOP2 for (int q = 0; q < nrow; q++) {
    const int row = adr + q;
    double* J = d.jpool + (con.jadr + q) * NCH;
    double pos = con.dist, margin = con.includemargin, diag; bool zk = false; int type;
    if (dim == 1) { type = EFC_CONTACT_FRICTIONLESS; diag = tran; PS_LOOP2 for (int k = 0; k < n; k++) J[k] = Jc[k]; }
    else if (m.cone == CONE_ELLIPTIC) {
      type = EFC_CONTACT_ELLIPTIC; diag = q < 3 ? tran : rot;
      if (q > 0) { pos = 0; margin = 0; zk = true; }
      PS_LOOP2 for (int k = 0; k < n; k++) J[k] = Jc[q * NCH + k];
    } else {
      type = EFC_CONTACT_PYRAMIDAL;
      const int r = 1 + q / 2; const double sg = (q & 1) ? -1.0 : 1.0, mu = con.friction[r - 1];
      diag = tran + mu * mu * (r - 1 < 2 ? tran : rot);
      PS_LOOP2 for (int k = 0; k < n; k++) J[k] = Jc[k] + sg * mu * Jc[r * NCH + k];
    }
    d.efc_type[row] = (unsigned char)type; d.efc_id[row] = (short)ci; d.efc_jrow[row] = (short)(con.jadr + q); d.efc_state[row] = 0;
    co_row_params(m.timestep, m.disableflags, con.solref, con.solimp, pos, margin, diag, zk, R + q, d.efc_force + row, d.efc_aref + row);
  }
  if (dim > 1) {
    if (m.cone == CONE_ELLIPTIC) {
      R[1] = R[0] / fmax(kMinVal, m.impratio);
      con.mu = con.friction[0] * sqrt(R[1] / R[0]);
      PS_LOOP2 for (int j = 1; j < dim - 1; j++) R[j + 1] = R[1] * con.friction[0] * con.friction[0] / (con.friction[j] * con.friction[j]);
    } else {
      const double R1 = R[0] / fmax(kMinVal, m.impratio);
      con.mu = con.friction[0] * sqrt(R1 / R[0]);
      const double Rpy = 2 * con.mu * con.mu * R[0];
      PS_LOOP2 for (int j = 0; j < 2 * (dim - 1); j++) R[j] = Rpy;
    }
  }
  PS_LOOP2 for (int q = 0; q < nrow; q++) d.efc_D[adr + q] = 1 / R[q];
}

template <class T, class LT>
PS_NI void co_make_constraints(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  if (m.disableflags & DSBL_CONSTRAINT) { if (L.lane == 0) { d.nefc = 0; d.ncon = 0; d.nlimit = 0; } L.sync(); return; }
  co_detect_contacts(L, m, d);
  const int ncon = d.ncon;
  // friction-loss rows: a fixed list
  const int nf = (m.disableflags & DSBL_FRICTIONLOSS) ? 0 : m.nfrictionrow;
  CO_FOR(k, nf) {
    const int i = m.frictionrow_dof[k];
    d.efc_type[k] = EFC_FRICTION_DOF; d.efc_id[k] = (short)i; d.efc_jrow[k] = 1; d.efc_state[k] = 0;
    co_row_params(m.timestep, m.disableflags, m.dof_solref + 2 * i, m.dof_solimp + 5 * i, 0, 0, m.dof_invweight0[i], true,
                  d.efc_R + k, d.efc_force + k, d.efc_aref + k);
    d.efc_D[k] = 1 / d.efc_R[k];
  }
  // joint limits, joint order, lower side first
  int nefc = nf;
  bool overflow = false;
  if (!(m.disableflags & DSBL_LIMIT)) {
    CO_ROUNDS(base, m.njnt) {
      const int j = base + L.lane; int cnt = 0; double dist[2]; int sgn[2];
      if (j < m.njnt && m.jnt_limited[j] && (m.jnt_type[j] == JNT_HINGE || m.jnt_type[j] == JNT_SLIDE)) {
        const double value = d.qpos[m.jnt_qposadr[j]], margin = m.jnt_margin[j];
        PS_LOOP2 for (int side = -1; side <= 1; side += 2) {
          const double ds = side * (m.jnt_range[2 * j + (side + 1) / 2] - value);
          if (ds < margin) { dist[cnt] = ds; sgn[cnt] = -side; cnt++; }
        }
      }
      int total; const int off = L.scan(cnt, &total);
      PS_LOOP2 for (int c = 0; c < cnt; c++) {
        const int row = nefc + off + c;
        if (row >= T::MAXEFC) break;
        d.efc_type[row] = EFC_LIMIT_JOINT; d.efc_id[row] = (short)j; d.efc_jrow[row] = (short)sgn[c]; d.efc_state[row] = 0;
        co_row_params(m.timestep, m.disableflags, m.jnt_solref + 2 * j, m.jnt_solimp + 5 * j, dist[c], m.jnt_margin[j],
                      m.dof_invweight0[m.jnt_dofadr[j]], false, d.efc_R + row, d.efc_force + row, d.efc_aref + row);
        d.efc_D[row] = 1 / d.efc_R[row];
      }
      nefc += total;
      if (nefc > T::MAXEFC) { nefc = T::MAXEFC; overflow = true; break; }
    }
  }
  const int nlimit = nefc - nf;
  // contact rows: the first contact that does not fit ends the list
  int jrows = 0;
  if (!overflow) {
    CO_ROUNDS(base, ncon) {
      const int ci = base + L.lane; int nr = 0;
      if (ci < ncon) {
        CoopContact& con = d.contact[ci];
        con.efc_address = -1;
        if (con.dist < con.includemargin) nr = con.dim == 1 ? 1 : (m.cone == CONE_ELLIPTIC ? con.dim : 2 * (con.dim - 1));
      }
      int total; const int off = L.scan(nr, &total);
      const bool fits = nefc + off + nr <= T::MAXEFC && jrows + off + nr <= T::NJROW;
      const int firstbad = L.min((nr > 0 && !fits) ? ci : 0x7fffffff);
      if (nr > 0 && ci < firstbad) {
        CoopContact& con = d.contact[ci];
        con.efc_address = (short)(nefc + off); con.jadr = (short)(jrows + off); con.nrow = (short)nr;
        co_contact_rows<T, LT>(m, d, ci);
      }
      if (firstbad != 0x7fffffff) { nefc += L.bcast(off, firstbad - base); overflow = true; break; }
      nefc += total; jrows += total;
    }
  }
  if (L.lane == 0) { d.nefc = nefc; d.nlimit = nlimit; if (overflow) d.warning |= WARN_CAPACITY; }
  L.sync();
}

// aref = -B (J v) - K imp (pos - margin); B and the position term were parked in efc_force / efc_aref
template <class T, class LT>
PS_NI void co_reference_acceleration(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  CO_FOR(i, d.nefc) {
    const double vel = co_row_dot(m, d, i, d.qvel);
    d.efc_aref[i] = -d.efc_force[i] * vel - d.efc_aref[i];
  }
  L.sync();
}

}  // namespace ps
