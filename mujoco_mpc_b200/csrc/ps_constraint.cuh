// Contacts and soft constraints of one rollout: primitive narrow-phase over the statically
// filtered geom pairs, friction-loss / joint-limit / contact rows with solref-solimp impedance,
// and the primal Newton solver with exact line search on the convex constraint cost.
//
// Device replacement for MuJoCo 3.1.4's mj_collision, mj_makeConstraint, mj_referenceConstraint
// and mj_fwdConstraint (Newton), which the reference's rollouts reach through mj_step
// (mjpc/trajectory.cc:158).
#pragma once
#include "ps_smooth.cuh"

namespace ps {

struct RawContact { double dist, pos[3], frame[6]; };

PS_HD int hit_plane_sphere(RawContact* c, const double* pp, const double* pm, const double* sp, double r, double margin) {
  const double n[3] = {pm[2], pm[5], pm[8]};
  const double dif[3] = {sp[0] - pp[0], sp[1] - pp[1], sp[2] - pp[2]};
  const double dist = dot3(dif, n) - r;
  if (dist > margin) return 0;
  c->dist = dist;
  for (int k = 0; k < 3; k++) { c->pos[k] = sp[k] + n[k] * (-dist * 0.5 - r); c->frame[k] = n[k]; c->frame[3 + k] = 0; }
  return 1;
}

PS_NT int hit_plane_capsule(RawContact* c, const double* pp, const double* pm, const double* cp, const double* cm,
                            const double* size, double margin) {
  const double axis[3] = {cm[2], cm[5], cm[8]};
  int cnt = 0;
  for (int s = 0; s < 2; s++) {
    const double sgn = s == 0 ? 1.0 : -1.0; double e[3];
    for (int k = 0; k < 3; k++) e[k] = cp[k] + sgn * axis[k] * size[1];
    if (hit_plane_sphere(c + cnt, pp, pm, e, size[0], margin)) { vcopy(c[cnt].frame + 3, axis, 3); cnt++; }
  }
  return cnt;
}

PS_NT int hit_plane_cylinder(RawContact* c, const double* pp, const double* pm, const double* cp, const double* cm,
                             const double* size, double margin) {
  const double n[3] = {pm[2], pm[5], pm[8]}; double axis[3] = {cm[2], cm[5], cm[8]};
  const double r = size[0], h = size[1];
  double prjaxis = dot3(n, axis);
  if (prjaxis > 0) { for (int k = 0; k < 3; k++) axis[k] = -axis[k]; prjaxis = -prjaxis; }
  const double dif[3] = {cp[0] - pp[0], cp[1] - pp[1], cp[2] - pp[2]};
  const double dist0 = dot3(dif, n);
  double vec[3]; for (int k = 0; k < 3; k++) vec[k] = axis[k] * prjaxis - n[k];
  const double len2 = dot3(vec, vec);
  if (len2 >= kMinVal * kMinVal) { const double s = r / sqrt(len2); for (int k = 0; k < 3; k++) vec[k] *= s; }
  else { vec[0] = cm[0] * r; vec[1] = cm[3] * r; vec[2] = cm[6] * r; }
  const double prjvec = dot3(vec, n);
  for (int k = 0; k < 3; k++) axis[k] *= h;
  prjaxis *= h;
  int cnt = 0;
  if (dist0 + prjaxis + prjvec <= margin) {
    const double dist = dist0 + prjaxis + prjvec; c[cnt].dist = dist;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = cp[k] + vec[k] + axis[k] - n[k] * dist * 0.5; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    cnt++;
  } else return 0;
  if (dist0 - prjaxis + prjvec <= margin) {
    const double dist = dist0 - prjaxis + prjvec; c[cnt].dist = dist;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = cp[k] + vec[k] - axis[k] - n[k] * dist * 0.5; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    cnt++;
  }
  const double prjvec1 = -prjvec * 0.5;
  if (dist0 + prjaxis + prjvec1 <= margin) {
    double vec1[3]; cross3(vec1, vec, axis); unit3(vec1);
    const double s = r * sqrt(3.0) * 0.5; for (int k = 0; k < 3; k++) vec1[k] *= s;
    const double dist = dist0 + prjaxis + prjvec1;
    for (int q = 0; q < 2; q++) {
      const double sg = q == 0 ? 1.0 : -1.0; c[cnt].dist = dist;
      for (int k = 0; k < 3; k++) {
        c[cnt].pos[k] = cp[k] + sg * vec1[k] + axis[k] - vec[k] * 0.5 - n[k] * dist * 0.5;
        c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0;
      }
      cnt++;
    }
  }
  return cnt;
}

PS_NT int hit_plane_box(RawContact* c, const double* pp, const double* pm, const double* bp, const double* bm,
                        const double* size, double margin) {
  const double n[3] = {pm[2], pm[5], pm[8]};
  const double dif[3] = {bp[0] - pp[0], bp[1] - pp[1], bp[2] - pp[2]};
  const double dist = dot3(dif, n);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    const double v[3] = {(i & 1 ? size[0] : -size[0]), (i & 2 ? size[1] : -size[1]), (i & 4 ? size[2] : -size[2])};
    double corner[3]; mat_vec3(corner, bm, v);
    const double ld = dot3(n, corner);
    if (dist + ld > margin || ld > 0) continue;
    c[cnt].dist = dist + ld;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = bp[k] + corner[k] - n[k] * c[cnt].dist * 0.5; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    if (++cnt >= 4) break;
  }
  return cnt;
}

PS_HD int hit_sphere_sphere(RawContact* c, const double* p1, double r1, const double* p2, double r2, double margin) {
  const double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  const double cd = norm3(dif);
  const double dist = cd - r1 - r2;
  if (dist > margin) return 0;
  double n[3];
  if (cd < kMinVal) { n[0] = 1; n[1] = 0; n[2] = 0; } else for (int k = 0; k < 3; k++) n[k] = dif[k] / cd;
  c->dist = dist;
  for (int k = 0; k < 3; k++) { c->pos[k] = p1[k] + n[k] * (r1 + 0.5 * dist); c->frame[k] = n[k]; c->frame[3 + k] = 0; }
  return 1;
}

PS_NT int hit_sphere_capsule(RawContact* c, const double* p1, double r1, const double* p2, const double* m2,
                             const double* size2, double margin) {
  const double axis[3] = {m2[2], m2[5], m2[8]};
  const double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  const double x = clampd(dot3(axis, dif), -size2[1], size2[1]);
  const double q[3] = {p2[0] + axis[0] * x, p2[1] + axis[1] * x, p2[2] + axis[2] * x};
  return hit_sphere_sphere(c, p1, r1, q, size2[0], margin);
}

PS_NT int hit_capsule_capsule(RawContact* c, const double* p1, const double* m1, const double* s1,
                              const double* p2, const double* m2, const double* s2, double margin) {
  const double a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]};
  const double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  const double ma = dot3(a1, a1), mb = -dot3(a1, a2), mc = dot3(a2, a2);
  const double u = -dot3(a1, dif), v = dot3(a2, dif);
  const double det = ma * mc - mb * mb;
  if (fabs(det) >= kMinVal) {
    double x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
    if (x1 > s1[1]) { x1 = s1[1]; x2 = (v - mb * s1[1]) / mc; }
    else if (x1 < -s1[1]) { x1 = -s1[1]; x2 = (v + mb * s1[1]) / mc; }
    if (x2 > s2[1]) { x2 = s2[1]; x1 = clampd((u - mb * s2[1]) / ma, -s1[1], s1[1]); }
    else if (x2 < -s2[1]) { x2 = -s2[1]; x1 = clampd((u + mb * s2[1]) / ma, -s1[1], s1[1]); }
    double q1[3], q2[3];
    for (int k = 0; k < 3; k++) { q1[k] = p1[k] + a1[k] * x1; q2[k] = p2[k] + a2[k] * x2; }
    return hit_sphere_sphere(c, q1, s1[0], q2, s2[0], margin);
  }
  int cnt = 0;
  for (int e = 0; e < 2; e++) {
    const double sg = e == 0 ? 1.0 : -1.0; double q1[3], q2[3];
    for (int k = 0; k < 3; k++) q1[k] = p1[k] + sg * a1[k] * s1[1];
    const double d2[3] = {q1[0] - p2[0], q1[1] - p2[1], q1[2] - p2[2]};
    const double x2 = clampd(dot3(a2, d2), -s2[1], s2[1]);
    for (int k = 0; k < 3; k++) q2[k] = p2[k] + a2[k] * x2;
    cnt += hit_sphere_sphere(c + cnt, q1, s1[0], q2, s2[0], margin);
    if (cnt >= 2) return cnt;
  }
  return cnt;
}

// complete an orthonormal contact frame from the normal (+ optional tangent hint)
PS_HD void complete_frame(double* f) {
  unit3(f);
  if (norm3(f + 3) < 0.5) { f[3] = 0; f[4] = 0; f[5] = 0; if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1; }
  const double dp = dot3(f, f + 3);
  for (int k = 0; k < 3; k++) f[3 + k] -= f[k] * dp;
  unit3(f + 3);
  cross3(f + 6, f, f + 3);
}

template <class T, class C>
PS_NT void mix_contact_params(const DevModel<T>& m, int g1, int g2, C* con) {
  const int p1 = m.geom_priority[g1], p2 = m.geom_priority[g2];
  double fri[3];
  if (p1 != p2) {
    const int g = p1 > p2 ? g1 : g2;
    con->dim = m.geom_condim[g];
    vcopy(con->solref, m.geom_solref + 2 * g, 2); vcopy(con->solimp, m.geom_solimp + 5 * g, 5);
    vcopy(fri, m.geom_friction + 3 * g, 3);
  } else {
    con->dim = m.geom_condim[g1] > m.geom_condim[g2] ? m.geom_condim[g1] : m.geom_condim[g2];
    const double s1 = m.geom_solmix[g1], s2 = m.geom_solmix[g2]; double mix;
    if (s1 >= kMinVal && s2 >= kMinVal) mix = s1 / (s1 + s2);
    else if (s1 < kMinVal && s2 < kMinVal) mix = 0.5;
    else mix = s1 < kMinVal ? 0.0 : 1.0;
    const double* r1 = m.geom_solref + 2 * g1; const double* r2 = m.geom_solref + 2 * g2;
    if (r1[0] > 0 && r2[0] > 0) for (int k = 0; k < 2; k++) con->solref[k] = mix * r1[k] + (1 - mix) * r2[k];
    else for (int k = 0; k < 2; k++) con->solref[k] = fmin(r1[k], r2[k]);
    for (int k = 0; k < 5; k++) con->solimp[k] = mix * m.geom_solimp[5 * g1 + k] + (1 - mix) * m.geom_solimp[5 * g2 + k];
    for (int k = 0; k < 3; k++) fri[k] = fmax(m.geom_friction[3 * g1 + k], m.geom_friction[3 * g2 + k]);
  }
  con->friction[0] = con->friction[1] = fri[0]; con->friction[2] = fri[1];
  con->friction[3] = con->friction[4] = fri[2];
}

template <class T>
PS_NT void detect_contacts(const DevModel<T>& m, RolloutData<T>& d) {
  d.ncon = 0;
  if (m.disableflags & (DSBL_CONTACT | DSBL_CONSTRAINT)) return;
  for (int p = 0; p < m.npair; p++) {
    const int g1 = m.pair_geom1[p], g2 = m.pair_geom2[p];   // type(g1) <= type(g2)
    const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
    const double margin = fmax(m.geom_margin[g1], m.geom_margin[g2]);
    const double gap = fmax(m.geom_gap[g1], m.geom_gap[g2]);
    const double* p1 = d.geom_xpos + 3 * g1; const double* m1 = d.geom_xmat + 9 * g1;
    const double* p2 = d.geom_xpos + 3 * g2; const double* m2 = d.geom_xmat + 9 * g2;
    const double* s1 = m.geom_size + 3 * g1; const double* s2 = m.geom_size + 3 * g2;
    const double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    if (t1 == GEOM_PLANE) {
      const double n[3] = {m1[2], m1[5], m1[8]};
      if (dot3(dif, n) > margin + m.geom_rbound[g2]) continue;
    } else {
      const double bound = m.geom_rbound[g1] + m.geom_rbound[g2] + margin;
      if (dot3(dif, dif) > bound * bound) continue;
    }
    RawContact rc[4]; int n = 0;
    if (t1 == GEOM_PLANE) {
      if (t2 == GEOM_SPHERE) n = hit_plane_sphere(rc, p1, m1, p2, s2[0], margin);
      else if (t2 == GEOM_CAPSULE) n = hit_plane_capsule(rc, p1, m1, p2, m2, s2, margin);
      else if (t2 == GEOM_CYLINDER) n = hit_plane_cylinder(rc, p1, m1, p2, m2, s2, margin);
      else if (t2 == GEOM_BOX) n = hit_plane_box(rc, p1, m1, p2, m2, s2, margin);
    } else if (t1 == GEOM_SPHERE) {
      if (t2 == GEOM_SPHERE) n = hit_sphere_sphere(rc, p1, s1[0], p2, s2[0], margin);
      else if (t2 == GEOM_CAPSULE) n = hit_sphere_capsule(rc, p1, s1[0], p2, m2, s2, margin);
    } else if (t1 == GEOM_CAPSULE && t2 == GEOM_CAPSULE) n = hit_capsule_capsule(rc, p1, m1, s1, p2, m2, s2, margin);
    for (int k = 0; k < n; k++) {
      if (d.ncon >= T::MAXCON) { d.warning |= WARN_CAPACITY; return; }
      Contact& con = d.contact[d.ncon++];
      con.dist = rc[k].dist; vcopy(con.pos, rc[k].pos, 3); vcopy(con.frame, rc[k].frame, 6);
      complete_frame(con.frame);
      con.includemargin = margin - gap; con.geom1 = g1; con.geom2 = g2; con.efc_address = -1; con.mu = 0;
      mix_contact_params(m, g1, g2, &con);
    }
  }
}

// ---------------------------------------------------------------- constraint rows
PS_HD void sanitize_sol(double timestep, int disableflags, const double* sr, const double* si, double* solref, double* solimp) {
  vcopy(solref, sr, 2); vcopy(solimp, si, 5);
  if (!(disableflags & DSBL_REFSAFE) && solref[0] > 0) solref[0] = fmax(solref[0], 2 * timestep);
  solimp[0] = clampd(solimp[0], 0.0001, 0.9999); solimp[1] = clampd(solimp[1], 0.0001, 0.9999);
  solimp[2] = fmax(0.0, solimp[2]); solimp[3] = clampd(solimp[3], 0.0001, 0.9999);
  solimp[4] = fmax(1.0, solimp[4]);
}

// x^p; the integer powers models actually use (solimp power 1..3) are exact IEEE products
PS_HD double pow_small(double x, double p) {
  if (p == 1) return x;
  if (p == 2) return x * x;
  if (p == 3) return x * x * x;
  return pow(x, p);
}

PS_NT double impedance_at(const double* si, double pos, double margin) {
  if (si[0] == si[1] || si[2] <= kMinVal) return 0.5 * (si[0] + si[1]);
  const double x = fabs((pos - margin) / si[2]);
  if (x >= 1) return si[1];
  if (x <= 0) return si[0];
  double y;
  if (si[4] == 1) y = x;
  else if (x <= si[3]) y = pow_small(x, si[4]) / pow_small(si[3], si[4] - 1);
  else y = 1 - pow_small(1 - x, si[4]) / pow_small(1 - si[3], si[4] - 1);
  return si[0] + y * (si[1] - si[0]);
}

template <class T>
PS_HD bool push_row(RolloutData<T>& d, double pos, double margin, double floss, int type, int id) {
  if (d.nefc >= T::MAXEFC) { d.warning |= WARN_CAPACITY; return false; }
  const int r = d.nefc++;
  d.efc_pos[r] = pos; d.efc_margin[r] = margin; d.efc_frictionloss[r] = floss; d.efc_type[r] = type; d.efc_id[r] = id;
  return true;
}

template <class T>
PS_NT void make_constraints(const DevModel<T>& m, RolloutData<T>& d) {
  constexpr int S = T::NV; const int nv = nv_of(m);
  d.nefc = 0; d.ncon = 0;
  if (m.disableflags & DSBL_CONSTRAINT) return;
  if (T::CONTACTS) detect_contacts(m, d);
  if (!(m.disableflags & DSBL_FRICTIONLOSS))
    PS_SPEC_UNROLL for (int i = 0; i < nv; i++) if (m.dof_frictionloss[i] > 0) {
      const int r = d.nefc;
      if (!push_row(d, 0, 0, m.dof_frictionloss[i], EFC_FRICTION_DOF, i)) break;
      vzero(d.efc_J + r * S, nv); d.efc_J[r * S + i] = 1;
    }
  if (!(m.disableflags & DSBL_LIMIT))
    PS_SPEC_UNROLL for (int j = 0; j < m.njnt; j++) {
      if (!m.jnt_limited[j]) continue;
      const int jt = m.jnt_type[j];
      if (jt != JNT_HINGE && jt != JNT_SLIDE) continue;
      const double value = d.qpos[m.jnt_qposadr[j]], margin = m.jnt_margin[j];
      PS_SPEC_UNROLL for (int side = -1; side <= 1; side += 2) {
        const double dist = side * (m.jnt_range[2 * j + (side + 1) / 2] - value);
        if (dist < margin) {
          const int r = d.nefc;
          if (!push_row(d, dist, margin, 0, EFC_LIMIT_JOINT, j)) break;
          vzero(d.efc_J + r * S, nv); d.efc_J[r * S + m.jnt_dofadr[j]] = -(double)side;
        }
      }
    }
  // contacts: Jacobian difference of the two bodies at the contact point, rotated into the frame
  if (T::CONTACTS) {
    double jp1[3 * S], jp2[3 * S], jr1[3 * S], jr2[3 * S], Jc[6 * S];
    for (int ci = 0; ci < d.ncon; ci++) {
      Contact& con = d.contact[ci];
      con.efc_address = -1;
      if (con.dist >= con.includemargin) continue;
      const int dim = con.dim;
      const bool needrot = dim > 3;
      point_jacobian(m, d, jp1, needrot ? jr1 : nullptr, con.pos, m.geom_bodyid[con.geom1]);
      point_jacobian(m, d, jp2, needrot ? jr2 : nullptr, con.pos, m.geom_bodyid[con.geom2]);
      for (int r = 0; r < dim; r++) {
        const double* f = con.frame + 3 * (r % 3);
        const double* a = r < 3 ? jp1 : jr1; const double* b = r < 3 ? jp2 : jr2;
        for (int i = 0; i < nv; i++)
          Jc[r * S + i] = f[0] * (b[i] - a[i]) + f[1] * (b[S + i] - a[S + i]) + f[2] * (b[2 * S + i] - a[2 * S + i]);
      }
      const int adr = d.nefc;
      bool ok = true;
      if (dim == 1) {
        ok = push_row(d, con.dist, con.includemargin, 0, EFC_CONTACT_FRICTIONLESS, ci);
        if (ok) vcopy(d.efc_J + adr * S, Jc, nv);
      } else if (m.cone == CONE_ELLIPTIC) {
        for (int r = 0; r < dim && ok; r++) {
          const int row = d.nefc;
          ok = push_row(d, r == 0 ? con.dist : 0, r == 0 ? con.includemargin : 0, 0, EFC_CONTACT_ELLIPTIC, ci);
          if (ok) vcopy(d.efc_J + row * S, Jc + r * S, nv);
        }
      } else {
        for (int r = 1; r < dim && ok; r++)
          for (int s = 0; s < 2 && ok; s++) {
            const double sg = s == 0 ? 1.0 : -1.0, mu = con.friction[r - 1];
            const int row = d.nefc;
            ok = push_row(d, con.dist, con.includemargin, 0, EFC_CONTACT_PYRAMIDAL, ci);
            if (ok) for (int i = 0; i < nv; i++) d.efc_J[row * S + i] = Jc[i] + sg * mu * Jc[r * S + i];
          }
      }
      if (!ok) { d.nefc = adr; break; }     // capacity: drop the partial contact (flagged in d.warning)
      con.efc_address = adr;
    }
  }
  // diagonal approximation of A = J M^-1 J', impedance -> R, D, (K, B, imp)
  const int nefc = d.nefc;
  for (int i = 0; i < nefc; i++) {
    const int id = d.efc_id[i], tp = d.efc_type[i];
    // efc_R temporarily holds diagApprox
    if (tp == EFC_FRICTION_DOF) d.efc_R[i] = m.dof_invweight0[id];
    else if (!T::CONTACTS || tp == EFC_LIMIT_JOINT) d.efc_R[i] = m.dof_invweight0[m.jnt_dofadr[id]];
    else {
      const Contact& con = d.contact[id];
      const int b1 = m.geom_bodyid[con.geom1], b2 = m.geom_bodyid[con.geom2];
      const double tran = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
      const double rot = m.body_invweight0[2 * b1 + 1] + m.body_invweight0[2 * b2 + 1];
      if (tp == EFC_CONTACT_FRICTIONLESS) d.efc_R[i] = tran;
      else if (tp == EFC_CONTACT_ELLIPTIC) { for (int j = 0; j < con.dim; j++) d.efc_R[i + j] = j < 3 ? tran : rot; i += con.dim - 1; }
      else {
        for (int j = 0; j < con.dim - 1; j++) {
          const double fri = con.friction[j];
          d.efc_R[i + 2 * j] = d.efc_R[i + 2 * j + 1] = tran + fri * fri * (j < 2 ? tran : rot);
        }
        i += 2 * (con.dim - 1) - 1;
      }
    }
  }
  for (int i = 0; i < nefc; i++) {
    const int id = d.efc_id[i], tp = d.efc_type[i];
    double solref[2], solimp[5];
    if (tp == EFC_FRICTION_DOF) sanitize_sol(m.timestep, m.disableflags, m.dof_solref + 2 * id, m.dof_solimp + 5 * id, solref, solimp);
    else if (!T::CONTACTS || tp == EFC_LIMIT_JOINT) sanitize_sol(m.timestep, m.disableflags, m.jnt_solref + 2 * id, m.jnt_solimp + 5 * id, solref, solimp);
    else sanitize_sol(m.timestep, m.disableflags, d.contact[id].solref, d.contact[id].solimp, solref, solimp);
    const double imp = impedance_at(solimp, d.efc_pos[i], d.efc_margin[i]);
    const double dmax = solimp[1]; double K, B;
    if (solref[0] > 0) {
      K = 1 / fmax(kMinVal, dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]);
      B = 2 / fmax(kMinVal, dmax * solref[0]);
    } else { K = -solref[0] / fmax(kMinVal, dmax * dmax); B = -solref[1] / fmax(kMinVal, dmax); }
    const bool ell_fric = T::ELLIPTIC && tp == EFC_CONTACT_ELLIPTIC && i > d.contact[id].efc_address;
    if (tp == EFC_FRICTION_DOF || ell_fric) K = 0;
    d.efc_KBI[3 * i] = K; d.efc_KBI[3 * i + 1] = B; d.efc_KBI[3 * i + 2] = imp;
    d.efc_R[i] = fmax(kMinVal, (1 - imp) * d.efc_R[i] / imp);
  }
  if (T::CONTACTS) for (int ci = 0; ci < d.ncon; ci++) {
    Contact& con = d.contact[ci];
    const int i = con.efc_address, dim = con.dim;
    if (i < 0 || dim == 1) continue;
    double* R = d.efc_R;
    if (m.cone == CONE_ELLIPTIC) {
      R[i + 1] = R[i] / fmax(kMinVal, m.impratio);
      con.mu = con.friction[0] * sqrt(R[i + 1] / R[i]);
      for (int j = 1; j < dim - 1; j++) R[i + j + 1] = R[i + 1] * con.friction[0] * con.friction[0] / (con.friction[j] * con.friction[j]);
    } else {
      const double R1 = R[i] / fmax(kMinVal, m.impratio);
      con.mu = con.friction[0] * sqrt(R1 / R[i]);
      const double Rpy = 2 * con.mu * con.mu * R[i];
      for (int j = 0; j < 2 * (dim - 1); j++) R[i + j] = Rpy;
    }
  }
  for (int i = 0; i < nefc; i++) d.efc_D[i] = 1 / d.efc_R[i];
}

// aref = -B (J v) - K imp (pos - margin)
template <class T>
PS_NT void reference_acceleration(const DevModel<T>& m, RolloutData<T>& d) {
  for (int i = 0; i < d.nefc; i++) {
    const double vel = vdot(d.efc_J + i * T::NV, d.qvel, nv_of(m));
    const double* k = d.efc_KBI + 3 * i;
    d.efc_aref[i] = -k[1] * vel - k[0] * k[2] * (d.efc_pos[i] - d.efc_margin[i]);
  }
}

// ---------------------------------------------------------------- Newton solver
// per-row state / force / cost for jar = J a - aref
template <class T>
PS_NT double update_rows(const DevModel<T>& m, RolloutData<T>& d, const double* jar, bool cone_hessian) {
  double cost = 0;
  const int nefc = d.nefc;
  for (int i = 0; i < nefc; i++) {
    const int tp = d.efc_type[i];
    const double D = d.efc_D[i], R = d.efc_R[i], x = jar[i];
    if (tp == EFC_FRICTION_DOF) {
      const double f = d.efc_frictionloss[i];
      if (x <= -R * f) { d.efc_state[i] = ST_LINEARNEG; d.efc_force[i] = f; cost += -0.5 * R * f * f - f * x; }
      else if (x >= R * f) { d.efc_state[i] = ST_LINEARPOS; d.efc_force[i] = -f; cost += -0.5 * R * f * f + f * x; }
      else { d.efc_state[i] = ST_QUADRATIC; d.efc_force[i] = -D * x; cost += 0.5 * D * x * x; }
    } else if (!T::ELLIPTIC || tp != EFC_CONTACT_ELLIPTIC) {
      if (x < 0) { d.efc_state[i] = ST_QUADRATIC; d.efc_force[i] = -D * x; cost += 0.5 * D * x * x; }
      else { d.efc_state[i] = ST_SATISFIED; d.efc_force[i] = 0; }
    } else {
      const int ci = d.efc_id[i];
      const Contact& con = d.contact[ci];
      const int dim = con.dim; const double mu = con.mu; const double* fr = con.friction;
      double U[6]; U[0] = jar[i] * mu;
      for (int j = 1; j < dim; j++) U[j] = jar[i + j] * fr[j - 1];
      const double N = U[0]; double TT = 0;
      for (int j = 1; j < dim; j++) TT += U[j] * U[j];
      const double Tn = sqrt(TT);
      int st;
      if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
        st = ST_SATISFIED; for (int j = 0; j < dim; j++) d.efc_force[i + j] = 0;
      } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
        st = ST_QUADRATIC;
        for (int j = 0; j < dim; j++) { d.efc_force[i + j] = -d.efc_D[i + j] * jar[i + j]; cost += 0.5 * d.efc_D[i + j] * jar[i + j] * jar[i + j]; }
      } else {
        st = ST_CONE;
        const double Dm = d.efc_D[i] / (mu * mu * (1 + mu * mu));
        const double NT = N - mu * Tn;
        cost += 0.5 * Dm * NT * NT;
        d.efc_force[i] = -Dm * NT * mu;
        for (int j = 1; j < dim; j++) d.efc_force[i + j] = -d.efc_force[i] / Tn * U[j] * fr[j - 1];
        if (cone_hessian && T::ELLIPTIC) {
          double* H = d.coneH + 36 * ci;
          H[0] = 1;
          for (int j = 1; j < dim; j++) H[j] = H[j * dim] = -mu * U[j] / Tn;
          const double s1 = mu * N / (Tn * Tn * Tn), s2 = mu * mu - mu * N / Tn;
          for (int k = 1; k < dim; k++) for (int j = 1; j < dim; j++) H[k * dim + j] = s1 * U[j] * U[k] + (k == j ? s2 : 0.0);
          for (int k = 0; k < dim; k++) for (int j = 0; j < dim; j++)
            H[k * dim + j] *= Dm * (k == 0 ? mu : fr[k - 1]) * (j == 0 ? mu : fr[j - 1]);
        }
      }
      for (int j = 0; j < dim; j++) d.efc_state[i + j] = st;
      i += dim - 1;
    }
  }
  return cost;
}

struct SolverScalars { double cost, gauss, prev_cost; };

template <class T>
PS_NT void update_cost(const DevModel<T>& m, RolloutData<T>& d, SolverScalars& c, bool coneH) {
  constexpr int S = T::NV; const int nv = nv_of(m);
  const double cc = update_rows(m, d, d.sv_Jaref, coneH);
  double g = 0;
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) g += 0.5 * (d.sv_Ma[i] - d.qfrc_smooth[i]) * (d.qacc[i] - d.qacc_smooth[i]);
  c.gauss = g; c.prev_cost = c.cost; c.cost = g + cc;
  vzero(d.qfrc_constraint, nv);
  for (int r = 0; r < d.nefc; r++) {
    const double f = d.efc_force[r];
    if (f != 0) PS_SPEC_UNROLL for (int i = 0; i < nv; i++) d.qfrc_constraint[i] += d.efc_J[r * S + i] * f;
  }
}

// grad = M a - qfrc_smooth - J' f
template <class T>
PS_FN void newton_gradient(const DevModel<T>& m, RolloutData<T>& d) {
  const int nv = nv_of(m);
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) d.sv_grad[i] = d.sv_Ma[i] - d.qfrc_smooth[i] - d.qfrc_constraint[i];
}

// H = M + J' D_active J (+ cone blocks), dense Cholesky, search = -H^-1 grad (grad: newton_gradient)
template <class T>
PS_NT void newton_direction(const DevModel<T>& m, RolloutData<T>& d) {
  constexpr int S = T::NV; const int nv = nv_of(m), nefc = d.nefc;
  double* H = d.sv_H;
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++) H[i * S + j] = d.qM[i * S + j];
  for (int r = 0; r < nefc; r++) {
    const int st = d.efc_state[r];
    if (st == ST_QUADRATIC) {
      const double D = d.efc_D[r]; const double* J = d.efc_J + r * S;
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) { if (J[i] == 0) continue; const double s = D * J[i]; PS_SPEC_UNROLL for (int j = 0; j <= i; j++) H[i * S + j] += s * J[j]; }
    } else if (T::ELLIPTIC && st == ST_CONE && d.efc_type[r] == EFC_CONTACT_ELLIPTIC) {
      // cone block J_c' (Hc J_c): T = Hc J_c first, then the symmetric update, one add per cone row
      const int ci = d.efc_id[r]; const int dim = d.contact[ci].dim; const double* CH = d.coneH + 36 * ci;
      double Tm[6 * S];
      for (int a = 0; a < dim; a++) PS_SPEC_UNROLL for (int j = 0; j < nv; j++) {
        double t = 0; for (int b = 0; b < dim; b++) t += CH[a * dim + b] * d.efc_J[(r + b) * S + j];
        Tm[a * S + j] = t;
      }
      PS_SPEC_UNROLL for (int i = 0; i < nv; i++) PS_SPEC_UNROLL for (int j = 0; j <= i; j++)
        for (int a = 0; a < dim; a++) H[i * S + j] += d.efc_J[(r + a) * S + i] * Tm[a * S + j];
      r += dim - 1;
    }
  }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++)
    PS_SPEC_UNROLL for (int j = 0; j <= i; j++) {
      double s = H[i * S + j];
      PS_SPEC_UNROLL for (int k = 0; k < j; k++) s -= H[i * S + k] * H[j * S + k];
      if (i == j) H[i * S + i] = sqrt(fmax(s, kMinVal)); else H[i * S + j] = s / H[j * S + j];
    }
  double* x = d.sv_search;
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) { double s = d.sv_grad[i]; PS_SPEC_UNROLL for (int k = 0; k < i; k++) s -= H[i * S + k] * x[k]; x[i] = s / H[i * S + i]; }
  PS_SPEC_UNROLL for (int i = nv - 1; i >= 0; i--) { double s = x[i]; PS_SPEC_UNROLL for (int k = nv - 1; k > i; k--) s -= H[k * S + i] * x[k]; x[i] = s / H[i * S + i]; }
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) x[i] = -x[i];
}

// value, slope and curvature of the total cost along the search direction at step alpha
template <class T>
PS_NT void line_eval(const DevModel<T>& m, const RolloutData<T>& d, const double* qg, double alpha,
                     double* cost, double* d1, double* d2) {
  double f = qg[0] + alpha * qg[1] + alpha * alpha * qg[2];
  double g = qg[1] + 2 * alpha * qg[2];
  double h = 2 * qg[2];
  const int nefc = d.nefc;
  for (int i = 0; i < nefc; i++) {
    const int tp = d.efc_type[i];
    const double D = d.efc_D[i], R = d.efc_R[i];
    const double x = d.sv_Jaref[i] + alpha * d.sv_Jv[i], v = d.sv_Jv[i];
    if (tp == EFC_FRICTION_DOF) {
      const double fl = d.efc_frictionloss[i];
      if (x <= -R * fl) { f += -0.5 * R * fl * fl - fl * x; g += -fl * v; }
      else if (x >= R * fl) { f += -0.5 * R * fl * fl + fl * x; g += fl * v; }
      else { f += 0.5 * D * x * x; g += D * x * v; h += D * v * v; }
    } else if (!T::ELLIPTIC || tp != EFC_CONTACT_ELLIPTIC) {
      if (x < 0) { f += 0.5 * D * x * x; g += D * x * v; h += D * v * v; }
    } else {
      const Contact& con = d.contact[d.efc_id[i]];
      const int dim = con.dim; const double mu = con.mu; const double* fr = con.friction;
      const double N = x * mu, Np = v * mu; double TT = 0, UV = 0, VV = 0;
      for (int j = 1; j < dim; j++) {
        const double u = (d.sv_Jaref[i + j] + alpha * d.sv_Jv[i + j]) * fr[j - 1], w = d.sv_Jv[i + j] * fr[j - 1];
        TT += u * u; UV += u * w; VV += w * w;
      }
      const double Tn = sqrt(TT);
      if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
      } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
        for (int j = 0; j < dim; j++) {
          const double xj = d.sv_Jaref[i + j] + alpha * d.sv_Jv[i + j], vj = d.sv_Jv[i + j], Dj = d.efc_D[i + j];
          f += 0.5 * Dj * xj * xj; g += Dj * xj * vj; h += Dj * vj * vj;
        }
      } else {
        const double Dm = D / (mu * mu * (1 + mu * mu));
        const double Tp = UV / Tn, Tpp = VV / Tn - UV * UV / (Tn * Tn * Tn);
        const double NT = N - mu * Tn, NTp = Np - mu * Tp;
        f += 0.5 * Dm * NT * NT; g += Dm * NT * NTp; h += Dm * (NTp * NTp + NT * (-mu * Tpp));
      }
      i += dim - 1;
    }
  }
  *cost = f; *d1 = g; *d2 = h;
}

template <class T>
PS_NT double line_search(const DevModel<T>& m, RolloutData<T>& d, const SolverScalars& c) {
  constexpr int S = T::NV; const int nv = nv_of(m), nefc = d.nefc;
  const double snorm = sqrt(vdot(d.sv_search, d.sv_search, nv));
  if (snorm < kMinVal) return 0;
  const double scale = 1.0 / (m.meaninertia * (nv > 1 ? nv : 1));
  const double gtol = m.tolerance * m.ls_tolerance * snorm / scale;
  mul_M(m, d, d.sv_Mv, d.sv_search);
  for (int r = 0; r < nefc; r++) d.sv_Jv[r] = vdot(d.efc_J + r * S, d.sv_search, nv);
  double qg[3] = {c.gauss, 0, 0};
  PS_SPEC_UNROLL for (int i = 0; i < nv; i++) { qg[1] += d.sv_search[i] * (d.sv_Ma[i] - d.qfrc_smooth[i]); qg[2] += 0.5 * d.sv_search[i] * d.sv_Mv[i]; }
  double c0, d0, h0; line_eval(m, d, qg, 0, &c0, &d0, &h0);
  if (!(h0 > 0) || d0 >= 0) return 0;
  double lo = 0, hi = -1;
  double alpha = -d0 / h0, cc, dd, hh;
  line_eval(m, d, qg, alpha, &cc, &dd, &hh);
  for (int it = 0; it < m.ls_iterations && fabs(dd) >= gtol; it++) {
    if (dd < 0) lo = alpha; else hi = alpha;
    double an = hh > 0 ? alpha - dd / hh : -1;
    if (hi < 0) { if (!(an > lo)) an = 2 * alpha; }
    else if (!(an > lo && an < hi)) an = 0.5 * (lo + hi);
    alpha = an;
    line_eval(m, d, qg, alpha, &cc, &dd, &hh);
  }
  if (!(cc < c0)) return 0;
  return alpha;
}

template <class T>
PS_NT void constraint_stage(const DevModel<T>& m, RolloutData<T>& d) {
  constexpr int S = T::NV; const int nv = nv_of(m), nefc = d.nefc;
  d.solver_iter = 0;
  if (!nefc) {
    PS_SPEC_UNROLL for (int i = 0; i < nv; i++) { d.qacc[i] = d.qacc_smooth[i]; d.qacc_warmstart[i] = d.qacc_smooth[i]; d.qfrc_constraint[i] = 0; }
    return;
  }
  SolverScalars c; c.cost = 0; c.prev_cost = 0; c.gauss = 0;
  const bool coneH = m.cone == CONE_ELLIPTIC;
  if (!(m.disableflags & DSBL_WARMSTART)) {
    // keep the warm start only if it beats the unconstrained acceleration
    vcopy(d.qacc, d.qacc_warmstart, nv);
    mul_M(m, d, d.sv_Ma, d.qacc);
    for (int r = 0; r < nefc; r++) d.sv_Jaref[r] = vdot(d.efc_J + r * S, d.qacc, nv) - d.efc_aref[r];
    double cw = update_rows(m, d, d.sv_Jaref, false);
    PS_SPEC_UNROLL for (int i = 0; i < nv; i++) cw += 0.5 * (d.sv_Ma[i] - d.qfrc_smooth[i]) * (d.qacc[i] - d.qacc_smooth[i]);
    for (int r = 0; r < nefc; r++) d.sv_Jv[r] = vdot(d.efc_J + r * S, d.qacc_smooth, nv) - d.efc_aref[r];
    const double cs = update_rows(m, d, d.sv_Jv, false);
    if (cw > cs) vcopy(d.qacc, d.qacc_smooth, nv);
  } else vcopy(d.qacc, d.qacc_smooth, nv);
  mul_M(m, d, d.sv_Ma, d.qacc);
  for (int r = 0; r < nefc; r++) d.sv_Jaref[r] = vdot(d.efc_J + r * S, d.qacc, nv) - d.efc_aref[r];
  update_cost(m, d, c, coneH);
  newton_gradient(m, d);
  newton_direction(m, d);
  const double scale = 1.0 / (m.meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  while (iter < m.iterations) {
    const double alpha = line_search(m, d, c);
    if (alpha == 0) break;
    PS_SPEC_UNROLL for (int i = 0; i < nv; i++) { d.qacc[i] += alpha * d.sv_search[i]; d.sv_Ma[i] += alpha * d.sv_Mv[i]; }
    for (int r = 0; r < nefc; r++) d.sv_Jaref[r] += alpha * d.sv_Jv[r];
    update_cost(m, d, c, coneH);
    // mj_solNewton factorises the Hessian and solves for the next direction before it tests for convergence; the direction
    // of the last iteration is never used, so the test comes first here and the last factorisation is skipped
    newton_gradient(m, d);
    iter++;
    const double improvement = scale * (c.prev_cost - c.cost);
    const double gradient = scale * sqrt(vdot(d.sv_grad, d.sv_grad, nv));
    if (improvement < m.tolerance || gradient < m.tolerance || iter >= m.iterations) break;
    newton_direction(m, d);
  }
  d.solver_iter = iter;
  vcopy(d.qacc_warmstart, d.qacc, nv);
}

}  // namespace ps
