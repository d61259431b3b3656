// ORACLE (test infrastructure).  Collision detection for primitive pairs (plane-sphere/capsule/
// cylinder/box, sphere-sphere, sphere-capsule, capsule-capsule).  Restates MuJoCo 3.1.4
// engine_collision_driver / engine_collision_primitive (third-party, not in /root/reference).
// Parity vs MuJoCo: UNPINNED.  Pairs outside this list are reported in Data::contact as skipped
// (see DESIGN.md "supported model class").
#include <algorithm>
#include "mm_engine.h"
#include "mm_math.h"

namespace mm {

static bool isStatic(const Model* m, int b) {
  while (b > 0) { if (m->body_dofnum[b] > 0) return false; b = m->body_parentid[b]; }
  return true;
}
// nearest ancestor-or-self with dofs (the "weld" body), 0 if static
static int weldBody(const Model* m, int b) {
  while (b > 0 && m->body_dofnum[b] == 0) b = m->body_parentid[b];
  return b;
}

static void makeFrame(double* f) {
  normalize3(f);
  if (norm3(f + 3) < 0.5) { zero(f + 3, 3); if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1; }
  double dp = dot3(f, f + 3);
  for (int k = 0; k < 3; k++) f[3 + k] -= f[k] * dp;
  normalize3(f + 3);
  cross(f + 6, f, f + 3);
}

struct RawCon { double dist, pos[3], frame[6]; };

static int planeSphere(RawCon* c, const double* p1, const double* m1, const double* p2, double r, double margin) {
  double n[3] = {m1[2], m1[5], m1[8]};
  double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  double dist = dot3(dif, n) - r;
  if (dist > margin) return 0;
  c->dist = dist;
  for (int k = 0; k < 3; k++) { c->pos[k] = p2[k] + n[k] * (-dist * 0.5 - r); c->frame[k] = n[k]; c->frame[3 + k] = 0; }
  return 1;
}

static int planeCapsule(RawCon* c, const double* p1, const double* m1, const double* p2, const double* m2,
                        const double* size, double margin) {
  double axis[3] = {m2[2], m2[5], m2[8]};
  int cnt = 0;
  for (int s = 0; s < 2; s++) {
    double sgn = s == 0 ? 1.0 : -1.0, e[3];
    for (int k = 0; k < 3; k++) e[k] = p2[k] + sgn * axis[k] * size[1];
    int n = planeSphere(c + cnt, p1, m1, e, size[0], margin);
    if (n) { copy(c[cnt].frame + 3, axis, 3); cnt++; }
  }
  return cnt;
}

static int planeCylinder(RawCon* c, const double* p1, const double* m1, const double* p2, const double* m2,
                         const double* size, double margin) {
  double n[3] = {m1[2], m1[5], m1[8]}, axis[3] = {m2[2], m2[5], m2[8]};
  double r = size[0], h = size[1];
  double prjaxis = dot3(n, axis);
  if (prjaxis > 0) { for (int k = 0; k < 3; k++) axis[k] = -axis[k]; prjaxis = -prjaxis; }
  double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  double dist0 = dot3(dif, n);
  double vec[3]; for (int k = 0; k < 3; k++) vec[k] = axis[k] * prjaxis - n[k];
  double len2 = dot3(vec, vec);
  if (len2 >= kMinVal * kMinVal) { double s = r / std::sqrt(len2); for (int k = 0; k < 3; k++) vec[k] *= s; }
  else { vec[0] = m2[0] * r; vec[1] = m2[3] * r; vec[2] = m2[6] * r; }
  double prjvec = dot3(vec, n);
  for (int k = 0; k < 3; k++) axis[k] *= h;
  prjaxis *= h;
  int cnt = 0;
  if (dist0 + prjaxis + prjvec <= margin) {
    double dist = dist0 + prjaxis + prjvec; c[cnt].dist = dist;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = p2[k] + vec[k] + axis[k] - n[k] * dist * 0.5; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    cnt++;
  } else return 0;
  if (dist0 - prjaxis + prjvec <= margin) {
    double dist = dist0 - prjaxis + prjvec; c[cnt].dist = dist;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = p2[k] + vec[k] - axis[k] - n[k] * dist * 0.5; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    cnt++;
  }
  double prjvec1 = -prjvec * 0.5;
  if (dist0 + prjaxis + prjvec1 <= margin) {
    double vec1[3]; cross(vec1, vec, axis); normalize3(vec1);
    double s = r * std::sqrt(3.0) * 0.5; for (int k = 0; k < 3; k++) vec1[k] *= s;
    double dist = dist0 + prjaxis + prjvec1;
    for (int q = 0; q < 2; q++) {
      double sg = q == 0 ? 1.0 : -1.0; c[cnt].dist = dist;
      for (int k = 0; k < 3; k++) {
        c[cnt].pos[k] = p2[k] + sg * vec1[k] + axis[k] - vec[k] * 0.5 - n[k] * dist * 0.5;
        c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0;
      }
      cnt++;
    }
  }
  return cnt;
}

static int planeBox(RawCon* c, const double* p1, const double* m1, const double* p2, const double* m2,
                    const double* size, double margin) {
  double n[3] = {m1[2], m1[5], m1[8]};
  double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  double dist = dot3(dif, n);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    double v[3] = {(i & 1 ? size[0] : -size[0]), (i & 2 ? size[1] : -size[1]), (i & 4 ? size[2] : -size[2])};
    double corner[3]; mulMatVec3(corner, m2, v);
    double ld = dot3(n, corner);
    if (dist + ld > margin || ld > 0) continue;
    c[cnt].dist = dist + ld;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = p2[k] + corner[k] - n[k] * c[cnt].dist * 0.5; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    if (++cnt >= 4) break;
  }
  return cnt;
}

static int sphereSphere(RawCon* c, const double* p1, double r1, const double* p2, double r2, double margin) {
  double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  double cd = norm3(dif);
  double dist = cd - r1 - r2;
  if (dist > margin) return 0;
  double n[3];
  if (cd < kMinVal) { n[0] = 1; n[1] = 0; n[2] = 0; } else for (int k = 0; k < 3; k++) n[k] = dif[k] / cd;
  c->dist = dist;
  for (int k = 0; k < 3; k++) { c->pos[k] = p1[k] + n[k] * (r1 + 0.5 * dist); c->frame[k] = n[k]; c->frame[3 + k] = 0; }
  return 1;
}

static int sphereCapsule(RawCon* c, const double* p1, double r1, const double* p2, const double* m2,
                         const double* size2, double margin) {
  double axis[3] = {m2[2], m2[5], m2[8]};
  double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  double x = clip(dot3(axis, dif), -size2[1], size2[1]);
  double q[3] = {p2[0] + axis[0] * x, p2[1] + axis[1] * x, p2[2] + axis[2] * x};
  return sphereSphere(c, p1, r1, q, size2[0], margin);
}

static int capsuleCapsule(RawCon* c, const double* p1, const double* m1, const double* s1,
                          const double* p2, const double* m2, const double* s2, double margin) {
  double a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]};
  double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  double ma = dot3(a1, a1), mb = -dot3(a1, a2), mc = dot3(a2, a2);
  double u = -dot3(a1, dif), v = dot3(a2, dif);
  double det = ma * mc - mb * mb;
  int cnt = 0;
  if (std::fabs(det) >= kMinVal) {   // general configuration
    double x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
    if (x1 > s1[1]) { x1 = s1[1]; x2 = (v - mb * s1[1]) / mc; }
    else if (x1 < -s1[1]) { x1 = -s1[1]; x2 = (v + mb * s1[1]) / mc; }
    if (x2 > s2[1]) { x2 = s2[1]; x1 = clip((u - mb * s2[1]) / ma, -s1[1], s1[1]); }
    else if (x2 < -s2[1]) { x2 = -s2[1]; x1 = clip((u + mb * s2[1]) / ma, -s1[1], s1[1]); }
    double q1[3], q2[3];
    for (int k = 0; k < 3; k++) { q1[k] = p1[k] + a1[k] * x1; q2[k] = p2[k] + a2[k] * x2; }
    return sphereSphere(c, q1, s1[0], q2, s2[0], margin);
  }
  // parallel axes: test both ends of each capsule against the other's segment
  for (int e = 0; e < 2; e++) {
    double sg = e == 0 ? 1.0 : -1.0, q1[3], q2[3];
    for (int k = 0; k < 3; k++) q1[k] = p1[k] + sg * a1[k] * s1[1];
    double d2[3] = {q1[0] - p2[0], q1[1] - p2[1], q1[2] - p2[2]};
    double x2 = clip(dot3(a2, d2), -s2[1], s2[1]);
    for (int k = 0; k < 3; k++) q2[k] = p2[k] + a2[k] * x2;
    cnt += sphereSphere(c + cnt, q1, s1[0], q2, s2[0], margin);
    if (cnt >= 2) return cnt;
  }
  return cnt;
}

// mix contact parameters of two geoms (mj_contactParam)
static void contactParam(const Model* m, int g1, int g2, Contact* con) {
  int p1 = m->geom_priority[g1], p2 = m->geom_priority[g2];
  double fri[3];
  if (p1 != p2) {
    int g = p1 > p2 ? g1 : g2;
    con->dim = m->geom_condim[g];
    copy(con->solref, m->geom_solref + 2 * g, 2); copy(con->solimp, m->geom_solimp + 5 * g, 5);
    copy(fri, m->geom_friction + 3 * g, 3);
  } else {
    con->dim = std::max(m->geom_condim[g1], m->geom_condim[g2]);
    double s1 = m->geom_solmix[g1], s2 = m->geom_solmix[g2], mix;
    if (s1 >= kMinVal && s2 >= kMinVal) mix = s1 / (s1 + s2);
    else if (s1 < kMinVal && s2 < kMinVal) mix = 0.5;
    else mix = s1 < kMinVal ? 0.0 : 1.0;
    const double* r1 = m->geom_solref + 2 * g1; const double* r2 = m->geom_solref + 2 * g2;
    if (r1[0] > 0 && r2[0] > 0) for (int k = 0; k < 2; k++) con->solref[k] = mix * r1[k] + (1 - mix) * r2[k];
    else for (int k = 0; k < 2; k++) con->solref[k] = std::fmin(r1[k], r2[k]);
    for (int k = 0; k < 5; k++) con->solimp[k] = mix * m->geom_solimp[5 * g1 + k] + (1 - mix) * m->geom_solimp[5 * g2 + k];
    for (int k = 0; k < 3; k++) fri[k] = std::fmax(m->geom_friction[3 * g1 + k], m->geom_friction[3 * g2 + k]);
  }
  con->friction[0] = con->friction[1] = fri[0]; con->friction[2] = fri[1];
  con->friction[3] = con->friction[4] = fri[2];
}

int collision(const Model* m, Data* d) {
  d->contact.clear();
  int unsupported = 0;
  if ((m->opt_disableflags & PS_DSBL_CONTACT) || (m->opt_disableflags & PS_DSBL_CONSTRAINT)) { d->ncon = 0; return 0; }
  for (int ga = 0; ga < m->ngeom; ga++)
    for (int gb = ga + 1; gb < m->ngeom; gb++) {
      int b1 = m->geom_bodyid[ga], b2 = m->geom_bodyid[gb];
      if (b1 == b2) continue;
      if (!((m->geom_contype[ga] & m->geom_conaffinity[gb]) || (m->geom_contype[gb] & m->geom_conaffinity[ga]))) continue;
      int w1 = weldBody(m, b1), w2 = weldBody(m, b2);
      if (w1 == w2) continue;                  // includes static-static
      if (!(m->opt_disableflags & PS_DSBL_FILTERPARENT) && w1 != 0 && w2 != 0) {
        if (weldBody(m, m->body_parentid[w1]) == w2 || weldBody(m, m->body_parentid[w2]) == w1) continue;
      }
      int g1 = ga, g2 = gb;
      if (m->geom_type[g1] > m->geom_type[g2]) std::swap(g1, g2);
      int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
      double margin = std::fmax(m->geom_margin[g1], m->geom_margin[g2]);
      double gap = std::fmax(m->geom_gap[g1], m->geom_gap[g2]);
      const double* p1 = &d->geom_xpos[3 * g1]; const double* m1 = &d->geom_xmat[9 * g1];
      const double* p2 = &d->geom_xpos[3 * g2]; const double* m2 = &d->geom_xmat[9 * g2];
      const double* s1 = m->geom_size + 3 * g1; const double* s2 = m->geom_size + 3 * g2;
      // bounding-sphere / plane filter
      if (t1 == PS_GEOM_PLANE) {
        double n[3] = {m1[2], m1[5], m1[8]}, dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
        if (t2 == PS_GEOM_PLANE) continue;
        if (dot3(dif, n) > margin + m->geom_rbound[g2]) continue;
      } else {
        double dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
        double bound = m->geom_rbound[g1] + m->geom_rbound[g2] + margin;
        if (dot3(dif, dif) > bound * bound) continue;
      }
      RawCon rc[8]; int n = 0;
      if (t1 == PS_GEOM_PLANE && t2 == PS_GEOM_SPHERE) n = planeSphere(rc, p1, m1, p2, s2[0], margin);
      else if (t1 == PS_GEOM_PLANE && t2 == PS_GEOM_CAPSULE) n = planeCapsule(rc, p1, m1, p2, m2, s2, margin);
      else if (t1 == PS_GEOM_PLANE && t2 == PS_GEOM_CYLINDER) n = planeCylinder(rc, p1, m1, p2, m2, s2, margin);
      else if (t1 == PS_GEOM_PLANE && t2 == PS_GEOM_BOX) n = planeBox(rc, p1, m1, p2, m2, s2, margin);
      else if (t1 == PS_GEOM_SPHERE && t2 == PS_GEOM_SPHERE) n = sphereSphere(rc, p1, s1[0], p2, s2[0], margin);
      else if (t1 == PS_GEOM_SPHERE && t2 == PS_GEOM_CAPSULE) n = sphereCapsule(rc, p1, s1[0], p2, m2, s2, margin);
      else if (t1 == PS_GEOM_CAPSULE && t2 == PS_GEOM_CAPSULE) n = capsuleCapsule(rc, p1, m1, s1, p2, m2, s2, margin);
      else { unsupported++; continue; }
      for (int k = 0; k < n; k++) {
        Contact con{};
        con.dist = rc[k].dist; copy(con.pos, rc[k].pos, 3); copy(con.frame, rc[k].frame, 6);
        makeFrame(con.frame);
        con.includemargin = margin - gap; con.geom1 = g1; con.geom2 = g2; con.efc_address = -1;
        contactParam(m, g1, g2, &con);
        d->contact.push_back(con);
      }
    }
  d->ncon = (int)d->contact.size();
  return unsupported;
}

}  // namespace mm
