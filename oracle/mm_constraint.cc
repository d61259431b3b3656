// ORACLE (test infrastructure).  Constraint construction: friction-loss, joint-limit and contact
// rows, impedance / reference acceleration.  Restates MuJoCo 3.1.4 engine_core_constraint
// (mj_makeConstraint, mj_makeImpedance, mj_referenceConstraint; third-party, not in
// /root/reference).  Parity vs MuJoCo: UNPINNED.
#include <algorithm>
#include "mm_engine.h"
#include "mm_math.h"

namespace mm {

int collision(const Model* m, Data* d);

constexpr double kMinImp = 0.0001, kMaxImp = 0.9999;

static void addRow(Data* d, int nv, const double* J, double pos, double margin, double floss, int type, int id) {
  d->efc_J.insert(d->efc_J.end(), J, J + nv);
  d->efc_pos.push_back(pos); d->efc_margin.push_back(margin); d->efc_frictionloss.push_back(floss);
  d->efc_type.push_back(type); d->efc_id.push_back(id);
  d->nefc++;
}

static void sanitizeSol(const Model* m, const double* sr, const double* si, double* solref, double* solimp) {
  copy(solref, sr, 2); copy(solimp, si, 5);
  if (!(m->opt_disableflags & PS_DSBL_REFSAFE) && solref[0] > 0) solref[0] = std::fmax(solref[0], 2 * m->opt_timestep);
  solimp[0] = clip(solimp[0], kMinImp, kMaxImp); solimp[1] = clip(solimp[1], kMinImp, kMaxImp);
  solimp[2] = std::fmax(0.0, solimp[2]); solimp[3] = clip(solimp[3], kMinImp, kMaxImp);
  solimp[4] = std::fmax(1.0, solimp[4]);
}

// x^p with the integer powers MuJoCo models actually use (solimp power 1 or 2) evaluated by exact
// IEEE multiplication, so the result does not depend on the libm at hand
static double powi(double x, double p) {
  if (p == 1) return x;
  if (p == 2) return x * x;
  if (p == 3) return x * x * x;
  return std::pow(x, p);
}

static double impedance(const double* si, double pos, double margin) {
  if (si[0] == si[1] || si[2] <= kMinVal) return 0.5 * (si[0] + si[1]);
  double x = std::fabs((pos - margin) / si[2]);
  if (x >= 1) return si[1];
  if (x <= 0) return si[0];
  double y;
  if (si[4] == 1) y = x;
  else if (x <= si[3]) y = powi(x, si[4]) / powi(si[3], si[4] - 1);
  else y = 1 - powi(1 - x, si[4]) / powi(1 - si[3], si[4] - 1);
  return si[0] + y * (si[1] - si[0]);
}

void makeConstraint(const Model* m, Data* d) {
  int nv = m->nv;
  d->nefc = 0;
  d->efc_J.clear(); d->efc_pos.clear(); d->efc_margin.clear(); d->efc_frictionloss.clear();
  d->efc_type.clear(); d->efc_id.clear();
  if (m->opt_disableflags & PS_DSBL_CONSTRAINT) { d->contact.clear(); d->ncon = 0; return; }
  collision(m, d);
  std::vector<double> J(nv);
  // friction loss
  if (!(m->opt_disableflags & PS_DSBL_FRICTIONLOSS))
    for (int i = 0; i < nv; i++) if (m->dof_frictionloss[i] > 0) {
      std::fill(J.begin(), J.end(), 0.0); J[i] = 1;
      addRow(d, nv, J.data(), 0, 0, m->dof_frictionloss[i], CNSTR_FRICTION_DOF, i);
    }
  // joint limits (hinge / slide)
  if (!(m->opt_disableflags & PS_DSBL_LIMIT))
    for (int j = 0; j < m->njnt; j++) {
      if (!m->jnt_limited[j]) continue;
      int t = m->jnt_type[j];
      if (t != PS_JNT_HINGE && t != PS_JNT_SLIDE) continue;
      double value = d->qpos[m->jnt_qposadr[j]], margin = m->jnt_margin[j];
      for (int side = -1; side <= 1; side += 2) {
        double dist = side * (m->jnt_range[2 * j + (side + 1) / 2] - value);
        if (dist < margin) {
          std::fill(J.begin(), J.end(), 0.0); J[m->jnt_dofadr[j]] = -(double)side;
          addRow(d, nv, J.data(), dist, margin, 0, CNSTR_LIMIT_JOINT, j);
        }
      }
    }
  // contacts
  std::vector<double> jp1(3 * nv), jp2(3 * nv), jr1(3 * nv), jr2(3 * nv), jd(6 * nv), Jc(6 * nv);
  for (int ci = 0; ci < d->ncon; ci++) {
    Contact& con = d->contact[ci];
    con.efc_address = -1;
    if (con.dist >= con.includemargin) continue;
    int b1 = m->geom_bodyid[con.geom1], b2 = m->geom_bodyid[con.geom2];
    jac(m, d, jp1.data(), jr1.data(), con.pos, b1);
    jac(m, d, jp2.data(), jr2.data(), con.pos, b2);
    for (int k = 0; k < 3 * nv; k++) { jd[k] = jp2[k] - jp1[k]; jd[3 * nv + k] = jr2[k] - jr1[k]; }
    // rotate into contact frame: Jc[r] = frame_r . jd(trans) for r<3, frame_{r-3} . jd(rot) for r>=3
    int dim = con.dim;
    for (int r = 0; r < dim; r++) {
      const double* f = con.frame + 3 * (r % 3); const double* src = jd.data() + (r < 3 ? 0 : 3 * nv);
      for (int i = 0; i < nv; i++) Jc[r * nv + i] = f[0] * src[i] + f[1] * src[nv + i] + f[2] * src[2 * nv + i];
    }
    con.efc_address = d->nefc;
    if (dim == 1) {
      addRow(d, nv, Jc.data(), con.dist, con.includemargin, 0, CNSTR_CONTACT_FRICTIONLESS, ci);
    } else if (m->opt_cone == PS_CONE_ELLIPTIC) {
      for (int r = 0; r < dim; r++)
        addRow(d, nv, Jc.data() + r * nv, r == 0 ? con.dist : double(0), r == 0 ? con.includemargin : double(0), 0, CNSTR_CONTACT_ELLIPTIC, ci);
    } else {
      for (int r = 1; r < dim; r++)
        for (int s = 0; s < 2; s++) {
          double sg = s == 0 ? 1.0 : -1.0, mu = con.friction[r - 1];
          for (int i = 0; i < nv; i++) J[i] = Jc[i] + sg * mu * Jc[r * nv + i];
          addRow(d, nv, J.data(), con.dist, con.includemargin, 0, CNSTR_CONTACT_PYRAMIDAL, ci);
        }
    }
  }
  // ---- diagApprox, impedance (R, D, KBIP)
  int nefc = d->nefc;
  d->efc_diagApprox.assign(nefc, 0); d->efc_R.assign(nefc, 0); d->efc_D.assign(nefc, 0);
  d->efc_KBIP.assign(4 * nefc, 0); d->efc_vel.assign(nefc, 0); d->efc_aref.assign(nefc, 0);
  d->efc_b.assign(nefc, 0); d->efc_force.assign(nefc, 0); d->efc_state.assign(nefc, 0);
  for (int i = 0; i < nefc; i++) {
    int id = d->efc_id[i], tp = d->efc_type[i];
    if (tp == CNSTR_FRICTION_DOF) d->efc_diagApprox[i] = m->dof_invweight0[id];
    else if (tp == CNSTR_LIMIT_JOINT) d->efc_diagApprox[i] = m->dof_invweight0[m->jnt_dofadr[id]];
    else {
      const Contact& con = d->contact[id];
      int b1 = m->geom_bodyid[con.geom1], b2 = m->geom_bodyid[con.geom2];
      double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
      double rot = m->body_invweight0[2 * b1 + 1] + m->body_invweight0[2 * b2 + 1];
      if (tp == CNSTR_CONTACT_FRICTIONLESS) d->efc_diagApprox[i] = tran;
      else if (tp == CNSTR_CONTACT_ELLIPTIC) { for (int j = 0; j < con.dim; j++) d->efc_diagApprox[i + j] = j < 3 ? tran : rot; i += con.dim - 1; }
      else {
        for (int j = 0; j < con.dim - 1; j++) {
          double fri = con.friction[j];
          d->efc_diagApprox[i + 2 * j] = d->efc_diagApprox[i + 2 * j + 1] = tran + fri * fri * (j < 2 ? tran : rot);
        }
        i += 2 * (con.dim - 1) - 1;
      }
    }
  }
  for (int i = 0; i < nefc; i++) {
    int id = d->efc_id[i], tp = d->efc_type[i];
    double solref[2], solimp[5];
    if (tp == CNSTR_FRICTION_DOF) sanitizeSol(m, m->dof_solref + 2 * id, m->dof_solimp + 5 * id, solref, solimp);
    else if (tp == CNSTR_LIMIT_JOINT) sanitizeSol(m, m->jnt_solref + 2 * id, m->jnt_solimp + 5 * id, solref, solimp);
    else sanitizeSol(m, d->contact[id].solref, d->contact[id].solimp, solref, solimp);
    double imp = impedance(solimp, d->efc_pos[i], d->efc_margin[i]);
    double dmax = solimp[1], K, B;
    if (solref[0] > 0) {
      K = 1 / std::fmax(kMinVal, dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]);
      B = 2 / std::fmax(kMinVal, dmax * solref[0]);
    } else { K = -solref[0] / std::fmax(kMinVal, dmax * dmax); B = -solref[1] / std::fmax(kMinVal, dmax); }
    bool ell_fric = tp == CNSTR_CONTACT_ELLIPTIC && i > d->contact[id].efc_address;
    if (tp == CNSTR_FRICTION_DOF || ell_fric) K = 0;
    d->efc_KBIP[4 * i] = K; d->efc_KBIP[4 * i + 1] = B; d->efc_KBIP[4 * i + 2] = imp; d->efc_KBIP[4 * i + 3] = 0;
    d->efc_R[i] = std::fmax(kMinVal, (1 - imp) * d->efc_diagApprox[i] / imp);
  }
  // frictional contacts: adjust R in friction dimensions, set regularized cone mu
  for (int ci = 0; ci < d->ncon; ci++) {
    Contact& con = d->contact[ci];
    int i = con.efc_address, dim = con.dim;
    if (i < 0 || dim == 1) continue;
    double* R = d->efc_R.data();
    if (m->opt_cone == PS_CONE_ELLIPTIC) {
      R[i + 1] = R[i] / std::fmax(kMinVal, m->opt_impratio);
      con.mu = con.friction[0] * std::sqrt(R[i + 1] / R[i]);
      for (int j = 1; j < dim - 1; j++) R[i + j + 1] = R[i + 1] * con.friction[0] * con.friction[0] / (con.friction[j] * con.friction[j]);
    } else {
      double R1 = R[i] / std::fmax(kMinVal, m->opt_impratio);   // R of first row plays the role of R[0]
      con.mu = con.friction[0] * std::sqrt(R1 / R[i]);
      double Rpy = 2 * con.mu * con.mu * R[i];
      for (int j = 0; j < 2 * (dim - 1); j++) R[i + j] = Rpy;
    }
  }
  for (int i = 0; i < nefc; i++) d->efc_D[i] = 1 / d->efc_R[i];
}

// mj_referenceConstraint: efc_vel = J qvel ; aref = -B vel - K imp (pos - margin)
void referenceConstraint(const Model* m, Data* d) {
  int nv = m->nv;
  for (int i = 0; i < d->nefc; i++) {
    d->efc_vel[i] = dot(&d->efc_J[i * nv], d->qvel.data(), nv);
    const double* k = &d->efc_KBIP[4 * i];
    d->efc_aref[i] = -k[1] * d->efc_vel[i] - k[0] * k[2] * (d->efc_pos[i] - d->efc_margin[i]);
  }
}

}  // namespace mm
