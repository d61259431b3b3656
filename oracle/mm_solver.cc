// ORACLE (test infrastructure).  Soft-constraint solver: primal Newton method with exact 1-D line
// search on the convex cost  0.5 (Ma - f_smooth).(a - a_smooth) + sum_i s_i(J_i a - aref_i).
// Restates MuJoCo 3.1.4 engine_solver (mj_solNewton) and engine_core_constraint
// (mj_constraintUpdate) -- third-party, not in /root/reference.  The line search is a safeguarded
// Newton iteration on the (monotone) directional derivative with MuJoCo's tolerance semantics; it
// is NOT claimed to reproduce MuJoCo's bracketing heuristics step for step.  Parity: UNPINNED.
#include <algorithm>
#include "mm_engine.h"
#include "mm_math.h"

namespace mm {

// per-row cost / force / state for a given jar = J a - aref   (mj_constraintUpdate)
static double constraintUpdate(const Model* m, Data* d, const double* jar, bool coneHessian) {
  double cost = 0;
  int nefc = d->nefc;
  for (int i = 0; i < nefc; i++) {
    int tp = d->efc_type[i];
    double D = d->efc_D[i], R = d->efc_R[i], x = jar[i];
    if (tp == CNSTR_FRICTION_DOF) {
      double f = d->efc_frictionloss[i];
      if (x <= -R * f) { d->efc_state[i] = STATE_LINEARNEG; d->efc_force[i] = f; cost += -0.5 * R * f * f - f * x; }
      else if (x >= R * f) { d->efc_state[i] = STATE_LINEARPOS; d->efc_force[i] = -f; cost += -0.5 * R * f * f + f * x; }
      else { d->efc_state[i] = STATE_QUADRATIC; d->efc_force[i] = -D * x; cost += 0.5 * D * x * x; }
    } else if (tp != CNSTR_CONTACT_ELLIPTIC) {
      if (x < 0) { d->efc_state[i] = STATE_QUADRATIC; d->efc_force[i] = -D * x; cost += 0.5 * D * x * x; }
      else { d->efc_state[i] = STATE_SATISFIED; d->efc_force[i] = 0; }
    } else {
      Contact& con = d->contact[d->efc_id[i]];
      int dim = con.dim; double mu = con.mu; const double* fr = con.friction;
      double U[6]; U[0] = jar[i] * mu;
      for (int j = 1; j < dim; j++) U[j] = jar[i + j] * fr[j - 1];
      double N = U[0], T = 0;
      for (int j = 1; j < dim; j++) T += U[j] * U[j];
      T = std::sqrt(T);
      int st;
      if (N >= mu * T || (T <= 0 && N >= 0)) {
        st = STATE_SATISFIED; for (int j = 0; j < dim; j++) d->efc_force[i + j] = 0;
      } else if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
        st = STATE_QUADRATIC;
        for (int j = 0; j < dim; j++) { d->efc_force[i + j] = -d->efc_D[i + j] * jar[i + j]; cost += 0.5 * d->efc_D[i + j] * jar[i + j] * jar[i + j]; }
      } else {
        st = STATE_CONE;
        double Dm = d->efc_D[i] / (mu * mu * (1 + mu * mu));
        double NT = N - mu * T;
        cost += 0.5 * Dm * NT * NT;
        d->efc_force[i] = -Dm * NT * mu;
        for (int j = 1; j < dim; j++) d->efc_force[i + j] = -d->efc_force[i] / T * U[j] * fr[j - 1];
        if (coneHessian) {
          double* H = con.H;
          H[0] = 1;
          for (int j = 1; j < dim; j++) H[j] = H[j * dim] = -mu * U[j] / T;
          double s1 = mu * N / (T * T * T), s2 = mu * mu - mu * N / T;
          for (int k = 1; k < dim; k++) for (int j = 1; j < dim; j++) H[k * dim + j] = s1 * U[j] * U[k] + (k == j ? s2 : double(0.0));
          for (int k = 0; k < dim; k++) for (int j = 0; j < dim; j++)
            H[k * dim + j] *= Dm * (k == 0 ? mu : fr[k - 1]) * (j == 0 ? mu : fr[j - 1]);
        }
      }
      for (int j = 0; j < dim; j++) d->efc_state[i + j] = st;
      i += dim - 1;
    }
  }
  return cost;
}

struct Ctx {
  int nv, nefc;
  std::vector<double> Ma, Jaref, grad, Mgrad, search, Mv, Jv, H;
  double cost, gauss, prev_cost;
};

static void updateCost(const Model* m, Data* d, Ctx& c, bool coneH) {
  int nv = c.nv;
  double cc = constraintUpdate(m, d, c.Jaref.data(), coneH);
  double g = 0;
  for (int i = 0; i < nv; i++) g += 0.5 * (c.Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
  c.gauss = g; c.prev_cost = c.cost; c.cost = g + cc;
  // qfrc_constraint = J' force
  for (int i = 0; i < nv; i++) d->qfrc_constraint[i] = 0;
  for (int r = 0; r < c.nefc; r++) { double f = d->efc_force[r]; if (f != 0) for (int i = 0; i < nv; i++) d->qfrc_constraint[i] += d->efc_J[r * nv + i] * f; }
}

// H = M + J' diag(D active) J + cone blocks; Cholesky; Mgrad = H^-1 grad
static void gradientAndDirection(const Model* m, Data* d, Ctx& c) {
  int nv = c.nv, nefc = c.nefc;
  for (int i = 0; i < nv; i++) c.grad[i] = c.Ma[i] - d->qfrc_smooth[i] - d->qfrc_constraint[i];
  double* H = c.H.data(); copy(H, d->qM.data(), nv * nv);
  for (int r = 0; r < nefc; r++) {
    int st = d->efc_state[r];
    if (st == STATE_QUADRATIC) {
      double D = d->efc_D[r]; const double* J = &d->efc_J[r * nv];
      for (int i = 0; i < nv; i++) { if (J[i] == 0) continue; double s = D * J[i]; for (int j = 0; j <= i; j++) H[i * nv + j] += s * J[j]; }
    } else if (st == STATE_CONE && d->efc_type[r] == CNSTR_CONTACT_ELLIPTIC) {
      // cone block J_c' (Hc J_c): T = Hc J_c first, then the symmetric update, one add per cone row
      const Contact& con = d->contact[d->efc_id[r]]; int dim = con.dim;
      double T[6 * 64];
      for (int a = 0; a < dim; a++) for (int j = 0; j < nv; j++) {
        double t = 0; for (int b = 0; b < dim; b++) t += con.H[a * dim + b] * d->efc_J[(r + b) * nv + j];
        T[a * nv + j] = t;
      }
      for (int i = 0; i < nv; i++) for (int j = 0; j <= i; j++)
        for (int a = 0; a < dim; a++) H[i * nv + j] += d->efc_J[(r + a) * nv + i] * T[a * nv + j];
      r += dim - 1;
    }
  }
  // Cholesky (lower), in place
  for (int i = 0; i < nv; i++) {
    for (int j = 0; j <= i; j++) {
      double s = H[i * nv + j];
      for (int k = 0; k < j; k++) s -= H[i * nv + k] * H[j * nv + k];
      if (i == j) H[i * nv + i] = std::sqrt(std::fmax(s, kMinVal));
      else H[i * nv + j] = s / H[j * nv + j];
    }
  }
  copy(c.Mgrad.data(), c.grad.data(), nv);
  for (int i = 0; i < nv; i++) { double s = c.Mgrad[i]; for (int k = 0; k < i; k++) s -= H[i * nv + k] * c.Mgrad[k]; c.Mgrad[i] = s / H[i * nv + i]; }
  for (int i = nv - 1; i >= 0; i--) { double s = c.Mgrad[i]; for (int k = nv - 1; k > i; k--) s -= H[k * nv + i] * c.Mgrad[k]; c.Mgrad[i] = s / H[i * nv + i]; }
  for (int i = 0; i < nv; i++) c.search[i] = -c.Mgrad[i];
}

// cost and its first two derivatives along the search direction at step alpha
// diagnostics (tools/solver_work_survey.py): line-search evaluations since process start; not thread-safe, single-thread use only
long long g_dbg_line_evals = 0;
static void evalLine(const Model* m, const Data* d, const Ctx& c, const double quadGauss[3], double alpha,
                     double* cost, double* d1, double* d2) {
  g_dbg_line_evals++;
  double f = quadGauss[0] + alpha * quadGauss[1] + alpha * alpha * quadGauss[2];
  double g = quadGauss[1] + 2 * alpha * quadGauss[2];
  double h = 2 * quadGauss[2];
  int nefc = c.nefc;
  for (int i = 0; i < nefc; i++) {
    int tp = d->efc_type[i];
    double D = d->efc_D[i], R = d->efc_R[i];
    double x = c.Jaref[i] + alpha * c.Jv[i], v = c.Jv[i];
    if (tp == CNSTR_FRICTION_DOF) {
      double fl = d->efc_frictionloss[i];
      if (x <= -R * fl) { f += -0.5 * R * fl * fl - fl * x; g += -fl * v; }
      else if (x >= R * fl) { f += -0.5 * R * fl * fl + fl * x; g += fl * v; }
      else { f += 0.5 * D * x * x; g += D * x * v; h += D * v * v; }
    } else if (tp != CNSTR_CONTACT_ELLIPTIC) {
      if (x < 0) { f += 0.5 * D * x * x; g += D * x * v; h += D * v * v; }
    } else {
      const Contact& con = d->contact[d->efc_id[i]];
      int dim = con.dim; double mu = con.mu; const double* fr = con.friction;
      double N = x * mu, Np = v * mu, TT = 0, UV = 0, VV = 0;
      for (int j = 1; j < dim; j++) {
        double u = (c.Jaref[i + j] + alpha * c.Jv[i + j]) * fr[j - 1], w = c.Jv[i + j] * fr[j - 1];
        TT += u * u; UV += u * w; VV += w * w;
      }
      double T = std::sqrt(TT);
      if (N >= mu * T || (T <= 0 && N >= 0)) {
      } else if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
        for (int j = 0; j < dim; j++) {
          double xj = c.Jaref[i + j] + alpha * c.Jv[i + j], vj = c.Jv[i + j], Dj = d->efc_D[i + j];
          f += 0.5 * Dj * xj * xj; g += Dj * xj * vj; h += Dj * vj * vj;
        }
      } else {
        double Dm = D / (mu * mu * (1 + mu * mu));
        double Tp = UV / T, Tpp = VV / T - UV * UV / (T * T * T);
        double NT = N - mu * T, NTp = Np - mu * Tp;
        f += 0.5 * Dm * NT * NT; g += Dm * NT * NTp; h += Dm * (NTp * NTp + NT * (-mu * Tpp));
      }
      i += dim - 1;
    }
  }
  *cost = f; *d1 = g; *d2 = h;
}

static double lineSearch(const Model* m, Data* d, Ctx& c) {
  int nv = c.nv, nefc = c.nefc;
  double snorm = std::sqrt(dot(c.search.data(), c.search.data(), nv));
  if (snorm < kMinVal) return 0;
  double scale = 1.0 / (m->stat_meaninertia * std::max(1, nv));
  double gtol = m->opt_tolerance * m->opt_ls_tolerance * snorm / scale;
  mulM(m, d, c.Mv.data(), c.search.data());
  for (int r = 0; r < nefc; r++) c.Jv[r] = dot(&d->efc_J[r * nv], c.search.data(), nv);
  double qg[3] = {c.gauss, 0, 0};
  for (int i = 0; i < nv; i++) { qg[1] += c.search[i] * (c.Ma[i] - d->qfrc_smooth[i]); qg[2] += 0.5 * c.search[i] * c.Mv[i]; }
  double c0, d0, h0; evalLine(m, d, c, qg, 0, &c0, &d0, &h0);
  if (!(h0 > 0) || d0 >= 0) return 0;
  double lo = 0, hi = -1;            // bracket: f'(lo) < 0 ; hi < 0 means "not yet found"
  double alpha = -d0 / h0, cc, dd, hh;
  evalLine(m, d, c, qg, alpha, &cc, &dd, &hh);
  for (int it = 0; it < m->opt_ls_iterations && std::fabs(dd) >= gtol; it++) {
    if (dd < 0) lo = alpha; else hi = alpha;
    double an = hh > 0 ? alpha - dd / hh : double(-1);
    if (hi < 0) { if (!(an > lo)) an = 2 * alpha; }
    else if (!(an > lo && an < hi)) an = 0.5 * (lo + hi);
    alpha = an;
    evalLine(m, d, c, qg, alpha, &cc, &dd, &hh);
  }
  if (!(cc < c0)) return 0;
  return alpha;
}

// mj_fwdConstraint with the Newton solver
void fwdConstraint(const Model* m, Data* d) {
  int nv = m->nv, nefc = d->nefc;
  d->solver_iter = 0;
  if (!nefc) {
    copy(d->qacc.data(), d->qacc_smooth.data(), nv);
    copy(d->qacc_warmstart.data(), d->qacc_smooth.data(), nv);
    zero(d->qfrc_constraint.data(), nv);
    return;
  }
  Ctx c; c.nv = nv; c.nefc = nefc;
  c.Ma.assign(nv, 0); c.Jaref.assign(nefc, 0); c.grad.assign(nv, 0); c.Mgrad.assign(nv, 0); c.search.assign(nv, 0);
  c.Mv.assign(nv, 0); c.Jv.assign(nefc, 0); c.H.assign(nv * nv, 0); c.cost = 0; c.prev_cost = 0;
  bool coneH = m->opt_cone == PS_CONE_ELLIPTIC;
  for (int r = 0; r < nefc; r++) d->efc_b[r] = dot(&d->efc_J[r * nv], d->qacc_smooth.data(), nv) - d->efc_aref[r];
  // warmstart: best of (qacc_warmstart, qacc_smooth)
  if (!(m->opt_disableflags & PS_DSBL_WARMSTART)) {
    copy(d->qacc.data(), d->qacc_warmstart.data(), nv);
    mulM(m, d, c.Ma.data(), d->qacc.data());
    for (int r = 0; r < nefc; r++) c.Jaref[r] = dot(&d->efc_J[r * nv], d->qacc.data(), nv) - d->efc_aref[r];
    double cw = constraintUpdate(m, d, c.Jaref.data(), false);
    for (int i = 0; i < nv; i++) cw += 0.5 * (c.Ma[i] - d->qfrc_smooth[i]) * (d->qacc[i] - d->qacc_smooth[i]);
    double cs = constraintUpdate(m, d, d->efc_b.data(), false);
    if (cw > cs) copy(d->qacc.data(), d->qacc_smooth.data(), nv);
  } else copy(d->qacc.data(), d->qacc_smooth.data(), nv);
  mulM(m, d, c.Ma.data(), d->qacc.data());
  for (int r = 0; r < nefc; r++) c.Jaref[r] = dot(&d->efc_J[r * nv], d->qacc.data(), nv) - d->efc_aref[r];
  updateCost(m, d, c, coneH);
  gradientAndDirection(m, d, c);
  double scale = 1.0 / (m->stat_meaninertia * std::max(1, nv));
  int iter = 0;
  while (iter < m->opt_iterations) {
    double alpha = lineSearch(m, d, c);
    if (alpha == 0) break;
    for (int i = 0; i < nv; i++) { d->qacc[i] += alpha * c.search[i]; c.Ma[i] += alpha * c.Mv[i]; }
    for (int r = 0; r < nefc; r++) c.Jaref[r] += alpha * c.Jv[r];
    updateCost(m, d, c, coneH);
    gradientAndDirection(m, d, c);
    iter++;
    double improvement = scale * (c.prev_cost - c.cost);
    double gradient = scale * std::sqrt(dot(c.grad.data(), c.grad.data(), nv));
    if (improvement < m->opt_tolerance || gradient < m->opt_tolerance) break;
  }
  d->solver_iter = iter;
  copy(d->qacc_warmstart.data(), d->qacc.data(), nv);
}

}  // namespace mm
