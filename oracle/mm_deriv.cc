// ORACLE (test infrastructure).  qDeriv = d(qfrc_smooth)/d(qvel) for the implicit integrators
// (MuJoCo 3.1.4 engine_derivative: mjd_smooth_vel = actuator + passive [- rne] derivatives;
// third-party, not in /root/reference).  Parity vs MuJoCo: UNPINNED.
#include "mm_engine.h"
#include "mm_math.h"

namespace mm {
void inertiaBox(const Model* m, int i, double box[3]);

// d(bias)/d(qvel): bias(v) = c(v) + g with c a homogeneous quadratic form in v, so the exact
// derivative column j is  c(v + e_j) - c(v) - c(e_j)  (polarisation identity, no truncation error).
static void rneVelDerivative(const Model* m, Data* d, double* Dm) {
  int nv = m->nv;
  std::vector<double> g(nv), cv(nv), ce(nv), cve(nv), v(nv), zerov(nv, 0.0);
  rne(m, d, zerov.data(), g.data());
  rne(m, d, d->qvel.data(), cv.data());
  for (int j = 0; j < nv; j++) {
    std::fill(v.begin(), v.end(), 0.0); v[j] = 1; rne(m, d, v.data(), ce.data());
    v = d->qvel; v[j] += 1; rne(m, d, v.data(), cve.data());
    for (int i = 0; i < nv; i++) Dm[i * nv + j] = (cve[i] - g[i]) - (cv[i] - g[i]) - (ce[i] - g[i]);
  }
}

void smoothVelDerivative(const Model* m, Data* d, bool flg_rne) {
  int nv = m->nv;
  double* D = d->qDeriv.data(); zero(D, nv * nv);
  // actuators: d force / d velocity through affine gain/bias
  for (int i = 0; i < m->nu; i++) {
    double bv = 0;
    if (m->actuator_biastype[i] == PS_BIAS_AFFINE) bv += m->actuator_biasprm[3 * i + 2];
    if (m->actuator_gaintype[i] == PS_GAIN_AFFINE) {
      double in = m->actuator_dyntype[i] == PS_DYN_NONE ? d->ctrl[i] : d->act[m->actuator_actadr[i]];
      bv += m->actuator_gainprm[3 * i + 2] * in;
    }
    if (bv != 0) { int dof = m->jnt_dofadr[m->actuator_trnid[2 * i]]; double g = m->actuator_gear[i]; D[dof * nv + dof] += g * g * bv; }
  }
  // passive: dampers
  if (!(m->opt_disableflags & PS_DSBL_PASSIVE)) {
    for (int i = 0; i < nv; i++) D[i * nv + i] -= m->dof_damping[i];
    // inertia-box fluid: B = d lfrc / d lvel (diagonal), qDeriv += J' R B R' J
    if (m->opt_viscosity > 0 || m->opt_density > 0) {
      std::vector<double> jp(3 * nv), jr(3 * nv), Jl(6 * nv);
      for (int b = 1; b < m->nbody; b++) {
        if (m->body_mass[b] < kMinVal) continue;
        double box[3]; inertiaBox(m, b, box);
        // local velocity
        const double* cv = &d->cvel[6 * b]; double dif[3], cr[3], lin[3], lvel[6];
        for (int k = 0; k < 3; k++) dif[k] = d->xipos[3 * b + k] - d->subtree_com[3 * m->body_rootid[b] + k];
        cross(cr, dif, cv); for (int k = 0; k < 3; k++) lin[k] = cv[3 + k] - cr[k];
        mulMatTVec3(lvel, &d->ximat[9 * b], cv); mulMatTVec3(lvel + 3, &d->ximat[9 * b], lin);
        double lw[3]; mulMatTVec3(lw, &d->ximat[9 * b], m->opt_wind);
        for (int k = 0; k < 3; k++) lvel[3 + k] -= lw[k];
        double B[6] = {0, 0, 0, 0, 0, 0}, dens = m->opt_density, visc = m->opt_viscosity;
        if (visc > 0) { double diam = (box[0] + box[1] + box[2]) / 3.0; for (int k = 0; k < 3; k++) { B[k] -= kPi * diam * diam * diam * visc; B[3 + k] -= 3.0 * kPi * diam * visc; } }
        if (dens > 0) {
          B[3] -= dens * box[1] * box[2] * std::fabs(lvel[3]); B[4] -= dens * box[0] * box[2] * std::fabs(lvel[4]); B[5] -= dens * box[0] * box[1] * std::fabs(lvel[5]);
          double b0 = box[0], b1 = box[1], b2 = box[2];
          B[0] -= dens * b0 * (b1 * b1 * b1 * b1 + b2 * b2 * b2 * b2) * 2 * std::fabs(lvel[0]) / 64.0;
          B[1] -= dens * b1 * (b0 * b0 * b0 * b0 + b2 * b2 * b2 * b2) * 2 * std::fabs(lvel[1]) / 64.0;
          B[2] -= dens * b2 * (b0 * b0 * b0 * b0 + b1 * b1 * b1 * b1) * 2 * std::fabs(lvel[2]) / 64.0;
        }
        jac(m, d, jp.data(), jr.data(), &d->xipos[3 * b], b);
        const double* R = &d->ximat[9 * b];
        for (int i = 0; i < nv; i++) for (int r = 0; r < 3; r++) {
          Jl[r * nv + i] = R[r] * jr[i] + R[3 + r] * jr[nv + i] + R[6 + r] * jr[2 * nv + i];
          Jl[(3 + r) * nv + i] = R[r] * jp[i] + R[3 + r] * jp[nv + i] + R[6 + r] * jp[2 * nv + i];
        }
        for (int r = 0; r < 6; r++) for (int i = 0; i < nv; i++) { double s = B[r] * Jl[r * nv + i]; if (s == 0) continue; for (int j = 0; j < nv; j++) D[i * nv + j] += s * Jl[r * nv + j]; }
      }
    }
  }
  if (flg_rne) {
    std::vector<double> Dr(nv * nv);
    rneVelDerivative(m, d, Dr.data());
    for (int i = 0; i < nv * nv; i++) D[i] -= Dr[i];
  }
}

}  // namespace mm
