// Warp-cooperative primal Newton solver with exact line search (mj_solNewton semantics as restated
// in ps_constraint.cuh): per-row work is spread over the lanes, every scalar (costs, step length,
// stopping tests) is recomputed redundantly by all lanes from shared memory in row order, so control
// flow stays warp-uniform and sums keep the sequential order.
#pragma once
#include "ps_coop_constraint.cuh"

namespace ps {

// state, force and cost term of every row for jar = J a - aref; returns the summed cost
template <class T, class LT>
PS_NI double co_update_rows(const LT& L, const DevModel<T>& m, CoopData<T>& d, const double* jar) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nefc = d.nefc;
  if constexpr (LT::single) {
    // ONE working lane: three homogeneous loops in row order (friction-loss rows, limit rows, contacts -- see co_line_eval),
    // the cost summed as the rows go by: the same terms in the same order as the loop over row indices below
    double cost = 0;
    const int nf = (m.disableflags & DSBL_FRICTIONLOSS) ? 0 : m.nfrictionrow;
    const int nl = d.nlimit, ncon = d.ncon;
    PS_LOOP2 for (int i = 0; i < nf; i++) {
      const double D = d.efc_D[i], R = d.efc_R[i], x = jar[i];
      const double f = m.dof_frictionloss[d.efc_id[i]];
      if (x <= -R * f) { d.efc_state[i] = ST_LINEARNEG; d.efc_force[i] = f; cost += -0.5 * R * f * f - f * x; }
      else if (x >= R * f) { d.efc_state[i] = ST_LINEARPOS; d.efc_force[i] = -f; cost += -0.5 * R * f * f + f * x; }
      else { d.efc_state[i] = ST_QUADRATIC; d.efc_force[i] = -D * x; cost += 0.5 * D * x * x; }
    }
    PS_LOOP2 for (int i = nf; i < nf + nl; i++) {
      const double D = d.efc_D[i], x = jar[i];
      if (x < 0) { d.efc_state[i] = ST_QUADRATIC; d.efc_force[i] = -D * x; cost += 0.5 * D * x * x; }
      else { d.efc_state[i] = ST_SATISFIED; d.efc_force[i] = 0; cost += 0.0; }
    }
    PS_LOOP2 for (int ci = 0; ci < ncon; ci++) {
      const CoopContact& con = d.contact[ci];
      const int i = con.efc_address;
      if (i < 0) continue;
      if (d.efc_type[i] != EFC_CONTACT_ELLIPTIC) {
        const int nrow = con.nrow;
        PS_LOOP2 for (int q = 0; q < nrow; q++) {
          const double D = d.efc_D[i + q], x = jar[i + q];
          if (x < 0) { d.efc_state[i + q] = ST_QUADRATIC; d.efc_force[i + q] = -D * x; cost += 0.5 * D * x * x; }
          else { d.efc_state[i + q] = ST_SATISFIED; d.efc_force[i + q] = 0; cost += 0.0; }
        }
        continue;
      }
      const int dim = con.dim; const double mu = con.mu; const double* fr = con.friction;
      double U[6]; U[0] = jar[i] * mu;
      PS_LOOP2 for (int j = 1; j < dim; j++) U[j] = jar[i + j] * fr[j - 1];
      const double N = U[0]; double TT = 0;
      PS_LOOP2 for (int j = 1; j < dim; j++) TT += U[j] * U[j];
      const double Tn = sqrt(TT);
      int st;
      if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
        st = ST_SATISFIED; PS_LOOP2 for (int j = 0; j < dim; j++) { d.efc_force[i + j] = 0; cost += 0.0; }
      } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
        st = ST_QUADRATIC;
        PS_LOOP2 for (int j = 0; j < dim; j++) { d.efc_force[i + j] = -d.efc_D[i + j] * jar[i + j]; cost += 0.5 * d.efc_D[i + j] * jar[i + j] * jar[i + j]; }
      } else {
        st = ST_CONE;
        const double Dm = d.efc_D[i] / (mu * mu * (1 + mu * mu));
        const double NT = N - mu * Tn;
        cost += 0.5 * Dm * NT * NT;
        d.efc_force[i] = -Dm * NT * mu;
        PS_LOOP2 for (int j = 1; j < dim; j++) { d.efc_force[i + j] = -d.efc_force[i] / Tn * U[j] * fr[j - 1]; cost += 0.0; }
      }
      PS_LOOP2 for (int j = 0; j < dim; j++) d.efc_state[i + j] = (unsigned char)st;
    }
    return cost;
  }
  CO_FOR(i, nefc) {
    const int tp = d.efc_type[i];
    const double D = d.efc_D[i], R = d.efc_R[i], x = jar[i];
    if (tp == EFC_FRICTION_DOF) {
      const double f = m.dof_frictionloss[d.efc_id[i]];
      if (x <= -R * f) { d.efc_state[i] = ST_LINEARNEG; d.efc_force[i] = f; d.sv_tmp[i] = -0.5 * R * f * f - f * x; }
      else if (x >= R * f) { d.efc_state[i] = ST_LINEARPOS; d.efc_force[i] = -f; d.sv_tmp[i] = -0.5 * R * f * f + f * x; }
      else { d.efc_state[i] = ST_QUADRATIC; d.efc_force[i] = -D * x; d.sv_tmp[i] = 0.5 * D * x * x; }
    } else if (tp != EFC_CONTACT_ELLIPTIC) {
      if (x < 0) { d.efc_state[i] = ST_QUADRATIC; d.efc_force[i] = -D * x; d.sv_tmp[i] = 0.5 * D * x * x; }
      else { d.efc_state[i] = ST_SATISFIED; d.efc_force[i] = 0; d.sv_tmp[i] = 0; }
    } else {
      const CoopContact& con = d.contact[d.efc_id[i]];
      if (i != con.efc_address) continue;              // the contact's first row does the whole cone
      const int dim = con.dim; const double mu = con.mu; const double* fr = con.friction;
      double U[6]; U[0] = jar[i] * mu;
      PS_LOOP2 for (int j = 1; j < dim; j++) U[j] = jar[i + j] * fr[j - 1];
      const double N = U[0]; double TT = 0;
      PS_LOOP2 for (int j = 1; j < dim; j++) TT += U[j] * U[j];
      const double Tn = sqrt(TT);
      int st;
      if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
        st = ST_SATISFIED; PS_LOOP2 for (int j = 0; j < dim; j++) { d.efc_force[i + j] = 0; d.sv_tmp[i + j] = 0; }
      } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
        st = ST_QUADRATIC;
        PS_LOOP2 for (int j = 0; j < dim; j++) { d.efc_force[i + j] = -d.efc_D[i + j] * jar[i + j]; d.sv_tmp[i + j] = 0.5 * d.efc_D[i + j] * jar[i + j] * jar[i + j]; }
      } else {
        st = ST_CONE;
        const double Dm = d.efc_D[i] / (mu * mu * (1 + mu * mu));
        const double NT = N - mu * Tn;
        d.sv_tmp[i] = 0.5 * Dm * NT * NT;
        d.efc_force[i] = -Dm * NT * mu;
        PS_LOOP2 for (int j = 1; j < dim; j++) { d.efc_force[i + j] = -d.efc_force[i] / Tn * U[j] * fr[j - 1]; d.sv_tmp[i + j] = 0; }
      }
      PS_LOOP2 for (int j = 0; j < dim; j++) d.efc_state[i + j] = (unsigned char)st;
    }
  }
  L.sync();
  double cost = 0;
  PS_LOOP2 for (int i = 0; i < nefc; i++) cost += d.sv_tmp[i];
  return cost;
}

struct CoSolverScalars { double cost, gauss, prev_cost; };

template <class T, class LT>
PS_NI void co_update_cost(const LT& L, const DevModel<T>& m, CoopData<T>& d, CoSolverScalars& c) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nv = m.nv;
  const double cc = co_update_rows(L, m, d, d.sv_Jaref);
  double g = 0;
  PS_LOOP2 for (int i = 0; i < nv; i++) g += 0.5 * (d.sv_Ma[i] - d.qfrc_smooth[i]) * (d.qacc[i] - d.qacc_smooth[i]);
  c.gauss = g; c.prev_cost = c.cost; c.cost = g + cc;
  // qfrc_constraint = J' f, one lane per dof, rows in order
  const int nf = (m.disableflags & DSBL_FRICTIONLOSS) ? 0 : m.nfrictionrow;
  const int nl = d.nlimit, ncon = d.ncon;
  if constexpr (LT::single) {
    // ONE working lane: unit rows dof by dof, then the contacts one after the other, each scattering into the dofs it touches.
    // A dof's sum still receives its friction row, the limit rows and the contacts' rows in that order (contacts and their
    // rows ascending), so it is the sum of the dof-major form below; a contact without force is skipped after nrow loads.
    PS_LOOP2 for (int i = 0; i < nv; i++) {
      double s = 0;
      const int fr = nf ? m.dof_frictionrow[i] : -1;
      if (fr >= 0) { const double f = d.efc_force[fr]; if (f != 0) s += f; }
      d.qfrc_constraint[i] = s;
    }
    PS_LOOP2 for (int r = nf; r < nf + nl; r++) {        // limit rows in order, each into the dof of its joint
      const double f = d.efc_force[r];
      if (f != 0) d.qfrc_constraint[m.jnt_dofadr[d.efc_id[r]]] += (double)d.efc_jrow[r] * f;
    }
    constexpr int MR = 6;
    PS_LOOP2 for (int ci = 0; ci < ncon; ci++) {
      const CoopContact& con = d.contact[ci];
      const int adr = con.efc_address, nrow = con.nrow;
      if (adr < 0) continue;
      const double* J = d.jpool + con.jadr * T::NCH;
      const signed char* dof = d.con_dof + ci * T::NCH;
      const int n = con.ndof;
      if (nrow <= MR) {
        double fq[MR]; unsigned nz = 0;
#pragma unroll
        for (int q = 0; q < MR; q++) if (q < nrow) { fq[q] = d.efc_force[adr + q]; if (fq[q] != 0) nz |= 1u << q; }
        if (!nz) continue;
        PS_LOOP2 for (int k = 0; k < n; k++) {
          double* dst = d.qfrc_constraint + dof[k];
          double s = *dst;
#pragma unroll
          for (int q = 0; q < MR; q++) if (nz >> q & 1u) s += J[q * T::NCH + k] * fq[q];
          *dst = s;
        }
      } else {
        PS_LOOP2 for (int k = 0; k < n; k++) {
          double* dst = d.qfrc_constraint + dof[k];
          double s = *dst;
          PS_LOOP2 for (int q = 0; q < nrow; q++) { const double f = d.efc_force[adr + q]; if (f != 0) s += J[q * T::NCH + k] * f; }
          *dst = s;
        }
      }
    }
    return;
  }
  CO_FOR(i, nv) {
    double s = 0;
    const int fr = nf ? m.dof_frictionrow[i] : -1;
    if (fr >= 0) { const double f = d.efc_force[fr]; if (f != 0) s += f; }
    PS_LOOP2 for (int r = nf; r < nf + nl; r++) {
      const double f = d.efc_force[r];
      if (f != 0 && m.jnt_dofadr[d.efc_id[r]] == i) s += (double)d.efc_jrow[r] * f;
    }
    PS_LOOP2 for (int ci = 0; ci < ncon; ci++) {
      const CoopContact& con = d.contact[ci];
      const int k = d.con_slot[ci * T::NV + i];
      if (con.efc_address < 0 || k < 0) continue;
      PS_LOOP2 for (int q = 0; q < con.nrow; q++) {
        const double f = d.efc_force[con.efc_address + q];
        if (f != 0) s += d.jpool[(con.jadr + q) * T::NCH + k] * f;
      }
    }
    d.qfrc_constraint[i] = s;
  }
  L.sync();
}

// grad = M a - qfrc_smooth - J' f
template <class T, class LT>
PS_NI void co_newton_gradient(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  CO_FOR(i, m.nv) d.sv_grad[i] = d.sv_Ma[i] - d.qfrc_smooth[i] - d.qfrc_constraint[i];
  L.sync();
}

// One contact in the cone state, ONE working lane (see co_newton_direction): local cone Hessian C (DIM x DIM) in registers,
// T = C J column by column, then the contact's block of H dof-slot row by row.  Products and summation orders are those of the
// cooperative form in co_newton_direction (b, a ascending), entry by entry.
template <int DIM, class T>
PS_FN void co_cone_block_single(CoopData<T>& d, const CoopContact& con, int adr, int n, const double* J, const signed char* dof, double* H) {
  constexpr int S = T::NV, NCH = T::NCH;
  const double mu = con.mu; const double* fr = con.friction;
  double U[DIM]; U[0] = d.sv_Jaref[adr] * mu;
#pragma unroll
  for (int j = 1; j < DIM; j++) U[j] = d.sv_Jaref[adr + j] * fr[j - 1];
  const double N = U[0]; double TT = 0;
#pragma unroll
  for (int j = 1; j < DIM; j++) TT += U[j] * U[j];
  const double Tn = sqrt(TT);
  const double Dm = d.efc_D[adr] / (mu * mu * (1 + mu * mu));
  const double s1 = mu * N / (Tn * Tn * Tn), s2 = mu * mu - mu * N / Tn;
  double hC[DIM * DIM];
#pragma unroll
  for (int k = 0; k < DIM; k++)
#pragma unroll
    for (int j = 0; j < DIM; j++) {
      double h;
      if (k == 0 && j == 0) h = 1;
      else if (k == 0 || j == 0) h = -mu * U[k + j] / Tn;
      else h = s1 * U[j] * U[k] + (k == j ? s2 : 0.0);
      hC[k * DIM + j] = h * (Dm * (k == 0 ? mu : fr[k - 1]) * (j == 0 ? mu : fr[j - 1]));
    }
  PS_LOOP2 for (int k = 0; k < n; k++) {
    double jc[DIM];
#pragma unroll
    for (int b = 0; b < DIM; b++) jc[b] = J[b * NCH + k];
#pragma unroll
    for (int a = 0; a < DIM; a++) {
      double t = 0;
#pragma unroll
      for (int b = 0; b < DIM; b++) t += hC[a * DIM + b] * jc[b];
      d.sv_T[a * NCH + k] = t;
    }
  }
  PS_LOOP2 for (int ki = 0; ki < n; ki++) {
    double jc[DIM];
#pragma unroll
    for (int a = 0; a < DIM; a++) jc[a] = J[a * NCH + ki];
    double* Hrow = H + dof[ki] * S;
    PS_LOOP2 for (int kj = 0; kj <= ki; kj++) {
      double tc[DIM];
#pragma unroll
      for (int a = 0; a < DIM; a++) tc[a] = d.sv_T[a * NCH + kj];
      double* h = Hrow + dof[kj];
      double acc = *h;
#pragma unroll
      for (int a = 0; a < DIM; a++) acc += jc[a] * tc[a];
      *h = acc;
    }
  }
}

// H = M + J' D_active J (+ cone blocks) -> Cholesky -> search = -H^-1 grad (grad: co_newton_gradient)
template <class T, class LT>
PS_NI void co_newton_direction(const LT& L, const DevModel<T>& m, CoopData<T>& d) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  constexpr int S = T::NV, NCH = T::NCH; const int nv = m.nv;
  double* H = d.qH;
  if constexpr (LT::single) {                  // nothing reads the upper triangle; a row's loads are in flight together
    // at the class capacity H already holds M: the M v that precedes every direction stored it (co_mul_M, co_H_prefilled) --
    // ncu had 8.7 % of the kernel's stall samples on this copy's loads, M having been evicted since the product read it
    if (!co_H_prefilled<T, LT>(m)) PS_LOOP2 for (int i = 0; i < nv; i++) {
      double r[S];
#pragma unroll
      for (int k = 0; k < S; k++) if (k <= i) r[k] = d.qM[i * S + k];
#pragma unroll
      for (int k = 0; k < S; k++) if (k <= i) H[i * S + k] = r[k];
    }
  }
  else CO_FOR(w, nv * S) H[w] = d.qM[w];
  L.sync();
  // unit rows: friction rows hit distinct diagonal entries; limit rows go through one lane, in order
  const int nf = (m.disableflags & DSBL_FRICTIONLOSS) ? 0 : m.nfrictionrow;
  CO_FOR(k, nf) if (d.efc_state[k] == ST_QUADRATIC) { const int i = d.efc_id[k]; H[i * S + i] += d.efc_D[k]; }
  L.sync();
  if (L.lane == 0)
    PS_LOOP2 for (int r = nf; r < nf + d.nlimit; r++) if (d.efc_state[r] == ST_QUADRATIC) { const int i = m.jnt_dofadr[d.efc_id[r]]; H[i * S + i] += d.efc_D[r]; }
  L.sync();
  // contacts one after the other; lanes own (dof slot, dof slot) pairs of the contact's block
  PS_LOOP2 for (int ci = 0; ci < d.ncon; ci++) {
    const CoopContact& con = d.contact[ci];
    const int adr = con.efc_address;
    if (adr < 0) continue;
    const int n = con.ndof, npair = n * (n + 1) / 2, nrow = con.nrow;
    const signed char* dof = d.con_dof + ci * NCH;
    const double* J = d.jpool + con.jadr * NCH;
    const bool cone = d.efc_type[adr] == EFC_CONTACT_ELLIPTIC && d.efc_state[adr] == ST_CONE;
    if constexpr (LT::single) {
      // ONE working lane: the contact's block is accumulated dof-slot row by dof-slot row with that row's Jacobian column
      // (times D, or the cone's T column) in registers.  Every H entry receives the very products of the pair loop below in
      // the same order -- (D J_ki) J_kj for q ascending, J_a,ki T_a,kj for a ascending -- so the sums are bit-identical; what
      // changes is the traffic to thread-private memory (ncu: this block was 40 % of the kernel's executed instructions,
      // 4 loads per row and pair): 1 load per row and pair, nothing at all for a contact with no active row.
      constexpr int MR = 6;                                  // rows kept in registers (elliptic: dim <= 6; pyramidal condim 3: 4)
      if (cone && con.dim == 3) { co_cone_block_single<3>(d, con, adr, n, J, dof, H); continue; }
      if (cone && con.dim == 6) { co_cone_block_single<6>(d, con, adr, n, J, dof, H); continue; }
      if (!cone && nrow <= MR) {
        double Dq[MR]; unsigned quad = 0;
#pragma unroll
        for (int q = 0; q < MR; q++) if (q < nrow && d.efc_state[adr + q] == ST_QUADRATIC) { quad |= 1u << q; Dq[q] = d.efc_D[adr + q]; }
        if (!quad) continue;                                  // no active row: the block adds nothing
        PS_LOOP2 for (int ki = 0; ki < n; ki++) {
          double DJ[MR];
#pragma unroll
          for (int q = 0; q < MR; q++) if (quad >> q & 1u) DJ[q] = Dq[q] * J[q * NCH + ki];
          double* Hrow = H + dof[ki] * S;
          PS_LOOP2 for (int kj = 0; kj <= ki; kj++) {
            double* h = Hrow + dof[kj];
            double acc = *h;
#pragma unroll
            for (int q = 0; q < MR; q++) if (quad >> q & 1u) acc += DJ[q] * J[q * NCH + kj];
            *h = acc;
          }
        }
        continue;
      }
    }
    if (cone) {
      const int dim = con.dim; const double mu = con.mu; const double* fr = con.friction;
      // local cone Hessian (dim x dim), recomputed from the row residuals
      double U[6]; U[0] = d.sv_Jaref[adr] * mu;
      PS_LOOP2 for (int j = 1; j < dim; j++) U[j] = d.sv_Jaref[adr + j] * fr[j - 1];
      const double N = U[0]; double TT = 0;
      PS_LOOP2 for (int j = 1; j < dim; j++) TT += U[j] * U[j];
      const double Tn = sqrt(TT);
      const double Dm = d.efc_D[adr] / (mu * mu * (1 + mu * mu));
      const double s1 = mu * N / (Tn * Tn * Tn), s2 = mu * mu - mu * N / Tn;
      CO_FOR2(w, dim, k, dim, j) {
        double h;
        if (k == 0 && j == 0) h = 1;
        else if (k == 0 || j == 0) h = -mu * U[k + j] / Tn;
        else h = s1 * U[j] * U[k] + (k == j ? s2 : 0.0);
        d.sv_coneH[w] = h * (Dm * (k == 0 ? mu : fr[k - 1]) * (j == 0 ? mu : fr[j - 1]));
      }
      L.sync();
      CO_FOR2(w, dim, a, n, k) {
        double t = 0;
        PS_LOOP2 for (int b = 0; b < dim; b++) t += d.sv_coneH[a * dim + b] * J[b * NCH + k];
        d.sv_T[a * NCH + k] = t;
      }
      L.sync();
    }
    CO_FOR(w, npair) {
      const int ki = m.tri_a[w], kj = m.tri_b[w];     // ki >= kj
      double* h = H + dof[ki] * S + dof[kj];
      double acc = *h;
      if (cone) { PS_LOOP2 for (int a = 0; a < con.dim; a++) acc += J[a * NCH + ki] * d.sv_T[a * NCH + kj]; }
      else PS_LOOP2 for (int q = 0; q < nrow; q++) if (d.efc_state[adr + q] == ST_QUADRATIC) acc += d.efc_D[adr + q] * J[q * NCH + ki] * J[q * NCH + kj];
      *h = acc;
    }
    L.sync();
  }
  if constexpr (LT::single) {
    // ONE working lane (thread-per-rollout kernel, host build): row-by-row Cholesky with the current row in registers.
    // Entry (i, k) = (H_ik - sum_{j<k} L_ij L_kj) / L_kk with j ascending -- the very operation sequence of the column
    // formulation below -- but row i is loaded once (up to 18 independent loads in flight), the finished rows are only
    // read, and nothing is stored until the row is complete: a third of the loads, a sixth of the stores, and no
    // load-after-store serialisation.  That kernel is bound by the latency of its thread-private memory (ncu:
    // long_scoreboard 21 cycles per issue, the trailing update of the column formulation alone 25 % of the stall samples).
    double r[S];
    PS_LOOP2 for (int i = 0; i < nv; i++) {
      const double* Hi = H + i * S;
#pragma unroll
      for (int k = 0; k < S; k++) if (k <= i) r[k] = Hi[k];
#pragma unroll
      for (int k = 0; k < S; k++) {
        if (k <= i) {
          const double* Lk = H + k * S;
          double acc = r[k];
#pragma unroll
          for (int j = 0; j < k; j++) acc -= r[j] * (k == i ? r[j] : Lk[j]);
          r[k] = (k == i) ? sqrt(fmax(acc, kMinVal)) : acc / Lk[k];
        }
      }
#pragma unroll
      for (int k = 0; k < S; k++) if (k <= i) H[i * S + k] = r[k];
    }
    // forward / backward substitution by columns, the vector in registers
#pragma unroll
    for (int i = 0; i < S; i++) r[i] = i < nv ? d.sv_grad[i] : 0.0;
#pragma unroll
    for (int k = 0; k < S; k++) {
      if (k < nv) {
        const double xk = r[k] / H[k * S + k];
        r[k] = xk;
#pragma unroll
        for (int i = k + 1; i < S; i++) if (i < nv) r[i] -= H[i * S + k] * xk;
      }
    }
#pragma unroll
    for (int k = S - 1; k >= 0; k--) {
      if (k < nv) {
        const double xk = r[k] / H[k * S + k];
        r[k] = xk;
#pragma unroll
        for (int i = 0; i < k; i++) r[i] -= H[k * S + i] * xk;
      }
    }
#pragma unroll
    for (int i = 0; i < S; i++) if (i < nv) d.sv_search[i] = -r[i];
    return;
  }
  // right-looking Cholesky: scale column j, then subtract its outer product from the trailing
  // triangle.  Entry (i, k) receives the products of columns 0 .. k-1 in that order, which is the
  // summation order of the row-by-row formulation (s -= L_ik' L_kk', k' ascending).
  PS_LOOP2 for (int j = 0; j < nv; j++) {
    const double dg = sqrt(fmax(H[j * S + j], kMinVal));
    L.sync();
    PS_LOOP2 for (int base = j; base < nv; base += L.width) {
      const int i = base + L.lane;
      if (i == j) H[j * S + j] = dg; else if (i < nv) H[i * S + j] = H[i * S + j] / dg;
    }
    L.sync();
    const int n = nv - 1 - j;
    CO_FOR(w, n * (n + 1) / 2) {
      const int i = j + 1 + m.tri_a[w], k = j + 1 + m.tri_b[w];
      H[i * S + k] -= H[i * S + j] * H[k * S + j];
    }
    L.sync();
  }
  // forward / backward substitution by columns
  double* x = d.sv_x;
  CO_FOR(i, nv) x[i] = d.sv_grad[i];
  L.sync();
  PS_LOOP2 for (int k = 0; k < nv; k++) {
    const double xk = x[k] / H[k * S + k];
    L.sync();
    if (L.lane == 0) x[k] = xk;
    PS_LOOP2 for (int base = k + 1; base < nv; base += L.width) { const int i = base + L.lane; if (i < nv) x[i] -= H[i * S + k] * xk; }
    L.sync();
  }
  PS_LOOP2 for (int k = nv - 1; k >= 0; k--) {
    const double xk = x[k] / H[k * S + k];
    L.sync();
    if (L.lane == 0) x[k] = xk;
    CO_FOR(i, k) x[i] -= H[k * S + i] * xk;
    L.sync();
  }
  CO_FOR(i, nv) d.sv_search[i] = -x[i];
  L.sync();
}

// value, slope and curvature of the total cost along the search direction at step alpha
template <class T, class LT>
PS_NI void co_line_eval(const LT& L, const DevModel<T>& m, CoopData<T>& d, const double* qg, double alpha,
                        double* cost, double* d1, double* d2) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nefc = d.nefc;
  if constexpr (LT::single) {
    // ONE working lane: the rows' terms go straight into the three sums, in row order (a cone's trailing rows add their
    // zeros like the two-pass form below does) -- no per-row stores and no second pass over them.  The rows are walked
    // as three homogeneous loops -- friction-loss rows, limit rows, contacts: that IS the row order (co_make_constraint) --
    // so that the threads of a warp run the same code at the same time; one loop over row indices with a branch on the row
    // type had them at 5 of 32 lanes (ncu), because rollouts with different limit counts hold different row types at one index.
    double f = qg[0] + alpha * qg[1] + alpha * alpha * qg[2], g = qg[1] + 2 * alpha * qg[2], h = 2 * qg[2];
    const int nf = (m.disableflags & DSBL_FRICTIONLOSS) ? 0 : m.nfrictionrow;
    const int nl = d.nlimit, ncon = d.ncon;
    PS_LOOP2 for (int i = 0; i < nf; i++) {
      const double D = d.efc_D[i], v = d.sv_Jv[i], x = d.sv_Jaref[i] + alpha * v;
      const double R = d.efc_R[i], fl = m.dof_frictionloss[d.efc_id[i]];
      if (x <= -R * fl) { f += -0.5 * R * fl * fl - fl * x; g += -fl * v; h += 0.0; }
      else if (x >= R * fl) { f += -0.5 * R * fl * fl + fl * x; g += fl * v; h += 0.0; }
      else { f += 0.5 * D * x * x; g += D * x * v; h += D * v * v; }
    }
    PS_LOOP2 for (int i = nf; i < nf + nl; i++) {
      const double D = d.efc_D[i], v = d.sv_Jv[i], x = d.sv_Jaref[i] + alpha * v;
      if (x < 0) { f += 0.5 * D * x * x; g += D * x * v; h += D * v * v; }
      else { f += 0.0; g += 0.0; h += 0.0; }
    }
    PS_LOOP2 for (int ci = 0; ci < ncon; ci++) {
      const CoopContact& con = d.contact[ci];
      const int i = con.efc_address;
      if (i < 0) continue;
      if (d.efc_type[i] != EFC_CONTACT_ELLIPTIC) {          // frictionless / pyramidal rows: plain inequalities
        const int nrow = con.nrow;
        PS_LOOP2 for (int q = 0; q < nrow; q++) {
          const double D = d.efc_D[i + q], v = d.sv_Jv[i + q], x = d.sv_Jaref[i + q] + alpha * v;
          if (x < 0) { f += 0.5 * D * x * x; g += D * x * v; h += D * v * v; }
          else { f += 0.0; g += 0.0; h += 0.0; }
        }
        continue;
      }
      const double D = d.efc_D[i], v = d.sv_Jv[i], x = d.sv_Jaref[i] + alpha * v;
      const int dim = con.dim; const double mu = con.mu; const double* fr = con.friction;
      const double N = x * mu, Np = v * mu; double TT = 0, UV = 0, VV = 0;
      PS_LOOP2 for (int j = 1; j < dim; j++) {
        const double u = (d.sv_Jaref[i + j] + alpha * d.sv_Jv[i + j]) * fr[j - 1], w = d.sv_Jv[i + j] * fr[j - 1];
        TT += u * u; UV += u * w; VV += w * w;
      }
      const double Tn = sqrt(TT);
      if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
        PS_LOOP2 for (int j = 0; j < dim; j++) { f += 0.0; g += 0.0; h += 0.0; }
      } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
        PS_LOOP2 for (int j = 0; j < dim; j++) {
          const double xj = d.sv_Jaref[i + j] + alpha * d.sv_Jv[i + j], vj = d.sv_Jv[i + j], Dj = d.efc_D[i + j];
          f += 0.5 * Dj * xj * xj; g += Dj * xj * vj; h += Dj * vj * vj;
        }
      } else {
        const double Dm = D / (mu * mu * (1 + mu * mu));
        const double Tp = UV / Tn, Tpp = VV / Tn - UV * UV / (Tn * Tn * Tn);
        const double NT = N - mu * Tn, NTp = Np - mu * Tp;
        f += 0.5 * Dm * NT * NT; g += Dm * NT * NTp; h += Dm * (NTp * NTp + NT * (-mu * Tpp));
        PS_LOOP2 for (int j = 1; j < dim; j++) { f += 0.0; g += 0.0; h += 0.0; }
      }
    }
    *cost = f; *d1 = g; *d2 = h;
    return;
  }
  double* tf = d.sv_tmp; double* tg = d.efc_force; double* th = d.efc_aref;     // both dead during the line search
  CO_FOR(i, nefc) {
    const int tp = d.efc_type[i];
    const double D = d.efc_D[i], R = d.efc_R[i];
    const double x = d.sv_Jaref[i] + alpha * d.sv_Jv[i], v = d.sv_Jv[i];
    if (tp == EFC_FRICTION_DOF) {
      const double fl = m.dof_frictionloss[d.efc_id[i]];
      if (x <= -R * fl) { tf[i] = -0.5 * R * fl * fl - fl * x; tg[i] = -fl * v; th[i] = 0; }
      else if (x >= R * fl) { tf[i] = -0.5 * R * fl * fl + fl * x; tg[i] = fl * v; th[i] = 0; }
      else { tf[i] = 0.5 * D * x * x; tg[i] = D * x * v; th[i] = D * v * v; }
    } else if (tp != EFC_CONTACT_ELLIPTIC) {
      if (x < 0) { tf[i] = 0.5 * D * x * x; tg[i] = D * x * v; th[i] = D * v * v; }
      else { tf[i] = 0; tg[i] = 0; th[i] = 0; }
    } else {
      const CoopContact& con = d.contact[d.efc_id[i]];
      if (i != con.efc_address) continue;
      const int dim = con.dim; const double mu = con.mu; const double* fr = con.friction;
      const double N = x * mu, Np = v * mu; double TT = 0, UV = 0, VV = 0;
      PS_LOOP2 for (int j = 1; j < dim; j++) {
        const double u = (d.sv_Jaref[i + j] + alpha * d.sv_Jv[i + j]) * fr[j - 1], w = d.sv_Jv[i + j] * fr[j - 1];
        TT += u * u; UV += u * w; VV += w * w;
      }
      const double Tn = sqrt(TT);
      PS_LOOP2 for (int j = 0; j < dim; j++) { tf[i + j] = 0; tg[i + j] = 0; th[i + j] = 0; }
      if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
      } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
        PS_LOOP2 for (int j = 0; j < dim; j++) {
          const double xj = d.sv_Jaref[i + j] + alpha * d.sv_Jv[i + j], vj = d.sv_Jv[i + j], Dj = d.efc_D[i + j];
          tf[i + j] = 0.5 * Dj * xj * xj; tg[i + j] = Dj * xj * vj; th[i + j] = Dj * vj * vj;
        }
      } else {
        const double Dm = D / (mu * mu * (1 + mu * mu));
        const double Tp = UV / Tn, Tpp = VV / Tn - UV * UV / (Tn * Tn * Tn);
        const double NT = N - mu * Tn, NTp = Np - mu * Tp;
        tf[i] = 0.5 * Dm * NT * NT; tg[i] = Dm * NT * NTp; th[i] = Dm * (NTp * NTp + NT * (-mu * Tpp));
      }
    }
  }
  L.sync();
  double f, g, h;
  if (LT::single || L.width == 1) {       // one working lane: plain sequential sums
    f = qg[0] + alpha * qg[1] + alpha * alpha * qg[2]; g = qg[1] + 2 * alpha * qg[2]; h = 2 * qg[2];
    PS_LOOP2 for (int i = 0; i < nefc; i++) { f += tf[i]; g += tg[i]; h += th[i]; }
  } else {                                   // three lanes add up one quantity each, in row order, then broadcast
    double acc = L.lane == 0 ? qg[0] + alpha * qg[1] + alpha * alpha * qg[2] : (L.lane == 1 ? qg[1] + 2 * alpha * qg[2] : 2 * qg[2]);
    const double* src = L.lane == 0 ? tf : (L.lane == 1 ? tg : th);
    if (L.lane < 3) PS_LOOP2 for (int i = 0; i < nefc; i++) acc += src[i];
    f = L.bcast(acc, 0); g = L.bcast(acc, 1); h = L.bcast(acc, 2);
  }
  L.sync();
  *cost = f; *d1 = g; *d2 = h;
}

// `active`: see co_constraint_stage (a thread without work passes through).  mj_solNewton's exact line search:
// Newton steps on the 1-D cost, bracketed, until |slope| < gtol or ls_iterations.
template <class T, class LT>
PS_NI double co_line_search(const LT& L, const DevModel<T>& m, CoopData<T>& d, const CoSolverScalars& c, bool active = true) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  const int nv = m.nv, nefc = active ? d.nefc : 0;
  double qg[3] = {c.gauss, 0, 0};
  double gtol = 0, c0 = 0, lo = 0, hi = -1, alpha = 0, cc = 0, dd = 0, hh = 0;
  bool go = false;
  if (active) {
    const double snorm = sqrt(vdot(d.sv_search, d.sv_search, nv));
    if (snorm < kMinVal) active = false;
    else {
      const double scale = 1.0 / (m.meaninertia * (nv > 1 ? nv : 1));
      gtol = m.tolerance * m.ls_tolerance * snorm / scale;
      co_mul_M(L, m, d, d.sv_Mv, d.sv_search, co_H_prefilled<T, LT>(m) ? d.qH : nullptr);     // + H := M for the direction that follows
      CO_FOR(r, nefc) d.sv_Jv[r] = co_row_dot(m, d, r, d.sv_search);
      L.sync();
      PS_LOOP2 for (int i = 0; i < nv; i++) { qg[1] += d.sv_search[i] * (d.sv_Ma[i] - d.qfrc_smooth[i]); qg[2] += 0.5 * d.sv_search[i] * d.sv_Mv[i]; }
      double d0, h0; co_line_eval(L, m, d, qg, 0, &c0, &d0, &h0);
      if (!(h0 > 0) || d0 >= 0) active = false;
      else {
        alpha = -d0 / h0;
        co_line_eval(L, m, d, qg, alpha, &cc, &dd, &hh);
        go = 0 < m.ls_iterations && fabs(dd) >= gtol;
      }
    }
  }
  int it = 0;
  // (under block lockstep the line search runs free: its code is small, and a barrier per line-search iteration cost 4.5 %)
  while (go) {
    if (go) {
      if (dd < 0) lo = alpha; else hi = alpha;
      double an = hh > 0 ? alpha - dd / hh : -1;
      if (hi < 0) { if (!(an > lo)) an = 2 * alpha; }
      else if (!(an > lo && an < hi)) an = 0.5 * (lo + hi);
      alpha = an;
      co_line_eval(L, m, d, qg, alpha, &cc, &dd, &hh);
      it++;
      go = it < m.ls_iterations && fabs(dd) >= gtol;
    }
  }
  if (!active || !(cc < c0)) return 0;
  return alpha;
}

template <class T, class LT>
PS_NI void co_constraint_stage(const LT& L, const DevModel<T>& m, CoopData<T>& d, bool active = true) { PS_ASSUME_SHARED(&m); if (LT::kSharedData) PS_ASSUME_SHARED(&d);
  // `active`: this rollout is still alive.  Under block lockstep (LT::kBlock) every thread of the block walks the Newton
  // loop together -- block_any() is a block-wide OR -- and a thread without work idles; for every other lane type
  // block_any(p) == p and the loop below is the plain `while (iter < iterations) { ... break ... }` of mj_solNewton.
  const int nv = m.nv, nefc = active ? d.nefc : 0;
  const bool run = active && nefc > 0;
  if (active && !nefc) {
    CO_FOR(i, nv) { d.qacc[i] = d.qacc_smooth[i]; d.qacc_warmstart[i] = d.qacc_smooth[i]; d.qfrc_constraint[i] = 0; }
    if (L.lane == 0) d.solver_iter = 0;
    L.sync();
  }
  CoSolverScalars c; c.cost = 0; c.prev_cost = 0; c.gauss = 0;
  if (run) {
    double* const Hcopy = co_H_prefilled<T, LT>(m) ? d.qH : nullptr;                 // + H := M for the first direction (co_mul_M)
    if (!(m.disableflags & DSBL_WARMSTART)) {
      // keep the warm start only if it beats the unconstrained acceleration
      CO_FOR(i, nv) d.qacc[i] = d.qacc_warmstart[i];
      L.sync();
      co_mul_M(L, m, d, d.sv_Ma, d.qacc, Hcopy);
      CO_FOR(r, nefc) d.sv_Jaref[r] = co_row_dot(m, d, r, d.qacc) - d.efc_aref[r];
      L.sync();
      double cw = co_update_rows(L, m, d, d.sv_Jaref);
      PS_LOOP2 for (int i = 0; i < nv; i++) cw += 0.5 * (d.sv_Ma[i] - d.qfrc_smooth[i]) * (d.qacc[i] - d.qacc_smooth[i]);
      L.sync();
      CO_FOR(r, nefc) d.sv_Jv[r] = co_row_dot(m, d, r, d.qacc_smooth) - d.efc_aref[r];
      L.sync();
      const double cs = co_update_rows(L, m, d, d.sv_Jv);
      L.sync();
      // M qacc and J qacc - aref of the chosen start: the warm start's are in sv_Ma / sv_Jaref already (computed above, not
      // touched since); the unconstrained acceleration's J a - aref is sv_Jv.  mj_solNewton computes both again -- the same
      // operations on the same operands, so the same bits -- at the price of one more pass over M and over every Jacobian row.
      if (cw > cs) {
        CO_FOR(i, nv) d.qacc[i] = d.qacc_smooth[i];
        CO_FOR(r, nefc) d.sv_Jaref[r] = d.sv_Jv[r];
        L.sync();
        co_mul_M(L, m, d, d.sv_Ma, d.qacc);
      }
      L.sync();
    } else {
      CO_FOR(i, nv) d.qacc[i] = d.qacc_smooth[i];
      L.sync();
      co_mul_M(L, m, d, d.sv_Ma, d.qacc, Hcopy);
      CO_FOR(r, nefc) d.sv_Jaref[r] = co_row_dot(m, d, r, d.qacc) - d.efc_aref[r];
      L.sync();
    }
    co_update_cost(L, m, d, c);
    co_newton_gradient(L, m, d);
    co_newton_direction(L, m, d);
  }
  const double scale = 1.0 / (m.meaninertia * (nv > 1 ? nv : 1));
  int iter = 0;
  bool more = run && iter < m.iterations;
  while (L.block_any(more)) {
    const double alpha = co_line_search(L, m, d, c, more);
    if (more && alpha == 0) more = false;
    if (more) {
      CO_FOR(i, nv) { d.qacc[i] += alpha * d.sv_search[i]; d.sv_Ma[i] += alpha * d.sv_Mv[i]; }
      CO_FOR(r, nefc) d.sv_Jaref[r] += alpha * d.sv_Jv[r];
      L.sync();
      co_update_cost(L, m, d, c);
    }
    L.block_sync();
    if (more) {
      // mj_solNewton factorises the Hessian and solves for the next direction before it tests for convergence; the
      // direction of the last iteration is never used, so the test comes first here and the last factorisation is skipped
      co_newton_gradient(L, m, d);
      iter++;
      const double improvement = scale * (c.prev_cost - c.cost);
      const double gradient = scale * sqrt(vdot(d.sv_grad, d.sv_grad, nv));
      if (improvement < m.tolerance || gradient < m.tolerance || iter >= m.iterations) more = false;
    }
    if (more) co_newton_direction(L, m, d);
  }
  if (run) {
    CO_FOR(i, nv) d.qacc_warmstart[i] = d.qacc[i];
    if (L.lane == 0) d.solver_iter = iter;
    L.sync();
  }
}

}  // namespace ps
