// Warp-cooperative rollout engine: ONE WARP advances ONE rollout, with the rollout's whole working
// set resident in shared memory.  This is the engine for the contact-rich models (walker,
// quadruped): per-rollout state there is 15-35 KB, far beyond what a single thread can keep in
// registers, and the thread-per-rollout engine (ps_rollout.cuh) degenerates into DRAM-latency-bound
// local-memory traffic with divergent lanes (profiles/r01_ncu_quadruped_thread_per_rollout.md).
//
// Every stage is written as a sequence of PHASES.  A phase is a parallel-for over independent
// output items (bodies of one tree level, dofs, constraint rows, matrix entries ...); each item is
// computed by one lane with exactly the arithmetic (and summation order) the sequential code uses,
// so results do not depend on the lane count.  tests/hostsim compiles this file with a lane count
// of 1 and diffs it against the oracle; the GPU runs it with 32 lanes.
#pragma once
#include "ps_math.cuh"

namespace ps {

// ---------------------------------------------------------------- lanes
// One working lane: the host build (tests/hostsim) and the thread-per-rollout instantiation of the
// same phases, where every "parallel-for" is an ordinary loop.
struct Lanes1 {
  static constexpr bool single = true, kSharedData = false;
  static constexpr int lane = 0, width = 1;
  PS_HD void sync() const {}
  PS_HD int bcast(int v, int) const { return v; }
  PS_HD double bcast(double v, int) const { return v; }
  PS_HD bool any(bool p) const { return p; }
  PS_HD int scan(int v, int* total) const { *total = v; return 0; }
  PS_HD int min(int v) const { return v; }
  // block-level lockstep hooks (Lanes1Block below): trivial for every other lane type
  static constexpr bool kBlock = false;
  PS_HD bool block_any(bool p) const { return p; }
  PS_HD void block_sync() const {}
};

#ifdef __CUDACC__
// The 32 lanes of a warp working on one rollout held in shared memory.
struct Lanes {
  static constexpr bool single = false, kSharedData = true;
  int lane;
  int width;      // 32; the diagnostics mode (DevModel::debug_serial) runs chosen stages with one working lane
  __device__ __forceinline__ void sync() const { __syncwarp(); }
  __device__ __forceinline__ int bcast(int v, int src) const { return __shfl_sync(0xffffffffu, v, src); }
  __device__ __forceinline__ double bcast(double v, int src) const { return __shfl_sync(0xffffffffu, v, src); }
  __device__ __forceinline__ bool any(bool p) const { return __any_sync(0xffffffffu, p); }
  // exclusive prefix sum over lanes; *total = sum over all lanes
  __device__ __forceinline__ int scan(int v, int* total) const {
    int incl = v; const int pl = threadIdx.x & 31;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (pl >= o) incl += t; }
    *total = __shfl_sync(0xffffffffu, incl, 31);
    return incl - v;
  }
  // smallest value over lanes
  __device__ __forceinline__ int min(int v) const {
    for (int o = 16; o > 0; o >>= 1) { const int t = __shfl_xor_sync(0xffffffffu, v, o); v = t < v ? t : v; }
    return v;
  }
  static constexpr bool kBlock = false;
  __device__ __forceinline__ bool block_any(bool p) const { return p; }
  __device__ __forceinline__ void block_sync() const {}
};
// One working lane per rollout (thread-per-rollout kernel) whose BLOCK walks the time steps, the stages of a step and the
// Newton iterations of the constraint solver in lockstep.  That kernel is bound by instruction delivery (ncu: GPC
// instruction-cache requests at 57 % of peak with 16 warps per SM each somewhere else in ~370 KB of code per step); warps
// that execute the same stage at the same time share the SM's instruction cache lines.  Only control flow changes: a
// rollout whose solver has converged idles at the barrier until the block's slowest one has, arithmetic is untouched.
struct Lanes1Block : Lanes1 {
  static constexpr bool kBlock = true;
  __device__ __forceinline__ bool block_any(bool p) const { return __syncthreads_or(p) != 0; }
  __device__ __forceinline__ void block_sync() const { __syncthreads(); }
};
#else
typedef Lanes1 Lanes;
#endif

// parallel-for over [0, n): item i is handled by lane i mod width
// PS_COMPACT_LOOPS (thread-per-rollout instantiation, ps_variant_legged_sparse.cu): the phase loops are NOT unrolled.  That
// kernel is bound by instruction delivery (ncu: GPC instruction-cache requests at 57 % of peak, `no_instruction` 8 cycles per
// issue -- its ~500 KB of code streams through the SMs' instruction caches), so a smaller loop body beats the ILP of an unrolled one.
#ifdef PS_COMPACT_LOOPS
#define PS_LOOP_PRAGMA _Pragma("unroll 1")
#else
#define PS_LOOP_PRAGMA
#endif
// PS_LOOP2: the plain loops inside the phases (row / dof / contact loops); compacted only at PS_COMPACT_LOOPS >= 2
#if defined(PS_COMPACT_LOOPS) && PS_COMPACT_LOOPS >= 2
#define PS_LOOP2 _Pragma("unroll 1")
#else
#define PS_LOOP2
#endif
#define CO_FOR(i, n) PS_LOOP_PRAGMA for (int i = L.lane; i < (n); i += L.width)
// parallel-for over the pairs (a, k) of [0, A) x [0, K), row-major (w = a K + k).  A single working lane walks them with
// counters -- no integer division per item (the thread-per-rollout kernel spent a visible share of its XU pipe on w / K, w % K)
#define CO_FOR2(w, A, a, K, k)                                                                                     \
  PS_LOOP_PRAGMA for (int w = L.lane, a = LT::single ? 0 : w / (K), k = LT::single ? 0 : w % (K); w < (A) * (K);           \
                      w += L.width, LT::single ? (++k == (K) ? (k = 0, ++a) : 0) : (a = w / (K), k = w % (K)))
// all lanes walk [0, n) in rounds of `width`; lanes past the end still take part in collectives
#define CO_ROUNDS(base, n) PS_LOOP_PRAGMA for (int base = 0; base < (n); base += L.width)

// ---------------------------------------------------------------- working set of one rollout
struct CoopContact {
  double dist, pos[3], frame[9], includemargin, friction[5], mu, solref[2], solimp[5];
  short dim, geom1, geom2, efc_address, ndof, jadr, nrow, pad_;
};

template <class T>
struct alignas(16) CoopData {
  static constexpr int NB = T::NBODY, NV = T::NV, NCH = T::NCH, NJROW = T::NJROW;
  static constexpr int NVEL = (6 * NB + 6 * NV) > 10 * NB ? (6 * NB + 6 * NV) : 10 * NB;
  double time;
  double qpos[T::NQ], qvel[NV], act[T::NA], ctrl[T::NU], qacc[NV], qacc_warmstart[NV], act_dot[T::NA];
  double mocap_pos[3 * T::NMOCAP], mocap_quat[4 * T::NMOCAP];
  // position stage
  double xpos[3 * NB], xquat[4 * NB], xmat[9 * NB], xipos[3 * NB], ximat[9 * NB];
  double xanchor[3 * T::NJNT], xaxis[3 * T::NJNT];
  double site_xpos[3 * T::NSITE], site_xmat[9 * T::NSITE];
  double track_xpos[3 * kMaxTrackGeom], track_xmat[9 * kMaxTrackGeom];
  double subtree_com[3 * NB], cinert[10 * NB], cdof[6 * NV];
  double velbuf[NVEL];                    // crb (position stage), then cvel | cdof_dot (velocity stage)
  double qM[NV * NV], qH[NV * NV], qLDiagInv[NV];   // qH: L'DL of M, then Newton Hessian / its Cholesky factor
  double actuator_length[T::NU], actuator_velocity[T::NU], actuator_force[T::NU];
  double qfrc_bias[NV], qfrc_passive[NV], qfrc_actuator[NV], qfrc_smooth[NV], qacc_smooth[NV], qfrc_constraint[NV];
  double subtree_linvel[3 * NB], subtree_angmom[3 * NB];
  // constraints: rows are friction-loss dofs, joint limits, then contacts (mj_makeConstraint order)
  CoopContact contact[T::MAXCON];
  signed char con_dof[T::MAXCON * NCH];   // dofs a contact's Jacobian rows touch, ascending
  signed char con_slot[T::MAXCON * NV];   // dof -> position in con_dof, or -1
  double jpool[NJROW * NCH];              // contact Jacobian rows, compressed to con_dof columns
  double efc_D[T::MAXEFC], efc_R[T::MAXEFC], efc_aref[T::MAXEFC], efc_force[T::MAXEFC];
  double sv_Jaref[T::MAXEFC], sv_Jv[T::MAXEFC], sv_tmp[T::MAXEFC];
  unsigned char efc_type[T::MAXEFC], efc_state[T::MAXEFC];
  short efc_id[T::MAXEFC], efc_jrow[T::MAXEFC];   // efc_jrow: jpool row (contacts) / sign of the unit row (limits)
  // solver scratch
  double sv_Ma[NV], sv_grad[NV], sv_search[NV], sv_Mv[NV], sv_x[NV];
  double sv_T[6 * NCH], sv_coneH[36];
  double scratch[12 * NB];                // cacc | cfrc (RNE), body velocities (subtree momentum)
  double residual[kMaxResidual];
  double sc_cost, sc_gauss, sc_prev_cost;
  int ncon, nefc, nlimit, solver_iter, warning, pad_;

  PS_HD double* crb() { return velbuf; }
  PS_HD double* cvel() { return velbuf; }
  PS_HD const double* cvel() const { return velbuf; }
  PS_HD double* cdof_dot() { return velbuf + 6 * NB; }
  PS_HD const double* cdof_dot() const { return velbuf + 6 * NB; }
};

// world frame of geom g, from its body frame (the cooperative engine does not keep geom frames)
template <class T, class D>
PS_HD void geom_pose(const DevModel<T>& m, const D& d, int g, double* pos, double* mat) {
  const int b = m.geom_bodyid[g]; double t[3], q[4];
  mat_vec3(t, d.xmat + 9 * b, m.geom_pos + 3 * g);
  for (int k = 0; k < 3; k++) pos[k] = d.xpos[3 * b + k] + t[k];
  if (mat) { quat_mul(q, d.xquat + 4 * b, m.geom_quat + 4 * g); quat_to_mat(mat, q); }
}

// accessors shared by the task residuals: geom world frames of either working-set flavour
template <class T> PS_HD const double* geom_world_pos(const DevModel<T>&, const RolloutData<T>& d, int g) { return d.geom_xpos + 3 * g; }
template <class T> PS_HD const double* geom_world_mat(const DevModel<T>&, const RolloutData<T>& d, int g) { return d.geom_xmat + 9 * g; }
template <class T> PS_HD const double* geom_world_pos(const DevModel<T>& m, const CoopData<T>& d, int g) { return d.track_xpos + 3 * m.geom_trackslot[g]; }
template <class T> PS_HD const double* geom_world_mat(const DevModel<T>& m, const CoopData<T>& d, int g) { return d.track_xmat + 9 * m.geom_trackslot[g]; }
template <class T> PS_HD const double* body_cvel(const RolloutData<T>& d) { return d.cvel; }
template <class T> PS_HD const double* body_cvel(const CoopData<T>& d) { return d.cvel(); }

}  // namespace ps
