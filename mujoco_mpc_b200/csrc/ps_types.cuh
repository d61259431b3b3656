// Device-side model / per-rollout data layout of the B200 Predictive Sampling rollout engine.
//
// The reference planner hands MuJoCo an mjModel + one mjData per worker thread
// (mjpc/planners/planner.cc:23-33, mjpc/planners/sampling/planner.cc:151-187).  Here the model
// constants live in ONE fixed-capacity POD (DevModel<T>) that every thread block pulls into
// shared memory with a single TMA bulk copy, and the per-rollout working set (RolloutData<T>) is a
// fixed-capacity struct private to the thread that owns the rollout.  Capacities are compile
// time (Traits) so every array has a constant stride; actual sizes are runtime fields.
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define PS_HD __host__ __device__ __forceinline__
#define PS_FN __host__ __device__ inline
// out-of-line on the device: the rollout kernels are hundreds of KB of SASS when everything is
// inlined, and a warp that walks that much code stalls on instruction fetch
#define PS_NI __host__ __device__ __noinline__ inline
#else
#define PS_HD inline
#define PS_FN inline
#define PS_NI inline
#endif
// stage functions of the thread-per-rollout engines: out of line except in the tiny instantiation,
// whose short steps lose more to call overhead than they gain in instruction-cache footprint
#ifdef PS_THREAD_INLINE
#define PS_NT PS_FN
#else
#define PS_NT PS_NI
#endif
// Model-specialised translation units (ps_variant_tiny_spec.cu) ask for full unrolling of the loops over bodies /
// joints / dofs / cost terms: their trip counts and index tables are compile-time there, so every array index
// becomes a constant and the per-rollout working set is promoted to registers.  Everywhere else the macro is empty
// and the generic kernels compile exactly as before.
#ifdef PS_SPEC_BUILD
#define PS_SPEC_UNROLL _Pragma("unroll")
#else
#define PS_SPEC_UNROLL
#endif
#if defined(__CUDA_ARCH__)
#define PS_ASSUME_SHARED(p) __builtin_assume(__isShared(p))
#else
#define PS_ASSUME_SHARED(p) ((void)0)
#endif

namespace ps {

constexpr double kMinVal = 1e-15;   // mjMINVAL
constexpr double kMaxVal = 1e10;    // mjMAXVAL
constexpr double kPi = 3.14159265358979323846;
constexpr int kMaxTerms = 16;       // cost terms carried on device (the six tasks use <= 9)
constexpr int kMaxResidual = 48;    // quadruped: 42
constexpr int kMaxTaskParam = 16;
constexpr int kMaxTaskAux = 96;
constexpr int kMaxRayGeom = 8;
constexpr int kMaxTrackGeom = 16;

// compile-time capacities of one engine instantiation
template <int NQ_, int NV_, int NU_, int NA_, int NBODY_, int NJNT_, int NGEOM_, int NSITE_,
          int NMOCAP_, int NPAIR_, int MAXCON_, int MAXEFC_, int NSENSOR_, bool ELLIPTIC_, int NCH_ = 0, int NJROW_ = 0>
struct Traits {
  static constexpr int NQ = NQ_, NV = NV_, NU = NU_, NA = (NA_ > 0 ? NA_ : 1), NBODY = NBODY_,
                       NJNT = NJNT_, NGEOM = NGEOM_, NSITE = (NSITE_ > 0 ? NSITE_ : 1),
                       NMOCAP = (NMOCAP_ > 0 ? NMOCAP_ : 1), NPAIR = (NPAIR_ > 0 ? NPAIR_ : 1),
                       MAXCON = (MAXCON_ > 0 ? MAXCON_ : 1), MAXEFC = (MAXEFC_ > 0 ? MAXEFC_ : 1),
                       NSENSOR = (NSENSOR_ > 0 ? NSENSOR_ : 1);
  static constexpr bool ELLIPTIC = ELLIPTIC_;
  static constexpr bool SPEC = false;           // model-specialised instantiation (SpecTraits below)
  // Feature switches derived from the capacities.  CONTACTS: the instantiation carries collision and
  // contact rows at all.  LEAN (the tiny thread-per-rollout engine): nq == nv == the capacity exactly,
  // hinge / slide joints only, no contacts, no activations, no fluid, Euler only -- the other code paths
  // are compiled out and the dof loops have compile-time trip counts, which keeps the per-step
  // instruction stream inside the SM's instruction cache (ncu: sm__icc_request_hit_rate 46 % and
  // `no_instruction` the top stall with everything compiled in).  fill_model<T> rejects models that
  // need a switched-off feature, so they fall through to the next larger instantiation.
  static constexpr bool CONTACTS = MAXCON_ > 0;
  static constexpr bool QUAT = NQ_ != NV_;      // free / ball joints possible at all
  static constexpr bool LEAN = (NV_ <= 2) && (NQ_ == NV_) && (NA_ == 0) && (MAXCON_ == 0) && (NCH_ == 0);
  // cooperative engine: max dofs one contact's Jacobian touches, rows of the compressed Jacobian pool
  static constexpr int NCH = NCH_ > 0 ? NCH_ : NV_, NJROW = NJROW_ > 0 ? NJROW_ : (MAXEFC_ > 0 ? MAXEFC_ : 1);
  static constexpr int NCONEH = ELLIPTIC_ ? MAXCON : 1;
  static constexpr int NTRI = NV_ * (NV_ - 1) / 2 + 1;                  // strictly-lower entries of an NV x NV matrix
  static constexpr int NFPAIR = NV_ * (NV_ + 1) * (NV_ + 2) / 6 + 1;    // worst case (serial chain) of the elimination pairs
};

enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { GEOM_PLANE = 0, GEOM_SPHERE = 2, GEOM_CAPSULE = 3, GEOM_CYLINDER = 5, GEOM_BOX = 6 };
enum { INT_EULER = 0, INT_RK4 = 1, INT_IMPLICIT = 2, INT_IMPLICITFAST = 3 };
enum { CONE_PYRAMIDAL = 0, CONE_ELLIPTIC = 1 };
enum { DSBL_CONSTRAINT = 1, DSBL_FRICTIONLOSS = 4, DSBL_LIMIT = 8, DSBL_CONTACT = 16, DSBL_PASSIVE = 32,
       DSBL_GRAVITY = 64, DSBL_CLAMPCTRL = 128, DSBL_WARMSTART = 256, DSBL_FILTERPARENT = 512,
       DSBL_ACTUATION = 1024, DSBL_REFSAFE = 2048, DSBL_EULERDAMP = 16384 };
enum { DYN_NONE = 0, DYN_INTEGRATOR = 1, DYN_FILTER = 2, DYN_FILTEREXACT = 3 };
enum { SENS_JOINTPOS = 8, SENS_JOINTVEL = 9, SENS_FRAMEPOS = 25, SENS_FRAMEXAXIS = 27, SENS_FRAMEYAXIS = 28,
       SENS_FRAMEZAXIS = 29, SENS_FRAMELINVEL = 30, SENS_FRAMEANGVEL = 31, SENS_SUBTREECOM = 34,
       SENS_SUBTREELINVEL = 35, SENS_SUBTREEANGMOM = 36 };
enum { OBJ_BODY = 1, OBJ_XBODY = 2, OBJ_GEOM = 5, OBJ_SITE = 6 };
enum { EFC_FRICTION_DOF = 0, EFC_LIMIT_JOINT, EFC_CONTACT_FRICTIONLESS, EFC_CONTACT_PYRAMIDAL, EFC_CONTACT_ELLIPTIC };
enum { ST_SATISFIED = 0, ST_QUADRATIC, ST_LINEARNEG, ST_LINEARPOS, ST_CONE };
enum { SPLINE_ZERO = 0, SPLINE_LINEAR = 1, SPLINE_CUBIC = 2 };

// ------------------------------------------------------------------------------------------
// Model constants (shared memory resident).  16-byte aligned, size a multiple of 16 so it can be
// moved with cp.async.bulk.
template <class T>
struct alignas(16) DevModel {
  // options
  double timestep, impratio, tolerance, ls_tolerance, density, viscosity, meaninertia;
  double gravity[3], wind[3];
  int integrator, cone, iterations, ls_iterations, disableflags;
  int nq, nv, nu, na, nbody, njnt, ngeom, nsite, nmocap, nsensor, npair;
  int any_damping;     // some dof has damping > 0 (Euler implicit-damping branch)
  int need_subtreevel; // a subtreelinvel / subtreeangmom sensor is read by the task
  int nray;            // group-0 scenery the Ground() ray cast visits (utilities.cc:538-556)
  int ray_geom[kMaxRayGeom];
  int ray_complete;    // every group-0 geom is static plane/sphere/box scenery and fits ray_geom
  int debug_serial;    // diagnostics: bit k runs stage k of the cooperative engine on one lane
  double qpos0[T::NQ], qpos_spring[T::NQ];
  // bodies
  int body_parentid[T::NBODY], body_rootid[T::NBODY], body_jntnum[T::NBODY], body_jntadr[T::NBODY],
      body_dofnum[T::NBODY], body_dofadr[T::NBODY], body_mocapid[T::NBODY], body_weldid[T::NBODY];
  double body_pos[T::NBODY * 3], body_quat[T::NBODY * 4], body_ipos[T::NBODY * 3], body_iquat[T::NBODY * 4],
      body_mass[T::NBODY], body_subtreemass[T::NBODY], body_inertia[T::NBODY * 3], body_invweight0[T::NBODY * 2];
  // joints
  int jnt_type[T::NJNT], jnt_qposadr[T::NJNT], jnt_dofadr[T::NJNT], jnt_bodyid[T::NJNT], jnt_limited[T::NJNT];
  double jnt_pos[T::NJNT * 3], jnt_axis[T::NJNT * 3], jnt_stiffness[T::NJNT], jnt_range[T::NJNT * 2],
      jnt_margin[T::NJNT], jnt_solref[T::NJNT * 2], jnt_solimp[T::NJNT * 5];
  // dofs
  int dof_bodyid[T::NV], dof_jntid[T::NV], dof_parentid[T::NV];
  double dof_armature[T::NV], dof_damping[T::NV], dof_frictionloss[T::NV], dof_solref[T::NV * 2],
      dof_solimp[T::NV * 5], dof_invweight0[T::NV];
  // geoms
  int geom_type[T::NGEOM], geom_condim[T::NGEOM], geom_bodyid[T::NGEOM], geom_priority[T::NGEOM],
      geom_group[T::NGEOM], geom_used[T::NGEOM];
  double geom_size[T::NGEOM * 3], geom_pos[T::NGEOM * 3], geom_quat[T::NGEOM * 4], geom_friction[T::NGEOM * 3],
      geom_solmix[T::NGEOM], geom_solref[T::NGEOM * 2], geom_solimp[T::NGEOM * 5], geom_margin[T::NGEOM],
      geom_gap[T::NGEOM], geom_rbound[T::NGEOM];
  // statically filtered collision candidates, in the order mj_collision would visit them
  int pair_geom1[T::NPAIR], pair_geom2[T::NPAIR];
  // sites
  int site_bodyid[T::NSITE];
  double site_pos[T::NSITE * 3], site_quat[T::NSITE * 4];
  // actuators (joint transmission)
  int act_dyntype[T::NU], act_gaintype[T::NU], act_biastype[T::NU], act_jntid[T::NU], act_actadr[T::NU],
      act_ctrllimited[T::NU], act_forcelimited[T::NU], act_actlimited[T::NU];
  double act_dynprm[T::NU * 3], act_gainprm[T::NU * 3], act_biasprm[T::NU * 3], act_ctrlrange[T::NU * 2],
      act_forcerange[T::NU * 2], act_actrange[T::NU * 2], act_gear[T::NU];
  // sensors the residuals may read
  int sensor_type[T::NSENSOR], sensor_objtype[T::NSENSOR], sensor_objid[T::NSENSOR];
  // ---- schedules for the warp-cooperative engine (ps_coop*.cuh), built once on the host
  // bodies grouped by tree depth; children of each body in DECREASING index order (the order in
  // which the sequential backward passes fold a child into its parent)
  int nlevel, ndoflevel, nmpair, nfrictionrow, ntrack, ntrack_model;
  short level_adr[T::NBODY + 1], level_body[T::NBODY], child_adr[T::NBODY + 1], child_list[T::NBODY];
  // dofs grouped by depth in the dof tree; per dof: descendants (decreasing), ancestors (parent first)
  short doflevel_adr[T::NV + 1], doflevel_dof[T::NV];
  short dofdesc_adr[T::NV + 1], dofdesc[T::NTRI], dofanc_adr[T::NV + 1], dofanc[T::NTRI];
  // L'DL elimination step k touches entries (i, j), i an ancestor of k, j an ancestor-or-self of i
  short fpair_adr[T::NV + 1]; unsigned char fpair_i[T::NFPAIR], fpair_j[T::NFPAIR];
  // nonzeros of the mass matrix: (i, j) with j an ancestor-or-self of i
  unsigned char mpair_i[T::NTRI + T::NV], mpair_j[T::NTRI + T::NV];
  short frictionrow_dof[T::NV], dof_frictionrow[T::NV];
  // w -> (a, b), a >= b: row-major enumeration of a lower triangle (entry w = a (a + 1) / 2 + b)
  unsigned char tri_a[T::NTRI + T::NV], tri_b[T::NTRI + T::NV];
  // geoms whose world frames the task reads (feet, ray scenery, sensor geoms)
  short track_geom[kMaxTrackGeom], geom_trackslot[T::NGEOM];
};

// ------------------------------------------------------------------------------------------
// Model-specialised instantiation.  `Spec` (csrc/ps_spec_generated.h, written by tools/gen_spec_header.py from a
// registered task's model + cost specification) carries every INTEGER of the model and of the cost specification --
// sizes, topology tables, joint / actuator / sensor types, norm types -- as compile-time constants; the doubles
// (inertias, offsets, gains, weights ...) stay run-time data in shared memory.  DevModel<SpecTraits<Spec>> keeps the
// field names of the generic DevModel, so the stage functions compile unchanged: `m.nbody` is a static constexpr
// and `m.body_parentid[b]` goes through a constexpr table proxy that folds once the loop over b is unrolled.  With
// every index constant the RolloutData of the rollout is promoted to registers and the branches on joint / sensor /
// norm types disappear.  fill_model rejects (falls through to the generic lean kernel) any model whose integers
// differ from the Spec's.
// PS_TAB(name, expr): table proxy whose operator[] is a pure expression in the index (a chain of conditionals written by
// the generator), so it folds to a constant wherever the index is one and never touches memory where it is not
#define PS_TAB(name, expr) struct name##_t { PS_HD constexpr int operator[](int i_) const { return expr; } } name

// Walk the dof tree upwards: i = first, parent[first], ... while i >= 0 (every index is < bound).  Model-specialised
// builds iterate i = bound-1 .. 0 with a compile-time trip count and skip the non-members, so that the walk unrolls and
// every visited index is a constant; the visiting order -- and hence the arithmetic -- is the same.
#ifdef PS_SPEC_BUILD
#define PS_CHAIN(i, first, bound, par) \
  _Pragma("unroll") for (int i = (bound) - 1, nx_##i = (first); i >= 0; i--) if (i == nx_##i && ((nx_##i = par[i]), true))
#else
#define PS_CHAIN(i, first, bound, par) for (int i = (first); i >= 0; i = par[i])
#endif

template <class Spec>
struct SpecTraits : Traits<Spec::NQ, Spec::NV, Spec::NU, 0, Spec::NBODY, Spec::NJNT, Spec::NGEOM, Spec::NSITE, Spec::NMOCAP, 0, 0, 4,
                           Spec::NSENSOR, false> {
  using Topo = Spec;
  using Base = Traits<Spec::NQ, Spec::NV, Spec::NU, 0, Spec::NBODY, Spec::NJNT, Spec::NGEOM, Spec::NSITE, Spec::NMOCAP, 0, 0, 4,
                      Spec::NSENSOR, false>;
  static constexpr bool SPEC = true;
  static_assert(Base::LEAN, "model specialisation is built on the lean instantiation (nv <= 2, no contacts, Euler)");
};

template <class Spec>
struct alignas(16) DevModel<SpecTraits<Spec>> : Spec::ModelInts {
  using T = SpecTraits<Spec>;
  double timestep, impratio, tolerance, ls_tolerance, density, viscosity, meaninertia;
  double gravity[3], wind[3];
  double qpos0[T::NQ], qpos_spring[T::NQ];
  double body_pos[T::NBODY * 3], body_quat[T::NBODY * 4], body_ipos[T::NBODY * 3], body_iquat[T::NBODY * 4],
      body_mass[T::NBODY], body_subtreemass[T::NBODY], body_inertia[T::NBODY * 3], body_invweight0[T::NBODY * 2];
  double jnt_pos[T::NJNT * 3], jnt_axis[T::NJNT * 3], jnt_stiffness[T::NJNT], jnt_range[T::NJNT * 2],
      jnt_margin[T::NJNT], jnt_solref[T::NJNT * 2], jnt_solimp[T::NJNT * 5];
  double dof_armature[T::NV], dof_damping[T::NV], dof_frictionloss[T::NV], dof_solref[T::NV * 2],
      dof_solimp[T::NV * 5], dof_invweight0[T::NV];
  double geom_size[T::NGEOM * 3], geom_pos[T::NGEOM * 3], geom_quat[T::NGEOM * 4], geom_friction[T::NGEOM * 3],
      geom_solmix[T::NGEOM], geom_solref[T::NGEOM * 2], geom_solimp[T::NGEOM * 5], geom_margin[T::NGEOM],
      geom_gap[T::NGEOM], geom_rbound[T::NGEOM];
  double site_pos[T::NSITE * 3], site_quat[T::NSITE * 4];
  double act_dynprm[T::NU * 3], act_gainprm[T::NU * 3], act_biasprm[T::NU * 3], act_ctrlrange[T::NU * 2],
      act_forcerange[T::NU * 2], act_actrange[T::NU * 2], act_gear[T::NU];
};

// Cost specification + residual snapshot (what Task::Residual()/BaseResidualFn::Update hands the
// planner every iteration, mjpc/task.cc:112-128), trimmed to device capacities.
struct alignas(16) DevTask {
  int task_id, num_residual, num_term, num_trace;
  int dim_norm_residual[kMaxTerms], num_norm_parameter[kMaxTerms], norm[kMaxTerms];
  int sensor_id[16];     // SensorByName resolved on the host (mjpc/utilities.cc:166-175)
  int trace_sensor[4];
  int iaux[16];
  double weight[kMaxTerms], norm_parameter[2 * kMaxTerms];
  double risk;
  double parameters[kMaxTaskParam];
  double aux[kMaxTaskAux];
};

// Cost specification as a model-specialised instantiation sees it: the integers (term layout, norm types, sensor ids)
// are the Spec's compile-time constants, the doubles are read from the DevTask in shared memory.  Same member names
// as DevTask, so cost_value / task_residual compile for either.
template <class Spec>
struct SpecTaskView : Spec::TaskInts {
  const double* weight; const double* norm_parameter; const double* parameters; const double* aux; const int* iaux; double risk;
  PS_HD explicit SpecTaskView(const DevTask& t)
      : weight(t.weight), norm_parameter(t.norm_parameter), parameters(t.parameters), aux(t.aux), iaux(t.iaux), risk(t.risk) {}
};

// Per-iteration inputs shared by all rollouts.
template <class T>
struct alignas(16) DevPlan {
  double state[T::NQ + T::NV + T::NA];
  double mocap[7 * T::NMOCAP];
  double time;
  double spline_times[36];   // kMaxSplinePoints-ish cap of the reference GUI (sampling/planner.h:36)
  int num_spline_points, interpolation, horizon, pad_;
  // qacc_warmstart a rollout's first constraint solve starts from.  Zero for planning rollouts (the pinned reading of
  // Trajectory::Rollout, whose per-thread mjData carries whatever that thread rolled last); the closed-loop plant sets it
  // to the value its previous launch ended with (ps_set_warmstart), as one mjData stepped by mj_step carries it over.
  double warmstart[T::NV];
};

// dof counts: compile-time in a LEAN instantiation (loops over dofs unroll, small arrays stay in registers)
template <class T> PS_HD int nv_of(const DevModel<T>& m) { if constexpr (T::LEAN) return T::NV; else return m.nv; }
template <class T> PS_HD int nq_of(const DevModel<T>& m) { if constexpr (T::LEAN) return T::NQ; else return m.nq; }

struct Contact {
  double dist, pos[3], frame[9], includemargin, friction[5], solref[2], solimp[5], mu;
  int dim, geom1, geom2, efc_address;
};

// Working set of ONE rollout (thread-private).
template <class T>
struct RolloutData {
  struct Rows {};                       // see the model-specialised RolloutData below
  PS_HD void bind(Rows&) {}
  double time;
  double qpos[T::NQ], qvel[T::NV], act[T::NA], ctrl[T::NU], qacc[T::NV], qacc_warmstart[T::NV], act_dot[T::NA];
  double mocap_pos[3 * T::NMOCAP], mocap_quat[4 * T::NMOCAP];
  // position stage
  double xpos[3 * T::NBODY], xquat[4 * T::NBODY], xmat[9 * T::NBODY], xipos[3 * T::NBODY], ximat[9 * T::NBODY];
  double xanchor[3 * T::NJNT], xaxis[3 * T::NJNT];
  double geom_xpos[3 * T::NGEOM], geom_xmat[9 * T::NGEOM];
  double site_xpos[3 * T::NSITE], site_xmat[9 * T::NSITE];
  double subtree_com[3 * T::NBODY], cinert[10 * T::NBODY], crb[10 * T::NBODY], cdof[6 * T::NV];
  double qM[T::NV * T::NV], qLD[T::NV * T::NV], qLDiagInv[T::NV];
  double actuator_length[T::NU], actuator_velocity[T::NU], actuator_force[T::NU];
  // velocity stage
  double cvel[6 * T::NBODY], cdof_dot[6 * T::NV];
  double qfrc_bias[T::NV], qfrc_passive[T::NV], qfrc_actuator[T::NV], qfrc_smooth[T::NV], qacc_smooth[T::NV],
      qfrc_constraint[T::NV];
  double subtree_linvel[3 * T::NBODY], subtree_angmom[3 * T::NBODY];
  // constraints
  int ncon, nefc, solver_iter, warning;
  Contact contact[T::MAXCON];
  double coneH[T::NCONEH * 36];
  int efc_type[T::MAXEFC], efc_id[T::MAXEFC], efc_state[T::MAXEFC];
  double efc_J[T::MAXEFC * T::NV], efc_pos[T::MAXEFC], efc_margin[T::MAXEFC], efc_frictionloss[T::MAXEFC],
      efc_D[T::MAXEFC], efc_R[T::MAXEFC], efc_aref[T::MAXEFC], efc_force[T::MAXEFC], efc_KBI[3 * T::MAXEFC];
  // solver scratch
  double sv_Ma[T::NV], sv_grad[T::NV], sv_search[T::NV], sv_Mv[T::NV], sv_H[T::NV * T::NV];
  double sv_Jaref[T::MAXEFC], sv_Jv[T::MAXEFC];
};

// Model-specialised working set.  Everything indexed by body / joint / dof lives in this struct and -- all indices
// being compile-time constants after unrolling -- is promoted to registers.  The constraint rows are the one part whose
// indices are data dependent (a joint-limit row exists only while the joint is near its limit), and a single
// variable-index access would keep the WHOLE struct in local memory; so the row arrays live in a separate object
// (Rows, a cold local-memory block touched only on steps with active rows) and this struct holds pointers to them.
template <class Spec>
struct RolloutData<SpecTraits<Spec>> {
  using T = SpecTraits<Spec>;
  struct Rows {
    int efc_type[T::MAXEFC], efc_id[T::MAXEFC], efc_state[T::MAXEFC];
    double efc_J[T::MAXEFC * T::NV], efc_pos[T::MAXEFC], efc_margin[T::MAXEFC], efc_frictionloss[T::MAXEFC],
        efc_D[T::MAXEFC], efc_R[T::MAXEFC], efc_aref[T::MAXEFC], efc_force[T::MAXEFC], efc_KBI[3 * T::MAXEFC];
    double sv_Jaref[T::MAXEFC], sv_Jv[T::MAXEFC];
  };
  PS_HD void bind(Rows& r) {
    efc_type = r.efc_type; efc_id = r.efc_id; efc_state = r.efc_state; efc_J = r.efc_J; efc_pos = r.efc_pos; efc_margin = r.efc_margin;
    efc_frictionloss = r.efc_frictionloss; efc_D = r.efc_D; efc_R = r.efc_R; efc_aref = r.efc_aref; efc_force = r.efc_force;
    efc_KBI = r.efc_KBI; sv_Jaref = r.sv_Jaref; sv_Jv = r.sv_Jv;
  }
  double time;
  double qpos[T::NQ], qvel[T::NV], act[T::NA], ctrl[T::NU], qacc[T::NV], qacc_warmstart[T::NV], act_dot[T::NA];
  double mocap_pos[3 * T::NMOCAP], mocap_quat[4 * T::NMOCAP];
  double xpos[3 * T::NBODY], xquat[4 * T::NBODY], xmat[9 * T::NBODY], xipos[3 * T::NBODY], ximat[9 * T::NBODY];
  double xanchor[3 * T::NJNT], xaxis[3 * T::NJNT];
  double geom_xpos[3 * T::NGEOM], geom_xmat[9 * T::NGEOM];
  double site_xpos[3 * T::NSITE], site_xmat[9 * T::NSITE];
  double subtree_com[3 * T::NBODY], cinert[10 * T::NBODY], crb[10 * T::NBODY], cdof[6 * T::NV];
  double qM[T::NV * T::NV], qLD[T::NV * T::NV], qLDiagInv[T::NV];
  double actuator_length[T::NU], actuator_velocity[T::NU], actuator_force[T::NU];
  double cvel[6 * T::NBODY], cdof_dot[6 * T::NV];
  double qfrc_bias[T::NV], qfrc_passive[T::NV], qfrc_actuator[T::NV], qfrc_smooth[T::NV], qacc_smooth[T::NV],
      qfrc_constraint[T::NV];
  double subtree_linvel[3 * T::NBODY], subtree_angmom[3 * T::NBODY];
  int ncon, nefc, solver_iter, warning;
  Contact contact[1];
  double coneH[1];
  int *efc_type, *efc_id, *efc_state;
  double *efc_J, *efc_pos, *efc_margin, *efc_frictionloss, *efc_D, *efc_R, *efc_aref, *efc_force, *efc_KBI;
  double sv_Ma[T::NV], sv_grad[T::NV], sv_search[T::NV], sv_Mv[T::NV], sv_H[T::NV * T::NV];
  double *sv_Jaref, *sv_Jv;
};

enum { WARN_BADQPOS = 1, WARN_BADQVEL = 2, WARN_BADQACC = 4, WARN_BADCTRL = 8, WARN_CAPACITY = 16 };

}  // namespace ps
