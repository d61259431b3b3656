// C ABI of the B200 Predictive Sampling rollout engine (include/ps_engine.h): host-side staging,
// kernel launches and read-back.  No CPU fallback exists: every entry point drives the CUDA
// kernels in ps_kernels.cuh, and ps_create() fails without a device.
//
// Boundary it replaces: the body of SamplingPlanner::OptimizePolicyCandidates / Rollouts
// (mjpc/planners/sampling/planner.cc:151-187,342-380) -- N ThreadPool closures each calling
// Trajectory::Rollout on a per-thread mjData -- becomes K1 + K2 + K3 on one stream.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <dlfcn.h>
#include <nccl.h>
#include "../../include/ps_engine.h"
#include "ps_variant.h"
#include "ps_kernels_aux.cuh"

using namespace ps;

static thread_local std::string g_error;
static int fail(const std::string& s) { g_error = s; return -1; }
#define CU(call)                                                                                   \
  do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

// Engine families, tried in order.  A family with two kernels shares one Constants layout: `small`
// (warp-per-rollout, shared-memory working set) serves batches up to `crossover` rollouts -- the
// planning-latency regime -- and `large` (thread-per-rollout over the same phases) the big batches.
struct Family { const Variant* small; const Variant* large; int crossover; bool large_required; };
static const Family kFamilies[] = {
    // model-specialised lean kernels of the registered tiny tasks: taken only when every integer of the model matches
    // the compiled-in description (else the next family); the generic lean kernel stays beside them for cost
    // specifications the specialisation was not compiled for and for the dense flavour
#define PS_SPEC_FAMILY(N) {variant_tiny(), variant_spec_##N(), 0, true},
    PS_SPEC_LIST(PS_SPEC_FAMILY)
#undef PS_SPEC_FAMILY
    {variant_tiny_coop(), variant_tiny(), 0},                 // nv ~ 2: the lean thread-per-rollout kernel wins at every batch size (1.4 vs 2.3 ms at N <= 1024);
                                                              // the warp kernel serves small models that need a feature the lean kernel compiles out
    {variant_tiny(), nullptr, 0},
    {variant_planar_coop(), variant_planar(), 2560},         // nv ~ 9: dense thread-per-rollout wins the big batches
    {variant_planar(), nullptr, 0},
    {variant_legged_coop(), variant_legged_sparse(), 6144},  // nv ~ 18: compressed-Jacobian thread-per-rollout (quadruped, 1 s horizon: warp kernel
                                                             // 617 / 1193 ms at 4096 / 8192 rollouts, thread kernel 816 / 1042 ms: profiles/r02_crossover_legged.txt)
    {variant_legged(), nullptr, 0},
};

// the cooperative engine carries semi-implicit Euler and no fluid forces
static bool coop_supported(const PsModel* m, std::string* why) {
  if (m->opt_integrator != PS_INT_EULER) { *why = "integrator is not Euler"; return false; }
  if (m->opt_density > 0 || m->opt_viscosity > 0) { *why = "fluid forces"; return false; }
  return true;
}

// ---------------------------------------------------------------- engine object
struct PsEngine_ {
  int device = 0; cudaStream_t stream = nullptr;
  const Variant* var = nullptr;
  void* h_const = nullptr; void* d_const = nullptr;             // pinned host mirror + device copy
  void* h_const2 = nullptr; void* d_const2 = nullptr;           // same, in the layout of the large-batch kernel
  int nq = 0, nv = 0, nu = 0, na = 0, nmocap = 0, nsensor = 0, ns = 0;
  std::vector<int> sensor_type; bool ray_ok = false; int* d_counter = nullptr;
  const Variant* var_large = nullptr; const Variant* last = nullptr; int crossover = 0; void* last_dconst = nullptr;
  bool large_ok = true;      // the large-batch kernel accepts the current cost specification (model-specialised kernels)
  int engine_override = 0;   // PS_ENGINE: 0 auto, 1 warp, 2 thread (compressed phases), 3 dense thread-per-rollout
  std::vector<double> ctrlrange;
  DevTask task{}; bool have_task = false, have_model = false;
  int P = 0, interp = 0, H = 0, N = 0, cap = 0, stride = 0, nelem = 0, nblocks = 0;
  long long offset = 0;
  double *d_nominal = nullptr, *d_ctrlrange = nullptr, *d_knots = nullptr, *d_returns = nullptr, *d_record = nullptr, *d_z = nullptr;
  unsigned char *d_failure = nullptr, *d_std2 = nullptr; MinLoc* d_blockmin = nullptr;
  int* d_fix = nullptr;                                           // [0] = count, [1 ..] = flagged samples (capacity-overflow fix-up)
  double* h_record = nullptr;                                     // pinned
  size_t knots_cap = 0, z_cap = 0; bool have_noise = false;
  unsigned long long seed = 1234, iteration = 0;
  // writeback buffers (parity / BestTrajectory), sized on demand
  Writeback wb{}; size_t wb_cap[7] = {0, 0, 0, 0, 0, 0, 0}; int wb_N = 0, wb_H = 0; bool wb_valid = false;
  double sigma0 = 0, sigma1 = 0;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  long long launches = 0;
  double* d_scratch = nullptr; size_t scratch_cap = 0;
  // cross-entropy mode: per-parameter noise std, elite ranks, update record
  double* d_param_std = nullptr; bool have_param_std = false; int* d_elite = nullptr; int elite_cap = 0;
  double* d_ce_out = nullptr; double* h_ce_out = nullptr; float ce_ms = 0;
  double* d_xfrc = nullptr; size_t xfrc_cap = 0;     // robust rollouts: standard-normal force noise [(t * 6 nbody + k)][stride]
  int nbody = 0; double timestep = 0; int integrator = 0;
  // host staging: the pinned mirrors are rewritten by the setters, so every H2D copy out of them is followed by
  // an event the next writer waits on (an async caller can never upload torn constants)
  cudaEvent_t ev_h2d = nullptr; bool h2d_pending = false; bool const_dirty = true, const2_dirty = true;
  double* h_nominal = nullptr;                                    // pinned staging of the nominal knots
  bool record_fresh = false;                                      // h_record holds the last iteration's record
  size_t last_upload = 0;
  // multi-GPU winner exchange (ps_comm_init): one ncclAllGather of the winner records + a min-loc kernel
  ncclComm_t comm = nullptr; int rank = 0, world = 1; double* d_gather = nullptr; double* d_global = nullptr; cudaEvent_t ev_x = nullptr;
};

// ---------------------------------------------------------------- NCCL, bound at run time
// The library does not link NCCL: a single-GPU user needs none, and a multi-GPU host process (torch, MPI launcher)
// normally has one loaded already -- RTLD_NOLOAD picks that copy up so that two NCCL versions never meet.
struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
  bool ok = false; std::string why;
};
static NcclApi& nccl_api() {
  static NcclApi api; static bool tried = false;
  if (tried) return api;
  tried = true;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { api.why = std::string("libnccl.so.2 not found: ") + dlerror(); return api; }
#define PS_NCCL_SYM(field, name) *(void**)(&api.field) = dlsym(h, name); if (!api.field) { api.why = std::string("missing symbol ") + name; return api; }
  PS_NCCL_SYM(GetUniqueId, "ncclGetUniqueId") PS_NCCL_SYM(CommInitRank, "ncclCommInitRank") PS_NCCL_SYM(CommDestroy, "ncclCommDestroy")
  PS_NCCL_SYM(AllGather, "ncclAllGather") PS_NCCL_SYM(GetErrorString, "ncclGetErrorString") PS_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef PS_NCCL_SYM
  api.ok = true;
  return api;
}
#define NC(call)                                                                                   \
  do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) return fail(std::string(#call) + ": " + nccl_api().GetErrorString(r_)); } while (0)

// the pinned mirrors may only be rewritten once the last copy out of them has completed
static void wait_upload(PsEngine_* e) {
  if (e->h2d_pending) { cudaEventSynchronize(e->ev_h2d); e->h2d_pending = false; }
}

static int ensure(double** p, size_t* cap, size_t need) {
  if (*cap >= need) return 0;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  CU(cudaMalloc(p, need * sizeof(double)));
  *cap = need; return 0;
}

extern "C" {

const char* ps_last_error(void) { return g_error.c_str(); }
/* other translation units of the library (ps_noalloc.cu) report through the same per-thread text */
void ps_set_last_error_(const char* s) { g_error = s ? s : ""; }
int ps_name(char* out, int cap) { std::snprintf(out, cap, "mujoco_mpc_b200 predictive-sampling engine (sm_100a)"); return 0; }

PsEngine* ps_create(int device_ordinal) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= device_ordinal) {
    g_error = "ps_create: no CUDA device " + std::to_string(device_ordinal) + " (this engine has no CPU path)"; return nullptr;
  }
  if (cudaSetDevice(device_ordinal) != cudaSuccess) { g_error = "cudaSetDevice failed"; return nullptr; }
  PsEngine* e = new PsEngine_;
  e->device = device_ordinal;
  if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) { g_error = "cudaStreamCreate failed"; delete e; return nullptr; }
  for (auto& ev : e->ev) cudaEventCreate(&ev);
  cudaEventCreateWithFlags(&e->ev_h2d, cudaEventDisableTiming); cudaEventCreate(&e->ev_x);
  cudaMallocHost(&e->h_nominal, 36 * 32 * sizeof(double));
  cudaMallocHost(&e->h_record, (3 + 36 * 32) * sizeof(double));
  cudaMalloc(&e->d_record, (3 + 36 * 32) * sizeof(double));
  cudaMalloc(&e->d_counter, sizeof(int));
  if (const char* env = std::getenv("PS_ENGINE")) {              // diagnostics / benchmarking: pin one kernel flavour
    const std::string v(env);
    e->engine_override = v == "warp" ? 1 : v == "thread" ? 2 : v == "dense" ? 3 : 0;
  }
  return e;
}

void ps_destroy(PsEngine* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  cudaStreamSynchronize(e->stream);
  void* bufs[] = {e->d_const, e->d_nominal, e->d_ctrlrange, e->d_knots, e->d_returns, e->d_record, e->d_z, e->d_failure, e->d_std2,
                  e->d_blockmin, e->d_counter, e->d_const2, e->wb.states, e->wb.actions, e->wb.residual, e->wb.costs, e->wb.times, e->wb.trace, e->wb.warmstart, e->d_scratch,
                  e->d_xfrc, e->d_param_std, e->d_elite, e->d_ce_out, e->d_fix};
  for (void* b : bufs) if (b) cudaFree(b);
  if (e->comm && nccl_api().ok) nccl_api().CommDestroy(e->comm);
  if (e->d_gather) cudaFree(e->d_gather);
  if (e->d_global) cudaFree(e->d_global);
  if (e->ev_h2d) cudaEventDestroy(e->ev_h2d);
  if (e->ev_x) cudaEventDestroy(e->ev_x);
  if (e->h_nominal) cudaFreeHost(e->h_nominal);
  if (e->h_const) cudaFreeHost(e->h_const);
  if (e->h_const2) cudaFreeHost(e->h_const2);
  if (e->h_record) cudaFreeHost(e->h_record);
  if (e->h_ce_out) cudaFreeHost(e->h_ce_out);
  for (auto& ev : e->ev) if (ev) cudaEventDestroy(ev);
  cudaStreamDestroy(e->stream);
  delete e;
}

int ps_set_model(PsEngine* e, const PsModel* m) {
  if (!e || !m) return fail("ps_set_model: null argument");
  CU(cudaSetDevice(e->device));
  wait_upload(e); e->const_dirty = e->const2_dirty = true;
  PairList pl = static_pairs(m);
  std::string why, all;
  const Variant* chosen = nullptr; void* host = nullptr; const Family* fam = nullptr;
  static const bool no_spec = std::getenv("PS_NO_SPEC") != nullptr;      // diagnostics: A/B against the generic lean kernel
  for (const Family& f : kFamilies) {
    const Variant& v = *f.small;
    if (f.large_required) {
      if (no_spec || e->engine_override == 1) continue;
      std::vector<char> probe(f.large->const_bytes + 128);
      void* pa = (void*)(((uintptr_t)probe.data() + 127) & ~(uintptr_t)127);
      if (f.large->fill(m, pl, pa, &why) != 0) { all += std::string(f.large->name) + ": " + why + "; "; continue; }
    }
    if (v.coop && (e->engine_override == 3 || !coop_supported(m, &why))) { all += std::string(v.name) + ": " + (e->engine_override == 3 ? "disabled" : why) + "; "; continue; }
    void* h = nullptr; CU(cudaMallocHost(&h, v.const_bytes)); std::memset(h, 0, v.const_bytes);
    if (v.fill(m, pl, h, &why) == 0) { chosen = &v; host = h; fam = &f; break; }
    all += std::string(v.name) + ": " + why + "; "; cudaFreeHost(h);
  }
  if (!chosen) return fail("ps_set_model: model outside the supported class (" + all + ")");
  if (e->h_const) cudaFreeHost(e->h_const);
  if (e->d_const) cudaFree(e->d_const);
  if (e->h_const2) cudaFreeHost(e->h_const2);
  if (e->d_const2) cudaFree(e->d_const2);
  e->h_const2 = nullptr; e->d_const2 = nullptr;
  e->var = chosen; e->h_const = host; e->var_large = nullptr; e->crossover = fam->crossover; e->last = chosen;
  CU(cudaMalloc(&e->d_const, chosen->const_bytes));
  const Variant* large = fam->large;
  // diagnostics / tuning: PS_PLANAR_LARGE=sparse serves the planar class's big batches with the single-lane phase-parallel
  // coding (the quadruped's large-batch kernel at planar capacities) instead of the dense thread-per-rollout kernel
  if (large == variant_planar() && std::getenv("PS_PLANAR_LARGE") && std::string(std::getenv("PS_PLANAR_LARGE")) == "sparse") large = variant_planar_sparse();
  if (large) {
    void* h2 = nullptr; CU(cudaMallocHost(&h2, large->const_bytes)); std::memset(h2, 0, large->const_bytes);
    if (large->fill(m, pl, h2, &why) == 0) { e->var_large = large; e->h_const2 = h2; CU(cudaMalloc(&e->d_const2, large->const_bytes)); }
    else cudaFreeHost(h2);     // the model fits only the small-batch kernel's capacities: it serves every batch size
  }
  e->nbody = m->nbody; e->timestep = m->opt_timestep; e->integrator = m->opt_integrator;
  e->nq = m->nq; e->nv = m->nv; e->nu = m->nu; e->na = m->na; e->nmocap = m->nmocap; e->nsensor = m->nsensor; e->ns = m->nq + m->nv + m->na;
  e->sensor_type.assign(m->sensor_type, m->sensor_type + m->nsensor);
  e->ray_ok = ray_scenery_ok(m);
  e->ctrlrange.assign(m->actuator_ctrlrange, m->actuator_ctrlrange + 2 * m->nu);
  if (e->d_ctrlrange) cudaFree(e->d_ctrlrange);
  CU(cudaMalloc(&e->d_ctrlrange, sizeof(double) * 2 * (m->nu > 0 ? m->nu : 1)));
  CU(cudaMemcpy(e->d_ctrlrange, e->ctrlrange.data(), sizeof(double) * 2 * m->nu, cudaMemcpyHostToDevice));
  // default mocap poses = body pos/quat (mj_resetData)
  std::vector<double> mocap(7 * (m->nmocap > 0 ? m->nmocap : 1), 0.0);
  for (int b = 0; b < m->nbody; b++) if (m->body_mocapid[b] >= 0) {
    const int id = m->body_mocapid[b];
    for (int k = 0; k < 3; k++) mocap[7 * id + k] = m->body_pos[3 * b + k];
    for (int k = 0; k < 4; k++) mocap[7 * id + 3 + k] = m->body_quat[4 * b + k];
  }
  std::vector<double> st(e->ns, 0.0); for (int i = 0; i < m->nq; i++) st[i] = m->qpos0[i];
  chosen->plan(e->h_const, st.data(), e->ns, mocap.data(), m->nmocap, 0.0, nullptr, 0, 0, 0);
  if (e->var_large) e->var_large->plan(e->h_const2, st.data(), e->ns, mocap.data(), m->nmocap, 0.0, nullptr, 0, 0, 0);
  if (const char* dbg = std::getenv("PS_COOP_SERIAL")) {
    chosen->debug(e->h_const, std::atoi(dbg));
    if (e->var_large && e->var_large->debug) e->var_large->debug(e->h_const2, std::atoi(dbg));
  }
  e->have_model = true; e->have_task = false; e->wb_valid = false; e->large_ok = true;
  return 0;
}

int ps_set_options(PsEngine* e, double timestep, int integrator) {
  if (!e || !e->have_model) return fail("ps_set_options: no model");
  if ((e->var->euler_only || (e->var_large && e->var_large->euler_only)) && integrator != PS_INT_EULER)
    return fail("ps_set_options: this model runs on an Euler-only kernel instantiation; call ps_set_model with the new integrator");
  if (integrator == PS_INT_RK4) return fail("ps_set_options: RK4 integrator is not supported");
  if (timestep == e->timestep && integrator == e->integrator) return 0;      // unchanged: nothing to re-upload
  wait_upload(e); e->const_dirty = e->const2_dirty = true;
  e->timestep = timestep; e->integrator = integrator;
  e->var->options(e->h_const, timestep, integrator);
  if (e->var_large) e->var_large->options(e->h_const2, timestep, integrator);
  return 0;
}

int ps_set_task(PsEngine* e, const PsTask* t) {
  if (!e || !e->have_model) return fail("ps_set_task: set the model first");
  DevTask d; int need_sub = 0; std::string why;
  if (convert_task(t, e->sensor_type.data(), e->nsensor, &d, &need_sub, &why)) return fail("ps_set_task: " + why);
  if (d.task_id == TASK_QUADRUPED) {
    if (!e->ray_ok) return fail("ps_set_task: Quadruped needs static plane/sphere/box group-0 scenery for its ground ray casts");
    if (e->nq != 19 || e->nu != 12) return fail("ps_set_task: Quadruped expects a free joint + 12 actuated joints");
    for (int k = 0; k < 3; k++) if (t->sensor_id[k] < 0) return fail("ps_set_task: Quadruped sensor ids (subtreecom, subtreelinvel, angmom) missing");
  }
  wait_upload(e); e->const_dirty = e->const2_dirty = true;
  e->task = d; e->have_task = true;
  e->large_ok = !(e->var_large && e->var_large->task_ok) || e->var_large->task_ok(d, need_sub) != 0;
  e->var->task(e->h_const, d, need_sub);
  if (e->var_large) e->var_large->task(e->h_const2, d, need_sub);
  return 0;
}

int ps_set_state(PsEngine* e, const double* state, const double* mocap, const double* userdata, double time) {
  (void)userdata;   // none of the supported tasks read userdata on the planning path
  if (!e || !e->have_model) return fail("ps_set_state: no model");
  wait_upload(e);
  e->var->plan(e->h_const, state, e->ns, mocap, e->nmocap, time, nullptr, 0, 0, 0);
  if (e->var_large) e->var_large->plan(e->h_const2, state, e->ns, mocap, e->nmocap, time, nullptr, 0, 0, 0);
  return 0;
}

int ps_set_nominal(PsEngine* e, const double* times, const double* knots, int P, int interpolation) {
  if (!e || !e->have_model) return fail("ps_set_nominal: no model");
  if (P < 1 || P > 36) return fail("ps_set_nominal: spline points must be in [1, 36]");
  CU(cudaSetDevice(e->device));
  const double nan = std::nan("");
  wait_upload(e);
  e->var->plan(e->h_const, nullptr, 0, nullptr, 0, nan, times, P, interpolation, 0);
  if (e->var_large) e->var_large->plan(e->h_const2, nullptr, 0, nullptr, 0, nan, times, P, interpolation, 0);
  e->P = P; e->interp = interpolation; e->nelem = P * e->nu;
  if (e->nelem > 36 * 32) return fail("ps_set_nominal: more than 1152 spline parameters");
  if (!e->d_nominal) CU(cudaMalloc(&e->d_nominal, sizeof(double) * 36 * 32));
  // knots is caller-owned pageable memory: stage it in the pinned mirror, no stream synchronisation
  std::memcpy(e->h_nominal, knots, sizeof(double) * e->nelem);
  CU(cudaMemcpyAsync(e->d_nominal, e->h_nominal, sizeof(double) * e->nelem, cudaMemcpyHostToDevice, e->stream));
  CU(cudaEventRecord(e->ev_h2d, e->stream)); e->h2d_pending = true;
  return 0;
}

int ps_set_noise(PsEngine* e, const double* z, const unsigned char* use_std2, int N) {
  if (!e || !e->have_model) return fail("ps_set_noise: no model");
  CU(cudaSetDevice(e->device));
  if (!z) { e->have_noise = false; return 0; }
  if (e->P <= 0) return fail("ps_set_noise: set the nominal plan first");
  const size_t need = (size_t)N * e->nelem;
  if (ensure(&e->d_z, &e->z_cap, need)) return -1;
  CU(cudaMemcpy(e->d_z, z, need * sizeof(double), cudaMemcpyHostToDevice));
  if (e->d_std2) { cudaFree(e->d_std2); e->d_std2 = nullptr; }
  if (use_std2) { CU(cudaMalloc(&e->d_std2, N)); CU(cudaMemcpy(e->d_std2, use_std2, N, cudaMemcpyHostToDevice)); }
  e->have_noise = true;
  return 0;
}

int ps_set_param_std(PsEngine* e, const double* param_std) {
  if (!e || !e->have_model) return fail("ps_set_param_std: no model");
  CU(cudaSetDevice(e->device));
  if (!param_std) { e->have_param_std = false; return 0; }
  if (e->P <= 0) return fail("ps_set_param_std: set the nominal plan first");
  if (!e->d_param_std) CU(cudaMalloc(&e->d_param_std, sizeof(double) * 36 * 32));
  CU(cudaMemcpyAsync(e->d_param_std, param_std, sizeof(double) * e->nelem, cudaMemcpyHostToDevice, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  e->have_param_std = true;
  return 0;
}

int ps_elite_update(PsEngine* e, int n_candidates, int n_elite, double* mean_knots, double* variance, double* avg_return,
                    double* best_return, int* order) {
  if (!e || e->N <= 0) return fail("ps_elite_update: no rollouts yet");
  if (n_candidates < 1 || n_candidates > e->N) return fail("ps_elite_update: n_candidates out of range");
  if (n_elite < 1 || n_elite > n_candidates) return fail("ps_elite_update: n_elite out of range");
  CU(cudaSetDevice(e->device));
  if (e->elite_cap < n_elite) {
    if (e->d_elite) cudaFree(e->d_elite);
    e->d_elite = nullptr; e->elite_cap = 0;
    CU(cudaMalloc(&e->d_elite, sizeof(int) * n_elite)); e->elite_cap = n_elite;
  }
  const int nout = 2 + 2 * e->nelem;
  if (!e->d_ce_out) { CU(cudaMalloc(&e->d_ce_out, sizeof(double) * (2 + 2 * 36 * 32))); CU(cudaMallocHost(&e->h_ce_out, sizeof(double) * (2 + 2 * 36 * 32))); }
  CU(cudaEventRecord(e->ev[0], e->stream));
  rank_kernel<<<(n_candidates + 255) / 256, 256, 0, e->stream>>>(e->d_returns, n_candidates, n_elite, e->d_elite);
  elite_stats_kernel<<<(e->nelem + 127) / 128, 128, 0, e->stream>>>(e->d_returns, e->d_knots, e->stride, e->nelem, e->d_elite, n_elite, e->d_ce_out);
  CU(cudaEventRecord(e->ev[1], e->stream));
  CU(cudaMemcpyAsync(e->h_ce_out, e->d_ce_out, sizeof(double) * nout, cudaMemcpyDeviceToHost, e->stream));
  std::vector<int> ord;
  if (order) CU(cudaMemcpyAsync(order, e->d_elite, sizeof(int) * n_elite, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  CU(cudaGetLastError());
  cudaEventElapsedTime(&e->ce_ms, e->ev[0], e->ev[1]);
  e->launches += 2;
  if (avg_return) *avg_return = e->h_ce_out[0];
  if (best_return) *best_return = e->h_ce_out[1];
  if (mean_knots) std::memcpy(mean_knots, e->h_ce_out + 2, sizeof(double) * e->nelem);
  if (variance) std::memcpy(variance, e->h_ce_out + 2 + e->nelem, sizeof(double) * e->nelem);
  return 0;
}

int ps_elite_update_ms(PsEngine* e, float* ms) { if (!e) return fail("null engine"); if (ms) *ms = e->ce_ms; return 0; }

int ps_set_seed(PsEngine* e, unsigned long long seed) { if (!e) return fail("null engine"); e->seed = seed; e->iteration = 0; return 0; }

static int alloc_samples(PsEngine* e, int N) {
  const int stride = (N + 31) & ~31;
  if (stride > e->cap) {
    if (e->d_returns) cudaFree(e->d_returns); if (e->d_failure) cudaFree(e->d_failure); if (e->d_blockmin) cudaFree(e->d_blockmin);
    if (e->d_fix) cudaFree(e->d_fix);
    e->d_returns = nullptr; e->d_failure = nullptr; e->d_blockmin = nullptr; e->d_fix = nullptr; e->cap = 0;
    CU(cudaMalloc(&e->d_fix, sizeof(int) * (stride + 1)));
    CU(cudaMalloc(&e->d_returns, sizeof(double) * stride));
    CU(cudaMalloc(&e->d_failure, stride));
    CU(cudaMalloc(&e->d_blockmin, sizeof(MinLoc) * (stride + 1)));   // one entry per block (a block is >= 1 rollout) + a spare
    e->cap = stride;
  }
  if (ensure(&e->d_knots, &e->knots_cap, (size_t)stride * e->nelem)) return -1;
  e->stride = stride;
  return 0;
}

static int prepare_writeback(PsEngine* e, int H, int stride) {
  const size_t ns = e->ns, nr = e->task.num_residual, nt = 3 * (size_t)e->task.num_trace;
  double** ptr[7] = {&e->wb.states, &e->wb.actions, &e->wb.residual, &e->wb.costs, &e->wb.times, &e->wb.trace, &e->wb.warmstart};
  const size_t need[7] = {H * ns * stride, (size_t)H * e->nu * stride, H * nr * stride, (size_t)H * stride, (size_t)H * stride, H * (nt ? nt : 1) * stride,
                          (size_t)(e->nv > 0 ? e->nv : 1) * stride};
  for (int k = 0; k < 7; k++) { if (ensure(ptr[k], &e->wb_cap[k], need[k])) return -1; CU(cudaMemsetAsync(*ptr[k], 0, need[k] * sizeof(double), e->stream)); }
  return 0;
}

static int run_iteration(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset, bool writeback) {
  if (!e || !e->have_model || !e->have_task) return fail("ps_rollouts: model and task must be set");
  if (e->P <= 0) return fail("ps_rollouts: nominal plan not set");
  if (N < 1 || H < 1) return fail("ps_rollouts: N and H must be positive");
  if (H > 512) return fail("ps_rollouts: horizon exceeds kMaxTrajectoryHorizon (512)");
  CU(cudaSetDevice(e->device));
  if (alloc_samples(e, N)) return -1;
  e->N = N; e->H = H; e->offset = sample_offset; e->sigma0 = sigma0; e->sigma1 = sigma1;
  const Variant* run = e->var;
  if (e->var_large && e->large_ok && (e->engine_override == 2 || (e->engine_override == 0 && N > e->crossover))) run = e->var_large;
  const bool large = run == e->var_large;
  char* hc = (char*)(large ? e->h_const2 : e->h_const); char* dc = (char*)(large ? e->d_const2 : e->d_const);
  wait_upload(e);
  e->var->plan(e->h_const, nullptr, 0, nullptr, 0, std::nan(""), nullptr, 0, 0, H);
  if (e->var_large) e->var_large->plan(e->h_const2, nullptr, 0, nullptr, 0, std::nan(""), nullptr, 0, 0, H);
  // model + task constants go up only when they changed; every iteration uploads the plan part (state, knot times)
  bool& dirty = large ? e->const2_dirty : e->const_dirty;
  if (dirty) { CU(cudaMemcpyAsync(dc, hc, run->const_bytes, cudaMemcpyHostToDevice, e->stream)); e->last_upload = run->const_bytes; dirty = false; }
  else { CU(cudaMemcpyAsync(dc + run->plan_offset, hc + run->plan_offset, run->plan_bytes, cudaMemcpyHostToDevice, e->stream)); e->last_upload = run->plan_bytes; }
  CU(cudaEventRecord(e->ev_h2d, e->stream)); e->h2d_pending = true;
  e->last = run; e->last_dconst = dc; e->record_fresh = false;
  const int stride = e->stride, nelem = e->nelem;
  if (writeback && prepare_writeback(e, H, stride)) return -1;
  CU(cudaEventRecord(e->ev[0], e->stream));
  dim3 pg((N + 255) / 256, nelem);
  perturb_kernel<<<pg, 256, 0, e->stream>>>(e->d_nominal, e->d_ctrlrange, N, nelem, e->nu, sigma0, sigma1, e->seed, e->iteration,
                                            (long long)sample_offset, e->have_noise ? e->d_z : nullptr,
                                            e->have_noise ? e->d_std2 : nullptr, e->d_knots, stride, e->have_param_std ? e->d_param_std : nullptr);
  CU(cudaEventRecord(e->ev[1], e->stream));
  e->nblocks = run->launch(e->stream, dc, e->d_knots, stride, 0, N, (long long)sample_offset, e->d_returns, e->d_failure, e->d_blockmin, e->wb, writeback, e->d_counter);
  if (run->fixup) {
    // capacity overflow never changes the answer: flagged rollouts are rolled again at the large capacities, then the
    // block minima are rebuilt from the (possibly rewritten) returns
    CU(cudaMemsetAsync(e->d_fix, 0, sizeof(int), e->stream));
    collect_flagged_kernel<<<(N + 255) / 256, 256, 0, e->stream>>>(e->d_failure, 0, N, e->d_fix + 1, e->d_fix);
    run->fixup(e->stream, dc, e->d_knots, stride, e->d_returns, e->d_failure, e->d_fix + 1, e->d_fix, e->d_counter, e->wb, writeback);
    e->nblocks = (N + 1023) / 1024;
    minloc_returns_kernel<<<e->nblocks, 256, 0, e->stream>>>(e->d_returns, N, (long long)sample_offset, e->d_blockmin);
    e->launches += 3;
  }
  CU(cudaEventRecord(e->ev[2], e->stream));
  argmin_kernel<<<1, 256, 0, e->stream>>>(e->d_blockmin, e->nblocks, e->d_returns, e->d_knots, stride, nelem, (long long)sample_offset, e->d_record);
  CU(cudaEventRecord(e->ev[3], e->stream));
  CU(cudaGetLastError());
  e->launches += 3; e->iteration++;
  if (e->comm) {
    // sharded batch: ONE all-gather of the per-rank winner records, then the min-loc (lowest global index on ties)
    // + knot select on every rank; the global record replaces the local one for the readers below
    const int count = 3 + nelem;
    NC(nccl_api().AllGather(e->d_record, e->d_gather, (size_t)count, ncclDouble, e->comm, e->stream));
    select_winner_kernel<<<1, 128, 0, e->stream>>>(e->d_gather, e->world, count, e->d_global);
    e->launches += 2;
  }
  CU(cudaEventRecord(e->ev_x, e->stream));
  e->wb_valid = writeback; e->wb_N = N; e->wb_H = H;
  return 0;
}

/* ---- multi-GPU: the winner exchange lives in the library */
int ps_comm_unique_id(void* out, int cap) {
  NcclApi& a = nccl_api();
  if (!a.ok) return fail("ps_comm_unique_id: " + a.why);
  if (cap < (int)sizeof(ncclUniqueId)) return fail("ps_comm_unique_id: buffer smaller than ncclUniqueId (128 bytes)");
  ncclUniqueId id; NC(a.GetUniqueId(&id));
  std::memcpy(out, &id, sizeof(id));
  return (int)sizeof(id);
}

int ps_comm_init(PsEngine* e, int rank, int world, const void* unique_id) {
  if (!e) return fail("null engine");
  if (world < 1 || rank < 0 || rank >= world) return fail("ps_comm_init: bad rank / world");
  NcclApi& a = nccl_api();
  if (!a.ok) return fail("ps_comm_init: " + a.why);
  CU(cudaSetDevice(e->device));
  if (e->comm) { a.CommDestroy(e->comm); e->comm = nullptr; }
  if (world == 1) { e->rank = 0; e->world = 1; return 0; }
  ncclUniqueId id; std::memcpy(&id, unique_id, sizeof(id));
  NC(a.CommInitRank(&e->comm, world, id, rank));
  e->rank = rank; e->world = world;
  if (e->d_gather) cudaFree(e->d_gather);
  if (e->d_global) cudaFree(e->d_global);
  CU(cudaMalloc(&e->d_gather, sizeof(double) * (3 + 36 * 32) * world));
  CU(cudaMalloc(&e->d_global, sizeof(double) * (3 + 36 * 32)));
  return 0;
}

int ps_comm_info(PsEngine* e, int* rank, int* world, int* nccl_version) {
  if (!e) return fail("null engine");
  if (rank) *rank = e->rank;
  if (world) *world = e->comm ? e->world : 1;
  if (nccl_version) { *nccl_version = 0; if (nccl_api().ok) nccl_api().GetVersion(nccl_version); }
  return 0;
}

int ps_set_engine_flavour(PsEngine* e, int flavour) {
  if (!e) return fail("null engine");
  if (flavour < 0 || flavour > 3) return fail("ps_set_engine_flavour: 0 auto, 1 warp-per-rollout, 2 thread-per-rollout, 3 dense thread-per-rollout");
  if (e->have_model) return fail("ps_set_engine_flavour: call before ps_set_model");
  e->engine_override = flavour;
  return 0;
}

int ps_rollout_candidates(PsEngine* e, const double* knots, int n, int repetitions, int H, double xfrc_std, double xfrc_rate,
                          const double* z, double* returns, unsigned char* failure) {
  if (!e || !e->have_model || !e->have_task) return fail("ps_rollout_candidates: model and task must be set");
  if (e->P <= 0) return fail("ps_rollout_candidates: set the nominal plan (knot times) first");
  if (n < 1 || repetitions < 1 || H < 2 || H > 512) return fail("ps_rollout_candidates: bad sizes");
  const Variant* run = e->var->robust ? e->var : (e->var_large && e->var_large->robust ? e->var_large : nullptr);
  if (!run) return fail("ps_rollout_candidates: force perturbations run on the dense thread-per-rollout kernel; create the engine with ps_set_engine_flavour(e, 3)");
  CU(cudaSetDevice(e->device));
  const int N = n * repetitions;
  if (alloc_samples(e, N)) return -1;
  const int stride = e->stride, nelem = e->nelem, n6 = 6 * e->nbody;
  e->N = N; e->H = H; e->offset = 0;
  const bool large = run == e->var_large;
  void* hc = large ? e->h_const2 : e->h_const; void* dc = large ? e->d_const2 : e->d_const;
  run->plan(hc, nullptr, 0, nullptr, 0, std::nan(""), nullptr, 0, 0, H);
  CU(cudaMemcpyAsync(dc, hc, run->const_bytes, cudaMemcpyHostToDevice, e->stream));
  e->last = run; e->last_dconst = dc;
  // candidate c fills columns c * repetitions .. (c + 1) * repetitions - 1
  std::vector<double> hk((size_t)nelem * stride, 0.0);
  for (int c = 0; c < n; c++) for (int j = 0; j < repetitions; j++) for (int k = 0; k < nelem; k++) hk[(size_t)k * stride + c * repetitions + j] = knots[(size_t)c * nelem + k];
  CU(cudaMemcpyAsync(e->d_knots, hk.data(), hk.size() * sizeof(double), cudaMemcpyHostToDevice, e->stream));
  const size_t zn = (size_t)(H - 1) * n6 * stride;
  std::vector<double> hz;
  XfrcNoise xf{nullptr, 0.0, 0.0};
  if (xfrc_std > 0) {
    if (ensure(&e->d_xfrc, &e->xfrc_cap, zn)) return -1;
    if (z) {      // injected [N][H-1][6 nbody] -> device layout
      hz.assign(zn, 0.0);
      for (int i = 0; i < N; i++) for (size_t r = 0; r < (size_t)(H - 1) * n6; r++) hz[r * stride + i] = z[(size_t)i * (H - 1) * n6 + r];
      CU(cudaMemcpyAsync(e->d_xfrc, hz.data(), zn * sizeof(double), cudaMemcpyHostToDevice, e->stream));
    } else {
      const size_t total = zn; const int threads = 256; const int blocks = (int)((total + threads - 1) / threads);
      normal_fill_kernel<<<blocks, threads, 0, e->stream>>>(e->d_xfrc, total, e->seed, e->iteration);
      e->launches++;
    }
    xf.z = e->d_xfrc;
    xf.rate = std::exp(-e->timestep / xfrc_rate);
    xf.scale = xfrc_std * std::sqrt(1 - xf.rate * xf.rate);
  }
  if (prepare_writeback(e, H, stride)) return -1;
  CU(cudaEventRecord(e->ev[0], e->stream)); CU(cudaEventRecord(e->ev[1], e->stream));
  run->robust(e->stream, dc, e->d_knots, stride, N, e->d_returns, e->d_failure, e->wb, xf);
  CU(cudaEventRecord(e->ev[2], e->stream)); CU(cudaEventRecord(e->ev[3], e->stream));
  CU(cudaStreamSynchronize(e->stream));       // hk / hz are pageable host buffers
  CU(cudaGetLastError());
  e->launches += 1; e->iteration++;
  e->wb_valid = true; e->wb_N = N; e->wb_H = H;
  if (returns) CU(cudaMemcpy(returns, e->d_returns, sizeof(double) * N, cudaMemcpyDeviceToHost));
  if (failure) CU(cudaMemcpy(failure, e->d_failure, N, cudaMemcpyDeviceToHost));
  return 0;
}

int ps_rollouts_async(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset) {
  return run_iteration(e, N, H, sigma0, sigma1, sample_offset, false);
}

static int fetch_record(PsEngine* e, int* winner, double* winner_return, double* nominal_return) {
  CU(cudaMemcpyAsync(e->h_record, e->comm ? e->d_global : e->d_record, sizeof(double) * (3 + e->nelem), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  e->record_fresh = true; e->h2d_pending = false;
  // sharded: the winner may live on another rank (then *winner = -1)
  const long long local = (long long)e->h_record[1] - e->offset;
  if (winner) *winner = (local >= 0 && local < e->N) ? (int)local : -1;
  if (winner_return) *winner_return = e->h_record[0];
  if (nominal_return) *nominal_return = e->h_record[2];
  return 0;
}

int ps_rollouts(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset, int* winner, double* winner_return,
                double* nominal_return) {
  if (run_iteration(e, N, H, sigma0, sigma1, sample_offset, false)) return -1;
  return fetch_record(e, winner, winner_return, nominal_return);
}

/* same as ps_rollouts but with full trajectory writeback of every sample (parity tests, traces) */
int ps_rollouts_writeback(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset, int* winner,
                          double* winner_return, double* nominal_return) {
  if (run_iteration(e, N, H, sigma0, sigma1, sample_offset, true)) return -1;
  return fetch_record(e, winner, winner_return, nominal_return);
}

int ps_get_returns(PsEngine* e, double* total_return, unsigned char* failure) {
  if (!e || e->N <= 0) return fail("ps_get_returns: nothing rolled out");
  CU(cudaSetDevice(e->device));
  CU(cudaStreamSynchronize(e->stream));
  if (total_return) CU(cudaMemcpy(total_return, e->d_returns, sizeof(double) * e->N, cudaMemcpyDeviceToHost));
  if (failure) CU(cudaMemcpy(failure, e->d_failure, e->N, cudaMemcpyDeviceToHost));
  return 0;
}

static int gather_column(PsEngine* e, const double* src, int rows, int i, double* out) {
  if (!out || rows == 0) return 0;
  if (ensure(&e->d_scratch, &e->scratch_cap, rows)) return -1;
  gather_column_kernel<<<(rows + 127) / 128, 128, 0, e->stream>>>(src, rows, e->stride, i, e->d_scratch);
  e->launches++;
  CU(cudaMemcpyAsync(out, e->d_scratch, sizeof(double) * rows, cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return 0;
}

int ps_get_knots(PsEngine* e, int sample, double* knots) {
  if (!e || e->N <= 0 || sample < 0 || sample >= e->N) return fail("ps_get_knots: bad sample");
  CU(cudaSetDevice(e->device));
  return gather_column(e, e->d_knots, e->nelem, sample, knots);
}

int ps_get_trajectory(PsEngine* e, int sample, double* states, double* actions, double* times, double* residual, double* costs,
                      double* trace, double* total_return, int* failure) {
  if (!e || e->N <= 0 || sample < 0 || sample >= e->N) return fail("ps_get_trajectory: bad sample");
  CU(cudaSetDevice(e->device));
  const int H = e->H, stride = e->stride;
  if (!e->wb_valid) {
    // re-roll just this sample with writeback (candidate knots are still resident)
    const size_t ns = e->ns, nr = e->task.num_residual, nt = 3 * (size_t)e->task.num_trace;
    double** ptr[7] = {&e->wb.states, &e->wb.actions, &e->wb.residual, &e->wb.costs, &e->wb.times, &e->wb.trace, &e->wb.warmstart};
    const size_t need[7] = {H * ns * stride, (size_t)H * e->nu * stride, H * nr * stride, (size_t)H * stride, (size_t)H * stride, H * (nt ? nt : 1) * stride,
                            (size_t)(e->nv > 0 ? e->nv : 1) * stride};
    for (int k = 0; k < 7; k++) if (ensure(ptr[k], &e->wb_cap[k], need[k])) return -1;
    // zero this sample's column lazily: a failed rollout leaves later steps untouched, so clear all
    for (int k = 0; k < 7; k++) CU(cudaMemsetAsync(*ptr[k], 0, need[k] * sizeof(double), e->stream));
    MinLoc* tmp = e->d_blockmin + e->cap;   // spare slot
    // single re-roll: whichever kernel scored the batch (its constants are still resident)
    e->last->launch(e->stream, e->last_dconst, e->d_knots, stride, sample, 1, e->offset, e->d_returns, e->d_failure, tmp, e->wb, true, e->d_counter);
    e->launches++;
    if (e->last->fixup) {      // the scoring kernel's capacity overflows for this sample again: take the large-capacity trajectory
      CU(cudaMemsetAsync(e->d_fix, 0, sizeof(int), e->stream));
      collect_flagged_kernel<<<1, 32, 0, e->stream>>>(e->d_failure, sample, 1, e->d_fix + 1, e->d_fix);
      e->last->fixup(e->stream, e->last_dconst, e->d_knots, stride, e->d_returns, e->d_failure, e->d_fix + 1, e->d_fix, e->d_counter, e->wb, true);
      e->launches += 2;
    }
    CU(cudaGetLastError());
  }
  const int ns = e->ns, nr = e->task.num_residual, nt = 3 * e->task.num_trace;
  if (gather_column(e, e->wb.states, H * ns, sample, states)) return -1;
  if (gather_column(e, e->wb.actions, H * e->nu, sample, actions)) return -1;
  if (gather_column(e, e->wb.times, H, sample, times)) return -1;
  if (gather_column(e, e->wb.residual, H * nr, sample, residual)) return -1;
  if (gather_column(e, e->wb.costs, H, sample, costs)) return -1;
  if (nt && gather_column(e, e->wb.trace, H * nt, sample, trace)) return -1;
  CU(cudaStreamSynchronize(e->stream));
  if (total_return) CU(cudaMemcpy(total_return, e->d_returns + sample, sizeof(double), cudaMemcpyDeviceToHost));
  if (failure) { unsigned char f = 0; CU(cudaMemcpy(&f, e->d_failure + sample, 1, cudaMemcpyDeviceToHost)); *failure = f & 1; }
  return 0;
}

// qacc_warmstart of the following launches' rollouts (DevPlan::warmstart); null: back to zeros
int ps_set_warmstart(PsEngine* e, const double* qacc_warmstart, int nv) {
  if (!e || !e->have_model) return fail("ps_set_warmstart: no model");
  if (qacc_warmstart && nv != e->nv) return fail("ps_set_warmstart: nv does not match the model");
  if (!e->var->warmstart) return fail("ps_set_warmstart: not supported by this kernel");
  e->var->warmstart(e->h_const, qacc_warmstart, e->nv);
  if (e->var_large && e->var_large->warmstart) e->var_large->warmstart(e->h_const2, qacc_warmstart, e->nv);
  return 0;
}

// qacc_warmstart rollout `sample` of the last writeback batch carried out of its last integrated step
int ps_get_warmstart(PsEngine* e, int sample, double* qacc_warmstart) {
  if (!e || !e->wb_valid || !e->wb.warmstart) return fail("ps_get_warmstart: call ps_rollouts_writeback first");
  if (sample < 0 || sample >= e->wb_N || !qacc_warmstart) return fail("ps_get_warmstart: bad sample");
  CU(cudaSetDevice(e->device));
  if (gather_column(e, e->wb.warmstart, e->nv, sample, qacc_warmstart)) return -1;
  CU(cudaStreamSynchronize(e->stream));
  return 0;
}

static int transpose_out(PsEngine* e, const double* src, int rows, double* out) {
  if (!out || rows == 0) return 0;
  const size_t total = (size_t)rows * e->N;
  if (ensure(&e->d_scratch, &e->scratch_cap, total)) return -1;
  dim3 grid((e->N + 31) / 32, (rows + 31) / 32), block(32, 8);
  transpose_kernel<<<grid, block, 0, e->stream>>>(src, rows, e->stride, e->N, e->d_scratch);
  e->launches++;
  CU(cudaMemcpyAsync(out, e->d_scratch, total * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  CU(cudaStreamSynchronize(e->stream));
  return 0;
}

int ps_get_all_trajectories(PsEngine* e, double* states, double* actions, double* residual, double* costs) {
  if (!e || !e->wb_valid) return fail("ps_get_all_trajectories: call ps_rollouts_writeback first");
  CU(cudaSetDevice(e->device));
  const int H = e->H;
  if (transpose_out(e, e->wb.states, H * e->ns, states)) return -1;
  if (transpose_out(e, e->wb.actions, H * e->nu, actions)) return -1;
  if (transpose_out(e, e->wb.residual, H * e->task.num_residual, residual)) return -1;
  if (transpose_out(e, e->wb.costs, H, costs)) return -1;
  return 0;
}

int ps_get_all_traces(PsEngine* e, double* trace) {
  if (!e || !e->wb_valid) return fail("ps_get_all_traces: call ps_rollouts_writeback first");
  if (e->task.num_trace <= 0) return fail("ps_get_all_traces: the task has no trace sensors");
  CU(cudaSetDevice(e->device));
  return transpose_out(e, e->wb.trace, e->H * 3 * e->task.num_trace, trace);
}

int ps_last_kernel_ms(PsEngine* e, float* perturb_ms, float* rollout_ms, float* reduce_ms) {
  if (!e) return fail("null engine");
  CU(cudaSetDevice(e->device));
  CU(cudaEventSynchronize(e->ev[3]));
  if (perturb_ms) CU(cudaEventElapsedTime(perturb_ms, e->ev[0], e->ev[1]));
  if (rollout_ms) CU(cudaEventElapsedTime(rollout_ms, e->ev[1], e->ev[2]));
  if (reduce_ms) CU(cudaEventElapsedTime(reduce_ms, e->ev[2], e->ev[3]));
  return 0;
}

/* device time of the cross-rank winner exchange of the last iteration (0 without a communicator) */
int ps_last_exchange_ms(PsEngine* e, float* ms) {
  if (!e) return fail("null engine");
  CU(cudaSetDevice(e->device));
  CU(cudaEventSynchronize(e->ev_x));
  if (ms) CU(cudaEventElapsedTime(ms, e->ev[3], e->ev_x));
  return 0;
}

/* bytes of constants the last iteration uploaded (the whole block after a model / task change, else the plan part) */
int ps_last_upload_bytes(PsEngine* e) { return e ? (int)e->last_upload : -1; }

int ps_kernel_launch_count(PsEngine* e, long long* launches) { if (!e) return fail("null engine"); *launches = e->launches; return 0; }

int ps_winner_record_device(PsEngine* e, void** device_ptr, int* count) {
  if (!e) return fail("null engine");
  if (device_ptr) *device_ptr = e->comm ? e->d_global : e->d_record;
  if (count) *count = 3 + e->nelem;
  return 0;
}

/* measured DFMA throughput of the device in TFLOP/s (best of `reps` launches) */
int ps_fp64_peak(PsEngine* e, int reps, double* tflops) {
  if (!e) return fail("null engine");
  CU(cudaSetDevice(e->device));
  int sms = 0; CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->device));
  const int blocks = sms * 8, threads = 256, iters = 1 << 16;
  double* out = nullptr; CU(cudaMalloc(&out, sizeof(double) * blocks * threads));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  double best = 0;
  for (int r = 0; r < reps + 1; r++) {
    cudaEventRecord(a, e->stream);
    dfma_probe_kernel<<<blocks, threads, 0, e->stream>>>(out, iters, 0.999999, 1e-6);
    cudaEventRecord(b, e->stream);
    CU(cudaEventSynchronize(b));
    float ms = 0; cudaEventElapsedTime(&ms, a, b);
    const double tf = 2.0 * 8 * (double)iters * blocks * threads / (ms * 1e-3) / 1e12;
    if (r > 0 && tf > best) best = tf;
  }
  cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(out);
  *tflops = best; e->launches += reps + 1;
  return 0;
}

/* host copy of the winner record fetched by the last ps_rollouts: [return, global index, nominal return, knots[P*nu]] */
int ps_get_winner_record(PsEngine* e, double* out, int cap) {
  if (!e || e->N <= 0) return fail("ps_get_winner_record: nothing rolled out");
  if (!e->record_fresh) return fail("ps_get_winner_record: the host record is stale (ps_rollouts_async does not fetch it; call ps_sync_record)");
  const int n = 3 + e->nelem;
  if (cap < n) return fail("ps_get_winner_record: buffer too small");
  for (int i = 0; i < n; i++) out[i] = e->h_record[i];
  return 0;
}

/* after ps_rollouts_async: wait for the iteration and fetch the winner record (what ps_rollouts does itself) */
int ps_sync_record(PsEngine* e, int* winner, double* winner_return, double* nominal_return) {
  if (!e || e->N <= 0) return fail("ps_sync_record: nothing rolled out");
  CU(cudaSetDevice(e->device));
  return fetch_record(e, winner, winner_return, nominal_return);
}

int ps_stream(PsEngine* e, void** stream) { if (!e) return fail("null engine"); *stream = (void*)e->stream; return 0; }

int ps_variant(PsEngine* e, char* out, int cap) {
  if (!e || !e->var) return fail("no model"); std::snprintf(out, cap, "%s", (e->last ? e->last : e->var)->name); return 0;
}

int ps_constants_bytes(PsEngine* e) { return (e && e->var) ? (int)(e->last ? e->last : e->var)->const_bytes : -1; }

}  // extern "C"
