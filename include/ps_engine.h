/* C ABI of the B200 Predictive Sampling rollout engine.
 *
 * Drop-in boundary for mjpc::SamplingPlanner (reference: mjpc/planners/sampling/planner.h:41-165,
 * interface mjpc/planners/planner.h:32-102).  A host-side planner class keeps the reference's
 * Initialize/Allocate/Reset/SetState/OptimizePolicy/ActionFromPolicy methods and forwards the
 * planning iteration to these entry points.  All pointers are HOST pointers owned by the caller;
 * every call is synchronous except ps_rollouts_async (results stay on the device until ps_sync_record); the
 * setters may be called at any time -- they wait for any upload still reading the engine's pinned staging.
 * Return value: 0 = ok, <0 = error (text in ps_last_error()).
 * There is no CPU fallback: without a CUDA device ps_create() fails.
 */
#ifndef PS_ENGINE_H_
#define PS_ENGINE_H_
#include "ps_model.h"

#ifdef __cplusplus
extern "C" {
#endif

#define PS_MAX_COST_TERMS 128     /* kMaxCostTerms, mjpc/task.h:32 */
#define PS_MAX_TASK_PARAMS 64
#define PS_MAX_TASK_AUX 128

/* residual functions implemented on device (Task::Name() of the reference task) */
enum { PS_TASK_PARTICLE = 0,      /* mjpc/tasks/particle/particle.cc:50-56 (time-varying goal) */
       PS_TASK_PARTICLE_FIXED,    /* particle.cc:73-78 (mocap goal) */
       PS_TASK_PARTICLE_TEST,     /* mjpc/test/testdata/particle_residual.h:33-43 */
       PS_TASK_CARTPOLE,          /* mjpc/tasks/cartpole/cartpole.cc:36-49 */
       PS_TASK_ACROBOT,           /* mjpc/tasks/acrobot/acrobot.cc:36-49 */
       PS_TASK_SWIMMER,           /* mjpc/tasks/swimmer/swimmer.cc:33-45 */
       PS_TASK_WALKER,            /* mjpc/tasks/walker/walker.cc:38-69 */
       PS_TASK_QUADRUPED,         /* mjpc/tasks/quadruped/quadruped.cc:33-221 */
       PS_NUM_TASKS };

enum { PS_SPLINE_ZERO = 0, PS_SPLINE_LINEAR = 1, PS_SPLINE_CUBIC = 2 }; /* spline.h SplineInterpolation */

/* Cost specification + residual snapshot: what Task::Residual()/BaseResidualFn::Update()
 * (mjpc/task.cc:112-128) hands the planner each iteration, flattened to a POD. */
typedef struct PsTask_ {
  int task_id;
  int num_residual, num_term, num_trace, num_parameters;
  int dim_norm_residual[PS_MAX_COST_TERMS];
  int num_norm_parameter[PS_MAX_COST_TERMS];
  int norm[PS_MAX_COST_TERMS];                 /* mjpc::NormType, mjpc/norm.h:24-35 */
  double weight[PS_MAX_COST_TERMS];
  double norm_parameter[2 * PS_MAX_COST_TERMS];
  double risk;
  double parameters[PS_MAX_TASK_PARAMS];       /* Task::parameters (residual_* numerics) */
  double aux[PS_MAX_TASK_AUX];                 /* task-specific residual state (quadruped gait/mode) */
  int trace_sensor[4];                         /* sensor ids of trace0.. (GetTraces, utilities.cc:250) */
  int sensor_id[16];                           /* SensorByName lookups resolved once on the host (utilities.cc:166-175) */
  int iaux[16];                                /* task-specific integer state (body / geom ids, modes) */
} PsTask;

typedef struct PsEngine_ PsEngine;

const char* ps_last_error(void);
/* device_ordinal: CUDA device; rank/world: sample sharding (global sample i lives on rank i*world/N) */
PsEngine* ps_create(int device_ordinal);
void ps_destroy(PsEngine* e);
int ps_set_model(PsEngine* e, const PsModel* m);
/* Agent::PlanIteration overrides (agent.cc:275-276): model->opt.timestep / opt.integrator */
int ps_set_options(PsEngine* e, double timestep, int integrator);
int ps_set_task(PsEngine* e, const PsTask* t);
/* SamplingPlanner::SetState, planner.cc:146-149 */
int ps_set_state(PsEngine* e, const double* state, const double* mocap, const double* userdata, double time);
/* mjData::qacc_warmstart.  The reference's plant is ONE mjData stepped by mj_step (testspeed.cc:95-113, agent.cc
 * plant step), so its constraint solver starts every step from the previous step's solution.  A closed-loop plant
 * run as one launch per planning period reproduces that by handing the value across launches:
 * ps_get_warmstart reads what rollout `sample` of the last writeback batch carried out of its last integrated
 * step (nv doubles), ps_set_warmstart makes the rollouts of the following launches start from it
 * (qacc_warmstart = NULL: back to zeros).  Planning rollouts start from zeros -- the pinned reading of
 * Trajectory::Rollout (trajectory.cc:100-130), whose per-thread mjData holds whatever that thread rolled last. */
int ps_set_warmstart(PsEngine* e, const double* qacc_warmstart, int nv);
int ps_get_warmstart(PsEngine* e, int sample, double* qacc_warmstart);
/* nominal plan (policy.plan): P knot times, knots[P][nu] */
int ps_set_nominal(PsEngine* e, const double* times, const double* knots, int P, int interpolation);
/* optional injected standard-normal noise z[N][P][nu] and per-sample second-std flags (parity
 * runs); pass NULL to use the on-device counter-based generator seeded by ps_set_seed */
int ps_set_noise(PsEngine* e, const double* z, const unsigned char* use_std2, int N);
int ps_set_seed(PsEngine* e, unsigned long long seed);
/* Cross-entropy mode (CrossEntropyPlanner::AddNoiseToPolicy, cross_entropy/planner.cc:343-375): per-parameter
 * noise standard deviations param_std[P][nu] = max(sqrt(variance_k), std_min), applied un-scaled to EVERY
 * candidate; the LAST column of the next ps_rollouts batch stays un-noised and carries the nominal rollout
 * the reference schedules beside the candidates (:416).  NULL returns to Predictive Sampling noise. */
int ps_set_param_std(PsEngine* e, const double* param_std);
/* Elite update of CrossEntropyPlanner::OptimizePolicy (cross_entropy/planner.cc:196-262) over candidates
 * [0, n_candidates) of the last batch: ranks them by (return, index) on the device, then mean_knots[P][nu] =
 * elite mean, variance[P][nu] (the reference's formula, see elite_stats_kernel), avg_return over the elites,
 * best_return; order[n_elite] (optional) = candidate indices best first. */
int ps_elite_update(PsEngine* e, int n_candidates, int n_elite, double* mean_knots, double* variance, double* avg_return,
                    double* best_return, int* order);
int ps_elite_update_ms(PsEngine* e, float* ms);
/* pin the kernel flavour before ps_set_model: 0 auto (by batch size), 1 warp-per-rollout, 2 thread-per-rollout,
 * 3 dense thread-per-rollout.  Force perturbations (ps_rollout_candidates) exist for 1 (walker / quadruped class) and 3 */
int ps_set_engine_flavour(PsEngine* e, int flavour);
/* RobustPlanner's re-rolls (mjpc/planners/robust/robust_planner.cc:108-130 + Trajectory::NoisyRollout,
 * mjpc/trajectory.cc:100-210): candidate policies knots[n][P][nu] (knot times from ps_set_nominal), each rolled
 * `repetitions` times under Ornstein-Uhlenbeck forces on every body, xfrc = rate xfrc + scale N(0,1) with
 * rate = exp(-dt / xfrc_rate), scale = xfrc_std sqrt(1 - rate^2).  z: optional injected N(0,1) tensor
 * [n*repetitions][H-1][6*nbody] (parity runs; NULL = device generator).  Outputs returns / failure
 * [n*repetitions], candidate-major.  Full writeback of the batch is kept for ps_get_trajectory. */
int ps_rollout_candidates(PsEngine* e, const double* knots, int n, int repetitions, int H, double xfrc_std, double xfrc_rate,
                          const double* z, double* returns, unsigned char* failure);
/* K1+K2+K3: perturb (sample_offset..sample_offset+N of a global batch), roll out H steps,
 * reduce.  Outputs: local winner index (lowest index on ties), its return, nominal return. */
int ps_rollouts(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset,
                int* winner, double* winner_return, double* nominal_return);
int ps_get_returns(PsEngine* e, double* total_return /*[N]*/, unsigned char* failure /*[N] or NULL*/);
int ps_get_knots(PsEngine* e, int sample, double* knots /*[P][nu]*/);
/* full trajectory of one sample (re-rolled with writeback): states[H][nq+nv+na], actions[H][nu],
 * times[H], residual[H][nr], costs[H], trace[H][3*num_trace]; any pointer may be NULL */
int ps_get_trajectory(PsEngine* e, int sample, double* states, double* actions, double* times,
                      double* residual, double* costs, double* trace, double* total_return, int* failure);
/* writeback of ALL samples for parity tests: states[N][H][ns], residual[N][H][nr] (may be NULL) */
int ps_get_all_trajectories(PsEngine* e, double* states, double* actions, double* residual, double* costs);
/* trace points of ALL samples of the last ps_rollouts_writeback: trace[N][H][3*num_trace] -- what
 * SamplingPlanner::Traces (planners/sampling/planner.cc:388-434) draws as line segments between steps */
int ps_get_all_traces(PsEngine* e, double* trace);
/* device time of the last ps_rollouts kernels in milliseconds (CUDA events) */
int ps_last_kernel_ms(PsEngine* e, float* perturb_ms, float* rollout_ms, float* reduce_ms);
int ps_kernel_launch_count(PsEngine* e, long long* launches);
/* Device-resident winner record of the last ps_rollouts: [return, global index, nominal return,
 * knots[P*nu]] as doubles.  Exposed so a multi-GPU host can feed it to an NCCL all-gather
 * (min-loc over ranks) without a host round trip.  *count = 3 + P*nu. */
int ps_winner_record_device(PsEngine* e, void** device_ptr, int* count);
/* host copy of the same record as fetched by the last ps_rollouts / ps_rollouts_writeback */
int ps_get_winner_record(PsEngine* e, double* out, int cap);
/* CUDA stream (cudaStream_t) all engine work is issued on */
int ps_stream(PsEngine* e, void** stream);
/* ps_rollouts without the final device->host read: results stay in the winner record */
int ps_rollouts_async(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset);
int ps_name(char* out, int cap);
/* kernel instantiation that served the last batch: "tiny", "planar", "legged" (thread-per-rollout),
   "planar_coop" / "legged_coop" (warp-per-rollout, small batches), "legged_sparse" (thread-per-rollout over
   the compressed-Jacobian phases, large batches) */
int ps_variant(PsEngine* e, char* out, int cap);
/* bytes of model + task + plan constants uploaded per planning iteration for that kernel */
int ps_constants_bytes(PsEngine* e);
/* ps_rollouts with full per-step writeback of every sample (parity tests, traces) */
int ps_rollouts_writeback(PsEngine* e, int N, int H, double sigma0, double sigma1, int sample_offset,
                          int* winner, double* winner_return, double* nominal_return);
/* Multi-GPU sample sharding (replaces the ThreadPool fan-out of mjpc/threadpool.cc:30-85 across
 * devices): each rank rolls its slice [sample_offset, sample_offset + N) of the global batch, then ONE ncclAllGather
 * of the winner records on the engine's stream + a min-loc kernel (lowest return, ties to the lowest global index)
 * leaves the GLOBAL record [return, global index, nominal return of global sample 0, winner knots] on every rank.
 * NCCL is bound at run time (dlopen of libnccl.so.2, re-using a copy the process already holds).
 *   ps_comm_unique_id: rank 0 creates the 128-byte ncclUniqueId; the host distributes it (MPI, torch.distributed, a file)
 *   ps_comm_init:      every rank joins (collective call).  world == 1 clears the communicator.
 * With a communicator set, ps_rollouts returns winner = -1 on ranks that do not own the global winner. */
int ps_comm_unique_id(void* out, int cap);
int ps_comm_init(PsEngine* e, int rank, int world, const void* unique_id);
int ps_comm_info(PsEngine* e, int* rank, int* world, int* nccl_version);
int ps_last_exchange_ms(PsEngine* e, float* ms);
/* after ps_rollouts_async: wait for the iteration and fetch the winner record (ps_rollouts does this itself) */
int ps_sync_record(PsEngine* e, int* winner, double* winner_return, double* nominal_return);
/* bytes of constants the last iteration uploaded: the whole block after a model / task / option change, else only
 * the plan part (state, mocap, time, knot times) */
int ps_last_upload_bytes(PsEngine* e);
/* measured FP64 FMA throughput of the device (TFLOP/s): roofline denominator of the rollout kernel */
int ps_fp64_peak(PsEngine* e, int reps, double* tflops);

/* Allocation-free winner selection.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * It is the engine's argmin (block-level min-loc + grid-level min-loc, lowest index on ties, NaN never wins) over a
 * caller-owned DEVICE array d_returns[n]; d_out (DEVICE, 2 doubles) receives [min return, sample_offset + index].
 * The per-block partial results live in a file-scope __device__ array of ps_noalloc_scratch_bytes() bytes (16.6 KB,
 * held by every GPU that loads the library until the process exits) instead of a cudaMalloc'ed buffer; the call
 * performs no cudaMalloc / cudaFree / allocating library call and is asynchronous on `stream` (cudaStream_t, NULL =
 * default stream).  n > PS_NOALLOC_MAX_RETURNS is rejected.  All calls on one device share the scratch: calls that
 * overlap on the GPU (different streams, no dependency) race -- serialise them. */
#define PS_NOALLOC_MAX_RETURNS (1 << 20)
int ps_argmin_noalloc(int device_ordinal, void* stream, const double* d_returns, int n, long long sample_offset, double* d_out);
int ps_noalloc_scratch_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* PS_ENGINE_H_ */
