"""ctypes binding of the host-side planner library (libps_planner.so: CudaSamplingPlanner, the
C++ drop-in for mjpc::SamplingPlanner).  Method names follow the reference's Planner interface."""
import ctypes
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libps_planner.so")
_dp = ctypes.POINTER(ctypes.c_double)
_LIB = None
ALLGATHER = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, _dp, _dp, ctypes.c_int)


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not built (python -c 'from mujoco_mpc_b200 import build; build.build(); build.build_host()')")
        ctypes.CDLL(os.path.join(_HERE, "libps_engine.so"), mode=ctypes.RTLD_GLOBAL)
        _LIB = prototype(ctypes.CDLL(LIB_PATH))
    return _LIB


def prototype(L):
    """return types of the psp_* wrappers (also used by the CPU test-suite on its host-compiled build of the classes)"""
    L.psp_create.restype = ctypes.c_void_p
    L.psp_engine.restype = ctypes.c_void_p
    L.psp_candidate_score.restype = ctypes.c_double
    L.psp_improvement.restype = ctypes.c_double
    return L


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _arr(x):
    return np.ascontiguousarray(np.asarray(x, dtype=np.float64))


def select_winner(records):
    """Cross-rank min-loc over winner records [world][3 + P*nu] (host logic, no GPU needed)."""
    r = _arr(records)
    return lib().psp_select_winner(_p(r), r.shape[0], r.shape[1])


class CudaSamplingPlanner:
    def __init__(self, model, task, device=0):
        self.L = lib()
        self.model, self.task = model, task
        self.h = ctypes.c_void_p(self.L.psp_create(ctypes.byref(model.c_struct()), ctypes.byref(task), device))
        self.ns = model.nq + model.nv + model.na
        self._cb = None
        self.horizon = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.psp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def Reset(self, horizon, initial_repeated_action=None):
        a = _arr(initial_repeated_action) if initial_repeated_action is not None else None
        self.L.psp_reset(self.h, horizon, _p(a))

    def configure(self, num_trajectory=-1, interpolation=-1, num_spline_points=-1, sigma0=-1.0, sigma1=-1.0, sliding=-1):
        self.L.psp_configure(self.h, num_trajectory, interpolation, num_spline_points, ctypes.c_double(sigma0),
                             ctypes.c_double(sigma1), sliding)

    def SetState(self, state, time, mocap=None, userdata=None):
        s = _arr(state); mo = _arr(mocap) if mocap is not None else None; ud = _arr(userdata) if userdata is not None else None
        self.L.psp_set_state(self.h, _p(s), _p(mo), _p(ud), ctypes.c_double(time))

    def SetTask(self, task):
        self.task = task
        self.L.psp_set_task(self.h, ctypes.byref(task))

    def SetNoise(self, z, use_std2=None):
        if z is None:
            self.L.psp_set_noise(self.h, None, None, 0, -1); return
        z = _arr(z)
        u = np.ascontiguousarray(use_std2, dtype=np.uint8) if use_std2 is not None else None
        z = z.reshape(z.shape[0], -1, self.model.nu)
        self.L.psp_set_noise(self.h, _p(z), u.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte)) if u is not None else None, z.shape[0], z.shape[1])

    def SetSeed(self, seed):
        self.L.psp_set_seed(self.h, ctypes.c_ulonglong(seed))

    def set_policy(self, times, knots, interp):
        """start from a given nominal spline instead of the Reset() one (parity runs)"""
        t = _arr(times); k = _arr(knots)
        self.L.psp_set_policy(self.h, _p(t), _p(k), len(t), int(interp))

    def SetSharding(self, rank, world, gather):
        """gather(mine: np.ndarray[count]) -> np.ndarray[world, count]"""
        def cb(ctx, mine, all_, count):
            try:
                m = np.ctypeslib.as_array(mine, shape=(count,)).copy()
                out = _arr(gather(m)).reshape(-1)
                ctypes.memmove(all_, out.ctypes.data, out.size * 8)
                return 0
            except Exception:
                return -1
        self._cb = ALLGATHER(cb)
        self.L.psp_set_sharding(self.h, rank, world, self._cb, None)

    def SetShardingNative(self, rank, world, dist):
        """sample sharding with the winner exchange inside the engine library (ncclAllGather + min-loc kernel);
        `dist`: an initialised torch.distributed module, used once to hand rank 0's ncclUniqueId to every rank"""
        from .engine import broadcast_unique_id
        uid = broadcast_unique_id(dist, rank)
        buf = (ctypes.c_char * 128).from_buffer_copy(uid)
        self.L.psp_set_sharding_native(self.h, int(rank), int(world), buf)

    def OptimizePolicy(self, horizon):
        self.horizon = horizon
        self.L.psp_optimize_policy(self.h, horizon)

    def OptimizePolicyCandidates(self, ncandidates, horizon):
        self.horizon = horizon
        return self.L.psp_optimize_policy_candidates(self.h, ncandidates, horizon)

    def CandidateScore(self, c):
        return self.L.psp_candidate_score(self.h, c)

    def ActionFromCandidatePolicy(self, c, time):
        a = np.zeros(max(self.model.nu, 1)); self.L.psp_action_from_candidate_policy(self.h, _p(a), c, ctypes.c_double(time))
        return a[:self.model.nu]

    def CopyCandidateToPolicy(self, c):
        self.L.psp_copy_candidate_to_policy(self.h, c)

    def NominalTrajectory(self, horizon):
        self.horizon = horizon
        self.L.psp_nominal_trajectory(self.h, horizon)

    def ActionFromPolicy(self, time, use_previous=False):
        a = np.zeros(max(self.model.nu, 1)); self.L.psp_action_from_policy(self.h, _p(a), ctypes.c_double(time), int(use_previous))
        return a[:self.model.nu]

    @property
    def winner(self):
        return self.L.psp_winner(self.h)

    @property
    def improvement(self):
        return self.L.psp_improvement(self.h)

    def engine_ptr(self):
        """PsEngine* behind the host class (returns of the whole batch, kernel timers ...)"""
        self.L.psp_engine.restype = ctypes.c_void_p
        return self.L.psp_engine(self.h)

    def NumParameters(self):
        return self.L.psp_num_parameters(self.h)

    def policy(self, previous=False):
        t = np.zeros(64); k = np.zeros(64 * max(self.model.nu, 1))
        P = self.L.psp_get_policy(self.h, int(previous), _p(t), _p(k))
        return t[:P].copy(), k[:P * self.model.nu].reshape(P, self.model.nu).copy()

    def timers(self):
        a, b, c = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
        self.L.psp_timers(self.h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
        return dict(noise_us=a.value, rollouts_us=b.value, policy_update_us=c.value)

    def BestTrajectory(self):
        H, m, t = self.horizon, self.model, self.task
        st = np.zeros((H, self.ns)); ac = np.zeros((H, m.nu)); tm = np.zeros(H); rs = np.zeros((H, t.num_residual)); cs = np.zeros(H)
        tr = np.zeros((H, max(1, 3 * t.num_trace))); ret = ctypes.c_double(); fail = ctypes.c_int()
        h = self.L.psp_best_trajectory(self.h, _p(st), _p(ac), _p(tm), _p(rs), _p(cs), _p(tr), ctypes.byref(ret), ctypes.byref(fail))
        if h == 0:
            return None
        return dict(states=st, actions=ac, times=tm, residual=rs, costs=cs, trace=tr[:, :3 * t.num_trace], total_return=ret.value,
                    failure=bool(fail.value))


class CudaCrossEntropyPlanner:
    """ctypes binding of mjpc_b200::CudaCrossEntropyPlanner (drop-in for mjpc::CrossEntropyPlanner,
    mjpc/planners/cross_entropy/planner.h): same rollout kernels, elite ranking + statistics on the device."""

    def __init__(self, model, task, device=0):
        self.L = lib()
        for f in ("pce_improvement", "pce_avg_return", "pce_best_return"):
            getattr(self.L, f).restype = ctypes.c_double
        self.L.pce_create.restype = ctypes.c_void_p
        self.L.pce_engine.restype = ctypes.c_void_p
        self.model, self.task = model, task
        self.h = ctypes.c_void_p(self.L.pce_create(ctypes.byref(model.c_struct()), ctypes.byref(task), device))
        self.ns = model.nq + model.nv + model.na
        self.horizon = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.pce_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def Reset(self, horizon, initial_repeated_action=None):
        a = _arr(initial_repeated_action) if initial_repeated_action is not None else None
        self.L.pce_reset(self.h, horizon, _p(a))

    def configure(self, num_trajectory=-1, n_elite=-1, interpolation=-1, num_spline_points=-1, std_initial=-1.0, std_min=-1.0):
        self.L.pce_configure(self.h, num_trajectory, n_elite, interpolation, num_spline_points, ctypes.c_double(std_initial),
                             ctypes.c_double(std_min))

    def SetState(self, state, time, mocap=None, userdata=None):
        s = _arr(state); mo = _arr(mocap) if mocap is not None else None; ud = _arr(userdata) if userdata is not None else None
        self.L.pce_set_state(self.h, _p(s), _p(mo), _p(ud), ctypes.c_double(time))

    def SetTask(self, task):
        self.task = task
        self.L.pce_set_task(self.h, ctypes.byref(task))

    def set_policy(self, times, knots, interp):
        t = _arr(times); k = _arr(knots)
        self.L.pce_set_policy(self.h, _p(t), _p(k), len(t), int(interp))

    def set_variance(self, v):
        v = _arr(v); self.L.pce_set_variance(self.h, _p(v), v.size)

    def variance(self, P):
        v = np.zeros(P * self.model.nu); self.L.pce_get_variance(self.h, _p(v), v.size)
        return v.reshape(P, self.model.nu)

    def SetNoise(self, z):
        if z is None:
            self.L.pce_set_noise(self.h, None, 0, -1); return
        z = _arr(z); z = z.reshape(z.shape[0], -1, self.model.nu); self.L.pce_set_noise(self.h, _p(z), z.shape[0], z.shape[1])

    def SetSeed(self, seed):
        self.L.pce_set_seed(self.h, ctypes.c_ulonglong(seed))

    def OptimizePolicy(self, horizon):
        self.horizon = horizon
        self.L.pce_optimize_policy(self.h, horizon)

    def NominalTrajectory(self, horizon):
        self.horizon = horizon
        self.L.pce_nominal_trajectory(self.h, horizon)

    def ActionFromPolicy(self, time, use_previous=False):
        a = np.zeros(max(self.model.nu, 1)); self.L.pce_action_from_policy(self.h, _p(a), ctypes.c_double(time), int(use_previous))
        return a[:self.model.nu]

    @property
    def improvement(self):
        return self.L.pce_improvement(self.h)

    @property
    def avg_return(self):
        return self.L.pce_avg_return(self.h)

    @property
    def best_return(self):
        return self.L.pce_best_return(self.h)

    def order(self, n):
        o = np.zeros(n, dtype=np.int32)
        k = self.L.pce_order(self.h, o.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), n)
        return o[:k]

    def policy(self, which=0):
        """which: 0 policy (elite mean), 1 previous_policy, 2 resampled_policy (the nominal that was rolled out)"""
        t = np.zeros(64); k = np.zeros(64 * max(self.model.nu, 1))
        P = self.L.pce_get_policy(self.h, which, _p(t), _p(k))
        return t[:P].copy(), k[:P * self.model.nu].reshape(P, self.model.nu).copy()

    def BestTrajectory(self):
        H, m, t = self.horizon, self.model, self.task
        st = np.zeros((H, self.ns)); ac = np.zeros((H, m.nu)); tm = np.zeros(H); rs = np.zeros((H, t.num_residual)); cs = np.zeros(H)
        ret = ctypes.c_double(); fail = ctypes.c_int()
        h = self.L.pce_best_trajectory(self.h, _p(st), _p(ac), _p(tm), _p(rs), _p(cs), ctypes.byref(ret), ctypes.byref(fail))
        if not h:
            return None
        return dict(states=st, actions=ac, times=tm, residual=rs, costs=cs, total_return=ret.value, failure=bool(fail.value))

    def engine_ptr(self):
        return self.L.pce_engine(self.h)


class _Delegate(CudaSamplingPlanner):
    """The CudaSamplingPlanner a CudaRobustPlanner wraps: same methods, handle owned by the robust planner."""

    def __init__(self, L, model, task, handle):
        self.L, self.model, self.task = L, model, task
        self.h = ctypes.c_void_p(handle)
        self.ns = model.nq + model.nv + model.na
        self._cb = None
        self.horizon = 0

    def close(self):
        self.h = None


class CudaRobustPlanner:
    """ctypes binding of mjpc_b200::CudaRobustPlanner (drop-in for mjpc::RobustPlanner,
    mjpc/planners/robust/robust_planner.h): the delegate's best candidates re-rolled under random body forces
    in one launch of the force-perturbed rollout kernel; the lowest mean return is kept."""

    def __init__(self, model, task, device=0, flavour=0):
        """flavour of the re-roll kernel: 0 = warp-per-rollout where the model class has one, 3 = dense thread-per-rollout"""
        self.L = lib()
        for f in ("prb_create", "prb_create_flavour", "prb_delegate", "prb_engine"):
            getattr(self.L, f).restype = ctypes.c_void_p
        self.L.prb_noisy_rollouts_us.restype = ctypes.c_double
        self.model, self.task = model, task
        self.h = ctypes.c_void_p(self.L.prb_create_flavour(ctypes.byref(model.c_struct()), ctypes.byref(task), device, int(flavour)))
        self.delegate = _Delegate(self.L, model, task, self.L.prb_delegate(self.h))
        self.horizon = 0

    def close(self):
        if getattr(self, "h", None):
            self.delegate.close()
            self.L.prb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def Reset(self, horizon, initial_repeated_action=None):
        a = _arr(initial_repeated_action) if initial_repeated_action is not None else None
        self.L.prb_reset(self.h, horizon, _p(a))

    def configure(self, ncandidates=-1, nrepetitions=-1, xfrc_std=-1.0, xfrc_rate=-1.0):
        self.L.prb_configure(self.h, ncandidates, nrepetitions, ctypes.c_double(xfrc_std), ctypes.c_double(xfrc_rate))

    def settings(self):
        a, b, c, d = ctypes.c_int(), ctypes.c_int(), ctypes.c_double(), ctypes.c_double()
        self.L.prb_get_settings(self.h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c), ctypes.byref(d))
        return dict(ncandidates=a.value, nrepetitions=b.value, xfrc_std=c.value, xfrc_rate=d.value)

    def SetState(self, state, time, mocap=None, userdata=None):
        s = _arr(state); mo = _arr(mocap) if mocap is not None else None; ud = _arr(userdata) if userdata is not None else None
        self.L.prb_set_state(self.h, _p(s), _p(mo), _p(ud), ctypes.c_double(time))

    def SetTask(self, task):
        self.task = task; self.delegate.task = task
        self.L.prb_set_task(self.h, ctypes.byref(task))

    def SetForceNoise(self, z, horizon):
        """z: N(0,1) tensor [ncandidates * nrepetitions][horizon - 1][6 * nbody] (parity runs); None = device generator"""
        if z is None:
            self.L.prb_set_force_noise(self.h, None, 0, 0); return
        z = _arr(z); self.L.prb_set_force_noise(self.h, _p(z), z.shape[0], horizon)

    def SetSeed(self, seed):
        self.L.prb_set_seed(self.h, ctypes.c_ulonglong(seed))

    def OptimizePolicy(self, horizon):
        self.horizon = self.delegate.horizon = horizon
        self.L.prb_optimize_policy(self.h, horizon)

    def NominalTrajectory(self, horizon):
        self.horizon = self.delegate.horizon = horizon
        self.L.prb_nominal_trajectory(self.h, horizon)

    def ActionFromPolicy(self, time, use_previous=False):
        a = np.zeros(max(self.model.nu, 1)); self.L.prb_action_from_policy(self.h, _p(a), ctypes.c_double(time), int(use_previous))
        return a[:self.model.nu]

    def BestTrajectory(self):
        return self.delegate.BestTrajectory()

    def NumParameters(self):
        return self.delegate.NumParameters()

    def policy(self, previous=False):
        return self.delegate.policy(previous)

    @property
    def best_candidate(self):
        return self.L.prb_best_candidate(self.h)

    def results(self):
        s = self.settings(); nc, rep = max(s["ncandidates"], 1), s["nrepetitions"]
        scores = np.zeros(nc); ret = np.zeros(nc * rep); fail = np.zeros(nc * rep, dtype=np.uint8)
        n = self.L.prb_results(self.h, _p(scores), _p(ret), fail.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte)), nc)
        return dict(scores=scores[:n], noisy_returns=ret[:n * rep].reshape(n, rep), noisy_failure=fail[:n * rep].reshape(n, rep),
                    noisy_rollouts_us=self.L.prb_noisy_rollouts_us(self.h))

    def reroll_kernel(self):
        """name of the engine variant the force-perturbed rollouts ran on"""
        buf = ctypes.create_string_buffer(64)
        eng = ctypes.CDLL(os.path.join(_HERE, "libps_engine.so"))
        eng.ps_variant(ctypes.c_void_p(self.L.prb_engine(self.h)), buf, 64)
        return buf.value.decode()
