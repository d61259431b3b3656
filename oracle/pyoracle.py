"""ORACLE (test infrastructure, not product): ctypes wrapper over oracle/libmm_oracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_dp = ctypes.POINTER(ctypes.c_double)


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "all"], stdout=subprocess.DEVNULL)


_REF = None


def ref_lib():
    """The reference's own norm.cc / spline.cc compiled unmodified (oracle/_ref); None when absent."""
    global _REF
    path = os.path.join(_HERE, "_ref", "libmjpc_ref.so")
    if _REF is None and os.path.exists(path):
        L = ctypes.CDLL(path)
        L.ref_norm.restype = ctypes.c_double
        _REF = L
    return _REF


def ref_norm(x, params, ntype):
    x = _arr(x); p = _arr(list(params) + [0.0, 0.0, 0.0])
    return ref_lib().ref_norm(_p(x), _p(p), ctypes.c_int(x.size), ctypes.c_int(ntype))


def ref_spline_sample(times, knots, interp, t):
    times = _arr(times); knots = _arr(knots).reshape(len(times), -1)
    out = np.zeros(knots.shape[1])
    ref_lib().ref_spline_sample(_p(times), _p(knots), len(times), out.size, interp, ctypes.c_double(t), _p(out))
    return out


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "libmm_oracle.so")
        if not os.path.exists(path):
            build()
        L = ctypes.CDLL(path)
        L.mmo_create.restype = ctypes.c_void_p
        L.mmo_cost.restype = ctypes.c_double
        L.mmo_norm.restype = ctypes.c_double
        L.mmo_improvement.restype = ctypes.c_double
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _arr(x):
    return np.ascontiguousarray(np.asarray(x, dtype=np.float64))


def norm(x, params, ntype):
    x = _arr(x); p = _arr(list(params) + [0.0, 0.0])
    return lib().mmo_norm(_p(x), _p(p), ctypes.c_int(x.size), ctypes.c_int(ntype))


def spline_sample(times, knots, interp, t):
    times = _arr(times); knots = _arr(knots).reshape(len(times), -1) if len(times) else np.zeros((0, 1))
    out = np.zeros(knots.shape[1] if len(times) else 1)
    lib().mmo_spline_sample(_p(times), _p(knots), len(times), out.size, interp, ctypes.c_double(t), _p(out))
    return out


class Oracle:
    def __init__(self, model, task):
        self.model, self.task = model, task
        self.L = lib()
        self.h = ctypes.c_void_p(self.L.mmo_create(ctypes.byref(model.c_struct()), ctypes.byref(task)))
        self.ns = model.nq + model.nv + model.na

    def __del__(self):
        try:
            self.L.mmo_destroy(self.h)
        except Exception:
            pass

    # ---- engine level
    def set(self, name, v):
        v = _arr(np.atleast_1d(v))
        r = self.L.mmo_set(self.h, name.encode(), _p(v), v.size)
        assert r >= 0, name

    def get(self, name, cap=65536):
        out = np.zeros(cap)
        n = self.L.mmo_get(self.h, name.encode(), _p(out), cap)
        assert n >= 0, name
        return out[:n].copy()

    def set_task(self, task):
        self.task = task; self.L.mmo_set_task(self.h, ctypes.byref(task))

    def reset_data(self): self.L.mmo_reset_data(self.h)
    def forward(self): self.L.mmo_forward(self.h)
    def step(self): self.L.mmo_step(self.h)

    def cost(self, residual):
        r = _arr(residual); return self.L.mmo_cost(self.h, _p(r))

    def cost_terms(self, residual):
        r = _arr(residual); out = np.zeros(self.task.num_term)
        self.L.mmo_cost_terms(self.h, _p(r), _p(out)); return out

    # ---- planner level
    def configure(self, num_spline_points, interpolation, sigma0, sigma1=0.0):
        self.P = num_spline_points
        self.L.mmo_planner_config(self.h, num_spline_points, interpolation, ctypes.c_double(sigma0), ctypes.c_double(sigma1))

    def set_sliding_plan(self, on):
        self.L.mmo_set_sliding_plan(self.h, int(on))

    def set_state(self, state, time, mocap=None, userdata=None):
        s = _arr(state); mo = _arr(mocap) if mocap is not None else None
        ud = _arr(userdata) if userdata is not None else None
        self.L.mmo_set_state(self.h, _p(s), _p(mo), _p(ud), ctypes.c_double(time))

    def set_policy(self, times, knots, interp):
        t = _arr(times); k = _arr(knots)
        self.L.mmo_set_policy(self.h, _p(t), _p(k), len(t), interp)

    def optimize(self, N, H, z, use_std2=None, nthread=1):
        z = _arr(z)
        u = np.ascontiguousarray(use_std2, dtype=np.uint8) if use_std2 is not None else None
        self.N, self.H = N, H
        self.L.mmo_optimize(self.h, N, H, _p(z), u.ctypes.data_as(ctypes.c_char_p) if u is not None else None, nthread)
        return self.L.mmo_winner(self.h)

    # ---- cross-entropy planner (mjpc/planners/cross_entropy/planner.cc)
    def ce_configure(self, std_min, n_elite, variance=None):
        self.L.mmo_ce_config(self.h, ctypes.c_double(std_min), int(n_elite))
        if variance is not None:
            v = _arr(variance); self.L.mmo_ce_set_variance(self.h, _p(v), v.size)

    def ce_optimize(self, N, H, z, nthread=1):
        z = _arr(z); self.N, self.H = N, H
        self.L.mmo_optimize_ce(self.h, N, H, _p(z), nthread)
        order = np.zeros(N, dtype=np.int32)
        n = self.L.mmo_ce_order(self.h, order.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), N)
        var = np.zeros(64 * max(self.model.nu, 1)); nv = self.L.mmo_ce_get_variance(self.h, _p(var), var.size)
        self.L.mmo_ce_avg_return.restype = ctypes.c_double; self.L.mmo_ce_nominal_return.restype = ctypes.c_double
        return dict(order=order[:n], variance=var[:nv].reshape(-1, self.model.nu).copy(), avg_return=self.L.mmo_ce_avg_return(self.h),
                    nominal_return=self.L.mmo_ce_nominal_return(self.h), improvement=self.improvement())

    # ---- robust planner (mjpc/planners/robust/robust_planner.cc) over the sampling planner
    def robust_optimize(self, N, H, z, ncandidates, repetitions, w, xfrc_std, xfrc_rate):
        z = _arr(z); w = _arr(w); self.N, self.H = N, H
        nc = min(ncandidates, N)
        scores = np.zeros(nc); order = np.zeros(nc, dtype=np.int32)
        nret = np.zeros(nc * repetitions); nfail = np.zeros(nc * repetitions, dtype=np.uint8)
        best = self.L.mmo_optimize_robust(self.h, N, H, _p(z), ncandidates, repetitions, _p(w), ctypes.c_double(xfrc_std), ctypes.c_double(xfrc_rate),
                                          _p(scores), order.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), _p(nret), nfail.ctypes.data_as(ctypes.c_char_p))
        return dict(best=best, scores=scores, order=order, noisy_returns=nret.reshape(nc, repetitions), noisy_failure=nfail.reshape(nc, repetitions))

    def policy(self, which=0):
        nu = self.model.nu
        t = np.zeros(64); k = np.zeros(64 * max(nu, 1))
        P = self.L.mmo_get_policy(self.h, which, _p(t), _p(k))
        return t[:P].copy(), k[:P * nu].reshape(P, nu).copy()

    def action(self, t, use_previous=False):
        out = np.zeros(max(self.model.nu, 1))
        self.L.mmo_action(self.h, ctypes.c_double(t), int(use_previous), _p(out)); return out[:self.model.nu]

    def returns(self):
        r = np.zeros(self.N); f = np.zeros(self.N, dtype=np.uint8)
        self.L.mmo_get_returns(self.h, _p(r), f.ctypes.data_as(ctypes.c_char_p)); return r, f

    def improvement(self): return self.L.mmo_improvement(self.h)

    def trajectory(self, i):
        H, m, t = self.H, self.model, self.task
        st = np.zeros((H, self.ns)); ac = np.zeros((H, m.nu)); tm = np.zeros(H)
        rs = np.zeros((H, t.num_residual)); cs = np.zeros(H); tr = np.zeros((H, 3 * t.num_trace))
        ret = ctypes.c_double(); fail = ctypes.c_int()
        self.L.mmo_get_trajectory(self.h, i, _p(st), _p(ac), _p(tm), _p(rs), _p(cs), _p(tr), ctypes.byref(ret), ctypes.byref(fail))
        return dict(states=st, actions=ac, times=tm, residual=rs, costs=cs, trace=tr, total_return=ret.value, failure=bool(fail.value))


# ---------------------------------------------------------------------------------------------------------------------
# The reference's OWN planner / rollout / cost / residual code (oracle/_ref/libmjpc_ref.so: mjpc/planners/sampling/*.cc,
# trajectory.cc, task.cc, utilities.cc, states/state.cc, threadpool.cc and the six tasks/*/*.cc compiled unmodified; the
# physics below mj_step is the restated step above).  Present only where /root/reference was available at build time.
class RefPlanner:
    def __init__(self, model, task_name):
        L = ref_lib()
        if L is None or not hasattr(L, "refp_create"):
            raise RuntimeError("oracle/_ref/libmjpc_ref.so with the reference planner is not built (needs /root/reference)")
        L.refp_create.restype = ctypes.c_void_p
        L.refp_improvement.restype = ctypes.c_double
        L.refp_task_residual.restype = ctypes.c_double
        self.L, self.model = L, model
        self._keep = model.c_struct()                      # the PsModel must outlive the mjModel view
        self.h = ctypes.c_void_p(L.refp_create(ctypes.byref(self._keep), task_name.encode()))
        if not self.h:
            raise RuntimeError(f"reference task {task_name!r} unknown")
        self.ns = model.nq + model.nv + model.na
        nr, nt, ntr, npar = ctypes.c_int(), ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        L.refp_task_dims(self.h, ctypes.byref(nr), ctypes.byref(nt), ctypes.byref(ntr), ctypes.byref(npar))
        self.num_residual, self.num_term, self.num_trace, self.num_parameters = nr.value, nt.value, ntr.value, npar.value

    def __del__(self):
        try:
            self.L.refp_destroy(self.h)
        except Exception:
            pass

    @staticmethod
    def zero_warmstart_per_rollout(on):
        """test-harness switch (ref_glue.cc): clear qacc_warmstart at the first step of every rollout, i.e. the engine's
        semantics; off = the reference's (the worker's mjData keeps the previous rollout's)"""
        ref_lib().ref_set_zero_warmstart_on_rewind(int(bool(on)))

    @staticmethod
    def tolerate_errors(on):
        """bench-harness switch (ref_planner_wrap.cc): count mju_error calls raised inside the reference's code and go on
        instead of aborting; off (default) = abort.  Returns nothing; see error_count()."""
        ref_lib().refp_tolerate_errors(int(bool(on)))

    @staticmethod
    def error_count(reset=False):
        L = ref_lib(); L.refp_error_count.restype = ctypes.c_long
        n = int(L.refp_error_count())
        if reset:
            L.refp_error_reset()
        return n

    def task_spec(self):
        nt = self.num_term
        dims = (ctypes.c_int * nt)(); norm = (ctypes.c_int * nt)(); npar = (ctypes.c_int * nt)()
        w = np.zeros(nt); prm = np.zeros(4 * nt + 4); risk = ctypes.c_double(); par = np.zeros(max(self.num_parameters, 1))
        self.L.refp_task_spec(self.h, dims, norm, npar, _p(w), _p(prm), ctypes.byref(risk), _p(par))
        return dict(dim_norm_residual=list(dims), norm=list(norm), num_norm_parameter=list(npar), weight=w, norm_parameter=prm[:sum(npar)],
                    risk=risk.value, parameters=par[:self.num_parameters])

    def task_set(self, mode=-1, parameters=None, weight=None):
        p = _arr(parameters) if parameters is not None else None; w = _arr(weight) if weight is not None else None
        self.L.refp_task_set(self.h, int(mode), _p(p), 0 if p is None else p.size, _p(w), 0 if w is None else w.size)

    def task_transition(self, state, mocap, time):
        s = _arr(state); mo = _arr(mocap).copy() if mocap is not None else None
        mode = self.L.refp_task_transition(self.h, _p(s), _p(mo), ctypes.c_double(time))
        return mode, mo

    def residual(self, state, ctrl, mocap, time):
        s = _arr(state); c = _arr(ctrl); mo = _arr(mocap) if mocap is not None else None
        r = np.zeros(self.num_residual); t = np.zeros(self.num_term)
        cost = self.L.refp_task_residual(self.h, _p(s), _p(c), _p(mo), ctypes.c_double(time), _p(r), _p(t))
        return r, t, cost

    def configure(self, num_spline_points, interpolation, sigma0, sigma1=0.0, num_trajectory=10, sliding_plan=0):
        self.P = num_spline_points
        self.L.refp_configure(self.h, int(num_spline_points), int(interpolation), ctypes.c_double(sigma0), ctypes.c_double(sigma1),
                              int(num_trajectory), int(sliding_plan))

    def set_state(self, state, time, mocap=None, userdata=None):
        s = _arr(state); mo = _arr(mocap) if mocap is not None else None; ud = _arr(userdata) if userdata is not None else None
        self.L.refp_set_state(self.h, _p(s), _p(mo), _p(ud), ctypes.c_double(time))

    def set_policy(self, times, knots, interp):
        t = _arr(times); k = _arr(knots)
        self.L.refp_set_policy(self.h, _p(t), _p(k), len(t), int(interp))

    def optimize(self, N, H, z=None, use_std2=None, nthread=1):
        zz = _arr(z) if z is not None else None
        u = np.ascontiguousarray(use_std2, dtype=np.uint8) if use_std2 is not None else None
        self.N, self.H = N, H
        w = self.L.refp_optimize(self.h, int(N), int(H), _p(zz), u.ctypes.data_as(ctypes.c_char_p) if u is not None else None, int(nthread))
        if w < 0 and N > self.L.refp_max_trajectory():
            raise ValueError(f"the reference planner rolls at most kMaxTrajectory = {self.L.refp_max_trajectory()} samples")
        return w

    def returns(self):
        r = np.zeros(self.N); f = np.zeros(self.N, dtype=np.uint8)
        self.L.refp_get_returns(self.h, self.N, _p(r), f.ctypes.data_as(ctypes.c_char_p)); return r, f

    def trajectory(self, i):
        H, m = self.H, self.model
        st = np.zeros((H, self.ns)); ac = np.zeros((H, m.nu)); tm = np.zeros(H)
        rs = np.zeros((H, self.num_residual)); cs = np.zeros(H); tr = np.zeros((H, max(3 * self.num_trace, 1)))
        ret = ctypes.c_double(); fail = ctypes.c_int()
        self.L.refp_get_trajectory(self.h, int(i), _p(st), _p(ac), _p(tm), _p(rs), _p(cs), _p(tr), ctypes.byref(ret), ctypes.byref(fail))
        return dict(states=st, actions=ac, times=tm, residual=rs, costs=cs, trace=tr[:, :3 * self.num_trace], total_return=ret.value,
                    failure=bool(fail.value))

    def policy(self, which=0):
        nu = self.model.nu
        t = np.zeros(600); k = np.zeros(600 * max(nu, 1))
        P = self.L.refp_get_policy(self.h, int(which), _p(t), _p(k))
        return t[:P].copy(), k[:P * nu].reshape(P, nu).copy()

    def action(self, t, use_previous=False):
        out = np.zeros(max(self.model.nu, 1))
        self.L.refp_action(self.h, ctypes.c_double(t), int(use_previous), _p(out)); return out[:self.model.nu]

    def improvement(self):
        return self.L.refp_improvement(self.h)

    def quadruped_state(self):
        """QuadrupedFlat::ResidualFn's Transition-managed state in the PsTask.aux layout (task.py A_MODE .. A_ORIENTATION + 4)"""
        aux = np.zeros(17)
        if self.L.refp_quadruped_state(self.h, _p(aux)) != 0:
            raise RuntimeError("not the quadruped task")
        return aux

    def export_task_state(self, task):
        """copy the reference task's current parameters / weights (and the quadruped's residual state) into a PsTask"""
        spec = self.task_spec()
        for k, v in enumerate(spec["parameters"]):
            task.parameters[k] = float(v)
        for k, v in enumerate(spec["weight"]):
            task.weight[k] = float(v)
        try:
            aux = self.quadruped_state()
            for k in range(17):
                task.aux[k] = float(aux[k])
        except RuntimeError:
            pass
        return task
