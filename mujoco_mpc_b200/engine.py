"""ctypes binding of the C-ABI engine library (include/ps_engine.h -> libps_engine.so).

This is plumbing for tests and the benchmark; the product is the shared library.  There is no
CPU fallback: if the library is missing or no CUDA device is present, construction raises.
"""
import ctypes
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libps_engine.so")
_dp = ctypes.POINTER(ctypes.c_double)
_LIB = None

EXPORTS = ["ps_last_error", "ps_create", "ps_destroy", "ps_set_model", "ps_set_options", "ps_set_task", "ps_set_state",
           "ps_set_nominal", "ps_set_noise", "ps_set_seed", "ps_rollouts", "ps_rollouts_async", "ps_rollouts_writeback",
           "ps_get_returns", "ps_get_knots", "ps_get_trajectory", "ps_get_all_trajectories", "ps_last_kernel_ms",
           "ps_kernel_launch_count", "ps_winner_record_device", "ps_stream", "ps_name", "ps_variant", "ps_fp64_peak", "ps_constants_bytes",
           "ps_set_param_std", "ps_elite_update", "ps_elite_update_ms", "ps_get_all_traces",
           "ps_set_engine_flavour", "ps_rollout_candidates", "ps_comm_unique_id", "ps_comm_init", "ps_comm_info",
           "ps_last_exchange_ms", "ps_sync_record", "ps_last_upload_bytes", "ps_argmin_noalloc", "ps_noalloc_scratch_bytes",
           "ps_set_warmstart", "ps_get_warmstart"]


class EngineError(RuntimeError):
    pass


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise EngineError(f"{LIB_PATH} not built (run python mujoco_mpc_b200/build.py); there is no CPU fallback")
        _LIB = prototype(ctypes.CDLL(LIB_PATH))
    return _LIB


def prototype(L):
    """pointer-typed signatures of the handle functions (also used by the CPU test-suite on its host stand-in)"""
    L.ps_last_error.restype = ctypes.c_char_p
    L.ps_create.restype = ctypes.c_void_p
    L.ps_create.argtypes = [ctypes.c_int]
    L.ps_destroy.argtypes = [ctypes.c_void_p]
    return L


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _arr(x):
    return np.ascontiguousarray(np.asarray(x, dtype=np.float64))


def comm_unique_id():
    """128-byte ncclUniqueId created by the engine library (rank 0 calls this, the host distributes it)"""
    buf = ctypes.create_string_buffer(128)
    if lib().ps_comm_unique_id(buf, 128) < 0:
        raise EngineError(lib().ps_last_error().decode())
    return buf.raw


def broadcast_unique_id(dist, rank):
    """rank 0's ncclUniqueId to every rank of an initialised torch.distributed group (any backend)"""
    import torch
    backend = dist.get_backend()
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    t = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        t = torch.frombuffer(bytearray(comm_unique_id()), dtype=torch.uint8).to(dev)
    dist.broadcast(t, src=0)
    return bytes(t.cpu().numpy().tobytes())


class Engine:
    """One engine handle on one CUDA device."""

    def __init__(self, model, task, device=0, flavour=0):
        self.L = lib()
        h = self.L.ps_create(device)
        if not h:
            raise EngineError(self.L.ps_last_error().decode())
        self.h = ctypes.c_void_p(h)
        self.model, self.task = model, task
        if flavour:          # 1 warp-per-rollout, 2 thread-per-rollout, 3 dense thread-per-rollout
            self._ck(self.L.ps_set_engine_flavour(self.h, int(flavour)))
        self._ck(self.L.ps_set_model(self.h, ctypes.byref(model.c_struct())))
        self._ck(self.L.ps_set_task(self.h, ctypes.byref(task)))
        self.ns = model.nq + model.nv + model.na
        self.N = self.H = self.P = 0

    def _ck(self, r):
        if r != 0:
            raise EngineError(self.L.ps_last_error().decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.ps_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def variant(self):
        b = ctypes.create_string_buffer(64)
        self._ck(self.L.ps_variant(self.h, b, 64))
        return b.value.decode()

    def constants_bytes(self):
        return int(self.L.ps_constants_bytes(self.h))

    def set_task(self, task):
        self.task = task
        self._ck(self.L.ps_set_task(self.h, ctypes.byref(task)))

    def set_options(self, timestep, integrator):
        self._ck(self.L.ps_set_options(self.h, ctypes.c_double(timestep), int(integrator)))

    def set_state(self, state, time, mocap=None):
        s = _arr(state)
        mo = _arr(mocap) if mocap is not None else None
        self._ck(self.L.ps_set_state(self.h, _p(s), _p(mo), None, ctypes.c_double(time)))

    def set_warmstart(self, qacc_warmstart=None):
        """mjData::qacc_warmstart of the following launches' rollouts (None: zeros, the planning default)"""
        w = _arr(qacc_warmstart) if qacc_warmstart is not None else None
        self._ck(self.L.ps_set_warmstart(self.h, _p(w), int(w.size) if w is not None else 0))

    def warmstart(self, sample=0):
        """qacc_warmstart rollout `sample` of the last writeback batch carried out of its last integrated step"""
        out = np.zeros(self.model.nv)
        self._ck(self.L.ps_get_warmstart(self.h, int(sample), _p(out)))
        return out

    def set_nominal(self, times, knots, interp):
        t = _arr(times)
        k = _arr(knots)
        self.P = len(t)
        self._ck(self.L.ps_set_nominal(self.h, _p(t), _p(k), len(t), int(interp)))

    def set_noise(self, z, use_std2=None):
        if z is None:
            self._ck(self.L.ps_set_noise(self.h, None, None, 0))
            return
        z = _arr(z)
        u = np.ascontiguousarray(use_std2, dtype=np.uint8) if use_std2 is not None else None
        self._ck(self.L.ps_set_noise(self.h, _p(z), u.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte)) if u is not None else None,
                                     z.shape[0]))

    def rollout_candidates(self, knots, repetitions, H, xfrc_std, xfrc_rate, z=None):
        """knots [n][P][nu] -> (returns [n][repetitions], failure [n][repetitions]) under OU force noise"""
        k = _arr(knots); n = k.shape[0]
        ret = np.zeros(n * repetitions); fl = np.zeros(n * repetitions, dtype=np.uint8)
        zz = _arr(z) if z is not None else None
        self._ck(self.L.ps_rollout_candidates(self.h, _p(k), n, int(repetitions), int(H), ctypes.c_double(xfrc_std), ctypes.c_double(xfrc_rate),
                                              _p(zz), _p(ret), fl.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte))))
        self.N, self.H = n * repetitions, H
        return ret.reshape(n, repetitions), fl.reshape(n, repetitions)

    def set_param_std(self, param_std):
        """cross-entropy mode: per-parameter noise std [P][nu]; None returns to Predictive Sampling noise"""
        if param_std is None:
            self._ck(self.L.ps_set_param_std(self.h, None))
            return
        self._ck(self.L.ps_set_param_std(self.h, _p(_arr(param_std))))

    def elite_update(self, n_candidates, n_elite):
        """rank candidates on the device and reduce the elites: (mean knots [P][nu], variance [P][nu], avg return,
        best return, order[n_elite])"""
        mean = np.zeros((self.P, self.model.nu)); var = np.zeros((self.P, self.model.nu))
        avg = ctypes.c_double(); best = ctypes.c_double(); order = np.zeros(n_elite, dtype=np.int32)
        self._ck(self.L.ps_elite_update(self.h, int(n_candidates), int(n_elite), _p(mean), _p(var), ctypes.byref(avg), ctypes.byref(best),
                                        order.ctypes.data_as(ctypes.POINTER(ctypes.c_int))))
        return mean, var, avg.value, best.value, order

    def set_seed(self, seed):
        self._ck(self.L.ps_set_seed(self.h, ctypes.c_ulonglong(seed)))

    def rollouts(self, N, H, sigma0, sigma1=0.0, sample_offset=0, writeback=False):
        w = ctypes.c_int()
        wr = ctypes.c_double()
        nr = ctypes.c_double()
        fn = self.L.ps_rollouts_writeback if writeback else self.L.ps_rollouts
        self._ck(fn(self.h, N, H, ctypes.c_double(sigma0), ctypes.c_double(sigma1), sample_offset, ctypes.byref(w),
                    ctypes.byref(wr), ctypes.byref(nr)))
        self.N, self.H = N, H
        return w.value, wr.value, nr.value

    def rollouts_async(self, N, H, sigma0, sigma1=0.0, sample_offset=0):
        self._ck(self.L.ps_rollouts_async(self.h, N, H, ctypes.c_double(sigma0), ctypes.c_double(sigma1), sample_offset))
        self.N, self.H = N, H

    def returns(self):
        r = np.zeros(self.N)
        f = np.zeros(self.N, dtype=np.uint8)
        self._ck(self.L.ps_get_returns(self.h, _p(r), f.ctypes.data_as(ctypes.POINTER(ctypes.c_ubyte))))
        return r, f

    def knots(self, sample):
        k = np.zeros((self.P, self.model.nu))
        self._ck(self.L.ps_get_knots(self.h, sample, _p(k)))
        return k

    def trajectory(self, i):
        H, m, t = self.H, self.model, self.task
        st = np.zeros((H, self.ns)); ac = np.zeros((H, m.nu)); tm = np.zeros(H)
        rs = np.zeros((H, t.num_residual)); cs = np.zeros(H); tr = np.zeros((H, max(1, 3 * t.num_trace)))
        ret = ctypes.c_double(); fail = ctypes.c_int()
        self._ck(self.L.ps_get_trajectory(self.h, i, _p(st), _p(ac), _p(tm), _p(rs), _p(cs), _p(tr), ctypes.byref(ret),
                                          ctypes.byref(fail)))
        return dict(states=st, actions=ac, times=tm, residual=rs, costs=cs, trace=tr[:, :3 * t.num_trace],
                    total_return=ret.value, failure=bool(fail.value))

    def all_trajectories(self):
        N, H, m, t = self.N, self.H, self.model, self.task
        st = np.zeros((N, H, self.ns)); ac = np.zeros((N, H, m.nu)); rs = np.zeros((N, H, t.num_residual)); cs = np.zeros((N, H))
        self._ck(self.L.ps_get_all_trajectories(self.h, _p(st), _p(ac), _p(rs), _p(cs)))
        return dict(states=st, actions=ac, residual=rs, costs=cs)

    def all_traces(self):
        """trace[N][H][3*num_trace] of the last writeback batch (SamplingPlanner::Traces, planner.cc:388-434)"""
        out = np.zeros((self.N, self.H, 3 * self.task.num_trace))
        self._ck(self.L.ps_get_all_traces(self.h, _p(out)))
        return out

    def kernel_ms(self):
        a, b, c = ctypes.c_float(), ctypes.c_float(), ctypes.c_float()
        self._ck(self.L.ps_last_kernel_ms(self.h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)))
        return a.value, b.value, c.value

    def launches(self):
        n = ctypes.c_longlong()
        self._ck(self.L.ps_kernel_launch_count(self.h, ctypes.byref(n)))
        return n.value

    def kernel_ms4(self):
        """(perturb, rollout, argmin, cross-rank exchange) device ms of the last iteration"""
        x = ctypes.c_float()
        self._ck(self.L.ps_last_exchange_ms(self.h, ctypes.byref(x)))
        return self.kernel_ms() + (x.value,)

    def upload_bytes(self):
        return int(self.L.ps_last_upload_bytes(self.h))

    def sync_record(self):
        w = ctypes.c_int(); wr = ctypes.c_double(); nr = ctypes.c_double()
        self._ck(self.L.ps_sync_record(self.h, ctypes.byref(w), ctypes.byref(wr), ctypes.byref(nr)))
        return w.value, wr.value, nr.value

    def winner_record(self):
        out = np.zeros(3 + self.P * self.model.nu)
        self._ck(self.L.ps_get_winner_record(self.h, _p(out), out.size))
        return out

    def comm_init(self, rank, world, unique_id):
        """join the engine's NCCL communicator (collective over all ranks); unique_id: 128 bytes from comm_unique_id()"""
        buf = (ctypes.c_char * 128).from_buffer_copy(bytes(unique_id))
        self._ck(self.L.ps_comm_init(self.h, int(rank), int(world), buf))

    def comm_init_torch(self, dist, rank, world):
        """distribute rank 0's ncclUniqueId over an existing torch.distributed group, then ps_comm_init"""
        self.comm_init(rank, world, broadcast_unique_id(dist, rank))

    def comm_info(self):
        r = ctypes.c_int(); w = ctypes.c_int(); v = ctypes.c_int()
        self._ck(self.L.ps_comm_info(self.h, ctypes.byref(r), ctypes.byref(w), ctypes.byref(v)))
        return r.value, w.value, v.value

    def winner_record_ptr(self):
        p = ctypes.c_void_p(); n = ctypes.c_int()
        self._ck(self.L.ps_winner_record_device(self.h, ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def fp64_peak(self, reps=5):
        v = ctypes.c_double()
        self._ck(self.L.ps_fp64_peak(self.h, reps, ctypes.byref(v)))
        return v.value

    def stream(self):
        p = ctypes.c_void_p()
        self._ck(self.L.ps_stream(self.h, ctypes.byref(p)))
        return p.value


def argmin_noalloc(d_returns_ptr, n, d_out_ptr, sample_offset=0, device=0, stream=None):
    """ps_argmin_noalloc: allocation-free min-loc over a caller-owned DEVICE array (raw pointers; asynchronous on `stream`)"""
    L = lib()
    L.ps_argmin_noalloc.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_void_p]
    r = L.ps_argmin_noalloc(int(device), ctypes.c_void_p(stream or 0), ctypes.c_void_p(d_returns_ptr), int(n), int(sample_offset),
                            ctypes.c_void_p(d_out_ptr))
    if r != 0:
        raise EngineError(L.ps_last_error().decode())
