"""ctypes binding of include/mpcb.h (libmpcb.so).  Host arrays are numpy float64, problem-major."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# MPCB_LIB selects a debug build of the same library (e.g. libmpcb_prof.so from `make prof`); default = the product
_LIB_PATH = os.environ.get("MPCB_LIB") or os.path.join(_HERE, "libmpcb.so")
_lib = None


class MpcbError(RuntimeError):
    pass


class Settings(C.Structure):
    _fields_ = [
        ("poly_order", C.c_int), ("num_seg", C.c_int), ("sqp_max_iter", C.c_int), ("qp_max_iter", C.c_int),
        ("line_search_max_iter", C.c_int), ("qp_check_termination", C.c_int), ("qp_adaptive_rho", C.c_int),
        ("qp_adaptive_rho_interval", C.c_int),
        ("qp_eps_abs", C.c_double), ("qp_eps_rel", C.c_double), ("qp_rho", C.c_double), ("qp_sigma", C.c_double),
        ("qp_alpha", C.c_double), ("qp_adaptive_rho_tolerance", C.c_double),
        ("sqp_eps_prim", C.c_double), ("sqp_eps_dual", C.c_double), ("sqp_tau", C.c_double), ("sqp_eta", C.c_double),
        ("hess_reg", C.c_double),
    ]


class Limits(C.Structure):
    _fields_ = [
        ("qmin", C.c_double * 7), ("qmax", C.c_double * 7), ("vmax", C.c_double * 7), ("amax", C.c_double * 7),
        ("taumax", C.c_double * 7), ("jmax", C.c_double * 7), ("margins", C.c_double * 5),
        ("min_height", C.c_double), ("final_eps", C.c_double), ("p_lb", C.c_double), ("p_ub", C.c_double),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("status", C.c_int), ("sqp_iters", C.c_int), ("qp_iters_total", C.c_int), ("qp_status", C.c_int * 4),
        ("qp_iters", C.c_int * 4), ("alpha", C.c_double * 4), ("ls_trials", C.c_int * 4),
        ("primal_step_norm", C.c_double), ("dual_step_norm", C.c_double), ("max_violation", C.c_double),
        ("l1_violation", C.c_double), ("cost", C.c_double),
    ]


class Account(C.Structure):
    _fields_ = [("flags", C.c_int * 4), ("mins", C.c_double * 28), ("maxs", C.c_double * 28), ("final_err", C.c_double * 14)]


STATS_DTYPE = np.dtype([("status", "i4"), ("sqp_iters", "i4"), ("qp_iters_total", "i4"), ("qp_status", "i4", 4), ("qp_iters", "i4", 4),
                        ("_pad0", "i4"), ("alpha", "f8", 4), ("ls_trials", "i4", 4), ("primal_step_norm", "f8"), ("dual_step_norm", "f8"),
                        ("max_violation", "f8"), ("l1_violation", "f8"), ("cost", "f8")])
ACCOUNT_DTYPE = np.dtype([("flags", "i4", 4), ("mins", "f8", 28), ("maxs", "f8", 28), ("final_err", "f8", 14)])
assert STATS_DTYPE.itemsize == C.sizeof(Stats), (STATS_DTYPE.itemsize, C.sizeof(Stats))
assert ACCOUNT_DTYPE.itemsize == C.sizeof(Account)

MEM_HOST, MEM_DEVICE = 0, 1
NOALLOC_MAX_OCPS = 8      # MPCB_NOALLOC_MAX_OCPS of include/mpcb.h (base shape)


def lib_path():
    return _LIB_PATH


def load_library():
    """Load libmpcb.so.  Fails loudly if it has not been built (python __graft_entry__.py / make)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise MpcbError(f"{_LIB_PATH} not found: build it with `make` (nvcc, sm_100a); there is no fallback")
        L = C.CDLL(_LIB_PATH)
        L.mpcb_last_error.restype = C.c_char_p
        L.mpcb_last_solve_ms.restype = C.c_double
        L.mpcb_measure_fp64_tflops.restype = C.c_double
        L.mpcb_launch_count.restype = C.c_longlong
        L.mpcb_last_kernel_ms.restype = C.c_double
        L.mpcb_last_kernel_ms.argtypes = [C.c_void_p, C.c_int]
        L.mpcb_last_kernel_launches.argtypes = [C.c_void_p, C.c_int]
        for name in ("mpcb_last_error", "mpcb_last_solve_ms", "mpcb_measure_fp64_tflops", "mpcb_launch_count", "mpcb_destroy",
                     "mpcb_set_limits"):
            getattr(L, name).argtypes = None
        _lib = L
    return _lib


def _p(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return a.ctypes.data_as(C.c_void_p)


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def default_settings(**kw):
    s = Settings()
    load_library().mpcb_default_settings(C.byref(s))
    for k, v in kw.items():
        setattr(s, k, v)
    return s


def default_limits(margins=None):
    l = Limits()
    load_library().mpcb_default_limits(C.byref(l))
    if margins is not None:
        for i, m in enumerate(margins):
            l.margins[i] = m
    return l


def time_nodes(s):
    n = load_library().mpcb_num_nodes(C.byref(s))
    t = np.zeros(n)
    rc = load_library().mpcb_time_nodes(C.byref(s), _p(t))
    if rc:
        raise MpcbError(f"mpcb_time_nodes failed ({rc})")
    return t


def generate_problems(lim, kind, seed, n):
    x0 = np.zeros((n, 14)); xT = np.zeros((n, 14))
    rc = load_library().mpcb_generate_problems(C.byref(lim), C.c_int(kind), C.c_ulonglong(seed), C.c_int(n), _p(x0), _p(xT))
    if rc:
        raise MpcbError(f"mpcb_generate_problems failed ({rc})")
    return x0, xT


def guess_batch(s, lim, x0, xT):
    L = load_library()
    x0 = _d(x0); xT = _d(xT)
    n = x0.shape[0]; N = L.mpcb_num_nodes(C.byref(s))
    gx = np.zeros((n, 14 * N)); gu = np.zeros((n, 7 * N)); gT = np.zeros(n)
    rc = L.mpcb_guess_batch(C.byref(s), C.byref(lim), C.c_int(n), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT))
    if rc:
        raise MpcbError(f"mpcb_guess_batch failed ({rc})")
    return gx, gu, gT


class OtgTraj(C.Structure):
    """mpcb_otg_traj (include/mpcb.h): the jerk-limited synchronised profile (Ruckig's Trajectory<7> in the reference)."""
    _fields_ = [("T", C.c_double), ("status", C.c_int), ("q0", C.c_double * 7), ("v0", C.c_double * 7), ("vc", C.c_double * 7),
                ("t", (C.c_double * 7) * 7), ("j", (C.c_double * 7) * 7)]


def otg_calculate(lim, x0, xT):
    """Ruckig::calculate (motionPlanner.cpp:149): returns (status, trajectory); status < 0 means no profile."""
    x0 = _d(x0); xT = _d(xT)
    tr = OtgTraj()
    rc = load_library().mpcb_otg_calculate(C.byref(lim), _p(x0), _p(xT), C.byref(tr))
    return rc, tr


def otg_at_time(tr, t):
    """Trajectory::at_time (motionPlanner.cpp:160): position, velocity, acceleration of the 7 joints at time t."""
    q = np.zeros(7); v = np.zeros(7); a = np.zeros(7)
    load_library().mpcb_otg_at_time(C.byref(tr), C.c_double(t), _p(q), _p(v), _p(a))
    return q, v, a


class Solver:
    """One handle on one CUDA device (replaces the `mpc_t mpc` member of the reference's MotionPlanner)."""

    def __init__(self, settings=None, limits=None, device=0):
        self.L = load_library()
        self.settings = settings if settings is not None else default_settings()
        self.limits = limits if limits is not None else default_limits()
        self.h = C.c_void_p()
        rc = self.L.mpcb_create(C.byref(self.settings), C.c_int(device), C.byref(self.h))
        if rc:
            msg = self.L.mpcb_last_error(None)
            raise MpcbError(f"mpcb_create failed ({rc}): {msg.decode() if msg else ''}")
        self.set_limits(self.limits)
        self.N = self.L.mpcb_num_nodes(C.byref(self.settings))
        self.nvar = self.L.mpcb_num_var(C.byref(self.settings))
        self.ndual = self.L.mpcb_num_dual(C.byref(self.settings))

    def _check(self, rc, what):
        if rc:
            msg = self.L.mpcb_last_error(self.h)
            raise MpcbError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")

    def set_limits(self, lim):
        self.limits = lim
        self._check(self.L.mpcb_set_limits(self.h, C.byref(lim)), "mpcb_set_limits")

    def close(self):
        if self.h:
            self.L.mpcb_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def solve(self, x0, xT, gx, gu, gT, glam=None, want_lam=False):
        """Host-memory call (copies inside).  Returns sol_x (n,14N), sol_u (n,7N), sol_T (n), sol_lam or None, stats (structured)."""
        x0 = _d(np.atleast_2d(x0)); xT = _d(np.atleast_2d(xT)); gx = _d(np.atleast_2d(gx)); gu = _d(np.atleast_2d(gu)); gT = _d(np.atleast_1d(gT))
        n = gT.shape[0]
        assert x0.shape == (n, 14) and xT.shape == (n, 14) and gx.shape == (n, 14 * self.N) and gu.shape == (n, 7 * self.N)
        glam = _d(np.atleast_2d(glam)) if glam is not None else None
        sx = np.zeros((n, 14 * self.N)); su = np.zeros((n, 7 * self.N)); sT = np.zeros(n)
        slam = np.zeros((n, self.ndual)) if want_lam else None
        st = np.zeros(n, dtype=STATS_DTYPE)
        rc = self.L.mpcb_solve_batch(self.h, C.c_int(n), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT), _p(glam), _p(sx), _p(su), _p(sT), _p(slam),
                                     _p(st), C.c_int(MEM_HOST), None)
        self._check(rc, "mpcb_solve_batch")
        return sx, su, sT, slam, st

    def solve_device(self, n, x0, xT, gx, gu, gT, glam, sx, su, sT, slam, stats, stream=0):
        """Device-pointer call (ints = raw device addresses, e.g. torch.Tensor.data_ptr()); only enqueues work."""
        rc = self.L.mpcb_solve_batch(self.h, C.c_int(n), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT), _p(glam) if glam else None, _p(sx), _p(su),
                                     _p(sT), _p(slam) if slam else None, _p(stats) if stats else None, C.c_int(MEM_DEVICE),
                                     C.c_void_p(stream) if stream else None)
        self._check(rc, "mpcb_solve_batch")

    def solve_device_noalloc(self, n, x0, xT, gx, gu, gT, glam, sx, su, sT, slam, stats, stream=0):
        """mpcb_solve_batch_noalloc: solve_device over the library's static device scratch (n <= NOALLOC_MAX_OCPS of the base shape);
        makes no allocation.  Calls that overlap on the GPU share (and race on) that scratch."""
        rc = self.L.mpcb_solve_batch_noalloc(self.h, C.c_int(n), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT), _p(glam) if glam else None, _p(sx),
                                             _p(su), _p(sT), _p(slam) if slam else None, _p(stats) if stats else None,
                                             C.c_void_p(stream) if stream else None)
        self._check(rc, "mpcb_solve_batch_noalloc")

    def guess(self, x0, xT):
        """warm_start_RK for a batch on the device (mpcb_guess_batch_gpu, host arrays in and out)."""
        x0 = _d(np.atleast_2d(x0)); xT = _d(np.atleast_2d(xT))
        n = x0.shape[0]
        gx = np.zeros((n, 14 * self.N)); gu = np.zeros((n, 7 * self.N)); gT = np.zeros(n)
        rc = self.L.mpcb_guess_batch_gpu(self.h, C.c_int(n), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT), C.c_int(MEM_HOST), None)
        self._check(rc, "mpcb_guess_batch_gpu")
        return gx, gu, gT

    def guess_device(self, n, x0, xT, gx, gu, gT, stream=0):
        """Device-pointer form (raw addresses); only enqueues the kernel."""
        rc = self.L.mpcb_guess_batch_gpu(self.h, C.c_int(n), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT), C.c_int(MEM_DEVICE),
                                         C.c_void_p(stream) if stream else None)
        self._check(rc, "mpcb_guess_batch_gpu")

    def generate(self, kind, seed, n, first=0):
        """Problem generation on the device (mpcb_generate_problems_gpu): problems [first, first + n) of the batch of `seed`."""
        x0 = np.zeros((n, 14)); xT = np.zeros((n, 14))
        rc = self.L.mpcb_generate_problems_gpu(self.h, C.c_int(kind), C.c_ulonglong(seed), C.c_ulonglong(first), C.c_int(n), _p(x0), _p(xT),
                                               C.c_int(MEM_HOST), None)
        self._check(rc, "mpcb_generate_problems_gpu")
        return x0, xT

    def generate_device(self, kind, seed, n, x0, xT, first=0, stream=0):
        """Device-pointer form (raw addresses); only enqueues the kernel."""
        rc = self.L.mpcb_generate_problems_gpu(self.h, C.c_int(kind), C.c_ulonglong(seed), C.c_ulonglong(first), C.c_int(n), _p(x0), _p(xT),
                                               C.c_int(MEM_DEVICE), C.c_void_p(stream) if stream else None)
        self._check(rc, "mpcb_generate_problems_gpu")

    def sample(self, sx, su, sT, target=None, npts=200, want_traj=True, want_acc=True):
        sx = _d(np.atleast_2d(sx)); su = _d(np.atleast_2d(su)); sT = _d(np.atleast_1d(sT))
        n = sT.shape[0]
        traj = np.zeros((n, npts + 1, 29)) if want_traj else None
        acc = np.zeros(n, dtype=ACCOUNT_DTYPE) if want_acc else None
        tg = _d(np.atleast_2d(target)) if target is not None else None
        rc = self.L.mpcb_sample_batch(self.h, C.c_int(n), _p(sx), _p(su), _p(sT), _p(tg), C.c_int(npts), _p(traj), _p(acc), C.c_int(MEM_HOST), None)
        self._check(rc, "mpcb_sample_batch")
        return traj, acc

    def sample_device(self, n, sx, su, sT, target, npts, traj, acc, stream=0):
        """Device-pointer form of the read-out + accounting (raw addresses; traj or acc may be 0); only enqueues the kernel."""
        rc = self.L.mpcb_sample_batch(self.h, C.c_int(n), _p(sx), _p(su), _p(sT), _p(target) if target else None, C.c_int(npts),
                                      _p(traj) if traj else None, _p(acc) if acc else None, C.c_int(MEM_DEVICE),
                                      C.c_void_p(stream) if stream else None)
        self._check(rc, "mpcb_sample_batch")

    def debug_node_constraints(self, qva):
        """Per-node device functions on n states (q, v, a): (g from the value pass, g from the derivative pass, Jg (n, 8, 22))."""
        qva = _d(np.atleast_2d(qva)); n = qva.shape[0]
        gv = np.zeros((n, 8)); gj = np.zeros((n, 8)); Jg = np.zeros((n, 8, 22))
        self._check(self.L.mpcb_debug_node_constraints(self.h, C.c_int(n), _p(qva), _p(gv), _p(gj), _p(Jg)), "mpcb_debug_node_constraints")
        return gv, gj, Jg

    def sample_profile(self, x0, xT, npts=200, want_traj=True, want_acc=True):
        """The warm-start profile sampled on the device with the benchmark's checks: (traj (n, npts+1, 29), account, durations)."""
        x0 = _d(np.atleast_2d(x0)); xT = _d(np.atleast_2d(xT)); n = x0.shape[0]
        traj = np.zeros((n, npts + 1, 29)) if want_traj else None
        acc = np.zeros(n, dtype=ACCOUNT_DTYPE) if want_acc else None
        T = np.zeros(n)
        rc = self.L.mpcb_sample_profile_batch(self.h, C.c_int(n), _p(x0), _p(xT), C.c_int(npts), _p(traj), _p(acc), _p(T), C.c_int(MEM_HOST), None)
        self._check(rc, "mpcb_sample_profile_batch")
        return traj, acc, T

    def last_solve_ms(self):
        return self.L.mpcb_last_solve_ms(self.h)

    def launch_count(self):
        return self.L.mpcb_launch_count(self.h)

    def last_kernel_ms(self, kind):
        """Device time of the last solve split by kernel kind: 0 = SQP stage kernels, 1 = ADMM kernel."""
        return self.L.mpcb_last_kernel_ms(self.h, kind)

    def last_kernel_launches(self, kind):
        return self.L.mpcb_last_kernel_launches(self.h, kind)

    def fp64_tflops(self):
        return self.L.mpcb_measure_fp64_tflops(self.h)
