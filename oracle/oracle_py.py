"""ctypes binding of the CPU oracle (oracle/liboracle_cpu.so).

TEST INFRASTRUCTURE -- may only be imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  PARITY UNPINNED for the SQP/ADMM part (see oracle.h).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle_cpu.so")


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


def build(force=False):
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "oracle.cpp")):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.orc_solve.restype = C.c_int
        _lib.orc_qp_admm.restype = C.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def default_settings(**kw):
    s = Settings()
    lib().orc_default_settings(C.byref(s))
    for k, v in kw.items():
        setattr(s, k, v)
    return s


def default_limits(margins=None):
    l = Limits()
    lib().orc_default_limits(C.byref(l))
    if margins is not None:
        for i, m in enumerate(margins):
            l.margins[i] = m
    return l


def fk(q):
    R = np.zeros((7, 3, 3)); p = np.zeros((7, 3))
    lib().orc_fk(_p(_d(q)), _p(R), _p(p))
    return R, p


def rnea(q, v, a, gravity=True):
    tau = np.zeros(7)
    (lib().orc_rnea if gravity else lib().orc_rnea_zero_gravity)(_p(_d(q)), _p(_d(v)), _p(_d(a)), _p(tau))
    return tau


def rnea_derivatives(q, v, a):
    tau = np.zeros(7); dq = np.zeros((7, 7)); dv = np.zeros((7, 7)); M = np.zeros((7, 7))
    lib().orc_rnea_derivatives(_p(_d(q)), _p(_d(v)), _p(_d(a)), _p(tau), _p(dq), _p(dv), _p(M))
    return tau, dq, dv, M


def crba(q):
    M = np.zeros((7, 7))
    lib().orc_crba(_p(_d(q)), _p(M))
    return M


def tool_height(q, jac=True):
    z = C.c_double(); Jz = np.zeros(7)
    lib().orc_tool_height(_p(_d(q)), C.byref(z), _p(Jz) if jac else None)
    return z.value, Jz


def frame_jacobian(q, offset):
    J = np.zeros((6, 7))
    lib().orc_frame_jacobian_world_aligned(_p(_d(q)), _p(_d(offset)), _p(J))
    return J


def node_constraints(q, v, a, jac=True):
    g = np.zeros(8); Jg = np.zeros((8, 22))
    lib().orc_node_constraints(_p(_d(q)), _p(_d(v)), _p(_d(a)), _p(g), _p(Jg) if jac else None)
    return g, Jg


def dims(s):
    L = lib()
    nn = L.orc_num_nodes(C.byref(s)); nvar = L.orc_num_var(C.byref(s)); neq = L.orc_num_eq(C.byref(s)); nineq = L.orc_num_ineq(C.byref(s))
    return dict(nn=nn, nvar=nvar, neq=neq, nineq=nineq, na=neq + nineq, ndual=neq + nineq + nvar)


def time_nodes(s):
    t = np.zeros(dims(s)["nn"])
    lib().orc_time_nodes(C.byref(s), _p(t))
    return t


def diff_matrix(s):
    D = np.zeros((s.poly_order + 1, s.poly_order + 1))
    lib().orc_diff_matrix(C.byref(s), _p(D))
    return D


def bounds(s, lim, x0, xT):
    d = dims(s)
    lbx = np.zeros(d["nvar"]); ubx = np.zeros(d["nvar"]); lbg = np.zeros(8); ubg = np.zeros(8)
    lib().orc_bounds(C.byref(s), C.byref(lim), _p(_d(x0)), _p(_d(xT)), _p(lbx), _p(ubx), _p(lbg), _p(ubg))
    return lbx, ubx, lbg, ubg


def eval_constraints(s, z):
    d = dims(s)
    c = np.zeros(d["neq"]); g = np.zeros(d["nineq"])
    lib().orc_eval_constraints(C.byref(s), _p(_d(z)), _p(c), _p(g))
    return c, g


def linearise(s, z, lam):
    d = dims(s)
    c = np.zeros(d["neq"]); g = np.zeros(d["nineq"]); Jg = np.zeros((d["nn"], 8, 22)); Hd = np.zeros(d["nvar"]); Ha = np.zeros(d["nvar"])
    lib().orc_linearise(C.byref(s), _p(_d(z)), _p(_d(lam)), _p(c), _p(g), _p(Jg), _p(Hd), _p(Ha))
    return c, g, Jg, Hd, Ha


def qp_dense(s, z, lam):
    d = dims(s)
    A = np.zeros((d["na"], d["nvar"])); H = np.zeros((d["nvar"], d["nvar"]))
    lib().orc_qp_dense(C.byref(s), _p(_d(z)), _p(_d(lam)), _p(A), _p(H))
    return A, H


def qp_admm(s, z_lin, Jg, Hd, Ha, h, lA, uA, lb, ub):
    d = dims(s)
    x = np.zeros(d["nvar"]); y = np.zeros(d["ndual"]); it = C.c_int(); res = np.zeros(4)
    st = lib().orc_qp_admm(C.byref(s), _p(_d(z_lin)), _p(_d(Jg)), _p(_d(Hd)), _p(_d(Ha)), _p(_d(h)), _p(_d(lA)), _p(_d(uA)),
                           _p(_d(lb)), _p(_d(ub)), _p(x), _p(y), C.byref(it), _p(res))
    return st, it.value, x, y, res


def solve(s, lim, x0, xT, gx, gu, gT, glam=None, want_lam=False):
    d = dims(s)
    sx = np.zeros(14 * d["nn"]); su = np.zeros(7 * d["nn"]); sT = C.c_double(); st = Stats()
    slam = np.zeros(d["ndual"]) if want_lam else None
    lib().orc_solve(C.byref(s), C.byref(lim), _p(_d(x0)), _p(_d(xT)), _p(_d(gx)), _p(_d(gu)), C.c_double(gT),
                    _p(_d(glam)) if glam is not None else None, _p(sx), _p(su), C.byref(sT), _p(slam) if want_lam else None, C.byref(st))
    return sx, su, sT.value, slam, st


def solve_batch(s, lim, x0, xT, gx, gu, gT, nthreads=1):
    d = dims(s)
    n = len(gT)
    x0 = _d(x0); xT = _d(xT); gx = _d(gx); gu = _d(gu); gT = _d(gT)
    sx = np.zeros((n, 14 * d["nn"])); su = np.zeros((n, 7 * d["nn"])); sT = np.zeros(n); st = (Stats * n)()
    lib().orc_solve_batch(C.byref(s), C.byref(lim), C.c_int(n), C.c_int(nthreads), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT),
                          _p(sx), _p(su), _p(sT), st)
    return sx, su, sT, st


def sample_solution(s, sx, su, T, npts=200):
    traj = np.zeros((npts + 1, 29))
    lib().orc_sample_solution(C.byref(s), _p(_d(sx)), _p(_d(su)), C.c_double(T), C.c_int(npts), _p(traj))
    return traj


def account(lim, traj, target):
    acc = Account()
    traj = _d(traj)
    lib().orc_account_trajectory(C.byref(lim), _p(traj), C.c_int(traj.shape[0] - 1), _p(_d(target)), C.byref(acc))
    return acc


def interp_state(s, sx, su, t01):
    q = np.zeros(7); v = np.zeros(7); a = np.zeros(7)
    lib().orc_interp_state(C.byref(s), _p(_d(sx)), _p(_d(su)), C.c_double(t01), _p(q), _p(v), _p(a))
    return q, v, a
