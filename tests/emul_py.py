"""ctypes binding of the TEST-ONLY host emulation of the kernel phases (tests/emul/libmpcb_emul.so)."""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle import oracle_py as O

_HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emul")
_lib = None


def lib():
    global _lib
    if _lib is None:
        subprocess.check_call(["make", "-s", "-C", _HERE])
        _lib = C.CDLL(os.path.join(_HERE, "libmpcb_emul.so"))
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def solve(s, lim, x0, xT, gx, gu, gT, glam=None, want_lam=False):
    d = O.dims(s)
    x0 = _d(np.atleast_2d(x0)); xT = _d(np.atleast_2d(xT)); gx = _d(np.atleast_2d(gx)); gu = _d(np.atleast_2d(gu)); gT = _d(np.atleast_1d(gT))
    n = len(gT)
    sx = np.zeros((n, 14 * d["nn"])); su = np.zeros((n, 7 * d["nn"])); sT = np.zeros(n); st = (O.Stats * n)()
    slam = np.zeros((n, d["ndual"])) if want_lam else None
    glam = _d(np.atleast_2d(glam)) if glam is not None else None
    rc = lib().emul_solve(C.byref(s), C.byref(lim), C.c_int(n), _p(x0), _p(xT), _p(gx), _p(gu), _p(gT), _p(glam), _p(sx), _p(su), _p(sT),
                          _p(slam), st)
    assert rc == 0, rc
    return sx, su, sT, slam, st


def kkt_apply(s, lim, x0, xT, z, lam, rhs):
    d = O.dims(s)
    out = np.zeros(d["nvar"]); rhoA = np.zeros(d["na"]); rhoB = np.zeros(d["nvar"])
    lib().emul_kkt_apply(C.byref(s), C.byref(lim), _p(_d(x0)), _p(_d(xT)), _p(_d(z)), _p(_d(lam)), _p(_d(rhs)), _p(out), _p(rhoA), _p(rhoB))
    return out, rhoA, rhoB
