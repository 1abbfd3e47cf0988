"""CPU ORACLE (test infrastructure) — ctypes front-end of oracle/_build/liborc.so (the restatement; no compiled reference exists).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.  It reuses the product's generic ctypes wrapper
(lidarslam_b200.capi.CApi) with the symbol prefix "orc_" so both sides are driven
by identical calls."""
import ctypes as C
import os
import subprocess

import numpy as np

from lidarslam_b200 import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def load():
    global _LIB
    if _LIB is None:
        p = os.path.join(_HERE, "_build", "liborc.so")
        if not os.path.exists(p):
            build()
        _LIB = C.CDLL(p)
    return _LIB


class Oracle(capi.CApi):
    def __init__(self, nb_threads=None):
        super().__init__(load(), "orc_")
        if nb_threads is not None:
            p = self.GetParams()
            p.nb_threads = nb_threads
            self.SetParams(p)


def max_threads():
    return int(load().orc_max_threads())


def default_params():
    return capi.default_params(load())


def default_ke_params():
    return capi.default_ke_params(load())


def detmath(fn, a, b=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b if b is not None else np.zeros_like(a), dtype=np.float64)
    out = np.zeros_like(a)
    load().orc_detmath(C.c_int(fn), capi._ptr(a), capi._ptr(b), capi._ptr(out), C.c_long(a.size))
    return out


def eigen33(mats, as_float=False):
    m = np.ascontiguousarray(mats, dtype=np.float64).reshape(-1, 9)
    ev = np.zeros((len(m), 3))
    vec = np.zeros((len(m), 9))
    load().orc_eigen33(capi._ptr(m), capi._ptr(ev), capi._ptr(vec), C.c_long(len(m)), C.c_int(1 if as_float else 0))
    return ev, vec.reshape(-1, 3, 3)


def lm_terms(A, P, X, w, x, sat):
    """28 normal-equation terms per record (common/lm_core.h::lm_eval_record)."""
    A = np.ascontiguousarray(A, dtype=np.float64).reshape(-1, 9)
    P = np.ascontiguousarray(P, dtype=np.float64).reshape(-1, 3)
    X = np.ascontiguousarray(X, dtype=np.float64).reshape(-1, 3)
    w = np.ascontiguousarray(w, dtype=np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.zeros((len(A), 28))
    load().orc_lm_terms(capi._ptr(A), capi._ptr(P), capi._ptr(X), capi._ptr(w), C.c_long(len(A)), capi._ptr(x), C.c_double(sat), capi._ptr(out))
    return out


def lm_solve(A, P, X, w, x0, sat, max_iter=15, two_d=False, threads=1, wheel=None, gravity=None):
    """The oracle's LocalOptimizer::Solve on raw residual records.
    wheel = (previous position xyz, distance, weight), gravity = (reference dir, current dir, weight): the extra
    residual blocks of Slam.cxx:1123-1131."""
    sensor = np.zeros(14)
    if wheel is not None:
        sensor[0], sensor[1:4], sensor[4], sensor[5] = 1, wheel[0], wheel[1], wheel[2]
    if gravity is not None:
        sensor[6], sensor[7:10], sensor[10:13], sensor[13] = 1, gravity[0], gravity[1], gravity[2]
    A = np.ascontiguousarray(A, dtype=np.float64).reshape(-1, 9)
    P = np.ascontiguousarray(P, dtype=np.float64).reshape(-1, 3)
    X = np.ascontiguousarray(X, dtype=np.float64).reshape(-1, 3)
    w = np.ascontiguousarray(w, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    x = np.zeros(6)
    cost, acc, term = C.c_double(), C.c_int(), C.c_int()
    it = load().orc_lm_solve_sensors(capi._ptr(A), capi._ptr(P), capi._ptr(X), capi._ptr(w), C.c_long(len(A)), capi._ptr(x0),
                                     C.c_double(sat), C.c_int(max_iter), C.c_int(1 if two_d else 0), C.c_int(threads), capi._ptr(sensor),
                                     capi._ptr(x), C.byref(cost), C.byref(acc), C.byref(term))
    return dict(x=x, cost=cost.value, iterations=it, accepted=acc.value, term=term.value)


def lm_solve_trace(A, P, X, w, x0, sat, max_iter=15, two_d=False, threads=1):
    """lm_solve plus the per-evaluation trace (rows: iteration, cost at x, cost at the candidate, radius, accepted) and the
    upper triangle of the normal matrix at the solution."""
    A = np.ascontiguousarray(A, dtype=np.float64).reshape(-1, 9)
    P = np.ascontiguousarray(P, dtype=np.float64).reshape(-1, 3)
    X = np.ascontiguousarray(X, dtype=np.float64).reshape(-1, 3)
    w = np.ascontiguousarray(w, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    x, H = np.zeros(6), np.zeros(21)
    trace = np.zeros((max_iter + 8, 5))
    cost, acc, term, rows = C.c_double(), C.c_int(), C.c_int(), C.c_int()
    it = load().orc_lm_solve_trace(capi._ptr(A), capi._ptr(P), capi._ptr(X), capi._ptr(w), C.c_long(len(A)), capi._ptr(x0), C.c_double(sat),
                                   C.c_int(max_iter), C.c_int(1 if two_d else 0), C.c_int(threads), capi._ptr(x), C.byref(cost),
                                   C.byref(acc), C.byref(term), capi._ptr(trace), C.c_int(len(trace)), C.byref(rows), capi._ptr(H))
    return dict(x=x, cost=cost.value, iterations=it, accepted=acc.value, term=term.value, trace=trace[:rows.value], H=H)


def registration_error(H21, two_d=False):
    """LocalOptimizer::EstimateRegistrationError on a given 6x6 normal matrix (21 upper-triangle entries, row-major)."""
    H = np.ascontiguousarray(H21, dtype=np.float64)
    err = capi.RegistrationError()
    load().orc_registration_error(capi._ptr(H), C.c_int(1 if two_d else 0), C.byref(err))
    return err
