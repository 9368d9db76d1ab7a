"""ORACLE -- TEST INFRASTRUCTURE ONLY.

ctypes front-end of oracle/_ref/liboracle.so (restated iDynTree subset + restated controller algebra linked to the
REAL vendored qpOASES 3.2.2).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product package never does.
Arrays are structure-of-arrays with the instance index fastest: shape (K, N) C-contiguous numpy arrays.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "liboracle.so")
_lib = None

PARAM_ORDER = ["friction", "loopRate", "maxTorque", "kpValue", "kpValueZ", "kdValue", "kiValue", "kpSwingValue",
               "kdSwingValue", "slackWeight"]
DEFAULT_PARAMS = dict(friction=1.0, loopRate=400.0, maxTorque=80.0, kpValue=6000.0, kpValueZ=10000.0, kdValue=1800.0,
                      kiValue=0.0, kpSwingValue=250.0, kdSwingValue=20.0, slackWeight=1000.0,
                      initialReferencePose=[0.0, 0.0, 0.5, 0.0, 0.0, 0.0])   # config/params_controller.yaml


def build():
    """compile the oracle (needs /root/reference for the qpOASES sources/headers unless prebuilt files exist)."""
    if os.path.isdir("/root/reference/lib/qpOASES/src"):
        subprocess.check_call(["make", "-s", "-C", _HERE, "CXX=g++"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    elif not os.path.exists(_LIB_PATH):
        raise RuntimeError("oracle/_ref/liboracle.so missing and the reference mount is not available to build it")
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_create.restype = C.c_void_p
        _lib.orc_create.argtypes = [C.c_int, C.c_int, C.c_void_p]
        _lib.orc_destroy.argtypes = [C.c_void_p]
        _lib.orc_step.restype = C.c_double
        _lib.orc_step.argtypes = [C.c_void_p, C.c_int, C.c_int]
        _lib.orc_max_threads.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def pack_params(params=None):
    P = dict(DEFAULT_PARAMS)
    if params:
        P.update(params)
    return np.array([P[k] for k in PARAM_ORDER] + list(P["initialReferencePose"]), dtype=np.float64)


class OracleBatch:
    """N independent reference-equivalent controllers (each with its own qpOASES SQProblem hot-start state)."""

    def __init__(self, N, params=None, lumped=False):
        self.N = N
        self._pp = pack_params(params)
        self._h = C.c_void_p(lib().orc_create(N, int(lumped), _p(self._pp)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_destroy(self._h)
            self._h = None

    def set_state(self, q, qd, base_pos, base_quat, base_twist):
        a = [_f64(x) for x in (q, qd, base_pos, base_quat, base_twist)]
        for x, k in zip(a, (12, 12, 3, 4, 6)):
            assert x.shape == (k, self.N), (x.shape, k)
        lib().orc_set_state(self._h, *[_p(x) for x in a])

    def set_reference(self, pose, vel, acc, sw_pos, sw_vel, sw_acc, contacts, switching=None):
        a = [_f64(x) for x in (pose, vel, acc, sw_pos, sw_vel, sw_acc)]
        c = np.ascontiguousarray(contacts, dtype=np.uint8)
        assert c.shape == (4, self.N)
        s = None if switching is None else np.ascontiguousarray(switching, dtype=np.uint8)
        lib().orc_set_reference(self._h, *[_p(x) for x in a], _p(c), _p(s))

    def step(self, nthreads=0, assemble_only=False):
        return lib().orc_step(self._h, nthreads, int(assemble_only))

    def outputs(self):
        N = self.N
        o = dict(tau=np.zeros((12, N)), grf=np.zeros((12, N)), x=np.zeros((42, N)), rc=np.zeros(N, np.int32),
                 nwsr=np.zeros(N, np.int32), ws=np.zeros((112, N), np.int8), y=np.zeros((112, N)))
        lib().orc_get_outputs(self._h, _p(o["tau"]), _p(o["grf"]), _p(o["x"]), _p(o["rc"]), _p(o["nwsr"]), _p(o["ws"]), _p(o["y"]))
        return o

    def qp(self, i):
        o = dict(H=np.zeros((42, 42)), g=np.zeros(42), A=np.zeros((70, 42)), lbA=np.zeros(70), ubA=np.zeros(70))
        lib().orc_get_qp(self._h, i, _p(o["H"]), _p(o["g"]), _p(o["A"]), _p(o["lbA"]), _p(o["ubA"]))
        return o

    def intermediates(self, i):
        o = dict(M=np.zeros((18, 18)), T=np.zeros((18, 18)), Mc=np.zeros((18, 18)), Jst=np.zeros((12, 18)), Jsw=np.zeros((12, 18)),
                 bias_c=np.zeros(18), coriolis=np.zeros(18), com=np.zeros(3), vcom=np.zeros(6), pose=np.zeros(6),
                 footJ=np.zeros((12, 18)), footPos=np.zeros(12), footVel=np.zeros(12), TdotInv=np.zeros((18, 18)), W=np.zeros(6))
        lib().orc_get_intermediates(self._h, i, *[_p(o[k]) for k in ("M", "T", "Mc", "Jst", "Jsw", "bias_c", "coriolis", "com",
                                                                     "vcom", "pose", "footJ", "footPos", "footVel", "TdotInv", "W")])
        return o


def rbd_eval(pos, quat, twist, q, dq, lumped=False, acc18=None):
    o = dict(M=np.zeros((18, 18)), J=np.zeros((4, 6, 18)), h=np.zeros(18), g=np.zeros(18), com=np.zeros(3), vcom=np.zeros(3),
             footPos=np.zeros(12), footVel=np.zeros((4, 6)), id=np.zeros(18))
    a = [_f64(x) for x in (pos, quat, twist, q, dq)]
    acc = None if acc18 is None else _f64(acc18)
    lib().orc_rbd_eval(int(lumped), *[_p(x) for x in a], _p(o["M"]), _p(o["J"]), _p(o["h"]), _p(o["g"]), _p(o["com"]), _p(o["vcom"]),
                       _p(o["footPos"]), _p(o["footVel"]), _p(acc), _p(o["id"]) if acc is not None else None)
    return o


def qp_sequence(H, g, A, lbA, ubA, nwsr_max=100):
    """real qpOASES SQProblem: init on QP 0 then hotstart on the rest, options exactly as wbc.cpp:517-521."""
    H, g, A, lbA, ubA = [_f64(x) for x in (H, g, A, lbA, ubA)]
    T = H.shape[0]
    o = dict(x=np.zeros((T, 42)), y=np.zeros((T, 112)), ws=np.zeros((T, 112), np.int8), rc=np.zeros(T, np.int32), nwsr=np.zeros(T, np.int32))
    lib().orc_qp_sequence(T, _p(H), _p(g), _p(A), _p(lbA), _p(ubA), nwsr_max, _p(o["x"]), _p(o["y"]), _p(o["ws"]), _p(o["rc"]), _p(o["nwsr"]))
    return o
