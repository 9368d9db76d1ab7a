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

    def rollout(self, ticks, cmd, planner_every=4, planner_params=None, nthreads=0, record=True):
        """closed loop of `ticks` control ticks from the current state: restated planner (constant /cmd_vel per robot, cmd (3, N)) ->
        controller -> plant (DEFINED in oracle_capi.cpp: no reference counterpart).  Returns wall seconds and per-tick records."""
        P = dict(DEFAULT_PLANNER_PARAMS)
        if planner_params:
            P.update(planner_params)
        pp = np.array([P[k] for k in PLANNER_PARAM_ORDER], dtype=np.float64)
        cmd = _f64(cmd); assert cmd.shape == (3, self.N)
        N, K = self.N, int(ticks)
        rec = dict(tau=np.zeros((K, 12, N)), rc=np.zeros((K, N), np.int32), nwsr=np.zeros((K, N), np.int32), q=np.zeros((K, 12, N)), pos=np.zeros((K, 3, N))) if record else {}
        f = lib().orc_rollout; f.restype = C.c_double
        f.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 5
        secs = f(self._h, K, int(planner_every), _p(pp), _p(cmd), int(nthreads), *[_p(rec.get(k)) for k in ("tau", "rc", "nwsr", "q", "pos")])
        rec["seconds"] = secs
        return rec

    def state(self):
        N = self.N
        o = dict(q=np.zeros((12, N)), qd=np.zeros((12, N)), base_pos=np.zeros((3, N)), base_quat=np.zeros((4, N)), base_twist=np.zeros((6, N)))
        lib().orc_get_state(self._h, *[_p(o[k]) for k in ("q", "qd", "base_pos", "base_quat", "base_twist")])
        return o

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


PLANNER_PARAM_ORDER = ["step_length", "height_control_point", "x_offset", "y_offset", "step_duration", "body_height", "body_final_velocity", "dt"]
DEFAULT_PLANNER_PARAMS = dict(step_length=0.1, height_control_point=0.1, x_offset=0.50, y_offset=0.33, step_duration=0.2, body_height=0.50,
                              body_final_velocity=0.40, dt=0.01)                       # config/params_planner.yaml


def planner_run(cmd, params=None):
    """restated MotionPlanner::plannerLoop for len(cmd) planner periods; cmd (K, 3) = /cmd_vel (linear.x, linear.y, angular.z) per period.
    Returns published (K,), and the latched message at the end of each period: pose/vel/acc (K, 6), sw_pos/sw_vel/sw_acc (K, 12), contacts (K, 4)."""
    P = dict(DEFAULT_PLANNER_PARAMS)
    if params:
        P.update(params)
    pp = np.array([P[k] for k in PLANNER_PARAM_ORDER], dtype=np.float64)
    cmd = _f64(cmd); K = cmd.shape[0]
    pub = np.zeros(K, np.uint8); msg = np.zeros((K, 54)); con = np.zeros((K, 4), np.uint8)
    lib().orc_planner_run(_p(pp), K, _p(cmd), _p(pub), _p(msg), _p(con))
    return dict(published=pub, pose=msg[:, 0:6], vel=msg[:, 6:12], acc=msg[:, 12:18], sw_pos=msg[:, 18:30], sw_vel=msg[:, 30:42], sw_acc=msg[:, 42:54], contacts=con)
