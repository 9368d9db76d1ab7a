"""Batched ANYmal whole-body controller for NVIDIA B200 (sm_100a) -- Python binding of the C ABI (include/wbc_b200.h).

The product is the CUDA shared library `libwbc_b200.so` (csrc/wbc_kernels.cu); this module is a thin ctypes front-end used by
the tests and bench.py.  The C++ mirror of the reference class is include/whole_body_controller_b200.hpp.
There is no CPU fallback: constructing a controller without the library or without a CUDA device raises.
"""
import ctypes as C
import os

import numpy as np

from .build import build_library, LIB_PATH, PROF_LIB_PATH, EXACT_LIB_PATH, F32_LIB_PATH

NUM_JOINTS, NUM_LEGS, QP_NV, QP_NC = 12, 4, 42, 70          # whole_body_controller.hpp:27-32 of the reference
_libs = {}

EXPORTS = ["wbc_default_params", "wbc_create", "wbc_destroy", "wbc_reset", "wbc_set_state", "wbc_set_reference", "wbc_update_state",
           "wbc_solve_qp", "wbc_solve_qp_static", "wbc_compute_joint_torques", "wbc_step", "wbc_synchronize", "wbc_get_outputs", "wbc_get_dense_qp",
           "wbc_planner_default_params", "wbc_planner_create", "wbc_planner_reset", "wbc_planner_set_cmd_vel", "wbc_planner_tick", "wbc_get_reference",
           "wbc_get_state", "wbc_plant_enable", "wbc_plant_step", "wbc_rollout", "wbc_rollout_stats", "wbc_num_robots", "wbc_launch_count", "wbc_last_kernel_ms", "wbc_profile", "wbc_measure_fp64_peak", "wbc_last_error", "wbc_get_duals", "wbc_warm_start_split", "wbc_set_option", "wbc_debug_counters", "wbc_measure_smem_peak",
           "wbc_multi_create", "wbc_multi_destroy", "wbc_multi_num_devices", "wbc_multi_num_robots", "wbc_multi_shard", "wbc_multi_handle", "wbc_multi_reset", "wbc_multi_set_state",
           "wbc_multi_set_reference", "wbc_multi_step", "wbc_multi_synchronize", "wbc_multi_get_outputs", "wbc_multi_last_gather_ms", "wbc_multi_last_error"]


class WbcParams(C.Structure):
    """wbc_params == WholeBodyController::params (reference whole_body_controller.hpp:91-104)."""
    _fields_ = [(k, C.c_double) for k in ("friction", "loopRate", "maxTorque", "kpValue", "kpValueZ", "kdValue", "kiValue",
                                          "kpSwingValue", "kdSwingValue", "slackWeight")] + [("initialReferencePose", C.c_double * 6)]


class WbcPlannerParams(C.Structure):
    """wbc_planner_params == MotionPlanner::params (reference motion_planner.hpp:40-51, config/params_planner.yaml)."""
    _fields_ = [(k, C.c_double) for k in ("step_length", "height_control_point", "x_offset", "y_offset", "step_duration", "body_height",
                                          "body_final_velocity", "dt")]


def load_library(profiling=False, exact=False, fp32=False):
    """dlopen the CUDA library (profiling=True: the -DWBC_PROFILE build of the same sources; exact=True: the -DWBC_EXACT verification build that
    reproduces qpOASES' arithmetic operation by operation); raises if it has not been built (no fallback of any kind)."""
    path = F32_LIB_PATH if fp32 else (EXACT_LIB_PATH if exact else (PROF_LIB_PATH if profiling else LIB_PATH))      # fp32=True: the QP solver in single precision
    if path not in _libs:
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a). "
                               "The controller has no CPU fallback.")
        lib = C.CDLL(path)
        lib.wbc_create.argtypes = [C.c_int, C.c_int, C.POINTER(WbcParams), C.c_void_p, C.POINTER(C.c_void_p)]
        for name in ("wbc_destroy", "wbc_update_state", "wbc_solve_qp", "wbc_solve_qp_static", "wbc_compute_joint_torques", "wbc_step", "wbc_synchronize"):
            getattr(lib, name).argtypes = [C.c_void_p]
        lib.wbc_reset.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.wbc_set_state.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        lib.wbc_set_reference.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        lib.wbc_get_outputs.argtypes = [C.c_void_p] + [C.c_void_p] * 7
        lib.wbc_get_dense_qp.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5
        lib.wbc_num_robots.argtypes = [C.c_void_p]
        lib.wbc_launch_count.argtypes = [C.c_void_p]; lib.wbc_launch_count.restype = C.c_longlong
        lib.wbc_last_kernel_ms.argtypes = [C.c_void_p] + [C.POINTER(C.c_float)] * 3
        lib.wbc_last_error.argtypes = [C.c_void_p]; lib.wbc_last_error.restype = C.c_char_p
        lib.wbc_default_params.argtypes = [C.POINTER(WbcParams)]
        lib.wbc_profile.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.wbc_measure_fp64_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        lib.wbc_measure_smem_peak.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        lib.wbc_planner_default_params.argtypes = [C.POINTER(WbcPlannerParams)]
        lib.wbc_planner_create.argtypes = [C.c_void_p, C.POINTER(WbcPlannerParams)]
        lib.wbc_planner_reset.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.wbc_planner_set_cmd_vel.argtypes = [C.c_void_p] + [C.c_void_p] * 3
        lib.wbc_get_reference.argtypes = [C.c_void_p] + [C.c_void_p] * 9
        lib.wbc_get_state.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        for name in ("wbc_planner_tick", "wbc_plant_enable", "wbc_plant_step"):
            getattr(lib, name).argtypes = [C.c_void_p]
        lib.wbc_rollout.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.wbc_get_duals.argtypes = [C.c_void_p, C.c_void_p]
        lib.wbc_warm_start_split.argtypes = [C.c_void_p]
        lib.wbc_rollout_stats.argtypes = [C.c_void_p] + [C.c_void_p] * 4 + [C.c_int]
        lib.wbc_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        lib.wbc_debug_counters.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.wbc_multi_create.argtypes = [C.c_int, C.POINTER(C.c_int), C.c_int, C.POINTER(WbcParams), C.POINTER(C.c_void_p)]
        for name in ("wbc_multi_destroy", "wbc_multi_num_devices", "wbc_multi_num_robots", "wbc_multi_reset", "wbc_multi_step", "wbc_multi_synchronize"):
            getattr(lib, name).argtypes = [C.c_void_p]
        lib.wbc_multi_shard.argtypes = [C.c_void_p, C.c_int] + [C.POINTER(C.c_int)] * 3
        lib.wbc_multi_handle.argtypes = [C.c_void_p, C.c_int]; lib.wbc_multi_handle.restype = C.c_void_p
        lib.wbc_multi_set_state.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        lib.wbc_multi_set_reference.argtypes = [C.c_void_p] + [C.c_void_p] * 8
        lib.wbc_multi_get_outputs.argtypes = [C.c_void_p] + [C.c_void_p] * 7
        lib.wbc_multi_last_gather_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
        lib.wbc_multi_last_error.argtypes = [C.c_void_p]; lib.wbc_multi_last_error.restype = C.c_char_p
        _libs[path] = lib
    return _libs[path]


def default_params():
    p = WbcParams(); load_library().wbc_default_params(C.byref(p)); return p


def _ptr(a):
    """numpy array -> host pointer; torch CUDA tensor -> device pointer; int -> raw pointer; None -> NULL."""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):
        assert a.is_contiguous()
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))


class WholeBodyController:
    """Batch of N independent reference-equivalent controllers on one GPU.

    Method names follow the reference class (updateState / solveQP / computeJointTorques); the three ROS callbacks become
    setJointAndBaseState / setReference.  Arrays are (K, N) C-contiguous float64 (uint8 for flags): numpy (host) or torch CUDA.
    """

    numberOfJoints, numberOfLegs, qpNumberOfVariables, qpNumberOfConstraints = NUM_JOINTS, NUM_LEGS, QP_NV, QP_NC

    def __init__(self, n_robots, device=0, params=None, stream=None, profiling=False, exact=False, fp32=False):
        lib = self._lib = load_library(profiling, exact, fp32)
        self.N = int(n_robots)
        self._h = C.c_void_p()
        p = params if params is not None else default_params()
        rc = lib.wbc_create(self.N, int(device), C.byref(p), C.c_void_p(stream) if stream else None, C.byref(self._h))
        if rc != 0:
            self._h = C.c_void_p()
            raise RuntimeError(f"wbc_create failed with {rc} (CUDA device {device} unusable? there is no CPU fallback)")

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(f"wbc error {rc}: {self._lib.wbc_last_error(self._h).decode()}")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.wbc_destroy(self._h); self._h = C.c_void_p()

    __del__ = close

    def setInitialState(self, first=0, count=None):
        self._ck(self._lib.wbc_reset(self._h, first, self.N - first if count is None else count))

    def setJointAndBaseState(self, q, qd, base_pos, base_quat, base_twist):
        self._ck(self._lib.wbc_set_state(self._h, _ptr(q), _ptr(qd), _ptr(base_pos), _ptr(base_quat), _ptr(base_twist)))

    def setReference(self, pose, vel, acc, sw_pos, sw_vel, sw_acc, contacts, switching=None):
        self._ck(self._lib.wbc_set_reference(self._h, _ptr(pose), _ptr(vel), _ptr(acc), _ptr(sw_pos), _ptr(sw_vel), _ptr(sw_acc),
                                                  _ptr(contacts), _ptr(switching)))

    def updateState(self):
        self._ck(self._lib.wbc_update_state(self._h))

    def solveQP(self):
        self._ck(self._lib.wbc_solve_qp(self._h))

    def solveQPStatic(self):
        """wbc_solve_qp_static: solveQP without any allocation (scratch in a static device array of the library; at most 1024 robots)"""
        self._ck(self._lib.wbc_solve_qp_static(self._h))

    def computeJointTorques(self):
        self._ck(self._lib.wbc_compute_joint_torques(self._h))

    def step(self):
        self._ck(self._lib.wbc_step(self._h))

    def synchronize(self):
        self._ck(self._lib.wbc_synchronize(self._h))

    def getOutputs(self, tau=None, grf=None, x=None, rc=None, nwsr=None, ws=None, com=None):
        """copies into the given buffers (host or device); with no argument returns fresh numpy arrays of everything."""
        if all(v is None for v in (tau, grf, x, rc, nwsr, ws, com)):
            N = self.N
            o = dict(tau=np.zeros((12, N)), grf=np.zeros((12, N)), x=np.zeros((42, N)), rc=np.zeros(N, np.int32), nwsr=np.zeros(N, np.int32),
                     ws=np.zeros((112, N), np.int8), com=np.zeros((3, N)))
            self._ck(self._lib.wbc_get_outputs(self._h, *[_ptr(o[k]) for k in ("tau", "grf", "x", "rc", "nwsr", "ws", "com")]))
            return o
        self._ck(self._lib.wbc_get_outputs(self._h, _ptr(tau), _ptr(grf), _ptr(x), _ptr(rc), _ptr(nwsr), _ptr(ws), _ptr(com)))

    def getDuals(self, y=None):
        """QProblem::getDualSolution of every robot, (112, N): 42 bound multipliers then 70 constraint multipliers"""
        out = np.zeros((112, self.N)) if y is None else y
        self._ck(self._lib.wbc_get_duals(self._h, _ptr(out)))
        return out

    def setOption(self, key, value):
        """explicit switches for tests / measurements (wbc_set_option): warm_start_split, cohort, setup_canary"""
        self._ck(self._lib.wbc_set_option(self._h, key.encode(), int(value)))

    def debugCounters(self, reset=False):
        """exact build only: (full LI tests run, inside the shortcut's undecided band, shortcut verdict wrong)"""
        out = np.zeros(3, np.uint64); self._ck(self._lib.wbc_debug_counters(self._h, _ptr(out), int(reset))); return out

    def warmStartSplit(self):
        return bool(self._lib.wbc_warm_start_split(self._h))

    def denseQP(self, robot):
        o = dict(H=np.zeros((42, 42)), g=np.zeros(42), A=np.zeros((70, 42)), lbA=np.zeros(70), ubA=np.zeros(70))
        self._ck(self._lib.wbc_get_dense_qp(self._h, robot, *[_ptr(o[k]) for k in ("H", "g", "A", "lbA", "ubA")]))
        return o

    def launchCount(self):
        return int(self._lib.wbc_launch_count(self._h))

    def profile(self, enable=True, fetch=False):
        """per-robot clock64 breakdown of the last solve, shape (N, 10) int64: 7 clock64 phase counters, iterations, algorithmic flops, cold flag"""
        out = np.zeros((self.N, 10), np.int64) if fetch else None
        self._ck(self._lib.wbc_profile(self._h, int(enable), _ptr(out)))
        return out

    def measureFp64Peak(self):
        g = C.c_double(); self._ck(self._lib.wbc_measure_fp64_peak(self._h, C.byref(g))); return g.value

    def measureSmemPeak(self):
        g = C.c_double(); self._ck(self._lib.wbc_measure_smem_peak(self._h, C.byref(g))); return g.value

    def lastKernelMs(self):
        a, b, c = C.c_float(), C.c_float(), C.c_float()
        self._ck(self._lib.wbc_last_kernel_ms(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return dict(update_state=a.value, solve_qp=b.value, compute_joint_torques=c.value)

    # ---- planner, plant, closed-loop rollout (reference src/motion_planner.cpp) ----
    def plannerCreate(self, params=None):
        p = params
        if p is None:
            p = WbcPlannerParams(); self._lib.wbc_planner_default_params(C.byref(p))
        self._ck(self._lib.wbc_planner_create(self._h, C.byref(p)))

    def plannerReset(self, first=0, count=None):
        self._ck(self._lib.wbc_planner_reset(self._h, first, self.N - first if count is None else count))

    def plannerSetCmdVel(self, vx=None, vy=None, yaw_rate=None):
        self._ck(self._lib.wbc_planner_set_cmd_vel(self._h, _ptr(vx), _ptr(vy), _ptr(yaw_rate)))

    def plannerTick(self):
        self._ck(self._lib.wbc_planner_tick(self._h))

    def getReference(self, with_published=True):
        N = self.N
        o = dict(pose=np.zeros((6, N)), vel=np.zeros((6, N)), acc=np.zeros((6, N)), sw_pos=np.zeros((12, N)), sw_vel=np.zeros((12, N)), sw_acc=np.zeros((12, N)),
                 contacts=np.zeros((4, N), np.uint8), switching=np.zeros(N, np.uint8))
        pub = np.zeros(N, np.uint8) if with_published else None
        self._ck(self._lib.wbc_get_reference(self._h, *[_ptr(o[k]) for k in ("pose", "vel", "acc", "sw_pos", "sw_vel", "sw_acc", "contacts", "switching")], _ptr(pub)))
        if with_published:
            o["published"] = pub
        return o

    def getState(self):
        N = self.N
        o = dict(q=np.zeros((12, N)), qd=np.zeros((12, N)), base_pos=np.zeros((3, N)), base_quat=np.zeros((4, N)), base_twist=np.zeros((6, N)))
        self._ck(self._lib.wbc_get_state(self._h, *[_ptr(o[k]) for k in ("q", "qd", "base_pos", "base_quat", "base_twist")]))
        return o

    def plantEnable(self):
        self._ck(self._lib.wbc_plant_enable(self._h))

    def plantStep(self):
        self._ck(self._lib.wbc_plant_step(self._h))

    def rollout(self, ticks, planner_every=4):
        self._ck(self._lib.wbc_rollout(self._h, int(ticks), int(planner_every)))

    def rolloutStats(self, reset=False):
        N = self.N
        o = dict(ticks=np.zeros(N, np.int32), failed_ticks=np.zeros(N, np.int32), nwsr_sum=np.zeros(N, np.int32), nwsr_max=np.zeros(N, np.int32))
        self._ck(self._lib.wbc_rollout_stats(self._h, *[_ptr(o[k]) for k in ("ticks", "failed_ticks", "nwsr_sum", "nwsr_max")], int(reset)))
        return o


class MultiGpuController:
    """One batch of N robots sharded by robot index over several GPUs of the box inside ONE process (wbc_multi_*, include/wbc_b200.h): device k
    owns a contiguous range with its own stream, no exchange on the hot path; getOutputs into device memory of one GPU is the optional single
    gather of the results.  Arrays are (K, N) over the WHOLE batch: numpy (host) or torch CUDA tensors of any device."""

    def __init__(self, n_robots, devices, params=None):
        lib = self._lib = load_library()
        self.N = int(n_robots); self.devices = [int(d) for d in devices]
        arr = (C.c_int * len(self.devices))(*self.devices)
        p = params if params is not None else default_params()
        self._m = C.c_void_p()
        rc = lib.wbc_multi_create(self.N, arr, len(self.devices), C.byref(p), C.byref(self._m))
        if rc != 0:
            self._m = C.c_void_p()
            raise RuntimeError(f"wbc_multi_create failed with {rc} (devices {self.devices}; there is no CPU fallback)")

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(f"wbc error {rc}: {self._lib.wbc_multi_last_error(self._m).decode()}")

    def close(self):
        if getattr(self, "_m", None) and self._m.value:
            self._lib.wbc_multi_destroy(self._m); self._m = C.c_void_p()

    __del__ = close

    def shards(self):
        out = []
        for k in range(self._lib.wbc_multi_num_devices(self._m)):
            d, f, c = C.c_int(), C.c_int(), C.c_int()
            self._ck(self._lib.wbc_multi_shard(self._m, k, C.byref(d), C.byref(f), C.byref(c))); out.append((d.value, f.value, c.value))
        return out

    def setInitialState(self):
        self._ck(self._lib.wbc_multi_reset(self._m))

    def setJointAndBaseState(self, q, qd, base_pos, base_quat, base_twist):
        self._ck(self._lib.wbc_multi_set_state(self._m, _ptr(q), _ptr(qd), _ptr(base_pos), _ptr(base_quat), _ptr(base_twist)))

    def setReference(self, pose, vel, acc, sw_pos, sw_vel, sw_acc, contacts, switching=None):
        self._ck(self._lib.wbc_multi_set_reference(self._m, _ptr(pose), _ptr(vel), _ptr(acc), _ptr(sw_pos), _ptr(sw_vel), _ptr(sw_acc), _ptr(contacts), _ptr(switching)))

    def step(self):
        self._ck(self._lib.wbc_multi_step(self._m))

    def synchronize(self):
        self._ck(self._lib.wbc_multi_synchronize(self._m))

    def getOutputs(self, tau=None, grf=None, x=None, rc=None, nwsr=None, ws=None, com=None):
        if all(v is None for v in (tau, grf, x, rc, nwsr, ws, com)):
            N = self.N
            o = dict(tau=np.zeros((12, N)), grf=np.zeros((12, N)), x=np.zeros((42, N)), rc=np.zeros(N, np.int32), nwsr=np.zeros(N, np.int32),
                     ws=np.zeros((112, N), np.int8), com=np.zeros((3, N)))
            self._ck(self._lib.wbc_multi_get_outputs(self._m, *[_ptr(o[k]) for k in ("tau", "grf", "x", "rc", "nwsr", "ws", "com")]))
            return o
        self._ck(self._lib.wbc_multi_get_outputs(self._m, _ptr(tau), _ptr(grf), _ptr(x), _ptr(rc), _ptr(nwsr), _ptr(ws), _ptr(com)))

    def lastGatherMs(self):
        t = C.c_float(); self._ck(self._lib.wbc_multi_last_gather_ms(self._m, C.byref(t))); return t.value
