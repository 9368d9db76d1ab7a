"""the ROS-facing drop-in of the reference's node class (include/anymal_wbc_b200/whole_body_controller_ros.hpp) is compiled against the
minimal roscpp / message stand-ins of tests/mock_ros (no ROS in the image) and linked with the CUDA library.
CPU: it compiles and links; without a GPU the constructor reports the failure the way the reference reports a failed model load (initStatus_ false).
GPU: the three callbacks are fed messages (joint names in Gazebo's alphabetical order, model index lookup on the first /gazebo/model_states), the body of
controlLoop() runs, and the torques / forces it PUBLISHES equal what the C ABI returns for the same robot state."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "quadrupedwholebodycontroller_b200")
EXE = os.path.join(ROOT, "examples", "ros_shim_test")


def _build(wbc):
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-DWBC_ROS_MOCK", "-I" + os.path.join(ROOT, "tests", "mock_ros"), "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "ros_shim_main.cpp"), "-L" + PKG, "-lwbc_b200", "-Wl,-rpath," + PKG, "-pthread", "-o", EXE])
    return EXE


def _state_file(tmp_path, st):
    p = os.path.join(tmp_path, "state.txt")
    with open(p, "w") as f:
        for k in ("q", "qd", "base_pos", "base_quat", "base_twist"):
            f.write(" ".join(repr(float(v)) for v in st[k][:, 0]) + "\n")
    return p


def test_ros_shim_compiles_and_reports_a_missing_gpu(wbc, wl, tmp_path):
    exe = _build(wbc)
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    st, _ = wl.nominal(1)
    r = subprocess.run([exe, _state_file(tmp_path, st), "1"], capture_output=True, text=True)
    assert r.returncode == 3 and "INIT_FAILED" in r.stdout and "no CPU fallback" in r.stderr, (r.returncode, r.stdout, r.stderr)


@pytest.mark.gpu
def test_ros_shim_publishes_what_the_c_abi_computes(wbc, wl, tmp_path):
    exe = _build(wbc)
    st, ref = wl.randomised(1, seed=77, contacts="stance")
    K = 3
    r = subprocess.run([exe, _state_file(tmp_path, st), str(K)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    out = {l.split()[0]: l.split()[1:] for l in r.stdout.splitlines() if l}
    assert out["rc"][0] == "0" and int(out["rc"][2]) == K and int(out["reset_msgs"][0]) in (200, 201)           # resetRobotSimState: 0.5 s at 400 Hz (wbc.cpp:589-606; 200 x 0.0025 accumulates to just above 0.5)
    tau = np.array([float(v) for v in out["tau"]]); grf = np.array([float(v) for v in out["grf"]])
    ctl = wbc.WholeBodyController(1); ctl.setJointAndBaseState(**st); ctl.setReference(**ref)
    for _ in range(K):
        ctl.step()
    o = ctl.getOutputs(); ctl.close()
    assert np.array_equal(tau, o["tau"][:, 0]) and np.array_equal(grf, o["grf"][:, 0])
