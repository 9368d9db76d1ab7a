"""The C++ host mirror of the reference class (include/whole_body_controller_b200.hpp) on top of the C ABI:
CPU: it compiles with g++ against the header and links to the CUDA library; without a GPU the node fails loudly (no fallback).
GPU: the ROS-free node (examples/whole_body_controller_node.cpp = reference node main + controlLoop body) runs config 1
(single ANYmal, nominal stance): tick 0 cold init, then hot starts, every QP solved, sum f_z = m g."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "quadrupedwholebodycontroller_b200")
EXE = os.path.join(ROOT, "examples", "wbc_node")


def _build(wbc, src="whole_body_controller_node.cpp", exe=EXE):
    cmd = ["g++", "-std=c++17", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", src),
           "-L" + PKG, "-lwbc_b200", "-Wl,-rpath," + PKG, "-o", exe]
    subprocess.check_call(cmd)
    return exe


def test_cpp_class_compiles_and_refuses_to_run_without_gpu(wbc):
    exe = _build(wbc)
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([exe, "1", "1"], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_cpp_node_runs_config1(wbc):
    exe = _build(wbc)
    r = subprocess.run([exe, "1", "20"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("tick")]
    assert len(lines) == 20
    fz = [float(re.search(r"sum f_z ([-0-9.e+]+)", l).group(1)) for l in lines]
    nwsr = [float(re.search(r"mean nWSR ([0-9.]+)", l).group(1)) for l in lines]
    # static equilibrium in the nominal stance: the pyramid forces carry the weight (51.63574 kg * 9.81) up to the PD pull towards the 0.5 m reference height
    assert all(abs(f - fz[-1]) < 1e-6 * abs(fz[-1]) for f in fz[3:]) and 300.0 < fz[-1] < 900.0, fz
    assert nwsr[0] > 40 and max(nwsr[3:]) <= 2, nwsr      # cold init, then hot starts on the same state


PLANNER_EXE = os.path.join(ROOT, "examples", "planner_node")


def test_cpp_planner_node_compiles_and_refuses_to_run_without_gpu(wbc):
    exe = _build(wbc, "motion_planner_node.cpp", PLANNER_EXE)
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([exe, "2", "8"], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr, (r.returncode, r.stderr)


@pytest.mark.gpu
def test_cpp_closed_loop_walks(wbc):
    """planner + controller + plant through the C++ mirrors: 8 robots walk for 2 s (two gait cycles), every QP solves, the base moves
    forward by about the commanded step_length * v per cycle and stays up"""
    exe = _build(wbc, "motion_planner_node.cpp", PLANNER_EXE)
    r = subprocess.run([exe, "8", "800"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    last = [l for l in r.stdout.splitlines() if l.startswith("tick 800")][0]
    x, y, z = [float(v) for v in re.search(r"base \(([-0-9.e+]+) ([-0-9.e+]+) ([-0-9.e+]+)\)", last).groups()]
    assert 0.4 < z < 0.7 and -0.05 < x < 0.2, last
    assert "failed solves 0" in r.stdout.splitlines()[-1]
