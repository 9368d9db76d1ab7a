#!/usr/bin/env python3
"""Generates tests/golden/planner_golden.npz with the oracle: the restated MotionPlanner (oracle/planner_oracle.cpp) and the closed
loop planner -> controller (+ real qpOASES) -> plant (oracle/oracle_capi.cpp).  Run in the build container; the vectors travel.

    python tools/make_golden_planner.py

Cases:
  p_*   planner alone, 140 periods each: straight walk, walk with yaw rate, stand -> walk -> stop (the /cmd_vel profile is stored)
  r_*   closed loop, 6 robots x 200 control ticks (50 planner periods: stand-up transient, first swing phase, first contact switch)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc                                   # noqa: E402
from quadrupedwholebodycontroller_b200 import workloads as wl      # noqa: E402

PLANNER_PERIODS = 140
ROLLOUT_N, ROLLOUT_TICKS = 6, 200


def planner_profiles():
    K = PLANNER_PERIODS
    a = np.tile([0.3, 0.0, 0.0], (K, 1))
    b = np.tile([0.2, -0.1, 0.3], (K, 1))
    c = np.zeros((K, 3)); c[5:60] = [0.25, 0.05, -0.2]
    return dict(straight=a, turning=b, startstop=c)


def rollout_inputs():
    """nominal stance with a small deterministic perturbation per robot, and one constant /cmd_vel per robot"""
    st, ref = wl.nominal(ROLLOUT_N)
    rs, _ = wl.randomised(ROLLOUT_N, seed=77, contacts="stance")
    st = {k: st[k] + 0.2 * (rs[k] - st[k]) for k in st}          # 20 % of the randomised offsets
    st["base_quat"] /= np.linalg.norm(st["base_quat"], axis=0, keepdims=True)
    cmd = np.zeros((3, ROLLOUT_N)); cmd[0] = [0.3, 0.3, 0.0, 0.2, -0.2, 0.1]; cmd[1] = [0.0, 0.1, 0.0, -0.1, 0.0, 0.2]; cmd[2] = [0.0, 0.0, 0.0, 0.2, 0.0, -0.3]
    return st, ref, cmd


def main():
    out = {}
    for name, cmd in planner_profiles().items():
        r = orc.planner_run(cmd)
        out[f"p_{name}_cmd"] = cmd
        for k, v in r.items():
            out[f"p_{name}_{k}"] = v
        print("planner", name, "published", int(r["published"].sum()), "of", len(cmd))
    st, ref, cmd = rollout_inputs()
    ob = orc.OracleBatch(ROLLOUT_N); ob.set_state(**st); ob.set_reference(**ref)
    devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(2); os.dup2(devnull, 2)
    try:
        r = ob.rollout(ROLLOUT_TICKS, cmd)
    finally:
        os.dup2(saved, 2); os.close(devnull); os.close(saved)
    for k in ("tau", "rc", "nwsr", "q", "pos"):
        out[f"r_{k}"] = r[k]
    for k, v in ob.state().items():
        out[f"r_final_{k}"] = v
    print("rollout: rc != 0 on", int((r["rc"] != 0).sum()), "ticks; nwsr mean", r["nwsr"].mean(), "max", r["nwsr"].max())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "planner_golden.npz"), **out)


if __name__ == "__main__":
    main()
