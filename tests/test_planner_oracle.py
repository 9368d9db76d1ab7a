"""CPU: the restated MotionPlanner (oracle/planner_oracle.cpp; reference src/motion_planner.cpp) against the committed golden vectors
(tests/golden/planner_golden.npz, made by tools/make_golden_planner.py) and against what the reference's code guarantees by construction.
The reference package ships no planner test; the properties below are read off its source (file:line in each test)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "tools"))
import make_golden_planner as mgp          # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "planner_golden.npz"))
KEYS = ("published", "pose", "vel", "acc", "sw_pos", "sw_vel", "sw_acc", "contacts")
SLOT = {0: 0, 1: 9, 2: 3, 3: 6}                    # step phase -> message offset: LH, RH, LF, RF (motion_planner.cpp:253-310)
CONTACTS = {0: (0, 1, 1, 1), 1: (1, 1, 1, 0), 2: (1, 0, 1, 1), 3: (1, 1, 0, 1)}


@pytest.mark.parametrize("name", ["straight", "turning", "startstop"])
def test_oracle_reproduces_golden(orc, name):
    r = orc.planner_run(GOLD[f"p_{name}_cmd"])
    for k in KEYS:
        assert np.array_equal(r[k], GOLD[f"p_{name}_{k}"]), (name, k)          # same machine arithmetic: bit-exact


def test_publish_pattern_and_contact_sequence(orc):
    """20 samples per 0.2 s phase at dt 0.01 (motion_planner.cpp:224-226), one silent period per phase change (:349-353), a second one
    at the end of a cycle (:357-376), phases LH, RH, LF, RF (:181)."""
    r = orc.planner_run(np.tile([0.3, 0.0, 0.0], (200, 1)))
    silent = np.nonzero(r["published"] == 0)[0]
    assert list(silent[:10]) == [20, 41, 62, 83, 84, 105, 126, 147, 168, 169]
    starts = [0, 21, 42, 63, 85, 106, 127, 148]
    for n, s in enumerate(starts):
        assert tuple(r["contacts"][s]) == CONTACTS[n % 4] and tuple(r["contacts"][s + 19]) == CONTACTS[n % 4]


def test_swing_curve_end_points_and_body_motion(orc):
    P = orc.DEFAULT_PLANNER_PARAMS
    r = orc.planner_run(np.tile([0.3, 0.0, 0.0], (200, 1)))
    lh0 = np.array([-P["x_offset"], P["y_offset"], 0.0])
    # first sample of phase 0: s = 0 -> Bezier starts at pi_foot_LH with zero velocity (quintic, vi = 0; :188)
    assert np.array_equal(r["sw_pos"][0, 0:3], lh0) and np.array_equal(r["sw_vel"][0, 0:3], np.zeros(3))
    # the foot target advances by step_length * cmd per cycle (:219-222); the last sample (t = 0.19) is close to it
    tgt = lh0 + P["step_length"] * np.array([0.3, 0.0, 0.0])
    assert np.linalg.norm(r["sw_pos"][19, 0:3] - tgt) < 2e-3
    assert r["sw_pos"][:20, 2].max() == pytest.approx(0.75 * P["height_control_point"], rel=0.05)      # cubic Bezier apex = 3/4 of the control height
    # cycle 0: pf_body_ was built from a zero command (:153) -> the body does not move; from cycle 1 on it advances by step_length * cmd per cycle
    assert np.array_equal(r["pose"][:84, 0], np.zeros(84))
    assert r["pose"][85:168, 0].max() < P["step_length"] * 0.3 + 1e-12 and r["pose"][167, 0] > 0.9 * P["step_length"] * 0.3
    assert np.all(r["pose"][:, 2] == P["body_height"])


def test_stand_still_publishes_every_period(orc):
    """zero /cmd_vel: `else` branch of plannerLoop (motion_planner.cpp:370-373) publishes the latched message with contacts 1111"""
    r = orc.planner_run(np.zeros((30, 3)))
    assert r["published"].all() and (r["contacts"] == 1).all()
    assert np.array_equal(r["pose"], np.tile([0, 0, 0.5, 0, 0, 0], (30, 1))) and not r["sw_pos"].any()


def test_command_change_takes_effect_at_cycle_boundaries(orc):
    """the rotated command is captured at the start of a cycle (:201-203); a stop request lets the running cycle finish (:224)"""
    cmd = GOLD["p_startstop_cmd"]; g = {k: GOLD[f"p_startstop_{k}"] for k in KEYS}
    assert (g["contacts"][:5] == 1).all()                       # standing
    assert tuple(g["contacts"][5]) == CONTACTS[0]                # walking starts in the period the command arrives
    last_walk = np.nonzero((g["contacts"] == 0).any(axis=1))[0].max()
    assert cmd[last_walk].any() is not None and last_walk >= 60    # command went to zero at period 60, the cycle still completes
    assert (g["contacts"][last_walk + 3:] == 1).all()


def test_rollout_golden_is_sane():
    assert (GOLD["r_rc"] == 0).all()
    tau = GOLD["r_tau"]
    assert np.abs(tau).max() <= 80.0 + 1e-9                      # torque limit (params_controller.yaml maxTorque)
    assert GOLD["r_nwsr"][0].min() >= 40                         # tick 0 is a cold start (SQProblem::init)
    assert np.median(GOLD["r_nwsr"][1:]) <= 2                    # later ticks are hot starts
    z = GOLD["r_pos"][:, 2, :]
    assert (z > 0.4).all() and (z < 0.7).all()                   # the base stays up
