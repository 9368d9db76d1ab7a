"""The log/dump format: the four topics the reference records (launch/whole_body_controller.launch:28-34) and plots
(scripts/plot_bag.py:34-132) as one table per (tick, robot).  CPU: layout and round trips on synthetic arrays; GPU: a recorded rollout."""
import csv
import os

import numpy as np
import pytest


def _fake(n, seed):
    rng = np.random.default_rng(seed)
    st = dict(q=rng.normal(size=(12, n)), qd=rng.normal(size=(12, n)), base_pos=rng.normal(size=(3, n)), base_quat=rng.normal(size=(4, n)), base_twist=rng.normal(size=(6, n)))
    out = dict(tau=rng.normal(size=(12, n)), grf=rng.normal(size=(12, n)), rc=rng.integers(0, 2, n).astype(np.int32) * 64, nwsr=rng.integers(0, 100, n).astype(np.int32))
    return st, out


def test_layout_and_round_trips(tmp_path):
    from quadrupedwholebodycontroller_b200 import recorder as rec
    assert rec.JOINT_NAMES[:3] == ["LH_HAA", "LH_HFE", "LH_KFE"] and rec.JOINT_NAMES[-1] == "RH_KFE" and len(rec.COLUMNS) == 3 + 12 + 7 + 12 + 12 + 2
    r = rec.RolloutRecorder([1, 5], loop_rate=400.0)
    frames = [_fake(8, s) for s in range(3)]
    for st, out in frames:
        r.record_arrays(st, out)
    t = r.table()
    assert t.shape == (6, len(rec.COLUMNS)) and list(t[:, 0]) == [1, 1, 2, 2, 3, 3] and list(t[:, 2]) == [1, 5, 1, 5, 1, 5]
    assert np.allclose(t[:, 1], t[:, 0] / 400.0)
    st, out = frames[1]
    row = t[3]                                              # tick 2, robot 5
    assert np.array_equal(row[3:15], st["q"][:, 5]) and np.array_equal(row[15:18], st["base_pos"][:, 5])
    assert np.array_equal(row[18:22], st["base_quat"][[1, 2, 3, 0], 5])          # ROS order x, y, z, w
    assert np.array_equal(row[22:34], out["tau"][:, 5]) and np.array_equal(row[34:46], out["grf"][:, 5]) and row[46] == out["rc"][5] and row[47] == out["nwsr"][5]
    p = os.path.join(tmp_path, "log.npz"); r.save_npz(p); z = rec.load_npz(p)
    assert z["columns"] == rec.COLUMNS and np.array_equal(z["data"], t) and z["loop_rate"] == 400.0
    c = os.path.join(tmp_path, "log.csv"); r.save_csv(c)
    rows = list(csv.reader(open(c)))
    assert rows[0] == rec.COLUMNS and len(rows) == 7
    assert np.array_equal(np.array([float(v) for v in rows[4]]), t[3])            # repr() round-trips doubles exactly
    tp = r.topics(5)
    assert np.array_equal(tp["/anymal/joint_states"]["RF_KFE"][1], t[1::2, 3 + 8])
    assert np.array_equal(tp["/anymal/desired_ground_reaction_forces"]["LF"][1], t[1::2, 34 + 5])
    assert tp["/gazebo/model_states"]["orientation_xyzw"].shape == (3, 4)


@pytest.mark.gpu
def test_recorded_rollout(wbc, wl, tmp_path):
    from quadrupedwholebodycontroller_b200 import recorder as rec
    n = 32
    st, ref, cmd = wl.rollout_start(n)
    ctl = wbc.WholeBodyController(n); ctl.plannerCreate(); ctl.plantEnable()
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
    r = rec.RolloutRecorder([0, 7, 31])
    for _ in range(40):
        ctl.rollout(1); r.record(ctl)
    ctl.close()
    t = r.table()
    assert t.shape == (120, len(rec.COLUMNS)) and (t[:, -2] == 0).all()
    tp = r.topics(7)
    fz = sum(tp["/anymal/desired_ground_reaction_forces"][leg][1] for leg in rec.LEG_ORDER)
    assert (fz[5:] > 300).all() and (fz[5:] < 900).all()                          # the feet carry the robot (51.6 kg)
    assert np.abs(np.stack([v[1] for v in tp["/anymal/joint_effort_group_controller/command"].values()])).max() <= 80.0 + 1e-9
