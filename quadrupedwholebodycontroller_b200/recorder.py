"""Per-tick log of selected robots of a batched run, in the layout of the four topics the reference records and plots
(launch/whole_body_controller.launch:28-34 `rosbag record`; scripts/plot_bag.py:34-132 reads exactly these):

    /anymal/joint_states                                -> joint positions[12] by joint name
    /gazebo/model_states (model "anymalModel")          -> base position xyz + orientation quaternion (x, y, z, w)
    /anymal/joint_effort_group_controller/command       -> commanded torques[12] in the order of config/controllers.yaml:8-19
    /anymal/desired_ground_reaction_forces              -> 12 values, 4 legs x (fx, fy, fz), leg order LH, LF, RF, RH (wbc.cpp:558-563)

There is no ROS here, so the bag becomes a compressed .npz (binary, also the regression/golden-vector container used under
tests/golden/) or a CSV with one row per (tick, robot).  `topics()` returns the same per-name time series plot_bag.py builds.
"""
import csv

import numpy as np

# model DOF order == controller command order (whole_body_controller.cpp:81,234; config/controllers.yaml:8-19)
JOINT_NAMES = [f"{leg}_{j}" for leg in ("LH", "LF", "RF", "RH") for j in ("HAA", "HFE", "KFE")]
LEG_ORDER = ["LH", "LF", "RF", "RH"]
COLUMNS = (["tick", "time", "robot"] + [f"q_{n}" for n in JOINT_NAMES] + ["base_x", "base_y", "base_z", "base_qx", "base_qy", "base_qz", "base_qw"]
           + [f"tau_{n}" for n in JOINT_NAMES] + [f"grf_{leg}_{a}" for leg in LEG_ORDER for a in "xyz"] + ["qp_return_value", "nwsr"])


class RolloutRecorder:
    """collects (q, base pose, tau, grf, rc, nWSR) of `robots` after every call of record()"""

    def __init__(self, robots, loop_rate=400.0):
        self.robots = np.asarray(robots, dtype=np.int64)
        self.loop_rate = float(loop_rate)
        self.rows = []
        self.tick = 0

    def record_arrays(self, state, outputs, ticks_elapsed=1):
        """state / outputs: the dicts of WholeBodyController.getState() / getOutputs() (arrays (K, N))"""
        self.tick += int(ticks_elapsed)
        r = self.robots
        quat = state["base_quat"][:, r]                                  # (w, x, y, z) on the controller side
        row = np.concatenate([np.full((1, len(r)), self.tick), np.full((1, len(r)), self.tick / self.loop_rate), r[None, :].astype(np.float64),
                              state["q"][:, r], state["base_pos"][:, r], quat[[1, 2, 3, 0]], outputs["tau"][:, r], outputs["grf"][:, r],
                              outputs["rc"][None, r].astype(np.float64), outputs["nwsr"][None, r].astype(np.float64)], axis=0)
        self.rows.append(row.T.copy())

    def record(self, controller, ticks_elapsed=1):
        self.record_arrays(controller.getState(), controller.getOutputs(), ticks_elapsed)

    def table(self):
        """(ticks * robots, len(COLUMNS)) float64"""
        return np.concatenate(self.rows, axis=0) if self.rows else np.zeros((0, len(COLUMNS)))

    def save_npz(self, path):
        np.savez_compressed(path, columns=np.array(COLUMNS), data=self.table(), joint_names=np.array(JOINT_NAMES), leg_order=np.array(LEG_ORDER),
                            loop_rate=self.loop_rate)

    def save_csv(self, path):
        with open(path, "w", newline="") as f:
            w = csv.writer(f); w.writerow(COLUMNS)
            for row in self.table():
                w.writerow([int(row[0]), repr(float(row[1])), int(row[2])] + [repr(float(v)) for v in row[3:-2]] + [int(row[-2]), int(row[-1])])

    def topics(self, robot):
        """the series scripts/plot_bag.py extracts from the bag, for one recorded robot"""
        t = self.table(); t = t[t[:, 2] == robot]
        col = {c: k for k, c in enumerate(COLUMNS)}
        time = t[:, col["time"]]
        return {
            "/anymal/joint_states": {n: (time, t[:, col[f"q_{n}"]]) for n in JOINT_NAMES},
            "/gazebo/model_states": dict(time=time, position=t[:, col["base_x"]:col["base_z"] + 1], orientation_xyzw=t[:, col["base_qx"]:col["base_qw"] + 1]),
            "/anymal/joint_effort_group_controller/command": {n: (time, t[:, col[f"tau_{n}"]]) for n in JOINT_NAMES},
            "/anymal/desired_ground_reaction_forces": {leg: (time, t[:, col[f"grf_{leg}_z"]]) for leg in LEG_ORDER},
        }


def load_npz(path):
    z = np.load(path)
    return dict(columns=[str(c) for c in z["columns"]], data=z["data"], joint_names=[str(c) for c in z["joint_names"]], leg_order=[str(c) for c in z["leg_order"]],
                loop_rate=float(z["loop_rate"]))
