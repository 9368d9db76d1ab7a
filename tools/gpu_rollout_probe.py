"""DEV: closed-loop rollout timing on the GPU: per-tick time over the gait, nWSR statistics, kernel split."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
T = int(sys.argv[2]) if len(sys.argv) > 2 else 400
st, ref, cmd = wl.rollout_start(N)
ctl = wbc.WholeBodyController(N, profiling=True); ctl.plannerCreate(); ctl.plantEnable(); ctl.profile(True)
ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
ctl.rollout(1); ctl.synchronize(); print("cold tick", ctl.lastKernelMs())
rows = []
for t in range(1, T):
    ctl.rollout(1); ms = ctl.lastKernelMs()
    if t % 4 == 0 or ms["solve_qp"] > 20:
        o = ctl.getOutputs(); rows.append((t, ms["solve_qp"], ms["update_state"], o["nwsr"].mean(), o["nwsr"].max(), int((o["rc"] != 0).sum())))
for r in rows[:: max(1, len(rows) // 60)]:
    print("tick %4d qp %7.3f ms rbd %6.3f ms nwsr mean %6.2f max %3d failed %d" % r)
ctl.rolloutStats(reset=True)
ctl.synchronize(); t0 = time.time(); ctl.rollout(200); ctl.synchronize(); dt = time.time() - t0
s = ctl.rolloutStats()
print(f"200 ticks: {dt * 1e3:.1f} ms wall -> {N * 200 / dt:.0f} solves/s ; failed ticks {int(s['failed_ticks'].sum())} ; nwsr mean {s['nwsr_sum'].sum() / s['ticks'].sum():.2f} max {s['nwsr_max'].max()}")
p = ctl.profile(True, fetch=True); tot = p[:, 6].sum()
names = ["shift", "stepdir", "step", "change", "chol", "rampdrift"]
print("last tick solver phases: " + " ".join(f"{n}={p[:, k].sum() / tot * 100:.0f}%" for k, n in enumerate(names)), "; setup+other =", f"{(1 - p[:, :6].sum() / tot) * 100:.0f}%")
ctl.close()
