"""DEV: closed-loop rollout of small batches (the regime where a control tick is shorter than the host's launch work): wall time per control tick of
wbc_rollout with the steady-state planner period replayed as a CUDA graph vs plain per-tick launches (wbc_set_option "rollout_graph")."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

T = int(sys.argv[1]) if len(sys.argv) > 1 else 400
for N in (1, 32, 148, 888, 4096):
    st, ref, cmd = wl.rollout_start(N)
    res = {}
    for graph in (0, 1):
        ctl = wbc.WholeBodyController(N); ctl.plannerCreate(); ctl.plantEnable(); ctl.setOption("rollout_graph", graph)
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
        ctl.rollout(120); ctl.synchronize()                      # cold tick, first contact switches, graph capture
        best = 1e30
        for rep in range(3):
            t0 = time.perf_counter(); ctl.rollout(T); t1 = time.perf_counter(); ctl.synchronize(); t2 = time.perf_counter()
            best = min(best, (t2 - t0) / T * 1e3); issue = (t1 - t0) / T * 1e3
        res[graph] = (best, issue, ctl.getOutputs()["tau"].copy())
        ctl.close()
    same = np.array_equal(res[0][2], res[1][2])
    print(f"N={N:5d}: plain launches {res[0][0]:.4f} ms/tick (host issue {res[0][1]:.4f}); graph replay {res[1][0]:.4f} ms/tick (host issue {res[1][1]:.4f}); "
          f"speed-up {res[0][0] / res[1][0]:.3f}x; same torques {same}", flush=True)
