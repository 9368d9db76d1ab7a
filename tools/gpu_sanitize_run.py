"""DEV: small hot path for compute-sanitizer (cold tick + two hot-started ticks incl. a contact change, planner + plant rollout), fp64 or fp32 solver."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl
fp32 = len(sys.argv) > 1 and sys.argv[1] == "fp32"
N = 200
st, ref = wl.randomised(N, seed=5, contacts="mixed", switching_frac=0.1)
ctl = wbc.WholeBodyController(N, fp32=fp32)
for t in range(3):
    if t: st = wl.advance_state(st)
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); ctl.synchronize()
o = ctl.getOutputs(); print("rc0", float((o["rc"] == 0).mean()), "nwsr", float(o["nwsr"].mean())); ctl.close()
st, ref, cmd = wl.rollout_start(N)
ctl = wbc.WholeBodyController(N, fp32=fp32); ctl.plannerCreate(); ctl.plantEnable()
ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
ctl.rollout(30); ctl.synchronize(); print("rollout ok", ctl.rolloutStats()["failed_ticks"].sum()); ctl.close()
