"""DEV: target of ncu captures of WARM closed-loop ticks: N robots, T ticks of the rollout (tick 0 cold, then hot starts)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
T = int(sys.argv[2]) if len(sys.argv) > 2 else 40
st, ref, cmd = wl.rollout_start(N)
ctl = wbc.WholeBodyController(N); ctl.plannerCreate(); ctl.plantEnable()
ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
for t in range(T):
    ctl.rollout(1)
    if t >= T - 3:
        ctl.synchronize(); o = ctl.getOutputs(); print(t, ctl.lastKernelMs(), "nwsr mean", o["nwsr"].mean())
ctl.close()
