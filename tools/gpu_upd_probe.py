"""DEV: update_state kernel time for the library selected by WBC_B200_LIB."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl
N = 16384
st, ref = wl.randomised(N, seed=1234, contacts="mixed")
ctl = wbc.WholeBodyController(N); ctl.plantEnable()
best = 1e9
for rep in range(6):
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.updateState(); ctl.synchronize()
    best = min(best, ctl.lastKernelMs()["update_state"])
print(os.environ.get("WBC_B200_LIB", "default"), "update_state", best, "ms for", N)
