"""DEV: one short pass of the hot path for ncu (cold solve of `N` stance robots, then one hot-started tick)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
contacts = sys.argv[2] if len(sys.argv) > 2 else "stance"
st, ref = wl.randomised(N, seed=1234, contacts=contacts)
ctl = wbc.WholeBodyController(N)
for kv in os.environ.get("VAR_OPTIONS", "").split(","):      # e.g. VAR_OPTIONS=park=0
    if kv:
        ctl.setOption(kv.split("=")[0], int(kv.split("=")[1]))
for t in range(2):
    if t:
        st = wl.advance_state(st)
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); ctl.synchronize()
    print(t, ctl.lastKernelMs())
ctl.close()
