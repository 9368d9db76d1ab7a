"""DEV: is the solver instruction-cache bound?  Identical robots (same control flow in every warp -> natural lockstep, shared I-cache lines)
vs randomised robots, at 1 warp/SM and at full occupancy."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

def run(N, ident):
    st, ref = wl.randomised(N, seed=1234, contacts="stance")
    if ident:
        st = {k: np.ascontiguousarray(np.repeat(v[:, :1], N, 1)) for k, v in st.items()}
    ctl = wbc.WholeBodyController(N); ctl.profile(True)
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); ctl.synchronize()
    ms = ctl.lastKernelMs(); p = ctl.profile(True, fetch=True); o = ctl.getOutputs()
    print(f"N={N} identical={ident}: qp {ms['solve_qp']:.2f} ms -> {N / ms['solve_qp'] * 1e3:.0f} solves/s; nwsr mean {o['nwsr'].mean():.1f}; cycles/iter {p[:, 6].sum() / max(1, p[:, 7].sum()):.0f}")
    ctl.close()

for N in (148, 148 * 5, 4096):
    for ident in (False, True):
        run(N, ident)
