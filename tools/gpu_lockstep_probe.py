"""DEV: would lock-step execution of the solver warps relieve the instruction supply?  Upper bound: every robot of the batch is a copy of
ONE robot, so all warps execute the very same instruction stream and stay roughly in phase on their own.  Reports working-set iterations/s of
the cold solve against resident warps per SM, for the normal (all different) and the identical batch."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
KMAX = int(sys.argv[2]) if len(sys.argv) > 2 else 5
st, ref = wl.randomised(N, seed=1234, contacts="stance")
for label, src in (("different", None), ("identical(3)", 3), ("identical(17)", 17)):
    s = st if src is None else {k: np.ascontiguousarray(np.repeat(v[:, [src]], N, axis=1)) for k, v in st.items()}
    for k in range(1, KMAX + 1):
        os.environ["WBC_QP_CTAS_PER_SM"] = str(k)
        ctl = wbc.WholeBodyController(N)
        best = 1e9
        for rep in range(2):
            ctl.setInitialState(); ctl.setJointAndBaseState(**s); ctl.setReference(**ref); ctl.step(); ctl.synchronize()
            best = min(best, ctl.lastKernelMs()["solve_qp"])
        it = ctl.getOutputs()["nwsr"].astype(np.float64).sum()
        print(f"{label:14s} warps/SM {k}: cold {best:8.2f} ms, {it / N:6.1f} iterations/robot -> {it / best / 1e3:8.1f} M iterations/s", flush=True)
        ctl.close()
