"""DEV: solver kernel time vs resident solver warps per SM (WBC_QP_CTAS_PER_SM knob), cold 4096 stance solves + one hot-started tick."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
st, ref = wl.randomised(N, seed=1234, contacts="stance")
st1 = wl.advance_state(st)
KMAX = int(sys.argv[2]) if len(sys.argv) > 2 else 5
print('library', os.environ.get('WBC_B200_LIB', 'default'), 'sm stride', os.environ.get('WBC_QP_SM_STRIDE', '1'), flush=True)
for k in range(1, KMAX + 1):
    os.environ["WBC_QP_CTAS_PER_SM"] = str(k)
    ctl = wbc.WholeBodyController(N)
    best = [1e9, 1e9]
    for rep in range(2):
        ctl.setInitialState()
        for t, s in enumerate((st, st1)):
            ctl.setJointAndBaseState(**s); ctl.setReference(**ref); ctl.step(); ctl.synchronize()
            best[t] = min(best[t], ctl.lastKernelMs()["solve_qp"])
    print(f"ctas/SM {k:2d} (granted {ctl.qpCtasPerSM() if hasattr(ctl, 'qpCtasPerSM') else '?'}): cold {best[0]:8.2f} ms  warm {best[1]:8.2f} ms  -> cold {N / best[0] * 1e3:9.0f} solves/s", flush=True)
    ctl.close()
