"""DEV: gang scheduling of kernel 2 on the closed-loop rollout: per-warp fetch (cohort 0, what hot-started launches use) against
gang fetch (cohort 1) tick by tick -- do the heavy ticks (contact switches, every robot takes 20-90 working-set changes) prefer the cohort?
usage: gpu_cohort_rollout_probe.py [robots] [ticks]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
T = int(sys.argv[2]) if len(sys.argv) > 2 else 240
res = {}
for mode in (0, 1):
    st, ref, cmd = wl.rollout_start(N)
    ctl = wbc.WholeBodyController(N); ctl.plannerCreate(); ctl.plantEnable()
    ctl.setOption("rollout_graph", 0); ctl.setOption("cohort", mode)
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
    ctl.rollout(8); ctl.synchronize()
    ms, nw = [], []
    for t in range(T):
        ctl.rollout(1); ctl.synchronize(); ms.append(ctl.lastKernelMs()["solve_qp"]); nw.append(float(ctl.getOutputs()["nwsr"].mean()))
    res[mode] = (np.array(ms), np.array(nw)); ctl.close()
a, b = res[0][0], res[1][0]; nw = res[0][1]
assert np.allclose(res[0][1], res[1][1]), "the two modes must take the same iterations"
heavy = nw > 5.0
print(f"{N} robots x {T} ticks; heavy ticks (mean nWSR > 5): {int(heavy.sum())}")
print(f"per-warp fetch: total {a.sum():8.1f} ms   quiet {a[~heavy].sum():8.1f}   heavy {a[heavy].sum():8.1f}")
print(f"gang fetch    : total {b.sum():8.1f} ms   quiet {b[~heavy].sum():8.1f}   heavy {b[heavy].sum():8.1f}")
print(f"best of both per tick class: {min(a[~heavy].sum(), b[~heavy].sum()) + min(a[heavy].sum(), b[heavy].sum()):8.1f} ms")
for t in np.nonzero(heavy)[0][:24]:
    print(f"  tick {t:4d} nwsr mean {nw[t]:6.2f}  per-warp {a[t]:7.2f} ms  gang {b[t]:7.2f} ms")
# threshold scan: at which mean nWSR does the gang fetch start to win?
for lo, hi in ((0, 0.5), (0.5, 2), (2, 5), (5, 15), (15, 40), (40, 200)):
    m = (nw >= lo) & (nw < hi)
    if m.any():
        print(f"  nwsr in [{lo},{hi}): {int(m.sum()):4d} ticks  per-warp {a[m].sum():8.1f}  gang {b[m].sum():8.1f}  ratio {a[m].sum() / b[m].sum():.3f}")
