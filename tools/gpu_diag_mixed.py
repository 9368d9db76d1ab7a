"""DEV: per-tick parity details of the seeded 'mixed' sequence of tests/test_gpu_parity.py."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl
from oracle import oracle as orc
contacts, n, ticks = "mixed", 128, 4
st, ref = wl.randomised(n, seed=99, contacts=contacts, switching_frac=0.1)
ctl = wbc.WholeBodyController(n); ob = orc.OracleBatch(n)
fd = os.open(os.devnull, os.O_WRONLY); sv = os.dup(2); os.dup2(fd, 2)
for t in range(ticks):
    if t: st = wl.advance_state(st)
    if t == 2:
        ref = dict(ref); ref["contacts"] = np.ascontiguousarray(np.roll(ref["contacts"], 1, axis=0)); ref["switching"] = None
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
    ob.set_state(**st); ob.set_reference(**ref); ob.step()
    g, w = ctl.getOutputs(), ob.outputs()
    ok = w["rc"] == 0
    rel = lambda a, b: np.abs(a - b).max(axis=0) / np.maximum(1.0, np.abs(b).max(axis=0))
    same = (g["ws"] == w["ws"]).all(axis=0)
    print(f"tick {t}: rc diff at {np.nonzero(g['rc'] != w['rc'])[0]} (mine {g['rc'][g['rc'] != w['rc']]}, oracle {w['rc'][g['rc'] != w['rc']]}); ws equal {same[ok].mean():.3f}; "
          f"nwsr equal {(g['nwsr'] == w['nwsr'])[ok].mean():.3f}; tau rel {rel(g['tau'], w['tau'])[ok].max():.2e} x rel {rel(g['x'], w['x'])[ok].max():.2e}", file=sys.stdout)
    bad = np.nonzero(ok & (rel(g["x"], w["x"]) > 1e-9))[0]
    for i in bad[:5]:
        print("   robot", i, "nwsr", g["nwsr"][i], w["nwsr"][i], "ws same", same[i], "x rel", rel(g["x"], w["x"])[i])
os.dup2(sv, 2)
