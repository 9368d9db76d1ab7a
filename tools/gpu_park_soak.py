"""DEV: soak of the parked cold path against the single launch of the general kernel: bigger batches, several chunks, every contact mix; all outputs incl. duals bit for bit."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

for contacts, n, chunk, seed in (("mixed", 20000, 4096, 5), ("diagonal", 12000, 5000, 6), ("stance", 9000, 0, 7)):
    st, ref = wl.randomised(n, seed=seed, contacts=contacts, switching_frac=0.1)
    outs = []
    for park in (1, 0):
        ctl = wbc.WholeBodyController(n); ctl.setOption("park", park)
        if park and chunk:
            ctl.setOption("park_chunk", chunk)
        rec = []; s = st
        for t in range(2):
            if t:
                s = wl.advance_state(s)
            ctl.setJointAndBaseState(**s); ctl.setReference(**ref); ctl.step(); o = ctl.getOutputs(); o["y"] = ctl.getDuals(); rec.append(o)
        ctl.close(); outs.append(rec)
    same = all(np.array_equal(outs[0][t][k], outs[1][t][k]) for t in range(2) for k in outs[0][t])
    o = outs[0][0]
    print(f"{contacts:9s} n={n} chunk={chunk or 'default'}: identical={same}  rc0 {np.mean(o['rc'] == 0):.4f} rc64 {np.mean(o['rc'] == 64):.4f} other {np.mean((o['rc'] != 0) & (o['rc'] != 64)):.4f}  nwsr mean {o['nwsr'].mean():.1f} max {o['nwsr'].max()}")
    assert same
print("OK")
