"""DEV: quick GPU probe -- parity on a small batch, kernel times and clock64 breakdown on larger ones."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

def run(N, contacts, ticks=2, check=False):
    st, ref = wl.randomised(N, seed=1234, contacts=contacts)
    ctl = wbc.WholeBodyController(N, profiling=True); ctl.profile(True)
    if check:
        from oracle import oracle as orc
        ob = orc.OracleBatch(N)
    for t in range(ticks):
        if t: st = wl.advance_state(st)
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); ctl.synchronize()
        ms = ctl.lastKernelMs(); o = ctl.getOutputs(); p = ctl.profile(True, fetch=True)
        tot = p[:, 6].sum()
        names = ["shift", "stepdir", "step", "change", "chol", "rampdrift"]
        br = " ".join(f"{n}={p[:, k].sum() / tot * 100:.0f}%" for k, n in enumerate(names))
        print(f"N={N} {contacts} tick {t}: qp {ms['solve_qp']:.2f} ms  rbd {ms['update_state']:.3f} ms  -> {N / ms['solve_qp'] * 1e3:.0f} solves/s; "
              f"nwsr mean {o['nwsr'].mean():.1f} rc!=0 {np.count_nonzero(o['rc'])}; cycles/iter {p[:, 6].sum() / max(1, p[:, 7].sum()):.0f}; {br}")
        if check:
            ob.set_state(**st); ob.set_reference(**ref); ob.step(); w = ob.outputs()
            ok = (w["rc"] == 0) & (o["rc"] == 0)
            sc = np.maximum(1.0, np.abs(w["tau"]).max(axis=0))
            print(f"   parity: rc equal {(o['rc'] == w['rc']).mean() * 100:.1f}%  nwsr equal {(o['nwsr'] == w['nwsr']).mean() * 100:.1f}%  ws equal "
                  f"{(o['ws'] == w['ws']).all(axis=0).mean() * 100:.1f}%  tau relerr {(np.abs(o['tau'] - w['tau']) / sc)[:, ok].max():.2e}")
    ctl.close()

if __name__ == "__main__":
    run(64, "mixed", check=True)
    run(4096, "stance")
