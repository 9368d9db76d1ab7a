"""DEV / measurement: the fp32 build of the solver (libwbc_b200_f32.so) against the fp64 product on the same batch: return codes, working sets,
tau / grf / x error relative to the fp64 result (north star: 1e-4), kernel times.  usage: gpu_fp32_probe.py [N] [contacts]"""
import sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
contacts = sys.argv[2] if len(sys.argv) > 2 else "stance"
st, ref = wl.randomised(N, seed=1234, contacts=contacts, switching_frac=0.1 if contacts == "mixed" else 0.0)
a = wbc.WholeBodyController(N); b = wbc.WholeBodyController(N, fp32=True)
rel = lambda x, y: np.abs(x - y).max(axis=0) / np.maximum(1.0, np.abs(y).max(axis=0))
out = []
s = st
for t in range(3):
    if t: s = wl.advance_state(s)
    ta = tb = 1e9
    for rep in range(2 if t == 0 else 1):
        for c in (a, b):
            if t == 0: c.setInitialState()
            c.setJointAndBaseState(**s); c.setReference(**ref); c.step(); c.synchronize()
        ta = min(ta, a.lastKernelMs()["solve_qp"]); tb = min(tb, b.lastKernelMs()["solve_qp"])
    oa, ob = a.getOutputs(), b.getOutputs()
    ok = (oa["rc"] == 0) & (ob["rc"] == 0)
    print('rc fp32', dict(zip(*[v.tolist() for v in np.unique(ob['rc'], return_counts=True)])), 'nwsr', ob['nwsr'][:8])
    if not ok.any(): continue
    r = dict(tick=t, n=N, contacts=contacts, fp64_ms=ta, fp32_ms=tb, speedup=ta / tb, rc_equal=float((oa["rc"] == ob["rc"]).mean()), fp32_rc0=float((ob["rc"] == 0).mean()), fp64_rc0=float((oa["rc"] == 0).mean()),
             ws_equal=float((oa["ws"] == ob["ws"]).all(axis=0)[ok].mean()), nwsr_mean_fp64=float(oa["nwsr"].mean()), nwsr_mean_fp32=float(ob["nwsr"].mean()),
             max_rel_tau=float(rel(ob["tau"], oa["tau"])[ok].max()), max_rel_grf=float(rel(ob["grf"], oa["grf"])[ok].max()), max_rel_x=float(rel(ob["x"], oa["x"])[ok].max()),
             frac_tau_within_1e4=float((rel(ob["tau"], oa["tau"])[ok] < 1e-4).mean()))
    print(json.dumps(r)); out.append(r)
a.close(); b.close()
