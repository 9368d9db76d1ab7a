"""DEV AID: host-emulated solver (one lane) vs the real qpOASES inside the oracle, tick by tick."""
import ctypes as C, sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from oracle import oracle as orc
from quadrupedwholebodycontroller_b200 import workloads as wl

rbd = C.CDLL("/tmp/libemul_rbd.so"); qp = C.CDLL("/tmp/libemul_qp.so")
qp.emulqp_create.restype = C.c_void_p
NS, NQ = rbd.emul_sizes(0), rbd.emul_sizes(1)
P = lambda a: a.ctypes.data_as(C.c_void_p)


def run(N, contacts, ticks=2, seed=11, verbose=True, nominal=False):
    st, ref = (wl.nominal(N) if nominal else wl.randomised(N, seed=seed, contacts=contacts))
    ob = orc.OracleBatch(N); pp = orc.pack_params()
    states = np.zeros((N, NS)); hs = [C.c_void_p(qp.emulqp_create()) for _ in range(N)]
    stats = dict(n=0, rc=0, nwsr=0, ws=0, xerr=0.0)
    for t in range(ticks):
        if t > 0: st = wl.advance_state(st)
        ob.set_state(**st); ob.set_reference(**ref); ob.step()
        o = ob.outputs()
        for i in range(N):
            qpd = np.zeros(NQ)
            a = [np.ascontiguousarray(st[k][:, i]) for k in ("q", "qd", "base_pos", "base_quat", "base_twist")]
            r = [np.ascontiguousarray(ref[k][:, i]) for k in ("pose", "vel", "acc", "sw_pos", "sw_vel", "sw_acc")]
            c = np.ascontiguousarray(ref["contacts"][:, i])
            rbd.emul_tick(P(pp), *[P(x) for x in a], *[P(x) for x in r], P(c), int(ref["switching"][i]), int(t == 0), P(states[i]), P(qpd))
            x = np.zeros(42); y = np.zeros(112); ws = np.zeros(112, np.int8); nw = C.c_int(100)
            rc = qp.emulqp_solve(hs[i], P(qpd), int(t == 0), C.byref(nw), P(x), P(y), P(ws))
            xe = np.abs(x - o["x"][:, i]).max() / max(1.0, np.abs(o["x"][:, i]).max())
            okrc, oknw, okws = rc == o["rc"][i], nw.value == o["nwsr"][i], (ws == o["ws"][:, i]).all()
            stats["n"] += 1; stats["rc"] += okrc; stats["nwsr"] += oknw; stats["ws"] += okws
            if o["rc"][i] == 0: stats["xerr"] = max(stats["xerr"], xe)
            if verbose and not (okrc and oknw and okws and xe < 1e-9):
                print(f"  tick {t} inst {i}: rc {rc}/{o['rc'][i]} nwsr {nw.value}/{o['nwsr'][i]} ws_equal {okws} xerr {xe:.2e}")
    return stats


if __name__ == "__main__":
    print("nominal", run(1, "stance", ticks=3, nominal=True))
    for c in ("stance", "diagonal", "mixed"):
        print(c, run(int(sys.argv[1]) if len(sys.argv) > 1 else 8, c))
