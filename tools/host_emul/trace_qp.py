"""DEV AID: print the iteration path of the emulated solver and of the real qpOASES (PL_MEDIUM) for one instance."""
import ctypes as C, sys, os
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..")); sys.path.insert(0, HERE)
from oracle import oracle as orc
from quadrupedwholebodycontroller_b200 import workloads as wl
import check_rbd
rbd = C.CDLL("/tmp/libemul_rbd.so"); qp = C.CDLL("/tmp/libemul_qp_trace.so"); qp.emulqp_create.restype = C.c_void_p
NS, NQ = rbd.emul_sizes(0), rbd.emul_sizes(1); P = lambda a: a.ctypes.data_as(C.c_void_p)
mode, seed, inst, T = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
N = inst + 1
st, ref = wl.randomised(N, seed=seed, contacts=mode, switching_frac=0.1); pp = orc.pack_params()
state = np.zeros(NS); h = C.c_void_p(qp.emulqp_create())
Hs, gs, As, ls, us = [], [], [], [], []
for t in range(T):
    if t > 0: st = wl.advance_state(st)
    qpd = np.zeros(NQ); i = inst
    a = [np.ascontiguousarray(st[k][:, i]) for k in ("q", "qd", "base_pos", "base_quat", "base_twist")]
    r = [np.ascontiguousarray(ref[k][:, i]) for k in ("pose", "vel", "acc", "sw_pos", "sw_vel", "sw_acc")]
    c = np.ascontiguousarray(ref["contacts"][:, i])
    rbd.emul_tick(P(pp), *[P(x) for x in a], *[P(x) for x in r], P(c), int(ref["switching"][i]), int(t == 0), P(state), P(qpd))
    d = check_rbd.qp_dense(qpd); Hs.append(d["H"]); gs.append(d["g"]); As.append(d["A"]); ls.append(d["lbA"]); us.append(d["ubA"])
    x = np.zeros(42); y = np.zeros(112); ws = np.zeros(112, np.int8); nw = C.c_int(100)
    print("=== MINE tick", t); sys.stdout.flush()
    rc = qp.emulqp_solve(h, P(qpd), int(t == 0), C.byref(nw), P(x), P(y), P(ws)); print("rc", rc, "nwsr", nw.value); sys.stdout.flush()
os.environ["ORC_PRINT"] = "1"
print("=== QPOASES"); sys.stdout.flush()
o = orc.qp_sequence(np.array(Hs), np.array(gs), np.array(As), np.array(ls), np.array(us))
print(o["rc"], o["nwsr"])
