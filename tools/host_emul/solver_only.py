"""DEV AID: solver-only comparison on the CPU -- the emulated solver (one lane) and the real qpOASES are given the SAME dense QPs (the emulated
assembly's), init then hotstart ticks.  usage: solver_only.py lib.so contacts N ticks [seed]"""
import ctypes as C, sys, os
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..")); sys.path.insert(0, HERE)
from oracle import oracle as orc
from quadrupedwholebodycontroller_b200 import workloads as wl
import check_rbd
rbd = C.CDLL("/tmp/libemul_rbd.so"); qp = C.CDLL(sys.argv[1]); qp.emulqp_create.restype = C.c_void_p
NS, NQ = rbd.emul_sizes(0), rbd.emul_sizes(1); P = lambda a: a.ctypes.data_as(C.c_void_p)
mode, N, T = sys.argv[2], int(sys.argv[3]), int(sys.argv[4]); seed = int(sys.argv[5]) if len(sys.argv) > 5 else 4242
st, ref = (wl.stress(N) if mode == "stress" else wl.randomised(N, seed=seed, contacts=mode, switching_frac=0.1)); pp = orc.pack_params()
states = np.zeros((N, NS)); hs = [C.c_void_p(qp.emulqp_create()) for _ in range(N)]
mine = [[] for _ in range(N)]; dens = [[] for _ in range(N)]
for t in range(T):
    if t: st = wl.advance_state(st)
    if t == 2 and mode == "mixed":
        ref = dict(ref); ref["contacts"] = np.ascontiguousarray(np.roll(ref["contacts"], 1, axis=0))
    for i in range(N):
        qpd = np.zeros(NQ)
        a = [np.ascontiguousarray(st[k][:, i]) for k in ("q", "qd", "base_pos", "base_quat", "base_twist")]
        r = [np.ascontiguousarray(ref[k][:, i]) for k in ("pose", "vel", "acc", "sw_pos", "sw_vel", "sw_acc")]
        c = np.ascontiguousarray(ref["contacts"][:, i])
        rbd.emul_tick(P(pp), *[P(x) for x in a], *[P(x) for x in r], P(c), int(ref["switching"][i]), int(t == 0), P(states[i]), P(qpd))
        dens[i].append(check_rbd.qp_dense(qpd))
        x = np.zeros(42); y = np.zeros(112); ws = np.zeros(112, np.int8); nw = C.c_int(100)
        rc = qp.emulqp_solve(hs[i], P(qpd), int(t == 0), C.byref(nw), P(x), P(y), P(ws))
        mine[i].append((rc, nw.value, ws.copy(), x.copy()))
fd = os.open(os.devnull, os.O_WRONLY); saved = os.dup(2); os.dup2(fd, 2)
bad = 0; badws = 0; tot = 0; xe = 0.0; hot_bad = 0
for i in range(N):
    o = orc.qp_sequence(*[np.array([d[k] for d in dens[i]]) for k in ("H", "g", "A", "lbA", "ubA")])
    for t in range(T):
        rc, nw, ws, x = mine[i][t]; tot += 1
        if rc != o["rc"][t] or nw != o["nwsr"][t]:
            bad += 1; hot_bad += t > 0
            if os.environ.get("VERBOSE"): os.write(saved, f"  robot {i} tick {t}: rc {rc}/{o["rc"][t]} nwsr {nw}/{o["nwsr"][t]}\n".encode())
        if not (ws == o["ws"][t]).all(): badws += 1
        if o["rc"][t] == 0: xe = max(xe, np.abs(x - o["x"][t]).max() / max(1.0, np.abs(o["x"][t]).max()))
os.dup2(saved, 2)
print(f"{sys.argv[1]} {mode} N={N} T={T}: rc/nWSR differ on {bad} of {tot} (tick, robot) pairs ({hot_bad} on hot-started ticks), working sets differ on {badws}, max x rel err {xe:.2e}")
