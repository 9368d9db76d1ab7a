"""DEV AID: compare the host-emulated RBD+assembly device code against the oracle (tools only; not shipped)."""
import ctypes as C, sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import oracle as orc
from quadrupedwholebodycontroller_b200 import workloads as wl

lib = C.CDLL("/tmp/libemul_rbd.so")
NS, NQ = lib.emul_sizes(0), lib.emul_sizes(1)
P = lambda a: a.ctypes.data_as(C.c_void_p)

def qp_dense(qp):
    """expand the compact QPData into the reference's dense H,g,A,lbA,ubA"""
    o = 0
    def take(n):
        nonlocal o
        v = qp[o:o + n]; o += n; return v
    Jc = take(216).reshape(12, 18); Mbb = take(36).reshape(6, 6); Mjj = take(144).reshape(12, 12); Hff = take(144).reshape(12, 12)
    gf = take(12); lbA = take(70); ubA = take(70); biasj = take(12); mu = take(1)[0]; rho = take(1)[0]; c = take(4)
    cm = np.repeat(c, 3)[:, None]
    Jst, Jsw = cm * Jc, (1 - cm) * Jc
    H = np.eye(42); H[18:30, 18:30] = Hff; H[30:, 30:] *= rho
    g = np.zeros(42); g[18:30] = gf
    A = np.zeros((70, 42))
    A[0:6, 0:6] = Mbb; A[0:6, 18:30] = -Jst[:, :6].T
    A[6:18, 0:6] = Jst[:, :6]; A[6:18, 6:18] = Jst[:, 6:]
    D = np.array([[1, 0, -mu], [-1, 0, -mu], [0, 1, -mu], [0, -1, -mu]])
    for l in range(4): A[18 + 4 * l:22 + 4 * l, 18 + 3 * l:21 + 3 * l] = D * c[l]
    A[34:46, 6:18] = Mjj; A[34:46, 18:30] = -Jst[:, 6:].T
    A[46:58, 0:18] = Jsw; A[46:58, 30:42] = -np.eye(12)
    A[58:70, 0:18] = Jsw; A[58:70, 30:42] = np.eye(12)
    return dict(H=H, g=g, A=A, lbA=lbA, ubA=ubA, biasj=biasj)

def run(N, contacts, ticks=3, seed=7):
    st, ref = wl.randomised(N, seed=seed, contacts=contacts, switching_frac=0.3)
    ob = orc.OracleBatch(N, lumped=False)
    pp = orc.pack_params()
    states = np.zeros((N, NS)); worst = {}
    for t in range(ticks):
        if t > 0:
            st = wl.advance_state(st)
        ob.set_state(**st); ob.set_reference(**ref); ob.step(assemble_only=True)
        for i in range(N):
            qp = np.zeros(NQ)
            a = [np.ascontiguousarray(st[k][:, i]) for k in ("q", "qd", "base_pos", "base_quat", "base_twist")]
            r = [np.ascontiguousarray(ref[k][:, i]) for k in ("pose", "vel", "acc", "sw_pos", "sw_vel", "sw_acc")]
            c = np.ascontiguousarray(ref["contacts"][:, i])
            lib.emul_tick(P(pp), *[P(x) for x in a], *[P(x) for x in r], P(c), int(ref["switching"][i]), int(t == 0), P(states[i]), P(qp))
            mine = qp_dense(qp); theirs = ob.qp(i)
            for k in ("H", "g", "A", "lbA", "ubA"):
                fin = np.abs(theirs[k]) < 1e19
                err = np.abs(mine[k] - theirs[k])[fin].max() / max(1.0, np.abs(theirs[k][fin]).max())
                assert (np.abs(mine[k][~fin]) > 1e19).all()
                worst[k] = max(worst.get(k, 0), err)
    return worst

if __name__ == "__main__":
    for c in ("stance", "diagonal", "mixed"):
        print(c, run(16, c))
