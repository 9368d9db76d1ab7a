"""CPU: the oracle against the committed golden vectors (tests/golden/wbc_golden.npz, made by tools/make_golden.py) and the
golden solutions against the KKT conditions of their own QPs -- the criterion qpOASES' unit tests use
(lib/qpOASES/testing/cpp/test_example1.cpp:77-111: stationarity / feasibility / complementarity residuals)."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "tools"))
import make_golden as mg          # noqa: E402

GOLD = np.load(os.path.join(HERE, "golden", "wbc_golden.npz"))


@pytest.mark.parametrize("name", list(mg.CASES))
def test_oracle_reproduces_golden(orc, wl, name):
    spec = mg.CASES[name]
    ob = orc.OracleBatch(spec["N"])
    st = ref = None
    for t in range(spec["ticks"]):
        st, ref = mg.case_inputs(spec, t, st, ref)
        if spec.get("swing") and t == 0:
            ref = dict(ref); ref["sw_pos"] = GOLD[f"{name}_swpos"]
        ob.set_state(**st); ob.set_reference(**ref); ob.step()
        o = ob.outputs()
        assert (o["rc"] == GOLD[f"{name}_t{t}_rc"]).all() and (o["nwsr"] == GOLD[f"{name}_t{t}_nwsr"]).all()
        assert (o["ws"] == GOLD[f"{name}_t{t}_ws"]).all()
        for k in ("tau", "grf", "x"):
            assert np.abs(o[k] - GOLD[f"{name}_t{t}_{k}"]).max() <= 1e-12 * max(1.0, np.abs(o[k]).max()), k


@pytest.mark.parametrize("name,t", [(n, t) for n, s in mg.CASES.items() for t in range(s["ticks"])])
def test_golden_solution_satisfies_kkt(name, t):
    H, g, A = GOLD[f"{name}_t{t}_qp0_H"], GOLD[f"{name}_t{t}_qp0_g"], GOLD[f"{name}_t{t}_qp0_A"]
    lbA, ubA = GOLD[f"{name}_t{t}_qp0_lbA"], GOLD[f"{name}_t{t}_qp0_ubA"]
    x, ws = GOLD[f"{name}_t{t}_x"][:, 0], GOLD[f"{name}_t{t}_ws"][:, 0]
    assert GOLD[f"{name}_t{t}_rc"][0] == 0
    Ax = A @ x
    scale = max(1.0, np.abs(Ax).max())
    assert (Ax >= lbA - 1e-8 * scale).all() and (Ax <= ubA + 1e-8 * scale).all()             # primal feasibility
    assert (ws[:42] == 0).all()                                                               # no far bound active at the solution
    act = np.nonzero(ws[42:])[0]
    # stationarity with multipliers supported on the active set (least squares; the working set may be degenerate)
    grad = H @ x + g
    lam = np.linalg.lstsq(A[act].T, grad, rcond=None)[0]
    assert np.abs(A[act].T @ lam - grad).max() < 1e-7 * max(1.0, np.abs(grad).max())
    # active constraints sit on their bound
    for i in act:
        b = lbA[i] if ws[42 + i] < 0 else ubA[i]
        assert abs(Ax[i] - b) < 1e-7 * scale


def test_nominal_stance_is_physically_plausible():
    f = GOLD["c1_t0_grf"][:, 0].reshape(4, 3)
    assert (f[:, 2] > 0).all()                                              # unilateral contacts
    assert (np.abs(f[:, 0]) <= f[:, 2] + 1e-9).all() and (np.abs(f[:, 1]) <= f[:, 2] + 1e-9).all()    # friction pyramid, mu = 1
    assert np.abs(GOLD["c1_t0_tau"]).max() <= 80.0 + 1e-9                   # torque limit
    assert 0.5 * 51.63574 * 9.81 < f[:, 2].sum() < 1.2 * 51.63574 * 9.81    # carries (most of) the weight
    assert GOLD["c1_t1_nwsr"][0] == 0                                       # same state again: hot start needs no working-set change
