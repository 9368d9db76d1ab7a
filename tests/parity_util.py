"""Shared checks of the GPU parity tests (and of bench.py's `parity` block): comparison against the oracle's outputs and a KKT check
of the device's own working set where it differs from the oracle's."""
import numpy as np

RTOL = 1e-9          # BASELINE.json north_star: tau, f (and x) within 1e-9 relative in fp64


def relerr(a, b):
    """|a - b| / max(1, max|b|) per robot (arrays (K, N))"""
    return np.abs(a - b).max(axis=0) / np.maximum(1.0, np.abs(b).max(axis=0))


def kkt_violation(qp, x, y, ws):
    """KKT residuals of (x, y, working set) for the dense QP `qp` in the qpOASES convention (lib/qpOASES/src/QProblem.cpp:2602-2641:
    g + H x = y_B + A' y_C;  y >= 0 at an active lower side, <= 0 at an active upper side, 0 when inactive).  Returns a dict of maxima,
    each scaled by max(1, |data|):  stationarity, feasibility of A x, activity (an active constraint sits on its bound), sign, inactive duals."""
    H, g, A, lbA, ubA = qp["H"], qp["g"], qp["A"], qp["lbA"], qp["ubA"]
    yB, yC = y[:42], y[42:]
    grad = H @ x + g
    stat = np.abs(grad - yB - A.T @ yC).max() / max(1.0, np.abs(grad).max(), np.abs(A.T @ yC).max())
    Ax = A @ x; sc = max(1.0, np.abs(Ax).max())
    lo, up = np.where(lbA > -1e19, lbA, -np.inf), np.where(ubA < 1e19, ubA, np.inf)
    feas = max(0.0, (lo - Ax).max(), (Ax - up).max()) / sc
    wsC = ws[42:]
    act = 0.0
    if (wsC == -1).any():
        act = max(act, np.abs(Ax - lbA)[wsC == -1].max() / sc)
    if (wsC == 1).any():
        act = max(act, np.abs(Ax - ubA)[wsC == 1].max() / sc)
    ysc = max(1.0, np.abs(y).max())
    sign = max(0.0, (-yC[wsC == -1]).max(initial=0.0), (yC[wsC == 1]).max(initial=0.0)) / ysc
    inactive = max(np.abs(yC[wsC == 0]).max(initial=0.0), np.abs(yB[ws[:42] == 0]).max(initial=0.0)) / ysc
    return dict(stationarity=stat, feasibility=feas, activity=act, sign=sign, inactive=inactive)


def compare(got, want, what, ctl=None, duals=None, min_ws_equal=0.95, kkt_tol=1e-7, check_failed_x=True):
    """rc identical for every robot; tau, grf, x within RTOL wherever the oracle solved (rc 0); robots whose QP did NOT solve must publish the
    same x / grf the oracle keeps (previous solution, wbc.cpp:536) and zero torques.  Working sets are compared exactly; a robot whose set differs
    from the oracle's must (i) still agree in x to RTOL (checked above) and (ii) satisfy the KKT conditions with the DEVICE's own set and duals
    (needs ctl + duals): two working sets can describe one degenerate vertex, but neither may carry a wrong-signed or spurious multiplier."""
    assert (got["rc"] == want["rc"]).all(), f"{what}: return codes differ at robots {np.nonzero(got['rc'] != want['rc'])[0][:10]}"
    ok = want["rc"] == 0
    worst = {}
    for k in ("tau", "grf", "x"):
        e = relerr(got[k], want[k]); worst[k] = float(e[ok].max(initial=0.0))
        assert worst[k] < RTOL, f"{what}: {k} rel err {worst[k]:.3e} at robot {int(np.argmax(np.where(ok, e, 0)))}"
    if check_failed_x and (~ok).any():
        assert not got["tau"][:, ~ok].any(), f"{what}: failed robots must publish zero torques"
        for k in ("x", "grf"):
            e = relerr(got[k], want[k])[~ok].max()
            assert e < RTOL, f"{what}: {k} of robots with rc != 0 differs from the oracle's kept solution ({e:.3e})"
    same_ws = (got["ws"] == want["ws"]).all(axis=0)
    frac = float(same_ws[ok].mean()) if ok.any() else 1.0
    assert frac >= min_ws_equal, f"{what}: working sets equal for only {frac * 100:.2f} %"
    checked = 0
    if ctl is not None and duals is not None:
        for i in np.nonzero(ok & ~same_ws)[0]:
            v = kkt_violation(ctl.denseQP(int(i)), got["x"][:, i], duals[:, i], got["ws"][:, i])
            assert max(v.values()) < kkt_tol, f"{what}: robot {i} has a working set different from the oracle's AND violates KKT: {v}"
            checked += 1
    return dict(ws_equal=frac, nwsr_equal=float((got["nwsr"] == want["nwsr"])[ok].mean()) if ok.any() else 1.0, ws_diff_kkt_checked=checked,
                n=int(ok.size), n_ok=int(ok.sum()), **{f"max_rel_{k}": v for k, v in worst.items()})
