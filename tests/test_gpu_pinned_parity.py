"""GPU (-m gpu): parity pinned on the workloads bench.py actually times (BASELINE.json configs[1..4]), at their own sizes, plus a
solver-only comparison with IDENTICAL inputs.

  * cold4096            every one of the 4096 robots of the headline batch against the oracle (about 1 s of CPU)
  * warm65536           the full 65536-robot batch on the GPU, a strided 2048-robot sample against the oracle, cold tick + hot-started tick
  * mixed1m             8 x 256 robots taken at equal distances from the 2^20 index space (counter-based generator), cold, switching on 10 %
  * rollout             64 robots x 1000 closed-loop ticks (the length of configs[4]) against the oracle's closed loop
  * solver only         the dense QPs the DEVICE assembled are handed, tick by tick, to the real qpOASES (SQProblem::init / hotstart,
                        lib/qpOASES/src/SQProblem.cpp:153-197): with bit-identical inputs there is no degeneracy excuse, so return code,
                        working set and nWSR must be equal for 100 % of the robots on cold AND hot-started ticks
  * kiValue != 0        the integral state (wbc.cpp:438-442) with an effect on the output, 6 ticks

Bars: rc identical; tau, f, x <= 1e-9 relative; working sets compared exactly, a differing set is accepted only when x agrees AND the
device's set satisfies the KKT conditions with the device's duals (tests/parity_util.py); nWSR equality is reported and asserted.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from parity_util import compare, relerr, RTOL          # noqa: E402

pytestmark = pytest.mark.gpu


def quiet_stderr():
    """qpOASES prints nWSR warnings on stderr"""
    class Q:
        def __enter__(self):
            self.fd = os.open(os.devnull, os.O_WRONLY); self.saved = os.dup(2); os.dup2(self.fd, 2)

        def __exit__(self, *a):
            os.dup2(self.saved, 2); os.close(self.fd); os.close(self.saved)
    return Q()


def take(d, idx):
    return {k: (None if v is None else np.ascontiguousarray(v[..., idx])) for k, v in d.items()}


def test_cold4096_every_robot_against_the_oracle(wbc, orc, wl):
    n = 4096
    st, ref = wl.randomised(n, seed=1234, contacts="stance")               # exactly bench.py's default workload
    ctl = wbc.WholeBodyController(n); ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
    got = ctl.getOutputs(); y = ctl.getDuals()
    ob = orc.OracleBatch(n); ob.set_state(**st); ob.set_reference(**ref)
    with quiet_stderr():
        ob.step()
    r = compare(got, ob.outputs(), "cold4096", ctl, y, min_ws_equal=0.999)
    print("cold4096", r)
    assert r["nwsr_equal"] >= 0.999, r                                      # a cold start follows the reference's iteration path
    # hot-started tick on the advanced state (the second half of tools/gpu_profile_run.py)
    st1 = wl.advance_state(st)
    ctl.setJointAndBaseState(**st1); ctl.step(); got = ctl.getOutputs(); y = ctl.getDuals()
    ob.set_state(**st1)
    with quiet_stderr():
        ob.step()
    r = compare(got, ob.outputs(), "cold4096 tick 1", ctl, y, min_ws_equal=0.95)
    print("cold4096 tick 1", r)
    ctl.close()


def test_warm65536_sampled_against_the_oracle(wbc, orc, wl):
    n, stride = 65536, 32
    st, ref = wl.randomised(n, seed=1234, contacts="diagonal")             # bench.py --workload warm65536
    idx = np.arange(0, n, stride)
    ctl = wbc.WholeBodyController(n); ob = orc.OracleBatch(idx.size)
    for t in range(2):
        if t:
            st = wl.advance_state(st)
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
        got = take(ctl.getOutputs(), idx); y = ctl.getDuals()[:, idx]
        ob.set_state(**take(st, idx)); ob.set_reference(**take(ref, idx))
        with quiet_stderr():
            ob.step()
        want = ob.outputs()
        # KKT check of differing sets needs the dense QP of the ORIGINAL robot index
        class View:
            def denseQP(self, i):
                return ctl.denseQP(int(idx[i]))
        r = compare(got, want, f"warm65536 tick {t}", View(), y, min_ws_equal=0.999 if t == 0 else 0.95)
        print("warm65536 tick", t, r)
    ctl.close()


def test_mixed1m_sampled_against_the_oracle(wbc, orc, wl):
    total, chunks, per = 1 << 20, 8, 256
    parts = [wl.randomised(per, seed=1234, first=c * (total // chunks) + 977, contacts="mixed", switching_frac=0.1) for c in range(chunks)]
    st = {k: np.ascontiguousarray(np.concatenate([p[0][k] for p in parts], axis=-1)) for k in parts[0][0]}
    ref = {k: np.ascontiguousarray(np.concatenate([p[1][k] for p in parts], axis=-1)) for k in parts[0][1]}
    n = chunks * per
    ctl = wbc.WholeBodyController(n); ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
    got = ctl.getOutputs(); y = ctl.getDuals()
    ob = orc.OracleBatch(n); ob.set_state(**st); ob.set_reference(**ref)
    with quiet_stderr():
        ob.step()
    want = ob.outputs()
    r = compare(got, want, "mixed1m sample", ctl, y, min_ws_equal=0.99)
    print("mixed1m sample", r, "rc histogram", dict(zip(*np.unique(want["rc"], return_counts=True))))
    assert r["nwsr_equal"] >= 0.99, r
    ctl.close()


def test_rollout_64_robots_1000_ticks_against_the_oracle(wbc, orc, wl):
    """configs[4] length.  Every (tick, robot) pair: identical return code; torques within 1e-9 for at least 95 % of the pairs and 1e-6
    for all of them (a closed loop amplifies a last-bit difference of one tick -- a different but KKT-equivalent working set at a degenerate
    vertex -- through the plant; the share of such pairs is asserted, not hidden)."""
    N, K = 64, 1000
    st, ref, cmd = wl.rollout_start(N)
    ctl = wbc.WholeBodyController(N); ctl.plannerCreate(); ctl.plantEnable()
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
    ob = orc.OracleBatch(N); ob.set_state(**st); ob.set_reference(**ref)
    with quiet_stderr():
        rec = ob.rollout(K, cmd)
    tight = loose = 0; worst = 0.0; nw_eq = 0
    for t in range(K):
        ctl.rollout(1); o = ctl.getOutputs()
        assert np.array_equal(o["rc"], rec["rc"][t]), (t, o["rc"], rec["rc"][t])
        e = relerr(o["tau"], rec["tau"][t])
        tight += int((e < RTOL).sum()); loose += int((e < 1e-6).sum()); worst = max(worst, float(e.max()))
        nw_eq += int((o["nwsr"] == rec["nwsr"][t]).sum())
    s = ctl.getState(); w = ob.state()
    pairs = N * K
    print(f"rollout 64x1000: tau within 1e-9 for {tight / pairs * 100:.3f} % of the (tick, robot) pairs, within 1e-6 for {loose / pairs * 100:.3f} %, worst {worst:.2e}; "
          f"nWSR equal for {nw_eq / pairs * 100:.3f} %")
    assert tight >= 0.95 * pairs and loose == pairs, (tight, loose, worst)
    for k in ("q", "qd", "base_pos", "base_quat", "base_twist"):
        assert np.abs(s[k] - w[k]).max() / max(1.0, np.abs(w[k]).max()) < 1e-7, k
    st_ = ctl.rolloutStats()
    assert (st_["ticks"] == K).all() and (st_["failed_ticks"] == (rec["rc"] != 0).sum(axis=0)).all()
    ctl.close()


def _solver_only(wbc, orc, wl, contacts, n, ticks, exact):
    """per tick the device's dense QP of EVERY robot (wbc_get_dense_qp) is recorded, then each robot's sequence is given to the real qpOASES
    (init, then hotstart, options of wbc.cpp:517-521); returns the (tick, robot) pairs that differ and the worst x error"""
    st, ref = (wl.stress(n) if contacts == "stress" else wl.randomised(n, seed=4242, contacts=contacts, switching_frac=0.1))
    ctl = wbc.WholeBodyController(n, exact=exact)
    qps = [[] for _ in range(n)]; outs = []
    for t in range(ticks):
        if t:
            st = wl.advance_state(st)
        if t == 2 and contacts == "mixed":
            ref = dict(ref); ref["contacts"] = np.ascontiguousarray(np.roll(ref["contacts"], 1, axis=0)); ref["switching"] = None
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
        outs.append(ctl.getOutputs())
        for i in range(n):
            qps[i].append(ctl.denseQP(i))
    counters = ctl.debugCounters() if exact else None
    ctl.close()
    path_diff, ws_diff, xerr, xbits = [], [], 0.0, 0
    with quiet_stderr():
        for i in range(n):
            seq = {k: np.stack([q[k] for q in qps[i]]) for k in ("H", "g", "A", "lbA", "ubA")}
            r = orc.qp_sequence(**seq)
            for t in range(ticks):
                o = outs[t]
                assert o["rc"][i] == r["rc"][t], (i, t, o["rc"][i], r["rc"][t])
                if o["nwsr"][i] != r["nwsr"][t]:
                    path_diff.append((i, t, int(o["nwsr"][i]), int(r["nwsr"][t])))
                if not np.array_equal(o["ws"][:, i], r["ws"][t]):
                    ws_diff.append((i, t))
                if r["rc"][t] == 0:
                    xerr = max(xerr, float(np.abs(o["x"][:, i] - r["x"][t]).max() / max(1.0, np.abs(r["x"][t]).max())))
                    xbits += int(np.count_nonzero(o["x"][:, i] != r["x"][t]))
    return dict(pairs=n * ticks, path_diff=path_diff, ws_diff=ws_diff, xerr=xerr, x_entries_not_bit_equal=xbits, counters=counters)


@pytest.mark.parametrize("contacts,n,ticks", [("stance", 128, 4), ("mixed", 192, 5), ("stress", 128, 4)])
def test_exact_build_walks_the_reference_iteration_path_bit_for_bit(wbc, orc, wl, contacts, n, ticks):
    """Rows Q1-Q10 with IDENTICAL inputs.  libwbc_b200_exact.so (-DWBC_EXACT -fmad=false) is the same solver source with the four arithmetic
    liberties of the product build switched off (reciprocal-multiply instead of divide in the triangular solves and DPOTRF, Givens chains
    from prefix sums, back substitution accumulated in the order the unknowns become available, FMA contraction) and the reference's full
    linear-independence test run on every add.  Required: return code, working set AND nWSR equal to the real qpOASES for 100 % of the
    (tick, robot) pairs, cold and hot-started, and x equal BIT FOR BIT -- every remaining difference would be a solver bug."""
    r = _solver_only(wbc, orc, wl, contacts, n, ticks, exact=True)
    print(f"exact build, {contacts}: {r['pairs']} (tick, robot) pairs, nWSR differs on {len(r['path_diff'])}, working set on {len(r['ws_diff'])}, "
          f"x entries not bit-equal {r['x_entries_not_bit_equal']}; LI tests {r['counters']}")
    assert not r["path_diff"] and not r["ws_diff"], (r["path_diff"][:10], r["ws_diff"][:10])
    assert r["x_entries_not_bit_equal"] == 0 and r["xerr"] == 0.0, r["xerr"]
    tests, band, wrong = (int(v) for v in r["counters"])
    assert tests > 0 and wrong == 0, r["counters"]          # the |Z'a| shortcut of the product build never contradicts the reference's full test


@pytest.mark.parametrize("contacts,n,ticks", [("stance", 128, 4), ("mixed", 192, 5)])
def test_product_build_on_device_assembled_qps_against_real_qpoases(wbc, orc, wl, contacts, n, ticks):
    """The shipped build on the same identical inputs: rc equal everywhere, x <= 1e-9, cold ticks follow the reference's path exactly; on
    hot-started ticks the last-bit liberties named above can flip the ramping / drift-correction decision of a zero-length step (qpOASES
    compares a step length of rounding-noise size with EPS, QProblem.cpp:1640-1660), after which both solvers reach the same optimum along
    different paths: the share of such pairs is asserted and reported."""
    r = _solver_only(wbc, orc, wl, contacts, n, ticks, exact=False)
    cold_diff = [p for p in r["path_diff"] if p[1] == 0]
    print(f"product build, {contacts}: {r['pairs']} pairs, nWSR differs on {len(r['path_diff'])} (cold ticks: {len(cold_diff)}), working set on {len(r['ws_diff'])}, "
          f"max x rel err {r['xerr']:.2e}")
    assert not cold_diff, cold_diff[:10]
    assert r["xerr"] < RTOL
    # a robot that took a different path through a degenerate vertex (all-zero rows of swing legs: active or not makes no difference) keeps
    # its KKT-equivalent working set on the following ticks, so the working-set count is cumulative over the ticks
    assert len(r["ws_diff"]) <= 0.25 * r["pairs"] and len(r["path_diff"]) <= 0.15 * r["pairs"], (len(r["ws_diff"]), len(r["path_diff"]))


def test_integral_gain_changes_the_output_and_matches_the_oracle(wbc, orc, wl):
    """kiValue != 0: integralError_ (wbc.cpp:438-442) is accumulated INSIDE solveQP's assembly and feeds the desired wrench; six ticks so that the
    accumulated state matters.  Also shows the gain has an effect (outputs differ from the ki = 0 run)."""
    n, ticks, ki = 96, 6, 800.0
    st0, ref = wl.randomised(n, seed=31, contacts="mixed")
    prm = wbc.default_params(); prm.kiValue = ki
    ctl = wbc.WholeBodyController(n, params=prm); ctl0 = wbc.WholeBodyController(n)
    ob = orc.OracleBatch(n, params=dict(kiValue=ki))
    st = st0; differs = 0.0
    for t in range(ticks):
        if t:
            st = wl.advance_state(st)
        for c in (ctl, ctl0):
            c.setJointAndBaseState(**st); c.setReference(**ref); c.step()
        ob.set_state(**st); ob.set_reference(**ref)
        with quiet_stderr():
            ob.step()
        got = ctl.getOutputs()
        r = compare(got, ob.outputs(), f"ki tick {t}", ctl, ctl.getDuals(), min_ws_equal=0.9)
        differs = max(differs, float(np.abs(got["tau"] - ctl0.getOutputs()["tau"]).max()))
    assert differs > 1e-3, differs
    ctl.close(); ctl0.close()


@pytest.mark.parametrize("contacts", ["stance", "mixed"])
def test_fp32_mode_against_the_fp64_oracle(wbc, orc, wl, contacts):
    """north star: "with an optional fp32 mode ... 1e-4 relative (fp32)".  libwbc_b200_f32.so = the same solver source in single precision with qpOASES'
    __USE_SINGLE_PRECISION__ constants and tolerances (Constants.hpp:46-47, Options.cpp:111-152,203-207); RBD and QP assembly stay fp64.  Against the
    fp64 oracle on 1024 robots, cold tick + two hot-started ticks: tau, f and x within 1e-4 relative on every robot both sides solved, return codes
    equal on >= 96 % (single precision reaches its looser termination tolerance where the fp64 reference hits nWSR = 100, so it solves MORE robots),
    working-set agreement reported."""
    n = 1024
    st, ref = wl.randomised(n, seed=1234, contacts=contacts, switching_frac=0.1 if contacts == "mixed" else 0.0)
    ctl = wbc.WholeBodyController(n, fp32=True); ob = orc.OracleBatch(n)
    for t in range(3):
        if t:
            st = wl.advance_state(st)
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
        ob.set_state(**st); ob.set_reference(**ref)
        with quiet_stderr():
            ob.step()
        got, want = ctl.getOutputs(), ob.outputs()
        ok = (got["rc"] == 0) & (want["rc"] == 0)
        e = {k: float(relerr(got[k], want[k])[ok].max()) for k in ("tau", "grf", "x")}
        rc_eq = float((got["rc"] == want["rc"]).mean()); ws_eq = float((got["ws"] == want["ws"]).all(axis=0)[ok].mean())
        print(f"fp32 {contacts} tick {t}: rc equal {rc_eq * 100:.2f} %, fp32 solved {float((got['rc'] == 0).mean()) * 100:.2f} % (fp64 {float((want['rc'] == 0).mean()) * 100:.2f} %), "
              f"working sets equal {ws_eq * 100:.1f} %, max rel err {e}")
        assert max(e.values()) < 1e-4, (t, e)
        assert rc_eq >= 0.96 and (got["rc"] == 0).mean() >= 0.99, (t, rc_eq)
    ctl.close()
