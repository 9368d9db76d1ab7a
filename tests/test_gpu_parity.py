"""GPU (-m gpu): the CUDA path, called through the C ABI, against (i) the committed golden vectors, (ii) the oracle on the
same seeded inputs at sizes the oracle finishes in seconds, and (iii) size-independent properties at full batch sizes.

Tolerance (BASELINE.json north_star): torques and contact forces within 1e-9 relative in fp64, identical qpOASES return
code and working set.  "Relative" = |a - b| / max(1, max|b|) per robot.  Working sets are compared exactly; where the
active set is degenerate (exactly zero rows of swing/stance legs) two different working sets can describe the same
vertex, so a mismatch is accepted only if the primal solution still agrees to 1e-9 and the device's set satisfies the KKT
conditions with the device's duals (tests/parity_util.py: stationarity, feasibility, activity, multiplier signs).
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "tools"))
import make_golden as mg          # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(HERE, "golden", "wbc_golden.npz"))
RTOL = 1e-9


sys.path.insert(0, HERE)
from parity_util import compare, relerr          # noqa: E402


def check_against(got, want, what, ctl=None, min_ws_equal=0.75):
    """tests/parity_util.py: rc identical, tau / grf / x <= 1e-9, robots with rc != 0 publish the oracle's kept solution, working sets compared
    exactly and every robot whose set differs from the oracle's must satisfy the KKT conditions with the DEVICE's set and duals.  The share of
    equal sets is bounded from below only loosely: at the degenerate vertices of this QP family (all-zero rows of swing legs)
    the path through a zero-length step is decided by rounding noise, and the oracle's QP differs from the device's in the last bits (dense
    18x18 LU vs closed form).  The sharp statement -- 100 % equal paths on identical inputs -- is tests/test_gpu_pinned_parity.py."""
    return compare(got, want, what, ctl, ctl.getDuals() if ctl is not None else None, min_ws_equal=min_ws_equal)


@pytest.mark.parametrize("name", list(mg.CASES))
def test_golden_vectors(wbc, name):
    spec = mg.CASES[name]
    ctl = wbc.WholeBodyController(spec["N"])
    st = ref = None
    for t in range(spec["ticks"]):
        st, ref = mg.case_inputs(spec, t, st, ref)
        if spec.get("swing") and t == 0:
            ref = dict(ref); ref["sw_pos"] = GOLD[f"{name}_swpos"]
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref)
        ctl.updateState(); ctl.solveQP(); ctl.computeJointTorques()
        got = ctl.getOutputs()
        want = {k: GOLD[f"{name}_t{t}_{k}"] for k in ("tau", "grf", "x", "rc", "nwsr", "ws")}
        check_against(got, want, f"{name} tick {t}", ctl)
        q = ctl.denseQP(0)
        for k in ("H", "g", "A", "lbA", "ubA"):
            w = GOLD[f"{name}_t{t}_qp0_{k}"]; fin = np.abs(w) < 1e19
            assert np.abs(q[k] - w)[fin].max() <= 1e-11 * max(1.0, np.abs(w[fin]).max()), (name, t, k)
            assert (np.abs(q[k][~fin]) >= 1e19).all()
    ctl.close()


@pytest.mark.parametrize("contacts,n,ticks", [("stance", 96, 3), ("diagonal", 96, 3), ("mixed", 128, 4)])
def test_against_oracle_on_seeded_batches(wbc, orc, wl, contacts, n, ticks):
    st, ref = wl.randomised(n, seed=99, contacts=contacts, switching_frac=0.1)
    ctl = wbc.WholeBodyController(n); ob = orc.OracleBatch(n)
    stats = []
    for t in range(ticks):
        if t:
            st = wl.advance_state(st)
        if t == 2 and contacts == "mixed":                      # a contact switch mid-sequence, `switching` derived like wbc.cpp:176-184
            ref = dict(ref); ref["contacts"] = np.ascontiguousarray(np.roll(ref["contacts"], 1, axis=0)); ref["switching"] = None
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
        ob.set_state(**st); ob.set_reference(**ref); ob.step()
        stats.append(check_against(ctl.getOutputs(), ob.outputs(), f"{contacts} tick {t}", ctl))
    assert stats[0]["nwsr_equal"] >= 0.9, stats        # cold solves follow the same iteration path
    ctl.close()


@pytest.mark.parametrize("max_torque", [80.0, 3.0])
def test_edge_cases_against_oracle(wbc, orc, wl, max_torque):
    """workloads.stress: flight (no stance foot), one stance foot, saturating references, fast joints, tilted base, far swing targets,
    one side in the air, deep crouch -- 4 ticks each.  With maxTorque = 3 Nm most QPs are infeasible: the cold start stops at nWSR = 100
    (rc 64) or fails outright (rc 37, RET_INIT_FAILED_INFEASIBILITY), later hot starts from the unfinished state stop with rc 61
    (RET_HOTSTART_STOPPED_INFEASIBILITY) and keep being called, like the oracle does.  Return codes must be identical tick by tick;
    failed robots publish zero torques (what terminate() sends, wbc.cpp:626-635)."""
    n = 128
    st, ref = wl.stress(n)
    prm = wbc.default_params(); prm.maxTorque = max_torque
    ctl = wbc.WholeBodyController(n, params=prm); ob = orc.OracleBatch(n, params=dict(maxTorque=max_torque))
    seen = set()
    for t in range(4):
        if t:
            st = wl.advance_state(st)
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step()
        ob.set_state(**st); ob.set_reference(**ref); ob.step()
        got, want = ctl.getOutputs(), ob.outputs()
        check_against(got, want, f"stress maxTorque {max_torque} tick {t}", ctl, min_ws_equal=0.7)       # incl. x / grf of the robots whose QP did not solve
        seen |= set(int(r) for r in want["rc"])
    assert seen >= ({0, 64} if max_torque == 80.0 else {0, 37, 61, 64}), seen
    ctl.close()


def test_three_methods_equal_step_and_reset_restarts(wbc, wl):
    st, ref = wl.randomised(32, seed=5, contacts="mixed")
    a = wbc.WholeBodyController(32); b = wbc.WholeBodyController(32)
    for c in (a, b):
        c.setJointAndBaseState(**st); c.setReference(**ref)
    a.step(); b.updateState(); b.solveQP(); b.computeJointTorques()
    oa, ob_ = a.getOutputs(), b.getOutputs()
    for k in oa:
        assert np.array_equal(oa[k], ob_[k]), k
    first = oa
    a.step()                                                      # hot start on the same state
    assert (a.getOutputs()["nwsr"] <= first["nwsr"]).all()
    a.setInitialState(); a.setJointAndBaseState(**st); a.setReference(**ref); a.step()     # reset => cold start again, bit-identical
    again = a.getOutputs()
    for k in first:
        assert np.array_equal(first[k], again[k]), k
    a.close(); b.close()


def test_two_kernel_warm_start_is_bit_identical_to_the_single_kernel_path(wbc, wl):
    """The TQ rebuild of the warm start runs as a kernel of its own ahead of the solve (wbc_warm_setup_kernel); wbc_set_option("warm_start_split", 0) keeps it
    inside the solve kernel.  Same functions, same arithmetic: every output of every tick must be bit-identical, also for robots that are
    reset (cold) in the middle of the sequence while the others stay warm."""
    n = 192
    st, ref = wl.randomised(n, seed=7, contacts="mixed", switching_frac=0.1)
    outs = []
    for split in (1, 0):
        ctl = wbc.WholeBodyController(n); ctl.setOption("warm_start_split", split); s = st; rec = []
        for t in range(4):
            if t:
                s = wl.advance_state(s)
            if t == 2:
                ctl.setInitialState(first=40, count=50)              # these 50 robots start cold again, the rest hot-start
            ctl.setJointAndBaseState(**s); ctl.setReference(**ref); ctl.step(); rec.append(ctl.getOutputs())
        ctl.close(); outs.append(rec)
    for t in range(4):
        for k in outs[0][t]:
            assert np.array_equal(outs[0][t][k], outs[1][t][k]), (t, k)
    assert (outs[0][2]["nwsr"][40:90] >= 40).all()                    # the reset robots did cold-start at tick 2


@pytest.mark.parametrize("n,chunk", [(1, 0), (5, 0), (97, 0), (300, 128), (1500, 0)])
def test_parked_cold_solves_are_bit_identical_to_the_single_launch(wbc, wl, n, chunk):
    """All-cold batches run as two launches of wbc_solve_cold_kernel: gang fetch until the first member of a gang is done, the others PARK (workspace, tensor-memory
    vectors and context to HBM), then per-warp fetch over the parked solves, each continued from the top of the iteration it stopped at.  wbc_set_option("park", 0)
    keeps the single launch of the general kernel.  Same functions, same arithmetic: every output must be bit-identical -- robots that hit the nWSR limit (rc 64)
    included, for batches smaller than a gang, not divisible by it, and larger than one chunk of the parking buffer -- and so must the hot-started tick that follows
    (the parked solves hand over the same solver state)."""
    st, ref = wl.randomised(n, seed=21, contacts="mixed", switching_frac=0.1)
    outs = []
    for park in (1, 0):
        ctl = wbc.WholeBodyController(n); ctl.setOption("park", park)
        if park and chunk:
            ctl.setOption("park_chunk", chunk)
        rec = []; s = st
        for t in range(3):
            if t == 1:
                s = wl.advance_state(s)
            if t == 2:
                ctl.setInitialState()                                # a second cold batch on the same handle (the parking buffer is reused)
            ctl.setJointAndBaseState(**s); ctl.setReference(**ref); ctl.step(); o = ctl.getOutputs(); o["y"] = ctl.getDuals(); rec.append(o)
        ctl.close(); outs.append(rec)
    for t in range(3):
        for k in outs[0][t]:
            assert np.array_equal(outs[0][t][k], outs[1][t][k]), (t, k)
    if n >= 97:
        assert (outs[0][0]["nwsr"] > 80).any()                       # some cold solves were long enough to be parked (wbc_qp.cuh: PARK_MIN_ITER)


def test_warm_setup_kernel_stays_inside_its_workspace_prefix(wbc, wl):
    """wbc_warm_setup_kernel allocates only the first QPWORK_TQ_BYTES of the solver workspace.  With wbc_set_option("setup_canary", 1) it is launched with the whole
    workspace, the part past the prefix is filled with a NaN pattern and checked at the end: a read from there poisons the factorisation, a write
    makes the kernel poison it on purpose -- either way the outputs would differ from the normal run (compute-sanitizer is not available on the pool)."""
    n = 256
    st, ref = wl.stress(n)
    outs = []
    for canary in (0, 1):
        ctl = wbc.WholeBodyController(n); ctl.setOption("setup_canary", canary); s = st; rec = []
        for t in range(3):
            if t:
                s = wl.advance_state(s)
            ctl.setJointAndBaseState(**s); ctl.setReference(**ref); ctl.step(); rec.append(ctl.getOutputs())
        ctl.close(); outs.append(rec)
    for t in range(3):
        assert np.isfinite(outs[1][t]["x"]).all(), t
        for k in outs[0][t]:
            assert np.array_equal(outs[0][t][k], outs[1][t][k]), (t, k)


def test_sharding_does_not_change_results(wbc, wl):
    n = 96
    st, ref = wl.randomised(n, seed=1234, contacts="mixed")
    full = wbc.WholeBodyController(n); full.setJointAndBaseState(**st); full.setReference(**ref); full.step(); fo = full.getOutputs()
    for rank in range(2):
        first, count = wl.shard_range(n, rank, 2)
        s2, r2 = wl.randomised(count, seed=1234, first=first, contacts="mixed")
        c = wbc.WholeBodyController(count); c.setJointAndBaseState(**s2); c.setReference(**r2); c.step(); o = c.getOutputs()
        for k in ("tau", "grf", "x", "rc", "nwsr", "ws"):
            assert np.array_equal(o[k], fo[k][..., first:first + count]), k
        c.close()
    full.close()


def test_device_pointers_and_host_pointers_agree(wbc, wl):
    import torch
    st, ref = wl.randomised(64, seed=3, contacts="stance")
    a = wbc.WholeBodyController(64); a.setJointAndBaseState(**st); a.setReference(**ref); a.step(); oa = a.getOutputs()
    dst = {k: torch.from_numpy(v).cuda() for k, v in st.items()}; dref = {k: torch.from_numpy(v).cuda() for k, v in ref.items()}
    b = wbc.WholeBodyController(64, stream=torch.cuda.current_stream().cuda_stream)
    b.setJointAndBaseState(**dst); b.setReference(**dref); b.step()
    tau = torch.zeros(12, 64, dtype=torch.float64, device="cuda"); b.getOutputs(tau=tau); torch.cuda.synchronize()
    assert np.array_equal(tau.cpu().numpy(), oa["tau"])
    a.close(); b.close()


@pytest.mark.parametrize("contacts,n", [("stance", 4096), ("mixed", 16384)])
def test_full_size_properties(wbc, wl, contacts, n):
    """BASELINE config sizes: no oracle here, only properties the domain offers for every solved robot --
    friction pyramid, torque limits, primal feasibility and stationarity of sampled dense QPs, determinism."""
    st, ref = wl.randomised(n, seed=1234, contacts=contacts)
    ctl = wbc.WholeBodyController(n); ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); o = ctl.getOutputs()
    ok = o["rc"] == 0
    assert set(np.unique(o["rc"])) <= {0, 64} and ok.mean() > 0.9
    f = o["grf"].reshape(4, 3, n); c = ref["contacts"].astype(bool)
    mu = 1.0; tol = 1e-7 * np.maximum(1.0, np.abs(f).max(axis=(0, 1)))
    assert ((np.abs(f[:, 0]) <= mu * f[:, 2] + tol) | ~c | ~ok).all() and ((np.abs(f[:, 1]) <= mu * f[:, 2] + tol) | ~c | ~ok).all()
    assert (np.abs(o["tau"])[:, ok] <= 80.0 * (1 + 1e-9)).all()
    assert (o["ws"][:42][:, ok] == 0).all()                        # no far bound is active at a solution
    for i in np.nonzero(ok)[0][:: max(1, ok.sum() // 12)]:
        q = ctl.denseQP(int(i)); x = o["x"][:, i]; Ax = q["A"] @ x; sc = max(1.0, np.abs(Ax).max())
        assert (Ax >= q["lbA"] - 1e-8 * sc).all() and (Ax <= q["ubA"] + 1e-8 * sc).all()
        act = np.nonzero(o["ws"][42:, i])[0]; grad = q["H"] @ x + q["g"]
        lam = np.linalg.lstsq(q["A"][act].T, grad, rcond=None)[0]
        assert np.abs(q["A"][act].T @ lam - grad).max() < 1e-6 * max(1.0, np.abs(grad).max())
    ctl.setInitialState(); ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); o2 = ctl.getOutputs()
    assert all(np.array_equal(o[k], o2[k]) for k in o)             # deterministic whatever the work-queue order was
    ctl.close()


def test_device_qp_blocks_against_the_independent_urdf_model(wbc):
    """The device's centroidal mass matrix and foot Jacobians (as they appear in the QP's A, wbc.cpp:492-497) against tests/golden/rbd_independent.npz:
    an independent xml.etree read of the URDF + geometric-Jacobian mass matrix + numpy controller algebra (tests/urdf_independent.py) -- nothing shared
    with tools/urdf_compiler.py, from which both the device table and the oracle tree come."""
    import urdf_independent as ui
    gold = np.load(os.path.join(HERE, "golden", "rbd_independent.npz")); st, ref = ui.fixture_states(); n = st["q"].shape[1]
    ctl = wbc.WholeBodyController(n); ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); o = ctl.getOutputs()
    rel = lambda a, b: np.abs(a - b).max() / max(1.0, np.abs(b).max())
    for i in range(n):
        A = ctl.denseQP(i)["A"]; Mc = gold[f"Mc_{i}"]
        assert rel(A[0:6, 0:6], Mc[0:6, 0:6]) < 1e-9 and rel(A[34:46, 6:18], Mc[6:18, 6:18]) < 1e-9, i
        assert rel(A[6:18, 0:18], gold[f"Jst_c_{i}"]) < 1e-9 and rel(A[46:58, 0:18], gold[f"Jsw_c_{i}"]) < 1e-9, i
        assert rel(A[0:6, 18:30], -gold[f"Jst_c_{i}"][:, 0:6].T) < 1e-9, i
        assert rel(o["com"][:, i], gold[f"com_{i}"]) < 1e-9, i
    ctl.close()
