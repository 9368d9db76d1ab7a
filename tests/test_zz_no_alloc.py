"""GPU (-m gpu), collected LAST on purpose: wbc_solve_qp_static, the no-allocation variant of solveQP (no counterpart in the reference).
Its warm-start scratch lives in a file-scope static __device__ array of the library instead of a cudaMalloc'd buffer,
so the same kernels see the same data and the outputs must be those of wbc_solve_qp -- compared bit for bit AND within the 1e-9 relative tolerance
the parity tests use (parity_util.RTOL), on a cold tick followed by two hot-started ticks in a row (the two calls that use the scratch), and once
more with a second handle, which finds the previous call's contents in the shared array."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from parity_util import RTOL, relerr          # noqa: E402

pytestmark = pytest.mark.gpu
KEYS = ("tau", "grf", "x", "rc", "nwsr", "ws")


def _ticks(wbc, wl, n, seed, static, ticks=3):
    st, ref = wl.randomised(n, seed=seed, contacts="mixed", switching_frac=0.1)
    ctl = wbc.WholeBodyController(n)
    outs = []
    for t in range(ticks):
        if t:
            st = wl.advance_state(st)
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref)
        ctl.updateState()
        (ctl.solveQPStatic if static else ctl.solveQP)()
        ctl.computeJointTorques(); ctl.synchronize()
        outs.append({k: v.copy() for k, v in ctl.getOutputs().items() if k in KEYS})
    ctl.close()
    return outs


def _same(a, b, what):
    for t, (oa, ob) in enumerate(zip(a, b)):
        assert (oa["rc"] == ob["rc"]).all() and (oa["nwsr"] == ob["nwsr"]).all() and (oa["ws"] == ob["ws"]).all(), f"{what}: tick {t} rc / nWSR / working set differ"
        for k in ("tau", "grf", "x"):
            e = float(relerr(oa[k], ob[k]).max())
            assert e < RTOL, f"{what}: tick {t} {k} rel err {e:.3e}"
            assert np.array_equal(oa[k], ob[k]), f"{what}: tick {t} {k} not bit-identical"


def test_solve_qp_static_equals_solve_qp_on_consecutive_calls(wbc, wl):
    n = 300
    want = _ticks(wbc, wl, n, 4242, static=False)
    got = _ticks(wbc, wl, n, 4242, static=True)
    assert (want[0]["rc"] == 0).mean() > 0.9 and want[2]["nwsr"].mean() < want[0]["nwsr"].mean()      # tick 0 cold, ticks 1-2 hot-started
    _same(got, want, "first handle")
    # a second handle and other data: the shared array still holds the previous calls' records
    want2 = _ticks(wbc, wl, 137, 99, static=False)
    got2 = _ticks(wbc, wl, 137, 99, static=True)
    _same(got2, want2, "second handle")


def test_solve_qp_static_refuses_batches_beyond_its_scratch(wbc, wl):
    n = 1025                                                   # WBC_STATIC_MAX_ROBOTS = 1024 (include/wbc_b200.h)
    st, ref = wl.randomised(n, seed=1, contacts="stance")
    ctl = wbc.WholeBodyController(n)
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.updateState()
    with pytest.raises(RuntimeError, match="exceed the static scratch of 1024"):
        ctl.solveQPStatic()
    ctl.solveQP(); ctl.synchronize()                           # the handle is still usable through the allocating entry point
    assert (ctl.getOutputs()["rc"] == 0).mean() > 0.9
    ctl.close()
