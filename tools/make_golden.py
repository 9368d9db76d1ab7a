#!/usr/bin/env python3
"""Generates tests/golden/wbc_golden.npz with the oracle (restated iDynTree + controller algebra + REAL vendored qpOASES),
run in the build container where /root/reference is mounted.  The vectors travel to the GPU box; the reference does not.

    python tools/make_golden.py

Cases (BASELINE.json configs made concrete in SURVEY.md section 8d, at sizes the CPU finishes in seconds):
  c1  config 1: single nominal stance robot, tick 0 cold (SQProblem::init) + 4 hot-start ticks on the same state
  c2  config 2: 24 randomised 4-contact states, cold
  c3  config 3: 24 diagonal-pair states with swing targets, tick 0 cold + tick 1 (state advanced by dt) hot start
  c4  config 4: 24 mixed-contact states, `switching` set on 30 %, cold + one hot start
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc                                   # noqa: E402
from quadrupedwholebodycontroller_b200 import workloads as wl      # noqa: E402

CASES = dict(c1=dict(kind="nominal", N=1, ticks=5, advance=False), c2=dict(kind="stance", N=24, ticks=1, advance=False, seed=1234),
             c3=dict(kind="diagonal", N=24, ticks=2, advance=True, seed=1234, swing=True),
             c4=dict(kind="mixed", N=24, ticks=2, advance=True, seed=4321, switching_frac=0.3))


def case_inputs(spec, tick, st=None, ref=None):
    """deterministic inputs of a case at a tick (shared by the generator and by the tests)"""
    if tick == 0:
        if spec["kind"] == "nominal":
            st, ref = wl.nominal(spec["N"])
        else:
            st, ref = wl.randomised(spec["N"], seed=spec["seed"], contacts=spec["kind"], switching_frac=spec.get("switching_frac", 0.0))
    elif spec["advance"]:
        st = wl.advance_state(st)
    return st, ref


def main():
    out = {}
    for name, spec in CASES.items():
        ob = orc.OracleBatch(spec["N"])
        st = ref = None
        for t in range(spec["ticks"]):
            st, ref = case_inputs(spec, t, st, ref)
            if spec.get("swing") and t == 0:
                probe = orc.OracleBatch(spec["N"]); probe.set_state(**st); probe.set_reference(**ref); probe.step(assemble_only=True)
                foot = np.stack([probe.intermediates(i)["footPos"] for i in range(spec["N"])], axis=1)
                ref = wl.with_swing_targets(ref, foot)
                out[f"{name}_swpos"] = ref["sw_pos"]
            ob.set_state(**st); ob.set_reference(**ref); ob.step()
            o = ob.outputs()
            for k in ("tau", "grf", "x", "rc", "nwsr", "ws"):
                out[f"{name}_t{t}_{k}"] = o[k]
            q = ob.qp(0)
            for k in ("H", "g", "A", "lbA", "ubA"):
                out[f"{name}_t{t}_qp0_{k}"] = q[k]
            print(name, "tick", t, "rc", np.bincount(o["rc"], minlength=1)[:1], "nonzero rc", np.count_nonzero(o["rc"]), "nwsr mean", o["nwsr"].mean())
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "wbc_golden.npz"), **out)


if __name__ == "__main__":
    main()
