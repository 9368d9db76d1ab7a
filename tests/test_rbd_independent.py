"""CPU: the oracle's rigid-body half (oracle/rbd_oracle.cpp on the 78-link tree compiled by tools/urdf_compiler.py) against an INDEPENDENT
model -- its own xml.etree read of urdf/anymal.urdf and geometric-Jacobian formulas (tests/urdf_independent.py).  A convention error in the URDF
compiler (rpy order, inertia frame, axis direction), which the device table and the oracle tree would share, shows up here.
Needs the reference mount for the URDF; the GPU side of the same pin uses the committed fixture tests/golden/rbd_independent.npz."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import urdf_independent as ui          # noqa: E402

needs_urdf = pytest.mark.skipif(not os.path.exists(ui.URDF), reason="reference mount (URDF) not available")


def rel(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() / max(1.0, np.abs(b).max())


@needs_urdf
def test_mass_matrix_jacobians_com_gravity_and_bias_against_the_oracle(orc, wl):
    mdl = ui.Model(); st, _ = wl.randomised(12, seed=5150, contacts="stance")
    st["qd"] *= 3.0; st["base_twist"] *= 4.0                                 # larger velocities: the Coriolis part of h matters
    for i in range(12):
        a = [st[k][:, i] for k in ("base_pos", "base_quat", "base_twist", "q", "qd")]
        e = mdl.evaluate(*a); o = orc.rbd_eval(*a)
        assert rel(e["M"], o["M"]) < 1e-9 and np.abs(e["M"] - e["M"].T).max() < 1e-12
        assert rel(e["J"], o["J"]) < 1e-9
        assert rel(e["com"], o["com"]) < 1e-9 and rel(e["vcom"], o["vcom"]) < 1e-9
        assert rel(e["footPos"], o["footPos"]) < 1e-9 and rel(e["footVel"], o["footVel"]) < 1e-9
        assert rel(e["g"], o["g"]) < 1e-9
        assert rel(mdl.bias(*a), o["h"]) < 1e-6                               # dJ nu by central differences: truncation error, not rounding


@needs_urdf
def test_known_answers_follow_from_the_urdf_alone(orc, wl):
    """what tests/test_oracle_rbd.py held to 1e-3, tightened to what the URDF determines"""
    mdl = ui.Model()
    assert abs(mdl.total_mass() - 51.63574) < 1e-9
    st, _ = wl.nominal(1); st["base_pos"][2, 0] = 0.60
    a = [st[k][:, 0] for k in ("base_pos", "base_quat", "base_twist", "q", "qd")]
    e = mdl.evaluate(*a); o = orc.rbd_eval(*a)
    assert np.abs(e["footPos"] - o["footPos"]).max() < 1e-12 and np.abs(e["com"] - o["com"]).max() < 1e-12
    assert np.abs(e["footPos"][2::3] - e["footPos"][2]).max() < 1e-12         # the four feet of the symmetric nominal stance are level
    assert abs(e["mass"] - o["M"][0, 0]) < 1e-9                               # the linear block of the MIXED mass matrix is m I


@needs_urdf
def test_committed_fixture_is_what_the_independent_model_gives(wl):
    """tests/golden/rbd_independent.npz (used by the GPU test, which has no URDF) is reproducible from the script that made it"""
    gold = np.load(os.path.join(HERE, "golden", "rbd_independent.npz")); mdl = ui.Model(); st, ref = ui.fixture_states()
    for i in range(st["q"].shape[1]):
        e = mdl.evaluate(st["base_pos"][:, i], st["base_quat"][:, i], st["base_twist"][:, i], st["q"][:, i], st["qd"][:, i])
        cb = ui.centroidal_blocks(e, ref["contacts"][:, i])
        assert np.array_equal(e["M"], gold[f"M_{i}"]) and np.array_equal(cb["Jst_c"], gold[f"Jst_c_{i}"])


def test_oracle_centroidal_blocks_match_the_fixture(orc):
    """independent M, J -> T, Mc, Jst_c (wbc.cpp:256-320 restated in numpy) against the oracle's controller algebra (dense 18x18 inverses)"""
    gold = np.load(os.path.join(HERE, "golden", "rbd_independent.npz")); st, ref = ui.fixture_states(); n = st["q"].shape[1]
    ob = orc.OracleBatch(n); ob.set_state(**st); ob.set_reference(**ref); ob.step(assemble_only=True)
    for i in range(n):
        m = ob.intermediates(i)
        assert rel(m["M"], gold[f"M_{i}"]) < 1e-9 and rel(m["Mc"], gold[f"Mc_{i}"]) < 1e-9
        assert rel(m["Jst"], gold[f"Jst_c_{i}"]) < 1e-9 and rel(m["Jsw"], gold[f"Jsw_c_{i}"]) < 1e-9
        assert rel(m["com"], gold[f"com_{i}"]) < 1e-9
