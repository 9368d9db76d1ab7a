"""CPU: the restated iDynTree subset (oracle/rbd_oracle.cpp) against the identities that iDynTree's own unit tests use
(lib/idyntree/src/high-level/tests/KinDynComputationsUnitTest.cpp:256-279 M*a + h == ID(a); :506-512 J*nu == frame velocity;
tolerance DEFAULT_TOL 1e-10) and against the URDF known answers."""
import json
import os

import numpy as np
import pytest

TOL = 1e-10       # iDynTree DEFAULT_TOL (lib/idyntree/src/core/src/Utils.cpp:15)


def _states(wl, n, seed=3):
    st, _ = wl.randomised(n, seed=seed)
    return st


def _eval(orc, st, i, lumped, acc=None):
    return orc.rbd_eval(st["base_pos"][:, i].copy(), st["base_quat"][:, i].copy(), st["base_twist"][:, i].copy(), st["q"][:, i].copy(),
                        st["qd"][:, i].copy(), lumped=lumped, acc18=acc)


@pytest.mark.parametrize("lumped", [False, True])
def test_mass_matrix_symmetric_positive_definite(orc, wl, lumped):
    st = _states(wl, 6)
    for i in range(6):
        M = _eval(orc, st, i, lumped)["M"]
        assert np.abs(M - M.T).max() < TOL
        assert np.linalg.eigvalsh(0.5 * (M + M.T)).min() > 1e-4
        assert abs(M[0, 0] - 51.63574) < 1e-9 and abs(M[1, 1] - 51.63574) < 1e-9      # total mass on the linear block (MIXED)


@pytest.mark.parametrize("lumped", [False, True])
def test_jacobian_times_nu_is_frame_velocity(orc, wl, lumped):
    st = _states(wl, 6, seed=5)
    for i in range(6):
        o = _eval(orc, st, i, lumped)
        nu = np.concatenate([st["base_twist"][:, i], st["qd"][:, i]])
        for leg in range(4):
            assert np.abs(o["J"][leg] @ nu - o["footVel"][leg]).max() < TOL


@pytest.mark.parametrize("lumped", [False, True])
def test_inverse_dynamics_identity(orc, wl, lumped):
    rng = np.random.default_rng(0)
    st = _states(wl, 6, seed=7)
    for i in range(6):
        acc = rng.uniform(-1, 1, 18)
        o = _eval(orc, st, i, lumped, acc=acc)
        assert np.abs(o["M"] @ acc + o["h"] - o["id"]).max() < 1e-9 * max(1.0, np.abs(o["id"]).max())


def test_gravity_forces_are_weight_at_rest(orc, wl):
    st, _ = wl.nominal(1)
    o = _eval(orc, st, 0, False)
    # gravity_ = (0,0,+9.81) in the reference (wbc.cpp:78): base force = -m*g_vector
    assert abs(o["g"][2] + 51.63574 * 9.81) < 1e-9 and np.abs(o["g"][:2]).max() < 1e-9
    assert np.abs(o["h"] - o["g"]).max() < 1e-10           # no velocity => bias == gravity


def test_lumped_13_body_model_equals_78_link_tree(orc, wl):
    """lumping fixed joints is value preserving: every hot-path quantity agrees to rounding"""
    st = _states(wl, 8, seed=11)
    for i in range(8):
        a, b = _eval(orc, st, i, False), _eval(orc, st, i, True)
        for k in ("M", "J", "h", "g", "com", "vcom", "footPos", "footVel"):
            assert np.abs(a[k] - b[k]).max() < 1e-11 * max(1.0, np.abs(a[k]).max()), k


def test_known_answers_of_the_urdf(orc, wl):
    here = os.path.dirname(os.path.abspath(__file__))
    model = json.load(open(os.path.join(here, "golden", "anymal_model.json")))
    assert len(model["full"]["links"]) == 78 and len(model["lumped"]["links"]) == 13 and len(model["fake_links"]) == 10
    assert abs(sum(l["m"] for l in model["full"]["links"]) - 51.63574) < 1e-9
    assert abs(sum(l["m"] for l in model["lumped"]["links"]) - 51.63574) < 1e-9
    assert [l["dof"] for l in model["lumped"]["links"]] == [-1] + list(range(12))
    # setInitialState configuration (wbc.cpp:81-82): base at z = 0.60 puts the feet at z = 0.015 and the CoM at (-0.008, 0.0003, 0.5227)
    st, _ = wl.nominal(1)
    st["base_pos"][2, 0] = 0.60
    o = _eval(orc, st, 0, False)
    assert np.abs(o["footPos"][2::3] - 0.015).max() < 1e-3
    assert np.abs(o["com"] - np.array([-0.008, 0.0003, 0.5227])).max() < 1e-3
