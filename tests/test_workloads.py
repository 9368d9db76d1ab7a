"""CPU: the synthetic workload generator (BASELINE.json configs) is counter based: sharding never changes a robot's data."""
import numpy as np


def test_shards_reproduce_the_full_batch(wl):
    st, ref = wl.randomised(64, seed=1234, contacts="mixed", switching_frac=0.1)
    for first, count in ((0, 16), (16, 16), (40, 24)):
        s2, r2 = wl.randomised(count, seed=1234, first=first, contacts="mixed", switching_frac=0.1)
        for k in st:
            assert np.array_equal(st[k][:, first:first + count], s2[k]), k
        assert np.array_equal(ref["contacts"][:, first:first + count], r2["contacts"])
        assert np.array_equal(ref["switching"][first:first + count], r2["switching"])


def test_contact_patterns(wl):
    _, ref = wl.randomised(4000, seed=1, contacts="mixed")
    n = ref["contacts"].sum(axis=0)
    assert 0.45 < (n == 4).mean() < 0.55 and 0.35 < (n == 3).mean() < 0.45 and 0.07 < (n == 2).mean() < 0.13
    _, ref = wl.randomised(256, seed=1, contacts="diagonal")
    c = ref["contacts"]
    assert ((c[0] == c[2]) & (c[1] == c[3]) & (c[0] != c[1])).all()


def test_layout_is_structure_of_arrays(wl):
    st, ref = wl.randomised(8)
    assert st["q"].shape == (12, 8) and st["base_quat"].shape == (4, 8) and ref["contacts"].shape == (4, 8) and ref["contacts"].dtype == np.uint8
    assert all(v.flags["C_CONTIGUOUS"] for v in list(st.values()) + list(ref.values()))
    assert np.allclose(np.linalg.norm(st["base_quat"], axis=0), 1.0)
    s2 = wl.advance_state(st)
    assert np.allclose(s2["q"], st["q"] + st["qd"] / 400.0) and np.allclose(np.linalg.norm(s2["base_quat"], axis=0), 1.0)


def test_rollout_start_is_sharding_invariant(wl):
    """config 5 inputs are keyed by the global robot index: any contiguous shard reproduces the same robots"""
    st, ref, cmd = wl.rollout_start(64)
    for rank in range(3):
        first, count = wl.shard_range(64, rank, 3)
        s2, r2, c2 = wl.rollout_start(count, first=first)
        for k in st:
            assert np.array_equal(s2[k], st[k][:, first:first + count]), k
        assert np.array_equal(c2, cmd[:, first:first + count])
    assert np.allclose(np.linalg.norm(st["base_quat"], axis=0), 1.0)
    assert (cmd[:, 15] == 0).all() and (cmd[0, :15] >= 0.1).all() and np.abs(cmd[2]).max() <= 0.3 + 1e-12


def test_oracle_closed_loop_is_stable(orc, wl):
    """the oracle's closed loop (restated planner + controller + real qpOASES + dense plant) walks for one second without a failed solve"""
    import os
    N = 4
    st, ref, cmd = wl.rollout_start(N)
    ob = orc.OracleBatch(N); ob.set_state(**st); ob.set_reference(**ref)
    fd = os.open(os.devnull, os.O_WRONLY); saved = os.dup(2); os.dup2(fd, 2)
    try:
        r = ob.rollout(400, cmd)
    finally:
        os.dup2(saved, 2); os.close(fd); os.close(saved)
    assert (r["rc"] == 0).all()
    assert r["nwsr"][0].min() > 40 and np.median(r["nwsr"][1:]) <= 2
    z = r["pos"][:, 2, :]
    assert (z > 0.4).all() and (z < 0.7).all()
    assert np.abs(r["tau"]).max() <= 80.0 + 1e-9


def test_stress_workload_kinds_and_sharding(wl):
    st, ref = wl.stress(64)
    c = ref["contacts"]; kind = np.arange(64) % 8
    assert (c[:, kind == 0] == 0).all() and (c[:, kind == 1].sum(axis=0) == 1).all() and (c[:, kind == 5].sum(axis=0) == 3).all()
    assert (c[:, kind == 6].sum(axis=0) == 2).all() and ((c[0] == c[1]) & (c[2] == c[3]))[kind == 6].all()
    assert np.abs(ref["acc"][:, kind == 2]).max() > 10 and not ref["acc"][:, kind != 2].any()
    assert np.abs(st["qd"][:, kind == 3]).max() > 3 and np.abs(st["qd"][:, kind != 3]).max() <= 0.5
    assert np.allclose(np.linalg.norm(st["base_quat"], axis=0), 1.0)
    assert (st["base_pos"][2, kind == 7] == 0.25).all()
    s2, r2 = wl.stress(24, first=40)
    for k in st:
        assert np.array_equal(s2[k], st[k][:, 40:64]), k
    for k in ref:
        assert np.array_equal(r2[k], ref[k][..., 40:64]), k
