"""GPU (-m gpu): the batched MotionPlanner, the hand-over of its messages to the controller, the plant and the closed-loop rollout,
called through the C ABI, against the committed golden vectors and the oracle (restated src/motion_planner.cpp; the plant has no
reference counterpart and is defined densely in oracle/oracle_capi.cpp).

Tolerances: planner outputs bit-exact when the yaw rate is zero (no FMA contraction on the device: __dmul_rn/__dadd_rn), <= 1e-12
relative when sin/cos of a non-zero yaw enter; discrete outputs (published, contacts, switching) always exact.  Closed loop: return
codes identical, torques within 1e-9 relative per tick as long as both sides took the same number of working-set changes, and within
1e-6 over the whole trajectory (a closed loop feeds rounding differences back through the state).
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "tools"))
import make_golden_planner as mgp          # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(HERE, "golden", "planner_golden.npz"))
MSG = ("pose", "vel", "acc", "sw_pos", "sw_vel", "sw_acc")


def run_device_planner(wbc, cmd_per_robot):
    """cmd_per_robot: (N, K, 3).  Returns per-period arrays (K, ..., N) of what the controller holds after each planner period."""
    N, K, _ = cmd_per_robot.shape
    ctl = wbc.WholeBodyController(N); ctl.plannerCreate()
    rec = {k: [] for k in MSG + ("contacts", "switching", "published")}
    for p in range(K):
        c = np.ascontiguousarray(cmd_per_robot[:, p, :].T)
        ctl.plannerSetCmdVel(c[0], c[1], c[2]); ctl.plannerTick()
        r = ctl.getReference()
        for k in rec:
            rec[k].append(r[k])
    ctl.close()
    return {k: np.stack(v) for k, v in rec.items()}


def test_planner_matches_golden_profiles(wbc):
    names = ["straight", "turning", "startstop"]
    cmd = np.stack([GOLD[f"p_{n}_cmd"] for n in names])
    dev = run_device_planner(wbc, cmd)
    for i, n in enumerate(names):
        pub = GOLD[f"p_{n}_published"].astype(bool)
        assert np.array_equal(dev["published"][:, i].astype(bool), pub), n
        # the controller holds the LAST PUBLISHED message (referenceCallback); the oracle records the latched ref_msg_ of every period,
        # which only changes in periods that publish -> equal in every period from the first publication on
        for k in MSG:
            want = GOLD[f"p_{n}_{k}"]; got = dev[k][:, :, i]
            if n == "turning":
                assert np.abs(got - want).max() <= 1e-12 * max(1.0, np.abs(want).max()), (n, k)
            else:
                assert np.array_equal(got, want), (n, k)
        assert np.array_equal(dev["contacts"][:, :, i], GOLD[f"p_{n}_contacts"]), n
        # isSwitchingFootState_ (whole_body_controller.cpp:176-184): set by the first message with new contacts, cleared by the next
        # message, latched over silent periods
        con = GOLD[f"p_{n}_contacts"]; sw = np.zeros(len(con), bool); prev = np.ones(4, np.uint8); cur = False
        for p in range(len(con)):
            if pub[p]:
                cur = bool((con[p] != prev).any()); prev = con[p]
            sw[p] = cur
        assert np.array_equal(dev["switching"][:, i].astype(bool), sw), n


def test_planner_matches_oracle_on_random_commands(wbc, orc):
    rng = np.random.default_rng(5)
    N, K = 64, 120
    cmd = np.zeros((N, K, 3)); cmd[:, :, 0] = rng.uniform(-0.4, 0.4, (N, 1)); cmd[:, :, 1] = rng.uniform(-0.2, 0.2, (N, 1))
    cmd[N // 2:, :, 2] = rng.uniform(-0.5, 0.5, (N // 2, 1))              # second half turns
    cmd[::8, :30, :] = 0.0                                                # some robots stand for a while first
    dev = run_device_planner(wbc, cmd)
    for i in range(N):
        r = orc.planner_run(cmd[i])
        assert np.array_equal(dev["published"][:, i], r["published"]), i
        assert np.array_equal(dev["contacts"][:, :, i], r["contacts"]), i
        for k in MSG:
            if i < N // 2:
                assert np.array_equal(dev[k][:, :, i], r[k]), (i, k)
            else:
                assert np.abs(dev[k][:, :, i] - r[k]).max() <= 1e-12 * max(1.0, np.abs(r[k]).max()), (i, k)


def rel(a, b):
    return np.abs(a - b).max() / max(1.0, np.abs(b).max())


def test_closed_loop_rollout_matches_golden(wbc):
    st, ref, cmd = mgp.rollout_inputs()
    N, K = mgp.ROLLOUT_N, mgp.ROLLOUT_TICKS
    ctl = wbc.WholeBodyController(N); ctl.plannerCreate(); ctl.plantEnable()
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
    tight, worst_all = 0, 0.0
    for t in range(K):
        ctl.rollout(1)
        o = ctl.getOutputs(); s = ctl.getState()
        assert np.array_equal(o["rc"], GOLD["r_rc"][t]), t
        for i in range(N):
            e = rel(o["tau"][:, i], GOLD["r_tau"][t][:, i]); worst_all = max(worst_all, e); tight += int(e < 1e-9)
        assert rel(s["q"], GOLD["r_q"][t]) < 1e-8 and rel(s["base_pos"], GOLD["r_pos"][t]) < 1e-8, t
    # every (tick, robot) pair within 1e-6, and the share within 1e-9 is asserted (a last-bit difference at a degenerate vertex is fed back through the plant)
    assert worst_all < 1e-6, worst_all
    assert tight >= 0.95 * N * K, (tight, N * K)
    fin = ctl.getState()
    for k in ("q", "qd", "base_pos", "base_quat", "base_twist"):
        assert rel(fin[k], GOLD[f"r_final_{k}"]) < 1e-8, k
    stats = ctl.rolloutStats()
    assert (stats["ticks"] == K).all() and (stats["failed_ticks"] == 0).all()
    assert (stats["nwsr_max"] >= 40).all()                      # the cold start of tick 0
    ctl.close()


def test_plant_step_matches_oracle_from_random_states(wbc, orc, wl):
    """one controller tick + one plant step from 64 randomised states with large velocities: closed-form T^-1 (device) vs dense 18x18
    inverse (oracle)"""
    N = 64
    st, ref = wl.randomised(N, seed=99, contacts="mixed")
    ctl = wbc.WholeBodyController(N); ctl.plannerCreate(); ctl.plantEnable()
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref)
    # two ticks: the second one has a non-trivial Tdot (first tick: old T = identity)
    for t in range(2):
        ctl.step(); ctl.plantStep()
    got = ctl.getState()
    ctl.close()
    # compare against the dense definition evaluated here in numpy from the oracle's intermediates
    ob3 = orc.OracleBatch(N); ob3.set_state(**st); ob3.set_reference(**ref)
    cur = {k: v.copy() for k, v in st.items()}
    dt = 1.0 / 400.0
    for t in range(2):
        ob3.set_state(**cur); ob3.step(); w = ob3.outputs()
        Told = [ob3.intermediates(i)["T"] for i in range(N)]
        if t == 0:
            Tprev = [np.eye(18)] * N
        for i in range(N):
            T = Told[i]; Tdot = (T - Tprev[i]) / dt if not ref["switching"][i] else np.zeros((18, 18))
            nu = np.concatenate([cur["base_twist"][:, i], cur["qd"][:, i]])
            nud = np.linalg.solve(T, w["x"][:18, i] - Tdot @ nu) if w["rc"][i] == 0 else np.zeros(18)
            cur["base_twist"][:, i] += nud[:6] * dt; cur["qd"][:, i] += nud[6:] * dt
            cur["q"][:, i] += cur["qd"][:, i] * dt; cur["base_pos"][:, i] += cur["base_twist"][:3, i] * dt
            rv = cur["base_twist"][3:, i] * dt; a = np.linalg.norm(rv)
            dq = np.concatenate([[np.cos(a / 2)], np.sin(a / 2) / a * rv]) if a > 1e-12 else np.concatenate([[1.0], 0.5 * rv])
            qw, qx, qy, qz = cur["base_quat"][:, i]; dw, dx, dy, dz = dq
            nq = np.array([dw * qw - dx * qx - dy * qy - dz * qz, dw * qx + dx * qw + dy * qz - dz * qy, dw * qy - dx * qz + dy * qw + dz * qx, dw * qz + dx * qy - dy * qx + dz * qw])
            cur["base_quat"][:, i] = nq / np.linalg.norm(nq)
        Tprev = Told
    for k in ("q", "qd", "base_pos", "base_quat", "base_twist"):
        assert rel(got[k], cur[k]) < 1e-9, k


def test_full_size_rollout_properties(wbc, wl):
    """BASELINE configs[4] size (16384 robots), 120 control ticks incl. the first contact switches: no oracle at this size, only what the
    domain guarantees -- every QP solves, torque limits and friction pyramids hold for stance legs, the base stays up, hot starts stay cheap,
    and the run is deterministic whatever order the work queue handed the robots out in."""
    n, T = 16384, 120
    st, ref, cmd = wl.rollout_start(n)

    def run():
        ctl = wbc.WholeBodyController(n); ctl.plannerCreate(); ctl.plantEnable()
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
        ctl.rollout(T)
        o, s, r, stats = ctl.getOutputs(), ctl.getState(), ctl.getReference(), ctl.rolloutStats()
        ctl.close()
        return o, s, r, stats
    o, s, r, stats = run()
    assert (stats["ticks"] == T).all() and stats["failed_ticks"].sum() == 0
    assert (np.abs(o["tau"]) <= 80.0 * (1 + 1e-9)).all()
    f = o["grf"].reshape(4, 3, n); c = r["contacts"].astype(bool)
    tol = 1e-7 * np.maximum(1.0, np.abs(f).max(axis=(0, 1)))
    assert ((np.abs(f[:, 0]) <= f[:, 2] + tol) | ~c).all() and ((np.abs(f[:, 1]) <= f[:, 2] + tol) | ~c).all()
    assert (np.abs(f.transpose(0, 2, 1)[~c]) <= 1e-6).all()               # swing legs carry no force (their H block + zero A columns)
    assert (s["base_pos"][2] > 0.4).all() and (s["base_pos"][2] < 0.7).all()
    assert np.allclose(np.linalg.norm(s["base_quat"], axis=0), 1.0, atol=1e-12)
    assert stats["nwsr_sum"].sum() / stats["ticks"].sum() < 6.0 and (stats["nwsr_max"] >= 40).all()
    walking = np.abs(cmd).sum(axis=0) > 0
    assert ((r["contacts"].sum(axis=0) == 3) == walking).all()            # tick 120 = planner period 30: second step phase of the crawl
    o2, s2, _, _ = run()
    assert all(np.array_equal(o[k], o2[k]) for k in o) and all(np.array_equal(s[k], s2[k]) for k in s)


def test_plant_invariants_stance_feet_and_power_balance(wbc, orc, wl):
    """what a rigid-body plant under (tau*, f*) must conserve.  32 standing robots (all four feet in stance, zero initial velocities, the
    controller pulls the body from its start posture to the reference pose), 0.5 s of closed loop on the device, at the reference's 400 Hz and at 1600 Hz:
      * the stance feet (almost) do not move: the QP's rows 6..17 demand zero foot acceleration and the plant integrates exactly the accelerations
        the QP returns.  "Almost": the reference forms the right-hand side -Jdot_c nu_c with nu_c = [v_com; omega_BASE; qdot] (wbc.cpp:504,511), whereas
        the centroidal velocity that belongs to its T (wbc.cpp:296-320) has the AVERAGE angular velocity in that slot; while the legs swing the body
        the two differ, and the feet creep (measured: 0.24 mm in 0.5 s at joint rates up to 0.7 rad/s).  That creep is a property of the reference's
        formulation, not of the integrator: it is the same at 400 Hz and at 1600 Hz (asserted), and it is small (asserted: < 1 mm in 0.5 s);
      * power balance: d/dt (kinetic + potential energy) = tau . qdot -- the contact forces do no work on feet that stand still -- again to first
        order in dt.
    Foot positions, the mass matrix and the CoM come from the oracle's rigid-body model evaluated on the DEVICE's states (checker only)."""
    n, seconds = 32, 0.5
    st, ref = wl.nominal(n)
    rs, _ = wl.randomised(n, seed=8, contacts="stance")
    st["q"] = st["q"] + 0.3 * (rs["q"] - st["q"]); st["base_pos"] = st["base_pos"] + 0.3 * (rs["base_pos"] - st["base_pos"])       # different postures, still at rest

    def model(s, i):
        return orc.rbd_eval(s["base_pos"][:, i], s["base_quat"][:, i], s["base_twist"][:, i], s["q"][:, i], s["qd"][:, i])

    def energy(s, i, o):
        nu = np.concatenate([s["base_twist"][:, i], s["qd"][:, i]])
        return 0.5 * nu @ o["M"] @ nu + o["M"][0, 0] * 9.81 * o["com"][2]

    def run(rate):
        prm = wbc.default_params(); prm.loopRate = rate
        pp = wbc.WbcPlannerParams(); wbc.load_library().wbc_planner_default_params(pp); pp.dt = 4.0 / rate      # one planner period every 4 control ticks, as in the reference
        ctl = wbc.WholeBodyController(n, params=prm); ctl.plannerCreate(pp); ctl.plantEnable()
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref)                   # /cmd_vel = 0: the planner keeps publishing "stand"
        T, dt = int(round(seconds * rate)), 1.0 / rate
        s = ctl.getState(); foot0 = np.stack([model(s, i)["footPos"] for i in range(n)]); E = np.array([energy(s, i, model(s, i)) for i in range(n)])
        drift, perr, pscale = 0.0, 0.0, 0.0
        for t in range(T):
            ctl.rollout(1); o = ctl.getOutputs(); s2 = ctl.getState()
            assert (o["rc"] == 0).all(), t
            if t % 4 == 3 or t == T - 1:                                          # the checker is the slow part: every 4th tick
                for i in range(n):
                    drift = max(drift, float(np.abs(model(s2, i)["footPos"] - foot0[i]).max()))
            for i in range(0, n, 4):
                m2 = model(s2, i); E2 = energy(s2, i, m2)
                # semi-implicit Euler: velocities first, so the work of this tick is done with the NEW joint rates
                P = float(o["tau"][:, i] @ s2["qd"][:, i]); r = (E2 - E[i]) / dt
                perr = max(perr, abs(r - P)); pscale = max(pscale, abs(P), abs(r)); E[i] = E2
        ctl.close()
        return drift, perr, pscale
    d400, p400, s400 = run(400.0); d1600, p1600, s1600 = run(1600.0)
    print(f"plant invariants over {seconds} s: stance-foot drift {d400:.3e} m at 400 Hz, {d1600:.3e} m at 1600 Hz; "
          f"power-balance error {p400:.3e} W (scale {s400:.3e}) at 400 Hz, {p1600:.3e} W (scale {s1600:.3e}) at 1600 Hz")
    assert d400 < 1e-3 and abs(d1600 - d400) < 0.1 * d400, (d400, d1600)
    assert p400 < 0.25 * s400 and p1600 < 0.6 * p400, (p400, s400, p1600, s1600)             # the energy rate matches tau . qdot to first order in dt


def test_rollout_graph_replay_equals_plain_launches(wbc, wl):
    """wbc_rollout replays a captured planner period (CUDA graph) once every robot is hot-started: outputs, states, per-robot statistics and
    the launch count must be those of the plain per-tick launches -- incl. a tick count that is not a multiple of the period and a call that
    starts in the middle of a period."""
    n = 600
    st, ref, cmd = wl.rollout_start(n)

    def run(graph):
        ctl = wbc.WholeBodyController(n); ctl.plannerCreate(); ctl.plantEnable(); ctl.setOption("rollout_graph", graph)
        ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
        ctl.rollout(3); ctl.rollout(58); ctl.rollout(40)          # 101 ticks: cold tick, part periods, 24 whole periods
        out = ctl.getOutputs(), ctl.getState(), ctl.rolloutStats(), ctl.launchCount()
        ctl.close()
        return out
    o0, s0, t0, l0 = run(0)
    o1, s1, t1, l1 = run(1)
    assert all(np.array_equal(o0[k], o1[k]) for k in o0) and all(np.array_equal(s0[k], s1[k]) for k in s0)
    assert all(np.array_equal(t0[k], t1[k]) for k in t0) and (t1["ticks"] == 101).all()
    assert l0 == l1
