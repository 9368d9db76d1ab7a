"""Synthetic workloads of BASELINE.json `configs` made concrete.

Every array is structure-of-arrays with the instance index fastest: shape (K, N), C-contiguous float64 / uint8,
exactly what the C-ABI (include/wbc_b200.h) takes.  The generator is counter based (numpy Philox, fixed number of
draws per instance, `advance` to the shard's first instance) so instance i gets the same numbers whatever the
sharding over GPUs is.
"""
import numpy as np

NOMINAL_Q = np.array([0.0, -0.4, 0.8, 0.0, 0.4, -0.8, 0.0, 0.4, -0.8, 0.0, -0.4, 0.8])   # wbc.cpp:81 (LH,LF,RF,RH)
NOMINAL_BASE_Z = 0.585          # feet at z ~ 0 in the nominal configuration (foot z = 0.015 at base z = 0.60)
REF_POSE = np.array([0.0, 0.0, 0.5, 0.0, 0.0, 0.0])                                       # params_controller.yaml:11
_DRAWS = 48                     # uniform draws reserved per instance (multiple of 4 = Philox block)

# contact patterns in leg order LH, LF, RF, RH (wbc.cpp:327-341)
STANCE = (1, 1, 1, 1)
SINGLE_SWING = ((0, 1, 1, 1), (1, 1, 1, 0), (1, 0, 1, 1), (1, 1, 0, 1))     # crawl order LH, RH, LF, RF (motion_planner.cpp:248-300)
DIAGONAL = ((1, 0, 1, 0), (0, 1, 0, 1))                                     # LH+RF / LF+RH


def rpy_to_quat_wxyz(r, p, y):
    cr, sr, cp, sp, cy, sy = np.cos(r / 2), np.sin(r / 2), np.cos(p / 2), np.sin(p / 2), np.cos(y / 2), np.sin(y / 2)
    return np.stack([cr * cp * cy + sr * sp * sy, sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy])


def _uniform(seed, first, count):
    bg = np.random.Philox(key=seed)
    bg.advance(first * (_DRAWS // 4))
    return np.random.Generator(bg).random((count, _DRAWS))


def nominal(N=1):
    """config 1: nominal stance, zero velocities, references = initialReferencePose."""
    z = np.zeros
    st = dict(q=np.repeat(NOMINAL_Q[:, None], N, 1), qd=z((12, N)), base_pos=np.repeat(np.array([[0.0], [0.0], [NOMINAL_BASE_Z]]), N, 1),
              base_quat=np.repeat(np.array([[1.0], [0], [0], [0]]), N, 1), base_twist=z((6, N)))
    ref = dict(pose=np.repeat(REF_POSE[:, None], N, 1), vel=z((6, N)), acc=z((6, N)), sw_pos=z((12, N)), sw_vel=z((12, N)),
               sw_acc=z((12, N)), contacts=np.ones((4, N), np.uint8), switching=np.zeros(N, np.uint8))
    return {k: np.ascontiguousarray(v) for k, v in st.items()}, {k: np.ascontiguousarray(v) for k, v in ref.items()}


def randomised(N, seed=1234, first=0, contacts="stance", switching_frac=0.0):
    """configs 2/3/4: randomised states around the nominal stance.

    q = nominal + U(-0.15,0.15); qd ~ U(-0.5,0.5); base rpy ~ U(-0.1,0.1); base xyz = (0,0,h) + U(-0.03,0.03);
    base twist ~ U(-0.2,0.2).  contacts: "stance" (1111), "diagonal" (1010 / 0101, p = 1/2) or "mixed"
    (50 % stance, 40 % single swing, 10 % diagonal pairs).  Swing references: p_d = (filled by the caller from the foot
    positions if wanted) left at zero here except "diagonal"/"mixed" where sw_pos is set by `with_swing_targets`.
    """
    u = _uniform(seed, first, N)
    s = lambda c0, c1, lo, hi: (lo + (hi - lo) * u[:, c0:c1]).T
    q = NOMINAL_Q[:, None] + s(0, 12, -0.15, 0.15)
    qd = s(12, 24, -0.5, 0.5)
    rpy = s(24, 27, -0.1, 0.1)
    pos = np.array([[0.0], [0.0], [NOMINAL_BASE_Z]]) + s(27, 30, -0.03, 0.03)
    twist = s(30, 36, -0.2, 0.2)
    quat = rpy_to_quat_wxyz(rpy[0], rpy[1], rpy[2])
    sel = u[:, 36]
    sub = u[:, 37]
    c = np.ones((4, N), np.uint8)
    if contacts == "diagonal":
        pat = np.array(DIAGONAL, np.uint8)[(sel >= 0.5).astype(int)]
        c = pat.T.copy()
    elif contacts == "mixed":
        single = np.array(SINGLE_SWING, np.uint8)[np.minimum((sub * 4).astype(int), 3)]
        diag = np.array(DIAGONAL, np.uint8)[(sub >= 0.5).astype(int)]
        pat = np.where((sel < 0.5)[:, None], np.ones((N, 4), np.uint8), np.where((sel < 0.9)[:, None], single, diag))
        c = pat.T.copy()
    elif contacts != "stance":
        raise ValueError(contacts)
    switching = (u[:, 38] < switching_frac).astype(np.uint8)
    z = np.zeros
    st = dict(q=q, qd=qd, base_pos=pos, base_quat=quat, base_twist=twist)
    ref = dict(pose=np.repeat(REF_POSE[:, None], N, 1), vel=z((6, N)), acc=z((6, N)), sw_pos=z((12, N)), sw_vel=z((12, N)),
               sw_acc=z((12, N)), contacts=c, switching=switching)
    return {k: np.ascontiguousarray(v, dtype=np.float64) for k, v in st.items()}, \
           {k: np.ascontiguousarray(v) for k, v in ref.items()}


def with_swing_targets(ref, foot_pos, lift=0.05):
    """config 3: swing reference p_d = current foot position + (0,0,lift), v_d = a_d = 0."""
    ref = dict(ref)
    sw = np.array(foot_pos, dtype=np.float64, copy=True)
    sw[2::3, :] += lift
    ref["sw_pos"] = np.ascontiguousarray(sw)
    return ref


def advance_state(state, dt=1.0 / 400.0):
    """config 3 tick 1: explicit Euler step of the state with its own velocities (world-frame base twist)."""
    s = {k: v.copy() for k, v in state.items()}
    s["q"] = state["q"] + dt * state["qd"]
    s["base_pos"] = state["base_pos"] + dt * state["base_twist"][0:3]
    w = state["base_twist"][3:6]
    qw, qx, qy, qz = state["base_quat"]
    # q_dot = 0.5 * (0, w_world) (x) q
    dqw = 0.5 * (-w[0] * qx - w[1] * qy - w[2] * qz)
    dqx = 0.5 * (w[0] * qw + w[1] * qz - w[2] * qy)
    dqy = 0.5 * (-w[0] * qz + w[1] * qw + w[2] * qx)
    dqz = 0.5 * (w[0] * qy - w[1] * qx + w[2] * qw)
    qn = np.stack([qw + dt * dqw, qx + dt * dqx, qy + dt * dqy, qz + dt * dqz])
    s["base_quat"] = np.ascontiguousarray(qn / np.linalg.norm(qn, axis=0, keepdims=True))
    return s


def shard_range(n_total, rank, world_size):
    """contiguous instance range [first, first+count) owned by `rank`: robots never migrate, no collective
    on the hot path; the remainder goes to the lowest ranks."""
    base, rem = divmod(int(n_total), int(world_size))
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def rollout_start(N, seed=2024, first=0):
    """config 5: start of a closed-loop rollout -- nominal stance with 20 % of the randomised offsets of `randomised`, references =
    initial pose, and one constant /cmd_vel per robot: forward 0.1..0.4 m/s, lateral +-0.1 m/s, every 4th robot turning (yaw rate
    +-0.3 rad/s), every 16th standing.  Keyed by the global robot index like `randomised`, so sharding does not change the data."""
    st, ref = nominal(N)
    rs, _ = randomised(N, seed=seed, first=first, contacts="stance")
    st = {k: st[k] + 0.2 * (rs[k] - st[k]) for k in st}
    st["base_quat"] /= np.linalg.norm(st["base_quat"], axis=0, keepdims=True)
    idx = np.arange(first, first + N)
    u = ((idx * 2654435761) % 1000) / 999.0                     # cheap counter-based uniform in [0,1]
    v = ((idx * 40503 + 17) % 1000) / 999.0
    cmd = np.zeros((3, N)); cmd[0] = 0.1 + 0.3 * u; cmd[1] = 0.2 * v - 0.1; cmd[2] = np.where(idx % 4 == 3, 0.6 * u - 0.3, 0.0)
    cmd[:, idx % 16 == 15] = 0.0
    return st, ref, cmd


STRESS_KINDS = ("flight", "one_foot", "saturating", "fast", "tilted", "far_swing", "two_feet_same_side", "crouched")


def stress(N, seed=77, first=0):
    """Edge cases of the path, one kind per robot (kind = global index mod 8), on top of `randomised`:
    0 flight (contacts 0000: no stance rows, all four friction blocks vanish), 1 a single stance foot, 2 desired CoM accelerations of
    +-30 m/s^2 / +-60 rad/s^2 (torque limits and friction pyramids saturate), 3 joint rates +-6 rad/s and base twist +-2, 4 base roll /
    pitch +-0.6 rad, 5 single swing with the swing target 0.4 m away (large swing accelerations -> slacks), 6 both feet of one side in
    the air, 7 deep crouch (knees +-1.8 rad, base 0.25 m).  Used by the parity tests only; not a BASELINE config."""
    st, ref = randomised(N, seed=seed, first=first, contacts="stance")
    u = _uniform(seed + 1, first, N)
    idx = np.arange(first, first + N); kind = idx % 8
    c = np.ones((4, N), np.uint8)
    c[:, kind == 0] = 0
    one = kind == 1; c[:, one] = 0; c[(idx[one] // 8) % 4, np.nonzero(one)[0]] = 1
    sat = kind == 2
    ref["acc"][0:3, sat] = (60.0 * u[sat, 0:3] - 30.0).T; ref["acc"][3:6, sat] = (120.0 * u[sat, 3:6] - 60.0).T
    fast = kind == 3
    st["qd"][:, fast] = (12.0 * u[fast, 6:18] - 6.0).T; st["base_twist"][:, fast] = (4.0 * u[fast, 18:24] - 2.0).T
    tilt = kind == 4
    rpy = (1.2 * u[tilt, 24:27] - 0.6).T; rpy[2] *= 0.2
    st["base_quat"][:, tilt] = rpy_to_quat_wxyz(rpy[0], rpy[1], rpy[2])
    far = kind == 5
    leg = (idx[far] // 8) % 4; c[leg, np.nonzero(far)[0]] = 0
    for k, i in zip(leg, np.nonzero(far)[0]):
        ref["sw_pos"][3 * k:3 * k + 3, i] = [0.4 * (1 if k in (1, 2) else -1), 0.3 * (1 if k < 2 else -1), 0.4]
        ref["sw_vel"][3 * k:3 * k + 3, i] = 2.0 * u[i, 27:30] - 1.0
    side = kind == 6
    left = ((idx[side] // 8) % 2 == 0)
    c[0, np.nonzero(side)[0]] = (~left).astype(np.uint8); c[1, np.nonzero(side)[0]] = (~left).astype(np.uint8)
    c[2, np.nonzero(side)[0]] = left.astype(np.uint8); c[3, np.nonzero(side)[0]] = left.astype(np.uint8)
    cr = kind == 7
    st["q"][:, cr] = (np.array([0.0, -1.0, 1.8, 0.0, 1.0, -1.8, 0.0, 1.0, -1.8, 0.0, -1.0, 1.8])[:, None] + 0.1 * (u[cr, 30:42].T - 0.5))
    st["base_pos"][2, cr] = 0.25
    ref["contacts"] = np.ascontiguousarray(c)
    return {k: np.ascontiguousarray(v) for k, v in st.items()}, {k: np.ascontiguousarray(v) for k, v in ref.items()}
