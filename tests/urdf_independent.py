"""An INDEPENDENT rigid-body model of the ANYmal URDF for the parity tests: its own xml.etree walk of
urdf/anymal.urdf -- nothing shared with tools/urdf_compiler.py, which produces BOTH the device's 13-body table and the oracle's 78-link
tree -- and a different formulation of every quantity:

    mass matrix     M = sum_i  m_i Jv_i' Jv_i + Jw_i' (R_i I_i R_i') Jw_i     from geometric link Jacobians (no CRBA)
    gravity term    g = - sum_i m_i Jv_i' g_vec
    bias forces     h = sum_i Jv_i' m_i (dJv_i nu) + Jw_i' (Ibar_i dJw_i nu + w_i x Ibar_i w_i) + g, with dJ nu by central differences along the motion (no RNEA)
    CoM, CoM velocity, foot positions and foot Jacobians from the same forward kinematics

Velocity convention = iDynTree MIXED (KinDynComputations.cpp:822-825): nu = [base linear velocity (world); base angular velocity (world); joint rates].
Run as a script (where /root/reference is mounted) it writes tests/golden/rbd_independent.npz: seeded states and this model's M, foot Jacobians, CoM
and the centroidal blocks of the QP (wbc.cpp:256-320, 479-497) for the GPU test, which has no access to the URDF.
"""
import os
import xml.etree.ElementTree as ET

import numpy as np

URDF = "/root/reference/urdf/anymal.urdf"
DOFS = [f"{leg}_{j}" for leg in ("LH", "LF", "RF", "RH") for j in ("HAA", "HFE", "KFE")]      # wbc.cpp:81,234
FEET = [f"{leg}_FOOT" for leg in ("LH", "LF", "RF", "RH")]                                    # wbc.cpp:327-341


def rpy_matrix(r, p, y):
    """URDF fixed-axis roll-pitch-yaw: R = Rz(y) Ry(p) Rx(r)"""
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]]); Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]]); Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _origin(el):
    o = el.find("origin") if el is not None else None
    xyz = np.array([float(v) for v in (o.get("xyz", "0 0 0") if o is not None else "0 0 0").split()])
    rpy = [float(v) for v in (o.get("rpy", "0 0 0") if o is not None else "0 0 0").split()]
    return rpy_matrix(*rpy), xyz


def skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])


def rodrigues(w):
    a = np.linalg.norm(w)
    if a < 1e-300:
        return np.eye(3)
    k = skew(w / a)
    return np.eye(3) + np.sin(a) * k + (1 - np.cos(a)) * (k @ k)


def quat_wxyz_to_R(q):
    """Eigen::Quaterniond::toRotationMatrix (not normalised), wbc.cpp:205-215"""
    w, x, y, z = q
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz, txx, txy, txz, tyy, tyz, tzz = tx * w, ty * w, tz * w, tx * x, ty * x, tz * x, ty * y, tz * y, tz * z
    return np.array([[1 - (tyy + tzz), txy - twz, txz + twy], [txy + twz, 1 - (txx + tzz), tyz - twx], [txz - twy, tyz + twx, 1 - (txx + tyy)]])


class Model:
    def __init__(self, path=URDF):
        root = ET.parse(path).getroot()
        self.links = {}
        for l in root.findall("link"):
            ine = l.find("inertial")
            if ine is None:
                self.links[l.get("name")] = None; continue
            R, p = _origin(ine); I = ine.find("inertia")
            g = lambda k: float(I.get(k))
            Im = np.array([[g("ixx"), g("ixy"), g("ixz")], [g("ixy"), g("iyy"), g("iyz")], [g("ixz"), g("iyz"), g("izz")]])
            self.links[l.get("name")] = dict(m=float(ine.find("mass").get("value")), c=p, I=R @ Im @ R.T)        # inertia about the CoM, in link axes
        self.children = {}
        for j in root.findall("joint"):
            R, p = _origin(j); ax = j.find("axis")
            a = np.array([float(v) for v in ax.get("xyz").split()]) if ax is not None else np.array([1.0, 0, 0])
            self.children.setdefault(j.find("parent").get("link"), []).append(dict(name=j.get("name"), type=j.get("type"), child=j.find("child").get("link"), R=R, p=p, axis=a))
        kids = {c["child"] for cs in self.children.values() for c in cs}
        self.root = [n for n in self.links if n not in kids][0]

    def total_mass(self):
        return sum(l["m"] for l in self.links.values() if l is not None)

    def kinematics(self, pos, R, q):
        """world pose of every link + the chain of (dof, world axis, world joint position) from the base to it"""
        out = {}
        stack = [(self.root, np.asarray(R, float), np.asarray(pos, float), [])]
        while stack:
            name, Rw, pw, chain = stack.pop()
            out[name] = (Rw, pw, chain)
            for j in self.children.get(name, []):
                Rj, pj = Rw @ j["R"], pw + Rw @ j["p"]
                ch = chain
                if j["type"] in ("revolute", "continuous"):
                    k = DOFS.index(j["name"]); aw = Rj @ j["axis"]
                    Rj = Rj @ rodrigues(j["axis"] * q[k]); ch = chain + [(k, aw, pj)]
                elif j["type"] != "fixed":
                    raise ValueError(j["type"])
                stack.append((j["child"], Rj, pj, ch))
        return out

    @staticmethod
    def point_jacobian(p, chain, pbase):
        """6x18 MIXED Jacobian of a point p rigidly attached to a link with joint chain `chain`: rows 0-2 linear, 3-5 angular velocity (world)"""
        J = np.zeros((6, 18)); J[0:3, 0:3] = np.eye(3); J[0:3, 3:6] = -skew(p - pbase); J[3:6, 3:6] = np.eye(3)
        for k, a, pj in chain:
            J[0:3, 6 + k] = np.cross(a, p - pj); J[3:6, 6 + k] = a
        return J

    def evaluate(self, pos, quat, twist, q, dq, gvec=(0.0, 0.0, 9.81)):
        """gvec = the world gravity vector AS THE REFERENCE PASSES IT to setRobotState (wbc.cpp:78,258: (0, 0, +9.81))"""
        pos, q, dq, twist = (np.asarray(v, float) for v in (pos, q, dq, twist))
        R = quat_wxyz_to_R(quat); nu = np.concatenate([twist, dq])
        kin = self.kinematics(pos, R, q)
        M = np.zeros((18, 18)); g = np.zeros(18); mtot = 0.0; com = np.zeros(3); per_link = []
        for name, l in self.links.items():
            if l is None:
                continue
            Rw, pw, chain = kin[name]; pc = pw + Rw @ l["c"]; J = self.point_jacobian(pc, chain, pos); Ibar = Rw @ l["I"] @ Rw.T
            M += l["m"] * J[0:3].T @ J[0:3] + J[3:6].T @ Ibar @ J[3:6]
            g -= l["m"] * (J[0:3].T @ np.asarray(gvec, float))             # generalized gravity forces: - Jv' (m gvec)
            mtot += l["m"]; com += l["m"] * pc; per_link.append((name, l, J, Ibar))
        com /= mtot
        Jcom = sum(l["m"] * J[0:3] for _, l, J, _ in per_link) / mtot
        feet = [self.point_jacobian(kin[f][1], kin[f][2], pos) for f in FEET]
        return dict(M=M, g=g, com=com, vcom=Jcom @ nu, J=np.stack(feet), footPos=np.concatenate([kin[f][1] for f in FEET]),
                    footVel=np.stack([Jf @ nu for Jf in feet]), mass=mtot, per_link=per_link, kin=kin, nu=nu, R=R, pbase=pos)

    def bias(self, pos, quat, twist, q, dq, gvec=(0.0, 0.0, 9.81), h=1e-5):
        """h(q, nu) = C nu + g without RNEA: per link, Jv' m (dJv nu) + Jw' (Ibar dJw nu + w x Ibar w), dJ nu by central differences along the motion"""
        e0 = self.evaluate(pos, quat, twist, q, dq, gvec); nu = e0["nu"]; R = e0["R"]
        def link_jacs(sign):
            p2 = np.asarray(pos, float) + sign * h * nu[0:3]; R2 = rodrigues(sign * h * nu[3:6]) @ R; q2 = np.asarray(q, float) + sign * h * nu[6:]
            kin = self.kinematics(p2, R2, q2); out = {}
            for name, l in self.links.items():
                if l is not None:
                    Rw, pw, chain = kin[name]; out[name] = self.point_jacobian(pw + Rw @ l["c"], chain, p2)
            return out
        Jp, Jm = link_jacs(+1), link_jacs(-1)
        hv = e0["g"].copy()
        for name, l, J, Ibar in e0["per_link"]:
            dJnu = (Jp[name] - Jm[name]) @ nu / (2 * h); w = J[3:6] @ nu
            hv += J[0:3].T @ (l["m"] * dJnu[0:3]) + J[3:6].T @ (Ibar @ dJnu[3:6] + np.cross(w, Ibar @ w))
        return hv


def centroidal_blocks(e, contacts):
    """wbc.cpp:256-320 in numpy: T = [Xinv Mbb^-1 [I 0] M; 0 I], Mc = T^-T M T^-1, Jst_c = (c . J_lin) T^-1 -- the blocks that appear in the QP's A"""
    M = e["M"]; d = e["com"] - e["pbase"]
    Xinv = np.eye(6); Xinv[0:3, 3:6] = -skew(d)
    T = np.eye(18); T[0:6, :] = Xinv @ np.linalg.solve(M[0:6, 0:6], M[0:6, :])
    Ti = np.linalg.inv(T); Mc = Ti.T @ M @ Ti
    Jlin = np.concatenate([e["J"][l, 0:3, :] for l in range(4)]); c = np.repeat(np.asarray(contacts, float), 3)[:, None]
    return dict(T=T, Mc=Mc, Jst_c=(c * Jlin) @ Ti, Jsw_c=((1 - c) * Jlin) @ Ti)


def fixture_states(n=16, seed=909):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from quadrupedwholebodycontroller_b200 import workloads as wl
    return wl.randomised(n, seed=seed, contacts="mixed")


if __name__ == "__main__":
    mdl = Model(); st, ref = fixture_states()
    out = dict(mass=mdl.total_mass())
    for i in range(st["q"].shape[1]):
        e = mdl.evaluate(st["base_pos"][:, i], st["base_quat"][:, i], st["base_twist"][:, i], st["q"][:, i], st["qd"][:, i])
        cb = centroidal_blocks(e, ref["contacts"][:, i])
        for k in ("M", "J", "com", "footPos"):
            out[f"{k}_{i}"] = e[k]
        for k in ("Mc", "Jst_c", "Jsw_c"):
            out[f"{k}_{i}"] = cb[k]
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rbd_independent.npz"), **out)
    print("written", len(out), "arrays; total mass", out["mass"])
