#!/usr/bin/env python3
"""Offline ANYmal model compiler: URDF -> constant tables (C headers + JSON).

Run ONCE in the build container (the URDF lives in the read-only reference
mount and does not travel to the GPU box); the generated files are committed:

    python tools/urdf_compiler.py /root/reference/urdf/anymal.urdf

Outputs
  oracle/anymal_tree_full.h                         78-link tree (iDynTree view after
                                                    removeFakeLinks) for the CPU oracle
  quadrupedwholebodycontroller_b200/csrc/anymal_model.h
                                                    13-body lumped tree for the CUDA kernels
                                                    (also consumed by the oracle to prove that
                                                    lumping preserves every hot-path output)
  tests/golden/anymal_model.json                    both tables for the python tests

Conventions restated from iDynTree 12.3.2 (paths relative to lib/idyntree/src):
  * URDF origin rpy = fixed-axis XYZ, R = Rz(y) Ry(p) Rx(r)   (core/src/Rotation.cpp RPY)
  * inertial: inertia given at the CoM in the inertial-origin frame, rotated into the link
    frame, then moved to the link origin with the parallel-axis theorem
    (model_io/codecs/src/InertialElement.cpp:91-99, core/src/SpatialInertia.cpp:42-67)
  * revolute joint: parent_X_child(q) = rest * Rot(axis_child, q), motion subspace of the
    child w.r.t. the parent in child coordinates = [0; axis]
    (model/src/RevoluteJoint.cpp:128-140,225-239)
  * removeFakeLinks: massless leaf links hanging on a fixed joint become frames
    (model/src/ModelTransformers.cpp:31-147)
  * DOF order is fixed BY NAME to LH,LF,RF,RH x HAA,HFE,KFE which is what
    src/whole_body_controller.cpp:81,234 and config/controllers.yaml:8-19 require.
"""
import json
import os
import sys
import xml.etree.ElementTree as ET

import numpy as np

LEGS = ["LH", "LF", "RF", "RH"]
DOF_NAMES = [f"{l}_{j}" for l in LEGS for j in ("HAA", "HFE", "KFE")]
FOOT_NAMES = [f"{l}_FOOT" for l in LEGS]


def rpy_to_R(r, p, y):
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def vec(s, default="0 0 0"):
    return np.array([float(x) for x in (s if s is not None else default).split()])


class Inertia:
    """mass, m*c and rotational inertia about the frame ORIGIN (iDynTree SpatialInertia)."""

    def __init__(self, m=0.0, mc=None, I=None):
        self.m = m
        self.mc = np.zeros(3) if mc is None else mc
        self.I = np.zeros((3, 3)) if I is None else I

    @staticmethod
    def from_com(m, c, Icom):
        c = np.asarray(c, float)
        I = Icom + m * (np.dot(c, c) * np.eye(3) - np.outer(c, c)) if m > 0 else Icom
        return Inertia(m, m * c, I)

    def transformed(self, R, p):
        """inertia expressed in frame A given A_X_B = (R, p) and self expressed in B."""
        if self.m == 0.0:
            return Inertia()
        c = self.mc / self.m
        Icom = self.I - self.m * (np.dot(c, c) * np.eye(3) - np.outer(c, c))
        return Inertia.from_com(self.m, R @ c + p, R @ Icom @ R.T)

    def __add__(self, o):
        return Inertia(self.m + o.m, self.mc + o.mc, self.I + o.I)


def parse(urdf_path):
    root = ET.parse(urdf_path).getroot()
    links = {}
    order = []
    for l in root.findall("link"):
        name = l.get("name")
        order.append(name)
        ine = l.find("inertial")
        if ine is None:
            links[name] = Inertia()
            continue
        o = ine.find("origin")
        xyz = vec(o.get("xyz") if o is not None else None)
        rpy = vec(o.get("rpy") if o is not None else None)
        m = float(ine.find("mass").get("value"))
        it = ine.find("inertia")
        g = lambda k: float(it.get(k))
        Ic = np.array([[g("ixx"), g("ixy"), g("ixz")], [g("ixy"), g("iyy"), g("iyz")], [g("ixz"), g("iyz"), g("izz")]])
        Rl = rpy_to_R(*rpy)
        links[name] = Inertia.from_com(m, xyz, Rl @ Ic @ Rl.T)
    joints = []
    for j in root.findall("joint"):
        o = j.find("origin")
        a = j.find("axis")
        ax = vec(a.get("xyz")) if a is not None else np.array([1.0, 0, 0])
        ax = ax / np.linalg.norm(ax)
        joints.append(dict(name=j.get("name"), type=j.get("type"), parent=j.find("parent").get("link"),
                           child=j.find("child").get("link"),
                           R=rpy_to_R(*vec(o.get("rpy") if o is not None else None)),
                           p=vec(o.get("xyz") if o is not None else None), axis=ax))
    return order, links, joints


def remove_fake_links(order, links, joints):
    """ModelTransformers.cpp:31-147 -- massless leaves on fixed joints become frames."""
    children = {n: [] for n in order}
    for j in joints:
        children[j["parent"]].append(j)
    fake = set()
    for j in joints:
        c = j["child"]
        if j["type"] == "fixed" and links[c].m == 0.0 and not children[c]:
            fake.add(c)
    root = (set(order) - {j["child"] for j in joints}).pop()
    fake.discard(root)
    keep = [n for n in order if n not in fake]
    return keep, [j for j in joints if j["child"] not in fake], sorted(fake)


def build_tree(order, links, joints):
    """depth-first traversal from the root; returns list of nodes with parent index."""
    root = (set(order) - {j["child"] for j in joints}).pop()
    by_parent = {}
    for j in joints:
        by_parent.setdefault(j["parent"], []).append(j)
    nodes = []

    def visit(name, parent_idx, joint):
        idx = len(nodes)
        dof = DOF_NAMES.index(joint["name"]) if joint is not None and joint["type"] in ("revolute", "continuous") else -1
        nodes.append(dict(name=name, parent=parent_idx, dof=dof,
                          R=(joint["R"] if joint else np.eye(3)), p=(joint["p"] if joint else np.zeros(3)),
                          axis=(joint["axis"] if joint else np.array([1.0, 0, 0])),
                          m=links[name].m, mc=links[name].mc, I=links[name].I))
        for j in by_parent.get(name, []):
            visit(j["child"], idx, j)

    visit(root, -1, None)
    return nodes


def lump(nodes):
    """merge every fixed-joint subtree into the nearest body that owns a DOF (or the base)."""
    body_of = {}          # full index -> lumped index
    X_in_body = {}        # full index -> (R,p) of the link frame in its lumped body's frame
    bodies = []
    for i, n in enumerate(nodes):
        if n["parent"] < 0 or n["dof"] >= 0:
            b = len(bodies)
            body_of[i] = b
            X_in_body[i] = (np.eye(3), np.zeros(3))
            if n["parent"] < 0:
                bodies.append(dict(name=n["name"], parent=-1, dof=-1, R=np.eye(3), p=np.zeros(3),
                                   axis=n["axis"], ine=Inertia(n["m"], n["mc"], n["I"])))
            else:
                Rp, pp = X_in_body[n["parent"]]
                bodies.append(dict(name=n["name"], parent=body_of[n["parent"]], dof=n["dof"],
                                   R=Rp @ n["R"], p=Rp @ n["p"] + pp, axis=n["axis"],
                                   ine=Inertia(n["m"], n["mc"], n["I"])))
        else:
            Rp, pp = X_in_body[n["parent"]]
            R, p = Rp @ n["R"], Rp @ n["p"] + pp
            b = body_of[n["parent"]]
            body_of[i] = b
            X_in_body[i] = (R, p)
            bodies[b]["ine"] = bodies[b]["ine"] + Inertia(n["m"], n["mc"], n["I"]).transformed(R, p)
    out = []
    for b in bodies:
        out.append(dict(name=b["name"], parent=b["parent"], dof=b["dof"], R=b["R"], p=b["p"], axis=b["axis"],
                        m=b["ine"].m, mc=b["ine"].mc, I=b["ine"].I))
    feet = []
    for f in FOOT_NAMES:
        i = [k for k, n in enumerate(nodes) if n["name"] == f][0]
        feet.append(dict(name=f, body=body_of[i], R=X_in_body[i][0], p=X_in_body[i][1]))
    # renumber so that body b >= 1 owns DOF b-1 (LH,LF,RF,RH x HIP,THIGH,SHANK): the CUDA kernels index by leg
    perm = sorted(range(len(out)), key=lambda k: out[k]["dof"])
    new_of = {old: new for new, old in enumerate(perm)}
    out = [dict(out[old], parent=(new_of[out[old]["parent"]] if out[old]["parent"] >= 0 else -1)) for old in perm]
    for f in feet:
        f["body"] = new_of[f["body"]]
    return out, feet


def fmt(x):
    return repr(float(x))


def emit_tree(nodes, feet, prefix, guard, comment):
    L = [f"// GENERATED by tools/urdf_compiler.py from urdf/anymal.urdf -- do not edit.", f"// {comment}",
         f"#ifndef {guard}", f"#define {guard}", ""]
    n = len(nodes)
    L.append(f"#define {prefix}_NLINKS {n}")
    L.append(f"#define {prefix}_NDOF 12")
    L.append(f"#define {prefix}_TOTAL_MASS {fmt(sum(x['m'] for x in nodes))}")
    arr = lambda ty, name, vals, per: L.append(
        f"static const {ty} {prefix}_{name}[{n * per if per else n}] = {{\n  " + (",\n  " if per else ", ").join(vals) + "\n};")
    arr("int", "parent", [str(x["parent"]) for x in nodes], 0)
    arr("int", "dof", [str(x["dof"]) for x in nodes], 0)
    arr("double", "restR", [", ".join(fmt(v) for v in x["R"].reshape(-1)) for x in nodes], 9)
    arr("double", "restP", [", ".join(fmt(v) for v in x["p"]) for x in nodes], 3)
    arr("double", "axis", [", ".join(fmt(v) for v in x["axis"]) for x in nodes], 3)
    arr("double", "mass", [fmt(x["m"]) for x in nodes], 0)
    arr("double", "mcom", [", ".join(fmt(v) for v in x["mc"]) for x in nodes], 3)
    arr("double", "inertia", [", ".join(fmt(v) for v in x["I"].reshape(-1)) for x in nodes], 9)
    L.append(f"static const int {prefix}_foot_link[4] = {{{', '.join(str(f['body']) for f in feet)}}};")
    L.append(f"static const double {prefix}_footR[36] = {{\n  " +
             ",\n  ".join(", ".join(fmt(v) for v in f["R"].reshape(-1)) for f in feet) + "\n};")
    L.append(f"static const double {prefix}_footP[12] = {{\n  " +
             ",\n  ".join(", ".join(fmt(v) for v in f["p"]) for f in feet) + "\n};")
    L.append("// link names: " + " ".join(x["name"] for x in nodes))
    L += ["", f"#endif  // {guard}", ""]
    return "\n".join(L)


def to_json(nodes, feet):
    return dict(links=[dict(name=x["name"], parent=x["parent"], dof=x["dof"], R=x["R"].tolist(), p=x["p"].tolist(),
                            axis=x["axis"].tolist(), m=x["m"], mc=x["mc"].tolist(), I=x["I"].tolist()) for x in nodes],
                feet=[dict(name=f["name"], link=f["body"], R=f["R"].tolist(), p=f["p"].tolist()) for f in feet])


def main():
    urdf = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/urdf/anymal.urdf"
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    order, links, joints = parse(urdf)
    keep, kjoints, fake = remove_fake_links(order, links, joints)
    full = build_tree(keep, links, kjoints)
    full_feet = [dict(name=f, body=[k for k, n in enumerate(full) if n["name"] == f][0], R=np.eye(3), p=np.zeros(3))
                 for f in FOOT_NAMES]
    lumped, feet = lump(full)
    assert len(lumped) == 13, len(lumped)
    print(f"URDF links {len(order)}, after removeFakeLinks {len(full)} (fake: {len(fake)}), lumped bodies {len(lumped)}")
    print("total mass", sum(x["m"] for x in full), sum(x["m"] for x in lumped))
    with open(os.path.join(repo, "oracle", "anymal_tree_full.h"), "w") as f:
        f.write(emit_tree(full, full_feet, "ANYMAL_FULL", "ANYMAL_TREE_FULL_H",
                          f"{len(full)}-link tree as iDynTree sees it after removeFakeLinks (oracle only)"))
    with open(os.path.join(repo, "quadrupedwholebodycontroller_b200", "csrc", "anymal_model.h"), "w") as f:
        f.write(emit_tree(lumped, feet, "ANYMAL", "ANYMAL_MODEL_H",
                          "13-body lumped model: base + 4 legs x {HIP,THIGH,SHANK}; DOF order LH,LF,RF,RH x HAA,HFE,KFE"))
    with open(os.path.join(repo, "tests", "golden", "anymal_model.json"), "w") as f:
        json.dump(dict(full=to_json(full, full_feet), lumped=to_json(lumped, feet), fake_links=fake), f)


if __name__ == "__main__":
    main()
