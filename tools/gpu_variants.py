"""DEV: same-box comparison of several builds / settings of the library.
usage: gpu_variants.py spec.json [reps]       spec = [{"label":..., "lib": path, "env": {...}}, ...]
Each variant runs in its own process (the library path and the DEV knobs are read at load / create time): the cold solve of 4096 stance
robots + its hot-started tick (tools/gpu_profile_run.py shape), optionally the quiet closed-loop ticks, and a digest of the outputs
(x, working set, rc, nWSR) so that a variant that changes a single bit of a result shows up."""
import hashlib, json, os, subprocess, sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
CHILD = r"""
import sys, os, json, hashlib
import numpy as np
sys.path.insert(0, os.path.dirname(%(here)r))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl
N = int(os.environ.get("VAR_N", "4096")); contacts = os.environ.get("VAR_CONTACTS", "stance"); reps = int(os.environ.get("VAR_REPS", "3"))
st, ref = wl.randomised(N, seed=1234, contacts=contacts); st1 = wl.advance_state(st)
ctl = wbc.WholeBodyController(N)
for kv in os.environ.get("VAR_OPTIONS", "").split(","):
    if kv:
        ctl.setOption(kv.split("=")[0], int(kv.split("=")[1]))
cold, tick1 = [], []
h = hashlib.sha256()
for rep in range(reps):
    ctl.setInitialState(); ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.step(); ctl.synchronize(); cold.append(ctl.lastKernelMs()["solve_qp"])
    if rep == 0:
        o = ctl.getOutputs()
        for k in ("x", "ws", "rc", "nwsr", "tau"):
            h.update(np.ascontiguousarray(o[k]).tobytes())
    ctl.setJointAndBaseState(**st1); ctl.step(); ctl.synchronize(); tick1.append(ctl.lastKernelMs()["solve_qp"])
    if rep == 0:
        o = ctl.getOutputs()
        for k in ("x", "ws", "rc", "nwsr", "tau"):
            h.update(np.ascontiguousarray(o[k]).tobytes())
ctl.close()
res = dict(cold=min(cold), cold_med=float(np.median(cold)), tick1=min(tick1), digest=h.hexdigest()[:16])
T = int(os.environ.get("VAR_ROLLOUT", "0"))
if T:
    NR = int(os.environ.get("VAR_ROLLOUT_N", "4096"))
    st, ref, cmd = wl.rollout_start(NR)
    ctl = wbc.WholeBodyController(NR); ctl.plannerCreate(); ctl.plantEnable()
    for kv in os.environ.get("VAR_OPTIONS", "").split(","):
        if kv:
            ctl.setOption(kv.split("=")[0], int(kv.split("=")[1]))
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref); ctl.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy())
    q = []
    for t in range(T):
        ctl.rollout(1)
        if t >= T - 3:
            ctl.synchronize(); q.append(ctl.lastKernelMs()["solve_qp"])
    o = ctl.getOutputs(); hh = hashlib.sha256()
    for k in ("x", "ws", "rc", "nwsr", "tau"):
        hh.update(np.ascontiguousarray(o[k]).tobytes())
    res["quiet3"] = sum(q); res["rdigest"] = hh.hexdigest()[:16]
    ctl.close()
print("RESULT " + json.dumps(res))
""" % dict(here=HERE)

spec = json.load(open(sys.argv[1])); reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
acc = {v["label"]: [] for v in spec}
for r in range(reps):
    for v in spec:
        env = dict(os.environ, **{k: str(x) for k, x in v.get("env", {}).items()})
        if v.get("lib"):
            env["WBC_B200_LIB"] = os.path.abspath(v["lib"])
        p = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=180)
        line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
        if not line:
            print(v["label"], "FAILED rc", p.returncode, p.stderr[-800:], flush=True); continue
        acc[v["label"]].append(json.loads(line[0][7:]))
base = None
for v in spec:
    rs = acc[v["label"]]
    if not rs:
        continue
    out = {k: round(min(r[k] for r in rs), 3) for k in rs[0] if not k.endswith("digest")}
    out.update({k: rs[0][k] for k in rs[0] if k.endswith("digest")})
    if base is None:
        base = out
    rel = {k: round(base[k] / out[k], 3) for k in ("cold", "tick1", "quiet3") if k in out and k in base}
    print(f"{v['label']:28s} {out}  speedup vs first {rel}  same_bits {out.get('digest') == base.get('digest')}", flush=True)
