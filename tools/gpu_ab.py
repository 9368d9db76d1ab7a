"""DEV: A/B two builds of the library on the same box: alternating runs of the cold 4096-robot solve, its hot-started tick and quiet closed-loop ticks.
usage: gpu_ab.py libA.so libB.so [reps]"""
import os, subprocess, sys, re
import numpy as np
libs = [os.path.abspath(sys.argv[1]), os.path.abspath(sys.argv[2])]; reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
here = os.path.dirname(os.path.abspath(__file__))
res = {l: {"cold": [], "tick1": [], "quiet": []} for l in libs}
for r in range(reps):
    for l in libs:
        env = dict(os.environ, WBC_B200_LIB=l)
        out = subprocess.run([sys.executable, os.path.join(here, "gpu_profile_run.py"), "4096", "stance"], env=env, capture_output=True, text=True).stdout
        v = [float(x) for x in re.findall(r"'solve_qp': ([0-9.]+)", out)]; res[l]["cold"].append(v[0]); res[l]["tick1"].append(v[1])
        out = subprocess.run([sys.executable, os.path.join(here, "gpu_profile_rollout.py"), "4096", "40"], env=env, capture_output=True, text=True).stdout
        v = [float(x) for x in re.findall(r"'solve_qp': ([0-9.]+)", out)]; res[l]["quiet"].append(sum(v))
for l in libs:
    print(os.path.basename(l), {k: f"{np.median(v):.3f} (min {min(v):.3f})" for k, v in res[l].items()})
a, b = libs
print("B/A median ratio", {k: round(float(np.median(res[b][k]) / np.median(res[a][k])), 4) for k in res[a]})
