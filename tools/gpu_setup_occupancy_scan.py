"""DEV: time of kernel 2a (warm-start TQ rebuild) against its residency: WBC_QP_SETUP_PAD adds shared memory per CTA.  Uses ncu-free timing: the solve_qp event pair
covers kernels 2a + 2, so the scan reports the difference to the 9-robots-per-SM run."""
import os, subprocess, sys, re
here = os.path.dirname(os.path.abspath(__file__))
base = None
for pad, rps in ((0, 9), (3500, 8), (7500, 7), (13000, 6), (20500, 5), (31000, 4), (50000, 3)):
    env = dict(os.environ, WBC_QP_SETUP_PAD=str(pad))
    out = subprocess.run([sys.executable, os.path.join(here, "gpu_profile_rollout.py"), "16384", "40"], env=env, capture_output=True, text=True).stdout
    v = [float(x) for x in re.findall(r"'solve_qp': ([0-9.]+)", out)]
    t = v[-1]
    if base is None: base = t
    print(f"pad {pad:6d} B -> {rps} robots per SM in kernel 2a: quiet tick solve (2a + 2) {t:.3f} ms, delta to 9/SM {t - base:+.3f} ms", flush=True)
