"""DEV: BASELINE configs[0] on the GPU -- ONE robot, nominal stance, cold init then hot starts on the same state: per-tick latency of the
three controller calls (host wall clock around step() + synchronize(), and the kernels' own time from CUDA events)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

for N in (1, 32, 148, 740, 1332):
    st, ref = wl.nominal(N)
    ctl = wbc.WholeBodyController(N)
    ctl.setJointAndBaseState(**st); ctl.setReference(**ref)
    wall, kern = [], []
    for t in range(300):
        t0 = time.perf_counter(); ctl.step(); ctl.synchronize(); t1 = time.perf_counter()
        k = ctl.lastKernelMs()
        if t == 0:
            cold = (t1 - t0) * 1e3, sum(k.values())
        elif t >= 50:
            wall.append((t1 - t0) * 1e3); kern.append(sum(k.values()))
    o = ctl.getOutputs()
    print(f"N={N:5d}: cold tick {cold[0]:.3f} ms wall ({cold[1]:.3f} ms kernels, nWSR {o['nwsr'].max()} on the last tick); hot-started tick p50 {np.median(wall):.3f} ms wall, "
          f"{np.median(kern):.3f} ms kernels, p99 {np.percentile(wall, 99):.3f} ms", flush=True)
    ctl.close()
