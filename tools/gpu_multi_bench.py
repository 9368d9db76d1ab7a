"""Library-level multi-GPU measurement (wbc_multi_*): ONE process, one host thread, D GPUs.  cold4096 per device (weak scaling), inputs from host
memory, device-side time of the step = max over the devices' kernel event pairs, wall time of the whole call sequence, and the optional
gather of (tau, grf, rc, nWSR) into device memory of GPU 0, timed separately (wbc_multi_last_gather_ms).  Prints one JSON line.
usage: gpu_multi_bench.py [n_devices] [robots_per_device] [steps]"""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

D = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
per = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
N = D * per
st, ref = wl.randomised(N, seed=1234, contacts="stance")
pst = {k: torch.from_numpy(v).pin_memory() for k, v in st.items()}; pref = {k: torch.from_numpy(v).pin_memory() for k, v in ref.items()}
m = wbc.MultiGpuController(N, list(range(D)))
tau = torch.zeros(12, N, dtype=torch.float64, device="cuda:0"); grf = torch.zeros(12, N, dtype=torch.float64, device="cuda:0")
rc = torch.zeros(N, dtype=torch.int32, device="cuda:0"); nw = torch.zeros(N, dtype=torch.int32, device="cuda:0")
wall, gather = [], []
for s in range(steps + 2):
    m.setInitialState(); m.synchronize()
    t0 = time.perf_counter()
    m.setJointAndBaseState(**pst); m.setReference(**pref); m.step(); m.synchronize()
    t1 = time.perf_counter()
    m.getOutputs(tau=tau, grf=grf, rc=rc, nwsr=nw)
    if s >= 2:
        wall.append(t1 - t0); gather.append(m.lastGatherMs())
o = m.getOutputs()
one = wbc.WholeBodyController(per, device=0); first = {k: np.ascontiguousarray(v[..., :per]) for k, v in st.items()}; rfirst = {k: np.ascontiguousarray(v[..., :per]) for k, v in ref.items()}
one.setJointAndBaseState(**first); one.setReference(**rfirst); one.step(); o1 = one.getOutputs()
same = all(np.array_equal(o[k][..., :per], o1[k]) for k in o1)
print(json.dumps(dict(what="library-level multi-GPU (wbc_multi_*), one process / one host thread", n_devices=D, robots_per_device=per, robots_total=N, steps=steps,
                      step_wall_ms=float(np.mean(wall) * 1e3), solves_per_s=float(N / np.mean(wall)), gather_ms_device=float(np.mean(gather)),
                      gather_bytes=int(N * (24 * 8 + 8)), gathered_equals_single_device=bool(same and np.array_equal(tau.cpu().numpy(), o["tau"])),
                      rc0_frac=float((o["rc"] == 0).mean()))))
m.close(); one.close()
