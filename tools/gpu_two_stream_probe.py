"""DEV: does sharding one GPU's batch over S independent handles / streams raise rollout throughput (kernels of one shard fill the gaps of the other)?
usage: gpu_two_stream_probe.py N ticks"""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import quadrupedwholebodycontroller_b200 as wbc
from quadrupedwholebodycontroller_b200 import workloads as wl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
T = int(sys.argv[2]) if len(sys.argv) > 2 else 100
for S, cps in ((1, None), (2, None), (2, "4"), (4, None), (1, None)):
    if cps: os.environ["WBC_QP_CTAS_PER_SM"] = cps
    else: os.environ.pop("WBC_QP_CTAS_PER_SM", None)
    streams = [torch.cuda.Stream() for _ in range(S)]; ctls = []
    for s in range(S):
        first, count = wl.shard_range(N, s, S)
        st, ref, cmd = wl.rollout_start(count, first=first)
        c = wbc.WholeBodyController(count, stream=streams[s].cuda_stream); c.plannerCreate(); c.plantEnable()
        c.setJointAndBaseState(**st); c.setReference(**ref); c.plannerSetCmdVel(cmd[0].copy(), cmd[1].copy(), cmd[2].copy()); ctls.append(c)
    for c in ctls: c.rollout(T)          # warm-up step (cold start + first gait phases)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(2):
        t0 = time.perf_counter()
        for k in range(T // 10):           # interleave the enqueueing so that neither stream runs ahead by more than 10 ticks
            for c in ctls: c.rollout(10)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0; best = min(best, dt)
    fails = sum(int(c.rolloutStats()["failed_ticks"].sum()) for c in ctls)
    print(f"shards {S} ctas/SM {cps or 'max'}: {T} ticks of {N} robots in {best * 1e3:8.1f} ms -> {N * T / best / 1e6:6.3f} M solves/s (failed ticks {fails})", flush=True)
    for c in ctls: c.close()
