#!/usr/bin/env python3
"""bench.py -- ANYmal whole-body-controller QP solves per second (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cold4096|warm65536]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

A "step" is one pass of the hot path (updateState -> solveQP -> computeJointTorques) over one batch of synthetic robot
states.  Default workload = BASELINE.json configs[1]: 4096 randomised 4-contact stance states per GPU, every solve a cold
start (SQProblem::init), so each step first resets the batch.  One process per GPU; the batch shards by instance index
(weak scaling, no collective on the hot path); timing is on the device with CUDA events, MAX over ranks.
One JSON line on rank 0.  --impl reference times the reference's CPU implementation of the path (the oracle: restated
iDynTree + controller algebra + the real vendored qpOASES) on the host cores, rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC, UNIT = "anymal_wbc_qp_solves_per_sec", "solves/s"
WORKLOADS = {
    # name: (robots per GPU, contacts, cold?)
    "cold4096": dict(n=4096, contacts="stance", cold=True, desc="BASELINE configs[1]: 4096 randomised 4-contact ANYmal stance states per GPU, 42-var/70-constraint QP, cold SQProblem::init every step"),
    "warm65536": dict(n=65536, contacts="diagonal", cold=False, desc="BASELINE configs[2]: 65536 diagonal-pair states per GPU, hot-started SQProblem::hotstart ticks"),
    "mixed1m": dict(n=1 << 20, contacts="mixed", cold=True, strong=True, switching_frac=0.1,
                    desc="BASELINE configs[3]: 2^20 randomised states in total, contacts 1111 / single swing / diagonal pairs (50/40/10 %), switching flag on 10 %, cold SQProblem::init, "
                         "sharded by robot index over the GPUs (strong scaling)"),
    "rollout16384": dict(n=16384, contacts="stance", cold=False, rollout=100,
                         desc="BASELINE configs[4]: closed loop of 16384 ANYmal WBC instances per GPU fed by the batched motion planner (crawl gait, /cmd_vel per robot), "
                              "ideal-actuator plant; one step = 100 consecutive 400 Hz control ticks (25 planner periods) of the running rollout, hot-started solves incl. contact switches"),
}
IN_BYTES_PER_ROBOT = (12 + 12 + 3 + 4 + 6 + 6 + 6 + 6 + 12 + 12 + 12) * 8 + 4 + 1      # state + reference + contacts + switching
OUT_BYTES_PER_ROBOT = (12 + 12) * 8 + 4 + 4                                            # tau, grf, rc, nwsr
# algorithmic HBM bytes of ONE solve in the solver kernel: compact QP in, warm-start state in/out, x + working set out
QPDATA_B, QPSTATE_B, XWS_B = 710 * 8, 3080, 42 * 8 + 112 + 8


def dist_env():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True); self.index = index; self.samples = []; self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        self.stop_flag = True; self.join(timeout=6)
        if not self.samples:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(len(s) > 2 + k and s[2 + k].lower().startswith("active") for s in self.samples)]
        return dict(sm_mhz=sm[len(sm) // 2] if sm else None, sm_max_mhz=float(self.samples[0][1]) if self.samples[0][1].replace(".", "").isdigit() else None,
                    reasons=reasons, samples=len(self.samples))


def cpu_reference(n_sample, contacts, cold, threads=0, rollout=0, switching_frac=0.0):
    """the reference's CPU path (oracle) on `n_sample` robots of the same workload; returns (solves/s, seconds, cores)"""
    from oracle import oracle as orc
    from quadrupedwholebodycontroller_b200 import workloads as wl
    if rollout:
        st, ref, cmd = wl.rollout_start(n_sample)
        ob = orc.OracleBatch(n_sample); cores = threads or orc.lib().orc_max_threads()
        ob.set_state(**st); ob.set_reference(**ref)
        devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(2); os.dup2(devnull, 2)
        try:
            secs = ob.rollout(rollout, cmd, nthreads=cores, record=False)["seconds"]        # includes the cold first tick (1 of `rollout` ticks)
        finally:
            os.dup2(saved, 2); os.close(devnull); os.close(saved)
        return n_sample * rollout / secs, secs, cores
    st, ref = wl.randomised(n_sample, seed=1234, contacts=contacts, switching_frac=switching_frac)
    ob = orc.OracleBatch(n_sample)
    cores = threads or orc.lib().orc_max_threads()
    ob.set_state(**st); ob.set_reference(**ref)
    if not cold:
        ob.step(nthreads=cores); st = wl.advance_state(st); ob.set_state(**st)
    devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(2); os.dup2(devnull, 2)      # qpOASES prints nWSR warnings on stderr
    try:
        secs = ob.step(nthreads=cores)
    finally:
        os.dup2(saved, 2); os.close(devnull); os.close(saved)
    return n_sample / secs, secs, cores


def run_reference(args, rank, world):
    if rank != 0:
        return
    spec = WORKLOADS[args.workload]
    ro = spec.get("rollout", 0)
    n_sample = 64 if ro else (512 if spec["cold"] else 4096)
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_reference(min(64, n_sample), spec["contacts"], spec["cold"], rollout=min(ro, 8))
    vals, secs = [], []
    for _ in range(args.steps):
        v, s, cores = cpu_reference(n_sample, spec["contacts"], spec["cold"], rollout=ro, switching_frac=spec.get("switching_frac", 0.0)); vals.append(v); secs.append(s)
    value = n_sample * max(1, ro) * len(secs) / sum(secs)
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup, ms_per_step=1e3 * sum(secs) / len(secs),
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64", data="synthetic", impl="reference",
                config=dict(workload=args.workload, description=spec["desc"], robots_per_step=n_sample, ticks_per_step=max(1, ro), l2="n/a (host)"),
                cpu_baseline=dict(value=value, unit=UNIT, cores=cores, kind="port", reference_parts="qpOASES 3.2.2 solver: the real, unmodified vendored library (oracle/_ref); iDynTree subset, controller algebra, planner, plant: restated",
                                  sample=f"{n_sample} robots of the {args.workload} workload per step: restated iDynTree/controller algebra (port) + the real vendored qpOASES 3.2.2 (reference), {cores} host threads"),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1); ap.add_argument("--steps", type=int, default=10); ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"]); ap.add_argument("--workload", default="cold4096", choices=list(WORKLOADS))
    ap.add_argument("--robots", type=int, default=0, help="override robots per GPU"); ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank, world, local = dist_env()
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    import quadrupedwholebodycontroller_b200 as wbc
    from quadrupedwholebodycontroller_b200 import workloads as wl
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the controller has no CPU fallback")
    args.warmup = max(args.warmup, 3)
    torch.cuda.set_device(local); dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    spec = WORKLOADS[args.workload]
    n = args.robots or spec["n"]
    if spec.get("strong"):
        first, n = wl.shard_range(n, rank, world)                      # strong scaling: the global batch is split over the ranks
    else:
        first = rank * n                                               # weak scaling: contiguous shard [rank*n, (rank+1)*n) of the global batch
    st, ref = wl.randomised(n, seed=1234, first=first, contacts=spec["contacts"], switching_frac=spec.get("switching_frac", 0.0))
    stream = torch.cuda.Stream(device=dev)                          # a real (non-NULL) stream shared by torch's events and the library
    torch.cuda.set_stream(stream)
    ctl = wbc.WholeBodyController(n, device=local, stream=stream.cuda_stream)
    ctl.profile(True)
    dst = {k: torch.from_numpy(v).to(dev) for k, v in st.items()}; dref = {k: torch.from_numpy(v).to(dev) for k, v in ref.items()}
    pst = {k: torch.from_numpy(v).pin_memory() for k, v in st.items()}; pref = {k: torch.from_numpy(v).pin_memory() for k, v in ref.items()}
    st1 = wl.advance_state(st); dst1 = {k: torch.from_numpy(v).to(dev) for k, v in st1.items()}; pst1 = {k: torch.from_numpy(v).pin_memory() for k, v in st1.items()}
    out_host = dict(tau=torch.zeros(12, n, dtype=torch.float64).pin_memory(), grf=torch.zeros(12, n, dtype=torch.float64).pin_memory(),
                    rc=torch.zeros(n, dtype=torch.int32).pin_memory(), nwsr=torch.zeros(n, dtype=torch.int32).pin_memory())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ro = spec.get("rollout", 0)
    if ro:
        rst, rref, rcmd = wl.rollout_start(n, first=first)
        pcmd = [torch.from_numpy(np.ascontiguousarray(rcmd[k])).pin_memory() for k in range(3)]
        ctl.plannerCreate(); ctl.plantEnable()
        ctl.setJointAndBaseState(**{k: torch.from_numpy(v).to(dev) for k, v in rst.items()}); ctl.setReference(**{k: torch.from_numpy(v).to(dev) for k, v in rref.items()})
        ctl.plannerSetCmdVel(*[c.to(dev) for c in pcmd])
        ctl.rollout(4)                                                 # untimed: the cold tick (SQProblem::init) and the first hot starts

    def prepare(host):
        """untimed: L2 flush; for the warm workload also the cold tick that the timed hot start follows"""
        if ro:
            flush.fill_(rank + 1); return
        if not spec["cold"]:
            ctl.setInitialState(); ctl.setJointAndBaseState(**(dst)); ctl.setReference(**dref); ctl.step()
        flush.fill_(rank + 1)

    def one_step(host):
        """timed: (reset for the cold workload) + inputs + the three hot methods (+ result read-back when host=True)"""
        if ro:                                                         # `ro` consecutive ticks of the running closed loop, all on the device
            if host:
                ctl.plannerSetCmdVel(*pcmd)                            # /cmd_vel from pinned host memory
            ctl.rollout(ro)
            if host:
                ctl.getOutputs(tau=out_host["tau"], grf=out_host["grf"], rc=out_host["rc"], nwsr=out_host["nwsr"])
            return
        s_in, r_in = ((pst, pref) if spec["cold"] else (pst1, pref)) if host else ((dst, dref) if spec["cold"] else (dst1, dref))
        if spec["cold"]:
            ctl.setInitialState()
        ctl.setJointAndBaseState(**s_in); ctl.setReference(**r_in)
        ctl.updateState(); ctl.solveQP(); ctl.computeJointTorques()
        if host:
            ctl.getOutputs(tau=out_host["tau"], grf=out_host["grf"], rc=out_host["rc"], nwsr=out_host["nwsr"])

    def timed(host, steps):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        kms = []
        for a, b in ev:
            prepare(host); barrier()
            a.record(stream); one_step(host); b.record(stream)
            torch.cuda.synchronize()
            kms.append(ctl.lastKernelMs())
        ms = torch.tensor([a.elapsed_time(b) for a, b in ev], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)                  # a step ends when the slowest rank ends
        return ms.cpu().numpy(), kms

    for _ in range(args.warmup):
        prepare(False); one_step(False)
    barrier()
    sampler = ClockSampler(local); sampler.start()
    l0 = ctl.launchCount()
    ms_dev, kms = timed(False, args.steps)
    launches = ctl.launchCount() - l0
    prof = ctl.profile(True, fetch=True)
    o = ctl.getOutputs()
    ms_e2e, _ = timed(True, args.steps)
    clocks = sampler.summary()
    total_robots = (args.robots or spec["n"]) if spec.get("strong") else n * world
    solves_per_step = total_robots * max(1, ro)
    value = solves_per_step * args.steps / (ms_dev.sum() * 1e-3)
    e2e = solves_per_step * args.steps / (ms_e2e.sum() * 1e-3)
    if ro:                                                             # per-launch kernel times / flop counts: one more step replayed tick by tick
        kms, fl, nw, it, cy = [], 0.0, 0.0, 0.0, 0.0
        for _ in range(ro):
            ctl.rollout(1); kms.append(ctl.lastKernelMs()); pp = ctl.profile(True, fetch=True)
            fl += float(pp[:, 8].sum()); it += float(pp[:, 7].sum()); cy += float(pp[:, 6].sum())
        prof = np.zeros((1, 10)); prof[0, 8] = fl / ro; prof[0, 7] = it; prof[0, 6] = cy
        o = ctl.getOutputs(); st_ro = ctl.rolloutStats()
    qp_ms = float(np.mean([k["solve_qp"] for k in kms])); rbd_ms = float(np.mean([k["update_state"] for k in kms]))
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    alg_bytes = n * (QPDATA_B + (QPSTATE_B if spec["cold"] else 2 * QPSTATE_B) + XWS_B)
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE solver launch from an `ncu --set full` capture of exactly this launch shape (profiles/): only for 4096 robots
    traffic = {True: 23.7e6 + 0.015e6, False: 36.5e6 + 0.07e6}[bool(spec["cold"])] if (n == 4096 and not ro) else None
    fp64_peak = ctl.measureFp64Peak()
    alg_flops = float(prof[:, 8].sum())
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup, ms_per_step=float(ms_dev.mean()),
                higher_is_better=True, scaling="strong" if spec.get("strong") else "weak", vs_baseline=None, dtype="f64", data="synthetic",
                config=dict(workload=args.workload, description=spec["desc"], robots_per_gpu=n, robots_total=total_robots, parallelism=f"instance-sharded x{world}",
                            l2="flushed between timed steps (256 MiB write)", peaks="measured (MEASURED_PEAKS.json)" if peaks else "fallback"),
                latency_ms=dict(p50=float(np.percentile(ms_dev, 50)), p99=float(np.percentile(ms_dev, 99)), min=float(ms_dev.min()), max=float(ms_dev.max())),
                e2e=dict(value=e2e, unit=UNIT, h2d_bytes_per_step=IN_BYTES_PER_ROBOT * n, d2h_bytes_per_step=OUT_BYTES_PER_ROBOT * n, ms_per_step=float(ms_e2e.mean())),
                gpu_launches=int(launches), clocks=clocks,
                kernels_ms=dict(update_state=rbd_ms, solve_qp=qp_ms, solve_qp_share=qp_ms * max(1, ro) / float(ms_dev.mean()),
                                note=("per launch; rollout: mean over the ticks of one extra step replayed tick by tick" if ro else "per launch, mean over the timed steps")
                                + "; solve_qp = wbc_warm_setup_kernel (hot-started robots only) + wbc_solve_qp_kernel, one event pair"),
                roofline=dict(bound="hbm", achieved=alg_bytes / (qp_ms * 1e-3) / 1e9, peak=hbm_peak, unit="GB/s", frac=alg_bytes / (qp_ms * 1e-3) / 1e9 / hbm_peak,
                              traffic=traffic, traffic_source="profiles/r01a_ncu_summary.md (cold launch) / profiles/r01f_ncu_summary.md (hot-started launch), 4096 robots" if traffic else None, kernel="wbc_solve_qp_kernel", peak_source="MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                              note="solver is FP64-latency / shared-memory bound, not HBM bound: see fp64_roofline and DESIGN.md section 4"),
                fp64_roofline=dict(achieved=alg_flops / (qp_ms * 1e-3) / 1e9, peak=fp64_peak, unit="GFLOP/s", frac=alg_flops / (qp_ms * 1e-3) / 1e9 / fp64_peak,
                                   algorithmic_mflop_per_solve=alg_flops / n / 1e6, peak_source="wbc_measure_fp64_peak (DFMA microbenchmark, this run)"),
                solver=dict(rc0_frac=float((o["rc"] == 0).mean()), rc64_frac=float((o["rc"] == 64).mean()), nwsr_mean=float(o["nwsr"].mean()),
                            cycles_per_iteration=float(prof[:, 6].sum() / max(1, prof[:, 7].sum()))))
    if ro:
        tick_ms = np.array([k["update_state"] + k["solve_qp"] + k["compute_joint_torques"] for k in kms])
        line["tick_latency_ms"] = dict(p50=float(np.percentile(tick_ms, 50)), p99=float(np.percentile(tick_ms, 99)), min=float(tick_ms.min()), max=float(tick_ms.max()),
                                       note="one control tick of the whole batch (updateState + solveQP + computeJointTorques kernels), replayed step")
        line["config"]["ticks_per_step"] = ro
        line["solver"].update(failed_ticks=int(st_ro["failed_ticks"].sum()), nwsr_mean_rollout=float(st_ro["nwsr_sum"].sum() / max(1, st_ro["ticks"].sum())),
                              nwsr_max_rollout=int(st_ro["nwsr_max"].max()), ticks_run_per_robot=int(st_ro["ticks"].max()))
    if rank == 0 and not args.no_cpu_baseline:
        n_sample = 128 if ro else (2048 if spec["cold"] else 8192)
        v, secs, cores = cpu_reference(n_sample, spec["contacts"], spec["cold"], rollout=ro, switching_frac=spec.get("switching_frac", 0.0))
        line["cpu_baseline"] = dict(value=v, unit=UNIT, cores=cores, kind="port", reference_parts="qpOASES 3.2.2 solver: the real, unmodified vendored library (oracle/_ref); iDynTree subset, controller algebra, planner, plant: restated", seconds=secs,
                                    sample=f"{n_sample} robots of the same workload, {'%d closed-loop ticks from the cold start' % ro if ro else 'one pass'}: restated iDynTree/controller algebra{' + planner + plant' if ro else ''} (port) + the real vendored qpOASES 3.2.2 (reference), {cores} host threads")
    if rank == 0:
        print(json.dumps(line), flush=True)
    ctl.close()
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
