"""DEV: condense `ncu --set full` reports into the figures DESIGN.md and bench.py quote.
usage: ncu_summary.py out.json label=report.ncu-rep[:launch_index[:kernel_substring]] ...
Per capture: duration, launch geometry, warp instructions, issue rate, cycles per issued instruction and its stall split, instruction-cache figures, FP64 pipe,
shared-memory data pipe, DRAM bytes.  The JSON is what bench.py reads (`profiles/r02_ncu_summary.json`, keys = workload names); a Markdown table goes to stdout."""
import csv, io, json, subprocess, sys

KEYS = {
    "duration_ms": ("gpu__time_duration.sum", 1.0),          # ncu prints ms in this page's units row -- checked below
    "grid": ("launch__grid_size", 1), "block": ("launch__block_size", 1), "regs": ("launch__registers_per_thread", 1),
    "smem_dyn_kb": ("launch__shared_mem_per_block_dynamic", 1.0),
    "warp_inst": ("smsp__inst_executed.sum", 1), "warps_per_sm": ("sm__warps_active.avg.per_cycle_active", 1.0),
    "issue_pct": ("smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0), "cyc_per_inst": ("smsp__average_warp_latency_per_inst_issued.ratio", 1.0),
    "icc_hit_pct": ("sm__icc_request_hit_rate.pct", 1.0), "gcc_inst_req_pct": ("gcc__cache_requests_type_instruction.sum.pct_of_peak_sustained_elapsed", 1.0),
    "fp64_pipe_pct": ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", 1.0),
    "smem_wavefronts": ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", 1), "smem_pipe_pct": ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", 1.0),
    "dram_read": ("dram__bytes_read.sum", 1.0), "dram_write": ("dram__bytes_write.sum", 1.0), "lanes_per_inst": ("smsp__thread_inst_executed_per_inst_executed.ratio", 1.0),
}
UNIT_SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3, "msecond": 1.0, "usecond": 1e-3, "nsecond": 1e-6, "second": 1e3}


def read(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    return rows[0], rows[1], rows[2:]


def num(v):
    try:
        return float(v.replace(",", ""))
    except Exception:
        return None


def capture(rep, idx=0, kern=None):
    hdr, units, rows = read(rep)
    if kern:
        rows = [r for r in rows if kern in r[hdr.index("Kernel Name")]]
    r = rows[idx]; d = dict(zip(hdr, r)); u = dict(zip(hdr, units))
    out = {"kernel": d["Kernel Name"].split("(")[0], "report": rep.split("/")[-1], "launch_index": idx}
    for k, (m, _) in KEYS.items():
        v = num(d.get(m, ""))
        if v is not None and u.get(m) in UNIT_SCALE and (k.startswith("dram") or k == "duration_ms" or k == "smem_dyn_kb"):
            v *= UNIT_SCALE[u[m]]
            if k == "smem_dyn_kb":
                v /= 1e3
        out[k] = v
    st = {}
    for h in hdr:
        if "issue_stalled" in h and "per_issue_active" in h and "not_issued" not in h:
            v = num(d[h])
            if v and v > 0.04:
                st[h.split("stalled_")[1].split("_per_issue")[0]] = round(v, 2)
    out["stall_cycles_per_issue"] = dict(sorted(st.items(), key=lambda kv: -kv[1]))
    out["dram_bytes"] = (out["dram_read"] or 0) + (out["dram_write"] or 0)
    return out


if __name__ == "__main__":
    res = {}
    for a in sys.argv[2:]:
        label, spec = a.split("=", 1); parts = spec.split(":")
        res[label] = capture(parts[0], int(parts[1]) if len(parts) > 1 else 0, parts[2] if len(parts) > 2 else None)
        res[label]["source"] = f"profiles/r02_ncu_summary.md ({res[label]['report']}, launch {res[label]['launch_index']})"
    json.dump(res, open(sys.argv[1], "w"), indent=1)
    cols = list(res)
    print("| metric | " + " | ".join(cols) + " |"); print("|---|" + "---|" * len(cols))
    def row(name, f):
        print(f"| {name} | " + " | ".join(f(res[c]) for c in cols) + " |")
    row("kernel", lambda r: r["kernel"])
    row("duration under ncu (ms)", lambda r: f"{r['duration_ms']:.3f}")
    row("grid x block, dynamic smem per CTA, registers", lambda r: f"{int(r['grid'])} x {int(r['block'])}, {r['smem_dyn_kb']:.1f} KB, {int(r['regs'])}")
    row("resident warps per SM", lambda r: f"{r['warps_per_sm']:.2f}")
    row("warp instructions", lambda r: f"{r['warp_inst'] / 1e6:.1f} M")
    row("active lanes per instruction", lambda r: f"{r['lanes_per_inst']:.1f}")
    row("smsp issue active", lambda r: f"{r['issue_pct']:.1f} %")
    row("cycles per issued instruction, per warp", lambda r: f"{r['cyc_per_inst']:.2f}: " + ", ".join(f"{k} {v}" for k, v in list(r["stall_cycles_per_issue"].items())[:7]))
    row("sm__icc_request_hit_rate", lambda r: f"{r['icc_hit_pct']:.1f} %")
    row("gcc instruction requests, % of peak", lambda r: f"{r['gcc_inst_req_pct']:.1f} %")
    row("FP64 pipe active (elapsed)", lambda r: f"{r['fp64_pipe_pct']:.2f} %")
    row("shared-memory data pipe", lambda r: f"{r['smem_pipe_pct']:.1f} % ({r['smem_wavefronts'] / 1e6:.0f} M wavefronts)")
    row("dram read / write", lambda r: f"{r['dram_read'] / 1e6:.1f} MB / {r['dram_write'] / 1e6:.1f} MB")
