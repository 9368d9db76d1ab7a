"""DEV: attribute an ncu SASS source page (ncu -i X.ncu-rep --page source --csv --print-source sass) to the device functions of the
library (symbol offsets from `cuobjdump -elf`): instructions executed, stall samples and the top stall reasons per function."""
import csv, collections, subprocess, sys, re, os
rep_csv, so, kernel = sys.argv[1], sys.argv[2], sys.argv[3]
which = int(sys.argv[4]) if len(sys.argv) > 4 else 1      # 1-based index of the kernel inside the report
elf = subprocess.run(["cuobjdump", "-elf", so], capture_output=True, text=True).stdout
syms = []
for line in elf.splitlines():
    m = re.match(r"\s*0x[0-9a-f]+\s+(0x[0-9a-f]+|0)\s+(0x[0-9a-f]+|0)\s+0x2\s+\S+\s+\S+\s+\$(\S+?)\$(\S+)", line)
    if m and kernel in m.group(3):
        name = subprocess.run(["c++filt", m.group(4)], capture_output=True, text=True).stdout.strip().split("(")[0]
        syms.append((int(m.group(1), 16), int(m.group(2), 16), name))
syms.sort()
def fn(off):
    for a, s, n in syms:
        if a <= off < a + s: return n
    return "kernel body / intrinsics"
rows = list(csv.reader(open(rep_csv)))
out = collections.defaultdict(lambda: collections.Counter()); base = None; nk = 0
for r in rows:
    if r and r[0] == "Kernel Name":
        nk += 1; base = None; continue
    if nk != which or not r or not r[0].startswith("0x"): 
        if r and r[0] == "Address": hdr = r; ix = {h: i for i, h in enumerate(hdr)}
        continue
    a = int(r[0], 16)
    if base is None: base = a
    f = fn(a - base); c = out[f]
    c["inst"] += int(r[ix["Instructions Executed"]]); c["samples"] += int(r[ix["# Samples"]]); c["static"] += 1
    for h in hdr:
        if h.startswith("stall_") and "Not Issued" not in h: c[h] += int(r[ix[h]] or 0)
    op = r[ix["Source"]].split(); op = (op[1] if op[0].startswith("@") else op[0]).split(".")[0]
    if op in ("DFMA", "DMUL", "DADD"): c["fp64"] += int(r[ix["Instructions Executed"]])
    if op in ("LDS", "STS"): c["lds"] += int(r[ix["Instructions Executed"]])
ti = sum(c["inst"] for c in out.values()); ts = sum(c["samples"] for c in out.values())
print(f"{'function':34s} {'static':>7s} {'inst%':>6s} {'samp%':>6s} {'fp64%':>6s} {'lds%':>5s}  top stalls")
for f, c in sorted(out.items(), key=lambda kv: -kv[1]["samples"]):
    st = sorted(((k, v) for k, v in c.items() if k.startswith("stall_")), key=lambda kv: -kv[1])[:3]
    print(f"{f[:34]:34s} {c['static']:7d} {c['inst'] / ti * 100:6.1f} {c['samples'] / ts * 100:6.1f} {c['fp64'] / max(1, c['inst']) * 100:6.1f} {c['lds'] / max(1, c['inst']) * 100:5.1f}  "
          + " ".join(f"{k[6:]}={v / max(1, c['samples']) * 100:.0f}%" for k, v in st))
print("total static", sum(c["static"] for c in out.values()), "inst", ti, "samples", ts)
