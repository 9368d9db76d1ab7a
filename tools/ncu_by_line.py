"""DEV: attribute an ncu SASS source page (ncu -i X.ncu-rep --page source --csv --print-source sass) to SOURCE LINES using the line
info of the library's cubin (nvdisasm -g): instructions executed and stall samples per file:line, top N.
usage: ncu_by_line.py sass.csv lib.so kernel_substring [topN]"""
import csv, collections, subprocess, sys, re, os, tempfile, glob
rep_csv, so, kernel = sys.argv[1], sys.argv[2], sys.argv[3]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
which = int(sys.argv[5]) if len(sys.argv) > 5 else 1      # 1-based index of the kernel header inside the report page
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
cubin = glob.glob(os.path.join(tmp, "*.cubin"))[0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
line_of = {}; cur = None; insec = False; fn_inline = None
for l in dis.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
    if m:
        insec = kernel in m.group(1); cur = None; continue
    if not insec: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m and cur: line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(rep_csv)))
agg = collections.defaultdict(lambda: [0, 0]); base = None; nk = 0
for r in rows:
    if r and r[0] == "Kernel Name": nk += 1; base = None; continue
    if r and r[0] == "Address": hdr = r; ix = {h: i for i, h in enumerate(hdr)}; continue
    if nk != which or not r or not r[0].startswith("0x"): continue
    a = int(r[0], 16)
    if base is None: base = a
    k = line_of.get(a - base, ("?", 0)); agg[k][0] += int(r[ix["Instructions Executed"]]); agg[k][1] += int(r[ix["# Samples"]])
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
src = {}
def text(f, n):
    if f not in src:
        p = os.path.join(os.path.dirname(os.path.abspath(so)), "csrc", f)
        src[f] = open(p).read().splitlines() if os.path.exists(p) else []
    return src[f][n - 1].strip()[:110] if 0 < n <= len(src[f]) else ""
print(f"total inst {ti} samples {ts}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:topn]:
    print(f"{v[0] / ti * 100:5.1f}% inst {v[1] / ts * 100:5.1f}% samp  {k[0]}:{k[1]:<5d} {text(*k)}")
