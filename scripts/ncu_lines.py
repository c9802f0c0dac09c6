#!/usr/bin/env python
"""Attribute executed SASS instructions of an ncu report to source lines (ncu's CSV export of the CUDA view carries no
metrics).  usage: ncu_lines.py report.ncu-rep object.o kernel_substring [top]"""
import collections, csv, os, re, subprocess, sys, tempfile
rep, obj, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, check=True, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], cwd=d, capture_output=True, text=True).stdout
lines, cur, insec = [], None, False
for ln in dis.splitlines():
    if ln.startswith("//---") and ".text." in ln:
        insec = kern in ln
        continue
    if not insec:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(sass.splitlines()))
h = rows[1]; ie = h.index("Instructions Executed"); isamp = h.index("# Samples")
data = rows[2:]
assert len(data) == len(lines), (len(data), len(lines))
agg = collections.Counter(); samp = collections.Counter()
for r, l in zip(data, lines):
    agg[l] += int(r[ie]); samp[l] += int(r[isamp])
tot = sum(agg.values()); ts = sum(samp.values())
src = {}
for (f, n), c in agg.most_common(top):
    if f not in src:
        p = os.path.join(os.path.dirname(os.path.abspath(obj)), f)
        src[f] = open(p).read().splitlines() if os.path.exists(p) else []
print(f"total warp instructions {tot}, samples {ts}")
for (f, n), c in sorted(agg.most_common(top), key=lambda kv: (kv[0][0], kv[0][1])):
    text = src[f][n - 1].strip()[:100] if src[f] and n <= len(src[f]) else ""
    print(f"{f}:{n:<4} inst {100*c/tot:5.1f}%  samples {100*samp[(f,n)]/ts:5.1f}%  {text}")
