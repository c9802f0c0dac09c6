#!/usr/bin/env python
"""Executed SASS opcode histogram of every kernel in an ncu report (source page, SASS view): share of the executed warp
instructions per opcode and the average number of active threads.  usage: ncu_ops.py report.ncu-rep [top]"""
import collections, csv, re, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 24
# one source page per kernel name (without a filter ncu repeats the first result)
raw = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
names = []
for r in raw[2:]:
    n = r[raw[0].index("Kernel Name")].split("(")[0].split("::")[-1].split("<")[0]
    if n not in names:
        names.append(n)
rows = []
for n in names:
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + n], capture_output=True, text=True).stdout
    part = list(csv.reader(out.splitlines()))
    # keep the first kernel section only
    k = [i for i, r in enumerate(part) if r and r[0] == "Kernel Name"]
    rows += part[k[0]:k[1]] if len(k) > 1 else part[k[0]:] if k else []
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]
        h = rows[i + 1]
        ie, te, src = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("Source")
        ops, thr = collections.Counter(), collections.Counter()
        j = i + 2
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            r = rows[j]
            if len(r) > max(ie, te):
                m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)", r[src])
                op = m.group(2) if m else "?"
                ops[op] += int(r[ie]); thr[op] += int(r[te])
            j += 1
        tot = sum(ops.values())
        print("==", name[:110])
        print(f"   executed warp instructions: {tot}, static SASS instructions: {j - i - 2}")
        for op, c in ops.most_common(top):
            print(f"   {op:10s} {100 * c / tot:5.1f} %   avg active threads {thr[op] / max(c, 1):5.1f}")
        i = j
    else:
        i += 1
