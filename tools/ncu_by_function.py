#!/usr/bin/env python3
"""Instruction and sample share per device function of an ncu report (source page), functions located by scanning the
csrc headers for their definition lines.  usage: ncu_by_function.py rep"""
import bisect, collections, csv, os, re, subprocess, sys
rep = sys.argv[1]
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "trajoptkp_b200", "csrc")
starts = {}
for fn in os.listdir(root):
    if not fn.endswith((".cuh", ".cu", ".h")):
        continue
    L = []
    for i, line in enumerate(open(os.path.join(root, fn)), 1):
        m = re.match(r"^(?:TK_DEV|__global__|__device__|static|inline|template)?.*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;]*$", line)
        if m and not line.startswith((" ", "\t", "//", "#", "}")) and "(" in line:
            L.append((i, m.group(1)))
    starts[fn] = L
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None
smp = collections.Counter(); ins = collections.Counter()
for r in csv.reader(src.splitlines()):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split('/')[-1]; continue
    if len(r) < 10 or r[0] in ("Line No", ""): continue
    try: ln = int(r[0]); s = int(r[6]); ie = int(r[7])
    except ValueError: continue
    L = starts.get(cur, [])
    k = bisect.bisect_right([a for a, _ in L], ln) - 1
    name = "%s:%s" % (cur, L[k][1] if k >= 0 else "?")
    smp[name] += s; ins[name] += ie
ts, ti = sum(smp.values()), sum(ins.values())
print("samples %d warp-inst %d" % (ts, ti))
for name, s in smp.most_common(40):
    print("%5.1f%% smp %5.1f%% inst  %s" % (100 * s / ts, 100 * ins[name] / ti, name))
