#!/usr/bin/env python3
"""Static SASS size of one kernel instantiation by source function (needs -lineinfo).
usage: sass_by_function.py obj.o kernel-substring [instantiation-substring]"""
import bisect, collections, os, re, subprocess, sys, tempfile
obj, kname = sys.argv[1], sys.argv[2]
inst = sys.argv[3] if len(sys.argv) > 3 else ""
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
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=d, capture_output=True)
cub = [os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
cur = None; fn = None
cnt = collections.Counter()
for line in sass.splitlines():
    m = re.match(r"\s*\.section\s+\.text\.(\S+)", line)
    if m: fn = m.group(1); continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", line) and fn and kname in fn and inst in fn and cur:
        cnt[cur] += 1
byfn = collections.Counter()
for (f, l), c in cnt.items():
    L = starts.get(f, [])
    k = bisect.bisect_right([a for a, _ in L], l) - 1
    byfn["%s:%s" % (f, L[k][1] if k >= 0 else "?")] += c
tot = sum(byfn.values())
print("total SASS instructions", tot, "=", tot * 16 // 1024, "KB")
for k, v in byfn.most_common(40):
    print("%6d %5.1f%%  %s" % (v, 100.0 * v / tot, k))
