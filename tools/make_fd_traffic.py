#!/usr/bin/env python3
"""profiles/fd_traffic.json from ncu --set full captures of the FD world kernels of ONE FD launch of the default bench
command (tools/capture_profiles.sh): dram__bytes_read.sum + dram__bytes_write.sum, summed over the kernels that make up
the launch.  The entry carries the hash of the kernel sources; bench.py reports `roofline.traffic` only while the
sources still have that hash (a capture describes the build it was taken on).
usage: make_fd_traffic.py task batch horizon kp rep-or-raw-page-csv [...]"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

task, batch, horizon, kp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
total, parts = 0.0, []
for rep in sys.argv[5:]:
    raw = open(rep).read() if rep.endswith(".csv") else \
        subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        b = 0.0
        for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(key)
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
            b += float(r[i].replace(",", "")) * scale
        parts.append({"kernel": r[hdr.index("Kernel Name")], "dram_bytes": b,
                      "ms": float(r[hdr.index("gpu__time_duration.sum")].replace(",", "")) *
                      {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[units[hdr.index("gpu__time_duration.sum")]]})
        total += b
path = os.path.join(ROOT, "profiles", "fd_traffic.json")
try:
    out = json.load(open(path))
    if "task" in out:      # round-1 single-entry layout
        out = {}
except Exception:
    out = {}
out[task] = {"batch": batch, "horizon": horizon, "kp": kp, "dram_bytes_per_launch": total, "kernels": parts,
             "source_hash": bench.source_hash(), "captures": [os.path.basename(r) for r in sys.argv[5:]]}
json.dump(out, open(path, "w"), indent=1)
print(json.dumps(out[task], indent=1))
