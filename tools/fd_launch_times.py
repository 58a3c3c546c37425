#!/usr/bin/env python3
"""Per-launch table of an `ncu --metrics gpu__time_duration.sum,... --csv` log.  usage: fd_launch_times.py log.csv [last N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 14
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]; ki = H.index('Kernel Name'); vi = H.index('Metric Value'); mi = H.index('Metric Name'); ii = H.index('ID')
cur, order = {}, []
for r in rows[hdr + 1:]:
    if len(r) <= vi: continue
    key = (r[ii], r[ki][:42])
    if key not in cur: cur[key] = {}; order.append(key)
    cur[key][r[mi]] = r[vi]
for k in order[-n:]:
    d = cur[k]
    print("%-44s %10.1f us  thr/inst %5s  grid %6s  smem %s" % (k[1], float(d.get('gpu__time_duration.sum', '0').replace(',', '')) / 1e3,
          d.get('smsp__thread_inst_executed_per_inst_executed.ratio'), d.get('launch__grid_size'), d.get('launch__shared_mem_per_block_dynamic')))
