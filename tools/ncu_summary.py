#!/usr/bin/env python3
"""Summarise an ncu report: key raw metrics per kernel + hottest source lines.  usage: ncu_summary.py rep [nlines]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
nl = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keys = ['Kernel Name', 'gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct', 'smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct',
        'smsp__warp_issue_stalled_barrier_per_warp_active.pct', 'smsp__warp_issue_stalled_wait_per_warp_active.pct',
        'smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct', 'smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct',
        'smsp__warp_issue_stalled_branch_resolving_per_warp_active.pct', 'smsp__warp_issue_stalled_no_instruction_per_warp_active.pct',
        'smsp__warp_issue_stalled_not_selected_per_warp_active.pct', 'smsp__warp_issue_stalled_dispatch_stall_per_warp_active.pct',
        'smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct', 'smsp__warp_issue_stalled_membar_per_warp_active.pct',
        'smsp__warp_issue_stalled_sleeping_per_warp_active.pct', 'smsp__warp_issue_stalled_drain_per_warp_active.pct',
        'smsp__warp_issue_stalled_imc_miss_per_warp_active.pct', 'smsp__warp_issue_stalled_selected_per_warp_active.pct',
        'smsp__warp_issue_stalled_misc_per_warp_active.pct', 'smsp__warp_issue_stalled_tex_throttle_per_warp_active.pct']
for r in rows[2:]:
    for k in keys:
        if k in hdr:
            i = hdr.index(k)
            print("%-75s %-14s %s" % (k, units[i], r[i]))
    # warp stall reasons, cycles per issued instruction (the metric names differ between ncu versions)
    stalls = [(float(r[i].replace(",", "")), k) for i, k in enumerate(hdr)
              if "issue_stalled" in k and k.endswith("_per_issue_active.ratio") and "not_issued" not in k and r[i] not in ("", "n/a")]
    for v, k in sorted(stalls, reverse=True)[:10]:
        print("%-75s %-14s %.3f" % (k, "per issue", v))
    for k in ("sm__icc_request_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__warps_active.avg.per_cycle_active",
              "smsp__issue_active.avg.per_cycle_active"):
        if k in hdr:
            print("%-75s %-14s %s" % (k, units[hdr.index(k)], r[hdr.index(k)]))
    print("---")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
cur = None
agg = collections.Counter(); inst = collections.Counter(); tinst = collections.Counter(); lines = {}
for r in csv.reader(src.splitlines()):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split('/')[-1]
        continue
    if len(r) < 10 or r[0] in ("Line No", ""):
        continue
    try:
        ln = int(r[0]); s = int(r[6]); ie = int(r[7]); te = int(r[8])
    except ValueError:
        continue
    agg[(cur, ln)] += s; inst[(cur, ln)] += ie; tinst[(cur, ln)] += te; lines[(cur, ln)] = r[1]
tot = sum(agg.values()); toti = sum(inst.values())
byfile = collections.Counter()
for (f, l), s in agg.items():
    byfile[f] += s
print("samples", tot, "warp-inst", toti, [(f, round(100 * s / max(tot, 1), 1)) for f, s in byfile.most_common()])
for (f, l), s in agg.most_common(nl):
    print("%5.1f%% smp %5.1f%% inst  %-14s %4d thr/inst=%4.1f %s" % (100 * s / tot, 100 * inst[(f, l)] / toti, f, l,
                                                                    tinst[(f, l)] / max(inst[(f, l)], 1), lines[(f, l)][:100]))
