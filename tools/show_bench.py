import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%.0f Jac/s  %.2f ms  fd %.2f ms x%d" % (d["value"], d["ms_per_step"], d["fd_kernel"]["avg_ms"], d["fd_kernel"]["launches"]), d["phase_ms"])
