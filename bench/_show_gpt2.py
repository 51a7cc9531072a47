import json
import sys
d = json.load(open(sys.argv[1]))
print(round(d["sum_kernel_ms_per_step"], 2), d["kernels_per_step"], round(d["graph_replay_ms_per_step"], 2))
for r in d["top"][:int(sys.argv[2]) if len(sys.argv) > 2 else 14]:
    print("  %7.3f %5.1f  %s" % (r["ms"], r["n"], r["name"][:105]))
