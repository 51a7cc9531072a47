#!/usr/bin/env python
"""Where does a service-path round spend its host time?  Runs ``service_bench.service_rounds`` with a sampling thread
that records, every 2 ms, the innermost repository frame (and the leaf function) of every other thread.  Development
tool: the output is a histogram, not a benchmark number."""
import collections
import json
import os
import sys
import threading
import time
import traceback

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import service_bench as sb  # noqa: E402

counts, stop = collections.Counter(), False


def sampler():
    me = threading.get_ident()
    while not stop:
        for tid, fr in sys._current_frames().items():
            if tid == me:
                continue
            st = traceback.extract_stack(fr)
            chain = [f for f in st if "metisfl_b200" in f.filename or "service_bench" in f.filename]
            if chain:
                key = " < ".join(f"{os.path.basename(f.filename)}:{f.lineno}:{f.name}" for f in reversed(chain[-3:]))
                counts[(key, st[-1].name)] += 1
        time.sleep(0.002)


if __name__ == "__main__":
    model = sys.argv[1] if len(sys.argv) > 1 else "resnet18"
    t = threading.Thread(target=sampler, daemon=True)
    t.start()
    r = sb.service_rounds(True, 20, 3, 4, 128, model)
    stop = True
    print(json.dumps(r))
    skip = ("_snapshot_loop", "wait_servicer", "monitor_federation")
    for (k, leaf), v in counts.most_common(60):
        if not any(s in k for s in skip):
            print(v, k, "|", leaf)
