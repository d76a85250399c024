"""GPU probe: H10 stage timings on one 4 MiB shard of each workload type."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
ctx = B.Context(0)
for name, data in (("text4M", W.text(1 << 22)), ("records4M", W.records(1 << 22, 3)), ("mixed6", W.mixed_segment(6)), ("rand1M", W.xorshift_bytes(1 << 20, 5))):
    for rep in range(2):
        t = time.perf_counter()
        off, ds, ln = ctx.stage_h10_matches(data, 22)
        dt = time.perf_counter() - t
    print(name, len(data), "matches/pos %.2f" % (len(ds) / len(data)), "wall %.3f s" % dt, ctx.timings(), flush=True)
