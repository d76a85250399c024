"""One-lane stage timings (CUDA events of bgpu_stage_timings) of one q5 job over n shards of config 2."""
import sys; sys.path.insert(0, '.')
import numpy as np, brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
q = int(sys.argv[2]) if len(sys.argv) > 2 else 5
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
ctx = B.Context()
big = W.text(n << 22) if len(sys.argv) > 4 and sys.argv[4] == "text" else np.concatenate([W.mixed_segment(i % 8, i // 8) for i in range(n)])
ctx.set_tuning(lanes=1, flags=flags)
for rep in range(3):
    out = ctx.encode(big, q, 24 if q >= 10 else 22, 0)
t = ctx.timings()
print(f"n={n} shards q{q}: out {len(out)} bytes; total {sum(ms for _, ms in t):.2f} ms")
print("  " + "  ".join(f"{k} {ms:.2f}" for k, ms in t))
