"""Stage timings of the config-5 batch (2000 FONT blobs, q11) through bgpu_encode_batch, second call."""
import sys, time; sys.path.insert(0, '.')
import brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ctx = B.Context(); ctx.set_tuning(flags=flags)
blobs = [b.tobytes() for b in W.font_blobs(n)]
for rep in range(3):
    t = time.perf_counter(); outs = ctx.encode_batch(blobs, 11, 22, 2); dt = time.perf_counter() - t
    tm = ctx.timings()
    print(f"rep {rep}: {dt*1e3:.1f} ms wall, stages sum {sum(ms for _, ms in tm):.1f} ms, out {sum(len(o) for o in outs)}")
print("  " + "  ".join(f"{k} {ms:.2f}" for k, ms in tm))
