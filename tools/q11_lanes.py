"""q11 (BASELINE config 3: text, lgwin 24) over n MiB, device-resident: one lane vs lanes x group sizes.
usage: q11_lanes.py [MiB] ; prints kernel ms and size per setting"""
import sys; sys.path.insert(0, '.')
import numpy as np, torch, brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 256
ctx = B.Context()
big = W.text(mb << 20)
d = torch.zeros(big.size + 4096, dtype=torch.uint8, device="cuda"); d[: big.size] = torch.from_numpy(big).cuda()
torch.cuda.synchronize()
for lanes, gs in [(1, 0), (2, 32), (2, 16), (4, 16), (3, 22), (4, 8), (2, 8), (1, 0)]:
    ctx.set_tuning(flags=1, lanes=lanes, group_shards=gs)
    best = None
    for rep in range(2):
        size, ms, _ = ctx.encode_device(d.data_ptr(), big.size, 11, 24, 0)
        best = ms if best is None else min(best, ms)
    print(f"lanes {lanes} group {gs}: {best:.1f} ms  ({big.size / best / 1e6:.3f} GB/s) size {size}", flush=True)
