"""A/B of tuning flags on one q5 job (device-resident, default lanes): same compressed bytes, kernel ms.
usage: ab_flags.py shards flagsA flagsB ..."""
import sys, hashlib; sys.path.insert(0, '.')
import numpy as np, torch, brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
n = int(sys.argv[1]); flags = [int(a) for a in sys.argv[2:]]
ctx = B.Context()
big = np.concatenate([W.mixed_segment(i % 8, i // 8) for i in range(n)])
d = torch.zeros(big.size + 4096, dtype=torch.uint8, device="cuda"); d[: big.size] = torch.from_numpy(big).cuda()
torch.cuda.synchronize()
for rnd in range(2):
    for f in flags:
        ctx.set_tuning(flags=f)
        best = 1e9
        for rep in range(3):
            size, ms, out = ctx.encode_device(d.data_ptr(), big.size, 5, 22, 0, fetch=(rep == 0 and rnd == 0))
            if out is not None: h = hashlib.md5(out).hexdigest()
            best = min(best, ms)
        print(f"flags {f}: {best:.2f} ms  size {size}  md5 {h if rnd == 0 else ''}", flush=True)
