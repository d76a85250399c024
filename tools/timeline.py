"""Per-lane stage timeline of one default (4-lane) q5 encode of n shards of config 2, input resident on the device
(BGPU_TIMELINE=1 makes the library print when each stage of each group ended).  usage: timeline.py [shards] [flags]"""
import os, sys; sys.path.insert(0, '.')
os.environ["BGPU_TIMELINE"] = "1"
import numpy as np, torch, brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
ctx = B.Context()
big = np.concatenate([W.mixed_segment(i % 8, i // 8) for i in range(n)])
d = torch.zeros(big.size + 4096, dtype=torch.uint8, device="cuda"); d[: big.size] = torch.from_numpy(big).cuda()
torch.cuda.synchronize()
for flags in [int(a) for a in sys.argv[2:]] or [0]:
    ctx.set_tuning(flags=flags | 1)
    for rep in range(3):
        sys.stderr.write(f"--- flags {flags} rep {rep}\n"); sys.stderr.flush()
        size, ms, _ = ctx.encode_device(d.data_ptr(), big.size, 5, 22, 0)
        sys.stderr.write(f"kernel_ms {ms:.1f} size {size}\n")
