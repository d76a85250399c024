"""Lane / group sweep of one q5 job (device-resident): usage: ab_lanes.py shards"""
import sys; sys.path.insert(0, '.')
import numpy as np, torch, brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
n = int(sys.argv[1]); fl = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ctx = B.Context()
big = np.concatenate([W.mixed_segment(i % 8, i // 8) for i in range(n)])
d = torch.zeros(big.size + 4096, dtype=torch.uint8, device="cuda"); d[: big.size] = torch.from_numpy(big).cuda()
torch.cuda.synchronize()
for lanes, gs in [(4, 0), (4, 32), (8, 32), (6, 43), (8, 16), (5, 52), (4, 0)]:
    ctx.set_tuning(flags=1 | fl, lanes=lanes, group_shards=gs)
    best = 1e9
    for rep in range(3):
        size, ms, _ = ctx.encode_device(d.data_ptr(), big.size, 5, 22, 0)
        best = min(best, ms)
    print(f"lanes {lanes} group {gs}: {best:.2f} ms size {size}", flush=True)
