"""q11 (config 3 text, 256 MiB): device-resident vs host-input call, with the lane timeline of the last host call."""
import os, sys, time; sys.path.insert(0, '.')
import numpy as np, torch, brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
ctx = B.Context()
big = W.text(256 << 20)
d = torch.zeros(big.size + 4096, dtype=torch.uint8, device="cuda"); d[: big.size] = torch.from_numpy(big).cuda()
hp = torch.from_numpy(big).pin_memory()
torch.cuda.synchronize()
for rep in range(2):
    size, ms, _ = ctx.encode_device(d.data_ptr(), big.size, 11, 24, 0)
    print(f"device-resident: {ms:.1f} ms", flush=True)
for rep in range(2):
    if rep == 1: os.environ["BGPU_TIMELINE"] = "1"
    t = time.perf_counter(); out = ctx.encode(hp.numpy(), 11, 24, 0); dt = time.perf_counter() - t
    print(f"host input (pinned): {dt*1e3:.1f} ms wall, {len(out)} bytes", flush=True)
