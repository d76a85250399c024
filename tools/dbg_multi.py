import sys, time, threading; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import numpy as np, ctypes as C, torch
import brotli_lib_b200 as B
from bench import gen_data
n = 512 << 20
host = gen_data("mixed", n, 0)
def run(devs, data, q=9, reps=2, label=""):
    ctx = B.Context(devices=devs) if len(devs) > 1 else B.Context(devs[0])
    L = ctx._L; cap = int(L.bgpu_max_compressed_size(data.size)); out = np.empty(cap, np.uint8); ln = C.c_size_t(0)
    for r in range(reps + 1):
        t = time.perf_counter()
        rc = L.bgpu_encode(ctx._h, data.ctypes.data, data.size, q, 24, 0, out.ctypes.data, cap, C.byref(ln))
        dt = time.perf_counter() - t
        assert rc == 0
        if r: print(label, devs, f"{data.size>>20} MiB q{q}: {dt*1e3:.1f} ms  slowest-device encode {ctx.counter(5)/1e3:.1f} ms  stages", [(k, round(v,1)) for k,v in ctx.timings() if v > 5][:8])
    ctx.close()
run([0], host[:256<<20], label="single")
run([0, 1], host, label="multi ")
run([0, 0], host, label="multi-same")
# two independent contexts from two python threads
def worker(dev, data, res):
    ctx = B.Context(dev); L = ctx._L; cap = int(L.bgpu_max_compressed_size(data.size)); out = np.empty(cap, np.uint8); ln = C.c_size_t(0)
    L.bgpu_encode(ctx._h, data.ctypes.data, data.size, 9, 24, 0, out.ctypes.data, cap, C.byref(ln))
    t = time.perf_counter(); L.bgpu_encode(ctx._h, data.ctypes.data, data.size, 9, 24, 0, out.ctypes.data, cap, C.byref(ln)); res[dev] = time.perf_counter() - t
    ctx.close()
res = {}
ths = [threading.Thread(target=worker, args=(d, host[d*(256<<20):(d+1)*(256<<20)], res)) for d in (0, 1)]
t = time.perf_counter(); [x.start() for x in ths]; [x.join() for x in ths]
print("two python threads, one context each:", {k: round(v*1e3,1) for k,v in res.items()})
run([0, 1], host, q=5, label="multi q5")
run([0], host[:256<<20], q=5, label="single q5")
