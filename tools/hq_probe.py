"""GPU probe: q10/q11 streams round-trip and size vs the oracle, with stage timings."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import brotli_lib_b200 as B
import _oracle as O
from conftest import load_input, CORPUS, SYNTH
from brotli_lib_b200 import workloads as W
ctx = B.Context(0)
names = sys.argv[1:] or ["alice29.txt", "enc-ttf.bin", "html-content.bin", "mapsdatazrh", "quickfox_repeated", "abc1000", "zeros", "random_org_10k.bin", "font4k", "text1M", "rec1M", "mixed6"]
tg = to = 0
for name in names:
    if name == "text1M": data = W.text(1 << 20).tobytes()
    elif name == "rec1M": data = W.records(1 << 20, 3).tobytes()
    elif name == "mixed6": data = W.mixed_segment(6).tobytes()
    elif name == "text4M": data = W.text(1 << 22).tobytes()
    else: data = load_input(name)
    for q, mode in ((10, 0), (11, 0), (11, 2)):
        for flags in (2, 0):
            ctx.set_tuning(flags=flags)
            t = time.perf_counter()
            s = ctx.encode(data, q, 22, mode)
            dt = time.perf_counter() - t
            ok = O.decode(s, len(data)) == data
            ref = len(O.encode(data, q, 22, mode, force_trivial=bool(flags)))
            if flags == 0: tg += len(s); to += ref
            print(f"{name:20s} q{q} mode{mode} {'triv' if flags else 'full'} n={len(data)} gpu={len(s)} oracle={ref} delta={100.0*(len(s)-ref)/ref:+.2f}% rt={ok} {dt*1e3:.0f} ms", flush=True)
            if not ok: print("ROUNDTRIP FAILED"); 
    print("   stages", [(n, round(m, 2)) for n, m in ctx.timings()], flush=True)
print(f"TOTAL full: gpu={tg} oracle={to} delta={100.0*(tg-to)/to:+.3f}%")
