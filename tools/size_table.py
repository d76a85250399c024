#!/usr/bin/env python
"""GPU stream size vs the oracle (reference restatement), per input.  Run on a GPU box."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import brotli_lib_b200 as B
import _oracle as O
from brotli_lib_b200 import workloads as W
from conftest import load_input

q = int(sys.argv[1]) if len(sys.argv) > 1 else 5
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
items = {"text": W.mixed_segment(0, 0, n).tobytes(), "records": W.mixed_segment(4, 0, n).tobytes(), "repeats": W.mixed_segment(6, 0, n).tobytes(),
         "random": W.mixed_segment(7, 0, min(n, 1 << 18)).tobytes()}
for name in ["enc-ttf.bin", "alice29.txt", "mapsdatazrh", "html-content.bin", "compressed_repeated"]:
    items[name] = load_input(name)
ctx = B.Context()
tg = to = 0
for name, d in items.items():
    for mode in ((0, 2) if "ttf" in name else (0,)):
        t = time.time(); s = ctx.encode(d, q, 22, mode); dt = time.time() - t
        ok = O.decode(s, len(d)) == d
        ref = len(O.encode(d, q, 22, mode)); triv = len(O.encode(d, q, 22, mode, force_trivial=True))
        tg += len(s); to += ref
        print(f"{name:20s} mode {mode} n={len(d):8d} gpu={len(s):8d} oracle={ref:8d} trivial={triv:8d} delta={100.0 * (len(s) - ref) / ref:+.2f}% ok={ok} {dt*1e3:.1f} ms", flush=True)
print(f"TOTAL gpu={tg} oracle={to} delta={100.0 * (tg - to) / to:+.3f}%")
print("stages", ctx.timings())
