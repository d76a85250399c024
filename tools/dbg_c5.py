import sys; sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import _oracle as O
import brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
ctx = B.Context()
blobs = [b.tobytes() for b in W.font_blobs(10, lo=4096, hi=40000)]
for q in (11, 10, 9, 5):
    for mode in (2, 0):
        tg = to = 0
        for b in blobs:
            s = ctx.encode(b, q, 22, mode); ref = len(O.encode(b, q, 22, mode))
            tg += len(s); to += ref
            if q == 11 and mode == 2: print(len(b), len(s), ref, f"{100.0*(len(s)-ref)/ref:+.2f}%")
        print("q", q, "mode", mode, tg, to, f"{100.0*(tg-to)/to:+.2f}%")
ctx.set_tuning(flags=2)
for q in (11,):
    tg = to = 0
    for b in blobs:
        s = ctx.encode(b, q, 22, 2); ref = len(O.encode(b, q, 22, 2, force_trivial=True)); tg += len(s); to += ref
    print("trivial q", q, tg, to, f"{100.0*(tg-to)/to:+.2f}%")
