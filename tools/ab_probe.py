"""A/B of the two q5-9 probe kernels (BGPU_F_PROBE_WINDOW = 64: the old kernel): identical best[] on every input, then stage timings."""
import sys; sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, brotli_lib_b200 as B
from brotli_lib_b200 import workloads as W
from conftest import load_input, CORPUS, SYNTH
ctx = B.Context()
inputs = [(n, load_input(n)) for n in CORPUS + SYNTH] + [(f"seg{i}", W.mixed_segment(i, 0).tobytes()) for i in range(8)]
bad = 0
for name, data in inputs:
    for q in (5, 6, 7, 9):
        ctx.set_tuning(flags=0); a = ctx.stage_match_candidates(data, q, 22)
        ctx.set_tuning(flags=64 | 2048); b = ctx.stage_match_candidates(data, q, 22)
        same = all(np.array_equal(x, y) for x, y in zip(a, b))
        if not same:
            bad += 1; print("MISMATCH", name, q, [int((x != y).sum()) for x, y in zip(a, b)])
print("mismatches:", bad)
big = np.concatenate([W.mixed_segment(i % 8, i // 8) for i in range(64)])   # 256 MiB of config 2
for q in (5, 9):
    for flags in (0, 2048, 64 | 2048, 0):
        ctx.set_tuning(flags=flags, lanes=1)
        ctx.encode(big, q, 22, 0)
        t = dict(ctx.timings())
        print(f"q{q} flags={flags}: k_probe {t.get('k_probe'):.2f} ms per 256 MiB; sort {t.get('k_sort_pass0'):.2f} + {t.get('k_sort_pass1'):.2f}")
