"""Per-shard size vs the oracle on the 8 config-2 shard types for a list of tuning flags (q5, q9)."""
import sys, time; sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np, brotli_lib_b200 as B, _oracle as O
from brotli_lib_b200 import workloads as W
flags_list = [int(a) for a in sys.argv[1:]] or [0]
ctx = B.Context()
sample = np.concatenate([W.mixed_segment(i, 0) for i in range(8)])
items = [sample[o:o + (1 << 22)] for o in range(0, sample.size, 1 << 22)]
for q in (5, 9):
    ref = [int(x) for x in O.encode_shards_sizes(sample, 1 << 22, q, 22, 0, 16)]
    for flags in flags_list:
        ctx.set_tuning(flags=flags)
        outs = ctx.encode_batch(items, q, 22, 0)
        g = [len(s) for s in outs]
        ok = all(O.decode(s, 1 << 22) == it.tobytes() for s, it in zip(outs[4:7], items[4:7]))
        print(f"q{q} flags={flags}: total {100*(sum(g)-sum(ref))/sum(ref):+.3f}% per-shard " + " ".join(f"{100*(a-b)/b:+.2f}" for a, b in zip(g, ref)) + f" ok={ok}", flush=True)
