"""Writes n 4 MiB shards of BASELINE config 2 (mixed corpus) to a file: input for tools/probe_bench."""
import sys; sys.path.insert(0, '.')
import numpy as np
from brotli_lib_b200 import workloads as W
n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
np.concatenate([W.mixed_segment(i % 8, i // 8) for i in range(n)]).tofile(sys.argv[1])
