"""GPU tier (-m gpu): the round-1 kernels kept behind tuning flags must produce the bytes of their replacements.

  BGPU_F_PROBE_WINDOW | BGPU_F_SORT_UNPACKED   k_probe_sorted + (position, key bytes) sort entries
  BGPU_F_SORT_UNPACKED                         k_probe_diag on unpacked entries
  BGPU_F_WRITER_TWO_PASS                       k_tilebits2 + k_write2 for every tile
against the defaults (packed sort entries at q5-6, lane-per-entry probe, k_tile_pack / k_tile_copy with the dense
by-kind token path).  BGPU_F_SPLIT_3ITER changes the model (three refinement iterations): valid stream, size within 0.2 %."""
import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu

PROBE_WINDOW, WRITER_TWO_PASS, SPLIT_3ITER, SORT_UNPACKED = 64, 256, 1024, 2048


def _inputs():
    from brotli_lib_b200 import workloads as W
    yield "mixed", np.concatenate([W.mixed_segment(i, 0, 700000) for i in (1, 4, 6, 7)]).tobytes()
    yield "text", W.text(1500000, 0x7E570002).tobytes()
    yield "font", b"".join(b.tobytes() for b in W.font_blobs(6, seed=0xF0470003, lo=60000, hi=200000))


@pytest.mark.parametrize("quality", [5, 6, 9])
def test_flagged_kernel_generations_give_identical_streams(gpu_ctx, quality):
    try:
        for name, data in _inputs():
            gpu_ctx.set_tuning(flags=0)
            base = gpu_ctx.encode(data, quality, 22, 0)
            assert O.decode(base, len(data)) == data, name
            for flags in (SORT_UNPACKED, PROBE_WINDOW | SORT_UNPACKED, WRITER_TWO_PASS, WRITER_TWO_PASS | PROBE_WINDOW | SORT_UNPACKED):
                gpu_ctx.set_tuning(flags=flags)
                assert gpu_ctx.encode(data, quality, 22, 0) == base, (name, quality, flags)
            gpu_ctx.set_tuning(flags=SPLIT_3ITER)
            s3 = gpu_ctx.encode(data, quality, 22, 0)
            assert O.decode(s3, len(data)) == data, name
            assert abs(len(s3) - len(base)) <= 0.002 * len(base), (name, len(s3), len(base))
    finally:
        gpu_ctx.set_tuning(flags=0)


def test_two_pass_writer_at_q11(gpu_ctx):
    from brotli_lib_b200 import workloads as W
    data = W.text(600000, 0x7E570009).tobytes()
    try:
        gpu_ctx.set_tuning(flags=0)
        base = gpu_ctx.encode(data, 11, 22, 0)
        gpu_ctx.set_tuning(flags=WRITER_TWO_PASS)
        assert gpu_ctx.encode(data, 11, 22, 0) == base
        assert O.decode(base, len(data)) == data
    finally:
        gpu_ctx.set_tuning(flags=0)
