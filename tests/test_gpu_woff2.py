"""GPU tier (-m gpu): the WOFF2 glyf/loca transform kernels (csrc/woff2.cuh) against the CPU
restatement of W3C WOFF 2.0 section 5.1 (oracle/woff2_glyf.py): byte-identical transformed tables, and the transformed
tables invert to the original glyphs.  Then the config-5 pipeline: transform -> FONT-mode q11 batch -> decode."""
import os
import sys

import pytest

import _oracle as O
from conftest import ROOT, load_input

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import woff2_glyf as WG   # noqa: E402
import _fontgen as FG     # noqa: E402

pytestmark = pytest.mark.gpu


def test_real_font_byte_identical(gpu_ctx):
    from brotli_lib_b200 import woff2
    font = load_input("enc-ttf.bin")
    glyf, loca, ng, fmt = WG.glyf_inputs(font)
    assert woff2.sfnt_glyf_tables(font) == (glyf, loca, ng, fmt)
    t = woff2.transform_font(font, gpu_ctx)
    assert t == WG.transform(glyf, loca, ng, fmt)
    assert WG.inverse(t) == WG.parse_glyf(glyf, loca, ng, fmt)


def test_synthetic_batch_byte_identical(gpu_ctx):
    from brotli_lib_b200 import woff2
    tables = [FG.make_font(seed, 50 + 37 * seed, seed % 2, big=seed % 3 == 0) for seed in range(16)]
    outs = woff2.transform_glyf_batch(tables, gpu_ctx)
    assert len(outs) == len(tables)
    for tb, got in zip(tables, outs):
        assert got == WG.transform(*tb)
        assert WG.inverse(got) == WG.parse_glyf(*tb)


def test_malformed_tables_are_rejected(gpu_ctx):
    import brotli_lib_b200._lib as L
    from brotli_lib_b200 import woff2
    glyf, loca, ng, fmt = FG.make_font(3, 60, 1)
    with pytest.raises(L.BgpuError):
        woff2.transform_glyf_batch([(glyf[: len(glyf) // 2], loca, ng, fmt)], gpu_ctx)   # loca points past the table
    with pytest.raises(L.BgpuError):
        woff2.transform_glyf_batch([(glyf, loca[:20], ng, fmt)], gpu_ctx)                # loca too short


def test_config5_pipeline_transform_then_font_batch(gpu_ctx):
    """Fonts -> transformed glyf blobs (device) -> brotliEncode FONT q11 batch (device) -> Google's decoder -> inverse."""
    from brotli_lib_b200 import woff2
    tables = [WG.glyf_inputs(load_input("enc-ttf.bin"))] + [FG.make_font(100 + s, 300, 1) for s in range(5)]
    blobs = woff2.transform_glyf_batch(tables, gpu_ctx)
    streams = gpu_ctx.encode_batch(blobs, 11, 22, 2)
    for tb, blob, s in zip(tables, blobs, streams):
        back = O.decode(s, len(blob))
        assert back == blob
        assert WG.inverse(back) == WG.parse_glyf(*tb)
    raw = gpu_ctx.encode(tables[0][0] + tables[0][1], 11, 22, 2)
    print(f"enc-ttf glyf+loca: raw {len(tables[0][0]) + len(tables[0][1])} -> {len(raw)} B; transformed {len(blobs[0])} -> {len(streams[0])} B")
    assert len(streams[0]) < len(raw)          # the point of the transform: the same outlines compress better
