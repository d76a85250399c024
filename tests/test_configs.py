"""BASELINE.json's five configurations as parity cases, at sizes the oracle finishes in
seconds: same generators, qualities, windows and modes as the bench lines; every stream decodes with
Google's C decoder and the compressed size is held against the oracle's at the same shard size
(north_star: within 0.5 % in total; negative = smaller than the reference is allowed)."""
import numpy as np
import pytest

import _oracle as O
from brotli_lib_b200 import workloads as W
from conftest import load_input

pytestmark = pytest.mark.gpu


def _rt(stream, data):
    assert O.decode(stream, len(data), "system") == data
    if O.have_ref_decoder():
        assert O.decode(stream, len(data), "ref") == data


def _oracle_sharded(data, quality, lgwin, mode, shard):
    """The oracle over the same independent windows the GPU uses (parts framing included)."""
    return len(O.encode_part(data, b"", True, True, quality, lgwin, mode, shard=shard))


def test_config1_enc_ttf_q11_lgwin22_generic(gpu_ctx):
    data = load_input("enc-ttf.bin")          # bench/fixtures/enc-ttf.bin, 312 668 bytes: one shard
    assert len(data) == 312668
    s = gpu_ctx.encode(data, 11, 22, 0)
    _rt(s, data)
    ref = len(O.encode(data, 11, 22, 0))
    print(f"config 1: gpu {len(s)} oracle {ref} ({100.0 * (len(s) - ref) / ref:+.2f} %)")
    assert len(s) <= ref * 1.005


def test_config2_q5_lgwin22_mixed_segments(gpu_ctx):
    seg = 1 << 18                               # 8 segment types x 256 KiB, one shard each
    data = np.concatenate([W.mixed_segment(i, 0, seg) for i in range(8)]).tobytes()
    gpu_ctx.set_tuning(shard_bytes=seg)
    try:
        s = gpu_ctx.encode(data, 5, 22, 0)
    finally:
        gpu_ctx.set_tuning()
    _rt(s, data)
    ref = _oracle_sharded(data, 5, 22, 0, seg)
    print(f"config 2: gpu {len(s)} oracle {ref} ({100.0 * (len(s) - ref) / ref:+.2f} %)")
    assert len(s) <= ref * 1.005


def test_config3_q11_lgwin24_text(gpu_ctx):
    data = W.text(3 << 16, 0x7E570002).tobytes()
    s = gpu_ctx.encode(data, 11, 24, 1)        # TEXT == GENERIC (enc-constants.ts:121-126)
    _rt(s, data)
    assert s == gpu_ctx.encode(data, 11, 24, 0)
    ref = len(O.encode(data, 11, 24, 1))
    print(f"config 3: gpu {len(s)} oracle {ref} ({100.0 * (len(s) - ref) / ref:+.2f} %)")
    assert len(s) <= ref * 1.005


def test_config4_q9_lgwin24_sharded_over_ranks(gpu_ctx):
    """Config 4 is config 2's generator at q9 / lgwin 24, cut into per-GPU parts: the parts of 1, 2 and 4
    'ranks' (one context here) stitch to streams of the same size as the one-context stream."""
    from brotli_lib_b200 import sharded as S
    seg = 1 << 17
    data = np.concatenate([W.mixed_segment(i, 0x100, seg) for i in range(8)]).tobytes()
    gpu_ctx.set_tuning(shard_bytes=seg)
    try:
        whole = gpu_ctx.encode(data, 9, 24, 0)
        _rt(whole, data)
        for world in (2, 4):
            enc = lambda c, p, f, l, q, w, m: gpu_ctx.encode_part(c, p, f, l, q, w, m)
            s = S.stitch([S.encode_my_part(data, g, world, 9, 24, 0, enc, seg) for g in range(world)])
            _rt(s, data)
            assert len(s) == len(whole)        # shard boundaries already carry the part framing
    finally:
        gpu_ctx.set_tuning()
    ref = _oracle_sharded(data, 9, 24, 0, seg)
    print(f"config 4: gpu {len(whole)} oracle {ref} ({100.0 * (len(whole) - ref) / ref:+.2f} %)")
    assert len(whole) <= ref * 1.005


def test_config5_font_mode_q11_batch_of_glyf_blobs(gpu_ctx):
    """Sizes follow the config's own distribution (log-uniform 4..300 KiB), so the total is dominated by the
    larger blobs as it is in the benchmark.  Per blob the GPU's modelling (block split, clustering, context
    maps; the parse alone is at +0.00 % with trivial metablocks on both sides) trails the reference on the
    smallest blobs (4-10 KiB: +2..3 %); the gate the north star states is the total."""
    blobs = [b.tobytes() for b in W.font_blobs(12)]
    outs = gpu_ctx.encode_batch(blobs, 11, 22, 2)
    tg = to = 0
    for b, s in zip(blobs, outs):
        _rt(s, b)
        assert s == gpu_ctx.encode(b, 11, 22, 2)      # a batch item is brotliEncode of that item
        ref = len(O.encode(b, 11, 22, 2))
        print(f"  blob {len(b)}: gpu {len(s)} oracle {ref} ({100.0 * (len(s) - ref) / ref:+.2f} %)")
        tg += len(s); to += ref
        assert len(s) <= ref * 1.04 + 48
    print(f"config 5: gpu {tg} oracle {to} ({100.0 * (tg - to) / to:+.2f} %)")
    assert tg <= to * 1.005
    # FONT changes the distance alphabet (npostfix 1, ndirect 12): the streams differ from GENERIC ones
    assert outs[0] != gpu_ctx.encode(blobs[0], 11, 22, 0)
