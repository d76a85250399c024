"""GPU tier (-m gpu): the on-device Brotli decoder (csrc/decode.cuh).

bgpu_decode (one device thread per stream) is checked against Google's C decoder on streams of this library, of the
oracle (the reference's encoder restated) and of Google's C encoder, and on the reference's own decoder vectors
(test/fixtures/vectors/*.compressed, copied under tests/golden/vectors): it must return exactly the bytes the C
decoder returns, or refuse (static-dictionary references are not supported) — never different bytes.
BGPU_F_VERIFY runs the same decoder shard-parallel over a job that is still on the device."""
import glob
import os

import numpy as np
import pytest

import _oracle as O
from conftest import CORPUS, SYNTH, ROOT, load_input

pytestmark = pytest.mark.gpu

VERIFY = 16   # BGPU_F_VERIFY


def _gpu_decode_or_refuse(ctx, stream, expected):
    import brotli_lib_b200._lib as L
    try:
        got = ctx.decode(stream, len(expected))
    except L.BgpuError as e:
        assert e.code == -9, e
        return False
    assert got == expected
    return True


@pytest.mark.parametrize("name", CORPUS + SYNTH)
def test_decodes_own_and_oracle_streams(gpu_ctx, name):
    data = load_input(name)
    if len(data) > 400000:
        data = data[:400000]
    for q, mode in ((0, 0), (1, 0), (3, 2), (5, 0), (9, 2), (11, 0)):
        s = gpu_ctx.encode(data, q, 22, mode)
        assert gpu_ctx.decode(s, len(data)) == data, (name, q)
    for q in (5, 11):
        s = O.encode(data, q, 22, 0)            # the reference's encoder (restated): full metablocks, no dictionary
        assert gpu_ctx.decode(s, len(data)) == data, (name, q, "oracle")
    s = O.ref_encode(data, 5, 22, 0)            # Google's C encoder: may use the static dictionary
    _gpu_decode_or_refuse(gpu_ctx, s, data)


def test_reference_decoder_vectors(gpu_ctx):
    """test/fixtures/vectors/*.compressed of the reference (Google's canonical decoder vectors)."""
    files = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "vectors", "*.compressed")))
    assert files
    ok = refused = 0
    for f in files:
        stream = open(f, "rb").read()
        try:
            expected = O.decode(stream, 4 << 20)
        except ValueError:
            continue
        if len(expected) > (2 << 20):
            continue
        if _gpu_decode_or_refuse(gpu_ctx, stream, expected):
            ok += 1
        else:
            refused += 1
    print(f"decoder vectors: {ok} decoded, {refused} refused (static dictionary)")
    assert ok >= 3


def test_multi_shard_streams_and_small_windows(gpu_ctx):
    parts = [load_input("alice29.txt"), bytes(70001), load_input("enc-ttf.bin")[:150003], load_input("random_org_10k.bin")]
    data = b"".join(parts)
    gpu_ctx.set_tuning(shard_bytes=65536, metablock_bytes=16384)
    try:
        for q in (2, 5, 10):
            s = gpu_ctx.encode(data, q, 22, 0)   # padding blocks between shards, unknown rings, several metablocks per shard
            assert gpu_ctx.decode(s, len(data)) == data
    finally:
        gpu_ctx.set_tuning()
    for lgwin in (10, 16, 17, 24):
        s = gpu_ctx.encode(data[:40000], 5, lgwin, 0)
        assert gpu_ctx.decode(s, 40000) == data[:40000]
    # raw (uncompressed-metablock) shards and the host-only forms
    rnd = np.random.default_rng(5).integers(0, 256, 5000, dtype=np.uint8).tobytes()
    gpu_ctx.set_tuning(shard_bytes=64)
    try:
        s = gpu_ctx.encode(rnd, 5, 22, 0)
    finally:
        gpu_ctx.set_tuning()
    assert gpu_ctx.decode(s, len(rnd)) == rnd
    for d in (b"", b"x", rnd[:63]):
        assert gpu_ctx.decode(gpu_ctx.encode(d, 5), len(d)) == d


def test_truncated_and_corrupt_streams_are_refused_or_detected(gpu_ctx):
    import brotli_lib_b200._lib as L
    data = load_input("alice29.txt")[:50000]
    s = gpu_ctx.encode(data, 5, 22, 0)
    with pytest.raises(L.BgpuError):
        gpu_ctx.decode(s[: len(s) // 2], len(data))
    with pytest.raises(L.BgpuError):
        gpu_ctx.decode(s, len(data) - 100)      # output buffer too small: reported, not overrun


@pytest.mark.parametrize("quality,mode", [(1, 0), (3, 0), (5, 0), (9, 2), (10, 0), (11, 2)])
def test_verify_flag_decodes_every_shard_on_the_device(gpu_ctx, quality, mode):
    from brotli_lib_b200 import workloads as W
    data = np.concatenate([W.mixed_segment(i, 0, 1 << 18) for i in range(8)] + [np.frombuffer(load_input("enc-ttf.bin"), dtype=np.uint8)]).tobytes()
    gpu_ctx.set_tuning(shard_bytes=1 << 18, flags=VERIFY)
    try:
        s = gpu_ctx.encode(data, quality, 22, mode)          # raises BGPU_E_VERIFY on any mismatch
        assert gpu_ctx.counter(4) == 0
        assert any(n == "k_verify" for n, _ in gpu_ctx.timings())
    finally:
        gpu_ctx.set_tuning()
    assert O.decode(s, len(data)) == data


def test_verify_flag_full_size_shards_with_lanes(gpu_ctx):
    from brotli_lib_b200 import workloads as W
    data = W.mixed(36 << 20).tobytes()
    gpu_ctx.set_tuning(flags=VERIFY, lanes=3, group_shards=3)
    try:
        s = gpu_ctx.encode(data, 5, 22, 0)
        assert gpu_ctx.counter(4) == 0
    finally:
        gpu_ctx.set_tuning()
    assert O.decode(s, len(data)) == data
