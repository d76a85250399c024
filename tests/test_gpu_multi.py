"""GPU tier (-m gpu): one brotliEncode call over several devices (bgpu_ctx_create_multi).

On a one-GPU box the device list repeats device 0 (several contexts on one GPU, one host thread each): the part
cutting, the per-device workers and the stitching are the same code; with >= 2 GPUs visible the real list is used too."""
import numpy as np
import pytest

import _oracle as O
from conftest import load_input

pytestmark = pytest.mark.gpu


def _rt(stream, data):
    assert O.decode(stream, len(data), "system") == data
    if O.have_ref_decoder():
        assert O.decode(stream, len(data), "ref") == data


def _ngpu():
    import torch
    return torch.cuda.device_count()


def _blob():
    from brotli_lib_b200 import workloads as W
    parts = [load_input("alice29.txt"), bytes(70001), load_input("enc-ttf.bin")[:150003], W.mixed_segment(4, 0, 300000).tobytes(),
             load_input("random_org_10k.bin"), W.mixed_segment(6, 0, 400000).tobytes()]
    return b"".join(parts)


@pytest.mark.parametrize("ndev", [2, 3, 8])
@pytest.mark.parametrize("quality,mode", [(5, 0), (9, 2), (3, 0), (11, 0)])
def test_multi_device_one_stream(gpu_ctx, ndev, quality, mode):
    import brotli_lib_b200 as B
    data = _blob()
    devs = [i % max(1, _ngpu()) for i in range(ndev)]
    ctx = B.Context(devices=devs)
    try:
        ctx.set_tuning(shard_bytes=65536)
        gpu_ctx.set_tuning(shard_bytes=65536)
        s = ctx.encode(data, quality, 22, mode)
        _rt(s, data)
        one = gpu_ctx.encode(data, quality, 22, mode)
        # parts only add a byte-aligning empty metadata block and an unknown distance ring at their first shard:
        # the single-context stream has the same per-shard rules, so the sizes are equal
        assert abs(len(s) - len(one)) <= 4 * ndev, (len(s), len(one))
        assert s == ctx.encode(data, quality, 22, mode)   # idempotent
        assert ctx.launch_count() > 0
        # fewer shards than devices, exactly one shard (not split), host-only inputs
        _rt(ctx.encode(data[:100000], quality, 22, mode), data[:100000])
        assert ctx.encode(data[:65536], quality, 22, mode) == gpu_ctx.encode(data[:65536], quality, 22, mode)
        assert ctx.encode(b"", quality) == gpu_ctx.encode(b"", quality)
        assert ctx.encode(data[:50], quality) == gpu_ctx.encode(data[:50], quality)
    finally:
        gpu_ctx.set_tuning()
        ctx.close()


def test_multi_device_default_shards_and_batch(gpu_ctx):
    """Default 4 MiB shards with lanes inside every device; a batch of independent streams dealt over the devices."""
    import brotli_lib_b200 as B
    from brotli_lib_b200 import workloads as W
    data = W.mixed(40 << 20).tobytes()[: (40 << 20) - 12345]
    devs = [i % max(1, _ngpu()) for i in range(4)]
    ctx = B.Context(devices=devs)
    try:
        s = ctx.encode(data, 5, 22, 0)
        _rt(s, data)
        one = gpu_ctx.encode(data, 5, 22, 0)
        assert abs(len(s) - len(one)) <= 64
        items = [b"", b"x", data[:63], data[1000:1000 + 4096], data[: 1 << 20], load_input("alice29.txt"), bytes(5000)] * 3
        outs = ctx.encode_batch(items, 5, 22, 2)
        for d, st in zip(items, outs):
            assert st == gpu_ctx.encode(d, 5, 22, 2)
    finally:
        ctx.close()


def test_two_contexts_on_different_devices_in_one_process(gpu_ctx):
    """kernel attributes (dynamic shared memory opt-in) are per device, not per process."""
    import brotli_lib_b200 as B
    if _ngpu() < 2:
        pytest.skip("needs >= 2 visible GPUs")
    data = _blob()
    a, b = B.Context(0), B.Context(1)
    try:
        for q in (5, 9, 11):
            sa, sb = a.encode(data, q, 22, 0), b.encode(data, q, 22, 0)
            assert sa == sb
            _rt(sb, data)
    finally:
        a.close(); b.close()
