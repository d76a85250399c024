"""GPU tier (-m gpu), collected last: bgpu_decode_noalloc, the decoder entry point that allocates nothing
(include/brotligpu.h; no counterpart in the reference).

Its scratch memory is a file-scope device array of BGPU_NOALLOC_SCRATCH_BYTES shared by every call, so the checks
are: same bytes as the existing entry point (bgpu_decode, which uses cudaMalloc'ed scratch) — bit-exact, the bar of
every decoder test here — on two calls in a row with different batches (the second must not see the first's tables,
jobs or results), the batch limit refused with a clear message, and no device allocation made by the call itself.
Device buffers for inputs and outputs are the CALLER's: the test makes them with torch before the calls."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_input

pytestmark = pytest.mark.gpu

MAX_STREAMS = 8              # BGPU_NOALLOC_MAX_STREAMS
SCRATCH_BYTES = 10 << 20     # BGPU_NOALLOC_SCRATCH_BYTES
_warm = {"done": False}      # the first launch of a kernel may make the driver load its code (lazy module loading)


def _noalloc(L, torch, streams, caps, cuda_stream=None):
    """Calls bgpu_decode_noalloc on device copies of `streams`; returns (rc, [decoded bytes], [status])."""
    n = len(streams)
    d_in = [torch.frombuffer(bytearray(s), dtype=torch.uint8).cuda() for s in streams]
    d_out = [torch.full((max(c, 1) + 16,), 0xAA, dtype=torch.uint8, device="cuda") for c in caps]
    torch.cuda.synchronize()
    in_ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in d_in])
    out_ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in d_out])
    lens = (C.c_size_t * n)(*[len(s) for s in streams])
    capv = (C.c_size_t * n)(*caps)
    got = (C.c_size_t * n)()
    status = (C.c_int * n)()
    free0 = torch.cuda.mem_get_info()[0]
    rc = L.bgpu_decode_noalloc(n, in_ptrs, lens, out_ptrs, capv, got, status, cuda_stream)
    if _warm["done"] or rc == -1:
        assert torch.cuda.mem_get_info()[0] == free0, "the call changed the device's free memory"
    _warm["done"] = _warm["done"] or rc != -1
    outs = []
    for k in range(n):
        host = d_out[k].cpu().numpy()
        assert (host[caps[k]:] == 0xAA).all(), "wrote beyond out_cap"
        outs.append(host[: got[k]].tobytes())
    return rc, outs, list(status)


def _streams(gpu_ctx):
    cases = [("alice29.txt", 5, 0), ("enc-ttf.bin", 9, 2), ("abc1000", 1, 0), ("mapsdatazrh", 11, 0), ("font4k", 11, 2), ("zeros", 5, 0),
             ("random_org_10k.bin", 3, 0), ("html-content.bin", 5, 0), ("quickfox_repeated", 9, 0), ("ramp", 0, 0), ("alice29.txt", 2, 0)]
    out = []
    for name, q, mode in cases:
        data = load_input(name)[:200000]
        out.append((data, gpu_ctx.encode(data, q, 22, mode)))
    return out


def test_noalloc_decoder_matches_bgpu_decode_on_two_calls_in_a_row(gpu_ctx):
    import torch
    import brotli_lib_b200._lib as lib
    L = lib.load()
    assert L.bgpu_noalloc_scratch_bytes() == SCRATCH_BYTES
    cases = _streams(gpu_ctx)
    want = [gpu_ctx.decode(s, len(d)) for d, s in cases]           # the existing entry point
    assert want == [d for d, _ in cases]
    # call 1: three streams on the default stream
    first = cases[:3]
    rc, outs, status = _noalloc(L, torch, [s for _, s in first], [len(d) for d, _ in first])
    assert rc == 0 and status == [0, 0, 0], lib.BgpuError(rc) if rc else None
    assert outs == want[:3]
    # call 2, straight after: a full batch of eight OTHER streams (every arena reused or first used), on a caller's stream
    second = cases[3:3 + MAX_STREAMS]
    st = torch.cuda.Stream()
    rc, outs, status = _noalloc(L, torch, [s for _, s in second], [len(d) for d, _ in second], C.c_void_p(st.cuda_stream))
    assert rc == 0 and status == [0] * MAX_STREAMS
    assert outs == want[3:3 + MAX_STREAMS]
    # call 3: the first batch again in reverse order, one output too small — per-stream status, others unaffected
    third = first[::-1]
    caps = [len(d) for d, _ in third]
    caps[1] = 100
    rc, outs, status = _noalloc(L, torch, [s for _, s in third], caps)
    assert rc == lib.E_OUTPUT_TOO_SMALL and status == [0, lib.E_OUTPUT_TOO_SMALL, 0]
    assert outs[0] == third[0][0] and outs[2] == third[2][0]
    # a corrupt stream is refused, as by bgpu_decode
    bad = bytearray(cases[0][1]); bad[len(bad) // 2] ^= 0x55; bad = bytes(bad[: len(bad) - 7])
    with pytest.raises(lib.BgpuError) as e:
        gpu_ctx.decode(bad, len(cases[0][0]))
    rc, outs, status = _noalloc(L, torch, [bad], [len(cases[0][0])])
    assert rc == e.value.code == status[0]


def test_noalloc_decoder_refuses_more_than_its_scratch_holds(gpu_ctx):
    import torch
    import brotli_lib_b200._lib as lib
    L = lib.load()
    data = load_input("abc1000")
    s = gpu_ctx.encode(data, 5, 22, 0)
    n = MAX_STREAMS + 1
    rc, outs, _ = _noalloc(L, torch, [s] * n, [len(data)] * n)
    assert rc == -1                                                  # BGPU_E_INVALID
    msg = L.bgpu_last_cuda_error().decode()
    assert "BGPU_NOALLOC_MAX_STREAMS = 8" in msg and str(SCRATCH_BYTES) in msg, msg
    assert all(o == b"" for o in outs)
    rc, outs, status = _noalloc(L, torch, [s] * MAX_STREAMS, [len(data)] * MAX_STREAMS)    # and the limit itself is fine
    assert rc == 0 and outs == [data] * MAX_STREAMS
    assert L.bgpu_decode_noalloc(0, None, None, None, None, None, None, None) == 0
