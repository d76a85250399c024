"""WOFF2 glyf/loca transform front end over the CUDA kernels (csrc/woff2.cuh; W3C WOFF 2.0 section 5.1).

Not part of the reference's API (brotli-lib only points at WOFF2, README.md:63): this is the data-parallel step directly
upstream of BASELINE config 5 — sfnt fonts in, transformed glyf blobs out, ready for `Context.encode_batch(..., mode=FONT)`.
Host side: sfnt table directory parsing (a few struct reads per font); the per-glyph work runs on the device.
"""
import ctypes as C
import struct

import numpy as np

from . import _lib
from .encode import default_context


def sfnt_glyf_tables(font):
    """(glyf, loca, numGlyphs, indexToLocFormat) of a TrueType sfnt file (bytes)."""
    n = struct.unpack(">H", font[4:6])[0]
    t = {}
    for i in range(n):
        tag, _, off, ln = struct.unpack(">4sIII", font[12 + 16 * i:28 + 16 * i])
        t[tag] = font[off:off + ln]
    for need in (b"glyf", b"loca", b"maxp", b"head"):
        if need not in t:
            raise ValueError(f"not a TrueType-outline sfnt: no {need.decode()} table")
    return t[b"glyf"], t[b"loca"], struct.unpack(">H", t[b"maxp"][4:6])[0], struct.unpack(">h", t[b"head"][50:52])[0]


def transform_glyf_batch(tables, context=None):
    """tables: [(glyf bytes, loca bytes, numGlyphs, indexToLocFormat)] -> [transformed glyf bytes] (one device job)."""
    ctx = context or default_context()
    L = ctx._L
    k = len(tables)
    if k == 0:
        return []
    keep, arr = [], (_lib.GlyfTable * k)()
    caps = (C.c_size_t * k)()
    for i, (glyf, loca, ng, fmt) in enumerate(tables):
        g = np.frombuffer(bytes(glyf), dtype=np.uint8)
        lo = np.frombuffer(bytes(loca), dtype=np.uint8)
        keep += [g, lo]
        arr[i].glyf = g.ctypes.data; arr[i].glyf_len = g.size
        arr[i].loca = lo.ctypes.data; arr[i].loca_len = lo.size
        arr[i].num_glyphs = int(ng); arr[i].index_format = int(fmt)
        caps[i] = L.bgpu_woff2_glyf_bound(g.size, int(ng))
    lens = (C.c_size_t * k)()
    for _ in range(2):   # a second pass only when a font needs more than the ordinary bound (the call reports the sizes)
        outs = [np.empty(caps[i], dtype=np.uint8) for i in range(k)]
        ptrs = (C.c_void_p * k)(*[o.ctypes.data for o in outs])
        rc = L.bgpu_woff2_transform_glyf(ctx._h, k, arr, ptrs, caps, lens)
        if rc != _lib.E_OUTPUT_TOO_SMALL:
            break
        for i in range(k):
            caps[i] = max(caps[i], lens[i])
    _lib.check(rc)
    return [outs[i][: lens[i]].tobytes() for i in range(k)]


def transform_font(font, context=None):
    """Transformed glyf table of one sfnt file."""
    return transform_glyf_batch([sfnt_glyf_tables(font)], context)[0]
