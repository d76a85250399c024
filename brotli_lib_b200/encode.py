"""Host-side mirror of brotli-lib's public encoder API over the CUDA hot path.

Mirrors /root/reference/src/encode/encode.ts:
  brotliEncode(input, {quality, lgwin, mode, sizeHint})      encode.ts:22-27, 50-90
  class BrotliEncoder { update(chunk); finish() }            encode.ts:290-409
  const enum EncoderMode { GENERIC, TEXT, FONT }             enc-constants.ts:56-60
Same names, argument meaning and error behaviour: options are clamped, never rejected
(encode.ts:57-62); TEXT == GENERIC; the result is a fresh bytes object; `update` returns the
bytes completed so far (possibly empty).  The work is done by libbrotligpu.so (include/brotligpu.h);
without the CUDA library or a device the calls raise — there is no CPU fallback.
"""
import ctypes as C
import enum

import numpy as np

from . import _lib


class EncoderMode(enum.IntEnum):
    GENERIC = 0
    TEXT = 1
    FONT = 2


class Context:
    """One bgpu_ctx (device, CUDA stream, scratch arenas).  Not thread-safe."""

    def __init__(self, device=-1, shard_bytes=0, segment_bytes=0, metablock_bytes=0, devices=None):
        """device: one GPU (-1 = current).  devices=[...]: one context over several GPUs of the box (bgpu_ctx_create_multi):
        encode() of a large input then runs one part per device and returns one stitched stream."""
        self._L = _lib.load()
        h = C.c_void_p()
        if devices is not None:
            ids = (C.c_int * len(devices))(*[int(d) for d in devices])
            _lib.check(self._L.bgpu_ctx_create_multi(ids, len(devices), C.byref(h)))
        else:
            _lib.check(self._L.bgpu_ctx_create(device, C.byref(h)))
        self._h = h
        if shard_bytes or segment_bytes or metablock_bytes:
            self.set_tuning(shard_bytes, segment_bytes, metablock_bytes)

    def set_tuning(self, shard_bytes=0, segment_bytes=0, metablock_bytes=0, flags=0, lanes=0, group_shards=0):
        t = _lib.Tuning(shard_bytes, segment_bytes, metablock_bytes, flags, lanes, group_shards)
        _lib.check(self._L.bgpu_ctx_set_tuning(self._h, C.byref(t)))

    def set_stream(self, cuda_stream):
        """Run on the caller's CUDA stream (an int handle, e.g. torch.cuda.current_stream().cuda_stream)."""
        _lib.check(self._L.bgpu_ctx_set_stream(self._h, C.c_void_p(cuda_stream or None)))

    def close(self):
        if getattr(self, "_h", None):
            self._L.bgpu_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- public API ----
    def encode(self, data, quality=11, lgwin=22, mode=EncoderMode.GENERIC):
        a = _as_u8(data)
        cap = self._L.bgpu_max_compressed_size(a.size)
        out = np.empty(cap, dtype=np.uint8)
        n = C.c_size_t(0)
        rc = self._L.bgpu_encode(self._h, _ptr(a), a.size, int(quality), int(lgwin), int(mode), out.ctypes.data, cap, C.byref(n))
        if rc == _lib.E_OUTPUT_TOO_SMALL:   # only with a custom shard_bytes below the default: the call reports the size needed
            cap = n.value
            out = np.empty(cap, dtype=np.uint8)
            rc = self._L.bgpu_encode(self._h, _ptr(a), a.size, int(quality), int(lgwin), int(mode), out.ctypes.data, cap, C.byref(n))
        _lib.check(rc)
        return out[: n.value].tobytes()

    def encode_batch(self, items, quality=11, lgwin=22, mode=EncoderMode.GENERIC):
        arrs = [_as_u8(x) for x in items]
        k = len(arrs)
        if k == 0:
            return []
        outs = [np.empty(self._L.bgpu_max_compressed_size(a.size), dtype=np.uint8) for a in arrs]
        inp = (C.c_void_p * k)(*[_ptr(a) for a in arrs])
        ns = (C.c_size_t * k)(*[a.size for a in arrs])
        outp = (C.c_void_p * k)(*[o.ctypes.data for o in outs])
        caps = (C.c_size_t * k)(*[o.size for o in outs])
        lens = (C.c_size_t * k)()
        _lib.check(self._L.bgpu_encode_batch(self._h, k, inp, ns, int(quality), int(lgwin), int(mode), outp, caps, lens))
        return [outs[i][: lens[i]].tobytes() for i in range(k)]

    def encode_part(self, data, prev2=b"", first=True, last=True, quality=11, lgwin=22, mode=EncoderMode.GENERIC):
        """One contiguous part of a stream whose other parts are encoded elsewhere (bgpu_encode_part).
        prev2: the two input bytes before `data` (needed unless first)."""
        a = _as_u8(data)
        cap = self._L.bgpu_max_compressed_size(a.size)
        out = np.empty(cap, dtype=np.uint8)
        n = C.c_size_t(0)
        pv = _as_u8(prev2)
        if not first and pv.size != 2:
            raise ValueError("prev2 must hold the two bytes before a non-first part")
        flags = (1 if first else 0) | (2 if last else 0)
        _lib.check(self._L.bgpu_encode_part(self._h, _ptr(a), a.size, _ptr(pv), flags, int(quality), int(lgwin), int(mode),
                                            out.ctypes.data, cap, C.byref(n)))
        return out[: n.value].tobytes()

    def encode_device(self, d_ptr, n, quality=11, lgwin=22, mode=EncoderMode.GENERIC, fetch=False):
        """Input already resident on the device (bench).  Returns (compressed_size, kernel_ms, bytes|None)."""
        ms = C.c_float(0)
        ln = C.c_size_t(0)
        out = None
        optr, cap = None, 0
        if fetch:
            cap = self._L.bgpu_max_compressed_size(n)
            out = np.empty(cap, dtype=np.uint8)
            optr = out.ctypes.data
        _lib.check(self._L.bgpu_encode_device(self._h, d_ptr, n, int(quality), int(lgwin), int(mode), optr, cap, C.byref(ln), C.byref(ms)))
        return ln.value, ms.value, (out[: ln.value].tobytes() if fetch else None)

    def decode(self, stream, max_out):
        """On-device Brotli decoder (bgpu_decode: one device thread, no static dictionary) — self-verification and tests."""
        a = _as_u8(stream)
        out = np.empty(max_out + 16, dtype=np.uint8)
        n = C.c_size_t(0)
        _lib.check(self._L.bgpu_decode(self._h, _ptr(a), a.size, out.ctypes.data, max_out, C.byref(n)))
        return out[: n.value].tobytes()

    def timings(self):
        names = (C.c_char_p * 64)()
        ms = (C.c_float * 64)()
        k = min(self._L.bgpu_stage_timings(self._h, names, ms, 64), 64)
        return [(names[i].decode(), ms[i]) for i in range(k)]

    def launch_count(self):
        return self._L.bgpu_last_launch_count(self._h)

    def counter(self, which=0):
        """Work counters of the last call (0: H10 match entries)."""
        return int(self._L.bgpu_last_counter(self._h, which))

    # ---- stage entry points (parity tests) ----
    def stage_match_candidates(self, data, quality=5, lgwin=22):
        a = _as_u8(data)
        ln = np.zeros(a.size, dtype=np.uint32)
        ds = np.zeros(a.size, dtype=np.uint32)
        _lib.check(self._L.bgpu_stage_match_candidates(self._h, _ptr(a), a.size, quality, lgwin, ln.ctypes.data, ds.ctypes.data))
        return ln, ds

    def stage_h10_matches(self, data, lgwin=22):
        """(offsets[n+1], dist, len): hash-binary-tree.ts findAllMatches for every position."""
        a = _as_u8(data)
        cap = a.size * 24 + 64
        off = np.zeros(a.size + 1, dtype=np.uint32)
        ds = np.zeros(cap, dtype=np.uint32)
        ln = np.zeros(cap, dtype=np.uint32)
        n = C.c_size_t(0)
        _lib.check(self._L.bgpu_stage_h10_matches(self._h, _ptr(a), a.size, lgwin, off.ctypes.data, ds.ctypes.data, ln.ctypes.data, cap, C.byref(n)))
        return off, ds[: n.value].copy(), ln[: n.value].copy()

    def stage_commands(self, data, quality=5, lgwin=22, mode=0):
        a = _as_u8(data)
        cap = a.size // 2 + 16
        out = np.zeros(cap, dtype=CMD_DTYPE)
        n = C.c_size_t(0)
        _lib.check(self._L.bgpu_stage_commands(self._h, _ptr(a), a.size, quality, lgwin, int(mode), out.ctypes.data, cap, C.byref(n)))
        return out[: n.value].copy()

    def stage_store_trivial(self, data, cmds, is_last=True, mode=0, start_nbits=0, start_value=0):
        a = _as_u8(data)
        cm = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
        cap = self._L.bgpu_max_compressed_size(a.size)
        out = np.zeros(cap, dtype=np.uint8)
        n = C.c_size_t(0)
        nb = C.c_uint64(0)
        tab = np.zeros(1, dtype=TRIVIAL_DTYPE)
        _lib.check(self._L.bgpu_stage_store_trivial(self._h, _ptr(a), a.size, cm.ctypes.data, cm.size, int(is_last), int(mode),
                                                    start_nbits, start_value, out.ctypes.data, cap, C.byref(n), C.byref(nb), tab.ctypes.data))
        return out[: n.value].tobytes(), tab[0]

    def stage_huffman(self, hists):
        h = np.ascontiguousarray(hists, dtype=np.uint32)
        if h.ndim == 1:
            h = h[None, :]
        count, alphabet = h.shape
        d = np.zeros((count, alphabet), dtype=np.uint8)
        b = np.zeros((count, alphabet), dtype=np.uint16)
        tb = np.zeros((count, 1024), dtype=np.uint8)
        tn = np.zeros(count, dtype=np.uint32)
        _lib.check(self._L.bgpu_stage_huffman(self._h, h.ctypes.data, count, alphabet, d.ctypes.data, b.ctypes.data, tb.ctypes.data, tn.ctypes.data))
        return d, b, tb, tn


    # full-metablock stages: `model` is a dict {"context_mode", "splits": [(num_types, types u8[], lengths u32[]) x 3],
    # "literal_map", "literal_clusters", "distance_map", "distance_clusters"} (include/brotligpu.h bgpu_mb_model)
    def _mb_model(self, model):
        keep = []
        m = _lib.MbModel()
        for k, (nt, types, lengths) in enumerate(model["splits"]):
            t = np.ascontiguousarray(types, dtype=np.uint8)
            ln = np.ascontiguousarray(lengths, dtype=np.uint32)
            keep += [t, ln]
            m.split[k].num_types = int(nt); m.split[k].num_blocks = int(t.size)
            m.split[k].types = _ptr(t); m.split[k].lengths = _ptr(ln)
        m.context_mode = int(model.get("context_mode", -1))
        lm = np.ascontiguousarray(model.get("literal_map", np.zeros(0)), dtype=np.uint32)
        dm = np.ascontiguousarray(model.get("distance_map", np.zeros(0)), dtype=np.uint32)
        keep += [lm, dm]
        m.literal_context_map = _ptr(lm); m.num_literal_clusters = int(model.get("literal_clusters", 1))
        m.distance_context_map = _ptr(dm); m.num_distance_clusters = int(model.get("distance_clusters", 1))
        return m, keep

    def stage_histograms(self, data, cmds, model, mode=0, prev2=None):
        """(lit [types*64][256], cmd [types][704], dist [types*4][A], context_mode): metablock.ts:558-637 on the device."""
        a = _as_u8(data)
        cm = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
        m, keep = self._mb_model(model)
        da = 124 if int(mode) == 2 else 64
        lh = np.zeros((m.split[0].num_types * 64, 256), dtype=np.uint32)
        ch = np.zeros((m.split[1].num_types, 704), dtype=np.uint32)
        dh = np.zeros((m.split[2].num_types * 4, da), dtype=np.uint32)
        pv = _as_u8(prev2) if prev2 is not None else None
        cmode = C.c_int(-1)
        _lib.check(self._L.bgpu_stage_histograms(self._h, _ptr(a), a.size, cm.ctypes.data, cm.size, int(mode), _ptr(pv) if pv is not None else None,
                                                 C.byref(m), lh.ctypes.data, ch.ctypes.data, dh.ctypes.data, C.byref(cmode)))
        return lh, ch, dh, cmode.value

    def stage_context_map(self, cmap, num_clusters):
        """(bytes, nbits) of encodeContextMap (context-map.ts:114-170) on the device."""
        mp = np.ascontiguousarray(cmap, dtype=np.uint32)
        out = np.zeros(mp.size * 4 + 4096, dtype=np.uint8)
        nb = C.c_uint64(0)
        _lib.check(self._L.bgpu_stage_context_map(self._h, mp.ctypes.data, mp.size, int(num_clusters), out.ctypes.data, out.size, C.byref(nb)))
        return out[: (nb.value + 7) // 8].tobytes(), nb.value

    def stage_store_metablock(self, data, cmds, model, is_last=True, mode=0, start_nbits=0, start_value=0, prev2=None):
        a = _as_u8(data)
        cm = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
        m, keep = self._mb_model(model)
        cap = self._L.bgpu_max_compressed_size(a.size) + 65536
        out = np.zeros(cap, dtype=np.uint8)
        n = C.c_size_t(0)
        pv = _as_u8(prev2) if prev2 is not None else None
        _lib.check(self._L.bgpu_stage_store_metablock(self._h, _ptr(a), a.size, cm.ctypes.data, cm.size, int(is_last), int(mode), start_nbits, start_value,
                                                      _ptr(pv) if pv is not None else None, C.byref(m), out.ctypes.data, cap, C.byref(n)))
        return out[: n.value].tobytes()


CMD_DTYPE = np.dtype([("insertLen", "<u4"), ("copyLen", "<u4"), ("distExtra", "<u4"), ("cmdPrefix", "<u2"), ("distPrefix", "<u2")])
TRIVIAL_DTYPE = np.dtype([("litHisto", "<u4", 256), ("cmdHisto", "<u4", 704), ("distHisto", "<u4", 544),
                          ("litDepths", "u1", 256), ("cmdDepths", "u1", 704), ("distDepths", "u1", 544),
                          ("litBits", "<u2", 256), ("cmdBits", "<u2", 704), ("distBits", "<u2", 544)])


def _as_u8(data):
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8)
    return np.frombuffer(bytes(data), dtype=np.uint8)


def _ptr(a):
    return a.ctypes.data if a.size else None


_default_ctx = None


def default_context():
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context()
    return _default_ctx


def brotliEncode(data, options=None, *, quality=None, lgwin=None, mode=None, sizeHint=None):
    """brotliEncode(input, {quality = 11, lgwin = 22, mode = GENERIC, sizeHint = 0}) -> bytes.

    encode.ts:50-90.  `options` may be a dict with the reference's option names; keyword
    arguments override it.  sizeHint is accepted and ignored (only dead code reads it)."""
    o = dict(options or {})
    q = quality if quality is not None else o.get("quality", 11)
    w = lgwin if lgwin is not None else o.get("lgwin", 22)
    m = mode if mode is not None else o.get("mode", EncoderMode.GENERIC)
    return default_context().encode(data, q, w, m)


class BrotliEncoder:
    """Streaming encoder (encode.ts:290-409): update(chunk) -> bytes, finish() -> bytes."""

    def __init__(self, options=None, *, quality=None, lgwin=None, mode=None, sizeHint=None, context=None):
        o = dict(options or {})
        q = quality if quality is not None else o.get("quality", 11)
        w = lgwin if lgwin is not None else o.get("lgwin", 22)
        m = mode if mode is not None else o.get("mode", EncoderMode.GENERIC)
        self._ctx = context or default_context()
        self._L = self._ctx._L
        h = C.c_void_p()
        _lib.check(self._L.bgpu_stream_create(self._ctx._h, int(q), int(w), int(m), C.byref(h)))
        self._h = h

    def update(self, chunk):
        a = _as_u8(chunk)
        o = C.c_void_p()
        n = C.c_size_t(0)
        _lib.check(self._L.bgpu_stream_update(self._h, _ptr(a), a.size, C.byref(o), C.byref(n)))
        return C.string_at(o, n.value) if n.value else b""

    def finish(self):
        o = C.c_void_p()
        n = C.c_size_t(0)
        _lib.check(self._L.bgpu_stream_finish(self._h, C.byref(o), C.byref(n)))
        return C.string_at(o, n.value) if n.value else b""

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._L.bgpu_stream_destroy(self._h)
                self._h = None
        except Exception:
            pass
