"""ctypes bindings for the oracle (test infrastructure) and the independent decoders.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_LIB = None


def build_oracle(force=False):
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    srcs = [os.path.join(ORACLE_DIR, "src", f) for f in os.listdir(os.path.join(ORACLE_DIR, "src"))]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", ORACLE_DIR, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


class Command(C.Structure):
    _fields_ = [("insertLen", C.c_uint32), ("copyLen", C.c_uint32), ("distExtra", C.c_uint32),
                ("cmdPrefix", C.c_uint16), ("distPrefix", C.c_uint16)]


CMD_DTYPE = np.dtype([("insertLen", "<u4"), ("copyLen", "<u4"), ("distExtra", "<u4"),
                      ("cmdPrefix", "<u2"), ("distPrefix", "<u2")])

TRIVIAL_DTYPE = np.dtype([("litHisto", "<u4", 256), ("cmdHisto", "<u4", 704), ("distHisto", "<u4", 544),
                          ("litDepths", "u1", 256), ("cmdDepths", "u1", 704), ("distDepths", "u1", 544),
                          ("litBits", "<u2", 256), ("cmdBits", "<u2", 704), ("distBits", "<u2", 544)])


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build_oracle())
        u8p, szp = C.c_void_p, C.POINTER(C.c_size_t)
        L.oracle_encode.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, u8p, C.c_size_t, szp]
        L.oracle_encode_stats.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, szp, C.POINTER(C.c_int)]
        L.oracle_encode_shards.argtypes = [u8p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, szp]
        L.oracle_encode_shards_sizes.argtypes = [u8p, C.c_size_t, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, szp, u8p]
        L.oracle_encode_part.argtypes = [u8p, C.c_size_t, u8p, C.c_int, C.c_int, C.c_int, C.c_int, u8p, C.c_size_t, szp]
        L.oracle_h5_chain_best.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, u8p, u8p]
        L.oracle_simple_table_best.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, u8p, u8p]
        L.oracle_h10_all_matches.argtypes = [u8p, C.c_size_t, C.c_int, u8p, u8p, u8p, C.c_size_t, szp]
        L.oracle_commands.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, u8p, C.c_size_t, szp]
        L.oracle_store_trivial.argtypes = [u8p, C.c_size_t, u8p, C.c_size_t, C.c_int, C.c_int, C.c_uint, C.c_uint,
                                           u8p, C.c_size_t, szp, C.POINTER(C.c_uint64), u8p]
        L.oracle_metablock_model.argtypes = [u8p, C.c_size_t, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, u8p, u8p, u8p, C.c_size_t, u8p, u8p]
        L.oracle_context_histograms.argtypes = [u8p, C.c_size_t, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, u8p, u8p, u8p, C.c_size_t, u8p, u8p, u8p]
        L.oracle_store_metablock.argtypes = [u8p, C.c_size_t, u8p, C.c_size_t, C.c_int, C.c_int, C.c_uint, C.c_uint, C.c_int, C.c_int,
                                             u8p, u8p, u8p, C.c_size_t, u8p, u8p, u8p, C.c_size_t, szp, C.POINTER(C.c_uint64)]
        L.oracle_encode_context_map.argtypes = [u8p, C.c_size_t, C.c_uint32, u8p, C.c_size_t, C.POINTER(C.c_uint64)]
        L.oracle_huffman.argtypes = [u8p, C.c_size_t, C.c_int, u8p, u8p]
        L.oracle_build_store_tree.argtypes = [u8p, C.c_size_t, u8p, u8p, u8p, C.c_size_t, C.POINTER(C.c_uint64)]
        L.oracle_context_lut.restype = C.POINTER(C.c_uint8)
        L.oracle_log2_table.restype = C.POINTER(C.c_double)
        L.oracle_stream_create.restype = C.c_void_p
        L.oracle_stream_create.argtypes = [C.c_int, C.c_int, C.c_int]
        L.oracle_stream_update.argtypes = [C.c_void_p, u8p, C.c_size_t, C.POINTER(C.c_void_p), szp]
        L.oracle_stream_finish.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), szp]
        L.oracle_stream_destroy.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def _buf(data):
    a = np.frombuffer(bytes(data), dtype=np.uint8) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data, dtype=np.uint8)
    return a, (a.ctypes.data if a.size else None)


def encode(data, quality=11, lgwin=22, mode=0, force_trivial=False):
    a, p = _buf(data)
    cap = a.size + a.size // 4 + 4096
    out = np.empty(cap, dtype=np.uint8)
    n = C.c_size_t(0)
    rc = lib().oracle_encode(p, a.size, quality, lgwin, mode, int(force_trivial), out.ctypes.data, cap, C.byref(n))
    if rc:
        raise RuntimeError(f"oracle_encode rc={rc}")
    return out[: n.value].tobytes()


def encode_part(data, prev2=b"", first=True, last=True, quality=11, lgwin=22, mode=0, shard=1 << 22):
    """One part of a stream (SURVEY.md §8e rules), itself cut into independent `shard`-byte windows the
    way the GPU path does; concatenating the parts of an input in order gives one valid stream."""
    L = lib()
    data = bytes(data)
    out = bytearray()
    n = len(data)
    for off in range(0, n, shard):
        chunk = data[off:off + shard]
        f = first and off == 0
        l = last and off + shard >= n
        pv = bytes(prev2) if off == 0 else data[off - 2:off]
        if off == 1:
            pv = bytes(prev2)[1:] + data[:1]
        pv = (b"\0\0" + pv)[-2:]
        cap = 2 * len(chunk) + 4096
        buf = np.empty(cap, dtype=np.uint8)
        ln = C.c_size_t(0)
        (ca, cp), (pa, pp) = _buf(chunk), _buf(pv)
        rc = L.oracle_encode_part(cp, ca.size, pp, (1 if f else 0) | (2 if l else 0), quality, lgwin, mode, buf.ctypes.data, cap, C.byref(ln))
        if rc:
            raise RuntimeError(f"oracle_encode_part rc={rc}")
        out += buf[: ln.value].tobytes()
    return bytes(out)


def encode_stats(data, quality=11, lgwin=22, mode=0):
    a, p = _buf(data)
    n = C.c_size_t(0)
    st = (C.c_int * 7)()
    rc = lib().oracle_encode_stats(p, a.size, quality, lgwin, mode, C.byref(n), st)
    if rc:
        raise RuntimeError(f"oracle_encode_stats rc={rc}")
    return n.value, list(st)


def encode_shards_total(data, shard, quality, lgwin=22, mode=0, threads=1):
    a, p = _buf(data)
    n = C.c_size_t(0)
    rc = lib().oracle_encode_shards(p, a.size, shard, quality, lgwin, mode, threads, C.byref(n))
    if rc:
        raise RuntimeError(f"oracle_encode_shards rc={rc}")
    return n.value


def encode_shards_sizes(data, shard, quality, lgwin=22, mode=0, threads=1):
    """Per-shard compressed sizes of `data` cut into independent `shard`-byte streams (threaded)."""
    a, p = _buf(data)
    n = C.c_size_t(0)
    sizes = np.zeros((a.size + shard - 1) // shard, dtype=np.uint64)
    rc = lib().oracle_encode_shards_sizes(p, a.size, shard, quality, lgwin, mode, threads, C.byref(n), sizes.ctypes.data)
    if rc:
        raise RuntimeError(f"oracle_encode_shards_sizes rc={rc}")
    return sizes


def h5_chain_best(data, quality=5, lgwin=22):
    a, p = _buf(data)
    ln = np.zeros(a.size, dtype=np.uint32)
    ds = np.zeros(a.size, dtype=np.uint32)
    lib().oracle_h5_chain_best(p, a.size, quality, lgwin, ln.ctypes.data, ds.ctypes.data)
    return ln, ds


def simple_table_best(data, quality=4, lgwin=22):
    """hash-simple.ts table candidate of every position (all earlier positions stored, empty distance cache)."""
    a, p = _buf(data)
    ln = np.zeros(a.size, dtype=np.uint32)
    ds = np.zeros(a.size, dtype=np.uint32)
    lib().oracle_simple_table_best(p, a.size, quality, lgwin, ln.ctypes.data, ds.ctypes.data)
    return ln, ds


def h10_all_matches(data, lgwin=22):
    """(offsets[n+1], dist, len) of BinaryTreeHasher.findAllMatches over every position."""
    a, p = _buf(data)
    cap = a.size * 24 + 64
    off = np.zeros(a.size + 1, dtype=np.uint32)
    ds = np.zeros(cap, dtype=np.uint32)
    ln = np.zeros(cap, dtype=np.uint32)
    n = C.c_size_t(0)
    rc = lib().oracle_h10_all_matches(p, a.size, lgwin, off.ctypes.data, ds.ctypes.data, ln.ctypes.data, cap, C.byref(n))
    if rc:
        raise RuntimeError(f"oracle_h10_all_matches rc={rc}")
    return off, ds[: n.value].copy(), ln[: n.value].copy()


def commands(data, quality=5, lgwin=22, mode=0):
    a, p = _buf(data)
    cap = a.size // 2 + 16
    out = np.zeros(cap, dtype=CMD_DTYPE)
    n = C.c_size_t(0)
    rc = lib().oracle_commands(p, a.size, quality, lgwin, mode, out.ctypes.data, cap, C.byref(n))
    if rc:
        raise RuntimeError(f"oracle_commands rc={rc}")
    return out[: n.value].copy()


def store_trivial(data, cmds, is_last=True, font=False, start_nbits=0, start_value=0):
    a, p = _buf(data)
    cmds = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
    cap = a.size + a.size // 4 + 4096
    out = np.zeros(cap, dtype=np.uint8)
    n = C.c_size_t(0)
    nb = C.c_uint64(0)
    tab = np.zeros(1, dtype=TRIVIAL_DTYPE)
    rc = lib().oracle_store_trivial(p, a.size, cmds.ctypes.data, cmds.size, int(is_last), int(font), start_nbits, start_value,
                                    out.ctypes.data, cap, C.byref(n), C.byref(nb), tab.ctypes.data)
    if rc:
        raise RuntimeError(f"oracle_store_trivial rc={rc}")
    return out[: n.value].tobytes(), nb.value, tab[0]


# ---- full-metablock stages (metablock.ts:504-761 cut at the model boundary) ----
MODEL_CAP_BLOCKS = 1 << 16


def _flat_model(model):
    """dict model -> the flat arrays of oracle/src/encode.cc (FlatToModel)."""
    hdr = np.zeros(16, dtype=np.uint32)
    types = np.zeros(3 * MODEL_CAP_BLOCKS, dtype=np.uint8)
    lengths = np.zeros(3 * MODEL_CAP_BLOCKS, dtype=np.uint32)
    hdr[0] = model["context_mode"]
    for k, (nt, t, ln) in enumerate(model["splits"]):
        hdr[1 + k] = nt; hdr[4 + k] = len(t)
        types[k * MODEL_CAP_BLOCKS:k * MODEL_CAP_BLOCKS + len(t)] = t
        lengths[k * MODEL_CAP_BLOCKS:k * MODEL_CAP_BLOCKS + len(ln)] = ln
    hdr[7] = model.get("literal_clusters", 1); hdr[8] = model.get("distance_clusters", 1)
    lm = np.ascontiguousarray(model.get("literal_map", np.zeros(int(hdr[1]) * 64)), dtype=np.uint32)
    dm = np.ascontiguousarray(model.get("distance_map", np.zeros(int(hdr[3]) * 4)), dtype=np.uint32)
    return hdr, types, lengths, lm, dm


def metablock_model(data, cmds, quality=5, font=False, prev=(0, 0)):
    """splitBlock + chooseContextMode + context-map clustering of one metablock = the whole input.
    Returns None when the reference stores it trivially, else the model dict (see Context._mb_model)."""
    a, p = _buf(data)
    cmds = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
    hdr = np.zeros(16, dtype=np.uint32)
    types = np.zeros(3 * MODEL_CAP_BLOCKS, dtype=np.uint8)
    lengths = np.zeros(3 * MODEL_CAP_BLOCKS, dtype=np.uint32)
    lm = np.zeros(256 * 64, dtype=np.uint32)
    dm = np.zeros(256 * 4, dtype=np.uint32)
    rc = lib().oracle_metablock_model(p, a.size, cmds.ctypes.data, cmds.size, quality, int(font), prev[0], prev[1], hdr.ctypes.data,
                                      types.ctypes.data, lengths.ctypes.data, MODEL_CAP_BLOCKS, lm.ctypes.data, dm.ctypes.data)
    if rc == 1:
        return None
    if rc:
        raise RuntimeError(f"oracle_metablock_model rc={rc}")
    splits = []
    for k in range(3):
        nb = int(hdr[4 + k])
        splits.append((int(hdr[1 + k]), types[k * MODEL_CAP_BLOCKS:k * MODEL_CAP_BLOCKS + nb].copy(), lengths[k * MODEL_CAP_BLOCKS:k * MODEL_CAP_BLOCKS + nb].copy()))
    return {"context_mode": int(hdr[0]), "splits": splits, "literal_map": lm[: int(hdr[1]) * 64].copy(), "literal_clusters": int(hdr[7]),
            "distance_map": dm[: int(hdr[3]) * 4].copy(), "distance_clusters": int(hdr[8])}


def context_histograms(data, cmds, model, font=False, prev=(0, 0)):
    a, p = _buf(data)
    cmds = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
    hdr, types, lengths, _, _ = _flat_model(model)
    da = 124 if font else 64
    lh = np.zeros((int(hdr[1]) * 64, 256), dtype=np.uint32)
    ch = np.zeros((int(hdr[2]), 704), dtype=np.uint32)
    dh = np.zeros((int(hdr[3]) * 4, da), dtype=np.uint32)
    rc = lib().oracle_context_histograms(p, a.size, cmds.ctypes.data, cmds.size, int(font), prev[0], prev[1], hdr.ctypes.data, types.ctypes.data,
                                         lengths.ctypes.data, MODEL_CAP_BLOCKS, lh.ctypes.data, ch.ctypes.data, dh.ctypes.data)
    if rc:
        raise RuntimeError(f"oracle_context_histograms rc={rc}")
    return lh, ch, dh


def store_metablock(data, cmds, model, is_last=True, font=False, start_nbits=0, start_value=0, prev=(0, 0)):
    a, p = _buf(data)
    cmds = np.ascontiguousarray(cmds, dtype=CMD_DTYPE)
    hdr, types, lengths, lm, dm = _flat_model(model)
    cap = a.size * 2 + (1 << 17)
    out = np.zeros(cap, dtype=np.uint8)
    n = C.c_size_t(0)
    nb = C.c_uint64(0)
    rc = lib().oracle_store_metablock(p, a.size, cmds.ctypes.data, cmds.size, int(is_last), int(font), start_nbits, start_value, prev[0], prev[1],
                                      hdr.ctypes.data, types.ctypes.data, lengths.ctypes.data, MODEL_CAP_BLOCKS, lm.ctypes.data, dm.ctypes.data,
                                      out.ctypes.data, cap, C.byref(n), C.byref(nb))
    if rc:
        raise RuntimeError(f"oracle_store_metablock rc={rc}")
    return out[: n.value].tobytes(), nb.value


def encode_context_map(cmap, num_clusters):
    mp = np.ascontiguousarray(cmap, dtype=np.uint32)
    out = np.zeros(mp.size * 4 + 4096, dtype=np.uint8)
    nb = C.c_uint64(0)
    rc = lib().oracle_encode_context_map(mp.ctypes.data, mp.size, int(num_clusters), out.ctypes.data, out.size, C.byref(nb))
    if rc:
        raise RuntimeError("oracle_encode_context_map failed")
    return out[: (nb.value + 7) // 8].tobytes(), nb.value


def set_h10_compat(bits):
    """0: H10 repairs on (default); bit 0: zero-initialised buckets, bit 1: dangling links at depth exhaustion (the TS as written)."""
    lib().oracle_set_h10_compat(int(bits))


def huffman(hist, tree_limit=15):
    h = np.ascontiguousarray(hist, dtype=np.uint32)
    d = np.zeros(h.size, dtype=np.uint8)
    b = np.zeros(h.size, dtype=np.uint16)
    lib().oracle_huffman(h.ctypes.data, h.size, tree_limit, d.ctypes.data, b.ctypes.data)
    return d, b


def build_store_tree(hist):
    h = np.ascontiguousarray(hist, dtype=np.uint32)
    d = np.zeros(h.size, dtype=np.uint8)
    b = np.zeros(h.size, dtype=np.uint16)
    out = np.zeros(4096, dtype=np.uint8)
    nb = C.c_uint64(0)
    rc = lib().oracle_build_store_tree(h.ctypes.data, h.size, d.ctypes.data, b.ctypes.data, out.ctypes.data, out.size, C.byref(nb))
    if rc:
        raise RuntimeError("oracle_build_store_tree failed")
    return d, b, out[: (nb.value + 7) // 8].tobytes(), nb.value


class StreamEncoder:
    def __init__(self, quality=11, lgwin=22, mode=0):
        self._h = lib().oracle_stream_create(quality, lgwin, mode)
        if not self._h:
            raise RuntimeError("unsupported quality for oracle streaming")

    def update(self, chunk):
        a, p = _buf(chunk)
        o = C.c_void_p()
        n = C.c_size_t(0)
        lib().oracle_stream_update(self._h, p, a.size, C.byref(o), C.byref(n))
        return C.string_at(o, n.value) if n.value else b""

    def finish(self):
        o = C.c_void_p()
        n = C.c_size_t(0)
        lib().oracle_stream_finish(self._h, C.byref(o), C.byref(n))
        return C.string_at(o, n.value) if n.value else b""

    def __del__(self):
        if getattr(self, "_h", None):
            lib().oracle_stream_destroy(self._h)
            self._h = None


# ---------------------------------------------------------------- independent decoders
_DEC = {}


def _decoder(which):
    if which not in _DEC:
        if which == "system":
            L = C.CDLL("libbrotlidec.so.1")
        else:
            path = os.path.join(ORACLE_DIR, "_ref", "libbrotli_ref.so")
            if not os.path.exists(path):
                if os.path.isdir("/root/reference/bench/vendor/brotli/c"):
                    subprocess.check_call(["make", "-C", ORACLE_DIR, "ref"], stdout=subprocess.DEVNULL)
                else:
                    return None
            L = C.CDLL(path)
        L.BrotliDecoderDecompress.argtypes = [C.c_size_t, C.c_void_p, C.POINTER(C.c_size_t), C.c_void_p]
        L.BrotliDecoderDecompress.restype = C.c_int
        _DEC[which] = L
    return _DEC[which]


def have_ref_decoder():
    return _decoder("ref") is not None


def decode(stream, max_out, which="system"):
    """Google C decoder (system libbrotlidec 1.1.0, or the vendored 1.2.0 build under oracle/_ref).
    Returns bytes or raises ValueError."""
    L = _decoder(which)
    if L is None:
        raise RuntimeError("decoder unavailable: " + which)
    s = np.frombuffer(bytes(stream), dtype=np.uint8)
    out = np.empty(max_out + 16, dtype=np.uint8)
    n = C.c_size_t(out.size)
    rc = L.BrotliDecoderDecompress(s.size, s.ctypes.data, C.byref(n), out.ctypes.data)
    if rc != 1:
        raise ValueError(f"BrotliDecoderDecompress failed rc={rc}")
    return out[: n.value].tobytes()


def ref_encode(data, quality, lgwin=22, mode=0, which="system"):
    """Google C encoder — context line only, never a parity target."""
    L = C.CDLL("libbrotlienc.so.1") if which == "system" else _decoder("ref")
    L.BrotliEncoderCompress.argtypes = [C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.POINTER(C.c_size_t), C.c_void_p]
    L.BrotliEncoderCompress.restype = C.c_int
    a, p = _buf(data)
    cap = a.size + a.size // 4 + 4096
    out = np.empty(cap, dtype=np.uint8)
    n = C.c_size_t(cap)
    rc = L.BrotliEncoderCompress(quality, lgwin, mode, a.size, p, C.byref(n), out.ctypes.data)
    if rc != 1:
        raise ValueError("BrotliEncoderCompress failed")
    return out[: n.value].tobytes()
