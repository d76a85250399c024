"""ctypes binding of libbrotligpu.so (include/brotligpu.h).  Fails loudly when the CUDA library
is missing or no device is usable — there is no CPU fallback on this path."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BGPU_LIB_PATH") or os.path.join(HERE, "libbrotligpu.so")   # override: kernel variants during development

E_OUTPUT_TOO_SMALL = -2
E_UNSUPPORTED = -3

SYMBOLS = [
    "bgpu_ctx_create", "bgpu_ctx_create_multi", "bgpu_ctx_destroy", "bgpu_ctx_set_tuning", "bgpu_ctx_set_stream", "bgpu_max_compressed_size", "bgpu_encode",
    "bgpu_encode_batch", "bgpu_encode_part", "bgpu_encode_device", "bgpu_decode", "bgpu_decode_noalloc", "bgpu_noalloc_scratch_bytes", "bgpu_woff2_glyf_bound", "bgpu_woff2_transform_glyf", "bgpu_stream_create", "bgpu_stream_update", "bgpu_stream_finish",
    "bgpu_stream_destroy", "bgpu_stage_match_candidates", "bgpu_stage_commands", "bgpu_stage_store_trivial",
    "bgpu_stage_huffman", "bgpu_stage_h10_matches", "bgpu_stage_histograms", "bgpu_stage_context_map", "bgpu_stage_store_metablock", "bgpu_stage_timings", "bgpu_last_launch_count", "bgpu_last_counter", "bgpu_strerror", "bgpu_last_cuda_error",
    "bgpu_version",
]


class Tuning(C.Structure):
    _fields_ = [("shard_bytes", C.c_uint32), ("segment_bytes", C.c_uint32), ("metablock_bytes", C.c_uint32),
                ("flags", C.c_uint32), ("lanes", C.c_uint32), ("group_shards", C.c_uint32)]


class BlockSplit(C.Structure):
    _fields_ = [("num_types", C.c_uint32), ("num_blocks", C.c_uint32), ("types", C.c_void_p), ("lengths", C.c_void_p)]


class MbModel(C.Structure):
    _fields_ = [("split", BlockSplit * 3), ("context_mode", C.c_int), ("literal_context_map", C.c_void_p),
                ("num_literal_clusters", C.c_uint32), ("distance_context_map", C.c_void_p), ("num_distance_clusters", C.c_uint32)]


class GlyfTable(C.Structure):
    _fields_ = [("glyf", C.c_void_p), ("glyf_len", C.c_size_t), ("loca", C.c_void_p), ("loca_len", C.c_size_t),
                ("num_glyphs", C.c_uint32), ("index_format", C.c_int)]


_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m brotli_lib_b200.build` (nvcc, sm_100a). "
                           "The encoder has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, sz, szp, i = C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_int
    L.bgpu_ctx_create.argtypes = [i, C.POINTER(vp)]
    L.bgpu_ctx_create_multi.argtypes = [C.POINTER(i), i, C.POINTER(vp)]
    L.bgpu_ctx_destroy.argtypes = [vp]
    L.bgpu_ctx_destroy.restype = None
    L.bgpu_ctx_set_tuning.argtypes = [vp, C.POINTER(Tuning)]
    L.bgpu_ctx_set_stream.argtypes = [vp, vp]
    L.bgpu_max_compressed_size.argtypes = [sz]
    L.bgpu_max_compressed_size.restype = sz
    L.bgpu_encode.argtypes = [vp, vp, sz, i, i, i, vp, sz, szp]
    L.bgpu_encode_batch.argtypes = [vp, i, C.POINTER(vp), szp, i, i, i, C.POINTER(vp), szp, szp]
    L.bgpu_encode_part.argtypes = [vp, vp, sz, vp, C.c_uint, i, i, i, vp, sz, szp]
    L.bgpu_encode_device.argtypes = [vp, vp, sz, i, i, i, vp, sz, szp, C.POINTER(C.c_float)]
    L.bgpu_decode.argtypes = [vp, vp, sz, vp, sz, szp]
    L.bgpu_decode_noalloc.argtypes = [i, vp, szp, vp, szp, szp, C.POINTER(C.c_int), vp]
    L.bgpu_noalloc_scratch_bytes.argtypes = []
    L.bgpu_noalloc_scratch_bytes.restype = sz
    L.bgpu_woff2_glyf_bound.argtypes = [sz, C.c_uint32]
    L.bgpu_woff2_glyf_bound.restype = sz
    L.bgpu_woff2_transform_glyf.argtypes = [vp, i, C.POINTER(GlyfTable), C.POINTER(vp), szp, szp]
    L.bgpu_stream_create.argtypes = [vp, i, i, i, C.POINTER(vp)]
    L.bgpu_stream_update.argtypes = [vp, vp, sz, C.POINTER(vp), szp]
    L.bgpu_stream_finish.argtypes = [vp, C.POINTER(vp), szp]
    L.bgpu_stream_destroy.argtypes = [vp]
    L.bgpu_stream_destroy.restype = None
    L.bgpu_stage_match_candidates.argtypes = [vp, vp, sz, i, i, vp, vp]
    L.bgpu_stage_commands.argtypes = [vp, vp, sz, i, i, i, vp, sz, szp]
    L.bgpu_stage_store_trivial.argtypes = [vp, vp, sz, vp, sz, i, i, C.c_uint, C.c_uint, vp, sz, szp,
                                           C.POINTER(C.c_uint64), vp]
    L.bgpu_stage_huffman.argtypes = [vp, vp, i, i, vp, vp, vp, vp]
    L.bgpu_stage_h10_matches.argtypes = [vp, vp, sz, i, vp, vp, vp, sz, szp]
    L.bgpu_stage_histograms.argtypes = [vp, vp, sz, vp, sz, i, vp, C.POINTER(MbModel), vp, vp, vp, C.POINTER(i)]
    L.bgpu_stage_context_map.argtypes = [vp, vp, sz, C.c_uint32, vp, sz, C.POINTER(C.c_uint64)]
    L.bgpu_stage_store_metablock.argtypes = [vp, vp, sz, vp, sz, i, i, C.c_uint, C.c_uint, vp, C.POINTER(MbModel), vp, sz, szp]
    L.bgpu_stage_timings.argtypes = [vp, C.POINTER(C.c_char_p), C.POINTER(C.c_float), i]
    L.bgpu_last_launch_count.argtypes = [vp]
    L.bgpu_last_counter.argtypes = [vp, i]
    L.bgpu_last_counter.restype = C.c_uint64
    L.bgpu_strerror.argtypes = [i]
    L.bgpu_strerror.restype = C.c_char_p
    L.bgpu_last_cuda_error.restype = C.c_char_p
    _lib = L
    return L


class BgpuError(RuntimeError):
    def __init__(self, code):
        L = load()
        self.code = code
        msg = L.bgpu_strerror(code).decode()
        detail = L.bgpu_last_cuda_error().decode()
        super().__init__(f"brotligpu: {msg} ({code})" + (f": {detail}" if detail and code in (-4, -5, -6, -8, -9) else ""))


def check(rc):
    if rc != 0:
        raise BgpuError(rc)
