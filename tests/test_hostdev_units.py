"""CPU tier: the __host__ __device__ format code the kernels run (brotli_lib_b200/csrc/codec_common.cuh),
compiled for the host, against the oracle; plus the C-ABI surface check (no compute without a GPU)."""
import ctypes as C
import os
import re

import numpy as np

import _oracle as O
from conftest import ROOT, load_input


def _hosttest():
    from brotli_lib_b200 import build
    return C.CDLL(build.build_hosttest())


def test_huffman_and_tree_serialisation_match_oracle():
    H = _hosttest()
    rng = np.random.default_rng(1)
    for trial in range(300):
        alpha = int(rng.choice([256, 704, 64, 124, 18, 26, 272, 10]))
        kind = trial % 6
        if kind == 0:
            h = rng.integers(0, 1000, alpha)
        elif kind == 1:
            h = (rng.random(alpha) < 0.1) * rng.integers(1, 100000, alpha)
        elif kind == 2:
            h = np.zeros(alpha, dtype=np.int64)
            idx = rng.choice(alpha, size=min(alpha, int(rng.integers(0, 6))), replace=False)
            h[idx] = rng.integers(1, 50, len(idx))
        elif kind == 3:
            h = 2 ** rng.integers(0, 31, alpha)
        elif kind == 4:
            h = np.ones(alpha, dtype=np.int64)
        else:
            h = np.array([int(1.6 ** i) % (2 ** 32 - 1) + 1 for i in range(alpha)])
        h = np.ascontiguousarray(h, dtype=np.uint32)
        d = np.zeros(alpha, np.uint8)
        b = np.zeros(alpha, np.uint16)
        tw = np.zeros(256, np.uint32)
        nb = H.bgpu_hosttest_huffman(h.ctypes.data, alpha, d.ctypes.data, b.ctypes.data, tw.ctypes.data)
        od, ob, obytes, onb = O.build_store_tree(h)
        assert nb == onb and tw.tobytes()[: (nb + 7) // 8] == obytes
        assert np.array_equal(d, od) and np.array_equal(b, ob)


def test_command_coding_matches_oracle_commands():
    H = _hosttest()
    out = np.zeros(1, dtype=O.CMD_DTYPE)
    val = C.c_uint64(0)
    H.bgpu_hosttest_cmd_extra.restype = C.c_uint32
    for name, q, mode in (("alice29.txt", 5, 0), ("enc-ttf.bin", 9, 2), ("enc-ttf.bin", 11, 0)):
        data = load_input(name)
        cm = O.commands(data, q, 22, mode)
        npostfix, ndirect = (1, 12) if mode == 2 else (0, 0)
        for c in cm[:: max(1, len(cm) // 3000)]:
            copy_len = int(c["copyLen"]) & 0x1FFFFFF
            if copy_len == 0:
                H.bgpu_hosttest_make_insert_cmd(int(c["insertLen"]), out.ctypes.data)
                assert out[0] == c
                continue
            # invert the distance prefix code (command.ts:225-248) to recover the distance code
            dcode = int(c["distPrefix"]) & 0x3FF
            if dcode < 16 + ndirect:
                code = dcode
            else:
                nbits = int(c["distPrefix"]) >> 10
                h = (dcode - ndirect - 16) >> npostfix
                lo = (dcode - ndirect - 16) & ((1 << npostfix) - 1)
                offset = ((2 + (h & 1)) << nbits) - 4
                code = ((offset + int(c["distExtra"])) << npostfix) + lo + ndirect + 16
            H.bgpu_hosttest_make_cmd(int(c["insertLen"]), copy_len, code, ndirect, npostfix, out.ctypes.data)
            assert out[0] == c
            assert H.bgpu_hosttest_cmd_extra(int(c["insertLen"]), int(c["copyLen"]), C.byref(val)) <= 48


def test_context_map_coder_matches_oracle():
    """encodeContextMap (context-map.ts:114-170) as k_mb_header runs it (codec_common.cuh encode_context_map), compiled
    for the host: move-to-front, zero-run RLE with every prefix size, the prefix code over clusters + RLE symbols."""
    H = _hosttest()
    H.bgpu_hosttest_context_map.restype = C.c_uint32
    rng = np.random.default_rng(9)
    cases = [(np.zeros(64, np.uint32), 1), (np.arange(64, dtype=np.uint32) % 2, 2),
             (np.concatenate([np.zeros(16000, np.uint32), np.arange(256, dtype=np.uint32)]), 256),
             (rng.integers(0, 64, 16384).astype(np.uint32), 64),
             (np.concatenate([np.arange(3), np.zeros(700), [2, 1, 0] * 5]).astype(np.uint32), 3)]
    for t in range(200):
        ncl = int(rng.integers(2, 65)) if t % 2 else 2 + t % 5
        n = int(rng.choice([4, 64, 256, 1024, 4096]))
        m = rng.integers(0, ncl, n).astype(np.uint32) if t % 3 else np.repeat(rng.integers(0, ncl, n // 8 + 1), 8)[:n].astype(np.uint32)
        k = min(ncl, n)
        m[:k] = np.arange(k)
        cases.append((m, int(m.max()) + 1))
    for cmap, ncl in cases:
        w = np.zeros(cmap.size + 1024, np.uint32)
        nb = H.bgpu_hosttest_context_map(cmap.ctypes.data, cmap.size, ncl, w.ctypes.data)
        obits, onb = O.encode_context_map(cmap, ncl)
        assert nb == onb and w.tobytes()[: (nb + 7) // 8] == obits, (cmap.size, ncl)


def test_c_abi_library_exports_every_declared_symbol():
    from brotli_lib_b200 import _lib
    header = open(os.path.join(ROOT, "include", "brotligpu.h")).read()
    declared = sorted(set(re.findall(r"\b(bgpu_[a-z0-9_]+)\s*\(", header)))
    assert set(declared) == set(_lib.SYMBOLS)
    assert os.path.exists(_lib.LIB_PATH), "run python -m brotli_lib_b200.build"
    L = C.CDLL(_lib.LIB_PATH)
    for sym in declared:
        assert hasattr(L, sym), sym
    L.bgpu_max_compressed_size.restype = C.c_size_t
    L.bgpu_max_compressed_size.argtypes = [C.c_size_t]
    assert (1 << 20) + 1024 <= L.bgpu_max_compressed_size(1 << 20) <= (1 << 20) * 1.11
    assert L.bgpu_max_compressed_size(0) >= 2 and L.bgpu_max_compressed_size(63) >= 63 + 6
    L.bgpu_strerror.restype = C.c_char_p
    assert b"quality" in L.bgpu_strerror(-3)


def test_node_addon_sources_are_real_and_well_formed():
    """node/addon.cc binds the C ABI 1:1.  No Node toolchain exists in this image, so the addon compiles
    to nothing here (guarded by __has_include(<node_api.h>)); against a declaration-only stub of the Node-API subset it
    uses, the whole file must pass the compiler's syntax and type checks, and it must name only declared C-ABI entries."""
    import subprocess
    addon = os.path.join(ROOT, "node", "addon.cc")
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-Wall", "-Werror", "-I", os.path.join(ROOT, "tests", "stubs"), "-I", os.path.join(ROOT, "include"), addon])
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", addon])   # without node_api.h: an empty translation unit
    header = open(os.path.join(ROOT, "include", "brotligpu.h")).read()
    declared = set(re.findall(r"\b(bgpu_[a-z0-9_]+)\s*\(", header))
    used = set(re.findall(r"\b(bgpu_[a-z0-9_]+)\s*\(", open(addon).read()))
    assert used and used <= declared, used - declared
    assert {"bgpu_encode", "bgpu_encode_batch", "bgpu_encode_part", "bgpu_stream_update", "bgpu_stream_finish", "bgpu_ctx_create_multi"} <= used
    ts = open(os.path.join(ROOT, "node", "encode.ts")).read()
    for name in ("export function brotliEncode", "export class BrotliEncoder", "export const enum EncoderMode", "export interface BrotliEncodeOptions"):
        assert name in ts, name


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "brotli_lib_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in text and "_oracle" not in text and "oracle/" not in text.replace("the oracle", ""), f


def test_distance_of_explicit_inverts_dist_prefix():
    """k_stitch's ring hand-over re-codes commands from their stored (prefix, extra bits): the inverse of
    command.ts:111-135 must return the distance make_cmd was given, in GENERIC and FONT distance parameters."""
    H = _hosttest()
    H.bgpu_hosttest_distance_of_explicit.restype = C.c_uint32
    out = np.zeros(1, dtype=O.CMD_DTYPE)
    rng = np.random.default_rng(3)
    dists = list(range(1, 300)) + [2 ** k + d for k in range(8, 22) for d in (-1, 0, 1)] + [int(x) for x in rng.integers(1, (1 << 22) - 16, 2000)] + [(1 << 22) - 16]
    for ndirect, npostfix in ((0, 0), (12, 1)):
        for d in dists:
            H.bgpu_hosttest_make_cmd(3, 7, d + 15, ndirect, npostfix, out.ctypes.data)
            got = H.bgpu_hosttest_distance_of_explicit(int(out["distPrefix"][0]), int(out["distExtra"][0]), ndirect, npostfix)
            assert got == d, (ndirect, npostfix, d, got)
