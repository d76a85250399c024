import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

DATA = os.path.join(ROOT, "tests", "golden", "data")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def corpus_file(name):
    with open(os.path.join(DATA, name), "rb") as f:
        return f.read()


class Xorshift32:
    """test/brotli.test.ts:247-255 of the reference (makeXorshift32)."""

    def __init__(self, seed):
        self.x = seed & 0xFFFFFFFF

    def next(self):
        x = self.x
        x ^= (x << 13) & 0xFFFFFFFF
        x ^= x >> 17
        x ^= (x << 5) & 0xFFFFFFFF
        self.x = x
        return x

    def bytes(self, n):  # test/brotli.test.ts:257-261 randomBytes
        return bytes(self.next() & 0xFF for _ in range(n))


FUZZ_SIZES = [0, 1, 2, 3, 4, 7, 15, 31, 63, 64, 65, 127, 255, 256, 257, 511, 1024, 2048]  # test/brotli.test.ts:266


def fuzz_inputs():
    """The reference's fuzz matrix (test/brotli.test.ts:264-281): one generator across all sizes."""
    g = Xorshift32(0xC0FFEE ^ 0xBEEF)
    out = []
    for size in FUZZ_SIZES:
        out.append((f"random{size}", g.bytes(size)))
        out.append((f"ramp{size}", bytes(i & 0xFF for i in range(size))))
    return out


def synthetic(name):
    if name == "zeros":
        return bytes(262144)
    if name == "quickfox_repeated":
        return (b"The quick brown fox jumps over the lazy dog" * 4096)[:176128]
    if name == "abc1000":
        return b"abcdefghij" * 1000
    if name == "ramp":
        return bytes(i & 0xFF for i in range(1024))
    if name == "font4k":  # test/brotli.test.ts:152-160: repetitive 4-byte records
        b = bytearray(4096)
        for i in range(0, 4096, 4):
            b[i] = i & 0xFF
            b[i + 1] = (i >> 8) & 0xFF
        return bytes(b)
    raise KeyError(name)


CORPUS = ["alice29.txt", "enc-ttf.bin", "html-content.bin", "mapsdatazrh", "backward65536", "random_org_10k.bin",
          "monkey", "ukkonooa", "cp852-utf8", "compressed_repeated"]
SYNTH = ["zeros", "quickfox_repeated", "abc1000", "ramp", "font4k"]


def load_input(name):
    return synthetic(name) if name in SYNTH else corpus_file(name)


@pytest.fixture(scope="session")
def gpu_ctx():
    import brotli_lib_b200 as B
    ctx = B.Context()
    yield ctx
    ctx.close()
