"""CPU tier: the WOFF2 glyf-transform restatement (oracle/woff2_glyf.py) — inverse(transform(table)) gives back every
glyph of the table, on the real font fixture (the reference's bench/fixtures/enc-ttf.bin) and on synthetic tables."""
import os
import struct
import sys

from conftest import ROOT, load_input

sys.path.insert(0, os.path.join(ROOT, "oracle"))
import woff2_glyf as WG   # noqa: E402
import _fontgen as FG     # noqa: E402


def _check(glyf, loca, ng, fmt):
    t = WG.transform(glyf, loca, ng, fmt)
    assert WG.inverse(t) == WG.parse_glyf(glyf, loca, ng, fmt)
    hdr = struct.unpack(">HHHHIIIIIII", t[:36])
    assert hdr[0] == 0 and hdr[2] == ng and hdr[3] == fmt and hdr[4] == 2 * ng
    assert 36 + sum(hdr[4:]) + ((ng + 7) // 8 if hdr[1] & 1 else 0) == len(t)
    return t


def test_real_font_round_trip():
    glyf, loca, ng, fmt = WG.glyf_inputs(load_input("enc-ttf.bin"))
    t = _check(glyf, loca, ng, fmt)
    assert ng == 2547 and len(t) < len(glyf) + len(loca)      # the transform also shrinks the tables before entropy coding


def test_synthetic_fonts_round_trip():
    for seed in range(12):
        _check(*FG.make_font(seed, 150, seed % 2, big=seed % 3 == 0))


def test_255u16_and_triplet_forms():
    for v in list(range(0, 800)) + [761, 762, 1000, 65535]:
        b = bytearray(); WG.put_255u16(b, v)
        assert WG.get_255u16(b, 0) == (v, len(b))
    for x in (-20000, -4096, -4095, -769, -768, -65, -64, -1, 0, 1, 64, 65, 768, 769, 1279, 1280, 4095, 4096, 30000):
        for y in (-30000, -1280, -1279, -769, -64, 0, 1, 64, 65, 768, 1279, 1280, 4095, 4096):
            for on in (0, 1):
                f, d = bytearray(), bytearray()
                WG.put_triplet(f, d, on, x, y)
                got = WG.get_triplet(f[0], d, 0)
                assert got == (bool(on), x, y, len(d)), (x, y, on, got)
