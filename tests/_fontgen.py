"""Synthetic glyf/loca tables for the WOFF2-transform tests: random simple glyphs (every coordinate-delta class of the
triplet encoding, repeated flags, short and long vectors, overlap bit), composites (every argument / transform form, with
and without instructions), empty glyphs, both loca formats.  Test infrastructure."""
import struct

import numpy as np


def _simple(rng, big):
    nc = int(rng.integers(1, 5))
    counts = [int(rng.integers(1, 40 if not big else 300)) for _ in range(nc)]
    if big and rng.random() < 0.3:
        counts[0] = int(rng.integers(760, 900))          # 255UInt16 three-byte form
    npts = sum(counts)
    ends, t = [], -1
    for c in counts:
        t += c; ends.append(t)
    cls = rng.integers(0, 7, npts)
    mag = np.array([0, 3, 60, 700, 1279, 4000, 20000])[cls]
    dx = (rng.integers(0, 2, npts) * 2 - 1) * rng.integers(0, mag + 1)
    dy = (rng.integers(0, 2, npts) * 2 - 1) * rng.integers(0, np.array([0, 3, 60, 700, 1279, 4000, 20000])[rng.integers(0, 7, npts)] + 1)
    dx[rng.random(npts) < 0.15] = 0
    dy[rng.random(npts) < 0.15] = 0
    xs, ys = np.cumsum(dx), np.cumsum(dy)
    # keep coordinates inside int16
    xs = np.clip(xs, -30000, 30000); ys = np.clip(ys, -30000, 30000)
    dx = np.diff(np.concatenate([[0], xs])); dy = np.diff(np.concatenate([[0], ys]))
    on = rng.integers(0, 2, npts)
    flags, xb, yb = [], bytearray(), bytearray()
    for i in range(npts):
        f = int(on[i])
        x, y = int(dx[i]), int(dy[i])
        if x == 0 and rng.random() < 0.8:
            f |= 16
        elif -255 <= x <= 255 and rng.random() < 0.8:
            f |= 2 | (16 if x >= 0 else 0); xb.append(abs(x))
        else:
            xb += struct.pack(">h", x)
        if y == 0 and rng.random() < 0.8:
            f |= 32
        elif -255 <= y <= 255 and rng.random() < 0.8:
            f |= 4 | (32 if y >= 0 else 0); yb.append(abs(y))
        else:
            yb += struct.pack(">h", y)
        flags.append(f)
    if rng.random() < 0.2:
        flags[0] |= 0x40                                  # OVERLAP_SIMPLE
    fb = bytearray()
    i = 0
    while i < npts:                                       # repeat-flag packing
        j = i
        while j + 1 < npts and flags[j + 1] == flags[i] and j - i < 255:
            j += 1
        if j > i and rng.random() < 0.9:
            fb += bytes([flags[i] | 8, j - i]); i = j + 1
        else:
            fb.append(flags[i]); i += 1
    instr = bytes(rng.integers(0, 256, int(rng.integers(0, 30)) if rng.random() < 0.5 else 0, dtype=np.uint8))
    box = (int(xs.min()), int(ys.min()), int(xs.max()), int(ys.max()))
    if rng.random() < 0.2:
        box = (box[0] - 1, box[1], box[2], box[3] + 5)    # a stored bbox that differs from the computed one -> explicit
    rec = struct.pack(">hhhhh", nc, *box) + struct.pack(f">{nc}H", *ends) + struct.pack(">H", len(instr)) + instr + bytes(fb) + bytes(xb) + bytes(yb)
    return rec


def _composite(rng):
    ncomp = int(rng.integers(1, 5))
    have_instr = rng.random() < 0.4
    body = bytearray()
    for k in range(ncomp):
        f = 0
        if rng.random() < 0.5:
            f |= 1
        form = int(rng.integers(0, 4))
        f |= [0, 8, 0x40, 0x80][form]
        if k + 1 < ncomp:
            f |= 0x20
        if have_instr and k + 1 == ncomp:
            f |= 0x100
        body += struct.pack(">HH", f, int(rng.integers(0, 500)))
        body += bytes(rng.integers(0, 256, 4 if f & 1 else 2, dtype=np.uint8))
        body += bytes(rng.integers(0, 256, [0, 2, 4, 8][form], dtype=np.uint8))
    rec = struct.pack(">hhhhh", -1, *[int(v) for v in rng.integers(-2000, 2000, 4)]) + bytes(body)
    if have_instr:
        ins = bytes(rng.integers(0, 256, int(rng.integers(0, 40)), dtype=np.uint8))
        rec += struct.pack(">H", len(ins)) + ins
    return rec


def make_font(seed, num_glyphs=200, index_format=1, big=False):
    """(glyf, loca, num_glyphs, index_format) of a synthetic TrueType outline set."""
    rng = np.random.default_rng(seed)
    glyf, offs = bytearray(), [0]
    for _ in range(num_glyphs):
        r = rng.random()
        rec = b"" if r < 0.1 else (_composite(rng) if r < 0.3 else _simple(rng, big))
        if r >= 0.1 and r < 0.12:
            rec = struct.pack(">hhhhh", 0, 0, 0, 0, 0)     # a record with zero contours
        glyf += rec
        pad = (-len(glyf)) % (2 if index_format == 0 else 4)
        glyf += bytes(pad)
        offs.append(len(glyf))
    if index_format == 0:
        if len(glyf) >= 131072:
            return make_font(seed, num_glyphs // 2, 0, big)
        loca = struct.pack(f">{num_glyphs + 1}H", *[o // 2 for o in offs])
    else:
        loca = struct.pack(f">{num_glyphs + 1}I", *offs)
    return bytes(glyf), loca, num_glyphs, index_format
