"""ORACLE (test infrastructure) — CPU restatement of the WOFF2 glyf/loca transform and of its inverse.

SURVEY.md §8f-4: the transform is outside the reference (brotli-lib only points at WOFF2: README.md:63, and its
bench fixtures came from a sibling repository, bench/create-fixtures.ts:17), so there is no reference code and no
golden vector to pin this row — "parity unpinned".  What this file restates is the published algorithm: W3C WOFF 2.0,
section 5.1 "Transformed glyf table format" (header of 36 bytes, seven sub-streams, 255UInt16, triplet encoding of
coordinate deltas), section 5.3 (loca is reconstructed, its transformed length is 0).  The checks the tests build on it:
the CUDA transform equals `transform()` byte for byte, and `inverse()` of either gives back every glyph of the original
table (contours, points, on-curve flags, instructions, bounding box, composite records).

Pure Python loops: fine for the font fixtures (hundreds of KB); only tests/, smoke() and bench.py's checker legs import it.
"""
import struct


# ---------------------------------------------------------------- sfnt helpers
def sfnt_tables(data):
    """{tag: bytes} of an sfnt (TrueType) file."""
    n = struct.unpack(">H", data[4:6])[0]
    out = {}
    for i in range(n):
        tag, _, off, ln = struct.unpack(">4sIII", data[12 + 16 * i:28 + 16 * i])
        out[tag.decode("latin1")] = data[off:off + ln]
    return out


def glyf_inputs(data):
    """(glyf bytes, loca bytes, numGlyphs, indexToLocFormat) of an sfnt file."""
    t = sfnt_tables(data)
    num_glyphs = struct.unpack(">H", t["maxp"][4:6])[0]
    index_format = struct.unpack(">h", t["head"][50:52])[0]
    return t["glyf"], t["loca"], num_glyphs, index_format


def loca_offsets(loca, num_glyphs, index_format):
    if index_format == 0:
        return [2 * x for x in struct.unpack(f">{num_glyphs + 1}H", loca[: 2 * (num_glyphs + 1)])]
    return list(struct.unpack(f">{num_glyphs + 1}I", loca[: 4 * (num_glyphs + 1)]))


# ---------------------------------------------------------------- glyph parsing (OpenType glyf)
def parse_glyph(g):
    """dict for one glyph record (bytes): kind 'empty' | 'simple' | 'composite'."""
    if len(g) == 0:
        return {"kind": "empty"}
    nc, x0, y0, x1, y1 = struct.unpack(">hhhhh", g[:10])
    if nc == 0:
        return {"kind": "empty"}   # a record with zero contours carries no outline
    if nc < 0:
        pos, have_instr = 10, False
        while True:
            flags = struct.unpack(">H", g[pos:pos + 2])[0]
            size = 4 + (4 if flags & 1 else 2)                      # ARG_1_AND_2_ARE_WORDS
            size += 2 if flags & 8 else (4 if flags & 0x40 else (8 if flags & 0x80 else 0))   # scale / x-y scale / 2x2
            pos += size
            have_instr |= bool(flags & 0x100)
            if not flags & 0x20:                                    # MORE_COMPONENTS
                break
        comp = g[10:pos]
        instr = b""
        if have_instr:
            n = struct.unpack(">H", g[pos:pos + 2])[0]
            instr = g[pos + 2:pos + 2 + n]
        return {"kind": "composite", "bbox": (x0, y0, x1, y1), "components": comp, "have_instr": have_instr, "instr": instr}
    ends = struct.unpack(f">{nc}H", g[10:10 + 2 * nc])
    npts = ends[-1] + 1
    pos = 10 + 2 * nc
    ilen = struct.unpack(">H", g[pos:pos + 2])[0]
    instr = g[pos + 2:pos + 2 + ilen]
    pos += 2 + ilen
    flags = []
    while len(flags) < npts:
        f = g[pos]; pos += 1
        flags.append(f)
        if f & 8:
            r = g[pos]; pos += 1
            flags.extend([f] * r)
    flags = flags[:npts]
    xs, ys, v = [], [], 0
    for f in flags:
        if f & 2:
            d = g[pos]; pos += 1
            v += d if f & 16 else -d
        elif not f & 16:
            v += struct.unpack(">h", g[pos:pos + 2])[0]; pos += 2
        xs.append(v)
    v = 0
    for f in flags:
        if f & 4:
            d = g[pos]; pos += 1
            v += d if f & 32 else -d
        elif not f & 32:
            v += struct.unpack(">h", g[pos:pos + 2])[0]; pos += 2
        ys.append(v)
    return {"kind": "simple", "bbox": (x0, y0, x1, y1), "ends": list(ends), "points": [(xs[i], ys[i], flags[i] & 1) for i in range(npts)],
            "instr": instr, "overlap": bool(flags and flags[0] & 0x40)}


def parse_glyf(glyf, loca, num_glyphs, index_format):
    off = loca_offsets(loca, num_glyphs, index_format)
    return [parse_glyph(glyf[off[i]:off[i + 1]]) for i in range(num_glyphs)]


# ---------------------------------------------------------------- WOFF2 section 5.1 primitives
def put_255u16(out, v):
    if v < 253:
        out.append(v)
    elif v < 506:
        out += bytes([255, v - 253])
    elif v < 762:
        out += bytes([254, v - 506])
    else:
        out += bytes([253, v >> 8, v & 0xFF])


def put_triplet(flags, data, on_curve, x, y):
    ax, ay = abs(x), abs(y)
    oc = 0 if on_curve else 128
    xs, ys = (0 if x < 0 else 1), (0 if y < 0 else 1)
    xy = xs + 2 * ys
    if x == 0 and ay < 1280:
        flags.append(oc + ((ay & 0xF00) >> 7) + ys); data.append(ay & 0xFF)
    elif y == 0 and ax < 1280:
        flags.append(oc + 10 + ((ax & 0xF00) >> 7) + xs); data.append(ax & 0xFF)
    elif ax < 65 and ay < 65:
        flags.append(oc + 20 + ((ax - 1) & 0x30) + (((ay - 1) & 0x30) >> 2) + xy); data.append((((ax - 1) & 0xF) << 4) | ((ay - 1) & 0xF))
    elif ax < 769 and ay < 769:
        flags.append(oc + 84 + 12 * (((ax - 1) & 0x300) >> 8) + (((ay - 1) & 0x300) >> 6) + xy); data += bytes([(ax - 1) & 0xFF, (ay - 1) & 0xFF])
    elif ax < 4096 and ay < 4096:
        flags.append(oc + 120 + xy); data += bytes([ax >> 4, ((ax & 0xF) << 4) | (ay >> 8), ay & 0xFF])
    else:
        flags.append(oc + 124 + xy); data += bytes([ax >> 8, ax & 0xFF, ay >> 8, ay & 0xFF])


def transform(glyf, loca, num_glyphs, index_format):
    """Transformed glyf table (WOFF2 section 5.1) of the given glyf + loca tables."""
    glyphs = parse_glyf(glyf, loca, num_glyphs, index_format)
    ncont, npts, flg, gly, comp, bbox, ins = (bytearray() for _ in range(7))
    bitmap = bytearray(4 * ((num_glyphs + 31) // 32))
    overlap = bytearray((num_glyphs + 7) // 8)
    any_overlap = False
    for gid, g in enumerate(glyphs):
        if g["kind"] == "empty":
            ncont += struct.pack(">h", 0)
            continue
        explicit_bbox = True
        if g["kind"] == "composite":
            ncont += struct.pack(">h", -1)
            comp += g["components"]
            if g["have_instr"]:
                put_255u16(gly, len(g["instr"])); ins += g["instr"]
        else:
            ncont += struct.pack(">h", len(g["ends"]))
            prev = -1
            for e in g["ends"]:
                put_255u16(npts, e - prev); prev = e
            lx = ly = 0
            for (x, y, on) in g["points"]:
                put_triplet(flg, gly, on, x - lx, y - ly); lx, ly = x, y
            put_255u16(gly, len(g["instr"])); ins += g["instr"]
            pts = g["points"]
            calc = (min(p[0] for p in pts), min(p[1] for p in pts), max(p[0] for p in pts), max(p[1] for p in pts)) if pts else (0, 0, 0, 0)
            explicit_bbox = calc != g["bbox"]
            if g["overlap"]:
                overlap[gid >> 3] |= 0x80 >> (gid & 7); any_overlap = True
        if explicit_bbox:
            bitmap[gid >> 3] |= 0x80 >> (gid & 7)
            bbox += struct.pack(">hhhh", *g["bbox"])
    bbox_stream = bytes(bitmap) + bytes(bbox)
    hdr = struct.pack(">HHHHIIIIIII", 0, 1 if any_overlap else 0, num_glyphs, index_format, len(ncont), len(npts), len(flg), len(gly), len(comp), len(bbox_stream), len(ins))
    return hdr + bytes(ncont) + bytes(npts) + bytes(flg) + bytes(gly) + bytes(comp) + bbox_stream + bytes(ins) + (bytes(overlap) if any_overlap else b"")


# ---------------------------------------------------------------- inverse (WOFF2 section 5.1, decoder side)
def get_255u16(buf, pos):
    c = buf[pos]
    if c == 253:
        return (buf[pos + 1] << 8) | buf[pos + 2], pos + 3
    if c == 255:
        return 253 + buf[pos + 1], pos + 2
    if c == 254:
        return 506 + buf[pos + 1], pos + 2
    return c, pos + 1


def _sign(flag, v):
    return v if flag & 1 else -v


def get_triplet(flag, data, pos):
    on = not flag & 128
    f = flag & 127
    if f < 10:
        return on, 0, _sign(f, ((f & 14) << 7) + data[pos]), pos + 1
    if f < 20:
        return on, _sign(f, (((f - 10) & 14) << 7) + data[pos]), 0, pos + 1
    if f < 84:
        b0, b1 = f - 20, data[pos]
        return on, _sign(f, 1 + (b0 & 0x30) + (b1 >> 4)), _sign(f >> 1, 1 + ((b0 & 0x0C) << 2) + (b1 & 0x0F)), pos + 1
    if f < 120:
        b0 = f - 84
        return on, _sign(f, 1 + ((b0 // 12) << 8) + data[pos]), _sign(f >> 1, 1 + (((b0 % 12) >> 2) << 8) + data[pos + 1]), pos + 2
    if f < 124:
        b2 = data[pos + 1]
        return on, _sign(f, (data[pos] << 4) + (b2 >> 4)), _sign(f >> 1, ((b2 & 0x0F) << 8) + data[pos + 2]), pos + 3
    return on, _sign(f, (data[pos] << 8) + data[pos + 1]), _sign(f >> 1, (data[pos + 2] << 8) + data[pos + 3]), pos + 4


def inverse(t):
    """Glyph records (as parse_glyph returns them) reconstructed from a transformed glyf table."""
    _, opt, num_glyphs, index_format, s_nc, s_np, s_fl, s_gl, s_co, s_bb, s_in = struct.unpack(">HHHHIIIIIII", t[:36])
    p = 36
    ncont = t[p:p + s_nc]; p += s_nc
    npts = t[p:p + s_np]; p += s_np
    flg = t[p:p + s_fl]; p += s_fl
    gly = t[p:p + s_gl]; p += s_gl
    comp = t[p:p + s_co]; p += s_co
    bb = t[p:p + s_bb]; p += s_bb
    ins = t[p:p + s_in]; p += s_in
    overlap = t[p:p + (num_glyphs + 7) // 8] if opt & 1 else None
    bitmap_len = 4 * ((num_glyphs + 31) // 32)
    bitmap, bbv = bb[:bitmap_len], bb[bitmap_len:]
    i_np = i_fl = i_gl = i_co = i_bb = i_in = 0
    out = []
    for gid in range(num_glyphs):
        nc = struct.unpack(">h", ncont[2 * gid:2 * gid + 2])[0]
        explicit = bool(bitmap[gid >> 3] & (0x80 >> (gid & 7)))
        if nc == 0:
            out.append({"kind": "empty"})
            continue
        if nc < 0:
            start, have_instr = i_co, False
            while True:
                flags = struct.unpack(">H", comp[i_co:i_co + 2])[0]
                size = 4 + (4 if flags & 1 else 2)
                size += 2 if flags & 8 else (4 if flags & 0x40 else (8 if flags & 0x80 else 0))
                i_co += size
                have_instr |= bool(flags & 0x100)
                if not flags & 0x20:
                    break
            instr = b""
            if have_instr:
                n, i_gl = get_255u16(gly, i_gl)
                instr = ins[i_in:i_in + n]; i_in += n
            assert explicit
            box = struct.unpack(">hhhh", bbv[i_bb:i_bb + 8]); i_bb += 8
            out.append({"kind": "composite", "bbox": box, "components": comp[start:i_co], "have_instr": have_instr, "instr": instr})
            continue
        ends, tot = [], 0
        for _ in range(nc):
            n, i_np = get_255u16(npts, i_np)
            tot += n; ends.append(tot - 1)
        pts, x, y = [], 0, 0
        for k in range(tot):
            on, dx, dy, i_gl = get_triplet(flg[i_fl + k], gly, i_gl)
            x += dx; y += dy
            pts.append((x, y, 1 if on else 0))
        i_fl += tot
        n, i_gl = get_255u16(gly, i_gl)
        instr = ins[i_in:i_in + n]; i_in += n
        if explicit:
            box = struct.unpack(">hhhh", bbv[i_bb:i_bb + 8]); i_bb += 8
        else:
            box = (min(q[0] for q in pts), min(q[1] for q in pts), max(q[0] for q in pts), max(q[1] for q in pts)) if pts else (0, 0, 0, 0)
        out.append({"kind": "simple", "bbox": box, "ends": ends, "points": pts, "instr": instr,
                    "overlap": bool(overlap and overlap[gid >> 3] & (0x80 >> (gid & 7)))})
    assert (i_np, i_fl, i_gl, i_co, i_bb, i_in) == (s_np, s_fl, s_gl, s_co, len(bbv), s_in), "streams not consumed exactly"
    return out
