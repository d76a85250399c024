"""Seeded synthetic inputs of the BASELINE.json configs.  numpy only.

  mixed(n, seed_offset)     config 2 / 4: 4 MiB segments, type by i mod 8:
                            0-3 English-like text, 4-5 binary records, 6 byte-repeat structure,
                            7 incompressible xorshift-like bytes
  text(n)                   config 3: English-like text (Zipf vocabulary of 50 000 words)
  font_blobs(count)         config 5: WOFF2-transformed-glyf-like blobs, log-uniform 4..300 KiB
"""
import numpy as np

SEG = 1 << 22
_LETTERS = np.frombuffer(b"etaoinshrdlcumwfgypbvkjxqz", dtype=np.uint8)
_LETTER_P = np.array([12.7, 9.1, 8.2, 7.5, 7.0, 6.7, 6.3, 6.1, 6.0, 4.3, 4.0, 2.8, 2.8, 2.4, 2.4, 2.2, 2.0, 2.0, 1.9, 1.5,
                      1.0, 0.8, 0.15, 0.15, 0.1, 0.07])
_VOCAB = None


def _vocab():
    """50 000 synthetic words, length 1 + Poisson(4.5), letters by English unigram frequency; 1 % are
    2-3 byte UTF-8 sequences (seed 0x7E570001)."""
    global _VOCAB
    if _VOCAB is None:
        rng = np.random.default_rng(0x7E570001)
        nwords = 50000
        lens = 1 + rng.poisson(4.5, nwords)
        total = int(lens.sum())
        flat = _LETTERS[rng.choice(26, size=total, p=_LETTER_P / _LETTER_P.sum())].copy()
        starts = np.concatenate(([0], np.cumsum(lens)[:-1]))
        utf = rng.random(nwords) < 0.01
        for w in np.nonzero(utf)[0]:
            k = 2 if lens[w] < 3 or rng.random() < 0.5 else 3
            s = starts[w]
            if k == 2:
                flat[s] = 0xC3
                flat[s + 1] = 0x80 + rng.integers(0, 64)
            else:
                flat[s] = 0xE2
                flat[s + 1] = 0x80 + rng.integers(0, 4)
                flat[s + 2] = 0x80 + rng.integers(0, 64)
            lens[w] = k if lens[w] >= k else lens[w]
        ranks = np.arange(1, nwords + 1, dtype=np.float64)
        cdf = np.cumsum(ranks ** -1.07)
        _VOCAB = (flat, starts.astype(np.int64), lens.astype(np.int64), cdf / cdf[-1])
    return _VOCAB


def text(n, seed=0x7E570002):
    """English-like text: Zipf(1.07) words, sentences of 4 + Poisson(12) words, capitalised first word,
    punctuation, paragraph break every ~6 sentences."""
    flat, starts, lens, cdf = _vocab()
    rng = np.random.default_rng(seed)
    out = np.empty(n + 64, dtype=np.uint8)
    pos = 0
    while pos < n:
        nw = max(1024, min(1 << 20, (n - pos) // 5 + 1024))
        w = np.searchsorted(cdf, rng.random(nw))
        wl = lens[w]
        # separators: 1 = space, sentence ends get ". " / "? " etc.
        sent_end = rng.random(nw) < (1.0 / 16.0)
        comma = (~sent_end) & (rng.random(nw) < 0.06)
        para = sent_end & (rng.random(nw) < (1.0 / 6.0))
        sep_len = 1 + sent_end.astype(np.int64) + comma.astype(np.int64) + para.astype(np.int64)
        tot = wl + sep_len
        ends = np.cumsum(tot)
        begins = ends - tot
        total = int(ends[-1])
        buf = np.full(total, 0x20, dtype=np.uint8)
        # word bytes
        idx = np.repeat(starts[w] - begins, wl) + np.arange(total)[_mask_from_runs(begins, wl, total)]
        buf[_mask_from_runs(begins, wl, total)] = flat[idx]
        # punctuation directly after the word
        pe = begins + wl
        buf[pe[comma]] = ord(",")
        punct = np.where(rng.random(nw) < 0.85, ord("."), np.where(rng.random(nw) < 0.5, ord("?"), ord(";")))
        buf[pe[sent_end]] = punct[sent_end]
        pp = pe[para]
        buf[pp + 1] = 0x0A
        buf[pp + 2] = 0x0A
        # capitalise the word after a sentence end
        cap = np.nonzero(sent_end[:-1])[0] + 1
        cb = begins[cap]
        lower = (buf[cb] >= 0x61) & (buf[cb] <= 0x7A)
        buf[cb[lower]] -= 0x20
        take = min(total, n - pos)
        out[pos:pos + take] = buf[:take]
        pos += take
    return out[:n]


def _mask_from_runs(begins, lens, total):
    m = np.zeros(total + 1, dtype=np.int32)
    np.add.at(m, begins, 1)
    np.add.at(m, begins + lens, -1)
    return np.cumsum(m[:-1]) > 0


def records(n, seed):
    """32-byte records: u32 counter, u16 Zipf(1.2) type over 64 values, 2 zero bytes, 3 f32 random-walk
    values (1e-3 steps), 12 bytes of a 4 096-entry string-table index pattern."""
    rng = np.random.default_rng(seed)
    k = n // 32 + 1
    rec = np.zeros((k, 32), dtype=np.uint8)
    rec[:, 0:4] = np.arange(k, dtype="<u4").view(np.uint8).reshape(k, 4)
    p = np.arange(1, 65, dtype=np.float64) ** -1.2
    t = rng.choice(64, size=k, p=p / p.sum()).astype("<u2")
    rec[:, 4:6] = t.view(np.uint8).reshape(k, 2)
    walk = np.cumsum(rng.normal(0, 1e-3, size=(k, 3)), axis=0).astype("<f4")
    rec[:, 8:20] = walk.view(np.uint8).reshape(k, 12)
    table = rng.integers(0, 256, size=(4096, 12), dtype=np.uint8)
    rec[:, 20:32] = table[(rng.zipf(1.3, size=k) - 1) % 4096]
    return rec.reshape(-1)[:n]


def xorshift_bytes(n, seed):
    """Incompressible bytes (vectorised stand-in for the reference's xorshift32 generator,
    test/brotli.test.ts:247-261: same distribution, numpy PCG stream)."""
    return np.random.default_rng(seed).integers(0, 256, size=n, dtype=np.uint8)


def repeats(n, seed):
    """Copy spans of 64-4096 bytes from up to 3.9 MiB back with probability 0.5, else fresh text."""
    rng = np.random.default_rng(seed)
    src = text(n, seed ^ 0x5A5A)
    out = np.empty(n, dtype=np.uint8)
    pos = 0
    while pos < n:
        span = int(rng.integers(64, 4097))
        span = min(span, n - pos)
        if pos > 4096 and rng.random() < 0.5:
            back = int(rng.integers(span, min(pos, int(3.9 * (1 << 20))) + 1))
            out[pos:pos + span] = out[pos - back:pos - back + span]
        else:
            out[pos:pos + span] = src[pos:pos + span]
        pos += span
    return out


def mixed_segment(i, seed_offset=0, n=SEG):
    t = i % 8
    if t < 4:
        return text(n, 0x5EED0000 + i + seed_offset)
    if t < 6:
        return records(n, 0xB1A50000 + i + seed_offset)
    if t == 6:
        return repeats(n, 0x4E9E0000 + i + seed_offset)
    return xorshift_bytes(n, (0xC0FFEE ^ 0xBEEF ^ i) + seed_offset)


def mixed(n, seed_offset=0, out=None):
    """Config 2 (and config 4 with seed_offset = k * 0x100): concatenation of 4 MiB segments."""
    if out is None:
        out = np.empty(n, dtype=np.uint8)
    for i, pos in enumerate(range(0, n, SEG)):
        take = min(SEG, n - pos)
        out[pos:pos + take] = mixed_segment(i, seed_offset, take)
    return out


def font_blobs(count, seed=0xF0470001, lo=4096, hi=300 * 1024):
    """Config 5: blobs shaped like a WOFF2-transformed glyf table (7 sub-streams after a 36-byte header)."""
    rng = np.random.default_rng(seed)
    sizes = np.exp(rng.uniform(np.log(lo), np.log(hi), count)).astype(np.int64)
    blobs = []
    for sz in sizes:
        sz = int(sz)
        body = sz - 36
        parts = (np.array([0.03, 0.06, 0.14, 0.62, 0.05, 0.02, 0.08]) * body).astype(np.int64)
        parts[3] += body - parts.sum()
        hdr = np.zeros(36, dtype=np.uint8)
        hdr[4:32] = parts.astype(">u4").view(np.uint8)
        ncont = np.where(rng.random(parts[0] // 2 + 1) < 0.02, -1, rng.integers(1, 5, parts[0] // 2 + 1)).astype(">i2").view(np.uint8)[:parts[0]]
        npts = np.minimum(rng.geometric(1 / 12.0, parts[1]), 255).astype(np.uint8)
        pz = np.arange(1, 129, dtype=np.float64) ** -1.3
        flags = rng.choice(128, size=parts[2], p=pz / pz.sum()).astype(np.uint8)
        m = int(parts[3])
        u = rng.random(m)
        trip = np.where(u < 0.45, rng.integers(0, 32, m), rng.integers(0, 256, m)).astype(np.uint8)
        rep = np.nonzero(u >= 0.8)[0]
        rep = rep[rep >= 8]
        trip[rep] = trip[rep - rng.integers(1, 9, rep.size)]
        comp = np.tile(rng.integers(0, 256, 8, dtype=np.uint8), parts[4] // 8 + 1)[:parts[4]].copy()
        comp[::8] = rng.integers(0, 8, comp[::8].size)
        bbox = rng.integers(0, 256, parts[5], dtype=np.uint8)
        pi = np.arange(1, 201, dtype=np.float64) ** -1.1
        instr = np.repeat(rng.choice(200, size=parts[6] // 2 + 1, p=pi / pi.sum()).astype(np.uint8), rng.integers(1, 4, parts[6] // 2 + 1))[:parts[6]]
        if instr.size < parts[6]:
            instr = np.concatenate([instr, np.zeros(parts[6] - instr.size, dtype=np.uint8)])
        blobs.append(np.concatenate([hdr, ncont, npts, flags, trip, comp, bbox, instr]))
    return blobs
