"""GPU tier (-m gpu): the integer stages of the DEFAULT (full-metablock) path, bit for bit against the oracle on
identical inputs.

The float-driven decisions of storeMetaBlock (block splits, context maps) are taken from the oracle and handed to the
device through the C ABI (bgpu_stage_histograms / bgpu_stage_context_map / bgpu_stage_store_metablock); everything
behind them — per-(block type, context) histograms (metablock.ts:558-637), chooseContextMode (context.ts:179-227),
encodeContextMap (context-map.ts:114-170), block-split codes and switches (metablock.ts:161-220), prefix codes,
header and the emit loop (metablock.ts:651-760) — must then produce the oracle's bytes.  The kernels are the ones the
default bgpu_encode path runs (k_mb_*, k_blk_expand, k_ctx_hist*, k_cluster_sum, k_huff_mb, k_mb_header, k_tile_model,
k_tilebits2, k_offsets2, k_write2), through the same host phases (model_host.inc full_*)."""
import numpy as np
import pytest

import _oracle as O
from conftest import load_input

pytestmark = pytest.mark.gpu

# (html-content, cp852-utf8, compressed_repeated, quickfox_repeated, random_org_10k: the reference stores them trivially at
# every quality — covered by test_store_trivial_bit_exact_given_oracle_commands)
INPUTS = ["alice29.txt", "enc-ttf.bin", "mapsdatazrh", "cfg2_text", "cfg2_records", "cfg2_repeat", "cfg3_text", "cfg5_font", "cfg5_font_small"]


def _input(name):
    from brotli_lib_b200 import workloads as W
    if name == "cfg2_text":        # BASELINE config 2, segment types 0..3
        return W.mixed_segment(1, 0, 300000).tobytes()
    if name == "cfg2_records":     # types 4, 5
        return W.mixed_segment(4, 0, 300000).tobytes()
    if name == "cfg2_repeat":      # type 6: byte-repeat structure
        return W.mixed_segment(6, 0, 400000).tobytes()
    if name == "cfg3_text":        # config 3 generator
        return W.text(500000, 0x7E570002).tobytes()
    if name == "cfg5_font":        # config 5: one WOFF2-glyf-like blob
        return W.font_blobs(1, seed=0xF0470001, lo=120000, hi=120001)[0].tobytes()
    if name == "cfg5_font_small":  # config 5's small end
        return W.font_blobs(1, seed=0xF0470002, lo=9000, hi=9001)[0].tobytes()
    return load_input(name)


def _rt(stream, data):
    assert O.decode(stream, len(data), "system") == data
    if O.have_ref_decoder():
        assert O.decode(stream, len(data), "ref") == data


def _case(name, quality, mode):
    data = _input(name)
    ocm = O.commands(data, quality, 22, mode)
    model = O.metablock_model(data, ocm, quality, mode == 2)
    return data, ocm, model


@pytest.mark.parametrize("name", INPUTS)
@pytest.mark.parametrize("quality,mode", [(5, 0), (9, 2), (11, 0)])
def test_context_histograms_bit_exact(gpu_ctx, name, quality, mode):
    """a12: metablock.ts:558-637 + chooseContextMode, under the oracle's commands and block splits."""
    data, ocm, model = _case(name, quality, mode)
    if model is None:
        pytest.skip("the reference stores this metablock trivially (metablock.ts:518, 531)")
    lh, ch, dh, cmode = gpu_ctx.stage_histograms(data, ocm, dict(model, context_mode=-1), mode)
    olh, och, odh = O.context_histograms(data, ocm, model, mode == 2)
    assert cmode == model["context_mode"]
    assert np.array_equal(ch, och)
    assert np.array_equal(dh, odh)
    assert np.array_equal(lh, olh)
    assert int(lh.sum()) == int(ocm["insertLen"].sum())


@pytest.mark.parametrize("name", INPUTS)
@pytest.mark.parametrize("quality,mode", [(5, 0), (9, 2), (11, 0)])
def test_store_metablock_bit_exact_given_oracle_model(gpu_ctx, name, quality, mode):
    """a15 + a16: same commands + splits + context maps in => the oracle's storeMetaBlock bytes out, which are also
    the bytes of the oracle's whole one-shot stream."""
    data, ocm, model = _case(name, quality, mode)
    if model is None:
        pytest.skip("the reference stores this metablock trivially (metablock.ts:518, 531)")
    for cmap, ncl in ((model["literal_map"], model["literal_clusters"]), (model["distance_map"], model["distance_clusters"])):
        bits, nb = gpu_ctx.stage_context_map(cmap, ncl)
        obits, onb = O.encode_context_map(cmap, ncl)
        assert nb == onb and bits == obits
    s = gpu_ctx.stage_store_metablock(data, ocm, model, True, mode, 4, 0b1011)
    ostream, _ = O.store_metablock(data, ocm, model, True, mode == 2, 4, 0b1011)
    assert s == ostream
    assert s == O.encode(data, quality, 22, mode)
    _rt(s, data)


@pytest.mark.parametrize("name", ["alice29.txt", "enc-ttf.bin", "cfg2_records"])
def test_store_metablock_not_last_and_mid_stream(gpu_ctx, name):
    """ISLAST = 0 (no alignment), and a metablock in mid-stream: literal contexts from the two bytes before it."""
    data, ocm, model = _case(name, 5, 0)
    assert model is not None
    s = gpu_ctx.stage_store_metablock(data, ocm, model, False, 0, 4, 0b1011)
    ostream, obits = O.store_metablock(data, ocm, model, False, False, 4, 0b1011)
    assert s == ostream
    prev = (0x65, 0x20)   # prev1 = byte at -1, prev2 = byte at -2
    lh, ch, dh, _ = gpu_ctx.stage_histograms(data, ocm, model, 0, prev2=bytes([prev[1], prev[0]]))
    olh, och, odh = O.context_histograms(data, ocm, model, False, prev)
    assert np.array_equal(lh, olh) and np.array_equal(ch, och) and np.array_equal(dh, odh)
    s = gpu_ctx.stage_store_metablock(data, ocm, model, True, 0, 0, 0, prev2=bytes([prev[1], prev[0]]))
    ostream, _ = O.store_metablock(data, ocm, model, True, False, 0, 0, prev)
    assert s == ostream


def test_store_metablock_synthetic_models(gpu_ctx):
    """Models the splitter would not produce: many tiny blocks, 256 block types, 64 clusters, types revisited in
    every order (exercises the last / second-last block-type codes and every block-length code range)."""
    rng = np.random.default_rng(11)
    data = _input("cfg2_text")[:200000]
    ocm = O.commands(data, 5, 22, 0)
    counts = [int(ocm["insertLen"].sum()), len(ocm), int(((ocm["copyLen"] & 0x1FFFFFF) > 0)[ocm["cmdPrefix"] >= 128].sum())]
    for trial in range(6):
        splits = []
        for k in range(3):
            nt = [2, 256, 17, 3, 64, 5][(trial + k) % 6]
            n = counts[k]
            if trial == 5 and k == 1:
                splits.append((1, np.zeros(0, np.uint8), np.zeros(0, np.uint32)))   # one implicit block
                continue
            nb = int(min(n, [3, 900, 40, 2, 300, 9][(trial + 2 * k) % 6]))
            cuts = np.sort(rng.choice(np.arange(1, n), nb - 1, replace=False)) if nb > 1 else np.zeros(0, np.int64)
            lengths = np.diff(np.concatenate([[0], cuts, [n]])).astype(np.uint32)
            types = np.zeros(nb, np.uint8)
            seen = 1
            for b in range(1, nb):   # a type may only appear once all smaller ones did (first-appearance numbering)
                r = rng.random()
                if seen < nt and r < 0.5:
                    types[b] = seen; seen += 1
                else:
                    t = int(rng.integers(0, seen))
                    types[b] = t if t != types[b - 1] or seen == 1 else (t + 1) % seen
            if seen == 1:   # one block type never has more than one block (clusterBlocks merges neighbours of equal type)
                types, lengths = np.zeros(1, np.uint8), np.array([n], np.uint32)
            splits.append((seen, types, lengths))
        ntl, ntd = splits[0][0], splits[2][0]
        ncl_l = int(min(64, [1, 64, 7, 2, 33, 64][trial])); ncl_d = int(min(ntd * 4, [1, 5, 64, 2, 3, 16][trial]))
        lm = rng.integers(0, ncl_l, ntl * 64).astype(np.uint32); lm[:ncl_l] = np.arange(ncl_l)
        dm = rng.integers(0, ncl_d, ntd * 4).astype(np.uint32); dm[:ncl_d] = np.arange(ncl_d)
        if trial % 2:
            lm[ntl * 16:] = lm[ntl * 16]    # long runs -> zero-run RLE with the largest prefixes
        model = {"context_mode": trial % 4, "splits": splits, "literal_map": lm, "literal_clusters": ncl_l, "distance_map": dm,
                 "distance_clusters": ncl_d}
        lh, ch, dh, _ = gpu_ctx.stage_histograms(data, ocm, model, 0)
        olh, och, odh = O.context_histograms(data, ocm, model, False)
        assert np.array_equal(lh, olh) and np.array_equal(ch, och) and np.array_equal(dh, odh), trial
        s = gpu_ctx.stage_store_metablock(data, ocm, model, True, 0, 4, 0b1011)
        ostream, _ = O.store_metablock(data, ocm, model, True, False, 4, 0b1011)
        assert s == ostream, trial
        _rt(s, data)


def test_context_map_edge_cases(gpu_ctx):
    rng = np.random.default_rng(3)
    cases = [(np.zeros(64, np.uint32), 1), (np.arange(64, dtype=np.uint32) % 2, 2), (np.repeat(np.arange(4, dtype=np.uint32), 64), 4),
             (np.concatenate([np.zeros(16000, np.uint32), np.arange(256, dtype=np.uint32)]), 256),
             (rng.integers(0, 64, 16384).astype(np.uint32), 64), (np.concatenate([np.arange(3), np.zeros(700), [2, 1, 0] * 5]).astype(np.uint32), 3),
             (np.arange(256, dtype=np.uint32)[::-1].copy(), 256)]
    for t in range(120):   # few clusters: simple (2..4 symbol) prefix codes over the context-map alphabet
        ncl = 2 + t % 5
        n = int(rng.choice([4, 64, 128, 700]))
        m = rng.integers(0, ncl, n).astype(np.uint32) if t % 3 else np.repeat(rng.integers(0, ncl, n // 4 + 1), 4)[:n].astype(np.uint32)
        k = min(ncl, n)
        m[:k] = np.arange(k)
        cases.append((m, int(m.max()) + 1))
    for cmap, ncl in cases:
        bits, nb = gpu_ctx.stage_context_map(cmap, ncl)
        obits, onb = O.encode_context_map(cmap, ncl)
        assert nb == onb and bits == obits, (len(cmap), ncl)


def test_store_metablock_full_size_shard(gpu_ctx):
    """One 4 MiB shard of BASELINE config 2 (the size the timed path runs): 4096 tiles, hundreds of blocks."""
    from brotli_lib_b200 import workloads as W
    for seg, q in ((2, 5), (5, 5), (6, 9)):
        data = W.mixed_segment(seg, 0).tobytes()
        ocm = O.commands(data, q, 22, 0)
        model = O.metablock_model(data, ocm, q, False)
        assert model is not None
        lh, ch, dh, cmode = gpu_ctx.stage_histograms(data, ocm, dict(model, context_mode=-1), 0)
        olh, och, odh = O.context_histograms(data, ocm, model, False)
        assert cmode == model["context_mode"] and np.array_equal(lh, olh) and np.array_equal(ch, och) and np.array_equal(dh, odh)
        s = gpu_ctx.stage_store_metablock(data, ocm, model, True, 0, 4, 0b1011)
        assert s == O.encode(data, q, 22, 0)
