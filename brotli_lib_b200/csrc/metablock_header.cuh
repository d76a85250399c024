// Metablock header of the full metablock: everything between MLEN and the prefix codes
// (metablock.ts:161-255, 663-707; context-map.ts:13-170, 215-264, 524-533).  Serial per
// metablock (a few thousand symbols at most), bit-exact given block lists and context maps.
#pragma once
#include "metablock_model.cuh"

namespace bgpu {

// metablock.ts:37-64
__constant__ uint16_t c_blen_offset[26] = {1, 5, 9, 13, 17, 25, 33, 41, 49, 65, 81, 97, 113, 145, 177, 209, 241, 305, 369, 497, 753, 1265, 2289, 4337, 8433, 16625};
__constant__ uint8_t c_blen_nbits[26] = {2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 6, 6, 7, 8, 9, 10, 11, 12, 13, 24};

__device__ __forceinline__ uint32_t block_len_code(uint32_t len) {   // metablock.ts:118-125
  uint32_t code = len >= 177 ? (len >= 753 ? 20 : 14) : (len >= 41 ? 7 : 0);
  while (code < 25 && len >= c_blen_offset[code + 1]) code++;
  return code;
}
__device__ __forceinline__ void put_varlen8(BitSink& w, uint32_t n) {  // context-map.ts:524-533
  if (n == 0) { w.put(1, 0); return; }
  const uint32_t nbits = log2_floor(n);
  w.put(1, 1); w.put(3, nbits); w.put(nbits, n - (1u << nbits));
}

constexpr int SH_MAX = 272;   // largest alphabet built serially: block types + 2 <= 258
struct SerialHuff {
  uint16_t sorted[SH_MAX]; uint32_t cnt[2 * SH_MAX + 2]; uint16_t lt[2 * SH_MAX + 2], rt[2 * SH_MAX + 2];
  uint8_t nd[2 * SH_MAX + 2]; uint8_t rs[SH_MAX + 8], re[SH_MAX + 8];
  uint32_t hist[SH_MAX]; uint8_t depth[SH_MAX]; uint16_t bits[SH_MAX];
};
// buildAndStoreHuffmanTree (context-map.ts:215-264) of S.hist, one thread.
__device__ void huff_serial_store(BitSink& w, SerialHuff& S, int alphabet) {
  uint16_t s4[4] = {0, 0, 0, 0}; int n = 0;
  for (int i = 0; i < alphabet; i++) { S.depth[i] = 0; S.bits[i] = 0; if (S.hist[i]) { if (n < 4) s4[n] = (uint16_t)i; n++; } }
  if (n <= 1) { tree_store_single(w, alphabet, s4[0]); return; }
  for (uint32_t limit = 1;; limit *= 2) {
    huff_sort_serial(S.hist, alphabet, limit, S.sorted);
    if (huff_depths_from_sorted(S.hist, S.sorted, n, limit, 15, S.cnt, S.lt, S.rt, S.nd, S.depth)) break;
    for (int i = 0; i < alphabet; i++) S.depth[i] = 0;
  }
  huff_codes(S.depth, alphabet, S.bits);
  tree_store(w, S.depth, alphabet, n, s4, S.rs, S.re);
}

// encodeContextMap (context-map.ts:114-170).  scratch: `size` u32.
__device__ void encode_context_map(BitSink& w, SerialHuff& S, const uint32_t* cmap, uint32_t size, uint32_t ncl, uint32_t* scratch) {
  put_varlen8(w, ncl - 1);
  if (ncl == 1) return;
  // moveToFrontTransform (:13-50)
  uint8_t mtf[256];
  uint32_t maxv = 0;
  for (uint32_t i = 0; i < size; i++) maxv = max(maxv, cmap[i]);
  for (uint32_t i = 0; i <= maxv; i++) mtf[i] = (uint8_t)i;
  for (uint32_t i = 0; i < size; i++) {
    const uint32_t v = cmap[i]; uint32_t idx = 0;
    while (mtf[idx] != v) idx++;
    scratch[i] = idx;
    for (uint32_t j = idx; j > 0; j--) mtf[j] = mtf[j - 1];
    mtf[0] = (uint8_t)v;
  }
  // runLengthCodeZeros (:55-111), in place (the output never overtakes the input)
  uint32_t maxReps = 0;
  for (uint32_t i = 0; i < size;) {
    uint32_t reps = 0;
    while (i < size && scratch[i] != 0) i++;
    while (i < size && scratch[i] == 0) { reps++; i++; }
    maxReps = max(maxReps, reps);
  }
  uint32_t maxPrefix = maxReps > 0 ? (uint32_t)log2_floor(maxReps) : 0;
  maxPrefix = min(maxPrefix, 6u);
  uint32_t nout = 0;
  for (uint32_t i = 0; i < size;) {
    if (scratch[i] != 0) { scratch[nout++] = scratch[i] + maxPrefix; i++; }
    else {
      uint32_t reps = 1;
      for (uint32_t k = i + 1; k < size && scratch[k] == 0; k++) reps++;
      i += reps;
      while (reps != 0) {
        if (reps < (2u << maxPrefix)) { const uint32_t pfx = log2_floor(reps); scratch[nout++] = pfx | ((reps - (1u << pfx)) << 9); break; }
        scratch[nout++] = maxPrefix | (((1u << maxPrefix) - 1) << 9);
        reps -= (2u << maxPrefix) - 1;
      }
    }
  }
  const int alphabet = (int)(ncl + maxPrefix);
  for (int i = 0; i < alphabet; i++) S.hist[i] = 0;
  for (uint32_t i = 0; i < nout; i++) S.hist[scratch[i] & 511]++;
  const bool useRle = maxPrefix > 0;
  w.put(1, useRle ? 1 : 0);
  if (useRle) w.put(4, maxPrefix - 1);
  huff_serial_store(w, S, alphabet);
  for (uint32_t i = 0; i < nout; i++) {
    const uint32_t sym = scratch[i] & 511, extra = scratch[i] >> 9;
    w.put(S.depth[sym], S.bits[sym]);
    if (sym > 0 && sym <= maxPrefix) w.put(sym, extra);
  }
  w.put(1, 1);   // IMTF
}

// One CTA per metablock; thread 0 does the serial work.  M.misc must be zeroed.
__global__ void __launch_bounds__(32) k_mb_header(const MbModel* __restrict__ models, Params P) {
  __shared__ SerialHuff S;
  __shared__ uint8_t s_tdepth[258]; __shared__ uint16_t s_tbits[258];
  if (threadIdx.x != 0) return;
  const MbModel& M = models[blockIdx.x];
  BitSink w{M.misc, 0};
  // storeCompressedMetaBlockHeader (metablock.ts:233-255)
  const bool isLast = M.is_last != 0;
  w.put(1, isLast ? 1 : 0);
  if (isLast) w.put(1, 0);
  {
    const int lg = M.len == 1 ? 1 : log2_floor(M.len - 1) + 1;
    const int mnibbles = (lg < 16 ? 16 : lg + 3) / 4;
    w.put(2, mnibbles - 4); w.put(mnibbles * 4, M.len - 1);
  }
  if (!isLast) w.put(1, 0);
  // buildAndStoreBlockSplitCode x 3 (metablock.ts:161-202, 690-692)
  for (int cat = 0; cat < 3; cat++) {
    const uint32_t ntypes = M.sstate[cat][2], nb = M.sstate[cat][1];
    if (ntypes <= 1) { w.put(1, 0); continue; }
    const uint32_t* bs = M.blk_start[cat]; const uint8_t* bt = M.blk_type[cat];
    uint32_t lenHist[26];
    for (int i = 0; i < 26; i++) lenHist[i] = 0;
    for (uint32_t i = 0; i < ntypes + 2; i++) S.hist[i] = 0;
    for (uint32_t b = 0; b < nb; b++) {
      if (b) {
        const uint32_t last = bt[b - 1], second = b >= 2 ? bt[b - 2] : 1, t = bt[b];
        S.hist[t == last + 1 ? 1 : (t == second ? 0 : t + 2)]++;
      }
      lenHist[block_len_code(bs[b + 1] - bs[b])]++;
    }
    put_varlen8(w, ntypes - 1);
    huff_serial_store(w, S, (int)ntypes + 2);
    for (uint32_t i = 0; i < ntypes + 2; i++) { s_tdepth[i] = S.depth[i]; s_tbits[i] = S.bits[i]; }
    for (int i = 0; i < 26; i++) S.hist[i] = lenHist[i];
    huff_serial_store(w, S, 26);
    // block switches (metablock.ts:204-220); the first one carries no type code
    for (uint32_t b = 0; b < nb; b++) {
      const uint32_t len = bs[b + 1] - bs[b], lc = block_len_code(len);
      uint64_t v = 0; uint32_t nbits = 0;
      if (b) {
        const uint32_t last = bt[b - 1], second = b >= 2 ? bt[b - 2] : 1, t = bt[b];
        const uint32_t tc = t == last + 1 ? 1 : (t == second ? 0 : t + 2);
        v = s_tbits[tc]; nbits = s_tdepth[tc];
      }
      v |= (uint64_t)S.bits[lc] << nbits; nbits += S.depth[lc];
      v |= (uint64_t)(len - c_blen_offset[lc]) << nbits; nbits += c_blen_nbits[lc];
      if (b == 0) { w.put(min(nbits, 32u), (uint32_t)(v & 0xFFFFFFFFu)); if (nbits > 32) w.put(nbits - 32, (uint32_t)(v >> 32)); }
      M.blk_tok[cat][b] = ((uint64_t)nbits << 56) | v;
    }
  }
  w.put(2, P.npostfix); w.put(4, P.ndirect >> P.npostfix);                       // metablock.ts:695-696
  for (uint32_t i = 0; i < M.sstate[LIT][2]; i++) w.put(2, M.cnt[3]);            // context mode per literal block type
  const bool trivial = M.cnt[4] != 0;
  encode_context_map(w, S, M.cmap[LIT], trivial ? 1 : M.sstate[LIT][2] * 64, M.cstate[LIT][2], M.scratch);
  encode_context_map(w, S, M.cmap[DIST], trivial ? 1 : M.sstate[DIST][2] * 4, M.cstate[DIST][2], M.scratch);
  M.cnt[5] = w.pos;
  M.cnt[6] = M.cstate[LIT][2]; M.cnt[7] = M.cstate[DIST][2];
}

}  // namespace bgpu
