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
// encodeContextMap alone (bgpu_stage_context_map): the same device function k_mb_header calls.  out must be zeroed.
// (no __restrict__ on out: the BitSink built over it is also written by non-inlined callees — with the qualifier the
// compiler kept words of `out` in registers across those calls and bits of the symbol list were lost, r4c)
__global__ void __launch_bounds__(32) k_stage_context_map(const uint32_t* cmap, uint32_t size, uint32_t ncl, uint32_t* scratch,
                                                          uint32_t* out, uint32_t* nbits) {
  __shared__ SerialHuff S;
  if (threadIdx.x != 0) return;
  BitSink w{out, 0};
  encode_context_map(w, S, cmap, size, ncl, scratch);
  *nbits = w.pos;
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
