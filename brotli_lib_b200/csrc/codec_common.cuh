// Format arithmetic shared by the kernels: command coding, Huffman code construction and
// prefix-code serialisation.  Everything here is __host__ __device__ so that the same code is
// exercised on the CPU (tests/test_hostdev_units.py via bgpu_hosttest_*) and inside the kernels.
//
// Reference semantics implemented (paths relative to the reference tree, src/encode/):
//   command.ts:11-135          insert/copy length codes, command code, distance prefix code
//   entropy-encode.ts:24-258   length-limited Huffman depths (two-queue merge, count_limit retry),
//                              canonical codes, bit reversal
//   context-map.ts:215-520     simple / complex prefix-code serialisation
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define HD __host__ __device__ __forceinline__
#define HDN __host__ __device__
#else
#define HD inline
#define HDN
#endif

namespace bgpu {

// 16-byte command record, identical to the reference's Command (command.ts:138-144).
struct Cmd {
  uint32_t insertLen;
  uint32_t copyLen;     // len | (delta & 0x7F) << 25
  uint32_t distExtra;
  uint16_t cmdPrefix;
  uint16_t distPrefix;  // code | nbits << 10
};

HD int log2_floor(uint32_t v) {
#ifdef __CUDA_ARCH__
  return 31 - __clz((int)v);
#else
  return 31 - __builtin_clz(v);
#endif
}

HD uint32_t ins_code(uint32_t insertLen) {  // command.ts:29-44
  if (insertLen < 6) return insertLen;
  if (insertLen < 130) { uint32_t nb = log2_floor(insertLen - 2) - 1; return (nb << 1) + ((insertLen - 2) >> nb) + 2; }
  if (insertLen < 2114) return log2_floor(insertLen - 66) + 10;
  if (insertLen < 6210) return 21;
  if (insertLen < 22594) return 22;
  return 23;
}
HD uint32_t copy_code(uint32_t copyLen) {  // command.ts:46-57
  if (copyLen < 10) return copyLen - 2;
  if (copyLen < 134) { uint32_t nb = log2_floor(copyLen - 6) - 1; return (nb << 1) + ((copyLen - 6) >> nb) + 4; }
  if (copyLen < 2118) return log2_floor(copyLen - 70) + 12;
  return 23;
}
// command.ts:11-27 as closed forms (base / extra-bit tables of RFC 7932 §5)
HD uint32_t ins_extra(uint32_t code) {
  const uint64_t lo = 0x5544332211000000ull;  // codes 0..15, 4 bits each
  if (code < 16) return (uint32_t)((lo >> (code * 4)) & 15);
  const uint8_t hi[8] = {6, 7, 8, 9, 10, 12, 14, 24};
  return hi[code - 16];
}
HD uint32_t ins_base(uint32_t code) {
  const uint32_t t[24] = {0, 1, 2, 3, 4, 5, 6, 8, 10, 14, 18, 26, 34, 50, 66, 98, 130, 194, 322, 578, 1090, 2114, 6210, 22594};
  return t[code];
}
HD uint32_t copy_extra(uint32_t code) {
  const uint64_t lo = 0x4433221100000000ull;
  if (code < 16) return (uint32_t)((lo >> (code * 4)) & 15);
  const uint8_t hi[8] = {5, 5, 6, 7, 8, 9, 10, 24};
  return hi[code - 16];
}
HD uint32_t copy_base(uint32_t code) {
  const uint32_t t[24] = {2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 18, 22, 30, 38, 54, 70, 102, 134, 198, 326, 582, 1094, 2118};
  return t[code];
}
HD uint32_t combine_codes(uint32_t ic, uint32_t cc, bool useLast) {  // command.ts:59-81
  uint32_t bits64 = (cc & 7) | ((ic & 7) << 3);
  if (useLast && ic < 8 && cc < 16) return cc < 8 ? bits64 : (bits64 | 64);
  uint32_t offset = 2 * ((cc >> 3) + 3 * (ic >> 3));
  offset = (offset << 5) + 0x40 + ((0x520D40u >> offset) & 0xC0);
  return offset | bits64;
}
// command.ts:111-135; returns code | nbits << 10, *extra = extra-bit value
HD uint32_t dist_prefix(uint32_t distanceCode, uint32_t ndirect, uint32_t npostfix, uint32_t* extra) {
  if (distanceCode < 16 + ndirect) { *extra = 0; return distanceCode; }
  uint32_t dist = (1u << (npostfix + 2)) + (distanceCode - 16 - ndirect);
  uint32_t bucket = log2_floor(dist) - 1;
  uint32_t postfix = dist & ((1u << npostfix) - 1);
  uint32_t prefix = (dist >> bucket) & 1;
  uint32_t offset = (2 + prefix) << bucket;
  uint32_t nbits = bucket - npostfix;
  *extra = (dist - offset) >> npostfix;
  return (nbits << 10) | (16 + ndirect + ((2 * (nbits - 1) + prefix) << npostfix) + postfix);
}
// Inverse of dist_prefix (command.ts:111-135): the backward distance a stored explicit distance code stands for.
HD uint32_t distance_of_explicit(uint32_t dp, uint32_t extra, uint32_t ndirect, uint32_t npostfix) {
  const uint32_t c = dp & 0x3FF;
  if (c < 16 + ndirect) return c - 15;
  const uint32_t v = c - 16 - ndirect;
  const uint32_t postfix = v & ((1u << npostfix) - 1), t = v >> npostfix, prefix = t & 1, nbits = (t >> 1) + 1;
  const uint32_t bucket = nbits + npostfix, offset = (2 + prefix) << bucket;
  const uint32_t dist = offset + (extra << npostfix) + postfix;
  return dist - (1u << (npostfix + 2)) + 16 + ndirect - 15;
}
// command.ts:146-179 (copyLenCodeDelta is always 0 on this path: no dictionary matches)
HD Cmd make_cmd(uint32_t insertLen, uint32_t copyLen, uint32_t distanceCode, uint32_t ndirect, uint32_t npostfix) {
  Cmd c;
  uint32_t extra;
  uint32_t dp = dist_prefix(distanceCode, ndirect, npostfix, &extra);
  c.insertLen = insertLen; c.copyLen = copyLen; c.distExtra = extra; c.distPrefix = (uint16_t)dp;
  c.cmdPrefix = (uint16_t)combine_codes(ins_code(insertLen), copy_code(copyLen), (dp & 0x3FF) == 0);
  return c;
}
// command.ts:183-208: insert-only command (copy length code 2, never decoded)
HD uint32_t insert_only_prefix(uint32_t insertLen) {
  uint32_t ic = ins_code(insertLen);
  return combine_codes(ic, 0, ic < 8);
}
HD Cmd make_insert_cmd(uint32_t insertLen) {
  Cmd c; c.insertLen = insertLen; c.copyLen = 2u << 25; c.distExtra = 0; c.distPrefix = 0;
  c.cmdPrefix = (uint16_t)insert_only_prefix(insertLen);
  return c;
}
// metablock.ts:273-287 storeCommandExtra: (nbits, value) of the packed insert+copy extra field
HD uint32_t cmd_extra(uint32_t insertLen, uint32_t copyLenField, uint64_t* value) {
  uint32_t mod = copyLenField >> 25;
  int32_t delta = (mod & 0x40) ? (int32_t)(mod | 0xFFFFFF80u) : (int32_t)mod;
  uint32_t lenCode = (uint32_t)((int32_t)(copyLenField & 0x1FFFFFF) + delta);
  uint32_t ic = ins_code(insertLen), cc = copy_code(lenCode);
  uint32_t ni = ins_extra(ic);
  *value = ((uint64_t)(lenCode - copy_base(cc)) << ni) | (uint64_t)(insertLen - ins_base(ic));
  return ni + copy_extra(cc);
}

// ------------------------------------------------------------------------------------------
// Bit sink over a zero-initialised u32 word array (LSB-first, bit-writer.ts:32-56).
struct BitSink {
  uint32_t* w; uint32_t pos;
  HD void put(uint32_t nbits, uint32_t value) {
    if (!nbits) return;
    uint32_t i = pos >> 5, o = pos & 31;
    w[i] |= value << o;
    if (o + nbits > 32) w[i + 1] |= value >> (32 - o);
    pos += nbits;
  }
};

// ------------------------------------------------------------------------------------------
// Huffman.  `sorted` holds the n non-zero symbols ordered by (max(count, count_limit) asc, symbol desc)
// — entropy-encode.ts:69-80,167-220 (total order, so any sort gives the reference's sequence).
// Scratch: cnt[2n+2] u32, lt/rt[2n+2] u16, nd[2n+2] u8.  Returns false if a leaf is deeper than
// tree_limit (entropy-encode.ts:118-124 then doubles count_limit and retries).
HDN inline bool huff_depths_from_sorted(const uint32_t* hist, const uint16_t* sorted, int n, uint32_t count_limit,
                                        int tree_limit, uint32_t* cnt, uint16_t* lt, uint16_t* rt, uint8_t* nd,
                                        uint8_t* depth_out) {
  for (int k = 0; k < n; k++) { uint32_t c = hist[sorted[k]]; cnt[k] = c > count_limit ? c : count_limit; }
  cnt[n] = 0xFFFFFFFFu; cnt[n + 1] = 0xFFFFFFFFu;
  int i = 0, j = n + 1;
  for (int k = n - 1; k > 0; k--) {
    int l, r;
    if (cnt[i] <= cnt[j]) l = i++; else l = j++;
    if (cnt[i] <= cnt[j]) r = i++; else r = j++;
    int jEnd = 2 * n - k;
    cnt[jEnd] = cnt[l] + cnt[r];  // u32 wrap == `>>> 0`
    lt[jEnd] = (uint16_t)l; rt[jEnd] = (uint16_t)r;
    cnt[jEnd + 1] = 0xFFFFFFFFu;
  }
  // entropy-encode.ts:133-165 setDepth: a leaf's depth is its level below the root (node 2n-1).
  // Children always have smaller indices than their parent, so one descending sweep suffices.
  int root = 2 * n - 1;
  nd[root] = 0;
  bool ok = true;
  for (int node = root; node > n; node--) {
    uint8_t d = nd[node] < 250 ? nd[node] + 1 : 250;  // saturate: only `> tree_limit` matters
    nd[lt[node]] = d; nd[rt[node]] = d;
  }
  for (int k = 0; k < n; k++) { if (nd[k] > tree_limit) ok = false; }
  if (!ok) return false;
  for (int k = 0; k < n; k++) depth_out[sorted[k]] = nd[k];
  return true;
}

// entropy-encode.ts:234-258 convertBitDepthsToSymbols
HDN inline void huff_codes(const uint8_t* depth, int len, uint16_t* bits) {
  uint16_t bl[16], next[16];
  for (int i = 0; i < 16; i++) bl[i] = 0;
  for (int i = 0; i < len; i++) bl[depth[i]]++;
  bl[0] = 0;
  uint32_t code = 0; next[0] = 0;
  for (int i = 1; i <= 15; i++) { code = (code + bl[i - 1]) << 1; next[i] = (uint16_t)code; }
  for (int i = 0; i < len; i++) {
    uint32_t d = depth[i];
    if (d) {
      uint32_t c = next[d]++;
      uint32_t r = 0;
      for (uint32_t b = 0; b < d; b++) r |= ((c >> b) & 1u) << (d - 1 - b);
      bits[i] = (uint16_t)r;
    } else bits[i] = 0;
  }
}

// Serial rank sort for small alphabets (code-length code, stage tests).
HDN inline int huff_sort_serial(const uint32_t* hist, int len, uint32_t count_limit, uint16_t* sorted) {
  int n = 0;
  for (int s = 0; s < len; s++) if (hist[s]) n++;
  for (int s = 0; s < len; s++) {
    if (!hist[s]) continue;
    uint32_t cs = hist[s] > count_limit ? hist[s] : count_limit;
    int rank = 0;
    for (int t = 0; t < len; t++) {
      if (!hist[t] || t == s) continue;
      uint32_t ct = hist[t] > count_limit ? hist[t] : count_limit;
      if (ct < cs || (ct == cs && t > s)) rank++;
    }
    sorted[rank] = (uint16_t)s;
  }
  return n;
}

// entropy-encode.ts:24-131 for the 18-symbol code-length code (tree_limit 5), fully serial.
HDN inline void huff_small(const uint32_t* hist, int len, int tree_limit, uint8_t* depth, uint16_t* bits) {
  uint16_t sorted[18]; uint32_t cnt[38]; uint16_t lt[38], rt[38]; uint8_t nd[38];
  for (int i = 0; i < len; i++) { depth[i] = 0; bits[i] = 0; }
  int nz = 0, last = 0;
  for (int i = 0; i < len; i++) if (hist[i]) { nz++; last = i; }
  if (nz == 0) return;
  if (nz == 1) { depth[last] = 1; return; }
  for (uint32_t limit = 1;; limit *= 2) {
    int n = huff_sort_serial(hist, len, limit, sorted);
    if (huff_depths_from_sorted(hist, sorted, n, limit, tree_limit, cnt, lt, rt, nd, depth)) break;
    for (int i = 0; i < len; i++) depth[i] = 0;
  }
  huff_codes(depth, len, bits);
}

// context-map.ts:350-473 RLE of code lengths into (symbol, extra) pairs.  Returns the count.
// rle_sym/rle_extra need alphabet + 8 entries.
HDN inline int tree_rle(const uint8_t* depth, int length, uint8_t* rle_sym, uint8_t* rle_extra) {
  int newLength = length;
  while (newLength > 0 && depth[newLength - 1] == 0) newLength--;
  int out = 0; uint32_t prev = 8; int i = 0;
  while (i < newLength) {
    uint32_t value = depth[i]; uint32_t reps = 1;
    while (i + (int)reps < newLength && depth[i + reps] == value) reps++;
    i += reps;
    if (value == 0) {
      if (reps == 11) { rle_sym[out] = 0; rle_extra[out++] = 0; reps--; }
      if (reps < 3) { for (uint32_t j = 0; j < reps; j++) { rle_sym[out] = 0; rle_extra[out++] = 0; } }
      else {
        int start = out; reps -= 3;
        for (;;) { rle_sym[out] = 17; rle_extra[out++] = reps & 7; reps >>= 3; if (reps == 0) break; reps--; }
        for (int a = start, b = out - 1; a < b; a++, b--) {
          uint8_t t = rle_sym[a]; rle_sym[a] = rle_sym[b]; rle_sym[b] = t;
          t = rle_extra[a]; rle_extra[a] = rle_extra[b]; rle_extra[b] = t;
        }
      }
    } else {
      if (prev != value) { rle_sym[out] = (uint8_t)value; rle_extra[out++] = 0; reps--; }
      if (reps == 7) { rle_sym[out] = (uint8_t)value; rle_extra[out++] = 0; reps--; }
      if (reps < 3) { for (uint32_t j = 0; j < reps; j++) { rle_sym[out] = (uint8_t)value; rle_extra[out++] = 0; } }
      else {
        int start = out; reps -= 3;
        for (;;) { rle_sym[out] = 16; rle_extra[out++] = reps & 3; reps >>= 2; if (reps == 0) break; reps--; }
        for (int a = start, b = out - 1; a < b; a++, b--) {
          uint8_t t = rle_sym[a]; rle_sym[a] = rle_sym[b]; rle_sym[b] = t;
          t = rle_extra[a]; rle_extra[a] = rle_extra[b]; rle_extra[b] = t;
        }
      }
      prev = value;
    }
  }
  return out;
}

// context-map.ts:215-347 buildAndStoreHuffmanTree, serialisation half.  `depth`/`bits` already hold
// the code for `alphabet` symbols with `count` (>= 2) non-zero histogram entries; s4 the first four.
// For count <= 1 use tree_store_single.  rle_* are scratch (alphabet + 8 bytes each).
template <class Sink>
HDN inline void tree_store(Sink& w, const uint8_t* depth, int alphabet, int count, const uint16_t* s4,
                           uint8_t* rle_sym, uint8_t* rle_extra) {
  uint32_t maxBits = 0; uint32_t c = (uint32_t)alphabet - 1;
  while (c) { c >>= 1; maxBits++; }
  if (count <= 4) {  // context-map.ts:266-291 (stable sort by depth)
    // The <= 4 used symbols are the ones with a non-zero depth (count >= 2: every used symbol has a code), in index
    // order; they are re-collected here into scalars and ranked (smaller depth first, ties by index) without any
    // indexed local array.  The earlier forms (insertion sort over a copy of s4; ranks over s4) read the caller's
    // 16-bit local s4[] and came out wrong on the device when inlined into the serial header coders (r4b-r4d:
    // 2- and 4-symbol codes lost or swapped symbols; correct on the host and in a printf build).
    (void)s4;
    uint32_t v0 = 0, v1 = 0, v2 = 0, v3 = 0, d0 = 255, d1 = 255, d2 = 255, d3 = 255; int m = 0;
    for (int i = 0; i < alphabet; i++) {
      const uint32_t d = depth[i];
      if (!d) continue;
      if (m == 0) { v0 = (uint32_t)i; d0 = d; } else if (m == 1) { v1 = (uint32_t)i; d1 = d; }
      else if (m == 2) { v2 = (uint32_t)i; d2 = d; } else if (m == 3) { v3 = (uint32_t)i; d3 = d; }
      m++;
    }
    // rank = number of symbols sorting before (index order is v0 < v1 < v2 < v3, so ties go to the earlier one); v3 takes the rank left over
    const uint32_t r0 = (d1 < d0) + (d2 < d0) + (d3 < d0);
    const uint32_t r1 = (d0 <= d1) + (d2 < d1) + (d3 < d1);
    const uint32_t r2 = (d0 <= d2) + (d1 <= d2) + (d3 < d2);
    w.put(2, 1); w.put(2, count - 1);
    uint32_t first_depth = 0;
    for (uint32_t r = 0; r < (uint32_t)count; r++) {
      const uint32_t sym = r0 == r ? v0 : r1 == r ? v1 : r2 == r ? v2 : v3;
      if (r == 0) first_depth = r0 == 0 ? d0 : r1 == 0 ? d1 : r2 == 0 ? d2 : d3;
      w.put(maxBits, sym);
    }
    if (count == 4) w.put(1, first_depth == 1 ? 1 : 0);
    return;
  }
  // context-map.ts:293-347
  int nr = tree_rle(depth, alphabet, rle_sym, rle_extra);
  uint32_t clh[18];
  for (int i = 0; i < 18; i++) clh[i] = 0;
  for (int i = 0; i < nr; i++) clh[rle_sym[i]]++;
  int numCodes = 0, firstCode = 0;
  for (int i = 0; i < 18; i++) if (clh[i]) { if (numCodes == 0) firstCode = i; numCodes++; }
  uint8_t cld[18]; uint16_t clb[18];
  huff_small(clh, 18, 5, cld, clb);
  {  // context-map.ts:485-520
    const uint8_t order[18] = {1, 2, 3, 4, 0, 5, 17, 6, 16, 7, 8, 9, 10, 11, 12, 13, 14, 15};
    const uint8_t sym[6] = {0, 7, 3, 2, 1, 15};
    const uint8_t len[6] = {2, 4, 3, 2, 2, 4};
    int codesToStore = 18;
    if (numCodes > 1) while (codesToStore > 0 && cld[order[codesToStore - 1]] == 0) codesToStore--;
    int skip = 0;
    if (cld[order[0]] == 0 && cld[order[1]] == 0) { skip = 2; if (cld[order[2]] == 0) skip = 3; }
    w.put(2, skip);
    for (int i = skip; i < codesToStore; i++) { uint32_t l = cld[order[i]]; w.put(len[l], sym[l]); }
  }
  if (numCodes == 1) cld[firstCode] = 0;
  for (int i = 0; i < nr; i++) {
    uint32_t code = rle_sym[i];
    w.put(cld[code], clb[code]);
    if (code == 16) w.put(2, rle_extra[i]);
    else if (code == 17) w.put(3, rle_extra[i]);
  }
}
// context-map.ts:236-243: zero or one used symbol
template <class Sink>
HD void tree_store_single(Sink& w, int alphabet, uint32_t symbol) {
  uint32_t maxBits = 0; uint32_t c = (uint32_t)alphabet - 1;
  while (c) { c >>= 1; maxBits++; }
  w.put(4, 1); w.put(maxBits, symbol);
}

// ------------------------------------------------------------------------------------------
// Serial coders of the metablock header (one thread per metablock in k_mb_header): small prefix codes
// (block types, block lengths, context-map symbols) and encodeContextMap.
template <class Sink>
HD void put_varlen8(Sink& w, uint32_t n) {  // context-map.ts:524-533
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
template <class Sink>
HDN inline void huff_serial_store(Sink& w, SerialHuff& S, int alphabet) {
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
HDN inline void encode_context_map(BitSink& w, SerialHuff& S, const uint32_t* cmap, uint32_t size, uint32_t ncl, uint32_t* scratch) {
  put_varlen8(w, ncl - 1);
  if (ncl == 1) return;
  // moveToFrontTransform (:13-50)
  uint8_t mtf[256];
  uint32_t maxv = 0;
  for (uint32_t i = 0; i < size; i++) maxv = cmap[i] > maxv ? cmap[i] : maxv;
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
    maxReps = reps > maxReps ? reps : maxReps;
  }
  uint32_t maxPrefix = maxReps > 0 ? (uint32_t)log2_floor(maxReps) : 0;
  maxPrefix = maxPrefix < 6u ? maxPrefix : 6u;
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

}  // namespace bgpu
