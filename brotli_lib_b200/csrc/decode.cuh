// On-device Brotli decoder for self-verification: RFC 7932 without the static dictionary
// (the encoder never emits a dictionary reference; one in the stream is reported as BGPU_DEC_DICTIONARY).
// Format source: RFC 7932 §§3-10; the reference's decoder is src/decode/engine.ts:1012-1518 (a port of
// Google's Java decoder) — this is written from the RFC, not from either.
//
// Decoding one stream is serial, so the unit of parallelism is the SHARD: the encoder's shards end byte-aligned
// and never reference bytes or distance-ring entries of another shard, so every shard of a
// device-resident stream is decoded by its own thread (one CTA each, lane 0 working; hundreds to thousands of
// shards per job keep the SMs busy with latency-bound serial work).  k_decode_shards either writes the bytes
// (bgpu_decode) or only compares them with the encoder's input, which then also serves as the copy history
// (BGPU_F_VERIFY: no output buffer, no device->host copy, no CPU decode).
#pragma once
#include "metablock_header.cuh"

namespace bgpu {

enum { BGPU_DEC_OK = 0, BGPU_DEC_TRUNCATED = 1, BGPU_DEC_FORMAT = 2, BGPU_DEC_DICTIONARY = 3, BGPU_DEC_MISMATCH = 4, BGPU_DEC_OUTPUT = 5 };

struct DecJob {
  uint64_t stream_off; uint64_t stream_len;   // bytes of this piece inside the stream buffer
  uint64_t out_off;    uint64_t out_len;      // where its decoded bytes belong / how many are expected (0 = unknown: decode to ISLAST)
  uint32_t flags;                             // 1: starts with the window header; 2: must end with ISLAST
  uint32_t pad;
};
struct DecResult { uint32_t status; uint32_t pad; uint64_t produced; uint64_t where; };

// Per-piece table arena: prefix codes as (first-level 8-bit table, canonical fallback) per tree.
constexpr int DEC_MAX_TREES = 256;
constexpr int DEC_LUT_BITS = 8;
struct DecTree { uint16_t count[16]; uint16_t first_sym_idx[16]; uint32_t sym_off; uint32_t alphabet; };
struct DecArena {
  uint16_t lut[3][DEC_MAX_TREES][1 << DEC_LUT_BITS];   // sym << 4 | len (len 0: not resolved in 8 bits)
  DecTree tree[3][DEC_MAX_TREES];
  uint16_t syms_lit[DEC_MAX_TREES * 256], syms_cmd[DEC_MAX_TREES * 704], syms_dist[DEC_MAX_TREES * 544];   // symbols sorted by (length, value), per tree at sym_off
  __device__ __forceinline__ uint16_t* syms(int c) { return c == 0 ? syms_lit : (c == 1 ? syms_cmd : syms_dist); }
  uint8_t cmap_lit[DEC_MAX_TREES * 64];
  uint8_t cmap_dist[DEC_MAX_TREES * 4];
  uint8_t lens[704 + 8];
  // block-switch codes: [category][0 = types, 1 = counts]
  uint16_t blut[3][2][1 << DEC_LUT_BITS]; DecTree btree[3][2]; uint16_t bsyms[3][2][258 + 2];
  uint16_t tlut[1 << DEC_LUT_BITS]; DecTree ttree; uint16_t tsyms[272 + 2];   // scratch code (context maps)
};

struct BitIn {
  const uint8_t* p; uint64_t n; uint64_t i; uint64_t val; uint32_t nb; bool over;
  __device__ __forceinline__ void init(const uint8_t* s, uint64_t len) { p = s; n = len; i = 0; val = 0; nb = 0; over = false; }
  __device__ __forceinline__ void fill() { while (nb <= 56) { const uint64_t b = i < n ? p[i] : 0; if (i >= n + 8) over = true; i++; val |= b << nb; nb += 8; } }
  __device__ __forceinline__ uint32_t peek(uint32_t k) { if (nb < k) fill(); return (uint32_t)(val & ((1ull << k) - 1)); }
  __device__ __forceinline__ void drop(uint32_t k) { val >>= k; nb -= k; }
  __device__ __forceinline__ uint32_t get(uint32_t k) { if (!k) return 0; const uint32_t v = peek(k); drop(k); return v; }
  __device__ __forceinline__ uint64_t consumed_bits() const { return i * 8 - nb; }
  __device__ __forceinline__ void align() { drop(nb & 7); }
  __device__ __forceinline__ bool truncated() const { return consumed_bits() > n * 8; }
};

// Canonical code from lens[0..alphabet): first-level table + (count, first index) per length + sorted symbols.
__device__ bool dec_build(const uint8_t* lens, uint32_t alphabet, uint16_t* lut, DecTree& T, uint16_t* syms) {
  for (int l = 0; l < 16; l++) T.count[l] = 0;
  for (uint32_t s = 0; s < alphabet; s++) T.count[lens[s]]++;
  T.count[0] = 0; T.alphabet = alphabet;
  uint32_t idx = 0, nz = 0;
  for (int l = 1; l < 16; l++) { T.first_sym_idx[l] = (uint16_t)idx; idx += T.count[l]; nz += T.count[l]; }
  uint16_t next[16];
  for (int l = 1; l < 16; l++) next[l] = T.first_sym_idx[l];
  for (uint32_t s = 0; s < alphabet; s++) if (lens[s]) syms[next[lens[s]]++] = (uint16_t)s;
  for (int i = 0; i < (1 << DEC_LUT_BITS); i++) lut[i] = 0;
  if (nz == 0) return false;
  if (nz == 1) { for (int i = 0; i < (1 << DEC_LUT_BITS); i++) lut[i] = (uint16_t)(syms[0] << 4); T.count[0] = 1; return true; }   // zero-bit code
  // canonical codes are MSB-first; the stream delivers them LSB-first, so table entries are indexed by the reversed code
  uint32_t code = 0; uint64_t space = 0;
  for (int l = 1; l < 16; l++) {
    for (uint32_t k = 0; k < T.count[l]; k++) {
      if (l <= DEC_LUT_BITS) {
        const uint32_t rev = __brev(code) >> (32 - l);
        for (uint32_t f = rev; f < (1u << DEC_LUT_BITS); f += 1u << l) lut[f] = (uint16_t)((syms[T.first_sym_idx[l] + k] << 4) | l);
      }
      code++;
    }
    space += (uint64_t)T.count[l] << (15 - l);
    code <<= 1;
  }
  return space == 32768;
}
__device__ __forceinline__ uint32_t dec_symbol(BitIn& br, const uint16_t* lut, const DecTree& T, const uint16_t* syms) {
  if (T.count[0]) return syms[0];                       // single-symbol code: no bits
  const uint32_t e = lut[br.peek(DEC_LUT_BITS)];
  if (e & 15) { br.drop(e & 15); return e >> 4; }
  // longer than 8 bits: canonical walk, one bit at a time (MSB-first code value)
  uint32_t code = 0, first = 0;
  for (int l = 1; l < 16; l++) {
    code |= br.get(1);
    const uint32_t c = T.count[l];
    if (code < first + c) return syms[T.first_sym_idx[l] + (code - first)];
    first = (first + c) << 1; code <<= 1;
  }
  return 0xFFFFu;
}

// RFC 7932 §3.4 / §3.5: reads one prefix code over `alphabet` symbols into (lut, T, syms).
__device__ bool dec_read_code(BitIn& br, uint32_t alphabet, uint8_t* lens, uint16_t* lut, DecTree& T, uint16_t* syms) {
  for (uint32_t s = 0; s < alphabet; s++) lens[s] = 0;
  const uint32_t hskip = br.get(2);
  if (hskip == 1) {
    const uint32_t nsym = br.get(2) + 1;
    uint32_t abits = 0; for (uint32_t c = alphabet - 1; c; c >>= 1) abits++;
    uint32_t sy[4] = {0, 0, 0, 0};
    for (uint32_t k = 0; k < nsym; k++) { sy[k] = br.get(abits); if (sy[k] >= alphabet) return false; }
    for (uint32_t a = 0; a < nsym; a++) for (uint32_t b = a + 1; b < nsym; b++) if (sy[a] == sy[b]) return false;
    if (nsym == 1) { lens[sy[0]] = 1; dec_build(lens, alphabet, lut, T, syms); T.count[0] = 1; return true; }
    if (nsym == 2) { lens[sy[0]] = 1; lens[sy[1]] = 1; }
    else if (nsym == 3) { lens[sy[0]] = 1; lens[sy[1]] = 2; lens[sy[2]] = 2; }
    else if (br.get(1)) { lens[sy[0]] = 1; lens[sy[1]] = 2; lens[sy[2]] = 3; lens[sy[3]] = 3; }
    else { lens[sy[0]] = 2; lens[sy[1]] = 2; lens[sy[2]] = 2; lens[sy[3]] = 2; }
    return dec_build(lens, alphabet, lut, T, syms);
  }
  // code-length code: 18 symbols in a fixed order, each with the fixed 2..4-bit code of §3.5
  const uint8_t order[18] = {1, 2, 3, 4, 0, 5, 17, 6, 16, 7, 8, 9, 10, 11, 12, 13, 14, 15};
  uint8_t cl[18]; for (int k = 0; k < 18; k++) cl[k] = 0;
  int space = 32, num = 0;
  for (int k = (int)hskip; k < 18 && space > 0; k++) {
    const uint32_t x = br.peek(4);
    uint32_t v, nb;
    if ((x & 3) == 0) { v = 0; nb = 2; } else if ((x & 3) == 2) { v = 3; nb = 2; } else if ((x & 3) == 1) { v = 4; nb = 2; }
    else if ((x & 7) == 3) { v = 2; nb = 3; } else if (x == 7) { v = 1; nb = 4; } else { v = 5; nb = 4; }
    br.drop(nb);
    cl[order[k]] = (uint8_t)v;
    if (v) { space -= 32 >> v; num++; }
  }
  if (num != 1 && space != 0) return false;
  // the code-length code itself (alphabet 18, max length 5): a 32-entry table is all it needs
  uint16_t clut[32];
  {
    DecTree C; uint16_t csy[18]; uint16_t big[1 << DEC_LUT_BITS];
    if (!dec_build(cl, 18, big, C, csy)) return false;
    for (int k = 0; k < 32; k++) clut[k] = big[k];
    if (C.count[0]) for (int k = 0; k < 32; k++) clut[k] = (uint16_t)(csy[0] << 4);
  }
  uint32_t symbol = 0, prev = 8, repeat = 0, repeat_len = 0; int sp = 32768;
  while (symbol < alphabet && sp > 0) {
    const uint32_t e = clut[br.peek(5)];
    br.drop(e & 15);
    const uint32_t c = e >> 4;
    if (c < 16) {
      repeat = 0; lens[symbol++] = (uint8_t)c;
      if (c) { prev = c; sp -= 32768 >> c; }
    } else {
      const uint32_t xb = c - 14, nl = c == 16 ? prev : 0;
      if (repeat_len != nl) { repeat = 0; repeat_len = nl; }
      const uint32_t old = repeat;
      if (repeat > 0) { repeat -= 2; repeat <<= xb; }
      repeat += br.get(xb) + 3;
      const uint32_t delta = repeat - old;
      if (symbol + delta > alphabet) return false;
      for (uint32_t k = 0; k < delta; k++) lens[symbol + k] = (uint8_t)repeat_len;
      symbol += delta;
      if (repeat_len) sp -= (int)(delta << (15 - repeat_len));
    }
    if (br.truncated()) return false;
  }
  if (sp != 0) {   // only a code with a single used symbol may leave the space open (zero-bit code)
    uint32_t nz = 0;
    for (uint32_t k = 0; k < alphabet; k++) nz += lens[k] ? 1 : 0;
    if (nz != 1) return false;
  }
  return dec_build(lens, alphabet, lut, T, syms);
}

__device__ __forceinline__ uint32_t dec_varlen8(BitIn& br) {   // 1..256 (§9.2 NBLTYPES / NTREES)
  if (!br.get(1)) return 1;
  const uint32_t n = br.get(3);
  return (1u << n) + 1 + br.get(n);
}

// §7.3 context map of `size` entries; returns the number of trees (0 on error).
__device__ uint32_t dec_context_map(BitIn& br, DecArena& A, uint8_t* cmap, uint32_t size) {
  const uint32_t ntrees = dec_varlen8(br);
  if (ntrees == 1) { for (uint32_t k = 0; k < size; k++) cmap[k] = 0; return 1; }
  uint32_t rlemax = 0;
  if (br.get(1)) rlemax = br.get(4) + 1;
  if (!dec_read_code(br, ntrees + rlemax, A.lens, A.tlut, A.ttree, A.tsyms)) return 0;
  for (uint32_t k = 0; k < size;) {
    const uint32_t s = dec_symbol(br, A.tlut, A.ttree, A.tsyms);
    if (s == 0xFFFFu || br.truncated()) return 0;
    if (s == 0) cmap[k++] = 0;
    else if (s <= rlemax) {
      uint32_t reps = (1u << s) + br.get(s);
      if (k + reps > size) return 0;
      while (reps--) cmap[k++] = 0;
    } else cmap[k++] = (uint8_t)(s - rlemax);
  }
  if (br.get(1)) {   // inverse move-to-front
    uint8_t mtf[256];
    for (int k = 0; k < 256; k++) mtf[k] = (uint8_t)k;
    for (uint32_t k = 0; k < size; k++) {
      const uint32_t idx = cmap[k]; const uint8_t v = mtf[idx];
      cmap[k] = v;
      for (uint32_t j = idx; j > 0; j--) mtf[j] = mtf[j - 1];
      mtf[0] = v;
    }
  }
  return ntrees;
}

struct DecBlocks { uint32_t ntypes, count, type, last2; };   // last2: second-to-last block type

__device__ __forceinline__ uint32_t dec_block_count(BitIn& br, DecArena& A, int cat) {
  const uint32_t c = dec_symbol(br, A.blut[cat][1], A.btree[cat][1], A.bsyms[cat][1]);
  if (c >= 26) return 0;
  return c_blen_offset[c] + br.get(c_blen_nbits[c]);
}
__device__ __forceinline__ bool dec_block_switch(BitIn& br, DecArena& A, int cat, DecBlocks& B) {
  const uint32_t tc = dec_symbol(br, A.blut[cat][0], A.btree[cat][0], A.bsyms[cat][0]);
  if (tc == 0xFFFFu) return false;
  uint32_t t = tc == 0 ? B.last2 : (tc == 1 ? B.type + 1 : tc - 2);
  if (t >= B.ntypes) t -= B.ntypes;
  if (t >= B.ntypes) return false;
  B.last2 = B.type; B.type = t;
  B.count = dec_block_count(br, A, cat);
  return B.count != 0;
}

// One piece (a whole stream, or one shard of a stream made by this library).  `out` may be NULL (verify only);
// `expected` may be NULL (decode only); the copy history is read from whichever exists (out first).
__device__ void dec_piece(const uint8_t* __restrict__ stream, const DecJob J, uint8_t* out, const uint8_t* expected, uint64_t out_cap,
                          DecArena& A, DecResult& R) {
  BitIn br; br.init(stream + J.stream_off, J.stream_len);
  const uint8_t* hist = out ? out : expected;
  uint64_t pos = J.out_off;             // absolute output position
  const uint64_t pos0 = J.out_off;
  int ring[4] = {16, 15, 11, 4}; int ridx = 0;   // ring[(ridx - 1 - k) & 3] = k-th last distance
  uint32_t status = BGPU_DEC_OK; uint32_t lgwin = 24;
  auto fail = [&](uint32_t s) { status = s; };
  if (J.flags & 1) {   // §9.1 WBITS
    if (!br.get(1)) lgwin = 16;
    else { uint32_t n = br.get(3); if (n) lgwin = 17 + n; else { n = br.get(3); if (n == 1) fail(BGPU_DEC_FORMAT); lgwin = n ? 8 + n : 17; } }
  }
  const uint64_t max_window = (1ull << lgwin) - 16;
  bool last = false;
  while (status == BGPU_DEC_OK && !last) {
    if (J.out_len && !(J.flags & 2) && pos - pos0 >= J.out_len && br.consumed_bits() + 8 > J.stream_len * 8) break;   // a shard that is not the last: all bytes done
    last = br.get(1);
    if (last && br.get(1)) break;                       // ISLASTEMPTY
    const uint32_t mnib = br.get(2);
    if (mnib == 3) {                                    // metadata block (the encoder's byte-aligning padding)
      if (br.get(1)) { fail(BGPU_DEC_FORMAT); break; }
      const uint32_t sb = br.get(2); uint32_t skip = 0;
      for (uint32_t k = 0; k < sb; k++) skip |= br.get(8) << (8 * k);
      if (sb) skip += 1;
      br.align();
      for (uint32_t k = 0; k < skip; k++) br.get(8);
      if (br.truncated()) fail(BGPU_DEC_TRUNCATED);
      if (J.out_len && !(J.flags & 2) && pos - pos0 >= J.out_len) break;
      continue;
    }
    uint32_t mlen = 0;
    for (uint32_t k = 0; k < mnib + 4; k++) mlen |= br.get(4) << (4 * k);
    mlen += 1;
    if (!last && br.get(1)) {                           // uncompressed
      br.align();
      for (uint32_t k = 0; k < mlen && status == BGPU_DEC_OK; k++) {
        const uint8_t b = (uint8_t)br.get(8);
        if (out) { if (pos >= out_cap) { fail(BGPU_DEC_OUTPUT); break; } out[pos] = b; }
        if (expected && expected[pos] != b) { fail(BGPU_DEC_MISMATCH); break; }
        pos++;
      }
      if (br.truncated()) fail(BGPU_DEC_TRUNCATED);
      continue;
    }
    // ---- §9.2 compressed metablock header
    DecBlocks B[3];
    for (int c = 0; c < 3 && status == BGPU_DEC_OK; c++) {
      B[c].ntypes = dec_varlen8(br); B[c].type = 0; B[c].last2 = 1; B[c].count = 1u << 24;
      if (B[c].ntypes >= 2) {
        if (!dec_read_code(br, B[c].ntypes + 2, A.lens, A.blut[c][0], A.btree[c][0], A.bsyms[c][0]) ||
            !dec_read_code(br, 26, A.lens, A.blut[c][1], A.btree[c][1], A.bsyms[c][1])) { fail(BGPU_DEC_FORMAT); break; }
        B[c].count = dec_block_count(br, A, c);
        if (!B[c].count) fail(BGPU_DEC_FORMAT);
      }
    }
    if (status != BGPU_DEC_OK) break;
    const uint32_t npostfix = br.get(2), ndirect = br.get(4) << npostfix;
    uint8_t cmode[DEC_MAX_TREES];
    for (uint32_t t = 0; t < B[0].ntypes; t++) cmode[t] = (uint8_t)br.get(2);
    const uint32_t ntl = dec_context_map(br, A, A.cmap_lit, B[0].ntypes * 64);
    const uint32_t ntd = ntl ? dec_context_map(br, A, A.cmap_dist, B[2].ntypes * 4) : 0;
    if (!ntl || !ntd) { fail(BGPU_DEC_FORMAT); break; }
    const uint32_t dalpha = 16 + ndirect + (48u << npostfix);
    const uint32_t ntree[3] = {ntl, B[1].ntypes, ntd}, alpha[3] = {256, 704, dalpha};
    for (int c = 0; c < 3 && status == BGPU_DEC_OK; c++) {
      uint32_t so = 0;
      for (uint32_t t = 0; t < ntree[c]; t++) {
        A.tree[c][t].sym_off = so;
        if (!dec_read_code(br, alpha[c], A.lens, A.lut[c][t], A.tree[c][t], A.syms(c) + so)) { fail(BGPU_DEC_FORMAT); break; }
        so += alpha[c];
      }
    }
    if (status != BGPU_DEC_OK) break;
    if (br.truncated()) { fail(BGPU_DEC_TRUNCATED); break; }
    // ---- §10 commands
    const uint64_t mb_end = pos + mlen;
    while (pos < mb_end && status == BGPU_DEC_OK) {
      if (B[1].count == 0 && !dec_block_switch(br, A, 1, B[1])) { fail(BGPU_DEC_FORMAT); break; }
      B[1].count--;
      const DecTree& CT = A.tree[1][B[1].type];
      const uint32_t cmd = dec_symbol(br, A.lut[1][B[1].type], CT, A.syms_cmd + CT.sym_off);
      if (cmd >= 704) { fail(BGPU_DEC_FORMAT); break; }
      const uint32_t cell = cmd >> 6;
      const uint8_t ibase[11] = {0, 0, 0, 0, 8, 8, 0, 16, 8, 16, 16}, cbase[11] = {0, 8, 0, 8, 0, 8, 16, 0, 16, 8, 16};
      const uint32_t icode = ibase[cell] + ((cmd >> 3) & 7), ccode = cbase[cell] + (cmd & 7);
      uint32_t ins = ins_base(icode) + br.get(ins_extra(icode));
      const uint32_t copy = copy_base(ccode) + br.get(copy_extra(ccode));
      if (pos + ins > mb_end) { fail(BGPU_DEC_FORMAT); break; }
      while (ins--) {
        if (B[0].count == 0 && !dec_block_switch(br, A, 0, B[0])) { fail(BGPU_DEC_FORMAT); break; }
        B[0].count--;
        // the stream starts with contexts (0, 0); a later shard's first literals see the previous shard's last bytes
        const bool start = (J.flags & 1) != 0;
        const uint32_t p1 = (start && pos == pos0) ? 0 : hist[pos - 1], p2 = (start && pos <= pos0 + 1) ? 0 : hist[pos - 2];
        const uint8_t* lut = c_ctx_lut + ((uint32_t)cmode[B[0].type] << 9);
        const uint32_t ctx = lut[p1] | lut[256 + p2];
        const uint32_t tr = A.cmap_lit[B[0].type * 64 + ctx];
        const DecTree& LT = A.tree[0][tr];
        const uint32_t lit = dec_symbol(br, A.lut[0][tr], LT, A.syms_lit + LT.sym_off);
        if (lit > 255) { fail(BGPU_DEC_FORMAT); break; }
        if (out) { if (pos >= out_cap) { fail(BGPU_DEC_OUTPUT); break; } out[pos] = (uint8_t)lit; }
        if (expected && expected[pos] != (uint8_t)lit) { fail(BGPU_DEC_MISMATCH); break; }
        pos++;
      }
      if (status != BGPU_DEC_OK) break;
      if (br.truncated()) { fail(BGPU_DEC_TRUNCATED); break; }
      if (pos >= mb_end) break;                         // the final command of a metablock may carry no copy
      uint32_t dist;
      if (cmd < 128) dist = (uint32_t)ring[(ridx - 1) & 3];
      else {
        if (B[2].count == 0 && !dec_block_switch(br, A, 2, B[2])) { fail(BGPU_DEC_FORMAT); break; }
        B[2].count--;
        const uint32_t dctx = copy > 4 ? 3 : copy - 2;
        const uint32_t tr = A.cmap_dist[B[2].type * 4 + dctx];
        const DecTree& DT = A.tree[2][tr];
        const uint32_t dc = dec_symbol(br, A.lut[2][tr], DT, A.syms_dist + DT.sym_off);
        if (dc >= dalpha) { fail(BGPU_DEC_FORMAT); break; }
        if (dc < 16) {
          const int8_t idx[16] = {0, 1, 2, 3, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1}, off[16] = {0, 0, 0, 0, -1, 1, -2, 2, -3, 3, -1, 1, -2, 2, -3, 3};
          const int d = ring[(ridx - 1 - idx[dc]) & 3] + off[dc];
          if (d <= 0) { fail(BGPU_DEC_FORMAT); break; }
          dist = (uint32_t)d;
        } else if (dc < 16 + ndirect) dist = dc - 15;
        else {
          const uint32_t v = dc - ndirect - 16, nb = 1 + (v >> (npostfix + 1)), h = v >> npostfix, lc = v & ((1u << npostfix) - 1);
          const uint32_t offs = ((2 + (h & 1)) << nb) - 4;
          dist = ((offs + br.get(nb)) << npostfix) + lc + ndirect + 1;
        }
        if (dc != 0) { ring[ridx & 3] = (int)dist; ridx++; }
      }
      const uint64_t avail = pos - pos0 < max_window ? pos - pos0 : max_window;   // a piece never reaches behind its own start
      if (dist > avail) { fail(BGPU_DEC_DICTIONARY); break; }   // beyond the data so far: a static-dictionary reference (not supported) or a corrupt stream
      if (pos + copy > mb_end) { fail(BGPU_DEC_FORMAT); break; }
      for (uint32_t k = 0; k < copy; k++) {
        const uint8_t b = hist[pos - dist];
        if (out) { if (pos >= out_cap) { fail(BGPU_DEC_OUTPUT); break; } out[pos] = b; }
        if (expected && expected[pos] != b) { fail(BGPU_DEC_MISMATCH); break; }
        pos++;
      }
      if (br.truncated()) fail(BGPU_DEC_TRUNCATED);
    }
  }
  if (status == BGPU_DEC_OK && J.out_len && pos - pos0 != J.out_len) status = BGPU_DEC_FORMAT;
  if (status == BGPU_DEC_OK && (J.flags & 2) && !last) status = BGPU_DEC_TRUNCATED;
  R.status = status; R.produced = pos - pos0; R.where = pos;
}

// grid = pieces; lane 0 of each CTA decodes one piece with its own table arena.
__global__ void __launch_bounds__(32) k_decode_shards(const uint8_t* __restrict__ stream, const DecJob* __restrict__ jobs, uint32_t n_jobs,
                                                      uint8_t* out, const uint8_t* expected, uint64_t out_cap, DecArena* arenas, DecResult* results) {
  const uint32_t j = blockIdx.x;
  if (j >= n_jobs || threadIdx.x != 0) return;
  dec_piece(stream, jobs[j], out, expected, out_cap, arenas[j], results[j]);
}

// Verification of a job still resident on the device (BGPU_F_VERIFY): shard j's compressed bytes (the encoder's own output
// region and byte count) are decoded and compared with the staged input, which also serves as the copy history.
__global__ void __launch_bounds__(32) k_verify_shards(const uint8_t* __restrict__ d_out, const ShardD* __restrict__ shards, const uint32_t* __restrict__ shard_bytes,
                                                      uint32_t n_shard, const uint8_t* d_in, DecArena* arenas, DecResult* results) {
  const uint32_t j = blockIdx.x;
  if (j >= n_shard || threadIdx.x != 0) return;
  const ShardD sh = shards[j];
  DecJob J; J.stream_off = sh.out_off; J.stream_len = shard_bytes[j]; J.out_off = sh.in_off + sh.hist; J.out_len = sh.len - sh.hist; J.flags = sh.flags & 3; J.pad = 0;
  dec_piece(d_out, J, nullptr, d_in, 0, arenas[j], results[j]);
}

}  // namespace bgpu
