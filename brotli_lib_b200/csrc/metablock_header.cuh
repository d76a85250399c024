// Metablock header of the full metablock: everything between MLEN and the prefix codes
// (metablock.ts:161-255, 663-707; context-map.ts:13-170, 215-264, 524-533).  Serial per
// metablock (a few thousand symbols at most), bit-exact given block lists and context maps.
#pragma once
#include "metablock_model.cuh"
#ifdef HDR_TIMING
#include <cstdio>
#endif

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

// Append-only bit writer for ONE thread: bits collect in a register and whole words leave with a single OR (shared
// memory: plain read-modify-write by the owning thread; global: atomicOr, fire-and-forget) — where BitSink does a
// dependent load + store per put.  `pos` may start anywhere (the first word is shared with whatever precedes it).
template <bool GLOBAL>
struct WordSink {
  uint32_t* w; uint32_t pos; uint64_t acc; uint32_t nacc, widx;
  __device__ __forceinline__ void init(uint32_t* dst, uint32_t bitpos) { w = dst; pos = bitpos; widx = bitpos >> 5; nacc = bitpos & 31; acc = 0; }
  __device__ __forceinline__ void store(uint32_t i, uint32_t v) { if (!v) return; if (GLOBAL) atomicOr(&w[i], v); else w[i] |= v; }
  __device__ __forceinline__ void put(uint32_t nbits, uint32_t value) {
    if (!nbits) return;
    acc |= (uint64_t)value << nacc; nacc += nbits; pos += nbits;
    if (nacc >= 32) { store(widx++, (uint32_t)acc); acc >>= 32; nacc -= 32; }
  }
  __device__ __forceinline__ void flush() { if (nacc) store(widx, (uint32_t)acc); }   // call once, after the last put
};

constexpr int HDR_WARPS = 5;            // block-split codes of the 3 categories, literal context map, distance context map
constexpr int HDR_PART_WORDS = 96;      // varlen8 + code over <= 258 types (<= 2150 bits) + code over 26 lengths (<= 290) + first token (<= 54)
constexpr int HDR_TREE_PART_WORDS = 80; // RLEMAX bits + one code over <= 272 symbols
constexpr uint32_t HDR_DSCR = 1024;     // distance context map entries (<= 256 types x 4)
constexpr uint32_t HDR_LSCR = 16384;    // literal context map entries (<= 256 types x 64)
struct HdrShared {
  SerialHuff S[HDR_WARPS];
  uint32_t part[3][HDR_PART_WORDS];           // block-split code of each category
  uint32_t tree[2][HDR_TREE_PART_WORDS];      // context maps: everything before the symbol list
  uint32_t part_bits[3], tree_bits[2], sym_bits[2], nout[2], max_prefix[2];
  uint32_t len_hist[3][26];
  uint8_t tdepth[3][258]; uint16_t tbits[3][258];
  uint16_t lscr[HDR_LSCR], dscr[HDR_DSCR];    // MTF / zero-run scratch of the two context maps
};

// encodeContextMap (context-map.ts:114-170) by one warp, in two steps.  cm_prepare: move-to-front (the list lives in
// registers, 32 entries per row; the longest zero run is tracked on the way), zero-run tokens + histogram + prefix code
// by lane 0 over shared memory; everything up to the symbol list goes to `tree` (shared memory) and the length of the
// symbol list is summed from the histogram, so the caller knows every bit offset before a single symbol is written.
// cm_emit: the warp writes the symbols at their final place, 32 per round (bit offsets by a warp scan).
// scratch[i] (16 bits, shared memory) ends as symbol | extra << 9.
template <int ROWS>
__device__ uint32_t cm_mtf(const uint32_t* __restrict__ cmap, uint32_t size, uint16_t* scratch) {
  const uint32_t lane = threadIdx.x & 31;
  uint32_t m[ROWS];
#pragma unroll
  for (int r = 0; r < ROWS; r++) m[r] = (uint32_t)r * 32 + lane;
  uint32_t run = 0, maxReps = 0;
  for (uint32_t i0 = 0; i0 < size; i0 += 32) {
    const uint32_t mine = i0 + lane < size ? cmap[i0 + lane] : 0;
    const uint32_t cnt = min(32u, size - i0);
    uint32_t res = 1;
    for (uint32_t k = 0; k < cnt; k++) {
      const uint32_t v = __shfl_sync(0xffffffffu, mine, k);
      uint32_t idx = 0xFFFFFFFFu;
#pragma unroll
      for (int r = 0; r < ROWS; r++) {
        if (idx == 0xFFFFFFFFu) {   // warp-uniform
          const uint32_t bal = __ballot_sync(0xffffffffu, m[r] == v);
          if (bal) idx = (uint32_t)r * 32 + (uint32_t)__ffs((int)bal) - 1;
        }
      }
      if (idx != 0) {   // warp-uniform; index 0 leaves the list as it is
        uint32_t carry = v;   // the value that enters position r * 32
#pragma unroll
        for (int r = 0; r < ROWS; r++) {
          if ((uint32_t)r * 32 <= idx) {   // warp-uniform
            const uint32_t up = __shfl_up_sync(0xffffffffu, m[r], 1);
            const uint32_t last = __shfl_sync(0xffffffffu, m[r], 31);
            if ((uint32_t)r * 32 + lane <= idx) m[r] = lane ? up : carry;
            carry = last;
          }
        }
      }
      if (lane == k) res = idx;
    }
    if (i0 + lane < size) scratch[i0 + lane] = (uint16_t)res;
    // zero runs of this chunk (lanes >= cnt hold 1): every lane walks the same mask
    uint32_t z = __ballot_sync(0xffffffffu, res == 0);
    uint32_t pos = 0;
    while (pos < 32) {
      const uint32_t rest = z >> pos;
      if (rest & 1u) {
        const uint32_t l = min((uint32_t)__ffs((int)~rest) - 1u, 32u - pos);   // ~rest has a set bit below 32 - pos or the run reaches the chunk end
        run += l; pos += l;
        if (pos < 32) { maxReps = max(maxReps, run); run = 0; }
      } else {
        maxReps = max(maxReps, run); run = 0;   // a run carried over from the previous chunk ends here
        if (!rest) { pos = 32; break; }
        pos += (uint32_t)__ffs((int)rest) - 1u;
      }
    }
    if (!(z >> 31)) { maxReps = max(maxReps, run); run = 0; }
  }
  return max(maxReps, run);
}

__device__ void cm_prepare(SerialHuff& S, const uint32_t* __restrict__ cmap, uint32_t size, uint32_t ncl, uint16_t* scratch,
                           uint32_t* tree, uint32_t& tree_bits, uint32_t& sym_bits, uint32_t& nout_o, uint32_t& max_prefix_o) {
  const uint32_t lane = threadIdx.x & 31;
  for (uint32_t i = lane; i < HDR_TREE_PART_WORDS; i += 32) tree[i] = 0;
  __syncwarp();
  if (ncl == 1) {
    if (lane == 0) { WordSink<false> w; w.init(tree, 0); put_varlen8(w, 0); w.flush(); tree_bits = w.pos; sym_bits = 0; nout_o = 0; max_prefix_o = 0; }
    return;
  }
  // moveToFrontTransform (:13-50)
  const uint32_t maxReps = ncl <= 32 ? cm_mtf<1>(cmap, size, scratch) : ncl <= 64 ? cm_mtf<2>(cmap, size, scratch)
                         : ncl <= 128 ? cm_mtf<4>(cmap, size, scratch) : cm_mtf<8>(cmap, size, scratch);
  __syncwarp();
  if (lane != 0) return;
  // runLengthCodeZeros (:55-111), in place (the output never overtakes the input)
  uint32_t maxPrefix = maxReps > 0 ? (uint32_t)log2_floor(maxReps) : 0;
  maxPrefix = maxPrefix < 6u ? maxPrefix : 6u;
  const int alphabet = (int)(ncl + maxPrefix);
  for (int i = 0; i < alphabet; i++) S.hist[i] = 0;
  uint32_t nout = 0;
  for (uint32_t i = 0; i < size;) {
    const uint32_t x = scratch[i];
    if (x != 0) { const uint32_t sy = x + maxPrefix; scratch[nout++] = (uint16_t)sy; S.hist[sy]++; i++; }
    else {
      uint32_t reps = 1;
      for (uint32_t k = i + 1; k < size && scratch[k] == 0; k++) reps++;
      i += reps;
      while (reps != 0) {
        if (reps < (2u << maxPrefix)) { const uint32_t pfx = log2_floor(reps); scratch[nout++] = (uint16_t)(pfx | ((reps - (1u << pfx)) << 9)); S.hist[pfx]++; break; }
        scratch[nout++] = (uint16_t)(maxPrefix | (((1u << maxPrefix) - 1) << 9)); S.hist[maxPrefix]++;
        reps -= (2u << maxPrefix) - 1;
      }
    }
  }
  WordSink<false> w; w.init(tree, 0);
  put_varlen8(w, ncl - 1);
  const bool useRle = maxPrefix > 0;
  w.put(1, useRle ? 1 : 0);
  if (useRle) w.put(4, maxPrefix - 1);
  huff_serial_store(w, S, alphabet);
  w.flush();
  uint32_t sb = 0;
  for (int sy = 0; sy < alphabet; sy++) sb += S.hist[sy] * (S.depth[sy] + ((uint32_t)sy > 0 && (uint32_t)sy <= maxPrefix ? (uint32_t)sy : 0u));
  tree_bits = w.pos; sym_bits = sb + 1; nout_o = nout; max_prefix_o = maxPrefix;   // + 1: the IMTF bit
}
__device__ void cm_emit(const SerialHuff& S, const uint16_t* scratch, uint32_t nout, uint32_t maxPrefix, uint32_t ncl, uint32_t* dst, uint32_t bitpos) {
  if (ncl == 1) return;
  const uint32_t lane = threadIdx.x & 31;
  uint32_t pos = bitpos;
  for (uint32_t i0 = 0; i0 < nout; i0 += 32) {
    uint64_t v = 0; uint32_t nb = 0;
    if (i0 + lane < nout) {
      const uint32_t t = scratch[i0 + lane], sym = t & 511, extra = t >> 9;
      v = S.bits[sym]; nb = S.depth[sym];
      if (sym > 0 && sym <= maxPrefix) { v |= (uint64_t)extra << nb; nb += sym; }
    }
    uint32_t incl = nb;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
    if (nb) {
      const uint32_t bp = pos + incl - nb, o = bp & 31;
      const uint64_t sv = v << o;   // nb <= 21, o <= 31: fits 64 bits
      atomicOr(&dst[bp >> 5], (uint32_t)sv);
      if ((uint32_t)(sv >> 32)) atomicOr(&dst[(bp >> 5) + 1], (uint32_t)(sv >> 32));
    }
    pos += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) atomicOr(&dst[pos >> 5], 1u << (pos & 31));   // IMTF
}

// OR-copy of nbits bits of src (shared memory) to bit position pos of dst (global, zeroed) by `nthr` threads.
__device__ __forceinline__ void hdr_copy(uint32_t* dst, uint32_t pos, const uint32_t* src, uint32_t nbits, uint32_t t, uint32_t nthr) {
  const uint32_t nw = (nbits + 31) >> 5;
  for (uint32_t k = t; k < nw; k += nthr) {
    uint32_t v = src[k];
    const uint32_t rem = nbits - k * 32;
    if (rem < 32) v &= (1u << rem) - 1u;
    if (!v) continue;
    const uint32_t bp = pos + k * 32, i = bp >> 5, o = bp & 31;
    atomicOr(&dst[i], v << o);
    if (o) atomicOr(&dst[i + 1], v >> (32 - o));
  }
}

// One CTA per metablock, five warps: warps 0-2 build the block-split code of one category each (histograms and
// block-switch tokens lane-parallel, the two small prefix codes by lane 0), warps 3 and 4 prepare the literal and the
// distance context map.  After one barrier every bit offset is known and the pieces go to M.misc (zeroed) at their
// final places.  (Round 1: one thread per metablock did all of it in sequence — 5.4 ms per job whatever its size, r5b.)
__global__ void __launch_bounds__(32 * HDR_WARPS) k_mb_header(const MbModel* __restrict__ models, Params P) {
  extern __shared__ __align__(16) uint8_t hdr_smem[];
  HdrShared& H = *reinterpret_cast<HdrShared*>(hdr_smem);
  const MbModel& M = models[blockIdx.x];
  const uint32_t tid = threadIdx.x, lane = tid & 31, wp = tid >> 5;
#ifdef HDR_TIMING
  const long long t_start = clock64(); long long t_a = 0, t_b = 0;
#endif
  const bool trivial = M.cnt[4] != 0;
  const uint32_t lit_size = trivial ? 1 : M.sstate[LIT][2] * 64, dist_size = trivial ? 1 : M.sstate[DIST][2] * 4;
  if (wp < 3) {
    const int cat = (int)wp;
    SerialHuff& S = H.S[cat];
    const uint32_t ntypes = M.sstate[cat][2], nb = M.sstate[cat][1];
    for (uint32_t i = lane; i < HDR_PART_WORDS; i += 32) H.part[cat][i] = 0;
    __syncwarp();
    if (ntypes <= 1) { if (lane == 0) { H.part[cat][0] = 0; H.part_bits[cat] = 1; } }
    else {
      // buildAndStoreBlockSplitCode (metablock.ts:161-202, 690-692)
      const uint32_t* bs = M.blk_start[cat]; const uint8_t* bt = M.blk_type[cat];
      for (uint32_t i = lane; i < ntypes + 2; i += 32) S.hist[i] = 0;
      if (lane < 26) H.len_hist[cat][lane] = 0;
      __syncwarp();
      for (uint32_t b = lane; b < nb; b += 32) {
        if (b) {
          const uint32_t last = bt[b - 1], second = b >= 2 ? bt[b - 2] : 1, t = bt[b];
          atomicAdd(&S.hist[t == last + 1 ? 1 : (t == second ? 0 : t + 2)], 1u);
        }
        atomicAdd(&H.len_hist[cat][block_len_code(bs[b + 1] - bs[b])], 1u);
      }
      __syncwarp();
      if (lane == 0) {
        WordSink<false> w; w.init(H.part[cat], 0);
        put_varlen8(w, ntypes - 1);
        huff_serial_store(w, S, (int)ntypes + 2);
        for (uint32_t i = 0; i < ntypes + 2; i++) { H.tdepth[cat][i] = S.depth[i]; H.tbits[cat][i] = S.bits[i]; }
        for (int i = 0; i < 26; i++) S.hist[i] = H.len_hist[cat][i];
        huff_serial_store(w, S, 26);
        {   // the first block switch carries no type code and is part of the header (metablock.ts:204-220)
          const uint32_t len = bs[1] - bs[0], lc = block_len_code(len);
          w.put(S.depth[lc], S.bits[lc]); w.put(c_blen_nbits[lc], len - c_blen_offset[lc]);
        }
        w.flush();
        H.part_bits[cat] = w.pos;
      }
      __syncwarp();
      for (uint32_t b = lane; b < nb; b += 32) {
        const uint32_t len = bs[b + 1] - bs[b], lc = block_len_code(len);
        uint64_t v = 0; uint32_t nbits = 0;
        if (b) {
          const uint32_t last = bt[b - 1], second = b >= 2 ? bt[b - 2] : 1, t = bt[b];
          const uint32_t tc = t == last + 1 ? 1 : (t == second ? 0 : t + 2);
          v = H.tbits[cat][tc]; nbits = H.tdepth[cat][tc];
        }
        v |= (uint64_t)S.bits[lc] << nbits; nbits += S.depth[lc];
        v |= (uint64_t)(len - c_blen_offset[lc]) << nbits; nbits += c_blen_nbits[lc];
        M.blk_tok[cat][b] = ((uint64_t)nbits << 56) | v;
      }
    }
  } else {
    const int k = (int)wp - 3;   // 0 literal, 1 distance
    const int cat = k == 0 ? LIT : DIST;
    cm_prepare(H.S[wp], M.cmap[cat], k == 0 ? lit_size : dist_size, M.cstate[cat][2], k == 0 ? H.lscr : H.dscr,
               H.tree[k], H.tree_bits[k], H.sym_bits[k], H.nout[k], H.max_prefix[k]);
  }
#ifdef HDR_TIMING
  t_a = clock64();
#endif
  __syncthreads();
#ifdef HDR_TIMING
  t_b = clock64();
#endif
  // storeCompressedMetaBlockHeader (metablock.ts:233-255) and the bits between the block-split codes and the context maps
  uint32_t head = 0, head_bits = 0;
  {
    const bool isLast = M.is_last != 0;
    head = isLast ? 1u : 0u; head_bits = 1;
    if (isLast) head_bits++;   // ISLASTEMPTY = 0
    const int lg = M.len == 1 ? 1 : log2_floor(M.len - 1) + 1;
    const uint32_t mnibbles = (uint32_t)((lg < 16 ? 16 : lg + 3) / 4);
    const uint64_t hv = (uint64_t)head | ((uint64_t)(mnibbles - 4) << head_bits) | ((uint64_t)(M.len - 1) << (head_bits + 2));
    head_bits += 2 + mnibbles * 4;
    if (!isLast) head_bits++;  // ISUNCOMPRESSED = 0
    if (tid == 0) { atomicOr(&M.misc[0], (uint32_t)hv); if (head_bits > 32) atomicOr(&M.misc[1], (uint32_t)(hv >> 32)); }
  }
  uint32_t pos = head_bits;
  for (int cat = 0; cat < 3; cat++) {
    if (wp == (uint32_t)cat) hdr_copy(M.misc, pos, H.part[cat], H.part_bits[cat], lane, 32);
    pos += H.part_bits[cat];
  }
  const uint32_t ntl = M.sstate[LIT][2];
  if (wp == 3) {   // metablock.ts:695-696, then the context mode of every literal block type
    if (lane == 0) { WordSink<true> w; w.init(M.misc, pos); w.put(2, P.npostfix); w.put(4, P.ndirect >> P.npostfix); w.flush(); }
    const uint32_t mode = M.cnt[3];
    for (uint32_t i = lane; i < ntl; i += 32) {
      const uint32_t bp = pos + 6 + 2 * i, o = bp & 31;
      if (mode) { atomicOr(&M.misc[bp >> 5], mode << o); if (o == 31 && (mode >> 1)) atomicOr(&M.misc[(bp >> 5) + 1], mode >> 1); }
    }
  }
  pos += 6 + 2 * ntl;
  const uint32_t lit_pos = pos, lit_sym = lit_pos + H.tree_bits[0], dist_pos = lit_sym + H.sym_bits[0], dist_sym = dist_pos + H.tree_bits[1];
  if (wp == 0) hdr_copy(M.misc, lit_pos, H.tree[0], H.tree_bits[0], lane, 32);
  if (wp == 1) hdr_copy(M.misc, dist_pos, H.tree[1], H.tree_bits[1], lane, 32);
  if (wp == 3) cm_emit(H.S[3], H.lscr, H.nout[0], H.max_prefix[0], M.cstate[LIT][2], M.misc, lit_sym);
  if (wp == 4) cm_emit(H.S[4], H.dscr, H.nout[1], H.max_prefix[1], M.cstate[DIST][2], M.misc, dist_sym);
#ifdef HDR_TIMING
  if (lane == 0 && (wp == 3 || wp == 0)) printf("mb %u warp %u: phase1 %lld barrier-wait %lld phase2 %lld cycles; types %u %u %u blocks %u %u %u ncl %u %u nout %u %u\n", blockIdx.x, wp,
      t_a - t_start, t_b - t_a, clock64() - t_b, M.sstate[0][2], M.sstate[1][2], M.sstate[2][2], M.sstate[0][1], M.sstate[1][1], M.sstate[2][1], M.cstate[LIT][2], M.cstate[DIST][2], H.nout[0], H.nout[1]);
#endif
  if (tid == 0) {
    M.cnt[5] = dist_sym + H.sym_bits[1];
    M.cnt[6] = M.cstate[LIT][2]; M.cnt[7] = M.cstate[DIST][2];
  }
}

}  // namespace bgpu
