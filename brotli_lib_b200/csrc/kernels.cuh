// sm_100a kernels of the Brotli encoder hot path (q5..q9 greedy path; K-numbers follow SURVEY.md §2).
//
//   K1-K3 match_h5.cuh  key sort (k_sort_*) + window probe (k_probe_sorted)
//                       hash-chains.ts:51-60,117-152 + match.ts:20-77,162-172
//   K5  k_parse     greedy parse              backward-references.ts:14-158
//       k_stitch    joins parse segments into whole-metablock command streams (no TS counterpart)
//   K7  k_hist      BuildHistograms           metablock.ts:305-325
//   K10 k_huff      Huffman + tree store      entropy-encode.ts:24-258, context-map.ts:215-520
//   K11 k_tilebits / k_offsets / k_write      metablock.ts:233-255, 300-388; bit-writer.ts:32-56
//
// Data layout in HBM (all indexed by absolute byte position in the staged input buffer):
//   in    u8 [N+pad]  input, every stream starts 256-byte aligned
//   sa/sb u32[N] x 4  key-sorted (position, 4 input bytes) entries, ping-pong of the two sort passes
//   best  u32[N]      chain-best match: len << 24 | distance   (len <= 128, distance < 2^22); aliases sa_pos
//   role  u8 [N]      1 literal | 2 copy start | 4 command start | 8 insert-only command
//   ins   u32[N>>sh]  insert length of the command starting at position p, slot p >> sh
//   cmds  Cmd[N>>sh]  command record, slot = (copy-start position) >> sh    (sh = 2: copies >= 4 bytes)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "codec_common.cuh"

namespace bgpu {

constexpr uint32_t NONE = 0xFFFFFFFFu;
constexpr uint32_t TILE = 1024;                 // positions per bit-writer tile
constexpr uint32_t MAX_BACKWARD = (1u << 22) - 16;  // backward-references.ts:31
constexpr int TREE_WORDS = 256;                 // serialised prefix code, u32 words per tree
constexpr int ALPHA_LIT = 256, ALPHA_CMD = 704, ALPHA_DIST_MAX = 128;
constexpr int CODE_STRIDE = ALPHA_LIT + ALPHA_CMD + ALPHA_DIST_MAX;  // per-metablock code table

struct ShardD {
  uint64_t in_off;    // absolute offset of the shard's first byte (multiple of 4)
  uint32_t len;
  uint32_t flags;     // 1 first shard of its stream, 2 last shard of its stream
  uint64_t out_off;   // byte offset of the shard's output region (multiple of 4)
  uint32_t out_cap;
  uint32_t first_mb, n_mb;
  uint32_t start_nbits, start_value;  // stream header bits (encodeWindowBits) for the first shard
  uint32_t first_stile;  // index of the shard's first sort tile (match_h5.cuh)
  uint32_t hist;         // leading bytes that are HISTORY (streaming, bgpu.cu stream_flush): matched against, not coded
};
struct MbD { uint32_t shard, rel_pos, len, first_seg, n_seg, first_tile, n_tile, is_last; };
struct SegD { uint32_t shard, rel_pos, len, mb, first_tile, ring_known, pad0, pad1; };

struct Params {
  int bucket_bits, depth, num_last;
  uint32_t npostfix, ndirect, dist_alphabet;
  int sh;  // command slot shift
  int hint;  // unused (the former BGPU_F_RING_HANDOVER guess; superseded by the REFINE pass of k_parse)
};

// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t ld32u(const uint8_t* base, uint64_t off) {
  // unaligned little-endian u32 from a 4-byte-aligned base (reads up to 7 bytes past off)
  const uint32_t* w = reinterpret_cast<const uint32_t*>(base) + (off >> 2);
  uint32_t lo = __ldg(w), hi = __ldg(w + 1);
  return __funnelshift_r(lo, hi, (uint32_t)(off & 3) * 8);
}
__device__ __forceinline__ uint32_t hash4(uint32_t v, int bits) { return (v * 0x1E35A7BDu) >> (32 - bits); }

// common prefix length of in[a..] and in[b..], limit-capped (match.ts:42-77)
__device__ __forceinline__ uint32_t match_len(const uint8_t* in, uint64_t a, uint64_t b, uint32_t limit) {
  uint32_t k = 0;
  while (k < limit) {
    uint32_t x = ld32u(in, a + k) ^ ld32u(in, b + k);
    if (x) { k += (__ffs((int)x) - 1) >> 3; break; }
    k += 4;
  }
  return k < limit ? k : limit;
}

}  // namespace bgpu
#include "match_h5.cuh"   // K1-K3: key sort + window probe
namespace bgpu {

// ---------------------------------------------------------------------------------------
// K5.  One warp per parse segment.  32 consecutive positions are evaluated speculatively with the
// current distance ring; the first lane holding an acceptable match is the reference's next
// command (the ring only changes at accepted matches, so the speculation is exact).
__device__ __forceinline__ uint32_t distance_to_code(uint32_t distance, const int* ring) {
  // backward-references.ts:146-158, plus a guard so that ring entries not yet learned in this
  // segment (0) are never referenced.  The reference tries 16 candidates in order: ring[0..3], then ring[0] -1 +1 -2 +2
  // -3 +3, then the same around ring[1]; the six offsets of a base are distinct, so at most one of them fits and
  // the order inside a base does not matter (r5m: the unrolled 16-way compare was 17 % of k_parse's instructions).
  const int d = (int)distance;
#pragma unroll
  for (int i = 0; i < 4; i++) if (ring[i] > 0 && d == ring[i]) return (uint32_t)i;
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const int delta = d - ring[k], a = delta < 0 ? -delta : delta;
    if (ring[k] > 0 && a >= 1 && a <= 3) return (uint32_t)(4 + 6 * k + 2 * (a - 1) + (delta > 0 ? 1 : 0));
  }
  return distance + 15;
}

// G lanes (a power of two <= 32) work on one segment; a warp carries 32 / G segments.  The parse of
// a segment is latency bound (a dependent chain of L2 reads per accepted match), so throughput
// comes from the number of segments in flight: G = 8 keeps one 32-byte sector of `best` and of the
// input per group read while quadrupling the segments per warp.
// Per-segment bookkeeping of the parse: seg_info bits 0..2 = min(ring updates made inside the segment, 7) — with >= 4 the
// segment's end view of the ring IS the decoder's ring there, whatever it started from; bit 8 = the segment's last copy
// was cut by the segment end and its source keeps matching for >= PARSE_TAIL_MIN bytes beyond it, i.e. the next segment
// starts inside a running copy (rare on text, where copies are short; the rule on copy-heavy data).  REFINE pass (second launch): only segments behind such a tail, whose predecessor's end view is
// complete, are parsed again, now starting from that ring — the last-distance probe then continues the running copy the
// way the reference's serial parse does (backward-references.ts:36-134), instead of re-finding it through the hash chain
// (DESIGN.md section 4 "Distance ring across parse segments": that re-finding cost ~20 bytes per 8 KiB boundary on copy-heavy
// data).  seg_assumed[seg] records the ring a segment was parsed from; k_ring_fixup reconciles it with the ring the decoder
// really has there.
constexpr uint32_t PARSE_TAIL_MIN = 8;
template <int G, bool REFINE>
__global__ void __launch_bounds__(128) k_parse(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                               const SegD* __restrict__ segs, uint32_t n_seg,
                                               const uint32_t* __restrict__ best, uint8_t* __restrict__ role,
                                               uint32_t* __restrict__ ins_at, Cmd* __restrict__ cmds,
                                               uint32_t* __restrict__ seg_ncmd, uint32_t* __restrict__ seg_trail, int4* __restrict__ seg_ring,
                                               uint32_t* __restrict__ seg_info, const int4* __restrict__ seg_ring0, const uint32_t* __restrict__ seg_info0,
                                               int4* __restrict__ seg_assumed, Params P) {
  const uint32_t seg = (blockIdx.x * 128 + threadIdx.x) / G;
  if (seg >= n_seg) return;
  const uint32_t wl = threadIdx.x & 31;
  const uint32_t lane = wl & (G - 1);
  const uint32_t gshift = wl & ~(uint32_t)(G - 1);
  const uint32_t gmask = (G == 32 ? 0xffffffffu : ((1u << G) - 1u)) << gshift;
  const SegD sg = segs[seg];
  const ShardD sh = shards[sg.shard];
  const uint8_t* sin = in + sh.in_off;
  const uint32_t* sbest = best + sh.in_off;
  uint8_t* srole = role + sh.in_off;
  const uint32_t seg_end = sg.rel_pos + sg.len;
  int ring[4];
  if (sg.ring_known) { ring[0] = 4; ring[1] = 11; ring[2] = 15; ring[3] = 16; }
  else { ring[0] = ring[1] = ring[2] = ring[3] = 0; }
  if (REFINE) {
    // (uniform over the G lanes of the segment: every value below is per segment)
    if (seg == 0 || sg.ring_known || segs[seg - 1].shard != sg.shard || segs[seg - 1].rel_pos + segs[seg - 1].len != sg.rel_pos) return;
    const uint32_t pi = seg_info0[seg - 1];
    if (!(pi & 0x100u) || (pi & 7u) < 4u) return;
    const int4 a = seg_ring0[seg - 1];
    ring[0] = a.x; ring[1] = a.y; ring[2] = a.z; ring[3] = a.w;
    if (lane == 0) seg_assumed[seg] = a;
    // the first pass marked this segment's positions: start from a clean slate (roles index everything else)
    for (uint32_t q = sg.rel_pos + lane * 4; q + 4 <= seg_end; q += G * 4) *reinterpret_cast<uint32_t*>(srole + q) = 0u;   // rel_pos is a multiple of 4
    if (lane == 0) for (uint32_t q = sg.rel_pos + ((seg_end - sg.rel_pos) & ~3u); q < seg_end; q++) srole[q] = 0;            // ragged shard end
    __syncwarp(gmask);
  }
  uint32_t pos = sg.rel_pos, ins_start = sg.rel_pos, ncmd = 0, npush = 0, last_dist = 0;
  while (pos < seg_end) {
    const uint32_t q = pos + lane;
    uint32_t blen = 0, bdist = 0; int bscore = 0;
    if (q + 4 <= seg_end) {
      const uint32_t maxLen = min(seg_end - q, 128u);
      const uint32_t maxBackward = min(q, MAX_BACKWARD);
      const uint32_t v = ld32u(sin, q);
      const uint32_t b = sbest[q];
      // hash-chains.ts:86-115: distance-cache candidates, all scored with the last-distance bonus
#pragma unroll
      for (int i = 0; i < 16; i++) {
        if (i >= P.num_last) break;
        // match.ts:128-148 prepareDistanceCache: ring[0] -+ 1,2,3 then ring[1] -+ 1,2,3
        const int base = i < 4 ? ring[i & 3] : (i < 10 ? ring[0] : ring[1]);
        const int d = i < 4 ? base : base + ((i & 1) ? 1 : -1) * (((i - 4) % 6) / 2 + 1);
        if (base <= 0 || d <= 0 || (uint32_t)d > maxBackward) continue;
        if (ld32u(sin, q - d) != v) continue;
        uint32_t len = 4 + match_len(sin, q + 4, q - d + 4, maxLen - 4);
        int score = 135 * (int)len + 975;
        if (score > bscore) { bscore = score; blen = len; bdist = (uint32_t)d; }
      }
      // chain best from K3 (truncated at the segment end)
      uint32_t clen = min(b >> 24, maxLen), cdist = b & 0xFFFFFFu;
      if (clen >= 4) {
        int score = 135 * (int)clen + 960 - 30 * log2_floor(cdist);
        if (score > bscore) { bscore = score; blen = clen; bdist = cdist; }
      }
    }
    const uint32_t ballot = (__ballot_sync(gmask, blen >= 4) & gmask) >> gshift;
    const uint32_t f = ballot ? (uint32_t)__ffs((int)ballot) - 1 : (uint32_t)G;
    if (lane < f && q < seg_end) srole[q] = 1 | (q == ins_start ? 4 : 0);
    if (f == (uint32_t)G) { pos += G; continue; }
    const uint32_t mlen = __shfl_sync(gmask, blen, gshift + f);
    const uint32_t mdist = __shfl_sync(gmask, bdist, gshift + f);
    const uint32_t qf = pos + f;
    const uint32_t code = distance_to_code(mdist, ring);
    if (lane == 0) {
      Cmd c = make_cmd(qf - ins_start, mlen, code, P.ndirect, P.npostfix);
      srole[qf] = 2 | (qf == ins_start ? 4 : 0);
      ins_at[(sh.in_off + ins_start) >> P.sh] = qf - ins_start;
      *reinterpret_cast<uint4*>(&cmds[(sh.in_off + qf) >> P.sh]) = *reinterpret_cast<uint4*>(&c);
    }
    if (code > 0) { ring[3] = ring[2]; ring[2] = ring[1]; ring[1] = ring[0]; ring[0] = (int)mdist; npush++; }
    pos = qf + mlen; ins_start = pos; ncmd++; last_dist = mdist;
  }
  if (lane == 0) {
    seg_ncmd[seg] = ncmd; seg_trail[seg] = ins_start; seg_ring[seg] = make_int4(ring[0], ring[1], ring[2], ring[3]);
    uint32_t tail = 0;
    if (ncmd && ins_start == seg_end && seg_end + PARSE_TAIL_MIN <= sh.len)
      tail = match_len(sin, seg_end, seg_end - last_dist, PARSE_TAIL_MIN) >= PARSE_TAIL_MIN ? 0x100u : 0u;
    seg_info[seg] = min(npush, 7u) | tail;
  }
}

// Joins the segments of a metablock: literals left over at the end of a segment become part of the
// next segment's first command; what remains at the metablock end is the insert-only command of
// encode.ts:253-267 / command.ts:183-208.
// One warp per metablock.  The pending literal run ("carry") after segment s is a function of the carry before it:
// a segment with commands replaces it by its own trailing literals (or none), a segment without commands keeps a
// pending run and otherwise starts one.  These functions compose ({keeps, own value}), so the carry entering every
// segment comes from a warp scan over 32 segments at a time, and the segments then apply their edits in parallel
// (each pending run is consumed by exactly one later segment).  Round 1 walked the 512 segments of a 4 MiB
// metablock with one thread: 1.3 ms per job whatever its size.
__global__ void __launch_bounds__(32) k_stitch(const ShardD* __restrict__ shards, const MbD* __restrict__ mbs, const SegD* __restrict__ segs,
                         uint32_t n_mb, uint8_t* __restrict__ role, uint32_t* __restrict__ ins_at, Cmd* __restrict__ cmds,
                         const uint32_t* __restrict__ seg_ncmd, const uint32_t* __restrict__ seg_trail, Params P) {
  const uint32_t m = blockIdx.x, lane = threadIdx.x;
  if (m >= n_mb) return;
  const MbD mb = mbs[m];
  const ShardD sh = shards[mb.shard];
  uint8_t* srole = role + sh.in_off;
  int64_t carry = -1;  // carry entering the current chunk of 32 segments (warp-uniform), shard-relative start of the pending run
  for (uint32_t s0 = 0; s0 < mb.n_seg; s0 += 32) {
    const uint32_t s = mb.first_seg + s0 + lane;
    const bool live = s0 + lane < mb.n_seg;
    SegD sg; sg.rel_pos = 0; sg.len = 0;
    uint32_t nc = 0, trail = 0;
    if (live) { sg = segs[s]; nc = seg_ncmd[s]; trail = seg_trail[s]; }
    const uint32_t seg_end = sg.rel_pos + sg.len;
    // f(x) = (keeps && x >= 0) ? x : own;  dead lanes are the identity (keeps, own = -1)
    bool keeps = !live || nc == 0;
    int64_t own = (live && trail < seg_end) ? (int64_t)trail : -1;
    // inclusive scan of the composition "earlier first"
    for (int o = 1; o < 32; o <<= 1) {
      const bool pk = __shfl_up_sync(0xffffffffu, keeps, o) != 0;
      const int64_t po = __shfl_up_sync(0xffffffffu, own, o);
      if (lane >= (uint32_t)o) { if (keeps) { own = po >= 0 ? po : own; keeps = pk; } }
    }
    const int64_t out = (keeps && carry >= 0) ? carry : own;            // carry after this lane's segment
    int64_t in_c = __shfl_up_sync(0xffffffffu, out, 1);                 // carry before it
    if (lane == 0) in_c = carry;
    if (live && in_c >= 0) {
      srole[sg.rel_pos] &= ~4;
      if (nc > 0) {
        const uint32_t ins_old = ins_at[(sh.in_off + sg.rel_pos) >> P.sh];
        const uint32_t ins_new = ins_old + (sg.rel_pos - (uint32_t)in_c);
        Cmd* c = &cmds[(sh.in_off + sg.rel_pos + ins_old) >> P.sh];
        c->insertLen = ins_new;
        c->cmdPrefix = (uint16_t)combine_codes(ins_code(ins_new), copy_code(c->copyLen & 0x1FFFFFF), (c->distPrefix & 0x3FF) == 0);
        ins_at[(sh.in_off + (uint32_t)in_c) >> P.sh] = ins_new;
      }
    }
    carry = __shfl_sync(0xffffffffu, out, 31);
  }
  if (lane == 0 && carry >= 0) {
    srole[carry] |= 8;
    ins_at[(sh.in_off + (uint32_t)carry) >> P.sh] = mb.rel_pos + mb.len - (uint32_t)carry;
  }
}


// Ring reconciliation (runs after k_stitch).  Segment s was parsed from the ring A0 = seg_assumed[s] (zeros = "unknown":
// nothing referenced until learned; or the first pass's end view of segment s - 1 when s was re-parsed by the REFINE pass).
// The ring the decoder really has at the start of s is R0 = the FINAL end view of segment s - 1 when that segment made
// >= 4 ring updates of its own (then its end view does not depend on where it started); otherwise R0 is taken as unknown.
// Walking the segment's commands, each distance is decoded under the A view and re-coded under the R view; both views
// are advanced by their own push rule (a command whose code is 0 = "last distance" pushes nothing), and the walk stops
// as soon as they agree — from there on the remaining commands mean the same thing to both.  With A0 = zeros and R0 known
// this turns explicit distances into short codes (the former BGPU_F_RING_HANDOVER); with A0 != R0 after a re-parse it keeps
// the stream valid.  One thread per segment.
__global__ void k_ring_fixup(const ShardD* __restrict__ shards, const SegD* __restrict__ segs, uint32_t n_seg,
                             const uint32_t* __restrict__ ins_at, Cmd* __restrict__ cmds,
                             const uint32_t* __restrict__ seg_ncmd, const int4* __restrict__ seg_ring, const uint32_t* __restrict__ seg_info,
                             const int4* __restrict__ seg_assumed, Params P) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_seg || s == 0) return;
  const SegD sg = segs[s];
  const uint32_t nc = seg_ncmd[s];
  if (sg.ring_known || nc == 0) return;
  const int4 a0 = seg_assumed[s];
  int A[4] = {a0.x, a0.y, a0.z, a0.w};
  int R[4] = {0, 0, 0, 0};
  const bool prev_ok = segs[s - 1].shard == sg.shard && segs[s - 1].rel_pos + segs[s - 1].len == sg.rel_pos && (seg_info[s - 1] & 7u) >= 4u;
  if (prev_ok) { const int4 e = seg_ring[s - 1]; R[0] = e.x; R[1] = e.y; R[2] = e.z; R[3] = e.w; }
  if (A[0] == R[0] && A[1] == R[1] && A[2] == R[2] && A[3] == R[3]) return;
  const ShardD sh = shards[sg.shard];
  const int8_t off[16] = {0, 0, 0, 0, -1, 1, -2, 2, -3, 3, -1, 1, -2, 2, -3, 3};
  uint32_t p = sg.rel_pos;
  for (uint32_t k = 0; k < nc; k++) {
    const uint32_t il = ins_at[(sh.in_off + p) >> P.sh];
    Cmd* c = &cmds[(sh.in_off + p + il) >> P.sh];
    const uint32_t dp = c->distPrefix, oc = dp & 0x3FF;
    const uint32_t dist = oc < 16 ? (uint32_t)((oc < 4 ? A[oc] : (oc < 10 ? A[0] : A[1])) + off[oc]) : distance_of_explicit(dp, c->distExtra, P.ndirect, P.npostfix);
    const uint32_t ncode = distance_to_code(dist, R);
    if (ncode != (oc < 16 ? oc : dist + 15)) {
      uint32_t extra; const uint32_t ndp = dist_prefix(ncode, P.ndirect, P.npostfix, &extra);
      c->distPrefix = (uint16_t)ndp; c->distExtra = extra;
      c->cmdPrefix = (uint16_t)combine_codes(ins_code(c->insertLen), copy_code(c->copyLen & 0x1FFFFFF), (ndp & 0x3FF) == 0);
    }
    if (ncode > 0) { R[3] = R[2]; R[2] = R[1]; R[1] = R[0]; R[0] = (int)dist; }
    if (oc > 0) { A[3] = A[2]; A[2] = A[1]; A[1] = A[0]; A[0] = (int)dist; }
    if (A[0] == R[0] && A[1] == R[1] && A[2] == R[2] && A[3] == R[3]) break;
    p += il + (c->copyLen & 0x1FFFFFF);
  }
}

// ---------------------------------------------------------------------------------------
// K7.  Literal / command / distance histograms of every metablock (metablock.ts:305-325) plus the
// per-tile command-independent bookkeeping.  One CTA per 8 tiles; shared-memory histograms are
// flushed with global atomics whenever the metablock changes.
struct MbHist { uint32_t lit[ALPHA_LIT]; uint32_t cmd[ALPHA_CMD]; uint32_t dist[ALPHA_DIST_MAX]; };

__global__ void __launch_bounds__(256) k_hist(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                              const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg, uint32_t n_tile,
                                              const uint8_t* __restrict__ role, const uint32_t* __restrict__ ins_at,
                                              const Cmd* __restrict__ cmds, MbHist* __restrict__ hist, Params P) {
  __shared__ uint32_t s_h[CODE_STRIDE];
  for (uint32_t i = threadIdx.x; i < CODE_STRIDE; i += 256) s_h[i] = 0;
  __syncthreads();
  uint32_t cur_mb = NONE;
  const uint32_t t0 = blockIdx.x * 8, t1 = min(t0 + 8, n_tile);
  for (uint32_t tile = t0; tile < t1; tile++) {
    const SegD sg = segs[tile_seg[tile]];
    if (sg.mb != cur_mb) {
      if (cur_mb != NONE) {
        __syncthreads();
        uint32_t* g = reinterpret_cast<uint32_t*>(&hist[cur_mb]);
        for (uint32_t i = threadIdx.x; i < CODE_STRIDE; i += 256) { uint32_t v = s_h[i]; if (v) atomicAdd(&g[i], v); s_h[i] = 0; }
        __syncthreads();
      }
      cur_mb = sg.mb;
    }
    const ShardD sh = shards[sg.shard];
    const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
    const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
    for (uint32_t rel = tile_rel + threadIdx.x; rel < tile_end; rel += 256) {
      const uint64_t p = sh.in_off + rel;
      const uint32_t r = role[p];
      if (r & 1) atomicAdd(&s_h[in[p]], 1u);
      if (r & 4) {
        uint32_t ins = ins_at[p >> P.sh];
        uint32_t prefix = (r & 8) ? insert_only_prefix(ins) : cmds[(p + ins) >> P.sh].cmdPrefix;
        atomicAdd(&s_h[ALPHA_LIT + prefix], 1u);
      }
      if (r & 2) {
        const Cmd c = cmds[p >> P.sh];
        if (c.cmdPrefix >= 128) atomicAdd(&s_h[ALPHA_LIT + ALPHA_CMD + (c.distPrefix & 0x3FF)], 1u);
      }
    }
  }
  __syncthreads();
  if (cur_mb != NONE) {
    uint32_t* g = reinterpret_cast<uint32_t*>(&hist[cur_mb]);
    for (uint32_t i = threadIdx.x; i < CODE_STRIDE; i += 256) { uint32_t v = s_h[i]; if (v) atomicAdd(&g[i], v); }
  }
}

// ---------------------------------------------------------------------------------------
// K10.  One CTA (128 threads) per prefix code.  Histogram -> depths (<= 15) -> canonical codes ->
// serialised tree bits.  The sort is a parallel rank-by-counting over the total order
// (count asc, symbol desc); the two-queue merge and the serialisation are run by thread 0.
// code table entry: bits << 8 | depth.
// CTA-cooperative (128 threads) body shared by k_huff and k_huff_mb (metablock_model.cuh).
__device__ __forceinline__ void huff_cta(const uint32_t* __restrict__ hist, const int alphabet, uint32_t* __restrict__ code_out,
                                         uint32_t* __restrict__ tree_bits_out, uint32_t* __restrict__ tree_nbits_out) {
  __shared__ uint32_t s_hist[ALPHA_CMD];
  __shared__ uint16_t s_sorted[ALPHA_CMD];
  __shared__ uint32_t s_cnt[2 * ALPHA_CMD + 2];
  __shared__ uint16_t s_lt[2 * ALPHA_CMD + 2], s_rt[2 * ALPHA_CMD + 2];
  __shared__ uint8_t s_nd[2 * ALPHA_CMD + 2];
  __shared__ uint8_t s_depth[ALPHA_CMD];
  __shared__ uint16_t s_bits[ALPHA_CMD];
  __shared__ uint8_t s_rle_sym[ALPHA_CMD + 8], s_rle_extra[ALPHA_CMD + 8];
  __shared__ uint32_t s_tree[TREE_WORDS];
  __shared__ int s_n, s_ok;
  __shared__ uint32_t s_limit;
  __shared__ uint16_t s_s4[4];
  const int tid = threadIdx.x;
  for (int i = tid; i < alphabet; i += 128) { s_hist[i] = hist[i]; s_depth[i] = 0; s_bits[i] = 0; }
  for (int i = tid; i < TREE_WORDS; i += 128) s_tree[i] = 0;
  __syncthreads();
  if (tid == 0) {
    int n = 0;
    for (int i = 0; i < alphabet; i++) if (s_hist[i]) { if (n < 4) s_s4[n] = (uint16_t)i; n++; }
    if (n == 0) s_s4[0] = 0;
    s_n = n; s_limit = 1; s_ok = 0;
  }
  __syncthreads();
  const int n = s_n;
  if (n >= 2) {
    for (;;) {
      const uint32_t limit = s_limit;
      for (int s = tid; s < alphabet; s += 128) {
        uint32_t hs = s_hist[s];
        if (!hs) continue;
        uint32_t cs = max(hs, limit);
        int rank = 0;
        for (int t = 0; t < alphabet; t++) {
          uint32_t ht = s_hist[t];
          if (!ht || t == s) continue;
          uint32_t ct = max(ht, limit);
          rank += (ct < cs || (ct == cs && t > s)) ? 1 : 0;
        }
        s_sorted[rank] = (uint16_t)s;
      }
      __syncthreads();
      if (tid == 0) {
        s_ok = huff_depths_from_sorted(s_hist, s_sorted, n, limit, 15, s_cnt, s_lt, s_rt, s_nd, s_depth) ? 1 : 0;
        if (!s_ok) s_limit = limit * 2;
      }
      __syncthreads();
      if (s_ok) break;
    }
  }
  if (tid == 0) {
    BitSink w{s_tree, 0};
    if (n <= 1) {
      // context-map.ts:236-243: depth stays 0 (the single symbol costs no bits)
      tree_store_single(w, alphabet, s_s4[0]);
    } else {
      huff_codes(s_depth, alphabet, s_bits);
      tree_store(w, s_depth, alphabet, n, s_s4, s_rle_sym, s_rle_extra);
    }
    *tree_nbits_out = w.pos;
  }
  __syncthreads();
  for (int i = tid; i < alphabet; i += 128) code_out[i] = ((uint32_t)s_bits[i] << 8) | s_depth[i];
  for (int i = tid; i < TREE_WORDS; i += 128) tree_bits_out[i] = s_tree[i];
}

__global__ void __launch_bounds__(128) k_huff(const uint32_t* __restrict__ hists, int hist_stride, int n_codes,
                                              const int* __restrict__ alphabets, const int* __restrict__ hist_offsets,
                                              uint32_t* __restrict__ codes, uint32_t* __restrict__ tree_bits,
                                              uint32_t* __restrict__ tree_nbits) {
  // code c of problem (blockIdx.x / n_codes) ; alphabets[c], hist_offsets[c] give its slice.
  const int prob = blockIdx.x / n_codes, which = blockIdx.x % n_codes;
  huff_cta(hists + (size_t)prob * hist_stride + hist_offsets[which], alphabets[which],
           codes + (size_t)prob * hist_stride + hist_offsets[which], tree_bits + (size_t)blockIdx.x * TREE_WORDS, tree_nbits + blockIdx.x);
}

// ---------------------------------------------------------------------------------------
// K11.  Tokens of one input position, in stream order (metablock.ts:355-383):
//   [command symbol + insert/copy extra bits]  if a command starts here
//   [literal symbol]                           if the byte is a literal
//   [distance symbol + extra bits]             if a copy starts here and the command codes a distance
struct Tok { uint32_t n0; uint64_t v0; uint32_t n1; uint64_t v1; };  // two fields of <= 63 bits

__device__ __forceinline__ Tok position_tokens(const uint8_t* __restrict__ in, uint64_t p, uint32_t r,
                                               const uint32_t* __restrict__ ins_at, const Cmd* __restrict__ cmds,
                                               const uint32_t* __restrict__ code /* shared-memory table of the metablock */, int sh) {
  Tok t; t.n0 = 0; t.v0 = 0; t.n1 = 0; t.v1 = 0;
  if (r & 4) {
    uint32_t ins = ins_at[p >> sh];
    uint32_t prefix, copyField;
    if (r & 8) { prefix = insert_only_prefix(ins); copyField = 2u << 25; }
    else { const Cmd c = cmds[(p + ins) >> sh]; prefix = c.cmdPrefix; copyField = c.copyLen; }
    uint32_t e = code[ALPHA_LIT + prefix];
    uint64_t xv; uint32_t xn = cmd_extra(ins, copyField, &xv);
    t.n0 = (e & 0xFF) + xn;
    t.v0 = (uint64_t)(e >> 8) | (xv << (e & 0xFF));
  }
  if (r & 1) {
    uint32_t e = code[in[p]];
    t.n1 = e & 0xFF; t.v1 = e >> 8;
  } else if (r & 2) {
    const Cmd c = cmds[p >> sh];
    if (c.cmdPrefix >= 128) {
      uint32_t e = code[ALPHA_LIT + ALPHA_CMD + (c.distPrefix & 0x3FF)];
      t.n1 = (e & 0xFF) + (c.distPrefix >> 10);
      t.v1 = (uint64_t)(e >> 8) | ((uint64_t)c.distExtra << (e & 0xFF));
    }
  }
  return t;
}

// per-tile bit totals
__global__ void __launch_bounds__(256) k_tilebits(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                  const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                  const uint8_t* __restrict__ role, const uint32_t* __restrict__ ins_at,
                                                  const Cmd* __restrict__ cmds, const uint32_t* __restrict__ codes,
                                                  uint32_t* __restrict__ tile_bits, Params P) {
  __shared__ uint32_t s_code[CODE_STRIDE];
  __shared__ uint32_t s_sum[8];
  const uint32_t tile = blockIdx.x;
  const SegD sg = segs[tile_seg[tile]];
  const ShardD sh = shards[sg.shard];
  for (uint32_t i = threadIdx.x; i < CODE_STRIDE; i += 256) s_code[i] = codes[(size_t)sg.mb * CODE_STRIDE + i];
  __syncthreads();
  const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
  const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
  uint32_t bits = 0;
  for (uint32_t rel = tile_rel + threadIdx.x; rel < tile_end; rel += 256) {
    const uint64_t p = sh.in_off + rel;
    const uint32_t r = role[p];
    if (r) { Tok t = position_tokens(in, p, r, ins_at, cmds, s_code, P.sh); bits += t.n0 + t.n1; }
  }
  for (int o = 16; o; o >>= 1) bits += __shfl_xor_sync(0xffffffffu, bits, o);
  if ((threadIdx.x & 31) == 0) s_sum[threadIdx.x >> 5] = bits;
  __syncthreads();
  if (threadIdx.x == 0) { uint32_t s = 0; for (int i = 0; i < 8; i++) s += s_sum[i]; tile_bits[tile] = s; }
}

__device__ __forceinline__ void gput(uint32_t* w, uint64_t& pos, uint32_t nbits, uint32_t value) {
  if (!nbits) return;
  uint64_t i = pos >> 5; uint32_t o = (uint32_t)pos & 31;
  atomicOr(&w[i], value << o);
  if (o + nbits > 32) atomicOr(&w[i + 1], value >> (32 - o));
  pos += nbits;
}

// One CTA per shard: metablock headers + prefix codes, bit offset of every tile, shard byte length.
// Header layout: metablock.ts:233-255 (ISLAST, [ISEMPTY], MNIBBLES, MLEN-1, [ISUNCOMPRESSED]) then
// metablock.ts:327-343 (three NBLTYPES = 1, NPOSTFIX/NDIRECT, NTREESL = 1 + context mode LSB6,
// NTREESD = 1), then the three prefix codes (literal, command, distance).
__global__ void __launch_bounds__(256) k_offsets(const ShardD* __restrict__ shards, const MbD* __restrict__ mbs,
                                                 const uint32_t* __restrict__ tile_bits, uint32_t* __restrict__ tile_off,
                                                 const uint32_t* __restrict__ tree_bits, const uint32_t* __restrict__ tree_nbits,
                                                 uint8_t* __restrict__ out, uint32_t* __restrict__ shard_bytes, Params P) {
  __shared__ uint32_t s_part[256];
  __shared__ uint64_t s_pos;
  const ShardD sh = shards[blockIdx.x];
  uint32_t* w = reinterpret_cast<uint32_t*>(out + sh.out_off);
  const int tid = threadIdx.x;
  if (tid == 0) {
    uint64_t pos = 0;
    if (sh.flags & 1) gput(w, pos, sh.start_nbits, sh.start_value);
    s_pos = pos;
  }
  __syncthreads();
  for (uint32_t m = sh.first_mb; m < sh.first_mb + sh.n_mb; m++) {
    const MbD mb = mbs[m];
    if (tid == 0) {
      uint64_t pos = s_pos;
      const bool isLast = mb.is_last != 0;
      gput(w, pos, 1, isLast ? 1 : 0);
      if (isLast) gput(w, pos, 1, 0);
      int lg = mb.len == 1 ? 1 : log2_floor(mb.len - 1) + 1;
      int mnibbles = (lg < 16 ? 16 : lg + 3) / 4;
      gput(w, pos, 2, mnibbles - 4);
      gput(w, pos, mnibbles * 4, mb.len - 1);
      if (!isLast) gput(w, pos, 1, 0);
      gput(w, pos, 3, 0);                                              // NBLTYPESL/I/D = 1
      gput(w, pos, 6, P.npostfix | ((P.ndirect >> P.npostfix) << 2));  // NPOSTFIX, NDIRECT
      gput(w, pos, 4, 0);                                              // NTREESL=1, mode LSB6, NTREESD=1
      for (int t = 0; t < 3; t++) {
        const uint32_t* tb = tree_bits + ((size_t)m * 3 + t) * TREE_WORDS;
        uint32_t nb = tree_nbits[m * 3 + t];
        for (uint32_t k = 0; k < nb; k += 32) gput(w, pos, min(32u, nb - k), nb - k >= 32 ? tb[k >> 5] : (tb[k >> 5] & ((1u << (nb - k)) - 1)));
      }
      s_pos = pos;
    }
    __syncthreads();
    // exclusive scan of the metablock's tile bit counts, 256 tiles per round
    for (uint32_t t0 = 0; t0 < mb.n_tile; t0 += 256) {
      uint32_t v = (t0 + tid < mb.n_tile) ? tile_bits[mb.first_tile + t0 + tid] : 0;
      s_part[tid] = v;
      __syncthreads();
      for (int o = 1; o < 256; o <<= 1) {
        uint32_t add = tid >= o ? s_part[tid - o] : 0;
        __syncthreads();
        s_part[tid] += add;
        __syncthreads();
      }
      uint32_t incl = s_part[tid];
      if (t0 + tid < mb.n_tile) tile_off[mb.first_tile + t0 + tid] = (uint32_t)s_pos + incl - v;
      __syncthreads();
      if (tid == 255) s_pos += incl;
      __syncthreads();
    }
  }
  if (tid == 0) {
    uint64_t pos = s_pos;
    if (!(sh.flags & 2)) gput(w, pos, 6, 6);  // empty metadata block: ISLAST=0, MNIBBLES=3 (11), reserved 0, MSKIPBYTES=0 (00)
    shard_bytes[blockIdx.x] = (uint32_t)((pos + 7) >> 3);
  }
}

// One CTA per tile: recompute the tokens, scan their lengths, assemble the tile's bit span in
// shared memory and store it (interior words with plain stores, the two boundary words with atomicOr).
constexpr uint32_t STAGE_WORDS = 4096;
__device__ __forceinline__ void sput(uint32_t* w, uint32_t pos, uint32_t nbits, uint64_t value) {
  if (!nbits) return;
  uint32_t i = pos >> 5, o = pos & 31;
  unsigned long long lo = (unsigned long long)value << o;   // bits [0,64) of value << o
  atomicOr(&w[i], (uint32_t)lo);
  if (o + nbits > 32) atomicOr(&w[i + 1], (uint32_t)(lo >> 32));
  if (o + nbits > 64) atomicOr(&w[i + 2], (uint32_t)(value >> (64 - o)));
}

__global__ void __launch_bounds__(256) k_write(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                               const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                               const uint8_t* __restrict__ role, const uint32_t* __restrict__ ins_at,
                                               const Cmd* __restrict__ cmds, const uint32_t* __restrict__ codes,
                                               const uint32_t* __restrict__ tile_bits, const uint32_t* __restrict__ tile_off,
                                               uint8_t* __restrict__ out, Params P) {
  __shared__ uint32_t s_code[CODE_STRIDE];
  __shared__ uint32_t s_stage[STAGE_WORDS];
  __shared__ uint32_t s_warp[8];
  const uint32_t tile = blockIdx.x;
  const uint32_t total = tile_bits[tile];
  if (total == 0) return;
  const SegD sg = segs[tile_seg[tile]];
  const ShardD sh = shards[sg.shard];
  const int tid = threadIdx.x;
  for (uint32_t i = tid; i < CODE_STRIDE; i += 256) s_code[i] = codes[(size_t)sg.mb * CODE_STRIDE + i];
  const uint32_t off = tile_off[tile];
  const uint32_t head = off & 31;
  const uint32_t nwords = (head + total + 31) >> 5;
  const bool staged = nwords + 2 <= STAGE_WORDS;
  if (staged) for (uint32_t i = tid; i < nwords + 2; i += 256) s_stage[i] = 0;
  __syncthreads();
  const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
  const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
  // thread t owns positions tile_rel + 4t .. +3
  Tok tk[4]; uint32_t mine = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    uint32_t rel = tile_rel + 4 * tid + k;
    tk[k].n0 = tk[k].n1 = 0; tk[k].v0 = tk[k].v1 = 0;
    if (rel < tile_end) {
      const uint64_t p = sh.in_off + rel;
      const uint32_t r = role[p];
      if (r) tk[k] = position_tokens(in, p, r, ins_at, cmds, s_code, P.sh);
    }
    mine += tk[k].n0 + tk[k].n1;
  }
  // block exclusive scan of `mine`
  uint32_t incl = mine;
  for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if ((tid & 31) >= o) incl += y; }
  if ((tid & 31) == 31) s_warp[tid >> 5] = incl;
  __syncthreads();
  uint32_t wbase = 0;
  for (int i = 0; i < (tid >> 5); i++) wbase += s_warp[i];
  uint32_t bitpos = head + wbase + incl - mine;
  uint32_t* gw = reinterpret_cast<uint32_t*>(out + sh.out_off) + (off >> 5);
#pragma unroll
  for (int k = 0; k < 4; k++) {
    if (staged) {
      sput(s_stage, bitpos, tk[k].n0, tk[k].v0); bitpos += tk[k].n0;
      sput(s_stage, bitpos, tk[k].n1, tk[k].v1); bitpos += tk[k].n1;
    } else {
      sput(gw, bitpos, tk[k].n0, tk[k].v0); bitpos += tk[k].n0;
      sput(gw, bitpos, tk[k].n1, tk[k].v1); bitpos += tk[k].n1;
    }
  }
  if (!staged) return;
  __syncthreads();
  for (uint32_t i = tid; i < nwords; i += 256) {
    uint32_t v = s_stage[i];
    if (i == 0 || i == nwords - 1) { if (v) atomicOr(&gw[i], v); }
    else gw[i] = v;
  }
}

}  // namespace bgpu
