// sm_100a match finder for the binary-tree hasher ("H10", q10/q11): FindAllMatches for every position.
//
// Reference semantics (paths relative to the reference tree, src/encode/):
//   hash-binary-tree.ts:19-55    17-bit hashBytes4 buckets, forest[2 * pos] / [2 * pos + 1] child links
//   hash-binary-tree.ts:57-153   storeAndFindMatches: walk <= 64 levels from the bucket root, compare up
//                                to 128 bytes in the tree, re-root the tree at the new position
//   hash-binary-tree.ts:156-227  findAllMatches: 63 short-distance probes (len >= 2, while best <= 2)
//                                followed by the tree walk; matches strictly increasing in length
// with two repairs: buckets start invalid as reset() would leave them (the TS
// constructor leaves them zero), and both dangling links are closed when the 64-level budget runs out.
//
// B200 formulation.  A position only ever meets positions of its own bucket, and a forest slot
// belongs to exactly one bucket's tree, so the trees of different buckets are independent serial
// problems: the positions of a shard are stably sorted by their 17-bit key (three LSD passes of the
// q5 key sort, match_h5.cuh) and every bucket's run is inserted oldest-first by ONE thread whose
// bucket root lives in a register (no `buckets` array).  131 072 buckets x shards independent
// chains; the candidates are exact because the insertion order inside a bucket is the reference's.
//
// Lengths are compared up to 128 bytes only (the tree never needs more: a candidate that reaches
// maxCompLen = 128 ends the walk).  A stored length of 128 means ">= 128": the consumer (k_h10_extend
// for the parity stage, the Zopfli pass for production) extends it to the true length on demand and
// re-applies the "strictly longer than the best so far" filter.
//
//   k_h10_short   per position: the <= 2 matches of the 63-position short scan
//   k_h10_runs    bucket run [start, end) inside the sorted array
//   k_h10_tree    one thread per bucket: tree walk + re-root for every position of the run
//   k_h10_tree_hot  one warp per long run (>= H10_HOT_MIN positions): 32 inserts pipelined through the tree
//   k_h10_extend  true lengths for the entries stored as 128 (stage API only)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bgpu {

constexpr int H10_BUCKET_BITS = 17;
constexpr uint32_t H10_BUCKETS = 1u << H10_BUCKET_BITS;
constexpr uint32_t H10_MAX_TREE_COMP = 128, H10_MAX_DEPTH = 64;
constexpr uint32_t H10_MAX_MATCHES = 68;   // 2 short + 64 tree levels (+ slack)
constexpr uint32_t H10_INVALID = 0xFFFFFFFFu;
constexpr uint32_t H10_PENDING = 0xFFFFFFFEu;   // forest slot owned by an insert in flight (k_h10_tree_hot)
constexpr uint32_t H10_HOT_MIN = 64;            // bucket runs of at least this many positions take the warp-pipelined kernel

// match entry: len << 24 | distance  (len <= 128 where 128 means ">= 128"; distance < 2^22)
__device__ __forceinline__ uint32_t h10_pack(uint32_t dist, uint32_t len) { return (len << 24) | dist; }

// common prefix of in[a + from ..] and in[b + from ..], result capped at `limit` (absolute length)
__device__ __forceinline__ uint32_t h10_match_from(const uint8_t* in, uint32_t a, uint32_t b, uint32_t from, uint32_t limit) {
  uint32_t k = from;
  while (k < limit) {
    const uint32_t x = ld32u(in, (uint64_t)a + k) ^ ld32u(in, (uint64_t)b + k);
    if (x) { k += (uint32_t)(__ffs((int)x) - 1) >> 3; break; }
    k += 4;
  }
  return k < limit ? k : limit;
}

// Same value as h10_match_from, compared 16 bytes per step with the loads of both sides in flight
// together (one memory round trip per step instead of one per 4 bytes); *right receives
// in[a + len] > in[b + len], taken from the bytes already loaded (meaningful when len < limit).
__device__ __forceinline__ uint32_t h10_match16(const uint8_t* in, uint32_t a, uint32_t b, uint32_t from, uint32_t limit, bool* right) {
  uint32_t k = from;
  while (k < limit) {
    const uint4 x = ld128u(in, (uint64_t)a + k), y = ld128u(in, (uint64_t)b + k);
    const uint32_t n = prefix16(x, y);
    if (n < 16) {
      const uint32_t wa = n < 4 ? x.x : (n < 8 ? x.y : (n < 12 ? x.z : x.w));
      const uint32_t wb = n < 4 ? y.x : (n < 8 ? y.y : (n < 12 ? y.z : y.w));
      const uint32_t sh = (n & 3) * 8;
      *right = ((wa >> sh) & 0xFFu) > ((wb >> sh) & 0xFFu);
      k += n;
      break;
    }
    k += 16;
  }
  return k < limit ? k : limit;
}

// Short scan (hash-binary-tree.ts:168-193).  out[p] = e0 | e1 << 14 with e = dist | len << 6
// (dist <= 63, len <= 128), e1 set only when e0 has length 2.
__global__ void __launch_bounds__(256) k_h10_short(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                   uint32_t* __restrict__ short_out) {
  const ShardD sh = shards[blockIdx.y];
  const uint32_t cur = blockIdx.x * 256 + threadIdx.x;
  if (cur >= sh.len) return;
  const uint8_t* sin = in + sh.in_off;
  uint32_t res = 0;
  if (cur + 4 <= sh.len) {
    const uint32_t maxLength = min(sh.len - cur, H10_MAX_TREE_COMP);
    const uint32_t v = ld32u(sin, cur) & 0xFFFFu;
    const uint32_t stop = cur > 64 ? cur - 64 : 0;
    uint32_t bestLen = 1, n = 0;
    // 4 candidate positions per 64-bit window: prev = i, i-1, i-2, i-3 share the bytes [i-3, i+1]
    for (uint32_t i = cur - 1; cur > 0 && i > stop && bestLen <= 2; i--) {
      if ((ld32u(sin, i) & 0xFFFFu) != v) continue;
      const uint32_t len = h10_match_from(sin, cur, i, 2, maxLength);
      if (len > bestLen) {
        bestLen = len;
        res |= ((cur - i) | (len << 6)) << (14 * n);
        n++;
      }
    }
  }
  short_out[sh.in_off + cur] = res;
}

// Bucket runs of the key-sorted entries.  grid (ceil(npos / 256), shards).  bstart/bend: [shard][bucket].
__global__ void __launch_bounds__(256) k_h10_runs(const ShardD* __restrict__ shards, const uint32_t* __restrict__ srt_v,
                                                  uint32_t* __restrict__ bstart, uint32_t* __restrict__ bend) {
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh);
  const uint32_t e = blockIdx.x * 256 + threadIdx.x;
  if (e >= npos) return;
  const uint32_t* sv = srt_v + sh.in_off;
  const uint32_t key = hash4(__ldg(sv + e), H10_BUCKET_BITS);
  const size_t b = (size_t)blockIdx.y * H10_BUCKETS + key;
  if (e == 0 || hash4(__ldg(sv + e - 1), H10_BUCKET_BITS) != key) bstart[b] = e;
  if (e + 1 == npos || hash4(__ldg(sv + e + 1), H10_BUCKET_BITS) != key) bend[b] = e + 1;
}

struct H10Out {
  uint32_t* m;         // match entries, region of shard s starts at s.in_off * cap_per_pos
  uint32_t* off;       // [pos] offset of the position's list inside the shard region
  uint8_t* cnt;        // [pos] number of entries
  uint32_t* cursor;    // [shard] bump allocator
  uint32_t* overflow;  // set when a shard region ran out
  uint32_t cap_per_pos;
};

// One thread per (shard, bucket).  forest: 2 u32 per position (left, right), shard-relative values.
__global__ void __launch_bounds__(128) k_h10_tree(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards, uint32_t n_shard,
                                                  const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ bstart,
                                                  const uint32_t* __restrict__ bend, const uint32_t* __restrict__ short_in,
                                                  uint32_t* __restrict__ forest, H10Out O) {
  const uint64_t t = (uint64_t)blockIdx.x * 128 + threadIdx.x;
  const uint32_t shard = (uint32_t)(t >> H10_BUCKET_BITS);
  if (shard >= n_shard) return;
  const uint32_t rs = bstart[t], re = bend[t];
  if (rs >= re || re - rs >= H10_HOT_MIN) return;   // long runs: k_h10_tree_hot
  const ShardD sh = shards[shard];
  const uint8_t* sin = in + sh.in_off;
  const uint32_t* sp = srt_pos + sh.in_off;
  uint32_t* fr = forest + 2 * sh.in_off;
  uint32_t* mreg = O.m + sh.in_off * O.cap_per_pos;
  const uint32_t region_cap = sh.len * O.cap_per_pos;
  uint32_t root = H10_INVALID;
  uint32_t lm[H10_MAX_MATCHES];
  for (uint32_t r = rs; r < re; r++) {
    const uint32_t cur = __ldg(sp + r);
    const uint32_t maxLength = sh.len - cur;
    const uint32_t maxComp = min(maxLength, H10_MAX_TREE_COMP);
    const bool reroot = maxLength >= H10_MAX_TREE_COMP;
    const uint32_t maxBackward = min(cur, MAX_BACKWARD);
    // short-scan results first
    const uint32_t sr = short_in[sh.in_off + cur];
    uint32_t n = 0, best = 1;
    if (sr & 0x3FFF) { lm[n++] = h10_pack(sr & 63, (sr >> 6) & 255); best = (sr >> 6) & 255; }
    if ((sr >> 14) & 0x3FFF) { lm[n++] = h10_pack((sr >> 14) & 63, (sr >> 20) & 255); best = (sr >> 20) & 255; }
    const bool short_capped = best >= H10_MAX_TREE_COMP && maxLength > H10_MAX_TREE_COMP;
    uint32_t prev = root;
    if (reroot) root = cur;
    uint32_t nodeL = 2 * cur, nodeR = 2 * cur + 1, bl = 0, br = 0;
    for (uint32_t depth = H10_MAX_DEPTH;; depth--) {
      if (depth == 0 || prev == H10_INVALID || cur - prev > maxBackward) {
        if (reroot) { fr[nodeL] = H10_INVALID; fr[nodeR] = H10_INVALID; }
        break;
      }
      // both children travel with the data compare: one dependent memory round trip per level
      const uint2 ch = *reinterpret_cast<const uint2*>(fr + 2 * prev);
      bool right = false;
      const uint32_t len = h10_match16(sin, cur, prev, min(bl, br), maxComp, &right);
      if (len > best || (short_capped && len >= H10_MAX_TREE_COMP)) {
        best = max(best, len);
        if (n < H10_MAX_MATCHES) lm[n++] = h10_pack(cur - prev, len);
      }
      if (len >= maxComp) {
        if (reroot) { fr[nodeL] = ch.x; fr[nodeR] = ch.y; }
        break;
      }
      if (right) {
        bl = len;
        if (reroot) fr[nodeL] = prev;
        nodeL = 2 * prev + 1;
        prev = ch.y;
      } else {
        br = len;
        if (reroot) fr[nodeR] = prev;
        nodeR = 2 * prev;
        prev = ch.x;
      }
    }
    uint32_t off = 0;
    if (n) {
      off = atomicAdd(&O.cursor[shard], n);
      if (off + n > region_cap) { atomicExch(O.overflow, 1u); n = 0; off = 0; }
      for (uint32_t k = 0; k < n; k++) mreg[off + k] = lm[k];
    }
    O.off[sh.in_off + cur] = off;
    O.cnt[sh.in_off + cur] = (uint8_t)n;
  }
}

// Long bucket runs.  hot_ctl[0] = number of listed buckets, hot_ctl[1] = work counter.
__global__ void __launch_bounds__(256) k_h10_hotlist(const uint32_t* __restrict__ bstart, const uint32_t* __restrict__ bend, uint64_t n_bucket,
                                                     uint32_t* __restrict__ hot_list, uint32_t* __restrict__ hot_ctl) {
  const uint64_t t = (uint64_t)blockIdx.x * 256 + threadIdx.x;
  if (t >= n_bucket) return;
  if (bend[t] - bstart[t] >= H10_HOT_MIN) hot_list[atomicAdd(&hot_ctl[0], 1u)] = (uint32_t)t;
}

// Warp-pipelined insertion for long bucket runs (a hot 4-gram bucket of a 4 MiB text shard holds
// tens of thousands of positions, and the thread-per-bucket chain of up to 64 dependent L2 round
// trips per insert is the critical path of the whole stage).  32 consecutive inserts of the run are
// in flight, one per lane, each one level per round.  The walk is top-down only: an insert owns at
// most two "holes" (the slots it will still write); every hole is marked H10_PENDING in the forest,
// and a later insert that needs a pending slot waits for it, so no insert ever overtakes an earlier
// one on a shared path and every read sees exactly the value the serial order produces.
// An insert whose first compare (against its predecessor, the bucket root at that time) already
// reaches maxCompLen ends at level 1 by copying the predecessor's two links, which makes the
// predecessor unreachable: inside a chain of such inserts (long runs, periodic data) only the last
// one has to own real links, copied from the chain's source node.
__global__ void __launch_bounds__(128) k_h10_tree_hot(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                      const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ bstart,
                                                      const uint32_t* __restrict__ bend, const uint32_t* __restrict__ short_in,
                                                      uint32_t* __restrict__ forest, H10Out O, const uint32_t* __restrict__ hot_list,
                                                      uint32_t* __restrict__ hot_ctl) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t FULL = 0xffffffffu;
  uint32_t lm[H10_MAX_MATCHES];
  for (;;) {
    uint32_t w = 0;
    if (lane == 0) w = atomicAdd(&hot_ctl[1], 1u);
    w = __shfl_sync(FULL, w, 0);
    if (w >= hot_ctl[0]) break;
    const uint32_t t = hot_list[w];
    const uint32_t shard = t >> H10_BUCKET_BITS;
    const ShardD sh = shards[shard];
    const uint8_t* sin = in + sh.in_off;
    const uint32_t* sp = srt_pos + sh.in_off;
    uint32_t* fr = forest + 2 * sh.in_off;
    const uint2* fr2 = reinterpret_cast<const uint2*>(fr);
    uint32_t* mreg = O.m + sh.in_off * O.cap_per_pos;
    const uint32_t region_cap = sh.len * O.cap_per_pos;
    const uint32_t rs = bstart[t], re = bend[t];
    uint32_t root = H10_INVALID;
    for (uint32_t r0 = rs; r0 < re; r0 += 32) {
      const bool active = r0 + lane < re;
      const uint32_t cur = active ? __ldg(sp + r0 + lane) : 0;
      const uint32_t maxLength = sh.len - cur;
      const uint32_t maxComp = min(maxLength, H10_MAX_TREE_COMP);
      const bool reroot = active && maxLength >= H10_MAX_TREE_COMP;
      const uint32_t maxBackward = min(cur, MAX_BACKWARD);
      // re-rooting positions are a prefix of the batch (only the last 127 positions of a shard do not re-root)
      const uint32_t nre = __popc(__ballot_sync(FULL, reroot));
      const uint32_t prevCur = __shfl_up_sync(FULL, cur, 1);
      const uint32_t lastRe = __shfl_sync(FULL, cur, nre ? nre - 1 : 0);
      const uint32_t startPrev = lane == 0 ? root : (lane - 1 < nre ? prevCur : (nre ? lastRe : root));
      const uint32_t next_root = nre ? lastRe : root;
      uint32_t n = 0, best = 1;
      bool short_capped = false;
      if (active) {
        const uint32_t sr = short_in[sh.in_off + cur];
        if (sr & 0x3FFF) { lm[n++] = h10_pack(sr & 63, (sr >> 6) & 255); best = (sr >> 6) & 255; }
        if ((sr >> 14) & 0x3FFF) { lm[n++] = h10_pack((sr >> 14) & 63, (sr >> 20) & 255); best = (sr >> 20) & 255; }
        short_capped = best >= H10_MAX_TREE_COMP && maxLength > H10_MAX_TREE_COMP;
      }
      // level-1 compare of every insert against its predecessor: independent of the tree state
      uint32_t len = 0; bool haveLen = false, l1t = false, right = false;
      if (active && startPrev != H10_INVALID && cur - startPrev <= maxBackward) {
        len = h10_match16(sin, cur, startPrev, 0, maxComp, &right);
        haveLen = true; l1t = len >= maxComp;
      }
      const bool next_l1t = __shfl_down_sync(FULL, l1t && reroot, 1) && lane < 31;
      const bool skippable = l1t && reroot && next_l1t;
      uint32_t src = startPrev;
      for (uint32_t k = 1; k < 32; k++) {
        const uint32_t s_up = __shfl_up_sync(FULL, src, 1);
        const bool k_up = __shfl_up_sync(FULL, skippable, 1);
        if (lane == k && l1t && k_up) src = s_up;
      }
      bool done = !active;
      uint32_t prev = startPrev, depth = H10_MAX_DEPTH, nodeL = 2 * cur, nodeR = 2 * cur + 1, bl = 0, br = 0;
      if (active && reroot && !skippable) { fr[2 * cur] = H10_PENDING; fr[2 * cur + 1] = H10_PENDING; }
      if (l1t) {
        if (len > best || short_capped) { best = max(best, len); lm[n++] = h10_pack(cur - startPrev, len); }
        if (!reroot || skippable) done = true;
      }
      __syncwarp();
      while (__ballot_sync(FULL, !done)) {
        if (!done) {
          if (l1t) {
            const uint2 ch = __ldcg(fr2 + src);
            if (ch.x != H10_PENDING && ch.y != H10_PENDING) { fr[2 * cur] = ch.x; fr[2 * cur + 1] = ch.y; done = true; }
          } else {
            for (;;) {
              if (depth == 0 || prev == H10_INVALID || cur - prev > maxBackward) {
                if (reroot) { fr[nodeL] = H10_INVALID; fr[nodeR] = H10_INVALID; }
                done = true; break;
              }
              const uint2 ch = __ldcg(fr2 + prev);
              if (!haveLen) { len = h10_match16(sin, cur, prev, min(bl, br), maxComp, &right); haveLen = true; }
              if (len >= maxComp) {
                if (ch.x == H10_PENDING || ch.y == H10_PENDING) break;
                if (len > best || short_capped) { best = max(best, len); if (n < H10_MAX_MATCHES) lm[n++] = h10_pack(cur - prev, len); }
                if (reroot) { fr[nodeL] = ch.x; fr[nodeR] = ch.y; }
                done = true; break;
              }
              const uint32_t child = right ? ch.y : ch.x;
              if (child == H10_PENDING) break;
              if (len > best) { best = len; if (n < H10_MAX_MATCHES) lm[n++] = h10_pack(cur - prev, len); }
              if (right) {
                bl = len;
                if (reroot) { fr[2 * prev + 1] = H10_PENDING; fr[nodeL] = prev; }
                nodeL = 2 * prev + 1;
              } else {
                br = len;
                if (reroot) { fr[2 * prev] = H10_PENDING; fr[nodeR] = prev; }
                nodeR = 2 * prev;
              }
              prev = child; depth--; haveLen = false;
              if (!(depth == 0 || prev == H10_INVALID || cur - prev > maxBackward)) break;   // next level in the next round
            }
          }
        }
        __syncwarp();
      }
      if (active) {
        uint32_t off = 0;
        if (n) {
          off = atomicAdd(&O.cursor[shard], n);
          if (off + n > region_cap) { atomicExch(O.overflow, 1u); n = 0; off = 0; }
          for (uint32_t k = 0; k < n; k++) mreg[off + k] = lm[k];
        }
        O.off[sh.in_off + cur] = off;
        O.cnt[sh.in_off + cur] = (uint8_t)n;
      }
      root = next_root;
      __syncwarp();
    }
  }
}

// Stage API only: true length of every entry (0 = dropped by the strictly-increasing filter).
// One thread per position.  full_len is parallel to the match entries.
__global__ void __launch_bounds__(256) k_h10_extend(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards, H10Out O,
                                                    uint32_t* __restrict__ full_len) {
  const ShardD sh = shards[blockIdx.y];
  const uint32_t cur = blockIdx.x * 256 + threadIdx.x;
  if (cur + 4 > sh.len) return;
  const uint8_t* sin = in + sh.in_off;
  const uint32_t n = O.cnt[sh.in_off + cur];
  const uint32_t* mr = O.m + sh.in_off * O.cap_per_pos + O.off[sh.in_off + cur];
  uint32_t* fl = full_len + sh.in_off * O.cap_per_pos + O.off[sh.in_off + cur];
  const uint32_t maxLength = sh.len - cur;
  uint32_t best = 1;
  for (uint32_t k = 0; k < n; k++) {
    const uint32_t e = mr[k];
    uint32_t len = e >> 24;
    if (len >= H10_MAX_TREE_COMP && maxLength > H10_MAX_TREE_COMP)
      len = h10_match_from(sin, cur, cur - (e & 0xFFFFFFu), H10_MAX_TREE_COMP, maxLength);
    if (len > best) { best = len; fl[k] = len; } else fl[k] = 0;
  }
}

}  // namespace bgpu
