// sm_100a match finder for the hash-chain hasher ("H5", q5..q9).
//
// Reference semantics (paths relative to the reference tree, src/encode/):
//   match.ts:160-172        hashBytes4: LE32(data[p..p+3]) * 0x1E35A7BD >>> (32 - bucketBits)
//   hash-chains.ts:51-60    store(): chains[p] = buckets[key]; buckets[key] = p
//   hash-chains.ts:117-152  chain walk of findLongestMatch: the <= 1 << (quality-1) most recent
//                           earlier positions with the same key, newest first, first strict
//                           maximum of backwardReferenceScore (match.ts:26-30)
//
// B200 formulation.  The reference's chain of position p is "the positions q < p with
// key[q] == key[p], newest first".  A per-position walk over a u32 link array costs one random
// 32-byte L2 sector per step (profiles/r1a: 26 sectors per input byte, L2-sector bound).  Here the
// positions of a shard are stably sorted by key instead (two LSD counting-sort passes, 7 low key
// bits then the 7 or 8 high bits), so the chain of the entry at sorted index r is the run
// r-1, r-2, ... r-depth inside its key run: contiguous, staged through shared memory and shared
// by neighbouring threads.  Entries carry the 4 input bytes, so a candidate is confirmed without
// touching the input; only real 4-byte matches go to the (L2-resident) input for extension.
//
//   k_sort_hist     per-tile digit histogram (pass 0 reads the input, pass 1 the pass-0 output)
//   k_sort_scan     exclusive scan of a shard's [digit][tile] table -> destination offsets
//   k_sort_scatter  stable rank inside the tile (match.any per warp round + per-warp counters),
//                   reorder through shared memory, coalesced run stores
//   k_probe_sorted  window probe over the sorted entries -> best[p] = len << 24 | distance
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bgpu {

constexpr uint32_t SORT_TILE = 4096;      // entries per sort tile (256 threads x 16)
constexpr uint32_t SORT_BINS = 256;       // bins reserved per tile in the histogram table
constexpr int SORT_LOW_BITS = 7;
constexpr uint32_t PROBE_TILE = 1024;     // sorted entries per probe CTA (256 threads x 4)
constexpr uint32_t PROBE_HALO = 256;      // >= max depth (quality 9)

__device__ __forceinline__ uint32_t sort_digit(uint32_t v, int bucket_bits, int pass) {
  const uint32_t key = hash4(v, bucket_bits);
  return pass == 0 ? (key & ((1u << SORT_LOW_BITS) - 1)) : (key >> SORT_LOW_BITS);
}
__device__ __forceinline__ uint32_t shard_npos(const ShardD& sh) { return sh.len >= 4 ? sh.len - 3 : 0; }
__device__ __forceinline__ uint32_t shard_nstile(const ShardD& sh) { return (shard_npos(sh) + SORT_TILE - 1) / SORT_TILE; }

// grid (max tiles per shard, shards).  hist layout per shard: [digit][tile], row length n_stile,
// at hist + first_stile * SORT_BINS.
template <int PASS>
__global__ void __launch_bounds__(256) k_sort_hist(const uint8_t* __restrict__ in, const uint32_t* __restrict__ src_v,
                                                   const ShardD* __restrict__ shards, uint32_t* __restrict__ hist, int bucket_bits) {
  __shared__ uint32_t s_h[SORT_BINS];
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh), n_stile = shard_nstile(sh), tile = blockIdx.x;
  if (tile >= n_stile) return;
  s_h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = tile * SORT_TILE;
  const uint8_t* sin = in + sh.in_off;
  const uint32_t* sv = src_v + sh.in_off;
#pragma unroll 4
  for (uint32_t k = 0; k < SORT_TILE / 256; k++) {
    const uint32_t e = base + k * 256 + threadIdx.x;
    if (e < npos) {
      const uint32_t v = PASS == 0 ? ld32u(sin, e) : __ldg(sv + e);
      atomicAdd(&s_h[sort_digit(v, bucket_bits, PASS)], 1u);
    }
  }
  __syncthreads();
  const uint32_t nb = PASS == 0 ? (1u << SORT_LOW_BITS) : (1u << (bucket_bits - SORT_LOW_BITS));
  if (threadIdx.x < nb) hist[(size_t)sh.first_stile * SORT_BINS + (size_t)threadIdx.x * n_stile + tile] = s_h[threadIdx.x];
}

// One CTA (1024 threads) per shard: in-place exclusive scan of nb * n_stile counters.
__global__ void __launch_bounds__(1024) k_sort_scan(const ShardD* __restrict__ shards, uint32_t* __restrict__ hist, int nb) {
  __shared__ uint32_t s_w[32];
  const ShardD sh = shards[blockIdx.x];
  const uint32_t n_stile = shard_nstile(sh);
  const uint32_t total = (uint32_t)nb * n_stile;
  if (total == 0) return;
  uint32_t* a = hist + (size_t)sh.first_stile * SORT_BINS;
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  // warp w owns the contiguous chunk [w * cw, (w + 1) * cw), cw a multiple of 32
  const uint32_t cw = ((total + 31) / 32 + 31) & ~31u;
  const uint32_t lo = w * cw, hi = min(lo + cw, total);
  uint32_t sum = 0;
  for (uint32_t i = lo + lane; i < hi; i += 32) sum += a[i];
  for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) s_w[w] = sum;
  __syncthreads();
  if (w == 0) {
    uint32_t x = s_w[lane], incl = x;
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
    s_w[lane] = incl - x;
  }
  __syncthreads();
  uint32_t carry = s_w[w];
  for (uint32_t i0 = lo; i0 < hi; i0 += 32) {
    const uint32_t i = i0 + lane;
    const uint32_t x = i < hi ? a[i] : 0;
    uint32_t incl = x;
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
    if (i < hi) a[i] = carry + incl - x;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
}

// Stable scatter of one tile.  PASS 0: entry e is position e (v read from the input); PASS 1:
// entries come from (src_pos, src_v).  256 threads = 8 warps; warp w ranks the contiguous slice
// [w * 512, (w + 1) * 512) of the tile in 16 rounds of 32 consecutive entries, which keeps equal
// digits in entry order.
template <int PASS>
__global__ void __launch_bounds__(256) k_sort_scatter(const uint8_t* __restrict__ in, const uint32_t* __restrict__ src_pos,
                                                      const uint32_t* __restrict__ src_v, const ShardD* __restrict__ shards,
                                                      const uint32_t* __restrict__ hist, uint32_t* __restrict__ dst_pos,
                                                      uint32_t* __restrict__ dst_v, int bucket_bits) {
  constexpr int ROUNDS = SORT_TILE / 256;  // 16
  __shared__ uint32_t s_cnt[8][SORT_BINS];
  __shared__ uint32_t s_base[SORT_BINS];   // tile-local exclusive offset of each digit
  __shared__ uint32_t s_goff[SORT_BINS];   // shard-level destination of the digit's first entry of this tile
  __shared__ uint32_t s_pos[SORT_TILE], s_v[SORT_TILE];
  __shared__ uint32_t s_ws[8];
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh), n_stile = shard_nstile(sh), tile = blockIdx.x;
  if (tile >= n_stile) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t nb = PASS == 0 ? (1u << SORT_LOW_BITS) : (1u << (bucket_bits - SORT_LOW_BITS));
  for (uint32_t i = tid; i < 8 * SORT_BINS; i += 256) (&s_cnt[0][0])[i] = 0;
  if (tid < nb) s_goff[tid] = hist[(size_t)sh.first_stile * SORT_BINS + (size_t)tid * n_stile + tile];
  __syncthreads();
  const uint32_t base = tile * SORT_TILE;
  const uint8_t* sin = in + sh.in_off;
  uint32_t v[ROUNDS]; uint16_t rk[ROUNDS];
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t e = base + w * (SORT_TILE / 8) + r * 32 + lane;
    const bool valid = e < npos;
    v[r] = valid ? (PASS == 0 ? ld32u(sin, e) : __ldg(src_v + sh.in_off + e)) : 0;
    const uint32_t d = valid ? sort_digit(v[r], bucket_bits, PASS) : (0x10000u | lane);
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const uint32_t before = __popc(peers & ((1u << lane) - 1));
    uint32_t b = 0;
    if (valid) b = s_cnt[w][d];
    __syncwarp();
    if (valid && (peers >> lane) == 1u) s_cnt[w][d] = b + __popc(peers);
    __syncwarp();
    rk[r] = (uint16_t)(b + before);
  }
  __syncthreads();
  // per-digit exclusive prefix over the 8 warps, then exclusive scan over digits
  uint32_t tot = 0;
  if (tid < nb) {
#pragma unroll
    for (int ww = 0; ww < 8; ww++) { uint32_t c = s_cnt[ww][tid]; s_cnt[ww][tid] = tot; tot += c; }
  }
  uint32_t incl = tot;
  for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
  if (lane == 31) s_ws[w] = incl;
  __syncthreads();
  uint32_t wb = 0;
  for (uint32_t i = 0; i < w; i++) wb += s_ws[i];
  s_base[tid] = wb + incl - tot;
  __syncthreads();
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t idx = w * (SORT_TILE / 8) + r * 32 + lane;
    const uint32_t e = base + idx;
    if (e < npos) {
      const uint32_t d = sort_digit(v[r], bucket_bits, PASS);
      const uint32_t lp = s_base[d] + s_cnt[w][d] + rk[r];
      s_v[lp] = v[r];
      s_pos[lp] = PASS == 0 ? e : __ldg(src_pos + sh.in_off + e);
    }
  }
  __syncthreads();
  const uint32_t nvalid = min(SORT_TILE, npos - base);
  uint32_t* dp = dst_pos + sh.in_off; uint32_t* dv = dst_v + sh.in_off;
#pragma unroll 4
  for (uint32_t j = tid; j < nvalid; j += 256) {
    const uint32_t vv = s_v[j];
    const uint32_t d = sort_digit(vv, bucket_bits, PASS);
    const uint32_t dest = s_goff[d] + (j - s_base[d]);
    dp[dest] = s_pos[j]; dv[dest] = vv;
  }
}

// Window probe over the key-sorted entries of a shard.  grid (ceil(4Mi / PROBE_TILE), shards).
__global__ void __launch_bounds__(256) k_probe_sorted(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                      const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ srt_v,
                                                      uint32_t* __restrict__ best, int depth, int bucket_bits) {
  __shared__ uint32_t s_p[PROBE_TILE + PROBE_HALO], s_v[PROBE_TILE + PROBE_HALO];
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh);
  const uint32_t g0 = blockIdx.x * PROBE_TILE;
  const uint32_t tid = threadIdx.x;
  uint32_t* sbest = best + sh.in_off;
  // the <= 3 tail positions of the shard have no 4-byte key: no candidate
  if (blockIdx.x == 0 && tid < sh.len - npos) sbest[npos + tid] = 0;
  if (g0 >= npos) return;
  const uint32_t* sp = srt_pos + sh.in_off; const uint32_t* sv = srt_v + sh.in_off;
  const uint32_t halo = (uint32_t)depth;  // <= PROBE_HALO
  for (uint32_t i = tid; i < PROBE_TILE + halo; i += 256) {
    const int64_t g = (int64_t)g0 - halo + i;
    const bool ok = g >= 0 && g < (int64_t)npos;
    s_p[i] = ok ? __ldg(sp + g) : 0;
    s_v[i] = ok ? __ldg(sv + g) : 0;
  }
  __syncthreads();
  const uint8_t* sin = in + sh.in_off;
#pragma unroll 1
  for (uint32_t k = 0; k < PROBE_TILE / 256; k++) {
    const uint32_t i = k * 256 + tid;
    const uint32_t g = g0 + i;
    if (g >= npos) break;
    const uint32_t v = s_v[halo + i], p = s_p[halo + i];
    const uint32_t key = hash4(v, bucket_bits);
    const uint32_t maxLen = min(sh.len - p, 128u);
    uint32_t bestLen = 0, bestDist = 0; int bestScore = 0;
    const uint32_t jmax = min((uint32_t)depth, g);
    for (uint32_t j = 1; j <= jmax; j++) {
      const uint32_t u = s_v[halo + i - j];
      if (u != v) {
        if (hash4(u, bucket_bits) != key) break;   // left the key run: end of the chain
        continue;                                  // hash collision: uses up one chain step
      }
      const uint32_t c = s_p[halo + i - j];
      const uint32_t backward = p - c;
      if (backward > MAX_BACKWARD) break;          // hash-chains.ts:124
      // hash-chains.ts:132-135 quick check on the byte at bestLen (prunes only losers)
      if (bestLen >= 4 && sin[c + bestLen] != sin[p + bestLen]) continue;
      const uint32_t len = 4 + match_len(sin, p + 4, c + 4, maxLen - 4);
      const int score = 135 * (int)len + 960 - 30 * log2_floor(backward);
      if (score > bestScore) { bestScore = score; bestLen = len; bestDist = backward; }
      if (bestLen == maxLen) break;
    }
    sbest[p] = (bestLen << 24) | bestDist;
  }
}

}  // namespace bgpu
