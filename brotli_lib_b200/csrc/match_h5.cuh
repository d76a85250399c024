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

// digit of one LSD pass: bits [shift, shift + nbits) of the bucket key
__device__ __forceinline__ uint32_t sort_digit(uint32_t v, int bucket_bits, int shift, int nbits) {
  return (hash4(v, bucket_bits) >> shift) & ((1u << nbits) - 1);
}
// Sort passes.  PASS 0 / 1: entries carry the 4 input bytes v, the key is hash4(v) (hash-chain and H10 hashers).
// PASS 2 / 3: SimpleHasher (hash-simple.ts:48-50, match.ts:174-189): the key hashes 5 bytes, so pass 2 computes
// it from the input once and the entries carry the KEY; later passes (3) take digits of it directly.
__device__ __forceinline__ uint32_t hash5(uint32_t v, uint32_t b4, int bits) { return ((v ^ (b4 << 24)) * 0x1E35A7BDu) >> (32 - bits); }
template <int PASS>
__device__ __forceinline__ uint32_t sort_load(const uint8_t* sin, const uint32_t* sv, uint32_t e, uint32_t shard_len, int bucket_bits) {
  if (PASS == 0) return ld32u(sin, e);
  if (PASS == 2) return hash5(ld32u(sin, e), e + 4 < shard_len ? (uint32_t)__ldg(sin + e + 4) : 0u, bucket_bits);   // data[pos + 4] | 0 past the end
  return __ldg(sv + e);
}
template <int PASS>
__device__ __forceinline__ uint32_t sort_digit_of(uint32_t v, int bucket_bits, int shift, int nbits) {
  return PASS >= 2 ? (v >> shift) & ((1u << nbits) - 1) : sort_digit(v, bucket_bits, shift, nbits);
}
__device__ __forceinline__ uint32_t shard_npos(const ShardD& sh) { return sh.len >= 4 ? sh.len - 3 : 0; }
__device__ __forceinline__ uint32_t shard_nstile(const ShardD& sh) { return (shard_npos(sh) + SORT_TILE - 1) / SORT_TILE; }

// grid (max tiles per shard, shards).  hist layout per shard: [digit][tile], row length n_stile,
// at hist + first_stile * SORT_BINS.
template <int PASS>
__global__ void __launch_bounds__(256) k_sort_hist(const uint8_t* __restrict__ in, const uint32_t* __restrict__ src_v,
                                                   const ShardD* __restrict__ shards, uint32_t* __restrict__ hist, int bucket_bits, int shift, int nbits) {
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
      const uint32_t v = sort_load<PASS>(sin, sv, e, sh.len, bucket_bits);
      atomicAdd(&s_h[sort_digit_of<PASS>(v, bucket_bits, shift, nbits)], 1u);
    }
  }
  __syncthreads();
  const uint32_t nb = 1u << nbits;
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
// entries come from (src_pos, src_v).  WARPS warps; warp w ranks the contiguous slice
// [w * SORT_TILE / WARPS, ...) of the tile in rounds of 32 consecutive entries, which keeps equal
// digits in entry order.  The rounds of a warp are a dependent chain through its counter row (read counter,
// match.any, leader adds the group size), so the kernel is latency-bound on that chain (r5d: 37 % of the stall
// samples on the leader's store, short scoreboard 12 per issue, warps active 49 %): more, shorter chains per tile
// (WARPS = 16: 8 rounds each) finish a tile sooner than 8 chains of 16 rounds.
constexpr uint32_t SORT_SCATTER_SMEM = SORT_TILE * 8;
template <int PASS, int WARPS = 8>
__global__ void __launch_bounds__(32 * WARPS, 2048 / (32 * WARPS) > 5 ? 5 : 2048 / (32 * WARPS)) k_sort_scatter(const uint8_t* __restrict__ in, const uint32_t* __restrict__ src_pos,
                                                      const uint32_t* __restrict__ src_v, const ShardD* __restrict__ shards,
                                                      const uint32_t* __restrict__ hist, uint32_t* __restrict__ dst_pos,
                                                      uint32_t* __restrict__ dst_v, int bucket_bits, int shift, int nbits) {
  constexpr int THREADS = 32 * WARPS;
  constexpr int ROUNDS = SORT_TILE / THREADS;
  constexpr uint32_t SLICE = SORT_TILE / WARPS;
  __shared__ uint16_t s_cnt[WARPS][SORT_BINS];
  __shared__ uint32_t s_base[SORT_BINS];   // tile-local exclusive offset of each digit
  __shared__ uint32_t s_goff[SORT_BINS];   // shard-level destination of the digit's first entry of this tile, minus s_base
  extern __shared__ __align__(16) uint32_t s_dyn[];   // SORT_SCATTER_SMEM bytes: the reorder buffers
  uint32_t* s_pos = s_dyn; uint32_t* s_v = s_dyn + SORT_TILE;
  __shared__ uint32_t s_ws[8];
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh), n_stile = shard_nstile(sh), tile = blockIdx.x;
  if (tile >= n_stile) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t nb = 1u << nbits;
  for (uint32_t i = tid; i < WARPS * SORT_BINS / 2; i += THREADS) reinterpret_cast<uint32_t*>(&s_cnt[0][0])[i] = 0;
  uint32_t goff = 0;
  if (tid < SORT_BINS) goff = tid < nb ? hist[(size_t)sh.first_stile * SORT_BINS + (size_t)tid * n_stile + tile] : 0;
  __syncthreads();
  const uint32_t base = tile * SORT_TILE;
  const uint8_t* sin = in + sh.in_off;
  uint32_t v[ROUNDS]; uint16_t rk[ROUNDS];
  // all loads of the thread first: the ranking rounds below are separated by __syncwarp, which keeps the
  // compiler from hoisting a round's load above the previous round
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t e = base + w * SLICE + r * 32 + lane;
    v[r] = e < npos ? sort_load<PASS>(sin, src_v + sh.in_off, e, sh.len, bucket_bits) : 0;
  }
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t e = base + w * SLICE + r * 32 + lane;
    const bool valid = e < npos;
    const uint32_t d = valid ? sort_digit_of<PASS>(v[r], bucket_bits, shift, nbits) : (0x10000u | lane);
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const uint32_t before = __popc(peers & ((1u << lane) - 1));
    uint32_t b = 0;
    if (valid) b = s_cnt[w][d];
    __syncwarp();
    if (valid && (peers >> lane) == 1u) s_cnt[w][d] = (uint16_t)(b + __popc(peers));
    __syncwarp();
    rk[r] = (uint16_t)(b + before);
  }
  __syncthreads();
  // per-digit exclusive prefix over the warps, then exclusive scan over digits (threads 0 .. 255 hold one digit each)
  uint32_t tot = 0;
  if (tid < nb) {
#pragma unroll
    for (int ww = 0; ww < WARPS; ww++) { uint32_t c = s_cnt[ww][tid]; s_cnt[ww][tid] = (uint16_t)tot; tot += c; }
  }
  if (tid < SORT_BINS) {
    uint32_t incl = tot;
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
    if (lane == 31) s_ws[w] = incl;
    tot = incl - tot;   // exclusive inside the warp
  }
  __syncthreads();
  if (tid < SORT_BINS) {
    uint32_t wb = 0;
    for (uint32_t i = 0; i < w; i++) wb += s_ws[i];
    s_base[tid] = wb + tot;
    s_goff[tid] = goff - (wb + tot);
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t idx = w * SLICE + r * 32 + lane;
    const uint32_t e = base + idx;
    if (e < npos) {
      const uint32_t d = sort_digit_of<PASS>(v[r], bucket_bits, shift, nbits);
      const uint32_t lp = s_base[d] + s_cnt[w][d] + rk[r];
      s_v[lp] = v[r];
      s_pos[lp] = (PASS == 0 || PASS == 2) ? e : __ldg(src_pos + sh.in_off + e);
    }
  }
  __syncthreads();
  const uint32_t nvalid = min(SORT_TILE, npos - base);
  uint32_t* dp = dst_pos + sh.in_off; uint32_t* dv = dst_v + sh.in_off;
#pragma unroll 4
  for (uint32_t j = tid; j < nvalid; j += THREADS) {
    const uint32_t vv = s_v[j];
    const uint32_t d = sort_digit_of<PASS>(vv, bucket_bits, shift, nbits);
    const uint32_t dest = s_goff[d] + j;
    dp[dest] = s_pos[j]; dv[dest] = vv;
  }
}

// ---- Packed form of the two passes for the hash-chain path (q5-9): ONE u32 per entry between the passes.
// A shard-relative position needs 22 bits (shards <= 4 MiB) and pass 1 only needs the key's high digit (<= 8 bits), so
// pass 0 writes  position | high digit << 22  and pass 1 writes the bare position; the probe takes the 4 key bytes from
// its own 16-byte gather.  Per entry and pass: 4 bytes written and read back instead of 8 (round 1: 49.8 GB of DRAM per
// GiB for this stage, profiles/traffic.json), and one reorder array in shared memory instead of two.
//   k_sort_hist_packed          digit histogram of pass 1 from the packed entries
//   k_sort_scatter_packed<0>    input -> packed entries, ordered by the low 7 key bits
//   k_sort_scatter_packed<1>    packed entries -> positions, ordered by the whole key
constexpr int SORT_POS_BITS = 22;
__global__ void __launch_bounds__(256) k_sort_hist_packed(const uint32_t* __restrict__ src, const ShardD* __restrict__ shards, uint32_t* __restrict__ hist, int nbits) {
  __shared__ uint32_t s_h[SORT_BINS];
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh), n_stile = shard_nstile(sh), tile = blockIdx.x;
  if (tile >= n_stile) return;
  s_h[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t base = tile * SORT_TILE;
  const uint32_t* sv = src + sh.in_off;
#pragma unroll 4
  for (uint32_t k = 0; k < SORT_TILE / 256; k++) {
    const uint32_t e = base + k * 256 + threadIdx.x;
    if (e < npos) atomicAdd(&s_h[__ldg(sv + e) >> SORT_POS_BITS], 1u);
  }
  __syncthreads();
  const uint32_t nb = 1u << nbits;
  if (threadIdx.x < nb) hist[(size_t)sh.first_stile * SORT_BINS + (size_t)threadIdx.x * n_stile + tile] = s_h[threadIdx.x];
}

template <int PASS>
__global__ void __launch_bounds__(256, 5) k_sort_scatter_packed(const uint8_t* __restrict__ in, const uint32_t* __restrict__ src,
                                                             const ShardD* __restrict__ shards, const uint32_t* __restrict__ hist,
                                                             uint32_t* __restrict__ dst, int bucket_bits, int nbits) {
  constexpr int WARPS = 8, ROUNDS = SORT_TILE / 256;
  constexpr uint32_t SLICE = SORT_TILE / WARPS;
  __shared__ uint16_t s_cnt[WARPS][SORT_BINS];
  __shared__ uint32_t s_base[SORT_BINS];
  __shared__ uint32_t s_goff[SORT_BINS];
  __shared__ uint32_t s_e[SORT_TILE];      // reorder buffer: the entries as they will be stored ...
  __shared__ uint8_t s_d[SORT_TILE];       // ... and their digit of this pass
  __shared__ uint32_t s_ws[8];
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh), n_stile = shard_nstile(sh), tile = blockIdx.x;
  if (tile >= n_stile) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t nb = 1u << nbits;
  for (uint32_t i = tid; i < WARPS * SORT_BINS / 2; i += 256) reinterpret_cast<uint32_t*>(&s_cnt[0][0])[i] = 0;
  const uint32_t goff = tid < nb ? hist[(size_t)sh.first_stile * SORT_BINS + (size_t)tid * n_stile + tile] : 0;
  __syncthreads();
  const uint32_t base = tile * SORT_TILE;
  const uint8_t* sin = in + sh.in_off;
  // v[r]: the entry as it leaves this pass; dg[r]: its digit in this pass
  uint32_t v[ROUNDS]; uint16_t rk[ROUNDS]; uint8_t dg[ROUNDS];
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t e = base + w * SLICE + r * 32 + lane;
    if (PASS == 0) {
      const uint32_t key = e < npos ? hash4(ld32u(sin, e), bucket_bits) : 0;
      dg[r] = (uint8_t)(key & ((1u << SORT_LOW_BITS) - 1));
      v[r] = e | ((key >> SORT_LOW_BITS) << SORT_POS_BITS);
    } else {
      const uint32_t x = e < npos ? __ldg(src + sh.in_off + e) : 0;
      dg[r] = (uint8_t)(x >> SORT_POS_BITS);
      v[r] = x & ((1u << SORT_POS_BITS) - 1);
    }
  }
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t e = base + w * SLICE + r * 32 + lane;
    const bool valid = e < npos;
    const uint32_t d = valid ? (uint32_t)dg[r] : (0x10000u | lane);
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    const uint32_t before = __popc(peers & ((1u << lane) - 1));
    uint32_t b = 0;
    if (valid) b = s_cnt[w][d];
    __syncwarp();
    if (valid && (peers >> lane) == 1u) s_cnt[w][d] = (uint16_t)(b + __popc(peers));
    __syncwarp();
    rk[r] = (uint16_t)(b + before);
  }
  __syncthreads();
  uint32_t tot = 0;
  if (tid < nb) {
#pragma unroll
    for (int ww = 0; ww < WARPS; ww++) { uint32_t c = s_cnt[ww][tid]; s_cnt[ww][tid] = (uint16_t)tot; tot += c; }
  }
  uint32_t incl = tot;
  for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
  if (lane == 31) s_ws[w] = incl;
  tot = incl - tot;
  __syncthreads();
  {
    uint32_t wb = 0;
    for (uint32_t i = 0; i < w; i++) wb += s_ws[i];
    s_base[tid] = wb + tot;
    s_goff[tid] = goff - (wb + tot);
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < ROUNDS; r++) {
    const uint32_t e = base + w * SLICE + r * 32 + lane;
    if (e < npos) {
      const uint32_t d = dg[r];
      const uint32_t lp = s_base[d] + s_cnt[w][d] + rk[r];
      s_e[lp] = v[r]; s_d[lp] = (uint8_t)d;
    }
  }
  __syncthreads();
  const uint32_t nvalid = min(SORT_TILE, npos - base);
  uint32_t* dp = dst + sh.in_off;
#pragma unroll 4
  for (uint32_t j = tid; j < nvalid; j += 256) dp[s_goff[s_d[j]] + j] = s_e[j];
}

// Window probe over the key-sorted entries of a shard.  grid (ceil(max npos / PROBE_TILE), shards).
// Every entry of the tile (and of the halo of `depth` entries before it) first fetches the 16 input
// bytes that follow its 4 key bytes into shared memory — the only random input reads of the common
// case.  Candidates are then confirmed (4 equal bytes) and extended to 20 bytes entirely in shared
// memory; only matches of >= 20 bytes go back to the (L2-resident) input.
__device__ __forceinline__ uint4 ld128u(const uint8_t* base, uint64_t off) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(base) + (off >> 2);
  const uint32_t sh = (uint32_t)(off & 3) * 8;
  const uint32_t a = __ldg(w), b = __ldg(w + 1), c = __ldg(w + 2), d = __ldg(w + 3), e = __ldg(w + 4);
  return make_uint4(__funnelshift_r(a, b, sh), __funnelshift_r(b, c, sh), __funnelshift_r(c, d, sh), __funnelshift_r(d, e, sh));
}
// 16 unaligned bytes as two aligned 16-byte loads (2 LSU instructions instead of the 5 of ld128u; the words are
// picked with selects).  `p` may have any alignment; reads stay inside the two 16-byte lines that hold p .. p + 15.
__device__ __forceinline__ uint4 ld128v(const uint8_t* p) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint4* q = reinterpret_cast<const uint4*>(a & ~(uintptr_t)15);
  const uint4 lo = __ldg(q), hi = __ldg(q + 1);
  const uint32_t k = (uint32_t)(a >> 2) & 3u, sh = (uint32_t)(a & 3) * 8;
  const uint32_t w0 = k == 0 ? lo.x : k == 1 ? lo.y : k == 2 ? lo.z : lo.w;
  const uint32_t w1 = k == 0 ? lo.y : k == 1 ? lo.z : k == 2 ? lo.w : hi.x;
  const uint32_t w2 = k == 0 ? lo.z : k == 1 ? lo.w : k == 2 ? hi.x : hi.y;
  const uint32_t w3 = k == 0 ? lo.w : k == 1 ? hi.x : k == 2 ? hi.y : hi.z;
  const uint32_t w4 = k == 0 ? hi.x : k == 1 ? hi.y : k == 2 ? hi.z : hi.w;
  return make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
}

// number of equal leading bytes of two 16-byte blocks (16 if identical)
__device__ __forceinline__ uint32_t prefix16(const uint4& x, const uint4& y) {
  // branch-free: the first differing word by selects, then its first differing byte
  const uint32_t d0 = x.x ^ y.x, d1 = x.y ^ y.y, d2 = x.z ^ y.z, d3 = x.w ^ y.w;
  const bool lo_any = (d0 | d1) != 0;
  const uint32_t d = lo_any ? (d0 ? d0 : d1) : (d2 ? d2 : d3);
  const uint32_t base = lo_any ? (d0 ? 0u : 4u) : (d2 ? 8u : 12u);
  return d ? base + ((uint32_t)(__ffs((int)d) - 1) >> 3) : 16u;
}

// Shared-memory plan of k_probe_sorted (dynamic, PROBE_SMEM bytes).
//   s_x        (TILE + HALO) uint4 the 16 input bytes after the key bytes
//   s_p, s_v   (TILE + HALO) u32   sorted positions / their 4 key bytes
constexpr uint32_t PROBE_N = PROBE_TILE + PROBE_HALO;
constexpr uint32_t PROBE_SMEM = PROBE_N * 16 + PROBE_N * 8;

// common prefix of in[a..] and in[b..] in 16-byte steps, limit-capped
__device__ __forceinline__ uint32_t match_len16(const uint8_t* in, uint64_t a, uint64_t b, uint32_t limit) {
  uint32_t k = 0;
  while (k < limit) {
    const uint32_t n = prefix16(ld128u(in, a + k), ld128u(in, b + k));
    k += n;
    if (n < 16) break;
  }
  return k < limit ? k : limit;
}

// The chain walk of hash-chains.ts:121-152 keeps the FIRST strict maximum of the score over the
// chain, newest candidate first.  One thread owns 4 consecutive sorted entries.  For each 16 chain
// steps it first builds the mask of confirmed candidates (same 4 bytes => same key run) from a
// register window over s_v, then visits only the set bits, so a warp iterates max-popcount times
// instead of 16 (profiles/r1c: 80 % of the 16 steps had a confirmed candidate in some lane and the
// whole warp paid for it).  Candidates that reach 20 bytes are deferred: their extension needs
// input reads, is rare outside repetitive data, and runs afterwards with the reference's pruning
// (a candidate that cannot beat the current best even at full length is skipped).
#ifndef PROBE_MIN_BLOCKS
#define PROBE_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(256, PROBE_MIN_BLOCKS) k_probe_sorted(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                      const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ srt_v,
                                                      uint32_t* __restrict__ best, int depth, int bucket_bits) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint4* s_x = reinterpret_cast<uint4*>(smem);
  uint32_t* s_p = reinterpret_cast<uint32_t*>(s_x + PROBE_N);
  uint32_t* s_v = s_p + PROBE_N;
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh);
  const uint32_t g0 = blockIdx.x * PROBE_TILE;
  const uint32_t tid = threadIdx.x;
  uint32_t* sbest = best + sh.in_off;
  // the <= 3 tail positions of the shard have no 4-byte key: no candidate
  if (blockIdx.x == 0 && tid < sh.len - npos) sbest[npos + tid] = 0;
  if (g0 >= npos) return;
  const uint32_t* sp = srt_pos + sh.in_off; const uint32_t* sv = srt_v + sh.in_off;
  const uint8_t* sin = in + sh.in_off;
  const uint32_t halo = (uint32_t)depth;  // multiple of 16, <= PROBE_HALO
  {
    // staging in three sweeps so that the loads of all (up to 5) entries of a thread are in flight together:
    // positions and key bytes, then the dependent 16 data bytes behind each position, then the stores
    constexpr int NST = (PROBE_TILE + PROBE_HALO + 255) / 256;
    uint32_t pp[NST], vv[NST]; uint4 xx[NST]; bool ok[NST];
#pragma unroll
    for (int t = 0; t < NST; t++) {
      const uint32_t i = tid + 256 * t;
      const int64_t g = (int64_t)g0 - halo + i;
      ok[t] = i < PROBE_TILE + halo && g >= 0 && g < (int64_t)npos;
      pp[t] = ok[t] ? __ldg(sp + g) : 0;
      vv[t] = ok[t] ? __ldg(sv + g) : 0;
    }
#pragma unroll
    // two aligned 16-byte loads per entry instead of five 4-byte ones (r3c A/B: k_probe 39.2 -> 37.4 ms per GiB)
#ifndef PROBE_LD128U
    for (int t = 0; t < NST; t++) xx[t] = ok[t] ? ld128v(sin + pp[t] + 4) : make_uint4(0, 0, 0, 0);   // reads <= 36 bytes past p: inside the padded input
#else
    for (int t = 0; t < NST; t++) xx[t] = ok[t] ? ld128u(sin, (uint64_t)pp[t] + 4) : make_uint4(0, 0, 0, 0);   // reads <= 24 bytes past p: inside the padded input
#endif
#pragma unroll
    for (int t = 0; t < NST; t++) {
      const uint32_t i = tid + 256 * t;
      if (i < PROBE_TILE + halo) { s_p[i] = pp[t]; s_v[i] = vv[t]; s_x[i] = xx[t]; }
    }
  }
  __syncthreads();
  // lanes take entries 4 apart so that a warp's shared-memory reads stay spread over the banks
  const uint32_t e0 = tid * 4;
  uint32_t bLen[4], bDist[4]; int bScore[4]; uint32_t longm[4];
#pragma unroll
  for (int k = 0; k < 4; k++) { bLen[k] = 0; bDist[k] = 0; bScore[k] = 0; }
  for (uint32_t jc = 0; jc < (uint32_t)depth; jc += 16) {
    // ---- candidate masks.  Steps jc+1 .. jc+16 of entry e look at entries e-jc-1 .. e-jc-16.
    uint32_t mask[4];
    {
      uint32_t win[20];   // s_v[halo + e0 - jc - 16 .. halo + e0 - jc + 3]
      const uint32_t* src = s_v + halo + e0 - jc - 16;
#pragma unroll
      for (int q = 0; q < 5; q++) {
        const uint4 t4 = *reinterpret_cast<const uint4*>(src + 4 * q);
        win[4 * q] = t4.x; win[4 * q + 1] = t4.y; win[4 * q + 2] = t4.z; win[4 * q + 3] = t4.w;
      }
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t v = s_v[halo + e0 + k];
        uint32_t m = 0;
#pragma unroll
        for (int j = 1; j <= 16; j++) m |= (win[16 + k - j] == v ? 1u : 0u) << (j - 1);
        // steps beyond the start of the shard / past the last entry do not exist
        const uint32_t g = g0 + e0 + k;
        if (g >= npos) m = 0;
        else if (g < jc + 16) m &= g > jc ? ((1u << (g - jc)) - 1u) : 0u;
        mask[k] = m;
      }
    }
    // ---- confirmed candidates, newest first; up to 20 bytes from shared memory.  PROBE_WIDE entries of the
    // thread advance together, one candidate each per round, so that their shared-memory gathers
    // (position + 16 data bytes of the candidate; profiles/r1t: short-scoreboard stalls on exactly these
    // loads are 35 % of the kernel's samples) are in flight at once instead of one after the other.
#ifndef PROBE_WIDE
#define PROBE_WIDE 2
#endif
#pragma unroll
    for (int k0 = 0; k0 < 4; k0 += PROBE_WIDE) {
      uint32_t pk[PROBE_WIDE], mlen[PROBE_WIDE], mk[PROBE_WIDE], lmk[PROBE_WIDE]; uint4 xk[PROBE_WIDE];
      uint32_t any = 0;
#pragma unroll
      for (int k = 0; k < PROBE_WIDE; k++) {
        pk[k] = s_p[halo + e0 + k0 + k];
        mlen[k] = min(sh.len - pk[k], 128u);
        xk[k] = s_x[halo + e0 + k0 + k];
        mk[k] = mask[k0 + k]; lmk[k] = 0; any |= mk[k];
      }
      while (any) {
        uint32_t cp[PROBE_WIDE], jj[PROBE_WIDE]; uint4 cx[PROBE_WIDE]; bool act[PROBE_WIDE];
#pragma unroll
        for (int k = 0; k < PROBE_WIDE; k++) {
          act[k] = mk[k] != 0;
          jj[k] = act[k] ? (uint32_t)__ffs((int)mk[k]) - 1 : 0u;
          mk[k] &= mk[k] - 1;                       // 0 stays 0
          const uint32_t ci = halo + e0 + k0 + k - jc - jj[k] - 1;
          cp[k] = act[k] ? s_p[ci] : 0u;
          cx[k] = act[k] ? s_x[ci] : make_uint4(0, 0, 0, 0);
        }
        any = 0;
#pragma unroll
        for (int k = 0; k < PROBE_WIDE; k++) {
          if (act[k]) {
            const uint32_t backward = pk[k] - cp[k];
            if (backward > MAX_BACKWARD) mk[k] = 0;   // hash-chains.ts:124: the walk ends here (only within 16 bytes of a 4 MiB shard's end)
            else {
              uint32_t len = 4 + prefix16(xk[k], cx[k]);
              if (len == 20 && mlen[k] > 20) lmk[k] |= 1u << jj[k];
              else {
                len = min(len, mlen[k]);
                const int score = 135 * (int)len + 960 - 30 * log2_floor(backward);
                // ties keep the earlier (nearer) candidate; a deferred long candidate is always nearer than
                // the ones after it and longer than 20 bytes, so ordering against it is settled below
                if (score > bScore[k0 + k]) { bScore[k0 + k] = score; bLen[k0 + k] = len; bDist[k0 + k] = backward; }
              }
            }
          }
          any |= mk[k];
        }
      }
#pragma unroll
      for (int k = 0; k < PROBE_WIDE; k++) longm[k0 + k] = lmk[k];
    }
    // ---- deferred candidates of >= 20 bytes (input reads), newest first
#pragma unroll
    for (int k = 0; k < 4; k++) {
      uint32_t lm = longm[k];
      if (!lm) continue;
      const uint32_t e = e0 + k;
      const uint32_t p = s_p[halo + e];
      const uint32_t maxLen = min(sh.len - p, 128u);
      while (lm) {
        const uint32_t jj = __ffs((int)lm) - 1; lm &= lm - 1;
        const uint32_t c = s_p[halo + e - jc - jj - 1];
        const uint32_t backward = p - c;
        const int dcost = 960 - 30 * log2_floor(backward);
        // Skip when even the full length cannot win.  Against an equal score the nearer candidate
        // wins, so ">=" when the current best is nearer, ">" when it is farther.
        const int ub = 135 * (int)maxLen + dcost;
        if (ub < bScore[k] || (ub == bScore[k] && bDist[k] <= backward)) continue;
        const uint32_t len = 20 + match_len16(sin, p + 20, c + 20, maxLen - 20);
        const int score = 135 * (int)len + dcost;
        if (score > bScore[k] || (score == bScore[k] && backward < bDist[k])) { bScore[k] = score; bLen[k] = len; bDist[k] = backward; }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint32_t e = e0 + k;
    if (g0 + e < npos) sbest[s_p[halo + e]] = (bLen[k] << 24) | bDist[k];
  }
}

// Lane-per-entry form of the same probe ("diagonal": at chain step j every lane of a warp looks at the entry j places
// before its own).  Consecutive lanes own consecutive sorted entries, so each step is one conflict-free row of shared
// memory per warp: the 4 key bytes of entries e - j (one wavefront), and for the lanes whose candidate is confirmed its
// position and 16 data bytes (contiguous 16-byte words, four wavefronts) — where k_probe_sorted gathers 16-byte words
// at random offsets (r3g: 3.8 G bank-conflict wavefronts per GiB on top of ~1.3 G useful ones; here 0.1 G, r5e).
// Same candidates, same order, same scores: bit-identical best[] (tools/probe_bench.cu, tools/ab_probe.py).
//   * no per-step chain-limit state: the walk's only early end (backward > MAX_BACKWARD, hash-chains.ts:124) equals
//     skipping every such candidate, because candidates come in order of growing distance; steps that would reach
//     before the shard's first entry exist only in the first tile of a shard (EDGE);
//   * candidates come in order of growing distance, so only a LONGER one can score higher (score = 135 len + 960 -
//     30 log2 distance): the score is computed only then;
//   * matches of >= 20 bytes are deferred to the input as in k_probe_sorted.
// Issue-bound (r5e: issue active 84 %, 28 warp instructions per position at depth 16).
// Shared memory: (PROBE_TILE + depth) * 24 bytes.
template <uint32_t HAVE>   // bytes known equal from shared memory: 20 (key bytes + the 16 after them) or 16 (packed sort: 16 from the position)
__device__ __forceinline__ void probe_deferred(const uint8_t* sin, const uint32_t* s_p, uint32_t i0, uint32_t jc, uint32_t p, uint32_t mlen, uint32_t lm,
                                               uint32_t& bLen, uint32_t& bDist, int& bScore) {
  while (lm) {   // newest first
    const uint32_t jj = __ffs((int)lm) - 1; lm &= lm - 1;
    const uint32_t c = s_p[i0 - jc - jj - 1];
    const uint32_t backward = p - c;
    const int dcost = 960 - 30 * log2_floor(backward);
    // Skip when even the full length cannot win.  Against an equal score the nearer candidate
    // wins, so ">=" when the current best is nearer, ">" when it is farther.
    const int ub = 135 * (int)mlen + dcost;
    if (ub < bScore || (ub == bScore && bDist <= backward)) continue;
    const uint32_t len = HAVE + match_len16(sin, p + HAVE, c + HAVE, mlen - HAVE);
    const int score = 135 * (int)len + dcost;
    if (score > bScore || (score == bScore && backward < bDist)) { bScore = score; bLen = len; bDist = backward; }
  }
}
template <bool EDGE, uint32_t HAVE>
__device__ __forceinline__ void probe_entry(const uint8_t* sin, const uint4* s_x, const uint32_t* s_p, const uint32_t* s_v, uint32_t i0, uint32_t g,
                                            uint32_t shard_len, int depth, uint32_t* sbest) {
  const uint32_t v = s_v[i0], p = s_p[i0];
  const uint4 x = s_x[i0];
  const uint32_t mlen = min(shard_len - p, 128u);
  uint32_t bLen = 0, bDist = 0; int bScore = 0;
  const uint32_t jlim = EDGE ? min((uint32_t)depth, g) : (uint32_t)depth;
  for (uint32_t jc = 0; jc < jlim; jc += 16) {
    uint32_t lm = 0;
#pragma unroll
    for (uint32_t jj = 0; jj < 16; jj++) {
      const uint32_t ci = i0 - jc - jj - 1;
      if ((!EDGE || jc + jj < jlim) && s_v[ci] == v) {
        const uint32_t cp = s_p[ci];
        const uint4 cx = s_x[ci];
        const uint32_t backward = p - cp;
        const bool near = backward <= MAX_BACKWARD;
        uint32_t len = (HAVE - 16) + prefix16(x, cx);
        if (len == HAVE && mlen > HAVE) { if (near) lm |= 1u << jj; }
        else {
          len = min(len, mlen);
          if (len > bLen) {
            const int score = 135 * (int)len + 960 - 30 * log2_floor(backward);
            if (score > bScore && near) { bScore = score; bLen = len; bDist = backward; }
          }
        }
      }
    }
    probe_deferred<HAVE>(sin, s_p, i0, jc, p, mlen, lm, bLen, bDist, bScore);
  }
  sbest[p] = (bLen << 24) | bDist;
}
// PACKED: the entries come from the packed sort (positions only); the 4 key bytes are the first word of the entry's own
// 16-byte gather, which then starts AT the position (matches of >= 16 bytes go to the input instead of >= 20).
template <bool PACKED>
__global__ void __launch_bounds__(256, 6) k_probe_diag(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                      const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ srt_v,
                                                      uint32_t* __restrict__ best, int depth, int bucket_bits) {
  extern __shared__ __align__(16) uint8_t smem[];
  const uint32_t halo = (uint32_t)depth;  // multiple of 16, <= PROBE_HALO
  const uint32_t n_ent = PROBE_TILE + halo;
  uint4* s_x = reinterpret_cast<uint4*>(smem);
  uint32_t* s_p = reinterpret_cast<uint32_t*>(s_x + n_ent);
  uint32_t* s_v = s_p + n_ent;
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh);
  const uint32_t g0 = blockIdx.x * PROBE_TILE;
  const uint32_t tid = threadIdx.x;
  uint32_t* sbest = best + sh.in_off;
  // the <= 3 tail positions of the shard have no 4-byte key: no candidate
  if (blockIdx.x == 0 && tid < sh.len - npos) sbest[npos + tid] = 0;
  if (g0 >= npos) return;
  const uint32_t* sp = srt_pos + sh.in_off; const uint32_t* sv = srt_v + sh.in_off;
  const uint8_t* sin = in + sh.in_off;
  {
    // staging in three sweeps so that the loads of all (up to 5) entries of a thread are in flight together:
    // positions and key bytes, then the dependent 16 data bytes behind each position, then the stores
    constexpr int NST = (PROBE_TILE + PROBE_HALO + 255) / 256;
    uint32_t pp[NST], vv[NST]; uint4 xx[NST]; bool ok[NST];
#pragma unroll
    for (int t = 0; t < NST; t++) {
      const uint32_t i = tid + 256 * t;
      const int64_t g = (int64_t)g0 - halo + i;
      ok[t] = i < n_ent && g >= 0 && g < (int64_t)npos;
      pp[t] = ok[t] ? __ldg(sp + g) : 0;
      vv[t] = (ok[t] && !PACKED) ? __ldg(sv + g) : 0;
    }
#pragma unroll
    for (int t = 0; t < NST; t++) xx[t] = ok[t] ? ld128v(sin + pp[t] + (PACKED ? 0 : 4)) : make_uint4(0, 0, 0, 0);   // reads <= 36 bytes past p: inside the padded input
#pragma unroll
    for (int t = 0; t < NST; t++) {
      const uint32_t i = tid + 256 * t;
      if (i < n_ent) { s_p[i] = pp[t]; s_v[i] = PACKED ? xx[t].x : vv[t]; s_x[i] = xx[t]; }
    }
  }
  __syncthreads();
  constexpr uint32_t HAVE = PACKED ? 16 : 20;
  if (g0 < halo) {
    for (uint32_t r = 0; r < PROBE_TILE / 256; r++) { const uint32_t e = r * 256 + tid; if (g0 + e < npos) probe_entry<true, HAVE>(sin, s_x, s_p, s_v, halo + e, g0 + e, sh.len, depth, sbest); }
  } else {
    for (uint32_t r = 0; r < PROBE_TILE / 256; r++) { const uint32_t e = r * 256 + tid; if (g0 + e < npos) probe_entry<false, HAVE>(sin, s_x, s_p, s_v, halo + e, g0 + e, sh.len, depth, sbest); }
  }
}


// SimpleHasher table candidate (hash-simple.ts:64-128, quality 1..4).  The table has ONE slot per key and every
// position is stored when it is visited or covered by a match (backward-references.ts:52-60, 99-110; exact for
// quality 3-4, a superset of the every-4th-position stores of quality 1-2), so the slot read at position p holds
// the newest earlier position with p's key = the entry before p in the key-sorted order; an empty slot reads as
// position 0 (the zero-initialised Uint32Array).  One thread per sorted entry.  The first 16 bytes of every
// position are fetched ONCE (a random 16-byte gather from the L2-resident input) and handed to the next lane,
// whose candidate it is; only matches that reach 16 bytes go back to the input.
// best[p] = len << 24 | distance, 0 when the candidate is shorter than 4.
__global__ void __launch_bounds__(256) k_probe_simple(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                      const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ srt_key,
                                                      uint32_t* __restrict__ best) {
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh);
  uint32_t* sbest = best + sh.in_off;
  if (blockIdx.x == 0 && threadIdx.x < sh.len - npos) sbest[npos + threadIdx.x] = 0;
  if (blockIdx.x * 256 >= npos) return;                 // whole CTA out of range (warps stay whole below)
  const uint32_t r = blockIdx.x * 256 + threadIdx.x;
  const bool valid = r < npos;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t* sp = srt_pos + sh.in_off; const uint32_t* sk = srt_key + sh.in_off;
  const uint8_t* sin = in + sh.in_off;
  const uint32_t p = valid ? __ldg(sp + r) : 0u;
  const uint32_t key = valid ? __ldg(sk + r) : 0xFFFFFFFFu;
  const uint4 xp = valid ? ld128v(sin + p) : make_uint4(0, 0, 0, 0);
  // the previous sorted entry: lane - 1, or a load for the first lane of the warp
  uint32_t pc = __shfl_up_sync(0xffffffffu, p, 1), kc = __shfl_up_sync(0xffffffffu, key, 1);
  uint4 xc;
  xc.x = __shfl_up_sync(0xffffffffu, xp.x, 1); xc.y = __shfl_up_sync(0xffffffffu, xp.y, 1);
  xc.z = __shfl_up_sync(0xffffffffu, xp.z, 1); xc.w = __shfl_up_sync(0xffffffffu, xp.w, 1);
  if (lane == 0) {
    if (valid && r > 0) { pc = __ldg(sp + r - 1); kc = __ldg(sk + r - 1); xc = ld128v(sin + pc); }
    else kc = 0xFFFFFFFEu;
  }
  if (!valid) return;
  uint32_t c = pc;
  if (kc != key) { c = 0; xc = ld128v(sin); }         // empty slot: position 0 (same address for every lane)
  const uint32_t backward = p - c;
  uint32_t out = 0;
  if (backward != 0 && backward <= MAX_BACKWARD) {     // hash-simple.ts:109-112 (maxBackward = min(p, 4 MiB - 16), c >= 0)
    const uint32_t limit = min(sh.len - p, 128u);
    uint32_t len = prefix16(xp, xc);
    if (len == 16 && limit > 16) len = 16 + match_len16(sin, (uint64_t)c + 16, (uint64_t)p + 16, limit - 16);
    len = min(len, limit);
    if (len >= 4) out = (len << 24) | backward;
  }
  sbest[p] = out;
}

}  // namespace bgpu
