// sm_100a Zopfli-style shortest-path parse for q10/q11, scanned per parse segment.
//
// Reference semantics (paths relative to the reference tree, src/encode/):
//   zopfli-cost-model.ts:29-65,163-189   pass-1 model: literal cost from the metablock's byte histogram,
//                                         cmd cost log2(11 + i), distance cost log2(20 + i)
//   zopfli-cost-model.ts:68-161          pass-2 model: Shannon costs from the previous pass's commands
//   backward-references-hq.ts:93-149     StartPosQueue (8 entries ordered by costdiff)
//   backward-references-hq.ts:168-190    computeMinimumCopyLength
//   backward-references-hq.ts:239-382    evaluateNode / updateNodes (<= 5 starts x 16 distance-cache
//                                         codes, <= 2 starts x the position's H10 matches)
//   backward-references-hq.ts:384-406, 610-673  backtrack + command creation
//
// B200 formulation.  The DP over positions is a dependent chain, so the unit of parallelism is the
// parse segment (north_star: "scanned per metablock segment"): one warp per segment, thousands of
// segments in flight.  Inside a position the warp is data-parallel:
//   * the <= 80 (start, distance-cache code) probes are one lane each (match length from the
//     L2-resident input), a 16-lane exclusive prefix-max turns them into disjoint length ranges
//     (the reference's `bestLen` bookkeeping), and every surviving range is relaxed with one lane
//     per copy length;
//   * the H10 match list is relaxed with one lane per copy length, starts 0 and 1 in the two half warps.
// A segment starts with an unknown distance ring (zeros: never referenced) unless it is the first of
// its stream, so the ring the encoder assumes is always a suffix-consistent view of the decoder's.
// Instead of walking `shortcut` links (computeDistanceCache), the ring of a node is derived when the
// node enters the queue from the ring stored for its start node (one 16-byte read): same values.
// Costs are f32 (the reference keeps f64 node costs over f32 tables); this stage is size-parity only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bgpu {

constexpr float ZOP_INF = 1.7e38f;
constexpr int ZOP_DIST_COSTS = 544;                 // zopfli-cost-model.ts: costDist length
constexpr int ZOP_MODEL_STRIDE = 256 + 704 + ZOP_DIST_COSTS + 16;   // lit | cmd | dist | [0] = minCostCmd

struct ZopParams { uint32_t max_zopfli_len, max_iters; };   // enc-constants.ts:103-109

// Byte histogram of every metablock (pass-1 literal model).  One CTA per bit-writer tile.
__global__ void __launch_bounds__(256) k_zop_bytehist(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                      const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                      uint32_t n_tile, uint32_t* __restrict__ hist /* [n_mb][256] */) {
  __shared__ uint32_t s_h[256];
  const uint32_t tile = blockIdx.x;
  if (tile >= n_tile) return;
  s_h[threadIdx.x] = 0;
  __syncthreads();
  const SegD sg = segs[tile_seg[tile]];
  const ShardD sh = shards[sg.shard];
  const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
  const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
  for (uint32_t rel = tile_rel + threadIdx.x; rel < tile_end; rel += 256) atomicAdd(&s_h[in[sh.in_off + rel]], 1u);
  __syncthreads();
  const uint32_t v = s_h[threadIdx.x];
  if (v) atomicAdd(&hist[(size_t)sg.mb * 256 + threadIdx.x], v);
}

// Cost tables of every metablock.  pass2 == 0: from the byte histogram; pass2 == 1: from the
// literal / command / distance histograms of the previous pass's commands (MbHist layout of k_hist).
__global__ void __launch_bounds__(256) k_zop_model(const uint32_t* __restrict__ bytehist, const uint32_t* __restrict__ mbhist,
                                                   const MbD* __restrict__ mbs, int pass2, float* __restrict__ model) {
  __shared__ double s_red[256];
  __shared__ double s_tot[3];
  __shared__ uint32_t s_nz[3];
  const uint32_t mb = blockIdx.x, tid = threadIdx.x;
  float* M = model + (size_t)mb * ZOP_MODEL_STRIDE;
  if (!pass2) {
    // zopfli-cost-model.ts:163-189
    const uint32_t h = bytehist[(size_t)mb * 256 + tid];
    const double l2t = log2((double)mbs[mb].len);
    double c = h == 0 ? l2t + 2.0 : fmax(l2t - log2((double)h), 1.0);
    M[tid] = (float)c;
    for (uint32_t i = tid; i < 704; i += 256) M[256 + i] = (float)log2(11.0 + i);
    for (uint32_t i = tid; i < ZOP_DIST_COSTS; i += 256) M[960 + i] = (float)log2(20.0 + i);
    if (tid == 0) M[960 + ZOP_DIST_COSTS] = (float)log2(11.0);
    return;
  }
  // zopfli-cost-model.ts:126-161 setCostFromHistogram for the three alphabets
  const uint32_t* H = mbhist + (size_t)mb * CODE_STRIDE;
  const int off[3] = {0, ALPHA_LIT, ALPHA_LIT + ALPHA_CMD}, cnt[3] = {256, 704, ZOP_DIST_COSTS}, avail[3] = {256, 704, ALPHA_DIST_MAX};
  for (int a = 0; a < 3; a++) {
    double sum = 0; uint32_t zeros = 0;
    for (int i = tid; i < cnt[a]; i += 256) { const uint32_t h = i < avail[a] ? H[off[a] + i] : 0; sum += h; zeros += h == 0; }
    s_red[tid] = sum;
    __syncthreads();
    for (int o = 128; o; o >>= 1) { if ((int)tid < o) s_red[tid] += s_red[tid + o]; __syncthreads(); }
    if (tid == 0) s_tot[a] = s_red[0];
    __syncthreads();
    s_red[tid] = zeros;
    __syncthreads();
    for (int o = 128; o; o >>= 1) { if ((int)tid < o) s_red[tid] += s_red[tid + o]; __syncthreads(); }
    if (tid == 0) s_nz[a] = (uint32_t)s_red[0];
    __syncthreads();
  }
  const int moff[3] = {0, 256, 960};
  for (int a = 0; a < 3; a++) {
    const double sum = s_tot[a];
    const double l2s = sum > 0 ? log2(sum) : 0.0;                 // fastLog2(0) == 0
    const double missingSum = a == 0 ? sum : sum + (double)s_nz[a];
    const double missing = (missingSum > 0 ? log2(missingSum) : 0.0) + 2.0;
    for (int i = tid; i < cnt[a]; i += 256) {
      const uint32_t h = i < avail[a] ? H[off[a] + i] : 0;
      float c = h == 0 ? (float)missing : (float)(l2s - log2((double)h));
      if (h != 0 && c < 1.0f) c = 1.0f;
      M[moff[a] + i] = c;
    }
  }
  __syncthreads();
  // minCostCmd (zopfli-cost-model.ts:113-118)
  float mn = ZOP_INF;
  for (int i = tid; i < 704; i += 256) mn = fminf(mn, M[256 + i]);
  s_red[tid] = mn;
  __syncthreads();
  for (int o = 128; o; o >>= 1) { if ((int)tid < o) s_red[tid] = fmin(s_red[tid], s_red[tid + o]); __syncthreads(); }
  if (tid == 0) M[960 + ZOP_DIST_COSTS] = (float)s_red[0];
}

// Exclusive prefix sums of the literal costs inside every segment: lc[p] = cost of bytes [seg start, p).
// One warp per segment.
__global__ void __launch_bounds__(128) k_zop_litcost(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                     const SegD* __restrict__ segs, uint32_t n_seg, const float* __restrict__ model,
                                                     float* __restrict__ lc) {
  const uint32_t seg = (blockIdx.x * 128 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (seg >= n_seg) return;
  const SegD sg = segs[seg];
  const ShardD sh = shards[sg.shard];
  const float* M = model + (size_t)sg.mb * ZOP_MODEL_STRIDE;
  const uint8_t* p = in + sh.in_off + sg.rel_pos;
  float* out = lc + sh.in_off + sg.rel_pos;
  float carry = 0.f;
  for (uint32_t base = 0; base < sg.len; base += 128) {
    // 4 consecutive bytes per lane
    float c[4]; float s = 0.f;
#pragma unroll
    for (int k = 0; k < 4; k++) { const uint32_t q = base + lane * 4 + k; c[k] = q < sg.len ? __ldg(M + p[q]) : 0.f; s += c[k]; }
    float incl = s;
    for (int o = 1; o < 32; o <<= 1) { const float y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
    float run = carry + incl - s;
#pragma unroll
    for (int k = 0; k < 4; k++) { const uint32_t q = base + lane * 4 + k; if (q < sg.len) out[q] = run; run += c[k]; }
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
}

}  // namespace bgpu
#include "zopfli_dp.cuh"
