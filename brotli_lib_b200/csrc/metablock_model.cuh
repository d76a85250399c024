// Parts C and D of the full metablock (see metablock.cuh): final block lists, context histograms,
// context maps, prefix codes and the metablock header.
#pragma once
#include "metablock.cuh"

namespace bgpu {

// ---------------------------------------------------------------------------------------
// C.  clusterBlocks hand-over: the clustering problem of (metablock, category) takes the
// per-block histograms when the split produced more than one block.  grid = problems.
__global__ void k_cb_prep(const MbModel* __restrict__ models, ClusterProb* __restrict__ cb, uint32_t n_prob) {
  const uint32_t pr = blockIdx.x * blockDim.x + threadIdx.x;
  if (pr >= n_prob) return;
  const MbModel& M = models[pr / 3];
  const uint32_t nb = M.sstate[pr % 3][1];
  cb[pr].m = nb > 1 ? nb : 0;
  cb[pr].max_clusters = M.max_types;
}

// Block types from the clustering (block-splitter.ts:366-389: merge neighbours of equal type,
// types numbered by first appearance) and the metablock gates of metablock.ts:518, 531.
// One warp per metablock; lanes 0..2 own one category each.
__global__ void __launch_bounds__(32) k_blocks_final(const MbModel* __restrict__ models, const ClusterProb* __restrict__ cb) {
  const MbModel& M = models[blockIdx.x];
  const uint32_t lane = threadIdx.x;
  if (lane < 3) {
    const ClusterProb& C = cb[blockIdx.x * 3 + lane];
    uint32_t* st = M.sstate[lane]; uint32_t* bs = M.blk_start[lane]; uint8_t* bt = M.blk_type[lane];
    const uint32_t nb = st[1], n = M.cnt[lane];
    uint32_t ntypes = 1;
    if (nb > 1) {
      const uint32_t ncl = C.state[2];
      uint32_t out = 0;
      for (uint32_t b = 0; b < nb; b++) {
        const uint32_t t = ncl ? min(C.sym[b], 255u) : 0;
        if (out && bt[out - 1] == t) continue;       // same type as the previous block: it simply grows
        bs[out] = bs[b]; bt[out] = (uint8_t)t; out++;
        if (t + 1 > ntypes) ntypes = t + 1;
      }
      bs[out] = n;
      st[1] = out;
    } else if (nb == 1) { bt[0] = 0; bs[0] = 0; bs[1] = n; }
    st[2] = ntypes;
  }
  __syncwarp();
  if (lane == 0) {
    const bool trivial = M.len < 128 || M.cnt[CMD] < 6 || (M.sstate[LIT][2] <= 1 && M.sstate[CMD][2] <= 1 && M.sstate[DIST][2] <= 1);
    M.cnt[4] = trivial ? 1 : 0;
    if (trivial) {
      M.cnt[3] = 0;   // storeMetaBlockTrivial writes context mode 0 (metablock.ts:338)
      for (int c = 0; c < 3; c++) {
        const uint32_t n = M.cnt[c];
        M.sstate[c][1] = n ? 1 : 0; M.sstate[c][2] = 1;
        M.blk_start[c][0] = 0; M.blk_start[c][1] = n; M.blk_type[c][0] = 0;
      }
    }
  }
}

// Block index of every symbol.  grid (chunks of 2048 symbols, metablocks * 3).
__global__ void __launch_bounds__(256) k_blk_expand(const MbModel* __restrict__ models) {
  const MbModel& M = models[blockIdx.y / 3];
  const int cat = blockIdx.y % 3;
  const uint32_t n = M.cnt[cat], nb = M.sstate[cat][1];
  const uint32_t i0 = blockIdx.x * 2048;
  if (i0 >= n) return;
  const uint32_t* bs = M.blk_start[cat];
  uint16_t* out = M.sym_blk[cat];
  for (uint32_t i = i0 + threadIdx.x; i < min(n, i0 + 2048); i += 256) {
    uint32_t lo = 0, hi = nb;   // last block with start <= i
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (bs[mid] <= i) lo = mid; else hi = mid; }
    out[i] = (uint16_t)lo;
  }
}

// ---------------------------------------------------------------------------------------
// D.  Zeroes the histogram rows the metablock will use.  grid (64, metablocks).
__global__ void __launch_bounds__(256) k_ctx_zero(const MbModel* __restrict__ models) {
  const MbModel& M = models[blockIdx.y];
  const bool trivial = M.cnt[4] != 0;
  const size_t nl = (size_t)(trivial ? 1 : M.sstate[LIT][2] * 64) * 256;
  const size_t nc = (size_t)M.sstate[CMD][2] * ALPHA_CMD;
  const size_t nd = (size_t)(trivial ? 1 : M.sstate[DIST][2] * 4) * ALPHA_DIST_MAX;
  const size_t step = (size_t)gridDim.x * 256, first = (size_t)blockIdx.x * 256 + threadIdx.x;
  for (size_t i = first; i < nl; i += step) M.ctx_hist[LIT][i] = 0;
  for (size_t i = first; i < nc; i += step) M.ctx_hist[CMD][i] = 0;
  for (size_t i = first; i < nd; i += step) M.ctx_hist[DIST][i] = 0;
  for (size_t i = first; i < (size_t)MAX_CTX_CLUSTERS * 256; i += step) M.cl_hist[LIT][i] = 0;
  for (size_t i = first; i < (size_t)MAX_CTX_CLUSTERS * ALPHA_DIST_MAX; i += step) M.cl_hist[DIST][i] = 0;
}

// Literal histograms of a chunk that lies inside ONE block (the common case): counted in a CTA-private
// [64 contexts][256 bytes] table of 16-bit counters in shared memory and flushed with one global atomic
// per non-zero counter.  The plain per-symbol global atomics below pile up on the few hot (context, byte)
// pairs of a text block (tens of thousands of increments of one address per metablock).
// grid (chunks of 4096 literals, metablocks).  Chunks that span a block boundary are left to k_ctx_hist.
__global__ void __launch_bounds__(256) k_ctx_hist_lit(const MbModel* __restrict__ models) {
  __shared__ uint32_t s_h[64 * 256 / 2];   // two 16-bit counters per word: a chunk holds <= 4096 symbols
  const MbModel& M = models[blockIdx.y];
  const uint32_t n = M.cnt[LIT];
  const uint32_t i0 = blockIdx.x * 4096;
  if (i0 >= n) return;
  const uint32_t i1 = min(n, i0 + 4096);
  const uint16_t* sb = M.sym_blk[LIT];
  const uint32_t b0 = sb[i0];
  if (b0 != sb[i1 - 1]) return;            // not uniform: k_ctx_hist counts this chunk
  const bool trivial = M.cnt[4] != 0;
  const uint32_t t = M.blk_type[LIT][b0];
  for (uint32_t i = threadIdx.x; i < 64 * 256 / 2; i += 256) s_h[i] = 0;
  __syncthreads();
  for (uint32_t i = i0 + threadIdx.x; i < i1; i += 256) {
    const uint32_t idx = (trivial ? 0u : (uint32_t)M.clctx[i]) * 256 + M.clit[i];
    atomicAdd(&s_h[idx >> 1], 1u << (16 * (idx & 1)));
  }
  __syncthreads();
  uint32_t* h = M.ctx_hist[LIT] + (size_t)(trivial ? 0 : t * 64) * 256;
  for (uint32_t i = threadIdx.x; i < 64 * 256 / 2; i += 256) {
    const uint32_t v = s_h[i];
    if (v & 0xFFFFu) atomicAdd(&h[2 * i], v & 0xFFFFu);
    if (v >> 16) atomicAdd(&h[2 * i + 1], v >> 16);
  }
}

// metablock.ts:558-637 in rank space.  grid (chunks of 4096 symbols, metablocks * 3).
__global__ void __launch_bounds__(256) k_ctx_hist(const MbModel* __restrict__ models) {
  const MbModel& M = models[blockIdx.y / 3];
  const int cat = blockIdx.y % 3;
  const uint32_t n = M.cnt[cat];
  const uint32_t i0 = blockIdx.x * 4096;
  if (i0 >= n) return;
  const uint32_t i1 = min(n, i0 + 4096);
  const uint16_t* sb = M.sym_blk[cat];
  const bool uniform = sb[i0] == sb[i1 - 1];
  if (cat == LIT && uniform) return;   // k_ctx_hist_lit took it
  const bool trivial = M.cnt[4] != 0;
  const uint8_t* bt = M.blk_type[cat];
  uint32_t* h = M.ctx_hist[cat];
  if (uniform) {
    // command / distance chunk inside one block: CTA-private counters (704, or 4 contexts x 128), one global
    // atomic per non-zero counter — the command alphabet has a handful of very hot symbols
    __shared__ uint32_t s_h[ALPHA_CMD > 4 * ALPHA_DIST_MAX ? ALPHA_CMD : 4 * ALPHA_DIST_MAX];
    const uint32_t rows = cat == CMD ? ALPHA_CMD : 4 * ALPHA_DIST_MAX;
    for (uint32_t i = threadIdx.x; i < rows; i += 256) s_h[i] = 0;
    __syncthreads();
    for (uint32_t i = i0 + threadIdx.x; i < i1; i += 256) {
      if (cat == CMD) atomicAdd(&s_h[M.csym[i]], 1u);
      else atomicAdd(&s_h[(trivial ? 0u : (uint32_t)M.ddctx[i]) * ALPHA_DIST_MAX + M.dsym[i]], 1u);
    }
    __syncthreads();
    const uint32_t t = bt[sb[i0]];
    uint32_t* hb = cat == CMD ? h + (size_t)t * ALPHA_CMD : h + (size_t)(trivial ? 0 : t * 4) * ALPHA_DIST_MAX;
    for (uint32_t i = threadIdx.x; i < rows; i += 256) { const uint32_t v = s_h[i]; if (v) atomicAdd(&hb[i], v); }
    return;
  }
  for (uint32_t i = i0 + threadIdx.x; i < i1; i += 256) {
    const uint32_t t = bt[sb[i]];
    if (cat == LIT) atomicAdd(&h[((size_t)(trivial ? 0 : t * 64 + M.clctx[i])) * 256 + M.clit[i]], 1u);
    else if (cat == CMD) atomicAdd(&h[(size_t)t * ALPHA_CMD + M.csym[i]], 1u);
    else atomicAdd(&h[((size_t)(trivial ? 0 : t * 4 + M.ddctx[i])) * ALPHA_DIST_MAX + M.dsym[i]], 1u);
  }
}

// Sizes of the two context-map clustering problems of every metablock (literal: index 2m,
// distance: 2m + 1).  metablock.ts:640-649.
__global__ void k_ctx_prep(const MbModel* __restrict__ models, ClusterProb* __restrict__ cx, uint32_t n_mb) {
  const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= n_mb) return;
  const MbModel& M = models[m];
  const bool trivial = M.cnt[4] != 0;
  cx[2 * m].m = trivial ? 1 : M.sstate[LIT][2] * 64;
  cx[2 * m + 1].m = trivial ? 1 : M.sstate[DIST][2] * 4;
  cx[2 * m].max_clusters = MAX_CTX_CLUSTERS; cx[2 * m + 1].max_clusters = MAX_CTX_CLUSTERS;
}

// buildClusteredHistograms (metablock.ts:799-819): out[sym[i]] += in[i].  grid (row chunks of 64, problems).
__global__ void __launch_bounds__(256) k_cluster_sum(const ClusterProb* __restrict__ probs, uint32_t* const* __restrict__ outs) {
  const ClusterProb p = probs[blockIdx.y];
  const uint32_t r0 = blockIdx.x * 64;
  if (r0 >= p.m) return;
  uint32_t* out = outs[blockIdx.y];
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (uint32_t r = r0 + w; r < min(p.m, r0 + 64); r += 8) {
    const uint32_t c = p.sym[r];
    const uint32_t* a = p.in + (size_t)r * p.stride;
    for (uint32_t k = lane; k < p.alpha; k += 32) { const uint32_t v = a[k]; if (v) atomicAdd(&out[(size_t)c * p.stride + k], v); }
  }
}

// SURVEY.md section 8d table bytes of the modelling stages, summed over the metablocks (BGPU_F_MODEL_STATS).
// out[0] += 4 A H_in, out[1] += 4 A H_out of the clustering problems (clusterBlocks per category, then the two context
// maps), out[2] += 7 A H of the prefix codes actually built.  One thread per metablock.
__global__ void k_model_stats(const MbModel* __restrict__ models, uint32_t n_mb, uint32_t dist_alphabet, unsigned long long* __restrict__ out) {
  const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= n_mb) return;
  const MbModel& M = models[m];
  const uint32_t A[3] = {ALPHA_LIT, ALPHA_CMD, dist_alphabet};
  unsigned long long in_b = 0, out_b = 0, huff = 0;
  const bool trivial = M.cnt[4] != 0;
  for (int c = 0; c < 3; c++) {
    const uint32_t nb = M.sstate[c][1], nt = M.sstate[c][2];
    if (nb > 1) { in_b += 4ull * A[c] * nb; out_b += 4ull * A[c] * nt; }
  }
  if (!trivial) {
    in_b += 4ull * 256 * M.sstate[LIT][2] * 64 + 4ull * dist_alphabet * M.sstate[DIST][2] * 4;
    out_b += 4ull * 256 * M.cstate[LIT][2] + 4ull * dist_alphabet * M.cstate[DIST][2];
  }
  huff = 7ull * (256ull * M.cstate[LIT][2] + (unsigned long long)ALPHA_CMD * M.sstate[CMD][2] + (unsigned long long)dist_alphabet * M.cstate[DIST][2]);
  atomicAdd(&out[0], in_b); atomicAdd(&out[1], out_b); atomicAdd(&out[2], huff);
}

// Prefix codes of a metablock.  grid (HDR_TREE_SLOTS, metablocks), 128 threads.
// slot < 64: literal cluster; 64..319: command block type; 320..383: distance cluster.
__global__ void __launch_bounds__(128) k_huff_mb(const MbModel* __restrict__ models, int dist_alphabet) {
  const MbModel& M = models[blockIdx.y];
  const uint32_t slot = blockIdx.x;
  const uint32_t ncl_l = M.cstate[LIT][2], ntc = M.sstate[CMD][2], ncl_d = M.cstate[DIST][2];
  const uint32_t* hist; uint32_t* code; int alphabet;
  if (slot < 64) { if (slot >= ncl_l) return; hist = M.cl_hist[LIT] + slot * 256; code = M.codes[LIT] + slot * 256; alphabet = 256; }
  else if (slot < 320) { const uint32_t t = slot - 64; if (t >= ntc) return; hist = M.ctx_hist[CMD] + (size_t)t * ALPHA_CMD; code = M.codes[CMD] + (size_t)t * ALPHA_CMD; alphabet = ALPHA_CMD; }
  else { const uint32_t c = slot - 320; if (c >= ncl_d) return; hist = M.cl_hist[DIST] + c * ALPHA_DIST_MAX; code = M.codes[DIST] + c * ALPHA_DIST_MAX; alphabet = dist_alphabet; }
  huff_cta(hist, alphabet, code, M.tree_bits + (size_t)slot * TREE_WORDS, M.tree_nbits + slot);
}

}  // namespace bgpu
