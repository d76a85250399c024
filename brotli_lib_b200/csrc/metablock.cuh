// sm_100a kernels of the full (non-trivial) metablock: storeMetaBlock of the reference
// (src/encode/metablock.ts:504-761) re-expressed as data-parallel passes.
//
//   A  k_mb_counts / k_mb_scan / k_mb_compact   literal / command / distance symbol streams of every
//                                               metablock in rank order (block-splitter.ts:405-427)
//      k_mb_ctxmode                             chooseContextMode (context.ts:179-227)
//   B  split_cluster.cuh                        splitBlock + clusterBlocks (K8, K9)
//   C  k_mb_gate / k_blocks_final / k_blk_expand  metablock.ts:518,531 gates, final block lists
//   D  k_ctx_hist_*                             per-(block type, context) histograms (metablock.ts:558-637)  K7
//      cluster kernels again                    context maps (metablock.ts:763-797)                    K9
//      k_cluster_sum, k_huff_mb                 clustered histograms, prefix codes                     K10
//      k_mb_header                              block-switch codes, context maps, header bits (metablock.ts:161-255, context-map.ts:114-170)
//   E  k_tilebits2 / k_offsets2 / k_write2      StoreMetaBlock emit loop (metablock.ts:714-756)         K11
#pragma once
#include "kernels.cuh"
#include "split_cluster.cuh"

namespace bgpu {

constexpr int LIT = 0, CMD = 1, DIST = 2;
constexpr int MAX_CTX_CLUSTERS = 64;          // metablock.ts:789
constexpr int HDR_TREE_SLOTS = 64 + 256 + 64; // literal clusters, command types, distance clusters

// Everything the kernels need to know about one metablock.  Pointers and capacities are set by
// the host (bgpu.cu plan_model); counts live in device memory because they come out of kernels.
struct MbModel {
  uint64_t abs;              // absolute offset of the metablock's first input byte
  uint32_t len, shard, first_tile, n_tile, is_last, stream_start;
  uint32_t max_blocks, max_types;
  uint32_t* cnt;             // [16] 0..2 symbol counts (L, C, D); 3 context mode; 4 trivial flag; 5 misc header bits;
                             //      6 literal clusters; 7 distance clusters
  uint32_t* tile_base;       // [n_tile][3] rank of the tile's first literal / command / distance
  // rank-space symbol streams
  uint8_t* clit;  uint8_t* clctx;  uint16_t* csym;  uint16_t* dsym;  uint8_t* ddctx;
  uint16_t* sym_blk[3];      // block index of every symbol
  // split state (also inside the SplitProb of the category)
  uint32_t* blk_start[3];    // [max_blocks + 1]
  uint8_t* blk_type[3];      // [max_blocks]
  uint32_t* sstate[3];       // SplitProb.state: 0 K, 1 number of blocks, 2 number of block types
  uint64_t* blk_tok[3];      // [max_blocks] block-switch token: bits << 56 | value
  // context modelling
  uint32_t* ctx_hist[3];     // LIT [types * 64][256]; CMD [types][704]; DIST [types * 4][128]
  uint32_t* cl_hist[3];      // LIT [64][256] clustered; CMD = ctx_hist[CMD]; DIST [64][128]
  uint32_t* cmap[3];         // LIT [types * 64], DIST [types * 4] cluster of every context (ClusterProb.sym)
  uint32_t* cstate[3];       // ClusterProb.state of the context problem (2: cluster count)
  uint32_t* codes[3];        // LIT [64][256]; CMD [types][704]; DIST [64][128]   bits << 8 | depth
  uint32_t* tree_bits;       // [HDR_TREE_SLOTS][TREE_WORDS]
  uint32_t* tree_nbits;      // [HDR_TREE_SLOTS]
  uint32_t* misc;            // header bits before the prefix codes
  uint32_t misc_words;
  uint32_t* scratch;         // [max_types * 64] context-map serialisation scratch
};

__constant__ uint8_t c_ctx_lut[2048];   // context.ts:12-162 (filled by the host from RFC 7932 §7.1 rules)

__device__ __forceinline__ bool cmd_has_dist(uint32_t role, uint32_t cmdPrefix) { return (role & 2) && cmdPrefix >= 128; }

// Roles of 4 consecutive positions as ONE aligned 32-bit load (shards are 256-byte aligned, tiles start at
// multiples of 1024) and the command prefixes of the copy starts among them fetched together rather than
// one dependent pair after the other.  Positions at or beyond tile_end read as role 0.
__device__ __forceinline__ void load_roles4(const uint8_t* __restrict__ role, const Cmd* __restrict__ cmds, uint64_t p0, uint32_t rel0,
                                            uint32_t tile_end, int shft, uint32_t r[4], uint32_t pre[4]) {
  uint32_t r4 = rel0 < tile_end ? *reinterpret_cast<const uint32_t*>(role + p0) : 0u;
  if (rel0 + 4 > tile_end && rel0 < tile_end) r4 &= (1u << (8 * (tile_end - rel0))) - 1u;   // ragged last tile of a segment
#pragma unroll
  for (int k = 0; k < 4; k++) {
    r[k] = (r4 >> (8 * k)) & 0xFFu;
    pre[k] = (r[k] & 2) ? cmds[(p0 + k) >> shft].cmdPrefix : 0u;
  }
}

// ---------------------------------------------------------------------------------------
// A.  Per-tile symbol counts.  grid = tiles, 256 threads (4 positions each).
__global__ void __launch_bounds__(256) k_mb_counts(const ShardD* __restrict__ shards, const SegD* __restrict__ segs,
                                                   const uint32_t* __restrict__ tile_seg, const MbModel* __restrict__ models,
                                                   const uint8_t* __restrict__ role, const Cmd* __restrict__ cmds, Params P) {
  __shared__ uint32_t s_c[3];
  const uint32_t tile = blockIdx.x;
  const SegD sg = segs[tile_seg[tile]];
  const ShardD sh = shards[sg.shard];
  const MbModel& M = models[sg.mb];
  if (threadIdx.x < 3) s_c[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
  const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
  uint32_t nl = 0, nc = 0, nd = 0;
  {
    uint32_t r[4], pre[4];
    const uint32_t rel0 = tile_rel + 4 * threadIdx.x;
    load_roles4(role, cmds, sh.in_off + rel0, rel0, tile_end, P.sh, r, pre);
#pragma unroll
    for (int k = 0; k < 4; k++) { nl += r[k] & 1; nc += (r[k] >> 2) & 1; nd += cmd_has_dist(r[k], pre[k]) ? 1 : 0; }
  }
  for (int o = 16; o; o >>= 1) { nl += __shfl_xor_sync(0xffffffffu, nl, o); nc += __shfl_xor_sync(0xffffffffu, nc, o); nd += __shfl_xor_sync(0xffffffffu, nd, o); }
  if ((threadIdx.x & 31) == 0) { atomicAdd(&s_c[0], nl); atomicAdd(&s_c[1], nc); atomicAdd(&s_c[2], nd); }
  __syncthreads();
  if (threadIdx.x < 3) M.tile_base[(tile - M.first_tile) * 3 + threadIdx.x] = s_c[threadIdx.x];
}

// Exclusive scan of the tile counts of one metablock + context mode.  One CTA per metablock.
__global__ void __launch_bounds__(256) k_mb_scan(const MbModel* __restrict__ models, const uint8_t* __restrict__ in) {
  __shared__ uint32_t s_part[256];
  __shared__ uint32_t s_run[3];
  const MbModel& M = models[blockIdx.x];
  const uint32_t tid = threadIdx.x;
  if (tid < 3) s_run[tid] = 0;
  __syncthreads();
  for (int c = 0; c < 3; c++) {
    for (uint32_t t0 = 0; t0 < M.n_tile; t0 += 256) {
      const uint32_t v = (t0 + tid < M.n_tile) ? M.tile_base[(t0 + tid) * 3 + c] : 0;
      s_part[tid] = v;
      __syncthreads();
      for (int o = 1; o < 256; o <<= 1) {
        const uint32_t add = tid >= (uint32_t)o ? s_part[tid - o] : 0;
        __syncthreads();
        s_part[tid] += add;
        __syncthreads();
      }
      const uint32_t incl = s_part[tid];
      if (t0 + tid < M.n_tile) M.tile_base[(t0 + tid) * 3 + c] = s_run[c] + incl - v;
      __syncthreads();
      if (tid == 255) s_run[c] += incl;
      __syncthreads();
    }
  }
  if (tid < 3) M.cnt[tid] = s_run[tid];
  if (tid == 0) {
    // context.ts:179-227 chooseContextMode over the first <= 4096 bytes of the metablock
    const uint8_t* d = in + M.abs;
    const uint32_t length = min(M.len, 4096u);
    uint32_t ascii = 0, lead = 0, cont = 0, signedPat = 0;
    const uint32_t sampleSize = min(length, 1024u);
    const uint32_t step = max(1u, sampleSize ? length / sampleSize : 1u);
    for (uint32_t i = 0; i < length; i += step) {
      const int byte = d[i];
      if (byte < 128) {
        ascii++;
        if (i > 0) { const int diff = abs(byte - (int)d[i - 1]); if (diff < 16) signedPat++; }
      } else if (byte >= 192) lead++;
      else cont++;
    }
    const uint32_t total = ascii + lead + cont;
    uint32_t mode = 0;
    if (total == 0) mode = 0;
    else if (lead > 0 && (float)cont > (float)lead * 0.5f) mode = 2;
    else if ((float)signedPat > (float)total * 0.3f) mode = 3;
    else if ((float)ascii > (float)total * 0.7f) mode = 2;
    M.cnt[3] = mode;
  }
}

// Writes the rank-ordered symbol streams.  grid = tiles, 256 threads; thread t owns positions 4t..4t+3.
__global__ void __launch_bounds__(256) k_mb_compact(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                    const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                    const MbModel* __restrict__ models, const uint8_t* __restrict__ role,
                                                    const uint32_t* __restrict__ ins_at, const Cmd* __restrict__ cmds, Params P) {
  __shared__ uint32_t s_w[8][3];
  const uint32_t tile = blockIdx.x;
  const SegD sg = segs[tile_seg[tile]];
  const ShardD sh = shards[sg.shard];
  const MbModel& M = models[sg.mb];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
  const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
  // the LUT is indexed by data bytes: divergent across a warp, which the constant cache would serialise
  __shared__ uint8_t lut[512];
  for (uint32_t i = tid; i < 512; i += 256) lut[i] = c_ctx_lut[(M.cnt[3] << 9) + i];
  uint32_t r[4], pre[4]; uint32_t nl = 0, nc = 0, nd = 0;
  load_roles4(role, cmds, sh.in_off + tile_rel + 4 * tid, tile_rel + 4 * tid, tile_end, P.sh, r, pre);
#pragma unroll
  for (int k = 0; k < 4; k++) { nl += r[k] & 1; nc += (r[k] >> 2) & 1; nd += cmd_has_dist(r[k], pre[k]) ? 1 : 0; }
  // block exclusive scan of the three counters (packed: 11 bits each is enough for 1024 positions)
  uint32_t packed = nl | (nc << 11) | (nd << 22), incl = packed;
  for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
  if (lane == 31) { s_w[w][0] = incl & 2047; s_w[w][1] = (incl >> 11) & 2047; s_w[w][2] = incl >> 22; }
  __syncthreads();
  uint32_t bl = 0, bc = 0, bd = 0;
  for (uint32_t i = 0; i < w; i++) { bl += s_w[i][0]; bc += s_w[i][1]; bd += s_w[i][2]; }
  const uint32_t ex = incl - packed;
  const uint32_t* tb = M.tile_base + (tile - M.first_tile) * 3;
  uint32_t il = tb[0] + bl + (ex & 2047), ic = tb[1] + bc + ((ex >> 11) & 2047), id = tb[2] + bd + (ex >> 22);
  // By kind, not by position (the lanes of a warp hold different roles at every position; r5m: 19 of 32 lanes active):
  // all literals of the thread from two aligned words, then its command starts, then its copy starts.  The streams are
  // indexed by rank, so the three kinds are independent.
  const uint64_t p0 = sh.in_off + tile_rel + 4 * tid;
  if ((r[0] | r[1] | r[2] | r[3]) & 1) {
    const uint32_t w1 = *reinterpret_cast<const uint32_t*>(in + p0);
    const uint32_t w0 = p0 >= 4 ? *reinterpret_cast<const uint32_t*>(in + p0 - 4) : 0u;
    const uint64_t six = ((uint64_t)w1 << 32) | w0;   // byte j = in[p0 - 4 + j]
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (r[k] & 1) {
        // literal context from the two previous bytes of the stream (context.ts:169-171); (0, 0) at the stream start
        const uint64_t p = p0 + k;
        const bool s1 = M.stream_start && p == M.abs, s2 = M.stream_start && p <= M.abs + 1;
        const uint32_t p1 = s1 ? 0 : (uint32_t)(six >> (8 * (k + 3))) & 0xFF, p2 = s2 ? 0 : (uint32_t)(six >> (8 * (k + 2))) & 0xFF;
        M.clit[il] = (uint8_t)(six >> (8 * (k + 4)));
        M.clctx[il] = lut[p1] | lut[256 + p2];
        il++;
      }
    }
  }
  // the first command start / copy start of the thread is handled outside the position loop, so that the body runs once
  // per thread whatever position it sits at; further ones (q10/11: copies of 2 bytes) take the loop
  if (nc) {
    const int kc = (r[0] & 4) ? 0 : (r[1] & 4) ? 1 : (r[2] & 4) ? 2 : 3;
    const uint32_t rc = kc == 0 ? r[0] : kc == 1 ? r[1] : kc == 2 ? r[2] : r[3];
    const uint64_t p = p0 + kc;
    const uint32_t ins = ins_at[p >> P.sh];
    M.csym[ic++] = (uint16_t)((rc & 8) ? insert_only_prefix(ins) : cmds[(p + ins) >> P.sh].cmdPrefix);
    if (nc > 1) {
#pragma unroll
      for (int k = 1; k < 4; k++) {
        if (k > kc && (r[k] & 4)) {
          const uint64_t q = p0 + k;
          const uint32_t ins2 = ins_at[q >> P.sh];
          M.csym[ic++] = (uint16_t)((r[k] & 8) ? insert_only_prefix(ins2) : cmds[(q + ins2) >> P.sh].cmdPrefix);
        }
      }
    }
  }
  if (nd) {
    const bool h0 = cmd_has_dist(r[0], pre[0]), h1 = cmd_has_dist(r[1], pre[1]), h2 = cmd_has_dist(r[2], pre[2]);
    const int kd = h0 ? 0 : h1 ? 1 : h2 ? 2 : 3;
    {
      const Cmd c = cmds[(p0 + kd) >> P.sh];
      const uint32_t copyLen = c.copyLen & 0x1FFFFFF;
      M.dsym[id] = c.distPrefix & 0x3FF;
      M.ddctx[id] = (uint8_t)(copyLen > 4 ? 3 : copyLen - 2);   // metablock.ts:622
      id++;
    }
    if (nd > 1) {
#pragma unroll
      for (int k = 1; k < 4; k++) {
        if (k > kd && cmd_has_dist(r[k], pre[k])) {
          const Cmd c = cmds[(p0 + k) >> P.sh];
          const uint32_t copyLen = c.copyLen & 0x1FFFFFF;
          M.dsym[id] = c.distPrefix & 0x3FF;
          M.ddctx[id] = (uint8_t)(copyLen > 4 ? 3 : copyLen - 2);
          id++;
        }
      }
    }
  }
}

}  // namespace bgpu
