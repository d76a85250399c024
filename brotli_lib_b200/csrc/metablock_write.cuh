// Part E of the full metablock: the emit loop of storeMetaBlock (metablock.ts:714-756) as a
// position-parallel bit writer.  Tokens of one input position, in stream order:
//   [command block switch] [command symbol + insert/copy extra bits]     if a command starts here
//   [literal block switch] [literal symbol, code chosen by block type and context]   if a literal
//   [distance block switch] [distance symbol + extra bits]               if a copy with an explicit distance starts here
#pragma once
#include "metablock_header.cuh"

namespace bgpu {

struct Tok4 { uint32_t n[4]; uint64_t v[4]; };

// Per-tile view of the metablock's model, in shared memory.  A tile of 1024 positions usually sits
// inside ONE block of each category (blocks hold hundreds to thousands of symbols): then the block
// type is a tile constant, no position emits a block switch, and the chain  rank -> block -> type ->
// context-map row  (three dependent global loads per symbol; profiles/r1k: long-scoreboard stalls
// 8.3 per issue in k_write2) collapses to a shared-memory row look-up.  Tiles that do contain a block
// boundary of a category keep the per-symbol look-ups for that category.  The context LUT row of the
// metablock's mode is staged here too: indexed by data bytes, it diverges across a warp, which the
// constant cache serialises.
struct TileModel {
  uint32_t uni[3], type[3];
  uint32_t cmap_lit[64], cmap_dist[4];
  uint8_t lut[512];
};

__device__ __forceinline__ void switch_token(const MbModel& M, const TileModel& T, int cat, uint32_t rank, uint32_t* type, uint32_t* n, uint64_t* v) {
  *n = 0; *v = 0;
  if (T.uni[cat]) { *type = T.type[cat]; return; }
  const uint32_t b = M.sym_blk[cat][rank];
  *type = M.blk_type[cat][b];
  if (b && M.blk_start[cat][b] == rank) { const uint64_t t = M.blk_tok[cat][b]; *n = (uint32_t)(t >> 56); *v = t & 0x00FFFFFFFFFFFFFFull; }
}

// il / ic / id: rank of this position's literal / command / distance symbol inside the metablock.
__device__ __forceinline__ Tok4 position_tokens2(const uint8_t* __restrict__ in, uint64_t p, uint32_t r, const uint32_t* __restrict__ ins_at,
                                                 const Cmd* __restrict__ cmds, const MbModel& M, uint32_t il, uint32_t ic, uint32_t id,
                                                 int sh, bool trivial, const TileModel& T) {
  Tok4 t;
#pragma unroll
  for (int k = 0; k < 4; k++) { t.n[k] = 0; t.v[k] = 0; }
  if (r & 4) {
    const uint32_t ins = ins_at[p >> sh];
    uint32_t prefix, copyField;
    if (r & 8) { prefix = insert_only_prefix(ins); copyField = 2u << 25; }
    else { const Cmd c = cmds[(p + ins) >> sh]; prefix = c.cmdPrefix; copyField = c.copyLen; }
    uint32_t type;
    switch_token(M, T, CMD, ic, &type, &t.n[0], &t.v[0]);
    const uint32_t e = M.codes[CMD][(size_t)type * ALPHA_CMD + prefix];
    uint64_t xv; const uint32_t xn = cmd_extra(ins, copyField, &xv);
    t.n[1] = (e & 0xFF) + xn;
    t.v[1] = (uint64_t)(e >> 8) | (xv << (e & 0xFF));
  }
  if (r & 1) {
    uint32_t type;
    switch_token(M, T, LIT, il, &type, &t.n[2], &t.v[2]);
    uint32_t ctx = 0;
    if (!trivial) {
      const bool s1 = M.stream_start && p == M.abs, s2 = M.stream_start && p <= M.abs + 1;
      ctx = T.lut[s1 ? 0 : in[p - 1]] | T.lut[256 + (s2 ? 0 : in[p - 2])];
    }
    const uint32_t cl = T.uni[LIT] ? T.cmap_lit[ctx] : M.cmap[LIT][type * 64 + ctx];
    const uint32_t e = M.codes[LIT][cl * 256 + in[p]];
    t.n[3] = e & 0xFF; t.v[3] = e >> 8;
  } else if (r & 2) {
    const Cmd c = cmds[p >> sh];
    if (c.cmdPrefix >= 128) {
      uint32_t type;
      switch_token(M, T, DIST, id, &type, &t.n[2], &t.v[2]);
      const uint32_t copyLen = c.copyLen & 0x1FFFFFF;
      const uint32_t dctx = trivial ? 0 : (copyLen > 4 ? 3 : copyLen - 2);
      const uint32_t cl = T.uni[DIST] ? T.cmap_dist[dctx] : M.cmap[DIST][type * 4 + dctx];
      const uint32_t e = M.codes[DIST][cl * ALPHA_DIST_MAX + (c.distPrefix & 0x3FF)];
      t.n[3] = (e & 0xFF) + (c.distPrefix >> 10);
      t.v[3] = (uint64_t)(e >> 8) | ((uint64_t)c.distExtra << (e & 0xFF));
    }
  }
  return t;
}

// Shared by k_tilebits2 and k_write2: roles of the thread's 4 positions and their symbol ranks.
struct TileRanks { uint32_t r[4]; uint32_t il, ic, id; };
__device__ __forceinline__ TileRanks tile_ranks(const MbModel& M, uint32_t tile, const ShardD& sh, uint32_t tile_rel, uint32_t tile_end,
                                                const uint8_t* __restrict__ role, const Cmd* __restrict__ cmds, int shft,
                                                uint32_t (*s_w)[3]) {
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  TileRanks R; uint32_t nl = 0, nc = 0, nd = 0;
  // the thread's 4 roles are one aligned 32-bit load (shards are 256-byte aligned, tiles start at multiples
  // of 1024); the command records of its copy starts are then fetched together, not one after the other
  const uint32_t rel0 = tile_rel + 4 * tid;
  const uint64_t p0 = sh.in_off + rel0;
  uint32_t r4 = rel0 < tile_end ? *reinterpret_cast<const uint32_t*>(role + p0) : 0u;
  if (rel0 + 4 > tile_end && rel0 < tile_end) r4 &= (1u << (8 * (tile_end - rel0))) - 1u;   // ragged last tile of a segment
  uint32_t pre[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    R.r[k] = (r4 >> (8 * k)) & 0xFFu;
    const bool copy_start = (R.r[k] & 2) && !(R.r[k] & 1);
    pre[k] = copy_start ? cmds[(p0 + k) >> shft].cmdPrefix : 0u;
  }
#pragma unroll
  for (int k = 0; k < 4; k++) {
    if (pre[k] >= 128) { nd++; R.r[k] |= 16; }   // bit 4: codes a distance
    nl += R.r[k] & 1; nc += (R.r[k] >> 2) & 1;
  }
  const uint32_t packed = nl | (nc << 11) | (nd << 22);
  uint32_t incl = packed;
  for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += y; }
  if (lane == 31) { s_w[w][0] = incl & 2047; s_w[w][1] = (incl >> 11) & 2047; s_w[w][2] = incl >> 22; }
  __syncthreads();
  uint32_t bl = 0, bc = 0, bd = 0;
  for (uint32_t i = 0; i < w; i++) { bl += s_w[i][0]; bc += s_w[i][1]; bd += s_w[i][2]; }
  const uint32_t ex = incl - packed;
  const uint32_t* tb = M.tile_base + (tile - M.first_tile) * 3;
  R.il = tb[0] + bl + (ex & 2047); R.ic = tb[1] + bc + ((ex >> 11) & 2047); R.id = tb[2] + bd + (ex >> 22);
  return R;
}

// One word per (tile, category), written by k_tile_model: bit 31 = one block covers the tile's symbols of
// the category and does not start inside the tile; low bits = that block's type.
__global__ void __launch_bounds__(256) k_tile_model(const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                    const MbModel* __restrict__ models, uint32_t n_tile, uint32_t* __restrict__ tile_model) {
  const uint32_t i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n_tile * 3) return;
  const uint32_t tile = i / 3, cat = i % 3;
  const MbModel& M = models[segs[tile_seg[tile]].mb];
  const uint32_t lt = tile - M.first_tile;
  const uint32_t first = M.tile_base[lt * 3 + cat];
  const uint32_t end = lt + 1 < M.n_tile ? M.tile_base[(lt + 1) * 3 + cat] : M.cnt[cat];
  uint32_t uni = 1, type = 0;
  if (end > first) {
    const uint32_t b = M.sym_blk[cat][first];
    const uint32_t* bs = M.blk_start[cat];
    uni = (end <= bs[b + 1] && (b == 0 || bs[b] < first)) ? 1u : 0u;   // block 0 never emits a switch
    type = M.blk_type[cat][b];
  }
  tile_model[i] = (uni << 31) | type;
}

// Fills the tile's TileModel.  No barrier inside: call it BEFORE tile_ranks, whose __syncthreads then
// publishes it, so that these loads travel together with the role / command loads of tile_ranks.
__device__ __forceinline__ void tile_model_load(const MbModel& M, uint32_t tile, const uint32_t* __restrict__ tile_model, bool trivial, TileModel& T) {
  const uint32_t tid = threadIdx.x;
  const uint32_t wl = tile_model[tile * 3 + LIT], wd = tile_model[tile * 3 + DIST];
  if (tid < 3) { const uint32_t w = tile_model[tile * 3 + tid]; T.uni[tid] = w >> 31; T.type[tid] = w & 0xFFFFu; }
  const uint8_t* lut = c_ctx_lut + (M.cnt[3] << 9);
  for (uint32_t i = tid; i < 512; i += 256) T.lut[i] = lut[i];
  if (tid < 64) T.cmap_lit[tid] = ((wl >> 31) && (!trivial || tid == 0)) ? M.cmap[LIT][(wl & 0xFFFFu) * 64 + tid] : 0;
  else if (tid < 68) T.cmap_dist[tid - 64] = ((wd >> 31) && (!trivial || tid == 64)) ? M.cmap[DIST][(wd & 0xFFFFu) * 4 + (tid - 64)] : 0;
}

__global__ void __launch_bounds__(256) k_tilebits2(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                   const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                   const MbModel* __restrict__ models, const uint8_t* __restrict__ role,
                                                   const uint32_t* __restrict__ ins_at, const Cmd* __restrict__ cmds,
                                                   const uint32_t* __restrict__ tile_model, uint32_t* __restrict__ tile_bits,
                                                   uint32_t* __restrict__ warp_base, uint16_t* __restrict__ thr_off, Params P) {
  __shared__ uint32_t s_w[8][3];
  __shared__ uint32_t s_sum[8];
  __shared__ TileModel s_T;
  const uint32_t tile = blockIdx.x;
  const SegD sg = segs[tile_seg[tile]];
  const ShardD sh = shards[sg.shard];
  const MbModel& M = models[sg.mb];
  const bool trivial = M.cnt[4] != 0;
  const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
  const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
  tile_model_load(M, tile, tile_model, trivial, s_T);
  TileRanks R = tile_ranks(M, tile, sh, tile_rel, tile_end, role, cmds, P.sh, s_w);   // contains a __syncthreads
  uint32_t bits = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint32_t r = R.r[k];
    if (!r) continue;
    const uint64_t p = sh.in_off + tile_rel + 4 * threadIdx.x + k;
    const Tok4 t = position_tokens2(in, p, r, ins_at, cmds, M, R.il, R.ic, R.id, P.sh, trivial, s_T);
    bits += t.n[0] + t.n[1] + t.n[2] + t.n[3];
    R.il += r & 1; R.ic += (r >> 2) & 1; R.id += (r >> 4) & 1;
  }
  // bit offset of the thread inside its warp (u16: 128 positions hold < 64 Ki bits) and of the warp inside
  // the tile: k_write2 then knows where every thread writes before it computes a single token
  uint32_t incl = bits;
  for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= (uint32_t)o) incl += y; }
  thr_off[(size_t)tile * 256 + threadIdx.x] = (uint16_t)(incl - bits);
  if ((threadIdx.x & 31) == 31) s_sum[threadIdx.x >> 5] = incl;
  __syncthreads();
  if (threadIdx.x < 8) {
    uint32_t base = 0;
    for (uint32_t i = 0; i < threadIdx.x; i++) base += s_sum[i];
    warp_base[(size_t)tile * 8 + threadIdx.x] = base;
    if (threadIdx.x == 7) tile_bits[tile] = base + s_sum[7];
  }
}

// atomicOr-copy of `nbits` bits from src (bit 0 of word 0 first) to bit position `pos` of w.
__device__ __forceinline__ void copy_bits(uint32_t* w, uint64_t pos, const uint32_t* __restrict__ src, uint32_t nbits, uint32_t tid, uint32_t nthr) {
  const uint32_t nw = (nbits + 31) >> 5;
  for (uint32_t k = tid; k < nw; k += nthr) {
    uint32_t v = src[k];
    const uint32_t rem = nbits - k * 32;
    if (rem < 32) v &= (1u << rem) - 1u;
    if (!v) continue;
    const uint64_t bp = pos + (uint64_t)k * 32;
    const uint64_t i = bp >> 5; const uint32_t o = (uint32_t)bp & 31;
    atomicOr(&w[i], v << o);
    if (o) atomicOr(&w[i + 1], v >> (32 - o));
  }
}

// One CTA per shard: headers + prefix codes of its metablocks, bit offset of every tile, shard length.
__global__ void __launch_bounds__(256) k_offsets2(const ShardD* __restrict__ shards, const MbD* __restrict__ mbs,
                                                  const MbModel* __restrict__ models, const uint32_t* __restrict__ tile_bits,
                                                  uint32_t* __restrict__ tile_off, uint8_t* __restrict__ out,
                                                  uint32_t* __restrict__ shard_bytes) {
  __shared__ uint32_t s_part[256];
  __shared__ uint64_t s_pos;
  __shared__ uint32_t s_toff[HDR_TREE_SLOTS]; __shared__ uint16_t s_tslot[HDR_TREE_SLOTS]; __shared__ uint32_t s_nt, s_ttot;
  const ShardD sh = shards[blockIdx.x];
  uint32_t* w = reinterpret_cast<uint32_t*>(out + sh.out_off);
  const uint32_t tid = threadIdx.x;
  if (tid == 0) {
    uint64_t pos = 0;
    if (sh.flags & 1) gput(w, pos, sh.start_nbits, sh.start_value);
    s_pos = pos;
  }
  __syncthreads();
  for (uint32_t m = sh.first_mb; m < sh.first_mb + sh.n_mb; m++) {
    const MbD mb = mbs[m];
    const MbModel& M = models[m];
    const uint32_t misc_bits = M.cnt[5];
    if (tid == 0) {
      uint32_t n = 0, off = 0;
      const uint32_t ncl_l = M.cnt[6], ntc = M.sstate[CMD][2], ncl_d = M.cnt[7];
      for (uint32_t s = 0; s < ncl_l; s++) { s_tslot[n] = (uint16_t)s; s_toff[n] = off; off += M.tree_nbits[s]; n++; }
      for (uint32_t s = 0; s < ntc; s++) { s_tslot[n] = (uint16_t)(64 + s); s_toff[n] = off; off += M.tree_nbits[64 + s]; n++; }
      for (uint32_t s = 0; s < ncl_d; s++) { s_tslot[n] = (uint16_t)(320 + s); s_toff[n] = off; off += M.tree_nbits[320 + s]; n++; }
      s_nt = n; s_ttot = off;
    }
    __syncthreads();
    const uint64_t base = s_pos;
    copy_bits(w, base, M.misc, misc_bits, tid, 256);
    for (uint32_t t = tid >> 5; t < s_nt; t += 8) {
      const uint32_t slot = s_tslot[t];
      copy_bits(w, base + misc_bits + s_toff[t], M.tree_bits + (size_t)slot * TREE_WORDS, M.tree_nbits[slot], tid & 31, 32);
    }
    __syncthreads();
    if (tid == 0) s_pos = base + misc_bits + s_ttot;
    __syncthreads();
    for (uint32_t t0 = 0; t0 < mb.n_tile; t0 += 256) {
      const uint32_t v = (t0 + tid < mb.n_tile) ? tile_bits[mb.first_tile + t0 + tid] : 0;
      s_part[tid] = v;
      __syncthreads();
      for (int o = 1; o < 256; o <<= 1) {
        const uint32_t add = tid >= (uint32_t)o ? s_part[tid - o] : 0;
        __syncthreads();
        s_part[tid] += add;
        __syncthreads();
      }
      const uint32_t incl = s_part[tid];
      if (t0 + tid < mb.n_tile) tile_off[mb.first_tile + t0 + tid] = (uint32_t)s_pos + incl - v;
      __syncthreads();
      if (tid == 255) s_pos += incl;
      __syncthreads();
    }
  }
  if (tid == 0) {
    uint64_t pos = s_pos;
    if (!(sh.flags & 2)) gput(w, pos, 6, 6);  // empty metadata block (RFC 7932 §9.2): byte-aligns a non-final shard
    shard_bytes[blockIdx.x] = (uint32_t)((pos + 7) >> 3);
  }
}

// Round-1 token writer, now the fall-back of k_tile_pack / k_tile_copy: emits the tiles listed in fb_list[0 .. *fb_count)
// (tiles whose bits did not fit the packed form) by recomputing their tokens.  A fixed grid walks the list.
__global__ void __launch_bounds__(256) k_write2(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                const MbModel* __restrict__ models, const uint8_t* __restrict__ role,
                                                const uint32_t* __restrict__ ins_at, const Cmd* __restrict__ cmds,
                                                const uint32_t* __restrict__ tile_model, const uint32_t* __restrict__ tile_bits,
                                                const uint32_t* __restrict__ tile_off, const uint32_t* __restrict__ warp_base,
                                                const uint16_t* __restrict__ thr_off, uint8_t* __restrict__ out,
                                                const uint32_t* __restrict__ fb_list, const uint32_t* __restrict__ fb_count, Params P) {
  __shared__ uint32_t s_stage[STAGE_WORDS];
  __shared__ uint32_t s_w[8][3];
  __shared__ TileModel s_T;
  const uint32_t n_fb = *fb_count;
  for (uint32_t idx = blockIdx.x; idx < n_fb; idx += gridDim.x) {
    const uint32_t tile = fb_list[idx];
    const uint32_t total = tile_bits[tile];
    if (total == 0) continue;
    const SegD sg = segs[tile_seg[tile]];
    const ShardD sh = shards[sg.shard];
    const MbModel& M = models[sg.mb];
    const bool trivial = M.cnt[4] != 0;
    const uint32_t tid = threadIdx.x;
    const uint32_t off = tile_off[tile];
    const uint32_t head = off & 31;
    const uint32_t nwords = (head + total + 31) >> 5;
    const bool staged = nwords + 2 <= STAGE_WORDS;
    __syncthreads();   // the previous tile of this CTA is done with the shared arrays
    if (staged) for (uint32_t i = tid; i < nwords + 2; i += 256) s_stage[i] = 0;
    const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
    const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
    tile_model_load(M, tile, tile_model, trivial, s_T);
    TileRanks R = tile_ranks(M, tile, sh, tile_rel, tile_end, role, cmds, P.sh, s_w);   // contains a __syncthreads
    uint32_t bitpos = head + warp_base[(size_t)tile * 8 + (tid >> 5)] + thr_off[(size_t)tile * 256 + tid];
    uint32_t* gw = reinterpret_cast<uint32_t*>(out + sh.out_off) + (off >> 5);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const uint32_t r = R.r[k];
      if (!r) continue;
      const uint64_t p = sh.in_off + tile_rel + 4 * tid + k;
      const Tok4 t = position_tokens2(in, p, r, ins_at, cmds, M, R.il, R.ic, R.id, P.sh, trivial, s_T);
      R.il += r & 1; R.ic += (r >> 2) & 1; R.id += (r >> 4) & 1;
#pragma unroll
      for (int f = 0; f < 4; f++) {
        if (staged) sput(s_stage, bitpos, t.n[f], t.v[f]); else sput(gw, bitpos, t.n[f], t.v[f]);
        bitpos += t.n[f];
      }
    }
    if (!staged) continue;
    __syncthreads();
    for (uint32_t i = tid; i < nwords; i += 256) {
      const uint32_t v = s_stage[i];
      if (i == 0 || i == nwords - 1) { if (v) atomicOr(&gw[i], v); }
      else gw[i] = v;
    }
  }
}

// ---- One-pass form of the bit writer: the tokens of a tile are computed ONCE.
//   k_tile_pack   k_tilebits2 + the token half of k_write2: every thread appends the tokens of its 4 positions to a
//                 private 256-bit slot in shared memory, a CTA scan of the bit counts gives each thread its offset in
//                 the tile, the slots are OR-ed into the tile's bit string (shared memory) and the string goes to
//                 tile_stage[tile] (PACK_WORDS words per tile) in coalesced stores.
//   k_offsets2    unchanged (bit offset of every tile inside its shard).
//   k_tile_copy   shifts each tile's string to its final bit offset: reads M bytes, writes M bytes; `in`, `role`,
//                 `ins`, `cmds`, the codes and the block model are not touched again (k_write2 re-read all of them:
//                 15.2 GB of DRAM traffic per GiB against 1.56 GB algorithmic, profiles/traffic.json).
// A tile that does not fit (a thread with > 256 bits, or > PACK_WORDS words in the tile: incompressible data with long
// codes) takes the round-1 path: k_tile_pack leaves thr_off / warp_base for it and appends it to fb_list, which
// k_write2 (a fixed grid) walks.
constexpr uint32_t PACK_WORDS = 640;          // 20 bits per position of a 1024-position tile
constexpr uint32_t PACK_SLOT_WORDS = 8;       // private bits per thread (4 positions)

__device__ __forceinline__ void pput(uint32_t* slot, uint32_t pos, uint32_t nbits, uint64_t value) {   // slot: word w at slot[w * 256]
  if (!nbits) return;
  const uint32_t i = pos >> 5, o = pos & 31;
  const unsigned long long lo = (unsigned long long)value << o;
  slot[i * 256] |= (uint32_t)lo;
  if (o + nbits > 32) slot[(i + 1) * 256] |= (uint32_t)(lo >> 32);
  if (o + nbits > 64) slot[(i + 2) * 256] |= (uint32_t)(value >> (64 - o));
}

__global__ void __launch_bounds__(256) k_tile_pack(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                   const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                   const MbModel* __restrict__ models, const uint8_t* __restrict__ role,
                                                   const uint32_t* __restrict__ ins_at, const Cmd* __restrict__ cmds,
                                                   const uint32_t* __restrict__ tile_model, uint32_t* __restrict__ tile_bits,
                                                   uint32_t* __restrict__ warp_base, uint16_t* __restrict__ thr_off,
                                                   uint32_t* __restrict__ tile_stage, uint8_t* __restrict__ tile_flag,
                                                   uint32_t* __restrict__ fb_list, uint32_t* __restrict__ fb_count, Params P) {
  __shared__ uint32_t s_w[8][3];
  __shared__ uint32_t s_sum[8];
  __shared__ uint32_t s_over;
  __shared__ TileModel s_T;
  __shared__ uint32_t s_slot[(PACK_SLOT_WORDS + 3) * 256];   // + 3: a put that overflows the slot stays inside the array
  __shared__ uint32_t s_str[PACK_WORDS + 2];
  const uint32_t tile = blockIdx.x, tid = threadIdx.x;
  const SegD sg = segs[tile_seg[tile]];
  const ShardD sh = shards[sg.shard];
  const MbModel& M = models[sg.mb];
  const bool trivial = M.cnt[4] != 0;
  const uint32_t tile_rel = sg.rel_pos + (tile - sg.first_tile) * TILE;
  const uint32_t tile_end = min(sg.rel_pos + sg.len, tile_rel + TILE);
#pragma unroll
  for (uint32_t w = 0; w < PACK_SLOT_WORDS + 3; w++) s_slot[w * 256 + tid] = 0;
  for (uint32_t i = tid; i < PACK_WORDS + 2; i += 256) s_str[i] = 0;
  if (tid == 0) s_over = 0;
  tile_model_load(M, tile, tile_model, trivial, s_T);
  TileRanks R = tile_ranks(M, tile, sh, tile_rel, tile_end, role, cmds, P.sh, s_w);   // contains a __syncthreads
  uint32_t bits = 0;
  uint32_t* slot = s_slot + tid;
  // Dense path.  In the positional loop below the lanes of a warp take different branches at every step (literal /
  // command start / copy start / covered by a copy): r5m measured 14 of 32 lanes active and 11.6 warp instructions per
  // position.  When one block of each category covers the tile (no block-switch tokens, block types and context-map rows
  // are tile constants) and no thread holds more than one command start and one distance (copies are >= 4 bytes at
  // q1-9, so 4 aligned positions cannot hold two), the tokens are computed BY KIND instead: all literals of the thread,
  // then its command, then its distance — each branch is taken once per thread, not once per position — and only then
  // appended in position order.
  uint32_t ncs = 0, nds = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) { ncs += (R.r[k] >> 2) & 1; nds += (R.r[k] >> 4) & 1; }
  const bool dense = __syncthreads_or((ncs > 1 || nds > 1) ? 1 : 0) == 0 && s_T.uni[LIT] && s_T.uni[CMD] && s_T.uni[DIST];
  if (dense) {
    const uint64_t p0 = sh.in_off + tile_rel + 4 * tid;
    uint32_t lt[4] = {0, 0, 0, 0};   // literal tokens: code bits << 8 | length (the layout of M.codes)
    const uint32_t any_lit = (R.r[0] | R.r[1] | R.r[2] | R.r[3]) & 1;
    if (any_lit) {
      // the 4 bytes of the thread and the 4 before them as two aligned words (contexts need the two previous bytes)
      const uint32_t w1 = *reinterpret_cast<const uint32_t*>(in + p0);
      const uint32_t w0 = p0 >= 4 ? *reinterpret_cast<const uint32_t*>(in + p0 - 4) : 0u;
      const uint64_t six = ((uint64_t)w1 << 32) | w0;   // byte j of `six` = in[p0 - 4 + j]
#pragma unroll
      for (int k = 0; k < 4; k++) {
        if (R.r[k] & 1) {
          const uint64_t p = p0 + k;
          uint32_t ctx = 0;
          if (!trivial) {
            const bool s1 = M.stream_start && p == M.abs, s2 = M.stream_start && p <= M.abs + 1;
            const uint32_t b1 = (uint32_t)(six >> (8 * (k + 3))) & 0xFF, b2 = (uint32_t)(six >> (8 * (k + 2))) & 0xFF;
            ctx = s_T.lut[s1 ? 0 : b1] | s_T.lut[256 + (s2 ? 0 : b2)];
          }
          lt[k] = M.codes[LIT][s_T.cmap_lit[ctx] * 256 + ((uint32_t)(six >> (8 * (k + 4))) & 0xFF)];
        }
      }
    }
    int kc = -1, kd = -1; uint32_t cn = 0, dn = 0; uint64_t cv = 0, dv = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) { if (R.r[k] & 4) kc = k; if (R.r[k] & 16) kd = k; }
    if (kc >= 0) {
      const uint64_t p = p0 + kc;
      const uint32_t r = R.r[0] * (kc == 0) + R.r[1] * (kc == 1) + R.r[2] * (kc == 2) + R.r[3] * (kc == 3);
      const uint32_t ins = ins_at[p >> P.sh];
      uint32_t prefix, copyField;
      if (r & 8) { prefix = insert_only_prefix(ins); copyField = 2u << 25; }
      else { const Cmd c = cmds[(p + ins) >> P.sh]; prefix = c.cmdPrefix; copyField = c.copyLen; }
      const uint32_t e = M.codes[CMD][(size_t)s_T.type[CMD] * ALPHA_CMD + prefix];
      uint64_t xv; const uint32_t xn = cmd_extra(ins, copyField, &xv);
      cn = (e & 0xFF) + xn; cv = (uint64_t)(e >> 8) | (xv << (e & 0xFF));
    }
    if (kd >= 0) {
      const Cmd c = cmds[(p0 + kd) >> P.sh];
      const uint32_t copyLen = c.copyLen & 0x1FFFFFF;
      const uint32_t dctx = trivial ? 0 : (copyLen > 4 ? 3 : copyLen - 2);
      const uint32_t e = M.codes[DIST][s_T.cmap_dist[dctx] * ALPHA_DIST_MAX + (c.distPrefix & 0x3FF)];
      dn = (e & 0xFF) + (c.distPrefix >> 10); dv = (uint64_t)(e >> 8) | ((uint64_t)c.distExtra << (e & 0xFF));
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (k == kc) { if (bits + cn <= PACK_SLOT_WORDS * 32) pput(slot, bits, cn, cv); bits += cn; }
      if (R.r[k] & 1) { const uint32_t n = lt[k] & 0xFF; if (bits + n <= PACK_SLOT_WORDS * 32) pput(slot, bits, n, lt[k] >> 8); bits += n; }
      else if (k == kd) { if (bits + dn <= PACK_SLOT_WORDS * 32) pput(slot, bits, dn, dv); bits += dn; }
    }
  } else {
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint32_t r = R.r[k];
    if (!r) continue;
    const uint64_t p = sh.in_off + tile_rel + 4 * tid + k;
    const Tok4 t = position_tokens2(in, p, r, ins_at, cmds, M, R.il, R.ic, R.id, P.sh, trivial, s_T);
    R.il += r & 1; R.ic += (r >> 2) & 1; R.id += (r >> 4) & 1;
#pragma unroll
    for (int f = 0; f < 4; f++) {
      if (bits + t.n[f] <= PACK_SLOT_WORDS * 32) pput(slot, bits, t.n[f], t.v[f]);
      bits += t.n[f];
    }
  }
  }
  if (bits > PACK_SLOT_WORDS * 32) s_over = 1;   // benign race: every writer stores 1
  uint32_t incl = bits;
  for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o); if ((tid & 31) >= (uint32_t)o) incl += y; }
  if ((tid & 31) == 31) s_sum[tid >> 5] = incl;
  __syncthreads();
  uint32_t wb = 0, total = 0;
  for (uint32_t i = 0; i < 8; i++) { if (i < (tid >> 5)) wb += s_sum[i]; total += s_sum[i]; }
  const bool packed = !s_over && total <= PACK_WORDS * 32;
  if (tid == 0) { tile_bits[tile] = total; tile_flag[tile] = packed ? 1 : 0; if (!packed && total) fb_list[atomicAdd(fb_count, 1u)] = tile; }
  if (!packed) {   // round-1 path: k_write2 recomputes the tokens of this tile
    thr_off[(size_t)tile * 256 + tid] = (uint16_t)(incl - bits);
    if ((tid & 31) == 0) warp_base[(size_t)tile * 8 + (tid >> 5)] = wb;
    return;
  }
  const uint32_t toff = wb + incl - bits;
  for (uint32_t w = 0; w * 32 < bits; w++) {
    const uint32_t v = slot[w * 256];
    if (!v) continue;
    const uint32_t d = toff + w * 32, i = d >> 5, o = d & 31;
    atomicOr(&s_str[i], v << o);
    if (o) atomicOr(&s_str[i + 1], v >> (32 - o));
  }
  __syncthreads();
  uint32_t* dst = tile_stage + (size_t)tile * PACK_WORDS;
  for (uint32_t i = tid; i * 32 < total; i += 256) dst[i] = s_str[i];
}

__global__ void __launch_bounds__(256) k_fb_all(uint32_t* __restrict__ fb_list, uint32_t* __restrict__ fb_count, uint32_t n_tile) {
  const uint32_t i = blockIdx.x * 256 + threadIdx.x;
  if (i < n_tile) fb_list[i] = i;
  if (i == 0) *fb_count = n_tile;
}

// One warp per tile: the tile's bit string moves to bit offset tile_off[tile] of its shard's output region
// (zeroed; neighbouring tiles share the first and last word, hence the OR there).
__global__ void __launch_bounds__(256) k_tile_copy(const ShardD* __restrict__ shards, const SegD* __restrict__ segs, const uint32_t* __restrict__ tile_seg,
                                                   const uint32_t* __restrict__ tile_bits, const uint32_t* __restrict__ tile_off,
                                                   const uint32_t* __restrict__ tile_stage, const uint8_t* __restrict__ tile_flag,
                                                   uint8_t* __restrict__ out, uint32_t n_tile) {
  const uint32_t tile = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (tile >= n_tile || !tile_flag[tile]) return;
  const uint32_t total = tile_bits[tile];
  if (total == 0) return;
  const ShardD sh = shards[segs[tile_seg[tile]].shard];
  const uint32_t off = tile_off[tile], head = off & 31;
  const uint32_t nsrc = (total + 31) >> 5, nwords = (head + total + 31) >> 5;
  const uint32_t* src = tile_stage + (size_t)tile * PACK_WORDS;
  uint32_t* gw = reinterpret_cast<uint32_t*>(out + sh.out_off) + (off >> 5);
  const uint32_t tail_mask = (total & 31) ? (1u << (total & 31)) - 1u : 0xFFFFFFFFu;
  for (uint32_t i = lane; i < nwords; i += 32) {
    uint32_t cur = i < nsrc ? src[i] : 0u, prev = (i > 0 && head) ? src[i - 1] : 0u;
    if (i == nsrc - 1) cur &= tail_mask;
    if (i > 0 && i - 1 == nsrc - 1) prev &= tail_mask;
    const uint32_t v = head ? (cur << head) | (prev >> (32 - head)) : cur;
    if (i == 0 || i == nwords - 1) { if (v) atomicOr(&gw[i], v); }
    else gw[i] = v;
  }
}

}  // namespace bgpu
