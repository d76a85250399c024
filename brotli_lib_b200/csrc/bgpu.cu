// libbrotligpu.so — C ABI (include/brotligpu.h) and host orchestration of the sm_100a kernels.
// The host side mirrors src/encode/encode.ts:50-287 of the reference: option clamping, the
// empty/uncompressed special cases (host only), shard splitting and stitching.
#include "../../include/brotligpu.h"
#include "metablock_write.cuh"
#include "h10.cuh"
#include "zopfli.cuh"
#include "decode.cuh"
#include "woff2.cuh"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

using namespace bgpu;

static thread_local std::string g_cuda_err;

#define CK(call)                                                                          \
  do {                                                                                    \
    cudaError_t e_ = (call);                                                              \
    if (e_ != cudaSuccess) {                                                              \
      char b_[512];                                                                       \
      snprintf(b_, sizeof b_, "%s at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      g_cuda_err = b_;                                                                    \
      return e_ == cudaErrorMemoryAllocation ? BGPU_E_NOMEM : BGPU_E_CUDA;                \
    }                                                                                     \
  } while (0)

namespace {

struct DBuf {
  void* p = nullptr; size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = n + n / 8 + 4096;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) { g_cuda_err = std::string("cudaMalloc: ") + cudaGetErrorString(e); p = nullptr; return BGPU_E_NOMEM; }
    cap = want;
    return 0;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct EncParams { int quality, lgwin, mode; Params k; };

// encode.ts:57-71, enc-constants.ts:111-127, hash-chains.ts:240-261
EncParams make_params(int quality, int lgwin, int mode) {
  EncParams p;
  p.quality = std::max(0, std::min(11, quality));
  p.lgwin = std::max(10, std::min(24, lgwin));
  p.mode = mode;
  p.k.npostfix = 0; p.k.ndirect = 0;
  if (p.quality >= 4 && mode == BGPU_MODE_FONT) { p.k.npostfix = 1; p.k.ndirect = 12; }
  p.k.dist_alphabet = 16 + p.k.ndirect + (48u << p.k.npostfix);   // encode.ts:283-287
  int q = std::max(5, std::min(9, p.quality));
  p.k.bucket_bits = q < 7 ? 14 : 15;
  p.k.depth = 1 << (q - 1);
  p.k.num_last = q < 7 ? 4 : (q < 9 ? 10 : 16);
  if (p.quality >= 1 && p.quality <= 4) {
    // SimpleHasher (hash-simple.ts:10-11, 134-150): one slot per key, 16 bits at quality 2, else 17; only
    // distanceCache[0] is probed (:79-97)
    p.k.bucket_bits = p.quality == 2 ? 16 : 17; p.k.depth = 1; p.k.num_last = 1;
  }
  p.k.sh = p.quality >= 10 ? 1 : 2;   // Zopfli emits 2- and 3-byte copies
  p.k.hint = 0;
  return p;
}

void window_bits(int lgwin, uint32_t* nbits, uint32_t* value) {  // bit-writer.ts:172-194
  if (lgwin == 16) { *value = 0; *nbits = 1; }
  else if (lgwin == 17) { *value = 1; *nbits = 7; }
  else if (lgwin > 17 && lgwin <= 24) { *value = ((lgwin - 17) << 1) | 1; *nbits = 4; }
  else { *value = ((lgwin - 8) << 4) | 1; *nbits = 7; }
}

struct StreamPlan { size_t in_off = 0, len = 0; uint32_t first_shard = 0, n_shard = 0; };

struct Job {
  std::vector<ShardD> shards; std::vector<MbD> mbs; std::vector<SegD> segs; std::vector<uint32_t> tile_seg;
  std::vector<StreamPlan> streams;
  size_t in_bytes = 0, out_bytes = 0;
  uint32_t n_stile = 0, max_shard_len = 0;
};

constexpr size_t IN_PAD = 1024;
#ifndef PARSE_G_BUILD
#define PARSE_G_BUILD 8
#endif
constexpr int PARSE_G = PARSE_G_BUILD;   // lanes per parse segment (kernels.cuh k_parse)

}  // namespace

struct bgpu_ctx {
  int device = 0;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  bgpu_tuning tune{0, 0, 0, 0, 0, 0};
  // lanes (lanes_host.inc): private sub-contexts that encode groups of shards concurrently
  std::vector<bgpu_ctx*> lanes; bool is_lane = false;
  cudaEvent_t lane_ev = nullptr, fork_ev = nullptr, h2d_ev = nullptr;
  cudaStream_t copy_stream = nullptr;   // parent only: the groups' host->device copies, in group order
  cudaStream_t hi_stream = nullptr; cudaEvent_t hi_ev = nullptr;   // lanes: high-priority side stream of the modelling phase (model_host.inc side_begin)
  uint32_t* h_sizes = nullptr; size_t h_sizes_cap = 0;   // pinned: shard sizes of the lane's current group
  uint8_t* pin = nullptr; size_t pin_cap = 0, pin_used = 0, pin_want = 0;   // pinned staging of descriptor uploads
  DBuf in, sa_pos, sa_v, sb_pos, sb_v, sort_hist, role, ins, cmds, out, shards, mbs, segs, tile_seg, seg_ncmd, seg_trail, seg_ring, hist, codes, tree_bits, tree_nbits,
      tile_bits, tile_off, tile_model, warp_base, thr_off, shard_bytes, huff_meta, model_arena, misc_arena, d_models, d_splits, d_cb, d_cx, d_bh_ptrs, d_bh_strides, d_cx_outs,
      h10_bstart, h10_bend, h10_short, h10_forest, h10_m, h10_off, h10_cnt, h10_cursor, h10_full, h10_hot,
      zop_bytehist, zop_model, zop_lc, zop_nodes, zop_ring, stats, dec_arena, dec_results, dec_jobs, seg_aux, tile_stage, tile_flag, fb_count;
  uint32_t h10_cap = 0;
  uint64_t counters[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // [0] H10 match entries; [1..3] model table bytes (BGPU_F_MODEL_STATS); [4] shards that failed BGPU_F_VERIFY
  bool lut_ready = false;
  std::vector<uint32_t> h_shard_bytes;
  // multi-device (multi_host.inc): one private context per device of bgpu_ctx_create_multi; parts of one input run on
  // them from one host thread each and are stitched into the caller's buffer
  std::vector<bgpu_ctx*> peers; bool is_peer = false;
  uint8_t* part_out = nullptr; size_t part_out_cap = 0;   // peer only: pinned staging of its compressed part
  // timings
  static const int kMaxStages = 40;
  cudaEvent_t ev[kMaxStages + 1]; const char* names[kMaxStages]; float ms[kMaxStages]; int n_stage = 0; int launches = 0;
  bool ev_ok = false;
};

namespace {

uint32_t eff(uint32_t v, uint32_t dflt) { return v ? v : dflt; }

// Splits every stream into shards / metablocks / parse segments / tiles.
// part_flags[s] (optional): BGPU_PART_FIRST | BGPU_PART_LAST of stream s; a part that is not FIRST carries no
// window header and starts with an unknown distance ring, a part that is not LAST ends with the
// byte-aligning empty metadata block instead of ISLAST.  `lead` bytes are reserved before the first
// stream (the two bytes that precede a non-first part live there: literal context, §8e rule 4).
// `hist` (streaming; one stream of one shard): the first hist bytes of the shard are the already coded window — they
// take part in match finding, metablocks / segments / tiles start behind them.  hist must be a multiple of 4.
void plan_job(const bgpu_ctx* ctx, const EncParams& P, const size_t* lens, int count, Job& job,
              const uint32_t* part_flags = nullptr, size_t lead = 0, size_t hist = 0) {
  uint32_t shard_bytes = std::min<uint32_t>(eff(ctx->tune.shard_bytes, 1u << 22), 1u << 22);
  if (P.lgwin < 22) shard_bytes = std::min<uint32_t>(shard_bytes, ((1u << P.lgwin) - 16) & ~3u);
  shard_bytes = std::max<uint32_t>(shard_bytes & ~3u, 64);
  // q5-9: 8 KiB measured best (parse 15.5 ms per GiB vs 17.4 at 16 KiB, ratio +0.0001; gpurun r2l).  q10/11: a Zopfli segment
  // is one warp's serial DP, so its size is the latency of a small job and irrelevant once ~8 K segments fill the machine:
  // 32 KiB for large jobs, down to 4 KiB for small ones (enc-ttf, 312 KB, one call: 340 ms / 138152 bytes at 32 KiB, 100 ms / 137689 at
  // 4 KiB, 75 ms / 137701 at 2 KiB; 2 KiB costs ~0.2 % on 5-20 KB font blobs, r4l/r4m)
  uint32_t seg_dflt = 1u << 13;
  if (P.quality >= 10) {
    size_t total = 0;
    for (int s = 0; s < count; s++) total += lens[s];
    seg_dflt = 1u << 15;
    while (seg_dflt > 4096 && total / seg_dflt < 8192) seg_dflt >>= 1;
  }
  uint32_t seg_bytes = eff(ctx->tune.segment_bytes, seg_dflt);
  seg_bytes = seg_bytes >= shard_bytes ? shard_bytes : std::max<uint32_t>(seg_bytes & ~(TILE - 1), TILE);
  uint32_t mb_bytes = eff(ctx->tune.metablock_bytes, shard_bytes);
  mb_bytes = mb_bytes >= shard_bytes ? shard_bytes : std::max<uint32_t>((mb_bytes / seg_bytes) * seg_bytes, seg_bytes);
  uint32_t wb_n, wb_v; window_bits(P.lgwin, &wb_n, &wb_v);
  size_t in_off = lead, out_off = 0;
  for (int s = 0; s < count; s++) {
    const bool part_first = !part_flags || (part_flags[s] & BGPU_PART_FIRST), part_last = !part_flags || (part_flags[s] & BGPU_PART_LAST);
    StreamPlan sp; sp.in_off = in_off; sp.len = lens[s]; sp.first_shard = (uint32_t)job.shards.size();
    size_t n = lens[s];
    for (size_t so = 0; so < n; so += shard_bytes) {
      ShardD sh; memset(&sh, 0, sizeof sh);
      sh.in_off = in_off + so; sh.len = (uint32_t)std::min<size_t>(shard_bytes, n - so);
      sh.flags = ((so == 0 && part_first) ? 1u : 0u) | ((so + shard_bytes >= n && part_last) ? 2u : 0u);
      sh.start_nbits = wb_n; sh.start_value = wb_v;
      sh.first_mb = (uint32_t)job.mbs.size();
      sh.first_stile = job.n_stile;
      job.n_stile += sh.len >= 4 ? (sh.len - 3 + SORT_TILE - 1) / SORT_TILE : 0;
      job.max_shard_len = std::max(job.max_shard_len, sh.len);
      uint32_t shard_idx = (uint32_t)job.shards.size();
      sh.hist = so == 0 ? (uint32_t)std::min<size_t>(hist, sh.len) : 0;
      for (uint32_t mo = sh.hist; mo < sh.len; mo += mb_bytes) {
        MbD mb; memset(&mb, 0, sizeof mb);
        mb.shard = shard_idx; mb.rel_pos = mo; mb.len = std::min(mb_bytes, sh.len - mo);
        mb.first_seg = (uint32_t)job.segs.size(); mb.first_tile = (uint32_t)job.tile_seg.size();
        mb.is_last = ((sh.flags & 2) && mo + mb_bytes >= sh.len) ? 1 : 0;
        uint32_t mb_idx = (uint32_t)job.mbs.size();
        for (uint32_t go = 0; go < mb.len; go += seg_bytes) {
          SegD sg; memset(&sg, 0, sizeof sg);
          sg.shard = shard_idx; sg.rel_pos = mo + go; sg.len = std::min(seg_bytes, mb.len - go); sg.mb = mb_idx;
          sg.first_tile = (uint32_t)job.tile_seg.size();
          sg.ring_known = (so == 0 && mo == 0 && go == 0 && part_first) ? 1 : 0;   // encode.ts:202 {4,11,15,16} only at stream start
          uint32_t seg_idx = (uint32_t)job.segs.size();
          for (uint32_t to = 0; to < sg.len; to += TILE) job.tile_seg.push_back(seg_idx);
          job.segs.push_back(sg);
        }
        mb.n_seg = (uint32_t)job.segs.size() - mb.first_seg; mb.n_tile = (uint32_t)job.tile_seg.size() - mb.first_tile;
        job.mbs.push_back(mb);
      }
      sh.n_mb = (uint32_t)job.mbs.size() - sh.first_mb;
      sh.out_off = out_off;
      sh.out_cap = (uint32_t)(2 * (size_t)sh.len + 4096 + (size_t)sh.n_mb * 2048);
      sh.out_cap = (sh.out_cap + 255) & ~255u;
      out_off += sh.out_cap;
      job.shards.push_back(sh);
    }
    sp.n_shard = (uint32_t)job.shards.size() - sp.first_shard;
    job.streams.push_back(sp);
    in_off += (n + 255) & ~(size_t)255;
  }
  job.in_bytes = in_off; job.out_bytes = out_off;
}

// Descriptor upload through the context's pinned staging arena: a cudaMemcpyAsync from pageable memory
// blocks the host until everything queued before it on the stream has run, which would stop the host
// from feeding the other lanes (lanes_host.inc).  The arena is rewound by timing_begin (start of a job,
// after the previous job's final synchronisation); when it is too small the copy falls back to the
// pageable path once and the arena grows at the next rewind.
int upload(bgpu_ctx* ctx, DBuf& b, const void* src, size_t bytes) {
  int rc = b.reserve(bytes ? bytes : 16);
  if (rc) return rc;
  if (!bytes) return 0;
  const size_t need = (bytes + 255) & ~(size_t)255;
  ctx->pin_want += need;
  if (ctx->pin && ctx->pin_used + need <= ctx->pin_cap) {
    uint8_t* stage = ctx->pin + ctx->pin_used; ctx->pin_used += need;
    memcpy(stage, src, bytes);
    CK(cudaMemcpyAsync(b.p, stage, bytes, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    CK(cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  }
  return 0;
}
void pin_rewind(bgpu_ctx* ctx) {
  if (ctx->pin_want > ctx->pin_cap) {
    if (ctx->pin) cudaFreeHost(ctx->pin);
    ctx->pin = nullptr; ctx->pin_cap = 0;
    const size_t want = ctx->pin_want + ctx->pin_want / 4 + (1u << 16);
    if (cudaMallocHost(&ctx->pin, want) == cudaSuccess) ctx->pin_cap = want; else { ctx->pin = nullptr; cudaGetLastError(); }
  }
  ctx->pin_used = 0; ctx->pin_want = 0;
}

void stage_mark(bgpu_ctx* ctx, const char* name) {
  if (!ctx->ev_ok || ctx->n_stage >= bgpu_ctx::kMaxStages) return;
  ctx->names[ctx->n_stage] = name;
  cudaEventRecord(ctx->ev[ctx->n_stage + 1], ctx->stream);
  ctx->n_stage++;
}

// Entropy back end (K7 -> K10 -> K11) over roles/commands already on the device.
int run_backend(bgpu_ctx* ctx, const Job& job, const EncParams& P, const uint8_t* d_in) {
  cudaStream_t st = ctx->stream;
  const uint32_t n_tile = (uint32_t)job.tile_seg.size(), n_mb = (uint32_t)job.mbs.size(), n_shard = (uint32_t)job.shards.size();
  int rc;
  if ((rc = ctx->hist.reserve((size_t)n_mb * sizeof(MbHist)))) return rc;
  if ((rc = ctx->codes.reserve((size_t)n_mb * CODE_STRIDE * 4))) return rc;
  if ((rc = ctx->tree_bits.reserve((size_t)n_mb * 3 * TREE_WORDS * 4))) return rc;
  if ((rc = ctx->tree_nbits.reserve((size_t)n_mb * 3 * 4))) return rc;
  if ((rc = ctx->tile_bits.reserve((size_t)n_tile * 4))) return rc;
  if ((rc = ctx->tile_off.reserve((size_t)n_tile * 4))) return rc;
  if ((rc = ctx->shard_bytes.reserve((size_t)n_shard * 4))) return rc;
  if ((rc = ctx->out.reserve(job.out_bytes + 64))) return rc;
  int meta[6] = {ALPHA_LIT, ALPHA_CMD, (int)P.k.dist_alphabet, 0, ALPHA_LIT, ALPHA_LIT + ALPHA_CMD};
  if ((rc = upload(ctx, ctx->huff_meta, meta, sizeof meta))) return rc;
  CK(cudaMemsetAsync(ctx->hist.p, 0, (size_t)n_mb * sizeof(MbHist), st));
  CK(cudaMemsetAsync(ctx->out.p, 0, job.out_bytes + 64, st));
  const ShardD* d_sh = ctx->shards.as<ShardD>(); const MbD* d_mb = ctx->mbs.as<MbD>(); const SegD* d_sg = ctx->segs.as<SegD>();
  const uint32_t* d_ts = ctx->tile_seg.as<uint32_t>();
  k_hist<<<(n_tile + 7) / 8, 256, 0, st>>>(d_in, d_sh, d_sg, d_ts, n_tile, ctx->role.as<uint8_t>(), ctx->ins.as<uint32_t>(),
                                           ctx->cmds.as<Cmd>(), ctx->hist.as<MbHist>(), P.k);
  ctx->launches++; stage_mark(ctx, "k_hist");
  k_huff<<<n_mb * 3, 128, 0, st>>>(ctx->hist.as<uint32_t>(), CODE_STRIDE, 3, ctx->huff_meta.as<int>(), ctx->huff_meta.as<int>() + 3,
                                   ctx->codes.as<uint32_t>(), ctx->tree_bits.as<uint32_t>(), ctx->tree_nbits.as<uint32_t>());
  ctx->launches++; stage_mark(ctx, "k_huff");
  k_tilebits<<<n_tile, 256, 0, st>>>(d_in, d_sh, d_sg, d_ts, ctx->role.as<uint8_t>(), ctx->ins.as<uint32_t>(), ctx->cmds.as<Cmd>(),
                                     ctx->codes.as<uint32_t>(), ctx->tile_bits.as<uint32_t>(), P.k);
  ctx->launches++; stage_mark(ctx, "k_tilebits");
  k_offsets<<<n_shard, 256, 0, st>>>(d_sh, d_mb, ctx->tile_bits.as<uint32_t>(), ctx->tile_off.as<uint32_t>(), ctx->tree_bits.as<uint32_t>(),
                                     ctx->tree_nbits.as<uint32_t>(), ctx->out.as<uint8_t>(), ctx->shard_bytes.as<uint32_t>(), P.k);
  ctx->launches++; stage_mark(ctx, "k_offsets");
  k_write<<<n_tile, 256, 0, st>>>(d_in, d_sh, d_sg, d_ts, ctx->role.as<uint8_t>(), ctx->ins.as<uint32_t>(), ctx->cmds.as<Cmd>(),
                                  ctx->codes.as<uint32_t>(), ctx->tile_bits.as<uint32_t>(), ctx->tile_off.as<uint32_t>(),
                                  ctx->out.as<uint8_t>(), P.k);
  ctx->launches++; stage_mark(ctx, "k_write");
  CK(cudaGetLastError());
  return 0;
}

int upload_plan(bgpu_ctx* ctx, const Job& job) {
  int rc;
  // shards, metablocks and 3 x metablocks are gridDim.y of the kernels (<= 65535).  Batches are split on the host and large
  // streams run as lane groups of <= 64 shards, so this is only reached with lanes off and tiny shards (lgwin 10: ~66 MB)
  if (job.shards.size() > 65535 || job.mbs.size() * 3 > 65535) {
    g_cuda_err = "one device job is limited to 65535 shards and 21845 metablocks (use the default lanes, or larger shards / metablocks)";
    return BGPU_E_INVALID;
  }
  if ((rc = upload(ctx, ctx->shards, job.shards.data(), job.shards.size() * sizeof(ShardD)))) return rc;
  if ((rc = upload(ctx, ctx->mbs, job.mbs.data(), job.mbs.size() * sizeof(MbD)))) return rc;
  if ((rc = upload(ctx, ctx->segs, job.segs.data(), job.segs.size() * sizeof(SegD)))) return rc;
  if ((rc = upload(ctx, ctx->tile_seg, job.tile_seg.data(), job.tile_seg.size() * 4))) return rc;
  return 0;
}

#include "model_host.inc"

// BGPU_F_VERIFY: every shard of the job decoded on the device (decode.cuh) and compared with the staged input.
int run_verify(bgpu_ctx* ctx, const Job& job, const uint8_t* d_in) {
  cudaStream_t st = ctx->stream;
  const uint32_t n_shard = (uint32_t)job.shards.size();
  int rc;
  if ((rc = ensure_lut(ctx))) return rc;
  if ((rc = ctx->dec_arena.reserve((size_t)n_shard * sizeof(DecArena)))) return rc;
  if ((rc = ctx->dec_results.reserve((size_t)n_shard * sizeof(DecResult)))) return rc;
  k_verify_shards<<<n_shard, 32, 0, st>>>(ctx->out.as<uint8_t>(), ctx->shards.as<ShardD>(), ctx->shard_bytes.as<uint32_t>(), n_shard, d_in,
                                         ctx->dec_arena.as<DecArena>(), ctx->dec_results.as<DecResult>());
  ctx->launches++; stage_mark(ctx, "k_verify");
  std::vector<DecResult> res(n_shard);
  CK(cudaMemcpyAsync(res.data(), ctx->dec_results.p, (size_t)n_shard * sizeof(DecResult), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  uint64_t bad = 0;
  for (uint32_t k = 0; k < n_shard; k++)
    if (res[k].status != BGPU_DEC_OK) {
      if (!bad) { char b[160]; snprintf(b, sizeof b, "verify: shard %u decode status %u at input offset %llu", k, res[k].status, (unsigned long long)res[k].where); g_cuda_err = b; }
      bad++;
    }
  ctx->counters[4] += bad;
  return bad ? BGPU_E_VERIFY : 0;
}

// storeMetaBlock (full modelling) unless the caller asked for storeMetaBlockTrivial everywhere.
int run_backend_any(bgpu_ctx* ctx, const Job& job, const EncParams& P, const uint8_t* d_in) {
  // metablock.ts:517-524: quality < 5 always stores trivial metablocks (as does encodeFast, encode.ts:169)
  int rc = ((ctx->tune.flags & BGPU_F_TRIVIAL_METABLOCKS) || P.quality < 5) ? run_backend(ctx, job, P, d_in) : run_backend_full(ctx, job, P, d_in);
  if (!rc && (ctx->tune.flags & BGPU_F_VERIFY)) rc = run_verify(ctx, job, d_in);
  return rc;
}

int reserve_front(bgpu_ctx* ctx, const Job& job, const EncParams& P) {
  int rc; size_t N = job.in_bytes + IN_PAD;
  if ((rc = ctx->role.reserve(N))) return rc;
  if ((rc = ctx->ins.reserve(((N >> P.k.sh) + 4) * 4))) return rc;
  if ((rc = ctx->cmds.reserve(((N >> P.k.sh) + 4) * sizeof(Cmd)))) return rc;
  return 0;
}

// Match finding + parse (K2, K3, K5, stitch).
int run_frontend(bgpu_ctx* ctx, const Job& job, const EncParams& P, const uint8_t* d_in) {
  cudaStream_t st = ctx->stream;
  int rc; size_t N = job.in_bytes + IN_PAD;
  const uint32_t n_tile = (uint32_t)job.tile_seg.size(), n_mb = (uint32_t)job.mbs.size(), n_shard = (uint32_t)job.shards.size(),
                 n_seg = (uint32_t)job.segs.size();
  if ((rc = ctx->sa_pos.reserve(N * 4))) return rc;   // pass-0 output, later `best`
  if ((rc = ctx->sa_v.reserve(N * 4))) return rc;
  if ((rc = ctx->sb_pos.reserve(N * 4))) return rc;
  if ((rc = ctx->sb_v.reserve(N * 4))) return rc;
  if ((rc = ctx->sort_hist.reserve(((size_t)job.n_stile + 1) * SORT_BINS * 4))) return rc;
  if ((rc = reserve_front(ctx, job, P))) return rc;
  if ((rc = ctx->seg_ncmd.reserve((size_t)n_seg * 4))) return rc;
  if ((rc = ctx->seg_trail.reserve((size_t)n_seg * 4))) return rc;
  if ((rc = ctx->seg_ring.reserve((size_t)n_seg * 16))) return rc;
  const ShardD* d_sh = ctx->shards.as<ShardD>(); const MbD* d_mb = ctx->mbs.as<MbD>(); const SegD* d_sg = ctx->segs.as<SegD>();
  uint32_t* sa_pos = ctx->sa_pos.as<uint32_t>(); uint32_t* sa_v = ctx->sa_v.as<uint32_t>();
  uint32_t* sb_pos = ctx->sb_pos.as<uint32_t>(); uint32_t* sb_v = ctx->sb_v.as<uint32_t>();
  uint32_t* sh_hist = ctx->sort_hist.as<uint32_t>();
  const bool simple = P.quality <= 4;
  uint32_t* best = simple ? sb_pos : sa_pos;     // the array the last sort pass did not write
  if (simple && job.max_shard_len >= 4) {
    // quality 1..4: 16/17-bit key of 5 bytes, three LSD passes (6 + 6 + 4/5 bits), one-slot table probe
    const uint32_t max_npos = job.max_shard_len - 3;
    const dim3 gs((max_npos + SORT_TILE - 1) / SORT_TILE, n_shard), gp((max_npos + 255) / 256, n_shard);
    const int bb = P.k.bucket_bits;
    k_sort_hist<2><<<gs, 256, 0, st>>>(d_in, nullptr, d_sh, sh_hist, bb, 0, 6);
    k_sort_scan<<<n_shard, 1024, 0, st>>>(d_sh, sh_hist, 1 << 6);
    k_sort_scatter<2><<<gs, 256, SORT_SCATTER_SMEM, st>>>(d_in, nullptr, nullptr, d_sh, sh_hist, sa_pos, sa_v, bb, 0, 6);
    ctx->launches += 3; stage_mark(ctx, "k_sort_pass0");
    k_sort_hist<3><<<gs, 256, 0, st>>>(d_in, sa_v, d_sh, sh_hist, bb, 6, 6);
    k_sort_scan<<<n_shard, 1024, 0, st>>>(d_sh, sh_hist, 1 << 6);
    k_sort_scatter<3><<<gs, 256, SORT_SCATTER_SMEM, st>>>(d_in, sa_pos, sa_v, d_sh, sh_hist, sb_pos, sb_v, bb, 6, 6);
    ctx->launches += 3; stage_mark(ctx, "k_sort_pass1");
    k_sort_hist<3><<<gs, 256, 0, st>>>(d_in, sb_v, d_sh, sh_hist, bb, 12, bb - 12);
    k_sort_scan<<<n_shard, 1024, 0, st>>>(d_sh, sh_hist, 1 << (bb - 12));
    k_sort_scatter<3><<<gs, 256, SORT_SCATTER_SMEM, st>>>(d_in, sb_pos, sb_v, d_sh, sh_hist, sa_pos, sa_v, bb, 12, bb - 12);
    ctx->launches += 3; stage_mark(ctx, "k_sort_pass2");
    k_probe_simple<<<gp, 256, 0, st>>>(d_in, d_sh, sa_pos, sa_v, best);
    ctx->launches++; stage_mark(ctx, "k_probe");
  } else if (job.max_shard_len >= 4) {
    const uint32_t max_npos = job.max_shard_len - 3;
    const dim3 gs((max_npos + SORT_TILE - 1) / SORT_TILE, n_shard), gp((max_npos + PROBE_TILE - 1) / PROBE_TILE, n_shard);
    const int bb = P.k.bucket_bits, nb0 = 1 << SORT_LOW_BITS, nb1 = 1 << (bb - SORT_LOW_BITS);
    // depth <= 32 (q5, q6): packed sort entries — one u32 per entry between the passes and positions only into the probe
    // (r5q: same order and same best[]; DRAM traffic of the stage 49.8 -> ~33 GB per GiB, time 45.6 -> 45.3 ms: the kernels
    // are not DRAM-bound).  Deeper chains (q7-9) keep the (position, key bytes) pairs: their probe resolves 16..19-byte
    // matches from shared memory, which the packed form sends to the input (q9: 233 -> 245 ms per GiB).
    const bool packed = P.k.depth <= 32 && !(ctx->tune.flags & BGPU_F_SORT_UNPACKED);
    const size_t sm = (size_t)(PROBE_TILE + P.k.depth) * 24;
    k_sort_hist<0><<<gs, 256, 0, st>>>(d_in, nullptr, d_sh, sh_hist, bb, 0, SORT_LOW_BITS);
    k_sort_scan<<<n_shard, 1024, 0, st>>>(d_sh, sh_hist, nb0);
    if (packed) k_sort_scatter_packed<0><<<gs, 256, 0, st>>>(d_in, nullptr, d_sh, sh_hist, sa_pos, bb, SORT_LOW_BITS);
    else k_sort_scatter<0><<<gs, 256, SORT_SCATTER_SMEM, st>>>(d_in, nullptr, nullptr, d_sh, sh_hist, sa_pos, sa_v, bb, 0, SORT_LOW_BITS);
    ctx->launches += 3; stage_mark(ctx, "k_sort_pass0");
    if (packed) k_sort_hist_packed<<<gs, 256, 0, st>>>(sa_pos, d_sh, sh_hist, bb - SORT_LOW_BITS);
    else k_sort_hist<1><<<gs, 256, 0, st>>>(d_in, sa_v, d_sh, sh_hist, bb, SORT_LOW_BITS, bb - SORT_LOW_BITS);
    k_sort_scan<<<n_shard, 1024, 0, st>>>(d_sh, sh_hist, nb1);
    if (packed) k_sort_scatter_packed<1><<<gs, 256, 0, st>>>(d_in, sa_pos, d_sh, sh_hist, sb_pos, bb, bb - SORT_LOW_BITS);
    else k_sort_scatter<1><<<gs, 256, SORT_SCATTER_SMEM, st>>>(d_in, sa_pos, sa_v, d_sh, sh_hist, sb_pos, sb_v, bb, SORT_LOW_BITS, bb - SORT_LOW_BITS);
    ctx->launches += 3; stage_mark(ctx, "k_sort_pass1");
    if (ctx->tune.flags & BGPU_F_PROBE_WINDOW) {   // the round-1 probe needs the key bytes: only with the unpacked sort
      if (packed) { g_cuda_err = "BGPU_F_PROBE_WINDOW needs BGPU_F_SORT_UNPACKED at quality 5-6"; return BGPU_E_INVALID; }
      CK(cudaFuncSetAttribute(k_probe_sorted, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PROBE_SMEM));
      k_probe_sorted<<<gp, 256, PROBE_SMEM, st>>>(d_in, d_sh, sb_pos, sb_v, best, P.k.depth, bb);
    } else if (packed) {
      CK(cudaFuncSetAttribute(k_probe_diag<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      k_probe_diag<true><<<gp, 256, sm, st>>>(d_in, d_sh, sb_pos, nullptr, best, P.k.depth, bb);
    } else {
      CK(cudaFuncSetAttribute(k_probe_diag<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
      k_probe_diag<false><<<gp, 256, sm, st>>>(d_in, d_sh, sb_pos, sb_v, best, P.k.depth, bb);
    }
    ctx->launches++; stage_mark(ctx, "k_probe");
  } else {
    CK(cudaMemsetAsync(best, 0, N * 4, st));
  }
  Params Pk = P.k; Pk.hint = 0;
  CK(cudaMemsetAsync(ctx->role.p, 0, N, st));
  // seg_aux: [0] seg_info (final), [1] seg_info of the first pass, then int4 x n_seg: first-pass end rings, assumed start rings
  if ((rc = ctx->seg_aux.reserve((size_t)n_seg * (8 + 32) + 64))) return rc;
  uint32_t* seg_info = ctx->seg_aux.as<uint32_t>(); uint32_t* seg_info0 = seg_info + n_seg;
  int4* seg_ring0 = reinterpret_cast<int4*>(ctx->seg_aux.as<uint8_t>() + (((size_t)n_seg * 8 + 15) & ~(size_t)15));
  int4* seg_assumed = seg_ring0 + n_seg;
  CK(cudaMemsetAsync(seg_assumed, 0, (size_t)n_seg * 16, st));
  const unsigned pgrid = (n_seg * PARSE_G + 127) / 128;
  k_parse<PARSE_G, false><<<pgrid, 128, 0, st>>>(d_in, d_sh, d_sg, n_seg, best, ctx->role.as<uint8_t>(), ctx->ins.as<uint32_t>(),
                                           ctx->cmds.as<Cmd>(), ctx->seg_ncmd.as<uint32_t>(), ctx->seg_trail.as<uint32_t>(), ctx->seg_ring.as<int4>(),
                                           seg_info, nullptr, nullptr, seg_assumed, Pk);
  ctx->launches++;
  if (n_seg > 1 && !(ctx->tune.flags & BGPU_F_NO_REFINE)) {
    // second pass over the segments that start inside a running copy, from the ring the first pass ended the previous segment with
    CK(cudaMemcpyAsync(seg_info0, seg_info, (size_t)n_seg * 4, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemcpyAsync(seg_ring0, ctx->seg_ring.p, (size_t)n_seg * 16, cudaMemcpyDeviceToDevice, st));
    k_parse<PARSE_G, true><<<pgrid, 128, 0, st>>>(d_in, d_sh, d_sg, n_seg, best, ctx->role.as<uint8_t>(), ctx->ins.as<uint32_t>(),
                                            ctx->cmds.as<Cmd>(), ctx->seg_ncmd.as<uint32_t>(), ctx->seg_trail.as<uint32_t>(), ctx->seg_ring.as<int4>(),
                                            seg_info, seg_ring0, seg_info0, seg_assumed, Pk);
    ctx->launches++;
  }
  stage_mark(ctx, "k_parse");
  k_stitch<<<n_mb, 32, 0, st>>>(d_sh, d_mb, d_sg, n_mb, ctx->role.as<uint8_t>(), ctx->ins.as<uint32_t>(), ctx->cmds.as<Cmd>(),
                                            ctx->seg_ncmd.as<uint32_t>(), ctx->seg_trail.as<uint32_t>(), Pk);
  ctx->launches++;
  if (n_seg > 1) {
    k_ring_fixup<<<(n_seg + 127) / 128, 128, 0, st>>>(d_sh, d_sg, n_seg, ctx->ins.as<uint32_t>(), ctx->cmds.as<Cmd>(), ctx->seg_ncmd.as<uint32_t>(),
                                                    ctx->seg_ring.as<int4>(), seg_info, seg_assumed, Pk);
    ctx->launches++;
  }
  stage_mark(ctx, "k_stitch");
  CK(cudaGetLastError());
  return 0;
}

#include "hq_host.inc"

int run_frontend_any(bgpu_ctx* ctx, const Job& job, const EncParams& P, const uint8_t* d_in) {
  return P.quality >= 10 ? run_frontend_hq(ctx, job, P, d_in) : run_frontend(ctx, job, P, d_in);
}

void timing_begin(bgpu_ctx* ctx) {
  pin_rewind(ctx);
  ctx->n_stage = 0; ctx->launches = 0; for (int i = 0; i < 8; i++) ctx->counters[i] = 0;
  if (ctx->ev_ok) cudaEventRecord(ctx->ev[0], ctx->stream);
}
void timing_end(bgpu_ctx* ctx) {
  if (!ctx->ev_ok) return;
  for (int i = 0; i < ctx->n_stage; i++) { float m = 0; cudaEventElapsedTime(&m, ctx->ev[i], ctx->ev[i + 1]); ctx->ms[i] = m; }
}

// ---- host-only paths (encode.ts:92-138; a18 in SURVEY.md §8a) ----
struct HostBits {
  std::vector<uint8_t> b; uint64_t pos = 0;
  void put(unsigned n, uint64_t v) {
    for (unsigned i = 0; i < n; i++, pos++) {
      if ((pos >> 3) >= b.size()) b.push_back(0);
      b[pos >> 3] |= (uint8_t)(((v >> i) & 1) << (pos & 7));
    }
  }
  void align() { pos = (pos + 7) & ~7ull; }
  void bytes(const uint8_t* p, size_t n) { b.resize(pos >> 3); b.insert(b.end(), p, p + n); pos += (uint64_t)n * 8; }
};
void encode_empty(std::vector<uint8_t>& out) {
  HostBits w; uint32_t n, v; window_bits(10, &n, &v);
  w.put(n, v); w.put(1, 1); w.put(1, 1);
  out = w.b;
}
void encode_uncompressed(const uint8_t* in, size_t n, std::vector<uint8_t>& out) {
  HostBits w; uint32_t wn, wv;
  int lgwin = std::max(10, std::min(24, n <= 1 ? 10 : (int)std::ceil(std::log2((double)n)) + 1));
  window_bits(lgwin, &wn, &wv);
  w.put(wn, wv);
  const size_t maxBlock = ((size_t)1 << 24) - 1;
  for (size_t pos = 0; pos < n;) {
    size_t bs = std::min(n - pos, maxBlock);
    w.put(1, 0);                                     // ISLAST = 0 (metablock.ts:257-271)
    int lg = bs == 1 ? 1 : (63 - __builtin_clzll((unsigned long long)(bs - 1))) + 1;
    int mnibbles = (lg < 16 ? 16 : lg + 3) / 4;
    w.put(2, mnibbles - 4); w.put(mnibbles * 4, bs - 1);
    w.put(1, 1);                                     // ISUNCOMPRESSED
    w.align();
    w.bytes(in + pos, bs);
    pos += bs;
  }
  w.put(1, 1); w.put(1, 1);                          // ISLAST + ISEMPTY (metablock.ts:848-853)
  out = w.b;
}

int host_special_case(const uint8_t* in, size_t n, const EncParams& P, std::vector<uint8_t>& v) {
  if (n == 0) { encode_empty(v); return 1; }
  if (P.quality == 0 || n < 64) { encode_uncompressed(in, n, v); return 1; }
  return 0;
}

// Raw form of a shard: its bytes as ONE uncompressed metablock (metablock.ts:257-271: ISLAST 0, MNIBBLES, MLEN - 1,
// ISUNCOMPRESSED 1, byte alignment, the bytes), after the window header when the shard opens the stream and followed
// by ISLAST + ISEMPTY (0x03, encode.ts:399-405) when it closes it.  The reference has no such fallback above quality 0;
// here it bounds the output (bgpu_max_compressed_size): a shard whose compressed form exceeds
// len + len / 32 + 64 bytes is replaced by its raw form — far beyond what the reference's own coding of incompressible
// bytes costs (<= 52 bytes per KiB-sized input, <= 270 bytes per MiB), so streams stay identical to the reference's.
struct RawForm { uint8_t hdr[8]; uint32_t hdr_bytes; uint32_t tail; };
uint32_t raw_shard_form(const ShardD& sh, RawForm& r) {
  HostBits w;
  if (sh.flags & 1) w.put(sh.start_nbits, sh.start_value);
  w.put(1, 0);
  const uint32_t len = sh.len - sh.hist;   // the coded bytes (all of the shard except a streaming history prefix)
  const int lg = len == 1 ? 1 : (31 - __builtin_clz(len - 1)) + 1;
  const int mnibbles = (lg < 16 ? 16 : lg + 3) / 4;
  w.put(2, mnibbles - 4); w.put(mnibbles * 4, len - 1);
  w.put(1, 1);
  w.align();
  r.hdr_bytes = (uint32_t)(w.pos >> 3); r.tail = (sh.flags & 2) ? 1 : 0;
  memcpy(r.hdr, w.b.data(), r.hdr_bytes);
  return r.hdr_bytes + len + r.tail;
}
inline bool shard_goes_raw(const ShardD& sh, uint32_t compressed_bytes) {
  const uint32_t len = sh.len - sh.hist;
  return len > 0 && (uint64_t)compressed_bytes > (uint64_t)len + (len >> 5) + 64;
}
// Bytes shard `sh` contributes to the stream (compressed, or raw when that is what will be written).
uint32_t shard_out_bytes(const ShardD& sh, uint32_t compressed_bytes) {
  if (!shard_goes_raw(sh, compressed_bytes)) return compressed_bytes;
  RawForm r; return raw_shard_form(sh, r);
}
// Queues the copies that put shard `sh` at dst (host): its compressed bytes from the device output region, or its raw
// form straight from the staged input.  Returns the bytes written.
int shard_to_host(const ShardD& sh, uint32_t compressed_bytes, const uint8_t* d_out, const uint8_t* d_in, uint8_t* dst, cudaStream_t st, uint32_t* written) {
  if (!shard_goes_raw(sh, compressed_bytes)) {
    CK(cudaMemcpyAsync(dst, d_out + sh.out_off, compressed_bytes, cudaMemcpyDeviceToHost, st));
    *written = compressed_bytes;
    return 0;
  }
  RawForm r; const uint32_t total = raw_shard_form(sh, r);
  memcpy(dst, r.hdr, r.hdr_bytes);
  CK(cudaMemcpyAsync(dst + r.hdr_bytes, d_in + sh.in_off + sh.hist, sh.len - sh.hist, cudaMemcpyDeviceToHost, st));
  if (r.tail) dst[r.hdr_bytes + sh.len - sh.hist] = 0x03;
  *written = total;
  return 0;
}

// Copies every stream's shards, concatenated, to the caller's buffers.  d_in: the staged input the job ran on.
int collect(bgpu_ctx* ctx, const Job& job, const uint8_t* d_in, uint8_t* const* out, const size_t* out_cap, size_t* out_len, const int* map, int count) {
  cudaStream_t st = ctx->stream;
  ctx->h_shard_bytes.resize(job.shards.size());
  CK(cudaMemcpyAsync(ctx->h_shard_bytes.data(), ctx->shard_bytes.p, job.shards.size() * 4, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  int rc = 0;
  for (int s = 0; s < count; s++) {
    const StreamPlan& sp = job.streams[s];
    size_t total = 0;
    for (uint32_t k = 0; k < sp.n_shard; k++) {
      uint32_t b = ctx->h_shard_bytes[sp.first_shard + k];
      if (b > job.shards[sp.first_shard + k].out_cap) { g_cuda_err = "shard output exceeded its device region"; return BGPU_E_CUDA; }
      total += shard_out_bytes(job.shards[sp.first_shard + k], b);
    }
    int dst = map ? map[s] : s;
    out_len[dst] = total;
    if (total > out_cap[dst] || !out[dst]) { rc = BGPU_E_OUTPUT_TOO_SMALL; continue; }
    size_t o = 0;
    for (uint32_t k = 0; k < sp.n_shard; k++) {
      const ShardD& sh = job.shards[sp.first_shard + k];
      uint32_t b = ctx->h_shard_bytes[sp.first_shard + k], wr = 0;
      int r2 = shard_to_host(sh, b, ctx->out.as<uint8_t>(), d_in, out[dst] + o, st, &wr);
      if (r2) return r2;
      o += wr;
    }
  }
  CK(cudaStreamSynchronize(st));
  return rc;
}

#include "lanes_host.inc"

}  // namespace
static int encode_part_host(bgpu_ctx* c, const uint8_t* in, size_t n, const uint8_t* prev2, unsigned part_flags, const EncParams& P,
                            uint8_t* out, size_t out_cap, size_t* out_len);
namespace {
#include "multi_host.inc"

}  // namespace

extern "C" {

int bgpu_ctx_create(int device_id, bgpu_ctx** out) {
  if (!out) return BGPU_E_INVALID;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) { g_cuda_err = e != cudaSuccess ? cudaGetErrorString(e) : "no CUDA device"; return BGPU_E_NO_DEVICE; }
  bgpu_ctx* c = new bgpu_ctx();
  if (device_id >= 0) { CK(cudaSetDevice(device_id)); c->device = device_id; } else { CK(cudaGetDevice(&c->device)); }
  CK(cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking));
  c->stream = c->own_stream;
  c->ev_ok = true;
  for (int i = 0; i <= bgpu_ctx::kMaxStages; i++) if (cudaEventCreate(&c->ev[i]) != cudaSuccess) c->ev_ok = false;
  *out = c;
  return 0;
}

int bgpu_ctx_create_multi(const int* device_ids, int n_dev, bgpu_ctx** out) {
  if (!out || !device_ids || n_dev < 1) return BGPU_E_INVALID;
  bgpu_ctx* c = nullptr;
  int rc = bgpu_ctx_create(device_ids[0], &c);
  if (rc) return rc;
  for (int i = 0; i < n_dev && n_dev > 1; i++) {
    bgpu_ctx* q = nullptr;
    if ((rc = bgpu_ctx_create(device_ids[i], &q))) { bgpu_ctx_destroy(c); return rc; }
    q->is_peer = true;
    c->peers.push_back(q);
  }
  cudaSetDevice(c->device);
  *out = c;
  return 0;
}

void bgpu_ctx_destroy(bgpu_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  DBuf* all[] = {&c->in, &c->sa_pos, &c->sa_v, &c->sb_pos, &c->sb_v, &c->sort_hist, &c->role, &c->ins, &c->cmds, &c->out, &c->shards, &c->mbs, &c->segs, &c->tile_seg, &c->seg_ncmd,
                 &c->seg_trail, &c->seg_ring, &c->hist, &c->codes, &c->tree_bits, &c->tree_nbits, &c->tile_bits, &c->tile_off, &c->tile_model, &c->warp_base, &c->thr_off, &c->shard_bytes, &c->huff_meta,
                 &c->model_arena, &c->misc_arena, &c->d_models, &c->d_splits, &c->d_cb, &c->d_cx, &c->d_bh_ptrs, &c->d_bh_strides, &c->d_cx_outs,
                 &c->h10_bstart, &c->h10_bend, &c->h10_short, &c->h10_forest, &c->h10_m, &c->h10_off, &c->h10_cnt, &c->h10_cursor, &c->h10_full, &c->h10_hot,
                 &c->zop_bytehist, &c->zop_model, &c->zop_lc, &c->zop_nodes, &c->zop_ring, &c->stats, &c->dec_arena, &c->dec_results, &c->dec_jobs, &c->seg_aux, &c->tile_stage, &c->tile_flag, &c->fb_count};
  for (DBuf* b : all) b->release();
  for (bgpu_ctx* l : c->lanes) bgpu_ctx_destroy(l);
  c->lanes.clear();
  for (bgpu_ctx* q : c->peers) bgpu_ctx_destroy(q);
  c->peers.clear();
  cudaSetDevice(c->device);
  if (c->part_out) cudaFreeHost(c->part_out);
  if (c->lane_ev) cudaEventDestroy(c->lane_ev);
  if (c->fork_ev) cudaEventDestroy(c->fork_ev);
  if (c->h2d_ev) cudaEventDestroy(c->h2d_ev);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  if (c->h_sizes) cudaFreeHost(c->h_sizes);
  if (c->pin) cudaFreeHost(c->pin);
  if (c->ev_ok) for (int i = 0; i <= bgpu_ctx::kMaxStages; i++) cudaEventDestroy(c->ev[i]);
  if (c->hi_stream) cudaStreamDestroy(c->hi_stream);
  if (c->hi_ev) cudaEventDestroy(c->hi_ev);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  delete c;
}

int bgpu_ctx_set_tuning(bgpu_ctx* c, const bgpu_tuning* t) {
  if (!c || !t) return BGPU_E_INVALID;
  c->tune = *t;
  for (bgpu_ctx* q : c->peers) q->tune = *t;
  return 0;
}

int bgpu_ctx_set_stream(bgpu_ctx* c, void* cuda_stream) {
  if (!c) return BGPU_E_INVALID;
  c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
  return 0;
}

// Per shard of len bytes the stream holds at most len + len / 32 + 64 bytes (shard_goes_raw) or the raw form
// (<= len + 6); shards are >= 1008 bytes for every lgwin at the default tuning (plan_shard_bytes), the last one may be
// shorter.  Host-only forms (encode.ts:92-138): n + 4 per 16 MiB block + <= 6.  (SURVEY.md section 8b asked for
// n + 4 ceil(n / 2^24) + 16 shards + 16, which presumes a break-even raw fallback; the threshold here is looser so that
// the reference's own coding of incompressible shards is never replaced.)
size_t bgpu_max_compressed_size(size_t n) { return n + (n >> 5) + 70 * (n / 1008 + 1) + 16; }

int bgpu_encode_batch(bgpu_ctx* c, int count, const uint8_t* const* in, const size_t* n, int quality, int lgwin, int mode,
                      uint8_t* const* out, const size_t* out_cap, size_t* out_len) {
  if (!c || count < 0 || (count && (!in || !n || !out || !out_cap || !out_len))) return BGPU_E_INVALID;
  EncParams P = make_params(quality, lgwin, mode);
  int rc;
  CK(cudaSetDevice(c->device));
  // host-only special cases first; the rest goes to the device in one job
  std::vector<int> map; std::vector<size_t> lens; std::vector<const uint8_t*> ptrs;
  int result = 0;
  for (int i = 0; i < count; i++) {
    if (n[i] && !in[i]) return BGPU_E_INVALID;
    std::vector<uint8_t> v;
    if (host_special_case(in[i], n[i], P, v)) {
      out_len[i] = v.size();
      if (v.size() > out_cap[i] || !out[i]) result = BGPU_E_OUTPUT_TOO_SMALL; else memcpy(out[i], v.data(), v.size());
    } else { map.push_back(i); lens.push_back(n[i]); ptrs.push_back(in[i]); }
  }
  if (map.empty()) return result;
  // Shards, metablocks and 3 x metablocks index gridDim.y (<= 65535) in the kernels: a batch of many streams goes to the
  // device as several jobs of at most BATCH_JOB_STREAMS streams, each holding at most BATCH_JOB_BYTES.
  if (lens.size() > 1) {
    constexpr size_t BATCH_JOB_STREAMS = 8192, BATCH_JOB_BYTES = (size_t)8 << 30;
    size_t first = 0, bytes = 0, cnt = 0; bool split = false;
    for (size_t s = 0; s < lens.size(); s++) { bytes += lens[s]; cnt++; if (cnt > BATCH_JOB_STREAMS || bytes > BATCH_JOB_BYTES) { split = true; break; } }
    if (split) {
      while (first < lens.size()) {
        size_t last = first, b = 0;
        while (last < lens.size() && last - first < BATCH_JOB_STREAMS && (last == first || b + lens[last] <= BATCH_JOB_BYTES)) b += lens[last++];
        const int k = (int)(last - first);
        std::vector<const uint8_t*> in_k(k); std::vector<size_t> n_k(k), cap_k(k), len_k(k, 0); std::vector<uint8_t*> out_k(k);
        for (int i = 0; i < k; i++) { const int o = map[first + i]; in_k[i] = ptrs[first + i]; n_k[i] = lens[first + i]; out_k[i] = out[o]; cap_k[i] = out_cap[o]; }
        rc = bgpu_encode_batch(c, k, in_k.data(), n_k.data(), quality, lgwin, mode, out_k.data(), cap_k.data(), len_k.data());
        for (int i = 0; i < k; i++) out_len[map[first + i]] = len_k[i];
        if (rc == BGPU_E_OUTPUT_TOO_SMALL) result = rc; else if (rc) return rc;
        first = last;
      }
      return result;
    }
  }
  if (lens.size() == 1 && multi_apply(c, P, lens[0])) {
    rc = encode_multi(c, ptrs[0], lens[0], P, out[map[0]], out_cap[map[0]], &out_len[map[0]]);
    return rc ? rc : result;
  }
  if (lens.size() > 1 && c->peers.size() > 1) {
    rc = encode_batch_multi(c, ptrs, lens, P, out, out_cap, out_len, map);
    return rc ? rc : result;
  }
  {
    uint32_t L; size_t gb;
    if (lens.size() == 1 && lanes_apply(c, P, lens[0], &L, &gb)) {
      rc = encode_lanes(c, ptrs[0], nullptr, lens[0], P, L, gb, out[map[0]], out_cap[map[0]], &out_len[map[0]], nullptr);
      if (!rc && !out[map[0]]) rc = BGPU_E_OUTPUT_TOO_SMALL;   // a size query, as on the one-job path (collect)
      return rc ? rc : result;
    }
  }
  Job job; plan_job(c, P, lens.data(), (int)lens.size(), job);
  if ((rc = c->in.reserve(job.in_bytes + IN_PAD))) return rc;
  timing_begin(c);
  for (size_t s = 0; s < lens.size(); s++) {
    CK(cudaMemcpyAsync(c->in.as<uint8_t>() + job.streams[s].in_off, ptrs[s], lens[s], cudaMemcpyHostToDevice, c->stream));
    size_t padded = ((lens[s] + 255) & ~(size_t)255) + (s + 1 == lens.size() ? IN_PAD : 0);
    CK(cudaMemsetAsync(c->in.as<uint8_t>() + job.streams[s].in_off + lens[s], 0, padded - lens[s], c->stream));
  }
  stage_mark(c, "h2d");
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = run_frontend_any(c, job, P, c->in.as<uint8_t>()))) return rc;
  if ((rc = run_backend_any(c, job, P, c->in.as<uint8_t>()))) return rc;
  rc = collect(c, job, c->in.as<uint8_t>(), out, out_cap, out_len, map.data(), (int)map.size());
  timing_end(c);
  return rc ? rc : result;
}

int bgpu_encode(bgpu_ctx* c, const uint8_t* in, size_t n, int quality, int lgwin, int mode, uint8_t* out, size_t out_cap, size_t* out_len) {
  if (!out_len) return BGPU_E_INVALID;
  const uint8_t* ins[1] = {in}; size_t ns[1] = {n}; uint8_t* outs[1] = {out}; size_t caps[1] = {out_cap};
  return bgpu_encode_batch(c, 1, ins, ns, quality, lgwin, mode, outs, caps, out_len);
}

// One contiguous piece of a stream that other contexts (other GPUs) encode the rest of.
int bgpu_encode_part(bgpu_ctx* c, const uint8_t* in, size_t n, const uint8_t* prev2, unsigned part_flags, int quality, int lgwin, int mode,
                     uint8_t* out, size_t out_cap, size_t* out_len) {
  if (!c || !out_len || (n && !in)) return BGPU_E_INVALID;
  const bool first = part_flags & BGPU_PART_FIRST;
  if (!first && n && !prev2) return BGPU_E_INVALID;
  EncParams P = make_params(quality, lgwin, mode);
  if (P.quality < 1) return BGPU_E_UNSUPPORTED;   // the host-only uncompressed path has no part form
  return encode_part_host(c, in, n, prev2, part_flags, P, out, out_cap, out_len);
}

}  // extern "C"

// The body of bgpu_encode_part (also the per-device worker of multi_host.inc): lanes when the part is large enough.
static int encode_part_host(bgpu_ctx* c, const uint8_t* in, size_t n, const uint8_t* prev2, unsigned part_flags, const EncParams& P,
                            uint8_t* out, size_t out_cap, size_t* out_len) {
  const bool first = part_flags & BGPU_PART_FIRST, last = part_flags & BGPU_PART_LAST;
  int rc;
  if (n == 0) {
    // nothing to code: the header (FIRST) and / or ISLAST + ISEMPTY (LAST), encode.ts:399-405
    HostBits w;
    if (first) { uint32_t wn, wv; window_bits(P.lgwin, &wn, &wv); w.put(wn, wv); }
    if (last) { w.put(1, 1); w.put(1, 1); }
    else if (first) { w.put(6, 6); }             // byte-align the header with an empty metadata block
    *out_len = w.b.size();
    if (w.b.size() > out_cap || (w.b.size() && !out)) return BGPU_E_OUTPUT_TOO_SMALL;
    if (!w.b.empty()) memcpy(out, w.b.data(), w.b.size());
    return 0;
  }
  CK(cudaSetDevice(c->device));
  {
    uint32_t L; size_t gb;
    if (lanes_apply(c, P, n, &L, &gb)) return encode_lanes(c, in, nullptr, n, P, L, gb, out, out_cap, out_len, nullptr, part_flags, prev2);
  }
  const size_t lead = first ? 0 : 256;
  uint32_t pf[1] = {part_flags}; size_t lens[1] = {n};
  Job job; plan_job(c, P, lens, 1, job, pf, lead);
  if ((rc = c->in.reserve(job.in_bytes + IN_PAD))) return rc;
  timing_begin(c);
  uint8_t* d_in = c->in.as<uint8_t>();
  if (lead) {
    CK(cudaMemsetAsync(d_in, 0, lead, c->stream));
    CK(cudaMemcpyAsync(d_in + lead - 2, prev2, 2, cudaMemcpyHostToDevice, c->stream));
  }
  CK(cudaMemcpyAsync(d_in + lead, in, n, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(d_in + lead + n, 0, job.in_bytes + IN_PAD - lead - n, c->stream));
  stage_mark(c, "h2d");
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = run_frontend_any(c, job, P, d_in))) return rc;
  if ((rc = run_backend_any(c, job, P, d_in))) return rc;
  uint8_t* outs[1] = {out}; size_t caps[1] = {out_cap};
  rc = collect(c, job, d_in, outs, caps, out_len, nullptr, 1);
  timing_end(c);
  return rc;
}

extern "C" {

int bgpu_encode_device(bgpu_ctx* c, const void* d_in, size_t n, int quality, int lgwin, int mode, uint8_t* out, size_t out_cap,
                       size_t* out_len, float* kernel_ms) {
  if (!c || !d_in || !out_len || n < 64) return BGPU_E_INVALID;
  EncParams P = make_params(quality, lgwin, mode);
  int rc;
  if (P.quality == 0) return BGPU_E_INVALID;
  CK(cudaSetDevice(c->device));
  {
    uint32_t L; size_t gb;
    if (lanes_apply(c, P, n, &L, &gb)) return encode_lanes(c, nullptr, (const uint8_t*)d_in, n, P, L, gb, out, out_cap, out_len, kernel_ms);
  }
  Job job; size_t lens[1] = {n}; plan_job(c, P, lens, 1, job);
  timing_begin(c);
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = run_frontend_any(c, job, P, (const uint8_t*)d_in))) return rc;
  if ((rc = run_backend_any(c, job, P, (const uint8_t*)d_in))) return rc;
  if (out) {
    uint8_t* outs[1] = {out}; size_t caps[1] = {out_cap};
    rc = collect(c, job, (const uint8_t*)d_in, outs, caps, out_len, nullptr, 1);
  } else {
    c->h_shard_bytes.resize(job.shards.size());
    CK(cudaMemcpyAsync(c->h_shard_bytes.data(), c->shard_bytes.p, job.shards.size() * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    size_t total = 0;
    for (size_t k = 0; k < job.shards.size(); k++) total += shard_out_bytes(job.shards[k], c->h_shard_bytes[k]);
    *out_len = total;
  }
  timing_end(c);
  if (kernel_ms) { float t = 0; for (int i = 0; i < c->n_stage; i++) t += c->ms[i]; *kernel_ms = t; }
  return rc;
}

// ------------------------------------------------------------------ streaming (encode.ts:290-409)
}  // extern "C"

struct bgpu_stream {
  bgpu_ctx* ctx; int quality, lgwin, mode;
  std::vector<uint8_t> pending;   // bytes not yet encoded (the reference copies each chunk, encode.ts:360-363)
  std::vector<uint8_t> outbuf;
  bool header_done = false, finished = false;
  size_t total_in = 0;
  // Device-resident window (the reference's ring buffer, encode.ts:411-489): win[cur] holds the last hist_len coded
  // bytes; a flush appends the new bytes behind them, encodes the shard [history | new] with metablocks only over
  // the new part, and the window slides (into the other buffer) when history + new would exceed the window.
  DBuf win[2]; int cur = 0; size_t hist_len = 0;
  uint8_t prev2[2] = {0, 0};   // the last two coded bytes (literal context of the next piece when no history is kept)
};

extern "C" {

int bgpu_stream_create(bgpu_ctx* c, int quality, int lgwin, int mode, bgpu_stream** s) {
  if (!c || !s) return BGPU_E_INVALID;
  EncParams P = make_params(quality, lgwin, mode);
  // streaming q0..4 uses SimpleHasher (encode.ts:343-344) and has no uncompressed path: quality 0 takes the
  // hasher of quality 1 (same 17-bit table, same sparse stores, lgblock = lgwin for both)
  if (P.quality < 1) P = make_params(1, lgwin, mode);
  bgpu_stream* st = new bgpu_stream(); st->ctx = c; st->quality = P.quality; st->lgwin = P.lgwin; st->mode = mode;
  *s = st;
  return 0;
}

}  // extern "C"

namespace {
// enc-constants.ts:129-147 computeLgBlock: the reference's update() emits one metablock per 1 << lgblock input bytes
uint32_t stream_block_bytes(const EncParams& P) {
  int lgblock;
  if (P.quality <= 1) lgblock = P.lgwin;
  else if (P.quality < 4) lgblock = 14;
  else { lgblock = 16; if (P.quality >= 9 && P.lgwin > lgblock) lgblock = std::min(18, P.lgwin); }
  return 1u << lgblock;
}
// Window W (= the shard size of the one-shot path), the flush unit (a block, but at most half the window so that at
// least as much history as new data fits) and the most new bytes one job takes.  Windows below 8 KiB keep no
// history: every flush is an independent shard, as in the one-shot path.
struct StreamGeom { uint32_t window, unit, max_new; bool history; };
StreamGeom stream_geom(const bgpu_ctx* c, const EncParams& P) {
  StreamGeom g; g.window = plan_shard_bytes(c, P);
  g.history = g.window >= 8192;
  if (!g.history) { g.unit = g.max_new = g.window; return g; }
  const uint32_t half = (g.window / 2) & ~4095u;
  g.unit = std::min(stream_block_bytes(P), half);
  g.max_new = (half / g.unit) * g.unit;
  return g;
}
}  // namespace

// Encodes the first `n` pending bytes as the next piece of the stream (first: carries the window header; last: ISLAST;
// otherwise byte-aligned by an empty metadata block): one job over [history | new bytes] on the device window.
static int stream_flush(bgpu_stream* s, size_t n, bool last) {
  bgpu_ctx* c = s->ctx;
  EncParams P = make_params(s->quality, s->lgwin, s->mode);
  CK(cudaSetDevice(c->device));
  const StreamGeom g = stream_geom(c, P);
  int rc;
  size_t H = g.history ? s->hist_len : 0;
  const size_t lead = (!g.history && s->header_done) ? 256 : 0;   // no history: the two previous bytes sit right before the piece
  const size_t cap = (size_t)g.window + IN_PAD + 4096 + 256;
  if ((rc = s->win[0].reserve(cap)) || (rc = s->win[1].reserve(cap))) return rc;
  if (H + n > g.window) {
    // slide: the newest (window - n) history bytes, rounded down to the 4-byte granularity of the tile loads
    const size_t keep = (g.window - n) & ~(size_t)3;
    CK(cudaMemcpyAsync(s->win[s->cur ^ 1].p, s->win[s->cur].as<uint8_t>() + (H - keep), keep, cudaMemcpyDeviceToDevice, c->stream));
    s->cur ^= 1; H = keep;
  }
  uint8_t* d_in = s->win[s->cur].as<uint8_t>();
  bgpu_tuning saved = c->tune;
  if (!c->tune.metablock_bytes && g.history) c->tune.metablock_bytes = g.unit;   // one metablock per block, as encode.ts:366-374
  Job job; size_t lens[1] = {H + n}; uint32_t pf[1] = {(s->header_done ? 0u : BGPU_PART_FIRST) | (last ? BGPU_PART_LAST : 0u)};
  plan_job(c, P, lens, 1, job, pf, lead, H);
  c->tune = saved;
  if (job.shards.size() != 1) return BGPU_E_INVALID;
  timing_begin(c);
  if (lead) {
    CK(cudaMemsetAsync(d_in, 0, lead, c->stream));
    CK(cudaMemcpyAsync(d_in + lead - 2, s->prev2, 2, cudaMemcpyHostToDevice, c->stream));
  }
  CK(cudaMemcpyAsync(d_in + lead + H, s->pending.data(), n, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(d_in + lead + H + n, 0, cap - lead - H - n, c->stream));
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = run_frontend_any(c, job, P, d_in))) return rc;
  if ((rc = run_backend_any(c, job, P, d_in))) return rc;
  const size_t ocap = n + (n >> 5) + 4096, old = s->outbuf.size();
  size_t len = 0;
  s->outbuf.resize(old + ocap);
  uint8_t* outs[1] = {s->outbuf.data() + old}; size_t caps[1] = {ocap};
  rc = collect(c, job, d_in, outs, caps, &len, nullptr, 1);
  s->outbuf.resize(old + (rc ? 0 : len));
  timing_end(c);
  if (rc) return rc;
  s->header_done = true;
  s->hist_len = g.history ? H + n : 0;
  if (n >= 2) { s->prev2[0] = s->pending[n - 2]; s->prev2[1] = s->pending[n - 1]; }
  else if (n == 1) { s->prev2[0] = s->prev2[1]; s->prev2[1] = s->pending[0]; }
  s->pending.erase(s->pending.begin(), s->pending.begin() + n);
  return 0;
}

extern "C" {

int bgpu_stream_update(bgpu_stream* s, const uint8_t* chunk, size_t n, const uint8_t** out, size_t* out_len) {
  if (!s || (n && !chunk) || !out || !out_len) return BGPU_E_INVALID;
  if (s->finished) return BGPU_E_STATE;
  s->outbuf.clear();
  s->pending.insert(s->pending.end(), chunk, chunk + n);
  s->total_in += n;
  EncParams P = make_params(s->quality, s->lgwin, s->mode);
  const StreamGeom g = stream_geom(s->ctx, P);
  // encode.ts:366-374: whole blocks are emitted as soon as they are complete — all the complete blocks of this call in
  // as few jobs as the window allows; at least one byte stays pending so that finish() can mark the final metablock ISLAST
  while (s->pending.size() > g.unit) {
    const size_t whole = ((s->pending.size() - 1) / g.unit) * g.unit;
    int rc = stream_flush(s, std::min<size_t>(whole, g.max_new), false);
    if (rc) return rc;
  }
  *out = s->outbuf.data(); *out_len = s->outbuf.size();
  return 0;
}

int bgpu_stream_finish(bgpu_stream* s, const uint8_t** out, size_t* out_len) {
  if (!s || !out || !out_len) return BGPU_E_INVALID;
  if (s->finished) return BGPU_E_STATE;
  s->outbuf.clear();
  if (!s->pending.empty()) {
    EncParams P = make_params(s->quality, s->lgwin, s->mode);
    const StreamGeom g = stream_geom(s->ctx, P);
    while (!s->pending.empty()) {
      const size_t n = std::min<size_t>(s->pending.size(), g.max_new);
      int rc = stream_flush(s, n, n == s->pending.size());
      if (rc) return rc;
    }
  } else {
    // encode.ts:399-405: ISLAST + ISEMPTY (after the window header if nothing was written yet)
    HostBits w;
    if (!s->header_done) { uint32_t wn, wv; window_bits(s->lgwin, &wn, &wv); w.put(wn, wv); }
    w.put(1, 1); w.put(1, 1);
    s->outbuf = w.b;
  }
  s->finished = true;
  *out = s->outbuf.data(); *out_len = s->outbuf.size();
  return 0;
}

void bgpu_stream_destroy(bgpu_stream* s) {
  if (!s) return;
  if (s->ctx) cudaSetDevice(s->ctx->device);
  s->win[0].release(); s->win[1].release();
  delete s;
}

// ------------------------------------------------------------------ stage entry points
int bgpu_stage_match_candidates(bgpu_ctx* c, const uint8_t* in, size_t n, int quality, int lgwin, uint32_t* len, uint32_t* dist) {
  if (!c || !in || !len || !dist || n == 0 || n > ((size_t)1 << 22)) return BGPU_E_INVALID;
  EncParams P = make_params(quality, lgwin, 0);
  CK(cudaSetDevice(c->device));
  bgpu_tuning saved = c->tune; c->tune.shard_bytes = 1u << 22;
  Job job; size_t lens[1] = {n}; plan_job(c, P, lens, 1, job);
  c->tune = saved;
  int rc;
  if ((rc = c->in.reserve(job.in_bytes + IN_PAD))) return rc;
  timing_begin(c);
  CK(cudaMemcpyAsync(c->in.p, in, n, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(c->in.as<uint8_t>() + n, 0, job.in_bytes + IN_PAD - n, c->stream));
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = run_frontend(c, job, P, c->in.as<uint8_t>()))) return rc;
  std::vector<uint32_t> b(n);
  CK(cudaMemcpyAsync(b.data(), P.quality <= 4 ? c->sb_pos.p : c->sa_pos.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  timing_end(c);
  for (size_t i = 0; i < n; i++) { len[i] = b[i] >> 24; dist[i] = (b[i] >> 24) ? (b[i] & 0xFFFFFF) : 0; }
  return 0;
}

// Rebuilds the command list from role / ins / cmds (test path; the production path never needs it).
static void rebuild_commands(const std::vector<uint8_t>& role, const std::vector<uint32_t>& ins, const std::vector<Cmd>& cmds, size_t n, int sh,
                             std::vector<Cmd>& out) {
  for (size_t p = 0; p < n; p++) {
    if (!(role[p] & 4)) continue;
    uint32_t il = ins[p >> sh];
    if (role[p] & 8) out.push_back(make_insert_cmd(il));
    else out.push_back(cmds[(p + il) >> sh]);
  }
}

int bgpu_stage_commands(bgpu_ctx* c, const uint8_t* in, size_t n, int quality, int lgwin, int mode, void* cmds_out, size_t cmd_cap, size_t* ncmd) {
  if (!c || !in || !ncmd || n == 0 || n > ((size_t)1 << 22)) return BGPU_E_INVALID;
  EncParams P = make_params(quality, lgwin, mode);
  int rc;
  CK(cudaSetDevice(c->device));
  Job job; size_t lens[1] = {n}; plan_job(c, P, lens, 1, job);
  if (job.shards.size() != 1) return BGPU_E_INVALID;
  if ((rc = c->in.reserve(job.in_bytes + IN_PAD))) return rc;
  timing_begin(c);
  CK(cudaMemcpyAsync(c->in.p, in, n, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(c->in.as<uint8_t>() + n, 0, job.in_bytes + IN_PAD - n, c->stream));
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = run_frontend_any(c, job, P, c->in.as<uint8_t>()))) return rc;
  size_t slots = (n >> P.k.sh) + 2;
  std::vector<uint8_t> role(n); std::vector<uint32_t> ins(slots); std::vector<Cmd> cm(slots);
  CK(cudaMemcpyAsync(role.data(), c->role.p, n, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(ins.data(), c->ins.p, slots * 4, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(cm.data(), c->cmds.p, slots * sizeof(Cmd), cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  timing_end(c);
  std::vector<Cmd> list; rebuild_commands(role, ins, cm, n, P.k.sh, list);
  *ncmd = list.size();
  if (list.size() > cmd_cap || !cmds_out) return BGPU_E_OUTPUT_TOO_SMALL;
  memcpy(cmds_out, list.data(), list.size() * sizeof(Cmd));
  return 0;
}

// roles / ins / cmds arrays (kernels.cuh layout) of a caller-supplied command list covering exactly n bytes.
// counts (optional): literals, commands, commands that code a distance.
static int roles_from_commands(const Cmd* cl, size_t ncmd, size_t n, int sh, std::vector<uint8_t>& role, std::vector<uint32_t>& ins,
                               std::vector<Cmd>& cm, uint32_t* counts) {
  const size_t slots = (n >> sh) + 2;
  role.assign(n + IN_PAD, 0); ins.assign(slots, 0); cm.resize(slots);
  memset(cm.data(), 0, slots * sizeof(Cmd));
  size_t pos = 0; uint32_t nl = 0, nd = 0;
  for (size_t i = 0; i < ncmd; i++) {
    const Cmd& k = cl[i];
    uint32_t copyLen = k.copyLen & 0x1FFFFFF;
    if (pos + k.insertLen + copyLen > n) return -1;
    size_t start = pos;
    for (uint32_t j = 0; j < k.insertLen; j++) role[pos + j] = 1;
    pos += k.insertLen; nl += k.insertLen;
    if (copyLen) { role[pos] |= 2; cm[pos >> sh] = k; if (k.cmdPrefix >= 128) nd++; }
    else if (i + 1 != ncmd) return -1;  // an insert-only command must be last
    role[start] |= 4 | (copyLen ? 0 : 8);
    ins[start >> sh] = k.insertLen;
    pos += copyLen;
  }
  if (pos != n) return -1;
  if (counts) { counts[0] = nl; counts[1] = (uint32_t)ncmd; counts[2] = nd; }
  return 0;
}

int bgpu_stage_store_trivial(bgpu_ctx* c, const uint8_t* in, size_t n, const void* cmds_in, size_t ncmd, int is_last, int mode,
                             unsigned start_nbits, unsigned start_value, uint8_t* out, size_t out_cap, size_t* out_len, uint64_t* out_bits,
                             void* tables) {
  if (!c || !in || !cmds_in || !out_len || n == 0 || n > ((size_t)1 << 24) || ncmd == 0) return BGPU_E_INVALID;
  EncParams P = make_params(9, 24, mode);
  P.k.sh = 1;  // caller-supplied commands may hold 2- and 3-byte copies (Zopfli)
  CK(cudaSetDevice(c->device));
  // one shard, one metablock, one segment covering the whole input
  bgpu_tuning saved = c->tune; c->tune.shard_bytes = 1u << 22; c->tune.segment_bytes = 0xFFFFFFFFu; c->tune.metablock_bytes = 0;
  Job job; size_t lens[1] = {std::min<size_t>(n, (size_t)1 << 22)}; plan_job(c, P, lens, 1, job);
  c->tune = saved;
  if (n > ((size_t)1 << 22)) return BGPU_E_INVALID;
  ShardD& sh = job.shards[0];
  sh.flags = (start_nbits ? 1u : 0u) | 2u;   // never append the padding block here
  sh.start_nbits = start_nbits; sh.start_value = start_value;
  job.mbs[0].is_last = is_last ? 1 : 0;
  size_t slots = (n >> P.k.sh) + 2;
  std::vector<uint8_t> role; std::vector<uint32_t> ins; std::vector<Cmd> cm;
  if (roles_from_commands((const Cmd*)cmds_in, ncmd, n, P.k.sh, role, ins, cm, nullptr)) return BGPU_E_INVALID;
  int rc;
  if ((rc = c->in.reserve(job.in_bytes + IN_PAD))) return rc;
  if ((rc = reserve_front(c, job, P))) return rc;
  timing_begin(c);
  CK(cudaMemcpyAsync(c->in.p, in, n, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(c->in.as<uint8_t>() + n, 0, job.in_bytes + IN_PAD - n, c->stream));
  CK(cudaMemcpyAsync(c->role.p, role.data(), n + IN_PAD, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(c->ins.p, ins.data(), slots * 4, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(c->cmds.p, cm.data(), slots * sizeof(Cmd), cudaMemcpyHostToDevice, c->stream));
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = run_backend(c, job, P, c->in.as<uint8_t>()))) return rc;
  uint8_t* outs[1] = {out}; size_t caps[1] = {out_cap};
  rc = collect(c, job, c->in.as<uint8_t>(), outs, caps, out_len, nullptr, 1);
  timing_end(c);
  if (out_bits) *out_bits = (uint64_t)*out_len * 8;
  if (tables) {
    // layout of oracle::TrivialTables: 3 histograms (256, 704, 544 u32), 3 depth arrays, 3 code arrays
    std::vector<uint32_t> h(CODE_STRIDE), cd(CODE_STRIDE);
    CK(cudaMemcpy(h.data(), c->hist.p, CODE_STRIDE * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(cd.data(), c->codes.p, CODE_STRIDE * 4, cudaMemcpyDeviceToHost));
    uint8_t* t = (uint8_t*)tables;
    uint32_t* th = (uint32_t*)t; uint8_t* td = t + (256 + 704 + 544) * 4; uint16_t* tb = (uint16_t*)(td + 256 + 704 + 544);
    memset(t, 0, (256 + 704 + 544) * 7);
    const int srcOff[3] = {0, ALPHA_LIT, ALPHA_LIT + ALPHA_CMD}, dstOff[3] = {0, 256, 960}, cnt[3] = {256, 704, (int)P.k.dist_alphabet};
    for (int a = 0; a < 3; a++)
      for (int i = 0; i < cnt[a]; i++) {
        th[dstOff[a] + i] = h[srcOff[a] + i];
        td[dstOff[a] + i] = (uint8_t)(cd[srcOff[a] + i] & 0xFF);
        tb[dstOff[a] + i] = (uint16_t)(cd[srcOff[a] + i] >> 8);
      }
  }
  return rc;
}

// Shared front half of bgpu_stage_histograms / bgpu_stage_store_metablock: one metablock = the whole input, roles from
// the caller's commands, symbol streams (part A), then the caller's block splits and context mode in place of parts B / C,
// then the per-context histograms (part D1) — the same host phases and kernels run_backend_full goes through.
static int stage_full_front(bgpu_ctx* c, const uint8_t* in, size_t n, const void* cmds_in, size_t ncmd, int is_last, int mode,
                            unsigned start_nbits, unsigned start_value, const uint8_t* prev2, const bgpu_mb_model* model,
                            Job& job, EncParams& P, FullBackend& F, uint8_t** d_in_out) {
  if (!c || !in || !cmds_in || !model || n == 0 || n > ((size_t)1 << 22) || ncmd == 0) return BGPU_E_INVALID;
  P = make_params(9, 24, mode);
  P.k.sh = 1;  // caller-supplied commands may hold 2- and 3-byte copies (Zopfli)
  CK(cudaSetDevice(c->device));
  const size_t lead = prev2 ? 256 : 0;
  bgpu_tuning saved = c->tune; c->tune.shard_bytes = 1u << 22; c->tune.segment_bytes = 0xFFFFFFFFu; c->tune.metablock_bytes = 0;
  size_t lens[1] = {n}; uint32_t pf[1] = {(prev2 ? 0u : BGPU_PART_FIRST) | BGPU_PART_LAST};
  plan_job(c, P, lens, 1, job, pf, lead);
  c->tune = saved;
  ShardD& sh = job.shards[0];
  sh.start_nbits = prev2 ? 0 : start_nbits; sh.start_value = start_value;   // a stage call with prev2 never opens the stream
  job.mbs[0].is_last = is_last ? 1 : 0;
  std::vector<uint8_t> role; std::vector<uint32_t> ins; std::vector<Cmd> cm; uint32_t counts[3];
  if (roles_from_commands((const Cmd*)cmds_in, ncmd, n, P.k.sh, role, ins, cm, counts)) return BGPU_E_INVALID;
  // validate the model against the symbol counts
  uint32_t max_blocks = 1, max_types = 1;
  for (int k = 0; k < 3; k++) {
    const bgpu_block_split& sp = model->split[k];
    if (sp.num_types < 1 || sp.num_types > 256 || sp.num_blocks > 65535u || (sp.num_blocks && (!sp.types || !sp.lengths))) return BGPU_E_INVALID;
    uint64_t sum = 0;
    for (uint32_t b = 0; b < sp.num_blocks; b++) { if (sp.types[b] >= sp.num_types || sp.lengths[b] == 0) return BGPU_E_INVALID; sum += sp.lengths[b]; }
    // block-splitter.ts:419-421 counts the distance symbol of a trailing insert-only command whose prefix is >= 128
    // although no distance is ever coded for it (metablock.ts:619, 741): the reference's last distance block is then
    // one symbol longer than what the emit loop consumes; its declared length is what the header codes.
    const Cmd& lastc = ((const Cmd*)cmds_in)[ncmd - 1];
    const bool phantom = k == 2 && (lastc.copyLen & 0x1FFFFFF) == 0 && lastc.cmdPrefix >= 128;
    if (sp.num_blocks && sum != counts[k] && !(phantom && sum == (uint64_t)counts[k] + 1)) return BGPU_E_INVALID;
    if (sp.num_types == 1 ? sp.num_blocks > 1 : sp.num_blocks < 2) return BGPU_E_INVALID;   // one type <=> at most one block (no split code)
    max_blocks = std::max(max_blocks, sp.num_blocks); max_types = std::max(max_types, sp.num_types);
  }
  if (model->context_mode > 3) return BGPU_E_INVALID;
  int rc;
  if ((rc = c->in.reserve(job.in_bytes + IN_PAD))) return rc;
  if ((rc = reserve_front(c, job, P))) return rc;
  timing_begin(c);
  uint8_t* d_in = c->in.as<uint8_t>();
  if (lead) {
    CK(cudaMemsetAsync(d_in, 0, lead, c->stream));
    CK(cudaMemcpyAsync(d_in + lead - 2, prev2, 2, cudaMemcpyHostToDevice, c->stream));
  }
  CK(cudaMemcpyAsync(d_in + lead, in, n, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(d_in + lead + n, 0, job.in_bytes + IN_PAD - lead - n, c->stream));
  const size_t slots = (n >> P.k.sh) + 2;
  CK(cudaMemsetAsync(c->role.p, 0, lead, c->stream));
  CK(cudaMemcpyAsync(c->role.as<uint8_t>() + lead, role.data(), n + IN_PAD - lead, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(c->ins.as<uint32_t>() + (lead >> P.k.sh), ins.data(), slots * 4, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(c->cmds.as<Cmd>() + (lead >> P.k.sh), cm.data(), slots * sizeof(Cmd), cudaMemcpyHostToDevice, c->stream));
  if ((rc = upload_plan(c, job))) return rc;
  if ((rc = full_setup(c, job, P, F, max_blocks + 1, max_types))) return rc;
  if ((rc = full_symbols(c, P, d_in, F, model->context_mode))) return rc;
  // the caller's block lists and context mode take the place of full_split (parts B and C)
  const MbModel& M = F.mp.models[0];
  for (int k = 0; k < 3; k++) {
    const bgpu_block_split& sp = model->split[k];
    const uint32_t nb = sp.num_blocks ? sp.num_blocks : (counts[k] ? 1u : 0u);   // lengths may sum to counts[k] + 1 (phantom distance, above)
    std::vector<uint32_t> bs(nb + 1, 0); std::vector<uint8_t> bt(std::max<uint32_t>(nb, 1), 0);
    for (uint32_t b = 0; b < nb; b++) { bs[b + 1] = bs[b] + (sp.num_blocks ? sp.lengths[b] : counts[k]); bt[b] = sp.num_blocks ? sp.types[b] : 0; }
    const uint32_t st3[3] = {0, nb, sp.num_types};
    CK(cudaMemcpyAsync(M.blk_start[k], bs.data(), (nb + 1) * 4, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(M.blk_type[k], bt.data(), bt.size(), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(M.sstate[k], st3, sizeof st3, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));   // the staging vectors die with this iteration
  }
  const uint32_t zero = 0;
  CK(cudaMemcpyAsync(M.cnt + 4, &zero, 4, cudaMemcpyHostToDevice, c->stream));   // never the trivial writer
  CK(cudaStreamSynchronize(c->stream));
  if ((rc = full_ctx_hist(c, F))) return rc;
  *d_in_out = d_in;
  return 0;
}

int bgpu_stage_histograms(bgpu_ctx* c, const uint8_t* in, size_t n, const void* cmds_in, size_t ncmd, int mode, const uint8_t* prev2,
                          const bgpu_mb_model* model, uint32_t* lit_hist, uint32_t* cmd_hist, uint32_t* dist_hist, int* context_mode_out) {
  Job job; EncParams P; FullBackend F; uint8_t* d_in = nullptr;
  int rc = stage_full_front(c, in, n, cmds_in, ncmd, 1, mode, 0, 0, prev2, model, job, P, F, &d_in);
  if (rc) return rc;
  const MbModel& M = F.mp.models[0];
  const uint32_t nt_l = model->split[0].num_types, nt_c = model->split[1].num_types, nt_d = model->split[2].num_types, da = P.k.dist_alphabet;
  if (lit_hist) CK(cudaMemcpyAsync(lit_hist, M.ctx_hist[LIT], (size_t)nt_l * 64 * 256 * 4, cudaMemcpyDeviceToHost, c->stream));
  if (cmd_hist) CK(cudaMemcpyAsync(cmd_hist, M.ctx_hist[CMD], (size_t)nt_c * ALPHA_CMD * 4, cudaMemcpyDeviceToHost, c->stream));
  std::vector<uint32_t> dh((size_t)nt_d * 4 * ALPHA_DIST_MAX);
  CK(cudaMemcpyAsync(dh.data(), M.ctx_hist[DIST], dh.size() * 4, cudaMemcpyDeviceToHost, c->stream));
  uint32_t cm_mode = 0;
  CK(cudaMemcpyAsync(&cm_mode, M.cnt + 3, 4, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  timing_end(c);
  if (dist_hist) for (size_t r = 0; r < (size_t)nt_d * 4; r++) memcpy(dist_hist + r * da, dh.data() + r * ALPHA_DIST_MAX, da * 4);
  if (context_mode_out) *context_mode_out = (int)cm_mode;
  return 0;
}

int bgpu_stage_store_metablock(bgpu_ctx* c, const uint8_t* in, size_t n, const void* cmds_in, size_t ncmd, int is_last, int mode,
                               unsigned start_nbits, unsigned start_value, const uint8_t* prev2, const bgpu_mb_model* model,
                               uint8_t* out, size_t out_cap, size_t* out_len) {
  if (!out_len || !model || !model->literal_context_map || !model->distance_context_map) return BGPU_E_INVALID;
  if (model->num_literal_clusters < 1 || model->num_literal_clusters > MAX_CTX_CLUSTERS || model->num_distance_clusters < 1 ||
      model->num_distance_clusters > MAX_CTX_CLUSTERS)
    return BGPU_E_INVALID;
  const size_t nlm = (size_t)model->split[0].num_types * 64, ndm = (size_t)model->split[2].num_types * 4;
  for (size_t i = 0; i < nlm && nlm <= 256 * 64; i++) if (model->literal_context_map[i] >= model->num_literal_clusters) return BGPU_E_INVALID;
  for (size_t i = 0; i < ndm && ndm <= 256 * 4; i++) if (model->distance_context_map[i] >= model->num_distance_clusters) return BGPU_E_INVALID;
  Job job; EncParams P; FullBackend F; uint8_t* d_in = nullptr;
  int rc = stage_full_front(c, in, n, cmds_in, ncmd, is_last, mode, start_nbits, start_value, prev2, model, job, P, F, &d_in);
  if (rc) return rc;
  // the caller's context maps take the place of full_ctx_cluster (part D2): ClusterProb.sym / state[2]
  const ClusterProb& XL = F.mp.cx[0]; const ClusterProb& XD = F.mp.cx[1];
  CK(cudaMemcpyAsync(XL.sym, model->literal_context_map, nlm * 4, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(XL.state + 2, &model->num_literal_clusters, 4, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(XD.sym, model->distance_context_map, ndm * 4, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemcpyAsync(XD.state + 2, &model->num_distance_clusters, 4, cudaMemcpyHostToDevice, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  if ((rc = full_emit(c, P, d_in, F))) return rc;
  uint8_t* outs[1] = {out}; size_t caps[1] = {out_cap};
  rc = collect(c, job, d_in, outs, caps, out_len, nullptr, 1);
  timing_end(c);
  return rc;
}

int bgpu_stage_context_map(bgpu_ctx* c, const uint32_t* map, size_t size, uint32_t num_clusters, uint8_t* bits, size_t cap_bytes, uint64_t* nbits) {
  if (!c || !map || !bits || !nbits || size == 0 || size > 256 * 64 || num_clusters < 1 || num_clusters > 256) return BGPU_E_INVALID;
  for (size_t i = 0; i < size; i++) if (map[i] >= num_clusters) return BGPU_E_INVALID;
  CK(cudaSetDevice(c->device));
  const size_t words = size + 1024;   // <= 8 + 6 bits per entry + one prefix code
  DBuf dm, ds, dout, dn;
  auto done = [&](int r) { dm.release(); ds.release(); dout.release(); dn.release(); return r; };
  int rc;
  if ((rc = dm.reserve(size * 4)) || (rc = ds.reserve(size * 4)) || (rc = dout.reserve(words * 4)) || (rc = dn.reserve(16))) return done(rc);
  cudaError_t e = cudaMemcpyAsync(dm.p, map, size * 4, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(dout.p, 0, words * 4, c->stream);
  uint32_t nb = 0; std::vector<uint8_t> host(words * 4);
  if (e == cudaSuccess) {
    k_stage_context_map<<<1, 32, 0, c->stream>>>(dm.as<uint32_t>(), (uint32_t)size, num_clusters, ds.as<uint32_t>(), dout.as<uint32_t>(), dn.as<uint32_t>());
    e = cudaMemcpyAsync(&nb, dn.p, 4, cudaMemcpyDeviceToHost, c->stream);
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(host.data(), dout.p, words * 4, cudaMemcpyDeviceToHost, c->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) { g_cuda_err = cudaGetErrorString(e); return done(BGPU_E_CUDA); }
  *nbits = nb;
  const size_t need = ((size_t)nb + 7) >> 3;
  if (need > cap_bytes) return done(BGPU_E_OUTPUT_TOO_SMALL);
  memcpy(bits, host.data(), need);
  return done(0);
}

int bgpu_stage_huffman(bgpu_ctx* c, const uint32_t* hist, int count, int alphabet, uint8_t* depths, uint16_t* bits, uint8_t* tree_bits,
                       uint32_t* tree_nbits) {
  if (!c || !hist || count <= 0 || alphabet < 2 || alphabet > ALPHA_CMD) return BGPU_E_INVALID;
  CK(cudaSetDevice(c->device));
  int rc; size_t total = (size_t)count * alphabet;
  DBuf dh, dc, dt, dn, dm;
  auto done = [&](int r) { dh.release(); dc.release(); dt.release(); dn.release(); dm.release(); return r; };
  if ((rc = dh.reserve(total * 4)) || (rc = dc.reserve(total * 4)) || (rc = dt.reserve((size_t)count * TREE_WORDS * 4)) ||
      (rc = dn.reserve((size_t)count * 4)) || (rc = dm.reserve(16)))
    return done(rc);
  int meta[2] = {alphabet, 0};
  cudaError_t e = cudaMemcpy(dh.p, hist, total * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(dm.p, meta, sizeof meta, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    k_huff<<<count, 128, 0, c->stream>>>(dh.as<uint32_t>(), alphabet, 1, dm.as<int>(), dm.as<int>() + 1, dc.as<uint32_t>(), dt.as<uint32_t>(), dn.as<uint32_t>());
    e = cudaStreamSynchronize(c->stream);
  }
  std::vector<uint32_t> cd(total);
  if (e == cudaSuccess) e = cudaMemcpy(cd.data(), dc.p, total * 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && tree_bits) e = cudaMemcpy(tree_bits, dt.p, (size_t)count * TREE_WORDS * 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess && tree_nbits) e = cudaMemcpy(tree_nbits, dn.p, (size_t)count * 4, cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) { g_cuda_err = cudaGetErrorString(e); return done(BGPU_E_CUDA); }
  for (size_t i = 0; i < total; i++) { if (depths) depths[i] = (uint8_t)(cd[i] & 0xFF); if (bits) bits[i] = (uint16_t)(cd[i] >> 8); }
  return done(0);
}

int bgpu_stage_h10_matches(bgpu_ctx* c, const uint8_t* in, size_t n, int lgwin, uint32_t* offsets, uint32_t* dist, uint32_t* len, size_t cap,
                           size_t* total) {
  if (!c || !in || !offsets || !total || n == 0 || n > ((size_t)1 << 22)) return BGPU_E_INVALID;
  EncParams P = make_params(11, lgwin, 0);
  CK(cudaSetDevice(c->device));
  bgpu_tuning saved = c->tune; c->tune.shard_bytes = 1u << 22;
  Job job; size_t lens[1] = {n}; plan_job(c, P, lens, 1, job);
  c->tune = saved;
  if (job.shards.size() != 1) return BGPU_E_INVALID;
  int rc;
  if ((rc = c->in.reserve(job.in_bytes + IN_PAD))) return rc;
  timing_begin(c);
  CK(cudaMemcpyAsync(c->in.p, in, n, cudaMemcpyHostToDevice, c->stream));
  CK(cudaMemsetAsync(c->in.as<uint8_t>() + n, 0, job.in_bytes + IN_PAD - n, c->stream));
  if ((rc = upload_plan(c, job))) return rc;
  uint32_t capu = 0;
  if ((rc = run_h10_retry(c, job, c->in.as<uint8_t>(), &capu))) return rc;
  const size_t N = job.in_bytes + IN_PAD;
  if ((rc = c->h10_full.reserve(N * 4 * capu))) return rc;
  H10Out O{c->h10_m.as<uint32_t>(), c->h10_off.as<uint32_t>(), c->h10_cnt.as<uint8_t>(), c->h10_cursor.as<uint32_t>(), c->h10_cursor.as<uint32_t>() + 1, capu};
  k_h10_extend<<<dim3((unsigned)((n + 255) / 256), 1), 256, 0, c->stream>>>(c->in.as<uint8_t>(), c->shards.as<ShardD>(), O, c->h10_full.as<uint32_t>());
  c->launches++; stage_mark(c, "k_h10_extend");
  uint32_t used = 0;
  CK(cudaMemcpyAsync(&used, c->h10_cursor.p, 4, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  std::vector<uint32_t> m(used), fl(used), off(n); std::vector<uint8_t> cnt(n);
  if (used) {
    CK(cudaMemcpyAsync(m.data(), c->h10_m.p, (size_t)used * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(fl.data(), c->h10_full.p, (size_t)used * 4, cudaMemcpyDeviceToHost, c->stream));
  }
  CK(cudaMemcpyAsync(off.data(), c->h10_off.p, n * 4, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaMemcpyAsync(cnt.data(), c->h10_cnt.p, n, cudaMemcpyDeviceToHost, c->stream));
  CK(cudaStreamSynchronize(c->stream));
  timing_end(c);
  size_t k = 0;
  for (size_t i = 0; i < n; i++) {
    offsets[i] = (uint32_t)k;
    for (uint32_t j = 0; j < cnt[i]; j++) {
      const uint32_t e = m[off[i] + j], L = fl[off[i] + j];
      if (!L) continue;
      if (k < cap && dist && len) { dist[k] = e & 0xFFFFFFu; len[k] = L; }
      k++;
    }
  }
  offsets[n] = (uint32_t)k;
  *total = k;
  return k > cap ? BGPU_E_OUTPUT_TOO_SMALL : 0;
}

int bgpu_stage_timings(bgpu_ctx* c, const char** names, float* ms, int cap) {
  if (!c) return 0;
  for (int i = 0; i < c->n_stage && i < cap; i++) { if (names) names[i] = c->names[i]; if (ms) ms[i] = c->ms[i]; }
  return c->n_stage;
}
int bgpu_last_launch_count(bgpu_ctx* c) { return c ? c->launches : 0; }
uint64_t bgpu_last_counter(bgpu_ctx* c, int which) { return (c && which >= 0 && which < 8) ? c->counters[which] : 0; }

int bgpu_decode(bgpu_ctx* c, const uint8_t* stream, size_t n, uint8_t* out, size_t out_cap, size_t* out_len) {
  if (!c || !stream || !out_len || n == 0 || (out_cap && !out)) return BGPU_E_INVALID;
  CK(cudaSetDevice(c->device));
  int rc;
  if ((rc = ensure_lut(c))) return rc;
  DBuf ds, dout;
  auto done = [&](int r) { ds.release(); dout.release(); return r; };
  if ((rc = ds.reserve(n + 16)) || (rc = dout.reserve(out_cap + 16)) || (rc = c->dec_arena.reserve(sizeof(DecArena))) ||
      (rc = c->dec_results.reserve(sizeof(DecResult))) || (rc = c->dec_jobs.reserve(sizeof(DecJob))))
    return done(rc);
  DecJob J; J.stream_off = 0; J.stream_len = n; J.out_off = 0; J.out_len = 0; J.flags = 3; J.pad = 0;
  DecResult R; memset(&R, 0, sizeof R);
  cudaError_t e = cudaMemcpyAsync(ds.p, stream, n, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(c->dec_jobs.p, &J, sizeof J, cudaMemcpyHostToDevice, c->stream);
  if (e == cudaSuccess) {
    k_decode_shards<<<1, 32, 0, c->stream>>>(ds.as<uint8_t>(), c->dec_jobs.as<DecJob>(), 1, dout.as<uint8_t>(), nullptr, out_cap, c->dec_arena.as<DecArena>(), c->dec_results.as<DecResult>());
    e = cudaMemcpyAsync(&R, c->dec_results.p, sizeof R, cudaMemcpyDeviceToHost, c->stream);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) { g_cuda_err = cudaGetErrorString(e); return done(BGPU_E_CUDA); }
  *out_len = R.produced;
  if (R.status == BGPU_DEC_OUTPUT) return done(BGPU_E_OUTPUT_TOO_SMALL);
  if (R.status != BGPU_DEC_OK) { char b[128]; snprintf(b, sizeof b, "decode status %u at output offset %llu", R.status, (unsigned long long)R.where); g_cuda_err = b; return done(BGPU_E_DECODE); }
  if (R.produced) { e = cudaMemcpy(out, dout.p, R.produced, cudaMemcpyDeviceToHost); if (e != cudaSuccess) { g_cuda_err = cudaGetErrorString(e); return done(BGPU_E_CUDA); } }
  return done(0);
}

}  // extern "C"
#include "woff2_host.inc"
extern "C" {

const char* bgpu_strerror(int code) {
  switch (code) {
    case 0: return "ok";
    case BGPU_E_INVALID: return "invalid argument";
    case BGPU_E_OUTPUT_TOO_SMALL: return "output buffer too small";
    case BGPU_E_UNSUPPORTED: return "quality not supported by the GPU hot path";
    case BGPU_E_NO_DEVICE: return "no CUDA device";
    case BGPU_E_CUDA: return "CUDA error";
    case BGPU_E_NOMEM: return "out of device memory";
    case BGPU_E_STATE: return "stream already finished";
    case BGPU_E_VERIFY: return "on-device verification failed: a shard does not decode back to its input";
    case BGPU_E_DECODE: return "corrupt or unsupported Brotli stream";
    default: return "unknown error";
  }
}
const char* bgpu_last_cuda_error(void) { return g_cuda_err.c_str(); }
int bgpu_version(void) { return 0x000100; }

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------
// bgpu_decode_noalloc: the on-device decoder for callers that may not allocate (include/brotligpu.h).  No
// counterpart in the reference.  The decoder's scratch memory (its
// per-stream table arena, the job descriptors and the results) lives in ONE file-scope device array instead of the
// context's cudaMalloc'ed buffers.  Every GPU that loads the library holds the array until the process exits.
//
// g_noalloc_scratch as written in HOST code is only a host-side stand-in for the symbol: its device address is taken
// with cudaGetSymbolAddress, once per device with that device current (noalloc_setup), and that address is what the
// kernel receives.  All calls on a device share the array and it keeps the previous call's contents: the job / result
// header is cleared at the start of every call (the arenas need no clearing: dec_read_code / dec_context_map write
// every table entry before dec_piece reads it), and two calls that overlap on the GPU race on it.
namespace {

constexpr size_t kNoallocArenaStride = (sizeof(DecArena) + 255) & ~(size_t)255;
constexpr size_t kNoallocHeader = 1024;   // NoallocJob[8] at 0, DecResult[8] at 512
constexpr size_t kNoallocScratchBytes = kNoallocHeader + (size_t)BGPU_NOALLOC_MAX_STREAMS * kNoallocArenaStride;
static_assert(kNoallocScratchBytes <= (size_t)BGPU_NOALLOC_SCRATCH_BYTES, "BGPU_NOALLOC_SCRATCH_BYTES in brotligpu.h is the stated limit");

static __device__ __align__(256) unsigned char g_noalloc_scratch[BGPU_NOALLOC_SCRATCH_BYTES];

struct NoallocJob { const uint8_t* stream; uint64_t stream_len; uint8_t* out; uint64_t out_cap; };
static_assert(sizeof(NoallocJob) * BGPU_NOALLOC_MAX_STREAMS <= 512 && sizeof(DecResult) * BGPU_NOALLOC_MAX_STREAMS <= 512, "header layout");

// grid = streams; lane 0 of each CTA decodes one whole stream (window header .. ISLAST) into its own output.
__global__ void __launch_bounds__(32) k_decode_noalloc(unsigned char* scratch, uint32_t n_streams) {
  const uint32_t j = blockIdx.x;
  if (j >= n_streams || threadIdx.x != 0) return;
  const NoallocJob N = reinterpret_cast<const NoallocJob*>(scratch)[j];
  DecResult* results = reinterpret_cast<DecResult*>(scratch + 512);
  DecArena* arena = reinterpret_cast<DecArena*>(scratch + kNoallocHeader + (size_t)j * kNoallocArenaStride);
  DecJob J; J.stream_off = 0; J.stream_len = N.stream_len; J.out_off = 0; J.out_len = 0; J.flags = 3; J.pad = 0;
  dec_piece(N.stream, J, N.out, nullptr, N.out_cap, *arena, results[j]);
}

constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[kNoallocMaxDevices];   // device address of g_noalloc_scratch, per device (nullptr: not set up)

// Once per device, with that device current: the array's device address and the context-lookup constants.
int noalloc_setup(int device, unsigned char** ptr) {
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    unsigned char* p = nullptr;
    CK(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    uint8_t lut[2048]; build_context_lut(lut);
    CK(cudaMemcpyToSymbol(c_ctx_lut, lut, sizeof lut));
    g_noalloc_ptr[device] = p;
  }
  *ptr = g_noalloc_ptr[device];
  return 0;
}

}  // namespace

extern "C" {

size_t bgpu_noalloc_scratch_bytes(void) { return sizeof g_noalloc_scratch; }

int bgpu_decode_noalloc(int n_streams, const void* const* d_stream, const size_t* stream_len, void* const* d_out, const size_t* out_cap,
                        size_t* out_len, int* status, void* cuda_stream) {
  if (n_streams < 0 || (n_streams && (!d_stream || !stream_len || !d_out || !out_cap || !out_len))) {
    g_cuda_err = "bgpu_decode_noalloc: null argument"; return BGPU_E_INVALID;
  }
  if (n_streams > BGPU_NOALLOC_MAX_STREAMS) {
    char b[160]; snprintf(b, sizeof b, "bgpu_decode_noalloc: %d streams in one call exceed BGPU_NOALLOC_MAX_STREAMS = %d (fixed scratch of %zu bytes)",
                          n_streams, BGPU_NOALLOC_MAX_STREAMS, sizeof g_noalloc_scratch);
    g_cuda_err = b; return BGPU_E_INVALID;
  }
  if (n_streams == 0) return 0;
  NoallocJob jobs[BGPU_NOALLOC_MAX_STREAMS]; DecResult res[BGPU_NOALLOC_MAX_STREAMS];
  for (int k = 0; k < n_streams; k++) {
    if (!d_stream[k] || stream_len[k] == 0 || (out_cap[k] && !d_out[k])) { g_cuda_err = "bgpu_decode_noalloc: empty stream or null buffer"; return BGPU_E_INVALID; }
    jobs[k].stream = (const uint8_t*)d_stream[k]; jobs[k].stream_len = stream_len[k]; jobs[k].out = (uint8_t*)d_out[k]; jobs[k].out_cap = out_cap[k];
    out_len[k] = 0; if (status) status[k] = 0;
  }
  int device = 0;
  CK(cudaGetDevice(&device));
  if (device < 0 || device >= kNoallocMaxDevices) { g_cuda_err = "bgpu_decode_noalloc: device ordinal out of range"; return BGPU_E_INVALID; }
  unsigned char* scratch = nullptr;
  int rc = noalloc_setup(device, &scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)cuda_stream;
  CK(cudaMemsetAsync(scratch, 0, kNoallocHeader, st));
  CK(cudaMemcpyAsync(scratch, jobs, sizeof(NoallocJob) * n_streams, cudaMemcpyHostToDevice, st));
  k_decode_noalloc<<<n_streams, 32, 0, st>>>(scratch, (uint32_t)n_streams);
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(res, scratch + 512, sizeof(DecResult) * n_streams, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  int result = 0;
  for (int k = 0; k < n_streams; k++) {
    out_len[k] = res[k].produced;
    int r = 0;
    if (res[k].status == BGPU_DEC_OUTPUT) r = BGPU_E_OUTPUT_TOO_SMALL;
    else if (res[k].status != BGPU_DEC_OK) r = BGPU_E_DECODE;
    if (status) status[k] = r;
    if (r && !result) {
      result = r;
      char b[160]; snprintf(b, sizeof b, "bgpu_decode_noalloc: stream %d: decode status %u at output offset %llu", k, res[k].status, (unsigned long long)res[k].where);
      g_cuda_err = b;
    }
  }
  return result;
}

}  // extern "C"
