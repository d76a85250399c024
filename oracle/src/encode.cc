// ORACLE (test infrastructure) — restates encode.ts and enc-constants.ts, and exports the
// C entry points the tests and bench.py's cpu_baseline bind with ctypes.
#include "common.h"
#include "bit_writer.h"
#include "command.h"
#include "entropy.h"
#include "metablock.h"
#include "greedy.h"
#include "zopfli.h"
#include <memory>
#include <thread>
#include <atomic>

namespace oracle {

// enc-constants.ts:86-97, 209-233 (only the fields the live code reads)
struct EncoderParams { int mode = 0, quality = 11, lgwin = 22, lgblock = 0; uint32_t npostfix = 0, ndirect = 0; };

// encode.ts:57-71 + enc-constants.ts:111-127 (clamp, never reject; FONT acts at q >= 4)
static EncoderParams MakeParams(int quality, int lgwin, int mode) {
  EncoderParams p;
  p.quality = std::max(0, std::min(11, quality));
  p.lgwin = std::max(10, std::min(24, lgwin));
  p.mode = mode;
  if (p.quality >= 4 && p.mode == 2) { p.npostfix = 1; p.ndirect = 12; }
  // enc-constants.ts:129-147 computeLgBlock (params.lgblock is always 0 on entry)
  if (p.quality == 0 || p.quality == 1) p.lgblock = p.lgwin;
  else if (p.quality < 4) p.lgblock = 14;
  else { p.lgblock = 16; if (p.quality >= 9 && p.lgwin > p.lgblock) p.lgblock = std::min(18, p.lgwin); }
  return p;
}
static inline uint32_t DistanceAlphabetSize(const EncoderParams& p) { return 16 + p.ndirect + (48u << p.npostfix); }  // encode.ts:283-287

// encode.ts:92-103
static std::vector<uint8_t> EncodeEmptyInput() {
  BitWriter w(16); unsigned bits, value;
  EncodeWindowBits(10, &bits, &value);
  w.writeBits(bits, value); w.writeBits(1, 1); w.writeBits(1, 1); w.alignToByte();
  return w.finish();
}
// encode.ts:105-138
static std::vector<uint8_t> EncodeUncompressed(const uint8_t* in, size_t n) {
  BitWriter w(n + 32);
  int lgwin = std::max(10, std::min(24, n <= 1 ? 10 : (int)std::ceil(std::log2((double)n)) + 1));
  unsigned bits, value; EncodeWindowBits(lgwin, &bits, &value);
  w.writeBits(bits, value);
  const size_t maxBlock = ((size_t)1 << 24) - 1;
  size_t pos = 0;
  while (pos < n) {
    size_t bs = std::min(n - pos, maxBlock);
    StoreUncompressedMetaBlock(w, in, pos, bs, pos + bs >= n);
    pos += bs;
  }
  return w.finish();
}

// encode.ts:250-267 trailing-insert merge
static void MergeTrailingInsert(std::vector<Command>& commands, uint32_t lastInsertLen, size_t metablockLen) {
  if (lastInsertLen > 0) {
    if (commands.empty()) commands.push_back(CreateInsertCommand((uint32_t)metablockLen));
    else {
      Command& last = commands.back();
      if (CommandCopyLen(last) == 0) last.insertLen += lastInsertLen;
      else commands.push_back(CreateInsertCommand(lastInsertLen));
    }
  } else if (commands.empty()) commands.push_back(CreateInsertCommand((uint32_t)metablockLen));
}

// Commands for one metablock of a one-shot encode (encode.ts:213-267; encode.ts:157-163 for quality 1), any quality >= 1.
static void OneShotCommands(const Bytes& data, size_t pos, size_t metablockLen, const EncoderParams& p, size_t mask,
                            HashChainHasher* hc, BinaryTreeHasher* bt, int32_t* distCache, std::vector<Command>& commands,
                            SimpleHasher* sh = nullptr) {
  size_t numLiterals = 0; uint32_t lastInsertLen = 0;
  if (p.quality <= 4) CreateBackwardReferences(metablockLen, pos, data, mask, *sh, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen,
                                               p.quality <= 2 ? 4 : 1);
  else if (p.quality >= 11) CreateHqZopfliBackwardReferences(metablockLen, pos, data, mask, *bt, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen);
  else if (p.quality >= 10) CreateZopfliBackwardReferences(metablockLen, pos, data, mask, p.quality, *bt, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen);
  else CreateBackwardReferences(metablockLen, pos, data, mask, *hc, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen);
  MergeTrailingInsert(commands, lastInsertLen, metablockLen);
}

// encode.ts:181-281.  forceTrivial selects storeMetaBlockTrivial for stage-parity tests.
// partFirst / partLast / prev1 / prev2 are NOT in the reference: they restate SURVEY.md §8e's stitching
// rules (header once, unknown distance ring and true literal context after a boundary, byte-aligning
// empty metadata block instead of ISLAST) so that parts encoded independently concatenate to one stream.
// Metablocks after the first (inputs > 16 MiB) take the true previous two bytes as literal context; the
// reference passes none (encode.ts:271-275 -> 0, 0), which desynchronises its own decoder there.
static std::vector<uint8_t> EncodeStandard(const uint8_t* in, size_t n, const EncoderParams& p, bool forceTrivial, MetaBlockStats* stats,
                                           bool partFirst = true, bool partLast = true, uint8_t prev1 = 0, uint8_t prev2 = 0) {
  BitWriter w(std::max<size_t>(1024, (size_t)((double)n * 1.2)));
  unsigned bits, value; EncodeWindowBits(p.lgwin, &bits, &value);
  if (partFirst) w.writeBits(bits, value);
  std::unique_ptr<HashChainHasher> hc; std::unique_ptr<BinaryTreeHasher> bt; std::unique_ptr<SimpleHasher> sh;
  if (p.quality <= 4) sh.reset(CreateSimpleHasher(p.quality));           // encode.ts:149, 192-193
  else if (p.quality <= 9) hc.reset(CreateHashChainHasher(p.quality, p.lgwin));
  else bt.reset(new BinaryTreeHasher(p.lgwin, n));
  int32_t distCache[16] = {4, 11, 15, 16};
  if (!partFirst) for (int i = 0; i < 16; i++) distCache[i] = 0;
  size_t mask = ((size_t)1 << p.lgwin) - 1;
  Bytes data{in, n};
  // quality 1 = encodeFast (encode.ts:140-179): blocks of 1 << lgblock, storeMetaBlockTrivial
  if (p.quality == 1) forceTrivial = true;
  const size_t maxMetablock = p.quality == 1 ? (size_t)1 << p.lgblock : (size_t)1 << 24;
  size_t pos = 0;
  while (pos < n) {
    size_t mlen = std::min(n - pos, maxMetablock);
    bool isLast = partLast && pos + mlen >= n;
    std::vector<Command> commands;
    OneShotCommands(data, pos, mlen, p, mask, hc.get(), bt.get(), distCache, commands, sh.get());
    if (forceTrivial) StoreMetaBlockTrivial(w, in, pos, mlen, mask, isLast, commands, DistanceAlphabetSize(p), p.npostfix, p.ndirect);
    else StoreMetaBlock(w, data, pos, mlen, mask, isLast, commands, DistanceAlphabetSize(p), p.quality, p.npostfix, p.ndirect,
                        pos ? in[pos - 1] : prev1, pos > 1 ? in[pos - 2] : (pos ? prev1 : prev2), stats);
    pos += mlen;
  }
  if (!partLast) { w.writeBits(6, 6); w.alignToByte(); }   // empty metadata block, RFC 7932 section 9.2
  return w.finish();
}

// encode.ts:50-90
static int BrotliEncode(const uint8_t* in, size_t n, int quality, int lgwin, int mode, bool forceTrivial,
                        std::vector<uint8_t>& out, MetaBlockStats* stats = nullptr) {
  EncoderParams p = MakeParams(quality, lgwin, mode);
  if (n == 0) { out = EncodeEmptyInput(); return 0; }
  if (p.quality == 0 || n < 64) { out = EncodeUncompressed(in, n); return 0; }
  out = EncodeStandard(in, n, p, forceTrivial, stats);
  return 0;
}

// encode.ts:290-490 streaming encoder (always trivial metablocks, :471-475)
class BrotliEncoder {
 public:
  EncoderParams p; std::vector<uint8_t> ring; size_t mask; int32_t distCache[16] = {4, 11, 15, 16};
  std::unique_ptr<HashChainHasher> hc; std::unique_ptr<BinaryTreeHasher> bt; std::unique_ptr<SimpleHasher> sh;
  size_t inputPos = 0, lastProcessedPos = 0; bool initialized = false, lastEmitted = false;
  BitWriter w; std::vector<uint8_t> outBuf;
  BrotliEncoder(int quality, int lgwin, int mode) : p(MakeParams(quality, lgwin, mode)), ring((size_t)1 << p.lgwin, 0), mask(ring.size() - 1), w(4096) {}
  void initialize() {
    if (initialized) return;
    unsigned bits, value; EncodeWindowBits(p.lgwin, &bits, &value);
    w.writeBits(bits, value);
    if (p.quality <= 4) sh.reset(CreateSimpleHasher(p.quality));          // encode.ts:343-344 (quality 0 included)
    else if (p.quality <= 9) hc.reset(CreateHashChainHasher(p.quality, p.lgwin));
    else bt.reset(new BinaryTreeHasher(p.lgwin, (size_t)1 << p.lgwin));
    initialized = true;
  }
  void append(const std::vector<uint8_t>& v) { outBuf.insert(outBuf.end(), v.begin(), v.end()); }
  void processBlock(size_t blockLen, bool isLast) {
    size_t pos = lastProcessedPos;
    Bytes data{ring.data(), ring.size()};
    std::vector<Command> commands; size_t numLiterals = 0; uint32_t lastInsertLen = 0;
    if (sh) CreateBackwardReferences(blockLen, pos, data, mask, *sh, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen, p.quality <= 2 ? 4 : 1);
    else if (hc) CreateBackwardReferences(blockLen, pos, data, mask, *hc, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen);
    else if (p.quality >= 11) CreateHqZopfliBackwardReferences(blockLen, pos, data, mask, *bt, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen);
    else CreateZopfliBackwardReferences(blockLen, pos, data, mask, p.quality, *bt, distCache, 0, p.npostfix, p.ndirect, commands, &numLiterals, &lastInsertLen);
    MergeTrailingInsert(commands, lastInsertLen, blockLen);
    StoreMetaBlockTrivial(w, ring.data(), pos, blockLen, mask, isLast, commands, DistanceAlphabetSize(p), p.npostfix, p.ndirect);
    lastProcessedPos += blockLen;
    if (isLast) lastEmitted = true;
    append(w.takeBytes());
  }
  void update(const uint8_t* in, size_t n) {
    initialize();
    for (size_t i = 0; i < n; i++) ring[(inputPos + i) & mask] = in[i];
    inputPos += n;
    size_t blockSize = (size_t)1 << p.lgblock;
    while (inputPos - lastProcessedPos >= blockSize) processBlock(blockSize, false);
  }
  void finish() {
    initialize();
    size_t remaining = inputPos - lastProcessedPos;
    if (remaining > 0) { processBlock(remaining, true); return; }
    if (!lastEmitted) { w.writeBits(1, 1); w.writeBits(1, 1); w.alignToByte(); lastEmitted = true; }
    append(w.takeBytes());
  }
};

}  // namespace oracle

using namespace oracle;

extern "C" {

int oracle_encode(const uint8_t* in, size_t n, int quality, int lgwin, int mode, int force_trivial,
                  uint8_t* out, size_t cap, size_t* out_len) {
  std::vector<uint8_t> v;
  int rc = BrotliEncode(in, n, quality, lgwin, mode, force_trivial != 0, v);
  if (rc) return rc;
  *out_len = v.size();
  if (v.size() > cap) return -1;
  memcpy(out, v.data(), v.size());
  return 0;
}

// stats[0..6] = trivial, litTypes, cmdTypes, distTypes, litClusters, distClusters, contextMode
int oracle_encode_stats(const uint8_t* in, size_t n, int quality, int lgwin, int mode, size_t* out_len, int* stats) {
  std::vector<uint8_t> v; MetaBlockStats s;
  int rc = BrotliEncode(in, n, quality, lgwin, mode, false, v, &s);
  if (rc) return rc;
  *out_len = v.size();
  stats[0] = s.trivial; stats[1] = s.litTypes; stats[2] = s.cmdTypes; stats[3] = s.distTypes;
  stats[4] = s.litClusters; stats[5] = s.distClusters; stats[6] = s.contextMode;
  return 0;
}

// One part of a stream (see EncodeStandard): flags bit 0 = first part, bit 1 = last part; prev = the two
// bytes before `in` in stream order (prev[0] at -2, prev[1] at -1), ignored for a first part.
int oracle_encode_part(const uint8_t* in, size_t n, const uint8_t* prev, int flags, int quality, int lgwin, int mode,
                       uint8_t* out, size_t cap, size_t* out_len) {
  EncoderParams p = MakeParams(quality, lgwin, mode);
  if (p.quality < 1 || n == 0) return -2;
  const bool first = flags & 1, last = flags & 2;
  std::vector<uint8_t> v = EncodeStandard(in, n, p, false, nullptr, first, last, first ? 0 : prev[1], first ? 0 : prev[0]);
  *out_len = v.size();
  if (v.size() > cap) return -1;
  memcpy(out, v.data(), v.size());
  return 0;
}

// Encode `count` independent shards on `threads` host threads (CPU-baseline helper): sizes[i] (may be NULL) = the
// compressed size of shard i, *total_out their sum; the streams themselves are discarded.
int oracle_encode_shards_sizes(const uint8_t* in, size_t n, size_t shard, int quality, int lgwin, int mode, int threads, size_t* total_out, size_t* sizes) {
  size_t count = (n + shard - 1) / shard;
  std::atomic<size_t> next(0), total(0); std::atomic<int> err(0);
  auto work = [&]() {
    for (;;) {
      size_t i = next.fetch_add(1);
      if (i >= count) break;
      size_t off = i * shard, len = std::min(shard, n - off);
      std::vector<uint8_t> v;
      int rc = BrotliEncode(in + off, len, quality, lgwin, mode, false, v);
      if (rc) err = rc;
      total += v.size();
      if (sizes) sizes[i] = v.size();
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < threads; t++) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
  *total_out = total;
  return err;
}
int oracle_encode_shards(const uint8_t* in, size_t n, size_t shard, int quality, int lgwin, int mode, int threads, size_t* total_out) {
  return oracle_encode_shards_sizes(in, n, shard, quality, lgwin, mode, threads, total_out, nullptr);
}

// hash-chains.ts chain part alone ("candidate-exact" stage): for every position p with at
// least 4 bytes left, the result of findLongestMatch with an empty distance cache, all
// earlier positions stored, maxLength = min(n-p, 128), maxBackward = min(p, 4M-16).
int oracle_h5_chain_best(const uint8_t* in, size_t n, int quality, int lgwin, uint32_t* len, uint32_t* dist) {
  std::unique_ptr<HashChainHasher> h(CreateHashChainHasher(quality, lgwin));
  h->numLastDistancesToCheck = 0;
  Bytes data{in, n}; size_t mask = ((size_t)1 << lgwin) - 1;
  int32_t dc[16] = {0};
  for (size_t p = 0; p < n; p++) { len[p] = 0; dist[p] = 0; }
  for (size_t p = 0; p + 4 <= n; p++) {
    HasherSearchResult r;
    h->findLongestMatch(data, mask, dc, p, (uint32_t)std::min<size_t>(n - p, 128), std::min<size_t>(p, ((size_t)1 << 22) - 16), r);
    len[p] = r.len; dist[p] = r.distance;
  }
  return 0;
}

// hash-simple.ts table part alone ("candidate-exact" stage, quality 2..4): findLongestMatch with an empty
// distance cache at every position in order (each call stores its position), same limits as above.
int oracle_simple_table_best(const uint8_t* in, size_t n, int quality, int lgwin, uint32_t* len, uint32_t* dist) {
  std::unique_ptr<SimpleHasher> h(CreateSimpleHasher(quality));
  Bytes data{in, n}; size_t mask = ((size_t)1 << lgwin) - 1;
  int32_t dc[16] = {0};
  for (size_t p = 0; p < n; p++) { len[p] = 0; dist[p] = 0; }
  for (size_t p = 0; p + 4 <= n; p++) {
    HasherSearchResult r;
    h->findLongestMatch(data, mask, dc, p, (uint32_t)std::min<size_t>(n - p, 128), std::min<size_t>(p, ((size_t)1 << 22) - 16), r);
    len[p] = r.len; dist[p] = r.distance;
  }
  return 0;
}

// hash-binary-tree.ts findAllMatches for EVERY position of a single-window input (no skipping of
// positions covered by long matches — that is caller logic, backward-references-hq.ts:520-540):
// offsets[i]..offsets[i+1] index (dist, len) of position i; maxLength = n - i, maxBackward = min(i, 4M-16).
int oracle_h10_all_matches(const uint8_t* in, size_t n, int lgwin, uint32_t* offsets, uint32_t* dist, uint32_t* len, size_t cap, size_t* total) {
  BinaryTreeHasher h(lgwin, n);
  Bytes data{in, n}; size_t mask = ((size_t)1 << lgwin) - 1;
  std::vector<BackwardMatch> m; size_t k = 0;
  for (size_t i = 0; i < n; i++) {
    offsets[i] = (uint32_t)k;
    if (i + 3 >= n) continue;
    h.findAllMatches(data, mask, i, n - i, std::min<size_t>(i, ((size_t)1 << 22) - 16), m);
    for (const BackwardMatch& b : m) { if (k < cap) { dist[k] = b.distance; len[k] = b.length; } k++; }
  }
  offsets[n] = (uint32_t)k;
  *total = k;
  return k > cap ? -1 : 0;
}

// Commands of the one-shot encode of a single-metablock input (encode.ts:213-267).
int oracle_commands(const uint8_t* in, size_t n, int quality, int lgwin, int mode, Command* cmds, size_t cap, size_t* ncmd) {
  EncoderParams p = MakeParams(quality, lgwin, mode);
  if (p.quality < 1 || n == 0 || n > ((size_t)1 << (p.quality == 1 ? p.lgblock : 24))) return -2;
  std::unique_ptr<HashChainHasher> hc; std::unique_ptr<BinaryTreeHasher> bt; std::unique_ptr<SimpleHasher> sh;
  if (p.quality <= 4) sh.reset(CreateSimpleHasher(p.quality));
  else if (p.quality <= 9) hc.reset(CreateHashChainHasher(p.quality, p.lgwin)); else bt.reset(new BinaryTreeHasher(p.lgwin, n));
  int32_t distCache[16] = {4, 11, 15, 16};
  std::vector<Command> commands;
  OneShotCommands(Bytes{in, n}, 0, n, p, ((size_t)1 << p.lgwin) - 1, hc.get(), bt.get(), distCache, commands, sh.get());
  *ncmd = commands.size();
  if (commands.size() > cap) return -1;
  memcpy(cmds, commands.data(), commands.size() * sizeof(Command));
  return 0;
}

// storeMetaBlockTrivial on caller-supplied commands, preceded by `start_nbits` header bits
// (start_value, LSB-first) — the window header for shard 0, nothing for later shards.
// tables (may be NULL): lit/cmd/dist histograms, depths and codes as laid out in TrivialTables.
int oracle_store_trivial(const uint8_t* in, size_t n, const Command* cmds, size_t ncmd, int is_last, int mode_font,
                         unsigned start_nbits, unsigned start_value, uint8_t* out, size_t cap, size_t* out_len, uint64_t* out_bits,
                         TrivialTables* tables) {
  uint32_t npostfix = mode_font ? 1 : 0, ndirect = mode_font ? 12 : 0;
  BitWriter w(n + 1024);
  w.writeBits(start_nbits, start_value);
  std::vector<Command> commands(cmds, cmds + ncmd);
  StoreMetaBlockTrivial(w, in, 0, n, ~(size_t)0, is_last != 0, commands, 16 + ndirect + (48u << npostfix), npostfix, ndirect, tables);
  *out_bits = w.pos;
  std::vector<uint8_t> v = w.finish();
  *out_len = v.size();
  if (v.size() > cap) return -1;
  memcpy(out, v.data(), v.size());
  return 0;
}

// ---- full-metablock stage entry points (metablock.ts:504-761 cut at the model boundary) ----
// Flat model layout shared with the tests: hdr[0] context mode, hdr[1..3] numTypes (literal, command,
// distance), hdr[4..6] numBlocks, hdr[7] literal clusters, hdr[8] distance clusters; types / lengths hold
// three runs of cap_blocks entries; lit_map has numTypes[0] * 64 entries, dist_map numTypes[2] * 4.
static void FlatToModel(const uint32_t* hdr, const uint8_t* types, const uint32_t* lengths, size_t cap, const uint32_t* lit_map,
                        const uint32_t* dist_map, MetaBlockModel& M) {
  BlockSplit* sp[3] = {&M.literalSplit, &M.commandSplit, &M.distanceSplit};
  M.contextMode = (int)hdr[0];
  for (int c = 0; c < 3; c++) {
    sp[c]->numTypes = hdr[1 + c]; sp[c]->numBlocks = hdr[4 + c];
    sp[c]->types.assign(types + c * cap, types + c * cap + hdr[4 + c]);
    sp[c]->lengths.assign(lengths + c * cap, lengths + c * cap + hdr[4 + c]);
  }
  M.numLiteralClusters = hdr[7]; M.numDistanceClusters = hdr[8];
  if (lit_map) M.literalContextMap.assign(lit_map, lit_map + (size_t)hdr[1] * 64);
  if (dist_map) M.distanceContextMap.assign(dist_map, dist_map + (size_t)hdr[3] * 4);
}

// splitBlock + chooseContextMode + histograms + context-map clustering of ONE metablock = the whole input.
// Returns 1 when the reference would store the metablock trivially (no model), -1 when cap_blocks is too small.
int oracle_metablock_model(const uint8_t* in, size_t n, const Command* cmds, size_t ncmd, int quality, int mode_font, int prev1, int prev2,
                           uint32_t* hdr, uint8_t* types, uint32_t* lengths, size_t cap_blocks, uint32_t* lit_map, uint32_t* dist_map) {
  uint32_t npostfix = mode_font ? 1 : 0, ndirect = mode_font ? 12 : 0;
  std::vector<Command> commands(cmds, cmds + ncmd);
  MetaBlockModel M;
  if (!ComputeMetaBlockModel(Bytes{in, n}, 0, n, ~(size_t)0, commands, 16 + ndirect + (48u << npostfix), quality, (uint8_t)prev1, (uint8_t)prev2, M)) return 1;
  const BlockSplit* sp[3] = {&M.literalSplit, &M.commandSplit, &M.distanceSplit};
  hdr[0] = (uint32_t)M.contextMode; hdr[7] = M.numLiteralClusters; hdr[8] = M.numDistanceClusters;
  for (int c = 0; c < 3; c++) {
    hdr[1 + c] = sp[c]->numTypes; hdr[4 + c] = sp[c]->numBlocks;
    if (sp[c]->numBlocks > cap_blocks) return -1;
    for (uint32_t b = 0; b < sp[c]->numBlocks; b++) { types[c * cap_blocks + b] = sp[c]->types[b]; lengths[c * cap_blocks + b] = sp[c]->lengths[b]; }
  }
  memcpy(lit_map, M.literalContextMap.data(), M.literalContextMap.size() * 4);
  memcpy(dist_map, M.distanceContextMap.data(), M.distanceContextMap.size() * 4);
  return 0;
}

// metablock.ts:558-637 alone, under caller-supplied splits and context mode.  lit_hist [numTypes[0] * 64][256],
// cmd_hist [numTypes[1]][704], dist_hist [numTypes[2] * 4][64 | 124].
int oracle_context_histograms(const uint8_t* in, size_t n, const Command* cmds, size_t ncmd, int mode_font, int prev1, int prev2,
                              const uint32_t* hdr, const uint8_t* types, const uint32_t* lengths, size_t cap_blocks,
                              uint32_t* lit_hist, uint32_t* cmd_hist, uint32_t* dist_hist) {
  uint32_t npostfix = mode_font ? 1 : 0, ndirect = mode_font ? 12 : 0, da = 16 + ndirect + (48u << npostfix);
  std::vector<Command> commands(cmds, cmds + ncmd);
  MetaBlockModel M; FlatToModel(hdr, types, lengths, cap_blocks, nullptr, nullptr, M);
  BuildContextHistograms(in, 0, n, ~(size_t)0, commands, da, (uint8_t)prev1, (uint8_t)prev2, M);
  for (size_t i = 0; i < M.literalHistograms.size(); i++) memcpy(lit_hist + i * 256, M.literalHistograms[i].data(), 256 * 4);
  for (size_t i = 0; i < M.commandHistograms.size(); i++) memcpy(cmd_hist + i * 704, M.commandHistograms[i].data(), 704 * 4);
  for (size_t i = 0; i < M.distanceHistograms.size(); i++) memcpy(dist_hist + i * da, M.distanceHistograms[i].data(), da * 4);
  return 0;
}

// metablock.ts:651-760: the full-metablock writer on caller-supplied commands, splits, context mode and context
// maps (histograms and prefix codes are re-derived from them, as storeMetaBlock does), preceded by start_nbits header bits.
int oracle_store_metablock(const uint8_t* in, size_t n, const Command* cmds, size_t ncmd, int is_last, int mode_font,
                           unsigned start_nbits, unsigned start_value, int prev1, int prev2,
                           const uint32_t* hdr, const uint8_t* types, const uint32_t* lengths, size_t cap_blocks,
                           const uint32_t* lit_map, const uint32_t* dist_map, uint8_t* out, size_t cap, size_t* out_len, uint64_t* out_bits) {
  uint32_t npostfix = mode_font ? 1 : 0, ndirect = mode_font ? 12 : 0, da = 16 + ndirect + (48u << npostfix);
  std::vector<Command> commands(cmds, cmds + ncmd);
  MetaBlockModel M; FlatToModel(hdr, types, lengths, cap_blocks, lit_map, dist_map, M);
  BuildContextHistograms(in, 0, n, ~(size_t)0, commands, da, (uint8_t)prev1, (uint8_t)prev2, M);
  BitWriter w(n + n / 2 + 65536);
  w.writeBits(start_nbits, start_value);
  EmitMetaBlock(w, in, 0, n, ~(size_t)0, is_last != 0, commands, da, npostfix, ndirect, (uint8_t)prev1, (uint8_t)prev2, M);
  *out_bits = w.pos;
  std::vector<uint8_t> v = w.finish();
  *out_len = v.size();
  if (v.size() > cap) return -1;
  memcpy(out, v.data(), v.size());
  return 0;
}

// encodeContextMap alone (context-map.ts:114-170).
int oracle_encode_context_map(const uint32_t* map, size_t size, uint32_t num_clusters, uint8_t* out, size_t cap, uint64_t* out_bits) {
  BitWriter w(size * 4 + 1024);
  EncodeContextMap(w, map, size, num_clusters);
  *out_bits = w.pos;
  std::vector<uint8_t> v = w.finish();
  if (v.size() > cap) return -1;
  memcpy(out, v.data(), v.size());
  return 0;
}

int oracle_huffman(const uint32_t* hist, size_t n, int tree_limit, uint8_t* depths, uint16_t* bits) {
  CreateHuffmanTree(hist, n, tree_limit, depths, bits);
  return 0;
}

// buildAndStoreHuffmanTree alone: serialised tree bits + the code it defines.
int oracle_build_store_tree(const uint32_t* hist, size_t n, uint8_t* depths, uint16_t* bits, uint8_t* out, size_t cap, uint64_t* out_bits) {
  BitWriter w(1024);
  memset(depths, 0, n); memset(bits, 0, 2 * n);
  BuildAndStoreHuffmanTree(w, hist, n, depths, bits);
  *out_bits = w.pos;
  std::vector<uint8_t> v = w.finish();
  if (v.size() > cap) return -1;
  memcpy(out, v.data(), v.size());
  return 0;
}

// test_h10_repairs_are_needed: 0 = repaired (default), bit 0 / bit 1 = the TS behaviours (zopfli.h g_h10_compat)
void oracle_set_h10_compat(int bits) { g_h10_compat = bits; }

const uint8_t* oracle_context_lut(void) { return ContextLookupTable(); }
const double* oracle_log2_table(void) { return kLog2Table(); }

void* oracle_stream_create(int quality, int lgwin, int mode) {
  return new BrotliEncoder(quality, lgwin, mode);
}
// Returns the bytes completed by this call; the pointer is valid until the next call.
int oracle_stream_update(void* s, const uint8_t* in, size_t n, const uint8_t** out, size_t* out_len) {
  BrotliEncoder* e = (BrotliEncoder*)s;
  e->outBuf.clear(); e->update(in, n);
  *out = e->outBuf.data(); *out_len = e->outBuf.size();
  return 0;
}
int oracle_stream_finish(void* s, const uint8_t** out, size_t* out_len) {
  BrotliEncoder* e = (BrotliEncoder*)s;
  e->outBuf.clear(); e->finish();
  *out = e->outBuf.data(); *out_len = e->outBuf.size();
  return 0;
}
void oracle_stream_destroy(void* s) { delete (BrotliEncoder*)s; }

}  // extern "C"
