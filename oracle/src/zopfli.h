// ORACLE (test infrastructure) — restates hash-binary-tree.ts:19-227, zopfli-cost-model.ts and
// backward-references-hq.ts.
#pragma once
#include "common.h"
#include "command.h"
#include "greedy.h"
#include <limits>

namespace oracle {

struct BackwardMatch { uint32_t distance, length; };

// hash-binary-tree.ts:19-227.  DEVIATION: the TS constructor leaves `buckets`
// zero-filled (only reset() writes the invalid marker and the factory never calls it), so
// position 0 becomes a node of every bucket's tree at once; its two forest slots are then shared
// between unrelated trees, the BST ordering breaks, and the "known common prefix" shortcut
// (:100-107) reports match lengths that were never compared -> corrupt streams (observed on
// bench/fixtures/enc-ttf.bin at q10/q11).  The restatement initialises buckets to invalidPos,
// as reset() and the C encoder do.
// Reproducer switches (tests/test_oracle.py::test_h10_repairs_are_needed): bit 0 = leave the buckets zero-initialised as the
// TS constructor does (hash-binary-tree.ts:30-45: only reset() fills them with the invalid marker, and nothing calls it);
// bit 1 = let the walk simply end when the 64-level budget runs out (:86), without closing the two dangling links.
// Default 0 = both repairs on.  Why they are needed: the length of a tree match is only COMPARED from
// curLen = min(bestLenLeft, bestLenRight) on (:103-109); the first curLen bytes are taken on trust from the search-tree
// order.  A stale link (bit 1) or a node reached through position 0's links, which belong to another bucket's tree and are
// then re-rooted into this one (bit 0), breaks that order, so lengths come out longer than the true common prefix and the
// Zopfli pass emits copies the decoder resolves to different bytes.
static int g_h10_compat = 0;
class BinaryTreeHasher {
 public:
  static const int kBucketBits = 17;
  static const uint32_t kMaxTreeCompLength = 128, kMaxTreeSearchDepth = 64;
  std::vector<uint32_t> buckets, forest; uint32_t windowMask, invalidPos;
  BinaryTreeHasher(int lgwin, size_t inputSize)
      : buckets((size_t)1 << kBucketBits, (g_h10_compat & 1) ? 0u : 0u - ((1u << lgwin) - 1)), windowMask((1u << lgwin) - 1), invalidPos(0u - ((1u << lgwin) - 1)) {
    size_t numNodes = std::min<size_t>(inputSize, (size_t)1 << lgwin);
    forest.assign(2 * numNodes, invalidPos);
  }
  size_t leftIdx(size_t pos) const { return 2 * (pos & windowMask); }
  size_t rightIdx(size_t pos) const { return 2 * (pos & windowMask) + 1; }
  // JS reads beyond the typed array give `undefined`; writes are dropped.
  uint32_t fget(size_t i, bool* undef) const { if (i < forest.size()) return forest[i]; *undef = true; return 0; }
  void fset(size_t i, uint32_t v) { if (i < forest.size()) forest[i] = v; }

  // :57-153
  void storeAndFindMatches(const Bytes& data, size_t curIx, size_t mask, size_t maxLength, size_t maxBackward,
                           std::vector<BackwardMatch>* matches) {
    size_t cur = curIx & mask;
    size_t maxCompLen = std::min<size_t>(maxLength, kMaxTreeCompLength);
    bool reroot = maxLength >= kMaxTreeCompLength;
    uint32_t key = HashBytes4(data, cur, kBucketBits);
    uint32_t prevIx = buckets[key];
    size_t nodeLeft = leftIdx(curIx), nodeRight = rightIdx(curIx);
    size_t bestLenLeft = 0, bestLenRight = 0;
    size_t bestLen = matches ? 1 : 0;
    if (reroot) buckets[key] = (uint32_t)curIx;
    // DEVIATION: the TS loop simply ends when the 64-level budget runs out (:86), leaving
    // forest[nodeLeft]/forest[nodeRight] pointing at stale subtrees; that breaks the BST ordering
    // the common-prefix shortcut relies on and yields unverified match lengths (corrupt streams on
    // bench/fixtures/enc-*.bin at q10/q11).  As in the C encoder, both dangling links are closed.
    for (uint32_t depthRemaining = kMaxTreeSearchDepth;; depthRemaining--) {
      if (depthRemaining == 0) { if (reroot && !(g_h10_compat & 2)) { fset(nodeLeft, invalidPos); fset(nodeRight, invalidPos); } break; }
      if (prevIx == invalidPos) { if (reroot) { fset(nodeLeft, invalidPos); fset(nodeRight, invalidPos); } break; }
      int64_t backward = (int64_t)curIx - (int64_t)prevIx;
      size_t prevMasked = prevIx & mask;
      if (backward <= 0 || backward > (int64_t)maxBackward) { if (reroot) { fset(nodeLeft, invalidPos); fset(nodeRight, invalidPos); } break; }
      size_t curLen = std::min(bestLenLeft, bestLenRight);
      size_t len = curLen;
      { const uint8_t* a = data.p + cur; const uint8_t* b = data.p + prevMasked;
        while (len < maxLength && a[len] == b[len]) len++; }
      if (matches && len > bestLen) { bestLen = len; matches->push_back({(uint32_t)backward, (uint32_t)len}); }
      if (len >= maxCompLen) {
        if (reroot) { bool u = false; uint32_t l = fget(leftIdx(prevIx), &u), r = fget(rightIdx(prevIx), &u); fset(nodeLeft, l); fset(nodeRight, r); }
        break;
      }
      if (data.at(cur + len) > data.at(prevMasked + len)) {
        bestLenLeft = len;
        if (reroot) fset(nodeLeft, prevIx);
        nodeLeft = rightIdx(prevIx);
        prevIx = forest[nodeLeft];
      } else {
        bestLenRight = len;
        if (reroot) fset(nodeRight, prevIx);
        nodeRight = leftIdx(prevIx);
        prevIx = forest[nodeRight];
      }
    }
  }
  // :156-227.  Result is sorted by strictly increasing length by construction.
  void findAllMatches(const Bytes& data, size_t mask, size_t curIx, size_t maxLength, size_t maxBackward,
                      std::vector<BackwardMatch>& matches) {
    matches.clear();
    size_t cur = curIx & mask;
    size_t bestLen = 1;
    size_t stop = curIx > 64 ? curIx - 64 : 0;
    for (size_t i = curIx - 1; curIx > 0 && i > stop && bestLen <= 2; i--) {
      size_t backward = curIx - i;
      if (backward > maxBackward) break;
      size_t prev = i & mask;
      if (data.at(cur) != data.at(prev) || data.at(cur + 1) != data.at(prev + 1)) continue;
      size_t len = 0;
      while (len < maxLength && data.p[prev + len] == data.p[cur + len]) len++;
      if (len > bestLen) { bestLen = len; matches.push_back({(uint32_t)backward, (uint32_t)len}); }
    }
    if (bestLen < maxLength) {
      std::vector<BackwardMatch> tree;
      storeAndFindMatches(data, curIx, mask, maxLength, maxBackward, &tree);
      for (const BackwardMatch& m : tree) if (m.length > bestLen) { bestLen = m.length; matches.push_back(m); }
    } else {
      storeAndFindMatches(data, curIx, mask, maxLength, maxBackward, nullptr);
    }
    std::stable_sort(matches.begin(), matches.end(), [](const BackwardMatch& a, const BackwardMatch& b) { return a.length < b.length; });
  }
};

// zopfli-cost-model.ts:12-206.  Float32Array stores are modelled with `float`; arithmetic is double.
static const double kInfinityCost = 1.7e38;
class ZopfliCostModel {
 public:
  std::vector<float> costCmd, costDist, literalCosts; double minCostCmd = kInfinityCost; size_t numBytes;
  ZopfliCostModel(size_t numBytes_, size_t distAlphabet) : costCmd(704, 0.f), costDist(distAlphabet, 0.f), literalCosts(numBytes_ + 2, 0.f), numBytes(numBytes_) {}
  void cumulate(const double* perByte, const float* perByteF, size_t position, const uint8_t* rb, size_t mask) {
    literalCosts[0] = 0;
    double carry = 0;
    for (size_t i = 0; i < numBytes; i++) {
      uint8_t b = rb[(position + i) & mask];
      carry += perByte ? perByte[b] : (double)perByteF[b];
      literalCosts[i + 1] = (float)((double)literalCosts[i] + carry);
      carry -= (double)literalCosts[i + 1] - (double)literalCosts[i];
    }
  }
  // :29-65, :163-189
  void setFromLiteralCosts(size_t position, const uint8_t* rb, size_t mask) {
    double costs[256]; uint32_t histogram[256] = {0};
    for (size_t i = 0; i < numBytes; i++) histogram[rb[(position + i) & mask]]++;
    double log2total = FastLog2((double)numBytes);
    for (int i = 0; i < 256; i++) {
      if (histogram[i] == 0) costs[i] = log2total + 2;
      else { costs[i] = log2total - FastLog2(histogram[i]); if (costs[i] < 1) costs[i] = 1; }
    }
    cumulate(costs, nullptr, position, rb, mask);
    for (int i = 0; i < 704; i++) costCmd[i] = (float)FastLog2(11 + i);
    for (size_t i = 0; i < costDist.size(); i++) costDist[i] = (float)FastLog2(20 + (double)i);
    minCostCmd = FastLog2(11);
  }
  // :126-161
  static void setCostFromHistogram(const uint32_t* histogram, size_t n, bool isLiteral, float* cost) {
    double sum = 0;
    for (size_t i = 0; i < n; i++) sum += histogram[i];
    double log2sum = FastLog2(sum);
    double missingSum = sum;
    if (!isLiteral) for (size_t i = 0; i < n; i++) if (histogram[i] == 0) missingSum++;
    double missingCost = FastLog2(missingSum) + 2;
    for (size_t i = 0; i < n; i++) {
      if (histogram[i] == 0) cost[i] = (float)missingCost;
      else { cost[i] = (float)(log2sum - FastLog2(histogram[i])); if (cost[i] < 1) cost[i] = 1; }
    }
  }
  // :68-124
  void setFromCommands(size_t position, const uint8_t* rb, size_t mask, const std::vector<Command>& commands, uint32_t lastInsertLen) {
    std::vector<uint32_t> hl(256, 0), hc(704, 0), hd(costDist.size(), 0); float costLiteral[256];
    size_t pos = position - lastInsertLen;
    for (const Command& cmd : commands) {
      uint32_t insLen = cmd.insertLen, copyLen = CommandCopyLen(cmd);
      hc[cmd.cmdPrefix]++;
      if (cmd.cmdPrefix >= 128) hd[cmd.distPrefix & 0x3FF]++;
      for (uint32_t j = 0; j < insLen; j++) hl[rb[(pos + j) & mask]]++;
      pos += insLen + copyLen;
    }
    setCostFromHistogram(hl.data(), 256, true, costLiteral);
    setCostFromHistogram(hc.data(), 704, false, costCmd.data());
    setCostFromHistogram(hd.data(), hd.size(), false, costDist.data());
    minCostCmd = kInfinityCost;
    for (int i = 0; i < 704; i++) if ((double)costCmd[i] < minCostCmd) minCostCmd = costCmd[i];
    cumulate(nullptr, costLiteral, position, rb, mask);
  }
  double getCommandCost(uint32_t c) const { return costCmd[c]; }
  // :195-197.  The TS indexes a 544-entry Float32Array with (dist+15)&0x3FF on the explicit
  // match path (backward-references-hq.ts:355-357); indices >= 544 read `undefined` and turn
  // the candidate's cost into NaN, which never wins a `<` comparison.  Reproduced with NaN.
  double getDistanceCost(uint32_t d) const { return d < costDist.size() ? (double)costDist[d] : std::numeric_limits<double>::quiet_NaN(); }
  double getLiteralCosts(size_t from, size_t to) const { return (double)literalCosts[to] - (double)literalCosts[from]; }
};

// backward-references-hq.ts:31-37
struct ZopfliNode { uint32_t length = 1, distance = 0, dcodeInsertLength = 0; double cost = kInfinityCost; uint32_t shortcut = 0; };
static inline uint32_t NodeCopyLength(const ZopfliNode& n) { return n.length & 0x1FFFFFF; }
static inline uint32_t NodeLengthCode(const ZopfliNode& n) { return NodeCopyLength(n) + 9 - (n.length >> 25); }
static inline uint32_t NodeInsertLength(const ZopfliNode& n) { return n.dcodeInsertLength & 0x7FFFFFF; }
static inline uint32_t NodeDistanceCode(const ZopfliNode& n) {
  uint32_t sc = n.dcodeInsertLength >> 27;
  return sc == 0 ? n.distance + 16 - 1 : sc - 1;
}
static inline uint32_t NodeCommandLength(const ZopfliNode& n) { return NodeCopyLength(n) + NodeInsertLength(n); }

// :93-149
struct PosData { size_t pos = 0; int32_t distanceCache[4] = {0, 0, 0, 0}; double costdiff = kInfinityCost, cost = kInfinityCost; };
struct StartPosQueue {
  PosData q[8]; uint32_t idx = 0;
  size_t size() const { return std::min<uint32_t>(idx, 8); }
  const PosData& at(size_t k) const { return q[(k - idx) & 7]; }
  void push(const PosData& pd) {
    uint32_t offset = (~idx++) & 7;
    q[offset] = pd;
    size_t len = size();
    for (size_t i = 1; i < len; i++) {
      uint32_t a = (offset + i - 1) & 7, b = (offset + i) & 7;
      if (q[a].costdiff > q[b].costdiff) std::swap(q[a], q[b]);
    }
  }
};

// :151-166
static inline void UpdateZopfliNode(std::vector<ZopfliNode>& nodes, size_t pos, size_t startPos, size_t len, size_t lenCode,
                                    uint32_t dist, uint32_t shortCode, double cost) {
  ZopfliNode& next = nodes[pos + len];
  next.length = (uint32_t)(len | ((len + 9 - lenCode) << 25));
  next.distance = dist;
  next.dcodeInsertLength = (shortCode << 27) | (uint32_t)(pos - startPos);
  next.cost = cost;
}
// :168-190
static inline size_t ComputeMinimumCopyLength(double startCost, const std::vector<ZopfliNode>& nodes, size_t numBytes, size_t pos) {
  double minCost = startCost; size_t len = 2, nextLenBucket = 4, nextLenOffset = 10;
  while (pos + len <= numBytes && nodes[pos + len].cost <= minCost) {
    len++;
    if (len == nextLenOffset) { minCost += 1.0; nextLenOffset += nextLenBucket; nextLenBucket *= 2; }
  }
  return len;
}
// :192-212
static inline uint32_t ComputeDistanceShortcut(size_t blockStart, size_t pos, size_t maxBackwardLimit, size_t gap, const std::vector<ZopfliNode>& nodes) {
  if (pos == 0) return 0;
  size_t cLen = NodeCopyLength(nodes[pos]), iLen = NodeInsertLength(nodes[pos]), dist = nodes[pos].distance;
  if (dist + cLen <= blockStart + pos + gap && dist <= maxBackwardLimit + gap && NodeDistanceCode(nodes[pos]) > 0) return (uint32_t)pos;
  return nodes[pos - cLen - iLen].shortcut;
}
// :214-234.  DEVIATION: the TS fills the tail from startingDistCache[idx-(4-idx)]
// (indices -4,-2,0,2), which desynchronises the encoder's distance ring from the decoder's and
// can corrupt streams; the restatement uses the C rule `dist_cache[idx] = *starting++`.
static inline void ComputeDistanceCache(size_t pos, const int32_t* starting, const std::vector<ZopfliNode>& nodes, int32_t* distCache) {
  int idx = 0; size_t p = nodes[pos].shortcut;
  while (idx < 4 && p > 0) {
    size_t iLen = NodeInsertLength(nodes[p]), cLen = NodeCopyLength(nodes[p]);
    distCache[idx++] = (int32_t)nodes[p].distance;
    p = nodes[p - cLen - iLen].shortcut;
  }
  for (; idx < 4; idx++) distCache[idx] = *starting++;
}
// :239-264
static inline void EvaluateNode(size_t blockStart, size_t pos, size_t maxBackwardLimit, size_t gap, const int32_t* starting,
                                const ZopfliCostModel& model, StartPosQueue& queue, std::vector<ZopfliNode>& nodes) {
  double nodeCost = nodes[pos].cost;
  nodes[pos].shortcut = ComputeDistanceShortcut(blockStart, pos, maxBackwardLimit, gap, nodes);
  if (nodeCost <= model.getLiteralCosts(0, pos)) {
    PosData pd; pd.pos = pos; pd.cost = nodeCost; pd.costdiff = nodeCost - model.getLiteralCosts(0, pos);
    ComputeDistanceCache(pos, starting, nodes, pd.distanceCache);
    queue.push(pd);
  }
}
// :267-382
static inline size_t UpdateNodes(size_t numBytes, size_t blockStart, size_t pos, const Bytes& rb, size_t mask, int quality,
                                 size_t maxBackwardLimit, const int32_t* starting, const std::vector<BackwardMatch>& matches,
                                 const ZopfliCostModel& model, StartPosQueue& queue, std::vector<ZopfliNode>& nodes) {
  static const uint8_t kIdx[16] = {0, 1, 2, 3, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1};
  static const int8_t kOff[16] = {0, 0, 0, 0, -1, 1, -2, 2, -3, 3, -1, 1, -2, 2, -3, 3};
  size_t curIx = blockStart + pos, cur = curIx & mask;
  size_t maxDistance = std::min(curIx, maxBackwardLimit);
  size_t maxLen = numBytes - pos;
  size_t maxZopfliLen = quality <= 10 ? 150 : 325;     // enc-constants.ts:103-105
  size_t maxIters = quality <= 10 ? 1 : 5;             // enc-constants.ts:107-109
  EvaluateNode(blockStart, pos, maxBackwardLimit, 0, starting, model, queue, nodes);
  const PosData& pd0 = queue.at(0);
  double minCost = pd0.cost + model.minCostCmd + model.getLiteralCosts(pd0.pos, pos);
  size_t minLen = ComputeMinimumCopyLength(minCost, nodes, numBytes, pos);
  size_t result = 0;
  for (size_t k = 0; k < maxIters && k < queue.size(); k++) {
    const PosData& pd = queue.at(k);
    size_t start = pd.pos;
    uint32_t insCode = GetInsertLengthCode((uint32_t)(pos - start));
    double baseCost = pd.costdiff + kInsExtra[insCode] + model.getLiteralCosts(0, pos);
    size_t bestLen = minLen - 1;
    for (int j = 0; j < 16 && bestLen < maxLen; j++) {
      int64_t backward = (int64_t)pd.distanceCache[kIdx[j]] + kOff[j];
      if (backward <= 0 || backward > (int64_t)maxDistance) continue;
      if (pd.distanceCache[kIdx[j]] <= 0) continue;   // unknown ring entry of a non-first part (never in the reference)
      size_t prevIx = (curIx - (size_t)backward) & mask;
      if (cur + bestLen > mask) break;
      if (rb.at(prevIx + bestLen) != rb.at(cur + bestLen)) continue;
      size_t len = 0;
      while (len < maxLen && rb.p[prevIx + len] == rb.p[cur + len]) len++;
      if (len >= 4) {
        double distCost = baseCost + model.getDistanceCost(j);
        for (size_t l = bestLen + 1; l <= len; l++) {
          uint32_t copyCode = GetCopyLengthCode((uint32_t)l);
          uint32_t cmdCode = CombineLengthCodes(insCode, copyCode, j == 0);
          double cost = (cmdCode < 128 ? baseCost : distCost) + kCopyExtra[copyCode] + model.getCommandCost(cmdCode);
          if (cost < nodes[pos + l].cost) { UpdateZopfliNode(nodes, pos, start, l, l, (uint32_t)backward, j + 1, cost); result = std::max(result, l); }
          bestLen = l;
        }
      }
    }
    if (k >= 2) continue;
    size_t matchLen = minLen;
    for (size_t j = 0; j < matches.size(); j++) {
      uint32_t dist = matches[j].distance;
      bool isDict = dist > maxDistance;
      uint32_t distCode = dist + 16 - 1;
      // JS number arithmetic: log2FloorNonZero(1) - 1 == -1 for distance 1 (backward-references-hq.ts:356-357)
      int nbits = distCode < 16 ? 0 : (int)Log2FloorNonZero(dist) - 1;
      double distCost = baseCost + (double)nbits + model.getDistanceCost(distCode & 0x3FF);
      size_t maxMatchLen = matches[j].length;
      if (matchLen < maxMatchLen && (isDict || maxMatchLen > maxZopfliLen)) matchLen = maxMatchLen;
      for (; matchLen <= maxMatchLen; matchLen++) {
        size_t lenCode = matchLen;
        uint32_t copyCode = GetCopyLengthCode((uint32_t)lenCode);
        uint32_t cmdCode = CombineLengthCodes(insCode, copyCode, false);
        double cost = distCost + kCopyExtra[copyCode] + model.getCommandCost(cmdCode);
        if (cost < nodes[pos + matchLen].cost) { UpdateZopfliNode(nodes, pos, start, matchLen, lenCode, dist, 0, cost); result = std::max(result, matchLen); }
      }
    }
  }
  return result;
}
// :384-406.  `cost` is reused as the `next` pointer.
static inline size_t ComputeShortestPathFromNodes(size_t numBytes, std::vector<ZopfliNode>& nodes) {
  size_t index = numBytes, numCommands = 0;
  while ((nodes[index].dcodeInsertLength & 0x7FFFFFF) == 0 && nodes[index].length == 1) index--;
  nodes[index].cost = 4294967295.0;
  while (index != 0) {
    size_t len = NodeCommandLength(nodes[index]);
    index -= len;
    nodes[index].cost = (double)len;
    numCommands++;
  }
  return numCommands;
}
// :610-673
static inline void CreateCommandsFromPath(size_t numBytes, size_t blockStart, const std::vector<ZopfliNode>& nodes, int32_t* distCache,
                                          uint32_t lastInsertLen, uint32_t npostfix, uint32_t ndirect, std::vector<Command>& commands,
                                          size_t* numLiteralsOut, uint32_t* lastInsertLenOut) {
  const size_t maxBackwardLimit = ((size_t)1 << 22) - 16;
  commands.clear();
  size_t numLiterals = 0, pos = 0; double offset = nodes[0].cost; bool isFirst = true;
  while (offset != 4294967295.0 && offset != 0) {
    const ZopfliNode& next = nodes[pos + (size_t)offset];
    uint32_t copyLen = NodeCopyLength(next), insertLen = NodeInsertLength(next);
    pos += insertLen;
    if (isFirst) { insertLen += lastInsertLen; isFirst = false; }
    uint32_t distance = next.distance, lenCode = NodeLengthCode(next), distCode = NodeDistanceCode(next);
    commands.push_back(CreateCommand(insertLen, copyLen, (int)lenCode - (int)copyLen, distCode, ndirect, npostfix));
    size_t dictionaryStart = std::min(blockStart + pos, maxBackwardLimit);
    bool isDictionary = distance > dictionaryStart;
    if (!isDictionary && distCode > 0) { distCache[3] = distCache[2]; distCache[2] = distCache[1]; distCache[1] = distCache[0]; distCache[0] = (int32_t)distance; }
    numLiterals += insertLen;
    pos += copyLen;
    offset = next.cost;
  }
  *numLiteralsOut = numLiterals; *lastInsertLenOut = (uint32_t)(numBytes - pos);
}

// :411-482 (quality 10)
static inline void CreateZopfliBackwardReferences(size_t numBytes, size_t position, const Bytes& rb, size_t mask, int quality,
                                                  BinaryTreeHasher& hasher, int32_t* distCache, uint32_t lastInsertLen,
                                                  uint32_t npostfix, uint32_t ndirect, std::vector<Command>& commands,
                                                  size_t* numLiterals, uint32_t* lastInsertLenOut) {
  const size_t maxBackwardLimit = ((size_t)1 << 22) - 16;
  const size_t maxZopfliLen = quality <= 10 ? 150 : 325;
  std::vector<ZopfliNode> nodes(numBytes + 1);
  nodes[0].length = 0; nodes[0].cost = 0;
  ZopfliCostModel model(numBytes, 544);
  model.setFromLiteralCosts(position, rb.p, mask);
  StartPosQueue queue;
  std::vector<BackwardMatch> matches;
  for (size_t i = 0; i + 3 < numBytes; i++) {
    size_t pos = position + i;
    hasher.findAllMatches(rb, mask, pos, numBytes - i, std::min(pos, maxBackwardLimit), matches);
    if (!matches.empty() && matches.back().length > maxZopfliLen) { BackwardMatch m = matches.back(); matches.clear(); matches.push_back(m); }
    size_t skip = UpdateNodes(numBytes, position, i, rb, mask, quality, maxBackwardLimit, distCache, matches, model, queue, nodes);
    if (skip >= 16384) i += skip - 1;
    else if (matches.size() == 1 && matches[0].length > maxZopfliLen) i += matches[0].length - 1;
  }
  ComputeShortestPathFromNodes(numBytes, nodes);
  CreateCommandsFromPath(numBytes, position, nodes, distCache, lastInsertLen, npostfix, ndirect, commands, numLiterals, lastInsertLenOut);
}

// Pass 0 of :485-545 — every position's match list (H10 candidate stage).
static inline void CollectHqMatches(size_t numBytes, size_t position, const Bytes& rb, size_t mask, BinaryTreeHasher& hasher,
                                    std::vector<std::vector<BackwardMatch>>& all) {
  const size_t maxBackwardLimit = ((size_t)1 << 22) - 16;
  all.assign(numBytes, {});
  std::vector<BackwardMatch> matches;
  for (size_t i = 0; i + 3 < numBytes; i++) {
    size_t pos = position + i;
    hasher.findAllMatches(rb, mask, pos, numBytes - i, std::min(pos, maxBackwardLimit), matches);
    if (!matches.empty() && matches.back().length > 325) {
      size_t skip = matches.back().length - 1;
      all[i] = {matches.back()};
      i += skip;
      continue;
    }
    all[i] = matches;
  }
}

// :485-608 (quality 11)
static inline void CreateHqZopfliBackwardReferences(size_t numBytes, size_t position, const Bytes& rb, size_t mask,
                                                    BinaryTreeHasher& hasher, int32_t* distCache, uint32_t lastInsertLen,
                                                    uint32_t npostfix, uint32_t ndirect, std::vector<Command>& commands,
                                                    size_t* numLiterals, uint32_t* lastInsertLenOut) {
  const int quality = 11;
  const size_t maxBackwardLimit = ((size_t)1 << 22) - 16;
  std::vector<std::vector<BackwardMatch>> all;
  CollectHqMatches(numBytes, position, rb, mask, hasher, all);
  int32_t orig[4] = {distCache[0], distCache[1], distCache[2], distCache[3]};
  ZopfliCostModel model(numBytes, 544);
  std::vector<ZopfliNode> nodes(numBytes + 1);
  for (int iteration = 0; iteration < 2; iteration++) {
    if (iteration > 0) std::fill(nodes.begin(), nodes.end(), ZopfliNode());
    nodes[0].length = 0; nodes[0].cost = 0;
    for (int k = 0; k < 4; k++) distCache[k] = orig[k];
    if (iteration == 0) model.setFromLiteralCosts(position, rb.p, mask);
    else model.setFromCommands(position, rb.p, mask, commands, lastInsertLen);
    StartPosQueue queue;
    for (size_t i = 0; i + 3 < numBytes; i++) {
      const std::vector<BackwardMatch>& m = all[i];
      size_t skip = UpdateNodes(numBytes, position, i, rb, mask, quality, maxBackwardLimit, distCache, m, model, queue, nodes);
      if (skip >= 16384) i += skip - 1;
      else if (m.size() == 1 && m[0].length > 325) i += m[0].length - 1;
    }
    ComputeShortestPathFromNodes(numBytes, nodes);
    CreateCommandsFromPath(numBytes, position, nodes, distCache, lastInsertLen, npostfix, ndirect, commands, numLiterals, lastInsertLenOut);
  }
}

}  // namespace oracle
