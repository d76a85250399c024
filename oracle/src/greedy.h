// ORACLE (test infrastructure) — restates match.ts, hash-chains.ts:19-153,240-261,
// hash-simple.ts:10-150 and backward-references.ts:14-158.
#pragma once
#include "common.h"
#include "command.h"

namespace oracle {

// match.ts:20-35
static inline int64_t BackwardReferenceScore(uint32_t copyLength, uint32_t backwardDistance) {
  return 30 * 8 * 4 + 135 * (int64_t)copyLength - 30 * (int64_t)Log2FloorNonZero(backwardDistance);
}
static inline int64_t BackwardReferenceScoreUsingLastDistance(uint32_t copyLength) {
  return 135 * (int64_t)copyLength + 30 * 8 * 4 + 15;
}
// match.ts:42-77 (bounded by limit, never reads out of range)
static inline uint32_t FindMatchLength(const uint8_t* data, size_t s1, size_t s2, uint32_t limit) {
  uint32_t matched = 0;
  while (matched < limit && data[s1 + matched] == data[s2 + matched]) matched++;
  return matched;
}
// match.ts:162-172
static inline uint32_t HashBytes4(const Bytes& d, size_t pos, int bucketBits) {
  uint32_t h32 = d.at0(pos) | (d.at0(pos + 1) << 8) | (d.at0(pos + 2) << 16) | (d.at0(pos + 3) << 24);
  return (h32 * 0x1E35A7BDu) >> (32 - bucketBits);
}
// match.ts:128-148.  The TS passes a 4-entry Int32Array, so for q7-9 these writes are
// dropped and entries 4..15 read as `undefined`.  The restatement keeps a real
// 16-entry cache (deviation "fix (b)").
static inline void PrepareDistanceCache(int32_t* dc, int numDistances) {
  if (numDistances > 4) {
    int32_t last = dc[0];
    dc[4] = last - 1; dc[5] = last + 1; dc[6] = last - 2; dc[7] = last + 2; dc[8] = last - 3; dc[9] = last + 3;
    if (numDistances > 10) {
      int32_t nl = dc[1];
      dc[10] = nl - 1; dc[11] = nl + 1; dc[12] = nl - 2; dc[13] = nl + 2; dc[14] = nl - 3; dc[15] = nl + 3;
    }
  }
}

struct HasherSearchResult { uint32_t len = 0, distance = 0; int64_t score = 0; int lenCodeDelta = 0; };

// hash-chains.ts:19-153
class HashChainHasher {
 public:
  std::vector<uint32_t> buckets, chains;
  int bucketBits, blockBits; uint32_t windowMask; int numLastDistancesToCheck;
  HashChainHasher(int bucketBits_, int blockBits_, int lgwin, int numLast)
      : buckets((size_t)1 << bucketBits_, 0), chains((size_t)1 << lgwin, 0), bucketBits(bucketBits_), blockBits(blockBits_),
        windowMask((1u << lgwin) - 1), numLastDistancesToCheck(numLast) {}
  // :51-60
  void store(const Bytes& data, size_t mask, size_t ix) {
    uint32_t key = HashBytes4(data, ix & mask, bucketBits);
    chains[ix & windowMask] = buckets[key];
    buckets[key] = (uint32_t)ix;
  }
  // :69-153
  void findLongestMatch(const Bytes& data, size_t ringBufferMask, int32_t* distanceCache, size_t curIx,
                        uint32_t maxLength, size_t maxBackward, HasherSearchResult& out) {
    size_t cur = curIx & ringBufferMask;
    uint32_t bestLen = out.len; int64_t bestScore = out.score;
    uint32_t key = HashBytes4(data, cur, bucketBits);
    size_t minorKey = curIx & windowMask;
    out.lenCodeDelta = 0;
    PrepareDistanceCache(distanceCache, numLastDistancesToCheck);
    for (int i = 0; i < numLastDistancesToCheck; i++) {
      int64_t cachedBackward = distanceCache[i];
      if (cachedBackward <= 0 || cachedBackward > (int64_t)maxBackward) continue;
      size_t prevIx = (curIx - (size_t)cachedBackward) & ringBufferMask;
      if (data.at(prevIx + bestLen) != data.at(cur + bestLen)) continue;
      uint32_t len = FindMatchLength(data.p, prevIx, cur, maxLength);
      if (len >= 4) {
        int64_t score = BackwardReferenceScoreUsingLastDistance(len);
        if (score > bestScore) { bestLen = len; out.len = len; out.distance = (uint32_t)cachedBackward; out.score = score; bestScore = score; }
      }
    }
    chains[minorKey] = buckets[key];
    buckets[key] = (uint32_t)curIx;
    size_t maxChainLength = (size_t)1 << blockBits;
    size_t prevIx = chains[minorKey];
    for (size_t chainLen = 0; chainLen < maxChainLength; chainLen++) {
      size_t backward = curIx - prevIx;
      if (backward == 0 || backward > maxBackward) break;
      size_t prevMasked = prevIx & ringBufferMask;
      if (data.at(prevMasked + bestLen) != data.at(cur + bestLen)) { prevIx = chains[prevIx & windowMask]; continue; }
      uint32_t len = FindMatchLength(data.p, prevMasked, cur, maxLength);
      if (len >= 4) {
        int64_t score = BackwardReferenceScore(len, (uint32_t)backward);
        if (score > bestScore) { bestLen = len; out.len = len; out.distance = (uint32_t)backward; out.score = score; bestScore = score; }
      }
      prevIx = chains[prevIx & windowMask];
    }
  }
};
// hash-chains.ts:240-261
static inline HashChainHasher* CreateHashChainHasher(int quality, int lgwin) {
  if (quality < 7) return new HashChainHasher(14, quality - 1, lgwin, 4);
  if (quality < 9) return new HashChainHasher(15, quality - 1, lgwin, 10);
  return new HashChainHasher(15, quality - 1, lgwin, 16);
}

// match.ts:174-189 hashBytes8, hashLen == 5 path (the only one hash-simple.ts uses)
static inline uint32_t HashBytes5(const Bytes& d, size_t pos, int bucketBits) {
  uint32_t h32 = d.at0(pos) | (d.at0(pos + 1) << 8) | (d.at0(pos + 2) << 16) | (d.at0(pos + 3) << 24);
  uint32_t b4 = d.at0(pos + 4);
  return ((h32 ^ (b4 << 24)) * 0x1E35A7BDu) >> (32 - bucketBits);
}

// hash-simple.ts:20-128: one slot per bucket, zero-initialised (so an empty bucket reads as position 0)
class SimpleHasher {
 public:
  std::vector<uint32_t> buckets; int bucketBits;
  explicit SimpleHasher(int bucketBits_) : buckets((size_t)1 << bucketBits_, 0), bucketBits(bucketBits_) {}
  // :52-55
  void store(const Bytes& data, size_t mask, size_t ix) { buckets[HashBytes5(data, ix & mask, bucketBits)] = (uint32_t)ix; }
  // :64-128
  void findLongestMatch(const Bytes& data, size_t ringBufferMask, int32_t* distanceCache, size_t curIx,
                        uint32_t maxLength, size_t maxBackward, HasherSearchResult& out) {
    size_t cur = curIx & ringBufferMask;
    uint32_t bestLen = out.len; int64_t bestScore = out.score;
    uint32_t key = HashBytes5(data, cur, bucketBits);
    out.lenCodeDelta = 0;
    int64_t cachedBackward = distanceCache[0];
    if (cachedBackward > 0 && cachedBackward <= (int64_t)maxBackward) {
      size_t prevIx = (curIx - (size_t)cachedBackward) & ringBufferMask;
      if (data.at(prevIx + bestLen) == data.at(cur + bestLen)) {
        uint32_t len = FindMatchLength(data.p, prevIx, cur, maxLength);
        if (len >= 4) {
          int64_t score = BackwardReferenceScoreUsingLastDistance(len);
          if (score > bestScore) { bestLen = len; out.len = len; out.distance = (uint32_t)cachedBackward; out.score = score; bestScore = score; }
        }
      }
    }
    size_t prevIx = buckets[key];
    buckets[key] = (uint32_t)curIx;
    size_t backward = curIx - prevIx;
    if (backward == 0 || backward > maxBackward) return;
    prevIx &= ringBufferMask;
    if (data.at(prevIx + bestLen) != data.at(cur + bestLen)) return;
    uint32_t len = FindMatchLength(data.p, prevIx, cur, maxLength);
    if (len >= 4) {
      int64_t score = BackwardReferenceScore(len, (uint32_t)backward);
      if (score > bestScore) { out.len = len; out.distance = (uint32_t)backward; out.score = score; }
    }
  }
};
// hash-simple.ts:134-150 (quality 2: 16 bits; 1, 3, 4: 17 bits)
static inline SimpleHasher* CreateSimpleHasher(int quality) { return new SimpleHasher(quality == 2 ? 16 : 17); }

// backward-references.ts:136-158
static inline uint32_t DistanceToCode(uint32_t distance, const int32_t* distCache) {
  static const uint8_t idx[16] = {0, 1, 2, 3, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 1};
  static const int8_t off[16] = {0, 0, 0, 0, -1, 1, -2, 2, -3, 3, -1, 1, -2, 2, -3, 3};
  for (int i = 0; i < 16; i++) {
    int64_t cached = (int64_t)distCache[idx[i]] + off[i];
    // distCache[..] > 0: always true in the reference; a part that does not open the stream starts with
    // an unknown (all-zero) ring whose entries must never be referenced (SURVEY.md §8e rule 3)
    if ((int64_t)distance == cached && cached > 0 && distCache[idx[i]] > 0) return i;
  }
  return distance + 16 - 1;
}

// backward-references.ts:14-134.  storeStride = 1 for quality > 2 (every covered position is stored),
// 4 for quality <= 2 (:99-104).  distCache must have 16 slots.
template <class Hasher>
static inline void CreateBackwardReferences(size_t numBytes, size_t position, const Bytes& rb, size_t mask,
                                            Hasher& hasher, int32_t* distCache, uint32_t lastInsertLen,
                                            uint32_t npostfix, uint32_t ndirect, std::vector<Command>& commands,
                                            size_t* numLiteralsOut, uint32_t* lastInsertLenOut, size_t storeStride = 1) {
  size_t numLiterals = 0; uint32_t insertLen = lastInsertLen;
  size_t pos = position; const size_t posEnd = position + numBytes;
  const size_t maxWindowBackward = ((size_t)1 << 22) - 16;
  HasherSearchResult result;
  while (pos < posEnd) {
    size_t maxLen = posEnd - pos;
    if (maxLen < 4) { insertLen += (uint32_t)maxLen; pos += maxLen; break; }
    size_t maxBackward = std::min(pos, maxWindowBackward);
    result = HasherSearchResult();
    hasher.findLongestMatch(rb, mask, distCache, pos, (uint32_t)std::min<size_t>(maxLen, 128), maxBackward, result);
    if (result.len >= 4 && result.score > 0 && result.distance > 0) {
      uint32_t distance = result.distance, matchLen = result.len;
      if (distance > pos) { insertLen++; pos++; continue; }
      uint32_t distCode = DistanceToCode(distance, distCache);
      commands.push_back(CreateCommand(insertLen, matchLen, result.lenCodeDelta, distCode, ndirect, npostfix));
      numLiterals += insertLen;
      if (distCode > 0) { distCache[3] = distCache[2]; distCache[2] = distCache[1]; distCache[1] = distCache[0]; distCache[0] = (int32_t)distance; }
      size_t storeEnd = std::min(pos + matchLen, posEnd >= 4 ? posEnd - 4 : 0);
      for (size_t i = pos + 1; i < storeEnd; i += storeStride) hasher.store(rb, mask, i);
      pos += matchLen;
      insertLen = 0;
    } else { insertLen++; pos++; }
  }
  if (insertLen > 0) { commands.push_back(CreateInsertCommand(insertLen)); numLiterals += insertLen; insertLen = 0; }
  *numLiteralsOut = numLiterals; *lastInsertLenOut = insertLen;
}

}  // namespace oracle
