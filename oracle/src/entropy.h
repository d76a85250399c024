// ORACLE (test infrastructure) — restates entropy-encode.ts:24-258 and context-map.ts.
#pragma once
#include "common.h"
#include "bit_writer.h"

namespace oracle {

// entropy-encode.ts:222-232
static inline uint32_t ReverseBits(unsigned numBits, uint32_t bits) {
  static const uint32_t LUT[16] = {0x0, 0x8, 0x4, 0xC, 0x2, 0xA, 0x6, 0xE, 0x1, 0x9, 0x5, 0xD, 0x3, 0xB, 0x7, 0xF};
  uint32_t retval = LUT[bits & 0xF];
  for (unsigned i = 4; i < numBits; i += 4) { retval <<= 4; bits >>= 4; retval |= LUT[bits & 0xF]; }
  retval >>= ((0u - numBits) & 0x3);
  return retval;
}

// entropy-encode.ts:234-258
static inline void ConvertBitDepthsToSymbols(const uint8_t* depths, size_t len, uint16_t* bits) {
  uint16_t blCount[16] = {0}, nextCode[16] = {0};
  for (size_t i = 0; i < len; i++) blCount[depths[i]]++;
  blCount[0] = 0;
  uint32_t code = 0;
  for (int i = 1; i <= 15; i++) { code = (code + blCount[i - 1]) << 1; nextCode[i] = (uint16_t)code; }
  for (size_t i = 0; i < len; i++)
    if (depths[i] > 0) bits[i] = (uint16_t)ReverseBits(depths[i], nextCode[depths[i]]++);
}

// entropy-encode.ts:133-165.  stack has 16 entries; level never exceeds maxDepth <= 15.
static inline bool SetDepth(int root, const int32_t* left, const int32_t* rightOrValue, uint8_t* depths, int maxDepth) {
  int32_t stack[16]; int level = 0; int p = root; stack[0] = -1;
  for (;;) {
    int32_t l = left[p];
    if (l >= 0) {
      level++;
      if (level > maxDepth) return false;
      stack[level] = rightOrValue[p];
      p = l;
      continue;
    } else {
      depths[rightOrValue[p]] = (uint8_t)level;
    }
    while (level >= 0 && stack[level] == -1) level--;
    if (level < 0) return true;
    p = stack[level];
    stack[level] = -1;
  }
}

// entropy-encode.ts:24-131.  `depths`/`bits` must hold `length` entries; they are zeroed first.
// Sort order (count asc, symbol desc) is a total order, so std::sort reproduces :167-220.
static inline void CreateHuffmanTree(const uint32_t* histogram, size_t length, int treeLimit,
                                     uint8_t* depths, uint16_t* bits) {
  memset(depths, 0, length); memset(bits, 0, length * 2);
  size_t nonZero = 0, lastNonZero = 0;
  for (size_t i = 0; i < length; i++) if (histogram[i] > 0) { nonZero++; lastNonZero = i; }
  if (nonZero == 0) return;
  if (nonZero == 1) { depths[lastNonZero] = 1; bits[lastNonZero] = 0; return; }
  size_t maxNodes = 2 * length + 2;
  std::vector<uint32_t> cnt(maxNodes); std::vector<int32_t> left(maxNodes), rv(maxNodes);
  struct Leaf { uint32_t c; int32_t v; };
  std::vector<Leaf> leaves; leaves.reserve(nonZero);
  for (uint32_t countLimit = 1;; countLimit *= 2) {
    leaves.clear();
    for (size_t i = length; i-- > 0;)
      if (histogram[i] > 0) leaves.push_back({std::max(histogram[i], countLimit), (int32_t)i});
    std::sort(leaves.begin(), leaves.end(), [](const Leaf& a, const Leaf& b) {
      if (a.c != b.c) return a.c < b.c;
      return a.v > b.v;
    });
    size_t n = leaves.size();
    for (size_t k = 0; k < n; k++) { cnt[k] = leaves[k].c; left[k] = -1; rv[k] = leaves[k].v; }
    cnt[n] = 0xFFFFFFFFu; left[n] = -1; rv[n] = -1;
    cnt[n + 1] = 0xFFFFFFFFu; left[n + 1] = -1; rv[n + 1] = -1;
    size_t i = 0, j = n + 1;
    for (size_t k = n - 1; k > 0; k--) {
      size_t l, r;
      if (cnt[i] <= cnt[j]) l = i++; else l = j++;
      if (cnt[i] <= cnt[j]) r = i++; else r = j++;
      size_t jEnd = 2 * n - k;
      cnt[jEnd] = cnt[l] + cnt[r];  // u32 wrap == `>>> 0`
      left[jEnd] = (int32_t)l; rv[jEnd] = (int32_t)r;
      cnt[jEnd + 1] = 0xFFFFFFFFu; left[jEnd + 1] = -1; rv[jEnd + 1] = -1;
    }
    if (SetDepth((int)(2 * n - 1), left.data(), rv.data(), depths, treeLimit)) break;
    memset(depths, 0, length);
  }
  ConvertBitDepthsToSymbols(depths, length, bits);
}

// context-map.ts:524-533
static inline void StoreVarLenUint8(BitWriter& w, uint32_t n) {
  if (n == 0) { w.writeBits(1, 0); return; }
  unsigned nbits = Log2FloorNonZero(n);
  w.writeBits(1, 1); w.writeBits(3, nbits); w.writeBits(nbits, n - (1u << nbits));
}

// context-map.ts:395-473 (RLE of code lengths; the TS always uses RLE)
static inline void WriteHuffmanTreeRepetitions(unsigned prevValue, unsigned value, unsigned reps,
                                               std::vector<uint8_t>& tree, std::vector<uint8_t>& extra) {
  if (prevValue != value) { tree.push_back(value); extra.push_back(0); reps--; }
  if (reps == 7) { tree.push_back(value); extra.push_back(0); reps--; }
  if (reps < 3) {
    for (unsigned j = 0; j < reps; j++) { tree.push_back(value); extra.push_back(0); }
  } else {
    size_t start = tree.size();
    reps -= 3;
    for (;;) {
      tree.push_back(16); extra.push_back(reps & 3);
      reps >>= 2;
      if (reps == 0) break;
      reps--;
    }
    std::reverse(tree.begin() + start, tree.end());
    std::reverse(extra.begin() + start, extra.end());
  }
}
static inline void WriteHuffmanTreeRepetitionsZeros(unsigned reps, std::vector<uint8_t>& tree, std::vector<uint8_t>& extra) {
  if (reps == 11) { tree.push_back(0); extra.push_back(0); reps--; }
  if (reps < 3) {
    for (unsigned j = 0; j < reps; j++) { tree.push_back(0); extra.push_back(0); }
  } else {
    size_t start = tree.size();
    reps -= 3;
    for (;;) {
      tree.push_back(17); extra.push_back(reps & 7);
      reps >>= 3;
      if (reps == 0) break;
      reps--;
    }
    std::reverse(tree.begin() + start, tree.end());
    std::reverse(extra.begin() + start, extra.end());
  }
}
// context-map.ts:350-393
static inline void WriteHuffmanTreeRepresentation(const uint8_t* depths, size_t length,
                                                  std::vector<uint8_t>& tree, std::vector<uint8_t>& extra) {
  size_t newLength = length;
  while (newLength > 0 && depths[newLength - 1] == 0) newLength--;
  unsigned prevValue = 8;
  size_t i = 0;
  while (i < newLength) {
    unsigned value = depths[i];
    unsigned reps = 1;
    while (i + reps < newLength && depths[i + reps] == value) reps++;
    i += reps;
    if (value == 0) WriteHuffmanTreeRepetitionsZeros(reps, tree, extra);
    else { WriteHuffmanTreeRepetitions(prevValue, value, reps, tree, extra); prevValue = value; }
  }
}
// context-map.ts:485-520
static inline void StoreHuffmanTreeOfHuffmanTree(BitWriter& w, int numCodes, const uint8_t* depths) {
  static const uint8_t storageOrder[18] = {1, 2, 3, 4, 0, 5, 17, 6, 16, 7, 8, 9, 10, 11, 12, 13, 14, 15};
  static const uint8_t symbols[6] = {0, 7, 3, 2, 1, 15};
  static const uint8_t bitLengths[6] = {2, 4, 3, 2, 2, 4};
  int codesToStore = 18;
  if (numCodes > 1) while (codesToStore > 0 && depths[storageOrder[codesToStore - 1]] == 0) codesToStore--;
  int skipSome = 0;
  if (depths[storageOrder[0]] == 0 && depths[storageOrder[1]] == 0) {
    skipSome = 2;
    if (depths[storageOrder[2]] == 0) skipSome = 3;
  }
  w.writeBits(2, skipSome);
  for (int i = skipSome; i < codesToStore; i++) {
    unsigned len = depths[storageOrder[i]];
    w.writeBits(bitLengths[len], symbols[len]);
  }
}
// context-map.ts:293-347
static inline void StoreComplexHuffmanTree(BitWriter& w, const uint8_t* depths, size_t length) {
  std::vector<uint8_t> tree, extra;
  WriteHuffmanTreeRepresentation(depths, length, tree, extra);
  uint32_t clHisto[18] = {0};
  for (uint8_t c : tree) clHisto[c]++;
  int numCodes = 0, firstCode = 0;
  for (int i = 0; i < 18; i++) if (clHisto[i]) { if (numCodes == 0) firstCode = i; numCodes++; }
  uint8_t clDepths[18]; uint16_t clBits[18];
  CreateHuffmanTree(clHisto, 18, 5, clDepths, clBits);
  StoreHuffmanTreeOfHuffmanTree(w, numCodes, clDepths);
  if (numCodes == 1) clDepths[firstCode] = 0;
  for (size_t i = 0; i < tree.size(); i++) {
    unsigned code = tree[i];
    w.writeBits(clDepths[code], clBits[code]);
    if (code == 16) w.writeBits(2, extra[i]);
    else if (code == 17) w.writeBits(3, extra[i]);
  }
}
// context-map.ts:266-291 (JS Array.sort is stable -> std::stable_sort)
static inline void StoreSimpleHuffmanTree(BitWriter& w, const uint8_t* depths, const uint32_t* s4, int numSymbols, unsigned maxBits) {
  uint32_t sorted[4];
  for (int i = 0; i < numSymbols; i++) sorted[i] = s4[i];
  std::stable_sort(sorted, sorted + numSymbols, [&](uint32_t a, uint32_t b) { return depths[a] < depths[b]; });
  w.writeBits(2, 1);
  w.writeBits(2, numSymbols - 1);
  for (int i = 0; i < numSymbols; i++) w.writeBits(maxBits, sorted[i]);
  if (numSymbols == 4) w.writeBits(1, depths[sorted[0]] == 1 ? 1 : 0);
}
// context-map.ts:215-264.  depths/bits must hold alphabetSize entries.
static inline void BuildAndStoreHuffmanTree(BitWriter& w, const uint32_t* histogram, size_t alphabetSize,
                                            uint8_t* depths, uint16_t* bits) {
  int count = 0; uint32_t s4[4] = {0, 0, 0, 0};
  for (size_t i = 0; i < alphabetSize; i++) if (histogram[i]) { if (count < 4) s4[count] = (uint32_t)i; count++; }
  unsigned maxBits = 0; size_t c = alphabetSize - 1;
  while (c) { c >>= 1; maxBits++; }
  if (count <= 1) {
    w.writeBits(4, 1);
    w.writeBits(maxBits, s4[0]);
    depths[s4[0]] = 0; bits[s4[0]] = 0;
    return;
  }
  CreateHuffmanTree(histogram, alphabetSize, 15, depths, bits);
  if (count <= 4) StoreSimpleHuffmanTree(w, depths, s4, count, maxBits);
  else StoreComplexHuffmanTree(w, depths, alphabetSize);
}

// context-map.ts:13-50
static inline std::vector<uint32_t> MoveToFrontTransform(const uint32_t* in, size_t size) {
  std::vector<uint32_t> out(size);
  if (size == 0) return out;
  uint32_t maxValue = in[0];
  for (size_t i = 1; i < size; i++) if (in[i] > maxValue) maxValue = in[i];
  std::vector<uint8_t> mtf(maxValue + 1);
  for (uint32_t i = 0; i <= maxValue; i++) mtf[i] = (uint8_t)i;
  for (size_t i = 0; i < size; i++) {
    uint32_t value = in[i]; uint32_t index = 0;
    while (mtf[index] != value) index++;
    out[i] = index;
    for (uint32_t j = index; j > 0; j--) mtf[j] = mtf[j - 1];
    mtf[0] = (uint8_t)value;
  }
  return out;
}
// context-map.ts:55-111
static inline void RunLengthCodeZeros(const std::vector<uint32_t>& in, unsigned maxRunLengthPrefix,
                                      std::vector<uint32_t>& out, unsigned* maxPrefixOut) {
  size_t size = in.size();
  uint32_t maxReps = 0;
  for (size_t i = 0; i < size;) {
    uint32_t reps = 0;
    while (i < size && in[i] != 0) i++;
    while (i < size && in[i] == 0) { reps++; i++; }
    if (reps > maxReps) maxReps = reps;
  }
  unsigned maxPrefix = maxReps > 0 ? Log2FloorNonZero(maxReps) : 0;
  maxPrefix = std::min(maxPrefix, maxRunLengthPrefix);
  out.clear();
  for (size_t i = 0; i < size;) {
    if (in[i] != 0) { out.push_back(in[i] + maxPrefix); i++; }
    else {
      uint32_t reps = 1;
      for (size_t k = i + 1; k < size && in[k] == 0; k++) reps++;
      i += reps;
      while (reps != 0) {
        if (reps < (2u << maxPrefix)) {
          unsigned p = Log2FloorNonZero(reps);
          out.push_back(p | ((reps - (1u << p)) << 9));
          break;
        } else {
          out.push_back(maxPrefix | (((1u << maxPrefix) - 1) << 9));
          reps -= (2u << maxPrefix) - 1;
        }
      }
    }
  }
  *maxPrefixOut = maxPrefix;
}
// context-map.ts:114-170
static inline void EncodeContextMap(BitWriter& w, const uint32_t* contextMap, size_t contextMapSize, uint32_t numClusters) {
  StoreVarLenUint8(w, numClusters - 1);
  if (numClusters == 1) return;
  std::vector<uint32_t> mtf = MoveToFrontTransform(contextMap, contextMapSize);
  std::vector<uint32_t> rle; unsigned maxPrefix;
  RunLengthCodeZeros(mtf, 6, rle, &maxPrefix);
  const uint32_t symbolMask = (1u << 9) - 1;
  size_t alphabetSize = numClusters + maxPrefix;
  std::vector<uint32_t> histogram(alphabetSize, 0);
  for (uint32_t s : rle) histogram[s & symbolMask]++;
  bool useRle = maxPrefix > 0;
  w.writeBits(1, useRle ? 1 : 0);
  if (useRle) w.writeBits(4, maxPrefix - 1);
  std::vector<uint8_t> depths(alphabetSize, 0); std::vector<uint16_t> bits(alphabetSize, 0);
  BuildAndStoreHuffmanTree(w, histogram.data(), alphabetSize, depths.data(), bits.data());
  for (uint32_t s : rle) {
    uint32_t sym = s & symbolMask, extraVal = s >> 9;
    w.writeBits(depths[sym], bits[sym]);
    if (sym > 0 && sym <= maxPrefix) w.writeBits(sym, extraVal);
  }
  w.writeBits(1, 1);
}

}  // namespace oracle
