// ORACLE (test infrastructure) — restates block-splitter.ts.
#pragma once
#include "common.h"
#include "cluster.h"

namespace oracle {

// block-splitter.ts:19-24.  The TS allocates 1024 fixed slots and silently drops writes
// beyond them; the restatement uses dynamic arrays ("use dynamic B").
struct BlockSplit {
  uint32_t numTypes = 1;
  std::vector<uint8_t> types;
  std::vector<uint32_t> lengths;
  uint32_t numBlocks = 0;
};

// block-splitter.ts:35-40  seed = ToInt32(seed * 16807); if (seed < 0) seed += 0x7FFFFFFF
static inline int32_t MyRand(int32_t* seed) {
  int64_t prod = (int64_t)*seed * 16807;          // exact in a double, so exact here
  int32_t s = (int32_t)(uint32_t)(uint64_t)prod;  // `| 0`
  if (s < 0) s += 0x7FFFFFFF;
  *seed = s;
  return s;
}

template <typename T>
static void InitialEntropyCodes(const T* data, size_t length, size_t stride, size_t numHistograms,
                                std::vector<ClusterHistogram>& histograms) {  // :44-74
  int32_t seed = 7;
  int64_t blockLength = (int64_t)(length / numHistograms);
  for (size_t i = 0; i < numHistograms; i++) histograms[i].clear();
  for (size_t i = 0; i < numHistograms; i++) {
    int64_t pos = (int64_t)(((unsigned __int128)length * i) / numHistograms);  // Math.floor(length*i/n)
    if (i != 0) pos += (int64_t)MyRand(&seed) % blockLength;
    if (pos + (int64_t)stride >= (int64_t)length) pos = (int64_t)length - (int64_t)stride - 1;
    for (size_t j = 0; j < stride && pos + (int64_t)j < (int64_t)length; j++) {
      // a negative pos reads `undefined` in JS -> data[undefined]++ is a no-op on a typed array
      if (pos + (int64_t)j < 0) { histograms[i].totalCount++; continue; }
      histograms[i].data[data[pos + j]]++;
      histograms[i].totalCount++;
    }
  }
}

template <typename T>
static void RefineEntropyCodes(const T* data, size_t length, size_t stride, size_t numHistograms,
                               std::vector<ClusterHistogram>& histograms, ClusterHistogram& tmp) {  // :76-110
  size_t iters = 2 * (length / stride) + 100;
  int32_t seed = 7;
  iters = ((iters + numHistograms - 1) / numHistograms) * numHistograms;
  for (size_t iter = 0; iter < iters; iter++) {
    tmp.clear();
    if (stride >= length) {
      for (size_t j = 0; j < length; j++) { tmp.data[data[j]]++; tmp.totalCount++; }
    } else {
      int64_t pos = (int64_t)MyRand(&seed) % (int64_t)(length - stride + 1);
      for (size_t j = 0; j < stride; j++) {
        if (pos + (int64_t)j >= 0) tmp.data[data[pos + j]]++;
        tmp.totalCount++;
      }
    }
    histograms[iter % numHistograms].add(tmp);
  }
}

// block-splitter.ts:112-114 (sic: + 2 for empty bins, unlike the C encoder's -2)
static inline double SplitBitCost(uint32_t count) { return count == 0 ? FastLog2(1) + 2 : FastLog2(count); }

template <typename T>
static size_t FindBlocks(const T* data, size_t length, double blockSwitchBitcost, size_t numHistograms,
                         const std::vector<ClusterHistogram>& histograms, uint8_t* blockId) {  // :117-211
  size_t alphabetSize = histograms[0].data.size();
  size_t bitmapLen = (numHistograms + 7) >> 3;
  if (numHistograms <= 1) { for (size_t i = 0; i < length; i++) blockId[i] = 0; return 1; }
  std::vector<double> insertCost(alphabetSize * numHistograms, 0.0);
  for (size_t i = 0; i < numHistograms; i++) insertCost[i] = FastLog2(histograms[i].totalCount);
  for (size_t i = alphabetSize; i-- > 0;)
    for (size_t j = 0; j < numHistograms; j++)
      insertCost[i * numHistograms + j] = insertCost[j] - SplitBitCost(histograms[j].data[i]);
  std::vector<double> cost(numHistograms, 0.0);
  std::vector<uint8_t> switchSignal(length * bitmapLen, 0);
  size_t numBlocks = 1;
  for (size_t byteIx = 0; byteIx < length; byteIx++) {
    size_t ix = byteIx * bitmapLen;
    size_t insertCostIx = (size_t)data[byteIx] * numHistograms;
    double minCost = 1e99;
    double blockSwitchCost = blockSwitchBitcost;
    if (byteIx < 2000) blockSwitchCost *= 0.77 + (0.07 / 2000) * (double)byteIx;
    for (size_t k = 0; k < numHistograms; k++) {
      cost[k] += insertCost[insertCostIx + k];
      if (cost[k] < minCost) { minCost = cost[k]; blockId[byteIx] = (uint8_t)k; }
    }
    for (size_t k = 0; k < numHistograms; k++) {
      cost[k] -= minCost;
      if (cost[k] >= blockSwitchCost) {
        cost[k] = blockSwitchCost;
        switchSignal[ix + (k >> 3)] |= (uint8_t)(1u << (k & 7));
      }
    }
  }
  size_t byteIx = length - 1;
  uint8_t curId = blockId[byteIx];
  while (byteIx > 0) {
    uint8_t mask = (uint8_t)(1u << (curId & 7));
    byteIx--;
    size_t ix = byteIx * bitmapLen;
    if (switchSignal[ix + (curId >> 3)] & mask) {
      if (curId != blockId[byteIx]) { curId = blockId[byteIx]; numBlocks++; }
    }
    blockId[byteIx] = curId;
  }
  return numBlocks;
}

// block-splitter.ts:213-234
static inline size_t RemapBlockIds(uint8_t* blockIds, size_t length, size_t numHistograms) {
  std::vector<uint16_t> newId(numHistograms, 256);
  size_t nextId = 0;
  for (size_t i = 0; i < length; i++) if (newId[blockIds[i]] == 256) newId[blockIds[i]] = (uint16_t)nextId++;
  for (size_t i = 0; i < length; i++) blockIds[i] = (uint8_t)newId[blockIds[i]];
  return nextId;
}

template <typename T>
static void BuildBlockHistograms(const T* data, size_t length, const uint8_t* blockIds, size_t numHistograms,
                                 std::vector<ClusterHistogram>& histograms) {  // :236-252
  for (size_t i = 0; i < numHistograms; i++) histograms[i].clear();
  for (size_t i = 0; i < length; i++) { ClusterHistogram& h = histograms[blockIds[i]]; h.data[data[i]]++; h.totalCount++; }
}

template <typename T>
static void ClusterBlocks(const T* data, size_t length, size_t numBlocks, const uint8_t* blockIds,
                          size_t alphabetSize, BlockSplit& split) {  // :316-390
  std::vector<uint32_t> blockLengths(numBlocks, 0);
  size_t blockIdx = 0;
  for (size_t i = 0; i < length; i++) {
    blockLengths[blockIdx]++;
    if (i + 1 == length || blockIds[i] != blockIds[i + 1]) blockIdx++;
  }
  std::vector<ClusterHistogram> blockHistograms; blockHistograms.reserve(numBlocks);
  std::vector<uint32_t> histogramSymbols(numBlocks);
  size_t pos = 0;
  for (size_t i = 0; i < numBlocks; i++) {
    ClusterHistogram h(alphabetSize);
    for (uint32_t j = 0; j < blockLengths[i]; j++) { h.data[data[pos++]]++; h.totalCount++; }
    h.bitCost = h.cost();
    blockHistograms.push_back(std::move(h));
    histogramSymbols[i] = (uint32_t)i;
  }
  std::vector<ClusterHistogram> out(numBlocks, ClusterHistogram(alphabetSize));
  ClusterHistograms(blockHistograms, numBlocks, 256, out, histogramSymbols.data());
  std::vector<uint32_t> newIndex(numBlocks, 0xFFFFFFFFu);
  uint32_t nextIndex = 0, curLength = 0;
  split.types.clear(); split.lengths.clear();
  for (size_t i = 0; i < numBlocks; i++) {
    curLength += blockLengths[i];
    if (i + 1 == numBlocks || histogramSymbols[i] != histogramSymbols[i + 1]) {
      uint32_t symbol = histogramSymbols[i];
      if (newIndex[symbol] == 0xFFFFFFFFu) newIndex[symbol] = nextIndex++;
      split.types.push_back((uint8_t)newIndex[symbol]);
      split.lengths.push_back(curLength);
      curLength = 0;
    }
  }
  split.numBlocks = (uint32_t)split.types.size();
  split.numTypes = nextIndex;
}

template <typename T>
static void SplitByteVector(const T* data, size_t length, size_t alphabetSize, size_t symbolsPerHistogram,
                            size_t maxHistograms, size_t stride, double blockSwitchCost, int quality,
                            BlockSplit& split) {  // :255-314
  size_t numHistograms = length / symbolsPerHistogram + 1;
  if (numHistograms > maxHistograms) numHistograms = maxHistograms;
  if (length == 0) { split.numTypes = 1; return; }
  if (length < 128) {
    split.numTypes = 1;
    split.types.push_back(0); split.lengths.push_back((uint32_t)length); split.numBlocks++;
    return;
  }
  std::vector<ClusterHistogram> histograms(numHistograms + 1, ClusterHistogram(alphabetSize));
  ClusterHistogram& tmp = histograms[numHistograms];
  InitialEntropyCodes(data, length, stride, numHistograms, histograms);
  RefineEntropyCodes(data, length, stride, numHistograms, histograms, tmp);
  std::vector<uint8_t> blockIds(length, 0);
  int iters = quality < 10 ? 3 : 10;
  size_t numBlocks = 0;
  for (int i = 0; i < iters; i++) {
    numBlocks = FindBlocks(data, length, blockSwitchCost, numHistograms, histograms, blockIds.data());
    numHistograms = RemapBlockIds(blockIds.data(), length, numHistograms);
    BuildBlockHistograms(data, length, blockIds.data(), numHistograms, histograms);
  }
  ClusterBlocks(data, length, numBlocks, blockIds.data(), alphabetSize, split);
}

// block-splitter.ts:394-464
static inline void SplitBlock(const std::vector<Command>& commands, const uint8_t* data, size_t offset, size_t mask,
                              int quality, BlockSplit& literalSplit, BlockSplit& cmdSplit, BlockSplit& distSplit) {
  std::vector<uint8_t> literals; std::vector<uint16_t> cmds, dists;
  size_t pos = offset;
  for (const Command& c : commands) {
    for (uint32_t i = 0; i < c.insertLen; i++) { literals.push_back(data[pos & mask]); pos++; }
    cmds.push_back(c.cmdPrefix);
    if (c.cmdPrefix >= 128) dists.push_back(c.distPrefix & 0x3FF);  // sic: counts a trailing insert-only command too
    pos += CommandCopyLen(c);
  }
  if (!literals.empty()) SplitByteVector(literals.data(), literals.size(), 256, 512, 64, 70, 26.0, quality, literalSplit);
  else { literalSplit.numTypes = 1; literalSplit.numBlocks = 0; }
  if (!cmds.empty()) SplitByteVector(cmds.data(), cmds.size(), 704, 1024, 64, 50, 28.1, quality, cmdSplit);
  else { cmdSplit.numTypes = 1; cmdSplit.numBlocks = 0; }
  if (!dists.empty()) SplitByteVector(dists.data(), dists.size(), 544, 512, 64, 50, 28.1, quality, distSplit);
  else { distSplit.numTypes = 1; distSplit.numBlocks = 0; }
}

}  // namespace oracle
