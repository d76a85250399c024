// ORACLE (test infrastructure) — restates metablock.ts.
#pragma once
#include "common.h"
#include "bit_writer.h"
#include "command.h"
#include "entropy.h"
#include "context.h"
#include "cluster.h"
#include "block_splitter.h"

namespace oracle {

// metablock.ts:37-64
static const struct { uint32_t offset, nbits; } kBlockLengthPrefix[26] = {
    {1, 2},    {5, 2},    {9, 2},    {13, 2},   {17, 3},   {25, 3},   {33, 3},    {41, 3},    {49, 4},
    {65, 4},   {81, 4},   {97, 4},   {113, 5},  {145, 5},  {177, 5},  {209, 5},   {241, 6},   {305, 6},
    {369, 7},  {497, 8},  {753, 9},  {1265, 10}, {2289, 11}, {4337, 12}, {8433, 13}, {16625, 24}};

// metablock.ts:118-125
static inline uint32_t BlockLengthPrefixCode(uint32_t len) {
  uint32_t code = len >= 177 ? (len >= 753 ? 20 : 14) : (len >= 41 ? 7 : 0);
  while (code < 25 && len >= kBlockLengthPrefix[code + 1].offset) code++;
  return code;
}
// metablock.ts:133-150
struct BlockTypeCodeCalculator {
  uint32_t lastType = 1, secondLastType = 0;
  uint32_t nextCode(uint32_t type) {
    uint32_t typeCode = type == lastType + 1 ? 1 : type == secondLastType ? 0 : type + 2;
    secondLastType = lastType; lastType = type;
    return typeCode;
  }
};
// metablock.ts:152-158
struct BlockSplitCode {
  std::vector<uint8_t> typeDepths, lengthDepths;
  std::vector<uint16_t> typeBits, lengthBits;
  BlockTypeCodeCalculator typeCalculator;
};
// metablock.ts:204-220
static inline void StoreBlockSwitch(BitWriter& w, BlockSplitCode& code, uint32_t blockLen, uint32_t blockType, bool isFirst) {
  uint32_t typeCode = code.typeCalculator.nextCode(blockType);
  if (!isFirst) w.writeBits(code.typeDepths[typeCode], code.typeBits[typeCode]);
  uint32_t lenCode = BlockLengthPrefixCode(blockLen);
  w.writeBits(code.lengthDepths[lenCode], code.lengthBits[lenCode]);
  w.writeBits(kBlockLengthPrefix[lenCode].nbits, blockLen - kBlockLengthPrefix[lenCode].offset);
}
// metablock.ts:161-202
static inline void BuildAndStoreBlockSplitCode(BitWriter& w, const uint8_t* types, const uint32_t* lengths, size_t numBlocks,
                                               uint32_t numTypes, BlockSplitCode& code) {
  code.typeDepths.assign(numTypes + 2, 0); code.typeBits.assign(numTypes + 2, 0);
  code.lengthDepths.assign(26, 0); code.lengthBits.assign(26, 0);
  std::vector<uint32_t> typeHisto(numTypes + 2, 0), lengthHisto(26, 0);
  BlockTypeCodeCalculator calc;
  for (size_t i = 0; i < numBlocks; i++) {
    uint32_t tc = calc.nextCode(types[i]);
    if (i != 0) typeHisto[tc]++;
    lengthHisto[BlockLengthPrefixCode(lengths[i])]++;
  }
  StoreVarLenUint8(w, numTypes - 1);
  if (numTypes > 1) {
    BuildAndStoreHuffmanTree(w, typeHisto.data(), numTypes + 2, code.typeDepths.data(), code.typeBits.data());
    BuildAndStoreHuffmanTree(w, lengthHisto.data(), 26, code.lengthDepths.data(), code.lengthBits.data());
    StoreBlockSwitch(w, code, lengths[0], types[0], true);
  }
}
// metablock.ts:222-255
static inline void StoreCompressedMetaBlockHeader(BitWriter& w, bool isLast, uint32_t length) {
  w.writeBits(1, isLast ? 1 : 0);
  if (isLast) w.writeBits(1, 0);
  int lg = length == 1 ? 1 : Log2FloorNonZero(length - 1) + 1;
  int mnibbles = (lg < 16 ? 16 : lg + 3) / 4;
  w.writeBits(2, mnibbles - 4);
  w.writeBits(mnibbles * 4, length - 1);
  if (!isLast) w.writeBits(1, 0);
}
// metablock.ts:257-271
static inline void StoreUncompressedMetaBlockHeader(BitWriter& w, uint32_t length) {
  w.writeBits(1, 0);
  int lg = length == 1 ? 1 : Log2FloorNonZero(length - 1) + 1;
  int mnibbles = (lg < 16 ? 16 : lg + 3) / 4;
  w.writeBits(2, mnibbles - 4);
  w.writeBits(mnibbles * 4, length - 1);
  w.writeBits(1, 1);
}
// metablock.ts:273-287 — packed in 64 bits (deviation §0.9c: the TS `<<` is 32-bit)
static inline void StoreCommandExtra(BitWriter& w, const Command& cmd) {
  uint32_t copyLenCode = CommandCopyLenCode(cmd);
  uint32_t insCode = GetInsertLengthCode(cmd.insertLen), copyCode = GetCopyLengthCode(copyLenCode);
  uint32_t insNumExtra = kInsExtra[insCode];
  uint64_t insExtraVal = cmd.insertLen - kInsBase[insCode];
  uint64_t copyExtraVal = copyLenCode - kCopyBase[copyCode];
  w.writeBits(insNumExtra + kCopyExtra[copyCode], (copyExtraVal << insNumExtra) | insExtraVal);
}

struct TrivialTables {  // exposed for stage-level parity tests
  uint32_t litHisto[256], cmdHisto[704], distHisto[544];
  uint8_t litDepths[256], cmdDepths[704], distDepths[544];
  uint16_t litBits[256], cmdBits[704], distBits[544];
};

// metablock.ts:290-389
static inline void StoreMetaBlockTrivial(BitWriter& w, const uint8_t* input, size_t startPos, size_t length, size_t mask,
                                         bool isLast, const std::vector<Command>& commands, uint32_t distAlphabetSize,
                                         uint32_t npostfix, uint32_t ndirect, TrivialTables* tablesOut = nullptr) {
  StoreCompressedMetaBlockHeader(w, isLast, (uint32_t)length);
  TrivialTables* T = tablesOut ? tablesOut : new TrivialTables;
  memset(T, 0, sizeof(*T));
  size_t pos = startPos;
  for (const Command& cmd : commands) {
    T->cmdHisto[cmd.cmdPrefix]++;
    for (uint32_t j = 0; j < cmd.insertLen; j++) T->litHisto[input[(pos + j) & mask]]++;
    pos += cmd.insertLen;
    uint32_t copyLen = CommandCopyLen(cmd);
    pos += copyLen;
    if (copyLen && cmd.cmdPrefix >= 128) T->distHisto[cmd.distPrefix & 0x3FF]++;
  }
  StoreVarLenUint8(w, 0); StoreVarLenUint8(w, 0); StoreVarLenUint8(w, 0);
  w.writeBits(2, npostfix);
  w.writeBits(4, ndirect >> npostfix);
  StoreVarLenUint8(w, 0);
  w.writeBits(2, 0);
  StoreVarLenUint8(w, 0);
  BuildAndStoreHuffmanTree(w, T->litHisto, 256, T->litDepths, T->litBits);
  BuildAndStoreHuffmanTree(w, T->cmdHisto, 704, T->cmdDepths, T->cmdBits);
  BuildAndStoreHuffmanTree(w, T->distHisto, distAlphabetSize, T->distDepths, T->distBits);
  pos = startPos;
  for (const Command& cmd : commands) {
    w.writeBits(T->cmdDepths[cmd.cmdPrefix], T->cmdBits[cmd.cmdPrefix]);
    StoreCommandExtra(w, cmd);
    for (uint32_t j = 0; j < cmd.insertLen; j++) {
      uint8_t lit = input[(pos + j) & mask];
      w.writeBits(T->litDepths[lit], T->litBits[lit]);
    }
    pos += cmd.insertLen;
    uint32_t copyLen = CommandCopyLen(cmd);
    pos += copyLen;
    if (copyLen && cmd.cmdPrefix >= 128) {
      uint32_t dc = cmd.distPrefix & 0x3FF;
      w.writeBits(T->distDepths[dc], T->distBits[dc]);
      w.writeBits(cmd.distPrefix >> 10, cmd.distExtra);
    }
  }
  if (isLast) w.alignToByte();
  if (!tablesOut) delete T;
}

// metablock.ts:392-501
class BlockEncoder {
 public:
  size_t histogramLength; uint32_t numBlockTypes; size_t numBlocks;
  const uint8_t* blockTypes; const uint32_t* blockLengths;
  bool hasSplitCode = false; BlockSplitCode splitCode;
  size_t blockIdx = 0; uint32_t blockLen = 0; size_t entropyIdx = 0;
  std::vector<uint8_t> depths; std::vector<uint16_t> bits;
  BlockEncoder(size_t hl, const BlockSplit& s)
      : histogramLength(hl), numBlockTypes(s.numTypes), numBlocks(s.numBlocks), blockTypes(s.types.data()),
        blockLengths(s.lengths.data()) {
    blockLen = numBlocks > 0 ? blockLengths[0] : 0;
  }
  void buildAndStoreEntropyCodes(BitWriter& w) {
    if (numBlockTypes > 1) { BuildAndStoreBlockSplitCode(w, blockTypes, blockLengths, numBlocks, numBlockTypes, splitCode); hasSplitCode = true; }
    else StoreVarLenUint8(w, 0);
  }
  void buildAndStoreHuffmanTrees(BitWriter& w, const std::vector<std::vector<uint32_t>>& histograms, size_t numHistograms) {
    depths.assign(numHistograms * histogramLength, 0); bits.assign(numHistograms * histogramLength, 0);
    for (size_t i = 0; i < numHistograms; i++)
      BuildAndStoreHuffmanTree(w, histograms[i].data(), histogramLength, &depths[i * histogramLength], &bits[i * histogramLength]);
  }
  void storeSymbol(BitWriter& w, uint32_t symbol) {
    if (blockLen == 0 && hasSplitCode && blockIdx + 1 < numBlocks) {
      blockIdx++;
      uint32_t bt = blockTypes[blockIdx];
      blockLen = blockLengths[blockIdx];
      entropyIdx = bt * histogramLength;
      StoreBlockSwitch(w, splitCode, blockLen, bt, false);
    }
    blockLen--;
    size_t ix = entropyIdx + symbol;
    w.writeBits(depths[ix], bits[ix]);
  }
  void storeSymbolWithContext(BitWriter& w, uint32_t symbol, uint32_t context, const uint32_t* contextMap, unsigned contextBits) {
    if (blockLen == 0 && hasSplitCode && blockIdx + 1 < numBlocks) {
      blockIdx++;
      uint32_t bt = blockTypes[blockIdx];
      blockLen = blockLengths[blockIdx];
      entropyIdx = (size_t)bt << contextBits;
      StoreBlockSwitch(w, splitCode, blockLen, bt, false);
    }
    blockLen--;
    size_t ix = (size_t)contextMap[entropyIdx + context] * histogramLength + symbol;
    w.writeBits(depths[ix], bits[ix]);
  }
};

// metablock.ts:763-797
static inline uint32_t ClusterAndBuildContextMap(const std::vector<std::vector<uint32_t>>& histograms, size_t numHistograms,
                                                 size_t alphabetSize, std::vector<uint32_t>& contextMap) {
  if (numHistograms <= 1) { contextMap[0] = 0; return 1; }
  std::vector<ClusterHistogram> in; in.reserve(numHistograms);
  for (size_t h = 0; h < numHistograms; h++) {
    ClusterHistogram ch(alphabetSize);
    for (size_t i = 0; i < alphabetSize; i++) { ch.data[i] = histograms[h][i]; ch.totalCount += histograms[h][i]; }
    ch.bitCost = ch.cost();
    in.push_back(std::move(ch));
  }
  std::vector<ClusterHistogram> out(numHistograms, ClusterHistogram(alphabetSize));
  ClusterHistograms(in, numHistograms, 64, out, contextMap.data());
  uint32_t maxCluster = 0;
  for (size_t i = 0; i < numHistograms; i++) if (contextMap[i] > maxCluster) maxCluster = contextMap[i];
  return maxCluster + 1;
}
// metablock.ts:799-819
static inline std::vector<std::vector<uint32_t>> BuildClusteredHistograms(const std::vector<std::vector<uint32_t>>& histograms,
                                                                          const std::vector<uint32_t>& contextMap,
                                                                          uint32_t numClusters, size_t alphabetSize) {
  std::vector<std::vector<uint32_t>> result(numClusters, std::vector<uint32_t>(alphabetSize, 0));
  for (size_t i = 0; i < histograms.size(); i++)
    for (size_t j = 0; j < alphabetSize; j++) result[contextMap[i]][j] += histograms[i][j];
  return result;
}

struct MetaBlockStats { bool trivial = true; uint32_t litTypes = 1, cmdTypes = 1, distTypes = 1, litClusters = 1, distClusters = 1; int contextMode = 0; };

// Everything storeMetaBlock decides before it writes a bit (metablock.ts:527-661): the three block
// splits, the context mode, the per-(block type, context) histograms and the two context maps.
// Stage-parity tests hand this to the GPU writer so that both sides emit from identical inputs.
struct MetaBlockModel {
  BlockSplit literalSplit, commandSplit, distanceSplit;
  int contextMode = 0;
  std::vector<std::vector<uint32_t>> literalHistograms, commandHistograms, distanceHistograms;
  std::vector<uint32_t> literalContextMap, distanceContextMap;
  uint32_t numLiteralClusters = 1, numDistanceClusters = 1;
};

// metablock.ts:558-637: per-context histograms of the commands under the model's splits and context mode.
static inline void BuildContextHistograms(const uint8_t* input, size_t startPos, size_t length, size_t mask,
                                          const std::vector<Command>& commands, uint32_t distAlphabetSize,
                                          uint8_t prev1, uint8_t prev2, MetaBlockModel& M) {
  const BlockSplit& literalSplit = M.literalSplit; const BlockSplit& commandSplit = M.commandSplit; const BlockSplit& distanceSplit = M.distanceSplit;
  const uint8_t* lut = GetContextLut(M.contextMode);
  size_t numLiteralContexts = (size_t)literalSplit.numTypes * 64;
  M.literalHistograms.assign(numLiteralContexts, std::vector<uint32_t>(256, 0));
  M.commandHistograms.assign(commandSplit.numTypes, std::vector<uint32_t>(704, 0));
  size_t numDistanceContexts = (size_t)distanceSplit.numTypes * 4;
  M.distanceHistograms.assign(numDistanceContexts, std::vector<uint32_t>(distAlphabetSize, 0));

  size_t pos = startPos;
  size_t litBlockIdx = 0; uint32_t litBlockLen = literalSplit.numBlocks > 0 ? literalSplit.lengths[0] : (uint32_t)length;
  uint32_t litBlockType = literalSplit.numBlocks > 0 ? literalSplit.types[0] : 0; uint32_t litCount = 0;
  size_t cmdBlockIdx = 0; uint32_t cmdBlockLen = commandSplit.numBlocks > 0 ? commandSplit.lengths[0] : (uint32_t)commands.size();
  uint32_t cmdBlockType = commandSplit.numBlocks > 0 ? commandSplit.types[0] : 0; uint32_t cmdCount = 0;
  size_t distBlockIdx = 0; uint32_t distBlockLen = distanceSplit.numBlocks > 0 ? distanceSplit.lengths[0] : (uint32_t)commands.size();
  uint32_t distBlockType = distanceSplit.numBlocks > 0 ? distanceSplit.types[0] : 0; uint32_t distCount = 0;
  uint8_t prevByte1 = prev1, prevByte2 = prev2;
  for (const Command& cmd : commands) {
    while (cmdCount >= cmdBlockLen && cmdBlockIdx + 1 < commandSplit.numBlocks) {
      cmdBlockIdx++; cmdBlockType = commandSplit.types[cmdBlockIdx]; cmdBlockLen = commandSplit.lengths[cmdBlockIdx]; cmdCount = 0;
    }
    M.commandHistograms[cmdBlockType][cmd.cmdPrefix]++;
    cmdCount++;
    for (uint32_t j = 0; j < cmd.insertLen; j++) {
      while (litCount >= litBlockLen && litBlockIdx + 1 < literalSplit.numBlocks) {
        litBlockIdx++; litBlockType = literalSplit.types[litBlockIdx]; litBlockLen = literalSplit.lengths[litBlockIdx]; litCount = 0;
      }
      uint8_t literal = input[(pos + j) & mask];
      uint32_t context = GetContext(prevByte1, prevByte2, lut);
      M.literalHistograms[litBlockType * 64 + context][literal]++;
      litCount++;
      prevByte2 = prevByte1; prevByte1 = literal;
    }
    pos += cmd.insertLen;
    uint32_t copyLen = CommandCopyLen(cmd);
    if (copyLen && cmd.cmdPrefix >= 128) {
      while (distCount >= distBlockLen && distBlockIdx + 1 < distanceSplit.numBlocks) {
        distBlockIdx++; distBlockType = distanceSplit.types[distBlockIdx]; distBlockLen = distanceSplit.lengths[distBlockIdx]; distCount = 0;
      }
      uint32_t distContext = copyLen > 4 ? 3 : copyLen - 2;
      M.distanceHistograms[distBlockType * 4 + distContext][cmd.distPrefix & 0x3FF]++;
      distCount++;
    }
    if (copyLen > 0) {
      size_t copyEnd = pos + copyLen;
      prevByte1 = input[(copyEnd - 1) & mask];
      prevByte2 = input[(copyEnd - 2) & mask];
    }
    pos += copyLen;
  }
}

// metablock.ts:527-661.  Returns false when the metablock takes the trivial writer (:518, :531).
static inline bool ComputeMetaBlockModel(const Bytes& data, size_t startPos, size_t length, size_t mask,
                                         const std::vector<Command>& commands, uint32_t distAlphabetSize, int quality,
                                         uint8_t prev1, uint8_t prev2, MetaBlockModel& M) {
  if (length < 128 || quality < 5 || commands.size() < 6) return false;
  SplitBlock(commands, data.p, startPos, mask, quality, M.literalSplit, M.commandSplit, M.distanceSplit);
  if (M.literalSplit.numTypes <= 1 && M.commandSplit.numTypes <= 1 && M.distanceSplit.numTypes <= 1) return false;
  M.contextMode = ChooseContextMode(data, startPos, std::min<size_t>(length, 4096));
  BuildContextHistograms(data.p, startPos, length, mask, commands, distAlphabetSize, prev1, prev2, M);
  size_t numLiteralContexts = (size_t)M.literalSplit.numTypes * 64, numDistanceContexts = (size_t)M.distanceSplit.numTypes * 4;
  M.literalContextMap.assign(numLiteralContexts, 0); M.distanceContextMap.assign(numDistanceContexts, 0);
  M.numLiteralClusters = ClusterAndBuildContextMap(M.literalHistograms, numLiteralContexts, 256, M.literalContextMap);
  M.numDistanceClusters = ClusterAndBuildContextMap(M.distanceHistograms, numDistanceContexts, distAlphabetSize, M.distanceContextMap);
  return true;
}

// metablock.ts:651-760: header, block-split codes, context maps, prefix codes and the emit loop, from a
// model whose histograms are already built.
static inline void EmitMetaBlock(BitWriter& w, const uint8_t* input, size_t startPos, size_t length, size_t mask, bool isLast,
                                 const std::vector<Command>& commands, uint32_t distAlphabetSize, uint32_t npostfix, uint32_t ndirect,
                                 uint8_t prev1, uint8_t prev2, const MetaBlockModel& M) {
  const BlockSplit& literalSplit = M.literalSplit; const BlockSplit& commandSplit = M.commandSplit; const BlockSplit& distanceSplit = M.distanceSplit;
  const uint8_t* lut = GetContextLut(M.contextMode);
  size_t numLiteralContexts = (size_t)literalSplit.numTypes * 64, numDistanceContexts = (size_t)distanceSplit.numTypes * 4;
  auto clusteredLit = BuildClusteredHistograms(M.literalHistograms, M.literalContextMap, M.numLiteralClusters, 256);
  auto clusteredDist = BuildClusteredHistograms(M.distanceHistograms, M.distanceContextMap, M.numDistanceClusters, distAlphabetSize);

  StoreCompressedMetaBlockHeader(w, isLast, (uint32_t)length);
  BlockEncoder literalEnc(256, literalSplit), commandEnc(704, commandSplit), distanceEnc(distAlphabetSize, distanceSplit);
  literalEnc.buildAndStoreEntropyCodes(w);
  commandEnc.buildAndStoreEntropyCodes(w);
  distanceEnc.buildAndStoreEntropyCodes(w);
  w.writeBits(2, npostfix);
  w.writeBits(4, ndirect >> npostfix);
  for (uint32_t i = 0; i < literalSplit.numTypes; i++) w.writeBits(2, M.contextMode);
  EncodeContextMap(w, M.literalContextMap.data(), numLiteralContexts, M.numLiteralClusters);
  EncodeContextMap(w, M.distanceContextMap.data(), numDistanceContexts, M.numDistanceClusters);
  literalEnc.buildAndStoreHuffmanTrees(w, clusteredLit, M.numLiteralClusters);
  commandEnc.buildAndStoreHuffmanTrees(w, M.commandHistograms, commandSplit.numTypes);
  distanceEnc.buildAndStoreHuffmanTrees(w, clusteredDist, M.numDistanceClusters);

  size_t pos = startPos; uint8_t prevByte1 = prev1, prevByte2 = prev2;
  for (const Command& cmd : commands) {
    commandEnc.storeSymbol(w, cmd.cmdPrefix);
    StoreCommandExtra(w, cmd);
    for (uint32_t j = 0; j < cmd.insertLen; j++) {
      uint8_t literal = input[(pos + j) & mask];
      uint32_t context = GetContext(prevByte1, prevByte2, lut);
      literalEnc.storeSymbolWithContext(w, literal, context, M.literalContextMap.data(), 6);
      prevByte2 = prevByte1; prevByte1 = literal;
    }
    pos += cmd.insertLen;
    uint32_t copyLen = CommandCopyLen(cmd);
    if (copyLen && cmd.cmdPrefix >= 128) {
      uint32_t distContext = copyLen > 4 ? 3 : copyLen - 2;
      distanceEnc.storeSymbolWithContext(w, cmd.distPrefix & 0x3FF, distContext, M.distanceContextMap.data(), 2);
      w.writeBits(cmd.distPrefix >> 10, cmd.distExtra);
    }
    if (copyLen > 0) {
      size_t copyEnd = pos + copyLen;
      prevByte1 = input[(copyEnd - 1) & mask];
      prevByte2 = input[(copyEnd - 2) & mask];
    }
    pos += copyLen;
  }
  if (isLast) w.alignToByte();
}

// metablock.ts:504-761.  prev1/prev2: literal-context bytes before startPos (0,0 in the TS;
// a stitched shard k>0 passes the true preceding bytes, SURVEY §8e rule 4).
static inline void StoreMetaBlock(BitWriter& w, const Bytes& data, size_t startPos, size_t length, size_t mask, bool isLast,
                                  const std::vector<Command>& commands, uint32_t distAlphabetSize, int quality,
                                  uint32_t npostfix, uint32_t ndirect, uint8_t prev1 = 0, uint8_t prev2 = 0,
                                  MetaBlockStats* stats = nullptr) {
  MetaBlockModel M;
  if (!ComputeMetaBlockModel(data, startPos, length, mask, commands, distAlphabetSize, quality, prev1, prev2, M)) {
    StoreMetaBlockTrivial(w, data.p, startPos, length, mask, isLast, commands, distAlphabetSize, npostfix, ndirect);
    return;
  }
  if (stats) {
    stats->trivial = false; stats->litTypes = M.literalSplit.numTypes; stats->cmdTypes = M.commandSplit.numTypes;
    stats->distTypes = M.distanceSplit.numTypes; stats->litClusters = M.numLiteralClusters; stats->distClusters = M.numDistanceClusters;
    stats->contextMode = M.contextMode;
  }
  EmitMetaBlock(w, data.p, startPos, length, mask, isLast, commands, distAlphabetSize, npostfix, ndirect, prev1, prev2, M);
}

// metablock.ts:821-855 (mask = all ones: the TS passes input.length-1)
static inline void StoreUncompressedMetaBlock(BitWriter& w, const uint8_t* input, size_t position, size_t length, bool isFinal) {
  StoreUncompressedMetaBlockHeader(w, (uint32_t)length);
  w.alignToByte();
  w.writeBytes(input + position, length);
  w.prepareStorage();
  if (isFinal) { w.writeBits(1, 1); w.writeBits(1, 1); w.alignToByte(); }
}

}  // namespace oracle
