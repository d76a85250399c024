// ORACLE (test infrastructure) — restates bit-cost.ts:14-135 and cluster.ts.
#pragma once
#include "common.h"

namespace oracle {

// bit-cost.ts:14-38
static inline double BitsEntropy(const uint32_t* histogram, size_t size) {
  double sum = 0, retval = 0;
  for (size_t i = 0; i < size; i++) {
    uint32_t p = histogram[i];
    if (p > 0) { sum += p; retval -= (double)p * FastLog2(p); }
  }
  if (sum > 0) retval += sum * FastLog2(sum);
  if (retval < sum) retval = sum;
  return retval;
}

// bit-cost.ts:44-135.  Math.round == floor(x + 0.5) for the non-negative values seen here.
static inline double PopulationCost(const uint32_t* data, size_t dataSize, double totalCount) {
  if (totalCount == 0) return 12;
  size_t s[5]; int count = 0;
  for (size_t i = 0; i < dataSize && count <= 4; i++) if (data[i] > 0) s[count++] = i;
  if (count == 1) return 12;
  if (count == 2) return 20 + totalCount;
  if (count == 3) {
    double h0 = data[s[0]], h1 = data[s[1]], h2 = data[s[2]];
    double hmax = std::max(h0, std::max(h1, h2));
    return 28 + 2 * (h0 + h1 + h2) - hmax;
  }
  if (count == 4) {
    double h[4] = {(double)data[s[0]], (double)data[s[1]], (double)data[s[2]], (double)data[s[3]]};
    std::sort(h, h + 4, [](double a, double b) { return a > b; });
    double h23 = h[2] + h[3];
    double hmax = std::max(h23, h[0]);
    return 37 + 3 * h23 + 2 * (h[0] + h[1]) - hmax;
  }
  double bits = 0; int maxDepth = 1;
  uint32_t depthHisto[18] = {0};
  double log2total = FastLog2(totalCount);
  for (size_t i = 0; i < dataSize;) {
    if (data[i] > 0) {
      double log2p = log2total - FastLog2(data[i]);
      int depth = (int)std::floor(log2p + 0.5);
      bits += (double)data[i] * log2p;
      if (depth > 15) depth = 15;
      if (depth > maxDepth) maxDepth = depth;
      // a negative depth cannot occur: data[i] <= totalCount for every caller
      depthHisto[depth < 0 ? 0 : depth]++;
      i++;
    } else {
      uint32_t reps = 1;
      for (size_t k = i + 1; k < dataSize && data[k] == 0; k++) reps++;
      i += reps;
      if (i == dataSize) break;
      if (reps < 3) depthHisto[0] += reps;
      else {
        reps -= 2;
        while (reps > 0) { depthHisto[17]++; bits += 3; reps >>= 3; }
      }
    }
  }
  bits += 18 + 2 * maxDepth;
  bits += BitsEntropy(depthHisto, 18);
  return bits;
}

// cluster.ts:7-11
struct ClusterHistogram {
  std::vector<uint32_t> data; double totalCount = 0; double bitCost = 0;
  explicit ClusterHistogram(size_t n = 0) : data(n, 0) {}
  void clear() { std::fill(data.begin(), data.end(), 0); totalCount = 0; bitCost = 0; }
  void add(const ClusterHistogram& b) {           // cluster.ts:54-59
    for (size_t i = 0; i < data.size(); i++) data[i] += b.data[i];
    totalCount += b.totalCount;
  }
  double cost() const { return PopulationCost(data.data(), data.size(), totalCount); }  // :61-63
};
struct HistogramPair { uint32_t idx1, idx2; double costCombo, costDiff; };

// cluster.ts:20-23
static inline double ClusterCostDiff(double a, double b) {
  double c = a + b;
  return c * FastLog2(c) - a * FastLog2(a) - b * FastLog2(b);
}
// cluster.ts:26-31
static inline bool HistogramPairIsLess(const HistogramPair& a, const HistogramPair& b) {
  if (a.costDiff != b.costDiff) return a.costDiff > b.costDiff;
  return (a.idx2 - a.idx1) > (b.idx2 - b.idx1);
}
// cluster.ts:65-132.  maxNumPairs is a double on purpose (cluster.ts:361 can yield x.5).
static inline void CompareAndPushToQueue(std::vector<ClusterHistogram>& out, ClusterHistogram& tmp,
                                         const uint32_t* clusterSize, uint32_t idx1, uint32_t idx2,
                                         double maxNumPairs, std::vector<HistogramPair>& pairs, size_t& numPairs) {
  if (idx1 == idx2) return;
  if (idx2 < idx1) std::swap(idx1, idx2);
  HistogramPair p; p.idx1 = idx1; p.idx2 = idx2; p.costCombo = 0;
  p.costDiff = 0.5 * ClusterCostDiff(clusterSize[idx1], clusterSize[idx2]);
  p.costDiff -= out[idx1].bitCost;
  p.costDiff -= out[idx2].bitCost;
  bool good = false;
  if (out[idx1].totalCount == 0) { p.costCombo = out[idx2].bitCost; good = true; }
  else if (out[idx2].totalCount == 0) { p.costCombo = out[idx1].bitCost; good = true; }
  else {
    double threshold = numPairs == 0 ? 1e99 : std::max(0.0, pairs[0].costDiff);
    tmp.data = out[idx1].data; tmp.totalCount = out[idx1].totalCount; tmp.bitCost = out[idx1].bitCost;
    tmp.add(out[idx2]);
    double costCombo = tmp.cost();
    if (costCombo < threshold - p.costDiff) { p.costCombo = costCombo; good = true; }
  }
  if (good) {
    p.costDiff += p.costCombo;
    if (pairs.size() < numPairs + 2) pairs.resize(numPairs + 2);
    if (numPairs > 0 && HistogramPairIsLess(pairs[0], p)) {
      if ((double)numPairs < maxNumPairs) { pairs[numPairs] = pairs[0]; numPairs++; }
      pairs[0] = p;
    } else if ((double)numPairs < maxNumPairs) {
      pairs[numPairs] = p; numPairs++;
    }
  }
}
// cluster.ts:135-228.  symbols/clusters are views (pointer + implicit length) as in the TS subarrays.
static inline size_t HistogramCombine(std::vector<ClusterHistogram>& out, ClusterHistogram& tmp, uint32_t* clusterSize,
                                      uint32_t* symbols, uint32_t* clusters, std::vector<HistogramPair>& pairs,
                                      size_t numClusters, size_t symbolsSize, size_t maxClusters, double maxNumPairs) {
  double costDiffThreshold = 0.0; size_t minClusterSize = 1; size_t numPairs = 0;
  for (size_t i1 = 0; i1 < numClusters; i1++)
    for (size_t i2 = i1 + 1; i2 < numClusters; i2++)
      CompareAndPushToQueue(out, tmp, clusterSize, clusters[i1], clusters[i2], maxNumPairs, pairs, numPairs);
  while (numClusters > minClusterSize) {
    if (pairs[0].costDiff >= costDiffThreshold) { costDiffThreshold = 1e99; minClusterSize = maxClusters; continue; }
    uint32_t b1 = pairs[0].idx1, b2 = pairs[0].idx2;
    out[b1].add(out[b2]);
    out[b1].bitCost = pairs[0].costCombo;
    clusterSize[b1] += clusterSize[b2];
    for (size_t i = 0; i < symbolsSize; i++) if (symbols[i] == b2) symbols[i] = b1;
    for (size_t i = 0; i < numClusters; i++) {
      if (clusters[i] == b2) {
        for (size_t j = i; j + 1 < numClusters; j++) clusters[j] = clusters[j + 1];
        break;
      }
    }
    numClusters--;
    size_t copyTo = 0;
    for (size_t i = 0; i < numPairs; i++) {
      HistogramPair p = pairs[i];
      if (p.idx1 == b1 || p.idx2 == b1 || p.idx1 == b2 || p.idx2 == b2) continue;
      if (HistogramPairIsLess(pairs[0], p)) { HistogramPair front = pairs[0]; pairs[0] = p; pairs[copyTo] = front; }
      else pairs[copyTo] = p;
      copyTo++;
    }
    numPairs = copyTo;
    for (size_t i = 0; i < numClusters; i++)
      CompareAndPushToQueue(out, tmp, clusterSize, b1, clusters[i], maxNumPairs, pairs, numPairs);
  }
  return numClusters;
}
// cluster.ts:230-242
static inline double HistogramBitCostDistance(const ClusterHistogram& h, const ClusterHistogram& cand, ClusterHistogram& tmp) {
  if (h.totalCount == 0) return 0.0;
  tmp.data = h.data; tmp.totalCount = h.totalCount; tmp.bitCost = h.bitCost;
  tmp.add(cand);
  return tmp.cost() - cand.bitCost;
}
// cluster.ts:244-275
static inline void HistogramRemap(const std::vector<ClusterHistogram>& in, size_t inSize, const uint32_t* clusters, size_t numClusters,
                                  std::vector<ClusterHistogram>& out, ClusterHistogram& tmp, uint32_t* symbols) {
  for (size_t i = 0; i < inSize; i++) {
    uint32_t bestOut = i == 0 ? symbols[0] : symbols[i - 1];
    double bestBits = HistogramBitCostDistance(in[i], out[bestOut], tmp);
    for (size_t j = 0; j < numClusters; j++) {
      double cur = HistogramBitCostDistance(in[i], out[clusters[j]], tmp);
      if (cur < bestBits) { bestBits = cur; bestOut = clusters[j]; }
    }
    symbols[i] = bestOut;
  }
  for (size_t i = 0; i < numClusters; i++) out[clusters[i]].clear();
  for (size_t i = 0; i < inSize; i++) out[symbols[i]].add(in[i]);
}
// cluster.ts:278-314
static inline size_t HistogramReindex(std::vector<ClusterHistogram>& out, uint32_t* symbols, size_t length) {
  const uint32_t INVALID = 0xFFFFFFFFu;
  std::vector<uint32_t> newIndex(length, INVALID);
  uint32_t next = 0;
  for (size_t i = 0; i < length; i++) if (newIndex[symbols[i]] == INVALID) newIndex[symbols[i]] = next++;
  std::vector<ClusterHistogram> tmp(next);
  next = 0;
  for (size_t i = 0; i < length; i++) {
    if (newIndex[symbols[i]] == next) { tmp[next] = out[symbols[i]]; next++; }
    symbols[i] = newIndex[symbols[i]];
  }
  for (size_t i = 0; i < tmp.size(); i++) out[i] = tmp[i];
  return tmp.size();
}
// cluster.ts:317-378.  `out` must hold inSize histograms; symbols holds inSize entries.
static inline size_t ClusterHistograms(const std::vector<ClusterHistogram>& in, size_t inSize, size_t maxHistograms,
                                       std::vector<ClusterHistogram>& out, uint32_t* symbols) {
  size_t dataSize = in[0].data.size();
  std::vector<uint32_t> clusterSize(inSize), clusters(inSize);
  const size_t maxInput = 64;
  const size_t pairsCapacity = maxInput * maxInput / 2;
  std::vector<HistogramPair> pairs(pairsCapacity + 1);
  ClusterHistogram tmp(dataSize);
  for (size_t i = 0; i < inSize; i++) {
    clusterSize[i] = 1;
    out[i] = in[i];
    out[i].bitCost = in[i].cost();
    symbols[i] = (uint32_t)i;
  }
  size_t numClusters = 0;
  for (size_t i = 0; i < inSize; i += maxInput) {
    size_t numToCombine = std::min(inSize - i, maxInput);
    for (size_t j = 0; j < numToCombine; j++) clusters[numClusters + j] = (uint32_t)(i + j);
    size_t nn = HistogramCombine(out, tmp, clusterSize.data(), symbols + i, clusters.data() + numClusters, pairs,
                                 numToCombine, numToCombine, maxHistograms, (double)pairsCapacity);
    numClusters += nn;
  }
  double maxNumPairs = std::min(64.0 * (double)numClusters, ((double)numClusters / 2.0) * (double)numClusters);
  numClusters = HistogramCombine(out, tmp, clusterSize.data(), symbols, clusters.data(), pairs,
                                 numClusters, inSize, maxHistograms, maxNumPairs);
  HistogramRemap(in, inSize, clusters.data(), numClusters, out, tmp, symbols);
  return HistogramReindex(out, symbols, inSize);
}

}  // namespace oracle
