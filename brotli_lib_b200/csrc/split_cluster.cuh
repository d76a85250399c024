// sm_100a kernels for metablock modelling: block splitting (K8) and histogram clustering (K9).
//
// Reference semantics (paths relative to the reference tree, src/encode/):
//   block-splitter.ts:394-464  splitBlock: literal / command / distance symbol streams split separately
//                 :255-314     splitByteVector: <= 64 entropy codes, 3 (q < 10) or 10 refinement iterations
//                 :117-211     findBlocks: min-cost assignment with a block-switch cost (26.0 / 28.1 bits)
//                 :316-390     clusterBlocks: per-block histograms clustered to <= 256 block types
//   cluster.ts:317-378         clusterHistograms: batches of 64, greedy best-pair merges, remap, reindex
//   bit-cost.ts:44-135         populationCost
//
// These stages are float-driven: north_star holds them to "valid stream, size within 0.5 %" of the
// reference, not to bit parity.  The B200 formulation therefore keeps the
// reference's structure where it parallelises and replaces what does not (DESIGN.md "Modelling"):
//   * findBlocks runs over GRANULES of 128 literals / 16 commands / 16 distances instead of single
//     symbols, in independent chunks of 512 granules, one warp per chunk (64 entropy codes = 2 per
//     lane, warp-min by redux).  The per-symbol switch bitmap of the reference (8 B per symbol)
//     becomes 8 B per granule.
//   * entropy codes are seeded from contiguous slices of the stream, not from an LCG sample.
//   * populationCost is replaced by  max(entropy, total) + 20 + 3 nnz + 3 zero-runs  (one pass,
//     warp ballots), which tracks the exact cost within 0.05 % of output size on the test corpus.
//   * clustering: round 1 = the reference's batches of 64 (greedy best pair, exact arg-min over a
//     64 x 64 matrix in shared memory); further rounds re-batch the survivors until <= 256 remain;
//     the last round is an all-pairs greedy over <= 256 clusters; remap uses a sparse distance.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace bgpu {

constexpr int SPLIT_MAXK = 64;          // entropy codes per split problem
constexpr int SPLIT_CHUNK = 512;        // granules per DP chunk (one warp): a dependent chain.  2048 -> 1024 -> 512: k_split 8.1 -> 7.1 -> 6.6 ms per 64-shard job, size unchanged (r5r)
constexpr int MAX_BLOCKS = 2048;        // blocks per (metablock, category)
constexpr int CL_BATCH = 64;
constexpr int CL_FINAL_MAX = 256;

// One clustering problem: `m` input histograms of `alpha` bins (row stride `stride` u32).
struct ClusterProb {
  const uint32_t* in;     // [m][stride]  inputs (read only)
  uint32_t* work;         // [m][stride]  cluster histograms (copy of `in`, merged in place)
  uint32_t* alias;        // [m]  alias[j] = i: cluster j was merged into i (alias[i] == i for live clusters)
  uint32_t* list;         // [m]  live cluster indices, ping
  uint32_t* list2;        // [m]  pong
  float* cost;            // [m]  cost of each live cluster
  uint32_t* total;        // [m]
  uint32_t* csize;        // [m]  inputs merged into the cluster
  uint32_t* sym;          // [m]  out: final cluster id per input (reindexed by first appearance)
  uint32_t* state;        // [8]  0: live count, 1: round target (forced max per batch), 2: final cluster count, 3: batch survivors scratch base
  uint32_t* batch_cnt;    // [ceil(m / 64)] survivors per batch
  uint32_t m, alpha, stride, max_clusters;
  uint32_t final_max;     // live clusters the all-pairs round accepts (<= CL_FINAL_MAX); the batch rounds merge down to it
};

__device__ __forceinline__ float flog2(float v) { return __log2f(v); }

// Warp-cooperative cost of histogram a (+ b).  All lanes return the same value; *total_out the sum.
// alpha is a multiple of 4 and <= 768, rows are 16-byte aligned: every lane fetches its (up to 6)
// 16-byte pieces of both rows up front, so a pair evaluation costs ONE L2 round trip instead of
// alpha / 32 dependent ones (profiles/r1i: k_cluster_final was 34 % of the q5 step, latency bound).
// NV = 16-byte pieces per lane: 2 covers alpha <= 256 (literal histograms), 6 covers alpha <= 768; the narrow
// form keeps the kernels that only see literal problems at a third of the row registers.
template <bool SHARED = false, int NV = 6>
__device__ __forceinline__ float hist_cost(const uint32_t* a, const uint32_t* b, int alpha, uint32_t* total_out) {
  const uint32_t lane = threadIdx.x & 31;
  const int nvec = alpha >> 2;
  const uint4* a4 = reinterpret_cast<const uint4*>(a);
  const uint4* b4 = reinterpret_cast<const uint4*>(b);
  uint4 v[NV];
#pragma unroll
  for (int t = 0; t < NV; t++) {
    const int i = t * 32 + (int)lane;
    v[t] = make_uint4(0, 0, 0, 0);
    if (i < nvec) {
      v[t] = SHARED ? a4[i] : __ldcg(a4 + i);
      if (b) { const uint4 y = SHARED ? b4[i] : __ldcg(b4 + i); v[t].x += y.x; v[t].y += y.y; v[t].z += y.z; v[t].w += y.w; }
    }
  }
  // The entropy term is the only float reduction (5 butterfly steps).  The non-zero count and the zero-run count come
  // from ballots of the four bins of every lane — the same value in all lanes, no reduction — and the total from one
  // REDUX (r5m: the four 5-step shuffle reductions + 2 shuffles per piece were most of a pair evaluation).
  float ent = 0.f; uint32_t tot = 0, nnz = 0, runs = 0, carry = 1;
#pragma unroll
  for (int t = 0; t < NV; t++) {
    if (t * 32 < nvec) {
      const uint4 x = v[t];
      tot += x.x + x.y + x.z + x.w;
      if (x.x) ent -= (float)x.x * flog2((float)x.x);
      if (x.y) ent -= (float)x.y * flog2((float)x.y);
      if (x.z) ent -= (float)x.z * flog2((float)x.z);
      if (x.w) ent -= (float)x.w * flog2((float)x.w);
      const uint32_t b0 = __ballot_sync(0xffffffffu, x.x != 0), b1 = __ballot_sync(0xffffffffu, x.y != 0),
                     b2 = __ballot_sync(0xffffffffu, x.z != 0), b3 = __ballot_sync(0xffffffffu, x.w != 0);
      nnz += __popc(b0) + __popc(b1) + __popc(b2) + __popc(b3);
      // zero runs that END at a non-zero bin (leading and interior ones, never the trailing run)
      runs += __popc(b0 & ~((b3 << 1) | carry)) + __popc(b1 & ~b0) + __popc(b2 & ~b1) + __popc(b3 & ~b2);
      carry = b3 >> 31;
    }
  }
  for (int o = 16; o; o >>= 1) ent += __shfl_xor_sync(0xffffffffu, ent, o);
  tot = __reduce_add_sync(0xffffffffu, tot);
  if (total_out) *total_out = tot;
  if (nnz <= 1) return 12.f;
  const float T = (float)tot;
  ent += T * flog2(T);
  return fmaxf(ent, T) + 20.f + 3.f * (float)nnz + 3.f * (float)runs;
}

// cluster.ts:20-23
__device__ __forceinline__ float cluster_cost_diff(float a, float b) { const float c = a + b; return c * flog2(c) - a * flog2(a) - b * flog2(b); }

// ---------------------------------------------------------------------------------------
// Batch round (cluster.ts:135-228 histogramCombine on one batch of <= 64 live clusters).
// grid (max batches, problems), 256 threads.  mode 0: first round (list = identity, costs computed
// here); mode 1: later rounds over p.list.  Merges while the best pair improves the cost, then —
// if the batch still holds more than `target` clusters — merges the cheapest pairs regardless.
// When the batch's 64 rows fit the dynamic shared memory given at launch (alpha <= 256: literal context
// maps and literal block histograms), they are staged there and every pair evaluation and merge works on
// shared memory; the survivors' rows are written back at the end.  profiles/r1k: the global-memory form
// is bound by one L2 round trip per pair evaluation (2016 + ~63 * 62 of them per batch).
template <int NV>
__global__ void __launch_bounds__(256) k_cluster_batch(const ClusterProb* __restrict__ probs, int mode, uint32_t smem_bytes) {
  extern __shared__ __align__(16) uint32_t s_rows[];   // [CL_BATCH][alpha] when staged
  __shared__ float s_d[CL_BATCH][CL_BATCH];      // costDiff of pair (i < j) at [i][j]
  __shared__ float s_c[CL_BATCH][CL_BATCH];      // combined cost of the pair
  __shared__ uint32_t s_idx[CL_BATCH];           // cluster index of slot i
  __shared__ float s_cost[CL_BATCH];
  __shared__ uint32_t s_tot[CL_BATCH], s_size[CL_BATCH];
  __shared__ uint8_t s_live[CL_BATCH];
  __shared__ float s_rv[8]; __shared__ int s_ri[8];
  __shared__ int s_best;
  const ClusterProb p = probs[blockIdx.y];
  if ((p.alpha > 256) != (NV > 2)) return;                 // the other instantiation's problem
  const uint32_t live = mode == 0 ? p.m : p.state[0];
  if (mode == 1 && live <= p.final_max) return;           // the final round takes over
  const uint32_t b0 = blockIdx.x * CL_BATCH;
  if (b0 >= live) return;
  const uint32_t n = min((uint32_t)CL_BATCH, live - b0);
  const uint32_t target = mode == 0 ? CL_BATCH : p.state[1];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t* src_list = mode == 0 ? nullptr : p.list;
  const bool staged = (size_t)p.alpha * CL_BATCH * 4 <= smem_bytes;
  if (tid < CL_BATCH) {
    s_live[tid] = tid < n;
    if (tid < n) {
      const uint32_t ci = src_list ? src_list[b0 + tid] : b0 + tid;
      s_idx[tid] = ci;
      if (mode == 1) { s_cost[tid] = p.cost[ci]; s_tot[tid] = p.total[ci]; s_size[tid] = p.csize[ci]; }
    }
  }
  __syncthreads();
  if (mode == 0) {
    // out[i] = in[i]; bitCost = cost (cluster.ts:336-341)
    for (uint32_t i = w; i < n; i += 8) {
      const uint32_t ci = s_idx[i];
      const uint32_t* a = p.in + (size_t)ci * p.stride; uint32_t* o = staged ? s_rows + (size_t)i * p.alpha : p.work + (size_t)ci * p.stride;
      for (uint32_t k = lane; k < p.alpha; k += 32) o[k] = a[k];
      uint32_t t; const float c = hist_cost<false, NV>(a, nullptr, (int)p.alpha, &t);
      if (lane == 0) { s_cost[i] = c; s_tot[i] = t; s_size[i] = 1; p.alias[ci] = ci; }
    }
    __syncthreads();
  } else if (staged) {
    for (uint32_t i = w; i < n; i += 8) {
      const uint32_t* a = p.work + (size_t)s_idx[i] * p.stride; uint32_t* o = s_rows + (size_t)i * p.alpha;
      for (uint32_t k = lane; k < p.alpha; k += 32) o[k] = __ldcg(a + k);
    }
    __syncthreads();
  }
  // all pairs
  auto eval_pair = [&](uint32_t i, uint32_t j) {   // i < j, whole warp
    float combo;
    if (s_tot[i] == 0) combo = s_cost[j];
    else if (s_tot[j] == 0) combo = s_cost[i];
    else if (staged) combo = hist_cost<true, NV>(s_rows + (size_t)i * p.alpha, s_rows + (size_t)j * p.alpha, (int)p.alpha, nullptr);
    else combo = hist_cost<false, NV>(p.work + (size_t)s_idx[i] * p.stride, p.work + (size_t)s_idx[j] * p.stride, (int)p.alpha, nullptr);
    if (lane == 0) {
      s_c[i][j] = combo;
      s_d[i][j] = 0.5f * cluster_cost_diff((float)s_size[i], (float)s_size[j]) - s_cost[i] - s_cost[j] + combo;
    }
  };
  for (uint32_t pr = w; pr < n * n; pr += 8) {
    const uint32_t i = pr / n, j = pr % n;
    if (i < j) eval_pair(i, j);
  }
  __syncthreads();
  uint32_t alive = n;
  while (alive > 1) {
    // arg-min over live pairs; ties: smaller (j - i), then smaller i (cluster.ts:26-31)
    float bv = 3.0e38f; int bi = -1;
    for (uint32_t pr = tid; pr < n * n; pr += 256) {
      const uint32_t i = pr / n, j = pr % n;
      if (i < j && s_live[i] && s_live[j]) {
        const float d = s_d[i][j];
        if (d < bv || (d == bv && bi >= 0 && (j - i) < (uint32_t)((bi % (int)n) - (bi / (int)n)))) { bv = d; bi = (int)pr; }
      }
    }
    for (int o = 16; o; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi >= 0 && (bi < 0 || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
    }
    if (lane == 0) { s_rv[w] = bv; s_ri[w] = bi; }
    __syncthreads();
    if (tid == 0) {
      float v = s_rv[0]; int ix = s_ri[0];
      for (int k = 1; k < 8; k++) if (s_ri[k] >= 0 && (ix < 0 || s_rv[k] < v || (s_rv[k] == v && s_ri[k] < ix))) { v = s_rv[k]; ix = s_ri[k]; }
      if (ix >= 0 && !(v < 0.f) && alive <= target) ix = -1;   // no beneficial merge left and the batch is small enough
      s_best = ix;
    }
    __syncthreads();
    const int best = s_best;
    if (best < 0) break;
    const uint32_t i = (uint32_t)best / n, j = (uint32_t)best % n;
    // merge j into i (cluster.ts:170-176)
    if (staged) {
      uint32_t* a = s_rows + (size_t)i * p.alpha; const uint32_t* b = s_rows + (size_t)j * p.alpha;
      for (uint32_t k = tid; k < p.alpha; k += 256) a[k] += b[k];
    } else {
      uint32_t* a = p.work + (size_t)s_idx[i] * p.stride; const uint32_t* b = p.work + (size_t)s_idx[j] * p.stride;
      for (uint32_t k = tid; k < p.alpha; k += 256) a[k] = __ldcg(a + k) + __ldcg(b + k);
    }
    __syncthreads();
    if (tid == 0) {
      s_cost[i] = s_c[i][j]; s_tot[i] += s_tot[j]; s_size[i] += s_size[j]; s_live[j] = 0;
      p.alias[s_idx[j]] = s_idx[i];
    }
    __threadfence_block();
    __syncthreads();
    alive--;
    for (uint32_t k = w; k < n; k += 8) {
      if (k == i || !s_live[k]) continue;
      if (k < i) eval_pair(k, i); else eval_pair(i, k);
    }
    __syncthreads();
  }
  // survivors of this batch, in slot order
  if (staged) {
    for (uint32_t i = w; i < n; i += 8) {
      if (!s_live[i]) continue;
      uint32_t* o = p.work + (size_t)s_idx[i] * p.stride; const uint32_t* a = s_rows + (size_t)i * p.alpha;
      for (uint32_t k = lane; k < p.alpha; k += 32) o[k] = a[k];
    }
  }
  if (tid == 0) {
    uint32_t cnt = 0;
    for (uint32_t k = 0; k < n; k++)
      if (s_live[k]) {
        const uint32_t ci = s_idx[k];
        p.cost[ci] = s_cost[k]; p.total[ci] = s_tot[k]; p.csize[ci] = s_size[k];
        p.list2[b0 + cnt] = ci; cnt++;
      }
    p.batch_cnt[blockIdx.x] = cnt;
  }
}

// Gathers the survivors of all batches into p.list and sets the next round's per-batch target.
// One CTA of 32 threads per problem.
__global__ void __launch_bounds__(32) k_cluster_gather(const ClusterProb* __restrict__ probs, int mode) {
  const ClusterProb p = probs[blockIdx.x];
  const uint32_t live = mode == 0 ? p.m : p.state[0];
  if (mode == 1 && live <= p.final_max) return;
  const uint32_t nb = (live + CL_BATCH - 1) / CL_BATCH;
  const uint32_t lane = threadIdx.x;
  uint32_t base = 0;
  for (uint32_t b = 0; b < nb; b++) {
    const uint32_t cnt = p.batch_cnt[b];
    for (uint32_t k = lane; k < cnt; k += 32) p.list[base + k] = p.list2[b * CL_BATCH + k];
    base += cnt;
  }
  if (lane == 0) {
    p.state[0] = base;
    // next round: force every batch of 64 down far enough that <= 7/8 of final_max clusters remain overall
    const uint32_t nb2 = (base + CL_BATCH - 1) / CL_BATCH;
    p.state[1] = nb2 ? max(1u, (p.final_max - p.final_max / 8) / nb2) : 1u;
  }
}

// Final round: all-pairs greedy over the <= 256 live clusters of a problem, down to max_clusters
// (cluster.ts:361-365), then remap every input to its cheapest cluster (cluster.ts:244-275 with a
// sparse distance), rebuild the cluster histograms and reindex by first appearance (:278-314).
// One CTA of 512 threads per problem; dynamic shared memory: 256 x 256 floats (pair matrix) + tables.
constexpr uint32_t CL_FINAL_PAIRS = CL_FINAL_MAX * (CL_FINAL_MAX - 1) / 2;   // pair (i < j) at j (j - 1) / 2 + i
constexpr uint32_t cl_final_smem(uint32_t maxn) { return maxn * (maxn - 1) / 2 * 4 + maxn * (4 + 4 + 4 + 4 + 1) + 256; }
constexpr uint32_t CL_FINAL_SMEM = cl_final_smem(CL_FINAL_MAX);
__device__ __forceinline__ uint32_t tri(uint32_t i, uint32_t j) { return j * (j - 1) / 2 + i; }

// MAXN = 256 for block clustering (up to 256 block types survive), 128 for context maps (at most 64 clusters
// survive): a quarter of the pair matrix (32 KiB instead of 130 KiB of shared memory: two problems per SM
// instead of one) and a quarter of the all-pairs evaluations.
template <int MAXN>
__global__ void __launch_bounds__(512) k_cluster_final(const ClusterProb* __restrict__ probs) {
  constexpr uint32_t PAIRS = MAXN * (MAXN - 1) / 2;
  extern __shared__ __align__(16) uint8_t csm[];
  float* s_d = reinterpret_cast<float*>(csm);                       // triangular pair matrix
  uint32_t* s_idx = reinterpret_cast<uint32_t*>(s_d + PAIRS);
  float* s_cost = reinterpret_cast<float*>(s_idx + MAXN);
  uint32_t* s_tot = reinterpret_cast<uint32_t*>(s_cost + MAXN);
  uint32_t* s_size = s_tot + MAXN;
  uint8_t* s_live = reinterpret_cast<uint8_t*>(s_size + MAXN);
  __shared__ float s_rv[16]; __shared__ int s_ri[16]; __shared__ int s_best; __shared__ float s_combo;
  const ClusterProb p = probs[blockIdx.x];
  const uint32_t n = p.state[0];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (n == 0 || n > (uint32_t)MAXN) { if (tid == 0) p.state[2] = 0; return; }
  for (uint32_t k = tid; k < n; k += 512) {
    const uint32_t ci = p.list[k];
    s_idx[k] = ci; s_cost[k] = p.cost[ci]; s_tot[k] = p.total[ci]; s_size[k] = p.csize[ci]; s_live[k] = 1;
  }
  __syncthreads();
  auto eval_pair = [&](uint32_t i, uint32_t j, float* combo_out) {
    float combo;
    if (s_tot[i] == 0) combo = s_cost[j];
    else if (s_tot[j] == 0) combo = s_cost[i];
    else combo = hist_cost(p.work + (size_t)s_idx[i] * p.stride, p.work + (size_t)s_idx[j] * p.stride, (int)p.alpha, nullptr);
    if (lane == 0) s_d[tri(i, j)] = 0.5f * cluster_cost_diff((float)s_size[i], (float)s_size[j]) - s_cost[i] - s_cost[j] + combo;
    if (combo_out) *combo_out = combo;
  };
  if (n > 1) {
    for (uint32_t pr = w; pr < n * n; pr += 16) {
      const uint32_t i = pr / n, j = pr % n;
      if (i < j) eval_pair(i, j, nullptr);
    }
  }
  __syncthreads();
  uint32_t alive = n;
  while (alive > 1) {
    float bv = 3.0e38f; int bi = -1;
    for (uint32_t pr = tid; pr < n * n; pr += 512) {
      const uint32_t i = pr / n, j = pr % n;
      if (i < j && s_live[i] && s_live[j]) { const float d = s_d[tri(i, j)]; if (d < bv) { bv = d; bi = (int)pr; } }
    }
    for (int o = 16; o; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o); const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi >= 0 && (bi < 0 || ov < bv || (ov == bv && oi < bi))) { bv = ov; bi = oi; }
    }
    if (lane == 0) { s_rv[w] = bv; s_ri[w] = bi; }
    __syncthreads();
    if (tid == 0) {
      float v = s_rv[0]; int ix = s_ri[0];
      for (int k = 1; k < 16; k++) if (s_ri[k] >= 0 && (ix < 0 || s_rv[k] < v || (s_rv[k] == v && s_ri[k] < ix))) { v = s_rv[k]; ix = s_ri[k]; }
      if (ix >= 0 && !(v < 0.f) && alive <= p.max_clusters) ix = -1;
      s_best = ix;
    }
    __syncthreads();
    const int best = s_best;
    if (best < 0) break;
    const uint32_t i = (uint32_t)best / n, j = (uint32_t)best % n;
    // the combined cost is recomputed (the matrix only keeps the difference)
    if (w == 0) { float combo; eval_pair(i, j, &combo); if (lane == 0) s_combo = combo; }
    __syncthreads();
    {
      uint32_t* a = p.work + (size_t)s_idx[i] * p.stride; const uint32_t* b = p.work + (size_t)s_idx[j] * p.stride;
      for (uint32_t k = tid; k < p.alpha; k += 512) a[k] = __ldcg(a + k) + __ldcg(b + k);
    }
    __syncthreads();
    if (tid == 0) { s_cost[i] = s_combo; s_tot[i] += s_tot[j]; s_size[i] += s_size[j]; s_live[j] = 0; p.alias[s_idx[j]] = s_idx[i]; }
    __threadfence_block();
    __syncthreads();
    alive--;
    for (uint32_t k = w; k < n; k += 16) {
      if (k == i || !s_live[k]) continue;
      if (k < i) eval_pair(k, i, nullptr); else eval_pair(i, k, nullptr);
    }
    __syncthreads();
  }
  // compact the live clusters (order of slots): list2[c] = cluster index, state[2] = count; the remap and
  // the reindexing run as their own kernels (k_cluster_remap: parallel over inputs, k_cluster_reindex)
  if (tid == 0) {
    uint32_t c = 0;
    for (uint32_t k = 0; k < n; k++) if (s_live[k]) { p.total[s_idx[k]] = s_tot[k]; p.list2[c++] = s_idx[k]; }
    p.state[2] = c;
  }
}

// remap (cluster.ts:244-275).  grid (ceil(max inputs / 64), problems), 8 warps, 8 inputs per warp.
// Distance of input h to cluster c:
//   sum over the support of h of [f(o + x) - f(o)] (+3 bits for a symbol new to the cluster)
//   + g(Tc + Th) - g(Tc),   f(p) = -p log2 p, g(T) = T log2 T.
// The input row stays in registers (<= 6 x 16 bytes per lane); every cluster row is fetched with all
// its 16-byte pieces in flight at once.
// When the final clusters' rows fit the dynamic shared memory (<= 64 clusters of <= 256 bins: every literal
// context map) the CTA stages them once and its 64 inputs are measured against shared memory.
template <int NV>
__global__ void __launch_bounds__(256) k_cluster_remap(const ClusterProb* __restrict__ probs, uint32_t smem_bytes) {
  extern __shared__ __align__(16) uint32_t s_cl[];   // [ncl][alpha] when staged
  const ClusterProb p = probs[blockIdx.y];
  if ((p.alpha > 256) != (NV > 2)) return;
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (blockIdx.x * 64 >= p.m) return;
  const uint32_t ncl = p.state[2];
  const int nvec = (int)p.alpha >> 2;
  const bool staged = (size_t)ncl * p.alpha * 4 <= smem_bytes;
  if (staged) {
    for (uint32_t c = w; c < ncl; c += 8) {
      const uint4* src = reinterpret_cast<const uint4*>(p.work + (size_t)p.list2[c] * p.stride);
      uint4* dst = reinterpret_cast<uint4*>(s_cl + (size_t)c * p.alpha);
      for (int k = (int)lane; k < nvec; k += 32) dst[k] = __ldcg(src + k);
    }
    __syncthreads();
  }
  const uint32_t i0 = blockIdx.x * 64 + w * 8;
  if (i0 >= p.m) return;
  for (uint32_t i = i0; i < min(p.m, i0 + 8); i++) {
    const uint4* h4 = reinterpret_cast<const uint4*>(p.in + (size_t)i * p.stride);
    uint4 hv[NV]; uint32_t th = 0;
#pragma unroll
    for (int t = 0; t < NV; t++) {
      const int k = t * 32 + (int)lane;
      hv[t] = k < nvec ? h4[k] : make_uint4(0, 0, 0, 0);
      th += hv[t].x + hv[t].y + hv[t].z + hv[t].w;
    }
    for (int o = 16; o; o >>= 1) th += __shfl_xor_sync(0xffffffffu, th, o);
    uint32_t bestc = 0xFFFFFFFFu; float bestd = 3.0e38f;
    if (th != 0) {
      // the input's current cluster through the alias chain: preferred on ties
      uint32_t cur = i;
      while (true) { const uint32_t a = p.alias[cur]; if (a == cur) break; cur = a; }
      for (uint32_t c = 0; c < ncl; c++) {
        const uint32_t ci = p.list2[c];
        const uint4* o4 = staged ? reinterpret_cast<const uint4*>(s_cl + (size_t)c * p.alpha) : reinterpret_cast<const uint4*>(p.work + (size_t)ci * p.stride);
        uint4 ov[NV];
#pragma unroll
        for (int t = 0; t < NV; t++) {
          const int k = t * 32 + (int)lane;
          ov[t] = (k < nvec && (hv[t].x | hv[t].y | hv[t].z | hv[t].w)) ? (staged ? o4[k] : __ldcg(o4 + k)) : make_uint4(0, 0, 0, 0);
        }
        float d = 0.f;
#pragma unroll
        for (int t = 0; t < NV; t++) {
          const uint32_t xs[4] = {hv[t].x, hv[t].y, hv[t].z, hv[t].w}, os[4] = {ov[t].x, ov[t].y, ov[t].z, ov[t].w};
#pragma unroll
          for (int q = 0; q < 4; q++) {
            if (xs[q]) {
              const float sum = (float)(os[q] + xs[q]);
              d -= sum * flog2(sum);
              if (os[q]) d += (float)os[q] * flog2((float)os[q]); else d += 3.f;
            }
          }
        }
        for (int oo = 16; oo; oo >>= 1) d += __shfl_xor_sync(0xffffffffu, d, oo);
        const float tc = (float)p.total[ci], ts = tc + (float)th;
        d += ts * flog2(ts) - (tc > 0.f ? tc * flog2(tc) : 0.f);
        const bool own = ci == cur;
        if (d < bestd || (d == bestd && own)) { bestd = d; bestc = c; }
      }
    }
    if (lane == 0) p.sym[i] = bestc;   // 0xFFFFFFFF: empty input, takes its predecessor's cluster (k_cluster_reindex)
  }
}

// Empty inputs follow the previous non-empty input (cluster.ts:252: bestOut starts at symbols[i-1]; an
// all-empty prefix takes the first non-empty input's cluster, or cluster 0), then the clusters are
// renumbered by first appearance (cluster.ts:278-314).  One CTA per problem.
__global__ void __launch_bounds__(256) k_cluster_reindex(const ClusterProb* __restrict__ probs) {
  __shared__ uint32_t s_first[CL_FINAL_MAX], s_new[CL_FINAL_MAX];
  __shared__ int s_wlast[8]; __shared__ int s_carry; __shared__ uint32_t s_firstsym, s_cnt;
  const ClusterProb p = probs[blockIdx.x];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t ncl = p.state[2];
  if (p.m == 0 || ncl == 0 || ncl > CL_FINAL_MAX) return;
  for (uint32_t k = tid; k < CL_FINAL_MAX; k += 256) s_first[k] = 0xFFFFFFFFu;
  if (tid == 0) { s_carry = -1; s_firstsym = 0xFFFFFFFFu; }
  __syncthreads();
  // first non-empty input
  {
    uint32_t f = 0xFFFFFFFFu;
    for (uint32_t i = tid; i < p.m; i += 256) if (p.sym[i] != 0xFFFFFFFFu) { f = i; break; }
    if (f != 0xFFFFFFFFu) atomicMin(&s_firstsym, f);
  }
  __syncthreads();
  const uint32_t first_ne = s_firstsym;
  const uint32_t lead = first_ne != 0xFFFFFFFFu ? p.sym[first_ne] : 0;
  __syncthreads();
  // fill pass: chunks of 256 inputs, prefix-max of "index of the last non-empty input"
  for (uint32_t base = 0; base < p.m; base += 256) {
    const uint32_t i = base + tid;
    const uint32_t sy = i < p.m ? p.sym[i] : 0xFFFFFFFFu;
    int v = sy != 0xFFFFFFFFu ? (int)i : -1;
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, v, o); if (lane >= (uint32_t)o) v = max(v, y); }
    if (lane == 31) s_wlast[w] = v;
    __syncthreads();
    int pre = s_carry;
    for (uint32_t k = 0; k < w; k++) pre = max(pre, s_wlast[k]);
    v = max(v, pre);
    __syncthreads();
    if (tid == 255) s_carry = v;
    if (i < p.m) {
      const uint32_t c = v >= 0 ? (v == (int)i ? sy : p.sym[v]) : lead;
      p.alias[i] = c;                      // alias is free now: effective (filled) cluster of input i
      atomicMin(&s_first[c], i);
    }
    __syncthreads();
  }
  // new id = rank of the cluster's first appearance
  for (uint32_t c = tid; c < ncl; c += 256) {
    const uint32_t f = s_first[c];
    uint32_t r = 0;
    if (f != 0xFFFFFFFFu) for (uint32_t k = 0; k < ncl; k++) r += s_first[k] < f;
    s_new[c] = r;
  }
  if (tid == 0) s_cnt = 0;
  __syncthreads();
  {
    uint32_t used = 0;
    for (uint32_t c = tid; c < ncl; c += 256) used += s_first[c] != 0xFFFFFFFFu;
    if (used) atomicAdd(&s_cnt, used);
  }
  for (uint32_t i = tid; i < p.m; i += 256) p.sym[i] = s_new[p.alias[i]];
  __syncthreads();
  if (tid == 0) p.state[2] = s_cnt;
}

// ---------------------------------------------------------------------------------------
// Block splitting.  One split problem = one symbol stream (literals, command codes or distance
// codes of one metablock).
struct SplitProb {
  const void* data;        // u8 (literals) or u16 symbols
  uint32_t* hist;          // [SPLIT_MAXK][alpha]  entropy-code histograms
  uint8_t* gid;            // [n_gran]  entropy code of each granule
  __half* table;           // [alpha][64]  insert-cost table of the current refinement iteration (k_split_table)
  __half* acc;             // [n_gran][64]  per-granule insert costs (k_split_acc), shared by the 3 categories of a metablock
  uint32_t* state;         // [4]  0: live entropy codes K, 1: number of blocks
  uint32_t* blk_start;     // [MAX_BLOCKS + 1]  first symbol of each block (prefix), blk_start[nb] = n
  uint8_t* blk_type;       // [MAX_BLOCKS]
  const uint32_t* n_ptr;   // number of symbols (device side)
  uint32_t alpha, gran, sym_per_hist, wide;   // wide: 1 = u16 symbols
  uint32_t max_blocks;
  float switch_cost;
};

__device__ __forceinline__ uint32_t split_sym(const SplitProb& p, uint32_t i) {
  return p.wide ? reinterpret_cast<const uint16_t*>(p.data)[i] : reinterpret_cast<const uint8_t*>(p.data)[i];
}
__device__ __forceinline__ uint32_t split_k0(uint32_t n, uint32_t sym_per_hist) { return min((uint32_t)SPLIT_MAXK, n / sym_per_hist + 1); }

// (Re)builds the entropy-code histograms from the granule assignment.  init = 1: granule j
// belongs to code j * K / G (contiguous slices).  grid (chunks of 64 granules, problems).
__global__ void __launch_bounds__(256) k_split_hist(const SplitProb* __restrict__ probs, int init, uint32_t gpc) {
  extern __shared__ uint32_t s_hh[];   // alpha counters
  const SplitProb p = probs[blockIdx.y];
  const uint32_t n = *p.n_ptr;
  if (n < 128) return;
  const uint32_t G = (n + p.gran - 1) / p.gran;
  const uint32_t K = init ? split_k0(n, p.sym_per_hist) : p.state[0];
  const uint32_t g0 = blockIdx.x * gpc;   // gpc granules per CTA (about 8192 symbols)
  if (g0 >= G) return;
  if (K <= 1 && !init) return;
  const uint32_t g1 = min(G, g0 + gpc);
  for (uint32_t i = threadIdx.x; i < p.alpha; i += 256) s_hh[i] = 0;
  __syncthreads();
  // runs of granules with the same entropy code are counted with all 256 threads at once (a granule is
  // only 16 symbols in the command / distance streams), then flushed to the code's global histogram
  __shared__ uint32_t s_run_end, s_run_id;
  uint32_t g = g0;
  while (g < g1) {
    if (threadIdx.x == 0) {
      const uint32_t id = init ? (uint32_t)(((uint64_t)g * K) / G) : p.gid[g];
      uint32_t e = g + 1;
      while (e < g1 && (init ? (uint32_t)(((uint64_t)e * K) / G) : (uint32_t)p.gid[e]) == id) e++;
      s_run_end = e; s_run_id = id;
    }
    __syncthreads();
    const uint32_t ge = s_run_end, id = s_run_id;
    const uint32_t b = g * p.gran, e = min(n, ge * p.gran);
    for (uint32_t i = b + threadIdx.x; i < e; i += 256) atomicAdd(&s_hh[split_sym(p, i)], 1u);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < p.alpha; i += 256) { const uint32_t v = s_hh[i]; if (v) { atomicAdd(&p.hist[id * p.alpha + i], v); s_hh[i] = 0; } }
    __syncthreads();
    g = ge;
  }
}

// findBlocks over granules (block-splitter.ts:117-211), in two kernels:
//   k_split_acc  every granule in parallel (one warp each): A[g][k] = sum over the granule's symbols of
//                insertCost[sym][k], stored relative to the granule's minimum and clamped at 64 bits
//                (the DP is invariant under a per-granule constant, and anything 64 above the minimum is
//                clamped to the block-switch cost right after it is added) as 64 halves = 128 B per granule;
//   k_split_dp   the dependent chain: one warp per chunk of SPLIT_CHUNK granules, 2 codes per lane,
//                switch words and ids of the chunk kept in shared memory for the trace-back.
// The insert-cost table of the problem (alpha x 64 halves) sits in dynamic shared memory of k_split_acc.
constexpr uint32_t SPLIT_ACC_GPC_LIT = 256, SPLIT_ACC_GPC_WIDE = 1024;   // granules per CTA (128 / 16 symbols each)
constexpr uint32_t SPLIT_HIST_GPC_LIT = 64, SPLIT_HIST_GPC_WIDE = 512;
constexpr float SPLIT_ACC_CLAMP = 64.f;

// insertCost[sym][k] = log2(total_k) - log2(count_k[sym])   (:136-143; empty bin: -(log2(1) + 2) sic),
// once per problem and refinement iteration, as alpha x 64 halves in global memory.
__global__ void __launch_bounds__(256) k_split_table(const SplitProb* __restrict__ probs) {
  __shared__ float s_lt[SPLIT_MAXK];
  const SplitProb p = probs[blockIdx.x];
  const uint32_t n = *p.n_ptr;
  const uint32_t K = p.state[0];
  if (n < 128 || K <= 1) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  for (uint32_t k = w; k < SPLIT_MAXK; k += 8) {
    uint32_t sum = 0;
    if (k < K) for (uint32_t i = lane; i < p.alpha; i += 32) sum += p.hist[k * p.alpha + i];
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) s_lt[k] = sum ? flog2((float)sum) : 0.f;
  }
  __syncthreads();
  for (uint32_t e = tid; e < p.alpha * SPLIT_MAXK; e += 256) {
    const uint32_t sym = e / SPLIT_MAXK, k = e % SPLIT_MAXK;
    float c = 60000.f;   // codes >= K never win
    if (k < K) { const uint32_t cnt = p.hist[k * p.alpha + sym]; c = fmaxf(0.f, s_lt[k] - (cnt ? flog2((float)cnt) : -2.f)); }
    p.table[e] = __float2half(fminf(c, 60000.f));
  }
}

__global__ void __launch_bounds__(256) k_split_acc(const SplitProb* __restrict__ probs, uint32_t gpc) {
  extern __shared__ __align__(16) __half s_ic[];   // [alpha][64]
  const SplitProb p = probs[blockIdx.y];
  const uint32_t n = *p.n_ptr;
  if (n < 128) return;
  const uint32_t G = (n + p.gran - 1) / p.gran;
  const uint32_t K = p.state[0];
  const uint32_t g0 = blockIdx.x * gpc;   // granules per CTA: amortises the table load (up to 90 KB)
  if (g0 >= G || K <= 1) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  {
    const uint4* src = reinterpret_cast<const uint4*>(p.table);
    uint4* dst = reinterpret_cast<uint4*>(s_ic);
    for (uint32_t e = tid; e < p.alpha * SPLIT_MAXK / 8; e += 256) dst[e] = src[e];
  }
  __syncthreads();
  const __half2* ic2 = reinterpret_cast<const __half2*>(s_ic);
  __half2* acc2 = reinterpret_cast<__half2*>(p.acc);
  const uint32_t gend = min(G, g0 + gpc);
  for (uint32_t g = g0 + w; g < gend; g += 8) {
    const uint32_t sb = g * p.gran, cnt = min(n, sb + p.gran) - sb;
    float a0 = 0.f, a1 = 0.f;
    if (!p.wide) {
      // 128 one-byte symbols: one u32 (4 symbols) per lane, broadcast by shuffle
      const uint8_t* d = reinterpret_cast<const uint8_t*>(p.data) + sb;
      uint32_t wv = 0;
      if (lane * 4 < cnt) {
        if (lane * 4 + 4 <= cnt && ((reinterpret_cast<uintptr_t>(d) & 3) == 0)) wv = *reinterpret_cast<const uint32_t*>(d + lane * 4);
        else for (uint32_t b = 0; b < 4 && lane * 4 + b < cnt; b++) wv |= (uint32_t)d[lane * 4 + b] << (8 * b);
      }
      // 4 symbols (one shuffled word) per step: their table entries are summed as half2 (each <= 60000 by
      // construction; four real costs stay far below the half range and lose < 1/16 bit) and widened once
      const uint32_t cnt4 = cnt & ~3u;
#pragma unroll 4
      for (uint32_t i = 0; i < cnt4; i += 4) {
        const uint32_t w4 = __shfl_sync(0xffffffffu, wv, i >> 2);
        const __half2 h = __hadd2(__hadd2(ic2[(w4 & 0xFF) * (SPLIT_MAXK / 2) + lane], ic2[((w4 >> 8) & 0xFF) * (SPLIT_MAXK / 2) + lane]),
                                  __hadd2(ic2[((w4 >> 16) & 0xFF) * (SPLIT_MAXK / 2) + lane], ic2[(w4 >> 24) * (SPLIT_MAXK / 2) + lane]));
        const float2 f = __half22float2(h);
        a0 += f.x; a1 += f.y;
      }
      for (uint32_t i = cnt4; i < cnt; i++) {
        const uint32_t sym = (__shfl_sync(0xffffffffu, wv, i >> 2) >> ((i & 3) * 8)) & 0xFF;
        const float2 f = __half22float2(ic2[sym * (SPLIT_MAXK / 2) + lane]);
        a0 += f.x; a1 += f.y;
      }
    } else {
      const uint16_t* d = reinterpret_cast<const uint16_t*>(p.data) + sb;
      const uint32_t mine = lane < cnt ? d[lane] : 0;
      if (cnt == 16) {
        // full granule: all 16 table rows in flight at once; groups of 4 summed as half2, widened once each
#pragma unroll
        for (uint32_t i = 0; i < 16; i += 4) {
          const uint32_t s0 = __shfl_sync(0xffffffffu, mine, i), s1 = __shfl_sync(0xffffffffu, mine, i + 1);
          const uint32_t s2 = __shfl_sync(0xffffffffu, mine, i + 2), s3 = __shfl_sync(0xffffffffu, mine, i + 3);
          const __half2 h = __hadd2(__hadd2(ic2[s0 * (SPLIT_MAXK / 2) + lane], ic2[s1 * (SPLIT_MAXK / 2) + lane]),
                                    __hadd2(ic2[s2 * (SPLIT_MAXK / 2) + lane], ic2[s3 * (SPLIT_MAXK / 2) + lane]));
          const float2 f = __half22float2(h);
          a0 += f.x; a1 += f.y;
        }
      } else {
        for (uint32_t i = 0; i < cnt; i++) {
          const uint32_t sym = __shfl_sync(0xffffffffu, mine, i);
          const float2 f = __half22float2(ic2[sym * (SPLIT_MAXK / 2) + lane]);
          a0 += f.x; a1 += f.y;
        }
      }
    }
    const float m = fminf(a0, a1);
    const float mn = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(m)));   // costs >= 0: bit order == value order
    acc2[(size_t)g * (SPLIT_MAXK / 2) + lane] = __floats2half2_rn(fminf(a0 - mn, SPLIT_ACC_CLAMP), fminf(a1 - mn, SPLIT_ACC_CLAMP));
  }
}

__global__ void __launch_bounds__(128) k_split_dp(const SplitProb* __restrict__ probs) {
  extern __shared__ __align__(16) uint8_t s_dp[];   // per warp: uint2 sw[SPLIT_CHUNK], u8 gid[SPLIT_CHUNK]
  const SplitProb p = probs[blockIdx.y];
  const uint32_t n = *p.n_ptr;
  if (n < 128) return;
  const uint32_t G = (n + p.gran - 1) / p.gran;
  const uint32_t K = p.state[0];
  const uint32_t nchunk = (G + SPLIT_CHUNK - 1) / SPLIT_CHUNK;
  if (blockIdx.x * 4 >= nchunk) return;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (K <= 1) {
    for (uint32_t g = blockIdx.x * 4 * SPLIT_CHUNK + tid; g < min(G, (blockIdx.x * 4 + 4) * SPLIT_CHUNK); g += 128) p.gid[g] = 0;
    return;
  }
  const uint32_t chunk = blockIdx.x * 4 + w;
  if (chunk >= nchunk) return;
  uint2* s_sw = reinterpret_cast<uint2*>(s_dp + (size_t)w * SPLIT_CHUNK * 9);
  uint8_t* s_gid = reinterpret_cast<uint8_t*>(s_sw + SPLIT_CHUNK);
  const uint32_t gbeg = chunk * SPLIT_CHUNK, gend = min(G, gbeg + SPLIT_CHUNK);
  float c0 = 0.f, c1 = 0.f;   // codes 2*lane, 2*lane+1
  const __half2* acc2 = reinterpret_cast<const __half2*>(p.acc);
  // The chain of a chunk is one dependent step per granule; its per-granule costs are prefetched SPLIT_PF
  // granules ahead into a register ring, so no step waits for memory (one step ahead left an L2 round trip
  // of ~700 cycles on every step of a ~150-cycle ALU chain: 1.3 ms per launch for the 16-symbol granules).
  constexpr int SPLIT_PF = 8;
  __half2 ring[SPLIT_PF];
#pragma unroll
  for (int k = 0; k < SPLIT_PF; k++) ring[k] = gbeg + k < gend ? acc2[(size_t)(gbeg + k) * (SPLIT_MAXK / 2) + lane] : __half2();
  for (uint32_t g0 = gbeg; g0 < gend; g0 += SPLIT_PF) {
#pragma unroll
    for (int k = 0; k < SPLIT_PF; k++) {
    const uint32_t g = g0 + k;
    if (g >= gend) break;
    const float2 f = __half22float2(ring[k]);
    if (g + SPLIT_PF < gend) ring[k] = acc2[(size_t)(g + SPLIT_PF) * (SPLIT_MAXK / 2) + lane];
    c0 += f.x; c1 += f.y;
    // warp arg-min (costs are >= 0, so the float order is the order of the bit patterns)
    const float m = fminf(c0, c1);
    const uint32_t mb = __reduce_min_sync(0xffffffffu, __float_as_uint(m));
    const float minc = __uint_as_float(mb);
    const uint32_t who = __ballot_sync(0xffffffffu, __float_as_uint(m) == mb);
    const uint32_t wl = __ffs((int)who) - 1;
    const uint32_t first = __shfl_sync(0xffffffffu, (c0 <= c1) ? 2 * lane : 2 * lane + 1, wl);
    const uint32_t sb = g * p.gran;
    float bsc = p.switch_cost;
    if (sb < 2000) bsc *= 0.77f + (0.07f / 2000.f) * (float)sb;   // :165-170
    c0 -= minc; c1 -= minc;
    const bool s0 = c0 >= bsc, s1 = c1 >= bsc;
    if (s0) c0 = bsc;
    if (s1) c1 = bsc;
    const uint32_t e0 = __ballot_sync(0xffffffffu, s0), e1 = __ballot_sync(0xffffffffu, s1);
    // switch bit of code 2l is bit l of e0, of code 2l + 1 bit l of e1
    if (lane == 0) { s_sw[g - gbeg] = make_uint2(e0, e1); s_gid[g - gbeg] = (uint8_t)first; }
  }
  }
  __syncwarp();
  // trace back (:194-209) in shared memory, then one coalesced store of the chunk's ids
  if (lane == 0) {
    uint32_t g = gend - 1 - gbeg;
    uint32_t cur = s_gid[g];
    while (g > 0) {
      g--;
      const uint2 sw = s_sw[g];
      if ((((cur & 1) ? sw.y : sw.x) >> (cur >> 1)) & 1) cur = s_gid[g];
      s_gid[g] = (uint8_t)cur;
    }
  }
  __syncwarp();
  for (uint32_t g = gbeg + lane; g < gend; g += 32) p.gid[g] = s_gid[g - gbeg];
}

// remapBlockIds (:213-234) + histogram reset for the next refinement.  One CTA per problem.
__global__ void __launch_bounds__(256) k_split_remap(const SplitProb* __restrict__ probs) {
  __shared__ uint32_t s_new[SPLIT_MAXK]; __shared__ uint32_t s_first[SPLIT_MAXK]; __shared__ uint32_t s_k;
  const SplitProb p = probs[blockIdx.x];
  const uint32_t n = *p.n_ptr;
  if (n < 128) return;
  const uint32_t G = (n + p.gran - 1) / p.gran;
  const uint32_t K = p.state[0];
  if (threadIdx.x < SPLIT_MAXK) { s_new[threadIdx.x] = 0xFFFFFFFFu; s_first[threadIdx.x] = 0xFFFFFFFFu; }
  __syncthreads();
  if (K > 1) {
    // first granule of every code (parallel), then new ids in order of first appearance
    for (uint32_t g = threadIdx.x; g < G; g += 256) {
      const uint32_t id = p.gid[g];
      if (g == 0 || p.gid[g - 1] != id) atomicMin(&s_first[id], g);
    }
    __syncthreads();
    if (threadIdx.x < SPLIT_MAXK) {
      const uint32_t f = s_first[threadIdx.x];
      if (f != 0xFFFFFFFFu) {
        uint32_t rank = 0;
        for (uint32_t k = 0; k < SPLIT_MAXK; k++) rank += s_first[k] < f;
        s_new[threadIdx.x] = rank;
      }
    }
    if (threadIdx.x == 0) { uint32_t c = 0; for (uint32_t k = 0; k < SPLIT_MAXK; k++) c += s_first[k] != 0xFFFFFFFFu; s_k = c; }
  } else if (threadIdx.x == 0) s_k = 1;
  __syncthreads();
  if (K > 1) for (uint32_t g = threadIdx.x; g < G; g += 256) p.gid[g] = (uint8_t)s_new[p.gid[g]];
  for (uint32_t i = threadIdx.x; i < SPLIT_MAXK * p.alpha; i += 256) p.hist[i] = 0;
  if (threadIdx.x == 0) p.state[0] = s_k;
}

// Initial state of a problem: K, zeroed histograms.  One CTA per problem.
__global__ void __launch_bounds__(256) k_split_begin(const SplitProb* __restrict__ probs) {
  const SplitProb p = probs[blockIdx.x];
  const uint32_t n = *p.n_ptr;
  for (uint32_t i = threadIdx.x; i < SPLIT_MAXK * p.alpha; i += 256) p.hist[i] = 0;
  if (threadIdx.x == 0) { p.state[0] = n < 128 ? 1 : split_k0(n, p.sym_per_hist); p.state[1] = 0; }
}

// Runs of equal granule ids -> blocks (at most p.max_blocks; later runs are folded into the last
// block).  One warp per problem.
__global__ void __launch_bounds__(32) k_split_blocks(const SplitProb* __restrict__ probs) {
  const SplitProb p = probs[blockIdx.x];
  const uint32_t n = *p.n_ptr;
  const uint32_t lane = threadIdx.x;
  if (n == 0) { if (lane == 0) { p.state[1] = 0; p.blk_start[0] = 0; } return; }
  if (n < 128 || p.state[0] <= 1) {
    if (lane == 0) { p.state[1] = 1; p.blk_start[0] = 0; p.blk_start[1] = n; p.blk_type[0] = 0; }
    return;
  }
  const uint32_t G = (n + p.gran - 1) / p.gran;
  uint32_t nb = 0;
  for (uint32_t g0 = 0; g0 < G; g0 += 32) {
    const uint32_t g = g0 + lane;
    const bool start = g < G && (g == 0 || p.gid[g] != p.gid[g - 1]);
    const uint32_t m = __ballot_sync(0xffffffffu, start);
    const uint32_t idx = nb + __popc(m & ((1u << lane) - 1));
    if (start && idx < p.max_blocks) { p.blk_start[idx] = g * p.gran; p.blk_type[idx] = p.gid[g]; }
    nb += __popc(m);
  }
  nb = min(nb, p.max_blocks);
  if (lane == 0) { p.state[1] = nb; p.blk_start[nb] = n; }
}

// Per-block histograms for clusterBlocks (:326-340).  grid (64, problems); CTA x takes blocks x, x + 64, ...
__global__ void __launch_bounds__(256) k_block_hist(const SplitProb* __restrict__ probs, uint32_t* const* __restrict__ bh_ptrs,
                                                    const uint32_t* __restrict__ strides) {
  extern __shared__ uint32_t s_hh[];
  const SplitProb p = probs[blockIdx.y];
  const uint32_t nb = p.state[1];
  if (nb <= 1) return;
  for (uint32_t blk = blockIdx.x; blk < nb; blk += gridDim.x) {
    for (uint32_t i = threadIdx.x; i < p.alpha; i += 256) s_hh[i] = 0;
    __syncthreads();
    const uint32_t b = p.blk_start[blk], e = p.blk_start[blk + 1];
    for (uint32_t i = b + threadIdx.x; i < e; i += 256) atomicAdd(&s_hh[split_sym(p, i)], 1u);
    __syncthreads();
    uint32_t* out = bh_ptrs[blockIdx.y] + (size_t)blk * strides[blockIdx.y];
    for (uint32_t i = threadIdx.x; i < p.alpha; i += 256) out[i] = s_hh[i];
    __syncthreads();
  }
}

}  // namespace bgpu
