// k_zop_dp: the shortest-path pass itself (see zopfli.cuh for the formulation).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef ZOP_DP_BLOCKS
#define ZOP_DP_BLOCKS 8
#endif
namespace bgpu {

constexpr uint32_t ZOP_W = 512;   // node costs of [i, i + ZOP_W) live in shared memory (power of two)

struct ZopQueue {           // StartPosQueue of one warp (backward-references-hq.ts:100-149)
  uint32_t pos[8]; float costdiff[8]; float cost[8]; float lc[8]; int ring[8][4];
};

// number of +1.0 steps computeMinimumCopyLength has applied when it tests length L
__device__ __forceinline__ float zop_min_adds(uint32_t L) { return L < 10 ? 0.f : (float)(log2_floor(L - 6) - 1); }

// Relaxes node `at` = i + len.  The cost of a node less than ZOP_W ahead is compared in the warp's
// shared-memory window (the global record is write-only on this path); farther nodes go to global
// memory and raise far_hi so that the window picks them up when it slides over them.
__device__ __forceinline__ void zop_relax(uint4* __restrict__ nodes, float* __restrict__ cw, uint32_t at, float cost, uint32_t len,
                                          uint32_t dist, uint32_t dcode, uint32_t& result, uint32_t& far_hi) {
  if (len < ZOP_W) {
    float* c = cw + (at & (ZOP_W - 1));
    if (cost < *c) { *c = cost; nodes[at] = make_uint4(__float_as_uint(cost), len, dist, dcode); result = max(result, len); }
  } else if (cost < __uint_as_float(nodes[at].x)) {
    nodes[at] = make_uint4(__float_as_uint(cost), len, dist, dcode);
    result = max(result, len); far_hi = max(far_hi, at);
  }
}

// One warp per parse segment; 4 warps per CTA.  nodes / ringarr are indexed by
// in_off + rel_pos + segment index (+ position inside the segment): n + 1 nodes per segment.
//
// Latency plan (a position is a chain of dependent steps, so every global round trip that can be
// taken off the chain is): node costs of the next ZOP_W positions sit in shared memory (minimum-
// length scan and every relaxation compare there); the node record, the literal-cost prefix and
// the H10 list header of a position come from registers loaded 32 positions at a time; the match
// entries of position i + 1 are fetched while position i is processed.  What stays on the chain is
// one L2 round trip for the distance-cache probes.
__global__ void __launch_bounds__(128, ZOP_DP_BLOCKS) k_zop_dp(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                const SegD* __restrict__ segs, uint32_t n_seg, H10Out O,
                                                const float* __restrict__ model, const float* __restrict__ lc,
                                                uint4* __restrict__ nodes, int4* __restrict__ ringarr, uint8_t* __restrict__ role,
                                                uint32_t* __restrict__ ins_at, Cmd* __restrict__ cmds, uint32_t* __restrict__ seg_ncmd,
                                                uint32_t* __restrict__ seg_trail, Params P, ZopParams Z) {
  __shared__ ZopQueue s_q[4];
  __shared__ float s_cw[4][ZOP_W];
  const uint32_t FULL = 0xffffffffu;
  const uint32_t seg = (blockIdx.x * 128 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (seg >= n_seg) return;
  ZopQueue& Q = s_q[threadIdx.x >> 5];
  float* cw = s_cw[threadIdx.x >> 5];
  const SegD sg = segs[seg];
  const ShardD sh = shards[sg.shard];
  const uint8_t* sin = in + sh.in_off;
  const float* M = model + (size_t)sg.mb * ZOP_MODEL_STRIDE;
  const float* lcs = lc + sh.in_off + sg.rel_pos;
  const uint64_t nb = sh.in_off + sg.rel_pos + seg;
  uint4* nd = nodes + nb;
  int4* rg = ringarr + nb;
  const uint32_t n = sg.len;
  const float minCostCmd = M[960 + ZOP_DIST_COSTS];
  const uint32_t* mreg_base = O.m + sh.in_off * O.cap_per_pos;
  const uint32_t* poff = O.off + sh.in_off + sg.rel_pos;
  const uint8_t* pcnt = O.cnt + sh.in_off + sg.rel_pos;

  for (uint32_t t = lane; t <= n; t += 32) nd[t] = make_uint4(t == 0 ? 0u : __float_as_uint(ZOP_INF), t == 0 ? 0u : 1u, 0u, 0u);
  for (uint32_t t = lane; t < ZOP_W; t += 32) cw[t] = t == 0 ? 0.f : ZOP_INF;
  __syncwarp();

  uint32_t qidx = 0;
  uint32_t far_max = 0;            // highest node index ever relaxed through global memory (warp uniform)
  uint32_t chunk_base = 0xFFFFFFFFu;
  float lc32 = 0.f; uint32_t off32 = 0, cnt32 = 0;   // lane j: position chunk_base + j
  uint32_t pre_pos = 0xFFFFFFFFu, pre_e = 0;         // lane t: entry t of position pre_pos (first 32 entries)
  uint32_t i = 0;
  while (i + 3 < n) {
    if ((i & ~31u) != chunk_base) {
      chunk_base = i & ~31u;
      const uint32_t q = chunk_base + lane;
      lc32 = q < n ? lcs[q] : 0.f;
      off32 = q < n ? poff[q] : 0; cnt32 = q < n ? pcnt[q] : 0;
    }
    const float lci = __shfl_sync(FULL, lc32, i & 31);
    const uint32_t cnt = __shfl_sync(FULL, cnt32, i & 31);
    const uint32_t moff = __shfl_sync(FULL, off32, i & 31);
    const uint32_t* mreg = mreg_base + moff;
    uint32_t my_e = 0;
    const bool have_pre = pre_pos == i;
    if (have_pre) my_e = pre_e; else if (lane < cnt) my_e = mreg[lane];
    // the match entries of the next position (same chunk only) travel while this one is processed
    const bool will_pre = ((i + 1) & 31) != 0 && i + 4 < n;
    if (will_pre) {
      const uint32_t c1 = __shfl_sync(FULL, cnt32, (i + 1) & 31), o1 = __shfl_sync(FULL, off32, (i + 1) & 31);
      pre_e = lane < c1 ? mreg_base[o1 + lane] : 0;
    }
    const float mycost = cw[i & (ZOP_W - 1)];
    // ---- evaluateNode (:239-264)
    if (mycost <= lci) {
      // ring of the node: the ring stored for its start node, with the node's distance pushed unless the
      // command used distance code 0 (computeDistanceCache :214-234 walks shortcut links to the same values)
      int4 ring;
      if (i == 0) ring = sg.ring_known ? make_int4(4, 11, 15, 16) : make_int4(0, 0, 0, 0);
      else {
        const uint4 me = nd[i];
        const uint32_t start = i - (me.y & 0x1FFFFFFu) - (me.w & 0x7FFFFFFu);
        const int4 rs = rg[start];
        ring = (me.w >> 27) == 1 ? rs : make_int4((int)me.z, rs.x, rs.y, rs.z);
      }
      if (lane == 0) rg[i] = ring;
      // StartPosQueue.push (:117-141): the queue is sorted by costdiff, the new entry replaces the worst
      // one and bubbles in front of every entry that is not smaller.  Entry k lives at Q.*[k]; lanes 0..7
      // move the entries behind the insertion point down by one slot, in parallel.
      {
        const float ncd = mycost - lci;
        const uint32_t keep = min(qidx, 7u);
        const bool mine = lane < keep;
        float e_cd = 0.f, e_cost = 0.f, e_lc = 0.f; uint32_t e_pos = 0; int e0 = 0, e1 = 0, e2 = 0, e3 = 0;
        if (mine) {
          e_pos = Q.pos[lane]; e_cd = Q.costdiff[lane]; e_cost = Q.cost[lane]; e_lc = Q.lc[lane];
          e0 = Q.ring[lane][0]; e1 = Q.ring[lane][1]; e2 = Q.ring[lane][2]; e3 = Q.ring[lane][3];
        }
        const uint32_t p = __popc(__ballot_sync(FULL, mine && e_cd < ncd));
        __syncwarp();
        if (mine && lane >= p) {
          Q.pos[lane + 1] = e_pos; Q.costdiff[lane + 1] = e_cd; Q.cost[lane + 1] = e_cost; Q.lc[lane + 1] = e_lc;
          Q.ring[lane + 1][0] = e0; Q.ring[lane + 1][1] = e1; Q.ring[lane + 1][2] = e2; Q.ring[lane + 1][3] = e3;
        }
        if (lane == 31) {
          Q.pos[p] = i; Q.costdiff[p] = ncd; Q.cost[p] = mycost; Q.lc[p] = lci;
          Q.ring[p][0] = ring.x; Q.ring[p][1] = ring.y; Q.ring[p][2] = ring.z; Q.ring[p][3] = ring.w;
        }
      }
      qidx++;
      __syncwarp();
    }
    const uint32_t K = min(Z.max_iters, min(qidx, 8u));
    const uint32_t maxLen = n - i;
    const uint32_t cur = sg.rel_pos + i;
    const uint32_t maxDistance = min(cur, MAX_BACKWARD);
    // ---- computeMinimumCopyLength (:168-190)
    uint32_t minLen;
    {
      const float minCost = Q.cost[0] + minCostCmd + (lci - Q.lc[0]);
      for (uint32_t base = 2;; base += 32) {
        const uint32_t L = base + lane;
        bool ok = i + L <= n;
        if (ok) {
          const float c = L < ZOP_W ? cw[(i + L) & (ZOP_W - 1)] : (i + L <= far_max ? __uint_as_float(nd[i + L].x) : ZOP_INF);
          ok = c <= minCost + zop_min_adds(L);
        }
        const uint32_t fail = __ballot_sync(FULL, !ok);
        if (fail) { minLen = base + __ffs((int)fail) - 1; break; }
      }
    }
    uint32_t result = 0, far_hi = 0;
    // ---- distance-cache candidates (:306-342): lane = (start k, code j)
    const uint32_t cur4 = ld32u(sin, cur);
    for (uint32_t r = 0; r * 2 < K; r++) {
      const uint32_t k = 2 * r + (lane >> 4), j = lane & 15;
      const uint32_t slot = k & 7;
      uint32_t len = 0; int backward = 0;
      if (k < K) {
        const uint32_t t6 = j >= 4 ? (j - 4) % 6 : 0;
        const int base = Q.ring[slot][j < 4 ? j : (j < 10 ? 0 : 1)];
        const int off = j < 4 ? 0 : ((t6 & 1) ? 1 : -1) * (int)(t6 / 2 + 1);
        backward = base + off;
        if (base > 0 && backward > 0 && (uint32_t)backward <= maxDistance) {
          if (ld32u(sin, cur - (uint32_t)backward) == cur4)
            len = min(4 + match_len16(sin, cur + 4, cur - (uint32_t)backward + 4, maxLen - 4), maxLen);
        }
      }
      // exclusive prefix max inside the 16-lane group = the reference's bestLen before candidate j
      uint32_t pm = len;
#pragma unroll
      for (int o = 1; o < 16; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, pm, o, 16); if (j >= (uint32_t)o) pm = max(pm, y); }
      pm = __shfl_up_sync(FULL, pm, 1, 16);
      if (j == 0) pm = 0;
      const uint32_t lo = max(pm, minLen - 1);
      uint32_t mask = __ballot_sync(FULL, len > lo);
      while (mask) {
        const int src = __ffs((int)mask) - 1; mask &= mask - 1;
        const uint32_t sk = 2 * r + ((uint32_t)src >> 4), sj = (uint32_t)src & 15;
        const uint32_t slo = __shfl_sync(FULL, lo, src), shi = __shfl_sync(FULL, len, src);
        const uint32_t sback = (uint32_t)__shfl_sync(FULL, backward, src);
        const uint32_t ss = sk;
        const uint32_t insLen = i - Q.pos[ss];
        const uint32_t ic = ins_code(insLen);
        const float baseCost = Q.costdiff[ss] + (float)ins_extra(ic) + lci;
        const float distCost = baseCost + M[960 + sj];
        for (uint32_t l = slo + 1 + lane; l <= shi; l += 32) {
          const uint32_t cc = copy_code(l), cmd = combine_codes(ic, cc, sj == 0);
          const float cost = (cmd < 128 ? baseCost : distCost) + (float)copy_extra(cc) + M[256 + cmd];
          zop_relax(nd, cw, i + l, cost, l, sback, ((sj + 1) << 27) | insLen, result, far_hi);
        }
        __syncwarp();
      }
    }
    // ---- H10 matches (:347-378): starts 0 and 1 in the two half warps, one lane per copy length
    uint32_t skipLen = 0;
    if (cnt) {
      const uint32_t shardMax = sh.len - cur;
      const uint32_t half = lane >> 4, hl = lane & 15;
      const bool kvalid = half < K;
      const uint32_t ss = half;
      const uint32_t insLen = i - Q.pos[ss];
      const uint32_t ic = ins_code(insLen);
      const float baseCost = Q.costdiff[ss] + (float)ins_extra(ic) + lci;
      uint32_t carry = 0;   // longest length among the entries already consumed
      // the longest match decides between "all matches" and "only the longest, then skip" (:528-535, :571-573)
      uint32_t lastLen;
      {
        const uint32_t e = cnt <= 32 ? __shfl_sync(FULL, my_e, (cnt - 1) & 31) : mreg[cnt - 1];
        lastLen = e >> 24;
        if (lastLen >= H10_MAX_TREE_COMP && shardMax > H10_MAX_TREE_COMP) {
          lastLen = H10_MAX_TREE_COMP + match_len16(sin, cur + H10_MAX_TREE_COMP, cur - (e & 0xFFFFFFu) + H10_MAX_TREE_COMP,
                                                    shardMax - H10_MAX_TREE_COMP);
          // a capped short-scan entry may be the longer of the two capped entries
          if (cnt >= 2) {
            const uint32_t e2 = mreg[cnt - 2];
            if ((e2 >> 24) >= H10_MAX_TREE_COMP)
              lastLen = max(lastLen, H10_MAX_TREE_COMP + match_len16(sin, cur + H10_MAX_TREE_COMP, cur - (e2 & 0xFFFFFFu) + H10_MAX_TREE_COMP,
                                                                     shardMax - H10_MAX_TREE_COMP));
          }
        }
        lastLen = min(lastLen, maxLen);
      }
      const bool onlyLast = lastLen > Z.max_zopfli_len;
      if (onlyLast) skipLen = lastLen;
      for (uint32_t cb = 0; cb < cnt; cb += 32) {
        const uint32_t t = cb + lane;
        uint32_t dist = 0, len = 0;
        if (t < cnt) {
          const uint32_t e = cb == 0 ? my_e : mreg[t];
          dist = e & 0xFFFFFFu; len = e >> 24;
          if (len >= H10_MAX_TREE_COMP && shardMax > H10_MAX_TREE_COMP)
            len = H10_MAX_TREE_COMP + match_len16(sin, cur + H10_MAX_TREE_COMP, cur - dist + H10_MAX_TREE_COMP, shardMax - H10_MAX_TREE_COMP);
          len = min(len, maxLen);
        }
        // running maximum of the lengths before this entry
        uint32_t pm = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(FULL, pm, o); if (lane >= (uint32_t)o) pm = max(pm, y); }
        const uint32_t incl_last = __shfl_sync(FULL, pm, 31);
        pm = __shfl_up_sync(FULL, pm, 1);
        if (lane == 0) pm = 0;
        pm = max(pm, carry);
        carry = max(carry, incl_last);
        uint32_t lo = max(minLen, pm + 1);
        if (onlyLast) lo = (len == lastLen && pm < len) ? max(minLen, len) : len + 1;
        uint32_t mask = __ballot_sync(FULL, t < cnt && len >= lo && len >= 2);
        while (mask) {
          const int src = __ffs((int)mask) - 1; mask &= mask - 1;
          const uint32_t sdist = __shfl_sync(FULL, dist, src), slo = __shfl_sync(FULL, lo, src), shi = __shfl_sync(FULL, len, src);
          const uint32_t didx = (sdist + 15) & 0x3FF;
          if (didx >= ZOP_DIST_COSTS) continue;           // costDist[idx] is undefined -> NaN cost, never taken (:355-357)
          const float distCost = baseCost + (float)(log2_floor(sdist) - 1) + M[960 + didx];
          for (uint32_t l0 = slo; l0 <= shi; l0 += 16) {
            const uint32_t l = l0 + hl;
            float cost = ZOP_INF;
            if (kvalid && l <= shi) {
              const uint32_t cc = copy_code(l), cmd = combine_codes(ic, cc, false);
              cost = distCost + (float)copy_extra(cc) + M[256 + cmd];
            }
            // both starts target the same node: the cheaper one (start 0 on ties) writes
            const float other = __shfl_xor_sync(FULL, cost, 16);
            const bool mine = half == 0 ? cost <= other : cost < other;
            if (mine && cost < ZOP_INF) zop_relax(nd, cw, i + l, cost, l, sdist, insLen, result, far_hi);
          }
          __syncwarp();
        }
      }
    }
    for (int o = 16; o; o >>= 1) { result = max(result, __shfl_xor_sync(FULL, result, o)); far_hi = max(far_hi, __shfl_xor_sync(FULL, far_hi, o)); }
    far_max = max(far_max, far_hi);
    uint32_t step = 1;
    if (result >= 16384) step = result;                 // :540-541
    else if (skipLen) step = skipLen;
    // slide the cost window: the slots of nodes [i, i + step) now stand for nodes [i + ZOP_W, i + step + ZOP_W)
    if (step == 1) {
      if (lane == 0) cw[i & (ZOP_W - 1)] = (i + ZOP_W <= far_max && i + ZOP_W <= n) ? __uint_as_float(nd[i + ZOP_W].x) : ZOP_INF;
    } else {
      const uint32_t lo_node = step >= ZOP_W ? i + step : i + ZOP_W, hi_node = i + step + ZOP_W;   // [lo_node, hi_node)
      for (uint32_t q = lo_node + lane; q < hi_node; q += 32)
        cw[q & (ZOP_W - 1)] = (q <= far_max && q <= n) ? __uint_as_float(nd[q].x) : ZOP_INF;
    }
    pre_pos = (will_pre && step == 1) ? i + 1 : 0xFFFFFFFFu;
    i += step;
    __syncwarp();
  }

  // ---- computeShortestPathFromNodes + createCommandsFromPath (:384-406, :610-673)
  uint32_t index = n;
  for (;;) {
    // 32 nodes per step, scanning back for the last node reached by a copy
    const int64_t q = (int64_t)index - lane;
    const bool set = q > 0 && nd[q].y != 1u;
    const uint32_t m = __ballot_sync(0xffffffffu, set);
    if (m) { index -= __ffs((int)m) - 1; break; }
    if (index <= 32) { index = 0; break; }
    index -= 32;
  }
  uint8_t* srole = role + sh.in_off + sg.rel_pos;
  const uint64_t abs0 = sh.in_off + sg.rel_pos;
  const uint32_t trail = index;
  for (uint32_t t = trail + lane; t < n; t += 32) srole[t] = 1 | (t == trail ? 4 : 0);
  uint32_t ncmd = 0, e = index;
  while (e > 0) {
    const uint4 x = nd[e];
    const uint32_t copy = x.y & 0x1FFFFFFu, ins = x.w & 0x7FFFFFFu, sc = x.w >> 27;
    const uint32_t cstart = e - copy, istart = cstart - ins;
    if (lane == 0) {
      const Cmd c = make_cmd(ins, copy, sc == 0 ? x.z + 15 : sc - 1, P.ndirect, P.npostfix);
      *reinterpret_cast<uint4*>(&cmds[(abs0 + cstart) >> P.sh]) = *reinterpret_cast<const uint4*>(&c);
      srole[cstart] = 2 | (ins == 0 ? 4 : 0);
      ins_at[(abs0 + istart) >> P.sh] = ins;
    }
    for (uint32_t t = lane; t < ins; t += 32) srole[istart + t] = 1 | (t == 0 ? 4 : 0);
    e = istart; ncmd++;
  }
  if (lane == 0) { seg_ncmd[seg] = ncmd; seg_trail[seg] = sg.rel_pos + trail; }
}

}  // namespace bgpu
