// k_zop_dp: the shortest-path pass itself (see zopfli.cuh for the formulation).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bgpu {

struct ZopQueue {           // StartPosQueue of one warp (backward-references-hq.ts:100-149)
  uint32_t pos[8]; float costdiff[8]; float cost[8]; int ring[8][4];
};

// number of +1.0 steps computeMinimumCopyLength has applied when it tests length L
__device__ __forceinline__ float zop_min_adds(uint32_t L) { return L < 10 ? 0.f : (float)(log2_floor(L - 6) - 1); }

__device__ __forceinline__ void zop_relax(uint4* __restrict__ nodes, uint32_t at, float cost, uint32_t len, uint32_t dist, uint32_t dcode,
                                          uint32_t& result) {
  if (cost < __uint_as_float(nodes[at].x)) {
    nodes[at] = make_uint4(__float_as_uint(cost), len, dist, dcode);
    result = max(result, len);
  }
}

// One warp per parse segment; 4 warps per CTA.  nodes / ringarr are indexed by
// in_off + rel_pos + segment index (+ position inside the segment): n + 1 nodes per segment.
__global__ void __launch_bounds__(128) k_zop_dp(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                const SegD* __restrict__ segs, uint32_t n_seg, H10Out O,
                                                const float* __restrict__ model, const float* __restrict__ lc,
                                                uint4* __restrict__ nodes, int4* __restrict__ ringarr, uint8_t* __restrict__ role,
                                                uint32_t* __restrict__ ins_at, Cmd* __restrict__ cmds, uint32_t* __restrict__ seg_ncmd,
                                                uint32_t* __restrict__ seg_trail, Params P, ZopParams Z) {
  __shared__ ZopQueue s_q[4];
  const uint32_t seg = (blockIdx.x * 128 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (seg >= n_seg) return;
  ZopQueue& Q = s_q[threadIdx.x >> 5];
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

  for (uint32_t t = lane; t <= n; t += 32) nd[t] = make_uint4(t == 0 ? 0u : __float_as_uint(ZOP_INF), t == 0 ? 0u : 1u, 0u, 0u);
  __syncwarp();

  uint32_t qidx = 0;
  for (uint32_t i = 0; i + 3 < n; i++) {
    // ---- evaluateNode (:239-264)
    const uint4 me = nd[i];
    const float lci = lcs[i];
    if (__uint_as_float(me.x) <= lci) {
      if (lane == 0) {
        int4 ring;
        if (i == 0) ring = sg.ring_known ? make_int4(4, 11, 15, 16) : make_int4(0, 0, 0, 0);
        else {
          const uint32_t start = i - (me.y & 0x1FFFFFFu) - (me.w & 0x7FFFFFFu);
          const int4 rs = rg[start];
          ring = (me.w >> 27) == 1 ? rs : make_int4((int)me.z, rs.x, rs.y, rs.z);
        }
        rg[i] = ring;
        // StartPosQueue.push (:117-141)
        const uint32_t offset = (~qidx) & 7;
        Q.pos[offset] = i; Q.costdiff[offset] = __uint_as_float(me.x) - lci; Q.cost[offset] = __uint_as_float(me.x);
        Q.ring[offset][0] = ring.x; Q.ring[offset][1] = ring.y; Q.ring[offset][2] = ring.z; Q.ring[offset][3] = ring.w;
        const uint32_t qlen = min(qidx + 1, 8u);
        for (uint32_t k = 1; k < qlen; k++) {
          const uint32_t a = (offset + k - 1) & 7, b = (offset + k) & 7;
          if (Q.costdiff[a] > Q.costdiff[b]) {
            uint32_t tp = Q.pos[a]; Q.pos[a] = Q.pos[b]; Q.pos[b] = tp;
            float tf = Q.costdiff[a]; Q.costdiff[a] = Q.costdiff[b]; Q.costdiff[b] = tf;
            tf = Q.cost[a]; Q.cost[a] = Q.cost[b]; Q.cost[b] = tf;
#pragma unroll
            for (int r = 0; r < 4; r++) { int ti = Q.ring[a][r]; Q.ring[a][r] = Q.ring[b][r]; Q.ring[b][r] = ti; }
          }
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
      const uint32_t s0 = (0u - qidx) & 7;
      const float minCost = Q.cost[s0] + minCostCmd + (lci - lcs[Q.pos[s0]]);
      for (uint32_t base = 2;; base += 32) {
        const uint32_t L = base + lane;
        const bool ok = i + L <= n && __uint_as_float(nd[i + L].x) <= minCost + zop_min_adds(L);
        const uint32_t fail = __ballot_sync(0xffffffffu, !ok);
        if (fail) { minLen = base + __ffs((int)fail) - 1; break; }
      }
    }
    uint32_t result = 0;
    // ---- distance-cache candidates (:306-342): lane = (start k, code j)
    for (uint32_t r = 0; r * 2 < K; r++) {
      const uint32_t k = 2 * r + (lane >> 4), j = lane & 15;
      const uint32_t slot = (k - qidx) & 7;
      uint32_t len = 0; int backward = 0;
      if (k < K) {
        const uint32_t t6 = j >= 4 ? (j - 4) % 6 : 0;
        const int base = Q.ring[slot][j < 4 ? j : (j < 10 ? 0 : 1)];
        const int off = j < 4 ? 0 : ((t6 & 1) ? 1 : -1) * (int)(t6 / 2 + 1);
        backward = base + off;
        if (base > 0 && backward > 0 && (uint32_t)backward <= maxDistance) {
          if (ld32u(sin, cur) == ld32u(sin, cur - backward) || maxLen < 4) {
            len = match_len16(sin, cur, cur - (uint32_t)backward, maxLen);
            if (len < 4) len = 0;
          }
        }
      }
      // exclusive prefix max inside the 16-lane group = the reference's bestLen before candidate j
      uint32_t pm = len;
#pragma unroll
      for (int o = 1; o < 16; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, pm, o, 16); if (j >= (uint32_t)o) pm = max(pm, y); }
      pm = __shfl_up_sync(0xffffffffu, pm, 1, 16);
      if (j == 0) pm = 0;
      const uint32_t lo = max(pm, minLen - 1);
      uint32_t mask = __ballot_sync(0xffffffffu, len > lo);
      while (mask) {
        const int src = __ffs((int)mask) - 1; mask &= mask - 1;
        const uint32_t sk = 2 * r + ((uint32_t)src >> 4), sj = (uint32_t)src & 15;
        const uint32_t slo = __shfl_sync(0xffffffffu, lo, src), shi = __shfl_sync(0xffffffffu, len, src);
        const uint32_t sback = (uint32_t)__shfl_sync(0xffffffffu, backward, src);
        const uint32_t ss = (sk - qidx) & 7;
        const uint32_t insLen = i - Q.pos[ss];
        const uint32_t ic = ins_code(insLen);
        const float baseCost = Q.costdiff[ss] + (float)ins_extra(ic) + lci;
        const float distCost = baseCost + M[960 + sj];
        for (uint32_t l = slo + 1 + lane; l <= shi; l += 32) {
          const uint32_t cc = copy_code(l), cmd = combine_codes(ic, cc, sj == 0);
          const float cost = (cmd < 128 ? baseCost : distCost) + (float)copy_extra(cc) + M[256 + cmd];
          zop_relax(nd, i + l, cost, l, sback, ((sj + 1) << 27) | insLen, result);
        }
        __syncwarp();
      }
    }
    // ---- H10 matches (:347-378): starts 0 and 1 in the two half warps, one lane per copy length
    const uint32_t cnt = O.cnt[sh.in_off + cur];
    uint32_t skipLen = 0;
    if (cnt) {
      const uint32_t* mreg = mreg_base + O.off[sh.in_off + cur];
      const uint32_t shardMax = sh.len - cur;
      const uint32_t half = lane >> 4, hl = lane & 15;
      const bool kvalid = half < K;
      const uint32_t ss = (half - qidx) & 7;
      const uint32_t insLen = i - Q.pos[ss];
      const uint32_t ic = ins_code(insLen);
      const float baseCost = Q.costdiff[ss] + (float)ins_extra(ic) + lci;
      uint32_t carry = 0;   // longest length among the entries already consumed
      // the longest match decides between "all matches" and "only the longest, then skip" (:528-535, :571-573)
      uint32_t lastLen;
      {
        const uint32_t e = mreg[cnt - 1];
        lastLen = e >> 24;
        if (lastLen >= H10_MAX_TREE_COMP && shardMax > H10_MAX_TREE_COMP)
          lastLen = H10_MAX_TREE_COMP + match_len16(sin, cur + H10_MAX_TREE_COMP, cur - (e & 0xFFFFFFu) + H10_MAX_TREE_COMP,
                                                    shardMax - H10_MAX_TREE_COMP);
        // a capped short-scan entry may be the longer of the two capped entries
        if (cnt >= 2 && (mreg[cnt - 2] >> 24) >= H10_MAX_TREE_COMP && shardMax > H10_MAX_TREE_COMP) {
          const uint32_t e2 = mreg[cnt - 2];
          const uint32_t l2 = H10_MAX_TREE_COMP + match_len16(sin, cur + H10_MAX_TREE_COMP, cur - (e2 & 0xFFFFFFu) + H10_MAX_TREE_COMP,
                                                              shardMax - H10_MAX_TREE_COMP);
          lastLen = max(lastLen, l2);
        }
        lastLen = min(lastLen, maxLen);
      }
      const bool onlyLast = lastLen > Z.max_zopfli_len;
      if (onlyLast) skipLen = lastLen;
      for (uint32_t cb = 0; cb < cnt; cb += 32) {
        const uint32_t t = cb + lane;
        uint32_t dist = 0, len = 0;
        if (t < cnt) {
          const uint32_t e = mreg[t];
          dist = e & 0xFFFFFFu; len = e >> 24;
          if (len >= H10_MAX_TREE_COMP && shardMax > H10_MAX_TREE_COMP)
            len = H10_MAX_TREE_COMP + match_len16(sin, cur + H10_MAX_TREE_COMP, cur - dist + H10_MAX_TREE_COMP, shardMax - H10_MAX_TREE_COMP);
          len = min(len, maxLen);
        }
        // running maximum of the lengths before this entry
        uint32_t pm = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, pm, o); if (lane >= (uint32_t)o) pm = max(pm, y); }
        const uint32_t incl_last = __shfl_sync(0xffffffffu, pm, 31);
        pm = __shfl_up_sync(0xffffffffu, pm, 1);
        if (lane == 0) pm = 0;
        pm = max(pm, carry);
        carry = max(carry, incl_last);
        uint32_t lo = max(minLen, pm + 1);
        if (onlyLast) lo = (len == lastLen && pm < len) ? max(minLen, len) : len + 1;
        uint32_t mask = __ballot_sync(0xffffffffu, t < cnt && len >= lo && len >= 2);
        while (mask) {
          const int src = __ffs((int)mask) - 1; mask &= mask - 1;
          const uint32_t sdist = __shfl_sync(0xffffffffu, dist, src), slo = __shfl_sync(0xffffffffu, lo, src),
                         shi = __shfl_sync(0xffffffffu, len, src);
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
            const float other = __shfl_xor_sync(0xffffffffu, cost, 16);
            const bool mine = half == 0 ? cost <= other : cost < other;
            if (mine && cost < ZOP_INF) zop_relax(nd, i + l, cost, l, sdist, insLen, result);
          }
          __syncwarp();
        }
      }
    }
    for (int o = 16; o; o >>= 1) result = max(result, __shfl_xor_sync(0xffffffffu, result, o));
    if (result >= 16384) i += result - 1;                 // :540-541
    else if (skipLen) i += skipLen - 1;
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
