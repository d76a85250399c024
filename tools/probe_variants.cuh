// Experimental variants of k_probe_diag for tools/probe_bench.cu (development only).
#pragma once
namespace bgpu {

template <int GATE>
__device__ __forceinline__ void pv_step(const uint4& x, uint32_t p, uint32_t mlen, uint32_t cp, const uint4& cx, uint32_t jj,
                                        uint32_t& jlim, uint32_t& lm, uint32_t& bLen, uint32_t& bDist, int& bScore) {
  const uint32_t backward = p - cp;
  if (backward > MAX_BACKWARD) { jlim = 0; return; }
  uint32_t len = 4 + prefix16(x, cx);
  if (len == 20 && mlen > 20) { lm |= 1u << jj; return; }
  len = min(len, mlen);
  if (GATE && len <= bLen) return;   // later candidates are farther: only a longer one can score higher
  const int score = 135 * (int)len + 960 - 30 * log2_floor(backward);
  if (score > bScore) { bScore = score; bLen = len; bDist = backward; }
}

__device__ __forceinline__ void pv_deferred(const uint8_t* sin, const uint32_t* s_p, uint32_t i0, uint32_t jc, uint32_t p, uint32_t mlen, uint32_t lm,
                                            uint32_t& bLen, uint32_t& bDist, int& bScore) {
  while (lm) {
    const uint32_t jj = __ffs((int)lm) - 1; lm &= lm - 1;
    const uint32_t c = s_p[i0 - jc - jj - 1];
    const uint32_t backward = p - c;
    const int dcost = 960 - 30 * log2_floor(backward);
    const int ub = 135 * (int)mlen + dcost;
    if (ub < bScore || (ub == bScore && bDist <= backward)) continue;
    const uint32_t len = 20 + match_len16(sin, p + 20, c + 20, mlen - 20);
    const int score = 135 * (int)len + dcost;
    if (score > bScore || (score == bScore && backward < bDist)) { bScore = score; bLen = len; bDist = backward; }
  }
}

template <int BLOCKS, int THREADS, int ROUNDS, int ILP, int GATE>
__global__ void __launch_bounds__(THREADS, BLOCKS) k_pv(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                        const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ srt_v,
                                                        uint32_t* __restrict__ best, int depth, int bucket_bits) {
  constexpr uint32_t TILE_N = THREADS * ROUNDS;
  extern __shared__ __align__(16) uint8_t smem[];
  const uint32_t halo = (uint32_t)depth;
  const uint32_t n_ent = TILE_N + halo;
  uint4* s_x = reinterpret_cast<uint4*>(smem);
  uint32_t* s_p = reinterpret_cast<uint32_t*>(s_x + n_ent);
  uint32_t* s_v = s_p + n_ent;
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh);
  const uint32_t g0 = blockIdx.x * TILE_N;
  const uint32_t tid = threadIdx.x;
  uint32_t* sbest = best + sh.in_off;
  if (blockIdx.x == 0 && tid < sh.len - npos) sbest[npos + tid] = 0;
  if (g0 >= npos) return;
  const uint32_t* sp = srt_pos + sh.in_off; const uint32_t* sv = srt_v + sh.in_off;
  const uint8_t* sin = in + sh.in_off;
  for (uint32_t i = tid; i < n_ent; i += THREADS * 4) {
    uint32_t pp[4], vv[4]; uint4 xx[4]; bool ok[4];
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const uint32_t ii = i + THREADS * t;
      const int64_t g = (int64_t)g0 - halo + ii;
      ok[t] = ii < n_ent && g >= 0 && g < (int64_t)npos;
      pp[t] = ok[t] ? __ldg(sp + g) : 0; vv[t] = ok[t] ? __ldg(sv + g) : 0;
    }
#pragma unroll
    for (int t = 0; t < 4; t++) xx[t] = ok[t] ? ld128v(sin + pp[t] + 4) : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int t = 0; t < 4; t++) { const uint32_t ii = i + THREADS * t; if (ii < n_ent) { s_p[ii] = pp[t]; s_v[ii] = vv[t]; s_x[ii] = xx[t]; } }
  }
  __syncthreads();
  for (uint32_t r = 0; r < ROUNDS; r += ILP) {
    uint32_t i0[ILP], v[ILP], p[ILP], mlen[ILP], bLen[ILP], bDist[ILP], jlim[ILP]; int bScore[ILP]; uint4 x[ILP]; bool live[ILP];
    uint32_t jmax = 0;
#pragma unroll
    for (int q = 0; q < ILP; q++) {
      const uint32_t e = (r + q) * THREADS + tid, g = g0 + e;
      live[q] = g < npos; i0[q] = halo + e;
      v[q] = s_v[i0[q]]; p[q] = s_p[i0[q]]; x[q] = s_x[i0[q]];
      mlen[q] = min(sh.len - p[q], 128u); bLen[q] = 0; bDist[q] = 0; bScore[q] = 0;
      jlim[q] = live[q] ? min((uint32_t)depth, g) : 0; jmax = max(jmax, jlim[q]);
    }
    for (uint32_t jc = 0; jc < jmax; jc += 16) {
      uint32_t lm[ILP];
#pragma unroll
      for (int q = 0; q < ILP; q++) lm[q] = 0;
#pragma unroll
      for (uint32_t jj = 0; jj < 16; jj++) {
        const uint32_t j = jc + jj + 1;
        bool c[ILP];
#pragma unroll
        for (int q = 0; q < ILP; q++) c[q] = j <= jlim[q] && s_v[i0[q] - j] == v[q];
#pragma unroll
        for (int q = 0; q < ILP; q++)
          if (c[q]) { const uint32_t ci = i0[q] - j; pv_step<GATE>(x[q], p[q], mlen[q], s_p[ci], s_x[ci], jj, jlim[q], lm[q], bLen[q], bDist[q], bScore[q]); }
      }
#pragma unroll
      for (int q = 0; q < ILP; q++) pv_deferred(sin, s_p, i0[q], jc, p[q], mlen[q], lm[q], bLen[q], bDist[q], bScore[q]);
    }
#pragma unroll
    for (int q = 0; q < ILP; q++) if (live[q]) sbest[p[q]] = (bLen[q] << 24) | bDist[q];
  }
}


// v2: no per-step chain-limit state.  The walk's only early end (backward > MAX_BACKWARD, hash-chains.ts:124) is the
// same as skipping every such candidate, because candidates come in order of growing distance; steps that reach
// before the shard's first entry exist only in the first tile (EDGE = true).
template <bool EDGE, int GATE>
__device__ __forceinline__ void pw_entry(const uint8_t* sin, const uint4* s_x, const uint32_t* s_p, const uint32_t* s_v, uint32_t i0, uint32_t g,
                                         uint32_t shard_len, int depth, uint32_t* sbest) {
  const uint32_t v = s_v[i0], p = s_p[i0];
  const uint4 x = s_x[i0];
  const uint32_t mlen = min(shard_len - p, 128u);
  uint32_t bLen = 0, bDist = 0; int bScore = 0;
  const uint32_t jlim = EDGE ? min((uint32_t)depth, g) : (uint32_t)depth;
  for (uint32_t jc = 0; jc < jlim; jc += 16) {
    uint32_t lm = 0;
#pragma unroll
    for (uint32_t jj = 0; jj < 16; jj++) {
      const uint32_t ci = i0 - jc - jj - 1;
      if ((!EDGE || jc + jj < jlim) && s_v[ci] == v) {
        const uint32_t cp = s_p[ci];
        const uint4 cx = s_x[ci];
        const uint32_t backward = p - cp;
        const bool near = backward <= MAX_BACKWARD;
        uint32_t len = 4 + prefix16(x, cx);
        if (len == 20 && mlen > 20) { if (near) lm |= 1u << jj; }
        else {
          len = min(len, mlen);
          if (!GATE || len > bLen) {
            const int score = 135 * (int)len + 960 - 30 * log2_floor(backward);
            if (score > bScore && near) { bScore = score; bLen = len; bDist = backward; }
          }
        }
      }
    }
    pv_deferred(sin, s_p, i0, jc, p, mlen, lm, bLen, bDist, bScore);
  }
  sbest[p] = (bLen << 24) | bDist;
}

template <int BLOCKS, int GATE>
__global__ void __launch_bounds__(256, BLOCKS) k_pw(const uint8_t* __restrict__ in, const ShardD* __restrict__ shards,
                                                    const uint32_t* __restrict__ srt_pos, const uint32_t* __restrict__ srt_v,
                                                    uint32_t* __restrict__ best, int depth, int bucket_bits) {
  extern __shared__ __align__(16) uint8_t smem[];
  const uint32_t halo = (uint32_t)depth;
  const uint32_t n_ent = PROBE_TILE + halo;
  uint4* s_x = reinterpret_cast<uint4*>(smem);
  uint32_t* s_p = reinterpret_cast<uint32_t*>(s_x + n_ent);
  uint32_t* s_v = s_p + n_ent;
  const ShardD sh = shards[blockIdx.y];
  const uint32_t npos = shard_npos(sh);
  const uint32_t g0 = blockIdx.x * PROBE_TILE;
  const uint32_t tid = threadIdx.x;
  uint32_t* sbest = best + sh.in_off;
  if (blockIdx.x == 0 && tid < sh.len - npos) sbest[npos + tid] = 0;
  if (g0 >= npos) return;
  const uint32_t* sp = srt_pos + sh.in_off; const uint32_t* sv = srt_v + sh.in_off;
  const uint8_t* sin = in + sh.in_off;
  {
    constexpr int NST = (PROBE_TILE + PROBE_HALO + 255) / 256;
    uint32_t pp[NST], vv[NST]; uint4 xx[NST]; bool ok[NST];
#pragma unroll
    for (int t = 0; t < NST; t++) {
      const uint32_t i = tid + 256 * t;
      const int64_t g = (int64_t)g0 - halo + i;
      ok[t] = i < n_ent && g >= 0 && g < (int64_t)npos;
      pp[t] = ok[t] ? __ldg(sp + g) : 0; vv[t] = ok[t] ? __ldg(sv + g) : 0;
    }
#pragma unroll
    for (int t = 0; t < NST; t++) xx[t] = ok[t] ? ld128v(sin + pp[t] + 4) : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int t = 0; t < NST; t++) { const uint32_t i = tid + 256 * t; if (i < n_ent) { s_p[i] = pp[t]; s_v[i] = vv[t]; s_x[i] = xx[t]; } }
  }
  __syncthreads();
  if (g0 < halo) {
    for (uint32_t r = 0; r < PROBE_TILE / 256; r++) { const uint32_t e = r * 256 + tid; if (g0 + e < npos) pw_entry<true, GATE>(sin, s_x, s_p, s_v, halo + e, g0 + e, sh.len, depth, sbest); }
  } else {
    for (uint32_t r = 0; r < PROBE_TILE / 256; r++) { const uint32_t e = r * 256 + tid; if (g0 + e < npos) pw_entry<false, GATE>(sin, s_x, s_p, s_v, halo + e, g0 + e, sh.len, depth, sbest); }
  }
}

template <int BLOCKS, int GATE, class T, class C>
void run_pw(const char* name, const uint8_t* d_in, const ShardD* d_sh, const uint32_t* sp, const uint32_t* sv, uint32_t* best, int depth, int bb,
            uint32_t max_npos, uint32_t n_shard, T& timed, C& check) {
  const size_t sm = (size_t)(PROBE_TILE + depth) * 24;
  auto kern = k_pw<BLOCKS, GATE>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess) { printf("%s: smem %zu refused\n", name, sm); cudaGetLastError(); return; }
  const dim3 gp((max_npos + PROBE_TILE - 1) / PROBE_TILE, n_shard);
  cudaMemset(best, 0xFF, 16);
  timed(name, [&] { kern<<<gp, 256, sm>>>(d_in, d_sh, sp, sv, best, depth, bb); });
  check(name, best);
}

template <int BLOCKS, int THREADS, int ROUNDS, int ILP, int GATE, class T, class C>
void run_pv(const char* name, const uint8_t* d_in, const ShardD* d_sh, const uint32_t* sp, const uint32_t* sv, uint32_t* best, int depth, int bb,
            uint32_t max_npos, uint32_t n_shard, T& timed, C& check) {
  constexpr uint32_t TILE_N = THREADS * ROUNDS;
  const size_t sm = (size_t)(TILE_N + depth) * 24;
  auto kern = k_pv<BLOCKS, THREADS, ROUNDS, ILP, GATE>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess) { printf("%s: smem %zu refused\n", name, sm); cudaGetLastError(); return; }
  const dim3 gp((max_npos + TILE_N - 1) / TILE_N, n_shard);
  cudaMemset(best, 0xFF, 16);
  timed(name, [&] { kern<<<gp, THREADS, sm>>>(d_in, d_sh, sp, sv, best, depth, bb); });
  check(name, best);
}

template <class T, class C>
void run_variants(const uint8_t* d_in, const ShardD* d_sh, const uint32_t* sp, const uint32_t* sv, uint32_t* best, int depth, int bb,
                  uint32_t max_npos, uint32_t n_shard, T& timed, C& check) {
#define PV(B, TH, R, I, G) run_pv<B, TH, R, I, G>("pv<" #B "," #TH "," #R "," #I "," #G ">", d_in, d_sh, sp, sv, best, depth, bb, max_npos, n_shard, timed, check)
  PV(6, 256, 4, 1, 0);
#define PW(B, G) run_pw<B, G>("pw<" #B "," #G ">", d_in, d_sh, sp, sv, best, depth, bb, max_npos, n_shard, timed, check)
  PW(6, 0); PW(6, 1); PW(5, 0); PW(5, 1); PW(4, 1); PW(7, 1);
#undef PW
#undef PV
}

}  // namespace bgpu
