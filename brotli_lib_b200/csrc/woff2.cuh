// WOFF2 glyf/loca transform front end: the step directly upstream of BASELINE config 5
// ("FONT mode q11 batch of WOFF2-transformed glyf/loca blobs").  Not part of the reference (brotli-lib only points at
// WOFF2: README.md:63; its bench fixtures came from a sibling repository, bench/create-fixtures.ts:17); the format is
// W3C WOFF 2.0 section 5.1 "Transformed glyf table format" and 5.3 (loca is reconstructed by the decoder).
//
// The transform is a data-parallel byte-stream rewrite: every glyph record is independent once loca gives its
// extent, and the position of its pieces in the seven output sub-streams is an exclusive prefix sum of per-glyph
// sizes.  Three launches over a batch of fonts:
//   k_glyf_measure  one thread per glyph: parse the record, sizes of its contribution to every sub-stream
//   k_glyf_scan     one CTA per font: prefix sums, stream bases, the 36-byte header
//   k_glyf_write    one thread per glyph: re-walk the record and write (255UInt16 point counts, flag bytes,
//                   coordinate triplets, composite records, explicit bounding boxes, instructions, bitmaps)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bgpu {

constexpr int GLYF_NS = 8;   // per-glyph size slots: 0 nPoints, 1 flags, 2 glyph, 3 composite, 4 bbox values, 5 instructions, 6 overlap bit, 7 explicit-bbox bit
struct GlyfFont {
  const uint8_t* glyf; const uint8_t* loca; uint8_t* out;
  uint32_t glyf_len, loca_len, num_glyphs, index_format;
  uint32_t* sizes;     // [num_glyphs][GLYF_NS]
  uint32_t* offs;      // [num_glyphs][GLYF_NS] exclusive prefix of sizes (slots 0..5)
  uint32_t* meta;      // [16]: 0..8 stream bases (nContour, nPoints, flags, glyph, composite, bbox bitmap, bbox values, instructions, overlap bitmap),
                       //       9 total bytes, 10 any-overlap, 11 status (0 ok, 1 malformed table)
};

__device__ __forceinline__ uint32_t be16(const uint8_t* p) { return ((uint32_t)p[0] << 8) | p[1]; }
__device__ __forceinline__ int be16s(const uint8_t* p) { return (int)(int16_t)be16(p); }
__device__ __forceinline__ uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }
__device__ __forceinline__ uint32_t size_255u16(uint32_t v) { return v < 253 ? 1u : (v < 762 ? 2u : 3u); }
__device__ __forceinline__ uint32_t put_255u16(uint8_t* o, uint32_t v) {   // WOFF2 section 4.1 / 5.1
  if (v < 253) { o[0] = (uint8_t)v; return 1; }
  if (v < 506) { o[0] = 255; o[1] = (uint8_t)(v - 253); return 2; }
  if (v < 762) { o[0] = 254; o[1] = (uint8_t)(v - 506); return 2; }
  o[0] = 253; o[1] = (uint8_t)(v >> 8); o[2] = (uint8_t)v; return 3;
}
// Triplet encoding of one point (WOFF2 section 5.2): flag byte + 1..4 data bytes.  Returns the data byte count.
template <bool WRITE>
__device__ __forceinline__ uint32_t triplet(bool on_curve, int x, int y, uint8_t* flag, uint8_t* d) {
  const uint32_t ax = (uint32_t)(x < 0 ? -x : x), ay = (uint32_t)(y < 0 ? -y : y);
  const uint32_t oc = on_curve ? 0u : 128u, xs = x < 0 ? 0u : 1u, ys = y < 0 ? 0u : 1u, xy = xs + 2 * ys;
  if (x == 0 && ay < 1280) { if (WRITE) { *flag = (uint8_t)(oc + ((ay & 0xF00) >> 7) + ys); d[0] = (uint8_t)ay; } return 1; }
  if (y == 0 && ax < 1280) { if (WRITE) { *flag = (uint8_t)(oc + 10 + ((ax & 0xF00) >> 7) + xs); d[0] = (uint8_t)ax; } return 1; }
  if (ax < 65 && ay < 65) {
    if (WRITE) { *flag = (uint8_t)(oc + 20 + ((ax - 1) & 0x30) + (((ay - 1) & 0x30) >> 2) + xy); d[0] = (uint8_t)((((ax - 1) & 0xF) << 4) | ((ay - 1) & 0xF)); }
    return 1;
  }
  if (ax < 769 && ay < 769) {
    if (WRITE) { *flag = (uint8_t)(oc + 84 + 12 * (((ax - 1) & 0x300) >> 8) + (((ay - 1) & 0x300) >> 6) + xy); d[0] = (uint8_t)(ax - 1); d[1] = (uint8_t)(ay - 1); }
    return 2;
  }
  if (ax < 4096 && ay < 4096) {
    if (WRITE) { *flag = (uint8_t)(oc + 120 + xy); d[0] = (uint8_t)(ax >> 4); d[1] = (uint8_t)(((ax & 0xF) << 4) | (ay >> 8)); d[2] = (uint8_t)ay; }
    return 3;
  }
  if (WRITE) { *flag = (uint8_t)(oc + 124 + xy); d[0] = (uint8_t)(ax >> 8); d[1] = (uint8_t)ax; d[2] = (uint8_t)(ay >> 8); d[3] = (uint8_t)ay; }
  return 4;
}

struct GlyphOut { uint8_t *npts, *flags, *glyph, *comp, *bbox, *instr; };

// Walks one glyph record g[0..len).  WRITE = false: fills sz[GLYF_NS]; WRITE = true: writes the pieces at O.
// Returns the contour count to store in nContourStream (0 empty, -1 composite), or -2 for a malformed record.
template <bool WRITE>
__device__ int glyph_walk(const uint8_t* g, uint32_t len, uint32_t* sz, GlyphOut O) {
  if (!WRITE) for (int k = 0; k < GLYF_NS; k++) sz[k] = 0;
  if (len == 0) return 0;
  if (len < 10) return -2;
  const int nc = be16s(g);
  if (nc == 0) return 0;
  if (nc < 0) {
    // composite: the component records are copied; instructions (if any) go to their own stream
    uint32_t pos = 10; bool have_instr = false;
    for (;;) {
      if (pos + 4 > len) return -2;
      const uint32_t f = be16(g + pos);
      uint32_t size = 4 + ((f & 1) ? 4 : 2);
      size += (f & 8) ? 2 : ((f & 0x40) ? 4 : ((f & 0x80) ? 8 : 0));
      pos += size;
      if (pos > len) return -2;
      have_instr |= (f & 0x100) != 0;
      if (!(f & 0x20)) break;
    }
    uint32_t ilen = 0;
    if (have_instr) { if (pos + 2 > len) return -2; ilen = be16(g + pos); if (pos + 2 + ilen > len) return -2; }
    if (!WRITE) { sz[3] = pos - 10; sz[2] = have_instr ? size_255u16(ilen) : 0; sz[5] = ilen; sz[4] = 8; sz[7] = 1; }
    else {
      for (uint32_t k = 0; k < pos - 10; k++) O.comp[k] = g[10 + k];
      if (have_instr) { put_255u16(O.glyph, ilen); for (uint32_t k = 0; k < ilen; k++) O.instr[k] = g[pos + 2 + k]; }
      for (int k = 0; k < 8; k++) O.bbox[k] = g[2 + k];
    }
    return -1;
  }
  // simple glyph
  uint32_t pos = 10 + 2 * (uint32_t)nc;
  if (pos + 2 > len) return -2;
  const uint32_t npts = be16(g + pos - 2) + 1;
  const uint32_t ilen = be16(g + pos);
  const uint8_t* instr = g + pos + 2;
  pos += 2 + ilen;
  if (pos > len) return -2;
  // pass A over the flags (with repeats): where the x and y coordinate arrays start
  uint32_t fpos = pos, xbytes = 0, ybytes = 0;
  for (uint32_t n = 0; n < npts;) {
    if (fpos >= len) return -2;
    const uint32_t f = g[fpos++];
    uint32_t reps = 1;
    if (f & 8) { if (fpos >= len) return -2; reps += g[fpos++]; }
    if (n + reps > npts) reps = npts - n;
    xbytes += reps * ((f & 2) ? 1u : ((f & 16) ? 0u : 2u));
    ybytes += reps * ((f & 4) ? 1u : ((f & 32) ? 0u : 2u));
    n += reps;
  }
  if (fpos + xbytes + ybytes > len) return -2;
  // pass B: points in order with three cursors
  uint32_t fc = pos, xc = fpos, yc = fpos + xbytes, rep = 0, f = 0;
  int x = 0, y = 0, lx = 0, ly = 0, x0 = 32767, y0 = 32767, x1 = -32768, y1 = -32768;
  uint32_t gbytes = 0, np_bytes = 0; bool overlap = false;
  if (WRITE || true) {
    int prev_end = -1;
    for (int c = 0; c < nc; c++) {
      const int e = (int)be16(g + 10 + 2 * c);
      const uint32_t cnt = (uint32_t)(e - prev_end) & 0xFFFFu; prev_end = e;
      if (WRITE) np_bytes += put_255u16(O.npts + np_bytes, cnt); else np_bytes += size_255u16(cnt);
    }
  }
  for (uint32_t n = 0; n < npts; n++) {
    if (rep) rep--;
    else { f = g[fc++]; if (f & 8) rep = g[fc++]; if (n == 0) overlap = (f & 0x40) != 0; }
    if (f & 2) { const int d = g[xc++]; x += (f & 16) ? d : -d; } else if (!(f & 16)) { x += be16s(g + xc); xc += 2; }
    if (f & 4) { const int d = g[yc++]; y += (f & 32) ? d : -d; } else if (!(f & 32)) { y += be16s(g + yc); yc += 2; }
    x = (int)(int16_t)x; y = (int)(int16_t)y;   // coordinates are 16-bit in the font
    uint8_t fl = 0;
    gbytes += triplet<WRITE>((f & 1) != 0, x - lx, y - ly, WRITE ? O.flags + n : &fl, WRITE ? O.glyph + gbytes : &fl);
    lx = x; ly = y;
    x0 = min(x0, x); y0 = min(y0, y); x1 = max(x1, x); y1 = max(y1, y);
  }
  const bool explicit_bbox = be16s(g + 2) != x0 || be16s(g + 4) != y0 || be16s(g + 6) != x1 || be16s(g + 8) != y1;
  if (!WRITE) {
    sz[0] = np_bytes; sz[1] = npts; sz[2] = gbytes + size_255u16(ilen); sz[5] = ilen; sz[4] = explicit_bbox ? 8 : 0; sz[6] = overlap ? 1 : 0; sz[7] = explicit_bbox ? 1 : 0;
  } else {
    put_255u16(O.glyph + gbytes, ilen);
    for (uint32_t k = 0; k < ilen; k++) O.instr[k] = instr[k];
    if (explicit_bbox) for (int k = 0; k < 8; k++) O.bbox[k] = g[2 + k];
  }
  return nc;
}

__device__ __forceinline__ bool glyph_extent(const GlyfFont& F, uint32_t gid, uint32_t* off, uint32_t* len) {
  uint32_t a, b;
  if (F.index_format == 0) { a = 2 * be16(F.loca + 2 * gid); b = 2 * be16(F.loca + 2 * gid + 2); }
  else { a = be32(F.loca + 4 * gid); b = be32(F.loca + 4 * gid + 4); }
  if (b < a || b > F.glyf_len) return false;
  *off = a; *len = b - a;
  return true;
}

// grid (ceil(max glyphs / 128), fonts)
__global__ void __launch_bounds__(128) k_glyf_measure(const GlyfFont* __restrict__ fonts) {
  const GlyfFont F = fonts[blockIdx.y];
  const uint32_t gid = blockIdx.x * 128 + threadIdx.x;
  if (gid >= F.num_glyphs) return;
  uint32_t off, len; uint32_t* sz = F.sizes + (size_t)gid * GLYF_NS;
  GlyphOut none{};
  int nc = -2;
  if (glyph_extent(F, gid, &off, &len)) nc = glyph_walk<false>(F.glyf + off, len, sz, none);
  if (nc == -2) { for (int k = 0; k < GLYF_NS; k++) sz[k] = 0; atomicExch(&F.meta[11], 1u); }
}

// one CTA (256 threads) per font: exclusive prefix sums of slots 0..5, stream bases, header
__global__ void __launch_bounds__(256) k_glyf_scan(const GlyfFont* __restrict__ fonts) {
  __shared__ uint32_t s_part[256]; __shared__ uint32_t s_run[6]; __shared__ uint32_t s_any;
  const GlyfFont F = fonts[blockIdx.x];
  const uint32_t tid = threadIdx.x, ng = F.num_glyphs;
  if (tid < 6) s_run[tid] = 0;
  if (tid == 0) s_any = 0;
  __syncthreads();
  for (int s = 0; s < 6; s++) {
    for (uint32_t g0 = 0; g0 < ng; g0 += 256) {
      const uint32_t g = g0 + tid;
      const uint32_t v = g < ng ? F.sizes[(size_t)g * GLYF_NS + s] : 0;
      s_part[tid] = v;
      __syncthreads();
      for (int o = 1; o < 256; o <<= 1) { const uint32_t add = tid >= (uint32_t)o ? s_part[tid - o] : 0; __syncthreads(); s_part[tid] += add; __syncthreads(); }
      if (g < ng) F.offs[(size_t)g * GLYF_NS + s] = s_run[s] + s_part[tid] - v;
      __syncthreads();
      if (tid == 255) s_run[s] += s_part[255];
      __syncthreads();
    }
  }
  for (uint32_t g = tid; g < ng; g += 256) if (F.sizes[(size_t)g * GLYF_NS + 6]) atomicOr(&s_any, 1u);
  __syncthreads();
  if (tid == 0) {
    const uint32_t bitmap = 4 * ((ng + 31) / 32), any = s_any;
    uint32_t* m = F.meta;
    m[0] = 36; m[1] = m[0] + 2 * ng; m[2] = m[1] + s_run[0]; m[3] = m[2] + s_run[1]; m[4] = m[3] + s_run[2]; m[5] = m[4] + s_run[3];
    m[6] = m[5] + bitmap; m[7] = m[6] + s_run[4]; m[8] = m[7] + s_run[5]; m[9] = m[8] + (any ? (ng + 7) / 8 : 0); m[10] = any;
  }
}

// grid (ceil(max glyphs / 128), fonts); F.out holds meta[9] bytes
__global__ void __launch_bounds__(128) k_glyf_write(const GlyfFont* __restrict__ fonts) {
  const GlyfFont F = fonts[blockIdx.y];
  const uint32_t gid = blockIdx.x * 128 + threadIdx.x;
  if (gid >= F.num_glyphs || F.meta[11]) return;
  const uint32_t* m = F.meta; uint8_t* o = F.out;
  if (gid == 0) {
    const uint32_t hdr[7] = {m[2] - m[1], m[3] - m[2], m[4] - m[3], m[5] - m[4], 2 * F.num_glyphs, m[7] - m[5], m[8] - m[7]};
    // reserved, optionFlags (bit 0: overlapSimpleBitmap), numGlyphs, indexFormat, then the seven stream sizes
    o[0] = 0; o[1] = 0; o[2] = 0; o[3] = (uint8_t)(m[10] ? 1 : 0); o[4] = (uint8_t)(F.num_glyphs >> 8); o[5] = (uint8_t)F.num_glyphs;
    o[6] = (uint8_t)(F.index_format >> 8); o[7] = (uint8_t)F.index_format;
    const uint32_t order[7] = {hdr[4], hdr[0], hdr[1], hdr[2], hdr[3], hdr[5], hdr[6]};   // nContour, nPoints, flag, glyph, composite, bbox, instruction
    for (int k = 0; k < 7; k++) { const uint32_t v = order[k]; o[8 + 4 * k] = (uint8_t)(v >> 24); o[9 + 4 * k] = (uint8_t)(v >> 16); o[10 + 4 * k] = (uint8_t)(v >> 8); o[11 + 4 * k] = (uint8_t)v; }
  }
  uint32_t off = 0, len = 0;
  glyph_extent(F, gid, &off, &len);
  const uint32_t* of = F.offs + (size_t)gid * GLYF_NS;
  GlyphOut O{o + m[1] + of[0], o + m[2] + of[1], o + m[3] + of[2], o + m[4] + of[3], o + m[6] + of[4], o + m[7] + of[5]};
  uint32_t dummy[GLYF_NS];
  const int nc = glyph_walk<true>(F.glyf + off, len, dummy, O);
  o[m[0] + 2 * gid] = (uint8_t)((uint32_t)nc >> 8); o[m[0] + 2 * gid + 1] = (uint8_t)nc;
  if ((gid & 7) == 0) {   // one bitmap byte per 8 glyphs, most significant bit first
    uint32_t bb = 0, ov = 0;
    for (uint32_t k = 0; k < 8 && gid + k < F.num_glyphs; k++) {
      const uint32_t* s = F.sizes + (size_t)(gid + k) * GLYF_NS;
      if (s[7]) bb |= 0x80u >> k;
      if (s[6]) ov |= 0x80u >> k;
    }
    o[m[5] + (gid >> 3)] = (uint8_t)bb;
    if (m[10]) o[m[8] + (gid >> 3)] = (uint8_t)ov;
  }
  // bitmap padding bytes beyond ceil(ng / 8) (the bbox bitmap is 4 * ceil(ng / 32) bytes)
  if (gid == 0) for (uint32_t k = (F.num_glyphs + 7) / 8; k < 4 * ((F.num_glyphs + 31) / 32); k++) o[m[5] + k] = 0;
}

}  // namespace bgpu
