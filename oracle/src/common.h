// ORACLE — TEST INFRASTRUCTURE ONLY.
// CPU restatement (C++17) of countertype/brotli-lib's TypeScript encoder, used as the
// parity checker for the CUDA path and as the timed CPU baseline.  Nothing under
// brotli_lib_b200/ (the product) may include, link or call this code.
//
// PARITY UNPINNED: the reference's tests hold no golden encoder bytes and
// no JavaScript runtime exists in the build container, so this restatement is pinned only
// by (1) following the TS source line by line (citations: file:line relative to
// /root/reference/src/encode/) and (2) the properties the reference's own tests assert
// (round-trip identity through an independent decoder, the <20 % ratio case, A1 01 for
// empty input).
#pragma once
#include <cstdint>
#include <cstddef>
#include <cstring>
#include <cmath>
#include <vector>
#include <algorithm>

namespace oracle {

// fast-log.ts:99-103  log2FloorNonZero = 31 - clz32(n | 0)
static inline int Log2FloorNonZero(uint32_t n) { return n == 0 ? -1 : 31 - __builtin_clz(n); }

// fast-log.ts:6-93 kLog2Table (doubles, log2(0) := 0) and :106-111 fastLog2.
// The table is regenerated with std::log2 (exact for the 256 integer arguments to the
// printed precision; tests/test_oracle_tables.py compares it with the TS literal when the
// reference tree is present).
struct Log2Table {
  double t[256];
  Log2Table() { t[0] = 0.0; for (int i = 1; i < 256; i++) t[i] = std::log2((double)i); }
};
inline const double* kLog2Table() { static Log2Table T; return T.t; }
static inline double FastLog2(double v) {
  if (v < 256.0) return kLog2Table()[(int)v];
  return std::log(v) * 1.4426950408889634;
}

// JS typed-array read: out-of-range index yields `undefined`; we model it as -1 so that
// comparisons against any real byte fail exactly as `undefined !== byte` does.
struct Bytes {
  const uint8_t* p; size_t n;
  inline int at(size_t i) const { return i < n ? (int)p[i] : -1; }
  // `data[i] << k` / `data[i] | 0` with undefined -> 0 (match.ts:164-167)
  inline uint32_t at0(size_t i) const { return i < n ? p[i] : 0u; }
};

// command.ts:138-144
struct Command {
  uint32_t insertLen;
  uint32_t copyLen;     // len | (delta & 0x7F) << 25
  uint32_t distExtra;
  uint16_t cmdPrefix;
  uint16_t distPrefix;  // code | nbits << 10
};
static_assert(sizeof(Command) == 16, "Command is 16 bytes (shared layout with the device)");

static inline uint32_t CommandCopyLen(const Command& c) { return c.copyLen & 0x1FFFFFF; }
// command.ts:214-219
static inline uint32_t CommandCopyLenCode(const Command& c) {
  uint32_t modifier = c.copyLen >> 25;
  int32_t delta = (modifier & 0x40) ? (int32_t)(modifier | 0xFFFFFF80u) : (int32_t)modifier;
  return (uint32_t)((int32_t)(c.copyLen & 0x1FFFFFF) + delta);
}

}  // namespace oracle
