// ORACLE (test infrastructure) — restates bit-writer.ts.
#pragma once
#include "common.h"

namespace oracle {

// bit-writer.ts:10-160.  LSB-first bit packer.  The TS writeBits is exact for nBits <= 32
// (32-bit shifts); callers that need more (storeCommandExtra) are given a
// 64-bit path here — the deliberate deviation "fix (c)".
class BitWriter {
 public:
  std::vector<uint8_t> buf;
  uint64_t pos = 0;            // bit position
  uint64_t flushedBytePos = 0;

  explicit BitWriter(size_t initial = 4096) { buf.assign(initial ? initial : 1, 0); }

  void ensure(uint64_t bits) {
    uint64_t need = ((pos + bits + 7) >> 3) + 9;
    if (need > buf.size()) buf.resize(std::max<uint64_t>(buf.size() * 2, need), 0);
  }
  // bit-writer.ts:32-56
  void writeBits(unsigned nBits, uint64_t value) {
    if (nBits == 0) return;
    ensure(nBits);
    uint64_t bytePos = pos >> 3; unsigned off = pos & 7;
    unsigned __int128 v = (unsigned __int128)value << off;
    unsigned total = off + nBits;
    buf[bytePos] |= (uint8_t)v;            // OR into the partially filled byte
    v >>= 8;
    for (unsigned w = 8, i = 1; w < total; w += 8, i++) { buf[bytePos + i] = (uint8_t)v; v >>= 8; }
    pos += nBits;
  }
  // bit-writer.ts:112-118
  unsigned alignToByte() {
    unsigned padding = (8 - (pos & 7)) & 7;
    if (padding) writeBits(padding, 0);
    return padding;
  }
  // bit-writer.ts:101-109
  void writeBytes(const uint8_t* p, size_t n) {
    ensure(n * 8);
    if (n) memcpy(&buf[pos >> 3], p, n);
    pos += (uint64_t)n * 8;
  }
  // bit-writer.ts:152-159
  void prepareStorage() { ensure(8); buf[pos >> 3] = 0; }
  // bit-writer.ts:138-144: only completely written bytes (streaming)
  std::vector<uint8_t> takeBytes() {
    uint64_t end = pos >> 3;
    if (end <= flushedBytePos) return {};
    std::vector<uint8_t> out(buf.begin() + flushedBytePos, buf.begin() + end);
    flushedBytePos = end;
    return out;
  }
  // bit-writer.ts:146-150
  std::vector<uint8_t> finish() const {
    return std::vector<uint8_t>(buf.begin(), buf.begin() + ((pos + 7) >> 3));
  }
};

// bit-writer.ts:172-194 encodeWindowBits (largeWindow is never set by the public API)
static inline void EncodeWindowBits(int lgwin, unsigned* bits, unsigned* value) {
  if (lgwin == 16) { *value = 0; *bits = 1; }
  else if (lgwin == 17) { *value = 1; *bits = 7; }
  else if (lgwin > 17 && lgwin <= 24) { *value = ((lgwin - 17) << 1) | 1; *bits = 4; }
  else { *value = ((lgwin - 8) << 4) | 1; *bits = 7; }
}

}  // namespace oracle
