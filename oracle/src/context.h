// ORACLE (test infrastructure) — restates context.ts.
#pragma once
#include "common.h"

namespace oracle {

// context.ts:12-162: 4 modes x (256 last-byte + 256 second-last-byte) entries.  This is the
// RFC 7932 §7.1 context lookup; it is regenerated here from the RFC's rules rather than
// carried as a literal (tests/test_oracle_tables.py compares it with the TS literal when the
// reference tree is present).  Mode ids: LSB6=0, MSB6=1, UTF8=2, SIGNED=3 (context.ts:4-9).
struct ContextLut {
  uint8_t t[2048];
  ContextLut() {
    for (int i = 0; i < 256; i++) {
      t[i] = i & 63; t[256 + i] = 0;                 // LSB6
      t[512 + i] = i >> 2; t[768 + i] = 0;           // MSB6
      int s = i == 0 ? 0 : i < 16 ? 1 : i < 64 ? 2 : i < 128 ? 3 : i < 192 ? 4 : i < 240 ? 5 : i < 255 ? 6 : 7;
      t[1536 + i] = s << 3; t[1792 + i] = s;         // SIGNED
    }
    // UTF8, last byte (Lut0): classes scaled by 4
    uint8_t* u0 = t + 1024; uint8_t* u1 = t + 1280;
    for (int i = 0; i < 256; i++) { u0[i] = 0; u1[i] = 0; }
    u0[9] = u0[10] = u0[13] = 4;
    static const uint8_t punct[33] = {8, 12, 16, 12, 12, 20, 12, 16, 24, 28, 12, 12, 32, 12, 36, 12,
                                      44, 44, 44, 44, 44, 44, 44, 44, 44, 44, 32, 32, 24, 40, 28, 12, 12};
    for (int i = 0; i < 33; i++) u0[32 + i] = punct[i];
    for (int c = 'A'; c <= 'Z'; c++) u0[c] = (c == 'A' || c == 'E' || c == 'I' || c == 'O' || c == 'U') ? 48 : 52;
    static const uint8_t p2[6] = {24, 12, 28, 12, 12, 12};
    for (int i = 0; i < 6; i++) u0[91 + i] = p2[i];
    for (int c = 'a'; c <= 'z'; c++) u0[c] = (c == 'a' || c == 'e' || c == 'i' || c == 'o' || c == 'u') ? 56 : 60;
    static const uint8_t p3[5] = {24, 12, 28, 12, 0};
    for (int i = 0; i < 5; i++) u0[123 + i] = p3[i];
    for (int i = 128; i < 192; i++) u0[i] = i & 1;
    for (int i = 192; i < 256; i++) u0[i] = 2 + (i & 1);
    // UTF8, second-last byte (Lut1)
    for (int i = 33; i < 127; i++) u1[i] = 1;
    for (int c = '0'; c <= '9'; c++) u1[c] = 2;
    for (int c = 'A'; c <= 'Z'; c++) u1[c] = 2;
    for (int c = 'a'; c <= 'z'; c++) u1[c] = 3;
    for (int i = 224; i < 256; i++) u1[i] = 2;
  }
};
inline const uint8_t* ContextLookupTable() { static ContextLut L; return L.t; }
inline const uint8_t* GetContextLut(int mode) { return ContextLookupTable() + (mode << 9); }  // context.ts:164-166
static inline uint32_t GetContext(uint8_t p1, uint8_t p2, const uint8_t* lut) { return lut[p1] | lut[256 + p2]; }  // :169-171

// context.ts:179-227 (the reference's own heuristic, not the C encoder's)
static inline int ChooseContextMode(const Bytes& data, size_t start, size_t length) {
  size_t ascii = 0, lead = 0, cont = 0, signedPat = 0;
  size_t sampleSize = std::min<size_t>(length, 1024);
  size_t step = std::max<size_t>(1, sampleSize ? length / sampleSize : 1);
  for (size_t i = 0; i < length; i += step) {
    int byte = data.at(start + i);
    if (byte < 0) { cont++; continue; }  // `undefined` fails both comparisons -> else branch
    if (byte < 128) {
      ascii++;
      if (i > 0) {
        int prev = data.at(start + i - 1);
        int diff = std::abs(byte - prev);
        if (diff < 16) signedPat++;
      }
    } else if (byte >= 192) lead++;
    else cont++;
  }
  size_t total = ascii + lead + cont;
  if (total == 0) return 0;
  if (lead > 0 && (double)cont > (double)lead * 0.5) return 2;
  if ((double)signedPat > (double)total * 0.3) return 3;
  if ((double)ascii > (double)total * 0.7) return 2;
  return 0;
}

}  // namespace oracle
