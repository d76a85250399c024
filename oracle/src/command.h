// ORACLE (test infrastructure) — restates command.ts.
#pragma once
#include "common.h"

namespace oracle {

// command.ts:11-27
static const uint32_t kInsBase[24] = {0, 1, 2, 3, 4, 5, 6, 8, 10, 14, 18, 26, 34, 50, 66, 98,
                                      130, 194, 322, 578, 1090, 2114, 6210, 22594};
static const uint32_t kInsExtra[24] = {0, 0, 0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5,
                                       6, 7, 8, 9, 10, 12, 14, 24};
static const uint32_t kCopyBase[24] = {2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 18, 22, 30, 38, 54,
                                       70, 102, 134, 198, 326, 582, 1094, 2118};
static const uint32_t kCopyExtra[24] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4,
                                        5, 5, 6, 7, 8, 9, 10, 24};

// command.ts:29-44
static inline uint32_t GetInsertLengthCode(uint32_t insertLen) {
  if (insertLen < 6) return insertLen;
  if (insertLen < 130) {
    uint32_t nbits = Log2FloorNonZero(insertLen - 2) - 1;
    return (nbits << 1) + ((insertLen - 2) >> nbits) + 2;
  }
  if (insertLen < 2114) return Log2FloorNonZero(insertLen - 66) + 10;
  if (insertLen < 6210) return 21;
  if (insertLen < 22594) return 22;
  return 23;
}
// command.ts:46-57
static inline uint32_t GetCopyLengthCode(uint32_t copyLen) {
  if (copyLen < 10) return copyLen - 2;
  if (copyLen < 134) {
    uint32_t nbits = Log2FloorNonZero(copyLen - 6) - 1;
    return (nbits << 1) + ((copyLen - 6) >> nbits) + 4;
  }
  if (copyLen < 2118) return Log2FloorNonZero(copyLen - 70) + 12;
  return 23;
}
// command.ts:59-81
static inline uint32_t CombineLengthCodes(uint32_t insCode, uint32_t copyCode, bool useLastDistance) {
  uint32_t bits64 = (copyCode & 7) | ((insCode & 7) << 3);
  if (useLastDistance && insCode < 8 && copyCode < 16) return copyCode < 8 ? bits64 : (bits64 | 64);
  uint32_t offset = 2 * ((copyCode >> 3) + 3 * (insCode >> 3));
  offset = (offset << 5) + 0x40 + ((0x520D40u >> offset) & 0xC0);
  return offset | bits64;
}
// command.ts:83-91
static inline uint32_t GetLengthCode(uint32_t insertLen, uint32_t copyLen, bool useLastDistance) {
  return CombineLengthCodes(GetInsertLengthCode(insertLen), GetCopyLengthCode(copyLen), useLastDistance);
}
// command.ts:111-135.  Returns code|nbits<<10 semantics split as in the TS tuple.
static inline void PrefixEncodeCopyDistance(uint32_t distanceCode, uint32_t ndirect, uint32_t npostfix,
                                            uint32_t* code, uint32_t* extraBits, uint32_t* nbitsOut) {
  if (distanceCode < 16 + ndirect) { *code = distanceCode; *extraBits = 0; *nbitsOut = 0; return; }
  uint32_t dist = (1u << (npostfix + 2)) + (distanceCode - 16 - ndirect);
  uint32_t bucket = Log2FloorNonZero(dist) - 1;
  uint32_t postfixMask = (1u << npostfix) - 1;
  uint32_t postfix = dist & postfixMask;
  uint32_t prefix = (dist >> bucket) & 1;
  uint32_t offset = (2 + prefix) << bucket;
  uint32_t nbits = bucket - npostfix;
  // NB: the TS packs nbits<<10 into the returned code as well; createCommand ORs nbits<<10
  // again, which is idempotent.
  *code = (nbits << 10) | (16 + ndirect + ((2 * (nbits - 1) + prefix) << npostfix) + postfix);
  *extraBits = (dist - offset) >> npostfix;
  *nbitsOut = nbits;
}
// command.ts:146-179
static inline Command CreateCommand(uint32_t insertLen, uint32_t copyLen, int copyLenCodeDelta,
                                    uint32_t distanceCode, uint32_t ndirect, uint32_t npostfix) {
  Command c;
  uint32_t delta = (uint32_t)copyLenCodeDelta & 0x7F;
  uint32_t code, extra, nbits;
  PrefixEncodeCopyDistance(distanceCode, ndirect, npostfix, &code, &extra, &nbits);
  c.insertLen = insertLen;
  c.copyLen = copyLen | (delta << 25);
  c.distExtra = extra;
  c.distPrefix = (uint16_t)(code | (nbits << 10));
  bool useLast = (code & 0x3FF) == 0;
  c.cmdPrefix = (uint16_t)GetLengthCode(insertLen, copyLen + copyLenCodeDelta, useLast);
  return c;
}
// command.ts:183-208
static inline Command CreateInsertCommand(uint32_t insertLen) {
  Command c;
  uint32_t insCode = GetInsertLengthCode(insertLen);
  c.cmdPrefix = (uint16_t)GetLengthCode(insertLen, 2, insCode < 8);
  c.insertLen = insertLen;
  c.copyLen = 0u | (2u << 25);
  c.distExtra = 0;
  c.distPrefix = 0;
  return c;
}

}  // namespace oracle
