// CPU build of the __host__ __device__ format code in codec_common.cuh, so that the exact
// functions the kernels run (Huffman construction, prefix-code serialisation, command coding)
// are checked against the oracle in the CPU-only test tier.  Not part of the product library.
#include "codec_common.cuh"
#include <cstring>
#include <vector>
using namespace bgpu;

extern "C" {

// Mirrors k_huff for one histogram: returns the serialised tree length in bits.
int bgpu_hosttest_huffman(const uint32_t* hist, int alphabet, uint8_t* depth, uint16_t* bits, uint32_t* tree_words /*256*/) {
  std::vector<uint16_t> sorted(alphabet), lt(2 * alphabet + 2), rt(2 * alphabet + 2);
  std::vector<uint32_t> cnt(2 * alphabet + 2);
  std::vector<uint8_t> nd(2 * alphabet + 2), rs(alphabet + 8), re(alphabet + 8);
  memset(depth, 0, alphabet); memset(bits, 0, 2 * alphabet); memset(tree_words, 0, 256 * 4);
  uint16_t s4[4] = {0, 0, 0, 0}; int n = 0;
  for (int i = 0; i < alphabet; i++) if (hist[i]) { if (n < 4) s4[n] = (uint16_t)i; n++; }
  BitSink w{tree_words, 0};
  if (n <= 1) { tree_store_single(w, alphabet, s4[0]); return (int)w.pos; }
  for (uint32_t limit = 1;; limit *= 2) {
    huff_sort_serial(hist, alphabet, limit, sorted.data());
    if (huff_depths_from_sorted(hist, sorted.data(), n, limit, 15, cnt.data(), lt.data(), rt.data(), nd.data(), depth)) break;
    memset(depth, 0, alphabet);
  }
  huff_codes(depth, alphabet, bits);
  tree_store(w, depth, alphabet, n, s4, rs.data(), re.data());
  return (int)w.pos;
}

void bgpu_hosttest_make_cmd(uint32_t insertLen, uint32_t copyLen, uint32_t distanceCode, uint32_t ndirect, uint32_t npostfix, Cmd* out) {
  *out = make_cmd(insertLen, copyLen, distanceCode, ndirect, npostfix);
}
void bgpu_hosttest_make_insert_cmd(uint32_t insertLen, Cmd* out) { *out = make_insert_cmd(insertLen); }
uint32_t bgpu_hosttest_distance_of_explicit(uint32_t distPrefix, uint32_t distExtra, uint32_t ndirect, uint32_t npostfix) {
  return distance_of_explicit(distPrefix, distExtra, ndirect, npostfix);
}
// encodeContextMap (context-map.ts:114-170) exactly as k_mb_header / k_stage_context_map run it.  words: zeroed, size + 1024.
uint32_t bgpu_hosttest_context_map(const uint32_t* cmap, uint32_t size, uint32_t ncl, uint32_t* words) {
  std::vector<uint32_t> scratch(size + 1);
  static SerialHuff S;
  BitSink w{words, 0};
  encode_context_map(w, S, cmap, size, ncl, scratch.data());
  return w.pos;
}
uint32_t bgpu_hosttest_cmd_extra(uint32_t insertLen, uint32_t copyLenField, uint64_t* value) { return cmd_extra(insertLen, copyLenField, value); }

}
