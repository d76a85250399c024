// Node-API addon: the C ABI of libbrotligpu.so (include/brotligpu.h) behind brotli-lib's encode API.
//
// Binds, 1:1:
//   encode(input, quality, lgwin, mode)            -> bgpu_encode            brotliEncode        src/encode/encode.ts:50-90
//   encodeBatch(inputs[], quality, lgwin, mode)    -> bgpu_encode_batch      (BASELINE config 5: independent streams)
//   encodePart(part, prev2, first, last, q, w, m)  -> bgpu_encode_part       one part of a stream (worker per GPU)
//   new Encoder(quality, lgwin, mode).update/finish-> bgpu_stream_*          BrotliEncoder       src/encode/encode.ts:290-409
//   init(deviceIds[])                               -> bgpu_ctx_create_multi one context over several GPUs
// node/encode.ts wraps these with the reference's names, option clamping and host-only cases.
//
// There is no Node toolchain (node, node-gyp, node_api.h) in the build image of this repository: the file is compiled
// only where <node_api.h> exists (node-gyp: node/binding.gyp); tests/test_hostdev_units.py checks its syntax against a
// declaration-only stub of the Node-API subset it uses (tests/stubs/node_api.h).  The same entry points are exercised
// end to end through ctypes (brotli_lib_b200/_lib.py).
#if defined(__has_include)
#if __has_include(<node_api.h>)
#define BGPU_HAVE_NODE_API 1
#endif
#endif

#ifdef BGPU_HAVE_NODE_API
#include <node_api.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "../include/brotligpu.h"

namespace {

bgpu_ctx* g_ctx = nullptr;   // one context on the main thread (the reference is synchronous and single-threaded)

napi_value throw_bgpu(napi_env env, int rc) {
  std::string msg = std::string("brotligpu: ") + bgpu_strerror(rc);
  const char* detail = bgpu_last_cuda_error();
  if (detail && *detail && (rc == BGPU_E_CUDA || rc == BGPU_E_NO_DEVICE || rc == BGPU_E_NOMEM)) msg += std::string(": ") + detail;
  napi_throw_error(env, nullptr, msg.c_str());
  return nullptr;
}

bool ensure_ctx(napi_env env) {
  if (g_ctx) return true;
  int rc = bgpu_ctx_create(-1, &g_ctx);
  if (rc) { throw_bgpu(env, rc); return false; }
  return true;
}

// Uint8Array (or any TypedArray / Buffer) -> (pointer, byte length); a zero-copy view of the caller's memory.
bool get_bytes(napi_env env, napi_value v, const uint8_t** p, size_t* n) {
  bool is_ta = false;
  if (napi_is_typedarray(env, v, &is_ta) != napi_ok || !is_ta) { napi_throw_type_error(env, nullptr, "expected a Uint8Array"); return false; }
  napi_typedarray_type t; size_t len; void* data; napi_value ab; size_t off;
  if (napi_get_typedarray_info(env, v, &t, &len, &data, &ab, &off) != napi_ok) return false;
  size_t elem = (t == napi_uint8_array || t == napi_int8_array || t == napi_uint8_clamped_array) ? 1 : (t == napi_uint16_array || t == napi_int16_array) ? 2 :
                (t == napi_float64_array || t == napi_bigint64_array || t == napi_biguint64_array) ? 8 : 4;
  *p = (const uint8_t*)data; *n = len * elem;
  return true;
}

bool get_i32(napi_env env, napi_value v, int32_t* out) { return napi_get_value_int32(env, v, out) == napi_ok; }

// A fresh, caller-owned Uint8Array holding exactly `len` bytes (bit-writer.ts:146-150 returns a slice).
napi_value make_u8(napi_env env, const uint8_t* src, size_t len) {
  void* data = nullptr; napi_value ab, arr;
  if (napi_create_arraybuffer(env, len, &data, &ab) != napi_ok) return nullptr;
  if (len) memcpy(data, src, len);
  if (napi_create_typedarray(env, napi_uint8_array, len, ab, 0, &arr) != napi_ok) return nullptr;
  return arr;
}

// encode(input: Uint8Array, quality, lgwin, mode): Uint8Array
napi_value Encode(napi_env env, napi_callback_info info) {
  size_t argc = 4; napi_value a[4];
  napi_get_cb_info(env, info, &argc, a, nullptr, nullptr);
  const uint8_t* in; size_t n; int32_t q = 11, w = 22, m = 0;
  if (argc < 1 || !get_bytes(env, a[0], &in, &n)) return nullptr;
  if (argc > 1) get_i32(env, a[1], &q);
  if (argc > 2) get_i32(env, a[2], &w);
  if (argc > 3) get_i32(env, a[3], &m);
  if (!ensure_ctx(env)) return nullptr;
  size_t cap = bgpu_max_compressed_size(n), len = 0;
  uint8_t* buf = (uint8_t*)malloc(cap ? cap : 1);
  if (!buf) { napi_throw_error(env, nullptr, "brotligpu: out of host memory"); return nullptr; }
  int rc = bgpu_encode(g_ctx, in, n, q, w, m, buf, cap, &len);
  if (rc == BGPU_E_OUTPUT_TOO_SMALL) {   // custom tuning below the default shard size: the call reports the size needed
    free(buf); cap = len; buf = (uint8_t*)malloc(cap ? cap : 1);
    if (!buf) { napi_throw_error(env, nullptr, "brotligpu: out of host memory"); return nullptr; }
    rc = bgpu_encode(g_ctx, in, n, q, w, m, buf, cap, &len);
  }
  napi_value out = rc ? throw_bgpu(env, rc) : make_u8(env, buf, len);
  free(buf);
  return out;
}

// encodeBatch(inputs: Uint8Array[], quality, lgwin, mode): Uint8Array[]
napi_value EncodeBatch(napi_env env, napi_callback_info info) {
  size_t argc = 4; napi_value a[4];
  napi_get_cb_info(env, info, &argc, a, nullptr, nullptr);
  uint32_t count = 0; int32_t q = 11, w = 22, m = 0;
  if (argc < 1 || napi_get_array_length(env, a[0], &count) != napi_ok) { napi_throw_type_error(env, nullptr, "expected an array of Uint8Array"); return nullptr; }
  if (argc > 1) get_i32(env, a[1], &q);
  if (argc > 2) get_i32(env, a[2], &w);
  if (argc > 3) get_i32(env, a[3], &m);
  if (!ensure_ctx(env)) return nullptr;
  std::vector<const uint8_t*> in(count); std::vector<size_t> n(count), cap(count), len(count, 0); std::vector<uint8_t*> out(count, nullptr);
  size_t total = 0;
  for (uint32_t i = 0; i < count; i++) {
    napi_value e; napi_get_element(env, a[0], i, &e);
    if (!get_bytes(env, e, &in[i], &n[i])) return nullptr;
    cap[i] = bgpu_max_compressed_size(n[i]); total += cap[i];
  }
  std::vector<uint8_t> arena(total ? total : 1);
  size_t o = 0;
  for (uint32_t i = 0; i < count; i++) { out[i] = arena.data() + o; o += cap[i]; }
  int rc = bgpu_encode_batch(g_ctx, (int)count, in.data(), n.data(), q, w, m, out.data(), cap.data(), len.data());
  if (rc) return throw_bgpu(env, rc);
  napi_value arr; napi_create_array_with_length(env, count, &arr);
  for (uint32_t i = 0; i < count; i++) napi_set_element(env, arr, i, make_u8(env, out[i], len[i]));
  return arr;
}

// encodePart(part, prev2, first, last, quality, lgwin, mode): Uint8Array     (INTEGRATION.md "One input over several GPUs")
napi_value EncodePart(napi_env env, napi_callback_info info) {
  size_t argc = 7; napi_value a[7];
  napi_get_cb_info(env, info, &argc, a, nullptr, nullptr);
  if (argc < 7) { napi_throw_type_error(env, nullptr, "encodePart(part, prev2, first, last, quality, lgwin, mode)"); return nullptr; }
  const uint8_t *in, *pv; size_t n, pn; bool first = true, last = true; int32_t q, w, m;
  if (!get_bytes(env, a[0], &in, &n) || !get_bytes(env, a[1], &pv, &pn)) return nullptr;
  napi_get_value_bool(env, a[2], &first); napi_get_value_bool(env, a[3], &last);
  get_i32(env, a[4], &q); get_i32(env, a[5], &w); get_i32(env, a[6], &m);
  if (!first && pn != 2) { napi_throw_range_error(env, nullptr, "prev2 must hold the two bytes before a non-first part"); return nullptr; }
  if (!ensure_ctx(env)) return nullptr;
  size_t cap = bgpu_max_compressed_size(n), len = 0;
  std::vector<uint8_t> buf(cap ? cap : 1);
  int rc = bgpu_encode_part(g_ctx, in, n, first ? nullptr : pv, (first ? BGPU_PART_FIRST : 0u) | (last ? BGPU_PART_LAST : 0u), q, w, m, buf.data(), cap, &len);
  return rc ? throw_bgpu(env, rc) : make_u8(env, buf.data(), len);
}

// init(deviceIds: number[]): one context over several GPUs; brotliEncode of a large input then returns one stitched stream
napi_value Init(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value a[1];
  napi_get_cb_info(env, info, &argc, a, nullptr, nullptr);
  uint32_t count = 0;
  if (argc < 1 || napi_get_array_length(env, a[0], &count) != napi_ok || count == 0) { napi_throw_type_error(env, nullptr, "init(deviceIds: number[])"); return nullptr; }
  std::vector<int> ids(count);
  for (uint32_t i = 0; i < count; i++) { napi_value e; napi_get_element(env, a[0], i, &e); int32_t v = 0; get_i32(env, e, &v); ids[i] = v; }
  if (g_ctx) { bgpu_ctx_destroy(g_ctx); g_ctx = nullptr; }
  int rc = bgpu_ctx_create_multi(ids.data(), (int)count, &g_ctx);
  if (rc) return throw_bgpu(env, rc);
  napi_value u; napi_get_undefined(env, &u);
  return u;
}

// ---- class Encoder { constructor(quality, lgwin, mode); update(chunk): Uint8Array; finish(): Uint8Array }
void EncoderFinalize(napi_env, void* data, void*) { bgpu_stream_destroy((bgpu_stream*)data); }

napi_value EncoderNew(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value a[3], self;
  napi_get_cb_info(env, info, &argc, a, &self, nullptr);
  int32_t q = 11, w = 22, m = 0;
  if (argc > 0) get_i32(env, a[0], &q);
  if (argc > 1) get_i32(env, a[1], &w);
  if (argc > 2) get_i32(env, a[2], &m);
  if (!ensure_ctx(env)) return nullptr;
  bgpu_stream* s = nullptr;
  int rc = bgpu_stream_create(g_ctx, q, w, m, &s);
  if (rc) return throw_bgpu(env, rc);
  napi_wrap(env, self, s, EncoderFinalize, nullptr, nullptr);
  return self;
}

napi_value EncoderUpdate(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value a[1], self; void* p = nullptr;
  napi_get_cb_info(env, info, &argc, a, &self, nullptr);
  if (napi_unwrap(env, self, &p) != napi_ok || !p) { napi_throw_error(env, nullptr, "Encoder not initialized"); return nullptr; }   // encode.ts:416
  const uint8_t* in; size_t n;
  if (argc < 1 || !get_bytes(env, a[0], &in, &n)) return nullptr;
  const uint8_t* out = nullptr; size_t len = 0;
  int rc = bgpu_stream_update((bgpu_stream*)p, in, n, &out, &len);   // copies the chunk (encode.ts:360-363)
  return rc ? throw_bgpu(env, rc) : make_u8(env, out, len);          // *out is shim-owned: copied into a fresh Uint8Array
}

napi_value EncoderFinish(napi_env env, napi_callback_info info) {
  size_t argc = 0; napi_value self; void* p = nullptr;
  napi_get_cb_info(env, info, &argc, nullptr, &self, nullptr);
  if (napi_unwrap(env, self, &p) != napi_ok || !p) { napi_throw_error(env, nullptr, "Encoder not initialized"); return nullptr; }
  const uint8_t* out = nullptr; size_t len = 0;
  int rc = bgpu_stream_finish((bgpu_stream*)p, &out, &len);
  return rc ? throw_bgpu(env, rc) : make_u8(env, out, len);
}

napi_value ModuleInit(napi_env env, napi_value exports) {
  const napi_property_descriptor fns[] = {
      {"encode", nullptr, Encode, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"encodeBatch", nullptr, EncodeBatch, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"encodePart", nullptr, EncodePart, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"init", nullptr, Init, nullptr, nullptr, nullptr, napi_default, nullptr},
  };
  napi_define_properties(env, exports, sizeof fns / sizeof fns[0], fns);
  const napi_property_descriptor methods[] = {
      {"update", nullptr, EncoderUpdate, nullptr, nullptr, nullptr, napi_default, nullptr},
      {"finish", nullptr, EncoderFinish, nullptr, nullptr, nullptr, napi_default, nullptr},
  };
  napi_value cls;
  napi_define_class(env, "Encoder", NAPI_AUTO_LENGTH, EncoderNew, nullptr, sizeof methods / sizeof methods[0], methods, &cls);
  napi_set_named_property(env, exports, "Encoder", cls);
  return exports;
}

}  // namespace

NAPI_MODULE(NODE_GYP_MODULE_NAME, ModuleInit)

#endif  // BGPU_HAVE_NODE_API
