/* Declaration-only stub of the Node-API subset node/addon.cc uses, so that its syntax and types can be checked in an
 * image without Node (tests/test_hostdev_units.py: g++ -fsyntax-only -I tests/stubs).  Not the real header: no
 * definitions, nothing links against it. */
#ifndef BGPU_TEST_NODE_API_STUB_H_
#define BGPU_TEST_NODE_API_STUB_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct napi_env__* napi_env;
typedef struct napi_value__* napi_value;
typedef struct napi_ref__* napi_ref;
typedef struct napi_callback_info__* napi_callback_info;
typedef enum { napi_ok, napi_invalid_arg, napi_generic_failure } napi_status;
typedef enum { napi_int8_array, napi_uint8_array, napi_uint8_clamped_array, napi_int16_array, napi_uint16_array, napi_int32_array,
               napi_uint32_array, napi_float32_array, napi_float64_array, napi_bigint64_array, napi_biguint64_array } napi_typedarray_type;
typedef enum { napi_default = 0 } napi_property_attributes;
typedef napi_value (*napi_callback)(napi_env env, napi_callback_info info);
typedef void (*napi_finalize)(napi_env env, void* finalize_data, void* finalize_hint);
typedef struct {
  const char* utf8name; napi_value name; napi_callback method; napi_callback getter; napi_callback setter; napi_value value;
  napi_property_attributes attributes; void* data;
} napi_property_descriptor;
#define NAPI_AUTO_LENGTH SIZE_MAX
#define NODE_GYP_MODULE_NAME brotligpu
#define NAPI_MODULE(name, init) extern "C" napi_value napi_register_module_v1(napi_env env, napi_value exports) { return init(env, exports); }
napi_status napi_get_cb_info(napi_env, napi_callback_info, size_t* argc, napi_value* argv, napi_value* this_arg, void** data);
napi_status napi_is_typedarray(napi_env, napi_value, bool* result);
napi_status napi_get_typedarray_info(napi_env, napi_value, napi_typedarray_type*, size_t* length, void** data, napi_value* arraybuffer, size_t* byte_offset);
napi_status napi_get_value_int32(napi_env, napi_value, int32_t*);
napi_status napi_get_value_bool(napi_env, napi_value, bool*);
napi_status napi_create_arraybuffer(napi_env, size_t, void** data, napi_value* result);
napi_status napi_create_typedarray(napi_env, napi_typedarray_type, size_t length, napi_value arraybuffer, size_t byte_offset, napi_value* result);
napi_status napi_get_array_length(napi_env, napi_value, uint32_t*);
napi_status napi_get_element(napi_env, napi_value, uint32_t, napi_value*);
napi_status napi_set_element(napi_env, napi_value, uint32_t, napi_value);
napi_status napi_create_array_with_length(napi_env, size_t, napi_value*);
napi_status napi_get_undefined(napi_env, napi_value*);
napi_status napi_throw_error(napi_env, const char* code, const char* msg);
napi_status napi_throw_type_error(napi_env, const char* code, const char* msg);
napi_status napi_throw_range_error(napi_env, const char* code, const char* msg);
napi_status napi_wrap(napi_env, napi_value js_object, void* native_object, napi_finalize, void* finalize_hint, napi_ref* result);
napi_status napi_unwrap(napi_env, napi_value js_object, void** result);
napi_status napi_define_properties(napi_env, napi_value object, size_t count, const napi_property_descriptor*);
napi_status napi_define_class(napi_env, const char* utf8name, size_t length, napi_callback constructor, void* data, size_t property_count,
                              const napi_property_descriptor* properties, napi_value* result);
napi_status napi_set_named_property(napi_env, napi_value object, const char* utf8name, napi_value value);
#ifdef __cplusplus
}
#endif
#endif
