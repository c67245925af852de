/*
 * node_api_min.h -- the handful of N-API declarations hlsmap_napi.c uses, so the addon can
 * be syntax-checked in an image without Node.  N-API is a stable C ABI; where the real
 * <node_api.h> exists it is used instead (see the __has_include in hlsmap_napi.c).
 */
#ifndef HLSMAP_NODE_API_MIN_H
#define HLSMAP_NODE_API_MIN_H
#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

typedef struct napi_env__ *napi_env;
typedef struct napi_value__ *napi_value;
typedef struct napi_callback_info__ *napi_callback_info;
typedef enum { napi_ok = 0 } napi_status;
typedef enum {
    napi_undefined, napi_null, napi_boolean, napi_number, napi_string, napi_symbol, napi_object,
    napi_function, napi_external, napi_bigint
} napi_valuetype;
typedef enum {
    napi_int8_array, napi_uint8_array, napi_uint8_clamped_array, napi_int16_array, napi_uint16_array,
    napi_int32_array, napi_uint32_array, napi_float32_array, napi_float64_array, napi_bigint64_array,
    napi_biguint64_array
} napi_typedarray_type;
typedef napi_value (*napi_callback)(napi_env env, napi_callback_info info);
typedef void (*napi_finalize)(napi_env env, void *data, void *hint);
#define NAPI_AUTO_LENGTH ((size_t)-1)

napi_status napi_get_cb_info(napi_env, napi_callback_info, size_t *argc, napi_value *argv, napi_value *this_arg, void **data);
napi_status napi_typeof(napi_env, napi_value, napi_valuetype *);
napi_status napi_is_typedarray(napi_env, napi_value, bool *);
napi_status napi_get_typedarray_info(napi_env, napi_value, napi_typedarray_type *, size_t *length, void **data, napi_value *arraybuffer, size_t *byte_offset);
napi_status napi_create_arraybuffer(napi_env, size_t byte_length, void **data, napi_value *result);
napi_status napi_create_typedarray(napi_env, napi_typedarray_type, size_t length, napi_value arraybuffer, size_t byte_offset, napi_value *result);
napi_status napi_create_object(napi_env, napi_value *);
napi_status napi_create_double(napi_env, double, napi_value *);
napi_status napi_create_external(napi_env, void *data, napi_finalize, void *hint, napi_value *);
napi_status napi_get_value_external(napi_env, napi_value, void **);
napi_status napi_get_value_int32(napi_env, napi_value, int32_t *);
napi_status napi_get_value_int64(napi_env, napi_value, int64_t *);
napi_status napi_create_external_arraybuffer(napi_env, void *external_data, size_t byte_length, napi_finalize, void *hint, napi_value *result);
napi_status napi_get_value_bool(napi_env, napi_value, bool *);
napi_status napi_get_named_property(napi_env, napi_value object, const char *name, napi_value *result);
napi_status napi_set_named_property(napi_env, napi_value object, const char *name, napi_value value);
napi_status napi_create_function(napi_env, const char *name, size_t length, napi_callback, void *data, napi_value *result);
napi_status napi_throw_error(napi_env, const char *code, const char *msg);
napi_status napi_throw_type_error(napi_env, const char *code, const char *msg);

typedef napi_value (*napi_addon_register_func)(napi_env env, napi_value exports);
#define NODE_GYP_MODULE_NAME hlsmap_napi
#define NAPI_MODULE(name, init) \
    napi_value napi_register_module_v1(napi_env env, napi_value exports) { return init(env, exports); }
#endif
