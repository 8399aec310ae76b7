/*
 * crt_addon.c -- N-API addon over the libcrt C ABI (include/crt.h).
 *
 * The thin Node.js face of the hot path: every export is one C-ABI call, ArrayBuffers in and out, no
 * arithmetic.  Uses only the stable node_api.h C API.  SOURCE ONLY in this repository: the build image has
 * no node and no node_api.h, so this file is never compiled or run here -- the tested boundary is the C ABI
 * underneath it (tests/ drive it through ctypes).  Build on a machine with Node >= 16:
 *
 *     cd js && npx node-gyp rebuild        # binding.gyp links ../computeraytracer_b200/libcrt.so
 *
 * JS surface (mirrors what src/main.js does against WebGPU):
 *     create(device) -> handle                 requestAdapter/requestDevice + pipeline  (main.js:8-9, 535-543)
 *     setBinding(handle, n, ArrayBuffer)       createBuffer/unmap + bind group entry    (main.js:139-393, 465-529)
 *     setParams(handle, {maxDepth, ...})       the shader's hard-coded consts           (ComputeShader.wgsl:11-20)
 *     reset(handle)                            zero accumulator, sample = 0             (main.js:298-311)
 *     dispatchFrames(handle, n)                n x frame()                              (main.js:597-611)
 *     readBinding(handle, n) -> ArrayBuffer    readback of bindings 0/1/2 (new)
 *     destroy(handle)                          frees the device memory now (pool, staging, scene), not at GC time
 *   multi-GPU (new; the reference is single-device):
 *     createMulti([devices], mode) -> handle   ONE handle over n GPUs of this process (crt_create_multi): every call above
 *                                              works on it, readBinding combines on the first device (fused peer kernel)
 *     commUniqueId() -> ArrayBuffer(128)       one process per GPU: rank 0 makes the id, the host carries it to the others
 *     commInit(handle, id, rank, world, mode)  ncclCommInitRank inside libcrt
 *     combine(handle, dst)                     ncclReduce of the partial sums / gather of the owned tiles to rank dst
 */
#include <node_api.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/crt.h"

#define NAPI_OK(call)                                             \
  do {                                                            \
    if ((call) != napi_ok) {                                      \
      napi_throw_error(env, NULL, "N-API call failed: " #call);   \
      return NULL;                                                \
    }                                                             \
  } while (0)

static napi_value throw_crt(napi_env env, crt_ctx* ctx, int rc) {
  char code[16];
  snprintf(code, sizeof code, "CRT_%d", rc);
  napi_throw_error(env, code, ctx ? crt_last_error(ctx) : "libcrt: no usable B200 (there is no CPU fallback)");
  return NULL;
}

/* The external wraps a small box, not the context itself, so that destroy() can free the context at once (12+ GB of pool
 * and staging at 1080p) and the finalizer -- whenever V8 gets to it -- finds nothing left to free twice. */
typedef struct { crt_ctx* ctx; } handle_box;

static void finalize_ctx(napi_env env, void* data, void* hint) {
  (void)env; (void)hint;
  handle_box* b = (handle_box*)data;
  if (b->ctx) crt_destroy(b->ctx);
  free(b);
}

static handle_box* unwrap_box(napi_env env, napi_value v) {
  void* p = NULL;
  if (napi_get_value_external(env, v, &p) != napi_ok || !p) {
    napi_throw_type_error(env, NULL, "expected a libcrt handle");
    return NULL;
  }
  return (handle_box*)p;
}

static crt_ctx* unwrap(napi_env env, napi_value v) {
  handle_box* b = unwrap_box(env, v);
  if (!b) return NULL;
  if (!b->ctx) { napi_throw_error(env, NULL, "libcrt handle has been destroyed"); return NULL; }
  return b->ctx;
}

static napi_value wrap_ctx(napi_env env, crt_ctx* ctx) {
  handle_box* b = (handle_box*)malloc(sizeof *b);
  if (!b) { crt_destroy(ctx); napi_throw_error(env, NULL, "out of memory"); return NULL; }
  b->ctx = ctx;
  napi_value out;
  if (napi_create_external(env, b, finalize_ctx, NULL, &out) != napi_ok) { crt_destroy(ctx); free(b); napi_throw_error(env, NULL, "napi_create_external failed"); return NULL; }
  return out;
}

static napi_value js_create(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  int32_t device = 0;
  if (argc >= 1) NAPI_OK(napi_get_value_int32(env, argv[0], &device));
  crt_ctx* ctx = NULL;
  int rc = crt_create(device, &ctx);
  if (rc) return throw_crt(env, NULL, rc);
  return wrap_ctx(env, ctx);
}

static napi_value js_create_multi(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  uint32_t n = 0, mode = CRT_SHARD_SPLIT_SPP;
  NAPI_OK(napi_get_array_length(env, argv[0], &n));
  if (argc >= 2) NAPI_OK(napi_get_value_uint32(env, argv[1], &mode));
  int devices[16];
  if (n < 1 || n > 16) { napi_throw_range_error(env, NULL, "1 .. 16 devices"); return NULL; }
  for (uint32_t i = 0; i < n; i++) {
    napi_value v; int32_t d = 0;
    NAPI_OK(napi_get_element(env, argv[0], i, &v));
    NAPI_OK(napi_get_value_int32(env, v, &d));
    devices[i] = d;
  }
  crt_ctx* ctx = NULL;
  int rc = crt_create_multi(devices, (int)n, mode, &ctx);
  if (rc) return throw_crt(env, NULL, rc);
  return wrap_ctx(env, ctx);
}

static napi_value js_comm_unique_id(napi_env env, napi_callback_info info) {
  (void)info;
  void* data = NULL; napi_value ab;
  NAPI_OK(napi_create_arraybuffer(env, CRT_COMM_ID_BYTES, &data, &ab));
  int rc = crt_comm_unique_id(data, CRT_COMM_ID_BYTES);
  if (rc) { napi_throw_error(env, "CRT_8", "crt_comm_unique_id: libnccl.so.2 could not be loaded or failed"); return NULL; }
  return ab;
}

static napi_value js_comm_init(napi_env env, napi_callback_info info) {
  size_t argc = 5; napi_value argv[5];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  void* id = NULL; size_t len = 0;
  NAPI_OK(napi_get_arraybuffer_info(env, argv[1], &id, &len));
  uint32_t rank = 0, world = 1, mode = CRT_SHARD_SPLIT_SPP;
  NAPI_OK(napi_get_value_uint32(env, argv[2], &rank));
  NAPI_OK(napi_get_value_uint32(env, argv[3], &world));
  NAPI_OK(napi_get_value_uint32(env, argv[4], &mode));
  int rc = crt_comm_init(ctx, id, len, rank, world, mode);
  if (rc) return throw_crt(env, ctx, rc);
  return NULL;
}

static napi_value js_combine(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  uint32_t dst = 0;
  if (argc >= 2) NAPI_OK(napi_get_value_uint32(env, argv[1], &dst));
  int rc = crt_combine(ctx, dst);   /* collective; asynchronous on the context's stream */
  if (rc) return throw_crt(env, ctx, rc);
  return NULL;
}

static napi_value js_set_binding(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  uint32_t binding = 0;
  NAPI_OK(napi_get_value_uint32(env, argv[1], &binding));
  void* data = NULL; size_t len = 0;
  bool is_ab = false;
  NAPI_OK(napi_is_arraybuffer(env, argv[2], &is_ab));
  if (is_ab) {
    NAPI_OK(napi_get_arraybuffer_info(env, argv[2], &data, &len));
  } else {   /* a typed array view */
    napi_typedarray_type type; size_t n; napi_value ab; size_t off;
    NAPI_OK(napi_get_typedarray_info(env, argv[2], &type, &n, &data, &ab, &off));
    static const size_t width[] = {1, 1, 1, 2, 2, 4, 4, 4, 8, 8, 8};
    len = n * width[type];
  }
  int rc = crt_set_binding(ctx, binding, data, len);
  if (rc) return throw_crt(env, ctx, rc);
  return NULL;
}

static int get_u32_prop(napi_env env, napi_value obj, const char* name, uint32_t* out) {
  bool has = false;
  if (napi_has_named_property(env, obj, name, &has) != napi_ok || !has) return 0;
  napi_value v;
  if (napi_get_named_property(env, obj, name, &v) != napi_ok) return 0;
  return napi_get_value_uint32(env, v, out) == napi_ok;
}

static int get_f32_prop(napi_env env, napi_value obj, const char* name, float* out) {
  bool has = false;
  if (napi_has_named_property(env, obj, name, &has) != napi_ok || !has) return 0;
  napi_value v; double d;
  if (napi_get_named_property(env, obj, name, &v) != napi_ok || napi_get_value_double(env, v, &d) != napi_ok) return 0;
  *out = (float)d;
  return 1;
}

static napi_value js_set_params(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  crt_params p;
  int rc = crt_get_params(ctx, &p);
  if (rc) return throw_crt(env, ctx, rc);
  get_u32_prop(env, argv[1], "maxDepth", &p.max_depth);
  get_u32_prop(env, argv[1], "gridSize", &p.grid_size);
  get_f32_prop(env, argv[1], "tMin", &p.t_min);
  get_f32_prop(env, argv[1], "exposure", &p.exposure);
  get_u32_prop(env, argv[1], "kernel", &p.kernel);
  get_u32_prop(env, argv[1], "framesPerLaunch", &p.frames_per_launch);
  get_u32_prop(env, argv[1], "wfSlots", &p.wf_slots);
  get_u32_prop(env, argv[1], "wfTuning", &p.wf_tuning);     /* CRT_WF_* bits, include/crt.h */
  get_u32_prop(env, argv[1], "bvhTuning", &p.bvh_tuning);   /* bins | CRT_BVH_EXACT_SLOTS | CRT_BVH_BUILDER_* */
  get_f32_prop(env, argv[1], "bvhPrimCost", &p.bvh_prim_cost);
  rc = crt_set_params(ctx, &p);
  if (rc) return throw_crt(env, ctx, rc);
  return NULL;
}

static napi_value js_reset(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  int rc = crt_reset(ctx);
  if (rc) return throw_crt(env, ctx, rc);
  return NULL;
}

static napi_value js_set_shard(napi_env env, napi_callback_info info) {
  size_t argc = 4; napi_value argv[4];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  uint32_t rank = 0, world = 1, mode = 0;
  NAPI_OK(napi_get_value_uint32(env, argv[1], &rank));
  NAPI_OK(napi_get_value_uint32(env, argv[2], &world));
  NAPI_OK(napi_get_value_uint32(env, argv[3], &mode));
  int rc = crt_set_shard(ctx, rank, world, mode);
  if (rc) return throw_crt(env, ctx, rc);
  return NULL;
}

static napi_value js_dispatch_frames(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  uint32_t n = 1;
  if (argc >= 2) NAPI_OK(napi_get_value_uint32(env, argv[1], &n));
  int rc = crt_dispatch_frames(ctx, n);   /* asynchronous on the context's stream, like queue.submit */
  if (rc) return throw_crt(env, ctx, rc);
  return NULL;
}

static napi_value js_read_binding(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  uint32_t binding = 0;
  NAPI_OK(napi_get_value_uint32(env, argv[1], &binding));
  size_t n = 4;
  if (binding <= 1) {
    void* dptr = NULL;
    int rc = crt_device_ptr(ctx, binding, &dptr, &n);   /* size of the binding for the current camera */
    if (rc) return throw_crt(env, ctx, rc);
  }
  void* data = NULL; napi_value ab;
  NAPI_OK(napi_create_arraybuffer(env, n, &data, &ab));
  int rc = crt_read_binding(ctx, binding, data, n);      /* synchronises */
  if (rc) return throw_crt(env, ctx, rc);
  return ab;
}

static napi_value js_counters(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  crt_ctx* ctx = unwrap(env, argv[0]);
  if (!ctx) return NULL;
  uint64_t c[CRT_NUM_COUNTERS];
  int rc = crt_get_counters(ctx, c, CRT_NUM_COUNTERS);
  if (rc) return throw_crt(env, ctx, rc);
  napi_value arr;
  NAPI_OK(napi_create_array_with_length(env, CRT_NUM_COUNTERS, &arr));
  for (uint32_t i = 0; i < CRT_NUM_COUNTERS; i++) {
    napi_value v;
    NAPI_OK(napi_create_double(env, (double)c[i], &v));
    NAPI_OK(napi_set_element(env, arr, i, v));
  }
  return arr;
}

static napi_value js_destroy(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  handle_box* b = unwrap_box(env, argv[0]);
  if (b && b->ctx) { crt_destroy(b->ctx); b->ctx = NULL; }   /* synchronises, frees everything now; the finalizer frees the box */
  return NULL;
}

static napi_value init(napi_env env, napi_value exports) {
  static const struct { const char* name; napi_callback fn; } fns[] = {
      {"create", js_create},           {"setBinding", js_set_binding},       {"setParams", js_set_params},
      {"reset", js_reset},             {"setShard", js_set_shard},           {"dispatchFrames", js_dispatch_frames},
      {"readBinding", js_read_binding}, {"counters", js_counters},           {"destroy", js_destroy},
      {"createMulti", js_create_multi}, {"commUniqueId", js_comm_unique_id}, {"commInit", js_comm_init},
      {"combine", js_combine},
  };
  for (size_t i = 0; i < sizeof fns / sizeof fns[0]; i++) {
    napi_value f;
    if (napi_create_function(env, fns[i].name, NAPI_AUTO_LENGTH, fns[i].fn, NULL, &f) != napi_ok) return NULL;
    if (napi_set_named_property(env, exports, fns[i].name, f) != napi_ok) return NULL;
  }
  return exports;
}

NAPI_MODULE(NODE_GYP_MODULE_NAME, init)
