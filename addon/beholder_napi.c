/*
 * beholder_napi.c — N-API shim over libbeholder_gpu.so (include/beholder_gpu.h).
 *
 * SOURCE ONLY: node, node-gyp and node_api.h are not present in the image this repository is built
 * and tested in, so this file has never been compiled there.  It uses only the stable N-API C ABI
 * and the C-ABI of the library, which IS tested (Python ctypes + GPU tests).  Build on a box with
 * Node >= 12:   node-gyp rebuild   (binding.gyp next to this file).
 *
 * JS surface (wrapped by beholder-gpu.js):
 *   const ctx = native.create(configBuffer)        // bh_create; configBuffer = a packed bh_config
 *   native.push(ctx, buffer, tsMs)                 // bh_push: rmsg.message.content (index.js:129)
 *   native.flush(ctx)      -> committedSeq         // bh_flush
 *   native.poll(ctx)       -> committedSeq         // bh_poll_committed (non-blocking; submits a batch past its deadline)
 *   native.submit(ctx)                             // bh_submit: hand the partial batch to the GPU now
 *   native.onCommit(ctx, fn)                       // bh_set_commit_callback: fn(committedSeq) on the JS thread
 *   native.snapshot(ctx, windowId, cap) -> Buffer  // bh_snapshot: packed bh_record rows
 *   native.counters(ctx)   -> Buffer               // bh_get_counters
 *   native.destroy(ctx)
 *   const g = native.createGroup(localCfg, ownedCfg, [dev0, dev1, ...], ownerStride)   // bh_create_group
 *   native.groupLocal(g, r) / native.groupOwned(g, r) -> ctx      // contexts owned by the group: never destroy them
 *   native.groupCloseWindow(g, windowId)           // bh_group_close_window (0xFFFFFFFFFFFFFFFFn = every open window)
 *   native.groupSync(g)                            // bh_group_sync
 */
#include <node_api.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/beholder_gpu.h"

#define NAPI_OK(call) do { if ((call) != napi_ok) { napi_throw_error(env, NULL, #call " failed"); return NULL; } } while (0)

static bh_ctx* unwrap(napi_env env, napi_value v) {
  void* p = NULL;
  if (napi_get_value_external(env, v, &p) != napi_ok) return NULL;
  return (bh_ctx*)p;
}

static void finalize_ctx(napi_env env, void* data, void* hint) { (void)env; (void)hint; bh_destroy((bh_ctx*)data); }

static napi_value throw_bh(napi_env env, bh_ctx* ctx, int rc) {
  char code[16]; snprintf(code, sizeof code, "%d", rc);
  napi_throw_error(env, code, bh_last_error(ctx));
  return NULL;
}

static napi_value Create(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1]; void* data; size_t len;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  NAPI_OK(napi_get_buffer_info(env, argv[0], &data, &len));
  if (len != sizeof(bh_config)) { napi_throw_error(env, NULL, "config buffer must be sizeof(bh_config) bytes"); return NULL; }
  bh_config cfg; memcpy(&cfg, data, sizeof cfg);
  bh_ctx* ctx = NULL;
  int rc = bh_create(&cfg, &ctx);
  if (rc != BH_OK) return throw_bh(env, NULL, rc);       /* no device -> BH_ERR_NO_DEVICE: there is no CPU path */
  napi_value ext;
  NAPI_OK(napi_create_external(env, ctx, finalize_ctx, NULL, &ext));
  return ext;
}

static napi_value Push(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3]; void* data; size_t len; int64_t ts = 0;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  bh_ctx* ctx = unwrap(env, argv[0]);
  NAPI_OK(napi_get_buffer_info(env, argv[1], &data, &len));
  if (argc > 2) napi_get_value_int64(env, argv[2], &ts);
  int rc = bh_push(ctx, (const uint8_t*)data, (uint32_t)len, (uint64_t)ts);   /* copied into the pinned ring */
  if (rc != BH_OK) return throw_bh(env, ctx, rc);
  return NULL;
}

static napi_value seq_result(napi_env env, uint64_t seq) {
  napi_value v; NAPI_OK(napi_create_bigint_uint64(env, seq, &v)); return v;
}

static napi_value Flush(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1]; uint64_t seq = 0;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  bh_ctx* ctx = unwrap(env, argv[0]);
  int rc = bh_flush(ctx, &seq);
  if (rc != BH_OK) return throw_bh(env, ctx, rc);
  return seq_result(env, seq);
}

static napi_value Poll(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1]; uint64_t seq = 0;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  bh_ctx* ctx = unwrap(env, argv[0]);
  int rc = bh_poll_committed(ctx, &seq);
  if (rc != BH_OK) return throw_bh(env, ctx, rc);
  return seq_result(env, seq);
}

static napi_value Submit(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  bh_ctx* ctx = unwrap(env, argv[0]);
  int rc = bh_submit(ctx);
  if (rc != BH_OK) return throw_bh(env, ctx, rc);
  return NULL;
}

/* ---- completion callback: library thread -> JS thread ----
 * bh_set_commit_callback calls on_commit_native on the library's completion thread; a threadsafe function
 * carries the sequence number to the JS thread, where fn(committedSeq) sends the acks (index.js:151,154).
 * The threadsafe function is unref'ed so that it does not keep the process alive by itself. */
static void call_js_commit(napi_env env, napi_value js_cb, void* context, void* data) {
  (void)context;
  uint64_t seq = (uint64_t)(uintptr_t)data;
  if (env == NULL || js_cb == NULL) return;
  napi_value undefined, arg;
  if (napi_get_undefined(env, &undefined) != napi_ok) return;
  if (napi_create_bigint_uint64(env, seq, &arg) != napi_ok) return;
  napi_call_function(env, undefined, js_cb, 1, &arg, NULL);
}
static void on_commit_native(void* user, uint64_t committed_seq) {
  napi_call_threadsafe_function((napi_threadsafe_function)user, (void*)(uintptr_t)committed_seq, napi_tsfn_nonblocking);
}
static napi_value OnCommit(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2], name; napi_threadsafe_function tsfn;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  bh_ctx* ctx = unwrap(env, argv[0]);
  NAPI_OK(napi_create_string_utf8(env, "beholder-gpu commit", NAPI_AUTO_LENGTH, &name));
  NAPI_OK(napi_create_threadsafe_function(env, argv[1], NULL, name, 0, 1, NULL, NULL, NULL, call_js_commit, &tsfn));
  NAPI_OK(napi_unref_threadsafe_function(env, tsfn));
  int rc = bh_set_commit_callback(ctx, on_commit_native, tsfn);
  if (rc != BH_OK) return throw_bh(env, ctx, rc);
  return NULL;
}

static napi_value Snapshot(napi_env env, napi_callback_info info) {
  size_t argc = 3; napi_value argv[3]; bool lossless; uint64_t window = 0; uint32_t cap = 4096;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  bh_ctx* ctx = unwrap(env, argv[0]);
  if (argc > 1) napi_get_value_bigint_uint64(env, argv[1], &window, &lossless);
  if (argc > 2) napi_get_value_uint32(env, argv[2], &cap);
  void* out; napi_value buf; uint64_t n = 0;
  NAPI_OK(napi_create_buffer(env, (size_t)cap * sizeof(bh_record), &out, &buf));
  int rc = bh_snapshot(ctx, window, (bh_record*)out, cap, &n);
  if (rc != BH_OK) return throw_bh(env, ctx, rc);
  napi_value res, nv;
  NAPI_OK(napi_create_object(env, &res));
  NAPI_OK(napi_create_bigint_uint64(env, n, &nv));
  NAPI_OK(napi_set_named_property(env, res, "rows", nv));
  NAPI_OK(napi_set_named_property(env, res, "data", buf));
  return res;
}

static napi_value Counters(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  bh_ctx* ctx = unwrap(env, argv[0]);
  void* out; napi_value buf;
  NAPI_OK(napi_create_buffer(env, sizeof(bh_counters), &out, &buf));
  int rc = bh_get_counters(ctx, (bh_counters*)out);
  if (rc != BH_OK && rc != BH_ERR_TABLE_FULL && rc != BH_ERR_WINDOW_RING) return throw_bh(env, ctx, rc);
  return buf;
}

/* ---- AMQP 0-9-1 delivery tap (bh_amqp_*): socket bytes in, bodies pushed as one batch, tags out ----
 * amqpCreate(consumerTag) -> external;  amqpFeed(tap, ctx, chunk: Buffer, tsMs, channelNumber) ->
 *   { consumed: Number, tags: BigUint64Array }   (bodies of the deliveries completed inside `chunk`
 * are already in the ingest ring when it returns; the caller keeps chunk.slice(consumed) for the next read).
 * Delivery tags are per channel: only deliveries that arrived on `channelNumber` are pushed and reported
 * (pass -1 for every channel). */
#define TAP_SLAB (4u << 20)
#define TAP_CAP 65536u
typedef struct tap_box { bh_amqp* tap; uint8_t* slab; uint32_t* offsets; bh_amqp_delivery* del; } tap_box;

static void tap_finalize(napi_env env, void* data, void* hint) {
  tap_box* b = (tap_box*)data;
  bh_amqp_destroy(b->tap); free(b->slab); free(b->offsets); free(b->del); free(b);
}

static napi_value AmqpCreate(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1]; char tag[256] = {0}; size_t tl = 0;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  if (argc) napi_get_value_string_utf8(env, argv[0], tag, sizeof tag, &tl);
  tap_box* b = (tap_box*)calloc(1, sizeof *b);
  b->slab = (uint8_t*)malloc(TAP_SLAB); b->offsets = (uint32_t*)malloc((TAP_CAP + 1) * 4);
  b->del = (bh_amqp_delivery*)malloc(TAP_CAP * sizeof(bh_amqp_delivery));
  if (bh_amqp_create(&b->tap, tl ? tag : NULL, 0) != BH_OK) { napi_throw_error(env, NULL, "bh_amqp_create failed"); return NULL; }
  napi_value ext;
  NAPI_OK(napi_create_external(env, b, tap_finalize, NULL, &ext));
  return ext;
}

static napi_value AmqpFeed(napi_env env, napi_callback_info info) {
  size_t argc = 5; napi_value argv[5]; void* data; size_t len; int64_t ts = 0; int32_t want_channel = -1; tap_box* b;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  NAPI_OK(napi_get_value_external(env, argv[0], (void**)&b));
  bh_ctx* ctx = unwrap(env, argv[1]);
  NAPI_OK(napi_get_buffer_info(env, argv[2], &data, &len));
  NAPI_OK(napi_get_value_int64(env, argv[3], &ts));
  if (argc > 4) napi_get_value_int32(env, argv[4], &want_channel);
  uint64_t pos = 0, total = 0; size_t tcap = 1024; uint64_t* tags = (uint64_t*)malloc(tcap * 8);
  for (;;) {
    uint64_t used = 0; uint32_t n = 0;
    int rc = bh_amqp_feed(b->tap, (const uint8_t*)data + pos, len - pos, b->slab, TAP_SLAB, b->offsets, b->del, TAP_CAP, &used, &n);
    pos += used;
    const uint32_t n_raw = n;
    if (n && want_channel >= 0) {   /* keep the wanted channel's deliveries: compact slab, offsets and tags in place */
      uint32_t k = 0, w = 0;
      for (uint32_t i = 0; i < n; i++) {
        if (b->del[i].channel != (uint16_t)want_channel) continue;
        uint32_t nb = b->offsets[i + 1] - b->offsets[i];
        memmove(b->slab + w, b->slab + b->offsets[i], nb);
        b->del[k] = b->del[i]; b->offsets[k] = w; w += nb; k++;
      }
      b->offsets[k] = w; n = k;
    }
    if (n) {   /* one timestamp for the batch: the arrival time of the socket read */
      uint64_t* tsv = (uint64_t*)malloc((size_t)n * 8);
      for (uint32_t i = 0; i < n; i++) tsv[i] = (uint64_t)ts;
      int r2 = bh_push_batch(ctx, b->slab, b->offsets, tsv, n);
      free(tsv);
      if (r2 != BH_OK) { free(tags); return throw_bh(env, ctx, r2); }
      if (total + n > tcap) { while (total + n > tcap) tcap *= 2; tags = (uint64_t*)realloc(tags, tcap * 8); }
      for (uint32_t i = 0; i < n; i++) tags[total + i] = b->del[i].delivery_tag;
      total += n;
    }
    if (rc != BH_OK) { free(tags); napi_throw_error(env, NULL, bh_amqp_last_error(b->tap)); return NULL; }
    if (!n_raw || !used) break;
  }
  napi_value res, cv, ab, ta; void* abdata;
  NAPI_OK(napi_create_object(env, &res));
  NAPI_OK(napi_create_double(env, (double)pos, &cv));
  NAPI_OK(napi_create_arraybuffer(env, total * 8, &abdata, &ab));
  memcpy(abdata, tags, total * 8); free(tags);
  NAPI_OK(napi_create_typedarray(env, napi_biguint64_array, total, ab, 0, &ta));
  NAPI_OK(napi_set_named_property(env, res, "consumed", cv));
  NAPI_OK(napi_set_named_property(env, res, "tags", ta));
  return res;
}

/* ---- several GPUs from this one process (bh_create_group): index.js:23-48 stays a single process ---- */
static void finalize_group(napi_env env, void* data, void* hint) { (void)env; (void)hint; bh_group_destroy((bh_group*)data); }

static napi_value CreateGroup(napi_env env, napi_callback_info info) {
  size_t argc = 4; napi_value argv[4]; void *lc, *oc; size_t ll, ol; uint32_t n = 0; int64_t stride = 0; int devices[64];
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  NAPI_OK(napi_get_buffer_info(env, argv[0], &lc, &ll));
  NAPI_OK(napi_get_buffer_info(env, argv[1], &oc, &ol));
  if (ll != sizeof(bh_config) || ol != sizeof(bh_config)) { napi_throw_error(env, NULL, "config buffers must be sizeof(bh_config) bytes"); return NULL; }
  NAPI_OK(napi_get_array_length(env, argv[2], &n));
  if (n == 0 || n > 64) { napi_throw_error(env, NULL, "1..64 devices"); return NULL; }
  for (uint32_t i = 0; i < n; i++) { napi_value e; int32_t d; NAPI_OK(napi_get_element(env, argv[2], i, &e)); NAPI_OK(napi_get_value_int32(env, e, &d)); devices[i] = d; }
  NAPI_OK(napi_get_value_int64(env, argv[3], &stride));
  bh_config lcfg, ocfg; memcpy(&lcfg, lc, sizeof lcfg); memcpy(&ocfg, oc, sizeof ocfg);
  bh_group* g = NULL;
  int rc = bh_create_group(&lcfg, &ocfg, devices, n, (uint64_t)stride, &g);
  if (rc != BH_OK) { napi_throw_error(env, NULL, bh_group_last_error()); return NULL; }
  napi_value ext;
  NAPI_OK(napi_create_external(env, g, finalize_group, NULL, &ext));
  return ext;
}

static napi_value group_ctx(napi_env env, napi_callback_info info, int owned) {
  size_t argc = 2; napi_value argv[2]; void* g; uint32_t r = 0;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  NAPI_OK(napi_get_value_external(env, argv[0], &g));
  NAPI_OK(napi_get_value_uint32(env, argv[1], &r));
  bh_ctx* c = owned ? bh_group_owned((bh_group*)g, r) : bh_group_local((bh_group*)g, r);
  if (!c) { napi_throw_error(env, NULL, "rank out of range"); return NULL; }
  napi_value ext;
  NAPI_OK(napi_create_external(env, c, NULL, NULL, &ext));     /* no finalizer: the group owns the context */
  return ext;
}
static napi_value GroupLocal(napi_env env, napi_callback_info info) { return group_ctx(env, info, 0); }
static napi_value GroupOwned(napi_env env, napi_callback_info info) { return group_ctx(env, info, 1); }

static napi_value GroupCloseWindow(napi_env env, napi_callback_info info) {
  size_t argc = 2; napi_value argv[2]; void* g; uint64_t w = 0; bool lossless;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  NAPI_OK(napi_get_value_external(env, argv[0], &g));
  NAPI_OK(napi_get_value_bigint_uint64(env, argv[1], &w, &lossless));
  if (bh_group_close_window((bh_group*)g, w) != BH_OK) { napi_throw_error(env, NULL, bh_group_last_error()); return NULL; }
  return NULL;
}

static napi_value GroupSync(napi_env env, napi_callback_info info) {
  size_t argc = 1; napi_value argv[1]; void* g;
  NAPI_OK(napi_get_cb_info(env, info, &argc, argv, NULL, NULL));
  NAPI_OK(napi_get_value_external(env, argv[0], &g));
  if (bh_group_sync((bh_group*)g) != BH_OK) { napi_throw_error(env, NULL, bh_group_last_error()); return NULL; }
  return NULL;
}

static napi_value Init(napi_env env, napi_value exports) {
  napi_property_descriptor props[] = {
    {"create", NULL, Create, NULL, NULL, NULL, napi_default, NULL},
    {"push", NULL, Push, NULL, NULL, NULL, napi_default, NULL},
    {"flush", NULL, Flush, NULL, NULL, NULL, napi_default, NULL},
    {"poll", NULL, Poll, NULL, NULL, NULL, napi_default, NULL},
    {"submit", NULL, Submit, NULL, NULL, NULL, napi_default, NULL},
    {"onCommit", NULL, OnCommit, NULL, NULL, NULL, napi_default, NULL},
    {"snapshot", NULL, Snapshot, NULL, NULL, NULL, napi_default, NULL},
    {"counters", NULL, Counters, NULL, NULL, NULL, napi_default, NULL},
    {"createGroup", NULL, CreateGroup, NULL, NULL, NULL, napi_default, NULL},
    {"groupLocal", NULL, GroupLocal, NULL, NULL, NULL, napi_default, NULL},
    {"groupOwned", NULL, GroupOwned, NULL, NULL, NULL, napi_default, NULL},
    {"groupCloseWindow", NULL, GroupCloseWindow, NULL, NULL, NULL, napi_default, NULL},
    {"groupSync", NULL, GroupSync, NULL, NULL, NULL, napi_default, NULL},
    {"amqpCreate", NULL, AmqpCreate, NULL, NULL, NULL, napi_default, NULL},
    {"amqpFeed", NULL, AmqpFeed, NULL, NULL, NULL, napi_default, NULL},
  };
  napi_define_properties(env, exports, sizeof props / sizeof props[0], props);
  return exports;
}
NAPI_MODULE(NODE_GYP_MODULE_NAME, Init)
