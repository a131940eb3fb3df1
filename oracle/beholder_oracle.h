/*
 * beholder_oracle.h — CPU oracle for the beholder progress-event path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under beholder_b200/ may include, link or
 * call this.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs use it, as the checker or the timed CPU baseline.
 *
 * PARITY UNPINNED: the reference (tritonmedia/beholder) has no tests, golden
 * vectors or fixtures, cannot run here (no node, triton-core not vendored), and
 * the arithmetic of its hot path lives in un-vendored third-party modules:
 *   triton-core 1.27.1  (yarn.lock:1991-2008)  -> proto.decode / enumToString
 *   protobufjs  6.8.8   (yarn.lock:1506-1518)  -> Reader / BufferReader / Type.decode
 *   prom-client 11.5.3  (yarn.lock:1494-1499)  -> Counter.inc
 * This file restates (a) the handler's control flow from index.js:127-155 and
 * (b) protobufjs 6.8.8's published reader algorithm, from its public sources as
 * recalled; the .proto (field numbers, enum list) ships in triton-core and is an
 * ASSUMPTION here, passed in as a runtime schema.
 *
 * Part 1 follows the reference.  Part 2 ("extension") has no reference
 * counterpart: windowed count/sum/min/max/log2-histogram
 * keyed by message fields, defined by this build.
 */
#ifndef BEHOLDER_ORACLE_H
#define BEHOLDER_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* declared protobuf scalar type of a known field (what protobufjs' generated
 * decoder would call for it: r.string(), r.bytes(), r.int32(), ...) */
enum {
  BHO_K_STRING = 1, BHO_K_BYTES = 2, BHO_K_INT32 = 3, BHO_K_UINT32 = 4,
  BHO_K_SINT32 = 5, BHO_K_BOOL = 6, BHO_K_FLOAT = 7, BHO_K_DOUBLE = 8
};
/* role bits */
enum { BHO_R_KEY = 1, BHO_R_STATUS = 2, BHO_R_VALUE = 4 };
/* ops bits (extension) */
enum { BHO_OP_COUNT = 1, BHO_OP_SUM = 2, BHO_OP_MIN = 4, BHO_OP_MAX = 8, BHO_OP_HIST = 16 };
/* per-event outcome */
enum {
  BHO_EV_OK = 0,
  BHO_EV_DROP_DECODE = 1,   /* proto.decode threw          (index.js:129 -> :149) */
  BHO_EV_DROP_ENUM = 2,     /* enumToString gave undefined (index.js:134,137 -> :149) */
  BHO_EV_DROP_NAN = 3       /* extension only: value is NaN */
};

#define BHO_MAX_FIELDS 16
#define BHO_MAX_ENUM 64
#define BHO_MAX_KEY_PARTS 3
#define BHO_HIST_MAX 64

typedef struct {
  uint32_t number, kind, role, reserved;
} bho_field;

typedef struct {
  uint32_t n_fields;
  bho_field fields[BHO_MAX_FIELDS];
  uint32_t n_enum;
  int32_t enum_number[BHO_MAX_ENUM];
  char enum_name[BHO_MAX_ENUM][32];   /* as in the .proto, e.g. "DEPLOYED" */
} bho_schema;

/* decoded message: what index.js:131 destructures, plus byte spans for keys */
typedef struct {
  int32_t outcome;                     /* BHO_EV_* (decode stage: OK or DROP_DECODE) */
  int32_t status;                      /* proto3 default 0 */
  double value;                        /* VALUE-role field as a JS number; default 0 */
  uint32_t n_parts;                    /* key parts, in schema order */
  uint32_t part_is_len[BHO_MAX_KEY_PARTS];
  uint32_t part_field[BHO_MAX_KEY_PARTS];
  uint32_t part_pos[BHO_MAX_KEY_PARTS];   /* LEN part: byte span in the payload */
  uint32_t part_len[BHO_MAX_KEY_PARTS];
  uint32_t part_val[BHO_MAX_KEY_PARTS];   /* varint part: 32-bit value */
} bho_msg;

/* ---- Part 1: the reference path ---- */

/* proto.decode(telemetryProgressProto, rmsg.message.content) — index.js:129 */
void bho_decode(const bho_schema* s, const uint8_t* buf, uint32_t len, bho_msg* out);

typedef struct bho_handler bho_handler;
/* log_mode: 0 = logger silenced, 1 = format+write the pino line of index.js:133
 * to /dev/null (cost model only) */
bho_handler* bho_handler_new(const bho_schema* s, int log_mode);
void bho_handler_free(bho_handler* h);
/* amqp.listen('v1.telemetry.progress', async rmsg => ...) body, index.js:127-155,
 * with db/trello stubbed.  Returns the BHO_EV_* outcome; the message is acked
 * in every case (index.js:151,154) — *acks is incremented. */
int bho_handle_progress(bho_handler* h, const uint8_t* buf, uint32_t len);
/* beholder_progress_updates_total{status=<lower-cased enum name>} — index.js:30-34 */
double bho_handler_count(const bho_handler* h, const char* status_lower);
/* by enum slot i (schema order) */
double bho_handler_count_at(const bho_handler* h, uint32_t i);
uint64_t bho_handler_acks(const bho_handler* h);
uint64_t bho_handler_dropped(const bho_handler* h, int outcome);
/* render the /metrics text for the counter (prom-client 11.5.3 text format) */
size_t bho_handler_metrics_text(const bho_handler* h, char* out, size_t cap);

/* ---- Part 2: extension (no reference counterpart) ---- */

typedef struct {
  uint32_t ops;
  uint32_t hist_buckets;      /* <= 64 */
  int32_t hist_e0;
  uint32_t reference_key;     /* 1: key = status number itself (identity), no hashing */
  uint64_t window_ms;         /* 0 = one cumulative window (id 0) */
  uint64_t hash_seed;
} bho_agg_config;

typedef struct {
  uint64_t key_hash, window_id, count;
  double sum, min, max;
} bho_record;

typedef struct bho_agg bho_agg;
bho_agg* bho_agg_new(const bho_schema* s, const bho_agg_config* c);
void bho_agg_free(bho_agg* a);
/* raw events: payload slab + n+1 offsets + n timestamps */
void bho_agg_push_raw(bho_agg* a, const uint8_t* slab, const uint32_t* offsets,
                      const uint64_t* ts_ms, uint64_t n);
/* pre-decoded columns */
void bho_agg_push_columns(bho_agg* a, const uint64_t* key_hash, const uint64_t* ts_ms,
                          const double* value, uint64_t n);
/* merge another aggregator's state into a (count/sum/hist add, min/max) */
void bho_agg_merge(bho_agg* a, const bho_agg* b);
uint64_t bho_agg_size(const bho_agg* a);
/* records sorted by (window_id, key_hash); hist_out may be NULL, else n*hist_buckets */
uint64_t bho_agg_snapshot(const bho_agg* a, bho_record* out, uint32_t* hist_out, uint64_t cap);
uint64_t bho_agg_dropped(const bho_agg* a, int outcome);
/* number of distinct full keys that share a (key_hash, window) with another key */
uint64_t bho_agg_hash_collisions(const bho_agg* a);

/* key hash of a decoded message (extension §B.2 key definition) */
uint64_t bho_key_hash(const bho_msg* m, const uint8_t* buf, uint64_t seed, int reference_key);
uint64_t bho_hash_bytes(const uint8_t* p, uint32_t len, uint32_t field, uint64_t seed);
uint32_t bho_hist_bucket(double v, uint32_t buckets, int32_t e0);
uint64_t bho_f64_order(double v);

/* decode-only columnar output for checking the GPU decode kernel */
void bho_decode_columns(const bho_schema* s, const bho_agg_config* c, const uint8_t* slab,
                        const uint32_t* offsets, uint64_t n, uint64_t* key_hash,
                        double* value, int32_t* status, uint8_t* outcome);

/* threaded baseline: shard events over n_threads private aggregators, then merge.
 * Returns a new aggregator. */
bho_agg* bho_agg_run_threads(const bho_schema* s, const bho_agg_config* c,
                             const uint8_t* slab, const uint32_t* offsets,
                             const uint64_t* ts_ms, uint64_t n, int n_threads);
/* threaded reference handler: per-thread handlers summed into counts_out[n_enum] */
void bho_handler_run_threads(const bho_schema* s, int log_mode, const uint8_t* slab,
                             const uint32_t* offsets, uint64_t n, int n_threads,
                             double* counts_out, uint64_t* dropped_out /*[4]*/);

#ifdef __cplusplus
}
#endif
#endif
