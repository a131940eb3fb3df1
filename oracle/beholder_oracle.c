/*
 * beholder_oracle.c — CPU oracle (TEST INFRASTRUCTURE ONLY; see beholder_oracle.h).
 * PARITY UNPINNED by the reference (no tests/golden vectors there; protobufjs,
 * prom-client and the .proto are un-vendored third-party code — restated here).
 *
 * Citations "index.js:N" are into /root/reference/index.js.  Citations
 * "protobufjs:<file>" are into the protobufjs 6.8.8 npm package (not on disk).
 */
#define _GNU_SOURCE
#include "beholder_oracle.h"

#include <ctype.h>
#include <fcntl.h>
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <unistd.h>

/* ===================================================================== */
/* protobufjs 6.8.8 Reader / BufferReader, restated                        */
/* ===================================================================== */

typedef struct {
  const uint8_t* buf;
  uint64_t pos;   /* JS number; may run past len inside uint32() */
  uint64_t len;
  int threw;
} rd_t;

/* this.buf[this.pos] where out-of-range indexes yield `undefined`:
 * (undefined & 127) === 0 and (undefined < 128) === false, i.e. it behaves
 * like a byte with no value bits and the continuation bit set. */
static inline uint32_t rd_peek(const rd_t* r, uint64_t pos) {
  return pos < r->len ? r->buf[pos] : 0x80u;
}

/* protobufjs:src/reader.js Reader.prototype.uint32 */
static uint32_t rd_uint32(rd_t* r) {
  uint32_t value, b;
  b = rd_peek(r, r->pos++); value = (b & 127u);              if (b < 128) return value;
  b = rd_peek(r, r->pos++); value |= (b & 127u) << 7;        if (b < 128) return value;
  b = rd_peek(r, r->pos++); value |= (b & 127u) << 14;       if (b < 128) return value;
  b = rd_peek(r, r->pos++); value |= (b & 127u) << 21;       if (b < 128) return value;
  b = rd_peek(r, r->pos++); value |= (b & 15u) << 28;        if (b < 128) return value;
  /* "if ((this.pos += 5) > this.len) { this.pos = this.len; throw indexOutOfRange }" */
  r->pos += 5;
  if (r->pos > r->len) { r->pos = r->len; r->threw = 1; }
  return value;
}

/* protobufjs:src/reader.js Reader.prototype.skip */
static void rd_skip_varint(rd_t* r) {
  for (;;) {
    if (r->pos >= r->len) { r->threw = 1; return; }
    if (!(r->buf[r->pos++] & 128)) return;
  }
}
static void rd_skip_n(rd_t* r, uint64_t n) {
  if (r->pos + n > r->len) { r->threw = 1; return; }
  r->pos += n;
}

/* protobufjs:src/reader.js Reader.prototype.skipType.  The recursion on wire
 * type 3 (start group) is unrolled into a depth counter: each nested start
 * group is closed by the next wire-type-4 tag, whatever its field number. */
static void rd_skip_type(rd_t* r, uint32_t wire) {
  uint32_t depth = 0;
  for (;;) {
    switch (wire) {
      case 0: rd_skip_varint(r); break;
      case 1: rd_skip_n(r, 8); break;
      case 2: { uint32_t n = rd_uint32(r); if (!r->threw) rd_skip_n(r, n); break; }
      case 3: depth++; break;
      case 5: rd_skip_n(r, 4); break;
      case 4:
        if (depth == 0) { r->threw = 1; return; }   /* default: throw "invalid wire type" */
        depth--;                                    /* closes the innermost group loop */
        break;
      default: r->threw = 1; return;                /* 6, 7 */
    }
    if (r->threw || depth == 0) return;
    wire = rd_uint32(r) & 7u;                       /* while ((wireType = this.uint32() & 7) !== 4) */
    if (r->threw) return;
  }
}

static const bho_field* schema_find(const bho_schema* s, uint32_t number) {
  for (uint32_t i = 0; i < s->n_fields; i++)
    if (s->fields[i].number == number) return &s->fields[i];
  return NULL;
}

static int key_part_index(const bho_schema* s, const bho_field* f) {
  int k = 0;
  for (uint32_t i = 0; i < s->n_fields; i++) {
    if (!(s->fields[i].role & BHO_R_KEY)) continue;
    if (&s->fields[i] == f) return k;
    k++;
  }
  return -1;
}

/* protobufjs:src/decoder.js generated Type.decode:
 *   while (r.pos < c) { var t = r.uint32(); switch (t >>> 3) { case N: m.f = r.<type>(); break;
 *                        default: r.skipType(t & 7); break } }
 * Known fields are read by their DECLARED type whatever the tag's wire type.
 * rmsg.message.content is a Node Buffer, so the reader is BufferReader, whose
 * string() clamps to the buffer end instead of throwing
 * (protobufjs:src/reader_buffer.js). */
void bho_decode(const bho_schema* s, const uint8_t* buf, uint32_t len, bho_msg* m) {
  rd_t r = {buf, 0, len, 0};
  memset(m, 0, sizeof *m);
  /* key parts default to "" / 0 (proto3 defaults) */
  for (uint32_t i = 0; i < s->n_fields; i++) {
    const bho_field* f = &s->fields[i];
    if (!(f->role & BHO_R_KEY) || m->n_parts >= BHO_MAX_KEY_PARTS) continue;
    uint32_t k = m->n_parts++;
    m->part_is_len[k] = (f->kind == BHO_K_STRING || f->kind == BHO_K_BYTES);
    m->part_field[k] = f->number;
  }
  while (r.pos < r.len) {
    uint32_t t = rd_uint32(&r);
    if (r.threw) break;
    const bho_field* f = schema_find(s, t >> 3);
    if (!f) { rd_skip_type(&r, t & 7u); if (r.threw) break; continue; }
    uint32_t u = 0, span_pos = 0, span_len = 0;
    double num = 0.0;
    switch (f->kind) {
      case BHO_K_STRING: {           /* BufferReader.string(): utf8Slice(pos, min(pos+len, this.len)) */
        uint32_t n = rd_uint32(&r);
        if (r.threw) break;
        uint64_t end = r.pos + n; if (end > r.len) end = r.len;
        span_pos = (uint32_t)r.pos; span_len = (uint32_t)(end - r.pos); r.pos = end;
        break;
      }
      case BHO_K_BYTES: {            /* Reader.bytes(): throws when end > this.len */
        uint32_t n = rd_uint32(&r);
        if (r.threw) break;
        if (r.pos + n > r.len) { r.threw = 1; break; }
        span_pos = (uint32_t)r.pos; span_len = n; r.pos += n;
        break;
      }
      case BHO_K_INT32: u = rd_uint32(&r); num = (double)(int32_t)u; break;     /* uint32() | 0 */
      case BHO_K_UINT32: u = rd_uint32(&r); num = (double)u; break;
      case BHO_K_SINT32: {           /* value >>> 1 ^ -(value & 1) | 0 */
        uint32_t v = rd_uint32(&r); u = (v >> 1) ^ (uint32_t)(-(int32_t)(v & 1u)); num = (double)(int32_t)u; break;
      }
      case BHO_K_BOOL: u = rd_uint32(&r) != 0; num = (double)u; break;
      case BHO_K_FLOAT: {
        if (r.pos + 4 > r.len) { r.threw = 1; break; }
        float fv; memcpy(&fv, buf + r.pos, 4); r.pos += 4; num = (double)fv; u = (uint32_t)(int32_t)0; break;
      }
      case BHO_K_DOUBLE: {
        if (r.pos + 8 > r.len) { r.threw = 1; break; }
        memcpy(&num, buf + r.pos, 8); r.pos += 8; break;
      }
      default: r.threw = 1; break;
    }
    if (r.threw) break;
    if (f->role & BHO_R_STATUS) m->status = (int32_t)u;
    if (f->role & BHO_R_VALUE) m->value = num;
    if (f->role & BHO_R_KEY) {
      int k = key_part_index(s, f);
      if (k >= 0 && k < BHO_MAX_KEY_PARTS) {
        m->part_pos[k] = span_pos; m->part_len[k] = span_len; m->part_val[k] = u;
      }
    }
  }
  m->outcome = r.threw ? BHO_EV_DROP_DECODE : BHO_EV_OK;
}

/* ===================================================================== */
/* Part 1: the handler of index.js:127-155                                 */
/* ===================================================================== */

struct bho_handler {
  bho_schema schema;
  int log_mode, log_fd;
  /* prom-client Counter hashMap: one JS double per label set, first-inc order */
  double count[BHO_MAX_ENUM];
  char label[BHO_MAX_ENUM][32];        /* statusText.toLowerCase() */
  uint32_t order[BHO_MAX_ENUM]; uint32_t n_order; uint8_t seen[BHO_MAX_ENUM];
  uint64_t acks, dropped[4];
};

bho_handler* bho_handler_new(const bho_schema* s, int log_mode) {
  bho_handler* h = calloc(1, sizeof *h);
  h->schema = *s; h->log_mode = log_mode; h->log_fd = -1;
  for (uint32_t i = 0; i < s->n_enum; i++) {
    size_t j = 0;
    for (; s->enum_name[i][j] && j < 31; j++) h->label[i][j] = (char)tolower((unsigned char)s->enum_name[i][j]);
    h->label[i][j] = 0;
  }
  if (log_mode) h->log_fd = open("/dev/null", O_WRONLY);
  return h;
}
void bho_handler_free(bho_handler* h) { if (h) { if (h->log_fd >= 0) close(h->log_fd); free(h); } }

/* proto.enumToString(type, 'TelemetryStatusEntry', status): reverse lookup of the
 * enum's valuesById; undefined when the number has no name.  index.js:134 */
static int enum_slot(const bho_schema* s, int32_t number) {
  for (uint32_t i = 0; i < s->n_enum; i++) if (s->enum_number[i] == number) return (int)i;
  return -1;
}

int bho_handle_progress(bho_handler* h, const uint8_t* buf, uint32_t len) {
  bho_msg m;
  int outcome = BHO_EV_OK;
  /* try { */
  bho_decode(&h->schema, buf, len, &m);                              /* index.js:129 */
  if (m.outcome != BHO_EV_OK) { outcome = BHO_EV_DROP_DECODE; goto caught; }
  if (h->log_mode) {                                                 /* index.js:133 */
    char line[512]; struct timeval tv; gettimeofday(&tv, NULL);
    uint32_t kl = m.n_parts && m.part_is_len[0] ? m.part_len[0] : 0; if (kl > 200) kl = 200;
    int n = snprintf(line, sizeof line,
      "{\"level\":30,\"time\":%lld,\"pid\":%d,\"hostname\":\"beholder\",\"name\":\"index.js\","
      "\"msg\":\"processing progress update on media %.*s status %d percent %.17g\",\"v\":1}\n",
      (long long)tv.tv_sec * 1000 + tv.tv_usec / 1000, 1, (int)kl,
      kl ? (const char*)buf + m.part_pos[0] : "", m.status, m.value);
    if (n > 0 && h->log_fd >= 0) { ssize_t w = write(h->log_fd, line, (size_t)n); (void)w; }
  }
  {
    int slot = enum_slot(&h->schema, m.status);                      /* index.js:134 */
    if (slot < 0) { outcome = BHO_EV_DROP_ENUM; goto caught; }       /* undefined.toLowerCase() throws, :137 */
    if (!h->seen[slot]) { h->seen[slot] = 1; h->order[h->n_order++] = (uint32_t)slot; }
    h->count[slot] += 1.0;                                           /* index.js:136-138 */
  }
  /* index.js:140-148: db.getByID + Trello comment — I/O, stubbed; they run after
   * the increment so their failure cannot un-count. */
  h->acks++;                                                         /* index.js:154 */
  return outcome;
caught:                                                              /* index.js:149-152 */
  h->dropped[outcome]++;
  h->acks++;                                                         /* index.js:151 */
  return outcome;
}

double bho_handler_count(const bho_handler* h, const char* status_lower) {
  for (uint32_t i = 0; i < h->schema.n_enum; i++)
    if (!strcmp(h->label[i], status_lower)) return h->count[i];
  return 0.0;
}
double bho_handler_count_at(const bho_handler* h, uint32_t i) { return i < BHO_MAX_ENUM ? h->count[i] : 0.0; }
uint64_t bho_handler_acks(const bho_handler* h) { return h->acks; }
uint64_t bho_handler_dropped(const bho_handler* h, int outcome) {
  return (outcome >= 0 && outcome < 4) ? h->dropped[outcome] : 0;
}

/* prom-client 11.5.3 registry.metrics() for one Counter (index.js:30-34) */
size_t bho_handler_metrics_text(const bho_handler* h, char* out, size_t cap) {
  size_t n = 0;
  n += (size_t)snprintf(out + n, n < cap ? cap - n : 0,
      "# HELP beholder_progress_updates_total Total number of messages processed in this processes lifetime\n"
      "# TYPE beholder_progress_updates_total counter\n");
  for (uint32_t j = 0; j < h->n_order; j++) {
    uint32_t i = h->order[j];
    n += (size_t)snprintf(out + n, n < cap ? cap - n : 0,
        "beholder_progress_updates_total{status=\"%s\"} %.17g\n", h->label[i], h->count[i]);
  }
  return n;
}

/* ===================================================================== */
/* Part 2: extension — no reference counterpart         */
/* ===================================================================== */

static inline uint64_t fmix64(uint64_t h) {
  h ^= h >> 33; h *= 0xFF51AFD7ED558CCDull; h ^= h >> 33; h *= 0xC4CEB9FE1A85EC53ull; h ^= h >> 33;
  return h;
}
/* BASELINE.md B.2 "Key": xor, 64-bit multiply, 32-bit multiply of (low half ^ rot16(high half)), swap halves */
static inline uint64_t absorb(uint64_t h, uint64_t w) {
  uint64_t m = (h ^ w) * 0x9E3779B97F4A7C15ull;
  uint32_t hi = (uint32_t)(m >> 32);
  uint32_t lo = ((uint32_t)m ^ ((hi << 16) | (hi >> 16))) * 0x85EBCA6Bu;
  return ((uint64_t)lo << 32) | hi;
}
static uint64_t absorb_bytes(uint64_t h, const uint8_t* p, uint32_t len, uint32_t field) {
  uint32_t i = 0;
  for (; i + 8 <= len; i += 8) { uint64_t w; memcpy(&w, p + i, 8); h = absorb(h, w); }
  if (i < len) { uint64_t w = 0; memcpy(&w, p + i, len - i); h = absorb(h, w); }
  return absorb(h, (uint64_t)len | ((uint64_t)field << 56));
}
uint64_t bho_hash_bytes(const uint8_t* p, uint32_t len, uint32_t field, uint64_t seed) {
  return fmix64(absorb_bytes(seed ^ 0x6A09E667F3BCC908ull, p, len, field));
}

uint64_t bho_key_hash(const bho_msg* m, const uint8_t* buf, uint64_t seed, int reference_key) {
  if (reference_key) return (uint64_t)(uint32_t)m->status;
  if (m->n_parts == 1 && !m->part_is_len[0]) return (uint64_t)m->part_val[0];
  uint64_t h = seed ^ 0x6A09E667F3BCC908ull;
  for (uint32_t k = 0; k < m->n_parts; k++) {
    if (m->part_is_len[k]) h = absorb_bytes(h, buf + m->part_pos[k], m->part_len[k], m->part_field[k]);
    else h = absorb(h, (uint64_t)m->part_val[k] | ((uint64_t)m->part_field[k] << 56) | (1ull << 55));
  }
  return fmix64(h);
}

uint64_t bho_f64_order(double v) {
  uint64_t b; memcpy(&b, &v, 8);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
static double f64_from_order(uint64_t o) {
  uint64_t b = (o >> 63) ? (o & 0x7FFFFFFFFFFFFFFFull) : ~o;
  double v; memcpy(&v, &b, 8); return v;
}

uint32_t bho_hist_bucket(double v, uint32_t buckets, int32_t e0) {
  uint64_t b; memcpy(&b, &v, 8);
  if ((b >> 63) || (b << 1) == 0) return 0;            /* v <= 0 (incl. -0.0) */
  int64_t k = (int64_t)((b >> 52) & 0x7FF) - (int64_t)e0;
  if (k < 1) k = 1;
  if (k > (int64_t)buckets - 1) k = (int64_t)buckets - 1;
  return (uint32_t)k;
}

/* --- group table keyed by the FULL key (canonical bytes) + window --- */
typedef struct {
  uint64_t key_hash, window_id, count;
  double sum;
  uint64_t min_ord, max_ord;
  uint32_t canon_off, canon_len;
  uint32_t hist_idx;
} grp_t;

struct bho_agg {
  bho_schema schema; bho_agg_config cfg;
  grp_t* g; uint64_t n_g, cap_g;
  uint64_t* slots; uint64_t n_slots;     /* open addressing: index+1 into g */
  uint8_t* arena; uint64_t arena_len, arena_cap;
  uint32_t* hist; uint64_t hist_cap;
  uint64_t dropped[4];
};

static uint64_t fnv1a(const uint8_t* p, uint32_t n, uint64_t w) {
  uint64_t h = 0xCBF29CE484222325ull ^ (w * 0x100000001B3ull);
  for (uint32_t i = 0; i < n; i++) { h ^= p[i]; h *= 0x100000001B3ull; }
  return h ^ (h >> 29);
}

bho_agg* bho_agg_new(const bho_schema* s, const bho_agg_config* c) {
  bho_agg* a = calloc(1, sizeof *a);
  if (s) a->schema = *s;
  a->cfg = *c;
  if (a->cfg.hist_buckets == 0 || a->cfg.hist_buckets > BHO_HIST_MAX) a->cfg.hist_buckets = BHO_HIST_MAX;
  a->n_slots = 1024; a->slots = calloc(a->n_slots, 8);
  return a;
}
void bho_agg_free(bho_agg* a) {
  if (!a) return;
  free(a->g); free(a->slots); free(a->arena); free(a->hist); free(a);
}

static void agg_rehash(bho_agg* a) {
  uint64_t ns = a->n_slots * 2;
  uint64_t* sl = calloc(ns, 8);
  for (uint64_t i = 0; i < a->n_g; i++) {
    grp_t* g = &a->g[i];
    uint64_t h = fnv1a(a->arena + g->canon_off, g->canon_len, g->window_id) & (ns - 1);
    while (sl[h]) h = (h + 1) & (ns - 1);
    sl[h] = i + 1;
  }
  free(a->slots); a->slots = sl; a->n_slots = ns;
}

static grp_t* agg_find(bho_agg* a, const uint8_t* canon, uint32_t clen, uint64_t win, uint64_t kh) {
  uint64_t h = fnv1a(canon, clen, win) & (a->n_slots - 1);
  for (;;) {
    uint64_t s = a->slots[h];
    if (!s) break;
    grp_t* g = &a->g[s - 1];
    if (g->window_id == win && g->canon_len == clen && !memcmp(a->arena + g->canon_off, canon, clen)) return g;
    h = (h + 1) & (a->n_slots - 1);
  }
  if (a->n_g == a->cap_g) { a->cap_g = a->cap_g ? a->cap_g * 2 : 1024; a->g = realloc(a->g, a->cap_g * sizeof(grp_t)); }
  if (a->arena_len + clen > a->arena_cap) {
    a->arena_cap = (a->arena_cap ? a->arena_cap * 2 : 65536) + clen; a->arena = realloc(a->arena, a->arena_cap);
  }
  grp_t* g = &a->g[a->n_g];
  memset(g, 0, sizeof *g);
  g->key_hash = kh; g->window_id = win; g->canon_off = (uint32_t)a->arena_len; g->canon_len = clen;
  g->min_ord = ~0ull; g->max_ord = 0;
  memcpy(a->arena + a->arena_len, canon, clen); a->arena_len += clen;
  if (a->cfg.ops & BHO_OP_HIST) {
    uint64_t need = (a->n_g + 1) * a->cfg.hist_buckets;
    if (need > a->hist_cap) {
      uint64_t nc = a->hist_cap ? a->hist_cap * 2 : 65536; if (nc < need) nc = need;
      a->hist = realloc(a->hist, nc * 4); memset(a->hist + a->hist_cap, 0, (nc - a->hist_cap) * 4); a->hist_cap = nc;
    }
    g->hist_idx = (uint32_t)a->n_g;
  }
  a->slots[h] = ++a->n_g;
  if (a->n_g * 2 > a->n_slots) { agg_rehash(a); g = &a->g[a->n_g - 1]; }
  return g;
}

static void agg_update(bho_agg* a, grp_t* g, double v) {
  uint32_t ops = a->cfg.ops;
  g->count++;
  if (ops & BHO_OP_SUM) g->sum += v;
  if (ops & (BHO_OP_MIN | BHO_OP_MAX)) {
    uint64_t o = bho_f64_order(v);
    if ((ops & BHO_OP_MIN) && o < g->min_ord) g->min_ord = o;
    if ((ops & BHO_OP_MAX) && o > g->max_ord) g->max_ord = o;
  }
  if (ops & BHO_OP_HIST)
    a->hist[(uint64_t)g->hist_idx * a->cfg.hist_buckets + bho_hist_bucket(v, a->cfg.hist_buckets, a->cfg.hist_e0)]++;
}

static int needs_value(uint32_t ops) { return (ops & (BHO_OP_SUM | BHO_OP_MIN | BHO_OP_MAX | BHO_OP_HIST)) != 0; }

/* one event through the windowed handler: same order as index.js:129-138
 * (decode -> enum check -> aggregate), the counter generalised per §B.2 */
static void agg_event_raw(bho_agg* a, const uint8_t* buf, uint32_t len, uint64_t ts) {
  bho_msg m;
  bho_decode(&a->schema, buf, len, &m);
  if (m.outcome != BHO_EV_OK) { a->dropped[BHO_EV_DROP_DECODE]++; return; }
  if (a->schema.n_enum && enum_slot(&a->schema, m.status) < 0) { a->dropped[BHO_EV_DROP_ENUM]++; return; }
  if (needs_value(a->cfg.ops) && m.value != m.value) { a->dropped[BHO_EV_DROP_NAN]++; return; }
  uint64_t kh = bho_key_hash(&m, buf, a->cfg.hash_seed, (int)a->cfg.reference_key);
  uint64_t win = a->cfg.window_ms ? ts / a->cfg.window_ms : 0;
  /* canonical full key: per part [u32 field][u32 len-or-~0][bytes or 4-byte value] */
  uint8_t canon[4096]; uint32_t cl = 0;
  if (a->cfg.reference_key) { memcpy(canon, &m.status, 4); cl = 4; }
  else for (uint32_t k = 0; k < m.n_parts; k++) {
    uint32_t f = m.part_field[k];
    if (m.part_is_len[k]) {
      uint32_t n = m.part_len[k];
      if (cl + 8 + n > sizeof canon) n = (uint32_t)sizeof canon - cl - 8;  /* keys > 4 KB: compared on a prefix + length */
      memcpy(canon + cl, &f, 4); memcpy(canon + cl + 4, &m.part_len[k], 4); memcpy(canon + cl + 8, buf + m.part_pos[k], n);
      cl += 8 + n;
    } else {
      uint32_t tag = 0xFFFFFFFFu;
      memcpy(canon + cl, &f, 4); memcpy(canon + cl + 4, &tag, 4); memcpy(canon + cl + 8, &m.part_val[k], 4);
      cl += 12;
    }
  }
  agg_update(a, agg_find(a, canon, cl, win, kh), m.value);
}

void bho_agg_push_raw(bho_agg* a, const uint8_t* slab, const uint32_t* off, const uint64_t* ts, uint64_t n) {
  for (uint64_t i = 0; i < n; i++) agg_event_raw(a, slab + off[i], off[i + 1] - off[i], ts ? ts[i] : 0);
}

void bho_agg_push_columns(bho_agg* a, const uint64_t* kh, const uint64_t* ts, const double* val, uint64_t n) {
  for (uint64_t i = 0; i < n; i++) {
    double v = val ? val[i] : 0.0;
    if (needs_value(a->cfg.ops) && v != v) { a->dropped[BHO_EV_DROP_NAN]++; continue; }
    uint64_t win = a->cfg.window_ms ? ts[i] / a->cfg.window_ms : 0;
    agg_update(a, agg_find(a, (const uint8_t*)&kh[i], 8, win, kh[i]), v);
  }
}

void bho_agg_merge(bho_agg* a, const bho_agg* b) {
  uint32_t B = a->cfg.hist_buckets;
  for (uint64_t i = 0; i < b->n_g; i++) {
    const grp_t* s = &b->g[i];
    grp_t* g = agg_find(a, b->arena + s->canon_off, s->canon_len, s->window_id, s->key_hash);
    g->count += s->count; g->sum += s->sum;
    if (s->min_ord < g->min_ord) g->min_ord = s->min_ord;
    if (s->max_ord > g->max_ord) g->max_ord = s->max_ord;
    if (a->cfg.ops & BHO_OP_HIST)
      for (uint32_t k = 0; k < B; k++) a->hist[(uint64_t)g->hist_idx * B + k] += b->hist[(uint64_t)s->hist_idx * B + k];
  }
  for (int k = 0; k < 4; k++) a->dropped[k] += b->dropped[k];
}

uint64_t bho_agg_size(const bho_agg* a) { return a->n_g; }
uint64_t bho_agg_dropped(const bho_agg* a, int o) { return (o >= 0 && o < 4) ? a->dropped[o] : 0; }

static const bho_agg* g_sort_ctx;
static int cmp_idx(const void* x, const void* y) {
  const grp_t* p = &g_sort_ctx->g[*(const uint64_t*)x];
  const grp_t* q = &g_sort_ctx->g[*(const uint64_t*)y];
  if (p->window_id != q->window_id) return p->window_id < q->window_id ? -1 : 1;
  if (p->key_hash != q->key_hash) return p->key_hash < q->key_hash ? -1 : 1;
  return 0;
}

static uint64_t* sorted_index(const bho_agg* a) {
  uint64_t* idx = malloc((a->n_g + 1) * 8);
  for (uint64_t i = 0; i < a->n_g; i++) idx[i] = i;
  g_sort_ctx = a;                       /* not thread-safe; tests call it from one thread */
  qsort(idx, a->n_g, 8, cmp_idx);
  return idx;
}

uint64_t bho_agg_snapshot(const bho_agg* a, bho_record* out, uint32_t* hist_out, uint64_t cap) {
  uint64_t* idx = sorted_index(a);
  uint64_t n = a->n_g < cap ? a->n_g : cap;
  uint32_t B = a->cfg.hist_buckets;
  for (uint64_t i = 0; i < n; i++) {
    const grp_t* g = &a->g[idx[i]];
    out[i].key_hash = g->key_hash; out[i].window_id = g->window_id; out[i].count = g->count;
    out[i].sum = g->sum;
    out[i].min = (a->cfg.ops & BHO_OP_MIN) && g->count ? f64_from_order(g->min_ord) : 0.0;
    out[i].max = (a->cfg.ops & BHO_OP_MAX) && g->count ? f64_from_order(g->max_ord) : 0.0;
    if (hist_out && (a->cfg.ops & BHO_OP_HIST)) memcpy(hist_out + i * B, a->hist + (uint64_t)g->hist_idx * B, B * 4);
  }
  free(idx);
  return a->n_g;
}

uint64_t bho_agg_hash_collisions(const bho_agg* a) {
  uint64_t* idx = sorted_index(a);
  uint64_t c = 0;
  for (uint64_t i = 1; i < a->n_g; i++) {
    const grp_t* p = &a->g[idx[i - 1]]; const grp_t* q = &a->g[idx[i]];
    if (p->window_id == q->window_id && p->key_hash == q->key_hash) c++;
  }
  free(idx);
  return c;
}

void bho_decode_columns(const bho_schema* s, const bho_agg_config* c, const uint8_t* slab,
                        const uint32_t* off, uint64_t n, uint64_t* key_hash, double* value,
                        int32_t* status, uint8_t* outcome) {
  for (uint64_t i = 0; i < n; i++) {
    bho_msg m; const uint8_t* buf = slab + off[i];
    bho_decode(s, buf, off[i + 1] - off[i], &m);
    int o = m.outcome;
    if (o == BHO_EV_OK && s->n_enum && enum_slot(s, m.status) < 0) o = BHO_EV_DROP_ENUM;
    if (o == BHO_EV_OK && needs_value(c->ops) && m.value != m.value) o = BHO_EV_DROP_NAN;
    outcome[i] = (uint8_t)o;
    int have = (m.outcome == BHO_EV_OK);
    key_hash[i] = have ? bho_key_hash(&m, buf, c->hash_seed, (int)c->reference_key) : 0;
    value[i] = have ? m.value : 0.0; status[i] = have ? m.status : 0;
  }
}

/* --- threaded baselines ("restatement, N cores") --- */
typedef struct {
  const bho_schema* s; const bho_agg_config* c; const uint8_t* slab; const uint32_t* off;
  const uint64_t* ts; uint64_t lo, hi; bho_agg* out; int log_mode; bho_handler* h;
} job_t;

static void* agg_job(void* p) {
  job_t* j = p;
  j->out = bho_agg_new(j->s, j->c);
  bho_agg_push_raw(j->out, j->slab, j->off + j->lo, j->ts ? j->ts + j->lo : NULL, j->hi - j->lo);
  return NULL;
}

bho_agg* bho_agg_run_threads(const bho_schema* s, const bho_agg_config* c, const uint8_t* slab,
                             const uint32_t* off, const uint64_t* ts, uint64_t n, int nt) {
  if (nt < 1) nt = 1;
  job_t* jobs = calloc((size_t)nt, sizeof *jobs); pthread_t* th = calloc((size_t)nt, sizeof *th);
  for (int t = 0; t < nt; t++) {
    jobs[t] = (job_t){s, c, slab, off, ts, n * (uint64_t)t / (uint64_t)nt, n * (uint64_t)(t + 1) / (uint64_t)nt, NULL, 0, NULL};
    pthread_create(&th[t], NULL, agg_job, &jobs[t]);
  }
  for (int t = 0; t < nt; t++) pthread_join(th[t], NULL);
  bho_agg* a = jobs[0].out;
  for (int t = 1; t < nt; t++) { bho_agg_merge(a, jobs[t].out); bho_agg_free(jobs[t].out); }
  free(jobs); free(th);
  return a;
}

static void* handler_job(void* p) {
  job_t* j = p;
  j->h = bho_handler_new(j->s, j->log_mode);
  for (uint64_t i = j->lo; i < j->hi; i++) bho_handle_progress(j->h, j->slab + j->off[i], j->off[i + 1] - j->off[i]);
  return NULL;
}

void bho_handler_run_threads(const bho_schema* s, int log_mode, const uint8_t* slab, const uint32_t* off,
                             uint64_t n, int nt, double* counts_out, uint64_t* dropped_out) {
  if (nt < 1) nt = 1;
  job_t* jobs = calloc((size_t)nt, sizeof *jobs); pthread_t* th = calloc((size_t)nt, sizeof *th);
  for (int t = 0; t < nt; t++) {
    jobs[t] = (job_t){s, NULL, slab, off, NULL, n * (uint64_t)t / (uint64_t)nt, n * (uint64_t)(t + 1) / (uint64_t)nt, NULL, log_mode, NULL};
    pthread_create(&th[t], NULL, handler_job, &jobs[t]);
  }
  for (uint32_t i = 0; i < s->n_enum; i++) counts_out[i] = 0.0;
  for (int k = 0; k < 4; k++) dropped_out[k] = 0;
  for (int t = 0; t < nt; t++) {
    pthread_join(th[t], NULL);
    for (uint32_t i = 0; i < s->n_enum; i++) counts_out[i] += jobs[t].h->count[i];
    for (int k = 0; k < 4; k++) dropped_out[k] += jobs[t].h->dropped[k];
    bho_handler_free(jobs[t].h);
  }
  free(jobs); free(th);
}
