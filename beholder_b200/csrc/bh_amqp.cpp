// bh_amqp.cpp — AMQP 0-9-1 delivery tap (include/beholder_gpu.h, "AMQP 0-9-1 delivery tap").
// Host code only.  Stands where amqplib 0.5.5's frame parser + triton-core's `amqp.listen` dispatch
// stand in the reference (index.js:43-44,127; neither is vendored under /root/reference): it turns the
// raw broker -> client byte stream into slab + offsets + delivery tags, the input format of
// bh_push_batch, without building an object per message.  Grammar: AMQP 0-9-1 specification, section
// 4.2 (framing) and the basic class (deliver = 60/60, return = 60/50, get-ok = 60/71).
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <unordered_map>

#include "../../include/beholder_gpu.h"

namespace {

enum ChanState { IDLE = 0, WANT_HEADER, WANT_BODY, SKIP_HEADER, SKIP_BODY };

struct Chan {
  int state = IDLE;
  bool filtered = false;          // a delivery for another consumer: followed, not reported
  uint64_t tag = 0, body_size = 0, got = 0;
  uint16_t redelivered = 0;
  std::string part;               // body bytes of earlier frames of the message in flight
};

inline uint32_t be16(const uint8_t* p) { return (uint32_t)p[0] << 8 | p[1]; }
inline uint32_t be32(const uint8_t* p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }
inline uint64_t be64(const uint8_t* p) { return (uint64_t)be32(p) << 32 | be32(p + 4); }

constexpr uint8_t kFrameEnd = 0xCE;
constexpr uint32_t kMaxFrame = 1u << 27;   // far above any negotiated frame-max (amqplib default 4096, RabbitMQ 131072)

}  // namespace

struct bh_amqp {
  std::string consumer_tag;
  uint64_t max_body = 16ull << 20;
  std::unordered_map<uint32_t, Chan> chans;
  bh_amqp_stats stats{};
  std::string err;
  bool dead = false;             // after a protocol error every further call fails
};

// protocol errors (BH_ERR_INVALID) are final: the byte stream can no longer be trusted
static int amqp_fail(bh_amqp* a, int code, const char* fmt, ...) {
  char buf[256];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  a->err = buf;
  if (code == BH_ERR_INVALID) a->dead = true;
  return code;
}

extern "C" {

int bh_amqp_create(bh_amqp** out, const char* consumer_tag, uint64_t max_body) {
  if (!out) return BH_ERR_INVALID;
  bh_amqp* a = new bh_amqp();
  if (consumer_tag) a->consumer_tag = consumer_tag;
  if (max_body) a->max_body = max_body;
  *out = a;
  return BH_OK;
}

void bh_amqp_destroy(bh_amqp* a) { delete a; }

const char* bh_amqp_last_error(const bh_amqp* a) { return a ? a->err.c_str() : "null tap"; }

int bh_amqp_get_stats(const bh_amqp* a, bh_amqp_stats* out) {
  if (!a || !out) return BH_ERR_INVALID;
  *out = a->stats;
  return BH_OK;
}

int bh_amqp_feed(bh_amqp* a, const uint8_t* bytes, uint64_t len, uint8_t* slab, uint64_t slab_cap,
                 uint32_t* offsets, bh_amqp_delivery* deliveries, uint32_t cap, uint64_t* consumed, uint32_t* n_out) {
  if (!a) return BH_ERR_INVALID;
  if ((!bytes && len) || !offsets || !consumed || !n_out || (cap && (!deliveries || !slab))) { a->err = "null argument"; return BH_ERR_INVALID; }
  *consumed = 0; *n_out = 0; offsets[0] = 0;
  if (a->dead) return BH_ERR_INVALID;     // bh_amqp_last_error still holds the protocol error
  uint64_t pos = 0, used = 0;
  uint32_t n = 0;
  int rc = BH_OK;
  while (len - pos >= 8) {
    const uint8_t* f = bytes + pos;
    const uint32_t type = f[0], channel = be16(f + 1), size = be32(f + 3);
    if (type != 1 && type != 2 && type != 3 && type != 8) { rc = amqp_fail(a, BH_ERR_INVALID, "unknown frame type %u at byte %llu", type, (unsigned long long)pos); break; }
    if (size > kMaxFrame) { rc = amqp_fail(a, BH_ERR_INVALID, "frame of %u bytes at byte %llu", size, (unsigned long long)pos); break; }
    if (len - pos < 8ull + size) break;                           // incomplete: wait for more bytes
    if (f[7 + size] != kFrameEnd) { rc = amqp_fail(a, BH_ERR_INVALID, "bad frame end marker at byte %llu", (unsigned long long)(pos + 7 + size)); break; }
    const uint8_t* p = f + 7;
    bool stop = false;
    if (type == 8) {
      a->stats.heartbeats++;
    } else if (type == 1) {
      Chan& c = a->chans[channel];
      if (c.state != IDLE) { rc = amqp_fail(a, BH_ERR_INVALID, "method frame on channel %u while content is expected", channel); break; }
      if (size < 4) { rc = amqp_fail(a, BH_ERR_INVALID, "method frame of %u bytes", size); break; }
      const uint32_t cls = be16(p), mth = be16(p + 2);
      if (cls == 60 && mth == 60) {                               // basic.deliver
        uint32_t q = 4;
        auto shortstr = [&](const uint8_t*& s, uint32_t& sl) { if (q + 1 > size) return false; sl = p[q]; if (q + 1 + sl > size) return false; s = p + q + 1; q += 1 + sl; return true; };
        const uint8_t *ct, *ex, *rk; uint32_t ctl, exl, rkl;
        if (!shortstr(ct, ctl) || q + 9 > size) { rc = amqp_fail(a, BH_ERR_INVALID, "truncated basic.deliver"); break; }
        c.tag = be64(p + q); c.redelivered = p[q + 8] & 1u; q += 9;
        if (!shortstr(ex, exl) || !shortstr(rk, rkl)) { rc = amqp_fail(a, BH_ERR_INVALID, "truncated basic.deliver"); break; }
        c.filtered = !a->consumer_tag.empty() && (ctl != a->consumer_tag.size() || memcmp(ct, a->consumer_tag.data(), ctl) != 0);
        c.state = WANT_HEADER;
      } else if (cls == 60 && (mth == 50 || mth == 71)) {          // basic.return / basic.get-ok carry content too
        c.state = SKIP_HEADER;
      } else {
        a->stats.other_methods++;
      }
    } else if (type == 2) {
      Chan& c = a->chans[channel];
      if (c.state != WANT_HEADER && c.state != SKIP_HEADER) { rc = amqp_fail(a, BH_ERR_INVALID, "content header on channel %u without a method", channel); break; }
      if (size < 14) { rc = amqp_fail(a, BH_ERR_INVALID, "content header of %u bytes", size); break; }
      const uint64_t body = be64(p + 4);
      if (body > a->max_body) { rc = amqp_fail(a, BH_ERR_INVALID, "announced body of %llu bytes exceeds max_body %llu", (unsigned long long)body, (unsigned long long)a->max_body); break; }
      c.body_size = body; c.got = 0; c.part.clear();
      if (c.state == SKIP_HEADER) {
        c.state = body ? SKIP_BODY : IDLE;
        if (!body) a->stats.skipped_content++;
      } else if (body) {
        c.state = WANT_BODY;
      } else if (c.filtered) {
        a->stats.filtered++; c.state = IDLE;
      } else if (n == cap) {
        stop = true;                                              // the empty message does not fit: leave the header unconsumed
      } else {
        deliveries[n] = bh_amqp_delivery{c.tag, 0u, (uint16_t)channel, c.redelivered};
        offsets[++n] = (uint32_t)used;
        a->stats.deliveries++; c.state = IDLE;
      }
    } else {                                                      // body
      Chan& c = a->chans[channel];
      if (c.state != WANT_BODY && c.state != SKIP_BODY) { rc = amqp_fail(a, BH_ERR_INVALID, "content body on channel %u without a header", channel); break; }
      if (c.got + size > c.body_size) { rc = amqp_fail(a, BH_ERR_INVALID, "body on channel %u exceeds the announced %llu bytes", channel, (unsigned long long)c.body_size); break; }
      const bool last = c.got + size == c.body_size;
      if (c.state == SKIP_BODY || c.filtered) {
        c.got += size;
        if (last) { if (c.state == SKIP_BODY) a->stats.skipped_content++; else a->stats.filtered++; c.state = IDLE; }
      } else if (!last) {
        c.part.append((const char*)p, size); c.got += size;
      } else if (n == cap || used + c.body_size > slab_cap || used + c.body_size > 0xFFFFFFFFull) {
        if (n == 0 && cap && c.body_size > slab_cap) { rc = amqp_fail(a, BH_ERR_TOO_LARGE, "message of %llu bytes exceeds the slab (%llu)", (unsigned long long)c.body_size, (unsigned long long)slab_cap); break; }
        stop = true;                                              // flush and call again: this frame stays unconsumed
      } else {
        if (!c.part.empty()) memcpy(slab + used, c.part.data(), c.part.size());
        memcpy(slab + used + c.part.size(), p, size);
        used += c.body_size;
        deliveries[n] = bh_amqp_delivery{c.tag, (uint32_t)c.body_size, (uint16_t)channel, c.redelivered};
        offsets[++n] = (uint32_t)used;
        a->stats.deliveries++; a->stats.body_bytes += c.body_size;
        c.part.clear(); c.state = IDLE;
      }
    }
    if (stop) break;
    pos += 8ull + size;
    a->stats.frames++;
  }
  *consumed = pos; *n_out = n;
  return rc;
}

}  // extern "C"
