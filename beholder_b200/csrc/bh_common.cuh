// bh_common.cuh — definitions shared by the kernels, the C-ABI layer and the synthetic
// stream generator of libbeholder_gpu.so.  Host+device inline functions only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/beholder_gpu.h"

#define BH_HD __host__ __device__ __forceinline__

namespace bh {

// ---------------------------------------------------------------------------------------
// key hash (BASELINE.md §B.2 "Key"): absorb 8-byte little-endian words, finalise with fmix64.
// The 32-bit rotate is a register swap on the GPU.
// ---------------------------------------------------------------------------------------
constexpr uint64_t kHashInit = 0x6A09E667F3BCC908ull;
constexpr uint64_t kHashMul = 0x9E3779B97F4A7C15ull;
// kHashMul^-1 mod 2^64 (Newton: x <- x * (2 - a x) doubles the correct low bits; a is odd, a*a = 1 mod 8)
constexpr uint64_t inv_mod_2_64(uint64_t a) {
  uint64_t x = a;
  for (int i = 0; i < 6; i++) x *= 2 - a * x;
  return x;
}
constexpr uint64_t kHashMulInv = inv_mod_2_64(kHashMul);
static_assert(kHashMul * kHashMulInv == 1ull, "modular inverse");

// One word into the state: xor, 64-bit multiply, then a second, 32-bit multiply of the product's low half
// mixed with its high half rotated by 16, halves swapped.  Why two multiplies: a bit that sits at
// position 63 when a product is formed changes nothing but bit 63 of it, a fixed one-bit difference that
// the NEXT word can cancel exactly.  The first version (one multiply + swap) let two keys differing in
// bit 63 of one word and bit 31 of the next collide; folding the halves before the multiply only moved
// the pattern to three bits (both found by tests: test_gpu_fuzz.py, test_oracle_extension.py).  Here the
// product's top bit enters the second multiply at bit 15, so every input difference has become a
// data-dependent many-bit difference before the next word arrives.  Each step is a bijection of the state.
constexpr uint32_t kHashMul32 = 0x85EBCA6Bu;
BH_HD uint64_t absorb(uint64_t h, uint64_t w) {
  const uint64_t m = (h ^ w) * kHashMul;
  const uint32_t hi = (uint32_t)(m >> 32);
  const uint32_t lo = ((uint32_t)m ^ ((hi << 16) | (hi >> 16))) * kHashMul32;
  return ((uint64_t)lo << 32) | hi;
}
BH_HD uint64_t fmix64(uint64_t h) {
  h ^= h >> 33; h *= 0xFF51AFD7ED558CCDull; h ^= h >> 33; h *= 0xC4CEB9FE1A85EC53ull; h ^= h >> 33;
  return h;
}
BH_HD uint64_t len_word(uint32_t len, uint32_t field) { return (uint64_t)len | ((uint64_t)field << 56); }
BH_HD uint64_t varint_word(uint32_t v, uint32_t field) {
  return (uint64_t)v | ((uint64_t)field << 56) | (1ull << 55);
}

// order-preserving map f64 -> u64 (total order on non-NaN; -0.0 < +0.0)
BH_HD uint64_t f64_order(uint64_t bits) { return (bits >> 63) ? ~bits : (bits | 0x8000000000000000ull); }
BH_HD uint64_t f64_unorder(uint64_t o) { return (o >> 63) ? (o & 0x7FFFFFFFFFFFFFFFull) : ~o; }

BH_HD uint32_t hist_bucket(uint64_t bits, uint32_t buckets, int32_t e0) {
  if ((bits >> 63) || (bits << 1) == 0) return 0;  // v <= 0
  int32_t k = (int32_t)((bits >> 52) & 0x7FF) - e0;
  k = k < 1 ? 1 : k;
  k = k > (int32_t)buckets - 1 ? (int32_t)buckets - 1 : k;
  return (uint32_t)k;
}

// ---------------------------------------------------------------------------------------
// exact u64 division by a run-time invariant (window_id = floor(ts_ms / window_ms))
// ---------------------------------------------------------------------------------------
struct Div64 {
  uint64_t magic;  // 0: power of two
  uint32_t shift;
  uint32_t add;
};

inline Div64 make_div64(uint64_t d) {
  Div64 r{0, 0, 0};
  uint32_t L = 63u - (uint32_t)__builtin_clzll(d);
  if ((d & (d - 1)) == 0) { r.shift = L; return r; }
  unsigned __int128 num = (unsigned __int128)1 << (64 + L);
  uint64_t m = (uint64_t)(num / d);
  uint64_t rem = (uint64_t)(num - (unsigned __int128)m * d);
  uint64_t e = d - rem;
  if (e < (1ull << L)) {
    r.shift = L;
  } else {
    m += m;
    uint64_t twice = rem + rem;
    if (twice >= d || twice < rem) m += 1;
    r.shift = L; r.add = 1;
  }
  r.magic = m + 1;
  return r;
}

BH_HD uint64_t mulhi64(uint64_t a, uint64_t b) {
#ifdef __CUDA_ARCH__
  return __umul64hi(a, b);
#else
  return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}

BH_HD uint64_t div64(uint64_t n, const Div64& d) {
  if (d.magic == 0) return n >> d.shift;
  uint64_t q = mulhi64(d.magic, n);
  if (d.add) { uint64_t t = ((n - q) >> 1) + q; return t >> d.shift; }
  return q >> d.shift;
}

// ---------------------------------------------------------------------------------------
// device-side view of the schema and the tables
// ---------------------------------------------------------------------------------------
struct DevSchema {
  uint32_t fld[32];     // by field number: kind | role << 8 | (key part index + 1) << 16; 0 = unknown field
  uint8_t part_field[4];
  uint8_t part_is_len[4];
  uint32_t n_parts;
  uint32_t identity_key;  // key_hash = the single varint key part itself (REFERENCE mode: status)
  uint32_t status_part;   // REFERENCE mode: use status as the identity key
  uint32_t has_enum;
  uint64_t enum_mask;     // valid status numbers 0..63
  uint32_t mode, ops;
  uint32_t hist_buckets;
  int32_t hist_e0;
  uint64_t hash_seed;
  uint64_t hash_mask;     // ANDed onto every key hash (all ones unless the debug_hash_mask test hook is set)
  uint64_t window_ms;     // the divisor behind wdiv (0 = no windows)
  // byte offsets a message of THIS launch may occupy (set per launch): offsets outside are treated as a message
  // that does not decode, and nothing outside is read, whatever the offsets table says
  uint32_t pay_lo, pay_hi;
  uint32_t has_window;
  uint32_t int_values;    // the VALUE field is int32 / sint32 / bool (or absent): integer shared-memory table
  // fields the fast decoder recognises by number (0xFFFFFFFF = not applicable)
  uint32_t fast_key0;        // key part 0 when it is a `string` with no other role
  uint32_t fast_status;      // status when it is int32/uint32/enum (roles: status, optionally key)
  uint32_t fast_status_part; // its key part index + 1, or 0
  uint32_t fast_value;       // value when its only role is value and it is int32 or double
  uint32_t fast_value_f64;   // ... and it is a double
  uint32_t tag_key0, tag_status, tag_value;   // their canonical one-byte tags (0xFFFFFFFF = none)
  uint32_t big_messages;     // messages above ~72 B expected: kernels that stage a batch in parts
  // by field number 0..15 (one-byte tags): bit set = a `string` with no role / a field the schema does not know
  uint32_t skip_str_mask, skip_unk_mask;
  Div64 wdiv;
  // array members through functions: the compile-time schema view (ProgressSchema) offers the same names
  __host__ __device__ __forceinline__ uint32_t field(uint32_t number) const { return fld[number]; }
  __host__ __device__ __forceinline__ uint32_t part_len(int k) const { return part_is_len[k]; }
  __host__ __device__ __forceinline__ uint32_t part_fld(int k) const { return part_field[k]; }
};

struct Stat {            // 32 B = one L2 sector; zero-initialised
  unsigned long long count;
  double sum;
  unsigned long long min_inv;  // ~order(min); 0 = none yet
  unsigned long long max_ord;  //  order(max); 0 = none yet
};

constexpr uint64_t kEmptyKey = 0xFFFFFFFFFFFFFFFFull;  // never stored as a key: handled by side slots
constexpr uint64_t kLockKey = 0xFFFFFFFFFFFFFFFEull;   // shared-memory table only
constexpr uint64_t kNoWindow = 0xFFFFFFFFFFFFFFFFull;

// BH_F_VERIFY_KEYS: per key hash, the key bytes first seen with it.  64 bytes = one entry.
struct DictEntry {
  unsigned long long kh;        // kEmptyKey: free
  unsigned int ready;           // 0 while the winner of the slot is still writing
  unsigned int len;             // bytes of the string key part
  unsigned int status;          // value of the status key part (0 when the key has none)
  unsigned int pad;
  unsigned char bytes[40];      // the key bytes; longer keys: the first 32 bytes + an 8-byte hash of all of them
};
static_assert(sizeof(DictEntry) == 64, "one line");

struct DevTables {
  unsigned long long* ring_win;  // [ring] window id held by each ring slot, kNoWindow = free
  unsigned long long* keys;      // [ring][stride]
  Stat* stats;                   // [ring][stride]
  uint32_t* hist;                // [ring][stride][B] or nullptr
  unsigned long long* counters;  // bh_counters as u64[8]
  uint64_t cap;                  // power of two
  uint64_t stride;               // cap + 2 (side slots for the two sentinel-valued keys)
  uint32_t ring;                 // power of two
  uint32_t hist_buckets;
  // ---- partitioned (deferred) aggregation for large key sets; part_lg == 0: off ----
  // An event that misses the per-CTA shared-memory table is not applied to the HBM tables at once (one random
  // DRAM line per table per event) but appended as a 16-byte tuple to the region [CTA][partition] of a log,
  // partition = top part_lg bits of key_hash * kHashMul = the top bits of its table slot, so that a
  // partition's slots are one contiguous slice of every window's table.  k_part_drain later gives each
  // partition to ONE CTA, which aggregates its tuples in shared memory and touches only that slice.
  uint4* part_tuples;            // [part_ctas][P][part_cap]: {hs low bits | ring slot << (64 - part_lg), value bits}
  uint32_t* part_fill;           // [part_ctas][P] tuples in each region (kept between the launches of an epoch)
  uint32_t part_lg;              // log2(P)
  uint32_t part_cap;             // tuples per region
  uint32_t part_ctas;            // regions per partition = the largest grid of the scattering kernels
  uint32_t part_smem_off;        // where a scattering kernel keeps its P cursors in dynamic shared memory (set per launch)
  DictEntry* dict;               // BH_F_VERIFY_KEYS, else nullptr
  uint64_t dict_cap;             // power of two
};

enum CounterIdx { C_OK = 0, C_DECODE = 1, C_ENUM = 2, C_NAN = 3, C_TABLE_FULL = 4, C_WINDOW = 5, C_SPILL = 6, C_SLOW = 7, C_COLLISION = 8 };
constexpr int kCounters = 9;       // sizeof(bh_counters) / 8
constexpr int kCounterSlots = 16;  // u64 slots allocated for them



// ---------------------------------------------------------------------------------------
// splitmix64 and the synthetic stream, one definition for host and device
// ---------------------------------------------------------------------------------------
BH_HD uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
BH_HD uint64_t synth_rnd(uint64_t seed, uint64_t g, uint32_t lane) {
  return splitmix64(seed ^ splitmix64(g * 8 + lane));
}

struct SynthEvent {
  uint64_t key_index;
  uint64_t ts_ms;
  uint32_t status;
  double value;
  uint32_t host_len;       // 0 = no host field
  char media_id[36];
  char host[16];
};

BH_HD char hex_digit(uint32_t v) { return (char)(v < 10 ? '0' + v : 'a' + (v - 10)); }

BH_HD void synth_media_id(uint64_t seed, uint64_t k, char* out) {
  uint64_t a = splitmix64(seed + 0x1000 + 2 * k), b = splitmix64(seed + 0x1000 + 2 * k + 1);
  int o = 0;
  for (int i = 0; i < 16; i++) {
    if (i == 8 || i == 12) out[o++] = '-';
    out[o++] = hex_digit((uint32_t)(a >> (60 - 4 * i)) & 15);
  }
  out[o++] = '-';
  for (int i = 0; i < 16; i++) {
    if (i == 4) out[o++] = '-';
    out[o++] = hex_digit((uint32_t)(b >> (60 - 4 * i)) & 15);
  }
}

BH_HD uint64_t synth_key_index(const bh_synth_config& c, uint64_t g, const uint64_t* cdf) {
  uint64_t u = synth_rnd(c.seed, g, 0);
  if (!c.zipf || !cdf) return mulhi64(u, c.n_keys);
  uint64_t lo = 0, hi = c.n_keys - 1;  // first k with cdf[k] > u; cdf[n_keys-1] = 2^64-1 catches all
  while (lo < hi) {
    uint64_t mid = (lo + hi) >> 1;
    if (cdf[mid] > u) hi = mid; else lo = mid + 1;
  }
  return lo;
}

BH_HD double synth_value(const bh_synth_config& c, uint64_t g) {
  uint64_t r = synth_rnd(c.seed, g, 2);
  if (!c.value_f64) return (double)(r % 101);
  uint64_t mant = r & 0xFFFFFFFFFFFFFull;
  uint64_t span = c.value_exp_span ? c.value_exp_span : 1;
  uint64_t e = (uint64_t)(1023 + c.value_exp_lo) + ((r >> 52) & 0x7FF) % span;
  uint64_t bits = (e << 52) | mant;
  if (c.value_signed && (r >> 63)) bits |= 0x8000000000000000ull;
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)bits);
#else
  double v; memcpy(&v, &bits, 8); return v;
#endif
}

BH_HD void synth_event_fields(const bh_synth_config& c, uint64_t g, const uint64_t* cdf, SynthEvent& e) {
  e.key_index = synth_key_index(c, g, cdf);
  uint32_t ns = c.n_status ? c.n_status : 1;
  e.status = (uint32_t)(e.key_index % ns);
  uint64_t r1 = synth_rnd(c.seed, g, 1);
  if (c.unknown_enum_ppm && (r1 % 1000000u) < c.unknown_enum_ppm) e.status = 1000u + (uint32_t)((r1 >> 32) % 1000u);
  e.value = synth_value(c, g);
  uint64_t r3 = synth_rnd(c.seed, g, 3);
  if ((r3 & 15) == 0) {
    e.host_len = 0;
  } else {
    e.host_len = 6 + (uint32_t)((r3 >> 4) % 11);
    e.host[0] = 'h'; e.host[1] = 'o'; e.host[2] = 's'; e.host[3] = 't'; e.host[4] = '-';
    for (uint32_t i = 5; i < e.host_len; i++) e.host[i] = hex_digit((uint32_t)(r3 >> (8 + 4 * (i - 5))) & 15);
  }
  uint64_t rate = c.rate_per_ms ? c.rate_per_ms : 1;
  uint64_t base = c.t0_ms + g / rate;
  if (c.jitter_ms) {
    uint64_t r4 = synth_rnd(c.seed, g, 4);
    uint64_t j = r4 % (2ull * c.jitter_ms + 1);
    base = base + j - c.jitter_ms;
  }
  e.ts_ms = base;
  synth_media_id(c.seed, e.key_index, e.media_id);
}

struct Emit {
  uint8_t* p;
  uint32_t n;
  uint32_t limit;   // bytes the event owns in the slab (a truncated event owns fewer than it emits)
  BH_HD void put(uint32_t b) { if (p && n < limit) p[n] = (uint8_t)b; n++; }
  BH_HD void varint(uint32_t v) { while (v >= 128) { put((v & 127) | 128); v >>= 7; } put(v); }
};

// Builds (out != nullptr, at most `limit` bytes written) or measures (out == nullptr) the wire bytes
// of event g.
BH_HD uint32_t synth_build(const bh_synth_config& c, uint64_t g, const uint64_t* cdf, uint8_t* out, uint64_t* ts,
                           uint32_t limit) {
  SynthEvent e;
  synth_event_fields(c, g, cdf, e);
  if (ts) *ts = e.ts_ms;
  uint32_t status_len = e.status < 128 ? 1u : (e.status < 16384 ? 2u : 3u);
  uint32_t natural = 38 + 1 + status_len + (c.value_f64 ? 9u : 2u) + (e.host_len ? 2u + e.host_len : 0u);
  Emit w{out, 0, limit};
  w.put(0x0A); w.put(36);                                  // field 1 mediaId: string
  for (int i = 0; i < 36; i++) w.put((uint8_t)e.media_id[i]);
  if (c.payload_bytes && natural + 2 <= c.payload_bytes) { // field 15: ignored pad the decoder must skip
    uint32_t R = c.payload_bytes - natural;
    w.put(0x7A);
    uint32_t p;
    if (R - 2 < 128) { p = R - 2; w.put(p); }
    else { p = R - 3; w.put((p & 127) | 128); w.put(p >> 7); }   // always 2 length bytes (non-minimal if p < 128)
    for (uint32_t i = 0; i < p; i++) w.put(0x2E);
  }
  uint32_t status_pos = w.n;
  w.put(0x10); w.varint(e.status);                         // field 2 status: enum
  uint32_t value_pos = w.n;
  if (c.value_f64) {                                       // field 3 progress
    w.put(0x19);
    uint64_t bits;
#ifdef __CUDA_ARCH__
    bits = (uint64_t)__double_as_longlong(e.value);
#else
    memcpy(&bits, &e.value, 8);
#endif
    for (int i = 0; i < 8; i++) w.put((uint32_t)(bits >> (8 * i)) & 255);
  } else {
    w.put(0x18); w.varint((uint32_t)e.value);
  }
  uint32_t host_pos = w.n;
  if (e.host_len) {                                        // field 4 host: string
    w.put(0x22); w.put(e.host_len);
    for (uint32_t i = 0; i < e.host_len; i++) w.put((uint8_t)e.host[i]);
  }
  uint32_t len = w.n;
  if (c.malformed_ppm) {
    uint64_t r5 = synth_rnd(c.seed, g, 5);
    if ((r5 % 1000000u) < c.malformed_ppm) {
      switch ((r5 >> 40) & 3) {
        case 0: len = len / 2; break;                                        // truncated delivery
        case 1: if (out && limit) out[0] = 0x0F; break;                      // known field, wire type 7
        case 2: {                                                            // unknown field 9, wire type 7 / 4
          uint32_t at = e.host_len ? host_pos : value_pos;
          if (out && at < limit) out[at] = e.host_len ? 0x4F : 0x4C;
          break;
        }
        default: if (out && status_pos + 1 < limit) out[status_pos + 1] |= 0x80; break;   // status varint runs on
      }
    }
  }
  return len;
}

}  // namespace bh
