// bh_kernels.cu — sm_100a kernels of the beholder progress-event path.
//
//   k_process_raw      proto.decode + enumToString drop rule + counter update, fused
//                      (index.js:129-138), one lane per message, payload staged per warp
//                      through shared memory with 128-bit loads, shared-memory group-by table
//                      in front of the HBM tables.
//   k_process_columns  the counter update alone on pre-decoded columns (index.js:136-138).
//   k_decode           proto.decode alone -> columns (index.js:129-134).
//   k_export_* / k_merge / k_close_*   window close: compact, merge partials, recycle.
//
// Nothing here is a dense contraction: no tensor cores.  All kernels are HBM / L2-atomic bound.
#include "bh_kernels.cuh"

namespace bh {

// =========================================================================================
// global (HBM) tables
// =========================================================================================
constexpr uint32_t kGslotFull = 0xFFFFFFFFu;
constexpr uint32_t kGslotWindow = 0xFFFFFFFEu;

__device__ __forceinline__ uint32_t ilog2_u64(uint64_t v) { return 63u - (uint32_t)__clzll((long long)v); }

// Find or claim the slot of (window, key).  Returns the slot inside the window's table, or
// kGslotFull / kGslotWindow.  *ring_slot receives the window's ring index.
__device__ __noinline__ uint32_t g_find_or_insert(const DevTables& T, uint64_t win, uint64_t kh, uint32_t* ring_slot) {
  uint32_t r = (uint32_t)win & (T.ring - 1);
  *ring_slot = r;
  unsigned long long w = T.ring_win[r];
  if (w != win) {
    if (w != kNoWindow) return kGslotWindow;
    unsigned long long old = atomicCAS(&T.ring_win[r], (unsigned long long)kNoWindow, (unsigned long long)win);
    if (old != kNoWindow && old != win) return kGslotWindow;
  }
  unsigned long long* keys = T.keys + (uint64_t)r * T.stride;
  if (kh >= kLockKey) return (uint32_t)T.cap + (uint32_t)(kh & 1);   // side slots for the sentinel-valued keys
  uint32_t shift = 64u - ilog2_u64(T.cap);
  uint64_t slot = (kh * kHashMul) >> shift;
  for (uint64_t i = 0; i < T.cap; i++) {
    unsigned long long k = keys[slot];
    if (k == kh) return (uint32_t)slot;
    if (k == kEmptyKey) {
      unsigned long long old = atomicCAS(&keys[slot], (unsigned long long)kEmptyKey, (unsigned long long)kh);
      if (old == kEmptyKey || old == kh) return (uint32_t)slot;
    }
    slot = (slot + 1) & (T.cap - 1);
  }
  return kGslotFull;
}

__device__ __forceinline__ void g_update(const DevTables& T, uint32_t ops, uint64_t gi, unsigned long long cnt,
                                         double sum, unsigned long long min_inv, unsigned long long max_ord) {
  Stat* s = T.stats + gi;
  atomicAdd(&s->count, cnt);
  if (ops & BH_OP_SUM) atomicAdd(&s->sum, sum);
  if ((ops & BH_OP_MIN) && min_inv > s->min_inv) atomicMax(&s->min_inv, min_inv);
  if ((ops & BH_OP_MAX) && max_ord > s->max_ord) atomicMax(&s->max_ord, max_ord);
}

// =========================================================================================
// shared-memory front table (per CTA)
// =========================================================================================
struct SmemTable {
  unsigned long long* key;
  double* sum;
  unsigned long long* mn;   // ~order(min)
  unsigned long long* mx;   //  order(max)
  uint32_t* cnt;
  uint32_t* wrel;
  uint32_t* gslot;
  uint32_t ns;              // power of two
};
constexpr uint32_t kSmemSlotBytes = 8 * 4 + 4 * 3;
constexpr int kSmemProbes = 12;

__device__ __forceinline__ SmemTable smem_table_carve(uint8_t* base, uint32_t ns) {
  SmemTable M;
  M.key = (unsigned long long*)base;
  M.sum = (double*)(base + (size_t)ns * 8);
  M.mn = (unsigned long long*)(base + (size_t)ns * 16);
  M.mx = (unsigned long long*)(base + (size_t)ns * 24);
  M.cnt = (uint32_t*)(base + (size_t)ns * 32);
  M.wrel = M.cnt + ns;
  M.gslot = M.wrel + ns;
  M.ns = ns;
  return M;
}

__device__ __forceinline__ void smem_table_clear(const SmemTable& M) {
  for (uint32_t i = threadIdx.x; i < M.ns; i += blockDim.x) {
    M.key[i] = kEmptyKey; M.sum[i] = 0.0; M.mn[i] = 0; M.mx[i] = 0; M.cnt[i] = 0; M.wrel[i] = 0; M.gslot[i] = 0;
  }
}

// returns the shared slot, or -1 when the event has to go to the HBM table directly
__device__ __forceinline__ int smem_find_or_insert(const SmemTable& M, const DevTables& T, uint64_t kh,
                                                   uint32_t wrel, uint64_t win) {
  uint32_t slot = (uint32_t)(((kh ^ ((uint64_t)wrel * 0xD6E8FEB86659FD93ull)) * kHashMul) >> 32) & (M.ns - 1);
  volatile unsigned long long* vkey = M.key;
  volatile uint32_t* vwrel = M.wrel;
#pragma unroll 1
  for (int p = 0; p < kSmemProbes; p++) {
    unsigned long long k = vkey[slot];
    if (k == kh) {
      if (vwrel[slot] == wrel) return (int)slot;
    } else if (k == kEmptyKey) {
      unsigned long long old = atomicCAS(&M.key[slot], (unsigned long long)kEmptyKey, (unsigned long long)kLockKey);
      if (old == kEmptyKey) {   // claimed: resolve the HBM slot once, then publish
        uint32_t r;
        uint32_t gs = g_find_or_insert(T, win, kh, &r);
        M.gslot[slot] = gs;
        vwrel[slot] = wrel;
        __threadfence_block();
        vkey[slot] = kh;
        return (int)slot;
      }
      if (old == kh && vwrel[slot] == wrel) return (int)slot;
    }
    slot = (slot + 1) & (M.ns - 1);
  }
  return -1;
}

struct LaneCounters {
  uint32_t c[7];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int i = 0; i < 7; i++) c[i] = 0;
  }
};

// One event into the tables.  `wbase` is the CTA's first window minus one.
__device__ __forceinline__ void aggregate_event(const DevSchema& S, const DevTables& T, const SmemTable& M,
                                                uint64_t kh, uint64_t win, uint64_t wbase, double v,
                                                LaneCounters& C) {
  const uint32_t ops = S.ops;
  uint64_t bits = (uint64_t)__double_as_longlong(v);
  uint64_t ord = f64_order(bits);
  uint64_t wr = win - wbase;
  int slot = -1;
  if (kh < kLockKey && wr < 0xFFFFFFFFull && M.ns) slot = smem_find_or_insert(M, T, kh, (uint32_t)wr, win);
  if (slot >= 0) {
    uint32_t gs = M.gslot[slot];
    if (gs >= kGslotWindow) { C.c[gs == kGslotFull ? C_TABLE_FULL : C_WINDOW]++; return; }
    atomicAdd(&M.cnt[slot], 1u);
    if (ops & BH_OP_SUM) atomicAdd(&M.sum[slot], v);
    if ((ops & BH_OP_MIN) && ~ord > M.mn[slot]) atomicMax(&M.mn[slot], (unsigned long long)~ord);
    if ((ops & BH_OP_MAX) && ord > M.mx[slot]) atomicMax(&M.mx[slot], (unsigned long long)ord);
    if (ops & BH_OP_HIST) {
      uint64_t gi = (uint64_t)((uint32_t)win & (T.ring - 1)) * T.stride + gs;
      atomicAdd(&T.hist[gi * T.hist_buckets + hist_bucket(bits, T.hist_buckets, S.hist_e0)], 1u);
    }
    C.c[C_OK]++;
    return;
  }
  // spill: straight to the HBM table
  uint32_t r;
  uint32_t gs = g_find_or_insert(T, win, kh, &r);
  if (gs >= kGslotWindow) { C.c[gs == kGslotFull ? C_TABLE_FULL : C_WINDOW]++; return; }
  uint64_t gi = (uint64_t)r * T.stride + gs;
  g_update(T, ops, gi, 1ull, v, ~ord, ord);
  if (ops & BH_OP_HIST) atomicAdd(&T.hist[gi * T.hist_buckets + hist_bucket(bits, T.hist_buckets, S.hist_e0)], 1u);
  C.c[C_OK]++; C.c[C_SPILL]++;
}

__device__ __forceinline__ void smem_table_flush(const DevSchema& S, const DevTables& T, const SmemTable& M, uint64_t wbase) {
  for (uint32_t i = threadIdx.x; i < M.ns; i += blockDim.x) {
    unsigned long long k = M.key[i];
    uint32_t c = M.cnt[i];
    if (k == kEmptyKey || c == 0) continue;
    uint32_t gs = M.gslot[i];
    if (gs >= kGslotWindow) continue;
    uint64_t win = wbase + M.wrel[i];
    uint64_t gi = (uint64_t)((uint32_t)win & (T.ring - 1)) * T.stride + gs;
    g_update(T, S.ops, gi, c, M.sum[i], M.mn[i], M.mx[i]);
  }
}

__device__ __forceinline__ void counters_flush(const DevTables& T, LaneCounters& C, uint32_t* s_cnt /*[8] zeroed*/) {
#pragma unroll
  for (int i = 0; i < 7; i++) {
    uint32_t v = __reduce_add_sync(0xFFFFFFFFu, C.c[i]);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_cnt[i], v);
  }
  __syncthreads();
  if (threadIdx.x < 7 && s_cnt[threadIdx.x]) atomicAdd(&T.counters[threadIdx.x], (unsigned long long)s_cnt[threadIdx.x]);
}

// =========================================================================================
// protobufjs 6.8.8 reader semantics on the device (what proto.decode does, index.js:129)
// =========================================================================================
struct SmemRd {           // event bytes staged in shared memory; 16 B of slack after the batch
  const uint8_t* p;
  __device__ __forceinline__ uint32_t b(uint32_t i) const { return p[i]; }
  __device__ __forceinline__ uint64_t le64(uint32_t i, uint32_t nvalid) const {
    uint32_t a = (uint32_t)__cvta_generic_to_shared(p + i);
    uint32_t sh = (a & 3u) * 8u;
    a &= ~3u;
    uint32_t w0, w1, w2;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w0) : "r"(a));
    asm volatile("ld.shared.u32 %0, [%1+4];" : "=r"(w1) : "r"(a));
    asm volatile("ld.shared.u32 %0, [%1+8];" : "=r"(w2) : "r"(a));
    uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
    uint64_t w = ((uint64_t)hi << 32) | lo;
    return nvalid >= 8 ? w : (w & ((1ull << (8 * nvalid)) - 1));
  }
};
struct GmemRd {           // slow path: event parsed straight from global memory
  const uint8_t* p;
  __device__ __forceinline__ uint32_t b(uint32_t i) const { return __ldg(p + i); }
  __device__ __forceinline__ uint64_t le64(uint32_t i, uint32_t nvalid) const {
    uint64_t w = 0;
    for (uint32_t j = 0; j < 8 && j < nvalid; j++) w |= (uint64_t)__ldg(p + i + j) << (8 * j);
    return w;
  }
};

// Reader.uint32(): 5 value bytes; a longer varint is assumed to be 10 bytes; bytes past the end read
// as `undefined` (no value bits, "continue").
template <class R>
__device__ __forceinline__ bool rd_uint32(const R& r, uint32_t& pos, uint32_t len, uint32_t& out) {
  uint32_t v = 0;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    uint32_t byte = pos < len ? r.b(pos) : 0x80u;
    pos++;
    v |= (byte & (i == 4 ? 15u : 127u)) << (7 * i);
    if (byte < 128) { out = v; return true; }
  }
  pos += 5;
  out = v;
  return pos <= len;
}

// Reader.skipType(): recursion on start-group unrolled into a depth counter
template <class R>
__device__ __forceinline__ bool rd_skip_type(const R& r, uint32_t& pos, uint32_t len, uint32_t wire) {
  uint32_t depth = 0;
  for (;;) {
    switch (wire) {
      case 0:
        for (;;) { if (pos >= len) return false; if (!(r.b(pos++) & 128)) break; }
        break;
      case 1: if ((uint64_t)pos + 8 > len) return false; pos += 8; break;
      case 2: { uint32_t n; if (!rd_uint32(r, pos, len, n)) return false;
                if ((uint64_t)pos + n > len) return false; pos += n; break; }
      case 3: depth++; break;
      case 5: if ((uint64_t)pos + 4 > len) return false; pos += 4; break;
      case 4: if (depth == 0) return false; depth--; break;
      default: return false;
    }
    if (depth == 0) return true;
    uint32_t t;
    if (!rd_uint32(r, pos, len, t)) return false;
    wire = t & 7u;
  }
}

struct Parsed {
  uint32_t status;
  double value;
  uint32_t ppos[BH_MAX_KEY_PARTS], plen[BH_MAX_KEY_PARTS], pval[BH_MAX_KEY_PARTS];
};

// Type.decode(): known fields are read by DECLARED type whatever the wire type; string() on a Node
// Buffer clamps to the end, bytes() throws.  Returns false where proto.decode would throw.
template <class R>
__device__ __forceinline__ bool decode_event(const DevSchema& S, const R& r, uint32_t len, Parsed& P) {
  P.status = 0; P.value = 0.0;
#pragma unroll
  for (int k = 0; k < BH_MAX_KEY_PARTS; k++) { P.ppos[k] = 0; P.plen[k] = 0; P.pval[k] = 0; }
  uint32_t pos = 0;
  while (pos < len) {
    uint32_t t;
    if (!rd_uint32(r, pos, len, t)) return false;
    uint32_t fn = t >> 3;
    uint32_t kind = fn < 32 ? S.kind[fn] : 0;
    if (kind == 0) { if (!rd_skip_type(r, pos, len, t & 7u)) return false; continue; }
    uint32_t role = S.role[fn];
    uint32_t u = 0, spos = 0, slen = 0;
    double num = 0.0;
    if (kind <= BH_K_BYTES) {
      uint32_t n;
      if (!rd_uint32(r, pos, len, n)) return false;
      uint64_t end = (uint64_t)pos + n;
      if (end > len) { if (kind == BH_K_BYTES) return false; end = len; }
      spos = pos; slen = (uint32_t)end - pos; pos = (uint32_t)end;
    } else if (kind <= BH_K_BOOL) {
      if (!rd_uint32(r, pos, len, u)) return false;
      if (kind == BH_K_SINT32) u = (u >> 1) ^ (0u - (u & 1u));
      else if (kind == BH_K_BOOL) u = u != 0;
      num = kind == BH_K_UINT32 || kind == BH_K_BOOL ? (double)u : (double)(int32_t)u;
    } else if (kind == BH_K_FLOAT) {
      if ((uint64_t)pos + 4 > len) return false;
      num = (double)__uint_as_float((uint32_t)r.le64(pos, 4)); pos += 4;
    } else {
      if ((uint64_t)pos + 8 > len) return false;
      num = __longlong_as_double((long long)r.le64(pos, 8)); pos += 8;
    }
    if (role & BH_R_STATUS) P.status = u;
    if (role & BH_R_VALUE) P.value = num;
    if (role & BH_R_KEY) {
      uint32_t part = S.part[fn];
#pragma unroll
      for (int k = 0; k < BH_MAX_KEY_PARTS; k++)
        if (part == (uint32_t)k + 1) { P.ppos[k] = spos; P.plen[k] = slen; P.pval[k] = u; }
    }
  }
  return true;
}

template <class R>
__device__ __forceinline__ uint64_t event_key_hash(const DevSchema& S, const R& r, const Parsed& P) {
  if (S.status_part) return (uint64_t)P.status;
  if (S.identity_key) return (uint64_t)P.pval[0];
  uint64_t h = S.hash_seed ^ kHashInit;
#pragma unroll
  for (int k = 0; k < BH_MAX_KEY_PARTS; k++) {
    if ((uint32_t)k >= S.n_parts) break;
    if (S.part_is_len[k]) {
      uint32_t len = P.plen[k], pos = P.ppos[k];
      for (uint32_t i = 0; i < len; i += 8) h = absorb(h, r.le64(pos + i, len - i));
      h = absorb(h, len_word(len, S.part_field[k]));
    } else {
      h = absorb(h, varint_word(P.pval[k], S.part_field[k]));
    }
  }
  return fmix64(h);
}

// decode + drop rules of one event.  Returns BH_EV_*.
template <class R>
__device__ __forceinline__ uint32_t decode_and_check(const DevSchema& S, const R& r, uint32_t len, uint64_t& kh,
                                                     double& value, uint32_t& status) {
  Parsed P;
  if (!decode_event(S, r, len, P)) { kh = 0; value = 0.0; status = 0; return BH_EV_DROP_DECODE; }
  kh = event_key_hash(S, r, P);
  value = P.value; status = P.status;
  if (S.has_enum && !(P.status < 64 && ((S.enum_mask >> P.status) & 1ull))) return BH_EV_DROP_ENUM;
  if ((S.ops & (BH_OP_SUM | BH_OP_MIN | BH_OP_MAX | BH_OP_HIST)) && P.value != P.value) return BH_EV_DROP_NAN;
  return BH_EV_OK;
}

// =========================================================================================
// raw path: stage a warp's batch of messages in shared memory, one lane per message
// =========================================================================================
constexpr uint32_t kWarpBuf = 2560;                 // payload bytes staged per warp batch
constexpr uint32_t kWarpBufAlloc = kWarpBuf + 32;   // + alignment head + over-read slack
constexpr int kRawThreads = 768;

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// Handler interface shared by the fused and the decode-only kernel.
template <class Sink>
__device__ __forceinline__ void raw_warp_loop(const DevSchema& S, const uint8_t* __restrict__ payload,
                                              const uint32_t* __restrict__ offsets, uint32_t cs, uint32_t ce,
                                              uint8_t* wbuf, Sink& sink) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const uint32_t total16 = (offsets[ce] + 15u) & ~15u;   // readable bound of this chunk
  for (uint32_t e0 = cs + warp * 32; e0 < ce; e0 += nwarps * 32) {
    const uint32_t e = e0 + lane;
    const bool valid = e < ce;
    uint32_t o0 = 0, o1 = 0;
    if (valid) { o0 = offsets[e]; o1 = offsets[e + 1]; }
    const uint32_t cnt = min(32u, ce - e0);
    uint32_t done = 0;
    while (done < cnt) {
      const uint32_t b0 = __shfl_sync(0xFFFFFFFFu, o0, done);
      const uint32_t A = b0 & ~15u;
      const bool fits = valid && lane >= done && o1 >= o0 && o0 >= A && (o1 - A) <= kWarpBuf;
      uint32_t mask = __ballot_sync(0xFFFFFFFFu, fits) >> done;
      uint32_t m = mask == 0xFFFFFFFFu ? 32u : (uint32_t)__ffs(~mask) - 1u;   // leading run of lanes that fit
      uint32_t outcome; uint64_t kh; double value; uint32_t status;
      bool mine;
      if (m == 0) {                    // a single oversized (or inconsistent) message: parse from global
        m = 1;
        mine = lane == done;
        if (mine) {
          if (o1 < o0) { outcome = BH_EV_DROP_DECODE; kh = 0; value = 0.0; status = 0; }
          else { GmemRd r{payload + o0}; outcome = decode_and_check(S, r, o1 - o0, kh, value, status); }
        }
      } else {
        uint32_t bend = __shfl_sync(0xFFFFFFFFu, o1, done + m - 1);
        uint32_t B = min((bend + 15u) & ~15u, total16);
        const uint4* src = reinterpret_cast<const uint4*>(payload + A);
        uint4* dst = reinterpret_cast<uint4*>(wbuf);
        uint32_t nvec = (B - A) >> 4;
        for (uint32_t i = lane; i < nvec; i += 32) dst[i] = ldg_stream(src + i);
        __syncwarp();
        mine = lane >= done && lane < done + m;
        if (mine) { SmemRd r{wbuf + (o0 - A)}; outcome = decode_and_check(S, r, o1 - o0, kh, value, status); }
      }
      if (mine) sink.event(e, outcome, kh, value, status);
      __syncwarp();
      done += m;
    }
  }
}

struct AggSink {
  const DevSchema& S; const DevTables& T; const SmemTable& M;
  const uint64_t* ts; uint64_t wbase; LaneCounters& C;
  __device__ __forceinline__ void event(uint32_t e, uint32_t outcome, uint64_t kh, double value, uint32_t) {
    if (outcome != BH_EV_OK) { C.c[outcome]++; return; }
    uint64_t win = 0;
    if (S.has_window) win = div64(ts ? ts[e] : 0ull, S.wdiv);
    aggregate_event(S, T, M, kh, win, wbase, value, C);
  }
};

__global__ void __launch_bounds__(kRawThreads, 1)
k_process_raw(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T,
              const uint8_t* __restrict__ payload, const uint32_t* __restrict__ offsets,
              const uint64_t* __restrict__ ts, uint32_t n, uint32_t chunk, uint32_t ns) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t s_cnt[8];
  SmemTable M = smem_table_carve(smem, ns);
  uint8_t* wbuf = smem + (size_t)ns * kSmemSlotBytes + (size_t)(threadIdx.x >> 5) * kWarpBufAlloc;
  smem_table_clear(M);
  if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint32_t cs = min(n, blockIdx.x * chunk), ce = min(n, cs + chunk);
  LaneCounters C; C.clear();
  uint64_t wbase = 0;
  if (cs < ce) {
    if (S.has_window) { uint64_t w0 = div64(ts ? ts[cs] : 0ull, S.wdiv); wbase = w0 ? w0 - 1 : 0; }
    AggSink sink{S, T, M, ts, wbase, C};
    raw_warp_loop(S, payload, offsets, cs, ce, wbuf, sink);
  }
  __syncthreads();
  smem_table_flush(S, T, M, wbase);
  counters_flush(T, C, s_cnt);
}

struct ColumnSink {
  uint64_t* kh; double* val; int32_t* status; uint8_t* outcome;
  __device__ __forceinline__ void event(uint32_t e, uint32_t oc, uint64_t k, double v, uint32_t st) {
    kh[e] = k; val[e] = v; status[e] = (int32_t)st; outcome[e] = (uint8_t)oc;
  }
};

__global__ void __launch_bounds__(kRawThreads, 1)
k_decode(const __grid_constant__ DevSchema S, const uint8_t* __restrict__ payload,
         const uint32_t* __restrict__ offsets, uint32_t n, uint32_t chunk, uint64_t* kh, double* val,
         int32_t* status, uint8_t* outcome) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* wbuf = smem + (size_t)(threadIdx.x >> 5) * kWarpBufAlloc;
  const uint32_t cs = min(n, blockIdx.x * chunk), ce = min(n, cs + chunk);
  if (cs >= ce) return;
  ColumnSink sink{kh, val, status, outcome};
  raw_warp_loop(S, payload, offsets, cs, ce, wbuf, sink);
}

// =========================================================================================
// columnar path
// =========================================================================================
constexpr int kColThreads = 1024;

__global__ void __launch_bounds__(kColThreads, 1)
k_process_columns(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T,
                  const uint64_t* __restrict__ kh, const uint64_t* __restrict__ ts,
                  const double* __restrict__ val, uint64_t n, uint64_t chunk, uint32_t ns) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ uint32_t s_cnt[8];
  SmemTable M = smem_table_carve(smem, ns);
  smem_table_clear(M);
  if (threadIdx.x < 8) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t cs = min(n, (uint64_t)blockIdx.x * chunk), ce = min(n, cs + chunk);
  LaneCounters C; C.clear();
  uint64_t wbase = 0;
  if (cs < ce && S.has_window) { uint64_t w0 = div64(ts[cs], S.wdiv); wbase = w0 ? w0 - 1 : 0; }
  const bool need_value = (S.ops & (BH_OP_SUM | BH_OP_MIN | BH_OP_MAX | BH_OP_HIST)) != 0;
  for (uint64_t i = cs + threadIdx.x; i < ce; i += blockDim.x) {
    uint64_t k = kh[i];
    double v = val ? val[i] : 0.0;
    uint64_t win = S.has_window ? div64(ts[i], S.wdiv) : 0ull;
    if (need_value && v != v) { C.c[C_NAN]++; continue; }
    aggregate_event(S, T, M, k, win, wbase, v, C);
  }
  __syncthreads();
  smem_table_flush(S, T, M, wbase);
  counters_flush(T, C, s_cnt);
}

// =========================================================================================
// window close: export (compact by owner), merge, recycle
// =========================================================================================
__device__ __forceinline__ bool export_row(const DevTables& T, uint64_t base, uint64_t slot, uint64_t* key) {
  const Stat* s = T.stats + base + slot;
  if (s->count == 0) return false;
  *key = slot >= T.cap ? (slot == T.cap ? kLockKey : kEmptyKey) : T.keys[base + slot];
  return true;
}

// win == kNoWindow exports every open window.  Rows are visited ring slot by ring slot.
__device__ __forceinline__ bool export_pick(const DevTables& T, uint64_t win, uint64_t row, uint64_t* base,
                                            uint64_t* slot, uint64_t* row_win) {
  uint32_t r = (uint32_t)(row / T.stride);
  *slot = row - (uint64_t)r * T.stride;
  *base = (uint64_t)r * T.stride;
  unsigned long long w = T.ring_win[r];
  *row_win = w;
  return w != kNoWindow && (win == kNoWindow || w == win);
}

__global__ void k_export_count(const __grid_constant__ DevTables T, uint64_t win, uint32_t owner_shift,
                               unsigned long long* owner_counts) {
  uint64_t rows = (uint64_t)T.ring * T.stride, r0 = 0;
  if (win != kNoWindow) { r0 = (uint64_t)((uint32_t)win & (T.ring - 1)) * T.stride; rows = r0 + T.stride; }
  for (uint64_t row = r0 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; row < rows; row += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t base, slot, w, key;
    if (!export_pick(T, win, row, &base, &slot, &w)) continue;
    if (!export_row(T, base, slot, &key)) continue;
    uint32_t o = owner_shift >= 64 ? 0u : (uint32_t)(key >> owner_shift);
    atomicAdd(&owner_counts[o], 1ull);
  }
}

// cursor[o] = first output row of owner o: packed (exclusive prefix) or o * owner_stride
__global__ void k_export_prefix(uint32_t n_owners, const unsigned long long* owner_counts, unsigned long long* cursor,
                                uint64_t owner_stride) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned long long acc = 0;
    for (uint32_t o = 0; o < n_owners; o++) {
      cursor[o] = owner_stride ? (unsigned long long)o * owner_stride : acc;
      acc += owner_counts[o];
    }
  }
}

__global__ void k_export_scatter(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T, uint64_t win,
                                 uint32_t owner_shift, bh_record* out, uint32_t* hist_out, uint64_t cap,
                                 unsigned long long* cursor, uint64_t owner_stride) {
  uint64_t rows = (uint64_t)T.ring * T.stride, r0 = 0;
  if (win != kNoWindow) { r0 = (uint64_t)((uint32_t)win & (T.ring - 1)) * T.stride; rows = r0 + T.stride; }
  for (uint64_t row = r0 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; row < rows; row += (uint64_t)gridDim.x * blockDim.x) {
    uint64_t base, slot, w, key;
    if (!export_pick(T, win, row, &base, &slot, &w)) continue;
    if (!export_row(T, base, slot, &key)) continue;
    uint32_t o = owner_shift >= 64 ? 0u : (uint32_t)(key >> owner_shift);
    unsigned long long pos = atomicAdd(&cursor[o], 1ull);
    if (pos >= cap || (owner_stride && pos >= (unsigned long long)(o + 1) * owner_stride)) continue;
    const Stat s = T.stats[base + slot];
    bh_record rec;
    rec.key_hash = key; rec.window_id = w; rec.count = s.count;
    rec.sum = (S.ops & BH_OP_SUM) ? s.sum : 0.0;
    rec.min = (S.ops & BH_OP_MIN) && s.min_inv ? __longlong_as_double((long long)f64_unorder(~s.min_inv)) : 0.0;
    rec.max = (S.ops & BH_OP_MAX) && s.max_ord ? __longlong_as_double((long long)f64_unorder(s.max_ord)) : 0.0;
    out[pos] = rec;
    if (hist_out && T.hist) {
      const uint32_t* h = T.hist + (base + slot) * T.hist_buckets;
      for (uint32_t b = 0; b < T.hist_buckets; b++) hist_out[pos * T.hist_buckets + b] = h[b];
    }
  }
}

__global__ void k_merge(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T,
                        const bh_record* __restrict__ recs, const uint32_t* __restrict__ hist, uint64_t n) {
  uint32_t dropped[2] = {0, 0};
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    bh_record rec = recs[i];
    if (rec.count == 0) continue;
    uint32_t r;
    uint32_t gs = g_find_or_insert(T, rec.window_id, rec.key_hash, &r);
    if (gs >= kGslotWindow) { dropped[gs == kGslotFull ? 0 : 1] += 1; continue; }
    uint64_t gi = (uint64_t)r * T.stride + gs;
    uint64_t omin = f64_order((uint64_t)__double_as_longlong(rec.min));
    uint64_t omax = f64_order((uint64_t)__double_as_longlong(rec.max));
    g_update(T, S.ops, gi, rec.count, rec.sum, ~omin, omax);
    if (hist && T.hist)
      for (uint32_t b = 0; b < T.hist_buckets; b++) {
        uint32_t h = hist[i * T.hist_buckets + b];
        if (h) atomicAdd(&T.hist[gi * T.hist_buckets + b], h);
      }
  }
  if (dropped[0]) atomicAdd(&T.counters[C_TABLE_FULL], (unsigned long long)dropped[0]);
  if (dropped[1]) atomicAdd(&T.counters[C_WINDOW], (unsigned long long)dropped[1]);
}

__global__ void k_close_clear(const __grid_constant__ DevTables T, uint64_t win) {
  uint32_t r = (uint32_t)win & (T.ring - 1);
  if (T.ring_win[r] != win) return;
  uint64_t base = (uint64_t)r * T.stride;
  for (uint64_t slot = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; slot < T.stride; slot += (uint64_t)gridDim.x * blockDim.x) {
    T.keys[base + slot] = kEmptyKey;
    T.stats[base + slot] = Stat{0ull, 0.0, 0ull, 0ull};
    if (T.hist) for (uint32_t b = 0; b < T.hist_buckets; b++) T.hist[(base + slot) * T.hist_buckets + b] = 0;
  }
}
__global__ void k_close_release(const __grid_constant__ DevTables T, uint64_t win) {
  uint32_t r = (uint32_t)win & (T.ring - 1);
  if (blockIdx.x == 0 && threadIdx.x == 0 && T.ring_win[r] == win) T.ring_win[r] = kNoWindow;
}

// =========================================================================================
// launchers
// =========================================================================================
// once per device, with that device current (bh_create)
cudaError_t configure_kernels(const LaunchGeom& G) {
  int dyn = G.smem_optin - 1024;   // leave room for the kernels' static shared memory
  cudaError_t e = cudaFuncSetAttribute(k_process_raw, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(k_decode, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(k_process_columns, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
}

static uint32_t pick_smem_slots(const DevTables& T, const LaunchGeom& G, size_t other_bytes) {
  // front table: as many slots as the window table has, capped by what fits beside the warp buffers
  uint64_t want = T.cap < 64 ? 64 : T.cap;
  uint32_t ns = 2048;
  while ((uint64_t)ns > want) ns >>= 1;
  while (ns > 64 && (size_t)ns * kSmemSlotBytes + other_bytes + 1024 > (size_t)G.smem_optin) ns >>= 1;
  return ns;
}

cudaError_t launch_process_raw(const DevSchema& S, const DevTables& T, const LaunchGeom& G,
                               const uint8_t* payload, const uint32_t* offsets, const uint64_t* ts,
                               uint32_t n, cudaStream_t st, int* launches) {
  if (n == 0) return cudaSuccess;
  size_t wbytes = (size_t)(kRawThreads / 32) * kWarpBufAlloc;
  uint32_t ns = pick_smem_slots(T, G, wbytes);
  size_t smem = (size_t)ns * kSmemSlotBytes + wbytes;
  uint32_t per = kRawThreads;  // events per CTA pass
  uint32_t grid = (uint32_t)G.sms;
  uint32_t need = (n + per - 1) / per;
  if (need < grid) grid = need;
  uint32_t chunk = (n + grid - 1) / grid;
  chunk = (chunk + 31u) & ~31u;
  k_process_raw<<<grid, kRawThreads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  if (launches) (*launches)++;
  return cudaGetLastError();
}

cudaError_t launch_decode(const DevSchema& S, const LaunchGeom& G, const uint8_t* payload,
                          const uint32_t* offsets, uint32_t n, uint64_t* kh, double* val,
                          int32_t* status, uint8_t* outcome, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  size_t smem = (size_t)(kRawThreads / 32) * kWarpBufAlloc;
  uint32_t grid = (uint32_t)G.sms;
  uint32_t need = (n + kRawThreads - 1) / kRawThreads;
  if (need < grid) grid = need;
  uint32_t chunk = (n + grid - 1) / grid;
  chunk = (chunk + 31u) & ~31u;
  k_decode<<<grid, kRawThreads, smem, st>>>(S, payload, offsets, n, chunk, kh, val, status, outcome);
  return cudaGetLastError();
}

cudaError_t launch_process_columns(const DevSchema& S, const DevTables& T, const LaunchGeom& G,
                                   const uint64_t* kh, const uint64_t* ts, const double* val,
                                   uint64_t n, cudaStream_t st, int* launches) {
  if (n == 0) return cudaSuccess;
  uint32_t ns = pick_smem_slots(T, G, 0);
  size_t smem = (size_t)ns * kSmemSlotBytes;
  uint64_t grid = (uint64_t)G.sms;
  uint64_t need = (n + kColThreads - 1) / kColThreads;
  if (need < grid) grid = need;
  uint64_t chunk = (n + grid - 1) / grid;
  k_process_columns<<<(uint32_t)grid, kColThreads, smem, st>>>(S, T, kh, ts, val, n, chunk, ns);
  if (launches) (*launches)++;
  return cudaGetLastError();
}

static uint32_t owner_shift_of(uint32_t n_owners) {
  uint32_t lg = 0;
  while ((1u << lg) < n_owners) lg++;
  return 64u - lg;   // 64 -> single owner
}

cudaError_t launch_export(const DevSchema& S, const DevTables& T, uint64_t window_id, uint32_t n_owners,
                          bh_record* out, uint32_t* hist_out, uint64_t cap, uint64_t owner_stride,
                          unsigned long long* owner_counts, unsigned long long* owner_cursor,
                          cudaStream_t st) {
  cudaError_t e = cudaMemsetAsync(owner_counts, 0, sizeof(unsigned long long) * n_owners, st);
  if (e != cudaSuccess) return e;
  uint32_t shift = owner_shift_of(n_owners);
  uint64_t rows = window_id == kNoWindow ? (uint64_t)T.ring * T.stride : T.stride;
  uint64_t blocks = (rows + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_export_count<<<(uint32_t)blocks, 256, 0, st>>>(T, window_id, shift, owner_counts);
  k_export_prefix<<<1, 32, 0, st>>>(n_owners, owner_counts, owner_cursor, owner_stride);
  k_export_scatter<<<(uint32_t)blocks, 256, 0, st>>>(S, T, window_id, shift, out, hist_out, cap, owner_cursor, owner_stride);
  return cudaGetLastError();
}

cudaError_t launch_merge(const DevSchema& S, const DevTables& T, const bh_record* recs,
                         const uint32_t* hist, uint64_t n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  uint64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_merge<<<(uint32_t)blocks, 256, 0, st>>>(S, T, recs, hist, n);
  return cudaGetLastError();
}

cudaError_t launch_close_window(const DevTables& T, uint64_t window_id, cudaStream_t st) {
  uint64_t blocks = (T.stride + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_close_clear<<<(uint32_t)blocks, 256, 0, st>>>(T, window_id);
  k_close_release<<<1, 32, 0, st>>>(T, window_id);
  return cudaGetLastError();
}

cudaError_t launch_init_tables(const DevTables& T, cudaStream_t st) {
  cudaError_t e;
  size_t rows = (size_t)T.ring * T.stride;
  if ((e = cudaMemsetAsync(T.keys, 0xFF, rows * 8, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(T.stats, 0, rows * sizeof(Stat), st)) != cudaSuccess) return e;
  if (T.hist && (e = cudaMemsetAsync(T.hist, 0, rows * T.hist_buckets * 4, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(T.ring_win, 0xFF, (size_t)T.ring * 8, st)) != cudaSuccess) return e;
  return cudaMemsetAsync(T.counters, 0, 8 * 8, st);
}

}  // namespace bh
