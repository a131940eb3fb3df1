// bh_kernels.cu — sm_100a kernels of the beholder progress-event path.
//
//   k_process_raw      proto.decode + enumToString drop rule + counter update, fused
//                      (index.js:129-138): one lane per message, a warp's batch of 32 messages is
//                      staged through shared memory with coalesced 128-bit loads, a per-CTA
//                      shared-memory group-by table sits in front of the HBM tables.
//   k_process_columns  the counter update alone on pre-decoded columns (index.js:136-138).
//   k_decode           proto.decode alone -> columns (index.js:129-134).
//   k_export_* / k_merge / k_close_*   window close: compact, merge partials, recycle.
//
// Nothing here is a dense contraction: no tensor cores.  The kernels are bound by HBM bandwidth,
// by instruction issue (the wire-format walk) or by L2 atomics (large key sets).
#include <atomic>

#include "bh_kernels.cuh"

namespace bh {

// =========================================================================================
// global (HBM) tables
// =========================================================================================
constexpr uint32_t kGslotFull = 0xFFFFFFFFu;
constexpr uint32_t kGslotWindow = 0xFFFFFFFEu;

__device__ __forceinline__ uint32_t ilog2_u64(uint64_t v) { return 63u - (uint32_t)__clzll((long long)v); }

// Probe `n` slots of ring slot r's key column for `kh`, starting at `slot`, stepping inside the aligned block of
// wrap + 1 slots that holds `slot`.  Returns the slot found or claimed, or kGslotFull.
__device__ __forceinline__ uint32_t g_probe_from(const DevTables& T, uint32_t r, uint64_t kh, uint64_t slot, uint64_t wrap, uint64_t n) {
  unsigned long long* keys = T.keys + (uint64_t)r * T.stride;
  const uint64_t hi = slot & ~wrap;
  for (uint64_t i = 0; i < n; i++) {
    unsigned long long k = keys[slot];
    if (k == kh) return (uint32_t)slot;
    if (k == kEmptyKey) {
      unsigned long long old = atomicCAS(&keys[slot], (unsigned long long)kEmptyKey, (unsigned long long)kh);
      if (old == kEmptyKey || old == kh) return (uint32_t)slot;
    }
    slot = hi | ((slot + 1) & wrap);
  }
  return kGslotFull;
}
// Linear probing wraps at the end of the table -- in partitioned mode at the end of the key's PARTITION SLICE
// (cap >> part_lg slots, the slots whose index starts with the partition's bits), so that a key never sits outside
// its partition's slice and the fold kernel, which holds one slice in shared memory, owns every slot a tuple of
// its partition can reach.  A slice fills up like a small table of its own: keys / P of them land in it, +- a few
// tens at 2048 slots, so the table's load factor decides, as before.
__device__ __forceinline__ uint64_t g_probe_wrap(const DevTables& T) {
  const uint64_t slice = T.part_lg ? T.cap >> T.part_lg : 0;
  return (slice ? slice : T.cap) - 1;
}

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
  if (kh >= kLockKey) return (uint32_t)T.cap + (uint32_t)(kh & 1);   // side slots for the sentinel-valued keys
  const uint32_t shift = 64u - ilog2_u64(T.cap);
  const uint64_t wrap = g_probe_wrap(T);
  return g_probe_from(T, r, kh, (kh * kHashMul) >> shift, wrap, wrap + 1);
}

// LOOK = true: the min / max atomics are skipped when a plain read already shows a value at least as good -- right for
// events going to the table one by one (a key's row settles after its first few events and then costs two L2 reads,
// no atomics).  LOOK = false: four fire-and-forget reductions, nothing to wait for -- right where a row is met once
// per launch (the shared-table flush, the merge of exported rows): there the reads were two more round trips per row.
template <bool LOOK = true>
__device__ __forceinline__ void g_update(const DevTables& T, uint32_t ops, uint64_t gi, unsigned long long cnt,
                                         double sum, unsigned long long min_inv, unsigned long long max_ord) {
  Stat* s = T.stats + gi;
  atomicAdd(&s->count, cnt);
  if (ops & BH_OP_SUM) atomicAdd(&s->sum, sum);
  if ((ops & BH_OP_MIN) && (!LOOK || min_inv > s->min_inv)) atomicMax(&s->min_inv, min_inv);
  if ((ops & BH_OP_MAX) && (!LOOK || max_ord > s->max_ord)) atomicMax(&s->max_ord, max_ord);
}
// shared-memory fetch-and-add by shared-window address (a file-scope __shared__ counter reached through a lambda's
// reference capture is otherwise treated as a generic address: a slower, long-scoreboard atomic)
__device__ __forceinline__ uint32_t atoms_add(uint32_t* p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;" : "=r"(old) : "r"((uint32_t)__cvta_generic_to_shared(p)), "r"(v) : "memory");
  return old;
}

// =========================================================================================
// shared-memory front table (per CTA)
//
// Two layouts, chosen by the declared type of the reduced field:
//   INTV  (int32 / sint32 / bool `progress`, the reference schema): sum as a 64-bit integer in two
//         32-bit words, min / max as int32 -> every update is a NATIVE 32-bit shared atomic
//         (ATOMS.ADD / ATOMS.MIN / ATOMS.MAX), no CAS loops, no divergence.  32 B per slot.
//   F64   (double / float / uint32 values, and the columnar path): sum f64, min / max as
//         order-preserving u64; shared memory has no native 64-bit add/max, the compiler's
//         ATOMS.CAST.SPIN loops are used.  40 B per slot.
// =========================================================================================
// Dynamic shared memory of the aggregating kernels and their 8 outcome counters, named at file scope so
// that every access is a plain shared-space access (32-bit address, no generic pointer kept in registers).
extern __shared__ __align__(16) uint8_t g_smem[];
__shared__ uint32_t g_cnt[8];

// Histogram rows of the entries that turn out hot (kHotThreshold events in one launch) are kept in shared memory:
// without them every event of a hot key is an L2 reduction on the same few sectors of its HBM row, and those
// serialise -- measured with Zipf keys over 1 M keys: the fused kernel took 858 us per 12.5 M events against 440 us
// with uniform keys, all of it same-address histogram reductions.
constexpr uint32_t kHotRows = 32, kHotThreshold = 16;
constexpr uint32_t kHotNone = 0xFFFF, kHotClaim = 0xFFFE, kHotFull = 0xFFFD;
constexpr uint32_t kHotBytes = kHotRows * BH_HIST_MAX * 4 + kHotRows * 4 + 16;
struct SmemTable {
  // Only the slot count is kept; the column arrays are found from it and the shared base at
  // each use: tag u32[ns] | key u64[ns] | INTV: sum_lo u32, sum_hi i32, min i32, max i32 / F64: sum f64,
  // min_inv u64, max_ord u64 | cnt u32[ns] | gslot u32[ns] (histograms only)
  // tag = window byte << 24 | the top 24 bits of the key hash (the slot index comes from a multiply of the
  // folded hash, so the two are as good as independent);
  // 0xFFFFFFFF = empty slot (the window byte is always < 255).
  uint32_t ns;              // power of two
  uint32_t shift;           // 32 - log2(ns)
  uint32_t hist;            // gslot and hot-row columns present
  uint32_t hot_on;          // the hot histogram rows follow the columns (not in the fold kernel)
  __device__ __forceinline__ uint32_t* tag() const { return (uint32_t*)g_smem; }
  __device__ __forceinline__ unsigned long long* key() const { return (unsigned long long*)(g_smem + (size_t)ns * 4); }
  template <bool INTV> __device__ __forceinline__ uint32_t* cnt() const { return (uint32_t*)(g_smem + (size_t)ns * (INTV ? 28 : 36)); }
  template <bool INTV> __device__ __forceinline__ uint32_t* gslot() const { return hist ? (uint32_t*)(g_smem + (size_t)ns * (INTV ? 32 : 40)) : nullptr; }
  // histograms only: hotrow u16[ns] (which hot row the entry owns, or kHotNone / kHotClaim / kHotFull), and after the
  // columns (16-byte aligned) the hot rows themselves: u32 [kHotRows][BH_HIST_MAX] | owner slot u32[kHotRows] | next u32
  template <bool INTV> __device__ __forceinline__ unsigned short* hotrow() const { return (unsigned short*)(g_smem + (size_t)ns * (INTV ? 36 : 44)); }
  template <bool INTV> __device__ __forceinline__ uint32_t* hot() const { return (uint32_t*)(g_smem + (((size_t)ns * (INTV ? 38 : 46) + 15) & ~(size_t)15)); }
  template <bool INTV> __device__ __forceinline__ uint32_t* hot_owner() const { return hot<INTV>() + kHotRows * BH_HIST_MAX; }
  template <bool INTV> __device__ __forceinline__ uint32_t* hot_next() const { return hot_owner<INTV>() + kHotRows; }
  // INTV value columns
  __device__ __forceinline__ uint32_t* sum_lo() const { return (uint32_t*)(g_smem + (size_t)ns * 12); }
  __device__ __forceinline__ int32_t* sum_hi() const { return (int32_t*)(g_smem + (size_t)ns * 16); }
  __device__ __forceinline__ int32_t* imin() const { return (int32_t*)(g_smem + (size_t)ns * 20); }
  __device__ __forceinline__ int32_t* imax() const { return (int32_t*)(g_smem + (size_t)ns * 24); }
  // F64 value columns
  __device__ __forceinline__ double* fsum() const { return (double*)(g_smem + (size_t)ns * 12); }
  __device__ __forceinline__ unsigned long long* fmin_inv() const { return (unsigned long long*)(g_smem + (size_t)ns * 20); }
  __device__ __forceinline__ unsigned long long* fmax_ord() const { return (unsigned long long*)(g_smem + (size_t)ns * 28); }
};
constexpr int kSmemProbes = 16;
constexpr int kSmemProbesPart = 6;   // scattering kernels: a miss only costs a 16-byte append, so give up sooner
constexpr uint32_t kTagEmpty = 0xFFFFFFFFu;
__device__ __forceinline__ uint32_t smem_tag(uint64_t kh, uint32_t wrel) { return (wrel << 24) | (uint32_t)(kh >> 40); }
__host__ __device__ constexpr uint32_t smem_slot_bytes(bool intv, bool hist) {
  return 4 + 8 + (intv ? 16 : 24) + 4 + (hist ? 4 + 2 : 0);
}
// the whole table: columns (16-byte aligned end) + the hot histogram rows
__host__ __device__ constexpr size_t smem_table_bytes(uint32_t ns, bool intv, bool hist, bool hot = true) {
  return (((size_t)ns * smem_slot_bytes(intv, hist) + 15) & ~(size_t)15) + (hist && hot ? kHotBytes : 0);
}

template <bool INTV>
__device__ __forceinline__ SmemTable smem_table_carve(uint32_t ns, bool hist, bool hot = true) {
  SmemTable M;
  M.ns = ns; M.hist = hist; M.hot_on = hist && hot;
  M.shift = 32u - (31u - (uint32_t)__clz((int)ns));
  return M;
}

template <bool INTV>
__device__ __forceinline__ void smem_table_clear(const SmemTable& M) {
  for (uint32_t i = threadIdx.x; i < M.ns; i += blockDim.x) {
    M.tag()[i] = kTagEmpty; M.key()[i] = kEmptyKey; M.cnt<INTV>()[i] = 0;
    if (M.hist) { M.gslot<INTV>()[i] = 0; M.hotrow<INTV>()[i] = (unsigned short)kHotNone; }
    if (INTV) { M.sum_lo()[i] = 0; M.sum_hi()[i] = 0; M.imin()[i] = 0x7FFFFFFF; M.imax()[i] = (int32_t)0x80000000; }
    else { M.fsum()[i] = 0.0; M.fmin_inv()[i] = 0; M.fmax_ord()[i] = 0; }
  }
  if (M.hot_on) for (uint32_t i = threadIdx.x; i < kHotRows * BH_HIST_MAX + kHotRows + 1; i += blockDim.x) M.hot<INTV>()[i] = 0;
}

// the entry `slot` has seen kHotThreshold events: give it a hot row if one is left (one winner per entry)
template <bool INTV>
__device__ __noinline__ void smem_hot_promote(const SmemTable M, int slot) {
  unsigned short* hp = M.hotrow<INTV>() + slot;
  if (atomicCAS(hp, (unsigned short)kHotNone, (unsigned short)kHotClaim) != (unsigned short)kHotNone) return;
  const uint32_t r = atomicAdd(M.hot_next<INTV>(), 1u);
  if (r < kHotRows) M.hot_owner<INTV>()[r] = (uint32_t)slot;
  *(volatile unsigned short*)hp = (unsigned short)(r < kHotRows ? r : kHotFull);
}

// returns the shared slot, or -1 when the event has to go to the HBM table directly.
// The look-up walks the 32-bit TAG column only: one load and two compares per probe; the 64-bit key is read
// once, when a tag matches.  Claiming a slot is ONE compare-and-swap on the tag (empty -> tag); the key
// follows with a plain store and readers that match the tag wait for it (all-ones = not written yet).  No
// fence is needed: only the one location is polled.  A reader that saw a torn or foreign key just walks on
// and may claim a second slot for the same (key, window): harmless, the flush merges both into the HBM
// entry.  With histograms on, the HBM slot must be visible before the key, which costs a block-level fence
// on that (rare) path only.
// Slow half: claim a slot (or find the key someone else is inserting right now), starting at `slot`.
// Kept out of line: the CAS and the wait loop make the compiler add yields and convergence barriers
// that must not sit in the hot look-up loop.
template <bool INTV>
__device__ __noinline__ int smem_insert(const SmemTable M, const DevTables& T, uint64_t kh, uint32_t tag,
                                        uint64_t win, uint32_t slot, int probes_left) {
  volatile unsigned long long* vkey = M.key();
  volatile uint32_t* vtag = M.tag();
#pragma unroll 1
  for (int p = 0; p < probes_left; p++) {
    uint32_t t = vtag[slot];
    if (t == kTagEmpty) {
      t = atomicCAS(&M.tag()[slot], kTagEmpty, tag);
      if (t == kTagEmpty) {     // claimed
        if (M.hist) { uint32_t r; M.gslot<INTV>()[slot] = g_find_or_insert(T, win, kh, &r); __threadfence_block(); }
        vkey[slot] = kh;
        return (int)slot;
      }
    }
    if (t == tag) {
      unsigned long long k = vkey[slot];
      while (k == kEmptyKey) k = vkey[slot];
      if (k == kh) return (int)slot;
    }
    slot = (slot + 1) & (M.ns - 1);
  }
  return -1;
}

// Hot half: read-only look-up (plain shared loads, no atomics, no waiting).
template <bool INTV, bool PART_OK>
// Called by every lane of `mask` together (there is a warp barrier inside); `eligible` = this lane's event may
// use the shared table at all.
__device__ __forceinline__ int smem_find_or_insert(const SmemTable& M, const DevTables& T, uint64_t kh,
                                                   uint32_t wrel, uint64_t win, uint32_t mask, bool eligible) {
  const uint32_t x = (uint32_t)kh ^ (uint32_t)(kh >> 32) ^ (wrel * 0x85EBCA6Bu);
  uint32_t slot = (x * 0x9E3779B1u) >> M.shift;
  const uint32_t tag = smem_tag(kh, wrel), smask = M.ns - 1;
  const uint32_t* tags = M.tag();
  uint32_t t = eligible ? tags[slot] : tag;
  int left = PART_OK && T.part_lg ? kSmemProbesPart : kSmemProbes;
#pragma unroll 1
  while (t != tag && t != kTagEmpty && --left != 0) {
    slot = (slot + 1) & smask;
    t = tags[slot];
  }
  // one exit, and the lanes that left the loop at different iterations are brought together again before the
  // key compare (measured: that compare ran with 3 of 32 lanes active, once per group of lanes)
  __syncwarp(mask);
  if (!eligible || left == 0) return -1;
  if (t == tag && M.key()[slot] == kh) return (int)slot;
  // empty slot, key not stored yet, or another key with the same tag: the out-of-line half sorts it out
  return smem_insert<INTV>(M, T, kh, tag, win, slot, left);
}

// Outcome counters of a CTA.  Only the count of aggregated events lives in a register; drops, spills
// and slow decodes are rare and go straight to a shared-memory array (keeps 7 registers out of the loop).
struct LaneCounters {
  uint32_t ok;
  __device__ __forceinline__ void bump(uint32_t which) { atomicAdd(&g_cnt[which], 1u); }   // g_cnt: zeroed by the kernel
};

// Partitioned mode (DevTables::part_*): append the event to this CTA's region of its partition.  Returns false
// when it has to take the direct path after all: region full (a skewed key distribution; the direct path is
// where hot keys are cheap, they stay in L2) or the window's ring slot is held by another window.
__device__ __noinline__ bool part_scatter(const DevTables& T, uint64_t kh, uint64_t win, uint64_t vbits) {
  const uint32_t r = (uint32_t)win & (T.ring - 1);
  unsigned long long w = T.ring_win[r];
  if (w != win) {                                       // claim the ring slot now: the drain resolves r -> window id
    if (w != kNoWindow) return false;
    const unsigned long long old = atomicCAS(&T.ring_win[r], (unsigned long long)kNoWindow, (unsigned long long)win);
    if (old != kNoWindow && old != win) return false;
  }
  const uint32_t sh = 64u - T.part_lg;
  const uint64_t hs = kh * kHashMul;
  const uint32_t p = (uint32_t)(hs >> sh);
  uint32_t* cur = (uint32_t*)(g_smem + T.part_smem_off);
  const uint32_t pos = atomicAdd(&cur[p], 1u);
  if (pos >= T.part_cap) return false;
  const uint64_t a = (hs & ((1ull << sh) - 1)) | ((uint64_t)r << sh);
  T.part_tuples[((uint64_t)blockIdx.x << T.part_lg | p) * T.part_cap + pos] =
      make_uint4((uint32_t)a, (uint32_t)(a >> 32), (uint32_t)vbits, (uint32_t)(vbits >> 32));
  return true;
}
// cursors of this CTA's regions: global <-> shared at the two ends of a scattering kernel
__device__ __forceinline__ void part_cursors_load(const DevTables& T) {
  if (!T.part_lg) return;
  uint32_t* cur = (uint32_t*)(g_smem + T.part_smem_off);
  for (uint32_t i = threadIdx.x; i < (1u << T.part_lg); i += blockDim.x) cur[i] = T.part_fill[((uint64_t)blockIdx.x << T.part_lg) + i];
}
__device__ __forceinline__ void part_cursors_store(const DevTables& T) {
  if (!T.part_lg) return;            // (uniform over the CTA)
  __syncthreads();                   // the flush just before still appends tuples
  const uint32_t* cur = (const uint32_t*)(g_smem + T.part_smem_off);
  for (uint32_t i = threadIdx.x; i < (1u << T.part_lg); i += blockDim.x) T.part_fill[((uint64_t)blockIdx.x << T.part_lg) + i] = min(cur[i], T.part_cap);
}

// one event's value into the shared-memory entry `slot`
template <bool INTV>
__device__ __forceinline__ void smem_apply(const SmemTable& M, int slot, uint32_t ops, double v, int32_t iv) {
  atomicAdd(&M.cnt<INTV>()[slot], 1u);
  if (INTV) {
    if (ops & BH_OP_SUM) {
      const uint32_t add = (uint32_t)iv;
      const uint32_t old = atomicAdd(M.sum_lo() + slot, add);
      const int32_t hi = (iv >> 31) + (int32_t)(old + add < old);   // sign extension + carry
      if (hi) atomicAdd(M.sum_hi() + slot, hi);
    }
    if (ops & BH_OP_MIN) atomicMin(M.imin() + slot, iv);
    if (ops & BH_OP_MAX) atomicMax(M.imax() + slot, iv);
  } else {
    const uint64_t ord = f64_order((uint64_t)__double_as_longlong(v));
    if (ops & BH_OP_SUM) atomicAdd(M.fsum() + slot, v);
    // both columns only ever grow: a plain read that already shows a value at least as large makes the
    // (CAS-loop) atomic unnecessary, and after a key's first few events that is the usual case
    if ((ops & BH_OP_MIN) && (unsigned long long)~ord > M.fmin_inv()[slot]) atomicMax(M.fmin_inv() + slot, (unsigned long long)~ord);
    if ((ops & BH_OP_MAX) && (unsigned long long)ord > M.fmax_ord()[slot]) atomicMax(M.fmax_ord() + slot, (unsigned long long)ord);
  }
}

// One event into the tables.  `wbase` is the CTA's first window minus one.
// INTV: `iv` is the value; otherwise `v`.  `mask` = the lanes that call it together.
// PART_OK = false (the drain kernel): a miss goes to the HBM table directly even in partitioned mode.
// HOT = false (the fold kernel): no hot histogram rows -- its tuples are what the scattering kernels' tables did NOT
// absorb, a partition's keys share them evenly, and the promotion path cost it 55 % (measured: 1.00 -> 1.55 ms per fold).
template <bool INTV, bool PART_OK = true, bool HOT = PART_OK>
__device__ __forceinline__ void aggregate_event(const DevSchema& S, const DevTables& T, const SmemTable& M,
                                                uint64_t kh, uint64_t win, uint64_t wbase, double v, int32_t iv,
                                                LaneCounters& C, uint32_t mask) {
  const uint32_t ops = S.ops;
  const uint64_t wr = win - wbase;
  const int slot = smem_find_or_insert<INTV, PART_OK>(M, T, kh, (uint32_t)wr, win, mask, kh < kLockKey && wr < 255ull);
  __syncwarp(mask);   // (the insert path is divergent) bring the lanes back before the updates
  const uint32_t in_table = __ballot_sync(mask, slot >= 0);
  if (slot >= 0) {
    bool refused = false;
    if (ops & BH_OP_HIST) {
      if (INTV) v = (double)iv;
      const uint32_t gs = M.gslot<INTV>()[slot];
      if (gs >= kGslotWindow) {
        C.bump(gs == kGslotFull ? C_TABLE_FULL : C_WINDOW);
        refused = true;
      } else {
        const uint32_t b = hist_bucket((uint64_t)__double_as_longlong(v), T.hist_buckets, S.hist_e0);
        const uint32_t hr = HOT ? M.hotrow<INTV>()[slot] : kHotFull;
        if (HOT && hr < kHotRows) {
          atomicAdd(&M.hot<INTV>()[hr * BH_HIST_MAX + b], 1u);
        } else {
          uint64_t gi = (uint64_t)((uint32_t)win & (T.ring - 1)) * T.stride + gs;
          atomicAdd(&T.hist[gi * T.hist_buckets + b], 1u);
          if (HOT && hr == kHotNone && M.cnt<INTV>()[slot] >= kHotThreshold) smem_hot_promote<INTV>(M, slot);
        }
      }
      if (HOT) __syncwarp(in_table);   // hot / not hot / the call: together again before the (CAS-loop) updates
    }
    if (!refused) { smem_apply<INTV>(M, slot, ops, v, iv); C.ok++; }
    return;
  }
  if (PART_OK && T.part_lg && kh < kLockKey &&
      part_scatter(T, kh, win, (uint64_t)__double_as_longlong(INTV ? (double)iv : v))) {   // always the f64 bits
    C.ok++; C.bump(C_SPILL);
    return;
  }
  // spill: straight to the HBM table
  if (INTV) v = (double)iv;
  const uint64_t bits = (uint64_t)__double_as_longlong(v);
  const uint64_t ord = f64_order(bits);
  uint32_t r;
  uint32_t gs = g_find_or_insert(T, win, kh, &r);
  if (gs >= kGslotWindow) { C.bump(gs == kGslotFull ? C_TABLE_FULL : C_WINDOW); return; }
  uint64_t gi = (uint64_t)r * T.stride + gs;
  g_update(T, ops, gi, 1ull, v, ~ord, ord);
  if (ops & BH_OP_HIST) atomicAdd(&T.hist[gi * T.hist_buckets + hist_bucket(bits, T.hist_buckets, S.hist_e0)], 1u);
  C.ok++; C.bump(C_SPILL);
}

// CTA end: every shared entry goes to its HBM slot.  Events of an entry whose window or key cannot
// be placed move from the "ok" counter to the drop counter that explains them.
// PART_OK (the scattering kernels in partitioned mode): an entry that absorbed a single event is that event -- its
// sum is its value -- so it is logged as a tuple like the events that missed the table, instead of costing a
// random HBM update now.  With large uniform key sets that is nearly every entry of the table.  (Not with
// histograms: there the entry's HBM slot was claimed and its bucket counted when the event arrived.)
template <bool INTV, bool PART_OK = false>
__device__ __forceinline__ void smem_table_flush(const DevSchema& S, const DevTables& T, const SmemTable& M, uint64_t wbase) {
  // every CTA hashes a key to the same shared slot: rotate the scan by CTA so that the CTAs do not all
  // hit the same HBM entry at the same moment (same-address L2 atomics serialise)
  const uint32_t rot = (blockIdx.x * 0x9E3779B1u) >> M.shift;
  for (uint32_t j = threadIdx.x; j < M.ns; j += blockDim.x) {
    const uint32_t i = (j + rot) & (M.ns - 1);
    unsigned long long k = M.key()[i];
    uint32_t c = M.cnt<INTV>()[i];
    if (k == kEmptyKey || c == 0) continue;
    uint64_t win = wbase + (M.tag()[i] >> 24);
    uint32_t r = (uint32_t)win & (T.ring - 1);
    if (PART_OK && T.part_lg && !M.hist && c == 1 && k < kLockKey) {
      // the event's value, from whichever column the configured ops keep (the max/min columns hold its exact
      // bits, -0.0 included; the sum column holds 0.0 + v)
      uint64_t vb;
      if (INTV) vb = (uint64_t)__double_as_longlong((double)((S.ops & BH_OP_SUM) ? (int32_t)M.sum_lo()[i] : (S.ops & BH_OP_MIN) ? M.imin()[i] : (S.ops & BH_OP_MAX) ? M.imax()[i] : 0));
      else vb = (S.ops & BH_OP_MAX) ? f64_unorder(M.fmax_ord()[i]) : (S.ops & BH_OP_MIN) ? f64_unorder(~M.fmin_inv()[i]) : (uint64_t)__double_as_longlong(M.fsum()[i]);
      if (part_scatter(T, k, win, vb)) continue;
    }
    uint32_t gs = M.hist ? M.gslot<INTV>()[i] : g_find_or_insert(T, win, k, &r);
    if (gs >= kGslotWindow) {
      if (!M.hist) {   // (with histograms on, the event was already refused when it arrived)
        atomicAdd(&T.counters[gs == kGslotFull ? C_TABLE_FULL : C_WINDOW], (unsigned long long)c);
        atomicAdd(&T.counters[C_OK], 0ull - (unsigned long long)c);
      }
      continue;
    }
    double sum; unsigned long long mn_inv, mx_ord;
    if (INTV) {
      long long s64 = (long long)(((unsigned long long)(uint32_t)M.sum_hi()[i] << 32) | M.sum_lo()[i]);
      sum = (double)s64;
      mn_inv = ~f64_order((uint64_t)__double_as_longlong((double)M.imin()[i]));
      mx_ord = f64_order((uint64_t)__double_as_longlong((double)M.imax()[i]));
    } else {
      sum = M.fsum()[i]; mn_inv = M.fmin_inv()[i]; mx_ord = M.fmax_ord()[i];
    }
    g_update<false>(T, S.ops, (uint64_t)r * T.stride + gs, c, sum, mn_inv, mx_ord);
  }
  if (M.hot_on) {   // the hot rows into their HBM rows (the owner's HBM slot was checked before it could turn hot)
    const uint32_t n = min(*M.hot_next<INTV>(), kHotRows);
    for (uint32_t i = threadIdx.x; i < n * BH_HIST_MAX; i += blockDim.x) {
      const uint32_t v = M.hot<INTV>()[i], b = i & (BH_HIST_MAX - 1);
      if (!v || b >= T.hist_buckets) continue;
      const uint32_t slot = M.hot_owner<INTV>()[i / BH_HIST_MAX];
      const uint64_t win = wbase + (M.tag()[slot] >> 24);
      const uint64_t gi = (uint64_t)((uint32_t)win & (T.ring - 1)) * T.stride + M.gslot<INTV>()[slot];
      atomicAdd(&T.hist[gi * T.hist_buckets + b], v);
    }
  }
}

__device__ __forceinline__ void counters_flush(const DevTables& T, LaneCounters& C) {
  uint32_t v = __reduce_add_sync(0xFFFFFFFFu, C.ok);
  if ((threadIdx.x & 31) == 0 && v) atomicAdd(&g_cnt[C_OK], v);
  __syncthreads();
  if (threadIdx.x < 8 && g_cnt[threadIdx.x]) atomicAdd(&T.counters[threadIdx.x], (unsigned long long)g_cnt[threadIdx.x]);
}

// shared-memory access by 32-bit shared-window address, for the staged payload bytes
// -DBH_BOUNDS_CHECK (debug builds; compute-sanitizer is not available on the test pool): every access to a
// warp's staging buffer is checked against that buffer, every byte the global-memory reader touches
// against the batch's readable range; a violation prints once and traps.
#ifdef BH_BOUNDS_CHECK
}  // namespace bh
#include <stdio.h>
namespace bh {
__shared__ uint32_t g_dbg_first;                 // shared-window address where the warp buffers start
__device__ const uint8_t* g_dbg_glo; __device__ const uint8_t* g_dbg_ghi;
__device__ __forceinline__ void dbg_smem(uint32_t a, uint32_t bytes) {
  const uint32_t lo = ((g_dbg_first + 67u) / 68u + (threadIdx.x >> 5) * 37u) * 68u, hi = lo + 37u * 68u;
  if (a < lo || a + bytes > hi) { printf("warp-buffer access %u+%u outside [%u,%u) block %u thread %u\n", a, bytes, lo, hi, blockIdx.x, threadIdx.x); __trap(); }
}
__device__ __forceinline__ void dbg_gmem(const void* p, uint32_t bytes) {
  const uint8_t* q = (const uint8_t*)p;
  if (q < g_dbg_glo || q + bytes > g_dbg_ghi) { printf("global read %p+%u outside [%p,%p)\n", p, bytes, g_dbg_glo, g_dbg_ghi); __trap(); }
}
#define BH_DBG_SMEM(a, n) dbg_smem(a, n)
#define BH_DBG_GMEM(p, n) dbg_gmem(p, n)
#else
#define BH_DBG_SMEM(a, n) ((void)0)
#define BH_DBG_GMEM(p, n) ((void)0)
#endif
__device__ __forceinline__ uint32_t lds8(uint32_t a) { BH_DBG_SMEM(a, 1); uint32_t v; asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32(uint32_t a) { BH_DBG_SMEM(a, 4); uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { BH_DBG_SMEM(a, 4); asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }

// =========================================================================================
// protobufjs 6.8.8 reader semantics on the device (what proto.decode does, index.js:129)
// =========================================================================================
// A warp's batch is staged in shared memory SKEWED: 4 bytes of padding after every 64, so that 32
// lanes walking 64-byte messages in lock step hit 32 different banks (stride 17 words) instead of 2.
// shift + multiply-add: two instructions (the plain C form is turned into shift, mask, add)
__device__ __forceinline__ uint32_t skew(uint32_t a) {
  uint32_t r;
  asm("{\n\t.reg .u32 t;\n\tshr.u32 t, %1, 6;\n\tmad.lo.u32 %0, t, 4, %1;\n\t}" : "=r"(r) : "r"(a));
  return r;
}

struct SmemRd {           // event bytes staged (skewed) in shared memory
  // Positions are in the warp buffer's "origin space": the buffer's shared-window address is a multiple
  // of 68, so a logical origin Q (a multiple of 64) exists with skew(Q + x) = address of logical byte x,
  // and a byte access is skew() plus the load, with no base register to add.
  uint32_t off;           // origin-space position of this event's first byte
  __device__ __forceinline__ uint32_t b(uint32_t i) const { return lds8(skew(off + i)); }
  __device__ __forceinline__ uint32_t ab(uint32_t lp) const { return lds8(skew(lp)); }   // absolute position
  __device__ __forceinline__ uint32_t w32(uint32_t a) const { return lds32(skew(a)); }   // aligned word
  __device__ __forceinline__ uint64_t le64(uint32_t i, uint32_t nvalid) const {
    uint32_t a = off + i;
    uint32_t sh = (a & 3u) * 8u;
    a &= ~3u;
    uint32_t w0 = w32(a), w1 = w32(a + 4), w2 = w32(a + 8);
    uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
    uint64_t w = ((uint64_t)hi << 32) | lo;
    return nvalid >= 8 ? w : (w & ((1ull << (8 * nvalid)) - 1));
  }
};
struct GmemRd {           // slow path: event parsed straight from global memory
  const uint8_t* p;
  __device__ __forceinline__ uint32_t b(uint32_t i) const { BH_DBG_GMEM(p + i, 1); return __ldg(p + i); }
  // up to 8 bytes from i, as one or two aligned 64-bit loads.  Only words that hold a wanted byte are
  // read, so the last access ends at most 7 bytes after the message: inside the buffer's 16-byte round-up.
  __device__ __forceinline__ uint64_t le64(uint32_t i, uint32_t nvalid) const {
    if (nvalid == 0) return 0ull;
    if (nvalid > 8) nvalid = 8;
    const uintptr_t q = (uintptr_t)(p + i);
    const unsigned long long* a = (const unsigned long long*)(q & ~(uintptr_t)7);
    const uint32_t sh = (uint32_t)(q & 7u) * 8u;
    BH_DBG_GMEM(a, 8);
    uint64_t w = __ldg(a) >> sh;
    if (sh + 8u * nvalid > 64u) { BH_DBG_GMEM(a + 1, 8); w |= (uint64_t)__ldg(a + 1) << (64u - sh); }
    return nvalid >= 8 ? w : (w & ((1ull << (8 * nvalid)) - 1));
  }
};

// Reader.uint32(): 5 value bytes; a longer varint is assumed to be 10 bytes; bytes past the end read
// as `undefined` (no value bits, "continue").
template <class R>
__device__ __noinline__ bool rd_uint32(const R& r, uint32_t& pos, uint32_t len, uint32_t& out) {
  uint32_t v = 0;
#pragma unroll 1
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
  uint32_t uval;      // VALUE field as decoded 32-bit integer (varint kinds)
  double value;       // VALUE field as a JS number (not filled by the fast path for integer kinds)
  uint32_t ppos[BH_MAX_KEY_PARTS], plen[BH_MAX_KEY_PARTS], pval[BH_MAX_KEY_PARTS];
  __device__ __forceinline__ void clear() {
    status = 0; uval = 0; value = 0.0;
#pragma unroll
    for (int k = 0; k < BH_MAX_KEY_PARTS; k++) { ppos[k] = 0; plen[k] = 0; pval[k] = 0; }
  }
};

// One known field read by its DECLARED type (protobufjs' generated decoder ignores the wire type),
// result stored by role.  string() on a Node Buffer clamps to the end, bytes() throws.
template <class R>
__device__ __forceinline__ bool decode_known_field(const R& r, uint32_t f, uint32_t& pos, uint32_t len, Parsed& P) {
  const uint32_t kind = f & 0xFFu, role = (f >> 8) & 0xFFu;
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
  if (role & BH_R_VALUE) { P.value = num; P.uval = u; }
  if (role & BH_R_KEY) {
    const uint32_t part = f >> 16;
#pragma unroll
    for (int k = 0; k < BH_MAX_KEY_PARTS; k++)
      if (part == (uint32_t)k + 1) { P.ppos[k] = spos; P.plen[k] = slen; P.pval[k] = u; }
  }
  return true;
}

// Type.decode(), complete.  Returns false where proto.decode would throw.
template <class R>
__device__ __noinline__ bool decode_event(const DevSchema& S, const R& r, uint32_t len, Parsed& P) {
  P.clear();
  uint32_t pos = 0;
  while (pos < len) {
    uint32_t t;
    if (!rd_uint32(r, pos, len, t)) return false;
    const uint32_t fn = t >> 3;
    const uint32_t f = fn < 32 ? S.fld[fn] : 0;
    if ((f & 0xFFu) == 0) { if (!rd_skip_type(r, pos, len, t & 7u)) return false; continue; }
    if (!decode_known_field(r, f, pos, len, P)) return false;
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
  return fmix64(h) & S.hash_mask;
}

// ---- fast path ---------------------------------------------------------------------------------
// Messages whose tags, lengths and varints are single bytes (two for the length of an unknown
// length-delimited field) — what well-formed api.TelemetryProgress events look like.  The three hot
// fields (first length-delimited key part, status, value) are recognised by comparing the field
// number with launch constants; every other field goes through the schema table.  Anything unusual
// makes it return false and the event is re-parsed by decode_event, so the outcome is decode_event's
// in every case; the fast path only has to agree with it on the inputs it accepts.
__device__ __forceinline__ bool decode_fast(const DevSchema& S, const SmemRd& r, uint32_t len, Parsed& P) {
  P.clear();
  uint32_t lp = r.off;
  const uint32_t end = r.off + len;
  while (lp < end) {
    if (lp + 2 > end) return false;
    const uint32_t t = r.ab(lp), b1 = r.ab(lp + 1);
    lp += 2;
    if (t & 0x80u) return false;
    const uint32_t fn = t >> 3;
    if (fn == S.fast_key0) {                 // string key part 0 (mediaId)
      if (b1 & 0x80u) return false;
      uint32_t e2 = lp + b1;
      e2 = e2 > end ? end : e2;              // BufferReader.string() clamps
      P.ppos[0] = lp - r.off; P.plen[0] = e2 - lp; lp = e2;
      continue;
    }
    if (fn == S.fast_status) {               // int32 / enum status, possibly also a key part
      if (b1 & 0x80u) return false;
      P.status = b1;
      if (S.fast_status_part == 1) P.pval[0] = b1;
      if (S.fast_status_part == 2) P.pval[1] = b1;
      if (S.fast_status_part == 3) P.pval[2] = b1;
      continue;
    }
    if (fn == S.fast_value) {                // int32 value as a one-byte varint, or a double
      if (S.fast_value_f64) {
        if (lp + 7 > end) return false;
        P.value = __longlong_as_double((long long)r.le64(lp - 1 - r.off, 8)); lp += 7;
      } else {
        if (b1 & 0x80u) return false;
        P.uval = b1;
      }
      continue;
    }
    const uint32_t f = S.fld[fn];
    if ((f & 0xFFu) == 0) {                  // unknown field: Reader.skipType
      const uint32_t wt = t & 7u;
      if (wt == 2) {
        uint32_t n = b1;
        if (b1 & 0x80u) {                    // two-byte length (pads of 128 B .. 16 KB)
          if (lp >= end) return false;
          const uint32_t b2 = r.ab(lp);
          if (b2 & 0x80u) return false;
          n = (b1 & 0x7Fu) | (b2 << 7);
          lp += 1;
        }
        if (lp + n > end) return false;      // decode_event will report the throw
        lp += n;
      } else if (wt == 0) {
        if (b1 & 0x80u) return false;
      } else {
        return false;
      }
      continue;
    }
    // other known fields (host, further key parts, ...): declared-type read through the table
    if (b1 & 0x80u) return false;
    const uint32_t kind = f & 0xFFu;
    if (kind > BH_K_BOOL) return false;
    if (kind <= BH_K_BYTES) {
      uint32_t e2 = lp + b1;
      if (e2 > end) { if (kind == BH_K_BYTES) return false; e2 = end; }
      const uint32_t part = f >> 16;
#pragma unroll
      for (int k = 0; k < BH_MAX_KEY_PARTS; k++)
        if (part == (uint32_t)k + 1) { P.ppos[k] = lp - r.off; P.plen[k] = e2 - lp; }
      lp = e2;
    } else {
      uint32_t pos = lp - 1 - r.off;         // back to the byte after the tag
      if (!decode_known_field(r, f, pos, len, P)) return false;
      lp = r.off + pos;
    }
  }
  return true;
}

// ---- straight path ------------------------------------------------------------------------------
// What encoders actually emit for api.TelemetryProgress: fields in number order, each at most once —
// [key part 0][unknown length-delimited pad]?[status]?[value]? then role-less strings (host) — all with
// single-byte tags and lengths.  Recognised with a handful of compares against launch constants, no
// loop over a schema table.  Anything else (other order, repeats, other wire types, long varints, a
// field with a role in the tail) returns false and the message goes to decode_fast / decode_event, so
// the outcome is always the general decoder's.  Bytes just past the message may be read (they are
// inside the warp buffer); every use is guarded by a position check.
// `mask` = the lanes that call it together.  The optional sections make lanes of one warp diverge (some
// messages carry a pad or a host field, some do not); the explicit __syncwarp(mask) after each section
// brings them back together — without it the compiler reconverges only at the end and the rest of the
// decoder plus the key hash run once per divergent group (measured: 40 % slower on mixed streams).
// Hence no early return before the last section: failures are carried in `ok`.
// Compile-time view of the reference's own message and grouping: api.TelemetryProgress { mediaId = 1 string,
// status = 2 enum, progress = 3 int32, host = 4 string } (index.js:46,131; numbers ASSUMED, see schema.py),
// grouped by (mediaId, status) in tumbling windows with the enum drop on.  Everything the decoder asks the
// schema is a constant here, so the straight-line decoder, the key hash and the drop rules lose their table
// look-ups and tests; what varies at run time (ops, enum mask, hash seed, window divisor) still comes from the
// DevSchema.  The host launches the kernel built on this view only when progress_schema_matches() says that
// the run-time schema is exactly this one; every other schema runs the same code on the DevSchema itself.
// REF = false: the windowed grouping above.  REF = true: the reference's literal counter (BH_MODE_REFERENCE:
// key = status, no key parts, no windows; mediaId is just another string to skip).
// F64 = true: the same message with `progress` declared double (BASELINE.json config 3's value type): field 3 is a
// 64-bit wire value (tag 0x19) and the f64 shared table is used.
template <bool REF, bool F64 = false>
struct ProgressSchemaT {
  const DevSchema& dyn;
  __device__ __forceinline__ operator const DevSchema&() const { return dyn; }
  static constexpr uint32_t kFld1 = REF ? (uint32_t)BH_K_STRING : (BH_K_STRING | (BH_R_KEY << 8) | (1u << 16));
  static constexpr uint32_t kFld2 = REF ? (BH_K_INT32 | (BH_R_STATUS << 8)) : (BH_K_INT32 | ((BH_R_KEY | BH_R_STATUS) << 8) | (2u << 16));
  static constexpr uint32_t kFld3 = (F64 ? BH_K_DOUBLE : BH_K_INT32) | (BH_R_VALUE << 8);
  static constexpr uint32_t kFld4 = BH_K_STRING;
  __host__ __device__ static constexpr uint32_t field(uint32_t n) {
    return n == 1 ? kFld1 : n == 2 ? kFld2 : n == 3 ? kFld3 : n == 4 ? kFld4 : 0u;
  }
  __host__ __device__ static constexpr uint32_t part_len(int k) { return !REF && k == 0 ? 1u : 0u; }
  __host__ __device__ static constexpr uint32_t part_fld(int k) { return REF ? 0u : k == 0 ? 1u : k == 1 ? 2u : 0u; }
  static constexpr uint32_t n_parts = REF ? 0 : 2, identity_key = 0, status_part = REF ? 1 : 0, has_enum = 1;
  static constexpr uint32_t has_window = REF ? 0 : 1, mode = REF ? BH_MODE_REFERENCE : BH_MODE_WINDOWED;
  static constexpr uint32_t int_values = F64 ? 0 : 1, fast_key0 = REF ? 0xFFFFFFFFu : 1u, fast_status = 2;
  static constexpr uint32_t fast_status_part = REF ? 0 : 2, fast_value = 3;
  static constexpr uint32_t fast_value_f64 = F64 ? 1 : 0, tag_key0 = REF ? 0xFFFFFFFFu : 0x0Au, tag_status = 0x10, tag_value = F64 ? 0x19 : 0x18;
  static constexpr uint32_t skip_str_mask = REF ? ((1u << 1) | (1u << 4)) : (1u << 4), skip_unk_mask = 0xFFFFu & ~0x1Eu;
  // run-time items
  uint32_t ops, pay_lo, pay_hi; uint64_t enum_mask, hash_seed, hash_mask, window_ms; Div64 wdiv;
  __device__ __forceinline__ explicit ProgressSchemaT(const DevSchema& S)
      : dyn(S), ops(S.ops), pay_lo(S.pay_lo), pay_hi(S.pay_hi), enum_mask(S.enum_mask), hash_seed(S.hash_seed), hash_mask(S.hash_mask),
        window_ms(S.window_ms), wdiv(S.wdiv) {}
};
typedef ProgressSchemaT<false> ProgressSchema;

template <bool REF, bool F64 = false>
static bool schema_matches(const DevSchema& S) {
  typedef ProgressSchemaT<REF, F64> P;
  for (uint32_t n = 0; n < 32; n++) if (S.fld[n] != P::field(n)) return false;
  for (int k = 0; k < BH_MAX_KEY_PARTS; k++) if (S.part_is_len[k] != P::part_len(k) || S.part_field[k] != P::part_fld(k)) return false;
  return S.n_parts == P::n_parts && S.identity_key == P::identity_key && S.status_part == P::status_part &&
         S.has_enum == P::has_enum && S.has_window == P::has_window && S.int_values == P::int_values &&
         S.fast_key0 == P::fast_key0 && S.fast_status == P::fast_status && S.fast_status_part == P::fast_status_part &&
         S.fast_value == P::fast_value && S.fast_value_f64 == P::fast_value_f64 && S.tag_key0 == P::tag_key0 &&
         S.tag_status == P::tag_status && S.tag_value == P::tag_value && S.skip_str_mask == P::skip_str_mask &&
         S.skip_unk_mask == P::skip_unk_mask && S.mode == P::mode;
}
// 0 = no compiled-in view fits, 1 = windowed (mediaId, status) grouping, 2 = reference mode, 3 = windowed with a double `progress`
int progress_schema_variant(const DevSchema& S) {
  return schema_matches<false>(S) ? 1 : schema_matches<true>(S) ? 2 : schema_matches<false, true>(S) ? 3 : 0;
}
bool progress_schema_matches(const DevSchema& S) { return progress_schema_variant(S) != 0; }

template <class SC>
__device__ __forceinline__ bool decode_canon(const SC& S, const SmemRd& r, uint32_t len, Parsed& P) {
  uint32_t lp = r.off;
  const uint32_t end = lp + len;
  uint32_t t = r.ab(lp), n = r.ab(lp + 1);
  bool ok = true;
  uint32_t kpos = 0, klen = 0, status = 0, uval = 0;
  double value = 0.0;
  // one optional role-less string (clamped to the end) or unknown length-delimited field (strict)
  auto plain = [&]() {
    const bool str = (S.skip_str_mask >> (t >> 3)) & 1u, unk = ((S.skip_unk_mask >> (t >> 3)) & 1u) && (t & 7u) == 2u;
    const bool here = ok && lp < end && (str || unk);
    const uint32_t b2 = r.ab(lp + 2);
    const bool two = (n & 0x80u) != 0;
    const bool good_len = lp + 2 <= end && (!two || (lp + 2 < end && !(b2 & 0x80u)));
    const uint32_t ln = two ? ((n & 0x7Fu) | (b2 << 7)) : n;
    const uint32_t after = lp + (two ? 3u : 2u) + ln;
    const bool good = good_len && !(after > end && unk);     // string() clamps, skip() throws
    ok = ok && (!here || good);
    if (here && good) lp = min(after, end);
  };
  if (S.tag_key0 != 0xFFFFFFFFu) {                           // (uniform) the first field is key part 0 ...
    const bool here = lp < end && t == S.tag_key0;
    const bool good = lp + 2 <= end && !(n & 0x80u);
    ok = !here || good;
    if (here && good) {
      const uint32_t e2 = min(lp + 2 + n, end);              // BufferReader.string() clamps
      kpos = lp + 2 - r.off; klen = e2 - (lp + 2); lp = e2;
    }
  } else {                                                   // ... or, with no string key (reference mode), one more plain field
    plain();
  }
  t = r.ab(lp); n = r.ab(lp + 1);
  plain();
  t = r.ab(lp); n = r.ab(lp + 1);
  {
    const bool here = ok && lp < end && t == S.tag_status;
    const bool good = lp + 2 <= end && !(n & 0x80u);
    ok = ok && (!here || good);
    if (here && good) { status = n; lp += 2; }
    t = r.ab(lp); n = r.ab(lp + 1);
  }
  {
    const bool here = ok && lp < end && t == S.tag_value;
    if (S.fast_value_f64) {
      const bool good = lp + 9 <= end;
      ok = ok && (!here || good);
      if (here && good) { value = __longlong_as_double((long long)r.le64(lp + 1 - r.off, 8)); lp += 9; }
    } else {
      const bool good = lp + 2 <= end && !(n & 0x80u);
      ok = ok && (!here || good);
      if (here && good) { uval = n; lp += 2; }
    }
    t = r.ab(lp); n = r.ab(lp + 1);
  }
  plain();
  ok = ok && lp == end;
  P.clear();
  P.ppos[0] = kpos; P.plen[0] = klen; P.status = status; P.uval = uval; P.value = value;
  if (S.fast_status_part == 1) P.pval[0] = status;
  if (S.fast_status_part == 2) P.pval[1] = status;
  if (S.fast_status_part == 3) P.pval[2] = status;
  return ok;
}

template <class SC>
__device__ __forceinline__ bool decode_straight(const SC& S, const SmemRd& r, uint32_t len, Parsed& P, uint32_t mask) {
  uint32_t lp = r.off;
  const uint32_t end = lp + len;
  uint32_t t = r.ab(lp), n = r.ab(lp + 1);
  uint32_t kpos = 0, klen = 0, status = 0, uval = 0;
  double value = 0.0;
  bool ok = true;
  // role-less strings (clamped) and unknown length-delimited fields (strict): skipped wherever they stand
  auto skip_plain = [&]() {
    while (ok && lp < end) {
      // one-byte tags only: the masks have no bit above field 15, so a tag with the continuation bit matches neither
      const bool str = (S.skip_str_mask >> (t >> 3)) & 1u, unk = ((S.skip_unk_mask >> (t >> 3)) & 1u) && (t & 7u) == 2u;
      if (!(str || unk)) break;                            // a field with a role (or anything unusual): caller decides
      if (lp + 2 > end) { ok = false; break; }
      lp += 2;
      if (n & 0x80u) {                                     // two-byte length
        const uint32_t b2 = r.ab(lp);
        if (lp >= end || (b2 & 0x80u)) { ok = false; break; }
        n = (n & 0x7Fu) | (b2 << 7);
        lp += 1;
      }
      if (lp + n > end) { if (unk) { ok = false; break; } lp = end; }   // string() clamps, skip() throws
      else lp += n;
      t = r.ab(lp); n = r.ab(lp + 1);
    }
  };
  if (lp < end && t == S.tag_key0) {
    if (lp + 2 > end || (n & 0x80u)) ok = false;
    else {
      lp += 2;
      const uint32_t e2 = min(lp + n, end);                // BufferReader.string() clamps
      kpos = lp - r.off; klen = e2 - lp; lp = e2;
      t = r.ab(lp); n = r.ab(lp + 1);
    }
  }
  __syncwarp(mask);
  skip_plain();
  __syncwarp(mask);
  // status and integer value: one- or two-byte varints (longer ones are left to the slower tiers)
  auto varint12 = [&](uint32_t& out) {
    if (lp + 2 > end) { ok = false; return; }
    if (!(n & 0x80u)) { out = n; lp += 2; return; }
    const uint32_t b2 = r.ab(lp + 2);
    if (lp + 3 > end || (b2 & 0x80u)) { ok = false; return; }
    out = (n & 0x7Fu) | (b2 << 7); lp += 3;
  };
  if (ok && lp < end && t == S.tag_status) {
    varint12(status);
    t = r.ab(lp); n = r.ab(lp + 1);
  }
  if (ok && lp < end && t == S.tag_value) {
    if (S.fast_value_f64) {
      if (lp + 9 > end) ok = false;
      else { value = __longlong_as_double((long long)r.le64(lp + 1 - r.off, 8)); lp += 9; }
    } else {
      varint12(uval);
    }
    t = r.ab(lp); n = r.ab(lp + 1);
  }
  __syncwarp(mask);
  skip_plain();
  __syncwarp(mask);
  P.clear();
  P.ppos[0] = kpos; P.plen[0] = klen; P.status = status; P.uval = uval; P.value = value;
  if (S.fast_status_part == 1) P.pval[0] = status;
  if (S.fast_status_part == 2) P.pval[1] = status;
  if (S.fast_status_part == 3) P.pval[2] = status;
  return ok && lp == end;                                  // something with a role is left: not the canonical order
}

__device__ __forceinline__ uint64_t hash_span_smem(const SmemRd& r, uint32_t lp, uint32_t n, uint32_t field, uint64_t h) {
  uint32_t a = lp & ~3u;
  const uint32_t sh = (lp & 3u) * 8u;
  uint32_t w0 = r.w32(a);
  a += 4;
  uint32_t i = 0;
#pragma unroll 1
  for (; i + 8 <= n; i += 8) {
    const uint32_t w1 = r.w32(a), w2 = r.w32(a + 4);
    const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
    h = absorb(h, ((uint64_t)hi << 32) | lo);
    w0 = w2; a += 8;
  }
  if (i < n) {
    const uint32_t rem = n - i;
    const uint32_t w1 = r.w32(a), w2 = r.w32(a + 4);   // may read past the event: masked below, inside the buffer's slack
    const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
    const uint64_t w = ((uint64_t)hi << 32) | lo;
    h = absorb(h, w & ((1ull << (8 * rem)) - 1));
  }
  return absorb(h, len_word(n, field));
}

template <class SC>
__device__ __forceinline__ uint64_t event_key_hash_smem(const SC& S, const SmemRd& r, const Parsed& P) {
  if (S.status_part) return (uint64_t)P.status;
  if (S.identity_key) return (uint64_t)P.pval[0];
  uint64_t h = S.hash_seed ^ kHashInit;
#pragma unroll
  for (int k = 0; k < BH_MAX_KEY_PARTS; k++) {
    if (k >= (int)S.n_parts) break;
    if (S.part_len(k)) h = hash_span_smem(r, r.off + P.ppos[k], P.plen[k], S.part_fld(k), h);
    else h = absorb(h, varint_word(P.pval[k], S.part_fld(k)));
  }
  return fmix64(h) & S.hash_mask;
}

// drop rules after a successful decode.  Returns BH_EV_*.  `ival` is the value for integer value
// kinds, `value` for the others.
template <class SC>
__device__ __forceinline__ uint32_t finish_event(const SC& S, const Parsed& P, double& value, int32_t& ival, uint32_t& status) {
  value = P.value; ival = (int32_t)P.uval; status = P.status;
  if (S.has_enum && !(P.status < 64 && ((S.enum_mask >> P.status) & 1ull))) return BH_EV_DROP_ENUM;
  if (!S.int_values && (S.ops & (BH_OP_SUM | BH_OP_MIN | BH_OP_MAX | BH_OP_HIST)) && P.value != P.value) return BH_EV_DROP_NAN;
  return BH_EV_OK;
}

constexpr uint32_t kSlowDecode = 0x100u;   // outcome flag: the message left the straight path (diagnostic counter)
__device__ __noinline__ bool decode_slow(const DevSchema& S, const SmemRd r, uint32_t len, Parsed& P) {
  return decode_fast(S, r, len, P) || decode_event(S, r, len, P);
}

// ---- block tier -----------------------------------------------------------------------------------
// A message that starts on a 4-byte boundary of the warp buffer and lies inside one 64-byte block of it is
// CONTIGUOUS in shared memory (the skew only moves whole blocks), so its bytes are at base + i: no skew
// arithmetic per access, aligned 32-bit loads, and the key bytes (at offset 2, after the one-byte tag and length)
// come out of consecutive words with a constant 16-bit funnel shift.  Fixed-size 64-byte events on a 64-byte grid
// (BASELINE.json config 2) are exactly that.  Accepted here: [string key or role-less string][one unknown
// length-delimited field with a one-byte length]?[status][value][one role-less string or unknown
// length-delimited field up to the end]? -- one-byte tags, lengths and varints, a 32-bit integer value, the key
// made of that string and, optionally, the status.  That is a subset of what decode_canon accepts, with the same
// field values by construction; anything else returns false with nothing decided and takes the tiers below.
template <class SC>
__device__ __forceinline__ bool decode_block(const SC& S, const SmemRd& r, uint32_t len, uint64_t& kh, uint32_t& status, uint32_t& uval,
                                             double& value, uint32_t& kloc) {
  // (f64 values were tried here too -- three more word pairs for the 8 value bytes -- and made the f64 kernels
  // slower, 0.443 -> 0.472 ms fused: they are at 72 registers already)
  if (S.fast_value_f64 || S.tag_status > 0x7Fu || S.tag_value > 0x7Fu || S.identity_key) return false;
  const bool key_is_string = S.tag_key0 <= 0x7Fu;
  if (key_is_string ? !(S.n_parts == 1 || (S.n_parts == 2 && S.fast_status_part == 2)) : !S.status_part) return false;
  const uint32_t x0 = r.off;
  if ((x0 & 3u) || (x0 & 63u) + len > 64u || len < 6u) return false;
  const uint32_t m = skew(x0);                               // bytes of the message: m + i
  auto rd4 = [&](uint32_t pos) {                              // 4 bytes at any offset: two aligned words, one funnel shift
    const uint32_t a = m + (pos & ~3u);
    return __funnelshift_r(lds32(a), lds32(a + 4), (pos & 3u) * 8u);
  };
  const uint32_t w0 = lds32(m);
  const uint32_t t0 = w0 & 0xFFu, klen = (w0 >> 8) & 0xFFu;
  bool ok = key_is_string ? t0 == S.tag_key0 : (t0 < 0x80u && ((S.skip_str_mask >> (t0 >> 3)) & 1u));
  ok = ok && klen < 0x80u;
  uint32_t d = 2u + klen;
  ok = ok && d + 4u <= len;
  if (!ok) return false;
  uint32_t x = rd4(d);
  {
    const uint32_t t = x & 0xFFu;                            // one unknown length-delimited field (the pad), strict
    if (t < 0x80u && ((S.skip_unk_mask >> (t >> 3)) & 1u) && (t & 7u) == 2u) {
      const uint32_t pl = (x >> 8) & 0xFFu;
      d += 2u + pl;
      if (pl >= 0x80u || d + 4u > len) return false;
      x = rd4(d);
    }
  }
  if ((x & 0x80FF80FFu) != (S.tag_status | (S.tag_value << 16))) return false;
  status = (x >> 8) & 0x7Fu; uval = x >> 24; value = 0.0;
  d += 4u;
  if (d != len) {                                            // one trailing field that ends the message
    if (d + 2u > len) return false;
    const uint32_t y = rd4(d), t = y & 0xFFu, n = (y >> 8) & 0xFFu;
    if (t >= 0x80u || n >= 0x80u) return false;
    const bool str = (S.skip_str_mask >> (t >> 3)) & 1u, unk = ((S.skip_unk_mask >> (t >> 3)) & 1u) && (t & 7u) == 2u;
    const uint32_t after = d + 2u + n;
    if (!(str ? after >= len : (unk && after == len))) return false;      // string() clamps to the end, skip() is strict
  }
  kloc = 2u | (klen << 16);
  if (!key_is_string) { kh = (uint64_t)status; return true; }            // reference mode: the key is the status itself
  // key hash (bh_common.cuh): 8-byte words of the key bytes, then the length word, then the status word
  uint64_t h = S.hash_seed ^ kHashInit;
  if (klen == 36u) {                                         // a UUID string: no loop, no masks
    uint32_t w[10];
    w[0] = w0;
#pragma unroll
    for (int j = 1; j < 10; j++) w[j] = lds32(m + 4u * j);
#pragma unroll
    for (int j = 0; j < 4; j++)
      h = absorb(h, ((uint64_t)__funnelshift_r(w[2 * j + 1], w[2 * j + 2], 16) << 32) | __funnelshift_r(w[2 * j], w[2 * j + 1], 16));
    h = absorb(h, (uint64_t)__funnelshift_r(w[8], w[9], 16));          // bytes 34..37: the last four key bytes
    h = absorb(h, len_word(36u, S.part_fld(0)));
    if (S.n_parts == 2) h = absorb(h, varint_word(status, S.part_fld(1)));
    kh = fmix64(h) & S.hash_mask;
    return true;
  }
  uint32_t wa = w0, a = m + 4u, i = 0;
#pragma unroll 1
  for (; i + 8u <= klen; i += 8u) {
    const uint32_t wb = lds32(a), wc = lds32(a + 4u);
    h = absorb(h, ((uint64_t)__funnelshift_r(wb, wc, 16) << 32) | __funnelshift_r(wa, wb, 16));
    wa = wc; a += 8u;
  }
  if (i < klen) {
    const uint32_t wb = lds32(a), wc = lds32(a + 4u);      // may read past the key: masked
    const uint64_t w = ((uint64_t)__funnelshift_r(wb, wc, 16) << 32) | __funnelshift_r(wa, wb, 16);
    h = absorb(h, w & ((1ull << (8u * (klen - i))) - 1ull));
  }
  h = absorb(h, len_word(klen, S.part_fld(0)));
  if (S.n_parts == 2) h = absorb(h, varint_word(status, S.part_fld(1)));
  kh = fmix64(h) & S.hash_mask;
  return true;
}

// kloc (for the key check of BH_F_VERIFY_KEYS; dead code elsewhere): position | length << 16 of the first key part's
// bytes inside the message
template <class SC>
__device__ __forceinline__ uint32_t decode_and_check_smem(const SC& S, const SmemRd& r, uint32_t len, uint64_t& kh,
                                                          double& value, int32_t& ival, uint32_t& status, uint32_t mask,
                                                          uint32_t& kloc) {
  Parsed P;
  uint32_t slow = 0;
  kloc = 0;
#ifndef BH_NO_BLOCK_TIER
  {
    uint32_t st = 0, uv = 0;
    double dv = 0.0;
    const bool blk = decode_block(S, r, len, kh, st, uv, dv, kloc);
    mask = __ballot_sync(mask, !blk);                      // the lanes that go on to the tiers below, together
    if (blk) {
      P.status = st; P.uval = uv; P.value = dv;
      return finish_event(S, P, value, ival, status);
    }
  }
#endif
#ifdef BH_NO_TIER0
  const bool canon = false;
#else
  const bool canon = decode_canon(S, r, len, P);
#endif
  const uint32_t rest = __ballot_sync(mask, !canon);      // the lanes that go on to the next tier, together
  if (!canon && !decode_straight(S, r, len, P, rest)) {
    Parsed Q;   // a separate object: P must not escape into the out-of-line slow path (it would live in local memory)
    if (!decode_slow(S, r, len, Q)) { kh = 0; value = 0.0; ival = 0; status = 0; return BH_EV_DROP_DECODE | kSlowDecode; }
    P = Q;
    slow = kSlowDecode;
  }
  kh = event_key_hash_smem(S, r, P);
  kloc = P.ppos[0] | (P.plen[0] << 16);
  return finish_event(S, P, value, ival, status) | slow;
}

// Out of line (rare): results come back through one small object, so that the caller's own
// kh/value/ival/status never have their address taken and stay in registers.
struct EventOut { uint64_t kh; double value; int32_t ival; uint32_t status; uint32_t kloc; };
__device__ __noinline__ uint32_t decode_and_check_gmem(const DevSchema& S, const uint8_t* p, uint32_t len, EventOut& out) {
  Parsed P;
  GmemRd r{p};
  out.kh = 0; out.value = 0.0; out.ival = 0; out.status = 0; out.kloc = 0;
  if (!decode_event(S, r, len, P)) return BH_EV_DROP_DECODE;
  out.kh = event_key_hash(S, r, P);
  out.kloc = P.ppos[0] | (P.plen[0] << 16);
  return finish_event(S, P, out.value, out.ival, out.status);
}

// =========================================================================================
// raw path: stage a warp's batch of messages in shared memory, one lane per message
// =========================================================================================
constexpr uint32_t kWarpBuf = 2304;                                  // payload bytes staged per warp batch (32 x 72)
constexpr uint32_t kWarpChunks = (kWarpBuf / 16 + 1 + 31) / 32;      // 16-byte chunks per lane (5)
constexpr uint32_t kWarpBufAlloc = ((kWarpBuf + kWarpBuf / 16 + 64 + 67) / 68) * 68;   // + skew padding + over-read slack, multiple of 68
static_assert(kWarpBufAlloc == 37u * 68u, "BH_BOUNDS_CHECK hard-codes the buffer size");
constexpr uint32_t kWarpBufSlack = 68;   // the first warp buffer is moved up to a multiple of 68

// origin (see SmemRd) of the calling warp's buffer; `first` = shared-window address where warp buffers may start
__device__ __forceinline__ uint32_t warp_buffer_origin(uint32_t first) {
  const uint32_t m = (first + 67u) / 68u + (threadIdx.x >> 5) * (kWarpBufAlloc / 68u);
  return m << 6;
}
constexpr int kRawThreads = 1024;      // decode-only kernel and the integer-table fused kernel
constexpr int kRawThreadsF64 = 768;    // fused kernel with the (larger) f64 table: fewer warp buffers, 4096 slots still fit

// streaming 128-bit load; NOT volatile so that the loads of one batch are issued back to back.  The payload is
// read once: its lines are marked evict-first in L2 so that 950 MB of input per launch do not push out what is
// re-used there -- the HBM table lines and, in partitioned mode, the half-written sectors at the ends of the
// tuple-log regions (measured at the C3 shape: see DESIGN).
__device__ __forceinline__ unsigned long long l2_evict_first_policy() {
  unsigned long long pol;
  asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
  uint4 v;
  const unsigned long long pol = l2_evict_first_policy();
  asm("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
  return v;
}
// the same load pinned to its place in the instruction stream (software pipelining: a plain asm may be sunk to its
// first use, which turns "load the next chunk, then work on this one" back into "wait for every chunk")
__device__ __forceinline__ uint4 ldg_stream_here(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
constexpr uint32_t kPrefetchAhead = 4;   // batches ahead whose cache lines are pulled into L2
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" :: "l"(p)); }

// coalesced copy of payload[A, B) (both multiples of 16) into the skewed warp buffer with origin `wb`
__device__ __forceinline__ void stage_range(const uint8_t* __restrict__ payload, uint32_t A, uint32_t B, uint32_t wb,
                                            uint32_t lane) {
  const uint4* src = reinterpret_cast<const uint4*>(payload + A);
  if (B <= A) return;
  const uint32_t nvec = (B - A) >> 4;
  // lane l copies chunks l, l+32, ...: 32 chunks = 512 B = 8 blocks of 64, so the skewed destination
  // advances by 512 + 8*4 = 544 bytes per step (a 16-byte chunk never straddles a 64-byte block)
  const uint32_t d0 = skew(wb + (lane << 4));
  // Every load is issued, past-the-end chunks re-reading the last one (same cache line as a neighbour's,
  // so no extra traffic): a load under `if` leaves registers that are undefined on one path, and the
  // compiler then carries them around the batch loop with moves and spills (20 registers' worth).
  uint4 v[kWarpChunks];
#pragma unroll
  for (uint32_t k = 0; k < kWarpChunks; k++) v[k] = ldg_stream(src + min(lane + 32 * k, nvec - 1));
#pragma unroll
  for (uint32_t k = 0; k < kWarpChunks; k++) {
    if (lane + 32 * k < nvec) {
      const uint32_t d = d0 + 544u * k;
      sts32(d, v[k].x); sts32(d + 4, v[k].y); sts32(d + 8, v[k].z); sts32(d + 12, v[k].w);
    }
  }
}

// Handler interface shared by the fused and the decode-only kernel.  Each warp takes batches of 32
// consecutive messages, copies their bytes (one contiguous range) into its shared buffer with
// coalesced 128-bit loads, then every lane parses its own message.  The offsets of the NEXT batch are
// loaded before the current one is parsed, and the lines the batch after that will touch are
// prefetched into L2, so a warp rarely waits for HBM.
template <bool BIG, class SC, class Sink>
__device__ __forceinline__ void raw_warp_loop(const SC& S, const uint8_t* __restrict__ payload,
                                              const uint32_t* __restrict__ offsets, uint32_t cs, uint32_t ce,
                                              uint32_t wb, Sink& sink) {
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const uint32_t total16 = min((offsets[ce] + 15u) & ~15u, S.pay_hi);   // readable bound of this chunk
  const uint32_t step = nwarps * 32;
  uint32_t e0 = cs + warp * 32;
  uint32_t n0 = 0, n1 = 0;                                // offsets of the batch about to be processed
  if (e0 + lane < ce) { n0 = offsets[e0 + lane]; n1 = offsets[e0 + lane + 1]; }
  for (; e0 < ce; e0 += step) {
    const uint32_t e = e0 + lane;
    const bool valid = e < ce;
    const uint32_t o0 = n0, o1 = n1;
    {                                                     // next batch's offsets; lines of the one after
      const uint32_t en = e + step;
      n0 = 0; n1 = 0;
      if (en < ce) { n0 = offsets[en]; n1 = offsets[en + 1]; }
      const uint32_t e2 = e0 + kPrefetchAhead * step;
      if (e2 < ce) {
        if (lane == 0) prefetch_l2(offsets + e2);
        sink.prefetch(e2, lane);
        // payload a few batches ahead: its byte range is not known yet; extrapolate from this batch's
        const uint32_t first = __shfl_sync(0xFFFFFFFFu, o0, 0), last = __shfl_sync(0xFFFFFFFFu, o1, 31);
        const uint32_t guess = first + kPrefetchAhead * (last - first) * nwarps + lane * 128u;
        if (lane < 18 && guess >= first && guess < total16) prefetch_l2(payload + guess);
      }
    }
    const uint64_t aux = sink.preload(e, valid);        // e.g. the event's timestamp: in flight while the batch is parsed
    // A batch is one contiguous byte range, staged in the warp buffer and parsed from shared memory.
    // BIG = false (messages up to ~72 B expected): the batch is staged whole; a batch that does not fit
    // is parsed from global memory, every lane its own message.  BIG = true: a batch may also be staged
    // in 2 parts of 16 messages (up to 144 B each) or 4 parts of 8 (up to 288 B) before falling back to
    // global memory.  Same results either way; the host picks the variant from the message-size hint.
    const bool sane = !valid || (o1 >= o0 && o0 >= S.pay_lo && o1 <= S.pay_hi);
    uint32_t outcome = BH_EV_DROP_DECODE; uint64_t kh = 0; double value = 0.0; int32_t ival = 0; uint32_t status = 0, kloc = 0;
    uint32_t parts = 0;
    if constexpr (!BIG) {                                 // small messages: the batch whole, or not at all
      const uint32_t A = __shfl_sync(0xFFFFFFFFu, o0, 0) & ~15u;
      if (__all_sync(0xFFFFFFFFu, !valid || (sane && o0 >= A && (o1 - A) <= kWarpBuf))) {
        const uint32_t bend = __shfl_sync(0xFFFFFFFFu, o1, min(32u, ce - e0) - 1);
        stage_range(payload, A, min((bend + 15u) & ~15u, total16), wb, lane);
        __syncwarp();
        const uint32_t mask = __ballot_sync(0xFFFFFFFFu, valid);
        if (valid) { SmemRd r{wb + (o0 - A)}; outcome = decode_and_check_smem(S, r, o1 - o0, kh, value, ival, status, mask, kloc); }
        __syncwarp();                               // the buffer is rewritten by the next batch
        parts = 1;
      }
    }
#pragma unroll
    for (uint32_t g = 32; BIG && g >= 8u && parts == 0; g >>= 1) {
      const uint32_t ga = __shfl_sync(0xFFFFFFFFu, o0, lane & ~(g - 1)) & ~15u;
      if (__all_sync(0xFFFFFFFFu, !valid || (sane && o0 >= ga && (o1 - ga) <= kWarpBuf))) parts = 32 / g;
    }
    if (BIG && parts) {
      const uint32_t g = 32 / parts, cnt = min(32u, ce - e0);
      for (uint32_t q = 0; q * g < cnt; q++) {
        const uint32_t A = __shfl_sync(0xFFFFFFFFu, o0, q * g) & ~15u;
        const uint32_t bend = __shfl_sync(0xFFFFFFFFu, o1, min(q * g + g, cnt) - 1);
        stage_range(payload, A, min((bend + 15u) & ~15u, total16), wb, lane);
        __syncwarp();
        const bool mine = valid && (!BIG || (lane / g) == q);
        const uint32_t mask = __ballot_sync(0xFFFFFFFFu, mine);
        if (mine) { SmemRd r{wb + (o0 - A)}; outcome = decode_and_check_smem(S, r, o1 - o0, kh, value, ival, status, mask, kloc); }
        __syncwarp();                               // the buffer is rewritten by the next part / batch
        if (!BIG) break;
      }
    } else if (!parts && valid && sane) {
      EventOut out;
      outcome = decode_and_check_gmem(S, payload + o0, o1 - o0, out);
      kh = out.kh; value = out.value; ival = out.ival; status = out.status; kloc = out.kloc;
    }
    sink.event(valid, e, outcome, kh, value, ival, status, aux, kloc);
  }
}

template <bool INTV, class SC>
struct AggSink {
  const SC& S; const DevTables& T; const SmemTable& M;
  const uint64_t* ts; uint64_t wbase; LaneCounters& C;
  uint64_t t_base;     // first millisecond of window `wbase`
  uint32_t w_ms;       // window_ms when the compare path may be used (it fits 30 bits), else 0
  __device__ __forceinline__ void prefetch(uint32_t e2, uint32_t lane) const {
    if (S.has_window && ts && lane < 2) prefetch_l2(ts + e2 + lane * 16);
  }
  __device__ __forceinline__ uint64_t preload(uint32_t e, bool valid) const {
    return (valid && S.has_window && ts) ? ts[e] : 0ull;
  }
  // called by the whole (converged) warp; `valid` marks the lanes that carry an event
  __device__ __forceinline__ void event(bool valid, uint32_t, uint32_t outcome, uint64_t kh, double value, int32_t ival, uint32_t, uint64_t ts_ms, uint32_t) {
    if (valid && (outcome & kSlowDecode)) C.bump(C_SLOW);
    outcome &= 0xFFu;
    if (valid && outcome != BH_EV_OK) C.bump(outcome);
    const bool ok = valid && outcome == BH_EV_OK;
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, ok);
    if (!ok) return;
    uint64_t win = 0;
    if (S.has_window) {
      // A CTA's events sit in a few consecutive windows from wbase on: the window is then three 32-bit compares on
      // the distance from wbase's start instead of a 64-bit division (any other timestamp takes the division).
      const uint64_t d = ts_ms - t_base;
      const uint32_t lo = (uint32_t)d;
      if (d < 4ull * w_ms) win = wbase + ((lo >= w_ms) + (lo >= 2u * w_ms) + (lo >= 3u * w_ms));   // (never taken when w_ms == 0)
      else win = div64(ts_ms, S.wdiv);
    }
    aggregate_event<INTV>(S, T, M, kh, win, wbase, value, ival, C, mask);
  }
};

template <bool INTV, bool BIG, class SC>
__device__ __forceinline__ void process_chunk(const SC& S, const DevTables& T, const SmemTable& M,
                                              const uint8_t* __restrict__ payload, const uint32_t* __restrict__ offsets,
                                              const uint64_t* __restrict__ ts, uint32_t cs, uint32_t ce, uint32_t wb,
                                              LaneCounters& C, uint64_t& wbase) {
  if (S.has_window) { uint64_t w0 = div64(ts ? ts[cs] : 0ull, S.wdiv); wbase = w0 ? w0 - 1 : 0; }
  AggSink<INTV, SC> sink{S, T, M, ts, wbase, C, wbase * S.window_ms, S.window_ms < (1ull << 30) ? (uint32_t)S.window_ms : 0u};
  raw_warp_loop<BIG>(S, payload, offsets, cs, ce, wb, sink);
}

// PROG != 0: the schema is the reference's own (ProgressSchemaT: 1 windowed grouping, 2 reference mode) and
// is read as compile-time constants
// MAXT: the f64 kernels exist twice -- built for 768 threads (72 registers; the 4 096-slot f64 table fits beside 24
// warp buffers) and for 1 024 threads (58 registers, no spills; used in partitioned mode, see launch_process_raw)
template <bool INTV, bool BIG, int PROG, int MAXT = kRawThreads>
__global__ void __launch_bounds__(MAXT, 1)
k_process_raw(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T,
              const uint8_t* __restrict__ payload, const uint32_t* __restrict__ offsets,
              const uint64_t* __restrict__ ts, uint32_t n, uint32_t chunk, uint32_t ns) {
  const bool hist = (S.ops & BH_OP_HIST) != 0;
  const SmemTable M = smem_table_carve<INTV>(ns, hist);
  const uint32_t wb = warp_buffer_origin((uint32_t)__cvta_generic_to_shared(g_smem) +
                                         (uint32_t)smem_table_bytes(ns, INTV, hist));
  smem_table_clear<INTV>(M);
  part_cursors_load(T);
  if (threadIdx.x < 8) g_cnt[threadIdx.x] = 0;
#ifdef BH_BOUNDS_CHECK
  if (threadIdx.x == 0) {
    g_dbg_first = (uint32_t)__cvta_generic_to_shared(g_smem) + (uint32_t)smem_table_bytes(ns, INTV, hist);
    g_dbg_glo = payload; g_dbg_ghi = payload + ((offsets[n] + 15u) & ~15u);    // every CTA writes the same values
  }
#endif
  __syncthreads();
  const uint32_t cs = min(n, blockIdx.x * chunk), ce = min(n, cs + chunk);
  LaneCounters C{0u};
  uint64_t wbase = 0;
  if (cs < ce) {
    if constexpr (PROG != 0) {
      const ProgressSchemaT<PROG == 2, PROG == 3> PS(S);
      process_chunk<INTV, BIG>(PS, T, M, payload, offsets, ts, cs, ce, wb, C, wbase);
    } else {
      process_chunk<INTV, BIG>(S, T, M, payload, offsets, ts, cs, ce, wb, C, wbase);
    }
  }
  __syncthreads();
  smem_table_flush<INTV, true>(S, T, M, wbase);
  part_cursors_store(T);
  counters_flush(T, C);
}

struct ColumnSink {
  const DevSchema& S;
  uint64_t* kh; double* val; int32_t* status; uint8_t* outcome;
  __device__ __forceinline__ void prefetch(uint32_t, uint32_t) const {}
  __device__ __forceinline__ uint64_t preload(uint32_t, bool) const { return 0ull; }
  __device__ __forceinline__ void event(bool valid, uint32_t e, uint32_t oc, uint64_t k, double v, int32_t iv, uint32_t st, uint64_t, uint32_t) {
    if (!valid) return;
    // integer value kinds: the fast path leaves `v` unset; (double)(int32) is the JS number
    oc &= 0xFFu;
    const bool have = oc != BH_EV_DROP_DECODE;
    kh[e] = k; val[e] = have && S.int_values ? (double)iv : v; status[e] = (int32_t)st; outcome[e] = (uint8_t)oc;
  }
};

template <bool BIG>
__global__ void __launch_bounds__(kRawThreads, 1)
k_decode(const __grid_constant__ DevSchema S, const uint8_t* __restrict__ payload,
         const uint32_t* __restrict__ offsets, uint32_t n, uint32_t chunk, uint64_t* kh, double* val,
         int32_t* status, uint8_t* outcome) {
  const uint32_t wb = warp_buffer_origin((uint32_t)__cvta_generic_to_shared(g_smem));
#ifdef BH_BOUNDS_CHECK
  if (threadIdx.x == 0) {
    g_dbg_first = (uint32_t)__cvta_generic_to_shared(g_smem);
    g_dbg_glo = payload; g_dbg_ghi = payload + ((offsets[n] + 15u) & ~15u);
  }
  __syncthreads();
#endif
  const uint32_t cs = min(n, blockIdx.x * chunk), ce = min(n, cs + chunk);
  if (cs >= ce) return;
  ColumnSink sink{S, kh, val, status, outcome};
  raw_warp_loop<BIG>(S, payload, offsets, cs, ce, wb, sink);
}

// =========================================================================================
// BH_F_VERIFY_KEYS: full-key identity check (collisions detected, never silently merged)
// =========================================================================================
// The tables identify a key by its 64-bit hash.  This kernel decodes a batch again and, per event, looks the hash
// up in a dictionary that keeps the key BYTES first seen with it (the winner of the slot writes them, everybody
// else compares): a mismatch is a hash collision between two different keys and is counted.  Key bytes are read
// from the payload in global memory (just staged, so from L2).
struct VerifySink {
  const DevSchema& S; const DevTables& T;
  const uint8_t* payload; const uint32_t* offsets;
  uint32_t collisions;
  __device__ __forceinline__ void prefetch(uint32_t, uint32_t) const {}
  __device__ __forceinline__ uint64_t preload(uint32_t, bool) const { return 0ull; }
  __device__ __forceinline__ void event(bool valid, uint32_t e, uint32_t oc, uint64_t kh, double, int32_t, uint32_t status, uint64_t, uint32_t kloc) {
    if (!valid || (oc & 0xFFu) != BH_EV_OK || kh >= kLockKey) return;
    const uint32_t klen = kloc >> 16;
    const GmemRd g{payload + offsets[e] + (kloc & 0xFFFFu)};
    const uint32_t st = S.n_parts == 2 ? status : 0u;
    // identity bytes: the key itself up to 40 bytes, else its first 32 bytes + a hash of all of it
    uint64_t w[5];
    if (klen <= 40u) {
#pragma unroll
      for (int i = 0; i < 5; i++) w[i] = g.le64(8u * i, klen > 8u * i ? klen - 8u * i : 0u);
    } else {
      uint64_t h = kHashInit;
      for (uint32_t i = 0; i < klen; i += 8) h = absorb(h, g.le64(i, klen - i));
#pragma unroll
      for (int i = 0; i < 4; i++) w[i] = g.le64(8u * i, 8u);
      w[4] = fmix64(h);
    }
    DictEntry* D = T.dict;
    uint64_t slot = (kh * kHashMul) >> (64u - ilog2_u64(T.dict_cap));
    for (uint64_t probes = 0; probes < T.dict_cap; probes++) {
      unsigned long long k = *(volatile unsigned long long*)&D[slot].kh;
      if (k == kEmptyKey) {
        k = atomicCAS(&D[slot].kh, (unsigned long long)kEmptyKey, (unsigned long long)kh);
        if (k == kEmptyKey) {                              // this lane owns the entry: write, then publish
          D[slot].len = klen; D[slot].status = st;
          unsigned long long* b = (unsigned long long*)D[slot].bytes;
#pragma unroll
          for (int i = 0; i < 5; i++) b[i] = w[i];
          __threadfence();
          *(volatile unsigned int*)&D[slot].ready = 1u;
          return;
        }
      }
      if (k == kh) {
        while (*(volatile unsigned int*)&D[slot].ready == 0u) {}
        __threadfence();
        const unsigned long long* b = (const unsigned long long*)D[slot].bytes;
        bool same = D[slot].len == klen && D[slot].status == st;
#pragma unroll
        for (int i = 0; i < 5; i++) same = same && __ldcg(b + i) == w[i];
        if (!same) collisions++;
        return;
      }
      slot = (slot + 1) & (T.dict_cap - 1);
    }
  }
};

template <bool BIG>
__global__ void __launch_bounds__(kRawThreads, 1)
k_verify_keys(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T, const uint8_t* __restrict__ payload,
              const uint32_t* __restrict__ offsets, uint32_t n, uint32_t chunk) {
  const uint32_t wb = warp_buffer_origin((uint32_t)__cvta_generic_to_shared(g_smem));
#ifdef BH_BOUNDS_CHECK
  if (threadIdx.x == 0) {
    g_dbg_first = (uint32_t)__cvta_generic_to_shared(g_smem);
    g_dbg_glo = payload; g_dbg_ghi = payload + ((offsets[n] + 15u) & ~15u);
  }
  __syncthreads();
#endif
  const uint32_t cs = min(n, blockIdx.x * chunk), ce = min(n, cs + chunk);
  if (cs >= ce) return;
  VerifySink sink{S, T, payload, offsets, 0u};
  raw_warp_loop<BIG>(S, payload, offsets, cs, ce, wb, sink);
  const uint32_t c = __reduce_add_sync(0xFFFFFFFFu, sink.collisions);
  if ((threadIdx.x & 31) == 0 && c) atomicAdd(&T.counters[C_COLLISION], (unsigned long long)c);
}

// =========================================================================================
// columnar path
// =========================================================================================
constexpr int kColThreads = 1024;

__global__ void __launch_bounds__(kColThreads, 1)
k_process_columns(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T,
                  const uint64_t* __restrict__ kh, const uint64_t* __restrict__ ts,
                  const double* __restrict__ val, uint64_t n, uint64_t chunk, uint32_t ns) {
  const SmemTable M = smem_table_carve<false>(ns, (S.ops & BH_OP_HIST) != 0);
  smem_table_clear<false>(M);
  part_cursors_load(T);
  if (threadIdx.x < 8) g_cnt[threadIdx.x] = 0;
  __syncthreads();
  const uint64_t cs = min(n, (uint64_t)blockIdx.x * chunk), ce = min(n, cs + chunk);
  LaneCounters C{0u};
  uint64_t wbase = 0;
  if (cs < ce && S.has_window) { uint64_t w0 = div64(ts[cs], S.wdiv); wbase = w0 ? w0 - 1 : 0; }
  const bool need_value = (S.ops & (BH_OP_SUM | BH_OP_MIN | BH_OP_MAX | BH_OP_HIST)) != 0;
  // one trip count per warp; the next iteration's three columns are loaded before this one is aggregated
  const uint32_t lane = threadIdx.x & 31;
  uint64_t base = cs + (threadIdx.x & ~31u);
  uint64_t nk = 0, nt = 0; double nv = 0.0;
  if (base + lane < ce) { nk = kh[base + lane]; nt = S.has_window ? ts[base + lane] : 0ull; nv = val ? val[base + lane] : 0.0; }
  for (; base < ce; base += blockDim.x) {
    const bool valid = base + lane < ce;
    const uint64_t k = nk, t = nt; const double v = nv;
    const uint64_t nxt = base + blockDim.x + lane;
    if (nxt < ce) { nk = kh[nxt]; nt = S.has_window ? ts[nxt] : 0ull; nv = val ? val[nxt] : 0.0; }
    const uint64_t win = valid && S.has_window ? div64(t, S.wdiv) : 0ull;
    const bool nan = need_value && v != v;
    if (valid && nan) C.bump(C_NAN);
    const bool ok = valid && !nan;
    const uint32_t mask = __ballot_sync(0xFFFFFFFFu, ok);
    if (ok) aggregate_event<false>(S, T, M, k, win, wbase, v, 0, C, mask);
  }
  __syncthreads();
  smem_table_flush<false, true>(S, T, M, wbase);
  part_cursors_store(T);
  counters_flush(T, C);
}

// =========================================================================================
// partitioned mode, second phase: one CTA per partition folds the logged tuples into the tables
// =========================================================================================
// Algorithmic bytes: 16 per logged tuple read once (coalesced, a region is contiguous), plus the partition's
// slice of every open window's table read and written once per drain -- which is why tuples are collected over
// many launches (an "epoch") before they are drained.  The tuples of one partition carry at most
// keys / P distinct keys per window, so the shared table absorbs nearly all of them and the HBM slice sees one
// update per (key, window), not one per event; what does not fit falls through to L2 atomics on a slice that
// only this CTA is touching.  Histogram buckets go to the HBM rows directly (one L2 atomic per tuple): the
// slice's rows stay in L2 while the partition is being drained.
constexpr int kDrainThreads = 1024;
constexpr uint32_t kDrainMaxCtas = 256;       // regions per partition (= SMs of the device that logged them)
__shared__ unsigned long long g_wmin;
__shared__ uint32_t g_fill[kDrainMaxCtas];
__shared__ uint32_t g_next;

template <bool INTV>
__global__ void __launch_bounds__(kDrainThreads, 1)
k_part_drain(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T, uint32_t ns) {
  const bool hist = (S.ops & BH_OP_HIST) != 0;
  const SmemTable M = smem_table_carve<INTV>(ns, hist, false);
  // window id of every ring slot, after the table (a tuple names its window by ring slot)
  unsigned long long* ringwin = (unsigned long long*)(g_smem + smem_table_bytes(ns, INTV, hist, false));
  if (threadIdx.x == 0) g_wmin = kNoWindow;
  if (threadIdx.x < 8) g_cnt[threadIdx.x] = 0;
  __syncthreads();
  for (uint32_t r = threadIdx.x; r < T.ring; r += blockDim.x) {
    const unsigned long long w = T.ring_win[r];
    ringwin[r] = w;
    if (w != kNoWindow) atomicMin(&g_wmin, w);
  }
  __syncthreads();
  const uint64_t wbase = g_wmin == kNoWindow || g_wmin == 0 ? 0 : g_wmin - 1;
  const uint32_t P = 1u << T.part_lg, sh = 64u - T.part_lg;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t ctas = min(T.part_ctas, kDrainMaxCtas);
  const uint64_t slice = T.cap >> T.part_lg;                     // table slots of one partition (0: table smaller than P)
  LaneCounters C{0u};
  // (Folding a large partition window by window, so that its pairs always fit the shared table, was measured at
  // 10 M keys: 2.24 vs 2.01 ms per drain.  With ~1.6 tuples per (key, window) there is nothing to absorb; the cost
  // is one HBM-table update per pair either way, and what overflows the shared table goes there directly.)
  for (uint32_t p = blockIdx.x; p < P; p += gridDim.x) {
    for (uint32_t c = threadIdx.x; c < ctas; c += blockDim.x) g_fill[c] = min(T.part_fill[((uint64_t)c << T.part_lg) | p], T.part_cap);
    // The partition's slice of every open window's key and stat columns is pulled into L2 now: the look-ups of
    // the flush (and of the inserts, with histograms) then cost an L2 hit instead of a DRAM round trip each.
    // A slice is contiguous: partition = the top bits of the slot index.
    {
      const uint64_t lines_k = (slice * 8 + 127) / 128, lines_s = (slice * sizeof(Stat) + 127) / 128;
      for (uint32_t r = 0; r < T.ring; r++) {
        if (ringwin[r] == kNoWindow) continue;
        const uint8_t* kb = (const uint8_t*)(T.keys + (uint64_t)r * T.stride + (uint64_t)p * slice);
        const uint8_t* sb = (const uint8_t*)(T.stats + (uint64_t)r * T.stride + (uint64_t)p * slice);
        for (uint64_t i = threadIdx.x; i < lines_k + lines_s; i += blockDim.x)
          prefetch_l2(i < lines_k ? kb + i * 128 : sb + (i - lines_k) * 128);
      }
    }
    {
      smem_table_clear<INTV>(M);
      if (threadIdx.x == 0) g_next = 0;
      __syncthreads();
      // Warps take whole regions, handed out by a shared counter (regions differ in fill), 32 tuples at a time;
      // the next 32 are already in flight while the current ones are aggregated.
      auto grab = [&]() { uint32_t v = 0; if (lane == 0) v = atoms_add(&g_next, 1u); return __shfl_sync(0xFFFFFFFFu, v, 0); };
      auto next_pos = [&](uint32_t& c, uint32_t& i0) { while (c < ctas && i0 >= g_fill[c]) { c = grab(); i0 = 0; } };
      auto load = [&](uint32_t c, uint32_t i0, bool& valid) {
        uint4 t = make_uint4(0, 0, 0, 0);
        valid = c < ctas && i0 + lane < g_fill[c];
        if (valid) t = ldg_stream_here(T.part_tuples + (((uint64_t)c << T.part_lg) | p) * T.part_cap + i0 + lane);
        return t;
      };
      uint32_t c = grab(), i0 = 0;
      next_pos(c, i0);
      bool valid;
      uint4 t = load(c, i0, valid);
      while (c < ctas) {
        uint32_t nc = c, ni = i0 + 32;
        next_pos(nc, ni);
        bool nvalid;
        const uint4 nt = load(nc, ni, nvalid);
        const uint64_t a = ((uint64_t)t.y << 32) | t.x, vbits = ((uint64_t)t.w << 32) | t.z;
        const uint32_t r = (uint32_t)(a >> sh);
        const uint32_t mask = __ballot_sync(0xFFFFFFFFu, valid);
        if (valid) {
          const uint64_t kh = ((a & ((1ull << sh) - 1)) | ((uint64_t)p << sh)) * kHashMulInv;
          const double v = __longlong_as_double((long long)vbits);
          aggregate_event<INTV, false>(S, T, M, kh, ringwin[r], wbase, v, INTV ? (int32_t)v : 0, C, mask);
        }
        __syncwarp();
        c = nc; i0 = ni; t = nt; valid = nvalid;
      }
      __syncthreads();
      smem_table_flush<INTV>(S, T, M, wbase);
      __syncthreads();
    }
    for (uint32_t cc = threadIdx.x; cc < T.part_ctas; cc += blockDim.x) T.part_fill[((uint64_t)cc << T.part_lg) | p] = 0;
    __syncthreads();
  }
  // the tuples were counted as aggregated when they were logged: only what was refused now moves to the drop counters
  if (threadIdx.x == C_TABLE_FULL || threadIdx.x == C_WINDOW) {
    const unsigned long long d = g_cnt[threadIdx.x];
    if (d) { atomicAdd(&T.counters[threadIdx.x], d); atomicAdd(&T.counters[C_OK], 0ull - d); }
  }
}

// =========================================================================================
// partitioned mode, second phase, slice-mirror form: the fold used whenever a slice (or a power-of-two part of
// it) fits shared memory
// =========================================================================================
// A partition's slots are one contiguous slice of every window's table and no key of the partition sits outside
// it (g_probe_wrap), and while the fold runs nobody else touches it.  So the CTA that owns the partition does not
// go through a hashed front table and per-entry HBM atomics at all: for each open window it loads the slice's KEY
// column into shared memory (coalesced), probes and claims slots there exactly as g_find_or_insert would in HBM
// (same home slot, same order), accumulates count / sum / min / max -- and the histogram rows, as 16-bit counters --
// per SLOT in shared memory, and adds the touched rows back with plain coalesced read-modify-writes.  HBM traffic
// per fold: the tuples once (16 B each), the key column of the open windows once, and 64 B (+ 2 x 4 B per live
// bucket) per touched (key, window) -- not one random sector per tuple and table.
// A slice larger than shared memory is taken in `sub`-slot parts; a tuple's probe chain that runs off the end of
// the part in shared memory goes on in the HBM table (those slots are this CTA's too, and not in shared memory at
// that moment).  Tuples are re-read once per (window, part) pass: after the first pass they come from L2.
// 16-bit histogram counters cannot overflow: with histograms a pass takes regions holding <= 65 535 tuples together
// (a "round"); a partition with more is folded in several rounds.
__shared__ unsigned long long g_seen;
__shared__ uint32_t g_round_end;

template <bool INTV>
__global__ void __launch_bounds__(kDrainThreads, 1)
k_part_fold(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T, uint32_t sub_lg) {
  const uint32_t ops = S.ops;
  const bool hist = (ops & BH_OP_HIST) != 0;
  const uint32_t B = T.hist_buckets, sub = 1u << sub_lg;
  const SmemTable M = smem_table_carve<INTV>(sub, false);
  uint8_t* after = g_smem + (((size_t)sub * smem_slot_bytes(INTV, false) + 15) & ~(size_t)15);
  unsigned long long* ringwin = (unsigned long long*)after;
  uint32_t* h32 = (uint32_t*)(after + (((size_t)T.ring * 8 + 15) & ~(size_t)15));     // u16 [sub][B], two to a word
  const uint32_t hwords = hist ? (sub * B) / 2 : 0;
  if (threadIdx.x < 8) g_cnt[threadIdx.x] = 0;
  for (uint32_t r = threadIdx.x; r < T.ring; r += blockDim.x) ringwin[r] = T.ring_win[r];
  __syncthreads();
  unsigned long long open = 0;
  for (uint32_t r = 0; r < T.ring; r++) if (ringwin[r] != kNoWindow) open |= 1ull << r;
  const uint32_t P = 1u << T.part_lg, sh = 64u - T.part_lg, cap_sh = 64u - ilog2_u64(T.cap);
  const uint32_t slice = (uint32_t)(T.cap >> T.part_lg), n_sub = slice >> sub_lg;
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t ctas = min(T.part_ctas, kDrainMaxCtas);
  volatile unsigned long long* vkey = M.key();
  LaneCounters C{0u};
  for (uint32_t p = blockIdx.x; p < P; p += gridDim.x) {
    for (uint32_t c = threadIdx.x; c < ctas; c += blockDim.x) g_fill[c] = min(T.part_fill[((uint64_t)c << T.part_lg) | p], T.part_cap);
    __syncthreads();
    for (uint32_t c0 = 0; c0 < ctas;) {
      if (threadIdx.x == 0) {
        uint32_t c1 = ctas;
        if (hist) { uint32_t sum = 0; c1 = c0; while (c1 < ctas && sum + g_fill[c1] <= 65535u) sum += g_fill[c1++]; if (c1 == c0) c1++; }
        g_round_end = c1; g_seen = 0;
      }
      __syncthreads();
      const uint32_t c1 = g_round_end;
      bool first = true;                     // the round's first pass also notes which windows have tuples at all
      unsigned long long todo = open;
      for (uint32_t r = 0; r < T.ring; r++) {
        for (uint32_t sb = 0; sb < n_sub && ((todo >> r) & 1); sb++) {
          const uint64_t g0 = (uint64_t)r * T.stride + (uint64_t)p * slice + ((uint64_t)sb << sub_lg);
          // ---- the part's key column in, everything else zero
          for (uint32_t i = threadIdx.x; i < sub; i += blockDim.x) {
            M.key()[i] = __ldcg(T.keys + g0 + i);
            M.cnt<INTV>()[i] = 0;
            if (INTV) { M.sum_lo()[i] = 0; M.sum_hi()[i] = 0; M.imin()[i] = 0x7FFFFFFF; M.imax()[i] = (int32_t)0x80000000; }
            else { M.fsum()[i] = 0.0; M.fmin_inv()[i] = 0; M.fmax_ord()[i] = 0; }
          }
          for (uint32_t i = threadIdx.x; i < hwords; i += blockDim.x) h32[i] = 0;
          if (threadIdx.x == 0) g_next = c0;
          __syncthreads();
          // ---- the round's tuples: warps take whole regions, 32 tuples at a time, the next 32 already in flight
          {
            auto grab = [&]() { uint32_t v = 0; if (lane == 0) v = atoms_add(&g_next, 1u); return __shfl_sync(0xFFFFFFFFu, v, 0); };
            // a region just taken is pulled into L2 whole (<= 128 B per tuple octet): the 32-tuple loads that follow
            // then wait for L2, not for DRAM
            auto next_pos = [&](uint32_t& c, uint32_t& i0) {
              while (c < c1 && i0 >= g_fill[c]) {
                c = grab(); i0 = 0;
                if (c < c1) {
                  const uint8_t* rb = (const uint8_t*)(T.part_tuples + (((uint64_t)c << T.part_lg) | p) * T.part_cap);
                  for (uint32_t o = lane * 128u; o < g_fill[c] * 16u; o += 32u * 128u) prefetch_l2(rb + o);
                }
              }
            };
            auto load = [&](uint32_t c, uint32_t i0, bool& valid) {
              uint4 t = make_uint4(0, 0, 0, 0);
              valid = c < c1 && i0 + lane < g_fill[c];
              if (valid) t = ldg_stream_here(T.part_tuples + (((uint64_t)c << T.part_lg) | p) * T.part_cap + i0 + lane);
              return t;
            };
            unsigned long long seen = 0;
            uint32_t c = grab(), i0 = 0;
            next_pos(c, i0);
            bool valid;
            uint4 t = load(c, i0, valid);
            while (c < c1) {
              uint32_t nc = c, ni = i0 + 32;
              next_pos(nc, ni);
              bool nvalid;
              const uint4 nt = load(nc, ni, nvalid);
              const uint64_t a = ((uint64_t)t.y << 32) | t.x, vbits = ((uint64_t)t.w << 32) | t.z;
              const uint32_t rr = (uint32_t)(a >> sh);
              const uint64_t hs = (a & ((1ull << sh) - 1)) | ((uint64_t)p << sh);
              const uint32_t off = (uint32_t)(hs >> cap_sh) & (slice - 1);
              if (valid && first) {
                seen |= 1ull << rr;
                if (!((open >> rr) & 1)) C.bump(C_WINDOW);          // (its window is gone: cannot happen, a close folds first)
              }
              const bool mine = valid && rr == r && (off >> sub_lg) == sb;
              const uint32_t mask = __ballot_sync(0xFFFFFFFFu, mine);
              if (mine) {
                const uint64_t kh = hs * kHashMulInv;
                const double v = __longlong_as_double((long long)vbits);
                uint32_t j = off & (sub - 1);
                int slot = -1;
                for (uint32_t left = n_sub == 1 ? sub : sub - j; left; left--) {
                  unsigned long long k = vkey[j];
                  if (k == kEmptyKey) {
                    k = atomicCAS(&M.key()[j], (unsigned long long)kEmptyKey, (unsigned long long)kh);
                    if (k == kEmptyKey) k = kh;
                  }
                  if (k == kh) { slot = (int)j; break; }
                  j = (j + 1) & (sub - 1);
                }
                __syncwarp(mask);          // the lanes left the probe loop at different turns
                if (slot >= 0) {
                  smem_apply<INTV>(M, slot, ops, v, INTV ? (int32_t)v : 0);
                  if (hist) {
                    const uint32_t hi = (uint32_t)slot * B + hist_bucket(vbits, B, S.hist_e0);
                    atomicAdd(&h32[hi >> 1], 1u << ((hi & 1u) * 16u));
                  }
                } else if (n_sub == 1) {
                  C.bump(C_TABLE_FULL);
                } else {
                  // off the end of the part held here: the rest of the chain is walked in the HBM table
                  const uint64_t start = (uint64_t)p * slice + ((((uint64_t)sb + 1) << sub_lg) & (slice - 1));
                  const uint32_t gs = g_probe_from(T, r, kh, start, slice - 1, slice - sub);
                  if (gs == kGslotFull) C.bump(C_TABLE_FULL);
                  else {
                    const uint64_t gi = (uint64_t)r * T.stride + gs, ord = f64_order(vbits);
                    g_update(T, ops, gi, 1ull, v, ~ord, ord);
                    if (hist) atomicAdd(&T.hist[gi * B + hist_bucket(vbits, B, S.hist_e0)], 1u);
                  }
                }
              }
              __syncwarp();
              c = nc; i0 = ni; t = nt; valid = nvalid;
            }
            if (first) {
              const uint32_t lo = __reduce_or_sync(0xFFFFFFFFu, (uint32_t)seen), hi = __reduce_or_sync(0xFFFFFFFFu, (uint32_t)(seen >> 32));
              if (lane == 0 && (lo | hi)) atomicOr(&g_seen, ((unsigned long long)hi << 32) | lo);
            }
          }
          __syncthreads();
          // ---- touched rows back: coalesced read-modify-write, nobody else is in this slice
          for (uint32_t i = threadIdx.x; i < sub; i += blockDim.x) {
            const uint32_t cnt = M.cnt<INTV>()[i];
            if (!cnt) continue;
            __stcg(T.keys + g0 + i, M.key()[i]);
            ulonglong2* sp = (ulonglong2*)(T.stats + g0 + i);
            ulonglong2 s0 = __ldcg(sp), s1 = __ldcg(sp + 1);          // {count, sum} {min_inv, max_ord}
            double sum; unsigned long long mn_inv, mx_ord;
            if (INTV) {
              sum = (double)(long long)(((unsigned long long)(uint32_t)M.sum_hi()[i] << 32) | M.sum_lo()[i]);
              mn_inv = ~f64_order((uint64_t)__double_as_longlong((double)M.imin()[i]));
              mx_ord = f64_order((uint64_t)__double_as_longlong((double)M.imax()[i]));
            } else {
              sum = M.fsum()[i]; mn_inv = M.fmin_inv()[i]; mx_ord = M.fmax_ord()[i];
            }
            s0.x += cnt;
            if (ops & BH_OP_SUM) s0.y = (unsigned long long)__double_as_longlong(__longlong_as_double((long long)s0.y) + sum);
            if (ops & BH_OP_MIN) s1.x = max(s1.x, mn_inv);
            if (ops & BH_OP_MAX) s1.y = max(s1.y, mx_ord);
            __stcg(sp, s0); __stcg(sp + 1, s1);
          }
          for (uint32_t i = threadIdx.x; i < hwords; i += blockDim.x) {
            const uint32_t w = h32[i];
            if (!w) continue;
            uint32_t* g = T.hist + g0 * B + 2ull * i;
            if (w & 0xFFFFu) __stcg(g, __ldcg(g) + (w & 0xFFFFu));
            if (w >> 16) __stcg(g + 1, __ldcg(g + 1) + (w >> 16));
          }
          __syncthreads();
          if (first) { todo &= g_seen; first = false; }
        }
      }
      c0 = c1;
    }
    for (uint32_t cc = threadIdx.x; cc < T.part_ctas; cc += blockDim.x) T.part_fill[((uint64_t)cc << T.part_lg) | p] = 0;
    __syncthreads();
  }
  // the tuples were counted as aggregated when they were logged: only what was refused now moves to the drop counters
  if (threadIdx.x == C_TABLE_FULL || threadIdx.x == C_WINDOW) {
    const unsigned long long d = g_cnt[threadIdx.x];
    if (d) { atomicAdd(&T.counters[threadIdx.x], d); atomicAdd(&T.counters[C_OK], 0ull - d); }
  }
}

// =========================================================================================
// window close: export (compact by owner), merge, recycle
// =========================================================================================
// Liveness is decided from the key column (8 bytes per slot, streamed) so that a close or an export of a large,
// sparsely filled table reads the 32-byte stat rows of live slots only.
__device__ __forceinline__ bool export_row(const DevTables& T, uint64_t base, uint64_t slot, uint64_t* key) {
  if (slot < T.cap) {
    const unsigned long long k = T.keys[base + slot];
    if (k == kEmptyKey) return false;
    *key = k;
  } else {
    *key = slot == T.cap ? kLockKey : kEmptyKey;        // the two side slots have no key cell
  }
  return T.stats[base + slot].count != 0;
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

// counts from the cursors of a single-pass export (cursor[o] started at o * owner_stride)
__global__ void k_export_counts(uint32_t n_owners, const unsigned long long* cursor, unsigned long long* owner_counts,
                                uint64_t owner_stride) {
  for (uint32_t o = blockIdx.x * blockDim.x + threadIdx.x; o < n_owners; o += gridDim.x * blockDim.x)
    owner_counts[o] = cursor[o] - (unsigned long long)o * owner_stride;
}

// One warp looks at 32 consecutive slots.  Output positions are claimed once per warp and owner (the lanes that
// export to the same owner elect a leader), records are written by their lanes, histogram rows are copied by the
// whole warp, one row at a time, so that both sides of the copy are coalesced.
__global__ void __launch_bounds__(256)
k_export_scatter(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T, uint64_t win,
                 uint32_t owner_shift, bh_record* out, uint32_t* hist_out, uint64_t cap,
                 unsigned long long* cursor, uint64_t owner_stride,
                 const unsigned long long* peer_base, const unsigned long long* peer_hist_base,
                 uint32_t my_rank) {
  uint64_t rows = (uint64_t)T.ring * T.stride, r0 = 0;
  if (win != kNoWindow) { r0 = (uint64_t)((uint32_t)win & (T.ring - 1)) * T.stride; rows = r0 + T.stride; }
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t step = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t row0 = r0 + blockIdx.x * (uint64_t)blockDim.x + (threadIdx.x & ~31u); row0 < rows; row0 += step) {
    const uint64_t row = row0 + lane;
    uint64_t base = 0, slot = 0, w = 0, key = 0;
    bool live = row < rows && export_pick(T, win, row, &base, &slot, &w) && export_row(T, base, slot, &key);
    const uint32_t livemask = __ballot_sync(0xFFFFFFFFu, live);
    if (!livemask) continue;
    const uint32_t o = live && owner_shift < 64 ? (uint32_t)(key >> owner_shift) : 0u;
    unsigned long long pos = 0;
    if (live) {
      const uint32_t peers = __match_any_sync(livemask, o);
      const uint32_t leader = __ffs(peers) - 1;
      unsigned long long first = 0;
      if (lane == leader) first = atomicAdd(&cursor[o], (unsigned long long)__popc(peers));
      first = __shfl_sync(peers, first, leader);
      pos = first + __popc(peers & ((1u << lane) - 1));
      if (pos >= cap || (owner_stride && pos >= (unsigned long long)(o + 1) * owner_stride)) live = false;   // cut: counted, not stored
    }
    uint32_t* hdst = nullptr;
    if (live) {
      const Stat s = T.stats[base + slot];
      bh_record rec;
      rec.key_hash = key; rec.window_id = w; rec.count = s.count;
      rec.sum = (S.ops & BH_OP_SUM) ? s.sum : 0.0;
      rec.min = (S.ops & BH_OP_MIN) && s.min_inv ? __longlong_as_double((long long)f64_unorder(~s.min_inv)) : 0.0;
      rec.max = (S.ops & BH_OP_MAX) && s.max_ord ? __longlong_as_double((long long)f64_unorder(s.max_ord)) : 0.0;
      // peer_base: owner o's receive buffer lives on GPU o (mapped peer memory, NVLink): the row goes straight
      // into block `my_rank` of it -- the export IS the exchange, no staging copy and no all-to-all
      const uint64_t peer_row = (uint64_t)my_rank * owner_stride + (pos - (unsigned long long)o * owner_stride);
      if (peer_base) ((bh_record*)peer_base[o])[peer_row] = rec;
      else out[pos] = rec;
      hdst = peer_base ? (peer_hist_base ? (uint32_t*)peer_hist_base[o] + peer_row * T.hist_buckets : nullptr)
                       : (hist_out ? hist_out + pos * T.hist_buckets : nullptr);
    }
    if (T.hist) {
      uint32_t todo = __ballot_sync(0xFFFFFFFFu, hdst != nullptr);
      while (todo) {
        const uint32_t l = __ffs(todo) - 1;
        todo &= todo - 1;
        uint32_t* d = (uint32_t*)__shfl_sync(0xFFFFFFFFu, (unsigned long long)hdst, l);
        const uint32_t* h = T.hist + (row0 + l) * T.hist_buckets;
        for (uint32_t b = lane; b < T.hist_buckets; b += 32) d[b] = h[b];
      }
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
    g_update<false>(T, S.ops, gi, rec.count, rec.sum, ~omin, omax);
    if (hist && T.hist)
      for (uint32_t b = 0; b < T.hist_buckets; b++) {
        uint32_t h = hist[i * T.hist_buckets + b];
        if (h) atomicAdd(&T.hist[gi * T.hist_buckets + b], h);
      }
  }
  if (dropped[0]) atomicAdd(&T.counters[C_TABLE_FULL], (unsigned long long)dropped[0]);
  if (dropped[1]) atomicAdd(&T.counters[C_WINDOW], (unsigned long long)dropped[1]);
}

// Rows received from other GPUs, one block per sender: block b holds min(counts[b], stride) valid rows at
// [b * stride, ...).  Nothing past a block's count is read, so the receive buffer never needs clearing.
// Loads bypass L1 (.cg): the rows were written by other GPUs over NVLink.
__global__ void k_merge_blocks(const __grid_constant__ DevSchema S, const __grid_constant__ DevTables T,
                               const bh_record* recs, const uint32_t* hist, uint32_t n_blocks, uint64_t stride,
                               const unsigned long long* counts) {
  uint32_t dropped[2] = {0, 0};
  for (uint32_t b = 0; b < n_blocks; b++) {
    const uint64_t n = min((uint64_t)__ldcg(&counts[b]), stride);
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
      const uint64_t row = (uint64_t)b * stride + i;
      const unsigned long long* q = reinterpret_cast<const unsigned long long*>(recs + row);
      bh_record rec;
      rec.key_hash = __ldcg(q); rec.window_id = __ldcg(q + 1); rec.count = __ldcg(q + 2);
      rec.sum = __longlong_as_double((long long)__ldcg(q + 3));
      rec.min = __longlong_as_double((long long)__ldcg(q + 4));
      rec.max = __longlong_as_double((long long)__ldcg(q + 5));
      if (rec.count == 0) continue;
      uint32_t r;
      const uint32_t gs = g_find_or_insert(T, rec.window_id, rec.key_hash, &r);
      if (gs >= kGslotWindow) { dropped[gs == kGslotFull ? 0 : 1] += 1; continue; }
      const uint64_t gi = (uint64_t)r * T.stride + gs;
      const uint64_t omin = f64_order((uint64_t)__double_as_longlong(rec.min));
      const uint64_t omax = f64_order((uint64_t)__double_as_longlong(rec.max));
      g_update<false>(T, S.ops, gi, rec.count, rec.sum, ~omin, omax);
      if (hist && T.hist)
        for (uint32_t k = 0; k < T.hist_buckets; k++) {
          const uint32_t h = __ldcg(&hist[row * T.hist_buckets + k]);
          if (h) atomicAdd(&T.hist[gi * T.hist_buckets + k], h);
        }
    }
  }
  if (dropped[0]) atomicAdd(&T.counters[C_TABLE_FULL], (unsigned long long)dropped[0]);
  if (dropped[1]) atomicAdd(&T.counters[C_WINDOW], (unsigned long long)dropped[1]);
}

// win == kNoWindow: every open window (bh_reset_device).  Only slots that hold a key are written: closing a
// window costs 8 bytes read per slot plus the rows that were in use, not the whole table (histogram rows
// included) every time.
__global__ void k_close_clear(const __grid_constant__ DevTables T, uint64_t win) {
  uint64_t rows = (uint64_t)T.ring * T.stride, r0 = 0;
  if (win != kNoWindow) {
    const uint32_t r = (uint32_t)win & (T.ring - 1);
    if (T.ring_win[r] != win) return;
    r0 = (uint64_t)r * T.stride; rows = r0 + T.stride;
  }
  for (uint64_t row = r0 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; row < rows; row += (uint64_t)gridDim.x * blockDim.x) {
    const uint32_t r = (uint32_t)(row / T.stride);
    const uint64_t slot = row - (uint64_t)r * T.stride;
    if (T.ring_win[r] == kNoWindow) continue;             // a free ring slot is clean already
    if (slot < T.cap) {
      if (T.keys[row] == kEmptyKey) continue;
      T.keys[row] = kEmptyKey;
    } else if (T.stats[row].count == 0) {
      continue;
    }
    T.stats[row] = Stat{0ull, 0.0, 0ull, 0ull};
    if (T.hist) {
      uint4* h = reinterpret_cast<uint4*>(T.hist + row * T.hist_buckets);
      if ((T.hist_buckets & 3u) == 0) for (uint32_t b = 0; b < T.hist_buckets / 4; b++) h[b] = make_uint4(0, 0, 0, 0);
      else for (uint32_t b = 0; b < T.hist_buckets; b++) T.hist[row * T.hist_buckets + b] = 0;
    }
  }
}
__global__ void k_close_release(const __grid_constant__ DevTables T, uint64_t win, bool reset) {
  if (win == kNoWindow) {                                 // every ring slot; reset: also the counters and the tuple log
    const uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x, step = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t r = i; r < T.ring; r += step) T.ring_win[r] = kNoWindow;
    if (!reset) return;
    if (i < (uint64_t)kCounterSlots) T.counters[i] = 0;
    if (T.part_lg) for (uint64_t q = i; q < ((uint64_t)T.part_ctas << T.part_lg); q += step) T.part_fill[q] = 0;
    return;
  }
  uint32_t r = (uint32_t)win & (T.ring - 1);
  if (blockIdx.x == 0 && threadIdx.x == 0 && T.ring_win[r] == win) T.ring_win[r] = kNoWindow;
}

// =========================================================================================
// launchers
// =========================================================================================
// once per device, with that device current (bh_create)
cudaError_t configure_kernels(const LaunchGeom& G) {
  const void* fns[] = {(const void*)k_process_raw<true, false, 0>, (const void*)k_process_raw<true, true, 0>,
                       (const void*)k_process_raw<false, false, 0>, (const void*)k_process_raw<false, true, 0>,
                       (const void*)k_process_raw<false, false, 0, kRawThreadsF64>, (const void*)k_process_raw<false, true, 0, kRawThreadsF64>,
                       (const void*)k_process_raw<true, false, 1>, (const void*)k_process_raw<true, false, 2>,
                       (const void*)k_process_raw<false, false, 3>, (const void*)k_process_raw<false, false, 3, kRawThreadsF64>,
                       (const void*)k_decode<false>, (const void*)k_decode<true>, (const void*)k_process_columns,
                       (const void*)k_part_drain<true>, (const void*)k_part_drain<false>,
                       (const void*)k_part_fold<true>, (const void*)k_part_fold<false>,
                       (const void*)k_verify_keys<false>, (const void*)k_verify_keys<true>};
  for (const void* f : fns) {
    cudaFuncAttributes at;
    cudaError_t e = cudaFuncGetAttributes(&at, f);
    if (e != cudaSuccess) return e;
    // everything the SM offers beyond the kernel's own static shared memory (pick_smem_slots keeps 1 KB + the
    // static arrays of the drain kernel clear of it)
    e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, G.smem_optin - (int)at.sharedSizeBytes);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

static uint32_t pick_smem_slots(bool intv, bool hist, const LaunchGeom& G, size_t other_bytes, uint64_t table_cap, uint32_t ring) {
  // front table: the largest power of two (<= 4096 slots) that fits beside the warp buffers
  uint32_t ns = 4096;
  if (const char* e = getenv("BEHOLDER_SMEM_SLOTS")) { int v = atoi(e); if (v >= 64 && v <= 4096 && !(v & (v - 1))) ns = (uint32_t)v; }   // experiments
  while (ns > 64 && smem_table_bytes(ns, intv, hist) + 16 + other_bytes + 1024 > (size_t)G.smem_optin) ns >>= 1;
  // No more slots than the HBM tables can hold pairs for (reference mode: 64 status values, one window).  Shared memory
  // a CTA does not ask for stays L1: the per-status kernel runs 8 % faster with 128 slots than with 4096 (0.196 -> 0.181
  // ms per 12.5 M events; carve-out 100 KB instead of 228 KB).
  const uint64_t pairs = 2 * table_cap * (ring < 4 ? ring : 4);
  while (ns > 64 && (uint64_t)(ns >> 1) >= pairs) ns >>= 1;
  return ns;
}

static DevSchema with_bounds(const DevSchema& S0, uint32_t lo, uint32_t hi) {
  DevSchema S = S0;
  S.pay_lo = lo & ~15u;
  S.pay_hi = hi > 0xFFFFFFF0u ? 0xFFFFFFF0u : (hi + 15u) & ~15u;
  return S;
}

cudaError_t launch_process_raw(const DevSchema& S0, const DevTables& T0, const LaunchGeom& G,
                               const uint8_t* payload, const uint32_t* offsets, const uint64_t* ts,
                               uint32_t n, cudaStream_t st, int* launches, int big_override, uint32_t lo, uint32_t hi) {
  if (n == 0) return cudaSuccess;
  const DevSchema S = with_bounds(S0, lo, hi);
  // reference mode only counts: whatever the declared type of `progress`, the (1 024-thread) integer-table
  // kernel does it; the value the decoder hands over is never used
  const bool hist = (S.ops & BH_OP_HIST) != 0, intv = S.int_values != 0 || S.mode == BH_MODE_REFERENCE;
  // f64 values: 768 threads leave room for the 4 096-slot f64 table; in partitioned mode the table is 2 048 slots
  // anyway (the cursors need room too) and the log phase wants warps, not slots: 1 024 threads (0.478 -> 0.440 ms
  // per 12.5 M events at the C3 shape)
  const uint32_t threads = intv || T0.part_lg ? kRawThreads : kRawThreadsF64;
  size_t wbytes = (size_t)(threads / 32) * kWarpBufAlloc + kWarpBufSlack;
  const size_t cur_bytes = T0.part_lg ? ((size_t)4 << T0.part_lg) : 0;     // partition cursors, after the warp buffers
  uint32_t ns = pick_smem_slots(intv, hist, G, wbytes + cur_bytes, T0.cap, T0.ring);
  size_t smem = smem_table_bytes(ns, intv, hist) + wbytes;
  DevTables T = T0;
  T.part_smem_off = (uint32_t)smem;
  smem += cur_bytes;
  uint32_t grid = (uint32_t)G.sms;
  uint32_t need = (n + threads - 1) / threads;
  if (need < grid) grid = need;
  uint32_t chunk = (n + grid - 1) / grid;
  chunk = (chunk + 31u) & ~31u;
  const bool big = big_override < 0 ? S.big_messages != 0 : big_override != 0;
  // the reference's own schema (small messages, integer values): the kernel with the schema folded in
  const int prog = !big && !G.generic_only ? progress_schema_variant(S) : 0;
  if (prog == 1) k_process_raw<true, false, 1><<<grid, kRawThreads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (prog == 2) k_process_raw<true, false, 2><<<grid, kRawThreads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (prog == 3 && threads == kRawThreads) k_process_raw<false, false, 3><<<grid, threads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (prog == 3) k_process_raw<false, false, 3, kRawThreadsF64><<<grid, threads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (intv && !big) k_process_raw<true, false, 0><<<grid, kRawThreads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (intv) k_process_raw<true, true, 0><<<grid, kRawThreads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (!big && threads == kRawThreads) k_process_raw<false, false, 0><<<grid, threads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (!big) k_process_raw<false, false, 0, kRawThreadsF64><<<grid, threads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else if (threads == kRawThreads) k_process_raw<false, true, 0><<<grid, threads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  else k_process_raw<false, true, 0, kRawThreadsF64><<<grid, threads, smem, st>>>(S, T, payload, offsets, ts, n, chunk, ns);
  if (launches) (*launches)++;
  return cudaGetLastError();
}

cudaError_t launch_decode(const DevSchema& S0, const LaunchGeom& G, const uint8_t* payload,
                          const uint32_t* offsets, uint32_t n, uint64_t* kh, double* val,
                          int32_t* status, uint8_t* outcome, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  const DevSchema S = with_bounds(S0, 0u, 0xFFFFFFFFu);
  size_t smem = (size_t)(kRawThreads / 32) * kWarpBufAlloc + kWarpBufSlack;
  uint32_t grid = (uint32_t)G.sms;
  uint32_t need = (n + kRawThreads - 1) / kRawThreads;
  if (need < grid) grid = need;
  uint32_t chunk = (n + grid - 1) / grid;
  chunk = (chunk + 31u) & ~31u;
  if (S.big_messages) k_decode<true><<<grid, kRawThreads, smem, st>>>(S, payload, offsets, n, chunk, kh, val, status, outcome);
  else k_decode<false><<<grid, kRawThreads, smem, st>>>(S, payload, offsets, n, chunk, kh, val, status, outcome);
  return cudaGetLastError();
}

cudaError_t launch_verify_keys(const DevSchema& S0, const DevTables& T, const LaunchGeom& G, const uint8_t* payload,
                               const uint32_t* offsets, uint32_t n, cudaStream_t st, int big_override, uint32_t lo, uint32_t hi) {
  if (n == 0 || !T.dict) return cudaSuccess;
  const DevSchema S = with_bounds(S0, lo, hi);
  size_t smem = (size_t)(kRawThreads / 32) * kWarpBufAlloc + kWarpBufSlack;
  uint32_t grid = (uint32_t)G.sms;
  uint32_t need = (n + kRawThreads - 1) / kRawThreads;
  if (need < grid) grid = need;
  uint32_t chunk = (n + grid - 1) / grid;
  chunk = (chunk + 31u) & ~31u;
  const bool big = big_override < 0 ? S.big_messages != 0 : big_override != 0;
  if (big) k_verify_keys<true><<<grid, kRawThreads, smem, st>>>(S, T, payload, offsets, n, chunk);
  else k_verify_keys<false><<<grid, kRawThreads, smem, st>>>(S, T, payload, offsets, n, chunk);
  return cudaGetLastError();
}

cudaError_t launch_process_columns(const DevSchema& S, const DevTables& T0, const LaunchGeom& G,
                                   const uint64_t* kh, const uint64_t* ts, const double* val,
                                   uint64_t n, cudaStream_t st, int* launches) {
  if (n == 0) return cudaSuccess;
  const bool hist = (S.ops & BH_OP_HIST) != 0;
  const size_t cur_bytes = T0.part_lg ? ((size_t)4 << T0.part_lg) : 0;
  uint32_t ns = pick_smem_slots(false, hist, G, cur_bytes, T0.cap, T0.ring);
  size_t smem = smem_table_bytes(ns, false, hist) + 16;
  DevTables T = T0;
  T.part_smem_off = (uint32_t)smem;
  smem += cur_bytes;
  uint64_t grid = (uint64_t)G.sms;
  uint64_t need = (n + kColThreads - 1) / kColThreads;
  if (need < grid) grid = need;
  uint64_t chunk = (n + grid - 1) / grid;
  k_process_columns<<<(uint32_t)grid, kColThreads, smem, st>>>(S, T, kh, ts, val, n, chunk, ns);
  if (launches) (*launches)++;
  return cudaGetLastError();
}

// partitioned mode: fold every logged tuple into the tables and empty the log
// intv: every logged value is an int32 (the log holds only events of an integer-valued raw schema): integer shared table
cudaError_t launch_part_drain(const DevSchema& S, const DevTables& T, const LaunchGeom& G, bool intv, cudaStream_t st, int* launches) {
  if (!T.part_lg) return cudaSuccess;
  const bool hist = (S.ops & BH_OP_HIST) != 0;
  const size_t ring_bytes = ((size_t)T.ring * 8 + 15) & ~(size_t)15;
  const uint32_t grid = min((uint32_t)G.sms, 1u << T.part_lg);
  // Which fold.  Measured on the many-key shapes (tools/foldbench.sh, ms per 12.5 M events, log + fold, hashed / mirror):
  //   1 M keys, 2^21 slots:  f64 + hist 0.86 / 1.17   f64 0.63 / 0.72   int 0.55 / 0.66
  //  10 M keys, 2^24 slots:  f64 + hist 2.44 / 4.07   f64 1.42 / 1.27   int 1.36 / 1.20
  // The mirror pays one walk over the tuples per (window, part) pass and wins once a partition's keys no longer
  // fit the hashed front table (4096 slots), where the hashed fold degrades to one HBM update per tuple; with
  // histograms its parts are 1024 slots and the passes cost more than they save.
  // BEHOLDER_FOLD=0 / 1 forces one of them, BEHOLDER_FOLD_SUB_LG caps the part size (tests, experiments).
  const char* e_fold = getenv("BEHOLDER_FOLD");
  const char* e_sub = getenv("BEHOLDER_FOLD_SUB_LG");
  const uint64_t slice = T.cap >> T.part_lg;
  // (slices above 16 384 slots would be walked in 8 and more parts, one pass over the tuples each: not measured, left
  // to the hashed fold)
  const int fold_on = e_fold ? atoi(e_fold) : (!hist && slice >= 8192 && slice <= 16384);
  const int fold_sub_lg_max = e_sub ? atoi(e_sub) : 30;
  if (fold_on && slice >= 64 && T.ring <= 64 && (!hist || T.part_cap <= 65535u)) {
    const size_t per_slot = smem_slot_bytes(intv, false) + (hist ? 2u * T.hist_buckets : 0u);
    uint32_t sub_lg = 0;
    while ((2ull << sub_lg) <= slice && (int)sub_lg + 1 <= fold_sub_lg_max &&
           (2ull << sub_lg) * per_slot + 16 + ring_bytes + 2048 <= (size_t)G.smem_optin) sub_lg++;
    if (sub_lg >= 6) {
      const size_t sub = (size_t)1 << sub_lg;
      const size_t smem = ((sub * smem_slot_bytes(intv, false) + 15) & ~(size_t)15) + ring_bytes + (hist ? sub * T.hist_buckets * 2 : 0);
      if (intv) k_part_fold<true><<<grid, kDrainThreads, smem, st>>>(S, T, sub_lg);
      else k_part_fold<false><<<grid, kDrainThreads, smem, st>>>(S, T, sub_lg);
      if (launches) (*launches)++;
      return cudaGetLastError();
    }
  }
  // (no hot histogram rows in the fold: see aggregate_event.  Their 16.6 KB would also push the CTA's shared memory from
  // the 196 KB to the 228 KB carve-out and halve what is left as L1, which the fold's calls spill through.)
  uint32_t ns = 4096;
  while (ns > 64 && smem_table_bytes(ns, intv, hist, false) + 16 + (size_t)T.ring * 8 + 2048 + 1024 > (size_t)G.smem_optin) ns >>= 1;
  const size_t smem = smem_table_bytes(ns, intv, hist, false) + (size_t)T.ring * 8;
  if (intv) k_part_drain<true><<<grid, kDrainThreads, smem, st>>>(S, T, ns);
  else k_part_drain<false><<<grid, kDrainThreads, smem, st>>>(S, T, ns);
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
                          cudaStream_t st, const unsigned long long* peer_base, const unsigned long long* peer_hist_base,
                          uint32_t my_rank) {
  cudaError_t e = cudaMemsetAsync(owner_counts, 0, sizeof(unsigned long long) * n_owners, st);
  if (e != cudaSuccess) return e;
  uint32_t shift = owner_shift_of(n_owners);
  uint64_t rows = window_id == kNoWindow ? (uint64_t)T.ring * T.stride : T.stride;
  uint64_t blocks = (rows + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  // rows packed by owner need every owner's count before the first row can be placed: two passes.  One owner, or
  // owners at a fixed stride: one pass, the counts are read off the cursors afterwards.
  const bool two_pass = n_owners > 1 && owner_stride == 0;
  if (two_pass) k_export_count<<<(uint32_t)blocks, 256, 0, st>>>(T, window_id, shift, owner_counts);
  k_export_prefix<<<1, 32, 0, st>>>(n_owners, owner_counts, owner_cursor, owner_stride);
  k_export_scatter<<<(uint32_t)blocks, 256, 0, st>>>(S, T, window_id, shift, out, hist_out, cap, owner_cursor, owner_stride,
                                                     peer_base, peer_hist_base, my_rank);
  if (!two_pass) k_export_counts<<<(n_owners + 255) / 256, 256, 0, st>>>(n_owners, owner_cursor, owner_counts, owner_stride);
  return cudaGetLastError();
}

// ---- allocation-free export --------------------------------------------------------------------
// Same kernels as launch_export, but the per-owner write cursors (the only scratch the export needs)
// live in this file-scope device array instead of memory obtained from cudaMalloc.  Each GPU that loads
// the library holds these 8 KB until the process exits.  In host code the NAME below is only a handle
// for the runtime, not a device address: the address is looked up once per device with
// cudaGetSymbolAddress (with that device current) and it is that pointer the kernels receive.
static __device__ __align__(256) unsigned long long g_static_export_cursor[BH_STATIC_EXPORT_MAX_OWNERS];
static std::atomic<unsigned long long*> g_static_export_cursor_addr[64];   // per device ordinal, filled on first use (two
                                                                           // threads racing here store the same address)

cudaError_t launch_export_static(const DevSchema& S, const DevTables& T, uint64_t window_id, uint32_t n_owners,
                                 bh_record* out, uint32_t* hist_out, uint64_t cap, uint64_t owner_stride,
                                 unsigned long long* owner_counts, cudaStream_t st) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  unsigned long long* cursor = g_static_export_cursor_addr[dev].load(std::memory_order_acquire);
  if (!cursor) {                                    // set-up: first call on this device
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_export_cursor);
    if (e != cudaSuccess) return e;
    cursor = (unsigned long long*)p;
    g_static_export_cursor_addr[dev].store(cursor, std::memory_order_release);
  }
  // k_export_prefix rewrites cursor[0 .. n_owners) on every call, so nothing of the previous call's
  // contents is read; owner_counts is zeroed by launch_export (a memset, not an allocation).
  return launch_export(S, T, window_id, n_owners, out, hist_out, cap, owner_stride, owner_counts, cursor, st);
}

cudaError_t launch_merge(const DevSchema& S, const DevTables& T, const bh_record* recs,
                         const uint32_t* hist, uint64_t n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  uint64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_merge<<<(uint32_t)blocks, 256, 0, st>>>(S, T, recs, hist, n);
  return cudaGetLastError();
}

cudaError_t launch_merge_blocks(const DevSchema& S, const DevTables& T, const bh_record* recs, const uint32_t* hist,
                                uint32_t n_blocks, uint64_t stride, const unsigned long long* counts, cudaStream_t st) {
  if (n_blocks == 0 || stride == 0) return cudaSuccess;
  uint64_t blocks = (stride + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_merge_blocks<<<(uint32_t)blocks, 256, 0, st>>>(S, T, recs, hist, n_blocks, stride, counts);
  return cudaGetLastError();
}

cudaError_t launch_close_window(const DevTables& T, uint64_t window_id, cudaStream_t st) {
  uint64_t blocks = ((window_id == kNoWindow ? T.stride * T.ring : T.stride) + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_close_clear<<<(uint32_t)blocks, 256, 0, st>>>(T, window_id);       // kNoWindow: every open window
  k_close_release<<<1, 256, 0, st>>>(T, window_id, false);
  return cudaGetLastError();
}

// bh_reset_device: close every open window, zero the counters, empty the tuple log -- O(slots that hold a key),
// where launch_init_tables (used once, on fresh memory) writes every byte of every table
cudaError_t launch_reset_tables(const DevTables& T, cudaStream_t st) {
  k_close_clear<<<148 * 8, 256, 0, st>>>(T, kNoWindow);
  k_close_release<<<148, 256, 0, st>>>(T, kNoWindow, true);
  return cudaGetLastError();
}

cudaError_t launch_init_tables(const DevTables& T, cudaStream_t st) {
  cudaError_t e;
  size_t rows = (size_t)T.ring * T.stride;
  if ((e = cudaMemsetAsync(T.keys, 0xFF, rows * 8, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(T.stats, 0, rows * sizeof(Stat), st)) != cudaSuccess) return e;
  if (T.hist && (e = cudaMemsetAsync(T.hist, 0, rows * T.hist_buckets * 4, st)) != cudaSuccess) return e;
  if ((e = cudaMemsetAsync(T.ring_win, 0xFF, (size_t)T.ring * 8, st)) != cudaSuccess) return e;
  if (T.part_lg && (e = cudaMemsetAsync(T.part_fill, 0, ((size_t)T.part_ctas << T.part_lg) * 4, st)) != cudaSuccess) return e;   // logged tuples are dropped with the tables
  if (T.dict && (e = cudaMemsetAsync(T.dict, 0xFF, T.dict_cap * sizeof(DictEntry), st)) != cudaSuccess) return e;
  if (T.dict && (e = cudaMemset2DAsync(&T.dict[0].ready, sizeof(DictEntry), 0, 4, T.dict_cap, st)) != cudaSuccess) return e;
  return cudaMemsetAsync(T.counters, 0, kCounterSlots * 8, st);
}

}  // namespace bh
