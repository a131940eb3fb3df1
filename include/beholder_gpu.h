/*
 * beholder_gpu.h — C-ABI of libbeholder_gpu.so: the B200 (sm_100a) implementation of
 * beholder's progress-event consume-and-aggregate path.
 *
 * This is the drop-in boundary: the N-API addon (addon/beholder_napi.c), the Python
 * host mirror (beholder_b200/) and the tests all bind exactly these symbols.  Plain
 * pointers and sizes only; no exceptions cross it; every function returns a bh_status
 * (0 = OK, negative = error, text via bh_last_error).  There is no CPU fallback: with no
 * sm_100 device bh_create fails with BH_ERR_NO_DEVICE.
 *
 * Citations "index.js:N" are into the reference, /root/reference/index.js.
 */
#ifndef BEHOLDER_GPU_H
#define BEHOLDER_GPU_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  BH_OK = 0,
  BH_ERR_INVALID = -1,       /* bad argument / config */
  BH_ERR_NO_DEVICE = -2,     /* no CUDA device of compute capability 10.x */
  BH_ERR_CUDA = -3,          /* a CUDA call failed; see bh_last_error */
  BH_ERR_NOMEM = -4,
  BH_ERR_TABLE_FULL = -5,    /* a window's key table overflowed: events were dropped */
  BH_ERR_WINDOW_RING = -6,   /* more simultaneously open windows than window_ring */
  BH_ERR_TOO_LARGE = -7,     /* batch exceeds a u32 offset / slab limit */
  BH_ERR_UNSUPPORTED = -8,   /* schema uses a field number or type the kernels do not take */
  BH_ERR_KEY_COLLISION = -9  /* BH_F_VERIFY_KEYS: two different keys share a 64-bit hash; their rows are merged */
} bh_status;

/* declared scalar type of a known field = the protobufjs reader call the generated
 * decoder makes for it (proto.decode, index.js:129) */
enum {
  BH_K_STRING = 1, BH_K_BYTES = 2, BH_K_INT32 = 3 /* int32, enum */, BH_K_UINT32 = 4,
  BH_K_SINT32 = 5, BH_K_BOOL = 6, BH_K_FLOAT = 7, BH_K_DOUBLE = 8
};
/* role bits of a field on this path */
enum {
  BH_R_KEY = 1,      /* part of the grouping key (windowed mode), schema order */
  BH_R_STATUS = 2,   /* `status`: the enum whose name is the reference's only label (index.js:134-138) */
  BH_R_VALUE = 4     /* `progress`: the reduced value (extension; index.js:143 only prints it) */
};
enum { BH_OP_COUNT = 1, BH_OP_SUM = 2, BH_OP_MIN = 4, BH_OP_MAX = 8, BH_OP_HIST = 16 };
enum {
  BH_MODE_REFERENCE = 0,  /* index.js:127-138 exactly: key = status, cumulative count, no window */
  BH_MODE_WINDOWED = 1    /* extension: key fields x tumbling window, ops */
};
/* per-event outcome written by bh_decode_device */
enum { BH_EV_OK = 0, BH_EV_DROP_DECODE = 1, BH_EV_DROP_ENUM = 2, BH_EV_DROP_NAN = 3 };

#define BH_MAX_FIELDS 16
#define BH_MAX_ENUM 64
#define BH_MAX_KEY_PARTS 3
#define BH_MAX_FIELD_NUMBER 31
#define BH_HIST_MAX 64

typedef struct { uint32_t number, kind, role, reserved; } bh_field;

/* Replaces the construction at index.js:27-48 (Prom.new, new AMQP, proto.load): everything the
 * device path needs to know about api.TelemetryProgress.  The .proto ships in triton-core, not in
 * the reference tree, so field numbers and the enum list are supplied by the caller. */
typedef struct bh_config {
  uint32_t struct_size;        /* sizeof(bh_config), checked */
  int32_t device;              /* CUDA ordinal */
  uint32_t mode;               /* BH_MODE_* */
  uint32_t ops;                /* BH_OP_* (REFERENCE mode forces COUNT) */
  uint64_t window_ms;          /* tumbling window; 0 = one cumulative window, id 0 */
  uint32_t window_ring;        /* simultaneously open windows, power of two, <= 1024 */
  uint32_t hist_buckets;       /* <= 64; 0 -> 64 */
  int32_t hist_e0;             /* bucket = clamp(biased_exponent(v) - e0, 1, B-1); 0 for v <= 0 */
  uint32_t n_fields;
  bh_field fields[BH_MAX_FIELDS];
  uint32_t n_enum;             /* valid numbers of TelemetryStatusEntry (each 0..63); 0 = no enum drop rule */
  int32_t enum_numbers[BH_MAX_ENUM];
  uint64_t table_capacity;     /* key slots per window, power of two */
  uint64_t hash_seed;
  uint64_t slab_bytes;         /* ingest ring: payload bytes per slab (0 -> 64 MiB) */
  uint32_t slab_events;        /* ingest ring: events per slab (0 -> 1 Mi) */
  uint32_t n_slabs;            /* ingest ring depth (0 -> 3) */
  uint32_t message_bytes_hint; /* typical serialized message size; above 72 selects the kernels that stage
                                * larger messages (results do not depend on it); 0 = small messages */
  uint32_t flags;              /* BH_F_* bits */
  uint32_t max_batch_latency_ms; /* a partly filled ingest slab is submitted once its oldest message is this old
                                * (checked by bh_poll_committed and by the completion thread); 0 -> 5 ms,
                                * BH_NO_BATCH_DEADLINE -> only when full / bh_submit / bh_flush.  Bounds the ack
                                * latency of index.js:151,154 under the prefetch limit of index.js:43. */
  uint32_t part_region_tuples; /* partitioned aggregation (BH_F_PARTITIONED): 16-byte tuples per (SM, partition) log
                                * region; 0 -> 2^20 / partitions.  The log takes SMs x partitions x this x 16 bytes of HBM
                                * (2.5 GB on a B200 by default). */
  uint64_t debug_hash_mask;    /* test hook: key hashes are ANDed with this (0 -> all ones) so that collisions can be made */
} bh_config;
#define BH_NO_BATCH_DEADLINE 0xFFFFFFFFu

/* bh_config.flags */
enum {
  BH_F_GENERIC_KERNELS = 1,    /* never launch the fused kernel that has the reference's own schema compiled in
                                * (api.TelemetryProgress keyed by mediaId + status, int32 progress): same results,
                                * slower; for tests.  The environment variable BEHOLDER_GENERIC_KERNELS=1 does the same. */
  BH_F_PARTITIONED = 4,        /* large key sets (BASELINE.json configs 3 and 4; no counterpart in the reference, whose
                                * counter has a handful of labels, index.js:136-138): events that miss the per-SM
                                * shared-memory table are logged as 16-byte tuples partitioned by table slot and folded
                                * into the HBM tables later, one partition per SM, instead of one random HBM update per
                                * event.  Same results.  On by itself when table_capacity >= 2^18; this flag forces it
                                * for smaller tables, BH_F_NO_PARTITION keeps it off.  While it is on, launches that feed
                                * one context must be issued on one stream at a time (the library re-orders itself
                                * across a change of stream), and reads (snapshot, export, counters) fold the log first.
                                * While it is on, a key's table slot is probed for inside its partition's slice of the
                                * table only (table_capacity / partitions slots, kept >= 256), so dropped_table_full can
                                * start a little before the whole table is full: size table_capacity for a load factor
                                * <= 0.75 as for any open-addressing table.  Forcing it needs
                                * table_capacity >= 256 x max(2, window_ring) (BH_ERR_INVALID otherwise). */
  BH_F_NO_PARTITION = 8,
  BH_F_VERIFY_KEYS = 16,       /* Full-key identity (collisions are detected, never silently merged).
                                * Rows are identified by the 64-bit hash of their key; with this flag every raw batch is
                                * also run through a check kernel that keeps, per hash, the key bytes first seen with it
                                * (a dictionary of 2 x table_capacity entries in HBM) and compares every later event's
                                * key against them.  A mismatch is counted in bh_counters.key_collisions and bh_flush
                                * returns BH_ERR_KEY_COLLISION.  Costs a second decode pass; keys must be one string field
                                * (<= 48 bytes) optionally followed by the status.  Not available for bh_push_columns
                                * (no key bytes there). */
  BH_F_NO_NUMA_BIND = 2        /* by default the pinned ingest ring (and bh_host_alloc buffers) are allocated while the
                                * calling thread is confined to the CPUs of the GPU's NUMA node, and the thread's
                                * affinity is restored afterwards; this flag (or BEHOLDER_NO_NUMA_BIND=1) turns it off */
};

/* One aggregate row.  REFERENCE mode: key_hash = status number, window_id = 0, count = the value of
 * beholder_progress_updates_total{status=<name>} (index.js:30-34). */
typedef struct bh_record {
  uint64_t key_hash, window_id, count;
  double sum, min, max;
} bh_record;

typedef struct bh_counters {
  uint64_t events_ok;          /* aggregated */
  uint64_t dropped_decode;     /* proto.decode threw -> warn + ack, not counted (index.js:149-152) */
  uint64_t dropped_enum;       /* enumToString undefined -> TypeError -> same catch (index.js:134,137) */
  uint64_t dropped_nan;        /* extension */
  uint64_t dropped_table_full; /* -> BH_ERR_TABLE_FULL */
  uint64_t dropped_window;     /* -> BH_ERR_WINDOW_RING */
  uint64_t smem_spills;        /* events that bypassed the shared-memory table (diagnostic) */
  uint64_t slow_decodes;       // messages that left the decoder's straight path (diagnostic)
  uint64_t key_collisions;     /* BH_F_VERIFY_KEYS: events whose key bytes differ from the key first seen with their hash */
} bh_counters;

typedef struct bh_ctx bh_ctx;

/* ---- lifecycle: replaces index.js:27-48 ---- */
int bh_create(const bh_config* cfg, bh_ctx** out);
void bh_destroy(bh_ctx* ctx);
const char* bh_last_error(bh_ctx* ctx);   /* ctx may be NULL: last bh_create error of this thread */
int bh_device_count(void);                /* CUDA devices of compute capability 10.x */

/* ---- ingest: replaces the body of `amqp.listen('v1.telemetry.progress', async rmsg => ...)`,
 * index.js:127-138.  The caller keeps rmsg and acks it once bh_poll_committed passes its sequence
 * number (index.js:151,154: every message is acked exactly once, dropped or not). ---- */
/* one message: rmsg.message.content (index.js:129).  Copied into the pinned ring before returning. */
int bh_push(bh_ctx* ctx, const uint8_t* payload, uint32_t len, uint64_t ts_ms);
/* n messages: slab + n+1 byte offsets (offsets[0] may be non-zero) + n timestamps (may be NULL = 0) */
int bh_push_batch(bh_ctx* ctx, const uint8_t* slab, const uint32_t* offsets, const uint64_t* ts_ms, uint32_t n);
/* same, but slab/offsets/ts_ms are in pinned host memory (bh_host_alloc) and stay valid until
 * committed: copied to the device straight from the caller's buffers. */
int bh_push_batch_pinned(bh_ctx* ctx, const uint8_t* slab, const uint32_t* offsets, const uint64_t* ts_ms, uint32_t n);
/* bh_push_batch_pinned with compact sidecars.  End to end the path is bound by the PCIe link, so what
 * travels with each message matters: a u16 LENGTH and an i16 timestamp DELTA instead of a u32 offset and a u64
 * timestamp take a 64-byte event from 76 to 68 bytes on the link; the offsets and timestamps are rebuilt on the
 * GPU.  Messages lie back to back from `slab` on (caller-pinned memory, as are all arrays; they must stay valid
 * until committed).  Per block of BH_COMPACT_BLOCK = 65 536 messages the caller gives block_offsets[k] = byte
 * offset of message k * 65 536 from `slab` (one extra entry at the end = total bytes) and ts_block[k] = a base
 * timestamp; ts_ms of message i = ts_block[i >> 16] + ts_delta[i].  A message longer than 65 535 bytes or a
 * timestamp further than +-32 767 ms from its block's base does not fit: use bh_push_batch_pinned for such a batch.
 * ts_block / ts_delta may be NULL where timestamps are not used.  slab_events must be at least 65 536. */
#define BH_COMPACT_BLOCK_LG 16
int bh_push_batch_compact(bh_ctx* ctx, const uint8_t* slab, const uint16_t* lengths, const uint64_t* block_offsets,
                          const uint64_t* ts_block, const int16_t* ts_delta, uint32_t n);
/* pre-decoded columns (the counter side alone: metrics.progress_updates_total.inc, index.js:136-138) */
int bh_push_columns(bh_ctx* ctx, const uint64_t* key_hash, const uint64_t* ts_ms, const double* value, uint64_t n);
/* submit the partial slab and wait until everything pushed so far is aggregated */
int bh_flush(bh_ctx* ctx, uint64_t* committed_seq);
/* non-blocking: number of events whose aggregation has completed on the device.  Also submits the partly
 * filled slab when its oldest message has waited max_batch_latency_ms. */
int bh_poll_committed(bh_ctx* ctx, uint64_t* seq);
/* submit the partly filled slab now, without waiting for its aggregation (it may wait for a free ring slab).
 * The handler calls it when half the broker's prefetch window (index.js:43) is waiting for an ack. */
int bh_submit(bh_ctx* ctx);
/* Completion thread: `cb(user, committed_seq)` is called from a thread the library owns whenever the
 * committed sequence number advances; the same thread enforces max_batch_latency_ms, so messages are acked
 * (index.js:151,154) even when the producer never calls in again.  cb runs with no library lock held and may call
 * any bh_* function except bh_destroy and bh_set_commit_callback.  cb = NULL stops the thread.  The N-API shim
 * forwards cb through napi_threadsafe_function so that rmsg.ack() runs on the JS thread. */
typedef void (*bh_commit_cb)(void* user, uint64_t committed_seq);
int bh_set_commit_callback(bh_ctx* ctx, bh_commit_cb cb, void* user);
int bh_host_alloc(void** out, size_t bytes);   /* pinned host memory for bh_push_batch_pinned */
int bh_host_free(void* p);
/* Page-lock memory the caller already owns (a pool of Node Buffers that is reused, an mmap'ed capture file), so
 * that bh_push_batch_pinned can take it: the GPU then reads it in place and the copy into the library's ring --
 * the limit of bh_push_batch from ordinary memory, one pass over host DRAM at memcpy speed -- disappears.
 * Registering costs on the order of a millisecond per 10 MB: do it once per buffer, not per batch.  The memory
 * must stay mapped until bh_host_unregister. */
int bh_host_register(void* p, size_t bytes);
int bh_host_unregister(void* p);

/* ---- read side: what Prom.expose() serves (index.js:28) ---- */
/* rows of one window (REFERENCE mode: window 0) in unspecified order; *n = rows available */
int bh_snapshot(bh_ctx* ctx, uint64_t window_id, bh_record* out, uint64_t cap, uint64_t* n);
/* same plus per-row log2 histograms, hist_out[cap * hist_buckets] */
int bh_snapshot_hist(bh_ctx* ctx, uint64_t window_id, bh_record* out, uint32_t* hist_out, uint64_t cap, uint64_t* n);
int bh_list_windows(bh_ctx* ctx, uint64_t* out, uint32_t cap, uint32_t* n);
/* drop a window's table so its ring slot can be reused */
int bh_close_window(bh_ctx* ctx, uint64_t window_id);
/* merge rows (and their histograms, may be NULL) taken earlier with bh_snapshot_hist back into the
 * tables: restart after a crash, or hand-over between processes.  The reference keeps its counters in
 * memory only and loses them on restart (index.js:32, "in this processes lifetime"). */
int bh_restore(bh_ctx* ctx, const bh_record* rows, const uint32_t* hist, uint64_t n);
/* drops / diagnostics; returns BH_ERR_TABLE_FULL / BH_ERR_WINDOW_RING if events were lost that way */
int bh_get_counters(bh_ctx* ctx, bh_counters* out);

/* ---- device-resident entry points (inputs/outputs already in HBM; asynchronous on `stream`,
 * a cudaStream_t passed as void*, NULL = the context's own stream).  Used by bench.py's
 * kernel-only measurement and by the multi-GPU exchange. ---- */
/* fused proto.decode + enumToString + counter update (index.js:129-138) over one batch.
 * d_payload must be 16-byte aligned and readable up to offsets[n] rounded up to 16. */
int bh_process_device(bh_ctx* ctx, const uint8_t* d_payload, const uint32_t* d_offsets,
                      const uint64_t* d_ts_ms, uint32_t n, void* stream);
int bh_process_columns_device(bh_ctx* ctx, const uint64_t* d_key_hash, const uint64_t* d_ts_ms,
                              const double* d_value, uint64_t n, void* stream);
/* proto.decode alone (index.js:129): payloads -> key-hash / value / status / outcome columns */
int bh_decode_device(bh_ctx* ctx, const uint8_t* d_payload, const uint32_t* d_offsets, uint32_t n,
                     uint64_t* d_key_hash, double* d_value, int32_t* d_status, uint8_t* d_outcome,
                     void* stream);
/* compact one window (window_id = BH_ALL_WINDOWS: every open window) into d_records grouped by
 * owner = key_hash >> (64 - log2(n_owners)).  d_owner_counts[n_owners] receives the rows each owner
 * has.  owner_stride = 0: rows packed, owner o starts at sum(counts[<o]).  owner_stride > 0: owner o
 * starts at o * owner_stride (fixed-size all-to-all; the caller zero-fills d_records first, rows with
 * count 0 are padding; counts[o] > owner_stride means rows were cut).  d_hist may be NULL.
 * No reference counterpart (the reference is one process). */
#define BH_ALL_WINDOWS 0xFFFFFFFFFFFFFFFFull
int bh_export_window_device(bh_ctx* ctx, uint64_t window_id, uint32_t n_owners, bh_record* d_records,
                            uint32_t* d_hist, uint64_t cap, uint64_t owner_stride,
                            uint64_t* d_owner_counts, void* stream);
/* Multi-GPU window close without a staging copy or an all-to-all (no counterpart in the reference, which is one
 * process): like bh_export_window_device with owner_stride > 0, but the rows of owner o are stored straight
 * into GPU o's receive buffer over NVLink.  d_peer_bases[o] = device-visible address of owner o's buffer
 * (n_owners blocks of owner_stride bh_record rows; peer-mapped memory, e.g. CUDA IPC / torch symmetric
 * memory); this rank writes block `my_rank` of every buffer.  The receive buffers must be zeroed (count 0 =
 * padding) before any rank exports, and all ranks must have finished before an owner merges: the caller puts
 * a cross-GPU barrier on the stream before and after (beholder_b200.dist.PeerWindowExchange).
 * d_peer_hist_bases[o] = owner o's histogram receive buffer (same row numbering, hist_buckets u32 per
 * row); required when the context keeps histograms, NULL otherwise. */
int bh_export_window_peers(bh_ctx* ctx, uint64_t window_id, uint32_t n_owners, const uint64_t* d_peer_bases,
                           const uint64_t* d_peer_hist_bases, uint32_t my_rank, uint64_t owner_stride,
                           uint64_t* d_owner_counts, void* stream);

/* bh_export_window_device for callers that must not allocate device memory.  It has no counterpart in
 * the reference.  Who needs it: a host that fixes its GPU memory budget at start-up and forbids the
 * driver allocator afterwards (a serving process sharing the GPU under a hard memory cap, a real-time
 * loop where cudaMalloc's implicit device synchronisation is unacceptable), or a caller that is inside
 * a phase where the allocator may not be entered at all.  Arguments, output layout and results are
 * those of bh_export_window_device; the difference is where the scratch (one write cursor per owner)
 * lives: in an array built into the library image, BH_STATIC_EXPORT_MAX_OWNERS * 8 bytes = 8 KB that
 * every GPU which loads the library keeps until the process exits.  The call allocates nothing: no
 * cudaMalloc / cudaFree, no allocating library routine.
 *   - n_owners above BH_STATIC_EXPORT_MAX_OWNERS is refused with BH_ERR_TOO_LARGE.
 *   - The first call on a device looks up the array's device address (cudaGetSymbolAddress); make that
 *     first call before starting any stream capture on that device.
 *   - All calls on one GPU share the one array.  Each call rewrites the cursors it uses before reading
 *     them, so earlier contents never leak in, but two calls that overlap on the GPU (different streams
 *     or contexts) race on it: keep them ordered. */
#define BH_STATIC_EXPORT_MAX_OWNERS 1024
int bh_export_window_static(bh_ctx* ctx, uint64_t window_id, uint32_t n_owners, bh_record* d_records,
                            uint32_t* d_hist, uint64_t cap, uint64_t owner_stride,
                            uint64_t* d_owner_counts, void* stream);
/* bh_close_window on the caller's stream (after the window has been exported); BH_ALL_WINDOWS closes every open one */
int bh_close_window_device(bh_ctx* ctx, uint64_t window_id, void* stream);
/* partitioned aggregation (BH_F_PARTITIONED): fold everything logged so far into the tables now.  Reads do this
 * by themselves; a caller that times or overlaps work can choose the moment.  No-op otherwise. */
int bh_drain_device(bh_ctx* ctx, void* stream);
/* drop every window and zero the drop counters (the counters of index.js:32 "reset on restart") */
int bh_reset_device(bh_ctx* ctx, void* stream);
/* merge partial rows (from other GPUs) into this context's tables */
int bh_merge_records_device(bh_ctx* ctx, const bh_record* d_records, const uint32_t* d_hist,
                            uint64_t n, void* stream);
int bh_sync(bh_ctx* ctx);
/* like bh_merge_records_device for rows that arrived in n_blocks blocks of block_stride rows (one per sending GPU):
 * block b holds min(d_block_counts[b], block_stride) valid rows; nothing beyond is read */
int bh_merge_blocks_device(bh_ctx* ctx, const bh_record* d_records, const uint32_t* d_hist, uint32_t n_blocks,
                           uint64_t block_stride, const uint64_t* d_block_counts, void* stream);
/* the context's own stream (a cudaStream_t): what `stream = NULL` means in the entry points above.  To name CUDA's
 * legacy default stream pass (void*)1 (cudaStreamLegacy), never NULL. */
void* bh_stream(bh_ctx* ctx);
/* make the context's own stream wait for everything queued on `stream` so far */
int bh_stream_wait(bh_ctx* ctx, void* stream);

/* ---- multi-GPU: key-hash sharded window close inside the library (SURVEY 8 b/e).  No counterpart in the
 * reference, which is the single process of index.js:23-48; this is what lets that one process drive several
 * GPUs.  Events are spread over the GPUs by arrival, `local` contexts aggregate what they see, a key's merged
 * row lives in the `owned` context of GPU (key_hash >> (64 - log2 n)).  bh_group_close_window runs, on every GPU
 * and asynchronously on its stream: export rows straight into the owners' receive buffers over NVLink ->
 * cross-GPU flag barrier -> merge -> recycle the window.  owner_stride = rows one GPU can send one owner per
 * close (exceeding it is reported by bh_group_sync as BH_ERR_TOO_LARGE, never silent). ---- */
typedef struct bh_group bh_group;
typedef struct bh_member bh_member;
int bh_create_group(const bh_config* local_cfg, const bh_config* owned_cfg, const int* devices, uint32_t n,
                    uint64_t owner_stride, bh_group** out);          /* the configs' `device` fields are ignored */
void bh_group_destroy(bh_group* g);
uint32_t bh_group_size(bh_group* g);
bh_ctx* bh_group_local(bh_group* g, uint32_t rank);   /* push / process this GPU's share of the events here */
bh_ctx* bh_group_owned(bh_group* g, uint32_t rank);   /* snapshot the keys this GPU owns here */
bh_member* bh_group_member(bh_group* g, uint32_t rank);
int bh_group_close_window(bh_group* g, uint64_t window_id);          /* BH_ALL_WINDOWS: every open window */
int bh_group_sync(bh_group* g);                        /* waits for every GPU; reports cut rows */
const char* bh_group_last_error(void);
/* One member per process (one process per GPU, e.g. under torchrun): create, exchange the 64-byte handles of
 * bh_member_ipc_handle between the ranks by any means, connect, then bh_member_close_window on every rank. */
#define BH_IPC_HANDLE_BYTES 64
int bh_member_create(bh_ctx* local, bh_ctx* owned, int device, uint32_t rank, uint32_t world, uint64_t owner_stride,
                     uint32_t hist_buckets, bh_member** out);
void bh_member_destroy(bh_member* m);
int bh_member_ipc_handle(bh_member* m, void* out64);
int bh_member_connect_ipc(bh_member* m, const void* handles /* world x 64 bytes, by rank */);
int bh_member_connect_ptrs(bh_member* m, void* const* areas /* world device pointers (same process) */);
void* bh_member_area(bh_member* m);
int bh_member_close_window(bh_member* m, uint64_t window_id, void* stream);
int bh_member_check(bh_member* m, uint64_t* owner_counts_out /* world values of the last close, may be NULL */);

/* ---- synthetic event streams (bench/test input) ---- */
typedef struct bh_synth_config {
  uint32_t struct_size;
  uint32_t n_status;           /* events use status numbers 0..n_status-1 */
  uint64_t seed;
  uint64_t n_keys;             /* distinct (mediaId, status) keys */
  uint64_t t0_ms;              /* first timestamp */
  uint32_t rate_per_ms;        /* events per millisecond of stream time */
  uint32_t jitter_ms;          /* ts = t0 + g / rate + U[-jitter, +jitter] */
  uint32_t payload_bytes;      /* pad every event to exactly this size when it fits; 0 = natural */
  uint32_t value_f64;          /* 0: progress = int32 varint 0..100; 1: double, random mantissa/exponent */
  int32_t value_exp_lo;        /* value_f64: exponent drawn from [lo, lo + span) */
  uint32_t value_exp_span;
  uint32_t value_signed;       /* value_f64: random sign */
  uint32_t malformed_ppm;      /* corrupted events per million */
  uint32_t unknown_enum_ppm;   /* events whose status has no enum name, per million */
  uint32_t zipf;               /* 0 uniform; 1: keys drawn through d_zipf_cdf */
} bh_synth_config;
/* bytes event g occupies (host helper, same function the device uses) */
uint32_t bh_synth_event_size(const bh_synth_config* c, uint64_t g, const uint64_t* zipf_cdf);
/* write event g into out (>= bh_synth_event_size bytes); returns its size and timestamp */
uint32_t bh_synth_event_host(const bh_synth_config* c, uint64_t g, const uint64_t* zipf_cdf,
                             uint8_t* out, uint64_t* ts_ms);
/* generate events [first, first+n) on the device: d_offsets[n+1], d_ts[n], payload written at
 * d_payload (capacity checked against the total, returned in *total_bytes). d_zipf_cdf[n_keys]
 * (u64 thresholds) is required when zipf != 0. */
int bh_synth_generate_device(int device, const bh_synth_config* c, uint64_t first, uint32_t n,
                             const uint64_t* d_zipf_cdf, uint8_t* d_payload, uint64_t payload_cap,
                             uint32_t* d_offsets, uint64_t* d_ts, uint64_t* total_bytes, void* stream);
/* columns of the same stream without the wire format: key hash as the fused path would compute it */
int bh_synth_columns_device(int device, const bh_synth_config* c, uint64_t hash_seed, uint64_t first,
                            uint64_t n, const uint64_t* d_zipf_cdf, uint64_t* d_key_hash,
                            uint64_t* d_ts, double* d_value, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * AMQP 0-9-1 delivery tap.  Replaces, for the progress queue, what happens between
 * the broker socket and the handler of index.js:127: amqplib 0.5.5 (yarn.lock:104-121, not vendored) parses
 * frames, builds one message object per delivery and triton-core calls `fn(rmsg)`; here the raw
 * broker -> client bytes of one connection are scanned once and the bodies of `basic.deliver` messages
 * land in a slab + offsets pair that bh_push_batch / bh_push_batch_pinned take as is -- no per-message
 * object.  Host code only (no GPU needed); frame grammar from the published AMQP 0-9-1 specification:
 *   frame   = type:u8 channel:u16 size:u32 payload[size] 0xCE          (integers big-endian)
 *   METHOD (1): class:u16 method:u16 args;  basic.deliver = (60,60): consumer-tag shortstr,
 *               delivery-tag u64, redelivered octet, exchange shortstr, routing-key shortstr
 *   HEADER (2): class:u16 weight:u16 body-size:u64 property flags/values
 *   BODY   (3): body bytes (a message's body may span several frames)      HEARTBEAT (8): empty
 * Content of basic.return (60,50) and basic.get-ok (60,71) is skipped; every other method and
 * heartbeats are skipped and counted.  The tap never answers the broker: acks stay with the AMQP
 * client, which uses the delivery tags reported here once bh_poll_committed has passed the batch
 * (`channel.ack({fields:{deliveryTag}}, allUpTo = true)`, the batched form of index.js:154). */
typedef struct bh_amqp bh_amqp;
typedef struct bh_amqp_delivery {
  uint64_t delivery_tag;
  uint32_t body_bytes;
  uint16_t channel;
  uint16_t redelivered;
} bh_amqp_delivery;
typedef struct bh_amqp_stats {
  uint64_t frames, deliveries, body_bytes, heartbeats, other_methods, skipped_content, filtered;
} bh_amqp_stats;
/* consumer_tag: only deliveries with this consumer tag are reported (NULL or "" = all of them);
 * max_body: a larger announced body size is a protocol error (0 = 16 MiB) */
int bh_amqp_create(bh_amqp** out, const char* consumer_tag, uint64_t max_body);
void bh_amqp_destroy(bh_amqp* a);
/* Scan bytes[0, len).  Whole frames are consumed; *consumed is where the caller's next call must start
 * (an incomplete frame at the end is left; partly received messages are remembered inside `a`).
 * Each completed delivery i gets deliveries[i] and its body at slab[offsets[i], offsets[i+1]);
 * offsets[0] = 0, offsets has room for cap + 1 entries.  The scan stops early, before the frame that
 * completes a message, when that message would not fit (slab_cap bytes, cap messages): flush what was
 * returned and call again.  BH_ERR_INVALID = protocol error (bad frame end, unknown frame type, content
 * without a method, body larger than announced): the connection must be dropped, as amqplib does. */
int bh_amqp_feed(bh_amqp* a, const uint8_t* bytes, uint64_t len, uint8_t* slab, uint64_t slab_cap,
                 uint32_t* offsets, bh_amqp_delivery* deliveries, uint32_t cap, uint64_t* consumed, uint32_t* n);
int bh_amqp_get_stats(const bh_amqp* a, bh_amqp_stats* out);
const char* bh_amqp_last_error(const bh_amqp* a);

#ifdef __cplusplus
}
#endif
#endif
