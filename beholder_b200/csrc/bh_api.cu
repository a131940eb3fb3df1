// bh_api.cu — the C-ABI of libbeholder_gpu.so (include/beholder_gpu.h): context, pinned ingest
// ring, read side.  Host code only; kernels are in bh_kernels.cu.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <ctype.h>
#ifdef __linux__
#include <sched.h>
#endif

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "bh_kernels.cuh"

using namespace bh;

namespace {

thread_local std::string g_create_error;

struct Slab {
  uint8_t* h_payload = nullptr;
  uint32_t* h_offsets = nullptr;
  uint64_t* h_ts = nullptr;
  uint8_t* d_payload = nullptr;
  uint32_t* d_offsets = nullptr;
  uint64_t* d_ts = nullptr;
  uint32_t n = 0;
  uint64_t bytes = 0;
  bool any_ts = false;
  bool in_flight = false;
  uint64_t seq_end = 0;
  cudaEvent_t copied = nullptr, done = nullptr;
  // compact sidecars (bh_push_batch_compact), allocated on first use
  uint16_t* d_len = nullptr; int16_t* d_tsd = nullptr;
};

}  // namespace

struct bh_ctx {
  bh_config cfg;
  int device = 0;
  DevSchema S;
  DevTables T;
  LaunchGeom G;
  cudaStream_t st_compute = nullptr;
  cudaStream_t st_copy[2] = {nullptr, nullptr};
  std::vector<Slab> slabs;
  uint32_t cur = 0;
  uint64_t slab_bytes = 0;
  uint32_t slab_events = 0;
  uint64_t pushed = 0, committed = 0;
  unsigned long long* d_owner_counts = nullptr;   // [1024]
  unsigned long long* d_owner_cursor = nullptr;   // [1024]
  bh_record* d_export = nullptr;
  uint32_t* d_export_hist = nullptr;
  uint64_t export_cap = 0;
  unsigned long long* h_small = nullptr;          // pinned, 1024 u64
  int launches = 0;
  // partitioned mode (DevTables::part_*): tuples logged since the last drain (upper bound), the fill at which a
  // drain is forced before the next launch, and the stream the log was last written on
  uint64_t part_pending = 0, part_limit = 0;
  bool part_f64_tuples = false;                   // the log holds values that did not come from an int32 field
  cudaStream_t part_stream = nullptr;
  cudaEvent_t part_ev = nullptr;
  cudaEvent_t wait_ev = nullptr;                  // bh_stream_wait
  std::string err;
  // One lock around every entry point: the producer thread (the Node main thread / the asyncio loop) and the
  // library's own completion thread are the two parties.  Recursive because entry points call each other.
  std::recursive_mutex mu;
  std::condition_variable_any cv;
  std::thread completer;                          // started by bh_set_commit_callback
  bool stop = false;
  bh_commit_cb cb = nullptr;
  void* cb_user = nullptr;
  uint64_t reported = 0;                          // last sequence number handed to cb
  std::chrono::steady_clock::time_point slab_t0;  // arrival of the first message of the partial slab
  std::chrono::microseconds max_latency{5000};
  bool no_deadline = false;
};
#define BH_LOCK(c) std::lock_guard<std::recursive_mutex> lock__((c)->mu)

namespace {

int fail(bh_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  if (c) c->err = buf; else g_create_error = buf;
  return code;
}

#define CU(c, call)                                                                                \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess) return fail((c), BH_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); \
  } while (0)

bool is_pow2(uint64_t v) { return v && !(v & (v - 1)); }

int build_schema(const bh_config& cfg, DevSchema& S) {
  memset(&S, 0, sizeof S);
  uint32_t parts = 0;
  bool have_status = false;
  S.int_values = 1;
  S.fast_key0 = S.fast_status = S.fast_value = 0xFFFFFFFFu;
  for (uint32_t i = 0; i < cfg.n_fields; i++) {
    const bh_field& f = cfg.fields[i];
    if (f.number < 1 || f.number > BH_MAX_FIELD_NUMBER) return fail(nullptr, BH_ERR_UNSUPPORTED, "field number %u outside 1..%d", f.number, BH_MAX_FIELD_NUMBER);
    if (f.kind < BH_K_STRING || f.kind > BH_K_DOUBLE) return fail(nullptr, BH_ERR_UNSUPPORTED, "field %u: unsupported declared type %u", f.number, f.kind);
    if (S.fld[f.number]) return fail(nullptr, BH_ERR_INVALID, "field %u declared twice", f.number);
    S.fld[f.number] = f.kind | ((f.role & 0xFFu) << 8);
    bool is_len = f.kind == BH_K_STRING || f.kind == BH_K_BYTES;
    if (f.role & BH_R_STATUS) {
      if (is_len || f.kind >= BH_K_FLOAT) return fail(nullptr, BH_ERR_INVALID, "status field %u must be a 32-bit varint type", f.number);
      have_status = true;
    }
    if ((f.role & BH_R_VALUE) && is_len) return fail(nullptr, BH_ERR_INVALID, "value field %u must be numeric", f.number);
    if ((f.role & BH_R_VALUE) && !(f.kind == BH_K_INT32 || f.kind == BH_K_SINT32 || f.kind == BH_K_BOOL)) S.int_values = 0;
    if (f.role & BH_R_KEY) {
      if (f.kind == BH_K_FLOAT || f.kind == BH_K_DOUBLE) return fail(nullptr, BH_ERR_UNSUPPORTED, "key field %u: float keys not supported", f.number);
      if (parts >= BH_MAX_KEY_PARTS) return fail(nullptr, BH_ERR_UNSUPPORTED, "more than %d key fields", BH_MAX_KEY_PARTS);
      S.fld[f.number] |= (parts + 1) << 16;
      if (parts == 0 && f.kind == BH_K_STRING && f.role == BH_R_KEY) S.fast_key0 = f.number;
      S.part_field[parts] = (uint8_t)f.number;
      S.part_is_len[parts] = is_len;
      parts++;
    }
  }
  for (uint32_t i = 0; i < cfg.n_fields; i++) {
    const bh_field& f = cfg.fields[i];
    if ((f.role & BH_R_STATUS) && !(f.role & BH_R_VALUE) && (f.kind == BH_K_INT32 || f.kind == BH_K_UINT32)) {
      S.fast_status = f.number;
      S.fast_status_part = (f.role & BH_R_KEY) ? (S.fld[f.number] >> 16) : 0;
    }
    if (f.role == BH_R_VALUE && (f.kind == BH_K_INT32 || f.kind == BH_K_DOUBLE)) {
      S.fast_value = f.number;
      S.fast_value_f64 = f.kind == BH_K_DOUBLE;
    }
  }
  S.tag_key0 = S.fast_key0 < 16 ? (S.fast_key0 << 3 | 2u) : 0xFFFFFFFFu;
  S.tag_status = S.fast_status < 16 ? (S.fast_status << 3) : 0xFFFFFFFFu;
  S.tag_value = S.fast_value < 16 ? (S.fast_value << 3 | (S.fast_value_f64 ? 1u : 0u)) : 0xFFFFFFFFu;
  S.big_messages = cfg.message_bytes_hint > 72;
  for (uint32_t n = 0; n < 16; n++) {
    if (S.fld[n] == (uint32_t)BH_K_STRING) S.skip_str_mask |= 1u << n;
    if (S.fld[n] == 0) S.skip_unk_mask |= 1u << n;
  }
  S.n_parts = parts;
  S.mode = cfg.mode;
  if (cfg.mode == BH_MODE_REFERENCE) {
    if (!have_status) return fail(nullptr, BH_ERR_INVALID, "REFERENCE mode needs a field with BH_R_STATUS");
    S.status_part = 1;
    S.ops = BH_OP_COUNT;
  } else if (cfg.mode == BH_MODE_WINDOWED) {
    S.ops = cfg.ops | BH_OP_COUNT;
    S.identity_key = parts == 1 && !S.part_is_len[0];
  } else {
    return fail(nullptr, BH_ERR_INVALID, "unknown mode %u", cfg.mode);
  }
  if (cfg.n_enum > BH_MAX_ENUM) return fail(nullptr, BH_ERR_INVALID, "n_enum > %d", BH_MAX_ENUM);
  for (uint32_t i = 0; i < cfg.n_enum; i++) {
    if (cfg.enum_numbers[i] < 0 || cfg.enum_numbers[i] > 63) return fail(nullptr, BH_ERR_UNSUPPORTED, "enum number %d outside 0..63", cfg.enum_numbers[i]);
    S.enum_mask |= 1ull << cfg.enum_numbers[i];
  }
  S.has_enum = cfg.n_enum > 0 && have_status;
  S.hist_buckets = cfg.hist_buckets ? cfg.hist_buckets : BH_HIST_MAX;
  if (S.hist_buckets < 2 || S.hist_buckets > BH_HIST_MAX) return fail(nullptr, BH_ERR_INVALID, "hist_buckets must be 2..%d", BH_HIST_MAX);
  S.hist_e0 = cfg.hist_e0;
  S.hash_seed = cfg.hash_seed;
  S.hash_mask = cfg.debug_hash_mask ? cfg.debug_hash_mask : ~0ull;
  S.has_window = cfg.mode == BH_MODE_WINDOWED && cfg.window_ms > 0;
  if (S.has_window) { S.wdiv = make_div64(cfg.window_ms); S.window_ms = cfg.window_ms; }
  return BH_OK;
}

int slab_alloc(bh_ctx* c, Slab& s) {
  CU(c, cudaHostAlloc((void**)&s.h_payload, c->slab_bytes + 64, cudaHostAllocDefault));
  CU(c, cudaHostAlloc((void**)&s.h_offsets, ((size_t)c->slab_events + 1) * 4, cudaHostAllocDefault));
  CU(c, cudaHostAlloc((void**)&s.h_ts, (size_t)c->slab_events * 8, cudaHostAllocDefault));
  CU(c, cudaMalloc((void**)&s.d_payload, c->slab_bytes + 64));
  CU(c, cudaMalloc((void**)&s.d_offsets, ((size_t)c->slab_events + 1) * 4));
  CU(c, cudaMalloc((void**)&s.d_ts, (size_t)c->slab_events * 8));
  CU(c, cudaEventCreateWithFlags(&s.copied, cudaEventDisableTiming));
  CU(c, cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
  s.h_offsets[0] = 0;
  return BH_OK;
}

void slab_free(Slab& s) {
  if (s.h_payload) cudaFreeHost(s.h_payload);
  if (s.h_offsets) cudaFreeHost(s.h_offsets);
  if (s.h_ts) cudaFreeHost(s.h_ts);
  if (s.d_payload) cudaFree(s.d_payload);
  if (s.d_offsets) cudaFree(s.d_offsets);
  if (s.d_ts) cudaFree(s.d_ts);
  if (s.d_len) cudaFree(s.d_len);
  if (s.d_tsd) cudaFree(s.d_tsd);

  if (s.copied) cudaEventDestroy(s.copied);
  if (s.done) cudaEventDestroy(s.done);
  s = Slab();
}

// wait until slab i is free for refilling
int slab_acquire(bh_ctx* c, uint32_t i) {
  Slab& s = c->slabs[i];
  if (s.in_flight) {
    CU(c, cudaEventSynchronize(s.done));
    s.in_flight = false;
    if (s.seq_end > c->committed) c->committed = s.seq_end;
  }
  return BH_OK;
}

// Partitioned mode.  part_drain folds the logged tuples into the tables on the stream that wrote them and makes
// `st` wait for it; every entry point that reads, exports, closes or resets tables calls it first.
int part_drain(bh_ctx* c, cudaStream_t st) {
  if (!c->T.part_lg || c->part_pending == 0) return BH_OK;
  cudaStream_t ps = c->part_stream ? c->part_stream : c->st_compute;
  CU(c, launch_part_drain(c->S, c->T, c->G, c->S.int_values != 0 && !c->part_f64_tuples, ps, &c->launches));
  c->part_pending = 0; c->part_f64_tuples = false;
  if (st != ps) {
    CU(c, cudaEventRecord(c->part_ev, ps));
    CU(c, cudaStreamWaitEvent(st, c->part_ev, 0));
  }
  return BH_OK;
}
// before a launch that may log up to n tuples on `st`
int part_reserve(bh_ctx* c, uint64_t n, cudaStream_t st, bool f64_values) {
  if (!c->T.part_lg) return BH_OK;
  if (c->part_pending && (c->part_pending + n > c->part_limit || (c->part_stream && c->part_stream != st))) {
    int rc = part_drain(c, st);
    if (rc) return rc;
  }
  c->part_pending += n; c->part_stream = st;
  c->part_f64_tuples = c->part_f64_tuples || f64_values;
  return BH_OK;
}

// launch the fused kernel over what slab `s` holds on the device; `virt` is the payload base the
// offsets are relative to
int slab_launch(bh_ctx* c, Slab& s, const uint8_t* virt, uint32_t n, bool with_ts, cudaStream_t copy_stream, uint64_t bytes,
                uint32_t first_off = 0) {
  CU(c, cudaEventRecord(s.copied, copy_stream));
  CU(c, cudaStreamWaitEvent(c->st_compute, s.copied, 0));
  { int rc = part_reserve(c, n, c->st_compute, false); if (rc) return rc; }
  // the ingest path knows the slab's byte count: pick the kernel variant from the real average size
  CU(c, launch_process_raw(c->S, c->T, c->G, virt, s.d_offsets, with_ts ? s.d_ts : nullptr, n, c->st_compute, &c->launches,
                           n && bytes / n > 72 ? 1 : 0, first_off, (uint32_t)(first_off + bytes)));
  CU(c, launch_verify_keys(c->S, c->T, c->G, virt, s.d_offsets, n, c->st_compute, n && bytes / n > 72 ? 1 : 0, first_off,
                           (uint32_t)(first_off + bytes)));
  CU(c, cudaEventRecord(s.done, c->st_compute));
  s.in_flight = true;
  s.seq_end = c->pushed;
  return BH_OK;
}

// submit the host-staged slab `cur`
int submit_current(bh_ctx* c) {
  Slab& s = c->slabs[c->cur];
  if (s.n == 0) return BH_OK;
  cudaStream_t cs = c->st_copy[c->cur & 1];
  CU(c, cudaMemcpyAsync(s.d_payload, s.h_payload, (s.bytes + 15) & ~15ull, cudaMemcpyHostToDevice, cs));
  CU(c, cudaMemcpyAsync(s.d_offsets, s.h_offsets, ((size_t)s.n + 1) * 4, cudaMemcpyHostToDevice, cs));
  bool with_ts = c->S.has_window;
  if (with_ts) CU(c, cudaMemcpyAsync(s.d_ts, s.h_ts, (size_t)s.n * 8, cudaMemcpyHostToDevice, cs));
  int rc = slab_launch(c, s, s.d_payload, s.n, with_ts, cs, s.bytes);
  if (rc) return rc;
  s.n = 0; s.bytes = 0; s.h_offsets[0] = 0;
  c->cur = (c->cur + 1) % (uint32_t)c->slabs.size();
  return slab_acquire(c, c->cur);
}

// Large host-to-ring copies run on a few threads: one core moves 10-15 GB/s, a PCIe Gen5 x16 link takes 55, so a
// single-threaded memcpy in front of the H2D copy would be the bottleneck of bh_push_batch from ordinary
// (pageable) memory.  Threads are started per copy (tens of microseconds against milliseconds of copying).
void ring_copy(uint8_t* dst, const uint8_t* src, size_t n) {
  constexpr size_t kParallelFrom = 8u << 20;
  static const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  static const unsigned cap = [] { const char* e = getenv("BEHOLDER_COPY_THREADS"); int v = e ? atoi(e) : 0; return v > 0 ? (unsigned)v : 8u; }();
  const unsigned nt = n < kParallelFrom ? 1u : std::min<unsigned>({cap, std::max(1u, hw / 2), (unsigned)(n / (2u << 20))});
  if (nt <= 1) { memcpy(dst, src, n); return; }
  std::vector<std::thread> th;
  const size_t part = ((n + nt - 1) / nt + 4095) & ~(size_t)4095;
  for (unsigned i = 1; i < nt; i++) {
    const size_t lo = std::min(n, (size_t)i * part), hi = std::min(n, lo + part);
    if (hi > lo) th.emplace_back([=] { memcpy(dst + lo, src + lo, hi - lo); });
  }
  memcpy(dst, src, std::min(n, part));
  for (auto& t : th) t.join();
}

// the partial slab just received its first message: start its deadline, wake the completion thread
void slab_first_message(bh_ctx* c) {
  c->slab_t0 = std::chrono::steady_clock::now();
  if (c->cb) c->cv.notify_one();
}

// index.js:151,154 acks every message, and the broker stops delivering at `prefetch` unacked ones
// (index.js:43), so a partly filled slab must not wait for more traffic: once its oldest message is older
// than max_batch_latency it is submitted as it is.  Never blocks: if the next ring slab is still busy the
// submit is left to the next call.
int submit_if_due(bh_ctx* c) {
  Slab& s = c->slabs[c->cur];
  if (s.n == 0 || c->no_deadline) return BH_OK;
  if (std::chrono::steady_clock::now() - c->slab_t0 < c->max_latency) return BH_OK;
  Slab& nx = c->slabs[(c->cur + 1) % (uint32_t)c->slabs.size()];
  if (nx.in_flight) {
    cudaError_t e = cudaEventQuery(nx.done);
    if (e == cudaErrorNotReady) { cudaGetLastError(); return BH_OK; }
  }
  return submit_current(c);
}

int poll_slabs(bh_ctx* c) {
  for (auto& s : c->slabs) {
    if (!s.in_flight) continue;
    cudaError_t e = cudaEventQuery(s.done);
    if (e == cudaSuccess) { s.in_flight = false; if (s.seq_end > c->committed) c->committed = s.seq_end; }
    else if (e != cudaErrorNotReady) return fail(c, BH_ERR_CUDA, "cudaEventQuery: %s", cudaGetErrorString(e));
    else cudaGetLastError();
  }
  return BH_OK;
}

// The completion thread: enforces the batch deadline and reports the committed sequence number
// through the callback, so that acks do not depend on the producer calling in again (a queue that goes quiet
// after a few messages still gets them acked).  The callback runs on this thread with no lock held; the
// N-API shim forwards it through napi_threadsafe_function, the Python mirror through call_soon_threadsafe.
void completer_main(bh_ctx* c) {
  cudaSetDevice(c->device);
  std::unique_lock<std::recursive_mutex> lk(c->mu);
  while (!c->stop) {
    submit_if_due(c);
    poll_slabs(c);
    if (c->committed > c->reported && c->cb) {
      const uint64_t seq = c->reported = c->committed;
      bh_commit_cb cb = c->cb; void* user = c->cb_user;
      lk.unlock();
      cb(user, seq);
      lk.lock();
      continue;
    }
    bool busy = false;
    for (auto& s : c->slabs) busy = busy || s.in_flight;
    if (busy) c->cv.wait_for(lk, std::chrono::microseconds(50));                  // a kernel is running: poll its event
    else if (c->slabs[c->cur].n && !c->no_deadline) c->cv.wait_until(lk, c->slab_t0 + c->max_latency);
    else c->cv.wait_for(lk, std::chrono::milliseconds(100));                      // idle: woken by the first message
  }
}

void stop_completer(bh_ctx* c) {
  std::thread t;
  {
    BH_LOCK(c);
    c->stop = true;
    c->cv.notify_all();
    t.swap(c->completer);
  }
  if (t.joinable()) t.join();
  BH_LOCK(c);
  c->stop = false;
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// Pinned host memory is placed on the NUMA node of the thread that allocates it.  While this guard lives,
// the calling thread may only run on the CPUs of the GPU's own node, so that the ingest ring (and buffers
// from bh_host_alloc) sit next to the PCIe root the copies go through: with 8 GPUs on two sockets the
// other placement sends half of the H2D traffic over the socket interconnect.  Best effort: no sysfs
// entry, a cpuset that excludes those CPUs, or BEHOLDER_NO_NUMA_BIND=1 leave the thread where it is.
struct NumaLocalGuard {
#ifdef __linux__
  cpu_set_t old; bool bound = false;
  explicit NumaLocalGuard(int device, bool enabled = true) {
    const char* off = getenv("BEHOLDER_NO_NUMA_BIND");
    if (!enabled || (off && off[0] == '1')) return;
    char bus[32] = {0};
    if (cudaDeviceGetPCIBusId(bus, sizeof bus, device) != cudaSuccess) { cudaGetLastError(); return; }
    for (char* q = bus; *q; q++) *q = (char)tolower(*q);
    char path[128];
    snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/numa_node", bus);
    FILE* f = fopen(path, "r");
    if (!f) return;
    int node = -1;
    if (fscanf(f, "%d", &node) != 1) node = -1;
    fclose(f);
    if (node < 0) return;
    snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
    f = fopen(path, "r");
    if (!f) return;
    char list[4096] = {0};
    if (!fgets(list, sizeof list, f)) list[0] = 0;
    fclose(f);
    if (sched_getaffinity(0, sizeof old, &old) != 0) return;
    cpu_set_t want; CPU_ZERO(&want);
    int n_want = 0;
    for (char* q = list; *q;) {                      // "0-31,64-95"
      char* e; long a = strtol(q, &e, 10);
      if (e == q) break;
      long b = a;
      if (*e == '-') { q = e + 1; b = strtol(q, &e, 10); }
      for (long c = a; c <= b && c < CPU_SETSIZE; c++) if (CPU_ISSET((int)c, &old)) { CPU_SET((int)c, &want); n_want++; }
      q = (*e == ',') ? e + 1 : e;
      if (*e != ',') break;
    }
    if (n_want && sched_setaffinity(0, sizeof want, &want) == 0) bound = true;
  }
  ~NumaLocalGuard() { if (bound) sched_setaffinity(0, sizeof old, &old); }
#else
  explicit NumaLocalGuard(int, bool = true) {}
#endif
};

int counters_status(const bh_counters& k) {
  if (k.dropped_table_full) return BH_ERR_TABLE_FULL;
  if (k.dropped_window) return BH_ERR_WINDOW_RING;
  if (k.key_collisions) return BH_ERR_KEY_COLLISION;
  return BH_OK;
}
static_assert(sizeof(bh_counters) == kCounters * 8, "bh_counters and the device counter array");

}  // namespace

extern "C" {

int bh_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  int ok = 0;
  for (int i = 0; i < n; i++) {
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, i) == cudaSuccess && major == 10) ok++;
  }
  return ok;
}

const char* bh_last_error(bh_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int bh_create(const bh_config* cfg, bh_ctx** out) {
  if (!cfg || !out) return fail(nullptr, BH_ERR_INVALID, "null argument");
  *out = nullptr;
  if (cfg->struct_size != sizeof(bh_config)) return fail(nullptr, BH_ERR_INVALID, "bh_config.struct_size %u != %zu", cfg->struct_size, sizeof(bh_config));
  DevSchema S;
  int rc = build_schema(*cfg, S);
  if (rc) return rc;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(nullptr, BH_ERR_NO_DEVICE, "no CUDA device (%s); this library has no CPU path", e != cudaSuccess ? cudaGetErrorString(e) : "count 0");
  }
  if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, BH_ERR_INVALID, "device %d out of range (0..%d)", cfg->device, ndev - 1);
  cudaDeviceProp prop;
  if ((e = cudaGetDeviceProperties(&prop, cfg->device)) != cudaSuccess) return fail(nullptr, BH_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
  if (prop.major != 10) return fail(nullptr, BH_ERR_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device, prop.major, prop.minor);

  uint64_t cap = cfg->table_capacity ? cfg->table_capacity : (cfg->mode == BH_MODE_REFERENCE ? 64 : (1ull << 16));
  if (!is_pow2(cap) || cap < 64 || cap > (1ull << 30)) return fail(nullptr, BH_ERR_INVALID, "table_capacity must be a power of two in [64, 2^30]");
  uint32_t ring = cfg->window_ring ? cfg->window_ring : (S.has_window ? 32 : 1);
  if (!is_pow2(ring) || ring > 1024) return fail(nullptr, BH_ERR_INVALID, "window_ring must be a power of two <= 1024");

  bh_ctx* c = new bh_ctx();
  c->cfg = *cfg; c->device = cfg->device; c->S = S;
  DeviceGuard guard(c->device);
  auto bail = [&](int code) { std::string m = c->err; bh_destroy(c); g_create_error = m; return code; };
#define CUB_(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { fail(c, BH_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); return bail(e__ == cudaErrorMemoryAllocation ? BH_ERR_NOMEM : BH_ERR_CUDA); } } while (0)
  c->G.sms = prop.multiProcessorCount;
  c->G.smem_optin = (int)prop.sharedMemPerBlockOptin;
  { const char* g = getenv("BEHOLDER_GENERIC_KERNELS"); c->G.generic_only = (g && g[0] == '1') || (cfg->flags & BH_F_GENERIC_KERNELS); }
  CUB_(cudaStreamCreateWithFlags(&c->st_compute, cudaStreamNonBlocking));
  CUB_(cudaStreamCreateWithFlags(&c->st_copy[0], cudaStreamNonBlocking));
  CUB_(cudaStreamCreateWithFlags(&c->st_copy[1], cudaStreamNonBlocking));
  DevTables& T = c->T;
  memset(&T, 0, sizeof T);
  T.cap = cap; T.stride = cap + 2; T.ring = ring; T.hist_buckets = S.hist_buckets;
  size_t rows = (size_t)ring * T.stride;
  CUB_(cudaMalloc((void**)&T.ring_win, (size_t)ring * 8));
  CUB_(cudaMalloc((void**)&T.keys, rows * 8));
  CUB_(cudaMalloc((void**)&T.stats, rows * sizeof(Stat)));
  if (S.ops & BH_OP_HIST) CUB_(cudaMalloc((void**)&T.hist, rows * S.hist_buckets * 4));
  CUB_(cudaMalloc((void**)&T.counters, kCounterSlots * 8));
  if (cfg->flags & BH_F_VERIFY_KEYS) {
    const bool ok_shape = cfg->mode == BH_MODE_WINDOWED && S.fast_key0 != 0xFFFFFFFFu &&
                          (S.n_parts == 1 || (S.n_parts == 2 && S.fast_status_part == 2));
    if (!ok_shape) { fail(c, BH_ERR_UNSUPPORTED, "BH_F_VERIFY_KEYS needs a windowed context keyed by one string field, optionally followed by the status"); return bail(BH_ERR_UNSUPPORTED); }
    T.dict_cap = 2 * cap;
    CUB_(cudaMalloc((void**)&T.dict, T.dict_cap * sizeof(DictEntry)));
  }
  // partitioned (deferred) aggregation: on request, or by itself once a window's table is too large for L2
  const bool part = cfg->mode == BH_MODE_WINDOWED && !(cfg->flags & BH_F_NO_PARTITION) &&
                    ((cfg->flags & BH_F_PARTITIONED) || cap >= (1ull << 18));
  uint32_t lg = 0;                                               // log2(partitions); 0 = not partitioned
  if (part) {
    while ((1ull << lg) < cap) lg++;
    // ~2048 table slots per partition, 64 .. 2048 partitions (at 2^24 slots 4 096 partitions are 5 % slower than
    // 1 024 or 2 048: the scatter of the log phase pays for every extra open region).  Measured at the C3 shape (2^21 slots): 512 / 1024 /
    // 2048 / 4096 partitions -> 164 / 152 / 159 / 178 DRAM bytes per event (log + fold) and 15.5 / 12.9 / 13.2 /
    // 16.0 ms per 200 M events: more partitions widen the front of half-written log sectors that L2 has to keep,
    // fewer crowd the fold's shared table.
    lg = lg > 22 ? 11 : (lg < 17 ? 6 : lg - 11);
    if (const char* e = getenv("BEHOLDER_PART_LG")) { int v = atoi(e); if (v >= 4 && v <= 14) lg = (uint32_t)v; }   // experiments
    while ((1u << lg) < ring) lg++;                              // a tuple carries its ring slot in the partition bits
    // probing wraps inside a partition's slice (g_probe_wrap): keep slices large enough to fill like the table
    // does (256 slots: +- 8 keys around the mean) -- fewer partitions for small tables, none if that cannot be
    uint32_t lgcap = 0;
    while ((1ull << lgcap) < cap) lgcap++;
    if (lgcap < lg + 8) {
      const uint32_t lg2 = lgcap >= 8 ? lgcap - 8 : 0;
      if (lg2 >= 1 && (1u << lg2) >= ring) lg = lg2;
      else if (cfg->flags & BH_F_PARTITIONED) {
        fail(c, BH_ERR_INVALID, "BH_F_PARTITIONED needs table_capacity >= 256 x max(2, window_ring) slots");
        return bail(BH_ERR_INVALID);
      } else lg = 0;
    }
  }
  if (lg) {
    T.part_lg = lg;
    // regions x tuples stays constant: 2^20 tuples per SM = 16 MB, 2.5 GB of log on 148 SMs, ~93 M events per fold
    T.part_cap = cfg->part_region_tuples ? cfg->part_region_tuples : std::max(128u, (1u << 20) >> lg);
    T.part_ctas = (uint32_t)prop.multiProcessorCount;
    const size_t regions = (size_t)T.part_ctas << lg;
    CUB_(cudaMalloc((void**)&T.part_tuples, regions * T.part_cap * sizeof(uint4)));
    CUB_(cudaMalloc((void**)&T.part_fill, regions * 4));
    CUB_(cudaEventCreateWithFlags(&c->part_ev, cudaEventDisableTiming));
    c->part_limit = (uint64_t)((double)regions * T.part_cap * 0.6);   // mean fill 0.6: a uniform stream stays > 10 sigma below a region's end
  }
  CUB_(cudaMalloc((void**)&c->d_owner_counts, 1024 * 8));
  CUB_(cudaMalloc((void**)&c->d_owner_cursor, 1024 * 8));
  CUB_(cudaHostAlloc((void**)&c->h_small, 1024 * 8, cudaHostAllocDefault));
  CUB_(launch_init_tables(T, c->st_compute));
  CUB_(configure_kernels(c->G));
  c->slab_bytes = cfg->slab_bytes ? cfg->slab_bytes : (64ull << 20);
  c->slab_events = cfg->slab_events ? cfg->slab_events : (1u << 20);
  if (c->slab_bytes > 0xFFFFFF00ull) { fail(c, BH_ERR_INVALID, "slab_bytes must stay below 4 GiB (u32 offsets)"); return bail(BH_ERR_INVALID); }
  if (cfg->max_batch_latency_ms == BH_NO_BATCH_DEADLINE) c->no_deadline = true;
  else c->max_latency = std::chrono::microseconds(1000ull * (cfg->max_batch_latency_ms ? cfg->max_batch_latency_ms : 5u));
  uint32_t n_slabs = cfg->n_slabs ? cfg->n_slabs : 3;
  c->slabs.resize(n_slabs);
  {
    NumaLocalGuard numa(c->device, !(cfg->flags & BH_F_NO_NUMA_BIND));   // pinned ring on the GPU's NUMA node
    for (auto& s : c->slabs) {
      int r2 = slab_alloc(c, s);
      if (r2) return bail(r2);
    }
  }
  CUB_(cudaStreamSynchronize(c->st_compute));
#undef CUB_
  *out = c;
  return BH_OK;
}

void bh_destroy(bh_ctx* c) {
  if (!c) return;
  stop_completer(c);
  DeviceGuard guard(c->device);
  if (c->st_compute) cudaStreamSynchronize(c->st_compute);
  for (auto& s : c->slabs) slab_free(s);
  if (c->T.ring_win) cudaFree(c->T.ring_win);
  if (c->T.keys) cudaFree(c->T.keys);
  if (c->T.stats) cudaFree(c->T.stats);
  if (c->T.hist) cudaFree(c->T.hist);
  if (c->T.counters) cudaFree(c->T.counters);
  if (c->T.dict) cudaFree(c->T.dict);
  if (c->T.part_tuples) cudaFree(c->T.part_tuples);
  if (c->T.part_fill) cudaFree(c->T.part_fill);
  if (c->part_ev) cudaEventDestroy(c->part_ev);
  if (c->wait_ev) cudaEventDestroy(c->wait_ev);
  if (c->d_owner_counts) cudaFree(c->d_owner_counts);
  if (c->d_owner_cursor) cudaFree(c->d_owner_cursor);
  if (c->d_export) cudaFree(c->d_export);
  if (c->d_export_hist) cudaFree(c->d_export_hist);
  if (c->h_small) cudaFreeHost(c->h_small);
  for (auto& s : c->st_copy) if (s) cudaStreamDestroy(s);
  if (c->st_compute) cudaStreamDestroy(c->st_compute);
  delete c;
}

int bh_host_alloc(void** out, size_t bytes) {
  if (!out) return BH_ERR_INVALID;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); dev = 0; }
  NumaLocalGuard numa(dev);          // on the current device's NUMA node
  cudaError_t e = cudaHostAlloc(out, bytes, cudaHostAllocPortable);
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); cudaGetLastError(); return e == cudaErrorMemoryAllocation ? BH_ERR_NOMEM : BH_ERR_CUDA; }
  return BH_OK;
}
int bh_host_free(void* p) { return cudaFreeHost(p) == cudaSuccess ? BH_OK : BH_ERR_CUDA; }

int bh_host_register(void* p, size_t bytes) {
  if (!p || !bytes) return BH_ERR_INVALID;
  cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); cudaGetLastError(); return e == cudaErrorMemoryAllocation ? BH_ERR_NOMEM : BH_ERR_CUDA; }
  return BH_OK;
}
int bh_host_unregister(void* p) {
  if (cudaHostUnregister(p) != cudaSuccess) { cudaGetLastError(); return BH_ERR_CUDA; }
  return BH_OK;
}

int bh_push(bh_ctx* c, const uint8_t* payload, uint32_t len, uint64_t ts_ms) {
  if (!c || (!payload && len)) return fail(c, BH_ERR_INVALID, "null argument");
  if ((uint64_t)len > c->slab_bytes) return fail(c, BH_ERR_TOO_LARGE, "message of %u bytes exceeds slab_bytes %llu", len, (unsigned long long)c->slab_bytes);
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  Slab* s = &c->slabs[c->cur];
  if (s->bytes + len > c->slab_bytes || s->n == c->slab_events) {
    int rc = submit_current(c);
    if (rc) return rc;
    s = &c->slabs[c->cur];
  }
  if (s->n == 0) slab_first_message(c);
  memcpy(s->h_payload + s->bytes, payload, len);
  s->h_ts[s->n] = ts_ms;
  s->bytes += len;
  s->n++;
  s->h_offsets[s->n] = (uint32_t)s->bytes;
  c->pushed++;
  return BH_OK;
}

int bh_push_batch(bh_ctx* c, const uint8_t* slab, const uint32_t* off, const uint64_t* ts, uint32_t n) {
  if (!c || (n && (!slab || !off))) return fail(c, BH_ERR_INVALID, "null argument");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  uint32_t i = 0;
  while (i < n) {
    Slab* s = &c->slabs[c->cur];
    if (off[i + 1] < off[i]) return fail(c, BH_ERR_INVALID, "offsets not monotone at %u", i);
    uint64_t len0 = off[i + 1] - off[i];
    if (len0 > c->slab_bytes) return fail(c, BH_ERR_TOO_LARGE, "message %u of %llu bytes exceeds slab_bytes", i, (unsigned long long)len0);
    if (s->bytes + len0 > c->slab_bytes || s->n == c->slab_events) {
      int rc = submit_current(c);
      if (rc) return rc;
      continue;
    }
    // largest j with events [i, j) fitting the current slab
    uint64_t room_b = c->slab_bytes - s->bytes;
    uint32_t room_e = c->slab_events - s->n;
    uint32_t hi = std::min<uint64_t>(n, (uint64_t)i + room_e);
    uint32_t j = (uint32_t)(std::upper_bound(off + i, off + hi + 1, (uint64_t)off[i] + room_b,
                                            [](uint64_t v, uint32_t o) { return v < o; }) - off) - 1;
    if (j <= i) j = i + 1;
    for (uint32_t k = i; k < j; k++)
      if (off[k + 1] < off[k]) return fail(c, BH_ERR_INVALID, "offsets not monotone at %u", k);
    uint64_t nb = off[j] - off[i];
    if (s->n == 0) slab_first_message(c);
    ring_copy(s->h_payload + s->bytes, slab + off[i], nb);
    for (uint32_t k = i; k < j; k++) {
      s->h_offsets[s->n + (k - i) + 1] = (uint32_t)(s->bytes + (off[k + 1] - off[i]));
      s->h_ts[s->n + (k - i)] = ts ? ts[k] : 0;
    }
    s->bytes += nb;
    s->n += j - i;
    c->pushed += j - i;
    i = j;
  }
  return BH_OK;
}

int bh_push_batch_pinned(bh_ctx* c, const uint8_t* slab, const uint32_t* off, const uint64_t* ts, uint32_t n) {
  if (!c || (n && (!slab || !off))) return fail(c, BH_ERR_INVALID, "null argument");
  if (c->S.has_window && !ts && n) return fail(c, BH_ERR_INVALID, "ts_ms required in windowed mode");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  // keep arrival order: anything staged through bh_push goes first
  int rc = submit_current(c);
  if (rc) return rc;
  uint32_t i = 0;
  while (i < n) {
    rc = slab_acquire(c, c->cur);
    if (rc) return rc;
    Slab& s = c->slabs[c->cur];
    uint32_t hi = (uint32_t)std::min<uint64_t>(n, (uint64_t)i + c->slab_events);
    uint32_t j = (uint32_t)(std::upper_bound(off + i, off + hi + 1, (uint64_t)off[i] + c->slab_bytes,
                                            [](uint64_t v, uint32_t o) { return v < o; }) - off) - 1;
    if (j <= i) return fail(c, BH_ERR_TOO_LARGE, "message %u exceeds slab_bytes", i);
    // The offsets are the caller's and are not scanned here (50 MB per 12.5 M events on the producer thread cost a
    // tenth of the ingest rate): the chunk's two ends are checked, and the kernel treats every offset outside
    // [off[i], off[j]] as a message that does not decode and reads nothing beyond the bytes copied below.
    if (off[j] < off[i] || (uint64_t)off[j] - off[i] > c->slab_bytes) return fail(c, BH_ERR_INVALID, "offsets must not decrease (at %u)", i);
    uint64_t nb = off[j] - off[i];
    uint32_t head = off[i] & 15u;
    cudaStream_t cs = c->st_copy[c->cur & 1];
    CU(c, cudaMemcpyAsync(s.d_payload + head, slab + off[i], nb, cudaMemcpyHostToDevice, cs));
    CU(c, cudaMemcpyAsync(s.d_offsets, off + i, ((size_t)(j - i) + 1) * 4, cudaMemcpyHostToDevice, cs));
    bool with_ts = c->S.has_window;
    if (with_ts) CU(c, cudaMemcpyAsync(s.d_ts, ts + i, (size_t)(j - i) * 8, cudaMemcpyHostToDevice, cs));
    c->pushed += j - i;
    // offsets stay absolute: hand the kernel the (16-byte aligned) address offset 0 would have
    const uint8_t* virt = s.d_payload + head - off[i];
    rc = slab_launch(c, s, virt, j - i, with_ts, cs, nb, off[i]);
    if (rc) return rc;
    c->cur = (c->cur + 1) % (uint32_t)c->slabs.size();
    i = j;
  }
  return slab_acquire(c, c->cur);
}

int bh_push_batch_compact(bh_ctx* c, const uint8_t* slab, const uint16_t* lengths, const uint64_t* boff, const uint64_t* ts_block,
                          const int16_t* ts_delta, uint32_t n) {
  if (!c || (n && (!slab || !lengths || !boff))) return fail(c, BH_ERR_INVALID, "null argument");
  const bool with_ts = c->S.has_window;
  if (with_ts && n && (!ts_block || !ts_delta)) return fail(c, BH_ERR_INVALID, "ts_block / ts_delta required in windowed mode");
  constexpr uint32_t B = 1u << BH_COMPACT_BLOCK_LG;
  if (c->slab_events < B) return fail(c, BH_ERR_INVALID, "bh_push_batch_compact needs slab_events >= %u", B);
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);          // keep arrival order: anything staged through bh_push goes first
  if (rc) return rc;
  const uint32_t n_blocks = (n + B - 1) / B;
  uint32_t k = 0;                      // first block of the next chunk
  while (k < n_blocks) {
    rc = slab_acquire(c, c->cur);
    if (rc) return rc;
    Slab& s = c->slabs[c->cur];
    // as many whole blocks as the ring slab takes (events and bytes)
    uint32_t k1 = std::min<uint32_t>(n_blocks, k + std::min<uint32_t>(c->slab_events / B, 64u));
    while (k1 > k && boff[k1] - boff[k] > c->slab_bytes) k1--;
    if (k1 == k) return fail(c, BH_ERR_TOO_LARGE, "block %u of %llu bytes exceeds slab_bytes", k, (unsigned long long)(boff[k + 1] - boff[k]));
    if (boff[k1] < boff[k]) return fail(c, BH_ERR_INVALID, "block_offsets must not decrease (at %u)", k);
    const uint32_t e0 = k * B, m = std::min<uint64_t>(n, (uint64_t)k1 * B) - e0, nb_blocks = k1 - k;
    const uint64_t nb = boff[k1] - boff[k];
    if (!s.d_len) {
      const size_t E = c->slab_events;
      CU(c, cudaMalloc((void**)&s.d_len, E * 2 + 16));
      CU(c, cudaMalloc((void**)&s.d_tsd, E * 2 + 16));
    }
    cudaStream_t cs = c->st_copy[c->cur & 1];
    CU(c, cudaMemcpyAsync(s.d_payload, slab + boff[k], nb, cudaMemcpyHostToDevice, cs));
    CU(c, cudaMemcpyAsync(s.d_len, lengths + e0, (size_t)m * 2, cudaMemcpyHostToDevice, cs));
    if (with_ts) CU(c, cudaMemcpyAsync(s.d_tsd, ts_delta + e0, (size_t)m * 2, cudaMemcpyHostToDevice, cs));
    // the few per-block offsets and base timestamps of this chunk go along as kernel arguments
    CU(c, launch_expand_compact(s.d_len, (const unsigned long long*)(boff + k), (const unsigned long long*)(with_ts ? ts_block + k : nullptr),
                                s.d_tsd, m, boff[k], s.d_offsets, with_ts ? (unsigned long long*)s.d_ts : nullptr, cs));
    (void)nb_blocks;
    c->pushed += m;
    // offsets are rebuilt from the lengths on the device; the kernel bounds every read by the nb bytes copied
    rc = slab_launch(c, s, s.d_payload, m, with_ts, cs, nb, 0);
    if (rc) return rc;
    c->cur = (c->cur + 1) % (uint32_t)c->slabs.size();
    k = k1;
  }
  return slab_acquire(c, c->cur);
}

int bh_push_columns(bh_ctx* c, const uint64_t* kh, const uint64_t* ts, const double* val, uint64_t n) {
  if (!c || (n && !kh)) return fail(c, BH_ERR_INVALID, "null argument");
  if (c->S.mode != BH_MODE_WINDOWED) return fail(c, BH_ERR_INVALID, "bh_push_columns needs BH_MODE_WINDOWED");
  if (c->T.dict) return fail(c, BH_ERR_UNSUPPORTED, "BH_F_VERIFY_KEYS: pre-hashed columns carry no key bytes to verify");
  if (c->S.has_window && !ts && n) return fail(c, BH_ERR_INVALID, "ts_ms required in windowed mode");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);
  if (rc) return rc;
  uint64_t per = std::min<uint64_t>(c->slab_bytes / 24, 0x7FFFFFFFull);
  if (per == 0) return fail(c, BH_ERR_INVALID, "slab_bytes too small");
  uint64_t i = 0;
  Slab* last = nullptr;
  while (i < n) {
    rc = slab_acquire(c, c->cur);
    if (rc) return rc;
    Slab& s = c->slabs[c->cur];
    uint64_t m = std::min(per, n - i);
    uint64_t* d_kh = (uint64_t*)s.d_payload;
    uint64_t* d_ts = d_kh + m;
    double* d_val = (double*)(d_ts + m);
    cudaStream_t cs = c->st_copy[c->cur & 1];
    CU(c, cudaMemcpyAsync(d_kh, kh + i, m * 8, cudaMemcpyHostToDevice, cs));
    if (c->S.has_window) CU(c, cudaMemcpyAsync(d_ts, ts + i, m * 8, cudaMemcpyHostToDevice, cs));
    if (val) CU(c, cudaMemcpyAsync(d_val, val + i, m * 8, cudaMemcpyHostToDevice, cs));
    CU(c, cudaEventRecord(s.copied, cs));
    CU(c, cudaStreamWaitEvent(c->st_compute, s.copied, 0));
    c->pushed += m;
    rc = part_reserve(c, m, c->st_compute, true);
    if (rc) return rc;
    CU(c, launch_process_columns(c->S, c->T, c->G, d_kh, d_ts, val ? d_val : nullptr, m, c->st_compute, &c->launches));
    CU(c, cudaEventRecord(s.done, c->st_compute));
    s.in_flight = true; s.seq_end = c->pushed;
    last = &s;
    c->cur = (c->cur + 1) % (uint32_t)c->slabs.size();
    i += m;
  }
  // the caller may reuse its buffers as soon as we return
  if (last) CU(c, cudaEventSynchronize(last->copied));
  return slab_acquire(c, c->cur);
}

int bh_poll_committed(bh_ctx* c, uint64_t* seq) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_if_due(c);
  if (rc) return rc;
  rc = poll_slabs(c);
  if (rc) return rc;
  if (seq) *seq = c->committed;
  return BH_OK;
}

int bh_submit(bh_ctx* c) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);
  if (c->cb) c->cv.notify_one();
  return rc;
}

int bh_set_commit_callback(bh_ctx* c, bh_commit_cb cb, void* user) {
  if (!c) return BH_ERR_INVALID;
  stop_completer(c);
  BH_LOCK(c);
  c->cb = cb; c->cb_user = user;
  c->reported = c->committed;
  if (cb) c->completer = std::thread(completer_main, c);
  return BH_OK;
}

static int ensure_export(bh_ctx* c, uint64_t cap, bool hist, bool all_windows);

// host-pointer merge: the restore side of bh_snapshot_hist
int bh_restore(bh_ctx* c, const bh_record* recs, const uint32_t* hist, uint64_t n) {
  if (!c || (n && !recs)) return fail(c, BH_ERR_INVALID, "null argument");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);
  if (rc) return rc;
  const uint64_t chunk = 1u << 16;
  bool with_hist = hist != nullptr && c->T.hist != nullptr;
  rc = ensure_export(c, std::min<uint64_t>(std::max<uint64_t>(n, 1), chunk), with_hist, false);
  if (rc) return rc;
  for (uint64_t i = 0; i < n; i += c->export_cap) {
    uint64_t m = std::min<uint64_t>(c->export_cap, n - i);
    CU(c, cudaMemcpyAsync(c->d_export, recs + i, m * sizeof(bh_record), cudaMemcpyHostToDevice, c->st_compute));
    if (with_hist) CU(c, cudaMemcpyAsync(c->d_export_hist, hist + i * c->T.hist_buckets, m * c->T.hist_buckets * 4, cudaMemcpyHostToDevice, c->st_compute));
    CU(c, launch_merge(c->S, c->T, c->d_export, with_hist ? c->d_export_hist : nullptr, m, c->st_compute));
    CU(c, cudaStreamSynchronize(c->st_compute));   // the caller's buffers are pageable: finish before the next chunk
  }
  bh_counters k;
  return bh_get_counters(c, &k);
}

int bh_sync(bh_ctx* c) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  { int rc = part_drain(c, c->st_compute); if (rc) return rc; }
  CU(c, cudaStreamSynchronize(c->st_compute));
  for (auto& s : c->slabs)
    if (s.in_flight) { s.in_flight = false; if (s.seq_end > c->committed) c->committed = s.seq_end; }
  return BH_OK;
}

int bh_get_counters(bh_ctx* c, bh_counters* out) {
  if (!c || !out) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  { int rc = part_drain(c, c->st_compute); if (rc) return rc; }
  CU(c, cudaMemcpyAsync(c->h_small, c->T.counters, kCounters * 8, cudaMemcpyDeviceToHost, c->st_compute));
  CU(c, cudaStreamSynchronize(c->st_compute));
  memcpy(out, c->h_small, sizeof(bh_counters));
  return counters_status(*out);
}

int bh_flush(bh_ctx* c, uint64_t* committed_seq) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);
  if (rc) return rc;
  rc = bh_sync(c);
  if (rc) return rc;
  c->committed = c->pushed;
  if (committed_seq) *committed_seq = c->committed;
  bh_counters k;
  rc = bh_get_counters(c, &k);
  if (rc == BH_ERR_TABLE_FULL) return fail(c, rc, "%llu events dropped: a window's key table is full (table_capacity %llu)", (unsigned long long)k.dropped_table_full, (unsigned long long)c->T.cap);
  if (rc == BH_ERR_WINDOW_RING) return fail(c, rc, "%llu events dropped: more than %u windows open (close windows or raise window_ring)", (unsigned long long)k.dropped_window, c->T.ring);
  if (rc == BH_ERR_KEY_COLLISION) return fail(c, rc, "%llu events carry a key whose 64-bit hash belongs to a different key: their rows are merged with that key's (change hash_seed)", (unsigned long long)k.key_collisions);
  return rc;
}

static int ensure_export(bh_ctx* c, uint64_t cap, bool hist, bool all_windows) {
  const uint64_t most = all_windows ? c->T.stride * c->T.ring : c->T.stride;    // rows a snapshot can return
  if (cap > most) cap = most;
  if (cap == 0) cap = 1;
  if (cap > c->export_cap) {
    if (c->d_export) cudaFree(c->d_export);
    if (c->d_export_hist) cudaFree(c->d_export_hist);
    c->d_export = nullptr; c->d_export_hist = nullptr; c->export_cap = 0;
    CU(c, cudaMalloc((void**)&c->d_export, cap * sizeof(bh_record)));
    c->export_cap = cap;      // (d_export_hist was freed above and is re-made at the new size below)
  }
  if (hist && c->T.hist && !c->d_export_hist)
    CU(c, cudaMalloc((void**)&c->d_export_hist, c->export_cap * c->T.hist_buckets * 4));
  return BH_OK;
}

int bh_snapshot_hist(bh_ctx* c, uint64_t window_id, bh_record* out, uint32_t* hist_out, uint64_t cap, uint64_t* n) {
  if (!c || !n || (cap && !out)) return fail(c, BH_ERR_INVALID, "null argument");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);
  if (rc) return rc;
  bool hist = hist_out != nullptr && c->T.hist != nullptr;
  rc = ensure_export(c, cap, hist, window_id == BH_ALL_WINDOWS);
  if (rc) return rc;
  uint64_t dcap = std::min<uint64_t>(cap, c->export_cap);
  rc = part_drain(c, c->st_compute);
  if (rc) return rc;
  CU(c, launch_export(c->S, c->T, window_id, 1, c->d_export, hist ? c->d_export_hist : nullptr, dcap, 0,
                      c->d_owner_counts, c->d_owner_cursor, c->st_compute));
  CU(c, cudaMemcpyAsync(c->h_small, c->d_owner_counts, 8, cudaMemcpyDeviceToHost, c->st_compute));
  CU(c, cudaStreamSynchronize(c->st_compute));
  uint64_t rows = c->h_small[0];
  *n = rows;
  uint64_t take = std::min(rows, dcap);
  if (take) {
    CU(c, cudaMemcpy(out, c->d_export, take * sizeof(bh_record), cudaMemcpyDeviceToHost));
    if (hist) CU(c, cudaMemcpy(hist_out, c->d_export_hist, take * c->T.hist_buckets * 4, cudaMemcpyDeviceToHost));
  }
  return BH_OK;
}

int bh_snapshot(bh_ctx* c, uint64_t window_id, bh_record* out, uint64_t cap, uint64_t* n) {
  return bh_snapshot_hist(c, window_id, out, nullptr, cap, n);
}

int bh_list_windows(bh_ctx* c, uint64_t* out, uint32_t cap, uint32_t* n) {
  if (!c || !n) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);
  if (rc) return rc;
  CU(c, cudaMemcpyAsync(c->h_small, c->T.ring_win, (size_t)c->T.ring * 8, cudaMemcpyDeviceToHost, c->st_compute));
  CU(c, cudaStreamSynchronize(c->st_compute));
  std::vector<uint64_t> w;
  for (uint32_t i = 0; i < c->T.ring; i++) if (c->h_small[i] != kNoWindow) w.push_back(c->h_small[i]);
  std::sort(w.begin(), w.end());
  *n = (uint32_t)w.size();
  for (uint32_t i = 0; i < w.size() && i < cap && out; i++) out[i] = w[i];
  return BH_OK;
}

int bh_close_window(bh_ctx* c, uint64_t window_id) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  int rc = submit_current(c);
  if (rc) return rc;
  rc = part_drain(c, c->st_compute);
  if (rc) return rc;
  CU(c, launch_close_window(c->T, window_id, c->st_compute));
  return BH_OK;
}

// ---- device-resident entry points ----
int bh_process_device(bh_ctx* c, const uint8_t* d_payload, const uint32_t* d_offsets, const uint64_t* d_ts,
                      uint32_t n, void* stream) {
  if (!c || (n && (!d_payload || !d_offsets))) return fail(c, BH_ERR_INVALID, "null argument");
  if (((uintptr_t)d_payload & 15) != 0) return fail(c, BH_ERR_INVALID, "d_payload must be 16-byte aligned");
  if (c->S.has_window && !d_ts && n) return fail(c, BH_ERR_INVALID, "d_ts_ms required in windowed mode");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  { int rc = part_reserve(c, n, st, false); if (rc) return rc; }
  CU(c, launch_process_raw(c->S, c->T, c->G, d_payload, d_offsets, d_ts, n, st, &c->launches));
  CU(c, launch_verify_keys(c->S, c->T, c->G, d_payload, d_offsets, n, st));
  return BH_OK;
}

int bh_process_columns_device(bh_ctx* c, const uint64_t* d_kh, const uint64_t* d_ts, const double* d_val,
                              uint64_t n, void* stream) {
  if (!c || (n && !d_kh)) return fail(c, BH_ERR_INVALID, "null argument");
  if (c->S.mode != BH_MODE_WINDOWED) return fail(c, BH_ERR_INVALID, "needs BH_MODE_WINDOWED");
  if (c->T.dict) return fail(c, BH_ERR_UNSUPPORTED, "BH_F_VERIFY_KEYS: pre-hashed columns carry no key bytes to verify");
  if (c->S.has_window && !d_ts && n) return fail(c, BH_ERR_INVALID, "d_ts_ms required in windowed mode");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  { int rc = part_reserve(c, n, st, true); if (rc) return rc; }
  CU(c, launch_process_columns(c->S, c->T, c->G, d_kh, d_ts, d_val, n, st, &c->launches));
  return BH_OK;
}

int bh_decode_device(bh_ctx* c, const uint8_t* d_payload, const uint32_t* d_offsets, uint32_t n,
                     uint64_t* d_kh, double* d_val, int32_t* d_status, uint8_t* d_outcome, void* stream) {
  if (!c || (n && (!d_payload || !d_offsets || !d_kh || !d_val || !d_status || !d_outcome))) return fail(c, BH_ERR_INVALID, "null argument");
  if (((uintptr_t)d_payload & 15) != 0) return fail(c, BH_ERR_INVALID, "d_payload must be 16-byte aligned");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  CU(c, launch_decode(c->S, c->G, d_payload, d_offsets, n, d_kh, d_val, d_status, d_outcome, st));
  return BH_OK;
}

int bh_export_window_device(bh_ctx* c, uint64_t window_id, uint32_t n_owners, bh_record* d_records,
                            uint32_t* d_hist, uint64_t cap, uint64_t owner_stride, uint64_t* d_owner_counts,
                            void* stream) {
  if (!c || !d_records || !d_owner_counts) return fail(c, BH_ERR_INVALID, "null argument");
  if (!is_pow2(n_owners) || n_owners > 1024) return fail(c, BH_ERR_INVALID, "n_owners must be a power of two <= 1024");
  if (owner_stride && owner_stride * n_owners > cap) return fail(c, BH_ERR_INVALID, "owner_stride * n_owners exceeds cap");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  { int rc = part_drain(c, st); if (rc) return rc; }
  CU(c, launch_export(c->S, c->T, window_id, n_owners, d_records, d_hist, cap, owner_stride,
                      (unsigned long long*)d_owner_counts, c->d_owner_cursor, st));
  return BH_OK;
}

int bh_export_window_peers(bh_ctx* c, uint64_t window_id, uint32_t n_owners, const uint64_t* d_peer_bases,
                           const uint64_t* d_peer_hist_bases, uint32_t my_rank, uint64_t owner_stride,
                           uint64_t* d_owner_counts, void* stream) {
  if (!c || !d_peer_bases || !d_owner_counts) return fail(c, BH_ERR_INVALID, "null argument");
  if (!is_pow2(n_owners) || n_owners > 1024) return fail(c, BH_ERR_INVALID, "n_owners must be a power of two <= 1024");
  if (my_rank >= n_owners || owner_stride == 0) return fail(c, BH_ERR_INVALID, "my_rank must be below n_owners and owner_stride positive");
  if ((c->S.ops & BH_OP_HIST) && !d_peer_hist_bases) return fail(c, BH_ERR_INVALID, "this context keeps histograms: d_peer_hist_bases is required");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  { int rc = part_drain(c, st); if (rc) return rc; }
  CU(c, launch_export(c->S, c->T, window_id, n_owners, nullptr, nullptr, (uint64_t)n_owners * owner_stride, owner_stride,
                      (unsigned long long*)d_owner_counts, c->d_owner_cursor, st,
                      (const unsigned long long*)d_peer_bases, (const unsigned long long*)d_peer_hist_bases, my_rank));
  return BH_OK;
}

int bh_export_window_static(bh_ctx* c, uint64_t window_id, uint32_t n_owners, bh_record* d_records,
                            uint32_t* d_hist, uint64_t cap, uint64_t owner_stride, uint64_t* d_owner_counts,
                            void* stream) {
  if (!c || !d_records || !d_owner_counts) return fail(c, BH_ERR_INVALID, "null argument");
  if (n_owners > BH_STATIC_EXPORT_MAX_OWNERS)
    return fail(c, BH_ERR_TOO_LARGE, "bh_export_window_static: n_owners %u exceeds the built-in scratch of %d owners; "
                "use bh_export_window_device", n_owners, BH_STATIC_EXPORT_MAX_OWNERS);
  if (!is_pow2(n_owners)) return fail(c, BH_ERR_INVALID, "n_owners must be a power of two");
  if (owner_stride && owner_stride * n_owners > cap) return fail(c, BH_ERR_INVALID, "owner_stride * n_owners exceeds cap");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  { int rc = part_drain(c, st); if (rc) return rc; }
  CU(c, launch_export_static(c->S, c->T, window_id, n_owners, d_records, d_hist, cap, owner_stride,
                             (unsigned long long*)d_owner_counts, st));
  return BH_OK;
}

int bh_close_window_device(bh_ctx* c, uint64_t window_id, void* stream) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  { int rc = part_drain(c, st); if (rc) return rc; }
  CU(c, launch_close_window(c->T, window_id, st));
  return BH_OK;
}

int bh_drain_device(bh_ctx* c, void* stream) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  return part_drain(c, stream ? (cudaStream_t)stream : c->st_compute);
}

int bh_reset_device(bh_ctx* c, void* stream) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  if (c->part_pending && c->part_stream && c->part_stream != st) {   // order the reset after the kernels that still write the log
    CU(c, cudaEventRecord(c->part_ev, c->part_stream));
    CU(c, cudaStreamWaitEvent(st, c->part_ev, 0));
  }
  c->part_pending = 0; c->part_f64_tuples = false;                  // the reset empties the log as well
  CU(c, launch_reset_tables(c->T, st));
  return BH_OK;
}

int bh_merge_records_device(bh_ctx* c, const bh_record* d_records, const uint32_t* d_hist, uint64_t n, void* stream) {
  if (!c || (n && !d_records)) return fail(c, BH_ERR_INVALID, "null argument");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  CU(c, launch_merge(c->S, c->T, d_records, d_hist, n, st));
  return BH_OK;
}

int bh_merge_blocks_device(bh_ctx* c, const bh_record* d_records, const uint32_t* d_hist, uint32_t n_blocks,
                           uint64_t block_stride, const uint64_t* d_block_counts, void* stream) {
  if (!c || (n_blocks && (!d_records || !d_block_counts))) return fail(c, BH_ERR_INVALID, "null argument");
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->st_compute;
  CU(c, launch_merge_blocks(c->S, c->T, d_records, d_hist, n_blocks, block_stride, (const unsigned long long*)d_block_counts, st));
  return BH_OK;
}

void* bh_stream(bh_ctx* c) { return c ? (void*)c->st_compute : nullptr; }

int bh_stream_wait(bh_ctx* c, void* stream) {
  if (!c) return BH_ERR_INVALID;
  BH_LOCK(c);
  DeviceGuard guard(c->device);
  if ((cudaStream_t)stream == c->st_compute) return BH_OK;
  if (!c->wait_ev) CU(c, cudaEventCreateWithFlags(&c->wait_ev, cudaEventDisableTiming));
  CU(c, cudaEventRecord(c->wait_ev, (cudaStream_t)stream));
  CU(c, cudaStreamWaitEvent(c->st_compute, c->wait_ev, 0));
  return BH_OK;
}

// test / bench hook: bh_push once per message, driven from C (what the N-API shim does for each rmsg.message.content)
int bh_debug_push_each(bh_ctx* c, const uint8_t* slab, const uint32_t* off, const uint64_t* ts, uint32_t n) {
  for (uint32_t i = 0; i < n; i++) {
    int rc = bh_push(c, slab + off[i], off[i + 1] - off[i], ts ? ts[i] : 0);
    if (rc != BH_OK) return rc;
  }
  return BH_OK;
}

// test hook: floor(n / d) through the same routine the kernels use for window ids
uint64_t bh_debug_div64(uint64_t n, uint64_t d) { return d ? div64(n, make_div64(d)) : 0; }

}  // extern "C"
