// bh_group.cu — multi-GPU window close inside libbeholder_gpu.so (SURVEY §8 b/e): no PyTorch, no NCCL.
//
// Events are spread over the GPUs by arrival and every GPU aggregates the keys it sees (its `local` context); a
// key's merged row lives on its OWNER (owner = top log2(world) bits of the key hash, the `owned` context of that
// GPU).  Closing a window is, on every GPU and on its own stream:
//     export kernel  : compacts the window's rows by owner and STORES them over NVLink straight into block
//                      `rank` of the owner's receive buffer (bh_export_window_peers) -- compaction and exchange are
//                      one kernel, there is no send buffer and no all-to-all
//     k_post_counts  : tells every owner how many rows this GPU sent it
//     k_group_barrier: one flag per (sender, receiver) in peer memory, release/acquire at system scope
//     merge kernel   : folds the received blocks into the owned table (bh_merge_blocks_device), reading exactly
//                      the announced rows, so receive buffers are never cleared
//     close          : recycles the window's slot in the local table
// Receive buffers are double-buffered by close number, which is why ONE barrier per close is enough: a GPU can
// store into buffer b again only two closes later, after a barrier that its owner reaches only once the merge of
// the earlier close is behind it on its stream.
//
// A member's peers are either contexts of the same process (bh_create_group: plain device pointers after
// cudaDeviceEnablePeerAccess) or live in other processes (one process per GPU, as `torchrun` starts them: the
// receive area travels as a CUDA IPC handle, exchanged by whatever the host has -- bench.py uses the
// torch.distributed store for that bootstrap and nothing else).
// Peer memory is only ever written with plain stores and read by its own GPU: no TMA, no multicast.
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "bh_common.cuh"

namespace {

constexpr uint32_t kMaxWorld = 64;
thread_local std::string g_group_error;

int gfail(int code, const char* what, cudaError_t e = cudaSuccess) {
  g_group_error = what;
  if (e != cudaSuccess) { g_group_error += ": "; g_group_error += cudaGetErrorString(e); }
  return code;
}
#define GCU(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return gfail(BH_ERR_CUDA, #call, e__); } while (0)

// rows this GPU sent to owner o (cut at the block size) -> slot `rank` of o's count array
__global__ void k_post_counts(const unsigned long long* owner_counts, const unsigned long long* peer_counts, uint32_t world,
                              uint32_t rank, uint64_t stride, unsigned int* overflow) {
  const uint32_t o = threadIdx.x;
  if (o >= world) return;
  const unsigned long long n = owner_counts[o];
  if (n > stride) atomicAdd(overflow, 1u);
  ((unsigned long long*)peer_counts[o])[rank] = n < stride ? n : stride;
}

// Cross-GPU barrier on `stream`.  flags[r] of GPU g = the last epoch GPU r has reached, written by r.  Everything
// this GPU stored into peer memory earlier on the stream is ordered before its flag by the system-scope fence and
// the release store; the acquire loads order the merge that follows after the peers' flags.
__global__ void k_group_barrier(const unsigned long long* peer_flags, unsigned int* my_flags, uint32_t world, uint32_t rank,
                                unsigned int epoch) {
  const uint32_t t = threadIdx.x;
  if (t >= world) return;
  __threadfence_system();
  unsigned int* dst = (unsigned int*)peer_flags[t] + rank;
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(dst), "r"(epoch) : "memory");
  unsigned int seen;
  do {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(my_flags + t) : "memory");
  } while ((int)(seen - epoch) < 0);
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

struct bh_member {
  bh_ctx *local = nullptr, *owned = nullptr;
  uint32_t rank = 0, world = 1, hist_buckets = 0;
  uint64_t stride = 0;
  int device = 0;
  // receive area of this GPU: 2 x { recs[world][stride] | hist[world][stride][B] | counts[world] } | flags[world]
  uint8_t* area = nullptr;
  size_t buf_bytes = 0, rec_bytes = 0, hist_bytes = 0, area_bytes = 0;
  uint8_t* peer_area[kMaxWorld] = {nullptr};
  bool ipc_open[kMaxWorld] = {false};
  bool connected = false;
  unsigned long long* d_tables = nullptr;   // device: [2][3][world] addresses (recs, hist, counts of every peer) + [world] flags
  unsigned long long* d_owner_counts = nullptr;
  unsigned int* d_overflow = nullptr;
  unsigned int epoch = 0;          // closes done; buffer = epoch & 1, barrier value = epoch + 1
  cudaEvent_t ev = nullptr;        // "my export of this close is done" (single-process groups)
  uint64_t closes = 0;
  uint8_t* buf(uint32_t peer, uint32_t b) const { return peer_area[peer] + (size_t)b * buf_bytes; }
};

struct bh_group {
  std::vector<bh_member*> members;
  std::vector<bh_ctx*> ctxs;     // owned by the group (local, owned per rank)
};

extern "C" {

const char* bh_group_last_error(void) { return g_group_error.c_str(); }

int bh_member_create(bh_ctx* local, bh_ctx* owned, int device, uint32_t rank, uint32_t world, uint64_t owner_stride,
                     uint32_t hist_buckets, bh_member** out) {
  if (!local || !owned || !out) return gfail(BH_ERR_INVALID, "null argument");
  if (world == 0 || world > kMaxWorld || (world & (world - 1)) || rank >= world) return gfail(BH_ERR_INVALID, "world must be a power of two <= 64 and rank below it");
  if (owner_stride == 0) return gfail(BH_ERR_INVALID, "owner_stride must be positive");
  bh_member* m = new bh_member();
  m->local = local; m->owned = owned; m->rank = rank; m->world = world; m->stride = owner_stride;
  m->hist_buckets = hist_buckets; m->device = device;
  m->rec_bytes = align_up((size_t)world * owner_stride * sizeof(bh_record), 256);
  m->hist_bytes = align_up((size_t)world * owner_stride * hist_buckets * 4, 256);
  m->buf_bytes = m->rec_bytes + m->hist_bytes + align_up((size_t)world * 8, 256);
  m->area_bytes = 2 * m->buf_bytes + align_up((size_t)world * 4, 256);
  int prev = 0;
  cudaGetDevice(&prev);
  cudaError_t e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaMalloc((void**)&m->area, m->area_bytes);
  if (e == cudaSuccess) e = cudaMemset(m->area, 0, m->area_bytes);
  if (e == cudaSuccess) e = cudaMalloc((void**)&m->d_tables, (size_t)(2 * 3 + 1) * world * 8);
  if (e == cudaSuccess) e = cudaMalloc((void**)&m->d_owner_counts, (size_t)world * 8);
  if (e == cudaSuccess) e = cudaMalloc((void**)&m->d_overflow, 4);
  if (e == cudaSuccess) e = cudaMemset(m->d_overflow, 0, 4);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&m->ev, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  cudaSetDevice(prev);
  if (e != cudaSuccess) { bh_member_destroy(m); return gfail(e == cudaErrorMemoryAllocation ? BH_ERR_NOMEM : BH_ERR_CUDA, "bh_member_create", e); }
  m->peer_area[rank] = m->area;
  *out = m;
  return BH_OK;
}

void bh_member_destroy(bh_member* m) {
  if (!m) return;
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(m->device);
  cudaDeviceSynchronize();
  for (uint32_t p = 0; p < m->world; p++) if (m->ipc_open[p] && m->peer_area[p]) cudaIpcCloseMemHandle(m->peer_area[p]);
  if (m->area) cudaFree(m->area);
  if (m->d_tables) cudaFree(m->d_tables);
  if (m->d_owner_counts) cudaFree(m->d_owner_counts);
  if (m->d_overflow) cudaFree(m->d_overflow);
  if (m->ev) cudaEventDestroy(m->ev);
  cudaSetDevice(prev);
  delete m;
}

int bh_member_ipc_handle(bh_member* m, void* out64) {
  if (!m || !out64) return gfail(BH_ERR_INVALID, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == BH_IPC_HANDLE_BYTES, "handle size");
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(m->device);
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, m->area);
  cudaSetDevice(prev);
  if (e != cudaSuccess) return gfail(BH_ERR_CUDA, "cudaIpcGetMemHandle", e);
  memcpy(out64, &h, sizeof h);
  return BH_OK;
}

static int member_finish_connect(bh_member* m) {
  const uint32_t W = m->world;
  std::vector<unsigned long long> t((size_t)(2 * 3 + 1) * W);
  for (uint32_t b = 0; b < 2; b++)
    for (uint32_t p = 0; p < W; p++) {
      t[(b * 3 + 0) * W + p] = (unsigned long long)(uintptr_t)m->buf(p, b);
      t[(b * 3 + 1) * W + p] = m->hist_buckets ? (unsigned long long)(uintptr_t)(m->buf(p, b) + m->rec_bytes) : 0ull;
      t[(b * 3 + 2) * W + p] = (unsigned long long)(uintptr_t)(m->buf(p, b) + m->rec_bytes + m->hist_bytes);
    }
  for (uint32_t p = 0; p < W; p++) t[6 * W + p] = (unsigned long long)(uintptr_t)(m->peer_area[p] + 2 * m->buf_bytes);
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(m->device);
  cudaError_t e = cudaMemcpy(m->d_tables, t.data(), t.size() * 8, cudaMemcpyHostToDevice);
  cudaSetDevice(prev);
  if (e != cudaSuccess) return gfail(BH_ERR_CUDA, "cudaMemcpy(peer tables)", e);
  m->connected = true;
  return BH_OK;
}

int bh_member_connect_ipc(bh_member* m, const void* handles) {
  if (!m || !handles) return gfail(BH_ERR_INVALID, "null argument");
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(m->device);
  for (uint32_t p = 0; p < m->world; p++) {
    if (p == m->rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, (const uint8_t*)handles + (size_t)p * BH_IPC_HANDLE_BYTES, sizeof h);
    void* ptr = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaSetDevice(prev); return gfail(BH_ERR_CUDA, "cudaIpcOpenMemHandle (peer GPU not reachable from this process?)", e); }
    m->peer_area[p] = (uint8_t*)ptr; m->ipc_open[p] = true;
  }
  cudaSetDevice(prev);
  return member_finish_connect(m);
}

int bh_member_connect_ptrs(bh_member* m, void* const* areas) {
  if (!m || !areas) return gfail(BH_ERR_INVALID, "null argument");
  for (uint32_t p = 0; p < m->world; p++) if (p != m->rank) m->peer_area[p] = (uint8_t*)areas[p];
  return member_finish_connect(m);
}

void* bh_member_area(bh_member* m) { return m ? m->area : nullptr; }

// A close in three steps on `st`: (1) rows and counts into the owners' buffers, (2) a barrier, (3) merge what
// arrived and recycle the window.  The barrier is the flag kernel when the peers are other processes, stream
// events when the whole group lives in this process (bh_group_close_window): a spinning kernel is only safe when
// its peers cannot be queued behind it on the same GPU.
static int member_export(bh_member* m, uint64_t window_id, cudaStream_t st) {
  const uint32_t W = m->world, b = m->epoch & 1u;
  const unsigned long long* tab = m->d_tables + (size_t)b * 3 * W;
  int rc = bh_export_window_peers(m->local, window_id, W, (const uint64_t*)tab, m->hist_buckets ? (const uint64_t*)(tab + W) : nullptr,
                                  m->rank, m->stride, (uint64_t*)m->d_owner_counts, st);
  if (rc != BH_OK) return gfail(rc, bh_last_error(m->local));
  k_post_counts<<<1, kMaxWorld, 0, st>>>(m->d_owner_counts, tab + 2 * W, W, m->rank, m->stride, m->d_overflow);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return gfail(BH_ERR_CUDA, "k_post_counts", e);
  return BH_OK;
}

static int member_finish(bh_member* m, uint64_t window_id, cudaStream_t st) {
  const uint32_t W = m->world, b = m->epoch & 1u;
  uint8_t* mine = m->buf(m->rank, b);
  int rc = bh_merge_blocks_device(m->owned, (const bh_record*)mine, m->hist_buckets ? (const uint32_t*)(mine + m->rec_bytes) : nullptr, W,
                                  m->stride, (const uint64_t*)(mine + m->rec_bytes + m->hist_bytes), st);
  if (rc == BH_OK) rc = bh_close_window_device(m->local, window_id, st);     // BH_ALL_WINDOWS: every open window
  if (rc == BH_OK) rc = bh_stream_wait(m->owned, st);      // reads of the owned table (snapshot ...) come after the merge
  if (rc != BH_OK) return gfail(rc, "merge / close after the exchange");
  m->epoch++;
  m->closes++;
  return BH_OK;
}

int bh_member_close_window(bh_member* m, uint64_t window_id, void* stream) {
  if (!m) return gfail(BH_ERR_INVALID, "null argument");
  if (!m->connected) return gfail(BH_ERR_INVALID, "member is not connected to its peers");
  cudaStream_t st = stream ? (cudaStream_t)stream : (cudaStream_t)bh_stream(m->local);
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(m->device);
  int rc = member_export(m, window_id, st);
  if (rc == BH_OK) {
    k_group_barrier<<<1, kMaxWorld, 0, st>>>(m->d_tables + 6 * m->world, (unsigned int*)(m->area + 2 * m->buf_bytes), m->world, m->rank,
                                             m->epoch + 1);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) rc = gfail(BH_ERR_CUDA, "k_group_barrier", e);
  }
  if (rc == BH_OK) rc = member_finish(m, window_id, st);
  cudaSetDevice(prev);
  return rc;
}

int bh_member_check(bh_member* m, uint64_t* owner_counts_out) {
  if (!m) return gfail(BH_ERR_INVALID, "null argument");
  int prev = 0;
  cudaGetDevice(&prev);
  cudaSetDevice(m->device);
  unsigned int ov = 0;
  std::vector<unsigned long long> cnt(m->world);
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemcpy(&ov, m->d_overflow, 4, cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(cnt.data(), m->d_owner_counts, (size_t)m->world * 8, cudaMemcpyDeviceToHost);
  cudaSetDevice(prev);
  if (e != cudaSuccess) return gfail(BH_ERR_CUDA, "bh_member_check", e);
  if (owner_counts_out) memcpy(owner_counts_out, cnt.data(), (size_t)m->world * 8);
  if (ov) return gfail(BH_ERR_TOO_LARGE, "a window close sent an owner more rows than owner_stride: rows were cut");
  return BH_OK;
}

// ---------------------------------------------------------------------------------------------------------
// single process, n GPUs
// ---------------------------------------------------------------------------------------------------------
int bh_create_group(const bh_config* local_cfg, const bh_config* owned_cfg, const int* devices, uint32_t n,
                    uint64_t owner_stride, bh_group** out) {
  if (!local_cfg || !owned_cfg || !devices || !out) return gfail(BH_ERR_INVALID, "null argument");
  *out = nullptr;
  bh_group* g = new bh_group();
  auto bail = [&](int code) { std::string msg = g_group_error; bh_group_destroy(g); g_group_error = msg; return code; };
  const uint32_t hist = (local_cfg->ops & BH_OP_HIST) ? (local_cfg->hist_buckets ? local_cfg->hist_buckets : BH_HIST_MAX) : 0;
  for (uint32_t r = 0; r < n; r++) {
    bh_config lc = *local_cfg, oc = *owned_cfg;
    lc.device = oc.device = devices[r];
    bh_ctx *l = nullptr, *o = nullptr;
    int rc = bh_create(&lc, &l);
    if (rc != BH_OK) { gfail(rc, bh_last_error(nullptr)); return bail(rc); }
    g->ctxs.push_back(l);
    rc = bh_create(&oc, &o);
    if (rc != BH_OK) { gfail(rc, bh_last_error(nullptr)); return bail(rc); }
    g->ctxs.push_back(o);
    bh_member* m = nullptr;
    rc = bh_member_create(l, o, devices[r], r, n, owner_stride, hist, &m);
    if (rc != BH_OK) return bail(rc);
    g->members.push_back(m);
  }
  int prev = 0;
  cudaGetDevice(&prev);
  for (uint32_t i = 0; i < n; i++)
    for (uint32_t j = 0; j < n; j++) {
      if (devices[i] == devices[j]) continue;
      cudaSetDevice(devices[i]);
      cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
      if (e != cudaSuccess) { cudaSetDevice(prev); gfail(BH_ERR_CUDA, "cudaDeviceEnablePeerAccess", e); return bail(BH_ERR_CUDA); }
    }
  cudaSetDevice(prev);
  std::vector<void*> areas(n);
  for (uint32_t r = 0; r < n; r++) areas[r] = g->members[r]->area;
  for (uint32_t r = 0; r < n; r++) {
    int rc = bh_member_connect_ptrs(g->members[r], areas.data());
    if (rc != BH_OK) return bail(rc);
  }
  *out = g;
  return BH_OK;
}

void bh_group_destroy(bh_group* g) {
  if (!g) return;
  for (bh_member* m : g->members) bh_member_destroy(m);
  for (bh_ctx* c : g->ctxs) bh_destroy(c);
  delete g;
}

uint32_t bh_group_size(bh_group* g) { return g ? (uint32_t)g->members.size() : 0; }
bh_ctx* bh_group_local(bh_group* g, uint32_t rank) { return g && rank < g->members.size() ? g->members[rank]->local : nullptr; }
bh_ctx* bh_group_owned(bh_group* g, uint32_t rank) { return g && rank < g->members.size() ? g->members[rank]->owned : nullptr; }
bh_member* bh_group_member(bh_group* g, uint32_t rank) { return g && rank < g->members.size() ? g->members[rank] : nullptr; }

int bh_group_close_window(bh_group* g, uint64_t window_id) {
  if (!g) return gfail(BH_ERR_INVALID, "null argument");
  // every GPU's sequence is queued on that GPU's own stream; the barriers inside meet on the devices
  for (bh_member* m : g->members) {
    int rc = bh_flush(m->local, nullptr);      // what was pushed through the ingest ring is aggregated first
    if (rc != BH_OK) return gfail(rc, bh_last_error(m->local));
  }
  int prev = 0;
  cudaGetDevice(&prev);
  int rc = BH_OK;
  for (bh_member* m : g->members) {                       // (1) every GPU exports to its peers
    cudaSetDevice(m->device);
    rc = member_export(m, window_id, (cudaStream_t)bh_stream(m->local));
    if (rc == BH_OK && cudaEventRecord(m->ev, (cudaStream_t)bh_stream(m->local)) != cudaSuccess) rc = gfail(BH_ERR_CUDA, "cudaEventRecord");
    if (rc != BH_OK) break;
  }
  for (bh_member* m : g->members) {                       // (2) every stream waits for every export, (3) merge
    if (rc != BH_OK) break;
    cudaSetDevice(m->device);
    for (bh_member* o : g->members)
      if (cudaStreamWaitEvent((cudaStream_t)bh_stream(m->local), o->ev, 0) != cudaSuccess) rc = gfail(BH_ERR_CUDA, "cudaStreamWaitEvent");
    if (rc == BH_OK) rc = member_finish(m, window_id, (cudaStream_t)bh_stream(m->local));
  }
  cudaSetDevice(prev);
  return rc;
}

int bh_group_sync(bh_group* g) {
  if (!g) return gfail(BH_ERR_INVALID, "null argument");
  for (bh_member* m : g->members) {
    int rc = bh_member_check(m, nullptr);
    if (rc != BH_OK) return rc;
  }
  return BH_OK;
}

}  // extern "C"
