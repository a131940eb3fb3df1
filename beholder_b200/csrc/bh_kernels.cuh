// bh_kernels.cuh — host-side launchers of the sm_100a kernels (defined in bh_kernels.cu).
#pragma once
#include "bh_common.cuh"

namespace bh {

struct LaunchGeom {
  int sms;            // SM count of the device
  int smem_optin;     // max opt-in dynamic shared memory per block
  int generic_only;   // 1 = never launch the kernel specialised on the reference's schema (env BEHOLDER_GENERIC_KERNELS=1; for tests)
};

// true when S is exactly the schema the specialised fused kernel has compiled in (ProgressSchema)
bool progress_schema_matches(const DevSchema& S);

// fused proto.decode + enum check + aggregate over one raw batch (index.js:129-138)
cudaError_t launch_process_raw(const DevSchema& S, const DevTables& T, const LaunchGeom& G,
                               const uint8_t* payload, const uint32_t* offsets, const uint64_t* ts,
                               uint32_t n, cudaStream_t st, int* launches, int big_override = -1,
                               uint32_t lo = 0, uint32_t hi = 0xFFFFFFFFu);
// big_override: -1 = follow the schema's message-size hint, 0 / 1 = small / large message kernels
// lo, hi: byte offsets the batch's messages may occupy; the kernel reads nothing outside, whatever the offsets say
// BH_F_VERIFY_KEYS: the key bytes of every event of the batch against the dictionary T.dict (no-op without one)
cudaError_t launch_verify_keys(const DevSchema& S, const DevTables& T, const LaunchGeom& G, const uint8_t* payload,
                               const uint32_t* offsets, uint32_t n, cudaStream_t st, int big_override = -1,
                               uint32_t lo = 0, uint32_t hi = 0xFFFFFFFFu);
// aggregate pre-decoded columns
cudaError_t launch_process_columns(const DevSchema& S, const DevTables& T, const LaunchGeom& G,
                                   const uint64_t* kh, const uint64_t* ts, const double* val,
                                   uint64_t n, cudaStream_t st, int* launches);
// partitioned mode (DevTables::part_*): fold the logged tuples into the tables, empty the log
cudaError_t launch_part_drain(const DevSchema& S, const DevTables& T, const LaunchGeom& G, bool intv, cudaStream_t st, int* launches);
// decode only -> columns
cudaError_t launch_decode(const DevSchema& S, const LaunchGeom& G, const uint8_t* payload,
                          const uint32_t* offsets, uint32_t n, uint64_t* kh, double* val,
                          int32_t* status, uint8_t* outcome, cudaStream_t st);
// compact one window into records grouped by owner (two passes: count, scatter)
cudaError_t launch_export(const DevSchema& S, const DevTables& T, uint64_t window_id, uint32_t n_owners,
                          bh_record* out, uint32_t* hist_out, uint64_t cap, uint64_t owner_stride,
                          unsigned long long* owner_counts, unsigned long long* owner_cursor,
                          cudaStream_t st, const unsigned long long* peer_base = nullptr,
                          const unsigned long long* peer_hist_base = nullptr, uint32_t my_rank = 0);
cudaError_t launch_export_static(const DevSchema& S, const DevTables& T, uint64_t window_id, uint32_t n_owners,
                                 bh_record* out, uint32_t* hist_out, uint64_t cap, uint64_t owner_stride,
                                 unsigned long long* owner_counts, cudaStream_t st);
cudaError_t launch_merge(const DevSchema& S, const DevTables& T, const bh_record* recs,
                         const uint32_t* hist, uint64_t n, cudaStream_t st);
cudaError_t launch_merge_blocks(const DevSchema& S, const DevTables& T, const bh_record* recs, const uint32_t* hist,
                                uint32_t n_blocks, uint64_t stride, const unsigned long long* counts, cudaStream_t st);
cudaError_t launch_close_window(const DevTables& T, uint64_t window_id, cudaStream_t st);
cudaError_t launch_reset_tables(const DevTables& T, cudaStream_t st);
// bh_compact.cu: u16 lengths + i16 timestamp deltas (+ per-block bases) -> u32 offsets, u64 timestamps
cudaError_t launch_expand_compact(const uint16_t* lengths, const unsigned long long* block_offsets, const unsigned long long* ts_block,
                                  const int16_t* ts_delta, uint32_t n, unsigned long long chunk_base, uint32_t* offsets,
                                  unsigned long long* ts, cudaStream_t st);
cudaError_t configure_kernels(const LaunchGeom& G);
cudaError_t launch_init_tables(const DevTables& T, cudaStream_t st);

}  // namespace bh
