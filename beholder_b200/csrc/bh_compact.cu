// bh_compact.cu — compact sidecars of an ingest batch (bh_push_batch_compact).
//
// End to end the path is bound by the PCIe link, so bytes per event over the link are what is left to gain: a
// 64-byte event travels with a u32 offset and a u64 timestamp = 76 bytes.  The compact form sends a u16 LENGTH and
// an i16 timestamp DELTA per event, plus one u64 byte offset and one u64 base timestamp per block of 65 536 events
// = 68 bytes per event; this kernel rebuilds the u32 offsets and u64 timestamps the processing kernels read, in
// HBM (6 bytes read and 12 written per event, a few microseconds per million).  One CTA per block of 65 536
// events: 1 024 threads x 64 consecutive events, a block-wide exclusive scan of the lengths.
#include "bh_common.cuh"

namespace bh {

constexpr uint32_t kCompactBlock = 1u << BH_COMPACT_BLOCK_LG;   // events per block
constexpr uint32_t kCompactMaxBlocks = 64;                      // blocks per ring chunk (4 Mi events)
constexpr int kExpandThreads = 1024;
constexpr uint32_t kPerThread = kCompactBlock / kExpandThreads;   // 64

// per-block byte offsets and base timestamps of ONE ring chunk travel as kernel arguments (two more tiny H2D copies
// per chunk cost 5 % of the link: each one stalls the copy engine for a few microseconds)
struct CompactBlocks {
  unsigned long long block_offsets[kCompactMaxBlocks + 1];
  unsigned long long ts_block[kCompactMaxBlocks];
};

__global__ void __launch_bounds__(kExpandThreads)
k_expand_compact(const uint16_t* __restrict__ lengths, const __grid_constant__ CompactBlocks blocks,
                 const int16_t* __restrict__ ts_delta, uint32_t n,
                 unsigned long long chunk_base, uint32_t* __restrict__ offsets, unsigned long long* __restrict__ ts) {
  const unsigned long long* block_offsets = blocks.block_offsets;
  const unsigned long long* ts_block = blocks.ts_block;
  __shared__ uint32_t warp_sum[kExpandThreads / 32];
  const uint32_t b = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t i0 = b * kCompactBlock + threadIdx.x * kPerThread;
  // this thread's 64 lengths -> local exclusive prefix
  uint32_t local[kPerThread];
  uint32_t sum = 0;
#pragma unroll
  for (uint32_t j = 0; j < kPerThread; j++) {
    const uint32_t i = i0 + j;
    local[j] = sum;
    sum += i < n ? (uint32_t)lengths[i] : 0u;
  }
  // exclusive scan of the per-thread sums over the CTA
  uint32_t incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if ((int)lane >= d) incl += v; }
  if (lane == 31) warp_sum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = warp_sum[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, w, d); if ((int)lane >= d) w += v; }
    warp_sum[lane] = w;      // inclusive over warps
  }
  __syncthreads();
  const uint32_t base = (uint32_t)(block_offsets[b] - chunk_base) + (warp ? warp_sum[warp - 1] : 0u) + (incl - sum);
  const unsigned long long tb = ts && b * kCompactBlock < n ? ts_block[b] : 0ull;
#pragma unroll
  for (uint32_t j = 0; j < kPerThread; j++) {
    const uint32_t i = i0 + j;
    if (i < n) {
      offsets[i] = base + local[j];
      if (ts) ts[i] = tb + (unsigned long long)(long long)ts_delta[i];
    } else if (i == n) {
      offsets[n] = base + local[j];        // the end of the last message
    }
  }
}

// offsets[0..n] and ts[0..n) of a chunk of n events that starts on a block boundary; block_offsets / ts_block point
// at the chunk's first block; chunk_base = byte offset of the chunk's first message in the caller's slab
// block_offsets / ts_block: HOST pointers to the chunk's first block (at most kCompactMaxBlocks blocks per chunk)
cudaError_t launch_expand_compact(const uint16_t* lengths, const unsigned long long* block_offsets, const unsigned long long* ts_block,
                                  const int16_t* ts_delta, uint32_t n, unsigned long long chunk_base, uint32_t* offsets,
                                  unsigned long long* ts, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  const uint32_t nb = (n + kCompactBlock - 1) / kCompactBlock;
  if (nb > kCompactMaxBlocks) return cudaErrorInvalidValue;
  CompactBlocks B;
  for (uint32_t k = 0; k <= nb; k++) B.block_offsets[k] = block_offsets[k];
  for (uint32_t k = nb + 1; k <= kCompactMaxBlocks; k++) B.block_offsets[k] = block_offsets[nb];
  for (uint32_t k = 0; k < kCompactMaxBlocks; k++) B.ts_block[k] = ts && k < nb ? ts_block[k] : 0ull;
  const uint32_t blocks = n / kCompactBlock + 1;     // (+1: the thread that owns index n writes offsets[n])
  k_expand_compact<<<blocks, kExpandThreads, 0, st>>>(lengths, B, ts_delta, n, chunk_base, offsets, ts);
  return cudaGetLastError();
}

}  // namespace bh
