// bh_synth.cu — seeded synthetic api.TelemetryProgress streams, generated on the
// device for the benchmarks and on the host (same inline functions) for CPU-side tests.
#include <cub/device/device_scan.cuh>

#include "bh_common.cuh"

using namespace bh;

namespace {

__global__ void k_synth_sizes(const __grid_constant__ bh_synth_config c, uint64_t first, uint32_t n,
                              const uint64_t* __restrict__ cdf, uint32_t* sizes, uint64_t* ts) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += gridDim.x * blockDim.x) {
    if (i == n) { sizes[i] = 0; continue; }
    uint64_t t;
    sizes[i] = synth_build(c, first + i, cdf, nullptr, &t, 0xFFFFFFFFu);
    ts[i] = t;
  }
}

__global__ void k_synth_build(const __grid_constant__ bh_synth_config c, uint64_t first, uint32_t n,
                              const uint64_t* __restrict__ cdf, const uint32_t* __restrict__ offsets,
                              uint8_t* payload) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    synth_build(c, first + i, cdf, payload + offsets[i], nullptr, offsets[i + 1] - offsets[i]);
}

__global__ void k_synth_columns(const __grid_constant__ bh_synth_config c, uint64_t hash_seed, uint64_t first,
                                uint64_t n, const uint64_t* __restrict__ cdf, uint64_t* kh, uint64_t* ts,
                                double* val) {
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
    SynthEvent e;
    synth_event_fields(c, first + i, cdf, e);
    uint64_t h = hash_seed ^ kHashInit;
    for (int j = 0; j < 36; j += 8) {
      uint64_t w = 0;
      for (int b = 0; b < 8 && j + b < 36; b++) w |= (uint64_t)(uint8_t)e.media_id[j + b] << (8 * b);
      h = absorb(h, w);
    }
    h = absorb(h, len_word(36, 1));
    h = absorb(h, varint_word(e.status, 2));
    kh[i] = fmix64(h);
    ts[i] = e.ts_ms;
    val[i] = e.value;
  }
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

uint32_t synth_max_event(const bh_synth_config& c) {
  uint32_t natural = 38 + 1 + 3 + (c.value_f64 ? 9u : 2u) + 18u;
  return natural > c.payload_bytes ? natural : c.payload_bytes;
}

}  // namespace

extern "C" {

uint32_t bh_synth_event_size(const bh_synth_config* c, uint64_t g, const uint64_t* zipf_cdf) {
  return synth_build(*c, g, zipf_cdf, nullptr, nullptr, 0xFFFFFFFFu);
}

uint32_t bh_synth_event_host(const bh_synth_config* c, uint64_t g, const uint64_t* zipf_cdf, uint8_t* out, uint64_t* ts_ms) {
  uint32_t len = synth_build(*c, g, zipf_cdf, nullptr, nullptr, 0xFFFFFFFFu);
  return synth_build(*c, g, zipf_cdf, out, ts_ms, len);
}

int bh_synth_generate_device(int device, const bh_synth_config* c, uint64_t first, uint32_t n,
                             const uint64_t* d_zipf_cdf, uint8_t* d_payload, uint64_t payload_cap,
                             uint32_t* d_offsets, uint64_t* d_ts, uint64_t* total_bytes, void* stream) {
  if (!c || c->struct_size != sizeof(bh_synth_config) || !d_payload || !d_offsets || !d_ts) return BH_ERR_INVALID;
  if (c->zipf && !d_zipf_cdf) return BH_ERR_INVALID;
  if (c->n_keys == 0) return BH_ERR_INVALID;
  if ((uint64_t)synth_max_event(*c) * n > 0xFFFFFFF0ull) return BH_ERR_TOO_LARGE;
  DeviceGuard guard(device);
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t blocks = (n + 256) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  k_synth_sizes<<<blocks, 256, 0, st>>>(*c, first, n, d_zipf_cdf, d_offsets, d_ts);
  size_t tmp_bytes = 0;
  if (cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, d_offsets, d_offsets, (int)(n + 1), st) != cudaSuccess) return BH_ERR_CUDA;
  void* tmp = nullptr;
  if (cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1) != cudaSuccess) return BH_ERR_NOMEM;
  cudaError_t e = cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, d_offsets, d_offsets, (int)(n + 1), st);
  uint32_t total = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&total, d_offsets + n, 4, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(tmp);
  if (e != cudaSuccess) return BH_ERR_CUDA;
  if (total_bytes) *total_bytes = total;
  if (((uint64_t)total + 15) / 16 * 16 > payload_cap) return BH_ERR_TOO_LARGE;
  k_synth_build<<<blocks, 256, 0, st>>>(*c, first, n, d_zipf_cdf, d_offsets, d_payload);
  return cudaGetLastError() == cudaSuccess ? BH_OK : BH_ERR_CUDA;
}

int bh_synth_columns_device(int device, const bh_synth_config* c, uint64_t hash_seed, uint64_t first,
                            uint64_t n, const uint64_t* d_zipf_cdf, uint64_t* d_key_hash, uint64_t* d_ts,
                            double* d_value, void* stream) {
  if (!c || c->struct_size != sizeof(bh_synth_config) || !d_key_hash || !d_ts || !d_value) return BH_ERR_INVALID;
  if (c->zipf && !d_zipf_cdf) return BH_ERR_INVALID;
  if (c->n_keys == 0) return BH_ERR_INVALID;
  DeviceGuard guard(device);
  uint64_t blocks = (n + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks == 0) return BH_OK;
  k_synth_columns<<<(uint32_t)blocks, 256, 0, (cudaStream_t)stream>>>(*c, hash_seed, first, n, d_zipf_cdf, d_key_hash, d_ts, d_value);
  return cudaGetLastError() == cudaSuccess ? BH_OK : BH_ERR_CUDA;
}

}  // extern "C"
