// microbenchmark: open-addressing probe loop in shared memory (lookups of existing keys), variants
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
constexpr int NS = 4096;
__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xFF51AFD7ED558CCDull; x ^= x >> 33; x *= 0xC4CEB9FE1A85EC53ull; x ^= x >> 33; return x; }
__device__ __forceinline__ uint32_t home(uint64_t kh) { uint32_t x = (uint32_t)kh ^ (uint32_t)(kh >> 32); return (x * 0x9E3779B1u) >> 20; }

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(uint32_t* out, int nkeys, int iters, long long* cyc, uint32_t* probes_out) {
  extern __shared__ __align__(16) uint8_t sm[];
  unsigned long long* key = (unsigned long long*)sm;
  uint8_t* wrel = sm + NS * 8;
  uint32_t* cnt = (uint32_t*)(sm + NS * 9);
  for (int i = threadIdx.x; i < NS; i += blockDim.x) { key[i] = ~0ull; wrel[i] = 0xFF; cnt[i] = 0; }
  __syncthreads();
  if (threadIdx.x == 0) {   // serial fill
    for (int j = 0; j < nkeys; j++) {
      uint64_t kh = mix(j + 1);
      uint32_t s = home(kh);
      while (key[s] != ~0ull) s = (s + 1) & (NS - 1);
      key[s] = kh; wrel[s] = 1;
    }
  }
  __syncthreads();
  uint32_t x = (threadIdx.x * 2654435761u) ^ (blockIdx.x * 40503u) ^ 12345u;
  uint32_t acc = 0, probes = 0;
  volatile unsigned long long* vkey = key; volatile uint8_t* vwrel = wrel;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
    x = x * 1664525u + 1013904223u;
    uint64_t kh = mix((x >> 8) % nkeys + 1);
    uint32_t slot = home(kh);
    int found = -1;
    if (MODE == 0) {            // volatile C++ loads, unroll 1 (as in the library)
#pragma unroll 1
      for (int p = 0; p < 16; p++) {
        probes++;
        unsigned long long kk = vkey[slot];
        if (kk == kh) { uint32_t w = vwrel[slot]; while (w == 0xFF) w = vwrel[slot]; if (w == 1) { found = slot; break; } }
        slot = (slot + 1) & (NS - 1);
      }
    } else if (MODE == 1) {     // plain loads
#pragma unroll 1
      for (int p = 0; p < 16; p++) {
        probes++;
        unsigned long long kk = key[slot];
        if (kk == kh && wrel[slot] == 1) { found = slot; break; }
        slot = (slot + 1) & (NS - 1);
      }
    } else if (MODE == 2) {     // warp-uniform loop: all lanes iterate until all found (no per-lane break)
      bool done = false;
#pragma unroll 1
      for (int p = 0; p < 16; p++) {
        probes++;
        unsigned long long kk = key[slot];
        if (!done && kk == kh && wrel[slot] == 1) { found = slot; done = true; }
        if (__all_sync(0xFFFFFFFFu, done)) break;
        if (!done) slot = (slot + 1) & (NS - 1);
      }
    } else if (MODE == 3) {     // two slots per probe step (LDS.128)
#pragma unroll 1
      for (int p = 0; p < 8; p++) {
        probes++;
        uint32_t s2 = slot & ~1u;
        ulonglong2 kk = *reinterpret_cast<ulonglong2*>(&key[s2]);
        if (kk.x == kh && s2 >= slot) { found = s2; break; }
        if (kk.y == kh) { found = s2 + 1; break; }
        slot = (s2 + 2) & (NS - 1);
      }
    } else if (MODE == 4) {     // no probing: direct slot (lower bound)
      found = slot; probes++;
    }
    if (found >= 0) atomicAdd(&cnt[found], 1u);
    acc += found;
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  probes = __reduce_add_sync(0xFFFFFFFFu, probes);
  if ((threadIdx.x & 31) == 0) atomicAdd(probes_out, probes);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc + cnt[threadIdx.x];
}
template <int MODE> void run(const char* name, int nkeys) {
  uint32_t* out; long long* cyc; uint32_t* pr; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8); cudaMalloc(&pr, 4);
  int iters = 2048;
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  k<MODE><<<148, 1024, 65536>>>(out, nkeys, iters, cyc, pr); cudaDeviceSynchronize();
  cudaMemset(pr, 0, 4);
  k<MODE><<<148, 1024, 65536>>>(out, nkeys, iters, cyc, pr); cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost); uint32_t p; cudaMemcpy(&p, pr, 4, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < 148; i++) avg += h[i]; avg /= 148;
  printf("%-34s keys=%5d  %8.1f cycles per warp-lookup per SM (latency per warp %.0f)  probes/lookup %.2f  err=%s\n", name, nkeys,
         avg / (32.0 * iters), avg / iters, (double)p / (148.0 * 1024 * iters), cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(cyc); cudaFree(pr);
}
int main() {
  for (int nk : {60, 1000, 1500, 2000}) {
    run<0>("volatile loads, per-lane break", nk); run<1>("plain loads, per-lane break", nk);
    run<2>("warp-uniform loop", nk); run<3>("LDS.128 two slots per step", nk); run<4>("no probing (direct)", nk);
  }
  return 0;
}
