// microbenchmark: how a warp's batch of 32 x 64-byte messages gets from global into shared memory and is read
// back by its lanes (one lane per message, lanes in lock step) -- the staging question of the fused kernel.
//   0  LDG.128 -> 4 x STS.32, 4-byte skew every 64 bytes (stride 68): lanes read 16 words with LDS.32   [the kernel's way]
//   1  cp.async.bulk (1-D TMA bulk copy, mbarrier) into an unskewed buffer: lanes read 16 words with LDS.32
//   2  cp.async.bulk, unskewed: lanes read 4 x LDS.128
//   3  LDG.128 -> STS.128, 16-byte skew (stride 80): lanes read 4 x LDS.128
//   4  LDG.128 -> STS.128, 16-byte skew (stride 80): lanes read 16 words with LDS.32
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o stage stage.cu && ./stage
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdlib>

constexpr int kWarps = 32, kBatchBytes = 2048, kBufBytes = 2048 + 2048 / 4 + 128;

__device__ __forceinline__ uint32_t lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint4 lds128(uint32_t a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void sts128(uint32_t a, uint4 v) { asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory"); }
__device__ __forceinline__ uint4 ldg(const uint4* p) { uint4 v; asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)); return v; }

template <int MODE, int WORK>
__global__ void __launch_bounds__(kWarps * 32, 1) k(const uint8_t* __restrict__ src, uint64_t batches, uint32_t* out) {
  extern __shared__ __align__(128) uint8_t sm[];
  __shared__ __align__(8) unsigned long long mbar[kWarps];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t buf = (uint32_t)__cvta_generic_to_shared(sm) + warp * kBufBytes;
  const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar[warp]);
  if (MODE == 1 || MODE == 2) {
    if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mb));
    __syncwarp();
  }
  uint32_t acc = 0, phase = 0;
  const uint64_t per = batches / gridDim.x, b0 = blockIdx.x * per;
  for (uint64_t b = b0 + warp; b < b0 + per; b += kWarps) {
    const uint8_t* g = src + b * kBatchBytes;
    if (MODE == 0 || MODE == 3 || MODE == 4) {
      uint4 v[4];
#pragma unroll
      for (int k = 0; k < 4; k++) v[k] = ldg((const uint4*)g + lane + 32 * k);
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t x = (lane + 32 * k) * 16;               // logical byte
        if (MODE == 0) { const uint32_t d = buf + x + 4 * (x >> 6); sts32(d, v[k].x); sts32(d + 4, v[k].y); sts32(d + 8, v[k].z); sts32(d + 12, v[k].w); }
        else sts128(buf + x + 16 * (x >> 6), v[k]);
      }
      __syncwarp();
    } else {
      if (lane == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mb), "r"(kBatchBytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(buf), "l"(g), "r"(kBatchBytes), "r"(mb) : "memory");
      }
      uint32_t done = 0;
      while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mb), "r"(phase) : "memory");
      phase ^= 1;
    }
    // one lane per 64-byte message: read all 16 words
    const uint32_t m = lane * 64;
    if (MODE == 0) { const uint32_t a = buf + m + 4 * lane;
#pragma unroll
      for (int j = 0; j < 16; j++) acc += lds32(a + 4 * j) * (j + 1); }
    if (MODE == 1) { const uint32_t a = buf + m;
#pragma unroll
      for (int j = 0; j < 16; j++) acc += lds32(a + 4 * j) * (j + 1); }
    if (MODE == 2 || MODE == 3) { const uint32_t a = buf + m + (MODE == 3 ? 16 * lane : 0);
#pragma unroll
      for (int j = 0; j < 4; j++) { const uint4 q = lds128(a + 16 * j); acc += q.x * (4 * j + 1) + q.y * (4 * j + 2) + q.z * (4 * j + 3) + q.w * (4 * j + 4); } }
    if (MODE == 4) { const uint32_t a = buf + m + 16 * lane;
#pragma unroll
      for (int j = 0; j < 16; j++) acc += lds32(a + 4 * j) * (j + 1); }
    // stand-in for decode + hash + table update: ~450 dependent integer instructions per lane, so that the
    // kernel is bound by instruction issue like the real one and the staging instructions compete for slots
    if (WORK) {
      uint32_t h = acc, g2 = acc ^ 0x9E3779B9u;
#pragma unroll 1
      for (int r = 0; r < WORK; r++) { h = (h ^ (g2 >> 15)) * 0x85EBCA6Bu; g2 = (g2 ^ (h >> 13)) * 0xC2B2AE35u + r; h += g2 << 3; g2 ^= h >> 7; }
      acc = h ^ g2;
    }
    __syncwarp();
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int MODE, int WORK> void run(const char* name, const uint8_t* src, uint64_t batches, uint32_t* out) {
  cudaFuncSetAttribute(k<MODE, WORK>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWarps * kBufBytes);
  k<MODE, WORK><<<148, kWarps * 32, kWarps * kBufBytes>>>(src, batches, out);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  for (int r = 0; r < 5; r++) k<MODE, WORK><<<148, kWarps * 32, kWarps * kBufBytes>>>(src, batches, out);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
  uint32_t h; cudaMemcpy(&h, out, 4, cudaMemcpyDeviceToHost);
  printf("work %2d  %-78s %7.3f ms  %6.0f GB/s  (check %08x) %s\n", WORK, name, ms, batches * (double)kBatchBytes / ms / 1e6, h, cudaGetErrorString(cudaGetLastError()));
}

int main() {
  const uint64_t batches = 148ull * 32 * 83;          // ~0.8 GB
  uint8_t* src; uint32_t* out;
  cudaMalloc(&src, batches * kBatchBytes + 4096); cudaMalloc(&out, 148 * 1024 * 4);
  cudaMemset(src, 0x5A, batches * kBatchBytes);
  // random bytes (constant data flatters the memory system)
  {
    uint32_t* h = (uint32_t*)malloc(1 << 24); uint32_t x = 12345;
    for (int i = 0; i < (1 << 22); i++) { x = x * 1664525u + 1013904223u; h[i] = x; }
    for (uint64_t o = 0; o < batches * kBatchBytes; o += 1 << 24) cudaMemcpy(src + o, h, std::min<uint64_t>(1 << 24, batches * kBatchBytes - o), cudaMemcpyHostToDevice);
    free(h);
  }
#define ALL(W) \
  run<0, W>("0 LDG.128 -> 4 x STS.32, stride 68; 16 x LDS.32 per lane", src, batches, out); \
  run<1, W>("1 cp.async.bulk, unskewed; 16 x LDS.32 per lane (16-way conflicts)", src, batches, out); \
  run<2, W>("2 cp.async.bulk, unskewed; 4 x LDS.128 per lane", src, batches, out); \
  run<3, W>("3 LDG.128 -> STS.128, stride 80; 4 x LDS.128 per lane", src, batches, out); \
  run<4, W>("4 LDG.128 -> STS.128, stride 80; 16 x LDS.32 per lane (4-way conflicts)", src, batches, out);
  ALL(0) ALL(40) ALL(56)
  return 0;
}
