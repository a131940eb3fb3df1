// microbenchmark: shared-memory atomic throughput on sm_100a (cycles per warp instruction per SM)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(uint32_t* out, int iters, uint32_t seed, long long* cyc) {
  extern __shared__ __align__(16) uint8_t sm[];
  uint32_t* s32 = (uint32_t*)sm; unsigned long long* s64 = (unsigned long long*)sm; double* sd = (double*)sm;
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) s32[i] = 0;
  __syncthreads();
  uint32_t x = seed ^ (threadIdx.x * 2654435761u) ^ (blockIdx.x * 40503u);
  uint32_t acc = 0;
  long long t0 = clock64();
  for (int i = 0; i < iters; i++) {
    x = x * 1664525u + 1013904223u;
    uint32_t idx = (x >> 8) & 2047;
    if (MODE == 0) atomicAdd(&s32[idx], 1u);                       // ATOMS.ADD no return use
    if (MODE == 1) acc += atomicAdd(&s32[idx], 1u);                // with return
    if (MODE == 2) atomicAdd(&sd[idx], 1.0);                       // f64 CAS loop
    if (MODE == 3) acc += (uint32_t)s64[idx];                      // LDS.64
    if (MODE == 4) atomicMax(&s64[idx], (unsigned long long)x);    // u64 CAS loop
    if (MODE == 5) { atomicAdd(&s32[idx], 1u); atomicAdd(&s32[idx + 2048], x & 127); }  // two adds
    if (MODE == 6) atomicAdd(&s32[(idx & 7) * 33], 1u);            // 8 hot addresses
    if (MODE == 7) atomicAdd((float*)&s32[idx], 1.0f);             // f32 add
    if (MODE == 8) atomicMax((int*)&s32[idx], (int)(x & 127));     // s32 max
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc + s32[threadIdx.x];
}
template <int MODE> void run(const char* name) {
  uint32_t* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  int iters = 4096;
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  k<MODE><<<148, 1024, 65536>>>(out, iters, 1, cyc); cudaDeviceSynchronize();
  k<MODE><<<148, 1024, 65536>>>(out, iters, 2, cyc); cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  double avg = 0; for (int i = 0; i < 148; i++) avg += h[i]; avg /= 148;
  // 32 warps x iters warp-instructions per SM
  printf("%-28s %8.2f cycles per warp-instr per SM (%.3f cyc/lane)  err=%s\n", name, avg / (32.0 * iters), avg / (32.0 * iters) / 32, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  run<0>("ATOMS.ADD u32 spread"); run<1>("ATOMS.ADD u32 +return"); run<2>("f64 add (CAS loop)"); run<3>("LDS.64");
  run<4>("u64 max (CAS loop)"); run<5>("2x ATOMS.ADD"); run<6>("ATOMS.ADD 8 hot addrs"); run<7>("f32 add"); run<8>("ATOMS.MAX s32");
  return 0;
}
