// microbenchmark: random read-modify-write in global memory on sm_100a, by footprint (L2-resident vs DRAM),
// and a partition scatter with shared-memory cursors.  Design input for the many-key aggregation path.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o gatoms gatoms.cu && ./gatoms
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 33; x *= 0xFF51AFD7ED558CCDull; x ^= x >> 33; x *= 0xC4CEB9FE1A85EC53ull; x ^= x >> 33; return x; }

struct Rec { unsigned long long key, count; double sum; unsigned long long mn, mx; unsigned long long pad[3]; };   // 64 B

// MODE 0: one u64 atomicAdd per op, random 8-byte word
// MODE 1: record update, 4 atomics in one 64-byte record (count, sum f64, max, max)
// MODE 2: record update as plain load + store (exclusive owner), 64 B
// MODE 3: three separate arrays (key read 8 B, stat 32 B: 4 atomics, hist u32 add in a 256 B row) = the round-1 layout
// MODE 4: record (4 atomics) + one u32 add in a 256-byte histogram row
// MODE 5: u32 hist add only
template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(uint8_t* base, uint64_t nrec, uint64_t ops_per_thread, uint64_t seed, uint32_t* hist,
                                             unsigned long long* keys, unsigned long long* stats) {
  uint64_t x = seed + blockIdx.x * 1024ull + threadIdx.x;
  Rec* recs = (Rec*)base;
  uint64_t acc = 0;
  for (uint64_t i = 0; i < ops_per_thread; i++) {
    x = mix(x + i);
    uint64_t r = __umul64hi(x, nrec);
    double v = (double)(x & 1023);
    if (MODE == 0) atomicAdd(&((unsigned long long*)base)[r], 1ull);
    if (MODE == 1 || MODE == 4) {
      Rec* q = recs + r;
      acc += q->key;
      atomicAdd(&q->count, 1ull); atomicAdd(&q->sum, v);
      if ((x & 1023) > q->mn) atomicMax(&q->mn, (unsigned long long)(x & 1023)); if ((x & 511) > q->mx) atomicMax(&q->mx, (unsigned long long)(x & 511));
      if (MODE == 4) atomicAdd(&hist[r * 64 + ((x >> 20) & 15) + 24], 1u);
    }
    if (MODE == 2) {
      Rec* q = recs + r;
      Rec t = *q;
      t.count += 1; t.sum += v; t.mn = max(t.mn, (unsigned long long)(x & 1023)); t.mx = max(t.mx, (unsigned long long)(x & 511));
      *q = t;
    }
    if (MODE == 3) {
      acc += keys[r];
      unsigned long long* s = stats + r * 4;
      atomicAdd(&s[0], 1ull); atomicAdd((double*)&s[1], v);
      if ((x & 1023) > s[2]) atomicMax(&s[2], (unsigned long long)(x & 1023)); if ((x & 511) > s[3]) atomicMax(&s[3], (unsigned long long)(x & 511));
      atomicAdd(&hist[r * 64 + ((x >> 20) & 15) + 24], 1u);
    }
    if (MODE == 5) atomicAdd(&hist[r * 64 + ((x >> 20) & 15) + 24], 1u);
  }
  if (acc == 0x1234567) hist[0] = 1;
}

// scatter: each thread appends 16-byte tuples to partition p's private region of its CTA; cursors in shared memory
template <int TB>
__global__ void __launch_bounds__(1024, 1) k_scatter(uint8_t* out, uint32_t P, uint32_t region_cap, uint64_t per_thread, uint64_t seed) {
  extern __shared__ uint32_t cur[];
  for (uint32_t i = threadIdx.x; i < P; i += blockDim.x) cur[i] = 0;
  __syncthreads();
  uint64_t x = seed + blockIdx.x * 1024ull + threadIdx.x;
  for (uint64_t i = 0; i < per_thread; i++) {
    x = mix(x + i);
    uint32_t p = (uint32_t)__umul64hi(x, P);
    uint32_t pos = atomicAdd(&cur[p], 1u);
    if (pos < region_cap) {
      uint8_t* dst = out + (((uint64_t)blockIdx.x * P + p) * region_cap + pos) * TB;
      if (TB == 16) *(uint4*)dst = make_uint4((uint32_t)x, (uint32_t)(x >> 32), (uint32_t)i, p);
      else { *(uint2*)dst = make_uint2((uint32_t)x, (uint32_t)(x >> 32)); *(uint2*)(dst + 8) = make_uint2((uint32_t)i, p); *(uint2*)(dst + 16) = make_uint2(p, pos); }
    }
  }
}

// scatter with ONE region per partition shared by all CTAs: cursor = a global atomic (with return) per tuple
__global__ void __launch_bounds__(1024, 1) k_scatter_g(uint8_t* out, uint32_t P, uint32_t region_cap, uint64_t per_thread, uint64_t seed,
                                                       unsigned int* gcur) {
  uint64_t x = seed + blockIdx.x * 1024ull + threadIdx.x;
  for (uint64_t i = 0; i < per_thread; i++) {
    x = mix(x + i);
    uint32_t p = (uint32_t)__umul64hi(x, P);
    uint32_t pos = atomicAdd(&gcur[p], 1u);
    if (pos < region_cap) *(uint4*)(out + ((uint64_t)p * region_cap + pos) * 16) = make_uint4((uint32_t)x, (uint32_t)(x >> 32), (uint32_t)i, p);
  }
}
void run_scatter_g(uint32_t P) {
  uint64_t per = 96;
  uint32_t region_cap = (uint32_t)(148ull * 1024 * per / P * 2 + 64);
  uint8_t* out; size_t bytes = (size_t)P * region_cap * 16;
  unsigned int* gcur;
  cudaMalloc(&out, bytes); cudaMalloc(&gcur, P * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaMemset(gcur, 0, P * 4);
  k_scatter_g<<<148, 1024>>>(out, P, region_cap, per, 1, gcur);
  float tot = 0;
  for (int r = 0; r < 4; r++) {
    cudaMemset(gcur, 0, P * 4);
    cudaEventRecord(e0);
    k_scatter_g<<<148, 1024>>>(out, P, region_cap, per, 5 + r, gcur);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); tot += ms;
  }
  double ops = 4.0 * 148 * 1024 * per;
  printf("scatter 16-byte tuples, GLOBAL cursors, P=%5u (buffer %5.0f MB)   %7.2f Gtuple/s  err=%s\n", P, bytes / 1048576.0, ops / tot / 1e6, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out); cudaFree(gcur);
}

template <int MODE> void run(const char* name, uint64_t bytes) {
  uint64_t nrec = MODE == 0 ? bytes / 8 : bytes / 64;
  if (MODE == 3) nrec = bytes / 40;
  if (MODE == 5) nrec = bytes / 256;
  uint8_t* base = nullptr; uint32_t* hist = nullptr; unsigned long long *keys = nullptr, *stats = nullptr;
  cudaMalloc(&base, bytes); cudaMemset(base, 0, bytes);
  if (MODE >= 3) { cudaMalloc(&hist, nrec * 256); cudaMemset(hist, 0, nrec * 256); } else cudaMalloc(&hist, 256);
  if (MODE == 3) { cudaMalloc(&keys, nrec * 8); cudaMalloc(&stats, nrec * 32); cudaMemset(stats, 0, nrec * 32); cudaMemset(keys, 0, nrec * 8); }
  uint64_t per = 128;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148, 1024>>>(base, nrec, per, 1, hist, keys, stats);
  cudaEventRecord(e0);
  for (int r = 0; r < 4; r++) k<MODE><<<148, 1024>>>(base, nrec, per, 77 + r, hist, keys, stats);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double ops = 4.0 * 148 * 1024 * per;
  printf("%-44s footprint %7.0f MB  %7.2f Gop/s  err=%s\n", name, bytes / 1048576.0 * (MODE == 4 ? 5 : 1), ops / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
  cudaFree(base); cudaFree(hist); if (keys) cudaFree(keys); if (stats) cudaFree(stats);
}

template <int TB> void run_scatter(uint32_t P) {
  uint64_t per = 96;                               // 148*1024*96 = 14.5 M tuples
  uint32_t region_cap = (uint32_t)(1024 * per / P * 2 + 64);
  uint8_t* out; size_t bytes = (size_t)148 * P * region_cap * TB;
  cudaMalloc(&out, bytes);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_scatter<TB><<<148, 1024, P * 4>>>(out, P, region_cap, per, 1);
  cudaEventRecord(e0);
  for (int r = 0; r < 4; r++) k_scatter<TB><<<148, 1024, P * 4>>>(out, P, region_cap, per, 5 + r);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  double ops = 4.0 * 148 * 1024 * per;
  printf("scatter %2d-byte tuples, P=%5u (buffer %5.0f MB)   %7.2f Gtuple/s = %6.0f GB/s  err=%s\n", TB, P, bytes / 1048576.0, ops / ms / 1e6, ops * TB / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}

int main(int argc, char** argv) {
  const uint64_t MB = 1048576;
  if (argc > 1) {
    for (uint32_t P : {256u, 1024u, 4096u, 16384u, 65536u}) { run_scatter_g(P); if (P <= 4096) run_scatter<16>(P); }
    return 0;
  }
  for (uint64_t f : {8 * MB, 32 * MB, 64 * MB, 128 * MB, 512 * MB, 2048 * MB}) {
    run<0>("u64 atomicAdd, random word", f);
    run<1>("64 B record: 4 atomics", f);
    run<2>("64 B record: plain load+store", f);
    run<3>("round-1 layout: key + 32 B stat + hist (x 7.4)", f);
    run<4>("64 B record + hist row (x 5)", f);
    run<5>("hist u32 add only (256 B rows)", f);
    printf("\n");
  }
  for (uint32_t P : {32u, 64u, 256u, 1024u, 4096u}) { run_scatter<16>(P); run_scatter<24>(P); }
  return 0;
}
