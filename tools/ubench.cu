// Micro-benchmarks that decide kernel design choices on B200 (run under gpurun):
//   HBM read / copy bandwidth with the kernels' own load pattern, shared-memory atomic rates
//   (conflict-free / random / hot bin), global RED rates (10^6 and 2*10^4 bins, 32/64 bit),
//   match.any vs the 8-ballot loop.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/ubench tools/ubench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e = (x);                                                       \
    if (e != cudaSuccess) {                                                    \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); \
      exit(1);                                                                 \
    }                                                                          \
  } while (0)

__device__ __forceinline__ int4 ldv4(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__global__ void k_read(const int4* __restrict__ a, size_t nvec, unsigned* sink) {
  unsigned acc = 0;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride * 4) {
    int4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) v[u] = ldv4(a + i + u * stride);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

__global__ void k_copy(const int4* __restrict__ a, int4* __restrict__ b, size_t nvec) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) b[i] = ldv4(a + i);
}

__device__ __forceinline__ uint32_t rnd(uint32_t& s) {
  s ^= s << 13;
  s ^= s >> 17;
  s ^= s << 5;
  return s;
}

// mode 0: lane-replicated (bin*32+lane), 1: random over nbins, 2: half of the hits on bin 0
template <int MODE>
__global__ void k_smem_atomic(int iters, int nbins, unsigned* sink) {
  extern __shared__ uint32_t sh[];
  for (int i = threadIdx.x; i < nbins * (MODE == 0 ? 32 : 1); i += blockDim.x) sh[i] = 0;
  __syncthreads();
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  const unsigned lane = threadIdx.x & 31;
  for (int i = 0; i < iters; ++i) {
    uint32_t r = rnd(s);
    uint32_t b = (r >> 4) % (uint32_t)nbins;
    if (MODE == 2 && (r & 1)) b = 0;
    if (MODE == 0)
      atomicAdd(sh + b * 32 + lane, 1u);
    else
      atomicAdd(sh + b, 1u);
  }
  __syncthreads();
  if (sh[threadIdx.x % nbins] == 0xffffffffu) *sink = 1;
}

// same loop without the atomic: cost of the RNG + modulo, to subtract
__global__ void k_smem_noatomic(int iters, int nbins, unsigned* sink) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u, acc = 0;
  for (int i = 0; i < iters; ++i) {
    uint32_t r = rnd(s);
    acc += (r >> 4) % (uint32_t)nbins;
  }
  if (acc == 0x12345678u) *sink = acc;
}

template <typename T>
__global__ void k_red(T* bins, uint32_t nbins, int iters) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 777u;
  for (int i = 0; i < iters; ++i) {
    const uint32_t b = (rnd(s) >> 3) % nbins;
    atomicAdd(bins + b, (T)1);
  }
}

__global__ void k_match_hw(int iters, unsigned* sink) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 99u, acc = 0;
  for (int i = 0; i < iters; ++i) {
    const uint32_t d = rnd(s) & 0xff;
    acc += __match_any_sync(0xffffffffu, d);
  }
  if (acc == 0x12345678u) *sink = acc;
}
__global__ void k_match_ballot(int iters, unsigned* sink) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 99u, acc = 0;
  for (int i = 0; i < iters; ++i) {
    const uint32_t d = rnd(s) & 0xff;
    uint32_t peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
      const bool bit = (d >> b) & 1;
      const uint32_t m = __ballot_sync(0xffffffffu, bit);
      peers &= bit ? m : ~m;
    }
    acc += peers;
  }
  if (acc == 0x12345678u) *sink = acc;
}
__global__ void k_match_none(int iters, unsigned* sink) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 99u, acc = 0;
  for (int i = 0; i < iters; ++i) acc += rnd(s) & 0xff;
  if (acc == 0x12345678u) *sink = acc;
}

template <typename F>
float time_ms(F f, int reps = 5) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  f();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(a));
    f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best;
}

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  printf("device %s sms %d max_clock_mhz %d\n", prop.name, sms, clk_khz / 1000);
  unsigned* sink;
  CK(cudaMalloc(&sink, 4));

  {  // bandwidth
    const size_t bytes = (size_t)4 << 30;
    int4 *a, *b;
    CK(cudaMalloc(&a, bytes));
    CK(cudaMalloc(&b, bytes));
    CK(cudaMemset(a, 1, bytes));
    CK(cudaMemset(b, 2, bytes));
    const size_t nvec = bytes / 16;
    for (int occ : {2, 4, 8}) {
      float ms = time_ms([&] { k_read<<<sms * occ, 256>>>(a, nvec, sink); });
      printf("read  4GiB grid=%dx%d  %.3f ms  %.1f GB/s\n", sms, occ, ms, bytes / ms / 1e6);
    }
    float ms = time_ms([&] { k_copy<<<sms * 8, 256>>>(a, b, nvec); });
    printf("copy  4GiB            %.3f ms  %.1f GB/s (r+w)\n", ms, 2.0 * bytes / ms / 1e6);
    CK(cudaFree(a));
    CK(cudaFree(b));
  }

  {  // smem atomics: per-SM rate
    const int iters = 4096, threads = 1024;
    const double total = (double)sms * threads * iters;
    float base = time_ms([&] { k_smem_noatomic<<<sms, threads>>>(iters, 256, sink); });
    printf("smem loop overhead (no atomic)         %.3f ms\n", base);
    CK(cudaFuncSetAttribute(k_smem_atomic<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
    CK(cudaFuncSetAttribute(k_smem_atomic<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
    CK(cudaFuncSetAttribute(k_smem_atomic<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024));
    float ms = time_ms([&] { k_smem_atomic<0><<<sms, threads, 256 * 32 * 4>>>(iters, 256, sink); });
    printf("ATOMS lane-replicated 256 bins         %.3f ms  %.1f G/s  (%.2f lanes/clk/SM at max clock)\n", ms,
           total / ms / 1e6, total / ms / 1e6 / sms / (clk_khz / 1e6));
    for (int nb : {256, 1024, 20002, 32768}) {
      ms = time_ms([&] { k_smem_atomic<1><<<sms, threads, nb * 4>>>(iters, nb, sink); });
      printf("ATOMS random %6d bins                %.3f ms  %.1f G/s  (%.2f lanes/clk/SM)\n", nb, ms,
             total / ms / 1e6, total / ms / 1e6 / sms / (clk_khz / 1e6));
    }
    ms = time_ms([&] { k_smem_atomic<2><<<sms, threads, 256 * 4>>>(iters, 256, sink); });
    printf("ATOMS 256 bins, half on one bin        %.3f ms  %.1f G/s\n", ms, total / ms / 1e6);
  }

  {  // global RED
    const int iters = 1024, threads = 256, grid = sms * 8;
    const double total = (double)grid * threads * iters;
    for (uint32_t nb : {1000000u, 20002u, 100u}) {
      uint32_t* b32;
      unsigned long long* b64;
      CK(cudaMalloc(&b32, (size_t)nb * 4));
      CK(cudaMalloc(&b64, (size_t)nb * 8));
      CK(cudaMemset(b32, 0, (size_t)nb * 4));
      CK(cudaMemset(b64, 0, (size_t)nb * 8));
      float ms = time_ms([&] { k_red<uint32_t><<<grid, threads>>>(b32, nb, iters); });
      printf("RED.32 random %8u bins             %.3f ms  %.1f G/s\n", nb, ms, total / ms / 1e6);
      ms = time_ms([&] { k_red<unsigned long long><<<grid, threads>>>(b64, nb, iters); });
      printf("RED.64 random %8u bins             %.3f ms  %.1f G/s\n", nb, ms, total / ms / 1e6);
      CK(cudaFree(b32));
      CK(cudaFree(b64));
    }
  }

  {  // match
    const int iters = 4096, threads = 512, grid = sms * 4;
    const double total = (double)grid * threads * iters;
    float b = time_ms([&] { k_match_none<<<grid, threads>>>(iters, sink); });
    float h = time_ms([&] { k_match_hw<<<grid, threads>>>(iters, sink); });
    float s = time_ms([&] { k_match_ballot<<<grid, threads>>>(iters, sink); });
    printf("match: baseline %.3f ms | match.any %.3f ms (%.1f G keys/s) | 8-ballot loop %.3f ms (%.1f G keys/s)\n", b,
           h, total / (h - b) / 1e6, s, total / (s - b) / 1e6);
  }
  return 0;
}
