// Shared device/host helpers for the torcheval_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/te_b200.h"

#define TE_CUDA_TRY(expr)                         \
  do {                                            \
    cudaError_t _e = (expr);                      \
    if (_e != cudaSuccess) return (int)_e;        \
  } while (0)

#define TE_LAUNCH_CHECK()                         \
  do {                                            \
    cudaError_t _e = cudaGetLastError();          \
    if (_e != cudaSuccess) return (int)_e;        \
  } while (0)

namespace te {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// Streaming 128-bit load: data is touched exactly once, keep it out of L1.
__device__ __forceinline__ int4 ld_stream_v4(const void* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t ld_stream_u32(const void* p) {
  uint32_t r;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream_v4(void* p, int4 v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// Flag/value words of the decoupled look-back chains are exchanged through L2.
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
  uint32_t r;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long r;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long* p) {
  unsigned long long r;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_release_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ unsigned lane_id() {
  unsigned r;
  asm("mov.u32 %0, %%laneid;" : "=r"(r));
  return r;
}
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned r;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(r));
  return r;
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}

// Order-preserving map float -> uint32 (ascending). -0.0 is folded onto +0.0 first so that
// the two compare equal, as `threshold.diff() != 0` does in the reference
// (functional/classification/auroc.py:122).
__device__ __forceinline__ uint32_t orderable_f32(float f) {
  uint32_t b = __float_as_uint(f);
  if (b == 0x80000000u) b = 0u;
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

// Inclusive block-wide sum scan (warp shuffles + one smem hop). `ws` needs BLOCK/32 slots.
// Returns the inclusive prefix of `v`; `total` is the block sum. Ends with a barrier, so `ws`
// can be reused right away.
template <typename T, int BLOCK>
__device__ __forceinline__ T block_scan_incl(T v, T* ws, T& total) {
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T u = __shfl_up_sync(kFull, v, o);
    if (lane >= (unsigned)o) v += u;
  }
  if (lane == 31) ws[w] = v;
  __syncthreads();
  T base = 0, tot = 0;
#pragma unroll
  for (int i = 0; i < BLOCK / 32; ++i) {
    const T x = ws[i];
    if (i < (int)w) base += x;
    tot += x;
  }
  total = tot;
  __syncthreads();
  return v + base;
}

inline int sm_count_of_current_device(int* out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  e = cudaDeviceGetAttribute(out, cudaDevAttrMultiProcessorCount, dev);
  return (int)e;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace te
