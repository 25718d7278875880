// Dense ranks of 64-bit scores (float64 probabilities, int64 counts): the road for inputs the packed
// 32-bit key cannot hold.  The reference sorts any dtype in place (functional/classification/
// auroc.py:121, torch.sort); all the ranking kernels here look at is the ORDER and the TIES of the
// scores, so a 64-bit score is replaced by its dense rank — #{distinct values below it} — and the
// rank goes through the 32-bit kernels as the bit pattern of a non-negative float.
//
//   split   : order-preserving 64-bit image of every score (sign handling, -0.0 folded into +0.0),
//             cut into a low and a high word; payload = the record's index
//   sort    : stable pair sort by the low word, gather the high words in that order, stable pair sort
//             by the high word (te_sort_u32's passes, twice): the payload is the permutation that
//             sorts the 64-bit images
//   rank    : a record opens a new tie group where its image differs from its predecessor's; chunked
//             count / one-CTA scan / write: rank_out[perm[i]] = #{group openings at or before i} - 1,
//             and, when asked for, values_out[rank] = the score that opened the group (the
//             thresholds of the precision-recall curve are the distinct scores themselves).
// Everything is index arithmetic on 4-byte words; bit-exact by construction.
#pragma once
#include "common.cuh"

namespace te {

constexpr int kDrBlock = 256;
constexpr int kDrIpt = 8;
constexpr int kDrChunk = kDrBlock * kDrIpt;  // records per CTA of the count / write kernels

__device__ __forceinline__ unsigned long long dense_image_f64(unsigned long long b) {
  if (b == 0x8000000000000000ull) b = 0ull;  // -0.0 ties with +0.0
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ unsigned long long dense_image_i64(unsigned long long b) {
  return b ^ 0x8000000000000000ull;
}

// IS_FLOAT: the 8-byte records are IEEE doubles, otherwise two's-complement integers
template <bool IS_FLOAT>
__global__ void __launch_bounds__(kDrBlock)
    dense_split_kernel(const unsigned long long* __restrict__ x, uint32_t n, uint32_t* __restrict__ lo,
                       uint32_t* __restrict__ hi, uint32_t* __restrict__ idx) {
  for (uint32_t i = blockIdx.x * kDrBlock + threadIdx.x; i < n; i += gridDim.x * kDrBlock) {
    const unsigned long long b = x[i];
    const unsigned long long u = IS_FLOAT ? dense_image_f64(b) : dense_image_i64(b);
    lo[i] = (uint32_t)u;
    hi[i] = (uint32_t)(u >> 32);
    idx[i] = i;
  }
}

__global__ void __launch_bounds__(kDrBlock)
    dense_gather_kernel(const uint32_t* __restrict__ src, const uint32_t* __restrict__ perm, uint32_t n,
                        uint32_t* __restrict__ out) {
  for (uint32_t i = blockIdx.x * kDrBlock + threadIdx.x; i < n; i += gridDim.x * kDrBlock) out[i] = src[perm[i]];
}

// does sorted position i open a new tie group?  (position 0 does)
__device__ __forceinline__ uint32_t dense_opens(const uint32_t* __restrict__ hi_s, const uint32_t* __restrict__ lo_s,
                                                uint32_t i) {
  return (i == 0u || hi_s[i] != hi_s[i - 1] || lo_s[i] != lo_s[i - 1]) ? 1u : 0u;
}

// group openings per chunk of kDrChunk sorted positions
__global__ void __launch_bounds__(kDrBlock)
    dense_count_kernel(const uint32_t* __restrict__ hi_s, const uint32_t* __restrict__ lo_s, uint32_t n,
                       uint32_t* __restrict__ chunk_count) {
  __shared__ uint32_t s_w[kDrBlock / kWarp];
  const uint32_t base = blockIdx.x * (uint32_t)kDrChunk + threadIdx.x * (uint32_t)kDrIpt;
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < kDrIpt; ++k)
    if (base + k < n) c += dense_opens(hi_s, lo_s, base + k);
  c = warp_sum(c);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int w = 0; w < kDrBlock / kWarp; ++w) t += s_w[w];
    chunk_count[blockIdx.x] = t;
  }
}

// one CTA: chunk_count -> exclusive prefix, in place; total[0] = number of distinct values
__global__ void __launch_bounds__(1024)
    dense_scan_kernel(uint32_t* chunk_count, uint32_t nchunks, uint32_t* total) {
  __shared__ uint32_t s_w[32];
  __shared__ uint32_t s_carry;
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0u;
  __syncthreads();
  for (uint32_t b0 = 0; b0 < nchunks; b0 += 1024) {
    const uint32_t i = b0 + tid;
    const uint32_t v = i < nchunks ? chunk_count[i] : 0u;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(kFull, incl, o);
      if (lane >= (unsigned)o) incl += u;
    }
    if (lane == 31) s_w[warp] = incl;
    __syncthreads();
    uint32_t wb = 0;
    for (unsigned w = 0; w < warp; ++w) wb += s_w[w];
    const uint32_t carry = s_carry;
    if (i < nchunks) chunk_count[i] = carry + wb + incl - v;
    __syncthreads();
    if (tid == 1023) s_carry = carry + wb + incl;
    __syncthreads();
  }
  if (tid == 0 && total) total[0] = s_carry;
}

// rank_out[perm[i]] = (group openings at or before i) - 1; values_out[rank] = the opening score
__global__ void __launch_bounds__(kDrBlock)
    dense_write_kernel(const uint32_t* __restrict__ hi_s, const uint32_t* __restrict__ lo_s,
                       const uint32_t* __restrict__ perm, uint32_t n, const uint32_t* __restrict__ chunk_excl,
                       uint32_t* __restrict__ rank_out, const unsigned long long* __restrict__ x,
                       unsigned long long* __restrict__ values_out) {
  __shared__ uint32_t s_w[kDrBlock / kWarp];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t base = blockIdx.x * (uint32_t)kDrChunk + tid * (uint32_t)kDrIpt;
  uint32_t f[kDrIpt];
  uint32_t c = 0;
#pragma unroll
  for (int k = 0; k < kDrIpt; ++k) {
    f[k] = base + k < n ? dense_opens(hi_s, lo_s, base + k) : 0u;
    c += f[k];
  }
  uint32_t incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(kFull, incl, o);
    if (lane >= (unsigned)o) incl += u;
  }
  if (lane == 31) s_w[warp] = incl;
  __syncthreads();
  uint32_t run = chunk_excl[blockIdx.x] + incl - c;  // openings before this thread's first record
  for (unsigned w = 0; w < warp; ++w) run += s_w[w];
#pragma unroll
  for (int k = 0; k < kDrIpt; ++k) {
    if (base + k < n) {
      run += f[k];
      const uint32_t src = perm[base + k];
      rank_out[src] = run - 1u;
      if (values_out && f[k]) values_out[run - 1u] = x[src];
    }
  }
}

}  // namespace te
