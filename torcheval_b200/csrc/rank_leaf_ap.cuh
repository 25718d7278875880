// Average precision (step integral, precision_recall_curve.py:208-231 + tensor_utils.py:12-16) on
// the counting stage of rank_leaf.cuh.
//
// AP = (1/P) * sum over distinct scores v of  pos_v * TP(v) / CNT(v),  TP(v) / CNT(v) = positives /
// records with a score >= v.  Unlike 2U this is not linear in what lies below a segment (a ratio
// of two prefixes), so the stage runs twice over the grouped keys:
//   phase A  = the AUROC counting stage as it is: {P, N} per segment (its 2U is simply not used);
//   prefix   = one CTA: positives below every segment; the records below it are seg_start[s];
//   phase B  = the same pieces, the same histograms, walked with
//                TP(v) = P - (positives below the segment) - (positives below v inside it)
//                CNT(v) = n - seg_start[s] - (records below v inside it),
//              one float64 partial per segment; a last CTA adds them in a fixed order.
// 4 more bytes per score than AUROC (the keys are streamed twice).
#pragma once
#include "rank_leaf.cuh"

namespace te {

struct ApTotals {
  unsigned long long P, N;
};

// positives below every segment (exclusive prefix of SegStat.P) and the totals
__global__ void __launch_bounds__(1024)
    leaf_ap_prefix_kernel(const SegStat* __restrict__ seg_stats, uint32_t* __restrict__ pos_below, ApTotals* totals) {
  __shared__ uint32_t wp[32], wn[32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int ROWS = kLeafSegs / 1024;
  uint32_t p = 0, nn = 0;
#pragma unroll 4
  for (int j = 0; j < ROWS; ++j) {
    const SegStat s = seg_stats[(warp * ROWS + j) * 32 + lane];
    p += (uint32_t)s.P, nn += (uint32_t)s.N;
  }
  p = warp_sum(p), nn = warp_sum(nn);
  if (lane == 0) wp[warp] = p, wn[warp] = nn;
  __syncthreads();
  uint32_t at = 0;
  unsigned long long tp = 0, tn = 0;
  for (int w = 0; w < 32; ++w) {
    if (w < (int)warp) at += wp[w];
    tp += wp[w], tn += wn[w];
  }
  for (int j = 0; j < ROWS; ++j) {
    const uint32_t s = (warp * ROWS + j) * 32 + lane;
    const uint32_t c = (uint32_t)seg_stats[s].P;
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(kFull, incl, o);
      if (lane >= (unsigned)o) incl += u;
    }
    pos_below[s] = at + incl - c;
    at += __shfl_sync(kFull, incl, 31);
  }
  if (tid == 0) totals->P = tp, totals->N = tn;
}

// Walk of one segment's histogram (layout of leaf_walk) for the AP partial of the segment.
struct ApWarpTot {
  uint32_t P, C;
};
__device__ __forceinline__ double leaf_walk_ap(const uint4 (&v)[kLeafVecPerThread], ApWarpTot* s_tot, double* s_part,
                                               double p_above_seg /* P - positives below the segment */,
                                               double c_above_seg /* n - records below the segment */) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // warp totals first: a value's TP / CNT need what every earlier warp of the CTA holds
  uint32_t P = 0, C = 0;
#pragma unroll
  for (int j = 0; j < kLeafVecPerThread; ++j) P += v[j].y + v[j].w, C += v[j].x + v[j].y + v[j].z + v[j].w;
  P = warp_sum(P), C = warp_sum(C);
  if (lane == 0) s_tot[warp] = ApWarpTot{P, C};
  __syncthreads();
  uint32_t pb = 0, cb = 0;  // positives / records below this warp's first value, inside the segment
  for (int w = 0; w < (int)warp; ++w) pb += s_tot[w].P, cb += s_tot[w].C;
  double acc = 0.0;
#pragma unroll
  for (int j = 0; j < kLeafVecPerThread; ++j) {
    const uint32_t pj = v[j].y + v[j].w, cj = v[j].x + v[j].y + v[j].z + v[j].w;
    uint32_t ip = pj, ic = cj;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t up = __shfl_up_sync(kFull, ip, o), uc = __shfl_up_sync(kFull, ic, o);
      if (lane >= (unsigned)o) ip += up, ic += uc;
    }
    const uint32_t p0 = pb + ip - pj, c0 = cb + ic - cj;  // below my first value
    if (v[j].y) acc += (double)v[j].y * ((p_above_seg - (double)p0) / (c_above_seg - (double)c0));
    if (v[j].w) {
      const uint32_t p1 = p0 + v[j].y, c1 = c0 + v[j].x + v[j].y;
      acc += (double)v[j].w * ((p_above_seg - (double)p1) / (c_above_seg - (double)c1));
    }
    pb += __shfl_sync(kFull, ip, 31);
    cb += __shfl_sync(kFull, ic, 31);
  }
  acc = warp_sum(acc);  // butterfly: the same bits in every lane, every run
  if (lane == 0) s_part[warp] = acc;
  __syncthreads();
  double tot = 0.0;
  if (threadIdx.x == 0)
    for (int w = 0; w < kLeafWarps; ++w) tot += s_part[w];
  return tot;  // valid in thread 0
}

// Phase B over the pieces of leaf_plan_kernel's work list (the ticket must have been reset).
__global__ void __launch_bounds__(kLeafBlock, 1)
    leaf_count_ap_kernel(const uint32_t* __restrict__ keys, const LeafItem* __restrict__ items, LeafHeader* hdr,
                         const uint32_t* __restrict__ seg_start, const uint32_t* __restrict__ pos_below,
                         const ApTotals* __restrict__ totals, double* __restrict__ ap_part,
                         uint32_t* __restrict__ slabs) {
  extern __shared__ __align__(16) uint32_t s_hist[];
  __shared__ uint32_t s_ticket;
  __shared__ LeafItem s_item;
  __shared__ ApWarpTot s_tot[kLeafWarps];
  __shared__ double s_part[kLeafWarps];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t n_items = hdr->n_items;
  const double Ptot = (double)totals->P, ntot = (double)(totals->P + totals->N);
  uint4* s_hist4 = reinterpret_cast<uint4*>(s_hist);
  if (tid == 0) {
    s_ticket = atomicAdd(&hdr->ticket, 1u);
    if (s_ticket < n_items) s_item = items[n_items - 1u - s_ticket];
  }
  for (;;) {
#pragma unroll
    for (int j = 0; j < kLeafVecPerThread; ++j) s_hist4[j * kLeafBlock + tid] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const uint32_t t = s_ticket;
    if (t >= n_items) break;
    const LeafItem it = s_item;
    uint32_t next_ticket = 0;
    if (tid == 0) next_ticket = atomicAdd(&hdr->ticket, 1u);
    const uint32_t cnt = it.end - it.begin;
    const uint32_t* __restrict__ k = keys + it.begin;
    const uint32_t head = min(cnt, (4u - (it.begin & 3u)) & 3u);
    if (tid < head) atomicAdd(&s_hist[k[tid] & kLeafMask], 1u);
    const uint32_t nvec = (cnt - head) >> 2;
    const int4* __restrict__ kv = reinterpret_cast<const int4*>(k + head);
    constexpr int U = 4;
    for (uint32_t i = tid; i < nvec; i += U * kLeafBlock) {
      int4 a[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * kLeafBlock < nvec) a[u] = ld_stream_v4(kv + i + u * kLeafBlock);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (i + u * kLeafBlock < nvec) {
          atomicAdd(&s_hist[(uint32_t)a[u].x & kLeafMask], 1u);
          atomicAdd(&s_hist[(uint32_t)a[u].y & kLeafMask], 1u);
          atomicAdd(&s_hist[(uint32_t)a[u].z & kLeafMask], 1u);
          atomicAdd(&s_hist[(uint32_t)a[u].w & kLeafMask], 1u);
        }
      }
    }
    const uint32_t done = head + 4u * nvec;
    if (tid < cnt - done) atomicAdd(&s_hist[k[done + tid] & kLeafMask], 1u);
    LeafItem next_item = it;
    if (tid == 0 && next_ticket < n_items) next_item = items[n_items - 1u - next_ticket];
    __syncthreads();
    if (tid == 0) s_ticket = next_ticket, s_item = next_item;
    if (it.slab >= 0) {
      uint4* dst = reinterpret_cast<uint4*>(slabs + (size_t)it.slab * kLeafBins);
#pragma unroll
      for (int j = 0; j < kLeafVecPerThread; ++j) dst[j * kLeafBlock + tid] = s_hist4[j * kLeafBlock + tid];
    } else {
      uint4 v[kLeafVecPerThread];
#pragma unroll
      for (int j = 0; j < kLeafVecPerThread; ++j) v[j] = s_hist4[(warp * kLeafVecPerThread + j) * 32 + lane];
      const double part = leaf_walk_ap(v, s_tot, s_part, Ptot - (double)pos_below[it.seg], ntot - (double)seg_start[it.seg]);
      if (tid == 0) ap_part[it.seg] = part;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kLeafBlock)
    leaf_merge_ap_kernel(const LeafMulti* __restrict__ multi, const LeafHeader* __restrict__ hdr,
                         const uint32_t* __restrict__ slabs, const uint32_t* __restrict__ seg_start,
                         const uint32_t* __restrict__ pos_below, const ApTotals* __restrict__ totals,
                         double* __restrict__ ap_part) {
  __shared__ ApWarpTot s_tot[kLeafWarps];
  __shared__ double s_part[kLeafWarps];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n_multi = hdr->n_multi;
  const double Ptot = (double)totals->P, ntot = (double)(totals->P + totals->N);
  for (uint32_t m = blockIdx.x; m < n_multi; m += gridDim.x) {
    const LeafMulti mm = multi[m];
    uint4 v[kLeafVecPerThread];
#pragma unroll
    for (int j = 0; j < kLeafVecPerThread; ++j) v[j] = make_uint4(0, 0, 0, 0);
    for (uint32_t q = 0; q < mm.npieces; ++q) {
      const uint4* src = reinterpret_cast<const uint4*>(slabs + (size_t)(mm.first_slab + q) * kLeafBins);
#pragma unroll
      for (int j = 0; j < kLeafVecPerThread; ++j) {
        const uint4 a = src[(warp * kLeafVecPerThread + j) * 32 + lane];
        v[j].x += a.x, v[j].y += a.y, v[j].z += a.z, v[j].w += a.w;
      }
    }
    const double part = leaf_walk_ap(v, s_tot, s_part, Ptot - (double)pos_below[mm.seg], ntot - (double)seg_start[mm.seg]);
    if (threadIdx.x == 0) ap_part[mm.seg] = part;
    __syncthreads();
  }
}

// Segments of at most kLeafTiny keys: one warp, four keys per lane, pairs counted directly.
// For an own positive key ki: positives / records with a smaller VALUE inside the segment.
__global__ void __launch_bounds__(256)
    leaf_tiny_ap_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ seg_start,
                        const uint32_t* __restrict__ pos_below, const ApTotals* __restrict__ totals,
                        double* __restrict__ ap_part) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (s >= kLeafSegs) return;
  const uint32_t b = seg_start[s], e = seg_start[s + 1];
  if (e <= b || e - b > kLeafTiny) return;
  const uint32_t cnt = e - b;
  constexpr int SLOTS = kLeafTiny / 32;
  uint32_t mine[SLOTS], src[SLOTS], pb[SLOTS], cb[SLOTS];
#pragma unroll
  for (int m = 0; m < SLOTS; ++m) {
    const uint32_t i = m * 32 + lane;
    const uint32_t k = i < cnt ? keys[b + i] : 0xffffffffu;  // (no key of this path reaches 2^31)
    src[m] = k;
    mine[m] = (i < cnt && (k & 1u)) ? (k >> 1) : 0u;  // value of an own positive; 0 compares below nothing
    pb[m] = cb[m] = 0u;
  }
#pragma unroll
  for (int m = 0; m < SLOTS; ++m) {
    if ((uint32_t)m * 32 < cnt) {  // uniform across the warp
      for (int j = 0; j < 32; ++j) {
        const uint32_t kj = __shfl_sync(kFull, src[m], j);
        const uint32_t vj = kj >> 1, pj = kj & 1u;  // padding: value 0x7fffffff, never below anything
#pragma unroll
        for (int q = 0; q < SLOTS; ++q) {
          const uint32_t lt = vj < mine[q] ? 1u : 0u;
          cb[q] += lt;
          pb[q] += lt & pj;
        }
      }
    }
  }
  const double p_above = (double)totals->P - (double)pos_below[s], c_above = (double)(totals->P + totals->N) - (double)b;
  double acc = 0.0;
#pragma unroll
  for (int m = 0; m < SLOTS; ++m) {
    const uint32_t i = m * 32 + lane;
    if (i < cnt && (src[m] & 1u)) acc += (p_above - (double)pb[m]) / (c_above - (double)cb[m]);
  }
  acc = warp_sum(acc);
  if (lane == 0) ap_part[s] = acc;
}

// Fixed-order sum of the 65 536 partials and the result block of te_binary_auprc.
__global__ void __launch_bounds__(1024)
    leaf_ap_reduce_kernel(const double* __restrict__ ap_part, const ApTotals* __restrict__ totals,
                          const uint32_t* __restrict__ flags_in, double* __restrict__ out) {
  __shared__ double s_d[32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int ROWS = kLeafSegs / 1024;
  double a = 0.0;
  for (int j = 0; j < ROWS; ++j) a += ap_part[(warp * ROWS + j) * 32 + lane];
  a = warp_sum(a);
  if (lane == 0) s_d[warp] = a;
  __syncthreads();
  if (tid == 0) {
    double td = 0.0;
    for (int w = 0; w < 32; ++w) td += s_d[w];
    const double P = (double)totals->P, N = (double)totals->N;
    out[0] = P == 0.0 ? -0.0 : td / P;  // no positives: recall is NaN -> 1.0, the last step contributes -0.0
    out[1] = P;
    out[2] = N;
    out[3] = td;
    out[4] = (double)(flags_in ? *flags_in : 0u);
    out[5] = out[6] = 0.0;
    out[7] = P + N;
  }
}

}  // namespace te
