// Exact binary AUROC without a full sort: two partition passes + one counting stage.
//
// AUROC only needs, for every distinct score, how many negatives and positives carry it and how
// many negatives lie below it — not the position of every record.  The packed key of a score
// below 2.0 has 31 significant bits, (float bits << 1) | label.  So the keys are ordered by their
// top 16 bits only (two 8-bit LSD passes of rank_sort.cuh, on bits 15..22 and 23..30), which
// leaves them grouped into 65 536 SEGMENTS of consecutive key values; inside a segment only 15
// bits vary (14 value bits and the label), and 2^15 32-bit counters are 128 KB: one CTA's shared
// memory.  The last stage therefore streams every segment once and counts its keys into a
// shared-memory histogram indexed by those 15 bits (one shared atomic per key, no ranking, no
// scatter, no write-back), then walks the histogram in value order:
//     2U_seg = sum over values v of pos_v * (2 * #neg below v inside the segment + neg_v)
// and leaves {2U_seg, P_seg, N_seg} per segment; AUROC is linear in what a segment does not know
// (the negatives below it), so one CTA finishes:  2U = sum_seg 2U_seg + 2 P_seg * N_below(seg).
//
// Bytes per score: 16 (ingest) + 2 * 8 (passes) + 4 (counting stage) = 36, against 52 for the
// four-pass sort + sweep, and two shared-memory-bound passes instead of four.
//
// Work decomposition of the counting stage: a segment of more than `piece` keys is cut into
// pieces that different CTAs count; each such piece stores its histogram to a slab in global
// memory (plain 16-byte stores, 128 KB per piece) and a second kernel adds up the slabs of a
// segment and walks the sum.  Segments that fit one piece never leave shared memory.  Pieces are
// handed out through a ticket, largest key range first (the heavy segments of [0, 1) scores are
// the top binades).
#pragma once
#include "common.cuh"
#include "rank_sort.cuh"  // batch_ptr

namespace te {

constexpr int kLeafBits = 15;
constexpr uint32_t kLeafBins = 1u << kLeafBits;     // counters per segment: (14 value bits, label)
constexpr uint32_t kLeafMask = kLeafBins - 1u;
constexpr uint32_t kLeafSegs = 1u << 16;            // key >> 15, keys < 2^31
constexpr int kLeafBlock = 1024;
constexpr size_t kLeafSmem = (size_t)kLeafBins * sizeof(uint32_t);  // 128 KB
constexpr int kLeafWarps = kLeafBlock / kWarp;
constexpr int kLeafVecPerThread = (int)(kLeafBins / 4 / kLeafBlock);  // uint4 per thread: 8
constexpr uint32_t kLeafMinPiece = 1u << 17, kLeafMaxPiece = 1u << 20;
constexpr uint32_t kLeafTiny = 128;  // segments of at most this many keys never see a histogram (leaf_tiny_kernel)

struct LeafItem {   // one piece of one segment
  uint32_t seg, begin, end;
  int32_t slab;     // >= 0: the segment has several pieces, this one stores its histogram there
};
struct LeafMulti {  // a segment with several pieces
  uint32_t seg, first_slab, npieces, pad;
};
struct alignas(16) SegStat {  // 16 bytes: the 65 536 of them are read by ONE CTA at the end, through one SM's L2 port
  unsigned long long A;        // 2U inside the segment
  uint32_t P, N;               // positives, negatives (a segment holds fewer than 2^31 keys)
};
struct LeafHeader {
  uint32_t ticket, n_items, n_multi, pad;
};

// seg_start[s] = first index whose key has (key >> 15) >= s, s = 0 .. 65536 (keys ordered by
// their top bits).  65 537 independent binary searches; the top levels share cache lines.
__global__ void __launch_bounds__(256)
    leaf_seg_bounds_kernel(const uint32_t* __restrict__ keys, uint32_t n, uint32_t* __restrict__ seg_start,
                           const uint32_t* __restrict__ n_dev = nullptr) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s > kLeafSegs) return;
  if (n_dev) n = *n_dev;  // (the record count lives in device memory)
  uint32_t lo = 0, hi = n;
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    if ((keys[mid] >> kLeafBits) < s)
      lo = mid + 1;
    else
      hi = mid;
  }
  seg_start[s] = lo;
}

// One CTA: the work list.  Warp w owns segments [2048 w, 2048 w + 2048) and reads them 32 at a
// time (coalesced); a first sweep gives every warp its totals, a second one hands every segment
// its first item, its first slab and its slot in the multi-piece list (warp scans + a carry).
__device__ __forceinline__ void leaf_plan_body(const uint32_t* __restrict__ seg_start, uint32_t piece, uint32_t max_items,
                                               uint32_t max_slabs, LeafItem* __restrict__ items,
                                               LeafMulti* __restrict__ multi, LeafHeader* hdr) {
  __shared__ uint32_t ws[3][32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int ROWS = kLeafSegs / 1024;
  auto pieces_of = [&](uint32_t s, uint32_t& b, uint32_t& e) -> uint32_t {
    b = seg_start[s], e = seg_start[s + 1];
    const uint32_t cnt = e > b ? e - b : 0u;
    return cnt <= kLeafTiny ? 0u : (cnt + piece - 1) / piece;  // (tiny segments: leaf_tiny_kernel)
  };
  uint32_t n_it = 0, n_sl = 0, n_mu = 0;
#pragma unroll 16
  for (int j = 0; j < ROWS; ++j) {
    uint32_t b, e;
    const uint32_t np = pieces_of((warp * ROWS + j) * 32 + lane, b, e);
    n_it += np;
    if (np > 1) n_sl += np, n_mu += 1;
  }
  n_it = warp_sum(n_it), n_sl = warp_sum(n_sl), n_mu = warp_sum(n_mu);
  if (lane == 0) ws[0][warp] = n_it, ws[1][warp] = n_sl, ws[2][warp] = n_mu;
  __syncthreads();
  uint32_t at_it = 0, at_sl = 0, at_mu = 0, tot_it = 0, tot_sl = 0, tot_mu = 0;
  for (int w = 0; w < 32; ++w) {
    if (w < (int)warp) at_it += ws[0][w], at_sl += ws[1][w], at_mu += ws[2][w];
    tot_it += ws[0][w], tot_sl += ws[1][w], tot_mu += ws[2][w];
  }
  // (the bounds hold by construction for keys ordered by their top bits; if the caller's keys are
  // not — scores >= 2.0, which the ingest flags — nothing may be written out of bounds)
  const bool fits = tot_it <= max_items && tot_sl <= max_slabs;
  if (tid == 0) {
    hdr->n_items = fits ? tot_it : 0u;
    hdr->n_multi = fits ? tot_mu : 0u;
  }
  if (!fits) return;
  for (int j0 = 0; j0 < ROWS; j0 += 8) {
   uint32_t b8[8], e8[8], np8[8];  // eight rows' loads in flight at once (one CTA: latency is all there is)
#pragma unroll
   for (int q = 0; q < 8; ++q) np8[q] = pieces_of((warp * ROWS + j0 + q) * 32 + lane, b8[q], e8[q]);
#pragma unroll
   for (int q = 0; q < 8; ++q) {
    const uint32_t s = (warp * ROWS + j0 + q) * 32 + lane;
    const uint32_t b = b8[q], e = e8[q], np = np8[q];
    const uint32_t sl = np > 1 ? np : 0u, mu = np > 1 ? 1u : 0u;
    uint32_t i_it = np, i_sl = sl, i_mu = mu;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u0 = __shfl_up_sync(kFull, i_it, o), u1 = __shfl_up_sync(kFull, i_sl, o),
                     u2 = __shfl_up_sync(kFull, i_mu, o);
      if (lane >= (unsigned)o) i_it += u0, i_sl += u1, i_mu += u2;
    }
    uint32_t my_it = at_it + i_it - np;
    const uint32_t my_sl = at_sl + i_sl - sl;
    if (np > 1) {
      LeafMulti m;
      m.seg = s, m.first_slab = my_sl, m.npieces = np, m.pad = 0;
      multi[at_mu + i_mu - mu] = m;
    }
    for (uint32_t q = 0; q < np; ++q) {
      LeafItem it;
      it.seg = s;
      it.begin = b + q * piece;
      it.end = min(e, it.begin + piece);
      it.slab = np > 1 ? (int32_t)(my_sl + q) : -1;
      items[my_it++] = it;
    }
    at_it += __shfl_sync(kFull, i_it, 31);
    at_sl += __shfl_sync(kFull, i_sl, 31);
    at_mu += __shfl_sync(kFull, i_mu, 31);
   }
  }
}

__global__ void __launch_bounds__(1024)
    leaf_plan_kernel(const uint32_t* __restrict__ seg_start, uint32_t piece, uint32_t max_items,
                     uint32_t max_slabs, LeafItem* __restrict__ items, LeafMulti* __restrict__ multi,
                     LeafHeader* hdr, size_t bstride = 0) {
  leaf_plan_body(batch_ptr(seg_start, bstride), piece, max_items, max_slabs, batch_ptr(items, bstride),
                 batch_ptr(multi, bstride), batch_ptr(hdr, bstride));
}

// Segments of at most kLeafTiny keys — for [0, 1) scores the thousands of segments of the low
// binades — are not worth clearing and walking 32 768 counters (~2 us of a whole CTA each): one
// WARP takes such a segment, four keys per lane, and counts pairs directly.  With packed keys
// (value << 1 | label) a positive ki and a negative kj contribute
//     2 [v_j < v_i] + [v_j == v_i]  =  [kj < ki] + [kj + 1 < ki],
// so sources that are not negatives are replaced by 0xfffffffe (no key of this path reaches 2^31)
// and own keys that are not positives by 0, and the inner loop is two compares per pair.
__device__ __forceinline__ void leaf_tiny_body(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ seg_start,
                                               SegStat* __restrict__ seg_stats, uint32_t s) {
  const unsigned lane = threadIdx.x & 31;
  if (s >= kLeafSegs) return;
  const uint32_t b = seg_start[s], e = seg_start[s + 1];
  if (e <= b || e - b > kLeafTiny) return;
  const uint32_t cnt = e - b;
  constexpr int SLOTS = kLeafTiny / 32;
  uint32_t mine[SLOTS], src[SLOTS];
  uint32_t npos = 0;
#pragma unroll
  for (int m = 0; m < SLOTS; ++m) {
    const uint32_t i = m * 32 + lane;
    const uint32_t k = i < cnt ? keys[b + i] : 0u;
    const bool pos = i < cnt && (k & 1u);
    npos += pos ? 1u : 0u;
    mine[m] = pos ? k : 0u;
    src[m] = (i < cnt && !(k & 1u)) ? k : 0xfffffffeu;
  }
  uint32_t acc = 0;  // <= 2 * 128 * 128
#pragma unroll
  for (int m = 0; m < SLOTS; ++m) {
    if ((uint32_t)m * 32 < cnt) {  // uniform across the warp
      for (int j = 0; j < 32; ++j) {
        const uint32_t kj = __shfl_sync(kFull, src[m], j);
#pragma unroll
        for (int q = 0; q < SLOTS; ++q) acc += (kj < mine[q] ? 1u : 0u) + (kj + 1u < mine[q] ? 1u : 0u);
      }
    }
  }
  acc = warp_sum(acc);
  npos = warp_sum(npos);
  if (lane == 0) {
    SegStat r;
    r.A = acc, r.P = npos, r.N = cnt - npos;
    seg_stats[s] = r;
  }
}

__global__ void __launch_bounds__(256)
    leaf_tiny_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ seg_start,
                     SegStat* __restrict__ seg_stats, size_t bstride = 0) {
  leaf_tiny_body(batch_ptr(keys, bstride), batch_ptr(seg_start, bstride), batch_ptr(seg_stats, bstride),
                 blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5));
}

// Both of the above in one launch: CTA 0 writes the work list (one CTA, ~50 us of latency) while the
// other CTAs take the tiny segments, a warp each — the two are independent, and on their own they
// were two launches in a row on the critical path.
__global__ void __launch_bounds__(1024)
    leaf_tiny_plan_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ seg_start,
                          SegStat* __restrict__ seg_stats, uint32_t piece, uint32_t max_items, uint32_t max_slabs,
                          LeafItem* __restrict__ items, LeafMulti* __restrict__ multi, LeafHeader* hdr, size_t bstride = 0) {
  seg_start = batch_ptr(seg_start, bstride);
  if (blockIdx.x == 0) {
    leaf_plan_body(seg_start, piece, max_items, max_slabs, batch_ptr(items, bstride), batch_ptr(multi, bstride),
                   batch_ptr(hdr, bstride));
    return;
  }
  leaf_tiny_body(batch_ptr(keys, bstride), seg_start, batch_ptr(seg_stats, bstride),
                 (blockIdx.x - 1u) * 32u + (threadIdx.x >> 5));
}

// Walk of one segment's histogram, held as kLeafVecPerThread uint4 per thread in the layout
// word = (warp * 8 + j) * 128 + lane * 4 + {0: neg_v, 1: pos_v, 2: neg_v+1, 3: pos_v+1}:
// value order is (warp, j, lane), so a warp scan per j plus a carry gives every lane the number of
// negatives below its two values; the 32 warp totals are combined by thread 0.
struct LeafWarpTot {
  unsigned long long A, P, N;
};
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void leaf_walk(const uint4 (&v)[kLeafVecPerThread], LeafWarpTot* s_tot,
                                          SegStat* __restrict__ out) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long A = 0;
  uint32_t carry = 0, P = 0;  // a segment holds fewer than 2^31 keys
#pragma unroll
  for (int j = 0; j < kLeafVecPerThread; ++j) {
    const uint32_t negs = v[j].x + v[j].z;
    uint32_t incl = negs;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(kFull, incl, o);
      if (lane >= (unsigned)o) incl += u;
    }
    const uint32_t below0 = carry + incl - negs;   // negatives below my first value (warp-relative)
    const uint32_t below1 = below0 + v[j].x;
    A += (unsigned long long)v[j].y * (2ull * below0 + v[j].x) + (unsigned long long)v[j].w * (2ull * below1 + v[j].z);
    P += v[j].y + v[j].w;
    carry += __shfl_sync(kFull, incl, 31);
  }
  A = warp_sum(A);
  const unsigned long long Pw = warp_sum((unsigned long long)P);
  if (lane == 0) {
    LeafWarpTot t;
    t.A = A, t.P = Pw, t.N = carry;
    s_tot[warp] = t;
  }
  __syncthreads();
  if (warp == 0) {  // lane w holds warp w's totals: negatives below it by a warp scan, the rest by warp sums
    static_assert(kLeafWarps == 32, "one lane per warp");
    const LeafWarpTot t = s_tot[lane];
    const uint32_t nw = (uint32_t)t.N;
    uint32_t incl = nw;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(kFull, incl, o);
      if (lane >= (unsigned)o) incl += u;
    }
    const unsigned long long a = warp_sum(t.A + 2ull * t.P * (unsigned long long)(incl - nw));
    const unsigned long long p = warp_sum(t.P);
    const uint32_t nn = __shfl_sync(kFull, incl, 31);
    if (lane == 0) {
      SegStat r;
      r.A = a, r.P = (uint32_t)p, r.N = nn;
      *out = r;
    }
  }
}

// Counting stage.  Persistent CTAs (one per SM, the histogram fills its shared memory) draw
// pieces from the work list.
__global__ void __launch_bounds__(kLeafBlock, 1)
    leaf_count_kernel(const uint32_t* __restrict__ keys, const LeafItem* __restrict__ items, LeafHeader* hdr,
                      SegStat* __restrict__ seg_stats, uint32_t* __restrict__ slabs, size_t bstride = 0) {
  keys = batch_ptr(keys, bstride), items = batch_ptr(items, bstride), hdr = batch_ptr(hdr, bstride);
  seg_stats = batch_ptr(seg_stats, bstride), slabs = batch_ptr(slabs, bstride);
  extern __shared__ __align__(16) uint32_t s_hist[];
  __shared__ uint32_t s_ticket;
  __shared__ LeafWarpTot s_tot[kLeafWarps];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ LeafItem s_item, s_next;
  __shared__ uint32_t s_next_ok;
  const uint32_t n_items = hdr->n_items;
  uint4* s_hist4 = reinterpret_cast<uint4*>(s_hist);
  // The ticket (a global atomic) and the item it names (a dependent load) are two round trips of
  // ~0.4 us each, and a light piece is counted in less than that.  Thread 0 therefore runs two rounds
  // ahead: in round r it draws the ticket of round r + 2 and loads the item of round r + 1, whose
  // ticket it drew in round r - 1 — neither wait ends up on the path of a round.
  uint32_t pend = 0;  // (thread 0) the ticket of the round after this one
  if (tid == 0) {
    s_ticket = atomicAdd(&hdr->ticket, 1u);
    pend = atomicAdd(&hdr->ticket, 1u);
    if (s_ticket < n_items) s_item = items[n_items - 1u - s_ticket];  // highest keys first: the heavy segments
  }
  // the histogram is cleared once here; every round leaves it cleared behind itself (the walk and the
  // slab copy write zeros where they have just read: one pass over the 128 KB per piece instead of two)
#pragma unroll
  for (int j = 0; j < kLeafVecPerThread; ++j) s_hist4[j * kLeafBlock + tid] = make_uint4(0, 0, 0, 0);
  for (;;) {
    __syncthreads();
    const uint32_t t = s_ticket;
    if (t >= n_items) break;
    const LeafItem it = s_item;
    uint32_t next_ticket = 0;
    LeafItem next_item = it;
    if (tid == 0) {
      next_ticket = pend;
      if (next_ticket < n_items) {
        pend = atomicAdd(&hdr->ticket, 1u);
        next_item = items[n_items - 1u - next_ticket];
      }
    }
    const uint32_t cnt = it.end - it.begin;
    const uint32_t* __restrict__ k = keys + it.begin;
    // scalar head up to the first 16-byte boundary, vector body, scalar tail
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
    if (tid == 0) {
      s_next = next_item;
      s_next_ok = next_ticket < n_items ? 1u : 0u;
    }
    __syncthreads();  // counts complete; everybody has read s_ticket / s_item of this round
    if (tid == 0) s_ticket = next_ticket, s_item = next_item;
    // the next piece's first lines on their way from DRAM to L2 while this one is walked (a light
    // piece is all latency: ~1 us of its ~4 is the first load)
    if (s_next_ok) {
      const uint32_t nb0 = s_next.begin, ne0 = s_next.end;
      if (tid < min((ne0 - nb0 + 31u) >> 5, (uint32_t)kLeafBlock)) prefetch_l2(keys + nb0 + 32u * tid);
    }
    if (it.slab >= 0) {
      uint4* dst = reinterpret_cast<uint4*>(slabs + (size_t)it.slab * kLeafBins);
#pragma unroll
      for (int j = 0; j < kLeafVecPerThread; ++j) {
        dst[j * kLeafBlock + tid] = s_hist4[j * kLeafBlock + tid];
        s_hist4[j * kLeafBlock + tid] = make_uint4(0, 0, 0, 0);
      }
    } else {
      uint4 v[kLeafVecPerThread];
#pragma unroll
      for (int j = 0; j < kLeafVecPerThread; ++j) {
        v[j] = s_hist4[(warp * kLeafVecPerThread + j) * 32 + lane];
        s_hist4[(warp * kLeafVecPerThread + j) * 32 + lane] = make_uint4(0, 0, 0, 0);
      }
      leaf_walk(v, s_tot, seg_stats + it.seg);
    }
    // (the sync at the top of the loop: the cleared histogram and the new s_ticket / s_item are visible)
  }
}

// Segments counted in several pieces: one CTA adds up the slabs of one segment and walks the sum.
__global__ void __launch_bounds__(kLeafBlock)
    leaf_merge_kernel(const LeafMulti* __restrict__ multi, const LeafHeader* __restrict__ hdr,
                      const uint32_t* __restrict__ slabs, SegStat* __restrict__ seg_stats, size_t bstride = 0) {
  multi = batch_ptr(multi, bstride), hdr = batch_ptr(hdr, bstride);
  slabs = batch_ptr(slabs, bstride), seg_stats = batch_ptr(seg_stats, bstride);
  __shared__ LeafWarpTot s_tot[kLeafWarps];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t n_multi = hdr->n_multi;
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
    leaf_walk(v, s_tot, seg_stats + mm.seg);
    __syncthreads();
  }
}

// One CTA: 2U = sum over segments (ascending) of A_seg + 2 P_seg * (negatives below the segment),
// and the result block of te_binary_auroc.  Warp w owns segments [2048 w, 2048 w + 2048), read 32
// at a time; all counts are below 2^31, only the 2U terms need 64 bits.  `flags_in` are the
// TE_RANK_FLAG_* bits collected on the way.
__global__ void __launch_bounds__(1024)
    leaf_combine_kernel(const SegStat* __restrict__ seg_stats, uint32_t n, const uint32_t* __restrict__ flags_in,
                        double* __restrict__ out, const uint32_t* __restrict__ n_dev = nullptr, size_t bstride = 0) {
  seg_stats = batch_ptr(seg_stats, bstride), flags_in = batch_ptr(flags_in, bstride);
  out += (size_t)blockIdx.y * 8;  // (batched: one 8-double row per problem)
  if (n_dev) n = *n_dev;
  __shared__ LeafWarpTot s_tot[32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int ROWS = kLeafSegs / 1024;
  unsigned long long a = 0;
  uint32_t p = 0, carry = 0;
  for (int j0 = 0; j0 < ROWS; j0 += 8) {
   SegStat s8[8];
#pragma unroll
   for (int q = 0; q < 8; ++q) s8[q] = seg_stats[(warp * ROWS + j0 + q) * 32 + lane];
#pragma unroll
   for (int q = 0; q < 8; ++q) {
    const SegStat s = s8[q];
    const uint32_t nn = (uint32_t)s.N;
    uint32_t incl = nn;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(kFull, incl, o);
      if (lane >= (unsigned)o) incl += u;
    }
    a += s.A + 2ull * s.P * (unsigned long long)(carry + incl - nn);
    p += (uint32_t)s.P;
    carry += __shfl_sync(kFull, incl, 31);
   }
  }
  a = warp_sum(a);
  const unsigned long long pw = warp_sum((unsigned long long)p);
  if (lane == 0) {
    LeafWarpTot t;
    t.A = a, t.P = pw, t.N = carry;
    s_tot[warp] = t;
  }
  __syncthreads();
  if (tid == 0) {
    unsigned long long tu = 0, tp = 0, tn = 0;
    for (int w = 0; w < 32; ++w) {
      const LeafWarpTot t = s_tot[w];
      tu += t.A + 2ull * t.P * tn;
      tp += t.P;
      tn += t.N;
    }
    const double P = (double)tp, N = (double)tn;
    out[0] = (P * N == 0.0) ? 0.5 : ((double)tu * 0.5) / (P * N);
    out[1] = P;
    out[2] = N;
    out[3] = (double)tu;
    out[4] = (double)(flags_in ? *flags_in : 0u);
    out[5] = (double)(tu >> 32);
    out[6] = (double)(tu & 0xffffffffull);
    out[7] = (double)n;
  }
}

}  // namespace te
