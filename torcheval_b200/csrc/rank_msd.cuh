// The two partition passes of the counting-stage AUROC path, most significant digit first, with no
// inter-CTA dependency in either of them.
//
// What the counting stage needs is the keys grouped by their top 16 bits (65 536 segments); the
// order inside a segment is irrelevant.  So neither pass has an order to preserve IF the second
// one works inside the buckets of the first: pass 1 partitions by key bits 23..30, pass 2
// partitions every bucket of pass 1 by bits 15..22.  Where a record goes is then a matter of
// counting, not of scanning over earlier tiles: the upsweep (which reads the raw input anyway)
// leaves the JOINT histogram of both digits — 65 536 counters, which is also the table of segment
// sizes the counting stage wants — and its exclusive prefix is the start of every (digit 2,
// digit 1) run.  A tile reserves room for each of its digit runs with ONE atomicAdd on a global
// cursor (256 per 24 576-record tile, their round trip overlapped with the shared-memory
// scatter).  No look-back, no flags, no stable ranking, no per-range histograms, any tile order:
// both passes are the kernel that took the first pass from 3.8 to 2.5 ms.
//
// The joint histogram is built in shared memory as 65 536 SIXTEEN-bit counters (128 KB), one
// shared atomic per record, and flushed into the global 32-bit table every 61 440 records of a CTA
// — before any counter can wrap.  A flush only touches non-zero counters, of which scores have few
// (uniform [0, 1): ~3 000 of 65 536 per window); a distribution that fills most of the 65 536
// prefixes in every 61 440-record window pays up to one L2 atomic per record there.
#pragma once
#include "rank_part.cuh"

namespace te {

constexpr int kUjBlock = 1024;
constexpr size_t kUjSmem = (size_t)kLeafSegs * sizeof(uint16_t);  // 128 KB
constexpr uint32_t kUjChunk = kUjBlock * 4;                       // records per CTA step
constexpr uint32_t kUjFlushEvery = 65535 / kUjChunk;              // 15 steps = 61 440 records

// Upsweep.  RAW: `score` holds packed keys (receive side of the sample sort): joint histogram and
// the wide-key flag only.  Else: (score, target) -> label bits (layout of rank_part.cuh) + joint
// histogram + TE_RANK_FLAG_* bits.  joint: uint32[65536], zeroed by the caller.
template <typename TargetT, bool RAW>
__global__ void __launch_bounds__(kUjBlock, 1)
    rank_upsweep_joint_kernel(const float* __restrict__ score, const TargetT* __restrict__ target, uint32_t n,
                              bool vec_ok, uint32_t* __restrict__ label_bits, uint32_t* joint, uint32_t* flags,
                              const SlabView view = SlabView{nullptr, 0, 0, 0}, size_t bstride = 0) {
  // bstride (RAW only): grid.y independent key arrays of n records each, one scratch slice apart
  if (RAW) score = batch_ptr(score, bstride);
  joint = batch_ptr(joint, bstride), flags = batch_ptr(flags, bstride);
  extern __shared__ __align__(16) uint32_t s_j[];  // [32768] words = 65536 16-bit counters
  __shared__ uint32_t s_best[2][kUjBlock / kWarp];
  __shared__ uint32_t s_hotseg[2], s_hotsum[2];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < (int)(kLeafSegs / 2); i += kUjBlock) s_j[i] = 0u;
  if (tid < 2) s_hotseg[tid] = 0xffffffffu, s_hotsum[tid] = 0u;
  __syncthreads();
  // Tied or quantised scores put most records of a window on one or two counters, and 32 lanes on one
  // shared address serialise (constant scores: 4.6 instead of 2.2 ms at 2^30).  Every flush therefore
  // also finds the two fullest counters of the window; if one holds >= 1/4 of it, records of those
  // two segments are from then on counted inside the warp — a ballot and a register — and added to
  // the global table once per flush.
  uint32_t hc0 = 0u, hc1 = 0u;  // this warp's hot records since the last flush (uniform over the lanes)
  uint32_t nflush = 0u;
  auto flush = [&]() {
    __syncthreads();
    const uint32_t h0 = s_hotseg[0], h1 = s_hotseg[1];
    if (lane == 0) {  // the warp's hot records: to the global table, and to the window totals of the two segments
      if (hc0) atomicAdd(joint + h0, hc0), atomicAdd(&s_hotsum[0], hc0);
      if (hc1) atomicAdd(joint + h1, hc1), atomicAdd(&s_hotsum[1], hc1);
    }
    hc0 = hc1 = 0u;
    // (the hot segments are re-chosen at the first and then every fourth flush: the choice costs
    // ~1.5 us of a whole CTA, 0.2 ms per 2^30 records if made every time)
    const bool pick = (nflush++ & 3u) == 0u;
    uint32_t best = 0u, second = 0u;  // (count << 16 | segment) of my fullest counters: a count fits 16 bits
    for (uint32_t w = tid; w < kLeafSegs / 2; w += kUjBlock) {
      const uint32_t v = s_j[w];
      if (v) {
        s_j[w] = 0u;
        const uint32_t lo = v & 0xffffu, hi = v >> 16;
        if (lo) atomicAdd(joint + 2 * w, lo);
        if (hi) atomicAdd(joint + 2 * w + 1, hi);
        if (!pick) continue;
        const uint32_t a = lo << 16 | (2 * w), c = hi << 16 | (2 * w + 1);
        const uint32_t big = max(a, c), small = min(a, c);
        if (big > best) {
          second = max(best, small);
          best = big;
        } else {
          second = max(second, big);
        }
      }
    }
    // top two of the warp, then of the CTA (all candidates are distinct segments)
    auto top2 = [&](uint32_t& b1, uint32_t& b2) {
      uint32_t wb = b1;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wb = max(wb, __shfl_xor_sync(kFull, wb, o));
      uint32_t ws = b1 == wb ? b2 : b1;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ws = max(ws, __shfl_xor_sync(kFull, ws, o));
      b1 = wb, b2 = ws;
    };
    if (!pick) {  // keep the current choice; only the window totals start over
      if (tid == 0) s_hotsum[0] = s_hotsum[1] = 0u;
      __syncthreads();
      return;
    }
    top2(best, second);
    if (lane == 0) s_best[0][warp] = best, s_best[1][warp] = second;
    __syncthreads();
    if (warp == 0) {
      uint32_t b1 = s_best[0][lane], b2 = s_best[1][lane];
      top2(b1, b2);
      if (lane == 0) {
        // hot segments never reach the counters: their window totals compete as they are
        unsigned long long c0 = (unsigned long long)(b1 >> 16) << 32 | (b1 & 0xffffu), c1 = (unsigned long long)(b2 >> 16) << 32 | (b2 & 0xffffu);
        auto offer = [&](unsigned long long c) {
          if (c > c0) {
            c1 = c0;
            c0 = c;
          } else if (c > c1) {
            c1 = c;
          }
        };
        if (h0 != 0xffffffffu) offer((unsigned long long)s_hotsum[0] << 32 | h0);
        if (h1 != 0xffffffffu) offer((unsigned long long)s_hotsum[1] << 32 | h1);
        const uint32_t need = kUjChunk * kUjFlushEvery / 4;
        const bool on = (uint32_t)(c0 >> 32) >= need;
        s_hotseg[0] = on ? (uint32_t)c0 : 0xffffffffu;
        s_hotseg[1] = on && (uint32_t)(c1 >> 32) >= need / 4 ? (uint32_t)c1 : 0xffffffffu;
        s_hotsum[0] = s_hotsum[1] = 0u;
      }
    }
    __syncthreads();
  };
  uint32_t bad_label = 0, neg_score = 0, wide = 0, steps = 0;
  // (slab view: chunks of the physical index space, a chunk never straddles slabs; the empty tails are
  // spread evenly over the CTAs by the strided assignment).  U chunks per round, all loads first: one
  // 16-byte load per thread in flight is far too little to keep HBM busy (0.31 -> 0.16 ms for 2^27 keys).
  constexpr int U = RAW ? 4 : 2;
  const uint32_t nchunks = view.counts ? view.nslabs * (view.cap / kUjChunk) : (n + kUjChunk - 1) / kUjChunk;
  // one chunk of 4 records per thread; FULL (the common case, uniform over the CTA): all 4096 records
  // exist and the 16-byte loads were issued, so there is not a single bounds check in the way
  auto count_chunk = [&](auto full_tag, uint64_t i64, uint32_t cnt, const int4& sq, const int4& t0, const int4& t1) {
    constexpr bool FULL = decltype(full_tag)::value;
    uint32_t b[4] = {0, 0, 0, 0}, lab[4] = {0, 0, 0, 0};
    if constexpr (FULL) {
      b[0] = (uint32_t)sq.x, b[1] = (uint32_t)sq.y, b[2] = (uint32_t)sq.z, b[3] = (uint32_t)sq.w;
      if constexpr (!RAW) {
        if constexpr (sizeof(TargetT) == 8) {
          TargetT tv[4];
          const unsigned long long u0 = ((unsigned long long)(uint32_t)t0.y << 32) | (uint32_t)t0.x;
          const unsigned long long u1 = ((unsigned long long)(uint32_t)t0.w << 32) | (uint32_t)t0.z;
          const unsigned long long u2 = ((unsigned long long)(uint32_t)t1.y << 32) | (uint32_t)t1.x;
          const unsigned long long u3 = ((unsigned long long)(uint32_t)t1.w << 32) | (uint32_t)t1.z;
          memcpy(&tv[0], &u0, 8), memcpy(&tv[1], &u1, 8), memcpy(&tv[2], &u2, 8), memcpy(&tv[3], &u3, 8);
#pragma unroll
          for (int j = 0; j < 4; ++j) lab[j] = LabelOf<TargetT>::get(tv[j], bad_label);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) lab[j] = LabelOf<TargetT>::get(target[i64 + j], bad_label);
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        b[j] = j < (int)cnt ? __float_as_uint(score[i64 + j]) : 0u;
        if (!RAW && j < (int)cnt) lab[j] = LabelOf<TargetT>::get(target[i64 + j], bad_label);
      }
    }
    if constexpr (!RAW) {
      uint32_t wj[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) wj[j] = __ballot_sync(kFull, lab[j] != 0u);  // (records past the end: 0)
      if (lane == 0 && (FULL || cnt))
        *reinterpret_cast<uint4*>(label_bits + (i64 >> 7) * 4) = make_uint4(wj[0], wj[1], wj[2], wj[3]);
    }
    const uint32_t h0 = s_hotseg[0], h1 = s_hotseg[1];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t k;
      if (RAW) {
        k = b[j];
        wide |= k >> 31;
      } else {
        const uint32_t bb = b[j] == 0x80000000u ? 0u : b[j];  // -0.0 ties with +0.0
        neg_score |= bb >> 31;
        wide |= (bb >> 30) & 1u;  // key bit 31: score >= 2.0
        k = bb << 1;              // (the label bit does not reach the digits)
      }
      const uint32_t seg = (k >> kLeafBits) & (kLeafSegs - 1u);
      bool cold = FULL || j < (int)cnt;
      if (h0 != 0xffffffffu) {  // uniform over the CTA
        const bool is0 = cold && seg == h0, is1 = cold && seg == h1;
        hc0 += __popc(__ballot_sync(kFull, is0));
        hc1 += __popc(__ballot_sync(kFull, is1));
        cold = cold && !is0 && !is1;
      }
      if (cold) atomicAdd(&s_j[seg >> 1], 1u << (16u * (seg & 1u)));
    }
  };
  for (uint32_t c0 = blockIdx.x; c0 < nchunks; c0 += U * gridDim.x) {
    int4 sv[U], t0v[U], t1v[U];
    uint32_t cvalid[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {  // all loads of the round first
      const uint32_t c = c0 + u * gridDim.x;
      cvalid[u] = c < nchunks ? view_valid_tile(view, c, kUjChunk, n) : 0u;
      if (cvalid[u] == kUjChunk && vec_ok) {
        const uint64_t i64 = (uint64_t)c * kUjChunk + 4ull * tid;
        sv[u] = ld_stream_v4(score + i64);
        if constexpr (!RAW && sizeof(TargetT) == 8) {
          t0v[u] = ld_stream_v4(target + i64);
          t1v[u] = ld_stream_v4(target + i64 + 2);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (cvalid[u] == 0u) continue;  // uniform over the CTA
      const uint64_t i64 = (uint64_t)(c0 + u * gridDim.x) * kUjChunk + 4ull * tid;
      if (cvalid[u] == kUjChunk && vec_ok) {
        count_chunk(std::true_type{}, i64, 4u, sv[u], t0v[u], t1v[u]);
      } else {
        const uint32_t cnt = 4u * tid < cvalid[u] ? min(4u, cvalid[u] - 4u * tid) : 0u;
        count_chunk(std::false_type{}, i64, cnt, sv[u], t0v[u], t1v[u]);
      }
      if (++steps == kUjFlushEvery) {
        flush();
        steps = 0;
      }
    }
  }
  flush();
  const uint32_t f = (__any_sync(kFull, bad_label != 0) ? TE_RANK_FLAG_BAD_LABEL : 0) |
                     (__any_sync(kFull, neg_score != 0) ? TE_RANK_FLAG_NEGATIVE_SCORE : 0) |
                     (__any_sync(kFull, wide != 0) ? TE_RANK_FLAG_WIDE_SCORE : 0);
  if (lane == 0 && f) atomicOr(flags, f);
}

// Tile table of the second pass: bucket b (= value of key bits 23..30) occupies
// [seg_start[256 b], seg_start[256 (b + 1)]) of the first pass's output; its tiles start at that
// start rounded down to a multiple of 4 records (16-byte copies), so the first tile of a bucket may
// begin with up to 3 records of the previous bucket, which are masked out.
struct MsdTables {
  uint32_t tile_prefix[kRadix + 1];  // tiles of the buckets before b; [256] = total
  // Which instantiation of a pass does the work (both are launched, the other one returns at once).
  // [1], pass 2: the one with eight counter copies when most records sit in buckets where one run
  // holds >= 3/8 (tied or quantised scores); [0], pass 1: always 0, it has a single instantiation.
  uint32_t use_hot[2];
};

// One CTA: exclusive prefix of the joint histogram -> seg_start[65537] (the segment bounds of the
// counting stage), the cursors of both passes, the tile table of the second pass.
template <int TILE>
__global__ void __launch_bounds__(1024)
    joint_scan_kernel(const uint32_t* __restrict__ joint, uint32_t* __restrict__ seg_start, uint32_t* __restrict__ cursor1,
                      uint32_t* __restrict__ cursor2, MsdTables* __restrict__ tables, size_t bstride = 0,
                      uint32_t hot8_top = 3, uint32_t hot8_low = 3) {
  joint = batch_ptr(joint, bstride), seg_start = batch_ptr(seg_start, bstride), cursor1 = batch_ptr(cursor1, bstride);
  cursor2 = batch_ptr(cursor2, bstride), tables = batch_ptr(tables, bstride);
  __shared__ uint32_t ws[32];
  __shared__ uint32_t s_tiles[kRadix], s_size[kRadix], s_hotsum[kRadix], s_bmax[kRadix], s_w8[3][8];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int ROWS = kLeafSegs / 1024;  // 64 rows of 32 segments per warp = 8 buckets of pass 1
  uint32_t tot = 0;
#pragma unroll 16
  for (int j = 0; j < ROWS; ++j) tot += joint[(warp * ROWS + j) * 32 + lane];
  tot = warp_sum(tot);
  if (lane == 0) ws[warp] = tot;
  __syncthreads();
  uint32_t at = 0;
  for (int w = 0; w < (int)warp; ++w) at += ws[w];
  uint32_t bmax = 0;  // fullest run of the bucket being scanned (lanes hold partial maxima)
  for (int j0 = 0; j0 < ROWS; j0 += 8) {  // one bucket of pass 1 per round
   uint32_t c8[8];  // eight independent loads in flight, then eight warp scans (one CTA: latency is all there is)
#pragma unroll
   for (int q = 0; q < 8; ++q) c8[q] = joint[(warp * ROWS + j0 + q) * 32 + lane];
#pragma unroll
   for (int q = 0; q < 8; ++q) {
    const uint32_t s = (warp * ROWS + j0 + q) * 32 + lane;
    const uint32_t c = c8[q];
    bmax = max(bmax, c);
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(kFull, incl, o);
      if (lane >= (unsigned)o) incl += u;
    }
    const uint32_t excl = at + incl - c;
    seg_start[s] = excl;
    cursor2[s] = excl;
    if ((s & (kRadix - 1)) == 0) cursor1[s >> kRadixBits] = excl;
    at += __shfl_sync(kFull, incl, 31);
    if (s == kLeafSegs - 1) seg_start[kLeafSegs] = excl + c;
   }
#pragma unroll
   for (int o = 16; o > 0; o >>= 1) bmax = max(bmax, __shfl_xor_sync(kFull, bmax, o));
   if (lane == 0) s_bmax[(warp * ROWS + j0) >> 3] = bmax;  // bucket = segment >> 8 = row >> 3
   bmax = 0;
  }
  __syncthreads();  // seg_start of every bucket boundary is in global memory, visible to this CTA
  // 256 threads, one per bucket: its tiles, its size, whether one run dominates it; then the
  // exclusive prefix of the tiles and the three totals by warp scans (8 warps)
  uint32_t my_tiles = 0, my_size = 0, my_hot = 0;
  if (tid < kRadix) {
    const uint32_t B = seg_start[tid * kRadix], E = seg_start[(tid + 1) * kRadix];
    const uint32_t A = B & ~3u;
    my_tiles = E > B ? (E - A + TILE - 1) / TILE : 0u;
    my_size = E - B;
    my_hot = (unsigned long long)s_bmax[tid] * 8 >= (unsigned long long)hot8_low * (E - B) ? E - B : 0u;  // dominated by one run
    uint32_t i_t = my_tiles, i_s = my_size, i_h = my_hot, mx = my_size;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u0 = __shfl_up_sync(kFull, i_t, o), u1 = __shfl_up_sync(kFull, i_s, o), u2 = __shfl_up_sync(kFull, i_h, o);
      if (lane >= (unsigned)o) i_t += u0, i_s += u1, i_h += u2;
      mx = max(mx, __shfl_xor_sync(kFull, mx, o));
    }
    if (lane == 31) s_w8[0][warp] = i_t, s_w8[1][warp] = i_s, s_w8[2][warp] = i_h;
    if (lane == 0) s_size[warp] = mx;  // (the largest bucket of this warp's 32)
    asm volatile("bar.sync 1, 256;" ::: "memory");
    uint32_t base_t = 0, n = 0, hot = 0, c0 = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      if (w < (int)warp) base_t += s_w8[0][w];
      n += s_w8[1][w];
      hot += s_w8[2][w];
      c0 = max(c0, s_size[w]);
    }
    tables->tile_prefix[tid] = base_t + i_t - my_tiles;
    if (tid == kRadix - 1) {
      tables->tile_prefix[kRadix] = base_t + i_t;
      // (what hurts is the multiplicity on ONE counter: two digits of 25 % each — uniform [0, 1)
      // scores — are served by plain atomics)
      tables->use_hot[0] = (unsigned long long)c0 * 8 >= (unsigned long long)hot8_top * n && n ? 1u : 0u;
      tables->use_hot[1] = (unsigned long long)hot * 2 >= n && n ? 1u : 0u;
    }
  }
  (void)s_tiles, (void)s_hotsum;
}

// One partition pass with cursor placement.  Persistent CTAs, tile t = blockIdx.x + k * gridDim.x.
//   BUCKETED == false (pass 1): one bucket = the whole input of n records; digit = key bits
//     23..30; cursors = cursor1[256].  PACK: keys_in holds the raw float32 scores and label_bits
//     the labels (one bit per record): the key is packed on the way into registers.
//   BUCKETED == true (pass 2): the input is the output of pass 1; tile t belongs to bucket b
//     (tables->tile_prefix), digit = key bits 15..22, cursors = cursor2 + 256 b.
// Slot i of thread t is record (i / 4 * BLOCK + t) * 4 + i % 4 of the tile (16-byte accesses); ranks
// inside a digit are handed out in arrival order (one shared atomic per record on a CTA-wide table).
// REPL: the CTA-wide counter table is kept in REPL copies, lane l counting in copy l % REPL, the
// copies 8 banks apart.  The digit of pass 1 is the float exponent: uniform [0, 1) scores put 25 % of
// the records on each of two counters, 8 lanes of every instruction on one address, and a shared
// atomic serialises them; with 4 copies it is 2 lanes per address.
// HOT: which of the two instantiations launched for pass 2 this is (tables->use_hot picks one: the
// plain one, or the one with more copies for buckets dominated by a single run; pass 1 has one).
template <bool BUCKETED, bool PACK, bool HOT, int BLOCK = kPrBlock, int IPT = kPrIpt, int REPL = 1>
__global__ void __launch_bounds__(BLOCK, 1024 / BLOCK)
    cursor_partition_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ label_bits,
                            uint32_t* __restrict__ keys_out, uint32_t n, const uint32_t* __restrict__ n_dev,
                            const uint32_t* __restrict__ seg_start, const MsdTables* __restrict__ tables,
                            uint32_t* cursors, const SlabView view = SlabView{nullptr, 0, 0, 0}, size_t bstride = 0) {
  // bstride: grid.y independent problems of n records each (one-vs-rest classes), one scratch slice apart
  keys_in = batch_ptr(keys_in, bstride), keys_out = batch_ptr(keys_out, bstride), label_bits = batch_ptr(label_bits, bstride);
  seg_start = batch_ptr(seg_start, bstride), tables = batch_ptr(tables, bstride), cursors = batch_ptr(cursors, bstride);
  // view (pass 1 only): the input is an inbox of slabs; tiles are tiles of the physical index space
  static_assert(!(BUCKETED && PACK), "only the first pass packs");
  extern __shared__ __align__(16) uint32_t s_buf[];  // [2][TILE]
  constexpr int TILE = BLOCK * IPT;
  static_assert(IPT % 4 == 0 && TILE <= 65536 && BLOCK >= kRadix, "16-byte slots; 16-bit ranks");
  constexpr int SHIFT = BUCKETED ? kLeafBits : kLeafBits + kRadixBits;
  constexpr int kCopy = kRadix + 32 / REPL;  // words per copy: copy c starts c * 32 / REPL banks further on
  static_assert(REPL == 1 || REPL == 2 || REPL == 4 || REPL == 8, "copies are picked by the low lane bits");
  __shared__ uint32_t s_cnt[REPL * kCopy], s_start[REPL * kCopy], s_goff[kRadix], s_wsum[kRadix / kWarp];
  __shared__ uint32_t s_tp[kRadix + 1];
  __shared__ uint32_t s_bs[BUCKETED ? kRadix + 1 : 1];  // where the buckets of pass 1 start (pass 2 looks them up per tile)
  __shared__ __align__(16) uint32_t s_bits[2][PACK ? TILE / 32 : 4];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if ((tables->use_hot[BUCKETED ? 1 : 0] != 0u) != HOT) return;  // the other instantiation's turn
  if (n_dev) n = *n_dev;  // (the record count lives in device memory: receive side of the sample sort)
  if (BUCKETED)
    for (int i = tid; i <= kRadix; i += BLOCK) s_tp[i] = tables->tile_prefix[i], s_bs[i] = seg_start[i * kRadix];
  for (int i = tid; i < REPL * kCopy; i += BLOCK) s_cnt[i] = 0u;
  __syncthreads();
  const uint32_t my_copy = (lane & (uint32_t)(REPL - 1)) * (uint32_t)kCopy;
  const bool slabs = !BUCKETED && view.counts != nullptr;
  const uint32_t ntiles = BUCKETED ? s_tp[kRadix] : (slabs ? view.nslabs * (view.cap / TILE) : (n + TILE - 1) / TILE);
  auto digit = [&](uint32_t k) -> uint32_t { return (k >> SHIFT) & (kRadix - 1); };
  auto rec = [&](int i) -> uint32_t { return (((uint32_t)i >> 2) * BLOCK + tid) * 4u + ((uint32_t)i & 3u); };
  // where tile t lives: first record (multiple of 4), the records [lo, hi) of it that are its own,
  // and its bucket
  struct Where {
    uint32_t base, lo, hi, bucket;
  };
  auto locate = [&](uint32_t t) -> Where {
    Where w;
    if (!BUCKETED) {
      w.base = t * (uint32_t)TILE, w.lo = 0u, w.bucket = 0u;
      w.hi = t < ntiles ? view_valid_tile(view, t, TILE, n) : 0u;
      return w;
    }
    uint32_t b = 0;  // last bucket whose tile_prefix <= t
#pragma unroll
    for (int step = kRadix / 2; step >= 1; step >>= 1)
      if (s_tp[b + step] <= t) b += step;
    const uint32_t B = s_bs[BUCKETED ? b : 0], E = s_bs[BUCKETED ? b + 1 : 0];
    w.bucket = b;
    w.base = (B & ~3u) + (t - s_tp[b]) * (uint32_t)TILE;
    w.lo = B > w.base ? B - w.base : 0u;
    w.hi = min((uint32_t)TILE, E - w.base);
    return w;
  };
  auto prefetch = [&](uint32_t t, uint32_t half) {
    if (t < ntiles) {
      const Where w = locate(t);
      if (slabs ? w.hi != 0u : (uint64_t)w.base + TILE <= n) {  // the whole tile is inside the array / the inbox (what is not the tile's own is masked later)
        uint32_t* dst = s_buf + half * TILE;
        const uint32_t* src = keys_in + w.base;
#pragma unroll
        for (int q = 0; q < IPT / 4; ++q) cp_async_16(dst + rec(4 * q), src + rec(4 * q));
        if (PACK && tid < TILE / 128) cp_async_16(&s_bits[half][tid * 4], label_bits + (size_t)(w.base >> 5) + tid * 4);
      }
    }
    cp_async_commit();  // (an empty group keeps the wait_group arithmetic uniform)
  };
  prefetch(blockIdx.x, 0u);
  uint32_t half = 0u;
  for (uint32_t t = blockIdx.x; t < ntiles; t += gridDim.x, half ^= 1u) {
    const Where w = locate(t);
    uint32_t* s_keys = s_buf + half * TILE;
    prefetch(t + gridDim.x, half ^ 1u);  // into the half whose tile left at the end of the previous round
    cp_async_wait<1>();                  // everything but the newest group: this tile has landed
    __syncthreads();
    if (w.hi <= w.lo) continue;  // (the empty tail of a slab)
    uint32_t key[IPT];
    if (slabs || (uint64_t)w.base + TILE <= n) {
#pragma unroll
      for (int q = 0; q < IPT / 4; ++q) {
        const uint4 a = *reinterpret_cast<const uint4*>(s_keys + rec(4 * q));
        key[4 * q] = a.x, key[4 * q + 1] = a.y, key[4 * q + 2] = a.z, key[4 * q + 3] = a.w;
        if (PACK) {  // the four label words of my warp's 128-record group (one broadcast load)
          const uint4 lw = *reinterpret_cast<const uint4*>(&s_bits[half][(q * (BLOCK / 32) + warp) * 4]);
          const uint32_t wj[4] = {lw.x, lw.y, lw.z, lw.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t bb = key[4 * q + j] == 0x80000000u ? 0u : key[4 * q + j];  // -0.0 ties with +0.0
            key[4 * q + j] = (bb << 1) | ((wj[j] >> lane) & 1u);
          }
        }
      }
    } else {  // the tile that runs past the end of the array
#pragma unroll
      for (int i = 0; i < IPT; ++i) {
        const uint32_t r = rec(i);
        key[i] = (r >= w.lo && r < w.hi) ? keys_in[w.base + r] : 0u;
        if (PACK && r < w.hi) {
          const size_t idx = (size_t)w.base + r;
          const uint32_t bb = key[i] == 0x80000000u ? 0u : key[i];
          key[i] = (bb << 1) | ((label_bits[(idx >> 7) * 4 + (idx & 3)] >> ((idx >> 2) & 31)) & 1u);
        }
      }
    }
    // ---- rank: arrival order inside (digit, copy), straight from the CTA-wide counters ----
    // Skewed digits (one binade = one top digit; tied or quantised scores = few low digits): up to 32
    // lanes of an instruction hit ONE counter and the atomic serialises them (two values 50 / 50:
    // 5.6 instead of 3.0 ms per pass with one table).  The copies spread them over REPL addresses.
    uint32_t rk[IPT / 2];
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
      const uint32_t r0 = rec(i);
      const uint32_t d = digit(key[i]);
      const bool in = r0 >= w.lo && r0 < w.hi;
      const uint32_t r = in ? atomicAdd(&s_cnt[my_copy + d], 1u) : 0u;
      if (i & 1)
        rk[i >> 1] |= r << 16;
      else
        rk[i >> 1] = r;
    }
    __syncthreads();  // all counts in; every thread holds its records, the buffer may be overwritten
    // ---- 256 digit threads: tile-local run starts, and room in the output from the global cursor
    //      (its round trip overlaps with the scatter) ----
    uint32_t g = 0u, excl = 0u;
    if (tid < kRadix) {
      uint32_t cc[REPL], c = 0u;
#pragma unroll
      for (int q = 0; q < REPL; ++q) cc[q] = s_cnt[q * kCopy + tid], c += cc[q];
      uint32_t incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(kFull, incl, o);
        if (lane >= (unsigned)o) incl += u;
      }
      if (lane == 31) s_wsum[warp] = incl;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      uint32_t wb = 0;
#pragma unroll
      for (int ww = 0; ww < kRadix / kWarp; ++ww)
        if (ww < (int)warp) wb += s_wsum[ww];
      excl = wb + incl - c;
      uint32_t at = excl;  // the copies of a digit follow one another inside its run
#pragma unroll
      for (int q = 0; q < REPL; ++q) {
        s_start[q * kCopy + tid] = at;
        at += cc[q];
        s_cnt[q * kCopy + tid] = 0u;  // for the next tile
      }
      if (c) g = atomicAdd(cursors + (BUCKETED ? w.bucket * kRadix : 0u) + tid, c);
    }
    __syncthreads();
    // ---- scatter into tile order ----
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
      const uint32_t r0 = rec(i);
      const uint32_t r = (i & 1) ? rk[i >> 1] >> 16 : rk[i >> 1] & 0xffffu;
      const uint32_t d = digit(key[i]);
      if (r0 >= w.lo && r0 < w.hi) s_keys[s_start[my_copy + d] + r] = key[i];
    }
    if (tid < kRadix) s_goff[tid] = g - excl;  // output index = s_goff[d] + tile position
    __syncthreads();
    // ---- output: every digit run leaves as one contiguous burst ----
    const uint32_t cnt = w.hi - w.lo;
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      const uint32_t p = tid + k * BLOCK;
      if (p < cnt) {
        const uint32_t kk = s_keys[p];
        keys_out[s_goff[digit(kk)] + p] = kk;
      }
    }
    __syncthreads();  // this half, s_start, s_goff are rewritten from the next round on
  }
  cp_async_wait<0>();
}

}  // namespace te
