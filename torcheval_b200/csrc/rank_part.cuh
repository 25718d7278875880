// First partition pass of the counting-stage AUROC path as an upsweep / downsweep pair without
// any traffic between thread blocks.
//
// The array is cut into R contiguous RANGES of whole 8192-record tiles, one per resident CTA.
//   upsweep   = the ingest kernel below: it reads (score, target), writes the packed keys, and
//               leaves, per range, the histogram of the digit the first pass partitions by (plus
//               the global histograms of both digits, as before);
//   range_base_kernel turns those into the output offset of every (range, digit) pair;
//   downsweep = partition_ranges_kernel: a persistent CTA walks the tiles of its range with the
//               running output offset of each digit in a register.  No look-back, no flags, no
//               spinning: nothing a CTA needs is produced while it runs.
// A first pass of a keys-only sort has no order to preserve, so ranks inside a tile come from ONE
// shared atomic per record on a CTA-wide table of 256 counters (arrival order): no per-warp
// tables to clear, sum and scan (rank_sort.cuh pays ~4 shared wavefronts per 32 records for
// those), no digit phase.
#pragma once
#include "rank_leaf.cuh"
#include "rank_sort.cuh"

namespace te {

// Downsweep geometry: 1024 threads x 24 records = 24 576-record tiles, one CTA per SM.  The runs a
// tile writes per digit are 96 records long on average; with 8192-record tiles (32-record runs, three
// CTAs per SM) the same kernel took 3.38 ms instead of 2.50 ms at 2^30 keys, with 4096-record tiles
// 4.33 ms (profiles/r02_part_block.txt): the length of the output bursts decides.
constexpr int kPrBlock = 1024;
constexpr int kPrIpt = 24;
constexpr int kPrBlockSmall = 512, kPrIptSmall = 16;  // the 8192-record geometry (TE_B200_PART_BLOCK=512)
constexpr int kIrBlock = 512;
constexpr size_t kIrSmem = (size_t)2 * kRadix * kWarp * sizeof(uint32_t);  // two lane-replicated histograms: 64 KB

struct RangeGeom {
  uint32_t ranges, range_records;  // records per range: a multiple of kRangeUnit
};
constexpr uint32_t kRangeUnit = 24576;  // a multiple of the tile sizes of both downsweep geometries (1024 x 24 and 512 x 16)
inline RangeGeom range_geom(uint32_t n, int resident_ctas) {
  RangeGeom g;
  uint32_t units = (n + kRangeUnit - 1) / kRangeUnit;
  if (units < 1) units = 1;
  const uint32_t want = (uint32_t)resident_ctas < units ? (uint32_t)resident_ctas : units;
  const uint32_t per = (units + want - 1) / want;
  g.range_records = per * kRangeUnit;
  g.ranges = (units + per - 1) / per;
  return g;
}

// Where the records of a first pass live: one contiguous array of n records, or — on the receive
// side of the distributed sample sort — `nslabs` slabs of `cap` records each (one per sending
// rank, cap a multiple of every tile size), of which only the first counts[s * count_stride]
// hold records.  The slab counts are device data (written by the senders), so nothing here
// needs the host to know how many records arrived.
struct SlabView {
  const uint32_t* counts;  // nullptr: contiguous
  uint32_t count_stride, cap, nslabs;
};
// number of records in tile `t` (tiles of `tile` records) of the physical index space; a tile never
// straddles slabs.  32-bit arithmetic only: this runs per tile / per 4096-record chunk in every
// thread, and a 64-bit division there was a third of the instructions of the slab upsweep.
__device__ __forceinline__ uint32_t view_valid_tile(const SlabView& v, uint32_t t, uint32_t tile, uint32_t n) {
  if (!v.counts) {
    const uint64_t base = (uint64_t)t * tile;
    return base < n ? (uint32_t)min((uint64_t)tile, (uint64_t)n - base) : 0u;
  }
  const uint32_t per = v.cap / tile;  // tiles per slab
  const uint32_t s = t / per;
  if (s >= v.nslabs) return 0u;
  const uint32_t off = (t - s * per) * tile, c = min(v.counts[s * v.count_stride], v.cap);
  return off < c ? min(tile, c - off) : 0u;
}
__device__ __forceinline__ uint32_t view_valid(const SlabView& v, uint64_t base, uint32_t tile, uint32_t n) {
  return view_valid_tile(v, (uint32_t)(base / tile), tile, n);
}

// Upsweep of the first pass, range by range.  flags: TE_RANK_FLAG_* seen.  hist: uint32[2][256]
// (+=), digits at key bits 15..22 and 23..30; range_hist: uint32[ranges][256] (=), the first digit
// per range.  What else it does depends on MODE:
//   kIrKeys : (score, target) -> packed keys written to keys_out (12 B read + 4 B written per record);
//   kIrBits : (score, target) -> the 0/1 labels only, ONE BIT per record, written to keys_out: the
//             first pass then packs keys itself from the scores and these bits, and the packed keys
//             never make a round trip through HBM (12 + 1/8 B per record here).  Layout: records
//             come in groups of 128; word j (0..3) of a group holds, in bit L, the label of record
//             4 L + j of the group — what a warp of the first pass reads with one 16-byte load;
//   kIrRaw  : `score` holds packed keys already (the receive side of the distributed sample sort):
//             histograms and the wide-key flag only, nothing is written.
enum IngestRangesMode { kIrKeys = 0, kIrRaw = 1, kIrBits = 2 };
template <typename TargetT, int MODE>
__global__ void __launch_bounds__(kIrBlock)
    rank_ingest_ranges_kernel(const float* __restrict__ score, const TargetT* __restrict__ target, uint32_t n,
                              bool vec_ok, uint32_t range_records, uint32_t* __restrict__ keys_out,
                              uint32_t* ghist, uint32_t* __restrict__ range_hist, uint32_t* flags) {
  constexpr bool RAW_KEYS = MODE == kIrRaw;
  extern __shared__ uint32_t sh[];  // [2][256][32 lanes]
  const unsigned tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < 2 * kRadix * kWarp; i += kIrBlock) sh[i] = 0;
  __syncthreads();
  const uint64_t e0 = (uint64_t)blockIdx.x * range_records;
  const uint32_t e1 = (uint32_t)min((uint64_t)n, e0 + (uint64_t)range_records);
  uint32_t bad_label = 0, neg_score = 0, wide = 0;
  // a warp takes 128 consecutive records per step; the loop bound is the warp's, so that every lane
  // is there for the ballots of kIrBits
  for (uint64_t w64 = e0 + 4ull * (tid - lane); w64 < e1; w64 += 4ull * kIrBlock) {
    const uint64_t i64 = w64 + 4ull * lane;
    const uint32_t i0 = (uint32_t)i64;
    const uint32_t cnt = i64 < e1 ? min(4u, e1 - i0) : 0u;
    uint32_t b[4] = {0, 0, 0, 0}, lab[4] = {0, 0, 0, 0};
    if (cnt == 4 && vec_ok) {
      const int4 a = ld_stream_v4(score + i0);
      b[0] = (uint32_t)a.x, b[1] = (uint32_t)a.y, b[2] = (uint32_t)a.z, b[3] = (uint32_t)a.w;
      if constexpr (!RAW_KEYS) {
        if constexpr (sizeof(TargetT) == 8) {
          const int4 t0 = ld_stream_v4(target + i0);
          const int4 t1 = ld_stream_v4(target + i0 + 2);
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
          for (int j = 0; j < 4; ++j) lab[j] = LabelOf<TargetT>::get(target[i0 + j], bad_label);
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        b[j] = j < (int)cnt ? __float_as_uint(score[i0 + j]) : 0u;
        if (!RAW_KEYS && j < (int)cnt) lab[j] = LabelOf<TargetT>::get(target[i0 + j], bad_label);
      }
    }
    uint32_t k[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (RAW_KEYS) {
        k[j] = b[j];
        wide |= b[j] >> 31;
      } else {
        const uint32_t bb = b[j] == 0x80000000u ? 0u : b[j];  // -0.0 ties with +0.0
        neg_score |= bb >> 31;
        k[j] = (bb << 1) | lab[j];
        wide |= (bb >> 30) & 1u;  // key bit 31: score >= 2.0
      }
    }
    if constexpr (MODE == kIrKeys) {
      if (cnt == 4) {
        st_stream_v4(keys_out + i0, make_int4((int)k[0], (int)k[1], (int)k[2], (int)k[3]));
      } else {
        for (uint32_t j = 0; j < cnt; ++j) keys_out[i0 + j] = k[j];
      }
    }
    if constexpr (MODE == kIrBits) {
      uint32_t wj[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) wj[j] = __ballot_sync(kFull, lab[j] != 0u);  // (records past the end: 0)
      if (lane == 0)
        *reinterpret_cast<uint4*>(keys_out + (w64 >> 7) * 4) = make_uint4(wj[0], wj[1], wj[2], wj[3]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (j < (int)cnt) {
        atomicAdd(sh + ((k[j] >> kLeafBits) & 0xffu) * kWarp + lane, 1u);
        atomicAdd(sh + (kRadix + ((k[j] >> (kLeafBits + 8)) & 0xffu)) * kWarp + lane, 1u);
      }
    }
  }
  __syncthreads();
  {
    uint32_t s = 0;  // thread t owns (digit position, digit) = t
#pragma unroll
    for (int r = 0; r < kWarp; ++r) s += sh[tid * kWarp + ((r + tid) & 31)];
    if (tid < kRadix) range_hist[(size_t)blockIdx.x * kRadix + tid] = s;
    if (s) atomicAdd(ghist + tid, s);
  }
  const uint32_t f = (__any_sync(kFull, bad_label != 0) ? TE_RANK_FLAG_BAD_LABEL : 0) |
                     (__any_sync(kFull, neg_score != 0) ? TE_RANK_FLAG_NEGATIVE_SCORE : 0) |
                     (__any_sync(kFull, wide != 0) ? TE_RANK_FLAG_WIDE_SCORE : 0);
  if (lane == 0 && f) atomicOr(flags, f);
}

// range_base[r][d] = digit_base[d] + sum over ranges r' < r of range_hist[r'][d]: where the records
// of range r with digit d start in the output of the first pass.
__global__ void __launch_bounds__(kRadix)
    range_base_kernel(const uint32_t* __restrict__ range_hist, uint32_t ranges,
                      const uint32_t* __restrict__ digit_base, uint32_t* __restrict__ range_base) {
  const unsigned d = threadIdx.x;
  uint32_t acc = digit_base[d];
#pragma unroll 8
  for (uint32_t r = 0; r < ranges; ++r) {
    const uint32_t c = range_hist[(size_t)r * kRadix + d];
    range_base[(size_t)r * kRadix + d] = acc;
    acc += c;
  }
}

// 16-byte asynchronous global -> shared copy (LDGSTS): no register, no scoreboard stall; completion
// is observed per thread with wait_group and made CTA-visible by the barrier that follows.
__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// Downsweep: one persistent CTA per range.  Slot i of thread t is record (i / 4 * 512 + t) * 4 + i % 4
// of the tile (16-byte accesses); any assignment will do, ranks are handed out in arrival order.
// The tiles of a range are known in advance, so tile k + 1 is copied into the other half of a
// two-tile shared buffer (cp.async) while tile k is ranked, scattered INTO ITS OWN BUFFER (every
// thread holds its records in registers by then) and written out: the kernel was bound by the
// latency of the loads at the head of every tile (ncu: long_scoreboard the top stall, LSU data
// pipe at 52 %: profiles/r02_ncu_summary_v1.txt), not by the pipe.
template <typename DigitOp, int BLOCK = kPrBlock, int IPT = kPrIpt>
__global__ void __launch_bounds__(BLOCK, BLOCK >= 1024 ? 1 : 3)
    partition_ranges_kernel(const uint32_t* __restrict__ keys_in, uint32_t* __restrict__ keys_out, uint32_t n,
                            const DigitOp dop, const uint32_t* __restrict__ range_base, uint32_t range_records,
                            uint32_t ranges_per_cta, const uint32_t* __restrict__ skip_flag,
                            const SlabView view = SlabView{nullptr, 0, 0, 0},
                            const uint32_t* __restrict__ label_bits = nullptr) {
  // label_bits given: keys_in holds the raw float32 SCORES and the labels come as one bit per
  // record (rank_ingest_ranges_kernel<kIrBits>); the key is packed here, on the way into registers
  extern __shared__ __align__(16) uint32_t s_buf[];  // [2][TILE]
  constexpr int TILE = BLOCK * IPT;
  static_assert(IPT % 4 == 0 && TILE <= 65536, "16-byte slots; ranks are packed as 16-bit halves");
  static_assert(kRangeUnit % TILE == 0 && BLOCK >= kRadix, "ranges are made of whole tiles");
  __shared__ uint32_t s_cnt[kRadix], s_start[kRadix], s_goff[kRadix], s_wsum[kRadix / kWarp];
  __shared__ __align__(16) uint32_t s_bits[2][TILE / 32];  // label bits of the two tiles in flight
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // Contiguous input: a CTA takes ranges_per_cta consecutive ranges — where one range ends in the
  // output of a digit is where the next one starts, so only the first range's bases are needed.
  // Slab input: range r is the tiles r, r + R, r + 2R, ... of the physical index space (R = grid
  // size), so that the empty tails of the slabs are spread evenly over the CTAs.
  const bool strided = view.counts != nullptr;
  const uint32_t ntiles = strided ? view.nslabs * (view.cap / TILE) : (n + TILE - 1) / TILE;
  const uint32_t tiles_per_cta = range_records / TILE * ranges_per_cta;
  const uint32_t tstep = strided ? gridDim.x : 1u;
  const uint32_t t0 = strided ? blockIdx.x : blockIdx.x * tiles_per_cta;
  const uint32_t t1 = strided ? ntiles : min(ntiles, t0 + tiles_per_cta);
  const bool skip_pass = skip_flag ? ld_relaxed_u32(skip_flag) != 0u : false;
  uint32_t run = tid < kRadix ? range_base[(size_t)blockIdx.x * ranges_per_cta * kRadix + tid] : 0u;  // output offset of digit tid
  if (tid < kRadix) s_cnt[tid] = 0u;
  auto rec = [&](int i) -> uint32_t { return (((uint32_t)i >> 2) * BLOCK + tid) * 4u + ((uint32_t)i & 3u); };
  // full tiles travel through cp.async; a partial one (the last of the array) is read directly
  auto prefetch = [&](uint32_t tile, uint32_t half) {
    if (tile < t1 && view_valid_tile(view, tile, TILE, n) == (uint32_t)TILE) {
      uint32_t* dst = s_buf + half * TILE;
      const uint32_t* src = keys_in + (size_t)tile * TILE;
#pragma unroll
      for (int q = 0; q < IPT / 4; ++q) cp_async_16(dst + rec(4 * q), src + rec(4 * q));
      if (label_bits && tid < TILE / 128) cp_async_16(&s_bits[half][tid * 4], label_bits + (size_t)tile * (TILE / 32) + tid * 4);
    }
    cp_async_commit();  // (an empty group keeps the wait_group arithmetic uniform)
  };
  prefetch(t0, 0u);
  uint32_t half = 0u;
  for (uint32_t tile = t0; tile < t1; tile += tstep, half ^= 1u) {
    const size_t base = (size_t)tile * TILE;
    const uint32_t valid = view_valid_tile(view, tile, TILE, n);
    uint32_t* s_keys = s_buf + half * TILE;
    prefetch(tile + tstep, half ^ 1u);   // into the half whose tile left at the end of the previous round
    cp_async_wait<1>();   // everything but the newest group: this tile has landed
    __syncthreads();
    uint32_t key[IPT];
    if (valid == (uint32_t)TILE) {
#pragma unroll
      for (int q = 0; q < IPT / 4; ++q) {
        const uint4 a = *reinterpret_cast<const uint4*>(s_keys + rec(4 * q));
        key[4 * q] = a.x, key[4 * q + 1] = a.y, key[4 * q + 2] = a.z, key[4 * q + 3] = a.w;
        if (label_bits) {  // the four label words of my warp's 128-record group (one broadcast load)
          const uint4 w = *reinterpret_cast<const uint4*>(&s_bits[half][(q * (BLOCK / 32) + warp) * 4]);
          const uint32_t wj[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t bb = key[4 * q + j] == 0x80000000u ? 0u : key[4 * q + j];  // -0.0 ties with +0.0
            key[4 * q + j] = (bb << 1) | ((wj[j] >> lane) & 1u);
          }
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < IPT; ++i) {
        key[i] = rec(i) < valid ? keys_in[base + rec(i)] : 0u;
        if (label_bits && rec(i) < valid) {
          const size_t idx = base + rec(i);
          const uint32_t bb = key[i] == 0x80000000u ? 0u : key[i];
          key[i] = (bb << 1) | ((label_bits[(idx >> 7) * 4 + (idx & 3)] >> ((idx >> 2) & 31)) & 1u);
        }
      }
    }
    if (valid == 0u) continue;  // (the empty tail of a slab)
    if (skip_pass) {  // one digit value only: the order is free, the records only have to be packed densely
      if (!strided) {
#pragma unroll
        for (int i = 0; i < IPT; ++i)
          if (rec(i) < valid) keys_out[base + rec(i)] = key[i];
        continue;
      }
    }
    // ---- rank: arrival order inside the digit, straight from the CTA-wide counters ----
    uint32_t rk[IPT / 2];
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
      const uint32_t r = rec(i) < valid ? atomicAdd(&s_cnt[dop(key[i])], 1u) : 0u;
      if (i & 1)
        rk[i >> 1] |= r << 16;
      else
        rk[i >> 1] = r;
    }
    __syncthreads();  // all counts in; every thread holds its records, the buffer may be overwritten
    // ---- 256 digit threads: where each digit's run starts in the tile and in the output ----
    if (tid < kRadix) {
      const uint32_t c = s_cnt[tid];
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
      for (int w = 0; w < kRadix / kWarp; ++w)
        if (w < (int)warp) wb += s_wsum[w];
      const uint32_t excl = wb + incl - c;
      s_start[tid] = excl;
      s_goff[tid] = run - excl;  // output index = s_goff[d] + tile position
      run += c;
      s_cnt[tid] = 0u;           // for the next tile
    }
    __syncthreads();
    // ---- scatter into tile order ----
#pragma unroll
    for (int i = 0; i < IPT; ++i) {
      const uint32_t r = (i & 1) ? rk[i >> 1] >> 16 : rk[i >> 1] & 0xffffu;
      if (rec(i) < valid) s_keys[s_start[dop(key[i])] + r] = key[i];
    }
    __syncthreads();
    // ---- output: every digit run leaves as one contiguous burst ----
#pragma unroll
    for (int k = 0; k < IPT; ++k) {
      const uint32_t p = tid + k * BLOCK;
      if (p < valid) {
        const uint32_t kk = s_keys[p];
        keys_out[s_goff[dop(kk)] + p] = kk;
      }
    }
    __syncthreads();  // this half, s_start, s_goff are rewritten from the next round on
  }
  cp_async_wait<0>();
}

}  // namespace te
