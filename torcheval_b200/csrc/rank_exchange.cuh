// Exchange step of the distributed sample sort with per-sender slabs: no rank needs anybody's
// counts before it sends, and the host never needs to know how many records arrived.
//
// Every rank owns an INBOX of W slabs of `cap` records (W = ranks; symmetric memory, peer-mapped),
// slab q reserved for what rank q sends.  Sender side, one kernel (pack_scatter_kernel): read
// (score, target), pack into keys, find the destination rank (binary search over the W - 1
// splitters in shared memory), group the tile by destination in shared memory, reserve room in the
// destination slab with one atomicAdd per (tile, destination) on a cursor in the sender's OWN
// memory, and store every group straight into the peer's slab (NVLink P2P stores, 16 bytes each).
// 12 bytes read from HBM and 4 bytes written over NVLink per record; the packed keys never touch
// the sender's HBM.  A last one-block kernel stores the W cursors (= slab fill counts) and the
// sender's TE_RANK_FLAG_* bits into the header of every receiver's inbox.  After the barrier the
// receiver runs the counting-stage path over its slabs (SlabView, rank_part.cuh): the histogram
// kernel below, the first partition pass, the look-back pass and the counting stage, all taking
// the fill counts from device memory.
//
// A slab that would overflow (a split more skewed than the slack allows) is not written past its
// end; the sender reports TE_RANK_FLAG_SLAB_OVERFLOW and the caller redoes the call with the
// count-first protocol (te_rank_pack + te_partition_scatter_u32).
#pragma once
#include "rank_part.cuh"

namespace te {

constexpr uint32_t kRankFlagSlabOverflow = 0x100u;  // internal: travels in the inbox header only
constexpr int kXsBlock = 512;
constexpr int kXsIpt = 16;
constexpr int kXsTile = kXsBlock * kXsIpt;  // 8192 records

struct PeerSlabs {
  unsigned long long slab[kMaxSplitters + 1];  // address of MY slab inside rank b's inbox
};

template <typename TargetT>
__global__ void __launch_bounds__(kXsBlock, 2)
    pack_scatter_kernel(const float* __restrict__ score, const TargetT* __restrict__ target, uint32_t n, bool vec_ok,
                        const uint32_t* __restrict__ splitters_dev, int nsp, const PeerSlabs peers, uint32_t cap,
                        uint32_t* cursors /*[W], zeroed*/, uint32_t* flags) {
  __shared__ __align__(16) uint32_t s_keys[kXsTile];
  // the per-destination counters in kXsRepl copies (lane l counts in copy l % kXsRepl, the copies in
  // different banks): with W destinations 32 / W lanes of every instruction hit one counter, and a
  // shared atomic on one address serialises them (2 ranks: 16 deep)
  constexpr int kXsRepl = 8, kXsCopy = 20;
  __shared__ uint32_t s_sp[16], s_cnt[kXsRepl * kXsCopy], s_startc[kXsRepl * kXsCopy], s_start[kMaxSplitters + 2],
      s_dst[kMaxSplitters + 1];
  __shared__ unsigned long long s_peer[kMaxSplitters + 1];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t nb = (uint32_t)nsp + 1u;
  if (tid < 16) s_sp[tid] = (int)tid < nsp ? splitters_dev[tid] : 0xffffffffu;
  if (tid <= (unsigned)kMaxSplitters) s_peer[tid] = peers.slab[tid];
  if (tid < kXsRepl * kXsCopy) s_cnt[tid] = 0u;
  const uint32_t my_copy = (lane & (uint32_t)(kXsRepl - 1)) * (uint32_t)kXsCopy;
  __syncthreads();
  auto bucket_of = [&](uint32_t k) -> uint32_t {  // #{splitters <= k}: branch-free binary search
    if (nsp == 1) return s_sp[0] <= k ? 1u : 0u;
    uint32_t b = 0;
#pragma unroll
    for (int step = 8; step >= 1; step >>= 1)
      if (step <= nsp && s_sp[b + step - 1] <= k) b += step;
    return min(b, (uint32_t)nsp);
  };
  auto rec = [&](int i) -> uint32_t { return (((uint32_t)i >> 2) * kXsBlock + tid) * 4u + ((uint32_t)i & 3u); };
  uint32_t bad_label = 0, neg_score = 0, wide = 0, overflow = 0;
  const uint32_t ntiles = (n + kXsTile - 1) / kXsTile;
  for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const uint32_t base = tile * (uint32_t)kXsTile;
    const uint32_t valid = min((uint32_t)kXsTile, n - base);
    uint32_t key[kXsIpt];
#pragma unroll
    for (int q = 0; q < kXsIpt / 4; ++q) {
      const uint32_t r0 = rec(4 * q);
      uint32_t b[4], lab[4];
      if (r0 + 4 <= valid && vec_ok) {
        const int4 a = ld_stream_v4(score + base + r0);
        b[0] = (uint32_t)a.x, b[1] = (uint32_t)a.y, b[2] = (uint32_t)a.z, b[3] = (uint32_t)a.w;
        if constexpr (sizeof(TargetT) == 8) {
          const int4 t0 = ld_stream_v4(target + base + r0);
          const int4 t1 = ld_stream_v4(target + base + r0 + 2);
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
          for (int j = 0; j < 4; ++j) lab[j] = LabelOf<TargetT>::get(target[base + r0 + j], bad_label);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const bool in = r0 + j < valid;
          b[j] = in ? __float_as_uint(score[base + r0 + j]) : 0u;
          lab[j] = in ? LabelOf<TargetT>::get(target[base + r0 + j], bad_label) : 0u;
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint32_t bb = b[j] == 0x80000000u ? 0u : b[j];  // -0.0 ties with +0.0
        neg_score |= bb >> 31;
        wide |= (bb >> 30) & 1u;
        key[4 * q + j] = (bb << 1) | lab[j];
      }
    }
    // ---- group by destination: arrival order inside a destination (the receiver orders them) ----
    uint32_t rk[kXsIpt / 2], bk[kXsIpt / 8];  // 16-bit ranks, 4-bit destinations
#pragma unroll
    for (int i = 0; i < kXsIpt; ++i) {
      const uint32_t d = bucket_of(key[i]);
      const uint32_t r = rec(i) < valid ? atomicAdd(&s_cnt[my_copy + d], 1u) : 0u;
      if (i & 1)
        rk[i >> 1] |= r << 16;
      else
        rk[i >> 1] = r;
      if (i & 7)
        bk[i >> 3] |= d << (4 * (i & 7));
      else
        bk[i >> 3] = d;
    }
    __syncthreads();
    // <= 16 destinations: warp 0 scans the counts and reserves room in the slabs, one cursor atomic
    // per destination with records; the atomics' round trip overlaps with the scatter below
    uint32_t my_off = 0u, my_cnt = 0u;
    if (warp == 0) {
      uint32_t cc[kXsRepl], c = 0u;
#pragma unroll
      for (int q = 0; q < kXsRepl; ++q) cc[q] = lane < nb ? s_cnt[q * kXsCopy + lane] : 0u, c += cc[q];
      uint32_t incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(kFull, incl, o);
        if (lane >= (unsigned)o) incl += u;
      }
      if (lane <= nb) s_start[lane] = incl - c;  // (lane == nb: the tile total)
      if (lane < nb) {
        uint32_t at = incl - c;  // the copies of a destination follow one another inside its group
#pragma unroll
        for (int q = 0; q < kXsRepl; ++q) {
          s_startc[q * kXsCopy + lane] = at;
          at += cc[q];
          s_cnt[q * kXsCopy + lane] = 0u;
        }
        my_cnt = c;
        my_off = c ? atomicAdd(&cursors[lane], c) : 0u;  // (first looked at after the scatter)
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kXsIpt; ++i) {
      const uint32_t d = (bk[i >> 3] >> (4 * (i & 7))) & 15u;
      const uint32_t r = (i & 1) ? rk[i >> 1] >> 16 : rk[i >> 1] & 0xffffu;
      if (rec(i) < valid) s_keys[s_startc[my_copy + d] + r] = key[i];
    }
    if (warp == 0 && lane < nb) {
      if (my_off + my_cnt > cap) {  // the slab is full: nothing of this group is written, the call is redone
        overflow = 1u;
        my_off = 0xffffffffu;
      }
      s_dst[lane] = my_off;
    }
    __syncthreads();
    // ---- every group leaves as 16-byte stores into the peer's slab (scalar head and tail); the 16
    //      warps are dealt out over the groups ----
    constexpr uint32_t kWarps = kXsBlock / kWarp;
    const uint32_t wpr = nb < kWarps ? kWarps / nb : 1u;  // warps per group
    for (uint32_t d = warp / wpr; d < nb; d += kWarps / wpr) {
      const uint32_t j = warp % wpr;
      const uint32_t rs = s_start[d], rc = s_start[d + 1] - rs, off = s_dst[d];
      if (rc == 0 || off == 0xffffffffu) continue;
      uint32_t* dst = reinterpret_cast<uint32_t*>(s_peer[d]) + off;  // dst[i] <- s_keys[rs + i]
      const uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3u);
      const uint32_t head = min(rc, (4u - mis) & 3u);
      const uint32_t body = (rc - head) >> 2;
      const uint32_t done = head + 4 * body;
      if (j == 0) {
        if (lane < head) dst[lane] = s_keys[rs + lane];
        if (lane >= 8 && lane - 8 < rc - done) dst[done + lane - 8] = s_keys[rs + done + lane - 8];
      }
      for (uint32_t q = j * kWarp + lane; q < body; q += wpr * kWarp) {
        const uint32_t p = head + 4 * q;
        uint4 v;
        v.x = s_keys[rs + p], v.y = s_keys[rs + p + 1], v.z = s_keys[rs + p + 2], v.w = s_keys[rs + p + 3];
        *reinterpret_cast<uint4*>(dst + p) = v;
      }
    }
    __syncthreads();  // s_keys, s_start, s_dst are rewritten by the next tile
  }
  const uint32_t f = (__any_sync(kFull, bad_label != 0) ? TE_RANK_FLAG_BAD_LABEL : 0) |
                     (__any_sync(kFull, neg_score != 0) ? TE_RANK_FLAG_NEGATIVE_SCORE : 0) |
                     (__any_sync(kFull, wide != 0) ? TE_RANK_FLAG_WIDE_SCORE : 0) |
                     (__any_sync(kFull, overflow != 0) ? kRankFlagSlabOverflow : 0);
  if (lane == 0 && f) atomicOr(flags, f);
}

// After the scatter: this sender's fill count of its slab in every receiver's inbox, and its flags,
// go into the receivers' headers (header slot of sender q: {count, flags} at hdr + 2 q).
struct PeerHeaders {
  unsigned long long hdr[kMaxSplitters + 1];  // address of MY slot in rank b's header
};
__global__ void publish_counts_kernel(const uint32_t* __restrict__ cursors, const uint32_t* __restrict__ flags,
                                      const PeerHeaders peers, int world, uint32_t cap) {
  const int b = threadIdx.x;
  if (b >= world) return;
  uint32_t* slot = reinterpret_cast<uint32_t*>(peers.hdr[b]);
  slot[0] = min(cursors[b], cap);
  slot[1] = *flags;
}

// Receive side, first kernel: n = sum of the slab counts, flags = OR of the senders' flags.
__global__ void slab_totals_kernel(const uint32_t* __restrict__ hdr, int world, uint32_t* n_out, uint32_t* flags_out) {
  if (threadIdx.x != 0) return;
  uint32_t n = 0, f = 0;
  for (int q = 0; q < world; ++q) n += hdr[2 * q], f |= hdr[2 * q + 1];
  *n_out = n;
  atomicOr(flags_out, f);
}

// Histograms of the two partition digits over a slab view: the upsweep of the first pass when the
// input is an inbox.  Range r = tiles r, r + R, r + 2R, ... (as partition_ranges_kernel walks them);
// `subs` CTAs share a range (every subs-th of its tiles each) and add into range_hist (zeroed).
template <int TILE>
__global__ void __launch_bounds__(kIrBlock)
    rank_hist_view_kernel(const uint32_t* __restrict__ keys, const SlabView view, uint32_t ranges, uint32_t subs,
                          uint32_t* ghist, uint32_t* range_hist, uint32_t* flags) {
  extern __shared__ uint32_t sh[];  // [2][256][32 lanes]
  const unsigned tid = threadIdx.x, lane = tid & 31;
  for (int i = tid; i < 2 * kRadix * kWarp; i += kIrBlock) sh[i] = 0;
  __syncthreads();
  const uint32_t r = blockIdx.x / subs, k = blockIdx.x % subs;
  const uint32_t ntiles = view.nslabs * (view.cap / TILE);
  uint32_t wide = 0;
  for (uint64_t tile = r + (uint64_t)k * ranges; tile < ntiles; tile += (uint64_t)subs * ranges) {
    const size_t base = (size_t)tile * TILE;
    const uint32_t valid = view_valid_tile(view, (uint32_t)tile, TILE, 0);
    for (uint32_t i0 = 4 * tid; i0 < valid; i0 += 4 * kIrBlock) {
      uint32_t kk[4];
      const uint32_t cnt = min(4u, valid - i0);
      if (cnt == 4) {
        const int4 a = ld_stream_v4(keys + base + i0);
        kk[0] = (uint32_t)a.x, kk[1] = (uint32_t)a.y, kk[2] = (uint32_t)a.z, kk[3] = (uint32_t)a.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) kk[j] = j < (int)cnt ? keys[base + i0 + j] : 0u;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j < (int)cnt) {
          wide |= kk[j] >> 31;
          atomicAdd(sh + ((kk[j] >> kLeafBits) & 0xffu) * kWarp + lane, 1u);
          atomicAdd(sh + (kRadix + ((kk[j] >> (kLeafBits + 8)) & 0xffu)) * kWarp + lane, 1u);
        }
      }
    }
  }
  __syncthreads();
  {
    uint32_t s = 0;  // thread t owns (digit position, digit) = t
#pragma unroll
    for (int q = 0; q < kWarp; ++q) s += sh[tid * kWarp + ((q + tid) & 31)];
    if (s) {
      if (tid < kRadix) atomicAdd(range_hist + (size_t)r * kRadix + tid, s);
      atomicAdd(ghist + tid, s);
    }
  }
  if (__any_sync(kFull, wide != 0) && lane == 0) atomicOr(flags, (uint32_t)TE_RANK_FLAG_WIDE_SCORE);
}


// ---- measurement only: what one GPU can push into its peers' memory -------------------------------
// Every CTA stores 8 KB chunks of a constant shared-memory buffer, chunk c to destination c % ndst.
// mode 0: one 16-byte st.global per lane (what pack_scatter_kernel does); mode 1: one bulk async
// copy shared -> global per chunk, issued by one thread (plain cp.async.bulk store, no multicast).
__global__ void __launch_bounds__(512, 2)
    p2p_store_bench_kernel(const PeerSlabs dst, int ndst, uint32_t words_per_dst, int mode) {
  constexpr uint32_t kChunk = 2048;  // words
  __shared__ __align__(128) uint32_t s_buf[kChunk];
  const unsigned tid = threadIdx.x;
  for (uint32_t i = tid; i < kChunk; i += blockDim.x) s_buf[i] = i;
  __syncthreads();
  if (mode == 1 && tid == 0) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const uint64_t total = (uint64_t)(words_per_dst / kChunk) * (uint64_t)ndst;
  const uint32_t s_addr = (uint32_t)__cvta_generic_to_shared(s_buf);
  for (uint64_t c = blockIdx.x; c < total; c += gridDim.x) {
    const int d = (int)(c % (uint64_t)ndst);
    uint32_t* p = reinterpret_cast<uint32_t*>(dst.slab[d]) + (size_t)(c / (uint64_t)ndst) * kChunk;
    if (mode == 0) {
      const uint4 v = *reinterpret_cast<const uint4*>(&s_buf[tid * 4]);
      *reinterpret_cast<uint4*>(p + tid * 4) = v;
    } else if (tid == 0) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p), "r"(s_addr), "r"(kChunk * 4u)
                   : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 8;" ::: "memory");  // (the source never changes: a bound on what is in flight)
    }
  }
  if (mode == 1 && tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace te
