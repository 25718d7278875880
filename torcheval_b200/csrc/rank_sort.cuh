// Building blocks of the exact ranking path (K1/K2): single-sweep LSD radix sort.
//
// One pass = one kernel: every CTA takes a tile of keys (ticket order), ranks them by the
// current 8-bit digit with warp-private shared-memory digit counters, learns where its digit
// runs start in the output through a decoupled look-back chain over the preceding tiles
// (one chain per digit), and writes the tile out through shared memory so that every digit run
// leaves as a contiguous burst.  The four digit histograms are produced up front by the ingest
// kernel, so each pass reads and writes every record exactly once (2*r bytes per record, the
// "bytes moved per sort pass" of BASELINE.json).
#pragma once
#include "common.cuh"

namespace te {

constexpr int kRadixBits = 8;
constexpr int kRadix = 1 << kRadixBits;

constexpr int kSortIptDefault = 16;
constexpr int kSortBlockDefault = 512;  // 8192 records per CTA; 256 (4096 records) is the alternative

// Look-back word of the sort passes (one per tile and digit), written at most twice:
//   0                     : tile has not published yet
//   v      (bit31 == 0)   : v - 1 = this tile's own count for the digit
//   v      (bit31 == 1)   : v & 0x7fffffff = count over this tile and every tile before it
// A word carries its own validity, so relaxed loads/stores are enough.  Limits n to 2^31 - 1.
constexpr uint32_t kLbInclusive = 0x80000000u;
constexpr int kLbWindowDefault = 8;   // predecessors fetched per look-back round trip

// What a pass sorts by: an 8-bit digit of the key ...
struct RadixDigit {
  int shift;
  template <typename KeyT>
  __device__ __forceinline__ uint32_t operator()(KeyT k) const {
    return (uint32_t)(k >> shift) & (kRadix - 1);
  }
};
// ... or the index of the range the key falls in, given up to 15 ascending splitters
// (bucket = #{splitters <= key}): the partition step of the distributed sample sort.
constexpr int kMaxSplitters = 15;
struct SplitterDigit {
  uint32_t sp[kMaxSplitters];
  int nsp;
  __device__ __forceinline__ uint32_t operator()(uint32_t k) const {
    uint32_t b = 0;
#pragma unroll
    for (int i = 0; i < kMaxSplitters; ++i) b += (i < nsp && sp[i] <= k) ? 1u : 0u;
    return b;
  }
};

template <typename KeyT, int kSortBlock, int kSortIpt = kSortIptDefault>
constexpr size_t onesweep_smem_bytes() {
  static_assert(sizeof(KeyT) * kSortBlock * kSortIpt >= sizeof(uint32_t) * (kSortBlock / kWarp) * kRadix,
                "the peer-mask tables alias the key staging area");
  return sizeof(KeyT) * kSortBlock * kSortIpt +
         sizeof(uint32_t) * ((kSortBlock / kWarp) * kRadix + kRadix + kRadix / kWarp + 4);
}

// kLbWindow > 0: every tile walks back over its predecessors' counts, kLbWindow at a time.
// kLbWindow == 0: "scan service" — the CTA that draws ticket 0 does no sorting; its 256 threads
// (one per digit) stream over the tiles in order, kServiceBatch tiles per L2 round trip, turning
// published tile counts (lookback[t][d]) into exclusive prefixes (prefix[t][d]).  A tile then
// reads exactly one word per digit instead of walking a chain whose length grows with the tile
// rate (tens of millions of tiles per second on B200) times the L2 latency.
constexpr int kServiceBatch = 32;

template <typename KeyT, bool HAS_VAL, int kLbWindow, int kSortBlock, typename DigitOp,
          int kSortIpt = kSortIptDefault>
__global__ void __launch_bounds__(kSortBlock, (kSortIpt <= 12 ? 1536 : 1024) / kSortBlock)
    onesweep_pass_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                         const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                         uint32_t n, const DigitOp dop, const uint32_t* __restrict__ digit_base,
                         uint32_t* lookback, uint32_t* prefix, uint32_t* ticket) {
  constexpr bool kService = kLbWindow == 0;
  static_assert(kSortBlock >= kRadix && kSortBlock % kRadix == 0, "one thread per digit is needed");
  constexpr int WARPS = kSortBlock / kWarp;
  constexpr int kSortTile = kSortBlock * kSortIpt;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw);
  uint32_t(*s_whist)[kRadix] = reinterpret_cast<uint32_t(*)[kRadix]>(s_keys + kSortTile);
  uint32_t* s_goff = &s_whist[WARPS][0];
  uint32_t* s_wsum = s_goff + kRadix;
  uint32_t& s_tile = s_wsum[kRadix / kWarp];

  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  for (int i = tid; i < WARPS * kRadix; i += kSortBlock) {
    (&s_whist[0][0])[i] = 0;
    reinterpret_cast<uint32_t*>(s_keys)[i] = 0;  // per-warp peer masks, see the ranking loop
  }
  __syncthreads();
  if (kService && s_tile == 0) {
    // ---- scan service (ticket 0 is always drawn by a resident CTA, and no tile ever waits
    //      before publishing its counts, so this loop cannot starve) ----
    if (tid >= kRadix) return;
    const uint32_t tiles = (n + kSortTile - 1) / kSortTile;
    uint32_t run = 0;  // keys with digit `tid` in all tiles before the current one
    for (uint32_t t0 = 0; t0 < tiles; t0 += kServiceBatch) {
      uint32_t a[kServiceBatch];
#pragma unroll
      for (int j = 0; j < kServiceBatch; ++j)
        a[j] = (t0 + j < tiles) ? ld_relaxed_u32(lookback + (size_t)(t0 + j) * kRadix + tid) : 1u;
#pragma unroll
      for (int j = 0; j < kServiceBatch; ++j) {
        if (t0 + j < tiles) {
          while (a[j] == 0) a[j] = ld_relaxed_u32(lookback + (size_t)(t0 + j) * kRadix + tid);
          st_relaxed_u32(prefix + (size_t)(t0 + j) * kRadix + tid, kLbInclusive | run);
          run += a[j] - 1u;
        }
      }
    }
    return;
  }
  const uint32_t tile = kService ? s_tile - 1u : s_tile;
  const uint32_t base = tile * (uint32_t)kSortTile;
  const uint32_t valid = min((uint32_t)kSortTile, n - base);
  const uint32_t wbase = warp * (kWarp * kSortIpt);  // tile-relative start of this warp's chunk

  // ---- load (warp-striped: slot i of lane l is record wbase + i*32 + l, so (i, l) order is
  //      array order and the ranking below is stable) ----
  KeyT key[kSortIpt];
#pragma unroll
  for (int i = 0; i < kSortIpt; ++i) {
    const uint32_t r = wbase + i * kWarp + lane;
    key[i] = r < valid ? keys_in[base + r] : ~(KeyT)0;  // padding sorts last, is never written
  }

  // ---- rank inside the warp ----
  // Peer discovery goes through shared memory instead of eight warp votes (measured on B200:
  // the ballot loop alone caps a pass at ~400 G keys/s, shared atomics run at >5 lanes/clk/SM):
  // every lane ORs its lane bit into a warp-private word per digit, reads the word back (= the
  // set of lanes holding the same digit), and the lowest of them bumps the warp's digit counter.
  uint32_t* wh = s_whist[warp];
  uint32_t* wm = reinterpret_cast<uint32_t*>(s_keys) + warp * kRadix;  // zeroed above, unused yet
  const uint32_t lane_bit = 1u << lane, lt = lanemask_lt();
  static_assert(kSortIpt % 4 == 0, "ranks and digits are packed 2 and 4 per register");
  uint32_t rk[kSortIpt / 2];  // two 16-bit ranks per register
#pragma unroll
  for (int i = 0; i < kSortIpt; ++i) {
    const uint32_t d = dop(key[i]);
    atomicOr(wm + d, lane_bit);
    __syncwarp();
    const uint32_t peers = wm[d];
    const uint32_t first = wh[d];
    __syncwarp();
    const uint32_t below = __popc(peers & lt);
    if (below == 0) {  // lowest lane of the group
      wh[d] = first + __popc(peers);
      wm[d] = 0;
    }
    __syncwarp();
    const uint32_t r = first + below;  // < 512
    if (i & 1)
      rk[i >> 1] |= r << 16;
    else
      rk[i >> 1] = r;
  }
  __syncthreads();

  // ---- per-digit: tile count, publish, local layout ----
  uint32_t my_cnt = 0, my_bin_start = 0;
  if (tid < kRadix) {
    const uint32_t d = tid;
    uint32_t cnt = 0;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) cnt += s_whist[w][d];
    st_relaxed_u32(lookback + (size_t)tile * kRadix + d, cnt + 1u);

    // exclusive scan of cnt over the 256 digits (8 warps, named barrier 1)
    uint32_t incl = cnt;
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
    const uint32_t bin_start = wb + incl - cnt;

    // warp-exclusive offsets inside the digit run, shifted to tile-local positions
    uint32_t run = bin_start;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = s_whist[w][d];
      s_whist[w][d] = run;
      run += c;
    }
    my_cnt = cnt;
    my_bin_start = bin_start;
  }
  __syncthreads();

  // ---- decoupled look-back over the preceding tiles (digit threads = warps 0..7), overlapped
  //      with the other warps already scattering their keys into tile order.  A B200 pass runs
  //      at tens of millions of tiles per second while one L2 round trip is ~0.5 us, so a tile
  //      routinely has to add up dozens of not-yet-prefixed predecessors: fetch them kLbWindow at
  //      a time with independent loads instead of one dependent load per predecessor. ----
  if (!kService && tid < kRadix) {
    const uint32_t d = tid;
    uint32_t excl = 0;
    int64_t t = (int64_t)tile - 1;
    constexpr int W = kLbWindow > 0 ? kLbWindow : 1;
    while (t >= 0) {
      uint32_t w[W];
#pragma unroll
      for (int j = 0; j < W; ++j)
        w[j] = (t - j >= 0) ? ld_relaxed_u32(lookback + (size_t)(t - j) * kRadix + d) : kLbInclusive;
      int used = 0;
      bool done = false;
#pragma unroll
      for (int j = 0; j < W; ++j) {
        if (!done && used == j) {
          if (w[j] & kLbInclusive) {
            excl += w[j] & ~kLbInclusive;
            done = true;
          } else if (w[j] != 0) {
            excl += w[j] - 1u;
            used = j + 1;
          }
        }
      }
      if (done) break;
      t -= used;  // the first unpublished predecessor is polled again
    }
    st_relaxed_u32(lookback + (size_t)tile * kRadix + d, kLbInclusive | (excl + my_cnt));
    s_goff[d] = digit_base[d] + excl - my_bin_start;  // output index = s_goff[d] + tile position
  }

  // ---- scatter into tile order through shared memory ----
#pragma unroll
  for (int i = 0; i < kSortIpt; ++i) {
    const uint32_t d = dop(key[i]);
    const uint32_t r = (i & 1) ? (rk[i >> 1] >> 16) : (rk[i >> 1] & 0xffffu);
    const uint32_t p = wh[d] + r;
    s_keys[p] = key[i];
    if (HAS_VAL) {
      if (i & 1)
        rk[i >> 1] = (rk[i >> 1] & 0xffffu) | (p << 16);
      else
        rk[i >> 1] = (rk[i >> 1] & 0xffff0000u) | p;
    }
  }
  if (kService && tid < kRadix) {
    // by now the service has usually been past this tile for a while: one load per digit
    const uint32_t* pp = prefix + (size_t)tile * kRadix + tid;
    uint32_t w;
    do {
      w = ld_relaxed_u32(pp);
    } while (w == 0);
    s_goff[tid] = digit_base[tid] + (w & ~kLbInclusive) - my_bin_start;
  }
  __syncthreads();

  uint32_t dpack[kSortIpt / 4];  // digit of tile position tid + k*BLOCK, four per register
#pragma unroll
  for (int k = 0; k < kSortIpt; ++k) {
    const uint32_t p = tid + k * kSortBlock;
    const KeyT kk = s_keys[p];
    const uint32_t d = dop(kk);
    if (p < valid) keys_out[s_goff[d] + p] = kk;
    if (HAS_VAL) {
      if ((k & 3) == 0)
        dpack[k >> 2] = d;
      else
        dpack[k >> 2] |= d << (8 * (k & 3));
    }
  }

  if (HAS_VAL) {
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys);
    uint32_t val[kSortIpt];
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) {
      const uint32_t r = wbase + i * kWarp + lane;
      val[i] = r < valid ? vals_in[base + r] : 0u;
    }
    __syncthreads();  // everybody is done reading s_keys
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) {
      const uint32_t p = (i & 1) ? (rk[i >> 1] >> 16) : (rk[i >> 1] & 0xffffu);
      s_vals[p] = val[i];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < kSortIpt; ++k) {
      const uint32_t p = tid + k * kSortBlock;
      const uint32_t d = (dpack[k >> 2] >> (8 * (k & 3))) & 0xffu;
      if (p < valid) vals_out[s_goff[d] + p] = s_vals[p];
    }
  }
}

// Exclusive scan of the four 256-bin digit histograms, in place (one warp-scan chain per pass).
__global__ void __launch_bounds__(1024) digit_scan_kernel(uint32_t* hist, int passes) {
  __shared__ uint32_t ws[32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int pass = tid >> 8;
  uint32_t v = pass < passes ? hist[tid] : 0u;
  uint32_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(kFull, incl, o);
    if (lane >= (unsigned)o) incl += u;
  }
  if (lane == 31) ws[warp] = incl;
  __syncthreads();
  uint32_t wb = 0;
  const unsigned w0 = (unsigned)pass * 8;  // first warp of this pass
  for (unsigned w = w0; w < warp; ++w) wb += ws[w];
  if (pass < passes) hist[tid] = wb + incl - v;
}

}  // namespace te
