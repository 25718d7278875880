// Building blocks of the exact ranking path (K1/K2): single-sweep LSD radix sort.
//
// One pass = one kernel: every CTA takes a tile of keys (tile = block index), ranks them by the
// current 8-bit digit with half-warp-private shared-memory tables (peer mask + running count in
// one word), learns where its digit runs start in the output through a two-level decoupled
// look-back over the preceding tiles, and writes the tile out through shared memory so that
// every digit run leaves as a contiguous burst.  grid.y batches independent sorts of equal length
// (one-vs-rest classes); the same kernel with "destination rank" as the digit is the partition
// pass of the distributed sample sort (optionally storing straight into peer GPUs' memory).  The four digit histograms are produced up front by the ingest
// kernel, so each pass reads and writes every record exactly once (2*r bytes per record, the
// "bytes moved per sort pass" of BASELINE.json).
#pragma once
#include <type_traits>

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

// What a pass sorts by: an 8-bit digit of the key ...
struct RadixDigit {
  static constexpr bool kCached = false;  // one shift + mask: recomputed wherever it is needed
  int shift;
  template <typename KeyT>
  __device__ __forceinline__ uint32_t operator()(KeyT k) const {
    return (uint32_t)(k >> shift) & (kRadix - 1);
  }
  __device__ __forceinline__ void stage(uint32_t*, unsigned) const {}
  template <typename KeyT>
  __device__ __forceinline__ uint32_t eval(KeyT k, const uint32_t*) const {
    return (*this)(k);
  }
};
// ... or the index of the range the key falls in, given up to 15 ascending splitters
// (bucket = #{splitters <= key}): the partition step of the distributed sample sort.
constexpr int kMaxSplitters = 15;
// Per-bucket output bases of a partition pass that scatters straight into other GPUs' memory:
// p[b] is where bucket b's run starts — an address in rank b's peer-mapped inbox, already
// shifted so that indexing with the pass's own output index lands in this rank's slot there.
struct PeerPtrs {
  unsigned long long p[kMaxSplitters + 1];
  int vec;  // 1: leave each run as 16-byte stores (head / body / tail), 0: one 4-byte store per record
};
struct SplitterDigit {
  // a compare per splitter against an array that lives in local memory: the pass evaluates it
  // once per record and keeps the 4-bit result in registers
  static constexpr bool kCached = true;
  uint32_t sp[kMaxSplitters];
  int nsp;
  __device__ __forceinline__ uint32_t operator()(uint32_t k) const {
    uint32_t b = 0;
#pragma unroll 1
    for (int i = 0; i < nsp; ++i) b += sp[i] <= k ? 1u : 0u;  // nsp = ranks - 1, usually 1..7
    return b;
  }
  // The pass kernel stages the splitters in shared memory (padded to 16 words with 0xffffffff) and
  // finds the bucket with a branch-free binary search: log2 steps, every step one shared load that
  // is conflict-free (<= 15 consecutive words, equal addresses broadcast).  The linear scan above
  // made the partition pass 2-4x slower than a radix pass at 8-16 ranks
  // (profiles/r01_part_bench.txt).
  __device__ __forceinline__ void stage(uint32_t* s, unsigned tid) const {
    if (tid < 16) s[tid] = (int)tid < nsp ? sp[tid] : 0xffffffffu;
  }
  __device__ __forceinline__ uint32_t eval(uint32_t k, const uint32_t* s) const {
    if (nsp == 1) return s[0] <= k ? 1u : 0u;  // two ranks: one compare (the staged copy: the
                                               // splitters may have come from device memory)
    uint32_t b = 0;
#pragma unroll
    for (int step = 8; step >= 1; step >>= 1) {
      if (step <= nsp) {  // uniform: 1 step for 2 ranks, 3 for 8, 4 for 16
        if (s[b + step - 1] <= k) b += step;
      }
    }
    return min(b, (uint32_t)nsp);  // (a padding word compares equal only to the key 0xffffffff)
  }
};

// 0/1 label of a target element; `bad` collects "not exactly 0 or 1".
template <typename T>
struct LabelOf;
template <>
struct LabelOf<long long> {
  static __device__ __forceinline__ uint32_t get(long long t, uint32_t& bad) {
    bad |= ((unsigned long long)t > 1ull);
    return (uint32_t)t & 1u;
  }
};
template <>
struct LabelOf<int> {
  static __device__ __forceinline__ uint32_t get(int t, uint32_t& bad) {
    bad |= ((unsigned)t > 1u);
    return (uint32_t)t & 1u;
  }
};
template <>
struct LabelOf<unsigned char> {
  static __device__ __forceinline__ uint32_t get(unsigned char t, uint32_t& bad) {
    bad |= (t > 1);
    return t & 1u;
  }
};
template <>
struct LabelOf<float> {
  static __device__ __forceinline__ uint32_t get(float t, uint32_t& bad) {
    bad |= !(t == 0.f || t == 1.f);
    return t == 1.f ? 1u : 0u;
  }
};
template <>
struct LabelOf<double> {
  static __device__ __forceinline__ uint32_t get(double t, uint32_t& bad) {
    bad |= !(t == 0.0 || t == 1.0);
    return t == 1.0 ? 1u : 0u;
  }
};

// Optional per-phase cycle accounting of the pass kernel (compile with -DTE_PHASE_STATS; read with
// te_debug_phase_stats).  Thread 0 of every CTA adds the clock64() deltas between phase boundaries.
__device__ unsigned long long g_phase_cycles[8];
#ifdef TE_PHASE_STATS
#define TE_PHASE_MARK(idx)                                              \
  do {                                                                  \
    if (threadIdx.x == 0) {                                             \
      const long long _now = clock64();                                 \
      atomicAdd(&g_phase_cycles[idx], (unsigned long long)(_now - _t)); \
      _t = _now;                                                        \
    }                                                                   \
  } while (0)
#define TE_PHASE_BEGIN() long long _t = clock64()
#else
#define TE_PHASE_MARK(idx) do { } while (0)
#define TE_PHASE_BEGIN() do { } while (0)
#endif

// ---- one LSD pass -----------------------------------------------------------------------------
// Work decomposition inside a CTA: the tile is cut into GROUPS of 16 * kSortIpt consecutive
// records, one per half-warp.  Slot i of sub-lane s in group g is record g*16*IPT + i*16 + s, so
// (i, s) order is array order and ranking round by round, lane by lane, is stable.
//
// The pass is bound by the L1/shared-memory data pipe (ncu: l1tex__data_pipe_lsu_wavefronts at
// ~80 % of peak with the first versions, ALU at 30 %), and every shared access indexed by a random
// digit costs ~3.4 wavefronts in bank conflicts.  Ranking with half-warp groups needs a 16-bit
// peer mask, which leaves room for the group's running digit count in the same 32-bit word:
// three 32-bit shared instructions per round (ATOMS.OR, LDS, STS) instead of five.
template <typename KeyT, int kSortBlock, int kSortIpt = kSortIptDefault>
constexpr size_t onesweep_stage_bytes() {
  // the key staging area doubles as the per-group {count << 16 | peer mask} tables of the ranking
  // loop, which are dead by the time keys are staged
  constexpr size_t keys = sizeof(KeyT) * kSortBlock * kSortIpt;
  constexpr size_t tables = sizeof(uint32_t) * (kSortBlock / 16) * kRadix;
  return keys > tables ? keys : tables;
}
template <typename KeyT, int kSortBlock, int kSortIpt = kSortIptDefault>
constexpr size_t onesweep_smem_bytes() {
  return onesweep_stage_bytes<KeyT, kSortBlock, kSortIpt>() +
         sizeof(uint32_t) * ((kSortBlock / 16) * kRadix + 2 * kRadix + kRadix / kWarp + 4 + 2 * (kMaxSplitters + 1) + 16 + 2 * (kMaxSplitters + 1));
}

// Cross-tile offsets: two-level decoupled look-back.  With ~30-60 million tiles per second per
// pass and an L2 round trip of ~500 cycles, a tile that walks back over its predecessors until
// it meets a finished one ends up adding ~100 of them, and the pass settles at exactly
// (window / round trip) tiles per cycle (profiles/r01_phase_stats.txt: look-back grew from 25 %
// to 33 % of the tile time as the other phases were shortened).  So tiles are grouped by
// kLbGroup = 32: a tile needs (a) the counts of the earlier tiles of its own group — independent
// loads, one round trip — and (b) ONE word per digit, the prefix over all earlier groups, which
// the last tile of the previous group publishes.  Only those last tiles form a chain, and it
// advances 32 tiles per round trip.  Tile order = block order (see below), so nothing waits on
// a tile that has not been dispatched.
// Batched launches (one-vs-rest: several independent sorts of the same length in one grid):
// blockIdx.y selects the problem, every scratch-resident pointer moves by blockIdx.y * bstride bytes.
template <typename T>
__device__ __forceinline__ T* batch_ptr(T* p, size_t bstride) {
  return p ? reinterpret_cast<T*>(reinterpret_cast<char*>(const_cast<typename std::remove_const<T>::type*>(p)) +
                                  (size_t)blockIdx.y * bstride)
           : p;
}

constexpr int kLbGroup = 32;
constexpr uint32_t kNoHotDigit = 0xffffffffu;

//
// kUnstable: ranks inside a round are handed out in arrival order (one ATOMS.ADD with return
// instead of ATOMS.OR + LDS + STS).  Only valid where the relative order of records with equal
// digits carries no information: the FIRST pass of a keys-only sort (records that agree in all
// 32 bits are indistinguishable, and later passes are stable), never for (key, value) pairs.
template <typename KeyT, bool HAS_VAL, int kSortBlock, typename DigitOp, int kSortIpt = kSortIptDefault,
          bool kPeerScatter = false, bool kUnstable = false>
__global__ void __launch_bounds__(kSortBlock, (kSortIpt <= 12 ? 1536 : 1024) / kSortBlock)
    onesweep_pass_kernel(const KeyT* __restrict__ keys_in_, KeyT* __restrict__ keys_out_,
                         const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                         uint32_t n, const DigitOp dop, const uint32_t* __restrict__ digit_base_,
                         uint32_t* lookback_, uint32_t* group_prefix_, const PeerPtrs peers,
                         const uint32_t* __restrict__ hot_info_, uint32_t* __restrict__ zero_next_,
                         size_t bstride, const uint32_t* __restrict__ skip_flag_ = nullptr,
                         const uint32_t* __restrict__ n_dev = nullptr) {
  // n_dev: the record count lives in device memory (the receive side of the sample sort); the grid
  // was sized for an upper bound and the tiles past the end leave at once — nobody waits on them
  if (n_dev) {
    n = *n_dev;
    if ((uint64_t)blockIdx.x * (kSortBlock * kSortIpt) >= n) return;
  }
  const uint32_t* __restrict__ skip_flag = batch_ptr(skip_flag_, bstride);
  if (HAS_VAL) vals_in = batch_ptr(vals_in, bstride), vals_out = batch_ptr(vals_out, bstride);
  const KeyT* __restrict__ keys_in = batch_ptr(keys_in_, bstride);
  KeyT* __restrict__ keys_out = batch_ptr(keys_out_, bstride);
  const uint32_t* __restrict__ digit_base = batch_ptr(digit_base_, bstride);
  uint32_t* lookback = batch_ptr(lookback_, bstride);
  uint32_t* group_prefix = batch_ptr(group_prefix_, bstride);
  const uint32_t* __restrict__ hot_info = batch_ptr(hot_info_, bstride);
  uint32_t* __restrict__ zero_next = batch_ptr(zero_next_, bstride);
  static_assert(kSortBlock >= kRadix && kSortBlock % kRadix == 0, "one thread per digit is needed");
  static_assert(kSortIpt % 4 == 0 && 16 * kSortIpt <= 256, "ranks are packed four per register");
  constexpr int GROUPS = kSortBlock / 16;
  constexpr int kGroupKeys = 16 * kSortIpt;
  constexpr int kSortTile = kSortBlock * kSortIpt;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  KeyT* s_keys = reinterpret_cast<KeyT*>(smem_raw);
  uint32_t* s_rank = reinterpret_cast<uint32_t*>(smem_raw);  // [GROUPS][256]; aliases s_keys
  uint32_t(*s_base)[kRadix] = reinterpret_cast<uint32_t(*)[kRadix]>(
      smem_raw + onesweep_stage_bytes<KeyT, kSortBlock, kSortIpt>());  // [GROUPS][256]
  uint32_t* s_goff = &s_base[GROUPS][0];       // 2*256 words: also the slice starts of the digit phase
  uint32_t* s_wsum = s_goff + 2 * kRadix;
  uint32_t* s_runs = s_wsum + kRadix / kWarp + 4;  // peer scatter: tile-local start, count per bucket
  uint32_t* s_sp = s_runs + 2 * (kMaxSplitters + 1);  // splitters (partition passes)
  unsigned long long* s_peer = reinterpret_cast<unsigned long long*>(s_sp + 16);  // inbox bases

  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  TE_PHASE_BEGIN();
  // Tile order = block order.  The look-back below waits on lower-numbered tiles, which is safe
  // because thread blocks are dispatched in increasing blockIdx order: a lower-numbered tile is
  // always resident or finished (the same assumption CUB's decoupled look-back scan makes).
  // Not drawing a ticket keeps a ~700-cycle global atomic off the head of every tile.
  const uint32_t tile = blockIdx.x;
  const uint32_t base = tile * (uint32_t)kSortTile;
  const uint32_t valid = min((uint32_t)kSortTile, n - base);
  const uint32_t sub = lane & 15u;
  const uint32_t group = tid >> 4;
  const uint32_t gbase = group * kGroupKeys;  // tile-relative start of this half-warp's records

  // "this pass is a no-op" flag (see below): requested ahead of the key loads so that it is back
  // before they are, and looked at only after the set-up work
  const uint32_t skip_pass = skip_flag ? ld_relaxed_u32(skip_flag) : 0u;
  // ---- load: a warp instruction fetches two 64-byte runs, one per half-warp.  The loads are
  //      issued first; the ranking tables are cleared while they are in flight. ----
  KeyT key[kSortIpt];
  // tile-relative index of the record in slot i of this thread
  auto rec = [&](int i) -> uint32_t { return gbase + i * 16 + sub; };
#pragma unroll
  for (int i = 0; i < kSortIpt; ++i) {
    const uint32_t r = rec(i);
    key[i] = r < valid ? keys_in[base + r] : ~(KeyT)0;  // padding sorts last, is never written
  }
  for (int i = tid; i < GROUPS * kRadix; i += kSortBlock) s_rank[i] = 0u;
  // the next pass's look-back row of this tile must read as "not published": clearing it here
  // (1 KB per tile, the pass has HBM bandwidth to spare) replaces a memset of the whole array
  if (zero_next && tid < kRadix) zero_next[(size_t)tile * kRadix + tid] = 0u;
  dop.stage(s_sp, tid);
  if (kPeerScatter && tid <= (unsigned)kMaxSplitters) s_peer[tid] = peers.p[tid];
  __syncthreads();
  // ---- a digit that is the same for every key (digit_scan_kernel found one full bin: low bits of
  //      bf16-valued scores, the exponent byte of one binade, constant scores) leaves the order
  //      as it is: the tile is copied through, no ranking, no look-back ----
  if (skip_pass) {
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) {
      const uint32_t r = rec(i);
      if (r < valid) {
        keys_out[base + r] = key[i];
        if (HAS_VAL) vals_out[base + r] = vals_in[base + r];
      }
    }
    return;
  }
  TE_PHASE_MARK(0);  // setup
#ifdef TE_PHASE_STATS
  {  // force the loads to complete here so that the load phase is visible on its own
    KeyT acc = 0;
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) acc |= key[i];
    if (acc == (KeyT)0x12345677) s_goff[0] = 1;
  }
  TE_PHASE_MARK(1);  // load (thread 0's own loads)
#endif

  // ---- rank inside the group: ranks are group-relative (< 256), four per register ----
  // Odd groups index their table with bit 4 of the digit flipped: the two half-warps of a warp
  // then hit different banks for the same digit, which matters when one digit dominates.
  const uint32_t rot = (group & 1u) << 4;
  uint32_t* gt = s_rank + group * kRadix;
  const uint32_t sub_bit = 1u << sub, sub_lt = sub_bit - 1u;
  uint32_t rk[kSortIpt / 4];
  // Skewed digits (constant exponent bits after a range partition, label-only low digits of
  // half-precision scores, two destination ranks): same-address shared atomics serialise, up to 16
  // deep.  digit_scan_kernel names up to two digits that hold >= 1/8 of the keys of this pass;
  // records with those digits are ranked with a ballot and a register counter instead.
  const uint32_t hot0 = hot_info ? hot_info[0] : kNoHotDigit;
  const uint32_t hot1 = hot_info ? hot_info[1] : kNoHotDigit;
  uint32_t dcache[DigitOp::kCached ? kSortIpt / 8 : 1];  // digits < 16, eight per register
  auto remember = [&](int i, uint32_t dd) {
    if (DigitOp::kCached) {
      if ((i & 7) == 0)
        dcache[i >> 3] = dd;
      else
        dcache[i >> 3] |= dd << (4 * (i & 7));
    }
  };
  auto digit_of = [&](int i) -> uint32_t {
    return DigitOp::kCached ? (dcache[i >> 3] >> (4 * (i & 7))) & 15u : dop.eval(key[i], s_sp);
  };
  if (hot0 == kNoHotDigit) {
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) {
      const uint32_t dd0 = dop.eval(key[i], s_sp);
      remember(i, dd0);
      const uint32_t d = dd0 ^ rot;
      uint32_t r;
      if (kUnstable) {
        static_assert(!(kUnstable && HAS_VAL), "payloads need a stable pass");
        // padding stays out of the arrival-order atomic: which of the lanes that hit one address
        // in the same instruction is served first is unspecified, and a padding record served
        // before a real one with the top digit would push the real one past `valid`
        r = rec(i) < valid ? atomicAdd(gt + d, 1u << 16) >> 16 : 0u;
      } else {
        atomicOr(gt + d, sub_bit);
        __syncwarp();
        const uint32_t e = gt[d];  // count << 16 | lanes of the group holding digit d this round
        __syncwarp();
        const uint32_t peers = e & 0xffffu, cnt = e >> 16;
        const uint32_t below = __popc(peers & sub_lt);
        if (below == 0) gt[d] = (cnt + __popc(peers)) << 16;  // lowest peer: bump count, clear mask
        __syncwarp();
        r = cnt + below;
      }
      if ((i & 3) == 0)
        rk[i >> 2] = r;
      else
        rk[i >> 2] |= r << (8 * (i & 3));
    }
  } else {
    uint32_t hc0 = 0, hc1 = 0;  // records of the group seen so far with digit hot0 / hot1
    const uint32_t hshift = lane & 16u;
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) {
      const uint32_t dd = dop.eval(key[i], s_sp);
      remember(i, dd);
      const bool live = !kUnstable || rec(i) < valid;  // (see the arrival-order atomic above)
      const bool h0 = live && dd == hot0, h1 = live && dd == hot1;
      const uint32_t b0 = (__ballot_sync(kFull, h0) >> hshift) & 0xffffu;
      const uint32_t b1 = (__ballot_sync(kFull, h1) >> hshift) & 0xffffu;
      const bool cold = !(h0 || h1);
      const uint32_t d = dd ^ rot;
      uint32_t r = 0;
      if (kUnstable) {
        if (cold && rec(i) < valid) r = atomicAdd(gt + d, 1u << 16) >> 16;
      } else {
        if (cold) atomicOr(gt + d, sub_bit);
        __syncwarp();
        const uint32_t e = cold ? gt[d] : 0u;
        __syncwarp();
        const uint32_t peers = e & 0xffffu, cnt = e >> 16;
        const uint32_t below = __popc(peers & sub_lt);
        if (cold && below == 0) gt[d] = (cnt + __popc(peers)) << 16;
        __syncwarp();
        r = cnt + below;
      }
      if (h0) r = hc0 + __popc(b0 & sub_lt);
      if (h1) r = hc1 + __popc(b1 & sub_lt);
      hc0 += __popc(b0);
      hc1 += __popc(b1);
      if ((i & 3) == 0)
        rk[i >> 2] = r;
      else
        rk[i >> 2] |= r << (8 * (i & 3));
    }
    if (sub == 0) {  // nobody touched the hot digits' table entries: they get their counts now
      gt[hot0 ^ rot] = hc0 << 16;
      if (hot1 != kNoHotDigit) gt[hot1 ^ rot] = hc1 << 16;
    }
  }
  TE_PHASE_MARK(2);  // rank (thread 0's warp)
  __syncthreads();
  TE_PHASE_MARK(3);  // wait for the other warps' load + rank

  // ---- per-digit: tile count, publish, local layout.  Thread (d, h) sums the groups of slice
  //      h = tid / 256 so that the 2*GROUPS shared-memory reads per digit are spread over the
  //      whole CTA; the digit threads (h == 0) combine the slices. ----
  constexpr int SLICES = kSortBlock / kRadix;
  constexpr int GPS = GROUPS / SLICES;  // groups per slice
  uint32_t* s_slice = s_goff;           // [SLICES][256] slice starts; s_goff proper is written later
  static_assert(SLICES <= 2, "s_goff has room for two slices");
  uint32_t my_cnt = 0, my_bin_start = 0;
  {
    const uint32_t d = tid & (kRadix - 1), h = tid >> 8;
    uint32_t part = 0;
    static_assert(GPS % 2 == 0, "group parity must follow the loop index");
#pragma unroll
    for (int g = 0; g < GPS; ++g) part += s_rank[(h * GPS + g) * kRadix + (d ^ ((g & 1) << 4))] >> 16;
    if (SLICES > 1 && h > 0) s_base[GROUPS - 1][d] = part;  // parked in the last row, rewritten below
    __syncthreads();
    uint32_t cnt = part;
    if (SLICES > 1 && h == 0) cnt += s_base[GROUPS - 1][d];
    uint32_t bin_start = 0;
    if (tid < kRadix) {
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
      bin_start = wb + incl - cnt;
      my_cnt = cnt;
      my_bin_start = bin_start;
      s_slice[d] = bin_start;                      // start of slice 0
      if (SLICES > 1) s_slice[kRadix + d] = bin_start + part;  // start of slice 1
    }
    __syncthreads();
    // group-exclusive offsets inside the digit run, shifted to tile-local positions
    uint32_t run = s_slice[h * kRadix + d];  // (s_slice overlays s_goff, which is written after the next barrier)
#pragma unroll
    for (int g = 0; g < GPS; ++g) {
      s_base[h * GPS + g][d] = run;
      run += s_rank[(h * GPS + g) * kRadix + (d ^ ((g & 1) << 4))] >> 16;
    }
  }
  __syncthreads();
  TE_PHASE_MARK(4);  // digit phase

  // ---- scatter into tile order through shared memory (needs only tile-local positions) ----
  const uint32_t* gb = s_base[group];
  uint32_t pos[HAS_VAL ? kSortIpt / 2 : 1];  // tile positions, two per register (pairs only)
#pragma unroll
  for (int i = 0; i < kSortIpt; ++i) {
    const uint32_t d = digit_of(i);
    const uint32_t r = (rk[i >> 2] >> (8 * (i & 3))) & 0xffu;
    const uint32_t p = gb[d] + r;
    if (!kUnstable || rec(i) < valid) s_keys[p] = key[i];  // (unranked padding has no slot)
    if (HAS_VAL) {
      if (i & 1)
        pos[i >> 1] |= p << 16;
      else
        pos[i >> 1] = p;
    }
  }
  TE_PHASE_MARK(5);  // scatter to smem

  // ---- two-level look-back: the whole CTA takes part, thread (d, h) fetches slice h of the
  //      in-group predecessors of digit d; the keys are in shared memory by now, so the loads
  //      have the registers to themselves ----
  {
    constexpr int JH = kLbGroup / SLICES;  // in-group predecessors per thread
    const uint32_t d = tid & (kRadix - 1), h = tid >> 8;
    const uint32_t grp = tile / kLbGroup, k = tile % kLbGroup;
    const uint32_t* arow = lookback + ((size_t)grp * kLbGroup + h * JH) * kRadix + d;
    const uint32_t* gp = group_prefix + (size_t)(grp - 1) * kRadix + d;  // used only if grp > 0
    uint32_t w[JH];
#pragma unroll
    for (int j = 0; j < JH; ++j) w[j] = (h * JH + j < k) ? ld_relaxed_u32(arow + (size_t)j * kRadix) : 1u;
    uint32_t gpre = (h == 0 && grp > 0) ? ld_relaxed_u32(gp) : kLbInclusive;
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < JH; ++j) {
      while (w[j] == 0) w[j] = ld_relaxed_u32(arow + (size_t)j * kRadix);
      sum += w[j] - 1u;
    }
    while (gpre == 0) gpre = ld_relaxed_u32(gp);
    sum += gpre & ~kLbInclusive;
    if (SLICES > 1 && h > 0) s_slice[kRadix + d] = sum;
    __syncthreads();  // also: every key is staged in s_keys
    if (tid < kRadix) {
      const uint32_t excl = sum + (SLICES > 1 ? s_slice[kRadix + d] : 0u);
      if (k == kLbGroup - 1)  // last tile of the group: extend the chain of group prefixes
        st_relaxed_u32(group_prefix + (size_t)grp * kRadix + d, kLbInclusive | (excl + my_cnt));
      s_goff[d] = digit_base[d] + excl - my_bin_start;  // output index = s_goff[d] + tile position
      if (kPeerScatter && d <= (uint32_t)kMaxSplitters) s_runs[d] = my_bin_start, s_runs[kMaxSplitters + 1 + d] = my_cnt;
    }
  }
  __syncthreads();
  TE_PHASE_MARK(6);  // look-back

  if constexpr (kPeerScatter && sizeof(KeyT) == 4 && !HAS_VAL) {
    // NVLink stores: a bucket's run is contiguous both in s_keys and in its inbox, so it leaves as
    // 16-byte stores (scalar head up to the first 16-byte boundary of the destination, scalar tail)
    if (peers.vec) {
      // the 16 warps are dealt out over the buckets (16 / buckets warps each); a warp walks its
      // share of the run with one 16-byte store per lane and step
      const uint32_t nb = (uint32_t)dop.nsp + 1u;
      constexpr uint32_t kWarps = kSortBlock / kWarp;
      const uint32_t wpr = nb < kWarps ? kWarps / nb : 1u;  // warps per run
      for (uint32_t d = warp / wpr; d < nb; d += kWarps / wpr) {
        const uint32_t j = warp % wpr;
        const uint32_t rs = s_runs[d];
        if (rs >= valid) break;  // runs are in tile order; padding sits at the very end
        const uint32_t rc = min(s_runs[kMaxSplitters + 1 + d], valid - rs);
        if (rc == 0) continue;
        uint32_t* dst = reinterpret_cast<uint32_t*>(s_peer[d]) + s_goff[d];  // dst[p], p = tile position
        const uint32_t mis = (uint32_t)((reinterpret_cast<uintptr_t>(dst + rs) >> 2) & 3u);
        const uint32_t head = min(rc, (4u - mis) & 3u);
        const uint32_t body = (rc - head) >> 2;
        const uint32_t done = head + 4 * body;
        if (j == 0) {  // scalar head and tail (<= 3 records each)
          if (lane < head) dst[rs + lane] = s_keys[rs + lane];
          if (lane >= 8 && lane - 8 < rc - done) dst[rs + done + lane - 8] = s_keys[rs + done + lane - 8];
        }
        for (uint32_t q = j * kWarp + lane; q < body; q += wpr * kWarp) {
          const uint32_t p = rs + head + 4 * q;
          uint4 v;
          v.x = s_keys[p], v.y = s_keys[p + 1], v.z = s_keys[p + 2], v.w = s_keys[p + 3];
          *reinterpret_cast<uint4*>(dst + p) = v;
        }
      }
      TE_PHASE_MARK(7);
      return;
    }
  }
  uint32_t dpack[kSortIpt / 4];  // digit of tile position tid + k*BLOCK, four per register
#pragma unroll
  for (int k = 0; k < kSortIpt; ++k) {
    const uint32_t p = tid + k * kSortBlock;
    const KeyT kk = s_keys[p];
    const uint32_t d = dop.eval(kk, s_sp);
    if (p < valid) {
      if (kPeerScatter)  // the digit is the destination rank: store over NVLink into its inbox
        reinterpret_cast<KeyT*>(s_peer[d])[s_goff[d] + p] = kk;
      else
        keys_out[s_goff[d] + p] = kk;
    }
    if (HAS_VAL) {
      if ((k & 3) == 0)
        dpack[k >> 2] = d;
      else
        dpack[k >> 2] |= d << (8 * (k & 3));
    }
  }
  TE_PHASE_MARK(7);  // output

  if (HAS_VAL) {
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys);
    uint32_t val[kSortIpt];
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) {
      const uint32_t r = gbase + i * 16 + sub;
      val[i] = r < valid ? vals_in[base + r] : 0u;
    }
    __syncthreads();  // everybody is done reading s_keys
#pragma unroll
    for (int i = 0; i < kSortIpt; ++i) {
      const uint32_t p = (i & 1) ? (pos[i >> 1] >> 16) : (pos[i >> 1] & 0xffffu);
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

// Exclusive scan of the four 256-bin digit histograms, in place (one warp-scan chain per pass),
// and the up-to-two "hot" digits of every pass: hot[2 * pass + {0, 1}] = digit holding at least
// 1/8 of the keys (largest first), kNoHotDigit otherwise.
__global__ void __launch_bounds__(1024) digit_scan_kernel(uint32_t* hist_, int passes, uint32_t* hot_,
                                                          size_t bstride, uint32_t* skip_ = nullptr) {
  uint32_t* hist = batch_ptr(hist_, bstride);
  uint32_t* hot = batch_ptr(hot_, bstride);
  uint32_t* skip = batch_ptr(skip_, bstride);  // skip[pass] = 1: one bin holds every key
  __shared__ uint32_t ws[32];
  __shared__ unsigned long long wmax[32];
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
  uint32_t wb = 0, total = 0;
  const unsigned w0 = (unsigned)pass * 8;  // first warp of this pass
  for (unsigned w = w0; w < w0 + 8; ++w) {
    if (w < warp) wb += ws[w];
    total += ws[w];
  }
  if (pass < passes) hist[tid] = wb + incl - v;
  if (skip && pass < passes && v == total && total > 0) skip[pass] = 1u;  // (zero-initialised by the caller)
  if (!hot) return;
  // top two bins of each pass: max-reduce (count << 8 | digit), twice
  unsigned long long mine = ((unsigned long long)v << 8) | (tid & 255u);
  unsigned long long first = 0;
  for (int round = 0; round < 2; ++round) {
    unsigned long long m = mine;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long u = __shfl_xor_sync(kFull, m, o);
      m = u > m ? u : m;
    }
    __syncthreads();
    if (lane == 0) wmax[warp] = m;
    __syncthreads();
    unsigned long long best = 0;
    for (unsigned w = w0; w < w0 + 8; ++w) best = wmax[w] > best ? wmax[w] : best;
    const uint32_t cnt = (uint32_t)(best >> 8), dig = (uint32_t)(best & 255u);
    const bool is_hot = pass < passes && total > 0 && cnt >= (total + 7) / 8 && (round == 0 || first != 0);
    if ((tid & 255u) == 0 && pass < passes) hot[2 * pass + round] = is_hot ? dig : kNoHotDigit;
    if (round == 0) {
      first = is_hot ? 1 : 0;
      if ((tid & 255u) == dig) mine = 0;  // leave the winner out of the second round
    }
  }
}

}  // namespace te
