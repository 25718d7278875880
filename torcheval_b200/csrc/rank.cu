// K1 — exact binary AUROC / AUPRC.
//
// Reference path: torcheval/metrics/functional/classification/auroc.py:115-152
// (_binary_auroc_compute_jit: sort(descending) + diff!=0 mask + gather + 2x cumsum +
// 2x masked_scatter_ + trapz) and precision_recall_curve.py:208-231 (_compute_for_each_class)
// + functional/tensor_utils.py:12-16 (_riemann_integral) for AUPRC.
//
// Here:  ingest (score -> order-preserving 32-bit key with the 0/1 label in bit 0, four digit
// histograms on the fly)  ->  4 single-sweep radix passes (rank_sort.cuh)  ->  ONE fused pass
// over the sorted keys that does the TP/FP running counts (block scan + decoupled look-back
// across tiles), the tie-group detection (`diff != 0` of the reference) and the trapezoid /
// step integration, leaving one partial per tile; a one-CTA reduce sums the partials in a fixed
// order.  AUROC is accumulated exactly in integers:
//     2*U = sum over positives of (#negatives with score <= s) + (#negatives with score < s)
//     AUROC = U / (P * N)          (identical to the reference's trapezoid sum over tie groups)
// so the result does not depend on launch geometry or timing.
//
// Bytes per score (SURVEY.md §8d formula, b_in = 12, r = 4, P = 4): 16 + 32 + 4 = 52.
#include <stdlib.h>

#include "rank_sort.cuh"
#include "rank_dense.cuh"
#include "rank_leaf.cuh"
#include "rank_part.cuh"
#include "rank_exchange.cuh"
#include "rank_msd.cuh"
#include "rank_leaf_ap.cuh"
#include "prof.cuh"

namespace te {
namespace {

// ------------------------------------------------------------------------------------------
// ingest: raw (score, target) -> keys (+ payload) and the four digit histograms
// ------------------------------------------------------------------------------------------
constexpr int kIngestBlock = 1024;
constexpr size_t kIngestSmem = (size_t)4 * kRadix * kWarp * sizeof(uint32_t);  // 128 KB

enum IngestMode { kPacked = 0, kIndexed = 1, kRawKeys = 2 };

struct RankScalars {            // lives in the zeroed part of the scratch buffer
  unsigned long long positives; // #label==1 (packed mode)
  uint32_t flags;               // TE_RANK_FLAG_*
  uint32_t pad;
};

// LEAF: the histograms the two-pass + counting-stage path needs (rank_leaf.cuh): digits at bits
// 15..22 and 23..30 instead of the four bytes, and a flag if any key uses bit 31 (score >= 2.0).
template <int MODE, typename TargetT, bool LEAF = false>
__global__ void __launch_bounds__(kIngestBlock, 1)
    rank_ingest_kernel(const float* __restrict__ score, const TargetT* __restrict__ target,
                       const uint32_t* __restrict__ raw_keys, uint32_t n, bool vec_ok,
                       uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                       uint32_t* ghist, RankScalars* scal, const uint32_t* __restrict__ payload_bad,
                       const double* __restrict__ weight = nullptr, size_t bstride = 0, size_t in_stride = 0) {
  extern __shared__ uint32_t sh[];  // [pass][digit][lane]
  // batched launches (grid.y = task): the inputs of task y start in_stride ELEMENTS further on,
  // everything that lives in the scratch buffer bstride BYTES further on
  if (in_stride) {
    if (score) score += (size_t)blockIdx.y * in_stride;
    if (target) target += (size_t)blockIdx.y * in_stride;
    if (weight) weight += (size_t)blockIdx.y * in_stride;
  }
  keys_out = batch_ptr(keys_out, bstride), vals_out = batch_ptr(vals_out, bstride);
  ghist = batch_ptr(ghist, bstride), scal = batch_ptr(scal, bstride), payload_bad = batch_ptr(payload_bad, bstride);
  // kIndexed: the 32-bit payload that travels with a key is its index — or, when every target is
  // exactly a float32 and there are no weights (payload_bad given and 0), the target value itself,
  // which spares the final sweep a random gather by index
  const bool as_payload = MODE == kIndexed && payload_bad && *payload_bad == 0u;
  // ... or, with weights that are all non-negative float32 values on 0/1 targets (rank_totals_kernel
  // checked), the weight's float32 bits with the label in the sign bit: no gather in the sweep either
  auto payload_of = [&](uint32_t i) -> uint32_t {
    if (weight) return __float_as_uint((float)weight[i]) | (target[i] == (TargetT)1 ? 0x80000000u : 0u);
    return __float_as_uint((float)target[i]);
  };
  const unsigned tid = threadIdx.x, lane = tid & 31;
  constexpr int kHists = LEAF ? 2 : 4;
  for (int i = tid; i < kHists * kRadix * kWarp; i += kIngestBlock) sh[i] = 0;
  __syncthreads();

  uint32_t npos = 0, bad_label = 0, neg_score = 0, wide = 0;
  const uint32_t ngroups = (n + 3) / 4;
  const uint32_t stride = gridDim.x * kIngestBlock;
  for (uint32_t g = blockIdx.x * kIngestBlock + tid; g < ngroups; g += stride) {
    const uint32_t i0 = g * 4;
    const uint32_t cnt = min(4u, n - i0);
    uint32_t k[4];
    if (MODE == kRawKeys) {
      if (cnt == 4 && vec_ok) {
        const int4 a = ld_stream_v4(raw_keys + i0);
        k[0] = a.x, k[1] = a.y, k[2] = a.z, k[3] = a.w;
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) k[j] = j < (int)cnt ? raw_keys[i0 + j] : 0u;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < (int)cnt) npos += k[j] & 1u;  // packed keys carry the label in bit 0
    } else {
      float x[4];
      if (cnt == 4 && vec_ok) {
        const int4 a = ld_stream_v4(score + i0);
        x[0] = __int_as_float(a.x), x[1] = __int_as_float(a.y);
        x[2] = __int_as_float(a.z), x[3] = __int_as_float(a.w);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) x[j] = j < (int)cnt ? score[i0 + j] : 0.f;
      }
      if (MODE == kPacked) {
        uint32_t lab[4];
        if constexpr (sizeof(TargetT) == 8) {
          if (cnt == 4 && vec_ok) {
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
            for (int j = 0; j < 4; ++j)
              lab[j] = j < (int)cnt ? LabelOf<TargetT>::get(target[i0 + j], bad_label) : 0u;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            lab[j] = j < (int)cnt ? LabelOf<TargetT>::get(target[i0 + j], bad_label) : 0u;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint32_t b = __float_as_uint(x[j]);
          if (b == 0x80000000u) b = 0u;  // -0.0 ties with +0.0
          neg_score |= b >> 31;          // sign bit: this path needs non-negative scores
          k[j] = (b << 1) | lab[j];
          if (j < (int)cnt) npos += lab[j];
        }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) k[j] = orderable_f32(x[j]);
      }
      if (cnt == 4) {
        st_stream_v4(keys_out + i0, make_int4((int)k[0], (int)k[1], (int)k[2], (int)k[3]));
        if (MODE == kIndexed) {
          if (as_payload)
            st_stream_v4(vals_out + i0, make_int4((int)payload_of(i0), (int)payload_of(i0 + 1), (int)payload_of(i0 + 2),
                                                  (int)payload_of(i0 + 3)));
          else
            st_stream_v4(vals_out + i0, make_int4((int)i0, (int)i0 + 1, (int)i0 + 2, (int)i0 + 3));
        }
      } else {
        for (uint32_t j = 0; j < cnt; ++j) {
          keys_out[i0 + j] = k[j];
          if (MODE == kIndexed) vals_out[i0 + j] = as_payload ? payload_of(i0 + j) : i0 + j;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (j < (int)cnt) {
#pragma unroll
        for (int p = 0; p < kHists; ++p) {
          const uint32_t d = (k[j] >> (LEAF ? kLeafBits + 8 * p : 8 * p)) & 0xffu;
          atomicAdd(sh + ((p << 8) + d) * kWarp + lane, 1u);
        }
        if (LEAF) wide |= k[j] >> 31;
      }
    }
  }
  __syncthreads();
  if (tid < kHists * kRadix) {
    uint32_t s = 0;  // thread t owns (pass, digit) = t
#pragma unroll
    for (int r = 0; r < kWarp; ++r) s += sh[tid * kWarp + ((r + tid) & 31)];
    if (s) atomicAdd(ghist + tid, s);
  }
  if (LEAF && __any_sync(kFull, wide != 0) && lane == 0) atomicOr(&scal->flags, (uint32_t)TE_RANK_FLAG_WIDE_SCORE);
  if (MODE == kRawKeys) {
    npos = warp_sum(npos);
    if (lane == 0 && npos) atomicAdd(&scal->positives, (unsigned long long)npos);
  }
  if (MODE == kPacked) {
    npos = warp_sum(npos);
    const uint32_t anybad = __any_sync(kFull, bad_label != 0);
    const uint32_t anyneg = __any_sync(kFull, neg_score != 0);
    if (lane == 0) {
      if (npos) atomicAdd(&scal->positives, (unsigned long long)npos);
      const uint32_t f = (anybad ? TE_RANK_FLAG_BAD_LABEL : 0) | (anyneg ? TE_RANK_FLAG_NEGATIVE_SCORE : 0);
      if (f) atomicOr(&scal->flags, f);
    }
  }
}

// ------------------------------------------------------------------------------------------
// fused scan / tie-dedup / integrate over sorted packed keys (ascending; bit 0 = label, so
// inside a tie group the negatives come first)
//
// Every CTA owns one contiguous CHUNK of the sorted array and streams it tile by tile, carrying
// its running state (negatives so far, negatives at the start of the current tie group, position
// of that start) in registers.  Chunks do not wait on one another:
//   * AUROC is linear in what a chunk does not know yet.  A positive contributes
//     (#neg before it) + (#neg before its tie group); with E = state of everything before the
//     chunk this is local + E.S (+ E.S, or + E.M if the group started before the chunk), so a
//     chunk reports {state, A = local sum, P = #positives, lead = #positives of the leading run}
//     and a one-CTA scan over the <= 1024 chunk records finishes the job:
//         2U = sum_c  A_c + E_c.S * (2 P_c - lead_c) + E_c.M * lead_c.
//   * AUPRC is not (precision is a ratio of prefix counts), so it takes two sweeps: the same
//     counting sweep, the scan publishes every chunk's E, and a second sweep integrates with
//     exact global counts.
// (A first version used a tile-granular decoupled look-back inside one pass; at B200 speeds a
// 16 KB tile lasts well under the L2 round trip that links neighbouring tiles, and the chain —
// not HBM — set the pace: profiles/r01_tune_rank_v2.txt, 21 % of the HBM roofline.)
// ------------------------------------------------------------------------------------------
constexpr int kFinBlock = 256;
constexpr int kFinIpt = 16;
constexpr int kFinStep = kWarp * kFinIpt;  // records a warp takes per step
constexpr int kFinWarpsPerSm = 32;         // chunks (= warps) per SM the sweep is cut into
constexpr int kMaxChunks = 8192;

struct FState {
  uint32_t S;   // negatives in the segment
  uint32_t M;   // negatives (from segment start) before the segment's last group start
  uint32_t Mp;  // global index of that group start
  uint32_t F;   // segment contains a group start
};
__device__ __forceinline__ FState f_identity() { return FState{0, 0, 0, 0}; }
__device__ __forceinline__ FState f_combine(const FState& a, const FState& b) {  // a then b
  FState r;
  r.S = a.S + b.S;
  r.F = a.F | b.F;
  r.M = b.F ? a.S + b.M : a.M;
  r.Mp = b.F ? b.Mp : a.Mp;
  return r;
}
template <bool WITH_POS>
__device__ __forceinline__ FState f_shfl_up(const FState& v, int o) {
  FState u;
  u.S = __shfl_up_sync(kFull, v.S, o);
  u.M = __shfl_up_sync(kFull, v.M, o);
  u.Mp = WITH_POS ? __shfl_up_sync(kFull, v.Mp, o) : 0u;
  u.F = __shfl_up_sync(kFull, v.F, o);
  return u;
}
template <bool WITH_POS>
__device__ __forceinline__ FState f_shfl(const FState& v, int src) {
  FState u;
  u.S = __shfl_sync(kFull, v.S, src);
  u.M = __shfl_sync(kFull, v.M, src);
  u.Mp = WITH_POS ? __shfl_sync(kFull, v.Mp, src) : 0u;
  u.F = __shfl_sync(kFull, v.F, src);
  return u;
}

// Exclusive warp scan of the lane states + the warp total, specialised for this monoid: S is a
// plain prefix sum; "does a group start in the lanes before me" is one ballot; and the (M, Mp) of
// the LAST such lane come from one indexed shuffle each — instead of five rounds of four shuffles
// and a conditional combine.  Relative to the start of the step, like f_combine over the lanes.
template <bool WITH_POS>
__device__ __forceinline__ void f_warp_scan(const FState& mine, unsigned lane, FState& excl, FState& total) {
  uint32_t s_incl = mine.S;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t u = __shfl_up_sync(kFull, s_incl, o);
    if (lane >= (unsigned)o) s_incl += u;
  }
  const uint32_t s_excl = s_incl - mine.S;
  const uint32_t fb = __ballot_sync(kFull, mine.F != 0);
  const uint32_t m_abs = s_excl + mine.M;  // negatives before my last group start, from the step start
  const uint32_t below = fb & ((1u << lane) - 1u);
  const int src = below ? 31 - __clz(below) : 0;
  const int last = fb ? 31 - __clz(fb) : 0;
  const uint32_t m_src = __shfl_sync(kFull, m_abs, src);
  const uint32_t m_last = __shfl_sync(kFull, m_abs, last);
  excl.S = s_excl;
  excl.F = below ? 1u : 0u;
  excl.M = below ? m_src : 0u;
  total.S = __shfl_sync(kFull, s_incl, 31);
  total.F = fb ? 1u : 0u;
  total.M = fb ? m_last : 0u;
  if (WITH_POS) {
    const uint32_t p_src = __shfl_sync(kFull, mine.Mp, src);
    const uint32_t p_last = __shfl_sync(kFull, mine.Mp, last);
    excl.Mp = below ? p_src : 0u;
    total.Mp = fb ? p_last : 0u;
  } else {
    excl.Mp = total.Mp = 0u;
  }
}

struct alignas(16) ChunkRec {
  FState st;                 // state of the whole chunk
  unsigned long long A;      // AUROC: local part of 2U
  unsigned long long P;      // positives in the chunk
  unsigned long long lead;   // positives whose tie group started before the chunk
  unsigned long long starts; // tie groups that start inside the chunk (counted when POS)
};

// The sorted array is cut into nchunks contiguous chunks of steps_per_chunk * 512 records, one per
// WARP: a warp carries its running state in registers from step to step, the only communication
// inside a step is one warp scan, and nothing is shared between warps (no shared memory, no
// block barrier).  The geometry depends on n and the SM count only, so separate launches (count
// sweep, integrate sweep, the two halves of the rank-sharded AUPRC) agree on it.
struct ChunkGeom {
  uint32_t nchunks, steps_per_chunk;
};
inline ChunkGeom chunk_geom(uint32_t n, int sms) {
  ChunkGeom g;
  uint32_t steps = (n + kFinStep - 1) / kFinStep;
  if (steps < 1) steps = 1;
  uint32_t want = (uint32_t)sms * (uint32_t)kFinWarpsPerSm;
  if (want > (uint32_t)kMaxChunks) want = kMaxChunks;
  if (want > steps) want = steps;
  if (want < 1) want = 1;
  g.steps_per_chunk = (steps + want - 1) / want;
  g.nchunks = (steps + g.steps_per_chunk - 1) / g.steps_per_chunk;
  return g;
}

// MODE 0: counting / AUROC sweep -> rec[chunk].   MODE 1: AUPRC integrate sweep, starts from
// excl[chunk] (global state before the chunk) -> partial[chunk] (double bits).
template <int MODE, bool POS = (MODE == 1)>
__global__ void __launch_bounds__(kFinBlock, kFinWarpsPerSm / (kFinBlock / kWarp))
    rank_chunk_kernel(const uint32_t* __restrict__ keys_, uint32_t n, uint32_t steps_per_chunk,
                      uint32_t nchunks, ChunkRec* rec_, const FState* __restrict__ excl_in_,
                      const RankScalars* __restrict__ scal_, unsigned long long* partial_,
                      double pos_above, double cnt_above, const double* __restrict__ above_dev,
                      size_t bstride) {
  const uint32_t* __restrict__ keys = batch_ptr(keys_, bstride);
  ChunkRec* rec = batch_ptr(rec_, bstride);
  const FState* __restrict__ excl_in = batch_ptr(excl_in_, bstride);
  const RankScalars* __restrict__ scal = batch_ptr(scal_, bstride);
  unsigned long long* partial = batch_ptr(partial_, bstride);
  // POS: also track WHERE the current tie group started (needed by the AUPRC integrate sweep and
  // by the count sweep that feeds it; plain AUROC leaves it out)
  const unsigned lane = threadIdx.x & 31;
  const uint32_t chunk = blockIdx.x * (kFinBlock / kWarp) + (threadIdx.x >> 5);
  if (chunk >= nchunks) return;
  const uint64_t begin64 = (uint64_t)chunk * steps_per_chunk * kFinStep;
  if (begin64 >= n) return;
  const uint32_t begin = (uint32_t)begin64;
  const uint32_t end = (uint32_t)min((uint64_t)n, begin64 + (uint64_t)steps_per_chunk * kFinStep);

  FState carry = MODE == 1 ? excl_in[chunk] : f_identity();
  unsigned long long accA = 0, accP = 0, accLead = 0;
  uint32_t accStarts = 0;
  double accD = 0.0;
  // pos_above / cnt_above: positives / samples with a score above everything in `keys` that live
  // on other ranks (0 when `keys` is the whole sample set); above_dev, if given, holds the same
  // two numbers in device memory
  if (MODE == 1 && above_dev) pos_above = above_dev[0], cnt_above = above_dev[1];
  const double ptot = MODE == 1 ? (double)scal->positives + pos_above : 0.0;

  // score of the record just before this lane's 16 (lane 0: carried from the previous step)
  uint32_t last = begin > 0 ? keys[begin - 1] >> 1 : 0u;

  auto load_step = [&](uint32_t at, uint32_t (&dst)[kFinIpt]) {
    if (at >= end) return;
    const uint32_t j0 = at + lane * kFinIpt;
    if (at + kFinStep <= end) {
#pragma unroll
      for (int q = 0; q < kFinIpt / 4; ++q) {
        const int4 a = ld_stream_v4(keys + j0 + 4 * q);
        dst[4 * q] = a.x, dst[4 * q + 1] = a.y, dst[4 * q + 2] = a.z, dst[4 * q + 3] = a.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < kFinIpt; ++j) dst[j] = (j0 + j < end) ? keys[j0 + j] : 0u;
    }
  };
  uint32_t knext[kFinIpt];
  load_step(begin, knext);

  for (uint32_t at = begin; at < end; at += kFinStep) {
    uint32_t k[kFinIpt];
#pragma unroll
    for (int j = 0; j < kFinIpt; ++j) k[j] = knext[j];
    load_step(at + kFinStep, knext);  // in flight while this step is processed
    const uint32_t i0 = at + lane * kFinIpt;
    const uint32_t cnt = i0 < end ? min((uint32_t)kFinIpt, end - i0) : 0u;

    // score before my first record: the previous lane's last one (a lane with fewer than 16
    // records is followed only by empty lanes, so its k[15] is never looked at)
    uint32_t prev = __shfl_up_sync(kFull, k[kFinIpt - 1] >> 1, 1);
    if (lane == 0) prev = last;
    const uint32_t tail = (end - at >= (uint32_t)kFinStep) ? k[kFinIpt - 1] >> 1 : 0u;
    last = __shfl_sync(kFull, tail, 31);  // only used if a full step follows, i.e. this one was full

    // one local pass: lane aggregate, group-start mask and (MODE 0) the local sums of the AUROC
    // integrand, all relative to the start of this lane's 16 records and in 32-bit
    FState mine = f_identity();
    uint32_t starts = 0;
    uint32_t n_pre = 0, n_post = 0;  // positives before / after the first group start seen here
    uint32_t s_lc = 0;               // sum over positives of (#negatives before it, local)
    uint32_t s_lgs = 0;              // sum over "post" positives of (#negatives before its group, local)
    if (MODE == 0 && cnt == (uint32_t)kFinIpt) {
      // full lane, counting sweep: straight-line integer arithmetic, no predicates per record
      uint32_t S = 0, M = 0, F = 0, npos = 0, Mp = 0;
#pragma unroll
      for (int j = 0; j < kFinIpt; ++j) {
        const uint32_t sc = k[j] >> 1, lab = k[j] & 1u;
        uint32_t st = sc != prev ? 1u : 0u;
        if (j == 0) st |= (i0 == 0) ? 1u : 0u;  // the first record of the array starts a group
        prev = sc;
        M = st ? S : M;
        F |= st;
        if (POS) {
          starts |= st << j;
          Mp = st ? i0 + j : Mp;
        }
        const uint32_t lf = lab & F;
        s_lc += lab * S;
        s_lgs += lf * M;
        npos += lab;
        n_post += lf;
        S += 1u - lab;
      }
      n_pre = npos - n_post;
      mine.S = S, mine.M = M, mine.F = F, mine.Mp = Mp;
    } else {
      uint32_t lgs = 0;
#pragma unroll
      for (int j = 0; j < kFinIpt; ++j) {
        if (j < (int)cnt) {
          const uint32_t sc = k[j] >> 1;
          const bool st = (j == 0 && i0 == 0) || sc != prev;
          prev = sc;
          if (st) {
            starts |= 1u << j;
            mine.F = 1;
            mine.M = mine.S;
            if (POS) mine.Mp = i0 + j;
            lgs = mine.S;
          }
          if (k[j] & 1u) {
            if (MODE == 0) {
              s_lc += mine.S;
              if (mine.F) {
                ++n_post;
                s_lgs += lgs;
              } else {
                ++n_pre;
              }
            }
          } else {
            ++mine.S;
          }
        }
      }
    }

    // warp scan of the lane aggregates
    FState excl, step_agg;
    f_warp_scan<POS>(mine, lane, excl, step_agg);
    const FState g = f_combine(carry, excl);

    if (MODE == 0) {
      // a positive contributes (#neg before it) + (#neg before its tie group): with g = state
      // of everything before this lane inside the chunk,
      //   "post" positives: (g.S + lc) + (g.S + lgs)
      //   "pre"  positives: (g.S + lc) + g.M      (group known inside the chunk: g.F)
      //                     (g.S + lc) + [E.M]    (group older than the chunk: counted in lead)
      const uint32_t npos = n_pre + n_post;
      if (POS) accStarts += __popc(starts);
      accP += npos;
      accA += (unsigned long long)g.S * (npos + n_post) + s_lc + s_lgs;
      if (g.F)
        accA += (unsigned long long)g.M * n_pre;
      else
        accLead += n_pre;
    } else {
      // AUPRC: needs exact global counts per positive
      uint32_t c = g.S, gs = g.M, gp = g.Mp;
#pragma unroll
      for (int j = 0; j < kFinIpt; ++j) {
        if (j < (int)cnt) {
          if (starts & (1u << j)) {
            gs = c;
            gp = i0 + j;
          }
          if (k[j] & 1u) {
            // precision at this tie group's threshold: TP / (TP + FP), counts of score >= s
            const double tp = ptot - (double)(gp - gs);
            accD += tp / ((double)(n - gp) + cnt_above);
          } else {
            ++c;
          }
        }
      }
    }
    carry = f_combine(carry, step_agg);
  }

  if (MODE == 0) {
    accA = warp_sum(accA);
    accP = warp_sum(accP);
    accLead = warp_sum(accLead);
    const unsigned long long nst = warp_sum((unsigned long long)accStarts);
    if (lane == 0) {
      ChunkRec r;
      r.st = carry;
      r.A = accA, r.P = accP, r.lead = accLead, r.starts = nst;
      rec[chunk] = r;
    }
  } else {
    accD = warp_sum(accD);  // butterfly: the same bits in every lane, every run
    if (lane == 0) partial[chunk] = (unsigned long long)__double_as_longlong(accD);
  }
}

// One CTA: exclusive scan of the chunk states, 1024 at a time.  MODE 0 finishes AUROC (out[] as
// documented at te_binary_auroc); MODE 1 publishes excl[chunk] for the AUPRC integrate sweep.
template <int MODE>
__global__ void __launch_bounds__(1024)
    rank_chunk_scan_kernel(const ChunkRec* __restrict__ rec_, uint32_t nchunks, uint32_t n,
                           const RankScalars* __restrict__ scal_, FState* excl_out_, double* out,
                           uint32_t* starts_out_ = nullptr, size_t bstride = 0) {
  const ChunkRec* __restrict__ rec = batch_ptr(rec_, bstride);
  const RankScalars* __restrict__ scal = batch_ptr(scal_, bstride);
  FState* excl_out = batch_ptr(excl_out_, bstride);
  uint32_t* starts_out = batch_ptr(starts_out_, bstride);
  if (out) out += 8 * (size_t)blockIdx.y;
  __shared__ FState s_warp[32];
  __shared__ unsigned long long s_u[32];
  __shared__ uint32_t s_ns[32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  FState carry = f_identity();
  unsigned long long u2 = 0;
  uint32_t starts_carry = 0;  // tie groups that start before the current strip
  for (uint32_t base = 0; base < nchunks; base += 1024) {
    const uint32_t i = base + tid;
    ChunkRec r;
    r.st = f_identity();
    r.A = r.P = r.lead = r.starts = 0;
    if (i < nchunks) r = rec[i];
    FState incl = r.st;
    uint32_t ns_incl = (uint32_t)r.starts;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const FState u = f_shfl_up<true>(incl, o);
      const uint32_t un = __shfl_up_sync(kFull, ns_incl, o);
      if (lane >= (unsigned)o) incl = f_combine(u, incl), ns_incl += un;
    }
    if (lane == 31) s_warp[warp] = incl, s_ns[warp] = ns_incl;
    FState excl = f_shfl_up<true>(incl, 1);
    if (lane == 0) excl = f_identity();
    __syncthreads();
    FState wprefix = f_identity(), strip = f_identity();
    uint32_t ns_prefix = 0, ns_strip = 0;
    for (int w = 0; w < 32; ++w) {
      const FState x = s_warp[w];
      if (w < (int)warp) wprefix = f_combine(wprefix, x), ns_prefix += s_ns[w];
      strip = f_combine(strip, x);
      ns_strip += s_ns[w];
    }
    excl = f_combine(carry, f_combine(wprefix, excl));
    if (MODE == 1) {
      if (i < nchunks) {
        excl_out[i] = excl;
        if (starts_out) starts_out[i] = starts_carry + ns_prefix + ns_incl - (uint32_t)r.starts;
      }
    } else if (i < nchunks) {
      u2 += r.A + (unsigned long long)excl.S * (2ull * r.P - r.lead) + (unsigned long long)excl.M * r.lead;
    }
    carry = f_combine(carry, strip);
    starts_carry += ns_strip;
    __syncthreads();  // s_warp is rewritten by the next strip
  }
  if (MODE == 1) {
    if (starts_out && tid == 0) starts_out[nchunks] = starts_carry;  // number of distinct scores
    return;
  }
  u2 = warp_sum(u2);
  if (lane == 0) s_u[warp] = u2;
  __syncthreads();
  if (tid == 0) {
    unsigned long long tu = 0;
    for (int w = 0; w < 32; ++w) tu += s_u[w];
    const double P = (double)scal->positives, N = (double)n - P;
    out[0] = (P * N == 0.0) ? 0.5 : ((double)tu * 0.5) / (P * N);
    out[1] = P;
    out[2] = N;
    out[3] = (double)tu;
    out[4] = (double)scal->flags;
    out[5] = (double)(tu >> 32);           // 2U exactly, as two 32-bit halves
    out[6] = (double)(tu & 0xffffffffull);
    out[7] = (double)n;
  }
}

// Precision-recall curve (precision_recall_curve.py:208-231): one point per distinct score, in
// ascending-threshold order — which is the order the keys are sorted in, so the "flip" of the
// reference costs nothing.  For the tie group that starts at sorted position gp, with gs
// negatives before it: TP = P - (gp - gs) samples are positive with score >= threshold, out of
// n - gp samples with score >= threshold.  The output index of a group is the number of groups
// that start before it: chunk offset (starts_excl) + a warp scan inside the step.
__global__ void __launch_bounds__(kFinBlock, kFinWarpsPerSm / (kFinBlock / kWarp))
    rank_pr_curve_kernel(const uint32_t* __restrict__ keys, uint32_t n, uint32_t steps_per_chunk,
                         uint32_t nchunks, const FState* __restrict__ excl_in,
                         const uint32_t* __restrict__ starts_excl, const RankScalars* __restrict__ scal,
                         float* __restrict__ precision, float* __restrict__ recall,
                         float* __restrict__ thresholds) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t chunk = blockIdx.x * (kFinBlock / kWarp) + (threadIdx.x >> 5);
  if (chunk >= nchunks) return;
  const uint64_t begin64 = (uint64_t)chunk * steps_per_chunk * kFinStep;
  if (begin64 >= n) return;
  const uint32_t begin = (uint32_t)begin64;
  const uint32_t end = (uint32_t)min((uint64_t)n, begin64 + (uint64_t)steps_per_chunk * kFinStep);
  const double ptot = (double)scal->positives;
  if (chunk == 0 && lane == 0) {  // the closing point without a threshold
    const uint32_t K = starts_excl[nchunks];
    precision[K] = 1.0f;
    recall[K] = 0.0f;
  }
  FState carry = excl_in[chunk];
  uint32_t out_at = starts_excl[chunk];
  uint32_t last = begin > 0 ? keys[begin - 1] >> 1 : 0u;
  for (uint32_t at = begin; at < end; at += kFinStep) {
    const uint32_t i0 = at + lane * kFinIpt;
    const uint32_t cnt = i0 < end ? min((uint32_t)kFinIpt, end - i0) : 0u;
    uint32_t k[kFinIpt];
    if (at + kFinStep <= end) {
#pragma unroll
      for (int q = 0; q < kFinIpt / 4; ++q) {
        const int4 a = ld_stream_v4(keys + i0 + 4 * q);
        k[4 * q] = a.x, k[4 * q + 1] = a.y, k[4 * q + 2] = a.z, k[4 * q + 3] = a.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < kFinIpt; ++j) k[j] = (i0 + j < end) ? keys[i0 + j] : 0u;
    }
    uint32_t prev = __shfl_up_sync(kFull, k[kFinIpt - 1] >> 1, 1);
    if (lane == 0) prev = last;
    const uint32_t tail = (end - at >= (uint32_t)kFinStep) ? k[kFinIpt - 1] >> 1 : 0u;
    last = __shfl_sync(kFull, tail, 31);
    FState mine = f_identity();
    uint32_t starts = 0;
#pragma unroll
    for (int j = 0; j < kFinIpt; ++j) {
      if (j < (int)cnt) {
        const uint32_t sc = k[j] >> 1;
        const bool st = (j == 0 && i0 == 0) || sc != prev;
        prev = sc;
        if (st) {
          starts |= 1u << j;
          mine.F = 1;
          mine.M = mine.S;
          mine.Mp = i0 + j;
        }
        if (!(k[j] & 1u)) ++mine.S;
      }
    }
    FState incl = mine;
    uint32_t ns_incl = __popc(starts);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const FState u = f_shfl_up<true>(incl, o);
      const uint32_t un = __shfl_up_sync(kFull, ns_incl, o);
      if (lane >= (unsigned)o) incl = f_combine(u, incl), ns_incl += un;
    }
    FState excl = f_shfl_up<true>(incl, 1);
    if (lane == 0) excl = f_identity();
    const FState step_agg = f_shfl<true>(incl, 31);
    const uint32_t step_starts = __shfl_sync(kFull, ns_incl, 31);
    const FState g = f_combine(carry, excl);
    uint32_t c = g.S, o = out_at + ns_incl - __popc(starts);
#pragma unroll
    for (int j = 0; j < kFinIpt; ++j) {
      if (j < (int)cnt) {
        if (starts & (1u << j)) {
          const uint32_t gp = i0 + j;
          const double tp = ptot - (double)(gp - c);
          precision[o] = (float)(tp / (double)(n - gp));
          recall[o] = ptot == 0.0 ? 1.0f : (float)(tp / ptot);  // no positives: NaN -> 1.0 (:228-229)
          thresholds[o] = __uint_as_float(k[j] >> 1);
          ++o;
        }
        if (!(k[j] & 1u)) ++c;
      }
    }
    carry = f_combine(carry, step_agg);
    out_at += step_starts;
  }
}

// Fixed-order sum of the per-tile partials and the final division.
//   out[0] = metric   out[1] = #positives   out[2] = #negatives   out[3] = raw sum
//   out[4] = TE_RANK_FLAG_* seen by the ingest kernel (non-zero: out[0] is not valid)
template <int MODE>
__global__ void __launch_bounds__(1024)
    rank_reduce_kernel(const unsigned long long* __restrict__ partial_, uint32_t ntiles, uint32_t n,
                       const RankScalars* __restrict__ scal_, double* out, size_t bstride = 0) {
  const unsigned long long* __restrict__ partial = batch_ptr(partial_, bstride);
  const RankScalars* __restrict__ scal = batch_ptr(scal_, bstride);
  out += 8 * (size_t)blockIdx.y;
  __shared__ unsigned long long s_u[32];
  __shared__ double s_d[32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long au = 0;
  double ad = 0.0;
  // contiguous slab per thread keeps the summation order independent of anything dynamic
  const uint32_t per = (ntiles + 1023) / 1024;
  const uint32_t b = min(ntiles, tid * per), e = min(ntiles, b + per);
  for (uint32_t i = b; i < e; ++i) {
    if (MODE == 0)
      au += partial[i];
    else
      ad += __longlong_as_double((long long)partial[i]);
  }
  au = warp_sum(au);
  // butterfly sums are order-symmetric, so `ad` is the same on every run as well
  ad = warp_sum(ad);
  if (lane == 0) s_u[warp] = au, s_d[warp] = ad;
  __syncthreads();
  if (tid == 0) {
    unsigned long long tu = 0;
    double td = 0.0;
    for (int w = 0; w < 32; ++w) tu += s_u[w], td += s_d[w];
    const double P = (double)scal->positives, N = (double)n - P;
    out[1] = P;
    out[2] = N;
    out[4] = (double)scal->flags;
    if (MODE == 0) {
      out[3] = (double)tu;
      out[0] = (P * N == 0.0) ? 0.5 : ((double)tu * 0.5) / (P * N);
    } else {
      out[3] = td;
      // no positives: recall is NaN -> 1.0, the last step then contributes -0.0
      // (precision_recall_curve.py:227-229)
      out[0] = P == 0.0 ? -0.0 : td / P;
    }
  }
}

// ------------------------------------------------------------------------------------------
// host side: scratch plan + launch sequences
// ------------------------------------------------------------------------------------------
inline int sort_block_from_env() {
  static const int b = [] {
    const char* e = getenv("TE_B200_SORT_BLOCK");
    const int v = e ? atoi(e) : kSortBlockDefault;
    return (v == 256 || v == 512) ? v : kSortBlockDefault;
  }();
  return b;
}

inline int sort_ipt_from_env() {
  static const int b = [] {
    const char* e = getenv("TE_B200_SORT_IPT");
    const int v = e ? atoi(e) : kSortIptDefault;
    return (v == 12 || v == 16) ? v : kSortIptDefault;
  }();
  return b;
}

inline bool hot_digits_from_env() {
  static const bool b = [] {
    const char* e = getenv("TE_B200_HOT_DIGITS");
    return e ? atoi(e) != 0 : false;  // measured a wash once the bank rotation is in (profiles/r01_skew_bench.txt)
  }();
  return b;
}

inline bool skip_passes_from_env() {
  static const bool b = [] {
    const char* e = getenv("TE_B200_SKIP_PASSES");
    return e ? atoi(e) != 0 : true;
  }();
  return b;
}

inline bool unstable_first_pass_from_env() {
  static const bool b = [] {
    const char* e = getenv("TE_B200_UNSTABLE_FIRST");
    return e ? atoi(e) != 0 : true;
  }();
  return b;
}

struct RankPlan {
  uint32_t n;
  int sort_block, sort_ipt;
  uint32_t sort_tiles;
  size_t off_keys_a, off_keys_b, off_vals_a, off_vals_b;
  size_t off_zero;      // start of the region that must be zero before launch
  size_t off_hist;      // 4*256 u32
  size_t off_ticket;    // 8 u32: words 0..3 = "pass p is a no-op" flags (the name is historical)
  size_t off_hot;       // 8 u32
  size_t off_scal;      // RankScalars
  size_t off_lookback;  // look-back counts of the first pass: sort_tiles * 256 u32 (zeroed by memset)
  size_t off_gprefix;   // group prefixes, 4 passes * ceil(sort_tiles / 32) * 256 u32 (zeroed by memset)
  size_t off_lookback_rest;  // counts of passes 1..3: each pass clears the rows of the next one
  size_t off_agg, off_inc;
  size_t off_starts;    // exclusive count of tie-group starts per chunk (+ total): PR curve
  size_t zero_bytes;
  size_t off_partial;
  // counting-stage path (rank_leaf.cuh)
  uint32_t leaf_piece, leaf_max_items, leaf_max_slabs;
  size_t off_leaf_hdr, off_seg_stats;  // inside the zeroed region
  size_t off_seg_start, off_leaf_items, off_leaf_multi, off_leaf_slabs;
  size_t off_range_hist, off_range_base;  // uint32[kMaxRanges][256] each (rank_part.cuh)
  size_t off_label_bits;                  // one bit per record, in 128-record groups of 16 bytes
  size_t off_joint;                       // uint32[65536] joint digit histogram (zeroed region), rank_msd.cuh
  size_t off_cursor1, off_cursor2, off_msd_tables;
  size_t off_pos_below, off_ap_part, off_ap_totals;  // average precision on the counting stage (rank_leaf_ap.cuh)
  size_t total;
};
constexpr uint32_t kMaxRanges = 1024;

inline uint32_t leaf_piece_of(int64_t n) {
  uint64_t m = ((uint64_t)n / 1024 + 4095) / 4096 * 4096;
  if (m < kLeafMinPiece) m = kLeafMinPiece;
  if (m > kLeafMaxPiece) m = kLeafMaxPiece;
  return (uint32_t)m;
}

RankPlan make_plan(int64_t n64, bool with_vals, bool with_keys, bool with_final, bool with_leaf = false) {
  RankPlan p{};
  p.n = (uint32_t)n64;
  p.sort_block = sort_block_from_env();
  p.sort_ipt = sort_ipt_from_env();
  const int64_t sort_tile = (int64_t)p.sort_block * p.sort_ipt;
  p.sort_tiles = (uint32_t)((n64 + sort_tile - 1) / sort_tile);
  if (p.sort_tiles == 0) p.sort_tiles = 1;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    const size_t at = o;
    o += align_up(bytes ? bytes : 1, 256);
    return at;
  };
  const size_t kb = (size_t)n64 * 4;
  p.off_keys_a = take(with_keys ? kb : 0);
  p.off_keys_b = take(with_keys ? kb : 0);
  p.off_vals_a = take(with_vals ? kb : 0);
  p.off_vals_b = take(with_vals ? kb : 0);
  p.off_zero = o;
  p.off_hist = take(4 * kRadix * 4);
  p.off_ticket = take(8 * 4);
  p.off_hot = take(8 * 4);  // hot digits: 2 per pass
  p.off_scal = take(sizeof(RankScalars));
  const size_t lb_groups = (p.sort_tiles + kLbGroup - 1) / kLbGroup;
  p.off_lookback = take((size_t)p.sort_tiles * kRadix * 4);
  p.off_gprefix = take((size_t)4 * lb_groups * kRadix * 4);
  p.off_agg = take(with_final ? (size_t)kMaxChunks * sizeof(ChunkRec) : 0);
  p.off_inc = take(with_final ? (size_t)kMaxChunks * sizeof(FState) : 0);
  p.off_starts = take(with_final ? (size_t)(kMaxChunks + 1) * 4 : 0);
  p.leaf_piece = leaf_piece_of(n64);
  p.leaf_max_items = with_leaf ? (uint32_t)(kLeafSegs + (uint64_t)n64 / p.leaf_piece + 1) : 0;
  p.leaf_max_slabs = with_leaf ? (uint32_t)(2 * ((uint64_t)n64 / p.leaf_piece) + 2) : 0;
  p.off_leaf_hdr = take(with_leaf ? sizeof(LeafHeader) : 0);
  p.off_seg_stats = take(with_leaf ? (size_t)kLeafSegs * sizeof(SegStat) : 0);
  p.off_joint = take(with_leaf ? (size_t)kLeafSegs * 4 : 0);
  p.zero_bytes = o - p.off_zero;
  p.off_lookback_rest = take((size_t)3 * p.sort_tiles * kRadix * 4);
  p.off_partial = take(with_final ? (size_t)kMaxChunks * 8 : 0);
  p.off_seg_start = take(with_leaf ? (size_t)(kLeafSegs + 1) * 4 : 0);
  p.off_leaf_items = take((size_t)p.leaf_max_items * sizeof(LeafItem));
  p.off_leaf_multi = take(with_leaf ? (size_t)kLeafSegs * sizeof(LeafMulti) : 0);
  p.off_leaf_slabs = take((size_t)p.leaf_max_slabs * kLeafBins * 4);
  p.off_range_hist = take(with_leaf ? (size_t)kMaxRanges * kRadix * 4 : 0);
  p.off_range_base = take(with_leaf ? (size_t)kMaxRanges * kRadix * 4 : 0);
  p.off_label_bits = take(with_leaf ? ((size_t)n64 + 127) / 128 * 16 : 0);
  p.off_cursor1 = take(with_leaf ? kRadix * 4 : 0);
  p.off_cursor2 = take(with_leaf ? (size_t)kLeafSegs * 4 : 0);
  p.off_msd_tables = take(with_leaf ? sizeof(MsdTables) : 0);
  p.off_pos_below = take(with_leaf ? (size_t)kLeafSegs * 4 : 0);
  p.off_ap_part = take(with_leaf ? (size_t)kLeafSegs * 8 : 0);
  p.off_ap_totals = take(with_leaf ? sizeof(ApTotals) : 0);
  p.total = o;
  return p;
}

template <typename T>
T* at(void* base, size_t off) {
  return reinterpret_cast<T*>(static_cast<char*>(base) + off);
}

// nb > 1: nb independent sorts of the same length whose scratch blocks (and key arrays) lie
// bstride bytes apart, done by the same launches (grid.y = nb).
int run_passes(const RankPlan& p, void* scratch, const uint32_t* k_in0, uint32_t* k_x, uint32_t* k_y,
               const uint32_t* v_in0, uint32_t* v_x, uint32_t* v_y, cudaStream_t st, int nb = 1,
               size_t bstride = 0, int npasses = 4, int shift0 = 0) {
  // npasses passes on the digits at bits shift0 + 8 * pass: in0 -> x -> y -> x -> y
  uint32_t* hist = at<uint32_t>(scratch, p.off_hist);
  uint32_t* hot = hot_digits_from_env() ? at<uint32_t>(scratch, p.off_hot) : nullptr;
  // words 0..3 of the (zeroed) ticket block: "pass p is a no-op" flags, set by digit_scan_kernel
  uint32_t* skip = skip_passes_from_env() ? at<uint32_t>(scratch, p.off_ticket) : nullptr;
  const size_t lb_groups = (p.sort_tiles + kLbGroup - 1) / kLbGroup;
  auto lb_counts = [&](int pass) {
    return pass == 0 ? at<uint32_t>(scratch, p.off_lookback)
                     : at<uint32_t>(scratch, p.off_lookback_rest) + (size_t)(pass - 1) * p.sort_tiles * kRadix;
  };
  {
    TE_PROF("digit_scan", st);
    digit_scan_kernel<<<dim3(1, nb), 1024, 0, st>>>(hist, npasses, hot, bstride, skip);
  }
  TE_LAUNCH_CHECK();
  const int block = p.sort_block;
  const size_t smem = block == 256 ? onesweep_smem_bytes<uint32_t, 256>() : onesweep_smem_bytes<uint32_t, 512>();
  using PassFn = void (*)(const uint32_t*, uint32_t*, const uint32_t*, uint32_t*, uint32_t,
                          const RadixDigit, const uint32_t*, uint32_t*, uint32_t*, const PeerPtrs,
                          const uint32_t*, uint32_t*, size_t, const uint32_t*, const uint32_t*);
  PassFn fn = nullptr;
  if (v_in0)
    fn = block == 256 ? onesweep_pass_kernel<uint32_t, true, 256, RadixDigit>
                      : onesweep_pass_kernel<uint32_t, true, 512, RadixDigit>;
  else
    fn = block == 256 ? onesweep_pass_kernel<uint32_t, false, 256, RadixDigit>
                      : onesweep_pass_kernel<uint32_t, false, 512, RadixDigit>;
  size_t smem_used = smem;
  if (p.sort_ipt == 12) {  // experimental geometry: 512 x 12 keys, 3 CTAs per SM
    if (block != 512) return TE_ERR_UNSUPPORTED;
    fn = v_in0 ? onesweep_pass_kernel<uint32_t, true, 512, RadixDigit, 12>
               : onesweep_pass_kernel<uint32_t, false, 512, RadixDigit, 12>;
    smem_used = onesweep_smem_bytes<uint32_t, 512, 12>();
  }
  TE_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_used));
  const uint32_t* kin = k_in0;
  const uint32_t* vin = v_in0;
  // keys only: the first pass may hand out ranks in arrival order (see kUnstable)
  PassFn fn_first = fn;
  if (!v_in0 && p.sort_ipt == kSortIptDefault && unstable_first_pass_from_env()) {
    fn_first = block == 256 ? onesweep_pass_kernel<uint32_t, false, 256, RadixDigit, kSortIptDefault, false, true>
                            : onesweep_pass_kernel<uint32_t, false, 512, RadixDigit, kSortIptDefault, false, true>;
    TE_CUDA_TRY(cudaFuncSetAttribute(fn_first, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_used));
  }
  for (int pass = 0; pass < npasses; ++pass) {
    uint32_t* kout = (pass & 1) ? k_y : k_x;
    uint32_t* vout = (pass & 1) ? v_y : v_x;
    uint32_t* lbp = lb_counts(pass);
    uint32_t* gpp = at<uint32_t>(scratch, p.off_gprefix) + (size_t)pass * lb_groups * kRadix;
    TE_PROF(v_in0 ? "onesweep_pass_pairs" : (pass == 0 && fn_first != fn ? "onesweep_pass_keys_first" : "onesweep_pass_keys"), st);
    (pass == 0 ? fn_first : fn)<<<dim3(p.sort_tiles, nb), block, smem_used, st>>>(kin, kout, vin, vout, p.n, RadixDigit{shift0 + 8 * pass},
                                               hist + pass * kRadix, lbp, gpp, PeerPtrs{},
                                               hot ? hot + 2 * pass : nullptr,
                                               pass + 1 < npasses ? lb_counts(pass + 1) : nullptr, bstride,
                                               skip ? skip + pass : nullptr, nullptr);
    TE_LAUNCH_CHECK();
    kin = kout;
    vin = vout;
  }
  return TE_OK;
}

template <int MODE, typename TargetT, bool LEAF = false>
int launch_ingest(const float* score, const void* target, const uint32_t* raw, const RankPlan& p,
                  uint32_t* keys_out, uint32_t* vals_out, void* scratch, int sms, cudaStream_t st,
                  const uint32_t* payload_bad = nullptr, const double* weight = nullptr, int nb = 1,
                  size_t bstride = 0, size_t in_stride = 0) {
  auto kern = rank_ingest_kernel<MODE, TargetT, LEAF>;
  TE_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)kIngestSmem));
  const uint32_t ngroups = (p.n + 3) / 4;
  uint32_t grid = (ngroups + kIngestBlock - 1) / kIngestBlock;
  if (grid > (uint32_t)sms) grid = sms;
  if (grid < 1) grid = 1;
  uintptr_t a = (uintptr_t)score | (uintptr_t)target | (uintptr_t)raw;
  const bool vec_ok = (a & 15) == 0 && in_stride % 16 == 0;  // (every task's rows stay 16-byte aligned)
  TE_PROF(MODE == kPacked ? (LEAF ? "rank_ingest_packed2" : "rank_ingest_packed")
                          : (MODE == kIndexed ? "rank_ingest_indexed" : "rank_hist_keys"), st);
  kern<<<dim3(grid, nb), kIngestBlock, kIngestSmem, st>>>(score, (const TargetT*)target, raw, p.n, vec_ok,
                                                           keys_out, vals_out, at<uint32_t>(scratch, p.off_hist),
                                                           at<RankScalars>(scratch, p.off_scal), payload_bad, weight,
                                                           bstride, in_stride);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

template <int MODE, bool LEAF = false>
int launch_ingest_dtype(const float* score, const void* target, int32_t dt, const RankPlan& p,
                        uint32_t* keys_out, uint32_t* vals_out, void* scratch, int sms,
                        cudaStream_t st) {
  switch (dt) {
    case TE_DTYPE_I64:
      return launch_ingest<MODE, long long, LEAF>(score, target, nullptr, p, keys_out, vals_out, scratch, sms, st);
    case TE_DTYPE_I32:
      return launch_ingest<MODE, int, LEAF>(score, target, nullptr, p, keys_out, vals_out, scratch, sms, st);
    case TE_DTYPE_U8:
      return launch_ingest<MODE, unsigned char, LEAF>(score, target, nullptr, p, keys_out, vals_out, scratch, sms, st);
    case TE_DTYPE_F32:
      return launch_ingest<MODE, float, LEAF>(score, target, nullptr, p, keys_out, vals_out, scratch, sms, st);
    case TE_DTYPE_F64:
      return launch_ingest<MODE, double, LEAF>(score, target, nullptr, p, keys_out, vals_out, scratch, sms, st);
    default:
      return TE_ERR_INVALID_ARGUMENT;
  }
}

// fused scan / integrate over the sorted packed keys + final division (both modes)
int finish_packed(const RankPlan& p, const uint32_t* ka, int mode, double* out, void* scratch, int sms,
                  cudaStream_t st, bool track_pos = false, int nb = 1, size_t bstride = 0) {
  const int64_t n = p.n;
  ChunkRec* rec = at<ChunkRec>(scratch, p.off_agg);
  FState* excl = at<FState>(scratch, p.off_inc);
  unsigned long long* partial = at<unsigned long long>(scratch, p.off_partial);
  const RankScalars* scal = at<RankScalars>(scratch, p.off_scal);
  const ChunkGeom cg = chunk_geom(p.n, sms);
  const uint32_t nchunks = n > 0 ? cg.nchunks : 0;
  const dim3 grid((nchunks + kFinBlock / kWarp - 1) / (kFinBlock / kWarp), nb);
  if (nchunks) {
    TE_PROF("rank_chunk_sweep", st);
    if (mode == 1 || track_pos)
      rank_chunk_kernel<0, true><<<grid, kFinBlock, 0, st>>>(ka, p.n, cg.steps_per_chunk, nchunks, rec, nullptr,
                                                             scal, nullptr, 0.0, 0.0, nullptr, bstride);
    else
      rank_chunk_kernel<0><<<grid, kFinBlock, 0, st>>>(ka, p.n, cg.steps_per_chunk, nchunks, rec, nullptr, scal,
                                                       nullptr, 0.0, 0.0, nullptr, bstride);
  }
  TE_LAUNCH_CHECK();
  if (mode == 0) {
    TE_PROF("rank_chunk_scan", st);
    rank_chunk_scan_kernel<0><<<dim3(1, nb), 1024, 0, st>>>(rec, nchunks, p.n, scal, nullptr, out, nullptr, bstride);
  } else {
    if (nchunks) {
      {
        TE_PROF("rank_chunk_scan", st);
        rank_chunk_scan_kernel<1><<<dim3(1, nb), 1024, 0, st>>>(rec, nchunks, p.n, scal, excl, nullptr, nullptr, bstride);
      }
      TE_PROF("rank_chunk_integrate", st);
      rank_chunk_kernel<1><<<grid, kFinBlock, 0, st>>>(ka, p.n, cg.steps_per_chunk, nchunks, nullptr, excl, scal,
                                                       partial, 0.0, 0.0, nullptr, bstride);
    }
    TE_PROF("rank_reduce", st);
    rank_reduce_kernel<1><<<dim3(1, nb), 1024, 0, st>>>(partial, nchunks, p.n, scal, out, bstride);
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

// Counting stage over keys ordered by their top 16 bits (rank_leaf.cuh): segment bounds -> work
// list -> histogram pieces -> merge of the segments counted in several pieces -> combine.
int finish_leaf(const RankPlan& p, const uint32_t* keys, double* out, void* scratch, int sms, cudaStream_t st,
                const uint32_t* n_dev = nullptr, bool have_seg_start = false, int nb = 1, size_t bstride = 0) {
  // nb > 1: nb independent problems in the same launches (grid.y), scratch slices bstride bytes apart,
  // out rows 8 doubles apart (needs have_seg_start: the batched callers come from the MSD passes)
  uint32_t* seg_start = at<uint32_t>(scratch, p.off_seg_start);
  LeafItem* items = at<LeafItem>(scratch, p.off_leaf_items);
  LeafMulti* multi = at<LeafMulti>(scratch, p.off_leaf_multi);
  LeafHeader* hdr = at<LeafHeader>(scratch, p.off_leaf_hdr);
  SegStat* stats = at<SegStat>(scratch, p.off_seg_stats);
  uint32_t* slabs = at<uint32_t>(scratch, p.off_leaf_slabs);
  if (!have_seg_start) {  // (the MSD passes leave the segment bounds: the prefix of their joint histogram)
    TE_PROF("leaf_seg_bounds", st);
    leaf_seg_bounds_kernel<<<(kLeafSegs + 1 + 255) / 256, 256, 0, st>>>(keys, p.n, seg_start, n_dev);
  }
  {
    TE_PROF("leaf_tiny_plan", st);  // CTA 0: the work list; the others: segments of <= 128 keys, a warp each
    leaf_tiny_plan_kernel<<<dim3(1 + kLeafSegs / 32, nb), 1024, 0, st>>>(keys, seg_start, stats, p.leaf_piece, p.leaf_max_items,
                                                                        p.leaf_max_slabs, items, multi, hdr, bstride);
  }
  TE_LAUNCH_CHECK();
  TE_CUDA_TRY(cudaFuncSetAttribute(leaf_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLeafSmem));
  {
    TE_PROF("leaf_count", st);
    leaf_count_kernel<<<dim3(sms, nb), kLeafBlock, kLeafSmem, st>>>(keys, items, hdr, stats, slabs, bstride);
  }
  {
    TE_PROF("leaf_merge", st);
    leaf_merge_kernel<<<dim3(sms, nb), kLeafBlock, 0, st>>>(multi, hdr, slabs, stats, bstride);
  }
  {
    TE_PROF("leaf_combine", st);
    leaf_combine_kernel<<<dim3(1, nb), 1024, 0, st>>>(stats, p.n, &at<RankScalars>(scratch, p.off_scal)->flags, out, n_dev,
                                                      bstride);
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

// Average precision on the counting stage: phase A = the AUROC stage without its combine ({P, N} per
// segment), prefix, phase B over the same work list (rank_leaf_ap.cuh).  seg_start must be there
// already (the MSD passes leave it).
int finish_leaf_ap(const RankPlan& p, const uint32_t* keys, double* out, void* scratch, int sms, cudaStream_t st) {
  uint32_t* seg_start = at<uint32_t>(scratch, p.off_seg_start);
  LeafItem* items = at<LeafItem>(scratch, p.off_leaf_items);
  LeafMulti* multi = at<LeafMulti>(scratch, p.off_leaf_multi);
  LeafHeader* hdr = at<LeafHeader>(scratch, p.off_leaf_hdr);
  SegStat* stats = at<SegStat>(scratch, p.off_seg_stats);
  uint32_t* slabs = at<uint32_t>(scratch, p.off_leaf_slabs);
  uint32_t* pos_below = at<uint32_t>(scratch, p.off_pos_below);
  double* ap_part = at<double>(scratch, p.off_ap_part);
  ApTotals* totals = at<ApTotals>(scratch, p.off_ap_totals);
  TE_CUDA_TRY(cudaFuncSetAttribute(leaf_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLeafSmem));
  TE_CUDA_TRY(cudaFuncSetAttribute(leaf_count_ap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kLeafSmem));
  TE_CUDA_TRY(cudaMemsetAsync(ap_part, 0, (size_t)kLeafSegs * 8, st));
  {
    TE_PROF("leaf_tiny", st);
    leaf_tiny_kernel<<<kLeafSegs / 8, 256, 0, st>>>(keys, seg_start, stats);
  }
  {
    TE_PROF("leaf_plan", st);
    leaf_plan_kernel<<<1, 1024, 0, st>>>(seg_start, p.leaf_piece, p.leaf_max_items, p.leaf_max_slabs, items, multi, hdr);
  }
  {
    TE_PROF("leaf_count", st);
    leaf_count_kernel<<<sms, kLeafBlock, kLeafSmem, st>>>(keys, items, hdr, stats, slabs);
  }
  {
    TE_PROF("leaf_merge", st);
    leaf_merge_kernel<<<sms, kLeafBlock, 0, st>>>(multi, hdr, slabs, stats);
  }
  {
    TE_PROF("leaf_ap_prefix", st);
    leaf_ap_prefix_kernel<<<1, 1024, 0, st>>>(stats, pos_below, totals);
  }
  TE_LAUNCH_CHECK();
  TE_CUDA_TRY(cudaMemsetAsync(&hdr->ticket, 0, sizeof(uint32_t), st));  // the same work list, from the top
  {
    TE_PROF("leaf_tiny_ap", st);
    leaf_tiny_ap_kernel<<<kLeafSegs / 8, 256, 0, st>>>(keys, seg_start, pos_below, totals, ap_part);
  }
  {
    TE_PROF("leaf_count_ap", st);
    leaf_count_ap_kernel<<<sms, kLeafBlock, kLeafSmem, st>>>(keys, items, hdr, seg_start, pos_below, totals, ap_part, slabs);
  }
  {
    TE_PROF("leaf_merge_ap", st);
    leaf_merge_ap_kernel<<<sms, kLeafBlock, 0, st>>>(multi, hdr, slabs, seg_start, pos_below, totals, ap_part);
  }
  {
    TE_PROF("leaf_ap_reduce", st);
    leaf_ap_reduce_kernel<<<1, 1024, 0, st>>>(ap_part, totals, &at<RankScalars>(scratch, p.off_scal)->flags, out);
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

// Which structure te_binary_auroc uses: TE_RANK_AUTO picks the counting-stage path from
// TE_B200_LEAF_MIN_N scores on (the fixed cost of walking up to 65 536 segment histograms only
// pays off for large inputs).
inline int64_t leaf_min_n_from_env() {
  static const int64_t v = [] {
    const char* e = getenv("TE_B200_LEAF_MIN_N");
    return e ? atoll(e) : (int64_t)1 << 25;  // crossover measured at ~2^24.7 (profiles/r02_leaf_bench_v5.txt)
  }();
  return v;
}

inline bool range_pass_from_env() {
  static const bool b = [] {
    const char* e = getenv("TE_B200_RANGE_PASS");
    return e ? atoi(e) != 0 : true;
  }();
  return b;
}

// ingest (per-range histograms) -> range bases -> first partition pass without look-back
// (rank_part.cuh) -> one stable look-back pass on the top digit.
//   raw_keys == nullptr: (score, target) are packed into `ka`;  ka -> kb -> ka, result in ka.
//   raw_keys given (packed already, the receive side of the sample sort): raw -> ka -> kb, result in kb.
template <typename TargetT>
int leaf_passes_ranges(const RankPlan& p, void* scratch, const float* score, const TargetT* target,
                       const uint32_t* raw_keys, uint32_t* ka, uint32_t* kb, int sms, cudaStream_t st) {
  static const bool small_tiles = [] {
    const char* e = getenv("TE_B200_PART_BLOCK");
    return e && atoi(e) == kPrBlockSmall;
  }();
  const int pblock = small_tiles ? kPrBlockSmall : kPrBlock;
  auto part = small_tiles ? partition_ranges_kernel<RadixDigit, kPrBlockSmall, kPrIptSmall>
                          : partition_ranges_kernel<RadixDigit, kPrBlock, kPrIpt>;
  const size_t part_smem = (size_t)2 * pblock * (small_tiles ? kPrIptSmall : kPrIpt) * sizeof(uint32_t);
  TE_CUDA_TRY(cudaFuncSetAttribute(part, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)part_smem));
  int per_sm = 0;
  TE_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, part, pblock, part_smem));
  if (per_sm < 1) return TE_ERR_UNSUPPORTED;
  // the ingest wants ~3 CTAs of 512 threads per SM to saturate HBM, the downsweep runs per_sm
  // (1 for the 1024-thread geometry): a downsweep CTA takes `rpc` consecutive ingest ranges
  const int rpc = per_sm >= 3 ? 1 : (per_sm == 2 ? 2 : 3);
  int resident = per_sm * sms * rpc;
  if (resident > (int)kMaxRanges) resident = kMaxRanges / rpc * rpc;
  RangeGeom rg = range_geom(p.n, resident);
  const uint32_t part_grid = (rg.ranges + rpc - 1) / rpc;
  uint32_t* hist = at<uint32_t>(scratch, p.off_hist);
  uint32_t* range_hist = at<uint32_t>(scratch, p.off_range_hist);
  uint32_t* range_base = at<uint32_t>(scratch, p.off_range_base);
  uint32_t* flags = &at<RankScalars>(scratch, p.off_scal)->flags;
  uint32_t* skip = skip_passes_from_env() ? at<uint32_t>(scratch, p.off_ticket) : nullptr;
  static const bool bits_ok = [] {
    const char* e = getenv("TE_B200_LABEL_BITS");
    return e ? atoi(e) != 0 : true;
  }();
  uint32_t* label_bits = (!raw_keys && bits_ok && ((uintptr_t)score & 15) == 0) ? at<uint32_t>(scratch, p.off_label_bits) : nullptr;
  if (raw_keys) {
    auto ing = rank_ingest_ranges_kernel<long long, kIrRaw>;
    TE_CUDA_TRY(cudaFuncSetAttribute(ing, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kIrSmem));
    TE_PROF("rank_hist_ranges", st);
    ing<<<rg.ranges, kIrBlock, kIrSmem, st>>>(reinterpret_cast<const float*>(raw_keys), nullptr, p.n,
                                              ((uintptr_t)raw_keys & 15) == 0, rg.range_records, nullptr, hist,
                                              range_hist, flags);
  } else {
    // label_bits: the upsweep leaves one bit per label instead of the packed keys, the first pass
    // packs from the scores and the bits (needs 16-byte aligned scores for its cp.async copies)
    auto ing_keys = rank_ingest_ranges_kernel<TargetT, kIrKeys>;
    auto ing_bits = rank_ingest_ranges_kernel<TargetT, kIrBits>;
    TE_CUDA_TRY(cudaFuncSetAttribute(ing_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kIrSmem));
    TE_CUDA_TRY(cudaFuncSetAttribute(ing_bits, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kIrSmem));
    const bool vec_ok = (((uintptr_t)score | (uintptr_t)target) & 15) == 0;
    TE_PROF(label_bits ? "rank_upsweep_bits" : "rank_ingest_ranges", st);
    if (label_bits)
      ing_bits<<<rg.ranges, kIrBlock, kIrSmem, st>>>(score, target, p.n, vec_ok, rg.range_records, label_bits, hist, range_hist, flags);
    else
      ing_keys<<<rg.ranges, kIrBlock, kIrSmem, st>>>(score, target, p.n, vec_ok, rg.range_records, ka, hist, range_hist, flags);
  }
  TE_LAUNCH_CHECK();
  {
    TE_PROF("digit_scan", st);
    digit_scan_kernel<<<1, 1024, 0, st>>>(hist, 2, nullptr, 0, skip);
    range_base_kernel<<<1, kRadix, 0, st>>>(range_hist, rg.ranges, hist, range_base);
  }
  TE_LAUNCH_CHECK();
  const uint32_t* in1 = raw_keys ? raw_keys : (label_bits ? reinterpret_cast<const uint32_t*>(score) : ka);
  uint32_t* out1 = raw_keys ? ka : kb;
  uint32_t* out2 = raw_keys ? kb : ka;
  {
    TE_PROF(label_bits ? "partition_ranges_pack" : "partition_ranges", st);
    part<<<part_grid, pblock, part_smem, st>>>(in1, out1, p.n, RadixDigit{kLeafBits}, range_base, rg.range_records,
                                               (uint32_t)rpc, skip, SlabView{nullptr, 0, 0, 0}, label_bits);
  }
  TE_LAUNCH_CHECK();
  // second pass: the look-back kernel, stable, on bits 23..30 (its look-back rows are the memset ones)
  if (p.sort_block != 512 || p.sort_ipt != kSortIptDefault) return TE_ERR_UNSUPPORTED;
  auto fn = onesweep_pass_kernel<uint32_t, false, 512, RadixDigit>;
  const size_t smem = onesweep_smem_bytes<uint32_t, 512>();
  TE_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  TE_PROF("onesweep_pass_keys", st);
  fn<<<dim3(p.sort_tiles, 1), 512, smem, st>>>(out1, out2, nullptr, nullptr, p.n, RadixDigit{kLeafBits + 8}, hist + kRadix,
                                              at<uint32_t>(scratch, p.off_lookback), at<uint32_t>(scratch, p.off_gprefix),
                                              PeerPtrs{}, nullptr, nullptr, 0, skip ? skip + 1 : nullptr, nullptr);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

// Upsweep (joint histogram of both digits + label bits) -> prefix -> two cursor-placed partition
// passes, most significant digit first (rank_msd.cuh).  seg_start is left for the counting stage.
//   raw_keys == nullptr: (score, target) in; pass 1 packs into kb, pass 2 -> ka; result in ka.
//   raw_keys given: raw -> ka -> kb; result in kb.
template <typename TargetT>
int leaf_passes_msd_geom(const RankPlan& p, void* scratch, const float* score, const TargetT* target,
                         const uint32_t* raw_keys, uint32_t* ka, uint32_t* kb, int sms, cudaStream_t st,
                         const SlabView view, const uint32_t* n_dev, int nb, size_t bstride);

// nb > 1 (raw_keys only): nb key arrays of p.n records each in the same launches (grid.y), every
// scratch-resident pointer bstride bytes apart — the classes of a one-vs-rest batch.
template <typename TargetT>
int leaf_passes_msd(const RankPlan& p, void* scratch, const float* score, const TargetT* target,
                    const uint32_t* raw_keys, uint32_t* ka, uint32_t* kb, int sms, cudaStream_t st,
                    const SlabView view = SlabView{nullptr, 0, 0, 0}, const uint32_t* n_dev = nullptr, int nb = 1,
                    size_t bstride = 0) {
  return leaf_passes_msd_geom<TargetT>(p, scratch, score, target, raw_keys, ka, kb, sms, st, view, n_dev, nb, bstride);
}

template <typename TargetT, int PBLOCK, int PIPT>
int leaf_passes_msd_impl(const RankPlan& p, void* scratch, const float* score, const TargetT* target,
                    const uint32_t* raw_keys, uint32_t* ka, uint32_t* kb, int sms, cudaStream_t st,
                    const SlabView view, const uint32_t* n_dev, int nb = 1, size_t bstride = 0) {
  constexpr int TILE = PBLOCK * PIPT;
  constexpr int kCtasPerSm = 1024 / PBLOCK;
#ifndef TE_TOP_REPL
#define TE_TOP_REPL 4
#endif
#ifndef TE_LOW_REPL
#define TE_LOW_REPL 1
#endif
  constexpr int kTopRepl = TE_TOP_REPL;  // counter copies of the first pass (its digit is the skewed one)
  constexpr int kLowRepl = TE_LOW_REPL;
  const size_t part_smem = (size_t)2 * TILE * sizeof(uint32_t);
  uint32_t* joint = at<uint32_t>(scratch, p.off_joint);
  uint32_t* seg_start = at<uint32_t>(scratch, p.off_seg_start);
  uint32_t* cursor1 = at<uint32_t>(scratch, p.off_cursor1);
  uint32_t* cursor2 = at<uint32_t>(scratch, p.off_cursor2);
  MsdTables* tables = at<MsdTables>(scratch, p.off_msd_tables);
  uint32_t* label_bits = at<uint32_t>(scratch, p.off_label_bits);
  uint32_t* flags = &at<RankScalars>(scratch, p.off_scal)->flags;
  const uint32_t nchunks = (p.n + kUjChunk - 1) / kUjChunk;  // (slab view: p.n is the capacity of the inbox)
  const uint32_t ugrid = nchunks < (uint32_t)sms ? nchunks : (uint32_t)sms;
  if (raw_keys) {
    auto up = rank_upsweep_joint_kernel<long long, true>;
    TE_CUDA_TRY(cudaFuncSetAttribute(up, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUjSmem));
    TE_PROF("rank_upsweep_joint_keys", st);
    up<<<dim3(ugrid, nb), kUjBlock, kUjSmem, st>>>(reinterpret_cast<const float*>(raw_keys), nullptr, p.n,
                                                   ((uintptr_t)raw_keys & 15) == 0, nullptr, joint, flags, view, bstride);
  } else {
    auto up = rank_upsweep_joint_kernel<TargetT, false>;
    TE_CUDA_TRY(cudaFuncSetAttribute(up, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUjSmem));
    TE_PROF("rank_upsweep_joint", st);
    up<<<ugrid, kUjBlock, kUjSmem, st>>>(score, target, p.n, (((uintptr_t)score | (uintptr_t)target) & 15) == 0,
                                         label_bits, joint, flags, SlabView{nullptr, 0, 0, 0}, (size_t)0);
  }
  TE_LAUNCH_CHECK();
  {
    TE_PROF("joint_scan", st);
    // one digit value holding >= hot8 / 8 of the records switches a pass to its hot-digit instantiation
    // (the first pass counts in four copies of its table and never needs the other instantiation:
    // 3.05 against 3.7 ms at 2^30 even when every record carries the same digit)
    const uint32_t hot8_top = 9u;
    static const uint32_t hot8_low = [] { const char* e = getenv("TE_B200_HOT8_LOW"); return e ? (uint32_t)atoi(e) : 3u; }();
    joint_scan_kernel<TILE><<<dim3(1, nb), 1024, 0, st>>>(joint, seg_start, cursor1, cursor2, tables, bstride, hot8_top, hot8_low);
  }
  TE_LAUNCH_CHECK();
  // both instantiations of a pass are launched; the device-side tables decide which one works
  // (the host never sees the histogram)
  const uint32_t tiles1 = (p.n + TILE - 1) / TILE;
  const uint32_t grid1 = tiles1 < (uint32_t)(sms * kCtasPerSm) ? tiles1 : (uint32_t)(sms * kCtasPerSm);
  uint32_t* out1 = raw_keys ? ka : kb;
  uint32_t* out2 = raw_keys ? kb : ka;
  const uint32_t* in1 = raw_keys ? raw_keys : reinterpret_cast<const uint32_t*>(score);
  const uint32_t* bits1 = raw_keys ? nullptr : label_bits;
  auto launch = [&](auto kern, const char* name, uint32_t grid, const uint32_t* in, const uint32_t* bits, uint32_t* outp,
                    const uint32_t* segs, uint32_t* cursors) -> int {
    TE_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)part_smem));
    TE_PROF(name, st);
    // (pass 1 reads through the view and needs no count; pass 2 and later take it from n_dev)
    kern<<<dim3(grid, nb), PBLOCK, part_smem, st>>>(in, bits, outp, p.n, segs ? n_dev : nullptr, segs, tables, cursors,
                                                      segs ? SlabView{nullptr, 0, 0, 0} : view, bstride);
    TE_LAUNCH_CHECK();
    return TE_OK;
  };
  int rc;
  if (raw_keys) {
    if ((rc = launch(cursor_partition_kernel<false, false, false, PBLOCK, PIPT, kTopRepl>, "cursor_pass_top", grid1, in1, bits1, out1, nullptr, cursor1))) return rc;
  } else {
    if ((rc = launch(cursor_partition_kernel<false, true, false, PBLOCK, PIPT, kTopRepl>, "cursor_pass_top_pack", grid1, in1, bits1, out1, nullptr, cursor1))) return rc;
  }
  if ((rc = launch(cursor_partition_kernel<true, false, false, PBLOCK, PIPT, kLowRepl>, "cursor_pass_low", (uint32_t)(sms * kCtasPerSm), out1, nullptr, out2, seg_start, cursor2))) return rc;
  if ((rc = launch(cursor_partition_kernel<true, false, true, PBLOCK, PIPT, 8>, "cursor_pass_low_hot", (uint32_t)(sms * kCtasPerSm), out1, nullptr, out2, seg_start, cursor2))) return rc;
  return TE_OK;
}

template <typename TargetT>
int leaf_passes_msd_geom(const RankPlan& p, void* scratch, const float* score, const TargetT* target,
                         const uint32_t* raw_keys, uint32_t* ka, uint32_t* kb, int sms, cudaStream_t st,
                         const SlabView view, const uint32_t* n_dev, int nb, size_t bstride) {
  static const int geom = [] {
    const char* e = getenv("TE_B200_PART_GEOM");
    return e ? atoi(e) : 0;
  }();
  if (geom == 1) return leaf_passes_msd_impl<TargetT, 512, 24>(p, scratch, score, target, raw_keys, ka, kb, sms, st, view, n_dev, nb, bstride);
  return leaf_passes_msd_impl<TargetT, kPrBlock, kPrIpt>(p, scratch, score, target, raw_keys, ka, kb, sms, st, view, n_dev, nb, bstride);
}

int leaf_passes_msd_dtype(const RankPlan& p, void* scratch, const float* score, const void* target, int32_t dt,
                          uint32_t* ka, uint32_t* kb, int sms, cudaStream_t st) {
  switch (dt) {
    case TE_DTYPE_I64: return leaf_passes_msd(p, scratch, score, (const long long*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_I32: return leaf_passes_msd(p, scratch, score, (const int*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_U8: return leaf_passes_msd(p, scratch, score, (const unsigned char*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_F32: return leaf_passes_msd(p, scratch, score, (const float*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_F64: return leaf_passes_msd(p, scratch, score, (const double*)target, nullptr, ka, kb, sms, st);
    default: return TE_ERR_INVALID_ARGUMENT;
  }
}

int leaf_passes_ranges_dtype(const RankPlan& p, void* scratch, const float* score, const void* target, int32_t dt,
                             uint32_t* ka, uint32_t* kb, int sms, cudaStream_t st) {
  switch (dt) {
    case TE_DTYPE_I64: return leaf_passes_ranges(p, scratch, score, (const long long*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_I32: return leaf_passes_ranges(p, scratch, score, (const int*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_U8: return leaf_passes_ranges(p, scratch, score, (const unsigned char*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_F32: return leaf_passes_ranges(p, scratch, score, (const float*)target, nullptr, ka, kb, sms, st);
    case TE_DTYPE_F64: return leaf_passes_ranges(p, scratch, score, (const double*)target, nullptr, ka, kb, sms, st);
    default: return TE_ERR_INVALID_ARGUMENT;
  }
}

int binary_auroc_leaf(const float* score, const void* target, int32_t dt, int64_t n, double* out, void* scratch,
                      size_t scratch_bytes, cudaStream_t st, int mode = 0) {
  const RankPlan p = make_plan(n, false, true, true, true);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  uint32_t* ka = at<uint32_t>(scratch, p.off_keys_a);
  uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
  static const bool msd = [] {
    const char* e = getenv("TE_B200_MSD_PASSES");
    return e ? atoi(e) != 0 : true;
  }();
  if (n > 0 && msd && ((uintptr_t)score & 15) == 0) {
    const int rc = leaf_passes_msd_dtype(p, scratch, score, target, dt, ka, kb, sms, st);
    if (rc) return rc;
    return mode == 1 ? finish_leaf_ap(p, ka, out, scratch, sms, st) : finish_leaf(p, ka, out, scratch, sms, st, nullptr, true);
  }
  if (mode == 1) return TE_ERR_UNSUPPORTED;  // (average precision rides on the MSD passes only)
  if (n > 0 && range_pass_from_env() && p.sort_block == 512 && p.sort_ipt == kSortIptDefault) {
    const int rc = leaf_passes_ranges_dtype(p, scratch, score, target, dt, ka, kb, sms, st);
    if (rc) return rc;
  } else if (n > 0) {
    int rc = launch_ingest_dtype<kPacked, true>(score, target, dt, p, ka, nullptr, scratch, sms, st);
    if (rc) return rc;
    // a -> b -> a: keys ordered by bits 15..30 end in `ka`
    rc = run_passes(p, scratch, ka, kb, ka, nullptr, nullptr, nullptr, st, 1, 0, 2, kLeafBits);
    if (rc) return rc;
  }
  return finish_leaf(p, ka, out, scratch, sms, st);
}

int binary_rank_packed(const float* score, const void* target, int32_t dt, int64_t n, int mode,
                       double* out, void* scratch, size_t scratch_bytes, cudaStream_t st,
                       int variant = TE_RANK_AUTO) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  static const bool msd_on = [] {
    const char* e = getenv("TE_B200_MSD_PASSES");
    return e ? atoi(e) != 0 : true;
  }();
  const bool counting_ok = mode == 0 || (msd_on && n > 0 && ((uintptr_t)score & 15) == 0);
  if (counting_ok && (variant == TE_RANK_COUNTING || (variant == TE_RANK_AUTO && n >= leaf_min_n_from_env())))
    return binary_auroc_leaf(score, target, dt, n, out, scratch, scratch_bytes, st, mode);
  const RankPlan p = make_plan(n, false, true, true);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  uint32_t* ka = at<uint32_t>(scratch, p.off_keys_a);
  uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
  if (n > 0) {
    int rc = launch_ingest_dtype<kPacked>(score, target, dt, p, ka, nullptr, scratch, sms, st);
    if (rc) return rc;
    // a -> b -> a -> b -> a : sorted keys end in `ka`
    rc = run_passes(p, scratch, ka, kb, ka, nullptr, nullptr, nullptr, st);
    if (rc) return rc;
  }
  return finish_packed(p, ka, mode, out, scratch, sms, st);
}


// ------------------------------------------------------------------------------------------
// general path: any float32 score (negative, mixed sign), soft labels, per-sample weights.
// Sort (key, index) pairs, then one fused pass that gathers target/weight by index and
// integrates with float64 running sums.  With tp_j = w_j * t_j and fp_j = w_j * (1 - t_j)
// (auroc.py:129-130) and TPb(j), FPb(j) the sums over all samples whose score is strictly below
// sample j's score, the reference's trapezoid area over tie groups equals
//     area = ( TP * FP + sum_j ( tp_j * FPb(j) - fp_j * TPb(j) ) ) / 2
// which needs only "running sum at the start of my tie group" — the same carry as the packed
// path.  AUPRC (positives are target == 1, precision_recall_curve.py:216):
//     AP = (1/P) * sum_{j positive} (P - TPb(j)) / (n - TPb(j) - FPb(j)).
// ------------------------------------------------------------------------------------------
constexpr int kGenBlock = 512;
constexpr int kGenIpt = 8;
constexpr int kGenTile = kGenBlock * kGenIpt;

struct GenScalars {
  double tp_total, fp_total;
  uint32_t payload_bad;  // != 0: gather target / weight by index; 0: the payload carries them (target value as
                         // float32, or — with weights — float32 weight bits + the 0/1 label in the sign bit)
  uint32_t pad;
};

struct GState {
  double Stp, Sfp;  // sums over the segment
  double Mtp, Mfp;  // sums from the segment start up to its last group start
  uint32_t F;
};
__device__ __forceinline__ GState g_identity() { return GState{0.0, 0.0, 0.0, 0.0, 0u}; }
__device__ __forceinline__ GState g_combine(const GState& a, const GState& b) {
  GState r;
  r.Stp = a.Stp + b.Stp;
  r.Sfp = a.Sfp + b.Sfp;
  r.F = a.F | b.F;
  r.Mtp = b.F ? a.Stp + b.Mtp : a.Mtp;
  r.Mfp = b.F ? a.Sfp + b.Mfp : a.Mfp;
  return r;
}
struct alignas(64) GenRec {
  unsigned long long flag;
  unsigned long long w[5];
  unsigned long long pad[2];
};
template <typename TargetT>
__device__ __forceinline__ void gen_terms(const TargetT* __restrict__ target,
                                          const double* __restrict__ weight, uint32_t idx, int mode,
                                          double& tp, double& fp) {
  const double t = (double)target[idx];
  if (mode == 0) {
    const double w = weight ? weight[idx] : 1.0;
    tp = w * t;
    fp = w * (1.0 - t);
  } else {
    tp = t == 1.0 ? 1.0 : 0.0;
    fp = 1.0 - tp;
  }
}

// the same from a payload that IS the target value (float32 bits), no weights
__device__ __forceinline__ void gen_terms_payload(uint32_t val, int mode, double& tp, double& fp, bool weighted = false) {
  if (weighted) {  // float32 weight bits, 0/1 label in the sign bit (mode 0 only: AUPRC takes no weights)
    const double w = (double)__uint_as_float(val & 0x7fffffffu);
    tp = (val >> 31) ? w : 0.0;
    fp = (val >> 31) ? 0.0 : w;
    return;
  }
  const double t = (double)__uint_as_float(val);
  if (mode == 0) {
    tp = t;
    fp = 1.0 - t;
  } else {
    tp = t == 1.0 ? 1.0 : 0.0;
    fp = 1.0 - tp;
  }
}

// totals needed up front by the AUPRC integrand (and reported for AUROC)
template <typename TargetT>
__global__ void __launch_bounds__(256)
    rank_totals_kernel(const TargetT* __restrict__ target, const double* __restrict__ weight,
                       uint32_t n, int mode, GenScalars* gs, size_t bstride = 0, size_t in_stride = 0) {
  __shared__ double ws[8];
  target += (size_t)blockIdx.y * in_stride;
  if (weight) weight += (size_t)blockIdx.y * in_stride;
  gs = batch_ptr(gs, bstride);
  double a = 0.0, b = 0.0;
  uint32_t inexact = 0u;
  for (uint32_t i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
    double tp, fp;
    gen_terms(target, weight, i, mode, tp, fp);
    a += tp;
    b += fp;
    const TargetT t = target[i];
    if (weight) {
      // weights ride along as float32 bits with the label in the sign bit: needs 0/1 targets and
      // weights that are non-negative (sign bit clear) float32 values
      const double w = weight[i];
      const float wf = (float)w;
      inexact |= ((double)wf != w || (__float_as_uint(wf) >> 31) || !(t == (TargetT)0 || t == (TargetT)1)) ? 1u : 0u;
    } else {
      inexact |= ((TargetT)(float)t != t) ? 1u : 0u;
    }
  }
  if (__any_sync(kFull, inexact != 0u) && (threadIdx.x & 31) == 0) atomicOr(&gs->payload_bad, 1u);
  double ta, tb;
  block_scan_incl<double, 256>(a, ws, ta);
  block_scan_incl<double, 256>(b, ws, tb);
  if (threadIdx.x == 0) {
    atomicAdd(&gs->tp_total, ta);
    atomicAdd(&gs->fp_total, tb);
  }
}

__global__ void __launch_bounds__(1024)
    rank_reduce_general_kernel(const double* __restrict__ partial, uint32_t ntiles, int mode,
                               const GenRec* __restrict__ inc, const GenScalars* __restrict__ gs,
                               double* out, size_t bstride = 0) {
  __shared__ double s_d[32];
  partial = batch_ptr(partial, bstride), inc = batch_ptr(inc, bstride), gs = batch_ptr(gs, bstride);
  out += (size_t)blockIdx.y * 8;
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double ad = 0.0;
  const uint32_t per = (ntiles + 1023) / 1024;
  const uint32_t b = min(ntiles, tid * per), e = min(ntiles, b + per);
  for (uint32_t i = b; i < e; ++i) ad += partial[i];
  ad = warp_sum(ad);
  if (lane == 0) s_d[warp] = ad;
  __syncthreads();
  if (tid == 0) {
    double td = 0.0;
    for (int w = 0; w < 32; ++w) td += s_d[w];
    double TP = gs->tp_total, FP = gs->fp_total;
    if (ntiles > 0 && mode == 0) {
      // take the totals from the scan itself so that area and factor are consistent
      TP = __longlong_as_double((long long)inc[ntiles - 1].w[0]);
      FP = __longlong_as_double((long long)inc[ntiles - 1].w[1]);
    }
    out[1] = TP;
    out[2] = FP;
    out[3] = td;
    out[4] = 0.0;
    if (mode == 0) {
      const double factor = TP * FP;
      out[0] = factor == 0.0 ? 0.5 : 0.5 * (factor + td) / factor;
    } else {
      out[0] = TP == 0.0 ? -0.0 : td / TP;
    }
  }
}

// ---- chunked form of the general AUROC sweep (mode 0) ----------------------------------------
// Same idea as rank_chunk_kernel: the area term  sum_j tp_j * FPb(j) - fp_j * TPb(j)  is LINEAR in
// the state E of everything before a chunk, so every warp sweeps one contiguous chunk with its
// state in registers and reports
//   A      the part that needs nothing from outside the chunk
//   TP2/FP2      sum of tp / fp over records whose tie group starts inside the chunk (they see E.S)
//   TPlead/FPlead  the same over the leading run whose group started before the chunk (they see E.M)
// and one CTA finishes:  sum_c A_c + E_c.Sfp*TP2_c - E_c.Stp*FP2_c + E_c.Mfp*TPlead_c - E_c.Mtp*FPlead_c.
constexpr int kGcIpt = 8;
constexpr int kGcStep = kWarp * kGcIpt;  // 256 records per warp step
constexpr int kGcBlock = 256;

struct alignas(16) GenChunkRec {
  double Stp, Sfp, Mtp, Mfp;
  double A, TP2, FP2, TPlead, FPlead;
  unsigned long long F;
};

inline ChunkGeom gen_chunk_geom(uint32_t n, int sms) {
  ChunkGeom g;
  uint32_t steps = (n + kGcStep - 1) / kGcStep;
  if (steps < 1) steps = 1;
  uint32_t want = (uint32_t)sms * 32u;
  if (want > (uint32_t)kMaxChunks) want = kMaxChunks;
  if (want > steps) want = steps;
  g.steps_per_chunk = (steps + want - 1) / want;
  g.nchunks = (steps + g.steps_per_chunk - 1) / g.steps_per_chunk;
  return g;
}

// exclusive warp scan of GState + warp total (prefix sums, one ballot, indexed shuffles)
__device__ __forceinline__ void g_warp_scan(const GState& mine, unsigned lane, GState& excl, GState& total) {
  double st = mine.Stp, sf = mine.Sfp;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double ut = __shfl_up_sync(kFull, st, o), uf = __shfl_up_sync(kFull, sf, o);
    if (lane >= (unsigned)o) st += ut, sf += uf;
  }
  const double et = __shfl_up_sync(kFull, st, 1), ef = __shfl_up_sync(kFull, sf, 1);
  excl.Stp = lane ? et : 0.0;
  excl.Sfp = lane ? ef : 0.0;
  const uint32_t fb = __ballot_sync(kFull, mine.F != 0);
  const double mt_abs = excl.Stp + mine.Mtp, mf_abs = excl.Sfp + mine.Mfp;
  const uint32_t below = fb & ((1u << lane) - 1u);
  const int src = below ? 31 - __clz(below) : 0, last = fb ? 31 - __clz(fb) : 0;
  const double mts = __shfl_sync(kFull, mt_abs, src), mfs = __shfl_sync(kFull, mf_abs, src);
  const double mtl = __shfl_sync(kFull, mt_abs, last), mfl = __shfl_sync(kFull, mf_abs, last);
  excl.F = below ? 1u : 0u;
  excl.Mtp = below ? mts : 0.0;
  excl.Mfp = below ? mfs : 0.0;
  total.Stp = __shfl_sync(kFull, st, 31);
  total.Sfp = __shfl_sync(kFull, sf, 31);
  total.F = fb ? 1u : 0u;
  total.Mtp = fb ? mtl : 0.0;
  total.Mfp = fb ? mfl : 0.0;
}

template <typename TargetT>
__global__ void __launch_bounds__(kGcBlock)
    gen_chunk_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                     const TargetT* __restrict__ target, const double* __restrict__ weight, uint32_t n,
                     uint32_t steps_per_chunk, uint32_t nchunks, const GenScalars* __restrict__ gs,
                     GenChunkRec* rec, size_t bstride = 0, size_t in_stride = 0) {
  keys = batch_ptr(keys, bstride), vals = batch_ptr(vals, bstride);
  gs = batch_ptr(gs, bstride), rec = batch_ptr(rec, bstride);
  target += (size_t)blockIdx.y * in_stride;
  if (weight) weight += (size_t)blockIdx.y * in_stride;
  const unsigned lane = threadIdx.x & 31;
  const uint32_t chunk = blockIdx.x * (kGcBlock / kWarp) + (threadIdx.x >> 5);
  if (chunk >= nchunks) return;
  const uint64_t begin64 = (uint64_t)chunk * steps_per_chunk * kGcStep;
  if (begin64 >= n) return;
  const uint32_t begin = (uint32_t)begin64;
  const uint32_t end = (uint32_t)min((uint64_t)n, begin64 + (uint64_t)steps_per_chunk * kGcStep);
  const bool as_payload = gs->payload_bad == 0u;  // what the ingest kernel stored in vals
  GState carry = g_identity();
  double A = 0.0, TP2 = 0.0, FP2 = 0.0, TPlead = 0.0, FPlead = 0.0;
  uint32_t last = begin > 0 ? keys[begin - 1] : 0u;
  for (uint32_t at = begin; at < end; at += kGcStep) {
    const uint32_t i0 = at + lane * kGcIpt;
    const uint32_t cnt = i0 < end ? min((uint32_t)kGcIpt, end - i0) : 0u;
    uint32_t k[kGcIpt], v[kGcIpt];
    if (cnt == (uint32_t)kGcIpt) {
#pragma unroll
      for (int q = 0; q < kGcIpt / 4; ++q) {
        const int4 a = ld_stream_v4(keys + i0 + 4 * q), b = ld_stream_v4(vals + i0 + 4 * q);
        k[4 * q] = a.x, k[4 * q + 1] = a.y, k[4 * q + 2] = a.z, k[4 * q + 3] = a.w;
        v[4 * q] = b.x, v[4 * q + 1] = b.y, v[4 * q + 2] = b.z, v[4 * q + 3] = b.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < kGcIpt; ++j) {
        k[j] = j < (int)cnt ? keys[i0 + j] : 0u;
        v[j] = j < (int)cnt ? vals[i0 + j] : 0u;
      }
    }
    // key before my first record: the previous lane's last one (only full lanes are followed by
    // non-empty ones), lane 0: carried from the previous step
    uint32_t prev = __shfl_up_sync(kFull, k[kGcIpt - 1], 1);
    if (lane == 0) prev = last;
    last = __shfl_sync(kFull, k[kGcIpt - 1], 31);

    GState mine = g_identity();
    double a_loc = 0.0, tp_post = 0.0, fp_post = 0.0, tp_pre = 0.0, fp_pre = 0.0;
#pragma unroll
    for (int j = 0; j < kGcIpt; ++j) {
      if (j < (int)cnt) {
        double tp, fp;
        if (as_payload)
          gen_terms_payload(v[j], 0, tp, fp, weight != nullptr);
        else
          gen_terms(target, weight, v[j], 0, tp, fp);
        const bool st = (j == 0 && i0 == 0) || k[j] != prev;
        prev = k[j];
        if (st) {
          mine.F = 1;
          mine.Mtp = mine.Stp;
          mine.Mfp = mine.Sfp;
        }
        if (mine.F) {
          a_loc += tp * mine.Mfp - fp * mine.Mtp;
          tp_post += tp;
          fp_post += fp;
        } else {
          tp_pre += tp;
          fp_pre += fp;
        }
        mine.Stp += tp;
        mine.Sfp += fp;
      }
    }
    GState excl, total;
    g_warp_scan(mine, lane, excl, total);
    const GState g = g_combine(carry, excl);  // everything before this lane, inside the chunk
    A += a_loc + g.Sfp * tp_post - g.Stp * fp_post;
    TP2 += tp_post;
    FP2 += fp_post;
    if (g.F) {
      A += g.Mfp * tp_pre - g.Mtp * fp_pre;
      TP2 += tp_pre;
      FP2 += fp_pre;
    } else {
      TPlead += tp_pre;
      FPlead += fp_pre;
    }
    carry = g_combine(carry, total);
  }
  A = warp_sum(A), TP2 = warp_sum(TP2), FP2 = warp_sum(FP2);  // butterflies: the same bits every run
  TPlead = warp_sum(TPlead), FPlead = warp_sum(FPlead);
  if (lane == 0) {
    GenChunkRec r;
    r.Stp = carry.Stp, r.Sfp = carry.Sfp, r.Mtp = carry.Mtp, r.Mfp = carry.Mfp, r.F = carry.F;
    r.A = A, r.TP2 = TP2, r.FP2 = FP2, r.TPlead = TPlead, r.FPlead = FPlead;
    rec[chunk] = r;
  }
}

// One CTA: exclusive scan of the chunk states (strips of 1024) and the final sum.  Writes the
// area term into partial[0] and the totals of the scan into inc[0] (where
// rank_reduce_general_kernel looks for them).
__global__ void __launch_bounds__(1024)
    gen_chunk_scan_kernel(GenChunkRec* rec, uint32_t nchunks, double* partial, GenRec* inc,
                          int write_excl, size_t bstride = 0) {
  rec = batch_ptr(rec, bstride), partial = batch_ptr(partial, bstride), inc = batch_ptr(inc, bstride);
  __shared__ GState s_warp[32];
  __shared__ double s_sum[32];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  GState carry = g_identity();
  double acc = 0.0;
  for (uint32_t base = 0; base < nchunks; base += 1024) {
    const uint32_t i = base + tid;
    GenChunkRec r{};
    if (i < nchunks) r = rec[i];
    GState mine;
    mine.Stp = r.Stp, mine.Sfp = r.Sfp, mine.Mtp = r.Mtp, mine.Mfp = r.Mfp, mine.F = (uint32_t)r.F;
    GState excl, wtotal;
    g_warp_scan(mine, lane, excl, wtotal);
    if (lane == 31) s_warp[warp] = wtotal;
    __syncthreads();
    GState wprefix = g_identity(), strip = g_identity();
    for (int w = 0; w < 32; ++w) {
      const GState x = s_warp[w];
      if (w < (int)warp) wprefix = g_combine(wprefix, x);
      strip = g_combine(strip, x);
    }
    const GState E = g_combine(carry, g_combine(wprefix, excl));
    if (i < nchunks) {
      acc += r.A + E.Sfp * r.TP2 - E.Stp * r.FP2;
      if (E.F) acc += E.Mfp * r.TPlead - E.Mtp * r.FPlead;
      if (write_excl) {  // AUPRC: the integrate sweep starts every chunk from E
        r.Stp = E.Stp, r.Sfp = E.Sfp, r.Mtp = E.Mtp, r.Mfp = E.Mfp, r.F = E.F;
        rec[i] = r;
      }
    }
    carry = g_combine(carry, strip);
    __syncthreads();
  }
  acc = warp_sum(acc);
  if (lane == 0) s_sum[warp] = acc;
  __syncthreads();
  if (tid == 0) {
    double t = 0.0;
    for (int w = 0; w < 32; ++w) t += s_sum[w];
    if (!write_excl) partial[0] = t;
    inc[0].w[0] = (unsigned long long)__double_as_longlong(carry.Stp);
    inc[0].w[1] = (unsigned long long)__double_as_longlong(carry.Sfp);
  }
}

// General AUPRC, second sweep: with E (the state of everything before the chunk, left in rec[] by
// the scan) every positive's precision at its tie group's threshold is known exactly:
//   (P - TPb) / (n - TPb - FPb),  TPb / FPb = sums over everything strictly below the group.
template <typename TargetT>
__global__ void __launch_bounds__(kGcBlock)
    gen_chunk_integrate_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                               const TargetT* __restrict__ target, uint32_t n, uint32_t steps_per_chunk,
                               uint32_t nchunks, const GenScalars* __restrict__ gs,
                               const GenChunkRec* __restrict__ rec, double* partial) {
  const unsigned lane = threadIdx.x & 31;
  const uint32_t chunk = blockIdx.x * (kGcBlock / kWarp) + (threadIdx.x >> 5);
  if (chunk >= nchunks) return;
  const uint64_t begin64 = (uint64_t)chunk * steps_per_chunk * kGcStep;
  if (begin64 >= n) return;
  const uint32_t begin = (uint32_t)begin64;
  const uint32_t end = (uint32_t)min((uint64_t)n, begin64 + (uint64_t)steps_per_chunk * kGcStep);
  const bool as_payload = gs->payload_bad == 0u;
  const double ptot = gs->tp_total, ntot = gs->tp_total + gs->fp_total;
  GState carry;
  {
    const GenChunkRec e = rec[chunk];
    carry.Stp = e.Stp, carry.Sfp = e.Sfp, carry.Mtp = e.Mtp, carry.Mfp = e.Mfp, carry.F = (uint32_t)e.F;
  }
  double acc = 0.0;
  uint32_t last = begin > 0 ? keys[begin - 1] : 0u;
  for (uint32_t at = begin; at < end; at += kGcStep) {
    const uint32_t i0 = at + lane * kGcIpt;
    const uint32_t cnt = i0 < end ? min((uint32_t)kGcIpt, end - i0) : 0u;
    uint32_t k[kGcIpt];
    double tp[kGcIpt], fp[kGcIpt];
#pragma unroll
    for (int j = 0; j < kGcIpt; ++j) {
      k[j] = j < (int)cnt ? keys[i0 + j] : 0u;
      tp[j] = fp[j] = 0.0;
      if (j < (int)cnt) {
        if (as_payload)
          gen_terms_payload(vals[i0 + j], 1, tp[j], fp[j]);
        else
          gen_terms(target, (const double*)nullptr, vals[i0 + j], 1, tp[j], fp[j]);
      }
    }
    uint32_t prev = __shfl_up_sync(kFull, k[kGcIpt - 1], 1);
    if (lane == 0) prev = last;
    last = __shfl_sync(kFull, k[kGcIpt - 1], 31);
    GState mine = g_identity();
    uint32_t starts = 0;
#pragma unroll
    for (int j = 0; j < kGcIpt; ++j) {
      if (j < (int)cnt) {
        const bool st = (j == 0 && i0 == 0) || k[j] != prev;
        prev = k[j];
        if (st) {
          starts |= 1u << j;
          mine.F = 1;
          mine.Mtp = mine.Stp;
          mine.Mfp = mine.Sfp;
        }
        mine.Stp += tp[j];
        mine.Sfp += fp[j];
      }
    }
    GState excl, total;
    g_warp_scan(mine, lane, excl, total);
    const GState g = g_combine(carry, excl);
    double ctp = g.Stp, cfp = g.Sfp, gtp = g.Mtp, gfp = g.Mfp;
#pragma unroll
    for (int j = 0; j < kGcIpt; ++j) {
      if (j < (int)cnt) {
        if (starts & (1u << j)) gtp = ctp, gfp = cfp;
        if (tp[j] != 0.0) acc += (ptot - gtp) / (ntot - gtp - gfp);
        ctp += tp[j];
        cfp += fp[j];
      }
    }
    carry = g_combine(carry, total);
  }
  acc = warp_sum(acc);
  if (lane == 0) partial[chunk] = acc;
}

struct GenPlan {
  RankPlan base;
  size_t off_gzero, gzero_bytes, off_gs, off_agg, off_inc, off_partial, off_chunks, total;
  uint32_t tiles;
};
GenPlan make_gen_plan(int64_t n) {
  GenPlan g{};
  g.base = make_plan(n, true, true, false);
  g.tiles = (uint32_t)((n + kGenTile - 1) / kGenTile);
  if (g.tiles == 0) g.tiles = 1;
  size_t o = g.base.total;
  auto take = [&](size_t bytes) {
    const size_t at = o;
    o += align_up(bytes, 256);
    return at;
  };
  g.off_gzero = o;
  g.off_gs = take(sizeof(GenScalars));
  g.off_agg = take(sizeof(GenRec));  // (unused since the sweeps became chunked; kept for layout)
  g.off_inc = take(sizeof(GenRec));  // totals of the scan, read by the final reduce
  g.gzero_bytes = o - g.off_gzero;
  g.off_partial = take((size_t)(g.tiles > (uint32_t)kMaxChunks ? g.tiles : (uint32_t)kMaxChunks) * 8);
  g.off_chunks = take((size_t)kMaxChunks * sizeof(GenChunkRec));
  g.total = o;
  return g;
}

// nb > 1: nb independent tasks in the same launches (grid.y = task): task y reads its inputs
// y * in_stride elements further on, works in the scratch slice y * bstride bytes further on and
// writes out[8 y ..] — WindowedBinaryAUROC.compute and (tasks, n) inputs with weights
// (window/auroc.py:176-185) are one launch train instead of a Python loop over tasks.
template <typename TargetT>
int binary_rank_general_t(const float* score, const TargetT* target, const double* weight, int64_t n,
                          int mode, double* out, void* scratch, const GenPlan& g, int sms,
                          cudaStream_t st, int nb = 1, size_t bstride = 0, size_t in_stride = 0) {
  const RankPlan& p = g.base;
  if (nb == 1) {
    TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
    TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, g.off_gzero), 0, g.gzero_bytes, st));
  } else {  // the same two regions of every task's slice
    TE_CUDA_TRY(cudaMemset2DAsync(at<char>(scratch, p.off_zero), bstride, 0, p.zero_bytes, nb, st));
    TE_CUDA_TRY(cudaMemset2DAsync(at<char>(scratch, g.off_gzero), bstride, 0, g.gzero_bytes, nb, st));
  }
  uint32_t* ka = at<uint32_t>(scratch, p.off_keys_a);
  uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
  uint32_t* va = at<uint32_t>(scratch, p.off_vals_a);
  uint32_t* vb = at<uint32_t>(scratch, p.off_vals_b);
  GenScalars* gs = at<GenScalars>(scratch, g.off_gs);
  GenRec* inc = at<GenRec>(scratch, g.off_inc);
  double* partial = at<double>(scratch, g.off_partial);
  uint32_t tiles = 0;
  if (n > 0) {
    tiles = g.tiles;
    {  // totals first: it also decides whether the target values can ride along as the payload
      TE_PROF("rank_totals", st);
      const uint32_t grid = (uint32_t)min((int64_t)sms * 8, (n + 255) / 256);
      rank_totals_kernel<TargetT><<<dim3(grid, nb), 256, 0, st>>>(target, weight, p.n, mode, gs, bstride, in_stride);
    }
    TE_LAUNCH_CHECK();
    int rc = launch_ingest<kIndexed, TargetT>(score, target, nullptr, p, ka, va, scratch, sms, st, &gs->payload_bad,
                                              weight, nb, bstride, in_stride);
    if (rc) return rc;
    rc = run_passes(p, scratch, ka, kb, ka, va, vb, va, st, nb, bstride);  // sorted pairs end in (ka, va)
    if (rc) return rc;
    {
      // AUROC is linear in the carried prefix: warp-private chunks + one scan CTA, no look-back.
      // AUPRC is not (a ratio of prefix sums): the same count sweep, the scan leaves every chunk's
      // starting state in place of its record, and a second sweep integrates.
      const ChunkGeom cg = gen_chunk_geom(p.n, sms);
      const uint32_t grid = (cg.nchunks + kGcBlock / kWarp - 1) / (kGcBlock / kWarp);
      GenChunkRec* crec = at<GenChunkRec>(scratch, g.off_chunks);
      {
        TE_PROF("gen_chunk_sweep", st);
        gen_chunk_kernel<TargetT><<<dim3(grid, nb), kGcBlock, 0, st>>>(ka, va, target, weight, p.n, cg.steps_per_chunk,
                                                                       cg.nchunks, gs, crec, bstride, in_stride);
      }
      {
        TE_PROF("gen_chunk_scan", st);
        gen_chunk_scan_kernel<<<dim3(1, nb), 1024, 0, st>>>(crec, cg.nchunks, partial, inc, mode == 0 ? 0 : 1, bstride);
      }
      tiles = 1;  // what the reduce below sums over
      if (mode != 0) {
        if (nb != 1) return TE_ERR_UNSUPPORTED;  // (only the AUROC sweep is batched)
        TE_PROF("gen_chunk_integrate", st);
        gen_chunk_integrate_kernel<TargetT><<<grid, kGcBlock, 0, st>>>(ka, va, target, p.n, cg.steps_per_chunk,
                                                                       cg.nchunks, gs, crec, partial);
        tiles = cg.nchunks;
      }
    }
    TE_LAUNCH_CHECK();
  }
  TE_PROF("rank_reduce_general", st);
  rank_reduce_general_kernel<<<dim3(1, nb), 1024, 0, st>>>(partial, tiles, mode, inc, gs, out, bstride);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

int binary_rank_general(const float* score, const void* target, int32_t dt, const double* weight,
                        int64_t n, int mode, double* out, void* scratch, size_t scratch_bytes,
                        cudaStream_t st, int64_t tasks = 1, int64_t ld = 0) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (tasks < 1 || tasks > 65535 || (tasks > 1 && (mode != 0 || ld < n))) return TE_ERR_UNSUPPORTED;
  const GenPlan g = make_gen_plan(n);
  const size_t bstride = tasks > 1 ? align_up(g.total, 256) : 0;
  if (!scratch || scratch_bytes < (tasks > 1 ? bstride * (size_t)tasks : g.total)) return TE_ERR_SCRATCH_TOO_SMALL;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  const int nb = (int)tasks;
  const size_t is = tasks > 1 ? (size_t)ld : 0;
  switch (dt) {
    case TE_DTYPE_I64:
      return binary_rank_general_t(score, (const long long*)target, weight, n, mode, out, scratch, g, sms, st, nb, bstride, is);
    case TE_DTYPE_I32:
      return binary_rank_general_t(score, (const int*)target, weight, n, mode, out, scratch, g, sms, st, nb, bstride, is);
    case TE_DTYPE_U8:
      return binary_rank_general_t(score, (const unsigned char*)target, weight, n, mode, out, scratch, g, sms, st, nb, bstride, is);
    case TE_DTYPE_F32:
      return binary_rank_general_t(score, (const float*)target, weight, n, mode, out, scratch, g, sms, st, nb, bstride, is);
    case TE_DTYPE_F64:
      return binary_rank_general_t(score, (const double*)target, weight, n, mode, out, scratch, g, sms, st, nb, bstride, is);
    default:
      return TE_ERR_INVALID_ARGUMENT;
  }
}


// ------------------------------------------------------------------------------------------
// pieces of the distributed sample sort (torcheval_b200/distributed.py)
// ------------------------------------------------------------------------------------------
__global__ void rank_info_kernel(const RankScalars* __restrict__ scal, long long* info) {
  if (threadIdx.x == 0) {
    info[0] = (long long)scal->positives;
    info[1] = (long long)scal->flags;
  }
}

// pack for the distributed path: (score, target) -> packed keys, #positives, flags, and — the
// splitters are already known — the sizes of the destination buckets, all in one streaming pass
// (12 B read + 4 B written per score).
// HIST: also the four radix-digit histograms of every destination bucket's keys, so that the
// receiving rank can start sorting without reading its inbox once more (the ranks sum the
// rows they are sent).  Bins per bucket in shared memory: digits 0-2 plain (uniform mantissa
// bits), digit 3 — the exponent byte, where half of the keys share a value — replicated 8 ways
// by lane so that a warp's same-address atomics stay short.
constexpr int kPackTopRepl = 8;
constexpr int kPackBinsPerBucket = 3 * kRadix + kRadix * kPackTopRepl;  // 2816 words = 11 KB

template <typename TargetT, bool HIST>
__global__ void __launch_bounds__(1024)
    pack_bucket_kernel(const float* __restrict__ score, const TargetT* __restrict__ target, uint32_t n,
                       bool vec_ok, const SplitterDigit dop_in, const uint32_t* __restrict__ splitters_dev,
                       uint32_t* __restrict__ keys_out, uint32_t* hist, RankScalars* scal,
                       uint32_t* __restrict__ digit_hist) {
  __shared__ uint32_t sh[16 * kWarp];
  __shared__ uint32_t s_sp[kMaxSplitters + 1];
  extern __shared__ uint32_t s_dh[];  // HIST: [buckets][kPackBinsPerBucket]
  const unsigned tid = threadIdx.x, lane = tid & 31;
  if (tid < 16 * kWarp) sh[tid] = 0;
  if (HIST)
    for (uint32_t i = tid; i < (uint32_t)(dop_in.nsp + 1) * kPackBinsPerBucket; i += blockDim.x) s_dh[i] = 0;
  // splitters either came with the launch (host array) or are read from device memory, so that a
  // caller can choose them on the device without a round trip through the host
  if (tid < kMaxSplitters + 1)
    s_sp[tid] = (int)tid < dop_in.nsp ? (splitters_dev ? splitters_dev[tid] : dop_in.sp[tid]) : 0xffffffffu;
  __syncthreads();
  auto dop = [&](uint32_t k) { return dop_in.eval(k, s_sp); };
  uint32_t npos = 0, bad_label = 0, neg_score = 0, wide = 0;
  const uint32_t ngroups = (n + 3) / 4;
  for (uint32_t g = blockIdx.x * blockDim.x + tid; g < ngroups; g += gridDim.x * blockDim.x) {
    const uint32_t i0 = g * 4, cnt = min(4u, n - i0);
    float x[4];
    uint32_t lab[4];
    if (cnt == 4 && vec_ok) {
      const int4 a = ld_stream_v4(score + i0);
      x[0] = __int_as_float(a.x), x[1] = __int_as_float(a.y), x[2] = __int_as_float(a.z), x[3] = __int_as_float(a.w);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) x[j] = j < (int)cnt ? score[i0 + j] : 0.f;
    }
    bool done = false;
    if constexpr (sizeof(TargetT) == 8) {
      if (cnt == 4 && vec_ok) {
        const int4 t0 = ld_stream_v4(target + i0), t1 = ld_stream_v4(target + i0 + 2);
        TargetT tv[4];
        const unsigned long long u0 = ((unsigned long long)(uint32_t)t0.y << 32) | (uint32_t)t0.x;
        const unsigned long long u1 = ((unsigned long long)(uint32_t)t0.w << 32) | (uint32_t)t0.z;
        const unsigned long long u2 = ((unsigned long long)(uint32_t)t1.y << 32) | (uint32_t)t1.x;
        const unsigned long long u3 = ((unsigned long long)(uint32_t)t1.w << 32) | (uint32_t)t1.z;
        memcpy(&tv[0], &u0, 8), memcpy(&tv[1], &u1, 8), memcpy(&tv[2], &u2, 8), memcpy(&tv[3], &u3, 8);
#pragma unroll
        for (int j = 0; j < 4; ++j) lab[j] = LabelOf<TargetT>::get(tv[j], bad_label);
        done = true;
      }
    }
    if (!done) {
#pragma unroll
      for (int j = 0; j < 4; ++j) lab[j] = j < (int)cnt ? LabelOf<TargetT>::get(target[i0 + j], bad_label) : 0u;
    }
    uint32_t k[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t b = __float_as_uint(x[j]);
      if (b == 0x80000000u) b = 0u;
      neg_score |= b >> 31;
      wide |= (b >> 30) & 1u;  // score >= 2.0: the key uses bit 31
      k[j] = (b << 1) | lab[j];
      if (j < (int)cnt) {
        npos += lab[j];
        const uint32_t bkt = dop(k[j]);
        if (!HIST) atomicAdd(sh + bkt * kWarp + lane, 1u);  // HIST: the size is the sum of the digit-0 bins
        if (HIST) {
          uint32_t* bins = s_dh + bkt * kPackBinsPerBucket;
          atomicAdd(bins + (k[j] & 255u), 1u);
          atomicAdd(bins + kRadix + ((k[j] >> 8) & 255u), 1u);
          atomicAdd(bins + 2 * kRadix + ((k[j] >> 16) & 255u), 1u);
          atomicAdd(bins + 3 * kRadix + (k[j] >> 24) * kPackTopRepl + (lane & (kPackTopRepl - 1)), 1u);
        }
      }
    }
    if (cnt == 4) {
      st_stream_v4(keys_out + i0, make_int4((int)k[0], (int)k[1], (int)k[2], (int)k[3]));
    } else {
      for (uint32_t j = 0; j < cnt; ++j) keys_out[i0 + j] = k[j];
    }
  }
  __syncthreads();
  if (!HIST && tid < 16) {
    uint32_t s = 0;
    for (int r = 0; r < kWarp; ++r) s += sh[tid * kWarp + r];
    if (s) atomicAdd(hist + tid, s);
  }
  if (HIST) {  // bucket sizes: warp w sums the 256 digit-0 bins of bucket w
    const unsigned w = tid >> 5;
    if (w <= (unsigned)dop_in.nsp) {
      unsigned long long s = 0;
      for (int d = lane; d < kRadix; d += kWarp) s += s_dh[w * kPackBinsPerBucket + d];
      s = warp_sum(s);
      if (lane == 0 && s) atomicAdd(hist + w, (uint32_t)s);
    }
  }
  if (HIST) {  // digit_hist[bucket][pass][digit]
    const uint32_t total = (uint32_t)(dop_in.nsp + 1) * 4 * kRadix;
    for (uint32_t i = tid; i < total; i += blockDim.x) {
      const uint32_t bkt = i / (4 * kRadix), r = i % (4 * kRadix);
      const uint32_t* bins = s_dh + bkt * kPackBinsPerBucket;
      uint32_t v;
      if (r < 3 * kRadix) {
        v = bins[r];
      } else {
        v = 0;
#pragma unroll
        for (int q = 0; q < kPackTopRepl; ++q) v += bins[3 * kRadix + (r - 3 * kRadix) * kPackTopRepl + q];
      }
      if (v) atomicAdd(digit_hist + i, v);
    }
  }
  npos = warp_sum(npos);
  const uint32_t anybad = __any_sync(kFull, bad_label != 0);
  const uint32_t anyneg = __any_sync(kFull, neg_score != 0);
  const uint32_t anywide = __any_sync(kFull, wide != 0);
  if (lane == 0) {
    if (npos) atomicAdd(&scal->positives, (unsigned long long)npos);
    const uint32_t f = (anybad ? TE_RANK_FLAG_BAD_LABEL : 0) | (anyneg ? TE_RANK_FLAG_NEGATIVE_SCORE : 0) |
                       (anywide ? TE_RANK_FLAG_WIDE_SCORE : 0);
    if (f) atomicOr(&scal->flags, f);
  }
}

// A rank's digit histograms came with the keys (te_rank_pack digit_hist rows, summed over the
// senders): install them and derive the number of positives (bit 0 of the key = odd digit 0).
__global__ void __launch_bounds__(1024)
    install_hist_kernel(const uint32_t* __restrict__ hist_in, uint32_t* __restrict__ hist, RankScalars* scal) {
  const unsigned tid = threadIdx.x;
  const uint32_t v = hist_in[tid];
  hist[tid] = v;
  unsigned long long odd = (tid < (unsigned)kRadix && (tid & 1u)) ? v : 0u;
  odd = warp_sum(odd);
  if ((tid & 31) == 0 && odd) atomicAdd(&scal->positives, odd);
}

__global__ void counts_to_hist_kernel(const long long* __restrict__ counts, uint32_t* hist, int nb) {
  if ((int)threadIdx.x < nb) hist[threadIdx.x] = (uint32_t)counts[threadIdx.x];
}

// bucket histogram for the splitter partition: hist[b] = #{keys with bucket b}, b < 16
__global__ void __launch_bounds__(1024)
    bucket_hist_kernel(const uint32_t* __restrict__ keys, uint32_t n, const SplitterDigit dop,
                       uint32_t* hist) {
  __shared__ uint32_t sh[16 * kWarp];
  const unsigned tid = threadIdx.x, lane = tid & 31;
  if (tid < 16 * kWarp) sh[tid] = 0;
  __syncthreads();
  const uint32_t ngroups = (n + 3) / 4;
  for (uint32_t g = blockIdx.x * blockDim.x + tid; g < ngroups; g += gridDim.x * blockDim.x) {
    const uint32_t i0 = g * 4, cnt = min(4u, n - i0);
    uint32_t k[4];
    if (cnt == 4) {
      const int4 a = ld_stream_v4(keys + i0);
      k[0] = a.x, k[1] = a.y, k[2] = a.z, k[3] = a.w;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) k[j] = j < (int)cnt ? keys[i0 + j] : 0u;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < (int)cnt) atomicAdd(sh + dop(k[j]) * kWarp + lane, 1u);
  }
  __syncthreads();
  if (tid < 16) {
    uint32_t s = 0;
    for (int r = 0; r < kWarp; ++r) s += sh[tid * kWarp + r];
    if (s) atomicAdd(hist + tid, s);
  }
}

__global__ void copy_counts_kernel(const uint32_t* __restrict__ hist, long long* counts, int nb) {
  if ((int)threadIdx.x < nb) counts[threadIdx.x] = (long long)hist[threadIdx.x];
}


// ------------------------------------------------------------------------------------------
// K2 — one-vs-rest (multiclass / multilabel) ranking: batched transposing ingest
// ------------------------------------------------------------------------------------------
// Reference: functional/classification/auroc.py:213-235 (input.T.sort(dim=1) + target[indices] ==
// arange[:, None]) and auprc.py:279-347 / precision_recall_curve.py:296-310 (Python loop over
// labels).  The (n, C) score matrix is row-major, so a class is a strided column.  One thread
// takes one row and kOvrBatch = 8 consecutive classes (32 contiguous bytes = one sector, no read
// amplification), builds the eight packed keys (label = target[i] == c, or target[i, c] == 1) and
// writes them to eight class-major key arrays — coalesced across the threads of a warp — while
// the per-class digit histograms accumulate in shared memory.  Each class array then goes
// through the same radix passes and fused sweep as the binary metric.
constexpr int kOvrBatch = 8;
constexpr int kOvrBlock = 512;

// HIST: the four digit histograms of the LSD passes; the counting-stage path builds its own joint
// histogram from the keys (rank_upsweep_joint_kernel) and skips them — four shared atomics per key
// are what bounds this kernel otherwise.
template <bool MULTILABEL, bool HIST = true>
__global__ void __launch_bounds__(kOvrBlock)
    ovr_ingest_kernel(const float* __restrict__ input, int64_t ld, const long long* __restrict__ target,
                      int64_t ld_t, uint32_t n, int c0, int nb, bool vec_ok, char* scratch0,
                      size_t plan_stride, size_t off_keys, size_t off_hist, size_t off_scal) {
  extern __shared__ uint32_t sh[];  // [nb][4][256]
  const unsigned tid = threadIdx.x, lane = tid & 31;
  if (HIST)
    for (int i = tid; i < kOvrBatch * 4 * kRadix; i += kOvrBlock) sh[i] = 0;
  __syncthreads();
  uint32_t npos[kOvrBatch];
  uint32_t negmask = 0, badmask = 0;
#pragma unroll
  for (int j = 0; j < kOvrBatch; ++j) npos[j] = 0;
  for (uint32_t i = blockIdx.x * kOvrBlock + tid; i < n; i += gridDim.x * kOvrBlock) {
    const float* row = input + (int64_t)i * ld + c0;
    float x[kOvrBatch];
    if (vec_ok && nb == kOvrBatch) {
      const int4 a = ld_stream_v4(row), b = ld_stream_v4(row + 4);
      x[0] = __int_as_float(a.x), x[1] = __int_as_float(a.y), x[2] = __int_as_float(a.z), x[3] = __int_as_float(a.w);
      x[4] = __int_as_float(b.x), x[5] = __int_as_float(b.y), x[6] = __int_as_float(b.z), x[7] = __int_as_float(b.w);
    } else {
#pragma unroll
      for (int j = 0; j < kOvrBatch; ++j) x[j] = j < nb ? __ldg(row + j) : 0.f;
    }
    long long tcls = 0;
    if (!MULTILABEL) tcls = __ldg(target + i);
#pragma unroll
    for (int j = 0; j < kOvrBatch; ++j) {
      if (j < nb) {
        uint32_t lab;
        if (MULTILABEL) {
          const long long t = __ldg(target + (int64_t)i * ld_t + c0 + j);
          lab = t == 1 ? 1u : 0u;
        } else {
          lab = tcls == (long long)(c0 + j) ? 1u : 0u;
        }
        uint32_t bits = __float_as_uint(x[j]);
        if (bits == 0x80000000u) bits = 0u;
        negmask |= (bits >> 31) << j;
        const uint32_t k = (bits << 1) | lab;
        npos[j] += lab;
        reinterpret_cast<uint32_t*>(scratch0 + (size_t)j * plan_stride + off_keys)[i] = k;
        if (HIST) {
          uint32_t* h = sh + j * 4 * kRadix;
#pragma unroll
          for (int p = 0; p < 4; ++p) atomicAdd(h + p * kRadix + ((k >> (8 * p)) & 0xffu), 1u);
        }
      }
    }
  }
  __syncthreads();
  for (int i = tid; HIST && i < nb * 4 * kRadix; i += kOvrBlock) {
    const uint32_t v = sh[i];
    if (v) {
      const int j = i / (4 * kRadix);
      atomicAdd(reinterpret_cast<uint32_t*>(scratch0 + (size_t)j * plan_stride + off_hist) + (i - j * 4 * kRadix), v);
    }
  }
#pragma unroll
  for (int j = 0; j < kOvrBatch; ++j) {
    if (j < nb) {
      const uint32_t s = warp_sum(npos[j]);
      const uint32_t anyneg = __any_sync(kFull, (negmask >> j) & 1u);
      if (lane == 0) {
        RankScalars* sc = reinterpret_cast<RankScalars*>(scratch0 + (size_t)j * plan_stride + off_scal);
        if (s) atomicAdd(&sc->positives, (unsigned long long)s);
        if (anyneg) atomicOr(&sc->flags, (uint32_t)TE_RANK_FLAG_NEGATIVE_SCORE);
      }
    }
  }
  (void)badmask;
}

}  // namespace
}  // namespace te

using namespace te;

extern "C" size_t te_binary_rank_scratch_bytes(int64_t n) {
  if (n < 0 || n > 0x7fffffffLL) return 0;
  return make_plan(n, false, true, true, true).total;  // (a superset of the plan without the counting stage)
}

extern "C" int te_binary_auroc(const float* score, const void* target, int32_t target_dtype,
                               int64_t n, double* out, void* scratch, size_t scratch_bytes,
                               void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  return binary_rank_packed(score, target, target_dtype, n, 0, out, scratch, scratch_bytes,
                            (cudaStream_t)stream);
}

extern "C" int te_binary_auroc_variant(const float* score, const void* target, int32_t target_dtype,
                                       int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                       int32_t variant, void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  if (variant != TE_RANK_AUTO && variant != TE_RANK_SORT && variant != TE_RANK_COUNTING) return TE_ERR_INVALID_ARGUMENT;
  return binary_rank_packed(score, target, target_dtype, n, 0, out, scratch, scratch_bytes,
                            (cudaStream_t)stream, variant);
}

extern "C" int te_binary_auprc(const float* score, const void* target, int32_t target_dtype,
                               int64_t n, double* out, void* scratch, size_t scratch_bytes,
                               void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  return binary_rank_packed(score, target, target_dtype, n, 1, out, scratch, scratch_bytes,
                            (cudaStream_t)stream);
}

extern "C" int te_binary_auprc_variant(const float* score, const void* target, int32_t target_dtype,
                                       int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                       int32_t variant, void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  if (variant != TE_RANK_AUTO && variant != TE_RANK_SORT && variant != TE_RANK_COUNTING) return TE_ERR_INVALID_ARGUMENT;
  return binary_rank_packed(score, target, target_dtype, n, 1, out, scratch, scratch_bytes,
                            (cudaStream_t)stream, variant);
}

extern "C" size_t te_sort_u32_scratch_bytes(int64_t n, int32_t with_values) {
  if (n < 0 || n > 0x7fffffffLL) return 0;
  return make_plan(n, with_values != 0, true, false).total;
}

extern "C" int te_sort_u32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in,
                           uint32_t* vals_out, int64_t n, void* scratch, size_t scratch_bytes,
                           void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (n > 0 && (!keys_in || !keys_out)) return TE_ERR_INVALID_ARGUMENT;
  if ((vals_in == nullptr) != (vals_out == nullptr)) return TE_ERR_INVALID_ARGUMENT;
  const bool wv = vals_in != nullptr;
  const RankPlan p = make_plan(n, wv, true, false);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  if (n == 0) return TE_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  int rc = launch_ingest<kRawKeys, long long>(nullptr, nullptr, keys_in, p, nullptr, nullptr, scratch, sms, st);
  if (rc) return rc;
  // in -> tmp -> out -> tmp -> out
  return run_passes(p, scratch, keys_in, at<uint32_t>(scratch, p.off_keys_a), keys_out, vals_in,
                    wv ? at<uint32_t>(scratch, p.off_vals_a) : nullptr, vals_out, st);
}

// ---- dense ranks of 64-bit scores (rank_dense.cuh) ------------------------------------------------
namespace {
struct DensePlan {
  RankPlan sort;
  size_t off_buf[6], off_chunks, off_sort, total;
  uint32_t n, nchunks;
};
DensePlan make_dense_plan(int64_t n64) {
  DensePlan d;
  d.n = (uint32_t)n64;
  d.nchunks = (uint32_t)((n64 + kDrChunk - 1) / kDrChunk);
  d.sort = make_plan(n64, true, true, false);
  size_t off = 0;
  for (int b = 0; b < 6; ++b) {
    d.off_buf[b] = off;
    off += align_up((size_t)n64 * sizeof(uint32_t), 256);
  }
  d.off_chunks = off;
  off += align_up(((size_t)d.nchunks + 1) * sizeof(uint32_t), 256);
  d.off_sort = off;
  off += d.sort.total;
  d.total = off;
  return d;
}
// stable pair sort through the scratch of one RankPlan: (k_in, v_in) -> (k_out, v_out)
int sort_pairs_u32(const RankPlan& p, void* scratch, const uint32_t* k_in, uint32_t* k_out, const uint32_t* v_in,
                   uint32_t* v_out, int sms, cudaStream_t st) {
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  const int rc = launch_ingest<kRawKeys, long long>(nullptr, nullptr, k_in, p, nullptr, nullptr, scratch, sms, st);
  if (rc) return rc;
  return run_passes(p, scratch, k_in, at<uint32_t>(scratch, p.off_keys_a), k_out, v_in,
                    at<uint32_t>(scratch, p.off_vals_a), v_out, st);
}
}  // namespace

extern "C" size_t te_dense_rank_scratch_bytes(int64_t n) {
  if (n < 0 || n > 0x7fffffffLL) return 0;
  return make_dense_plan(n).total;
}

extern "C" int te_dense_rank(const void* x, int32_t dtype, int64_t n, uint32_t* rank_out, void* values_out,
                             uint32_t* n_distinct, void* scratch, size_t scratch_bytes, void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (dtype != TE_DTYPE_F64 && dtype != TE_DTYPE_I64) return TE_ERR_UNSUPPORTED;
  if (n > 0 && (!x || !rank_out)) return TE_ERR_INVALID_ARGUMENT;
  const DensePlan d = make_dense_plan(n);
  if (!scratch || scratch_bytes < d.total) return TE_ERR_SCRATCH_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    if (n_distinct) TE_CUDA_TRY(cudaMemsetAsync(n_distinct, 0, sizeof(uint32_t), st));
    return TE_OK;
  }
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  uint32_t* lo = at<uint32_t>(scratch, d.off_buf[0]);
  uint32_t* hi = at<uint32_t>(scratch, d.off_buf[1]);
  uint32_t* idx = at<uint32_t>(scratch, d.off_buf[2]);   // iota, later the final permutation
  uint32_t* tmp = at<uint32_t>(scratch, d.off_buf[3]);   // sorted low words / gathered words
  uint32_t* perm1 = at<uint32_t>(scratch, d.off_buf[4]);
  uint32_t* hi_s = at<uint32_t>(scratch, d.off_buf[5]);
  uint32_t* chunks = at<uint32_t>(scratch, d.off_chunks);
  void* sort_scratch = at<char>(scratch, d.off_sort);
  const uint32_t un = (uint32_t)n;
  const int64_t want_blocks = (n + kDrBlock - 1) / kDrBlock;
  const int grid = (int)(want_blocks < (int64_t)sms * 16 ? want_blocks : (int64_t)sms * 16);
  const unsigned long long* xx = (const unsigned long long*)x;
  {
    TE_PROF("dense_split", st);
    if (dtype == TE_DTYPE_F64)
      dense_split_kernel<true><<<grid, kDrBlock, 0, st>>>(xx, un, lo, hi, idx);
    else
      dense_split_kernel<false><<<grid, kDrBlock, 0, st>>>(xx, un, lo, hi, idx);
  }
  TE_LAUNCH_CHECK();
  int rc = sort_pairs_u32(d.sort, sort_scratch, lo, tmp, idx, perm1, sms, st);  // by the low word
  if (rc) return rc;
  {
    TE_PROF("dense_gather", st);
    dense_gather_kernel<<<grid, kDrBlock, 0, st>>>(hi, perm1, un, tmp);
  }
  TE_LAUNCH_CHECK();
  rc = sort_pairs_u32(d.sort, sort_scratch, tmp, hi_s, perm1, idx, sms, st);  // stable, by the high word
  if (rc) return rc;
  {
    TE_PROF("dense_gather", st);
    dense_gather_kernel<<<grid, kDrBlock, 0, st>>>(lo, idx, un, tmp);  // low words in sorted order
  }
  TE_LAUNCH_CHECK();
  {
    TE_PROF("dense_count", st);
    dense_count_kernel<<<d.nchunks, kDrBlock, 0, st>>>(hi_s, tmp, un, chunks);
  }
  TE_LAUNCH_CHECK();
  {
    TE_PROF("dense_scan", st);
    dense_scan_kernel<<<1, 1024, 0, st>>>(chunks, d.nchunks, n_distinct);
  }
  TE_LAUNCH_CHECK();
  {
    TE_PROF("dense_write", st);
    dense_write_kernel<<<d.nchunks, kDrBlock, 0, st>>>(hi_s, tmp, idx, un, chunks, rank_out, xx,
                                                       (unsigned long long*)values_out);
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" size_t te_binary_rank_general_scratch_bytes(int64_t n) {
  if (n < 0 || n > 0x7fffffffLL) return 0;
  return make_gen_plan(n).total;
}

extern "C" size_t te_binary_auroc_general_batched_scratch_bytes(int64_t tasks, int64_t n) {
  if (n < 0 || n > 0x7fffffffLL || tasks < 1 || tasks > 65535) return 0;
  return align_up(make_gen_plan(n).total, 256) * (size_t)tasks;
}

extern "C" int te_binary_auroc_general_batched(const float* score, const void* target, int32_t target_dtype,
                                               const double* weight, int64_t tasks, int64_t n, int64_t ld,
                                               double* out, void* scratch, size_t scratch_bytes, void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  if (tasks == 1)
    return binary_rank_general(score, target, target_dtype, weight, n, 0, out, scratch, scratch_bytes, (cudaStream_t)stream);
  return binary_rank_general(score, target, target_dtype, weight, n, 0, out, scratch, scratch_bytes,
                             (cudaStream_t)stream, tasks, ld);
}

extern "C" int te_binary_auroc_general(const float* score, const void* target, int32_t target_dtype,
                                       const double* weight, int64_t n, double* out, void* scratch,
                                       size_t scratch_bytes, void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  return binary_rank_general(score, target, target_dtype, weight, n, 0, out, scratch, scratch_bytes,
                             (cudaStream_t)stream);
}

extern "C" int te_binary_auprc_general(const float* score, const void* target, int32_t target_dtype,
                                       int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                       void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  return binary_rank_general(score, target, target_dtype, nullptr, n, 1, out, scratch, scratch_bytes,
                             (cudaStream_t)stream);
}

extern "C" size_t te_rank_pack_scratch_bytes(int64_t n) {
  if (n < 0 || n > 0x7fffffffLL) return 0;
  return make_plan(n, false, false, false).total;
}

static int fill_splitters(SplitterDigit& dop, const uint32_t* splitters_host, int32_t num_splitters) {
  if (num_splitters < 0 || num_splitters > kMaxSplitters) return TE_ERR_INVALID_ARGUMENT;
  if (num_splitters > 0 && !splitters_host) return TE_ERR_INVALID_ARGUMENT;
  dop = SplitterDigit{};
  dop.nsp = num_splitters;
  for (int i = 0; i < num_splitters; ++i) {
    if (i > 0 && splitters_host[i] < splitters_host[i - 1]) return TE_ERR_INVALID_ARGUMENT;
    dop.sp[i] = splitters_host[i];
  }
  return TE_OK;
}

template <typename TargetT>
static void launch_pack(const float* score, const void* target, const RankPlan& p, const SplitterDigit& dop,
                        const uint32_t* splitters_dev, uint32_t* keys_out, void* scratch, int sms,
                        cudaStream_t st, uint32_t* digit_hist) {
  const uint32_t ngroups = (p.n + 3) / 4;
  uint32_t grid = (ngroups + 1023) / 1024;
  if (grid > (uint32_t)sms * 2) grid = sms * 2;
  const bool vec_ok = ((((uintptr_t)score) | ((uintptr_t)target)) & 15) == 0;
  if (digit_hist) {
    const size_t smem = (size_t)(dop.nsp + 1) * kPackBinsPerBucket * sizeof(uint32_t);
    cudaFuncSetAttribute(pack_bucket_kernel<TargetT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (smem > 100 * 1024 && grid > (uint32_t)sms) grid = sms;  // one CTA per SM fits then
    pack_bucket_kernel<TargetT, true><<<grid, 1024, smem, st>>>(score, (const TargetT*)target, p.n, vec_ok, dop,
                                                                splitters_dev, keys_out, at<uint32_t>(scratch, p.off_hist),
                                                                at<RankScalars>(scratch, p.off_scal), digit_hist);
  } else {
    pack_bucket_kernel<TargetT, false><<<grid, 1024, 0, st>>>(score, (const TargetT*)target, p.n, vec_ok, dop,
                                                              splitters_dev, keys_out, at<uint32_t>(scratch, p.off_hist),
                                                              at<RankScalars>(scratch, p.off_scal), nullptr);
  }
}

static int rank_pack_impl(const float* score, const void* target, int32_t target_dtype, int64_t n,
                          const uint32_t* splitters_host, const uint32_t* splitters_dev, int32_t num_splitters,
                          uint32_t* keys_out, int64_t* info, int64_t* counts, uint32_t* digit_hist,
                          void* scratch, size_t scratch_bytes, void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (!info || !counts || (n > 0 && (!score || !target || !keys_out))) return TE_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)keys_out & 15) != 0) return TE_ERR_INVALID_ARGUMENT;
  SplitterDigit dop;
  if (splitters_dev) {
    if (num_splitters < 0 || num_splitters > kMaxSplitters) return TE_ERR_INVALID_ARGUMENT;
    dop = SplitterDigit{};
    dop.nsp = num_splitters;
  } else {
    const int rs = fill_splitters(dop, splitters_host, num_splitters);
    if (rs) return rs;
  }
  const RankPlan p = make_plan(n, false, false, false);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.off_scal + 256 - p.off_zero, st));
  if (digit_hist)
    TE_CUDA_TRY(cudaMemsetAsync(digit_hist, 0, (size_t)(num_splitters + 1) * 4 * kRadix * sizeof(uint32_t), st));
  if (n > 0) {
    TE_PROF("pack_bucket", st);
    switch (target_dtype) {
      case TE_DTYPE_I64: launch_pack<long long>(score, target, p, dop, splitters_dev, keys_out, scratch, sms, st, digit_hist); break;
      case TE_DTYPE_I32: launch_pack<int>(score, target, p, dop, splitters_dev, keys_out, scratch, sms, st, digit_hist); break;
      case TE_DTYPE_U8: launch_pack<unsigned char>(score, target, p, dop, splitters_dev, keys_out, scratch, sms, st, digit_hist); break;
      case TE_DTYPE_F32: launch_pack<float>(score, target, p, dop, splitters_dev, keys_out, scratch, sms, st, digit_hist); break;
      case TE_DTYPE_F64: launch_pack<double>(score, target, p, dop, splitters_dev, keys_out, scratch, sms, st, digit_hist); break;
      default: return TE_ERR_INVALID_ARGUMENT;
    }
    TE_LAUNCH_CHECK();
  }
  rank_info_kernel<<<1, 32, 0, st>>>(at<RankScalars>(scratch, p.off_scal), (long long*)info);
  copy_counts_kernel<<<1, 32, 0, st>>>(at<uint32_t>(scratch, p.off_hist), (long long*)counts, num_splitters + 1);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_rank_pack(const float* score, const void* target, int32_t target_dtype, int64_t n,
                            const uint32_t* splitters_host, int32_t num_splitters, uint32_t* keys_out,
                            int64_t* info, int64_t* counts, uint32_t* digit_hist, void* scratch,
                            size_t scratch_bytes, void* stream) {
  return rank_pack_impl(score, target, target_dtype, n, splitters_host, nullptr, num_splitters, keys_out, info,
                        counts, digit_hist, scratch, scratch_bytes, stream);
}

extern "C" int te_rank_pack_dsplit(const float* score, const void* target, int32_t target_dtype, int64_t n,
                                   const uint32_t* splitters_dev, int32_t num_splitters, uint32_t* keys_out,
                                   int64_t* info, int64_t* counts, uint32_t* digit_hist, void* scratch,
                                   size_t scratch_bytes, void* stream) {
  if (num_splitters > 0 && !splitters_dev) return TE_ERR_INVALID_ARGUMENT;
  static const uint32_t none = 0;  // num_splitters == 0: nothing is read through the pointer
  return rank_pack_impl(score, target, target_dtype, n, nullptr, splitters_dev ? splitters_dev : &none,
                        num_splitters, keys_out, info, counts, digit_hist, scratch, scratch_bytes, stream);
}

extern "C" size_t te_partition_u32_scratch_bytes(int64_t n) { return te_rank_pack_scratch_bytes(n); }

static int partition_impl(const uint32_t* keys_in, int64_t n, const uint32_t* splitters_host,
                          int32_t num_splitters, const int64_t* counts_in, uint32_t* keys_out,
                          const uint64_t* peer_ptrs_host, int64_t* counts, void* scratch,
                          size_t scratch_bytes, void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (!counts) return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!keys_in || (!keys_out && !peer_ptrs_host))) return TE_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)keys_in & 15) != 0) return TE_ERR_INVALID_ARGUMENT;
  SplitterDigit dop;
  const int rs = fill_splitters(dop, splitters_host, num_splitters);
  if (rs) return rs;
  PeerPtrs pp{};
  if (peer_ptrs_host) {
    for (int b = 0; b <= num_splitters; ++b) pp.p[b] = peer_ptrs_host[b];
    static const int vec = [] {
      const char* e = getenv("TE_B200_PEER_VEC");
      return e ? atoi(e) : 1;
    }();
    pp.vec = vec;
  }
  const RankPlan p = make_plan(n, false, false, false);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  uint32_t* hist = at<uint32_t>(scratch, p.off_hist);
  uint32_t* hot = hot_digits_from_env() ? at<uint32_t>(scratch, p.off_hot) : nullptr;
  if (counts_in) {  // bucket sizes already known (te_rank_pack computed them)
    counts_to_hist_kernel<<<1, 32, 0, st>>>((const long long*)counts_in, hist, num_splitters + 1);
  } else if (n > 0) {
    TE_PROF("bucket_hist", st);
    const uint32_t ngroups = (p.n + 3) / 4;
    uint32_t grid = (ngroups + 1023) / 1024;
    if (grid > (uint32_t)sms * 2) grid = sms * 2;
    bucket_hist_kernel<<<grid, 1024, 0, st>>>(keys_in, p.n, dop, hist);
  }
  TE_LAUNCH_CHECK();
  copy_counts_kernel<<<1, 32, 0, st>>>(hist, (long long*)counts, num_splitters + 1);
  TE_LAUNCH_CHECK();
  if (n == 0) return TE_OK;
  {
    TE_PROF("digit_scan", st);
    digit_scan_kernel<<<1, 1024, 0, st>>>(hist, 1, hot, 0);
  }
  TE_LAUNCH_CHECK();
  uint32_t* lb = at<uint32_t>(scratch, p.off_lookback);
  uint32_t* gp = at<uint32_t>(scratch, p.off_gprefix);
  const bool peer = peer_ptrs_host != nullptr;
  // kUnstable: the keys of a bucket are sorted afterwards, so their order inside it is free and
  // ranks can be handed out in arrival order (one ATOMS.ADD instead of ATOMS.OR + LDS + STS)
  TE_PROF(peer ? "onesweep_partition_peer" : "onesweep_partition", st);
#define TE_LAUNCH_PART(B, PEER)                                                                          \
  do {                                                                                                   \
    auto fn = onesweep_pass_kernel<uint32_t, false, B, SplitterDigit, kSortIptDefault, PEER, true>;      \
    const size_t smem = onesweep_smem_bytes<uint32_t, B>();                                              \
    TE_CUDA_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
    fn<<<p.sort_tiles, B, smem, st>>>(keys_in, keys_out, nullptr, nullptr, p.n, dop, hist, lb, gp, pp, hot, \
                                      nullptr, 0, nullptr, nullptr);                                     \
  } while (0)
  if (p.sort_ipt != kSortIptDefault) return TE_ERR_UNSUPPORTED;
  if (p.sort_block == 256) {
    if (peer) TE_LAUNCH_PART(256, true); else TE_LAUNCH_PART(256, false);
  } else {
    if (peer) TE_LAUNCH_PART(512, true); else TE_LAUNCH_PART(512, false);
  }
#undef TE_LAUNCH_PART
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_partition_u32(const uint32_t* keys_in, int64_t n, const uint32_t* splitters_host,
                                int32_t num_splitters, const int64_t* counts_in, uint32_t* keys_out,
                                int64_t* counts, void* scratch, size_t scratch_bytes, void* stream) {
  if (n > 0 && !keys_out) return TE_ERR_INVALID_ARGUMENT;
  return partition_impl(keys_in, n, splitters_host, num_splitters, counts_in, keys_out, nullptr, counts,
                        scratch, scratch_bytes, stream);
}

extern "C" int te_partition_scatter_u32(const uint32_t* keys_in, int64_t n, const uint32_t* splitters_host,
                                        int32_t num_splitters, const int64_t* counts_in,
                                        const uint64_t* dest_ptrs_host, int64_t* counts, void* scratch,
                                        size_t scratch_bytes, void* stream) {
  if (!dest_ptrs_host) return TE_ERR_INVALID_ARGUMENT;
  return partition_impl(keys_in, n, splitters_host, num_splitters, counts_in, nullptr, dest_ptrs_host, counts,
                        scratch, scratch_bytes, stream);
}

// mode 0: AUROC of the keys.  mode 1: the same, and the chunk states the AUPRC integrate sweep
// starts from are left in `scratch` next to the sorted keys (te_binary_auprc_from_keys_finish).
static int rank_from_keys(const uint32_t* keys, int64_t n, int mode, const uint32_t* digit_hist, double* out,
                          void* scratch, size_t scratch_bytes, void* stream, int variant = TE_RANK_SORT) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (!out || (n > 0 && !keys)) return TE_ERR_INVALID_ARGUMENT;
  if (((uintptr_t)keys & 15) != 0) return TE_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  if (mode == 0 && (variant == TE_RANK_COUNTING || (variant == TE_RANK_AUTO && n >= leaf_min_n_from_env()))) {
    // counting-stage path (rank_leaf.cuh); the senders' byte histograms are not the ones it needs
    const RankPlan p = make_plan(n, false, true, true, true);
    if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
    if (p.sort_block != 512 || p.sort_ipt != kSortIptDefault) return TE_ERR_UNSUPPORTED;
    TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
    uint32_t* ka = at<uint32_t>(scratch, p.off_keys_a);
    uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
    if (n > 0) {
      const int rc = leaf_passes_ranges<long long>(p, scratch, nullptr, nullptr, keys, ka, kb, sms, st);
      if (rc) return rc;
    }
    return finish_leaf(p, kb, out, scratch, sms, st);
  }
  const RankPlan p = make_plan(n, false, true, true);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  uint32_t* ka = at<uint32_t>(scratch, p.off_keys_a);
  uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
  if (n > 0) {
    int rc = TE_OK;
    if (digit_hist) {  // the histograms travelled with the keys: no extra read of the inbox
      TE_PROF("install_hist", st);
      install_hist_kernel<<<1, 4 * kRadix, 0, st>>>(digit_hist, at<uint32_t>(scratch, p.off_hist),
                                                    at<RankScalars>(scratch, p.off_scal));
    } else {
      rc = launch_ingest<kRawKeys, long long>(nullptr, nullptr, keys, p, nullptr, nullptr, scratch, sms, st);
    }
    if (rc) return rc;
    rc = run_passes(p, scratch, keys, ka, kb, nullptr, nullptr, nullptr, st);  // sorted keys end in kb
    if (rc) return rc;
  }
  const int rc = finish_packed(p, kb, 0, out, scratch, sms, st, mode == 1);
  if (rc || mode == 0 || n == 0) return rc;
  const ChunkGeom cg = chunk_geom(p.n, sms);
  TE_PROF("rank_chunk_scan", st);
  rank_chunk_scan_kernel<1><<<1, 1024, 0, st>>>(at<ChunkRec>(scratch, p.off_agg), cg.nchunks, p.n,
                                                at<RankScalars>(scratch, p.off_scal),
                                                at<FState>(scratch, p.off_inc), nullptr);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_binary_auroc_from_keys(const uint32_t* keys, int64_t n, const uint32_t* digit_hist,
                                         double* out, void* scratch, size_t scratch_bytes, void* stream) {
  return rank_from_keys(keys, n, 0, digit_hist, out, scratch, scratch_bytes, stream, TE_RANK_AUTO);
}

extern "C" int te_binary_auroc_from_keys_variant(const uint32_t* keys, int64_t n, const uint32_t* digit_hist,
                                                 double* out, void* scratch, size_t scratch_bytes, int32_t variant,
                                                 void* stream) {
  if (variant != TE_RANK_AUTO && variant != TE_RANK_SORT && variant != TE_RANK_COUNTING) return TE_ERR_INVALID_ARGUMENT;
  return rank_from_keys(keys, n, 0, digit_hist, out, scratch, scratch_bytes, stream, variant);
}

extern "C" int te_binary_auprc_from_keys_begin(const uint32_t* keys, int64_t n, const uint32_t* digit_hist,
                                               double* out, void* scratch, size_t scratch_bytes,
                                               void* stream) {
  return rank_from_keys(keys, n, 1, digit_hist, out, scratch, scratch_bytes, stream);
}

extern "C" int te_binary_auprc_from_keys_finish(int64_t n, double positives_above, double samples_above,
                                                const double* above_dev, double* out, void* scratch,
                                                size_t scratch_bytes, void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (!out) return TE_ERR_INVALID_ARGUMENT;
  if (!above_dev && (positives_above < 0.0 || samples_above < positives_above)) return TE_ERR_INVALID_ARGUMENT;
  const RankPlan p = make_plan(n, false, true, true);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  const ChunkGeom cg = chunk_geom(p.n, sms);
  const uint32_t nchunks = n > 0 ? cg.nchunks : 0;
  const uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
  unsigned long long* partial = at<unsigned long long>(scratch, p.off_partial);
  const RankScalars* scal = at<RankScalars>(scratch, p.off_scal);
  if (nchunks) {
    TE_PROF("rank_chunk_integrate", st);
    rank_chunk_kernel<1><<<(nchunks + kFinBlock / kWarp - 1) / (kFinBlock / kWarp), kFinBlock, 0, st>>>(
        kb, p.n, cg.steps_per_chunk, nchunks, nullptr, at<FState>(scratch, p.off_inc), scal, partial,
        positives_above, samples_above, above_dev, 0);
  }
  TE_PROF("rank_reduce", st);
  rank_reduce_kernel<1><<<1, 1024, 0, st>>>(partial, nchunks, p.n, scal, out);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

__global__ void pr_info_kernel(const uint32_t* starts, uint32_t nchunks, double* out) {
  if (threadIdx.x == 0) out[0] = (double)starts[nchunks];
}

extern "C" int te_binary_pr_curve_begin(const float* score, const void* target, int32_t target_dtype,
                                        int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                        void* stream) {
  if (!out || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  const RankPlan p = make_plan(n, false, true, true);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  uint32_t* ka = at<uint32_t>(scratch, p.off_keys_a);
  uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
  if (n > 0) {
    int rc = launch_ingest_dtype<kPacked>(score, target, target_dtype, p, ka, nullptr, scratch, sms, st);
    if (rc) return rc;
    rc = run_passes(p, scratch, ka, kb, ka, nullptr, nullptr, nullptr, st);  // sorted keys end in ka
    if (rc) return rc;
  }
  const int rc = finish_packed(p, ka, 0, out, scratch, sms, st, true);  // out[] as te_binary_auroc
  if (rc) return rc;
  const ChunkGeom cg = chunk_geom(p.n, sms);
  const uint32_t nchunks = n > 0 ? cg.nchunks : 0;
  uint32_t* starts = at<uint32_t>(scratch, p.off_starts);
  rank_chunk_scan_kernel<1><<<1, 1024, 0, st>>>(at<ChunkRec>(scratch, p.off_agg), nchunks, p.n,
                                                at<RankScalars>(scratch, p.off_scal),
                                                at<FState>(scratch, p.off_inc), nullptr, starts);
  pr_info_kernel<<<1, 32, 0, st>>>(starts, nchunks, out);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_binary_pr_curve_finish(int64_t n, float* precision, float* recall, float* thresholds,
                                         void* scratch, size_t scratch_bytes, void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (!precision || !recall || (n > 0 && !thresholds)) return TE_ERR_INVALID_ARGUMENT;
  const RankPlan p = make_plan(n, false, true, true);
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  if (n == 0) {  // no scores: the curve is the closing point alone
    const float one = 1.0f, zero = 0.0f;
    TE_CUDA_TRY(cudaMemcpyAsync(precision, &one, 4, cudaMemcpyHostToDevice, st));
    TE_CUDA_TRY(cudaMemcpyAsync(recall, &zero, 4, cudaMemcpyHostToDevice, st));
    return TE_OK;
  }
  const ChunkGeom cg = chunk_geom(p.n, sms);
  TE_PROF("rank_pr_curve", st);
  rank_pr_curve_kernel<<<(cg.nchunks + kFinBlock / kWarp - 1) / (kFinBlock / kWarp), kFinBlock, 0, st>>>(
      at<uint32_t>(scratch, p.off_keys_a), p.n, cg.steps_per_chunk, cg.nchunks, at<FState>(scratch, p.off_inc),
      at<uint32_t>(scratch, p.off_starts), at<RankScalars>(scratch, p.off_scal), precision, recall, thresholds);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_debug_phase_stats(uint64_t* out8, int reset) {
  unsigned long long h[8] = {};
  if (out8) {
    TE_CUDA_TRY(cudaMemcpyFromSymbol(h, g_phase_cycles, sizeof(h)));
    for (int i = 0; i < 8; ++i) out8[i] = h[i];
  }
  if (reset) {
    unsigned long long z[8] = {};
    TE_CUDA_TRY(cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z)));
  }
  return TE_OK;
}

extern "C" size_t te_ovr_rank_scratch_bytes(int64_t n) {
  if (n < 0 || n > 0x7fffffffLL) return 0;
  const size_t a = align_up(make_plan(n, false, true, true).total, 256);
  const size_t b = align_up(make_plan(n, false, true, true, true).total, 256);  // counting-stage path
  return (a > b ? a : b) * kOvrBatch;
}

extern "C" int te_ovr_rank(const float* input, int64_t ld, const int64_t* target, int64_t target_ld,
                           int64_t n, int32_t num_classes, int32_t multilabel, int32_t mode, double* out,
                           void* scratch, size_t scratch_bytes, void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (num_classes < 1 || !out || (mode != 0 && mode != 1) || ld < num_classes) return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!input || !target)) return TE_ERR_INVALID_ARGUMENT;
  if (multilabel && target_ld < num_classes) return TE_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  // AUROC of long columns: the counting-stage path of the binary metric, the classes of a batch in
  // the same launches (grid.y).  Its fixed work per class (65 536 segments to plan and walk) pays
  // off from TE_B200_OVR_LEAF_MIN_N samples on.
  static const int64_t ovr_leaf_min_n = [] {
    const char* e = getenv("TE_B200_OVR_LEAF_MIN_N");
    return e ? atoll(e) : (int64_t)1 << 25;  // 2^24: 22.6 vs 21.5 ms, 2^25: 8.1 vs 9.8, 2^26: 13.5 vs 19.4 (profiles/r02_ovr_bench.txt)
  }();
  if (mode == 0 && n >= ovr_leaf_min_n && (((uintptr_t)scratch) & 255) == 0) {
    const RankPlan p = make_plan(n, false, true, true, true);
    const size_t stride = align_up(p.total, 256);
    if (!scratch || scratch_bytes < stride * kOvrBatch) return TE_ERR_SCRATCH_TOO_SMALL;
    char* base = static_cast<char*>(scratch);
    for (int c0 = 0; c0 < num_classes; c0 += kOvrBatch) {
      const int nb = num_classes - c0 < kOvrBatch ? num_classes - c0 : kOvrBatch;
      TE_CUDA_TRY(cudaMemset2DAsync(base + p.off_zero, stride, 0, p.zero_bytes, nb, st));
      const bool vec_ok = (((uintptr_t)input) & 15) == 0 && (ld % 4) == 0 && (c0 % 4) == 0;
      int64_t want = (n + kOvrBlock - 1) / kOvrBlock;
      const int grid = (int)(want < (int64_t)sms * 4 ? want : (int64_t)sms * 4);
      {
        TE_PROF("ovr_ingest_keys", st);  // packed keys of class c0 + j -> keys_b of slice j
        if (multilabel)
          ovr_ingest_kernel<true, false><<<grid, kOvrBlock, 0, st>>>(input, ld, (const long long*)target, target_ld, p.n, c0, nb,
                                                                     vec_ok, base, stride, p.off_keys_b, p.off_hist, p.off_scal);
        else
          ovr_ingest_kernel<false, false><<<grid, kOvrBlock, 0, st>>>(input, ld, (const long long*)target, target_ld, p.n, c0, nb,
                                                                      vec_ok, base, stride, p.off_keys_b, p.off_hist, p.off_scal);
      }
      TE_LAUNCH_CHECK();
      uint32_t* ka = at<uint32_t>(base, p.off_keys_a);
      uint32_t* kb = at<uint32_t>(base, p.off_keys_b);
      // raw keys in kb -> ka -> kb (grouped by their top 16 bits), then the counting stage
      int rc = leaf_passes_msd<long long>(p, base, nullptr, nullptr, kb, ka, kb, sms, st, SlabView{nullptr, 0, 0, 0}, nullptr,
                                          nb, stride);
      if (rc) return rc;
      rc = finish_leaf(p, kb, out + (size_t)c0 * 8, base, sms, st, nullptr, true, nb, stride);
      if (rc) return rc;
    }
    return TE_OK;
  }
  const RankPlan p = make_plan(n, false, true, true);
  const size_t stride = align_up(p.total, 256);
  if (!scratch || scratch_bytes < stride * kOvrBatch) return TE_ERR_SCRATCH_TOO_SMALL;
  const size_t smem = (size_t)kOvrBatch * 4 * kRadix * sizeof(uint32_t);
  char* base = static_cast<char*>(scratch);
  for (int c0 = 0; c0 < num_classes; c0 += kOvrBatch) {
    const int nb = num_classes - c0 < kOvrBatch ? num_classes - c0 : kOvrBatch;
    TE_CUDA_TRY(cudaMemset2DAsync(base + p.off_zero, stride, 0, p.zero_bytes, nb, st));  // every slab's control block
    if (n > 0) {
      const bool vec_ok = (((uintptr_t)input) & 15) == 0 && (ld % 4) == 0 && (c0 % 4) == 0;
      int64_t want = (n + kOvrBlock - 1) / kOvrBlock;
      const int grid = (int)(want < (int64_t)sms * 2 ? want : (int64_t)sms * 2);
      TE_PROF("ovr_ingest", st);
      if (multilabel) {
        TE_CUDA_TRY(cudaFuncSetAttribute(ovr_ingest_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ovr_ingest_kernel<true><<<grid, kOvrBlock, smem, st>>>(input, ld, (const long long*)target, target_ld, p.n, c0,
                                                                 nb, vec_ok, base, stride, p.off_keys_a, p.off_hist, p.off_scal);
      } else {
        TE_CUDA_TRY(cudaFuncSetAttribute(ovr_ingest_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ovr_ingest_kernel<false><<<grid, kOvrBlock, smem, st>>>(input, ld, (const long long*)target, target_ld, p.n, c0,
                                                                  nb, vec_ok, base, stride, p.off_keys_a, p.off_hist, p.off_scal);
      }
      TE_LAUNCH_CHECK();
    }
    // the nb classes of the batch are sorted and swept by the same launches (grid.y = nb)
    uint32_t* ka = at<uint32_t>(base, p.off_keys_a);
    uint32_t* kb = at<uint32_t>(base, p.off_keys_b);
    if (n > 0) {
      const int rc = run_passes(p, base, ka, kb, ka, nullptr, nullptr, nullptr, st, nb, stride);
      if (rc) return rc;
    }
    const int rc = finish_packed(p, ka, mode, out + (size_t)c0 * 8, base, sms, st, false, nb, stride);
    if (rc) return rc;
  }
  return TE_OK;
}

// ------------------------------------------------------------------------------------------
// Small single-CTA helpers of the distributed sample sort.  Each replaces a dozen or more tiny
// framework kernels whose launch gaps, not their work, were 10 % of an 8-GPU step
// (profiles/r01_dist_trace_4gpu.txt).
// ------------------------------------------------------------------------------------------
namespace te {
namespace {

// samples: float bit patterns of regularly sampled scores from every rank (negative = "this rank
// has no data").  Bitonic sort in shared memory, then W-1 quantile splitters in KEY space
// ((bits << 1), label bit clear), as uint32 and again as int64 (for the host message).
__global__ void __launch_bounds__(1024)
    select_splitters_kernel(const int32_t* __restrict__ samples, uint32_t total, uint32_t padded, int world,
                            uint32_t* __restrict__ sp32, long long* __restrict__ sp64) {
  extern __shared__ uint32_t s_pool[];
  __shared__ uint32_t s_valid;
  const unsigned tid = threadIdx.x;
  if (tid == 0) s_valid = 0;
  __syncthreads();
  uint32_t mine = 0;
  for (uint32_t i = tid; i < padded; i += blockDim.x) {
    const int32_t v = i < total ? samples[i] : -1;
    s_pool[i] = v >= 0 ? (uint32_t)v : 0xffffffffu;
    mine += v >= 0 ? 1u : 0u;
  }
  mine = (uint32_t)warp_sum((unsigned long long)mine);
  if ((tid & 31) == 0 && mine) atomicAdd(&s_valid, mine);
  __syncthreads();
  for (uint32_t k = 2; k <= padded; k <<= 1) {
    for (uint32_t j = k >> 1; j > 0; j >>= 1) {
      for (uint32_t t = tid; t < padded / 2; t += blockDim.x) {
        const uint32_t i = 2 * t - (t & (j - 1));  // element whose partner is i + j
        const uint32_t a = s_pool[i], b = s_pool[i + j];
        const bool up = (i & k) == 0;
        if ((a > b) == up) s_pool[i] = b, s_pool[i + j] = a;
      }
      __syncthreads();
    }
  }
  const uint32_t m = s_valid;
  if (tid + 1 < (unsigned)world) {
    uint32_t v = 0;
    if (m) {
      uint32_t at = (uint32_t)(((unsigned long long)(tid + 1) * m) / (unsigned)world);
      if (at > m - 1) at = m - 1;
      v = (s_pool[at] << 1) & 0xfffffffeu;
    }
    sp32[tid] = v;
    sp64[tid] = (long long)v;
  }
}

// rows: the float64[8] results of te_binary_auroc_from_keys of every rank, in rank (= ascending
// score range) order.  2U = sum_r (2U_r + 2 P_r sum_{q<r} N_q), exact in uint64.
__global__ void combine_auroc_kernel(const double* __restrict__ rows, int world, double n_recv, double* out) {
  if (threadIdx.x != 0) return;
  unsigned long long u2 = 0, neg_below = 0, P = 0, N = 0;
  for (int r = 0; r < world; ++r) {
    const double* w = rows + 8 * r;
    const unsigned long long p = (unsigned long long)w[1], n = (unsigned long long)w[2];
    const unsigned long long u = ((unsigned long long)w[5] << 32) | (unsigned long long)w[6];
    u2 += u + 2ull * p * neg_below;
    neg_below += n;
    P += p;
    N += n;
  }
  out[0] = (P == 0 || N == 0) ? 0.5 : ((double)u2 * 0.5) / ((double)P * (double)N);
  out[1] = (double)P;
  out[2] = (double)N;
  out[3] = (double)u2;
  out[4] = 0.0;
  out[5] = n_recv;
  out[6] = (double)(u2 >> 32);            // 2U exactly, as two 32-bit halves
  out[7] = (double)(u2 & 0xffffffffull);
}

}  // namespace
}  // namespace te

// ------------------------------------------------------------------------------------------
// Slab exchange of the distributed sample sort (rank_exchange.cuh)
// ------------------------------------------------------------------------------------------
extern "C" size_t te_slab_inbox_bytes(int32_t world, int64_t cap) {
  if (world < 1 || world > kMaxSplitters + 1 || cap < 0) return 0;
  return 1024 + (size_t)world * (size_t)cap * 4;  // header (world x {count, flags}, padded) + the slabs
}

extern "C" int64_t te_slab_capacity(int64_t n_max_per_rank, int32_t world) {
  // room for 1.3x an even split of the largest shard + a tile, rounded up to whole downsweep tiles
  if (world < 1 || n_max_per_rank < 0) return 0;
  const int64_t want = (int64_t)(1.3 * (double)n_max_per_rank / world) + 2 * kRangeUnit;
  return (want + kRangeUnit - 1) / kRangeUnit * kRangeUnit;
}

template <typename TargetT>
static int pack_scatter_impl(const float* score, const TargetT* target, int64_t n, const uint32_t* splitters_dev,
                             int32_t world, int32_t rank, const uint64_t* inbox_ptrs_host, int64_t cap,
                             uint32_t* work /*[32] device words*/, cudaStream_t st) {
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(work, 0, 32 * sizeof(uint32_t), st));
  uint32_t* cursors = work;       // [16]
  uint32_t* flags = work + 16;    // [1]
  PeerSlabs slabs{};
  PeerHeaders hdrs{};
  for (int b = 0; b < world; ++b) {
    slabs.slab[b] = inbox_ptrs_host[b] + 1024 + (unsigned long long)rank * (unsigned long long)cap * 4ull;
    hdrs.hdr[b] = inbox_ptrs_host[b] + 8ull * (unsigned long long)rank;
  }
  if (n > 0) {
    const uint32_t tiles = (uint32_t)((n + kXsTile - 1) / kXsTile);
    int per_sm = 0;
    TE_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pack_scatter_kernel<TargetT>, kXsBlock, 0));
    if (per_sm < 1) per_sm = 1;
    const uint32_t grid = tiles < (uint32_t)(per_sm * sms) ? tiles : (uint32_t)(per_sm * sms);
    const bool vec_ok = (((uintptr_t)score | (uintptr_t)target) & 15) == 0;
    TE_PROF("pack_scatter_peer", st);
    pack_scatter_kernel<TargetT><<<grid, kXsBlock, 0, st>>>(score, target, (uint32_t)n, vec_ok, splitters_dev, world - 1,
                                                            slabs, (uint32_t)cap, cursors, flags);
  }
  TE_LAUNCH_CHECK();
  publish_counts_kernel<<<1, 32, 0, st>>>(cursors, flags, hdrs, world, (uint32_t)cap);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_rank_pack_scatter(const float* score, const void* target, int32_t target_dtype, int64_t n,
                                    const uint32_t* splitters_dev, int32_t world, int32_t rank,
                                    const uint64_t* inbox_ptrs_host, int64_t cap, void* work, void* stream) {
  if (n < 0 || n > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (world < 1 || world > kMaxSplitters + 1 || rank < 0 || rank >= world || !inbox_ptrs_host || !work)
    return TE_ERR_INVALID_ARGUMENT;
  if (cap <= 0 || cap % kRangeUnit != 0 || cap > 0x7fffffffLL || (n > 0 && (!score || !target))) return TE_ERR_INVALID_ARGUMENT;
  if (world > 1 && !splitters_dev) return TE_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* w = static_cast<uint32_t*>(work);
  switch (target_dtype) {
    case TE_DTYPE_I64: return pack_scatter_impl(score, (const long long*)target, n, splitters_dev, world, rank, inbox_ptrs_host, cap, w, st);
    case TE_DTYPE_I32: return pack_scatter_impl(score, (const int*)target, n, splitters_dev, world, rank, inbox_ptrs_host, cap, w, st);
    case TE_DTYPE_U8: return pack_scatter_impl(score, (const unsigned char*)target, n, splitters_dev, world, rank, inbox_ptrs_host, cap, w, st);
    case TE_DTYPE_F32: return pack_scatter_impl(score, (const float*)target, n, splitters_dev, world, rank, inbox_ptrs_host, cap, w, st);
    case TE_DTYPE_F64: return pack_scatter_impl(score, (const double*)target, n, splitters_dev, world, rank, inbox_ptrs_host, cap, w, st);
    default: return TE_ERR_INVALID_ARGUMENT;
  }
}

// measurement only (tools/p2p_store_bench.py): raw store bandwidth into up to 16 destinations
extern "C" int te_debug_p2p_store_bench(const uint64_t* dst_ptrs_host, int32_t ndst, int64_t words_per_dst, int32_t mode,
                                        int32_t grid, void* stream) {
  if (!dst_ptrs_host || ndst < 1 || ndst > kMaxSplitters + 1 || words_per_dst < 2048 || grid < 1) return TE_ERR_INVALID_ARGUMENT;
  PeerSlabs ps{};
  for (int b = 0; b < ndst; ++b) ps.slab[b] = dst_ptrs_host[b];
  p2p_store_bench_kernel<<<grid, 512, 0, (cudaStream_t)stream>>>(ps, ndst, (uint32_t)words_per_dst, mode);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" size_t te_binary_auroc_from_slabs_scratch_bytes(int32_t world, int64_t cap) {
  if (world < 1 || world > kMaxSplitters + 1 || cap <= 0 || (int64_t)world * cap > 0x7fffffffLL) return 0;
  return make_plan((int64_t)world * cap, false, true, true, true).total;
}

extern "C" int te_binary_auroc_from_slabs(const void* inbox, int32_t world, int64_t cap, double* out, void* scratch,
                                          size_t scratch_bytes, void* stream) {
  if (!inbox || !out || world < 1 || world > kMaxSplitters + 1 || cap <= 0 || cap % kRangeUnit != 0)
    return TE_ERR_INVALID_ARGUMENT;
  const int64_t n_cap = (int64_t)world * cap;
  if (n_cap > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  const RankPlan p = make_plan(n_cap, false, true, true, true);  // sized for full slabs
  if (!scratch || scratch_bytes < p.total) return TE_ERR_SCRATCH_TOO_SMALL;
  if (p.sort_block != 512 || p.sort_ipt != kSortIptDefault) return TE_ERR_UNSUPPORTED;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  TE_CUDA_TRY(cudaMemsetAsync(at<char>(scratch, p.off_zero), 0, p.zero_bytes, st));
  const uint32_t* hdr = static_cast<const uint32_t*>(inbox);
  const uint32_t* keys_in = reinterpret_cast<const uint32_t*>(static_cast<const char*>(inbox) + 1024);
  uint32_t* ka = at<uint32_t>(scratch, p.off_keys_a);
  uint32_t* kb = at<uint32_t>(scratch, p.off_keys_b);
  uint32_t* hist = at<uint32_t>(scratch, p.off_hist);
  uint32_t* range_hist = at<uint32_t>(scratch, p.off_range_hist);
  uint32_t* range_base = at<uint32_t>(scratch, p.off_range_base);
  RankScalars* scal = at<RankScalars>(scratch, p.off_scal);
  uint32_t* flags = &scal->flags;
  uint32_t* n_dev = &scal->pad;  // the number of records received, summed on the device
  uint32_t* skip = skip_passes_from_env() ? at<uint32_t>(scratch, p.off_ticket) : nullptr;
  slab_totals_kernel<<<1, 32, 0, st>>>(hdr, world, n_dev, flags);
  TE_LAUNCH_CHECK();
  const SlabView view{hdr, 2u, (uint32_t)cap, (uint32_t)world};
  // upsweep over the slabs (joint histogram) -> cursors -> two cursor-placed passes -> counting stage
  const int rc = leaf_passes_msd<long long>(p, scratch, nullptr, nullptr, keys_in, ka, kb, sms, st, view, n_dev);
  if (rc) return rc;
  return finish_leaf(p, kb, out, scratch, sms, st, n_dev, true);
}

extern "C" int te_select_splitters(const int32_t* samples, int64_t total, int32_t world, uint32_t* splitters,
                                   int64_t* splitters_i64, void* stream) {
  if (world < 1 || world > kMaxSplitters + 1 || total < 0 || total > 32768) return TE_ERR_INVALID_ARGUMENT;
  if (world > 1 && (!splitters || !splitters_i64 || (total > 0 && !samples))) return TE_ERR_INVALID_ARGUMENT;
  if (world == 1) return TE_OK;
  uint32_t padded = 2;
  while (padded < (uint32_t)total) padded <<= 1;
  const size_t smem = (size_t)padded * sizeof(uint32_t);
  TE_CUDA_TRY(cudaFuncSetAttribute(select_splitters_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaStream_t st = (cudaStream_t)stream;
  TE_PROF("select_splitters", st);
  select_splitters_kernel<<<1, 1024, smem, st>>>(samples, (uint32_t)total, padded, world, splitters,
                                                 (long long*)splitters_i64);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_combine_auroc_stats(const double* rows, int32_t world, double n_received, double* out,
                                      void* stream) {
  if (!rows || !out || world < 1) return TE_ERR_INVALID_ARGUMENT;
  combine_auroc_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(rows, world, n_received, out);
  TE_LAUNCH_CHECK();
  return TE_OK;
}
