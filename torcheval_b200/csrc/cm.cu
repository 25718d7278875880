// K4 — confusion-matrix counting: cm[target[i], pred[i]] += 1.
//
// Replaces the reference's sparse_coo_tensor(...).to_dense() coalesce
// (torcheval/metrics/functional/classification/confusion_matrix.py:219-234) and the two
// torch.max(...) range checks with their host syncs (:271-282) by one streaming pass:
// 128-bit loads of the two int64 label streams, a fused range check, and either
//   * a shared-memory privatised histogram (small C: bins live in smem, lane-replicated so
//     that a warp never collides with itself), or
//   * fire-and-forget RED.ADD into the L2-resident C*C matrix (large C: 10^6 bins at C=1000).
// Algorithmic traffic: 16 B / sample.
#include "common.cuh"
#include "prof.cuh"

namespace te {
namespace {

constexpr int kCmBlock = 256;

struct RangeReport {
  long long* range;  // {min_pred, max_pred, min_target, max_target} over OUT-OF-RANGE values only
  __device__ __forceinline__ void bad(long long p, long long t, unsigned long long C) const {
    if ((unsigned long long)p >= C) {
      atomicMin(range + 0, p);
      atomicMax(range + 1, p);
    }
    if ((unsigned long long)t >= C) {
      atomicMin(range + 2, t);
      atomicMax(range + 3, t);
    }
  }
};

__global__ void cm_init_range_kernel(long long* range) {
  if (threadIdx.x == 0) {
    range[0] = 0x7fffffffffffffffLL;
    range[1] = (long long)0x8000000000000000ULL;
    range[2] = 0x7fffffffffffffffLL;
    range[3] = (long long)0x8000000000000000ULL;
  }
}

// Sink policies: where one (target, pred) hit goes.
template <typename BinT>
struct GlobalSink {
  BinT* bins;
  __device__ __forceinline__ void add(uint32_t bin) const { atomicAdd(bins + bin, (BinT)1); }
};
template <int REPL>
struct SmemSink {
  uint32_t* sh;
  unsigned lane;
  __device__ __forceinline__ void add(uint32_t bin) const {
    atomicAdd(sh + (size_t)bin * REPL + (REPL > 1 ? lane : 0u), 1u);
  }
};

template <int VEC, int UNROLL, typename Sink>
__device__ __forceinline__ void cm_stream(const long long* __restrict__ pred,
                                          const long long* __restrict__ target, int64_t n,
                                          uint32_t C, const Sink& sink, const RangeReport& rr) {
  const int64_t nvec = n / VEC;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (int64_t base = first; base < nvec; base += stride * UNROLL) {
    long long p[UNROLL][VEC], t[UNROLL][VEC];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t idx = base + (int64_t)u * stride;
      if (idx < nvec) {
        if (VEC == 2) {
          int4 a = ld_stream_v4(pred + 2 * idx);
          int4 b = ld_stream_v4(target + 2 * idx);
          p[u][0] = ((long long)(uint32_t)a.y << 32) | (uint32_t)a.x;
          p[u][VEC - 1] = ((long long)(uint32_t)a.w << 32) | (uint32_t)a.z;
          t[u][0] = ((long long)(uint32_t)b.y << 32) | (uint32_t)b.x;
          t[u][VEC - 1] = ((long long)(uint32_t)b.w << 32) | (uint32_t)b.z;
        } else {
          p[u][0] = __ldg(pred + idx);
          t[u][0] = __ldg(target + idx);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t idx = base + (int64_t)u * stride;
      if (idx < nvec) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
          const unsigned long long pu = (unsigned long long)p[u][v];
          const unsigned long long tu = (unsigned long long)t[u][v];
          if (pu < C && tu < C)
            sink.add((uint32_t)tu * C + (uint32_t)pu);
          else
            rr.bad(p[u][v], t[u][v], C);
        }
      }
    }
  }
  // scalar tail (n % VEC elements)
  if (VEC > 1 && first == 0) {
    for (int64_t i = nvec * VEC; i < n; ++i) {
      const long long pv = pred[i], tv = target[i];
      if ((unsigned long long)pv < C && (unsigned long long)tv < C)
        sink.add((uint32_t)tv * C + (uint32_t)pv);
      else
        rr.bad(pv, tv, C);
    }
  }
}

template <int VEC, typename BinT>
__global__ void __launch_bounds__(kCmBlock)
    cm_global_kernel(const long long* __restrict__ pred, const long long* __restrict__ target,
                     int64_t n, uint32_t C, BinT* bins, long long* range) {
  GlobalSink<BinT> sink{bins};
  RangeReport rr{range};
  cm_stream<VEC, 4>(pred, target, n, C, sink, rr);
}

template <int VEC, int REPL>
__global__ void __launch_bounds__(kCmBlock)
    cm_smem_kernel(const long long* __restrict__ pred, const long long* __restrict__ target,
                   int64_t n, uint32_t C, unsigned long long* cm, long long* range) {
  extern __shared__ uint32_t sh[];
  const uint32_t nbins = C * C;
  for (uint32_t i = threadIdx.x; i < nbins * REPL; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  SmemSink<REPL> sink{sh, lane_id()};
  RangeReport rr{range};
  cm_stream<VEC, 4>(pred, target, n, C, sink, rr);
  __syncthreads();
  for (uint32_t b = threadIdx.x; b < nbins; b += blockDim.x) {
    uint32_t s = 0;
#pragma unroll
    for (int r = 0; r < REPL; ++r) s += sh[(size_t)b * REPL + ((r + threadIdx.x) & (REPL - 1))];
    if (s) atomicAdd(cm + b, (unsigned long long)s);
  }
}

// Large C on B200: the bins do not fit one SM's shared memory, and RED into L2 tops out near
// 190 G updates/s chip-wide (profiles/r01_ubench_b200.txt).  A thread-block cluster of S CTAs owns
// S * kClusterBinsPerCta bins in its *distributed* shared memory: bin b lives in the smem of
// cluster rank b / kClusterBinsPerCta and is updated with red.shared::cluster; only bins beyond
// the cluster's capacity fall back to RED in L2.  Each cluster keeps private counts for the
// samples it streams and folds them into `cm` once at the end.
constexpr uint32_t kClusterBinsPerCta = 57344;  // 224 KB of u32 counters per CTA
constexpr int kClusterBlock = 1024;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

struct ClusterSink {
  uint32_t smem_base;   // shared::cta address of this CTA's bin array
  uint32_t dsm_bins;    // bins held in distributed shared memory (multiple of kClusterBinsPerCta)
  unsigned long long* cm;
  __device__ __forceinline__ void add(uint32_t bin) const {
    if (bin < dsm_bins) {
      const uint32_t owner = bin / kClusterBinsPerCta;
      const uint32_t off = bin - owner * kClusterBinsPerCta;
      uint32_t raddr;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_base + off * 4u), "r"(owner));
      asm volatile("red.relaxed.cluster.shared::cluster.add.u32 [%0], %1;" ::"r"(raddr), "r"(1u) : "memory");
    } else {
      atomicAdd(cm + bin, 1ull);
    }
  }
};

template <int VEC>
__global__ void __launch_bounds__(kClusterBlock, 1)
    cm_cluster_kernel(const long long* __restrict__ pred, const long long* __restrict__ target,
                      int64_t n, uint32_t C, unsigned long long* cm, long long* range) {
  extern __shared__ __align__(16) uint32_t shbins[];
  const uint32_t rank = cluster_ctarank(), csize = cluster_nctarank();
  const uint32_t nbins = C * C;
  for (uint32_t i = threadIdx.x; i < kClusterBinsPerCta; i += blockDim.x) shbins[i] = 0;
  cluster_sync_all();  // every CTA's bins are zero before anybody adds remotely
  uint32_t dsm = csize * kClusterBinsPerCta;
  ClusterSink sink{(uint32_t)__cvta_generic_to_shared(shbins), dsm, cm};
  RangeReport rr{range};
  cm_stream<VEC, 2>(pred, target, n, C, sink, rr);
  cluster_sync_all();  // all remote adds into my bins have landed
  const uint32_t first = rank * kClusterBinsPerCta;
  for (uint32_t i = threadIdx.x; i < kClusterBinsPerCta; i += blockDim.x) {
    const uint32_t v = shbins[i];
    if (v && first + i < nbins) atomicAdd(cm + first + i, (unsigned long long)v);
  }
}

__global__ void cm_fold32_kernel(const uint32_t* __restrict__ bins, unsigned long long* cm,
                                 uint32_t nbins) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nbins; i += gridDim.x * blockDim.x)
    cm[i] += bins[i];
}

// ---- logits (N, C) -> argmax -> count ---------------------------------------------------
// torch.argmax semantics (confusion_matrix.py:225-226): first maximal index, NaN counts as max.
struct Best {
  float v;
  int i;
};
__device__ __forceinline__ bool better(float av, int ai, float bv, int bi) {
  const bool an = av != av, bn = bv != bv;
  if (an || bn) return an && (!bn || ai < bi);
  return av > bv || (av == bv && ai < bi);
}

// One warp per row; lanes stride over the columns (coalesced 128 B per instruction).
template <typename BinT>
__global__ void __launch_bounds__(kCmBlock)
    cm_logits_warp_kernel(const float* __restrict__ logits, int64_t ld,
                          const long long* __restrict__ target, int64_t n, uint32_t C, BinT* bins,
                          long long* range) {
  const unsigned lane = lane_id();
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x / kWarp);
  RangeReport rr{range};
  for (int64_t row = (int64_t)blockIdx.x * (blockDim.x / kWarp) + threadIdx.x / kWarp; row < n;
       row += warps) {
    const float* r = logits + row * ld;
    float bv = 0.f;
    int bi = -1;
    for (uint32_t c = lane; c < C; c += kWarp) {
      const float v = __ldg(r + c);
      if (bi < 0 || better(v, (int)c, bv, bi)) {
        bv = v;
        bi = (int)c;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(kFull, bv, o);
      const int oi = __shfl_xor_sync(kFull, bi, o);
      if (oi >= 0 && (bi < 0 || better(ov, oi, bv, bi))) {
        bv = ov;
        bi = oi;
      }
    }
    if (lane == 0) {
      const long long t = target[row];
      if ((unsigned long long)t < C)
        atomicAdd(bins + (uint32_t)t * C + (uint32_t)bi, (BinT)1);
      else
        rr.bad(0, t, C);
    }
  }
}

// Small C: a CTA stages a tile of whole rows in smem with coalesced loads, then one thread
// scans one row.
constexpr int kLogitsTileFloats = 8192;
template <typename BinT>
__global__ void __launch_bounds__(kCmBlock)
    cm_logits_tile_kernel(const float* __restrict__ logits, int64_t ld,
                          const long long* __restrict__ target, int64_t n, uint32_t C, BinT* bins,
                          long long* range) {
  __shared__ float tile[kLogitsTileFloats];
  const int rows_per_tile = kLogitsTileFloats / (int)C;
  const int64_t ntiles = (n + rows_per_tile - 1) / rows_per_tile;
  RangeReport rr{range};
  for (int64_t tl = blockIdx.x; tl < ntiles; tl += gridDim.x) {
    const int64_t row0 = tl * rows_per_tile;
    const int rows = (int)min((int64_t)rows_per_tile, n - row0);
    if (ld == (int64_t)C) {
      const float* src = logits + row0 * ld;
      for (int i = threadIdx.x; i < rows * (int)C; i += blockDim.x) tile[i] = __ldg(src + i);
    } else {
      for (int i = threadIdx.x; i < rows * (int)C; i += blockDim.x) {
        const int r = i / (int)C, c = i - r * (int)C;
        tile[i] = __ldg(logits + (row0 + r) * ld + c);
      }
    }
    __syncthreads();
    for (int r = threadIdx.x; r < rows; r += blockDim.x) {
      const float* q = tile + r * (int)C;
      float bv = q[0];
      int bi = 0;
      for (int c = 1; c < (int)C; ++c) {
        const float v = q[c];
        if (better(v, c, bv, bi)) {
          bv = v;
          bi = c;
        }
      }
      const long long t = target[row0 + r];
      if ((unsigned long long)t < C)
        atomicAdd(bins + (uint32_t)t * C + (uint32_t)bi, (BinT)1);
      else
        rr.bad(0, t, C);
    }
    __syncthreads();
  }
}

// ---- per-class counts for accuracy / F1 -------------------------------------------------
// num_label[c] = #{target == c}, num_pred[c] = #{pred == c}, num_tp[c] = #{pred == target == c}
// i.e. the row sums, column sums and diagonal of the confusion matrix, which is what the
// reference's scatter_(reduce="add") calls compute (accuracy.py:275-277, f1_score.py:184-192) —
// but 3*C bins fit in shared memory even at C = 1000, so this runs at streaming speed where the
// full C*C matrix is bound by L2 atomics.
struct ClassSink {
  uint32_t* sh;  // [3][C]: tp, label, pred
  uint32_t C;
  __device__ __forceinline__ void add(uint32_t p, uint32_t t) const {
    atomicAdd(sh + C + t, 1u);
    atomicAdd(sh + 2 * C + p, 1u);
    if (p == t) atomicAdd(sh + t, 1u);
  }
};

template <int VEC>
__global__ void __launch_bounds__(kCmBlock)
    class_counts_kernel(const long long* __restrict__ pred, const long long* __restrict__ target,
                        int64_t n, uint32_t C, unsigned long long* out3, long long* range) {
  extern __shared__ uint32_t sh[];
  for (uint32_t i = threadIdx.x; i < 3 * C; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  ClassSink sink{sh, C};
  RangeReport rr{range};
  const int64_t nvec = n / VEC;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  constexpr int UNROLL = 4;
  for (int64_t base = first; base < nvec; base += stride * UNROLL) {
    long long p[UNROLL][VEC], t[UNROLL][VEC];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t idx = base + (int64_t)u * stride;
      if (idx < nvec) {
        if (VEC == 2) {
          const int4 a = ld_stream_v4(pred + 2 * idx);
          const int4 b = ld_stream_v4(target + 2 * idx);
          p[u][0] = ((long long)(uint32_t)a.y << 32) | (uint32_t)a.x;
          p[u][VEC - 1] = ((long long)(uint32_t)a.w << 32) | (uint32_t)a.z;
          t[u][0] = ((long long)(uint32_t)b.y << 32) | (uint32_t)b.x;
          t[u][VEC - 1] = ((long long)(uint32_t)b.w << 32) | (uint32_t)b.z;
        } else {
          p[u][0] = __ldg(pred + idx);
          t[u][0] = __ldg(target + idx);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t idx = base + (int64_t)u * stride;
      if (idx < nvec) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
          if ((unsigned long long)p[u][v] < C && (unsigned long long)t[u][v] < C)
            sink.add((uint32_t)p[u][v], (uint32_t)t[u][v]);
          else
            rr.bad(p[u][v], t[u][v], C);
        }
      }
    }
  }
  if (VEC > 1 && first == 0)
    for (int64_t i = nvec * VEC; i < n; ++i) {
      const long long pv = pred[i], tv = target[i];
      if ((unsigned long long)pv < C && (unsigned long long)tv < C)
        sink.add((uint32_t)pv, (uint32_t)tv);
      else
        rr.bad(pv, tv, C);
    }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < 3 * C; i += blockDim.x) {
    const uint32_t v = sh[i];
    if (v) atomicAdd(out3 + i, (unsigned long long)v);
  }
}

// logits variant: one warp per row, argmax as in cm_logits_warp_kernel
__global__ void __launch_bounds__(kCmBlock)
    class_counts_logits_kernel(const float* __restrict__ logits, int64_t ld,
                               const long long* __restrict__ target, int64_t n, uint32_t C,
                               unsigned long long* out3, long long* range) {
  extern __shared__ uint32_t sh[];
  for (uint32_t i = threadIdx.x; i < 3 * C; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  ClassSink sink{sh, C};
  RangeReport rr{range};
  const unsigned lane = lane_id();
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x / kWarp);
  for (int64_t row = (int64_t)blockIdx.x * (blockDim.x / kWarp) + threadIdx.x / kWarp; row < n;
       row += warps) {
    const float* r = logits + row * ld;
    float bv = 0.f;
    int bi = -1;
    for (uint32_t c = lane; c < C; c += kWarp) {
      const float v = __ldg(r + c);
      if (bi < 0 || better(v, (int)c, bv, bi)) {
        bv = v;
        bi = (int)c;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(kFull, bv, o);
      const int oi = __shfl_xor_sync(kFull, bi, o);
      if (oi >= 0 && (bi < 0 || better(ov, oi, bv, bi))) {
        bv = ov;
        bi = oi;
      }
    }
    if (lane == 0) {
      const long long t = target[row];
      if ((unsigned long long)t < C)
        sink.add((uint32_t)bi, (uint32_t)t);
      else
        rr.bad(0, t, C);
    }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < 3 * C; i += blockDim.x) {
    const uint32_t v = sh[i];
    if (v) atomicAdd(out3 + i, (unsigned long long)v);
  }
}

// #{pred[i] == target[i]} — micro accuracy / micro F1 when the number of classes is not given
// (accuracy.py:269-272, f1_score.py:176-180).
__global__ void __launch_bounds__(kCmBlock)
    match_count_kernel(const long long* __restrict__ pred, const long long* __restrict__ target,
                       int64_t n, bool vec, unsigned long long* out) {
  __shared__ unsigned long long ws[kCmBlock / 32];
  unsigned long long acc = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (vec) {
    const int64_t nvec = n / 2;
    for (int64_t i = first; i < nvec; i += stride) {
      const int4 a = ld_stream_v4(pred + 2 * i), b = ld_stream_v4(target + 2 * i);
      acc += (a.x == b.x && a.y == b.y) + (a.z == b.z && a.w == b.w);
    }
    if (first == 0 && (n & 1)) acc += pred[n - 1] == target[n - 1];
  } else {
    for (int64_t i = first; i < n; i += stride) acc += __ldg(pred + i) == __ldg(target + i);
  }
  unsigned long long tot;
  block_scan_incl<unsigned long long, kCmBlock>(acc, ws, tot);
  if (threadIdx.x == 0 && tot) atomicAdd(out, tot);
}

// ---- TE_CM_BUCKETED: two passes, every increment is a SHARED-memory atomic ------------------
// 10^6 bins (C = 1000) do not fit one SM, and one RED per sample into L2 is capped by the SM<->L2
// request rate at ~146 G samples/s = 36 % of the HBM roofline (profiles/r01_tune_cm_v2.txt).  But a
// SLICE of the matrix — R consecutive target rows, R*C <= 50 000 bins — does fit.  So:
//   pass 1 (route)  streams the int64 pairs once (16 B/sample), turns every sample into a 16-bit
//                   index inside its slice (local_row * C + pred) and writes the indices back
//                   grouped by slice WITHIN each 8192-sample tile (unordered multisplit over <= 31
//                   slices: ballots + one shared atomic per distinct slice and warp round), plus a
//                   33-entry offset table per tile and global per-slice totals;
//   pass 2 (count)  one CTA per SM holds one slice's bins in shared memory (int32), walks the
//                   runs of that slice tile by tile (2 B/sample) with shared atomics, and adds
//                   its bins into the int64 matrix at the end.  CTAs are dealt out over the
//                   slices in proportion to the slice totals, so a skewed class distribution
//                   still keeps every SM busy.
// Traffic: 16 + 2 + 2 = 20 B/sample instead of 16 B + one L2 atomic.
constexpr int kRtBlock = 512;
constexpr int kRtIpt = 16;
constexpr int kRtTile = kRtBlock * kRtIpt;  // 8192 samples
constexpr int kRtMaxSlices = 31;            // slice 31 = "not counted" (out of range / past the end)
constexpr uint32_t kRtMaxSliceBins = 50000; // 200 KB of int32 bins per CTA
constexpr int kCtBlock = 1024;

struct BucketGeom {
  uint32_t rows, slices;  // target rows per slice, number of slices
};
inline bool bucket_geom(int32_t C, BucketGeom* g) {
  if (C < 1 || (uint32_t)C > kRtMaxSliceBins) return false;
  const uint32_t rows = kRtMaxSliceBins / (uint32_t)C;
  const uint32_t slices = ((uint32_t)C + rows - 1) / rows;
  if (rows < 1 || slices > (uint32_t)kRtMaxSlices || rows * (uint32_t)C > 65536u) return false;
  g->rows = rows;
  g->slices = slices;
  return true;
}
struct BucketPlan {
  uint64_t tiles;
  size_t off_keys, off_table, off_totals, total;
};
inline BucketPlan bucket_plan(int64_t n) {
  BucketPlan p{};
  p.tiles = ((uint64_t)n + kRtTile - 1) / kRtTile;
  if (p.tiles == 0) p.tiles = 1;
  size_t o = 0;
  p.off_keys = o;
  o += align_up((size_t)p.tiles * kRtTile * 2, 256);
  p.off_table = o;
  o += align_up((size_t)p.tiles * 32 * 2, 256);  // 32 x uint16 run starts per tile
  p.off_totals = o;
  o += 256;  // 32 x uint64 slice totals (zeroed per call)
  p.total = o;
  return p;
}

constexpr int kRtParts = 4;  // slice counters are split over 4 interleaved thread classes

__global__ void __launch_bounds__(kRtBlock, 2)  // pins 64 registers = two CTAs per SM (one more kernel argument took
                                                // ptxas to 69 registers and one CTA: 0.81 -> 1.30 ms, profiles/r02_cm_batches_overlap.txt)
    cm_route_kernel(const long long* __restrict__ pred, const long long* __restrict__ target, int64_t n,
                    uint32_t C, uint32_t rows, uint32_t inv_rows, bool vec_ok,
                    uint16_t* __restrict__ keys, uint16_t* __restrict__ table,
                    unsigned long long* __restrict__ totals, long long* range) {
  __shared__ __align__(16) uint16_t s_keys[kRtTile];
  // position counters: [slice][part]; a thread uses part = tid % 4, so the lanes of a warp that
  // hit the same slice are spread over four addresses (the ranking is unordered: ANY position
  // inside the slice's run will do, so one shared atomic with return per sample is all it takes)
  __shared__ uint32_t s_cnt[32 * kRtParts];
  const unsigned tid = threadIdx.x, lane = tid & 31;
  const uint64_t tile = blockIdx.x;
  const int64_t base = (int64_t)(tile * kRtTile);
  if (tid < 32 * kRtParts) s_cnt[tid] = 0;
  __syncthreads();
  RangeReport rr{range};
  const uint32_t part = tid & (kRtParts - 1);
  uint32_t rec[kRtIpt];      // (slice * 4 + part) << 16 | index inside the slice
  uint32_t rk[kRtIpt / 2];   // arrival number inside (slice, part), two per register
  const bool full = vec_ok && base + kRtTile <= n;
#pragma unroll
  for (int h = 0; h < kRtIpt / 4; ++h) {
    // 4 samples: two 128-bit loads per stream; sample 2*(tid + k*512) + v of the tile
    int4 a[2], b[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int64_t i = base + 2 * ((int64_t)tid + (int64_t)(2 * h + k) * kRtBlock);
      if (full) {
        a[k] = ld_stream_v4(pred + i);
        b[k] = ld_stream_v4(target + i);
      } else {
        long long pv[2], tv[2];
#pragma unroll
        for (int v = 0; v < 2; ++v) {
          pv[v] = i + v < n ? pred[i + v] : -1;  // -1: past the end, never reported (see below)
          tv[v] = i + v < n ? target[i + v] : -1;
        }
        a[k] = make_int4((int)(uint32_t)pv[0], (int)(pv[0] >> 32), (int)(uint32_t)pv[1], (int)(pv[1] >> 32));
        b[k] = make_int4((int)(uint32_t)tv[0], (int)(tv[0] >> 32), (int)(uint32_t)tv[1], (int)(tv[1] >> 32));
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int s = 4 * h + j;
      const int4 pa = a[j >> 1], ta = b[j >> 1];
      const uint32_t plo = (j & 1) ? (uint32_t)pa.z : (uint32_t)pa.x, phi = (j & 1) ? (uint32_t)pa.w : (uint32_t)pa.y;
      const uint32_t tlo = (j & 1) ? (uint32_t)ta.z : (uint32_t)ta.x, thi = (j & 1) ? (uint32_t)ta.w : (uint32_t)ta.y;
      uint32_t sp = 31u * kRtParts, idx = 0u, r = 0u;
      if ((phi | thi) == 0u && plo < C && tlo < C) {
        const uint32_t slice = __umulhi(tlo, inv_rows);
        idx = (tlo - slice * rows) * C + plo;
        sp = slice * kRtParts + part;
        r = atomicAdd(&s_cnt[sp], 1u);
      } else if (full || base + 2 * ((int64_t)tid + (int64_t)(2 * h + (j >> 1)) * kRtBlock) + (j & 1) < n) {
        rr.bad((long long)(((unsigned long long)phi << 32) | plo), (long long)(((unsigned long long)thi << 32) | tlo), C);
      }
      rec[s] = (sp << 16) | idx;
      if ((s & 1) == 0)
        rk[s >> 1] = r;
      else
        rk[s >> 1] |= r << 16;
    }
  }
  __syncthreads();
  // exclusive scan over the 31 * 4 (slice, part) counters in slice-major order -> run starts
  if (tid < 32 * kRtParts) {
    const uint32_t c = tid < 31 * kRtParts ? s_cnt[tid] : 0u;
    uint32_t incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t u = __shfl_up_sync(kFull, incl, o);
      if (lane >= (unsigned)o) incl += u;
    }
    __shared__ uint32_t s_w[kRtParts];
    if (lane == 31) s_w[tid >> 5] = incl;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    uint32_t wb = 0;
    for (unsigned w = 0; w < (tid >> 5); ++w) wb += s_w[w];
    const uint32_t excl = wb + incl - c;
    asm volatile("bar.sync 1, 128;" ::: "memory");
    s_cnt[tid] = excl;  // now the start of (slice, part) inside the tile
    if ((tid & (kRtParts - 1)) == 0) table[tile * 32 + tid / kRtParts] = (uint16_t)excl;  // entry 31 = counted total
  }
  __syncthreads();
  if (tid < 31) {  // slice totals for the count pass's work split
    const uint32_t c = s_cnt[(tid + 1) * kRtParts] - s_cnt[tid * kRtParts];
    if (c) atomicAdd(totals + tid, (unsigned long long)c);
  }
#pragma unroll
  for (int s = 0; s < kRtIpt; ++s) {
    const uint32_t sp = rec[s] >> 16;
    const uint32_t r = (s & 1) ? (rk[s >> 1] >> 16) : (rk[s >> 1] & 0xffffu);
    if (sp < 31u * kRtParts) s_keys[s_cnt[sp] + r] = (uint16_t)(rec[s] & 0xffffu);
  }
  __syncthreads();
  // the tile's indices, grouped by slice, as 16-byte stores (positions past the counted total
  // are never read)
  int4* dst = reinterpret_cast<int4*>(keys + tile * kRtTile);
  const int4* src = reinterpret_cast<const int4*>(s_keys);
#pragma unroll
  for (int k = 0; k < kRtTile * 2 / 16 / kRtBlock; ++k) dst[tid + k * kRtBlock] = src[tid + k * kRtBlock];
}

__global__ void __launch_bounds__(kCtBlock, 1)
    cm_count_kernel(const uint16_t* __restrict__ keys, const uint16_t* __restrict__ table,
                    const unsigned long long* __restrict__ totals, uint64_t tiles, uint32_t C,
                    uint32_t rows, uint32_t slices, unsigned long long* cm) {
  extern __shared__ uint32_t sh_bins[];
  __shared__ unsigned long long s_cum[33];
  const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    unsigned long long c = 0;
    for (uint32_t b = 0; b < 32; ++b) {
      s_cum[b] = c;
      c += b < slices ? totals[b] : 0ull;
    }
    s_cum[32] = c;
  }
  __syncthreads();
  const unsigned long long total = s_cum[32];
  if (total == 0) return;
  // this CTA's share of the counted samples, in slice-major order
  const unsigned long long lo = total * blockIdx.x / gridDim.x, hi = total * (blockIdx.x + 1) / gridDim.x;
  for (uint32_t b = 0; b < slices; ++b) {
    const unsigned long long c0 = s_cum[b], c1 = s_cum[b + 1];
    if (c1 <= lo || c0 >= hi || c1 == c0) continue;
    // the part of slice b that is mine, mapped onto a range of tiles (consecutive CTAs share the
    // boundary value, so the tiles of a slice are partitioned exactly)
    const unsigned long long a0 = (lo > c0 ? lo - c0 : 0ull), a1 = (hi < c1 ? hi : c1) - c0, cnt = c1 - c0;
    const uint64_t t0 = (uint64_t)((__uint128_t)a0 * tiles / cnt);
    const uint64_t t1 = a1 == cnt ? tiles : (uint64_t)((__uint128_t)a1 * tiles / cnt);
    if (t1 <= t0) continue;
    const uint32_t row0 = b * rows;
    const uint32_t nbins = (min(row0 + rows, C) - row0) * C;
    for (uint32_t i = tid; i < nbins; i += kCtBlock) sh_bins[i] = 0u;
    __syncthreads();
    // A warp takes the runs of slice b in two tiles at a time.  A run is ~400 indices (8192 / slices):
    // read with 2-byte loads it was 13 dependent-looking rounds behind a table lookup, and the kernel
    // sat at 26 % issue / 28 % DRAM waiting for them.  Now: both tiles' table entries, then every
    // 16-byte piece of both runs (aligned body; <= 7 indices of head and of tail as scalars) are in
    // flight before the first shared atomic.
    constexpr int kU = 2;
    constexpr uint32_t kWarps = kCtBlock / kWarp;
    auto count8 = [&](const int4& v) {
      const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        atomicAdd(&sh_bins[w[j] & 0xffffu], 1u);
        atomicAdd(&sh_bins[w[j] >> 16], 1u);
      }
    };
    for (uint64_t t = t0 + warp; t < t1; t += kU * kWarps) {
      uint32_t rs[kU], re[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const uint64_t tt = t + (uint64_t)u * kWarps;
        rs[u] = tt < t1 ? table[tt * 32 + b] : 0u;
        re[u] = tt < t1 ? table[tt * 32 + b + 1] : 0u;
      }
      int4 v[kU][2];
      uint32_t hk[kU], tk[kU], a8[kU], nv[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const uint16_t* run = keys + (t + (uint64_t)u * kWarps) * kRtTile;  // 16 KB aligned
        const uint32_t a = min((rs[u] + 7u) & ~7u, re[u]);                   // end of the scalar head
        const uint32_t z = max(re[u] & ~7u, a);                              // start of the scalar tail
        a8[u] = a >> 3, nv[u] = (z - a) >> 3;
        hk[u] = rs[u] + lane < a ? run[rs[u] + lane] : 0xffffffffu;
        tk[u] = z + lane < re[u] ? run[z + lane] : 0xffffffffu;
        const int4* body = reinterpret_cast<const int4*>(run) + a8[u];
#pragma unroll
        for (int j = 0; j < 2; ++j)
          if (lane + 32u * j < nv[u]) v[u][j] = ld_stream_v4(body + lane + 32u * j);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (hk[u] != 0xffffffffu) atomicAdd(&sh_bins[hk[u]], 1u);
        if (tk[u] != 0xffffffffu) atomicAdd(&sh_bins[tk[u]], 1u);
#pragma unroll
        for (int j = 0; j < 2; ++j)
          if (lane + 32u * j < nv[u]) count8(v[u][j]);
        if (nv[u] > 64u) {  // a run longer than 512 indices (skewed targets): the rest, one piece at a time
          const int4* body = reinterpret_cast<const int4*>(keys + (t + (uint64_t)u * kWarps) * kRtTile) + a8[u];
          for (uint32_t q = 64u + lane; q < nv[u]; q += kWarp) count8(ld_stream_v4(body + q));
        }
      }
    }
    __syncthreads();
    unsigned long long* out = cm + (size_t)row0 * C;
    for (uint32_t i = tid; i < nbins; i += kCtBlock) {
      const uint32_t v = sh_bins[i];
      if (v) atomicAdd(out + i, (unsigned long long)v);
    }
    __syncthreads();
  }
}

int cm_bucketed(const long long* pred, const long long* target, int64_t n, int32_t C,
                unsigned long long* cm, long long* range, void* scratch, int sms, cudaStream_t st) {
  BucketGeom g;
  if (!bucket_geom(C, &g)) return TE_ERR_UNSUPPORTED;
  const BucketPlan p = bucket_plan(n);
  char* base = static_cast<char*>(scratch);
  uint16_t* keys = reinterpret_cast<uint16_t*>(base + p.off_keys);
  uint16_t* table = reinterpret_cast<uint16_t*>(base + p.off_table);
  unsigned long long* totals = reinterpret_cast<unsigned long long*>(base + p.off_totals);
  TE_CUDA_TRY(cudaMemsetAsync(totals, 0, 256, st));
  const uint32_t inv_rows = (uint32_t)((((uint64_t)1 << 32) + g.rows - 1) / g.rows);
  const bool vec_ok = (((uintptr_t)pred | (uintptr_t)target) & 15) == 0;
  {
    TE_PROF("cm_route", st);
    cm_route_kernel<<<(unsigned)p.tiles, kRtBlock, 0, st>>>(pred, target, n, (uint32_t)C, g.rows, inv_rows, vec_ok,
                                                            keys, table, totals, range);
  }
  TE_LAUNCH_CHECK();
  const size_t sh = (size_t)g.rows * C * 4;
  TE_CUDA_TRY(cudaFuncSetAttribute(cm_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
  TE_PROF("cm_count", st);
  cm_count_kernel<<<sms, kCtBlock, sh, st>>>(keys, table, totals, p.tiles, (uint32_t)C, g.rows, g.slices, cm);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

constexpr int64_t kBucketedMinN = (int64_t)1 << 22;  // below this the extra pass does not pay

int pick_variant(int32_t C, int32_t variant) {
  if (variant != TE_CM_AUTO) return variant;
  const int64_t nbins = (int64_t)C * C;
  if (nbins * 32 * 4 <= 32 * 1024) return TE_CM_SMEM_REPL;   // C <= 16
  if (nbins * 4 <= 48 * 1024) return TE_CM_SMEM;             // C <= 110
  return TE_CM_GLOBAL64;
}

}  // namespace
}  // namespace te

using namespace te;

extern "C" size_t te_cm_scratch_bytes(int32_t num_classes, int32_t variant) {
  if (num_classes < 1) return 0;
  const int v = pick_variant(num_classes, variant);
  if (v == TE_CM_GLOBAL32) return align_up((size_t)num_classes * num_classes * 4, 256);
  return 0;
}

extern "C" size_t te_cm_scratch_bytes_n(int32_t num_classes, int32_t variant, int64_t n) {
  if (num_classes < 1 || n < 0) return 0;
  BucketGeom g;
  const bool can = bucket_geom(num_classes, &g) && n < ((int64_t)1 << 32);
  if (variant == TE_CM_BUCKETED) return can ? bucket_plan(n).total : 0;
  if (variant == TE_CM_AUTO && can && pick_variant(num_classes, variant) == TE_CM_GLOBAL64 && n >= kBucketedMinN)
    return bucket_plan(n).total;
  return te_cm_scratch_bytes(num_classes, variant);
}

template <int VEC>
static int cm_launch(const long long* pred, const long long* target, int64_t n, int32_t C,
                     unsigned long long* cm, long long* range, void* scratch, int v, int sms,
                     cudaStream_t st) {
  const uint32_t nbins = (uint32_t)C * (uint32_t)C;
  const int64_t nvec = (n + VEC - 1) / VEC;
  int64_t want = (nvec + kCmBlock * 4 - 1) / (kCmBlock * 4);
  if (want < 1) want = 1;
  const int grid = (int)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
  switch (v) {
    case TE_CM_GLOBAL64: {
      TE_PROF("cm_global64", st);
      cm_global_kernel<VEC, unsigned long long>
          <<<grid, kCmBlock, 0, st>>>(pred, target, n, (uint32_t)C, cm, range);
      break;
    }
    case TE_CM_GLOBAL32: {
      // int32 bins hold at most 2^31 hits: feed the stream in slices.
      const int64_t slice = (int64_t)1 << 30;
      for (int64_t off = 0; off < n; off += slice) {
        const int64_t len = (n - off) < slice ? (n - off) : slice;
        TE_CUDA_TRY(cudaMemsetAsync(scratch, 0, (size_t)nbins * 4, st));
        {
          TE_PROF("cm_global32", st);
          cm_global_kernel<VEC, uint32_t>
              <<<grid, kCmBlock, 0, st>>>(pred + off, target + off, len, (uint32_t)C,
                                          (uint32_t*)scratch, range);
        }
        TE_PROF("cm_fold32", st);
        cm_fold32_kernel<<<(nbins + 255) / 256 < 1184 ? (nbins + 255) / 256 : 1184, 256, 0, st>>>(
            (const uint32_t*)scratch, cm, nbins);
      }
      break;
    }
    case TE_CM_SMEM_REPL: {
      const size_t sh = (size_t)nbins * 32 * 4;
      if (sh > 48 * 1024)
        TE_CUDA_TRY(cudaFuncSetAttribute(cm_smem_kernel<VEC, 32>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
      const int g = grid < sms * 4 ? grid : sms * 4;
      TE_PROF("cm_smem_repl", st);
      cm_smem_kernel<VEC, 32><<<g, kCmBlock, sh, st>>>(pred, target, n, (uint32_t)C, cm, range);
      break;
    }
    case TE_CM_SMEM: {
      const size_t sh = (size_t)nbins * 4;
      if (sh > 48 * 1024)
        TE_CUDA_TRY(cudaFuncSetAttribute(cm_smem_kernel<VEC, 1>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
      const int g = grid < sms * 4 ? grid : sms * 4;
      TE_PROF("cm_smem", st);
      cm_smem_kernel<VEC, 1><<<g, kCmBlock, sh, st>>>(pred, target, n, (uint32_t)C, cm, range);
      break;
    }
    case TE_CM_CLUSTER: {
      const size_t sh = (size_t)kClusterBinsPerCta * 4;
      auto kern = cm_cluster_kernel<VEC>;
      TE_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
      TE_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      cudaLaunchConfig_t cfg{};
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      cfg.blockDim = dim3(kClusterBlock);
      cfg.dynamicSmemBytes = sh;
      cfg.stream = st;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      // largest cluster the device can co-schedule: 16, else 8
      int csize = 16, nclusters = 0;
      for (; csize >= 8; csize /= 2) {
        attr[0].val.clusterDim.x = csize;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3(csize);
        if (cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg) == cudaSuccess && nclusters > 0) break;
        (void)cudaGetLastError();
        nclusters = 0;
      }
      if (nclusters <= 0) return TE_ERR_UNSUPPORTED;
      cfg.gridDim = dim3(nclusters * csize);
      const long long* a0 = pred;
      const long long* a1 = target;
      int64_t a2 = n;
      uint32_t a3 = (uint32_t)C;
      unsigned long long* a4 = cm;
      long long* a5 = range;
      TE_PROF(csize == 16 ? "cm_cluster16" : "cm_cluster8", st);
      TE_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, a0, a1, a2, a3, a4, a5));
      break;
    }
    default:
      return TE_ERR_INVALID_ARGUMENT;
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_cm_update(const int64_t* pred, const int64_t* target, int64_t n,
                            int32_t num_classes, int64_t* cm, int64_t* range, void* scratch,
                            size_t scratch_bytes, int32_t variant, void* stream) {
  if (n < 0 || num_classes < 1 || num_classes > 46340 || !cm || !range) return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!pred || !target)) return TE_ERR_INVALID_ARGUMENT;
  int v = pick_variant(num_classes, variant);
  {
    // large C: the bucketed two-pass variant when the caller's scratch has room for it
    // (te_cm_scratch_bytes_n), else RED into L2
    BucketGeom g;
    const bool can = bucket_geom(num_classes, &g) && n < ((int64_t)1 << 32);
    if (variant == TE_CM_BUCKETED && !can) return TE_ERR_UNSUPPORTED;
    if (variant == TE_CM_AUTO && v == TE_CM_GLOBAL64 && can && n >= kBucketedMinN && scratch &&
        scratch_bytes >= bucket_plan(n).total)
      v = TE_CM_BUCKETED;
    if (v == TE_CM_BUCKETED && (!scratch || scratch_bytes < bucket_plan(n).total)) return TE_ERR_SCRATCH_TOO_SMALL;
  }
  if ((v == TE_CM_SMEM_REPL && (size_t)num_classes * num_classes * 128 > 200 * 1024) ||
      (v == TE_CM_SMEM && (size_t)num_classes * num_classes * 4 > 200 * 1024))
    return TE_ERR_INVALID_ARGUMENT;
  if (v != TE_CM_BUCKETED && (scratch_bytes < te_cm_scratch_bytes(num_classes, v) || (v == TE_CM_GLOBAL32 && !scratch)))
    return TE_ERR_SCRATCH_TOO_SMALL;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  cm_init_range_kernel<<<1, 32, 0, st>>>((long long*)range);
  TE_LAUNCH_CHECK();
  if (n == 0) return TE_OK;
  if (v == TE_CM_BUCKETED)
    return cm_bucketed((const long long*)pred, (const long long*)target, n, num_classes,
                       (unsigned long long*)cm, (long long*)range, scratch, sms, st);
  const bool aligned = (((uintptr_t)pred | (uintptr_t)target) & 15) == 0;
  if (aligned)
    return cm_launch<2>((const long long*)pred, (const long long*)target, n, num_classes,
                        (unsigned long long*)cm, (long long*)range, scratch, v, sms, st);
  return cm_launch<1>((const long long*)pred, (const long long*)target, n, num_classes,
                      (unsigned long long*)cm, (long long*)range, scratch, v, sms, st);
}

extern "C" int te_cm_update_logits(const float* logits, int64_t ld, const int64_t* target,
                                   int64_t n, int32_t num_classes, int64_t* cm, int64_t* range,
                                   void* stream) {
  if (n < 0 || num_classes < 1 || num_classes > 46340 || !cm || !range || ld < num_classes)
    return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!logits || !target)) return TE_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  cm_init_range_kernel<<<1, 32, 0, st>>>((long long*)range);
  TE_LAUNCH_CHECK();
  if (n == 0) return TE_OK;
  const uint32_t C = (uint32_t)num_classes;
  if (C <= 64) {
    const int rows_per_tile = kLogitsTileFloats / (int)C;
    const int64_t ntiles = (n + rows_per_tile - 1) / rows_per_tile;
    const int grid = (int)(ntiles < (int64_t)sms * 8 ? ntiles : (int64_t)sms * 8);
    TE_PROF("cm_logits_tile", st);
    cm_logits_tile_kernel<unsigned long long><<<grid, kCmBlock, 0, st>>>(
        logits, ld, (const long long*)target, n, C, (unsigned long long*)cm, (long long*)range);
  } else {
    const int64_t want = (n + (kCmBlock / kWarp) - 1) / (kCmBlock / kWarp);
    const int grid = (int)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
    TE_PROF("cm_logits_warp", st);
    cm_logits_warp_kernel<unsigned long long><<<grid, kCmBlock, 0, st>>>(
        logits, ld, (const long long*)target, n, C, (unsigned long long*)cm, (long long*)range);
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_class_counts(const int64_t* pred, const float* logits, int64_t ld,
                               const int64_t* target, int64_t n, int32_t num_classes,
                               int64_t* counts3, int64_t* range, void* stream) {
  if (n < 0 || num_classes < 1 || num_classes > 18000 || !counts3 || !range)
    return TE_ERR_INVALID_ARGUMENT;
  if ((pred == nullptr) == (logits == nullptr) && n > 0) return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && !target) return TE_ERR_INVALID_ARGUMENT;
  if (logits && ld < num_classes) return TE_ERR_INVALID_ARGUMENT;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  cm_init_range_kernel<<<1, 32, 0, st>>>((long long*)range);
  TE_LAUNCH_CHECK();
  if (n == 0) return TE_OK;
  const uint32_t C = (uint32_t)num_classes;
  const size_t sh = (size_t)3 * C * 4;
  unsigned long long* out3 = (unsigned long long*)counts3;
  if (pred) {
    const bool aligned = (((uintptr_t)pred | (uintptr_t)target) & 15) == 0;
    const int64_t nvec = aligned ? (n + 1) / 2 : n;
    int64_t want = (nvec + kCmBlock * 4 - 1) / (kCmBlock * 4);
    const int occ = sh > 100 * 1024 ? 1 : (sh > 48 * 1024 ? 2 : 4);
    const int grid = (int)(want < (int64_t)sms * occ ? (want < 1 ? 1 : want) : (int64_t)sms * occ);
    TE_PROF("class_counts", st);
    if (aligned) {
      if (sh > 48 * 1024)
        TE_CUDA_TRY(cudaFuncSetAttribute(class_counts_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
      class_counts_kernel<2><<<grid, kCmBlock, sh, st>>>((const long long*)pred, (const long long*)target, n, C, out3, (long long*)range);
    } else {
      if (sh > 48 * 1024)
        TE_CUDA_TRY(cudaFuncSetAttribute(class_counts_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
      class_counts_kernel<1><<<grid, kCmBlock, sh, st>>>((const long long*)pred, (const long long*)target, n, C, out3, (long long*)range);
    }
  } else {
    const int64_t want = (n + (kCmBlock / kWarp) - 1) / (kCmBlock / kWarp);
    const int occ = sh > 100 * 1024 ? 1 : (sh > 48 * 1024 ? 2 : 8);
    const int grid = (int)(want < (int64_t)sms * occ ? want : (int64_t)sms * occ);
    if (sh > 48 * 1024)
      TE_CUDA_TRY(cudaFuncSetAttribute(class_counts_logits_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
    TE_PROF("class_counts_logits", st);
    class_counts_logits_kernel<<<grid, kCmBlock, sh, st>>>(logits, ld, (const long long*)target, n, C, out3, (long long*)range);
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_match_count(const int64_t* pred, const int64_t* target, int64_t n, int64_t* out,
                              void* stream) {
  if (n < 0 || !out || (n > 0 && (!pred || !target))) return TE_ERR_INVALID_ARGUMENT;
  if (n == 0) return TE_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  const bool vec = (((uintptr_t)pred | (uintptr_t)target) & 15) == 0;
  int64_t want = (n / 2 + kCmBlock * 4 - 1) / (kCmBlock * 4);
  const int grid = (int)(want < 1 ? 1 : (want < (int64_t)sms * 8 ? want : (int64_t)sms * 8));
  TE_PROF("match_count", st);
  match_count_kernel<<<grid, kCmBlock, 0, st>>>((const long long*)pred, (const long long*)target, n,
                                                vec, (unsigned long long*)out);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

// ------------------------------------------------------------------------------------------
// te_cm_update_noalloc — the TE_CM_GLOBAL32 variant with its int32 bins in a static device array
// ------------------------------------------------------------------------------------------
// Who needs this: hosts that hand this library raw device pointers but cannot give it a scratch
// buffer because they are not allowed to allocate at the point of the call — an inference server
// whose whole HBM budget was carved up at start-up and whose request path is forbidden from
// touching the allocator (a cudaMalloc there can stall on the device or simply fail), or a plugin
// called back from someone else's runtime that owns all device memory.  They still pass the
// inputs and the output matrix; the only thing this entry point supplies itself is the 4-byte
// bin array that te_cm_update(TE_CM_GLOBAL32) takes as `scratch`.
//
// The bins live in a file-scope __device__ array of TE_CM_NOALLOC_MAX_CLASSES^2 counters
// (4 MiB), present on every GPU that loads the library for as long as the process lives.
//   * Nothing is allocated or freed here: no cudaMalloc/cudaFree, no allocating library call.
//   * The array is shared by every call on the device and keeps whatever the last call left in
//     it, so each call first clears the C*C counters it is going to use (cudaMemsetAsync on the
//     caller's stream).  Two calls whose GPU work overlaps (different streams, or different
//     threads) would count into the same bins and both results would be wrong: serialise them.
//   * `g_cm_static_bins` names a device symbol; in host code that name is not a device address.
//     The address is fetched once per device with cudaGetSymbolAddress, with that device
//     current, and cached.  The first call on a device does this set-up — make that first call
//     before any stream capture is started on the device.
static __device__ __align__(256) uint32_t g_cm_static_bins[TE_CM_NOALLOC_MAX_CLASSES * TE_CM_NOALLOC_MAX_CLASSES];

namespace {
constexpr int kMaxDevices = 64;
// Device address of g_cm_static_bins per device, filled on first use.  Written without a lock:
// two host threads making the first call on one device at the same time both look the symbol up
// and store the SAME address, so the race is benign (an aligned pointer store is not torn).
uint32_t* g_static_bins_ptr[kMaxDevices] = {};
}  // namespace

extern "C" int te_cm_update_noalloc(const int64_t* pred, const int64_t* target, int64_t n,
                                    int32_t num_classes, int64_t* cm, int64_t* range, void* stream) {
  if (n < 0 || num_classes < 1 || !cm || !range) return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!pred || !target)) return TE_ERR_INVALID_ARGUMENT;
  if (num_classes > TE_CM_NOALLOC_MAX_CLASSES) return TE_ERR_STATIC_SCRATCH_TOO_SMALL;
  int dev = -1;
  TE_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return TE_ERR_UNSUPPORTED;
  uint32_t* bins = g_static_bins_ptr[dev];
  if (!bins) {
    // once per device, with that device current: turn the symbol into a device address
    void* p = nullptr;
    TE_CUDA_TRY(cudaGetSymbolAddress(&p, g_cm_static_bins));
    if (!p) return TE_ERR_UNSUPPORTED;
    bins = static_cast<uint32_t*>(p);
    g_static_bins_ptr[dev] = bins;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int sms = 0;
  TE_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  cm_init_range_kernel<<<1, 32, 0, st>>>((long long*)range);
  TE_LAUNCH_CHECK();
  if (n == 0) return TE_OK;
  const uint32_t C = (uint32_t)num_classes, nbins = C * C;
  const bool aligned = (((uintptr_t)pred | (uintptr_t)target) & 15) == 0;
  const int64_t slice = (int64_t)1 << 30;  // int32 bins hold at most 2^31 hits per fold
  for (int64_t off = 0; off < n; off += slice) {
    const int64_t len = (n - off) < slice ? (n - off) : slice;
    const int64_t nvec = aligned ? (len + 1) / 2 : len;
    int64_t want = (nvec + kCmBlock * 4 - 1) / (kCmBlock * 4);
    const int grid = (int)(want < 1 ? 1 : (want < (int64_t)sms * 8 ? want : (int64_t)sms * 8));
    // the array still holds the previous call's counts: clear what this call will use
    TE_CUDA_TRY(cudaMemsetAsync(bins, 0, (size_t)nbins * sizeof(uint32_t), st));
    if (aligned)
      cm_global_kernel<2, uint32_t><<<grid, kCmBlock, 0, st>>>(
          (const long long*)pred + off, (const long long*)target + off, len, C, bins, (long long*)range);
    else
      cm_global_kernel<1, uint32_t><<<grid, kCmBlock, 0, st>>>(
          (const long long*)pred + off, (const long long*)target + off, len, C, bins, (long long*)range);
    TE_LAUNCH_CHECK();
    const uint32_t fgrid = (nbins + 255) / 256 < 1184 ? (nbins + 255) / 256 : 1184;
    cm_fold32_kernel<<<fgrid, 256, 0, st>>>(bins, (unsigned long long*)cm, nbins);
    TE_LAUNCH_CHECK();
  }
  return TE_OK;
}
