// K3 — binned (thresholded) counting for BinaryBinnedAUROC / AUPRC / precision-recall curve.
//
// The reference computes, per threshold t_j, TP_j = #{target & x >= t_j} and FP_j either with
// searchsorted + float64 histc + flip-cumsum-flip
// (torcheval/metrics/functional/classification/binned_precision_recall_curve.py:84-110) or with
// a (T, tasks, N) boolean tensor (functional/classification/binned_auroc.py:117-125).  Both are
// count-identical to: bucket(x) = #{j : t_j <= x}, a (bucket, label) histogram, and suffix sums.
// Here the histogram is privatised per CTA in shared memory (2*(T+1) u32 counters, the threshold
// table and a cell->first-candidate lookup table all live in smem), the score/label streams are
// read once with 128-bit loads (12 B / score) and the per-CTA counters are merged
// into the caller's int64 histogram with one RED per non-empty bin.
#include "common.cuh"
#include "prof.cuh"

namespace te {
namespace {

constexpr int kBinBlock = 1024;
constexpr size_t kBinSmemBudget = 220 * 1024;

struct BinGeom {
  int cells;        // G, power of two
  size_t smem;      // bytes for the smem-resident variant (0: does not fit)
};

inline BinGeom binned_geom(int32_t T) {
  BinGeom g;
  int G = 1024;
  while (G < 2 * T && G < 32768) G <<= 1;
  auto need = [&](int cells) {
    return (size_t)8 * (T + 1) + (size_t)4 * T + (size_t)2 * (cells + 1) + 16;
  };
  while (need(G) > kBinSmemBudget && G > 1024) G >>= 1;
  g.cells = G;
  g.smem = (T < 65536 && need(G) <= kBinSmemBudget) ? need(G) : 0;
  return g;
}

// Same expression in the table builder and in the sample loop: it only has to be monotone.
__device__ __forceinline__ int cell_of(float v, float lo, float scale, int G) {
  const float f = __fmul_rn(__fsub_rn(v, lo), scale);
  int c = __float2int_rz(f);
  c = c < 0 ? 0 : c;
  return c > G - 1 ? G - 1 : c;
}
__device__ __forceinline__ float cell_scale(float lo, float hi, int G) {
  return hi > lo ? __fdiv_rn((float)G, __fsub_rn(hi, lo)) : 0.f;
}

// lut[c] = #{j : cell_of(thr[j]) < c}, c in [0, G].
__global__ void binned_lut_kernel(const float* __restrict__ thr, int T, uint32_t* lut, int G) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c > G) return;
  const float lo = thr[0], hi = thr[T - 1];
  const float scale = cell_scale(lo, hi, G);
  int b = 0, e = T;
  while (b < e) {
    const int m = (b + e) >> 1;
    if (cell_of(thr[m], lo, scale, G) < c)
      b = m + 1;
    else
      e = m;
  }
  lut[c] = (uint32_t)b;
}

template <typename LutT>
__device__ __forceinline__ int bucket_of(float x, float lo, float hi, float scale, int G, int T,
                                         const float* __restrict__ TH,
                                         const LutT* __restrict__ L) {
  if (!(x >= lo)) return 0;  // below the first threshold (or NaN): counted nowhere
  if (x >= hi) return T;
  const int c = cell_of(x, lo, scale, G);
  int b = (int)L[c], e = (int)L[c + 1];
  while (e - b > 4) {
    const int m = (b + e) >> 1;
    if (TH[m] <= x)
      b = m + 1;
    else
      e = m;
  }
  while (b < e && TH[b] <= x) ++b;
  return b;
}

template <int VEC, bool SMEM>
__global__ void __launch_bounds__(kBinBlock, 1)
    binned_hist_kernel(const float* __restrict__ x, int64_t x_stride,
                       const long long* __restrict__ target, int64_t t_stride, int64_t n,
                       const float* __restrict__ thr, int T, const uint32_t* __restrict__ lut,
                       int G, unsigned long long* hist) {
  extern __shared__ __align__(16) unsigned char smraw[];
  const int task = blockIdx.y;
  x += (int64_t)task * x_stride;
  target += (int64_t)task * t_stride;
  hist += (size_t)task * 2 * (T + 1);

  uint32_t* sh = reinterpret_cast<uint32_t*>(smraw);
  float* sthr = reinterpret_cast<float*>(sh + 2 * (T + 1));
  uint16_t* slut = reinterpret_cast<uint16_t*>(sthr + T);
  if (SMEM) {
    for (int i = threadIdx.x; i < 2 * (T + 1); i += blockDim.x) sh[i] = 0;
    for (int i = threadIdx.x; i < T; i += blockDim.x) sthr[i] = thr[i];
    for (int i = threadIdx.x; i <= G; i += blockDim.x) slut[i] = (uint16_t)lut[i];
    __syncthreads();
  }
  const float lo = thr[0], hi = thr[T - 1];
  const float scale = cell_scale(lo, hi, G);

  auto hit = [&](float xv, long long tv) {
    int b;
    if (SMEM)
      b = bucket_of<uint16_t>(xv, lo, hi, scale, G, T, sthr, slut);
    else
      b = bucket_of<uint32_t>(xv, lo, hi, scale, G, T, thr, lut);
    const int bin = 2 * b + (tv != 0 ? 1 : 0);
    if (SMEM)
      atomicAdd(sh + bin, 1u);
    else
      atomicAdd(hist + bin, 1ull);
  };

  const int64_t nvec = n / VEC;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t first = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  constexpr int UNROLL = VEC == 4 ? 3 : 2;
  for (int64_t base = first; base < nvec; base += stride * UNROLL) {
    float xv[UNROLL][VEC];
    long long tv[UNROLL][VEC];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t idx = base + (int64_t)u * stride;
      if (idx < nvec) {
        if (VEC == 4) {
          const int4 a = ld_stream_v4(x + 4 * idx);
          const int4 t0 = ld_stream_v4(target + 4 * idx);
          const int4 t1 = ld_stream_v4(target + 4 * idx + 2);
          xv[u][0] = __int_as_float(a.x);
          xv[u][1 % VEC] = __int_as_float(a.y);
          xv[u][2 % VEC] = __int_as_float(a.z);
          xv[u][3 % VEC] = __int_as_float(a.w);
          tv[u][0] = (long long)(((unsigned long long)(uint32_t)t0.y << 32) | (uint32_t)t0.x);
          tv[u][1 % VEC] = (long long)(((unsigned long long)(uint32_t)t0.w << 32) | (uint32_t)t0.z);
          tv[u][2 % VEC] = (long long)(((unsigned long long)(uint32_t)t1.y << 32) | (uint32_t)t1.x);
          tv[u][3 % VEC] = (long long)(((unsigned long long)(uint32_t)t1.w << 32) | (uint32_t)t1.z);
        } else {
          xv[u][0] = __ldg(x + idx);
          tv[u][0] = __ldg(target + idx);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t idx = base + (int64_t)u * stride;
      if (idx < nvec) {
#pragma unroll
        for (int v = 0; v < VEC; ++v) hit(xv[u][v], tv[u][v]);
      }
    }
  }
  if (VEC > 1 && first == 0)
    for (int64_t i = nvec * VEC; i < n; ++i) hit(x[i], target[i]);

  if (SMEM) {
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * (T + 1); i += blockDim.x) {
      const uint32_t c = sh[i];
      if (c) atomicAdd(hist + i, (unsigned long long)c);
    }
  }
}

// Suffix sums of the (bucket, label) histogram: x >= thr[j]  <=>  bucket >= j + 1.
__global__ void __launch_bounds__(kBinBlock)
    binned_counts_kernel(const long long* __restrict__ hist, int T, long long* num_tp,
                         long long* num_fp, long long* num_fn) {
  __shared__ long long ws[kBinBlock / 32];
  const int task = blockIdx.x;
  hist += (size_t)task * 2 * (T + 1);
  num_tp += (size_t)task * T;
  num_fp += (size_t)task * T;
  if (num_fn) num_fn += (size_t)task * T;
  long long carry_tp = 0, carry_fp = 0;
  for (int top = T; top > 0; top -= kBinBlock) {
    const int j = top - 1 - (int)threadIdx.x;
    const long long a = j >= 0 ? hist[2 * (j + 1) + 1] : 0;
    const long long b = j >= 0 ? hist[2 * (j + 1) + 0] : 0;
    long long tot_a, tot_b;
    const long long sa = block_scan_incl<long long, kBinBlock>(a, ws, tot_a);
    const long long sb = block_scan_incl<long long, kBinBlock>(b, ws, tot_b);
    if (j >= 0) {
      num_tp[j] = carry_tp + sa;
      num_fp[j] = carry_fp + sb;
    }
    carry_tp += tot_a;
    carry_fp += tot_b;
  }
  if (num_fn) {
    const long long total_pos = carry_tp + hist[1];
    for (int top = T; top > 0; top -= kBinBlock) {
      const int j = top - 1 - (int)threadIdx.x;
      if (j >= 0) num_fn[j] = total_pos - num_tp[j];
    }
  }
}

// mode 0: trapezoid AUROC over the thresholds (binned_auroc.py:120-136), 0.5 when degenerate.
// mode 1: step-integrated AUPRC (binned_precision_recall_curve.py:113-129 + tensor_utils.py:12-16,
//         NaN -> 0 as binned_auprc.py:110).
__global__ void __launch_bounds__(kBinBlock)
    binned_integrate_kernel(const long long* __restrict__ num_tp,
                            const long long* __restrict__ num_fp,
                            const long long* __restrict__ num_fn, int T, int mode, double* out) {
  __shared__ double ws[kBinBlock / 32];
  const int task = blockIdx.x;
  num_tp += (size_t)task * T;
  num_fp += (size_t)task * T;
  if (num_fn) num_fn += (size_t)task * T;
  double acc = 0.0;
  for (int j = threadIdx.x; j < T; j += blockDim.x) {
    const double tp = (double)num_tp[j], fp = (double)num_fp[j];
    const double tpn = j + 1 < T ? (double)num_tp[j + 1] : 0.0;
    const double fpn = j + 1 < T ? (double)num_fp[j + 1] : 0.0;
    if (mode == 0) {
      acc += (fp - fpn) * (tp + tpn) * 0.5;
    } else {
      const double pos = tp + (double)num_fn[j];  // total positives, same for every j
      const double prec = (tp + fp) > 0.0 ? tp / (tp + fp) : 1.0;
      const double rec = tp / pos;
      const double recn = j + 1 < T ? tpn / pos : 0.0;
      acc += (rec - recn) * prec;
    }
  }
  double total;
  block_scan_incl<double, kBinBlock>(acc, ws, total);
  if (threadIdx.x == 0) {
    if (mode == 0) {
      const double factor = (double)num_tp[0] * (double)num_fp[0];
      out[task] = factor == 0.0 ? 0.5 : total / factor;
    } else {
      out[task] = total == total ? total : 0.0;
    }
  }
}


// ---- (N, C) inputs: multiclass (label = target[n] == c) and multilabel (label = target[n, c])
// binned counting (binned_precision_recall_curve.py:239-291 and :434-486, the "memory"
// formulations; the "vectorized" ones :214-236 / :406-431 count the same thing).
// hist[c][bucket][label]; thresholds and lookup table in smem; bins in smem when C*(T+1)*2 u32
// fit, otherwise RED into the int64 histogram.
template <bool MULTILABEL, bool SMEM_HIST>
__global__ void __launch_bounds__(kBinBlock, 1)
    binned_hist2d_kernel(const float* __restrict__ x, const long long* __restrict__ target,
                         int64_t n, int C, const float* __restrict__ thr, int T,
                         const uint32_t* __restrict__ lut, int G, bool vec_ok,
                         unsigned long long* hist) {
  extern __shared__ __align__(16) unsigned char smraw[];
  float* sthr = reinterpret_cast<float*>(smraw);
  uint16_t* slut = reinterpret_cast<uint16_t*>(sthr + T);
  uint32_t* sh = reinterpret_cast<uint32_t*>(smraw + (((size_t)4 * T + (size_t)2 * (G + 1) + 15) & ~(size_t)15));
  const int nbins = C * (T + 1) * 2;
  for (int i = threadIdx.x; i < T; i += blockDim.x) sthr[i] = thr[i];
  for (int i = threadIdx.x; i <= G; i += blockDim.x) slut[i] = (uint16_t)lut[i];
  if (SMEM_HIST)
    for (int i = threadIdx.x; i < nbins; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  const float lo = sthr[0], hi = sthr[T - 1];
  const float scale = cell_scale(lo, hi, G);

  const int64_t total = n * C;
  const int64_t ngroups = (total + 3) / 4;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g < ngroups) {
    // (row, col) of the first element of my group, advanced incrementally
    int64_t row = (g * 4) / C;
    int col = (int)((g * 4) - row * C);
    const int64_t step = stride * 4;
    const int64_t step_rows = step / C;
    const int step_cols = (int)(step - step_rows * C);
    for (; g < ngroups; g += stride) {
      const int64_t e0 = g * 4;
      const int cnt = (int)min((int64_t)4, total - e0);
      float xv[4];
      if (cnt == 4 && vec_ok) {
        const int4 a = ld_stream_v4(x + e0);
        xv[0] = __int_as_float(a.x), xv[1] = __int_as_float(a.y);
        xv[2] = __int_as_float(a.z), xv[3] = __int_as_float(a.w);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) xv[j] = j < cnt ? x[e0 + j] : 0.f;
      }
      int64_t r = row;
      int c = col;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (j < cnt) {
          const long long tv = MULTILABEL ? target[e0 + j] : target[r];
          const int label = MULTILABEL ? (tv != 0 ? 1 : 0) : (tv == (long long)c ? 1 : 0);
          const int b = bucket_of<uint16_t>(xv[j], lo, hi, scale, G, T, sthr, slut);
          const int bin = (c * (T + 1) + b) * 2 + label;
          if (SMEM_HIST)
            atomicAdd(sh + bin, 1u);
          else
            atomicAdd(hist + bin, 1ull);
          if (++c == C) {
            c = 0;
            ++r;
          }
        }
      }
      row += step_rows;
      col += step_cols;
      if (col >= C) {
        col -= C;
        ++row;
      }
    }
  }
  if (SMEM_HIST) {
    __syncthreads();
    for (int i = threadIdx.x; i < nbins; i += blockDim.x) {
      const uint32_t v = sh[i];
      if (v) atomicAdd(hist + i, (unsigned long long)v);
    }
  }
}

}  // namespace
}  // namespace te

using namespace te;

extern "C" size_t te_binned_scratch_bytes(int32_t num_thresholds) {
  if (num_thresholds < 1) return 0;
  const BinGeom g = binned_geom(num_thresholds);
  return align_up((size_t)(g.cells + 1) * 4, 256);
}

extern "C" int te_binned_hist_update(const float* x, int64_t x_task_stride, const int64_t* target,
                                     int64_t target_task_stride, int64_t n, int32_t num_tasks,
                                     const float* threshold, int32_t num_thresholds,
                                     int64_t* hist, void* scratch, size_t scratch_bytes,
                                     void* stream) {
  if (n < 0 || num_tasks < 1 || num_tasks > 65535 || num_thresholds < 1 || !threshold || !hist)
    return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!x || !target)) return TE_ERR_INVALID_ARGUMENT;
  if (!scratch || scratch_bytes < te_binned_scratch_bytes(num_thresholds))
    return TE_ERR_SCRATCH_TOO_SMALL;
  if (n == 0) return TE_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int T = num_thresholds;
  const BinGeom g = binned_geom(T);
  uint32_t* lut = (uint32_t*)scratch;
  {
    TE_PROF("binned_lut", st);
    binned_lut_kernel<<<(g.cells + 1 + 255) / 256, 256, 0, st>>>(threshold, T, lut, g.cells);
  }
  TE_LAUNCH_CHECK();
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  const bool vec = ((((uintptr_t)x | (uintptr_t)target) & 15) == 0) &&
                   (num_tasks == 1 || (x_task_stride % 4 == 0 && target_task_stride % 2 == 0));
  const int64_t nvec = vec ? (n + 3) / 4 : n;
  int64_t want = (nvec + kBinBlock * 2 - 1) / (kBinBlock * 2);
  if (want < 1) want = 1;
  int gx = (int)(want < sms ? want : sms);
  if (num_tasks > 1) {
    // keep roughly one wave of CTAs in total
    int per = sms / num_tasks;
    if (per < 1) per = 1;
    if (gx > per) gx = per;
  }
  dim3 grid(gx, num_tasks);
  unsigned long long* h = (unsigned long long*)hist;
  const long long* tg = (const long long*)target;
  TE_PROF(g.smem ? "binned_hist_smem" : "binned_hist_global", st);
  if (g.smem) {
    if (vec) {
      TE_CUDA_TRY(cudaFuncSetAttribute(binned_hist_kernel<4, true>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
      binned_hist_kernel<4, true><<<grid, kBinBlock, g.smem, st>>>(
          x, x_task_stride, tg, target_task_stride, n, threshold, T, lut, g.cells, h);
    } else {
      TE_CUDA_TRY(cudaFuncSetAttribute(binned_hist_kernel<1, true>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
      binned_hist_kernel<1, true><<<grid, kBinBlock, g.smem, st>>>(
          x, x_task_stride, tg, target_task_stride, n, threshold, T, lut, g.cells, h);
    }
  } else {
    if (vec)
      binned_hist_kernel<4, false><<<grid, kBinBlock, 0, st>>>(
          x, x_task_stride, tg, target_task_stride, n, threshold, T, lut, g.cells, h);
    else
      binned_hist_kernel<1, false><<<grid, kBinBlock, 0, st>>>(
          x, x_task_stride, tg, target_task_stride, n, threshold, T, lut, g.cells, h);
  }
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_binned_counts(const int64_t* hist, int32_t num_thresholds, int32_t num_tasks,
                                int64_t* num_tp, int64_t* num_fp, int64_t* num_fn, void* stream) {
  if (!hist || !num_tp || !num_fp || num_thresholds < 1 || num_tasks < 1)
    return TE_ERR_INVALID_ARGUMENT;
  TE_PROF("binned_counts", (cudaStream_t)stream);
  binned_counts_kernel<<<num_tasks, kBinBlock, 0, (cudaStream_t)stream>>>(
      (const long long*)hist, num_thresholds, (long long*)num_tp, (long long*)num_fp,
      (long long*)num_fn);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_binned_integrate(const int64_t* num_tp, const int64_t* num_fp,
                                   const int64_t* num_fn, int32_t num_thresholds,
                                   int32_t num_tasks, int32_t mode, double* out, void* stream) {
  if (!num_tp || !num_fp || !out || num_thresholds < 1 || num_tasks < 1 ||
      (mode != TE_BINNED_AUROC && mode != TE_BINNED_AUPRC) ||
      (mode == TE_BINNED_AUPRC && !num_fn))
    return TE_ERR_INVALID_ARGUMENT;
  TE_PROF("binned_integrate", (cudaStream_t)stream);
  binned_integrate_kernel<<<num_tasks, kBinBlock, 0, (cudaStream_t)stream>>>(
      (const long long*)num_tp, (const long long*)num_fp, (const long long*)num_fn,
      num_thresholds, mode, out);
  TE_LAUNCH_CHECK();
  return TE_OK;
}

extern "C" int te_binned_hist2d_update(const float* x, const int64_t* target, int64_t n,
                                       int32_t num_classes, int32_t multilabel,
                                       const float* threshold, int32_t num_thresholds,
                                       int64_t* hist, void* scratch, size_t scratch_bytes,
                                       void* stream) {
  if (n < 0 || num_classes < 1 || num_thresholds < 1 || !threshold || !hist)
    return TE_ERR_INVALID_ARGUMENT;
  if (n > 0 && (!x || !target)) return TE_ERR_INVALID_ARGUMENT;
  if (!scratch || scratch_bytes < te_binned_scratch_bytes(num_thresholds))
    return TE_ERR_SCRATCH_TOO_SMALL;
  const int T = num_thresholds;
  const BinGeom g = binned_geom(T);
  const size_t tab = (((size_t)4 * T + (size_t)2 * (g.cells + 1) + 15) & ~(size_t)15);
  if (T >= 65536 || tab > kBinSmemBudget) return TE_ERR_UNSUPPORTED;
  if ((int64_t)num_classes * (T + 1) * 2 > 0x7fffffffLL) return TE_ERR_UNSUPPORTED;
  if (n == 0) return TE_OK;
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* lut = (uint32_t*)scratch;
  {
    TE_PROF("binned_lut", st);
    binned_lut_kernel<<<(g.cells + 1 + 255) / 256, 256, 0, st>>>(threshold, T, lut, g.cells);
  }
  TE_LAUNCH_CHECK();
  int sms = 0;
  TE_CUDA_TRY((cudaError_t)sm_count_of_current_device(&sms));
  const size_t hist_bytes = (size_t)num_classes * (T + 1) * 2 * 4;
  const bool smem_hist = tab + hist_bytes <= kBinSmemBudget;
  const size_t smem = tab + (smem_hist ? hist_bytes : 0);
  const int64_t ngroups = (n * num_classes + 3) / 4;
  int64_t want = (ngroups + kBinBlock * 2 - 1) / (kBinBlock * 2);
  const int grid = (int)(want < 1 ? 1 : (want < sms ? want : sms));
  const bool vec_ok = ((uintptr_t)x & 15) == 0;
  unsigned long long* h = (unsigned long long*)hist;
  const long long* tg = (const long long*)target;
  TE_PROF(smem_hist ? "binned_hist2d_smem" : "binned_hist2d_global", st);
#define TE_LAUNCH_2D(ML, SH)                                                                     \
  do {                                                                                           \
    TE_CUDA_TRY(cudaFuncSetAttribute(binned_hist2d_kernel<ML, SH>,                               \
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   \
    binned_hist2d_kernel<ML, SH><<<grid, kBinBlock, smem, st>>>(x, tg, n, num_classes, threshold, \
                                                                 T, lut, g.cells, vec_ok, h);    \
  } while (0)
  if (multilabel) {
    if (smem_hist)
      TE_LAUNCH_2D(true, true);
    else
      TE_LAUNCH_2D(true, false);
  } else {
    if (smem_hist)
      TE_LAUNCH_2D(false, true);
    else
      TE_LAUNCH_2D(false, false);
  }
#undef TE_LAUNCH_2D
  TE_LAUNCH_CHECK();
  return TE_OK;
}
