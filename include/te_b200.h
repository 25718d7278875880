/* te_b200.h — C ABI of libte_b200.so: the sm_100a kernels behind torcheval_b200.
 *
 * This is the drop-in boundary for torcheval's ranking & counting metric hot path.  The
 * reference (Connor-Guo/torcheval 0.0.6) is pure Python over ATen ops, so there is no native
 * FFI to bind to: each entry point below replaces the ATen op sequence of the cited reference
 * function (paths relative to the reference tree, torcheval/metrics/...).  INTEGRATION.md shows
 * the ctypes stub a maintainer of the reference would add at each call site.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is DEVICE memory of the current CUDA device
 *     unless stated otherwise; `stream` is a cudaStream_t passed as void* (NULL = default);
 *   - calls only enqueue work on `stream`; nothing synchronises, nothing is allocated or freed:
 *     the caller asks te_*_scratch_bytes() and passes a scratch buffer (256-byte aligned);
 *   - return value: 0 = TE_OK; negative = TE_ERR_* (argument rejected, nothing enqueued);
 *     positive = the cudaError_t reported by the runtime;
 *   - counters are int64 and ACCUMULATE (+=) so that a metric state can be updated in place.
 */
#ifndef TE_B200_H_
#define TE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TE_B200_ABI_VERSION 1

#if defined(__GNUC__)
#define TE_API __attribute__((visibility("default")))
#else
#define TE_API
#endif

enum {
  TE_OK = 0,
  TE_ERR_INVALID_ARGUMENT = -1,
  TE_ERR_SCRATCH_TOO_SMALL = -2,
  TE_ERR_UNSUPPORTED = -3,
  TE_ERR_STATIC_SCRATCH_TOO_SMALL = -4 /* size exceeds the built-in scratch of a *_noalloc entry */
};

/* ABI version of the loaded library (TE_B200_ABI_VERSION at build time). */
TE_API int te_abi_version(void);
/* Static description of a status code returned by any entry point. */
TE_API const char* te_status_string(int status);

/* Per-kernel timing for bench.py (no counterpart in the reference; off by default).  While
 * enabled, every kernel launch made by this library is bracketed by CUDA events recorded on the
 * launching stream.  te_profile_read() waits for those events and returns the number of distinct
 * kernel names n (<= max_entries): names is filled with n newline-terminated names,
 * total_ms[i] / launches[i] with the summed duration and launch count of name i.
 * te_profile_enable(0 or 1) also discards what was recorded so far. */
TE_API int te_profile_enable(int on);
TE_API int te_profile_read(char* names, size_t names_bytes, double* total_ms, int64_t* launches,
                           int32_t max_entries);

/* Debug aid, meaningful only in a build with -DTE_PHASE_STATS (python -m torcheval_b200._build
 * --stats): out8 (HOST memory) receives the clock cycles that thread 0 of every radix-pass CTA
 * spent in {setup, load, rank, wait, digit phase, look-back, scatter, output}; reset != 0 clears
 * the counters.  Synchronises the device. */
TE_API int te_debug_phase_stats(uint64_t* out8, int reset);

/* ------------------------------------------------------------------------------------------
 * K4  confusion-matrix counting
 * replaces functional/classification/confusion_matrix.py:219-234 (`_update`: vstack +
 * sparse_coo_tensor(...).to_dense()) and the torch.max range checks of :271-282.
 * ------------------------------------------------------------------------------------------ */
enum {
  TE_CM_AUTO = 0,      /* pick by num_classes */
  TE_CM_GLOBAL64 = 1,  /* RED.ADD.64 straight into `cm` (L2 resident)            */
  TE_CM_GLOBAL32 = 2,  /* RED.ADD.32 into scratch bins, folded into `cm`         */
  TE_CM_SMEM_REPL = 3, /* per-CTA smem bins, replicated per lane (tiny C)        */
  TE_CM_SMEM = 4,      /* per-CTA smem bins (C*C*4 bytes must fit in smem)       */
  TE_CM_CLUSTER = 5,   /* bins spread over the distributed smem of a 16/8-CTA cluster */
  TE_CM_BUCKETED = 6   /* two passes: route 16-bit indices by matrix slice, count in smem  */
};

TE_API size_t te_cm_scratch_bytes(int32_t num_classes, int32_t variant);
/* Scratch for n samples.  TE_CM_BUCKETED (and TE_CM_AUTO, which picks it for large num_classes
 * and n >= 2^22 when the scratch passed to te_cm_update is at least this big) needs ~2 bytes per
 * sample; every other variant needs what te_cm_scratch_bytes says.  0 for TE_CM_BUCKETED means
 * the variant does not apply (num_classes^2 / 31 slices must fit 50 000 bins: num_classes <= 1240). */
TE_API size_t te_cm_scratch_bytes_n(int32_t num_classes, int32_t variant, int64_t n);

/* cm[target[i] * C + pred[i]] += 1 for i < n.
 * range[4] is overwritten with {min_pred, max_pred, min_target, max_target} taken over the
 * OUT-OF-RANGE entries only (untouched sentinels INT64_MAX / INT64_MIN mean "all in range");
 * out-of-range samples are not counted.  The host raises the reference's ValueError
 * (confusion_matrix.py:271-282) from these four values. */
TE_API int te_cm_update(const int64_t* pred, const int64_t* target, int64_t n, int32_t num_classes,
                 int64_t* cm, int64_t* range, void* scratch, size_t scratch_bytes,
                 int32_t variant, void* stream);

/* No counterpart in the reference.  te_cm_update for callers that are not allowed to allocate
 * device memory where they call it and therefore have no scratch buffer to pass (a serving
 * process whose HBM was fully handed out at start-up, a plugin inside a runtime that owns all
 * device memory): same inputs, outputs and semantics as te_cm_update(..., TE_CM_GLOBAL32, ...),
 * but the int32 bins live in a static device array inside the library
 * (TE_CM_NOALLOC_MAX_CLASSES^2 counters = 4 MiB, resident on every GPU that loads the library
 * until the process exits).  num_classes above TE_CM_NOALLOC_MAX_CLASSES is refused with
 * TE_ERR_STATIC_SCRATCH_TOO_SMALL and nothing is enqueued.  The call performs no cudaMalloc /
 * cudaFree and calls no allocating routine.  All calls on a device share the one array and it
 * keeps the previous call's counts, so every call clears the part it uses first; calls whose GPU
 * work overlaps in time race on it — issue them one after the other.  The first call on each
 * device looks up the array's address (cudaGetSymbolAddress); make that first call before any
 * stream capture begins on that device. */
#define TE_CM_NOALLOC_MAX_CLASSES 1024
TE_API int te_cm_update_noalloc(const int64_t* pred, const int64_t* target, int64_t n,
                                int32_t num_classes, int64_t* cm, int64_t* range, void* stream);

/* Same with pred = argmax(logits[i, :]) (first maximal index, NaN counts as maximal —
 * torch.argmax, confusion_matrix.py:225-226).  logits is (n, num_classes) float32 with row
 * stride `ld` elements. */
TE_API int te_cm_update_logits(const float* logits, int64_t ld, const int64_t* target, int64_t n,
                        int32_t num_classes, int64_t* cm, int64_t* range, void* stream);

/* Per-class counts for accuracy and F1: counts3 is int64 [3][C] (+=) holding
 * num_tp[c] = #{pred == target == c}, num_label[c] = #{target == c}, num_pred[c] = #{pred == c}
 * — the diagonal, row sums and column sums of the confusion matrix, i.e. the three
 * scatter_(reduce="add") of functional/classification/f1_score.py:184-192 and the two of
 * accuracy.py:275-277.  Exactly one of `pred` (int64 classes) and `logits` ((n, C) float32, row
 * stride ld, argmax'ed) is non-NULL.  `range` as in te_cm_update. */
TE_API int te_class_counts(const int64_t* pred, const float* logits, int64_t ld, const int64_t* target,
                           int64_t n, int32_t num_classes, int64_t* counts3, int64_t* range,
                           void* stream);

/* *out (int64, +=) gets #{pred[i] == target[i]}: micro accuracy / micro F1 without a class
 * count (accuracy.py:269-272, f1_score.py:176-180). */
TE_API int te_match_count(const int64_t* pred, const int64_t* target, int64_t n, int64_t* out,
                          void* stream);

/* ------------------------------------------------------------------------------------------
 * K3  binned (thresholded) counting
 * replaces functional/classification/binned_precision_recall_curve.py:84-110 (`_update`:
 * searchsorted + histc + flip-cumsum-flip) and functional/classification/binned_auroc.py:117-125
 * (the (T, tasks, N) boolean formulation), which count the same thing.
 * ------------------------------------------------------------------------------------------ */
TE_API size_t te_binned_scratch_bytes(int32_t num_thresholds);

/* hist[task][bucket][label] += 1 with bucket = #{j : threshold[j] <= x} in [0, T] and
 * label = (target != 0); hist is int64 [num_tasks][T + 1][2].  `threshold` is float32, sorted
 * ascending (binned_auroc.py:78-82).  x/target rows of task k start at k * *_task_stride
 * elements.  NaN scores fall in bucket 0 (counted at no threshold). */
TE_API int te_binned_hist_update(const float* x, int64_t x_task_stride, const int64_t* target,
                          int64_t target_task_stride, int64_t n, int32_t num_tasks,
                          const float* threshold, int32_t num_thresholds, int64_t* hist,
                          void* scratch, size_t scratch_bytes, void* stream);

/* (n, C) row-major float32 scores.  multilabel == 0: target is (n,) int64 and
 * label = (target[i] == c) (binned_precision_recall_curve.py:239-291); multilabel != 0: target
 * is (n, C) int64 and label = (target[i, c] != 0) (:434-486).  hist is int64 [C][T + 1][2]. */
TE_API int te_binned_hist2d_update(const float* x, const int64_t* target, int64_t n,
                                   int32_t num_classes, int32_t multilabel, const float* threshold,
                                   int32_t num_thresholds, int64_t* hist, void* scratch,
                                   size_t scratch_bytes, void* stream);

/* Suffix sums: num_tp[task][j] = #{label 1, x >= threshold[j]}, num_fp likewise for label 0,
 * num_fn = positives - num_tp (binned_precision_recall_curve.py:102-110).  num_fn may be NULL.
 * Outputs are overwritten. */
TE_API int te_binned_counts(const int64_t* hist, int32_t num_thresholds, int32_t num_tasks,
                     int64_t* num_tp, int64_t* num_fp, int64_t* num_fn, void* stream);

enum { TE_BINNED_AUROC = 0, TE_BINNED_AUPRC = 1 };
/* out[task] (float64) = trapezoid AUROC over the thresholds (binned_auroc.py:120-136; 0.5 when
 * TP*FP == 0) or step-integrated AUPRC (binned_precision_recall_curve.py:113-129 +
 * functional/tensor_utils.py:12-16; NaN -> 0 as binned_auprc.py:110). */
TE_API int te_binned_integrate(const int64_t* num_tp, const int64_t* num_fp, const int64_t* num_fn,
                        int32_t num_thresholds, int32_t num_tasks, int32_t mode, double* out,
                        void* stream);

/* ------------------------------------------------------------------------------------------
 * K1  exact binary AUROC / AUPRC (sort + fused scan / tie dedup / integrate)
 * replaces functional/classification/auroc.py:115-152 (`_binary_auroc_compute_jit`) and
 * functional/classification/precision_recall_curve.py:208-231 (`_compute_for_each_class`) +
 * functional/tensor_utils.py:12-16 (`_riemann_integral`).
 * ------------------------------------------------------------------------------------------ */
enum { TE_DTYPE_I64 = 0, TE_DTYPE_I32 = 1, TE_DTYPE_U8 = 2, TE_DTYPE_F32 = 3, TE_DTYPE_F64 = 4 };

/* Conditions the packed path cannot represent; reported in out[4], the caller then uses the
 * general entry point instead. */
enum {
  TE_RANK_FLAG_BAD_LABEL = 1,      /* a target is not exactly 0 or 1 (soft labels)   */
  TE_RANK_FLAG_NEGATIVE_SCORE = 2, /* a score has its sign bit set (other than -0.0) */
  TE_RANK_FLAG_WIDE_SCORE = 4      /* TE_RANK_COUNTING only: a score is >= 2.0 (its packed key
                                      needs 32 bits); call again with TE_RANK_SORT              */
};

/* How te_binary_auroc orders the keys.
 *  TE_RANK_SORT     : four 8-bit radix passes (a full sort) + one fused scan/integrate sweep;
 *  TE_RANK_COUNTING : two 8-bit passes order the keys by their top 16 bits, then every segment of
 *                     2^15 consecutive key values is counted into a shared-memory histogram and
 *                     integrated from the counts (36 instead of 52 bytes per score).  Needs every
 *                     score < 2.0, else out[4] carries TE_RANK_FLAG_WIDE_SCORE;
 *  TE_RANK_AUTO     : TE_RANK_COUNTING from 2^25 scores on (env TE_B200_LEAF_MIN_N), else TE_RANK_SORT. */
enum { TE_RANK_AUTO = 0, TE_RANK_SORT = 1, TE_RANK_COUNTING = 2 };

TE_API size_t te_binary_rank_scratch_bytes(int64_t n);

/* Unweighted binary AUROC of n float32 scores against 0/1 targets of `target_dtype`.
 * out (float64[8], device): [0] AUROC (0.5 when there are no positives or no negatives,
 * auroc.py:146-151)  [1] #positives  [2] #negatives  [3] 2*U (exact integer, as double)
 * [4] TE_RANK_FLAG_* (non-zero: [0..3] are not valid).  n <= 2^31 - 1. */
TE_API int te_binary_auroc(const float* score, const void* target, int32_t target_dtype, int64_t n,
                           double* out, void* scratch, size_t scratch_bytes, void* stream);

/* te_binary_auroc with the structure chosen by the caller (te_binary_auroc itself = TE_RANK_AUTO). */
TE_API int te_binary_auroc_variant(const float* score, const void* target, int32_t target_dtype,
                                   int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                   int32_t variant, void* stream);

/* Same contract for AUPRC (average precision, step integration): out[0] = AP, -0.0 when there
 * are no positives (precision_recall_curve.py:227-229); out[3] = sum of per-positive precisions.
 * A target counts as positive iff it equals 1. */
TE_API int te_binary_auprc(const float* score, const void* target, int32_t target_dtype, int64_t n,
                           double* out, void* scratch, size_t scratch_bytes, void* stream);

/* te_binary_auprc with the structure chosen by the caller (te_binary_auprc itself = TE_RANK_AUTO).
 * TE_RANK_COUNTING runs the counting stage twice over the grouped keys (counts, then the integral
 * with exact prefix counts); it needs scores < 2.0 (else TE_RANK_FLAG_WIDE_SCORE in out[4]) and a
 * 16-byte aligned score array (else it falls back to TE_RANK_SORT by itself). */
TE_API int te_binary_auprc_variant(const float* score, const void* target, int32_t target_dtype,
                                   int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                   int32_t variant, void* stream);

/* General variants: any float32 score (negative values, mixed signs), soft labels (AUROC uses
 * w*t and w*(1-t) as auroc.py:129-130), optional float64 per-sample weights (NULL = 1).  Same
 * out[] layout with [1] = sum w*t, [2] = sum w*(1-t), [3] = the signed cross term; [4] = 0.
 * Sorts (key, index) pairs and gathers target/weight by index in the fused pass. */
TE_API size_t te_binary_rank_general_scratch_bytes(int64_t n);
TE_API int te_binary_auroc_general(const float* score, const void* target, int32_t target_dtype,
                                   const double* weight, int64_t n, double* out, void* scratch,
                                   size_t scratch_bytes, void* stream);

/* `tasks` independent rankings of equal length in ONE launch train (grid.y = task): row y of
 * score / target / weight starts y * ld elements after row 0 (ld >= n), out is [tasks][8] with row y
 * as te_binary_auroc_general writes it.  What WindowedBinaryAUROC.compute and (tasks, n) inputs with
 * weights need (torcheval/metrics/window/auroc.py:176-185 -> _binary_auroc_compute on 2-D tensors);
 * tasks <= 65535. */
TE_API size_t te_binary_auroc_general_batched_scratch_bytes(int64_t tasks, int64_t n);
TE_API int te_binary_auroc_general_batched(const float* score, const void* target, int32_t target_dtype,
                                           const double* weight, int64_t tasks, int64_t n, int64_t ld,
                                           double* out, void* scratch, size_t scratch_bytes, void* stream);
TE_API int te_binary_auprc_general(const float* score, const void* target, int32_t target_dtype,
                                   int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                   void* stream);

/* K2 — one-vs-rest ranking over an (n, C) row-major float32 score matrix (row stride ld):
 * replaces functional/classification/auroc.py:213-235 (multiclass AUROC) and auprc.py:279-347 +
 * precision_recall_curve.py:296-310 (multiclass / multilabel AUPRC, a Python loop over labels).
 * multilabel == 0: target is (n,) int64 class ids, positives of class c are target == c;
 * multilabel != 0: target is (n, C) int64 with row stride target_ld, positives are == 1.
 * mode 0 = AUROC, 1 = AUPRC.  out is float64 [C][8], row c laid out as te_binary_auroc's out
 * (out[c][4] != 0: class c has negative scores, use the general entry point for that column). */
TE_API size_t te_ovr_rank_scratch_bytes(int64_t n);
TE_API int te_ovr_rank(const float* input, int64_t ld, const int64_t* target, int64_t target_ld, int64_t n,
                       int32_t num_classes, int32_t multilabel, int32_t mode, double* out, void* scratch,
                       size_t scratch_bytes, void* stream);

/* Pieces of the distributed sample sort of rank-sharded scores (torcheval_b200/distributed.py;
 * replaces dist.all_gather_object of the whole sample set at toolkit.py:388).
 *  te_rank_pack: score/target -> packed keys ((float bits << 1) | label); info (int64[2], device)
 *    = {#positives, TE_RANK_FLAG_*}; counts (int64[num_splitters + 1], device) = sizes of the
 *    buckets the keys fall in, bucket = #{splitters <= key}, for up to 15 ascending splitters
 *    given in HOST memory (num_splitters may be 0).  keys_out must be 16-byte aligned.
 *    digit_hist (device, may be NULL): uint32[num_splitters + 1][4][256], the four radix-digit
 *    histograms of the keys of every bucket; the rank that receives bucket b sums the rows b of all
 *    senders and hands the result to te_binary_auroc_from_keys / te_binary_auprc_from_keys_begin as
 *    digit_hist (uint32[4][256]; NULL there = histogram the keys first), which saves one read of
 *    the received keys.
 *  te_partition_u32: partition of keys by the same kind of splitters (the order of the keys INSIDE
 *    a bucket is unspecified: they are sorted next); counts_in (device,
 *    may be NULL) are bucket sizes already known from te_rank_pack, counts receives them again,
 *    keys_out the keys grouped by bucket.
 *  te_binary_auroc_from_keys: sort packed keys and integrate; out[] as te_binary_auroc plus
 *    out[5], out[6] = high / low 32 bits of the exact integer 2U and out[7] = n. */
TE_API size_t te_rank_pack_scratch_bytes(int64_t n);
TE_API int te_rank_pack(const float* score, const void* target, int32_t target_dtype, int64_t n,
                        const uint32_t* splitters_host, int32_t num_splitters, uint32_t* keys_out,
                        int64_t* info, int64_t* counts, uint32_t* digit_hist, void* scratch,
                        size_t scratch_bytes, void* stream);
/* te_rank_pack with the splitters in DEVICE memory (uint32[num_splitters], ascending: not
 * checked), for callers that pick them on the device and do not want a host round trip. */
TE_API int te_rank_pack_dsplit(const float* score, const void* target, int32_t target_dtype, int64_t n,
                               const uint32_t* splitters_dev, int32_t num_splitters, uint32_t* keys_out,
                               int64_t* info, int64_t* counts, uint32_t* digit_hist, void* scratch,
                               size_t scratch_bytes, void* stream);
TE_API size_t te_partition_u32_scratch_bytes(int64_t n);
TE_API int te_partition_u32(const uint32_t* keys_in, int64_t n, const uint32_t* splitters_host,
                            int32_t num_splitters, const int64_t* counts_in, uint32_t* keys_out,
                            int64_t* counts, void* scratch, size_t scratch_bytes, void* stream);
/* Partition fused with the exchange: bucket b is not written to a local array but straight to
 * dest_ptrs_host[b] (HOST array of num_splitters + 1 device addresses), typically the peer-mapped
 * inbox of rank b (NVLink P2P stores).  dest_ptrs_host[b] must be the address such that
 * element (exclusive bucket offset of b in a contiguous layout + i) is the i-th key of bucket b,
 * i.e. inbox_b + slot_of_this_rank - sum(counts[0..b-1]) elements.  The caller synchronises the
 * ranks before the inboxes are read. */
TE_API int te_partition_scatter_u32(const uint32_t* keys_in, int64_t n, const uint32_t* splitters_host,
                                    int32_t num_splitters, const int64_t* counts_in,
                                    const uint64_t* dest_ptrs_host, int64_t* counts, void* scratch,
                                    size_t scratch_bytes, void* stream);
TE_API int te_binary_auroc_from_keys(const uint32_t* keys, int64_t n, const uint32_t* digit_hist,
                                     double* out, void* scratch, size_t scratch_bytes, void* stream);
/* te_binary_auroc_from_keys with the structure chosen by the caller (TE_RANK_*; the plain entry
 * point = TE_RANK_AUTO).  TE_RANK_COUNTING ignores digit_hist and reports keys >= 2^31 through
 * TE_RANK_FLAG_WIDE_SCORE in out[4]. */
TE_API int te_binary_auroc_from_keys_variant(const uint32_t* keys, int64_t n, const uint32_t* digit_hist,
                                             double* out, void* scratch, size_t scratch_bytes,
                                             int32_t variant, void* stream);
/* Average precision of rank-sharded scores (precision is a ratio of prefix
 * counts, so the step integral of precision_recall_curve.py:208-231 + tensor_utils.py:12-16 takes
 * two phases around one exchange of (P_r, N_r)):
 *  _begin : sort the keys this rank received and count; out[] as te_binary_auroc_from_keys.  The
 *           sorted keys and the chunk states stay in `scratch`, which the caller must leave
 *           untouched until _finish (same n, same scratch, same device).
 *  _finish: out[3] = sum over this rank's positives of precision at their threshold, with
 *           positives_above / samples_above = positives / samples held by the ranks whose score
 *           ranges lie above this one (exact integers passed as doubles; if above_dev is not
 *           NULL the two numbers are read from that device array instead, so that they can come
 *           straight out of a collective).  AP = sum over ranks of out[3], divided by the global
 *           number of positives; out[1], out[2] = local P, N. */
TE_API int te_binary_auprc_from_keys_begin(const uint32_t* keys, int64_t n, const uint32_t* digit_hist,
                                           double* out, void* scratch, size_t scratch_bytes, void* stream);
TE_API int te_binary_auprc_from_keys_finish(int64_t n, double positives_above, double samples_above,
                                            const double* above_dev, double* out, void* scratch,
                                            size_t scratch_bytes, void* stream);

/* Slab exchange of the distributed sample sort (no counterpart in the reference: toolkit.py:371-391
 * gathers everything everywhere).  Every rank owns an inbox of te_slab_inbox_bytes(world, cap)
 * bytes in peer-mapped memory: a 1 KiB header (world x {fill count, flags} uint32 pairs) followed by
 * `world` slabs of `cap` packed keys, slab q reserved for sender q; cap = te_slab_capacity(largest
 * shard, world).
 *  te_rank_pack_scatter: (score, target) -> packed keys, stored straight into the owning rank's
 *    inbox (rank b owns the keys in [splitter b-1, splitter b); splitters_dev: uint32[world - 1]
 *    in device memory).  inbox_ptrs_host[b] = address of rank b's inbox as mapped into THIS
 *    process (HOST array).  `work`: 128 bytes of device memory (cleared by the call).  The fill
 *    counts and this sender's TE_RANK_FLAG_* bits go into every receiver's header; bit 0x100 there
 *    = a slab overflowed, the result of this exchange must be discarded.  The caller synchronises
 *    the ranks (barrier) before anybody reads an inbox.
 *  te_binary_auroc_from_slabs: AUROC of the keys in this rank's inbox, counts read from its
 *    header on the device (no host read); out[] as te_binary_auroc_from_keys, out[4] = OR of all
 *    senders' flags (and TE_RANK_FLAG_WIDE_SCORE if a key needs 32 bits). */
TE_API size_t te_slab_inbox_bytes(int32_t world, int64_t cap);
TE_API int64_t te_slab_capacity(int64_t n_max_per_rank, int32_t world);
TE_API int te_rank_pack_scatter(const float* score, const void* target, int32_t target_dtype, int64_t n,
                                const uint32_t* splitters_dev, int32_t world, int32_t rank,
                                const uint64_t* inbox_ptrs_host, int64_t cap, void* work, void* stream);
/* Measurement only (tools/p2p_store_bench.py): every CTA of `grid` stores 8 KB chunks into the ndst
 * (<= 16) buffers of words_per_dst uint32 each; mode 0 = 16-byte st.global per lane, mode 1 = one bulk
 * async copy (shared -> global, no multicast) per chunk.  Nothing is read back. */
TE_API int te_debug_p2p_store_bench(const uint64_t* dst_ptrs_host, int32_t ndst, int64_t words_per_dst, int32_t mode,
                                    int32_t grid, void* stream);
TE_API size_t te_binary_auroc_from_slabs_scratch_bytes(int32_t world, int64_t cap);
TE_API int te_binary_auroc_from_slabs(const void* inbox, int32_t world, int64_t cap, double* out, void* scratch,
                                      size_t scratch_bytes, void* stream);

/* Two one-CTA helpers of the distributed sample sort (no counterpart in the reference, which has no
 * sharded path: toolkit.py:371-391 gathers everything everywhere).
 *  te_select_splitters: `samples` = int32[total] float bit patterns of sampled scores from all
 *    ranks (negative = no sample), total <= 32768; writes the world - 1 quantile splitters in key
 *    space ((bits << 1), ascending) as uint32 and as int64.  world <= 16.
 *  te_combine_auroc_stats: `rows` = float64[world][8], the te_binary_auroc_from_keys results of
 *    the ranks in ascending score-range order; out[] = [auroc, P, N, 2U, 0, n_received, 2U >> 32, 2U & 0xffffffff] with
 *    2U = sum_r (2U_r + 2 P_r sum_{q<r} N_q) in exact 64-bit integers. */
TE_API int te_select_splitters(const int32_t* samples, int64_t total, int32_t world, uint32_t* splitters,
                               int64_t* splitters_i64, void* stream);
TE_API int te_combine_auroc_stats(const double* rows, int32_t world, double n_received, double* out,
                                  void* stream);

/* Exact precision-recall curve of (score, target == 1) — precision_recall_curve.py:208-231
 * (_compute_for_each_class: sort, tie mask, cumsum, flip, append (1, 0)).  Two calls around one
 * host read, because the number of points K (= distinct scores) sizes the outputs:
 *  _begin : sort + count; out[0] = K, out[1] = #positives, out[2] = #negatives,
 *           out[4] = TE_RANK_FLAG_* (non-zero: the packed representation does not apply, see
 *           te_binary_auroc).  `scratch` (te_binary_rank_scratch_bytes(n)) keeps the sorted keys and
 *           must reach _finish untouched.
 *  _finish: precision[K + 1], recall[K + 1], thresholds[K] (float32, device), thresholds ascending,
 *           closing point (precision 1, recall 0) last; recall is 1 where there are no positives. */
TE_API int te_binary_pr_curve_begin(const float* score, const void* target, int32_t target_dtype,
                                    int64_t n, double* out, void* scratch, size_t scratch_bytes,
                                    void* stream);
TE_API int te_binary_pr_curve_finish(int64_t n, float* precision, float* recall, float* thresholds,
                                     void* scratch, size_t scratch_bytes, void* stream);

/* The sort on its own: ascending, stable, 32-bit keys with an optional 32-bit payload
 * (vals_in/vals_out both NULL or both set).  Replaces torch.sort at auroc.py:121. */
TE_API size_t te_sort_u32_scratch_bytes(int64_t n, int32_t with_values);
TE_API int te_sort_u32(const uint32_t* keys_in, uint32_t* keys_out, const uint32_t* vals_in,
                       uint32_t* vals_out, int64_t n, void* scratch, size_t scratch_bytes,
                       void* stream);

/* Dense ranks of 64-bit scores: rank_out[i] = #{distinct values < x[i]} (uint32, device), the road
 * for scores the packed 32-bit key cannot hold (float64 probabilities, int64 beyond 2^24).  The
 * reference sorts any dtype in place (functional/classification/auroc.py:121); the ranking kernels
 * here only look at order and ties, which the dense rank carries (as the bit pattern of a
 * non-negative float it goes through te_binary_auroc and friends unchanged).  dtype: TE_DTYPE_F64
 * (-0.0 ties with +0.0; NaNs order above +inf by bit pattern) or TE_DTYPE_I64.  values_out (NULL or
 * n elements of the input's type): values_out[r] = the score of rank r — the thresholds of
 * precision_recall_curve.py:89-98; n_distinct (NULL or one uint32, device) = number of ranks.
 * Two stable pair sorts (low word, high word) with te_sort_u32's passes + a chunked scan. */
TE_API size_t te_dense_rank_scratch_bytes(int64_t n);
TE_API int te_dense_rank(const void* x, int32_t dtype, int64_t n, uint32_t* rank_out, void* values_out,
                         uint32_t* n_distinct, void* scratch, size_t scratch_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TE_B200_H_ */
