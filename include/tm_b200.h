/*
 * tm_b200.h — C ABI of the B200-native metric-state hot path (libtm_b200.so).
 *
 * The reference (torchmetrics 1.4.0dev) is pure Python: its "FFI" for this path is the set of private
 * `_*_update` functions in src/torchmetrics/functional/classification/ that the Metric classes call.
 * Each entry point below replaces the device work of one of those functions; the citation names it
 * (paths relative to /root/reference/src/torchmetrics).  INTEGRATION.md shows the ctypes binding a
 * maintainer of the reference would add at each call site.
 *
 * Conventions
 *   - plain pointers and sizes only; no torch types.  `stream` is a cudaStream_t passed as void*
 *     (NULL = the legacy default stream).  All work is enqueued on `stream`; nothing synchronises
 *     unless the function name ends in `_host`.
 *   - every pointer named *_dev is device memory of the CURRENT device (cudaSetDevice by the caller).
 *   - count outputs are int64 and are ACCUMULATED INTO (state += update), never overwritten, so a
 *     metric state can be passed directly; pass a zeroed buffer to get the bare update.
 *   - return value: TMB_OK or a negative TMB_ERR_* (argument / launch errors, detected on the host).
 *     Data errors (labels out of range, unsorted thresholds) are detected on the device and reported
 *     through `status_dev`, TWO int32 words: [0] collects error bits (LABEL_RANGE, THRESHOLD_ORDER,
 *     OUT_OF_UNIT_RANGE), [1] the fused-validation findings (TARGET_VALUES, PREDS_VALUES); the caller reads them when it wants
 *     to pay for the synchronisation (the Python layer does so lazily, on compute()).
 *   - vector loads need natural alignment of the element type times the vector width: every data
 *     pointer must be 32-byte aligned (torch allocations are; offset views must be copied first).
 *   - the streaming kernels of tmb_confmat_update_ws and tmb_binned_curve_update* are launched with the
 *     programmatic-stream-serialization attribute: their set-up (shared-memory clears, barrier init)
 *     may run under the tail of the kernel that precedes them in `stream`; they execute
 *     griddepcontrol.wait before the first access to global memory, so the results are those of plain
 *     stream order whatever that preceding kernel is.
 */
#ifndef TM_B200_H
#define TM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TMB_ABI_VERSION 1

/* the library is built with -fvisibility=hidden; only these entry points are exported */
#if defined(__GNUC__)
#define TMB_API __attribute__((visibility("default")))
#else
#define TMB_API
#endif

/* return codes */
#define TMB_OK 0
#define TMB_ERR_ARG (-1)        /* bad size / null pointer / unsupported dtype            */
#define TMB_ERR_ALIGN (-2)      /* a data pointer is not 32-byte aligned                  */
#define TMB_ERR_CUDA (-3)       /* a CUDA runtime call failed; see tmb_last_cuda_error()  */
#define TMB_ERR_TOO_LARGE (-4)  /* problem does not fit this entry point's stated limits  */

/* bits written (OR-ed) into status_dev[0] by kernels */
#define TMB_STATUS_LABEL_RANGE 1      /* key = target*C + pred outside [0, C*C): the reference raises here */
#define TMB_STATUS_THRESHOLD_ORDER 2  /* thresholds not ascending                                         */
#define TMB_STATUS_OUT_OF_UNIT_RANGE 4 /* TMB_PROB_CHECK: a score outside [0,1] was seen, update discarded   */
/* Fused input validation (what the reference's `_*_tensor_validation` spends torch.unique sorts on): every kernel
 * also reports, for free, whether a label it read lies outside the legal set.  A clear bit proves the reference's
 * value checks would pass; a set bit sends the caller to the exact (slow) check for the error message. */
#define TMB_STATUS_TARGET_VALUES 8    /* a non-ignored target outside [0, C) (K1, class-index K3) or {0, 1} (K2, K3) */
#define TMB_STATUS_PREDS_VALUES 16    /* an integer pred outside [0, C) (K1) or {0, 1} (K2)                         */
#define TMB_STATUS_PEER_TIMEOUT 32    /* word 0: tmb_oneshot_allreduce gave up waiting for a peer (result invalid)  */

/* element types (labels: I64, I32, U8; scores: F32, F16, BF16, F64 — cast anything else first) */
#define TMB_I64 0
#define TMB_I32 1
#define TMB_U8 4  /* also torch.bool */
#define TMB_F32 8
#define TMB_F16 9
#define TMB_BF16 10
#define TMB_F64 11

/* how float scores are mapped to probabilities before thresholding */
#define TMB_PROB_AS_IS 0 /* use the scores unchanged                                                              */
#define TMB_PROB_AUTO 1  /* reference rule: sigmoid the whole batch iff any (non-ignored) score lies outside [0,1] */
#define TMB_PROB_SIGMOID 2 /* always apply sigmoid                                                                 */
#define TMB_PROB_CHECK 3 /* as-is, but if any score lies outside [0,1] discard the update and set
                            TMB_STATUS_OUT_OF_UNIT_RANGE (multiclass: the caller then applies softmax itself)     */

/* reductions of tmb_oneshot_allreduce */
#define TMB_REDUCE_SUM 0
#define TMB_REDUCE_MAX 1
#define TMB_REDUCE_MIN 2
#define TMB_ONESHOT_SLOT_BYTES 65536 /* largest arena one call can reduce */
#define TMB_ONESHOT_MAX_SEGMENTS 16  /* source tensors one packed call can take */

TMB_API int tmb_abi_version(void);
TMB_API const char* tmb_last_cuda_error(void);
/* number of SMs of the current device (grid sizing; 148 on B200) */
TMB_API int tmb_sm_count(void);

/* ---------------------------------------------------------------------------------------------
 * K1  confusion-matrix update from label tensors.
 * Replaces: functional/classification/confusion_matrix.py:333-337 `_multiclass_confusion_matrix_update`
 *           (key = target*C + pred; `_bincount`, utilities/data.py:179-207) fused with the ignore_index
 *           filter of `_multiclass_confusion_matrix_format` (:325-328); same kernel serves the
 *           confusion-matrix branch of `_multiclass_stat_scores_update` (stat_scores.py:405-418) and
 *           `_binary_confusion_matrix_update` (:149-153, C = 2).
 * preds_dev/target_dev: n labels of the given integer dtype.  confmat_dev: C*C int64, row = target.
 * Keys outside [0, C*C) set TMB_STATUS_LABEL_RANGE and are not counted (the reference raises).
 * algo: 0 = choose by C; 1 = global 64-bit RED straight into confmat_dev (lanes holding the same cell are merged
 *       first when a warp looks concentrated); 2 = shared-memory privatised; 3/4/5 = RED variants for tuning.
 * ------------------------------------------------------------------------------------------- */
TMB_API int tmb_confmat_update(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype, int64_t n,
                       int64_t num_classes, int has_ignore, int64_t ignore_index, int64_t* confmat_dev,
                       int32_t* status_dev, int algo, void* stream);

/* K1 for large C through a bucket partition (phase A: queue 12-bit local bins per bucket, phase B: per-bucket
 * shared-memory histograms) instead of one L2 RED per sample; same contract and result as tmb_confmat_update.
 * Serves ceil(C*C / 4096) <= 1024 buckets (C <= 2048) and n < 2^32.  workspace_dev: at least
 * tmb_confmat_workspace_bytes(n, C) bytes (more = roomier queues), contents irrelevant; samples that do not fit a
 * bucket's queue are still counted, through the slower direct RED.  Not for launches that overlap on one workspace. */
TMB_API size_t tmb_confmat_workspace_bytes(int64_t n, int64_t num_classes);
TMB_API int tmb_confmat_update_ws(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype,
                                  int64_t n, int64_t num_classes, int has_ignore, int64_t ignore_index,
                                  int64_t* confmat_dev, int32_t* status_dev, void* workspace_dev,
                                  size_t workspace_bytes, void* stream);

/* Plain histogram: out_dev[x[i]] += 1 for x[i] in [0, num_bins); values outside set TMB_STATUS_LABEL_RANGE.
 * Replaces: utilities/data.py:179-207 `_bincount` (torch.bincount and its deterministic-mode mesh fallback). */
TMB_API int tmb_bincount_update(const void* x_dev, int x_dtype, int64_t n, int64_t num_bins, int64_t* out_dev,
                                int32_t* status_dev, void* stream);

/* Multiclass micro statistics: n_valid and n_equal (tp = n_equal, fp = fn = n_valid - n_equal).
 * Replaces: stat_scores.py:394-404 (the `average="micro"` branch).  out2_dev: int64[2] += {n_valid, n_equal}. */
TMB_API int tmb_match_count_update(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype, int64_t n,
                           int has_ignore, int64_t ignore_index, int64_t* out2_dev, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K3  binned (multi-threshold) curve state update.
 * Replaces: functional/classification/precision_recall_curve.py:190-250
 *           `_binary_precision_recall_curve_update` (+ `_update_vectorized`, `_update_loop`) fused with the
 *           per-element steps of `_binary_precision_recall_curve_format` (:176-184: ignore filter and the
 *           all-or-nothing sigmoid), and, with num_columns > 1, `_multilabel_..._update` (:769-793) /
 *           `_multiclass_..._update` (:458-527).
 * preds_dev: n_rows*num_columns scores, row-major.  target_dev: num_columns == 1 or !target_is_class_index:
 * same shape, positive iff == 1; target_is_class_index: n_rows class ids (one-hot on the fly, multiclass).
 * thresholds_dev: T ascending float32 values (the caller's own tensor: never recomputed).
 * confmat_dev: (T, num_columns, 2, 2) int64, [t][c][target][pred] as in the reference, accumulated.
 * workspace_dev: tmb_binned_curve_workspace_bytes() bytes, ZERO on first use; the kernels leave it zero.
 * One workspace must not be shared by calls that may overlap on the GPU.
 * ------------------------------------------------------------------------------------------- */
TMB_API size_t tmb_binned_curve_workspace_bytes(int num_thresholds, int64_t num_columns);
TMB_API int tmb_binned_curve_update(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype,
                            int64_t n_rows, int64_t num_columns, int target_is_class_index,
                            const float* thresholds_dev, int num_thresholds, int has_ignore, int64_t ignore_index,
                            int prob_mode, int64_t* confmat_dev, void* workspace_dev, size_t workspace_bytes,
                            int32_t* status_dev, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K3 without any allocation.  This entry point has no counterpart in the reference.
 * Same computation and arguments as tmb_binned_curve_update, minus the workspace: the scratch is a fixed
 * TMB_BINNED_STATIC_SCRATCH_BYTES array built into the library (each GPU that loads the library keeps it until the
 * process exits), and the call performs no cudaMalloc / cudaFree and uses no allocating routine.  It is for callers
 * that cannot allocate device memory when they update a curve: loops that forbid allocation after warm-up because
 * cudaMalloc / cudaFree stall the whole device, processes whose device memory already belongs to another pool, and
 * bindings that only ever receive pointers and have no device allocator to call.
 * Limit: num_columns * 2 * (num_thresholds + 1) * 8 + 32 bytes must fit the array (binary curve: up to 4093
 * thresholds; T = 200: up to 20 columns); larger problems return TMB_ERR_TOO_LARGE with an explanatory
 * tmb_last_cuda_error() text.
 * The first call on a device looks up the array's device address (cudaGetSymbolAddress); make that first call
 * before starting any stream capture on the device.  All calls on a device share the one array and the call clears
 * the part it uses before launching, so two calls that may overlap on the GPU (different streams, different host
 * threads) race with each other: serialise them, or give the overlapping ones tmb_binned_curve_update with their
 * own workspaces.
 * ------------------------------------------------------------------------------------------- */
#define TMB_BINNED_STATIC_SCRATCH_BYTES 65536
TMB_API int tmb_binned_curve_update_static(const void* preds_dev, int preds_dtype, const void* target_dev,
                                           int target_dtype, int64_t n_rows, int64_t num_columns,
                                           int target_is_class_index, const float* thresholds_dev, int num_thresholds,
                                           int has_ignore, int64_t ignore_index, int prob_mode, int64_t* confmat_dev,
                                           int32_t* status_dev, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K4  exact binary curve: descending sort of the scores + cumulative counts at every distinct score.
 * Replaces: functional/classification/precision_recall_curve.py:28-80 `_binary_clf_curve` (no sample weights):
 *           torch.argsort, two gathers, the `torch.where` compaction with its host sync, and the float32 cumsum.
 * preds_dev: n float32 scores (already formatted: ignored samples removed, sigmoid applied if needed).
 * target_dev: n labels; positive iff == pos_label.
 * fps_dev / tps_dev (int64) and thresholds_dev (float32) need capacity n; the first D entries are written, D =
 * number of distinct scores.  counts2_dev: int64[2] = {number of positives, D}.  Counts are exact integers (the
 * reference holds them in float32 and rounds above 2^24).  n < 2^32 - 4096.
 * workspace_dev: tmb_exact_curve_workspace_bytes(n) bytes, contents irrelevant.
 * ------------------------------------------------------------------------------------------- */
TMB_API size_t tmb_exact_curve_workspace_bytes(int64_t n);
TMB_API int tmb_exact_curve(const float* preds_dev, const void* target_dev, int target_dtype, int64_t n,
                            int64_t pos_label, int64_t* fps_dev, int64_t* tps_dev, float* thresholds_dev,
                            int64_t* counts2_dev, void* workspace_dev, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K4b  exact AUROC and average precision without materialising the curve.
 * Replaces: `_binary_clf_curve` (precision_recall_curve.py:28-80) FOLLOWED BY the reductions of
 *           functional/classification/roc.py:60-78 + utilities/compute.py `_auc_compute_without_check`
 *           (trapezoid area under (fpr, tpr) with the (0, 0) point prepended) and
 *           functional/classification/average_precision.py:39-46 (-sum(diff(recall) * precision[:-1]))
 *           for thresholds=None, max_fpr=None.  Same sort as tmb_exact_curve; the three n-sized outputs and the
 *           torch passes over them are replaced by two sums over the run ends, exact integer terms added in fp64.
 * out4_dev: double[4] = { sum_k (FP_k-FP_{k-1})(TP_k+TP_{k-1}),  sum_k (TP_k-TP_{k-1}) TP_k/(TP_k+FP_k),
 *                         number of positives,  number of distinct scores }
 *           AUROC = out4[0] / (2 * Npos * Nneg),  AP = out4[1] / Npos   (Nneg = n - Npos); the caller handles
 *           Npos == 0 / Nneg == 0 the way the reference does (warnings, zeros / nan).
 * workspace_dev: tmb_exact_curve_workspace_bytes(n) bytes, contents irrelevant.  n < 2^32 - 4096.
 * ------------------------------------------------------------------------------------------- */
TMB_API int tmb_exact_auroc_ap(const float* preds_dev, const void* target_dev, int target_dtype, int64_t n,
                               int64_t pos_label, double* out4_dev, void* workspace_dev, size_t workspace_bytes,
                               void* stream);

/* K4b for one SHARD of a score range partition (distributed exact curve, SURVEY.md section 8(f) N4: no reference
 * counterpart does this step - the reference gathers every sample on every rank, metric.py:427-457, and sorts the
 * whole set there).  The shard holds all samples of a contiguous score range; tp_offset / fp_offset are the numbers
 * of positives / negatives ranked above it (summed over the shards with higher scores).  The two sums of out4 are
 * then this shard's terms of the GLOBAL sums, so adding out4[0] and out4[1] over the shards gives exactly what
 * tmb_exact_auroc_ap returns for the concatenated input; out4[2], out4[3] stay local (positives, distinct scores). */
TMB_API int tmb_exact_auroc_ap_offset(const float* preds_dev, const void* target_dev, int target_dtype, int64_t n,
                                      int64_t pos_label, int64_t tp_offset, int64_t fp_offset, double* out4_dev,
                                      void* workspace_dev, size_t workspace_bytes, void* stream);

/* Regroup (score, target) pairs by score range for the same distributed curve (no reference counterpart: the
 * reference never partitions, it gathers).  splitters_dev: num_ranges - 1 ascending float32 boundaries; a sample goes
 * to range (num_ranges - 1) - #{splitters < score}, NaN to range 0, so range 0 holds the highest scores and equal
 * scores always share a range.  out_scores_dev / out_positive_dev (n entries each) receive the samples grouped by
 * range, range 0 first; the label becomes one byte (target == pos_label).  totals2_dev: int64[2 * num_ranges] =
 * samples per range, then positives per range.  num_ranges <= 64, n < 2^32.
 * workspace_dev: tmb_range_partition_workspace_bytes(n, num_ranges) bytes, contents irrelevant. */
TMB_API size_t tmb_range_partition_workspace_bytes(int64_t n, int num_ranges);
TMB_API int tmb_range_partition(const float* scores_dev, const void* target_dev, int target_dtype, int64_t n,
                                int64_t pos_label, const float* splitters_dev, int num_ranges, float* out_scores_dev,
                                uint8_t* out_positive_dev, int64_t* totals2_dev, void* workspace_dev,
                                size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K1c  per-class tp / fp / tn / fn of multiclass label pairs, without the C x C matrix.
 * Replaces: functional/classification/stat_scores.py:395-419, the non-one-hot branch of
 *           `_multiclass_stat_scores_update` (multidim_average="global", top_k=1, average != "micro"):
 *           bincount(target * C + preds, minlength=C*C).reshape(C, C) followed by diag / sum(0) / sum(1) / sum().
 *           The four vectors are three histograms of length C (tp, column sums, row sums) and a count.
 * preds_dev/target_dev: n integer labels each.  A target equal to ignore_index (when has_ignore) drops the sample,
 * as the reference's `_format` + update do.  Labels outside [0, C) follow the reference's key arithmetic: the pair
 * is counted in cell (key / C, key % C) when key = target*C + pred is inside [0, C*C), otherwise
 * TMB_STATUS_LABEL_RANGE is raised; TMB_STATUS_TARGET_VALUES / _PREDS_VALUES are reported as for K1.
 * out_dev: int64[4 * C] laid out [tp | fp | tn | fn], accumulated.
 * workspace_dev: tmb_class_stats_workspace_bytes(C) bytes, zero on first use, left zero.  n < 2^32.
 * ------------------------------------------------------------------------------------------- */
TMB_API size_t tmb_class_stats_workspace_bytes(int64_t num_classes);
TMB_API int tmb_class_stats_update(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype,
                                   int64_t n, int64_t num_classes, int has_ignore, int64_t ignore_index,
                                   int64_t* out_dev, void* workspace_dev, size_t workspace_bytes,
                                   int32_t* status_dev, void* stream);

/* K1c with top_k > 1.  Replaces the one-hot branch of `_multiclass_stat_scores_update` (stat_scores.py:363-393,
 * multidim_average="global"): `select_topk` one-hot of the scores, one-hot of the target and four masked sums over two
 * (N, C) tensors.  topk_dev: (n, top_k) class ids, row-major — the indices of `preds.topk(top_k, dim=1)`, which is what
 * `select_topk` (utilities/data.py:113-149) uses, so ties resolve identically.  target_dev: n class ids.
 * tp[c] counts samples whose target c is among their k ids, fp[c] samples that list c without it being the target,
 * fn[c] samples with target c not listed.  A target outside [0, C) that is not ignore_index raises
 * TMB_STATUS_LABEL_RANGE (the reference's one_hot fails on it).  Same out / workspace contract as above. */
TMB_API int tmb_class_stats_topk_update(const void* topk_dev, int topk_dtype, const void* target_dev, int target_dtype,
                                        int64_t n, int64_t num_classes, int64_t top_k, int has_ignore,
                                        int64_t ignore_index, int64_t* out_dev, void* workspace_dev,
                                        size_t workspace_bytes, int32_t* status_dev, void* stream);

/* K1c fused with the argmax of `_multiclass_stat_scores_format` (stat_scores.py:333-334) for score inputs, top_k = 1:
 * scores_dev is (n, C) row-major float32 / float16 / bfloat16 / float64; the predicted class of a row is the index of
 * its first maximum with NaN counting as the maximum (torch.argmax's rule).  One warp per row; the (n,) label tensor
 * the reference materialises is never written.  target_dev, out_dev, workspace_dev, status_dev as above. */
TMB_API int tmb_class_stats_argmax_update(const void* scores_dev, int scores_dtype, const void* target_dev,
                                          int target_dtype, int64_t n, int64_t num_classes, int has_ignore,
                                          int64_t ignore_index, int64_t* out_dev, void* workspace_dev,
                                          size_t workspace_bytes, int32_t* status_dev, void* stream);

/* Row argmax with the class ids written out.  Replaces `preds.argmax(dim=1)` in `_multiclass_confusion_matrix_format`
 * (confusion_matrix.py:318-319) / `_multiclass_stat_scores_format` (stat_scores.py:333-334) for (n, C) row-major
 * scores: index of the first maximum per row, NaN counting as the maximum (torch.argmax's rule).
 * out_dev: int32[n] — one of the label dtypes the K1 entry points read. */
TMB_API int tmb_rows_argmax(const void* scores_dev, int scores_dtype, int64_t n, int64_t num_classes, int32_t* out_dev,
                            void* stream);

/* ---------------------------------------------------------------------------------------------
 * K2  tp / fp / tn / fn per label.
 * Replaces: functional/classification/stat_scores.py:120-131 `_binary_stat_scores_update` and :672-681
 *           `_multilabel_stat_scores_update` (multidim_average="global"), fused with `_format`
 *           (:90-117, :647-669: all-or-nothing sigmoid, strict `preds > threshold`, ignore_index -> no count).
 * preds_dev/target_dev: (n_rows, num_labels) row-major (binary: num_labels = 1, n_rows = numel).
 * Float preds are thresholded (strict >, threshold rounded to the preds dtype as torch does for a Python
 * scalar); integer preds are compared to target as labels.
 * range_check_ignored: 1 = ignored positions take part in the TMB_PROB_AUTO range check (stat-scores and
 * multilabel confusion-matrix `_format` order), 0 = they are skipped (binary confusion matrix, :133-142).
 * out_dev: int64[4 * num_labels] laid out [tp | fp | tn | fn], accumulated.
 * workspace_dev: tmb_stat_scores_workspace_bytes() bytes, zero on first use, left zero.
 * ------------------------------------------------------------------------------------------- */
TMB_API size_t tmb_stat_scores_workspace_bytes(int64_t num_labels);
TMB_API int tmb_stat_scores_update(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype,
                           int64_t n_rows, int64_t num_labels, double threshold, int prob_mode, int has_ignore,
                           int64_t ignore_index, int range_check_ignored, int64_t* out_dev, void* workspace_dev,
                           size_t workspace_bytes, int32_t* status_dev /* may be NULL */, void* stream);

/* K2, row-wise form: the four counts of every ROW of (n_rows, width) row-major tensors.
 * Replaces: the `multidim_average="samplewise"` branches of `_binary_stat_scores_update`
 *           (stat_scores.py:127-130; rows = samples, width = the flattened extra dimensions) and
 *           `_multilabel_stat_scores_update` (:677-680; rows = (sample, label) pairs, width = extra dimensions), and
 *           the legacy `_stat_scores` (:861-906) with reduce="samples" / the per-sample forms of (N, C, X) inputs.
 * Float preds are thresholded (strict >) after an optional sigmoid (prob_mode: TMB_PROB_AS_IS or TMB_PROB_SIGMOID —
 * the all-or-nothing range decision is the caller's, it spans the whole batch); integer preds are labels.
 * out_dev: int64[4 * n_rows] laid out [tp | fp | tn | fn], WRITTEN (per-sample rows are cat states, not sums). */
TMB_API int tmb_stat_scores_rows(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype,
                                 int64_t n_rows, int64_t width, double threshold, int prob_mode, int has_ignore,
                                 int64_t ignore_index, int64_t* out_dev, void* stream);

/* K1c, row-wise form: per sample and class tp / fp / tn / fn of (n_rows, width) label tensors.
 * Replaces: the samplewise one-hot branch of `_multiclass_stat_scores_update` (stat_scores.py:363-393, top_k = 1):
 *           no (N, X, C) one-hot tensors, three length-C shared-memory histograms per sample instead.
 * Targets equal to ignore_index are skipped; a label outside [0, num_classes) raises TMB_STATUS_LABEL_RANGE (the
 * reference's one_hot raises).  out_dev: int64[4 * n_rows * num_classes] laid out [tp | fp | tn | fn][row][class],
 * WRITTEN.  TMB_ERR_TOO_LARGE when 3 * num_classes counters exceed one CTA's shared memory. */
TMB_API int tmb_class_stats_rows(const void* preds_dev, int preds_dtype, const void* target_dev, int target_dtype,
                                 int64_t n_rows, int64_t width, int64_t num_classes, int has_ignore,
                                 int64_t ignore_index, int64_t* out_dev, int32_t* status_dev, void* stream);
/* The same with top_k > 1 (`select_topk` + one-hot in the reference, stat_scores.py:363-393): topk_ids_dev holds the
 * class ids of the k highest scores of every element, (n_rows, width, top_k) row-major
 * (`preds.topk(k, dim=1).indices` moved last); a class among them that is not the target counts as a false positive of
 * that class, the target as a hit when it is among them and as a miss otherwise. */
TMB_API int tmb_class_stats_rows_topk(const void* topk_ids_dev, int ids_dtype, const void* target_dev, int target_dtype,
                                      int64_t n_rows, int64_t width, int top_k, int64_t num_classes, int has_ignore,
                                      int64_t ignore_index, int64_t* out_dev, int32_t* status_dev, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused reducer: count states -> score, one launch.
 * Replaces: the chain of elementwise / reduction ops behind `compute()` of the stat-score family:
 *           `_safe_divide` + `_adjust_weights_safe_divide` (utilities/compute.py:46-69) as used by
 *           `_precision_recall_reduce` (precision_recall.py:37-57), `_fbeta_reduce` (f_beta.py:37-57),
 *           `_accuracy_reduce` (accuracy.py:37-86), `_specificity_reduce` (specificity.py:37-58).
 * tp / fp / tn / fn: int64[rows * num_labels], row-major (rows = 1 for `multidim_average="global"`).
 * kind: TMB_STAT_*; accuracy of a multiclass task is TMB_STAT_RECALL (tp / (tp + fn)), of a binary / multilabel task
 *   TMB_STAT_ACCURACY ((tp + tn) / all).  beta is only read for TMB_STAT_FBETA.
 * average: TMB_AVERAGE_NONE (also the reference's "binary"): one score per element, out: float32[rows * num_labels];
 *   _MICRO: the counts are summed over the labels first, out: float32[rows];
 *   _MACRO, _WEIGHTED: out: float32[rows * num_labels] holds the terms w_l * score_l / sum_l w_l whose sum over the
 *   labels is the average (the caller adds them with the same torch.sum the reference uses, so the last bit agrees).
 *   multilabel == 0 gives a class with tp + fp + fn == 0 weight 0 in the macro average.
 * Every value written is bit-identical to the reference's float32 arithmetic. */
#define TMB_STAT_PRECISION 0
#define TMB_STAT_RECALL 1
#define TMB_STAT_SPECIFICITY 2
#define TMB_STAT_ACCURACY 3
#define TMB_STAT_FBETA 4
#define TMB_AVERAGE_NONE 0
#define TMB_AVERAGE_MICRO 1
#define TMB_AVERAGE_MACRO 2
#define TMB_AVERAGE_WEIGHTED 3
TMB_API int tmb_stat_reduce(const int64_t* tp_dev, const int64_t* fp_dev, const int64_t* tn_dev, const int64_t* fn_dev,
                            int64_t rows, int64_t num_labels, int kind, int average, int multilabel, double beta,
                            float* out_dev, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K5  one-shot all-reduce of a small state arena over NVLink peer memory.
 * Replaces: the collective behind `Metric._sync_dist` (metric.py:427-457: one `gather_all_tensors`,
 *           utilities/distributed.py:97-147, per state, then stack + sum) for arenas up to TMB_ONESHOT_SLOT_BYTES:
 *           the packed sum / max / min states of a metric or a whole MetricCollection are a few KB.
 * peer_buffers_dev: DEVICE array of `world` pointers; entry r is rank r's buffer of tmb_oneshot_buffer_bytes() bytes
 *   as mapped into THIS process (symmetric memory / IPC / cuMemMap; entry `rank` is this rank's own buffer).  The
 *   buffers must be zero before the first call.
 * Collective: every rank calls it with the same count / dtype / op and the same `epoch`, a counter that starts at 1
 *   and is incremented by one per call on that set of buffers.  src_dev may equal dst_dev.  dtype: TMB_I64, TMB_I32,
 *   TMB_F32, TMB_F64.  The result is reduced in rank order on every rank (bit-identical everywhere).
 * One kernel (one CTA) per rank: publish the arena in the own buffer, raise a flag on every peer, wait for all peers'
 *   flags, read all peers' arenas with P2P loads.  Calls on one set of buffers must be stream-ordered on every rank.
 *   A peer that does not arrive within ~2 s raises TMB_STATUS_PEER_TIMEOUT in status_dev[0] instead of hanging.
 * ------------------------------------------------------------------------------------------- */
TMB_API size_t tmb_oneshot_buffer_bytes(void);
TMB_API int tmb_oneshot_allreduce(void* const* peer_buffers_dev, int world, int rank, const void* src_dev,
                                  void* dst_dev, int64_t count, int dtype, int op, uint32_t epoch,
                                  int32_t* status_dev, void* stream);
/* The same with the local arena scattered over up to TMB_ONESHOT_MAX_SEGMENTS device tensors of one dtype (the
 * tp / fp / tn / fn states of a metric, say): srcs / counts are HOST arrays; the segments are packed on the way into
 * the rank's slot and dst_dev receives the reduced arena in segment order (sum of counts elements). */
TMB_API int tmb_oneshot_allreduce_packed(void* const* peer_buffers_dev, int world, int rank,
                                         const void* const* srcs_dev_ptrs, const int64_t* counts, int num_segments,
                                         void* dst_dev, int dtype, int op, uint32_t epoch, int32_t* status_dev,
                                         void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TM_B200_H */
