/*
 * oracle.c — plain C restatement (OpenMP) of the reference's CPU algorithms for the metric-state hot path.
 *
 * TEST INFRASTRUCTURE ONLY: used by tests/ as a second checker and by bench.py as the timed CPU baseline
 * (`cpu_baseline`, `--impl reference`).  Nothing under torchmetrics_b200/ links or calls it.
 * Parity pinned: tests/test_oracle_golden.py checks these functions against the golden vectors generated from the
 * unmodified reference (tests/golden/make_golden.py).
 *
 * Each function follows the reference's own formulation, pass structure included, so that its timing is a fair
 * stand-in for the reference on CPU (paths relative to /root/reference/src/torchmetrics):
 *   orc_confmat_update      functional/classification/confusion_matrix.py:333-337  (key tensor, then bincount)
 *   orc_binned_update_loop  functional/classification/precision_recall_curve.py:228-250 (one pass per threshold)
 *   orc_multilabel_stats    functional/classification/stat_scores.py:647-681 (threshold, then 4 masked sums)
 */
#include <omp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

int orc_num_threads(void) { return omp_get_max_threads(); }
void orc_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

/* returns 0, or -1 when a key falls outside [0, C*C) (the reference raises there) */
int orc_confmat_update(const int64_t* preds, const int64_t* target, int64_t n, int64_t C, int64_t* confmat /* C*C, += */) {
  const int64_t CC = C * C;
  /* step 1 (:335): unique_mapping = target * C + preds, materialised like the reference does */
  int64_t* key = (int64_t*)malloc((size_t)(n > 0 ? n : 1) * sizeof(int64_t));
  if (!key) return -2;
  int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad)
  for (int64_t i = 0; i < n; ++i) {
    key[i] = target[i] * C + preds[i];
    bad |= (key[i] < 0 || key[i] >= CC);
  }
  if (bad) {
    free(key);
    return -1;
  }
  /* step 2 (:336, utilities/data.py:207): bincount — private histogram per thread, merged at the end */
  const int nt = omp_get_max_threads();
  int64_t* priv = (int64_t*)calloc((size_t)nt * (size_t)CC, sizeof(int64_t));
  if (!priv) {
    free(key);
    return -2;
  }
#pragma omp parallel
  {
    int64_t* h = priv + (size_t)omp_get_thread_num() * (size_t)CC;
#pragma omp for schedule(static)
    for (int64_t i = 0; i < n; ++i) h[key[i]]++;
  }
#pragma omp parallel for schedule(static)
  for (int64_t b = 0; b < CC; ++b) {
    int64_t s = 0;
    for (int t = 0; t < nt; ++t) s += priv[(size_t)t * (size_t)CC + b];
    confmat[b] += s;
  }
  free(priv);
  free(key);
  return 0;
}

/* (T,2,2) state, [i][target][pred]; positives are target == 1, everything else negative (:241) */
void orc_binned_update_loop(const float* preds, const int64_t* target, int64_t n, const float* thr, int T,
                            int64_t* confmat /* T*4, += */) {
  for (int i = 0; i < T; ++i) { /* "Iterate one threshold at a time to conserve memory" (:243) */
    const float th = thr[i];
    int64_t tp = 0, fp = 0, fn = 0;
#pragma omp parallel for schedule(static) reduction(+ : tp, fp, fn)
    for (int64_t k = 0; k < n; ++k) {
      const int above = preds[k] >= th;
      const int pos = target[k] == 1;
      tp += pos & above;
      fp += (!pos) & above;
      fn += pos & (!above);
    }
    confmat[i * 4 + 3] += tp;
    confmat[i * 4 + 1] += fp;
    confmat[i * 4 + 2] += fn;
    confmat[i * 4 + 0] += n - tp - fp - fn;
  }
}

/* preds/target (n, L) row-major, probabilities in [0,1]; out[4][L] = tp|fp|tn|fn, += ; ignored targets (== ignore) skipped */
void orc_multilabel_stats(const float* preds, const int64_t* target, int64_t n, int64_t L, float threshold,
                          int has_ignore, int64_t ignore, int64_t* out) {
  const int nt = omp_get_max_threads();
  int64_t* priv = (int64_t*)calloc((size_t)nt * 4 * (size_t)L, sizeof(int64_t));
#pragma omp parallel
  {
    int64_t* h = priv + (size_t)omp_get_thread_num() * 4 * (size_t)L;
#pragma omp for schedule(static)
    for (int64_t r = 0; r < n; ++r) {
      for (int64_t c = 0; c < L; ++c) {
        const int64_t t = target[r * L + c];
        if (has_ignore && t == ignore) continue;
        const int p = preds[r * L + c] > threshold; /* strict (:660) */
        if (t == 1) {
          if (p) h[0 * L + c]++; else h[3 * L + c]++;
        } else if (t == 0) {
          if (p) h[1 * L + c]++; else h[2 * L + c]++;
        }
      }
    }
  }
  for (int t = 0; t < nt; ++t)
    for (int64_t b = 0; b < 4 * L; ++b) out[b] += priv[(size_t)t * 4 * (size_t)L + b];
  free(priv);
}
