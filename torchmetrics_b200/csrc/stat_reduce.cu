// Fused reducers: tp / fp / tn / fn count states -> precision, recall, specificity, accuracy, F-beta in ONE launch
// (contract: tmb_stat_reduce in include/tm_b200.h).
//
// The reference turns the states into a score with a chain of ~15 tiny elementwise / reduction ops per metric
// (utilities/compute.py:46-69 `_safe_divide`, `_adjust_weights_safe_divide`; precision_recall.py:37-57,
// f_beta.py:37-57, accuracy.py:37-86, specificity.py:37-58), i.e. ~15 kernel launches on a few hundred elements.
// Here a warp takes one row of the (rows, L) states and does the whole chain in registers.  The arithmetic is the
// reference's, operation by operation, in float32 with round-to-nearest (no fused multiply-add), so every value written
// is the same bits torch computes.  For the macro / weighted averages the kernel writes the L terms w_l * score_l / W
// and the caller adds them up with torch.sum: the order in which torch adds float32 terms is its own, and the last bit
// of e.g. the README's 0.19858768582344055 depends on it.
#include "common.cuh"

namespace tmb {

constexpr int kSrThreads = 128;
constexpr int kSrWarps = kSrThreads / 32;

struct SrParams {
  const long long* tp;
  const long long* fp;
  const long long* tn;
  const long long* fn;
  long long rows, L;
  int kind, average, multilabel;
  float one_plus_b2, b2;
  float* out;
};

// numerator and denominator of one score, as the float32 values torch arrives at
__device__ __forceinline__ void sr_num_den(int kind, long long tp, long long fp, long long tn, long long fn,
                                           float opb2, float b2, float& num, float& den) {
  switch (kind) {
    case TMB_STAT_PRECISION:
      num = __ll2float_rn(tp);
      den = __ll2float_rn(tp + fp);
      break;
    case TMB_STAT_RECALL:
      num = __ll2float_rn(tp);
      den = __ll2float_rn(tp + fn);
      break;
    case TMB_STAT_SPECIFICITY:
      num = __ll2float_rn(tn);
      den = __ll2float_rn(tn + fp);
      break;
    case TMB_STAT_ACCURACY:
      num = __ll2float_rn(tp + tn);
      den = __ll2float_rn(tp + tn + fp + fn);
      break;
    default:  // TMB_STAT_FBETA: (1 + b^2) tp / ((1 + b^2) tp + b^2 fn + fp), every step rounded to float32
      num = __fmul_rn(opb2, __ll2float_rn(tp));
      den = __fadd_rn(__fadd_rn(num, __fmul_rn(b2, __ll2float_rn(fn))), __ll2float_rn(fp));
      break;
  }
}
// _safe_divide: a zero denominator is replaced by one
__device__ __forceinline__ float sr_divide(float num, float den) { return __fdiv_rn(num, den == 0.f ? 1.f : den); }

__device__ __forceinline__ long long warp_sum(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__global__ void __launch_bounds__(kSrThreads) stat_reduce_kernel(const SrParams P) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * kSrWarps + (threadIdx.x >> 5);
  if (row >= P.rows) return;
  const long long* tp = P.tp + row * P.L;
  const long long* fp = P.fp + row * P.L;
  const long long* tn = P.tn + row * P.L;
  const long long* fn = P.fn + row * P.L;
  if (P.average == TMB_AVERAGE_NONE) {  // also "binary": one score per element
    for (long long l = lane; l < P.L; l += 32) {
      float num, den;
      sr_num_den(P.kind, tp[l], fp[l], tn[l], fn[l], P.one_plus_b2, P.b2, num, den);
      P.out[row * P.L + l] = sr_divide(num, den);
    }
    return;
  }
  if (P.average == TMB_AVERAGE_MICRO) {  // the counts are summed over the labels first
    long long s_tp = 0, s_fp = 0, s_tn = 0, s_fn = 0;
    for (long long l = lane; l < P.L; l += 32) {
      s_tp += tp[l];
      s_fp += fp[l];
      s_tn += tn[l];
      s_fn += fn[l];
    }
    s_tp = warp_sum(s_tp);
    s_fp = warp_sum(s_fp);
    s_tn = warp_sum(s_tn);
    s_fn = warp_sum(s_fn);
    if (lane == 0) {
      float num, den;
      sr_num_den(P.kind, s_tp, s_fp, s_tn, s_fn, P.one_plus_b2, P.b2, num, den);
      P.out[row] = sr_divide(num, den);
    }
    return;
  }
  // macro / weighted: the terms (w_l * score_l) / sum_l w_l, w_l = tp + fn (weighted) or 1, and 0 for a multiclass
  // class that never occurs (tp + fp + fn == 0)
  const bool weighted = P.average == TMB_AVERAGE_WEIGHTED;
  long long wsum = 0;
  for (long long l = lane; l < P.L; l += 32) {
    if (weighted) wsum += tp[l] + fn[l];
    else wsum += (P.multilabel || tp[l] + fp[l] + fn[l] != 0) ? 1 : 0;
  }
  wsum = warp_sum(wsum);
  const float W = wsum == 0 ? 1.f : __ll2float_rn(wsum);
  for (long long l = lane; l < P.L; l += 32) {
    float num, den;
    sr_num_den(P.kind, tp[l], fp[l], tn[l], fn[l], P.one_plus_b2, P.b2, num, den);
    const float score = sr_divide(num, den);
    const float w = weighted ? __ll2float_rn(tp[l] + fn[l])
                             : ((P.multilabel || tp[l] + fp[l] + fn[l] != 0) ? 1.f : 0.f);
    P.out[row * P.L + l] = __fdiv_rn(__fmul_rn(w, score), W);
  }
}

}  // namespace tmb

using namespace tmb;

extern "C" int tmb_stat_reduce(const int64_t* tp, const int64_t* fp, const int64_t* tn, const int64_t* fn,
                               int64_t rows, int64_t num_labels, int kind, int average, int multilabel, double beta,
                               float* out, void* stream) {
  if (rows < 0 || num_labels < 0 || kind < TMB_STAT_PRECISION || kind > TMB_STAT_FBETA ||
      average < TMB_AVERAGE_NONE || average > TMB_AVERAGE_WEIGHTED)
    return TMB_ERR_ARG;
  if (rows == 0 || (num_labels == 0 && average != TMB_AVERAGE_MICRO)) return TMB_OK;
  if (!tp || !fp || !tn || !fn || !out) return TMB_ERR_ARG;
  if (rows > (int64_t)0x7fffffff * kSrWarps) return TMB_ERR_TOO_LARGE;
  SrParams P{};
  P.tp = (const long long*)tp;
  P.fp = (const long long*)fp;
  P.tn = (const long long*)tn;
  P.fn = (const long long*)fn;
  P.rows = rows;
  P.L = num_labels;
  P.kind = kind;
  P.average = average;
  P.multilabel = multilabel ? 1 : 0;
  const double b2 = beta * beta;  // the reference's Python floats, rounded to float32 when they meet the tensor
  P.one_plus_b2 = (float)(1.0 + b2);
  P.b2 = (float)b2;
  P.out = out;
  const unsigned grid = (unsigned)((rows + kSrWarps - 1) / kSrWarps);
  stat_reduce_kernel<<<grid, kSrThreads, 0, (cudaStream_t)stream>>>(P);
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}
