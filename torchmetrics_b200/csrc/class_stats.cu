// K1c — per-class tp / fp / tn / fn of multiclass label pairs without the C x C matrix (contract:
// tmb_class_stats_update in include/tm_b200.h).
//
// The reference derives them from a full confusion matrix (functional/classification/stat_scores.py:395-419:
// bincount(target * C + preds, minlength=C*C).reshape(C, C), then diag / column sums / row sums) — C*C counters
// zeroed, filled and reduced per update (8 MB at C = 1000) to produce 4*C numbers.  The same numbers are three
// histograms of length C:
//     tp[c] = #(target == c and pred == c)      colsum[c] = #(pred == c)      rowsum[c] = #(target == c)
//     fp = colsum - tp,   fn = rowsum - tp,   tn = counted - tp - fp - fn
// Labels outside [0, C) follow the reference's arithmetic exactly: a pair whose key target*C + pred still lands in
// [0, C*C) is counted in cell (key / C, key % C); any other key raises the label-range status bit.
#include "common.cuh"

namespace tmb {

#ifndef TMB_K1C_THREADS
#define TMB_K1C_THREADS 512
#endif
#ifndef TMB_K1C_SMEM_KB
#define TMB_K1C_SMEM_KB 48
#endif
#ifndef TMB_K1C_PER_SM
#define TMB_K1C_PER_SM 3
#endif
#ifndef TMB_K1C_AM_PER_SM
#define TMB_K1C_AM_PER_SM 4
#endif
constexpr int kClassThreads = TMB_K1C_THREADS;
constexpr size_t kClassSmemBudget = TMB_K1C_SMEM_KB * 1024;  // bins per CTA: 3 * C u32 per copy
// CTAs per SM of the label / top-k kernels.  Three, not the four that would fit: measured at C = 1000 on 2^24 / 2^26
// int64 label pairs, 4 per SM 70.0 / 214 us, 3 per SM 53.6 / 162 us (58 -> 76 % and 76 -> 101 % of the measured copy
// peak; profiles/r2_k1c_variants.log) - the fourth CTA's 48 KB of shared memory come out of the L1 the loads stream
// through, and every CTA ends with 3 C atomics on the same 3 C global counters.
constexpr int kClassCtasPerSm = TMB_K1C_PER_SM;
constexpr int kClassArgmaxCtasPerSm = TMB_K1C_AM_PER_SM;

struct ClassStatsWs {
  int ticket;
  int pad;
  u64 total;  // counted samples of the update in flight
  // followed by u64 hist[3 * C]: tp | colsum | rowsum
};

struct ClassParams {
  const void* preds;
  const void* target;
  long long n;
  long long C;
  long long CC;
  int has_ignore;
  long long ignore;
  int copies;  // shared-memory copies of the bins per CTA (0: straight to the global scratch)
  int top_k;   // predicted classes per sample (preds is (n, top_k)); 1 for plain label predictions
  int group;   // argmax form: lanes that share a row (power of two)
  u64* out;    // [4][C] tp | fp | tn | fn, accumulated
  ClassStatsWs* ws;
  int* status;
};

// Shared tail of the two kernels: per-CTA bins -> global scratch, last CTA turns the three histograms into
// tp / fp / tn / fn, adds them to the caller's accumulator and leaves the scratch zero.
template <bool SMEM>
__device__ __forceinline__ void finish(const ClassParams& P, u32* s_bins, u32* s_valid, int* s_last, u32 valid,
                                       int bad) {
  const int tid = threadIdx.x;
  const u32 C = (u32)P.C;
  u64* const hist = reinterpret_cast<u64*>(P.ws + 1);
  if (bad) report_status(P.status, bad);
  valid = __reduce_add_sync(0xffffffffu, valid);
  if ((tid & 31) == 0 && valid) atomicAdd(s_valid, valid);
  __syncthreads();
  if constexpr (SMEM) {
    for (u32 b = tid; b < 3u * C; b += kClassThreads) {
      u64 v = 0;
      for (int w = 0; w < P.copies; ++w) v += s_bins[(u32)w * 3u * C + b];
      if (v) atomicAdd(hist + b, v);
    }
  }
  if (tid == 0 && (*s_valid)) atomicAdd(&P.ws->total, (u64)(*s_valid));
  __threadfence();
  __syncthreads();
  if (tid == 0) *s_last = (atomicAdd(&P.ws->ticket, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!*s_last) return;
  // last CTA: histograms -> tp / fp / tn / fn, added to the caller's accumulator; scratch left zero
  __threadfence();
  const u64 total = __ldcg(&P.ws->total);
  for (u32 c = tid; c < C; c += kClassThreads) {
    const u64 tp = __ldcg(hist + c), col = __ldcg(hist + C + c), row = __ldcg(hist + 2 * C + c);
    hist[c] = 0;
    hist[C + c] = 0;
    hist[2 * C + c] = 0;
    const u64 fp = col - tp, fn = row - tp;
    P.out[c] += tp;
    P.out[C + c] += fp;
    P.out[2 * (u64)C + c] += total - tp - fp - fn;
    P.out[3 * (u64)C + c] += fn;
  }
  __syncthreads();
  if (tid == 0) {
    P.ws->total = 0;
    P.ws->ticket = 0;
  }
}

template <typename PT, typename TT, bool HAS_IGNORE, bool SMEM>
__global__ void __launch_bounds__(kClassThreads) class_stats_kernel(const ClassParams P) {
  extern __shared__ u32 s_bins[];  // [copies][3][C]
  __shared__ int s_last;
  __shared__ u32 s_valid;
  const int tid = threadIdx.x, warp = tid >> 5;
  const u32 C = (u32)P.C;
  u64* const hist = reinterpret_cast<u64*>(P.ws + 1);
  if constexpr (SMEM)
    for (u32 i = tid; i < (u32)P.copies * 3u * C; i += kClassThreads) s_bins[i] = 0;
  if (tid == 0) s_valid = 0;
  __syncthreads();
  u32* const my = SMEM ? s_bins + (u32)(warp % P.copies) * 3u * C : nullptr;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  u32 valid = 0;
  int bad = 0;

  auto count = [&](PT pv, TT tv) {
    const long long tl = as_i64(tv), pl = as_i64(pv);
    if constexpr (HAS_IGNORE)
      if (tl == P.ignore) return;
    u32 r = (u32)tl, c = (u32)pl;
    if (((u64)tl >= (u64)P.C) | ((u64)pl >= (u64)P.C)) {  // rare: the reference still counts a key that aliases
      const long long k = tl * P.C + pl;
      bad |= ((u64)tl >= (u64)P.C ? TMB_STATUS_TARGET_VALUES : 0) | ((u64)pl >= (u64)P.C ? TMB_STATUS_PREDS_VALUES : 0);
      if ((u64)k >= (u64)P.CC) {
        bad |= TMB_STATUS_LABEL_RANGE;
        return;
      }
      r = (u32)(k / P.C);
      c = (u32)(k - (long long)r * P.C);
    }
    ++valid;
    if constexpr (SMEM) {
      atomicAdd(my + C + c, 1u);
      atomicAdd(my + 2 * C + r, 1u);
      if (r == c) atomicAdd(my + c, 1u);
    } else {
      atomicAdd(hist + C + c, 1ULL);
      atomicAdd(hist + 2 * C + r, 1ULL);
      if (r == c) atomicAdd(hist + c, 1ULL);
    }
  };

  const long long nvec = P.n >> 2;
  const long long stride = (long long)gridDim.x * kClassThreads;
  for (long long g = (long long)blockIdx.x * kClassThreads + tid; g < nvec; g += stride) {
    const Vec4<PT> p = load4(preds + 4 * g);
    const Vec4<TT> t = load4(target + 4 * g);
#pragma unroll
    for (int j = 0; j < 4; ++j) count(p.v[j], t.v[j]);
  }
  if (blockIdx.x == 0 && tid < (int)(P.n & 3)) {
    const long long i = (nvec << 2) + tid;
    count(preds[i], target[i]);
  }
  finish<SMEM>(P, s_bins, &s_valid, &s_last, valid, bad);
}

// top_k > 1 (reference: the one-hot branch, stat_scores.py:363-393, with select_topk): every sample carries the k
// class ids of its highest scores.  The same three histograms with "c is among the k" in place of "pred == c":
//     tp[c] = #(target == c and c in topk)   colsum[c] = #(c in topk)   rowsum[c] = #(target == c)
// so a sample costs k + 1 (+1) increments instead of two C-wide one-hot rows.  A target outside [0, C) that is not the
// ignore value is what makes the reference's one_hot fail: label-range status bit.
template <typename PT, typename TT, bool HAS_IGNORE, bool SMEM>
__global__ void __launch_bounds__(kClassThreads) class_stats_topk_kernel(const ClassParams P) {
  extern __shared__ u32 s_bins[];
  __shared__ int s_last;
  __shared__ u32 s_valid;
  const int tid = threadIdx.x, warp = tid >> 5;
  const u32 C = (u32)P.C;
  const int k = P.top_k;
  u64* const hist = reinterpret_cast<u64*>(P.ws + 1);
  if constexpr (SMEM)
    for (u32 i = tid; i < (u32)P.copies * 3u * C; i += kClassThreads) s_bins[i] = 0;
  if (tid == 0) s_valid = 0;
  __syncthreads();
  u32* const my = SMEM ? s_bins + (u32)(warp % P.copies) * 3u * C : nullptr;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  u32 valid = 0;
  int bad = 0;
  auto bump = [&](u32 idx) {
    if constexpr (SMEM) atomicAdd(my + idx, 1u); else atomicAdd(hist + idx, 1ULL);
  };
  const long long stride = (long long)gridDim.x * kClassThreads;
  for (long long i = (long long)blockIdx.x * kClassThreads + tid; i < P.n; i += stride) {
    const long long tl = as_i64(target[i]);
    if constexpr (HAS_IGNORE)
      if (tl == P.ignore) continue;
    if ((u64)tl >= (u64)P.C) {
      bad |= TMB_STATUS_LABEL_RANGE | TMB_STATUS_TARGET_VALUES;
      continue;
    }
    ++valid;
    bump(2 * C + (u32)tl);
    const PT* row = preds + i * k;
    for (int j = 0; j < k; ++j) {
      const long long pl = as_i64(row[j]);
      if ((u64)pl >= (u64)P.C) {  // cannot come out of a top-k over C scores
        bad |= TMB_STATUS_PREDS_VALUES;
        continue;
      }
      bump(C + (u32)pl);
      if (pl == tl) bump((u32)pl);
    }
  }
  finish<SMEM>(P, s_bins, &s_valid, &s_last, valid, bad);
}

// Score inputs with top_k == 1 (the usual "logits in, accuracy out" call): the reference's `_format` runs
// `preds.argmax(dim=1)` (stat_scores.py:333-334) and the update re-reads the labels.  Here one warp takes one row of C
// scores, finds the first maximum (torch.argmax's rule: the lowest index among equal maxima, NaN counts as the
// maximum) and lane 0 bumps the three histograms — the (N,) label tensor is never written.
template <typename CT>
__device__ __forceinline__ bool beats(CT v, u32 i, CT bv, u32 bi) {
  const bool vn = v != v, bn = bv != bv;
  if (bn) return vn && i < bi;
  if (vn) return true;
  return v > bv || (v == bv && i < bi);
}

// lane-local scan visits increasing indices, so "strictly better" is enough there
template <typename CT>
__device__ __forceinline__ bool beats_later(CT v, CT bv) {
  return v > bv || (v != v && bv == bv);
}

template <typename ST, typename TT, bool HAS_IGNORE, bool SMEM>
__global__ void __launch_bounds__(kClassThreads) class_stats_argmax_kernel(const ClassParams P) {
  using CT = typename CmpType<ST>::type;
  extern __shared__ u32 s_bins[];
  __shared__ int s_last;
  __shared__ u32 s_valid;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const u32 C = (u32)P.C;
  u64* const hist = reinterpret_cast<u64*>(P.ws + 1);
  if constexpr (SMEM)
    for (u32 i = tid; i < (u32)P.copies * 3u * C; i += kClassThreads) s_bins[i] = 0;
  if (tid == 0) s_valid = 0;
  __syncthreads();
  u32* const my = SMEM ? s_bins + (u32)(warp % P.copies) * 3u * C : nullptr;
  const ST* __restrict__ scores = static_cast<const ST*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  u32 valid = 0;
  int bad = 0;
  auto bump = [&](u32 idx) {
    if constexpr (SMEM) atomicAdd(my + idx, 1u); else atomicAdd(hist + idx, 1ULL);
  };
  // g lanes share a row (a power of two, ~8 elements per lane), so a
  // warp works on 32 / g rows at once: short rows do not leave lanes idle
  const u32 g = (u32)P.group, sub = (u32)lane & (g - 1u), rpw = 32u / g;
  const bool vec = (C & 3u) == 0;  // rows start on a 4-element boundary: vector loads
  const long long wstride = (long long)gridDim.x * (kClassThreads / 32) * rpw;
  for (long long rbase = ((long long)blockIdx.x * (kClassThreads / 32) + warp) * rpw; rbase < P.n; rbase += wstride) {
    const long long row = rbase + (u32)lane / g;
    bool active = row < P.n;
    long long tl = 0;
    if (active) {
      tl = as_i64(target[row]);
      if constexpr (HAS_IGNORE) active = tl != P.ignore;
    }
    CT bv = (CT)0;
    u32 bi = 0;
    if (active) {
      const ST* r = scores + row * (long long)C;
      bv = to_cmp(r[0]);  // every lane starts from element 0: a lane without elements of its own holds a valid pair
      if (vec) {
        for (u32 j = sub * 4; j < C; j += g * 4) {
          const Vec4<ST> v = load4(r + j);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const CT x = to_cmp(v.v[q]);
            if (beats_later(x, bv)) { bv = x; bi = j + q; }
          }
        }
      } else {
        for (u32 j = sub; j < C; j += g) {
          const CT x = to_cmp(r[j]);
          if (beats_later(x, bv)) { bv = x; bi = j; }
        }
      }
    }
    for (u32 o = g >> 1; o; o >>= 1) {  // stays inside the aligned group of g lanes
      const CT ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const u32 oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (beats(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (active && sub == 0) {
      if ((u64)tl >= (u64)P.C) {  // pred is in range by construction: the key arithmetic of the label kernel
        const long long k = tl * P.C + (long long)bi;
        bad |= TMB_STATUS_TARGET_VALUES;
        if ((u64)k >= (u64)P.CC) {
          bad |= TMB_STATUS_LABEL_RANGE;
        } else {
          const u32 rr = (u32)(k / P.C), cc = (u32)(k - (long long)rr * P.C);
          ++valid;
          bump(C + cc);
          bump(2 * C + rr);
          if (rr == cc) bump(cc);
        }
      } else {
        ++valid;
        bump(C + bi);
        bump(2 * C + (u32)tl);
        if ((u32)tl == bi) bump(bi);
      }
    }
  }
  finish<SMEM>(P, s_bins, &s_valid, &s_last, valid, bad);
}

// The same row scan with the class ids written out (int32): the argmax of `_multiclass_confusion_matrix_format`
// (confusion_matrix.py:318-319) for the callers that need the labels themselves (K1's C x C matrix).
template <typename ST>
__global__ void __launch_bounds__(kClassThreads) rows_argmax_kernel(const ST* __restrict__ scores, long long n, u32 C,
                                                                    u32 g, int* __restrict__ out) {
  using CT = typename CmpType<ST>::type;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const u32 sub = (u32)lane & (g - 1u), rpw = 32u / g;
  const bool vec = (C & 3u) == 0;
  const long long wstride = (long long)gridDim.x * (kClassThreads / 32) * rpw;
  for (long long rbase = ((long long)blockIdx.x * (kClassThreads / 32) + warp) * rpw; rbase < n; rbase += wstride) {
    const long long row = rbase + (u32)lane / g;
    const bool active = row < n;
    CT bv = (CT)0;
    u32 bi = 0;
    if (active) {
      const ST* r = scores + row * (long long)C;
      bv = to_cmp(r[0]);
      if (vec) {
        for (u32 j = sub * 4; j < C; j += g * 4) {
          const Vec4<ST> v = load4(r + j);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const CT x = to_cmp(v.v[q]);
            if (beats_later(x, bv)) { bv = x; bi = j + q; }
          }
        }
      } else {
        for (u32 j = sub; j < C; j += g) {
          const CT x = to_cmp(r[j]);
          if (beats_later(x, bv)) { bv = x; bi = j; }
        }
      }
    }
    for (u32 o = g >> 1; o; o >>= 1) {
      const CT ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const u32 oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (beats(ov, oi, bv, bi)) { bv = ov; bi = oi; }
    }
    if (active && sub == 0) out[row] = (int)bi;
  }
}

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_class_stats_workspace_bytes(int64_t num_classes) {
  if (num_classes < 1) return 0;
  return sizeof(ClassStatsWs) + (size_t)num_classes * 3 * sizeof(u64);
}

static int class_stats_launch(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n,
                              int64_t num_classes, int64_t top_k, int has_ignore, int64_t ignore_index, int64_t* out,
                              void* workspace, size_t workspace_bytes, int32_t* status, void* stream) {
  if (top_k < 1 || top_k > num_classes) return TMB_ERR_ARG;
  if (n < 0 || num_classes < 1 || num_classes > (1LL << 31) - 1 || !out || !workspace || !status) return TMB_ERR_ARG;
  if (workspace_bytes < tmb_class_stats_workspace_bytes(num_classes)) return TMB_ERR_ARG;
  if (n == 0) return TMB_OK;
  if (!preds || !target) return TMB_ERR_ARG;
  if (n >= (1LL << 32)) return TMB_ERR_TOO_LARGE;  // u32 per-CTA counters
  if (!aligned32(preds) || !aligned32(target) || !aligned32(workspace)) return TMB_ERR_ALIGN;
  ClassParams P{};
  P.preds = preds;
  P.target = target;
  P.n = n;
  P.C = num_classes;
  P.CC = (long long)num_classes * num_classes;
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  P.out = (u64*)out;
  P.ws = (ClassStatsWs*)workspace;
  P.status = status;
  P.top_k = (int)top_k;
  const size_t per_copy = (size_t)num_classes * 3 * 4;
  size_t copies = per_copy <= kClassSmemBudget ? kClassSmemBudget / per_copy : 0;
  if (copies > kClassThreads / 32) copies = kClassThreads / 32;
  P.copies = (int)copies;
  const size_t smem = copies * per_copy;
  long long want = ((top_k > 1 ? n : (n >> 2)) + kClassThreads - 1) / kClassThreads;
  const long long cap = (long long)sm_count() * kClassCtasPerSm;
  const int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
  cudaStream_t st = (cudaStream_t)stream;
  TMB_DISPATCH_INT(preds_dtype, PT, {
    TMB_DISPATCH_INT(target_dtype, TT, {
      if (top_k > 1) {
        if (copies) {
          if (P.has_ignore) class_stats_topk_kernel<PT, TT, true, true><<<grid, kClassThreads, smem, st>>>(P);
          else class_stats_topk_kernel<PT, TT, false, true><<<grid, kClassThreads, smem, st>>>(P);
        } else {
          if (P.has_ignore) class_stats_topk_kernel<PT, TT, true, false><<<grid, kClassThreads, 0, st>>>(P);
          else class_stats_topk_kernel<PT, TT, false, false><<<grid, kClassThreads, 0, st>>>(P);
        }
      } else if (copies) {
        if (P.has_ignore) class_stats_kernel<PT, TT, true, true><<<grid, kClassThreads, smem, st>>>(P);
        else class_stats_kernel<PT, TT, false, true><<<grid, kClassThreads, smem, st>>>(P);
      } else {
        if (P.has_ignore) class_stats_kernel<PT, TT, true, false><<<grid, kClassThreads, 0, st>>>(P);
        else class_stats_kernel<PT, TT, false, false><<<grid, kClassThreads, 0, st>>>(P);
      }
      TMB_CUDA_CHECK(cudaGetLastError());
      return TMB_OK;
    })
  })
  return TMB_ERR_ARG;
}

extern "C" int tmb_class_stats_update(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                      int64_t n, int64_t num_classes, int has_ignore, int64_t ignore_index,
                                      int64_t* out, void* workspace, size_t workspace_bytes, int32_t* status,
                                      void* stream) {
  return class_stats_launch(preds, preds_dtype, target, target_dtype, n, num_classes, 1, has_ignore, ignore_index, out,
                            workspace, workspace_bytes, status, stream);
}

extern "C" int tmb_class_stats_topk_update(const void* topk, int topk_dtype, const void* target, int target_dtype,
                                           int64_t n, int64_t num_classes, int64_t top_k, int has_ignore,
                                           int64_t ignore_index, int64_t* out, void* workspace, size_t workspace_bytes,
                                           int32_t* status, void* stream) {
  return class_stats_launch(topk, topk_dtype, target, target_dtype, n, num_classes, top_k, has_ignore, ignore_index,
                            out, workspace, workspace_bytes, status, stream);
}


extern "C" int tmb_class_stats_argmax_update(const void* scores, int scores_dtype, const void* target, int target_dtype,
                                             int64_t n, int64_t num_classes, int has_ignore, int64_t ignore_index,
                                             int64_t* out, void* workspace, size_t workspace_bytes, int32_t* status,
                                             void* stream) {
  if (n < 0 || num_classes < 1 || num_classes > (1LL << 31) - 1 || !out || !workspace || !status) return TMB_ERR_ARG;
  if (workspace_bytes < tmb_class_stats_workspace_bytes(num_classes)) return TMB_ERR_ARG;
  if (n == 0) return TMB_OK;
  if (!scores || !target) return TMB_ERR_ARG;
  if (n >= (1LL << 32)) return TMB_ERR_TOO_LARGE;
  if (!aligned32(scores) || !aligned32(target) || !aligned32(workspace)) return TMB_ERR_ALIGN;
  ClassParams P{};
  P.preds = scores;
  P.target = target;
  P.n = n;
  P.C = num_classes;
  P.CC = (long long)num_classes * num_classes;
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  P.out = (u64*)out;
  P.ws = (ClassStatsWs*)workspace;
  P.status = status;
  P.top_k = 1;
  const size_t per_copy = (size_t)num_classes * 3 * 4;
  size_t copies = per_copy <= kClassSmemBudget ? kClassSmemBudget / per_copy : 0;
  if (copies > kClassThreads / 32) copies = kClassThreads / 32;
  P.copies = (int)copies;
  const size_t smem = copies * per_copy;
  int group = 1;
  while (group < 32 && (long long)group * 8 < num_classes) group <<= 1;
  P.group = group;
  const long long rows_per_cta = (long long)(kClassThreads / 32) * (32 / group);
  const long long want = (n + rows_per_cta - 1) / rows_per_cta;
  const long long cap = (long long)sm_count() * kClassArgmaxCtasPerSm;
  const int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
  cudaStream_t st = (cudaStream_t)stream;
  TMB_DISPATCH_FLOAT(scores_dtype, ST, {
    TMB_DISPATCH_INT(target_dtype, TT, {
      if (copies) {
        if (P.has_ignore) class_stats_argmax_kernel<ST, TT, true, true><<<grid, kClassThreads, smem, st>>>(P);
        else class_stats_argmax_kernel<ST, TT, false, true><<<grid, kClassThreads, smem, st>>>(P);
      } else {
        if (P.has_ignore) class_stats_argmax_kernel<ST, TT, true, false><<<grid, kClassThreads, 0, st>>>(P);
        else class_stats_argmax_kernel<ST, TT, false, false><<<grid, kClassThreads, 0, st>>>(P);
      }
      TMB_CUDA_CHECK(cudaGetLastError());
      return TMB_OK;
    })
  })
  return TMB_ERR_ARG;
}


extern "C" int tmb_rows_argmax(const void* scores, int scores_dtype, int64_t n, int64_t num_classes, int32_t* out,
                               void* stream) {
  if (n < 0 || num_classes < 1 || num_classes > (1LL << 31) - 1 || !out) return TMB_ERR_ARG;
  if (n == 0) return TMB_OK;
  if (!scores) return TMB_ERR_ARG;
  if (!aligned32(scores)) return TMB_ERR_ALIGN;
  int group = 1;
  while (group < 32 && (long long)group * 8 < num_classes) group <<= 1;
  const long long rows_per_cta = (long long)(kClassThreads / 32) * (32 / group);
  const long long want = (n + rows_per_cta - 1) / rows_per_cta;
  const long long cap = (long long)sm_count() * kClassArgmaxCtasPerSm;
  const int grid = (int)(want < cap ? want : cap);
  cudaStream_t st = (cudaStream_t)stream;
  TMB_DISPATCH_FLOAT(scores_dtype, ST, {
    rows_argmax_kernel<ST><<<grid, kClassThreads, 0, st>>>((const ST*)scores, n, (u32)num_classes, (u32)group, out);
    TMB_CUDA_CHECK(cudaGetLastError());
    return TMB_OK;
  })
  return TMB_ERR_ARG;
}
