// K2 — tp / fp / tn / fn per label in one pass (contract in include/tm_b200.h).
//
// The reference runs ~20 elementwise / reduction kernels over bool and int64 temporaries
// (stat_scores.py:90-131, :647-681).  Here each (pred, target) element is read once with 128/256-bit loads.
// Threads are mapped so that a thread always sees the same 4 labels (its position in a row is invariant under
// the grid stride), so the four statistics of each label live in registers for the whole kernel; they are
// merged through shared memory once per CTA and through one 64-bit RED per (label, statistic) per CTA.
//
// The reference's all-or-nothing sigmoid is handled as in K3: pass 1 thresholds the raw scores and raises a
// flag when a score lies outside [0,1]; pass 2 always launches, exits unless the flag is up, and then
// recounts with sigmoid(p).  Partial counts of an abandoned pass 1 are dropped by the finalising CTA.
#include "common.cuh"

namespace tmb {

struct StatWs {
  int flag_oor;
  int ticket;
  int pad[6];
  // followed by u64 acc[4 * L]  ([tp|fp|tn|fn][label])
};
static_assert(sizeof(StatWs) == 32, "workspace header");

constexpr int kStatThreads = 512;

struct StatParams {
  const void* preds;
  const void* target;
  long long n_rows;
  long long L;
  double threshold;
  int has_ignore;
  long long ignore;
  int check_range;
  int only_if_flag;
  int range_check_ignored;  // 1: ignored positions take part in the [0,1] check (stat-scores _format order)
  int* status;              // fused validation bits (may be null)
  u64* out;  // [4][L]
  StatWs* ws;
};

template <typename PT>
struct IsFloatScore {
  static constexpr bool value = false;
};
template <> struct IsFloatScore<float> { static constexpr bool value = true; };
template <> struct IsFloatScore<double> { static constexpr bool value = true; };
template <> struct IsFloatScore<__half> { static constexpr bool value = true; };
template <> struct IsFloatScore<__nv_bfloat16> { static constexpr bool value = true; };

// the comparison `preds > threshold` happens in the preds dtype (a Python scalar does not promote a tensor)
template <typename PT>
__device__ __forceinline__ typename CmpType<PT>::type threshold_as(double thr) {
  if constexpr (!IsFloatScore<PT>::value) return 0.f;
  else if constexpr (sizeof(PT) == 8) return thr;
  else if constexpr (sizeof(PT) == 4) return (float)thr;
  else return to_cmp(PT((float)thr));
}

struct Counts {
  u32 tp, fp, tn, fn;
};
// per-thread bookkeeping of the fused validation
struct Seen {
  u32 elements;  // elements handed to tally()
  u32 ignored;   // of which equal to ignore_index
  u64 preds_or;  // OR of the label preds (integer preds only)
};

template <typename PT, typename TT, bool SIGMOID>
__device__ __forceinline__ void tally(PT pv, TT tv, const StatParams& P, typename CmpType<PT>::type thr, Counts& c,
                                      int& oor, Seen& seen) {
  long long t = as_i64(tv);
  // Fused validation costs nothing per element on the common path: a target that is neither 0, 1 nor the ignore
  // value lands in none of the four counters, so "counted + ignored != elements seen" at the end of the thread
  // says one was there.
  if constexpr (IsFloatScore<PT>::value) {
    using CT = typename CmpType<PT>::type;
    CT p = to_cmp(pv);
    // stat-scores _format range-checks the whole preds tensor, ignored positions included (stat_scores.py:104,
    // :657); the binary confusion-matrix _format drops ignored samples first (confusion_matrix.py:133-142)
    const bool ign = P.has_ignore && t == P.ignore;
    if (P.check_range && (P.range_check_ignored || !ign)) oor |= !(p >= (CT)0 && p <= (CT)1);
    if (ign) {
      ++seen.ignored;
      return;
    }
    if constexpr (SIGMOID) p = round_to_storage<PT>(sigmoid_ref(p));
    const bool pb = p > thr;
    c.tp += (t == 1) & pb;
    c.fn += (t == 1) & !pb;
    c.fp += (t == 0) & pb;
    c.tn += (t == 0) & !pb;
  } else {
    long long p = as_i64(pv);
    seen.preds_or |= (u64)p;  // label preds outside {0, 1} leave a bit above bit 0
    if (P.has_ignore && t == P.ignore) {
      ++seen.ignored;
      return;
    }
    const bool eq = (t == p);
    c.tp += (t == 1) & eq;
    c.fn += (t == 1) & !eq;
    c.fp += (t == 0) & !eq;
    c.tn += (t == 0) & eq;
  }
}

#ifndef TMB_K2_TILED_PER_SM
#define TMB_K2_TILED_PER_SM 3
#endif
// resident CTAs per SM: four for the flat (L == 1) form, which needs 32 registers - binary stat scores went from
// 64.7 / 86.6 / 94.4 % of the measured copy peak at 2^24 / 2^26 / 2^28 samples with two CTAs per SM to 70.6 / 95.7 /
// 105.6 % (profiles/r2_k2flat_variants.log); the tiled form, whose per-label counters live in registers, runs three
// (40 registers and 40 bytes of spills instead of 60 registers: L = 100 on 2^23 rows 1678 -> 1627 us, 91.5 -> 94.4 %,
// profiles/r2_k2tiled_variants.log).
constexpr int kStatFlatCtasPerSm = 4;
constexpr int kStatTiledCtasPerSm = TMB_K2_TILED_PER_SM;
// V = 4: L % 4 == 0, a thread owns 4 consecutive labels.  V = 1: any L, a thread owns one label.
// FLAT (L == 1): every element belongs to label 0 and is read 4 at a time.
template <typename PT, typename TT, bool SIGMOID, int V, bool FLAT>
__global__ void __launch_bounds__(kStatThreads, FLAT ? kStatFlatCtasPerSm : kStatTiledCtasPerSm)
stat_scores_kernel(const StatParams P) {
  using CT = typename CmpType<PT>::type;
  StatWs* ws = P.ws;
  if (P.only_if_flag && ws->flag_oor == 0) return;
  extern __shared__ u32 s_cnt[];  // [4][CW*V] for this CTA's column tile
  __shared__ int s_last;
  const int NT = blockDim.x, tid = threadIdx.x;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const CT thr = threshold_as<PT>(P.threshold);
  const long long L = P.L;
  u64* acc = reinterpret_cast<u64*>(ws + 1);
  int oor = 0;
  Seen seen{0, 0, 0};
  Counts c[V];
#pragma unroll
  for (int j = 0; j < V; ++j) c[j] = Counts{0, 0, 0, 0};

  const long long colgroups = FLAT ? 1 : L / V;
  const int CW = (int)(colgroups < NT ? colgroups : NT);  // column groups per CTA tile
  const int RP = FLAT ? NT : NT / CW;                     // rows per pass
  const int ncols_smem = FLAT ? 1 : CW * V;
  for (int i = tid; i < 4 * ncols_smem; i += NT) s_cnt[i] = 0;
  __syncthreads();

  long long cg = 0;
  bool active;
  if constexpr (FLAT) {
    active = true;
    const long long n = P.n_rows;
    const long long nvec = n >> 2;
    const long long stride = (long long)gridDim.x * NT;
    long long it = 0;
    for (long long g = (long long)blockIdx.x * NT + tid; g < nvec; g += stride, ++it) {
      Vec4<PT> p = load4(preds + 4 * g);
      Vec4<TT> t = load4(target + 4 * g);
#pragma unroll
      for (int j = 0; j < 4; ++j) tally<PT, TT, SIGMOID>(p.v[j], t.v[j], P, thr, c[0], oor, seen);
      if (P.check_range && (it & 63) == 63) {
        if (oor) atomicOr(&ws->flag_oor, 1);
        if (*(volatile int*)&ws->flag_oor) {
          ++it;
          break;
        }
      }
    }
    seen.elements = (u32)it * 4;
    if (blockIdx.x == 0 && tid < (n & 3)) {
      long long i = (nvec << 2) + tid;
      tally<PT, TT, SIGMOID>(preds[i], target[i], P, thr, c[0], oor, seen);
      seen.elements += 1;
    }
  } else {
    const int r0 = tid / CW;
    cg = (long long)blockIdx.y * CW + (tid - r0 * CW);
    active = r0 < RP && cg < colgroups;
    if (active) {
      const long long rstride = (long long)gridDim.x * RP;
      long long it = 0;
      for (long long r = (long long)blockIdx.x * RP + r0; r < P.n_rows; r += rstride, ++it) {
        const long long e = r * L + cg * V;
        if constexpr (V == 4) {
          Vec4<PT> p = load4(preds + e);
          Vec4<TT> t = load4(target + e);
#pragma unroll
          for (int j = 0; j < 4; ++j) tally<PT, TT, SIGMOID>(p.v[j], t.v[j], P, thr, c[j], oor, seen);
        } else {
          tally<PT, TT, SIGMOID>(preds[e], target[e], P, thr, c[0], oor, seen);
        }
        if (P.check_range && (it & 63) == 63) {
          if (oor) atomicOr(&ws->flag_oor, 1);
          if (*(volatile int*)&ws->flag_oor) {
            ++it;
            break;
          }
        }
      }
      seen.elements = (u32)it * V;
    }
  }
  if (P.check_range && oor) atomicOr(&ws->flag_oor, 1);
  if (P.status) {
    u32 counted = seen.ignored;
#pragma unroll
    for (int j = 0; j < V; ++j) counted += c[j].tp + c[j].fp + c[j].tn + c[j].fn;
    const int vbits = (counted != seen.elements ? TMB_STATUS_TARGET_VALUES : 0) |
                      (seen.preds_or > 1ULL ? TMB_STATUS_PREDS_VALUES : 0);
    if (vbits) report_status(P.status, vbits);
  }

  // merge: registers -> shared (per CTA) -> global scratch
  if (active) {
    const int lc = FLAT ? 0 : (int)(cg - (long long)blockIdx.y * CW) * V;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      if (c[j].tp) atomicAdd(&s_cnt[0 * ncols_smem + lc + j], c[j].tp);
      if (c[j].fp) atomicAdd(&s_cnt[1 * ncols_smem + lc + j], c[j].fp);
      if (c[j].tn) atomicAdd(&s_cnt[2 * ncols_smem + lc + j], c[j].tn);
      if (c[j].fn) atomicAdd(&s_cnt[3 * ncols_smem + lc + j], c[j].fn);
    }
  }
  __syncthreads();
  const long long col0 = FLAT ? 0 : (long long)blockIdx.y * CW * V;
  for (int i = tid; i < 4 * ncols_smem; i += NT) {
    const int s = i / ncols_smem, lc = i - s * ncols_smem;
    u32 v = s_cnt[i];
    if (v && col0 + lc < L) atomicAdd(acc + s * L + col0 + lc, (u64)v);
  }
  __threadfence();
  __syncthreads();
  const int nblocks = gridDim.x * gridDim.y;
  if (tid == 0) s_last = (atomicAdd(&ws->ticket, 1) == nblocks - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int flag = *(volatile int*)&ws->flag_oor;
  const bool discard = P.check_range && flag != 0;
  for (long long i = tid; i < 4 * L; i += NT) {
    u64 v = __ldcg(acc + i);
    acc[i] = 0;
    if (!discard && v) P.out[i] += v;
  }
  if (tid == 0) {
    ws->ticket = 0;
    if (!P.check_range && P.only_if_flag) ws->flag_oor = 0;
  }
}

template <typename PT, typename TT, bool SIGMOID>
static int launch_one(const StatParams& P, cudaStream_t st) {
  const long long L = P.L;
  if (L == 1) {
    long long nvec = P.n_rows >> 2;
    long long want = (nvec + kStatThreads - 1) / kStatThreads;
    long long cap = (long long)sm_count() * kStatFlatCtasPerSm;
    int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
    stat_scores_kernel<PT, TT, SIGMOID, 1, true><<<grid, kStatThreads, 4 * sizeof(u32), st>>>(P);
  } else {
    const int V = (L % 4 == 0) ? 4 : 1;
    const long long colgroups = L / V;
    const int CW = (int)(colgroups < kStatThreads ? colgroups : kStatThreads);
    const int RP = kStatThreads / CW;
    const long long gy = (colgroups + CW - 1) / CW;
    if (gy > 65535) return TMB_ERR_TOO_LARGE;
    long long want = (P.n_rows + RP - 1) / RP;
    long long cap = ((long long)sm_count() * kStatTiledCtasPerSm + gy - 1) / gy;
    if (cap < 1) cap = 1;
    int gx = (int)(want < 1 ? 1 : want < cap ? want : cap);
    const size_t smem = (size_t)4 * CW * V * sizeof(u32);
    dim3 grid(gx, (unsigned)gy);
    if (V == 4) stat_scores_kernel<PT, TT, SIGMOID, 4, false><<<grid, kStatThreads, smem, st>>>(P);
    else stat_scores_kernel<PT, TT, SIGMOID, 1, false><<<grid, kStatThreads, smem, st>>>(P);
  }
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}

template <typename PT, typename TT>
static int launch_stat(StatParams P, int prob_mode, cudaStream_t st) {
  if constexpr (!IsFloatScore<PT>::value) {
    P.check_range = 0; P.only_if_flag = 0;
    return launch_one<PT, TT, false>(P, st);
  } else {
    int rc;
    switch (prob_mode) {
      case TMB_PROB_AS_IS:
        P.check_range = 0; P.only_if_flag = 0;
        return launch_one<PT, TT, false>(P, st);
      case TMB_PROB_SIGMOID:
        P.check_range = 0; P.only_if_flag = 0;
        return launch_one<PT, TT, true>(P, st);
      case TMB_PROB_AUTO:
        P.check_range = 1; P.only_if_flag = 0;
        rc = launch_one<PT, TT, false>(P, st);
        if (rc != TMB_OK) return rc;
        P.check_range = 0; P.only_if_flag = 1;
        return launch_one<PT, TT, true>(P, st);
      default:
        return TMB_ERR_ARG;
    }
  }
}


// ---------------------------------------------------------------------------------------------------------------
// Row-wise statistics (multidim_average="samplewise"; the legacy `_stat_scores` with reduce="samples"): the four
// counts of every row of an (n_rows, width) pair of tensors.  The reference builds bool / int64 temporaries of the
// input size and calls four `sum(dim)` reductions (stat_scores.py:127-130, :677-680, :891-906); here a row is read
// once by a group of 1..32 lanes (the group size follows the row width, so short rows do not idle a warp and
// long rows are read with coalesced 16/32-byte loads), reduced with shuffles and written as four int64.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kRowsThreads = 256;

template <typename PT, typename TT, bool SIGMOID, int V>
__global__ void __launch_bounds__(kRowsThreads) stat_scores_rows_kernel(const StatParams P, long long width,
                                                                         int lpr_log2) {
  using CT = typename CmpType<PT>::type;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const CT thr = threshold_as<PT>(P.threshold);
  const int lane = threadIdx.x & 31;
  const int lpr = 1 << lpr_log2;
  const int sub = lane & (lpr - 1);
  const long long rows_per_warp = 32 >> lpr_log2;
  const long long warp0 = ((long long)blockIdx.x * kRowsThreads + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * kRowsThreads) >> 5;
  long long* out = reinterpret_cast<long long*>(P.out);
  for (long long row0 = warp0 * rows_per_warp; row0 < P.n_rows; row0 += nwarps * rows_per_warp) {
    const long long row = row0 + (lane >> lpr_log2);
    Counts c{0, 0, 0, 0};
    int oor = 0;
    Seen seen{0, 0, 0};
    if (row < P.n_rows) {
      const long long base = row * width;
      if constexpr (V == 4) {
        for (long long j = (long long)sub * 4; j < width; j += (long long)lpr * 4) {
          const Vec4<PT> pv = load4(preds + base + j);
          const Vec4<TT> tv = load4(target + base + j);
#pragma unroll
          for (int q = 0; q < 4; ++q) tally<PT, TT, SIGMOID>(pv.v[q], tv.v[q], P, thr, c, oor, seen);
        }
      } else {
        for (long long j = sub; j < width; j += lpr) tally<PT, TT, SIGMOID>(preds[base + j], target[base + j], P, thr, c, oor, seen);
      }
    }
    for (int o = lpr >> 1; o; o >>= 1) {
      c.tp += __shfl_xor_sync(0xffffffffu, c.tp, o);
      c.fp += __shfl_xor_sync(0xffffffffu, c.fp, o);
      c.tn += __shfl_xor_sync(0xffffffffu, c.tn, o);
      c.fn += __shfl_xor_sync(0xffffffffu, c.fn, o);
    }
    if (row < P.n_rows && sub == 0) {
      out[row] = c.tp;
      out[P.n_rows + row] = c.fp;
      out[2 * P.n_rows + row] = c.tn;
      out[3 * P.n_rows + row] = c.fn;
    }
  }
}

template <typename PT, typename TT>
static int launch_rows(StatParams P, long long width, int prob_mode, cudaStream_t st) {
  P.check_range = 0;
  P.only_if_flag = 0;
  const bool v4 = (width % 4) == 0;
  const long long per = v4 ? 8 : 2;  // elements a lane should have to itself before another lane joins the row
  int lpr_log2 = 0;
  while (lpr_log2 < 5 && ((long long)per << lpr_log2) < width) ++lpr_log2;
  const long long rows_per_warp = 32 >> lpr_log2;
  const long long warps = (P.n_rows + rows_per_warp - 1) / rows_per_warp;
  const long long want = (warps + kRowsThreads / 32 - 1) / (kRowsThreads / 32);
  const long long cap = (long long)sm_count() * 16;
  const int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
  bool sigmoid = false;
  if constexpr (IsFloatScore<PT>::value) {
    if (prob_mode == TMB_PROB_SIGMOID) sigmoid = true;
    else if (prob_mode != TMB_PROB_AS_IS) return TMB_ERR_ARG;
  }
  if constexpr (IsFloatScore<PT>::value) {
    if (sigmoid) {
      if (v4) stat_scores_rows_kernel<PT, TT, true, 4><<<grid, kRowsThreads, 0, st>>>(P, width, lpr_log2);
      else stat_scores_rows_kernel<PT, TT, true, 1><<<grid, kRowsThreads, 0, st>>>(P, width, lpr_log2);
      TMB_CUDA_CHECK(cudaGetLastError());
      return TMB_OK;
    }
  }
  if (v4) stat_scores_rows_kernel<PT, TT, false, 4><<<grid, kRowsThreads, 0, st>>>(P, width, lpr_log2);
  else stat_scores_rows_kernel<PT, TT, false, 1><<<grid, kRowsThreads, 0, st>>>(P, width, lpr_log2);
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Row-wise multiclass statistics: per sample n and class c, tp / fp / tn / fn of the (X,) label vectors of that sample.
// The reference one-hots both tensors to (N, X, C) and takes four masked sums (stat_scores.py:363-393); here a warp
// keeps three length-C histograms of one sample in shared memory (hits of c, predictions of c, targets of c) and
// converts them on the way out.
// ---------------------------------------------------------------------------------------------------------------
struct ClassRowsParams {
  const void* preds;
  const void* target;
  long long n_rows, width, C;
  int top_k;       // class ids per element of `preds` ((n_rows, width, top_k) row-major); 1 = plain labels
  int has_ignore;
  long long ignore;
  long long* out;  // [4][n_rows][C]
  int* status;
};

template <typename PT, typename TT>
__global__ void __launch_bounds__(kRowsThreads) class_stats_rows_kernel(const ClassRowsParams P, int warps_per_cta) {
  extern __shared__ u32 s_hist[];  // [warps_per_cta][3][C]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp >= warps_per_cta) return;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long C = P.C;
  u32* hit = s_hist + (size_t)warp * 3 * C;
  u32* pred_of = hit + C;
  u32* targ_of = pred_of + C;
  int bad = 0;
  for (long long row = (long long)blockIdx.x * warps_per_cta + warp; row < P.n_rows;
       row += (long long)gridDim.x * warps_per_cta) {
    for (long long c = lane; c < 3 * C; c += 32) hit[c] = 0;
    __syncwarp();
    const long long base = row * P.width;
    u32 valid = 0;
    for (long long j = lane; j < P.width; j += 32) {
      const long long t = as_i64(target[base + j]);
      if (P.has_ignore && t == P.ignore) continue;
      if ((u64)t >= (u64)C) {
        bad = 1;  // the reference's one_hot raises on such a label
        continue;
      }
      // top_k class ids per element (the k highest scores; one id for label inputs): a class among them that is not
      // the target is a false positive of that class, the target is a hit if it is among them and a miss otherwise
      const PT* ids = preds + (base + j) * P.top_k;
      bool in_range = true;
      for (int q = 0; q < P.top_k; ++q) in_range &= (u64)as_i64(ids[q]) < (u64)C;
      if (!in_range) {
        bad = 1;
        continue;
      }
      ++valid;
      bool hit_target = false;
      for (int q = 0; q < P.top_k; ++q) {
        const long long p = as_i64(ids[q]);
        if (p == t) hit_target = true;
        else atomicAdd(&pred_of[p], 1u);
      }
      atomicAdd(hit_target ? &hit[t] : &targ_of[t], 1u);
    }
    valid = __reduce_add_sync(0xffffffffu, valid);
    __syncwarp();
    long long* o = P.out + row * C;
    const long long plane = P.n_rows * C;
    for (long long c = lane; c < C; c += 32) {
      const long long tp = hit[c], fp = pred_of[c], fn = targ_of[c];
      o[c] = tp;
      o[plane + c] = fp;
      o[2 * plane + c] = (long long)valid - tp - fp - fn;
      o[3 * plane + c] = fn;
    }
    __syncwarp();
  }
  if (bad) report_status(P.status, TMB_STATUS_LABEL_RANGE);
}

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_stat_scores_workspace_bytes(int64_t num_labels) {
  if (num_labels < 1) return 0;
  return sizeof(StatWs) + (size_t)4 * num_labels * sizeof(u64);
}

extern "C" int tmb_stat_scores_update(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                      int64_t n_rows, int64_t num_labels, double threshold, int prob_mode,
                                      int has_ignore, int64_t ignore_index, int range_check_ignored, int64_t* out,
                                      void* workspace, size_t workspace_bytes, int32_t* status, void* stream) {
  if (n_rows < 0 || num_labels < 1 || !out || !workspace) return TMB_ERR_ARG;
  if (workspace_bytes < tmb_stat_scores_workspace_bytes(num_labels)) return TMB_ERR_ARG;
  if (n_rows == 0) return TMB_OK;
  if (!preds || !target) return TMB_ERR_ARG;
  if (!aligned32(preds) || !aligned32(target) || !aligned32(workspace)) return TMB_ERR_ALIGN;
  StatParams P{};
  P.preds = preds;
  P.target = target;
  P.n_rows = n_rows;
  P.L = num_labels;
  P.threshold = threshold;
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  P.range_check_ignored = range_check_ignored ? 1 : 0;
  P.out = (u64*)out;
  P.status = status;
  P.ws = (StatWs*)workspace;
  cudaStream_t st = (cudaStream_t)stream;
  if (is_float_dtype(preds_dtype)) {
    TMB_DISPATCH_FLOAT(preds_dtype, PT, {
      TMB_DISPATCH_INT(target_dtype, TT, { return launch_stat<PT, TT>(P, prob_mode, st); })
    })
  } else {
    TMB_DISPATCH_INT(preds_dtype, PT, {
      TMB_DISPATCH_INT(target_dtype, TT, { return launch_stat<PT, TT>(P, prob_mode, st); })
    })
  }
  return TMB_ERR_ARG;
}

extern "C" int tmb_stat_scores_rows(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                    int64_t n_rows, int64_t width, double threshold, int prob_mode, int has_ignore,
                                    int64_t ignore_index, int64_t* out, void* stream) {
  if (n_rows < 0 || width < 0 || !out) return TMB_ERR_ARG;
  if (n_rows == 0) return TMB_OK;
  if (width >= (1LL << 31)) return TMB_ERR_TOO_LARGE;  // 32-bit counters per row
  cudaStream_t st = (cudaStream_t)stream;
  if (width == 0) {
    TMB_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)4 * n_rows * sizeof(int64_t), st));
    return TMB_OK;
  }
  if (!preds || !target) return TMB_ERR_ARG;
  if (!aligned32(preds) || !aligned32(target)) return TMB_ERR_ALIGN;
  StatParams P{};
  P.preds = preds;
  P.target = target;
  P.n_rows = n_rows;
  P.L = 1;
  P.threshold = threshold;
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  P.out = (u64*)out;
  if (is_float_dtype(preds_dtype)) {
    TMB_DISPATCH_FLOAT(preds_dtype, PT, {
      TMB_DISPATCH_INT(target_dtype, TT, { return launch_rows<PT, TT>(P, width, prob_mode, st); })
    })
  } else {
    TMB_DISPATCH_INT(preds_dtype, PT, {
      TMB_DISPATCH_INT(target_dtype, TT, { return launch_rows<PT, TT>(P, width, prob_mode, st); })
    })
  }
  return TMB_ERR_ARG;
}

static int class_stats_rows_impl(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                 int64_t n_rows, int64_t width, int top_k, int64_t num_classes, int has_ignore,
                                 int64_t ignore_index, int64_t* out, int32_t* status, void* stream) {
  if (n_rows < 0 || width < 0 || num_classes < 1 || top_k < 1 || top_k > num_classes || !out || !status)
    return TMB_ERR_ARG;
  if (n_rows == 0) return TMB_OK;
  if (width >= (1LL << 31)) return TMB_ERR_TOO_LARGE;
  cudaStream_t st = (cudaStream_t)stream;
  if (width == 0) {
    TMB_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)4 * n_rows * num_classes * sizeof(int64_t), st));
    return TMB_OK;
  }
  if (!preds || !target) return TMB_ERR_ARG;
  const size_t per_warp = (size_t)3 * num_classes * sizeof(u32);
  const size_t room = (size_t)max_optin_smem() - 1024;
  if (per_warp > room) {
    set_error_text("tmb_class_stats_rows: three per-sample histograms of num_classes bins do not fit shared memory");
    return TMB_ERR_TOO_LARGE;
  }
  int warps = (int)(room / per_warp);
  if (warps > kRowsThreads / 32) warps = kRowsThreads / 32;
  const size_t smem = per_warp * warps;
  ClassRowsParams P{};
  P.preds = preds;
  P.target = target;
  P.n_rows = n_rows;
  P.width = width;
  P.C = num_classes;
  P.top_k = top_k;
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  P.out = (long long*)out;
  P.status = status;
  const long long want = (n_rows + warps - 1) / warps;
  const long long cap = (long long)sm_count() * 8;
  const int grid = (int)(want < cap ? want : cap);
  TMB_DISPATCH_INT(preds_dtype, PT, {
    TMB_DISPATCH_INT(target_dtype, TT, {
      auto k = class_stats_rows_kernel<PT, TT>;
      TMB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, kRowsThreads, smem, st>>>(P, warps);
      TMB_CUDA_CHECK(cudaGetLastError());
      return TMB_OK;
    })
  })
  return TMB_ERR_ARG;
}

extern "C" int tmb_class_stats_rows(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                    int64_t n_rows, int64_t width, int64_t num_classes, int has_ignore,
                                    int64_t ignore_index, int64_t* out, int32_t* status, void* stream) {
  return class_stats_rows_impl(preds, preds_dtype, target, target_dtype, n_rows, width, 1, num_classes, has_ignore,
                               ignore_index, out, status, stream);
}

extern "C" int tmb_class_stats_rows_topk(const void* topk_ids, int ids_dtype, const void* target, int target_dtype,
                                         int64_t n_rows, int64_t width, int top_k, int64_t num_classes,
                                         int has_ignore, int64_t ignore_index, int64_t* out, int32_t* status,
                                         void* stream) {
  return class_stats_rows_impl(topk_ids, ids_dtype, target, target_dtype, n_rows, width, top_k, num_classes,
                               has_ignore, ignore_index, out, status, stream);
}
