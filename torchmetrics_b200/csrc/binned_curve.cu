// K3 — binned (multi-threshold) curve state update (contract in include/tm_b200.h).
//
// The reference makes 3*T reduction passes over the scores (precision_recall_curve.py:228-250).  Here the
// batch is read ONCE: each score is mapped to bin = #{i : thr[i] <= p} (exactly, against the caller's own
// float32 threshold values), a histogram over (T+1) x {neg,pos} bins is accumulated on chip, and a suffix
// sum over bins in the last CTA turns it into the (T,2,2) state, which is added in place.
//
// Two accumulation strategies, both on shared-memory atomics (measured at >1500 G increments/s chip-wide on B200):
//   * warp histograms (one column of float32 scores — the BinaryAUROC/AP(thresholds=200) case): one u32 copy of
//     the bins per warp, bins found arithmetically for evenly spaced thresholds;
//   * generic (any number of columns, any score dtype): one copy of all columns' bins per CTA in shared memory, or
//     u64 bins in global scratch when they do not fit.
//
// The all-or-nothing sigmoid of the reference's `_format` step (":183-184: if not torch.all((preds >= 0) *
// (preds <= 1)): preds = preds.sigmoid()") is resolved on the device without a host round trip: pass 1
// bins the raw scores and raises a flag on the first score outside [0,1]; pass 2 is always enqueued but
// exits at once unless the flag is up, in which case it redoes the batch with sigmoid(p) and pass 1's
// partial histogram is discarded.
#include <cstdlib>
#include <mutex>
#include <string>
#include <type_traits>

#include "common.cuh"

namespace tmb {

struct BinnedWs {
  int flag_oor;  // pass 1 saw a score outside [0,1]
  int ticket;    // CTAs finished (last one finalises)
  int pad[6];
  // followed by u64 hist[num_columns * 2 * (T+1)], layout [column][pos][bin]
};
static_assert(sizeof(BinnedWs) == 32, "workspace header");

constexpr int kLutCells = 1024;

struct BinnedParams {
  const void* preds;
  const void* target;
  long long n_rows;
  long long num_columns;
  int target_is_class_index;
  const float* thr;
  int T;
  int has_ignore;
  long long ignore;
  int check_range;    // raise ws->flag_oor on scores outside [0,1] and stop early
  int only_if_flag;   // run only when ws->flag_oor is already up (auto pass 2)
  int oor_counts_ignored;  // range check also looks at ignored entries (multilabel _format order)
  u64* confmat;       // (T, C, 2, 2) int64
  BinnedWs* ws;
  int* status;
  int status_bit_on_oor;  // TMB_PROB_CHECK semantics: report instead of redoing
  int hist_copies;        // warp-histogram kernel: copies of the bins per CTA
};

// ---- exact binning ------------------------------------------------------------------------------
// cell() is a monotone map of a score onto kLutCells uniform cells spanning [thr[0], thr[T-1]].  lut[c] is
// the number of thresholds whose cell is < c: all of those are <= any score that falls in cell c, the ones
// in the same cell are compared one by one, the ones in later cells are larger.  Monotonicity of cell() is
// all that exactness needs, so float rounding inside it is harmless.
struct BinMap {
  float lo, scale;
};

__device__ __forceinline__ int cell_of(float p, const BinMap& m) {
  if (!(p >= m.lo)) return 0;  // also NaN
  float x = (p - m.lo) * m.scale;
  int c = (int)x;  // cvt.rzi saturates, NaN -> 0
  return c < kLutCells - 1 ? c : kLutCells - 1;
}
__device__ __forceinline__ int cell_of(double p, const BinMap& m) { return cell_of((float)p, m); }

template <typename CT>
__device__ __forceinline__ int bin_of(CT p, const BinMap& m, const float* __restrict__ s_thr,
                                      const unsigned short* __restrict__ s_lut) {
  int b = s_lut[cell_of(p, m)];
  // s_thr[T] is NaN, so the walk stops there
  while (p >= (CT)s_thr[b]) ++b;
  return b;
}

// copies the thresholds into s_thr (T+1 floats, NaN sentinel) for the calling CTA; returns false if unsorted
__device__ __forceinline__ bool load_thresholds(const float* __restrict__ thr, int T, float* s_thr, BinMap& m) {
  int unsorted = 0;
  for (int i = threadIdx.x; i <= T; i += blockDim.x) {
    float v = i < T ? thr[i] : __int_as_float(0x7fc00000);
    s_thr[i] = v;
    if (i > 0 && i < T && !(thr[i - 1] <= v)) unsorted = 1;
  }
  unsorted = __syncthreads_or(unsorted);
  float lo = s_thr[0], hi = s_thr[T - 1];
  float range = hi - lo;
  m.lo = lo;
  m.scale = (range > 0.f && range < 3.0e38f) ? (float)kLutCells / range : 0.f;
  if (!(m.scale < 3.0e38f)) m.scale = 0.f;
  return !unsorted;
}

// the cell -> first-candidate table of bin_of(); only needed when the thresholds are not evenly spaced
__device__ __forceinline__ void build_lut(int T, const float* s_thr, unsigned short* s_lut, const BinMap& m) {
  for (int c = threadIdx.x; c < kLutCells; c += blockDim.x) {
    // lower_bound over the non-decreasing sequence cell_of(thr[i]) for value c
    int l = 0, r = T;
    while (l < r) {
      int mid = (l + r) >> 1;
      if (cell_of(s_thr[mid], m) < c) l = mid + 1; else r = mid;
    }
    s_lut[c] = (unsigned short)l;
  }
  __syncthreads();
}

__device__ __forceinline__ bool build_binning(const float* __restrict__ thr, int T, float* s_thr,
                                              unsigned short* s_lut, BinMap& m) {
  const bool sorted = load_thresholds(thr, T, s_thr, m);
  build_lut(T, s_thr, s_lut, m);
  return sorted;
}

template <typename CT>
__device__ __forceinline__ bool out_of_unit(CT p) { return !(p >= (CT)0 && p <= (CT)1); }

// last CTA: hist -> suffix sums -> confmat += ; leaves the workspace zeroed.  s_scan: 2*(T+1) u64 of smem.
// Warps 0 and 1 scan the negative / positive histogram of a column, 32 bins per step from the top bin down (a single
// thread walking the 201 bins of T = 200 cost ~3 us of every update).
__device__ void finalize_columns(const BinnedParams& P, u64* hist, u64* s_scan, bool discard) {
  const int NB = P.T + 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (long long c = 0; c < P.num_columns; ++c) {
    u64* h = hist + c * 2 * NB;
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * NB; i += blockDim.x) {
      s_scan[i] = __ldcg(h + i);
      h[i] = 0;
    }
    __syncthreads();
    if (!discard) {
      if (warp < 2) {
        u64* s = s_scan + (warp ? NB : 0);
        u64 carry = 0;
        for (int top = NB - 1; top >= 0; top -= 32) {
          const int b = top - lane;  // lane 0 holds the highest bin of the step
          u64 v = b >= 0 ? s[b] : 0;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const u64 y = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += y;
          }
          v += carry;
          carry = __shfl_sync(0xffffffffu, v, 31);
          if (b >= 0) s[b] = v;
        }
      }
      __syncthreads();
      const u64 totneg = s_scan[0], totpos = s_scan[NB];
      for (int i = threadIdx.x; i < P.T; i += blockDim.x) {
        u64 fp = s_scan[i + 1], tp = s_scan[NB + i + 1];
        u64* o = P.confmat + ((long long)i * P.num_columns + c) * 4;
        o[0] += totneg - fp;  // [0][0] tn
        o[1] += fp;           // [0][1]
        o[2] += totpos - tp;  // [1][0] fn
        o[3] += tp;           // [1][1]
      }
    }
  }
}

// =================================================================================================
// shared pieces of the one-column kernels
// =================================================================================================
enum : int { kModeAsIs = 0, kModeAsIsCheck = 1, kModeSigmoid = 2 };

// Evenly spaced thresholds (what torch.linspace / an int `thresholds` argument produce) need no table walk:
// with j = clamp(round((p - lo) / step)) the stored values satisfy thr[j-1] <= p < thr[j+1], so
// bin = j + (p >= thr[j]) — ONE compare against the caller's own value, still exact.  The spacing is verified
// here, per CTA, against the actual tensor (every thr[i] within step/8 of lo + i*step, computed in double);
// x = (p-lo)*inv_step+0.5 is off by at most T*2^-22 <= 1/64 for T <= 65536, which leaves the needed margin.
struct UniformMap {
  float lo, inv_step;
  unsigned jmax;
  bool ok;
};

__device__ __forceinline__ UniformMap detect_uniform(const float* s_thr, int T) {
  UniformMap u;
  const double lo = s_thr[0], hi = s_thr[T - 1];
  const double step = T >= 2 ? (hi - lo) / (double)(T - 1) : 0.0;
  int bad = !(T >= 2 && T <= 65536 && step > 0.0 && step < 1e30 && fabs(lo) < 1e30);
  if (!bad)
    for (int i = threadIdx.x; i < T; i += blockDim.x) {
      double d = fabs((double)s_thr[i] - (lo + (double)i * step));
      if (!(d <= 0.125 * step)) bad = 1;
    }
  bad = __syncthreads_or(bad);
  u.ok = !bad;
  u.lo = (float)lo;
  u.inv_step = bad ? 0.f : (float)(1.0 / step);
  u.jmax = (unsigned)(T - 1);
  return u;
}

// =================================================================================================
// warp-histogram kernel (one column, float32 scores): shared-memory atomics on per-warp copies of the bins
// =================================================================================================
// Measured on B200 (profiles/r1_red_bench.log): shared-memory atomics sustain >1500 G increments/s chip-wide, four
// times what this path needs.  One u32 copy of the 2*(T+1) bins per WARP keeps contention to same-bin lanes of one
// warp, costs ~1.6 KB per warp at T = 200 and leaves room for 64 resident warps per SM, which is what hides the HBM
// latency (a first version with per-THREAD 8-bit counters in 202 KB of shared memory ran 16 warps per SM and stopped
// at 75 % of the HBM peak; this one reads at 97-106 % of the measured copy bandwidth from 2^26 samples up).
// For large T the warps share fewer copies (copy = warp % copies) so that two CTAs still fit an SM.
constexpr int kWarpHistThreads = 1024;
constexpr size_t kWarpHistSmemBudget = 100 * 1024;

static size_t warphist_fixed_bytes(int T) {
  const int NB2 = 2 * (T + 1);
  return (size_t)NB2 * 8 + (size_t)(T + 1) * 4 + kLutCells * 2 + 16;
}
// number of histogram copies per CTA (<= one per warp); 0 if not even one fits
static int warphist_copies(int T) {
  const size_t per = (size_t)2 * (T + 1) * 4;
  if (warphist_fixed_bytes(T) + per > kWarpHistSmemBudget) return 0;
  size_t c = (kWarpHistSmemBudget - warphist_fixed_bytes(T)) / per;
  return (int)(c > kWarpHistThreads / 32 ? kWarpHistThreads / 32 : c);
}

template <typename TT, int MODE, bool HAS_IGNORE>
__global__ void __launch_bounds__(kWarpHistThreads, 2) binned_warphist_kernel(const BinnedParams P) {
  constexpr bool CHECK = MODE == kModeAsIsCheck;
  constexpr bool SIGMOID = MODE == kModeSigmoid;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BinnedWs* ws = P.ws;
  const int NT = kWarpHistThreads, tid = threadIdx.x, warp = tid >> 5;
  const int T = P.T, NB = T + 1, NB2 = 2 * NB;
  const int copies = P.hist_copies;
  u32* s_hist = reinterpret_cast<u32*>(smem_raw);                       // [copies][NB2]
  u64* s_acc = reinterpret_cast<u64*>(smem_raw + (((size_t)copies * NB2 * 4 + 7) & ~(size_t)7));  // finalize scratch
  float* s_thr = reinterpret_cast<float*>(s_acc + NB2);
  unsigned short* s_lut = reinterpret_cast<unsigned short*>(s_thr + NB);
  __shared__ int s_last;

  for (int i = tid; i < copies * NB2; i += NT) s_hist[i] = 0;
  // Launched with programmatic stream serialisation: the launch and everything above overlap the tail of the kernel
  // before this one in the stream (the previous update, or the first pass of an AUTO update); nothing in global
  // memory - not even the thresholds - is read before that kernel has completed.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (P.only_if_flag && *(volatile int*)&ws->flag_oor == 0) return;
  BinMap m;
  const bool sorted = load_thresholds(P.thr, T, s_thr, m);
  if (!sorted) {
    if (blockIdx.x == 0 && tid == 0) atomicOr(P.status, TMB_STATUS_THRESHOLD_ORDER);
    return;
  }
  const UniformMap um = detect_uniform(s_thr, T);
  if (!um.ok) build_lut(T, s_thr, s_lut, m);  // evenly spaced thresholds never consult the table

  const float* __restrict__ preds = static_cast<const float*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long n = P.n_rows;
  const long long nvec = n >> 2;
  const long long stride = (long long)gridDim.x * NT;
  const long long ignore = P.ignore;
  u32* const my = s_hist + (warp % copies) * NB2;
  int oor = 0, vbits = 0;

  auto count = [&](float p, TT tv, auto uniform_tag) {
    const long long tl = as_i64(tv);
    bool valid = true;
    if constexpr (HAS_IGNORE) valid = tl != ignore;
    vbits |= (valid && (u64)tl > 1ULL) ? TMB_STATUS_TARGET_VALUES : 0;  // fused validation: target not in {0, 1}
    if constexpr (CHECK) oor |= (int)(valid & !(p >= 0.f && p <= 1.f));
    if constexpr (SIGMOID) p = sigmoid_ref(p);
    int cnt;
    if constexpr (decltype(uniform_tag)::value) {
      const float x = fmaf(p - um.lo, um.inv_step, 0.5f);
      const unsigned j = min(__float2uint_rz(x), um.jmax);
      cnt = (int)j + (int)(p >= s_thr[j]);
    } else {
      cnt = bin_of(p, m, s_thr, s_lut);
    }
    if (valid) atomicAdd(my + cnt + (tl == 1 ? NB : 0), 1u);
  };

  auto run = [&](auto uniform_tag) {
    long long it = 0;
    for (long long g = (long long)blockIdx.x * NT + tid; g < nvec; g += stride, ++it) {
      const Vec4<float> p = load4(preds + 4 * g);
      const Vec4<TT> t = load4(target + 4 * g);
#pragma unroll
      for (int j = 0; j < 4; ++j) count(p.v[j], t.v[j], uniform_tag);
      if constexpr (CHECK) {
        if ((it & 31) == 31) {
          if (oor) atomicOr(&ws->flag_oor, 1);
          if (*(volatile int*)&ws->flag_oor) break;  // the pass is void: every thread leaves at its next check
        }
      }
    }
    if (blockIdx.x == 0 && tid < (int)(n & 3)) {
      const long long i = (nvec << 2) + tid;
      count(preds[i], target[i], uniform_tag);
    }
  };
  if (um.ok) run(std::true_type{}); else run(std::false_type{});
  // every CTA of this grid is resident (at most two per SM are launched), so the next kernel of the stream may start
  // taking the places of the CTAs that leave; it waits for this grid to complete before it touches global memory
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (CHECK && oor) atomicOr(&ws->flag_oor, 1);
  if (vbits) report_status(P.status, vbits);
  __syncthreads();

  // copies -> global scratch (u32 counts cannot overflow: the launcher keeps a launch below 2^32 samples)
  u64* hist = reinterpret_cast<u64*>(ws + 1);
  for (int b = tid; b < NB2; b += NT) {
    u64 c = 0;
    for (int w = 0; w < copies; ++w) c += s_hist[w * NB2 + b];
    if (c) atomicAdd(hist + b, c);
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(&ws->ticket, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int flag = *(volatile int*)&ws->flag_oor;
  const bool discard = CHECK && flag != 0;
  finalize_columns(P, hist, s_acc, discard);
  if (tid == 0) {
    ws->ticket = 0;
    if (CHECK) {
      if (discard && P.status_bit_on_oor) {
        atomicOr(P.status, P.status_bit_on_oor);
        ws->flag_oor = 0;
      }
    } else if (P.only_if_flag) {
      ws->flag_oor = 0;
    }
  }
}

template <typename TT>
static void (*pick_warphist(int mode, bool has_ignore))(const BinnedParams) {
#define TMB_PICKW(MODE) \
  return has_ignore ? binned_warphist_kernel<TT, MODE, true> : binned_warphist_kernel<TT, MODE, false>;
  if (mode == kModeAsIs) { TMB_PICKW(kModeAsIs) }
  if (mode == kModeAsIsCheck) { TMB_PICKW(kModeAsIsCheck) }
  TMB_PICKW(kModeSigmoid)
#undef TMB_PICKW
}

// =================================================================================================
// multi-column kernel (multilabel / multiclass, float32 scores): the warp-histogram kernel with one block of bins
// per column, plus a finalize kernel that gives every column to a warp
// =================================================================================================
// Shared memory holds `copies` u32 copies of the C * 2 * (T+1) bins (as many as fit, at most one per warp; one copy
// for C = 100, T = 200).  Neighbouring lanes work on neighbouring columns, so even a single copy sees little
// same-address contention.  The per-CTA bins go to the global u64 scratch with REDs (zeros skipped), and a second
// kernel turns every column's histogram into suffix sums with one warp per column — the single-CTA, one-column-at-
// a-time finalize of the generic kernel cost 2.5 us per column, 250 us at C = 100.
constexpr int kMultiThreads = 1024;

template <typename TT, int MODE, bool HAS_IGNORE, bool CLASS_INDEX>
__global__ void __launch_bounds__(kMultiThreads, 2) binned_multicol_kernel(const BinnedParams P) {
  constexpr bool CHECK = MODE == kModeAsIsCheck;
  constexpr bool SIGMOID = MODE == kModeSigmoid;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BinnedWs* ws = P.ws;
  if (P.only_if_flag && ws->flag_oor == 0) return;

  const int NT = kMultiThreads, tid = threadIdx.x, warp = tid >> 5;
  const int T = P.T, NB = T + 1, NB2 = 2 * NB;
  const u32 C = (u32)P.num_columns;
  const u32 nbins = C * (u32)NB2;
  const int copies = P.hist_copies;
  u32* s_hist = reinterpret_cast<u32*>(smem_raw);  // [copies][C][2][NB]
  float* s_thr = reinterpret_cast<float*>(smem_raw + (size_t)copies * nbins * 4);
  unsigned short* s_lut = reinterpret_cast<unsigned short*>(s_thr + NB);

  for (u32 i = tid; i < (u32)copies * nbins; i += NT) s_hist[i] = 0;
  BinMap m;
  const bool sorted = load_thresholds(P.thr, T, s_thr, m);
  if (!sorted) {
    if (blockIdx.x == 0 && tid == 0) atomicOr(P.status, TMB_STATUS_THRESHOLD_ORDER);
    return;
  }
  const UniformMap um = detect_uniform(s_thr, T);
  if (!um.ok) build_lut(T, s_thr, s_lut, m);  // evenly spaced thresholds never consult the table

  const float* __restrict__ preds = static_cast<const float*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long total = P.n_rows * (long long)C;  // < 2^32 (launcher)
  const long long nvec = total >> 2;
  const long long stride = (long long)gridDim.x * NT;
  const long long ignore = P.ignore;
  const bool range_sees_ignored = P.oor_counts_ignored != 0;
  u32* const my = s_hist + (u32)(warp % copies) * nbins;
  int oor = 0, vbits = 0;

  // tl: the element's own target (multilabel) or its row's class id (CLASS_INDEX)
  auto count = [&](float p, long long tl, u32 col, auto uniform_tag) {
    bool valid = true;
    if constexpr (HAS_IGNORE) valid = tl != ignore;
    vbits |= (valid && (u64)tl >= (u64)(CLASS_INDEX ? C : 2u)) ? TMB_STATUS_TARGET_VALUES : 0;
    if constexpr (CHECK) oor |= (int)((valid | range_sees_ignored) & !(p >= 0.f && p <= 1.f));
    if constexpr (SIGMOID) p = sigmoid_ref(p);
    int cnt;
    if constexpr (decltype(uniform_tag)::value) {
      const float x = fmaf(p - um.lo, um.inv_step, 0.5f);
      const unsigned j = min(__float2uint_rz(x), um.jmax);
      cnt = (int)j + (int)(p >= s_thr[j]);
    } else {
      cnt = bin_of(p, m, s_thr, s_lut);
    }
    const bool pos = CLASS_INDEX ? (tl == (long long)col) : (tl == 1);
    if (valid) atomicAdd(my + col * (u32)NB2 + (u32)cnt + (pos ? (u32)NB : 0u), 1u);
  };

  auto run = [&](auto uniform_tag) {
    // the column (and row) of a thread's vector advance by a fixed step per grid stride: no division in the loop
    const long long g0 = (long long)blockIdx.x * NT + tid;
    u32 col = (u32)((4 * g0) % C), row = (u32)((4 * g0) / C);
    const u32 step_col = (u32)((4 * stride) % C), step_row = (u32)((4 * stride) / C);
    long long it = 0;
    for (long long g = g0; g < nvec; g += stride, ++it) {
      const Vec4<float> p = load4(preds + 4 * g);
      u32 c = col;
      if constexpr (CLASS_INDEX) {
        u32 r = row;
        long long tl = as_i64(target[r]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          count(p.v[j], tl, c, uniform_tag);
          if (++c == C) {
            c = 0;
            ++r;
            if (j < 3 && r < (u32)P.n_rows) tl = as_i64(target[r]);
          }
        }
      } else {
        const Vec4<TT> t = load4(target + 4 * g);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          count(p.v[j], as_i64(t.v[j]), c, uniform_tag);
          if (++c == C) c = 0;
        }
      }
      col += step_col;
      row += step_row;
      if (col >= C) {
        col -= C;
        ++row;
      }
      if constexpr (CHECK) {
        if ((it & 31) == 31) {
          if (oor) atomicOr(&ws->flag_oor, 1);
          if (*(volatile int*)&ws->flag_oor) break;  // the pass is void: every thread leaves at its next check
        }
      }
    }
    if (blockIdx.x == 0 && tid < (int)(total & 3)) {
      const long long e = (nvec << 2) + tid;
      const u32 r = (u32)(e / C), c = (u32)(e % C);
      count(preds[e], as_i64(CLASS_INDEX ? target[r] : target[e]), c, uniform_tag);
    }
  };
  if (um.ok) run(std::true_type{}); else run(std::false_type{});
  if (CHECK && oor) atomicOr(&ws->flag_oor, 1);
  if (vbits) report_status(P.status, vbits);
  __syncthreads();

  // copies -> global scratch (u32 counts cannot overflow: the launcher keeps a launch below 2^32 elements)
  u64* hist = reinterpret_cast<u64*>(ws + 1);
  for (u32 b = tid; b < nbins; b += NT) {
    u64 c = 0;
    for (int w = 0; w < copies; ++w) c += s_hist[(u32)w * nbins + b];
    if (c) atomicAdd(hist + b, c);
  }
}

// One warp per column: suffix sums of the column's negative / positive histograms, added to confmat[:, c];
// leaves the scratch zeroed.  The last CTA settles the pass flags.
constexpr int kFinalizeThreads = 256;

__global__ void __launch_bounds__(kFinalizeThreads) binned_finalize_kernel(const BinnedParams P) {
  BinnedWs* ws = P.ws;
  const int flag = *(volatile int*)&ws->flag_oor;
  if (P.only_if_flag && flag == 0) return;  // pass 2 of AUTO did not run
  const bool discard = P.check_range && flag != 0;
  const int NB = P.T + 1;
  const u32 C = (u32)P.num_columns;
  const int lane = threadIdx.x & 31;
  u64* hist = reinterpret_cast<u64*>(ws + 1);
  const u32 warps = gridDim.x * (kFinalizeThreads / 32);
  for (u32 c = blockIdx.x * (kFinalizeThreads / 32) + (threadIdx.x >> 5); c < C; c += warps) {
    u64* hn = hist + (size_t)c * 2 * NB;
    u64* hp = hn + NB;
    if (discard) {
      for (int b = lane; b < NB; b += 32) {
        hn[b] = 0;
        hp[b] = 0;
      }
      continue;
    }
    // totals first, then a running suffix from the top bin down, 32 bins per step
    u64 tn = 0, tp = 0;
    for (int b = lane; b < NB; b += 32) {
      tn += __ldcg(hn + b);
      tp += __ldcg(hp + b);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      tn += __shfl_xor_sync(0xffffffffu, tn, o);
      tp += __shfl_xor_sync(0xffffffffu, tp, o);
    }
    u64 carry_n = 0, carry_p = 0;
    for (int top = NB - 1; top >= 0; top -= 32) {
      const int b = top - lane;  // lane 0 holds the highest bin of the step
      u64 vn = 0, vp = 0;
      if (b >= 0) {
        vn = __ldcg(hn + b);
        vp = __ldcg(hp + b);
        hn[b] = 0;
        hp[b] = 0;
      }
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {  // inclusive scan over lanes 0..lane == bins top..b
        const u64 yn = __shfl_up_sync(0xffffffffu, vn, o), yp = __shfl_up_sync(0xffffffffu, vp, o);
        if (lane >= o) {
          vn += yn;
          vp += yp;
        }
      }
      vn += carry_n;
      vp += carry_p;
      carry_n = __shfl_sync(0xffffffffu, vn, 31);
      carry_p = __shfl_sync(0xffffffffu, vp, 31);
      if (b >= 1) {  // bin b counts scores >= thr[b-1]: suffix sum from b = predicted positive at threshold b-1
        u64* o = P.confmat + ((size_t)(b - 1) * C + c) * 4;
        o[0] += tn - vn;  // [0][0] tn
        o[1] += vn;       // [0][1] fp
        o[2] += tp - vp;  // [1][0] fn
        o[3] += vp;       // [1][1] tp
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&ws->ticket, 1) == (int)gridDim.x - 1) {
      ws->ticket = 0;
      if (P.check_range) {
        if (discard && P.status_bit_on_oor) {
          atomicOr(P.status, P.status_bit_on_oor);
          ws->flag_oor = 0;
        }
      } else if (P.only_if_flag) {
        ws->flag_oor = 0;
      }
    }
  }
}

template <typename TT>
static void (*pick_multicol(int mode, bool has_ignore, bool class_index))(const BinnedParams) {
#define TMB_PICKM(MODE)                                                                                        \
  return class_index ? (has_ignore ? binned_multicol_kernel<TT, MODE, true, true>                              \
                                   : binned_multicol_kernel<TT, MODE, false, true>)                            \
                     : (has_ignore ? binned_multicol_kernel<TT, MODE, true, false>                             \
                                   : binned_multicol_kernel<TT, MODE, false, false>);
  if (mode == kModeAsIs) { TMB_PICKM(kModeAsIs) }
  if (mode == kModeAsIsCheck) { TMB_PICKM(kModeAsIsCheck) }
  TMB_PICKM(kModeSigmoid)
#undef TMB_PICKM
}

// =================================================================================================
// atomic kernel (any number of columns, any T)
// =================================================================================================
constexpr int kAtomicThreads = 512;

template <typename PT, typename TT, bool SIGMOID, bool SMEM_BINS>
__global__ void __launch_bounds__(kAtomicThreads) binned_atomic_kernel(const BinnedParams P) {
  using CT = typename CmpType<PT>::type;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  BinnedWs* ws = P.ws;
  if (P.only_if_flag && ws->flag_oor == 0) return;

  const int NT = blockDim.x, tid = threadIdx.x;
  const int T = P.T, NB = T + 1;
  const long long C = P.num_columns;
  const long long nbins = C * 2 * NB;
  float* s_thr = reinterpret_cast<float*>(smem_raw);
  unsigned short* s_lut = reinterpret_cast<unsigned short*>(s_thr + NB);
  // scan buffer for finalize (2*NB u64), 8-byte aligned after the LUT
  size_t off = ((size_t)NB * 4 + kLutCells * 2 + 7) & ~(size_t)7;
  u64* s_scan = reinterpret_cast<u64*>(smem_raw + off);
  u32* s_bins = reinterpret_cast<u32*>(s_scan + 2 * NB);
  __shared__ int s_last;

  if constexpr (SMEM_BINS)
    for (long long i = tid; i < nbins; i += NT) s_bins[i] = 0;
  BinMap m;
  const bool sorted = build_binning(P.thr, T, s_thr, s_lut, m);
  if (!sorted) {
    if (blockIdx.x == 0 && tid == 0) atomicOr(P.status, TMB_STATUS_THRESHOLD_ORDER);
    return;
  }
  u64* hist = reinterpret_cast<u64*>(ws + 1);
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long total = P.n_rows * C;
  const long long nvec = total >> 2;
  const long long stride = (long long)gridDim.x * NT;
  int oor = 0, vbits = 0;

  auto count = [&](PT pv, long long tl, long long col) {
    // tl: multilabel -> the element's own target; class-index mode -> the row's class id
    const bool ign = P.has_ignore && tl == P.ignore;
    CT p = to_cmp(pv);
    if (P.check_range && (!ign || P.oor_counts_ignored)) oor |= out_of_unit(p);
    if (ign) return;
    // fused validation: class id outside [0, C) / binary target outside {0, 1}
    vbits |= ((u64)tl >= (u64)(P.target_is_class_index ? C : 2)) ? TMB_STATUS_TARGET_VALUES : 0;
    if constexpr (SIGMOID) p = round_to_storage<PT>(sigmoid_ref(p));
    const bool pos = P.target_is_class_index ? (tl == col) : (tl == 1);
    long long idx = (col * 2 + (pos ? 1 : 0)) * NB + bin_of(p, m, s_thr, s_lut);
    if constexpr (SMEM_BINS) atomicAdd(s_bins + idx, 1u); else atomicAdd(hist + idx, 1ULL);
  };

  long long it = 0;
  for (long long g = (long long)blockIdx.x * NT + tid; g < nvec; g += stride, ++it) {
    const long long e = g << 2;
    Vec4<PT> p = load4(preds + e);
    long long row = e / C, col = e - row * C;
    if (P.target_is_class_index) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        count(p.v[j], as_i64(target[row]), col);
        if (++col == C) { col = 0; ++row; }
      }
    } else {
      Vec4<TT> t = load4(target + e);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        count(p.v[j], as_i64(t.v[j]), col);
        if (++col == C) col = 0;
      }
    }
    if (P.check_range && (it & 63) == 63) {
      if (oor) atomicOr(&ws->flag_oor, 1);
      if (*(volatile int*)&ws->flag_oor) break;
    }
  }
  if (blockIdx.x == 0 && tid < (total & 3)) {
    long long e = (nvec << 2) + tid;
    long long row = e / C, col = e - row * C;
    count(preds[e], as_i64(P.target_is_class_index ? target[row] : target[e]), col);
  }
  if (P.check_range && oor) atomicOr(&ws->flag_oor, 1);
  if (vbits) report_status(P.status, vbits);
  __syncthreads();
  if constexpr (SMEM_BINS) {
    for (long long i = tid; i < nbins; i += NT) {
      u32 c = s_bins[i];
      if (c) atomicAdd(hist + i, (u64)c);
    }
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) s_last = (atomicAdd(&ws->ticket, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const int flag = *(volatile int*)&ws->flag_oor;
  const bool discard = P.check_range && flag != 0;
  finalize_columns(P, hist, s_scan, discard);
  if (tid == 0) {
    ws->ticket = 0;
    if (P.check_range) {
      if (discard && P.status_bit_on_oor) {
        atomicOr(P.status, P.status_bit_on_oor);
        ws->flag_oor = 0;
      }
    } else if (P.only_if_flag) {
      ws->flag_oor = 0;
    }
  }
}

// ---- host -----------------------------------------------------------------------------------------
template <typename PT, typename TT>
static int launch_binned(BinnedParams P, int prob_mode, cudaStream_t st) {
  const long long total = P.n_rows * P.num_columns;
  const int NB = P.T + 1;
  const long long nbins = P.num_columns * 2 * NB;
  // one column of float32 scores: warp-histogram kernel; everything else: the generic atomic kernel
  const int copies = (std::is_same<PT, float>::value && P.num_columns == 1 && !P.target_is_class_index &&
                      total < (1LL << 32)) ? warphist_copies(P.T) : 0;

  // several columns of float32 scores: per-column shared-memory bins + warp-per-column finalize
  const size_t multi_fixed = (size_t)NB * 4 + kLutCells * 2 + 16;
  const bool multicol = std::is_same<PT, float>::value && !copies && total < (1LL << 32) && P.T <= 65535 &&
                        multi_fixed + (size_t)nbins * 4 <= (size_t)max_optin_smem() - 1024;

  auto run = [&](bool sigmoid) -> int {
    if (multicol) {
      // two CTAs per SM (64 resident warps) when at least four copies of the bins fit in half an SM's shared memory
      const bool two = multi_fixed + (size_t)nbins * 4 * 4 <= kWarpHistSmemBudget;
      size_t room = (two ? kWarpHistSmemBudget : (size_t)max_optin_smem() - 1024) - multi_fixed;
      size_t c = room / ((size_t)nbins * 4);
      P.hist_copies = (int)(c > kMultiThreads / 32 ? kMultiThreads / 32 : c);
      const size_t smem = multi_fixed + (size_t)P.hist_copies * nbins * 4;
      long long want = ((total >> 2) + kMultiThreads - 1) / kMultiThreads;
      long long cap = (long long)sm_count() * (two ? 2 : 1);
      int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
      const int mode = sigmoid ? kModeSigmoid : (P.check_range ? kModeAsIsCheck : kModeAsIs);
      auto k = pick_multicol<TT>(mode, P.has_ignore != 0, P.target_is_class_index != 0);
      TMB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, kMultiThreads, smem, st>>>(P);
      const int fgrid = (int)((P.num_columns + kFinalizeThreads / 32 - 1) / (kFinalizeThreads / 32));
      binned_finalize_kernel<<<fgrid, kFinalizeThreads, 0, st>>>(P);
    } else if (copies) {
      P.hist_copies = copies;
      const size_t smem = warphist_fixed_bytes(P.T) + (size_t)copies * 2 * NB * 4 + 8;
      long long want = ((total >> 2) + kWarpHistThreads - 1) / kWarpHistThreads;
      long long cap = (long long)sm_count() * 2;
      int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
      const int mode = sigmoid ? kModeSigmoid : (P.check_range ? kModeAsIsCheck : kModeAsIs);
      auto k = pick_warphist<TT>(mode, P.has_ignore != 0);
      TMB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      cudaLaunchConfig_t cfg{};
      cfg.gridDim = dim3((unsigned)grid);
      cfg.blockDim = dim3(kWarpHistThreads);
      cfg.dynamicSmemBytes = smem;
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      TMB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, k, P));
    } else {
      const size_t fixed = (((size_t)NB * 4 + kLutCells * 2 + 7) & ~(size_t)7) + (size_t)2 * NB * 8;
      const size_t with_bins = fixed + (size_t)nbins * 4;
      const bool smem_bins = with_bins <= (size_t)max_optin_smem() - 256;
      const size_t smem = smem_bins ? with_bins : fixed;
      if (fixed > (size_t)max_optin_smem() - 256) return TMB_ERR_TOO_LARGE;
      long long nvec = total >> 2;
      long long want = (nvec + kAtomicThreads - 1) / kAtomicThreads;
      int per_sm = smem > 100 * 1024 ? 1 : smem > 48 * 1024 ? 2 : 4;
      long long cap = (long long)sm_count() * per_sm;
      int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
      void (*k)(const BinnedParams);
      if (smem_bins) k = sigmoid ? binned_atomic_kernel<PT, TT, true, true> : binned_atomic_kernel<PT, TT, false, true>;
      else k = sigmoid ? binned_atomic_kernel<PT, TT, true, false> : binned_atomic_kernel<PT, TT, false, false>;
      TMB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      k<<<grid, kAtomicThreads, smem, st>>>(P);
    }
    TMB_CUDA_CHECK(cudaGetLastError());
    return TMB_OK;
  };

  int rc;
  switch (prob_mode) {
    case TMB_PROB_AS_IS:
      P.check_range = 0; P.only_if_flag = 0; P.status_bit_on_oor = 0;
      return run(false);
    case TMB_PROB_SIGMOID:
      P.check_range = 0; P.only_if_flag = 0; P.status_bit_on_oor = 0;
      return run(true);
    case TMB_PROB_AUTO:
      P.check_range = 1; P.only_if_flag = 0; P.status_bit_on_oor = 0;
      rc = run(false);
      if (rc != TMB_OK) return rc;
      P.check_range = 0; P.only_if_flag = 1;
      return run(true);
    case TMB_PROB_CHECK:
      P.check_range = 1; P.only_if_flag = 0; P.status_bit_on_oor = TMB_STATUS_OUT_OF_UNIT_RANGE;
      return run(false);
    default:
      return TMB_ERR_ARG;
  }
}

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_binned_curve_workspace_bytes(int num_thresholds, int64_t num_columns) {
  if (num_thresholds < 1 || num_columns < 1) return 0;
  return sizeof(BinnedWs) + (size_t)num_columns * 2 * ((size_t)num_thresholds + 1) * sizeof(u64);
}

// argument checks shared by the two entry points (nothing is enqueued before they pass)
static int binned_check_args(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n_rows,
                             int64_t num_columns, int target_is_class_index, const float* thresholds,
                             int num_thresholds, int prob_mode, const int64_t* confmat, const int32_t* status) {
  if (n_rows < 0 || num_columns < 1 || num_thresholds < 1 || num_thresholds > 65000 || !thresholds || !confmat ||
      !status)
    return TMB_ERR_ARG;
  if (!is_float_dtype(preds_dtype) || !is_int_dtype(target_dtype)) return TMB_ERR_ARG;
  if (prob_mode < TMB_PROB_AS_IS || prob_mode > TMB_PROB_CHECK) return TMB_ERR_ARG;
  // multiclass scores are normalised by a row softmax, not a sigmoid: only AS_IS / CHECK make sense there
  if (target_is_class_index && (prob_mode == TMB_PROB_AUTO || prob_mode == TMB_PROB_SIGMOID)) return TMB_ERR_ARG;
  if (n_rows == 0) return TMB_OK;
  if (!preds || !target) return TMB_ERR_ARG;
  if (!aligned32(preds) || !aligned32(target)) return TMB_ERR_ALIGN;
  return TMB_OK;
}

extern "C" int tmb_binned_curve_update(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                       int64_t n_rows, int64_t num_columns, int target_is_class_index,
                                       const float* thresholds, int num_thresholds, int has_ignore,
                                       int64_t ignore_index, int prob_mode, int64_t* confmat, void* workspace,
                                       size_t workspace_bytes, int32_t* status, void* stream) {
  if (!workspace || workspace_bytes < tmb_binned_curve_workspace_bytes(num_thresholds, num_columns))
    return TMB_ERR_ARG;
  const int rc = binned_check_args(preds, preds_dtype, target, target_dtype, n_rows, num_columns,
                                   target_is_class_index, thresholds, num_thresholds, prob_mode, confmat, status);
  if (rc != TMB_OK) return rc;
  if (n_rows == 0) return TMB_OK;
  if (!aligned32(workspace)) return TMB_ERR_ALIGN;
  BinnedParams P{};
  P.preds = preds;
  P.target = target;
  P.n_rows = n_rows;
  P.num_columns = num_columns;
  P.target_is_class_index = target_is_class_index ? 1 : 0;
  P.thr = thresholds;
  P.T = num_thresholds;
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  // the multilabel _format step range-checks before masking ignored entries (precision_recall_curve.py:756-766),
  // the binary and multiclass ones after dropping them (:176-184, :441-448)
  P.oor_counts_ignored = (num_columns > 1 && !target_is_class_index) ? 1 : 0;
  P.confmat = (u64*)confmat;
  P.ws = (BinnedWs*)workspace;
  P.status = status;
  cudaStream_t st = (cudaStream_t)stream;
  TMB_DISPATCH_FLOAT(preds_dtype, PT, {
    TMB_DISPATCH_INT(target_dtype, TT, { return launch_binned<PT, TT>(P, prob_mode, st); })
  })
  return TMB_ERR_ARG;
}

// =================================================================================================
// Entry point that allocates nothing (no counterpart in the reference)
// =================================================================================================
// Some callers of this library may not allocate device memory at the moment they want a curve update: an
// evaluation or serving loop that forbids allocation after warm-up (cudaMalloc / cudaFree synchronise the whole
// device and would stall every other stream), a process whose device memory is entirely owned by another pool so a
// raw cudaMalloc would fail or fragment it, or a thin foreign-language binding that is only ever handed pointers and
// has no device allocator of its own.  For them the kernel's scratch (flag, ticket, histogram) lives in this
// fixed array, which every GPU that loads the library holds until the process exits.
static __device__ __align__(256) unsigned char g_binned_static_scratch[TMB_BINNED_STATIC_SCRATCH_BYTES];

namespace {
// In host code the array's name is only a stand-in: the device address has to be asked for, once per device,
// with that device current.
std::mutex g_static_mu;
void* g_static_addr[64] = {};
}  // namespace

extern "C" int tmb_binned_curve_update_static(const void* preds, int preds_dtype, const void* target,
                                              int target_dtype, int64_t n_rows, int64_t num_columns,
                                              int target_is_class_index, const float* thresholds, int num_thresholds,
                                              int has_ignore, int64_t ignore_index, int prob_mode, int64_t* confmat,
                                              int32_t* status, void* stream) {
  if (num_thresholds < 1 || num_columns < 1) return TMB_ERR_ARG;
  const size_t need = tmb_binned_curve_workspace_bytes(num_thresholds, num_columns);
  if (need > (size_t)TMB_BINNED_STATIC_SCRATCH_BYTES) {
    std::string msg = "tmb_binned_curve_update_static: num_columns * 2 * (num_thresholds + 1) counters need " +
                      std::to_string(need) + " bytes of scratch but the built-in array holds " +
                      std::to_string((size_t)TMB_BINNED_STATIC_SCRATCH_BYTES) +
                      "; use tmb_binned_curve_update with a caller-provided workspace";
    set_error_text(msg.c_str());
    return TMB_ERR_TOO_LARGE;
  }
  // every check that can fail comes before the first thing put on the stream: a rejected call leaves the shared
  // array as the previous call left it
  const int rc = binned_check_args(preds, preds_dtype, target, target_dtype, n_rows, num_columns,
                                   target_is_class_index, thresholds, num_thresholds, prob_mode, confmat, status);
  if (rc != TMB_OK) return rc;
  if (n_rows == 0) return TMB_OK;
  int dev = 0;
  TMB_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error_text("tmb_binned_curve_update_static: only device ordinals 0..63 are supported");
    return TMB_ERR_ARG;
  }
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_addr[dev]) {
      // first call on this device: resolve the device address of the array (do this before any stream capture)
      void* p = nullptr;
      TMB_CUDA_CHECK(cudaGetSymbolAddress(&p, g_binned_static_scratch));
      g_static_addr[dev] = p;
    }
    scratch = g_static_addr[dev];
  }
  // the array keeps whatever the previous call left behind; flag, ticket and histogram must start at zero
  TMB_CUDA_CHECK(cudaMemsetAsync(scratch, 0, need, (cudaStream_t)stream));
  return tmb_binned_curve_update(preds, preds_dtype, target, target_dtype, n_rows, num_columns, target_is_class_index,
                                 thresholds, num_thresholds, has_ignore, ignore_index, prob_mode, confmat, scratch,
                                 (size_t)TMB_BINNED_STATIC_SCRATCH_BYTES, status, stream);
}
