// K1 — confusion-matrix update from label tensors (see include/tm_b200.h for the contract).
//
// HBM-bound streaming histogram: every (pred, target) pair is read once with 128/256-bit loads and turned
// into one integer increment of confmat[target*C + pred].
//   * large C (C*C bins do not fit shared memory; C = 1000 -> 10^6 bins = 8 MB, L2-resident on B200):
//     one 64-bit RED (fire-and-forget atomic) per sample straight into the int64 state. No key tensor,
//     no scratch, no second pass.
//   * small C: bins privatised in shared memory as u32 (one copy per warp when they are few), shared-memory
//     atomics, one global RED per non-zero bin per CTA.
#include "common.cuh"

namespace tmb {

constexpr int kConfmatThreads = 256;

template <typename PT, typename TT>
__device__ __forceinline__ bool make_key(PT p, TT t, long long C, long long CC, bool has_ignore, long long ignore,
                                         long long& key, int& bad) {
  const long long tl = as_i64(t), pl = as_i64(p);
  if (has_ignore && tl == ignore) return false;
  // fused validation: labels outside [0, C) (unsigned compare catches negatives too)
  bad |= ((u64)tl >= (u64)C ? TMB_STATUS_TARGET_VALUES : 0) | ((u64)pl >= (u64)C ? TMB_STATUS_PREDS_VALUES : 0);
  key = tl * C + pl;
  if (key < 0 || key >= CC) {
    bad |= TMB_STATUS_LABEL_RANGE;
    return false;
  }
  return true;
}

// ---- large C: global RED ---------------------------------------------------------------------
// one RED per sample; AGG: lanes of a warp holding the same key are merged first (match.any), which pays when
// the data are concentrated on few cells (diagonal-heavy confusion matrices) and costs ~10% when they are not
// Default policy: plain REDs, unless the warp looks concentrated.  The probe compares every lane's key with its
// neighbour's (one shuffle): uniform data almost never has 4 equal neighbours among 32 lanes, while a cell holding a
// third of the samples has ~3.5 on average and a dominant one far more; only then is the match.any merge paid.
__device__ __forceinline__ void global_count_adaptive(u64* confmat, long long key, bool valid) {
  const unsigned lane = threadIdx.x & 31;
  const long long nb = __shfl_sync(0xffffffffu, key, (lane + 1) & 31);
  const unsigned eq = __ballot_sync(0xffffffffu, valid && nb == key);
  if (__popc(eq) >= 4) {
    const unsigned active = __ballot_sync(0xffffffffu, valid);
    if (valid) {
      const unsigned peers = __match_any_sync(active, (unsigned long long)key);
      if ((int)lane == __ffs(peers) - 1) atomicAdd(confmat + key, (u64)__popc(peers));
    }
  } else if (valid) {
    atomicAdd(confmat + key, 1ULL);
  }
}

template <bool AGG>
__device__ __forceinline__ void global_count(u64* confmat, long long key, bool valid) {
  if constexpr (AGG) {
    unsigned active = __ballot_sync(0xffffffffu, valid);
    if (valid) {
      unsigned peers = __match_any_sync(active, (unsigned long long)key);
      if ((int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(confmat + key, (u64)__popc(peers));
    }
  } else {
    if (valid) atomicAdd(confmat + key, 1ULL);
  }
}

template <typename PT, typename TT, int UNROLL, int MINB, int AGG>  // AGG: 0 never, 1 always, 2 adaptive
__global__ void __launch_bounds__(kConfmatThreads, MINB)
confmat_global_kernel(const PT* __restrict__ preds, const TT* __restrict__ target, long long n, long long C,
                      int has_ignore, long long ignore, u64* __restrict__ confmat, int* __restrict__ status) {
  const long long CC = C * C;
  const long long nvec = n >> 2;  // groups of 4 samples
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int bad = 0;
  // every lane of a warp runs the same number of iterations (match.any needs converged warps)
  const long long per = (long long)UNROLL * stride;
  const long long iters = (nvec + per - 1) / per;
  for (long long it = 0; it < iters; ++it, g += per) {
    Vec4<PT> p[UNROLL];
    Vec4<TT> t[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      if (g + u * stride < nvec) {
        p[u] = load4(preds + 4 * (g + u * stride));
        t[u] = load4(target + 4 * (g + u * stride));
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const bool in = g + u * stride < nvec;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        long long key = 0;
        const bool valid = in && make_key(p[u].v[j], t[u].v[j], C, CC, has_ignore, ignore, key, bad);
        if constexpr (AGG == 2) global_count_adaptive(confmat, key, valid);
        else global_count<AGG == 1>(confmat, key, valid);
      }
    }
  }
  // ragged tail (< 4 samples)
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    long long i = (nvec << 2) + threadIdx.x;
    long long key;
    if (make_key(preds[i], target[i], C, CC, has_ignore, ignore, key, bad)) atomicAdd(confmat + key, 1ULL);
  }
  if (bad) report_status(status, bad);
}

// ---- small C: shared-memory privatised ----------------------------------------------------------
// u32 bins in shared memory, updated with shared-memory atomics (>1500 G/s chip-wide on B200, so no warp-level
// aggregation: match.any costs ten times more than the atomics it would save).  When C*C is small the CTA keeps
// several copies of the bins (copy = warp % copies) so that only lanes of one warp contend for a cell; 1024
// threads per CTA, two CTAs per SM when they fit = 64 resident warps to hide the HBM latency.
constexpr int kSmemThreads = 1024;

template <typename PT, typename TT>
__global__ void __launch_bounds__(kSmemThreads)
confmat_smem_kernel(const PT* __restrict__ preds, const TT* __restrict__ target, long long n, long long C,
                    int has_ignore, long long ignore, int copies, u64* __restrict__ confmat,
                    int* __restrict__ status) {
  extern __shared__ u32 hist[];
  const int CC = (int)(C * C);
  for (int i = threadIdx.x; i < copies * CC; i += blockDim.x) hist[i] = 0;
  __syncthreads();
  u32* const my = hist + ((threadIdx.x >> 5) % copies) * CC;
  const long long nvec = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  int bad = 0;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < nvec; g += stride) {
    const Vec4<PT> p = load4(preds + 4 * g);
    const Vec4<TT> t = load4(target + 4 * g);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      long long key = 0;
      if (make_key(p.v[j], t.v[j], C, (long long)CC, has_ignore, ignore, key, bad)) atomicAdd(my + key, 1u);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (nvec << 2) + threadIdx.x;
    long long key = 0;
    if (make_key(preds[i], target[i], C, (long long)CC, has_ignore, ignore, key, bad)) atomicAdd(my + key, 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < CC; i += blockDim.x) {
    u64 c = 0;
    for (int w = 0; w < copies; ++w) c += hist[w * CC + i];
    if (c) atomicAdd(confmat + i, c);
  }
  if (bad) report_status(status, bad);
}

// plain histogram of one label tensor (utilities/data.py:179-207 `_bincount`): global RED per element
template <typename PT>
__global__ void __launch_bounds__(kConfmatThreads)
bincount_kernel(const PT* __restrict__ x, long long n, long long nbins, u64* __restrict__ out, int* __restrict__ status) {
  const long long nvec = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  int bad = 0;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < nvec; g += stride) {
    Vec4<PT> v = load4(x + 4 * g);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      long long k = as_i64(v.v[j]);
      if (k < 0 || k >= nbins) bad = 1; else atomicAdd(out + k, 1ULL);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    long long k = as_i64(x[(nvec << 2) + threadIdx.x]);
    if (k < 0 || k >= nbins) bad = 1; else atomicAdd(out + k, 1ULL);
  }
  if (bad) report_status(status, bad);
}

// n_valid / n_equal (multiclass micro statistics)
template <typename PT, typename TT>
__global__ void __launch_bounds__(kConfmatThreads)
match_count_kernel(const PT* __restrict__ preds, const TT* __restrict__ target, long long n, int has_ignore,
                   long long ignore, u64* __restrict__ out2) {
  const long long nvec = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  u32 valid = 0, equal = 0;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < nvec; g += stride) {
    Vec4<PT> p = load4(preds + 4 * g);
    Vec4<TT> t = load4(target + 4 * g);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      long long tl = as_i64(t.v[j]);
      bool ok = !(has_ignore && tl == ignore);
      valid += ok;
      equal += ok && (tl == as_i64(p.v[j]));
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    long long i = (nvec << 2) + threadIdx.x;
    long long tl = as_i64(target[i]);
    bool ok = !(has_ignore && tl == ignore);
    valid += ok;
    equal += ok && (tl == as_i64(preds[i]));
  }
  valid = __reduce_add_sync(0xffffffffu, valid);
  equal = __reduce_add_sync(0xffffffffu, equal);
  __shared__ u32 s[2];
  if (threadIdx.x < 2) s[threadIdx.x] = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&s[0], valid);
    atomicAdd(&s[1], equal);
  }
  __syncthreads();
  if (threadIdx.x < 2 && s[threadIdx.x]) atomicAdd(out2 + threadIdx.x, (u64)s[threadIdx.x]);
}

static int grid_for(long long nvec, int threads, int ctas_per_sm) {
  long long want = (nvec + threads - 1) / threads;
  long long cap = (long long)sm_count() * ctas_per_sm;
  if (want < 1) want = 1;
  return (int)(want < cap ? want : cap);
}

template <typename PT, typename TT>
static int launch_confmat(const PT* preds, const TT* target, long long n, long long C, int has_ignore,
                          long long ignore, u64* confmat, int* status, int algo, cudaStream_t st) {
  const long long CC = C * C;
  const size_t smem = (size_t)CC * sizeof(u32);
  const bool fits = smem <= (size_t)max_optin_smem() - 1024;
  if (algo == 0) algo = (fits && CC <= 51200) ? 2 : 1;  // <= 200 KB of u32 bins: one privatised copy per CTA
  if (algo == 2 && !fits) return TMB_ERR_TOO_LARGE;
  const long long nvec = n >> 2;
  if (algo == 1) {
    // 4 resident CTAs of 256 threads per SM, 2 groups of 4 samples in flight per thread
    int grid = grid_for((nvec + 1) / 2, kConfmatThreads, 4);
    confmat_global_kernel<PT, TT, 2, 4, 2><<<grid, kConfmatThreads, 0, st>>>(preds, target, n, C, has_ignore,
                                                                                 ignore, confmat, status);
  } else if (algo == 3) {
    int grid = grid_for((nvec + 1) / 2, kConfmatThreads, 4);
    confmat_global_kernel<PT, TT, 2, 4, 0><<<grid, kConfmatThreads, 0, st>>>(preds, target, n, C, has_ignore,
                                                                                 ignore, confmat, status);
  } else if (algo == 4) {
    int grid = grid_for(nvec, kConfmatThreads, 8);
    confmat_global_kernel<PT, TT, 1, 8, 0><<<grid, kConfmatThreads, 0, st>>>(preds, target, n, C, has_ignore,
                                                                                 ignore, confmat, status);
  } else if (algo == 5) {
    int grid = grid_for((nvec + 1) / 2, kConfmatThreads, 4);
    confmat_global_kernel<PT, TT, 2, 4, 1><<<grid, kConfmatThreads, 0, st>>>(preds, target, n, C, has_ignore,
                                                                                ignore, confmat, status);
  } else if (algo != 2) {
    return TMB_ERR_ARG;
  } else {
    // One CTA of 1024 threads per SM.  Copies of the bins: one per warp while 32 of them fit 200 KB (C <= 39), else
    // as many as fit 100 KB - more copies than that cost more to clear and merge than they save in contention.
    // Measured against two CTAs per SM with 100 KB each (profiles/r2_k1s_variants.log, 2^24 / 2^26 int64 pairs):
    // C=10 89.8 / 101 -> 90 / 105 % of the measured copy peak, C=32 82.5 / 94.9 -> 86.2 / 103.6 %, C=100 79.4 / 99
    // -> 86 / 103.5 %, C=150 73.4 -> 82.5 % at 2^24: half as many CTAs end with C*C atomics on the same C*C cells.
    int copies;
    if (smem * (kSmemThreads / 32) <= (size_t)200 * 1024) copies = kSmemThreads / 32;
    else copies = smem <= (size_t)100 * 1024 ? (int)((size_t)100 * 1024 / smem) : 1;
    const size_t dyn = smem * copies;
    const int per_sm = 1;
    int grid = grid_for(nvec, kSmemThreads, per_sm);
    auto k = confmat_smem_kernel<PT, TT>;
    TMB_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
    k<<<grid, kSmemThreads, dyn, st>>>(preds, target, n, C, has_ignore, ignore, copies, confmat, status);
  }
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}

}  // namespace tmb

using namespace tmb;

extern "C" int tmb_confmat_update(const void* preds, int preds_dtype, const void* target, int target_dtype, int64_t n,
                                  int64_t num_classes, int has_ignore, int64_t ignore_index, int64_t* confmat,
                                  int32_t* status, int algo, void* stream) {
  if (n < 0 || num_classes < 1 || num_classes > 3000000 || !confmat || !status) return TMB_ERR_ARG;
  if (n == 0) return TMB_OK;
  if (!preds || !target) return TMB_ERR_ARG;
  if (!aligned32(preds) || !aligned32(target)) return TMB_ERR_ALIGN;
  cudaStream_t st = (cudaStream_t)stream;
  TMB_DISPATCH_INT(preds_dtype, PT, {
    TMB_DISPATCH_INT(target_dtype, TT, {
      return launch_confmat<PT, TT>((const PT*)preds, (const TT*)target, n, num_classes, has_ignore, ignore_index,
                                    (u64*)confmat, status, algo, st);
    })
  })
  return TMB_ERR_ARG;
}

extern "C" int tmb_bincount_update(const void* x, int x_dtype, int64_t n, int64_t num_bins, int64_t* out,
                                   int32_t* status, void* stream) {
  if (n < 0 || num_bins < 1 || !out || !status) return TMB_ERR_ARG;
  if (n == 0) return TMB_OK;
  if (!x) return TMB_ERR_ARG;
  if (!aligned32(x)) return TMB_ERR_ALIGN;
  cudaStream_t st = (cudaStream_t)stream;
  TMB_DISPATCH_INT(x_dtype, PT, {
    int grid = grid_for(n >> 2, kConfmatThreads, 8);
    bincount_kernel<PT><<<grid, kConfmatThreads, 0, st>>>((const PT*)x, n, num_bins, (u64*)out, status);
    TMB_CUDA_CHECK(cudaGetLastError());
    return TMB_OK;
  })
  return TMB_ERR_ARG;
}

extern "C" int tmb_match_count_update(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                      int64_t n, int has_ignore, int64_t ignore_index, int64_t* out2, void* stream) {
  if (n < 0 || !out2) return TMB_ERR_ARG;
  if (n == 0) return TMB_OK;
  if (!preds || !target) return TMB_ERR_ARG;
  if (!aligned32(preds) || !aligned32(target)) return TMB_ERR_ALIGN;
  cudaStream_t st = (cudaStream_t)stream;
  TMB_DISPATCH_INT(preds_dtype, PT, {
    TMB_DISPATCH_INT(target_dtype, TT, {
      int grid = grid_for(n >> 2, kConfmatThreads, 8);
      match_count_kernel<PT, TT><<<grid, kConfmatThreads, 0, st>>>((const PT*)preds, (const TT*)target, n, has_ignore,
                                                                   ignore_index, (u64*)out2);
      TMB_CUDA_CHECK(cudaGetLastError());
      return TMB_OK;
    })
  })
  return TMB_ERR_ARG;
}
