// Score-range partition for the distributed exact curve (contract: tmb_range_partition in include/tm_b200.h).
//
// Every sample is assigned to one of R score ranges (range 0 = the highest scores) by R-1 ascending splitters and the
// samples are regrouped by range: scores as float32, the label as one byte "is positive".  Equal scores always land
// in the same range, so runs of ties never straddle two ranges (= two ranks after the all-to-all that follows).
//
// Two passes over the input, both coalesced; the unit of work is a WARP with a contiguous slice of the input:
//   count    per warp and range: samples and positives (ballots, no atomics)
//   offsets  one small kernel: exclusive prefix over (range, warp) -> where each warp's share of each range starts
//   scatter  the same walk again; a lane's slot = the warp's cursor of its range + the lower lanes of the same range
// The regrouping is stable, although nothing downstream needs that.
#include "common.cuh"

namespace tmb {

constexpr int kRpThreads = 256;
constexpr int kRpWarps = kRpThreads / 32;
constexpr int kRpMaxRanges = 64;

struct RpParams {
  const float* scores;
  const void* target;
  long long n;
  long long pos_label;
  const float* splitters;  // [R - 1] ascending
  int R;
  long long per_warp;      // elements per warp slice (multiple of 32)
  long long total_warps;
  u32* counts;             // [total_warps][R] samples per (warp, range); after `offsets`: exclusive start positions
  u32* poscnt;             // [total_warps][R] positives per (warp, range)
  float* out_scores;
  unsigned char* out_pos;
  long long* totals;       // [2][R] samples / positives per range
};

__device__ __forceinline__ int range_of(float x, const float* s_split, int R) {
  if (x != x) return 0;  // NaN ranks above everything in the descending order of the curve
  // number of splitters strictly below x (torch.bucketize, right=False); R is small, the loop is uniform
  int below = 0;
  for (int i = 0; i < R - 1; ++i) below += (s_split[i] < x) ? 1 : 0;
  return (R - 1) - below;
}

// RR > 0: at most RR ranges, the per-range cursors live in registers (every lane keeps the same copy), the loop over
// the ranges is unrolled and touches no shared memory.  RR == 0: any R <= 64, cursors in shared memory.
template <typename TT, bool SCATTER, int RR>
__global__ void __launch_bounds__(kRpThreads) range_partition_kernel(const RpParams P) {
  __shared__ float s_split[kRpMaxRanges];
  __shared__ u32 s_cur[kRpWarps][kRpMaxRanges];
  __shared__ u32 s_pos[kRpWarps][kRpMaxRanges];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = P.R;
  for (int i = threadIdx.x; i < R - 1; i += kRpThreads) s_split[i] = P.splitters[i];
  const long long gw = (long long)blockIdx.x * kRpWarps + warp;
  if (lane < R) {
    s_cur[warp][lane] = (SCATTER && gw < P.total_warps) ? P.counts[gw * R + lane] : 0u;
    s_pos[warp][lane] = 0;
  }
  if (lane + 32 < R) {
    s_cur[warp][lane + 32] = (SCATTER && gw < P.total_warps) ? P.counts[gw * R + lane + 32] : 0u;
    s_pos[warp][lane + 32] = 0;
  }
  __syncthreads();
  if (gw >= P.total_warps) return;
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long begin = gw * P.per_warp;
  const long long end = begin + P.per_warp < P.n ? begin + P.per_warp : P.n;
  const u32 lt = (1u << lane) - 1u;
  constexpr int NR = RR > 0 ? RR : 1;
  u32 cur[NR], pcnt[NR];
#pragma unroll
  for (int r = 0; r < NR; ++r) {
    cur[r] = (RR > 0 && r < R) ? s_cur[warp][r] : 0u;
    pcnt[r] = 0;
  }
  constexpr int kBatch = 4;  // steps whose loads are issued together
  for (long long i0 = begin; i0 < end; i0 += 32 * kBatch) {
    float xs[kBatch];
    bool ps[kBatch];
    int ds[kBatch];
#pragma unroll
    for (int k = 0; k < kBatch; ++k) {
      const long long i = i0 + 32 * k + lane;
      xs[k] = 0.f;
      ps[k] = false;
      if (i < end) {
        xs[k] = __ldg(P.scores + i);
        ps[k] = as_i64(__ldg(target + i)) == P.pos_label;
      }
    }
#pragma unroll
    for (int k = 0; k < kBatch; ++k) ds[k] = (i0 + 32 * k + lane < end) ? range_of(xs[k], s_split, R) : -1;
#pragma unroll
    for (int k = 0; k < kBatch; ++k) {
      const int d = ds[k];
      const float x = xs[k];
      const bool pos = ps[k];
      if constexpr (RR > 0) {
#pragma unroll
        for (int r = 0; r < RR; ++r) {
          const u32 m = __ballot_sync(0xffffffffu, d == r);
          if constexpr (SCATTER) {
            if (d == r) {
              const u32 slot = cur[r] + __popc(m & lt);
              P.out_scores[slot] = x;
              P.out_pos[slot] = pos ? 1 : 0;
            }
          } else {
            pcnt[r] += __popc(__ballot_sync(0xffffffffu, d == r && pos));
          }
          cur[r] += __popc(m);
        }
      } else {
        for (int r = 0; r < R; ++r) {
          const u32 m = __ballot_sync(0xffffffffu, d == r);
          if (m == 0) continue;  // uniform
          if constexpr (SCATTER) {
            if (d == r) {
              const u32 slot = s_cur[warp][r] + __popc(m & lt);
              P.out_scores[slot] = x;
              P.out_pos[slot] = pos ? 1 : 0;
            }
            __syncwarp();
            if (lane == 0) s_cur[warp][r] += __popc(m);
          } else {
            const u32 mp = __ballot_sync(0xffffffffu, d == r && pos);
            if (lane == 0) {
              s_cur[warp][r] += __popc(m);
              s_pos[warp][r] += __popc(mp);
            }
          }
          __syncwarp();
        }
      }
    }
  }
  if constexpr (!SCATTER) {
    if constexpr (RR > 0) {
#pragma unroll
      for (int r = 0; r < RR; ++r) {
        if (lane == 0 && r < R) {
          P.counts[gw * R + r] = cur[r];
          P.poscnt[gw * R + r] = pcnt[r];
        }
      }
    } else {
      __syncwarp();
      for (int r = lane; r < R; r += 32) {
        P.counts[gw * R + r] = s_cur[warp][r];
        P.poscnt[gw * R + r] = s_pos[warp][r];
      }
    }
  }
}

// One CTA of 1024 threads: for every range, an exclusive scan of the warps' counts (the (range, warp) order of the
// output) carried from range to range, and the range's totals.  The table is a few thousand warps x a few ranges.
__global__ void __launch_bounds__(1024) range_offsets_kernel(const RpParams P) {
  __shared__ u32 s_wsum[32], s_psum[32];
  __shared__ u32 s_carry;
  __shared__ unsigned long long s_ptotal;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0;
  __syncthreads();
  for (int r = 0; r < P.R; ++r) {
    const u32 range_base = s_carry;
    if (tid == 0) s_ptotal = 0;
    __syncthreads();
    for (long long start = 0; start < P.total_warps; start += 1024) {
      const long long i = start + tid;
      const u32 v = i < P.total_warps ? P.counts[i * P.R + r] : 0u;
      u32 pv = i < P.total_warps ? P.poscnt[i * P.R + r] : 0u;
      u32 x = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u32 y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
      }
      pv = __reduce_add_sync(0xffffffffu, pv);
      if (lane == 31) s_wsum[warp] = x;
      if (lane == 0) s_psum[warp] = pv;
      __syncthreads();
      u32 before = 0;
      for (int w = 0; w < warp; ++w) before += s_wsum[w];
      const u32 carry = s_carry;
      if (i < P.total_warps) P.counts[i * P.R + r] = carry + before + x - v;  // n < 2^32 (checked by the entry point)
      __syncthreads();
      if (tid == 1023) s_carry = carry + before + x;
      if (tid == 0) {
        unsigned long long ps = 0;
        for (int w = 0; w < 32; ++w) ps += s_psum[w];
        s_ptotal += ps;
      }
      __syncthreads();
    }
    if (tid == 0) {
      P.totals[r] = (long long)(s_carry - range_base);
      P.totals[P.R + r] = (long long)s_ptotal;
    }
    __syncthreads();
  }
}

static void rp_geometry(long long n, long long* per_warp, long long* total_warps) {
  long long warps = (long long)sm_count() * 4 * kRpWarps;  // four CTAs per SM
  long long per = (n + warps - 1) / warps;
  per = (per + 31) & ~31LL;
  if (per < 1024) per = 1024;
  *per_warp = per;
  *total_warps = (n + per - 1) / per;
}

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_range_partition_workspace_bytes(int64_t n, int num_ranges) {
  if (n < 1 || num_ranges < 1 || num_ranges > kRpMaxRanges) return 0;
  long long per, tw;
  rp_geometry(n, &per, &tw);
  return (size_t)2 * tw * num_ranges * sizeof(u32) + 256;
}

extern "C" int tmb_range_partition(const float* scores, const void* target, int target_dtype, int64_t n,
                                   int64_t pos_label, const float* splitters, int num_ranges, float* out_scores,
                                   uint8_t* out_positive, int64_t* totals2, void* workspace, size_t workspace_bytes,
                                   void* stream) {
  if (n < 0 || num_ranges < 1 || num_ranges > kRpMaxRanges || !totals2) return TMB_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    TMB_CUDA_CHECK(cudaMemsetAsync(totals2, 0, (size_t)2 * num_ranges * sizeof(int64_t), st));
    return TMB_OK;
  }
  if (n >= (1LL << 32)) return TMB_ERR_TOO_LARGE;
  if (!scores || !target || !out_scores || !out_positive || !workspace || (num_ranges > 1 && !splitters))
    return TMB_ERR_ARG;
  if (workspace_bytes < tmb_range_partition_workspace_bytes(n, num_ranges)) return TMB_ERR_ARG;
  RpParams P{};
  P.scores = scores;
  P.target = target;
  P.n = n;
  P.pos_label = pos_label;
  P.splitters = splitters;
  P.R = num_ranges;
  rp_geometry(n, &P.per_warp, &P.total_warps);
  P.counts = (u32*)workspace;
  P.poscnt = P.counts + P.total_warps * num_ranges;
  P.out_scores = out_scores;
  P.out_pos = out_positive;
  P.totals = (long long*)totals2;
  const int grid = (int)((P.total_warps + kRpWarps - 1) / kRpWarps);
  TMB_DISPATCH_INT(target_dtype, TT, {
    if (num_ranges <= 8) {
      range_partition_kernel<TT, false, 8><<<grid, kRpThreads, 0, st>>>(P);
      range_offsets_kernel<<<1, 1024, 0, st>>>(P);
      range_partition_kernel<TT, true, 8><<<grid, kRpThreads, 0, st>>>(P);
    } else {
      range_partition_kernel<TT, false, 0><<<grid, kRpThreads, 0, st>>>(P);
      range_offsets_kernel<<<1, 1024, 0, st>>>(P);
      range_partition_kernel<TT, true, 0><<<grid, kRpThreads, 0, st>>>(P);
    }
    TMB_CUDA_CHECK(cudaGetLastError());
    return TMB_OK;
  })
  return TMB_ERR_ARG;
}
