// K4 — exact binary curve: radix sort of the scores (descending) with a 1-byte label payload, then a scan that
// emits cumulative true / false positives at the last element of every run of equal scores
// (contract in include/tm_b200.h; reference: functional/classification/precision_recall_curve.py:28-80).
//
// Sort: least-significant-digit radix sort, 4 passes of 8 bits over order-preserving 32-bit keys, "onesweep" style:
// the digit histograms of all four passes are counted once up front, and every pass is ONE kernel whose tiles
// chain their digit counts through status words (decoupled look-back) instead of a histogram + scan + scatter
// triple.  The first pass reads the caller's scores and targets directly.  Ties keep their input order, which the
// curve does not depend on (counts are read only at run boundaries).
// Scan: tile sums of (positives, run ends) -> exclusive scan of tile sums -> per-tile rescan that writes
// (fps, tps, threshold) at the compacted position of every run end.  Counts are exact int64.
#include <cstdlib>

#include "common.cuh"

namespace tmb {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;  // 4096 keys per CTA
constexpr int kRadix = 256;
constexpr int kSortWarps = kSortThreads / 32;

// float -> u32 whose ascending order is the DESCENDING order of the floats (NaN with sign bit clear sorts first,
// as torch.argsort(descending=True) places it)
__device__ __forceinline__ u32 desc_key(float f) {
  u32 b = __float_as_uint(f);
  u32 asc = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  return ~asc;
}
__device__ __forceinline__ float key_to_float(u32 k) {
  u32 asc = ~k;
  u32 b = (asc & 0x80000000u) ? (asc & 0x7fffffffu) : ~asc;
  return __uint_as_float(b);
}

// ---- sort ---------------------------------------------------------------------------------------------------
// k4_digit_hist: one read of the scores -> the four 256-bin digit histograms of the whole input (the digit bases of
// every pass are known before the first pass starts).
__global__ void __launch_bounds__(512) k4_digit_hist(const float* __restrict__ preds, long long n,
                                                     u32* __restrict__ ghist) {
  __shared__ u32 h[4 * kRadix];
  for (int i = threadIdx.x; i < 4 * kRadix; i += blockDim.x) h[i] = 0;
  __syncthreads();
  const long long nvec = n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < nvec; g += stride) {
    const Vec4<float> p = load4(preds + 4 * g);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const u32 k = desc_key(p.v[q]);
      atomicAdd(&h[k & 0xff], 1u);
      atomicAdd(&h[kRadix + ((k >> 8) & 0xff)], 1u);
      atomicAdd(&h[2 * kRadix + ((k >> 16) & 0xff)], 1u);
      atomicAdd(&h[3 * kRadix + (k >> 24)], 1u);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const u32 k = desc_key(preds[(nvec << 2) + threadIdx.x]);
    atomicAdd(&h[k & 0xff], 1u);
    atomicAdd(&h[kRadix + ((k >> 8) & 0xff)], 1u);
    atomicAdd(&h[2 * kRadix + ((k >> 16) & 0xff)], 1u);
    atomicAdd(&h[3 * kRadix + (k >> 24)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 4 * kRadix; i += blockDim.x) {
    const u32 c = h[i];
    if (c) atomicAdd(ghist + i, c);
  }
}

// Tile status words of the decoupled look-back: {flag : 2 | count}.  flag 0 = nothing yet, 1 = the tile's own count
// of the digit, 2 = the count summed over this and all earlier tiles.  32-bit words hold inputs below 2^30 keys,
// 64-bit words the rest.
template <typename ST>
struct StatusWord;
template <>
struct StatusWord<u32> {
  static constexpr int kFlagShift = 30;
};
template <>
struct StatusWord<u64> {
  static constexpr int kFlagShift = 62;
};

// status words travel through the L2 with GPU-scope relaxed accesses (a word carries its own flag, so no fence)
#ifdef TMB_K4_VOLATILE_STATUS
__device__ __forceinline__ u32 ld_status(const u32* p) { return *reinterpret_cast<const volatile u32*>(p); }
__device__ __forceinline__ u64 ld_status(const u64* p) { return *reinterpret_cast<const volatile u64*>(p); }
#else
__device__ __forceinline__ u32 ld_status(const u32* p) {
  u32 v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ u64 ld_status(const u64* p) {
  u64 v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
#endif
__device__ __forceinline__ void st_status(u32* p, u32 v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_status(u64* p, u64 v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// peers &= lanes whose key agrees with mine in the bit(s) of `bitmask`: test, ballot, complement the ballot on the
// lanes where the bit is clear, and - four instructions per digit bit, written out because the compiler's version of
// the same expression takes six
__device__ __forceinline__ u32 match_bit(u32 peers, u32 key, u32 bitmask) {
  u32 out;
  asm volatile(
      "{\n"
      " .reg .pred p;\n"
      " .reg .b32 v, t;\n"
      " and.b32 t, %2, %3;\n"
      " setp.ne.u32 p, t, 0;\n"
      " vote.sync.ballot.b32 v, p, 0xffffffff;\n"
      " @!p not.b32 v, v;\n"
      " and.b32 %0, %1, v;\n"
      "}"
      : "=r"(out)
      : "r"(peers), "r"(key), "r"(bitmask));
  return out;
}

#ifndef TMB_K4_WINDOW
#define TMB_K4_WINDOW 2  // status words of this many earlier tiles are fetched per look-back step
#endif

// One radix pass in ONE kernel ("onesweep"): a CTA takes the next tile of 4096 keys, ranks them by the pass's 8-bit
// digit, publishes its 256 digit counts, adds up the counts of the tiles before it by walking their status words
// backwards until it meets an inclusive one, and writes its keys to their final places of this pass.  Tile ids come
// from a counter, so a tile only ever waits for tiles that are already running.
//
// The kernel is bound by the shared-memory / load-store pipe (one wavefront per clock and SM), not by HBM, so the
// design spends ALU instructions to save shared-memory traffic:
//  * Ranking is stable (least-significant-digit passes depend on it): warp w owns keys [512 w, 512 w + 512) of the
//    tile and takes them 32 per round.  The lanes holding the same digit are found with eight ballots (one per
//    digit bit, no memory traffic); a key's rank is the warp's running count of its digit plus the lower lanes of
//    its set, and the highest lane of the set stores the new count.  (A shared-memory formulation - OR the lane
//    bit into a mask word per digit, read it back - measured 15 wavefronts per round against 6 for this one.)
//  * The tile is staged in digit order with the label riding in the key: inside a digit run the lowest digit bit
//    is known from the position, so that bit carries the label while the key sits in shared memory.
// SHIFT == 0: the pass reads the caller's scores and targets and makes keys / labels on the fly.
#ifndef TMB_K4_OS_THREADS
#define TMB_K4_OS_THREADS 256
#endif
#ifndef TMB_K4_MIN_CTAS
#define TMB_K4_MIN_CTAS 4
#endif
constexpr int kOsThreads = TMB_K4_OS_THREADS;       // threads of a onesweep CTA (at least one per digit)
constexpr int kOsWarps = kOsThreads / 32;
constexpr int kOsItems = kSortTile / kOsThreads;    // keys per thread
static_assert(kOsThreads >= kRadix && kOsItems % 2 == 0 && (kOsWarps * kRadix / 4) % kOsThreads == 0, "geometry");
template <int SHIFT, typename TT, typename ST>
__global__ void __launch_bounds__(kOsThreads, TMB_K4_MIN_CTAS)
k4_onesweep(const u32* __restrict__ keys_in, const unsigned char* __restrict__ labs_in,
            const float* __restrict__ preds, const TT* __restrict__ target, long long pos_label, long long n,
            const u32* __restrict__ ghist, ST* __restrict__ status, u32* __restrict__ tile_counter,
            u32* __restrict__ keys_out, unsigned char* __restrict__ labs_out) {
  constexpr bool FIRST = SHIFT == 0;  // the pass over the lowest digit reads the caller's scores and targets
  constexpr int SH = StatusWord<ST>::kFlagShift;
  constexpr ST kValMask = (ST(1) << SH) - 1;
  __shared__ __align__(16) u32 s_keys[kSortTile];
  __shared__ __align__(16) u32 s_cnt[kOsWarps][kRadix];  // per-warp digit counts, then staging bases
  __shared__ u32 s_start[kRadix];                          // first position of each digit run in the staged tile
  __shared__ u32 s_dst[kRadix];                            // destination of staged position 0, per digit run
  __shared__ u32 s_wsum[kRadix / 32], s_gsum[kRadix / 32];
  __shared__ u32 s_tile;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  {
    uint4* z = reinterpret_cast<uint4*>(&s_cnt[0][0]);
#pragma unroll
    for (int i = 0; i < kOsWarps * kRadix / 4 / kOsThreads; ++i) z[i * kOsThreads + tid] = make_uint4(0, 0, 0, 0);
  }
  if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
  __syncthreads();
  const u32 tile = s_tile;
  const long long base = (long long)tile * kSortTile;
  const long long left = n - base;
  const int valid = left < kSortTile ? (int)left : kSortTile;
  const int wofs = warp * (32 * kOsItems) + lane;

  u32 key[kOsItems];
  u32 labbits = 0;
  if constexpr (FIRST) {
#pragma unroll
    for (int r = 0; r < kOsItems; ++r) {
      const int j = wofs + r * 32;
      float p = 0.f;
      long long t = 0;
      if (j < valid) {
        p = __ldg(preds + base + j);
        t = as_i64(__ldg(target + base + j));
      }
      key[r] = desc_key(p);
      labbits |= (u32)(t == pos_label) << r;
    }
  } else {
    // the key / label buffers are padded to whole tiles: loads past the end are harmless and never used
#pragma unroll
    for (int r = 0; r < kOsItems; ++r) key[r] = keys_in[base + wofs + r * 32];
#pragma unroll
    for (int r = 0; r < kOsItems; ++r) labbits |= (u32)(labs_in[base + wofs + r * 32] & 1u) << r;
  }

  u32 rank2[kOsItems / 2];  // two 16-bit ranks per register
  {
    u32* cnt = s_cnt[warp];
    const u32 lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kOsItems; ++r) {
      const bool in = wofs + r * 32 < valid;
      const u32 d = (key[r] >> SHIFT) & 0xffu;
      u32 peers = __ballot_sync(0xffffffffu, in);
#pragma unroll
      for (int bit = 0; bit < 8; ++bit) {
        peers = match_bit(peers, key[r], 1u << (SHIFT + bit));
      }
      const u32 before = cnt[d];
      const u32 rk = before + __popc(peers & lt);
      if (r & 1) rank2[r >> 1] |= rk << 16;
      else rank2[r >> 1] = rk;
      __syncwarp();
      if (in && (peers >> lane) == 1u) cnt[d] = before + __popc(peers);  // highest lane of the set
      __syncwarp();
    }
  }
  __syncthreads();

  // thread d <-> digit d (the first 256 threads)
  const bool digit_thread = tid < kRadix;
  u32 my_cnt = 0, my_start = 0, my_gbase = 0;
  u32 x = 0, gx = 0, g = 0;
  if (digit_thread) {
    const int d = tid;
    u32 run = 0;
#pragma unroll
    for (int w = 0; w < kOsWarps; ++w) run += s_cnt[w][d];
    my_cnt = run;
    // the first tile has nothing before it: its own counts are inclusive already
    st_status(status + (size_t)tile * kRadix + d, (ST(tile == 0 ? 2 : 1) << SH) | (ST)run);
    // exclusive scans over the digits: this tile's counts (start of the digit's run in the staged tile) and the
    // whole input's counts (start of the digit's region in the output)
    g = ghist[d];
    x = run;
    gx = g;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const u32 y = __shfl_up_sync(0xffffffffu, x, o), gy = __shfl_up_sync(0xffffffffu, gx, o);
      if (lane >= o) {
        x += y;
        gx += gy;
      }
    }
    if (lane == 31) {
      s_wsum[warp] = x;
      s_gsum[warp] = gx;
    }
  }
  __syncthreads();
  if (digit_thread) {
    const int d = tid;
    u32 add = 0, gadd = 0;
    for (int w = 0; w < warp; ++w) {
      add += s_wsum[w];
      gadd += s_gsum[w];
    }
    my_start = add + x - my_cnt;
    my_gbase = gadd + gx - g;
    s_start[d] = my_start;
    u32 at = my_start;  // staging base of (warp, digit): the digit's run start + the counts of the warps before
#pragma unroll
    for (int w = 0; w < kOsWarps; ++w) {
      const u32 c = s_cnt[w][d];
      s_cnt[w][d] = at;
      at += c;
    }
  }
  __syncthreads();

  // stage the tile in digit order; the lowest digit bit of the staged word is the label
  {
    const u32* bases = s_cnt[warp];
#pragma unroll
    for (int r = 0; r < kOsItems; ++r) {
      if (wofs + r * 32 < valid) {
        const u32 d = (key[r] >> SHIFT) & 0xffu;
        const u32 rk = (r & 1) ? (rank2[r >> 1] >> 16) : (rank2[r >> 1] & 0xffffu);
        s_keys[bases[d] + rk] = (key[r] & ~(1u << SHIFT)) | (((labbits >> r) & 1u) << SHIFT);
      }
    }
  }

  // look back over the earlier tiles' counts of digit d
  if (digit_thread) {
    const int d = tid;
    ST excl = 0;
#ifdef TMB_K4_EXPERIMENT_NO_LOOKBACK
    if (false) {
#else
    if (tile > 0) {
#endif
      // The tiles of a wave reach this point together, so the nearest inclusive word is often several tiles back:
      // fetch a window of words per step (independent loads, one L2 round trip) and consume the ready prefix.
      long long j = (long long)tile - 1;  // nearest tile not yet added
      bool done = false;
      while (!done) {
        ST sw[TMB_K4_WINDOW];
#pragma unroll
        for (int q = 0; q < TMB_K4_WINDOW; ++q)
          sw[q] = j - q >= 0 ? ld_status(status + (size_t)(j - q) * kRadix + d) : (ST(2) << SH);
        int used = 0;
#pragma unroll
        for (int q = 0; q < TMB_K4_WINDOW; ++q) {
          const ST f = sw[q] >> SH;
          if (!done && used == q && f != 0) {
            excl += sw[q] & kValMask;
            used = q + 1;
            done = f == 2;
          }
        }
        j -= used;
#ifdef TMB_K4_BACKOFF_NS
        if (used == 0) __nanosleep(TMB_K4_BACKOFF_NS);
#endif
      }
      st_status(status + (size_t)tile * kRadix + d, (ST(2) << SH) | (excl + (ST)my_cnt));
    }
    // destinations stay below 2^32 (checked by the entry point), so 32-bit wrap-around arithmetic is exact
    s_dst[d] = my_gbase + (u32)excl - my_start;
  }
  __syncthreads();

  // coalesced write-out: consecutive threads hold consecutive positions of one digit run.  The digit of a staged
  // word: its upper seven digit bits, and the lowest from which of the two candidate runs the position lies in.
#pragma unroll 4
  for (int pos = tid; pos < valid; pos += kOsThreads) {
    const u32 w = s_keys[pos];
    const u32 dodd = ((w >> SHIFT) & 0xfeu) | 1u;
    const u32 low = (u32)pos >= s_start[dodd] ? 1u : 0u;
    const u32 dst = s_dst[dodd - 1u + low] + (u32)pos;
    keys_out[dst] = (w & ~(1u << SHIFT)) | (low << SHIFT);
    labs_out[dst] = (unsigned char)((w >> SHIFT) & 1u);
  }
}

// ---- scan over the sorted sequence ------------------------------------------------------------------
// run end at i  <=>  i == n-1 or score[i] != score[i+1] (float compare: -0.0 and 0.0 are one run, NaN - x != 0 is
// true: exactly what the reference's `torch.where(preds[1:] - preds[:-1])` sees)

// A thread's 16 consecutive items of a tile, loaded with 16-byte loads (the buffers are padded to whole tiles and the
// element after the tile is readable: it lies in the workspace and only matters when it belongs to the input).
struct Items16 {
  u32 labels;  // bit r: item r is a positive
  u32 ends;    // bit r: item r is a run end
};
__device__ __forceinline__ Items16 load_items16(const u32* __restrict__ keys, const unsigned char* __restrict__ labs,
                                                long long first, long long n, u32* keys_out = nullptr) {
  u32 k[kSortItems + 1];
  const uint4* kv = reinterpret_cast<const uint4*>(keys + first);
#pragma unroll
  for (int q = 0; q < kSortItems / 4; ++q) {
    const uint4 v = kv[q];
    k[4 * q] = v.x; k[4 * q + 1] = v.y; k[4 * q + 2] = v.z; k[4 * q + 3] = v.w;
  }
  k[kSortItems] = keys[first + kSortItems];
  const uint4 lv = *reinterpret_cast<const uint4*>(labs + first);
  const u32 lw[4] = {lv.x, lv.y, lv.z, lv.w};
  Items16 it{0u, 0u};
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    const long long i = first + r;
    const bool in = i < n;
    const u32 lab = (lw[r >> 2] >> (8 * (r & 3))) & 0xffu;
    // run end  <=>  last element, or score differs from the next one (float compare: -0.0 == 0.0, NaN != x)
    const bool end = in && (i == n - 1 || (key_to_float(k[r]) - key_to_float(k[r + 1])) != 0.0f);
    it.labels |= (u32)(in && lab) << r;
    it.ends |= (u32)end << r;
  }
  if (keys_out) {
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) keys_out[r] = k[r];
  }
  return it;
}

__global__ void __launch_bounds__(kSortThreads) k4_tile_sums(const u32* __restrict__ keys,
                                                             const unsigned char* __restrict__ labs, long long n,
                                                             u64* __restrict__ tile_pos, u64* __restrict__ tile_ends) {
  const long long first = (long long)blockIdx.x * kSortTile + (long long)threadIdx.x * kSortItems;
  const Items16 it = load_items16(keys, labs, first, n);
  u32 pos = __popc(it.labels), ends = __popc(it.ends);
  pos = __reduce_add_sync(0xffffffffu, pos);
  ends = __reduce_add_sync(0xffffffffu, ends);
  __shared__ u32 sp[kSortWarps], se[kSortWarps];
  if ((threadIdx.x & 31) == 0) {
    sp[threadIdx.x >> 5] = pos;
    se[threadIdx.x >> 5] = ends;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    u64 a = 0, b = 0;
    for (int w = 0; w < kSortWarps; ++w) {
      a += sp[w];
      b += se[w];
    }
    tile_pos[blockIdx.x] = a;
    tile_ends[blockIdx.x] = b;
  }
}

// single CTA: exclusive scan of both tile arrays in place; totals to out2 = {num_positive, num_distinct}
__global__ void __launch_bounds__(1024) k4_scan_tiles(u64* __restrict__ tile_pos, u64* __restrict__ tile_ends,
                                                      int ntiles, long long* __restrict__ out2) {
  __shared__ u64 wa[32], wb[32];
  __shared__ u64 ca, cb;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { ca = 0; cb = 0; }
  __syncthreads();
  for (int start = 0; start < ntiles; start += 1024) {
    int i = start + threadIdx.x;
    u64 va = i < ntiles ? tile_pos[i] : 0, vb = i < ntiles ? tile_ends[i] : 0;
    u64 xa = va, xb = vb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      u64 ya = __shfl_up_sync(0xffffffffu, xa, o), yb = __shfl_up_sync(0xffffffffu, xb, o);
      if (lane >= o) { xa += ya; xb += yb; }
    }
    if (lane == 31) { wa[warp] = xa; wb[warp] = xb; }
    __syncthreads();
    if (warp == 0) {
      u64 a = wa[lane], b = wb[lane], sa = a, sb = b;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        u64 ya = __shfl_up_sync(0xffffffffu, sa, o), yb = __shfl_up_sync(0xffffffffu, sb, o);
        if (lane >= o) { sa += ya; sb += yb; }
      }
      wa[lane] = sa - a;
      wb[lane] = sb - b;
    }
    __syncthreads();
    const u64 baseA = ca + wa[warp], baseB = cb + wb[warp];
    if (i < ntiles) {
      tile_pos[i] = baseA + xa - va;
      tile_ends[i] = baseB + xb - vb;
    }
    __syncthreads();
    if (threadIdx.x == 1023) { ca = baseA + xa; cb = baseB + xb; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out2[0] = (long long)ca;
    out2[1] = (long long)cb;
  }
}

// per tile: inclusive scans with the tile's base, emit at run ends.  A thread owns 16 consecutive items (16-byte
// loads); the tile's run ends are compacted in shared memory (local position, positives so far, score) so that the
// three output arrays are written with consecutive threads on consecutive rows.
__global__ void __launch_bounds__(kSortThreads) k4_emit(const u32* __restrict__ keys,
                                                        const unsigned char* __restrict__ labs, long long n,
                                                        const u64* __restrict__ tile_pos, const u64* __restrict__ tile_ends,
                                                        long long* __restrict__ fps, long long* __restrict__ tps,
                                                        float* __restrict__ thr) {
  __shared__ u32 wp[kSortWarps], we[kSortWarps];
  __shared__ unsigned short s_idx[kSortTile];
  __shared__ unsigned short s_cum[kSortTile];
  __shared__ float s_thr[kSortTile];
  __shared__ u32 s_nend;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long tbase = (long long)blockIdx.x * kSortTile;
  const int j0 = tid * kSortItems;
  u32 k[kSortItems];
  const Items16 it = load_items16(keys, labs, tbase + j0, n, k);
  const u32 tp = __popc(it.labels), te = __popc(it.ends);
  u32 xp = tp, xe = te;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u32 yp = __shfl_up_sync(0xffffffffu, xp, o), ye = __shfl_up_sync(0xffffffffu, xe, o);
    if (lane >= o) { xp += yp; xe += ye; }
  }
  if (lane == 31) { wp[warp] = xp; we[warp] = xe; }
  __syncthreads();
  u32 addp = 0, adde = 0;
  for (int w = 0; w < warp; ++w) { addp += wp[w]; adde += we[w]; }
  const u32 pos_before = addp + xp - tp;  // positives of this tile strictly before this thread's first item
  u32 slot = adde + xe - te;              // run ends of this tile before this thread
#pragma unroll
  for (int r = 0; r < kSortItems; ++r) {
    if ((it.ends >> r) & 1u) {
      s_idx[slot] = (unsigned short)(j0 + r);
      s_cum[slot] = (unsigned short)(pos_before + __popc(it.labels & (0xffffffffu >> (31 - r))));  // <= 4096 fits
      s_thr[slot] = key_to_float(k[r]);
      ++slot;
    }
  }
  if (tid == kSortThreads - 1) s_nend = slot;
  __syncthreads();
  const u32 nend = s_nend;
  const u64 pos0 = tile_pos[blockIdx.x], out0 = tile_ends[blockIdx.x];
  for (u32 e = tid; e < nend; e += kSortThreads) {
    const long long tpv = (long long)(pos0 + s_cum[e]);
    tps[out0 + e] = tpv;
    fps[out0 + e] = tbase + s_idx[e] + 1 - tpv;
    thr[out0 + e] = s_thr[e];
  }
}

// ---- scalar reductions over the sorted sequence: exact AUROC and average precision without the curve -----------
// With run ends k = 1..K (TP_k, FP_k cumulative, TP_0 = FP_0 = 0):
//   trapezoid ROC area * 2 * Npos * Nneg  =  sum_k (FP_k - FP_{k-1}) * (TP_k + TP_{k-1})          (roc.py:60-78 + auc)
//   average precision * Npos              =  sum_k (TP_k - TP_{k-1}) * TP_k / (TP_k + FP_k)       (average_precision.py:39-46)
// Every term needs the previous run end.  Inside a tile it is the neighbour in the compacted list; across tiles it
// comes from a max-scan (cumulative counts are monotone, so "latest" == "largest").

// per tile: positives, run ends, and (index + 1, positives up to it) of the tile's last run end (0, 0 if none)
__global__ void __launch_bounds__(kSortThreads) k4_tile_sums_last(const u32* __restrict__ keys,
                                                                  const unsigned char* __restrict__ labs, long long n,
                                                                  u64* __restrict__ tile_pos, u64* __restrict__ tile_ends,
                                                                  u64* __restrict__ tile_last_idx1,
                                                                  u64* __restrict__ tile_last_tp) {
  const long long base = (long long)blockIdx.x * kSortTile;
  const int j0 = threadIdx.x * kSortItems;  // local index of this thread's first item
  const Items16 it = load_items16(keys, labs, base + j0, n);
  u32 pos = __popc(it.labels), ends = __popc(it.ends);
  u32 last1 = it.ends ? (u32)j0 + (32u - __clz(it.ends)) : 0u;  // (local index of this thread's last run end) + 1
  pos = __reduce_add_sync(0xffffffffu, pos);
  ends = __reduce_add_sync(0xffffffffu, ends);
  last1 = __reduce_max_sync(0xffffffffu, last1);
  __shared__ u32 sp[kSortWarps], se[kSortWarps], sl[kSortWarps], sq[kSortWarps];
  __shared__ u32 s_last1;
  if ((threadIdx.x & 31) == 0) {
    sp[threadIdx.x >> 5] = pos;
    se[threadIdx.x >> 5] = ends;
    sl[threadIdx.x >> 5] = last1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    u32 m = 0;
    for (int w = 0; w < kSortWarps; ++w) m = sl[w] > m ? sl[w] : m;
    s_last1 = m;
  }
  __syncthreads();
  const u32 lastidx1 = s_last1;
  // positives of the tile at local indices < lastidx1: this thread's items r with j0 + r < lastidx1
  const int keep = (int)lastidx1 - j0;
  const u32 mask = keep <= 0 ? 0u : keep >= kSortItems ? 0xffffu : ((1u << keep) - 1u);
  u32 upto = __popc(it.labels & mask);
  upto = __reduce_add_sync(0xffffffffu, upto);
  if ((threadIdx.x & 31) == 0) sq[threadIdx.x >> 5] = upto;
  __syncthreads();
  if (threadIdx.x == 0) {
    u64 a = 0, b = 0, c = 0;
    for (int w = 0; w < kSortWarps; ++w) {
      a += sp[w];
      b += se[w];
      c += sq[w];
    }
    tile_pos[blockIdx.x] = a;
    tile_ends[blockIdx.x] = b;
    tile_last_idx1[blockIdx.x] = lastidx1 ? (u64)base + lastidx1 : 0;  // global index + 1
    tile_last_tp[blockIdx.x] = c;                                       // local; made global by the scan
  }
}

// single CTA: exclusive sum-scan of tile_pos / tile_ends, then exclusive max-scan of the last-run-end pair
__global__ void __launch_bounds__(1024) k4_scan_tiles_last(u64* __restrict__ tile_pos, u64* __restrict__ tile_ends,
                                                           u64* __restrict__ tile_last_idx1,
                                                           u64* __restrict__ tile_last_tp, int ntiles,
                                                           double* __restrict__ out4) {
  __shared__ u64 wa[32], wb[32], wc[32], wd[32];
  __shared__ u64 ca, cb, cc, cd;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { ca = 0; cb = 0; cc = 0; cd = 0; }
  __syncthreads();
  for (int start = 0; start < ntiles; start += 1024) {
    const int i = start + threadIdx.x;
    const u64 va = i < ntiles ? tile_pos[i] : 0, vb = i < ntiles ? tile_ends[i] : 0;
    u64 xa = va, xb = vb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const u64 ya = __shfl_up_sync(0xffffffffu, xa, o), yb = __shfl_up_sync(0xffffffffu, xb, o);
      if (lane >= o) { xa += ya; xb += yb; }
    }
    if (lane == 31) { wa[warp] = xa; wb[warp] = xb; }
    __syncthreads();
    if (warp == 0) {
      const u64 a = wa[lane], b = wb[lane];
      u64 sa = a, sb = b;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u64 ya = __shfl_up_sync(0xffffffffu, sa, o), yb = __shfl_up_sync(0xffffffffu, sb, o);
        if (lane >= o) { sa += ya; sb += yb; }
      }
      wa[lane] = sa - a;
      wb[lane] = sb - b;
    }
    __syncthreads();
    const u64 baseA = ca + wa[warp], baseB = cb + wb[warp];
    const u64 exclA = baseA + xa - va;  // positives before tile i
    // the tile's last run end in global terms (0 = the tile has none)
    const u64 vi = i < ntiles ? tile_last_idx1[i] : 0;
    const u64 vt = (i < ntiles && vi) ? exclA + tile_last_tp[i] : 0;
    u64 xc = vi, xd = vt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const u64 yc = __shfl_up_sync(0xffffffffu, xc, o), yd = __shfl_up_sync(0xffffffffu, xd, o);
      if (lane >= o) { xc = yc > xc ? yc : xc; xd = yd > xd ? yd : xd; }
    }
    if (lane == 31) { wc[warp] = xc; wd[warp] = xd; }
    __syncthreads();
    if (warp == 0) {
      u64 sc = wc[lane], sd = wd[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const u64 yc = __shfl_up_sync(0xffffffffu, sc, o), yd = __shfl_up_sync(0xffffffffu, sd, o);
        if (lane >= o) { sc = yc > sc ? yc : sc; sd = yd > sd ? yd : sd; }
      }
      // exclusive over warps
      const u64 pc = __shfl_up_sync(0xffffffffu, sc, 1), pd = __shfl_up_sync(0xffffffffu, sd, 1);
      wc[lane] = lane ? pc : 0;
      wd[lane] = lane ? pd : 0;
    }
    __syncthreads();
    // exclusive max within the warp, combined with earlier warps and earlier chunks
    u64 ec = __shfl_up_sync(0xffffffffu, xc, 1), ed = __shfl_up_sync(0xffffffffu, xd, 1);
    if (lane == 0) { ec = 0; ed = 0; }
    const u64 prevC = max(max(ec, wc[warp]), cc), prevD = max(max(ed, wd[warp]), cd);
    if (i < ntiles) {
      tile_pos[i] = exclA;
      tile_ends[i] = baseB + xb - vb;
      tile_last_idx1[i] = prevC;  // (index + 1) of the last run end before the tile, 0 if none
      tile_last_tp[i] = prevD;    // cumulative positives there
    }
    __syncthreads();
    if (threadIdx.x == 1023) {
      ca = baseA + xa;
      cb = baseB + xb;
      cc = max(prevC, xc);
      cd = max(prevD, xd);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    out4[2] = (double)ca;  // positives
    out4[3] = (double)cb;  // distinct scores
  }
}

// per tile: the two sums over the tile's run ends, written to tile_auc / tile_ap (deterministic: no float atomics).
// A thread walks the run ends among its 16 consecutive items with the running (TP, FP) in registers; what it needs
// from the threads before it — positives so far and the previous run end — comes from two block scans (a sum and a
// "latest" = max, cumulative counts being monotone).
__global__ void __launch_bounds__(kSortThreads) k4_reduce_tiles(const u32* __restrict__ keys,
                                                                const unsigned char* __restrict__ labs, long long n,
                                                                const u64* __restrict__ tile_pos,
                                                                const u64* __restrict__ tile_last_idx1,
                                                                const u64* __restrict__ tile_last_tp,
                                                                double* __restrict__ tile_auc,
                                                                double* __restrict__ tile_ap, u64 tp_off, u64 fp_off) {
  __shared__ u32 wp[kSortWarps], wl[kSortWarps], wt[kSortWarps];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long tbase = (long long)blockIdx.x * kSortTile;
  const int j0 = tid * kSortItems;
  const Items16 it = load_items16(keys, labs, tbase + j0, n);
  const u32 npos = __popc(it.labels);
  // this thread's last run end: (local index + 1, positives of the thread up to and including it); 0 = none
  const u32 hi = it.ends ? 32u - __clz(it.ends) : 0u;  // 1-based position of the highest set bit
  const u32 my_last1 = it.ends ? (u32)j0 + hi : 0u;
  const u32 my_last_pos = it.ends ? __popc(it.labels & (0xffffffffu >> (32u - hi))) : 0u;
  // inclusive scans over the warp: positives (sum); last run end and the tile-local TP there (max)
  u32 xp = npos, xl = my_last1;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u32 yp = __shfl_up_sync(0xffffffffu, xp, o), yl = __shfl_up_sync(0xffffffffu, xl, o);
    if (lane >= o) {
      xp += yp;
      xl = yl > xl ? yl : xl;
    }
  }
  if (lane == 31) {
    wp[warp] = xp;
    wl[warp] = xl;
  }
  __syncthreads();
  u32 addp = 0, prevl_w = 0;
  for (int w = 0; w < warp; ++w) {
    addp += wp[w];
    prevl_w = wl[w] > prevl_w ? wl[w] : prevl_w;
  }
  const u32 tp_before = addp + xp - npos;            // positives of the tile before this thread's first item
  const u32 my_last_tp = tp_before + my_last_pos;    // tile-local TP at this thread's last run end (if any)
  u32 xt = it.ends ? my_last_tp : 0u;                // max-scan: TP is monotone, so the latest is the largest
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const u32 yt = __shfl_up_sync(0xffffffffu, xt, o);
    if (lane >= o) xt = yt > xt ? yt : xt;
  }
  if (lane == 31) wt[warp] = xt;
  __syncthreads();
  u32 prevt_w = 0;
  for (int w = 0; w < warp; ++w) prevt_w = wt[w] > prevt_w ? wt[w] : prevt_w;
  // exclusive: the lanes before this one
  u32 pl = __shfl_up_sync(0xffffffffu, xl, 1), pt = __shfl_up_sync(0xffffffffu, xt, 1);
  if (lane == 0) {
    pl = 0;
    pt = 0;
  }
  const u32 prev_l = pl > prevl_w ? pl : prevl_w;  // local (index + 1) of the previous run end in this tile, 0 = none
  const u32 prev_t = pt > prevt_w ? pt : prevt_w;
  const u64 pos0 = tile_pos[blockIdx.x];
  u64 tpp, fpp;  // cumulative counts at the previous run end
  if (prev_l) {
    tpp = pos0 + prev_t;
    fpp = (u64)tbase + prev_l - tpp;
  } else {
    tpp = tile_last_tp[blockIdx.x];
    fpp = tile_last_idx1[blockIdx.x] - tpp;
  }
  // counts of a score range sorted elsewhere (distributed curve: everything ranked above this shard)
  tpp += tp_off;
  fpp += fp_off;
  double auc = 0.0, ap = 0.0;
  u32 e = it.ends;
  while (e) {
    const u32 r = __ffs(e) - 1;
    e &= e - 1;
    const u64 tpk = tp_off + pos0 + tp_before + __popc(it.labels & (0xffffffffu >> (31u - r)));
    const u64 idx1 = tp_off + fp_off + (u64)tbase + j0 + r + 1;  // TP_k + FP_k
    const u64 fpk = idx1 - tpk;
    auc += (double)(fpk - fpp) * (double)(tpk + tpp);
    if (tpk != tpp) ap += (double)(tpk - tpp) * ((double)tpk / (double)idx1);
    tpp = tpk;
    fpp = fpk;
  }
  // fixed-order block reduction
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    auc += __shfl_xor_sync(0xffffffffu, auc, o);
    ap += __shfl_xor_sync(0xffffffffu, ap, o);
  }
  __shared__ double ra[kSortWarps], rb[kSortWarps];
  if (lane == 0) {
    ra[warp] = auc;
    rb[warp] = ap;
  }
  __syncthreads();
  if (tid == 0) {
    double x = 0.0, y = 0.0;
    for (int w = 0; w < kSortWarps; ++w) {
      x += ra[w];
      y += rb[w];
    }
    tile_auc[blockIdx.x] = x;
    tile_ap[blockIdx.x] = y;
  }
}

// single CTA, fixed order: out4[0] = sum tile_auc, out4[1] = sum tile_ap
__global__ void __launch_bounds__(1024) k4_sum_tiles(const double* __restrict__ tile_auc,
                                                     const double* __restrict__ tile_ap, int ntiles,
                                                     double* __restrict__ out4) {
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < ntiles; i += 1024) {
    a += tile_auc[i];
    b += tile_ap[i];
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  __shared__ double ra[32], rb[32];
  if ((threadIdx.x & 31) == 0) { ra[threadIdx.x >> 5] = a; rb[threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double x = 0.0, y = 0.0;
    for (int w = 0; w < 32; ++w) { x += ra[w]; y += rb[w]; }
    out4[0] = x;
    out4[1] = y;
  }
}

struct K4Ws {
  u32* keys[2];
  unsigned char* labs[2];
  unsigned char* zeroed;   // status words of the four passes, digit histograms, tile counters: cleared per call
  size_t zeroed_bytes;
  void* status;            // [4][ntiles][256] status words (u32 or u64)
  u32* ghist;              // [4][256]
  u32* counters;           // [4] + padding
  u64* tile_pos;
  u64* tile_ends;
  u64* tile_last_idx1;
  u64* tile_last_tp;
  double* tile_auc;
  double* tile_ap;
};

static size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// inputs of 2^30 keys and more need 64-bit status words; TMB_K4_WIDE_STATUS=1 forces them (tests)
static bool k4_wide_status(long long n) {
  static const bool forced = [] {
    const char* e = getenv("TMB_K4_WIDE_STATUS");
    return e && e[0] == '1';
  }();
  return forced || n >= (1LL << 30);
}

static size_t k4_layout(long long n, unsigned char* base, K4Ws* ws) {
  const long long ntiles = (n + kSortTile - 1) / kSortTile;
  const size_t npad = (size_t)ntiles * kSortTile;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    unsigned char* p = base ? base + off : nullptr;
    off += align256(bytes);
    return p;
  };
  u32* k0 = (u32*)take(npad * 4);
  u32* k1 = (u32*)take(npad * 4 + 256);  // + the one key the scan kernels read past the last tile
  unsigned char* l0 = take(npad);
  unsigned char* l1 = take(npad);
  const size_t zero_begin = off;
  void* status = take((size_t)4 * ntiles * kRadix * (k4_wide_status(n) ? 8 : 4));
  u32* ghist = (u32*)take(4 * kRadix * 4);
  u32* counters = (u32*)take(256);
  const size_t zero_end = off;
  u64* tp = (u64*)take((size_t)ntiles * 8);
  u64* te = (u64*)take((size_t)ntiles * 8);
  u64* tli = (u64*)take((size_t)ntiles * 8);
  u64* tlt = (u64*)take((size_t)ntiles * 8);
  double* tau = (double*)take((size_t)ntiles * 8);
  double* tap = (double*)take((size_t)ntiles * 8);
  if (ws) {
    ws->tile_last_idx1 = tli; ws->tile_last_tp = tlt; ws->tile_auc = tau; ws->tile_ap = tap;
    ws->keys[0] = k0; ws->keys[1] = k1; ws->labs[0] = l0; ws->labs[1] = l1;
    ws->zeroed = base + zero_begin; ws->zeroed_bytes = zero_end - zero_begin;
    ws->status = status; ws->ghist = ghist; ws->counters = counters; ws->tile_pos = tp; ws->tile_ends = te;
  }
  return off;
}

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_exact_curve_workspace_bytes(int64_t n) {
  if (n < 1) return 256;
  return k4_layout(n, nullptr, nullptr);
}

template <typename TT, typename ST>
static void k4_launch_passes(const float* preds, const TT* target, long long n, long long pos_label, const K4Ws& ws,
                             int ntiles, cudaStream_t st) {
  ST* status = (ST*)ws.status;
  const size_t per_pass = (size_t)ntiles * kRadix;
  using U8 = unsigned char;
  k4_onesweep<0, TT, ST><<<ntiles, kOsThreads, 0, st>>>(nullptr, nullptr, preds, target, pos_label, n, ws.ghist,
                                                          status, ws.counters, ws.keys[0], ws.labs[0]);
  k4_onesweep<8, U8, ST><<<ntiles, kOsThreads, 0, st>>>(ws.keys[0], ws.labs[0], nullptr, nullptr, 0, n,
                                                          ws.ghist + kRadix, status + per_pass, ws.counters + 1,
                                                          ws.keys[1], ws.labs[1]);
  k4_onesweep<16, U8, ST><<<ntiles, kOsThreads, 0, st>>>(ws.keys[1], ws.labs[1], nullptr, nullptr, 0, n,
                                                           ws.ghist + 2 * kRadix, status + 2 * per_pass,
                                                           ws.counters + 2, ws.keys[0], ws.labs[0]);
  k4_onesweep<24, U8, ST><<<ntiles, kOsThreads, 0, st>>>(ws.keys[0], ws.labs[0], nullptr, nullptr, 0, n,
                                                           ws.ghist + 3 * kRadix, status + 3 * per_pass,
                                                           ws.counters + 3, ws.keys[1], ws.labs[1]);
}

// digit histograms + 4 onesweep passes; the sorted keys / labels end up in buffer pair 1
static int k4_sort(const float* preds, const void* target, int target_dtype, long long n, long long pos_label,
                   const K4Ws& ws, int ntiles, cudaStream_t st, int* cur_out) {
  TMB_CUDA_CHECK(cudaMemsetAsync(ws.zeroed, 0, ws.zeroed_bytes, st));
  {
    const long long want = ((n >> 2) + 511) / 512;
    const long long cap = (long long)sm_count() * 4;
    const int grid = (int)(want < 1 ? 1 : want < cap ? want : cap);
    k4_digit_hist<<<grid, 512, 0, st>>>(preds, n, ws.ghist);
  }
  const bool wide = k4_wide_status(n);
  TMB_DISPATCH_INT(target_dtype, TT, {
    if (wide) k4_launch_passes<TT, u64>(preds, (const TT*)target, n, pos_label, ws, ntiles, st);
    else k4_launch_passes<TT, u32>(preds, (const TT*)target, n, pos_label, ws, ntiles, st);
  })
  *cur_out = 1;
  return TMB_OK;
}

static int k4_check(const float* preds, const void* target, long long n, void* workspace, size_t workspace_bytes,
                    int* ntiles) {
  if (!preds || !target) return TMB_ERR_ARG;
  if (n >= (1LL << 32) - kSortTile) return TMB_ERR_TOO_LARGE;  // per-digit tile offsets are 32-bit
  if (workspace_bytes < tmb_exact_curve_workspace_bytes(n)) return TMB_ERR_ARG;
  if (!aligned32(preds) || !aligned32(target) || !aligned32(workspace)) return TMB_ERR_ALIGN;
  const long long ntiles_ll = (n + kSortTile - 1) / kSortTile;
  if (ntiles_ll > 0x7fffffff) return TMB_ERR_TOO_LARGE;
  *ntiles = (int)ntiles_ll;
  return TMB_OK;
}

extern "C" int tmb_exact_curve(const float* preds, const void* target, int target_dtype, int64_t n, int64_t pos_label,
                               int64_t* fps, int64_t* tps, float* thresholds, int64_t* counts2, void* workspace,
                               size_t workspace_bytes, void* stream) {
  if (n < 0 || !counts2 || !workspace) return TMB_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    TMB_CUDA_CHECK(cudaMemsetAsync(counts2, 0, 16, st));
    return TMB_OK;
  }
  if (!fps || !tps || !thresholds) return TMB_ERR_ARG;
  int ntiles = 0;
  int rc = k4_check(preds, target, n, workspace, workspace_bytes, &ntiles);
  if (rc != TMB_OK) return rc;
  K4Ws ws;
  k4_layout(n, (unsigned char*)workspace, &ws);
  int cur = 0;
  rc = k4_sort(preds, target, target_dtype, n, pos_label, ws, ntiles, st, &cur);
  if (rc != TMB_OK) return rc;
  k4_tile_sums<<<ntiles, kSortThreads, 0, st>>>(ws.keys[cur], ws.labs[cur], n, ws.tile_pos, ws.tile_ends);
  k4_scan_tiles<<<1, 1024, 0, st>>>(ws.tile_pos, ws.tile_ends, ntiles, (long long*)counts2);
  k4_emit<<<ntiles, kSortThreads, 0, st>>>(ws.keys[cur], ws.labs[cur], n, ws.tile_pos, ws.tile_ends,
                                           (long long*)fps, (long long*)tps, thresholds);
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}

extern "C" int tmb_exact_auroc_ap_offset(const float* preds, const void* target, int target_dtype, int64_t n,
                                         int64_t pos_label, int64_t tp_offset, int64_t fp_offset, double* out4,
                                         void* workspace, size_t workspace_bytes, void* stream);

extern "C" int tmb_exact_auroc_ap(const float* preds, const void* target, int target_dtype, int64_t n,
                                  int64_t pos_label, double* out4, void* workspace, size_t workspace_bytes,
                                  void* stream) {
  return tmb_exact_auroc_ap_offset(preds, target, target_dtype, n, pos_label, 0, 0, out4, workspace, workspace_bytes,
                                   stream);
}

extern "C" int tmb_exact_auroc_ap_offset(const float* preds, const void* target, int target_dtype, int64_t n,
                                         int64_t pos_label, int64_t tp_offset, int64_t fp_offset, double* out4,
                                         void* workspace, size_t workspace_bytes, void* stream) {
  if (tp_offset < 0 || fp_offset < 0) return TMB_ERR_ARG;
  if (n < 0 || !out4 || !workspace) return TMB_ERR_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (n == 0) {
    TMB_CUDA_CHECK(cudaMemsetAsync(out4, 0, 32, st));
    return TMB_OK;
  }
  int ntiles = 0;
  int rc = k4_check(preds, target, n, workspace, workspace_bytes, &ntiles);
  if (rc != TMB_OK) return rc;
  K4Ws ws;
  k4_layout(n, (unsigned char*)workspace, &ws);
  int cur = 0;
  rc = k4_sort(preds, target, target_dtype, n, pos_label, ws, ntiles, st, &cur);
  if (rc != TMB_OK) return rc;
  k4_tile_sums_last<<<ntiles, kSortThreads, 0, st>>>(ws.keys[cur], ws.labs[cur], n, ws.tile_pos, ws.tile_ends,
                                                     ws.tile_last_idx1, ws.tile_last_tp);
  k4_scan_tiles_last<<<1, 1024, 0, st>>>(ws.tile_pos, ws.tile_ends, ws.tile_last_idx1, ws.tile_last_tp, ntiles, out4);
  k4_reduce_tiles<<<ntiles, kSortThreads, 0, st>>>(ws.keys[cur], ws.labs[cur], n, ws.tile_pos, ws.tile_last_idx1,
                                                   ws.tile_last_tp, ws.tile_auc, ws.tile_ap, (u64)tp_offset,
                                                   (u64)fp_offset);
  k4_sum_tiles<<<1, 1024, 0, st>>>(ws.tile_auc, ws.tile_ap, ntiles, out4);
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}
