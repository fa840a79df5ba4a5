// K1, large-C path: confusion matrix through a one-pass bucket partition (contract: tmb_confmat_update_ws in
// include/tm_b200.h).
//
// One 64-bit RED per sample into a 10^6-bin table tops out at ~190 G increments/s on B200 whatever the kernel
// around it looks like (measured: L2 tag / die-to-die fabric bound, profiles/r1_confmat_ncu_raw.csv,
// gpurun_out/red_bench.log) — 46 % of what the HBM stream of (pred, target) pairs could feed.  Shared-memory
// atomics run >1500 G/s chip-wide, so the increments are moved there:
//
//   phase A  stream the batch once; key = target*C + pred; bucket = key % NB, local bin = key / NB (< 4096).
//            Every tile ranks its samples per bucket with shared-memory atomics and appends the 12-bit local bins
//            as u16, in coalesced runs, to queue segments that belong to this CTA alone (one segment per bucket per
//            CTA, fill counts kept in shared memory): no global atomics — a shared cursor per bucket serialises
//            thousands of reservations on one L2 address and was measured to dominate the kernel.
//            2 B/sample written, L2-resident for the batch sizes this path is used for.
//   phase B  a few CTAs per bucket: histogram the bucket's segments into 4096 shared-memory bins, then add the
//            non-zero bins to the int64 state (128-byte runs).
//
// bucket = key % NB interleaves neighbouring cells over different buckets, so a heavy class (a whole row or column of
// the matrix) spreads evenly.  Segments have a fixed capacity (the caller sizes the workspace, 2x the uniform share
// by default); entries that do not fit are added straight to the state with a global RED — never lost, just slower.
#include "common.cuh"

namespace tmb {

constexpr int kPartThreads = 512;
constexpr int kPartItems = 16;                              // samples per thread per tile (4 vectors of 4)
constexpr int kPartTile = kPartThreads * kPartItems;        // 8192
constexpr int kPartCtasPerSm = 2;
constexpr int kScanPer = 2;                                 // buckets scanned per thread (kMaxBuckets / kPartThreads)
constexpr int kBinsPerBucket = 4096;
constexpr int kMaxBuckets = 1024;
constexpr int kHistThreads = 512;
constexpr int kHistSplit = 4;                               // CTAs per bucket in phase B

struct PartParams {
  const void* preds;
  const void* target;
  long long n;
  long long C;
  long long CC;
  u32 nb;          // number of buckets
  u32 magic;       // ceil(2^32 / nb): key / nb == __umulhi(key, magic) for key < 2^22, nb <= 1024
  u32 lognb;       // streaming phase A: nb is a power of two
  u32 cap;         // streaming phase A: FIFO entries per bucket
  int has_ignore;
  long long ignore;
  u32 G;           // CTAs of phase A == segments per bucket
  u32* fill;       // [nb][G] entries each CTA produced for each bucket (may run past segcap)
  unsigned short* queue;  // [nb][G][segcap]
  u32 segcap;
  u64* confmat;
  int* status;
};

// Cell -> (bucket, local bin).  Groups of 16 consecutive cells (128 B of the int64 state) stay together so that phase
// B flushes 128-byte runs; successive groups go round-robin over the buckets, so a heavy row or column of the matrix
// (1000 cells = 63 groups) still spreads over many buckets.
//   group = key / 16,  bucket = group % nb,  local bin = (group / nb) * 16 + key % 16   (< 4096)
__device__ __forceinline__ u32 pack_slot(u32 key, const PartParams& P) {
  const u32 group = key >> 4;
  const u32 q = __umulhi(group, P.magic);  // group / nb (nb >= 2, group < 2^18)
  const u32 bucket = group - q * P.nb;
  return (bucket << 12) | (q << 4) | (key & 15u);
}
// inverse: flat cell index of (bucket, local bin)
__device__ __forceinline__ u64 cell_of(u32 bucket, u32 lbin, u32 nb) {
  return ((u64)(lbin >> 4) * nb + bucket) * 16 + (lbin & 15u);
}

// Slot of one sample, or kNoSlot when it is ignored / out of range.  Written without branches: everything is
// computed and then selected, so the sixteen samples of a thread form one straight-line block.
constexpr u32 kNoSlot = (u32)kMaxBuckets << 12;  // "bucket" kMaxBuckets: a dummy counter and a dummy staging cell

template <typename PT, typename TT, bool HAS_IGNORE>
__device__ __forceinline__ u32 slot_of(PT pv, TT tv, const PartParams& P, int& bad) {
  const long long tl = as_i64(tv), pl = as_i64(pv);
  const long long k = tl * P.C + pl;
  const bool in_range = (u64)k < (u64)P.CC;
  // fused validation bits (TMB_STATUS_*): key range, target range, pred range
  int flags = (in_range ? 0 : TMB_STATUS_LABEL_RANGE) | ((u64)tl >= (u64)P.C ? TMB_STATUS_TARGET_VALUES : 0) |
              ((u64)pl >= (u64)P.C ? TMB_STATUS_PREDS_VALUES : 0);
  bool counted = in_range;
  if constexpr (HAS_IGNORE) {
    const bool ign = tl == P.ignore;
    flags = ign ? 0 : flags;
    counted = in_range & !ign;
  }
  bad |= flags;
  const u32 packed = pack_slot((u32)k, P);
  return counted ? packed : kNoSlot;
}

template <typename PT, typename TT, bool HAS_IGNORE>
__global__ void __launch_bounds__(kPartThreads, kPartCtasPerSm) confmat_partition_kernel(const PartParams P) {
  __shared__ u32 s_cnt[kMaxBuckets + 1]; // per-bucket count of this tile (+1: dummy bucket of uncounted samples)
  __shared__ unsigned short s_start[kMaxBuckets + 2];  // exclusive prefix: first staged slot of each bucket
  __shared__ u32 s_delta[kMaxBuckets];   // (offset in this CTA's segment of the bucket) - s_start
  __shared__ u32 s_fill[kMaxBuckets];    // entries this CTA has appended to each bucket so far
  __shared__ u32 s_stage[kPartTile + 1]; // the tile in bucket order (+1: dummy cell)
  __shared__ u32 s_warp[32];
  __shared__ u32 s_total;
  static_assert(kMaxBuckets == kScanPer * kPartThreads, "each thread scans kScanPer buckets");
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long n = P.n;
  const long long ntiles = (n + kPartTile - 1) / kPartTile;
  int bad = 0;
#pragma unroll
  for (int k = 0; k < kScanPer; ++k) {
    s_cnt[kScanPer * tid + k] = 0;
    s_fill[kScanPer * tid + k] = 0;
  }
  __syncthreads();

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long base = tile * kPartTile;
    u32 slot[kPartItems];            // kNoSlot = nothing to count
    unsigned short rank[kPartItems];
    // ---- load 4 vectors of 4 samples per thread, rank every counted sample inside its bucket
    if (base + kPartTile <= n) {
      // two vectors in flight per thread: the next one loads while this one is ranked
      const PT* pp = preds + base + (long long)tid * 4;
      const TT* tp = target + base + (long long)tid * 4;
      Vec4<PT> pn = load4(pp);
      Vec4<TT> tn = load4(tp);
#pragma unroll
      for (int v = 0; v < kPartItems / 4; ++v) {
        const Vec4<PT> pc = pn;
        const Vec4<TT> tc = tn;
        if (v + 1 < kPartItems / 4) {
          pn = load4(pp + (long long)(v + 1) * kPartThreads * 4);
          tn = load4(tp + (long long)(v + 1) * kPartThreads * 4);
        }
        u32 r4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {  // the four atomics are in flight together; results are consumed below
          const u32 sl = slot_of<PT, TT, HAS_IGNORE>(pc.v[j], tc.v[j], P, bad);
          slot[v * 4 + j] = sl;
          r4[j] = atomicAdd(&s_cnt[sl >> 12], 1u);  // uncounted samples bump the dummy counter
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) rank[v * 4 + j] = (unsigned short)r4[j];
      }
    } else {
      // last, ragged tile
#pragma unroll
      for (int i = 0; i < kPartItems; ++i) {
        const long long e = base + ((long long)(i >> 2) * kPartThreads + tid) * 4 + (i & 3);
        u32 sl = kNoSlot;
        if (e < n) sl = slot_of<PT, TT, HAS_IGNORE>(preds[e], target[e], P, bad);
        slot[i] = sl;
        rank[i] = (unsigned short)atomicAdd(&s_cnt[sl >> 12], 1u);
      }
    }
    __syncthreads();
    // ---- exclusive scan of the bucket counts (kScanPer buckets per thread); advance this CTA's segment fills
    u32 c[kScanPer], gq[kScanPer];
    u32 mine = 0;
#pragma unroll
    for (int k = 0; k < kScanPer; ++k) {
      c[k] = s_cnt[kScanPer * tid + k];
      gq[k] = s_fill[kScanPer * tid + k];
      s_fill[kScanPer * tid + k] = gq[k] + c[k];
      mine += c[k];
      s_cnt[kScanPer * tid + k] = 0;  // ready for the next tile (only this thread reads these slots here)
    }
    u32 x = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      u32 y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) s_warp[warp] = x;
    __syncthreads();
    u32 wbase = 0;
#pragma unroll
    for (int w = 0; w < kPartThreads / 32; ++w) wbase += (w < warp) ? s_warp[w] : 0;
    u32 run = wbase + x - mine;
#pragma unroll
    for (int k = 0; k < kScanPer; ++k) {
      s_start[kScanPer * tid + k] = (unsigned short)run;
      s_delta[kScanPer * tid + k] = gq[k] - run;
      run += c[k];
    }
    if (tid == kPartThreads - 1) {
      s_total = run;
      s_start[kMaxBuckets] = (unsigned short)kPartTile;  // uncounted samples are staged onto the dummy cell
      s_cnt[kMaxBuckets] = 0;
    }
    __syncthreads();
    // ---- stage the tile in bucket order
#pragma unroll
    for (int i = 0; i < kPartItems; ++i) {
      const u32 bk = slot[i] >> 12;
      const u32 pos = bk == (u32)kMaxBuckets ? (u32)kPartTile : s_start[bk] + rank[i];
      s_stage[pos] = slot[i];
    }
    __syncthreads();
    // ---- write out: consecutive threads hold consecutive slots of the same bucket's run.  The next tile touches
    // s_stage / s_start / s_delta only after its own barriers, which every thread reaches after this loop.
    const u32 ntile = s_total;
    unsigned short* const myq = P.queue + (size_t)blockIdx.x * P.segcap;
    const u32 bstride = P.G * P.segcap;  // nb * G * segcap < 2^31 entries (entry point clamps segcap)
    auto emit = [&](u32 pos, u32 sl, u32 delta) {
      const u32 b = sl >> 12, v = sl & 0xfffu;
      const u32 off = delta + pos;
      if (off < P.segcap) {
        myq[b * bstride + off] = (unsigned short)v;
      } else {
        // segment full (a bucket far heavier than its share, typically one dominant cell): count directly, but let
        // the lanes that hold the same cell merge first so a hot cell costs one RED per warp, not one per sample
        const unsigned peers = __match_any_sync(__activemask(), sl);
        if ((threadIdx.x & 31) == (u32)(__ffs(peers) - 1))
          atomicAdd(P.confmat + cell_of(b, v, P.nb), (u64)__popc(peers));
      }
    };
    if (ntile == kPartTile) {  // the common case: every slot of the tile is occupied; batch the shared-memory reads
#pragma unroll
      for (int i0 = 0; i0 < kPartItems; i0 += 4) {
        u32 sl[4], dl[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) sl[i] = s_stage[tid + (i0 + i) * kPartThreads];
#pragma unroll
        for (int i = 0; i < 4; ++i) dl[i] = s_delta[sl[i] >> 12];
#pragma unroll
        for (int i = 0; i < 4; ++i) emit(tid + (i0 + i) * kPartThreads, sl[i], dl[i]);
      }
    } else {
      for (u32 pos = tid; pos < ntile; pos += kPartThreads) {
        const u32 sl = s_stage[pos];
        emit(pos, sl, s_delta[sl >> 12]);
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < kScanPer; ++k) {
    const u32 b = kScanPer * tid + k;
    if (b < P.nb) P.fill[(size_t)b * P.G + blockIdx.x] = s_fill[b];
  }
  if (bad) report_status(P.status, bad);
}

// ------------------------------------------------------------------------------------------------------------
// Phase A, streaming variant: per-bucket FIFOs in shared memory instead of rank / scan / stage.
//   A sample reserves a FIFO slot with one shared-memory atomic on its bucket's counter and stores its 12-bit local
//   bin there; nothing is kept in registers across a barrier.  After every tile one thread per bucket copies the
//   bucket's full 8-entry chunks (16-byte vector stores) to this CTA's segment of the bucket and moves the < 8
//   leftover entries to the front of the FIFO, so segments only ever grow by aligned chunks.  An entry that finds
//   its FIFO full (a bucket far heavier than the others) is counted with a direct RED, lanes merged first.
//   Here nb is a power of two: bucket = group & (nb - 1), local bin = ((group >> lognb) << 4) | (key & 15).
// ------------------------------------------------------------------------------------------------------------
constexpr int kFifoThreads = 512;
constexpr int kFifoItems = 16;
constexpr int kFifoTile = kFifoThreads * kFifoItems;
constexpr int kFifoCtasPerSm = 2;
constexpr int kFifoEntries = 24576;  // u16 FIFO entries per CTA, split evenly over the buckets

template <typename PT, typename TT, bool HAS_IGNORE>
__global__ void __launch_bounds__(kFifoThreads, kFifoCtasPerSm) confmat_fifo_partition_kernel(const PartParams P) {
  extern __shared__ uint4 smem_dyn[];
  u32* const s_cnt = reinterpret_cast<u32*>(smem_dyn);               // [nb] entries waiting in each FIFO
  u32* const s_fill = s_cnt + kMaxBuckets;                           // [nb] entries appended to each segment so far
  unsigned short* const s_fifo = reinterpret_cast<unsigned short*>(s_fill + kMaxBuckets);  // [nb][cap]
  const int tid = threadIdx.x;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long n = P.n;
  const long long ntiles = (n + kFifoTile - 1) / kFifoTile;
  const u32 nb = P.nb, lognb = P.lognb, cap = P.cap, nbmask = nb - 1;
  const u32 C32 = (u32)P.C;
  int bad = 0;
  const u64 pol_stream = l2_policy_evict_first(), pol_keep = l2_policy_evict_last();
  for (u32 b = tid; b < nb; b += kFifoThreads) {
    s_cnt[b] = 0;
    s_fill[b] = 0;
  }
  __syncthreads();

  auto push = [&](PT pv, TT tv) {
    const long long tl = as_i64(tv), pl = as_i64(pv);
    bool counted = true;
    if constexpr (HAS_IGNORE) counted = tl != P.ignore;
    u32 key = (u32)tl * C32 + (u32)pl;
    if (((u64)tl >= (u64)P.C) | ((u64)pl >= (u64)P.C)) {  // rare: the reference still counts a key that aliases
      if (counted) {
        const long long k = tl * P.C + pl;
        const bool in_range = (u64)k < (u64)P.CC;
        bad |= (in_range ? 0 : TMB_STATUS_LABEL_RANGE) | ((u64)tl >= (u64)P.C ? TMB_STATUS_TARGET_VALUES : 0) |
               ((u64)pl >= (u64)P.C ? TMB_STATUS_PREDS_VALUES : 0);
        counted = in_range;
        key = (u32)k;
      }
    }
    if (counted) {
      const u32 group = key >> 4;
      const u32 b = group & nbmask;
      const u32 lbin = ((group >> lognb) << 4) | (key & 15u);
      const u32 pos = atomicAdd(&s_cnt[b], 1u);
      if (pos < cap) {
        s_fifo[b * cap + pos] = (unsigned short)lbin;
      } else {
        const unsigned peers = __match_any_sync(__activemask(), key);
        if ((threadIdx.x & 31) == (u32)(__ffs(peers) - 1)) atomicAdd(P.confmat + key, (u64)__popc(peers));
      }
    }
  };

  unsigned short* const myq = P.queue + (size_t)blockIdx.x * P.segcap;
  const u32 bstride = P.G * P.segcap;
  // copy the full chunks of every FIFO to the segments; `final` also writes the partial chunk
  auto flush = [&](bool final) {
    for (u32 b = tid; b < nb; b += kFifoThreads) {
      u32 c = s_cnt[b];
      c = c < cap ? c : cap;
      u32 off = s_fill[b];
      const uint4* src = reinterpret_cast<const uint4*>(s_fifo + b * cap);
      const u32 nchunks = final ? (c + 7) >> 3 : c >> 3;
      unsigned short* seg = myq + b * bstride;
      for (u32 k = 0; k < nchunks; ++k) {
        const uint4 w = src[k];
        const u32 cnt = (k * 8 + 8 <= c) ? 8u : (c - k * 8);
        if (off + 8 <= P.segcap) {
          store16_hint(seg + off, w, pol_keep);
          off += cnt;
        } else {  // segment full: this CTA saw far more than its share of the bucket
          const u32 e[4] = {w.x, w.y, w.z, w.w};
          for (u32 i = 0; i < cnt; ++i) {
            const u32 lbin = (e[i >> 1] >> ((i & 1) * 16)) & 0xffffu;
            atomicAdd(P.confmat + cell_of(b, lbin, nb), 1ULL);
          }
        }
      }
      if (!final) {
        if (nchunks && (c & 7u)) reinterpret_cast<uint4*>(s_fifo + b * cap)[0] = src[nchunks];  // leftovers to the front
        s_cnt[b] = c & 7u;
      }
      s_fill[b] = off;
    }
  };

  Vec4<PT> pn;
  Vec4<TT> tn;
  long long tile = blockIdx.x;
  if (tile < ntiles && tile * kFifoTile + kFifoTile <= n) {
    pn = load4_hint(preds + tile * kFifoTile + (long long)tid * 4, pol_stream);
    tn = load4_hint(target + tile * kFifoTile + (long long)tid * 4, pol_stream);
  }
  for (; tile < ntiles; tile += gridDim.x) {
    const long long base = tile * kFifoTile;
    if (base + kFifoTile <= n) {
      const PT* pp = preds + base + (long long)tid * 4;
      const TT* tp = target + base + (long long)tid * 4;
#pragma unroll
      for (int v = 0; v < kFifoItems / 4; ++v) {
        const Vec4<PT> pc = pn;
        const Vec4<TT> tc = tn;
        if (v + 1 < kFifoItems / 4) {
          pn = load4_hint(pp + (long long)(v + 1) * kFifoThreads * 4, pol_stream);
          tn = load4_hint(tp + (long long)(v + 1) * kFifoThreads * 4, pol_stream);
        } else {
          // first vector of this CTA's next tile: in flight while the FIFOs are flushed
          const long long nbase = (tile + gridDim.x) * kFifoTile;
          if (nbase + kFifoTile <= n) {
            pn = load4_hint(preds + nbase + (long long)tid * 4, pol_stream);
            tn = load4_hint(target + nbase + (long long)tid * 4, pol_stream);
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) push(pc.v[j], tc.v[j]);
      }
    } else {
      for (long long e = base + tid; e < n; e += kFifoThreads) push(preds[e], target[e]);
    }
    __syncthreads();
    flush(false);
    __syncthreads();
  }
  flush(true);
  for (u32 b = tid; b < nb; b += kFifoThreads) P.fill[(size_t)b * P.G + blockIdx.x] = s_fill[b];
  if (bad) report_status(P.status, bad);
}

__global__ void __launch_bounds__(kHistThreads) confmat_bucket_hist_kernel(const PartParams P) {
  __shared__ u32 h[kBinsPerBucket];
  const u32 b = blockIdx.x;
  const u64 pol_done = l2_policy_evict_first();  // the queue is dead once read
  for (int i = threadIdx.x; i < kBinsPerBucket; i += kHistThreads) h[i] = 0;
  __syncthreads();
  auto count8 = [&](const uint4& w) {
    atomicAdd(&h[w.x & 0xffff], 1u); atomicAdd(&h[w.x >> 16], 1u);
    atomicAdd(&h[w.y & 0xffff], 1u); atomicAdd(&h[w.y >> 16], 1u);
    atomicAdd(&h[w.z & 0xffff], 1u); atomicAdd(&h[w.z >> 16], 1u);
    atomicAdd(&h[w.w & 0xffff], 1u); atomicAdd(&h[w.w >> 16], 1u);
  };
  // the bucket's G segments are dealt round-robin to the warps of the kHistSplit CTAs that share the bucket; a
  // segment is a few hundred entries, i.e. about one 16-byte vector per lane
  const u32 lane = threadIdx.x & 31;
  const u32 wid = blockIdx.y * (kHistThreads / 32) + (threadIdx.x >> 5);
  const u32 nw = kHistSplit * (kHistThreads / 32);
  const u32* fill_b = P.fill + (size_t)b * P.G;
  u32 next_fill = wid < P.G ? fill_b[wid] : 0;
  for (u32 seg = wid; seg < P.G; seg += nw) {
    const u32 produced = next_fill;
    if (seg + nw < P.G) next_fill = fill_b[seg + nw];  // overlaps this segment's work
    const u32 len = produced < P.segcap ? produced : P.segcap;
    const unsigned short* q = P.queue + ((size_t)b * P.G + seg) * P.segcap;  // 16-byte aligned (segcap % 8 == 0)
    const uint4* qv = reinterpret_cast<const uint4*>(q);
    const u32 nvec = len >> 3;
    for (u32 i = lane; i < nvec; i += 32) count8(load16_hint(qv + i, pol_done));
    for (u32 k = (nvec << 3) + lane; k < len; k += 32) atomicAdd(&h[q[k]], 1u);
  }
  __syncthreads();
  for (u32 lbin = threadIdx.x; lbin < kBinsPerBucket; lbin += kHistThreads) {
    const u32 c = h[lbin];
    if (c) atomicAdd(P.confmat + cell_of(b, lbin, P.nb), (u64)c);
  }
}

static int part_mode() {
  static const int m = getenv("TMB_PARTA") ? atoi(getenv("TMB_PARTA")) : 1;
  return m;
}
static u32 part_ctas() { return (u32)sm_count() * (part_mode() ? kFifoCtasPerSm : kPartCtasPerSm); }
// buckets for a CC-cell matrix; 0 = not served by the partition path
static long long part_buckets(long long CC) {
  const long long need = (((CC + 15) >> 4) + (kBinsPerBucket / 16) - 1) / (kBinsPerBucket / 16);
  if (need > kMaxBuckets) return 0;
  if (!part_mode() || need < 2) return need;
  long long nb = 256;  // few buckets mean many lanes of a warp on one counter
  while (nb < need) nb <<= 1;
  return nb;
}
static size_t part_header_bytes(u32 nb, u32 G) { return ((size_t)nb * G * 4 + 255) & ~(size_t)255; }

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_confmat_workspace_bytes(int64_t n, int64_t num_classes) {
  if (n < 1 || num_classes < 1) return 0;
  const long long CC = (long long)num_classes * num_classes;
  const long long nb = part_buckets(CC);
  if (nb == 0) return 0;  // not served by the partition path
  const long long G = part_ctas();
  // per (bucket, CTA) segment: twice the uniform share plus slack, rounded to 8 entries (16 bytes)
  long long segcap = 2 * ((n + nb * G - 1) / (nb * G)) + 64;
  segcap = (segcap + 7) & ~7LL;
  return part_header_bytes((u32)nb, (u32)G) + (size_t)nb * (size_t)G * (size_t)segcap * 2;
}

extern "C" int tmb_confmat_update_ws(const void* preds, int preds_dtype, const void* target, int target_dtype,
                                     int64_t n, int64_t num_classes, int has_ignore, int64_t ignore_index,
                                     int64_t* confmat, int32_t* status, void* workspace, size_t workspace_bytes,
                                     void* stream) {
  if (n < 0 || num_classes < 1 || !confmat || !status) return TMB_ERR_ARG;
  if (n == 0) return TMB_OK;
  if (!preds || !target || !workspace) return TMB_ERR_ARG;
  if (!aligned32(preds) || !aligned32(target) || !aligned32(workspace)) return TMB_ERR_ALIGN;
  const long long CC = (long long)num_classes * num_classes;
  const long long nb = part_buckets(CC);
  if (nb == 0 || n >= (1LL << 32)) return TMB_ERR_TOO_LARGE;
  if (nb < 2)  // everything fits one shared-memory histogram: the direct kernels are the right tool
    return tmb_confmat_update(preds, preds_dtype, target, target_dtype, n, num_classes, has_ignore, ignore_index,
                              confmat, status, 0, stream);
  const u32 G = part_ctas();
  const size_t header = part_header_bytes((u32)nb, G);
  if (workspace_bytes < header + (size_t)nb * G * 16) return TMB_ERR_ARG;
  size_t segcap = (workspace_bytes - header) / ((size_t)nb * G * 2);
  const size_t max_entries = ((size_t)1 << 31) - 8;  // 32-bit indexing inside the queue
  if (segcap * nb * G > max_entries) segcap = max_entries / ((size_t)nb * G);
  segcap &= ~(size_t)7;
  PartParams P{};
  P.preds = preds;
  P.target = target;
  P.n = n;
  P.C = num_classes;
  P.CC = CC;
  P.nb = (u32)nb;
  P.magic = (u32)(((1ULL << 32) + (u64)nb - 1) / (u64)nb);
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  P.G = G;
  P.fill = (u32*)workspace;
  P.queue = (unsigned short*)((unsigned char*)workspace + header);
  P.segcap = (u32)segcap;
  P.confmat = (u64*)confmat;
  P.status = status;
  cudaStream_t st = (cudaStream_t)stream;
  if (part_mode()) {
    u32 lg = 0;
    while ((1u << lg) < (u32)nb) ++lg;
    P.lognb = lg;
    P.cap = ((u32)kFifoEntries / (u32)nb) & ~7u;
    const size_t smem = (size_t)kMaxBuckets * 8 + (size_t)kFifoEntries * 2 + 16;
    TMB_DISPATCH_INT(preds_dtype, PT, {
      TMB_DISPATCH_INT(target_dtype, TT, {
        auto launch = [&](auto kern) -> int {
          TMB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
          kern<<<G, kFifoThreads, smem, st>>>(P);
          return TMB_OK;
        };
        const int rc = P.has_ignore ? launch(confmat_fifo_partition_kernel<PT, TT, true>)
                                    : launch(confmat_fifo_partition_kernel<PT, TT, false>);
        if (rc != TMB_OK) return rc;
        confmat_bucket_hist_kernel<<<dim3((unsigned)nb, kHistSplit), kHistThreads, 0, st>>>(P);
        TMB_CUDA_CHECK(cudaGetLastError());
        return TMB_OK;
      })
    })
    return TMB_ERR_ARG;
  }
  // always G CTAs: the ones without a tile still publish their (zero) fill counts
  TMB_DISPATCH_INT(preds_dtype, PT, {
    TMB_DISPATCH_INT(target_dtype, TT, {
      if (P.has_ignore) confmat_partition_kernel<PT, TT, true><<<G, kPartThreads, 0, st>>>(P);
      else confmat_partition_kernel<PT, TT, false><<<G, kPartThreads, 0, st>>>(P);
    })
  })
  confmat_bucket_hist_kernel<<<dim3((unsigned)nb, kHistSplit), kHistThreads, 0, st>>>(P);
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}
