// K1, large-C path: confusion matrix through a one-pass bucket partition (contract: tmb_confmat_update_ws in
// include/tm_b200.h).
//
// One 64-bit RED per sample into a 10^6-bin table tops out at ~190 G increments/s on B200 whatever the kernel
// around it looks like (measured: L2 tag / die-to-die fabric bound, profiles/r1_confmat_ncu_raw.csv,
// profiles/r1_red_bench.log) — 46 % of what the HBM stream of (pred, target) pairs could feed.  Shared-memory
// atomics run >1500 G/s chip-wide, so the increments are moved there:
//
//   phase A  stream the batch once; key = target*C + pred; groups of 16 neighbouring cells (128 B of the int64
//            state) are dealt round-robin over NB buckets (NB a power of two, >= 128):
//                group = key >> 4,  bucket = group & (NB-1),  local bin = ((group >> log2 NB) << 4) | (key & 15) < 8192
//            Every bucket has a FIFO in shared memory.  A sample reserves a FIFO slot with one shared-memory atomic
//            on the bucket's counter and stores its 13-bit local bin there (u16); nothing is kept in registers
//            across a barrier.  After every 8192-sample tile the full 8-entry chunks of each FIFO are appended
//            (16-byte vector stores, L2 "evict last") to a queue segment that belongs to this CTA and bucket alone
//            — no global atomics, a shared cursor per bucket serialises thousands of reservations on one L2
//            address — and the < 8 leftovers move to the front of the FIFO.  2 B/sample written.
//   phase B  one to four CTAs per bucket: histogram the bucket's segments into 8192 shared-memory bins, then add the
//            non-zero bins to the int64 state (128-byte runs of REDs).
//
// A heavy class (a whole row or column of the matrix) spreads evenly over the buckets because neighbouring groups
// go to different buckets.  An entry that finds its FIFO or its segment full (one cell far heavier than the rest)
// is merged with the other lanes holding the same cell and tallied in a 64-slot per-CTA table that is added to the
// state once at the end — never lost, and a dominant cell costs one RED per CTA.
#include <cmath>

#include "common.cuh"

namespace tmb {

constexpr int kBinsPerBucket = 8192;
constexpr int kMaxBuckets = 1024;
constexpr int kMinBuckets = 128;     // fewer buckets put many lanes of a warp on one FIFO counter
constexpr int kHistThreads = 512;
constexpr int kHistSplit = 4;        // most CTAs per bucket in phase B
constexpr int kHotSlots = 64;        // per-CTA table of cells whose FIFO overflowed (dominant cells)
constexpr int kTile = 8192;          // samples between FIFO flushes

struct PartParams {
  const void* preds;
  const void* target;
  long long n;
  long long C;
  long long CC;
  u32 nb;          // number of buckets (power of two)
  u32 lognb;
  u32 cap;         // FIFO entries per bucket (multiple of 8)
  int has_ignore;
  long long ignore;
  u32 G;           // CTAs of phase A == segments per bucket
  u32* fill;       // [nb][G] entries each CTA appended to its segment of each bucket
  unsigned short* queue;  // [nb][G][segcap]
  u32 segcap;      // multiple of 8
  int stream_evict_first;  // read the batch with the L2 "evict first" hint (queues small enough to stay in L2)
  u64* confmat;
  int* status;
};

// flat cell index of (bucket, local bin)
__device__ __forceinline__ u64 cell_of(u32 bucket, u32 lbin, u32 nb) {
  return ((u64)(lbin >> 4) * nb + bucket) * 16 + (lbin & 15u);
}

// ------------------------------------------------------------------------------------------------------------
// The per-CTA FIFO machinery shared by the two phase-A kernels.
// ------------------------------------------------------------------------------------------------------------
struct Fifos {
  u32* cnt;              // [nb] entries waiting in each FIFO (may run past cap: those entries went to the hot table)
  u32* fill;             // [nb] entries appended to this CTA's segment of each bucket so far
  unsigned short* fifo;  // [nb][cap]
  u32* hot_key;          // [kHotSlots]
  u32* hot_cnt;
  u32 nb, lognb, nbmask, cap, C32;
  u32 cnt_addr, fifo_addr;  // 32-bit shared-window addresses of cnt / fifo
  int bad;
  u64 pol_keep;

  __device__ __forceinline__ void init(const PartParams& P, int tid, int nthreads) {
    nb = P.nb;
    lognb = P.lognb;
    nbmask = P.nb - 1;
    cap = P.cap;
    C32 = (u32)P.C;
    cnt_addr = (u32)__cvta_generic_to_shared(cnt);
    fifo_addr = (u32)__cvta_generic_to_shared(fifo);
    bad = 0;
    pol_keep = l2_policy_evict_last();
    for (u32 b = tid; b < nb; b += nthreads) {
      cnt[b] = 0;
      fill[b] = 0;
    }
    if (tid < 32) cnt[nb + tid] = cap;  // dummy counters: a position taken from them never fits
    if (tid < kHotSlots) {
      hot_key[tid] = 0xffffffffu;
      hot_cnt[tid] = 0;
    }
  }

  // FIFO or segment full: a cell (or bucket) far heavier than the rest.  Lanes holding the same cell merge, then the
  // cell is tallied in the per-CTA table; a table slot taken by another cell falls through to a RED.
  __device__ __forceinline__ void overflow(const PartParams& P, u32 key) {
    const unsigned peers = __match_any_sync(__activemask(), key);
    if ((threadIdx.x & 31) == (u32)(__ffs(peers) - 1)) {
      const u32 h = (key * 2654435761u) >> 26;
      const u32 old = atomicCAS(&hot_key[h], 0xffffffffu, key);
      if (old == 0xffffffffu || old == key) atomicAdd(&hot_cnt[h], (u32)__popc(peers));
      else atomicAdd(P.confmat + key, (u64)__popc(peers));
    }
  }

  __device__ __forceinline__ void put(const PartParams& P, u32 key) {
    const u32 group = key >> 4;
    const u32 b = group & nbmask;
    const u32 pos = atomicAdd(&cnt[b], 1u);
    if (pos < cap) fifo[b * cap + pos] = (unsigned short)(((group >> lognb) << 4) | (key & 15u));
    else overflow(P, key);
  }

  // The general path for one (not ignored) sample whose target or pred is outside [0, C): the reference still counts
  // it when target*C + pred lands inside the matrix (it only fails when its bincount cannot be reshaped).
  __device__ __forceinline__ void put_checked(const PartParams& P, long long tl, long long pl) {
    const long long k = tl * P.C + pl;
    const bool in_range = (u64)k < (u64)P.CC;
    bad |= (in_range ? 0 : TMB_STATUS_LABEL_RANGE) | ((u64)tl >= (u64)P.C ? TMB_STATUS_TARGET_VALUES : 0) |
           ((u64)pl >= (u64)P.C ? TMB_STATUS_PREDS_VALUES : 0);
    if (in_range) put(P, (u32)k);
  }

  // K samples of one thread: keys first, then the K atomics back to back, then the K stores, so the atomics'
  // latencies overlap.  The common path has no branches: a sample that is not counted here (ignored, or outside
  // [0, C) and left to the rare path) bumps one of 32 per-lane dummy counters that start at `cap`, so its
  // "position" never fits and nothing is stored; everything rare is re-derived after the batch.
  template <int K, typename PT, typename TT, bool HAS_IGNORE>
  __device__ __forceinline__ void push(const PartParams& P, const PT (&pv)[K], const TT (&tv)[K]) {
    u32 key[K], pos[K];
    bool ok[K];
    bool any_rare = false;
    const u32 dummy = nb + (threadIdx.x & 31u);
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const long long tl = as_i64(tv[j]), pl = as_i64(pv[j]);
      // outside [0, C) <=> a high word is set or a low word is >= C (C <= 2048 here)
      const u32 lo_max = max((u32)tl, (u32)pl);
      u32 hi = 0;
      if constexpr (sizeof(TT) == 8) hi |= (u32)((u64)tl >> 32);
      if constexpr (sizeof(PT) == 8) hi |= (u32)((u64)pl >> 32);
      const bool oob = (lo_max >= C32) | (hi != 0);
      key[j] = (u32)tl * C32 + (u32)pl;
      if constexpr (HAS_IGNORE) {
        const bool counted = tl != P.ignore;
        ok[j] = counted & !oob;
        any_rare |= counted & oob;
      } else {
        ok[j] = !oob;
        any_rare |= oob;
      }
    }
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const u32 b = ok[j] ? ((key[j] >> 4) & nbmask) : dummy;
      pos[j] = atomicAdd(&cnt[b], 1u);
    }
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const u32 group = key[j] >> 4;
      const u32 b = group & nbmask;
      const bool fits = pos[j] < cap;  // never for a dummy counter
      if (fits) fifo[b * cap + pos[j]] = (unsigned short)(((group >> lognb) << 4) | (key[j] & 15u));
      any_rare |= ok[j] & !fits;
    }
    if (any_rare) {
#pragma unroll
      for (int j = 0; j < K; ++j) {
        const long long tl = as_i64(tv[j]), pl = as_i64(pv[j]);
        const bool counted = HAS_IGNORE ? tl != P.ignore : true;
        if (counted && !ok[j]) put_checked(P, tl, pl);
        if (ok[j] && pos[j] >= cap) overflow(P, key[j]);
      }
    }
  }

  // Append the full chunks of every FIFO to this CTA's segments (`final`: the partial chunk too) and move the
  // leftovers to the front.  THREADS / nb consecutive lanes share a bucket; chunk k goes to lane k % share.
  template <int THREADS>
  __device__ __forceinline__ void flush(const PartParams& P, int tid, bool final) {
    const u32 share_log = THREADS > nb ? (u32)(31 - __clz(THREADS)) - lognb : 0u;  // THREADS, nb: powers of two
    const u32 share = 1u << share_log;
    unsigned short* const myq = P.queue + (size_t)blockIdx.x * P.segcap;
    const u32 bstride = P.G * P.segcap;  // nb * G * segcap < 2^31 entries (entry point clamps segcap)
    for (u32 w = tid; w < (nb << share_log); w += THREADS) {  // whole warps enter and leave together
      const u32 b = w >> share_log, sub = w & (share - 1u);
      u32 c = cnt[b];
      c = c < cap ? c : cap;
      const u32 off0 = fill[b];
      __syncwarp();  // every sharer has read the counters before lane `sub == 0` rewrites them
      const uint4* src = reinterpret_cast<const uint4*>(fifo + b * cap);
      const u32 nchunks = final ? (c + 7) >> 3 : c >> 3;
      unsigned short* seg = myq + b * bstride;
      for (u32 k = sub; k < nchunks; k += share) {
        const uint4 wv = src[k];
        const u32 off = off0 + k * 8;
        if (off + 8 <= P.segcap) {
          store16_hint(seg + off, wv, pol_keep);
        } else {  // segment full: this CTA saw far more than its share of the bucket
          const u32 m = (k * 8 + 8 <= c) ? 8u : (c - k * 8);
          const u32 e[4] = {wv.x, wv.y, wv.z, wv.w};
          for (u32 i = 0; i < m; ++i) {
            const u32 lbin = (e[i >> 1] >> ((i & 1) * 16)) & 0xffffu;
            atomicAdd(P.confmat + cell_of(b, lbin, nb), 1ULL);
          }
        }
      }
      __syncwarp();  // every chunk of the bucket has been read before its leftovers move to the front
      if (sub == 0) {
        const u32 stored = final ? c : (c & ~7u);
        const u32 room = P.segcap - off0;  // multiple of 8
        if (!final) {
          if (nchunks && (c & 7u)) reinterpret_cast<uint4*>(fifo + b * cap)[0] = src[nchunks];
          cnt[b] = c & 7u;
        }
        fill[b] = off0 + (stored < room ? stored : room);
      }
    }
  }

  __device__ __forceinline__ void finish(const PartParams& P, int tid, int nthreads) {
    if (tid < kHotSlots && hot_cnt[tid]) atomicAdd(P.confmat + hot_key[tid], (u64)hot_cnt[tid]);
    for (u32 b = tid; b < nb; b += nthreads) P.fill[(size_t)b * P.G + blockIdx.x] = fill[b];
    if (bad) report_status(P.status, bad);
  }
};

// Dynamic shared memory of a phase-A kernel after its own leading bytes: [cnt | fill | fifo]
__device__ __forceinline__ void carve(Fifos& F, unsigned char* base, u32* hot_key, u32* hot_cnt) {
  F.cnt = reinterpret_cast<u32*>(base);
  F.fill = F.cnt + kMaxBuckets + 32;  // 32 per-lane dummy counters follow the bucket counters
  F.fifo = reinterpret_cast<unsigned short*>(F.fill + kMaxBuckets);
  F.hot_key = hot_key;
  F.hot_cnt = hot_cnt;
}
static size_t fifo_smem_bytes(u32 entries) { return (size_t)(2 * kMaxBuckets + 32) * 4 + (size_t)entries * 2 + 16; }

// ------------------------------------------------------------------------------------------------------------
// Phase A, register-fed: 512 threads, two CTAs per SM, each thread loads 4 vectors of 4 samples per tile with the
// next vector in flight while the current one is pushed.
// ------------------------------------------------------------------------------------------------------------
constexpr int kRegThreads = 512;
constexpr int kRegItems = kTile / kRegThreads;  // 16
constexpr int kRegCtasPerSm = 2;
constexpr int kRegFifoEntries = 36864;          // u16 FIFO entries per CTA, split evenly over the buckets

template <typename PT, typename TT, bool HAS_IGNORE>
__global__ void __launch_bounds__(kRegThreads, kRegCtasPerSm) confmat_partition_reg_kernel(const PartParams P) {
  extern __shared__ __align__(16) unsigned char smem_reg[];
  __shared__ u32 s_hot_key[kHotSlots], s_hot_cnt[kHotSlots];
  Fifos F;
  carve(F, smem_reg, s_hot_key, s_hot_cnt);
  const int tid = threadIdx.x;
  F.init(P, tid, kRegThreads);
  // launched with programmatic stream serialisation: the launch and the set-up above overlap the tail of the kernel
  // before this one in the stream (phase B of the previous update, which signals early); global memory - the batch,
  // the queues that phase B may still be reading - is only touched after that kernel has completed
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long n = P.n;
  const long long ntiles = (n + kTile - 1) / kTile;
  const u64 pol_stream = P.stream_evict_first ? l2_policy_evict_first() : l2_policy_evict_normal();
  __syncthreads();

  Vec4<PT> pn;
  Vec4<TT> tn;
  long long tile = blockIdx.x;
  if (tile < ntiles && tile * kTile + kTile <= n) {
    pn = load4_hint(preds + tile * kTile + (long long)tid * 4, pol_stream);
    tn = load4_hint(target + tile * kTile + (long long)tid * 4, pol_stream);
  }
  for (; tile < ntiles; tile += gridDim.x) {
    const long long base = tile * kTile;
    if (base + kTile <= n) {
      const PT* pp = preds + base + (long long)tid * 4;
      const TT* tp = target + base + (long long)tid * 4;
#pragma unroll
      for (int v = 0; v < kRegItems / 4; ++v) {
        const Vec4<PT> pc = pn;
        const Vec4<TT> tc = tn;
        if (v + 1 < kRegItems / 4) {
          pn = load4_hint(pp + (long long)(v + 1) * kRegThreads * 4, pol_stream);
          tn = load4_hint(tp + (long long)(v + 1) * kRegThreads * 4, pol_stream);
        } else {
          // first vector of this CTA's next tile: in flight while the FIFOs are flushed
          const long long nbase = (tile + gridDim.x) * kTile;
          if (nbase + kTile <= n) {
            pn = load4_hint(preds + nbase + (long long)tid * 4, pol_stream);
            tn = load4_hint(target + nbase + (long long)tid * 4, pol_stream);
          }
        }
        F.push<4, PT, TT, HAS_IGNORE>(P, pc.v, tc.v);
      }
    } else {
      for (long long e = base + tid; e < n; e += kRegThreads) {
        const PT pv[1] = {preds[e]};
        const TT tv[1] = {target[e]};
        F.push<1, PT, TT, HAS_IGNORE>(P, pv, tv);
      }
    }
    __syncthreads();
    F.flush<kRegThreads>(P, tid, false);
    __syncthreads();
  }
  F.flush<kRegThreads>(P, tid, true);
  __syncthreads();
  F.finish(P, tid, kRegThreads);
}

// ------------------------------------------------------------------------------------------------------------
// Phase A, TMA-fed: the batch arrives through a ring of shared-memory stages filled by bulk asynchronous copies
// (cp.async.bulk + mbarrier), so the bytes in flight per SM are set by the ring (4 x 32 KB), not by how many
// registers the threads can park loads in.  One CTA of 1024 threads per SM; thread 0 is the producer.  A tile is
// the four stages of the ring, so slots and barrier parities are compile-time / per-tile constants.
// ------------------------------------------------------------------------------------------------------------
constexpr int kTmaThreads = 1024;
constexpr int kTmaPer = 2;                                  // samples per thread per stage
constexpr int kTmaStageSamples = kTmaThreads * kTmaPer;     // 2048
constexpr int kTmaStages = 5;                               // FIFOs are flushed once per trip round the ring
constexpr int kTmaStageBytes = kTmaStageSamples * 16;       // room for two int64 arrays
constexpr int kTmaFifoEntries = 24576;

__device__ __forceinline__ u32 smem_addr(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u32 bar, u32 count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u32 bar, u32 bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(u32 bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(u32 bar, u32 parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(u32 dst, const void* src, u32 bytes, u32 bar, u64 policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(policy) : "memory");
}

template <typename T>
struct alignas(2 * sizeof(T)) Pair2 {
  T v[2];
};

template <typename PT, typename TT, bool HAS_IGNORE>
__global__ void __launch_bounds__(kTmaThreads, 1) confmat_partition_tma_kernel(const PartParams P) {
  extern __shared__ __align__(128) unsigned char smem_tma[];
  unsigned char* const ring = smem_tma;  // [stages][preds half | target half]
  __shared__ u32 s_hot_key[kHotSlots], s_hot_cnt[kHotSlots];
  __shared__ __align__(8) u64 s_full[kTmaStages], s_empty[kTmaStages];
  Fifos F;
  carve(F, smem_tma + kTmaStages * kTmaStageBytes, s_hot_key, s_hot_cnt);
  const int tid = threadIdx.x, lane = tid & 31;
  const PT* __restrict__ preds = static_cast<const PT*>(P.preds);
  const TT* __restrict__ target = static_cast<const TT*>(P.target);
  const long long n = P.n;
  const u64 pol_stream = P.stream_evict_first ? l2_policy_evict_first() : l2_policy_evict_normal();
  // stages are dealt round-robin to the CTAs (finer than whole tiles: the CTAs finish within one stage of each
  // other); the ragged end of the batch is pushed straight from global memory
  const u32 nstages = (u32)(n / kTmaStageSamples);
  const u32 mine = nstages > blockIdx.x ? (nstages - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  auto issue = [&](u32 k, u32 slot) {  // producer: this CTA's k-th stage into ring slot `slot`
    const long long s0 = (long long)(blockIdx.x + k * gridDim.x) * kTmaStageSamples;
    const u32 bar = smem_addr(&s_full[slot]);
    mbar_expect_tx(bar, (u32)(kTmaStageSamples * (sizeof(PT) + sizeof(TT))));
    bulk_load(smem_addr(ring + slot * kTmaStageBytes), preds + s0, (u32)(kTmaStageSamples * sizeof(PT)), bar,
              pol_stream);
    bulk_load(smem_addr(ring + slot * kTmaStageBytes + kTmaStageBytes / 2), target + s0,
              (u32)(kTmaStageSamples * sizeof(TT)), bar, pol_stream);
  };
  if (tid == 0) {
    for (int i = 0; i < kTmaStages; ++i) {
      mbar_init(smem_addr(&s_full[i]), 1);
      mbar_init(smem_addr(&s_empty[i]), kTmaThreads / 32);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  F.init(P, tid, kTmaThreads);
  // launched with programmatic stream serialisation: the launch and the set-up above overlap the tail of the kernel
  // before this one in the stream (phase B of the previous update, which signals early); global memory - the batch,
  // the queues that phase B may still be reading - is only touched after that kernel has completed
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (tid == 0)
    for (u32 slot = 0; slot < kTmaStages && slot < mine; ++slot) issue(slot, slot);
  __syncthreads();

  for (u32 k0 = 0; k0 < mine; k0 += kTmaStages) {
    const u32 phase = (k0 / kTmaStages) & 1u;
#pragma unroll
    for (int slot = 0; slot < kTmaStages; ++slot) {
      if (k0 + slot < mine) {
        mbar_wait(smem_addr(&s_full[slot]), phase);
        const Pair2<PT> pp = reinterpret_cast<const Pair2<PT>*>(ring + slot * kTmaStageBytes)[tid];
        const Pair2<TT> tt =
            reinterpret_cast<const Pair2<TT>*>(ring + slot * kTmaStageBytes + kTmaStageBytes / 2)[tid];
        F.push<2, PT, TT, HAS_IGNORE>(P, pp.v, tt.v);
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_addr(&s_empty[slot]));
      }
      if (tid == 0 && slot >= 1 && k0 + kTmaStages + slot - 1 < mine) {
        // refill the slot consumed one step ago: by now (almost) every warp has released it
        mbar_wait(smem_addr(&s_empty[slot - 1]), phase);
        issue(k0 + kTmaStages + slot - 1, slot - 1);
      }
    }
    __syncthreads();
    if (tid == 0 && k0 + 2 * kTmaStages - 1 < mine) {
      mbar_wait(smem_addr(&s_empty[kTmaStages - 1]), phase);
      issue(k0 + 2 * kTmaStages - 1, kTmaStages - 1);
    }
    F.flush<kTmaThreads>(P, tid, false);
    __syncthreads();
  }
  if (blockIdx.x == gridDim.x - 1) {
    for (long long e = (long long)nstages * kTmaStageSamples + tid; e < n; e += kTmaThreads) {
      const PT pv[1] = {preds[e]};
      const TT tv[1] = {target[e]};
      F.push<1, PT, TT, HAS_IGNORE>(P, pv, tv);
    }
  }
  __syncthreads();
  F.flush<kTmaThreads>(P, tid, true);
  __syncthreads();
  F.finish(P, tid, kTmaThreads);
}

// ------------------------------------------------------------------------------------------------------------
// Phase B: kHistSplit CTAs per bucket histogram the bucket's segments in shared memory and add the result to the
// state.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kHistThreads) confmat_bucket_hist_kernel(const PartParams P) {
  __shared__ u32 h[kBinsPerBucket];
  const u32 b = blockIdx.x;
  const u64 pol_done = l2_policy_evict_first();  // the queue is dead once read
  for (int i = threadIdx.x; i < kBinsPerBucket; i += kHistThreads) h[i] = 0;
  // launched with programmatic stream serialisation: everything above overlaps the tail of phase A
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __syncthreads();
  auto count8 = [&](const uint4& w) {
    atomicAdd(&h[w.x & 0xffff], 1u); atomicAdd(&h[w.x >> 16], 1u);
    atomicAdd(&h[w.y & 0xffff], 1u); atomicAdd(&h[w.y >> 16], 1u);
    atomicAdd(&h[w.z & 0xffff], 1u); atomicAdd(&h[w.z >> 16], 1u);
    atomicAdd(&h[w.w & 0xffff], 1u); atomicAdd(&h[w.w >> 16], 1u);
  };
  // the bucket's G segments are dealt round-robin to the warps of the kHistSplit CTAs that share the bucket; a
  // segment is a few hundred entries.  The loop is software-pipelined: the (up to four) 16-byte vectors a lane takes
  // from its warp's NEXT segment are in flight while it counts the current one, and the fill count two segments
  // ahead is loaded alongside - the plain loop waited for one queue load per eight atomics (48 % of its stall samples,
  // profiles/r2_confmat_hot.txt).
  const u32 lane = threadIdx.x & 31;
  const u32 wid = blockIdx.y * (kHistThreads / 32) + (threadIdx.x >> 5);
  const u32 nw = gridDim.y * (kHistThreads / 32);
  const u32* fill_b = P.fill + (size_t)b * P.G;
  auto seg_vecs = [&](u32 seg) {
    return reinterpret_cast<const uint4*>(P.queue + ((size_t)b * P.G + seg) * P.segcap);  // segcap % 8 == 0
  };
  constexpr int kVecs = 4;  // covers segments of up to 1024 entries; longer ones finish in the tail loop
  uint4 cur[kVecs];
  u32 len = wid < P.G ? fill_b[wid] : 0;
  u32 len1 = wid + nw < P.G ? fill_b[wid + nw] : 0;
  if (wid < P.G) {
    const uint4* qv = seg_vecs(wid);
#pragma unroll
    for (int k = 0; k < kVecs; ++k)
      if (lane + 32 * k < (len >> 3)) cur[k] = load16_hint(qv + lane + 32 * k, pol_done);
  }
  for (u32 seg = wid; seg < P.G; seg += nw) {
    uint4 nxt[kVecs];
    const u32 len2 = seg + 2 * nw < P.G ? fill_b[seg + 2 * nw] : 0;
    if (seg + nw < P.G) {
      const uint4* qn = seg_vecs(seg + nw);
#pragma unroll
      for (int k = 0; k < kVecs; ++k)
        if (lane + 32 * k < (len1 >> 3)) nxt[k] = load16_hint(qn + lane + 32 * k, pol_done);
    }
    const uint4* qv = seg_vecs(seg);
    const u32 nvec = len >> 3;
#pragma unroll
    for (int k = 0; k < kVecs; ++k)
      if (lane + 32 * k < nvec) count8(cur[k]);
    for (u32 i = lane + 32 * kVecs; i < nvec; i += 32) count8(load16_hint(qv + i, pol_done));
    const unsigned short* q = reinterpret_cast<const unsigned short*>(qv);
    for (u32 k = (nvec << 3) + lane; k < len; k += 32) atomicAdd(&h[q[k]], 1u);
#pragma unroll
    for (int k = 0; k < kVecs; ++k) cur[k] = nxt[k];
    len = len1;
    len1 = len2;
  }
  // all CTAs of this grid are resident (at most four per SM are launched): the next kernel of the stream may begin
  // its set-up where CTAs leave; it waits for this grid to complete before it touches global memory
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  __syncthreads();
  for (u32 lbin = threadIdx.x; lbin < kBinsPerBucket; lbin += kHistThreads) {
    const u32 c = h[lbin];
    if (c) atomicAdd(P.confmat + cell_of(b, lbin, P.nb), (u64)c);
  }
}

// 8-byte labels (the reference's default dtype) make phase A a bandwidth problem: TMA-fed kernel, one CTA per SM.
// Narrower labels leave it bound by the FIFO pushes, where the register-fed kernel's four pushes in flight per
// thread win (measured).
static bool part_use_tma(int preds_dtype, int target_dtype) {
  return dtype_size(preds_dtype) == 8 && dtype_size(target_dtype) == 8;
}
static u32 part_ctas(bool tma) { return (u32)sm_count() * (tma ? 1 : kRegCtasPerSm); }
static size_t part_header_bytes(u32 nb, u32 G) { return ((size_t)nb * G * 4 + 255) & ~(size_t)255; }
// buckets for a CC-cell matrix: 0 = not served by the partition path, 1 = small enough for the direct kernels
static long long part_buckets(long long CC) {
  const long long need = (((CC + 15) >> 4) + (kBinsPerBucket / 16) - 1) / (kBinsPerBucket / 16);
  if (need > kMaxBuckets) return 0;
  if (need < 2) return 1;
  long long nb = kMinBuckets;
  while (nb < need) nb <<= 1;
  return nb;
}

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_confmat_workspace_bytes(int64_t n, int64_t num_classes) {
  if (n < 1 || num_classes < 1) return 0;
  const long long nb = part_buckets((long long)num_classes * num_classes);
  if (nb == 0) return 0;  // not served by the partition path
  const long long G = part_ctas(false);  // the larger of the two grids; the entry point re-derives segcap from G
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
  const bool tma = part_use_tma(preds_dtype, target_dtype);
  const u32 G = part_ctas(tma);
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
  while ((1u << P.lognb) < P.nb) ++P.lognb;
  P.has_ignore = has_ignore ? 1 : 0;
  P.ignore = ignore_index;
  P.G = G;
  P.fill = (u32*)workspace;
  P.queue = (unsigned short*)((unsigned char*)workspace + header);
  P.segcap = (u32)segcap;
  P.confmat = (u64*)confmat;
  P.status = status;
  // The queues (2 B/sample) are handed from phase A to phase B through the L2.  While they fit comfortably the
  // batch is read "evict first" so that it does not push them out (2^24 samples: 71 vs 73 us); the hint costs ~6 % of
  // streaming bandwidth, so larger batches, whose queues spill anyway, are read normally (2^26: 242 vs 258 us).
  P.stream_evict_first = (size_t)n * 2 <= ((size_t)40 << 20);
  cudaStream_t st = (cudaStream_t)stream;
  const u32 entries = tma ? kTmaFifoEntries : kRegFifoEntries;
  P.cap = (entries / P.nb) & ~7u;
  const size_t smem = fifo_smem_bytes(entries) + (tma ? (size_t)kTmaStages * kTmaStageBytes : 0);
  // always G CTAs: the ones without a tile still publish their (zero) fill counts
  TMB_DISPATCH_INT(preds_dtype, PT, {
    TMB_DISPATCH_INT(target_dtype, TT, {
      auto launch = [&](auto kern, int threads) -> int {
        TMB_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaLaunchConfig_t acfg{};
        acfg.gridDim = dim3(G);
        acfg.blockDim = dim3((unsigned)threads);
        acfg.dynamicSmemBytes = smem;
        acfg.stream = st;
        cudaLaunchAttribute aattr[1];
        aattr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        aattr[0].val.programmaticStreamSerializationAllowed = 1;
        acfg.attrs = aattr;
        acfg.numAttrs = 1;
        TMB_CUDA_CHECK(cudaLaunchKernelEx(&acfg, kern, P));
        return TMB_OK;
      };
      int rc;
      if (tma)
        rc = P.has_ignore ? launch(confmat_partition_tma_kernel<PT, TT, true>, kTmaThreads)
                          : launch(confmat_partition_tma_kernel<PT, TT, false>, kTmaThreads);
      else
        rc = P.has_ignore ? launch(confmat_partition_reg_kernel<PT, TT, true>, kRegThreads)
                          : launch(confmat_partition_reg_kernel<PT, TT, false>, kRegThreads);
      if (rc != TMB_OK) return rc;
      cudaLaunchConfig_t cfg{};
      // CTAs per bucket: enough CTAs to fill the chip, no more (every extra CTA re-adds the bucket's bins with REDs;
      // measured at 2^24 samples: C=1000 68 us with 2, 70 with 4, 74 with 1; C=2000 74 us with 1, 82 with 4)
      const int split = nb >= 256 ? 1 : (int)(256 / nb) < kHistSplit ? (int)(256 / nb) : kHistSplit;
      cfg.gridDim = dim3((unsigned)nb, split);
      cfg.blockDim = dim3(kHistThreads);
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      TMB_CUDA_CHECK(cudaLaunchKernelEx(&cfg, confmat_bucket_hist_kernel, P));
      return TMB_OK;
    })
  })
  return TMB_ERR_ARG;
}
