// K5 — one-shot all-reduce of a small state arena over NVLink peer memory (contract: tmb_oneshot_allreduce in
// include/tm_b200.h).  Replaces, for arenas of a few KB, the NCCL all_reduce behind `Metric._sync_dist`
// (reference: metric.py:427-457 gathers every state with utilities/distributed.py:97-147 and sums the stack).
//
// Every rank owns one buffer in symmetric memory (mapped into every peer's address space): two data slots and two
// rows of arrival flags.  One call = one kernel per rank, one CTA:
//   1. copy the local arena into this rank's slot (epoch parity picks the slot);
//   2. system-scope fence, then store the epoch number into this rank's flag on EVERY peer (release);
//   3. wait until all W flags of the local row carry the epoch (acquire): every peer's slot is complete;
//   4. load all W slots - P2P loads over NVLink / NVSwitch - and reduce them in rank order, so every rank gets the
//      same bits, floating point included.
// Two slots are enough: a rank can enter epoch e+2 (and overwrite slot e) only after every peer has raised its flag for
// e+1, which a peer does after it finished reading epoch e.  No NCCL call, no intermediate kernel: the latency is one
// launch plus two NVLink round trips.  A wait that lasts longer than ~2 s gives up and raises
// TMB_STATUS_PEER_TIMEOUT in the status word instead of hanging the GPU.
#include "common.cuh"

namespace tmb {

constexpr int kOsaThreads = 1024;
constexpr int kOsaMaxWorld = 64;

__device__ __forceinline__ void st_flag_release_sys(u32* p, u32 v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ u32 ld_flag_acquire_sys(const u32* p) {
  u32 v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
template <typename T>
__device__ __forceinline__ T ld_peer(const T* p) {
  return *reinterpret_cast<const volatile T*>(p);  // system-scope relaxed load: never served from a stale L1 line
}

template <typename T, int OP>
__device__ __forceinline__ T combine(T a, T b) {
  if constexpr (OP == TMB_REDUCE_SUM) return a + b;
  else if constexpr (OP == TMB_REDUCE_MAX) return b > a ? b : a;
  else return b < a ? b : a;
}

// the local arena may be scattered over several state tensors: they are packed on the way into the slot
struct OsaSegments {
  const void* ptr[TMB_ONESHOT_MAX_SEGMENTS];
  long long count[TMB_ONESHOT_MAX_SEGMENTS];
  int n;
};

template <typename T, int OP>
__global__ void __launch_bounds__(kOsaThreads) oneshot_allreduce_kernel(unsigned char* const* __restrict__ peers,
                                                                        int world, int rank, const OsaSegments segs,
                                                                        T* __restrict__ dst, long long count, u32 epoch,
                                                                        int* status) {
  __shared__ int s_timeout;
  const int tid = threadIdx.x;
  const size_t slot_off = (size_t)(epoch & 1u) * TMB_ONESHOT_SLOT_BYTES;
  const size_t flag_off = (size_t)2 * TMB_ONESHOT_SLOT_BYTES + (size_t)(epoch & 1u) * kOsaMaxWorld * sizeof(u32);
  if (tid == 0) s_timeout = 0;
  // 1. publish the local arena in this rank's slot
  T* mine = reinterpret_cast<T*>(peers[rank] + slot_off);
  {
    long long base = 0;
    for (int k = 0; k < segs.n; ++k) {
      const T* __restrict__ src = static_cast<const T*>(segs.ptr[k]);
      for (long long i = tid; i < segs.count[k]; i += kOsaThreads) mine[base + i] = src[i];
      base += segs.count[k];
    }
  }
  __threadfence_system();
  __syncthreads();
  // 2. tell every peer (and this rank) that the slot of this epoch is complete
  if (tid < world) st_flag_release_sys(reinterpret_cast<u32*>(peers[tid] + flag_off) + rank, epoch);
  // 3. wait for every peer's slot
  if (tid < world) {
    const u32* flag = reinterpret_cast<const u32*>(peers[rank] + flag_off) + tid;
    const long long t0 = clock64();
    while (ld_flag_acquire_sys(flag) != epoch) {
      if (clock64() - t0 > 4000000000LL) {  // ~2 s at 2 GHz: a peer never arrived
        s_timeout = 1;
        break;
      }
    }
  }
  __syncthreads();
  if (s_timeout) {
    if (tid == 0) atomicOr(status, TMB_STATUS_PEER_TIMEOUT);
    return;
  }
  // 4. reduce the W slots in rank order
  for (long long i = tid; i < count; i += kOsaThreads) {
    T acc = ld_peer(reinterpret_cast<const T*>(peers[0] + slot_off) + i);
    for (int r = 1; r < world; ++r) acc = combine<T, OP>(acc, ld_peer(reinterpret_cast<const T*>(peers[r] + slot_off) + i));
    dst[i] = acc;
  }
}

template <typename T>
static int launch_osa(unsigned char* const* peers, int world, int rank, const OsaSegments& src, void* dst,
                      long long count, int op, u32 epoch, int* status, cudaStream_t st) {
  switch (op) {
    case TMB_REDUCE_SUM:
      oneshot_allreduce_kernel<T, TMB_REDUCE_SUM><<<1, kOsaThreads, 0, st>>>(peers, world, rank, src, (T*)dst,
                                                                             count, epoch, status);
      break;
    case TMB_REDUCE_MAX:
      oneshot_allreduce_kernel<T, TMB_REDUCE_MAX><<<1, kOsaThreads, 0, st>>>(peers, world, rank, src, (T*)dst,
                                                                             count, epoch, status);
      break;
    case TMB_REDUCE_MIN:
      oneshot_allreduce_kernel<T, TMB_REDUCE_MIN><<<1, kOsaThreads, 0, st>>>(peers, world, rank, src, (T*)dst,
                                                                             count, epoch, status);
      break;
    default:
      return TMB_ERR_ARG;
  }
  TMB_CUDA_CHECK(cudaGetLastError());
  return TMB_OK;
}

}  // namespace tmb

using namespace tmb;

extern "C" size_t tmb_oneshot_buffer_bytes(void) {
  return (size_t)2 * TMB_ONESHOT_SLOT_BYTES + (size_t)2 * kOsaMaxWorld * sizeof(u32);
}

static int osa_dispatch(void* const* peer_buffers_dev, int world, int rank, const OsaSegments& segs, void* dst,
                        int64_t count, int dtype, int op, uint32_t epoch, int32_t* status, void* stream) {
  if (!peer_buffers_dev || world < 1 || world > kOsaMaxWorld || rank < 0 || rank >= world || count < 0 || !status ||
      epoch == 0)
    return TMB_ERR_ARG;
  if (count == 0) return TMB_OK;
  if (!dst) return TMB_ERR_ARG;
  const int esize = dtype_size(dtype);
  if (esize == 0) return TMB_ERR_ARG;
  if ((size_t)count * esize > (size_t)TMB_ONESHOT_SLOT_BYTES) return TMB_ERR_TOO_LARGE;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* const* peers = reinterpret_cast<unsigned char* const*>(peer_buffers_dev);
  switch (dtype) {
    case TMB_I64: return launch_osa<long long>(peers, world, rank, segs, dst, count, op, epoch, status, st);
    case TMB_I32: return launch_osa<int>(peers, world, rank, segs, dst, count, op, epoch, status, st);
    case TMB_F32: return launch_osa<float>(peers, world, rank, segs, dst, count, op, epoch, status, st);
    case TMB_F64: return launch_osa<double>(peers, world, rank, segs, dst, count, op, epoch, status, st);
    default: return TMB_ERR_ARG;
  }
}

extern "C" int tmb_oneshot_allreduce(void* const* peer_buffers_dev, int world, int rank, const void* src, void* dst,
                                     int64_t count, int dtype, int op, uint32_t epoch, int32_t* status, void* stream) {
  if (count > 0 && !src) return TMB_ERR_ARG;
  OsaSegments segs{};
  segs.ptr[0] = src;
  segs.count[0] = count;
  segs.n = 1;
  return osa_dispatch(peer_buffers_dev, world, rank, segs, dst, count, dtype, op, epoch, status, stream);
}

extern "C" int tmb_oneshot_allreduce_packed(void* const* peer_buffers_dev, int world, int rank,
                                            const void* const* srcs, const int64_t* counts, int num_segments,
                                            void* dst, int dtype, int op, uint32_t epoch, int32_t* status,
                                            void* stream) {
  if (!srcs || !counts || num_segments < 1 || num_segments > TMB_ONESHOT_MAX_SEGMENTS) return TMB_ERR_ARG;
  OsaSegments segs{};
  int64_t total = 0;
  for (int k = 0; k < num_segments; ++k) {
    if (counts[k] < 0 || (counts[k] > 0 && !srcs[k])) return TMB_ERR_ARG;
    segs.ptr[k] = srcs[k];
    segs.count[k] = counts[k];
    total += counts[k];
  }
  segs.n = num_segments;
  return osa_dispatch(peer_buffers_dev, world, rank, segs, dst, total, dtype, op, epoch, status, stream);
}
