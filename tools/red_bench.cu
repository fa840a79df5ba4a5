// Micro-benchmark (development tool): throughput of scattered integer increments on B200 by memory level.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/red_bench tools/red_bench.cu && /tmp/red_bench
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
namespace cg = cooperative_groups;
typedef unsigned long long u64;
typedef unsigned int u32;

__device__ __forceinline__ u32 hash32(u32 x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}

template <int MODE>  // 0 red64, 1 red32, 2 atom64 (with return), 3 no-op (hash only)
__global__ void k_global(void* table, u32 mask, long long n, u64* sink) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  u64 acc = 0;
  for (; i < n; i += stride) {
    u32 k = hash32((u32)i) & mask;
    if (MODE == 0) atomicAdd((u64*)table + k, 1ULL);
    if (MODE == 1) atomicAdd((u32*)table + k, 1u);
    if (MODE == 2) acc += atomicAdd((u64*)table + k, 1ULL);
    if (MODE == 3) acc += k;
  }
  if (acc == 0x123456789ULL) *sink = acc;
}

// shared-memory atomics: each CTA increments its own 4096-bin table
template <int MODE>  // 0 atomicAdd no return, 1 with return, 2 match_any + leader add, 3 plain RMW (racy, timing only)
__global__ void k_smem(long long n, u64* sink) {
  __shared__ u32 h[4096];
  for (int i = threadIdx.x; i < 4096; i += blockDim.x) h[i] = 0;
  __syncthreads();
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  u32 acc = 0;
  for (; i < n; i += stride) {
    u32 k = hash32((u32)i) & 4095;
    if (MODE == 0) atomicAdd(&h[k], 1u);
    if (MODE == 1) acc += atomicAdd(&h[k], 1u);
    if (MODE == 2) {
      unsigned peers = __match_any_sync(0xffffffffu, k & 255);
      acc += __popc(peers);
    }
    if (MODE == 3) h[k] += 1;
  }
  __syncthreads();
  if (acc == 0x12345678u || h[threadIdx.x] == 0xffffffffu) *sink = acc;
}

// distributed shared memory: cluster of CS CTAs, table of CS*BINS u32 spread over the cluster, random remote REDs
template <int CS, int BINS>
__global__ void k_dsmem(long long n, u64* sink) {
  extern __shared__ u32 hd[];
  cg::cluster_group cluster = cg::this_cluster();
  for (int i = threadIdx.x; i < BINS; i += blockDim.x) hd[i] = 0;
  cluster.sync();
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    u32 k = hash32((u32)i);
    u32 owner = (k >> 20) % CS;
    u32 bin = k % BINS;
    u32* remote = cluster.map_shared_rank(hd, owner);
    atomicAdd(remote + bin, 1u);
  }
  cluster.sync();
  if (hd[threadIdx.x] == 0xffffffffu) *sink = 1;
}

template <typename F>
float timeit(F f, int reps = 5) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    cudaEventRecord(a); f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("  CUDA error: %s\n", cudaGetErrorString(e));
  return best;
}

int main() {
  const long long n = 1LL << 26;
  void* table; cudaMalloc(&table, 1ULL << 28); cudaMemset(table, 0, 1ULL << 28);
  u64* sink; cudaMalloc(&sink, 8);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  printf("SMs %d, n = 2^26 increments per test\n", sms);
  for (int ctas_per_sm : {4, 8}) {
    int grid = sms * ctas_per_sm;
    for (u32 bits : {10u, 16u, 20u, 24u}) {
      u32 mask = (1u << bits) - 1;
      float t0 = timeit([&] { k_global<0><<<grid, 256>>>(table, mask, n, sink); });
      float t1 = timeit([&] { k_global<1><<<grid, 256>>>(table, mask, n, sink); });
      float t2 = timeit([&] { k_global<2><<<grid, 256>>>(table, mask, n, sink); });
      printf("global table 2^%u bins, %d CTA/SM: RED.64 %7.1f G/s  RED.32 %7.1f G/s  ATOM.64 %7.1f G/s\n", bits,
             ctas_per_sm, n / t0 / 1e6, n / t1 / 1e6, n / t2 / 1e6);
    }
  }
  {
    float t = timeit([&] { k_global<3><<<sms * 8, 256>>>(table, 0xfffff, n, sink); });
    printf("hash only (no memory op): %7.1f G/s\n", n / t / 1e6);
  }
  for (int ctas_per_sm : {2, 4, 8}) {
    int grid = sms * ctas_per_sm;
    float t0 = timeit([&] { k_smem<0><<<grid, 256>>>(n, sink); });
    float t1 = timeit([&] { k_smem<1><<<grid, 256>>>(n, sink); });
    float t2 = timeit([&] { k_smem<2><<<grid, 256>>>(n, sink); });
    float t3 = timeit([&] { k_smem<3><<<grid, 256>>>(n, sink); });
    printf("smem 4096 bins, %d CTA/SM: ATOMS(no ret) %7.1f G/s  ATOMS(ret) %7.1f G/s  match_any %7.1f G/s  plain RMW %7.1f G/s\n",
           ctas_per_sm, n / t0 / 1e6, n / t1 / 1e6, n / t2 / 1e6, n / t3 / 1e6);
  }
  {
    constexpr int CS = 8, BINS = 32768;
    cudaFuncSetAttribute(k_dsmem<CS, BINS>, cudaFuncAttributeMaxDynamicSharedMemorySize, BINS * 4);
    cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3((sms / CS) * CS); cfg.blockDim = dim3(512);
    cfg.dynamicSmemBytes = BINS * 4;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim = {CS, 1, 1};
    cfg.attrs = at; cfg.numAttrs = 1;
    float t = timeit([&] { cudaLaunchKernelEx(&cfg, k_dsmem<CS, BINS>, n, sink); });
    printf("DSMEM cluster=%d, %d bins/CTA, random remote RED.32: %7.1f G/s\n", CS, BINS, n / t / 1e6);
  }
  {
    constexpr int CS = 16, BINS = 32768;
    cudaFuncSetAttribute(k_dsmem<CS, BINS>, cudaFuncAttributeMaxDynamicSharedMemorySize, BINS * 4);
    cudaFuncSetAttribute(k_dsmem<CS, BINS>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cudaLaunchConfig_t cfg{}; cfg.gridDim = dim3((sms / CS) * CS); cfg.blockDim = dim3(512);
    cfg.dynamicSmemBytes = BINS * 4;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim = {CS, 1, 1};
    cfg.attrs = at; cfg.numAttrs = 1;
    float t = timeit([&] { cudaLaunchKernelEx(&cfg, k_dsmem<CS, BINS>, n, sink); });
    printf("DSMEM cluster=%d, %d bins/CTA, random remote RED.32: %7.1f G/s\n", CS, BINS, n / t / 1e6);
  }
  return 0;
}
