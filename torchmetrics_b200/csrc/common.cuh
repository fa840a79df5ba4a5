// Shared device/host helpers for libtm_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>

#include "../../include/tm_b200.h"

namespace tmb {

using u64 = unsigned long long;
using u32 = unsigned int;

// ---- host side -------------------------------------------------------------------------------
void set_cuda_error(cudaError_t e, const char* where);
void set_error_text(const char* text);  // message returned by tmb_last_cuda_error() for non-CUDA failures
int sm_count();            // cached per device
int max_optin_smem();      // cached per device (bytes)

#define TMB_CUDA_CHECK(expr)                          \
  do {                                                \
    cudaError_t _e = (expr);                          \
    if (_e != cudaSuccess) {                          \
      ::tmb::set_cuda_error(_e, #expr);               \
      return TMB_ERR_CUDA;                            \
    }                                                 \
  } while (0)

inline bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31u) == 0; }

inline int dtype_size(int dt) {
  switch (dt) {
    case TMB_I64: case TMB_F64: return 8;
    case TMB_I32: case TMB_F32: return 4;
    case TMB_F16: case TMB_BF16: return 2;
    case TMB_U8: return 1;
    default: return 0;
  }
}
inline bool is_int_dtype(int dt) { return dt == TMB_I64 || dt == TMB_I32 || dt == TMB_U8; }
inline bool is_float_dtype(int dt) { return dt >= TMB_F32 && dt <= TMB_F64; }

// ---- device side: streaming vector loads ---------------------------------------------------------
// Inputs are read exactly once, so they go through the non-coherent path without allocating in L1.
// Vec4<T> is 4 consecutive elements; one load instruction of 4*sizeof(T) bytes (32..256 bit).
template <typename T>
struct Vec4 {
  T v[4];
};

__device__ __forceinline__ void ldg_stream_256(const void* p, u64& a, u64& b, u64& c, u64& d) {
  asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
               : "=l"(a), "=l"(b), "=l"(c), "=l"(d)
               : "l"(p));
}
__device__ __forceinline__ void ldg_stream_128(const void* p, u32& a, u32& b, u32& c, u32& d) {
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(a), "=r"(b), "=r"(c), "=r"(d)
               : "l"(p));
}
__device__ __forceinline__ void ldg_stream_64(const void* p, u32& a, u32& b) {
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p));
}
__device__ __forceinline__ u32 ldg_stream_32(const void* p) {
  u32 a;
  asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(a) : "l"(p));
  return a;
}

template <typename T>
__device__ __forceinline__ Vec4<T> load4(const T* p) {
  Vec4<T> r;
  if constexpr (sizeof(T) == 8) {
    u64 a, b, c, d;
    ldg_stream_256(p, a, b, c, d);
    r.v[0] = *reinterpret_cast<T*>(&a);
    r.v[1] = *reinterpret_cast<T*>(&b);
    r.v[2] = *reinterpret_cast<T*>(&c);
    r.v[3] = *reinterpret_cast<T*>(&d);
  } else if constexpr (sizeof(T) == 4) {
    u32 a, b, c, d;
    ldg_stream_128(p, a, b, c, d);
    r.v[0] = *reinterpret_cast<T*>(&a);
    r.v[1] = *reinterpret_cast<T*>(&b);
    r.v[2] = *reinterpret_cast<T*>(&c);
    r.v[3] = *reinterpret_cast<T*>(&d);
  } else if constexpr (sizeof(T) == 2) {
    u32 a, b;
    ldg_stream_64(p, a, b);
    unsigned short s0 = a & 0xffff, s1 = a >> 16, s2 = b & 0xffff, s3 = b >> 16;
    r.v[0] = *reinterpret_cast<T*>(&s0);
    r.v[1] = *reinterpret_cast<T*>(&s1);
    r.v[2] = *reinterpret_cast<T*>(&s2);
    r.v[3] = *reinterpret_cast<T*>(&s3);
  } else {
    u32 a = ldg_stream_32(p);
    unsigned char c0 = a & 0xff, c1 = (a >> 8) & 0xff, c2 = (a >> 16) & 0xff, c3 = a >> 24;
    r.v[0] = *reinterpret_cast<T*>(&c0);
    r.v[1] = *reinterpret_cast<T*>(&c1);
    r.v[2] = *reinterpret_cast<T*>(&c2);
    r.v[3] = *reinterpret_cast<T*>(&c3);
  }
  return r;
}

// ---- L2 residency hints -------------------------------------------------------------------------------------
// A batch streamed once should not push out the few MB a kernel pair hands over through the L2 (K1's queues):
// the stream is read "evict first", the hand-over is written "evict last" and read back "evict first".
__device__ __forceinline__ u64 l2_policy_evict_first() {
  u64 pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ u64 l2_policy_evict_normal() {
  u64 pol;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ u64 l2_policy_evict_last() {
  u64 pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
template <typename T>
__device__ __forceinline__ Vec4<T> load4_hint(const T* p, u64 pol) {
  Vec4<T> r;
  if constexpr (sizeof(T) == 8) {
    u64 a, b, c, d;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u64 {%0,%1,%2,%3}, [%4], %5;"
                 : "=l"(a), "=l"(b), "=l"(c), "=l"(d)
                 : "l"(p), "l"(pol));
    r.v[0] = *reinterpret_cast<T*>(&a);
    r.v[1] = *reinterpret_cast<T*>(&b);
    r.v[2] = *reinterpret_cast<T*>(&c);
    r.v[3] = *reinterpret_cast<T*>(&d);
  } else if constexpr (sizeof(T) == 4) {
    u32 a, b, c, d;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(a), "=r"(b), "=r"(c), "=r"(d)
                 : "l"(p), "l"(pol));
    r.v[0] = *reinterpret_cast<T*>(&a);
    r.v[1] = *reinterpret_cast<T*>(&b);
    r.v[2] = *reinterpret_cast<T*>(&c);
    r.v[3] = *reinterpret_cast<T*>(&d);
  } else {
    r = load4(p);
  }
  return r;
}
__device__ __forceinline__ void store16_hint(void* p, const uint4& w, u64 pol) {
  asm volatile("st.global.L2::cache_hint.v4.b32 [%0], {%1,%2,%3,%4}, %5;"
               :: "l"(p), "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w), "l"(pol)
               : "memory");
}
__device__ __forceinline__ uint4 load16_hint(const void* p, u64 pol) {
  uint4 r;
  asm volatile("ld.global.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol));
  return r;
}

// Scores are compared in float (f32/f16/bf16 convert exactly) or in double (f64), which is what torch's
// type promotion does for `preds >= thresholds` with float32 thresholds.
template <typename T>
struct CmpType {
  using type = float;
};
template <>
struct CmpType<double> {
  using type = double;
};

__device__ __forceinline__ float to_cmp(float x) { return x; }
__device__ __forceinline__ double to_cmp(double x) { return x; }
__device__ __forceinline__ float to_cmp(__half x) { return __half2float(x); }
__device__ __forceinline__ float to_cmp(__nv_bfloat16 x) { return __bfloat162float(x); }

// Same expression ATen evaluates for sigmoid on CUDA: 1 / (1 + exp(-x)) in the op-math type (float for
// f32/f16/bf16), full-precision expf and IEEE division (this file is built without fast-math).
__device__ __forceinline__ float sigmoid_ref(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ double sigmoid_ref(double x) { return 1.0 / (1.0 + exp(-x)); }

// ATen rounds the sigmoid result back to the tensor dtype before anything else sees it.
template <typename T>
__device__ __forceinline__ typename CmpType<T>::type round_to_storage(typename CmpType<T>::type x) {
  if constexpr (sizeof(T) == 2) {
    T t = T(x);
    return to_cmp(t);
  } else {
    return x;
  }
}

__device__ __forceinline__ long long as_i64(long long x) { return x; }
__device__ __forceinline__ long long as_i64(int x) { return x; }
__device__ __forceinline__ long long as_i64(short x) { return x; }
__device__ __forceinline__ long long as_i64(signed char x) { return x; }
__device__ __forceinline__ long long as_i64(unsigned char x) { return x; }

// status words: [0] errors that invalidate the update, [1] fused-validation findings (informational)
__device__ __forceinline__ void report_status(int* status, int bits) {
  const int err = bits & (TMB_STATUS_LABEL_RANGE | TMB_STATUS_THRESHOLD_ORDER | TMB_STATUS_OUT_OF_UNIT_RANGE);
  const int val = bits & (TMB_STATUS_TARGET_VALUES | TMB_STATUS_PREDS_VALUES);
  if (err) atomicOr(status, err);
  if (val) atomicOr(status + 1, val);
}

// Dispatch helpers ------------------------------------------------------------------------------
#define TMB_DISPATCH_INT(dt, T, ...)                                    \
  switch (dt) {                                                         \
    case TMB_I64: { using T = long long; __VA_ARGS__; break; }          \
    case TMB_I32: { using T = int; __VA_ARGS__; break; }                \
    case TMB_U8:  { using T = unsigned char; __VA_ARGS__; break; }      \
    default: return TMB_ERR_ARG;                                        \
  }

#define TMB_DISPATCH_FLOAT(dt, T, ...)                                  \
  switch (dt) {                                                         \
    case TMB_F32:  { using T = float; __VA_ARGS__; break; }             \
    case TMB_F16:  { using T = __half; __VA_ARGS__; break; }            \
    case TMB_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }     \
    case TMB_F64:  { using T = double; __VA_ARGS__; break; }            \
    default: return TMB_ERR_ARG;                                        \
  }

}  // namespace tmb
