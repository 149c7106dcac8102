// common.cuh -- shared helpers for libradnerf_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/radnerf_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libradnerf_b200 is written for sm_100a (B200) only"
#endif

namespace rnb {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs; grids are sized in multiples of this
constexpr unsigned kFullMask = 0xffffffffu;

void set_error(const char* fmt, ...);

inline cudaStream_t as_stream(rnb_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

inline int check_launch_at(const char* what, const char* file, int line) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();  // clear the (non-sticky) error
    set_error("%s: %s [%s:%d]", what, cudaGetErrorString(e), file, line);
    return RNB_ERR_LAUNCH;
  }
  return RNB_OK;
}

#define check_launch(what) check_launch_at(what, __FILE__, __LINE__)

#define RNB_REQUIRE(cond, what)                         \
  do {                                                  \
    if (!(cond)) {                                      \
      ::rnb::set_error("%s: bad argument (%s)", what, #cond); \
      return RNB_ERR_BADARG;                            \
    }                                                   \
  } while (0)

#define RNB_CUDA(call, what)                                              \
  do {                                                                    \
    cudaError_t e__ = (call);                                             \
    if (e__ != cudaSuccess) {                                             \
      ::rnb::set_error("%s: %s [%s:%d]", what, cudaGetErrorString(e__), __FILE__, __LINE__); \
      return RNB_ERR_LAUNCH;                                              \
    }                                                                     \
  } while (0)

inline int ceil_div(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// ---- device helpers ----------------------------------------------------------------
__device__ __forceinline__ float ld_nc_f32(const float* p) { return __ldg(p); }

// streaming (read-once) loads / write-once stores: do not pollute L1
__device__ __forceinline__ float ld_stream_f32(const float* p) {
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ float4 ld_stream_f32x4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream_f32(float* p, float v) {
  asm volatile("st.global.L1::no_allocate.f32 [%0], %1;" ::"l"(p), "f"(v));
}
__device__ __forceinline__ void st_stream_f32x4(float4* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w));
}

// uniform [0,1) from a 64-bit counter (splitmix64 finaliser; 24 mantissa bits): the library's own generator where the caller
// supplies no random inputs (marching noise) -- no ATen launch inside the step, so the step can be captured in a CUDA graph
__device__ __forceinline__ float counter_uniform(uint64_t seed, uint64_t ctr) {
  uint64_t z = seed + (ctr + 1) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(uint32_t)(z >> 40) * (1.0f / 16777216.0f);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
  return v;
}
// inclusive scan across the warp
__device__ __forceinline__ float warp_incl_scan(float v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    float n = __shfl_up_sync(kFullMask, v, o);
    if (lane >= o) v += n;
  }
  return v;
}
__device__ __forceinline__ int warp_incl_scan_i(int v, int lane) {
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int n = __shfl_up_sync(kFullMask, v, o);
    if (lane >= o) v += n;
  }
  return v;
}

}  // namespace rnb
