// Shared helpers of libfedb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define FEDB_MAX_PEERS 16

// ---- error plumbing: every extern "C" entry returns 0 or a cudaError_t-like code and
// ---- leaves a message readable through fedb_last_error().
extern thread_local char g_fedb_err[512];
int fedb_fail(int code, const char* file, int line, const char* what);

#define FEDB_CHECK(expr)                                                      \
  do {                                                                        \
    cudaError_t _e = (expr);                                                  \
    if (_e != cudaSuccess) return fedb_fail((int)_e, __FILE__, __LINE__, cudaGetErrorString(_e)); \
  } while (0)

#define FEDB_REQUIRE(cond, msg)                                               \
  do {                                                                        \
    if (!(cond)) return fedb_fail(-1, __FILE__, __LINE__, msg);               \
  } while (0)

#define FEDB_LAUNCH_OK() FEDB_CHECK(cudaGetLastError())

static inline int fedb_num_sms() {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

static inline int fedb_ceil_div(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- device helpers ---------------------------------------------------------------------
__device__ __forceinline__ float4 ld_nc_f4(const float4* p) {
  // streaming 128-bit load, no L1 allocation (also the right thing for peer addresses: peer
  // data bypasses the local L2 and is only L1-cacheable)
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float4* p) {
  // coherent (non-.nc) 128-bit load for data another GPU has just written
  float4 r;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_na_f4(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
