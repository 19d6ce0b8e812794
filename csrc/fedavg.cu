// K-AGG / K-BCAST / K-REPL and the device-flag barrier (§7.2).
//
// One launch per rank performs, over its shard [begin, end) of the flat fp32 arena:
//   (A) "my arena is final" -> every peer; wait for every peer             (replaces join #1)
//   (B) acc = ((s0 + s1) + s2 ...) / N with sources read through NVLink peer mappings
//       (or decoded from fp8 block-scaled deltas, or summed in the switch by multimem.ld_reduce)
//   (C) result -> every destination arena (clients, server global, backup), optionally as one
//       multimem.st through the NVLS multicast alias, optionally with bf16 shadow copies
//   (D) "my shard is written" -> every peer; wait for every peer          (replaces join #2)
// so that the reference's upload + allreduce() + SendModel fan-out + backup push
// (src/server.py:126-153, 155-179) collapses into a single kernel.  Spins are bounded.
#include "common.cuh"
#include "../include/fedb200.h"
#include <cuda_fp8.h>

namespace {

constexpr int kThreads = 256;
constexpr int kUnroll = 2;

struct SpinResult { bool ok; };

__device__ __forceinline__ bool spin_until_ge(const uint32_t* flag, uint32_t want, const uint32_t* abort_flag,
                                              unsigned long long timeout_ns, uint32_t* status, uint32_t code) {
  unsigned long long t0 = globaltimer_ns();
  unsigned int backoff = 32;
  while ((int32_t)(ld_acquire_sys_u32(flag) - want) < 0) {
    if (ld_acquire_sys_u32(abort_flag) != 0) { atomicMax(status, 3u); return false; }
    if (*reinterpret_cast<volatile uint32_t*>(status) != 0) return false;   // another waiter of this rank gave up
    if (globaltimer_ns() - t0 > timeout_ns) { atomicMax(status, code); return false; }
    __nanosleep(backoff);
    if (backoff < 1024) backoff <<= 1;
  }
  return true;
}

__device__ __forceinline__ float4 f4_add(float4 a, float4 b) {
  return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
__device__ __forceinline__ float4 f4_fma_s(float4 a, float s, float4 acc) {
  // product rounded before the add (matches "acc + s_i * w_i" evaluated by the oracle in fp32)
  return make_float4(acc.x + __fmul_rn(a.x, s), acc.y + __fmul_rn(a.y, s), acc.z + __fmul_rn(a.z, s),
                     acc.w + __fmul_rn(a.w, s));
}
__device__ __forceinline__ float4 f4_div(float4 a, float d) {
  return make_float4(__fdiv_rn(a.x, d), __fdiv_rn(a.y, d), __fdiv_rn(a.z, d), __fdiv_rn(a.w, d));
}
__device__ __forceinline__ float4 ld_src_f4(const float* base, long long i4) {
  float4 r;
  const float4* p = reinterpret_cast<const float4*>(base) + i4;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* mc, long long i4) {
  float4 r;
  const float4* p = reinterpret_cast<const float4*>(mc) + i4;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void mc_st_f4(float* mc, long long i4, float4 v) {
  float4* p = reinterpret_cast<float4*>(mc) + i4;
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 dequant4(const uint8_t* q, const float* scale, long long i4) {
  // 4 consecutive e4m3 values share one 32-element block scale
  uint32_t packed = *reinterpret_cast<const uint32_t*>(q + i4 * 4);
  float s = scale[(i4 * 4) >> 5];
  __nv_fp8x4_e4m3 v;
  v.__x = packed;
  float4 f = static_cast<float4>(v);
  return make_float4(__fmul_rn(f.x, s), __fmul_rn(f.y, s), __fmul_rn(f.z, s), __fmul_rn(f.w, s));   // no fma contraction
}
__device__ __forceinline__ uint2 pack_bf16x4(float4 v) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
  __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&lo);
  r.y = *reinterpret_cast<uint32_t*>(&hi);
  return r;
}

// kSrc > 0: the number of sources is a compile-time constant, so ALL kSrc * kUnroll 128-bit loads of a thread are issued
// before the first add (kSrc * kUnroll * 16 B in flight per thread hides the NVLink round trip); kSrc == 0: run-time count.
template <bool kNvls, bool kDecode, int kSrc>
__global__ void __launch_bounds__(kThreads, (kSrc > 4 ? 2 : 4)) fedavg_kernel(const __grid_constant__ fedb_agg_desc_t p) {
  const int tid = threadIdx.x;
  __shared__ uint32_t s_epoch;
  __shared__ int s_last;
  bool ok = p.gate == nullptr || ld_acquire_sys_u32(p.gate) == 0;

  // ---------------------------------------------------------------- (A) arrive + wait
  if (p.sync_mode == 1) {
    if (tid == 0) s_epoch = ld_acquire_sys_u32(&p.ctrl_local->epoch) + 1;
    __syncthreads();
    const uint32_t epoch = s_epoch;
    if (blockIdx.x == 0 && tid < p.world) {
      __threadfence_system();
      st_release_sys_u32(&p.ctrl_peer[tid]->arrive[p.rank], epoch);
    }
    bool mine = true;
    if (tid < p.world)
      mine = spin_until_ge(&p.ctrl_local->arrive[tid], epoch, &p.ctrl_local->abort_flag, p.timeout_ns,
                           &p.ctrl_local->status, 1u);
    // A failed wait (dead or late peer, abort) must leave EVERY arena untouched: the round did not happen.  The outcome is
    // block-wide (one vote), and a block that got through still yields to a failure another block of this rank recorded.
    ok = __syncthreads_and(mine ? 1 : 0) != 0;
    if (ok && ld_acquire_sys_u32(&p.ctrl_local->status) != 0) ok = false;
    if (!ok && blockIdx.x == 0 && tid < p.world) {
      // withdraw my announcement, so that a peer which shows up later cannot complete the round on its own
      st_release_sys_u32(&p.ctrl_peer[tid]->arrive[p.rank], epoch - 1);
    }
  }

  // ---------------------------------------------------------------- (B)+(C) reduce + fan-out
  const long long v_begin = p.begin >> 2, v_end = p.end >> 2;
  const long long stride = (long long)gridDim.x * kThreads;
  const int n_src = kSrc > 0 ? kSrc : p.n_src;
  const float n_div = (float)n_src;
  float wsum = 0.f;
  if (!p.uniform)
    for (int k = 0; k < n_src; ++k) wsum += p.weight[k];

  for (long long i0 = v_begin + (long long)blockIdx.x * kThreads + tid; ok && i0 < v_end; i0 += stride * kUnroll) {
    float4 acc[kUnroll];
    bool live[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) live[u] = i0 + u * stride < v_end;
    if constexpr (kSrc > 0 && !kNvls && !kDecode) {
      float4 v[kSrc][kUnroll];
#pragma unroll
      for (int k = 0; k < kSrc; ++k)
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
          if (live[u]) v[k][u] = ld_src_f4(p.src[k], i0 + u * stride);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        if (!live[u]) continue;
        acc[u] = p.uniform ? v[0][u] : f4_fma_s(v[0][u], p.weight[0], make_float4(0.f, 0.f, 0.f, 0.f));
#pragma unroll
        for (int k = 1; k < kSrc; ++k)
          acc[u] = p.uniform ? f4_add(acc[u], v[k][u]) : f4_fma_s(v[k][u], p.weight[k], acc[u]);
      }
    } else {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const long long i = i0 + u * stride;
        if (!live[u]) continue;
        if constexpr (kNvls) {
          acc[u] = mc_ld_reduce_f4(p.mc_src, i);
        } else if constexpr (kDecode) {
          acc[u] = dequant4(p.src_q[0], p.src_scale[0], i);
          if (!p.uniform) acc[u] = f4_fma_s(acc[u], p.weight[0], make_float4(0.f, 0.f, 0.f, 0.f));
        } else {
          acc[u] = ld_src_f4(p.src[0], i);
          if (!p.uniform) acc[u] = f4_fma_s(acc[u], p.weight[0], make_float4(0.f, 0.f, 0.f, 0.f));
        }
      }
      if constexpr (!kNvls) {
        for (int k = 1; k < n_src; ++k) {
          float4 v[kUnroll];
#pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            if (!live[u]) continue;
            const long long i = i0 + u * stride;
            if constexpr (kDecode) v[u] = dequant4(p.src_q[k], p.src_scale[k], i);
            else v[u] = ld_src_f4(p.src[k], i);
          }
#pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            if (!live[u]) continue;
            acc[u] = p.uniform ? f4_add(acc[u], v[u]) : f4_fma_s(v[u], p.weight[k], acc[u]);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (!live[u]) continue;
      const long long i = i0 + u * stride;
      float4 r = f4_div(acc[u], p.uniform ? n_div : wsum);
      if constexpr (kDecode) r = f4_add(ld_src_f4(p.base, i), r);   // mean(base + d_k) = base + mean(d_k)
      if (p.mc_dst != nullptr) mc_st_f4(p.mc_dst, i, r);
      for (int d = 0; d < p.n_dst; ++d) {
        st_na_f4(reinterpret_cast<float4*>(p.dst[d]) + i, r);
        if (p.dst_bf16[d] != nullptr) reinterpret_cast<uint2*>(p.dst_bf16[d])[i] = pack_bf16x4(r);
      }
    }
  }

  // int64 counters (N3): tiny, one block.  sum in int64, fp32 true division, truncate.
  if (ok && p.do_i64 && blockIdx.x == 0) {
    for (long long i = tid; i < p.n_i64; i += kThreads) {
      long long s = p.src_i64[0][i];
      for (int k = 1; k < n_src; ++k) s += p.src_i64[k][i];
      long long r = (long long)truncf(__fdiv_rn(__ll2float_rn(s), n_div));
      for (int d = 0; d < p.n_dst_i64; ++d)
        if (p.dst_i64[d] != nullptr) p.dst_i64[d][i] = r;
    }
  }

  // ---------------------------------------------------------------- (D) done + wait
  __threadfence_system();
  __syncthreads();
  if (tid == 0) {
    uint32_t prev = atomicAdd(&p.ctrl_local->blocks_done, 1u);
    s_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  // last block of this rank: every store of the grid is globally performed
  if (ok && p.sync_mode == 1 && ld_acquire_sys_u32(&p.ctrl_local->status) != 0) ok = false;
  if (ok && p.version != 0 && tid < p.n_dst && p.dst_ctrl[tid] != nullptr)
    st_release_sys_u32(&p.dst_ctrl[tid]->version, p.version);
  if (ok && p.sync_mode == 1) {
    const uint32_t epoch = s_epoch;
    if (tid < p.world) {
      st_release_sys_u32(&p.ctrl_peer[tid]->done[p.rank], epoch);
      spin_until_ge(&p.ctrl_local->done[tid], epoch, &p.ctrl_local->abort_flag, p.timeout_ns,
                    &p.ctrl_local->status, 2u);
    }
    __syncthreads();
    // the epoch only advances over a completed round; after a failure the host picks a fresh one (reset_sync)
    if (tid == 0 && ld_acquire_sys_u32(&p.ctrl_local->status) == 0) st_release_sys_u32(&p.ctrl_local->epoch, epoch);
  }
  if (tid == 0) {
    p.ctrl_local->blocks_done = 0;
    __threadfence();
  }
}

// ---------------------------------------------------------------------------- K-REPL
__global__ void __launch_bounds__(kThreads, 4)
replicate_kernel(float* __restrict__ dst, const float* __restrict__ src, long long n4,
                 int64_t* __restrict__ dst_i64, const int64_t* __restrict__ src_i64, long long n_i64,
                 fedb_ctrl_t* dst_ctrl, uint32_t version, uint32_t* counter) {
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n4; i += stride)
    st_na_f4(reinterpret_cast<float4*>(dst) + i, ld_src_f4(src, i));
  if (blockIdx.x == 0)
    for (long long i = threadIdx.x; i < n_i64; i += kThreads) dst_i64[i] = src_i64[i];
  __threadfence_system();
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1);
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    *counter = 0;
    if (dst_ctrl != nullptr) st_release_sys_u32(&dst_ctrl->version, version);   // replica is whole
  }
}

struct BarrierArgs { fedb_ctrl_t* peers[FEDB_MAX_PEERS]; };
__global__ void flag_barrier_kernel_v(fedb_ctrl_t* local, const __grid_constant__ BarrierArgs a, int rank,
                                      int world, unsigned long long timeout_ns) {
  const int tid = threadIdx.x;
  if (tid >= world) return;
  const uint32_t epoch = ld_acquire_sys_u32(&local->epoch) + 1;
  __threadfence_system();
  st_release_sys_u32(&a.peers[tid]->arrive[rank], epoch);
  spin_until_ge(&local->arrive[tid], epoch, &local->abort_flag, timeout_ns, &local->status, 1u);
  st_release_sys_u32(&a.peers[tid]->done[rank], epoch);
  spin_until_ge(&local->done[tid], epoch, &local->abort_flag, timeout_ns, &local->status, 2u);
  __syncwarp();
  if (tid == 0) st_release_sys_u32(&local->epoch, epoch);
}

}  // namespace

extern "C" int fedb_fedavg(const fedb_agg_desc_t* d, int max_blocks, void* stream) {
  FEDB_REQUIRE(d != nullptr, "fedb_fedavg: null descriptor");
  FEDB_REQUIRE(d->n_src >= 1 && d->n_src <= FEDB_MAX_PEERS, "fedb_fedavg: n_src out of range");
  FEDB_REQUIRE(d->n_dst >= 0 && d->n_dst <= FEDB_MAX_DST, "fedb_fedavg: n_dst out of range");
  FEDB_REQUIRE(d->n_dst_i64 >= 0 && d->n_dst_i64 <= FEDB_MAX_DST, "fedb_fedavg: n_dst_i64 out of range");
  FEDB_REQUIRE((d->begin % 4) == 0 && (d->end % 4) == 0 && d->begin <= d->end,
               "fedb_fedavg: range must be 4-element aligned");
  FEDB_REQUIRE(d->ctrl_local != nullptr, "fedb_fedavg: ctrl_local is required (grid completion counter)");
  FEDB_REQUIRE(d->sync_mode == 0 || (d->world >= 1 && d->world <= FEDB_MAX_PEERS), "fedb_fedavg: bad world");
  const bool decode = d->src_q[0] != nullptr;
  const bool nvls = d->mc_src != nullptr;
  FEDB_REQUIRE(!(decode && nvls), "fedb_fedavg: NVLS reduction needs raw fp32 sources");
  FEDB_REQUIRE(!nvls || d->uniform, "fedb_fedavg: NVLS reduction supports the uniform mean only");
  FEDB_REQUIRE(!decode || d->base != nullptr, "fedb_fedavg: compressed uploads need a base model");
  const long long n4 = (d->end - d->begin) / 4;
  int blocks = fedb_ceil_div(n4, (long long)kThreads * kUnroll);
  const bool plain = !nvls && !decode;
  const int per_sm = (plain && d->n_src > 4 && d->n_src <= 8) ? 2 : 4;   // matches the kernels' launch bounds
  const int cap = max_blocks > 0 ? max_blocks : fedb_num_sms() * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  cudaStream_t s = (cudaStream_t)stream;
  if (nvls) fedavg_kernel<true, false, 0><<<blocks, kThreads, 0, s>>>(*d);
  else if (decode) fedavg_kernel<false, true, 0><<<blocks, kThreads, 0, s>>>(*d);
  else {
    switch (d->n_src) {
#define FEDB_AGG_CASE(N) case N: fedavg_kernel<false, false, N><<<blocks, kThreads, 0, s>>>(*d); break;
      FEDB_AGG_CASE(1) FEDB_AGG_CASE(2) FEDB_AGG_CASE(3) FEDB_AGG_CASE(4)
      FEDB_AGG_CASE(5) FEDB_AGG_CASE(6) FEDB_AGG_CASE(7) FEDB_AGG_CASE(8)
#undef FEDB_AGG_CASE
      default: fedavg_kernel<false, false, 0><<<blocks, kThreads, 0, s>>>(*d);
    }
  }
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_replicate(float* dst, const float* src, int64_t n_f32, int64_t* dst_i64,
                              const int64_t* src_i64, int64_t n_i64, fedb_ctrl_t* dst_ctrl, uint32_t version,
                              void* stream) {
  FEDB_REQUIRE((n_f32 % 4) == 0, "fedb_replicate: n_f32 must be a multiple of 4");
  FEDB_REQUIRE(dst_ctrl != nullptr, "fedb_replicate: dst_ctrl required (completion counter + version)");
  int blocks = fedb_ceil_div(n_f32 / 4, kThreads * 4);
  if (blocks > fedb_num_sms() * 2) blocks = fedb_num_sms() * 2;
  if (blocks < 1) blocks = 1;
  // the replica's own blocks_done word doubles as this kernel's completion counter
  replicate_kernel<<<blocks, kThreads, 0, (cudaStream_t)stream>>>(dst, src, n_f32 / 4, dst_i64, src_i64, n_i64,
                                                                  dst_ctrl, version, &dst_ctrl->blocks_done);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_flag_barrier(fedb_ctrl_t* local, fedb_ctrl_t* const* peers, int rank, int world,
                                 uint64_t timeout_ns, void* stream) {
  FEDB_REQUIRE(world >= 1 && world <= FEDB_MAX_PEERS, "fedb_flag_barrier: bad world");
  BarrierArgs a;
  memset(&a, 0, sizeof(a));
  for (int i = 0; i < world; ++i) a.peers[i] = peers[i];
  flag_barrier_kernel_v<<<1, 32, 0, (cudaStream_t)stream>>>(local, a, rank, world, timeout_ns);
  FEDB_LAUNCH_OK();
  return 0;
}

// ---------------------------------------------------------------------------------------------- K-PUSH
// Bit-exact fan-out of a byte range into peer memory: dst[k][i] = src[i] for every destination, or ONE multimem.st per
// 16 bytes through the NVLS multicast alias (the switch replicates it, the bytes cross this GPU's links once).  No
// arithmetic touches the data, so it carries uint8 images as well as floats.  Used to all-gather the evaluation set:
// every rank uploads 1/N of it from its host and pushes that slice to everybody over NVLink, instead of N ranks
// pulling the whole set through their own PCIe links every round.  The caller orders the reads of the peers (K-AGG's
// arrive / done flags of the same round do that in the federation round).
struct PushArgs {
  const float4* src;
  float4* dst[FEDB_MAX_DST];
  float4* mc_dst;
  int n_dst;
  long long n4;
};

__global__ void __launch_bounds__(256) push_copy_kernel(PushArgs a) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < a.n4; i += (long long)gridDim.x * 256) {
    const float4 v = a.src[i];
    if (a.mc_dst != nullptr) {
      mc_st_f4(reinterpret_cast<float*>(a.mc_dst), i, v);
    } else {
      for (int k = 0; k < a.n_dst; ++k) a.dst[k][i] = v;
    }
  }
}

extern "C" int fedb_push_copy(const void* src, void* const* dst, int n_dst, void* mc_dst, uint64_t bytes, void* stream) {
  FEDB_REQUIRE(src != nullptr && bytes % 16 == 0, "fedb_push_copy: bytes must be a multiple of 16");
  FEDB_REQUIRE(mc_dst != nullptr || (n_dst >= 1 && n_dst <= FEDB_MAX_DST), "fedb_push_copy: 1..FEDB_MAX_DST destinations");
  FEDB_REQUIRE((((uintptr_t)src | (uintptr_t)mc_dst) & 15) == 0, "fedb_push_copy: pointers must be 16-byte aligned");
  if (bytes == 0) return 0;
  PushArgs a;
  a.src = (const float4*)src; a.mc_dst = (float4*)mc_dst; a.n_dst = mc_dst ? 0 : n_dst; a.n4 = (long long)(bytes / 16);
  for (int k = 0; k < FEDB_MAX_DST; ++k) {
    a.dst[k] = (!mc_dst && k < n_dst) ? (float4*)dst[k] : nullptr;
    FEDB_REQUIRE(((uintptr_t)a.dst[k] & 15) == 0, "fedb_push_copy: destinations must be 16-byte aligned");
  }
  long long blocks = (a.n4 + 255) / 256;
  if (blocks > fedb_num_sms() * 4) blocks = fedb_num_sms() * 4;
  push_copy_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  FEDB_LAUNCH_OK();
  return 0;
}
