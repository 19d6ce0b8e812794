// Depthwise RxR convolution, R in {3, 5, 7} (groups == Cin == Cout), forward / data gradient / weight gradient, NHWC fp32.
// MobileNet, MobileNetV2, EfficientNet, ShuffleNet(V2) spend most of their layers here (reference:
// src/models/mobilenet.py:14-17, mobilenetv2.py:20-22, efficientnet.py:60-67, shufflenetv2.py:33-36 - nn.Conv2d with
// groups=planes).  There is no reduction over channels, so this is a streaming op: the arithmetic intensity is 9 FMA
// per loaded element and the bound is memory / L1 bandwidth, not the tensor cores.
//
// Layout: a thread owns FOUR consecutive channels (one 128-bit load per tap) and walks a run of output pixels along
// the row with the nine filter taps of its channels in registers; a warp covers 32 consecutive channel quads (or several
// pixels when C < 128), so every global access is a whole 512-byte line.  The weight gradient keeps 9 x 4 accumulators
// per thread over a chunk of pixels, reduces the threads of a block through shared memory in a fixed order and leaves
// one partial row per block; a second kernel adds the partial rows in block order (deterministic, no atomics).
#include "common.cuh"
#include "../include/fedb200.h"

namespace {

struct DwShape {
  int N, H, W, C, P, Q, sh, sw, ph, pw;
};

// Filter taps of four channels.  3x3 and 5x5 filters live in registers for the whole pixel run; a 7x7 filter (49 x 4
// values) would not fit, so its taps are re-read through L1 at every use (the filter is a few KB and stays cached).
template <int R, bool kInRegs = (R <= 5)>
struct Taps;

template <int R>
struct Taps<R, true> {
  float4 t[R * R];
  // the taps of four consecutive channels are 4 * R * R CONTIGUOUS floats of the OIHW filter, starting at a multiple of
  // 16 bytes (c % 4 == 0): R * R 128-bit loads and a register transpose instead of 4 * R * R scalar loads (on the small
  // late maps a thread computes only a few pixels, so the tap loads are a large part of its work)
  __device__ __forceinline__ void load(const float* __restrict__ w, int c) {
    float f[4 * R * R];
    const float4* p = reinterpret_cast<const float4*>(w + (long long)c * R * R);
#pragma unroll
    for (int i = 0; i < R * R; ++i) {
      const float4 v = __ldg(p + i);
      f[4 * i] = v.x; f[4 * i + 1] = v.y; f[4 * i + 2] = v.z; f[4 * i + 3] = v.w;
    }
#pragma unroll
    for (int k = 0; k < R * R; ++k) t[k] = make_float4(f[k], f[R * R + k], f[2 * R * R + k], f[3 * R * R + k]);
  }
  __device__ __forceinline__ float4 get(int k) const { return t[k]; }
};

template <int R>
struct Taps<R, false> {
  const float* w_;
  int c_;
  __device__ __forceinline__ void load(const float* __restrict__ w, int c) { w_ = w; c_ = c; }
  __device__ __forceinline__ float4 get(int k) const {
    return make_float4(__ldg(w_ + (c_ + 0) * R * R + k), __ldg(w_ + (c_ + 1) * R * R + k), __ldg(w_ + (c_ + 2) * R * R + k),
                       __ldg(w_ + (c_ + 3) * R * R + k));
  }
};

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void fma4(float4& a, const float4& x, const float4& w) {
  a.x = fmaf(x.x, w.x, a.x);
  a.y = fmaf(x.y, w.y, a.y);
  a.z = fmaf(x.z, w.z, a.z);
  a.w = fmaf(x.w, w.w, a.w);
}

constexpr int kRun = 8;      // output pixels per thread along the row

// y[n,p,q,c] = bias[c] + sum_{r,s} x[n, p*sh - ph + r, q*sw - pw + s, c] * w[c][r][s]
template <int R>
__global__ void __launch_bounds__(256) dw_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w,
                                                    const float* __restrict__ bias, float* __restrict__ y, DwShape d) {
  const int C4 = d.C >> 2, runs = (d.Q + kRun - 1) / kRun;
  const long long total = (long long)d.N * d.P * runs * C4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int c = (int)(i % C4) << 2;
    long long t = i / C4;
    const int q0 = (int)(t % runs) * kRun;
    t /= runs;
    const int p = (int)(t % d.P), n = (int)(t / d.P);
    Taps<R> tap;
    tap.load(w, c);
    const float4 b = bias != nullptr ? ld4(bias + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    const int qe = min(q0 + kRun, d.Q);
    for (int q = q0; q < qe; ++q) {
      float4 acc = b;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int ih = p * d.sh - d.ph + r;
        if (ih < 0 || ih >= d.H) continue;
#pragma unroll
        for (int s = 0; s < R; ++s) {
          const int iw = q * d.sw - d.pw + s;
          if (iw < 0 || iw >= d.W) continue;
          fma4(acc, ld4(x + (((long long)n * d.H + ih) * d.W + iw) * d.C + c), tap.get(r * R + s));
        }
      }
      *reinterpret_cast<float4*>(y + (((long long)n * d.P + p) * d.Q + q) * d.C + c) = acc;
    }
  }
}

// dx[n,h,w,c] = sum_{r,s : (h+ph-r) % sh == 0, (w+pw-s) % sw == 0} dy[n,(h+ph-r)/sh,(w+pw-s)/sw,c] * w[c][r][s]
template <int R>
__global__ void __launch_bounds__(256) dw_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ w,
                                                      float* __restrict__ dx, DwShape d) {
  const int C4 = d.C >> 2, runs = (d.W + kRun - 1) / kRun;
  const long long total = (long long)d.N * d.H * runs * C4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int c = (int)(i % C4) << 2;
    long long t = i / C4;
    const int w0 = (int)(t % runs) * kRun;
    t /= runs;
    const int h = (int)(t % d.H), n = (int)(t / d.H);
    Taps<R> tap;
    tap.load(w, c);
    const int we = min(w0 + kRun, d.W);
    for (int wx = w0; wx < we; ++wx) {
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int r = 0; r < R; ++r) {
        const int hp = h + d.ph - r;
        if (hp < 0 || hp % d.sh != 0) continue;
        const int p = hp / d.sh;
        if (p >= d.P) continue;
#pragma unroll
        for (int s = 0; s < R; ++s) {
          const int wq = wx + d.pw - s;
          if (wq < 0 || wq % d.sw != 0) continue;
          const int q = wq / d.sw;
          if (q >= d.Q) continue;
          fma4(acc, ld4(dy + (((long long)n * d.P + p) * d.Q + q) * d.C + c), tap.get(r * R + s));
        }
      }
      *reinterpret_cast<float4*>(dx + (((long long)n * d.H + h) * d.W + wx) * d.C + c) = acc;
    }
  }
}

// dw[c][r][s] = sum_{n,p,q} dy[n,p,q,c] * x[n, p*sh - ph + r, q*sw - pw + s, c]
// grid = (channel tiles, pixel chunks, filter-row strips); block = TX channel quads x TY pixel lanes (TX * TY == 256).
// A block accumulates RB filter rows (RB * R taps x 4 channels in registers: 3x3 and 5x5 in one strip, 7x7 in two).
// part[chunk][tap][C]
template <int TX, int R, int RB>
__global__ void __launch_bounds__(256) dw_wgrad_partial_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                              float* __restrict__ part, DwShape d, long long pix_per_chunk) {
  constexpr int TY = 256 / TX;
  constexpr int T = R * R;
  __shared__ float4 red[TY][TX + 1];
  const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
  const int c4 = blockIdx.x * TX + tx;
  const bool live = c4 < (d.C >> 2);
  const int c = c4 << 2;
  const int r0 = blockIdx.z * RB;                                   // first filter row of this strip
  const long long npix = (long long)d.N * d.P * d.Q;
  const long long p0 = (long long)blockIdx.y * pix_per_chunk;
  const long long p1 = min(p0 + pix_per_chunk, npix);
  float4 acc[RB * R];
#pragma unroll
  for (int k = 0; k < RB * R; ++k) acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (live) {
    for (long long pix = p0 + ty; pix < p1; pix += TY) {
      const int q = (int)(pix % d.Q), p = (int)((pix / d.Q) % d.P), n = (int)(pix / ((long long)d.Q * d.P));
      const float4 g = ld4(dy + pix * d.C + c);
#pragma unroll
      for (int rr = 0; rr < RB; ++rr) {
        const int r = r0 + rr;
        const int ih = p * d.sh - d.ph + r;
        if (r >= R || ih < 0 || ih >= d.H) continue;
#pragma unroll
        for (int s = 0; s < R; ++s) {
          const int iw = q * d.sw - d.pw + s;
          if (iw < 0 || iw >= d.W) continue;
          fma4(acc[rr * R + s], ld4(x + (((long long)n * d.H + ih) * d.W + iw) * d.C + c), g);
        }
      }
    }
  }
  // fixed-order reduction over the TY pixel lanes, one tap at a time
#pragma unroll
  for (int k = 0; k < RB * R; ++k) {
    const int tap = r0 * R + k;
    red[ty][tx] = acc[k];
    __syncthreads();
    if (ty == 0 && live && tap < T) {
      float4 sum = red[0][tx];
      for (int j = 1; j < TY; ++j) {
        const float4 v = red[j][tx];
        sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
      }
      *reinterpret_cast<float4*>(part + ((long long)blockIdx.y * T + tap) * d.C + c) = sum;
    }
    __syncthreads();
  }
}

// dw[c][tap] += sum_chunk part[chunk][tap][c].  Block = 32 elements x 8 groups of chunks: group w adds chunks w, w+8, ...
// (four independent loads in flight), the eight group sums are combined in group order, so the result does not depend on
// scheduling.  (One thread per element walking all chunks - up to 592 - was a 29 us latency chain per layer.)
__global__ void __launch_bounds__(256) dw_wgrad_final_kernel(const float* __restrict__ part, float* __restrict__ dw, int chunks,
                                                            int C, int T) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;              // i = tap * C + c
  const long long TC = (long long)T * C;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  if (i < TC) {
    int k = w;
    for (; k + 24 < chunks; k += 32) {
      a0 += part[(long long)(k + 0) * TC + i];
      a1 += part[(long long)(k + 8) * TC + i];
      a2 += part[(long long)(k + 16) * TC + i];
      a3 += part[(long long)(k + 24) * TC + i];
    }
    for (; k < chunks; k += 8) a0 += part[(long long)k * TC + i];
  }
  red[w][lane] = (a0 + a1) + (a2 + a3);
  __syncthreads();
  if (w == 0 && i < TC) {
    float sum = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) sum += red[q][lane];
    const int tap = i / C, c = i - tap * C;
    dw[c * T + tap] += sum;
  }
}

template <int TX, int R, int RB>
void launch_wgrad(const float* x, const float* dy, float* scratch, const DwShape& d, long long ppc, dim3 grid, cudaStream_t s) {
  grid.z = (R + RB - 1) / RB;
  dw_wgrad_partial_kernel<TX, R, RB><<<grid, 256, 0, s>>>(x, dy, scratch, d, ppc);
}

template <int R, int RB>
void launch_wgrad_tx(int TX, const float* x, const float* dy, float* scratch, const DwShape& d, long long ppc, dim3 grid,
                     cudaStream_t s) {
  if (TX == 32) launch_wgrad<32, R, RB>(x, dy, scratch, d, ppc, grid, s);
  else if (TX == 16) launch_wgrad<16, R, RB>(x, dy, scratch, d, ppc, grid, s);
  else launch_wgrad<8, R, RB>(x, dy, scratch, d, ppc, grid, s);
}

DwShape mk(const int* s) {
  DwShape d;
  d.N = s[0]; d.H = s[1]; d.W = s[2]; d.C = s[3]; d.P = s[4]; d.Q = s[5]; d.sh = s[6]; d.sw = s[7]; d.ph = s[8]; d.pw = s[9];
  return d;
}

int grid_for(long long n) {
  long long b = (n + 255) / 256;
  const long long cap = (long long)fedb_num_sms() * 16;
  return (int)(b > cap ? cap : (b < 1 ? 1 : b));
}

}  // namespace

// shape10 = {N, H, W, C, P, Q, stride_h, stride_w, pad_h, pad_w}; fs x fs taps, fs in {3, 5, 7}; C % 4 == 0;
// w is [C][1][fs][fs]
extern "C" int fedb_dw_conv_fwd_k(const float* x, const float* w, const float* bias, float* y, const int* shape10, int fs,
                                  void* stream) {
  const DwShape d = mk(shape10);
  FEDB_REQUIRE(d.C % 4 == 0, "fedb_dw_conv: C must be a multiple of 4");
  FEDB_REQUIRE(fs == 3 || fs == 5 || fs == 7, "fedb_dw_conv: 3x3, 5x5 or 7x7");
  const long long items = (long long)d.N * d.P * ((d.Q + kRun - 1) / kRun) * (d.C / 4);
  cudaStream_t s = (cudaStream_t)stream;
  if (fs == 3) dw_fwd_kernel<3><<<grid_for(items), 256, 0, s>>>(x, w, bias, y, d);
  else if (fs == 5) dw_fwd_kernel<5><<<grid_for(items), 256, 0, s>>>(x, w, bias, y, d);
  else dw_fwd_kernel<7><<<grid_for(items), 256, 0, s>>>(x, w, bias, y, d);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_dw_conv_dgrad_k(const float* dy, const float* w, float* dx, const int* shape10, int fs, void* stream) {
  const DwShape d = mk(shape10);
  FEDB_REQUIRE(d.C % 4 == 0, "fedb_dw_conv: C must be a multiple of 4");
  FEDB_REQUIRE(fs == 3 || fs == 5 || fs == 7, "fedb_dw_conv: 3x3, 5x5 or 7x7");
  const long long items = (long long)d.N * d.H * ((d.W + kRun - 1) / kRun) * (d.C / 4);
  cudaStream_t s = (cudaStream_t)stream;
  if (fs == 3) dw_dgrad_kernel<3><<<grid_for(items), 256, 0, s>>>(dy, w, dx, d);
  else if (fs == 5) dw_dgrad_kernel<5><<<grid_for(items), 256, 0, s>>>(dy, w, dx, d);
  else dw_dgrad_kernel<7><<<grid_for(items), 256, 0, s>>>(dy, w, dx, d);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" long long fedb_dw_conv_wgrad_scratch_k(const int* shape10, int fs) {
  const DwShape d = mk(shape10);
  return (long long)fedb_num_sms() * 4 * fs * fs * d.C;       // upper bound on chunks * taps * C floats
}

// dw (+)= the weight gradient; scratch: fedb_dw_conv_wgrad_scratch_k(shape, fs) floats owned by the caller
extern "C" int fedb_dw_conv_wgrad_k(const float* x, const float* dy, float* dw, float* scratch, const int* shape10, int fs,
                                    void* stream) {
  const DwShape d = mk(shape10);
  FEDB_REQUIRE(d.C % 4 == 0 && scratch != nullptr, "fedb_dw_conv_wgrad: C % 4 and a scratch buffer are required");
  FEDB_REQUIRE(fs == 3 || fs == 5 || fs == 7, "fedb_dw_conv: 3x3, 5x5 or 7x7");
  const int C4 = d.C / 4;
  const int TX = C4 >= 32 ? 32 : (C4 >= 16 ? 16 : 8);
  const int ctiles = (C4 + TX - 1) / TX;
  const long long npix = (long long)d.N * d.P * d.Q;
  const int TY = 256 / TX;
  int chunks = fedb_num_sms() * 4 / ctiles;
  if (chunks < 1) chunks = 1;
  long long ppc = (npix + chunks - 1) / chunks;
  if (ppc < 4 * TY) ppc = 4 * TY;                          // at least four pixels per thread
  chunks = (int)((npix + ppc - 1) / ppc);
  dim3 grid(ctiles, chunks, 1);
  cudaStream_t s = (cudaStream_t)stream;
  if (fs == 3) launch_wgrad_tx<3, 3>(TX, x, dy, scratch, d, ppc, grid, s);
  else if (fs == 5) launch_wgrad_tx<5, 5>(TX, x, dy, scratch, d, ppc, grid, s);
  else launch_wgrad_tx<7, 4>(TX, x, dy, scratch, d, ppc, grid, s);
  dw_wgrad_final_kernel<<<fedb_ceil_div(fs * fs * d.C, 32), 256, 0, s>>>(scratch, dw, chunks, d.C, fs * fs);
  FEDB_LAUNCH_OK();
  return 0;
}

// 3x3 forms of the four entry points above
extern "C" int fedb_dw_conv_fwd(const float* x, const float* w, const float* bias, float* y, const int* shape10, void* stream) {
  return fedb_dw_conv_fwd_k(x, w, bias, y, shape10, 3, stream);
}
extern "C" int fedb_dw_conv_dgrad(const float* dy, const float* w, float* dx, const int* shape10, void* stream) {
  return fedb_dw_conv_dgrad_k(dy, w, dx, shape10, 3, stream);
}
extern "C" long long fedb_dw_conv_wgrad_scratch(const int* shape10) { return fedb_dw_conv_wgrad_scratch_k(shape10, 3); }
extern "C" int fedb_dw_conv_wgrad(const float* x, const float* dy, float* dw, float* scratch, const int* shape10, void* stream) {
  return fedb_dw_conv_wgrad_k(x, dy, dw, scratch, shape10, 3, stream);
}
