// Generic fp32 NHWC operators for the whole model zoo: every op the 44
// constructors use has a hand-written forward and backward here, so any zoo model trains on the GPU
// without a library kernel.  These are the BREADTH kernels - simple, correct for every
// (kernel, stride, padding, groups) combination in the zoo (1x1/3x3/5x5/7x7, depthwise, grouped,
// bias) - and the numerical cross-check for the tcgen05 paths (gemm_tc.cu / conv_tc.cu), which take
// over the GEMM-shaped layers.  Activations are NHWC (channels_last), weights stay OIHW as in the
// checkpoint; small transposed weight copies are made per call so that every global access is
// coalesced along the channel dimension.
#include "common.cuh"
#include "../include/fedb200.h"

namespace {

struct ConvShape {
  int N, H, W, C, K, R, S, P, Q, sh, sw, ph, pw, groups;
};

__device__ __forceinline__ float act_fwd(float z, int act) {
  if (act == 1) return fmaxf(z, 0.f);
  if (act == 2) return z / (1.f + expf(-z));          // swish
  if (act == 3) return 1.f / (1.f + expf(-z));        // sigmoid
  return z;
}
__device__ __forceinline__ float act_bwd(float z, int act) {   // d act / d z
  if (act == 1) return z > 0.f ? 1.f : 0.f;
  if (act == 2) { float s = 1.f / (1.f + expf(-z)); return s * (1.f + z * (1.f - s)); }
  if (act == 3) { float s = 1.f / (1.f + expf(-z)); return s * (1.f - s); }
  return 1.f;
}

// w[K][Cg][R][S] -> wt[R][S][Cg][K] (forward) and wt2[R][S][K][Cg] (dgrad)
__global__ void weight_transpose_kernel(const float* __restrict__ w, float* __restrict__ wt, float* __restrict__ wt2,
                                        int K, int Cg, int R, int S) {
  const int total = K * Cg * R * S;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int s = i % S, r = (i / S) % R, c = (i / (S * R)) % Cg, k = i / (S * R * Cg);
    const float v = w[i];
    if (wt) wt[((r * S + s) * Cg + c) * K + k] = v;
    if (wt2) wt2[((r * S + s) * K + k) * Cg + c] = v;
  }
}

__global__ void __launch_bounds__(256) conv_fwd_kernel(const float* __restrict__ x, const float* __restrict__ wt,
                                                      const float* __restrict__ bias, float* __restrict__ y,
                                                      ConvShape c) {
  const int Cg = c.C / c.groups, Kg = c.K / c.groups;
  const long long total = (long long)c.N * c.P * c.Q * c.K;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int k = (int)(i % c.K);
    long long t = i / c.K;
    const int q = (int)(t % c.Q); t /= c.Q;
    const int p = (int)(t % c.P);
    const int n = (int)(t / c.P);
    const int g = k / Kg;
    float acc = bias ? bias[k] : 0.f;
    for (int r = 0; r < c.R; ++r) {
      const int ih = p * c.sh - c.ph + r;
      if (ih < 0 || ih >= c.H) continue;
      for (int s = 0; s < c.S; ++s) {
        const int iw = q * c.sw - c.pw + s;
        if (iw < 0 || iw >= c.W) continue;
        const float* xp = x + (((long long)n * c.H + ih) * c.W + iw) * c.C + g * Cg;
        const float* wp = wt + ((long long)(r * c.S + s) * Cg) * c.K + k;
#pragma unroll 4
        for (int cc = 0; cc < Cg; ++cc) acc = fmaf(xp[cc], wp[(long long)cc * c.K], acc);
      }
    }
    y[i] = acc;
  }
}

__global__ void __launch_bounds__(256) conv_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ wt2,
                                                        float* __restrict__ dx, ConvShape c) {
  const int Cg = c.C / c.groups, Kg = c.K / c.groups;
  const long long total = (long long)c.N * c.H * c.W * c.C;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ch = (int)(i % c.C);
    long long t = i / c.C;
    const int w = (int)(t % c.W); t /= c.W;
    const int h = (int)(t % c.H);
    const int n = (int)(t / c.H);
    const int g = ch / Cg, cl = ch % Cg;
    float acc = 0.f;
    for (int r = 0; r < c.R; ++r) {
      const int pp = h + c.ph - r;
      if (pp < 0 || pp % c.sh) continue;
      const int p = pp / c.sh;
      if (p >= c.P) continue;
      for (int s = 0; s < c.S; ++s) {
        const int qq = w + c.pw - s;
        if (qq < 0 || qq % c.sw) continue;
        const int q = qq / c.sw;
        if (q >= c.Q) continue;
        const float* dyp = dy + (((long long)n * c.P + p) * c.Q + q) * c.K + g * Kg;
        const float* wp = wt2 + ((long long)(r * c.S + s) * c.K + g * Kg) * Cg + cl;
#pragma unroll 4
        for (int kk = 0; kk < Kg; ++kk) acc = fmaf(dyp[kk], wp[(long long)kk * Cg], acc);
      }
    }
    dx[i] = acc;
  }
}

// grid.x: chunks of kPix output pixels, grid.y: tiles of 256 weight elements; partial sums are
// combined with atomicAdd (dw is zeroed by the caller).
constexpr int kPix = 256;
__global__ void __launch_bounds__(256) conv_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                        float* __restrict__ dw, float* __restrict__ dbias,
                                                        ConvShape c) {
  const int Cg = c.C / c.groups, Kg = c.K / c.groups;
  const int nW = c.K * Cg * c.R * c.S;
  const int wi = blockIdx.y * 256 + threadIdx.x;
  const long long npix = (long long)c.N * c.P * c.Q;
  const long long p0 = (long long)blockIdx.x * kPix;
  const long long p1 = p0 + kPix < npix ? p0 + kPix : npix;
  if (wi < nW) {
    const int s = wi % c.S, r = (wi / c.S) % c.R, cl = (wi / (c.S * c.R)) % Cg, k = wi / (c.S * c.R * Cg);
    const int g = k / Kg;
    float acc = 0.f;
    for (long long pix = p0; pix < p1; ++pix) {
      const int q = (int)(pix % c.Q);
      const int p = (int)((pix / c.Q) % c.P);
      const int n = (int)(pix / ((long long)c.Q * c.P));
      const int ih = p * c.sh - c.ph + r, iw = q * c.sw - c.pw + s;
      if (ih < 0 || ih >= c.H || iw < 0 || iw >= c.W) continue;
      acc = fmaf(dy[pix * c.K + k], x[(((long long)n * c.H + ih) * c.W + iw) * c.C + g * Cg + cl], acc);
    }
    atomicAdd(dw + wi, acc);
  }
  if (dbias != nullptr && blockIdx.y == 0) {
    for (int k = threadIdx.x; k < c.K; k += 256) {
      float acc = 0.f;
      for (long long pix = p0; pix < p1; ++pix) acc += dy[pix * c.K + k];
      atomicAdd(dbias + k, acc);
    }
  }
}

// ------------------------------------------------------------------------------------ batch norm
// NHWC: element (row, c) at row*C + c.  Block = 32 channels x 8 row-lanes over ONE chunk of rows; grid =
// (C/32, chunks).  Partial sums go to scratch[chunk][2][C]; a finalize kernel combines the chunks in a fixed order
// (deterministic, no atomics) - with one block per 32 channels only C/32 SMs would stream the whole tensor.
__global__ void __launch_bounds__(256) bn_partial_kernel(const float* __restrict__ x, float* __restrict__ scratch, long long rows,
                                                        int C, long long rows_per_chunk) {
  __shared__ float s_sum[8][33], s_sq[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int ch = blockIdx.x * 32 + cx;
  const long long r0 = (long long)blockIdx.y * rows_per_chunk;
  const long long r1 = r0 + rows_per_chunk < rows ? r0 + rows_per_chunk : rows;
  float sum = 0.f, sq = 0.f;
  if (ch < C)
    for (long long r = r0 + ry; r < r1; r += 8) {
      const float v = x[r * C + ch];
      sum += v;
      sq = fmaf(v, v, sq);
    }
  s_sum[ry][cx] = sum;
  s_sq[ry][cx] = sq;
  __syncthreads();
  if (ry == 0 && ch < C) {
#pragma unroll
    for (int j = 1; j < 8; ++j) { sum += s_sum[j][cx]; sq += s_sq[j][cx]; }
    scratch[((long long)blockIdx.y * 2 + 0) * C + ch] = sum;
    scratch[((long long)blockIdx.y * 2 + 1) * C + ch] = sq;
  }
}

// 128-bit forms of the four streaming BN kernels (C % 4 == 0, which is every zoo layer): a thread owns four consecutive
// channels, a block is TX channel quads x 256/TX row lanes, every access is a whole 16-byte word and two rows are in
// flight per thread.  BatchNorm is pure memory traffic (8 passes over the activation per training step), so this is
// what decides its share of a zoo step.
__device__ __forceinline__ float4 ldf4(const float* p) { return *reinterpret_cast<const float4*>(p); }
// optional bf16 shadow of a result (the operand format of the tensor-core weight-gradient GEMMs): written by the
// kernel that PRODUCES the activation / the output gradient, so no separate cast pass reads it again
__device__ __forceinline__ void st_bf16x4(uint16_t* base, long long i4, float a, float b, float c, float d) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  uint2 r;
  r.x = *reinterpret_cast<uint32_t*>(&lo);
  r.y = *reinterpret_cast<uint32_t*>(&hi);
  reinterpret_cast<uint2*>(base)[i4] = r;
}

// What the LAST block of a channel group does with the reduced sums when the second pass is fused into this kernel (few
// chunks: the small late layers, where a separate finalize launch costs more than its work): forward -> mean, invstd and the
// running statistics; backward -> dbeta, dgamma.
struct BnFinal {
  int fused;                 // 0: the caller launches bn_finalize_kernel / bn_bwd_finalize_kernel
  float* out_a;              // fwd: mean      bwd: dbeta
  float* out_b;              // fwd: invstd    bwd: dgamma
  float* run_mean;
  float* run_var;
  float eps, momentum;
};
// completion counters of the fused second pass, one per channel group; zero between launches (the last block resets its own).
// One training stream per device uses them at a time (the engine's); evaluation-mode BatchNorm never does.
__device__ unsigned int g_bn_group_done[256];

template <int TX, bool kBwd>
__global__ void __launch_bounds__(256) bn_partial4_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                         const float* __restrict__ mean, const float* __restrict__ invstd,
                                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                                         float* __restrict__ scratch, long long rows, int C, int act,
                                                         long long rows_per_chunk, BnFinal fin) {
  constexpr int TY = 256 / TX;
  __shared__ float4 s_a[TY][TX], s_b[TY][TX];
  const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
  const int c = (blockIdx.x * TX + tx) * 4;
  const bool live = c < C;
  const long long r0 = (long long)blockIdx.y * rows_per_chunk;
  const long long r1 = r0 + rows_per_chunk < rows ? r0 + rows_per_chunk : rows;
  float4 sa = make_float4(0.f, 0.f, 0.f, 0.f), sb = sa;
  if (live) {
    float4 m = sa, is = sa, g = sa, b = sa;
    if (kBwd) { m = ldf4(mean + c); is = ldf4(invstd + c); g = ldf4(gamma + c); b = ldf4(beta + c); }
    auto add = [&](const float4& xv, const float4& dv) {
      if (!kBwd) {
        sa.x += xv.x; sa.y += xv.y; sa.z += xv.z; sa.w += xv.w;
        sb.x = fmaf(xv.x, xv.x, sb.x); sb.y = fmaf(xv.y, xv.y, sb.y); sb.z = fmaf(xv.z, xv.z, sb.z); sb.w = fmaf(xv.w, xv.w, sb.w);
      } else {
        const float xh[4] = {(xv.x - m.x) * is.x, (xv.y - m.y) * is.y, (xv.z - m.z) * is.z, (xv.w - m.w) * is.w};
        float d[4] = {dv.x, dv.y, dv.z, dv.w};
        if (act) {
          d[0] *= act_bwd(fmaf(xh[0], g.x, b.x), act); d[1] *= act_bwd(fmaf(xh[1], g.y, b.y), act);
          d[2] *= act_bwd(fmaf(xh[2], g.z, b.z), act); d[3] *= act_bwd(fmaf(xh[3], g.w, b.w), act);
        }
        sa.x += d[0]; sa.y += d[1]; sa.z += d[2]; sa.w += d[3];
        sb.x = fmaf(d[0], xh[0], sb.x); sb.y = fmaf(d[1], xh[1], sb.y); sb.z = fmaf(d[2], xh[2], sb.z); sb.w = fmaf(d[3], xh[3], sb.w);
      }
    };
    long long r = r0 + ty;
    for (; r + 3 * TY < r1; r += 4 * TY) {                  // four rows (eight loads in the backward form) in flight
      const float4 x0 = ldf4(x + r * C + c), x1 = ldf4(x + (r + TY) * C + c);
      const float4 x2 = ldf4(x + (r + 2 * TY) * C + c), x3 = ldf4(x + (r + 3 * TY) * C + c);
      float4 d0 = sa, d1 = sa, d2 = sa, d3 = sa;
      if (kBwd) {
        d0 = ldf4(dy + r * C + c); d1 = ldf4(dy + (r + TY) * C + c);
        d2 = ldf4(dy + (r + 2 * TY) * C + c); d3 = ldf4(dy + (r + 3 * TY) * C + c);
      }
      add(x0, d0);
      add(x1, d1);
      add(x2, d2);
      add(x3, d3);
    }
    for (; r < r1; r += TY) add(ldf4(x + r * C + c), kBwd ? ldf4(dy + r * C + c) : sa);
  }
  s_a[ty][tx] = sa;
  s_b[ty][tx] = sb;
  __syncthreads();
  if (ty == 0 && live) {
    for (int j = 1; j < TY; ++j) {
      const float4 a = s_a[j][tx], b2 = s_b[j][tx];
      sa.x += a.x; sa.y += a.y; sa.z += a.z; sa.w += a.w;
      sb.x += b2.x; sb.y += b2.y; sb.z += b2.z; sb.w += b2.w;
    }
    *reinterpret_cast<float4*>(scratch + ((long long)blockIdx.y * 2 + 0) * C + c) = sa;
    *reinterpret_cast<float4*>(scratch + ((long long)blockIdx.y * 2 + 1) * C + c) = sb;
  }
  if (!fin.fused) return;
  // ---- fused second pass: the block that finishes last for this channel group adds the chunks in chunk order
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&g_bn_group_done[blockIdx.x], 1u) == gridDim.y - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  constexpr int NCH = TX * 4, PER = 256 / NCH;                 // channels of the group, threads per channel
  float* ra = reinterpret_cast<float*>(&s_a[0][0]);            // [PER][NCH] partial sums (the arrays are free again)
  float* rb = reinterpret_cast<float*>(&s_b[0][0]);
  const int chl = threadIdx.x % NCH, part = threadIdx.x / NCH;
  const int ch = blockIdx.x * NCH + chl;
  const int chunks = gridDim.y;
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (ch < C) {
    int k = part;
    for (; k + PER < chunks; k += 2 * PER) {
      const float x0 = __ldcg(scratch + ((long long)k * 2 + 0) * C + ch), y0 = __ldcg(scratch + ((long long)k * 2 + 1) * C + ch);
      const float x1 = __ldcg(scratch + ((long long)(k + PER) * 2 + 0) * C + ch);
      const float y1 = __ldcg(scratch + ((long long)(k + PER) * 2 + 1) * C + ch);
      a0 += x0; b0 += y0; a1 += x1; b1 += y1;
    }
    if (k < chunks) { a0 += __ldcg(scratch + ((long long)k * 2 + 0) * C + ch); b0 += __ldcg(scratch + ((long long)k * 2 + 1) * C + ch); }
  }
  ra[part * NCH + chl] = a0 + a1;
  rb[part * NCH + chl] = b0 + b1;
  __syncthreads();
  if (part == 0 && ch < C) {
    float sum = 0.f, sq = 0.f;
#pragma unroll
    for (int q = 0; q < PER; ++q) { sum += ra[q * NCH + chl]; sq += rb[q * NCH + chl]; }
    if (!kBwd) {
      const float m = sum / (float)rows;
      float var = sq / (float)rows - m * m;
      var = var < 0.f ? 0.f : var;
      fin.out_a[ch] = m;
      fin.out_b[ch] = rsqrtf(var + fin.eps);
      if (fin.run_mean != nullptr) {
        const float unbiased = rows > 1 ? var * (float)rows / (float)(rows - 1) : var;
        fin.run_mean[ch] = (1.f - fin.momentum) * fin.run_mean[ch] + fin.momentum * m;
        fin.run_var[ch] = (1.f - fin.momentum) * fin.run_var[ch] + fin.momentum * unbiased;
      }
    } else {
      fin.out_a[ch] = sum;       // dbeta
      fin.out_b[ch] = sq;        // dgamma
    }
  }
  if (threadIdx.x == 0) g_bn_group_done[blockIdx.x] = 0;
}

__global__ void __launch_bounds__(256) bn_apply4_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                       const float* __restrict__ mean, const float* __restrict__ invstd,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       long long total4, int C, int act, uint16_t* __restrict__ y16) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total4; i += (long long)gridDim.x * 256) {
    const int c = (int)((i * 4) % C);
    const float4 v = ldf4(x + i * 4), m = ldf4(mean + c), is = ldf4(invstd + c), g = ldf4(gamma + c), b = ldf4(beta + c);
    float4 o;
    o.x = act_fwd(fmaf((v.x - m.x) * is.x, g.x, b.x), act);
    o.y = act_fwd(fmaf((v.y - m.y) * is.y, g.y, b.y), act);
    o.z = act_fwd(fmaf((v.z - m.z) * is.z, g.z, b.z), act);
    o.w = act_fwd(fmaf((v.w - m.w) * is.w, g.w, b.w), act);
    *reinterpret_cast<float4*>(y + i * 4) = o;
    if (y16 != nullptr) st_bf16x4(y16, i, o.x, o.y, o.z, o.w);
  }
}

__global__ void __launch_bounds__(256) bn_bwd_apply4_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                           float* __restrict__ dx, const float* __restrict__ mean,
                                                           const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                           const float* __restrict__ beta, const float* __restrict__ dgamma,
                                                           const float* __restrict__ dbeta, long long total4, int C,
                                                           float inv_rows, int act, int training, uint16_t* __restrict__ dx16) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total4; i += (long long)gridDim.x * 256) {
    const int c = (int)((i * 4) % C);
    const float4 xv = ldf4(x + i * 4), dv = ldf4(dy + i * 4);
    const float4 m = ldf4(mean + c), is = ldf4(invstd + c), g = ldf4(gamma + c), b = ldf4(beta + c);
    const float4 dg = ldf4(dgamma + c), db = ldf4(dbeta + c);
    const float xs[4] = {xv.x, xv.y, xv.z, xv.w}, ds[4] = {dv.x, dv.y, dv.z, dv.w};
    const float ms[4] = {m.x, m.y, m.z, m.w}, iss[4] = {is.x, is.y, is.z, is.w}, gs[4] = {g.x, g.y, g.z, g.w};
    const float bs[4] = {b.x, b.y, b.z, b.w}, dgs[4] = {dg.x, dg.y, dg.z, dg.w}, dbs[4] = {db.x, db.y, db.z, db.w};
    float o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float xh = (xs[k] - ms[k]) * iss[k];
      float d = ds[k];
      if (act) d *= act_bwd(fmaf(xh, gs[k], bs[k]), act);
      o[k] = training ? gs[k] * iss[k] * (d - dbs[k] * inv_rows - xh * dgs[k] * inv_rows) : gs[k] * iss[k] * d;
    }
    *reinterpret_cast<float4*>(dx + i * 4) = make_float4(o[0], o[1], o[2], o[3]);
    if (dx16 != nullptr) st_bf16x4(dx16, i, o[0], o[1], o[2], o[3]);
  }
}

// Second pass of the chunked reductions: block = 32 channels x 8 groups of chunks.  Group w adds chunks w, w+8, ... (four
// independent loads in flight), the eight group sums are combined in group order: the result does not depend on
// scheduling.  Returns the two per-channel sums to the threads of group 0.
__device__ __forceinline__ bool chunk_reduce2(const float* __restrict__ scratch, int chunks, int C, int* ch_out, float* a_out,
                                              float* b_out) {
  __shared__ float sa[8][32], sb[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int ch = blockIdx.x * 32 + lane;
  float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
  if (ch < C) {
    int k = w;
    for (; k + 24 < chunks; k += 32) {                     // four chunks = eight loads in flight per thread
      float xs[4], ys[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        xs[u] = scratch[((long long)(k + 8 * u) * 2 + 0) * C + ch];
        ys[u] = scratch[((long long)(k + 8 * u) * 2 + 1) * C + ch];
      }
      a0 += xs[0]; b0 += ys[0]; a1 += xs[1]; b1 += ys[1]; a0 += xs[2]; b0 += ys[2]; a1 += xs[3]; b1 += ys[3];
    }
    for (; k < chunks; k += 8) { a0 += scratch[((long long)k * 2 + 0) * C + ch]; b0 += scratch[((long long)k * 2 + 1) * C + ch]; }
  }
  sa[w][lane] = a0 + a1;
  sb[w][lane] = b0 + b1;
  __syncthreads();
  if (w != 0 || ch >= C) return false;
  float a = 0.f, b = 0.f;
#pragma unroll
  for (int q = 0; q < 8; ++q) { a += sa[q][lane]; b += sb[q][lane]; }
  *ch_out = ch; *a_out = a; *b_out = b;
  return true;
}

__global__ void __launch_bounds__(256) bn_finalize_kernel(const float* __restrict__ scratch, int chunks, float* __restrict__ mean,
                                                         float* __restrict__ invstd, float* __restrict__ run_mean,
                                                         float* __restrict__ run_var, long long rows, int C, float eps,
                                                         float momentum) {
  int ch;
  float sum, sq;
  if (!chunk_reduce2(scratch, chunks, C, &ch, &sum, &sq)) return;
  const float m = sum / (float)rows;
  float var = sq / (float)rows - m * m;
  var = var < 0.f ? 0.f : var;
  mean[ch] = m;
  invstd[ch] = rsqrtf(var + eps);
  if (run_mean != nullptr) {
    const float unbiased = rows > 1 ? var * (float)rows / (float)(rows - 1) : var;
    run_mean[ch] = (1.f - momentum) * run_mean[ch] + momentum * m;
    run_var[ch] = (1.f - momentum) * run_var[ch] + momentum * unbiased;
  }
}

// y = act(gamma * (x - mean) * invstd + beta); in eval mode mean/invstd come from running stats
__global__ void __launch_bounds__(256) bn_apply_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                      const float* __restrict__ mean, const float* __restrict__ invstd,
                                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                                      long long total, int C, int act) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ch = (int)(i % C);
    const float z = fmaf((x[i] - mean[ch]) * invstd[ch], gamma[ch], beta[ch]);
    y[i] = act_fwd(z, act);
  }
}

__global__ void bn_eval_prep_kernel(const float* __restrict__ run_mean, const float* __restrict__ run_var,
                                    float* __restrict__ mean, float* __restrict__ invstd, int C, float eps) {
  const int ch = blockIdx.x * blockDim.x + threadIdx.x;
  if (ch < C) { mean[ch] = run_mean[ch]; invstd[ch] = rsqrtf(run_var[ch] + eps); }
}

// pass 1 of the backward: per-channel sum(dz) and sum(dz * xhat), dz = dy * act'(z); chunked like the forward stats
__global__ void __launch_bounds__(256) bn_bwd_partial_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                            const float* __restrict__ mean, const float* __restrict__ invstd,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            float* __restrict__ scratch, long long rows, int C, int act,
                                                            long long rows_per_chunk) {
  __shared__ float s_a[8][33], s_b[8][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int ch = blockIdx.x * 32 + cx;
  const long long r0 = (long long)blockIdx.y * rows_per_chunk;
  const long long r1 = r0 + rows_per_chunk < rows ? r0 + rows_per_chunk : rows;
  float sa = 0.f, sb = 0.f;
  if (ch < C) {
    const float m = mean[ch], is = invstd[ch], g = gamma[ch], b = beta[ch];
    for (long long r = r0 + ry; r < r1; r += 8) {
      const float xh = (x[r * C + ch] - m) * is;
      float d = dy[r * C + ch];
      if (act) d *= act_bwd(fmaf(xh, g, b), act);
      sa += d;
      sb = fmaf(d, xh, sb);
    }
  }
  s_a[ry][cx] = sa;
  s_b[ry][cx] = sb;
  __syncthreads();
  if (ry == 0 && ch < C) {
#pragma unroll
    for (int j = 1; j < 8; ++j) { sa += s_a[j][cx]; sb += s_b[j][cx]; }
    scratch[((long long)blockIdx.y * 2 + 0) * C + ch] = sa;
    scratch[((long long)blockIdx.y * 2 + 1) * C + ch] = sb;
  }
}

__global__ void __launch_bounds__(256) bn_bwd_finalize_kernel(const float* __restrict__ scratch, int chunks,
                                                             float* __restrict__ dgamma, float* __restrict__ dbeta, int C) {
  int ch;
  float sa, sb;
  if (!chunk_reduce2(scratch, chunks, C, &ch, &sa, &sb)) return;
  dbeta[ch] = sa;
  dgamma[ch] = sb;
}

// pass 2: dx = gamma * invstd * (dz - mean(dz) - xhat * mean(dz * xhat))
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                          float* __restrict__ dx, const float* __restrict__ mean,
                                                          const float* __restrict__ invstd, const float* __restrict__ gamma,
                                                          const float* __restrict__ beta, const float* __restrict__ dgamma,
                                                          const float* __restrict__ dbeta, long long total, int C,
                                                          float inv_rows, int act, int training) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ch = (int)(i % C);
    const float is = invstd[ch], g = gamma[ch];
    const float xh = (x[i] - mean[ch]) * is;
    float d = dy[i];
    if (act) d *= act_bwd(fmaf(xh, g, beta[ch]), act);
    // eval mode: the statistics are constants, BN is a per-channel affine map
    dx[i] = training ? g * is * (d - dbeta[ch] * inv_rows - xh * dgamma[ch] * inv_rows) : g * is * d;
  }
}

// ------------------------------------------------------------------------------------ elementwise
__global__ void __launch_bounds__(256) act_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n, int act) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) y[i] = act_fwd(x[i], act);
}
__global__ void __launch_bounds__(256) act_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                     float* __restrict__ dx, long long n, int act) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256)
    dx[i] = dy[i] * act_bwd(x[i], act);
}
__global__ void __launch_bounds__(256) add_act_kernel(const float* __restrict__ a, const float* __restrict__ b,
                                                     float* __restrict__ y, long long n, int act) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256)
    y[i] = act_fwd(a[i] + b[i], act);
}
// 128-bit forms (n % 4 == 0, 16-byte aligned pointers): the elementwise ops are pure memory traffic
__global__ void __launch_bounds__(256) act_fwd4_kernel(const float4* __restrict__ x, float4* __restrict__ y, long long n4, int act) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = x[i];
    y[i] = make_float4(act_fwd(v.x, act), act_fwd(v.y, act), act_fwd(v.z, act), act_fwd(v.w, act));
  }
}
__global__ void __launch_bounds__(256) act_bwd4_kernel(const float4* __restrict__ x, const float4* __restrict__ dy,
                                                      float4* __restrict__ dx, long long n4, int act) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = x[i], d = dy[i];
    dx[i] = make_float4(d.x * act_bwd(v.x, act), d.y * act_bwd(v.y, act), d.z * act_bwd(v.z, act), d.w * act_bwd(v.w, act));
  }
}
__global__ void __launch_bounds__(256) add_act4_kernel(const float4* __restrict__ a, const float4* __restrict__ b,
                                                      float4* __restrict__ y, long long n4, int act,
                                                      uint16_t* __restrict__ y16) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 u = a[i], v = b[i];
    const float4 o = make_float4(act_fwd(u.x + v.x, act), act_fwd(u.y + v.y, act), act_fwd(u.z + v.z, act), act_fwd(u.w + v.w, act));
    y[i] = o;
    if (y16 != nullptr) st_bf16x4(y16, i, o.x, o.y, o.z, o.w);
  }
}
__device__ __host__ inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// y[n,hw,c] = x[n,hw,c] * s[n,c] (mode 0) ; per-sample scalar s[n] (mode 1: drop-connect / dropout masks)
__global__ void __launch_bounds__(256) scale_kernel(const float* __restrict__ x, const float* __restrict__ s,
                                                   float* __restrict__ y, long long total, long long per_sample, int C, int mode) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const long long n = i / per_sample;
    y[i] = x[i] * (mode == 0 ? s[n * C + (i % C)] : s[n]);
  }
}
__global__ void __launch_bounds__(256) scale4_kernel(const float* __restrict__ x, const float* __restrict__ s,
                                                    float* __restrict__ y, long long total4, long long per_sample, int C, int mode) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total4; i += (long long)gridDim.x * 256) {
    const long long e = i * 4, n = e / per_sample;
    const float4 v = *reinterpret_cast<const float4*>(x + e);
    float4 g;
    if (mode == 0) g = *reinterpret_cast<const float4*>(s + n * C + (e % C));
    else { const float f = s[n]; g = make_float4(f, f, f, f); }
    *reinterpret_cast<float4*>(y + e) = make_float4(v.x * g.x, v.y * g.y, v.z * g.z, v.w * g.w);
  }
}
// Per-(sample, channel) reductions over the HW positions of an NHWC map: block = 32 channels x 8 position lanes of one
// sample (grid = (N, C/32)), lanes stride over the positions, the eight lane sums are added in lane order.
//   kGate: ds[n,c] = sum_hw dy*x   (SE excitation backward wrt the gate)
//   else:  out[n,c] = scale * sum_hw x   (global average pooling)
template <bool kGate>
__global__ void __launch_bounds__(256) hw_reduce_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                       float* __restrict__ out, int HW, int C, float scale) {
  __shared__ float s_acc[8][33];
  const int cx = threadIdx.x & 31, ly = threadIdx.x >> 5;
  const int n = blockIdx.x, ch = blockIdx.y * 32 + cx;
  float acc = 0.f;
  if (ch < C) {
    const float* xp = x + (long long)n * HW * C + ch;
    const float* dp = kGate ? dy + (long long)n * HW * C + ch : nullptr;
    for (int p = ly; p < HW; p += 8) acc = kGate ? fmaf(xp[(long long)p * C], dp[(long long)p * C], acc) : acc + xp[(long long)p * C];
  }
  s_acc[ly][cx] = acc;
  __syncthreads();
  if (ly == 0 && ch < C) {
#pragma unroll
    for (int j = 1; j < 8; ++j) acc += s_acc[j][cx];
    out[(long long)n * C + ch] = acc * scale;
  }
}
// Philox-free counter hash: mask[i] = (u < keep) / keep
__global__ void __launch_bounds__(256) bernoulli_mask_kernel(float* __restrict__ m, long long n, float keep,
                                                            unsigned long long seed, const unsigned long long* __restrict__ ctr) {
  const unsigned long long base = seed ^ (ctr ? *ctr * 0x9E3779B97F4A7C15ull : 0ull);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
    unsigned long long z = base + (unsigned long long)i * 0xD1B54A32D192ED03ull + 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const float u = (float)(z >> 40) * (1.f / 16777216.f);
    m[i] = u < keep ? 1.f / keep : 0.f;
  }
}

// ------------------------------------------------------------------------------------ pooling
struct PoolShape { int N, H, W, C, P, Q, kh, kw, sh, sw, ph, pw; };

__global__ void __launch_bounds__(256) pool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                      int* __restrict__ arg, PoolShape s, int is_max) {
  const long long total = (long long)s.N * s.P * s.Q * s.C;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ch = (int)(i % s.C);
    long long t = i / s.C;
    const int q = (int)(t % s.Q); t /= s.Q;
    const int p = (int)(t % s.P);
    const int n = (int)(t / s.P);
    float best = -INFINITY, sum = 0.f;
    int besti = -1;
    for (int r = 0; r < s.kh; ++r) {
      const int ih = p * s.sh - s.ph + r;
      if (ih < 0 || ih >= s.H) continue;
      for (int c = 0; c < s.kw; ++c) {
        const int iw = q * s.sw - s.pw + c;
        if (iw < 0 || iw >= s.W) continue;
        const float v = x[(((long long)n * s.H + ih) * s.W + iw) * s.C + ch];
        sum += v;
        if (v > best) { best = v; besti = ih * s.W + iw; }
      }
    }
    if (is_max) { y[i] = best; if (arg) arg[i] = besti; }
    else y[i] = sum / (float)(s.kh * s.kw);          // count_include_pad=True, torch's default
  }
}

__global__ void __launch_bounds__(256) pool_bwd_kernel(const float* __restrict__ dy, const int* __restrict__ arg,
                                                      float* __restrict__ dx, PoolShape s, int is_max) {
  // gather form: one thread per INPUT element, deterministic
  const long long total = (long long)s.N * s.H * s.W * s.C;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ch = (int)(i % s.C);
    long long t = i / s.C;
    const int w = (int)(t % s.W); t /= s.W;
    const int h = (int)(t % s.H);
    const int n = (int)(t / s.H);
    float acc = 0.f;
    for (int r = 0; r < s.kh; ++r) {
      const int pp = h + s.ph - r;
      if (pp < 0 || pp % s.sh) continue;
      const int p = pp / s.sh;
      if (p >= s.P) continue;
      for (int c = 0; c < s.kw; ++c) {
        const int qq = w + s.pw - c;
        if (qq < 0 || qq % s.sw) continue;
        const int q = qq / s.sw;
        if (q >= s.Q) continue;
        const long long o = (((long long)n * s.P + p) * s.Q + q) * s.C + ch;
        if (is_max) { if (arg[o] == h * s.W + w) acc += dy[o]; }
        else acc += dy[o] / (float)(s.kh * s.kw);
      }
    }
    dx[i] = acc;
  }
}

// 128-bit forms of the two pooling kernels (C % 4 == 0): four channels per thread
__global__ void __launch_bounds__(256) pool_fwd4_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                       int* __restrict__ arg, PoolShape s, int is_max) {
  const int C4 = s.C >> 2;
  const long long total = (long long)s.N * s.P * s.Q * C4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ch = (int)(i % C4) << 2;
    long long t = i / C4;
    const int q = (int)(t % s.Q); t /= s.Q;
    const int p = (int)(t % s.P);
    const int n = (int)(t / s.P);
    float best[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY}, sum[4] = {0.f, 0.f, 0.f, 0.f};
    int besti[4] = {-1, -1, -1, -1};
    for (int r = 0; r < s.kh; ++r) {
      const int ih = p * s.sh - s.ph + r;
      if (ih < 0 || ih >= s.H) continue;
      for (int c = 0; c < s.kw; ++c) {
        const int iw = q * s.sw - s.pw + c;
        if (iw < 0 || iw >= s.W) continue;
        const float4 v4 = *reinterpret_cast<const float4*>(x + (((long long)n * s.H + ih) * s.W + iw) * s.C + ch);
        const float v[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          sum[k] += v[k];
          if (v[k] > best[k]) { best[k] = v[k]; besti[k] = ih * s.W + iw; }
        }
      }
    }
    const long long o = i * 4;
    if (is_max) {
      *reinterpret_cast<float4*>(y + o) = make_float4(best[0], best[1], best[2], best[3]);
      if (arg) *reinterpret_cast<int4*>(arg + o) = make_int4(besti[0], besti[1], besti[2], besti[3]);
    } else {
      const float inv = 1.f / (float)(s.kh * s.kw);          // count_include_pad=True, torch's default
      *reinterpret_cast<float4*>(y + o) = make_float4(sum[0] * inv, sum[1] * inv, sum[2] * inv, sum[3] * inv);
    }
  }
}

__global__ void __launch_bounds__(256) pool_bwd4_kernel(const float* __restrict__ dy, const int* __restrict__ arg,
                                                       float* __restrict__ dx, PoolShape s, int is_max) {
  const int C4 = s.C >> 2;
  const long long total = (long long)s.N * s.H * s.W * C4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int ch = (int)(i % C4) << 2;
    long long t = i / C4;
    const int w = (int)(t % s.W); t /= s.W;
    const int h = (int)(t % s.H);
    const int n = (int)(t / s.H);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const int me = h * s.W + w;
    for (int r = 0; r < s.kh; ++r) {
      const int pp = h + s.ph - r;
      if (pp < 0 || pp % s.sh) continue;
      const int p = pp / s.sh;
      if (p >= s.P) continue;
      for (int c = 0; c < s.kw; ++c) {
        const int qq = w + s.pw - c;
        if (qq < 0 || qq % s.sw) continue;
        const int q = qq / s.sw;
        if (q >= s.Q) continue;
        const long long o = (((long long)n * s.P + p) * s.Q + q) * s.C + ch;
        const float4 d = *reinterpret_cast<const float4*>(dy + o);
        if (is_max) {
          const int4 a = *reinterpret_cast<const int4*>(arg + o);
          if (a.x == me) acc[0] += d.x;
          if (a.y == me) acc[1] += d.y;
          if (a.z == me) acc[2] += d.z;
          if (a.w == me) acc[3] += d.w;
        } else {
          const float inv = 1.f / (float)(s.kh * s.kw);
          acc[0] += d.x * inv; acc[1] += d.y * inv; acc[2] += d.z * inv; acc[3] += d.w * inv;
        }
      }
    }
    *reinterpret_cast<float4*>(dx + i * 4) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  }
}

// ------------------------------------------------------------------------------------ channel moves (K11)
// out[row, dst_off + j] = in[row, src_off + map(j)] for j < width; map = identity or shuffle
__global__ void __launch_bounds__(256) channel_copy_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                          long long rows, int Cin, int Cout, int src_off, int dst_off,
                                                          int width, int shuffle_groups, int accumulate) {
  const long long total = rows * width;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int j = (int)(i % width);
    const long long row = i / width;
    int sj = j;
    if (shuffle_groups > 1) {         // out channel j = (cpg_idx * g + grp)  <-  in channel grp * cpg + cpg_idx
      const int cpg = width / shuffle_groups;
      sj = (j % shuffle_groups) * cpg + j / shuffle_groups;
    }
    const float v = in[row * Cin + src_off + sj];
    float* o = out + row * Cout + dst_off + j;
    *o = accumulate ? *o + v : v;
  }
}

// the same for channel windows that are whole 16-byte words on both sides (no shuffle): cat / slice of DenseNet, GoogLeNet,
// DPN, DLA move activations at memory speed
__global__ void __launch_bounds__(256) channel_copy4_kernel(const float* __restrict__ in, float* __restrict__ out, long long rows,
                                                           int Cin, int Cout, int src_off, int dst_off, int width4,
                                                           int accumulate) {
  const long long total = rows * width4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int j = (int)(i % width4);
    const long long row = i / width4;
    const float4 v = *reinterpret_cast<const float4*>(in + row * Cin + src_off + 4 * j);
    float4* o = reinterpret_cast<float4*>(out + row * Cout + dst_off + 4 * j);
    if (accumulate) {
      const float4 w = *o;
      *o = make_float4(w.x + v.x, w.y + v.y, w.z + v.z, w.w + v.w);
    } else {
      *o = v;
    }
  }
}

// ------------------------------------------------------------------------------------ small linear (classifier)
// y[b][j] = act(bias[j] + sum_k x[b][k] W[j][k]); one warp per output
__global__ void __launch_bounds__(256) linear_fwd_kernel(const float* __restrict__ x, const float* __restrict__ W,
                                                        const float* __restrict__ bias, float* __restrict__ y, int B,
                                                        int K, int J, int act) {
  const int lane = threadIdx.x & 31;
  const long long o = ((long long)blockIdx.x * 256 + threadIdx.x) >> 5;
  if (o >= (long long)B * J) return;
  const int j = (int)(o % J), b = (int)(o / J);
  float acc = 0.f;
  for (int k = lane; k < K; k += 32) acc = fmaf(x[(long long)b * K + k], W[(long long)j * K + k], acc);
  acc = warp_sum(acc);
  if (lane == 0) y[o] = act_fwd(acc + (bias ? bias[j] : 0.f), act);
}
// dx[b][k] = sum_j dy[b][j] W[j][k]
__global__ void __launch_bounds__(256) linear_dgrad_kernel(const float* __restrict__ dy, const float* __restrict__ W,
                                                          float* __restrict__ dx, int B, int K, int J) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i >= (long long)B * K) return;
  const int k = (int)(i % K), b = (int)(i / K);
  float acc = 0.f;
  for (int j = 0; j < J; ++j) acc = fmaf(dy[(long long)b * J + j], W[(long long)j * K + k], acc);
  dx[i] = acc;
}
// dW[j][k] = sum_b dy[b][j] x[b][k];  db[j] = sum_b dy[b][j]   (deterministic: fixed order over b)
__global__ void __launch_bounds__(256) linear_wgrad_kernel(const float* __restrict__ dy, const float* __restrict__ x,
                                                          float* __restrict__ dW, float* __restrict__ db, int B, int K, int J) {
  const long long i = (long long)blockIdx.x * 256 + threadIdx.x;
  if (i < (long long)J * K) {
    const int k = (int)(i % K), j = (int)(i / K);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;          // four independent chains: the loads of a batch chunk overlap
    int b = 0;
    for (; b + 3 < B; b += 4) {
      a0 = fmaf(dy[(long long)(b + 0) * J + j], x[(long long)(b + 0) * K + k], a0);
      a1 = fmaf(dy[(long long)(b + 1) * J + j], x[(long long)(b + 1) * K + k], a1);
      a2 = fmaf(dy[(long long)(b + 2) * J + j], x[(long long)(b + 2) * K + k], a2);
      a3 = fmaf(dy[(long long)(b + 3) * J + j], x[(long long)(b + 3) * K + k], a3);
    }
    for (; b < B; ++b) a0 = fmaf(dy[(long long)b * J + j], x[(long long)b * K + k], a0);
    dW[i] += (a0 + a1) + (a2 + a3);
  } else if (db != nullptr && i < (long long)J * K + J) {
    const int j = (int)(i - (long long)J * K);
    float acc = 0.f;
    for (int b = 0; b < B; ++b) acc += dy[(long long)b * J + j];
    db[j] += acc;
  }
}

inline int grid_for(long long n, int cap_mult = 16) {
  long long b = (n + 255) / 256;
  const long long cap = (long long)fedb_num_sms() * cap_mult;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace

#define S_ ((cudaStream_t)stream)

extern "C" int fedb_nn_weight_transpose(const float* w, float* wt, float* wt2, int K, int Cg, int R, int S, void* stream) {
  weight_transpose_kernel<<<grid_for((long long)K * Cg * R * S), 256, 0, S_>>>(w, wt, wt2, K, Cg, R, S);
  FEDB_LAUNCH_OK();
  return 0;
}

static ConvShape mk(const int* d) {
  ConvShape c;
  c.N = d[0]; c.H = d[1]; c.W = d[2]; c.C = d[3]; c.K = d[4]; c.R = d[5]; c.S = d[6]; c.P = d[7]; c.Q = d[8];
  c.sh = d[9]; c.sw = d[10]; c.ph = d[11]; c.pw = d[12]; c.groups = d[13];
  return c;
}

extern "C" int fedb_nn_conv_fwd(const float* x, const float* wt, const float* bias, float* y, const int* shape14, void* stream) {
  const ConvShape c = mk(shape14);
  FEDB_REQUIRE(c.C % c.groups == 0 && c.K % c.groups == 0, "conv: channels not divisible by groups");
  conv_fwd_kernel<<<grid_for((long long)c.N * c.P * c.Q * c.K), 256, 0, S_>>>(x, wt, bias, y, c);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_conv_dgrad(const float* dy, const float* wt2, float* dx, const int* shape14, void* stream) {
  const ConvShape c = mk(shape14);
  conv_dgrad_kernel<<<grid_for((long long)c.N * c.H * c.W * c.C), 256, 0, S_>>>(dy, wt2, dx, c);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_conv_wgrad(const float* x, const float* dy, float* dw, float* dbias, const int* shape14, void* stream) {
  const ConvShape c = mk(shape14);
  const int nW = c.K * (c.C / c.groups) * c.R * c.S;
  dim3 grid(fedb_ceil_div((long long)c.N * c.P * c.Q, kPix), fedb_ceil_div(nW, 256));
  conv_wgrad_kernel<<<grid, 256, 0, S_>>>(x, dy, dw, dbias, c);
  FEDB_LAUNCH_OK();
  return 0;
}

// how many row chunks for a [rows x C] reduction so that ~4 blocks land on every SM (bounded by the scratch size)
static int bn_tx(int C) { const int c4 = C / 4; return c4 >= 32 ? 32 : (c4 >= 16 ? 16 : 8); }

constexpr int kBnFuseMaxChunks = 64;      // up to this many chunks the last block of a channel group does the second pass itself

// returns true when the second pass ran inside the kernel (the caller then skips its finalize launch)
template <bool kBwd>
static bool bn_partial4(const float* x, const float* dy, const float* mean, const float* invstd, const float* gamma,
                        const float* beta, float* scratch, long long rows, int C, int act, long long rpc, int chunks,
                        BnFinal fin, cudaStream_t s) {
  const int tx = bn_tx(C);
  dim3 grid(fedb_ceil_div(C / 4, tx), chunks);
  fin.fused = (chunks <= kBnFuseMaxChunks && grid.x <= 256 && fin.out_a != nullptr) ? 1 : 0;
  if (tx == 32) bn_partial4_kernel<32, kBwd><<<grid, 256, 0, s>>>(x, dy, mean, invstd, gamma, beta, scratch, rows, C, act, rpc, fin);
  else if (tx == 16) bn_partial4_kernel<16, kBwd><<<grid, 256, 0, s>>>(x, dy, mean, invstd, gamma, beta, scratch, rows, C, act, rpc, fin);
  else bn_partial4_kernel<8, kBwd><<<grid, 256, 0, s>>>(x, dy, mean, invstd, gamma, beta, scratch, rows, C, act, rpc, fin);
  return fin.fused != 0;
}

static int bn_chunks(long long rows, int C, long long scratch_floats, long long* rows_per_chunk) {
  const int cgroups = C % 4 == 0 ? fedb_ceil_div(C / 4, bn_tx(C)) : fedb_ceil_div(C, 32);
  long long want = (long long)fedb_num_sms() * 4 / cgroups;
  if (want < 1) want = 1;
  const long long by_rows = (rows + 63) / 64;                  // at least 64 rows per chunk
  if (want > by_rows) want = by_rows;
  const long long by_scratch = scratch_floats / (2LL * C);
  if (want > by_scratch) want = by_scratch;
  if (want < 1) want = 1;
  *rows_per_chunk = (rows + want - 1) / want;
  return (int)((rows + *rows_per_chunk - 1) / *rows_per_chunk);
}

extern "C" int fedb_nn_bn_fwd16(const float* x, float* y, float* mean, float* invstd, float* run_mean, float* run_var,
                                const float* gamma, const float* beta, long long rows, int C, float eps, float momentum,
                                int training, int act, float* scratch, long long scratch_floats, uint16_t* y_bf16, void* stream) {
  FEDB_REQUIRE(y_bf16 == nullptr || C % 4 == 0, "fedb_nn_bn_fwd16: the bf16 shadow needs C % 4 == 0");
  if (training) {
    FEDB_REQUIRE(scratch != nullptr && scratch_floats >= 2LL * C, "fedb_nn_bn_fwd: scratch of at least 2*C floats required");
    long long rpc = 0;
    const int chunks = bn_chunks(rows, C, scratch_floats, &rpc);
    bool done = false;
    if (C % 4 == 0) {
      BnFinal fin = {0, mean, invstd, run_mean, run_var, eps, momentum};
      done = bn_partial4<false>(x, nullptr, nullptr, nullptr, nullptr, nullptr, scratch, rows, C, 0, rpc, chunks, fin, S_);
    } else bn_partial_kernel<<<dim3(fedb_ceil_div(C, 32), chunks), 256, 0, S_>>>(x, scratch, rows, C, rpc);
    if (!done)
      bn_finalize_kernel<<<fedb_ceil_div(C, 32), 256, 0, S_>>>(scratch, chunks, mean, invstd, run_mean, run_var, rows, C, eps, momentum);
  }
  else bn_eval_prep_kernel<<<fedb_ceil_div(C, 128), 128, 0, S_>>>(run_mean, run_var, mean, invstd, C, eps);
  if (C % 4 == 0) bn_apply4_kernel<<<grid_for(rows * C / 4), 256, 0, S_>>>(x, y, mean, invstd, gamma, beta, rows * C / 4, C, act, y_bf16);
  else bn_apply_kernel<<<grid_for(rows * C), 256, 0, S_>>>(x, y, mean, invstd, gamma, beta, rows * C, C, act);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_bn_fwd(const float* x, float* y, float* mean, float* invstd, float* run_mean, float* run_var,
                              const float* gamma, const float* beta, long long rows, int C, float eps, float momentum,
                              int training, int act, float* scratch, long long scratch_floats, void* stream) {
  return fedb_nn_bn_fwd16(x, y, mean, invstd, run_mean, run_var, gamma, beta, rows, C, eps, momentum, training, act, scratch,
                          scratch_floats, nullptr, stream);
}
extern "C" int fedb_nn_bn_bwd16(const float* x, const float* dy, float* dx, const float* mean, const float* invstd,
                                const float* gamma, const float* beta, float* dgamma, float* dbeta, long long rows, int C,
                                int act, int training, float* scratch, long long scratch_floats, uint16_t* dx_bf16, void* stream) {
  FEDB_REQUIRE(dx_bf16 == nullptr || C % 4 == 0, "fedb_nn_bn_bwd16: the bf16 shadow needs C % 4 == 0");
  FEDB_REQUIRE(scratch != nullptr && scratch_floats >= 2LL * C, "fedb_nn_bn_bwd: scratch of at least 2*C floats required");
  long long rpc = 0;
  const int chunks = bn_chunks(rows, C, scratch_floats, &rpc);
  bool done = false;
  if (C % 4 == 0) {
    BnFinal fin = {0, dbeta, dgamma, nullptr, nullptr, 0.f, 0.f};
    done = bn_partial4<true>(x, dy, mean, invstd, gamma, beta, scratch, rows, C, act, rpc, chunks, fin, S_);
  } else bn_bwd_partial_kernel<<<dim3(fedb_ceil_div(C, 32), chunks), 256, 0, S_>>>(x, dy, mean, invstd, gamma, beta, scratch, rows, C, act, rpc);
  if (!done) bn_bwd_finalize_kernel<<<fedb_ceil_div(C, 32), 256, 0, S_>>>(scratch, chunks, dgamma, dbeta, C);
  if (C % 4 == 0)
    bn_bwd_apply4_kernel<<<grid_for(rows * C / 4), 256, 0, S_>>>(x, dy, dx, mean, invstd, gamma, beta, dgamma, dbeta, rows * C / 4,
                                                                C, 1.f / (float)rows, act, training, dx_bf16);
  else
    bn_bwd_apply_kernel<<<grid_for(rows * C), 256, 0, S_>>>(x, dy, dx, mean, invstd, gamma, beta, dgamma, dbeta, rows * C, C,
                                                           1.f / (float)rows, act, training);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_nn_bn_bwd(const float* x, const float* dy, float* dx, const float* mean, const float* invstd,
                              const float* gamma, const float* beta, float* dgamma, float* dbeta, long long rows, int C,
                              int act, int training, float* scratch, long long scratch_floats, void* stream) {
  return fedb_nn_bn_bwd16(x, dy, dx, mean, invstd, gamma, beta, dgamma, dbeta, rows, C, act, training, scratch, scratch_floats,
                          nullptr, stream);
}

extern "C" int fedb_nn_act_fwd(const float* x, float* y, long long n, int act, void* stream) {
  if (n % 4 == 0 && aligned16(x) && aligned16(y))
    act_fwd4_kernel<<<grid_for(n / 4), 256, 0, S_>>>((const float4*)x, (float4*)y, n / 4, act);
  else
    act_fwd_kernel<<<grid_for(n), 256, 0, S_>>>(x, y, n, act);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_act_bwd(const float* x, const float* dy, float* dx, long long n, int act, void* stream) {
  if (n % 4 == 0 && aligned16(x) && aligned16(dy) && aligned16(dx))
    act_bwd4_kernel<<<grid_for(n / 4), 256, 0, S_>>>((const float4*)x, (const float4*)dy, (float4*)dx, n / 4, act);
  else
    act_bwd_kernel<<<grid_for(n), 256, 0, S_>>>(x, dy, dx, n, act);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_add_act16(const float* a, const float* b, float* y, long long n, int act, uint16_t* y_bf16, void* stream) {
  const bool vec = n % 4 == 0 && aligned16(a) && aligned16(b) && aligned16(y);
  FEDB_REQUIRE(y_bf16 == nullptr || vec, "fedb_nn_add_act16: the bf16 shadow needs n % 4 == 0 and 16-byte aligned tensors");
  if (vec)
    add_act4_kernel<<<grid_for(n / 4), 256, 0, S_>>>((const float4*)a, (const float4*)b, (float4*)y, n / 4, act, y_bf16);
  else
    add_act_kernel<<<grid_for(n), 256, 0, S_>>>(a, b, y, n, act);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_add_act(const float* a, const float* b, float* y, long long n, int act, void* stream) {
  return fedb_nn_add_act16(a, b, y, n, act, nullptr, stream);
}
extern "C" int fedb_nn_scale(const float* x, const float* s, float* y, long long total, long long per_sample, int C, int mode,
                             void* stream) {
  if (C % 4 == 0 && per_sample % 4 == 0 && total % 4 == 0 && aligned16(x) && aligned16(y) && (mode != 0 || aligned16(s)))
    scale4_kernel<<<grid_for(total / 4), 256, 0, S_>>>(x, s, y, total / 4, per_sample, C, mode);
  else
    scale_kernel<<<grid_for(total), 256, 0, S_>>>(x, s, y, total, per_sample, C, mode);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_scale_bwd_gate(const float* x, const float* dy, float* ds, int HW, int C, int N, void* stream) {
  hw_reduce_kernel<true><<<dim3(N, fedb_ceil_div(C, 32)), 256, 0, S_>>>(x, dy, ds, HW, C, 1.f);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_bernoulli_mask(float* m, long long n, float keep, uint64_t seed, const uint64_t* ctr_dev, void* stream) {
  bernoulli_mask_kernel<<<grid_for(n), 256, 0, S_>>>(m, n, keep, seed, (const unsigned long long*)ctr_dev);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_pool_fwd(const float* x, float* y, int* arg, const int* shape12, int is_max, void* stream) {
  PoolShape s;
  const int* d = shape12;
  s.N = d[0]; s.H = d[1]; s.W = d[2]; s.C = d[3]; s.P = d[4]; s.Q = d[5]; s.kh = d[6]; s.kw = d[7]; s.sh = d[8]; s.sw = d[9];
  s.ph = d[10]; s.pw = d[11];
  if (!is_max && s.P == 1 && s.Q == 1 && s.ph == 0 && s.pw == 0 && s.kh == s.H && s.kw == s.W && s.H * s.W >= 16)
    // global average pooling (SE squeeze): a parallel reduction over the positions instead of one serial window per thread
    hw_reduce_kernel<false><<<dim3(s.N, fedb_ceil_div(s.C, 32)), 256, 0, S_>>>(x, nullptr, y, s.H * s.W, s.C, 1.f / (float)(s.H * s.W));
  else if (s.C % 4 == 0 && aligned16(x) && aligned16(y) && aligned16(arg))
    pool_fwd4_kernel<<<grid_for((long long)s.N * s.P * s.Q * (s.C / 4)), 256, 0, S_>>>(x, y, arg, s, is_max);
  else
    pool_fwd_kernel<<<grid_for((long long)s.N * s.P * s.Q * s.C), 256, 0, S_>>>(x, y, arg, s, is_max);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_pool_bwd(const float* dy, const int* arg, float* dx, const int* shape12, int is_max, void* stream) {
  PoolShape s;
  const int* d = shape12;
  s.N = d[0]; s.H = d[1]; s.W = d[2]; s.C = d[3]; s.P = d[4]; s.Q = d[5]; s.kh = d[6]; s.kw = d[7]; s.sh = d[8]; s.sw = d[9];
  s.ph = d[10]; s.pw = d[11];
  if (s.C % 4 == 0 && aligned16(dy) && aligned16(dx) && aligned16(arg))
    pool_bwd4_kernel<<<grid_for((long long)s.N * s.H * s.W * (s.C / 4)), 256, 0, S_>>>(dy, arg, dx, s, is_max);
  else
    pool_bwd_kernel<<<grid_for((long long)s.N * s.H * s.W * s.C), 256, 0, S_>>>(dy, arg, dx, s, is_max);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_channel_copy(const float* in, float* out, long long rows, int Cin, int Cout, int src_off, int dst_off,
                                    int width, int shuffle_groups, int accumulate, void* stream) {
  if (shuffle_groups <= 1 && ((Cin | Cout | src_off | dst_off | width) & 3) == 0 && aligned16(in) && aligned16(out))
    channel_copy4_kernel<<<grid_for(rows * (width / 4)), 256, 0, S_>>>(in, out, rows, Cin, Cout, src_off, dst_off, width / 4,
                                                                      accumulate);
  else
    channel_copy_kernel<<<grid_for(rows * width), 256, 0, S_>>>(in, out, rows, Cin, Cout, src_off, dst_off, width,
                                                               shuffle_groups, accumulate);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_linear_fwd(const float* x, const float* W, const float* bias, float* y, int B, int K, int J, int act,
                                  void* stream) {
  linear_fwd_kernel<<<fedb_ceil_div((long long)B * J * 32, 256), 256, 0, S_>>>(x, W, bias, y, B, K, J, act);
  FEDB_LAUNCH_OK();
  return 0;
}
extern "C" int fedb_nn_linear_bwd(const float* dy, const float* x, const float* W, float* dx, float* dW, float* db, int B, int K,
                                  int J, void* stream) {
  if (dx) linear_dgrad_kernel<<<fedb_ceil_div((long long)B * K, 256), 256, 0, S_>>>(dy, W, dx, B, K, J);
  if (dW) linear_wgrad_kernel<<<fedb_ceil_div((long long)J * K + J, 256), 256, 0, S_>>>(dy, x, dW, db, B, K, J);
  FEDB_LAUNCH_OK();
  return 0;
}
