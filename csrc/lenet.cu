// LeNet training step in TWO kernels (the minimum training slice; K1,K7,K8,K9,K10,K13,K14).
//
//   S  sample kernel   one CTA per sample, the WHOLE forward and backward of that sample in shared memory:
//                      u8 image -> normalise (+crop/flip) -> conv1+ReLU+pool -> conv2+ReLU+pool -> fc1/fc2/fc3 ->
//                      softmax-CE -> backward through the three linears (data path) -> un-pool -> conv2 wgrad +
//                      sparse dgrad -> conv1 wgrad.  Leaves per-sample rows in scratch: conv partial gradients,
//                      (flat, h1, h2) and (dh1, dh2, dlogits).  No cross-sample dependency, no atomics.
//   U  update kernel   fc weight gradients as fixed-order batch reductions dW = sum_s dout[s] (x) in[s] (8 x 128 tiles,
//                      4 rows per thread), conv / bias partial reductions, then the SGD update (K10) in place.
//   E  eval kernel     forward + loss / correct of five samples per CTA, two CTAs per SM (the per-round evaluation).
//
// fp32 on CUDA cores on purpose: N_out is 6/16/10 and K is 75/150/400, far below a tcgen05 tile, and fp32 keeps
// the step comparable bit-for-tolerance with the reference's CPU run (src/models/lenet.py:5-23,
// src/main.py:142-151).  Every summation order is fixed, so a step is reproducible run to run.
// Weight reads of the classifier come from L2 (192 KB fc1 matrix, read twice per sample); all loads of a phase
// are issued before their first use so one phase costs about one memory latency, not one per element.
#include "common.cuh"
#include "../include/fedb200.h"
#include "augment.cuh"

namespace {

constexpr int C1 = 6, C2 = 16;
constexpr int P1 = 14, P2 = 5;                 // pooled sizes
constexpr int NP1 = C1 * P1 * P1;              // 1176
constexpr int NFLAT = C2 * P2 * P2;            // 400
constexpr int H1 = 120, H2 = 84, NCLS = 10;
constexpr int NW1 = C1 * 3 * 25, NW2 = C2 * C1 * 25;   // 450, 2400
constexpr int CONV_PART = 2880;                // [w1 450 | b1 6 | w2 2400 | b2 16] padded
constexpr int OFF_B1 = 450, OFF_W2 = 456, OFF_B2 = 2856;
// Image planes of the TRAINING kernel are padded in shared memory: row stride 38, plane stride 1218.  The weight
// gradient of conv1 reads img[ci][y+r][x+s] with lanes spread over (r, ci): banks 6r + 2ci are all distinct, whereas
// the natural 32/1024 strides put every (r, ci) of a tap column into the same bank (measured: 35 % of the kernel).
constexpr int IW = 38, IPLANE = 32 * IW + 2;
// The pooled conv1 maps (conv2's input) are padded likewise: row stride 22, plane stride 334.  Both the 6x6 patch loads
// of conv2's forward (lanes = pooled positions) and the filter-row loads of its weight gradient (lanes = (r, ci)) then
// cost two wavefronts instead of three with the dense 14 / 196 strides.
constexpr int P1S = 22, P1PLANE = 334;
constexpr int kThreads = 512;          // sample kernel: 16 warps per sample keep the issue slots fed
constexpr int kWarps = kThreads / 32;

struct Workspace {
  float* flat;      // [B][400]
  float* h1;        // [B][120]
  float* h2;        // [B][84]
  float* dlogits;   // [B][10]
  float* dh2;       // [B][84]
  float* dh1;       // [B][120]
  float* loss_rows; // [B]
  float* corr_rows; // [B]
  float* conv_part; // [B][2880]
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__host__ size_t carve(Workspace* w, void* base, int B) {
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  char* b = (char*)base;
#define FEDB_TAKE_F(field, n) { size_t o = take(sizeof(float) * (size_t)(n)); if (w) w->field = (float*)(b + o); }
  FEDB_TAKE_F(flat, (size_t)B * NFLAT)
  FEDB_TAKE_F(h1, (size_t)B * H1)
  FEDB_TAKE_F(h2, (size_t)B * H2)
  FEDB_TAKE_F(dlogits, (size_t)B * NCLS)
  FEDB_TAKE_F(dh2, (size_t)B * H2)
  FEDB_TAKE_F(dh1, (size_t)B * H1)
  FEDB_TAKE_F(loss_rows, (size_t)B)
  FEDB_TAKE_F(corr_rows, (size_t)B)
  FEDB_TAKE_F(conv_part, (size_t)B * CONV_PART)
#undef FEDB_TAKE_F
  return off;
}

struct Params {
  const float *w1, *b1, *w2, *b2, *fw1, *fb1, *fw2, *fb2, *fw3, *fb3;
};

// shared-memory plan of the sample kernel (floats)
struct Sm {
  static constexpr int img = 0;                       // 3 padded planes
  static constexpr int w1 = img + 3 * IPLANE + 2;     // 18 filters x 28 (25 used: rows padded for 128-bit loads)
  static constexpr int b1 = w1 + 18 * 28;             // 8
  static constexpr int w2 = b1 + 8;                   // 96 filters x 28
  static constexpr int b2 = w2 + 96 * 28;             // 16
  static constexpr int p1 = b2 + 16;                  // 6 padded planes
  static constexpr int flat = p1 + C1 * P1PLANE;      // 400
  static constexpr int h1 = flat + NFLAT;             // 120
  static constexpr int h2 = h1 + H1;                  // 84
  static constexpr int lg = h2 + H2;                  // 16
  static constexpr int dlg = lg + 16;                 // 16
  static constexpr int dh2 = dlg + 16;                // 84
  static constexpr int dh1 = dh2 + H2;                // 120
  static constexpr int dflat = dh1 + H1;              // 400
  static constexpr int fwd_end = dflat + NFLAT;
  // backward only
  static constexpr int g2 = fwd_end;                  // 400
  static constexpr int g1 = g2 + NFLAT;               // 1176
  static constexpr int dp1 = g1 + NP1;                // 8 x 1176 (also the 8 x 400 partials of d(flat), and 2 x 120 of dh1)
  static constexpr int half = dp1 + 8 * NP1;          // 5 x 450
  static constexpr int pos2 = half + 5 * NW1 + 2;     // 400 ints: arg-max offsets in a dense 14x14 plane (data gradient)
  static constexpr int pos2p = pos2 + NFLAT;          // 400 ints: the same offsets in the padded p1 planes (weight gradient)
  static constexpr int pos1 = pos2p + NFLAT;          // 1176 ints
  static constexpr int idx1 = pos1 + NP1;             // 1176 bytes = 294 words
  static constexpr int idx2 = idx1 + 296;             // 400 bytes = 100 words
  static constexpr int total = idx2 + 100;
};
constexpr size_t kSmemBytes = sizeof(float) * (size_t)Sm::total;

// conv (valid, 5x5) + ReLU + 2x2 max-pool for one output position and CO output channels,
// input planes in shared memory.  Returns pooled values and the argmax slot (0..3).
template <int CI, int CO_PER_THREAD, int IN_W, int PLANE = IN_W * IN_W>
__device__ __forceinline__ void conv5_relu_pool(const float* __restrict__ in, const float* __restrict__ w,
                                                const float* __restrict__ bias, int co0, int py, int px,
                                                float (&out)[CO_PER_THREAD], uint8_t (&arg)[CO_PER_THREAD]) {
  float acc[CO_PER_THREAD][4];
#pragma unroll
  for (int c = 0; c < CO_PER_THREAD; ++c) {
    const float b = bias[co0 + c];
    acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = b;
  }
  const int y0 = 2 * py, x0 = 2 * px;
#pragma unroll 1
  for (int ci = 0; ci < CI; ++ci) {
    float patch[6][6];
    const float* src = in + ci * PLANE + y0 * IN_W + x0;           // even offset: rows are 8-byte aligned
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int s = 0; s < 6; s += 2) {
        const float2 v = *reinterpret_cast<const float2*>(src + r * IN_W + s);
        patch[r][s] = v.x;
        patch[r][s + 1] = v.y;
      }
#pragma unroll
    for (int c = 0; c < CO_PER_THREAD; ++c) {
      float wreg[28];                                               // one padded 5x5 filter: 7 x LDS.128
      const float4* wk4 = reinterpret_cast<const float4*>(w + ((co0 + c) * CI + ci) * 28);
#pragma unroll
      for (int t = 0; t < 7; ++t) {
        const float4 v = wk4[t];
        wreg[4 * t] = v.x; wreg[4 * t + 1] = v.y; wreg[4 * t + 2] = v.z; wreg[4 * t + 3] = v.w;
      }
#pragma unroll
      for (int r = 0; r < 5; ++r)
#pragma unroll
        for (int s = 0; s < 5; ++s) {
          const float wv = wreg[r * 5 + s];
          acc[c][0] = fmaf(wv, patch[r][s], acc[c][0]);
          acc[c][1] = fmaf(wv, patch[r][s + 1], acc[c][1]);
          acc[c][2] = fmaf(wv, patch[r + 1][s], acc[c][2]);
          acc[c][3] = fmaf(wv, patch[r + 1][s + 1], acc[c][3]);
        }
    }
  }
#pragma unroll
  for (int c = 0; c < CO_PER_THREAD; ++c) {
    // first maximum wins, scanning (0,0),(0,1),(1,0),(1,1) like torch's max_pool2d
    float m = acc[c][0];
    uint8_t a = 0;
#pragma unroll
    for (int k = 1; k < 4; ++k)
      if (acc[c][k] > m) { m = acc[c][k]; a = (uint8_t)k; }
    out[c] = fmaxf(m, 0.f);
    arg[c] = a;
  }
}

__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float dot4(float4 a, float4 b, float acc) {
  acc = fmaf(a.x, b.x, acc);
  acc = fmaf(a.y, b.y, acc);
  acc = fmaf(a.z, b.z, acc);
  return fmaf(a.w, b.w, acc);
}

// y[j] = act(b[j] + sum_k W[j][k] x[k]) for one sample: warp w owns rows w, w+kWarps, ...; lanes own float4 slices of k.
// ROWS rows are in flight at once (all their loads are issued before the first FMA).
template <int K, int J, int ROWS, bool kRelu>
__device__ __forceinline__ void matvec(const float* __restrict__ W, const float* __restrict__ bias,
                                       const float* __restrict__ x_s, float* __restrict__ y_s, int warp, int lane) {
  constexpr int K4 = K / 4;                       // K is a multiple of 4 for 400 / 120 / 84
  constexpr int T = (K4 + 31) / 32;               // float4 slices per lane
  float4 xv[T];
#pragma unroll
  for (int t = 0; t < T; ++t) {
    const int k4 = lane + 32 * t;
    xv[t] = k4 < K4 ? reinterpret_cast<const float4*>(x_s)[k4] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (int j0 = warp; j0 < J; j0 += kWarps * ROWS) {
    float4 wv[ROWS][T];
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int j = j0 + kWarps * r;
#pragma unroll
      for (int t = 0; t < T; ++t) {
        const int k4 = lane + 32 * t;
        wv[r][t] = (j < J && k4 < K4) ? ldg4(W + (size_t)j * K + 4 * k4) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int r = 0; r < ROWS; ++r) {
      const int j = j0 + kWarps * r;
      float acc = 0.f;
#pragma unroll
      for (int t = 0; t < T; ++t) acc = dot4(wv[r][t], xv[t], acc);
      acc = warp_sum(acc);
      if (lane == 0 && j < J) {
        const float v = acc + __ldg(bias + j);
        y_s[j] = kRelu ? fmaxf(v, 0.f) : v;
      }
    }
  }
}

template <bool kU8, bool kTrain>
__global__ void __launch_bounds__(kThreads, 1) lenet_sample_kernel(const void* __restrict__ images, Params P, Workspace W,
                                                               const int64_t* __restrict__ labels, int B, int augment,
                                                               unsigned long long seed,
                                                               const unsigned long long* __restrict__ seed_dev,
                                                               float* __restrict__ logits_out,
                                                               long long* __restrict__ phase_clk = nullptr) {
  // phase_clk (debug, normally null): CTA 0 stamps clock64() at every phase boundary (tools/prof_lenet.py --phases)
#define FEDB_PHASE_MARK(i) do { if (phase_clk != nullptr && blockIdx.x == 0 && threadIdx.x == 0) phase_clk[i] = clock64(); } while (0)
  extern __shared__ __align__(16) float sm[];
  float* s_img = sm + Sm::img;
  float* s_w1 = sm + Sm::w1;
  float* s_b1 = sm + Sm::b1;
  float* s_w2 = sm + Sm::w2;
  float* s_b2 = sm + Sm::b2;
  float* s_p1 = sm + Sm::p1;
  float* s_flat = sm + Sm::flat;
  float* s_h1 = sm + Sm::h1;
  float* s_h2 = sm + Sm::h2;
  float* s_lg = sm + Sm::lg;
  float* s_dlg = sm + Sm::dlg;
  uint8_t* s_idx1 = reinterpret_cast<uint8_t*>(sm + Sm::idx1);
  uint8_t* s_idx2 = reinterpret_cast<uint8_t*>(sm + Sm::idx2);
  const int n = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (n >= B) return;

  FEDB_PHASE_MARK(0);
  // ------------------------------------------------------------------ phase 0: stage image + conv weights
  {
    constexpr int NW2V = (NW2 / 4 + kThreads - 1) / kThreads;     // float4 loads of conv2.weight per thread
    constexpr int NPX = 3072 / kThreads;                           // pixels per thread
    float4 w2v[NW2V];
#pragma unroll
    for (int j = 0; j < NW2V; ++j) {
      const int i4 = tid + kThreads * j;
      w2v[j] = i4 < NW2 / 4 ? ldg4(P.w2 + 4 * i4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const float w1v = tid < NW1 ? __ldg(P.w1 + tid) : 0.f;
    if constexpr (kU8) {
      const uint8_t* img = (const uint8_t*)images + (size_t)n * 3072;
      const fedb::Crop cr = fedb::draw_crop(seed ^ (seed_dev ? *seed_dev : 0ull), (unsigned long long)n, augment);
      float px[NPX];
#pragma unroll
      for (int j = 0; j < NPX; ++j) {                      // destination-centric: all byte loads in flight
        const int e = tid + kThreads * j;
        const int x = e & 31, y = (e >> 5) & 31, c = e >> 10;
        const int sx = (cr.flip ? 31 - x : x) + cr.dx, sy = y + cr.dy;
        const bool in = sx >= 0 && sx < 32 && sy >= 0 && sy < 32;
        px[j] = in ? (float)__ldg(img + (sy * 32 + sx) * 3 + c) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < NPX; ++j) {
        const int c = (tid + kThreads * j) >> 10;
        const float mean = c == 0 ? 0.4914f : (c == 1 ? 0.4822f : 0.4465f);
        const float stdv = c == 0 ? 0.2023f : (c == 1 ? 0.1994f : 0.2010f);
        const int e = tid + kThreads * j;
        s_img[c * IPLANE + ((e >> 5) & 31) * IW + (e & 31)] = (px[j] / 255.f - mean) / stdv;
      }
    } else {
      const float* img = (const float*)images + (size_t)n * 3072;
      float4 v[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) v[j] = tid + kThreads * j < 768 ? ldg4(img + 4 * (tid + kThreads * j)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int e = 4 * (tid + kThreads * j);                  // 4 consecutive pixels of one row
        if (e < 3072) {
          float* d = s_img + (e >> 10) * IPLANE + ((e >> 5) & 31) * IW + (e & 31);
          reinterpret_cast<float2*>(d)[0] = make_float2(v[j].x, v[j].y);
          reinterpret_cast<float2*>(d)[1] = make_float2(v[j].z, v[j].w);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < NW2V; ++j) {
      const int i = 4 * (tid + kThreads * j);
      if (i < NW2) {
        const float v[4] = {w2v[j].x, w2v[j].y, w2v[j].z, w2v[j].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) s_w2[((i + q) / 25) * 28 + (i + q) % 25] = v[q];
      }
    }
    if (tid < NW1) s_w1[(tid / 25) * 28 + tid % 25] = w1v;
    if (tid < C1) s_b1[tid] = __ldg(P.b1 + tid);
    if (tid < C2) s_b2[tid] = __ldg(P.b2 + tid);
  }
  __syncthreads();

  FEDB_PHASE_MARK(1);
  // ------------------------------------------------------------------ phase 1: conv1 + ReLU + pool
  if (tid < 2 * P1 * P1) {                     // (position, half of the output channels) per thread
    const int pos = tid % 196, hf = tid / 196;
    const int py = pos / P1, px = pos % P1;
    float o[3];
    uint8_t a[3];
    conv5_relu_pool<3, 3, IW, IPLANE>(s_img, s_w1, s_b1, 3 * hf, py, px, o, a);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      s_p1[(3 * hf + c) * P1PLANE + py * P1S + px] = o[c];
      s_idx1[(3 * hf + c) * 196 + pos] = a[c];
    }
  }
  __syncthreads();

  FEDB_PHASE_MARK(2);
  // ------------------------------------------------------------------ phase 2: conv2 + ReLU + pool
  if (tid < NFLAT / 2) {                       // (position, pair of output channels) per thread: the phase is bound
    const int pos = tid % 25, cg = tid / 25;   // by shared-memory loads, and a patch now feeds two channels
    const int py = pos / P2, px = pos % P2;
    float o[2];
    uint8_t a[2];
    conv5_relu_pool<C1, 2, P1S, P1PLANE>(s_p1, s_w2, s_b2, 2 * cg, py, px, o, a);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      s_flat[(2 * cg + c) * 25 + pos] = o[c];  // co * 25 + pos: NCHW flatten order == fc1's input order
      s_idx2[(2 * cg + c) * 25 + pos] = a[c];
    }
  }
  __syncthreads();

  FEDB_PHASE_MARK(3);
  // ------------------------------------------------------------------ phase 3: classifier forward
  matvec<NFLAT, H1, 5, true>(P.fw1, P.fb1, s_flat, s_h1, warp, lane);
  __syncthreads();
  matvec<H1, H2, 6, true>(P.fw2, P.fb2, s_h1, s_h2, warp, lane);
  __syncthreads();
  matvec<H2, NCLS, 1, false>(P.fw3, P.fb3, s_h2, s_lg, warp, lane);
  __syncthreads();

  FEDB_PHASE_MARK(4);
  // ------------------------------------------------------------------ phase 4: softmax cross-entropy
  if (warp == 0) {
    const float z = lane < NCLS ? s_lg[lane] : -INFINITY;
    float mx = z;
    int arg = lane < NCLS ? lane : 1 << 20;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    const float se = warp_sum(lane < NCLS ? expf(z - mx) : 0.f);
    const float lse = logf(se) + mx;
    const int t = labels != nullptr ? (int)labels[n] : 0;
    if (lane < NCLS) {
      const float smx = expf(z - lse);
      s_dlg[lane] = (smx - (lane == t ? 1.f : 0.f)) / (float)B;
      if (logits_out) logits_out[(size_t)n * NCLS + lane] = z;
    }
    if (lane == 0) {
      W.loss_rows[n] = lse - s_lg[t];
      W.corr_rows[n] = (arg == t) ? 1.f : 0.f;
    }
  }
  if constexpr (!kTrain) return;
  __syncthreads();

  float* s_dh2 = sm + Sm::dh2;
  float* s_dh1 = sm + Sm::dh1;
  float* s_dflat = sm + Sm::dflat;
  float* s_g2 = sm + Sm::g2;
  float* s_g1 = sm + Sm::g1;
  float* s_dp1 = sm + Sm::dp1;
  float* s_half = sm + Sm::half;
  int* s_pos2 = reinterpret_cast<int*>(sm + Sm::pos2);
  int* s_pos2p = reinterpret_cast<int*>(sm + Sm::pos2p);
  int* s_pos1 = reinterpret_cast<int*>(sm + Sm::pos1);

  FEDB_PHASE_MARK(5);
  // ------------------------------------------------------------------ phase 5: classifier backward (data path)
  // dh2[j] = relu'(h2[j]) * sum_c dlg[c] W3[c][j]
  if (tid < H2) {
    float wv[NCLS];
#pragma unroll
    for (int c = 0; c < NCLS; ++c) wv[c] = __ldg(P.fw3 + c * H2 + tid);
    float acc = 0.f;
#pragma unroll
    for (int c = 0; c < NCLS; ++c) acc = fmaf(s_dlg[c], wv[c], acc);
    s_dh2[tid] = s_h2[tid] > 0.f ? acc : 0.f;
  }
  __syncthreads();
  // dh1[i] = relu'(h1[i]) * sum_j dh2[j] W2[j][i]; four threads per output (j quarters), all 21 loads in flight
  if (tid < 4 * H1) {
    const int i = tid % H1, qt = tid / H1;
    float wv[21];
#pragma unroll
    for (int j = 0; j < 21; ++j) wv[j] = __ldg(P.fw2 + (qt * 21 + j) * H1 + i);
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 21; ++j) acc = fmaf(s_dh2[qt * 21 + j], wv[j], acc);
    s_dp1[tid] = acc;
  }
  __syncthreads();
  if (tid < H1)
    s_dh1[tid] = s_h1[tid] > 0.f ? ((s_dp1[tid] + s_dp1[H1 + tid]) + (s_dp1[2 * H1 + tid] + s_dp1[3 * H1 + tid])) : 0.f;
  __syncthreads();
  FEDB_PHASE_MARK(6);
  // dflat[k] = sum_j dh1[j] W1[j][k]: warps 0..14 own rows [8w, 8w+8), lanes own float4 slices of k
  if (warp < 15) {
    float4 acc[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) acc[t] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
    for (int jj = 0; jj < 8; jj += 4) {
      float4 wv[4][4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int j = warp * 8 + jj + r;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          const int k4 = lane + 32 * t;
          wv[r][t] = k4 < NFLAT / 4 ? ldg4(P.fw1 + (size_t)j * NFLAT + 4 * k4) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float d = s_dh1[warp * 8 + jj + r];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
          acc[t].x = fmaf(d, wv[r][t].x, acc[t].x);
          acc[t].y = fmaf(d, wv[r][t].y, acc[t].y);
          acc[t].z = fmaf(d, wv[r][t].z, acc[t].z);
          acc[t].w = fmaf(d, wv[r][t].w, acc[t].w);
        }
      }
    }
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int k4 = lane + 32 * t;
      if (k4 < NFLAT / 4) reinterpret_cast<float4*>(s_dp1 + warp * NFLAT)[k4] = acc[t];
    }
  }
  __syncthreads();
  for (int k = tid; k < NFLAT; k += kThreads) {
    float d = 0.f;
#pragma unroll
    for (int w = 0; w < 15; ++w) d += s_dp1[w * NFLAT + k];
    s_dflat[k] = d;
  }
  __syncthreads();

  FEDB_PHASE_MARK(7);
  // publish the rows kernel U needs (coalesced)
  for (int e = tid; e < NFLAT; e += kThreads) W.flat[(size_t)n * NFLAT + e] = s_flat[e];
  if (tid < H1) { W.h1[(size_t)n * H1 + tid] = s_h1[tid]; W.dh1[(size_t)n * H1 + tid] = s_dh1[tid]; }
  if (tid < H2) { W.h2[(size_t)n * H2 + tid] = s_h2[tid]; W.dh2[(size_t)n * H2 + tid] = s_dh2[tid]; }
  if (tid < NCLS) W.dlogits[(size_t)n * NCLS + tid] = s_dlg[tid];

  FEDB_PHASE_MARK(8);
  // ------------------------------------------------------------------ phase 6: feature extractor backward
  for (int e = tid; e < 8 * NP1; e += kThreads) s_dp1[e] = 0.f;
  for (int e = tid; e < NFLAT; e += kThreads) {
    const int k = s_idx2[e];
    const int pos = e % 25, py = pos / 5, px = pos % 5;
    s_g2[e] = s_flat[e] > 0.f ? s_dflat[e] : 0.f;
    s_pos2[e] = (2 * py + (k >> 1)) * 14 + (2 * px + (k & 1));       // offset of the arg-max inside a 14x14 p1 plane
    s_pos2p[e] = (2 * py + (k >> 1)) * P1S + (2 * px + (k & 1));     // ... and inside the padded plane
  }
  __syncthreads();

  FEDB_PHASE_MARK(9);
  float* part = W.conv_part + (size_t)n * CONV_PART;
  // conv2 weight gradient: dW2[co][ci][r][s] = sum_pos g2[co][pos] * p1[ci][y+r][x+s]
  // one thread per (co, ci, filter row): the five taps of the row share the loads of g2 and of the arg-max offset
  if (tid < NW2 / 5) {
    const int r = tid % 5, ci = (tid / 5) % C1, co = tid / 30;
    const float* src = s_p1 + ci * P1PLANE + r * P1S;
    float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 5
    for (int pos = 0; pos < 25; ++pos) {
      const float g = s_g2[co * 25 + pos];
      const float* q = src + s_pos2p[co * 25 + pos];
#pragma unroll
      for (int t = 0; t < 5; ++t) acc[t] = fmaf(g, q[t], acc[t]);
    }
#pragma unroll
    for (int t = 0; t < 5; ++t) part[OFF_W2 + tid * 5 + t] = acc[t];
  }
  if (tid < C2) {
    float acc = 0.f;
    for (int pos = 0; pos < 25; ++pos) acc += s_g2[tid * 25 + pos];
    part[OFF_B2 + tid] = acc;
  }
  FEDB_PHASE_MARK(10);
  // conv2 data gradient, sparse scatter: each non-zero of d(conv2 out) touches a 5x5 patch of every input channel.
  // Warps w and w+8 share accumulator copy w%8 but own disjoint input channels (0-2 / 3-5); each copy's owners
  // walk non-zeros w%8, w%8+8, ...; lanes are the 25 taps.
  {
    float* acc = s_dp1 + (warp & 7) * NP1;
    const int ci0 = (warp >> 3) * 3;
    const int r = lane / 5, s = lane % 5;
    for (int e = warp & 7; e < NFLAT; e += 8) {
      const float g = s_g2[e];
      if (g != 0.f && lane < 25) {
        const int co = e / 25;
        const int base = s_pos2[e] + r * 14 + s;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const int ci = ci0 + c;
          acc[ci * 196 + base] = fmaf(g, s_w2[(co * C1 + ci) * 28 + lane], acc[ci * 196 + base]);
        }
      }
      __syncwarp();
    }
  }
  __syncthreads();
  FEDB_PHASE_MARK(11);
  for (int e = tid; e < NP1; e += kThreads) {
    float d = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) d += s_dp1[w * NP1 + e];
    const int k = s_idx1[e];
    const int pos = e % 196, py = pos / 14, px = pos % 14;
    s_g1[e] = s_p1[(e / 196) * P1PLANE + py * P1S + px] > 0.f ? d : 0.f;
    s_pos1[e] = (2 * py + (k >> 1)) * IW + (2 * px + (k & 1));       // offset of the arg-max inside a padded image plane
  }
  __syncthreads();
  FEDB_PHASE_MARK(12);
  // conv1 weight gradient: one thread per (fifth of the positions, co, ci, filter row); the five taps of the row
  // share g1 and the arg-max offset, which come in as 128-bit loads for four positions at a time.  Lanes run over
  // (r, ci) first: with the padded planes their addresses fall into distinct banks.
  if (tid < 5 * 90) {
    const int o = tid % 90, q5 = tid / 90;
    const int r = o % 5, ci = (o / 5) % 3, co = o / 15;
    const float* src = s_img + ci * IPLANE + r * IW;
    const int lo = q5 * 40, n4 = q5 < 4 ? 10 : 9;                   // 4 x 40 + 36 positions
    const float4* g4 = reinterpret_cast<const float4*>(s_g1 + co * 196 + lo);
    const int4* p4 = reinterpret_cast<const int4*>(s_pos1 + co * 196 + lo);
    float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 2
    for (int i = 0; i < n4; ++i) {
      const float4 g = g4[i];
      const int4 pp = p4[i];
      const float gv[4] = {g.x, g.y, g.z, g.w};
      const int pv[4] = {pp.x, pp.y, pp.z, pp.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float* q = src + pv[k];
#pragma unroll
        for (int t = 0; t < 5; ++t) acc[t] = fmaf(gv[k], q[t], acc[t]);
      }
    }
#pragma unroll
    for (int t = 0; t < 5; ++t) s_half[q5 * NW1 + o * 5 + t] = acc[t];   // o * 5 + t == co*75 + ci*25 + r*5 + s
  }
  __syncthreads();
  FEDB_PHASE_MARK(13);
  for (int o = tid; o < NW1; o += kThreads)
    part[o] = ((s_half[o] + s_half[NW1 + o]) + (s_half[2 * NW1 + o] + s_half[3 * NW1 + o])) + s_half[4 * NW1 + o];
  if (tid < C1) {
    float acc = 0.f;
    for (int pos = 0; pos < 196; ++pos) acc += s_g1[tid * 196 + pos];
    part[OFF_B1 + tid] = acc;
  }
  FEDB_PHASE_MARK(14);
#undef FEDB_PHASE_MARK
}

// ------------------------------------------------------------------------------------ kernel U
struct UpdateArgs {
  float* param;
  float* momentum;
  float* dparam;           // grad-only mode (tests): write gradient here, do not update
  long long off[10];
  float lr, mu, wd;
  int B;
  float* stats;            // [loss_sum_of_means, correct, samples, steps]
};

// Block roles (grid = kUpdateBlocks):
//   [0, 60)    fc1.weight tiles: 8 rows (j) x 128 cols (k) of dW1 = dh1^T flat      (15 x 4 tiles)
//   [60, 71)   fc2.weight tiles: 8 x 128 of dW2 = dh2^T h1                           (11 x 1)
//   [71, 73)   fc3.weight tiles: 8 x 128 of dW3 = dlogits^T h2                       (2 x 1)
//   [73, 170)  32 of the remaining parameters per block (fc biases = column sums of dout, conv partial sums):
//              warp w sums samples [16w, 16w+16) of a 128-sample chunk, lanes are parameters (coalesced rows)
//   170        loss / accuracy counters (K13)
// The whole kernel is a handful of dependent memory round trips, so every block issues ALL loads of a 128-sample
// chunk (and of the parameter / momentum words it will update) before it touches the first one.  Summation orders
// are fixed (sample order inside a group, groups in order), so a step is reproducible run to run.
constexpr int kRestParams = H1 + H2 + NCLS + NW1 + C1 + NW2 + C2;   // 3086
constexpr int kRestBlocks = (kRestParams + 31) / 32;                 // 97
constexpr int kUpdateBlocks = 73 + kRestBlocks + 1;
constexpr int kUChunk = 128;                                         // samples staged per pass
constexpr size_t kUpdateSmemBytes = sizeof(float) * (size_t)(kUChunk * 128 + kUChunk * 8);

template <bool kApply>
__device__ __forceinline__ void apply_update(const UpdateArgs& U, long long dst, float g, float p, float m) {
  if constexpr (kApply) {
    const float g1 = __fadd_rn(g, __fmul_rn(U.wd, p));
    const float v = __fadd_rn(__fmul_rn(U.mu, m), g1);
    U.momentum[dst] = v;
    U.param[dst] = __fadd_rn(p, __fmul_rn(-U.lr, v));
  } else {
    U.dparam[dst] = g;
  }
}

template <bool kApply>
__global__ void __launch_bounds__(256) update_kernel(UpdateArgs U, Workspace W) {
  extern __shared__ __align__(16) float usm[];
  const int B = U.B, b = blockIdx.x, tid = threadIdx.x;

  if (b < 73) {
    float* s_in = usm;                                          // [128 samples][128 cols]
    float* s_dout = usm + kUChunk * 128;                        // [128 samples][8 rows]
    int J, K, jt, kt;
    const float *dout, *in;
    long long off;
    if (b < 60)      { J = H1; K = NFLAT; jt = b / 4; kt = b % 4; dout = W.dh1; in = W.flat; off = U.off[4]; }
    else if (b < 71) { J = H2; K = H1; jt = b - 60; kt = 0; dout = W.dh2; in = W.h1; off = U.off[6]; }
    else             { J = NCLS; K = H2; jt = b - 71; kt = 0; dout = W.dlogits; in = W.h2; off = U.off[8]; }
    // compute roles: warp = (row quad rq, sample quarter sq); a thread owns rows 4rq..4rq+3 x columns k..k+3 and sums the
    // samples of its quarter.  Four rows per loaded x quarter the shared-memory wavefronts of the FMA loop (the loop is
    // bound by them); the four sample quarters are added in a fixed order at the end.
    const int wp = tid >> 5, lane = tid & 31;
    const int rq = wp & 1, sq = wp >> 1;
    const int j0 = jt * 8 + 4 * rq, k = kt * 128 + lane * 4;
    const bool col_ok = k < K;
    const long long dst0 = off + (long long)j0 * K + k;         // multiple of 4 floats: segments are 32-float aligned
    float4 pw[4], mw[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      pw[r] = mw[r] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (kApply && sq == 0 && col_ok && j0 + r < J) {
        pw[r] = *reinterpret_cast<const float4*>(U.param + dst0 + (long long)r * K);
        mw[r] = *reinterpret_cast<const float4*>(U.momentum + dst0 + (long long)r * K);
      }
    }
    // loader roles: float4 column lk4 of sample rows ls, ls+8, ...; dout element (ds, dj)
    const int lk4 = tid & 31, ls = tid >> 5;
    const int dj = tid & 7, ds = tid >> 3;
    const bool lk_ok = kt * 128 + lk4 * 4 < K;                  // K is a multiple of 4
    const bool dj_ok = jt * 8 + dj < J;
    float4 acc[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) acc[r] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int s0 = 0; s0 < B; s0 += kUChunk) {
      float4 vin[16];
      float vd[4];
#pragma unroll
      for (int r = 0; r < 16; ++r) {
        const int sidx = s0 + ls + 8 * r;
        vin[r] = (lk_ok && sidx < B) ? __ldg(reinterpret_cast<const float4*>(in + (size_t)sidx * K + kt * 128) + lk4)
                                     : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const int sidx = s0 + ds + 32 * r;
        vd[r] = (dj_ok && sidx < B) ? __ldg(dout + (size_t)sidx * J + jt * 8 + dj) : 0.f;
      }
      if (s0 > 0) __syncthreads();                              // previous chunk fully consumed
#pragma unroll
      for (int r = 0; r < 16; ++r) reinterpret_cast<float4*>(s_in)[(ls + 8 * r) * 32 + lk4] = vin[r];
#pragma unroll
      for (int r = 0; r < 4; ++r) s_dout[(ds + 32 * r) * 8 + dj] = vd[r];
      __syncthreads();
      const int ns = min(kUChunk, B - s0);
      const int sb = sq * (kUChunk / 4), se = min(sb + kUChunk / 4, ns);
#pragma unroll 8
      for (int sidx = sb; sidx < se; ++sidx) {
        const float4 d = *reinterpret_cast<const float4*>(s_dout + sidx * 8 + 4 * rq);
        const float4 x = reinterpret_cast<const float4*>(s_in)[sidx * 32 + lane];
        const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          acc[r].x = fmaf(dv[r], x.x, acc[r].x);
          acc[r].y = fmaf(dv[r], x.y, acc[r].y);
          acc[r].z = fmaf(dv[r], x.z, acc[r].z);
          acc[r].w = fmaf(dv[r], x.w, acc[r].w);
        }
      }
    }
    __syncthreads();                                            // every warp is done with the staged tile
    float4* s_part4 = reinterpret_cast<float4*>(s_in);           // [3 quarters][8 rows][32 lanes]
    if (sq > 0) {
#pragma unroll
      for (int r = 0; r < 4; ++r) s_part4[((sq - 1) * 8 + 4 * rq + r) * 32 + lane] = acc[r];
    }
    __syncthreads();
    if (sq == 0 && col_ok) {
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        if (j0 + r >= J) break;
        float4 g = acc[r];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const float4 h = s_part4[(q * 8 + 4 * rq + r) * 32 + lane];
          g.x += h.x; g.y += h.y; g.z += h.z; g.w += h.w;
        }
        const long long dst = dst0 + (long long)r * K;
        apply_update<kApply>(U, dst + 0, g.x, pw[r].x, mw[r].x);
        apply_update<kApply>(U, dst + 1, g.y, pw[r].y, mw[r].y);
        apply_update<kApply>(U, dst + 2, g.z, pw[r].z, mw[r].z);
        apply_update<kApply>(U, dst + 3, g.w, pw[r].w, mw[r].w);
      }
    }
    return;
  }

  if (b == kUpdateBlocks - 1) {                      // K13: counters stay on the device
    if (tid < 32 && U.stats != nullptr) {
      float l = 0.f, c = 0.f;
      for (int sidx = tid; sidx < B; sidx += 32) { l += W.loss_rows[sidx]; c += W.corr_rows[sidx]; }
      l = warp_sum(l);
      c = warp_sum(c);
      if (tid == 0) {
        U.stats[0] += l / (float)B;
        U.stats[1] += c;
        U.stats[2] += (float)B;
        U.stats[3] += 1.f;
      }
    }
    return;
  }

  // ---- remaining parameters: lane = parameter, warp = group of 16 samples
  float* s_part = usm;                                           // [8 warps][32]
  const int lane = tid & 31, w = tid >> 5;
  const int i = (b - 73) * 32 + lane;
  const bool ok = i < kRestParams;
  const float* src = nullptr;
  size_t stride = 0;
  long long dst = 0;
  if (ok) {
    if (i < H1)                  { src = W.dh1 + i; stride = H1; dst = U.off[5] + i; }
    else if (i < H1 + H2)        { src = W.dh2 + (i - H1); stride = H2; dst = U.off[7] + (i - H1); }
    else if (i < H1 + H2 + NCLS) { src = W.dlogits + (i - H1 - H2); stride = NCLS; dst = U.off[9] + (i - H1 - H2); }
    else {
      const int c = i - (H1 + H2 + NCLS);                        // position inside the conv partial row
      src = W.conv_part + c;
      stride = CONV_PART;
      if (c < OFF_B1)      dst = U.off[0] + c;
      else if (c < OFF_W2) dst = U.off[1] + (c - OFF_B1);
      else if (c < OFF_B2) dst = U.off[2] + (c - OFF_W2);
      else                 dst = U.off[3] + (c - OFF_B2);
    }
  }
  float p = 0.f, m = 0.f;
  if (kApply && ok && w == 0) { p = U.param[dst]; m = U.momentum[dst]; }
  float g = 0.f;
  for (int s0 = 0; s0 < B; s0 += kUChunk) {
    float v[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int sidx = s0 + 16 * w + r;
      v[r] = (ok && sidx < B) ? __ldg(src + (size_t)sidx * stride) : 0.f;
    }
    float part = 0.f;
#pragma unroll
    for (int r = 0; r < 16; ++r) part += v[r];
    if (s0 > 0) __syncthreads();
    s_part[w * 32 + lane] = part;
    __syncthreads();
    if (w == 0) {
#pragma unroll
      for (int q = 0; q < 8; ++q) g += s_part[q * 32 + lane];
    }
  }
  if (w == 0 && ok) apply_update<kApply>(U, dst, g, p, m);
}

// ---------------------------------------------------------------------------------------------- evaluation kernel
// Forward + loss/correct of kEvalS samples per CTA (the per-round evaluation of the global model on the test set, N8).
// Compared with one sample per CTA: the classifier weights cross L2 once per kEvalS samples, every phase has kEvalS
// times more independent work between two barriers (the phases of a single sample are latency-bound), and each
// conv1 work item produces all six channels from one 6x6x3 patch.
constexpr int kEvalS = 5;
struct SmE {
  static constexpr int img = 0;                                // S x 3072
  static constexpr int w1 = img + kEvalS * 3072;
  static constexpr int b1 = w1 + 18 * 28;
  static constexpr int w2 = b1 + 8;
  static constexpr int b2 = w2 + 96 * 28;
  static constexpr int p1 = b2 + 16;                           // S x 1176
  static constexpr int flat = p1 + kEvalS * NP1;               // S x 400
  static constexpr int h1 = flat + kEvalS * NFLAT;             // S x 120
  static constexpr int h2 = h1 + kEvalS * H1;                  // S x 84
  static constexpr int lg = h2 + kEvalS * H2;                  // S x 16
  static constexpr int total = lg + kEvalS * 16;
};
constexpr size_t kEvalSmemBytes = sizeof(float) * (size_t)SmE::total;

// y[s][j] = act(b[j] + sum_k W[j][k] x[s][k]) for kEvalS samples at once.  LPR lanes share a row (32/LPR rows per warp
// and iteration): a weight slice is loaded once and used for every sample, and a row sum costs log2(LPR) shuffles.
template <int K, int J, int LPR, bool kRelu>
__device__ __forceinline__ void matvec_multi(const float* __restrict__ W, const float* __restrict__ bias,
                                             const float* __restrict__ x_s, float* __restrict__ y_s, int warp, int lane) {
  constexpr int K4 = K / 4;
  constexpr int RPW = 32 / LPR;
  constexpr int T = (K4 + LPR - 1) / LPR;         // float4 slices per lane
  constexpr int TC = T < 7 ? T : 7;               // slices in flight (register budget)
  const int sub = lane / LPR, l = lane % LPR;
  for (int j0 = warp * RPW; j0 < J; j0 += kWarps * RPW) {
    const int j = j0 + sub;
    float acc[kEvalS];
#pragma unroll
    for (int smp = 0; smp < kEvalS; ++smp) acc[smp] = 0.f;
#pragma unroll
    for (int t0 = 0; t0 < T; t0 += TC) {
      float4 wv[TC];
#pragma unroll
      for (int t = 0; t < TC; ++t) {
        const int k4 = l + LPR * (t0 + t);
        wv[t] = (t0 + t < T && j < J && k4 < K4) ? ldg4(W + (size_t)j * K + 4 * k4) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int smp = 0; smp < kEvalS; ++smp)
#pragma unroll
        for (int t = 0; t < TC; ++t) {
          const int k4 = l + LPR * (t0 + t);
          if (t0 + t < T && k4 < K4) acc[smp] = dot4(wv[t], reinterpret_cast<const float4*>(x_s + smp * K)[k4], acc[smp]);
        }
    }
#pragma unroll
    for (int smp = 0; smp < kEvalS; ++smp) {
      float v = acc[smp];
#pragma unroll
      for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (l == 0 && j < J) {
        v += __ldg(bias + j);
        y_s[smp * J + j] = kRelu ? fmaxf(v, 0.f) : v;
      }
    }
  }
}

// kQueue: the CTAs are persistent and pull chunks of kEvalS samples from an atomic counter until the test set is done.  Two
// launches may share one counter: a small "background" grid that runs next to the training kernels of the following round
// (the step kernel leaves 148 - 128 SMs idle) and a full-size "finisher" grid that takes whatever is left when the round
// needs the result - every chunk is scored exactly once, by whoever gets it first.
// (the chunk body is a separate, non-inlined function: inlined into the queue loop it spilled 400 bytes at 64 registers)
__device__ __noinline__ void lenet_eval_chunk(const uint8_t* __restrict__ images, const Params& P,
                                              const int64_t* __restrict__ labels, int B, float* __restrict__ loss_rows,
                                              float* __restrict__ corr_rows, int chunk) {
  extern __shared__ __align__(16) float sm[];
  float* s_img = sm + SmE::img;
  float* s_w1 = sm + SmE::w1;
  float* s_b1 = sm + SmE::b1;
  float* s_w2 = sm + SmE::w2;
  float* s_b2 = sm + SmE::b2;
  float* s_p1 = sm + SmE::p1;
  float* s_flat = sm + SmE::flat;
  float* s_h1 = sm + SmE::h1;
  float* s_h2 = sm + SmE::h2;
  float* s_lg = sm + SmE::lg;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = chunk * kEvalS;
  const int ns = min(kEvalS, B - n0);

  // ---- phase 0: conv weights into padded rows; images as 128-bit loads (HWC bytes), scattered into CHW planes
  {
    constexpr int NW2V = (NW2 / 4 + kThreads - 1) / kThreads;
    constexpr int NV = (kEvalS * 192 + kThreads - 1) / kThreads;      // 192 x 16 bytes per image
    float4 w2v[NW2V];
    uint4 pv[NV];
#pragma unroll
    for (int j = 0; j < NW2V; ++j) {
      const int i4 = tid + kThreads * j;
      w2v[j] = i4 < NW2 / 4 ? ldg4(P.w2 + 4 * i4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const uint4* src = reinterpret_cast<const uint4*>(images + (size_t)n0 * 3072);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = tid + kThreads * j;
      pv[j] = i < ns * 192 ? __ldg(src + i) : make_uint4(0u, 0u, 0u, 0u);
    }
    const float w1v = tid < NW1 ? __ldg(P.w1 + tid) : 0.f;
    // samples that do not exist (last CTA) are zero images: their results are never written
    for (int i = ns * 3072 + tid; i < kEvalS * 3072; i += kThreads) s_img[i] = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = tid + kThreads * j;
      if (i < ns * 192) {
        const int smp = i / 192, b0 = (i % 192) * 16;
        const unsigned wds[4] = {pv[j].x, pv[j].y, pv[j].z, pv[j].w};
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const int b = b0 + k, pix = b / 3, c = b - 3 * pix;
          const float v = (float)((wds[k >> 2] >> (8 * (k & 3))) & 0xffu);
          const float mean = c == 0 ? 0.4914f : (c == 1 ? 0.4822f : 0.4465f);
          const float stdv = c == 0 ? 0.2023f : (c == 1 ? 0.1994f : 0.2010f);
          s_img[smp * 3072 + c * 1024 + pix] = (v / 255.f - mean) / stdv;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < NW2V; ++j) {
      const int i = 4 * (tid + kThreads * j);
      if (i < NW2) {
        const float v[4] = {w2v[j].x, w2v[j].y, w2v[j].z, w2v[j].w};
#pragma unroll
        for (int q = 0; q < 4; ++q) s_w2[((i + q) / 25) * 28 + (i + q) % 25] = v[q];
      }
    }
    if (tid < NW1) s_w1[(tid / 25) * 28 + tid % 25] = w1v;
    if (tid < C1) s_b1[tid] = __ldg(P.b1 + tid);
    if (tid < C2) s_b2[tid] = __ldg(P.b2 + tid);
  }
  __syncthreads();

  // ---- phase 1: conv1 + ReLU + pool, (sample, position) per item, all six channels from one patch
  for (int item = tid; item < kEvalS * 196; item += kThreads) {
    const int smp = item / 196, pos = item - smp * 196;
    float o[C1];
    uint8_t a[C1];
    conv5_relu_pool<3, C1, 32>(s_img + smp * 3072, s_w1, s_b1, 0, pos / P1, pos % P1, o, a);
#pragma unroll
    for (int c = 0; c < C1; ++c) s_p1[smp * NP1 + c * 196 + pos] = o[c];
  }
  __syncthreads();

  // ---- phase 2: conv2 + ReLU + pool, (sample, channel pair, position) per item
  for (int item = tid; item < kEvalS * 200; item += kThreads) {
    const int smp = item / 200, r = item - smp * 200;
    const int cg = r / 25, pos = r - cg * 25;
    float o[2];
    uint8_t a[2];
    conv5_relu_pool<C1, 2, P1>(s_p1 + smp * NP1, s_w2, s_b2, 2 * cg, pos / P2, pos % P2, o, a);
    s_flat[smp * NFLAT + (2 * cg) * 25 + pos] = o[0];
    s_flat[smp * NFLAT + (2 * cg + 1) * 25 + pos] = o[1];
  }
  __syncthreads();

  // ---- phase 3: classifier
  matvec_multi<NFLAT, H1, 8, true>(P.fw1, P.fb1, s_flat, s_h1, warp, lane);
  __syncthreads();
  matvec_multi<H1, H2, 8, true>(P.fw2, P.fb2, s_h1, s_h2, warp, lane);
  __syncthreads();
  for (int o = tid; o < kEvalS * NCLS; o += kThreads) {               // 84 x 10: one thread per logit
    const int smp = o / NCLS, c = o - smp * NCLS;
    float acc = __ldg(P.fb3 + c);
#pragma unroll 4
    for (int k = 0; k < H2; ++k) acc = fmaf(__ldg(P.fw3 + c * H2 + k), s_h2[smp * H2 + k], acc);
    s_lg[smp * 16 + c] = acc;
  }
  __syncthreads();

  // ---- phase 4: loss and correctness, one warp per sample
  if (warp < ns) {
    const int n = n0 + warp;
    const float z = lane < NCLS ? s_lg[warp * 16 + lane] : -INFINITY;
    float mx = z;
    int arg = lane < NCLS ? lane : 1 << 20;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    const float se = warp_sum(lane < NCLS ? expf(z - mx) : 0.f);
    const float lse = logf(se) + mx;
    if (lane == 0) {
      const int t = (int)labels[n];
      loss_rows[n] = lse - s_lg[warp * 16 + t];
      corr_rows[n] = (arg == t) ? 1.f : 0.f;
    }
  }
}

template <bool kQueue>
__global__ void __launch_bounds__(kThreads, 2) lenet_eval_kernel(const uint8_t* __restrict__ images, const __grid_constant__ Params P,
                                                               const int64_t* __restrict__ labels, int B,
                                                               float* __restrict__ loss_rows, float* __restrict__ corr_rows,
                                                               unsigned int* __restrict__ counter) {
  const int nchunks = (B + kEvalS - 1) / kEvalS;
  if constexpr (!kQueue) {
    if ((int)blockIdx.x < nchunks) lenet_eval_chunk(images, P, labels, B, loss_rows, corr_rows, blockIdx.x);
  } else {
    __shared__ unsigned int s_chunk;
    for (;;) {
      __syncthreads();                                 // everybody is done with the previous chunk's shared memory
      if (threadIdx.x == 0) s_chunk = atomicAdd(counter, 1u);
      __syncthreads();
      const int chunk = (int)s_chunk;
      if (chunk >= nchunks) return;
      lenet_eval_chunk(images, P, labels, B, loss_rows, corr_rows, chunk);
    }
  }
}

__global__ void __launch_bounds__(256) stats_kernel(Workspace W, int B, float* __restrict__ stats) {
  __shared__ float sl[256], sc[256];
  float l = 0.f, c = 0.f;
  for (int i = threadIdx.x; i < B; i += 256) { l += W.loss_rows[i]; c += W.corr_rows[i]; }
  sl[threadIdx.x] = l;
  sc[threadIdx.x] = c;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { sl[threadIdx.x] += sl[threadIdx.x + s]; sc[threadIdx.x] += sc[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    stats[0] += sl[0] / (float)B;
    stats[1] += sc[0];
    stats[2] += (float)B;
    stats[3] += 1.f;
  }
}

Params make_params(const fedb_lenet_t* net) {
  Params P;
  const float* a = net->param;
  P.w1 = a + net->off[0]; P.b1 = a + net->off[1]; P.w2 = a + net->off[2]; P.b2 = a + net->off[3];
  P.fw1 = a + net->off[4]; P.fb1 = a + net->off[5]; P.fw2 = a + net->off[6]; P.fb2 = a + net->off[7];
  P.fw3 = a + net->off[8]; P.fb3 = a + net->off[9];
  return P;
}

int ensure_smem() {
  static bool done = false;
  if (!done) {
    FEDB_CHECK(cudaFuncSetAttribute(lenet_sample_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(lenet_sample_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(lenet_sample_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(lenet_sample_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(lenet_eval_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kEvalSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(lenet_eval_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kEvalSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(update_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUpdateSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(update_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kUpdateSmemBytes));
    done = true;
  }
  return 0;
}

int check_offsets(const fedb_lenet_t* net) {
  for (int i = 0; i < 10; ++i)
    FEDB_REQUIRE(net->off[i] % 4 == 0, "fedb_lenet: parameter offsets must be multiples of 4 elements (16-byte loads)");
  FEDB_REQUIRE(((uintptr_t)net->param & 15) == 0, "fedb_lenet: arena must be 16-byte aligned");
  FEDB_REQUIRE(((uintptr_t)net->momentum & 15) == 0, "fedb_lenet: momentum buffer must be 16-byte aligned");
  return 0;
}

constexpr int kNumParams = 62006;
constexpr size_t kFwdSmemBytes = sizeof(float) * (size_t)Sm::fwd_end + 0;   // eval uses the forward part only...
// ...but idx1/idx2 live at the end of the plan, so the full plan is always requested (89 KB, 2 CTAs per SM).

}  // namespace

extern "C" size_t fedb_lenet_workspace_bytes(int B) { return carve(nullptr, nullptr, B); }
extern "C" size_t fedb_lenet_eval_bytes(int B) { return 2 * align_up(sizeof(float) * (size_t)B, 256); }

extern "C" int fedb_lenet_train_step(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                                     void* workspace, float* stats, uint64_t step_seed, void* stream) {
  FEDB_REQUIRE(net && images_u8 && labels && workspace && B > 0, "fedb_lenet_train_step: bad arguments");
  if (int rc = ensure_smem()) return rc;
  if (int rc = check_offsets(net)) return rc;
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  lenet_sample_kernel<true, true><<<B, kThreads, kSmemBytes, s>>>(images_u8, P, W, labels, B, net->augment, step_seed,
                                                                  (const unsigned long long*)net->seed_dev, nullptr);
  UpdateArgs U;
  U.param = net->param; U.momentum = net->momentum; U.dparam = nullptr;
  for (int i = 0; i < 10; ++i) U.off[i] = net->off[i];
  U.lr = net->lr; U.mu = net->mu; U.wd = net->weight_decay; U.B = B; U.stats = stats;
  update_kernel<true><<<kUpdateBlocks, 256, kUpdateSmemBytes, s>>>(U, W);
  FEDB_LAUNCH_OK();
  return 0;
}

// Debug: runs the sample kernel of a training step once and leaves CTA 0's clock64() stamps of the 15 phase
// boundaries in clk_dev[0..14] (SM cycles).  Used by tools/prof_lenet.py --phases.
extern "C" int fedb_lenet_phase_clocks(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                                       void* workspace, long long* clk_dev, void* stream) {
  FEDB_REQUIRE(net && images_u8 && labels && workspace && clk_dev && B > 0, "fedb_lenet_phase_clocks: bad arguments");
  if (int rc = ensure_smem()) return rc;
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  lenet_sample_kernel<true, true><<<B, kThreads, kSmemBytes, (cudaStream_t)stream>>>(
      images_u8, P, W, labels, B, 0, 0ull, nullptr, nullptr, clk_dev);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_lenet_eval(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                               void* workspace, float* stats, void* stream) {
  FEDB_REQUIRE(net && images_u8 && labels && workspace && B > 0, "fedb_lenet_eval: bad arguments");
  if (int rc = ensure_smem()) return rc;
  if (int rc = check_offsets(net)) return rc;
  // only the two per-sample rows are needed: [loss B | correct B], fedb_lenet_eval_bytes(B) bytes
  Workspace W;
  memset(&W, 0, sizeof(W));
  W.loss_rows = (float*)workspace;
  W.corr_rows = (float*)((char*)workspace + align_up(sizeof(float) * (size_t)B, 256));
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  FEDB_REQUIRE(((uintptr_t)images_u8 & 15) == 0, "fedb_lenet_eval: images must be 16-byte aligned");
  lenet_eval_kernel<false><<<(B + kEvalS - 1) / kEvalS, kThreads, kEvalSmemBytes, s>>>(images_u8, P, labels, B, W.loss_rows,
                                                                                       W.corr_rows, nullptr);
  if (stats) stats_kernel<<<1, 256, 0, s>>>(W, B, stats);
  FEDB_LAUNCH_OK();
  return 0;
}

// Evaluation through a work queue (see lenet_eval_kernel<true>).  `counter` is one device word that the caller zeroes
// before the FIRST launch of an evaluation; every launch that shares it scores the chunks it manages to take.  ctas: grid
// size of this launch (a handful for the background launch, the whole machine for the finisher).  No statistics here:
// call fedb_lenet_eval_stats once every launch of the evaluation has finished.
extern "C" int fedb_lenet_eval_queue(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                                     void* workspace, uint32_t* counter, int ctas, void* stream) {
  FEDB_REQUIRE(net && images_u8 && labels && workspace && counter && B > 0 && ctas > 0, "fedb_lenet_eval_queue: bad arguments");
  if (int rc = ensure_smem()) return rc;
  if (int rc = check_offsets(net)) return rc;
  FEDB_REQUIRE(((uintptr_t)images_u8 & 15) == 0, "fedb_lenet_eval_queue: images must be 16-byte aligned");
  float* loss_rows = (float*)workspace;
  float* corr_rows = (float*)((char*)workspace + align_up(sizeof(float) * (size_t)B, 256));
  const Params P = make_params(net);
  const int nchunks = (B + kEvalS - 1) / kEvalS;
  if (ctas > nchunks) ctas = nchunks;
  lenet_eval_kernel<true><<<ctas, kThreads, kEvalSmemBytes, (cudaStream_t)stream>>>(images_u8, P, labels, B, loss_rows, corr_rows,
                                                                                    counter);
  FEDB_LAUNCH_OK();
  return 0;
}

// stats[0] += mean loss, stats[1] += correct, stats[2] += B, stats[3] += 1 from the per-sample rows of the workspace
extern "C" int fedb_lenet_eval_stats(void* workspace, int B, float* stats, void* stream) {
  FEDB_REQUIRE(workspace && stats && B > 0, "fedb_lenet_eval_stats: bad arguments");
  Workspace W;
  memset(&W, 0, sizeof(W));
  W.loss_rows = (float*)workspace;
  W.corr_rows = (float*)((char*)workspace + align_up(sizeof(float) * (size_t)B, 256));
  stats_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(W, B, stats);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_lenet_forward(const fedb_lenet_t* net, const uint8_t* images_u8, const float* images_f32, int B,
                                  void* workspace, float* logits, void* stream) {
  FEDB_REQUIRE(net && (images_u8 || images_f32) && workspace && logits && B > 0, "fedb_lenet_forward: bad arguments");
  if (int rc = ensure_smem()) return rc;
  if (int rc = check_offsets(net)) return rc;
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  if (images_u8)
    lenet_sample_kernel<true, false><<<B, kThreads, kSmemBytes, s>>>(images_u8, P, W, nullptr, B, net->augment, net->seed,
                                                                     nullptr, logits);
  else
    lenet_sample_kernel<false, false><<<B, kThreads, kSmemBytes, s>>>(images_f32, P, W, nullptr, B, 0, 0ull, nullptr, logits);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_lenet_grad(const fedb_lenet_t* net, const float* images_f32, const int64_t* labels, int B,
                               void* workspace, float* dparam, float* loss_out, void* stream) {
  FEDB_REQUIRE(net && images_f32 && labels && workspace && dparam && B > 0, "fedb_lenet_grad: bad arguments");
  if (int rc = ensure_smem()) return rc;
  if (int rc = check_offsets(net)) return rc;
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  lenet_sample_kernel<false, true><<<B, kThreads, kSmemBytes, s>>>(images_f32, P, W, labels, B, 0, 0ull, nullptr, nullptr);
  UpdateArgs U;
  U.param = net->param; U.momentum = net->momentum; U.dparam = dparam;
  for (int i = 0; i < 10; ++i) U.off[i] = net->off[i];
  U.lr = 0.f; U.mu = 0.f; U.wd = 0.f; U.B = B; U.stats = loss_out;   // loss_out: 4 floats like stats
  update_kernel<false><<<kUpdateBlocks, 256, kUpdateSmemBytes, s>>>(U, W);
  FEDB_LAUNCH_OK();
  return 0;
}
