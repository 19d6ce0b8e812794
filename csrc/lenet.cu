// LeNet training step as four fused kernels (the minimum training slice).
//
//   A  features_fwd   one CTA per sample: u8 image -> normalise(+crop/flip) -> conv1+ReLU+pool ->
//                     conv2+ReLU+pool, everything in shared memory; saves pooled maps + argmax
//   B  classifier     one CTA per kS samples: fc1/fc2/fc3 forward, softmax-CE, backward through the
//                     three linears down to d(flat); keeps dh1/dh2/dlogits for the weight gradients
//   C  features_bwd   one CTA per sample: un-pool, conv2 wgrad + sparse dgrad, conv1 wgrad;
//                     per-sample partial gradients go to scratch (deterministic, no atomics)
//   D  update         one thread per parameter: fc weight gradients as batch reductions, conv
//                     partial reduction, then the SGD update (K10) in the same kernel
//
// All fp32 on CUDA cores: N_out is 6/16/10 and K is 75/150, far below a tcgen05 tile, and fp32
// keeps the step bit-comparable with the reference's CPU run (src/models/lenet.py:5-23,
// src/main.py:142-151).  Summation order over the batch is fixed, so a step is reproducible.
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
constexpr int kS = 4;                          // samples per classifier CTA

struct Workspace {
  float* p1;        // [B][1176]
  float* flat;      // [B][400]
  float* h1;        // [B][120]
  float* h2;        // [B][84]
  float* logits;    // [B][10]
  float* dlogits;   // [B][10]
  float* dh2;       // [B][84]
  float* dh1;       // [B][120]
  float* dflat;     // [B][400]
  float* loss_rows; // [B]
  float* corr_rows; // [B]
  float* conv_part; // [B][2880]
  uint8_t* idx1;    // [B][1176]
  uint8_t* idx2;    // [B][400]
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__host__ size_t carve(Workspace* w, void* base, int B) {
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; };
  char* b = (char*)base;
#define FEDB_TAKE_F(field, n) { size_t o = take(sizeof(float) * (size_t)(n)); if (w) w->field = (float*)(b + o); }
  FEDB_TAKE_F(p1, (size_t)B * NP1)
  FEDB_TAKE_F(flat, (size_t)B * NFLAT)
  FEDB_TAKE_F(h1, (size_t)B * H1)
  FEDB_TAKE_F(h2, (size_t)B * H2)
  FEDB_TAKE_F(logits, (size_t)B * NCLS)
  FEDB_TAKE_F(dlogits, (size_t)B * NCLS)
  FEDB_TAKE_F(dh2, (size_t)B * H2)
  FEDB_TAKE_F(dh1, (size_t)B * H1)
  FEDB_TAKE_F(dflat, (size_t)B * NFLAT)
  FEDB_TAKE_F(loss_rows, (size_t)B)
  FEDB_TAKE_F(corr_rows, (size_t)B)
  FEDB_TAKE_F(conv_part, (size_t)B * CONV_PART)
#undef FEDB_TAKE_F
  { size_t o = take((size_t)B * NP1); if (w) w->idx1 = (uint8_t*)(b + o); }
  { size_t o = take((size_t)B * NFLAT); if (w) w->idx2 = (uint8_t*)(b + o); }
  return off;
}

struct Params {
  const float *w1, *b1, *w2, *b2, *fw1, *fb1, *fw2, *fb2, *fw3, *fb3;
};

// ------------------------------------------------------------------------------------ kernel A
// conv (valid, 5x5) + ReLU + 2x2 max-pool for one output position and CO output channels,
// input planes in shared memory.  Returns pooled values and the argmax slot (0..3).
template <int CI, int CO_PER_THREAD, int IN_W>
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
    const float* src = in + ci * IN_W * IN_W + y0 * IN_W + x0;
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
      for (int s = 0; s < 6; ++s) patch[r][s] = src[r * IN_W + s];
#pragma unroll
    for (int c = 0; c < CO_PER_THREAD; ++c) {
      const float* wk = w + ((co0 + c) * CI + ci) * 25;
#pragma unroll
      for (int r = 0; r < 5; ++r)
#pragma unroll
        for (int s = 0; s < 5; ++s) {
          const float wv = wk[r * 5 + s];
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

template <bool kU8>
__global__ void __launch_bounds__(256) features_fwd_kernel(const void* __restrict__ images, Params P, Workspace W,
                                                          int B, int augment, unsigned long long seed,
                                                          const unsigned long long* __restrict__ seed_dev) {
  __shared__ float s_img[3 * 32 * 32];
  __shared__ float s_w1[NW1], s_b1[C1], s_w2[NW2], s_b2[C2];
  __shared__ float s_p1[NP1];
  const int n = blockIdx.x, tid = threadIdx.x;
  if (n >= B) return;

  if constexpr (kU8) {
    const uint8_t* img = (const uint8_t*)images + (size_t)n * 3072;
    const fedb::Crop cr = fedb::draw_crop(seed ^ (seed_dev ? *seed_dev : 0ull), (unsigned long long)n, augment);
    for (int e = tid; e < 3072; e += 256) {
      const int x = e & 31, y = (e >> 5) & 31, c = e >> 10;
      s_img[e] = fedb::fetch_pixel(img, cr, c, y, x);
    }
  } else {
    const float4* img = (const float4*)((const float*)images + (size_t)n * 3072);
    for (int e = tid; e < 768; e += 256) reinterpret_cast<float4*>(s_img)[e] = img[e];
  }
  for (int e = tid; e < NW1; e += 256) s_w1[e] = P.w1[e];
  for (int e = tid; e < NW2; e += 256) s_w2[e] = P.w2[e];
  if (tid < C1) s_b1[tid] = P.b1[tid];
  if (tid < C2) s_b2[tid] = P.b2[tid];
  __syncthreads();

  // conv1: 196 pooled positions x 6 channels; one position per thread, all channels
  if (tid < P1 * P1) {
    const int py = tid / P1, px = tid % P1;
    float o[C1];
    uint8_t a[C1];
    conv5_relu_pool<3, C1, 32>(s_img, s_w1, s_b1, 0, py, px, o, a);
#pragma unroll
    for (int c = 0; c < C1; ++c) {
      s_p1[c * 196 + tid] = o[c];
      W.p1[(size_t)n * NP1 + c * 196 + tid] = o[c];
      W.idx1[(size_t)n * NP1 + c * 196 + tid] = a[c];
    }
  }
  __syncthreads();

  // conv2: 25 pooled positions x 16 channels; (position, channel pair) per thread -> 200 threads
  if (tid < 200) {
    const int pos = tid % 25, cg = tid / 25;   // cg in 0..7 -> channels 2cg, 2cg+1
    const int py = pos / P2, px = pos % P2;
    float o[2];
    uint8_t a[2];
    conv5_relu_pool<C1, 2, P1>(s_p1, s_w2, s_b2, 2 * cg, py, px, o, a);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const int f = (2 * cg + c) * 25 + pos;     // NCHW flatten order == fc1's input order
      W.flat[(size_t)n * NFLAT + f] = o[c];
      W.idx2[(size_t)n * NFLAT + f] = a[c];
    }
  }
}

// ------------------------------------------------------------------------------------ kernel B
// y[s][j] = act(b[j] + sum_k x[s][k] W[j][k]); one warp per output j, lanes over k.
template <int K, int J, bool kRelu>
__device__ __forceinline__ void fc_forward(const float* __restrict__ Wt, const float* __restrict__ bias,
                                           const float* __restrict__ x_s /*[kS][K] smem*/,
                                           float* __restrict__ y_s /*[kS][J] smem*/, int warp, int lane) {
  for (int j = warp; j < J; j += 8) {
    float acc[kS];
#pragma unroll
    for (int s = 0; s < kS; ++s) acc[s] = 0.f;
    const float* wr = Wt + (size_t)j * K;
    for (int k = lane; k < K; k += 32) {
      const float wv = wr[k];
#pragma unroll
      for (int s = 0; s < kS; ++s) acc[s] = fmaf(wv, x_s[s * K + k], acc[s]);
    }
#pragma unroll
    for (int s = 0; s < kS; ++s) acc[s] = warp_sum(acc[s]);
    if (lane == 0) {
      const float b = bias[j];
#pragma unroll
      for (int s = 0; s < kS; ++s) {
        float v = acc[s] + b;
        y_s[s * J + j] = kRelu ? fmaxf(v, 0.f) : v;
      }
    }
  }
}

// dx[s][k] = (sum_j dy[s][j] W[j][k]) * (x[s][k] > 0 if masked); one thread per k
template <int K, int J, bool kMask>
__device__ __forceinline__ void fc_backward_data(const float* __restrict__ Wt, const float* __restrict__ dy_s,
                                                 const float* __restrict__ x_s, float* __restrict__ dx_s, int tid) {
  for (int k = tid; k < K; k += 256) {
    float acc[kS];
#pragma unroll
    for (int s = 0; s < kS; ++s) acc[s] = 0.f;
#pragma unroll 4
    for (int j = 0; j < J; ++j) {
      const float wv = Wt[(size_t)j * K + k];
#pragma unroll
      for (int s = 0; s < kS; ++s) acc[s] = fmaf(wv, dy_s[s * J + j], acc[s]);
    }
#pragma unroll
    for (int s = 0; s < kS; ++s) dx_s[s * K + k] = (kMask && !(x_s[s * K + k] > 0.f)) ? 0.f : acc[s];
  }
}

template <bool kTrain>
__global__ void __launch_bounds__(256) classifier_kernel(Params P, Workspace W, const int64_t* __restrict__ labels,
                                                        int B, float* __restrict__ logits_out) {
  __shared__ float s_flat[kS * NFLAT];
  __shared__ float s_h1[kS * H1], s_h2[kS * H2], s_lg[kS * NCLS];
  __shared__ float s_dlg[kS * NCLS], s_dh2[kS * H2], s_dh1[kS * H1];
  __shared__ float s_dflat[kS * NFLAT];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * kS;
  const int ns = min(kS, B - n0);
  for (int e = tid; e < kS * NFLAT; e += 256) {
    const int s = e / NFLAT;
    s_flat[e] = s < ns ? W.flat[(size_t)n0 * NFLAT + e] : 0.f;
  }
  __syncthreads();
  fc_forward<NFLAT, H1, true>(P.fw1, P.fb1, s_flat, s_h1, warp, lane);
  __syncthreads();
  fc_forward<H1, H2, true>(P.fw2, P.fb2, s_h1, s_h2, warp, lane);
  __syncthreads();
  fc_forward<H2, NCLS, false>(P.fw3, P.fb3, s_h2, s_lg, warp, lane);
  __syncthreads();

  // softmax-CE per sample: warp s handles sample s
  if (warp < ns) {
    const int n = n0 + warp;
    const float z = lane < NCLS ? s_lg[warp * NCLS + lane] : -INFINITY;
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
    const int t = (int)labels[n];
    if (lane < NCLS) {
      const float sm = expf(z - lse);
      s_dlg[warp * NCLS + lane] = (sm - (lane == t ? 1.f : 0.f)) / (float)B;
      if (logits_out) logits_out[(size_t)n * NCLS + lane] = z;
    }
    if (lane == 0) {
      W.loss_rows[n] = lse - s_lg[warp * NCLS + t];
      W.corr_rows[n] = (arg == t) ? 1.f : 0.f;
    }
  } else if (warp < kS && lane < NCLS) {
    s_dlg[warp * NCLS + lane] = 0.f;
  }
  if constexpr (!kTrain) return;
  __syncthreads();

  fc_backward_data<H2, NCLS, true>(P.fw3, s_dlg, s_h2, s_dh2, tid);
  __syncthreads();
  fc_backward_data<H1, H2, true>(P.fw2, s_dh2, s_h1, s_dh1, tid);
  __syncthreads();
  fc_backward_data<NFLAT, H1, false>(P.fw1, s_dh1, s_flat, s_dflat, tid);
  __syncthreads();

  // publish what kernels C and D need
  for (int e = tid; e < ns * NFLAT; e += 256) W.dflat[(size_t)n0 * NFLAT + e] = s_dflat[e];
  for (int e = tid; e < ns * H1; e += 256) {
    W.h1[(size_t)n0 * H1 + e] = s_h1[e];
    W.dh1[(size_t)n0 * H1 + e] = s_dh1[e];
  }
  for (int e = tid; e < ns * H2; e += 256) {
    W.h2[(size_t)n0 * H2 + e] = s_h2[e];
    W.dh2[(size_t)n0 * H2 + e] = s_dh2[e];
  }
  for (int e = tid; e < ns * NCLS; e += 256) W.dlogits[(size_t)n0 * NCLS + e] = s_dlg[e];
}

// ------------------------------------------------------------------------------------ kernel C
template <bool kU8>
__global__ void __launch_bounds__(256) features_bwd_kernel(const void* __restrict__ images, Params P, Workspace W,
                                                          int B, int augment, unsigned long long seed,
                                                          const unsigned long long* __restrict__ seed_dev) {
  extern __shared__ __align__(16) float smem[];
  float* s_img = smem;                    // 3072
  float* s_p1 = s_img + 3072;             // 1176
  float* s_w2 = s_p1 + NP1;               // 2400
  float* s_g2 = s_w2 + NW2;               // 400  masked d(pool2)
  float* s_g1 = s_g2 + NFLAT;             // 1176 masked d(pool1)
  float* s_dp1 = s_g1 + NP1;              // 8 x 1176 private accumulators
  int* s_pos2 = reinterpret_cast<int*>(s_dp1 + 8 * NP1);   // 400: argmax position y*10+x inside conv2 output
  int* s_pos1 = s_pos2 + NFLAT;                            // 1176: y*28+x inside conv1 output
  const int n = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (n >= B) return;

  if constexpr (kU8) {
    const uint8_t* img = (const uint8_t*)images + (size_t)n * 3072;
    const fedb::Crop cr = fedb::draw_crop(seed ^ (seed_dev ? *seed_dev : 0ull), (unsigned long long)n, augment);
    for (int e = tid; e < 3072; e += 256) {
      const int x = e & 31, y = (e >> 5) & 31, c = e >> 10;
      s_img[e] = fedb::fetch_pixel(img, cr, c, y, x);
    }
  } else {
    const float4* img = (const float4*)((const float*)images + (size_t)n * 3072);
    for (int e = tid; e < 768; e += 256) reinterpret_cast<float4*>(s_img)[e] = img[e];
  }
  for (int e = tid; e < NP1; e += 256) s_p1[e] = W.p1[(size_t)n * NP1 + e];
  for (int e = tid; e < NW2; e += 256) s_w2[e] = P.w2[e];
  for (int e = tid; e < 8 * NP1; e += 256) s_dp1[e] = 0.f;
  for (int e = tid; e < NFLAT; e += 256) {
    const float a = W.flat[(size_t)n * NFLAT + e];
    const int k = W.idx2[(size_t)n * NFLAT + e];
    const int pos = e % 25, py = pos / 5, px = pos % 5;
    s_g2[e] = a > 0.f ? W.dflat[(size_t)n * NFLAT + e] : 0.f;
    s_pos2[e] = (2 * py + (k >> 1)) * 10 + (2 * px + (k & 1));
  }
  __syncthreads();

  float* part = W.conv_part + (size_t)n * CONV_PART;

  // conv2 weight gradient: dW2[co][ci][r][s] = sum_pos g2[co][pos] * p1[ci][y+r][x+s]
  for (int e = tid; e < NW2; e += 256) {
    const int s = e % 5, r = (e / 5) % 5, ci = (e / 25) % C1, co = e / 150;
    const float* src = s_p1 + ci * 196 + r * 14 + s;
    float acc = 0.f;
#pragma unroll 5
    for (int pos = 0; pos < 25; ++pos) {
      const int q = s_pos2[co * 25 + pos];
      acc = fmaf(s_g2[co * 25 + pos], src[(q / 10) * 14 + (q % 10)], acc);
    }
    part[OFF_W2 + e] = acc;
  }
  if (tid < C2) {
    float acc = 0.f;
    for (int pos = 0; pos < 25; ++pos) acc += s_g2[tid * 25 + pos];
    part[OFF_B2 + tid] = acc;
  }

  // conv2 data gradient, sparse scatter: each non-zero of d(conv2 out) touches a 5x5 patch of every
  // input channel.  Warp w owns non-zeros w, w+8, ... and its own accumulator copy; lanes are taps.
  {
    float* acc = s_dp1 + warp * NP1;
    const int r = lane / 5, s = lane % 5;
    for (int e = warp; e < NFLAT; e += 8) {
      const float g = s_g2[e];
      if (g != 0.f && lane < 25) {           // warp-uniform test on g, lane test for the 25 taps
        const int co = e / 25, q = s_pos2[e];
        const int base = ((q / 10) + r) * 14 + (q % 10) + s;
#pragma unroll
        for (int ci = 0; ci < C1; ++ci)
          acc[ci * 196 + base] = fmaf(g, s_w2[(co * C1 + ci) * 25 + lane], acc[ci * 196 + base]);
      }
      __syncwarp();
    }
  }
  __syncthreads();
  for (int e = tid; e < NP1; e += 256) {
    float d = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) d += s_dp1[w * NP1 + e];
    const int k = W.idx1[(size_t)n * NP1 + e];
    const int pos = e % 196, py = pos / 14, px = pos % 14;
    s_g1[e] = s_p1[e] > 0.f ? d : 0.f;
    s_pos1[e] = (2 * py + (k >> 1)) * 28 + (2 * px + (k & 1));
  }
  __syncthreads();

  // conv1 weight gradient: dW1[co][ci][r][s] = sum_pos g1[co][pos] * img[ci][y+r][x+s]; 450 outputs,
  // each split in two halves of the 196 positions so that all 256 threads have work
  __shared__ float s_half[2 * NW1];
  for (int e = tid; e < 2 * NW1; e += 256) {
    const int half = e / NW1, o = e % NW1;
    const int s = o % 5, r = (o / 5) % 5, ci = (o / 25) % 3, co = o / 75;
    const float* src = s_img + ci * 1024 + r * 32 + s;
    float acc = 0.f;
    const int lo = half * 98;
#pragma unroll 7
    for (int pos = lo; pos < lo + 98; ++pos) {
      const int q = s_pos1[co * 196 + pos];
      acc = fmaf(s_g1[co * 196 + pos], src[(q / 28) * 32 + (q % 28)], acc);
    }
    s_half[e] = acc;
  }
  __syncthreads();
  for (int o = tid; o < NW1; o += 256) part[o] = s_half[o] + s_half[NW1 + o];
  if (tid < C1) {
    float acc = 0.f;
    for (int pos = 0; pos < 196; ++pos) acc += s_g1[tid * 196 + pos];
    part[OFF_B1 + tid] = acc;
  }
}

constexpr size_t kBwdSmemBytes = sizeof(float) * (3072 + NP1 + NW2 + NFLAT + NP1 + 8 * NP1) + sizeof(int) * (NFLAT + NP1);

// ------------------------------------------------------------------------------------ kernel D
struct UpdateArgs {
  float* param;
  float* momentum;
  float* dparam;           // grad-only mode (tests): write gradient here, do not update
  long long off[10];
  float lr, mu, wd;
  int B;
  float* stats;            // [loss_sum_of_means, correct, samples, steps]
};

template <bool kApply>
__global__ void __launch_bounds__(256) update_kernel(UpdateArgs U, Workspace W) {
  // parameter index space, in this order: fc1.w fc1.b fc2.w fc2.b fc3.w fc3.b | conv1.w conv1.b conv2.w conv2.b
  constexpr int N_FC1W = H1 * NFLAT, N_FC2W = H2 * H1, N_FC3W = NCLS * H2;
  constexpr int E0 = N_FC1W, E1 = E0 + H1, E2 = E1 + N_FC2W, E3 = E2 + H2, E4 = E3 + N_FC3W, E5 = E4 + NCLS;
  constexpr int E6 = E5 + NW1, E7 = E6 + C1, E8 = E7 + NW2, E9 = E8 + C2;
  const int B = U.B;
  const int i = blockIdx.x * 256 + threadIdx.x;

  if (blockIdx.x == 0 && threadIdx.x < 32 && U.stats != nullptr) {   // K13: counters stay on the device
    float l = 0.f, c = 0.f;
    for (int s = threadIdx.x; s < B; s += 32) { l += W.loss_rows[s]; c += W.corr_rows[s]; }
    l = warp_sum(l);
    c = warp_sum(c);
    if (threadIdx.x == 0) {
      U.stats[0] += l / (float)B;
      U.stats[1] += c;
      U.stats[2] += (float)B;
      U.stats[3] += 1.f;
    }
  }
  if (i >= E9) return;

  float g = 0.f;
  long long dst;
  if (i < E5) {
    const float *dout, *in;
    int J, K, j, k;
    bool is_bias = false;
    if (i < E0)      { J = H1; K = NFLAT; j = i / NFLAT; k = i % NFLAT; dout = W.dh1; in = W.flat; dst = U.off[4] + i; }
    else if (i < E1) { J = H1; K = 0; j = i - E0; k = 0; dout = W.dh1; in = nullptr; is_bias = true; dst = U.off[5] + j; }
    else if (i < E2) { const int e = i - E1; J = H2; K = H1; j = e / H1; k = e % H1; dout = W.dh2; in = W.h1; dst = U.off[6] + e; }
    else if (i < E3) { J = H2; K = 0; j = i - E2; k = 0; dout = W.dh2; in = nullptr; is_bias = true; dst = U.off[7] + j; }
    else if (i < E4) { const int e = i - E3; J = NCLS; K = H2; j = e / H2; k = e % H2; dout = W.dlogits; in = W.h2; dst = U.off[8] + e; }
    else             { J = NCLS; K = 0; j = i - E4; k = 0; dout = W.dlogits; in = nullptr; is_bias = true; dst = U.off[9] + j; }
    if (is_bias) {
#pragma unroll 8
      for (int s = 0; s < B; ++s) g += dout[(size_t)s * J + j];
    } else {
#pragma unroll 8
      for (int s = 0; s < B; ++s) g = fmaf(dout[(size_t)s * J + j], in[(size_t)s * K + k], g);
    }
  } else {
    int c;
    if (i < E6)      { c = i - E5; dst = U.off[0] + c; }
    else if (i < E7) { c = OFF_B1 + (i - E6); dst = U.off[1] + (i - E6); }
    else if (i < E8) { c = OFF_W2 + (i - E7); dst = U.off[2] + (i - E7); }
    else             { c = OFF_B2 + (i - E8); dst = U.off[3] + (i - E8); }
#pragma unroll 8
    for (int s = 0; s < B; ++s) g += W.conv_part[(size_t)s * CONV_PART + c];
  }

  if constexpr (kApply) {
    const float p = U.param[dst];
    const float g1 = __fadd_rn(g, __fmul_rn(U.wd, p));
    const float v = __fadd_rn(__fmul_rn(U.mu, U.momentum[dst]), g1);
    U.momentum[dst] = v;
    U.param[dst] = __fadd_rn(p, __fmul_rn(-U.lr, v));
  } else {
    U.dparam[dst] = g;
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

int ensure_bwd_smem() {
  static bool done[2] = {false, false};
  if (!done[0]) {
    FEDB_CHECK(cudaFuncSetAttribute(features_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kBwdSmemBytes));
    FEDB_CHECK(cudaFuncSetAttribute(features_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)kBwdSmemBytes));
    done[0] = true;
  }
  return 0;
}

constexpr int kNumParams = 62006;

}  // namespace

extern "C" size_t fedb_lenet_workspace_bytes(int B) { return carve(nullptr, nullptr, B); }

extern "C" int fedb_lenet_train_step(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                                     void* workspace, float* stats, uint64_t step_seed, void* stream) {
  FEDB_REQUIRE(net && images_u8 && labels && workspace && B > 0, "fedb_lenet_train_step: bad arguments");
  if (int rc = ensure_bwd_smem()) return rc;
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  features_fwd_kernel<true><<<B, 256, 0, s>>>(images_u8, P, W, B, net->augment, step_seed, (const unsigned long long*)net->seed_dev);
  classifier_kernel<true><<<fedb_ceil_div(B, kS), 256, 0, s>>>(P, W, labels, B, nullptr);
  features_bwd_kernel<true><<<B, 256, kBwdSmemBytes, s>>>(images_u8, P, W, B, net->augment, step_seed, (const unsigned long long*)net->seed_dev);
  UpdateArgs U;
  U.param = net->param; U.momentum = net->momentum; U.dparam = nullptr;
  for (int i = 0; i < 10; ++i) U.off[i] = net->off[i];
  U.lr = net->lr; U.mu = net->mu; U.wd = net->weight_decay; U.B = B; U.stats = stats;
  update_kernel<true><<<fedb_ceil_div(kNumParams, 256), 256, 0, s>>>(U, W);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_lenet_eval(const fedb_lenet_t* net, const uint8_t* images_u8, const int64_t* labels, int B,
                               void* workspace, float* stats, void* stream) {
  FEDB_REQUIRE(net && images_u8 && labels && workspace && B > 0, "fedb_lenet_eval: bad arguments");
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  features_fwd_kernel<true><<<B, 256, 0, s>>>(images_u8, P, W, B, 0, 0ull, nullptr);
  classifier_kernel<false><<<fedb_ceil_div(B, kS), 256, 0, s>>>(P, W, labels, B, nullptr);
  if (stats) stats_kernel<<<1, 256, 0, s>>>(W, B, stats);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_lenet_forward(const fedb_lenet_t* net, const uint8_t* images_u8, const float* images_f32, int B,
                                  void* workspace, float* logits, void* stream) {
  FEDB_REQUIRE(net && (images_u8 || images_f32) && workspace && logits && B > 0, "fedb_lenet_forward: bad arguments");
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  // labels are only needed for the loss rows; point them at a zeroed scratch (dlogits area reused as int64 zeros)
  FEDB_CHECK(cudaMemsetAsync(W.dflat, 0, sizeof(int64_t) * (size_t)B, s));
  if (images_u8) features_fwd_kernel<true><<<B, 256, 0, s>>>(images_u8, P, W, B, net->augment, net->seed, nullptr);
  else features_fwd_kernel<false><<<B, 256, 0, s>>>(images_f32, P, W, B, 0, 0ull, nullptr);
  classifier_kernel<false><<<fedb_ceil_div(B, kS), 256, 0, s>>>(P, W, (const int64_t*)W.dflat, B, logits);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_lenet_grad(const fedb_lenet_t* net, const float* images_f32, const int64_t* labels, int B,
                               void* workspace, float* dparam, float* loss_out, void* stream) {
  FEDB_REQUIRE(net && images_f32 && labels && workspace && dparam && B > 0, "fedb_lenet_grad: bad arguments");
  if (int rc = ensure_bwd_smem()) return rc;
  Workspace W;
  carve(&W, workspace, B);
  const Params P = make_params(net);
  cudaStream_t s = (cudaStream_t)stream;
  features_fwd_kernel<false><<<B, 256, 0, s>>>(images_f32, P, W, B, 0, 0ull, nullptr);
  classifier_kernel<true><<<fedb_ceil_div(B, kS), 256, 0, s>>>(P, W, labels, B, nullptr);
  features_bwd_kernel<false><<<B, 256, kBwdSmemBytes, s>>>(images_f32, P, W, B, 0, 0ull, nullptr);
  UpdateArgs U;
  U.param = net->param; U.momentum = net->momentum; U.dparam = dparam;
  for (int i = 0; i < 10; ++i) U.off[i] = net->off[i];
  U.lr = 0.f; U.mu = 0.f; U.wd = 0.f; U.B = B; U.stats = loss_out;   // loss_out: 4 floats like stats
  update_kernel<false><<<fedb_ceil_div(kNumParams, 256), 256, 0, s>>>(U, W);
  FEDB_LAUNCH_OK();
  return 0;
}
