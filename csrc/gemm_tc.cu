// tcgen05 / TMEM / TMA GEMM for sm_100a (north star: "conv/GEMM as tcgen05/TMEM tiles
// fed by TMA").  D[M,N] (+)= A[M,K] * B[N,K]^T with fp32 operands fed to the 5th-generation tensor
// cores as TF32 (kind::tf32, fp32 accumulation in tensor memory): activations and weights of the
// framework are fp32, so TMA streams them straight from the NHWC tensors / OIHW weights - no
// conversion pass - and the result stays within tf32 rounding of the reference's fp32 math.
//
// One kernel template covers the three GEMMs of a 1x1 convolution / linear layer:
//   forward   Y[pix,Cout]  = X[pix,Cin]   * W[Cout,Cin]^T        TF32, A and B K-major (as stored)
//   dgrad     dX[pix,Cin]  = dY[pix,Cout] * Wt[Cin,Cout]^T       TF32, K-major with a transposed copy of the (small) weight
//   wgrad     dW[Cout,Cin] = dY[pix,Cout]^T * X[pix,Cin]         BF16 (kind::f16), A and B MN-major exactly as the NHWC
//                                                                 tensors lie in memory, split-K over the pixels
// (TF32 operands in MN-major order need the 32-byte-atom swizzle variant; the weight gradient instead reads
//  bf16 copies of dY and X, which is also what mixed-precision training stores for the backward pass.)
//
// The same kernel runs convolutions as IMPLICIT GEMMs (no im2col buffer): every filter tap is a shifted 4-D TMA box of
// the NHWC activation (TMA's out-of-bounds zero fill is the padding and the channel-tail padding, TMA element strides
// are the stride-2 sampling), the stride-2 data gradient is four parity-class GEMMs with a scattering epilogue, the
// weight gradient is the bf16 MN-major form over shifted boxes, and grouped convolutions run all groups in one launch
// (the group index shifts the channel coordinates).  See the host entry points at the end of the file.
//
// Structure (one CTA per 128 x BN output tile, 6 warps, two CTAs per SM):
//   warp 0      TMA producer: cp.async.bulk.tensor.2d/4d (128B swizzle) into a 3-4 stage smem ring, mbarrier tx-count
//   warp 1      TMEM allocation + MMA issue: one elected lane issues tcgen05.mma.cta_group::1.kind::tf32 / kind::f16,
//               tcgen05.commit releases smem stages and finally signals the epilogue
//   warps 2..5  epilogue: tcgen05.ld (32 lanes x 32 bit) -> registers -> bias / ReLU -> staging through the (now idle)
//               pipeline smem -> whole 512-byte rows to global memory (red.global.add for split-K partial tiles)
// SASS evidence (profiles/sass_evidence.md): UTCHMMA for the MMA, LDTM for tcgen05.ld, UTMALDG.2D/4D for TMA.
#include "common.cuh"
#include "../include/fedb200.h"
#include <cuda.h>

namespace {

constexpr int BM = 128;          // UMMA M
constexpr int STAGES = 4;   // BN <= 64; a 128-wide tile takes 3 so that two CTAs (2 x 97 KB) share an SM: one CTA's
                            // epilogue and pipeline ramp then overlap the other's main loop
// KIND 0: tf32, both operands K-major.   stage = [rows][32 fp32] (128 B rows), 4 MMAs of K=8 per stage
// KIND 1: bf16, both operands MN-major.  stage = chunks of 64 bf16 (128 B) x 64 K-rows, 4 MMAs of K=16 per stage
template <int KIND> struct Kind;
template <> struct Kind<0> { static constexpr int kElem = 4, kBK = 32, kUmmaK = 8, kSteps = 4, kChunk = 32; };
template <> struct Kind<1> { static constexpr int kElem = 2, kBK = 64, kUmmaK = 16, kSteps = 4, kChunk = 64; };
constexpr int kThreads = 192;

// ------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::
          "r"(smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_slot, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
template <int KIND>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (KIND == 0) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// 64-bit shared-memory matrix descriptor, 128-byte swizzle, Blackwell version field = 1
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;          // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;          // SWIZZLE_128B
  return d;
}
// instruction descriptor: D=F32, dense, M=128; KIND 0: A=B=TF32 K-major, KIND 1: A=B=BF16 MN-major
__host__ __device__ constexpr uint32_t make_idesc(int n, int kind) {
  const uint32_t fmt = kind == 0 ? 2u : 1u;
  const uint32_t mn = kind == 0 ? 0u : 1u;
  return (1u << 4) | (fmt << 7) | (fmt << 10) | (mn << 15) | (mn << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(BM >> 4) << 24);
}

struct GemmArgs {
  float* D;
  const float* bias;   // per output column, may be null
  long long ldd;
  int M, N, K;
  int k_per_split;     // K elements handled by one blockIdx.z
  int relu;
  int atomic_out;      // 1: red.add into D (split-K / accumulate)
  // implicit-GEMM 3x3 convolution (stride 1, pad 1), KIND 0 only: A tiles are SHIFTED 4-D TMA boxes of the NHWC
  // activation, one per filter tap, with out-of-bounds rows/columns zero-filled by the TMA unit (= the padding)
  int conv;            // 0: plain GEMM
  int Cin, H, W;       // activation geometry
  int bw, bh, bn;      // box: bw*bh*bn == 128 output pixels = one M tile (bw == W); H, W are the OUTPUT map
  int cstride, cpad, fs;   // convolution stride, padding, taps per filter ROW (tap t -> row t / fs, column t % fs)
  // output scatter of the parity-decomposed data gradient of a stride-2 convolution: tile row (n, p, q) of the
  // [N, H, W] sub-grid is written to pixel (n, p*ostride + oy, q*ostride + ox) of the [N, H*ostride, W*ostride] map
  int ostride, oy, ox;     // ostride <= 1: rows are written in place
  // grouped convolution in ONE launch: the tensor maps cover the whole tensors, the group index (blockIdx.z for the
  // forward / data-gradient form, folded into blockIdx.y for the weight gradient) shifts the channel coordinate of the
  // activation by a_goff, the row coordinate of the other operand by b_goff and the output by d_goff elements.  Tile
  // rows / columns that spill into the next group are multiplied by explicit zeros of the packed filter (forward) or
  // masked by the M / N guards of the epilogue.
  int groups, a_goff, b_goff;
  long long d_goff;
  int gx, gy, gz;          // logical tile grid (M tiles, N tiles [x taps x groups], splits / groups), walked persistently
};

__device__ __forceinline__ long long out_row(const GemmArgs& g, int row) {
  if (g.ostride <= 1) return row;
  const int q = row % g.W, p = (row / g.W) % g.H, n = row / (g.W * g.H);
  return ((long long)(n * g.H + p) * g.ostride + g.oy) * (g.W * g.ostride) + q * g.ostride + g.ox;
}

// smem layout: NST pipeline stages (A tile then B tile, each a multiple of the 1024 B swizzle atom), then a staging tile of
// the epilogue (128 rows x (BN + 4) floats: its own memory, because the main loop of the NEXT tile runs while a tile is
// written out), then the barriers.  One CTA per SM: the stages take what is left of the 227 KB.
#ifndef FEDB_GEMM_CTAS_SMALL
#define FEDB_GEMM_CTAS_SMALL 2      // 2: tiles of at most 64 columns run two CTAs per SM (eight epilogue warps draining the
                                    // memory-bound narrow layers), three stages each; 1: one CTA per SM everywhere.
                                    // Measured on B200: ResNet-18 step 4.10 -> 3.88 ms, VGG16 3.11 -> 3.06 ms with 2.
#endif
template <int BN, int KIND>
struct Smem {
  static constexpr int kCtas = (BN <= 64 && FEDB_GEMM_CTAS_SMALL == 2) ? 2 : 1;
  static constexpr int kBN = (KIND == 1 && BN < 64) ? 64 : BN;                 // MN-major chunks are 64 wide
  static constexpr int kABytes = BM * Kind<KIND>::kBK * Kind<KIND>::kElem;     // 16 KiB
  static constexpr int kBBytes = kBN * Kind<KIND>::kBK * Kind<KIND>::kElem;
  static constexpr int kStage = kABytes + kBBytes;
  static constexpr int kPitch = BN + 4;                                        // floats; +4 keeps 128-bit accesses conflict-free
  static constexpr int kStaging = (BM * kPitch * 4 + 1023) / 1024 * 1024;
  static constexpr int kBudget = (kCtas == 2 ? 110 : 200) * 1024 - kStaging;
  static constexpr int kStages = kBudget / kStage > 8 ? 8 : kBudget / kStage;
  static constexpr int kTotal = kStages * kStage + kStaging + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct TileCoord { int bx, by, bz; };
// tile order: by fastest (CTAs that run at the same time share the activation rows of one M tile through L2), then bz, then bx
__device__ __forceinline__ TileCoord tile_coord(const GemmArgs& g, int tile) {
  TileCoord t;
  t.by = tile % g.gy;
  const int r = tile / g.gy;
  t.bz = r % g.gz;
  t.bx = r / g.gz;
  return t;
}

// PERSISTENT kernel: gridDim.x CTAs (one per SM) walk the tiles tile = blockIdx.x, + gridDim.x, ...  The TMA producer runs
// ahead across tile boundaries (the smem ring never drains), the accumulator is double-buffered in TMEM (2 x BN columns), so
// the MMA warp starts tile i+1 while warps 2..5 drain tile i.
template <int BN, int KIND>
__global__ void __launch_bounds__(kThreads, Smem<BN, KIND>::kCtas)
gemm_tc_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, GemmArgs g) {
  constexpr int NST = Smem<BN, KIND>::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  float* staging = reinterpret_cast<float*>(smem + NST * Smem<BN, KIND>::kStage);
  uint64_t* full = (uint64_t*)(smem + NST * Smem<BN, KIND>::kStage + Smem<BN, KIND>::kStaging);
  uint64_t* empty = full + NST;
  uint64_t* accum_full = empty + NST;       // [2] MMA -> epilogue: accumulator buffer complete
  uint64_t* accum_empty = accum_full + 2;   // [2] epilogue -> MMA: accumulator buffer drained
  uint32_t* tmem_slot = (uint32_t*)(accum_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // conv wgrad (g.conv == 2): by enumerates (group, filter tap, N tile); the tap selects the spatial shift of the
  // X operand and the column block [tap*Cin, (tap+1)*Cin) of the packed gradient dWk[Cout][9*Cin]
  const int n_tiles = (g.N + BN - 1) / BN;
  const bool grouped = g.groups > 1;
  const int ntaps = g.fs * g.fs;
  constexpr int BK = Kind<KIND>::kBK;
  constexpr int kChunkBytes = BK * 128;          // one MN-major chunk: BK K-rows of 128 bytes
  constexpr uint32_t kTmemCols = BN < 32 ? 32 : BN;   // power of two >= 32, per accumulator buffer
  const int total_tiles = g.gx * g.gy * g.gz;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&accum_full[b], 1); mbar_init(&accum_empty[b], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // per tile, identical in every role
  auto k_range = [&](const TileCoord& t, int& k_begin, int& num_k) {
    k_begin = (grouped && g.conv == 1) ? 0 : t.bz * g.k_per_split;
    const int k_end = min(g.K, k_begin + g.k_per_split);
    num_k = (k_end - k_begin + BK - 1) / BK;
    if (num_k < 0) num_k = 0;
  };

  if (warp == 0) {
    // ===================================================================== TMA producer
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
      uint32_t it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const TileCoord tc = tile_coord(g, tile);
        const int ytap = g.conv == 2 ? tc.by / n_tiles : 0;
        const int tap = grouped ? ytap % ntaps : ytap;
        const int grp = !grouped ? 0 : (g.conv == 2 ? ytap / ntaps : tc.bz);
        const int m0 = tc.bx * BM, n0 = (g.conv == 2 ? tc.by % n_tiles : tc.by) * BN;
        int k_begin, num_k;
        k_range(tc, k_begin, num_k);
        for (int i = 0; i < num_k; ++i, ++it) {
          const int s = it % NST;
          const uint32_t ph = (it / NST) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          uint8_t* a_dst = smem + s * Smem<BN, KIND>::kStage;
          uint8_t* b_dst = a_dst + Smem<BN, KIND>::kABytes;
          mbar_expect_tx(&full[s], Smem<BN, KIND>::kStage);
          const int k0 = k_begin + i * BK;
          if constexpr (KIND == 0) {
            if (g.conv) {
              // K index = (tap, channel chunk); the weight matrix is [Cout][9*Cin] with the same ordering
              const int per_tap = g.Cin / BK;
              const int kk = k0 / BK;
              const int ktap = kk / per_tap, c0 = (kk % per_tap) * BK;
              const int r = ktap / g.fs, sx = ktap % g.fs;
              int n_img, h0;
              if (g.bn == 1) { const int tpi = g.H / g.bh; n_img = tc.bx / tpi; h0 = (tc.bx % tpi) * g.bh; }
              else           { n_img = tc.bx * g.bn; h0 = 0; }
              // box [32 ch x bw x bh x bn] output pixels; with stride 2 the tensor map itself skips every other input
              // pixel (elementStrides), so the same 128 x 32 tile lands in shared memory
              tma_load_4d(a_dst, &mapA, &full[s], c0 + grp * g.a_goff, sx - g.cpad, h0 * g.cstride + r - g.cpad, n_img);
              tma_load_2d(b_dst, &mapB, &full[s], k0, n0 + grp * g.b_goff);
            } else {
              tma_load_2d(a_dst, &mapA, &full[s], k0, m0);                   // box [32 fp32 (K) x BM rows]
              tma_load_2d(b_dst, &mapB, &full[s], k0, n0);                   // box [32 fp32 (K) x BN rows]
            }
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c)                                // box [64 bf16 (M) x BK K-rows] per chunk
              tma_load_2d(a_dst + c * kChunkBytes, &mapA, &full[s], m0 + c * 64 + grp * g.b_goff, k0);
            if (g.conv == 2) {
              // K runs over output pixels in tiles of 64 = one [bw x bh x bn] box of the bf16 NHWC activation, shifted
              // by the filter tap (zero fill outside the image = the padding)
              const int pt = k0 / BK;
              int n_img, h0;
              if (g.bn == 1) { const int tpi = g.H / g.bh; n_img = pt / tpi; h0 = (pt % tpi) * g.bh; }
              else           { n_img = pt * g.bn; h0 = 0; }
#pragma unroll
              for (int c = 0; c < Smem<BN, KIND>::kBN / 64; ++c)
                tma_load_4d(b_dst + c * kChunkBytes, &mapB, &full[s], n0 + c * 64 + grp * g.a_goff, tap % g.fs - g.cpad,
                            h0 * g.cstride + tap / g.fs - g.cpad, n_img);
            } else {
#pragma unroll
              for (int c = 0; c < Smem<BN, KIND>::kBN / 64; ++c)
                tma_load_2d(b_dst + c * kChunkBytes, &mapB, &full[s], n0 + c * 64, k0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(BN, KIND);
      uint32_t it = 0, t = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const TileCoord tc = tile_coord(g, tile);
        int k_begin, num_k;
        k_range(tc, k_begin, num_k);
        if (num_k == 0) continue;
        const uint32_t acc = t & 1;
        mbar_wait(&accum_empty[acc], ((t >> 1) & 1) ^ 1);      // the epilogue has drained this buffer (two tiles ago)
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * kTmemCols;
        for (int i = 0; i < num_k; ++i, ++it) {
          const int s = it % NST;
          const uint32_t ph = (it / NST) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint32_t a_base = smem_u32(smem + s * Smem<BN, KIND>::kStage);
          const uint32_t b_base = a_base + Smem<BN, KIND>::kABytes;
#pragma unroll
          for (int k = 0; k < Kind<KIND>::kSteps; ++k) {
            // K-major (tf32): 8 rows x 128 B swizzle atoms, SBO = 1024 B between 8-row groups; one MMA consumes
            //   32 B of K, so the start address advances 32 B inside the swizzled row.
            // MN-major (bf16): 64-element chunks of M/N lie kChunkBytes apart (LBO); 8 K-rows x 128 B form one
            //   1024 B atom (SBO); one MMA consumes 16 K-rows = 2048 B.
            const uint64_t ad = KIND == 0 ? make_smem_desc(a_base + k * 32, 16, 1024)
                                          : make_smem_desc(a_base + k * 2048, kChunkBytes, 1024);
            const uint64_t bd = KIND == 0 ? make_smem_desc(b_base + k * 32, 16, 1024)
                                          : make_smem_desc(b_base + k * 2048, kChunkBytes, 1024);
            umma<KIND>(tmem_d, ad, bd, idesc, (i | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty[s]);          // frees the smem stage once these MMAs have read it
        }
        umma_commit(&accum_full[acc]);     // accumulator complete -> epilogue
        ++t;
      }
    }
  } else {
    // ===================================================================== epilogue (warps 2..5)
    const int q = warp & 3;                                  // TMEM lane quarter this warp may access
    constexpr int PITCH = Smem<BN, KIND>::kPitch;
    // Each lane owns one accumulator ROW (tcgen05.ld 32x32b), but a row-per-lane global store is 32 scattered 64-byte
    // pieces; staging through smem turns the write-out into whole rows.  A warp only ever touches its own 32 rows.
    float* stage = staging + (size_t)(q * 32) * PITCH;
    const bool vec_ok = !g.atomic_out && (g.ldd % 4 == 0) && (g.N % 4 == 0) && (g.d_goff % 4 == 0) &&
                        ((reinterpret_cast<uintptr_t>(g.D) & 15) == 0);
    uint32_t t = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const TileCoord tc = tile_coord(g, tile);
      int k_begin, num_k;
      k_range(tc, k_begin, num_k);
      if (num_k == 0) continue;
      const int ytap = g.conv == 2 ? tc.by / n_tiles : 0;
      const int tap = grouped ? ytap % ntaps : ytap;
      const int grp = !grouped ? 0 : (g.conv == 2 ? ytap / ntaps : tc.bz);
      const int m0 = tc.bx * BM, n0 = (g.conv == 2 ? tc.by % n_tiles : tc.by) * BN;
      const int row = m0 + q * 32 + lane;
      const uint32_t acc = t & 1;
      mbar_wait(&accum_full[acc], (t >> 1) & 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * kTmemCols;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 16) {
        float v[16];
        tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = n0 + c0 + j;
          if (g.bias != nullptr && (grouped || tc.bz == 0) && col < g.N) v[j] += g.bias[grp * g.N + col];
          if (g.relu) v[j] = fmaxf(v[j], 0.f);
        }
        if (vec_ok) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            *reinterpret_cast<float4*>(stage + lane * PITCH + c0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        } else if (row < g.M) {
          float* drow = g.D + grp * g.d_goff + out_row(g, row) * g.ldd + (long long)tap * g.Cin + n0 + c0;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (n0 + c0 + j < g.N) {
              if (g.atomic_out) atomicAdd(drow + j, v[j]);
              else drow[j] = v[j];
            }
          }
        }
      }
      // the accumulator buffer is in registers / smem now: hand it back so the MMA warp can start the tile after next
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&accum_empty[acc]);
      if (vec_ok) {
        constexpr int PER_ROW = BN / 4;                        // float4 per tile row
        constexpr int ROWS_IT = 32 / PER_ROW > 0 ? 32 / PER_ROW : 1;   // rows one warp pass writes (BN = 64: two half-warps)
        const int c4 = lane % PER_ROW, rsub = lane / PER_ROW;
        for (int r0 = 0; r0 < 32; r0 += ROWS_IT) {
          const int r = r0 + rsub;
          const int grow = m0 + q * 32 + r;
          if (r < 32 && grow < g.M) {
            const long long orow = out_row(g, grow);
#pragma unroll
            for (int cc = c4; cc < PER_ROW; cc += 32) {
              const int col = n0 + 4 * cc;
              if (col < g.N)
                *reinterpret_cast<float4*>(g.D + grp * g.d_goff + orow * g.ldd + col) =
                    *reinterpret_cast<const float4*>(stage + r * PITCH + 4 * cc);
            }
          }
        }
        __syncwarp();                                          // the staged rows are consumed before the next tile overwrites them
      }
      ++t;
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * kTmemCols);
  }
}

// ------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int get_encode(EncodeTiledFn* out) {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || p == nullptr) return fedb_fail((int)e ? (int)e : -2, __FILE__, __LINE__, "cuTensorMapEncodeTiled not available");
    fn = (EncodeTiledFn)p;
  }
  *out = fn;
  return 0;
}

// 2-D tensor [rows][cols] (elements of `elem` bytes: 4 = fp32, 2 = bf16) with row pitch ld (elements);
// box = [box_cols x box_rows] with box_cols * elem == 128 bytes, 128B swizzle
int make_map_2d(CUtensorMap* map, const void* base, int elem, long long rows, long long cols, long long ld, int box_cols,
                int box_rows) {
  EncodeTiledFn enc;
  if (int rc = get_encode(&enc)) return rc;
  FEDB_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * elem) % 16 == 0, "gemm_tc: operand base/pitch must be 16-byte aligned");
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * elem};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, elem == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)base, dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fedb_fail((int)r, __FILE__, __LINE__, "cuTensorMapEncodeTiled failed");
  return 0;
}

// 4-D NHWC activation [N][H][W][C] (fp32: elem 4, bf16: elem 2); box = [128 bytes of channels x bw x bh x bn],
// 128B swizzle, zero fill outside
int make_map_nhwc(CUtensorMap* map, const void* base, int elem, int N, int H, int W, int C, int bw, int bh, int bn,
                  int estride = 1, long long pitch = 0 /* elements between pixels; 0 = C (dense) */) {
  if (pitch <= 0) pitch = C;
  EncodeTiledFn enc;
  if (int rc = get_encode(&enc)) return rc;
  FEDB_REQUIRE(((uintptr_t)base & 15) == 0 && (pitch * elem) % 16 == 0, "conv_tc: activation base/pitch must be 16-byte aligned");
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)pitch * elem, (cuuint64_t)W * pitch * elem, (cuuint64_t)H * W * pitch * elem};
  // with a traversal stride the box SPANS bw*stride input pixels and picks every stride-th one
  cuuint32_t box[4] = {(cuuint32_t)(128 / elem), (cuuint32_t)(bw * estride), (cuuint32_t)(bh * estride), (cuuint32_t)bn};
  cuuint32_t estr[4] = {1, (cuuint32_t)estride, (cuuint32_t)estride, 1};
  CUresult r = enc(map, elem == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fedb_fail((int)r, __FILE__, __LINE__, "cuTensorMapEncodeTiled (NHWC) failed");
  return 0;
}

// mode 0 (forward):  Wk[k][(r*3+s)*C + c] = W[k][c][r][s]
// mode 1 (dgrad):    Wd[c][(r*3+s)*K + k] = W[k][c][2-r][2-s]     (transposed + spatially flipped filter)
// mode 2 (unpack):   dW[k][c][r][s] += dWk[k][(r*3+s)*C + c]      (packed weight gradient back to OIHW)
// Rp = reduction channels rounded up to the K tile (32): the packed matrices of modes 0 / 1 carry zeros in the tail of
// every tap so that a tap always starts on a K-tile boundary (the activation side gets its zeros from TMA).
// Grouped filters (OIHW with I = C channels per group, K output channels over all groups): modes 0 and 2 are unchanged
// (every output row only knows its own group's inputs); mode 1 emits one row per INPUT channel of the whole tensor,
// [groups * C][flipped tap][K/groups padded], with the output channels of that input channel's group.
__global__ void __launch_bounds__(256) pack3x3_kernel(const float* __restrict__ w, float* __restrict__ out, int K, int C, int mode,
                                                     int T /* taps: 9 or 1 */, int Rp, int groups = 1) {
  const int total = mode == 0 ? K * T * Rp : (mode == 1 ? groups * C * T * Rp : K * C * T);
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    if (mode == 0) {
      const int c = i % Rp, tap = (i / Rp) % T, k = i / (T * Rp);
      out[i] = c < C ? w[(k * C + c) * T + tap] : 0.f;
    } else if (mode == 2) {
      const int c = i % C, tap = (i / C) % T, k = i / (T * C);
      out[(k * C + c) * T + tap] += w[i];
    } else {
      const int k = i % Rp, tap = (i / Rp) % T, cg = i / (T * Rp);       // cg: input channel of the whole tensor
      const int Kg = K / groups, grp = cg / C, c = cg - grp * C;
      out[i] = k < Kg ? w[((grp * Kg + k) * C + c) * T + (T - 1 - tap)] : 0.f;
    }
  }
}

// ---- many small layout jobs in ONE launch.  A training step needs, per convolution, the forward packing of the filter, the
// data-gradient packing and the un-packing of the weight gradient; as one launch each they are ~60 latency-bound kernels
// of a ResNet-18 step.  The jobs of a whole model live in a device table (built once, the addresses never change), the
// step runs "pack everything" once after the weights changed and "un-pack every gradient" once after the backward pass.
struct PackJob {
  const float* src;
  float* dst;
  long long total;      // elements of the iteration space
  int kind;             // 0 / 1: pack3x3_kernel modes; 2: dst(OIHW) += src(packed), src = 0; 3: stride-2 data-gradient
                        // classes (pack_dgrad_s2_kernel); 4: transpose src[K][C] -> dst[C][K]
  int K, C, T, Rp, groups;
  int chunk0;           // index of this job's first chunk of kPackChunk elements in the flat grid (prefix sum over the jobs)
  int pad_;
};
constexpr int kPackChunk = 9216;   // elements per CTA: the flat grid gives every job CTAs in proportion to its size;
                                   // 9216 = 32 filters x 32 channels x 9 taps, the smem tile of the coalesced forms below

// value of element i of a gather job (kinds 0, 1, 3, 4: dst[i] = ...)
__device__ __forceinline__ float pack_job_value(const PackJob& j, long long i) {
  const int K = j.K, C = j.C, T = j.T, Rp = j.Rp;
  if (j.kind == 0) {
    const int c = (int)(i % Rp), tap = (int)((i / Rp) % T), k = (int)(i / ((long long)T * Rp));
    return c < C ? j.src[((long long)k * C + c) * T + tap] : 0.f;
  }
  if (j.kind == 1) {
    const int k = (int)(i % Rp), tap = (int)((i / Rp) % T), cg = (int)(i / ((long long)T * Rp));
    const int Kg = K / j.groups, grp = cg / C, c = cg - grp * C;
    return k < Kg ? j.src[(((long long)grp * Kg + k) * C + c) * T + (T - 1 - tap)] : 0.f;
  }
  if (j.kind == 3) {
    const int Cw = C * j.groups;
    const long long KC = (long long)Rp * Cw;
    const int unit = (int)(i / KC);
    const int cls = unit < 1 ? 0 : (unit < 3 ? 1 : (unit < 5 ? 2 : 3));
    const int base = cls == 0 ? 0 : (cls == 1 ? 1 : (cls == 2 ? 3 : 5));
    const int a = cls >> 1, b = cls & 1;
    const int Tc = (a + 1) * (b + 1), fc = b + 1;
    const long long jj = i - base * KC;
    const int k = (int)(jj % Rp), t = (int)((jj / Rp) % Tc), c = (int)(jj / ((long long)Rp * Tc));
    const int rr = t / fc, ss = t % fc;
    const int r = a == 0 ? 1 : (rr == 0 ? 2 : 0), sx = b == 0 ? 1 : (ss == 0 ? 2 : 0);
    const int grp = c / C, cl = c - grp * C;
    return k < K ? j.src[((((long long)grp * K + k) * C + cl) * 3 + r) * 3 + sx] : 0.f;
  }
  const int k = (int)(i % K), c = (int)(i / K);
  return j.src[(long long)k * C + c];
}

__device__ __forceinline__ void pack_job_elem(const PackJob& j, long long i) {
  const int K = j.K, C = j.C, T = j.T, Rp = j.Rp;
  if (j.kind == 0) {
    const int c = (int)(i % Rp), tap = (int)((i / Rp) % T), k = (int)(i / ((long long)T * Rp));
    j.dst[i] = c < C ? j.src[((long long)k * C + c) * T + tap] : 0.f;
  } else if (j.kind == 1) {
    const int k = (int)(i % Rp), tap = (int)((i / Rp) % T), cg = (int)(i / ((long long)T * Rp));
    const int Kg = K / j.groups, grp = cg / C, c = cg - grp * C;
    j.dst[i] = k < Kg ? j.src[(((long long)grp * Kg + k) * C + c) * T + (T - 1 - tap)] : 0.f;
  } else if (j.kind == 2) {
    const int c = (int)(i % C), tap = (int)((i / C) % T), k = (int)(i / ((long long)T * C));
    float* srcw = const_cast<float*>(j.src);
    j.dst[((long long)k * C + c) * T + tap] += srcw[i];
    srcw[i] = 0.f;                                       // the split-K partial sums of the next step start from zero
  } else if (j.kind == 3) {
    // K = output channels per group (real), Rp = K rounded up to 32, C = input channels per group
    const int Cw = C * j.groups;
    const long long KC = (long long)Rp * Cw;
    const int unit = (int)(i / KC);
    const int cls = unit < 1 ? 0 : (unit < 3 ? 1 : (unit < 5 ? 2 : 3));
    const int base = cls == 0 ? 0 : (cls == 1 ? 1 : (cls == 2 ? 3 : 5));
    const int a = cls >> 1, b = cls & 1;
    const int Tc = (a + 1) * (b + 1), fc = b + 1;
    const long long jj = i - base * KC;
    const int k = (int)(jj % Rp), t = (int)((jj / Rp) % Tc), c = (int)(jj / ((long long)Rp * Tc));
    const int rr = t / fc, ss = t % fc;
    const int r = a == 0 ? 1 : (rr == 0 ? 2 : 0), sx = b == 0 ? 1 : (ss == 0 ? 2 : 0);
    const int grp = c / C, cl = c - grp * C;
    j.dst[i] = k < K ? j.src[((((long long)grp * K + k) * C + cl) * 3 + r) * 3 + sx] : 0.f;
  } else {
    const int k = (int)(i % K), c = (int)(i / K);
    j.dst[i] = j.src[(long long)k * C + c];
  }
}

// 3x3 filters with K % 32 == 0 and C % 32 == 0 (kinds 5, 6, 7 = tiled forms of kinds 0, 1, 2): one CTA moves a tile of
// 32 filters x 32 channels x 9 taps through shared memory, so that both the OIHW side ([k][c][tap]: 288 contiguous floats
// per filter) and the packed side (32 contiguous channels / filters per tap) are read and written in whole 128-byte lines.
// The element-wise forms gather with a stride of 9 floats and were most of the step's packing time.
__device__ __forceinline__ void pack_tile_3x3(const PackJob& j, int tile, float* __restrict__ sm /* [32][289] */) {
  // Every warp owns four of the tile's 32 rows and walks them with compile-time trip counts: no per-element index division
  // (the first version spent its time on e / 288 and % 9), 36 independent loads per thread and phase.
  const int K = j.K, C = j.C, cb = C / 32;
  const int k0 = (tile / cb) * 32, c0 = (tile % cb) * 32;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float v[36];
  if (j.kind == 7) {
    // src = dWk [k][tap][c] (cleared), dst = dW OIHW (+=)
    float* __restrict__ srcw = const_cast<float*>(j.src);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float* row = srcw + ((long long)(k0 + warp * 4 + q) * 9) * C + c0 + lane;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) v[q * 9 + tap] = row[(long long)tap * C];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float* row = srcw + ((long long)(k0 + warp * 4 + q) * 9) * C + c0 + lane;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        sm[(warp * 4 + q) * 289 + lane * 9 + tap] = v[q * 9 + tap];
        row[(long long)tap * C] = 0.f;
      }
    }
    __syncthreads();
    float* __restrict__ dst = j.dst;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float* row = dst + ((long long)(k0 + warp * 4 + q) * C + c0) * 9 + lane;
#pragma unroll
      for (int i = 0; i < 9; ++i) v[q * 9 + i] = row[32 * i];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      float* row = dst + ((long long)(k0 + warp * 4 + q) * C + c0) * 9 + lane;
#pragma unroll
      for (int i = 0; i < 9; ++i) row[32 * i] = v[q * 9 + i] + sm[(warp * 4 + q) * 289 + lane + 32 * i];
    }
    return;
  }
  const float* __restrict__ src = j.src;
#pragma unroll
  for (int q = 0; q < 4; ++q) {                                   // OIHW tile, 288 contiguous floats per filter
    const float* row = src + ((long long)(k0 + warp * 4 + q) * C + c0) * 9 + lane;
#pragma unroll
    for (int i = 0; i < 9; ++i) v[q * 9 + i] = row[32 * i];
  }
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int i = 0; i < 9; ++i) sm[(warp * 4 + q) * 289 + lane + 32 * i] = v[q * 9 + i];
  __syncthreads();
  float* __restrict__ dst = j.dst;
  if (j.kind == 5) {                                              // dst [k][tap][c]   (Rp == C): lane = channel
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int k = warp * 4 + q;
      float* row = dst + ((long long)(k0 + k) * 9) * C + c0 + lane;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) row[(long long)tap * C] = sm[k * 289 + lane * 9 + tap];
    }
  } else {                                                        // kind 6: dst [c][flipped tap][k]   (Rp == K): lane = filter
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int c = warp * 4 + q;
      float* row = dst + ((long long)(c0 + c) * 9) * K + k0 + lane;
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) row[(long long)tap * K] = sm[lane * 289 + c * 9 + (8 - tap)];
    }
  }
}

__global__ void __launch_bounds__(256) pack_multi_kernel(const PackJob* __restrict__ jobs, int njobs) {
  __shared__ float sm_tile[32 * 289];
  __shared__ int s_job;
  // the job of this chunk: the last one whose first chunk is <= blockIdx.x.  Every thread looks at one or a few table
  // entries at once (one memory round trip; a binary search by every thread was six dependent loads per CTA)
  for (int t = threadIdx.x; t < njobs; t += 256) {
    const int first = jobs[t].chunk0;
    const int next = t + 1 < njobs ? jobs[t + 1].chunk0 : 0x7fffffff;
    if (first <= (int)blockIdx.x && (int)blockIdx.x < next) s_job = t;
  }
  __syncthreads();
  const PackJob j = jobs[s_job];
  if (j.kind >= 5) {                     // tiled forms: one chunk == one tile
    pack_tile_3x3(j, blockIdx.x - j.chunk0, sm_tile);
    return;
  }
  const long long begin = (long long)(blockIdx.x - j.chunk0) * kPackChunk;
  const long long end = begin + kPackChunk < j.total ? begin + kPackChunk : j.total;
  if (j.kind == 2) {
    for (long long i = begin + threadIdx.x; i < end; i += 256) pack_job_elem(j, i);
    return;
  }
  // pure gathers (dst[i] = f(src)): four independent elements per thread in flight
  for (long long i0 = begin + threadIdx.x; i0 < end; i0 += 4 * 256) {
    float v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * 256;
      v[u] = i < end ? pack_job_value(j, i) : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = i0 + u * 256;
      if (i < end) j.dst[i] = v[u];
    }
  }
}

template <int BN, int KIND>
int launch(const CUtensorMap& ma, const CUtensorMap& mb, const GemmArgs& g, int splits, cudaStream_t s) {
  static bool attr_set = false;
  if (!attr_set) {
    FEDB_CHECK(cudaFuncSetAttribute(gemm_tc_kernel<BN, KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Smem<BN, KIND>::kTotal));
    attr_set = true;
  }
  GemmArgs gp = g;
  gp.gx = fedb_ceil_div(g.M, BM); gp.gy = fedb_ceil_div(g.N, BN); gp.gz = splits;
  const long long tiles = (long long)gp.gx * gp.gy * gp.gz;
  const int slots = fedb_num_sms() * Smem<BN, KIND>::kCtas;
  const int ctas = (int)(tiles < slots ? tiles : slots);
  gemm_tc_kernel<BN, KIND><<<ctas, kThreads, Smem<BN, KIND>::kTotal, s>>>(ma, mb, gp);
  FEDB_LAUNCH_OK();
  return 0;
}

template <int KIND>
int dispatch_bn(int bn, const CUtensorMap& ma, const CUtensorMap& mb, const GemmArgs& g, int splits, cudaStream_t s) {
  switch (bn) {
    case 128: return launch<128, KIND>(ma, mb, g, splits, s);
    case 64: return launch<64, KIND>(ma, mb, g, splits, s);
    case 32: return launch<32, KIND>(ma, mb, g, splits, s);
    default: return fedb_fail(-1, __FILE__, __LINE__, "gemm_tc: unsupported BN");
  }
}

int run_gemm(int kind, const void* A, const void* B, float* D, const float* bias, int M, int N, int K, long long lda,
             long long ldb, long long ldd, int relu, int splits, int accumulate, cudaStream_t s) {
  FEDB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_tc: empty problem");
  FEDB_REQUIRE(splits >= 1, "gemm_tc: splits >= 1");
  FEDB_REQUIRE(!(relu && (splits > 1 || accumulate)), "gemm_tc: ReLU cannot be fused into a partial sum");
  int bn = N >= 128 ? 128 : (N > 32 ? 64 : 32);
  if (kind == 0 && bn == 128 && splits == 1 && (long long)fedb_ceil_div(M, BM) * fedb_ceil_div(N, 128) < fedb_num_sms()) bn = 64;
  const int bk = kind == 0 ? Kind<0>::kBK : Kind<1>::kBK;
  CUtensorMap ma, mb;
  if (kind == 0) {       // A [M][K], B [N][K], fp32
    if (int rc = make_map_2d(&ma, A, 4, M, K, lda, 32, BM)) return rc;
    if (int rc = make_map_2d(&mb, B, 4, N, K, ldb, 32, bn)) return rc;
  } else {               // A [K][M], B [K][N], bf16
    if (int rc = make_map_2d(&ma, A, 2, K, M, lda, 64, bk)) return rc;
    if (int rc = make_map_2d(&mb, B, 2, K, N, ldb, 64, bk)) return rc;
  }
  GemmArgs g;
  g.D = D; g.bias = bias; g.ldd = ldd; g.M = M; g.N = N; g.K = K;
  int kps = (K + splits - 1) / splits;
  kps = (kps + bk - 1) / bk * bk;
  g.k_per_split = kps;
  splits = (K + kps - 1) / kps;
  g.relu = relu;
  g.atomic_out = (splits > 1 || accumulate) ? 1 : 0;
  g.conv = 0; g.Cin = g.H = g.W = g.bw = g.bh = g.bn = 0; g.cstride = 1; g.cpad = 0; g.fs = 1;
  g.ostride = g.oy = g.ox = 0;
  g.groups = 1; g.a_goff = g.b_goff = 0; g.d_goff = 0;
  return kind == 0 ? dispatch_bn<0>(bn, ma, mb, g, splits, s) : dispatch_bn<1>(bn, ma, mb, g, splits, s);
}

__global__ void __launch_bounds__(256) cast_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n4) {
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n4; i += (long long)gridDim.x * 256) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 r;
    r.x = *reinterpret_cast<uint32_t*>(&lo);
    r.y = *reinterpret_cast<uint32_t*>(&hi);
    reinterpret_cast<uint2*>(dst)[i] = r;
  }
}

// out[c][r] = in[r][c]  (small weight matrices only)
__global__ void __launch_bounds__(256) transpose_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols) {
  __shared__ float tile[32][33];
  const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8)
    if (by + j < rows && bx + tx < cols) tile[j][tx] = in[(long long)(by + j) * cols + bx + tx];
  __syncthreads();
  for (int j = ty; j < 32; j += 8)
    if (bx + j < cols && by + tx < rows) out[(long long)(bx + j) * rows + by + tx] = tile[tx][j];
}

}  // namespace

// D[M,N] (+)= A[M,K] * B[N,K]^T, fp32 operands as TF32, both K-contiguous.
extern "C" int fedb_gemm_tf32(const float* A, const float* B, float* D, const float* bias, int M, int N, int K, long long lda,
                              long long ldb, long long ldd, int relu, int splits, int accumulate, void* stream) {
  return run_gemm(0, A, B, D, bias, M, N, K, lda, ldb, ldd, relu, splits, accumulate, (cudaStream_t)stream);
}

// D[M,N] (+)= At[K,M]^T * Bt[K,N], bf16 operands stored with the REDUCTION dimension outermost (the layout of
// NHWC activations / output gradients when the reduction runs over pixels), fp32 accumulation and output.
extern "C" int fedb_gemm_bf16_tn(const uint16_t* At, const uint16_t* Bt, float* D, int M, int N, int K, long long lda,
                                 long long ldb, long long ldd, int splits, int accumulate, void* stream) {
  return run_gemm(1, At, Bt, D, nullptr, M, N, K, lda, ldb, ldd, 0, splits, accumulate, (cudaStream_t)stream);
}

// Y[N,P,Q,Cout] = conv(X[N,H,W,Cin]; fs x fs filter, stride 1|2, padding) as an implicit GEMM on the tensor cores.
// Wk is the packed filter [Cout][fs*fs*Cin] (fedb_pack_filter mode 0; for 3x3/stride 1, mode 1 + swapped channel
// roles gives the data gradient).  P = H/stride, Q = W/stride ("same"-style geometry of the CIFAR zoo).
// Grouped convolutions run one group per call: X / Y point at the group's first channel and xpitch / ypitch are the
// channel counts of the WHOLE tensors (elements between consecutive pixels).
// groups > 1: X / Y are the WHOLE tensors (xpitch / ypitch their channel counts), Cin / Cout are per group, Wk holds the
// filters of all groups ([groups * Cout][taps][Cin padded to 32]); one launch, blockIdx.z = group.
static int conv_tf32_impl(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int Cin, int Cout,
                          int fs, int stride, int pad, int relu, long long xpitch, long long ypitch, int groups, void* stream) {
  FEDB_REQUIRE(Cin % 4 == 0, "fedb_conv_tf32: Cin must be a multiple of 4 (16-byte pixel pitch)");
  FEDB_REQUIRE((fs == 3 && pad == 1) || (fs == 1 && pad == 0), "fedb_conv_tf32: 3x3/pad 1 or 1x1/pad 0");
  FEDB_REQUIRE(stride == 1 || stride == 2, "fedb_conv_tf32: stride 1 or 2");
  const int P = H / stride, Q = W / stride;
  FEDB_REQUIRE(Q <= 32 && Q >= 2 && (Q & (Q - 1)) == 0 && (P & (P - 1)) == 0 && P * stride == H && Q * stride == W,
               "fedb_conv_tf32: output maps must be powers of two, width <= 32");
  const int bw = Q;
  int bh = 128 / bw;
  if (bh > P) bh = P;
  const int bn = 128 / (bw * bh);
  FEDB_REQUIRE(bw * bh * bn == 128 && (bn == 1 || bh == P), "fedb_conv_tf32: unsupported tile geometry");
  // every tap occupies ceil(Cin / 32) K tiles: the activation tail comes back as zeros from TMA (channel index out of
  // bounds), the packed filter carries explicit zeros (fedb_pack_filter); a 1x1 filter needs no packing at all
  const int cin_pad = (Cin + 31) / 32 * 32;
  const int M = N * P * Q, K = fs * fs * cin_pad;
  int bnt = Cout >= 128 ? 128 : (Cout > 32 ? 64 : 32);
  // few pixels, many channels (the 8x8 / 4x4 stages): 128-wide tiles would leave most SMs without a tile while each tile walks
  // a K loop of thousands; 64-wide tiles double the tile count and halve the MMA time per tile
  if (bnt == 128 && (long long)fedb_ceil_div(M, BM) * fedb_ceil_div(Cout, 128) * groups < fedb_num_sms()) bnt = 64;
  CUtensorMap ma, mb;
  if (int rc = make_map_nhwc(&ma, X, 4, N, H, W, groups > 1 ? (int)xpitch : Cin, bw, bh, bn, stride, xpitch)) return rc;
  if (fs == 1) { if (int rc = make_map_2d(&mb, Wk, 4, (long long)Cout * groups, Cin, Cin, 32, bnt)) return rc; }
  else if (int rc = make_map_2d(&mb, Wk, 4, (long long)Cout * groups, K, K, 32, bnt)) return rc;
  GemmArgs g;
  g.D = Y; g.bias = bias; g.ldd = ypitch > 0 ? ypitch : Cout; g.M = M; g.N = Cout; g.K = K; g.k_per_split = K; g.relu = relu;
  g.atomic_out = 0;
  g.conv = 1; g.Cin = cin_pad; g.H = P; g.W = Q; g.bw = bw; g.bh = bh; g.bn = bn; g.cstride = stride; g.cpad = pad; g.fs = fs;
  g.ostride = g.oy = g.ox = 0;
  g.groups = groups; g.a_goff = groups > 1 ? Cin : 0; g.b_goff = groups > 1 ? Cout : 0; g.d_goff = groups > 1 ? Cout : 0;
  return dispatch_bn<0>(bnt, ma, mb, g, groups, (cudaStream_t)stream);
}

extern "C" int fedb_conv_tf32_g(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int Cin,
                                int Cout, int fs, int stride, int pad, int relu, long long xpitch, long long ypitch, void* stream) {
  return conv_tf32_impl(X, Wk, Y, bias, N, H, W, Cin, Cout, fs, stride, pad, relu, xpitch, ypitch, 1, stream);
}

// All groups of a grouped convolution in one launch.  C, K: channel counts of the whole tensors.
extern "C" int fedb_conv_tf32_grouped(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int C,
                                      int K, int groups, int fs, int stride, int pad, int relu, void* stream) {
  FEDB_REQUIRE(groups >= 1 && C % groups == 0 && K % groups == 0, "fedb_conv_tf32_grouped: channels must divide into the groups");
  return conv_tf32_impl(X, Wk, Y, bias, N, H, W, C / groups, K / groups, fs, stride, pad, relu, C, K, groups, stream);
}

// Filters of the four output-parity classes of a stride-2 3x3/pad-1 data gradient, K-major for the GEMM:
//   class (a, b), a = h % 2, b = w % 2 of the dX pixel; rows: a == 0 -> filter row 1 (dY row p'), a == 1 -> filter row 2
//   (dY row p') then row 0 (dY row p'+1); columns likewise.  out = [class][c][tap][k], classes (0,0),(0,1),(1,0),(1,1)
//   with 1, 2, 2, 4 taps: class offsets 0, 1, 3, 5 in units of C*K floats.
__global__ void __launch_bounds__(256) pack_dgrad_s2_kernel(const float* __restrict__ w, float* __restrict__ out, int Kreal, int Cg,
                                                           int K /* Kreal rounded up to 32 */, int groups) {
  // Kreal: output channels PER GROUP, Cg: input channels per group, C: input channels of the whole tensor (rows)
  const int C = Cg * groups;
  const int total = 9 * K * C;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    const int unit = i / (K * C);                       // 0..8: which (class, tap)
    const int cls = unit < 1 ? 0 : (unit < 3 ? 1 : (unit < 5 ? 2 : 3));
    const int base = cls == 0 ? 0 : (cls == 1 ? 1 : (cls == 2 ? 3 : 5));
    const int a = cls >> 1, b = cls & 1;
    const int T = (a + 1) * (b + 1), fc = b + 1;
    const int j = i - base * K * C;                     // index inside the class: (c * T + t) * K + k
    const int k = j % K, t = (j / K) % T, c = j / (K * T);
    const int rr = t / fc, ss = t % fc;
    const int r = a == 0 ? 1 : (rr == 0 ? 2 : 0), sx = b == 0 ? 1 : (ss == 0 ? 2 : 0);
    const int grp = c / Cg, cl = c - grp * Cg;
    out[i] = k < Kreal ? w[(((grp * Kreal + k) * Cg + cl) * 3 + r) * 3 + sx] : 0.f;
  }
}

extern "C" int fedb_pack_dgrad_s2_grouped(const float* w_oihw, float* out, int K, int C, int groups, void* stream);
extern "C" int fedb_pack_dgrad_s2(const float* w_oihw, float* out, int K, int C, void* stream) {
  return fedb_pack_dgrad_s2_grouped(w_oihw, out, K, C, 1, stream);
}

// K: output channels over all groups, C: input channels per group; out holds 9 * groups * C * pad32(K / groups) floats
extern "C" int fedb_pack_dgrad_s2_grouped(const float* w_oihw, float* out, int K, int C, int groups, void* stream) {
  FEDB_REQUIRE(groups >= 1 && K % groups == 0, "fedb_pack_dgrad_s2: K must divide into the groups");
  const int Kg = K / groups, Kp = (Kg + 31) / 32 * 32;
  int blocks = fedb_ceil_div((long long)Kp * C * groups * 9, 256);
  if (blocks > fedb_num_sms() * 8) blocks = fedb_num_sms() * 8;
  pack_dgrad_s2_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w_oihw, out, Kg, C, Kp, groups);
  FEDB_LAUNCH_OK();
  return 0;
}

// Data gradient of a stride-2 convolution as stride-1 implicit GEMMs over the four parity classes of dX (each class
// sees a dense 1x1 / 1x2 / 2x1 / 2x2 filter over dY), written straight into the interleaved positions of dX by the
// epilogue.  dY [N,P,Q,Cout], dX [N,2P,2Q,Cin].  fs == 3: Wcls from fedb_pack_dgrad_s2.  fs == 1: Wcls = W^T
// ([Cin][Cout]); only class (0,0) receives gradient, the caller zero-fills dX first.
static int conv_dgrad_s2_impl(const float* dY, const float* Wcls, float* dX, int N, int P, int Q, int Cout, int Cin, int fs,
                              long long dypitch, long long dxpitch, int groups, void* stream) {
  FEDB_REQUIRE(Cout % 4 == 0, "fedb_conv_dgrad_s2_tf32: Cout must be a multiple of 4 (16-byte pixel pitch)");
  const int cout_pad = (Cout + 31) / 32 * 32;
  FEDB_REQUIRE(fs == 1 || fs == 3, "fedb_conv_dgrad_s2_tf32: 1x1 or 3x3");
  FEDB_REQUIRE(Q <= 32 && Q >= 2 && (Q & (Q - 1)) == 0 && (P & (P - 1)) == 0, "fedb_conv_dgrad_s2_tf32: dY maps must be powers of two");
  const int bw = Q;
  int bh = 128 / bw;
  if (bh > P) bh = P;
  const int bn = 128 / (bw * bh);
  FEDB_REQUIRE(bw * bh * bn == 128 && (bn == 1 || bh == P), "fedb_conv_dgrad_s2_tf32: unsupported tile geometry");
  int bnt = Cin >= 128 ? 128 : (Cin > 32 ? 64 : 32);
  if (bnt == 128 && (long long)fedb_ceil_div(N * P * Q, BM) * fedb_ceil_div(Cin, 128) * groups < fedb_num_sms()) bnt = 64;
  CUtensorMap ma, mb;
  if (int rc = make_map_nhwc(&ma, dY, 4, N, P, Q, groups > 1 ? (int)dypitch : Cout, bw, bh, bn, 1, dypitch)) return rc;
  const int ncls = fs == 3 ? 4 : 1;
  const int base[4] = {0, 1, 3, 5};
  for (int cls = 0; cls < ncls; ++cls) {
    const int a = cls >> 1, b = cls & 1;
    const int taps = fs == 3 ? (a + 1) * (b + 1) : 1;
    const int K = taps * cout_pad;
    // filters of all groups: [class][groups * Cin rows][taps][Cout padded]
    if (fs == 1) { if (int rc = make_map_2d(&mb, Wcls, 4, (long long)Cin * groups, Cout, Cout, 32, bnt)) return rc; }
    else if (int rc = make_map_2d(&mb, Wcls + (size_t)base[cls] * Cin * groups * cout_pad, 4, (long long)Cin * groups, K, K, 32, bnt)) return rc;
    GemmArgs g;
    g.D = dX; g.bias = nullptr; g.ldd = dxpitch > 0 ? dxpitch : Cin; g.M = N * P * Q; g.N = Cin; g.K = K; g.k_per_split = K; g.relu = 0; g.atomic_out = 0;
    g.conv = 1; g.Cin = cout_pad; g.H = P; g.W = Q; g.bw = bw; g.bh = bh; g.bn = bn; g.cstride = 1; g.cpad = 0; g.fs = b + 1;
    g.ostride = 2; g.oy = a; g.ox = b;
    g.groups = groups; g.a_goff = groups > 1 ? Cout : 0; g.b_goff = groups > 1 ? Cin : 0; g.d_goff = groups > 1 ? Cin : 0;
    if (int rc = dispatch_bn<0>(bnt, ma, mb, g, groups, (cudaStream_t)stream)) return rc;
  }
  return 0;
}

extern "C" int fedb_conv_dgrad_s2_tf32_g(const float* dY, const float* Wcls, float* dX, int N, int P, int Q, int Cout, int Cin,
                                         int fs, long long dypitch, long long dxpitch, void* stream) {
  return conv_dgrad_s2_impl(dY, Wcls, dX, N, P, Q, Cout, Cin, fs, dypitch, dxpitch, 1, stream);
}

extern "C" int fedb_conv_dgrad_s2_tf32(const float* dY, const float* Wcls, float* dX, int N, int P, int Q, int Cout, int Cin,
                                       int fs, void* stream) {
  return conv_dgrad_s2_impl(dY, Wcls, dX, N, P, Q, Cout, Cin, fs, 0, 0, 1, stream);
}

// K, C: channel counts of the whole dY / dX tensors; Wcls from fedb_pack_dgrad_s2_grouped (fs == 3) or the per-group
// transposes [groups * Cg][Kg] (fs == 1, dX zero-filled by the caller).
extern "C" int fedb_conv_dgrad_s2_tf32_grouped(const float* dY, const float* Wcls, float* dX, int N, int P, int Q, int K, int C,
                                               int groups, int fs, void* stream) {
  FEDB_REQUIRE(groups >= 1 && C % groups == 0 && K % groups == 0, "fedb_conv_dgrad_s2_tf32_grouped: channels must divide into the groups");
  return conv_dgrad_s2_impl(dY, Wcls, dX, N, P, Q, K / groups, C / groups, fs, K, C, groups, stream);
}

extern "C" int fedb_conv_tf32(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int Cin,
                              int Cout, int fs, int stride, int pad, int relu, void* stream) {
  return fedb_conv_tf32_g(X, Wk, Y, bias, N, H, W, Cin, Cout, fs, stride, pad, relu, 0, 0, stream);
}

extern "C" int fedb_conv3x3_tf32(const float* X, const float* Wk, float* Y, const float* bias, int N, int H, int W, int Cin,
                                 int Cout, int relu, void* stream) {
  return fedb_conv_tf32(X, Wk, Y, bias, N, H, W, Cin, Cout, 3, 1, 1, relu, stream);
}

// dWk[Cout][fs*fs*Cin] (+)= sum over output pixels of dY[pix][Cout]^T * shift_tap(X)[pix][Cin]: bf16 copies of dY and X
// (NHWC), reduction over the pixels split across CTAs, partial tiles combined with red.global.add (dWk zeroed by the
// caller).  Strided convolutions read X through a tensor map with element strides.
static int conv_wgrad_impl(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int Cin, int Cout, int fs,
                           int stride, int pad, int splits, long long dypitch, long long xpitch, int groups, void* stream) {
  FEDB_REQUIRE((xpitch > 0 ? xpitch : Cin) % 8 == 0 && (dypitch > 0 ? dypitch : Cout) % 8 == 0,
               "fedb_conv_wgrad_bf16: bf16 pixel pitches must be multiples of 8 channels (16 bytes); pad with fedb_cast_bf16_pad");
  FEDB_REQUIRE((fs == 3 && pad == 1) || (fs == 1 && pad == 0), "fedb_conv_wgrad_bf16: 3x3/pad 1 or 1x1/pad 0");
  FEDB_REQUIRE(stride == 1 || stride == 2, "fedb_conv_wgrad_bf16: stride 1 or 2");
  const int P = H / stride, Q = W / stride;
  FEDB_REQUIRE(Q <= 32 && Q >= 2 && (Q & (Q - 1)) == 0 && (P & (P - 1)) == 0 && P * stride == H && Q * stride == W,
               "fedb_conv_wgrad_bf16: output maps must be powers of two, width <= 32");
  const int bw = Q;
  int bh = 64 / bw;
  if (bh > P) bh = P;
  const int bn = 64 / (bw * bh);
  FEDB_REQUIRE(bw * bh * bn == 64 && (bn == 1 || bh == P), "fedb_conv_wgrad_bf16: unsupported tile geometry");
  const int pix = N * P * Q, taps = fs * fs;
  const int bnt = Cin >= 128 ? 128 : 64;
  CUtensorMap ma, mb;
  if (int rc = make_map_2d(&ma, dY, 2, pix, groups > 1 ? dypitch : Cout, dypitch > 0 ? dypitch : Cout, 64, 64)) return rc;
  if (int rc = make_map_nhwc(&mb, X, 2, N, H, W, groups > 1 ? (int)xpitch : Cin, bw, bh, bn, stride, xpitch)) return rc;
  GemmArgs g;
  g.D = dWk; g.bias = nullptr; g.ldd = (long long)taps * Cin; g.M = Cout; g.N = Cin;
  const int ktiles = (pix + 63) / 64;
  if (splits < 1) splits = 1;
  if (splits > ktiles) splits = ktiles;
  const int tiles_per_split = (ktiles + splits - 1) / splits;
  g.K = ktiles * 64; g.k_per_split = tiles_per_split * 64;
  splits = (ktiles + tiles_per_split - 1) / tiles_per_split;
  g.relu = 0; g.atomic_out = 1;
  g.conv = 2; g.Cin = Cin; g.H = P; g.W = Q; g.bw = bw; g.bh = bh; g.bn = bn; g.cstride = stride; g.cpad = pad; g.fs = fs;
  g.ostride = g.oy = g.ox = 0;
  g.groups = groups; g.a_goff = groups > 1 ? Cin : 0; g.b_goff = groups > 1 ? Cout : 0;
  g.d_goff = groups > 1 ? (long long)Cout * taps * Cin : 0;        // dWk of group g: [Cout][taps][Cin] after g such blocks
  static bool attr128 = false, attr64 = false;
  cudaStream_t s = (cudaStream_t)stream;
  g.gx = fedb_ceil_div(Cout, BM); g.gy = fedb_ceil_div(Cin, bnt) * taps * groups; g.gz = splits;
  const long long tiles = (long long)g.gx * g.gy * g.gz;
  const int slots = fedb_num_sms() * (bnt == 128 ? 1 : Smem<64, 1>::kCtas);
  const int ctas = (int)(tiles < slots ? tiles : slots);
  if (bnt == 128) {
    if (!attr128) { FEDB_CHECK(cudaFuncSetAttribute(gemm_tc_kernel<128, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<128, 1>::kTotal)); attr128 = true; }
    gemm_tc_kernel<128, 1><<<ctas, kThreads, Smem<128, 1>::kTotal, s>>>(ma, mb, g);
  } else {
    if (!attr64) { FEDB_CHECK(cudaFuncSetAttribute(gemm_tc_kernel<64, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<64, 1>::kTotal)); attr64 = true; }
    gemm_tc_kernel<64, 1><<<ctas, kThreads, Smem<64, 1>::kTotal, s>>>(ma, mb, g);
  }
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_conv_wgrad_bf16_g(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int Cin, int Cout,
                                      int fs, int stride, int pad, int splits, long long dypitch, long long xpitch, void* stream) {
  return conv_wgrad_impl(dY, X, dWk, N, H, W, Cin, Cout, fs, stride, pad, splits, dypitch, xpitch, 1, stream);
}

extern "C" int fedb_conv_wgrad_bf16(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int Cin, int Cout,
                                    int fs, int stride, int pad, int splits, void* stream) {
  return conv_wgrad_impl(dY, X, dWk, N, H, W, Cin, Cout, fs, stride, pad, splits, 0, 0, 1, stream);
}

// C, K: channel counts of the whole tensors; dWk = [K][taps][C / groups] (the packed OIHW of a grouped filter), zeroed
// by the caller
extern "C" int fedb_conv_wgrad_bf16_grouped(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int C, int K,
                                            int groups, int fs, int stride, int pad, int splits, void* stream) {
  FEDB_REQUIRE(groups >= 1 && C % groups == 0 && K % groups == 0, "fedb_conv_wgrad_bf16_grouped: channels must divide into the groups");
  return conv_wgrad_impl(dY, X, dWk, N, H, W, C / groups, K / groups, fs, stride, pad, splits, K, C, groups, stream);
}

extern "C" int fedb_conv3x3_wgrad_bf16(const uint16_t* dY, const uint16_t* X, float* dWk, int N, int H, int W, int Cin, int Cout,
                                       int splits, void* stream) {
  return fedb_conv_wgrad_bf16(dY, X, dWk, N, H, W, Cin, Cout, 3, 1, 1, splits, stream);
}

// filter (un)packing for fs*fs = taps (9 or 1); see pack3x3_kernel for the modes
// K: output channels over all groups, C: input channels PER GROUP (groups == 1: the plain filter)
extern "C" int fedb_pack_filter_grouped(const float* w_oihw, float* out, int K, int C, int groups, int mode, int taps, void* stream) {
  FEDB_REQUIRE(taps == 9 || taps == 1, "fedb_pack_filter: 9 taps or 1");
  FEDB_REQUIRE(groups >= 1 && K % groups == 0, "fedb_pack_filter: K must divide into the groups");
  // a 1x1 filter is its own K-major matrix: mode 1 with one tap is the plain per-group transpose, unpadded
  const int Rp = mode == 0 ? (C + 31) / 32 * 32 : (mode == 1 ? (taps == 1 ? K / groups : (K / groups + 31) / 32 * 32) : 0);
  const long long total = mode == 0 ? (long long)K * taps * Rp
                                    : (mode == 1 ? (long long)groups * C * taps * Rp : (long long)K * C * taps);
  int blocks = fedb_ceil_div(total, 256);
  if (blocks > fedb_num_sms() * 8) blocks = fedb_num_sms() * 8;
  pack3x3_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w_oihw, out, K, C, mode, taps, Rp, groups);
  FEDB_LAUNCH_OK();
  return 0;
}

// jobs_dev: device array of njobs 56-byte records {src, dst, total, kind, K, C, T, Rp, groups, chunk0, pad}; chunk0 is the
// running sum of ceil(total / FEDB_PACK_CHUNK) over the preceding jobs and total_chunks the sum over all of them (the
// caller computes both once, when it builds the table)
extern "C" int fedb_pack_multi(const void* jobs_dev, int njobs, int total_chunks, void* stream) {
  static_assert(sizeof(PackJob) == 56, "PackJob layout is part of the ABI (fedb200/ops/cuda_gemm.py)");
  static_assert(kPackChunk == FEDB_PACK_CHUNK, "FEDB_PACK_CHUNK in include/fedb200.h");
  if (njobs <= 0 || total_chunks <= 0) return 0;
  FEDB_REQUIRE(jobs_dev != nullptr, "fedb_pack_multi: null job table");
  pack_multi_kernel<<<total_chunks, 256, 0, (cudaStream_t)stream>>>((const PackJob*)jobs_dev, njobs);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_pack_filter(const float* w_oihw, float* out, int K, int C, int mode, int taps, void* stream) {
  return fedb_pack_filter_grouped(w_oihw, out, K, C, 1, mode, taps, stream);
}

extern "C" int fedb_pack3x3(const float* w_oihw, float* out, int K, int C, int mode, void* stream) {
  int blocks = fedb_ceil_div((long long)K * C * 9, 256);
  if (blocks > fedb_num_sms() * 8) blocks = fedb_num_sms() * 8;
  FEDB_REQUIRE((mode == 0 ? C : (mode == 1 ? K : 32)) % 32 == 0, "fedb_pack3x3: unpadded form, use fedb_pack_filter");
  pack3x3_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(w_oihw, out, K, C, mode, 9, mode == 0 ? C : K);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_cast_bf16(const float* src, uint16_t* dst, long long n, void* stream) {
  FEDB_REQUIRE(n % 4 == 0, "fedb_cast_bf16: n must be a multiple of 4");
  long long blocks = (n / 4 + 255) / 256;
  if (blocks > fedb_num_sms() * 16) blocks = fedb_num_sms() * 16;
  if (blocks < 1) blocks = 1;
  cast_bf16_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n / 4);
  FEDB_LAUNCH_OK();
  return 0;
}

// Explicit im2col for convolutions whose reduction is tiny (Cin * fs * fs <= 32: the 3-channel stem of every zoo model).
// Column j = c * fs*fs + r * fs + s, the OIHW flattening of the filter, padded with zeros to 32 (fp32 copy, forward
// GEMM) and to 64 (bf16 copy, weight-gradient GEMM).  Either output may be null.
__global__ void __launch_bounds__(256) im2col_small_kernel(const float* __restrict__ x, float* __restrict__ o32,
                                                          __nv_bfloat16* __restrict__ o16, int N, int H, int W, int C, int fs,
                                                          int stride, int pad, int P, int Q) {
  const long long total = (long long)N * P * Q * 32;
  const int taps = fs * fs, kreal = C * taps;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int j = (int)(i & 31);
    const long long pix = i >> 5;
    float v = 0.f;
    if (j < kreal) {
      const int c = j / taps, t = j - c * taps, r = t / fs, sx = t - r * fs;
      const int q = (int)(pix % Q), p = (int)((pix / Q) % P), n = (int)(pix / ((long long)P * Q));
      const int ih = p * stride - pad + r, iw = q * stride - pad + sx;
      if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = x[(((long long)n * H + ih) * W + iw) * C + c];
    }
    if (o32 != nullptr) o32[i] = v;
    if (o16 != nullptr) {
      o16[pix * 64 + j] = __float2bfloat16(v);
      o16[pix * 64 + 32 + j] = __float2bfloat16(0.f);
    }
  }
}

extern "C" int fedb_im2col_small(const float* x, float* out_f32, uint16_t* out_bf16, int N, int H, int W, int C, int fs,
                                 int stride, int pad, void* stream) {
  FEDB_REQUIRE(C * fs * fs <= 32 && stride >= 1, "fedb_im2col_small: Cin * fs * fs must be <= 32");
  const int P = (H + 2 * pad - fs) / stride + 1, Q = (W + 2 * pad - fs) / stride + 1;
  long long blocks = ((long long)N * P * Q * 32 + 255) / 256;
  if (blocks > fedb_num_sms() * 16) blocks = fedb_num_sms() * 16;
  im2col_small_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(x, out_f32, (__nv_bfloat16*)out_bf16, N, H, W, C, fs, stride,
                                                                     pad, P, Q);
  FEDB_LAUNCH_OK();
  return 0;
}

// fp32 [rows][C] -> bf16 [rows][Cpad] with zero columns: gives activations whose channel count is not a multiple of 8 the
// 16-byte row pitch the bf16 TMA maps of the weight-gradient GEMMs need (DenseNet's 12-channel growth, ShuffleNetV2's 116)
__global__ void __launch_bounds__(256) cast_bf16_pad_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                                           long long rows, int C, int Cpad, long long src_pitch) {
  const long long total = rows * Cpad;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const long long r = i / Cpad;
    const int c = (int)(i - r * Cpad);
    dst[i] = __float2bfloat16(c < C ? src[r * src_pitch + c] : 0.f);
  }
}

// src rows are src_pitch floats apart (0 = C, dense): with src pointing at a group's first channel of an NHWC tensor this
// extracts that group's channels into a dense, zero-padded bf16 matrix
extern "C" int fedb_cast_bf16_pad_p(const float* src, uint16_t* dst, long long rows, int C, int Cpad, long long src_pitch,
                                    void* stream) {
  FEDB_REQUIRE(Cpad >= C && Cpad % 8 == 0, "fedb_cast_bf16_pad: Cpad must be a multiple of 8 and >= C");
  long long blocks = (rows * Cpad + 255) / 256;
  if (blocks > fedb_num_sms() * 16) blocks = fedb_num_sms() * 16;
  if (blocks < 1) blocks = 1;
  cast_bf16_pad_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, rows, C, Cpad,
                                                                      src_pitch > 0 ? src_pitch : C);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_cast_bf16_pad(const float* src, uint16_t* dst, long long rows, int C, int Cpad, void* stream) {
  return fedb_cast_bf16_pad_p(src, dst, rows, C, Cpad, 0, stream);
}

extern "C" int fedb_transpose(const float* in, float* out, int rows, int cols, void* stream) {
  dim3 grid(fedb_ceil_div(cols, 32), fedb_ceil_div(rows, 32));
  transpose_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in, out, rows, cols);
  FEDB_LAUNCH_OK();
  return 0;
}
