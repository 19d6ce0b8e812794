// K9 + K13: softmax cross-entropy forward+backward with device-side accuracy bookkeeping.
// The reference does loss.item() and .sum().item() every step - two host syncs per step
// (src/main.py:153-156); here the counters stay on the device and are read once per pass.
#include "common.cuh"
#include "../include/fedb200.h"

namespace {

// one warp per sample (the entry points accept C <= 4096).  A label outside [0, C) is never dereferenced: the row's loss
// becomes NaN (torch raises for such a label; a kernel cannot, so the poison value makes it visible in the counters).
__global__ void __launch_bounds__(128) ce_kernel(const float* __restrict__ logits, const int64_t* __restrict__ target,
                                                float* __restrict__ dlogits, float* __restrict__ loss_rows,
                                                float* __restrict__ correct_rows, int B, int C) {
  const int lane = threadIdx.x & 31;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (row >= B) return;
  const float* z = logits + (long long)row * C;
  float mx = -INFINITY;
  int arg = 0;
  for (int c = lane; c < C; c += 32) {
    float v = z[c];
    if (v > mx) { mx = v; arg = c; }
  }
  // argmax with first-index tie break, like torch.max(1)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float om = __shfl_xor_sync(0xffffffffu, mx, o);
    int oa = __shfl_xor_sync(0xffffffffu, arg, o);
    if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
  }
  float se = 0.f;
  for (int c = lane; c < C; c += 32) se += expf(z[c] - mx);
  se = warp_sum(se);
  const float lse = logf(se) + mx;
  const long long t64 = target[row];
  const bool t_ok = t64 >= 0 && t64 < C;
  const int t = t_ok ? (int)t64 : -1;
  if (dlogits != nullptr) {
    const float invB = 1.f / (float)B;
    for (int c = lane; c < C; c += 32) {
      float sm = expf(z[c] - lse);
      dlogits[(long long)row * C + c] = (sm - (c == t ? 1.f : 0.f)) * invB;
    }
  }
  if (lane == 0) {
    loss_rows[row] = t_ok ? lse - z[t] : __int_as_float(0x7fc00000);
    correct_rows[row] = (arg == t) ? 1.f : 0.f;
  }
}

// deterministic finish: one block sums the rows in a fixed tree
__global__ void __launch_bounds__(256) ce_finish_kernel(const float* __restrict__ loss_rows,
                                                       const float* __restrict__ correct_rows,
                                                       float* __restrict__ loss_out, float* __restrict__ stats, int B) {
  __shared__ float sl[256], sc[256];
  float l = 0.f, c = 0.f;
  for (int i = threadIdx.x; i < B; i += 256) { l += loss_rows[i]; c += correct_rows[i]; }
  sl[threadIdx.x] = l;
  sc[threadIdx.x] = c;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { sl[threadIdx.x] += sl[threadIdx.x + s]; sc[threadIdx.x] += sc[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (loss_out) *loss_out = sl[0] / (float)B;
    if (stats) { stats[0] += sl[0] / (float)B; stats[1] += sc[0]; stats[2] += (float)B; stats[3] += 1.f; }
  }
}

}  // namespace

// scratch: two float rows of B, kept per device and grown on demand (never inside a capture the
// first time: callers warm up once before capturing).
static float* g_ce_rows[16] = {nullptr};
static int g_ce_cap[16] = {0};

extern "C" int fedb_cross_entropy(const float* logits, const int64_t* target, float* dlogits, float* loss_out,
                                  float* stats, int B, int C, void* stream) {
  FEDB_REQUIRE(B > 0 && C > 0 && C <= 4096, "fedb_cross_entropy: bad shape");
  int dev = 0;
  FEDB_CHECK(cudaGetDevice(&dev));
  FEDB_REQUIRE(dev < 16, "fedb_cross_entropy: device index too large");
  if (g_ce_cap[dev] < B) {
    if (g_ce_rows[dev]) FEDB_CHECK(cudaFree(g_ce_rows[dev]));
    int cap = B < 4096 ? 4096 : B;
    FEDB_CHECK(cudaMalloc(&g_ce_rows[dev], sizeof(float) * 2 * cap));
    g_ce_cap[dev] = cap;
  }
  float* rows = g_ce_rows[dev];
  cudaStream_t s = (cudaStream_t)stream;
  ce_kernel<<<fedb_ceil_div(B, 4), 128, 0, s>>>(logits, target, dlogits, rows, rows + g_ce_cap[dev], B, C);
  FEDB_LAUNCH_OK();
  ce_finish_kernel<<<1, 256, 0, s>>>(rows, rows + g_ce_cap[dev], loss_out, stats, B);
  FEDB_LAUNCH_OK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of fedb_cross_entropy.
//
// Who needs it: hosts that must not (or cannot) ask the driver for memory at the moment they compute
// a loss.  Two kinds show up around this library.  (1) A process that has already handed every byte
// of the GPU to a pool of its own - torch's caching allocator running at its high-water mark next to
// the model it trains, or a serving runtime that carves one fixed arena at start-up: there a
// cudaMalloc from a library either fails with out-of-memory although the pool has room, or it
// fragments the little that is left.  (2) A latency-critical loop (a worker's evaluation inside
// SendModel, a coordinator's round loop): cudaMalloc and cudaFree synchronise the whole device, so a
// lazily growing scratch buffer turns into a stall at an unpredictable step.
// fedb_cross_entropy grows its row scratch with cudaMalloc/cudaFree when the batch grows; this entry
// point never does.  Its scratch is the fixed array below, which lives in the library image: every
// GPU that loads libfedb200.so holds these 128 KiB from load until the process exits.
//
// Rules that follow from the fixed array:
//  * B may not exceed FEDB_CE_STATIC_MAX_ROWS; larger batches are refused with an error (split the
//    batch, or use fedb_cross_entropy).
//  * All calls on one GPU share the array and it keeps whatever the last call left in it.  Nothing in
//    it has to start at zero: each call writes rows [0, B) of both halves before its second kernel
//    reads exactly those rows, so stale rows are never observed and there is nothing to clear.
//    (`stats` accumulates, but that buffer is the caller's, as in fedb_cross_entropy.)
//  * Two calls whose kernels can be in flight at the same time on one GPU (different streams, or
//    different host threads on different streams) race on the array.  Serialise them - same stream,
//    or an event between them.
//  * The FIRST call on each device looks up the array's device address (cudaGetSymbolAddress, with
//    that device current) and caches it per device; make that first call before starting any stream
//    capture on the device.  The name of the array is only a host-side handle for that lookup - the
//    kernels are always given the looked-up pointer.
// FEDB_CE_STATIC_MAX_ROWS (16384) comes from include/fedb200.h: 2 x 16384 floats = 128 KiB
#define FEDB_CE_STATIC_MAX_DEVICES 64
#define FEDB_STR2(x) #x
#define FEDB_STR(x) FEDB_STR2(x)

static __device__ __align__(256) float g_ce_static_rows[2 * FEDB_CE_STATIC_MAX_ROWS];
static float* g_ce_static_addr[FEDB_CE_STATIC_MAX_DEVICES] = {nullptr};

extern "C" int fedb_cross_entropy_static(const float* logits, const int64_t* target, float* dlogits, float* loss_out,
                                         float* stats, int B, int C, void* stream) {
  FEDB_REQUIRE(logits != nullptr && target != nullptr, "fedb_cross_entropy_static: null input");
  FEDB_REQUIRE(B > 0 && C > 0 && C <= 4096, "fedb_cross_entropy_static: bad shape");
  if (B > FEDB_CE_STATIC_MAX_ROWS)
    return fedb_fail(-2, __FILE__, __LINE__,
                     "fedb_cross_entropy_static: batch exceeds FEDB_CE_STATIC_MAX_ROWS (" FEDB_STR(FEDB_CE_STATIC_MAX_ROWS) " rows); "
                     "split the batch or call fedb_cross_entropy");
  int dev = -1;
  FEDB_CHECK(cudaGetDevice(&dev));
  FEDB_REQUIRE(dev >= 0 && dev < FEDB_CE_STATIC_MAX_DEVICES, "fedb_cross_entropy_static: device index out of range");
  if (g_ce_static_addr[dev] == nullptr) {
    // once per device, with that device current: the symbol has one instance per GPU
    float* p = nullptr;
    FEDB_CHECK(cudaGetSymbolAddress((void**)&p, g_ce_static_rows));
    FEDB_REQUIRE(p != nullptr, "fedb_cross_entropy_static: symbol lookup returned null");
    g_ce_static_addr[dev] = p;
  }
  float* loss_rows = g_ce_static_addr[dev];
  float* corr_rows = loss_rows + FEDB_CE_STATIC_MAX_ROWS;
  cudaStream_t s = (cudaStream_t)stream;
  ce_kernel<<<fedb_ceil_div(B, 4), 128, 0, s>>>(logits, target, dlogits, loss_rows, corr_rows, B, C);
  FEDB_LAUNCH_OK();
  ce_finish_kernel<<<1, 256, 0, s>>>(loss_rows, corr_rows, loss_out, stats, B);
  FEDB_LAUNCH_OK();
  return 0;
}
