// K14: the input pipeline on the device.  uint8 HWC images (as CIFAR-10 is stored) become
// normalised fp32 tensors, optionally with RandomCrop(32, padding=4) + RandomHorizontalFlip.
// The reference does this per sample on the CPU, also for batches it then skips
// (src/main.py:36-46,142-145).  The random draw is a pure function of (seed, sample index) so the
// forward and backward kernels of a fused step can recompute it instead of storing the crop.
#include "common.cuh"
#include "../include/fedb200.h"
#include "augment.cuh"

namespace {

__global__ void __launch_bounds__(256) prepare_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, int B,
                                                     int nhwc_out, int augment, unsigned long long seed,
                                                     const unsigned long long* __restrict__ seed_dev) {
  const int n = blockIdx.x;
  if (n >= B) return;
  // the device word is the pass counter: a captured graph bakes `seed` in, the word still changes from replay to replay
  if (seed_dev != nullptr) seed ^= *seed_dev * 0x9E3779B97F4A7C15ull;
  const fedb::Crop cr = fedb::draw_crop(seed, (unsigned long long)n, augment);
  const uint8_t* img = src + (long long)n * 3072;
  float* out = dst + (long long)n * 3072;
  for (int e = threadIdx.x; e < 3072; e += 256) {
    int c, y, x;
    if (nhwc_out) { c = e % 3; x = (e / 3) & 31; y = e / 96; }
    else { x = e & 31; y = (e >> 5) & 31; c = e >> 10; }
    out[e] = fedb::fetch_pixel(img, cr, c, y, x);
  }
}

}  // namespace

extern "C" int fedb_prepare_batch(const uint8_t* src_u8, float* dst, int B, int nhwc_out, int augment, uint64_t seed,
                                  void* stream) {
  FEDB_REQUIRE(B > 0, "fedb_prepare_batch: empty batch");
  prepare_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(src_u8, dst, B, nhwc_out, augment, seed, nullptr);
  FEDB_LAUNCH_OK();
  return 0;
}

extern "C" int fedb_prepare_batch_dev(const uint8_t* src_u8, float* dst, int B, int nhwc_out, int augment, uint64_t seed,
                                      const uint64_t* seed_dev, void* stream) {
  FEDB_REQUIRE(B > 0, "fedb_prepare_batch_dev: empty batch");
  prepare_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(src_u8, dst, B, nhwc_out, augment, seed,
                                                      (const unsigned long long*)seed_dev);
  FEDB_LAUNCH_OK();
  return 0;
}
