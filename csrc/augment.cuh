// Device-side crop/flip draw + normalised pixel fetch shared by augment.cu and lenet.cu.
#pragma once
#include <stdint.h>

namespace fedb {

struct Crop { int dx, dy, flip; };   // dx,dy in [-4,4]; zero padding outside the image

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;                       // splitmix64
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ Crop draw_crop(unsigned long long seed, unsigned long long sample, int augment) {
  Crop c{0, 0, 0};
  if (augment) {
    const unsigned long long r = mix64(seed ^ mix64(sample));
    c.dx = (int)(r % 9) - 4;
    c.dy = (int)((r >> 16) % 9) - 4;
    c.flip = (int)((r >> 40) & 1);
  }
  return c;
}

// (pixel/255 - mean)/std of the reference's Normalize; constants from src/main.py:40,45
__device__ __forceinline__ float fetch_pixel(const uint8_t* __restrict__ img_hwc, const Crop& cr, int c, int y, int x) {
  const float mean[3] = {0.4914f, 0.4822f, 0.4465f};
  const float stdv[3] = {0.2023f, 0.1994f, 0.2010f};
  int sx = (cr.flip ? 31 - x : x) + cr.dx;
  int sy = y + cr.dy;
  float v = 0.f;   // the padding is added BEFORE ToTensor/Normalize: a padded pixel is raw 0
  if (sx >= 0 && sx < 32 && sy >= 0 && sy < 32) v = (float)img_hwc[(sy * 32 + sx) * 3 + c];
  return (v / 255.f - mean[c]) / stdv[c];
}

}  // namespace fedb
