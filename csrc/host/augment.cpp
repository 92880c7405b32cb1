// Batched image augmentation kernels on NCHW float32 host batches (ref include/data_augmentation/*.hpp).
// The random decisions (which samples, angles, offsets, factors) are drawn by the caller and passed in as arrays, so the
// native path and the pure-torch path are bit-comparable.  Samples are processed in parallel (OpenMP); the reference
// loops over them serially.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "host_common.h"

static inline float clamp01(float v) { return v < 0.f ? 0.f : (v > 1.f ? 1.f : v); }

DCNN_HOST_API void dcnn_aug_flip(float* data, int n, int c, int h, int w, const uint8_t* mask, int vertical) {
#pragma omp parallel for schedule(static)
  for (int b = 0; b < n; ++b) {
    if (!mask[b]) continue;
    float* img = data + (size_t)b * c * h * w;
    for (int ch = 0; ch < c; ++ch) {
      float* p = img + (size_t)ch * h * w;
      if (vertical) {
        for (int y = 0; y < h / 2; ++y)
          for (int x = 0; x < w; ++x) std::swap(p[y * w + x], p[(h - 1 - y) * w + x]);
      } else {
        for (int y = 0; y < h; ++y) std::reverse(p + y * w, p + y * w + w);
      }
    }
  }
}

// data = clamp(data * mul[b] + add[b], 0, 1) for masked samples (brightness: mul=1; contrast: add=0)
DCNN_HOST_API void dcnn_aug_affine_clamp(float* data, int n, size_t per_sample, const uint8_t* mask, const float* mul,
                                         const float* add) {
#pragma omp parallel for schedule(static)
  for (int b = 0; b < n; ++b) {
    if (!mask[b]) continue;
    float* p = data + (size_t)b * per_sample;
    const float m = mul[b], a = add[b];
    for (size_t i = 0; i < per_sample; ++i) p[i] = clamp01(p[i] * m + a);
  }
}

// data = clamp(data + noise, 0, 1) for masked samples; noise has the same shape as data
DCNN_HOST_API void dcnn_aug_add_noise_clamp(float* data, const float* noise, int n, size_t per_sample,
                                            const uint8_t* mask) {
#pragma omp parallel for schedule(static)
  for (int b = 0; b < n; ++b) {
    if (!mask[b]) continue;
    float* p = data + (size_t)b * per_sample;
    const float* q = noise + (size_t)b * per_sample;
    for (size_t i = 0; i < per_sample; ++i) p[i] = clamp01(p[i] + q[i]);
  }
}

// zero-pad by `pad` then take the h x w window starting at (sy[b], sx[b]) of the padded image
DCNN_HOST_API void dcnn_aug_random_crop(float* data, int n, int c, int h, int w, int pad, const uint8_t* mask,
                                        const int* sx, const int* sy) {
#pragma omp parallel
  {
    std::vector<float> tmp((size_t)h * w);
#pragma omp for schedule(static)
    for (int b = 0; b < n; ++b) {
      if (!mask[b]) continue;
      const int dx = sx[b] - pad, dy = sy[b] - pad;  // source = dest + (dx, dy)
      for (int ch = 0; ch < c; ++ch) {
        float* p = data + ((size_t)b * c + ch) * h * w;
        for (int y = 0; y < h; ++y) {
          const int yy = y + dy;
          for (int x = 0; x < w; ++x) {
            const int xx = x + dx;
            tmp[(size_t)y * w + x] = (yy >= 0 && yy < h && xx >= 0 && xx < w) ? p[(size_t)yy * w + xx] : 0.f;
          }
        }
        memcpy(p, tmp.data(), sizeof(float) * h * w);
      }
    }
  }
}

DCNN_HOST_API void dcnn_aug_cutout(float* data, int n, int c, int h, int w, int size, const uint8_t* mask,
                                   const int* x0, const int* y0) {
#pragma omp parallel for schedule(static)
  for (int b = 0; b < n; ++b) {
    if (!mask[b]) continue;
    for (int ch = 0; ch < c; ++ch) {
      float* p = data + ((size_t)b * c + ch) * h * w;
      for (int y = y0[b]; y < y0[b] + size && y < h; ++y)
        for (int x = x0[b]; x < x0[b] + size && x < w; ++x)
          if (y >= 0 && x >= 0) p[(size_t)y * w + x] = 0.f;
    }
  }
}

// bilinear rotation about the image centre, zero fill where any of the four taps falls outside (rotation.hpp:52-97)
DCNN_HOST_API void dcnn_aug_rotate(float* data, int n, int c, int h, int w, const uint8_t* mask,
                                   const float* angle_deg) {
#pragma omp parallel
  {
    std::vector<float> tmp((size_t)h * w);
#pragma omp for schedule(static)
    for (int b = 0; b < n; ++b) {
      if (!mask[b]) continue;
      const float rad = angle_deg[b] * (float)M_PI / 180.0f;
      const float ca = cosf(rad), sa = sinf(rad);
      const float cx = w / 2.0f, cy = h / 2.0f;
      for (int ch = 0; ch < c; ++ch) {
        float* p = data + ((size_t)b * c + ch) * h * w;
        for (int y = 0; y < h; ++y) {
          for (int x = 0; x < w; ++x) {
            const float sx = (x - cx) * ca - (y - cy) * sa + cx;
            const float sy = (x - cx) * sa + (y - cy) * ca + cy;
            const int x1 = (int)floorf(sx), y1 = (int)floorf(sy);
            const int x2 = x1 + 1, y2 = y1 + 1;
            float v = 0.f;
            if (x1 >= 0 && x2 < w && y1 >= 0 && y2 < h) {
              const float wx = sx - x1, wy = sy - y1;
              v = p[(size_t)y1 * w + x1] * (1 - wx) * (1 - wy) + p[(size_t)y1 * w + x2] * wx * (1 - wy) +
                  p[(size_t)y2 * w + x1] * (1 - wx) * wy + p[(size_t)y2 * w + x2] * wx * wy;
            }
            tmp[(size_t)y * w + x] = v;
          }
        }
        memcpy(p, tmp.data(), sizeof(float) * h * w);
      }
    }
  }
}

// per-channel (x - mean) / std; 1-channel images use mean[0]/std[0] (normalization.hpp:22-84)
DCNN_HOST_API void dcnn_aug_normalize(float* data, int n, int c, int h, int w, const float* mean, const float* stdv) {
#pragma omp parallel for schedule(static)
  for (int b = 0; b < n; ++b) {
    for (int ch = 0; ch < c; ++ch) {
      const float m = c == 3 ? mean[ch] : mean[0];
      const float inv = 1.0f / (c == 3 ? stdv[ch] : stdv[0]);
      float* p = data + ((size_t)b * c + ch) * h * w;
      for (size_t i = 0; i < (size_t)h * w; ++i) p[i] = (p[i] - m) * inv;
    }
  }
}
