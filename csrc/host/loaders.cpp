// Dataset file parsers (ref include/data_loading/mnist_data_loader.hpp:58-106, cifar10_data_loader.hpp:83-120,
// cifar100_data_loader.hpp) -- whole file in one pass, no per-cell std::string / std::stod.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "host_common.h"

// MNIST CSV: header line, then "label,p0,...,p783" per row; pixels scaled by 1/255.
// Returns the number of rows parsed (<= max_rows), or -1 if the file cannot be opened.
DCNN_HOST_API long long dcnn_load_mnist_csv(const char* path, float* images, int* labels, long long max_rows,
                                            int pixels_per_row) {
  FILE* f = fopen(path, "rb");
  if (!f) return -1;
  fseek(f, 0, SEEK_END);
  const long size = ftell(f);
  fseek(f, 0, SEEK_SET);
  std::vector<char> buf((size_t)size + 1);
  if (fread(buf.data(), 1, (size_t)size, f) != (size_t)size) { fclose(f); return -1; }
  fclose(f);
  buf[size] = 0;
  char* p = buf.data();
  char* end = p + size;
  while (p < end && *p != '\n') ++p;  // header
  if (p < end) ++p;
  long long rows = 0;
  while (p < end && rows < max_rows) {
    while (p < end && (*p == '\r' || *p == '\n')) ++p;
    if (p >= end) break;
    char* q;
    const long lab = strtol(p, &q, 10);
    if (q == p) break;
    p = q;
    float* dst = images + rows * (long long)pixels_per_row;
    int got = 0;
    while (p < end && *p == ',' && got < pixels_per_row) {
      ++p;
      const double v = strtod(p, &q);
      p = q;
      dst[got++] = (float)(v / 255.0);
    }
    while (p < end && *p != '\n') ++p;  // ignore surplus cells
    if (got != pixels_per_row) continue;  // invalid row: skip (the reference warns and skips)
    labels[rows++] = (int)lab;
  }
  return rows;
}

// CIFAR binary records: `label_bytes` label bytes (the LAST one is used: CIFAR-10 has 1, CIFAR-100 has coarse+fine) followed
// by `pixels` bytes (3072, channel-major).  Pixels scaled by 1/255.
DCNN_HOST_API long long dcnn_load_cifar_bin(const char* path, int label_bytes, int pixels, float* images, int* labels,
                                            int* coarse_labels, long long max_records) {
  FILE* f = fopen(path, "rb");
  if (!f) return -1;
  const int rec = label_bytes + pixels;
  std::vector<unsigned char> buf((size_t)rec);
  long long n = 0;
  while (n < max_records && fread(buf.data(), 1, (size_t)rec, f) == (size_t)rec) {
    labels[n] = buf[label_bytes - 1];
    if (coarse_labels && label_bytes > 1) coarse_labels[n] = buf[0];
    float* dst = images + n * (long long)pixels;
    for (int i = 0; i < pixels; ++i) dst[i] = buf[label_bytes + i] * (1.0f / 255.0f);
    ++n;
  }
  fclose(f);
  return n;
}

// Gather `count` samples (by index) into a contiguous batch and write one-hot labels.
DCNN_HOST_API void dcnn_gather_batch(const float* images, const int* labels, const long long* indices, int count,
                                     size_t per_sample, int num_classes, float* out_images, float* out_onehot) {
#pragma omp parallel for schedule(static)
  for (int i = 0; i < count; ++i) {
    memcpy(out_images + (size_t)i * per_sample, images + (size_t)indices[i] * per_sample, per_sample * sizeof(float));
    float* oh = out_onehot + (size_t)i * num_classes;
    memset(oh, 0, sizeof(float) * num_classes);
    const int l = labels[indices[i]];
    if (l >= 0 && l < num_classes) oh[l] = 1.0f;
  }
}
