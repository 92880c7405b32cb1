// Stand-alone driver for dcnn_tma_stream_probe (no Python start-up: the whole sweep runs in about a second).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Icsrc scripts/native/tma_probe_main.cu -o scripts/native/tma_probe -lcuda
#include <stdio.h>

#include "../../csrc/conv_tc.cu"
extern "C" const char* dcnn_last_error(void);

static float* A = nullptr;
static float* B = nullptr;
// A: activations-like [2M pixels][128 channels] (1 GiB); B: weights-like [16384 rows][128] (8 MiB, L2 resident)
static const long long kARows = 1 << 21, kBRows = 1 << 14, kCols = 128;
static FILE* out = nullptr;

static void run(int a_rows, int b_rows, int stages, int shared, int ctas, int iters = 1500) {
  long long bpi = 0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int rc = dcnn_tma_stream_probe(A, kARows, kCols, a_rows, B, kBRows, kCols, b_rows, 50, stages, shared, ctas, &bpi, 0);
  if (rc != 0 || cudaDeviceSynchronize() != cudaSuccess) {
    fprintf(out, "A %d B %d stages %d: failed (%s)\n", a_rows, b_rows, stages, dcnn_last_error());
    fflush(out);
    return;
  }
  cudaEventRecord(e0, 0);
  dcnn_tma_stream_probe(A, kARows, kCols, a_rows, B, kBRows, kCols, b_rows, iters, stages, shared, ctas, &bpi, 0);
  cudaEventRecord(e1, 0);
  cudaEventSynchronize(e1);
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const int n = ctas > 0 ? ctas : num_sms();
  const double tot = (double)bpi * iters * n;
  fprintf(out, "A %3d rows  B %3d rows (%s)  stages %d  CTAs %3d: %7.3f ms  %7.2f TB/s chip  %6.1f B/clk/SM @1.965GHz\n",
          a_rows, b_rows, shared ? "shared" : "unique", stages, n, ms, tot / ms / 1e9, tot / n / (ms * 1e-3) / 1.965e9);
  fflush(out);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
}

int main(int argc, char** argv) {
  out = argc > 1 ? fopen(argv[1], "w") : stdout;
  if (!out) out = stdout;
  if (cudaMalloc(&A, kARows * kCols * 4) != cudaSuccess || cudaMalloc(&B, kBRows * kCols * 4) != cudaSuccess) {
    fprintf(out, "cudaMalloc failed\n");
    return 1;
  }
  cudaMemset(A, 0, kARows * kCols * 4);
  cudaMemset(B, 0, kBRows * kCols * 4);
  const int sms = num_sms();
  fprintf(out, "%d SMs\n", sms);
  const int counts[] = {1, 8, 37, 74, 148};
  for (int c : counts) run(128, 128, 4, 1, c < sms ? c : sms);
  const int st[] = {2, 3, 6};
  for (int s : st) run(128, 128, s, 1, sms);
  const int cfg[][3] = {{128, 128, 4}, {256, 128, 3}, {256, 256, 2}, {256, 64, 4}, {128, 0, 6}, {256, 0, 4}, {64, 64, 6}};
  for (auto& c : cfg) {
    run(c[0], c[1], c[2], 1, sms);
    if (c[1] > 0) run(c[0], c[1], c[2], 0, sms);
  }
  if (out != stdout) fclose(out);
  return 0;
}
