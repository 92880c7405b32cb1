// Loss and optimizer kernels.
//   * softmax_ce_kernel: loss + gradient + correct-count in ONE pass, one warp per sample, results stay on the
//     device (the reference recomputes max/sum-exp per (sample,class) thread and does 3 cudaMallocAsync + a D2H +
//     cudaStreamSynchronize per call: loss_ops.cu:164-190, 379-419; accuracy.cu:94-111).
//   * elementwise regression losses (MSE / MAE / Huber) and probability cross-entropy.
//   * multi-tensor SGD / SGD-momentum / Adam / AdamW over a device-resident pointer table: one launch for the
//     whole model instead of one per parameter tensor (optimizers.hpp:208-239), with optional gradient zeroing.
#include "common.cuh"

namespace dcnn {

// out[0] += mean loss contribution, out[1] += correct count.  grad may be null.  mode 0: logits (softmax-CE /
// logsoftmax-CE, identical maths in the reference), mode 1: probabilities (CrossEntropyLoss with epsilon clamp).
__global__ void __launch_bounds__(256) softmax_ce_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                         float* __restrict__ grad, float* __restrict__ out, int B,
                                                         int C, int mode, float eps) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= B) return;
  const int lane = threadIdx.x & 31;
  const float* p = pred + (long long)row * C;
  const float* t = tgt + (long long)row * C;
  // argmax of prediction and target in one sweep
  float pm = -INFINITY, tm = -INFINITY;
  int pi = 0x7fffffff, ti = 0x7fffffff;
  for (int c = lane; c < C; c += 32) {
    const float a = p[c], b = t[c];
    if (a > pm) { pm = a; pi = c; }
    if (b > tm) { tm = b; ti = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float a = __shfl_xor_sync(0xffffffffu, pm, o); const int ai = __shfl_xor_sync(0xffffffffu, pi, o);
    if (a > pm || (a == pm && ai < pi)) { pm = a; pi = ai; }
    const float b = __shfl_xor_sync(0xffffffffu, tm, o); const int bi = __shfl_xor_sync(0xffffffffu, ti, o);
    if (b > tm || (b == tm && bi < ti)) { tm = b; ti = bi; }
  }
  const float invB = 1.f / (float)B;
  float loss;
  if (mode == 0) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += expf(p[c] - pm);
    s = warp_sum(s);
    const float lse = logf(s) + pm;
    loss = lse - p[ti];
    if (grad != nullptr) {
      const float inv = 1.f / s;
      float* g = grad + (long long)row * C;
      for (int c = lane; c < C; c += 32) g[c] = (expf(p[c] - pm) * inv - t[c]) * invB;
    }
  } else {
    const float pc = fminf(fmaxf(p[ti], eps), 1.f - eps);
    loss = -logf(pc);
    if (grad != nullptr) {
      float* g = grad + (long long)row * C;
      for (int c = lane; c < C; c += 32) g[c] = (p[c] - t[c]) * invB;
    }
  }
  if (lane == 0) {
    atomicAdd(out + 0, loss * invB);
    atomicAdd(out + 1, pi == ti ? 1.f : 0.f);
  }
}

// kind 0: MSE, 1: MAE, 2: Huber(delta).  out[0] += mean loss.
__global__ void __launch_bounds__(256) regression_loss_kernel(const float* __restrict__ pred,
                                                              const float* __restrict__ tgt, float* __restrict__ grad,
                                                              float* __restrict__ out, long long n, int kind,
                                                              float delta) {
  __shared__ float red[8];
  const float invn = 1.f / (float)n;
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float d = pred[i] - tgt[i];
    float l, g;
    if (kind == 0) { l = d * d; g = 2.f * d * invn; }
    else if (kind == 1) { l = fabsf(d); g = d > 0.f ? invn : -invn; }
    else {
      const float a = fabsf(d);
      if (a <= delta) { l = 0.5f * a * a; g = d * invn; }
      else { l = delta * a - 0.5f * delta * delta; g = (d > 0.f ? delta : -delta) * invn; }
    }
    acc += l;
    if (grad != nullptr) grad[i] = g;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += red[i];
    atomicAdd(out, s * invn);
  }
}

// ---------------------------------------------------------------------------------------------- optimizers
// Pointer table on the device: for tensor t: param[t], grad[t], s1[t], s2[t], numel[t]; block_tensor[b] / block_off[b]
// map each CTA to a (tensor, chunk) so one launch covers every parameter tensor.
struct MtTable {
  float* const* param;
  float* const* grad;
  float* const* s1;  // velocity or Adam m
  float* const* s2;  // Adam v
  const long long* numel;
  const int* block_tensor;
  const long long* block_off;
};
constexpr int kMtChunk = 256 * 4 * 4;  // elements per CTA

// mode 0: SGD, 1: SGD momentum, 2: Adam (L2), 3: AdamW (decoupled)
template <int MODE>
__global__ void __launch_bounds__(256) multi_tensor_update_kernel(MtTable tb, float lr, float a, float b, float eps,
                                                                  float bc1, float bc2, float wd, int zero_grads,
                                                                  const float* __restrict__ hyper) {
  // step-varying scalars may come from device memory so that a captured CUDA graph can be replayed every step
  if (hyper != nullptr) { lr = hyper[0]; bc1 = hyper[1]; bc2 = hyper[2]; }
  const int t = tb.block_tensor[blockIdx.x];
  const long long off = tb.block_off[blockIdx.x];
  const long long n = tb.numel[t];
  float* p = tb.param[t];
  float* g = tb.grad[t];
  float* s1 = MODE >= 1 ? tb.s1[t] : nullptr;
  float* s2 = MODE >= 2 ? tb.s2[t] : nullptr;
  const long long end = min(n, off + kMtChunk);
  for (long long i = off + threadIdx.x; i < end; i += 256) {
    const float gi = g[i];
    float pi = p[i];
    if (MODE == 0) {
      pi -= lr * gi;
    } else if (MODE == 1) {
      const float v = a * s1[i] - lr * gi;
      s1[i] = v;
      pi += v;
    } else {
      const float m = a * s1[i] + (1.f - a) * gi;
      const float v = b * s2[i] + (1.f - b) * gi * gi;
      s1[i] = m;
      s2[i] = v;
      float upd = lr * (m / bc1) / (sqrtf(v / bc2) + eps);
      if (wd > 0.f) {
        if (MODE == 3) pi -= wd * lr * pi;
        else upd += wd * lr * pi;
      }
      pi -= upd;
    }
    p[i] = pi;
    if (zero_grads) g[i] = 0.f;
  }
}

__global__ void fill_kernel(float* p, long long n, float v) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    p[i] = v;
}

}  // namespace dcnn
using namespace dcnn;

DCNN_API int dcnn_softmax_ce(const float* pred, const float* tgt, float* grad, float* out2, int B, int C, int mode,
                             float eps, cudaStream_t stream) {
  DCNN_CHECK_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(float), stream));
  softmax_ce_kernel<<<ceil_div(B, 8), 256, 0, stream>>>(pred, tgt, grad, out2, B, C, mode, eps);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_regression_loss(const float* pred, const float* tgt, float* grad, float* out1, long long n, int kind,
                                  float delta, cudaStream_t stream) {
  DCNN_CHECK_CUDA(cudaMemsetAsync(out1, 0, sizeof(float), stream));
  regression_loss_kernel<<<ew_grid(n, 256, 2), 256, 0, stream>>>(pred, tgt, grad, out1, n, kind, delta);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

// All table pointers are DEVICE pointers (built once per optimizer.attach()).
DCNN_API int dcnn_multi_tensor_update(int mode, float* const* param, float* const* grad, float* const* s1,
                                      float* const* s2, const long long* numel, const int* block_tensor,
                                      const long long* block_off, int num_blocks, float lr, float a, float b, float eps,
                                      float bc1, float bc2, float wd, int zero_grads, const float* hyper,
                                      cudaStream_t stream) {
  if (num_blocks <= 0) return 0;
  MtTable tb{param, grad, s1, s2, numel, block_tensor, block_off};
  switch (mode) {
    case 0: multi_tensor_update_kernel<0><<<num_blocks, 256, 0, stream>>>(tb, lr, a, b, eps, bc1, bc2, wd, zero_grads, hyper); break;
    case 1: multi_tensor_update_kernel<1><<<num_blocks, 256, 0, stream>>>(tb, lr, a, b, eps, bc1, bc2, wd, zero_grads, hyper); break;
    case 2: multi_tensor_update_kernel<2><<<num_blocks, 256, 0, stream>>>(tb, lr, a, b, eps, bc1, bc2, wd, zero_grads, hyper); break;
    case 3: multi_tensor_update_kernel<3><<<num_blocks, 256, 0, stream>>>(tb, lr, a, b, eps, bc1, bc2, wd, zero_grads, hyper); break;
    default: DCNN_REQUIRE(false, "bad optimizer mode %d", mode);
  }
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}

DCNN_API int dcnn_mt_chunk() { return kMtChunk; }

DCNN_API int dcnn_fill(float* p, long long n, float v, cudaStream_t stream) {
  fill_kernel<<<ew_grid(n, 256), 256, 0, stream>>>(p, n, v);
  DCNN_CHECK_LAUNCH();
  DCNN_COUNT_LAUNCH(1);
  return 0;
}
