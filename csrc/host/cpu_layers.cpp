// Native CPU layer kernels (NCHW fp32, OpenMP): BatchNorm statistics / apply / backward, max- and average pooling
// forward / backward.  Together with cpu_math.cpp (GEMM-based conv and Dense) they are the counterpart of the
// reference's CPU layer ops (src/nn/layers_impl/cpu/batchnorm_ops.cpp, maxpool_ops.cpp, avgpool_ops.cpp) for the
// DCNN_CPU_BACKEND=native path; semantics follow SURVEY appendix A.3 (biased variance for normalisation, avg-pool
// divisor always kh*kw, max-pool ties -> first maximum in scan order, gradients accumulate).
#include <math.h>
#include <string.h>

#include <vector>

#include "host_common.h"

// per-channel mean and biased variance over (N, H, W); FP64 accumulation, two passes
DCNN_HOST_API void dcnn_cpu_bn_stats(const float* x, int N, int C, int HW, float* mean, float* var) {
#pragma omp parallel for schedule(static)
  for (int c = 0; c < C; ++c) {
    double s = 0;
    for (int n = 0; n < N; ++n) {
      const float* p = x + ((size_t)n * C + c) * HW;
      double t = 0;
#pragma omp simd reduction(+ : t)
      for (int i = 0; i < HW; ++i) t += p[i];
      s += t;
    }
    const double m = s / ((double)N * HW);
    double q = 0;
    for (int n = 0; n < N; ++n) {
      const float* p = x + ((size_t)n * C + c) * HW;
      double t = 0;
#pragma omp simd reduction(+ : t)
      for (int i = 0; i < HW; ++i) t += ((double)p[i] - m) * ((double)p[i] - m);
      q += t;
    }
    mean[c] = (float)m;
    var[c] = (float)(q / ((double)N * HW));
  }
}

// y = x * scale[c] + shift[c] (+ residual) (ReLU)
DCNN_HOST_API void dcnn_cpu_bn_apply(const float* x, const float* scale, const float* shift, const float* residual,
                                     float* y, int N, int C, int HW, int relu) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int c = 0; c < C; ++c) {
      const size_t o = ((size_t)n * C + c) * HW;
      const float sc = scale[c], sh = shift[c];
      const float* r = residual ? residual + o : nullptr;
#pragma omp simd
      for (int i = 0; i < HW; ++i) {
        float v = x[o + i] * sc + sh;
        if (r) v += r[i];
        y[o + i] = (relu && v < 0.f) ? 0.f : v;
      }
    }
  }
}

// Backward of (BN -> [+residual] -> [ReLU]).  g = dout masked by (out > 0) when relu.  Emits the per-channel sums
// (sum_dy, sum_dy_xhat) and, in a second call (phase 1) after the caller merged them across replicas if it needs to,
// dx = (g - sum_dy/m - xhat * sum_dy_xhat/m) * scale.  `masked` (optional) receives g.
DCNN_HOST_API void dcnn_cpu_bn_bwd_sums(const float* dout, const float* x, const float* out, const float* mean,
                                        const float* invstd, int N, int C, int HW, int relu, double* sum_dy,
                                        double* sum_dy_xhat) {
#pragma omp parallel for schedule(static)
  for (int c = 0; c < C; ++c) {
    double a = 0, b = 0;
    const float mu = mean[c], is = invstd[c];
    for (int n = 0; n < N; ++n) {
      const size_t o = ((size_t)n * C + c) * HW;
      // one (n, c) plane at a time in float lanes (a plane is at most a few thousand elements), planes summed in FP64
      for (int i0 = 0; i0 < HW; i0 += 4096) {
        const int len = HW - i0 < 4096 ? HW - i0 : 4096;
        const float* pd = dout + o + i0;
        const float* px = x + o + i0;
        float ta = 0.f, tb = 0.f;
        if (relu) {
          const float* po = out + o + i0;
#pragma omp simd reduction(+ : ta, tb)
          for (int i = 0; i < len; ++i) {
            const float g = po[i] > 0.f ? pd[i] : 0.f;
            ta += g;
            tb += g * ((px[i] - mu) * is);
          }
        } else {
#pragma omp simd reduction(+ : ta, tb)
          for (int i = 0; i < len; ++i) {
            ta += pd[i];
            tb += pd[i] * ((px[i] - mu) * is);
          }
        }
        a += ta;
        b += tb;
      }
    }
    sum_dy[c] = a;
    sum_dy_xhat[c] = b;
  }
}

DCNN_HOST_API void dcnn_cpu_bn_bwd_dx(const float* dout, const float* x, const float* out, const float* mean,
                                      const float* invstd, const float* scale, const double* sum_dy,
                                      const double* sum_dy_xhat, double count, int N, int C, int HW, int relu, float* dx,
                                      float* masked) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int c = 0; c < C; ++c) {
      const size_t o = ((size_t)n * C + c) * HW;
      const float mu = mean[c], is = invstd[c], sc = scale[c];
      const float ma = (float)(sum_dy[c] / count), mb = (float)(sum_dy_xhat[c] / count);
#pragma omp simd
      for (int i = 0; i < HW; ++i) {
        const float g = (relu && !(out[o + i] > 0.f)) ? 0.f : dout[o + i];
        const float xhat = (x[o + i] - mu) * is;
        dx[o + i] = (g - ma - xhat * mb) * sc;
        if (masked) masked[o + i] = g;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------ pooling
namespace {
// One (n, c) plane whose windows all lie inside the image.  KH/KW > 0 fix the window at compile time (the loops unroll
// into compare/select chains without branches); 0 takes the run-time size.  First maximum in scan order wins; a NaN in
// the window makes the output NaN at the first NaN's position (ATen's convention), handled off the fast path.
template <int KH, int KW>
inline void pool_plane_inside(const float* p, float* q, long long* qi, int W, int OH, int OW, int kh_rt, int kw_rt, int sh,
                              int sw) {
  const int kh = KH ? KH : kh_rt, kw = KW ? KW : kw_rt;
  for (int oh = 0; oh < OH; ++oh) {
    for (int ow = 0; ow < OW; ++ow) {
      const float* win = p + (size_t)oh * sh * W + (size_t)ow * sw;
      float best = win[0];
      int off = 0;
      bool nan = false;
      for (int r = 0; r < kh; ++r)
        for (int s = 0; s < kw; ++s) {
          const float v = win[r * W + s];
          const bool take = v > best;
          nan |= v != v;
          best = take ? v : best;
          off = take ? r * W + s : off;
        }
      if (nan) {
        for (int r = 0, found = 0; r < kh && !found; ++r)
          for (int s = 0; s < kw; ++s)
            if (win[r * W + s] != win[r * W + s]) {
              best = win[r * W + s];
              off = r * W + s;
              found = 1;
              break;
            }
      }
      q[oh * OW + ow] = best;
      qi[oh * OW + ow] = (long long)(win - p) + off;
    }
  }
}
}  // namespace

// idx[n,c,oh,ow] = flat input index (ih*W + iw) of the first maximum in scan order (torch.max_pool2d's convention too)
DCNN_HOST_API void dcnn_cpu_maxpool_fwd(const float* x, float* y, long long* idx, int N, int C, int H, int W, int OH,
                                        int OW, int kh, int kw, int sh, int sw, int ph, int pw) {
  // windows that never leave the image (no padding, complete windows): no bounds tests in the inner loops
  const bool inside = ph == 0 && pw == 0 && (OH - 1) * sh + kh <= H && (OW - 1) * sw + kw <= W;
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int c = 0; c < C; ++c) {
      const float* p = x + ((size_t)n * C + c) * H * W;
      float* q = y + ((size_t)n * C + c) * OH * OW;
      long long* qi = idx + ((size_t)n * C + c) * OH * OW;
      if (inside) {
        if (kh == 2 && kw == 2) pool_plane_inside<2, 2>(p, q, qi, W, OH, OW, kh, kw, sh, sw);
        else if (kh == 3 && kw == 3) pool_plane_inside<3, 3>(p, q, qi, W, OH, OW, kh, kw, sh, sw);
        else pool_plane_inside<0, 0>(p, q, qi, W, OH, OW, kh, kw, sh, sw);
        continue;
      }
      for (int oh = 0; oh < OH; ++oh) {
        for (int ow = 0; ow < OW; ++ow) {
          float best = -INFINITY;
          long long bi = -1;
          for (int r = 0; r < kh; ++r) {
            const int ih = oh * sh - ph + r;
            if (ih < 0 || ih >= H) continue;
            for (int s = 0; s < kw; ++s) {
              const int iw = ow * sw - pw + s;
              if (iw < 0 || iw >= W) continue;
              const float v = p[(size_t)ih * W + iw];
              if (bi < 0 || v > best || (v != v && best == best)) {   // NaN propagates like ATen
                best = v;
                bi = (long long)ih * W + iw;
              }
            }
          }
          q[oh * OW + ow] = best;
          qi[oh * OW + ow] = bi;
        }
      }
    }
  }
}

DCNN_HOST_API void dcnn_cpu_maxpool_bwd(const float* dy, const long long* idx, float* dx, int N, int C, int HW,
                                        int OHW) {
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int c = 0; c < C; ++c) {
      float* d = dx + ((size_t)n * C + c) * HW;
      memset(d, 0, sizeof(float) * HW);
      const float* g = dy + ((size_t)n * C + c) * OHW;
      const long long* qi = idx + ((size_t)n * C + c) * OHW;
      for (int i = 0; i < OHW; ++i)
        if (qi[i] >= 0) d[qi[i]] += g[i];
    }
  }
}

// divisor is always kh*kw (zero padding counts)
DCNN_HOST_API void dcnn_cpu_avgpool_fwd(const float* x, float* y, int N, int C, int H, int W, int OH, int OW, int kh,
                                        int kw, int sh, int sw, int ph, int pw) {
  const float inv = 1.f / (float)(kh * kw);
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int c = 0; c < C; ++c) {
      const float* p = x + ((size_t)n * C + c) * H * W;
      float* q = y + ((size_t)n * C + c) * OH * OW;
      for (int oh = 0; oh < OH; ++oh)
        for (int ow = 0; ow < OW; ++ow) {
          float acc = 0.f;
          for (int r = 0; r < kh; ++r) {
            const int ih = oh * sh - ph + r;
            if (ih < 0 || ih >= H) continue;
            for (int s = 0; s < kw; ++s) {
              const int iw = ow * sw - pw + s;
              if (iw >= 0 && iw < W) acc += p[(size_t)ih * W + iw];
            }
          }
          q[oh * OW + ow] = acc * inv;
        }
    }
  }
}

DCNN_HOST_API void dcnn_cpu_avgpool_bwd(const float* dy, float* dx, int N, int C, int H, int W, int OH, int OW, int kh,
                                        int kw, int sh, int sw, int ph, int pw) {
  const float inv = 1.f / (float)(kh * kw);
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; ++n) {
    for (int c = 0; c < C; ++c) {
      float* d = dx + ((size_t)n * C + c) * H * W;
      memset(d, 0, sizeof(float) * H * W);
      const float* g = dy + ((size_t)n * C + c) * OH * OW;
      for (int oh = 0; oh < OH; ++oh)
        for (int ow = 0; ow < OW; ++ow) {
          const float v = g[oh * OW + ow] * inv;
          for (int r = 0; r < kh; ++r) {
            const int ih = oh * sh - ph + r;
            if (ih < 0 || ih >= H) continue;
            for (int s = 0; s < kw; ++s) {
              const int iw = ow * sw - pw + s;
              if (iw >= 0 && iw < W) d[(size_t)ih * W + iw] += v;
            }
          }
        }
    }
  }
}

// ------------------------------------------------------------------------------------------------ optimizers
// One pass per tensor (ref src/nn/optimizers_impl/cpu: sgd / adam kernels; maths as adam_kernels.cu:24-60).
// v = mu*v - lr*g; p += v   (mu == 0: p -= lr*g, `vel` may be NULL)
DCNN_HOST_API void dcnn_cpu_sgd_update(float* p, float* g, float* vel, long long n, float lr, float mu, int zero_grads) {
#pragma omp parallel for simd schedule(static) if (n > 32768)
  for (long long i = 0; i < n; ++i) {
    if (vel != nullptr && mu > 0.f) {
      const float v = mu * vel[i] - lr * g[i];
      vel[i] = v;
      p[i] += v;
    } else {
      p[i] -= lr * g[i];
    }
    if (zero_grads) g[i] = 0.f;
  }
}

// Adam (L2 weight decay folded into the update) / AdamW (decoupled), bias corrections passed in
DCNN_HOST_API void dcnn_cpu_adam_update(float* p, float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                                        float eps, float bc1, float bc2, float wd, int decoupled, int zero_grads) {
#pragma omp parallel for simd schedule(static) if (n > 32768)
  for (long long i = 0; i < n; ++i) {
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    float pi = p[i];
    float upd = lr * (mi / bc1) / (sqrtf(vi / bc2) + eps);
    if (wd > 0.f) {
      if (decoupled) pi -= wd * lr * pi;
      else upd += wd * lr * pi;
    }
    p[i] = pi - upd;
    if (zero_grads) g[i] = 0.f;
  }
}

// ------------------------------------------------------------------------------------------------ softmax cross-entropy
// pred / target: [B][C] (targets one-hot).  out[0] = mean loss, out[1] = number of rows whose argmax matches the target's.
// grad (optional) = (softmax(pred) - target) / B.
DCNN_HOST_API void dcnn_cpu_softmax_ce(const float* pred, const float* target, float* grad, double* out, int B, int C) {
  double loss = 0, correct = 0;
#pragma omp parallel for schedule(static) reduction(+ : loss, correct) if ((long long)B * C > 16384)
  for (int b = 0; b < B; ++b) {
    const float* p = pred + (size_t)b * C;
    const float* t = target + (size_t)b * C;
    int pi = 0, ti = 0;
    for (int c = 1; c < C; ++c) {
      if (p[c] > p[pi]) pi = c;
      if (t[c] > t[ti]) ti = c;
    }
    double s = 0;
    for (int c = 0; c < C; ++c) s += exp((double)p[c] - (double)p[pi]);
    loss += log(s) + (double)p[pi] - (double)p[ti];
    correct += pi == ti ? 1.0 : 0.0;
    if (grad != nullptr) {
      float* g = grad + (size_t)b * C;
      const double inv = 1.0 / s;
      for (int c = 0; c < C; ++c) g[c] = (float)((exp((double)p[c] - (double)p[pi]) * inv - (double)t[c]) / B);
    }
  }
  out[0] = loss / B;
  out[1] = correct;
}
