// Reference arm harness: drives the UNMODIFIED reference library (tungphambasement/DCNN, copied verbatim under
// baseline/_ref/src) through its own public C++ API -- model zoo (create_resnet50_tiny_imagenet), Sequential
// forward/backward, LogSoftmaxCrossEntropyLoss, compute_class_corrects, Adam -- with exactly the per-batch
// sequence of train_class_epoch (include/nn/train.hpp:108-170) and the optimizer settings of
// examples/tiny_imagenet_resnet50.cpp:90, on synthetic data.  (include/nn/train.hpp itself cannot be included
// offline: it needs TBB headers, which are not on this machine; the loop body below restates it call for call.)
//
// usage: ref_bench <batch> <steps> <warmup> [model]   -> prints one JSON line
// model (default resnet50_tiny_imagenet, the headline): resnet9_cifar10 | resnet18_tiny_imagenet |
// resnet34_tiny_imagenet | resnet50_tiny_imagenet, each with the optimizer of the reference's example for it.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "nn/example_models.hpp"
#include "nn/loss.hpp"
#include "nn/optimizers.hpp"
#include "nn/sequential.hpp"
#include "utils/utils_extended.hpp"

using namespace tnn;

int main(int argc, char **argv) {
  const size_t batch = argc > 1 ? std::atoi(argv[1]) : 256;
  const int steps = argc > 2 ? std::atoi(argv[2]) : 20;
  const int warmup = argc > 3 ? std::atoi(argv[3]) : 5;
  const std::string which = argc > 4 ? argv[4] : "resnet50_tiny_imagenet";
  size_t classes = 200, side = 64;
  const char *model_label = "ResNet-50-Tiny-ImageNet", *opt_label = "Adam(lr 1e-4, eps 1e-3, wd 1e-3)";
  Sequential<float> model = create_resnet50_tiny_imagenet();
  std::unique_ptr<Adam<float>> optimizer;
  if (which == "resnet50_tiny_imagenet") {
    optimizer = std::make_unique<Adam<float>>(1e-4f, 0.9f, 0.999f, 1e-3f, 1e-3f, false);   // tiny_imagenet_resnet50.cpp:90
  } else if (which == "resnet18_tiny_imagenet" || which == "resnet34_tiny_imagenet") {
    model = which == "resnet18_tiny_imagenet" ? create_resnet18_tiny_imagenet() : create_resnet34_tiny_imagenet();
    model_label = which == "resnet18_tiny_imagenet" ? "resnet18_tiny_imagenet" : "resnet34_tiny_imagenet";
    opt_label = "Adam(lr 1e-4, eps 1e-7)";
    optimizer = std::make_unique<Adam<float>>(1e-4f, 0.9f, 0.999f, 1e-7f);
  } else if (which == "resnet9_cifar10") {
    model = create_resnet9_cifar10();
    classes = 10;
    side = 32;
    model_label = "resnet9_cifar10";
    opt_label = "Adam(lr 1e-3, eps 1e-8, wd 5e-4)";
    optimizer = std::make_unique<Adam<float>>(1e-3f, 0.9f, 0.999f, 1e-8f, 5e-4f, false);   // cifar10_resnet9.cpp
  } else {
    std::fprintf(stderr, "unknown model %s\n", which.c_str());
    return 2;
  }
  const bool headline = which == "resnet50_tiny_imagenet";
  model.set_device(DeviceType::GPU);
  model.initialize();
  auto loss_function = LossFactory<float>::create_logsoftmax_crossentropy();
  optimizer->attach(model.parameters(), model.gradients());
  model.set_training(true);
  const Device *model_device = model.get_device();

  // synthetic batches on the host (the reference loaders hand out CPU tensors)
  const int nbuf = 4;
  std::mt19937 rng(1234);
  std::normal_distribution<float> nd(0.f, 1.f);
  std::vector<Tensor<float>> xs, ts;
  for (int b = 0; b < nbuf; ++b) {
    Tensor<float> x({batch, 3, side, side});
    for (size_t i = 0; i < x.size(); ++i) x.data()[i] = nd(rng);
    Tensor<float> t({batch, classes, 1, 1});
    t.fill(0.f);
    for (size_t n = 0; n < batch; ++n) t.data()[n * classes + (rng() % classes)] = 1.f;
    xs.push_back(std::move(x));
    ts.push_back(std::move(t));
  }

  Tensor<float> device_batch_data(model_device), device_batch_labels(model_device), loss_gradient(model_device);
  float last_loss = 0.f;
  auto one_step = [&](int i) {
    device_batch_data = xs[i % nbuf].to_device(model_device);
    Tensor<float> predictions = model.forward(device_batch_data);
    device_batch_labels = ts[i % nbuf].to_device(model_device);
    float loss;
    loss_function->compute_loss(predictions, device_batch_labels, loss);
    int corrects = compute_class_corrects(predictions, device_batch_labels);
    (void)corrects;
    loss_function->compute_gradient(predictions, device_batch_labels, loss_gradient);
    model.backward(loss_gradient);
    optimizer->update();
    optimizer->clear_gradients();
    last_loss = loss;
  };

  for (int i = 0; i < warmup; ++i) one_step(i);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0, 0);
  auto w0 = std::chrono::high_resolution_clock::now();
  for (int i = 0; i < steps; ++i) one_step(i);
  cudaEventRecord(e1, 0);
  cudaDeviceSynchronize();
  auto w1 = std::chrono::high_resolution_clock::now();
  float ms_ev = 0.f;
  cudaEventElapsedTime(&ms_ev, e0, e1);
  const double ms_wall = std::chrono::duration<double, std::milli>(w1 - w0).count();
  // the reference's step already copies each batch host->device and reads loss/accuracy back every step, so its
  // kernel-only and end-to-end numbers are the same measurement
  const double ms = ms_wall > ms_ev ? ms_wall : ms_ev;
  const double ips = batch * steps / (ms / 1000.0);
  char metric[160];
  if (headline) std::snprintf(metric, sizeof(metric), "train images/sec ResNet-50 64x64 (Tiny-ImageNet shape), max over ranks");
  else std::snprintf(metric, sizeof(metric), "train images/sec %s %zux%zu, max over ranks", model_label, side, side);
  std::printf(
      "{\"metric\": \"%s\", \"value\": %.2f, "
      "\"unit\": \"images/s\", \"n_gpus\": 1, \"steps\": %d, \"warmup\": %d, \"ms_per_step\": %.4f, "
      "\"higher_is_better\": true, \"scaling\": \"weak\", \"vs_baseline\": null, \"dtype\": \"fp32 (cuDNN default "
      "math: TF32 tensor cores allowed)\", \"data\": \"synthetic\", \"impl\": \"reference\", \"config\": {\"model\": "
      "\"%s\", \"global_batch\": %zu, \"per_gpu_batch\": %zu, \"image\": \"3x%zux%zu\", "
      "\"classes\": %zu, \"parallelism\": \"single\", \"optimizer\": \"%s\", "
      "\"loss\": \"LogSoftmaxCrossEntropy\"}, \"e2e\": {\"value\": %.2f, \"unit\": \"images/s\", "
      "\"h2d_bytes_per_step\": %zu, \"d2h_bytes_per_step\": 8}, \"ms_cuda_events\": %.3f, \"ms_wall\": %.3f, "
      "\"final_loss\": %.5f}\n",
      metric, ips, steps, warmup, ms / steps, model_label, batch, batch, side, side, classes, opt_label, ips,
      batch * (3 * side * side + classes) * 4, ms_ev, ms_wall, last_loss);
  return 0;
}
