// Reference arm, BASELINE config 1: the UNMODIFIED reference's MNIST CNN (create_mnist_trainer) trained on the CPU with
// synthetic 28x28 data, batch 64, Adam(0.01, eps 1e-8) + LogSoftmaxCrossEntropy as in examples/mnist_cnn_trainer.cpp,
// per-batch sequence of train_class_epoch (include/nn/train.hpp:108-170).  The offline subset build has no TBB and the
// reference's OpenMP path does not compile with this GCC (schedule(affinity)), so its parallel_for runs serially: this
// is a SINGLE-THREAD number for the reference's hand-written AVX2 kernels.
//
// usage: ref_mnist_cpu <batch> <steps> <warmup>   -> prints one JSON line
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "nn/example_models.hpp"
#include "nn/loss.hpp"
#include "nn/optimizers.hpp"
#include "nn/sequential.hpp"
#include "utils/utils_extended.hpp"

using namespace tnn;

int main(int argc, char **argv) {
  const size_t batch = argc > 1 ? std::atoi(argv[1]) : 64;
  const int steps = argc > 2 ? std::atoi(argv[2]) : 30;
  const int warmup = argc > 3 ? std::atoi(argv[3]) : 5;
  const size_t classes = 10;
  auto model = create_mnist_trainer();
  model.set_device(DeviceType::CPU);
  model.initialize();
  auto optimizer = std::make_unique<Adam<float>>(0.01f, 0.9f, 0.999f, 1e-8f);
  auto loss_function = LossFactory<float>::create_logsoftmax_crossentropy();
  optimizer->attach(model.parameters(), model.gradients());
  model.set_training(true);
  std::mt19937 rng(1234);
  std::uniform_real_distribution<float> ud(0.f, 1.f);
  Tensor<float> x({batch, 1, 28, 28});
  for (size_t i = 0; i < x.size(); ++i) x.data()[i] = ud(rng);
  Tensor<float> t({batch, classes, 1, 1});
  t.fill(0.f);
  for (size_t n = 0; n < batch; ++n) t.data()[n * classes + (rng() % classes)] = 1.f;
  Tensor<float> loss_gradient;
  float last_loss = 0.f;
  auto one_step = [&]() {
    Tensor<float> predictions = model.forward(x);
    float loss;
    loss_function->compute_loss(predictions, t, loss);
    int corrects = compute_class_corrects(predictions, t);
    (void)corrects;
    loss_function->compute_gradient(predictions, t, loss_gradient);
    model.backward(loss_gradient);
    optimizer->update();
    optimizer->clear_gradients();
    last_loss = loss;
  };
  for (int i = 0; i < warmup; ++i) one_step();
  auto w0 = std::chrono::high_resolution_clock::now();
  for (int i = 0; i < steps; ++i) one_step();
  auto w1 = std::chrono::high_resolution_clock::now();
  const double ms = std::chrono::duration<double, std::milli>(w1 - w0).count();
  std::printf("{\"impl\": \"reference\", \"config\": \"MNIST CNN, CPU, batch %zu, serial parallel_for (1 thread)\", "
              "\"value\": %.1f, \"unit\": \"images/s\", \"ms_per_step\": %.3f, \"steps\": %d, \"final_loss\": %.5f}\n",
              batch, batch * steps / (ms / 1000.0), ms / steps, steps, last_loss);
  return 0;
}
