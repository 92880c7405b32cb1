// Snapshot interoperability harness over the UNMODIFIED reference (tests/test_reference_interop.py):
//
//   ref_snapshot write <model> <prefix> [train steps]
//       builds the zoo model with the reference's own builder (example_models.hpp), optionally takes a few SGD steps on
//       deterministic data (so that weights, BatchNorm parameters and running statistics are not at their initial
//       values), saves <prefix>.json + <prefix>.bin with Sequential::save_to_file, and writes <prefix>.io.bin =
//       { u64 in_shape[4], float in[], u64 out_shape[4], float out_train[], float out_eval[] } where out_train is the
//       training-mode forward (batch statistics) and out_eval the evaluation-mode forward of a fixed input.
//   ref_snapshot train <prefix> <input io.bin> <out prefix> <steps> <sgd|sgdm|adam|adamw> <softmax_ce|logsoftmax_ce>
//       loads <prefix>, takes <steps> optimisation steps on the io file's input with one-hot targets (class of sample
//       n = (3n) mod classes), prints one loss per step (train.hpp's sequence: forward, compute_loss, compute_gradient,
//       backward, update, clear_gradients) and saves the result as <out prefix>.json/.bin.
//   ref_snapshot schedulers <steps> <unused>
//       every learning-rate scheduler of nn/schedulers.hpp on an SGD optimizer with base lr 0.1: prints
//       "<name> lr0 lr1 ... lr<steps>" (ReduceLROnPlateau is fed the metric 1 / (1 + step / 7), floor-quantised to 0.05).
//   ref_snapshot losses <io file with predictions and targets> <prob|logits>
//       value and gradient of every loss of nn/loss.hpp (plain cross-entropy only for probability inputs).
//   ref_snapshot read <prefix> <input io.bin> <output file>
//       loads <prefix>.json/.bin with Sequential::from_file (a snapshot written by dcnn_b200), runs the training-mode
//       forward of the input stored in the io file and writes { u64 out_shape[4], float out[] }.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>
#include <vector>

#include "nn/example_models.hpp"
#include "nn/loss.hpp"
#include "nn/optimizers.hpp"
#include "nn/schedulers.hpp"
#include "nn/sequential.hpp"
#include "partitioner/naive_partitioner.hpp"

using namespace tnn;

static Sequential<float> make(const std::string &name, std::vector<size_t> *shape, size_t *classes) {
  if (name == "mnist_cnn") { *shape = {4, 1, 28, 28}; *classes = 10; return create_mnist_trainer(); }
  if (name == "cifar10_cnn_v1") { *shape = {4, 3, 32, 32}; *classes = 10; return create_cifar10_trainer_v1(); }
  if (name == "resnet9_cifar10") { *shape = {4, 3, 32, 32}; *classes = 10; return create_resnet9_cifar10(); }
  if (name == "resnet18_tiny_imagenet") { *shape = {2, 3, 64, 64}; *classes = 200; return create_resnet18_tiny_imagenet(); }
  // architecture-only entries (partition / init modes): the rest of the reference's zoo
  if (name == "cifar10_cnn_v2") { *shape = {2, 3, 32, 32}; *classes = 10; return create_cifar10_trainer_v2(); }
  if (name == "resnet18_cifar10") { *shape = {2, 3, 32, 32}; *classes = 10; return create_resnet18_cifar10(); }
  if (name == "resnet20_cifar10") { *shape = {2, 3, 32, 32}; *classes = 10; return create_resnet20_cifar10(); }
  if (name == "resnet50_cifar10") { *shape = {2, 3, 32, 32}; *classes = 10; return create_resnet50_cifar10(); }
  if (name == "resnet9_tiny_imagenet") { *shape = {2, 3, 64, 64}; *classes = 200; return create_resnet9_tiny_imagenet(); }
  if (name == "cnn_tiny_imagenet") { *shape = {2, 3, 64, 64}; *classes = 200; return create_cnn_tiny_imagenet(); }
  if (name == "resnet34_tiny_imagenet") { *shape = {2, 3, 64, 64}; *classes = 200; return create_resnet34_tiny_imagenet(); }
  if (name == "resnet50_tiny_imagenet") { *shape = {2, 3, 64, 64}; *classes = 200; return create_resnet50_tiny_imagenet(); }
  if (name == "resnet50_imagenet") { *shape = {1, 3, 224, 224}; *classes = 1000; return create_resnet50_imagenet(); }
  if (name == "all_layers") {   // every layer kind the zoo models do not cover: GroupNorm, average pooling with padding,
    *shape = {3, 3, 12, 12};    // strided / asymmetric convolution, the smooth activations, dropout (identity in eval)
    *classes = 6;
    return SequentialBuilder<float>("all_layers")
        .input({3, 12, 12})
        .conv2d(8, 3, 5, 1, 2, 1, 2, true, "c_asym")
        .groupnorm(4, 1e-5f, true, "gn")
        .activation("elu", "a_elu")
        .avgpool2d(3, 3, 2, 2, 1, 1, "ap_pad")
        .conv2d(12, 1, 1, 1, 1, 0, 0, false, "c_1x1")
        .batchnorm(1e-3f, 0.2f, true, "bn")
        .activation("leaky_relu", "a_leaky")
        .basic_residual_block(12, 12, 1, "rb_id")
        .maxpool2d(2, 2, 1, 1, 0, 0, "mp_overlap")
        .activation("tanh", "a_tanh")
        .flatten("fl")
        .dense(16, true, "d1")
        .activation("sigmoid", "a_sig")
        .dense(6, false, "d2")
        .build();
  }
  std::fprintf(stderr, "unknown model %s\n", name.c_str());
  std::exit(2);
}

static void fill_input(Tensor<float> &x) {
  for (size_t i = 0; i < x.size(); ++i) x.data()[i] = 0.5f + 0.5f * std::sin(0.37f * (float)i) * std::cos(0.011f * (float)i);
}

static void put_shape(std::ofstream &f, const Tensor<float> &t) {
  auto s = t.shape();
  for (int i = 0; i < 4; ++i) {
    uint64_t d = i < (int)s.size() ? s[i] : 1;
    f.write(reinterpret_cast<const char *>(&d), 8);
  }
}

int main(int argc, char **argv) {
  if (argc < 4) return 2;
  const std::string mode = argv[1];
  if (mode == "write") {
    const std::string prefix = argv[3];
    const int steps = argc > 4 ? std::atoi(argv[4]) : 2;
    std::vector<size_t> shape;
    size_t classes = 0;
    auto model = make(argv[2], &shape, &classes);
    model.set_device(DeviceType::CPU);
    model.initialize();
    Tensor<float> x(shape);
    fill_input(x);
    if (steps > 0) {
      auto opt = std::make_unique<SGD<float>>(0.05f, 0.9f);
      auto loss = LossFactory<float>::create_softmax_crossentropy();
      opt->attach(model.parameters(), model.gradients());
      model.set_training(true);
      Tensor<float> t({shape[0], classes, 1, 1});
      t.fill(0.f);
      for (size_t n = 0; n < shape[0]; ++n) t.data()[n * classes + (n * 3) % classes] = 1.f;
      Tensor<float> g;
      for (int i = 0; i < steps; ++i) {
        Tensor<float> y = model.forward(x);
        loss->compute_gradient(y, t, g);
        model.backward(g);
        opt->update();
        opt->clear_gradients();
      }
    }
    model.save_to_file(prefix);
    model.set_training(true);
    Tensor<float> y_train = model.forward(x);
    // NOTE: that forward moved the BatchNorm running statistics once more; they are not part of the reference's
    // snapshot, so the evaluation-mode output is only comparable through the sidecar below.
    model.set_training(false);
    Tensor<float> y_eval = model.forward(x);
    std::ofstream f(prefix + ".io.bin", std::ios::binary);
    put_shape(f, x);
    f.write(reinterpret_cast<const char *>(x.data()), x.size() * 4);
    put_shape(f, y_train);
    f.write(reinterpret_cast<const char *>(y_train.data()), y_train.size() * 4);
    f.write(reinterpret_cast<const char *>(y_eval.data()), y_eval.size() * 4);
    std::printf("wrote %s.{json,bin,io.bin}: %zu outputs\n", prefix.c_str(), y_train.size());
    return 0;
  }
  if (mode == "schedulers") {
    const int steps = std::atoi(argv[2]);
    auto run = [&](const char *name, auto make) {
      SGD<float> opt(0.1f, 0.0f);
      auto sch = make(&opt);
      std::printf("%s %.9g", name, opt.get_learning_rate());
      for (int i = 0; i < steps; ++i) {
        sch->step();
        std::printf(" %.9g", opt.get_learning_rate());
      }
      std::printf("\n");
    };
    run("step", [](Optimizer<float> *o) { return std::make_unique<StepLR<float>>(o, 4, 0.5f); });
    run("multistep", [](Optimizer<float> *o) { return std::make_unique<MultiStepLR<float>>(o, std::vector<size_t>{3, 7, 8}, 0.3f); });
    run("exponential", [](Optimizer<float> *o) { return std::make_unique<ExponentialLR<float>>(o, 0.9f); });
    run("cosine", [](Optimizer<float> *o) { return std::make_unique<CosineAnnealingLR<float>>(o, 10, 0.001f); });
    run("cosine_restarts", [](Optimizer<float> *o) { return std::make_unique<CosineAnnealingWarmRestarts<float>>(o, 3, 2, 0.001f); });
    run("linear_warmup", [](Optimizer<float> *o) { return std::make_unique<LinearWarmup<float>>(o, 5, 0.01f); });
    run("warmup_cosine", [](Optimizer<float> *o) { return std::make_unique<WarmupCosineAnnealing<float>>(o, 4, 16, 0.01f, 0.002f); });
    run("polynomial", [](Optimizer<float> *o) { return std::make_unique<PolynomialLR<float>>(o, 12, 2.0f, 0.005f); });
    run("onecycle", [](Optimizer<float> *o) { return std::make_unique<OneCycleLR<float>>(o, 0.5f, 16, 0.25f, 10.0f, 100.0f); });
    {
      SGD<float> opt(0.1f, 0.0f);
      ReduceLROnPlateau<float> sch(&opt, ReduceLROnPlateau<float>::Mode::MIN, 0.5f, 2, 1e-4f, 0.01f);
      std::printf("plateau %.9g", opt.get_learning_rate());
      for (int i = 0; i < steps; ++i) {
        const float metric = std::floor(20.f / (1.f + (float)i / 7.f)) * 0.05f;
        sch.step(metric);
        std::printf(" %.9g", opt.get_learning_rate());
      }
      std::printf("\n");
    }
    return 0;
  }
  if (mode == "configs") {
    // ref_snapshot configs x x: optimizer configurations as the coordinator ships them inside a StageConfig
    std::printf("optimizer %s\n", SGD<float>(0.05f, 0.9f).get_config().to_json().dump().c_str());
    std::printf("optimizer %s\n", Adam<float>(0.01f, 0.8f, 0.95f, 1e-6f, 1e-3f, false).get_config().to_json().dump().c_str());
    std::printf("optimizer %s\n", Adam<float>(0.02f, 0.9f, 0.999f, 1e-8f, 1e-2f, true).get_config().to_json().dump().c_str());
    return 0;
  }
  if (mode == "evalafter") {
    // ref_snapshot evalafter <prefix> <io> <out> <k>: load a snapshot (fresh BatchNorm running statistics), run k
    // training-mode forwards of the io input, then write the evaluation-mode forward -> running-statistics convention
    auto model = Sequential<float>::from_file(argv[2]);
    model.set_device(DeviceType::CPU);
    std::ifstream in(argv[3], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> x({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(x.data()), x.size() * 4);
    const int k = argc > 5 ? std::atoi(argv[5]) : 3;
    model.set_training(true);
    for (int i = 0; i < k; ++i) model.forward(x);
    model.set_training(false);
    Tensor<float> y = model.forward(x);
    std::ofstream f(argv[4], std::ios::binary);
    put_shape(f, y);
    f.write(reinterpret_cast<const char *>(y.data()), y.size() * 4);
    return 0;
  }
  if (mode == "fwdbwd") {
    // ref_snapshot fwdbwd <pool_tie|dropout> <io> <out>: tiny single-layer models; writes y then dx for dy = 1
    std::ifstream in(argv[3], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> x({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(x.data()), x.size() * 4);
    const std::string which = argv[2];
    SequentialBuilder<float> b(which);
    b.input(std::vector<size_t>{(size_t)s[1], (size_t)s[2], (size_t)s[3]});
    if (which == "pool_tie") b.maxpool2d(2, 2, 2, 2, 0, 0, "mp");
    else if (which == "pool_overlap_tie") b.maxpool2d(3, 3, 1, 1, 1, 1, "mp");
    else b.dropout(0.25f, "do");
    auto model = b.build();
    model.set_device(DeviceType::CPU);
    model.initialize();
    model.set_training(true);
    Tensor<float> y = model.forward(x);
    Tensor<float> g(y.shape());
    g.fill(1.f);
    Tensor<float> dx = model.backward(g);
    std::ofstream f(argv[4], std::ios::binary);
    put_shape(f, y);
    f.write(reinterpret_cast<const char *>(y.data()), y.size() * 4);
    put_shape(f, dx);
    f.write(reinterpret_cast<const char *>(dx.data()), dx.size() * 4);
    return 0;
  }
  if (mode == "init") {
    // ref_snapshot init <model> <unused>: per parameter tensor "param <numel> <min> <max> <mean> <std>" after initialize()
    std::vector<size_t> shape;
    size_t classes = 0;
    auto model = make(argv[2], &shape, &classes);
    model.set_device(DeviceType::CPU);
    model.initialize();
    for (auto *p : model.parameters()) {
      double mn = 1e30, mx = -1e30, sum = 0, sq = 0;
      for (size_t i = 0; i < p->size(); ++i) {
        const double v = p->data()[i];
        mn = std::min(mn, v); mx = std::max(mx, v); sum += v; sq += v * v;
      }
      const double mean = sum / p->size();
      std::printf("param %zu %.9g %.9g %.9g %.9g\n", p->size(), mn, mx, mean, std::sqrt(std::max(0.0, sq / p->size() - mean * mean)));
    }
    return 0;
  }
  if (mode == "partition") {
    // ref_snapshot partition <model> <stages>: NaivePartitioner ranges, then Sequential::split -> one config per line
    std::vector<size_t> shape;
    size_t classes = 0;
    auto model = make(argv[2], &shape, &classes);
    NaivePartitioner<float> part;
    auto parts = part.get_partitions(model.get_layers(), (size_t)std::atoi(argv[3]));
    std::printf("partitions");
    for (auto &p : parts) std::printf(" %zu:%zu", p.start_layer, p.end_layer);
    std::printf("\n");
    auto stages = model.split(parts);
    for (auto &st : stages) std::printf("stage %s\n", st.get_config().dump().c_str());
    std::vector<size_t> in = shape;
    std::printf("forward_complexity");
    for (uint64_t v : model.forward_complexity(in)) std::printf(" %llu", (unsigned long long)v);
    std::printf("\nbackward_complexity");
    for (uint64_t v : model.backward_complexity(in)) std::printf(" %llu", (unsigned long long)v);
    std::printf("\noutput_shape");
    for (size_t d : model.compute_output_shape(in)) std::printf(" %zu", d);
    std::printf("\n");
    return 0;
  }
  if (mode == "losses") {
    // argv[2] = io file {shape, predictions, shape, targets(as "out_train")}; prints "<loss> value g0 g1 ... g(n-1)"
    std::ifstream in(argv[2], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> pred({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(pred.data()), pred.size() * 4);
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> tgt({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(tgt.data()), tgt.size() * 4);
    auto run = [&](const char *name, std::unique_ptr<Loss<float>> loss) {
      float l = 0.f;
      Tensor<float> g;
      loss->compute_loss(pred, tgt, l);
      loss->compute_gradient(pred, tgt, g);
      std::printf("%s %.9g", name, l);
      for (size_t i = 0; i < g.size(); ++i) std::printf(" %.9g", g.data()[i]);
      std::printf("\n");
    };
    run("softmax_ce", LossFactory<float>::create_softmax_crossentropy());
    run("logsoftmax_ce", LossFactory<float>::create_logsoftmax_crossentropy());
    run("mse", LossFactory<float>::create_mse());
    run("mae", LossFactory<float>::create_mae());
    run("huber", LossFactory<float>::create_huber(0.7f));
    if (std::string(argv[3]) == "prob") run("crossentropy", LossFactory<float>::create_crossentropy());
    return 0;
  }
  if (mode == "train") {
    if (argc < 8) return 2;
    auto model = Sequential<float>::from_file(argv[2]);
    model.set_device(DeviceType::CPU);
    std::ifstream in(argv[3], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> x({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(x.data()), x.size() * 4);
    const int steps = std::atoi(argv[5]);
    const std::string on = argv[6], ln = argv[7];
    std::unique_ptr<Optimizer<float>> opt;
    if (on == "sgd") opt = std::make_unique<SGD<float>>(0.05f, 0.0f);
    else if (on == "sgdm") opt = std::make_unique<SGD<float>>(0.05f, 0.9f);
    else if (on == "adam") opt = std::make_unique<Adam<float>>(0.01f, 0.9f, 0.999f, 1e-8f, 1e-3f, false);
    else opt = std::make_unique<Adam<float>>(0.01f, 0.9f, 0.999f, 1e-8f, 1e-2f, true);
    auto loss = ln == "softmax_ce" ? LossFactory<float>::create_softmax_crossentropy()
                                   : LossFactory<float>::create_logsoftmax_crossentropy();
    opt->attach(model.parameters(), model.gradients());
    model.set_training(true);
    Tensor<float> y0 = model.forward(x);
    const size_t classes = y0.size() / s[0];
    Tensor<float> t({(size_t)s[0], classes, 1, 1});
    t.fill(0.f);
    for (size_t n = 0; n < s[0]; ++n) t.data()[n * classes + (n * 3) % classes] = 1.f;
    Tensor<float> g;
    for (int i = 0; i < steps; ++i) {
      Tensor<float> y = model.forward(x);
      float l = 0.f;
      loss->compute_loss(y, t, l);
      loss->compute_gradient(y, t, g);
      model.backward(g);
      opt->update();
      opt->clear_gradients();
      std::printf("loss %.9g\n", l);
    }
    model.save_to_file(argv[4]);
    return 0;
  }
  if (mode == "read") {
    auto model = Sequential<float>::from_file(argv[2]);
    model.set_device(DeviceType::CPU);
    std::ifstream in(argv[3], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> x({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(x.data()), x.size() * 4);
    model.set_training(true);
    Tensor<float> y = model.forward(x);
    std::ofstream f(argv[4], std::ios::binary);
    put_shape(f, y);
    f.write(reinterpret_cast<const char *>(y.data()), y.size() * 4);
    std::printf("read %s: %zu outputs\n", argv[2], y.size());
    return 0;
  }
  return 2;
}
