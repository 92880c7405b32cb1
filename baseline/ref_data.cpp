// Data-path interoperability harness over the UNMODIFIED reference's loaders and augmentations
// (tests/test_reference_interop.py):
//
//   ref_data mnist <csv> <out>                     MNISTDataLoader::load_data + prepare_batches(all) -> first batch
//   ref_data cifar10 <bin> <out> | cifar100 <bin> <out>
//   ref_data wifi <csv> <feat_end> <tgt_end> <out> WiFiDataLoader::load_data(+normalize_data) -> first batch
//   ref_data normalize <io file> <out>             NormalizationAugmentation(mean .1 .2 .3, std .5 .6 .7) on a batch
//   ref_data hflip|vflip <io file> <out>           Horizontal / VerticalFlipAugmentation(probability 1) on a batch
//   ref_data accuracy <io file> <unused>           io = predictions, targets -> prints "corrects N accuracy A"
// <out> = { u64 shape[4], float data[], u64 shape[4], float labels[] } (labels omitted for normalize).
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <string>

#include "data_augmentation/augmentation.hpp"
#include "data_loading/cifar100_data_loader.hpp"
#include "data_loading/cifar10_data_loader.hpp"
#include "data_loading/mnist_data_loader.hpp"
#include "data_loading/wifi_data_loader.hpp"
#include "utils/utils_extended.hpp"

using namespace tnn;

static void put(std::ofstream &f, const Tensor<float> &t) {
  auto s = t.shape();
  for (int i = 0; i < 4; ++i) {
    uint64_t d = i < (int)s.size() ? s[i] : 1;
    f.write(reinterpret_cast<const char *>(&d), 8);
  }
  f.write(reinterpret_cast<const char *>(t.data()), t.size() * 4);
}

template <class L> static int image_batch(L &loader, const char *src, const char *out) {
  if (!loader.load_data(src)) return 1;
  loader.prepare_batches(loader.size());
  Tensor<float> x, y;
  if (!loader.get_next_batch(x, y)) return 1;
  std::ofstream f(out, std::ios::binary);
  put(f, x);
  put(f, y);
  std::printf("samples %zu\n", (size_t)loader.size());
  return 0;
}

int main(int argc, char **argv) {
  if (argc < 4) return 2;
  const std::string mode = argv[1];
  if (mode == "mnist") { MNISTDataLoader<float> l; return image_batch(l, argv[2], argv[3]); }
  if (mode == "cifar10") { CIFAR10DataLoader<float> l; return image_batch(l, argv[2], argv[3]); }
  if (mode == "cifar100") { CIFAR100DataLoader<float> l; return image_batch(l, argv[2], argv[3]); }
  if (mode == "cifar100coarse") { CIFAR100DataLoader<float> l(true); return image_batch(l, argv[2], argv[3]); }
  if (mode == "wifi") {
    WiFiDataLoader l(true);
    const size_t fe = std::atoi(argv[3]), te = std::atoi(argv[4]);
    if (!l.load_data(argv[2], 0, fe, fe, te, true)) return 1;
    l.normalize_data();
    l.prepare_batches(l.size());
    Tensor<float> x, y;
    if (!l.get_next_batch(x, y)) return 1;
    std::ofstream f(argv[5], std::ios::binary);
    put(f, x);
    put(f, y);
    return 0;
  }
  if (mode == "hflip" || mode == "vflip") {
    std::ifstream in(argv[2], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> x({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]}), labels({(size_t)s[0], 1, 1, 1});
    in.read(reinterpret_cast<char *>(x.data()), x.size() * 4);
    if (mode == "hflip") {
      HorizontalFlipAugmentation<float> aug(1.0f);
      aug.apply(x, labels);
    } else {
      VerticalFlipAugmentation<float> aug(1.0f);
      aug.apply(x, labels);
    }
    std::ofstream f(argv[3], std::ios::binary);
    put(f, x);
    return 0;
  }
  if (mode == "accuracy") {
    std::ifstream in(argv[2], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> p({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(p.data()), p.size() * 4);
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> t({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]});
    in.read(reinterpret_cast<char *>(t.data()), t.size() * 4);
    std::printf("corrects %d accuracy %.9g\n", compute_class_corrects(p, t), compute_class_accuracy(p, t));
    return 0;
  }
  if (mode == "normalize") {
    std::ifstream in(argv[2], std::ios::binary);
    uint64_t s[4];
    in.read(reinterpret_cast<char *>(s), 32);
    Tensor<float> x({(size_t)s[0], (size_t)s[1], (size_t)s[2], (size_t)s[3]}), labels({(size_t)s[0], 1, 1, 1});
    in.read(reinterpret_cast<char *>(x.data()), x.size() * 4);
    NormalizationAugmentation<float> aug({0.1f, 0.2f, 0.3f}, {0.5f, 0.6f, 0.7f});
    aug.apply(x, labels);
    std::ofstream f(argv[3], std::ios::binary);
    put(f, x);
    return 0;
  }
  return 2;
}
