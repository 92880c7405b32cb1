// libdcnn_host.so: native CPU-side runtime pieces (data loading, augmentation, wire/record codecs).
// The reference implements these in C++ (include/data_loading, include/data_augmentation, include/pipeline/
// binary_serializer.hpp + tbuffer.hpp, include/tensor/tensor.hpp save/load); so do we, behind a C ABI.
#pragma once
#include <stddef.h>
#include <stdint.h>
#define DCNN_HOST_API extern "C" __attribute__((visibility("default")))
