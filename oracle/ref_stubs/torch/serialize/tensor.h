// Minimal stand-in for <torch/extension.h> so that the reference's iou3d_cpu.cpp compiles with plain g++
// (the oracle only calls its inline iou_bev(); at::Tensor is never instantiated with real data).
#pragma once
#include <stdlib.h>
namespace at {
struct TensorType { bool is_cuda() const { return false; } };
struct Tensor {
  bool is_contiguous() const { return true; }
  long size(int) const { return 0; }
  TensorType type() const { return TensorType(); }
  template <typename T> T* data() const { return nullptr; }
};
}  // namespace at
