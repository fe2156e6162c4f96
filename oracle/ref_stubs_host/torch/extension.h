// Stand-in for <torch/extension.h> that lets the reference's iou3d_nms.cpp (the host side of pcdet_nms: cudaMalloc, kernel
// launcher, blocking copy, greedy host loop, cudaFree) compile with nvcc's host compiler and RUN without linking torch:
// a Tensor here is a raw pointer + two sizes + a device flag, i.e. everything that file reads from an at::Tensor.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
namespace at {
struct TensorType {
  bool cuda;
  bool is_cuda() const { return cuda; }
};
struct Tensor {
  void* ptr;
  long n0, n1;
  bool cuda;
  bool is_contiguous() const { return true; }
  long size(int d) const { return d == 0 ? n0 : n1; }
  TensorType type() const { return TensorType{cuda}; }
  template <typename T> T* data() const { return static_cast<T*>(ptr); }
};
}  // namespace at
