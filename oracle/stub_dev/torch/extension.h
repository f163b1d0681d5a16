// TEST INFRASTRUCTURE ONLY. Minimal stand-in for <torch/extension.h> used to compile the REFERENCE's own
// arcnerf/ops CUDA kernels (which touch tensors only through data_ptr() and size()) into a plain shared
// library without libtorch: a tensor here is a device pointer plus up to four extents.
#pragma once
#include <cstdint>
namespace torch {
struct Tensor {
    void *p = nullptr;
    int64_t s[4] = {0, 0, 0, 0};
    Tensor() {}
    Tensor(const void *ptr, int64_t s0, int64_t s1 = 1, int64_t s2 = 1, int64_t s3 = 1) : p(const_cast<void *>(ptr)), s{s0, s1, s2, s3} {}
    void *data_ptr() const { return p; }
    int64_t size(int i) const { return s[i]; }
};
}  // namespace torch
