// Adam update arithmetic shared by the single-GPU kernel (optim.cu) and the peer-memory kernel (peer.cu).
// torch.optim.Adam semantics (the reference's optimiser, common/trainer/optimizer.py).
#pragma once
#include "common.cuh"
#include <math.h>

namespace {

struct AdamHyper {
    float lr, beta1, beta2, eps, weight_decay;
    float bias1, bias2_sqrt;   // 1 - beta1^t, sqrt(1 - beta2^t)
    float grad_scale;          // multiplies the incoming gradient (e.g. 1/world_size after an all-reduce sum)
};

__device__ __forceinline__ void bias_from_step(AdamHyper &h, int step) {
    h.bias1 = (float)(1.0 - pow((double)h.beta1, (double)step));
    h.bias2_sqrt = (float)sqrt(1.0 - pow((double)h.beta2, (double)step));
}

// Written with explicit fused multiply-adds so that every kernel that includes this header rounds identically
// (left to the compiler, which product of `a*b + c*d` gets contracted differs from one kernel to the next).
__device__ __forceinline__ void adam_one(float &p, float &m, float &v, float g, const AdamHyper &h) {
    g = __fmaf_rn(h.weight_decay, p, __fmul_rn(g, h.grad_scale));
    m = __fmaf_rn(h.beta1, m, __fmul_rn(1.0f - h.beta1, g));
    v = __fmaf_rn(h.beta2, v, __fmul_rn(__fmul_rn(1.0f - h.beta2, g), g));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), h.bias2_sqrt), h.eps);
    p = __fmaf_rn(-__fdiv_rn(h.lr, h.bias1), __fdiv_rn(m, denom), p);
}

inline AdamHyper make_adam_hyper(float lr, float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                                 int step) {
    AdamHyper h;
    h.lr = lr; h.beta1 = beta1; h.beta2 = beta2; h.eps = eps; h.weight_decay = weight_decay;
    h.bias1 = (float)(1.0 - pow((double)beta1, (double)step));
    h.bias2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    h.grad_scale = grad_scale;
    return h;
}

}  // namespace
