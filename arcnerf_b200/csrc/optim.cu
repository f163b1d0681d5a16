// Step-level elementwise kernels around the hot path: fused Adam (torch.optim.Adam semantics, with the fp16
// working copy of the parameters written in the same pass and the gradient buffer zeroed for the next step) and
// the image loss used by the NGP config (Huber, delta 0.1, weight 3000: configs/expr/NeRF/lego/
// nerf_lego_nerf_ngp.yaml:192-197, arcnerf/loss/img_loss.py:80-98) with its gradient.
// Both are pure HBM streams: Adam moves 34 B per parameter (p, m, v, g in; p, m, v, g=0, fp16 p out).
#include "common.cuh"
#include "adam.cuh"

namespace {

__global__ void adam_kernel(size_t n, float *__restrict__ p, float *__restrict__ m, float *__restrict__ v,
                            float *__restrict__ g, __half *__restrict__ p_half, AdamHyper h, int zero_grad,
                            const int32_t *__restrict__ step_dev) {
    arc_pdl_sync();
    if (step_dev) bias_from_step(h, max(*step_dev, 1));    // step count kept on the device (replayable launches)
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t n4 = n / 4;
    // p, m, v and g are touched exactly once per step: streaming (evict-first) loads and stores keep them from pushing
    // the fp16 working copy - which the next forward gathers from - out of L2
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += stride) {
        float4 pp = __ldcs(reinterpret_cast<const float4 *>(p) + k), mm = __ldcs(reinterpret_cast<const float4 *>(m) + k),
               vv = __ldcs(reinterpret_cast<const float4 *>(v) + k), gg = __ldcs(reinterpret_cast<const float4 *>(g) + k);
        adam_one(pp.x, mm.x, vv.x, gg.x, h); adam_one(pp.y, mm.y, vv.y, gg.y, h);
        adam_one(pp.z, mm.z, vv.z, gg.z, h); adam_one(pp.w, mm.w, vv.w, gg.w, h);
        __stcs(reinterpret_cast<float4 *>(p) + k, pp);
        __stcs(reinterpret_cast<float4 *>(m) + k, mm);
        __stcs(reinterpret_cast<float4 *>(v) + k, vv);
        if (zero_grad) reinterpret_cast<float4 *>(g)[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p_half) {
            __half2 a = __floats2half2_rn(pp.x, pp.y), b = __floats2half2_rn(pp.z, pp.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t *>(&a);
            pk.y = *reinterpret_cast<uint32_t *>(&b);
            reinterpret_cast<uint2 *>(p_half)[k] = pk;
        }
    }
    for (size_t k = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        float pp = p[k], mm = m[k], vv = v[k];
        adam_one(pp, mm, vv, g[k], h);
        p[k] = pp; m[k] = mm; v[k] = vv;
        if (zero_grad) g[k] = 0.f;
        if (p_half) p_half[k] = __float2half_rn(pp);
    }
}

__global__ void counter_add_kernel(int32_t *c, int32_t v) {
    arc_pdl_sync();
    if (threadIdx.x == 0 && blockIdx.x == 0) *c += v;
}

__global__ void fill_f32_kernel(float *dst, float value, size_t n) {
    arc_pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = value;
}

__global__ void scale_f32_kernel(float *x, float a, size_t n) {
    arc_pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) x[i] *= a;
}

// loss_sum[0] += weight/(3R) * sum huber(rgb - target);  g_rgb = d loss / d rgb
__global__ void huber_kernel(uint32_t n_elem, const float *__restrict__ rgb, const float *__restrict__ target,
                             float delta, float scale, float *__restrict__ g_rgb, float *__restrict__ loss_sum) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float l = 0.f;
    if (i < n_elem) {
        const float d = rgb[i] - target[i];
        const float a = fabsf(d);
        if (a < delta) { l = 0.5f / delta * (a * a); if (g_rgb) g_rgb[i] = scale * d / delta; }
        else { l = a - 0.5f * delta; if (g_rgb) g_rgb[i] = scale * (d > 0.f ? 1.f : -1.f); }
    }
    #pragma unroll
    for (int s = 16; s > 0; s >>= 1) l += __shfl_xor_sync(ARC_FULL_MASK, l, s);
    __shared__ float ws[32];
    const uint32_t lane = threadIdx.x & 31u, w = threadIdx.x >> 5;
    if (lane == 0) ws[w] = l;
    __syncthreads();
    if (w == 0) {
        l = lane < (blockDim.x >> 5) ? ws[lane] : 0.f;
        #pragma unroll
        for (int s = 16; s > 0; s >>= 1) l += __shfl_xor_sync(ARC_FULL_MASK, l, s);
        if (lane == 0 && loss_sum) atomicAdd(loss_sum, l * scale);
    }
}

}  // namespace

extern "C" {

// One Adam step over a flat fp32 parameter vector (torch.optim.Adam arithmetic: L2 weight decay added to the
// gradient, bias correction by step count `step` >= 1).  p_half (optional): fp16 copy of the updated
// parameters, written in the same pass.  zero_grad != 0 clears g after use.  All buffers 16-byte aligned.
// step_dev (device int32, optional) overrides `step`: the bias corrections are then derived on the device, so the
// launch can be replayed (CUDA graph) while a counter kernel advances the step.
// The reference runs torch.optim.Adam (common/trainer/optimizer.py) and re-casts fp32->fp16 at every forward
// (tinycudann/modules.py:156).
int arcb200_adam_step(float *p, float *m, float *v, float *g, void *p_half, int64_t n, int step, float lr,
                      float beta1, float beta2, float eps, float weight_decay, float grad_scale, int zero_grad,
                      const int32_t *step_dev, void *stream) {
    ARC_REQUIRE(n >= 0, "adam_step: negative n");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(p && m && v && g && (step >= 1 || step_dev), "adam_step: null pointer or step < 1");
    if (step < 1) step = 1;
    ARC_REQUIRE((((uintptr_t)p | (uintptr_t)m | (uintptr_t)v | (uintptr_t)g) & 15u) == 0 && ((uintptr_t)p_half & 7u) == 0,
                "adam_step: misaligned buffer");
    const AdamHyper h = make_adam_hyper(lr, beta1, beta2, eps, weight_decay, grad_scale, step);
    long long blocks = (n / 4 + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    arc_launch(adam_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, (size_t)n, p, m, v, g, (__half *)p_half, h, zero_grad, step_dev);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// counter[0] += value on the stream (device-resident step counter for replayed launch sequences)
int arcb200_counter_add(int32_t *counter, int value, void *stream) {
    ARC_REQUIRE(counter != nullptr, "counter_add: null pointer");
    arc_launch(counter_add_kernel, 1, 32, 0, (cudaStream_t)stream, counter, value);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// dst[0..n) = value on the stream (accumulators that a later kernel adds into, e.g. the loss)
int arcb200_fill_f32(float *dst, float value, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && (dst || n == 0), "fill_f32: bad argument");
    if (n == 0) return ARCB200_OK;
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    arc_launch(fill_f32_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, dst, value, (size_t)n);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// x[0..n) *= a on the stream (density -> opacity = density * dt of the occupancy refresh, base_3d_model.py:368-389)
int arcb200_scale_f32(float *x, float a, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && (x || n == 0), "scale_f32: bad argument");
    if (n == 0) return ARCB200_OK;
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    arc_launch(scale_f32_kernel, (unsigned)blocks, 256, 0, (cudaStream_t)stream, x, a, (size_t)n);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// loss_sum[0] += weight * mean(huber_delta(rgb - target)) over n_rays*3 entries; g_rgb (n_rays,3) = its gradient.
int arcb200_huber_loss(const float *rgb, const float *target, float delta, float weight, float *g_rgb,
                       float *loss_sum, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_rays * 3 < (1ll << 32), "huber_loss: n_rays out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(rgb && target && delta > 0.f, "huber_loss: null pointer or non-positive delta");
    const uint32_t n_elem = (uint32_t)(n_rays * 3);
    arc_launch(huber_kernel, arc_blocks(n_elem, 256), 256, 0, (cudaStream_t)stream, n_elem, rgb, target, delta,
                                                                           weight / (float)n_elem, g_rgb, loss_sum);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
