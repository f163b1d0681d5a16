// Real spherical-harmonics encoding of view directions, degree 1..5 (1, 4, 9, 16 or 25 outputs).
//
// Semantics: tiny-cuda-nn's SphericalHarmonics encoding as ArcNerf's SHEmbedder(backend='tcnn') drives it
// (arcnerf/models/base_modules/encoding/sh_encoder.py:59-65,116-118 -> tcnn encodings/spherical_harmonics.h:46-152):
// the embedder maps a direction d to (d+1)/2, the encoding maps it back with 2x-1 and evaluates the real SH
// basis (Sloan, "Stupid Spherical Harmonics Tricks", appendix A) in Cartesian form.  With remap != 0 this
// kernel takes the raw direction and applies both maps (same roundings); with remap == 0 the input is
// already in [0,1]^3 (the encoding's own convention).  Streaming kernel: 12 B in, 2-4 B per output out.
#include "common.cuh"

namespace {

template <typename OUT> __device__ __forceinline__ OUT cvt(float v);
template <> __device__ __forceinline__ float cvt<float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt<__half>(float v) { return __float2half_rn(v); }

__device__ __forceinline__ void sh_basis(float x, float y, float z, int degree, float *o) {
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    o[0] = 0.28209479177387814f;
    if (degree <= 1) return;
    o[1] = -0.48860251190291987f * y;
    o[2] = 0.48860251190291987f * z;
    o[3] = -0.48860251190291987f * x;
    if (degree <= 2) return;
    o[4] = 1.0925484305920792f * xy;
    o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz;
    o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    if (degree <= 3) return;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
    o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
    o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
    o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
    if (degree <= 4) return;
    const float x4 = x2 * x2, y4 = y2 * y2, z4 = z2 * z2;
    o[16] = 2.5033429417967046f * xy * (x2 - y2);
    o[17] = 1.7701307697799304f * yz * (-3.0f * x2 + y2);
    o[18] = 0.94617469575756008f * xy * (7.0f * z2 - 1.0f);
    o[19] = 0.66904654355728921f * yz * (3.0f - 7.0f * z2);
    o[20] = -3.1735664074561294f * z2 + 3.7024941420321507f * z4 + 0.31735664074561293f;
    o[21] = 0.66904654355728921f * xz * (3.0f - 7.0f * z2);
    o[22] = 0.47308734787878004f * (x2 - y2) * (7.0f * z2 - 1.0f);
    o[23] = 1.7701307697799304f * xz * (-x2 + 3.0f * y2);
    o[24] = -3.7550144126950569f * x2 * y2 + 0.62583573544917614f * x4 + 0.62583573544917614f * y4;
}

// One thread per direction (grid-stride); dirs may be indexed through `index` (e.g. the ray id of a packed sample,
// so that per-ray directions need not be repeated per sample as fg_model.py:289 does).  fp16 rows whose start is
// 16-byte aligned are written with 16-byte stores.
template <typename OUT>
__global__ void sh_fwd_kernel(uint32_t n, const int32_t *__restrict__ n_dev, const float *__restrict__ dirs,
                              const int32_t *__restrict__ index, int degree, int remap, OUT *__restrict__ out,
                              uint32_t out_stride, int vec) {
    arc_pdl_sync();
    if (n_dev) n = min(n, (uint32_t)max(*n_dev, 0));
    const int nd = degree * degree;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const size_t src = index ? (size_t)index[i] : (size_t)i;
        float x = __ldg(dirs + 3 * src), y = __ldg(dirs + 3 * src + 1), z = __ldg(dirs + 3 * src + 2);
        if (remap) { x = (x + 1.0f) / 2.0f; y = (y + 1.0f) / 2.0f; z = (z + 1.0f) / 2.0f; }
        x = fmaf(x, 2.f, -1.f); y = fmaf(y, 2.f, -1.f); z = fmaf(z, 2.f, -1.f);
        float o[32];
        sh_basis(x, y, z, degree, o);
        OUT *dst = out + (size_t)i * out_stride;
        if (sizeof(OUT) == 2 && vec && nd == 16) {
            uint4 w[2];
            uint32_t *pw = reinterpret_cast<uint32_t *>(w);
            #pragma unroll
            for (int k = 0; k < 8; ++k) {
                __half2 h = __floats2half2_rn(o[2 * k], o[2 * k + 1]);
                pw[k] = *reinterpret_cast<uint32_t *>(&h);
            }
            reinterpret_cast<uint4 *>(dst)[0] = w[0];
            reinterpret_cast<uint4 *>(dst)[1] = w[1];
        } else {
            #pragma unroll
            for (int k = 0; k < 25; ++k) if (k < nd) dst[k] = cvt<OUT>(o[k]);
        }
    }
}

}  // namespace

extern "C" {

// dirs (n_dirs,3) fp32; index (n) int32 or NULL (then n_dirs == n); out: n rows of `out_stride` elements
// (fp16 when out_is_half != 0, else fp32), the first degree^2 of each row are written.
// Replaces tcnn Module.fwd for a SphericalHarmonics encoding (bindings.cpp:78-109 -> kernel_sh,
// spherical_harmonics.h:46-152) as called from SHEmbedder.forward (sh_encoder.py:104-123).
int arcb200_sh_fwd(const float *dirs, const int32_t *index, int degree, int remap, void *out, int out_is_half,
                   int64_t out_stride, int64_t n, const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(n >= 0 && n < (1ll << 32), "sh_fwd: n out of range");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(dirs && out, "sh_fwd: null pointer");
    ARC_REQUIRE(degree >= 1 && degree <= 5, "sh_fwd: degree must be 1..5");
    ARC_REQUIRE(out_stride >= degree * degree, "sh_fwd: out_stride smaller than degree^2");
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    const int vec = out_is_half && (((uintptr_t)out & 15u) == 0) && ((out_stride * 2) % 16 == 0);
    if (out_is_half)
        arc_launch(sh_fwd_kernel<__half>, (unsigned)blocks, 256, 0, (cudaStream_t)stream, (uint32_t)n, n_dev, dirs, index, degree, remap,
                                                                                (__half *)out, (uint32_t)out_stride, vec);
    else
        arc_launch(sh_fwd_kernel<float>, (unsigned)blocks, 256, 0, (cudaStream_t)stream, (uint32_t)n, n_dev, dirs, index, degree, remap,
                                                                               (float *)out, (uint32_t)out_stride, 0);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
