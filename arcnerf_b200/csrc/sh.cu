// Real spherical-harmonics encoding of view directions, degree 1..5 (1, 4, 9, 16 or 25 outputs).
//
// Semantics: tiny-cuda-nn's SphericalHarmonics encoding as ArcNerf's SHEmbedder(backend='tcnn') drives it
// (arcnerf/models/base_modules/encoding/sh_encoder.py:59-65,116-118 -> tcnn encodings/spherical_harmonics.h:46-152):
// the embedder maps a direction d to (d+1)/2, the encoding maps it back with 2x-1 and evaluates the real SH
// basis (Sloan, "Stupid Spherical Harmonics Tricks", appendix A) in Cartesian form.  With remap != 0 this
// kernel takes the raw direction and applies both maps (same roundings); with remap == 0 the input is
// already in [0,1]^3 (the encoding's own convention).  Streaming kernel: 12 B in, 2-4 B per output out.
#include "common.cuh"
#include "sh.cuh"

namespace {

template <typename OUT> __device__ __forceinline__ OUT cvt(float v);
template <> __device__ __forceinline__ float cvt<float>(float v) { return v; }
template <> __device__ __forceinline__ __half cvt<__half>(float v) { return __float2half_rn(v); }

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
        sh_input(x, y, z, remap);
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
