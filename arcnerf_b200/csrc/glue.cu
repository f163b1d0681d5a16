// Elementwise glue between the geometry head, the density activation and the compositor:
//   * TruncExp (arcnerf/ops/trunc_exp.py:7-37): forward exp(x), backward g * exp(clamp(x, -clip, clip));
//   * the gradient that enters the geometry head, assembled from the density gradient (through TruncExp, channel
//     0) and the feature gradient coming back from the radiance head (FusedMLPGeoNet.handle_output_combine,
//     tcnn_fusedmlp_module.py:99-111: sigma = act(out[:,0]), feature = all W_feat outputs).
// Strided, type-generic loads so that they can read a column of a wider fp16 activation row in place.
#include "common.cuh"

namespace {

unsigned stride_blocks(long long n, int threads) {
    long long want = (n + threads - 1) / threads;
    const long long cap = (long long)arc_sm_count() * 8;
    return (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
}

template <typename T> __device__ __forceinline__ float ldv(const T *p);
template <> __device__ __forceinline__ float ldv<float>(const float *p) { return *p; }
template <> __device__ __forceinline__ float ldv<__half>(const __half *p) { return __half2float(*p); }

template <typename T>
__global__ void trunc_exp_fwd_kernel(int64_t n, const int32_t *__restrict__ n_dev, const T *__restrict__ x, int64_t ldx,
                                     float *__restrict__ y) {
    arc_pdl_sync();
    if (n_dev) { const int64_t nd = *n_dev; n = nd < n ? (nd > 0 ? nd : 0) : n; }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = expf(ldv<T>(x + i * ldx));
}

template <typename T>
__global__ void trunc_exp_bwd_kernel(int64_t n, const int32_t *__restrict__ n_dev, const T *__restrict__ x, int64_t ldx,
                                     const float *__restrict__ g, float clip, float *__restrict__ gx) {
    arc_pdl_sync();
    if (n_dev) { const int64_t nd = *n_dev; n = nd < n ? (nd > 0 ? nd : 0) : n; }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        gx[i] = g[i] * expf(fminf(fmaxf(ldv<T>(x + i * ldx), -clip), clip));
}

// dY (n, width) fp32 = d_feat (n, width; leading dim ldf; type TF; NULL = zeros); dY[:,0] += g_sigma * exp(clamp(x0))
template <typename TX, typename TF>
__global__ void geo_head_grad_kernel(int64_t n, const int32_t *__restrict__ n_dev, int width, const TX *__restrict__ x0,
                                     int64_t ldx, const float *__restrict__ g_sigma, float clip,
                                     const TF *__restrict__ d_feat, int64_t ldf, float *__restrict__ dY) {
    arc_pdl_sync();
    if (n_dev) { const int64_t nd = *n_dev; n = nd < n ? (nd > 0 ? nd : 0) : n; }
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n * width; k += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = k / width;
        const int c = (int)(k - i * width);
        float v = d_feat ? ldv<TF>(d_feat + i * ldf + c) : 0.0f;
        if (c == 0) v += g_sigma[i] * expf(fminf(fmaxf(ldv<TX>(x0 + i * ldx), -clip), clip));
        dY[k] = v;
    }
}

// world position -> unit-cube coordinate of the hash grid: (p - min) / (max - min) per axis
__global__ void world_to_unit_kernel(size_t n3, const float *__restrict__ pts, const float *__restrict__ aabb, float *__restrict__ unit) {
    arc_pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n3; i += (size_t)gridDim.x * blockDim.x) {
        const int a = (int)(i % 3);
        unit[i] = (pts[i] - aabb[a]) / (aabb[3 + a] - aabb[a]);
    }
}

}  // namespace

extern "C" {

// y[i] = exp(x[i*ldx]); x fp16 (x_is_half) or fp32.  Replaces TruncExpOps.forward (arcnerf/ops/trunc_exp.py:10-24).
int arcb200_trunc_exp_fwd(const void *x, int x_is_half, int64_t ldx, float *y, int64_t n, const int32_t *n_dev,
                          void *stream) {
    ARC_REQUIRE(n >= 0, "trunc_exp_fwd: negative n");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(x && y && ldx >= 1, "trunc_exp_fwd: null pointer / bad stride");
    if (x_is_half) arc_launch(trunc_exp_fwd_kernel<__half>, stride_blocks(n, 256), 256, 0, (cudaStream_t)stream, n, n_dev, (const __half *)x, ldx, y);
    else arc_launch(trunc_exp_fwd_kernel<float>, stride_blocks(n, 256), 256, 0, (cudaStream_t)stream, n, n_dev, (const float *)x, ldx, y);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// gx[i] = g[i] * exp(clamp(x[i*ldx], -clip, clip)).  Replaces TruncExpOps.backward (trunc_exp.py:26-37).
int arcb200_trunc_exp_bwd(const void *x, int x_is_half, int64_t ldx, const float *g, float clip, float *gx, int64_t n,
                          const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(n >= 0, "trunc_exp_bwd: negative n");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(x && g && gx && ldx >= 1, "trunc_exp_bwd: null pointer / bad stride");
    if (x_is_half) arc_launch(trunc_exp_bwd_kernel<__half>, stride_blocks(n, 256), 256, 0, (cudaStream_t)stream, n, n_dev, (const __half *)x, ldx, g, clip, gx);
    else arc_launch(trunc_exp_bwd_kernel<float>, stride_blocks(n, 256), 256, 0, (cudaStream_t)stream, n, n_dev, (const float *)x, ldx, g, clip, gx);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Gradient w.r.t. the geometry head's output row (n, width) fp32: the feature gradient d_feat (fp16/fp32, optional)
// plus, on channel 0, the density gradient pulled back through TruncExp.  x0 points at channel 0 of the head's
// stored output (stride ldx).  (handle_output_combine, tcnn_fusedmlp_module.py:99-111, differentiated.)
int arcb200_geo_head_grad(const void *x0, int x_is_half, int64_t ldx, const float *g_sigma, float clip,
                          const void *d_feat, int f_is_half, int64_t ldf, int width, float *dY, int64_t n,
                          const int32_t *n_dev, void *stream) {
    ARC_REQUIRE(n >= 0 && width >= 1, "geo_head_grad: bad sizes");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(x0 && g_sigma && dY, "geo_head_grad: null pointer");
    const unsigned blocks = stride_blocks(n * width, 256);
    cudaStream_t st = (cudaStream_t)stream;
#define GO(TX, TF) arc_launch(geo_head_grad_kernel<TX, TF>, blocks, 256, 0, st, n, n_dev, width, (const TX *)x0, ldx, g_sigma, clip, (const TF *)d_feat, ldf, dY)
    if (x_is_half) { if (f_is_half) GO(__half, __half); else GO(__half, float); }
    else { if (f_is_half) GO(float, __half); else GO(float, float); }
#undef GO
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// unit (n,3) = (pts - aabb_min) / (aabb_max - aabb_min); aabb (2,3) = min; max.  The normalisation the reference does
// inside HashGridEmbedder.forward before calling tcnn (hashgrid_encoder.py:150-160); out may alias pts.
int arcb200_world_to_unit(const float *pts, const float *aabb, float *unit, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0, "world_to_unit: negative n");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(pts && aabb && unit, "world_to_unit: null pointer");
    arc_launch(world_to_unit_kernel, stride_blocks(n * 3, 256), 256, 0, (cudaStream_t)stream, (size_t)n * 3, pts, aabb, unit);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
