// Depth-sample generation around the march: near/far from a bounding sphere, uniform / inverse /
// fixed-step depth samples with interval jitter, sample points, and the sort that merges hierarchical samples.
// Reference: arcnerf/geometry/ray.py:11-30 (get_ray_points_by_zvals), :180-255 (sphere_ray_intersection),
// arcnerf/render/ray_helper.py:231-263 (get_zvals_from_near_far), :266-315 (get_zvals_from_near_far_fix_step),
// :318-358 (get_zvals_outside_sphere / get_zvals_from_sphere_radius), :361-407 (perturb_interval[_with_mask]),
// arcnerf/models/nerf_model.py:93-117 (upsample_zvals: sample_pdf + torch.sort of the concatenation).
// Compiled without FMA contraction so that every value is the fp32 expression tree of the reference's eager ops.
// The reference draws its jitter with torch.rand; here the caller passes the uniforms (any generator), which is
// what makes the kernels checkable against the reference formulae.
#include "common.cuh"
#include <math.h>

namespace {

__device__ __forceinline__ float zero_small(float v) { return fabsf(v) < 1e-5f ? 0.0f : v; }   // set_tensor_to_zeros

// one thread per (ray, sphere)
__global__ void sphere_intersect_kernel(uint32_t n, uint32_t n_r, const float *__restrict__ rays_o,
                                        const float *__restrict__ rays_d, const float *__restrict__ radius, float cx,
                                        float cy, float cz, float *__restrict__ near, float *__restrict__ far,
                                        float *__restrict__ pts, uint8_t *__restrict__ mask) {
    arc_pdl_sync();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * n_r) return;
    const uint32_t ray = i / n_r, s = i - ray * n_r;
    const float ox = rays_o[3 * ray], oy = rays_o[3 * ray + 1], oz = rays_o[3 * ray + 2];
    const float dx = rays_d[3 * ray], dy = rays_d[3 * ray + 1], dz = rays_d[3 * ray + 2];
    const float r = radius[s];
    const float ocx = cx - ox, ocy = cy - oy, ocz = cz - oz;
    float z_half = zero_small(ocx * dx + ocy * dy + ocz * dz);
    const float oc2 = ocx * ocx + ocy * ocy + ocz * ocz;
    const bool inside = sqrtf(oc2) <= r;
    bool ok = (z_half > 0.0f) || inside;
    const float d2 = zero_small(oc2 - z_half * z_half);
    ok = ok && (d2 >= 0.0f);
    float z_off = zero_small(r * r - d2);
    ok = ok && (z_off >= 0.0f);
    z_off = sqrtf(z_off);
    float zn = fmaxf(z_half - z_off, 0.0f), zf = fmaxf(z_half + z_off, 0.0f);
    if (!ok) { zn = 0.0f; zf = 0.0f; }
    if (near) near[i] = zn;
    if (far) far[i] = zf;
    if (mask) mask[i] = ok ? 1 : 0;
    if (pts) {
        float *p = pts + (size_t)i * 6;
        p[0] = ox + dx * zn; p[1] = oy + dy * zn; p[2] = oz + dz * zn;
        p[3] = ox + dx * zf; p[4] = oy + dy * zf; p[5] = oz + dz * zf;
    }
}

// torch.linspace(0, 1, steps)[k] in fp32: start + step*k for the first half, end - step*(steps-1-k) for the second
__device__ __forceinline__ float linspace01(int k, int steps) {
    if (steps <= 1) return 0.0f;
    const float step = 1.0f / (float)(steps - 1);
    return k < steps / 2 ? step * (float)k : 1.0f - step * (float)(steps - 1 - k);
}

__device__ __forceinline__ float zval_at(float near, float far, int k, int n_pts, int inclusive, int inverse_linear) {
    const float t = inclusive ? linspace01(k, n_pts) : linspace01(k + 1, n_pts + 2);
    if (inverse_linear) return 1.0f / (1.0f / (near + 1e-8f) * (1.0f - t) + 1.0f / (far + 1e-8f) * t);
    return near + (far - near) * t;
}

// one thread per (ray, sample); jitter needs the two neighbours, which are recomputed
__global__ void zvals_near_far_kernel(uint32_t n_rays, int n_pts, const float *__restrict__ near,
                                      const float *__restrict__ far, int inclusive, int inverse_linear,
                                      const float *__restrict__ u, float *__restrict__ zvals) {
    arc_pdl_sync();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)n_rays * n_pts) return;
    const uint32_t ray = (uint32_t)(i / n_pts);
    const int k = (int)(i - (size_t)ray * n_pts);
    const float a = near[ray], b = far[ray];
    const float z = zval_at(a, b, k, n_pts, inclusive, inverse_linear);
    if (!u) { zvals[i] = z; return; }
    const float lo = k > 0 ? 0.5f * (z + zval_at(a, b, k - 1, n_pts, inclusive, inverse_linear)) : z;
    const float hi = k + 1 < n_pts ? 0.5f * (zval_at(a, b, k + 1, n_pts, inclusive, inverse_linear) + z) : z;
    zvals[i] = lo + (hi - lo) * u[i];
}

// one warp per ray: fixed-step samples clamped to [near, far], duplicates of the far end masked out, jitter on the
// valid prefix, everything clamped to [first, last valid] (perturb_interval_with_mask)
__global__ void zvals_fix_step_kernel(uint32_t n_rays, int n_pts, const float *__restrict__ near,
                                      const float *__restrict__ far, float fix_t, int inclusive,
                                      const float *__restrict__ u, float *__restrict__ zvals, uint8_t *__restrict__ mask) {
    arc_pdl_sync();
    const uint32_t ray = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (ray >= n_rays) return;
    const float a = near[ray], b = far[ray];
    const float start = inclusive ? a : a + fix_t;
    auto raw = [&](int k) { return fminf(fmaxf(start + (float)k * fix_t, a), b); };
    auto valid = [&](int k) { return k == 0 || (raw(k) - raw(k - 1)) != 0.0f; };
    auto jittered = [&](int k) {
        const float z = raw(k);
        if (!u) return z;
        const float lo = k > 0 ? 0.5f * (z + raw(k - 1)) : z;
        const float hi = k + 1 < n_pts ? 0.5f * (raw(k + 1) + z) : z;
        return lo + (hi - lo) * u[(size_t)ray * n_pts + k];
    };
    // number of valid samples (the valid ones form a prefix unless near == far, where only sample 0 is valid)
    int n_valid = 0;
    for (int k0 = 0; k0 < n_pts; k0 += 32) {
        const int k = k0 + (int)lane;
        n_valid += __popc(__ballot_sync(ARC_FULL_MASK, k < n_pts && valid(k)));
    }
    // value at column (n_valid - 1) after the masked overwrite, and at column 0
    const int last = n_valid - 1;
    const float v_last = valid(last) ? jittered(last) : raw(last);
    const float v_first = jittered(0);
    for (int k = (int)lane; k < n_pts; k += 32) {
        const bool ok = valid(k);
        const float v = ok ? jittered(k) : raw(k);
        zvals[(size_t)ray * n_pts + k] = u ? fminf(fmaxf(v, v_first), v_last) : v;
        if (mask) mask[(size_t)ray * n_pts + k] = ok ? 1 : 0;
    }
}

__global__ void ray_points_kernel(size_t n_rays, int n_pts, const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                                  const float *__restrict__ zvals, float *__restrict__ pts) {
    arc_pdl_sync();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays * (size_t)n_pts) return;
    const size_t ray = i / n_pts;
    const float z = zvals[i];
    pts[3 * i] = rays_o[3 * ray] + rays_d[3 * ray] * z;
    pts[3 * i + 1] = rays_o[3 * ray + 1] + rays_d[3 * ray + 1] * z;
    pts[3 * i + 2] = rays_o[3 * ray + 2] + rays_d[3 * ray + 2] * z;
}

// the reference's `assert torch.all(deltas >= 0)` of ray_marching (ray_helper.py:532-534) as a device-side flag: deltas below 1e-5 in
// magnitude count as zero; a negative (or NaN) interval sets flag[0].  The host reads the flag without stalling the stream.
__global__ void check_increasing_kernel(size_t n_rays, int p, const float *__restrict__ zvals, int32_t *__restrict__ flag) {
    arc_pdl_sync();
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < 2 || k >= n_rays * (size_t)(p - 1)) return;
    const size_t r = k / (p - 1), c = k - r * (p - 1);
    float d = zvals[r * p + c + 1] - zvals[r * p + c];
    if (fabsf(d) < 1e-5f) d = 0.0f;
    if (!(d >= 0.0f)) atomicOr(flag, 1);
}

// one CTA per row: ascending bitonic sort of `n` floats in shared memory (n_pow2 = next power of two, padded with +inf)
__global__ void sort_rows_kernel(int n, int n_pow2, const float *in, float *out) {
    arc_pdl_sync();
    extern __shared__ float s[];
    const size_t row = blockIdx.x;
    for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) s[i] = i < n ? in[row * n + i] : INFINITY;
    __syncthreads();
    for (int k = 2; k <= n_pow2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
                const int p = i ^ j;
                if (p > i) {
                    const float a = s[i], b = s[p];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) { s[i] = b; s[p] = a; }
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) out[row * n + i] = s[i];
}

}  // namespace

extern "C" {

// near/far (n_rays, n_r) of rays against n_r concentric spheres (radius[n_r], centre origin[3] on the host); pts
// (n_rays, n_r, 2, 3) and mask (n_rays, n_r) bytes optional.  Six-case semantics of the reference (miss -> 0/0).
int arcb200_sphere_ray_intersection(const float *rays_o, const float *rays_d, const float *radius, const float *origin,
                                    float *near, float *far, float *pts, uint8_t *mask, int64_t n_rays, int64_t n_r,
                                    void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_r >= 1 && n_rays * n_r < (1ll << 31), "sphere_ray_intersection: sizes out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(rays_o && rays_d && radius && origin, "sphere_ray_intersection: null pointer");
    arc_launch(sphere_intersect_kernel, arc_blocks(n_rays * n_r, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n_rays,
               (uint32_t)n_r, rays_o, rays_d, radius, origin[0], origin[1], origin[2], near, far, pts, mask);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// zvals (n_rays, n_pts) between near and far (n_rays each): uniform in depth or in inverse depth, end points included
// or not; u (n_rays, n_pts) uniforms in [0,1) or NULL: jitter inside the half-way intervals (perturb_interval).
int arcb200_zvals_from_near_far(const float *near, const float *far, int n_pts, int inclusive, int inverse_linear,
                                const float *u, float *zvals, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_pts >= 1, "zvals_from_near_far: sizes out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(near && far && zvals, "zvals_from_near_far: null pointer");
    arc_launch(zvals_near_far_kernel, arc_blocks(n_rays * n_pts, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n_rays, n_pts,
               near, far, inclusive, inverse_linear, u, zvals);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Fixed-step variant: zvals[k] = clamp(start + k*fix_t, near, far), mask[k] = 0 for repeats of the far end; with u the
// valid samples are jittered and every value is clamped to [first, last valid] (the reference always jitters here).
int arcb200_zvals_fix_step(const float *near, const float *far, float fix_t, int n_pts, int inclusive, const float *u,
                           float *zvals, uint8_t *mask, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_pts >= 1, "zvals_fix_step: sizes out of range");
    ARC_REQUIRE(fix_t > 0.0f, "zvals_fix_step: only a positive step is allowed");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(near && far && zvals, "zvals_fix_step: null pointer");
    arc_launch(zvals_fix_step_kernel, arc_blocks(n_rays * 32, 256), 256, 0, (cudaStream_t)stream, (uint32_t)n_rays, n_pts, near,
               far, fix_t, inclusive, u, zvals, mask);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// pts (n_rays, n_pts, 3) = rays_o + rays_d * zvals
int arcb200_ray_points(const float *rays_o, const float *rays_d, const float *zvals, float *pts, int64_t n_rays, int n_pts,
                       void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_pts >= 1, "ray_points: sizes out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(rays_o && rays_d && zvals && pts, "ray_points: null pointer");
    arc_launch(ray_points_kernel, arc_blocks(n_rays * n_pts, 256), 256, 0, (cudaStream_t)stream, (size_t)n_rays, n_pts, rays_o,
               rays_d, zvals, pts);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// flag[0] |= 1 when some row of zvals (n_rays, p) decreases by 1e-5 or more (or holds a NaN): ray_marching's assertion
// (ray_helper.py:532-534) evaluated on the device; the caller clears the flag and reads it when convenient.
int arcb200_check_increasing(const float *zvals, int64_t n_rays, int p, int32_t *flag, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && p >= 0 && flag, "check_increasing: bad argument");
    if (n_rays == 0 || p < 2) return ARCB200_OK;
    ARC_REQUIRE(zvals != nullptr, "check_increasing: null zvals");
    arc_launch(check_increasing_kernel, arc_blocks(n_rays * (p - 1), 256), 256, 0, (cudaStream_t)stream, (size_t)n_rays, p, zvals, flag);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

// Ascending sort of every row of vals (n_rows, n), n <= 8192: the merge of coarse and importance samples
// (torch.sort(torch.cat([zvals, new], -1), -1), nerf_model.py:112).  out may alias in.
int arcb200_sort_rows(const float *in, float *out, int64_t n_rows, int n, void *stream) {
    ARC_REQUIRE(n_rows >= 0 && n >= 1, "sort_rows: sizes out of range");
    if (n > 8192) { arcb200_set_error("sort_rows: rows of %d values, at most 8192 supported", n); return ARCB200_ERR_LIMIT; }
    if (n_rows == 0) return ARCB200_OK;
    ARC_REQUIRE(in && out, "sort_rows: null pointer");
    int p2 = 1;
    while (p2 < n) p2 <<= 1;
    const int threads = p2 >= 512 ? 256 : (p2 >= 64 ? p2 / 2 : 32);
    arc_launch(sort_rows_kernel, (unsigned)n_rows, threads, (size_t)p2 * sizeof(float), (cudaStream_t)stream, n, p2, in, out);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
