// Surface-ray intersection for SDF / density fields: the per-ray state machines of
// arcnerf/geometry/ray.py:403-601 (sphere_tracing, secant_root_finding) as device kernels around the caller's field
// evaluation.  The reference runs these loops with boolean-mask gathers / scatters (index_put, nonzero, a host
// synchronisation per `torch.all(...)` test) - about a dozen eager kernels and two syncs per iteration.  Here an iteration is
// "points -> the caller's field kernels -> ONE update kernel", every ray is carried through every iteration (a ray that has
// converged or left [near, far] ignores its new value, exactly what the reference's masks do), and the early exits of the
// reference become device-side flags, so the loop never returns to the host.
//
// Arithmetic is one rounded fp32 operation at a time (-fmad=false), in the order of the torch expressions.
// Reproduced quirk (ray.py:571-573): `if ~ind_low.sum() > 0` never holds (bitwise not of a non-negative count), so the secant
// iteration only ever moves its LOW end; the high end keeps its bracket value.
#include "common.cuh"

namespace {

// sphere tracing, one iteration after the field was evaluated at pts = o + z d   (ray.py:441-461)
__global__ void trace_update_kernel(size_t n, const float *__restrict__ sdf_new, float *__restrict__ zvals, float *__restrict__ sdf,
                                    const float *__restrict__ near, const float *__restrict__ far, int near_far_per_ray,
                                    uint8_t *__restrict__ mask, uint8_t *__restrict__ obj, float threshold) {
    arc_pdl_sync();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool m = mask[i] != 0, o = obj[i] != 0;
    float s = sdf[i];
    if (!o && m) s = sdf_new[i];                          // sdf[valid_mask] = sdf_func(pts)
    if (fabsf(s) < threshold) o = true;                   // obj_mask[|sdf| < threshold] = True
    float z = zvals[i];
    if (!o && m) z = z + s;                               // zvals[~obj & mask] += sdf
    const float nr = near[near_far_per_ray ? i : 0], fr = far[near_far_per_ray ? i : 0];
    if (z > fr) m = false;
    if (z < nr) m = false;
    sdf[i] = s; zvals[i] = z; mask[i] = m; obj[i] = o;
}

// zvals[zvals <= near] = 0, then pts = o + z d   (ray.py:463-467, 596-600)
__global__ void surface_finish_kernel(size_t n, const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                                      float *__restrict__ zvals, const float *__restrict__ near, int near_per_ray,
                                      float *__restrict__ pts) {
    arc_pdl_sync();
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float z = zvals[i];
    if (z <= near[near_per_ray ? i : 0]) z = 0.0f;
    zvals[i] = z;
    pts[3 * i] = rays_o[3 * i] + rays_d[3 * i] * z;
    pts[3 * i + 1] = rays_o[3 * i + 1] + rays_d[3 * i + 1] * z;
    pts[3 * i + 2] = rays_o[3 * i + 2] + rays_d[3 * i + 2] * z;
}

// secant method, the uniform scan: step[r][j] = near (1 - t_j) + far t_j with t = linspace(0, 1, n_step), and its points
__global__ void secant_steps_kernel(size_t n_rays, int n_step, const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                                    const float *__restrict__ near, const float *__restrict__ far, int near_far_per_ray,
                                    float *__restrict__ step, float *__restrict__ pts) {
    arc_pdl_sync();
    const size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_rays * (size_t)n_step) return;
    const size_t r = k / n_step;
    const int j = (int)(k - r * n_step);
    // torch.linspace(0, 1, n): one fused multiply-add per value from the nearer end (see csrc/mcubes.cu)
    const float dt = n_step > 1 ? 1.0f / (float)(n_step - 1) : 0.0f;
    const float t = j < n_step / 2 ? fmaf(dt, (float)j, 0.0f) : fmaf(-dt, (float)(n_step - 1 - j), 1.0f);
    const float nr = near[near_far_per_ray ? r : 0], fr = far[near_far_per_ray ? r : 0];
    const float a = nr * (1.0f - t), b = fr * t;
    const float z = a + b;
    step[k] = z;
    pts[3 * k] = rays_o[3 * r] + rays_d[3 * r] * z;
    pts[3 * k + 1] = rays_o[3 * r + 1] + rays_d[3 * r + 1] * z;
    pts[3 * k + 2] = rays_o[3 * r + 2] + rays_d[3 * r + 2] * z;
}

__device__ __forceinline__ float sgn(float x) { return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f); }

// secant method, bracket: first sign change + -> - along the scan   (ray.py:515-553)
__global__ void secant_init_kernel(size_t n_rays, int n_step, const float *__restrict__ geo, const float *__restrict__ step,
                                   float level, int descent, uint8_t *__restrict__ not_occ, uint8_t *__restrict__ mask,
                                   float *__restrict__ z_high, float *__restrict__ z_low, float *__restrict__ g_high,
                                   float *__restrict__ g_low, float *__restrict__ z_mid, float *__restrict__ z_mid_init) {
    arc_pdl_sync();
    const size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rays) return;
    const float *g = geo + r * n_step;
    const float *st = step + r * n_step;
    auto diff = [&](int j) { const float d = g[j] - level; return descent ? d * -1.0f : d; };
    // cost_j = sign(d_j d_{j+1}) (n_step - j) for j < n_step - 1, cost_{n-1} = 1; torch.min: the smallest, first index on ties
    float best = 0.f;
    int index = 0;
    float prev = diff(0);
    const float d0 = prev;
    for (int j = 0; j < n_step; ++j) {
        float cost;
        if (j + 1 < n_step) {
            const float nxt = diff(j + 1);
            cost = sgn(prev * nxt) * (float)(n_step - j);
            prev = nxt;
        } else {
            cost = 1.0f;
        }
        if (j == 0 || cost < best) { best = cost; index = j; }
    }
    const bool m_not_occ = d0 > 0.f;
    const bool m_change = best < 0.f;
    const float d_idx = diff(index);
    const bool m_pos = d_idx > 0.f;
    const bool m = m_not_occ && m_change && m_pos;
    const int i2 = index + 1 < n_step ? index + 1 : n_step - 1;
    const float zh = st[index], gh = d_idx, zl = st[i2], gl = diff(i2);
    // z_mid = -geo_low * (z_high - z_low) / (geo_high - geo_low) + z_low
    const float zm = (-gl) * (zh - zl) / (gh - gl) + zl;
    not_occ[r] = m_not_occ; mask[r] = m;
    z_high[r] = zh; g_high[r] = gh; z_low[r] = zl; g_low[r] = gl; z_mid[r] = zm; z_mid_init[r] = zm;
}

// secant method, one iteration after the field was evaluated at z_mid   (ray.py:557-579).
// flags[it] counts the rays whose estimate after iteration it - 1 is at least `threshold` away from their first estimate; the
// reference stops before iteration it > 0 when that count is zero.  A skipped iteration counts nothing, so every later
// iteration finds a zero as well: the stop is sticky without further state.
__global__ void secant_update_kernel(size_t n_rays, const float *__restrict__ geo_mid, float level, int descent,
                                     const uint8_t *__restrict__ mask, float *__restrict__ z_low, float *__restrict__ g_low,
                                     const float *__restrict__ z_high, const float *__restrict__ g_high, float *__restrict__ z_mid,
                                     const float *__restrict__ z_mid_init, float threshold, int32_t *__restrict__ flags, int it) {
    arc_pdl_sync();
    const size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rays) return;
    if (it > 0 && flags[it] == 0) return;                                 // every ray within `threshold` of its first estimate: stop now
    if (!mask[r]) return;
    float d = geo_mid[r] - level;
    if (descent) d = d * -1.0f;
    float zl = z_low[r], gl = g_low[r];
    const float zm = z_mid[r];
    if (d < 0.f) { zl = zm; gl = d; z_low[r] = zl; g_low[r] = gl; }
    // the high end is never moved (see the header)
    const float zh = z_high[r], gh = g_high[r];
    const float nz = (-gl) * (zh - zl) / (gh - gl) + zl;
    z_mid[r] = nz;
    if (!(fabsf(z_mid_init[r] - nz) < threshold)) atomicAdd(&flags[it + 1], 1);
}

// zvals[mask] = z_mid; zvals[~mask] = far; zvals[~not_occ] = 0   (ray.py:591-594); the near clamp and the points follow in
// surface_finish_kernel
__global__ void secant_select_kernel(size_t n_rays, const uint8_t *__restrict__ mask, const uint8_t *__restrict__ not_occ,
                                     const float *__restrict__ z_mid, const float *__restrict__ far, int far_per_ray,
                                     float *__restrict__ zvals) {
    arc_pdl_sync();
    const size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rays) return;
    float z = mask[r] ? z_mid[r] : far[far_per_ray ? r : 0];
    if (!not_occ[r]) z = 0.0f;
    zvals[r] = z;
}

}  // namespace

extern "C" {

int arcb200_trace_update(const float *sdf_new, float *zvals, float *sdf, const float *near, const float *far, int near_far_per_ray,
                         uint8_t *mask, uint8_t *obj, float threshold, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0, "trace_update: negative size");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(sdf_new && zvals && sdf && near && far && mask && obj, "trace_update: null pointer");
    arc_launch(trace_update_kernel, arc_blocks(n_rays, 256), 256, 0, (cudaStream_t)stream, (size_t)n_rays, sdf_new, zvals, sdf, near, far,
               near_far_per_ray, mask, obj, threshold);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_surface_finish(const float *rays_o, const float *rays_d, float *zvals, const float *near, int near_per_ray, float *pts,
                           int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0, "surface_finish: negative size");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(rays_o && rays_d && zvals && near && pts, "surface_finish: null pointer");
    arc_launch(surface_finish_kernel, arc_blocks(n_rays, 256), 256, 0, (cudaStream_t)stream, (size_t)n_rays, rays_o, rays_d, zvals, near,
               near_per_ray, pts);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_secant_steps(const float *rays_o, const float *rays_d, const float *near, const float *far, int near_far_per_ray,
                         int n_step, float *step, float *pts, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_step >= 1 && n_rays * (int64_t)n_step < (1ll << 40), "secant_steps: sizes out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(rays_o && rays_d && near && far && step && pts, "secant_steps: null pointer");
    arc_launch(secant_steps_kernel, arc_blocks(n_rays * n_step, 256), 256, 0, (cudaStream_t)stream, (size_t)n_rays, n_step, rays_o, rays_d,
               near, far, near_far_per_ray, step, pts);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_secant_init(const float *geo, const float *step, int n_step, float level, int descent, uint8_t *not_occ, uint8_t *mask,
                        float *z_high, float *z_low, float *g_high, float *g_low, float *z_mid, float *z_mid_init, int32_t *flags,
                        int n_iter, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && n_step >= 1 && n_iter >= 0, "secant_init: sizes out of range");
    ARC_REQUIRE(flags, "secant_init: null flags");
    cudaStream_t st = (cudaStream_t)stream;
    ARC_REQUIRE(cudaMemsetAsync(flags, 0, sizeof(int32_t) * (size_t)(n_iter + 1), st) == cudaSuccess, "secant_init: memset failed");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(geo && step && not_occ && mask && z_high && z_low && g_high && g_low && z_mid && z_mid_init, "secant_init: null pointer");
    arc_launch(secant_init_kernel, arc_blocks(n_rays, 128), 128, 0, st, (size_t)n_rays, n_step, geo, step, level, descent, not_occ, mask,
               z_high, z_low, g_high, g_low, z_mid, z_mid_init);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_secant_update(const float *geo_mid, float level, int descent, const uint8_t *mask, float *z_low, float *g_low,
                          const float *z_high, const float *g_high, float *z_mid, const float *z_mid_init, float threshold,
                          int32_t *flags, int it, int n_iter, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0 && it >= 0 && it < n_iter, "secant_update: iteration out of range");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(geo_mid && mask && z_low && g_low && z_high && g_high && z_mid && z_mid_init && flags, "secant_update: null pointer");
    arc_launch(secant_update_kernel, arc_blocks(n_rays, 256), 256, 0, (cudaStream_t)stream, (size_t)n_rays, geo_mid, level, descent, mask,
               z_low, g_low, z_high, g_high, z_mid, z_mid_init, threshold, flags, it);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

int arcb200_secant_select(const uint8_t *mask, const uint8_t *not_occ, const float *z_mid, const float *far, int far_per_ray,
                          float *zvals, int64_t n_rays, void *stream) {
    ARC_REQUIRE(n_rays >= 0, "secant_select: negative size");
    if (n_rays == 0) return ARCB200_OK;
    ARC_REQUIRE(mask && not_occ && z_mid && far && zvals, "secant_select: null pointer");
    arc_launch(secant_select_kernel, arc_blocks(n_rays, 256), 256, 0, (cudaStream_t)stream, (size_t)n_rays, mask, not_occ, z_mid, far,
               far_per_ray, zvals);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
