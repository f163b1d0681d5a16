// Camera rays on the device: what the reference builds on the host / in eager torch before every forward.
// Reference: arcnerf/render/ray_helper.py:12-119 `get_rays` (pixel grid in (W, H) "wh" order or row-major, optional half-pixel
// centre, pixel -> camera with fx, fy, cx, cy and skew (geometry/projection.py:8-34), camera -> world with c2w, direction =
// world point - camera centre, optional normalisation, mip-NeRF ray radius for full images) and the training pipeline's batch
// gather (arcnerf/trainer/pipeline.py:204-300: a random batch of rays out of all training images).
// Here one thread makes one ray from (camera id, pixel id): a batch of rays from MANY cameras is one launch, no pixel grid, no
// per-image tensors and no gather.  Arithmetic order follows the reference expression by expression (fp32, this file is compiled
// without FMA contraction), so the results match its CPU output to the last bit or two.
#include "common.cuh"

namespace {

struct Cam { float fx, fy, cx, cy, s; float c2w[12]; };      // c2w: first three rows of the 4x4 pose, row-major

__device__ __forceinline__ void make_dir(const Cam &c, float i, float j, int normalize, float &dx, float &dy, float &dz) {
    // pixel_to_cam with z = 1, then cam_to_world (rotation + translation), then minus the camera centre
    const float yc = (j - c.cy) / c.fy;
    const float xc = (i - (c.s * (j - c.cy) / c.fy) - c.cx) / c.fx;
    const float wx = c.c2w[0] * xc + c.c2w[1] * yc + c.c2w[2] * 1.0f + c.c2w[3];
    const float wy = c.c2w[4] * xc + c.c2w[5] * yc + c.c2w[6] * 1.0f + c.c2w[7];
    const float wz = c.c2w[8] * xc + c.c2w[9] * yc + c.c2w[10] * 1.0f + c.c2w[11];
    dx = wx - c.c2w[3]; dy = wy - c.c2w[7]; dz = wz - c.c2w[11];
    if (normalize) {
        const float nrm = sqrtf(dx * dx + dy * dy + dz * dz);
        dx /= nrm; dy /= nrm; dz /= nrm;
    }
}

__global__ void gen_rays_kernel(int64_t n, const float *__restrict__ cams, int n_cams, const int32_t *__restrict__ cam_id,
                                const int64_t *__restrict__ pix, int W, int H, int wh_order, int center_pixel, int normalize,
                                float *__restrict__ rays_o, float *__restrict__ rays_d, float *__restrict__ rays_r) {
    arc_pdl_sync();
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (int64_t)gridDim.x * blockDim.x) {
        const int ci = cam_id ? __ldg(cam_id + k) : 0;
        Cam c;
        const float *p = cams + (size_t)(ci < n_cams ? ci : 0) * 17;
        c.fx = p[0]; c.fy = p[1]; c.cx = p[2]; c.cy = p[3]; c.s = p[4];
        #pragma unroll
        for (int q = 0; q < 12; ++q) c.c2w[q] = p[5 + q];
        // pixel of ray k: an explicit (W, H)-order index, or the k-th pixel of the full image in the requested order
        int64_t id = pix ? __ldg(pix + k) : k;
        int ix, iy;
        if (pix || wh_order) { ix = (int)(id / H); iy = (int)(id - (int64_t)ix * H); }
        else { iy = (int)(id / W); ix = (int)(id - (int64_t)iy * W); }
        const float off = center_pixel ? 0.5f : 0.0f;
        float dx, dy, dz;
        make_dir(c, (float)ix + off, (float)iy + off, normalize, dx, dy, dz);
        rays_o[3 * k] = c.c2w[3]; rays_o[3 * k + 1] = c.c2w[7]; rays_o[3 * k + 2] = c.c2w[11];
        rays_d[3 * k] = dx; rays_d[3 * k + 1] = dy; rays_d[3 * k + 2] = dz;
        if (rays_r) {
            // mip-NeRF radius: distance to the direction of the next pixel along x (the last column takes the one before it),
            // times 2 / sqrt(12)
            int nx = ix + 1 < W ? ix + 1 : ix - 1, bx = ix + 1 < W ? ix : ix - 2;
            if (W < 3) { nx = W > 1 ? 1 : 0; bx = 0; }
            float ax, ay, az, ex, ey, ez;
            make_dir(c, (float)bx + off, (float)iy + off, normalize, ax, ay, az);
            make_dir(c, (float)nx + off, (float)iy + off, normalize, ex, ey, ez);
            const float d2 = (ax - ex) * (ax - ex) + (ay - ey) * (ay - ey) + (az - ez) * (az - ez);
            rays_r[k] = sqrtf(d2) * 2.0f / sqrtf(12.0f);
        }
    }
}

}  // namespace

extern "C" {

// cams: DEVICE array of n_cams x 17 floats: fx, fy, cx, cy, skew, then the first three rows of c2w (row-major).
// cam_id (n) int32 or NULL (camera 0); pix (n) int64 pixel index i * H + j (the reference's unrolled (i, j) index) or NULL:
// ray k is pixel k of the full W x H image in wh order (column-major, wh_order != 0) or row-major; then n must be W * H.
// rays_o, rays_d (n,3) fp32; rays_r (n) fp32 or NULL (full image only, like the reference).
int arcb200_gen_rays(const float *cams, int n_cams, const int32_t *cam_id, const int64_t *pix, int W, int H, int wh_order,
                     int center_pixel, int normalize, float *rays_o, float *rays_d, float *rays_r, int64_t n, void *stream) {
    ARC_REQUIRE(n >= 0 && W > 0 && H > 0 && n_cams > 0, "gen_rays: bad sizes");
    if (n == 0) return ARCB200_OK;
    ARC_REQUIRE(cams && rays_o && rays_d, "gen_rays: null pointer");
    ARC_REQUIRE(pix || n == (int64_t)W * H, "gen_rays: without a pixel list n must be W * H");
    ARC_REQUIRE(!rays_r || !pix, "gen_rays: the ray radius is defined for full images only");
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)arc_sm_count() * 16;
    arc_launch(gen_rays_kernel, (unsigned)(blocks < cap ? blocks : cap), 256, 0, (cudaStream_t)stream, n, cams, n_cams, cam_id, pix, W, H,
               wh_order, center_pixel, normalize, rays_o, rays_d, rays_r);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
