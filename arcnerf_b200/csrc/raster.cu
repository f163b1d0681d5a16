// Triangle-mesh rasteriser for the extracted-mesh preview: replaces the pytorch3d renderer the reference drives
// from render_mesh_images (arcnerf/geometry/mesh.py:174-240, arcnerf/render/render_pytorch3d.py: MeshRasterizer with blur_radius 0,
// one face per pixel, vertex or face colours).  No graphics API: two compute passes over a 64-bit z-buffer.
//   pass 1  one thread per (view, face): project, walk the pixel centres of the triangle's bounding box, atomicMin of
//           (depth bits << 32 | face id) - the nearest face wins, ties go to the lower face id (deterministic)
//   pass 2  one thread per pixel: recompute the perspective-correct barycentrics of the winning face, blend the vertex colours
//           (or take the face colour), write uint8 RGB; pixels no face covers keep the background colour
// Pinhole camera as in ray_helper.get_rays: x_cam = w2c . X, u = fx x / z + cx, v = fy y / z + cy, pixel (row v, column u), sampled
// at pixel centres.  Faces with a vertex at z <= near are skipped (no clipping); both windings are drawn (pytorch3d's default
// cull_backfaces = False).  pytorch3d's Phong lighting is NOT reproduced: the image is the unlit colour (parity unpinned, pytorch3d
// is not in this image; the kernels are tested bit for bit against a numpy restatement, oracle/raster.py).
// Compiled with -fmad=false.
#include "common.cuh"

namespace {

struct Cam {
    float r[12];            // rows 0..2 of w2c
    float fx, fy, cx, cy;
};

__device__ __forceinline__ Cam load_cam(const float *__restrict__ w2c, int view, float4 k) {
    Cam c;
    #pragma unroll
    for (int q = 0; q < 12; ++q) c.r[q] = w2c[16 * (size_t)view + q];
    c.fx = k.x; c.fy = k.y; c.cx = k.z; c.cy = k.w;
    return c;
}

__device__ __forceinline__ void project(const Cam &c, const float *__restrict__ v, float &u, float &w, float &z) {
    const float x = c.r[0] * v[0] + c.r[1] * v[1] + c.r[2] * v[2] + c.r[3];
    const float y = c.r[4] * v[0] + c.r[5] * v[1] + c.r[6] * v[2] + c.r[7];
    z = c.r[8] * v[0] + c.r[9] * v[1] + c.r[10] * v[2] + c.r[11];
    u = c.fx * x / z + c.cx;
    w = c.fy * y / z + c.cy;
}

// signed doubled area of (a, b, p)
__device__ __forceinline__ float edge(float ax, float ay, float bx, float by, float px, float py) {
    return (bx - ax) * (py - ay) - (by - ay) * (px - ax);
}

struct Tri {
    float u[3], v[3], z[3], area;
    bool ok;
};

__device__ __forceinline__ Tri setup(const Cam &c, const float *__restrict__ verts, const int32_t *__restrict__ face, float near) {
    Tri t;
    t.ok = true;
    #pragma unroll
    for (int k = 0; k < 3; ++k) {
        project(c, verts + 3 * (size_t)face[k], t.u[k], t.v[k], t.z[k]);
        if (!(t.z[k] > near)) t.ok = false;
    }
    t.area = edge(t.u[0], t.v[0], t.u[1], t.v[1], t.u[2], t.v[2]);
    if (t.area == 0.f || !isfinite(t.area)) t.ok = false;
    return t;
}

// barycentrics of the pixel centre; inside when all three have the sign of the area
__device__ __forceinline__ bool inside(const Tri &t, float px, float py, float &b0, float &b1, float &b2) {
    const float w0 = edge(t.u[1], t.v[1], t.u[2], t.v[2], px, py);
    const float w1 = edge(t.u[2], t.v[2], t.u[0], t.v[0], px, py);
    const float w2 = edge(t.u[0], t.v[0], t.u[1], t.v[1], px, py);
    const bool in = t.area > 0.f ? (w0 >= 0.f && w1 >= 0.f && w2 >= 0.f) : (w0 <= 0.f && w1 <= 0.f && w2 <= 0.f);
    b0 = w0 / t.area; b1 = w1 / t.area; b2 = w2 / t.area;
    return in;
}

__device__ __forceinline__ float depth_at(const Tri &t, float b0, float b1, float b2) {
    return 1.0f / (b0 / t.z[0] + b1 / t.z[1] + b2 / t.z[2]);          // 1/z is linear in screen space
}

__global__ void raster_depth_kernel(const float *__restrict__ verts, const int32_t *__restrict__ faces, int64_t n_f,
                                    const float *__restrict__ w2c, float4 intr, int n_views, int H, int W, float near,
                                    unsigned long long *__restrict__ zbuf) {
    arc_pdl_sync();
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_f * n_views) return;
    const int view = (int)(k / n_f);
    const int64_t f = k - (int64_t)view * n_f;
    const Cam c = load_cam(w2c, view, intr);
    const Tri t = setup(c, verts, faces + 3 * f, near);
    if (!t.ok) return;
    const float umin = fminf(t.u[0], fminf(t.u[1], t.u[2])), umax = fmaxf(t.u[0], fmaxf(t.u[1], t.u[2]));
    const float vmin = fminf(t.v[0], fminf(t.v[1], t.v[2])), vmax = fmaxf(t.v[0], fmaxf(t.v[1], t.v[2]));
    if (!(umax >= 0.f && vmax >= 0.f && umin <= (float)W && vmin <= (float)H)) return;
    const int c0 = max(0, (int)floorf(umin - 0.5f)), c1 = min(W - 1, (int)ceilf(umax - 0.5f));
    const int r0 = max(0, (int)floorf(vmin - 0.5f)), r1 = min(H - 1, (int)ceilf(vmax - 0.5f));
    unsigned long long *zb = zbuf + (size_t)view * H * W;
    for (int r = r0; r <= r1; ++r) {
        for (int cc = c0; cc <= c1; ++cc) {
            float b0, b1, b2;
            if (!inside(t, (float)cc + 0.5f, (float)r + 0.5f, b0, b1, b2)) continue;
            const float z = depth_at(t, b0, b1, b2);
            if (!(z > near)) continue;
            const unsigned long long key = ((unsigned long long)__float_as_uint(z) << 32) | (unsigned long long)(uint32_t)f;
            atomicMin(zb + (size_t)r * W + cc, key);
        }
    }
}

__global__ void raster_resolve_kernel(const float *__restrict__ verts, const int32_t *__restrict__ faces,
                                      const float *__restrict__ vert_colors, const float *__restrict__ face_colors,
                                      const float *__restrict__ w2c, float4 intr, int n_views, int H, int W, float near, float bkg,
                                      const unsigned long long *__restrict__ zbuf, uint8_t *__restrict__ image,
                                      int32_t *__restrict__ face_id, float *__restrict__ depth) {
    arc_pdl_sync();
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n_pix = (int64_t)H * W;
    if (k >= n_pix * n_views) return;
    const int view = (int)(k / n_pix);
    const int64_t p = k - (int64_t)view * n_pix;
    const unsigned long long key = zbuf[k];
    float col[3] = {bkg, bkg, bkg};
    int32_t fid = -1;
    float z = 0.f;
    if (key != ~0ull) {
        fid = (int32_t)(uint32_t)(key & 0xffffffffull);
        z = __uint_as_float((uint32_t)(key >> 32));
        if (face_colors) {
            #pragma unroll
            for (int q = 0; q < 3; ++q) col[q] = face_colors[3 * (size_t)fid + q];
        } else if (vert_colors) {
            const Cam c = load_cam(w2c, view, intr);
            const int32_t *face = faces + 3 * (size_t)fid;
            const Tri t = setup(c, verts, face, near);
            const int r = (int)(p / W), cc = (int)(p - (int64_t)r * W);
            float b0, b1, b2;
            inside(t, (float)cc + 0.5f, (float)r + 0.5f, b0, b1, b2);
            // perspective-correct weights: b_i / z_i, renormalised
            const float q0 = b0 / t.z[0], q1 = b1 / t.z[1], q2 = b2 / t.z[2];
            const float s = q0 + q1 + q2;
            #pragma unroll
            for (int q = 0; q < 3; ++q)
                col[q] = (q0 * vert_colors[3 * (size_t)face[0] + q] + q1 * vert_colors[3 * (size_t)face[1] + q] +
                          q2 * vert_colors[3 * (size_t)face[2] + q]) / s;
        } else {
            col[0] = col[1] = col[2] = 0.f;              // silhouette: covered pixels black on the background
        }
    }
    if (image) {
        #pragma unroll
        for (int q = 0; q < 3; ++q) image[3 * k + q] = (uint8_t)(fminf(fmaxf(col[q], 0.f), 1.f) * 255.0f + 0.5f);
    }
    if (face_id) face_id[k] = fid;
    if (depth) depth[k] = z;
}

}  // namespace

extern "C" {

int arcb200_raster_scratch_bytes(int n_views, int H, int W, int64_t *bytes) {
    if (!bytes || n_views < 0 || H < 0 || W < 0) return ARCB200_ERR_ARG;
    *bytes = (int64_t)n_views * H * W * 8;
    return ARCB200_OK;
}

// w2c: DEVICE (n_views, 4, 4) row-major; intrinsic: HOST 3 x 3 row-major.  scratch: arcb200_raster_scratch_bytes (the z-buffer).
int arcb200_raster_mesh(const float *verts, const int32_t *faces, const float *vert_colors, const float *face_colors, const float *w2c,
                        const float *intrinsic, int n_views, int H, int W, float near, float bkg, void *scratch, uint8_t *image,
                        int32_t *face_id, float *depth, int64_t n_v, int64_t n_f, void *stream) {
    ARC_REQUIRE(n_views >= 0 && H >= 0 && W >= 0 && n_v >= 0 && n_f >= 0 && n_f < (1ll << 31), "raster_mesh: sizes out of range");
    if (n_views == 0 || H == 0 || W == 0) return ARCB200_OK;
    ARC_REQUIRE(w2c && intrinsic && scratch && (n_f == 0 || (verts && faces)), "raster_mesh: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long *zbuf = (unsigned long long *)scratch;
    const size_t n_pix = (size_t)n_views * H * W;
    ARC_REQUIRE(cudaMemsetAsync(zbuf, 0xff, n_pix * 8, st) == cudaSuccess, "raster_mesh: memset failed");
    const float4 intr = make_float4(intrinsic[0], intrinsic[4], intrinsic[2], intrinsic[5]);
    if (n_f > 0)
        arc_launch(raster_depth_kernel, arc_blocks(n_f * n_views, 128), 128, 0, st, verts, faces, n_f, w2c, intr, n_views, H, W, near, zbuf);
    arc_launch(raster_resolve_kernel, arc_blocks((int64_t)n_pix, 256), 256, 0, st, verts, faces, vert_colors, face_colors, w2c, intr,
               n_views, H, W, near, bkg, (const unsigned long long *)zbuf, image, face_id, depth);
    ARC_LAUNCH_CHECK();
    return ARCB200_OK;
}

}  // extern "C"
