// TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
//
// Host-side driver around the REFERENCE's own HOST_DEVICE sampler functions. The reference
// sources are #included from where they lie under /root/reference (see oracle/Makefile);
// nothing of them is copied into this repository. Each driver below restates, as a host
// `for` loop over work items, the body of one reference __global__ kernel that cannot itself
// run without a GPU, and calls the reference's inline functions for all the arithmetic:
//   volume_func.h:17-56   aabb_ray_intersect          volume_func.h:72-88   density_grid_occupied_at
//   volume_func.h:100-134 distance/advance_to_next_voxel
//   volume_func.h:142-194 morton3D / density_grid_occupied_at_bit
//   volume_func.h:225-297 multivol helpers            pcg32.h               pcg32
// The kernel bodies driven are cited at each function. Compiled by nvcc as host code
// (x86-64 baseline: no FMA contraction), output goes to oracle/_ref/libarcref.so.
#include <cstdint>
#include <cmath>
#include <algorithm>
#include "common.h"
#include "volume_func.h"

using std::min;
using std::max;

// volume_func.h:201-223 is __device__-only (frexpf); same arithmetic for the host.
static int host_mip_from_pos(const Vector3f &pos, const Vector3f xyz_min, const Vector3f xyz_max,
                             const uint32_t n_cascades) {
    int ex, ey, ez;
    Vector3f c = (xyz_min + xyz_max) / 2.0f;
    Vector3f h = (xyz_max - xyz_min) / 2.0f;
    Vector3f a = (pos - c).cwiseAbs().cwiseProduct(h.cwiseInverse());
    frexpf(a.x(), &ex); frexpf(a.y(), &ey); frexpf(a.z(), &ez);
    int e = max(max(ex, ey), ez);
    return min((int)n_cascades - 1, max(0, e));
}

extern "C" {

// volume_func_kernel.cu:74-123
void ref_aabb_intersection(int n_rays, int n_v, const float *o, const float *d, const float *aabb,
                           float *near, float *far, float *pts, uint8_t *mask) {
    const Vector3f *ro = (const Vector3f *)o, *rd = (const Vector3f *)d, *ab = (const Vector3f *)aabb;
    Vector3f *p = (Vector3f *)pts;
    for (int i = 0; i < n_rays * n_v; ++i) {
        int r = i / n_v, v = i % n_v;
        Vector2f tm = aabb_ray_intersect(ro[r], rd[r], ab[v * 2], ab[v * 2 + 1]);
        if (tm.x() > 0) { near[i] = tm.x(); far[i] = tm.y(); mask[i] = 1; }
        else { near[i] = 0.0f; far[i] = 0.0f; mask[i] = 0; }
        p[i * 2] = ro[r] + near[i] * rd[r];
        p[i * 2 + 1] = ro[r] + far[i] * rd[r];
    }
}

// volume_func_kernel.cu:16-37
void ref_check_pts_in_occ_voxel(int n, const float *xyz, const uint8_t *bitfield, const float *aabb,
                                int n_grid, uint8_t *out) {
    const Vector3f *x = (const Vector3f *)xyz, *ab = (const Vector3f *)aabb;
    for (int i = 0; i < n; ++i)
        out[i] = density_grid_occupied_at(x[i], (const bool *)bitfield, ab[0], ab[1], (uint32_t)n_grid);
}

static void fill_tail(float *z, uint32_t j, uint32_t n_pts) {
    float last = 0.f;
    if (j > 0 && j < n_pts) last = z[j - 1];
    while (j > 0 && j < n_pts) { z[j] = last; ++j; }
}

// volume_func_kernel.cu:174-236 (bit=0) and bitfield_func_kernel.cu:20-82 (bit=1).
// rng_state/rng_inc: the module-level pcg32 as it is when the launcher passes it by value.
void ref_sparse_volume_sampling(int bit, int n_rays, const float *o, const float *d, const float *near,
                                const float *far, const float *aabb, const uint8_t *bitfield, int n_grid,
                                int n_pts, float dt, float near_distance, uint64_t rng_state,
                                uint64_t rng_inc, float *zvals, uint8_t *mask) {
    const Vector3f *ro = (const Vector3f *)o, *rd = (const Vector3f *)d, *ab = (const Vector3f *)aabb;
    const Vector3f xyz_min = ab[0], xyz_max = ab[1];
    for (int i = 0; i < n_rays; ++i) {
        pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
        rng.advance((int64_t)((uint32_t)i * 8u));
        Vector3f _o = ro[i], _d = rd[i];
        float startt = fmaxf(near[i], near_distance);
        float far_end = far[i];
        startt += dt * random_val(rng);
        uint32_t j = 0; float t = startt; Vector3f pos;
        float *z = zvals + (size_t)i * n_pts; uint8_t *m = mask + (size_t)i * n_pts;
        while (t <= far_end && j < (uint32_t)n_pts && check_pts_in_aabb(pos = _o + _d * t, xyz_min, xyz_max)) {
            bool occ = bit ? density_grid_occupied_at_bit(pos, bitfield, xyz_min, xyz_max, n_grid)
                           : density_grid_occupied_at(pos, (const bool *)bitfield, xyz_min, xyz_max, n_grid);
            if (occ) { z[j] = t; m[j] = 1; ++j; t += dt; }
            else { t = advance_to_next_voxel(t, dt, pos, _d, xyz_min, xyz_max, n_grid); }
        }
        fill_tail(z, j, n_pts);
    }
}

// multivol_func_kernel.cu:19-106
void ref_sparse_sampling_in_multivol_bitfield(int n_rays, const float *o, const float *d, const float *near,
        const float *far, const float *min_aabb, const float *aabb, const uint8_t *bitfield, int n_grid,
        int n_cascade, int n_pts, float cone_angle, float min_step, float max_step, float near_distance,
        int inclusive, uint64_t rng_state, uint64_t rng_inc, float *zvals, uint8_t *mask) {
    const Vector3f *ro = (const Vector3f *)o, *rd = (const Vector3f *)d;
    const Vector3f xyz_min = ((const Vector3f *)aabb)[0], xyz_max = ((const Vector3f *)aabb)[1];
    const Vector3f mv_min = ((const Vector3f *)min_aabb)[0], mv_max = ((const Vector3f *)min_aabb)[1];
    for (int i = 0; i < n_rays; ++i) {
        pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
        rng.advance((int64_t)((uint32_t)i * 8u));
        Vector3f _o = ro[i], _d = rd[i];
        float startt = fmaxf(near[i], near_distance);
        float far_end = far[i];
        startt += calc_dt(startt, cone_angle, min_step, max_step) * random_val(rng);
        uint32_t j = 0; float t = startt; Vector3f pos;
        float *z = zvals + (size_t)i * n_pts; uint8_t *m = mask + (size_t)i * n_pts;
        while (t <= far_end && j < (uint32_t)n_pts && check_pts_in_aabb(pos = _o + _d * t, xyz_min, xyz_max)) {
            float dt = calc_dt(t, cone_angle, min_step, max_step);
            uint32_t mip = host_mip_from_pos(pos, mv_min, mv_max, n_cascade);
            if (mip == 0 && !inclusive) {
                while (j > 0) { z[j] = 0.0f; m[j] = 0; j--; }
                z[j] = 0.0f; m[j] = 0;
                t = advance_to_next_voxel(t, dt, pos, _d, mv_min, mv_max, n_grid);
            } else {
                if (density_grid_occupied_at_multivol(pos, bitfield, mip, mv_min, mv_max, n_grid, inclusive)) {
                    z[j] = t; m[j] = 1; ++j; t += dt;
                } else {
                    t = advance_to_next_voxel_multivol(t, cone_angle, min_step, max_step, pos, _d, mv_min, mv_max, n_grid);
                }
            }
        }
        fill_tail(z, j, n_pts);
    }
}

// bitfield_func_kernel.cu:143-185
void ref_generate_grid_samples(int n_elements, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                               const float *grid_in, float *out, uint32_t *indices, uint32_t n_grid,
                               float thresh) {
    Vector3f *po = (Vector3f *)out;
    for (uint32_t i = 0; i < (uint32_t)n_elements; ++i) {
        pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
        rng.advance((int64_t)(i * 4u));
        uint32_t n3 = n_grid * n_grid * n_grid;
        uint32_t idx = 0;
        for (uint32_t j = 0; j < 10; ++j) {
            idx = ((i + step * (uint32_t)n_elements) * 56924617u + j * 19349663u + 96925573u) % n3;
            if (grid_in[idx] > thresh) break;
        }
        uint32_t pi = idx % n3;
        uint32_t x = morton3D_invert(pi >> 0), y = morton3D_invert(pi >> 1), z = morton3D_invert(pi >> 2);
        Eigen::Vector3f pos = (Eigen::Vector3f{(float)x, (float)y, (float)z} + random_val_3d(rng)) / n_grid;
        po[i] = pos; indices[i] = idx;
    }
}

// multivol_func_kernel.cu:177-240
void ref_generate_grid_samples_multivol(int n_elements, const float *aabb, uint64_t rng_state, uint64_t rng_inc,
        uint32_t step, const float *grid_in, float *out, uint32_t *indices, uint32_t n_cascades,
        uint32_t n_grid, float thresh, int inclusive) {
    Vector3f *po = (Vector3f *)out; const Vector3f *ab = (const Vector3f *)aabb;
    for (uint32_t i = 0; i < (uint32_t)n_elements; ++i) {
        pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
        rng.advance((int64_t)(i * 4u));
        uint32_t level = 0;
        if (inclusive) level = (uint32_t)(random_val(rng) * n_cascades) % n_cascades;
        else while (level == 0) level = (uint32_t)(random_val(rng) * n_cascades) % n_cascades;
        uint32_t n3 = n_grid * n_grid * n_grid;
        uint32_t idx = 0;
        for (uint32_t j = 0; j < 10; ++j) {
            idx = ((i + step * (uint32_t)n_elements) * 56924617u + j * 19349663u + 96925573u) % n3;
            if (inclusive) idx += level * n3; else idx += (level - 1) * n3;
            if (grid_in[idx] > thresh) break;
        }
        uint32_t pi = idx % n3;
        uint32_t x = morton3D_invert(pi >> 0), y = morton3D_invert(pi >> 1), z = morton3D_invert(pi >> 2);
        Vector3f c = (ab[0] + ab[1]) / 2.0f; Vector3f len = ab[1] - ab[0];
        Eigen::Vector3f pos = (Eigen::Vector3f{(float)x, (float)y, (float)z} + random_val_3d(rng)) / n_grid;
        pos = pos - Eigen::Vector3f::Constant(0.5f);
        pos = pos.cwiseProduct(len) * scalbnf(1.0f, level) + c;
        po[i] = pos; indices[i] = idx;
    }
}

// pcg32.h: module rng is `static pcg32 rng{9121}` (common.h:23); launchers call rng.advance() after use.
void ref_rng_init(uint64_t seed, uint64_t *state, uint64_t *inc) { pcg32 r{seed}; *state = r.state; *inc = r.inc; }
void ref_rng_advance(uint64_t *state, uint64_t inc, int64_t delta) {
    pcg32 r; r.state = *state; r.inc = inc; r.advance(delta); *state = r.state;
}
float ref_rng_next_float(uint64_t *state, uint64_t inc) {
    pcg32 r; r.state = *state; r.inc = inc; float f = r.next_float(); *state = r.state; return f;
}
uint32_t ref_morton3D(uint32_t x, uint32_t y, uint32_t z) { return morton3D(x, y, z); }
uint32_t ref_morton3D_invert(uint32_t x) { return morton3D_invert(x); }

}  // extern "C"
