// TEST INFRASTRUCTURE ONLY -- never linked into or called from the product path.
//
// The REFERENCE's own arcnerf/ops CUDA kernels compiled for sm_100 exactly as nvcc compiles them for the
// reference's extension (default flags: FMA contraction ON), from the sources where they lie under
// /root/reference (see oracle/Makefile target `ref_dev`; nothing is copied).  One translation unit per reference
// module (-DARCREF_MODULE=1 volume_func, 2 bitfield_func, 3 multivol_func): each module keeps its own
// `static pcg32 rng{9121}` (arcnerf/ops/include/common.h:23), as in the reference's three pybind modules.
// Used on the GPU box to (a) report how often the reference's device build differs from its host build (the pinned
// canonical semantics) and (b) time the reference's kernels on the same B200.
#include <cstdint>
#include <torch/extension.h>

#if ARCREF_MODULE == 1
#include "volume_func/volume_func_kernel.cu"      // -I$(REF)/arcnerf/ops/src
#elif ARCREF_MODULE == 2
#include "bitfield_func/bitfield_func_kernel.cu"
#elif ARCREF_MODULE == 3
#include "multivol_func/multivol_func_kernel.cu"
#else
#error "ARCREF_MODULE must be 1, 2 or 3"
#endif

using torch::Tensor;

extern "C" {

// re-seed this module's generator (a fresh process of the reference starts at pcg32{9121})
void arcrefdev_reset_rng() { rng = pcg32{9121}; }

#if ARCREF_MODULE == 1
// volume_func.cpp:95-130 -> volume_func_kernel.cu:134-170
void arcrefdev_aabb_intersection(int n_rays, int n_v, const float *o, const float *d, const float *aabb, float *near,
                                 float *far, float *pts, uint8_t *mask) {
    aabb_intersection_cuda(Tensor(o, n_rays, 3), Tensor(d, n_rays, 3), Tensor(aabb, n_v, 2, 3), Tensor(near, n_rays, n_v),
                           Tensor(far, n_rays, n_v), Tensor(pts, n_rays, n_v, 2, 3), Tensor(mask, n_rays, n_v));
}
// volume_func.cpp:150-215 -> volume_func_kernel.cu:253-291
void arcrefdev_sparse_volume_sampling(int n_rays, const float *o, const float *d, const float *near, const float *far,
                                      int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *volume,
                                      float near_distance, float *zvals, uint8_t *mask) {
    sparse_volume_sampling_cuda(Tensor(o, n_rays, 3), Tensor(d, n_rays, 3), Tensor(near, n_rays, 1), Tensor(far, n_rays, 1), n_pts,
                                dt, Tensor(aabb, 2, 3), n_grid, Tensor(volume, n_grid, n_grid, n_grid), near_distance,
                                Tensor(zvals, n_rays, n_pts), Tensor(mask, n_rays, n_pts));
}
#elif ARCREF_MODULE == 2
// bitfield_func.cpp -> bitfield_func_kernel.cu:99-137
void arcrefdev_sparse_volume_sampling_bit(int n_rays, const float *o, const float *d, const float *near, const float *far,
                                          int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *bits,
                                          float near_distance, float *zvals, uint8_t *mask) {
    sparse_volume_sampling_bit_cuda(Tensor(o, n_rays, 3), Tensor(d, n_rays, 3), Tensor(near, n_rays, 1), Tensor(far, n_rays, 1),
                                    n_pts, dt, Tensor(aabb, 2, 3), n_grid, Tensor(bits, (int64_t)n_grid * n_grid * n_grid / 8),
                                    near_distance, Tensor(zvals, n_rays, n_pts), Tensor(mask, n_rays, n_pts));
}
#else
// multivol_func.cpp -> multivol_func_kernel.cu:128-171
void arcrefdev_sparse_sampling_in_multivol_bitfield(int n_rays, const float *o, const float *d, const float *near,
                                                    const float *far, int n_pts, float cone_angle, float min_step,
                                                    float max_step, const float *min_aabb, const float *aabb, int n_grid,
                                                    int n_cascade, const uint8_t *bits, int64_t n_bytes, float near_distance,
                                                    int inclusive, float *zvals, uint8_t *mask) {
    sparse_sampling_in_multivol_bitfield_cuda(Tensor(o, n_rays, 3), Tensor(d, n_rays, 3), Tensor(near, n_rays, 1),
                                              Tensor(far, n_rays, 1), n_pts, cone_angle, min_step, max_step, Tensor(min_aabb, 2, 3),
                                              Tensor(aabb, 2, 3), n_grid, n_cascade, Tensor(bits, n_bytes), near_distance,
                                              inclusive != 0, Tensor(zvals, n_rays, n_pts), Tensor(mask, n_rays, n_pts));
}
#endif

}  // extern "C"
