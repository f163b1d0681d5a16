/*
 * libarcb200 - C ABI of the B200-native volumetric-rendering hot path for ArcNerf.
 *
 * Every entry point takes plain pointers and sizes.  Unless a parameter is documented as "host",
 * pointers are DEVICE pointers on the current CUDA device.  `stream` is a cudaStream_t passed as
 * void* (NULL = the legacy default stream).  No entry point allocates device memory, synchronises
 * the device or the stream, or touches the host copy of any tensor: the caller owns every buffer
 * (as in the reference, whose Python wrappers allocate all outputs: arcnerf/ops/volume_func.py:60-63,
 * 100-101) and all work is enqueued on `stream`.
 *
 * Return value: 0 on success, negative on error (ARCB200_ERR_*); arcb200_last_error() returns the
 * text for the calling thread.  A reference-side binding turns a negative return into the
 * RuntimeError its pybind layer raises today.
 *
 * "Replaces" lines cite the reference interface (file:line under /root/reference) each entry point
 * stands in for.
 */
#ifndef ARCB200_H
#define ARCB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ARCB200_OK 0
#define ARCB200_ERR_ARG (-1)
#define ARCB200_ERR_CUDA (-2)
#define ARCB200_ERR_LIMIT (-3)

/* ---------------------------------------------------------------- library ------------------- */
int arcb200_version(void);
const char *arcb200_last_error(void);
/* CUDA devices visible to the library; negative when the CUDA runtime cannot start (no GPU). */
int arcb200_device_count(void);

/* The reference keeps one generator per extension module, `static pcg32 rng{9121}`
 * (arcnerf/ops/include/common.h:21-23), passes it by value to each sampling launch and advances it
 * by 2^32 afterwards (volume_func_kernel.cu:288).  Here the caller holds that state as two
 * integers, passes them to the sampling calls and advances them itself (both are host pointers). */
void arcb200_rng_init(uint64_t seed, uint64_t *state, uint64_t *inc);
void arcb200_rng_advance(uint64_t *state, uint64_t inc, int64_t delta);

/* ---------------------------------------------------------------- arcnerf/ops/volume_func ---- */
/* Replaces _volume_func.aabb_intersection (src/volume_func/volume_func.cpp:61-118 ->
 * volume_func_kernel.cu:74-170; python arcnerf/ops/volume_func.py:49-84).
 * rays_o, rays_d (n_rays,3); aabb (n_v,2,3) = [min xyz; max xyz]; near, far, mask (n_rays,n_v);
 * pts (n_rays,n_v,2,3) or NULL.  mask is one byte per entry (torch.bool storage). */
int arcb200_aabb_intersection(const float *rays_o, const float *rays_d, const float *aabb, float *near,
                              float *far, float *pts, uint8_t *mask, int64_t n_rays, int64_t n_v,
                              void *stream);

/* Replaces _volume_func.check_pts_in_occ_voxel (volume_func.cpp:20-59 -> volume_func_kernel.cu:16-68;
 * python volume_func.py:15-43).  xyz (n,3); bitfield (n_grid^3) bool bytes, flat index x*n^2+y*n+z;
 * aabb (2,3); out (n) bool bytes. */
int arcb200_check_pts_in_occ_voxel(const float *xyz, const uint8_t *bitfield, const float *aabb, int n_grid,
                                   uint8_t *out, int64_t n, void *stream);

/* Replaces _volume_func.sparse_volume_sampling (volume_func.cpp:120-214 -> volume_func_kernel.cu:174-291;
 * python volume_func.py:90-134).  near, far (n_rays,1); aabb (2,3); bitfield (n_grid^3) bool bytes;
 * zvals (n_rays,n_pts) float, mask (n_rays,n_pts) bool bytes: EVERY entry is written (the reference
 * relies on zero-filled outputs; this call does not), valid samples first, then the last valid distance
 * repeated with mask 0, all zeros for a ray without samples.  (rng_state, rng_inc): see arcb200_rng_*. */
int arcb200_sparse_volume_sampling(const float *rays_o, const float *rays_d, const float *near, const float *far,
                                   int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *bitfield,
                                   float near_distance, float *zvals, uint8_t *mask, int64_t n_rays,
                                   uint64_t rng_state, uint64_t rng_inc, void *stream);

/* Replaces _volume_func.tensor_reduce_max (volume_func.cpp:216-275 -> volume_func_kernel.cu:297-337;
 * python volume_func.py:140-167).  uni_tensor (n_group) must be zero-filled by the caller; values are
 * compared through their bit patterns (non-negative floats), as in the reference. */
int arcb200_tensor_reduce_max(const float *full_tensor, const int64_t *group_idx, int n_group, float *uni_tensor,
                              int64_t n, void *stream);

/* ---------------------------------------------------------------- arcnerf/ops/bitfield_func -- */
/* Replaces _bitfield_func.sparse_volume_sampling_bit (bitfield_func.cpp:20-84 ->
 * bitfield_func_kernel.cu:20-137; python bitfield_func.py:17-61).  As arcb200_sparse_volume_sampling but
 * bitfield is (n_grid^3/8) bytes, bit (m%8) of byte (m/8) for Morton code m. */
int arcb200_sparse_volume_sampling_bit(const float *rays_o, const float *rays_d, const float *near,
                                       const float *far, int n_pts, float dt, const float *aabb, int n_grid,
                                       const uint8_t *bitfield, float near_distance, float *zvals, uint8_t *mask,
                                       int64_t n_rays, uint64_t rng_state, uint64_t rng_inc, void *stream);

/* Replaces _bitfield_func.generate_grid_samples (bitfield_func.cpp:86-117 -> bitfield_func_kernel.cu:143-211;
 * python bitfield_func.py:67-88).  positions (n_elements,3) in [0,1)^3, indices (n_elements) uint32
 * (the reference stores them in an int32 tensor). */
int arcb200_generate_grid_samples(const float *density_grid, int ema_step, int64_t n_elements, int n_grid,
                                  float thresh, float *positions, uint32_t *indices, uint64_t rng_state,
                                  uint64_t rng_inc, void *stream);

/* Replaces _bitfield_func.splat_grid_samples (bitfield_func.cpp:119-147 -> bitfield_func_kernel.cu:217-254;
 * python bitfield_func.py:94-106). */
int arcb200_splat_grid_samples(const float *density, const uint32_t *indices, int64_t n_samples,
                               float *density_grid_tmp, void *stream);

/* Replaces _bitfield_func.ema_grid_samples_nerf (bitfield_func.cpp:149-173 -> bitfield_func_kernel.cu:260-296;
 * python bitfield_func.py:112-121). */
int arcb200_ema_grid_samples_nerf(const float *density_grid_tmp, int64_t n_elements, float decay,
                                  float *density_grid, void *stream);

/* Replaces _bitfield_func.update_bitfield (bitfield_func.cpp:175-204 -> bitfield_func_kernel.cu:302-346;
 * python bitfield_func.py:127-139) and _multivol_func.update_bitfield_multivol (multivol_func.cpp:170-207 ->
 * multivol_func_kernel.cu:278-330; python multivol_func.py:131-143): n_cells = n_grid^3 for the former,
 * n_grid^3 * (n_cascade or n_cascade-1) for the latter.  The threshold is min(opa_thres, mean);
 * the mean comes from mean_host, or from device memory when mean_dev is not NULL (no host round trip). */
int arcb200_update_bitfield(const float *density_grid, float mean_host, const float *mean_dev, uint8_t *bitfield,
                            float opa_thres, int64_t n_cells, void *stream);

/* Replaces _bitfield_func.count_bitfield (bitfield_func.cpp:206-232 -> bitfield_func_kernel.cu:352-389;
 * python bitfield_func.py:145-159).  counter[0] += count.  faithful != 0 reproduces the reference's
 * `byte && (1<<j)` test (bitfield_func_kernel.cu:365: 8 per non-zero byte); faithful == 0 counts set bits. */
int arcb200_count_bitfield(const uint8_t *bitfield, float *counter, int64_t n_cells, int faithful, void *stream);

/* ---------------------------------------------------------------- arcnerf/ops/multivol_func -- */
/* Replaces _multivol_func.sparse_sampling_in_multivol_bitfield (multivol_func.cpp:20-112 ->
 * multivol_func_kernel.cu:19-171; python multivol_func.py:17-89).  min_aabb, aabb (2,3). */
int arcb200_sparse_sampling_in_multivol_bitfield(const float *rays_o, const float *rays_d, const float *near,
                                                 const float *far, int n_pts, float cone_angle, float min_step,
                                                 float max_step, const float *min_aabb, const float *aabb,
                                                 int n_grid, int n_cascade, const uint8_t *bitfield,
                                                 float near_distance, int inclusive, float *zvals, uint8_t *mask,
                                                 int64_t n_rays, uint64_t rng_state, uint64_t rng_inc,
                                                 void *stream);

/* Replaces _multivol_func.generate_grid_samples_multivol (multivol_func.cpp:114-168 ->
 * multivol_func_kernel.cu:177-271; python multivol_func.py:95-125). */
int arcb200_generate_grid_samples_multivol(const float *density_grid, int ema_step, int64_t n_elements,
                                           const float *aabb, int n_cascade, int n_grid, float thresh,
                                           int inclusive, float *positions, uint32_t *indices,
                                           uint64_t rng_state, uint64_t rng_inc, void *stream);

/* ---------------------------------------------------------------- packed (ragged) sampling --- */
/* The same march as arcb200_sparse_volume_sampling[_bit] (bit = 0 / 1), delivered as a packed sample list
 * instead of the padded (n_rays,n_pts) matrix that FgModel re-compacts with boolean indexing
 * (arcnerf/models/fg_model.py:251-318).  Outputs: counts (n_rays) int32; offsets (n_rays+1) int32 exclusive
 * prefix of counts; meta[0] = total samples, meta[1] = largest count; t (capacity) sample distances;
 * xyz (capacity,3) world positions or NULL; ray_id (capacity) or NULL.  Samples of one ray are contiguous,
 * rays in index order; samples beyond `capacity` are dropped (offsets still describe the full list).
 * scratch: arcb200_march_packed_scratch_bytes(n_rays, n_pts) bytes of device memory. */
int64_t arcb200_march_packed_scratch_bytes(int64_t n_rays, int n_pts);
int arcb200_march_packed(int bit, const float *rays_o, const float *rays_d, const float *near, const float *far,
                         int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *field,
                         float near_distance, int64_t n_rays, uint64_t rng_state, uint64_t rng_inc,
                         void *scratch, int32_t *counts, int32_t *offsets, int32_t *meta, int64_t capacity,
                         float *t_out, float *xyz_out, int32_t *ray_id_out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ARCB200_H */
