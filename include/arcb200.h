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
 * xyz (capacity,3) world positions or NULL; ray_id (capacity) or NULL; unit (capacity,3) positions normalised to the
 * box ((p-min)/(max-min), the encoder input, hashgrid_encoder.py:183) or NULL.  Samples of one ray are contiguous,
 * rays in index order; samples beyond `capacity` are dropped (offsets still describe the full list).
 * scratch: arcb200_march_packed_scratch_bytes(n_rays, n_pts) bytes of device memory. */
int64_t arcb200_march_packed_scratch_bytes(int64_t n_rays, int n_pts);
int arcb200_march_packed(int bit, const float *rays_o, const float *rays_d, const float *near, const float *far,
                         int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *field,
                         float near_distance, int64_t n_rays, uint64_t rng_state, uint64_t rng_inc,
                         void *scratch, int32_t *counts, int32_t *offsets, int32_t *meta, int64_t capacity,
                         float *t_out, float *xyz_out, int32_t *ray_id_out, float *unit_out, void *stream);


/* No counterpart in the reference.  arcb200_march_packed for hosts that cannot bring scratch memory: a sampler
 * called from another engine's per-frame callback (a cudaMalloc there would synchronise the whole device and stall
 * the engine's other streams), or a plug-in inside a host that owns all device memory through its own arena and
 * passes in nothing but the call's input and output buffers.  The scratch is a fixed device array inside the
 * library: 16 MiB + 4 KiB of device memory on every GPU that loads the library, until the process exits.  Same
 * arguments (minus `scratch`) and same outputs as arcb200_march_packed.  Limits: n_rays * n_pts <= 4 194 304
 * (arcb200_march_packed_static_max_slots) and n_rays <= 262 144; larger requests return ARCB200_ERR_LIMIT.
 * The call allocates nothing.  Its first use on a device looks up the array's device address; make that first call
 * before any stream capture begins.  All calls on a device share the array (stale contents are harmless, nothing
 * in it needs clearing), so calls whose kernels could overlap on the GPU race: keep them ordered. */
int64_t arcb200_march_packed_static_max_slots(void);
int arcb200_march_packed_static(int bit, const float *rays_o, const float *rays_d, const float *near, const float *far,
                                int n_pts, float dt, const float *aabb, int n_grid, const uint8_t *field,
                                float near_distance, int64_t n_rays, uint64_t rng_state, uint64_t rng_inc,
                                int32_t *counts, int32_t *offsets, int32_t *meta, int64_t capacity, float *t_out,
                                float *xyz_out, int32_t *ray_id_out, float *unit_out, void *stream);

/* Row counts on the device: the entry points below that take `n_dev` (device int32 or NULL) process only
 * min(n, *n_dev) rows, so the sample count produced by arcb200_march_packed (meta[0]) never has to travel to
 * the host between the sampler and the field kernels. */

/* ---------------------------------------------------------------- tiny-cuda-nn binding: HashGrid ------
 * ArcNerf reaches these through tinycudann.Encoding / tinycudann.Network (HashGridEmbedder
 * hashgrid_encoder.py:102-113,182-184; SHEmbedder sh_encoder.py:59-65,116-118; FusedMLP*
 * tcnn_fusedmlp_module.py:66-77,162-173).  "tcnn" below = /root/reference/include/tiny-cuda-nn. */

/* Level geometry on the host (tcnn include/tiny-cuda-nn/encodings/grid.h:195-204,982-1012).  offsets
 * (n_levels+1, in table entries), scales, resolutions (n_levels) are HOST arrays.  Returns the number of
 * table entries (parameters = entries * n_features) or a negative error. */
int64_t arcb200_grid_geometry(int n_levels, int log2_hashmap_size, int base_resolution, float per_level_scale,
                              uint32_t *offsets, float *scales, uint32_t *resolutions);

/* fp32 master parameters -> fp16 working copy (tcnn bindings/torch/tinycudann/modules.py:156 does this cast
 * on every forward; callers of this library do it only after the parameters changed). */
int arcb200_cast_f32_to_f16(const float *src, void *dst_half, int64_t n, void *stream);

/* Replaces tcnn Module.fwd for a HashGrid encoding (bindings/torch/tinycudann/bindings.cpp:78-109 ->
 * grid.h:220-392,1034-1105).  xyz (n,3) fp32 in [0,1]; table: fp16, flat [level][entry][feature];
 * out (n, n_levels*n_features) fp16 row-major; dy_dx (n, n_levels*n_features, 3) fp32 or NULL.
 * offsets/scales/resolutions: HOST arrays from arcb200_grid_geometry. */
int arcb200_grid_fwd(const float *xyz, const void *table_half, int n_levels, int n_features, const uint32_t *offsets,
                     const float *scales, const uint32_t *resolutions, void *out_half, float *dy_dx, int64_t n,
                     const int32_t *n_dev, void *stream);

/* Replaces the parameter-gradient half of tcnn Module.bwd for a HashGrid (bindings.cpp:111-170 ->
 * kernel_grid_backward, grid.h:394-517).  grad_table (fp32, table layout) += ...; the caller zeroes it.
 * dL_dy (n, L*F) fp16 (dy_is_half) or fp32, multiplied by gscale on the fly. */
int arcb200_grid_bwd(const float *xyz, const void *dL_dy, int dy_is_half, float gscale, int n_levels, int n_features,
                     const uint32_t *offsets, const float *scales, const uint32_t *resolutions, float *grad_table,
                     int64_t n, const int32_t *n_dev, void *stream);

/* Replaces the input-gradient half of tcnn Module.bwd (kernel_grid_backward_input, grid.h:549-576):
 * dL_dx (n,3) = dL_dy . dy_dx. */
int arcb200_grid_bwd_input(const void *dL_dy, int dy_is_half, float gscale, const float *dy_dx, int n_levels,
                           int n_features, float *dL_dx, int64_t n, void *stream);

/* Replaces tcnn Module.bwd_bwd_input for a HashGrid (bindings.cpp:172-241 -> grid.h:578-895): double backward
 * of the input gradient (eikonal loss).  v = dL2/d(dL/dx) (n,3).  Optional outputs: grad_table (accumulated),
 * d_dLdy (n, L*F) fp32, d_x (n,3) fp32 (zeroed by the call). */
int arcb200_grid_bwd_bwd_input(const float *xyz, const void *table_half, const void *dL_dy, int dy_is_half,
                               float gscale, const float *v, int n_levels, int n_features, const uint32_t *offsets,
                               const float *scales, const uint32_t *resolutions, float *grad_table, float *d_dLdy,
                               float *d_x, int64_t n, void *stream);

/* ---------------------------------------------------------------- tiny-cuda-nn binding: SH -------------
 * Replaces tcnn Module.fwd for a SphericalHarmonics encoding (kernel_sh, encodings/spherical_harmonics.h:46-152)
 * as SHEmbedder.forward calls it (sh_encoder.py:104-123).  dirs (n_dirs,3) fp32; index (n) int32 or NULL
 * (per-sample row of dirs, so that per-ray directions need not be repeated per sample); remap != 0: input is a
 * direction in [-1,1] and the embedder's (d+1)/2 map is applied first; out: n rows of out_stride elements, fp16
 * (out_is_half) or fp32, first degree^2 written.  degree 1..5. */
int arcb200_sh_fwd(const float *dirs, const int32_t *index, int degree, int remap, void *out, int out_is_half,
                   int64_t out_stride, int64_t n, const int32_t *n_dev, void *stream);

/* ---------------------------------------------------------------- tiny-cuda-nn binding: FullyFusedMLP --
 * Network: in_dim -> width (x n_hidden, hidden_act) -> out_dim (out_act), no biases; flat fp32 parameters in
 * tcnn order (fully_fused_mlp.cu:658-671): W_0 (width x in_pad), W_h (width x width) ..., W_last
 * (out_pad x width), row-major, in_pad/out_pad = dims rounded up to 16; input pad columns are ONES
 * (encodings/identity.h:63-64).  Activation codes: 0 none, 1 ReLU, 2 sigmoid, 3 softplus (beta 10), 4 exp. */
int arcb200_mlp_n_params(int in_dim, int width, int n_hidden, int out_dim);
int arcb200_mlp_blob_bytes(int in_dim, int width, int n_hidden, int out_dim);
/* fp32 parameters -> fp16 tensor-core operand tiles (blob, 16-byte aligned device buffer).  Call after every
 * parameter update; the forward/backward take the blob. */
int arcb200_mlp_pack_weights(const float *params, void *blob, int in_dim, int width, int n_hidden, int out_dim,
                             void *stream);
/* Replaces tcnn Module.fwd for a FullyFusedMLP (bindings.cpp:78-109 -> fully_fused_mlp.cu:499-700).
 * X (n,in_dim) fp32 or fp16, leading dimension ldx; Y (n,out_dim) fp32 or fp16, leading dimension ldy. */
int arcb200_mlp_fwd(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                    int out_dim, int hidden_act, int out_act, void *Y, int y_is_half, int64_t ldy, int64_t n,
                    const int32_t *n_dev, void *stream);
/* Forward with extras.  act_save (optional; arcb200_mlp_act_save_bytes(n, width, n_hidden) bytes, 16-byte aligned)
 * receives the hidden activations for arcb200_mlp_bwd_ex, as tcnn keeps them in training mode
 * (fully_fused_mlp.cu:702-760, `forward.hidden`); exp0_out (optional, n floats) receives exp(output channel 0) taken
 * from the value as stored in Y: FusedMLPGeoNet's TruncExp density (tcnn_fusedmlp_module.py:99-111,
 * arcnerf/ops/trunc_exp.py:10-24) without a separate pass. */
int64_t arcb200_mlp_act_save_bytes(int64_t n, int width, int n_hidden);
int arcb200_mlp_fwd_ex(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, void *Y, int y_is_half, int64_t ldy, void *act_save,
                       float *exp0_out, int64_t n, const int32_t *n_dev, void *stream);
/* arcb200_mlp_fwd_ex with an fp16 output, plus the degree-4 SH encoding of a per-row direction (row i: dirs[index[i]] or
 * dirs[i]) in output columns [sh_col, sh_col+16): FusedMLPGeoNet -> SHEmbedder (sh_encoder.py:104-123) -> the `[feat, sh]`
 * concatenation of fuse_radiance_inputs (encoder_mlp_network.py:93-118) in one launch.  zero_scalar (optional): one
 * float cleared by the launch (the loss accumulator of the step). */
int arcb200_mlp_fwd_sh(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, void *Y, int64_t ldy, void *act_save, float *exp0_out,
                       const float *sh_dirs, const int32_t *sh_index, int sh_col, int sh_remap, float *zero_scalar,
                       int64_t n, const int32_t *n_dev, void *stream);
/* Backward with extras.  act_saved: buffer written by arcb200_mlp_fwd_ex for the same X / weights / n - the forward is
 * then not re-computed, and y_saved (n,out_dim; fp16 if ys_is_half else fp32; leading dimension ldys) must hold the
 * network outputs whenever out_act != 0.  g0/x0: extra gradient on output channel 0 that arrives through exp():
 * dY[:,0] += g0[row] * exp(clamp(x0[row*ldx0], +-clip))  (TruncExp backward, trunc_exp.py:26-37, fused). */
int arcb200_mlp_bwd_ex(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                       int out_dim, int hidden_act, int out_act, const float *dY, int64_t ldy, float loss_scale,
                       void *dX, int dx_is_half, int64_t ldgx, float *dW, const void *act_saved, const void *y_saved,
                       int ys_is_half, int64_t ldys, const float *g0, const void *x0, int x0_is_half, int64_t ldx0,
                       float clip, int64_t n, const int32_t *n_dev, void *stream);

/* Replaces tcnn Module.bwd for a FullyFusedMLP (bindings.cpp:111-170 -> fully_fused_mlp.cu:150-259,702-857).
 * dY (n,out_dim) fp32; dX (n,in_dim) fp32/fp16 or NULL; dW flat fp32 in the parameter layout, ACCUMULATED.
 * loss_scale: internal fp16 gradient scale (128 in the reference, modules.py:141).  width must be 64 or 128. */
int arcb200_mlp_bwd(const void *X, int x_is_half, int64_t ldx, const void *blob, int in_dim, int width, int n_hidden,
                    int out_dim, int hidden_act, int out_act, const float *dY, int64_t ldy, float loss_scale,
                    void *dX, int dx_is_half, int64_t ldgx, float *dW, int64_t n, const int32_t *n_dev, void *stream);

/* ---------------------------------------------------------------- arcnerf/render/ray_helper ------------
 * Rays are segments of flat per-sample arrays: offsets (n_rays+1) int32, or NULL for dense (n_rays, dense_p)
 * tensors.  capacity = length of the per-sample arrays (packed samples beyond it were dropped by the sampler).
 * pmax_dev: device int32, the row length P_max the equivalent padded call of the reference would use
 * (meta+1 of arcb200_march_packed), or NULL. */

/* Replaces ray_marching + alpha_to_weights (arcnerf/render/ray_helper.py:476-620).  geo = sigma, or alpha when
 * is_alpha; rad (N,3) or NULL; noise (N) or NULL (already scaled by noise_std); bkg (n_rays,3) if bkg_per_ray else
 * (1,3), or NULL.  Outputs: rgb (n_rays,3), depth, acc (n_rays); per-sample weights / trans_shift / alpha or NULL
 * (the closing sample of a segment gets 0). */
int arcb200_composite_fwd(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev,
                          const float *geo, int is_alpha, const float *rad, const float *t, const float *noise,
                          int add_inf_z, int white_bkg, const float *bkg, int bkg_per_ray, float *rgb, float *depth,
                          float *acc, float *w_out, float *trans_out, float *alpha_out, int64_t n_rays,
                          void *stream);
/* The same with one more output: t_last_out (n_rays) or NULL, the transmittance that weights the background - the reference's
 * trans_shift[:, -1], which FullModel.blend_bkg_rgb multiplies the background model's colour and depth with
 * (arcnerf/models/full_model.py:296-330). */
int arcb200_composite_fwd_ex(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev, const float *geo,
                             int is_alpha, const float *rad, const float *t, const float *noise, int add_inf_z, int white_bkg,
                             const float *bkg, int bkg_per_ray, float *rgb, float *depth, float *acc, float *w_out,
                             float *trans_out, float *alpha_out, float *t_last_out, int64_t n_rays, void *stream);
/* Gradient of the above (the reference uses autograd through ~15 eager ops): d_geo (N), d_rad (N,3) or NULL from
 * g_rgb (n_rays,3), g_depth, g_acc (n_rays), g_w (N), each optional.  max_samples_per_ray <= 4096. */
int arcb200_composite_bwd(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev,
                          const float *geo, int is_alpha, const float *rad, const float *t, const float *noise,
                          int add_inf_z, int white_bkg, const float *bkg, int bkg_per_ray, const float *g_rgb,
                          const float *g_depth, const float *g_acc, const float *g_w, float *d_geo, float *d_rad,
                          int max_samples_per_ray, int64_t n_rays, void *stream);

/* Fused training pass: arcb200_composite_fwd (density input) -> arcb200_huber_loss -> arcb200_composite_bwd in one
 * kernel; same arithmetic as the three calls.  rgb and g_rgb_out (n_rays,3) are optional outputs; step_counter (optional
 * device int32) is incremented by one (the optimiser step count read through arcb200_adam_step's step_dev). */
int arcb200_composite_huber_train(const int32_t *offsets, int dense_p, int64_t capacity, const int32_t *pmax_dev,
                                  const float *geo, const float *rad, const float *t, const float *noise, int add_inf_z,
                                  int white_bkg, const float *bkg, int bkg_per_ray, const float *target, float delta,
                                  float weight, float *rgb, float *g_rgb_out, float *loss_sum, float *d_geo, float *d_rad,
                                  int32_t *step_counter, int max_samples_per_ray, int64_t n_rays, void *stream);
/* Replaces sdf_to_alpha (arcnerf/models/neus_model.py:242-265).  sdf, slope, alpha (n_rays,p); zvals (n_rays,p+1);
 * scale s from s_dev (device float) when not NULL, else s_host. */
int arcb200_sdf_to_alpha(const float *sdf, const float *zvals, const float *slope, const float *s_dev, float s_host,
                         int clip, float *alpha, int64_t n_rays, int p, void *stream);
/* Its gradient; d_s (1 float, optional) is accumulated atomically (zero it first). */
int arcb200_sdf_to_alpha_bwd(const float *sdf, const float *zvals, const float *slope, const float *s_dev,
                             float s_host, int clip, const float *g_alpha, float *d_sdf, float *d_slope, float *d_s,
                             int64_t n_rays, int p, void *stream);

/* Replaces the searchsorted/gather/interpolate core of sample_cdf (arcnerf/render/ray_helper.py:434-473).
 * bins, cdf (n_rays,n_bins); u (n_rays,n_sample) if u_per_ray else (n_sample); out (n_rays,n_sample) unsorted. */
int arcb200_sample_cdf(const float *bins, const float *cdf, const float *u, int u_per_ray, float eps, float *out,
                       int64_t n_rays, int n_bins, int n_sample, void *stream);

/* ---------------------------------------------------------------- arcnerf/ops/trunc_exp --------------
 * Replaces TruncExpOps.forward / .backward (arcnerf/ops/trunc_exp.py:10-37): y = exp(x), gx = g * exp(clamp(x, +-clip)).
 * x is read with stride ldx (fp16 or fp32) so that a column of a wider activation row can be used in place. */
int arcb200_trunc_exp_fwd(const void *x, int x_is_half, int64_t ldx, float *y, int64_t n, const int32_t *n_dev,
                          void *stream);
int arcb200_trunc_exp_bwd(const void *x, int x_is_half, int64_t ldx, const float *g, float clip, float *gx, int64_t n,
                          const int32_t *n_dev, void *stream);
/* Gradient entering FusedMLPGeoNet's network from its two consumers (handle_output_combine,
 * tcnn_fusedmlp_module.py:99-111, differentiated): dY (n,width) fp32 = d_feat (optional) with channel 0 increased by
 * g_sigma * exp(clamp(x0)) where x0 is channel 0 of the stored head output. */
/* unit (n,3) = (pts - aabb_min) / (aabb_max - aabb_min), aabb (2,3) = min; max: the normalisation in front of the hash
 * grid (hashgrid_encoder.py:150-160).  unit may alias pts. */
int arcb200_world_to_unit(const float *pts, const float *aabb, float *unit, int64_t n, void *stream);
int arcb200_geo_head_grad(const void *x0, int x_is_half, int64_t ldx, const float *g_sigma, float clip,
                          const void *d_feat, int f_is_half, int64_t ldf, int width, float *dY, int64_t n,
                          const int32_t *n_dev, void *stream);

/* ---------------------------------------------------------------- training-step helpers ----------------
 * torch.optim.Adam arithmetic (common/trainer/optimizer.py -> torch) in one pass that also writes the fp16
 * working copy (p_half, optional) and clears the gradient (zero_grad).  step >= 1.
 * step_dev (device int32 or NULL) overrides `step` so that the launch can be replayed from a CUDA graph;
 * arcb200_counter_add advances such a counter on the stream. */
int arcb200_adam_step(float *p, float *m, float *v, float *g, void *p_half, int64_t n, int step, float lr,
                      float beta1, float beta2, float eps, float weight_decay, float grad_scale, int zero_grad,
                      const int32_t *step_dev, void *stream);
int arcb200_counter_add(int32_t *counter, int value, void *stream);
/* dst[0..n) = value on the stream (clears accumulators such as the loss; no counterpart in the reference, where torch does it). */
int arcb200_fill_f32(float *dst, float value, int64_t n, void *stream);
/* x[0..n) *= a  (opacity = density * dt in the occupancy refresh, arcnerf/models/base_3d_model.py:368-389) */
int arcb200_scale_f32(float *x, float a, int64_t n, void *stream);
/* Adam + fp16 operand re-packing of n_mlps (<= 4) small networks in ONE launch (per network: arcb200_adam_step followed
 * by arcb200_mlp_pack_weights).  Reference: torch.optim.Adam on tcnn's `params` (common/trainer/optimizer.py) and tcnn's
 * fp32->fp16 parameter cast at the next forward (tinycudann/modules.py:156).  shapes: host int[n_mlps*4] = in_dim,
 * width, n_hidden, out_dim; p, m, v, blobs: host arrays of n_mlps device pointers; grads: host array of n_mlps*world
 * device pointers, grads[j*world + w] = gradient source w of network j (world 1 = the local gradient; several = every
 * rank's copy through peer memory), summed in index order and scaled by grad_scale; zero_grad clears source 0. */
int arcb200_mlp_adam_pack(int n_mlps, const int *shapes, float *const *p, float *const *m, float *const *v,
                          void *const *grads, int world, void *const *blobs, int step, float lr, float beta1,
                          float beta2, float eps, float weight_decay, float grad_scale, int zero_grad,
                          const int32_t *step_dev, void *stream);
/* ImgLoss with HuberLoss (arcnerf/loss/img_loss.py:48-98): loss_sum[0] += weight * mean(huber), g_rgb = gradient. */
int arcb200_huber_loss(const float *rgb, const float *target, float delta, float weight, float *g_rgb,
                       float *loss_sum, int64_t n_rays, void *stream);

/* ---------------------------------------------------------------- frequency encoding (a11) ----
 * FreqEmbedder.forward (arcnerf/models/base_modules/encoding/freq_encoder.py:66-88): out = [x | sin(x f_0) | cos(x f_0) |
 * ...], bands given by the caller as a HOST array (2^k or the reference's linear rule), and its input gradient. */
int arcb200_freq_embed_fwd(const float *x, int d, const float *freqs, int n_freqs, int include_input, float *out,
                           int64_t ld_out, int64_t n, void *stream);
int arcb200_freq_embed_bwd(const float *x, int d, const float *freqs, int n_freqs, int include_input, const float *g_out,
                           int64_t ld_g, float *g_x, int64_t n, void *stream);

/* ---------------------------------------------------------------- occupancy refresh of the volume bound (a6) ----
 * VolumeBound.optimize (obj_bound/volume_bound.py:160-212) + Volume.update_opafield_by_voxel_idx /
 * update_bitfield_by_opafield (arcnerf/geometry/volume.py:983-1018) without torch.randperm / where / unique / host
 * synchronisation.  candidates: n_uniform distinct voxels (keyed permutation of the flat index = sampling without
 * replacement) + the first n_occ_max occupied voxels in flat order, one pcg32-jittered point each (idx -1 = unused
 * slot); the caller evaluates opacity = density * dt at pts; apply: per-voxel max, opa = opa >= 0 ? max(opa*ema, new)
 * : opa for the sampled voxels only, volume byte = opa >= min(mean(max(opa,0)), threshold).  scratch:
 * arcb200_volume_refresh_scratch_bytes(n_grid), shared by both calls. */
int64_t arcb200_volume_refresh_scratch_bytes(int n_grid);
int arcb200_volume_refresh_candidates(const uint8_t *volume, int n_grid, int64_t n_uniform, int64_t n_occ_max,
                                      uint32_t key, uint64_t rng_state, uint64_t rng_inc, const float *aabb, void *scratch,
                                      int32_t *idx_out, float *pts_out, void *stream);
int arcb200_volume_refresh_apply(const int32_t *idx, const float *opacity, int64_t n_cand, float ema, float threshold,
                                 int n_grid, void *scratch, float *opafield, uint8_t *volume, void *stream);

/* ---------------------------------------------------------------- depth samples around the march (a15) ----
 * arcnerf/geometry/ray.py:180-255 sphere_ray_intersection (near/far/pts/mask against n_r concentric spheres; used by
 * get_near_far_from_rays, ray_helper.py:175-228, and the NeRF++ shells, :318-358). */
int arcb200_sphere_ray_intersection(const float *rays_o, const float *rays_d, const float *radius, const float *origin,
                                    float *near, float *far, float *pts, uint8_t *mask, int64_t n_rays, int64_t n_r,
                                    void *stream);
/* ray_helper.py:231-263 get_zvals_from_near_far (+ :361-378 perturb_interval when u, the caller's uniforms, is given). */
int arcb200_zvals_from_near_far(const float *near, const float *far, int n_pts, int inclusive, int inverse_linear,
                                const float *u, float *zvals, int64_t n_rays, void *stream);
/* ray_helper.py:266-315 get_zvals_from_near_far_fix_step (+ :381-407 perturb_interval_with_mask when u is given). */
int arcb200_zvals_fix_step(const float *near, const float *far, float fix_t, int n_pts, int inclusive, const float *u,
                           float *zvals, uint8_t *mask, int64_t n_rays, void *stream);
/* arcnerf/geometry/ray.py:11-30 get_ray_points_by_zvals. */
int arcb200_ray_points(const float *rays_o, const float *rays_d, const float *zvals, float *pts, int64_t n_rays, int n_pts,
                       void *stream);
/* ray_marching's `assert torch.all(deltas >= 0)` (ray_helper.py:532-534) as a device-side flag: flag[0] |= 1 when a row of zvals
 * (n_rays, p) decreases by 1e-5 or more or holds a NaN.  The caller clears the flag and reads it without stalling the stream. */
int arcb200_check_increasing(const float *zvals, int64_t n_rays, int p, int32_t *flag, void *stream);
/* torch.sort(torch.cat([zvals, new_zvals], -1), -1) of NeRF/NeuS upsample_zvals (nerf_model.py:112, neus_model.py:171):
 * ascending sort of every row, n <= 8192, out may alias in. */
int arcb200_sort_rows(const float *in, float *out, int64_t n_rows, int n, void *stream);

/* ---------------------------------------------------------------- data-parallel step over peer memory ----
 * Replaces, for one node, what the reference gets from torch DDP + torch.optim.Adam: the bucketed all-reduce of the
 * gradients hooked into backward (common/trainer/basic_trainer.py:192-202, set_model_to_ddp) followed by the same
 * Adam step on every rank (common/trainer/optimizer.py).  One process per GPU; every rank allocates an arena with
 * arcb200_peer_alloc, sends the 64-byte handle to the others (any host channel), and maps theirs with
 * arcb200_peer_open.  Gradients, the fp16 working copy of the sharded parameters and the barrier flags live in the
 * arena.  Per step: barrier, arcb200_adam_step_peer per parameter vector, barrier, then clear the own gradients. */
int arcb200_peer_alloc(int64_t bytes, void **ptr_out, unsigned char *handle_out /* 64 bytes */);
int arcb200_peer_open(const unsigned char *handle /* 64 bytes */, void **ptr_out);
int arcb200_peer_close(void *ptr);
int arcb200_peer_free(void *ptr);
/* Flag barrier between ranks on `stream`: flags[w] = rank w's flag array (16 uint64, zeroed) as mapped in this
 * process; epoch = private device uint64 (starts at 0); status = device int32 set to 1 if a wait exceeded the
 * timeout.  timeout_ms > 0: fatal after that many ms (status is raised, then the kernel traps - a step built on a peer's
 * unfinished gradient must not be applied, the job ends with a launch failure at the next synchronisation);
 * timeout_ms == 0: default = ARCB200_PEER_TIMEOUT_MS or 600 000 ms, fatal; timeout_ms < 0: |timeout_ms| ms, status only. */
int arcb200_peer_barrier(void *const *flags, int world, int rank, void *epoch, int32_t *status, int timeout_ms,
                         void *stream);
/* Adam over elements [lo, hi) (lo % 4 == 0) of a flat parameter vector on the gradient summed over grads[0..world)
 * in rank order (peer loads), times grad_scale; the updated fp16 copy of the range is stored into half_out[0..world)
 * (peer stores; half_out may be NULL).  p, m, v are this rank's full-length vectors.  Gradients are not cleared.
 * Same arithmetic as arcb200_adam_step.  world <= 16. */
int arcb200_adam_step_peer(const void *const *grads, void *const *half_out, int world, int64_t lo, int64_t hi,
                           float *p, float *m, float *v, int step, float lr, float beta1, float beta2, float eps,
                           float weight_decay, float grad_scale, const int32_t *step_dev, void *stream);
/* Half-precision gradient exchange (what travels over NVLink is a loss-scaled fp16 copy, as precise as the fp16 gradients
 * tcnn hands to DDP in the reference, tinycudann/modules.py:93-96,141): arcb200_grad_to_half makes the copy of everything
 * outside the slice this rank owns (and clears the fp32 gradient there); arcb200_adam_step_peer16 is
 * arcb200_adam_step_peer on  g_own (fp32, cleared over the range)  +  sum over the other ranks of grads16[w] / scale. */
int arcb200_grad_to_half(float *g32, void *g16, int64_t n, int64_t keep_lo, int64_t keep_hi, float scale, void *stream);
int arcb200_adam_step_peer16(float *g_own, const void *const *grads16, void *const *half_out, int world, int rank,
                             int64_t lo, int64_t hi, float *p, float *m, float *v, float scale, int step, float lr,
                             float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                             const int32_t *step_dev, void *stream);
/* The same exchange carried out by the NVSwitch (NVLS): mc_grad16 and mc_half_out are MULTICAST addresses (one mapping bound to
 * the same offset of every rank's arena; arcnerf_b200/peer.py SymmArena obtains it from torch's symmetric memory) of the fp16
 * gradient copies - made by arcb200_grad_to_half over the whole vector (keep_lo = keep_hi = 0) - and of the fp16 working table.
 * One multimem.ld_reduce (fp32 accumulation in the fabric) replaces the world-1 peer loads, one multimem.st the world peer
 * stores.  lo, hi multiples of 8.  Replaces the same DDP all-reduce + Adam (common/trainer/basic_trainer.py:192-202). */
int arcb200_adam_step_mc16(const void *mc_grad16, void *mc_half_out, int64_t lo, int64_t hi, float *p, float *m, float *v,
                           float scale, int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                           float grad_scale, const int32_t *step_dev, void *stream);

/* ---- fused field forward (csrc/field_fused.cuh) -----------------------------------------------------------------------
 * The NGP foreground model's per-sample network in one launch: hash-grid encode (16 levels x 2 features) -> geometry head
 * 32-64-16 (TruncExp density on channel 0) -> [feature | SH degree 4 of the view direction] -> radiance head 32-64-64-3
 * (sigmoid).  Replaces Base3dModel._forward_pts_dir for the NGP configuration (arcnerf/models/base_3d_model.py:233-254 ->
 * FusedMLPGeoNet.forward / FusedMLPRadianceNet.forward, tcnn_fusedmlp_module.py:81-111,177-192); same values as
 * arcb200_grid_fwd + arcb200_mlp_fwd_sh + arcb200_mlp_fwd_ex in sequence.  unit (n,3) fp32; dirs (n_rays,3) with ray_id (n)
 * int32, or ray_id NULL and dirs (n,3); outputs sigma (n), rgb (n,3) fp32.  Optional (NULL to skip) outputs kept for the
 * backward kernels: enc_out (n,32) fp16, rad_in_out (n,32) fp16, geo_act / rad_act (arcb200_mlp_act_save_bytes), and
 * zero_scalar, one float the launch clears. */
int arcb200_field_fwd(const float *unit, const void *table_half, int n_levels, int n_features, const uint32_t *offsets,
                      const float *scales, const uint32_t *resolutions, const void *geo_blob, const void *rad_blob,
                      const float *dirs, const int32_t *ray_id, int sh_remap, void *enc_out, void *rad_in_out, void *geo_act,
                      void *rad_act, float *sigma, float *rgb, float *zero_scalar, int64_t n, const int32_t *n_dev,
                      void *stream);

/* ---- vanilla-NeRF network pair as one dense-layer chain (csrc/dense_chain.cu) ----------------------------------------------
 * GeoNet (D hidden layers of width W = 256 with ReLU, input embedding re-concatenated after layer `skip` (-1: none), last
 * layer 1 + W outputs) + RadianceNet (mode 'vf': [view embedding | feature] -> W_rad -> 3, sigmoid) with the frequency
 * embeddings (d_x inputs x n_freq_x bands, 3 x n_freq_v, log sampling, input included) computed inside the kernels.
 * Replaces arcnerf/models/base_modules/geo_rad_model/linear_network_module.py:17-197,234-335 (torch.nn.Linear stacks) and
 * encoding/freq_encoder.py:10-88 as used by NeRF (nerf_model.py:52-91) and the NeRF++ background (nerfpp_bkg_model.py:51-114).
 * Flat fp32 parameter vector: geometry layers 0..D, each weight (out x in, row-major, the layer's full input width) followed
 * by its bias, then the two radiance layers likewise - the reference's `layers.{i}.weight / bias` tensors in order.
 * arcb200_nerf_mlp_sizes: out[0] parameters, out[1] / out[2] bytes of the forward / backward operand blobs, out[3] floats of
 * the vector buffer, out[4] / out[5] fp16 columns per point of the forward / backward save buffers.
 * arcb200_nerf_mlp_pack: parameters -> blobs + vectors (after every update).  _fwd: x (n,d_x), dirs (n,3) -> sigma (n) raw,
 * rgb (n,3); save = NULL for inference or (out[4] x n_cap) fp16.  _bwd: d_sigma (n), d_rgb (n,3) -> d_params ACCUMULATED;
 * bwd_save (out[5] x n_cap) fp16 and dsig16 (n_cap x 16) fp16 are scratch; loss_scale multiplies the gradients inside the fp16
 * tiles and is divided out again (128 in the reference's tcnn path).  n_cap % 128 == 0; the save buffers keep the kernels'
 * operand-tile layout (128-row tiles), they are opaque to the caller. */
int arcb200_nerf_mlp_sizes(int d_x, int n_freq_x, int n_freq_v, int W, int D, int skip, int W_rad, int64_t *out);
int arcb200_nerf_mlp_pack(int d_x, int n_freq_x, int n_freq_v, int W, int D, int skip, int W_rad, const float *params,
                          void *blob_fwd, void *blob_bwd, float *vec, void *stream);
int arcb200_nerf_mlp_fwd(int d_x, int n_freq_x, int n_freq_v, int W, int D, int skip, int W_rad, const float *x,
                         const float *dirs, const void *blob_fwd, const float *vec, void *save, int64_t n_cap, float *sigma,
                         float *rgb, int64_t n, void *stream);
int arcb200_nerf_mlp_bwd(int d_x, int n_freq_x, int n_freq_v, int W, int D, int skip, int W_rad, const void *blob_bwd,
                         const float *vec, const void *fwd_save, const float *rgb, const float *d_sigma, const float *d_rgb,
                         float loss_scale, void *bwd_save, void *dsig16, int64_t n_cap, float *d_params, int64_t n, void *stream);

/* ---- camera rays on the device (csrc/raygen.cu) -----------------------------------------------------------------------
 * One ray per (camera, pixel): replaces arcnerf/render/ray_helper.py:12-119 get_rays (pixel grid, pixel_to_world, direction,
 * normalisation, mip-NeRF radius) and the per-batch ray gather of arcnerf/trainer/pipeline.py:204-300.  cams: DEVICE array of
 * n_cams x 17 floats (fx, fy, cx, cy, skew, rows 0..2 of c2w row-major); cam_id (n) int32 or NULL (camera 0); pix (n) int64
 * pixel index i * H + j, or NULL for the full W x H image (n = W * H) in wh order (wh_order != 0) or row-major; rays_r (n) or
 * NULL (full image only). */
int arcb200_gen_rays(const float *cams, int n_cams, const int32_t *cam_id, const int64_t *pix, int W, int H, int wh_order,
                     int center_pixel, int normalize, float *rays_o, float *rays_d, float *rays_r, int64_t n, void *stream);

/* ---- mesh extraction on the device (csrc/mcubes.cu) ---------------------------------------------------------------------
 * arcb200_volume_pts: voxel-centre points of x slabs [x0, x1) of an n_grid^3 lattice over the box [lo, hi] (HOST float[3] each),
 * pts ((x1 - x0) * n_grid^2, 3), x outermost - Volume.cal_volume_pts, arcnerf/geometry/volume.py:225-232 (symmetric linspace per
 * axis), generated slab by slab so that the full lattice never has to be stored.
 *
 * Marching cubes over vol (nx, ny, nz) fp32, z fastest - extract_mesh, arcnerf/geometry/mesh.py:12-47 (which runs
 * skimage.measure.marching_cubes on the host): ascent = 0 for grad_dir 'descent' (inside: value > level), 1 for 'ascent'.
 *   1. arcb200_mc_scratch_sizes -> out[0..2] = int32 entries of point_blocks, cell_blocks, vbase;
 *   2. arcb200_mc_count fills point_blocks / cell_blocks with exclusive block offsets; their LAST entries are the vertex and the
 *      face count (read them back to size the outputs);
 *   3. arcb200_mc_emit writes verts (V, 3) = index * spacing - offset (HOST float[3] each; extract_mesh's offset is
 *      (volume_len - voxel_size) / 2), unit vertex normals (V, 3) pointing outside (NULL to skip) and faces (F, 3) int32 whose
 *      winding has the normal pointing outside (what extract_mesh returns after its faces[:, ::-1]).
 * Vertices are welded by construction (one per crossing lattice edge, in lattice order); the order of vertices and faces and the
 * resolution of ambiguous cells follow this library's generated case table, not skimage's. */
int arcb200_volume_pts(const float *lo, const float *hi, int n_grid, int x0, int x1, float *pts, void *stream);
int arcb200_mc_scratch_sizes(int nx, int ny, int nz, int64_t *out);
int arcb200_mc_count(const float *vol, int nx, int ny, int nz, float level, int ascent, int32_t *point_blocks,
                     int32_t *cell_blocks, void *stream);
int arcb200_mc_emit(const float *vol, int nx, int ny, int nz, float level, int ascent, const float *spacing, const float *offset,
                    const int32_t *point_blocks, const int32_t *cell_blocks, int32_t *vbase, float *verts, float *normals,
                    int64_t v_cap, int32_t *faces, int64_t f_cap, void *stream);
/* Mesh components of get_normals / get_face_centers (arcnerf/geometry/mesh.py:88-125, trimesh on the host in the reference): unit
 * face normals (F, 3) by the right-hand rule, face centres (F, 3), unit vertex normals (V, 3) = normalised sum of the unit
 * normals of the vertex's faces.  Any output may be NULL. */
int arcb200_mesh_components(const float *verts, const int32_t *faces, int64_t n_v, int64_t n_f, float *face_normals,
                            float *face_centers, float *vert_normals, void *stream);

/* ---- surface-ray intersection (csrc/surface.cu) -------------------------------------------------------------------------
 * The per-ray state machines of arcnerf/geometry/ray.py:403-601 around the caller's field evaluation; no host synchronisation
 * inside the loops (the reference's `torch.all(...)` early exits become device-side counters).  near / far: (n_rays) when
 * near_far_per_ray, else one value.  mask / obj / not_occ: uint8 per ray.
 * sphere_tracing (:403-467): zvals = near, mask = 1, obj = 0, sdf = 0; repeat n_iter times { pts = arcb200_ray_points(zvals);
 *   sdf_new = field(pts); arcb200_trace_update }; arcb200_surface_finish (zvals <= near -> 0, pts).
 * secant_root_finding (:470-601): arcb200_secant_steps (uniform scan + its points) -> geo = field(pts) -> arcb200_secant_init
 *   (bracket of the first + -> - sign change, first estimate; clears flags, n_iter + 1 int32) -> repeat for it = 0..n_iter-1
 *   { pts = arcb200_ray_points(z_mid); arcb200_secant_update } -> arcb200_secant_select -> arcb200_surface_finish.
 *   As in the reference, only the low end of the bracket ever moves (`~ind_low.sum() > 0` never holds, ray.py:571). */
int arcb200_trace_update(const float *sdf_new, float *zvals, float *sdf, const float *near, const float *far, int near_far_per_ray,
                         uint8_t *mask, uint8_t *obj, float threshold, int64_t n_rays, void *stream);
int arcb200_surface_finish(const float *rays_o, const float *rays_d, float *zvals, const float *near, int near_per_ray, float *pts,
                           int64_t n_rays, void *stream);
int arcb200_secant_steps(const float *rays_o, const float *rays_d, const float *near, const float *far, int near_far_per_ray,
                         int n_step, float *step, float *pts, int64_t n_rays, void *stream);
int arcb200_secant_init(const float *geo, const float *step, int n_step, float level, int descent, uint8_t *not_occ, uint8_t *mask,
                        float *z_high, float *z_low, float *g_high, float *g_low, float *z_mid, float *z_mid_init, int32_t *flags,
                        int n_iter, int64_t n_rays, void *stream);
int arcb200_secant_update(const float *geo_mid, float level, int descent, const uint8_t *mask, float *z_low, float *g_low,
                          const float *z_high, const float *g_high, float *z_mid, const float *z_mid_init, float threshold,
                          int32_t *flags, int it, int n_iter, int64_t n_rays, void *stream);
int arcb200_secant_select(const uint8_t *mask, const uint8_t *not_occ, const float *z_mid, const float *far, int far_per_ray,
                          float *zvals, int64_t n_rays, void *stream);

/* ---- mesh preview rasteriser (csrc/raster.cu) --------------------------------------------------------------------------------
 * Replaces the pytorch3d renderer behind render_mesh_images (arcnerf/geometry/mesh.py:174-240, arcnerf/render/
 * render_pytorch3d.py; blur radius 0, one face per pixel): n_views pinhole views (w2c DEVICE (n_views,4,4); intrinsic HOST 3x3)
 * of a mesh with vertex colours (V,3), face colours (F,3) or neither (silhouette), image (n_views,H,W,3) uint8 on background
 * `bkg` (0..1), optional winning face id (n_views,H,W) int32 (-1: none) and depth.  Nearest face per pixel centre through a 64-bit
 * atomicMin z-buffer (scratch: arcb200_raster_scratch_bytes); faces with a vertex at z <= near are skipped; unlit colours
 * (pytorch3d's Phong lighting is not reproduced). */
int arcb200_raster_scratch_bytes(int n_views, int H, int W, int64_t *bytes);
int arcb200_raster_mesh(const float *verts, const int32_t *faces, const float *vert_colors, const float *face_colors, const float *w2c,
                        const float *intrinsic, int n_views, int H, int W, float near, float bkg, void *scratch, uint8_t *image,
                        int32_t *face_id, float *depth, int64_t n_v, int64_t n_f, void *stream);

/* ---- narrow dense layers in fp32 (csrc/small_gemm.cu) -------------------------------------------------------------------
 * The SDF network of the NeuS-NGP configuration (32 -> 64 softplus(100) -> 17, weight-normed, GeoNet of
 * geo_rad_model/linear_network_module.py:17-197 with DenseLayer = nn.Linear + activation, base_modules/linear.py) and the
 * derivative passes autograd builds for forward_with_grad and the eikonal loss (base_network.py:30-44).
 * arcb200_gemm_nt: C (n x N) = A (n x K) . B^T (+ bias), B (N x K), N, K <= 128.   arcb200_gemm_tn: C (N x K) += A^T . B with
 * A (n x N), B (n x K), N, K <= 128 (caller zeroes C).   arcb200_softplus: order-th derivative (0, 1, 2) of softplus_beta. */
int arcb200_gemm_nt(const float *A, int64_t lda, const float *B, int64_t ldb, const float *bias, float *C, int64_t ldc, int64_t n,
                    int N, int K, void *stream);
int arcb200_gemm_tn(const float *A, int64_t lda, const float *B, int64_t ldb, float *C, int64_t ldc, int64_t n, int N, int K,
                    void *stream);
int arcb200_softplus(const float *x, float beta, int order, float *y, int64_t n, void *stream);

/* ---- measurement aid (csrc/probe.cu) ---------------------------------------------------------------------------------
 * Random-gather bandwidth of a device table: n_threads (rounded up to 256) threads issue trips * 8 independent loads of
 * req_bytes (4, 8 or 16) at uniformly random entries of `table` (n_bytes) and write one checksum each to sink
 * (n_threads uint32).  The roof the hash-grid kernels are compared with; no counterpart in the reference.
 * Returns the number of requests issued, negative on error. */
long long arcb200_probe_gather(const void *table, int64_t n_bytes, int req_bytes, int n_threads, int trips, uint32_t seed,
                               uint32_t *sink, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ARCB200_H */
