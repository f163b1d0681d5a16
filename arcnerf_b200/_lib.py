"""ctypes binding of libarcb200.so (include/arcb200.h).

The product path has NO fallback: if the shared library is missing or a call fails, a RuntimeError is
raised.  Nothing here imports oracle/.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libarcb200.so')

P = C.c_void_p      # device or host pointer
I = C.c_int
L = C.c_int64
F = C.c_float
U64 = C.c_uint64

# name -> argtypes (all return int unless listed in _RESTYPE)
_SIGS = {
    'arcb200_version': [],
    'arcb200_last_error': [],
    'arcb200_device_count': [],
    'arcb200_rng_init': [U64, C.POINTER(U64), C.POINTER(U64)],
    'arcb200_rng_advance': [C.POINTER(U64), U64, L],
    'arcb200_aabb_intersection': [P, P, P, P, P, P, P, L, L, P],
    'arcb200_check_pts_in_occ_voxel': [P, P, P, I, P, L, P],
    'arcb200_sparse_volume_sampling': [P, P, P, P, I, F, P, I, P, F, P, P, L, U64, U64, P],
    'arcb200_sparse_volume_sampling_bit': [P, P, P, P, I, F, P, I, P, F, P, P, L, U64, U64, P],
    'arcb200_tensor_reduce_max': [P, P, I, P, L, P],
    'arcb200_generate_grid_samples': [P, I, L, I, F, P, P, U64, U64, P],
    'arcb200_splat_grid_samples': [P, P, L, P, P],
    'arcb200_ema_grid_samples_nerf': [P, L, F, P, P],
    'arcb200_update_bitfield': [P, F, P, P, F, L, P],
    'arcb200_count_bitfield': [P, P, L, I, P],
    'arcb200_sparse_sampling_in_multivol_bitfield': [P, P, P, P, I, F, F, F, P, P, I, I, P, F, I, P, P, L, U64, U64, P],
    'arcb200_generate_grid_samples_multivol': [P, I, L, P, I, I, F, I, P, P, U64, U64, P],
    'arcb200_march_packed_scratch_bytes': [L, I],
    'arcb200_march_packed': [I, P, P, P, P, I, F, P, I, P, F, L, U64, U64, P, P, P, P, L, P, P, P, P, P],
    'arcb200_march_packed_static_max_slots': [],
    'arcb200_march_packed_static': [I, P, P, P, P, I, F, P, I, P, F, L, U64, U64, P, P, P, L, P, P, P, P, P],
    'arcb200_grid_geometry': [I, I, I, F, P, P, P],
    'arcb200_cast_f32_to_f16': [P, P, L, P],
    'arcb200_grid_fwd': [P, P, I, I, P, P, P, P, P, L, P, P],
    'arcb200_grid_bwd': [P, P, I, F, I, I, P, P, P, P, L, P, P],
    'arcb200_grid_bwd_input': [P, I, F, P, I, I, P, L, P],
    'arcb200_grid_bwd_bwd_input': [P, P, P, I, F, P, I, I, P, P, P, P, P, P, L, P],
    'arcb200_sh_fwd': [P, P, I, I, P, I, L, L, P, P],
    'arcb200_mlp_n_params': [I, I, I, I],
    'arcb200_mlp_blob_bytes': [I, I, I, I],
    'arcb200_mlp_pack_weights': [P, P, I, I, I, I, P],
    'arcb200_mlp_fwd': [P, I, L, P, I, I, I, I, I, I, P, I, L, L, P, P],
    'arcb200_mlp_act_save_bytes': [L, I, I],
    'arcb200_mlp_fwd_ex': [P, I, L, P, I, I, I, I, I, I, P, I, L, P, P, L, P, P],
    'arcb200_mlp_fwd_sh': [P, I, L, P, I, I, I, I, I, I, P, L, P, P, P, P, I, I, P, L, P, P],
    'arcb200_mlp_bwd_ex': [P, I, L, P, I, I, I, I, I, I, P, L, F, P, I, L, P, P, P, I, L, P, P, I, L, F, L, P, P],
    'arcb200_mlp_bwd': [P, I, L, P, I, I, I, I, I, I, P, L, F, P, I, L, P, L, P, P],
    'arcb200_composite_fwd': [P, I, L, P, P, I, P, P, P, I, I, P, I, P, P, P, P, P, P, L, P],
    'arcb200_composite_fwd_ex': [P, I, L, P, P, I, P, P, P, I, I, P, I, P, P, P, P, P, P, P, L, P],
    'arcb200_composite_bwd': [P, I, L, P, P, I, P, P, P, I, I, P, I, P, P, P, P, P, P, I, L, P],
    'arcb200_composite_huber_train': [P, I, L, P, P, P, P, P, I, I, P, I, P, F, F, P, P, P, P, P, P, I, L, P],
    'arcb200_sdf_to_alpha': [P, P, P, P, F, I, P, L, I, P],
    'arcb200_sdf_to_alpha_bwd': [P, P, P, P, F, I, P, P, P, P, L, I, P],
    'arcb200_sample_cdf': [P, P, P, I, F, P, L, I, I, P],
    'arcb200_trunc_exp_fwd': [P, I, L, P, L, P, P],
    'arcb200_trunc_exp_bwd': [P, I, L, P, F, P, L, P, P],
    'arcb200_world_to_unit': [P, P, P, L, P],
    'arcb200_geo_head_grad': [P, I, L, P, F, P, I, L, I, P, L, P, P],
    'arcb200_adam_step': [P, P, P, P, P, L, I, F, F, F, F, F, F, I, P, P],
    'arcb200_counter_add': [P, I, P],
    'arcb200_fill_f32': [P, F, L, P],
    'arcb200_scale_f32': [P, F, L, P],
    'arcb200_huber_loss': [P, P, F, F, P, P, L, P],
    'arcb200_mlp_adam_pack': [I, P, P, P, P, P, I, P, I, F, F, F, F, F, F, I, P, P],
    'arcb200_freq_embed_fwd': [P, I, P, I, I, P, L, L, P],
    'arcb200_freq_embed_bwd': [P, I, P, I, I, P, L, P, L, P],
    'arcb200_volume_refresh_scratch_bytes': [I],
    'arcb200_volume_refresh_candidates': [P, I, L, L, C.c_uint32, U64, U64, P, P, P, P, P],
    'arcb200_volume_refresh_apply': [P, P, L, F, F, I, P, P, P, P],
    'arcb200_sphere_ray_intersection': [P, P, P, P, P, P, P, P, L, L, P],
    'arcb200_zvals_from_near_far': [P, P, I, I, I, P, P, L, P],
    'arcb200_zvals_fix_step': [P, P, F, I, I, P, P, P, L, P],
    'arcb200_ray_points': [P, P, P, P, L, I, P],
    'arcb200_sort_rows': [P, P, L, I, P],
    'arcb200_peer_alloc': [L, P, P],
    'arcb200_peer_open': [P, P],
    'arcb200_peer_close': [P],
    'arcb200_peer_free': [P],
    'arcb200_peer_barrier': [P, I, I, P, P, I, P],
    'arcb200_grad_to_half': [P, P, L, L, L, F, P],
    'arcb200_adam_step_peer16': [P, P, P, I, I, L, L, P, P, P, F, I, F, F, F, F, F, F, P, P],
    'arcb200_adam_step_mc16': [P, P, L, L, P, P, P, F, I, F, F, F, F, F, F, P, P],
    'arcb200_adam_step_peer': [P, P, I, L, L, P, P, P, I, F, F, F, F, F, F, P, P],
    'arcb200_probe_gather': [P, L, I, I, I, C.c_uint32, P, P],
    'arcb200_nerf_mlp_sizes': [I, I, I, I, I, I, I, P],
    'arcb200_nerf_mlp_pack': [I, I, I, I, I, I, I, P, P, P, P, P],
    'arcb200_nerf_mlp_fwd': [I, I, I, I, I, I, I, P, P, P, P, P, L, P, P, L, P],
    'arcb200_nerf_mlp_bwd': [I, I, I, I, I, I, I, P, P, P, P, P, P, F, P, P, L, P, L, P],
    'arcb200_gen_rays': [P, I, P, P, I, I, I, I, I, P, P, P, L, P],
    'arcb200_volume_pts': [P, P, I, I, I, P, P],
    'arcb200_mc_scratch_sizes': [I, I, I, P],
    'arcb200_mc_count': [P, I, I, I, F, I, P, P, P],
    'arcb200_mc_emit': [P, I, I, I, F, I, P, P, P, P, P, P, P, L, P, L, P],
    'arcb200_check_increasing': [P, L, I, P, P],
    'arcb200_mesh_components': [P, P, L, L, P, P, P, P],
    'arcb200_raster_scratch_bytes': [I, I, I, P],
    'arcb200_raster_mesh': [P, P, P, P, P, P, I, I, I, F, F, P, P, P, P, L, L, P],
    'arcb200_trace_update': [P, P, P, P, P, I, P, P, F, L, P],
    'arcb200_surface_finish': [P, P, P, P, I, P, L, P],
    'arcb200_secant_steps': [P, P, P, P, I, I, P, P, L, P],
    'arcb200_secant_init': [P, P, I, F, I, P, P, P, P, P, P, P, P, P, I, L, P],
    'arcb200_secant_update': [P, F, I, P, P, P, P, P, P, P, F, P, I, I, L, P],
    'arcb200_secant_select': [P, P, P, P, I, P, L, P],
    'arcb200_gemm_nt': [P, L, P, L, P, P, L, L, I, I, P],
    'arcb200_gemm_tn': [P, L, P, L, P, L, L, I, I, P],
    'arcb200_softplus': [P, F, I, P, L, P],
    'arcb200_field_fwd': [P, P, I, I, P, P, P, P, P, P, P, I, P, P, P, P, P, P, P, L, P, P],
}
_RESTYPE = {
    'arcb200_probe_gather': C.c_longlong,
    'arcb200_last_error': C.c_char_p,
    'arcb200_rng_init': None,
    'arcb200_rng_advance': None,
    'arcb200_march_packed_scratch_bytes': L,
    'arcb200_march_packed_static_max_slots': L,
    'arcb200_grid_geometry': L,
    'arcb200_mlp_act_save_bytes': L,
    'arcb200_volume_refresh_scratch_bytes': L,
}

_lib = None


def load():
    """Load libarcb200.so; raises RuntimeError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            'libarcb200.so not found at %s: build it with `python -m arcnerf_b200.build` '
            '(there is no CPU or PyTorch fallback for this path)' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, args in _SIGS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue  # entry point of a later source file that is not in this build
        fn.argtypes = args
        fn.restype = _RESTYPE.get(name, I)
    _lib = lib
    return lib


def register(name, argtypes, restype=I):
    """Used by host modules that add entry points (keeps the table next to its caller)."""
    _SIGS[name] = argtypes
    if restype is not I:
        _RESTYPE[name] = restype
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.argtypes = argtypes
        fn.restype = restype


def last_error():
    msg = load().arcb200_last_error()
    return msg.decode() if msg else ''


def call(name, *args):
    """Invoke an int-returning entry point; a negative return raises RuntimeError (the reference's
    pybind layer raises RuntimeError from TORCH_CHECK / std::runtime_error, arcnerf/ops/include/utils.h:17-27)."""
    lib = load()
    fn = getattr(lib, name, None)
    if fn is None:
        raise RuntimeError('libarcb200.so does not export %s (stale build?)' % name)
    rc = fn(*args)
    if rc is not None and rc < 0:
        raise RuntimeError('%s failed (%d): %s' % (name, rc, last_error()))
    return rc


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def current_stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


class Rng:
    """Caller-held pcg32 state matching the reference's per-module `static pcg32 rng{9121}`."""
    SEED = 9121

    def __init__(self, seed=SEED):
        s, i = U64(), U64()
        load().arcb200_rng_init(U64(seed), C.byref(s), C.byref(i))
        self.state, self.inc = s.value, i.value

    def advance(self, delta=1 << 32):
        s = U64(self.state)
        load().arcb200_rng_advance(C.byref(s), U64(self.inc), L(delta))
        self.state = s.value
