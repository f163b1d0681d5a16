"""Edge cases of the C ABI added late in round 1: empty inputs return OK without touching memory, bad arguments come
back as negative codes with a message (-> RuntimeError in the Python layer), limits are reported as limits."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def test_empty_inputs_are_accepted():
    from arcnerf_b200 import _lib
    st = _lib.current_stream()
    z = torch.zeros(8, device='cuda')
    f3 = (C.c_float * 3)(0.0, 0.0, 0.0)
    assert _lib.call('arcb200_sphere_ray_intersection', None, None, None, f3, None, None, None, None, 0, 1, st) == 0
    assert _lib.call('arcb200_zvals_from_near_far', None, None, 8, 1, 0, None, None, 0, st) == 0
    assert _lib.call('arcb200_zvals_fix_step', None, None, 0.1, 8, 1, None, None, None, 0, st) == 0
    assert _lib.call('arcb200_ray_points', None, None, None, None, 0, 4, st) == 0
    assert _lib.call('arcb200_sort_rows', None, None, 0, 16, st) == 0
    assert _lib.call('arcb200_freq_embed_fwd', None, 3, None, 0, 1, None, 3, 0, st) == 0
    assert _lib.call('arcb200_freq_embed_bwd', None, 3, None, 0, 1, None, 3, None, 0, st) == 0
    assert _lib.call('arcb200_world_to_unit', None, None, None, 0, st) == 0
    assert _lib.call('arcb200_fill_f32', None, 0.0, 0, st) == 0
    assert _lib.call('arcb200_grad_to_half', None, None, 0, 0, 0, 128.0, st) == 0
    assert _lib.call('arcb200_composite_huber_train', None, 4, 0, None, None, None, None, None, 0, 0, None, 0, None, 0.1, 1.0,
                     None, None, None, None, None, None, 4, 0, st) == 0
    assert _lib.call('arcb200_adam_step_peer', None, None, 1, 0, 0, None, None, None, 1, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1.0, None, st) == 0
    torch.cuda.synchronize()
    assert float(z.sum()) == 0.0


def test_bad_arguments_raise_with_a_message():
    from arcnerf_b200 import _lib
    st = _lib.current_stream()
    x = torch.zeros((16, 3), device='cuda')
    out = torch.zeros((16, 64), device='cuda')
    fr = (C.c_float * 40)(*([1.0] * 40))
    with pytest.raises(RuntimeError, match='frequency bands'):
        _lib.call('arcb200_freq_embed_fwd', x.data_ptr(), 3, fr, 33, 1, out.data_ptr(), 64, 16, st)
    with pytest.raises(RuntimeError, match='leading dimension'):
        _lib.call('arcb200_freq_embed_fwd', x.data_ptr(), 3, fr, 10, 1, out.data_ptr(), 62, 16, st)
    with pytest.raises(RuntimeError, match='nothing to write'):
        _lib.call('arcb200_freq_embed_fwd', x.data_ptr(), 3, fr, 0, 0, out.data_ptr(), 64, 16, st)
    with pytest.raises(RuntimeError, match='positive step'):
        _lib.call('arcb200_zvals_fix_step', x.data_ptr(), x.data_ptr(), 0.0, 8, 1, None, out.data_ptr(), None, 16, st)
    with pytest.raises(RuntimeError, match='8192'):
        _lib.call('arcb200_sort_rows', out.data_ptr(), out.data_ptr(), 1, 9000, st)
    with pytest.raises(RuntimeError, match='n_grid'):
        _lib.call('arcb200_volume_refresh_candidates', x.data_ptr(), 0, 1, 1, 0, 1, 1, x.data_ptr(), x.data_ptr(), x.data_ptr(),
                  x.data_ptr(), st)
    with pytest.raises(RuntimeError, match='candidate counts'):
        _lib.call('arcb200_volume_refresh_candidates', x.data_ptr(), 4, 65, 0, 0, 1, 1, x.data_ptr(), x.data_ptr(), x.data_ptr(),
                  x.data_ptr(), st)
    with pytest.raises(RuntimeError, match='1..4 networks'):
        _lib.call('arcb200_mlp_adam_pack', 5, None, None, None, None, None, 1, None, 1, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1.0, 1, None, st)
    with pytest.raises(RuntimeError, match='bad range'):
        _lib.call('arcb200_grad_to_half', x.data_ptr(), out.data_ptr(), 48, 4, 16, 128.0, st)
    with pytest.raises(RuntimeError, match='non-positive delta'):
        _lib.call('arcb200_composite_huber_train', None, 4, 16, None, x.data_ptr(), x.data_ptr(), x.data_ptr(), None, 0, 0, None, 0,
                  x.data_ptr(), 0.0, 1.0, None, None, None, out.data_ptr(), None, None, 4, 4, st)
    with pytest.raises(RuntimeError, match='samples per ray'):
        _lib.call('arcb200_composite_huber_train', None, 5000, 16, None, x.data_ptr(), x.data_ptr(), x.data_ptr(), None, 0, 0, None, 0,
                  x.data_ptr(), 0.1, 1.0, None, None, None, out.data_ptr(), None, None, 5000, 1, st)
    with pytest.raises(RuntimeError, match='fp16 output'):
        _lib.call('arcb200_mlp_fwd_sh', x.data_ptr(), 1, 32, x.data_ptr(), 32, 64, 1, 16, 1, 0, out.data_ptr(), 24, None, None,
                  x.data_ptr(), None, 16, 1, None, 16, None, st)
    with pytest.raises(RuntimeError, match='rank'):
        _lib.call('arcb200_peer_barrier', (C.c_void_p * 1)(x.data_ptr()), 1, 3, x.data_ptr(), x.data_ptr(), 10, st)
    assert 'rank' in _lib.last_error()
