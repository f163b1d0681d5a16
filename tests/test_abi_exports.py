"""CPU: libarcb200.so loads without a GPU and exports every symbol include/arcb200.h declares."""
import ctypes
import os
import re

from arcnerf_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'arcb200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(arcb200_\w+)\s*\(', text)))


def test_library_loads_and_exports_header():
    assert os.path.exists(build.LIB_PATH)
    lib = ctypes.CDLL(build.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 19
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_table_covers_header():
    names = declared_symbols()
    _lib.load()
    # host modules register their own entry points on import
    import arcnerf_b200.plugin  # noqa: F401
    import importlib
    for mod in ('volume_func', 'bitfield_func', 'multivol_func', 'hashgrid_encoder', 'sh_encoder',
                'tcnn_fusedmlp_module', 'ray_helper', 'trunc_exp'):
        try:
            importlib.import_module('arcnerf_b200.plugin.' + mod)
        except ModuleNotFoundError:
            pass
    missing = [n for n in names if n not in _lib._SIGS]
    assert not missing, missing


def test_version_and_error_text():
    lib = _lib.load()
    assert lib.arcb200_version() >= 100
    # no compute without a GPU; but argument errors are reported without touching the device
    try:
        _lib.call('arcb200_aabb_intersection', None, None, None, None, None, None, None, 4, 1, None)
    except RuntimeError as e:
        assert 'null pointer' in str(e)
    else:
        raise AssertionError('expected RuntimeError')


def test_rng_matches_reference_seed():
    r = _lib.Rng()
    assert (r.state, r.inc) == (0xc8e29e7d80289fd7, 0x3)
