"""Name -> class registries with the reference's interface (common/utils/registry.py:8-86): `@REG.register()`
decorates a class, `REG.get(name)` returns it; components are instantiated as `REG.get(cfg.type)(**cfg.__dict__)`
(arcnerf/models/base_modules/encoding/__init__.py:26-51, base_modules/__init__.py:28-61)."""


class Registry:
    def __init__(self, name):
        self._name = name
        self._obj_map = {}

    def register(self, obj=None):
        if obj is None:
            def deco(cls):
                self._do_register(cls.__name__, cls)
                return cls
            return deco
        self._do_register(obj.__name__, obj)
        return obj

    def _do_register(self, name, obj):
        assert name not in self._obj_map, "An object named '{}' was already registered in '{}' registry!".format(name, self._name)
        self._obj_map[name] = obj

    def get(self, name):
        ret = self._obj_map.get(name)
        if ret is None:
            raise KeyError("No object named '{}' found in '{}' registry!".format(name, self._name))
        return ret

    def __contains__(self, name):
        return name in self._obj_map


ENCODER_REGISTRY = Registry('encoder')
MODULE_REGISTRY = Registry('module')
BOUND_REGISTRY = Registry('bound')
MODEL_REGISTRY = Registry('model')          # arcnerf/models/__init__.py:19-41 (NeRF, NeRFPP, Neus ... selected by cfgs.model.type)


def cfg_to_dict(cfg):
    """Accepts the reference's nested `Obj` config nodes (common/utils/cfgs_utils.py:10-31) or plain dicts."""
    if cfg is None:
        return {}
    if isinstance(cfg, dict):
        return dict(cfg)
    return dict(cfg.__dict__)


def build_encoder(cfg):
    """arcnerf/models/base_modules/encoding/__init__.py:26-51: -> (encoder module, output dim)"""
    d = cfg_to_dict(cfg)
    if not d or d.get('type') is None:
        return None, None
    from . import hashgrid_encoder, sh_encoder  # noqa: F401  (registration)
    enc = ENCODER_REGISTRY.get(d['type'])(**{k: v for k, v in d.items() if k != 'type'})
    return enc, enc.get_output_dim()
