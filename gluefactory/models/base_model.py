"""BaseModel contract (reference: gluefactory/models/base_model.py:13-157).

* `default_conf` dictionaries are merged down the class hierarchy by the metaclass;
* the constructor merges the user conf (dict or DictConfig) over the defaults, then freezes it;
* `forward` checks `required_data_keys` recursively and defers to `_forward`;
* `load_state_dict` marks the model initialised (`is_initialized`).
"""
from abc import ABCMeta, abstractmethod
from copy import copy

from torch import nn

from glue_factory_b200._conf import get_conf_backend

omegaconf, OmegaConf = get_conf_backend()


class MetaModel(ABCMeta):
    def __prepare__(name, bases, **kwds):
        merged = OmegaConf.create()
        for base in bases:
            for attr in ("base_default_conf", "default_conf"):
                piece = getattr(base, attr, {})
                if isinstance(piece, dict):
                    piece = OmegaConf.create(piece)
                merged = OmegaConf.merge(merged, piece)
        return dict(base_default_conf=merged)


class BaseModel(nn.Module, metaclass=MetaModel):
    default_conf = {
        "name": None,
        "trainable": True,
        "freeze_batch_normalization": False,
        "timeit": False,
    }
    required_data_keys = []
    strict_conf = False
    are_weights_initialized = False

    def __init__(self, conf):
        super().__init__()
        defaults = OmegaConf.merge(self.base_default_conf, OmegaConf.create(self.default_conf))
        if self.strict_conf:
            OmegaConf.set_struct(defaults, True)
        if "pad" in conf and "pad" not in defaults:  # legacy key, kept for old configs
            with omegaconf.read_write(conf):
                with omegaconf.open_dict(conf):
                    conf["interpolation"] = {"pad": conf.pop("pad")}
        if isinstance(conf, dict):
            conf = OmegaConf.create(conf)
        self.conf = conf = OmegaConf.merge(defaults, conf)
        OmegaConf.set_readonly(conf, True)
        OmegaConf.set_struct(conf, True)
        self.required_data_keys = copy(self.required_data_keys)
        self._init(conf)
        if not conf.trainable:
            for p in self.parameters():
                p.requires_grad = False

    def train(self, mode=True):
        super().train(mode)
        if self.conf.freeze_batch_normalization:
            for module in self.modules():
                if isinstance(module, nn.modules.batchnorm._BatchNorm):
                    module.eval()
        return self

    def forward(self, data):
        def check(expected, given):
            for key in expected:
                assert key in given, f"Missing key {key} in data"
                if isinstance(expected, dict):
                    check(expected[key], given[key])

        check(self.required_data_keys, data)
        return self._forward(data)

    @abstractmethod
    def _init(self, conf):
        raise NotImplementedError

    @abstractmethod
    def _forward(self, data):
        raise NotImplementedError

    @abstractmethod
    def loss(self, pred, data):
        raise NotImplementedError

    def load_state_dict(self, *args, **kwargs):
        result = super().load_state_dict(*args, **kwargs)
        self.set_initialized()
        return result

    def is_initialized(self):
        ok = True
        for _, child in self.named_children():
            if isinstance(child, BaseModel):
                ok = ok and child.is_initialized()
            else:
                has_params = len(list(child.parameters())) > 0
                ok = ok and (not has_params or self.are_weights_initialized)
        return ok

    def set_initialized(self, to: bool = True):
        self.are_weights_initialized = to
        for _, w in self.named_parameters():
            if isinstance(w, BaseModel):
                w.set_initialized(to)
