"""A small OmegaConf-compatible configuration subset.

The reference configures every model through OmegaConf
(reference: gluefactory/models/base_model.py:13-22,65-86, matchers/lightglue.py:335,
models/two_view_pipeline.py:18).  `omegaconf` is not installable on the build image, so
`get_conf_backend()` returns the real package when it is importable and this subset otherwise.
Only the calls that occur on the extraction+matching path are provided:
create, merge, set_struct, set_readonly, to_container, load, from_cli/from_dotlist,
is_config, read_write, open_dict, and DictConfig attribute / item / mapping access.
"""
from __future__ import annotations

import contextlib
import copy
from typing import Any


class MissingMandatoryValue(Exception):
    pass


class ConfigAttributeError(AttributeError):
    pass


class ConfigKeyError(KeyError):
    pass


class ReadonlyConfigError(Exception):
    pass


def _wrap(value, parent_flags=None):
    if isinstance(value, DictConfig):
        return DictConfig(value._content_plain())
    if isinstance(value, dict):
        return DictConfig(value)
    if isinstance(value, (list, tuple)):
        return [_wrap(v) if isinstance(v, (dict, DictConfig)) else copy.deepcopy(v) for v in value]
    return value


class DictConfig:
    """Nested mapping with attribute access, `struct` and `readonly` flags."""

    __slots__ = ("_content", "_struct", "_readonly")

    def __init__(self, content=None):
        object.__setattr__(self, "_content", {})
        object.__setattr__(self, "_struct", False)
        object.__setattr__(self, "_readonly", False)
        if content is not None:
            for k, v in dict(content).items():
                self._content[str(k) if not isinstance(k, str) else k] = _wrap(v)

    # -- flags ---------------------------------------------------------------------------
    def _set_flag(self, name, value):
        object.__setattr__(self, name, value)
        for v in self._content.values():
            if isinstance(v, DictConfig):
                v._set_flag(name, value)

    def _content_plain(self):
        out = {}
        for k, v in self._content.items():
            if isinstance(v, DictConfig):
                out[k] = v._content_plain()
            elif isinstance(v, list):
                out[k] = [x._content_plain() if isinstance(x, DictConfig) else copy.deepcopy(x) for x in v]
            else:
                out[k] = v
        return out

    # -- access --------------------------------------------------------------------------
    def _get(self, key, exc):
        try:
            value = self._content[key]
        except KeyError:
            raise exc(f"Missing key {key}") from None
        if isinstance(value, str) and value == "???":
            raise MissingMandatoryValue(f"Missing mandatory value: {key}")
        return value

    def __getattr__(self, key):
        if key.startswith("__"):
            raise AttributeError(key)
        return self._get(key, ConfigAttributeError)

    def __getitem__(self, key):
        return self._get(key, ConfigKeyError)

    def _set(self, key, value):
        if self._readonly:
            raise ReadonlyConfigError(f"Cannot change read-only config container: {key}")
        if self._struct and key not in self._content:
            raise ConfigAttributeError(f"Key '{key}' is not in struct")
        value = _wrap(value)
        if isinstance(value, DictConfig):
            value._set_flag("_struct", self._struct)
            value._set_flag("_readonly", self._readonly)
        self._content[key] = value

    def __setattr__(self, key, value):
        self._set(key, value)

    def __setitem__(self, key, value):
        self._set(key, value)

    def __delitem__(self, key):
        if self._readonly:
            raise ReadonlyConfigError(key)
        del self._content[key]

    def __contains__(self, key):
        return key in self._content

    def __iter__(self):
        return iter(self._content)

    def __len__(self):
        return len(self._content)

    def __eq__(self, other):
        if isinstance(other, DictConfig):
            return self._content_plain() == other._content_plain()
        if isinstance(other, dict):
            return self._content_plain() == other
        return NotImplemented

    def __repr__(self):
        return repr(self._content_plain())

    def __deepcopy__(self, memo):
        new = DictConfig(self._content_plain())
        new._set_flag("_struct", self._struct)
        new._set_flag("_readonly", self._readonly)
        return new

    def __getstate__(self):
        return (self._content_plain(), self._struct, self._readonly)

    def __setstate__(self, state):
        content, struct, readonly = state
        DictConfig.__init__(self, content)
        self._set_flag("_struct", struct)
        self._set_flag("_readonly", readonly)

    def keys(self):
        return self._content.keys()

    def values(self):
        return [self[k] for k in self._content]

    def items(self):
        return [(k, self[k]) for k in self._content]

    def get(self, key, default=None):
        if key not in self._content:
            return default
        value = self._content[key]
        if isinstance(value, str) and value == "???":
            return default
        return value

    def pop(self, key, *default):
        if self._readonly:
            raise ReadonlyConfigError(key)
        if self._struct:
            raise ConfigAttributeError("DictConfig in struct mode does not support pop")
        return self._content.pop(key, *default)


def _merge_into(dst: DictConfig, src) -> None:
    items = src._content.items() if isinstance(src, DictConfig) else dict(src).items()
    for k, v in items:
        if k not in dst._content and dst._struct:
            raise ConfigKeyError(f"Key '{k}' not in struct config")
        cur = dst._content.get(k)
        if isinstance(cur, DictConfig) and isinstance(v, (dict, DictConfig)):
            _merge_into(cur, v)
        else:
            new = _wrap(v)
            if isinstance(new, DictConfig):
                new._set_flag("_struct", dst._struct)
            dst._content[k] = new


def _parse_scalar(text: str) -> Any:
    import yaml

    try:
        return yaml.safe_load(text)
    except yaml.YAMLError:
        return text


class OmegaConf:
    """Static API mirror of `omegaconf.OmegaConf` (subset)."""

    @staticmethod
    def create(obj=None):
        if obj is None:
            return DictConfig()
        if isinstance(obj, str):
            import yaml

            obj = yaml.safe_load(obj) or {}
        if isinstance(obj, (list, tuple)):
            return _wrap(list(obj))
        return DictConfig(obj)

    @staticmethod
    def merge(*confs):
        if not confs:
            return DictConfig()
        first = confs[0]
        out = DictConfig(first._content_plain() if isinstance(first, DictConfig) else first)
        if isinstance(first, DictConfig):
            out._set_flag("_struct", first._struct)
        for c in confs[1:]:
            if c is None:
                continue
            _merge_into(out, c)
        return out

    @staticmethod
    def set_struct(conf, value):
        conf._set_flag("_struct", bool(value))

    @staticmethod
    def set_readonly(conf, value):
        conf._set_flag("_readonly", bool(value))

    @staticmethod
    def is_config(obj):
        return isinstance(obj, DictConfig)

    @staticmethod
    def is_dict(obj):
        return isinstance(obj, DictConfig)

    @staticmethod
    def to_container(conf, resolve=False, **_):
        if isinstance(conf, DictConfig):
            return conf._content_plain()
        return copy.deepcopy(conf)

    @staticmethod
    def to_yaml(conf):
        import yaml

        return yaml.safe_dump(OmegaConf.to_container(conf), sort_keys=False)

    @staticmethod
    def load(path):
        import yaml

        with open(path, "r") as fid:
            return DictConfig(yaml.safe_load(fid) or {})

    @staticmethod
    def save(conf, path):
        with open(path, "w") as fid:
            fid.write(OmegaConf.to_yaml(conf))

    @staticmethod
    def from_dotlist(dotlist):
        out = DictConfig()
        for entry in dotlist:
            key, _, value = entry.partition("=")
            node = out
            parts = key.split(".")
            for p in parts[:-1]:
                if p not in node._content:
                    node._content[p] = DictConfig()
                node = node._content[p]
            node._content[parts[-1]] = _wrap(_parse_scalar(value))
        return out

    @staticmethod
    def from_cli(args_list=None):
        import sys

        return OmegaConf.from_dotlist(sys.argv[1:] if args_list is None else args_list)


@contextlib.contextmanager
def read_write(conf):
    prev = conf._readonly if isinstance(conf, DictConfig) else False
    if isinstance(conf, DictConfig):
        conf._set_flag("_readonly", False)
    try:
        yield conf
    finally:
        if isinstance(conf, DictConfig):
            conf._set_flag("_readonly", prev)


@contextlib.contextmanager
def open_dict(conf):
    prev = conf._struct if isinstance(conf, DictConfig) else False
    if isinstance(conf, DictConfig):
        conf._set_flag("_struct", False)
    try:
        yield conf
    finally:
        if isinstance(conf, DictConfig):
            conf._set_flag("_struct", prev)


def get_conf_backend():
    """Return (module, OmegaConf): the real omegaconf when installed, else this subset."""
    try:
        import omegaconf as _real  # noqa: WPS433

        if getattr(_real, "__b200glue_shim__", False):
            raise ImportError
        return _real, _real.OmegaConf
    except ImportError:
        import sys

        return sys.modules[__name__], OmegaConf
