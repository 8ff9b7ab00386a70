"""Hierarchical configuration with the surface of the reference's deepclr/config/config.py that the model path
touches: declared / dynamic parameters, attribute + item access, mapping protocol (`**cfg.params`), yaml input,
freeze.  Training-only features (list overrides, dumps of docs) are not rebuilt."""
from __future__ import annotations

from collections import OrderedDict
from enum import Enum
from typing import Any, Dict, Optional, Union

import yaml


class ConfigEnum(Enum):
    @classmethod
    def create(cls, x: Union[str, 'ConfigEnum']) -> Any:
        if isinstance(x, cls):
            return x
        return cls[str(x).upper()]


class Config(OrderedDict):
    """Ordered group of parameters; nested dicts read from yaml become nested Config groups."""
    _RESERVED = ('_dynamic', '_frozen', '_required')

    def __init__(self, allow_dynamic_params: Optional[bool] = None):
        super().__init__()
        object.__setattr__(self, '_dynamic', bool(allow_dynamic_params))
        object.__setattr__(self, '_frozen', False)
        object.__setattr__(self, '_required', set())

    # -- access
    def __getattr__(self, key: str) -> Any:
        try:
            return self[key]
        except KeyError:
            raise AttributeError(key)

    def __setattr__(self, key: str, value: Any) -> None:
        self[key] = value

    def __setitem__(self, key: str, value: Any) -> None:
        if self._frozen:
            raise RuntimeError("Config is frozen")
        if key not in self and not self._dynamic:
            raise KeyError("Parameter '{}' is not defined".format(key))
        if isinstance(value, dict) and not isinstance(value, Config):
            value = Config._from_dict(value)
        super().__setitem__(key, value)

    @staticmethod
    def _from_dict(data: Dict) -> 'Config':
        grp = Config(allow_dynamic_params=True)
        for k, v in data.items():
            grp[k] = v
        return grp

    # -- schema
    def define_group(self, name: str, parent: Optional['Config'] = None) -> 'Config':
        grp = Config(allow_dynamic_params=self._dynamic)
        OrderedDict.__setitem__(parent if parent is not None else self, name, grp)
        return grp

    def define_param(self, name: str, parent: Optional['Config'] = None, default: Any = None,
                     required: Optional[bool] = None, doc: Optional[str] = None) -> None:
        target = parent if parent is not None else self
        OrderedDict.__setitem__(target, name, default)
        if required:
            target._required.add(name)

    add_internal_group = define_group

    def add_internal_param(self, name: str, parent: Optional['Config'] = None, value: Any = None) -> None:
        OrderedDict.__setitem__(parent if parent is not None else self, name, value)

    # -- io
    def read_dict(self, data: Dict) -> None:
        for k, v in data.items():
            if isinstance(v, dict) and isinstance(self.get(k), Config):
                self[k].read_dict(v)
            else:
                self[k] = v

    def read_file(self, f: str) -> None:
        with open(f, 'r') as fh:
            self.read_dict(yaml.safe_load(fh) or {})

    def read_str(self, s: str) -> None:
        self.read_dict(yaml.safe_load(s) or {})

    def dict(self) -> OrderedDict:
        return OrderedDict((k, v.dict() if isinstance(v, Config) else v) for k, v in self.items())

    # -- state
    def is_valid(self) -> bool:
        ok = all(self.get(k) is not None for k in self._required)
        return ok and all(v.is_valid() for v in self.values() if isinstance(v, Config))

    def freeze(self) -> None:
        object.__setattr__(self, '_frozen', True)
        for v in self.values():
            if isinstance(v, Config):
                v.freeze()

    def defrost(self) -> None:
        object.__setattr__(self, '_frozen', False)
        for v in self.values():
            if isinstance(v, Config):
                v.defrost()

    def is_frozen(self) -> bool:
        return self._frozen

    def copy(self, defrost: bool = True) -> 'Config':
        new = Config._from_dict(self.dict())
        object.__setattr__(new, '_dynamic', self._dynamic)
        if not defrost and self._frozen:
            new.freeze()
        return new

    def __reduce__(self):
        return (Config._from_dict, (self.dict(),))
