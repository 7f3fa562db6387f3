"""Configuration objects with the reference's field names and defaults.

``ConfigDict`` is a small attribute-access nested dict (the role ``ml_collections.ConfigDict``
plays in data_paral.py:39-72 and param_sharding.py:32-55,244-246); ``dp_config()`` /
``fsdp_config()`` return the reference's literal values. ``apply_overrides`` adds the
``--key=value`` overrides the reference lacks.
"""
from __future__ import annotations

import copy
from typing import Any, Iterable

import torch


class ConfigDict(dict):
    def __init__(self, initial=None, **kw):
        super().__init__()
        for k, v in dict(initial or {}, **kw).items():
            self[k] = v

    def __setitem__(self, k, v):
        if isinstance(v, dict) and not isinstance(v, ConfigDict):
            v = ConfigDict(v)
        super().__setitem__(k, v)

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(f"config has no field {k!r}; fields: {sorted(self)}") from e

    def __setattr__(self, k, v):
        self[k] = v

    def copy_and_resolve(self) -> "ConfigDict":
        return copy.deepcopy(self)

    def to_dict(self) -> dict:
        return {k: (v.to_dict() if isinstance(v, ConfigDict) else v) for k, v in self.items()}


def data_config() -> ConfigDict:
    return ConfigDict(batch_size=128, num_classes=10, input_size=784)


def dp_config() -> ConfigDict:
    """data_paral.py:39-72."""
    data = data_config()
    model = ConfigDict(hidden_size=512, dropout_rate=0.1, dtype=torch.bfloat16,
                       num_classes=data.num_classes, data_axis_name="data")
    optimizer = ConfigDict(learning_rate=1e-3, num_minibatches=4)
    return ConfigDict(model=model, optimizer=optimizer, data=data,
                      data_axis_name=model.data_axis_name, seed=69)


def fsdp_config() -> ConfigDict:
    """param_sharding.py:32-55 plus the ``min_weight_size = 2**4`` mutation at :244-246."""
    data = data_config()
    model = ConfigDict(hidden_size=512, dropout_rate=0.1, dtype=torch.bfloat16,
                       num_classes=data.num_classes, data_axis_name="data", lr=1e-4,
                       min_weight_size=2 ** 4)
    return ConfigDict(model=model, data=data, seed=6969, num_minibatches=4,
                      data_axis_name=model.data_axis_name)


def _parse(v: str) -> Any:
    for cast in (int, float):
        try:
            return cast(v)
        except ValueError:
            pass
    if v.lower() in ("true", "false"):
        return v.lower() == "true"
    return v


def apply_overrides(cfg: ConfigDict, overrides: Iterable[str]) -> ConfigDict:
    """``--model.hidden_size=1024`` style overrides (dotted path into the nested config)."""
    for o in overrides:
        o = o.lstrip("-")
        if "=" not in o:
            raise ValueError(f"override {o!r} is not key=value")
        path, val = o.split("=", 1)
        node = cfg
        *parents, last = path.split(".")
        for p in parents:
            node = node[p]
        if last not in node:
            raise KeyError(f"unknown config field {path!r}")
        node[last] = _parse(val)
    return cfg
