"""Config plumbing with the reference's semantics (util/util.py:8-54): nested dict configs whose
``'module'`` key is a dotted path that is imported and called with the remaining keys as kwargs.

Reference module paths are accepted unchanged: ``network.*`` / ``instancer.instancer.*`` /
``util.*`` are mapped onto this package, so shipped config files work without edits.
"""

from __future__ import annotations

import importlib
import math
from typing import Any

_ALIASES = {
    "network.model.": "nerf_tex_b200.model.",
    "network.layer.": "nerf_tex_b200.layer.",
    "network.renderer.": "nerf_tex_b200.renderer.",
    "network.ray_sampler.": "nerf_tex_b200.ray_sampler.",
    "network.proxy.": "nerf_tex_b200.proxy.",
    "network.loss.": "nerf_tex_b200.loss.",
    "network.render.": "nerf_tex_b200.render.",
    "network.pixel_sampler.": "nerf_tex_b200.pixel_sampler.",
    "network.dataset.": "nerf_tex_b200.dataset.",
    "network.logger.": "nerf_tex_b200.logger.",
    "network.train.": "nerf_tex_b200.train.",
    "instancer.instancer.": "nerf_tex_b200.instancer.",
}


class EasyDict(dict):
    """dict with attribute access, nested (util/util.py:8-29)."""

    def __init__(self, other: dict = ()) -> None:
        super().__init__()
        for key in dict(other):
            value = other[key]
            if isinstance(value, dict):
                value = EasyDict(value)
            self[key] = value

    def __getattr__(self, key: str) -> Any:
        try:
            return self[key]
        except KeyError:
            raise AttributeError(key)

    def __setattr__(self, key: str, value: Any) -> None:
        self[key] = value

    def __delattr__(self, key: str) -> None:
        del self[key]


def resolve_path(path: str) -> str:
    for old, new in _ALIASES.items():
        if path.startswith(old):
            return new + path[len(old):]
    return path


def get_attr_from_module(module_name: str, attr_name: str) -> Any:
    return getattr(importlib.import_module(module_name), attr_name)


def get_attr_from_path(path: str) -> Any:
    """util/util.py:36-42."""
    path = resolve_path(path)
    module_name, _, attr_name = path.rpartition(".")
    return get_attr_from_module(module_name, attr_name)


def instantiate(config) -> Any:
    """util/util.py:44-54."""
    if config is None:
        return None
    args = EasyDict(config)
    module = args.module
    del args.module
    return get_attr_from_path(module)(**args)


def format_name(prefix: str, idx: int, max_idx: int, suffix: str) -> str:
    """util/util.py:56-62."""
    n_chars = math.ceil(math.log10(max_idx + 1))
    return prefix + ("{:0" + str(n_chars) + "d}").format(idx) + suffix
