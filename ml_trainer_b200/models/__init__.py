"""Model zoo + registry (``--model`` flag of main.py / bench.py)."""
from __future__ import annotations

from .mlmodel import MLModel

_REGISTRY = {"mlmodel": lambda **kw: MLModel(**kw)}


def register(name):
    def deco(fn):
        _REGISTRY[name] = fn
        return fn
    return deco


def build_model(name: str, **kw):
    name = name.lower()
    name = {"resnet-50": "resnet50", "bert-base": "bert_base", "bert": "bert_base"}.get(name, name)
    if name not in _REGISTRY:                  # the larger model families register themselves on import
        import importlib

        for mod in ("bert", "resnet"):
            if name.startswith(mod):
                try:
                    importlib.import_module(f"{__name__}.{mod}")
                except ImportError:
                    pass
    if name not in _REGISTRY:
        raise KeyError(f"unknown model {name!r}; known: {sorted(_REGISTRY)}")
    return _REGISTRY[name](**kw)


def available_models():
    return sorted(set(_REGISTRY) | {"resnet50", "bert_base"})


__all__ = ["MLModel", "build_model", "available_models", "register"]
