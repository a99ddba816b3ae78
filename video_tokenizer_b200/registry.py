"""Model registry with the reference's plug-in contract (models/models.py:1-27).

``register(name)`` / ``make(spec, args, load_sd)`` behave exactly like the reference's:
``make`` deep-copies ``spec['args']``, overlays ``args``, drops keys the constructor does
not name unless it takes ``**kwargs``, and optionally loads ``spec['sd']`` strictly.

``install(models_module)`` is the drop-in switch: it re-registers ``"vq"`` and
``"bottleneck"`` in the REFERENCE's registry dict (plain dict assignment, later
registration wins), so ``models.make({'name': 'bottleneck', ...})`` inside
``LARPTokenizer.__init__`` (models/larp_tokenizer.py:216-217) builds the B200 modules with
zero edits to the reference.
"""
from __future__ import annotations

import copy
import inspect

import torch

models: dict = {}


def register(name):
    def decorator(cls):
        models[name] = cls
        return cls
    return decorator


def make(model_spec, args=None, load_sd=False) -> torch.nn.Module:
    if args is not None:
        model_args = copy.deepcopy(model_spec["args"])
        model_args.update(args)
    else:
        model_args = model_spec["args"]
    cls = models[model_spec["name"]]
    params = inspect.signature(cls).parameters
    if "kwargs" not in params:
        model_args = {k: v for k, v in model_args.items() if k in params}
    model = cls(**model_args)
    if load_sd:
        model.load_state_dict(model_spec["sd"], strict=True)
    return model


def install(reference_models_module) -> None:
    """Swap the reference's "vq" / "bottleneck" entries for the B200 implementations.
    ``reference_models_module`` is the reference's ``models`` package (or its
    ``models.models`` module): anything exposing the registry dict as ``.models``."""
    from . import bottleneck as _b  # noqa: F401  (registers into our own dict)
    reg = getattr(reference_models_module, "models", None)
    if isinstance(reg, dict):
        target = reg
    elif hasattr(reg, "models") and isinstance(reg.models, dict):
        target = reg.models
    else:
        raise TypeError("expected the reference `models` package or its `models.models` module")
    target["vq"] = models["vq"]
    target["bottleneck"] = models["bottleneck"]
    # the B200 Bottleneck resolves its regularizer through the same (reference) registry
    _b._external_make = getattr(reference_models_module, "make", None) or getattr(reg, "make", None)
