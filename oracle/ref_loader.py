"""Import the UNMODIFIED reference ``models/bottleneck.py`` (build container only).

TEST INFRASTRUCTURE.  ``/root/reference`` does not exist on the GPU box; there the two files
come from ``baseline/_ref`` (staged by ``oracle/stage_ref.py``, git-ignored, shipped with the
snapshot).  Used by ``oracle/make_golden.py`` to produce ``tests/golden/*.npz``, by the
cross-check in ``tests/test_oracle_golden.py`` and by the CPU legs of ``bench.py`` (the
reference arm times this module itself: ``cpu_baseline.kind = "reference"``).

``import models`` on the reference fails (timm/easydict/vjepa2 absent),
so the two files the path needs are executed under a stub ``models`` package that
exposes only ``register`` / ``make`` from ``models/models.py``.  No reference file is
copied or modified.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types


def reference_root() -> str | None:
    staged = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")
    for cand in (os.environ.get("VT_REFERENCE_ROOT"), "/root/reference", staged):
        if cand and os.path.isfile(os.path.join(cand, "models", "bottleneck.py")):
            return cand
    return None


def load_reference():
    """Returns (bottleneck_module, registry_module) of the reference, or raises."""
    root = reference_root()
    if root is None:
        raise FileNotFoundError("reference tree not available (expected on the build container only)")
    if "ref_models_shell" in sys.modules:
        shell = sys.modules["ref_models_shell"]
        return shell._bottleneck, shell._registry

    saved = {k: sys.modules.get(k) for k in ("models", "models.models", "models.bottleneck")}

    def _exec(name, path):
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod

    try:
        registry = _exec("models.models", os.path.join(root, "models", "models.py"))
        pkg = types.ModuleType("models")
        pkg.__path__ = []
        pkg.register, pkg.make, pkg.models = registry.register, registry.make, registry.models
        sys.modules["models"] = pkg
        bott = _exec("models.bottleneck", os.path.join(root, "models", "bottleneck.py"))
    finally:
        # leave no stub 'models' package behind for unrelated importers
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    shell = types.ModuleType("ref_models_shell")
    shell._bottleneck, shell._registry = bott, registry
    sys.modules["ref_models_shell"] = shell
    return bott, registry
