"""Config-2 harness: the UNMODIFIED reference `LARPTokenizer` (models/larp_tokenizer.py) under stub dependencies.

TEST INFRASTRUCTURE.  `import models` on the reference needs timm, easydict, vjepa2, flash_attn,
lpips, ... none of which exist in this image.  This loader pre-seeds `sys.modules` with minimal shells for exactly
the names `models/larp_tokenizer.py` imports at module level, exposes the reference's real registry
(`models/models.py`) through a `models` shell package whose `__path__` is the reference's `models/` directory
(so `models/__init__.py` is never executed), and imports `models.transformer`, `models.bottleneck`,
`models.larp_tokenizer` from the reference tree unmodified.

The reference tree is found at $VT_REFERENCE_ROOT, /root/reference (build container) or baseline/_ref
(staged by oracle/stage_ref.py; the only copy the GPU box has).  Nothing here is imported by the package.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def reference_root(need=("models/larp_tokenizer.py", "models/transformer.py", "models/bottleneck.py")):
    for cand in (os.environ.get("VT_REFERENCE_ROOT"), "/root/reference", os.path.join(ROOT, "baseline", "_ref")):
        if cand and all(os.path.isfile(os.path.join(cand, f)) for f in need):
            return cand
    return None


# ---- stubs -----------------------------------------------------------------------------------------
class _EasyDict(dict):
    """attribute-access dict (easydict.EasyDict as used at models/larp_tokenizer.py:232)"""

    def __init__(self, d=None, **kw):
        super().__init__()
        for k, v in dict(d or {}, **kw).items():
            self[k] = v

    def __setitem__(self, k, v):
        if isinstance(v, dict) and not isinstance(v, _EasyDict):
            v = _EasyDict(v)
        super().__setitem__(k, v)

    __setattr__ = __setitem__

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:
            raise AttributeError(k) from e


class _Attention(nn.Module):
    def __init__(self, dim, num_heads, qkv_bias, attn_drop, proj_drop):
        super().__init__()
        self.num_heads = num_heads
        self.qkv = nn.Linear(dim, dim * 3, bias=qkv_bias)
        self.proj = nn.Linear(dim, dim)
        self.attn_drop, self.proj_drop = attn_drop, proj_drop

    def forward(self, x):
        B, n, C = x.shape
        q, k, v = self.qkv(x).reshape(B, n, 3, self.num_heads, C // self.num_heads).permute(2, 0, 3, 1, 4).unbind(0)
        x = F.scaled_dot_product_attention(q, k, v, dropout_p=self.attn_drop if self.training else 0.0)
        return F.dropout(self.proj(x.transpose(1, 2).reshape(B, n, C)), self.proj_drop, self.training)


class _Block(nn.Module):
    """timm.models.vision_transformer.Block as the reference builds it (models/transformer.py:17-24):
    pre-norm LayerNorm -> MHA -> LayerNorm -> GELU MLP, residual adds."""

    def __init__(self, dim, num_heads, mlp_ratio=4.0, qkv_bias=False, proj_drop=0.0, attn_drop=0.0, **_):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.attn = _Attention(dim, num_heads, qkv_bias, attn_drop, proj_drop)
        self.norm2 = nn.LayerNorm(dim)
        hidden = int(dim * mlp_ratio)
        self.mlp = nn.Sequential(nn.Linear(dim, hidden), nn.GELU(), nn.Linear(hidden, dim))

    def forward(self, x):
        x = x + self.attn(self.norm1(x))
        return x + self.mlp(self.norm2(x))


class _PatchEmbed(nn.Module):
    """timm PatchEmbed surface that models/embed.py:18-35 subclasses (2-D patches, flattened tokens)."""

    def __init__(self, img_size=224, patch_size=16, in_chans=3, embed_dim=768, norm_layer=None, flatten=True, bias=True, **_):
        super().__init__()
        self.img_size, self.patch_size, self.flatten = (img_size, img_size), (patch_size, patch_size), flatten
        self.grid_size = (img_size // patch_size, img_size // patch_size)
        self.num_patches = self.grid_size[0] * self.grid_size[1]
        self.proj = nn.Conv2d(in_chans, embed_dim, kernel_size=patch_size, stride=patch_size, bias=bias)
        self.norm = norm_layer(embed_dim) if norm_layer else nn.Identity()

    def forward(self, x):
        x = self.proj(x)
        if self.flatten:
            x = x.flatten(2).transpose(1, 2)
        return self.norm(x)


def _shell(name, **attrs):
    m = types.ModuleType(name)
    m.__path__ = []   # behaves as a package for dotted imports
    for k, v in attrs.items():
        setattr(m, k, v)
    return m


def _flash_attn_func(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False, **_):
    # (b, s, h, d) layout like flash_attn.flash_attn_func
    o = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), dropout_p=dropout_p,
                                       is_causal=causal, scale=softmax_scale)
    return o.transpose(1, 2)


_STUBS = {
    "easydict": lambda: _shell("easydict", EasyDict=_EasyDict),
    "timm": lambda: _shell("timm"),
    "timm.models": lambda: _shell("timm.models"),
    "timm.models.vision_transformer": lambda: _shell("timm.models.vision_transformer", Block=_Block, PatchEmbed=_PatchEmbed),
    "vjepa2": lambda: _shell("vjepa2"),
    "vjepa2.src": lambda: _shell("vjepa2.src"),
    "vjepa2.src.models": lambda: _shell("vjepa2.src.models"),
    "vjepa2.src.models.vision_transformer": lambda: _shell("vjepa2.src.models.vision_transformer", vit_large_rope=None),
    "vjepa2.src.datasets": lambda: _shell("vjepa2.src.datasets"),
    "vjepa2.src.datasets.utils": lambda: _shell("vjepa2.src.datasets.utils"),
    "vjepa2.src.datasets.utils.video": lambda: _shell("vjepa2.src.datasets.utils.video"),
    "vjepa2.src.datasets.utils.video.transforms": lambda: _shell("vjepa2.src.datasets.utils.video.transforms"),
    "vjepa2.src.datasets.utils.video.volume_transforms": lambda: _shell("vjepa2.src.datasets.utils.video.volume_transforms"),
    "flash_attn": lambda: _shell("flash_attn", flash_attn_func=_flash_attn_func),
}

_loaded = None


def load_tokenizer_package():
    """-> the reference `models` shell package (attributes: register, make, models = the registry dict,
    larp_tokenizer, bottleneck, transformer submodules).  Raises FileNotFoundError when no reference tree exists."""
    global _loaded
    if _loaded is not None:
        return _loaded
    root = reference_root()
    if root is None:
        raise FileNotFoundError("no reference tree (VT_REFERENCE_ROOT, /root/reference, baseline/_ref): run "
                                "`python -m oracle.stage_ref` in the build container")
    for name, mk in _STUBS.items():
        if name == "flash_attn":
            try:   # a real flash_attn serves as well; the tokenizer path under test never calls it
                importlib.import_module("flash_attn")
                continue
            except Exception:
                pass
        if name not in sys.modules or name != "flash_attn":
            sys.modules[name] = mk()
    for a, b in (("timm", "models"), ("timm.models", "vision_transformer")):
        setattr(sys.modules[a], b, sys.modules[a + "." + b])
    if "utils" not in sys.modules or not hasattr(sys.modules["utils"], "__b200vq_stub__"):
        u = _shell("utils")
        u.__b200vq_stub__ = True
        sys.modules["utils"] = u
    for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
        del sys.modules[k]
    spec = importlib.util.spec_from_file_location("models.models", os.path.join(root, "models", "models.py"))
    registry = importlib.util.module_from_spec(spec)
    sys.modules["models.models"] = registry
    spec.loader.exec_module(registry)
    pkg = types.ModuleType("models")
    pkg.__path__ = [os.path.join(root, "models")]
    pkg.register, pkg.make, pkg.models = registry.register, registry.make, registry.models
    sys.modules["models"] = pkg
    for sub in ("transformer", "bottleneck", "larp_tokenizer"):
        setattr(pkg, sub, importlib.import_module("models." + sub))
    pkg.__reference_root__ = root
    _loaded = pkg
    return pkg


def bottleneck_spec(dim=24, codebook_size=8192, stochastic=True, temperature=0.03):
    """the `bottleneck` block of cfgs/larp_tokenizer.yaml:39-54"""
    return {"name": "bottleneck", "args": {"bottleneck_dim": dim, "norm": "none", "regularizer": {"name": "vq", "args": {
        "codebook_size": codebook_size, "commitment_loss_weight": 0.25, "codebook_loss_weight": 1.0,
        "entropy_loss_weight": 0.0, "entropy_loss_temperature": 0.01, "l2_normalized": True, "stochastic": stochastic,
        "stochastic_temperature": temperature}}}}


def build_tokenizer(pkg, *, seed=0, depth=6, hidden=768, heads=12, input_size=128, frame_num=16, token_num=1024,
                    patch=16, temporal_patch=2, bottleneck=None):
    """LARPTokenizer(bottleneck_type='vq') on 16 x 128 x 128 clips (BASELINE configs[1]): patch 16 / temporal 2 ->
    512 context tokens + 1024 latent queries.  `final_layer.linear` is zero-initialised by the
    reference (models/larp_tokenizer.py:326-328), which would make the reconstruction gradient into the bottleneck
    exactly 0, so it is re-randomised here (seeded)."""
    torch.manual_seed(seed)
    tok = pkg.larp_tokenizer.LARPTokenizer(
        bottleneck=bottleneck or bottleneck_spec(), prior_model={"name": "none"}, bottleneck_token_num=token_num,
        input_size=input_size, frame_num=frame_num, temporal_patch_size=temporal_patch, patch_size=patch,
        decoder_temporal_patch_size=temporal_patch, decoder_patch_size=patch, bottleneck_type="vq",
        encoder_hidden_size=hidden, decoder_hidden_size=hidden, encoder_num_heads=heads, decoder_num_heads=heads,
        encoder_depth=depth, decoder_depth=depth)
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        tok.final_layer.linear.weight.copy_(torch.randn(tok.final_layer.linear.weight.shape, generator=g) * 0.02)
    return tok
