"""video_tokenizer_b200 -- B200-native quantization bottleneck of the LARP video tokenizer.

Drop-in for the hot path of zhxie0117/video-tokenizer (models/bottleneck.py as driven by
models/larp_tokenizer.py and trainers/larp_tokenizer_trainer.py): hand-written sm_100a CUDA
behind a C ABI (include/b200vq.h, libb200vq.so), bound with ctypes, wrapped in the
reference's own module API and registry names.  Importing this package does not need a
GPU; calling it does (there is no CPU fallback).
"""
from . import _lib, registry
from .bottleneck import AbstractRegularizer, Bottleneck, SimpleVectorQuantizer, entropy_loss
from .graphed import GraphedBottleneckStep
from .functional import codebook_lookup, prepare_codebook, usage_stats, vq_quantize
from .registry import install, make, register
from .simvq import SimVQ
from .sq import VectorQuantizer

__all__ = ["Bottleneck", "SimpleVectorQuantizer", "VectorQuantizer", "SimVQ", "AbstractRegularizer", "entropy_loss", "vq_quantize",
           "codebook_lookup", "prepare_codebook", "usage_stats", "install", "make", "register", "registry",
           "load_library", "GraphedBottleneckStep"]
__version__ = "0.1.0"


def load_library():
    """Load libb200vq.so (building it in-tree if nvcc is available and sources changed)."""
    return _lib.load()
