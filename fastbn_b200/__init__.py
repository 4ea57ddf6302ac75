"""fastbn_b200 — B200-native junction-tree exact inference (the `-a 2` hot path of Tonanguyxiro/FastBN).

The compute path is the in-tree CUDA library fastbn_b200/libfastbn_b200.so (sm_100a kernels behind the C ABI of
include/fastbn_b200.h). There is no CPU fallback: importing works without a GPU (so the ABI can be inspected), but any
compute call raises when the library or a CUDA device is missing."""
from . import capi  # noqa: F401
from .capi import Plan, FastBNError, load_library, tree_from_npz  # noqa: F401

__all__ = ["capi", "Plan", "FastBNError", "load_library", "tree_from_npz"]
