"""acceleratedvolrenderer_b200 — B200-native volumetric hot path of tsvdh/AcceleratedVolRenderer.

The product is libavr_b200.so (hand-written sm_100a CUDA behind the C ABI of include/avr.h).  This
package is the thin Python host mirror used by tests and bench: it loads the library with ctypes
and fails loudly if it is missing — there is no CPU fallback anywhere in the package.
"""
from ._abi import NSPEC, GridDesc, NanoVDBDesc, NanoVDBGrid, RGBGridDesc, RenderDesc  # noqa: F401

__all__ = ["NSPEC", "GridDesc", "NanoVDBDesc", "NanoVDBGrid", "RGBGridDesc", "RenderDesc", "load_library", "Context"]


def load_library():
    from .api import load_library as _l
    return _l()


def __getattr__(name):
    if name in ("Context", "Medium", "Graph"):
        from . import api
        return getattr(api, name)
    raise AttributeError(name)
