"""zero-tig-b200: B200-native (sm_100a) implementation of the Zero-TIG per-frame-pair hot
path (RAFT flow -> backward warp -> illumination-guided enhance/denoise nets) behind the
reference's ``model/model.py`` interface.  See DESIGN.md / INTEGRATION.md at the repo root.
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401  (ctypes binding; loading is lazy)


def build(force=False, verbose=False):
    """Compile csrc/*.cu into csrc/libzerotig.so with nvcc for sm_100a."""
    from .build import build as _b
    return _b(force=force, verbose=verbose)
