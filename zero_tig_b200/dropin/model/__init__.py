"""Stand-in for the reference's ``model`` package (model/__init__.py is empty upstream).

Self-locating: the real implementation is the ``zero_tig_b200`` package two directories up; its parent
is appended to ``sys.path`` when the package is not importable yet (PYTHONPATH-only use)."""
import importlib.util as _u
import os as _os
import sys as _sys

if _u.find_spec("zero_tig_b200") is None:
    _sys.path.append(_os.path.normpath(_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "..", "..", "..")))
