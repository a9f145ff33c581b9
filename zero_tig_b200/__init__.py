"""Import shim: the product package lives in the directory ``zero-tig_b200/`` (a
hyphen is not a legal Python identifier), so ``import zero_tig_b200`` resolves here
and forwards to that directory."""
import os as _os

_real = _os.path.normpath(_os.path.join(_os.path.dirname(_os.path.abspath(__file__)), "..", "zero-tig_b200"))
__path__ = [_real]
__file__ = _os.path.join(_real, "__init__.py")
with open(__file__, "r") as _f:
    exec(compile(_f.read(), __file__, "exec"))
