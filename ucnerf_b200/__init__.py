"""Import alias for the product package, whose directory is named ``uc-nerf_b200`` (not a valid
Python identifier).  ``import ucnerf_b200`` resolves every submodule from ``uc-nerf_b200/``."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "uc-nerf_b200")
if not _os.path.isdir(_real):  # pragma: no cover
    raise ImportError(f"product package directory missing: {_real}")
__path__.append(_real)
__version__ = "0.1.0"
