"""path-tracer_b200 — host-side mirror of the reference's render interface over the C ABI.

Filled in below by `api.py`; this file only re-exports.
"""
from . import abi  # noqa: F401
try:
    from .api import *  # noqa: F401,F403
except ImportError:  # api.py is added with the CUDA library
    pass
