"""Import stub for `from ipdb.__main__ import set_trace` (utils.py:3)."""
from . import set_trace  # noqa: F401
