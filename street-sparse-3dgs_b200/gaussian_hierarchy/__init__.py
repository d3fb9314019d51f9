"""Drop-in for the reference package ``gaussian_hierarchy`` (GH/setup.py:17-35, GH/ext.cpp:15-21)."""
from . import _C  # noqa: F401
