"""Test / benchmark harness: seeded synthetic scenes, cameras and hierarchies,
and loaders for the reference builds under oracle/_ref.  Not part of the product path."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "street-sparse-3dgs_b200")


def add_paths():
    """Make the drop-in packages importable under the reference's own names."""
    if PKG not in sys.path:
        sys.path.insert(0, PKG)
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
