"""Shim of the reference's ``inference`` package (inference/__init__.py): LoadedModel is the B200 drop-in; Camera,
Orientation, Material and the GVDB renderers fall through to the reference tree when it is on sys.path (the renderers
need the reference's Windows DLL; ``DenseGridRenderer`` is this repo's dense-brick stand-in with the same interface)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _overlay import overlay  # noqa: E402

sys.path.pop(0)
overlay("inference", __path__)

from .loadedmodel import LoadedModel  # noqa: E402
from isosurfacesuperresolution_b200.renderer import DenseGridRenderer  # noqa: E402,F401

try:
    from .camera import Camera, Orientation  # noqa: F401
except ImportError:
    pass
try:
    from .renderer import Material, Renderer, DirectRenderer, DirectVolumeRenderer  # noqa: F401
except (ImportError, OSError):
    pass
