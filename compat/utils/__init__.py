"""Shim of the reference's ``utils`` package (utils/__init__.py): ScreenSpaceShading and initialImage are the B200
drop-ins; MeanVariance / SSIM / MSSSIM / PSNR fall through to the reference tree when it is on sys.path."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _overlay import overlay  # noqa: E402

sys.path.pop(0)
overlay("utils", __path__)

from .shading import ScreenSpaceShading  # noqa: E402
from .initial_image import initialImage  # noqa: E402

for _mod, _names in (("mv", ("MeanVariance",)), ("ssim", ("SSIM", "MSSSIM")), ("psnr", ("PSNR",))):
    try:
        _m = __import__(__name__ + "." + _mod, fromlist=list(_names))
        for _n in _names:
            globals()[_n] = getattr(_m, _n)
    except ImportError:
        pass
