"""Shim of the reference's ``models`` package (models/__init__.py:16-49): EnhanceNet, VideoTools and createNetwork are the
B200 drop-ins; the other generators fall through to the reference tree when it is on sys.path (compat/_overlay.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from _overlay import overlay  # noqa: E402

sys.path.pop(0)
overlay("models", __path__)

from .enhancenet import EnhanceNet  # noqa: E402
from .videotools import VideoTools  # noqa: E402

_OTHERS = {}
for _mod, _cls in (("subpixelnet", "SubpixelNet"), ("tecogan", "TecoGAN"), ("rcan", "RCAN")):
    try:
        _OTHERS[_cls.lower()] = getattr(__import__(__name__ + "." + _mod, fromlist=[_cls]), _cls)
        globals()[_cls] = _OTHERS[_cls.lower()]
    except ImportError:
        pass


def createNetwork(name, upscale_factor, input_channels, channel_mask, output_channels, additional_opt):
    """models.createNetwork (models/__init__.py:21-49)."""
    if name.lower() == "enhancenet":
        return EnhanceNet(upscale_factor, input_channels, channel_mask, output_channels, additional_opt)
    if name.lower() in _OTHERS:                      # outside the B200 hot path: the reference's own class
        return _OTHERS[name.lower()](upscale_factor, input_channels, channel_mask, output_channels, additional_opt)
    if name.lower() in ("subpixelnet", "tecogan", "rcan"):
        raise NotImplementedError(name + " is outside the B200 hot path and the reference tree is not on sys.path")
    raise ValueError("Unknown model %s" % name)
