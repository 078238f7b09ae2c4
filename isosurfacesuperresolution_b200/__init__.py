"""isosurfacesuperresolution_b200 -- B200 (sm_100a) implementation of the frame-recurrent 4x EnhanceNet inference path of
shamanDevel/IsosurfaceSuperresolution, behind the reference's own Python API:

    createNetwork / EnhanceNet   (models/__init__.py:21-49, models/enhancenet.py)
    VideoTools                   (models/videotools.py)
    initialImage                 (utils/initial_image.py)
    ScreenSpaceShading           (utils/shading.py)
    LoadedModel                  (inference/loadedmodel.py)

The arithmetic lives in libisr.so (csrc/, C ABI in include/isr.h); importing this package needs the built library.
"""
from ._lib import lib as _load_lib

_load_lib()   # fail loudly when the CUDA extension has not been built

from .enhancenet import EnhanceNet          # noqa: E402
from .initial_image import initialImage     # noqa: E402
from .shading import ScreenSpaceShading     # noqa: E402
from .videotools import VideoTools          # noqa: E402
from .inference import LoadedModel, SequenceRunner, compose_frame, temporal_smoothing   # noqa: E402
from .dataset import SequenceDataset, evaluate, masked_psnr           # noqa: E402
from .renderer import DenseGridRenderer, camera_matrices              # noqa: E402


def createNetwork(name, upscale_factor, input_channels, channel_mask, output_channels, additional_opt):
    """models.createNetwork (models/__init__.py:21-49); only EnhanceNet is on the accelerated path."""
    if name.lower() == "enhancenet":
        return EnhanceNet(upscale_factor, input_channels, channel_mask, output_channels, additional_opt)
    if name.lower() in ("subpixelnet", "tecogan", "rcan"):
        raise NotImplementedError(name + " is outside the B200 hot path (SURVEY.md section 8); use the reference model")
    raise ValueError("Unknown model %s" % name)


__all__ = ["createNetwork", "EnhanceNet", "VideoTools", "initialImage", "ScreenSpaceShading", "LoadedModel",
           "SequenceRunner", "compose_frame", "temporal_smoothing", "SequenceDataset", "evaluate", "masked_psnr"]
