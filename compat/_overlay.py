"""Shared helper of the compat/ shim packages (models, utils, inference).

``sys.path.insert(0, "<repo>/compat")`` in front of the reference's SuperresolutionNetwork directory is the ONLY change a
reference script needs: the packages here carry the reference's package names, so ``import models``, ``from utils import
ScreenSpaceShading, initialImage`` and ``from inference import LoadedModel`` resolve to the B200 drop-ins, and whole-module
checkpoints un-pickle ``models.enhancenet.EnhanceNet`` (inference/loadedmodel.py:16-22, written at
mainVideoUnshaded.py:799-811) into the drop-in class.

Everything OUTSIDE the hot path (SubpixelNet / TecoGAN / RCAN, SSIM / PSNR / MeanVariance, Camera / Renderer ...) is
not re-implemented: each shim package appends the reference's directory of the same name -- if one is found further
down ``sys.path`` -- to its ``__path__``, so ``from .ssim import SSIM`` falls through to the reference's own file."""
import os
import sys

_COMPAT = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_COMPAT)
if _REPO not in sys.path:
    sys.path.append(_REPO)          # for `import isosurfacesuperresolution_b200`


def overlay(package_name, package_path):
    """Appends <entry>/<package_name> of later sys.path entries (the reference tree) to the shim package's __path__."""
    for entry in sys.path:
        cand = os.path.join(os.path.abspath(entry or "."), package_name)
        if os.path.abspath(entry or ".") == _COMPAT or cand in package_path:
            continue
        if os.path.isfile(os.path.join(cand, "__init__.py")):
            package_path.append(cand)
    return package_path
