"""``inference.loadedmodel.LoadedModel`` (inference/loadedmodel.py:8-121) -> the B200 drop-in."""
from isosurfacesuperresolution_b200.inference import LoadedModel  # noqa: F401
