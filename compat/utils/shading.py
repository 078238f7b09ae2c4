"""``utils.shading.ScreenSpaceShading`` (utils/shading.py:7-207) -> the B200 drop-in."""
from isosurfacesuperresolution_b200.shading import ScreenSpaceShading  # noqa: F401
