"""``utils.initial_image.initialImage`` (utils/initial_image.py:5-54) -> the B200 drop-in."""
from isosurfacesuperresolution_b200.initial_image import initialImage  # noqa: F401
