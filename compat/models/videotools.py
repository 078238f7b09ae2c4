"""``models.videotools.VideoTools`` (models/videotools.py:6-87) -> the B200 drop-in."""
from isosurfacesuperresolution_b200.videotools import VideoTools  # noqa: F401
