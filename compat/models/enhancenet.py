"""``models.enhancenet.EnhanceNet`` (models/enhancenet.py:7-145) -> the B200 drop-in.  Defined as a subclass HERE so that
its pickled class path is ``models.enhancenet.EnhanceNet`` in both directions: reference checkpoints load into it, and
checkpoints written by a script running on the shim load in the stock reference (same module tree and attributes)."""
from isosurfacesuperresolution_b200.enhancenet import EnhanceNet as _EnhanceNet


class EnhanceNet(_EnhanceNet):
    pass
