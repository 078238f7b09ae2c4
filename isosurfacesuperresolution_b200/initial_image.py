"""Drop-in replacement of the reference's ``utils.initial_image.initialImage``
(/root/reference/SuperresolutionNetwork/utils/initial_image.py:5-54)."""
from . import ops


def initialImage(current_input, channels, mode, aoInverted, upscaling=4):
    """Frame-0 substitute for the previous high-res output: 'zero', 'unshaded' or 'input'."""
    if mode not in ("zero", "unshaded", "input"):
        raise ValueError("unknown input mode: " + str(mode))
    if not current_input.is_cuda:
        raise RuntimeError("initialImage (B200 build): needs a CUDA tensor; there is no CPU fallback")
    return ops.initial_image(current_input, channels, mode, aoInverted, upscaling)
