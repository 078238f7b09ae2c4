"""Drop-in replacement of the reference's ``models.videotools.VideoTools``
(/root/reference/SuperresolutionNetwork/models/videotools.py:6-87).  CUDA tensors without autograd run the libisr.so
kernels; grad-enabled calls (training back-propagates through the warp, mainVideoUnshaded.py:429-433,467) use torch ops.
"""
import torch
import torch.nn.functional as F

from . import ops


class VideoTools:
    @staticmethod
    def flatten_high(image_high, upscale_factor):
        """B x C x H*r x W*r -> B x C*r*r x H x W (inverse PixelShuffle), videotools.py:8-25."""
        if torch.is_grad_enabled() and image_high.requires_grad:
            return F.pixel_unshuffle(image_high, upscale_factor)
        if not image_high.is_cuda:
            raise RuntimeError("VideoTools.flatten_high (B200 build): needs a CUDA tensor; there is no CPU fallback")
        return ops.flatten_high(image_high, upscale_factor)

    @staticmethod
    def warp_upscale(image_high, flow_low, upscale_factor, special_mask=False, align_corners=False):
        """videotools.py:49-87.  ``align_corners=False`` is what the reference does under the installed torch."""
        B, C, H, W = image_high.shape
        assert flow_low.shape[1] == 2 and H == flow_low.shape[2] * upscale_factor \
            and W == flow_low.shape[3] * upscale_factor                                   # videotools.py:63
        if torch.is_grad_enabled() and (image_high.requires_grad or flow_low.requires_grad):
            return VideoTools._warp_autograd(image_high, flow_low, upscale_factor, special_mask, align_corners)
        if not image_high.is_cuda:
            raise RuntimeError("VideoTools.warp_upscale (B200 build): needs CUDA tensors; there is no CPU fallback")
        return ops.warp_upscale(image_high, flow_low, upscale_factor, special_mask, align_corners)

    @staticmethod
    def _warp_autograd(image_high, flow_low, r, special_mask, align_corners):
        B, C, H, W = image_high.shape
        f2 = torch.cat((flow_low[:, 0:1] * -2.0, flow_low[:, 1:2] * 2.0), dim=1)
        fh = F.interpolate(f2, scale_factor=r, mode="bilinear").permute(0, 2, 3, 1)
        gh = torch.linspace(-1, 1, H, device=image_high.device, dtype=image_high.dtype).unsqueeze(1)
        gw = torch.linspace(-1, 1, W, device=image_high.device, dtype=image_high.dtype).unsqueeze(0)
        grid = torch.stack(torch.broadcast_tensors(gw, gh), dim=2).unsqueeze(0) + fh
        img = image_high
        if special_mask:
            img = torch.cat([img[:, 0:1] * 0.5 + 0.5, img[:, 1:]], dim=1)
        out = F.grid_sample(img, grid, mode="bilinear", padding_mode="zeros", align_corners=align_corners)
        if special_mask:
            out = torch.cat([out[:, 0:1] * 2 - 1, out[:, 1:]], dim=1)
        return out
