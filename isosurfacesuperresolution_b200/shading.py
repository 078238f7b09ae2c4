"""Drop-in replacement of the reference's ``utils.shading.ScreenSpaceShading``
(/root/reference/SuperresolutionNetwork/utils/shading.py:7-207): same fluent setters, public fields and
``forward(input) -> B x 3 x H x W``.  CUDA tensors under ``no_grad`` run the single-pass kernel ``isr_shade``;
grad-enabled calls (the loss shades with autograd on, losses/lossnet_unshaded.py:228-230) use differentiable torch ops.
"""
import math

import numpy as np
import torch
import torch.nn as nn

from . import ops


def _vec3(color):
    assert isinstance(color, np.ndarray)
    assert color.shape == (3,)
    return [float(v) for v in np.asarray(color, dtype=np.float32)]


class ScreenSpaceShading(nn.Module):
    def __init__(self, device):
        super().__init__()
        self._device = device
        self.enable_specular = True
        self.inverse_ao = False
        self._bg = [0.0, 0.0, 0.0]
        self._ambient = self._diffuse = self._specular = self._light = self._material = None
        self._ao = 1.0
        self._specular_exponent = 1
        self._fov = 30.0

    def fov(self, fov):
        assert isinstance(fov, (float, int))
        assert 0 < fov < 90, "fov has to be in (0,90)"
        self._fov = float(fov)          # the eye direction is the constant (0,0,1) (shading.py:141-145)
        return self

    def get_fov(self):
        return self._fov

    def ambient_light_color(self, color):
        self._ambient = _vec3(color)
        return self

    def diffuse_light_color(self, color):
        self._diffuse = _vec3(color)
        return self

    def specular_light_color(self, color):
        self._specular = _vec3(color)
        return self

    def specular_exponent(self, exponent):
        assert isinstance(exponent, (int, float))
        if isinstance(exponent, float):
            assert exponent.is_integer()
            exponent = int(exponent)
        assert exponent > 0
        self._specular_exponent = exponent

    def light_direction(self, dir):
        assert isinstance(dir, np.ndarray)
        assert dir.shape == (3,)
        self._light = [float(v) for v in (dir / np.linalg.norm(dir)).astype(np.float32)]   # shading.py:77
        return self

    def material_color(self, color):
        self._material = _vec3(color)
        return self

    def ambient_occlusion(self, ao):
        self._ao = float(ao)
        return self

    def background(self, color):
        self._bg = _vec3(color)
        return self

    def params22(self):
        """The 22-float block of include/isr.h isr_shade."""
        for name in ("_ambient", "_diffuse", "_specular", "_light", "_material"):
            if getattr(self, name) is None:
                raise AttributeError("ScreenSpaceShading: " + name[1:] + " has not been set")
        return (self._ambient + self._diffuse + self._specular + self._light + self._material + self._bg
                + [self._ao, float(self._specular_exponent), float(bool(self.enable_specular)),
                   float(bool(self.inverse_ao))])

    def forward(self, input):
        assert input.shape[1] >= 5                                  # shading.py:150
        if torch.is_grad_enabled() and input.requires_grad:
            return self._forward_autograd(input)
        if not input.is_cuda:
            # several reference scripts shade `.cpu()` copies with ScreenSpaceShading('cpu') (mainPSNR2_AllAngles.py:188,
            # mainDatasetViewer.py:115, mainComparisonImages.py:189).  There is still no CPU arithmetic
            # here: the tensor is shaded by the CUDA kernel and the result is returned on the caller's device.
            if not torch.cuda.is_available():
                raise RuntimeError("ScreenSpaceShading (B200 build): needs a CUDA device; there is no CPU fallback")
            return ops.shade(input.cuda(), self.params22()).to(input.device)
        return ops.shade(input, self.params22())

    def _forward_autograd(self, x):
        # shading.py:148-191 restated with differentiable ops (training only)
        p = self.params22()
        t = lambda v: torch.tensor(v, device=x.device, dtype=x.dtype).view(1, 3, 1, 1)
        amb, dif, spc, light, mat, bg = t(p[0:3]), t(p[3:6]), t(p[6:9]), t(p[9:12]), t(p[12:15]), t(p[15:18])
        mask = torch.clamp(x[:, 0:1] * 0.5 + 0.5, 0, 1)
        normal = x[:, 1:4]
        if x.shape[1] >= 6:
            ao = 1 - x[:, 5:6] if self.inverse_ao else x[:, 5:6]
            ao = self._ao * torch.clamp(ao, 0, 1) + (1 - self._ao)
        else:
            ao = torch.ones_like(x[:, 0:1])
        ldn = torch.sum(light * normal, dim=1, keepdim=True)
        col = amb * mat + (dif * mat) * torch.abs(ldn)
        if self.enable_specular:
            rz = torch.clamp(2 * ldn * normal[:, 2:3] - light[:, 2:3], 0, 1)
            col = col + spc * (((self._specular_exponent + 2) / (2 * math.pi)) * rz ** self._specular_exponent)
        col = col * ao
        return torch.clamp(bg + mask * (col - bg), 0, 1)

    @staticmethod
    def normalize(input, dim):
        """shading.py:193-207: v / max(||v||_2, 1e-7); torch ops (cheap, used by callers on arbitrary tensors)."""
        norm = torch.clamp(torch.norm(input, p=2, dim=dim, keepdim=True), min=1e-7)
        return input / norm
