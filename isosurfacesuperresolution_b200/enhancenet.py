"""Drop-in replacement of the reference's ``models.enhancenet.EnhanceNet``
(/root/reference/SuperresolutionNetwork/models/enhancenet.py:7-145).

Same constructor, attributes, module tree and ``state_dict`` keys (``preblock.0``, ``blocks.{0..9}.{0,2}``,
``postblock.{1,4,6,8}``), same ``forward(inputs) -> (outputs, residual)`` contract.  Under ``torch.no_grad()`` on a CUDA
device the forward runs libisr.so's sm_100a kernels (tcgen05 implicit-GEMM convolutions); the parameters stay ordinary
``nn.Parameter``s and are re-packed when they change.  With autograd enabled (training scripts back-propagate through
ten frames, mainVideoUnshaded.py:447,467) the same parameters are applied with differentiable torch ops -- that is the
training path, not an inference fallback: inference on a non-CUDA device raises.

Extension (north star): ``forward(inputs, previous_output, flow)`` fuses warp + space-to-depth + concat
(videotools.py:8-25,49-87) into the network-input packing, so ``inputs`` is then the bare low-res G-buffer.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops

STATE_DICT_KEYS = (
    ["preblock.0.weight", "preblock.0.bias"]
    + [f"blocks.{i}.{j}.{k}" for i in range(10) for j in (0, 2) for k in ("weight", "bias")]
    + [f"postblock.{j}.{k}" for j in (1, 4, 6, 8) for k in ("weight", "bias")]
)


class EnhanceNet(nn.Module):
    # tensor-core operand scheme of the CUDA path (include/isr.h); "hybrid" meets the 1e-2 / 50 dB parity bar
    precision = "hybrid"

    def __init__(self, upscale_factor, input_channels, channel_mask, output_channels, opt):
        super().__init__()
        assert upscale_factor == 4
        self.upscale_factor = upscale_factor
        self.upsample = opt.upsample
        self.recon_type = opt.reconType
        self.use_bn = opt.useBN
        self.channel_mask = channel_mask
        self.input_channels = input_channels
        self.output_channels = output_channels
        if self.use_bn:
            raise NotImplementedError("useBN=True is not supported by the B200 path: fold BatchNorm into the convs first")
        self._build(input_channels, output_channels)
        self._init_weights()
        self._handle = None
        self._handle_key = None

    # ---- module tree (enhancenet.py:92-125) -------------------------------------------------------------------
    def _up(self, factor):
        if self.upsample in ("nearest", "bilinear", "bicubic"):
            return nn.Upsample(scale_factor=factor, mode=self.upsample)
        raise ValueError("unsupported upsample mode " + str(self.upsample))

    def _build(self, cin, cout):
        self.preblock = nn.Sequential(nn.Conv2d(cin, 64, 3, padding=1), nn.ReLU())
        self.blocks = nn.ModuleList(
            [nn.Sequential(nn.Conv2d(64, 64, 3, padding=1), nn.ReLU(), nn.Conv2d(64, 64, 3, padding=1)) for _ in range(10)])
        self.postblock = nn.Sequential(
            self._up(2), nn.Conv2d(64, 64, 3, padding=1), nn.ReLU(),
            self._up(2), nn.Conv2d(64, 64, 3, padding=1), nn.ReLU(),
            nn.Conv2d(64, 64, 3, padding=1), nn.ReLU(),
            nn.Conv2d(64, cout, 3, padding=1))

    def _init_weights(self):
        # enhancenet.py:127-133: orthogonal init with ReLU gain on the residual-block convs only
        gain = nn.init.calculate_gain("relu")
        for block in self.blocks:
            for m in block:
                if isinstance(m, nn.Conv2d):
                    nn.init.orthogonal_(m.weight, gain)

    # ---- pickling: drop the native handle -------------------------------------------------------------------
    def __getstate__(self):
        st = dict(self.__dict__)
        st["_handle"] = None
        st["_handle_key"] = None
        return st

    @classmethod
    def from_reference(cls, module):
        """Builds a drop-in from a reference ``EnhanceNet`` instance (e.g. un-pickled from a checkpoint)."""
        from argparse import Namespace
        opt = Namespace(upsample=module.upsample, reconType=module.recon_type, useBN=module.use_bn)
        mask = getattr(module, "channel_mask", None)
        cin = module.preblock[0].in_channels
        cout = module.postblock[8].out_channels
        if mask is None:                                   # old checkpoints (enhancenet.py:57-61)
            mask = list(range(min(cin, cout)))
        net = cls(module.upscale_factor, cin, mask, cout, opt)
        net.load_state_dict(module.state_dict())
        return net.to(module.preblock[0].weight.device).train(module.training)

    # ---- CUDA path ------------------------------------------------------------------------------------------
    def _params(self):
        sd = dict(self.named_parameters())
        return [sd[k] for k in STATE_DICT_KEYS]

    def _native(self):
        params = self._params()
        dev = params[0].device
        key = (str(dev), self.upsample, self.recon_type, tuple(p._version for p in params),
               tuple(p.data_ptr() for p in params))
        if self._handle is None or self._handle_key != key:
            if self._handle is None or self._handle_key[:3] != key[:3]:
                n_mask = len(self.channel_mask) if self.recon_type == "residual" else 0
                self._handle = ops.NetHandle(self.input_channels, self.output_channels, n_mask, self.upsample,
                                             self.recon_type == "residual", dev)
            self._handle.load_weights(params)
            self._handle_key = key
        return self._handle

    def _check_recon(self):
        if self.recon_type not in ("residual", "direct"):
            raise ValueError("Unknown Reconstruction Type: " + str(self.recon_type))     # enhancenet.py:80
        if self.recon_type == "residual" and len(self.channel_mask) > self.output_channels:
            raise ValueError("number of output channels must be at least the number of masked input channels")

    def forward(self, inputs, previous_output=None, flow=None, special_mask=None, return_state=False):
        self._check_recon()
        needs_grad = torch.is_grad_enabled() and (
            inputs.requires_grad or any(p.requires_grad for p in self.parameters()))
        if needs_grad:
            return self._forward_autograd(inputs, previous_output, flow, special_mask)
        if not inputs.is_cuda:
            raise RuntimeError("EnhanceNet (B200 build): inference needs CUDA tensors; there is no CPU fallback")
        h = self._native()
        unshaded = self.output_channels == 6
        if previous_output is None:
            recon, raw, state = h.forward(inputs, self.precision, want_raw=True, want_state=return_state,
                                          state_unshaded=unshaded)
        else:
            if special_mask is None:
                special_mask = unshaded                    # mainVideoUnshaded.py:429-433 vs mainVideo.py:408
            recon, raw, state = h.forward_recurrent(inputs, previous_output, flow, special_mask, self.precision,
                                                    want_raw=True, want_state=return_state, state_unshaded=unshaded)
        if return_state:
            return recon, raw, state
        return recon, raw

    # ---- training path (differentiable torch ops on the same parameters) --------------------------------------
    def _forward_autograd(self, inputs, previous_output, flow, special_mask):
        if previous_output is not None:
            from .videotools import VideoTools
            if flow is not None:
                sm = (self.output_channels == 6) if special_mask is None else special_mask
                previous_output = VideoTools.warp_upscale(previous_output, flow, 4, special_mask=sm)
            inputs = torch.cat((inputs, VideoTools.flatten_high(previous_output, 4)), dim=1)
        features = self.preblock(inputs)
        for block in self.blocks:
            features = features + block(features)
        outputs = self.postblock(features)
        if self.recon_type == "residual":
            n = len(self.channel_mask)
            resized = F.interpolate(inputs[:, 0:n, :, :], size=[outputs.shape[2], outputs.shape[3]], mode=self.upsample)
            if n == self.output_channels:
                return resized + outputs, outputs
            return torch.cat([resized + outputs[:, 0:n, :, :], outputs[:, n:, :, :]], dim=1), outputs
        return outputs, outputs
