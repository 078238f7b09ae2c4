"""Inference facade: drop-in for the reference's ``inference.loadedmodel.LoadedModel``
(/root/reference/SuperresolutionNetwork/inference/loadedmodel.py:8-121) plus ``SequenceRunner``, the forward-only frame
recurrence of the training / PSNR scripts (mainVideoUnshaded.py:417-465, mainPSNR1.py:148-180) on the fused kernels.
"""
import os.path

import numpy as np
import torch

from .enhancenet import EnhanceNet
from .initial_image import initialImage
from .shading import ScreenSpaceShading


def _checkpoint_pickle_module():
    """``pickle`` with one change: ``models.enhancenet.EnhanceNet`` -- the class path inside the reference's whole-module
    checkpoints (mainVideoUnshaded.py:799-811) -- resolves to the drop-in when no ``models`` package is importable (with
    compat/ or the reference tree on sys.path the normal lookup wins)."""
    import pickle
    import types

    class Unpickler(pickle.Unpickler):
        def find_class(self, module, name):
            if module == "models.enhancenet" and name == "EnhanceNet":
                try:
                    return super().find_class(module, name)
                except (ImportError, AttributeError):
                    return EnhanceNet
            return super().find_class(module, name)

    mod = types.ModuleType("isr_checkpoint_pickle")
    mod.__dict__.update({k: getattr(pickle, k) for k in dir(pickle) if not k.startswith("__")})
    mod.Unpickler = Unpickler
    mod.load = lambda f, **kw: Unpickler(f, **kw).load()
    return mod


def _as_dropin(model):
    if isinstance(model, EnhanceNet):
        return model
    if model.__class__.__name__ == "EnhanceNet":        # a reference module, e.g. from a whole-module pickle
        return EnhanceNet.from_reference(model)
    raise TypeError("only EnhanceNet generators are on the B200 path, got " + model.__class__.__name__)


class LoadedModel:
    """``LoadedModel(name, device, upscale_factor)``; ``name`` is a checkpoint path as in the reference, or a dict
    ``{'model': module, 'parameters': {...}}`` / a module (convenience for tests)."""

    def __init__(self, name, device, upscale_factor):
        self.device = device
        self.upscale_factor = upscale_factor
        if isinstance(name, (str, bytes, os.PathLike)):
            self.name = os.path.splitext(os.path.basename(name))[0]
            # whole-module pickles (mainVideoUnshaded.py:799-811) need weights_only=False; only for trusted files
            checkpoint = torch.load(name, map_location="cpu", weights_only=False,
                                    pickle_module=_checkpoint_pickle_module())
        else:
            self.name = "model"
            checkpoint = name if isinstance(name, dict) else {"model": name}
        self.parameters = checkpoint.get("parameters", dict())
        if not isinstance(self.parameters, dict):
            self.parameters = vars(self.parameters)
        self.model = _as_dropin(checkpoint["model"]).to(device)
        self.model.train(False)

        # first leaf module decides the channel layout (loadedmodel.py:26-55)
        first_module = self.model
        while True:
            children = list(first_module.children())
            if not children:
                break
            first_module = children[0]
        self.input_channels = first_module.in_channels
        if self.input_channels == 5 + 6 * (self.upscale_factor ** 2) or self.parameters.get("unshaded", False):
            self.unshaded = True
        else:
            self.unshaded = False
            self.input_single_channels = self.input_channels - 3 * (self.upscale_factor ** 2)
            self.has_normal = self.input_single_channels >= 7
            if self.has_normal:
                self.input_single_channels -= 3
            self.has_depth = self.input_single_channels >= 5
            if self.has_depth:
                self.input_single_channels -= 1
        if "initialImage" not in self.parameters:
            self.initial_image_mode = "input" if self.unshaded else "zero"
        else:
            self.initial_image_mode = self.parameters["initialImage"]
        self.inverse_ao = self.parameters.get("aoInverted", False)

    @staticmethod
    def _inpaint_flow(current_low):
        """loadedmodel.py:77-82: Navier-Stokes in-painting of the flow where mask == 0, on the host (OpenCV), as the
        reference does.  The device replacement is ops.inpaint_flow (inference(..., inpaint="gpu"))."""
        import cv2 as cv
        cpu = current_low.detach().cpu().numpy()
        out = np.empty((cpu.shape[0], 2) + cpu.shape[2:], np.float32)
        for b in range(cpu.shape[0]):
            hole = np.uint8(cpu[b, 3] == 0)
            out[b, 0] = cv.inpaint(cpu[b, 8], hole, 3, cv.INPAINT_NS)
            out[b, 1] = cv.inpaint(cpu[b, 9], hole, 3, cv.INPAINT_NS)
        return torch.from_numpy(out).to(current_low.device)

    def _variant(self):
        if self.unshaded:
            return "unshaded"
        if self.has_normal and self.has_depth:
            return "shaded_normal_depth"
        if self.has_normal:
            return "shaded_normal"
        if self.has_depth:
            return "shaded_depth"
        return "shaded"

    def inference(self, current_low, prev_high, inpaint_flow=True, inpaint=None):
        """current_low: B x >=10 x h x w renderer output (RGB, mask, normal, depth, flow, ...) on the GPU -- or the
        renderer's own interleaved B x h x w x 12 / h x w x 12 frame; prev_high: previous high-res result or None.
        Returns the prediction B x Cout x 4h x 4w.

        ``inpaint``: "opencv" (the reference's host-side cv.inpaint, loadedmodel.py:77-82; default, needs a device
        synchronisation and a CPU round trip per frame) or "gpu" (push-pull fill on the device, ops.inpaint_flow:
        nothing leaves the GPU).  ``self.inpaint`` sets the default.  Channel slicing / mask remap / clamps
        (loadedmodel.py:84-105) are one ingest kernel either way."""
        from . import ops
        mode = inpaint or getattr(self, "inpaint", "opencv")
        if mode not in ("opencv", "gpu"):
            raise ValueError("unknown inpaint mode: " + str(mode))
        with torch.no_grad():
            current_low = current_low.to(self.device)
            interleaved = current_low.dim() == 3 or (current_low.shape[-1] in (10, 12) and current_low.shape[1] not in (10, 12))
            inp, flow, hole, _ = ops.ingest_gbuffer(current_low, self._variant(), interleaved=interleaved)
            if inpaint_flow:
                if mode == "gpu":
                    flow = ops.inpaint_flow(flow, hole)
                else:
                    planar = current_low if not interleaved else (
                        current_low if current_low.dim() == 4 else current_low.unsqueeze(0)).permute(0, 3, 1, 2)
                    flow = self._inpaint_flow(planar)
            if self.unshaded:
                if prev_high is None:
                    prev = initialImage(inp, 6, self.initial_image_mode, self.inverse_ao, self.upscale_factor)
                    prediction, _ = self.model(inp, prev, None)
                else:
                    prediction, _ = self.model(inp, prev_high.to(self.device), flow, special_mask=True)
            else:
                if prev_high is None:
                    # loadedmodel.py:110 passes upscale_factor as aoInverted (harmless for 'zero'); reproduced
                    prev_high = initialImage(inp, 3, self.initial_image_mode, self.upscale_factor)
                prediction, _ = self.model(inp, prev_high.to(self.device), flow, special_mask=False)
        return prediction


def compose_frame(image_low, image_raw, shading=None, masking=False, background=(0.0, 0.0, 0.0)):
    """The composition step of ``GUI.performSuperresolution`` (mainGUI.py:603-608,615-636; mainComparisonVideo3.py:540-556)
    as ONE kernel.  ``image_low``: B x 12 x h x w renderer output (mask already in [-1, 1] as the GUI keeps it);
    ``image_raw``: the post-processed network output B x 6 x 4h x 4w (unshaded networks, needs ``shading``) or the RGB
    prediction B x 3 x 4h x 4w (shaded networks).  Returns the B x 12 x 4h x 4w display tensor."""
    from . import ops
    unshaded = image_raw.shape[1] >= 5
    if unshaded and shading is None:
        raise ValueError("compose_frame: unshaded composition needs a ScreenSpaceShading object")
    return ops.compose(image_low, image_raw, shading.params22() if unshaded else None, unshaded, masking,
                       float(background[0]))


def temporal_smoothing(previous_rgb, flow, image_rgb, factor, upscale_factor=4):
    """The GUI's temporal post-smoothing (mainGUI.py:834-849) as one kernel: the previously DISPLAYED RGB frame is warped
    by the (in-painted) flow with ``special_mask=False`` and blended, ``factor * warped + (1 - factor) * image_rgb``
    (``factor`` = slider / 100).  ``previous_rgb is None`` or ``factor == 0`` returns ``image_rgb`` (mainGUI.py:835-838)."""
    from . import ops
    if previous_rgb is None or factor == 0:
        return image_rgb
    return ops.temporal_blend(previous_rgb, flow, image_rgb, float(factor), upscale_factor)


class SequenceRunner:
    """Forward-only frame recurrence over a batch of independent sequences (mainVideoUnshaded.py:417-465).

    ``low``: B x T x c_low x h x w, ``flow``: B x T x 2 x h x w (CUDA).  Per frame: fused warp + space-to-depth + concat
    -> EnhanceNet -> fused clamp/normalise state; optional shading of every frame."""

    def __init__(self, model, initial_image="input", shading=None):
        self.model = _as_dropin(model)
        self.initial_image = initial_image
        self.shading = shading
        self.unshaded = self.model.output_channels == 6

    @torch.no_grad()
    def run(self, low, flow, keep_outputs=True, max_batch=None):
        """``max_batch``: process the sequences in chunks of at most that many (they are independent, and a sequence's
        result does not depend on its batch, bit for bit), which bounds the workspace: BASELINE.json configs[4] (64
        sequences of 512^2 -> 2048^2 per GPU) needs 32 GiB per 64-channel 4x tensor in one piece, 2 GiB with max_batch=4."""
        if max_batch is not None and low.shape[0] > max_batch:
            parts = [self.run(low[i:i + max_batch], flow[i:i + max_batch], keep_outputs)
                     for i in range(0, low.shape[0], max_batch)]
            preds = [torch.cat([p[0][j] for p in parts]) for j in range(len(parts[0][0]))]
            rgbs = [torch.cat([p[1][j] for p in parts]) for j in range(len(parts[0][1]))]
            return preds, rgbs, torch.cat([p[2] for p in parts])
        B, T = low.shape[:2]
        c_prev = self.model.output_channels
        preds, rgbs = [], []
        state = None
        for j in range(T):
            inp = low[:, j].contiguous()
            if j == 0:
                prev = initialImage(inp, c_prev, self.initial_image, False, 4)
                pred, _, state = self.model(inp, prev, None, return_state=True)
            else:
                pred, _, state = self.model(inp, state, flow[:, j - 1].contiguous(), special_mask=self.unshaded,
                                            return_state=True)
            if self.shading is not None:
                rgb = self.shading(state)
                if keep_outputs:
                    rgbs.append(rgb)
            if keep_outputs:
                preds.append(pred)
        return preds, rgbs, state


__all__ = ["LoadedModel", "SequenceRunner", "ScreenSpaceShading", "compose_frame", "temporal_smoothing"]
