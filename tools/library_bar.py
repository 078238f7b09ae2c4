"""The "library bar" of SURVEY.md 8(d): the same recurrent step (warp + space-to-depth + concat, EnhanceNet, post-processing,
shading) written with stock PyTorch ops (cuDNN convolutions, grid_sample, interpolate) on the same B200, i.e. what the
reference's own GPU path does on this hardware.  Workload: BASELINE.json configs[1] (16 sequences x 256^2 -> 1024^2).
Prints one JSON object; measurement only -- nothing here is on the product path, and nothing of oracle/ is imported.

    python tools/library_bar.py [--steps 10] [--warmup 3]
"""
import argparse
import json
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F


class TorchEnhanceNet(nn.Module):
    """Architecture of models/enhancenet.py:92-145 (bilinear upsampling, residual reconstruction), stock modules."""

    def __init__(self, cin=101, cout=6):
        super().__init__()
        self.pre = nn.Conv2d(cin, 64, 3, padding=1)
        self.blocks = nn.ModuleList(nn.ModuleList([nn.Conv2d(64, 64, 3, padding=1), nn.Conv2d(64, 64, 3, padding=1)])
                                    for _ in range(10))
        self.post = nn.ModuleList([nn.Conv2d(64, 64, 3, padding=1) for _ in range(3)])
        self.final = nn.Conv2d(64, cout, 3, padding=1)
        self.cout = cout

    def forward(self, x):
        f = F.relu(self.pre(x))
        for a, b in self.blocks:
            f = f + b(F.relu(a(f)))
        f = F.relu(self.post[0](F.interpolate(f, scale_factor=2, mode="bilinear", align_corners=False)))
        f = F.relu(self.post[1](F.interpolate(f, scale_factor=2, mode="bilinear", align_corners=False)))
        f = F.relu(self.post[2](f))
        res = self.final(f)
        return res + self._skip(x)

    def _skip(self, x):
        up = F.interpolate(x[:, :5], scale_factor=4, mode="bilinear", align_corners=False)
        return torch.cat((up, torch.ones_like(up[:, :1])), 1)      # mask, normal, depth from the input; AO starts at 1


def step(net, low, flow, prev, dtype, channels_last):
    B, _, h, w = low.shape
    # warp the previous high-res output by the upscaled flow (videotools.py:51-87), then space-to-depth (:8-25)
    f = F.interpolate(flow, scale_factor=4, mode="bilinear", align_corners=False)
    ys, xs = torch.meshgrid(torch.linspace(-1 + 1 / (4 * h), 1 - 1 / (4 * h), 4 * h, device=low.device),
                            torch.linspace(-1 + 1 / (4 * w), 1 - 1 / (4 * w), 4 * w, device=low.device), indexing="ij")
    grid = torch.stack((xs[None] - 2.0 * f[:, 0], ys[None] + 2.0 * f[:, 1]), -1)
    warped = F.grid_sample(prev, grid, mode="bilinear", padding_mode="zeros", align_corners=False)
    x = torch.cat((low, F.pixel_unshuffle(warped, 4)), 1)
    if dtype != torch.float32:
        x = x.to(dtype)
    if channels_last:
        x = x.contiguous(memory_format=torch.channels_last)
    y = net(x).float()
    # post-processing (mainVideoUnshaded.py:460-465) and Phong shading of the prediction (utils/shading.py:148-191)
    mask = y[:, 0:1].clamp(-1, 1)
    normal = F.normalize(y[:, 1:4], dim=1)
    depth, ao = y[:, 4:5].clamp(0, 1), y[:, 5:6].clamp(0, 1)
    out = torch.cat((mask, normal, depth, ao), 1)
    light = torch.tensor([0.1, 0.1, 1.0], device=y.device)
    light = (light / light.norm()).view(1, 3, 1, 1)
    ndl = (normal * light).sum(1, keepdim=True).abs()
    rgb = (0.1 + ndl) * torch.tensor([1.0, 0.3, 0.0], device=y.device).view(1, 3, 1, 1) * ao
    rgb = rgb * (mask * 0.5 + 0.5)
    return out, rgb


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    if not torch.cuda.is_available():
        print(json.dumps({"unavailable": "no CUDA device"}))
        return 0
    dev = torch.device("cuda")
    B, h, w = 16, 256, 256
    g = torch.Generator(device=dev).manual_seed(4322)
    low = torch.rand(B, 5, h, w, device=dev, generator=g) * 2 - 1
    flow = (torch.rand(B, 2, h, w, device=dev, generator=g) - 0.5) * 0.04
    results = {}
    for name, dtype, cl, tf32 in (("fp32", torch.float32, False, False), ("tf32", torch.float32, False, True),
                                  ("bf16_channels_last", torch.bfloat16, True, True),
                                  ("fp16_channels_last", torch.float16, True, True)):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        torch.backends.cudnn.benchmark = True
        torch.manual_seed(1234)
        net = TorchEnhanceNet().to(dev).eval()
        if dtype != torch.float32:
            net = net.to(dtype)
        if cl:
            net = net.to(memory_format=torch.channels_last)
        prev = torch.zeros(B, 6, 4 * h, 4 * w, device=dev)
        try:
            with torch.no_grad():
                for _ in range(args.warmup):
                    prev, rgb = step(net, low, flow, prev, dtype, cl)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(args.steps):
                    prev, rgb = step(net, low, flow, prev, dtype, cl)
                e1.record()
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.steps
            results[name] = {"ms_per_step": ms, "mpix_per_s": B * 16 * h * w / ms / 1e3,
                             "finite": bool(torch.isfinite(prev).all())}
        except RuntimeError as e:                                  # e.g. out of memory for a variant
            results[name] = {"error": str(e).splitlines()[0][:200]}
        del net, prev
        torch.cuda.empty_cache()
    print(json.dumps({"workload": "unshaded 4x, 16 x 256^2 -> 1024^2, one recurrent step (stock PyTorch ops / cuDNN)",
                      "torch": torch.__version__, "cudnn": torch.backends.cudnn.version(), "steps": args.steps,
                      "warmup": args.warmup, "results": results}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
