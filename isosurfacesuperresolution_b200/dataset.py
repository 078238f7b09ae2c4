"""Reader for the reference's on-disk sequence format and a forward-only evaluation loop.

Format (DataGenerator/DataGeneratorVideo2.py:46-90, read by datasetVideo.py:104-170): per sequence i three fp32 files
``low_%05d.npy`` (T x 5 x h x w: mask in [-1,1], normal, depth), ``high_%05d.npy`` (T x 6 x 4h x 4w: + AO) and
``flow_%05d.npy`` (T x 2 x h x w, already in-painted); the dataset path is a folder, or a text file whose lines are
sub-folders.  Host side only: files are memory-mapped, batches are staged in pinned memory and handed to the CUDA path
(SequenceRunner); nothing here computes on the CPU besides the PSNR bookkeeping of the result.
"""
import os

import numpy as np
import torch

from . import ops
from .inference import SequenceRunner


def _names(folder, i):
    return tuple(os.path.join(folder, "%s_%05d.npy" % (m, i)) for m in ("high", "low", "flow"))


class SequenceDataset:
    """``SequenceDataset(path, number_of_images=None)``: same discovery rule as datasetVideo.py:114-146 (count upwards
    from 0 until ``low_%05d.npy`` is missing; a file path lists sub-folders, one per line)."""

    def __init__(self, path, number_of_images=None, mmap=True):
        folders = [path]
        if os.path.isfile(path):
            with open(path, "r") as fp:
                folders = [os.path.join(os.path.dirname(path), line.strip()) for line in fp if line.strip()]
        self.paths = []
        for folder in folders:
            i = 0
            while os.path.exists(_names(folder, i)[1]):
                self.paths.append(_names(folder, i))
                i += 1
        if not self.paths:
            raise ValueError("No image found")                      # datasetVideo.py:149
        if number_of_images is not None and number_of_images > 0:
            self.paths = self.paths[:number_of_images]
        self._mmap = "r" if mmap else None
        low0 = np.load(self.paths[0][1], mmap_mode="r")
        self.num_frames, self.input_channels = int(low0.shape[0]), int(low0.shape[1])
        self.low_size = tuple(int(v) for v in low0.shape[2:])
        high0 = np.load(self.paths[0][0], mmap_mode="r")
        self.output_channels = int(high0.shape[1])
        self.upscale_factor = int(high0.shape[2]) // self.low_size[0]

    def __len__(self):
        return len(self.paths)

    def __getitem__(self, i):
        """-> (low T x 5 x h x w, flow T x 2 x h x w, high T x 6 x H x W) numpy views"""
        hp, lp, fp = self.paths[i]
        return np.load(lp, mmap_mode=self._mmap), np.load(fp, mmap_mode=self._mmap), np.load(hp, mmap_mode=self._mmap)

    def batches(self, batch_size, pin=True, with_high=True):
        """Yields (low B x T x C x h x w, flow B x T x 2 x h x w, high or None) as (pinned) torch tensors; the last batch
        may be ragged."""
        for start in range(0, len(self), batch_size):
            items = [self[i] for i in range(start, min(start + batch_size, len(self)))]
            out = []
            for k in range(3 if with_high else 2):
                t = torch.from_numpy(np.stack([np.asarray(it[k], dtype=np.float32) for it in items]))
                out.append(t.pin_memory() if pin and torch.cuda.is_available() else t)
            yield (out[0], out[1], out[2] if with_high else None)


def masked_psnr(img1, img2, mask=None, epsilon=1e-7):
    """utils/psnr.py:10-22 (PSNR.forward): per-sample PSNR, optionally restricted to mask > 0 pixels."""
    if mask is None:
        return 10 * torch.log10(1 / (epsilon + torch.mean((img1 - img2) ** 2, dim=[1, 2, 3])))
    img1, img2 = mask * img1, mask * img2
    B, C, H, W = mask.shape
    factor = (H * W) / torch.sum(mask, dim=[1, 2, 3])
    return 10 * factor * torch.log10(1 / (epsilon + torch.mean((img1 - img2) ** 2, dim=[1, 2, 3])))


def evaluate(model, dataset, batch_size=16, initial_image="zero", shading=None, device="cuda"):
    """Forward-only pass over a dataset (the loop of mainPSNR3_AllStats.py:330-360 on the fused kernels): per frame the
    masked PSNR of normal / depth / AO and the PSNR of the mask against ``high``.  Returns a dict of means."""
    runner = SequenceRunner(model, initial_image=initial_image, shading=shading)
    sums = {"psnr_mask": 0.0, "psnr_normal": 0.0, "psnr_depth": 0.0, "psnr_ao": 0.0}
    n = 0
    for low, flow, high in dataset.batches(batch_size):
        low, flow, high = (t.to(device, non_blocking=True) for t in (low, flow, high))
        preds, _, _ = runner.run(low, flow)
        for j, pred in enumerate(preds):
            gt = high[:, j]
            m = (gt[:, 0:1] * 0.5 + 0.5).clamp(0, 1)
            if pred.shape[1] >= 5:      # mainPSNR3_AllStats.py:352-354: clamp mask, normalise normal, clamp depth / AO
                pred = ops.postprocess(pred, unshaded=True)
            sums["psnr_mask"] += float(masked_psnr(pred[:, 0:1], gt[:, 0:1]).sum())
            sums["psnr_normal"] += float(masked_psnr(pred[:, 1:4], gt[:, 1:4], m).sum())
            sums["psnr_depth"] += float(masked_psnr(pred[:, 4:5], gt[:, 4:5], m).sum())
            if pred.shape[1] > 5 and gt.shape[1] > 5:
                sums["psnr_ao"] += float(masked_psnr(pred[:, 5:6], gt[:, 5:6], m).sum())
            n += pred.shape[0]
    return {k: v / max(n, 1) for k, v in sums.items()} | {"frames": n}


__all__ = ["SequenceDataset", "masked_psnr", "evaluate"]
