"""CPU: the .npy sequence reader (datasetVideo.py:104-170 format) and the masked PSNR (utils/psnr.py:10-22)."""
import os

import numpy as np
import pytest
import torch


def _write(folder, n, T=3, h=4, w=6, seed=0):
    os.makedirs(folder, exist_ok=True)
    rng = np.random.RandomState(seed)
    data = []
    for i in range(n):
        low = rng.rand(T, 5, h, w).astype(np.float32)
        high = rng.rand(T, 6, 4 * h, 4 * w).astype(np.float32)
        flow = rng.rand(T, 2, h, w).astype(np.float32)
        np.save(os.path.join(folder, "low_%05d.npy" % i), low)
        np.save(os.path.join(folder, "high_%05d.npy" % i), high)
        np.save(os.path.join(folder, "flow_%05d.npy" % i), flow)
        data.append((low, flow, high))
    return data


def test_folder_listfile_and_batches(tmp_path):
    from isosurfacesuperresolution_b200.dataset import SequenceDataset
    a = _write(str(tmp_path / "setA"), 3, seed=1)
    b = _write(str(tmp_path / "setB"), 2, seed=2)
    ds = SequenceDataset(str(tmp_path / "setA"))
    assert len(ds) == 3 and ds.num_frames == 3 and ds.input_channels == 5 and ds.output_channels == 6
    assert ds.low_size == (4, 6) and ds.upscale_factor == 4
    for i in range(3):
        low, flow, high = ds[i]
        assert np.array_equal(low, a[i][0]) and np.array_equal(flow, a[i][1]) and np.array_equal(high, a[i][2])
    # a list file names sub-folders, one per line (datasetVideo.py:118-136)
    with open(tmp_path / "sets.txt", "w") as fp:
        fp.write("setA\nsetB\n")
    both = SequenceDataset(str(tmp_path / "sets.txt"))
    assert len(both) == 5 and np.array_equal(both[3][0], b[0][0])
    assert len(SequenceDataset(str(tmp_path / "sets.txt"), number_of_images=4)) == 4
    # batches: ragged tail, order preserved
    got = list(both.batches(2, pin=False))
    assert [g[0].shape[0] for g in got] == [2, 2, 1] and got[0][0].shape == (2, 3, 5, 4, 6)
    assert torch.equal(got[2][2][0], torch.from_numpy(b[1][2]))
    assert list(both.batches(8, pin=False, with_high=False))[0][2] is None
    with pytest.raises(ValueError):
        SequenceDataset(str(tmp_path / "empty"))


def test_masked_psnr_matches_reference_formula():
    from isosurfacesuperresolution_b200.dataset import masked_psnr
    g = torch.Generator().manual_seed(0)
    x, y = torch.rand(2, 3, 8, 8, generator=g), torch.rand(2, 3, 8, 8, generator=g)
    m = (torch.rand(2, 1, 8, 8, generator=g) > 0.5).float()
    ref = 10 * torch.log10(1 / (1e-7 + torch.mean((x - y) ** 2, dim=[1, 2, 3])))
    assert torch.allclose(masked_psnr(x, y), ref)
    f = 64 / m.sum(dim=[1, 2, 3])
    ref_m = 10 * f * torch.log10(1 / (1e-7 + torch.mean((m * x - m * y) ** 2, dim=[1, 2, 3])))
    assert torch.allclose(masked_psnr(x, y, m), ref_m)
