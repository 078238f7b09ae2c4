"""CPU: the oracle's restatement of the G-buffer ingest (inference/loadedmodel.py:84-105, mainGUI.py:713-716) and of the
push-pull flow in-painting that replaces cv.inpaint (loadedmodel.py:77-82).  The ingest is checked against the reference's
own slicing expressions evaluated with torch; the in-painter (a different algorithm from OpenCV's: parity unpinned)
through its defining properties and against OpenCV on the reference's call."""
import numpy as np
import pytest
import torch

from oracle import oracle as orc


def _render(B=2, h=19, w=27, seed=3):
    rng = np.random.RandomState(seed)
    r = rng.uniform(-0.2, 1.2, (B, h, w, 12)).astype(np.float32)
    r[..., 3] = (rng.rand(B, h, w) > 0.4).astype(np.float32)          # mask in {0, 1}
    r[..., 8:10] = rng.uniform(-0.02, 0.02, (B, h, w, 2)).astype(np.float32)
    return r


@pytest.mark.parametrize("variant", [0, 1, 2, 3, 4])
def test_ingest_matches_reference_slicing(variant):
    r = _render()
    cur = torch.from_numpy(r).permute(0, 3, 1, 2).contiguous()        # the tensor LoadedModel.inference receives
    if variant == 0:
        ref = torch.cat((cur[:, 3:4] * 2 - 1, cur[:, 4:8]), dim=1)    # loadedmodel.py:85
    elif variant == 4:
        ref = torch.clamp(cur[:, 0:8], 0, 1)                          # :98
    elif variant == 3:
        ref = cur[:, 0:7]                                             # :100
    elif variant == 2:
        ref = torch.cat((cur[:, 0:4], cur[:, 7:8]), dim=1)            # :102
    else:
        ref = cur[:, 0:4]                                             # :104
    for interleaved, src in ((True, r), (False, cur.numpy())):
        low, flow, hole, img = orc.ingest_gbuffer(src, variant, interleaved=interleaved)
        assert low.shape[1] == orc.INGEST_CHANNELS[variant]
        assert np.array_equal(low, ref.numpy())
        assert np.array_equal(flow, cur[:, 8:10].numpy())
        assert np.array_equal(hole, np.uint8(cur[:, 3].numpy() == 0))     # loadedmodel.py:80
        gui = torch.cat((cur[:, 0:3], cur[:, 3:4] * 2 - 1, cur[:, 4:]), dim=1)   # mainGUI.py:713-716
        assert np.array_equal(img, gui.numpy())


def _smooth_flow(h, w):
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    return np.stack([0.01 * np.sin(xx / 9) + 0.002 * yy / h, 0.01 * np.cos(yy / 7)], 0)[None].astype(np.float32)


def test_inpaint_properties():
    rng = np.random.RandomState(0)
    h, w = 37, 53                                                     # odd sizes: ragged pyramid levels
    flow = _smooth_flow(h, w)
    hole = (rng.rand(1, h, w) < 0.3).astype(np.uint8)
    hole[0, 10:25, 20:40] = 1
    out = orc.inpaint_flow_pushpull(flow, hole)
    assert out.dtype == np.float32 and out.shape == flow.shape
    assert np.array_equal(out[:, :, hole[0] == 0], flow[:, :, hole[0] == 0])      # valid pixels are kept exactly
    # the fill is an average of valid values: it stays inside their range
    for c in range(2):
        valid = flow[0, c][hole[0] == 0]
        assert out[0, c].min() >= valid.min() - 1e-7 and out[0, c].max() <= valid.max() + 1e-7
    # no hole: identity; all hole: zeros; constant field: the constant
    assert np.array_equal(orc.inpaint_flow_pushpull(flow, np.zeros_like(hole)), flow)
    assert not orc.inpaint_flow_pushpull(flow, np.ones_like(hole)).any()
    const = np.full((1, 2, 8, 9), 0.37, np.float32)
    hh = np.ones((1, 8, 9), np.uint8)
    hh[0, 3, 4] = 0
    assert np.array_equal(np.unique(orc.inpaint_flow_pushpull(const, hh)), np.array([0.37], np.float32))
    # 1x1 and single-row images
    assert orc.inpaint_flow_pushpull(np.ones((1, 2, 1, 1), np.float32), np.ones((1, 1, 1), np.uint8)).shape == (1, 2, 1, 1)
    row = orc.inpaint_flow_pushpull(np.arange(14, dtype=np.float32).reshape(1, 2, 1, 7), np.uint8([[[0, 1, 1, 0, 1, 1, 0]]]))
    assert row[0, 0, 0, 0] == 0 and row[0, 0, 0, 3] == 3 and row[0, 0, 0, 6] == 6


def test_inpaint_against_opencv_on_the_reference_call():
    """loadedmodel.py:79-81 with cv.INPAINT_NS, radius 3.  Different algorithm: the fills agree to within the variation of
    the flow across the hole's neighbourhood (here: small holes as at silhouette edges, < 10 % of the flow range)."""
    cv = pytest.importorskip("cv2")
    rng = np.random.RandomState(1)
    h, w = 64, 96
    flow = _smooth_flow(h, w)
    hole = (rng.rand(1, h, w) < 0.15).astype(np.uint8)
    ref = np.stack([cv.inpaint(flow[0, c], hole[0], 3, cv.INPAINT_NS) for c in range(2)])[None]
    out = orc.inpaint_flow_pushpull(flow, hole)
    rng_ = float(flow.max() - flow.min())
    assert np.array_equal(out[:, :, hole[0] == 0], ref[:, :, hole[0] == 0])
    assert float(np.abs(out - ref).max()) <= 0.1 * rng_
