"""Pins the oracle (both backends) against vectors produced by the live reference
(oracle/gen_golden.py).  CPU only."""
import numpy as np
import pytest

from oracle import oracle as orc


def maxabs(a, b):
    return float(np.max(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))))


@pytest.mark.parametrize("backend", ["c", "torch"])
def test_flatten_high_bit_exact(golden, backend):
    g = golden("flatten_high")
    y = orc.flatten_high(g["x"], 4, backend=backend)
    assert y.shape == g["y"].shape
    assert np.array_equal(y, g["y"])           # indexing: bit-exact


@pytest.mark.parametrize("backend", ["c", "torch"])
@pytest.mark.parametrize("flow", ["zero", "const", "smooth", "wild"])
@pytest.mark.parametrize("special_mask", [False, True])
def test_warp_upscale(golden, backend, flow, special_mask):
    g = golden("warp_upscale")
    out = orc.warp_upscale(g["img"], g[f"flow_{flow}"], 4, special_mask=special_mask, backend=backend)
    assert maxabs(out, g[f"{flow}_{int(special_mask)}"]) <= 1e-5   # north_star warp tolerance


def test_warp_upscale_rgb_nonsquare(golden):
    g = golden("warp_upscale_rgb")
    assert maxabs(orc.warp_upscale(g["img"], g["flow"], 4), g["out"]) <= 1e-5


def test_initial_image(golden):
    g = golden("initial_image")
    for mode in ("zero", "unshaded", "input"):
        for ch in (5, 6):
            assert maxabs(orc.initial_image(g["cur"], ch, mode, False, 4), g[f"{mode}_{ch}"]) <= 1e-6
    assert maxabs(orc.initial_image(g["cur"], 6, "unshaded", True, 4), g["unshaded_6_inv"]) == 0
    assert maxabs(orc.initial_image(g["cur"], 3, "input", False, 4), g["input_3"]) <= 1e-6
    with pytest.raises(ValueError):
        orc.initial_image(g["cur"], 4, "unshaded", False, 4)
    with pytest.raises(ValueError):
        orc.initial_image(g["cur"], 6, "bogus", False, 4)


@pytest.mark.parametrize("cfg", ["train", "gui", "nospec"])
def test_shading(golden, cfg):
    g = golden("shading")
    assert maxabs(orc.shade(g["gbuffer"], g[f"{cfg}_params"]), g[f"{cfg}_c6"]) <= 1e-5
    assert maxabs(orc.shade(g["gbuffer"][:, 0:5], g[f"{cfg}_params"]), g[f"{cfg}_c5"]) <= 1e-5


def test_postprocess_and_normalize(golden):
    g = golden("shading")
    assert maxabs(orc.postprocess(g["gbuffer"], unshaded=True), g["postprocess"]) <= 1e-6
    x = g["normalize_in"]
    pad = np.concatenate([np.zeros_like(x[:, 0:1]), x, np.zeros_like(x[:, 0:2])], axis=1)
    out = orc.postprocess(pad, unshaded=True)[:, 1:4]
    assert maxabs(out, g["normalize_out"]) <= 1e-6
    assert np.all(out[..., -1] == 0)           # zero vectors stay zero


@pytest.mark.parametrize("backend", ["c", "torch"])
@pytest.mark.parametrize("ups", ["bilinear", "nearest"])
def test_enhancenet_forward(golden, weights, backend, ups):
    g = golden("enhancenet_forward")
    recon, raw = orc.enhancenet_forward(weights, g["x"], ups, "residual", 5, backend=backend)
    tol = 2e-4 if backend == "c" else 1e-5      # fp32 summation-order noise of the scalar conv
    assert maxabs(raw, g[f"raw_{ups}"]) <= tol
    assert maxabs(recon, g[f"recon_{ups}"]) <= tol


def test_enhancenet_direct(golden, weights):
    g = golden("enhancenet_forward")
    recon, raw = orc.enhancenet_forward(weights, g["x"], "bilinear", "direct", 5, backend="torch")
    assert maxabs(recon, g["recon_direct"]) <= 1e-5
    assert np.array_equal(recon, raw)


def test_sequence_recurrence(golden, weights):
    g = golden("sequence_unshaded")
    preds, rgbs = orc.unshaded_sequence(weights, g["low"], g["flow"], "bilinear", "input", backend="torch",
                                        shade_params=g["shade_params"])
    for j in range(3):
        assert maxabs(preds[j], g[f"pred_{j}"]) <= 1e-4
        assert maxabs(rgbs[j], g[f"rgb_{j}"]) <= 1e-4


# ---- round-2 fixtures (oracle/gen_golden_r02.py, live reference) -----------------------------------------------------
def test_sequence_recurrence_10_frames(golden, weights):
    """configs[1] runs 10-frame sequences: the oracle follows the live reference over all ten (free-running on both
    sides; fp32 vs fp32, so the feedback loop only amplifies summation-order noise)."""
    g = golden("sequence_unshaded_10")
    preds, _ = orc.unshaded_sequence(weights, g["low"], g["flow"], "bilinear", "input", backend="torch")
    assert len(preds) == 10
    for j in range(10):
        assert maxabs(preds[j], g[f"pred_{j}"]) <= 2e-3, j
    assert maxabs(preds[0], g["pred_0"]) <= 1e-5


def _seeded_reference_weights(cin, cout, seed):
    """The reference constructor's weights under torch.manual_seed(seed), rebuilt without the reference: the same
    sequence of nn.Conv2d constructions (default init) and the orthogonal re-init of the 20 block convs
    (models/enhancenet.py:92-133) draws the same random numbers.  Pinned by the fixtures' checksums."""
    import torch
    import torch.nn as nn
    torch.manual_seed(seed)
    convs = [nn.Conv2d(cin, 64, 3, padding=1)] + [nn.Conv2d(64, 64, 3, padding=1) for _ in range(20)]
    convs += [nn.Conv2d(64, 64, 3, padding=1) for _ in range(3)] + [nn.Conv2d(64, cout, 3, padding=1)]
    for c in convs[1:21]:
        nn.init.orthogonal_(c.weight, nn.init.calculate_gain("relu"))
    sd = {}
    for key, c in zip(orc.STATE_DICT_KEYS[0::2], convs):
        sd[key] = c.weight.detach().numpy()
        sd[key.replace("weight", "bias")] = c.bias.detach().numpy()
    return sd


def _weight_sum(sd):
    return [float(sum(np.asarray(v, np.float64).sum() for v in sd.values())),
            float(sum(np.abs(np.asarray(v, np.float64)).sum() for v in sd.values()))]


def test_sequence_shaded(golden):
    """Shaded (RGB) recurrence, mainVideo.py:383-434 / loadedmodel.py:97-119: 56 -> 3, special_mask=False, clamp state."""
    g = golden("sequence_shaded")
    sd = _seeded_reference_weights(56, 3, 1234)
    assert np.allclose(_weight_sum(sd), g["weight_sum"], rtol=1e-9)
    preds = orc.shaded_sequence(sd, g["input"], g["flow"], "bilinear", backend="torch")
    for j in range(g["input"].shape[1]):
        assert maxabs(preds[j], g[f"pred_{j}"]) <= 1e-4, j


@pytest.mark.parametrize("seed", [1, 7, 42, 2024, 31337])
def test_enhancenet_more_seeds(golden, seed):
    g = golden("enhancenet_seeds")
    sd = _seeded_reference_weights(101, 6, seed)
    assert np.allclose(_weight_sum(sd), g[f"weight_sum_{seed}"], rtol=1e-9)
    for kind in ("uniform", "gbuffer"):
        recon, _ = orc.enhancenet_forward(sd, g[f"x_{kind}"], "bilinear", "residual", 5, backend="torch")
        assert maxabs(recon, g[f"recon_{kind}_{seed}"]) <= 1e-5


def test_loadedmodel_facade_oracle(golden, weights):
    """The oracle's restatement of LoadedModel.inference against the reference's own facade (un-pickled checkpoint,
    OpenCV in-painting included), unshaded and shaded, first and second frame."""
    g = golden("loadedmodel")
    sd_s = _seeded_reference_weights(56, 3, 1234)
    for tag, sd, unshaded, init in (("unshaded", weights, True, "input"), ("shaded", sd_s, False, "zero")):
        p0 = orc.loadedmodel_inference(sd, g["render_0"], None, unshaded, init)
        assert maxabs(p0, g[f"{tag}_p0"]) <= 1e-5
        assert maxabs(orc.postprocess(g[f"{tag}_p0"], unshaded=unshaded), g[f"{tag}_state0"]) <= 1e-6
        p1 = orc.loadedmodel_inference(sd, g["render_1"], g[f"{tag}_state0"], unshaded, init)
        assert maxabs(p1, g[f"{tag}_p1"]) <= 1e-5
