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
