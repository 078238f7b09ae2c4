"""
gen_golden_r02.py -- round-2 fixtures from the LIVE reference (imported unmodified from
/root/reference/SuperresolutionNetwork).  Run in the build container only (the GPU box has no /root/reference):

    python oracle/gen_golden_r02.py        # writes tests/golden/{sequence_unshaded_10,sequence_shaded,loadedmodel,
                                           #                      enhancenet_seeds}.npz and checkpoint_unshaded_ref.pth

* sequence_unshaded_10: the 10-frame unshaded recurrence of mainVideoUnshaded.py:419-465 (configs[1] has 10 frames); its
  first three frames equal tests/golden/sequence_unshaded.npz.
* sequence_shaded: the shaded (RGB) recurrence of mainVideo.py:383-434 in the form the inference facade runs it
  (loadedmodel.py:97-119): 56 -> 3 network, previous RGB frame warped with special_mask=False, clamp(pred, 0, 1) as the
  state, zero initial image.
* loadedmodel: the reference's own LoadedModel.inference (checkpoint un-pickle, layout sniffing, OpenCV flow
  in-painting, initial image / warp, flatten, concat, model) for an unshaded and a shaded checkpoint, two frames each.
* enhancenet_seeds: forward passes of reference-constructed networks for more weight seeds (the weights themselves are
  rebuilt by the drop-in's seeded constructor and pinned by their checksums).
* checkpoint_unshaded_ref.pth: a whole-module checkpoint of the REFERENCE class, written the way
  mainVideoUnshaded.py:799-811 writes it; un-pickling it resolves `models.enhancenet.EnhanceNet`.
"""
import os
import sys
import tempfile
import warnings
from argparse import Namespace

os.environ["TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD"] = "1"     # the reference calls a bare torch.load(name) (loadedmodel.py:16)

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference/SuperresolutionNetwork"
sys.path.insert(0, REF)
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != HERE]
sys.path.insert(0, os.path.dirname(HERE))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

import models  # noqa: E402  (reference)
from models import VideoTools  # noqa: E402
from utils import ScreenSpaceShading, initialImage  # noqa: E402
from inference.loadedmodel import LoadedModel  # noqa: E402  (reference; needs cv2)
from oracle import oracle as orc  # noqa: E402  (synthetic input generator only)

warnings.filterwarnings("ignore")
SEEDS = (1, 7, 42, 2024, 31337)


def ref_net(cin, cout, mask, upsample, seed):
    torch.manual_seed(seed)
    opt = Namespace(upsample=upsample, reconType="residual", useBN=False)
    return models.createNetwork("EnhanceNet", 4, cin, mask, cout, opt).eval()


def checksum(net):
    sd = net.state_dict()
    return np.array([float(sum(v.double().sum() for v in sd.values())),
                     float(sum(v.double().abs().sum() for v in sd.values()))])


def shaded_input(low, ao):
    """SURVEY 8d shaded variant: RGB = reference shading of the G-buffer, mask in [0,1], normal, depth, all clamped."""
    shd = ScreenSpaceShading(torch.device("cpu"))
    shd.fov(30); shd.ambient_light_color(np.array([0.1, 0.1, 0.1])); shd.diffuse_light_color(np.array([1.0, 1.0, 1.0]))
    shd.specular_light_color(np.array([0.2, 0.2, 0.2])); shd.specular_exponent(16)
    shd.light_direction(np.array([0.1, 0.1, 1.0])); shd.material_color(np.array([1.0, 0.3, 0.0]))
    shd.ambient_occlusion(0.5); shd.background(np.array([0.0, 0.0, 0.0]))
    gb = torch.from_numpy(np.concatenate([low, ao], axis=1))
    rgb = shd(gb)
    return torch.clamp(torch.cat([rgb, gb[:, 0:1] * 0.5 + 0.5, gb[:, 1:4], gb[:, 4:5]], dim=1), 0, 1)


def main():
    torch.set_num_threads(8)
    g = torch.Generator().manual_seed(98765)

    # ---- 10-frame unshaded recurrence (mainVideoUnshaded.py:419-465) -------------------------------------------------
    net = ref_net(101, 6, [0, 1, 2, 3, 4], "bilinear", 1234)
    low, flow, _ = orc.synthetic_gbuffer(1, 16, 24, 21, T=10)
    tl, tf = torch.from_numpy(low), torch.from_numpy(flow)
    seq = {"low": low, "flow": flow}
    prev = None
    with torch.no_grad():
        for j in range(10):
            inp = tl[:, j]
            if j == 0:
                pw = initialImage(inp, 6, "input", False, 4)
            else:
                VideoTools._offset_cache.clear()
                pw = VideoTools.warp_upscale(prev, tf[:, j - 1], 4, special_mask=True)
            pred, _ = net(torch.cat((inp, VideoTools.flatten_high(pw, 4)), dim=1))
            prev = torch.cat([torch.clamp(pred[:, 0:1], -1, 1), ScreenSpaceShading.normalize(pred[:, 1:4], dim=1),
                              torch.clamp(pred[:, 4:5], 0, 1), torch.clamp(pred[:, 5:6], 0, 1)], dim=1)
            seq[f"pred_{j}"] = pred.numpy()
    old = dict(np.load(os.path.join(OUT, "sequence_unshaded.npz")))
    for j in range(3):
        assert np.array_equal(old[f"pred_{j}"], seq[f"pred_{j}"]), "10-frame fixture must extend the 3-frame one"
    np.savez_compressed(os.path.join(OUT, "sequence_unshaded_10.npz"), **seq)

    # ---- whole-module checkpoint of the reference class (mainVideoUnshaded.py:799-811) ---------------------------------
    params = {"upsample": "bilinear", "reconType": "residual", "useBN": False, "initialImage": "input", "unshaded": True,
              "aoInverted": False, "upscale_factor": 4, "model": "EnhanceNet"}
    ckpt_path = os.path.join(OUT, "checkpoint_unshaded_ref.pth")
    torch.save({"epoch": 1, "model": net, "parameters": params}, ckpt_path)

    # ---- shaded recurrence (mainVideo.py:383-434 as run by loadedmodel.py:97-119) ------------------------------------
    net_s = ref_net(56, 3, [0, 1, 2], "bilinear", 1234)
    T = 4
    low, flow, ao = orc.synthetic_gbuffer(1, 16, 24, 23, T=T)
    tf = torch.from_numpy(flow)
    inputs = torch.stack([shaded_input(low[:, j], ao[:, j]) for j in range(T)], dim=1)     # 1 x T x 8 x h x w
    sq = {"input": inputs.numpy(), "flow": flow, "weight_sum": checksum(net_s)}
    prev = None
    with torch.no_grad():
        for j in range(T):
            inp = inputs[:, j]
            if j == 0:
                pw = torch.zeros(1, 3, 64, 96)                                              # mainVideo.py:385-388
            else:
                VideoTools._offset_cache.clear()
                pw = VideoTools.warp_upscale(prev, tf[:, j - 1], 4, special_mask=False)     # loadedmodel.py:111-115
            pred, _ = net_s(torch.cat((inp, VideoTools.flatten_high(pw, 4)), dim=1))
            prev = torch.clamp(pred, 0, 1)                                                  # mainVideo.py:416,434
            sq[f"pred_{j}"] = pred.numpy()
    np.savez_compressed(os.path.join(OUT, "sequence_shaded.npz"), **sq)

    # ---- the reference's LoadedModel.inference, unshaded and shaded ----------------------------------------------------
    low, flow, ao = orc.synthetic_gbuffer(1, 16, 24, 25, T=2)
    lm_out = {}
    renders = []
    for j in range(2):
        r = torch.rand(1, 12, 16, 24, generator=g)
        sin = shaded_input(low[:, j], ao[:, j])
        m01 = (torch.from_numpy(low[:, j, 0:1]) > 0).float()
        r[:, 0:3] = sin[:, 0:3]
        r[:, 3:4] = m01                                        # renderer mask: exactly 0 outside the surface
        r[:, 4:8] = torch.from_numpy(low[:, j, 1:5])
        r[:, 8:10] = torch.from_numpy(flow[:, j]) * m01        # no flow outside the surface -> cv.inpaint fills it
        r[:, 10:11] = torch.from_numpy(ao[:, j])
        renders.append(r)
        lm_out[f"render_{j}"] = r.numpy()
    with tempfile.TemporaryDirectory() as tmp:
        p_s = os.path.join(tmp, "shaded.pth")
        torch.save({"epoch": 1, "model": net_s, "parameters": {"initialImage": "zero"}}, p_s)
        for tag, path in (("unshaded", ckpt_path), ("shaded", p_s)):
            lm = LoadedModel(path, torch.device("cpu"), 4)
            assert lm.unshaded == (tag == "unshaded")
            if tag == "shaded":
                assert lm.has_normal and lm.has_depth and lm.initial_image_mode == "zero"
            VideoTools._offset_cache.clear()
            p0 = lm.inference(renders[0], None)
            if tag == "unshaded":                              # mainGUI.py:594-599
                st = torch.cat([torch.clamp(p0[:, 0:1], -1, 1), ScreenSpaceShading.normalize(p0[:, 1:4], dim=1),
                                torch.clamp(p0[:, 4:], 0, 1)], dim=1)
            else:                                              # mainVideo.py:416
                st = torch.clamp(p0, 0, 1)
            VideoTools._offset_cache.clear()
            p1 = lm.inference(renders[1], st)
            lm_out[f"{tag}_p0"], lm_out[f"{tag}_state0"], lm_out[f"{tag}_p1"] = p0.numpy(), st.numpy(), p1.numpy()
    np.savez_compressed(os.path.join(OUT, "loadedmodel.npz"), **lm_out)

    # ---- more weight seeds: forward of the reference-constructed network ----------------------------------------------
    xin = torch.rand(1, 101, 16, 20, generator=g) * 2 - 1
    low, _, ao = orc.synthetic_gbuffer(1, 16, 20, 27)
    xg = xin.clone()
    xg[:, 0:5] = torch.from_numpy(low[:, 0])                   # a G-buffer-like input next to the uniform stress input
    sd = {"x_uniform": xin.numpy(), "x_gbuffer": xg.numpy(), "seeds": np.array(SEEDS)}
    with torch.no_grad():
        for seed in SEEDS:
            n = ref_net(101, 6, [0, 1, 2, 3, 4], "bilinear", seed)
            sd[f"weight_sum_{seed}"] = checksum(n)
            sd[f"recon_uniform_{seed}"] = n(xin)[0].numpy()
            sd[f"recon_gbuffer_{seed}"] = n(xg)[0].numpy()
    np.savez_compressed(os.path.join(OUT, "enhancenet_seeds.npz"), **sd)

    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
