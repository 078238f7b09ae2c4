"""
gen_golden.py -- generates tests/golden/*.npz from the LIVE reference (imported from
/root/reference/SuperresolutionNetwork, unmodified).  Run in the build container only; the
GPU box has no /root/reference, which is why the vectors are committed.

    python oracle/gen_golden.py            # rewrites tests/golden/

Everything is seeded; torch 2.11.0+cu128 CPU produced the committed files.
"""
import os
import sys
import warnings
from argparse import Namespace

import numpy as np
import torch

REF = "/root/reference/SuperresolutionNetwork"
sys.path.insert(0, REF)
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != HERE]
sys.path.insert(0, os.path.dirname(HERE))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

import models  # noqa: E402  (reference)
from models import VideoTools  # noqa: E402
from utils import ScreenSpaceShading, initialImage  # noqa: E402
from oracle import oracle as orc  # noqa: E402  (only for the synthetic input generator)

warnings.filterwarnings("ignore")


def ref_net(cin, cout, mask, upsample, seed):
    torch.manual_seed(seed)
    opt = Namespace(upsample=upsample, reconType="residual", useBN=False)
    return models.createNetwork("EnhanceNet", 4, cin, mask, cout, opt).eval()


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    g = torch.Generator().manual_seed(4321)

    # ---- flatten_high (bit-exact) -----------------------------------------------------------
    x = torch.randn(2, 6, 24, 40, generator=g)
    np.savez_compressed(os.path.join(OUT, "flatten_high.npz"), x=x.numpy(),
                        y=VideoTools.flatten_high(x, 4).numpy())

    # ---- warp_upscale ------------------------------------------------------------------------
    cases = {}
    B, C, h, w = 2, 6, 18, 30
    img = torch.rand(B, C, 4 * h, 4 * w, generator=g) * 2 - 1
    flows = {
        "zero": torch.zeros(B, 2, h, w),
        "const": torch.full((B, 2, h, w), 0.013),
        "smooth": torch.from_numpy(orc.synthetic_gbuffer(B, h, w, 7)[1][:, 0]),
        "wild": (torch.rand(B, 2, h, w, generator=g) - 0.5) * 1.5,   # far out of range -> padding
    }
    for name, fl in flows.items():
        for sm in (False, True):
            VideoTools._offset_cache.clear()
            cases[f"{name}_{int(sm)}"] = VideoTools.warp_upscale(img, fl, 4, special_mask=sm).numpy()
        cases[f"flow_{name}"] = fl.numpy()
    np.savez_compressed(os.path.join(OUT, "warp_upscale.npz"), img=img.numpy(), **cases)
    # non-square 3-channel (shaded variant) + 270x480-like aspect
    img3 = torch.rand(1, 3, 4 * 27, 4 * 48, generator=g)
    fl3 = torch.from_numpy(orc.synthetic_gbuffer(1, 27, 48, 9)[1][:, 0])
    VideoTools._offset_cache.clear()
    np.savez_compressed(os.path.join(OUT, "warp_upscale_rgb.npz"), img=img3.numpy(), flow=fl3.numpy(),
                        out=VideoTools.warp_upscale(img3, fl3, 4, special_mask=False).numpy())

    # ---- initialImage ------------------------------------------------------------------------
    cur = torch.rand(2, 5, 9, 11, generator=g) * 2 - 1
    ii = {"cur": cur.numpy()}
    for mode in ("zero", "unshaded", "input"):
        for ch in (5, 6):
            ii[f"{mode}_{ch}"] = initialImage(cur, ch, mode, False, 4).contiguous().numpy()
    ii["unshaded_6_inv"] = initialImage(cur, 6, "unshaded", True, 4).contiguous().numpy()
    ii["input_3"] = initialImage(cur, 3, "input", False, 4).contiguous().numpy()
    np.savez_compressed(os.path.join(OUT, "initial_image.npz"), **ii)

    # ---- shading + normalize -----------------------------------------------------------------
    low, _, ao = orc.synthetic_gbuffer(2, 40, 56, 11)
    gb = np.concatenate([low[:, 0], ao[:, 0]], axis=1)
    gb = gb + 0.2 * torch.randn(*gb.shape, generator=g).numpy()      # out-of-range mask / ao, unnormalised normals
    gbt = torch.from_numpy(gb.astype(np.float32))
    sh = {"gbuffer": gbt.numpy()}
    configs = {
        "train": dict(ambient=(0.1,) * 3, diffuse=(1.0,) * 3, specular=(0.2,) * 3, exponent=16, light=(0.1, 0.1, 1.0),
                      material=(1.0, 1.0, 1.0), ao=1.0, background=(0.0, 0.0, 0.0), spec=True, inv=False),
        "gui": dict(ambient=(0.1,) * 3, diffuse=(0.9, 0.8, 1.0), specular=(0.2, 0.3, 0.1), exponent=8, light=(0.3, -0.2, 0.9),
                    material=(1.0, 0.3, 0.0), ao=0.5, background=(1.0, 1.0, 1.0), spec=True, inv=True),
        "nospec": dict(ambient=(0.05, 0.1, 0.2), diffuse=(1.0,) * 3, specular=(0.0,) * 3, exponent=4, light=(0.0, 0.0, 1.0),
                       material=(0.2, 0.9, 0.4), ao=0.8, background=(0.2, 0.4, 0.6), spec=False, inv=False),
    }
    for name, c in configs.items():
        s = ScreenSpaceShading(torch.device("cpu"))
        s.fov(30)
        s.ambient_light_color(np.array(c["ambient"]))
        s.diffuse_light_color(np.array(c["diffuse"]))
        s.specular_light_color(np.array(c["specular"]))
        s.specular_exponent(c["exponent"])
        s.light_direction(np.array(c["light"]))
        s.material_color(np.array(c["material"]))
        s.ambient_occlusion(c["ao"])
        s.background(np.array(c["background"]))
        s.enable_specular = c["spec"]
        s.inverse_ao = c["inv"]
        sh[f"{name}_c6"] = s(gbt).numpy()
        sh[f"{name}_c5"] = s(gbt[:, 0:5]).numpy()
        sh[f"{name}_params"] = orc.shading_params(c["ambient"], c["diffuse"], c["specular"], c["light"], c["material"],
                                                  c["background"], c["ao"], c["exponent"], c["spec"], c["inv"])
    nrm = torch.cat([gbt[:, 1:4], torch.zeros(2, 3, 40, 1)], dim=3)   # includes exact zero vectors
    sh["normalize_in"] = nrm.numpy()
    sh["normalize_out"] = ScreenSpaceShading.normalize(nrm, dim=1).numpy()
    post = torch.cat([torch.clamp(gbt[:, 0:1], -1, 1), ScreenSpaceShading.normalize(gbt[:, 1:4], dim=1),
                      torch.clamp(gbt[:, 4:5], 0, 1), torch.clamp(gbt[:, 5:6], 0, 1)], dim=1)
    sh["postprocess"] = post.numpy()
    np.savez_compressed(os.path.join(OUT, "shading.npz"), **sh)

    # ---- EnhanceNet forward ------------------------------------------------------------------
    # weights: reference constructor under torch.manual_seed(1234); the full state_dict is stored once
    # (unshaded/bilinear); the other variants store a checksum and are rebuilt by the drop-in ctor.
    net = ref_net(101, 6, [0, 1, 2, 3, 4], "bilinear", 1234)
    sd = {k: v.numpy() for k, v in net.state_dict().items()}
    np.savez_compressed(os.path.join(OUT, "enhancenet_weights_unshaded_seed1234.npz"), **sd)
    xin = torch.rand(1, 101, 16, 20, generator=g) * 2 - 1
    en = {"x": xin.numpy()}
    with torch.no_grad():
        for ups in ("bilinear", "nearest"):
            net.upsample = ups
            net.postblock[0].mode = ups
            net.postblock[3].mode = ups
            recon, raw = net(xin)
            en[f"recon_{ups}"], en[f"raw_{ups}"] = recon.numpy(), raw.numpy()
        net.upsample = "bilinear"; net.postblock[0].mode = "bilinear"; net.postblock[3].mode = "bilinear"
        net.recon_type = "direct"
        en["recon_direct"] = net(xin)[0].numpy()
        net.recon_type = "residual"
    # shaded variant (56 -> 3), weights by seed; checksum only
    net_s = ref_net(56, 3, [0, 1, 2], "bilinear", 1234)
    xs = torch.rand(1, 56, 12, 16, generator=g)
    with torch.no_grad():
        rs, ws = net_s(xs)
    en["shaded_x"], en["shaded_recon"], en["shaded_raw"] = xs.numpy(), rs.numpy(), ws.numpy()
    en["shaded_weight_sum"] = np.array([float(sum(v.double().sum() for v in net_s.state_dict().values())),
                                        float(sum(v.double().abs().sum() for v in net_s.state_dict().values()))])
    en["unshaded_weight_sum"] = np.array([float(sum(v.double().sum() for v in net.state_dict().values())),
                                          float(sum(v.double().abs().sum() for v in net.state_dict().values()))])
    np.savez_compressed(os.path.join(OUT, "enhancenet_forward.npz"), **en)

    # ---- 3-frame unshaded recurrence (mainVideoUnshaded.py:417-465) + shading -----------------
    low, flow, _ = orc.synthetic_gbuffer(1, 16, 24, 21, T=3)
    tl, tf = torch.from_numpy(low), torch.from_numpy(flow)
    shd = ScreenSpaceShading(torch.device("cpu"))
    shd.fov(30); shd.ambient_light_color(np.array([0.1, 0.1, 0.1])); shd.diffuse_light_color(np.array([1.0, 1.0, 1.0]))
    shd.specular_light_color(np.array([0.2, 0.2, 0.2])); shd.specular_exponent(16)
    shd.light_direction(np.array([0.1, 0.1, 1.0])); shd.material_color(np.array([1.0, 0.3, 0.0]))
    shd.ambient_occlusion(0.5); shd.background(np.array([1.0, 1.0, 1.0]))
    seq = {"low": low, "flow": flow}
    prev = None
    with torch.no_grad():
        for j in range(3):
            inp = tl[:, j]
            if j == 0:
                pw = initialImage(inp, 6, "input", False, 4)
            else:
                VideoTools._offset_cache.clear()
                pw = VideoTools.warp_upscale(prev, tf[:, j - 1], 4, special_mask=True)
            single = torch.cat((inp, VideoTools.flatten_high(pw, 4)), dim=1)
            pred, _ = net(single)
            prev = torch.cat([torch.clamp(pred[:, 0:1], -1, 1), ScreenSpaceShading.normalize(pred[:, 1:4], dim=1),
                              torch.clamp(pred[:, 4:5], 0, 1), torch.clamp(pred[:, 5:6], 0, 1)], dim=1)
            seq[f"pred_{j}"] = pred.numpy()
            seq[f"rgb_{j}"] = shd(prev).numpy()
    seq["shade_params"] = orc.shading_params(material=(1.0, 0.3, 0.0), ao_strength=0.5, background=(1.0, 1.0, 1.0))
    np.savez_compressed(os.path.join(OUT, "sequence_unshaded.npz"), **seq)

    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
