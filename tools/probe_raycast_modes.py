"""Ray-caster: one-pass against two-pass ambient occlusion (tunable raycast_two_pass 0 / 2) over frame sizes and ray counts."""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import isosurfacesuperresolution_b200 as isr
from interactive_loop import blob_volume
for n in (128, 256):
    r = isr.DenseGridRenderer("cuda"); r.load(blob_volume(n))
    r.send_command("isovalue", "0.1"); r.send_command("cameraOrigin", "0.3,0.4,-1.8"); r.send_command("aoradius", "0.05")
    for (W, H) in ((480, 270), (960, 540), (1920, 1080)):
        for ao in (4, 8, 16, 32, 64):
            r.send_command("resolution", "%d,%d" % (W, H)); r.send_command("aosamples", str(ao))
            f = torch.empty(H, W, 12, device="cuda")
            res = []
            for mode in (0, 2):
                isr._lib.set_tunable("raycast_two_pass", mode)
                r.render_direct(f); res.append(min(r.render_direct(f) for _ in range(3)) * 1e3)
                chk = float(f.double().sum())
            isr._lib.set_tunable("raycast_two_pass", 1)
            print(n, W, H, ao, "one-pass %.2f ms  two-pass %.2f ms" % tuple(res), "sum %.4f" % chk, flush=True)
