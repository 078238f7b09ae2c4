"""GPU probe: which UMMA shared-memory addressing variant of the conv kernel is correct on this part?
Runs every (halo_pitch, base_offset_mode) variant in its own subprocess (a trapped kernel poisons the CUDA context) and
prints max-abs error against torch's fp32 conv2d.  Usage: python tools/probe_conv.py [--out gpurun_out/probe.json]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r"""
import sys, json, torch
sys.path.insert(0, %r)
from isosurfacesuperresolution_b200 import ops, _lib
pitch, bo = int(sys.argv[1]), int(sys.argv[2])
_lib.set_tunable("halo_pitch", pitch); _lib.set_tunable("base_offset_mode", bo)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
g = torch.Generator(device="cpu").manual_seed(5)
res = {}
for (B, H, W) in [(1, 16, 8), (1, 32, 24), (2, 37, 29)]:
    x = (torch.rand(B, 64, H, W, generator=g) * 2 - 1).cuda()
    w = (torch.randn(64, 64, 3, 3, generator=g) * 0.05).cuda()
    b = torch.randn(64, generator=g).cuda()
    ref = torch.nn.functional.conv2d(x.double(), w.double(), b.double(), padding=1).float()
    for fmt, na, nw in [("bf16", 1, 1), ("fp16", 1, 1), ("bf16", 2, 2), ("fp16", 2, 2), ("fp16", 2, 1)]:
        out = ops.debug_conv3x3(x, w, b, fmt=fmt, n_act_parts=na, n_w_parts=nw)
        res["%%dx%%dx%%d_%%s_%%d%%d" %% (B, H, W, fmt, na, nw)] = float((out - ref).abs().max())
    # epilogue variants: relu + addend, final (cout=6)
    add = torch.randn(B, 64, H, W, generator=g).cuda()
    out = ops.debug_conv3x3(x, w, b, addend=add, relu=True, fmt="bf16", n_act_parts=2, n_w_parts=2)
    res["%%dx%%dx%%d_add_relu" %% (B, H, W)] = float((out - torch.relu(ref + add)).abs().max())
    w6 = (torch.randn(6, 64, 3, 3, generator=g) * 0.05).cuda(); b6 = torch.randn(6, generator=g).cuda()
    ref6 = torch.nn.functional.conv2d(x.double(), w6.double(), b6.double(), padding=1).float()
    out6 = ops.debug_conv3x3(x, w6, b6, fmt="fp16", n_act_parts=1, n_w_parts=1)
    res["%%dx%%dx%%d_final6" %% (B, H, W)] = float((out6 - ref6).abs().max())
torch.cuda.synchronize()
print("RESULT " + json.dumps(res))
""" % ROOT


def main():
    out_path = None
    if "--out" in sys.argv:
        out_path = sys.argv[sys.argv.index("--out") + 1]
    summary = {}
    for pitch in (10, 16):
        for bo in (0, 1):
            key = f"pitch{pitch}_bo{bo}"
            try:
                r = subprocess.run([sys.executable, "-c", CHILD, str(pitch), str(bo)], capture_output=True, text=True,
                                   timeout=240)
                line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
                if line:
                    summary[key] = json.loads(line[-1][7:])
                else:
                    summary[key] = {"error": (r.stderr or r.stdout)[-600:]}
            except subprocess.TimeoutExpired:
                summary[key] = {"error": "timeout"}
            print(key, json.dumps(summary[key]), flush=True)
    if out_path:
        os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
        with open(out_path, "w") as f:
            json.dump(summary, f, indent=1)


if __name__ == "__main__":
    main()
