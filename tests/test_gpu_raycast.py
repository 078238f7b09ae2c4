"""GPU parity of the dense-grid isosurface ray-caster (isr_raycast_iso through the C ABI) against the CPU restatement
oracle/raycast_oracle.c on the same seeded inputs, and of the DirectRenderer-style front end.  Both sides evaluate the same
fp32 expressions in the same order (CUDA -fmad=false, gcc -ffp-contract=off): the bar is BIT-EXACT for every channel."""
from argparse import Namespace

import numpy as np
import pytest
import torch

from oracle import oracle as orc
from test_raycast_oracle import make_params, sphere_volume

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def isr():
    if not torch.cuda.is_available():
        pytest.fail("GPU tests need a CUDA device")
    import isosurfacesuperresolution_b200 as pkg
    return pkg


def smooth_volume(n, seed):
    """band-limited random field in [-1, 1]: several blobs and valleys, concave regions for the AO rays"""
    rng = np.random.default_rng(seed)
    ax = (np.arange(n) + 0.5) / n - 0.5
    z, y, x = np.meshgrid(ax, ax, ax, indexing="ij")
    v = np.zeros((n, n, n))
    for _ in range(6):
        k = rng.uniform(-9, 9, 3)
        v += rng.uniform(0.3, 1.0) * np.sin(k[0] * x + k[1] * y + k[2] * z + rng.uniform(0, 6.28))
    v -= 6.0 * (x * x + y * y + z * z)          # keep the surface inside the brick
    return (v / np.abs(v).max()).astype(np.float32)


def run_gpu(isr, vol, P, W, H, skip=False, **kw):
    out = torch.full((H, W, 12), float("nan"), device="cuda")
    v = torch.from_numpy(vol).cuda()
    isr.ops.raycast_iso(v, P, out, skip=isr.ops.raycast_build_skip(v) if skip else None, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("scene", ["sphere", "two_spheres", "field"])
def test_raycast_bit_exact_against_oracle(isr, scene):
    n, W, H = 40, 96, 64                          # W, H not multiples of the 16 x 8 block on purpose below
    if scene == "sphere":
        vol, iso = sphere_volume(n, [(0.02, -0.03, 0.01)], [0.27]), 0.0
    elif scene == "two_spheres":
        vol, iso = sphere_volume(n, [(-0.17, 0, 0), (0.17, 0.02, 0)], [0.18, 0.18]), 0.0
    else:
        vol, iso = smooth_volume(n, 7), 0.1
        W, H = 90, 61
    P, _, _ = make_params(n, [0.25, 0.3, -1.4], [0.0, 0.0, 0.0], W, H, prev=([0.2, 0.33, -1.42], [0.01, 0.0, 0.0]), iso=iso,
                          step=0.05, ao_radius=0.08)
    hemi, rot = isr.ops.raycast_ao_tables()
    oh, orot = orc.raycast_ao_tables()
    assert np.array_equal(hemi, oh) and np.array_equal(rot, orot)
    ref = orc.raycast_iso(vol, P, W, H, ao_samples=8, hemisphere=hemi, rotations=rot)
    got = run_gpu(isr, vol, P, W, H, ao_samples=8, hemisphere=hemi, rotations=rot)
    assert not np.isnan(got).any()
    assert 0.05 < ref[..., 3].mean() < 0.95
    assert np.array_equal(got[..., 3], ref[..., 3]), "hit decisions differ"
    for c, name in enumerate(["r", "g", "b", "mask", "nx", "ny", "nz", "depth", "flow_x", "flow_y", "ao", "shadow"]):
        assert np.array_equal(got[..., c], ref[..., c]), f"channel {name}: max diff {np.abs(got[..., c] - ref[..., c]).max()}"
    # the empty-space grid changes which samples are evaluated, never their positions or the frame
    fast = run_gpu(isr, vol, P, W, H, skip=True, ao_samples=8, hemisphere=hemi, rotations=rot)
    assert np.array_equal(fast, ref)
    # one-pass form (a thread traces all AO rays of its pixel) and two-pass form (one thread per (pixel, ray), summed in
    # sample order; chosen automatically for small frames with many rays): same frame
    try:
        for mode in (0, 2):
            isr._lib.set_tunable("raycast_two_pass", mode)
            for skip in (False, True):
                one = run_gpu(isr, vol, P, W, H, skip=skip, ao_samples=8, hemisphere=hemi, rotations=rot)
                assert np.array_equal(one, ref), f"two_pass mode {mode}"
    finally:
        assert isr._lib.set_tunable("raycast_two_pass", 1) == 2
    many = run_gpu(isr, vol, P, W, H, skip=True, ao_samples=40, hemisphere=hemi, rotations=rot)     # auto -> two-pass
    assert np.array_equal(many, orc.raycast_iso(vol, P, W, H, ao_samples=40, hemisphere=hemi, rotations=rot))
    if scene != "sphere":
        assert ref[..., 10][ref[..., 3] == 1].min() < (0.9 if scene == "two_spheres" else 1.0)   # AO rays did hit something


def test_raycast_viewport_no_ao_and_coarse_steps(isr):
    n, W, H = 32, 64, 48
    vol = smooth_volume(n, 3)
    P, _, _ = make_params(n, [0.0, 0.2, -1.5], [0, 0, 0], W, H, iso=0.05, step=0.5)
    for kw in (dict(viewport=[5, 7, 50, 40]), dict(binary_steps=0), dict(binary_steps=3)):
        ref = orc.raycast_iso(vol, P, W, H, **kw)
        got = run_gpu(isr, vol, P, W, H, **kw)
        assert np.array_equal(got, ref)
        assert np.array_equal(run_gpu(isr, vol, P, W, H, skip=True, **kw), ref)


def test_raycast_skip_grid_is_the_cell_maximum(isr):
    vol = np.ascontiguousarray(smooth_volume(21, 11)[:, :19, :])    # ragged: 21 x 19 x 21 voxels
    nz, ny, nx = vol.shape
    skip = isr.ops.raycast_build_skip(torch.from_numpy(vol).cuda()).cpu().numpy()
    off = 0
    for cell in (8, 2):                                             # coarse grid, then the fine one
        dims = [(n + cell - 1) // cell for n in (nz, ny, nx)]
        grid = skip[off:off + int(np.prod(dims))].reshape(dims)
        off += int(np.prod(dims))
        for cz in range(dims[0]):
            for cy in range(dims[1]):
                for cx in range(dims[2]):
                    sl = tuple(slice(max(cell * c - 1, 0), min(cell * c + cell + 1, n)) for c, n in ((cz, nz), (cy, ny), (cx, nx)))
                    assert grid[cz, cy, cx] == vol[sl].max()
    dual = skip[off:off + vol.size].reshape(vol.shape)             # third level: max of each interpolation cell
    off += vol.size
    for z in range(nz):
        for y in range(ny):
            for x in range(nx):
                assert dual[z, y, x] == vol[z:z + 2, y:y + 2, x:x + 2].max()
    assert off == skip.size


def test_raycast_argument_errors(isr):
    vol = torch.zeros(8, 8, 8, device="cuda")
    out = torch.zeros(16, 16, 12, device="cuda")
    P, _, _ = make_params(8, [0, 0, -1.5], [0, 0, 0], 16, 16)
    bad = P.copy()
    bad[17] = 0.0                                                # step
    with pytest.raises(isr.ops.IsrError):
        isr.ops.raycast_iso(vol, bad, out)
    with pytest.raises(isr.ops.IsrError):
        isr.ops.raycast_iso(vol, P, out, ao_samples=4)           # tables missing
    with pytest.raises(RuntimeError):
        isr.ops.raycast_iso(vol.cpu(), P, out)                   # no CPU path


def test_dense_grid_renderer_feeds_the_network(isr, weights):
    """DirectRenderer-style front end: two frames with a camera move -> flow, then ingest + super-resolution."""
    n, W, H = 48, 64, 64
    r = isr.DenseGridRenderer("cuda")
    r.load(sphere_volume(n, [(-0.15, 0, 0), (0.15, 0.05, 0.05)], [0.17, 0.15]))
    r.send_command("resolution", "%d,%d" % (W, H))
    r.send_command("cameraOrigin", "0.1,0.3,-1.6")
    r.send_command("cameraLookAt", "0,0,0")
    r.send_command("cameraUp", "0,1,0")
    r.send_command("cameraFoV", "45")
    r.send_command("isovalue", "0.0")
    r.send_command("aosamples", "8")
    r.send_command("aoradius", "0.1")
    with pytest.raises(ValueError):
        r.send_command("nonsense", "1")
    f0 = torch.empty(H, W, 12, device="cuda")
    assert r.render_direct(f0) > 0 and r.get_time() > 0
    assert torch.all(f0[..., 8:10] == 0)                         # first frame: previous camera = current camera
    r.send_command("cameraOrigin", "0.16,0.28,-1.6")
    P1 = r.params()
    f1 = torch.empty(H, W, 12, device="cuda")
    r.render_direct(f1)
    ref = orc.raycast_iso(r.volume.cpu().numpy(), P1, W, H, ao_samples=8, hemisphere=r.hemisphere, rotations=r.rotations)
    assert np.array_equal(f1.cpu().numpy(), ref)
    m = f1[..., 3] == 1
    assert 0.05 < m.float().mean() < 0.9 and f1[..., 8:10][m].abs().max() > 1e-3
    # warping the previous mask by the flow (grid = identity + flow * (-2, +2), videotools.py:66-81) reproduces the current
    # mask away from the silhouettes: the flow convention of the renderer is the one the network's warp expects
    ys, xs = torch.meshgrid((torch.arange(H, device="cuda") + 0.5) / H * 2 - 1, (torch.arange(W, device="cuda") + 0.5) / W * 2 - 1,
                            indexing="ij")
    grid = torch.stack((xs - 2.0 * f1[..., 8], ys + 2.0 * f1[..., 9]), -1)[None]
    warped = torch.nn.functional.grid_sample(f0[..., 3][None, None], grid, mode="bilinear", padding_mode="zeros",
                                             align_corners=False)[0, 0]
    interior = (torch.nn.functional.avg_pool2d(m[None, None].float(), 7, 1, 3)[0, 0] == 1)
    assert interior.sum() > 50 and (warped[interior] - 1).abs().max() < 1e-4
    # renderer frame -> G-buffer ingest -> device flow in-painting -> recurrent network step
    opt = Namespace(upsample="bilinear", reconType="residual", useBN=False)
    net = isr.createNetwork("EnhanceNet", 4, 101, [0, 1, 2, 3, 4], 6, opt)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in weights.items()})
    lm = isr.LoadedModel({"model": net.cuda().eval(), "parameters": {"initialImage": "input"}}, torch.device("cuda"), 4)
    prev = None
    for f in (f0, f1):
        prev = lm.inference(f[None].contiguous(), prev, inpaint="gpu")
    assert prev.shape == (1, 6, 4 * H, 4 * W) and torch.isfinite(prev).all()
