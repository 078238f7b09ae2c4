"""CPU checks of the dense-grid ray-caster restatement (oracle/raycast_oracle.c) against ANALYTIC answers on a sphere, and
of the host-side camera (renderer.camera_matrices).  The reference's renderer cannot run here (GVDB binary, no volumes):
parity of this row is unpinned by the reference; these tests pin the restatement to geometry instead."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle  # noqa: E402
from isosurfacesuperresolution_b200._lib import (RC_AMBIENT, RC_AO_BIAS, RC_AO_RADIUS, RC_BOX_MIN, RC_CAM_POS, RC_CAM_S,  # noqa: E402
                                                 RC_CAM_U, RC_CAM_V, RC_DIFFUSE, RC_ISO, RC_LIGHT, RC_NORMAL_MAT, RC_PARAMS,
                                                 RC_SPEC_EXP, RC_SPECULAR, RC_STEP, RC_VIEW_CUR, RC_VIEW_NEXT, RC_VOXEL)
from isosurfacesuperresolution_b200.renderer import camera_matrices  # noqa: E402


def sphere_volume(n, centres, radii):
    """value = max_k (R_k - |x - c_k|) in world units on an n^3 brick spanning [-0.5, 0.5]^3; surface at iso 0."""
    voxel = 1.0 / n
    ax = (np.arange(n, dtype=np.float64) + 0.5) * voxel - 0.5
    z, y, x = np.meshgrid(ax, ax, ax, indexing="ij")
    v = np.full((n, n, n), -1e9)
    for c, r in zip(centres, radii):
        v = np.maximum(v, r - np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2))
    return v.astype(np.float32)


def make_params(n, origin, look_at, W, H, prev=None, fov=45.0, step=0.05, iso=0.0, ao_radius=0.05):
    P = np.zeros(RC_PARAMS, np.float32)
    up = [0.0, 1.0, 0.0]
    view, proj, S, U, V = camera_matrices(origin, look_at, up, fov, W / H)
    po, pl = prev if prev is not None else (origin, look_at)
    view_n, proj_n, _, _, _ = camera_matrices(po, pl, up, fov, W / H)
    P[RC_CAM_POS:RC_CAM_POS + 3] = origin
    P[RC_CAM_U:RC_CAM_U + 3], P[RC_CAM_V:RC_CAM_V + 3], P[RC_CAM_S:RC_CAM_S + 3] = U, V, S
    P[RC_BOX_MIN:RC_BOX_MIN + 3] = -0.5
    P[RC_VOXEL], P[RC_ISO], P[RC_STEP] = 1.0 / n, iso, step
    P[RC_VIEW_CUR:RC_VIEW_CUR + 16] = (proj @ view).reshape(-1)
    P[RC_VIEW_NEXT:RC_VIEW_NEXT + 16] = (proj_n @ view_n).reshape(-1)
    P[RC_NORMAL_MAT:RC_NORMAL_MAT + 16] = view.reshape(-1)
    light = np.asarray(look_at, np.float64) - np.asarray(origin, np.float64)
    P[RC_LIGHT:RC_LIGHT + 3] = light / np.linalg.norm(light)
    P[RC_AMBIENT:RC_AMBIENT + 3], P[RC_DIFFUSE:RC_DIFFUSE + 3], P[RC_SPECULAR:RC_SPECULAR + 3] = 0.1, 0.7, 1.0
    P[RC_SPEC_EXP], P[RC_AO_RADIUS], P[RC_AO_BIAS] = 32, ao_radius, 1e-3
    return P, (view, proj, S, U, V), (view_n, proj_n)


def analytic_sphere(origin, S, U, V, W, H, R):
    ys, xs = np.meshgrid((np.arange(H) + 0.5) / H, (np.arange(W) + 0.5) / W, indexing="ij")
    d = xs[..., None] * U + ys[..., None] * V + S
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    o = np.asarray(origin, np.float64)
    b = d @ o
    disc = b * b - (o @ o - R * R)
    t = -b - np.sqrt(np.maximum(disc, 0))
    hit = o + t[..., None] * d
    return disc, hit


def test_camera_rays_and_projection_agree():
    W, H = 64, 48
    view, proj, S, U, V = camera_matrices([0.3, 0.4, -1.2], [0.05, 0.0, 0.1], [0, 1, 0], 40.0, W / H)
    vp = proj @ view
    for (x, y) in [(0, 0), (63, 0), (10, 40), (31, 23)]:
        sx, sy = (x + 0.5) / W, (y + 0.5) / H
        d = sx * U + sy * V + S
        X = np.array([0.3, 0.4, -1.2]) + 0.9 * d / np.linalg.norm(d)
        q = vp @ np.append(X, 1.0)
        assert np.allclose(q[:2] / q[3], [2 * sx - 1, 1 - 2 * sy], atol=1e-9)   # image row 0 is the top
    assert np.allclose(view[:3, :3] @ view[:3, :3].T, np.eye(3), atol=1e-12)


def test_sphere_mask_depth_normal_flow_against_geometry():
    n, W, H, R = 48, 64, 48, 0.2
    vol = sphere_volume(n, [(0, 0, 0)], [R])
    origin, look = [0.2, 0.3, -1.4], [0.0, 0.0, 0.0]
    prev = ([0.25, 0.28, -1.4], [0.0, 0.02, 0.0])
    P, (view, proj, S, U, V), (view_n, proj_n) = make_params(n, origin, look, W, H, prev=prev)
    out = oracle.raycast_iso(vol, P, W, H)
    disc, hit = analytic_sphere(origin, S, U, V, W, H, R)
    mask = out[..., 3]
    assert set(np.unique(mask)) == {0.0, 1.0}
    inside = disc > (2.0 / n) ** 2 * 0 + 0.01          # well inside the silhouette
    outside = disc < -0.01
    assert np.all(mask[inside] == 1) and np.all(mask[outside] == 0)
    assert 0.02 < mask.mean() < 0.8
    # misses keep the defaults of render_kernel.cu:224-230
    miss = mask == 0
    assert np.all(out[miss][:, [0, 1, 2, 4, 5, 6, 7, 8, 9]] == 0) and np.all(out[miss][:, 10:] == 1)
    # depth, normal, flow at pixels well inside (the trilinear distance field is accurate to O(voxel^2 / R))
    hh = np.concatenate([hit, np.ones(hit.shape[:2] + (1,))], -1)
    q = hh @ (proj @ view).T
    qn = hh @ (proj_n @ view_n).T
    depth_a = q[..., 2] / q[..., 3]
    flow_a = 0.5 * (q[..., :2] / q[..., 3:4] - qn[..., :2] / qn[..., 3:4])
    normal_a = (hit / np.linalg.norm(hit, axis=-1, keepdims=True)) @ view[:3, :3].T
    assert np.abs(out[..., 7] - depth_a)[inside].max() < 2e-3
    assert np.abs(out[..., 8:10] - flow_a)[inside].max() < 2e-3
    assert np.abs(out[..., 4:7] - normal_a)[inside].max() < 0.06
    assert np.abs(np.linalg.norm(out[..., 4:7], axis=-1)[mask == 1] - 1).max() < 1e-5
    assert np.abs(flow_a[inside]).max() > 0.01          # the camera did move
    # colour: ambient + diffuse |n.l| + specular term, all within [ambient, ambient + diffuse + specular * 34 / 6.82]
    rgb = out[..., :3][mask == 1]
    assert rgb.min() >= 0.1 - 1e-6 and np.all(rgb[:, 0] == rgb[:, 1])


def test_static_camera_has_zero_flow_and_viewport_clips():
    n, W, H = 32, 40, 40
    vol = sphere_volume(n, [(0, 0, 0)], [0.3])
    P, _, _ = make_params(n, [0.0, 0.0, -1.5], [0, 0, 0], W, H)
    full = oracle.raycast_iso(vol, P, W, H)
    assert np.all(full[..., 8:10] == 0) and full[..., 3].sum() > 20
    part = oracle.raycast_iso(vol, P, W, H, viewport=[10, 12, 30, 28])
    win = np.zeros((H, W), bool)
    win[12:28, 10:30] = True
    assert np.array_equal(part[win], full[win])
    assert np.all(part[~win][:, :10] == 0) and np.all(part[~win][:, 10:] == 1)


def test_ambient_occlusion_convex_is_open_and_contact_is_dark():
    n, W, H = 48, 48, 48
    hemi, rot = oracle.raycast_ao_tables()
    assert hemi.shape == (512, 4) and np.all(hemi[:, 2] >= 0) and np.allclose(np.linalg.norm(rot[:, :2], axis=1), 1, atol=1e-6)
    P, _, _ = make_params(n, [0.0, 0.0, -1.5], [0, 0, 0], W, H, step=0.1, ao_radius=0.1)
    one = oracle.raycast_iso(sphere_volume(n, [(0, 0, 0)], [0.3]), P, W, H, ao_samples=16, hemisphere=hemi, rotations=rot)
    m = one[..., 3] == 1
    assert one[..., 10][m].mean() > 0.97                 # a convex body does not occlude itself
    two = oracle.raycast_iso(sphere_volume(n, [(-0.17, 0, 0), (0.17, 0, 0)], [0.18, 0.18]), P, W, H, ao_samples=16,
                             hemisphere=hemi, rotations=rot)
    m2 = two[..., 3] == 1
    ao = two[..., 10]
    assert ao[m2].min() < 0.8 and ao[m2].max() == 1.0
    # the darkest pixels sit in the crease between the two spheres (image centre column)
    ys, xs = np.nonzero(m2 & (ao < 0.85))
    assert len(xs) > 0 and np.abs(xs - (W - 1) / 2).mean() < W * 0.12
    assert np.all((ao >= 0) & (ao <= 1))


def test_binary_search_refines_the_hit():
    n, W, H, R = 32, 32, 32, 0.3
    vol = sphere_volume(n, [(0, 0, 0)], [R])
    origin = [0.0, 0.0, -1.5]
    P, (view, proj, S, U, V), _ = make_params(n, origin, [0, 0, 0], W, H, step=0.5)
    disc, hit = analytic_sphere(origin, S, U, V, W, H, R)
    hh = np.concatenate([hit, np.ones(hit.shape[:2] + (1,))], -1)
    q = hh @ (proj @ view).T
    depth_a = q[..., 2] / q[..., 3]
    inside = disc > 0.02
    err = {}
    for steps in (0, 10):
        out = oracle.raycast_iso(vol, P, W, H, binary_steps=steps)
        err[steps] = np.abs(out[..., 7] - depth_a)[inside & (out[..., 3] == 1)].mean()
    assert err[10] < 0.25 * err[0]
