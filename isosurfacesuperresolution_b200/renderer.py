"""Headless dense-grid isosurface renderer with the reference's ``DirectRenderer`` interface
(SuperresolutionNetwork/inference/renderer.py:78-117: ``load`` / ``send_command(cmd, value)`` / ``render_direct(tensor)``
/ ``get_time`` / ``close``; commands of GPURendererDirect.cpp:395-426), on ``isr_raycast_iso``.

The reference renders sparse GVDB volumes (closed Windows binary, no volume shipped); this one renders ONE dense fp32
brick (a ``.npy`` file or an array, nz x ny x nx), centred at the origin with its longest side ``world_size`` long.  The
camera follows GVDB's ``Camera3D`` (third-party/include/gvdb/gvdb_camera.cpp:425-485: gluLookAt basis, a projection with
``proj[0][0] = 2 near / (tan(fov/2) near)``, near 0.1, far 5000; ray corners :598-601), the per-frame uploads follow
``render`` in GPURendererDirect.cpp:309-365: flow is measured against the camera of the PREVIOUS ``render_direct`` call
(:432-443).  The frame is written into a caller-allocated H x W x 12 CUDA tensor; the call synchronises and returns
seconds, as the reference's ``render(CUdeviceptr)`` does.
"""
import math
import time

import numpy as np
import torch

from . import ops
from ._lib import (RC_AMBIENT, RC_AO_BIAS, RC_AO_RADIUS, RC_BOX_MIN, RC_CAM_POS, RC_CAM_S, RC_CAM_U, RC_CAM_V, RC_DIFFUSE,
                   RC_ISO, RC_LIGHT, RC_NORMAL_MAT, RC_PARAMS, RC_SPEC_EXP, RC_SPECULAR, RC_STEP, RC_VIEW_CUR,
                   RC_VIEW_NEXT, RC_VOXEL)

NEAR, FAR = 0.1, 5000.0   # gvdb_camera.cpp:59-60


def _unit(v):
    v = np.asarray(v, np.float64)
    return v / np.linalg.norm(v)


def camera_matrices(origin, look_at, up, fov_deg, aspect):
    """(view 4x4, proj 4x4, S, U, V): GVDB's view and projection matrices (math convention, column vectors) and the ray
    basis ``ray(sx, sy) = normalize(sx * U + sy * V + S)`` with (0, 0) the TOP-left corner of the image."""
    origin = np.asarray(origin, np.float64)
    fwd = _unit(np.asarray(look_at, np.float64) - origin)
    side = _unit(np.cross(fwd, np.asarray(up, np.float64)))
    upv = _unit(np.cross(side, fwd))
    view = np.eye(4)
    view[0, :3], view[1, :3], view[2, :3] = side, upv, -fwd
    view[:3, 3] = -view[:3, :3] @ origin
    sx = math.tan(math.radians(fov_deg) / 2.0) * NEAR
    sy = sx / aspect
    proj = np.zeros((4, 4))
    proj[0, 0] = 2.0 * NEAR / sx
    proj[1, 1] = 2.0 * NEAR / sy
    proj[2, 2] = -(FAR + NEAR) / (FAR - NEAR)
    proj[2, 3] = -(2.0 * FAR * NEAR) / (FAR - NEAR)
    proj[3, 2] = -1.0
    hx, hy = 1.0 / proj[0, 0], 1.0 / proj[1, 1]          # view-space half extents at unit distance
    S = fwd - hx * side + hy * upv
    U = 2.0 * hx * side
    V = -2.0 * hy * upv
    return view, proj, S, U, V


class DenseGridRenderer:
    def __init__(self, device="cuda", world_size=1.0):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DenseGridRenderer needs a CUDA device; there is no CPU fallback")
        self.world_size = float(world_size)
        self.volume, self.skip, self.use_skip = None, None, True
        self.time = 0.0
        # Args defaults, GPURendererDirect.cpp:100-128
        self.resolution = (512, 512)
        self.fov = 45.0
        self.origin, self.look_at, self.up = [0.0, 0.0, -1.0], [0.0, 0.0, 0.0], [0.0, 1.0, 0.0]
        self.viewport = None
        self.isovalue = 0.0
        self.diffuse, self.specular, self.ambient = [0.7, 0.7, 0.7], [1.0, 1.0, 1.0], [0.1, 0.1, 0.1]
        self.specular_exponent = 32
        self.camera_light, self.light_direction = True, [0.0, 0.0, 0.0]
        self.ao_samples, self.ao_radius, self.ao_bias = 32, 0.01, 1e-3
        self.step, self.binary_steps = 0.05, 10           # SetSteps(0.05, ...), :365; binarySearchSteps, render_kernel.cu:21
        self._last = None                                 # proj @ view of the previous frame (flow is computed against it)
        self.hemisphere, self.rotations = ops.raycast_ao_tables()

    # -- DirectRenderer interface ---------------------------------------------------------------------------------
    def load(self, volume):
        if isinstance(volume, str):
            volume = np.load(volume)
        v = torch.as_tensor(np.ascontiguousarray(volume, dtype=np.float32) if not isinstance(volume, torch.Tensor) else volume)
        if v.dim() != 3 or min(v.shape) < 2:
            raise ValueError("volume: nz x ny x nx with at least 2 voxels per axis expected")
        self.volume = v.to(self.device, torch.float32).contiguous()
        self.skip = ops.raycast_build_skip(self.volume) if self.use_skip else None

    def send_command(self, cmd, value):
        assert isinstance(cmd, str)
        assert isinstance(value, str)
        f = [float(t) for t in value.replace(";", ",").split(",") if t.strip() != ""]
        if cmd in ("fov", "cameraFoV"):
            self.fov = f[0]
        elif cmd == "cameraOrigin":
            self.origin = f[:3]
        elif cmd == "cameraLookAt":
            self.look_at = f[:3]
        elif cmd == "cameraUp":
            self.up = f[:3]
        elif cmd == "resolution":
            self.resolution = (int(f[0]), int(f[1]))
        elif cmd == "isovalue":
            self.isovalue = f[0]
        elif cmd == "unshaded":
            pass                                          # parsed and unused by the reference's iso kernel as well
        elif cmd == "aosamples":
            self.ao_samples = int(f[0])
        elif cmd == "aoradius":
            self.ao_radius = f[0]
        elif cmd == "viewport":
            self.viewport = [int(t) for t in f[:4]]
        else:
            raise ValueError(f"Unknown command: '{cmd}'")  # the reference returns -1 (GPURendererDirect.cpp:422-425)

    def params(self):
        """The ISR_RC_* block for the current state (previous camera = the last rendered one, or the current one)."""
        if self.volume is None:
            raise RuntimeError("no volume loaded")
        W, H = self.resolution
        nz, ny, nx = self.volume.shape
        voxel = self.world_size / max(nx, ny, nz)
        P = np.zeros(RC_PARAMS, np.float32)
        view, proj, S, U, V = camera_matrices(self.origin, self.look_at, self.up, self.fov, W / H)
        # the previous frame's full view-projection matrix: a change of cameraUp, fov or resolution between two frames
        # moves the projection as much as a change of origin / look-at does
        vp_cur = proj @ view
        vp_last = self._last if self._last is not None else vp_cur
        P[RC_CAM_POS:RC_CAM_POS + 3] = self.origin
        P[RC_CAM_U:RC_CAM_U + 3], P[RC_CAM_V:RC_CAM_V + 3], P[RC_CAM_S:RC_CAM_S + 3] = U, V, S
        P[RC_BOX_MIN:RC_BOX_MIN + 3] = [-0.5 * nx * voxel, -0.5 * ny * voxel, -0.5 * nz * voxel]
        P[RC_VOXEL], P[RC_ISO], P[RC_STEP] = voxel, self.isovalue, self.step
        P[RC_VIEW_CUR:RC_VIEW_CUR + 16] = vp_cur.reshape(-1)
        P[RC_VIEW_NEXT:RC_VIEW_NEXT + 16] = vp_last.reshape(-1)
        P[RC_NORMAL_MAT:RC_NORMAL_MAT + 16] = view.reshape(-1)
        light = (np.asarray(self.look_at, np.float64) - np.asarray(self.origin, np.float64)) if self.camera_light \
            else np.asarray(self.light_direction, np.float64)
        P[RC_LIGHT:RC_LIGHT + 3] = _unit(light)
        P[RC_AMBIENT:RC_AMBIENT + 3], P[RC_DIFFUSE:RC_DIFFUSE + 3] = self.ambient, self.diffuse
        P[RC_SPECULAR:RC_SPECULAR + 3] = self.specular
        P[RC_SPEC_EXP], P[RC_AO_RADIUS], P[RC_AO_BIAS] = self.specular_exponent, self.ao_radius, self.ao_bias
        return P

    def render_direct(self, tensor):
        W, H = self.resolution
        assert tuple(tensor.shape) == (H, W, 12), "target tensor must be resolutionY x resolutionX x 12"
        P = self.params()
        n_ao = max(0, min(512, int(self.ao_samples)))
        torch.cuda.synchronize(self.device)
        t0 = time.perf_counter()
        ops.raycast_iso(self.volume, P, tensor, ao_samples=n_ao, hemisphere=self.hemisphere, rotations=self.rotations,
                        binary_steps=self.binary_steps, viewport=self.viewport, skip=self.skip)
        torch.cuda.synchronize(self.device)
        self.time = time.perf_counter() - t0
        self._last = P[RC_VIEW_CUR:RC_VIEW_CUR + 16].reshape(4, 4).copy()
        return self.time

    def get_time(self):
        """Returns the time of the last render pass in seconds"""
        return self.time

    def close(self):
        pass  # No-op
