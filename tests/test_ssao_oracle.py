"""CPU: the restatement of ShadowFX::screenSpaceAmbientOcclusion (oracle/ssao_oracle.cpp) against the reference's own
ShadowFX.cpp compiled in place (oracle/_ref, built where /root/reference is present), and the random tables of the
product library against the oracle's."""
import ctypes

import numpy as np
import pytest

from oracle import oracle as orc


def ssao_frame(h=48, w=64, seed=5):
    """A 12-channel interleaved frame (GPURendererDirect wire format) with a sphere-ish depth field and unit normals."""
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    f = rng.rand(h, w, 12).astype(np.float32)
    r2 = ((xx - w / 2) / (w / 3)) ** 2 + ((yy - h / 2) / (h / 3)) ** 2
    inside = r2 < 1
    depth = np.where(inside, 0.5 - 0.2 * np.sqrt(np.clip(1 - r2, 0, 1)) + 0.01 * rng.rand(h, w), 0).astype(np.float32)
    depth[0:4, 0:4] = 1.0                                           # depth == 1 is skipped like depth == 0
    n = np.stack([(xx - w / 2) / (w / 3), (yy - h / 2) / (h / 3), np.sqrt(np.clip(1 - r2, 0.01, 1))], -1)
    f[..., 4:7] = (n * (1 + 0.3 * rng.rand(h, w, 1))).astype(np.float32)      # not normalised: the function normalises
    f[..., 7] = depth
    return f


@pytest.mark.skipif(not orc.ssao_ref_available(), reason="oracle/_ref is built only where /root/reference is present")
@pytest.mark.parametrize("cfg", [dict(), dict(samples=32, sample_radius=0.05), dict(noise_samples=2, bias=0.0)])
def test_restatement_equals_reference_build(cfg):
    f = ssao_frame()
    a = orc.ssao(f, use_reference_build=True, **cfg)
    b = orc.ssao(f, **cfg)
    assert np.array_equal(a, b)                                     # every channel, bit for bit
    ao = b[..., 10]
    assert ao.min() >= 0 and ao.max() <= 1 and np.all(ao[f[..., 7] == 0] == 0) and 0.2 < ao[f[..., 7] > 0].mean() < 1


def test_library_tables_equal_oracle_tables():
    """isr_ssao_tables runs on the host (no GPU needed): same std::default_random_engine(42) sequence as the oracle."""
    from isosurfacesuperresolution_b200 import _lib
    for samples, ns in ((64, 4), (32, 2), (128, 8)):
        k, n = np.zeros(3 * samples, np.float32), np.zeros(3 * ns * ns, np.float32)
        _lib.check(_lib.lib().isr_ssao_tables(samples, ns, k.ctypes.data_as(ctypes.c_void_p), n.ctypes.data_as(ctypes.c_void_p)))
        ok, on = orc.ssao_tables(samples, ns)
        assert np.array_equal(k, ok) and np.array_equal(n, on)
        assert np.all(k.reshape(-1, 3)[:, 1] >= 0) and np.all(n.reshape(-1, 3)[:, 1] == 0)      # hemisphere, y-free noise
    with pytest.raises(_lib.IsrError):
        _lib.check(_lib.lib().isr_ssao_tables(129, 4, None, None))
