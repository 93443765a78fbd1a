"""Tier 2 parity of the wavefront renderer: per-path agreement at 1 spp, same-sample
images at higher spp, full-size C2 properties, sample sharding."""
import numpy as np
import pytest

from rtw_types import SCENE_CASES

pytestmark = pytest.mark.gpu

# Tolerances (stated).  The device and glibc libm differ by <= 2 ulp in
# sinf/cosf/logf/powf/atan2f/asinf, so a path can differ in the last bits and — rarely —
# take a different branch.  Per path: relative 2e-4 on >= 99.5 % of the pixels at 1 spp.
# Image: relMSE = mean((I-R)^2 / (R^2 + 1e-2)) <= 1e-3 (target of BASELINE.json) — measured
# values are orders of magnitude below because both sides draw the SAME samples.
PATH_RTOL = 2e-4
PATH_FRACTION = 0.995
RELMSE_MAX = 1e-3


def relmse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


@pytest.mark.parametrize("sid,Nx,Ny", SCENE_CASES)
def test_one_spp_per_path_agreement(director, orc, sid, Nx, Ny):
    director.init(Nx, Ny, 1)
    director.createScene(sid)
    o = orc.OracleScene(director.scene.desc, director.scene.desc.contents.n_prims)
    ref, c = o.render(Nx, Ny, 1, mode=1)
    out = director.renderFrame(1)
    lin = director.linear()
    a, b = ref[..., :3], lin[..., :3]
    tol = PATH_RTOL * np.maximum(np.abs(a).max(axis=2), 1.0)
    frac = (np.abs(a - b).max(axis=2) <= tol).mean()
    assert frac >= PATH_FRACTION, frac
    assert np.all(lin[..., 3] == 1.0) and np.all(out[..., 3] == 1.0)
    st = director.stats()
    assert st["paths"] == Nx * Ny
    # ray counts: identical up to the rare diverged paths
    assert abs(st["rays_closest"] - c["rays_closest"]) <= 0.01 * c["rays_closest"] + 16
    assert abs(st["rays_shadow"] - c["rays_shadow"]) <= 0.01 * c["rays_shadow"] + 16
    assert st["kernel_launches"] > 0
    # gamma: out = sqrt(linear)
    assert np.allclose(out[..., :3], np.sqrt(np.maximum(lin[..., :3], 0)), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("sid,Nx,Ny,spp", [(0, 96, 96, 64), (1, 96, 54, 32), (4, 64, 64, 32)])
def test_same_sample_images(director, orc, sid, Nx, Ny, spp):
    director.init(Nx, Ny, spp)
    director.createScene(sid)
    o = orc.OracleScene(director.scene.desc, director.scene.desc.contents.n_prims)
    ref, _ = o.render(Nx, Ny, spp, mode=1)
    director.renderFrame(spp)
    lin = director.linear()
    assert np.all(lin[..., 3] == spp)
    r = ref[..., :3] / spp
    e = relmse(lin[..., :3], r)
    assert e <= RELMSE_MAX, e
    assert e <= 1e-4, e      # same samples on both sides: far below the target
    assert abs(lin[..., :3].mean() - r.mean()) <= 2e-3 * r.mean()


def test_pool_size_and_lag_do_not_change_the_image(director, orc):
    Nx = Ny = 64
    director.init(Nx, Ny, 8)
    director.createScene(0)
    imgs = []
    for pool, lag in ((1024, 1), (4096, 3), (1 << 16, 7)):
        director.set_option("pool_paths", pool)
        director.set_option("lag", lag)
        director.renderFrame(8)
        imgs.append(director.linear())
    director.set_option("pool_paths", 1 << 21)
    director.set_option("lag", 3)
    for im in imgs[1:]:
        assert np.allclose(imgs[0], im, rtol=1e-5, atol=1e-6)   # fp32 atomic order only


def test_sample_sharding_matches_single_render(director, rtw):
    """Rank g of G renders samples k = g (mod G): two 'ranks' on one GPU."""
    torch = pytest.importorskip("torch")
    Nx = Ny = 64
    director.init(Nx, Ny, 8)
    director.createScene(0)
    director.renderFrame(8)
    whole = director.linear()
    acc = torch.zeros(Ny, Nx, 4, device="cuda:0")
    torch.cuda.synchronize()
    for g in range(2):
        director.render_accum(acc.data_ptr(), g, 2, 4)
    out = torch.empty_like(acc)
    director.tonemap(acc.data_ptr(), out.data_ptr(), 8)
    torch.cuda.synchronize()
    a = acc.cpu().numpy()
    assert np.all(a[..., 3] == 8)
    assert np.allclose(a[..., :3] / 8, whole[..., :3], rtol=1e-5, atol=1e-6)
    assert np.allclose(out.cpu().numpy()[..., :3], np.sqrt(a[..., :3] / 8), rtol=1e-6, atol=1e-7)


def test_c2_full_size_properties(director, orc):
    """C2 (Cornell 600x600) at 64 spp — size-independent checks at the benchmark size."""
    Nx = Ny = 600
    spp = 64
    director.init(Nx, Ny, spp)
    director.createScene(0)
    out = director.renderFrame(spp)
    lin = director.linear()
    st = director.stats()
    assert st["paths"] == Nx * Ny * spp
    assert np.all(lin[..., 3] == spp) and np.isfinite(lin).all() and (lin[..., :3] >= 0).all()
    # the oracle at 4 spp estimates the same image mean (same estimator)
    o = orc.OracleScene(director.scene.desc, 13)
    ref, c = o.render(Nx, Ny, 4, mode=0)
    m_ref = ref[..., :3].mean() / 4
    assert abs(lin[..., :3].mean() - m_ref) <= 0.02 * m_ref
    # rays per path agree with the oracle's counters
    assert abs(st["rays_closest"] / st["paths"] - c["rays_closest"] / c["paths"]) < 0.02
    assert abs(st["rays_shadow"] / st["paths"] - c["rays_shadow"] / c["paths"]) < 0.02
    # left wall is green-ish, right wall red-ish (scene sanity)
    assert out[300, 590, 0] > out[300, 590, 1] and out[300, 10, 1] > out[300, 10, 0]


def test_profile_option_reports_stage_times(director):
    director.init(64, 64, 2)
    director.createScene(0)
    director.set_option("profile", 1)
    director.renderFrame(2)
    st = director.stats()
    director.set_option("profile", 0)
    assert st["traverse_ms"] > 0 and st["shade_ms"] > 0 and st["logic_ms"] > 0
