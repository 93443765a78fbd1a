"""the denoise STAND-IN (rtw_denoise, csrc/kernels/denoise.cu) — a documented, different filter in
the place of the reference's OptiX AI denoiser (Director.cpp:887-949,988-1000), excluded from parity with the reference.
Its own parity: the CUDA filter == oracle.cpp orc_denoise bit for bit; plus the properties any such filter must have."""
import numpy as np
import pytest


def test_oracle_filter_properties(orc):
    rng = np.random.RandomState(5)
    flat = np.full((24, 40, 4), 0.37, np.float32)
    flat[..., 3] = 1.0
    assert np.allclose(orc.denoise(flat, 4, 0.2), flat, rtol=0, atol=2e-7)   # a constant image is a fixed point (up to fp32 rounding of the weighted mean)
    img = rng.rand(33, 47, 4).astype(np.float32)
    img[..., 3] = 1.0
    out = orc.denoise(img, 3, 0.25)
    assert out.shape == img.shape and np.all(out[..., 3] == 1.0)
    assert out[..., :3].min() >= img[..., :3].min() - 1e-6 and out[..., :3].max() <= img[..., :3].max() + 1e-6   # convex combination
    assert out[..., :3].var() < img[..., :3].var()                   # it smooths
    edge = np.zeros((16, 32, 4), np.float32)
    edge[:, 16:, :3] = 1.0
    edge[..., 3] = 1.0
    sharp = orc.denoise(edge, 3, 0.05)
    assert abs(sharp[8, 15, 0]) < 0.01 and abs(sharp[8, 16, 0] - 1.0) < 0.01      # edge-stopping keeps a step edge


@pytest.mark.gpu
@pytest.mark.parametrize("shape,its,sigma", [((33, 47), 3, 0.25), ((1, 1), 1, 0.1), ((64, 5), 8, 0.5), ((200, 320), 4, 0.1)])
def test_cuda_filter_equals_the_oracle(director, orc, shape, its, sigma):
    rng = np.random.RandomState(shape[0] * 7 + its)
    img = rng.rand(shape[0], shape[1], 4).astype(np.float32)
    img[..., 3] = 1.0
    assert np.array_equal(director.denoise(img, its, sigma).view(np.uint32), orc.denoise(img, its, sigma).view(np.uint32))


@pytest.mark.gpu
def test_denoised_render_is_closer_to_the_converged_frame(director, orc):
    director.init(160, 120, 8)
    director.createScene(0)
    noisy = director.renderFrame(8).copy()
    ref = director.renderFrame(2048).copy()
    den = director.denoise(noisy, 3, 0.3)
    assert np.mean((den[..., :3] - ref[..., :3]) ** 2) < 0.6 * np.mean((noisy[..., :3] - ref[..., :3]) ** 2)
