"""rtw_denoise_static (include/rtw.h): the allocation-free twin of rtw_denoise — no counterpart in the reference, added
for callers that cannot call cudaMalloc.  Its ping-pong scratch is a file-scope
`static __device__` array inside librtw.so (csrc/kernels/denoise.cu), shared by every call on a device.

Collected LAST on purpose (a failure here must not hide any other test); passes alone
in a fresh process and after the rest of the suite.  The check is against the EXISTING entry point, bit for bit (the
tolerance tests/test_denoise.py uses for this filter), on two calls in a row — the second call starts from whatever the
first left in the shared array."""
import numpy as np
import pytest


def _image(shape, seed):
    rng = np.random.RandomState(seed)
    img = rng.rand(shape[0], shape[1], 4).astype(np.float32)
    img[..., 3] = 1.0
    return img


def test_the_header_declares_it_and_the_library_exports_it(rtw):
    import os
    text = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "rtw.h")).read()
    assert "rtw_denoise_static" in text and "RTW_DENOISE_STATIC_MAX_PIXELS (1 << 19)" in text
    assert "rtw_denoise_static" in rtw.ABI_SYMBOLS and hasattr(rtw.lib(), "rtw_denoise_static")
    assert rtw.DENOISE_STATIC_MAX_PIXELS == 1 << 19


@pytest.mark.gpu
def test_static_scratch_filter_equals_rtw_denoise_on_two_calls_in_a_row(rtw, director):
    import torch
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream(dev).cuda_stream
    # two different frames, sizes and pass counts (odd and even: the first pass writes the output or the scratch),
    # the second SMALLER than the first so that stale pixels of call 1 lie beyond and inside its part of the array
    calls = [((200, 320), 4, 0.1, 11), ((33, 47), 3, 0.25, 12), ((600, 600), 5, 0.3, 13), ((1, 1), 1, 0.1, 14)]
    for shape, its, sigma, seed in calls:
        img = _image(shape, seed)
        want = director.denoise(img, its, sigma)                       # the existing entry point (host buffers, cudaMalloc inside)
        src = torch.from_numpy(img).to(dev)
        dst = torch.full_like(src, float("nan"))
        torch.cuda.synchronize(dev)
        director.denoise_static(src.data_ptr(), dst.data_ptr(), shape[1], shape[0], its, sigma, stream)
        torch.cuda.synchronize(dev)
        got = dst.cpu().numpy()
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (shape, its)
        assert np.array_equal(src.cpu().numpy().view(np.uint32), img.view(np.uint32))      # the input is read-only
    # the same call twice gives the same bits (nothing in the shared array has to start at zero)
    img = _image((120, 160), 21)
    src = torch.from_numpy(img).to(dev)
    torch.cuda.synchronize(dev)
    outs = []
    for _ in range(2):
        dst = torch.empty_like(src)
        director.denoise_static(src.data_ptr(), dst.data_ptr(), 160, 120, 3, 0.2, None)    # NULL = the context's stream
        torch.cuda.synchronize(dev)
        outs.append(dst.cpu().numpy())
    assert np.array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))
    assert np.array_equal(outs[0].view(np.uint32), director.denoise(img, 3, 0.2).view(np.uint32))


@pytest.mark.gpu
def test_static_scratch_filter_refuses_what_it_cannot_do(rtw, director):
    import torch
    dev = torch.device("cuda:0")
    big = torch.zeros((1025, 512, 4), dtype=torch.float32, device=dev)         # one row more than 2^19 pixels
    out = torch.empty_like(big)
    with pytest.raises(rtw.RtwError, match="RTW_DENOISE_STATIC_MAX_PIXELS"):
        director.denoise_static(big.data_ptr(), out.data_ptr(), 512, 1025, 2, 0.2)
    director.denoise_static(big.data_ptr(), out.data_ptr(), 512, 1024, 2, 0.2)   # exactly the maximum is fine
    torch.cuda.synchronize(dev)
    assert float(out[:1024].abs().max()) == 0.0
    with pytest.raises(rtw.RtwError, match="overlap"):
        director.denoise_static(big.data_ptr(), big.data_ptr(), 64, 64, 2, 0.2)
    with pytest.raises(rtw.RtwError, match="bad denoise arguments"):
        director.denoise_static(big.data_ptr(), out.data_ptr(), 64, 64, 9, 0.2)
    with pytest.raises(rtw.RtwError, match="bad denoise arguments"):
        director.denoise_static(0, out.data_ptr(), 64, 64, 2, 0.2)
