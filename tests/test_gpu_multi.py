"""-m gpu: several GPUs behind the C-ABI (rtw_create_multi).

On a one-GPU box the device is listed TWICE: two contexts, two worker threads, the peer-copy exchange
(NCCL refuses a communicator with a duplicate GPU).  With two or more GPUs the same tests run over
NCCL (ncclReduce for sample shards, ncclSend/ncclRecv for tiles).  Either way:
  - sample sharding: the G-context frame equals the one-context frame up to fp32 summation order;
  - tile sharding: bit-identical to the one-context frame at 1 spp, same tolerance above;
  - the frame at 1 spp is the oracle's, bit for bit."""
import os
import subprocess

import numpy as np
import pytest

# PyTorch first: rtw_create_multi dlopens NCCL, and on a box with two or more GPUs the system's libnccl.so.2 would then
# sit under the soname that torch's own, newer copy needs (ImportError: undefined symbol ncclDevCommCreate).
torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices(rtw, n):
    have = rtw.device_count()
    return list(range(n)) if have >= n else [0] * n


@pytest.mark.parametrize("n", [2, 3])
@pytest.mark.parametrize("sid,Nx,Ny", [(0, 96, 64), (4, 64, 48)])
def test_tiles_are_bit_identical_at_one_sample(director, rtw, orc, sid, Nx, Ny, n):
    m = rtw.MultiDirector(_devices(rtw, n))
    try:
        m.init(Nx, Ny, 1)
        m.set_option("shard", 2)
        m.createScene(sid)
        img = m.renderFrame(1)
        st = m.stats()
        assert st["shard"] == "tiles" and st["paths"] == Nx * Ny and len(st["per_device"]) == n
        o = orc.OracleScene(m.scene.desc, m.scene.desc.contents.n_prims)
        ref, c = o.render(Nx, Ny, 1, mode=0)
        assert np.array_equal(orc.tonemap(ref, 1).view(np.uint32), img.view(np.uint32))
        assert np.array_equal(m.linear().view(np.uint32), ref.view(np.uint32))
        assert st["rays_closest"] == c["rays_closest"] and st["rays_shadow"] == c["rays_shadow"]
    finally:
        m.destroy()


@pytest.mark.parametrize("shard", [1, 2])
def test_sharded_frame_equals_the_one_gpu_frame(director, rtw, orc, shard):
    Nx, Ny, spp = 80, 60, 9            # 9 samples over 2 contexts: 5 + 4
    director.init(Nx, Ny, spp)
    director.createScene(0)
    director.renderFrame(spp)
    one = director.linear()
    m = rtw.MultiDirector(_devices(rtw, 2))
    try:
        m.init(Nx, Ny, spp)
        m.set_option("shard", shard)
        m.createScene(0)
        m.renderFrame(spp)
        two = m.linear()
        st = m.stats()
    finally:
        m.destroy()
    assert st["shard"] == ("samples" if shard == 1 else "tiles")
    assert np.all(two[..., 3] == spp)
    assert np.allclose(one[..., :3], two[..., :3], rtol=1e-5, atol=1e-6)
    assert st["paths"] == Nx * Ny * spp and st["rays_closest"] == director.stats()["rays_closest"]
    if shard == 1:
        assert [p["paths"] for p in st["per_device"]] == [Nx * Ny * 5, Nx * Ny * 4]
    else:
        assert [p["paths"] for p in st["per_device"]] == [Nx * 30 * spp, Nx * 30 * spp]


def test_rows_entry_point_renders_only_its_rows(director, rtw, orc):
    """rtw_render_accum_rows: a tile leaves every other row of the accumulator untouched."""
    Nx, Ny = 64, 40
    director.init(Nx, Ny, 1)
    director.createScene(3)
    acc = torch.full((Ny, Nx, 4), -7.0, device="cuda:0")   # the context of the `director` fixture is on GPU 0
    L = rtw.lib()
    assert L.rtw_render_accum_rows(director._c, Nx, Ny, 10, 17, 0, 1, 1, 20, acc.data_ptr()) == rtw.RTW_OK
    got = acc.cpu().numpy()
    assert np.all(got[:10] == -7.0) and np.all(got[27:] == -7.0)
    o = orc.OracleScene(director.scene.desc, director.scene.desc.contents.n_prims)
    ref, _ = o.render(Nx, Ny, 1, mode=0)
    assert np.allclose(got[10:27] + 7.0, ref[10:27], rtol=1e-5, atol=1e-5)
    assert L.rtw_render_accum_rows(director._c, Nx, Ny, 30, 11, 0, 1, 1, 20, acc.data_ptr()) == rtw.RTW_ERR_INVALID   # rows past the image


def test_nccl_path_when_two_gpus_are_present(rtw, orc):
    if rtw.device_count() < 2:
        pytest.skip("one GPU: the NCCL exchange is covered by the 2-GPU run (profiles/r02) and bench.py --gpus N")
    m = rtw.MultiDirector([0, 1])
    try:
        assert m.uses_nccl()
        Nx, Ny, spp = 200, 120, 16
        m.init(Nx, Ny, spp)
        for shard in (1, 2):
            m.set_option("shard", shard)
            m.createScene(0)
            m.renderFrame(spp)
            lin = m.linear()
            o = orc.OracleScene(m.scene.desc, 13)
            ref, c = o.render(Nx, Ny, spp, mode=0)
            assert np.allclose(lin[..., :3], ref[..., :3] / spp, rtol=1e-5, atol=1e-6)
            assert m.stats()["rays_closest"] == c["rays_closest"]
    finally:
        m.destroy()


def test_cli_gpus_flag(rtw, tmp_path):
    if rtw.device_count() < 2:
        pytest.skip("one GPU")
    cli = os.path.join(ROOT, "raytracing-weekend_b200", "rtw_cli")
    a, b = tmp_path / "a.ppm", tmp_path / "b.ppm"
    subprocess.check_call([cli, "-s", "0", "-dx", "320", "-dy", "200", "-ns", "1", "-o", str(a)])
    subprocess.check_call([cli, "-s", "0", "-dx", "320", "-dy", "200", "-ns", "1", "-gpus", "2", "-o", str(b)])
    assert a.read_bytes() == b.read_bytes()      # 1 spp: tile mode, bit-identical
