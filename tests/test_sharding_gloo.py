"""Host logic of the N>1 path on CPU: two gloo ranks, each adds its sample shard
into its own accumulator (the per-rank renderer here is the ORACLE, this is a test), one reduce
onto rank 0; the result must be the single-process frame up to fp32 summation order."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NX, NY, SPP = 40, 40, 7          # odd spp: the shards differ in size (4 and 3 samples)


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rtw = importlib.import_module("raytracing-weekend_b200")
    sharding = importlib.import_module("raytracing-weekend_b200.sharding")
    from oracle import oracle_py as orc
    s = rtw.Scene(0, NX, NY)
    o = orc.OracleScene(s.desc, s.desc.contents.n_prims)
    accum = torch.zeros(NY, NX, 4)

    def render_shard(first, stride, count):
        img, _ = o.render(NX, NY, count, mode=0, sample_first=first, sample_stride=stride)
        accum.add_(torch.from_numpy(img))

    n = sharding.render_frame_sharded(render_shard, accum, SPP)
    assert n == len(range(rank, SPP, world))
    if rank == 0:
        np.save(out_path, accum.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sample_sharding_matches_single_render(tmp_path, rtw, orc):
    out = str(tmp_path / "frame.npy")
    port = 29600 + os.getpid() % 300
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    s = rtw.Scene(0, NX, NY)
    o = orc.OracleScene(s.desc, s.desc.contents.n_prims)
    ref, _ = o.render(NX, NY, SPP, mode=0)
    assert np.all(got[..., 3] == SPP)
    assert np.allclose(got[..., :3], ref[..., :3], rtol=1e-5, atol=1e-6)
    s.close()


def test_shard_samples_partition():
    sharding = importlib.import_module("raytracing-weekend_b200.sharding")
    for spp in (0, 1, 7, 1024, 10240):
        for world in (1, 2, 3, 8):
            ks = []
            for r in range(world):
                first, stride, count = sharding.shard_samples(spp, r, world)
                ks += [first + i * stride for i in range(count)]
            assert sorted(ks) == list(range(spp))
    with pytest.raises(ValueError):
        sharding.shard_samples(4, 2, 2)
