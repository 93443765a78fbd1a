"""Known-answer vectors committed under tests/golden/ (made by tests/golden/make_golden.py): the oracle must
reproduce them on the CPU (`-m "not gpu"`), the CUDA path must reproduce the very same digests through the
C-ABI on the GPU (`-m gpu`).  Integer / bit-exact: SHA-256 of the raw arrays."""
import importlib
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "oracle_kat.json")))
IDS = ["scene%d" % c["scene"] for c in KAT["cases"]]


def _golden():
    import sys
    sys.path.insert(0, os.path.join(HERE, "golden"))
    return importlib.import_module("make_golden")


@pytest.mark.parametrize("case", KAT["cases"], ids=IDS)
def test_oracle_reproduces_the_committed_vectors(rtw, orc, case):
    import rayset
    g = _golden()
    now, _ = g.known_answers(rtw, orc, rayset, case["scene"], case["Nx"], case["Ny"])
    assert json.loads(json.dumps(now)) == case


@pytest.mark.gpu
@pytest.mark.parametrize("case", KAT["cases"], ids=IDS)
def test_cuda_path_reproduces_the_committed_vectors(director, rtw, orc, case):
    import rayset
    g = _golden()
    _, (scene, rays, shadow) = g.known_answers(rtw, orc, rayset, case["scene"], case["Nx"], case["Ny"])
    assert g.digest(rays) == case["rays_sha256"]
    director.init(case["Nx"], case["Ny"], 1)
    director.createScene(case["scene"])
    hits = director.trace_closest(rays)
    hits["t"][hits["prim"] < 0] = 0
    assert g.digest(hits) == case["hits_sha256"] and int((hits["prim"] >= 0).sum()) == case["n_hit"]
    occ = director.trace_any(shadow)
    assert g.digest(occ) == case["occluded_sha256"] and int(occ.sum()) == case["n_occluded"]
    director.renderFrame(1)
    img = director.linear()
    assert g.digest(img) == case["image_sha256"]
    st = director.stats()
    assert st["rays_closest"] == case["rays_closest"] and st["rays_shadow"] == case["rays_shadow"]
