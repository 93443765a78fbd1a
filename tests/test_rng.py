"""RNG streams: the oracle against the reference's own lib/random.cuh (golden KATs
produced by oracle/make_rng_kat.sh from the reference source) and against the published
SDK lcg recurrence; the product's device bodies (host-compiled) against the oracle."""
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(__file__), "golden", "rng_kat.json")


def test_oracle_tea_matches_reference_kat(orc):
    kat = json.load(open(GOLD))
    for row in kat["tea"]:
        assert orc.tea(64, row["s0"], row["s1"]) == row["tea64"]
        assert orc.tea(16, row["s0"], row["s1"]) == row["tea16"]
        assert orc.tea(4, row["s0"], row["s1"]) == row["tea4"]
    # value measured in the survey session
    assert orc.tea(64, 0, 0) == 1132204115


def _xs_stream(seed, n):
    out, bits = [], []
    s = seed
    for _ in range(n):
        s ^= (s << 13) & 0xffffffff
        s ^= s >> 17
        s ^= (s << 5) & 0xffffffff
        out.append(s)
    return out


def test_oracle_randf_matches_reference_kat(orc):
    kat = json.load(open(GOLD))
    # the oracle's stream hook starts from tea<64>(seed0, seed1); check the generators
    # through pure-python xorshift (pinned by the KAT) and the oracle's float conversion
    for row in kat["randf_streams"]:
        assert _xs_stream(row["seed"], 16) == row["xorshift32"]
        f = (np.array(row["xorshift32"], dtype=np.uint32).astype(np.float32) / np.float32(4294967296.0))
        assert f.view(np.uint32).tolist() == row["randf_bits"]
    for row in kat["randf_one_branch"]:
        # (float)0x3F7FFFFF, not 0.99999994 (lib/random.cuh:31-38)
        assert row["randf_bits"] == int(np.float32(0x3F7FFFFF).view(np.uint32))


def test_oracle_streams_follow_the_recurrences(orc):
    for s0, s1 in [(0, 0), (359999, 3), (12345, 678)]:
        t, lcg, rnd, xs, rf = orc.rng_stream(s0, s1, 64)
        assert t == orc.tea(64, s0, s1)
        st = t
        for i in range(64):
            st = (1664525 * st + 1013904223) & 0xffffffff
            assert lcg[i] == st & 0x00FFFFFF
            assert rnd[i] == np.float32(st & 0x00FFFFFF) / np.float32(0x01000000)
        assert xs.tolist() == _xs_stream(t, 64)
        expect = xs.astype(np.float32) / np.float32(4294967296.0)
        expect[expect == 1.0] = np.float32(0x3F7FFFFF)
        assert np.array_equal(rf, expect)


def test_randf_stream_from_survey(orc):
    # randf stream from seed 0x6314759 measured on the reference source
    kat = json.load(open(GOLD))
    row = [r for r in kat["randf_streams"] if r["seed"] == 0x6314759][0]
    f = np.array(row["randf_bits"][:3], dtype=np.uint32).view(np.float32)
    assert np.allclose(f, [0.959296644, 0.0746408552, 0.638831079], rtol=0, atol=1e-9)


def test_product_device_rng_bodies_match_oracle(orc, emu):
    for s0, s1 in [(0, 0), (1, 0), (359999, 1023), (0xffffffff, 7)]:
        a = orc.rng_stream(s0, s1, 256)
        b = emu.rng_stream(s0, s1, 256)
        assert a[0] == b[0]
        for x, y in zip(a[1:], b[1:]):
            assert np.array_equal(x.view(np.uint32), y.view(np.uint32))
