"""Arithmetic contract: the path's transcendental functions are defined as fixed
fp32 operation sequences (oracle/detmath.h, csrc/device/detmath.h).
 - CPU: the oracle's statement is pinned against float64 libm (accuracy in ulp), and the
   product's header compiled for the host (tests/emu) gives the same bits.
 - GPU: the device gives the same bits as the oracle on every input.
"""
import numpy as np
import pytest

SIN, COS, LOG, POW5, ATAN2, ASIN = range(6)


def ulp_err(got, want64):
    got = got.astype(np.float64)
    w32 = want64.astype(np.float32)
    e = np.frexp(np.where(w32 == 0, np.float32(1e-45), w32))[1].astype(np.float64)
    ulp = np.ldexp(1.0, (e - 24).astype(np.int64))
    err = np.abs(got - want64) / ulp
    both_nan = np.isnan(got) & np.isnan(want64)
    same_inf = np.isinf(want64) & (got == want64)
    return np.where(both_nan | same_inf, 0.0, err)


def inputs():
    """Argument sets the path actually produces, plus edge cases."""
    rng = np.random.default_rng(12345)
    lattice = (np.arange(0, 1 << 24, 61, dtype=np.uint32).astype(np.float32) / np.float32(16777216.0))
    phi = (np.float32(2.0) * np.float32(3.141592654)) * lattice          # randomCosineDirection, lib/sampling.cuh:52
    wide = rng.uniform(-2.0e4, 2.0e4, 400000).astype(np.float32)          # noise / checker arguments
    huge = np.array([1.0e5, 1.0000001e5, 3.0e6, 1.0e9, -7.5e12, 3.0e38, np.inf, -np.inf, np.nan, 0.0, -0.0], np.float32)
    u_rnd = lattice[1:]                                                   # rnd(): k / 2^24
    u_randf = (rng.integers(1, 1 << 32, 400000, dtype=np.uint64).astype(np.float32) / np.float32(4294967296.0))
    log_edge = np.array([0.0, -0.0, -1.0, np.inf, np.nan, 1.0, 1e-40, 1.1754944e-38, 3.0e38, 0.70710677, 0.7071068], np.float32)
    unit = rng.uniform(-1, 1, 400000).astype(np.float32)
    unit2 = rng.uniform(-1, 1, 400000).astype(np.float32)
    asin_edge = np.array([1.0, -1.0, 0.5, -0.5, 0.50000006, 1.0000001, -1.0000001, 0.0, -0.0, np.nan, 1e-20], np.float32)
    at_y = np.array([0, 0, -0.0, 1, -1, np.inf, np.inf, -np.inf, 1, np.nan, 1e-30, 0.0, -0.0], np.float32)
    at_x = np.array([0, -1, -1, 0, 0, np.inf, -np.inf, np.inf, np.inf, 1, -1e30, -0.0, -0.0], np.float32)
    return {
        SIN: (np.concatenate([phi, wide, huge]), None),
        COS: (np.concatenate([phi, wide, huge]), None),
        LOG: (np.concatenate([u_rnd, u_randf, log_edge]), None),
        POW5: (np.concatenate([unit, np.array([0, -0.0, 1, -1e-8, 2, np.nan, np.inf], np.float32)]), None),
        ATAN2: (np.concatenate([unit, at_y]), np.concatenate([unit2, at_x])),
        ASIN: (np.concatenate([unit, asin_edge]), None),
    }


def truth(fn, x, y):
    x64 = x.astype(np.float64)
    with np.errstate(all="ignore"):
        if fn == SIN:
            return np.sin(x64)
        if fn == COS:
            return np.cos(x64)
        if fn == LOG:
            return np.log(x64)
        if fn == POW5:
            return x64 ** 5
        if fn == ATAN2:
            return np.arctan2(x64, y.astype(np.float64))
        return np.arcsin(x64)


# stated accuracy of the contract (CUDA libm documents sinf/cosf 2, logf 1, powf 4, atan2f 3, asinf 2 ulp: same class)
MAX_ULP = {SIN: 2.0, COS: 2.0, LOG: 1.0, POW5: 4.0, ATAN2: 4.0, ASIN: 3.0}


@pytest.mark.parametrize("fn", [SIN, COS, LOG, POW5, ATAN2, ASIN])
def test_oracle_detmath_against_float64(orc, fn):
    x, y = inputs()[fn]
    got = orc.math_eval(fn, x, y)
    want = truth(fn, x, y)
    fin = np.isfinite(x) & (np.abs(x) <= 1e13)
    if fn == LOG:
        fin &= ~((x > 0) & (x < 1.2e-38))      # subnormal inputs: result checked to 1e-6 relative below
    if fn == POW5:
        fin &= np.abs(x) > 1e-6                  # below that x^5 is subnormal in fp32
    with np.errstate(all="ignore"):
        err = ulp_err(got, want)
    assert np.nanmax(err[fin]) <= MAX_ULP[fn], (fn, float(np.nanmax(err[fin])), x[fin][np.nanargmax(err[fin])])
    # non-finite conventions follow libm
    assert np.array_equal(np.isnan(got), np.isnan(want.astype(np.float32)))
    inf = np.isinf(want)
    assert np.array_equal(got[inf].astype(np.float64), want[inf])


def test_oracle_detmath_special_values(orc):
    f = np.float32
    assert orc.math_eval(LOG, [1e-40])[0] == pytest.approx(np.log(1e-40), rel=1e-6)
    assert orc.math_eval(LOG, [1.0])[0] == 0.0
    a = orc.math_eval(ATAN2, [0.0, 0.0, -0.0, 1.0, -1.0], [0.0, -1.0, -1.0, 0.0, 0.0])
    assert a[0] == 0.0 and a[1] == f(np.pi) and a[2] == -f(np.pi) and a[3] == f(np.pi / 2) and a[4] == -f(np.pi / 2)
    s = orc.math_eval(ASIN, [1.0, -1.0, 1.0000001])
    assert s[0] == f(np.pi / 2) and s[1] == -f(np.pi / 2) and np.isnan(s[2])
    # sin/cos stay in [-1, 1] for every finite argument, NaN otherwise
    x = np.array([1e9, -7.5e12, 3e38, 1e5, -1e5], f)
    assert np.all(np.abs(orc.math_eval(SIN, x)) <= 1.0) and np.all(np.abs(orc.math_eval(COS, x)) <= 1.0)
    assert np.isnan(orc.math_eval(SIN, [np.inf, np.nan])).all()


@pytest.mark.parametrize("fn", [SIN, COS, LOG, POW5, ATAN2, ASIN])
def test_product_header_on_host_matches_oracle_bits(orc, emu, fn):
    """csrc/device/detmath.h (compiled by g++ in the test harness) == oracle/detmath.h."""
    x, y = inputs()[fn]
    a = orc.math_eval(fn, x, y)
    b = emu.math_eval(fn, x, y)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("fn", [SIN, COS, LOG, POW5, ATAN2, ASIN])
def test_device_detmath_matches_oracle_bits(director, orc, fn):
    x, y = inputs()[fn]
    a = orc.math_eval(fn, x, y)
    b = director.math_eval(fn, x, y)
    nan = np.isnan(a)
    assert np.array_equal(nan, np.isnan(b))
    assert np.array_equal(a.view(np.uint32)[~nan], b.view(np.uint32)[~nan])
