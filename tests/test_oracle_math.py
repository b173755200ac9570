"""oracle/orc_math.h (fdlibm = Java StrictMath algorithms) against numpy/glibc: <= 1 ulp, exact special cases."""
import numpy as np


def _ulps(a, b):
    ia = a.view(np.int64).copy(); ib = b.view(np.int64).copy()
    ia[ia < 0] = np.iinfo(np.int64).min - ia[ia < 0]
    ib[ib < 0] = np.iinfo(np.int64).min - ib[ib < 0]
    return np.abs(ia - ib)


def test_atan2_within_1ulp(orc):
    r = np.random.default_rng(1)
    y = r.uniform(-1, 1, 400000) * 10.0 ** r.uniform(-6, 6, 400000)
    x = r.uniform(-1, 1, 400000) * 10.0 ** r.uniform(-6, 6, 400000)
    assert _ulps(orc.atan2(y, x), np.arctan2(y, x)).max() <= 1


def test_sincos_within_1ulp(orc):
    a = np.random.default_rng(2).uniform(-np.pi, np.pi, 400000)
    s, c = orc.sincos(a)
    assert _ulps(s, np.sin(a)).max() <= 1
    assert _ulps(c, np.cos(a)).max() <= 1
    a = np.random.default_rng(3).uniform(-1000, 1000, 100000)     # medium-range reduction
    s, c = orc.sincos(a)
    assert _ulps(s, np.sin(a)).max() <= 1
    assert _ulps(c, np.cos(a)).max() <= 1


def test_special_cases(orc):
    L = orc.lib()
    assert L.orc_atan2(0.0, 1.0) == 0.0 and np.signbit(L.orc_atan2(-0.0, 1.0))
    assert L.orc_atan2(0.0, 0.0) == 0.0                       # Math.atan2(0,0) = 0 (constant image => angle 0)
    assert L.orc_atan2(0.0, -1.0) == np.pi and L.orc_atan2(-0.0, -1.0) == -np.pi
    assert L.orc_atan2(1.0, 0.0) == np.pi / 2 and L.orc_atan2(-1.0, 0.0) == -np.pi / 2
    assert L.orc_atan2(3.0, 3.0) == np.pi / 4
    assert L.orc_cos(0.0) == 1.0 and L.orc_sin(0.0) == 0.0
    assert np.isnan(L.orc_atan2(np.nan, 1.0))
