"""Pins the oracle's colour SURF (DescribePointSurfPlanar / DetectDescribeSurfPlanar / SurfPlanar_to_DetectDescribePoint) with
the reference's own tests:

  boofcv-feature/src/test/java/boofcv/alg/feature/describe/TestDescribePointSurfPlanar.java:47-92 (compareToSingleBand: every
      band block of the colour descriptor is the single-band descriptor up to one common scale; Laplacian sign from grey)
  boofcv-feature/src/test/java/boofcv/alg/feature/detdesc/TestDetectDescribeSurfPlanar.java:47-74 (same points as the detector
      run on the band average)
"""
import numpy as np

from synth import blob_image


def _planar(w, h, nb=3, seed=234):
    return np.stack([blob_image(w, h, seed + 17 * b) for b in range(nb)])


def test_average_is_sum_over_bands_divided(orc):
    rng = np.random.default_rng(5)
    b = (rng.random((3, 40, 50)) * 200).astype(np.float32)
    g = orc.planar_average(b)
    assert np.array_equal(g, ((b[0] + b[1]) + b[2]) / np.float32(3))
    u = rng.integers(0, 256, (3, 40, 50)).astype(np.uint8)
    assert np.array_equal(orc.planar_average(u), (u.astype(np.int32).sum(0) // 3).astype(np.uint8))
    u4 = rng.integers(0, 256, (4, 9, 7)).astype(np.uint8)
    assert np.array_equal(orc.planar_average(u4), (u4.astype(np.int32).sum(0) // 4).astype(np.uint8))


def test_compare_to_single_band(orc):
    # TestDescribePointSurfPlanar.compareToSingleBand — the inputs there are not integral images either
    rng = np.random.default_rng(234)
    w, h = 200, 150
    bands = (rng.random((3, h, w)) * 200).astype(np.float32)
    gray = orc.planar_average(bands)
    n = 100
    x = rng.random(n) * w; y = rng.random(n) * h; ang = rng.random(n) * 2 * np.pi; sc = rng.random(n) * 10 + 0.9
    raw = [orc.describe_raw(bands[b], x, y, ang, sc) for b in range(3)]
    found = np.concatenate(raw, axis=1)
    for i in range(n):
        v = found[i].copy()
        nrm = 0.0
        for t in v:
            nrm += t * t
        if nrm != 0:
            v = v / np.sqrt(nrm)
        for b in range(3):
            single, _ = orc.describe(bands[b], x[i:i + 1], y[i:i + 1], ang[i:i + 1], sc[i:i + 1])
            blk = v[64 * b:64 * (b + 1)]
            bn = np.sqrt((blk * blk).sum())
            if bn > 0:
                assert np.allclose(single[0], blk / bn, atol=1e-8)
        _, white = orc.describe(gray, x[i:i + 1], y[i:i + 1], ang[i:i + 1], sc[i:i + 1])
        assert orc.laplace_sign(gray, int(x[i] + 0.5), int(y[i] + 0.5), sc[i]) == bool(white[0])


def test_detects_the_points_of_the_band_average(orc):
    bands = _planar(320, 240)
    r = orc.surf_color_fast(bands)
    g = orc.surf_fast(orc.planar_average(bands))
    assert len(r["x"]) == len(g["x"]) > 50
    assert np.array_equal(r["x"], g["x"]) and np.array_equal(r["y"], g["y"]) and np.array_equal(r["scale"], g["scale"])
    assert np.array_equal(r["white"], g["white"])                  # Laplacian sign comes from the grey integral
    assert r["desc"].shape[1] == 192
    assert np.allclose(np.sqrt((r["desc"] ** 2).sum(1)), 1.0, atol=1e-12)
    # orientation uses objectRadius = scale (DetectDescribeSurfPlanar.java:108), not 2*scale: the angles differ from grey SURF
    assert not np.array_equal(r["angle"], g["angle"])
    gi = orc.integral(orc.planar_average(bands))
    assert np.array_equal(r["angle"], orc.orientation(gi, r["x"], r["y"], r["scale"] * 0.5))


def test_color_stable_and_u8(orc):
    bands = _planar(240, 200, nb=3, seed=9)
    r = orc.surf_color_stable(bands)
    assert r["desc"].shape[1] == 192 and len(r["x"]) > 30
    u = np.clip(bands, 0, 255).astype(np.uint8)
    ru = orc.surf_color_fast(u)
    gu = orc.surf_fast(orc.planar_average(u))
    assert np.array_equal(ru["x"], gu["x"]) and len(ru["x"]) > 30
