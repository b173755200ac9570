"""Pins oracle NMS + Fast-Hessian detector with the reference's tests, re-expressed.

  boofcv-feature/src/test/java/boofcv/alg/feature/detect/extract/GenericNonMaxTests.java:94-124 (border influence),
      :152-174 (strict rule), :200-245 (same point set as NonMaxExtractorNaive, Gaussian noise sigma 3, thr 0.6)
  boofcv-feature/src/test/java/boofcv/alg/feature/detect/interest/GenericFeatureDetectorTests.java:52-124 (properties)
  boofcv-feature/src/main/java/boofcv/alg/feature/detect/interest/FastHessianFeatureDetector.java:336-350 (polyPeak)
"""
import numpy as np
import pytest

from synth import blob_image


def _set(xy):
    return set(map(tuple, xy.tolist()))


def test_strict_rule_equal_values_knock_each_other_out(orc):
    img = np.zeros((20, 30), np.float32)           # GenericNonMaxTests.testStrictRule: 30 wide, 20 tall
    for x, y in ((3, 5), (5, 7), (7, 7)):
        img[y, x] = 30
    assert len(orc.nonmax(img, radius=2, thresh=5, border=0)) == 0


def test_border_pixels_still_suppress(orc):
    img = np.zeros((20, 30), np.float32)
    img[1, 0] = 90
    img[1, 1] = 30
    assert _set(orc.nonmax(img, 1, 5, 0)) == {(0, 1)}
    assert len(orc.nonmax(img, 1, 5, 1)) == 0       # (1,1) is inside, but (0,1) in the ignored border beats it


@pytest.mark.parametrize("radius", [1, 2, 3, 4])
@pytest.mark.parametrize("border", [0, 1, 3])
def test_block_nms_equals_naive(orc, radius, border):
    r = np.random.default_rng(radius * 10 + border)
    img = (r.standard_normal((61, 73)) * 3).astype(np.float32)
    big = np.zeros((80, 90), np.float32); big[5:66, 4:77] = img
    for view in (img, big[5:66, 4:77]):             # also through a sub-image
        a = orc.nonmax(view, radius, 0.6, border)
        b = orc.nonmax(view, radius, 0.6, border, naive=True)
        assert len(a) == len(_set(a)) and _set(a) == _set(b) and len(a) > 0


def test_block_nms_order_is_block_raster(orc):
    img = (np.random.default_rng(3).standard_normal((50, 64)) * 3).astype(np.float32)
    xy = orc.nonmax(img, 2, 0.6, 4).astype(int)
    key = ((xy[:, 1] - 4) // 3) * 1000 + (xy[:, 0] - 4) // 3
    assert np.all(np.diff(key) > 0)                 # at most one per block, block-row-major (NonMaxBlock.java:84-93)


def test_nms_ignores_flt_max_and_nan(orc):
    img = np.zeros((12, 12), np.float32)
    img[5, 5] = np.finfo(np.float32).max
    assert len(orc.nonmax(img, 1, 0.5, 0)) == 0     # peakVal != Float.MAX_VALUE (NonMaxBlockSearchStrict.java:75)
    img[5, 5] = np.nan; img[8, 8] = 2
    assert _set(orc.nonmax(img, 1, 0.5, 0)) == {(8, 8)}


def test_poly_peak(orc):
    assert orc.poly_peak(1, 2, 1) == 0.0
    assert orc.poly_peak(2, 2, 2) == 0.0            # a == 0 guard
    assert orc.poly_peak(1, 3, 2) == pytest.approx(1.0 / 6.0, rel=1e-6)
    lo, mid, hi = np.float32(0.3), np.float32(1.7), np.float32(1.1)
    a = np.float32(np.float32(np.float32(0.5) * lo - mid) + np.float32(0.5) * hi)
    b = np.float32(np.float32(0.5) * hi - np.float32(0.5) * lo)
    assert orc.poly_peak(lo, mid, hi) == np.float32(-b / np.float32(np.float32(2) * a))


def test_detector_properties(orc):
    """GenericFeatureDetectorTests: finds points, all inside the image, same twice, scales in range."""
    img = blob_image(320, 240, seed=11)
    ii = orc.integral(img)
    a = orc.fh_detect(ii); b = orc.fh_detect(ii)
    n = len(a["x"])
    assert n > 20
    assert all(np.array_equal(a[k], b[k]) for k in a)
    assert (a["x"] > 0).all() and (a["x"] < 320).all() and (a["y"] > 0).all() and (a["y"] < 240).all()
    assert (a["scale"] >= 1.2 * 9 / 9 - 1e-9).all() and (a["scale"] <= 1.2 * 195 / 9).all()
    # emission order: octave, then level, then block raster
    assert np.all(np.diff(a["octave"] * 10 + a["level"]) >= 0)
    # every emitted point is a strict scale-space maximum above threshold
    assert (a["response"] >= 1.0).all()


def test_detector_threads_equal_single(orc):
    ii = orc.integral(blob_image(300, 200, seed=5))
    a = orc.fh_detect(ii, nthreads=1); b = orc.fh_detect(ii, nthreads=4)
    assert all(np.array_equal(a[k], b[k]) for k in a)


def test_detector_matches_stage_composition(orc):
    """detect == hessian x3 -> nonmax(border=size/(2 skip)) -> border reject -> 3x3 checks -> polyPeak, first octave level 1."""
    img = blob_image(200, 160, seed=7)
    ii = orc.integral(img)
    det = orc.fh_detect(ii)
    sel = (det["octave"] == 0) & (det["level"] == 1)
    l0, l1, l2 = (orc.hessian(ii, 1, s) for s in (9, 15, 21))
    cand = orc.nonmax(l1, 2, 1.0, 15 // 2)
    ig = 15 // 2 + 2
    out = []
    for x, y in cand.astype(int):
        if x < ig or x >= 200 - ig or y < ig or y >= 160 - ig:
            continue
        v = l1[y, x]
        if (l0[y - 1:y + 2, x - 1:x + 2] >= v).any() or (l2[y - 1:y + 2, x - 1:x + 2] >= v).any():
            continue
        px = orc.poly_peak(l1[y, x - 1], v, l1[y, x + 1]); py = orc.poly_peak(l1[y - 1, x], v, l1[y + 1, x])
        ps = orc.poly_peak(l0[y, x], v, l2[y, x])
        ix = np.float32(np.float32(np.float32(x) + np.float32(px)) * np.float32(1))
        iy = np.float32(np.float32(np.float32(y) + np.float32(py)) * np.float32(1))
        isz = np.float32(np.float32(15) + np.float32(np.float32(ps) * np.float32(6)))
        out.append((float(ix), float(iy), 1.2 * float(isz) / 9.0))
    got = list(zip(det["x"][sel], det["y"][sel], det["scale"][sel]))
    assert len(out) > 3 and got == out


def test_detector_stops_when_kernel_exceeds_image(orc):
    """FastHessianFeatureDetector.java:177-179: an octave whose largest size exceeds the image is skipped (and all after)."""
    ii = orc.integral(blob_image(60, 60, seed=2))
    det = orc.fh_detect(ii)
    assert set(det["octave"].tolist()) <= {0, 1}     # sizes 27 and 51 fit in 60, 99 does not


def test_max_features_per_scale(orc):
    ii = orc.integral(blob_image(320, 240, seed=11))
    full = orc.fh_detect(ii)
    lim = orc.fh_detect(ii, dict(max_features_per_scale=5))
    for o in range(4):
        for l in (1, 2):
            assert ((lim["octave"] == o) & (lim["level"] == l)).sum() <= 5
    assert _set(np.stack([lim["x"], lim["y"]], 1)) <= _set(np.stack([full["x"], full["y"]], 1))
