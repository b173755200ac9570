"""Pins oracle weights / sparse gradient / orientation / descriptor with the reference's tests, re-expressed.

  boofcv-feature/src/test/java/boofcv/alg/feature/describe/BaseTestDescribeSurf.java:62-209 (sub-image, scale/rotation
      change, border no-throw, constant image -> 0, x-ramp -> v[i]==v[i+1]>0 and v[i+2]=v[i+3]=0, rotated by pi/2, fraction)
  boofcv-feature/src/test/java/boofcv/alg/feature/orientation/GenericOrientationIntegralTests.java:76-199
      (ramp image at angle theta -> theta recovered; tolerance 0.01 for average, pi/9 for sliding window)
  boofcv-ip/src/test/java/boofcv/alg/transform/ii/TestIntegralImageOps.java:90-108 (isInBounds)
  boofcv-feature/src/test/java/boofcv/abst/feature/detdesc/GenericTestsDetectDescribePoint.java:83-211 (interface checks)
"""
import numpy as np
import pytest

from synth import blob_image

W, H = 50, 60


def _ramp(theta, w=W, h=H, mul=1.0):
    """TestImplSurfDescribeOps.createGradient: (float)(cos*x + sin*y)."""
    c, s = np.cos(theta), np.sin(theta)
    y, x = np.mgrid[0:h, 0:w]
    return (mul * (c * x + s * y)).astype(np.float32)


def _noise_ii(seed=234, w=W, h=H):
    return np.random.default_rng(seed).uniform(0, 100, (h, w)).astype(np.float32)


def test_gaussian_tables(orc):
    k = orc.gaussian2d(-1, 6)                        # orientation weights: sigma = (2*6+1)/5 = 2.6
    assert k.shape == (13, 13) and abs(k.sum() - 1) < 1e-12 and np.allclose(k, k.T) and k[6, 6] == k.max()
    g = np.exp(-np.arange(-6, 7) ** 2 / (2 * 2.6 ** 2)); g2 = np.outer(g, g)
    assert np.allclose(k, g2 / g2.sum(), rtol=1e-12)
    e = orc.gaussian_width(4.5, 20)                  # descriptor weights: even width, samples at +-0.5, +-1.5 ...
    d = np.abs(np.arange(20) - 9.5)
    ex = np.exp(-(d[None, :] ** 2 + d[:, None] ** 2) / (2 * 4.5 ** 2))
    assert np.allclose(e, ex / ex.sum(), rtol=1e-12)
    o = orc.gaussian_width(2.5, 9)                   # odd width -> gaussian2D(sigma, width/2)
    assert np.allclose(o, orc.gaussian2d(2.5, 4))


def test_gradient_definition_and_bounds(orc):
    """SparseIntegralGradient_NoBorder_F32.compute == difference of half-window sums; SparseGradientSafe -> 0 outside."""
    img = np.random.default_rng(1).integers(0, 50, (40, 40)).astype(np.float32)   # small ints: float sums exact
    ii = orc.integral(img)
    for width in (2.0, 4.0, 6.6):
        r = max(1, int(width + 0.5) // 2)
        xy = np.array([[20, 20], [r + 1, r + 1], [39 - r, 39 - r], [r, 20], [20, 40 - r], [0, 0]], np.int32)
        g = orc.gradient(ii, width, xy, safe=True)
        for (x, y), (dx, dy) in zip(xy, g):
            inb = x - r - 1 >= 0 and y - r - 1 >= 0 and x + r < 40 and y + r < 40
            if not inb:
                assert dx == 0 and dy == 0
                continue
            win = img[y - r:y + r + 1, x - r:x + r + 1]
            assert dx == win[:, r + 1:].sum() - win[:, :r].sum()
            assert dy == win[r + 1:, :].sum() - win[:r, :].sum()


def test_haar_definition_and_bounds(orc):
    """SparseIntegralHaar_NoBorder_F32.compute :46-73 == right half minus left half (bottom minus top) of the 2r x 2r window whose
    first pixel is (x-r+1, y-r+1); bounds x0 = y0 = -r, x1 = y1 = r (setWidth :37-43); SparseGradientSafe -> 0 outside."""
    img = np.random.default_rng(2).integers(0, 50, (40, 40)).astype(np.float32)
    ii = orc.integral(img)
    for width in (2.0, 4.0, 6.6):
        r = max(1, int(width + 0.5) // 2)
        xy = np.array([[20, 20], [r, r], [39 - r, 39 - r], [r - 1, 20], [20, 40 - r], [0, 0]], np.int32)
        g = orc.gradient(ii, width, xy, safe=True, haar=True)
        for (x, y), (dx, dy) in zip(xy, g):
            inb = x - r >= 0 and y - r >= 0 and x + r < 40 and y + r < 40
            if not inb:
                assert dx == 0 and dy == 0
                continue
            win = img[y - r + 1:y + r + 1, x - r + 1:x + r + 1]
            assert dx == win[:, r:].sum() - win[:, :r].sum()
            assert dy == win[r:, :].sum() - win[:r, :].sum()


def test_is_inside(orc):
    assert orc.is_inside_rect(100, 100, 10, 10, 20, 4)
    assert not orc.is_inside_rect(100, 100, 2, 10, 20, 4)        # x0 = 2 - 2 - 1 < 0
    assert not orc.is_inside_rect(100, 100, 10, 10, 88, 4)       # x1 = 98 + 2 >= 100
    assert orc.is_inside_rotated(100, 100, 50, 50, 10, 3, 1.0, 1.0, 0.0)
    assert not orc.is_inside_rotated(100, 100, 12, 50, 10, 3, 1.0, 1.0, 0.0)   # 12 - (10+2+1) < 0
    assert orc.is_inside_rotated(100, 100, 13, 50, 10, 3, 1.0, 1.0, 0.0)
    c = s = np.sqrt(0.5)
    assert not orc.is_inside_rotated(100, 100, 17, 50, 10, 3, 1.0, c, s)       # rotation grows the padding by ceil(13*(sqrt2-1))=6


@pytest.mark.parametrize("cfg,tol", [("ORI_AVERAGE", 0.01), ("ORI_SLIDING", np.pi / 9)])
def test_orientation_recovers_ramp_angle(orc, cfg, tol):
    cfg = getattr(orc, cfg)
    w, h = 130, 140            # large enough to hold 13 (17) samples at period 5 (3.25) plus the kernel
    n = 2 * int(np.pi / tol)
    for i in range(0, n, max(1, n // 60)):
        angle = np.arctan2(np.sin(i * tol), np.cos(i * tol))
        ii = orc.integral(_ramp(angle, w, h, 10.0))
        for radius in (10, 15, 7.5) if i == 0 else (10,):
            found = orc.orientation(ii, [w // 2], [h // 2], [radius / 2.0], cfg)[0]   # scale = radius * 0.5
            d = abs(found - angle); d = min(d, 2 * np.pi - d)
            assert d < tol, (i, angle, found)


@pytest.mark.parametrize("cfg", ["ORI_AVERAGE", "ORI_SLIDING"])
def test_orientation_border_and_constant(orc, cfg):
    cfg = getattr(orc, cfg)
    ii = orc.integral(_ramp(0.0, 30, 40, 10.0))
    xs = [0] * 40 + [29] * 40 + list(range(30)) * 2
    ys = list(range(40)) * 2 + [0] * 30 + [39] * 30
    a = orc.orientation(ii, xs, ys, [5.0] * len(xs), cfg)        # checkBorderExplode: no crash, finite
    assert np.isfinite(a).all()
    flat = orc.integral(np.full((40, 30), 50, np.float32))
    assert orc.orientation(flat, [15], [20], [2.0], cfg)[0] == 0.0   # atan2(0,0) = 0


@pytest.mark.parametrize("desc", ["DESC_SPEED", "DESC_STABILITY"])
class TestDescribe:
    def test_sub_image(self, orc, desc):
        ii = _noise_ii()
        big = np.zeros((H + 9, W + 6), np.float32); big[4:4 + H, 3:3 + W] = ii
        a, wa = orc.describe(ii, [W // 2], [H // 2], [0.0], [1.0], getattr(orc, desc))
        b, wb = orc.describe(big[4:4 + H, 3:3 + W], [W // 2], [H // 2], [0.0], [1.0], getattr(orc, desc))
        assert np.array_equal(a, b) and wa == wb

    def test_scale_and_rotation_change_it(self, orc, desc):
        ii = _noise_ii()
        cfg = getattr(orc, desc)
        a, _ = orc.describe(ii, [W // 2], [H // 2], [0.0], [1.0], cfg)
        b, _ = orc.describe(ii, [W // 2], [H // 2], [0.0], [1.5], cfg)
        c, _ = orc.describe(ii, [W // 2], [H // 2], [1.0], [1.0], cfg)
        assert np.abs(a - b).max() > 1e-4 and np.abs(a - c).max() > 1e-4

    def test_border_does_not_blow_up(self, orc, desc):
        ii = _noise_ii()
        ang = [2 * np.pi * i / 10 for i in range(10)]
        for x, y in ((0, 0), (W - 1, H - 1)):
            d, _ = orc.describe(ii, [x] * 10, [y] * 10, ang, [1.0] * 10, getattr(orc, desc))
            assert np.isfinite(d).all()

    def test_constant_image_gives_zero(self, orc, desc):
        ii = orc.integral(np.full((H, W), 50, np.float32))
        d, _ = orc.describe(ii, [20], [20], [0.75], [1.0], getattr(orc, desc))
        assert np.abs(d).max() <= 1e-4

    @pytest.mark.parametrize("x,y,scale", [(15, 15, 1.0), (25, 25, 1.5)])
    def test_x_ramp(self, orc, desc, x, y, scale):
        ii = orc.integral(_ramp(0.0))
        d, _ = orc.describe(ii, [x], [y], [0.0], [scale], getattr(orc, desc))
        d = d[0].reshape(16, 4)
        assert np.allclose(d[:, 0], d[:, 1], atol=1e-4) and (d[:, 0] > 0).all()
        assert np.abs(d[:, 2:]).max() <= 1e-4
        if scale == 1.0:
            d, _ = orc.describe(ii, [x], [y], [np.pi / 2], [scale], getattr(orc, desc))
            d = d[0].reshape(16, 4)
            assert np.allclose(-d[:, 2], d[:, 3], atol=1e-4) and (d[:, 2] < 0).all()
            assert np.abs(d[:, :2]).max() <= 1e-4

    def test_unit_norm_and_laplace_sign(self, orc, desc):
        img = np.zeros((80, 80), np.float32); img[36:45, 36:45] = 100     # bright blob => negative laplacian => not white
        ii = orc.integral(img)
        d, white = orc.describe(ii, [40], [40], [0.3], [1.2], getattr(orc, desc))
        assert abs((d * d).sum() - 1) < 1e-12 and not white[0]
        d, white = orc.describe(orc.integral(100 - img), [40], [40], [0.3], [1.2], getattr(orc, desc))
        assert white[0]


def test_describe_fast_matches_python_restatement(orc):
    """DescribePointSurf.features :235-295 written out independently in Python for one interior feature."""
    ii = _noise_ii(7, 90, 90)
    x, y, angle, scale = 44.3, 45.6, 0.4, 1.3
    got, _ = orc.describe(ii, [x], [y], [angle], [scale], orc.DESC_SPEED)
    s_, c_ = orc.sincos(np.array([angle])); c, s = c_[0], s_[0]
    wt = orc.gaussian_width(4.5, 20); wt = wt / wt[10, 10]
    r = max(1, int(3 * scale + 0.5) // 2)
    cx, cy = x + 0.5, y + 0.5
    feat = []
    for rY in range(-10, 10, 5):
        for rX in range(-10, 10, 5):
            acc = [0.0, 0.0, 0.0, 0.0]
            for i in range(5):
                regionY = (rY + i) * scale
                for j in range(5):
                    w = wt[10 + rY + i, 10 + rX + j]
                    regionX = (rX + j) * scale
                    px = int(cx + c * regionX - s * regionY); py = int(cy + s * regionX + c * regionY)
                    g = orc.gradient(ii, 3 * scale, np.array([[px, py]], np.int32), safe=False)[0]
                    dx, dy = w * g[0], w * g[1]
                    pdx = c * dx + s * dy; pdy = -s * dx + c * dy
                    acc[0] += pdx; acc[1] += abs(pdx); acc[2] += pdy; acc[3] += abs(pdy)
            feat += acc
    feat = np.array(feat)
    norm = 0.0
    for v in feat:
        norm += v * v
    assert np.array_equal(got[0], feat / np.sqrt(norm))


def test_detect_describe_interface_checks(orc):
    """GenericTestsDetectDescribePoint: 100x120 noise -> N>5, finite, in-bounds non-zero locations, sub-image and repeat identical."""
    img = np.random.default_rng(234).uniform(0, 100, (120, 100)).astype(np.float32)
    for fn in (orc.surf_fast, orc.surf_stable):
        a = fn(img)
        n = len(a["x"])
        assert n > 5 and a["desc"].shape == (n, 64)
        assert np.isfinite(a["desc"]).all() and np.isfinite(a["angle"]).all()
        assert (a["x"] != 0).all() and (a["y"] != 0).all()
        assert (a["x"] >= 0).all() and (a["x"] < 100).all() and (a["y"] >= 0).all() and (a["y"] < 120).all()
        assert (np.abs(a["angle"]) <= np.pi).all() and (a["scale"] > 0).all()
        b = fn(img)
        big = np.zeros((130, 111), np.float32); big[6:126, 5:105] = img
        c = fn(big[6:126, 5:105])
        for k in a:
            assert np.array_equal(a[k], b[k]) and np.array_equal(a[k], c[k])


def test_mt_equals_st(orc):
    """TestWrapDetectDescribeSurf_MT.java:54-101: 400x300 noise, >200 features, threaded == single-thread."""
    img = np.random.default_rng(234).uniform(0, 255, (300, 400)).astype(np.float32)
    for fn in (orc.surf_fast, orc.surf_stable):
        a = fn(img, nthreads=1); b = fn(img, nthreads=4)
        assert len(a["x"]) > 200
        for k in a:
            assert np.array_equal(a[k], b[k])


def test_u8_pipeline_runs(orc):
    img = np.clip(blob_image(160, 120, 3), 0, 255).astype(np.uint8)
    a = orc.surf(img)
    assert len(a["x"]) > 5 and np.isfinite(a["desc"]).all()
