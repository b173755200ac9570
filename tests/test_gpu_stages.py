"""GPU parity, stage by stage, through the C ABI vs the CPU oracle.  Bit-exact everywhere (the kernels keep the reference's
evaluation order); the reference's own tolerance for these stages is 1e-4 (BoofDefaults.TEST_FLOAT_TOL), stated where used.
Mirrors: TestImplIntegralImageOps, TestIntegralImageFeatureIntensity, GenericNonMaxTests / GenericFeatureDetectorTests,
GenericOrientationIntegralTests, BaseTestDescribeSurf."""
import numpy as np
import pytest

from synth import blob_image

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def st():
    from boofcv_b200 import stages
    return stages


def _noise(w, h, seed=234, lo=0.0, hi=100.0):
    return np.random.default_rng(seed).uniform(lo, hi, (h, w)).astype(np.float32)


# ------------------------------------------------------------------ integral image
@pytest.mark.parametrize("shape", [(20, 30), (640, 480), (1920, 1080), (33, 65), (1, 1), (1, 77), (77, 1), (257, 3)])
def test_integral_f32_bit_exact(st, orc, shape):
    w, h = shape
    img = _noise(w, h, 1, 0, 255)
    assert np.array_equal(st.integral(img), orc.integral(img))


@pytest.mark.parametrize("shape", [(20, 30), (641, 479), (1920, 1080), (1, 1), (1, 40), (7, 33), (1023, 31), (1025, 65), (2047, 40), (2049, 33),
                                   (3840, 64), (4100, 35), (8192, 33), (8200, 34)])
def test_integral_u8_s32_bit_exact(st, orc, shape):
    # integer images take the parallel scan (bands of 32 rows, 256 threads x 4 / 8 / 16 / 32 columns, vector loads when aligned);
    # widths beyond 8192 fall back to the serial passes; values wrap like Java int
    w, h = shape
    rng = np.random.default_rng(2)
    u8 = rng.integers(0, 256, (h, w), dtype=np.uint8)
    assert np.array_equal(st.integral(u8), orc.integral(u8))
    s32 = rng.integers(-2**31, 2**31 - 1, (h, w), dtype=np.int64).astype(np.int32)     # wraps like Java int
    assert np.array_equal(st.integral(s32), orc.integral(s32))


def test_integral_brute_force_definition(st):
    # TestImplIntegralImageOps.java:69-88 — integral = sum over the rectangle, tol 1e-1, 20x30
    img = _noise(20, 30, 3)
    ii = st.integral(img)
    ref = np.cumsum(np.cumsum(img.astype(np.float64), 0), 1)
    assert np.abs(ii - ref).max() < 1e-1


def test_integral_sub_image(st, orc):
    big = _noise(100, 80, 4)
    sub = big[10:50, 20:90]                   # row-strided view == BoofCV sub-image
    assert np.array_equal(st.integral(sub), orc.integral(np.ascontiguousarray(sub)))
    rng = np.random.default_rng(5)
    for dt, lo, hi in ((np.uint8, 0, 256), (np.int32, -2**31, 2**31 - 1)):      # misaligned rows for the scan's vector loads
        bigi = rng.integers(lo, hi, (90, 301), dtype=np.int64).astype(dt)
        for x0 in (0, 1, 3, 4):
            subi = bigi[5:77, x0:x0 + 263]
            assert np.array_equal(st.integral(subi), orc.integral(np.ascontiguousarray(subi))), (dt, x0)


def test_integral_empty(st):
    assert st.integral(np.zeros((0, 0), np.float32)).shape == (0, 0)


# ------------------------------------------------------------------ stand-alone NonMaxBlock (bs_nonmax)
@pytest.mark.parametrize("shape,radius,thr,border", [((120, 90), 2, 5.0, 0), ((121, 91), 2, 5.0, 3), ((640, 480), 1, 0.5, 1), ((77, 50), 5, -1e30, 0),
                                                     ((1920, 1080), 2, 50.0, 4), ((10, 10), 3, 0.0, 6)])
def test_nonmax_block_strict_equals_oracle(st, orc, shape, radius, thr, border):
    # GenericNonMaxTests / TestNonMaxBlockStrict: same points, same order as NonMaxBlock + NonMaxBlockSearchStrict.Max; the oracle's
    # block search is itself pinned against NonMaxExtractorNaive.strictRule (tests/test_oracle_detect.py)
    w, h = shape
    rng = np.random.default_rng(w + h)
    img = rng.uniform(0, 100, (h, w)).astype(np.float32)
    img[h // 2, w // 2] = 500.0; img[h // 2, w // 2 + 1] = 500.0          # equal neighbours: the strict rule rejects both
    got = st.nonmax(img, radius, thr, border)
    ref = orc.nonmax(img, radius, thr, border)
    assert np.array_equal(got, ref) and (len(ref) > 0 or w <= 2 * border)
    sub = np.zeros((h + 7, w + 9), np.float32); sub[3:3 + h, 5:5 + w] = img
    assert np.array_equal(st.nonmax(sub[3:3 + h, 5:5 + w], radius, thr, border), ref)      # strided view == sub-image


# ------------------------------------------------------------------ Fast-Hessian intensity
@pytest.mark.parametrize("skip,size", [(1, 9), (2, 9), (3, 9), (4, 9), (1, 15), (1, 27), (2, 51), (4, 99), (8, 195)])
def test_hessian_bit_exact_f32(st, orc, skip, size):
    img = _noise(400, 300, 5, 0, 200)
    ii = orc.integral(img)
    g = st.hessian(ii, skip, size)
    assert np.array_equal(g, orc.hessian(ii, skip, size))
    # the reference's own check: fast == naive within 1e-4 (TestIntegralImageFeatureIntensity.java:44-89)
    assert np.abs(g - orc.hessian(ii, skip, size, naive=True)).max() <= 1e-4 * max(1.0, np.abs(g).max())


@pytest.mark.parametrize("skip,size", [(1, 9), (2, 15), (4, 27)])
def test_hessian_bit_exact_s32(st, orc, skip, size):
    u8 = np.random.default_rng(6).integers(0, 256, (120, 160), dtype=np.uint8)
    ii = orc.integral(u8)
    assert np.array_equal(st.hessian(ii, skip, size), orc.hessian(ii, skip, size))


def test_hessian_small_image(st, orc):
    ii = orc.integral(_noise(12, 10, 7))
    assert np.array_equal(st.hessian(ii, 1, 9), orc.hessian(ii, 1, 9))


@pytest.mark.parametrize("shape", [(400, 300), (644, 196), (250, 197), (64, 48), (1920, 1080)])
def test_scale_space_layers_bit_exact(st, orc, shape):
    """Every layer of the scale space as the batch path's kernels write it (TMA-staged tile kernels for octaves 0-1 when the
    width is a multiple of 4, manual staging otherwise, the generic kernel for the sparse octaves): border ring in
    block_zero's summation order, inner region in hessianInner's, both bit for bit."""
    ii = orc.integral(_noise(shape[0], shape[1], 21, 0, 255))
    layers = st.fh_intensity(ii)
    sizes = [(L["skip"], L["size"]) for L in layers]
    assert sizes[:4] == [(1, 9), (1, 15), (1, 21), (1, 27)]
    for L in layers:
        o = orc.hessian(ii, L["skip"], L["size"], nthreads=8)
        assert L["intensity"].shape == o.shape
        assert np.array_equal(L["intensity"].view(np.uint32), o.view(np.uint32)), (L["skip"], L["size"])


def test_scale_space_integer_integral(st, orc):
    u8 = np.random.default_rng(8).integers(0, 256, (120, 160), dtype=np.uint8)
    ii = orc.integral(u8)
    for L in st.fh_intensity(ii):
        assert np.array_equal(L["intensity"], orc.hessian(ii, L["skip"], L["size"])), (L["skip"], L["size"])


# ------------------------------------------------------------------ detector
def _same_detection(g, o):
    assert len(g["x"]) == len(o["x"])
    for k in ("x", "y", "scale", "octave", "level", "px", "py", "response"):
        assert np.array_equal(g[k], o[k]), k


@pytest.mark.parametrize("shape,seed", [((400, 300), 234), ((640, 480), 11), ((250, 197), 12)])
def test_fh_detect_same_points_same_order(st, orc, shape, seed):
    img = blob_image(shape[0], shape[1], seed)
    ii = orc.integral(img)
    g, o = st.fh_detect(ii), orc.fh_detect(ii)
    assert len(o["x"]) > 50
    _same_detection(g, o)


def test_fh_detect_noise_image(st, orc):
    # the reference's detector tests use uniform noise (GenericTestsDetectDescribePoint.java:83-211, Random(234))
    ii = orc.integral(_noise(400, 300, 234))
    g, o = st.fh_detect(ii), orc.fh_detect(ii)
    assert len(o["x"]) > 200
    _same_detection(g, o)


@pytest.mark.parametrize("cfg", [dict(extract_radius=1), dict(extract_radius=3, detect_threshold=10.0),
                                 dict(octaves=2, scales_per_octave=5), dict(initial_sample_size=2, initial_size=15)])
def test_fh_detect_configs(st, orc, cfg):
    ii = orc.integral(blob_image(320, 240, 5))
    _same_detection(st.fh_detect(ii, cfg), orc.fh_detect(ii, cfg))


def test_fh_detect_integer_integral(st, orc):
    u8 = np.clip(blob_image(320, 240, 8), 0, 255).astype(np.uint8)
    ii = orc.integral(u8)
    g, o = st.fh_detect(ii), orc.fh_detect(ii)
    assert len(o["x"]) > 20
    _same_detection(g, o)


def test_fh_detect_too_small_for_any_octave(st):
    ii = np.ones((20, 20), np.float32).cumsum(0).cumsum(1).astype(np.float32)
    assert len(st.fh_detect(ii)["x"]) == 0


def test_fh_detect_constant_image_has_no_points(st, orc):
    ii = orc.integral(np.full((200, 200), 50, np.float32))
    assert len(st.fh_detect(ii)["x"]) == len(orc.fh_detect(ii)["x"]) == 0


@pytest.mark.parametrize("n", [1, 5, 40, 300])
def test_fh_detect_n_best_per_scale(st, orc, n):
    # SelectNBestFeatures (SelectNBestFeatures.java:51-93) through ConfigFastHessian.maxFeaturesPerScale
    ii = orc.integral(blob_image(640, 480, 21))
    cfg = dict(max_features_per_scale=n)
    g, o = st.fh_detect(ii, cfg), orc.fh_detect(ii, cfg)
    assert len(o["x"]) > 0
    _same_detection(g, o)
    full = orc.fh_detect(ii)
    if n <= 40:
        assert len(o["x"]) < len(full["x"])           # the limit really cut something


def test_fh_detect_n_best_with_equal_responses(st, orc):
    # a periodic pattern gives many maxima with bit-identical responses: the tie rule (earlier block wins) is exercised
    tile = blob_image(48, 48, 3)
    ii = orc.integral(np.tile(tile, (6, 8)).astype(np.float32))
    for n in (3, 17, 60):
        cfg = dict(max_features_per_scale=n, octaves=2)
        _same_detection(st.fh_detect(ii, cfg), orc.fh_detect(ii, cfg))
    u8 = np.tile(np.clip(tile, 0, 255).astype(np.uint8), (6, 8))
    ii = orc.integral(u8)                              # integer integral: exact ties everywhere
    for n in (2, 25):
        cfg = dict(max_features_per_scale=n)
        g, o = st.fh_detect(ii, cfg), orc.fh_detect(ii, cfg)
        _same_detection(g, o)


# ------------------------------------------------------------------ orientation / descriptor
def _points(orc, img):
    ii = orc.integral(img)
    d = orc.fh_detect(ii)
    return ii, d


@pytest.mark.parametrize("variant", ["average", "sliding"])
def test_orientation_bit_exact(st, orc, variant):
    ii, d = _points(orc, blob_image(400, 300, 21))
    cfg = orc.ORI_AVERAGE if variant == "average" else orc.ORI_SLIDING
    g = st.orientation(ii, d["x"], d["y"], d["scale"], cfg)
    o = orc.orientation(ii, d["x"], d["y"], d["scale"], cfg)
    assert len(o) > 50 and np.array_equal(g, o)


def test_orientation_unweighted_and_border_points(st, orc):
    ii = orc.integral(_noise(120, 100, 22))
    x = np.array([2.0, 117.5, 60.0, 0.0, 119.0]); y = np.array([3.0, 50.0, 98.0, 0.0, 99.0]); s = np.array([1.2, 2.5, 4.0, 1.0, 6.0])
    for base in (orc.ORI_AVERAGE, orc.ORI_SLIDING):
        for sigma in (-1.0, 0.0, 2.0):
            cfg = dict(base, weight_sigma=sigma)
            assert np.array_equal(st.orientation(ii, x, y, s, cfg), orc.orientation(ii, x, y, s, cfg))


def test_orientation_ramp_recovers_angle(st, orc):
    # GenericOrientationIntegralTests.java:102-143 — intensity ramp along theta -> that angle
    yy, xx = np.mgrid[0:100, 0:100].astype(np.float64)
    for theta in np.linspace(-np.pi, np.pi, 9)[1:]:
        img = (np.cos(theta) * xx + np.sin(theta) * yy + 200).astype(np.float32)
        ii = orc.integral(img)
        a = st.orientation(ii, [50.0], [50.0], [2.0])[0]
        assert abs(np.arctan2(np.sin(a - theta), np.cos(a - theta))) < 0.05


@pytest.mark.parametrize("variant", ["speed", "stability"])
def test_describe_bit_exact(st, orc, variant):
    ii, d = _points(orc, blob_image(400, 300, 23))
    ang = orc.orientation(ii, d["x"], d["y"], d["scale"])
    cfg = orc.DESC_SPEED if variant == "speed" else orc.DESC_STABILITY
    gd, gw = st.describe(ii, d["x"], d["y"], ang, d["scale"], cfg)
    od, ow = orc.describe(ii, d["x"], d["y"], ang, d["scale"], cfg)
    assert len(od) > 50 and np.array_equal(gd, od) and np.array_equal(gw, ow)
    assert np.allclose(np.linalg.norm(gd, axis=1), 1.0, atol=1e-12)


@pytest.mark.parametrize("variant", ["speed", "stability"])
def test_describe_haar_bit_exact(st, orc, variant):
    # ConfigSurfDescribe.useHaar = true: SparseIntegralHaar_NoBorder_F32 samples the descriptor (border features included, where
    # SparseGradientSafe returns zero for samples whose 2r x 2r window leaves the image)
    ii, d = _points(orc, blob_image(400, 300, 23))
    ang = orc.orientation(ii, d["x"], d["y"], d["scale"])
    cfg = dict(orc.DESC_SPEED if variant == "speed" else orc.DESC_STABILITY); cfg["use_haar"] = 1
    gd, gw = st.describe(ii, d["x"], d["y"], ang, d["scale"], cfg)
    od, ow = orc.describe(ii, d["x"], d["y"], ang, d["scale"], cfg)
    assert len(od) > 50 and np.array_equal(gd, od) and np.array_equal(gw, ow)
    plain, _ = orc.describe(ii, d["x"], d["y"], ang, d["scale"], orc.DESC_SPEED if variant == "speed" else orc.DESC_STABILITY)
    assert not np.array_equal(od, plain)                         # the switch really changes the sampler
    x = np.array([2.0, 397.0, 200.0]); y = np.array([2.0, 297.0, 1.0]); s = np.array([1.5, 2.0, 3.0]); a = np.array([0.0, 1.0, -2.0])
    gd, gw = st.describe(ii, x, y, a, s, cfg)
    od, ow = orc.describe(ii, x, y, a, s, cfg)
    assert np.array_equal(gd, od) and np.array_equal(gw, ow)


def test_describe_border_and_constant(st, orc):
    # BaseTestDescribeSurf.java:125-151 — features at the border do not blow up; constant image -> all zeros
    ii = orc.integral(_noise(80, 60, 24))
    x = np.array([0.0, 79.0, 40.0, 5.0]); y = np.array([0.0, 59.0, 30.0, 55.0]); s = np.array([1.5, 2.0, 3.0, 4.0])
    a = np.array([0.0, 1.0, -2.0, 3.0])
    for cfg in (orc.DESC_SPEED, orc.DESC_STABILITY):
        gd, gw = st.describe(ii, x, y, a, s, cfg)
        od, ow = orc.describe(ii, x, y, a, s, cfg)
        assert np.array_equal(gd, od) and np.array_equal(gw, ow) and np.isfinite(gd).all()
    flat = orc.integral(np.full((100, 100), 50, np.float32))
    gd, _ = st.describe(flat, [50.0], [50.0], [0.3], [2.0])
    assert np.all(gd == 0)


def test_describe_ramp_pattern(st, orc):
    # BaseTestDescribeSurf.java:159-186 — x-ramp: v[i]==v[i+1]>0, v[i+2]==v[i+3]==0 per sub-region
    yy, xx = np.mgrid[0:100, 0:100]
    ii = orc.integral((xx * 2.0).astype(np.float32))
    gd, _ = st.describe(ii, [50.0], [50.0], [0.0], [1.0])
    v = gd[0].reshape(16, 4)
    assert np.allclose(v[:, 0], v[:, 1]) and (v[:, 0] > 0).all() and np.allclose(v[:, 2:], 0)


def test_describe_integer_integral(st, orc):
    u8 = np.clip(blob_image(320, 240, 25), 0, 255).astype(np.uint8)
    ii = orc.integral(u8)
    d = orc.fh_detect(ii)
    for ocfg, dcfg in ((orc.ORI_AVERAGE, orc.DESC_SPEED), (orc.ORI_SLIDING, orc.DESC_STABILITY)):
        ga = st.orientation(ii, d["x"], d["y"], d["scale"], ocfg)
        assert np.array_equal(ga, orc.orientation(ii, d["x"], d["y"], d["scale"], ocfg))
        gd, gw = st.describe(ii, d["x"], d["y"], ga, d["scale"], dcfg)
        od, ow = orc.describe(ii, d["x"], d["y"], ga, d["scale"], dcfg)
        assert np.array_equal(gd, od) and np.array_equal(gw, ow)
