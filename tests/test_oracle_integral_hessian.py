"""Pins oracle integral-image / block / Fast-Hessian functions with the reference's own tests, re-expressed.

Reference tests restated (paths under /root/reference/main):
  boofcv-ip/src/test/java/boofcv/alg/transform/ii/impl/TestImplIntegralImageOps.java:55-88   (transform = brute-force sum, 20x30, tol 1e-1, sub-images)
  ...TestImplIntegralImageOps.java:131-178                                                     (block_unsafe / block_zero known answers)
  boofcv-feature/src/test/java/boofcv/alg/feature/detect/intensity/impl/TestImplIntegralImageFeatureIntensity.java:43-100
                                                                                               (hessianInner == hessianNaive inside the border, 60x70, skip 1..4, size 9, tol 1e-4)
  boofcv-feature/src/test/java/boofcv/alg/feature/detect/intensity/TestIntegralImageFeatureIntensity.java:44-89
                                                                                               (hessian == hessianNaive on the whole image)
"""
import numpy as np
import pytest

W, H = 20, 30


def _rand_img(dtype, w=W, h=H, hi=100, seed=234):
    r = np.random.default_rng(seed)
    if dtype == np.float32:
        return r.uniform(0, hi, (h, w)).astype(np.float32)
    return r.integers(0, hi, (h, w)).astype(dtype)


@pytest.mark.parametrize("dtype", [np.float32, np.uint8, np.int32])
@pytest.mark.parametrize("sub", [False, True])
def test_transform_is_brute_force_sum(orc, dtype, sub):
    img = _rand_img(dtype)
    if sub:  # BoofTesting.checkSubImage: same pixels viewed inside a larger array (startIndex/stride != width)
        big = np.zeros((H + 7, W + 5), dtype)
        big[3:3 + H, 2:2 + W] = img
        view = big[3:3 + H, 2:2 + W]
        assert view.strides[0] != W * view.itemsize
    else:
        view = img
    ii = orc.integral(view)
    expected = np.cumsum(np.cumsum(img.astype(np.float64), axis=0), axis=1)
    assert np.abs(ii.astype(np.float64) - expected).max() <= 1e-1
    if dtype != np.float32:
        assert ii.dtype == np.int32 and np.array_equal(ii.astype(np.int64), expected.astype(np.int64))


def test_transform_f32_summation_order(orc):
    """ImplIntegralImageOps.java:48-65: serial row total, then prev row + total, all in float."""
    img = _rand_img(np.float32, 37, 23, hi=255.0, seed=5)
    ii = orc.integral(img)
    exp = np.zeros_like(img)
    for y in range(img.shape[0]):
        total = np.float32(0)
        for x in range(img.shape[1]):
            total = np.float32(total + img[y, x])
            exp[y, x] = total if y == 0 else np.float32(exp[y - 1, x] + total)
    assert np.array_equal(ii, exp)


def test_transform_s32_wraps_like_java_int(orc):
    img = np.full((4, 4), 2 ** 30, np.int32)
    ii = orc.integral(img)
    exp = (np.cumsum(np.cumsum(img.astype(np.int64), 0), 1) & 0xFFFFFFFF).astype(np.uint32).view(np.int32)
    assert np.array_equal(ii, exp.reshape(4, 4))


@pytest.mark.parametrize("dtype", [np.float32, np.uint8])
def test_block_known_answers(orc, dtype):
    ii = orc.integral(np.ones((H, W), dtype))
    assert orc.block_unsafe(ii, 4, 5, 8, 8) == 12
    assert orc.block_zero(ii, 4, 5, 8, 8) == 12
    assert orc.block_zero(ii, -1, -2, 2, 3) == 12
    assert orc.block_zero(ii, W - 2, H - 3, W + 1, H + 3) == 2
    assert orc.block_zero(ii, 3, -4, -1, -1) == 0
    assert orc.block_zero(ii, W + 1, H + 2, W + 6, H + 8) == 0


def test_derivative_kernels(orc):
    """DerivativeIntegralImage.java:102-158 for size 9: lobes of 3, band of 5."""
    assert orc.kernel_deriv("xx", 9) == [(-5, -3, 4, 2, 1), (-2, -3, 1, 2, -3)]
    assert orc.kernel_deriv("yy", 9) == [(-3, -5, 2, 4, 1), (-3, -2, 2, 1, -3)]
    assert orc.kernel_deriv("xy", 9) == [(-4, -4, -1, -1, 1), (0, -4, 3, -1, -1), (0, 0, 3, 3, 1), (-4, 0, -1, 3, -1)]


@pytest.mark.parametrize("dtype", [np.float32, np.int32])
@pytest.mark.parametrize("skip", [1, 2, 3, 4])
def test_hessian_equals_naive(orc, dtype, skip):
    w, h, size = 60, 70, 9
    img = _rand_img(np.float32 if dtype == np.float32 else np.uint8, w, h, hi=100)
    ii = orc.integral(img)
    fast = orc.hessian(ii, skip, size)
    naive = orc.hessian(ii, skip, size, naive=True)
    assert fast.shape == (h // skip, w // skip)
    assert np.abs(fast - naive).max() <= 1e-4 * max(1.0, np.abs(naive).max())


def _inner_numpy(ii, skip, size):
    """Independent float32 restatement of hessianInner (ImplIntegralImageFeatureIntensity.java:132-213) in numpy,
    keeping Java's operand order.  NOTE Dyy is br-bl-tr+tl there (:187-188) while block_zero/convolveSparse give
    br-tr-bl+tl (ImplIntegralImageOps.java:195-214): inner and border pixels are NOT the same float expression."""
    H, W = ii.shape
    w, h = W // skip, H // skip
    bS = size // 3; bL = size - bS - 1; rF = size // 2; rS = bL // 2
    borderOrig = rF + 1 + (skip - (rF + 1) % skip); border = borderOrig // skip
    ys = np.arange(border, h - border) * skip; xs = np.arange(border, w - border) * skip
    Y, X = np.meshgrid(ys, xs, indexing="ij")
    g = lambda yy, xx: ii[yy, xx]
    f = np.float32
    top, bot, left = Y - rS - 1, Y - rS - 1 + bL, X - rF - 1
    Dxx = g(bot, left + 3 * bS) - g(top, left + 3 * bS) - g(bot, left) + g(top, left)
    Dxx = Dxx - f(3) * (g(bot, left + 2 * bS) - g(top, left + 2 * bS) - g(bot, left + bS) + g(top, left + bS))
    L, R, T = X - rS - 1, X - rS - 1 + bL, Y - rF - 1
    Dyy = g(T + 3 * bS, R) - g(T + 3 * bS, L) - g(T, R) + g(T, L)
    Dyy = Dyy - f(3) * (g(T + 2 * bS, R) - g(T + 2 * bS, L) - g(T + bS, R) + g(T + bS, L))
    y1, y2, y3, y4 = Y - bS - 1, Y - 1, Y, Y + bS
    x0 = X - bS - 1; x3 = x0 + bS + 1; x4 = x3 + bS
    Dxy = g(y2, x0 + bS) - g(y1, x0 + bS) - g(y2, x0) + g(y1, x0)
    Dxy = Dxy - (g(y2, x4) - g(y1, x4) - g(y2, x3) + g(y1, x3))
    Dxy = Dxy + (g(y4, x4) - g(y3, x4) - g(y4, x3) + g(y3, x3))
    Dxy = Dxy - (g(y4, x0 + bS) - g(y3, x0 + bS) - g(y4, x0) + g(y3, x0))
    norm = f(1.0) / f(size * size)
    Dxx = Dxx * norm; Dxy = Dxy * norm; Dyy = Dyy * norm
    assert Dxx.dtype == np.float32
    return border, Dxx * Dyy - f(0.81) * Dxy * Dxy


@pytest.mark.parametrize("size,skip", [(9, 1), (15, 2), (27, 1), (51, 4)])
def test_hessian_inner_bit_exact_vs_numpy_restatement(orc, size, skip):
    ii = orc.integral(_rand_img(np.float32, 260, 230, hi=255, seed=size))
    border, exp = _inner_numpy(ii, skip, size)
    got = orc.hessian(ii, skip, size)
    h, w = got.shape
    assert np.array_equal(got[border:h - border, border:w - border], exp)
    naive = orc.hessian(ii, skip, size, naive=True)
    m = np.ones_like(got, bool); m[border:h - border, border:w - border] = False
    assert np.array_equal(got[m], naive[m])          # border pixels are the convolveSparse expression


@pytest.mark.parametrize("size,skip", [(15, 1), (27, 2), (51, 4), (99, 8), (21, 1), (195, 8)])
def test_hessian_equals_naive_all_default_sizes(orc, size, skip):
    img = _rand_img(np.float32, 320, 240, hi=255, seed=size)
    ii = orc.integral(img)
    a, b = orc.hessian(ii, skip, size), orc.hessian(ii, skip, size, naive=True)
    assert np.abs(a - b).max() <= 1e-4 * max(1.0, np.abs(b).max())


def test_hessian_threads_equal_single(orc):
    ii = orc.integral(_rand_img(np.float32, 200, 150, hi=255))
    assert np.array_equal(orc.hessian(ii, 2, 27, nthreads=1), orc.hessian(ii, 2, 27, nthreads=4))


def test_hessian_blob_sign(orc):
    """A bright blob on a dark background gives a positive determinant at its centre (both curvatures negative)."""
    img = np.zeros((60, 60), np.float32)
    img[29:32, 29:32] = 100
    inten = orc.hessian(orc.integral(img), 1, 9)
    assert inten[30, 30] > 0 and inten[30, 30] == inten.max()
