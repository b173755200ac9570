"""Tensor-core candidate filter (tcgen05, fp16 operands scaled per row to norm 1..2, fp16 accumulators, K = 64) + exact fp64 re-score — bs_assoc mode 1, through the C ABI.

The contract is the reference's (AssociateGreedy.java:65-118): pairs and fit must be BIT-IDENTICAL to the exact kernel and to
the oracle, whatever the tensor cores round.  The filter only has to keep the exact minimum (and its exact ties) among the
candidates, which holds while its estimate stays inside the error bound it assumes; the first test measures that."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bs():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import boofcv_b200
    return boofcv_b200


def _unit(rng, n, dof=64):
    x = rng.normal(size=(n, dof))
    return x / np.linalg.norm(x, axis=1, keepdims=True)


def _exact_scores(a, b):
    s = np.zeros((len(a), len(b)))
    for k in range(a.shape[1]):
        d = a[:, k][:, None] - b[:, k][None, :]
        s = s + d * d
    return s


def _greedy(bs, backwards, max_fit=np.finfo(np.float64).max, mode=1):
    g = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), max_fit, backwards)
    g.setMode(mode)
    return g


def test_filter_estimate_stays_inside_its_error_budget(bs):
    # 300 x 517: ragged against the 256-row / 112-column tiles; norm regimes: unit, small, large (beyond fp16's range without the per-row
    # scaling), and rows of very different magnitude in one problem
    rng = np.random.default_rng(1)
    for scale in (1.0, 0.05, 300.0, None):
        if scale is None:                                       # every row its own magnitude, 1e-4 .. 1e4
            a = _unit(rng, 300) * 10.0 ** rng.uniform(-4, 4, size=(300, 1)); b = _unit(rng, 517) * 10.0 ** rng.uniform(-4, 4, size=(517, 1))
        else:
            a = _unit(rng, 300) * scale; b = _unit(rng, 517) * scale
        b[7] = a[3]                                             # an exact zero distance
        g = _greedy(bs, False)
        scores, cnt, _ = g.filterScores(a, b)
        exact = _exact_scores(a, b)
        na = (a * a).sum(1)[:, None]; nb = (b * b).sum(1)[None, :]
        budget = 6.4e-3 * np.sqrt(na * nb) + 1e-6 * (na + nb)   # bs_assoc_tc.cu tolerance_of() is twice this
        err = np.abs(scores.astype(np.float64) - exact)
        assert np.isfinite(scores).all()
        assert (err <= 0.75 * budget).all(), float((err / budget).max())     # the budget is a worst-case bound
        # (with mixed magnitudes the tolerance, which is relative to the LARGEST column norm, admits hundreds of small columns: such a
        # problem ends in the exact kernel — test_mode1_rows_of_mixed_magnitude)
        assert (cnt >= 1).all() and (scale is None or cnt.max() <= 128)
        g.close()


@pytest.mark.parametrize("n,m,backwards", [(300, 517, True), (1000, 777, False), (2048, 2304, True), (129, 4100, True),
                                            (600, 100, True), (600, 224, False), (70, 112, True), (1, 113, True)])   # one / two column tiles per row block
def test_mode1_bit_exact_vs_exact_kernel_and_oracle(bs, orc, n, m, backwards):
    rng = np.random.default_rng(n + m)
    dst = _unit(rng, m); src = _unit(rng, n)
    k = min(n, m) // 2
    src[:k] = dst[rng.permutation(m)[:k]] + 0.02 * rng.normal(size=(k, 64))      # planted matches with noise
    tc, ex = _greedy(bs, backwards), _greedy(bs, backwards, mode=0)
    for g in (tc, ex):
        g.setSource(src); g.setDestination(dst); g.associate()
    assert np.array_equal(tc.getPairs(), ex.getPairs())
    assert np.array_equal(tc.getFitQuality(), ex.getFitQuality())
    op, of = orc.associate(src, dst, np.finfo(np.float64).max, backwards)
    assert np.array_equal(tc.getPairs(), op) and np.array_equal(tc.getFitQuality(), of)
    assert (tc.getPairs() >= 0).sum() >= k // 2
    _, _, (runs, fallbacks) = tc.filterScores(src[:1], dst[:1], want_scores=False)
    assert runs >= 1 and fallbacks == 0                        # the tensor-core path really produced the result
    tc.close(); ex.close()


def test_mode1_rows_of_mixed_magnitude(bs, orc):
    # descriptors that are NOT normalised, magnitudes from 1e-3 to 1e3 inside one problem (the operands are scaled per row before the
    # fp16 rounding: bs_assoc_tc.cu split_kernel), with planted matches at every magnitude
    rng = np.random.default_rng(77)
    n, m = 900, 1300
    dst = _unit(rng, m) * 10.0 ** rng.uniform(-3, 3, size=(m, 1))
    src = _unit(rng, n) * 10.0 ** rng.uniform(-3, 3, size=(n, 1))
    k = 400
    pick = rng.permutation(m)[:k]
    src[:k] = dst[pick] * (1.0 + 0.01 * rng.normal(size=(k, 64)))
    for backwards in (False, True):
        tc, ex = _greedy(bs, backwards), _greedy(bs, backwards, mode=0)
        for g in (tc, ex):
            g.setSource(src); g.setDestination(dst); g.associate()
        op, of = orc.associate(src, dst, np.finfo(np.float64).max, backwards)
        assert np.array_equal(tc.getPairs(), op) and np.array_equal(tc.getFitQuality(), of)
        assert np.array_equal(ex.getPairs(), op) and np.array_equal(ex.getFitQuality(), of)
        tc.close(); ex.close()


def test_mode1_norms_outside_the_filter_range_fall_back(bs, orc):
    # rows of norm 1e-20 and 1e+20 (squared norms beyond 1e-30 .. 1e30): outside what the fp16 operands and the float scales cover ->
    # the whole problem is redone with the exact kernel, results unchanged
    rng = np.random.default_rng(21)
    for scale in (1e-20, 1e20):
        dst = _unit(rng, 400); src = _unit(rng, 300)
        src[:100] = dst[rng.permutation(400)[:100]] + 0.01 * rng.normal(size=(100, 64))
        dst[7] *= scale; src[3] *= scale
        tc = _greedy(bs, True)
        tc.setSource(src); tc.setDestination(dst); tc.associate()
        op, of = orc.associate(src, dst, np.finfo(np.float64).max, True)
        assert np.array_equal(tc.getPairs(), op) and np.array_equal(tc.getFitQuality(), of), scale
        _, _, (runs, fallbacks) = tc.filterScores(src[:1], dst[:1], want_scores=False)
        assert fallbacks >= 1
        tc.close()


def test_mode1_near_ties_duplicates_and_threshold(bs, orc):
    # adversarial for a low-precision filter: destinations that differ by 1e-9 .. 1e-4 (far below fp16 resolution), exact
    # duplicates (tie -> later index forward, rejected backward), duplicate sources, and an inclusive threshold
    rng = np.random.default_rng(5)
    m = 640
    dst = _unit(rng, m)
    for j, eps in zip(range(100, 140), np.logspace(-9, -4, 40)):
        dst[j + 200] = dst[j] + eps * rng.normal(size=64)       # near-duplicates of dst[100..139]
    dst[500] = dst[20]; dst[501] = dst[20]                     # exact triplicate
    src = np.concatenate([dst[100:140] + 1e-6 * rng.normal(size=(40, 64)), dst[20:21], dst[20:21], _unit(rng, 300)])
    for backwards in (False, True):
        for max_fit in (np.finfo(np.float64).max, 0.5, 1e-9):
            tc = _greedy(bs, backwards, max_fit)
            tc.setSource(src); tc.setDestination(dst); tc.associate()
            op, of = orc.associate(src, dst, max_fit, backwards)
            assert np.array_equal(tc.getPairs(), op), (backwards, max_fit)
            assert np.array_equal(tc.getFitQuality(), of), (backwards, max_fit)
            tc.close()
    assert op[40] == -1 or not backwards


def test_mode1_candidate_overflow_falls_back_to_exact(bs, orc):
    # 200 identical destinations: every one of them is a candidate of every source -> lists overflow (cap 128) -> exact redo
    rng = np.random.default_rng(9)
    dst = np.repeat(_unit(rng, 1), 200, axis=0); src = _unit(rng, 50)
    tc = _greedy(bs, True)
    tc.setSource(src); tc.setDestination(dst); tc.associate()
    op, of = orc.associate(src, dst, np.finfo(np.float64).max, True)
    assert np.array_equal(tc.getPairs(), op) and np.array_equal(tc.getFitQuality(), of)
    _, _, (runs, fallbacks) = tc.filterScores(src[:1], dst[:1], want_scores=False)
    assert fallbacks >= 1
    tc.close()


def test_mode1_rejects_other_descriptor_lengths_and_auto_mode_picks_by_size(bs, orc):
    rng = np.random.default_rng(3)
    g = _greedy(bs, True)
    g.setSource(rng.normal(size=(10, 32))); g.setDestination(rng.normal(size=(10, 32)))
    with pytest.raises(ValueError):
        g.associate()
    g.close()
    auto = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), np.finfo(np.float64).max, True)     # mode -1
    src, dst = _unit(rng, 2100), _unit(rng, 2100)              # 2100 * 2100 >= 2^22 -> tensor-core path
    auto.setSource(src); auto.setDestination(dst); auto.associate()
    op, of = orc.associate(src, dst, np.finfo(np.float64).max, True)
    assert np.array_equal(auto.getPairs(), op) and np.array_equal(auto.getFitQuality(), of)
    _, _, (runs, _) = auto.filterScores(src[:1], dst[:1], want_scores=False)
    assert runs >= 1
    auto.close()


def test_mode1_large_property_checks(bs):
    # BASELINE-sized shapes are checked through properties: planted permutation recovered exactly, backward validation
    # symmetric, tensor-core and exact kernels agree on a 20k x 20k instance
    import torch
    rng = np.random.default_rng(11)
    n = 20000
    dst = _unit(rng, n)
    perm = rng.permutation(n)
    src = dst[perm] + 0.01 * rng.normal(size=(n, 64))
    d_src, d_dst = torch.from_numpy(src).cuda(), torch.from_numpy(dst).cuda()
    tc, ex = _greedy(bs, True), _greedy(bs, True, mode=0)
    p1, f1 = tc.associateDevice(d_src, d_dst)
    p0, f0 = ex.associateDevice(d_src, d_dst)
    torch.cuda.synchronize()
    assert torch.equal(p1, p0) and torch.equal(f1, f0)
    assert np.array_equal(p1.cpu().numpy(), perm)
    tc.close(); ex.close()
