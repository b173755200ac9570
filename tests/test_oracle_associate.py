"""Pins oracle greedy association with the reference's known-answer vectors.

  boofcv-feature/src/test/java/boofcv/alg/feature/associate/TestAssociateGreedy.java:38-102 (basic / maxError / backwards)
  boofcv-feature/src/test/java/boofcv/abst/feature/associate/StandardAssociateDescriptionChecks.java:76-239
      (inclusive threshold, uniqueDestination, unassociated lists)
  boofcv-feature/src/test/java/boofcv/alg/feature/associate/TestAssociateGreedy_MT.java:32-59 (threaded == single)
Scores in those tests are ScoreAssociateEuclidean_F64 (|a-b| for 1-d), hence score="euclidean".
"""
import numpy as np

from synth import descriptors, planted_descriptors

DMAX = np.finfo(np.float64).max


def test_basic(orc):
    p, f = orc.associate([1, 2, 3, 4], [3, 4, 1, 40], max_fit_error=0.5, backwards=False, score="euclidean")
    assert p.tolist() == [2, -1, 0, 1]
    assert f[0] == 0 and f[2] == 0 and f[3] == 0


def test_max_error(orc):
    p, _ = orc.associate([1, 2, 3, 4], [3, 4, 1.1, 40], max_fit_error=10, score="euclidean")
    assert p[1] == 2
    p, _ = orc.associate([1, 2, 3, 4], [3, 4, 1.1, 40], max_fit_error=0.1, score="euclidean")
    assert p[1] == -1


def test_backwards(orc):
    p, f = orc.associate([1, 2, 3, 8], [3, 4, 1, 10], max_fit_error=10, backwards=True, score="euclidean")
    assert p.tolist() == [2, -1, 0, 3]
    assert f[0] == 0 and f[2] == 0 and f[3] == 2 and f[1] == DMAX     # rejected rows get Double.MAX_VALUE (:110)


def test_threshold_is_inclusive(orc):
    t = 1.1 - 1
    p, _ = orc.associate([1.0], [1.1], max_fit_error=t, score="euclidean")
    assert p[0] == 0
    p, _ = orc.associate([1.0], [1.1], max_fit_error=np.nextafter(t, 0), score="euclidean")
    assert p[0] == -1


def test_unique_destination_with_backwards(orc):
    p, _ = orc.associate([1.0, 1.001], [1.0], backwards=True, score="euclidean")
    assert (p >= 0).sum() == 1 and p[0] == 0
    p, _ = orc.associate([1.0, 1.0], [1.0], backwards=True, score="euclidean")   # exact tie: "<=" rejects both (:109)
    assert (p >= 0).sum() == 0


def test_ties_pick_last_index(orc):
    p, _ = orc.associate([5.0], [7.0, 3.0, 7.0, 3.0], score="euclidean")         # "fit <= bestScore" (:88)
    assert p[0] == 3


def test_empty_lists(orc):
    p, f = orc.associate(np.zeros((3, 64)), np.zeros((0, 64)))
    assert p.tolist() == [-1, -1, -1] and (f == DMAX).all()
    p, f = orc.associate(np.zeros((0, 64)), np.zeros((4, 64)), backwards=True)
    assert len(p) == 0


def test_matches_and_unassociated_lists(orc):
    p, f = orc.associate([1, 2, 3, 8], [3, 4, 1, 10], max_fit_error=10, backwards=True, score="euclidean")
    m, us, ud = orc.matches_from_pairs(p, f, 4)
    assert [(a, b) for a, b, _ in m] == [(0, 2), (2, 0), (3, 3)] and us.tolist() == [1] and ud.tolist() == [1]


def test_mt_equals_st_and_nobuffer(orc):
    r = np.random.default_rng(234)
    a = r.uniform(0, 1, (200, 1)); b = r.uniform(0, 1, (200, 1))             # TestAssociateGreedy_MT: 200 x 200, 1-d
    for bw in (False, True):
        p1, f1 = orc.associate(a, b, backwards=bw, nthreads=1)
        p2, f2 = orc.associate(a, b, backwards=bw, nthreads=4)
        p3, f3 = orc.associate(a, b, backwards=bw, nobuffer=True, nthreads=3)
        assert np.array_equal(p1, p2) and np.array_equal(f1, f2) and np.array_equal(p1, p3) and np.array_equal(f1, f3)


def test_euclidean_sq_is_sequential_sum(orc):
    a, b = descriptors(2, seed=9)
    exp = 0.0
    for k in range(64):
        d = a[k] - b[k]
        exp += d * d
    assert orc.lib().orc_euclidean_sq(a.ctypes.data, b.ctypes.data, 64) == exp


def test_planted_matches_recovered(orc):
    src, dst, perm = planted_descriptors(500, noise=0.02)
    p, f = orc.associate(src, dst, backwards=True, nthreads=4)
    assert np.array_equal(p, perm.astype(np.int32)) and (f < 0.1).all()
