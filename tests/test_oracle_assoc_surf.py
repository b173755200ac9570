"""Pins the oracle's AssociateSurfBasic with the reference's known-answer tests.

  boofcv-feature/src/test/java/boofcv/alg/feature/associate/TestAssociateSurfBasic.java:44-162
      (checkAssociateByIntensity, basicAssociation, checkUnassociated; greedy(ScoreAssociateEuclidean_F64, 20, true))
Descriptors there are 64-d with only value[0] set, so Euclidean^2 with the threshold squared gives the same answers.
"""
import numpy as np


def _desc(vals):
    d = np.zeros((len(vals), 64)); d[:, 0] = vals
    return d


def test_never_across_laplacian_signs(orc):
    m, us, ud = orc.associate_surf_basic(_desc([10]), [1], _desc([0, 10]), [1, 0], 20.0 ** 2, True)
    assert len(m) == 1 and m[0][1] == 0          # the worse descriptor with the same sign wins


def test_basic_association(orc):
    src, sw = _desc([10, 12, 5, 2344, 1000]), [1, 1, 0, 0, 0]
    dst, dw = _desc([0, 10.1, 13, 0.1, 7]), [1, 1, 1, 0, 0]
    m, us, ud = orc.associate_surf_basic(src, sw, dst, dw, 20.0 ** 2, True)
    assert [(a, b) for a, b, _ in m] == [(0, 1), (1, 2), (2, 4)]
    assert all(f != 0 for _, _, f in m)
    assert sorted(us.tolist()) == [3, 4]


def test_check_unassociated(orc):
    src, sw = _desc([10, 12, 5]), [1, 1, 0]
    dst, dw = _desc([0, 10.1, 13, 0.1, 7]), [1, 1, 1, 0, 0]
    m, us, ud = orc.associate_surf_basic(src, sw, dst, dw, 20.0 ** 2, True)
    assert len(m) == 3 and len(us) == 0 and ud.tolist() == [0, 3]


def test_order_and_empty_lists(orc):
    # matches: white class in ascending source index, then the other class (:94-121)
    src, sw = _desc([1, 2, 3, 4]), [0, 1, 0, 1]
    dst, dw = _desc([4, 3, 2, 1]), [1, 0, 1, 0]
    m, us, ud = orc.associate_surf_basic(src, sw, dst, dw)
    assert [(a, b) for a, b, _ in m] == [(1, 2), (3, 0), (0, 3), (2, 1)]
    # nothing is reported when either list is empty (:89-92), not even unassociated sources
    m, us, ud = orc.associate_surf_basic(src, sw, np.zeros((0, 64)), [])
    assert m == [] and len(us) == 0 and len(ud) == 0
    # a class without destinations: its sources are unassociated
    m, us, ud = orc.associate_surf_basic(src, sw, _desc([2.1]), [1], backwards=True)
    assert [(a, b) for a, b, _ in m] == [(1, 0)] and us.tolist() == [3, 0, 2]
