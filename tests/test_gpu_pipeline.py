"""GPU parity of the drop-in interfaces (DetectDescribePoint / AssociateDescription mirrors) vs the CPU oracle.
Templates: TestWrapDetectDescribeSurf_MT.java:54-101 (same features, here also in the same order and bit for bit),
GenericTestsDetectDescribePoint.java:83-211, TestAssociateGreedy.java:38-102, StandardAssociateDescriptionChecks.java:76-239,
TestAssociateGreedy_MT.java:32-59."""
import numpy as np
import pytest

from synth import blob_image, shifted_pair, descriptors, planted_descriptors

pytestmark = pytest.mark.gpu
DMAX = float(np.finfo(np.float64).max)


@pytest.fixture(scope="module")
def bs():
    import boofcv_b200
    return boofcv_b200


def _same(r, o):
    assert len(r["x"]) == len(o["x"])
    for k in ("x", "y", "scale", "angle", "desc", "white"):
        assert np.array_equal(r[k], o[k]), k


# ------------------------------------------------------------------ detect + describe
@pytest.mark.parametrize("which", ["fast", "stable"])
@pytest.mark.parametrize("shape,seed", [((400, 300), 234), ((640, 480), 3)])
def test_surf_matches_oracle_bit_exact(bs, orc, which, shape, seed):
    img = blob_image(shape[0], shape[1], seed)
    det = getattr(bs.FactoryDetectDescribe, "surfFast" if which == "fast" else "surfStable")(None, None, None, np.float32)
    det.detect(img)
    o = (orc.surf_fast if which == "fast" else orc.surf_stable)(img)
    assert len(o["x"]) > 100
    _same(det.readFrame(0), o)
    det.close()


def test_surf_noise_image_interface_contract(bs, orc):
    # GenericTestsDetectDescribePoint: 100x120 uniform noise, N>5, finite, in bounds, repeatable, sub-image identical
    rng = np.random.default_rng(234)
    img = rng.uniform(0, 100, (120, 100)).astype(np.float32)
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.detect(img)
    n = det.getNumberOfFeatures()
    assert n > 5 and det.hasScale() and det.hasOrientation() and det.getDescriptionType() is bs.BrightFeature
    assert det.createDescription().size() == 64
    first = det.readFrame(0)
    for i in range(n):
        p = det.getLocation(i)
        assert 0 <= p.x < 100 and 0 <= p.y < 120 and (p.x != 0 or p.y != 0)
        assert np.isfinite(det.getRadius(i)) and det.getRadius(i) > 0 and -np.pi <= det.getOrientation(i) <= np.pi
        d = det.getDescription(i)
        assert np.isfinite(d.value).all() and d.size() == 64
    _same(first, orc.surf_fast(img))
    det.detect(img)                                  # two calls -> identical
    _same(det.readFrame(0), first)
    big = np.zeros((150, 140), np.float32); big[20:140, 30:130] = img
    det.detect(big[20:140, 30:130])                  # sub-image -> identical
    _same(det.readFrame(0), first)
    det.close()


def test_surf_u8_input(bs, orc):
    img = np.clip(blob_image(320, 240, 9), 0, 255).astype(np.uint8)
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.uint8)
    det.detect(img)
    o = orc.surf(img)
    assert len(o["x"]) > 20
    _same(det.readFrame(0), o)
    with pytest.raises(ValueError):
        det.detect(img.astype(np.float32))
    det.close()


def test_surf_configs(bs, orc):
    img = blob_image(320, 240, 10)
    cfg = bs.ConfigFastHessian(10.0, 3, -1, 2, 15, 5, 3)
    det = bs.FactoryDetectDescribe.surfStable(cfg, None, None, np.float32)
    det.detect(img)
    o = orc.surf(img, dict(detect_threshold=10.0, extract_radius=3, initial_sample_size=2, initial_size=15,
                           scales_per_octave=5, octaves=3), orc.ORI_SLIDING, orc.DESC_STABILITY)
    _same(det.readFrame(0), o)
    det.close()
    bad = bs.ConfigFastHessian(); bad.extractRadius = 0
    with pytest.raises(ValueError):
        bs.FactoryDetectDescribe.surfFast(bad, None, None, np.float32)


def test_surf_use_haar_end_to_end(bs, orc):
    # the Java config object -> bs_desc_cfg.use_haar -> describe_kernel, through the drop-in interface (integer images too)
    img = blob_image(320, 240, 12)
    d = bs.ConfigSurfDescribe.Speed(); d.useHaar = True
    det = bs.FactoryDetectDescribe.surfFast(None, d, None, np.float32)
    det.detect(img)
    cfg = dict(orc.DESC_SPEED); cfg["use_haar"] = 1
    _same(det.readFrame(0), orc.surf(img, None, orc.ORI_AVERAGE, cfg))
    det.close()
    u8 = np.clip(img, 0, 255).astype(np.uint8)
    det = bs.FactoryDetectDescribe.surfFast(None, d, None, np.uint8)
    det.detect(u8)
    _same(det.readFrame(0), orc.surf(u8, None, orc.ORI_AVERAGE, cfg))
    det.close()


def test_retained_features_frame_at_a_time_tracker(bs, orc):
    # FactoryPointTracker.dda_FH_SURF_Fast (FactoryPointTracker.java:169-175): detect frame t, keep its features, detect frame t+1
    # on the SAME detector, associate previous -> current.  The previous frame's descriptors must survive the second detect.
    a, b = shifted_pair(320, 240)
    c = blob_image(320, 240, 99)
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    oa, ob, oc = orc.surf_fast(a), orc.surf_fast(b), orc.surf_fast(c)
    det.detect(a); prev = det.retainFrame(0)
    det.detect(b); cur = det.retainFrame(0)
    assert prev.count() == (len(oa["x"]), 64) and cur.count() == (len(ob["x"]), 64)
    det.detect(c)                                    # overwrites the detector's pool twice over
    for backwards, basic in ((True, False), (False, False), (True, True)):
        assoc = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, backwards)
        p, f = assoc.associateRetained(prev, cur, surfBasic=basic)
        if basic:                                    # the oracle's matches, in AssociateSurfBasic's order (white class first)
            matches, _, _ = orc.associate_surf_basic(oa["desc"], oa["white"], ob["desc"], ob["white"], 0.1, backwards)
            pn, fn = p.cpu().numpy(), f.cpu().numpy()
            got = sorted((i, int(pn[i]), float(fn[i])) for i in range(len(pn)) if pn[i] >= 0)
            assert got == sorted(matches) and len(matches) > 10
        else:
            op, of = orc.associate(oa["desc"], ob["desc"], 0.1, backwards)
            assert np.array_equal(p.cpu().numpy(), op) and np.array_equal(f.cpu().numpy(), of), (backwards, basic)
        assoc.close()
    _same(det.readFrame(0), oc)
    prev.close(); cur.close(); det.close()


def test_surf_max_features_per_scale(bs, orc):
    # ExampleFeatureSurf.java:62 style configuration: ConfigFastHessian(0, 2, 200, 2, 9, 4, 4); here 25 per scale, batch of 3
    import torch
    from boofcv_b200.synth import blob_images_torch
    cfg = bs.ConfigFastHessian(1.0, 2, 25, 1, 9, 4, 4)
    det = bs.FactoryDetectDescribe.surfFast(cfg, None, None, np.float32)
    frames = blob_images_torch(400, 300, 3, seed=70, device="cuda")
    det.setWorkspace(2, 0)
    det.detectBatch(frames)
    host = frames.cpu().numpy()
    for f in range(3):
        o = orc.surf(host[f], dict(max_features_per_scale=25), orc.ORI_AVERAGE, orc.DESC_SPEED)
        assert 0 < len(o["x"]) <= 25 * 8
        _same(det.readFrame(f), o)
    det.close()


def test_surf_device_batch_equals_single_frames(bs, orc):
    import torch
    from boofcv_b200.synth import blob_images_torch
    frames = blob_images_torch(320, 240, 5, seed=40, device="cuda")
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.setWorkspace(2, 0)                           # 5 frames in chunks of 2: exercises the chunk loop and the carry
    det.detectBatch(frames)
    host = frames.cpu().numpy()
    total = 0
    for f in range(5):
        assert np.array_equal(host[f], blob_image(320, 240, 40 + f))      # torch and numpy generators agree
        r = det.readFrame(f)
        _same(r, orc.surf_fast(host[f]))
        total += len(r["x"])
    assert det.totalFeatures() == total
    v = det.deviceView()
    assert v.total == total and v.batch == 5 and v.dof == 64
    det.close()


def test_surf_device_batch_on_several_streams(bs, orc):
    """More frames than one integral group (32): the batch is spread over three streams that write one pool in frame order.
    Same features as on one stream, and as the oracle finds frame by frame."""
    from boofcv_b200.synth import blob_images_torch
    n = 75
    frames = blob_images_torch(200, 152, n, seed=900, device="cuda")
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.setWorkspace(8, 0)
    res = {}
    for ways in (3, 1, 4):
        det.setWays(ways)
        det.detectBatch(frames)
        res[ways] = [det.readFrame(f) for f in range(n)]
    for f in range(n):
        for k in ("x", "y", "scale", "angle", "desc", "white"):
            assert np.array_equal(res[3][f][k], res[1][f][k]) and np.array_equal(res[4][f][k], res[1][f][k]), (f, k)
    host = frames.cpu().numpy()
    for f in (0, 31, 32, 33, 74):
        _same(res[3][f], orc.surf_fast(host[f]))
    assert sum(len(r["x"]) for r in res[3]) == det.totalFeatures() > n
    with pytest.raises(ValueError):                   # IllegalArgumentException on the Java side
        det.setWays(0)
    with pytest.raises(ValueError):
        det.setWays(5)
    det.close()


def test_surf_pool_overflow_is_reported(bs):
    from boofcv_b200._native import BoofSurfError, BS_ERR_CAPACITY
    img = blob_image(320, 240, 41)
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.setWorkspace(0, 8)
    with pytest.raises(BoofSurfError) as e:
        det.detect(img)
        det.readFrame(0)
    assert e.value.code == BS_ERR_CAPACITY
    det.close()


# ------------------------------------------------------------------ association
def _assoc(bs, src, dst, max_err, backwards):
    a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), max_err, backwards)
    a.setSource(src); a.setDestination(dst); a.associate()
    return a


def test_associate_known_answers(bs):
    # TestAssociateGreedy.java:38-102 (1-d descriptors; thresholds squared because the GPU score is Euclidean^2)
    a = _assoc(bs, np.array([1.0, 2, 3, 4]), np.array([3.0, 4, 1, 40]), 0.5 ** 2, False)
    assert list(a.getPairs()) == [2, -1, 0, 1] and a.getFitQuality()[0] == 0 and a.getFitQuality()[2] == 0
    assert list(_assoc(bs, np.array([1.0, 2, 3, 4]), np.array([3.0, 4, 1.1, 40]), 10.0 ** 2, False).getPairs())[1] == 2
    assert list(_assoc(bs, np.array([1.0, 2, 3, 4]), np.array([3.0, 4, 1.1, 40]), 0.1 ** 2, False).getPairs())[1] == -1
    a = _assoc(bs, np.array([1.0, 2, 3, 8]), np.array([3.0, 4, 1, 10]), 10.0 ** 2, True)
    assert list(a.getPairs()) == [2, -1, 0, 3] and a.getFitQuality()[3] == 4.0 and a.getFitQuality()[1] == DMAX
    # inclusive threshold (StandardAssociateDescriptionChecks.java:158-161) and unique destination (:221-239)
    assert list(_assoc(bs, np.array([1.0]), np.array([1.5]), 0.25, False).getPairs()) == [0]
    a = _assoc(bs, np.array([1.0, 1.001]), np.array([1.0]), DMAX, True)
    assert len(a.getMatches()) == 1 and a.uniqueDestination() and a.uniqueSource()
    assert a.getScoreType() == bs.MatchScoreType.NORM_ERROR


def test_associate_errors_and_empty(bs):
    a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), DMAX, True)
    with pytest.raises(ValueError):
        a.associate()
    a.setSource(np.zeros((0, 64)))
    with pytest.raises(ValueError):
        a.associate()
    a.setDestination(descriptors(5))
    a.associate()
    assert len(a.getMatches()) == 0 and list(a.getUnassociatedDestination()) == [0, 1, 2, 3, 4]
    a.setSource(descriptors(3)); a.setDestination(np.zeros((0, 64))); a.associate()
    assert list(a.getPairs()) == [-1, -1, -1] and list(a.getUnassociatedSource()) == [0, 1, 2]


@pytest.mark.parametrize("n,m", [(200, 200), (1000, 777), (65, 3000), (3000, 65), (1, 1)])
@pytest.mark.parametrize("backwards", [False, True])
def test_associate_random_bit_exact(bs, orc, n, m, backwards):
    src, dst = descriptors(n, seed=1234), descriptors(m, seed=4321)
    a = _assoc(bs, src, dst, DMAX, backwards)
    pairs, fit = orc.associate(src, dst, DMAX, backwards)
    assert np.array_equal(a.getPairs(), pairs) and np.array_equal(a.getFitQuality(), fit)
    matches, us, ud = orc.matches_from_pairs(pairs, fit, m)
    assert [(x.src, x.dst, x.fitScore) for x in a.getMatches()] == matches
    assert np.array_equal(a.getUnassociatedSource(), us) and np.array_equal(a.getUnassociatedDestination(), ud)


def test_associate_ties_and_duplicates(bs, orc):
    # duplicate rows on both sides: forward picks the LAST equal minimum, backward rejects non-unique column minima
    base = descriptors(50, seed=5)
    src = np.concatenate([base, base[:10], base[20:25]])
    dst = np.concatenate([base[::-1], base[5:15]])
    for backwards in (False, True):
        for thr in (DMAX, 0.5, 0.0):
            a = _assoc(bs, src, dst, thr, backwards)
            pairs, fit = orc.associate(src, dst, thr, backwards)
            assert np.array_equal(a.getPairs(), pairs) and np.array_equal(a.getFitQuality(), fit)


def test_associate_planted_and_threshold(bs, orc):
    src, dst, perm = planted_descriptors(1500, seed=77)
    a = _assoc(bs, src, dst, 0.5, True)
    pairs, fit = orc.associate(src, dst, 0.5, True)
    assert np.array_equal(a.getPairs(), pairs) and np.array_equal(a.getFitQuality(), fit)
    assert (a.getPairs() == perm).mean() > 0.9
    a.setMaxScoreThreshold(0.15); a.associate()
    pairs, fit = orc.associate(src, dst, 0.15, True)
    assert 0 < (pairs >= 0).sum() < len(pairs)
    assert np.array_equal(a.getPairs(), pairs) and np.array_equal(a.getFitQuality(), fit)


def test_associate_other_dof(bs, orc):
    for dof in (1, 7, 64, 100, 192):
        src, dst = descriptors(300, dof, 1), descriptors(257, dof, 2)
        a = _assoc(bs, src, dst, DMAX, True)
        pairs, fit = orc.associate(src, dst, DMAX, True)
        assert np.array_equal(a.getPairs(), pairs) and np.array_equal(a.getFitQuality(), fit)


def test_config1_shifted_pair_end_to_end(bs, orc):
    # BASELINE config 1: surfFast on 640x480 + greedy vs a shifted copy
    A, B = shifted_pair(640, 480)
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.detect(A); ra = det.readFrame(0)
    det.detect(B); rb = det.readFrame(0)
    oa, ob = orc.surf_fast(A), orc.surf_fast(B)
    _same(ra, oa); _same(rb, ob)
    a = _assoc(bs, ra["desc"], rb["desc"], DMAX, True)
    pairs, fit = orc.associate(oa["desc"], ob["desc"], DMAX, True)
    assert np.array_equal(a.getPairs(), pairs) and np.array_equal(a.getFitQuality(), fit)
    ms = a.getMatches()
    dx = np.array([ra["x"][k.src] - rb["x"][k.dst] for k in ms]); dy = np.array([ra["y"][k.src] - rb["y"][k.dst] for k in ms])
    good = (np.abs(dx - 7) < 1.0) & (np.abs(dy - 3) < 1.0)
    assert len(ms) > 100 and good.mean() > 0.8
    det.close()


@pytest.mark.parametrize("mode", [0, 1])
def test_sharded_association_device_steps(bs, orc, mode):
    # the three C-ABI steps of a row-sharded association, three shards simulated on one GPU: forward per
    # shard with global row offsets, merge of the stacked column partials, validation per shard == single-process result
    import torch
    from boofcv_b200 import parallel
    rng = np.random.default_rng(77)
    n, m, dof = 301, 257, 64
    dst = rng.normal(size=(m, dof)); src = rng.normal(size=(n, dof))
    src[:100] = dst[rng.permutation(m)[:100]] + 0.01 * rng.normal(size=(100, dof))
    src[7] = src[3]; dst[11] = dst[10]
    op, of = orc.associate(src, dst, 40.0, True)
    assoc = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 40.0, True)
    assoc.setMode(mode)                               # 0: exact fp64 kernels, 1: tensor-core filter + exact re-score
    be = parallel.GpuAssocBackend(assoc)
    d_dst = torch.from_numpy(dst).cuda()
    spans = [parallel.shard_range(n, r, 3) for r in range(3)]
    fwd = [be.forward(torch.from_numpy(src[b:e].copy()).cuda(), b, d_dst) for b, e in spans]
    cmin, cfirst, clast = be.merge(torch.stack([f[2] for f in fwd]), torch.stack([f[3] for f in fwd]), torch.stack([f[4] for f in fwd]))
    pairs, fit = [], []
    for (b, e), f in zip(spans, fwd):
        p, q = be.validate(f[0], f[1], b, cmin, cfirst, clast)
        pairs.append(p.cpu().numpy()); fit.append(q.cpu().numpy())
    torch.cuda.synchronize()
    assert np.array_equal(np.concatenate(pairs), op) and np.array_equal(np.concatenate(fit), of)
    # world size 1 through the orchestrator
    sa = parallel.ShardedAssociate(be, True)
    p, q = sa.associate(torch.from_numpy(src).cuda(), d_dst)
    assert np.array_equal(p.cpu().numpy(), op) and np.array_equal(q.cpu().numpy(), of)
    assoc.close()


def test_surf_host_batch_equals_single_frames(bs, orc):
    # bs_surf_detect_host_batch: 70 host frames (two full sub-batches of 32 and a ragged one) through the two-lane copy/compute
    # pipeline; every frame's list must equal the single-image result, in order
    import torch
    frames = np.stack([blob_image(160, 120, 500 + (f % 5)) for f in range(70)])
    ref = [orc.surf_fast(frames[f]) for f in range(5)]
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    cap = sum(len(ref[f % 5]["x"]) for f in range(70))
    out = bs.HostFeatureBuffers(70, cap + 10, 64, pinned=True)
    det.detectHostBatch(torch.from_numpy(frames).pin_memory(), out)
    assert out.total == cap
    for f in range(70):
        r, o = out.frame(f), ref[f % 5]
        for k in ("x", "y", "scale", "angle", "desc", "white"):
            assert np.array_equal(r[k], o[k]), (f, k)
    # pageable numpy input, strided rows (sub-image), second call on the same handle
    wide = np.zeros((3, 120, 200), np.float32); wide[:, :, 20:180] = frames[:3]
    t = torch.from_numpy(wide)[:, :, 20:180]
    out2 = bs.HostFeatureBuffers(3, cap, 64, pinned=False)
    det.detectHostBatch(t, out2)
    for f in range(3):
        assert np.array_equal(out2.frame(f)["desc"], ref[f]["desc"])
    # capacity error is loud
    small = bs.HostFeatureBuffers(70, 5, 64, pinned=False)
    with pytest.raises(Exception):
        det.detectHostBatch(frames, small)
    out.close(); det.close()


# ------------------------------------------------------------------ AssociateSurfBasic (split by Laplacian sign)
def _desc1(vals):
    d = np.zeros((len(vals), 64)); d[:, 0] = vals
    return d


def _surf_basic(bs, src, sw, dst, dw, max_err, backwards):
    alg = bs.AssociateSurfBasic(bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), max_err, backwards))
    w = bs.WrapAssociateSurfBasic(alg)
    w.setSource((src, np.asarray(sw, np.uint8))); w.setDestination((dst, np.asarray(dw, np.uint8))); w.associate()
    return w


def test_assoc_surf_basic_known_answers(bs):
    # TestAssociateSurfBasic.java:44-162 (threshold squared: the GPU score is Euclidean^2)
    w = _surf_basic(bs, _desc1([10]), [1], _desc1([0, 10]), [1, 0], 400.0, True)
    assert [(m.src, m.dst) for m in w.getMatches()] == [(0, 0)]
    w = _surf_basic(bs, _desc1([10, 12, 5, 2344, 1000]), [1, 1, 0, 0, 0], _desc1([0, 10.1, 13, 0.1, 7]), [1, 1, 1, 0, 0], 400.0, True)
    assert [(m.src, m.dst) for m in w.getMatches()] == [(0, 1), (1, 2), (2, 4)] and all(m.fitScore != 0 for m in w.getMatches())
    assert sorted(w.getUnassociatedSource().tolist()) == [3, 4]
    w = _surf_basic(bs, _desc1([10, 12, 5]), [1, 1, 0], _desc1([0, 10.1, 13, 0.1, 7]), [1, 1, 1, 0, 0], 400.0, True)
    assert len(w.getMatches()) == 3 and len(w.getUnassociatedSource()) == 0 and w.getUnassociatedDestination().tolist() == [0, 3]
    assert w.uniqueSource() and w.uniqueDestination() and w.getScoreType() == bs.MatchScoreType.NORM_ERROR


@pytest.mark.parametrize("n,m", [(300, 280), (3000, 2500), (1, 5), (40, 0)])
@pytest.mark.parametrize("backwards", [False, True])
def test_assoc_surf_basic_matches_oracle(bs, orc, n, m, backwards):
    rng = np.random.default_rng(n * 7 + m)
    src, dst = descriptors(n, seed=11), descriptors(m, seed=12) if m else np.zeros((0, 64))
    sw, dw = rng.integers(0, 2, n), rng.integers(0, 2, m)
    w = _surf_basic(bs, src, sw, dst, dw, 1.5, backwards)
    matches, us, ud = orc.associate_surf_basic(src, sw, dst, dw, 1.5, backwards)
    assert [(x.src, x.dst, x.fitScore) for x in w.getMatches()] == matches
    assert np.array_equal(w.getUnassociatedSource(), us) and np.array_equal(w.getUnassociatedDestination(), ud)


def test_assoc_surf_basic_one_sided_classes_and_device(bs, orc):
    import torch
    src, dst = descriptors(500, seed=21), descriptors(400, seed=22)
    sw = np.ones(500, np.uint8); dw = np.zeros(400, np.uint8); dw[:7] = 1          # few white destinations, no dark sources
    w = _surf_basic(bs, src, sw, dst, dw, DMAX, True)
    matches, us, ud = orc.associate_surf_basic(src, sw, dst, dw, DMAX, True)
    assert [(x.src, x.dst, x.fitScore) for x in w.getMatches()] == matches and np.array_equal(w.getUnassociatedSource(), us)
    # device-resident call on real SURF features of a shifted pair: same pairs as the host call
    a, b = shifted_pair(320, 240)
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.detect(a); ra = det.readFrame(0)
    det.detect(b); rb = det.readFrame(0)
    alg = bs.AssociateSurfBasic(bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, True))
    alg.setSrc(ra); alg.setDst(rb); alg.associate()
    matches, us, ud = orc.associate_surf_basic(ra["desc"], ra["white"], rb["desc"], rb["white"], 0.1, True)
    assert [(x.src, x.dst, x.fitScore) for x in alg.getMatches()] == matches and len(matches) > 10
    t = lambda x, dt: torch.as_tensor(np.ascontiguousarray(x), dtype=dt, device="cuda")
    p, f = alg.associateDevice(t(ra["desc"], torch.float64), t(ra["white"], torch.uint8), t(rb["desc"], torch.float64), t(rb["white"], torch.uint8))
    assert np.array_equal(p.cpu().numpy(), alg.getPairs()) and np.array_equal(f.cpu().numpy(), alg.getFitQuality())


# ------------------------------------------------------------------ colour SURF (surfColorFast / surfColorStable)
def _planar(w, h, nb=3, seed=234):
    return np.stack([blob_image(w, h, seed + 17 * b) for b in range(nb)])


@pytest.mark.parametrize("which", ["fast", "stable"])
@pytest.mark.parametrize("nb,shape", [(3, (320, 240)), (2, (400, 300)), (1, (200, 160))])
def test_surf_color_matches_oracle_bit_exact(bs, orc, which, nb, shape):
    bands = _planar(shape[0], shape[1], nb)
    it = bs.ImageType.pl(nb, np.float32)
    det = (bs.FactoryDetectDescribe.surfColorFast if which == "fast" else bs.FactoryDetectDescribe.surfColorStable)(None, None, None, it)
    det.detect(bands)
    o = (orc.surf_color_fast if which == "fast" else orc.surf_color_stable)(bands)
    r = det.readFrame(0)
    assert len(o["x"]) > 30 and r["desc"].shape[1] == 64 * nb
    _same(r, o)
    assert det.getRadius(0) == o["scale"][0] and det.createDescription().size() == 64 * nb
    assert det.getDescription(1).white == bool(o["white"][1])


def test_surf_color_u8_and_errors(bs, orc):
    bands = np.clip(_planar(320, 240, 3, seed=5), 0, 255).astype(np.uint8)
    det = bs.FactoryDetectDescribe.surfColorFast(None, None, None, bs.ImageType.pl(3, np.uint8))
    det.detect(bands)
    _same(det.readFrame(0), orc.surf_color_fast(bands))
    with pytest.raises(ValueError):
        det.detect(bands[:2])                                    # SurfPlanar_to_DetectDescribePoint.java:59-61
    with pytest.raises(ValueError):
        bs.FactoryDetectDescribe.surfColorFast(None, None, None, bs.ImageType.single(np.float32))   # "Image type not supported"
    grey = bs.FactoryDetectDescribe.surfFast(None, None, None, np.uint8)
    grey.detect(bands[0])                                        # the grey handle is unaffected by the colour one


def test_surf_color_device_batch(bs, orc):
    import torch
    frames = np.stack([_planar(320, 240, 3, seed=s) for s in (1, 2, 3, 4, 5)])
    det = bs.FactoryDetectDescribe.surfColorFast(None, None, None, bs.ImageType.pl(3, np.float32))
    det.setWorkspace(2, 0)                                       # several chunks and a partial one
    det.detectBatch(torch.as_tensor(frames, device="cuda"))
    for f in range(len(frames)):
        _same(det.readFrame(f), orc.surf_color_fast(frames[f]))


# ------------------------------------------------------------------ detect -> associate on the device (stereo / frame-to-frame)
@pytest.mark.parametrize("surf_basic", [False, True])
def test_associate_frames_on_device(bs, orc, surf_basic):
    import torch
    a0, b0 = shifted_pair(320, 240)
    a1, b1 = shifted_pair(320, 240, seed=77)
    frames = torch.as_tensor(np.stack([a0, b0, a1, b1]), device="cuda")
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.detectBatch(frames)
    assoc = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, True)
    for fa, fb, ia, ib in ((0, 1, a0, b0), (2, 3, a1, b1), (3, 0, b1, a0)):
        p, f = assoc.associateFrames(det, fa, fb, surfBasic=surf_basic)
        oa, ob = orc.surf_fast(ia), orc.surf_fast(ib)
        if surf_basic:
            matches, _, _ = orc.associate_surf_basic(oa["desc"], oa["white"], ob["desc"], ob["white"], 0.1, True)
            got = {(i, int(d)): float(s) for i, (d, s) in enumerate(zip(p.cpu().numpy(), f.cpu().numpy())) if d >= 0}
            assert got == {(s, d): sc for s, d, sc in matches} and (fa == 3 or len(matches) > 10)
        else:
            pairs, fit = orc.associate(oa["desc"], ob["desc"], 0.1, True)
            assert np.array_equal(p.cpu().numpy(), pairs) and np.array_equal(f.cpu().numpy(), fit)
            if fa < 3:
                assert (pairs >= 0).sum() > 10


@pytest.mark.parametrize("mode", [-1, 1])
def test_associate_frame_pairs_batched(bs, orc, mode):
    """bs_assoc_frames_batch_device: a list of frame pairs queued back to back == the same pairs one call at a time == the oracle."""
    import torch
    a0, b0 = shifted_pair(320, 240)
    a1, b1 = shifted_pair(320, 240, seed=77)
    imgs = [a0, b0, a1, b1]
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.detectBatch(torch.as_tensor(np.stack(imgs), device="cuda"))
    assoc = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), 0.1, True)
    assoc.setMode(mode)
    src, dst = [0, 2, 3, 1], [1, 3, 0, 1]
    p, f, off = assoc.associateFramesBatch(det, src, dst)
    assert off[0] == 0 and len(off) == 5 and off[-1] == p.numel() == f.numel()
    feats = [orc.surf_fast(im) for im in imgs]
    for k, (fa, fb) in enumerate(zip(src, dst)):
        assert off[k + 1] - off[k] == det.countFrame(fa)
        pk, fk = p[off[k]:off[k + 1]].cpu().numpy(), f[off[k]:off[k + 1]].cpu().numpy()
        p1, f1 = assoc.associateFrames(det, fa, fb)
        assert np.array_equal(pk, p1.cpu().numpy()) and np.array_equal(fk, f1.cpu().numpy())
        pairs, fit = orc.associate(feats[fa]["desc"], feats[fb]["desc"], 0.1, True)
        assert np.array_equal(pk, pairs) and np.array_equal(fk, fit)
    # an empty list and a capacity that is too small
    p, f, off = assoc.associateFramesBatch(det, [], [])
    assert p.numel() == 0 and list(off) == [0]
    with pytest.raises(Exception):
        assoc.associateFramesBatch(det, [0], [9])
    det.close(); assoc.close()
