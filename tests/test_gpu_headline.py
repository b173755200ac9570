"""Parity at the sizes the bench numbers are quoted on: the FULL pipeline against the CPU oracle,
bit for bit, on 1920x1080 batches shaped like bench.py's (chunk 32, three streams, the host-batch lanes), on a 3840x2160
surfStable frame, with the sliding-window walk scratch forced small (the in-kernel walk branch), and the tensor-core
association against the oracle's greedy association at 20k x 20k, single call and row-sharded.
Templates: TestWrapDetectDescribeSurf_MT.java:54-101, TestAssociateGreedy_MT.java:32-59 (same results as the single-thread
reference, at a size where the work really is split)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
NT = max(1, os.cpu_count() or 1)
DMAX = float(np.finfo(np.float64).max)


@pytest.fixture(scope="module")
def bs():
    import boofcv_b200
    return boofcv_b200


def _same(r, o, what=""):
    assert len(r["x"]) == len(o["x"]), (what, len(r["x"]), len(o["x"]))
    for k in ("x", "y", "scale", "angle", "desc", "white"):
        assert np.array_equal(r[k], o[k]), (what, k)


def test_surf_fast_1080p_bench_shaped_batch(bs, orc):
    # 70 frames of 1920x1080 (three distinct ones, cycled): chunk 32 + three ways = three integral groups on three streams,
    # exactly bench.py's device-resident configuration; every frame's list == the oracle's for that frame
    import torch
    from boofcv_b200.synth import blob_images_torch
    distinct = blob_images_torch(1920, 1080, 3, seed=234, device="cuda")
    ref = [orc.surf_fast(distinct[i].cpu().numpy(), nthreads=NT) for i in range(3)]
    assert all(len(r["x"]) > 5000 for r in ref)
    B = 70
    frames = distinct[torch.arange(B, device="cuda") % 3].contiguous()
    det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32)
    det.setWorkspace(32, 0)
    det.setWays(3)
    det.detectBatch(frames)
    assert det.totalFeatures() == sum(len(ref[f % 3]["x"]) for f in range(B))
    for f in (0, 1, 2, 31, 32, 33, 63, 64, 65, 69):
        _same(det.readFrame(f), ref[f % 3], "frame %d" % f)
    # counts of every frame (cheap), so that no chunk/way boundary goes unchecked
    for f in range(B):
        assert det.countFrame(f) == len(ref[f % 3]["x"]), f
    # the host-facing call (what bench.py's e2e times): pinned frames in, pinned feature lists out, six lanes of 16 frames
    host = frames.cpu().pin_memory()
    cap = det.totalFeatures() + 1000
    out = bs.HostFeatureBuffers(B, cap, 64, pinned=True)
    det.detectHostBatch(host, out)
    assert out.total == sum(len(ref[f % 3]["x"]) for f in range(B))
    for f in (0, 15, 16, 17, 47, 48, 69):
        b, e = out.frame_offset[f], out.frame_offset[f + 1]
        o = ref[f % 3]
        assert e - b == len(o["x"])
        assert np.array_equal(out.xy[b:e, 0], o["x"]) and np.array_equal(out.xy[b:e, 1], o["y"])
        assert np.array_equal(out.scale[b:e], o["scale"]) and np.array_equal(out.angle[b:e], o["angle"])
        assert np.array_equal(out.desc[b:e], o["desc"]) and np.array_equal(out.white[b:e].astype(bool), o["white"])
    out.close()
    det.close()


def test_surf_stable_2160p_frame(bs, orc):
    # BASELINE configs[2]: surfStable (4 octaves, sliding-window orientation, DescribePointSurfMod) on 3840x2160
    from boofcv_b200.synth import blob_image
    img = blob_image(3840, 2160, 77)
    o = orc.surf_stable(img, nthreads=NT)
    assert len(o["x"]) > 20000
    det = bs.FactoryDetectDescribe.surfStable(None, None, None, np.float32)
    det.detect(img)
    _same(det.readFrame(0), o, "2160p stable")
    det.close()


def test_surf_stable_walk_scratch_overflow_branch(bs, orc):
    # the walk scratch holds 50 features: the remaining ~900 sliding-window walks run inside describe_kernel (bs_capi.cu
    # reserve_walk / describe_kernel's walk branch) and must give the same angles and descriptors
    from boofcv_b200.synth import blob_image
    img = blob_image(640, 480, 5)
    o = orc.surf_stable(img)
    assert len(o["x"]) > 200
    det = bs.FactoryDetectDescribe.surfStable(None, None, None, np.float32)
    det.setWalkScratch(50)
    det.detect(img)
    _same(det.readFrame(0), o, "small walk scratch")
    det.detect(img)
    _same(det.readFrame(0), o, "small walk scratch, second call")
    det.close()
    # batch of 3 with the small scratch (the scratch is per launch, the features of a chunk share it)
    import torch
    det = bs.FactoryDetectDescribe.surfStable(None, None, None, np.float32)
    det.setWalkScratch(300)
    frames = torch.from_numpy(np.stack([img, blob_image(640, 480, 6), img])).cuda()
    det.detectBatch(frames)
    _same(det.readFrame(0), o, "batch 0"); _same(det.readFrame(2), o, "batch 2")
    _same(det.readFrame(1), orc.surf_stable(blob_image(640, 480, 6)), "batch 1")
    det.close()


def _unit(rng, n):
    x = rng.normal(size=(n, 64))
    return x / np.linalg.norm(x, axis=1, keepdims=True)


def test_mode1_association_20k_vs_oracle_and_sharded(bs, orc):
    # 20 000 x 20 480 unit descriptors, 12 000 planted noisy matches, duplicates on both sides; oracle = the reference's loops in
    # fp64 on all host threads.  Then the same problem row-sharded three ways through the C-ABI steps.
    import torch
    from boofcv_b200 import parallel
    rng = np.random.default_rng(2024)
    n, m = 20000, 20480
    dst = _unit(rng, m); src = _unit(rng, n)
    k = 12000
    src[:k] = dst[rng.permutation(m)[:k]] + 0.02 * rng.normal(size=(k, 64))
    src[:k] /= np.linalg.norm(src[:k], axis=1, keepdims=True)
    src[15000] = src[3]; dst[20001] = dst[17]                   # duplicate source row, duplicate destination row
    for backwards in (True, False):
        op, of = orc.associate(src, dst, DMAX, backwards, nthreads=NT)
        a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), DMAX, backwards)
        a.setMode(1)
        p, f = a.associateDevice(torch.from_numpy(src).cuda(), torch.from_numpy(dst).cuda())
        assert np.array_equal(p.cpu().numpy(), op) and np.array_equal(f.cpu().numpy(), of)
        _, _, (runs, fallbacks) = a.filterScores(src[:1], dst[:1], want_scores=False)
        assert runs >= 1 and fallbacks == 0
        if backwards:
            assert (op >= 0).sum() >= k // 2
            be = parallel.GpuAssocBackend(a)
            d_dst = torch.from_numpy(dst).cuda()
            spans = [parallel.shard_range(n, r, 3) for r in range(3)]
            fwd = [be.forward(torch.from_numpy(src[b:e].copy()).cuda(), b, d_dst) for b, e in spans]
            cmin, cfirst, clast = be.merge(torch.stack([x[2] for x in fwd]), torch.stack([x[3] for x in fwd]), torch.stack([x[4] for x in fwd]))
            ps, fs = [], []
            for (b, e), x in zip(spans, fwd):
                pp, qq = be.validate(x[0], x[1], b, cmin, cfirst, clast)
                ps.append(pp.cpu().numpy()); fs.append(qq.cpu().numpy())
            assert np.array_equal(np.concatenate(ps), op) and np.array_equal(np.concatenate(fs), of)
        a.close()


def test_second_device_in_one_process(bs, orc):
    # function attributes and SM counts are per device; handles on cuda:1 after cuda:0 in ONE process
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from boofcv_b200.synth import blob_image
    img = blob_image(400, 300, 234)
    o = orc.surf_fast(img)
    rng = np.random.default_rng(3)
    src, dst = _unit(rng, 2100), _unit(rng, 2300)
    op, of = orc.associate(src, dst, DMAX, True, nthreads=NT)
    for dev in (0, 1):
        det = bs.FactoryDetectDescribe.surfFast(None, None, None, np.float32, device=dev)
        det.detect(img)
        _same(det.readFrame(0), o, "device %d" % dev)
        det.close()
        for mode in (0, 1):
            a = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), DMAX, True, device=dev)
            a.setMode(mode)
            p, f = a.associateDevice(torch.from_numpy(src).cuda(dev), torch.from_numpy(dst).cuda(dev))
            assert np.array_equal(p.cpu().numpy(), op) and np.array_equal(f.cpu().numpy(), of), (dev, mode)
            a.close()
