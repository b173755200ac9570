"""world_size-2 gloo tests (CPU) of the multi-GPU host logic in boofcv_b200/parallel.py: the frame split and the sharded
association orchestration (row offsets, the all-gather of the column partials, merge, validation).

The device steps are replaced by a numpy checker backend with the same contract as the C-ABI calls
(bs_assoc_forward_device / _merge_columns_device / _validate_device); the expected answer is the oracle's single-process
AssociateGreedy.  On the GPU box tests/test_gpu_pipeline.py runs the same orchestration with the real backend."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from boofcv_b200 import parallel  # noqa: E402

DMAX = float(np.finfo(np.float64).max)


def _scores(a, b):
    """DescriptorDistance.euclideanSq for every pair, summed over k in index order (no FMA): bit-exact with the reference."""
    s = np.zeros((len(a), len(b)), np.float64)
    for k in range(a.shape[1]):
        d = a[:, k][:, None] - b[:, k][None, :]
        s = s + d * d
    return s


class NumpyBackend:
    """Contract of the three device steps, restated with numpy on CPU tensors (test double, not a product path)."""

    def __init__(self, max_fit):
        self.max_fit = max_fit

    def forward(self, src, row_offset, dst):
        a, b = src.numpy(), dst.numpy()
        n, m = len(a), len(b)
        w = _scores(a, b)
        pairs = np.full(n, -1, np.int32); fit = np.full(n, self.max_fit, np.float64)
        for i in range(n):
            best = self.max_fit
            for j in range(m):
                if w[i, j] <= best:           # AssociateGreedy.java:80-91: later equal scores replace earlier ones
                    best = w[i, j]; pairs[i] = j
            fit[i] = best
        cmin = np.full(m, np.inf); cfirst = np.full(m, -1, np.int32); clast = np.full(m, -1, np.int32)
        for j in range(m):
            if n == 0:
                continue
            col = w[:, j]
            cmin[j] = col.min()
            hits = np.nonzero(col == cmin[j])[0]
            cfirst[j] = hits[0] + row_offset; clast[j] = hits[-1] + row_offset
        return tuple(torch.from_numpy(x) for x in (pairs, fit, cmin, cfirst, clast))

    def merge(self, all_min, all_first, all_last):
        amin, afirst, alast = all_min.numpy(), all_first.numpy(), all_last.numpy()
        ranks, m = amin.shape
        cmin = np.full(m, np.inf); cfirst = np.full(m, -1, np.int32); clast = np.full(m, -1, np.int32)
        for j in range(m):
            for r in range(ranks):
                if afirst[r, j] < 0:
                    continue
                if cfirst[j] < 0 or amin[r, j] < cmin[j]:
                    cmin[j] = amin[r, j]; cfirst[j] = afirst[r, j]; clast[j] = alast[r, j]
                elif amin[r, j] == cmin[j]:
                    cfirst[j] = min(cfirst[j], afirst[r, j]); clast[j] = max(clast[j], alast[r, j])
        return tuple(torch.from_numpy(x) for x in (cmin, cfirst, clast))

    def validate(self, pairs, fit, row_offset, cmin, cfirst, clast):
        p, f = pairs.numpy().copy(), fit.numpy().copy()
        for i in range(len(p)):
            j = p[i]
            if j < 0:
                continue
            g = row_offset + i
            if not (cmin[j] == f[i] and cfirst[j] == g and clast[j] == g):
                p[i] = -1; f[i] = DMAX
        return torch.from_numpy(p), torch.from_numpy(f)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _case(seed, n, m, dof=8):
    rng = np.random.default_rng(seed)
    dst = rng.normal(size=(m, dof))
    src = rng.normal(size=(n, dof))
    k = min(n, m) // 2
    src[:k] = dst[rng.permutation(m)[:k]] + 0.01 * rng.normal(size=(k, dof))     # planted matches
    if n > 6 and m > 3:
        src[5] = src[4]                    # duplicate source rows: a tie in a column, both must be rejected backwards
        dst[2] = dst[1]                    # duplicate destination rows: forward tie, the later index wins
    return src, dst


def _worker(rank, world, port, cases, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        res = []
        for (seed, n, m, backwards, max_fit) in cases:
            src, dst = _case(seed, n, m)
            b, e = parallel.shard_range(n, rank, world)
            sa = parallel.ShardedAssociate(NumpyBackend(max_fit), backwards)
            # the caller may pass the global index of its first row (shard_range knows it) and save the all-gather of the shard sizes
            pairs, fit = sa.associate(torch.from_numpy(src[b:e].copy()), torch.from_numpy(dst), row_offset=b if seed % 2 else None)
            gp, gf = parallel.gather_matches(pairs, fit)
            res.append((gp, gf))
        if rank == 0:
            out.put(res)
        dist.barrier()
    finally:
        dist.destroy_process_group()


CASES = [(1, 37, 29, True, DMAX), (2, 37, 29, False, DMAX), (3, 64, 80, True, 0.5), (4, 5, 40, True, DMAX), (5, 1, 3, True, DMAX),
         (6, 41, 1, True, DMAX)]


@pytest.mark.timeout(300)
def test_sharded_association_world2_matches_single_process(orc):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, CASES, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for (seed, n, m, backwards, max_fit), (gp, gf) in zip(CASES, res):
        src, dst = _case(seed, n, m)
        op, of = orc.associate(src, dst, max_fit, backwards)
        assert np.array_equal(gp, op), (seed, n, m)
        assert np.array_equal(gf, of), (seed, n, m)


def test_single_process_orchestration_equals_oracle(orc):
    # world size 1 (no process group): forward + validate only
    for (seed, n, m, backwards, max_fit) in CASES:
        src, dst = _case(seed, n, m)
        sa = parallel.ShardedAssociate(NumpyBackend(max_fit), backwards)
        pairs, fit = sa.associate(torch.from_numpy(src), torch.from_numpy(dst))
        op, of = orc.associate(src, dst, max_fit, backwards)
        assert np.array_equal(pairs.numpy(), op) and np.array_equal(fit.numpy(), of)


def test_shard_range_covers_everything_once():
    for total in (0, 1, 7, 8, 1024, 1025):
        for world in (1, 2, 3, 8):
            spans = [parallel.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for (b0, e0), (b1, e1) in zip(spans, spans[1:]):
                assert e0 == b1 and e0 - b0 >= e1 - b1 >= 0
            assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1
    with pytest.raises(ValueError):
        parallel.shard_range(4, 2, 2)
    assert parallel.frame_shard(1024, 3, 8) == (384, 512)
