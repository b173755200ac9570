"""bs_assoc_forward_static — the allocation-free forward association (static device scratch), through the C ABI.

No reference counterpart (see include/boofsurf.h).  Checked against the existing
entry point bs_assoc_run_device (exact mode, no backward validation) bit for bit — the tolerance every association test here
uses — on two calls in a row that share the static scratch, the second smaller than the first so that stale contents of the
first would show.  Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DMAX = float(np.finfo(np.float64).max)


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import boofcv_b200
    from boofcv_b200 import _native
    return torch, boofcv_b200, _native


def _unit(rng, n, dof=64):
    x = rng.normal(size=(n, dof))
    return x / np.linalg.norm(x, axis=1, keepdims=True)


def _static(env, src, dst, max_fit, stream=None):
    torch, _, N = env
    n, dof = src.shape
    m = dst.shape[0]
    pairs = torch.full((n,), -7, dtype=torch.int32, device=src.device)
    fit = torch.full((n,), -7.0, dtype=torch.float64, device=src.device)
    st = C.c_void_p(stream.cuda_stream if stream is not None else torch.cuda.current_stream(src.device).cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr() if t.numel() else 0)
    N.check(N.lib().bs_assoc_forward_static(src.device.index or 0, p(src), n, p(dst), m, dof, float(max_fit), p(pairs), p(fit), st))
    return pairs, fit


def _existing(env, src, dst, max_fit):
    _, bs, _ = env
    g = bs.FactoryAssociation.greedy(bs.ScoreAssociateEuclideanSq_F64(), max_fit, False)
    g.setMode(0)                                                   # exact fp64 kernels: the path the static variant mirrors
    pairs, fit = g.associateDevice(src, dst)
    pairs, fit = pairs.clone(), fit.clone()
    g.close()
    return pairs, fit


def test_static_forward_matches_existing_entry_point_on_two_calls_in_a_row(env, orc):
    torch, _, _ = env
    rng = np.random.default_rng(77)
    cases = []
    for n, m, max_fit in ((3000, 5000, DMAX), (700, 901, 0.9)):       # second call: fewer rows and columns than the first
        dst = _unit(rng, m); src = _unit(rng, n)
        k = n // 2
        src[:k] = dst[rng.permutation(m)[:k]] + 0.02 * rng.normal(size=(k, 64))
        dst[m - 1] = dst[3]                                           # an exact tie: the later index must win
        cases.append((torch.from_numpy(src).cuda(), torch.from_numpy(dst).cuda(), max_fit, src, dst))
    expect = [_existing(env, s, d, mf) for s, d, mf, _, _ in cases]
    torch.cuda.synchronize()
    got = [_static(env, s, d, mf) for s, d, mf, _, _ in cases]        # back to back on one stream, sharing the scratch
    torch.cuda.synchronize()
    for (gp, gf), (ep, ef), (_, _, mf, src, dst) in zip(got, expect, cases):
        assert torch.equal(gp, ep) and torch.equal(gf, ef)
        op, of = orc.associate(src, dst, mf, False)
        assert np.array_equal(gp.cpu().numpy(), op) and np.array_equal(gf.cpu().numpy(), of)
    assert (got[0][0] >= 0).all() and (got[1][0] < 0).any() and (got[1][0] >= 0).any()


def test_static_forward_allocates_no_device_memory_and_runs_on_a_side_stream(env):
    torch, _, _ = env
    rng = np.random.default_rng(5)
    src = torch.from_numpy(_unit(rng, 1500, 32)).cuda(); dst = torch.from_numpy(_unit(rng, 2100, 32)).cuda()
    ep, ef = _existing(env, src, dst, DMAX)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    _static(env, src, dst, DMAX, side)                                # first call on this device: set-up (symbol address, attribute)
    torch.cuda.synchronize()
    n, m = src.shape[0], dst.shape[0]
    pairs = torch.empty(n, dtype=torch.int32, device="cuda"); fit = torch.empty(n, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    N = env[2]
    p = lambda t: C.c_void_p(t.data_ptr())
    N.check(N.lib().bs_assoc_forward_static(0, p(src), n, p(dst), m, 32, DMAX, p(pairs), p(fit), C.c_void_p(side.cuda_stream)))
    side.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0                                             # the call took no device memory
    assert torch.equal(pairs, ep) and torch.equal(fit, ef)


def test_static_forward_limits_and_edge_cases(env):
    torch, _, N = env
    L = N.lib()
    one = torch.zeros((1, 64), dtype=torch.float64, device="cuda")
    pairs = torch.zeros(4, dtype=torch.int32, device="cuda"); fit = torch.zeros(4, dtype=torch.float64, device="cuda")
    p = lambda t: C.c_void_p(t.data_ptr())
    # beyond the stated maximum: rejected with a clear error before anything is touched
    rc = L.bs_assoc_forward_static(0, p(one), 131072 + 1, p(one), 1, 64, DMAX, p(pairs), p(fit), None)
    assert rc == N.BS_ERR_ARG and b"BS_ASSOC_STATIC_MAX_ROWS" in L.bs_last_error()
    assert L.bs_assoc_forward_static(0, p(one), 1, p(one), 1, 300, DMAX, p(pairs), p(fit), None) == N.BS_ERR_ARG
    assert L.bs_assoc_forward_static(99, p(one), 1, p(one), 1, 64, DMAX, p(pairs), p(fit), None) == N.BS_ERR_ARG
    assert L.bs_assoc_forward_static(0, None, 1, p(one), 1, 64, DMAX, p(pairs), p(fit), None) == N.BS_ERR_ARG
    assert L.bs_assoc_forward_static(0, None, 0, None, 0, 64, DMAX, None, None, None) == 0             # no source rows: nothing to do
    # no destination: every source unmatched with the untouched running best (AssociateGreedy.java:76-95)
    src = torch.from_numpy(_unit(np.random.default_rng(1), 4)).cuda()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert L.bs_assoc_forward_static(0, p(src), 4, None, 0, 64, 0.5, p(pairs), p(fit), st) == 0
    torch.cuda.synchronize()
    assert pairs.tolist() == [-1] * 4 and fit.tolist() == [0.5] * 4
    # a single pair, threshold inclusive
    assert L.bs_assoc_forward_static(0, p(src), 1, p(src), 1, 64, 0.0, p(pairs), p(fit), st) == 0
    torch.cuda.synchronize()
    assert pairs[0].item() == 0 and fit[0].item() == 0.0
