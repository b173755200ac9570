"""CPU-side checks of the boundary: the C-ABI library builds, loads and exports every symbol include/boofsurf.h declares
(no compute calls: there is no GPU here), the ctypes structs match the header, and host-only helpers behave."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def N():
    from boofcv_b200 import _native
    _native.build()
    return _native


def _header_functions():
    src = open(os.path.join(ROOT, "include", "boofsurf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bs_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound(N):
    names = _header_functions()
    assert len(names) >= 30
    lib = N.lib()
    for n in names:
        assert hasattr(lib, n), "missing export: " + n
        assert n in N.SIGNATURES, "no ctypes signature for " + n
    assert sorted(N.SIGNATURES) == names


def test_defaults_match_reference_configs(N):
    fh = N.FhCfg(); N.lib().bs_fh_cfg_default(C.byref(fh))
    # ConfigFastHessian.java:33-70
    assert (fh.detect_threshold, fh.extract_radius, fh.max_features_per_scale, fh.initial_sample_size, fh.initial_size,
            fh.scales_per_octave, fh.octaves, fh.scale_step_size) == (1.0, 2, -1, 1, 9, 4, 4, 6)
    o = N.OriCfg(); N.lib().bs_ori_cfg_default(C.byref(o), 0)
    assert (o.variant, o.object_radius_to_scale, o.radius, o.sample_period, o.sample_width, o.weight_sigma) == (0, 0.5, 6, 1.0, 6, -1.0)
    N.lib().bs_ori_cfg_default(C.byref(o), 1)
    assert (o.variant, o.radius, o.sample_period) == (1, 8, 0.65) and abs(o.window_size - np.pi / 3) < 1e-15
    d = N.DescCfg(); N.lib().bs_desc_cfg_default(C.byref(d), 1)
    assert (d.variant, d.width_large_grid, d.width_sub_region, d.width_sample, d.use_haar, d.weight_sigma, d.overlap,
            d.sigma_large_grid, d.sigma_sub_region) == (1, 4, 5, 3, 0, 4.5, 2, 2.5, 2.5)
    assert N.lib().bs_version().startswith(b"boofsurf")


def test_matches_bookkeeping_host_helper(N, orc):
    # WrapAssociateGreedy.associate :85-92 + FindUnassociated.checkDestination — pure host code, runs without a GPU
    pairs = np.array([2, -1, 0, 3, -1], np.int32); fit = np.array([0.5, 9, 0.25, 2, 9], np.float64)
    n, m = 5, 6
    ms = np.zeros(n, np.int32); md = np.zeros(n, np.int32); sc = np.zeros(n); us = np.zeros(n, np.int32); ud = np.zeros(m, np.int32)
    nm, nus, nud = C.c_int(), C.c_int(), C.c_int()
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    assert N.lib().bs_assoc_matches(p(pairs), p(fit), n, m, p(ms), p(md), p(sc), C.byref(nm), p(us), C.byref(nus), p(ud), C.byref(nud)) == 0
    matches, o_us, o_ud = orc.matches_from_pairs(pairs, fit, m)
    assert [(int(ms[k]), int(md[k]), float(sc[k])) for k in range(nm.value)] == matches
    assert list(us[:nus.value]) == list(o_us) and list(ud[:nud.value]) == list(o_ud)


def test_no_gpu_means_loud_failure_not_fallback(N):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = N.lib().bs_surf_create(0, None, None, None, 0, C.byref(h))
    assert rc != 0 and not h.value and N.lib().bs_last_error()
    from boofcv_b200 import FactoryDetectDescribe
    with pytest.raises(Exception):
        FactoryDetectDescribe.surfFast(None, None, None, np.float32)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "boofcv_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                s = open(os.path.join(dirpath, f), errors="replace").read()
                assert "bsoracle" not in s and "orc_" not in s and "oracle/" not in s, os.path.join(dirpath, f)
