"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes wrapper around oracle/libbsoracle.so, the CPU restatement of BoofCV's SURF path (see oracle/orc_surf.h for the
reference file:line each function follows).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product package (boofcv_b200/) never does.

Parity status: the reference is Java and cannot run in this image, so end-to-end outputs are "parity unpinned";
the stage functions are pinned by the reference's own known-answer / differential tests, re-expressed in
tests/test_oracle_*.py.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

FH_DEFAULT = dict(detect_threshold=1.0, extract_radius=2, max_features_per_scale=-1, initial_sample_size=1,
                  initial_size=9, scales_per_octave=4, octaves=4, scale_step_size=6)
ORI_AVERAGE = dict(variant=0, object_radius_to_scale=0.5, radius=6, sample_period=1.0, sample_width=6,
                   weight_sigma=-1.0, window_size=np.pi / 3.0)
ORI_SLIDING = dict(variant=1, object_radius_to_scale=0.5, radius=8, sample_period=0.65, sample_width=6,
                   weight_sigma=-1.0, window_size=np.pi / 3.0)
DESC_SPEED = dict(variant=0, width_large_grid=4, width_sub_region=5, width_sample=3, use_haar=0, weight_sigma=4.5,
                  overlap=2, sigma_large_grid=2.5, sigma_sub_region=2.5)
DESC_STABILITY = dict(DESC_SPEED, variant=1)


def build(force=False):
    so = os.path.join(_HERE, "libbsoracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("orc_surf.cpp", "orc_capi.cpp", "orc_surf.h", "orc_math.h", "Makefile")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libbsoracle.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        for name, res in (("orc_atan2", C.c_double), ("orc_sin", C.c_double), ("orc_cos", C.c_double),
                          ("orc_block_unsafe_f32", C.c_float), ("orc_block_zero_f32", C.c_float),
                          ("orc_poly_peak", C.c_float), ("orc_euclidean_sq", C.c_double)):
            getattr(_LIB, name).restype = res
        _LIB.orc_laplace_sign_f32.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_double]
        _LIB.orc_atan2.argtypes = [C.c_double, C.c_double]
        _LIB.orc_sin.argtypes = [C.c_double]
        _LIB.orc_cos.argtypes = [C.c_double]
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _img(a, dtype):
    a = np.asarray(a)
    assert a.dtype == dtype and a.ndim == 2 and a.strides[1] == a.itemsize, (a.dtype, a.shape, a.strides)
    return a, a.strides[0] // a.itemsize


def _fh(cfg):
    c = dict(FH_DEFAULT, **(cfg or {}))
    iv = np.array([c["extract_radius"], c["max_features_per_scale"], c["initial_sample_size"], c["initial_size"],
                   c["scales_per_octave"], c["octaves"], c["scale_step_size"]], dtype=np.int32)
    return iv, C.c_float(c["detect_threshold"])


def _ori(cfg):
    c = dict(ORI_AVERAGE, **(cfg or {}))
    return np.array([c["variant"], c["object_radius_to_scale"], c["radius"], c["sample_period"], c["sample_width"],
                     c["weight_sigma"], c["window_size"]], dtype=np.float64)


def _desc(cfg):
    c = dict(DESC_SPEED, **(cfg or {}))
    return np.array([c["variant"], c["width_large_grid"], c["width_sub_region"], c["width_sample"], c["use_haar"],
                     c["weight_sigma"], c["overlap"], c["sigma_large_grid"], c["sigma_sub_region"]], dtype=np.float64)


# ------------------------------------------------------------------ math
def atan2(y, x):
    y = np.ascontiguousarray(y, np.float64); x = np.ascontiguousarray(x, np.float64)
    out = np.empty_like(y)
    lib().orc_atan2_v(_p(y), _p(x), _p(out), C.c_long(y.size))
    return out


def sincos(a):
    a = np.ascontiguousarray(a, np.float64)
    s = np.empty_like(a); c = np.empty_like(a)
    lib().orc_sincos_v(_p(a), _p(s), _p(c), C.c_long(a.size))
    return s, c


# ------------------------------------------------------------------ integral image
def integral(img):
    """GIntegralImageOps.transform: F32->F32, U8->S32, S32->S32.  Accepts row-strided views (sub-images)."""
    img = np.asarray(img)
    if img.dtype == np.float32:
        a, st = _img(img, np.float32); out = np.empty(a.shape, np.float32)
        lib().orc_integral_f32(_p(a), st, _p(out), a.shape[1], a.shape[1], a.shape[0])
    elif img.dtype == np.uint8:
        a, st = _img(img, np.uint8); out = np.empty(a.shape, np.int32)
        lib().orc_integral_u8(_p(a), st, _p(out), a.shape[1], a.shape[1], a.shape[0])
    elif img.dtype == np.int32:
        a, st = _img(img, np.int32); out = np.empty(a.shape, np.int32)
        lib().orc_integral_s32(_p(a), st, _p(out), a.shape[1], a.shape[1], a.shape[0])
    else:
        raise TypeError("Unknown input type: %s" % img.dtype)
    return out


def block_unsafe(ii, x0, y0, x1, y1):
    a, st = _img(ii, ii.dtype)
    f = lib().orc_block_unsafe_f32 if a.dtype == np.float32 else lib().orc_block_unsafe_s32
    return f(_p(a), st, a.shape[1], a.shape[0], x0, y0, x1, y1)


def block_zero(ii, x0, y0, x1, y1):
    a, st = _img(ii, ii.dtype)
    f = lib().orc_block_zero_f32 if a.dtype == np.float32 else lib().orc_block_zero_s32
    return f(_p(a), st, a.shape[1], a.shape[0], x0, y0, x1, y1)


def kernel_deriv(which, size):
    out = np.zeros(21, np.int32)
    lib().orc_kernel_deriv({"xx": 0, "yy": 1, "xy": 2}[which], size, _p(out))
    n = out[0]
    return [tuple(out[1 + 5 * i:6 + 5 * i]) for i in range(n)]


# ------------------------------------------------------------------ Fast-Hessian intensity / NMS / detector
def hessian(ii, skip, size, naive=False, nthreads=1):
    a, st = _img(ii, ii.dtype)
    h, w = a.shape
    out = np.zeros((h // skip, w // skip), np.float32)
    f = lib().orc_hessian_f32 if a.dtype == np.float32 else lib().orc_hessian_s32
    f(_p(a), st, w, h, skip, size, _p(out), int(naive), nthreads)
    return out


def nonmax(inten, radius, thresh, border, naive=False):
    a, st = _img(inten, np.float32)
    h, w = a.shape
    cap = max(1, w * h)
    xy = np.zeros((cap, 2), np.int16)
    n = lib().orc_nonmax(_p(a), st, w, h, radius, C.c_float(thresh), border, int(naive), _p(xy), cap)
    return xy[:n].copy()


def poly_peak(lo, mid, hi):
    return lib().orc_poly_peak(C.c_float(lo), C.c_float(mid), C.c_float(hi))


def fh_detect(ii, cfg=None, nthreads=1, max_out=None):
    """FastHessianFeatureDetector.detect on an integral image.  Returns dict(x, y, scale, octave, level, px, py, response)
    in the reference's single-thread emission order."""
    a, st = _img(ii, ii.dtype)
    h, w = a.shape
    iv, thr = _fh(cfg)
    cap = max_out or max(1024, w * h // 8)
    xys = np.zeros((cap, 3), np.float64); meta = np.zeros((cap, 4), np.int32); resp = np.zeros(cap, np.float32)
    f = lib().orc_fh_detect_f32 if a.dtype == np.float32 else lib().orc_fh_detect_s32
    n = f(_p(a), st, w, h, _p(iv), thr, cap, _p(xys), _p(meta), _p(resp), nthreads)
    assert n <= cap
    return dict(x=xys[:n, 0].copy(), y=xys[:n, 1].copy(), scale=xys[:n, 2].copy(), octave=meta[:n, 0].copy(),
                level=meta[:n, 1].copy(), px=meta[:n, 2].copy(), py=meta[:n, 3].copy(), response=resp[:n].copy())


# ------------------------------------------------------------------ weights, gradient, containment
def gaussian2d(sigma, radius):
    out = np.zeros((2 * radius + 1) ** 2, np.float64)
    lib().orc_gaussian(0, C.c_double(sigma), radius, _p(out))
    return out.reshape(2 * radius + 1, -1)


def gaussian_width(sigma, width):
    out = np.zeros(width * width, np.float64)
    lib().orc_gaussian(1, C.c_double(sigma), width, _p(out))
    return out.reshape(width, width)


def gradient(ii, width, xy, safe=True, haar=False):
    a, st = _img(ii, np.float32)
    xy = np.ascontiguousarray(xy, np.int32)
    out = np.zeros((len(xy), 2), np.float64)
    lib().orc_gradient_f32(_p(a), st, a.shape[1], a.shape[0], C.c_double(width), _p(xy), len(xy), int(safe), _p(out), int(haar))
    return out


def is_inside_rotated(w, h, X, Y, radius_regions, kernel_size, scale, c, s):
    return bool(lib().orc_is_inside_rotated(w, h, C.c_double(X), C.c_double(Y), radius_regions, kernel_size,
                                            C.c_double(scale), C.c_double(c), C.c_double(s)))


def is_inside_rect(w, h, tl_x, tl_y, region_size, sample_size):
    return bool(lib().orc_is_inside_rect(w, h, C.c_double(tl_x), C.c_double(tl_y), C.c_double(region_size),
                                         C.c_double(sample_size)))


# ------------------------------------------------------------------ orientation / describe
def _xys(x, y, scale):
    return np.ascontiguousarray(np.stack([np.asarray(x, np.float64), np.asarray(y, np.float64),
                                          np.asarray(scale, np.float64)], axis=1))


def orientation(ii, x, y, scale, cfg=None):
    a, st = _img(ii, ii.dtype)
    xys = _xys(x, y, scale)
    out = np.zeros(len(xys), np.float64)
    f = lib().orc_orientation_f32 if a.dtype == np.float32 else lib().orc_orientation_s32
    f(_p(a), st, a.shape[1], a.shape[0], _p(_ori(cfg)), _p(xys), len(xys), _p(out))
    return out


def describe(ii, x, y, angle, scale, cfg=None):
    a, st = _img(ii, ii.dtype)
    xys = _xys(x, y, scale)
    ang = np.ascontiguousarray(angle, np.float64)
    dv = _desc(cfg)
    dof = int(dv[1]) ** 2 * 4
    desc = np.zeros((len(xys), dof), np.float64); white = np.zeros(len(xys), np.uint8)
    f = lib().orc_describe_f32 if a.dtype == np.float32 else lib().orc_describe_s32
    f(_p(a), st, a.shape[1], a.shape[0], _p(dv), _p(xys), _p(ang), len(xys), _p(desc), _p(white))
    return desc, white.astype(bool)


# ------------------------------------------------------------------ full pipeline
def surf(image, fh=None, ori=None, desc=None, nthreads=1, max_out=None):
    """WrapDetectDescribeSurf.detect.  ori/desc default to surfFast's (average orientation, DescribePointSurf)."""
    img = np.asarray(image)
    a, st = _img(img, img.dtype)
    h, w = a.shape
    iv, thr = _fh(fh)
    dv = _desc(desc)
    dof = int(dv[1]) ** 2 * 4
    cap = max_out or max(1024, w * h // 8)
    while True:
        xys = np.zeros((cap, 3), np.float64); angle = np.zeros(cap, np.float64)
        d = np.zeros((cap, dof), np.float64); white = np.zeros(cap, np.uint8)
        f = {np.dtype(np.float32): lib().orc_surf_f32, np.dtype(np.uint8): lib().orc_surf_u8}[a.dtype]
        n = f(_p(a), st, w, h, _p(iv), thr, _p(_ori(ori)), _p(dv), cap, _p(xys), _p(angle), _p(d), _p(white), nthreads)
        if n <= cap:
            break
        cap = n
    return dict(x=xys[:n, 0].copy(), y=xys[:n, 1].copy(), scale=xys[:n, 2].copy(), angle=angle[:n].copy(),
                desc=d[:n].copy(), white=white[:n].astype(bool))


def surf_fast(image, fh=None, nthreads=1):
    return surf(image, fh, ORI_AVERAGE, DESC_SPEED, nthreads)


def surf_stable(image, fh=None, nthreads=1):
    return surf(image, fh, ORI_SLIDING, DESC_STABILITY, nthreads)


# ------------------------------------------------------------------ association
def associate(src, dst, max_fit_error=np.finfo(np.float64).max, backwards=False, score="euclidean_sq", nthreads=1,
              nobuffer=False):
    """AssociateGreedy.associate.  Returns (pairs int32[N], fit float64[N])."""
    src = np.ascontiguousarray(src, np.float64); dst = np.ascontiguousarray(dst, np.float64)
    if src.ndim == 1: src = src[:, None]
    if dst.ndim == 1: dst = dst[:, None]
    n, dof = src.shape; m = dst.shape[0]
    pairs = np.full(n, -1, np.int32); fit = np.zeros(n, np.float64)
    if nobuffer:
        assert score == "euclidean_sq"
        lib().orc_associate_nobuffer(_p(src), n, _p(dst), m, dof, C.c_double(max_fit_error), int(backwards), _p(pairs),
                                     _p(fit), nthreads)
    else:
        lib().orc_associate(_p(src), n, _p(dst), m, dof, C.c_double(max_fit_error), int(backwards),
                            {"euclidean_sq": 0, "euclidean": 1}[score], _p(pairs), _p(fit), nthreads)
    return pairs, fit


def matches_from_pairs(pairs, fit, n_dst):
    """WrapAssociateGreedy.associate :85-92 + FindUnassociated.checkDestination :56-72."""
    src_idx = np.nonzero(pairs >= 0)[0].astype(np.int32)
    matches = [(int(i), int(pairs[i]), float(fit[i])) for i in src_idx]
    unassoc_src = np.nonzero(pairs < 0)[0].astype(np.int32)
    hit = np.zeros(n_dst, bool); hit[pairs[src_idx]] = True
    return matches, unassoc_src, np.nonzero(~hit)[0].astype(np.int32)


def associate_surf_basic(src, src_white, dst, dst_white, max_fit_error=np.finfo(np.float64).max, backwards=False):
    """AssociateSurfBasic.sort + associate (F/alg/feature/associate/AssociateSurfBasic.java:83-143) around the greedy
    oracle, WrapAssociateSurfBasic.getUnassociatedDestination (F/abst/feature/associate/WrapAssociateSurfBasic.java:71-73).
    Returns (matches [(src, dst, score)], unassociated_src, unassociated_dst) in the reference's order."""
    src = np.ascontiguousarray(src, np.float64); dst = np.ascontiguousarray(dst, np.float64)
    sw = np.asarray(src_white) != 0; dw = np.asarray(dst_white) != 0
    matches, unassoc = [], []
    if len(src) and len(dst):                                   # :89-92
        for cls in (True, False):                               # positive lists first (:94-107), then negative (:108-121)
            si = np.nonzero(sw == cls)[0]; di = np.nonzero(dw == cls)[0]
            # WrapAssociateGreedy.associate accepts empty lists: every source is then unassociated
            if len(si) == 0:
                continue
            if len(di) == 0:
                unassoc += [int(i) for i in si]
                continue
            pairs, fit = associate(src[si], dst[di], max_fit_error, backwards)
            m, us, _ = matches_from_pairs(pairs, fit, len(di))
            matches += [(int(si[a]), int(di[b]), f) for a, b, f in m]
            unassoc += [int(si[i]) for i in us]
    hit = np.zeros(len(dst), bool)
    for _, b, _ in matches:
        hit[b] = True
    return matches, np.asarray(unassoc, np.int32), np.nonzero(~hit)[0].astype(np.int32)


# ------------------------------------------------------------------ colour SURF (surfColorFast / surfColorStable)
def planar_average(bands):
    """ConvertImage.average(Planar, gray) (core/image/impl/ImplConvertPlanarToGray.java:44-84 U8, :296-336 F32)."""
    b = np.ascontiguousarray(bands)
    nb, h, w = b.shape
    out = np.zeros((h, w), b.dtype)
    f = {np.dtype(np.float32): lib().orc_planar_average_f32, np.dtype(np.uint8): lib().orc_planar_average_u8}[b.dtype]
    f(_p(b), nb, w, h, _p(out))
    return out


def describe_raw(ii, x, y, angle, scale, cfg=None):
    """DescribePointSurf.describe(x,y,angle,scale,TupleDesc_F64) :195-213: the sums only, not normalised."""
    a, st = _img(ii, ii.dtype)
    assert a.dtype == np.float32
    xys = _xys(x, y, scale)
    ang = np.ascontiguousarray(angle, np.float64)
    dv = _desc(cfg)
    desc = np.zeros((len(xys), int(dv[1]) ** 2 * 4), np.float64)
    lib().orc_describe_raw_f32(_p(a), st, a.shape[1], a.shape[0], _p(dv), _p(xys), _p(ang), len(xys), _p(desc))
    return desc


def laplace_sign(ii, x, y, scale, cfg=None):
    a, st = _img(ii, ii.dtype)
    return bool(lib().orc_laplace_sign_f32(_p(a), st, a.shape[1], a.shape[0], _p(_desc(cfg)), int(x), int(y), float(scale)))


def surf_planar(bands, fh=None, ori=None, desc=None, nthreads=1):
    """SurfPlanar_to_DetectDescribePoint.detect -> DetectDescribeSurfPlanar.detect -> DescribePointSurfPlanar.describe.
    bands: [nb][h][w] float32 or uint8.  desc rows hold nb*64 doubles."""
    b = np.ascontiguousarray(bands)
    nb, h, w = b.shape
    iv, thr = _fh(fh)
    dv = _desc(desc)
    dof = int(dv[1]) ** 2 * 4 * nb
    cap = max(1024, w * h // 8)
    while True:
        xys = np.zeros((cap, 3), np.float64); angle = np.zeros(cap, np.float64)
        d = np.zeros((cap, dof), np.float64); white = np.zeros(cap, np.uint8)
        f = {np.dtype(np.float32): lib().orc_surf_planar_f32, np.dtype(np.uint8): lib().orc_surf_planar_u8}[b.dtype]
        n = f(_p(b), nb, w, h, _p(iv), thr, _p(_ori(ori)), _p(dv), cap, _p(xys), _p(angle), _p(d), _p(white), nthreads)
        if n <= cap:
            break
        cap = n
    return dict(x=xys[:n, 0].copy(), y=xys[:n, 1].copy(), scale=xys[:n, 2].copy(), angle=angle[:n].copy(),
                desc=d[:n].copy(), white=white[:n].astype(bool))


def surf_color_fast(bands, fh=None, nthreads=1):
    return surf_planar(bands, fh, ORI_AVERAGE, DESC_SPEED, nthreads)


def surf_color_stable(bands, fh=None, nthreads=1):
    return surf_planar(bands, fh, ORI_SLIDING, DESC_STABILITY, nthreads)
