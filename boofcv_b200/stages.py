"""Stage-level calls through the C ABI (host arrays in, host arrays out).  Each runs exactly the kernel the full pipeline
uses; the parity tests compare them with the oracle stage by stage.  Names follow the reference:
GIntegralImageOps.transform, IntegralImageFeatureIntensity.hessian, FastHessianFeatureDetector.detect,
OrientationIntegral.compute, DescribePointSurf.describe."""
import ctypes as C

import numpy as np

from . import _native as N

FH_DEFAULT = dict(detect_threshold=1.0, extract_radius=2, max_features_per_scale=-1, initial_sample_size=1,
                  initial_size=9, scales_per_octave=4, octaves=4, scale_step_size=6)
ORI_AVERAGE = dict(variant=0, object_radius_to_scale=0.5, radius=6, sample_period=1.0, sample_width=6,
                   weight_sigma=-1.0, window_size=np.pi / 3.0)
ORI_SLIDING = dict(variant=1, object_radius_to_scale=0.5, radius=8, sample_period=0.65, sample_width=6,
                   weight_sigma=-1.0, window_size=np.pi / 3.0)
DESC_SPEED = dict(variant=0, width_large_grid=4, width_sub_region=5, width_sample=3, use_haar=0, weight_sigma=4.5,
                  overlap=2, sigma_large_grid=2.5, sigma_sub_region=2.5)
DESC_STABILITY = dict(DESC_SPEED, variant=1)

_PIXEL = {np.dtype(np.float32): N.PIXEL_F32, np.dtype(np.uint8): N.PIXEL_U8, np.dtype(np.int32): N.PIXEL_S32}


def fh_cfg(cfg=None):
    c = dict(FH_DEFAULT, **(cfg or {}))
    return N.FhCfg(**c)


def ori_cfg(cfg=None):
    c = dict(ORI_AVERAGE, **(cfg or {}))
    return N.OriCfg(**c)


def desc_cfg(cfg=None):
    c = dict(DESC_SPEED, **(cfg or {}))
    return N.DescCfg(**c)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def pixel_type(dtype):
    try:
        return _PIXEL[np.dtype(dtype)]
    except KeyError:
        raise ValueError("Unknown input type: %s" % dtype)     # GIntegralImageOps.getIntegralType


def _strided(img):
    """(array, stride in elements) of a 2-D row-strided view (BoofCV sub-image); copies only if columns are strided."""
    a = np.asarray(img)
    if a.ndim != 2:
        raise ValueError("expected a 2-D image")
    if a.size and (a.strides[1] != a.itemsize or a.strides[0] % a.itemsize or a.strides[0] < a.shape[1] * a.itemsize):
        a = np.ascontiguousarray(a)
    return a, (a.strides[0] // a.itemsize if a.size else max(a.shape[1], 1))


def _ii(ii):
    a = np.ascontiguousarray(ii)
    if a.dtype not in (np.float32, np.int32) or a.ndim != 2:
        raise ValueError("integral image must be 2-D float32 or int32")
    return a, (N.PIXEL_F32 if a.dtype == np.float32 else N.PIXEL_S32)


def integral(img, device=0):
    a, stride = _strided(img)
    pt = pixel_type(a.dtype)
    h, w = a.shape
    out = np.empty((h, w), np.float32 if pt == N.PIXEL_F32 else np.int32)
    if a.size == 0:
        return out
    N.check(N.lib().bs_integral(device, pt, _p(a), stride, w, h, _p(out)))
    return out


def hessian(ii, skip, size, device=0):
    a, t = _ii(ii)
    h, w = a.shape
    out = np.zeros((h // skip, w // skip), np.float32)
    N.check(N.lib().bs_hessian(device, t, _p(a), w, h, skip, size, _p(out)))
    return out


def nonmax(intensity, radius, threshold, ignore_border=0, device=0):
    """NonMaxBlock + NonMaxBlockSearchStrict.Max on one intensity image (bs_nonmax): int16 (x, y) in the reference's order."""
    a = np.asarray(intensity, np.float32)
    if a.ndim != 2:
        raise ValueError("intensity must be [height][width]")
    if a.strides[1] != 4:
        a = np.ascontiguousarray(a)
    h, w = a.shape
    stride = a.strides[0] // 4 if h > 1 else max(w, 1)
    cap = max(1, w * h)
    xy = np.zeros((cap, 2), np.int16)
    n = C.c_int(0)
    N.check(N.lib().bs_nonmax(device, _p(a), stride, w, h, int(radius), C.c_float(threshold), int(ignore_border), cap, _p(xy), C.byref(n)))
    return xy[:n.value].copy()


def fh_intensity(ii, cfg=None, device=0):
    """Every layer of the Fast-Hessian scale space, computed by the kernels the batch path uses (FastHessianFeatureDetector
    .detectOctave's intensity[] images).  Returns a list of dicts {size, skip, intensity[h,w]} in (octave, scale) order."""
    a, t = _ii(ii)
    h, w = a.shape
    c = fh_cfg(cfg)
    nl, nf = C.c_int(0), C.c_long(0)
    layers = np.zeros((64, 5), np.int64)
    N.check(N.lib().bs_fh_intensity(device, t, _p(a), w, h, C.byref(c), C.byref(nl), _p(layers), C.byref(nf), None, 0))
    out = np.zeros(max(nf.value, 1), np.float32)
    N.check(N.lib().bs_fh_intensity(device, t, _p(a), w, h, C.byref(c), C.byref(nl), _p(layers), C.byref(nf), _p(out), out.size))
    res = []
    for k in range(nl.value):
        size, skip, lw, lh, off = (int(v) for v in layers[k])
        res.append(dict(size=size, skip=skip, intensity=out[off:off + lw * lh].reshape(lh, lw).copy()))
    return res


def fh_detect(ii, cfg=None, device=0, max_out=None):
    a, t = _ii(ii)
    h, w = a.shape
    cap = max_out or max(1024, w * h // 8)
    c = fh_cfg(cfg)
    while True:
        xys = np.zeros((cap, 3), np.float64); meta = np.zeros((cap, 4), np.int32); resp = np.zeros(cap, np.float32)
        n = C.c_int(0)
        N.check(N.lib().bs_fh_detect(device, t, _p(a), w, h, C.byref(c), cap, C.byref(n), _p(xys), _p(meta), _p(resp)))
        if n.value <= cap:
            break
        cap = n.value
    n = n.value
    return dict(x=xys[:n, 0].copy(), y=xys[:n, 1].copy(), scale=xys[:n, 2].copy(), octave=meta[:n, 0].copy(),
                level=meta[:n, 1].copy(), px=meta[:n, 2].copy(), py=meta[:n, 3].copy(), response=resp[:n].copy())


def _xys(x, y, scale):
    return np.ascontiguousarray(np.stack([np.asarray(x, np.float64), np.asarray(y, np.float64),
                                          np.asarray(scale, np.float64)], axis=1))


def orientation(ii, x, y, scale, cfg=None, device=0):
    a, t = _ii(ii)
    xys = _xys(x, y, scale)
    out = np.zeros(len(xys), np.float64)
    c = ori_cfg(cfg)
    N.check(N.lib().bs_orientation(device, t, _p(a), a.shape[1], a.shape[0], C.byref(c), _p(xys), len(xys), _p(out)))
    return out


def describe(ii, x, y, angle, scale, cfg=None, device=0):
    a, t = _ii(ii)
    xys = _xys(x, y, scale)
    ang = np.ascontiguousarray(angle, np.float64)
    c = desc_cfg(cfg)
    dof = c.width_large_grid ** 2 * 4
    desc = np.zeros((len(xys), dof), np.float64); white = np.zeros(len(xys), np.uint8)
    N.check(N.lib().bs_describe(device, t, _p(a), a.shape[1], a.shape[0], C.byref(c), _p(xys), _p(ang), len(xys),
                                _p(desc), _p(white)))
    return desc, white.astype(bool)
