"""Synthetic inputs for tests and bench (there is no network for datasets; BASELINE.json's configs are all synthetic).

blob_image: deterministic multi-scale value noise — a weighted sum of bilinearly interpolated random lattices with cell
sizes 8,16,..,128 px (smooth blobs at several SURF scales), then a 5x5 box filter, range ~[45,210].  Every step is exact
int64 arithmetic followed by ONE int->float32 conversion and ONE float32 multiply, so the numpy (CPU) and torch (GPU)
generators produce bit-identical frames: the CPU baseline and the GPU path see the same pixels.  At 1920x1080 the default
Fast-Hessian configuration finds ~10k features per frame (5k/4k/1.2k/0.3k in octaves 0..3).

descriptors: unit-norm 64-d Gaussian vectors, optionally with planted near-duplicates.
"""
import numpy as np

_CELLS = (8, 16, 32, 64, 128)
_WEIGHTS = (2, 3, 4, 4, 4)
_CMAX = max(_CELLS)
_BOX = 5
_SCALE = np.float32(1.0 / (_BOX * _BOX * _CMAX * _CMAX * sum(_WEIGHTS)))
_OFF = 1024          # keeps lattice coordinates positive


def _hash_np(v):
    """32-bit integer finaliser (xorshift-multiply) on uint64 lanes masked to 32 bits -> int64 in [0,255]."""
    m = np.uint64(0xFFFFFFFF)
    v = v & m
    v ^= v >> np.uint64(16); v = (v * np.uint64(0x7FEB352D)) & m
    v ^= v >> np.uint64(15); v = (v * np.uint64(0x846CA68B)) & m
    v ^= v >> np.uint64(16)
    return (v >> np.uint64(24)).astype(np.int64)


def blob_image(width, height, seed=234):
    """float32 [height, width] frame; identical to blob_images_torch(...)[i] for seed+i."""
    p = _BOX // 2
    ys = np.arange(-p, height + p, dtype=np.int64)[:, None] + _OFF
    xs = np.arange(-p, width + p, dtype=np.int64)[None, :] + _OFF
    raw = np.zeros((height + 2 * p, width + 2 * p), np.int64)

    def lattice(cy, cx, k):
        return _hash_np((cy * 65537 + cx * 257 + k * 1000003 + int(seed) * 7919).astype(np.uint64))

    for k, (c, wgt) in enumerate(zip(_CELLS, _WEIGHTS)):
        cy, fy = ys // c, ys % c
        cx, fx = xs // c, xs % c
        v = ((c - fy) * (c - fx) * lattice(cy, cx, k) + (c - fy) * fx * lattice(cy, cx + 1, k)
             + fy * (c - fx) * lattice(cy + 1, cx, k) + fy * fx * lattice(cy + 1, cx + 1, k))
        raw += (wgt * (_CMAX * _CMAX // (c * c))) * v
    cs = np.cumsum(np.cumsum(np.pad(raw, ((1, 0), (1, 0))), 0), 1)
    box = cs[_BOX:, _BOX:] - cs[:-_BOX, _BOX:] - cs[_BOX:, :-_BOX] + cs[:-_BOX, :-_BOX]
    return box.astype(np.float32) * _SCALE


def blob_images_torch(width, height, n, seed=234, device="cuda", chunk=8):
    """[n, height, width] float32 frames on `device`; frame i == blob_image(width, height, seed + i) bit for bit."""
    import torch
    M = 0xFFFFFFFF
    p = _BOX // 2
    out = torch.empty((n, height, width), device=device, dtype=torch.float32)
    ys = (torch.arange(-p, height + p, device=device, dtype=torch.int64) + _OFF)[None, :, None]
    xs = (torch.arange(-p, width + p, device=device, dtype=torch.int64) + _OFF)[None, None, :]

    def lattice(cy, cx, k, sd):
        v = (cy * 65537 + cx * 257 + k * 1000003 + sd) & M
        v = v ^ (v >> 16); v = (v * 0x7FEB352D) & M
        v = v ^ (v >> 15); v = (v * 0x846CA68B) & M
        v = v ^ (v >> 16)
        return v >> 24

    for i0 in range(0, n, chunk):
        m = min(chunk, n - i0)
        sd = (torch.arange(i0, i0 + m, device=device, dtype=torch.int64) + int(seed))[:, None, None] * 7919
        raw = torch.zeros((m, height + 2 * p, width + 2 * p), device=device, dtype=torch.int64)
        for k, (c, wgt) in enumerate(zip(_CELLS, _WEIGHTS)):
            cy, fy = torch.div(ys, c, rounding_mode="floor"), ys % c
            cx, fx = torch.div(xs, c, rounding_mode="floor"), xs % c
            v = ((c - fy) * (c - fx) * lattice(cy, cx, k, sd) + (c - fy) * fx * lattice(cy, cx + 1, k, sd)
                 + fy * (c - fx) * lattice(cy + 1, cx, k, sd) + fy * fx * lattice(cy + 1, cx + 1, k, sd))
            raw += (wgt * (_CMAX * _CMAX // (c * c))) * v
        cs = torch.nn.functional.pad(raw, (1, 0, 1, 0)).cumsum(1).cumsum(2)
        box = cs[:, _BOX:, _BOX:] - cs[:, :-_BOX, _BOX:] - cs[:, _BOX:, :-_BOX] + cs[:, :-_BOX, :-_BOX]
        out[i0:i0 + m] = box.to(torch.float32) * float(_SCALE)
    return out


def shifted_pair(width, height, dx=7, dy=3, seed=234):
    """(A, B) with B = A shifted by (dx, dy): B[y, x] = A'[y + dy, x + dx] cropped from one larger frame (config 1)."""
    big = blob_image(width + dx, height + dy, seed)
    return np.ascontiguousarray(big[:height, :width]), np.ascontiguousarray(big[dy:dy + height, dx:dx + width])


def descriptors(n, dof=64, seed=1234):
    v = np.random.default_rng(seed).standard_normal((n, dof))
    return v / np.sqrt((v * v).sum(1, keepdims=True))


def planted_descriptors(n, dof=64, seed=1234, noise=0.05):
    """(src, dst, perm): dst[perm[i]] = normalise(src[i] + noise*N(0,1)) so true matches exist."""
    r = np.random.default_rng(seed + 1)
    src = descriptors(n, dof, seed)
    perm = r.permutation(n)
    d = src + noise * r.standard_normal((n, dof))
    d /= np.sqrt((d * d).sum(1, keepdims=True))
    dst = np.empty_like(d); dst[perm] = d
    return src, dst, perm
